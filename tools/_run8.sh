cd $GRAFT_REPO_ROOT
N=$(nvidia-smi -L | wc -l)
echo "gpus: $N" > gpurun_out/t8.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu 2>&1 | grep '^{' | cut -c1-260 >> gpurun_out/t8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 tools/banded_dist.py 16384 16384 2>&1 | grep "^{" >> gpurun_out/t8.log
cat gpurun_out/t8.log
