"""CPU-side tests: tables against the reference's own host code output (golden, dumped from the
reference build), geometry, config plumbing, the C-ABI surface, and the N>1 bench plumbing."""
import ctypes, os, re, subprocess, sys
import numpy as np
import pytest
import refio
from popsift_b200 import capi
from tools import golden_cases as gc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLE_CASES = [c for c, (_, _, planes) in gc.CASES.items() if planes and os.path.exists(os.path.join(refio.GOLD, "ref", c + ".tables.bin"))]


@pytest.mark.parametrize("case", TABLE_CASES)
def test_gauss_tables_bitexact_vs_reference(case):
    """All four Gauss table families and spans, bit for bit (reference gauss_filter.cu:127-257)."""
    t = refio.RefTables(os.path.join(refio.GOLD, "ref", case + ".tables.bin"))
    cfg = capi.config_from_flags(gc.config_flags(gc.CASES[case][1]))
    for which, name in enumerate(["inc", "abs_o0", "abs_oN", "dd"]):
        ref = getattr(t, name)
        for l in range(ref["filter"].shape[0]):
            f, s, sp = capi.gauss_table(cfg, which, l)
            assert np.array_equal(f.view(np.uint32), ref["filter"][l].view(np.uint32)), (name, l)
            assert np.float32(s).view(np.uint32) == ref["sigma"][l].view(np.uint32), (name, l)
            assert sp == ref["span"][l], (name, l)


@pytest.mark.parametrize("case", TABLE_CASES)
def test_desc_tables_and_scalars_vs_reference(case):
    t = refio.RefTables(os.path.join(refio.GOLD, "ref", case + ".tables.bin"))
    g, tile, sc = capi.desc_tables(capi.config_from_flags(gc.config_flags(gc.CASES[case][1])))
    assert np.array_equal(g.view(np.uint32), t.desc_gauss.view(np.uint32))
    assert np.array_equal(tile.view(np.uint32), t.desc_tile.view(np.uint32))
    assert sc[0] == np.float32(t.sigma_k) and sc[1] == np.float32(t.threshold)


def test_default_spans_match_survey_crib():
    """default spans 6 | 6,8,9,11,14; threshold 1.7."""
    cfg = capi.default_config()
    assert [capi.gauss_table(cfg, 0, l)[2] for l in range(6)] == [6, 6, 8, 9, 11, 14]
    assert capi.gauss_table(capi.default_config(downsampling=0), 0, 0)[2] == 8
    assert [capi.gauss_table(capi.default_config(gauss_mode="opencv"), 0, l)[2] for l in range(6)] == [6, 6, 7, 9, 11, 14]
    assert abs(capi.desc_tables(cfg)[2][1] - 1.7) < 1e-6


@pytest.mark.parametrize("w,h,up,expect", [(640, 480, 1, (1280, 960, 7)), (1920, 1080, 1, (3840, 2160, 9)),
                                            (3840, 2160, 1, (7680, 4320, 10)), (16384, 16384, 0, (16384, 16384, 12)),
                                            (64, 48, 1, (128, 96, 4)), (17, 17, 1, (34, 34, 3)), (256, 192, -1, (128, 96, 4))])
def test_geometry_matches_reference_formula(w, h, up, expect):
    """octave-0 size and octave count (reference popsift.cpp:97-115);
    the small cases are what the reference build reported on B200 (tests/golden/ref/log.json)."""
    assert capi.geometry(capi.default_config(downsampling=-up), w, h) == expect


def test_golden_headers_agree_with_geometry():
    for case, (img, flags, _) in gc.CASES.items():
        p = os.path.join(refio.GOLD, "ref", case + ".feat.bin")
        if not os.path.exists(p): continue
        r = refio.RefFeatures(p)
        w, h, _ = gc.IMAGES[img]
        assert capi.geometry(capi.config_from_flags(gc.config_flags(flags)), w, h) == (r.w0, r.h0, r.n_octaves), case
        assert r.ext_ct.sum() == r.n_feat and r.ori_ct.sum() == r.n_desc


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "psift.h")).read()
    declared = set(re.findall(r"\b(psift_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)
    L = ctypes.CDLL(capi.LIB_PATH)
    for name in declared:
        assert getattr(L, name) is not None


def test_struct_sizes_match_header():
    assert capi.FEATURE_DTYPE.itemsize == 52 and capi.EXTREMUM_DTYPE.itemsize == 44
    assert ctypes.sizeof(capi.Config) == 19 * 4


def test_no_device_fails_loudly():
    """Without a GPU the product path must fail, never fall back (there is no CPU path)."""
    if capi.lib().psift_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.PsiftError):
        capi.Context(capi.default_config())


def test_bad_config_rejected():
    for kw in ({"gauss_mode": "fixed9"}, {"sigma": 2.5}, {"levels": 12}):
        cfg = capi.default_config(**kw)
        f = np.zeros(32, np.float32)
        if "gauss_mode" in kw:
            continue   # tables exist for every span mode; the pipeline rejects it at create time
        rc = capi.lib().psift_debug_gauss_table(ctypes.byref(cfg), 0, 0, f.ctypes.data, None, None)
        assert rc != 0


def test_shard_is_a_partition():
    from popsift_b200 import dist as pdist
    for n in (0, 1, 7, 1000):
        for world in (1, 2, 3, 8):
            spans = [pdist.shard(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1


_GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import torch.distributed as dist
from popsift_b200 import dist as pdist
rank = int(os.environ["RANK"])
dist.init_process_group("gloo")
pdist.barrier()
pdist.cpu_barrier()                        # the CPU-side wait used while rank 0 alone drives all GPUs
m = pdist.max_over_ranks(1.0 + rank)       # elapsed time: max over ranks
s = pdist.sum_over_ranks(10 * (rank + 1))  # frames: sum over ranks
b, e = pdist.shard(101, rank, 2)
tot = pdist.sum_over_ranks(e - b)
assert m == 2.0 and s == 30.0 and tot == 101.0, (m, s, tot)
dist.destroy_process_group()
print("ok", rank)
"""


def test_two_rank_gloo_timing_and_sharding(tmp_path):
    """The N>1 plumbing of bench.py on CPU: world_size 2 over gloo, rendezvous on 127.0.0.1."""
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", str(script)],
                       capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


def test_band_plan_partitions_rows():
    """psift_band_plan (host logic of the banded mode): bands partition the rows of octave 0, start on
    multiples of 2^(split+1), and the seed rows the bands contribute partition the seed plane."""
    from popsift_b200 import banded
    cfg = capi.default_config(downsampling=0)
    for (w, h, parts) in [(16384, 16384, 8), (8192, 8192, 2), (4000, 3001, 3), (1024, 1536, 4)]:
        split, b = capi.band_plan(cfg, w, h, parts)
        w0, h0, n_oct = capi.geometry(cfg, w, h)
        assert b[0] == 0 and b[-1] == h0 and all(b[i] < b[i + 1] for i in range(parts))
        assert all(v % (1 << (split + 1)) == 0 for v in b[:-1])
        assert 0 <= split < n_oct - 1
        sh = h0
        for _ in range(split + 1): sh = (sh + 1) // 2
        rows = [banded._own_seed_rows(b, r, split, sh) for r in range(parts)]
        assert rows[0][0] == 0 and rows[-1][1] == sh and all(rows[i][1] == rows[i + 1][0] for i in range(parts - 1))
    split, _ = capi.band_plan(cfg, 16384, 16384, 8)
    assert split == 2                                   # octave 3 of a 16384^2 image is 2048^2: finished un-split


def test_demo_cli_fails_loudly_without_gpu(tmp_path):
    """popsift-demo (apps/popsift_demo.cpp) parses the reference's flags; with no CUDA device it must say so and
    fail, not fall back to anything."""
    import subprocess
    exe = os.path.join(ROOT, "popsift_b200", "popsift-demo")
    if not os.path.exists(exe): pytest.skip("popsift-demo not built")
    r = subprocess.run([exe, "--no-such-flag"], capture_output=True, text=True)
    assert r.returncode != 0 and "unrecognised option" in r.stderr
    if capi.lib().psift_device_count() > 0: return
    p = tmp_path / "a.pgm"; p.write_bytes(b"P5\n4 4\n255\n" + bytes(16))
    r = subprocess.run([exe, "-i", str(p), "--desc-mode", "grid"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode != 0 and "no CUDA device" in r.stderr


def test_cmake_package_installs_and_is_found(tmp_path):
    """The drop-in's packaging (reference src/CMakeLists.txt, src/cmake/Config.cmake.in:33-43): the project installs
    libpopsift + PopSiftConfig.cmake, and a caller that only says find_package(PopSift) / PopSift::popsift builds
    against the installed headers and library (nvcc cross-compiles sm_100a here; nothing is run)."""
    import shutil
    if not shutil.which("cmake"): pytest.skip("cmake not installed")
    b, inst, cons = tmp_path / "b", tmp_path / "inst", tmp_path / "cons"
    def run(*cmd):
        r = subprocess.run(list(cmd), capture_output=True, text=True, timeout=1500)
        assert r.returncode == 0, " ".join(cmd) + "\n" + r.stdout[-1500:] + r.stderr[-1500:]
    run("cmake", "-S", ROOT, "-B", str(b), "-DCMAKE_BUILD_TYPE=Release")
    run("cmake", "--build", str(b), "-j8")
    run("cmake", "--install", str(b), "--prefix", str(inst))
    assert (inst / "lib" / "cmake" / "PopSift" / "PopSiftConfig.cmake").exists()
    assert (inst / "include" / "popsift" / "popsift.h").exists() and (inst / "include" / "psift.h").exists()
    assert any(f.name.startswith("libpopsift.so") for f in (inst / "lib").iterdir())
    run("cmake", "-S", os.path.join(ROOT, "tests", "cmake_consumer"), "-B", str(cons), "-DCMAKE_PREFIX_PATH=" + str(inst))
    run("cmake", "--build", str(cons))
    assert (cons / "consumer").exists()
