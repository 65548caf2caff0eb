"""GPU parity tests of the matcher and the device-resident results: psift_match /
psift_wait_dev through the C ABI against
  (1) the reference's own compute_distance kernel (features.cu:184-225) run on a B200 on the same descriptor sets
      (tests/golden/ref/*.match.bin, made by tools/run_ref_golden.py), integers bit-exact,
  (2) the CPU restatement oracle/cpu_match on seeded random sets with duplicates and ragged sizes,
  (3) size-independent properties at full size (permutation round trip, self match)."""
import os, subprocess
import numpy as np
import pytest
import refio
from popsift_b200 import capi, synth
from tools import golden_cases as gc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPU_MATCH = os.path.join(ROOT, "oracle", "cpu_match")
GOLDEN = [c for c in gc.MATCH_CASES if os.path.exists(os.path.join(refio.GOLD, "ref", c + ".match.bin"))]


def golden_desc(case):
    return refio.RefFeatures(os.path.join(refio.GOLD, "ref", case + ".feat.bin")).desc


def gpu_match(l, r):
    with capi.DevFeatures.from_host(l) as a, capi.DevFeatures.from_host(r) as b:
        return a.match(b)


@pytest.mark.parametrize("case", GOLDEN)
def test_match_equals_reference_kernel(case):
    lc, rc = gc.MATCH_CASES[case]
    l, r = golden_desc(lc), golden_desc(rc)
    want = np.fromfile(os.path.join(refio.GOLD, "ref", case + ".match.bin"), "<i4").reshape(-1, 3)
    m, d = gpu_match(l, r)
    assert m.shape == want.shape
    assert np.array_equal(m, want), np.flatnonzero((m != want).any(1))[:10]


def cpu_match(tmp_path, l, r):
    lf, rf, of = tmp_path / "l.f32", tmp_path / "r.f32", tmp_path / "m.bin"
    l.astype("<f4").tofile(lf); r.astype("<f4").tofile(rf)
    subprocess.run([CPU_MATCH, str(lf), str(rf), str(of)], check=True, timeout=300)
    return np.fromfile(of, "<i4").reshape(-1, 3)


def rootsift_like(rng, n):
    v = rng.gamma(0.6, 1.0, (n, 128)).astype(np.float32)
    v = np.sqrt(v / v.sum(1, keepdims=True)).astype(np.float32)
    return v


@pytest.mark.skipif(not os.path.exists(CPU_MATCH), reason="oracle/cpu_match not built")
@pytest.mark.parametrize("nl,nr,seed", [(1, 1, 1), (33, 65, 2), (257, 31, 3), (700, 900, 4), (64, 64, 5)])
def test_match_equals_cpu_oracle(tmp_path, nl, nr, seed):
    rng = np.random.default_rng(seed)
    l, r = rootsift_like(rng, nl), rootsift_like(rng, nr)
    if nr > 8:                                            # duplicates on the right: ties must go to the lower index
        r[nr // 2] = r[1]; r[nr - 1] = r[1]
    if nl > 4 and nr > 4:
        l[3] = r[1]                                       # exact hit on a duplicated descriptor
        l[nl - 1] = r[nr - 2] + np.float32(1e-4)
    m, d = gpu_match(l, r)
    want = cpu_match(tmp_path, l, r)
    assert np.array_equal(m, want), np.flatnonzero((m != want).any(1))[:10]
    if nl > 4 and nr > 8:
        assert m[3, 0] == 1 and m[3, 1] == nr // 2 and d[3, 0] == 0.0 and d[3, 1] == 0.0


def test_match_full_size_properties():
    """25 000 x 25 000 descriptors (a 4K frame's worth): right = permuted left."""
    rng = np.random.default_rng(9)
    n = 25000
    l = rootsift_like(rng, n)
    perm = rng.permutation(n)
    r = l[perm]
    m, d = gpu_match(l, r)
    inv = np.empty(n, np.int64); inv[perm] = np.arange(n)
    assert np.array_equal(m[:, 0], inv)                   # every left descriptor finds itself
    assert np.all(d[:, 0] == 0.0) and np.all(d[:, 1] > 0.0) and np.all(m[:, 2] == 1)
    assert np.all(m[:, 1] != m[:, 0])


def test_device_resident_result_equals_host_result():
    """psift_wait_dev (reference clone_device_descriptors, sift_pyramid.cu:323-361): same features, descriptors and
    descriptor->feature map as the host result; the 72-byte reference-layout records point into the device block."""
    img = synth.gen(640, 480, 3)
    with capi.Context(capi.default_config()) as ctx:
        host = ctx.extract(img)
        with ctx.extract_dev(img) as dev:
            assert (dev.n_feat, dev.n_desc) == (host.n_feat, host.n_desc)
            assert np.array_equal(dev.features(), host.feat)
            assert np.array_equal(dev.descriptors(), host.desc)
            rev = dev.reverse_map()
            for i in (0, host.n_feat // 2, host.n_feat - 1):
                for k in range(host.feat["num_ori"][i]):
                    assert rev[host.feat["desc_idx"][i][k]] == i
            ref = dev.features_ref_layout()
            assert ref[:, :36].tobytes() == np.ascontiguousarray(host.feat.view(np.uint8).reshape(host.n_feat, -1)[:, :36]).tobytes()
            ptr = ref[:, 40:72].copy().view("<u8").reshape(-1, 4)
            idx = host.feat["desc_idx"]
            base = dev.info.desc
            assert np.array_equal(ptr[idx >= 0], base + 512 * idx[idx >= 0].astype(np.uint64))
            assert np.all(ptr[idx < 0] == 0)
            # two images of the same scene match mostly onto each other
            img2 = np.roll(img, 3, axis=1)
            with ctx.extract_dev(img2) as dev2:
                m, d = dev.match(dev2)
                assert m[:, 2].mean() > 0.5


def test_cpp_matching_mode(tmp_path):
    """PopSift(config, MatchingMode) + SiftJob::getDev + FeaturesDev::match (reference src/application/match.cpp:255-268)."""
    exe = os.path.join(ROOT, "tests", "cpp", "match_demo")
    if not os.path.exists(exe): pytest.skip("tests/cpp/match_demo not built")
    wa, ha, _ = gc.IMAGES["g256"]; wb, hb, _ = gc.IMAGES["g640"]
    out = tmp_path / "m.bin"
    r = subprocess.run([exe, gc.image_path("g256"), str(wa), str(ha), gc.image_path("g640"), str(wb), str(hb), str(out)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-800:]
    b = out.read_bytes()
    nl, nr = np.frombuffer(b, "<i4", 2)
    m = np.frombuffer(b, "<i4", 3 * nl, 8).reshape(nl, 3)
    a, c = golden_desc("g256_default"), golden_desc("g640_default")
    assert (nl, nr) == (len(a), len(c))
    img_a = np.fromfile(gc.image_path("g256"), np.uint8).reshape(ha, wa)
    img_b = np.fromfile(gc.image_path("g640"), np.uint8).reshape(hb, wb)
    with capi.Context(capi.default_config()) as ctx, ctx.extract_dev(img_a) as da, ctx.extract_dev(img_b) as db:
        want, _ = da.match(db)
    assert np.array_equal(m, want)
    lines = [l for l in r.stdout.splitlines() if l.startswith(("accept", "reject"))]
    assert len(lines) == nl
    assert sum(l.startswith("accept") for l in lines) == int(m[:, 2].sum())
