"""CPU tests of the matcher's checker (no GPU): oracle/cpu_match (the CPU restatement of the reference's
compute_distance, features.cu:163-225) against the goldens produced by the reference's own kernel on a B200
(tests/golden/ref/*.match.bin), and against an independent float64 brute force where the answer is unambiguous."""
import os, subprocess
import numpy as np
import pytest
import refio
from tools import golden_cases as gc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPU_MATCH = os.path.join(ROOT, "oracle", "cpu_match")
GOLDEN = [c for c in gc.MATCH_CASES if os.path.exists(os.path.join(refio.GOLD, "ref", c + ".match.bin"))]
pytestmark = pytest.mark.skipif(not os.path.exists(CPU_MATCH), reason="oracle/cpu_match not built (make -C oracle)")


def run(tmp_path, l, r):
    lf, rf, of = tmp_path / "l.f32", tmp_path / "r.f32", tmp_path / "m.bin"
    np.ascontiguousarray(l, "<f4").tofile(lf); np.ascontiguousarray(r, "<f4").tofile(rf)
    subprocess.run([CPU_MATCH, str(lf), str(rf), str(of)], check=True, timeout=300)
    return np.fromfile(of, "<i4").reshape(-1, 3)


@pytest.mark.parametrize("case", GOLDEN)
def test_cpu_match_equals_reference_kernel(tmp_path, case):
    lc, rc = gc.MATCH_CASES[case]
    desc = lambda c: refio.RefFeatures(os.path.join(refio.GOLD, "ref", c + ".feat.bin")).desc
    want = np.fromfile(os.path.join(refio.GOLD, "ref", case + ".match.bin"), "<i4").reshape(-1, 3)
    got = run(tmp_path, desc(lc), desc(rc))
    assert np.array_equal(got, want), np.flatnonzero((got != want).any(1))[:10]


def test_cpu_match_against_float64_brute_force(tmp_path):
    rng = np.random.default_rng(11)
    l = rng.random((90, 128)).astype(np.float32); r = rng.random((130, 128)).astype(np.float32)
    got = run(tmp_path, l, r)
    d = ((l[:, None, :].astype(np.float64) - r[None, :, :].astype(np.float64)) ** 2).sum(2)
    order = np.argsort(d, axis=1, kind="stable")
    assert np.array_equal(got[:, 0], order[:, 0]) and np.array_equal(got[:, 1], order[:, 1])
    ratio = d[np.arange(90), order[:, 0]] / d[np.arange(90), order[:, 1]]
    clear = np.abs(ratio - 0.8) > 1e-4
    assert np.array_equal(got[clear, 2], (ratio[clear] < 0.8).astype(np.int32))


def test_cpu_match_edge_cases(tmp_path):
    one = np.ones((1, 128), np.float32)
    assert run(tmp_path, one, np.zeros((0, 128), np.float32)).tolist() == [[0, 0, 0]]      # inf/inf is not < 0.8
    assert run(tmp_path, one, one * 2).tolist() == [[0, 0, 1]]                              # d/inf = 0
    assert run(tmp_path, np.zeros((0, 128), np.float32), one).shape == (0, 3)
    r = np.stack([one[0] * 3, one[0] * 2, one[0] * 2])                                      # tie: lower index wins
    assert run(tmp_path, one, r).tolist() == [[1, 2, 0]]
