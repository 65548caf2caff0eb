"""Readers for the binary files written by oracle/ref_driver.cu (reference CUDA build) and the
canonical ordering used to compare feature lists.  TEST INFRASTRUCTURE."""
import os
import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

FEATREC = np.dtype([("octave", "<i4"), ("x", "<f4"), ("y", "<f4"), ("sigma", "<f4"), ("num_ori", "<i4"),
                    ("ori", "<f4", (4,)), ("desc_idx", "<i4", (4,))])
EXTREMUM = np.dtype([("x", "<f4"), ("y", "<f4"), ("lpos", "<i4"), ("sigma", "<f4"), ("octave", "<i4"),
                     ("num_ori", "<i4"), ("idx_ori", "<i4"), ("ori", "<f4", (4,))])


class RefFeatures:
    def __init__(self, path):
        b = open(path, "rb").read()
        hdr = np.frombuffer(b, "<i4", 6, 0)
        assert hdr[0] == 0x54465350, "bad magic"
        self.n_octaves, self.n_levels, self.w0, self.h0 = int(hdr[2]), int(hdr[3]), int(hdr[4]), int(hdr[5])
        off = 24
        self.ext_ct = np.frombuffer(b, "<i4", 20, off).copy(); off += 80
        self.ori_ct = np.frombuffer(b, "<i4", 20, off).copy(); off += 80
        self.n_feat, self.n_desc = [int(v) for v in np.frombuffer(b, "<i4", 2, off)]; off += 8
        self.feat = np.frombuffer(b, FEATREC, self.n_feat, off).copy(); off += FEATREC.itemsize * self.n_feat
        self.desc = np.frombuffer(b, "<f4", self.n_desc * 128, off).reshape(self.n_desc, 128).copy(); off += 512 * self.n_desc
        self.ext = np.frombuffer(b, EXTREMUM, self.n_feat, off).copy(); off += EXTREMUM.itemsize * self.n_feat
        assert off == len(b), (off, len(b))


class RefPlanes:
    def __init__(self, path):
        b = open(path, "rb").read()
        hdr = np.frombuffer(b, "<i4", 3, 0)
        assert hdr[0] == 0x4e4c5050
        self.n_octaves, self.n_levels = int(hdr[1]), int(hdr[2])
        off = 12
        self.gauss, self.dog, self.size = [], [], []
        for o in range(self.n_octaves):
            w, h = [int(v) for v in np.frombuffer(b, "<i4", 2, off)]; off += 8
            self.size.append((w, h))
            g = np.frombuffer(b, "<f4", self.n_levels * w * h, off).reshape(self.n_levels, h, w).copy(); off += 4 * self.n_levels * w * h
            d = np.frombuffer(b, "<f4", (self.n_levels - 1) * w * h, off).reshape(self.n_levels - 1, h, w).copy(); off += 4 * (self.n_levels - 1) * w * h
            self.gauss.append(g); self.dog.append(d)
        assert off == len(b)


def _gauss_table(b, off, L):
    n = L * 32
    t = {"filter": np.frombuffer(b, "<f4", n, off).reshape(L, 32).copy()}
    off += 4 * n
    off += 4 * n  # i_filter (hardware-interpolation variant, unused here)
    t["sigma"] = np.frombuffer(b, "<f4", L, off).copy(); off += 4 * L
    t["span"] = np.frombuffer(b, "<i4", L, off).copy(); off += 4 * L
    off += 4 * L  # i_span
    return t, off


class RefTables:
    """h_gauss (reference gauss_filter.h:49-105 GaussInfo) + h_consts (sift_constants.h:56-67 ConstInfo)."""
    def __init__(self, path):
        b = open(path, "rb").read()
        self.stages = int(np.frombuffer(b, "<i4", 1, 0)[0])
        off = 4
        self.inc, off = _gauss_table(b, off, 12)
        self.abs_o0, off = _gauss_table(b, off, 12)
        self.abs_oN, off = _gauss_table(b, off, 12)
        self.dd, off = _gauss_table(b, off, 20)
        off += 4  # _span_mode
        assert off == 15016, off
        ci = np.frombuffer(b, "<i4", 2, off); cf = np.frombuffer(b, "<f4", 4, off + 8)
        self.max_extrema, self.max_orientations = int(ci[0]), int(ci[1])
        self.sigma0, self.sigma_k, self.edge_limit, self.threshold = [float(v) for v in cf]
        self.norm_multi = int(np.frombuffer(b, "<i4", 1, off + 24)[0])
        self.desc_gauss = np.frombuffer(b, "<f4", 1600, off + 28).reshape(40, 40).copy()
        self.desc_tile = np.frombuffer(b, "<f4", 16, off + 28 + 6400).copy()
        assert off + 28 + 6400 + 64 == len(b), len(b)


def canonical_order(ext):
    """Sort keypoints by (octave, lpos, round(y), round(x), y, x): the reference's output order is
    nondeterministic (atomics), its own regression script sorts too (testOxfordDataset.sh.in:138-155)."""
    return np.lexsort((ext["x"], ext["y"], np.rint(ext["x"]), np.rint(ext["y"]), ext["lpos"], ext["octave"]))
