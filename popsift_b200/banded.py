"""One giant image in row bands.

The CUDA side is psift_submit_band_u8 / psift_band_seed_copy / psift_submit_seed_dev_f32 (include/psift.h):
every context builds octaves 0..split for its band (+halo) in GLOBAL coordinates, the rows of level 0 of
octave split+1 that the bands produce are gathered into one plane -- the one exchange step, an NCCL
all-gather over NVLink when the bands live on different GPUs -- and every context finishes the small octaves
from that plane, keeping the keypoints whose detection row is in its band.  The union over the bands equals
the un-split result bit for bit (tests/test_parity_gpu.py::test_banded_equals_whole).

    extract_banded_local(ctx, img, parts)      all bands one after the other on ONE context (test / fallback)
    extract_banded_dist(ctx, img_dev, w, h)    one band per rank of torch.distributed (rank r owns band r)
"""
import numpy as np
import torch

from . import capi


class Merged:
    """Concatenation of several Extractions (descriptor indices re-based)."""
    def __init__(self, parts):
        feats, exts, descs = [], [], []
        nd = 0
        for p in parts:
            f, e = p.feat.copy(), p.ext.copy()
            if len(f):
                di = f["desc_idx"]; di[di >= 0] += nd; f["desc_idx"] = di
                e["idx_ori"] += nd
            feats.append(f); exts.append(e); descs.append(p.desc)
            nd += p.n_desc
        self.feat = np.concatenate(feats) if feats else np.zeros(0, capi.FEATURE_DTYPE)
        self.ext = np.concatenate(exts) if exts else np.zeros(0, capi.EXTREMUM_DTYPE)
        self.desc = np.concatenate(descs) if descs else np.zeros((0, 128), np.float32)
        self.n_feat, self.n_desc = len(self.feat), len(self.desc)
        self.ext_ct = sum((np.asarray(p.ext_ct) for p in parts), np.zeros_like(np.asarray(parts[0].ext_ct)))
        self.ori_ct = sum((np.asarray(p.ori_ct) for p in parts), np.zeros_like(np.asarray(parts[0].ori_ct)))
        self.n_octaves, self.n_levels, self.w0, self.h0 = parts[0].n_octaves, parts[0].n_levels, parts[0].w0, parts[0].h0


def digest(ex):
    """Order-independent checksum of an Extraction / Merged: (sum of CRC32 over the feature records without their
    descriptor indices, sum of CRC32 over the descriptor rows).  Sums over disjoint parts add up to the whole's."""
    import zlib
    f = ex.feat
    rec = np.zeros(len(f), dtype=[("o", "<i4"), ("x", "<f4"), ("y", "<f4"), ("s", "<f4"), ("n", "<i4"), ("a", "<f4", (4,))])
    rec["o"], rec["x"], rec["y"], rec["s"], rec["n"], rec["a"] = f["octave"], f["x"], f["y"], f["sigma"], f["num_ori"], f["ori"]
    fb = rec.tobytes(); n = rec.dtype.itemsize
    fs = sum(zlib.crc32(fb[i * n:(i + 1) * n]) for i in range(len(rec)))
    db = np.ascontiguousarray(ex.desc, np.float32).tobytes()
    ds = sum(zlib.crc32(db[i * 512:(i + 1) * 512]) for i in range(len(ex.desc)))
    return fs, ds


def _own_seed_rows(bounds, r, split, seed_h):
    """rows of octave split+1 that band r contributes to the gathered plane: a partition of [0, seed_h)"""
    lo = bounds[r] >> (split + 1)
    hi = seed_h if r + 1 == len(bounds) - 1 else min(seed_h, bounds[r + 1] >> (split + 1))
    return lo, hi


def extract_banded_local(ctx, img, parts, cfg=None, max_seed_pixels=0):
    """All bands on one context, sequentially.  img: host uint8 array."""
    cfg = cfg if cfg is not None else capi.default_config()
    h, w = img.shape
    split, bounds = capi.band_plan(cfg, w, h, parts, max_seed_pixels)
    out, plane, n_oct = [], None, capi.geometry(cfg, w, h)[2]
    for r in range(parts):
        j = ctx.submit_band(img, bounds[r], bounds[r + 1], split)
        out.append(ctx.wait(j))
        if split + 1 < n_oct:
            _, _, lo, hi, sw, sh = ctx.band_seed(j)
            if plane is None:
                plane = torch.empty((sh, sw), dtype=torch.float32, device="cuda")
            a, b = _own_seed_rows(bounds, r, split, sh)
            assert lo <= a and b <= hi, (lo, a, b, hi)
            ctx.band_seed_copy(j, plane[a].data_ptr(), sw * 4, a, b)
        ctx.free(j)
    if plane is not None:
        for r in range(parts):
            j = ctx.submit_seed(w, h, split + 1, plane.data_ptr(), plane.shape[1] * 4, bounds[r], bounds[r + 1])
            out.append(ctx.wait(j)); ctx.free(j)
    return Merged(out), split, bounds


def extract_banded_dist(ctx, img, cfg=None, max_seed_pixels=0):
    """One band per rank (rank r owns band r).  img: host uint8 array, the same on every rank.
    Returns this rank's keypoints (band pass + its share of the upper octaves) as a Merged."""
    import torch.distributed as dist
    cfg = cfg if cfg is not None else capi.default_config()
    rank, world = dist.get_rank(), dist.get_world_size()
    h, w = img.shape
    split, bounds = capi.band_plan(cfg, w, h, world, max_seed_pixels)
    n_oct = capi.geometry(cfg, w, h)[2]
    j = ctx.submit_band(img, bounds[rank], bounds[rank + 1], split)
    out = [ctx.wait(j)]
    if split + 1 < n_oct:
        _, _, lo, hi, sw, sh = ctx.band_seed(j)
        rows = [_own_seed_rows(bounds, r, split, sh) for r in range(world)]
        most = max(b - a for a, b in rows)
        a, b = rows[rank]
        mine = torch.zeros((most, sw), dtype=torch.float32, device="cuda")
        ctx.band_seed_copy(j, mine.data_ptr(), sw * 4, a, b)
        ctx.free(j)
        allrows = torch.empty((world * most, sw), dtype=torch.float32, device="cuda")
        dist.all_gather_into_tensor(allrows, mine)                  # the one exchange step
        plane = torch.empty((sh, sw), dtype=torch.float32, device="cuda")
        for r, (ra, rb) in enumerate(rows):
            plane[ra:rb] = allrows[r * most: r * most + (rb - ra)]
        torch.cuda.synchronize()
        j2 = ctx.submit_seed(w, h, split + 1, plane.data_ptr(), sw * 4, bounds[rank], bounds[rank + 1])
        out.append(ctx.wait(j2)); ctx.free(j2)
    else:
        ctx.free(j)
    return Merged(out), split, bounds
