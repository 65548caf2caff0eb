"""Comparison of an extraction (popsift_b200.capi.Extraction) with the reference's output for the
same image and Config (tests/refio.RefFeatures).  TEST INFRASTRUCTURE.

Rules (BASELINE.json north_star):
  * integer outputs bit-exact: per-octave keypoint counts, totals, (octave, lpos) and num_ori of
    every keypoint after canonical ordering;
  * floats within tolerance: |dx|,|dy| < 1e-3 px (octave coordinates), |dsigma|/sigma < 1e-5,
    |dtheta| < 1e-3 rad (mod 2pi), descriptor L2 distance < 1e-3 on unit-scaled descriptors.
"""
import numpy as np
import refio

TOL_XY = 1e-3
TOL_SIGMA_REL = 1e-5
TOL_THETA = 1e-3
TOL_DESC = 1e-3


def ang_diff(a, b):
    d = np.abs(a - b) % (2 * np.pi)
    return np.minimum(d, 2 * np.pi - d)


def compare(mine, ref, norm_multi=0, skip_last_desc=True):
    """Returns a dict of findings; 'ok_int' / 'ok_float' summarise the two gates."""
    rep = {}
    rep["n_feat"] = (mine.n_feat, ref.n_feat)
    rep["n_desc"] = (mine.n_desc, ref.n_desc)
    rep["ext_ct_equal"] = bool(np.array_equal(mine.ext_ct, ref.ext_ct))
    rep["ori_ct_equal"] = bool(np.array_equal(mine.ori_ct, ref.ori_ct))
    ok_int = mine.n_feat == ref.n_feat and rep["ext_ct_equal"]
    rep["ok_float"] = False
    if not ok_int:
        rep["ok_int"] = False
        return rep
    om, orf = refio.canonical_order(mine.ext), refio.canonical_order(ref.ext)
    me, re_ = mine.ext[om], ref.ext[orf]
    mf, rf = mine.feat[om], ref.feat[orf]
    rep["octave_equal"] = bool(np.array_equal(me["octave"], re_["octave"]))
    rep["lpos_equal"] = bool(np.array_equal(me["lpos"], re_["lpos"]))
    rep["num_ori_equal"] = bool(np.array_equal(me["num_ori"], re_["num_ori"]))
    rep["num_ori_mismatches"] = int((me["num_ori"] != re_["num_ori"]).sum())
    rep["xy_bitexact"] = bool(np.array_equal(me["x"].view(np.uint32), re_["x"].view(np.uint32)) and
                              np.array_equal(me["y"].view(np.uint32), re_["y"].view(np.uint32)))
    rep["sigma_bitexact"] = bool(np.array_equal(me["sigma"].view(np.uint32), re_["sigma"].view(np.uint32)))
    rep["max_dxy"] = float(max(np.abs(me["x"] - re_["x"]).max(initial=0), np.abs(me["y"] - re_["y"]).max(initial=0)))
    rep["max_dsigma_rel"] = float((np.abs(me["sigma"] - re_["sigma"]) / np.maximum(re_["sigma"], 1e-30)).max(initial=0))
    rep["feat_xy_max"] = float(max(np.abs(mf["x"] - rf["x"]).max(initial=0), np.abs(mf["y"] - rf["y"]).max(initial=0)))
    ok_int = ok_int and rep["octave_equal"] and rep["lpos_equal"] and rep["num_ori_equal"] and mine.n_desc == ref.n_desc
    rep["ok_int"] = bool(ok_int)
    if not rep["num_ori_equal"]:
        return rep
    # orientations are stored in descending peak order on both sides (vectorised: config 5 has 250 k keypoints)
    n = len(me)
    kk = np.arange(4)[None, :]
    valid = kk < me["num_ori"][:, None]                         # (n, 4)
    if n:
        dth_all = ang_diff(me["ori"].astype(np.float32), re_["ori"].astype(np.float32))
        ia = mf["desc_idx"][valid]; jr = rf["desc_idx"][valid]
        same_ori = (me["ori"].view(np.uint32) == re_["ori"].view(np.uint32))[valid]
        dth = dth_all[valid]
        rep["ori_bitexact"] = (int(same_ori.sum()), int(valid.sum()))
    else:
        ia = jr = np.zeros(0, np.int64); same_ori = np.zeros(0, bool); dth = np.zeros(0, np.float32)
        rep["ori_bitexact"] = (0, 0)
    keep = np.ones(len(ia), bool)
    if skip_last_desc and ref.n_desc % 32 != 0:
        keep = jr != ref.n_desc - 1        # reference race on its last descriptor (s_desc_normalize.h:21-32)
    ia, jr, same_ori = ia[keep], jr[keep], same_ori[keep]
    s = 2.0 ** (-norm_multi)
    dd = np.zeros(0); nbit = 0; same_ori_diff_desc = 0
    if len(ia):
        dmax, eqs = [], []
        for c0 in range(0, len(ia), 65536):                     # chunks keep the float64 temporaries small
            a = mine.desc[ia[c0:c0 + 65536]]; b = ref.desc[jr[c0:c0 + 65536]]
            dmax.append(np.linalg.norm((a.astype(np.float64) - b.astype(np.float64)) * s, axis=1))
            eqs.append((a.view(np.uint32) == b.view(np.uint32)).all(axis=1))
        dd = np.concatenate(dmax); eq = np.concatenate(eqs)
        nbit = int(eq.sum())
        same_ori_diff_desc = int((~eq & same_ori).sum())
    dth = list(dth) if len(dth) else []
    dd = list(dd)
    rep["max_dtheta"] = float(max(dth)) if dth else 0.0
    rep["max_desc_l2"] = float(max(dd)) if dd else 0.0
    rep["desc_bitexact"] = (nbit, len(dd))
    rep["same_ori_diff_desc"] = same_ori_diff_desc
    rep["ok_float"] = bool(rep["max_dxy"] < TOL_XY and rep["max_dsigma_rel"] < TOL_SIGMA_REL and
                           rep["max_dtheta"] < TOL_THETA and rep["max_desc_l2"] < TOL_DESC)
    return rep


def cell_histogram(ext, w0, h0, grid):
    """Keypoints per grid-filter cell (reference s_extrema.cu: cell = floor(y/(h/g))*g + floor(x/(w/g)),
    octave coordinates; octave o is ceil-halved o times from w0 x h0)."""
    hist = np.zeros(grid * grid, np.int64)
    for e in ext:
        w, h = w0, h0
        for _ in range(int(e["octave"])):
            w, h = (w + 1) // 2, (h + 1) // 2
        wd, hd = np.float32(w) / np.float32(grid), np.float32(h) / np.float32(grid)
        hist[int(np.floor(np.float32(e["y"]) / hd)) * grid + int(np.floor(np.float32(e["x"]) / wd))] += 1
    return hist


def filter_params(flags):
    g = int(flags[flags.index("--filter-grid") + 1]) if "--filter-grid" in flags else 2
    srt = flags[flags.index("--filter-sort") + 1] if "--filter-sort" in flags else "random"
    return g, srt


def compare_planes(ctx, planes):
    """Gaussian and DoG planes vs the reference dump: returns (n_planes, n_bitexact, max_abs_diff)."""
    n = nb = 0
    mx = 0.0
    for o in range(planes.n_octaves):
        for l in range(planes.n_levels):
            a = ctx.level(o, l); b = planes.gauss[o][l]
            n += 1; eq = np.array_equal(a.view(np.uint32), b.view(np.uint32)); nb += int(eq)
            if not eq: mx = max(mx, float(np.abs(a - b).max()))
        for l in range(planes.n_levels - 1):
            a = ctx.dog(o, l); b = planes.dog[o][l]
            n += 1; eq = np.array_equal(a.view(np.uint32), b.view(np.uint32)); nb += int(eq)
            if not eq: mx = max(mx, float(np.abs(a - b).max()))
    return n, nb, mx
