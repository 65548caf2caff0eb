"""Golden-fixture case list shared by tools/run_ref_golden.py (runs the reference CUDA
build, oracle/_ref/ref_driver, on a GPU box) and tests/ (which compare against the
committed outputs).  TEST INFRASTRUCTURE."""
import sys, os as _os
sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import os
import numpy as np
from popsift_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

# name -> (w, h, generator)
IMAGES = {
    "g64":    (64, 48,   lambda: synth.gen(64, 48, 11)),
    "g33x65": (33, 65,   lambda: synth.gen(33, 65, 13)),
    "g17":    (17, 17,   lambda: synth.gen(17, 17, 14)),
    "g256":   (256, 192, lambda: synth.gen(256, 192, 12)),
    "blob":   (96, 80,   lambda: synth.single_blob(96, 80, 41.3, 37.6, 4.0)),
    "const":  (64, 64,   lambda: synth.constant(64, 64)),
    "check":  (128, 96,  lambda: synth.checkerboard(128, 96)),
    "g640":   (640, 480, lambda: synth.gen(640, 480, 1)),
}

# case name -> (image, ref_driver flags, dump planes?)
CASES = {
    "g64_default":      ("g64",    [], True),
    "g33x65_default":   ("g33x65", [], True),
    "g17_default":      ("g17",    [], True),
    "g64_down0":        ("g64",    ["--downsampling", "0"], True),
    "g64_opencvgauss":  ("g64",    ["--gauss-mode", "opencv", "--mode", "opencv"], True),
    "blob_default":     ("blob",   [], False),
    "const_default":    ("const",  [], False),
    "check_default":    ("check",  [], False),
    "g256_default":     ("g256",   [], False),
    "g256_default_b":   ("g256",   [], False),   # second run: the reference's own run-to-run spread
    "g256_default_c":   ("g256",   [], False),
    "g256_l2":          ("g256",   ["--norm-mode", "classic"], False),
    "g256_multi9":      ("g256",   ["--norm-multi", "9"], False),
    "g256_vlfeat_igrid": ("g256",  ["--mode", "vlfeat", "--gauss-mode", "vlfeat", "--downsampling", "-1", "--desc-mode", "igrid"], False),
    "g256_notile":      ("g256",   ["--desc-mode", "notile"], False),
    "g256_grid":        ("g256",   ["--desc-mode", "grid"], False),
    "g256_iloop":       ("g256",   ["--desc-mode", "iloop"], False),
    "g256_opencv":      ("g256",   ["--mode", "opencv", "--gauss-mode", "opencv", "--norm-mode", "RootSift"], False),
    "g256_down0":       ("g256",   ["--downsampling", "0"], False),
    "g256_down1":       ("g256",   ["--downsampling", "1"], False),
    "g256_levels4":     ("g256",   ["--levels", "4"], False),
    "g256_filter":      ("g256",   ["--filter-max-extrema", "40", "--filter-grid", "4", "--filter-sort", "down"], False),
    "g256_filter_up":   ("g256",   ["--filter-max-extrema", "40", "--filter-grid", "2", "--filter-sort", "up"], False),
    "g640_filter":      ("g640",   ["--filter-max-extrema", "200", "--filter-grid", "4", "--filter-sort", "down"], False),
    "g640_filter_up":   ("g640",   ["--filter-max-extrema", "120", "--filter-grid", "3", "--filter-sort", "up"], False),
    "g640_filter_rand": ("g640",   ["--filter-max-extrema", "150", "--filter-grid", "3", "--filter-sort", "random"], False),
    "g640_default":     ("g640",   [], False),
    "g64_direct":       ("g64",    ["--gauss-mode", "vlfeat-direct"], True),
    "g256_direct":      ("g256",   ["--gauss-mode", "vlfeat-direct"], False),
    "g64_relative":     ("g64",    ["--gauss-mode", "relative"], True),
    "g64_scaledirect":  ("g64",    ["--scaling", "direct"], True),
    "g256_scaledirect": ("g256",   ["--scaling", "direct"], False),
    "g64_scaledirect_rel": ("g64", ["--scaling", "direct", "--gauss-mode", "relative"], True),
    "g256_relative":    ("g256",   ["--gauss-mode", "relative"], False),
    "g64_scaledirect_fixed9": ("g64", ["--scaling", "direct", "--gauss-mode", "fixed9"], True),
    "g64_fixed9":       ("g64",    ["--gauss-mode", "fixed9"], True),
    "g64_fixed15":      ("g64",    ["--gauss-mode", "fixed15"], True),
    # float input images (reference ImageFloat, s_image.cu:175-293; PopSift::enqueue(int,int,const float*)): the
    # same pictures as float32 u8/255, see float_image()
    "g64_float":        ("g64",    ["--float-mode"], True),
    "g256_float":       ("g256",   ["--float-mode"], False),
    "g640_float":       ("g640",   ["--float-mode", "--downsampling", "0"], False),
}
VALUELESS = ("--float-mode",)


def config_flags(flags):
    """the flags that describe the Config (drops the ones that describe the input)"""
    return [f for f in flags if f not in VALUELESS]


def is_float(flags):
    return "--float-mode" in flags


def float_image(name):
    """the float32 version of a golden input: u8 / 255 in float32, what both the reference run and the tests feed"""
    w, h, _ = IMAGES[name]
    return np.fromfile(image_path(name), np.uint8).reshape(h, w).astype(np.float32) / np.float32(255.0)


# Matcher goldens (reference compute_distance, features.cu:184-225, run by `ref_driver --match-desc`): the left and
# right descriptor sets are those of two golden extraction cases above; output tests/golden/ref/<name>.match.bin.
MATCH_CASES = {
    "m_g640_vs_filter":   ("g640_default", "g640_filter"),      # right is a subset of left: exact zeros
    "m_g256_vs_opencv":   ("g256_default", "g256_opencv"),      # same image, other mode: near matches
    "m_g256_vs_g640":     ("g256_default", "g640_default"),     # unrelated images
    "m_g640_vs_g256":     ("g640_default", "g256_default"),
    "m_g256_self":        ("g256_default", "g256_default"),
    "m_check_self":       ("check_default", "check_default"),   # checkerboard: repeated descriptors, ties
    "m_multi9_vs_l2":     ("g256_multi9", "g256_l2"),           # scaled descriptors
    "m_tiny":             ("g17_default", "blob_default"),      # 2 x 3
    "m_one_right":        ("g64_default", "g17_default"),       # r_len 2
    "m_empty_right":      ("g64_default", "const_default"),     # r_len 0
    "m_empty_left":       ("const_default", "g64_default"),     # l_len 0
}


def image_path(name):
    return os.path.join(GOLD, "inputs", name + ".u8")


def write_inputs():
    os.makedirs(os.path.join(GOLD, "inputs"), exist_ok=True)
    for name, (w, h, fn) in IMAGES.items():
        img = fn()
        assert img.shape == (h, w) and img.dtype == np.uint8
        img.tofile(image_path(name))


if __name__ == "__main__":
    write_inputs()
    print("wrote", len(IMAGES), "inputs to", os.path.join(GOLD, "inputs"))
