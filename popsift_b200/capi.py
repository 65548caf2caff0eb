"""ctypes binding of the C ABI in include/psift.h (libpopsift_b200.so).

This is plumbing for tests and bench.py; the product is the shared library.  Nothing here
computes: if the library (or a CUDA device) is missing the calls raise.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpopsift_b200.so")

MAX_OCT = 20
MAX_ORI = 4


class Config(C.Structure):
    _fields_ = [("octaves", C.c_int), ("levels", C.c_int), ("sigma", C.c_float), ("edge_limit", C.c_float),
                ("threshold", C.c_float), ("upscale_factor", C.c_float), ("gauss_mode", C.c_int),
                ("sift_mode", C.c_int), ("desc_mode", C.c_int), ("norm_mode", C.c_int), ("norm_multi", C.c_int),
                ("max_extrema", C.c_int), ("filter_max_extrema", C.c_int), ("filter_grid_size", C.c_int),
                ("filter_sorting", C.c_int), ("assume_initial_blur", C.c_int), ("initial_blur", C.c_float),
                ("deterministic_order", C.c_int), ("scaling_mode", C.c_int)]


FEATURE_DTYPE = np.dtype([("octave", "<i4"), ("x", "<f4"), ("y", "<f4"), ("sigma", "<f4"), ("num_ori", "<i4"),
                          ("ori", "<f4", (4,)), ("desc_idx", "<i4", (4,))])
EXTREMUM_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("lpos", "<i4"), ("sigma", "<f4"), ("octave", "<i4"),
                           ("num_ori", "<i4"), ("idx_ori", "<i4"), ("ori", "<f4", (4,))])


class Result(C.Structure):
    _fields_ = [("n_feat", C.c_int), ("n_desc", C.c_int), ("feat", C.c_void_p), ("desc", C.c_void_p),
                ("ext", C.c_void_p), ("ext_ct", C.c_int * MAX_OCT), ("ori_ct", C.c_int * MAX_OCT),
                ("n_octaves", C.c_int), ("n_levels", C.c_int), ("w0", C.c_int), ("h0", C.c_int)]


GAUSS = {"vlfeat": 0, "vlfeat-hw-interpolated": 1, "relative": 1, "vlfeat-direct": 2, "opencv": 3, "fixed9": 4, "fixed15": 5}
MODE = {"popsift": 0, "opencv": 1, "vlfeat": 2}
DESC = {"loop": 0, "iloop": 1, "grid": 2, "igrid": 3, "notile": 4}
NORM = {"RootSift": 0, "classic": 1}
FSORT = {"random": 0, "down": 1, "up": 2}   # reference sift_conf.cu:131-141 (up = smallest scale first)

class BandedResult(C.Structure):
    """psift_banded_result (psift.h)"""
    _fields_ = [("n_feat", C.c_int), ("n_desc", C.c_int), ("split_octave", C.c_int),
                ("feat", C.c_void_p), ("ext", C.c_void_p), ("desc", C.c_void_p), ("halo_exchange", C.c_int), ("block", C.c_void_p)]


class DevInfo(C.Structure):
    """psift_dev_info (psift.h)"""
    _fields_ = [("device", C.c_int), ("n_feat", C.c_int), ("n_desc", C.c_int),
                ("feat", C.c_void_p), ("feat_ref", C.c_void_p), ("desc", C.c_void_p), ("rev", C.c_void_p)]


EXPORTS = ["psift_config_default", "psift_last_error", "psift_device_count", "psift_create", "psift_configure",
           "psift_destroy", "psift_submit_u8", "psift_submit_f32", "psift_submit_dev_u8", "psift_submit_pinned_u8", "psift_submit_pinned_f32", "psift_wait_take", "psift_result_block_free", "psift_host_alloc_pinned", "psift_host_free_pinned", "psift_wait",
           "psift_job_free", "psift_octave_size", "psift_debug_read_level", "psift_debug_read_dog",
           "psift_debug_gauss_table", "psift_debug_desc_tables", "psift_pyramid_geometry", "psift_last_timing",
           "psift_launch_count", "psift_slot_count", "psift_cxx_stream", "psift_d2h_bytes", "psift_debug_mark", "psift_debug_mark_elapsed", "psift_debug_time_levels", "psift_last_level_timing",
           "psift_extract_banded", "psift_banded_free", "psift_band_plan", "psift_submit_band_u8", "psift_band_seed", "psift_band_seed_copy", "psift_submit_seed_dev_f32",
           "psift_wait_dev", "psift_dev_features_alloc", "psift_dev_features_free", "psift_dev_features_info", "psift_match",
           "psift_match_dev", "psift_match_dev_noalloc", "psift_dev_download", "psift_dev_upload"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libpopsift_b200.so is not built: run `make` (or __graft_entry__.build()) first")
        L = C.CDLL(LIB_PATH)
        L.psift_last_error.restype = C.c_char_p
        L.psift_launch_count.restype = C.c_longlong
        L.psift_d2h_bytes.restype = C.c_longlong
        L.psift_d2h_bytes.argtypes = [C.c_void_p]
        L.psift_launch_count.argtypes = [C.c_void_p]
        L.psift_slot_count.argtypes = [C.c_void_p]
        L.psift_cxx_stream.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.psift_create.argtypes = [C.c_int, C.POINTER(Config), C.POINTER(C.c_void_p)]
        L.psift_configure.argtypes = [C.c_void_p, C.POINTER(Config)]
        L.psift_destroy.argtypes = [C.c_void_p]
        L.psift_submit_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.psift_submit_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.psift_submit_dev_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.psift_submit_pinned_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.psift_wait.argtypes = [C.c_void_p, C.POINTER(Result)]
        L.psift_wait_take.argtypes = [C.c_void_p, C.POINTER(Result), C.POINTER(C.c_void_p)]
        L.psift_result_block_free.argtypes = [C.c_void_p]
        L.psift_result_block_free.restype = None
        L.psift_submit_pinned_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.psift_job_free.argtypes = [C.c_void_p]
        L.psift_octave_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psift_debug_read_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.psift_debug_read_dog.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.psift_debug_gauss_table.argtypes = [C.POINTER(Config), C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.psift_debug_desc_tables.argtypes = [C.POINTER(Config), C.c_void_p, C.c_void_p, C.c_void_p]
        L.psift_pyramid_geometry.argtypes = [C.POINTER(Config), C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psift_last_timing.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        L.psift_band_plan.argtypes = [C.POINTER(Config), C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psift_submit_band_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.psift_band_seed.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psift_band_seed_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int]
        L.psift_submit_seed_dev_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.psift_debug_mark.argtypes = [C.c_void_p, C.c_int]
        L.psift_extract_banded.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.POINTER(BandedResult)]
        L.psift_banded_free.argtypes = [C.POINTER(BandedResult)]
        L.psift_banded_free.restype = None
        L.psift_debug_time_levels.argtypes = [C.c_void_p, C.c_int]
        L.psift_last_level_timing.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_int)]
        L.psift_wait_dev.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.psift_dev_features_alloc.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.psift_dev_features_free.argtypes = [C.c_void_p]
        L.psift_dev_features_info.argtypes = [C.c_void_p, C.POINTER(DevInfo)]
        L.psift_match.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.psift_match_dev.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.psift_match_dev_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.psift_dev_download.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.psift_dev_upload.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.psift_debug_mark_elapsed.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        _lib = L
    return _lib


class PsiftError(RuntimeError):
    pass


def _chk(rc):
    if rc != 0:
        raise PsiftError("psift error %d: %s" % (rc, lib().psift_last_error().decode()))


def default_config(**kw):
    """psift_config with the reference defaults; keyword arguments use the reference's spellings
    (mode='vlfeat', gauss_mode='opencv', desc_mode='igrid', norm_mode='classic', downsampling=-1 ...)."""
    cfg = Config()
    lib().psift_config_default(C.byref(cfg))
    for k, v in kw.items():
        if k == "mode": cfg.sift_mode = MODE[v]
        elif k == "gauss_mode": cfg.gauss_mode = GAUSS[v]
        elif k == "desc_mode": cfg.desc_mode = DESC[v]
        elif k == "norm_mode": cfg.norm_mode = NORM[v]
        elif k == "filter_sort": cfg.filter_sorting = FSORT[v]
        elif k == "downsampling": cfg.upscale_factor = -float(v)
        elif k == "scaling": cfg.scaling_mode = {"default": 0, "direct": 1}[v]
        elif k == "initial_blur":
            cfg.initial_blur = float(v); cfg.assume_initial_blur = 0 if v == 0 else 1
        elif hasattr(cfg, k): setattr(cfg, k, v)
        else: raise KeyError(k)
    return cfg


def config_from_flags(flags):
    """Translate ref_driver / popsift-demo style flags (tools/golden_cases.py) into a psift_config."""
    kw = {}
    it = iter(flags)
    for f in it:
        v = next(it)
        key = f.lstrip("-").replace("-", "_")
        if key in ("mode", "gauss_mode", "desc_mode", "norm_mode", "filter_sort", "scaling"): kw[key] = v
        elif key == "downsampling": kw["downsampling"] = float(v)
        elif key in ("octaves", "levels", "norm_multi", "filter_max_extrema", "max_extrema"): kw[key] = int(v)
        elif key == "filter_grid": kw["filter_grid_size"] = int(v)
        elif key in ("sigma", "threshold", "initial_blur"): kw[key] = float(v)
        elif key == "edge_threshold": kw["edge_limit"] = float(v)
        else: raise KeyError(f)
    return default_config(**kw)


class Extraction:
    """Features of one image, copied out of the job's pinned buffers."""
    def __init__(self, r):
        self.n_feat, self.n_desc = r.n_feat, r.n_desc
        self.feat = np.ctypeslib.as_array(C.cast(r.feat, C.POINTER(C.c_uint8)), (r.n_feat * FEATURE_DTYPE.itemsize,)).view(FEATURE_DTYPE).copy() if r.n_feat else np.zeros(0, FEATURE_DTYPE)
        self.ext = np.ctypeslib.as_array(C.cast(r.ext, C.POINTER(C.c_uint8)), (r.n_feat * EXTREMUM_DTYPE.itemsize,)).view(EXTREMUM_DTYPE).copy() if r.n_feat else np.zeros(0, EXTREMUM_DTYPE)
        self.desc = np.ctypeslib.as_array(C.cast(r.desc, C.POINTER(C.c_float)), (r.n_desc, 128)).copy() if r.n_desc else np.zeros((0, 128), np.float32)
        self.ext_ct = np.array(r.ext_ct[:], dtype=np.int32)
        self.ori_ct = np.array(r.ori_ct[:], dtype=np.int32)
        self.n_octaves, self.n_levels, self.w0, self.h0 = r.n_octaves, r.n_levels, r.w0, r.h0


class Context:
    """One extraction context on one CUDA device (psift_create / psift_destroy)."""
    def __init__(self, cfg=None, device=0):
        self._h = C.c_void_p()
        cfg = cfg if cfg is not None else default_config()
        _chk(lib().psift_create(device, C.byref(cfg), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().psift_destroy(self._h); self._h = C.c_void_p()

    def __enter__(self): return self
    def __exit__(self, *a): self.close()

    def configure(self, cfg):
        _chk(lib().psift_configure(self._h, C.byref(cfg)))

    def submit(self, img):
        img = np.ascontiguousarray(img)
        h, w = img.shape
        job = C.c_void_p()
        if img.dtype == np.uint8:
            _chk(lib().psift_submit_u8(self._h, w, h, img.ctypes.data, w, C.byref(job)))
        elif img.dtype == np.float32:
            _chk(lib().psift_submit_f32(self._h, w, h, img.ctypes.data, w * 4, C.byref(job)))
        else:
            raise TypeError("image must be uint8 or float32")
        return job

    def submit_pinned_u8(self, ptr, w, h, pitch):
        """image in page-locked host memory that stays untouched until wait(): uploaded without a private copy"""
        job = C.c_void_p()
        _chk(lib().psift_submit_pinned_u8(self._h, w, h, C.c_void_p(ptr), pitch, C.byref(job)))
        return job

    def submit_dev_u8(self, ptr, w, h, pitch):
        job = C.c_void_p()
        _chk(lib().psift_submit_dev_u8(self._h, w, h, C.c_void_p(ptr), pitch, C.byref(job)))
        return job

    # ---- one giant image in row bands (psift.h, "one giant image over several GPUs") ----
    def submit_band(self, img, y0, y1, split, w=None, h=None, pitch=None):
        """img: host uint8 array (whole image) or a device pointer (then give w, h, pitch)"""
        job = C.c_void_p()
        if isinstance(img, np.ndarray):
            img = np.ascontiguousarray(img); hh, ww = img.shape
            _chk(lib().psift_submit_band_u8(self._h, ww, hh, img.ctypes.data, ww, 0, y0, y1, split, C.byref(job)))
        else:
            _chk(lib().psift_submit_band_u8(self._h, w, h, C.c_void_p(img), pitch, 1, y0, y1, split, C.byref(job)))
        return job

    def band_seed(self, job):
        p = C.c_void_p(); pitch, lo, hi, sw, sh = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _chk(lib().psift_band_seed(job, C.byref(p), C.byref(pitch), C.byref(lo), C.byref(hi), C.byref(sw), C.byref(sh)))
        return p.value, pitch.value, lo.value, hi.value, sw.value, sh.value

    def band_seed_copy(self, job, dst_ptr, dst_pitch_bytes, row_from, row_to):
        _chk(lib().psift_band_seed_copy(job, C.c_void_p(dst_ptr), dst_pitch_bytes, row_from, row_to))

    def submit_seed(self, w, h, first_octave, plane_ptr, pitch_bytes, y0, y1):
        job = C.c_void_p()
        _chk(lib().psift_submit_seed_dev_f32(self._h, w, h, first_octave, C.c_void_p(plane_ptr), pitch_bytes, y0, y1, C.byref(job)))
        return job

    def wait(self, job, copy=True):
        r = Result()
        _chk(lib().psift_wait(job, C.byref(r)))
        return Extraction(r) if copy else (r.n_feat, r.n_desc)

    def free(self, job):
        lib().psift_job_free(job)

    def wait_dev(self, job):
        """device-resident result of a job (reference MatchingMode): a DevFeatures"""
        h = C.c_void_p()
        _chk(lib().psift_wait_dev(job, C.byref(h)))
        return DevFeatures(h)

    def extract_dev(self, img):
        job = self.submit(img)
        try:
            return self.wait_dev(job)
        finally:
            self.free(job)

    def extract(self, img):
        job = self.submit(img)
        try:
            return self.wait(job)
        finally:
            self.free(job)

    def level(self, octave, level):
        w, h = C.c_int(), C.c_int()
        _chk(lib().psift_octave_size(self._h, octave, C.byref(w), C.byref(h)))
        out = np.empty((h.value, w.value), np.float32)
        _chk(lib().psift_debug_read_level(self._h, octave, level, out.ctypes.data))
        return out

    def dog(self, octave, level):
        w, h = C.c_int(), C.c_int()
        _chk(lib().psift_octave_size(self._h, octave, C.byref(w), C.byref(h)))
        out = np.empty((h.value, w.value), np.float32)
        _chk(lib().psift_debug_read_dog(self._h, octave, level, out.ctypes.data))
        return out

    def timing(self):
        ms = (C.c_float * 8)()
        _chk(lib().psift_last_timing(self._h, ms))
        return dict(zip(["upload", "pyramid", "extrema", "orientation", "descriptor", "download", "total"], [float(x) for x in ms[:7]]))

    def time_levels(self, on):
        """bracket every octave-0 blur launch with CUDA events on its own stream (psift_debug_time_levels)"""
        _chk(lib().psift_debug_time_levels(self._h, 1 if on else 0))

    def level_timing(self):
        """durations (ms) of the octave-0 Gaussian-level launches of the last finished job, level 0 first"""
        ms = (C.c_float * 16)(); n = C.c_int()
        _chk(lib().psift_last_level_timing(self._h, ms, 16, C.byref(n)))
        return [float(ms[i]) for i in range(n.value)]

    def mark(self, which):
        """device-side stopwatch (psift_debug_mark): 0 = start, 1 = stop"""
        _chk(lib().psift_debug_mark(self._h, which))

    def mark_elapsed_ms(self):
        ms = C.c_float()
        _chk(lib().psift_debug_mark_elapsed(self._h, C.byref(ms)))
        return float(ms.value)

    def launch_count(self):
        return int(lib().psift_launch_count(self._h))

    def d2h_bytes(self):
        return int(lib().psift_d2h_bytes(self._h))


def cxx_stream(imgs, warm_frames, timed_frames, first_device=0, n_devices=1, n_producers=1, window=8):
    """psift_cxx_stream: the frames through PopSift::enqueue()/SiftJob::get() (the C++ classes); returns a dict"""
    imgs = [np.ascontiguousarray(im, np.uint8) for im in imgs]
    h, w = imgs[0].shape
    ptrs = (C.c_void_p * len(imgs))(*[im.ctypes.data for im in imgs])
    sec, nf, nd, cs, dev = C.c_double(), C.c_longlong(), C.c_longlong(), C.c_double(), C.c_int()
    _chk(lib().psift_cxx_stream(first_device, n_devices, n_producers, w, h, ptrs, len(imgs), warm_frames, timed_frames, window,
                                C.byref(sec), C.byref(nf), C.byref(nd), C.byref(cs), C.byref(dev)))
    return {"seconds": sec.value, "n_feat": nf.value, "n_desc": nd.value, "checksum": cs.value, "devices": dev.value,
            "images_per_sec": timed_frames / sec.value if sec.value > 0 else 0.0}


def gauss_table(cfg, which, level):
    f = np.zeros(32, np.float32); s = C.c_float(); sp = C.c_int()
    _chk(lib().psift_debug_gauss_table(C.byref(cfg), which, level, f.ctypes.data, C.byref(s), C.byref(sp)))
    return f, s.value, sp.value


def desc_tables(cfg):
    g = np.zeros((40, 40), np.float32); t = np.zeros(16, np.float32); sc = np.zeros(2, np.float32)
    _chk(lib().psift_debug_desc_tables(C.byref(cfg), g.ctypes.data, t.ctypes.data, sc.ctypes.data))
    return g, t, sc


def band_plan(cfg, w, h, parts, max_seed_pixels=0):
    """(split octave, [parts+1 band boundaries in rows of octave 0])"""
    split = C.c_int(); b = (C.c_int * (parts + 1))()
    _chk(lib().psift_band_plan(C.byref(cfg), w, h, parts, max_seed_pixels, C.byref(split), b))
    return split.value, [int(v) for v in b]


class BandedExtraction:
    """Result of extract_banded: same fields as Extraction / banded.Merged (feat, ext, desc, n_feat, n_desc)."""
    def __init__(self, r):
        self.n_feat, self.n_desc, self.split, self.halo_exchange = r.n_feat, r.n_desc, r.split_octave, r.halo_exchange
        def arr(ptr, n, dt):
            if n == 0: return np.zeros(0, dt)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), (n * dt.itemsize,)).view(dt).copy()
        self.feat = arr(r.feat, r.n_feat, FEATURE_DTYPE)
        self.ext = arr(r.ext, r.n_feat, EXTREMUM_DTYPE)
        self.desc = arr(r.desc, r.n_desc * 128, np.dtype("<f4")).reshape(-1, 128) if r.n_desc else np.zeros((0, 128), np.float32)


def extract_banded(contexts, img, max_seed_pixels=0, copy=True):
    """One giant image over several contexts (psift_extract_banded): band pass with halo exchange, seed gather, upper
    octaves, merge.  copy=False: only (n_feat, n_desc, split, halo_exchange), for timing."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    hs = (C.c_void_p * len(contexts))(*[c._h for c in contexts])
    r = BandedResult()
    _chk(lib().psift_extract_banded(hs, len(contexts), w, h, img.ctypes.data, w, max_seed_pixels, C.byref(r)))
    try:
        return BandedExtraction(r) if copy else (r.n_feat, r.n_desc, r.split_octave, r.halo_exchange)
    finally:
        lib().psift_banded_free(C.byref(r))


def geometry(cfg, w, h):
    a, b, n = C.c_int(), C.c_int(), C.c_int()
    _chk(lib().psift_pyramid_geometry(C.byref(cfg), w, h, C.byref(a), C.byref(b), C.byref(n)))
    return a.value, b.value, n.value


class DevFeatures:
    """psift_dev_features: features, descriptors and the descriptor->feature map of one image in device memory
    (reference FeaturesDev, features.h:100-118)."""
    def __init__(self, handle):
        self._h = handle
        self.info = DevInfo()
        _chk(lib().psift_dev_features_info(self._h, C.byref(self.info)))
        self.n_feat, self.n_desc, self.device = self.info.n_feat, self.info.n_desc, self.info.device

    @classmethod
    def from_host(cls, desc, device=0):
        """upload an (n,128) float32 array as a descriptor set (no features)"""
        desc = np.ascontiguousarray(desc, np.float32).reshape(-1, 128)
        h = C.c_void_p()
        _chk(lib().psift_dev_features_alloc(device, 0, len(desc), C.byref(h)))
        f = cls(h)
        _chk(lib().psift_dev_upload(device, f.info.desc, desc.ctypes.data, desc.nbytes))
        return f

    def _download(self, ptr, shape, dtype):
        out = np.empty(shape, dtype)
        _chk(lib().psift_dev_download(self.device, ptr, out.ctypes.data, out.nbytes))
        return out

    def descriptors(self): return self._download(self.info.desc, (self.n_desc, 128), np.float32)
    def features(self):    return self._download(self.info.feat, (self.n_feat,), FEATURE_DTYPE)
    def reverse_map(self): return self._download(self.info.rev, (self.n_desc,), np.int32)
    def features_ref_layout(self): return self._download(self.info.feat_ref, (self.n_feat, 72), np.uint8)

    def match(self, other):
        """FeaturesDev::match (features.cu:268-302): (match[n,3] int32 = best, second, accept; dist[n,2] = squared L2)"""
        m = np.zeros((self.n_desc, 3), np.int32); d = np.zeros((self.n_desc, 2), np.float32)
        _chk(lib().psift_match(self._h, other._h, m.ctypes.data, d.ctypes.data))
        return m, d

    def close(self):
        if self._h: lib().psift_dev_features_free(self._h); self._h = None
    def __enter__(self): return self
    def __exit__(self, *a): self.close()
    def __del__(self):
        try: self.close()
        except Exception: pass
