/*
 * psift.h -- C ABI of the B200-native SIFT extraction path (libpopsift_b200.so).
 *
 * Plain pointers and sizes only.  This is the boundary a foreign-language binding
 * (ctypes, cgo, JNI ...) or the C++ drop-in classes in <popsift/popsift.h> call.
 * Each entry point names the reference interface it stands in for
 * (paths relative to the reference tree, /root/reference).
 *
 * Error convention: every int-returning function returns PSIFT_OK (0) or a negative
 * PSIFT_E_* code; psift_last_error() returns a thread-local message.  Nothing in
 * this library calls exit() (the reference does: common/debug_macros.h:119-146).
 * There is no CPU fallback: without a CUDA device psift_create fails.
 */
#ifndef PSIFT_H
#define PSIFT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PSIFT_OK            0
#define PSIFT_E_INVALID    -1   /* bad argument / unsupported config */
#define PSIFT_E_CUDA       -2   /* CUDA runtime error (message has the detail) */
#define PSIFT_E_NODEVICE   -3
#define PSIFT_E_STATE      -4   /* call order violated */

#define PSIFT_MAX_OCTAVES  20   /* reference sift_conf.h:13 MAX_OCTAVES */
#define PSIFT_MAX_ORI       4   /* reference sift_constants.h:51 ORIENTATION_MAX_COUNT */

/* enums mirror popsift::Config (reference sift_conf.h:28-120) value for value */
enum { PSIFT_GAUSS_VLFEAT_COMPUTE = 0, PSIFT_GAUSS_VLFEAT_RELATIVE = 1, PSIFT_GAUSS_VLFEAT_RELATIVE_ALL = 2,
       PSIFT_GAUSS_OPENCV_COMPUTE = 3, PSIFT_GAUSS_FIXED9 = 4, PSIFT_GAUSS_FIXED15 = 5 };
enum { PSIFT_MODE_POPSIFT = 0, PSIFT_MODE_OPENCV = 1, PSIFT_MODE_VLFEAT = 2 };
enum { PSIFT_DESC_LOOP = 0, PSIFT_DESC_ILOOP = 1, PSIFT_DESC_GRID = 2, PSIFT_DESC_IGRID = 3, PSIFT_DESC_NOTILE = 4 };
enum { PSIFT_NORM_ROOTSIFT = 0, PSIFT_NORM_CLASSIC = 1 };
enum { PSIFT_FILTER_RANDOM = 0, PSIFT_FILTER_LARGEST_FIRST = 1, PSIFT_FILTER_SMALLEST_FIRST = 2 };

/* POD mirror of the state of popsift::Config (reference sift_conf.h:270-320, defaults
 * sift_conf.cu:17-40).  psift_config_default() fills the reference defaults. */
typedef struct psift_config {
    int   octaves;              /* -1: derive from the image size (popsift.cpp:107-111) */
    int   levels;               /* 3 */
    float sigma;                /* 1.6 */
    float edge_limit;           /* 10 */
    float threshold;            /* 0.04 */
    float upscale_factor;       /* 1: input is upscaled by 2^upscale_factor (= -downsampling) */
    int   gauss_mode;           /* PSIFT_GAUSS_* */
    int   sift_mode;            /* PSIFT_MODE_* */
    int   desc_mode;            /* PSIFT_DESC_* */
    int   norm_mode;            /* PSIFT_NORM_* */
    int   norm_multi;           /* descriptors are scaled by 2^norm_multi */
    int   max_extrema;          /* 100000 per octave */
    int   filter_max_extrema;   /* -1: no grid filter */
    int   filter_grid_size;     /* 2 */
    int   filter_sorting;       /* PSIFT_FILTER_* */
    int   assume_initial_blur;  /* 1 */
    float initial_blur;         /* 0.5 */
    int   deterministic_order;  /* 1 (default): features come back sorted (octave, level, y, x) -- an extension;
                                   the reference's order is whatever its atomics produced */
    int   scaling_mode;         /* 0: every octave is seeded by the previous one (Config::ScaleDefault);
                                   1: level 0 of EVERY octave is filtered straight from the input image
                                   (Config::ScaleDirect, s_pyramid_build.cu:499-516) */
} psift_config;

/* One returned keypoint: field for field popsift::Feature (reference features.h:22-35), with
 * descriptor INDICES instead of pointers so that the record is position independent. */
typedef struct psift_feature {
    int   debug_octave;
    float xpos, ypos, sigma;             /* input-image coordinates */
    int   num_ori;
    float orientation[PSIFT_MAX_ORI];
    int   desc_idx[PSIFT_MAX_ORI];       /* index into psift_result.desc (units of 128 floats); -1 if unused */
} psift_feature;

/* Internal keypoint record in octave coordinates (reference sift_extremum.h:37-49 Extremum);
 * exposed for parity checks of the integer fields (octave, lpos). */
typedef struct psift_extremum {
    float xpos, ypos;
    int   lpos;
    float sigma;
    int   octave, num_ori, idx_ori;
    float orientation[PSIFT_MAX_ORI];
} psift_extremum;

typedef struct psift_result {
    int                  n_feat;
    int                  n_desc;
    const psift_feature* feat;        /* n_feat records, pinned host memory owned by the job */
    const float*         desc;        /* n_desc x 128 floats */
    const psift_extremum* ext;        /* n_feat records (octave coordinates) */
    int                  ext_ct[PSIFT_MAX_OCTAVES];   /* keypoints per octave */
    int                  ori_ct[PSIFT_MAX_OCTAVES];   /* descriptors per octave */
    int                  n_octaves;
    int                  n_levels;                     /* Gaussian levels per octave = levels+3 */
    int                  w0, h0;                       /* octave-0 size */
} psift_result;

typedef struct psift_ctx psift_ctx;
typedef struct psift_job psift_job;

void        psift_config_default( psift_config* cfg );              /* popsift::Config::Config(), sift_conf.cu:17-40 */
const char* psift_last_error( void );
int         psift_device_count( void );                              /* common/device_prop.cu:28-50 */

/* PopSift::PopSift + configure (popsift.cpp:16-87): bind to a CUDA device, build the Gauss and
 * descriptor tables (gauss_filter.cu:127-257, sift_constants.cu:22-53). */
int  psift_create( int device, const psift_config* cfg, psift_ctx** out );
int  psift_configure( psift_ctx* ctx, const psift_config* cfg );    /* PopSift::configure, popsift.cpp:63-87 */
void psift_destroy( psift_ctx* ctx );                               /* PopSift::uninit, popsift.cpp:122-137 */

/* PopSift::enqueue (popsift.cpp:139-169) + the worker's step1/step2/get_descriptors
 * (popsift.cpp:182-213).  The image is copied before the call returns (popsift.cpp:245-247).
 * The work is queued on the context's streams; psift_wait blocks until the features are on the host.  A frame
 * with more keypoints than the slot's buffers hold (they start at max_extrema, sift_pyramid.cu:148-159) is re-run
 * inside psift_wait with larger buffers (Pyramid::reallocExtrema, sift_pyramid.cu:179-215): a caller-owned device
 * image therefore has to stay valid until psift_wait returns. */
int  psift_submit_u8 ( psift_ctx* ctx, int w, int h, const uint8_t* host, size_t pitch_bytes, psift_job** out );
int  psift_submit_f32( psift_ctx* ctx, int w, int h, const float*   host, size_t pitch_bytes, psift_job** out );
/* same, but the image is in PAGE-LOCKED host memory (cudaHostAlloc / cudaHostRegister) that the caller leaves untouched
 * until psift_wait returns: it is uploaded straight from there, without the private copy that psift_submit_u8 makes
 * (SiftJob::SiftJob + Image::load copy twice, popsift.cpp:238-253, s_image.cu:71-79).  An extension: enqueue()'s
 * "the caller may free its buffer at once" does not hold for this entry. */
int  psift_submit_pinned_u8( psift_ctx* ctx, int w, int h, const uint8_t* pinned_host, size_t pitch_bytes, psift_job** out );
int  psift_submit_pinned_f32( psift_ctx* ctx, int w, int h, const float* pinned_host, size_t pitch_bytes, psift_job** out );   /* the same for float images */
/* page-locked host memory for such images (cudaHostAlloc / cudaFreeHost behind the C ABI, for callers without a CUDA binding);
 * fails without a CUDA device */
int  psift_host_alloc_pinned( void** out, size_t bytes );
void psift_host_free_pinned( void* p );
/* same, input already resident in device memory (no H2D inside) */
int  psift_submit_dev_u8( psift_ctx* ctx, int w, int h, const uint8_t* dev, size_t pitch_bytes, psift_job** out );
int  psift_wait( psift_job* job, psift_result* out );               /* SiftJob::get, popsift.cpp:300-313 */
/* psift_wait, and the caller TAKES the job's result buffers: out->feat / out->ext / out->desc stay valid after
 * psift_job_free, until psift_result_block_free( *block ).  The block is page-locked memory from a process-wide pool of
 * recycled blocks; the slot picks up another one for its next frame.  This is how PopSift's worker fills a FeaturesHost
 * without copying the descriptors (the reference downloads straight into the FeaturesHost it returns, after a
 * per-image cudaHostRegister: sift_pyramid.cu:281-321, features.cu:84-108). */
typedef struct psift_result_block psift_result_block;
int  psift_wait_take( psift_job* job, psift_result* out, psift_result_block** block );
void psift_result_block_free( psift_result_block* block );
void psift_job_free( psift_job* job );                              /* delete SiftJob */

/* ---- Device-resident results and matching (reference Config::MatchingMode) ----------------------------------
 * psift_wait_dev replaces PopSift::matchPrepareLoop + Pyramid::clone_device_descriptors (popsift.cpp:215-236,
 * sift_pyramid.cu:323-361): the job's features, descriptors and descriptor->feature map are cloned into device
 * buffers owned by the returned object; nothing is copied to the host.  May be called instead of, or in addition
 * to, psift_wait; the job is still released with psift_job_free.  psift_dev_features_alloc/free are FeaturesDev's
 * constructor and destructor (features.cu:126-161).
 * psift_match replaces FeaturesDev::match's compute_distance (features.cu:163-225, 268-302): for every left
 * descriptor, match3[3*i+0..2] = {index of the nearest right descriptor, index of the second nearest, accept}, accept
 * = d1/d2 < 0.8 on SQUARED L2 distances; dist2[2*i+0..1] = {d1, d2} (may be NULL).  Bit-identical to the reference's
 * arithmetic, so the integers agree exactly.  Operands on different GPUs: the right one is copied over. */
typedef struct psift_dev_features psift_dev_features;
typedef struct psift_dev_info {
    int            device, n_feat, n_desc;
    psift_feature* feat;        /* device: n_feat records with descriptor indices */
    void*          feat_ref;    /* device: n_feat popsift::Feature records (72 bytes, device Descriptor pointers, features.h:22-35) */
    float*         desc;        /* device: n_desc x 128 */
    int*           rev;         /* device: descriptor -> feature index (FeaturesDev::getReverseMap) */
} psift_dev_info;
int  psift_wait_dev( psift_job* job, psift_dev_features** out );
int  psift_dev_features_alloc( int device, int n_feat, int n_desc, psift_dev_features** out );
void psift_dev_features_free( psift_dev_features* f );
int  psift_dev_features_info( const psift_dev_features* f, psift_dev_info* info );
int  psift_match( const psift_dev_features* l, const psift_dev_features* r, int* match3_host, float* dist2_host );
int  psift_match_dev( int device, const float* l_desc_dev, int l_len, const float* r_desc_dev, int r_len,
                      int* match3_host, float* dist2_host );
/* The matcher without any allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  It has no counterpart in the reference.
 * Same arguments, same results as psift_match_dev; the device-side result scratch that psift_match_dev
 * cudaMallocs per call is a static device array here (PSIFT_NOALLOC_MATCH_MAX x 20 bytes = 1.25 MiB, held by every GPU
 * that loads the library until exit).  l_len > PSIFT_NOALLOC_MATCH_MAX returns PSIFT_E_INVALID with a message.  All calls
 * on a GPU share the array (it keeps the last call's contents; the rows a call uses are cleared first): calls that
 * overlap on the same GPU race - serialise them.  Calls no cudaMalloc / cudaFree / allocating routine. */
#define PSIFT_NOALLOC_MATCH_MAX 65536
int  psift_match_dev_noalloc( int device, const float* l_desc_dev, int l_len, const float* r_desc_dev, int r_len,
                              int* match3_host, float* dist2_host );
int  psift_dev_download( int device, const void* dev, void* host, size_t bytes );
int  psift_dev_upload( int device, void* dev, const void* host, size_t bytes );

/* Parity / profiling hooks (the reference's LogMode::All dumps, sift_octave.cu:110-187). Valid after
 * psift_wait and before the next submit on the same context slot. */
int  psift_octave_size( psift_ctx* ctx, int octave, int* w, int* h );
int  psift_debug_read_level( psift_ctx* ctx, int octave, int level, float* host_out );   /* Gaussian level, w*h floats */
int  psift_debug_read_dog  ( psift_ctx* ctx, int octave, int level, float* host_out );   /* G[level+1]-G[level] */
/* Host-only table builders (no device needed): the Gauss tables exactly as they are uploaded
 * (gauss_filter.cu:127-257): which = 0 inc, 1 abs_o0, 2 abs_oN, 3 dd; 32 taps per level; and the
 * scalar/descriptor constants (sift_constants.cu:22-53): scalars = sigma_k, threshold. */
int  psift_debug_gauss_table( const psift_config* cfg, int which, int level, float* filter32, float* sigma, int* span );
int  psift_debug_desc_tables( const psift_config* cfg, float* desc_gauss_1600, float* desc_tile_16, float* scalars2 );
/* octave-0 size and octave count for an input size (popsift.cpp:89-120) */
int  psift_pyramid_geometry( const psift_config* cfg, int w, int h, int* w0, int* h0, int* n_octaves );

/* Device-side timing of the last finished job on this context, milliseconds (CUDA events on
 * the context's stream): [0] upload, [1] pyramid, [2] extrema, [3] orientation, [4] descriptor+norm,
 * [5] download, [6] total.  For bench.py's roofline figures. */
int  psift_last_timing( psift_ctx* ctx, float ms[8] );
/* Per-launch timing of the dominant kernel: with time_levels on, every Gaussian-level launch of octave 0 (the six
 * large blur launches of a frame) is bracketed by a CUDA-event pair on the stream it is launched on; after psift_wait,
 * psift_last_level_timing returns their durations in milliseconds (ms[l] = level l, *n = levels).  The event records
 * sit between programmatically dependent launches, so frames run a little slower while this is on: bench.py switches
 * it on only for the frames it measures the roofline on. */
int  psift_debug_time_levels( psift_ctx* ctx, int on );
int  psift_last_level_timing( psift_ctx* ctx, float* ms, int cap, int* n );
/* Measurement / test hook for the C++ classes (host/bench_shim.cpp): one PopSift object on GPUs first_device ..
 * first_device + n_devices - 1 (n_devices 0: all from first_device), default Config; n_producers threads call
 * PopSift::enqueue( w, h, frames[i % n_distinct] ) -- pageable caller memory -- and SiftJob::get(), each keeping at most
 * `window` jobs outstanding; warm_frames untimed frames, then timed_frames frames between two wall-clock stamps
 * (*seconds).  n_feat / n_desc / checksum are summed over the timed frames (positions and descriptor bytes are read). */
int psift_cxx_stream( int first_device, int n_devices, int n_producers, int w, int h,
                      const unsigned char* const* frames, int n_distinct, int warm_frames, int timed_frames,
                      int window, double* seconds, long long* n_feat, long long* n_desc, double* checksum,
                      int* devices_used );
/* frame slots of the context = the most jobs that can be in flight on it (6, or PSIFT_SLOTS); a submit beyond that
 * returns PSIFT_E_STATE until a job has been waited for and freed */
int psift_slot_count( psift_ctx* ctx );
/* number of kernels this library launched since the context was created */
long long psift_launch_count( psift_ctx* ctx );
/* result bytes (counters, keypoint records, descriptors) queued for download to the host since the context was created */
long long psift_d2h_bytes( psift_ctx* ctx );
/* ---- one giant image over several GPUs (the reference has no such mode: it runs
 * config 5 un-tiled on one GPU, sift_pyramid.cu:108-165) -------------------------------------------
 * The image is cut into row bands, one per context (GPU).  Every context gets the WHOLE u8 image (it is
 * the float pyramid that does not fit or takes the time, not the bytes) and builds octaves 0..split for
 * its band plus a halo wide enough that everything the band owns is computed from exactly the data the
 * un-split run uses; all coordinates stay global, so the results are bit-identical and the union over
 * the bands is the un-split keypoint set.  The fused downscale leaves each context its rows of level 0 of
 * octave split+1 (psift_band_seed); the caller gathers those rows into one plane (the ONE exchange step:
 * all-gather over NVLink) and every context finishes the small octaves from it with
 * psift_submit_seed_dev_f32, keeping the keypoints whose detection row falls in its band.
 * psift_band_plan: split octave and band boundaries (rows of octave 0, bounds[parts+1]).  */
int  psift_band_plan( const psift_config* cfg, int w, int h, int parts, int max_seed_pixels, int* split_octave, int* bounds );
int  psift_submit_band_u8( psift_ctx* ctx, int w, int h, const uint8_t* img, size_t pitch, int on_device,
                           int y0, int y1, int split_octave, psift_job** out );
/* rows [row_lo,row_hi) of level 0 of octave split+1 (size seed_w x seed_h), device memory of the job's slot */
int  psift_band_seed( psift_job* job, const float** dev_rows, int* pitch_floats, int* row_lo, int* row_hi, int* seed_w, int* seed_h );
/* copy seed rows [row_from,row_to) to dst (device memory, row row_from first); waits for the job's stream */
int  psift_band_seed_copy( psift_job* job, float* dst_dev, size_t dst_pitch_bytes, int row_from, int row_to );
/* octaves first_octave.. of a w x h image from the given level-0 plane of octave first_octave */
int  psift_submit_seed_dev_f32( psift_ctx* ctx, int w, int h, int first_octave, const float* dev_plane, size_t pitch_bytes,
                                int y0, int y1, psift_job** out );

/* The whole of the above in one call: band pass on every context, exchange of the seed rows (device-to-device; peer
 * copies over NVLink between GPUs), seeded pass, merge.  ctxs: n_ctx contexts with the SAME configuration, normally one
 * per GPU (several on one GPU work too).  The result arrays are page-locked memory owned by the result: psift_banded_free. */
typedef struct psift_banded_result {
    int n_feat, n_desc, split_octave;
    psift_feature*  feat;
    psift_extremum* ext;
    float*          desc;
    int             halo_exchange;   /* 1: halos were exchanged per level between neighbouring bands; 0: every band recomputed its halo */
    void*           block;           /* owner of the three arrays (page-locked, recycled): released by psift_banded_free */
} psift_banded_result;
int  psift_extract_banded( psift_ctx** ctxs, int n_ctx, int w, int h, const uint8_t* host_img, size_t pitch,
                           int max_seed_pixels, psift_banded_result* out );
void psift_banded_free( psift_banded_result* r );

/* Device-side stopwatch over everything queued on the context between two marks: mark 0 is
 * stamped once all work queued so far has finished, mark 1 likewise; elapsed() waits for mark 1
 * and returns the CUDA-event time between them.  bench.py times its K steps with it. */
int  psift_debug_mark( psift_ctx* ctx, int which );
int  psift_debug_mark_elapsed( psift_ctx* ctx, float* ms );

#ifdef __cplusplus
}
#endif
#endif /* PSIFT_H */
