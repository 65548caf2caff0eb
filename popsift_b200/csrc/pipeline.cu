/* pipeline.cu -- per-device context, frame slots, launch sequence and the C ABI (include/psift.h).
 *
 * Replaces the reference's Pyramid/Octave/Image host objects (sift_pyramid.cu:108-239,
 * sift_octave.cu:33-89, s_image.cu:64-169) and its process-global device symbols
 * (sift_pyramid.cu:38-49) by a context that owns everything it touches, so several
 * contexts (one per GPU, or several per GPU) can live in one process.
 *
 * A frame runs entirely stream-ordered on its slot's stream: upload -> pyramid ->
 * DoG/extrema -> order -> orientation -> prefix -> descriptors(+norm) -> features ->
 * counter download.  Later stages read their work counts on the device, so the host
 * never has to wait in the middle of a frame (the reference synchronises the device
 * at least five times per image, see s_orientation.cu:351,427; sift_desc.cu:61,109).
 * psift_wait() then copies exactly n_feat features and n_desc descriptors.
 */
#include "common.h"
#include "kernels.h"
#include <cuda.h>                 /* CUtensorMap: the driver entry point is fetched through the runtime, nothing links libcuda */
#include <cstdio>
#include <cstring>
#include <cstdarg>
#include <vector>
#include <mutex>
#include <chrono>
#include <thread>
#include <string>
#include <cmath>
#include <new>
#include <cstdint>
#if defined( __x86_64__ ) && defined( __SSE2__ )
#include <emmintrin.h>
#endif

using namespace psift;

static thread_local char g_err[512] = "";
static int fail( int code, const char* fmt, ... )
{
    va_list ap; va_start( ap, fmt ); vsnprintf( g_err, sizeof g_err, fmt, ap ); va_end( ap );
    return code;
}
#define CU( call ) do { cudaError_t e_ = ( call ); if( e_ != cudaSuccess ) \
    return fail( PSIFT_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString( e_ ), __FILE__, __LINE__ ); } while( 0 )

namespace {

constexpr int N_EVENTS = 7;

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t  ev[N_EVENTS] = {};
    cudaEvent_t  ev_counters = nullptr;
    PyrStreams   ps = {};                                   /* chain = stream, two side streams, fork/join events */
    cudaEvent_t  lvl_ev[2 * MAX_LVL] = {};                  /* psift_debug_time_levels */
    /* the frame as a CUDA graph: captured from the very launches run_frame() makes, replayed for every later frame of
     * the same shape (the ~60 launches, event forks and joins of a frame cost the submitting thread 0.26 ms as stream
     * calls, 0.02 ms as one graph launch).  epoch counts every change of a pointer or size baked into the graph. */
    cudaGraphExec_t graph = nullptr;
    unsigned long long epoch = 1, graph_epoch = 0;
    long long graph_launches = 0;
    int  frames_direct = 0;
    /* results are downloaded by the frame itself, without a round trip through the waiting thread: the first pre_ext
     * keypoints and pre_desc descriptors (sized from the frames seen so far, +25 %) are copied right after the counters;
     * psift_wait fetches the rest only when a frame has more.  run_pre_*: what the frame in flight was queued with. */
    int  pre_ext = 0, pre_desc = 0, run_pre_ext = 0, run_pre_desc = 0;
    bool         busy = false;

    /* input */
    void*  d_img = nullptr; size_t img_pitch = 0; int img_w = 0, img_h = 0, img_elem = 0; size_t img_cap_w = 0, img_cap_h = 0;
    cudaTextureObject_t img_tex = 0;
    const void* tex_bound_ptr = nullptr; size_t tex_bound_pitch = 0;
    void*  h_img = nullptr; size_t h_img_cap = 0;          /* pinned private copy of the caller's image */

    /* pyramid */
    float* arena = nullptr; size_t arena_floats = 0;
    PyrLayout L;
    int in_w = 0, in_h = 0;
    int reg_kind = 0, reg_a = 0, reg_b = 0, reg_c = 0, reg_exchange = 0;   /* region the layout was built for (see Region) */
    std::vector<cudaTextureObject_t> lintex_h;
    cudaTextureObject_t* lintex_d = nullptr;
    void*  tmaps_d = nullptr;            /* MAX_OCT tensor maps (TMA descriptors) of the octaves' level stacks, for k_extrema */
    std::vector<cudaTextureObject_t> scratch_tex;          /* Gauss mode "relative": the scratch plane, one view per octave */
    float* scratch = nullptr;

    /* keypoints */
    Cand* cand = nullptr; Cand* sorted = nullptr; Cand* sort_tmp = nullptr;
    int cand_stride = 0, ext_cap = 0, desc_cap = 0;
    psift_extremum* ext = nullptr;
    psift_feature*  feat = nullptr;
    int*   desc_map = nullptr;
    DescWork* desc_work = nullptr;
    float* desc = nullptr;
    Counters* ct = nullptr;
    void*  filt = nullptr;                                  /* grid filter scratch */

    /* host results (pinned) */
    Counters* h_ct = nullptr;
    psift_extremum* h_ext = nullptr; psift_feature* h_feat = nullptr; float* h_desc = nullptr;
    int h_ext_cap = 0, h_desc_cap = 0;
};

} // namespace

struct psift_ctx {
    int device = 0;
    int n_sm = 148;                                         /* queried at create (grid sizes are multiples of it) */
    psift_config cfg;
    GaussTables  gt;
    Consts       K;
    DescTables   dt;
    DescTables*  d_dt = nullptr;
    std::vector<Slot> slots;
    int next_slot = 0;
    long long launches = 0;
    float last_ms[8] = { 0 };
    long long d2h_bytes = 0;                                /* result bytes queued for download so far */
    float last_lvl_ms[MAX_LVL] = { 0 };                     /* octave-0 launches of the last finished job, when timed */
    int   last_lvl_n = 0;
    bool  time_levels = false;
    int last_slot = -1;
    cudaStream_t mark_stream = nullptr;                     /* psift_debug_mark */
    cudaEvent_t  mark_ev[2] = { nullptr, nullptr };
    std::vector<cudaEvent_t> mark_join;
    std::mutex mu;
};

struct psift_job {
    psift_ctx* ctx;
    int slot;
    bool waited;
    int status;
};

static bool linear_needed( int desc_mode ) { return desc_mode == PSIFT_DESC_IGRID || desc_mode == PSIFT_DESC_NOTILE || desc_mode == PSIFT_DESC_ILOOP; }

static void free_slot_pyramid( Slot& s )
{
    for( auto t : s.lintex_h ) if( t ) cudaDestroyTextureObject( t );
    s.lintex_h.clear();
    for( auto t : s.scratch_tex ) if( t ) cudaDestroyTextureObject( t );
    s.scratch_tex.clear();
    if( s.lintex_d ) { cudaFree( s.lintex_d ); s.lintex_d = nullptr; }
    if( s.tmaps_d ) { cudaFree( s.tmaps_d ); s.tmaps_d = nullptr; }
    if( s.arena ) { cudaFree( s.arena ); s.arena = nullptr; s.arena_floats = 0; }
    s.in_w = s.in_h = 0;
}

static void free_slot( Slot& s )
{
    free_slot_pyramid( s );
    if( s.img_tex ) cudaDestroyTextureObject( s.img_tex );
    if( s.d_img ) cudaFree( s.d_img );
    if( s.h_img ) cudaFreeHost( s.h_img );
    cudaFree( s.cand ); cudaFree( s.sorted ); cudaFree( s.sort_tmp ); cudaFree( s.ext ); cudaFree( s.feat ); cudaFree( s.desc_map ); cudaFree( s.desc_work ); cudaFree( s.desc ); cudaFree( s.ct ); cudaFree( s.filt );
    if( s.h_ct ) cudaFreeHost( s.h_ct );
    if( s.h_ext ) cudaFreeHost( s.h_ext );
    if( s.h_feat ) cudaFreeHost( s.h_feat );
    if( s.h_desc ) cudaFreeHost( s.h_desc );
    for( auto& e : s.ev ) if( e ) cudaEventDestroy( e );
    for( auto& e : s.lvl_ev ) if( e ) cudaEventDestroy( e );
    if( s.graph ) { cudaGraphExecDestroy( s.graph ); s.graph = nullptr; }
    if( s.ev_counters ) cudaEventDestroy( s.ev_counters );
    for( auto& e : s.ps.seeded ) if( e ) cudaEventDestroy( e );
    for( int k = 0; k < 2; k++ ) {
        if( s.ps.side_done[k] ) cudaEventDestroy( s.ps.side_done[k] );
        if( s.ps.side[k] ) cudaStreamDestroy( s.ps.side[k] );
    }
    if( s.stream ) cudaStreamDestroy( s.stream );
    s = Slot();
}

/* (re)build the pyramid geometry of a slot for an input size (Pyramid::Pyramid / resetDimensions,
 * sift_pyramid.cu:108-177; octave sizes by repeated ceil(/2), :132-134) */
/* What a job covers: the whole image (the normal case), a row band of a giant image up to a
 * split octave, or the octaves above the split from a given level-0 plane. */
struct Region {
    int kind = 0;            /* 0 whole image, 1 band, 2 seeded upper octaves */
    int y0 = 0, y1 = 0;      /* owned rows of octave 0 */
    int split = 0;           /* band: last octave built; seeded: first octave built */
    int exchange = 0;        /* band: 1 = the band computes its own rows only; `halo` rows above and below arrive from the
                                neighbouring bands after every level (psift_extract_banded) */
};

/* rows of octave 0 a band must hold so that everything it owns, in every octave up to `split`, is
 * computed from exact data: the recurrence of SURVEY appendix A with this configuration's spans */
static int band_halo_rows( const psift_ctx* c, int split )
{
    const int n_lvl = c->cfg.levels + 3, ds_level = n_lvl - 3;
    int s_all = 0, s_ds = 0;
    for( int l = 1; l < n_lvl; l++ ) { s_all += c->gt.inc[l].span; if( l <= ds_level ) s_ds += c->gt.inc[l].span; }
    const int s_0 = c->gt.dd[0].span + c->gt.inc[0].span;
    /* window of the orientation / descriptor / refinement stages around a keypoint, octave pixels */
    const int win = (int)ceilf( 3.6f * 3.0f * c->cfg.sigma * powf( 2.0f, (float)n_lvl / (float)c->cfg.levels ) ) + 16;
    int e = win + s_all;
    for( int o = split; o > 0; o-- ) e = std::max( 2 * e + s_ds + 2, win + s_all );
    return e + s_0 + 2;
}

static int setup_pyramid_impl( psift_ctx* c, Slot& s, int w, int h, const Region& reg );
static int setup_pyramid( psift_ctx* c, Slot& s, int w, int h, const Region& reg )
{
    const int rc = setup_pyramid_impl( c, s, w, h, reg );
    /* a failure after the geometry was committed must not leave a slot that looks ready for this size */
    if( rc != PSIFT_OK ) { s.in_w = s.in_h = 0; s.epoch++; }
    return rc;
}
/* rows a band keeps of its neighbours' planes when halos are exchanged per level: the window of the orientation /
 * descriptor / refinement stages around a keypoint (octave pixels, the same in every octave), which is more than any
 * filter's reach */
static int band_exchange_rows( const psift_ctx* c )
{
    const int n_lvl = c->cfg.levels + 3;
    int s_max = 0;
    for( int l = 0; l < n_lvl; l++ ) s_max = std::max( s_max, c->gt.inc[l].span );
    s_max = std::max( s_max, c->gt.dd[0].span );
    const int win = (int)ceilf( 3.6f * 3.0f * c->cfg.sigma * powf( 2.0f, (float)n_lvl / (float)c->cfg.levels ) ) + 16;
    return ( std::max( win, s_max ) + 2 + 1 ) / 2 * 2;
}

static int setup_pyramid_impl( psift_ctx* c, Slot& s, int w, int h, const Region& reg )
{
    if( s.in_w == w && s.in_h == h && s.arena && s.reg_exchange == reg.exchange && s.reg_kind == reg.kind && s.reg_a == reg.y0 && s.reg_b == reg.y1 && s.reg_c == reg.split ) return PSIFT_OK;
    /* everything that can be refused is refused before the slot's state is touched */
    if( reg.kind == 1 && c->cfg.gauss_mode == PSIFT_GAUSS_VLFEAT_RELATIVE )
        return fail( PSIFT_E_INVALID, "the relative Gauss mode is not available for banded extraction" );
    int w0, h0;
    octave0_size( c->cfg, w, h, w0, h0 );
    const int n_oct = octave_count( c->cfg, w, h );
    const int n_lvl = c->cfg.levels + 3;
    PyrLayout L;
    memset( &L, 0, sizeof L );
    L.n_oct = n_oct; L.n_lvl = n_lvl;
    L.oct_begin = 0; L.oct_end = n_oct; L.alloc_end = n_oct;
    if( reg.kind == 1 ) {
        if( reg.split < 0 || reg.split >= n_oct ) return fail( PSIFT_E_INVALID, "split octave %d out of range (the image has %d octaves)", reg.split, n_oct );
        L.oct_end = reg.split + 1; L.alloc_end = std::min( n_oct, reg.split + 2 );
    } else if( reg.kind == 2 ) {
        if( reg.split < 1 || reg.split >= n_oct ) return fail( PSIFT_E_INVALID, "first octave %d out of range (the image has %d octaves)", reg.split, n_oct );
        L.oct_begin = reg.split;
    }
    if( reg.kind != 0 && ( reg.y0 < 0 || reg.y1 <= reg.y0 ) ) return fail( PSIFT_E_INVALID, "bad band rows [%d, %d)", reg.y0, reg.y1 );
    size_t total = 0;
    int ow = w0, oh = h0;
    std::vector<size_t> off( n_oct, 0 );
    int lo0 = 0, hi0 = h0;
    if( reg.kind == 1 ) {
        const int halo = band_halo_rows( c, reg.split ), al = 1 << ( reg.split + 1 );
        lo0 = std::max( 0, reg.y0 - halo ) / al * al;
        hi0 = std::min( h0, ( std::min( h0, reg.y1 ) + halo + al - 1 ) / al * al );
    }
    for( int o = 0; o < n_oct; o++ ) {
        if( ow < 1 || oh < 1 ) return fail( PSIFT_E_INVALID, "image %dx%d too small for %d octaves", w, h, n_oct );
        L.w[o] = ow; L.h[o] = oh;
        L.row_lo[o] = 0; L.row_hi[o] = oh; L.own_lo[o] = 0; L.own_hi[o] = oh;
        if( reg.kind == 1 && !reg.exchange ) {
            L.row_lo[o] = lo0 >> o;
            L.row_hi[o] = std::min( oh, o == 0 ? hi0 : ( L.row_hi[o - 1] + 1 ) >> 1 );
        }
        if( reg.kind != 0 ) {
            L.own_lo[o] = std::min( oh, reg.y0 >> o );
            L.own_hi[o] = reg.y1 >= h0 ? oh : std::min( oh, reg.y1 >> o );
        }
        L.cmp_lo[o] = L.row_lo[o]; L.cmp_hi[o] = L.row_hi[o];
        if( reg.kind == 1 && reg.exchange ) {
            /* own rows are computed here (band starts are multiples of 2^(split+1): even in every octave up to the split),
             * the halo rows are copied in; the seed octave holds the band's own rows of level 0 */
            const int E = o <= reg.split ? band_exchange_rows( c ) : 0;
            L.cmp_lo[o] = L.own_lo[o]; L.cmp_hi[o] = L.own_hi[o];
            L.row_lo[o] = std::max( 0, L.own_lo[o] - E ); L.row_hi[o] = std::min( oh, L.own_hi[o] + E );
        }
        L.pitch[o] = ( ow + 31 ) / 32 * 32;
        const bool held = o >= L.oct_begin && o < L.alloc_end;
        L.lstride[o] = held ? ( (size_t)L.pitch[o] * ( L.row_hi[o] - L.row_lo[o] ) + 127 ) / 128 * 128 : 0;   /* keep planes 512-byte aligned (textures) */
        off[o] = total;
        total += L.lstride[o] * ( o < L.oct_end ? n_lvl : 1 );          /* the seed octave of a band holds level 0 only */
        ow = (int)ceilf( ow / 2.0f );
        oh = (int)ceilf( oh / 2.0f );
    }
    const bool relative = c->cfg.gauss_mode == PSIFT_GAUSS_VLFEAT_RELATIVE;
    const size_t scratch_off = total;
    if( relative ) total += L.lstride[0];                                   /* horizontal-pass plane, octave-0 sized */
    if( total > s.arena_floats ) {
        free_slot_pyramid( s );
        CU( cudaMalloc( &s.arena, total * sizeof(float) ) );
        s.arena_floats = total;
    } else {
        for( auto t : s.lintex_h ) if( t ) cudaDestroyTextureObject( t );
        s.lintex_h.clear();
        for( auto t : s.scratch_tex ) if( t ) cudaDestroyTextureObject( t );
        s.scratch_tex.clear();
    }
    s.scratch = relative ? s.arena + scratch_off : nullptr;
    /* base[o] is where row 0 of the plane WOULD be: kernels address with global rows */
    for( int o = 0; o < n_oct; o++ )
        L.base[o] = reinterpret_cast<float*>( reinterpret_cast<uintptr_t>( s.arena + off[o] ) - sizeof(float) * (size_t)L.row_lo[o] * L.pitch[o] );
    s.L = L; s.in_w = w; s.in_h = h; s.epoch++;
    s.reg_kind = reg.kind; s.reg_a = reg.y0; s.reg_b = reg.y1; s.reg_c = reg.split; s.reg_exchange = reg.exchange;

    /* TMA descriptors: every octave's stack of levels as a 3-D tensor (column, stored row, level), read by k_extrema in
     * boxes of one tile plus halo.  Columns go up to the pitch (the padding is readable), rows are the stored ones:
     * a box that sticks out is filled with zeros, which no candidate ever looks at. */
    {
        typedef CUresult ( *EncodeFn )( CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill );
        static EncodeFn encode = nullptr;
        if( !encode ) {
            void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
            CU( cudaGetDriverEntryPoint( "cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr ) );
            if( !fn || qr != cudaDriverEntryPointSuccess ) return fail( PSIFT_E_CUDA, "cuTensorMapEncodeTiled is not available in this driver" );
            encode = (EncodeFn)fn;
        }
        std::vector<CUtensorMap> maps( MAX_OCT );
        memset( maps.data(), 0, sizeof(CUtensorMap) * MAX_OCT );
        int box_w = 0, box_h = 0; extrema_box( &box_w, &box_h );
        for( int o = L.oct_begin; o < L.oct_end; o++ ) {
            const cuuint64_t dims[3]    = { (cuuint64_t)L.pitch[o], (cuuint64_t)( L.row_hi[o] - L.row_lo[o] ), (cuuint64_t)n_lvl };
            const cuuint64_t strides[2] = { (cuuint64_t)L.pitch[o] * sizeof(float), (cuuint64_t)L.lstride[o] * sizeof(float) };
            const cuuint32_t box[3]     = { (cuuint32_t)box_w, (cuuint32_t)box_h, (cuuint32_t)n_lvl };
            const cuuint32_t estr[3]    = { 1, 1, 1 };
            void* first = L.base[o] + (size_t)L.row_lo[o] * L.pitch[o];
            const CUresult r = encode( &maps[o], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, first, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE );
            if( r != CUDA_SUCCESS ) return fail( PSIFT_E_CUDA, "cuTensorMapEncodeTiled failed (%d) for octave %d", (int)r, o );
        }
        if( !s.tmaps_d ) CU( cudaMalloc( &s.tmaps_d, sizeof(CUtensorMap) * MAX_OCT ) );
        CU( cudaMemcpy( s.tmaps_d, maps.data(), sizeof(CUtensorMap) * MAX_OCT, cudaMemcpyHostToDevice ) );
    }

    if( linear_needed( c->cfg.desc_mode ) || relative ) {
        /* bilinear point-of-view of every Gaussian level (Octave::alloc_data_tex, sift_octave.cu:211-248) */
        s.lintex_h.assign( (size_t)n_oct * n_lvl, 0 );
        for( int o = L.oct_begin; o < L.oct_end; o++ )
            for( int l = 0; l < n_lvl; l++ ) {
                cudaResourceDesc rd; memset( &rd, 0, sizeof rd );
                rd.resType = cudaResourceTypePitch2D;
                /* a band binds the rows it stores; the kernels shift the sample row by row_lo (gradient_lin) */
                rd.res.pitch2D.devPtr = L.base[o] + (size_t)l * L.lstride[o] + (size_t)L.row_lo[o] * L.pitch[o];
                rd.res.pitch2D.desc = cudaCreateChannelDesc( 32, 0, 0, 0, cudaChannelFormatKindFloat );
                rd.res.pitch2D.width = L.w[o]; rd.res.pitch2D.height = L.row_hi[o] - L.row_lo[o];
                rd.res.pitch2D.pitchInBytes = (size_t)L.pitch[o] * sizeof(float);
                cudaTextureDesc td; memset( &td, 0, sizeof td );
                td.normalizedCoords = 0;
                td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
                td.readMode = cudaReadModeElementType;
                td.filterMode = cudaFilterModeLinear;
                CU( cudaCreateTextureObject( &s.lintex_h[(size_t)o * n_lvl + l], &rd, &td, nullptr ) );
            }
        if( s.lintex_d ) cudaFree( s.lintex_d );
        CU( cudaMalloc( &s.lintex_d, s.lintex_h.size() * sizeof(cudaTextureObject_t) ) );
        CU( cudaMemcpy( s.lintex_d, s.lintex_h.data(), s.lintex_h.size() * sizeof(cudaTextureObject_t), cudaMemcpyHostToDevice ) );
    }
    if( relative ) {
        /* the scratch plane seen with every octave's geometry (the reference's intermediate array, sift_octave.cu:250-330) */
        s.scratch_tex.assign( (size_t)n_oct, 0 );
        for( int o = 0; o < n_oct; o++ ) {
            cudaResourceDesc rd; memset( &rd, 0, sizeof rd );
            rd.resType = cudaResourceTypePitch2D;
            rd.res.pitch2D.devPtr = s.scratch;
            rd.res.pitch2D.desc = cudaCreateChannelDesc( 32, 0, 0, 0, cudaChannelFormatKindFloat );
            rd.res.pitch2D.width = L.w[o]; rd.res.pitch2D.height = L.h[o];
            rd.res.pitch2D.pitchInBytes = (size_t)L.pitch[o] * sizeof(float);
            cudaTextureDesc td; memset( &td, 0, sizeof td );
            td.normalizedCoords = 0;
            td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
            td.readMode = cudaReadModeElementType;
            td.filterMode = cudaFilterModeLinear;
            CU( cudaCreateTextureObject( &s.scratch_tex[o], &rd, &td, nullptr ) );
        }
    }
    return PSIFT_OK;
}

static int bind_input_texture( Slot& s, const void* dptr, size_t pitch, int w, int h, int elem )
{
    if( s.img_tex && s.tex_bound_ptr == dptr && s.tex_bound_pitch == pitch && s.img_w == w && s.img_h == h && s.img_elem == elem ) return PSIFT_OK;
    if( s.img_tex ) { cudaDestroyTextureObject( s.img_tex ); s.img_tex = 0; }
    /* Image::createTexture / ImageFloat::createTexture, s_image.cu:140-169, 264-293 */
    cudaResourceDesc rd; memset( &rd, 0, sizeof rd );
    rd.resType = cudaResourceTypePitch2D;
    rd.res.pitch2D.devPtr = const_cast<void*>( dptr );
    rd.res.pitch2D.desc = ( elem == 1 ) ? cudaCreateChannelDesc( 8, 0, 0, 0, cudaChannelFormatKindUnsigned )
                                        : cudaCreateChannelDesc( 32, 0, 0, 0, cudaChannelFormatKindFloat );
    rd.res.pitch2D.pitchInBytes = pitch; rd.res.pitch2D.width = w; rd.res.pitch2D.height = h;
    cudaTextureDesc td; memset( &td, 0, sizeof td );
    td.normalizedCoords = 1;
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.readMode = ( elem == 1 ) ? cudaReadModeNormalizedFloat : cudaReadModeElementType;
    td.filterMode = cudaFilterModeLinear;
    CU( cudaCreateTextureObject( &s.img_tex, &rd, &td, nullptr ) );
    s.epoch++;
    s.tex_bound_ptr = dptr; s.tex_bound_pitch = pitch; s.img_w = w; s.img_h = h; s.img_elem = elem;
    return PSIFT_OK;
}

/* Buffers sized by the number of keypoints / descriptors of a frame.  The reference starts from the same sizes
 * (max_extrema total, sift_pyramid.cu:148-159) and reallocates when a frame needs more
 * (Pyramid::reallocExtrema, sift_pyramid.cu:179-215); psift_wait() does the same and re-runs the frame. */
static int alloc_keypoint_buffers( Slot& s, int ext_cap, int desc_cap )
{
    cudaFree( s.sorted ); cudaFree( s.sort_tmp ); cudaFree( s.ext ); cudaFree( s.feat ); cudaFree( s.desc_map ); cudaFree( s.desc_work ); cudaFree( s.desc ); cudaFree( s.filt );
    s.sort_tmp = nullptr; s.sorted = nullptr; s.ext = nullptr; s.feat = nullptr; s.desc_map = nullptr; s.desc_work = nullptr; s.desc = nullptr; s.filt = nullptr;
    s.ext_cap = ext_cap; s.desc_cap = desc_cap; s.epoch++;
    CU( cudaMalloc( &s.sorted, sizeof(Cand) * (size_t)ext_cap ) );
    CU( cudaMalloc( &s.sort_tmp, sizeof(Cand) * (size_t)ext_cap ) );
    CU( cudaMalloc( &s.ext,    sizeof(psift_extremum) * (size_t)ext_cap ) );
    CU( cudaMalloc( &s.feat,   sizeof(psift_feature) * (size_t)ext_cap ) );
    CU( cudaMalloc( &s.desc_map, sizeof(int) * (size_t)desc_cap ) );
    CU( cudaMalloc( &s.desc_work, sizeof(DescWork) * (size_t)desc_cap ) );
    CU( cudaMalloc( &s.desc,   sizeof(float) * 128 * (size_t)desc_cap ) );
    CU( cudaMalloc( &s.filt, filter_scratch_bytes( ext_cap ) ) );
    return PSIFT_OK;
}

static int init_slot( psift_ctx* c, Slot& s )
{
    CU( cudaStreamCreateWithFlags( &s.stream, cudaStreamNonBlocking ) );
    for( auto& e : s.ev ) CU( cudaEventCreate( &e ) );
    CU( cudaEventCreateWithFlags( &s.ev_counters, cudaEventDisableTiming ) );
    s.ps.chain = s.stream;
    if( !getenv( "PSIFT_SERIAL_PYRAMID" ) ) {
        for( int k = 0; k < 2; k++ ) {
            CU( cudaStreamCreateWithFlags( &s.ps.side[k], cudaStreamNonBlocking ) );
            CU( cudaEventCreateWithFlags( &s.ps.side_done[k], cudaEventDisableTiming ) );
        }
        for( auto& e : s.ps.seeded ) CU( cudaEventCreateWithFlags( &e, cudaEventDisableTiming ) );
    }
    const int me = c->cfg.max_extrema;
    s.cand_stride = me;
    CU( cudaMalloc( &s.cand,   sizeof(Cand) * (size_t)me * MAX_OCT ) );
    int rc = alloc_keypoint_buffers( s, me, std::max( 2 * me, me + me / 4 ) );   /* sift_pyramid.cu:148-159 */
    if( rc ) return rc;
    CU( cudaMalloc( &s.ct, sizeof(CountersAndSort) ) );           /* the counters, and behind them the sort's bucket counters */
    CU( cudaMallocHost( &s.h_ct, sizeof(Counters) ) );
    return PSIFT_OK;
}

/* Host result buffers are page-locked BLOCKS (keypoint records + descriptors) that can leave the slot: psift_wait_take
 * hands the block of a finished job to the caller (the C++ FeaturesHost keeps it: no copy of the descriptors), and the
 * slot picks up another one from a process-wide pool of recycled blocks for its next frame.  The downloads into a block
 * are plain stream copies queued behind the frame's graph, so a new block needs no re-capture. */
struct psift_result_block { psift_extremum* ext; psift_feature* feat; float* desc; int ext_cap, desc_cap; };
namespace {
struct ResultPool {
    std::mutex mu;
    std::vector<psift_result_block*> idle;
    psift_result_block* get( int n_ext, int n_desc )
    {
        {
            std::lock_guard<std::mutex> lk( mu );
            size_t best = idle.size();                     /* best fit: a frame must not walk off with a giant image's block */
            for( size_t i = 0; i < idle.size(); i++ )
                if( idle[i]->ext_cap >= n_ext && idle[i]->desc_cap >= n_desc && ( best == idle.size() || idle[i]->desc_cap < idle[best]->desc_cap ) ) best = i;
            if( best < idle.size() ) { psift_result_block* b = idle[best]; idle[best] = idle.back(); idle.pop_back(); return b; }
        }
        psift_result_block* b = new( std::nothrow ) psift_result_block();
        if( !b ) return nullptr;
        b->ext_cap  = std::max( 4096, n_ext + n_ext / 2 );
        b->desc_cap = std::max( 8192, n_desc + n_desc / 2 );
        /* portable: a block may come back from one GPU's context and serve another's */
        if( cudaHostAlloc( &b->ext,  sizeof(psift_extremum) * (size_t)b->ext_cap, cudaHostAllocPortable ) != cudaSuccess ||
            cudaHostAlloc( &b->feat, sizeof(psift_feature) * (size_t)b->ext_cap, cudaHostAllocPortable ) != cudaSuccess ||
            cudaHostAlloc( &b->desc, sizeof(float) * 128 * (size_t)b->desc_cap, cudaHostAllocPortable ) != cudaSuccess ) {
            cudaGetLastError();
            destroy( b );
            return nullptr;
        }
        return b;
    }
    static void destroy( psift_result_block* b )
    {
        if( !b ) return;
        if( b->ext ) cudaFreeHost( b->ext );
        if( b->feat ) cudaFreeHost( b->feat );
        if( b->desc ) cudaFreeHost( b->desc );
        delete b;
    }
    void put( psift_result_block* b )
    {
        if( !b ) return;
        psift_result_block* victim = nullptr;
        {
            std::lock_guard<std::mutex> lk( mu );
            idle.push_back( b );
            if( idle.size() > 64 ) {
                /* full: the SMALLEST block goes (64 frame-sized blocks left behind by a stream must not push out the one
                 * block of a giant image, which costs 100 ms to page-lock again) */
                size_t k = 0;
                for( size_t i = 1; i < idle.size(); i++ )
                    if( (long long)idle[i]->desc_cap * 512 + (long long)idle[i]->ext_cap * 96 < (long long)idle[k]->desc_cap * 512 + (long long)idle[k]->ext_cap * 96 ) k = i;
                victim = idle[k]; idle[k] = idle.back(); idle.pop_back();
            }
        }
        destroy( victim );
    }
};
ResultPool& result_pool() { static ResultPool* p = new ResultPool; return *p; }    /* never torn down: results may outlive every context */
} // namespace

static psift_result_block* detach_block( Slot& s )
{
    if( !s.h_ext ) return nullptr;
    psift_result_block* b = new psift_result_block();
    b->ext = s.h_ext; b->feat = s.h_feat; b->desc = s.h_desc; b->ext_cap = s.h_ext_cap; b->desc_cap = s.h_desc_cap;
    s.h_ext = nullptr; s.h_feat = nullptr; s.h_desc = nullptr; s.h_ext_cap = s.h_desc_cap = 0;
    return b;
}

static int ensure_host_results( Slot& s, int n_ext, int n_desc )
{
    if( s.h_ext && n_ext <= s.h_ext_cap && n_desc <= s.h_desc_cap ) return PSIFT_OK;
    result_pool().put( detach_block( s ) );
    psift_result_block* b = result_pool().get( std::max( n_ext, s.pre_ext ), std::max( n_desc, s.pre_desc ) );
    if( !b ) return fail( PSIFT_E_CUDA, "out of page-locked host memory for %d keypoints / %d descriptors", n_ext, n_desc );
    s.h_ext = b->ext; s.h_feat = b->feat; s.h_desc = b->desc; s.h_ext_cap = b->ext_cap; s.h_desc_cap = b->desc_cap;
    delete b;
    return PSIFT_OK;
}

/* the frame's own download of its results: the first pre_ext keypoints and pre_desc descriptors, right behind the kernels */
static int queue_result_copies( Slot& s )
{
    cudaStream_t st = s.stream;
    if( s.pre_ext > 0 ) {
        CU( cudaMemcpyAsync( s.h_feat, s.feat, sizeof(psift_feature)  * (size_t)s.pre_ext, cudaMemcpyDeviceToHost, st ) );
        CU( cudaMemcpyAsync( s.h_ext,  s.ext,  sizeof(psift_extremum) * (size_t)s.pre_ext, cudaMemcpyDeviceToHost, st ) );
    }
    if( s.pre_desc > 0 )
        CU( cudaMemcpyAsync( s.h_desc, s.desc, sizeof(float) * 128 * (size_t)s.pre_desc, cudaMemcpyDeviceToHost, st ) );
    return PSIFT_OK;
}

static bool graphs_enabled()
{
    static int v = -1;
    if( v < 0 ) { const char* e = getenv( "PSIFT_GRAPH" ); v = e ? atoi( e ) : 1; }
    return v != 0;
}

/* the launches of one frame, onto the slot's stream (directly, or into a stream capture) */
enum { FRAME_BEGIN = 1, FRAME_PYRAMID = 2, FRAME_REST = 4, FRAME_ALL = 7 };
static int enqueue_frame( psift_ctx* c, Slot& s, bool capturing, int parts = FRAME_ALL )
{
    cudaStream_t st = s.stream;
    const PyrLayout& L = s.L;
    /* stage stamps: inside a graph they must be external event-record nodes to be readable afterwards */
    auto stamp = [&]( cudaEvent_t e ) { return capturing ? cudaEventRecordWithFlags( e, st, cudaEventRecordExternal ) : cudaEventRecord( e, st ); };
    if( parts & FRAME_BEGIN ) {
    CU( stamp( s.ev[1] ) );
    CU( cudaMemsetAsync( s.ct, 0, sizeof(CountersAndSort), st ) );           /* reset_extrema_mgmt, sift_pyramid.cu:363-370 */
    }
    float shift = 0.5f;                                              /* s_pyramid_build.cu:110-114 */
    if( c->cfg.sift_mode == PSIFT_MODE_POPSIFT || c->cfg.sift_mode == PSIFT_MODE_VLFEAT )
        shift = 0.5f * powf( 2.0f, c->cfg.upscale_factor );
    s.ps.level_tex   = s.lintex_h.empty() ? nullptr : s.lintex_h.data();
    s.ps.scratch_tex = s.scratch_tex.empty() ? nullptr : s.scratch_tex.data();
    s.ps.scratch     = s.scratch;
    s.ps.lvl_ev = nullptr;
    if( !capturing && c->time_levels && L.oct_begin == 0 ) {
        for( int k = 0; k < 2 * L.n_lvl; k++ ) if( !s.lvl_ev[k] ) CU( cudaEventCreate( &s.lvl_ev[k] ) );
        s.ps.lvl_ev = s.lvl_ev;
    }
    if( parts & FRAME_PYRAMID )
    CU( launch_pyramid( L, s.img_tex, shift, powf( 2.0f, c->cfg.upscale_factor ), c->cfg.gauss_mode, c->cfg.scaling_mode, c->gt, s.ps, &c->launches ) );
    if( !( parts & FRAME_REST ) ) return PSIFT_OK;
    CU( stamp( s.ev[2] ) );
    CU( launch_extrema( L, c->K, s.tmaps_d, s.cand, s.cand_stride, s.sorted, s.sort_tmp, s.ext_cap, s.ct, &reinterpret_cast<CountersAndSort*>( s.ct )->ss, st, &c->launches ) );
    CU( launch_grid_filter( L, c->cfg, s.sorted, s.cand, s.ext_cap, s.ct, s.filt, st, &c->launches ) );
    CU( stamp( s.ev[3] ) );
    CU( launch_orientation( L, s.sorted, s.ct, s.ext, s.desc_map, s.desc_work, s.desc_cap, c->n_sm * orientation_blocks_per_sm(), st, &c->launches ) );
    CU( stamp( s.ev[4] ) );
    CU( launch_descriptors( L, c->K, s.ext, s.ct, s.desc_work, s.lintex_d, c->d_dt, s.desc, s.feat, c->n_sm * 4, st, &c->launches ) );
    CU( stamp( s.ev[5] ) );
    CU( cudaMemcpyAsync( s.h_ct, s.ct, sizeof(Counters), cudaMemcpyDeviceToHost, st ) );
    if( !capturing ) return queue_result_copies( s );        /* a graph leaves them to run_frame: their target may change */
    return PSIFT_OK;
}

/* queue one frame on the slot's stream; the input texture is already bound.  Whole-image frames run as a CUDA graph
 * from their second use on (the first one runs as plain launches: it also sets the kernels' attributes, which must
 * not happen inside a capture); banded / seeded passes and per-level timing keep plain launches. */
static int run_frame( psift_ctx* c, Slot& s, bool pyramid_done = false )
{
    s.pre_ext = std::min( s.pre_ext, s.ext_cap ); s.pre_desc = std::min( s.pre_desc, s.desc_cap );
    if( s.reg_kind != 0 ) s.pre_ext = s.pre_desc = 0;      /* parts of a banded image: whoever merges them places the download */
    if( s.pre_ext > 0 || s.pre_desc > 0 ) { const int rc = ensure_host_results( s, s.pre_ext, s.pre_desc ); if( rc ) return rc; }
    s.run_pre_ext = s.pre_ext; s.run_pre_desc = s.pre_desc;
    c->d2h_bytes += (long long)s.pre_ext * ( sizeof(psift_feature) + sizeof(psift_extremum) ) + (long long)s.pre_desc * 512 + sizeof(Counters);
    if( pyramid_done ) {
        /* the arena still holds this frame's planes (a re-run with larger keypoint buffers, or a band whose pyramid was
         * built level by level together with its neighbours): everything from the DoG stage on */
        const int rc = enqueue_frame( c, s, false, FRAME_BEGIN | FRAME_REST );
        if( rc ) return rc;
        CU( cudaEventRecord( s.ev_counters, s.stream ) );
        return PSIFT_OK;
    }
    const bool want_graph = graphs_enabled() && s.reg_kind == 0 && !c->time_levels && s.frames_direct > 0;
    if( want_graph && ( !s.graph || s.graph_epoch != s.epoch ) ) {
        if( s.graph ) { cudaGraphExecDestroy( s.graph ); s.graph = nullptr; }
        const long long l0 = c->launches;
        cudaGraph_t g = nullptr;
        CU( cudaStreamBeginCapture( s.stream, cudaStreamCaptureModeRelaxed ) );
        const int rc = enqueue_frame( c, s, true );
        const cudaError_t ee = cudaStreamEndCapture( s.stream, &g );
        s.graph_launches = c->launches - l0;
        c->launches = l0;
        if( rc != PSIFT_OK || ee != cudaSuccess || !g ) {
            if( g ) cudaGraphDestroy( g );
            cudaGetLastError();
            return rc != PSIFT_OK ? rc : fail( PSIFT_E_CUDA, "frame capture: %s", cudaGetErrorString( ee ) );
        }
        const cudaError_t ei = cudaGraphInstantiate( &s.graph, g, 0 );
        cudaGraphDestroy( g );
        if( ei != cudaSuccess ) { s.graph = nullptr; return fail( PSIFT_E_CUDA, "frame graph: %s", cudaGetErrorString( ei ) ); }
        s.graph_epoch = s.epoch;
    }
    if( want_graph ) {
        CU( cudaGraphLaunch( s.graph, s.stream ) );
        c->launches += s.graph_launches;
        const int rcq = queue_result_copies( s );
        if( rcq ) return rcq;
    } else {
        const int rc = enqueue_frame( c, s, false );
        if( rc ) return rc;
        s.frames_direct++;
    }
    CU( cudaEventRecord( s.ev_counters, s.stream ) );
    return PSIFT_OK;
}

/* Wait for a frame's counters.  A frame with more keypoints than the buffers hold: grow them and run it again (the
 * input texture, or the seed plane in the arena, is still in place; a caller-owned device image must stay valid
 * until here).  A clamped keypoint list under-counts the descriptors, hence the loop.  Idempotent. */
static int settle_frame( psift_ctx* c, Slot& s )
{
    CU( cudaEventSynchronize( s.ev_counters ) );
    for( int pass = 0; s.h_ct->overflow && pass < 4; pass++ ) {
        const int ne = std::max( s.ext_cap,  s.h_ct->need_ext  + s.h_ct->need_ext / 8 );
        const int nd = std::max( std::max( s.desc_cap, s.h_ct->need_desc + s.h_ct->need_desc / 4 ), ne + ne / 4 );
        CU( cudaStreamSynchronize( s.stream ) );
        int rc = alloc_keypoint_buffers( s, ne, nd );
        if( rc ) return rc;
        rc = run_frame( c, s, true );
        if( rc ) return rc;
        CU( cudaEventSynchronize( s.ev_counters ) );
    }
    if( s.h_ct->overflow )
        return fail( PSIFT_E_STATE, "keypoint buffers still too small after 4 passes (%d keypoints, %d descriptors needed)", s.h_ct->need_ext, s.h_ct->need_desc );
    return PSIFT_OK;
}

/* ------------------------------------------------------------------------------------------ */
/*                                           C ABI                                            */
/* ------------------------------------------------------------------------------------------ */

namespace psift {
int set_error( int code, const char* msg ) { return fail( code, "%s", msg ); }
__global__ void k_dog_plane( const float* lo, const float* hi, float* out, int w, int h, int pitch )
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if( x < w && y < h ) out[(size_t)y * w + x] = __fsub_rn( hi[(size_t)y * pitch + x], lo[(size_t)y * pitch + x] );
}
cudaError_t launch_dog_plane( const float* lo, const float* hi, float* out, int w, int h, int pitch, cudaStream_t st )
{
    k_dog_plane<<<dim3( ( w + 127 ) / 128, h ), 128, 0, st>>>( lo, hi, out, w, h, pitch );
    return cudaGetLastError();
}
}

extern "C" {

const char* psift_last_error( void ) { return g_err; }

void psift_config_default( psift_config* c )
{
    /* popsift::Config::Config(), sift_conf.cu:17-40 */
    c->octaves = -1; c->levels = 3; c->sigma = 1.6f; c->edge_limit = 10.0f; c->threshold = 0.04f;
    c->upscale_factor = 1.0f;
    c->gauss_mode = PSIFT_GAUSS_VLFEAT_COMPUTE; c->sift_mode = PSIFT_MODE_POPSIFT;
    c->desc_mode = PSIFT_DESC_LOOP; c->norm_mode = PSIFT_NORM_ROOTSIFT; c->norm_multi = 0;
    c->max_extrema = 100000; c->filter_max_extrema = -1; c->filter_grid_size = 2;
    c->filter_sorting = PSIFT_FILTER_RANDOM;
    c->assume_initial_blur = 1; c->initial_blur = 0.5f;
    c->deterministic_order = 1;
    c->scaling_mode = 0;
}

int psift_device_count( void )
{
    int n = 0;
    if( cudaGetDeviceCount( &n ) != cudaSuccess ) { cudaGetLastError(); return 0; }
    return n;
}

static int check_config( const psift_config* cfg )
{
    if( !cfg ) return fail( PSIFT_E_INVALID, "null config" );
    const bool fixed = cfg->gauss_mode == PSIFT_GAUSS_FIXED9 || cfg->gauss_mode == PSIFT_GAUSS_FIXED15;
    if( cfg->gauss_mode != PSIFT_GAUSS_VLFEAT_COMPUTE && cfg->gauss_mode != PSIFT_GAUSS_OPENCV_COMPUTE &&
        cfg->gauss_mode != PSIFT_GAUSS_VLFEAT_RELATIVE_ALL && cfg->gauss_mode != PSIFT_GAUSS_VLFEAT_RELATIVE && !fixed )
        return fail( PSIFT_E_INVALID, "unknown gauss mode %d", cfg->gauss_mode );
    if( cfg->scaling_mode != 0 && cfg->scaling_mode != 1 ) return fail( PSIFT_E_INVALID, "bad scaling mode" );
    if( fixed && cfg->levels != 3 )           /* s_pyramid_fixed.cu:270-289: "Unsupported number of levels for making all octaves at once" */
        return fail( PSIFT_E_INVALID, "the fixed9/fixed15 Gauss modes need levels = 3 (6 Gaussian levels), as in the reference" );
    if( cfg->desc_mode < 0 || cfg->desc_mode > PSIFT_DESC_NOTILE ) return fail( PSIFT_E_INVALID, "bad desc mode" );
    if( cfg->sift_mode < 0 || cfg->sift_mode > PSIFT_MODE_VLFEAT ) return fail( PSIFT_E_INVALID, "bad sift mode" );
    if( cfg->max_extrema < 1 ) return fail( PSIFT_E_INVALID, "max_extrema must be positive" );
    if( cfg->upscale_factor != floorf( cfg->upscale_factor ) || cfg->upscale_factor > 4 || cfg->upscale_factor < -8 )
        return fail( PSIFT_E_INVALID, "downsampling must be an integer in [-4, 8]" );
    if( cfg->filter_grid_size < 1 || cfg->filter_grid_size > 32 ) return fail( PSIFT_E_INVALID, "filter grid size must be in 1..32" );
    return PSIFT_OK;
}

int psift_debug_gauss_table( const psift_config* cfg_in, int which, int level, float* filter32, float* sigma, int* span )
{
    if( !cfg_in ) return fail( PSIFT_E_INVALID, "null config" );
    psift_config cfg = *cfg_in;
    cfg.levels = std::max( 2, cfg.levels );                      /* popsift.cpp:71 */
    GaussTables gt; const char* err = "";
    if( !build_gauss_tables( cfg, gt, &err ) ) return fail( PSIFT_E_INVALID, "%s", err );
    const Taps* t = nullptr; float sg = 0.f;
    const int lim = ( which == 3 ) ? MAX_OCT : MAX_LVL;
    if( level < 0 || level >= lim ) return fail( PSIFT_E_INVALID, "level out of range" );
    switch( which ) {
    case 0: t = &gt.inc[level];    sg = gt.inc_sigma[level]; break;
    case 1: t = &gt.abs_o0[level]; sg = gt.abs_o0_sigma[level]; break;
    case 2: t = &gt.abs_oN[level]; sg = gt.abs_oN_sigma[level]; break;
    case 3: t = &gt.dd[level];     sg = gt.dd_sigma[level]; break;
    default: return fail( PSIFT_E_INVALID, "which must be 0..3" );
    }
    if( filter32 ) memcpy( filter32, t->g, sizeof(float) * TAP_ALIGN );
    if( sigma ) *sigma = sg;
    if( span ) *span = t->span;
    return PSIFT_OK;
}

int psift_debug_desc_tables( const psift_config* cfg_in, float* desc_gauss_1600, float* desc_tile_16, float* scalars2 )
{
    if( !cfg_in ) return fail( PSIFT_E_INVALID, "null config" );
    psift_config cfg = *cfg_in;
    cfg.levels = std::max( 2, cfg.levels );
    Consts K; DescTables* dt = new DescTables;
    build_consts( cfg, K, *dt );
    if( desc_gauss_1600 ) memcpy( desc_gauss_1600, dt->desc_gauss, sizeof dt->desc_gauss );
    if( desc_tile_16 ) memcpy( desc_tile_16, dt->desc_tile, sizeof dt->desc_tile );
    if( scalars2 ) { scalars2[0] = K.sigma_k; scalars2[1] = K.threshold; }
    delete dt;
    return PSIFT_OK;
}

int psift_pyramid_geometry( const psift_config* cfg, int w, int h, int* w0, int* h0, int* n_octaves )
{
    if( !cfg || w < 1 || h < 1 ) return fail( PSIFT_E_INVALID, "bad arguments" );
    int a, b; octave0_size( *cfg, w, h, a, b );
    if( w0 ) *w0 = a;
    if( h0 ) *h0 = b;
    if( n_octaves ) *n_octaves = octave_count( *cfg, w, h );
    return PSIFT_OK;
}

int psift_configure( psift_ctx* c, const psift_config* cfg_in )
{
    if( !c ) return fail( PSIFT_E_INVALID, "null context" );
    int rc = check_config( cfg_in );
    if( rc ) return rc;
    std::lock_guard<std::mutex> lk( c->mu );
    for( auto& s : c->slots ) if( s.busy ) return fail( PSIFT_E_STATE, "configure while jobs are in flight" );
    psift_config cfg = *cfg_in;
    cfg.levels = std::max( 2, cfg.levels );                      /* popsift.cpp:71 */
    GaussTables gt; const char* err = "";
    if( !build_gauss_tables( cfg, gt, &err ) ) return fail( PSIFT_E_INVALID, "%s", err );
    CU( cudaSetDevice( c->device ) );
    const bool realloc_slots = c->slots.empty() || cfg.max_extrema != c->cfg.max_extrema;
    c->cfg = cfg; c->gt = gt;
    build_consts( cfg, c->K, c->dt );
    if( !c->d_dt ) CU( cudaMalloc( &c->d_dt, sizeof(DescTables) ) );
    CU( cudaMemcpy( c->d_dt, &c->dt, sizeof(DescTables), cudaMemcpyHostToDevice ) );
    if( realloc_slots ) {
        for( auto& s : c->slots ) free_slot( s );
        /* frames in flight per context: each slot owns a pyramid arena (265 MB at 1080p), keypoint buffers and streams.
         * Four keep the GPU busy when the inputs are already on the device; with host inputs every frame also spends
         * about 0.3 ms outside the GPU (pinned copy, upload, the waiter's wake-up, result download), which more slots hide */
        int n_slots = 6;
        if( const char* e = getenv( "PSIFT_SLOTS" ) ) n_slots = std::min( 16, std::max( 1, atoi( e ) ) );
        c->slots.assign( (size_t)n_slots, Slot() );
        for( auto& s : c->slots ) { rc = init_slot( c, s ); if( rc ) return rc; }
    } else {
        for( auto& s : c->slots ) free_slot_pyramid( s );      /* geometry may change with the config */
    }
    return PSIFT_OK;
}

int psift_create( int device, const psift_config* cfg, psift_ctx** out )
{
    if( !out ) return fail( PSIFT_E_INVALID, "null out pointer" );
    *out = nullptr;
    int n = psift_device_count();
    if( n <= 0 ) return fail( PSIFT_E_NODEVICE, "no CUDA device: this library has no CPU path" );
    if( device < 0 || device >= n ) return fail( PSIFT_E_INVALID, "device %d out of range (0..%d)", device, n - 1 );
    psift_config def;
    if( !cfg ) { psift_config_default( &def ); cfg = &def; }
    psift_ctx* c = new( std::nothrow ) psift_ctx;
    if( !c ) return fail( PSIFT_E_INVALID, "out of memory" );
    c->device = device;
    if( cudaDeviceGetAttribute( &c->n_sm, cudaDevAttrMultiProcessorCount, device ) != cudaSuccess || c->n_sm < 1 ) { cudaGetLastError(); c->n_sm = 148; }
    int rc = psift_configure( c, cfg );
    if( rc ) { psift_destroy( c ); return rc; }
    *out = c;
    return PSIFT_OK;
}

void psift_destroy( psift_ctx* c )
{
    if( !c ) return;
    cudaSetDevice( c->device );
    for( auto& s : c->slots ) { if( s.stream ) cudaStreamSynchronize( s.stream ); free_slot( s ); }
    if( c->d_dt ) cudaFree( c->d_dt );
    for( auto e : c->mark_join ) cudaEventDestroy( e );
    for( auto e : c->mark_ev ) if( e ) cudaEventDestroy( e );
    if( c->mark_stream ) cudaStreamDestroy( c->mark_stream );
    delete c;
}

/* rows of a host image into the page-locked staging buffer; large images are copied by several threads (one core moves
 * about 5 GB/s, a 16384 x 16384 image is 268 MB) */
/* Copy into the page-locked staging block: read once by the DMA engine and never by this CPU again, so large runs are
 * written with streaming stores (no read-for-ownership of the destination, the caller's working set stays cached). */
static void stage_copy( char* dst, const char* src, size_t n )
{
#if defined( __x86_64__ ) && defined( __SSE2__ )
    if( n >= ( 256u << 10 ) && ( ( (uintptr_t)dst & 15u ) == 0 ) ) {
        size_t i = 0;
        for( ; i + 64 <= n; i += 64 ) {
            const __m128i a = _mm_loadu_si128( (const __m128i*)( src + i ) ),      b = _mm_loadu_si128( (const __m128i*)( src + i + 16 ) );
            const __m128i c = _mm_loadu_si128( (const __m128i*)( src + i + 32 ) ), d = _mm_loadu_si128( (const __m128i*)( src + i + 48 ) );
            _mm_stream_si128( (__m128i*)( dst + i ), a );      _mm_stream_si128( (__m128i*)( dst + i + 16 ), b );
            _mm_stream_si128( (__m128i*)( dst + i + 32 ), c ); _mm_stream_si128( (__m128i*)( dst + i + 48 ), d );
        }
        _mm_sfence();
        if( i < n ) memcpy( dst + i, src + i, n - i );
        return;
    }
#endif
    memcpy( dst, src, n );
}

static void stage_rows( char* dst, size_t dpitch, const char* src, size_t spitch, size_t row_bytes, int nrows )
{
    auto part = [=]( int y0, int y1 ) {
        if( spitch == dpitch && y1 > y0 ) stage_copy( dst + (size_t)y0 * dpitch, src + (size_t)y0 * spitch, spitch * (size_t)( y1 - y0 - 1 ) + row_bytes );
        else for( int y = y0; y < y1; y++ ) memcpy( dst + (size_t)y * dpitch, src + (size_t)y * spitch, row_bytes );
    };
    const size_t bytes = row_bytes * (size_t)nrows;
    int nt = (int)std::min<size_t>( 8, bytes >> 24 );                       /* one thread per 16 MB, at most 8 */
    nt = std::min<int>( nt, std::max( 1u, std::thread::hardware_concurrency() ) );
    if( nt < 2 ) { part( 0, nrows ); return; }
    std::vector<std::thread> th;
    for( int t = 0; t < nt; t++ ) th.emplace_back( part, (int)( (long long)nrows * t / nt ), (int)( (long long)nrows * ( t + 1 ) / nt ) );
    for( auto& t : th ) t.join();
}

static thread_local bool g_pinned_source = false;          /* set by psift_submit_pinned_u8 around submit_common */

/* defer: set the slot up and bring the input in, but queue no kernels: the caller (psift_extract_banded) drives the
 * launches itself and may pass src == nullptr for a seeded pass whose plane it fills in place */
static int submit_common( psift_ctx* c, int w, int h, const void* src, size_t pitch, int elem, bool on_device, psift_job** out,
                          const Region& reg = Region(), bool defer = false )
{
    if( !c || !out || ( !src && !( defer && reg.kind == 2 ) ) ) return fail( PSIFT_E_INVALID, "null argument" );
    if( reg.kind != 0 && ( c->cfg.gauss_mode == PSIFT_GAUSS_FIXED9 || c->cfg.gauss_mode == PSIFT_GAUSS_FIXED15 ) )
        return fail( PSIFT_E_INVALID, "banded extraction is not available for the fixed9/fixed15 Gauss modes" );
    if( reg.kind != 0 && c->cfg.scaling_mode == 1 )
        return fail( PSIFT_E_INVALID, "banded extraction is not available with ScaleDirect scaling (every octave is filtered from the whole input image)" );
    if( reg.kind != 0 && c->cfg.filter_max_extrema > 0 ) return fail( PSIFT_E_INVALID, "the grid filter needs all keypoints of the image: not available for banded extraction" );
    *out = nullptr;
    if( w < 1 || h < 1 ) return fail( PSIFT_E_INVALID, "bad image size %dx%d", w, h );
    if( reg.kind != 2 ) {                 /* (a seeded pass checks its plane's pitch once the octave sizes are known) */
        if( pitch == 0 ) pitch = (size_t)w * elem;
        if( pitch < (size_t)w * elem ) return fail( PSIFT_E_INVALID, "pitch smaller than a row" );
    }
    std::lock_guard<std::mutex> lk( c->mu );
    CU( cudaSetDevice( c->device ) );
    int si = -1;
    /* the lowest free slot: a caller with one frame in flight keeps using one arena, staging buffer and result block
     * (a 16384 x 16384 frame owns 9 GB of them) instead of allocating them afresh on every slot in turn */
    for( size_t k = 0; k < c->slots.size(); k++ )
        if( !c->slots[k].busy ) { si = (int)k; break; }
    if( si < 0 ) return fail( PSIFT_E_STATE, "all %d frame slots are in flight: wait for and free a job first", (int)c->slots.size() );
    Slot& s = c->slots[si];
    int rc = setup_pyramid( c, s, w, h, reg );
    if( rc ) return rc;
    CU( cudaEventRecord( s.ev[0], s.stream ) );
    if( reg.kind == 2 ) {
        /* src: level 0 of octave reg.split, whole plane, device memory, `pitch` bytes per row */
        const PyrLayout& L = s.L;
        const int o = reg.split;
        if( !src ) { /* filled in place by the caller */ } else
        if( pitch < (size_t)L.w[o] * 4 ) return fail( PSIFT_E_INVALID, "seed plane pitch smaller than a row of octave %d (%d pixels)", o, L.w[o] );
        if( src )
        CU( cudaMemcpy2DAsync( L.base[o], (size_t)L.pitch[o] * 4, src, pitch, (size_t)L.w[o] * 4, L.h[o], cudaMemcpyDeviceToDevice, s.stream ) );
    } else if( on_device ) {
        if( pitch % 32 || ( (uintptr_t)src % 512 ) )
            return fail( PSIFT_E_INVALID, "device image needs a 512-byte aligned base and a pitch that is a multiple of 32" );
        if( graphs_enabled() && reg.kind == 0 ) {
            /* the frame graph has the input texture baked in: keep the texture on the slot's own image buffer and
             * bring the caller's image there (2 MB device-to-device at 1080p, about a microsecond) */
            const size_t row = (size_t)w * elem;
            if( !s.d_img || s.img_cap_w < row || s.img_cap_h < (size_t)h ) {
                if( s.img_tex ) { cudaDestroyTextureObject( s.img_tex ); s.img_tex = 0; }
                if( s.d_img ) cudaFree( s.d_img );
                s.img_cap_w = std::max( s.img_cap_w, row ); s.img_cap_h = std::max( s.img_cap_h, (size_t)h );
                CU( cudaMallocPitch( &s.d_img, &s.img_pitch, s.img_cap_w, s.img_cap_h ) );
            }
            CU( cudaMemcpy2DAsync( s.d_img, s.img_pitch, src, pitch, row, h, cudaMemcpyDeviceToDevice, s.stream ) );
            rc = bind_input_texture( s, s.d_img, s.img_pitch, w, h, elem );
        } else
            rc = bind_input_texture( s, src, pitch, w, h, elem );
        if( rc ) return rc;
    } else {
        /* private pinned copy (SiftJob::SiftJob, popsift.cpp:238-253; Image::load, s_image.cu:71-79) */
        const size_t row = (size_t)w * elem;
        if( !s.d_img || s.img_cap_w < row || s.img_cap_h < (size_t)h ) {
            if( s.img_tex ) { cudaDestroyTextureObject( s.img_tex ); s.img_tex = 0; }
            if( s.d_img ) cudaFree( s.d_img );
            s.img_cap_w = std::max( s.img_cap_w, row ); s.img_cap_h = std::max( s.img_cap_h, (size_t)h );
            CU( cudaMallocPitch( &s.d_img, &s.img_pitch, s.img_cap_w, s.img_cap_h ) );
        }
        /* a band samples only some rows of the input: the bilinear footprint of its stored rows of octave 0 (+ a margin) */
        int ra = 0, rb = h;
        if( reg.kind == 1 ) {
            float shift = 0.5f;
            if( c->cfg.sift_mode == PSIFT_MODE_POPSIFT || c->cfg.sift_mode == PSIFT_MODE_VLFEAT ) shift = 0.5f * powf( 2.0f, c->cfg.upscale_factor );
            const double sc = (double)h / (double)s.L.h[0];
            ra = std::max( 0, (int)floor( ( s.L.row_lo[0] + shift ) * sc - 0.5 ) - 2 );
            rb = std::min( h, (int)floor( ( s.L.row_hi[0] - 1 + shift ) * sc - 0.5 ) + 4 );
        }
        const char* src_rows = (const char*)src + (size_t)ra * pitch;
        char*       dev_rows = (char*)s.d_img + (size_t)ra * s.img_pitch;
        const int   nrows = rb - ra;
        /* the pinned copy has the DEVICE pitch, so that the upload is one linear transfer and not a row-by-row 2-D copy */
        const size_t need = s.img_pitch * nrows;
        if( !g_pinned_source && need > s.h_img_cap ) {
            if( s.h_img ) cudaFreeHost( s.h_img );
            CU( cudaMallocHost( &s.h_img, need ) );
            s.h_img_cap = need;
        }
        if( g_pinned_source ) {
            /* psift_submit_pinned_u8: the caller's buffer is page-locked and stays put until psift_wait: no private copy */
            CU( cudaMemcpy2DAsync( dev_rows, s.img_pitch, src_rows, pitch, row, nrows, cudaMemcpyHostToDevice, s.stream ) );
        } else {
            stage_rows( (char*)s.h_img, s.img_pitch, src_rows, pitch, row, nrows );
            CU( cudaMemcpyAsync( dev_rows, s.h_img, need, cudaMemcpyHostToDevice, s.stream ) );
        }
        rc = bind_input_texture( s, s.d_img, s.img_pitch, w, h, elem );
        if( rc ) return rc;
    }
    if( !defer ) {
        rc = run_frame( c, s );
        if( rc ) return rc;
    }
    s.busy = true;
    psift_job* j = new psift_job;
    j->ctx = c; j->slot = si; j->waited = false; j->status = PSIFT_OK;
    *out = j;
    return PSIFT_OK;
}

int psift_submit_u8( psift_ctx* c, int w, int h, const uint8_t* host, size_t pitch, psift_job** out )
{ return submit_common( c, w, h, host, pitch, 1, false, out ); }
int psift_submit_f32( psift_ctx* c, int w, int h, const float* host, size_t pitch, psift_job** out )
{ return submit_common( c, w, h, host, pitch, 4, false, out ); }
int psift_host_alloc_pinned( void** out, size_t bytes )
{
    if( !out ) return fail( PSIFT_E_INVALID, "null argument" );
    *out = nullptr;
    int n = 0;
    if( cudaGetDeviceCount( &n ) != cudaSuccess || n <= 0 ) { cudaGetLastError(); return fail( PSIFT_E_CUDA, "no CUDA device" ); }
    const cudaError_t e = cudaHostAlloc( out, bytes ? bytes : 1, cudaHostAllocPortable );
    if( e != cudaSuccess ) { cudaGetLastError(); *out = nullptr; return fail( PSIFT_E_CUDA, "cudaHostAlloc: %s", cudaGetErrorString( e ) ); }
    return PSIFT_OK;
}
void psift_host_free_pinned( void* p ) { if( p ) cudaFreeHost( p ); }

int psift_submit_pinned_u8( psift_ctx* c, int w, int h, const uint8_t* pinned_host, size_t pitch, psift_job** out )
{
    g_pinned_source = true;
    const int rc = submit_common( c, w, h, pinned_host, pitch, 1, false, out );
    g_pinned_source = false;
    return rc;
}
int psift_submit_pinned_f32( psift_ctx* c, int w, int h, const float* pinned_host, size_t pitch, psift_job** out )
{
    g_pinned_source = true;
    const int rc = submit_common( c, w, h, pinned_host, pitch, 4, false, out );
    g_pinned_source = false;
    return rc;
}
int psift_submit_dev_u8( psift_ctx* c, int w, int h, const uint8_t* dev, size_t pitch, psift_job** out )
{ return submit_common( c, w, h, dev, pitch, 1, true, out ); }

int psift_band_plan( const psift_config* cfg, int w, int h, int parts, int max_seed_pixels, int* split_octave, int* bounds )
{
    if( !cfg || parts < 1 || !split_octave || !bounds ) return fail( PSIFT_E_INVALID, "null argument" );
    int w0, h0;
    octave0_size( *cfg, w, h, w0, h0 );
    const int n_oct = octave_count( *cfg, w, h );
    if( n_oct < 2 ) return fail( PSIFT_E_INVALID, "an image with one octave cannot be split" );
    if( max_seed_pixels <= 0 ) max_seed_pixels = 2048 * 2048;
    /* the octaves above the split are finished un-split: pick the first split whose seed is small */
    int k = 0, ow = w0, oh = h0;
    for( ;; ) {
        const int nw = ( ow + 1 ) / 2, nh = ( oh + 1 ) / 2;
        if( (long long)nw * nh <= max_seed_pixels || k + 2 >= n_oct ) break;
        ow = nw; oh = nh; k++;
    }
    *split_octave = k;
    const int al = 1 << ( k + 1 );
    for( int p = 0; p <= parts; p++ ) {
        long long y = (long long)h0 * p / parts;
        y = ( y + al / 2 ) / al * al;
        bounds[p] = (int)std::min<long long>( y, h0 );
    }
    bounds[0] = 0; bounds[parts] = h0;
    for( int p = 0; p < parts; p++ ) if( bounds[p + 1] <= bounds[p] ) return fail( PSIFT_E_INVALID, "image too small for %d bands", parts );
    return PSIFT_OK;
}

int psift_submit_band_u8( psift_ctx* c, int w, int h, const uint8_t* img, size_t pitch, int on_device,
                          int y0, int y1, int split_octave, psift_job** out )
{
    Region r; r.kind = 1; r.y0 = y0; r.y1 = y1; r.split = split_octave;
    if( split_octave >= 0 && ( ( y0 & ( ( 1 << ( split_octave + 1 ) ) - 1 ) ) != 0 ) )
        return fail( PSIFT_E_INVALID, "band start %d must be a multiple of 2^(split+1)", y0 );
    return submit_common( c, w, h, img, pitch, 1, on_device != 0, out, r );
}

int psift_band_seed( psift_job* j, const float** dev_rows, int* pitch_floats, int* row_lo, int* row_hi, int* seed_w, int* seed_h )
{
    if( !j || !dev_rows ) return fail( PSIFT_E_INVALID, "null argument" );
    const Slot& s = j->ctx->slots[j->slot];
    const PyrLayout& L = s.L;
    if( s.reg_kind != 1 || L.oct_end >= L.alloc_end ) return fail( PSIFT_E_STATE, "not a band job, or no octave above the split" );
    const int o = L.oct_end;
    *dev_rows = L.base[o] + (size_t)L.row_lo[o] * L.pitch[o];
    if( pitch_floats ) *pitch_floats = L.pitch[o];
    if( row_lo ) *row_lo = L.row_lo[o];
    if( row_hi ) *row_hi = L.row_hi[o];
    if( seed_w ) *seed_w = L.w[o];
    if( seed_h ) *seed_h = L.h[o];
    return PSIFT_OK;
}

int psift_band_seed_copy( psift_job* j, float* dst_dev, size_t dst_pitch_bytes, int row_from, int row_to )
{
    if( !j || !dst_dev ) return fail( PSIFT_E_INVALID, "null argument" );
    const Slot& s = j->ctx->slots[j->slot];
    const PyrLayout& L = s.L;
    if( s.reg_kind != 1 || L.oct_end >= L.alloc_end ) return fail( PSIFT_E_STATE, "not a band job, or no octave above the split" );
    const int o = L.oct_end;
    if( row_from < L.row_lo[o] || row_to > L.row_hi[o] || row_to <= row_from ) return fail( PSIFT_E_INVALID, "rows [%d,%d) outside the band's seed rows [%d,%d)", row_from, row_to, L.row_lo[o], L.row_hi[o] );
    CU( cudaSetDevice( j->ctx->device ) );
    CU( cudaMemcpy2DAsync( dst_dev, dst_pitch_bytes, L.base[o] + (size_t)row_from * L.pitch[o], (size_t)L.pitch[o] * 4,
                           (size_t)L.w[o] * 4, row_to - row_from, cudaMemcpyDefault, s.stream ) );   /* dst may live on another GPU */
    CU( cudaStreamSynchronize( s.stream ) );
    return PSIFT_OK;
}

int psift_submit_seed_dev_f32( psift_ctx* c, int w, int h, int first_octave, const float* dev_plane, size_t pitch_bytes,
                               int y0, int y1, psift_job** out )
{
    Region r; r.kind = 2; r.y0 = y0; r.y1 = y1; r.split = first_octave;
    return submit_common( c, w, h, dev_plane, pitch_bytes, 4, true, out, r );
}

int psift_wait( psift_job* j, psift_result* r )
{
    if( !j || !r ) return fail( PSIFT_E_INVALID, "null argument" );
    psift_ctx* c = j->ctx;
    Slot& s = c->slots[j->slot];
    CU( cudaSetDevice( c->device ) );
    if( !j->waited ) {
        int rc0 = settle_frame( c, s );
        if( rc0 ) return rc0;
        const int n_ext = s.h_ct->ext_total, n_desc = s.h_ct->ori_total;
        if( n_ext > s.run_pre_ext || n_desc > s.run_pre_desc ) {
            /* more results than the frame downloaded by itself: fetch them all (the host buffers may have to grow) */
            int rc = ensure_host_results( s, n_ext, n_desc );
            if( rc ) return rc;
            if( n_ext > 0 ) {
                CU( cudaMemcpyAsync( s.h_feat, s.feat, sizeof(psift_feature) * (size_t)n_ext, cudaMemcpyDeviceToHost, s.stream ) );
                CU( cudaMemcpyAsync( s.h_ext,  s.ext,  sizeof(psift_extremum) * (size_t)n_ext, cudaMemcpyDeviceToHost, s.stream ) );
            }
            if( n_desc > 0 )
                CU( cudaMemcpyAsync( s.h_desc, s.desc, sizeof(float) * 128 * (size_t)n_desc, cudaMemcpyDeviceToHost, s.stream ) );
            c->d2h_bytes += (long long)n_ext * ( sizeof(psift_feature) + sizeof(psift_extremum) ) + (long long)n_desc * 512;
        }
        CU( cudaEventRecord( s.ev[6], s.stream ) );
        CU( cudaStreamSynchronize( s.stream ) );
        j->waited = true;
        /* size the next frames' own download from what this one produced */
        const int want_ext  = std::min( s.ext_cap,  ( n_ext  + n_ext  / 4 + 1023 ) / 1024 * 1024 );
        const int want_desc = std::min( s.desc_cap, ( n_desc + n_desc / 4 + 1023 ) / 1024 * 1024 );
        if( want_ext > s.pre_ext || want_desc > s.pre_desc ) {
            s.pre_ext = std::max( s.pre_ext, want_ext ); s.pre_desc = std::max( s.pre_desc, want_desc );
            s.epoch++;
        }
        std::lock_guard<std::mutex> lk( c->mu );
        float ms = 0.f;
        for( int k = 0; k < 6; k++ ) { if( cudaEventElapsedTime( &ms, s.ev[k], s.ev[k + 1] ) == cudaSuccess ) c->last_ms[k] = ms; }
        if( cudaEventElapsedTime( &ms, s.ev[0], s.ev[6] ) == cudaSuccess ) c->last_ms[6] = ms;
        c->last_lvl_n = 0;
        if( s.ps.lvl_ev ) {
            for( int l = 0; l < s.L.n_lvl; l++ )
                c->last_lvl_ms[l] = cudaEventElapsedTime( &ms, s.lvl_ev[2 * l], s.lvl_ev[2 * l + 1] ) == cudaSuccess ? ms : -1.0f;
            c->last_lvl_n = s.L.n_lvl;
        }
        c->last_slot = j->slot;
    }
    memset( r, 0, sizeof *r );
    if( !s.h_ext && ( s.h_ct->ext_total > 0 || s.h_ct->ori_total > 0 ) )
        return fail( PSIFT_E_STATE, "the results of this job were taken (psift_wait_take)" );
    r->n_feat = s.h_ct->ext_total; r->n_desc = s.h_ct->ori_total;
    r->feat = s.h_feat; r->desc = s.h_desc; r->ext = s.h_ext;
    for( int o = 0; o < MAX_OCT; o++ ) { r->ext_ct[o] = s.h_ct->ext_ct[o]; r->ori_ct[o] = s.h_ct->ori_ct[o]; }
    r->n_octaves = s.L.n_oct; r->n_levels = s.L.n_lvl; r->w0 = s.L.w[0]; r->h0 = s.L.h[0];
    return PSIFT_OK;
}

int psift_wait_take( psift_job* j, psift_result* r, psift_result_block** block )
{
    if( !block ) return fail( PSIFT_E_INVALID, "null argument" );
    *block = nullptr;
    const int rc = psift_wait( j, r );
    if( rc ) return rc;
    Slot& s = j->ctx->slots[j->slot];
    /* r's pointers now belong to the block; the slot takes another one from the pool for its next frame */
    *block = detach_block( s );
    return PSIFT_OK;
}

void psift_result_block_free( psift_result_block* b ) { result_pool().put( b ); }

void psift_job_free( psift_job* j )
{
    if( !j ) return;
    psift_ctx* c = j->ctx;
    Slot& s = c->slots[j->slot];
    if( !j->waited ) { cudaSetDevice( c->device ); cudaStreamSynchronize( s.stream ); }
    { std::lock_guard<std::mutex> lk( c->mu ); s.busy = false; }
    delete j;
}

int psift_octave_size( psift_ctx* c, int octave, int* w, int* h )
{
    if( !c || c->last_slot < 0 ) return fail( PSIFT_E_STATE, "no finished job" );
    const Slot& s = c->slots[c->last_slot];
    if( octave < 0 || octave >= s.L.n_oct ) return fail( PSIFT_E_INVALID, "octave out of range" );
    if( w ) *w = s.L.w[octave];
    if( h ) *h = s.L.h[octave];
    return PSIFT_OK;
}

int psift_debug_read_level( psift_ctx* c, int octave, int level, float* out )
{
    if( !c || c->last_slot < 0 || !out ) return fail( PSIFT_E_STATE, "no finished job" );
    const Slot& s = c->slots[c->last_slot];
    if( octave < 0 || octave >= s.L.n_oct || level < 0 || level >= s.L.n_lvl ) return fail( PSIFT_E_INVALID, "octave/level out of range" );
    CU( cudaSetDevice( c->device ) );
    const float* p = s.L.base[octave] + (size_t)level * s.L.lstride[octave];
    CU( cudaMemcpy2D( out, (size_t)s.L.w[octave] * 4, p, (size_t)s.L.pitch[octave] * 4, (size_t)s.L.w[octave] * 4, s.L.h[octave], cudaMemcpyDeviceToHost ) );
    return PSIFT_OK;
}

int psift_debug_read_dog( psift_ctx* c, int octave, int level, float* out )
{
    if( !c || c->last_slot < 0 || !out ) return fail( PSIFT_E_STATE, "no finished job" );
    const Slot& s = c->slots[c->last_slot];
    if( octave < 0 || octave >= s.L.n_oct || level < 0 || level >= s.L.n_lvl - 1 ) return fail( PSIFT_E_INVALID, "octave/level out of range" );
    CU( cudaSetDevice( c->device ) );
    const int w = s.L.w[octave], h = s.L.h[octave];
    float* tmp = nullptr;
    CU( cudaMalloc( &tmp, sizeof(float) * (size_t)w * h ) );
    const float* lo = s.L.base[octave] + (size_t)level * s.L.lstride[octave];
    cudaError_t e = launch_dog_plane( lo, lo + s.L.lstride[octave], tmp, w, h, s.L.pitch[octave], 0 );
    if( e == cudaSuccess ) e = cudaMemcpy( out, tmp, sizeof(float) * (size_t)w * h, cudaMemcpyDeviceToHost );
    cudaFree( tmp );
    if( e != cudaSuccess ) return fail( PSIFT_E_CUDA, "dog readback: %s", cudaGetErrorString( e ) );
    return PSIFT_OK;
}

int psift_last_timing( psift_ctx* c, float ms[8] )
{
    if( !c || !ms ) return fail( PSIFT_E_INVALID, "null argument" );
    std::lock_guard<std::mutex> lk( c->mu );
    memcpy( ms, c->last_ms, sizeof c->last_ms );
    return PSIFT_OK;
}

int psift_debug_time_levels( psift_ctx* c, int on )
{
    if( !c ) return fail( PSIFT_E_INVALID, "null argument" );
    std::lock_guard<std::mutex> lk( c->mu );
    c->time_levels = on != 0;
    return PSIFT_OK;
}

int psift_last_level_timing( psift_ctx* c, float* ms, int cap, int* n )
{
    if( !c || !ms || !n ) return fail( PSIFT_E_INVALID, "null argument" );
    std::lock_guard<std::mutex> lk( c->mu );
    *n = std::min( cap, c->last_lvl_n );
    for( int l = 0; l < *n; l++ ) ms[l] = c->last_lvl_ms[l];
    return PSIFT_OK;
}

int psift_slot_count( psift_ctx* c ) { return c ? (int)c->slots.size() : 0; }
long long psift_launch_count( psift_ctx* c ) { return c ? c->launches : 0; }
long long psift_d2h_bytes( psift_ctx* c ) { return c ? c->d2h_bytes : 0; }

int psift_debug_mark( psift_ctx* c, int which )
{
    if( !c || which < 0 || which > 1 ) return fail( PSIFT_E_INVALID, "bad mark" );
    std::lock_guard<std::mutex> lk( c->mu );
    CU( cudaSetDevice( c->device ) );
    if( !c->mark_stream ) {
        CU( cudaStreamCreateWithFlags( &c->mark_stream, cudaStreamNonBlocking ) );
        for( auto& e : c->mark_ev ) CU( cudaEventCreate( &e ) );
        c->mark_join.resize( c->slots.size(), nullptr );
        for( auto& e : c->mark_join ) CU( cudaEventCreateWithFlags( &e, cudaEventDisableTiming ) );
    }
    for( size_t i = 0; i < c->slots.size(); i++ ) {
        CU( cudaEventRecord( c->mark_join[i], c->slots[i].stream ) );
        CU( cudaStreamWaitEvent( c->mark_stream, c->mark_join[i], 0 ) );
    }
    CU( cudaEventRecord( c->mark_ev[which], c->mark_stream ) );
    return PSIFT_OK;
}

int psift_debug_mark_elapsed( psift_ctx* c, float* ms )
{
    if( !c || !ms || !c->mark_stream ) return fail( PSIFT_E_INVALID, "no marks set" );
    CU( cudaEventSynchronize( c->mark_ev[1] ) );
    CU( cudaEventElapsedTime( ms, c->mark_ev[0], c->mark_ev[1] ) );
    return PSIFT_OK;
}

/* ---------------------------------------------------------------------------------------------
 * Device-resident results and matching (reference Config::MatchingMode: popsift.cpp:215-236,
 * Pyramid::clone_device_descriptors sift_pyramid.cu:323-361, FeaturesDev features.cu:126-302)
 * ------------------------------------------------------------------------------------------- */
} // extern "C"  (the struct below is C++)

struct psift_dev_features {
    int device = 0, n_feat = 0, n_desc = 0;
    psift_feature* feat = nullptr;      /* n_feat records, descriptor INDICES */
    void*  feat_ref = nullptr;          /* n_feat records in the reference's 72-byte layout, device descriptor POINTERS */
    float* desc = nullptr;              /* n_desc x 128 */
    int*   rev = nullptr;               /* descriptor -> feature (feat_to_ext_map) */
};

extern "C" {

int psift_dev_features_alloc( int device, int n_feat, int n_desc, psift_dev_features** out )
{
    if( !out || n_feat < 0 || n_desc < 0 ) return fail( PSIFT_E_INVALID, "bad argument" );
    *out = nullptr;
    CU( cudaSetDevice( device ) );
    psift_dev_features* f = new psift_dev_features;
    f->device = device; f->n_feat = n_feat; f->n_desc = n_desc;
    cudaError_t e = cudaMalloc( &f->feat, sizeof(psift_feature) * (size_t)std::max( n_feat, 1 ) );
    if( e == cudaSuccess ) e = cudaMalloc( &f->feat_ref, (size_t)72 * std::max( n_feat, 1 ) );
    if( e == cudaSuccess ) e = cudaMalloc( &f->desc, sizeof(float) * 128 * (size_t)std::max( n_desc, 1 ) );
    if( e == cudaSuccess ) e = cudaMalloc( &f->rev, sizeof(int) * (size_t)std::max( n_desc, 1 ) );
    if( e != cudaSuccess ) { psift_dev_features_free( f ); return fail( PSIFT_E_CUDA, "device features: %s", cudaGetErrorString( e ) ); }
    *out = f;
    return PSIFT_OK;
}

void psift_dev_features_free( psift_dev_features* f )
{
    if( !f ) return;
    cudaSetDevice( f->device );
    cudaFree( f->feat ); cudaFree( f->feat_ref ); cudaFree( f->desc ); cudaFree( f->rev );
    delete f;
}

int psift_dev_features_info( const psift_dev_features* f, psift_dev_info* info )
{
    if( !f || !info ) return fail( PSIFT_E_INVALID, "null argument" );
    info->device = f->device; info->n_feat = f->n_feat; info->n_desc = f->n_desc;
    info->feat = f->feat; info->feat_ref = f->feat_ref; info->desc = f->desc; info->rev = f->rev;
    return PSIFT_OK;
}

int psift_wait_dev( psift_job* j, psift_dev_features** out )
{
    if( !j || !out ) return fail( PSIFT_E_INVALID, "null argument" );
    *out = nullptr;
    psift_ctx* c = j->ctx;
    Slot& s = c->slots[j->slot];
    CU( cudaSetDevice( c->device ) );
    int rc = settle_frame( c, s );
    if( rc ) return rc;
    const int n_ext = s.h_ct->ext_total, n_desc = s.h_ct->ori_total;
    psift_dev_features* f = nullptr;
    rc = psift_dev_features_alloc( c->device, n_ext, n_desc, &f );
    if( rc ) return rc;
    cudaError_t e = cudaSuccess;
    if( n_ext > 0 ) e = cudaMemcpyAsync( f->feat, s.feat, sizeof(psift_feature) * (size_t)n_ext, cudaMemcpyDeviceToDevice, s.stream );
    if( e == cudaSuccess && n_desc > 0 ) e = cudaMemcpyAsync( f->desc, s.desc, sizeof(float) * 128 * (size_t)n_desc, cudaMemcpyDeviceToDevice, s.stream );
    if( e == cudaSuccess && n_desc > 0 ) e = cudaMemcpyAsync( f->rev, s.desc_map, sizeof(int) * (size_t)n_desc, cudaMemcpyDeviceToDevice, s.stream );
    if( e == cudaSuccess ) e = launch_ref_features( f->feat, n_ext, f->desc, f->feat_ref, s.stream );
    if( e == cudaSuccess ) e = cudaStreamSynchronize( s.stream );
    if( e != cudaSuccess ) { psift_dev_features_free( f ); return fail( PSIFT_E_CUDA, "clone to device features: %s", cudaGetErrorString( e ) ); }
    *out = f;
    return PSIFT_OK;
}

int psift_match_dev( int device, const float* l_desc, int l_len, const float* r_desc, int r_len, int* match3, float* dist2 )
{
    if( l_len < 0 || r_len < 0 || ( l_len > 0 && ( !l_desc || !match3 ) ) || ( r_len > 0 && !r_desc ) ) return fail( PSIFT_E_INVALID, "bad argument" );
    if( l_len == 0 ) return PSIFT_OK;
    CU( cudaSetDevice( device ) );
    int* d_m = nullptr; float* d_d = nullptr;
    CU( cudaMalloc( &d_m, sizeof(int) * 3 * (size_t)l_len ) );
    cudaError_t e = dist2 ? cudaMalloc( &d_d, sizeof(float) * 2 * (size_t)l_len ) : cudaSuccess;
    if( e == cudaSuccess ) e = launch_match( l_desc, l_len, r_desc, r_len, d_m, d_d, 0 );
    if( e == cudaSuccess ) e = cudaMemcpy( match3, d_m, sizeof(int) * 3 * (size_t)l_len, cudaMemcpyDeviceToHost );
    if( e == cudaSuccess && dist2 ) e = cudaMemcpy( dist2, d_d, sizeof(float) * 2 * (size_t)l_len, cudaMemcpyDeviceToHost );
    cudaFree( d_m ); cudaFree( d_d );
    if( e != cudaSuccess ) return fail( PSIFT_E_CUDA, "match: %s", cudaGetErrorString( e ) );
    return PSIFT_OK;
}

int psift_match( const psift_dev_features* l, const psift_dev_features* r, int* match3, float* dist2 )
{
    if( !l || !r ) return fail( PSIFT_E_INVALID, "null argument" );
    if( l->device == r->device ) return psift_match_dev( l->device, l->desc, l->n_desc, r->desc, r->n_desc, match3, dist2 );
    /* the right-hand descriptors live on another GPU: bring them over (NVLink peer copy) */
    CU( cudaSetDevice( l->device ) );
    float* tmp = nullptr;
    CU( cudaMalloc( &tmp, sizeof(float) * 128 * (size_t)std::max( r->n_desc, 1 ) ) );
    cudaError_t e = cudaMemcpyPeer( tmp, l->device, r->desc, r->device, sizeof(float) * 128 * (size_t)r->n_desc );
    int rc = e == cudaSuccess ? psift_match_dev( l->device, l->desc, l->n_desc, tmp, r->n_desc, match3, dist2 )
                              : fail( PSIFT_E_CUDA, "peer copy: %s", cudaGetErrorString( e ) );
    cudaFree( tmp );
    return rc;
}

int psift_dev_download( int device, const void* dev, void* host, size_t bytes )
{
    if( bytes == 0 ) return PSIFT_OK;
    if( !dev || !host ) return fail( PSIFT_E_INVALID, "null argument" );
    CU( cudaSetDevice( device ) );
    CU( cudaMemcpy( host, dev, bytes, cudaMemcpyDeviceToHost ) );
    return PSIFT_OK;
}

int psift_dev_upload( int device, void* dev, const void* host, size_t bytes )
{
    if( bytes == 0 ) return PSIFT_OK;
    if( !dev || !host ) return fail( PSIFT_E_INVALID, "null argument" );
    CU( cudaSetDevice( device ) );
    CU( cudaMemcpy( dev, host, bytes, cudaMemcpyHostToDevice ) );
    return PSIFT_OK;
}

/* One giant image over the given contexts (one per GPU; several on one GPU work too), start to finish.
 *
 *   1. every context gets the rows of the u8 image that its band samples (staged and uploaded by one thread per context);
 *   2. octaves 0..split are built LEVEL BY LEVEL on all contexts together: a context computes its own rows of a level,
 *      then pulls the `halo` rows above and below from its two neighbours (device-to-device / NVLink peer copies on its
 *      own stream, behind an event of the neighbour's stream) - SURVEY 8e's per-level halo exchange.  The halo is as wide
 *      as the orientation / descriptor windows, so the later stages find every row they read;
 *   3. the band's rows of the seed plane (level 0 of octave split+1, written by the fused downscale) go straight into
 *      every context's second slot, which finishes the small octaves un-split, keeping the keypoints it owns;
 *   4. DoG / extrema / orientation / descriptors per band; results merged.
 * Every pixel is computed by the same arithmetic from the same inputs as in the un-split run, so the union of the bands'
 * keypoints is that run's result bit for bit.  PSIFT_BAND_EXCHANGE=0 (or a band thinner than the halo, or a Gauss mode
 * that cannot be driven level by level) falls back to bands that recompute their halo (band_halo_rows). */
namespace {
/* rows [a, b) of level l of octave o: from the neighbour's arena into mine, on my stream */
static cudaError_t pull_rows( const PyrLayout& mine, const PyrLayout& theirs, int o, int l, int a, int b, cudaStream_t st )
{
    if( b <= a ) return cudaSuccess;
    const size_t off_m = (size_t)l * mine.lstride[o]   + (size_t)a * mine.pitch[o];
    const size_t off_t = (size_t)l * theirs.lstride[o] + (size_t)a * theirs.pitch[o];
    return cudaMemcpyAsync( mine.base[o] + off_m, theirs.base[o] + off_t, sizeof(float) * (size_t)( b - a ) * mine.pitch[o], cudaMemcpyDefault, st );
}
}

int psift_extract_banded( psift_ctx** ctxs, int n_ctx, int w, int h, const uint8_t* host_img, size_t pitch,
                          int max_seed_pixels, psift_banded_result* out )
{
    if( !ctxs || n_ctx < 1 || !host_img || !out ) return fail( PSIFT_E_INVALID, "null argument" );
    memset( out, 0, sizeof *out );
    for( int r = 0; r < n_ctx; r++ ) if( !ctxs[r] ) return fail( PSIFT_E_INVALID, "null context" );
    const psift_config cfg = ctxs[0]->cfg;
    const int n_oct = octave_count( cfg, w, h );
    int split = 0;
    std::vector<int> bounds( n_ctx + 1, 0 );
    int rc = psift_band_plan( &cfg, w, h, n_ctx, max_seed_pixels, &split, bounds.data() );
    if( rc ) return rc;
    out->split_octave = split;
    const bool have_seed = split + 1 < n_oct;

    /* halo exchange needs a Gauss mode that is built level by level, and bands at least as thick as the halo */
    bool exchange = true;
    if( const char* e = getenv( "PSIFT_BAND_EXCHANGE" ) ) exchange = atoi( e ) != 0;
    if( cfg.gauss_mode != PSIFT_GAUSS_VLFEAT_COMPUTE && cfg.gauss_mode != PSIFT_GAUSS_OPENCV_COMPUTE && cfg.gauss_mode != PSIFT_GAUSS_VLFEAT_RELATIVE_ALL ) exchange = false;
    const int E = band_exchange_rows( ctxs[0] );
    for( int r = 0; r < n_ctx && n_ctx > 1; r++ ) if( ( ( bounds[r + 1] - bounds[r] ) >> split ) < E ) exchange = false;
    out->halo_exchange = exchange ? 1 : 0;

    static const bool trace = getenv( "PSIFT_BAND_TRACE" ) != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    auto lap = [&]( const char* what ) {
        if( trace ) fprintf( stderr, "[banded] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now() - t_begin ).count() );
    };
    std::vector<psift_job*> jobs( n_ctx, nullptr ), seeded( n_ctx, nullptr );
    std::vector<cudaEvent_t> ev( n_ctx, nullptr ), ev_seed( n_ctx, nullptr );
    auto cleanup = [&]() {
        for( int r = 0; r < n_ctx; r++ ) {
            if( jobs[r] ) { psift_job_free( jobs[r] ); jobs[r] = nullptr; }
            if( seeded[r] ) { psift_job_free( seeded[r] ); seeded[r] = nullptr; }
            if( ev[r] || ev_seed[r] ) cudaSetDevice( ctxs[r]->device );
            if( ev[r] ) { cudaEventDestroy( ev[r] ); ev[r] = nullptr; }
            if( ev_seed[r] ) { cudaEventDestroy( ev_seed[r] ); ev_seed[r] = nullptr; }
        }
    };
    for( int r = 0; r < n_ctx; r++ )
        for( int q = 0; q < n_ctx; q++ )                                /* peer access where the hardware has it; else copies are staged */
            if( ctxs[q]->device != ctxs[r]->device ) { cudaSetDevice( ctxs[r]->device ); cudaDeviceEnablePeerAccess( ctxs[q]->device, 0 ); cudaGetLastError(); }

    lap( "plan, peer access" );
    /* 1. upload: one thread per context stages its band's rows into page-locked memory and queues the copy */
    {
        std::vector<int> rcs( n_ctx, PSIFT_OK );
        std::vector<std::string> msgs( n_ctx );
        std::vector<std::thread> th;
        for( int r = 0; r < n_ctx; r++ )
            th.emplace_back( [&, r]() {
                Region reg; reg.kind = 1; reg.y0 = bounds[r]; reg.y1 = bounds[r + 1]; reg.split = split; reg.exchange = exchange ? 1 : 0;
                rcs[r] = submit_common( ctxs[r], w, h, host_img, pitch, 1, false, &jobs[r], reg, exchange );
                if( rcs[r] != PSIFT_OK ) msgs[r] = g_err;
            } );
        for( auto& t : th ) t.join();
        for( int r = 0; r < n_ctx; r++ ) if( rcs[r] != PSIFT_OK ) { cleanup(); return fail( rcs[r], "band %d: %s", r, msgs[r].c_str() ); }
    }
    lap( "band slots, staging, upload" );
    auto slot_of = [&]( psift_job* j ) -> Slot& { return j->ctx->slots[j->slot]; };
#define BX( call ) do { cudaError_t e_ = ( call ); if( e_ != cudaSuccess ) { cleanup(); \
    return fail( PSIFT_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString( e_ ), __FILE__, __LINE__ ); } } while( 0 )
    for( int r = 0; r < n_ctx; r++ ) {
        BX( cudaSetDevice( ctxs[r]->device ) );
        BX( cudaEventCreateWithFlags( &ev[r], cudaEventDisableTiming ) );
        BX( cudaEventCreateWithFlags( &ev_seed[r], cudaEventDisableTiming ) );
    }

    /* 3a. the second slot of every context: the octaves above the split, from a seed plane that is filled in place */
    if( have_seed )
        for( int r = 0; r < n_ctx; r++ ) {
            Region reg; reg.kind = 2; reg.y0 = bounds[r]; reg.y1 = bounds[r + 1]; reg.split = split + 1;
            if( ( rc = submit_common( ctxs[r], w, h, nullptr, 0, 4, true, &seeded[r], reg, true ) ) != PSIFT_OK ) { cleanup(); return rc; }
        }

    lap( "events, seeded slots" );
    /* 2. octaves 0..split, level by level, halos pulled from the neighbours after every level */
    if( exchange ) {
        float shift = 0.5f;
        if( cfg.sift_mode == PSIFT_MODE_POPSIFT || cfg.sift_mode == PSIFT_MODE_VLFEAT ) shift = 0.5f * powf( 2.0f, cfg.upscale_factor );
        const int n_lvl = cfg.levels + 3, ds_level = n_lvl - 3;
        for( int r = 0; r < n_ctx; r++ ) {                               /* stage stamps and counters, as enqueue_frame does */
            Slot& s = slot_of( jobs[r] );
            BX( cudaSetDevice( ctxs[r]->device ) );
            BX( cudaEventRecord( s.ev[1], s.stream ) );
        }
        for( int o = 0; o <= split; o++ )
            for( int l = ( o == 0 ? 0 : 1 ); l < n_lvl; l++ ) {
                for( int r = 0; r < n_ctx; r++ ) {
                    Slot& s = slot_of( jobs[r] );
                    BX( cudaSetDevice( ctxs[r]->device ) );
                    BX( launch_pyramid_level( s.L, o, l, s.img_tex, shift, cfg.gauss_mode, ctxs[r]->gt, s.stream ) );
                    ctxs[r]->launches++;
                    BX( cudaEventRecord( ev[r], s.stream ) );
                }
                for( int r = 0; r < n_ctx; r++ ) {
                    Slot& s = slot_of( jobs[r] );
                    BX( cudaSetDevice( ctxs[r]->device ) );
                    for( int d = -1; d <= 1; d += 2 ) {
                        const int q = r + d;
                        if( q < 0 || q >= n_ctx ) continue;
                        const PyrLayout& T = slot_of( jobs[q] ).L;
                        BX( cudaStreamWaitEvent( s.stream, ev[q], 0 ) );
                        if( d < 0 ) BX( pull_rows( s.L, T, o, l, s.L.row_lo[o], s.L.cmp_lo[o], s.stream ) );
                        else        BX( pull_rows( s.L, T, o, l, s.L.cmp_hi[o], s.L.row_hi[o], s.stream ) );
                        if( l == ds_level && o < split ) {               /* the fused downscale wrote level 0 of the next octave */
                            if( d < 0 ) BX( pull_rows( s.L, T, o + 1, 0, s.L.row_lo[o + 1], s.L.cmp_lo[o + 1], s.stream ) );
                            else        BX( pull_rows( s.L, T, o + 1, 0, s.L.cmp_hi[o + 1], s.L.row_hi[o + 1], s.stream ) );
                        }
                    }
                }
                /* 3b. the seed rows exist as soon as level `ds_level` of the split octave does: send them on, ahead of the
                 * remaining levels */
                if( have_seed && o == split && l == ds_level )
                    for( int r = 0; r < n_ctx; r++ ) {
                        Slot& s = slot_of( jobs[r] );
                        const int so = split + 1;
                        BX( cudaSetDevice( ctxs[r]->device ) );
                        for( int q = 0; q < n_ctx; q++ )
                            BX( pull_rows( slot_of( seeded[q] ).L, s.L, so, 0, s.L.cmp_lo[so], s.L.cmp_hi[so], s.stream ) );
                        BX( cudaEventRecord( ev_seed[r], s.stream ) );
                    }
            }
        for( int r = 0; r < n_ctx; r++ ) {
            BX( cudaSetDevice( ctxs[r]->device ) );
            if( ( rc = run_frame( ctxs[r], slot_of( jobs[r] ), true ) ) != PSIFT_OK ) { cleanup(); return rc; }
        }
    } else if( have_seed ) {
        /* bands that recomputed their halo ran as ordinary jobs: their seed rows follow everything else on their stream */
        for( int r = 0; r < n_ctx; r++ ) {
            Slot& s = slot_of( jobs[r] );
            const int so = split + 1;
            const int a = bounds[r] >> so, b = ( r + 1 == n_ctx ) ? s.L.h[so] : std::min( s.L.h[so], bounds[r + 1] >> so );
            BX( cudaSetDevice( ctxs[r]->device ) );
            for( int q = 0; q < n_ctx; q++ )
                BX( pull_rows( slot_of( seeded[q] ).L, s.L, so, 0, a, b, s.stream ) );
            BX( cudaEventRecord( ev_seed[r], s.stream ) );
        }
    }
    lap( "levels + exchange queued" );
    /* 3c. the small octaves, once every band's seed rows have arrived */
    if( have_seed )
        for( int r = 0; r < n_ctx; r++ ) {
            Slot& s = slot_of( seeded[r] );
            BX( cudaSetDevice( ctxs[r]->device ) );
            for( int q = 0; q < n_ctx; q++ ) BX( cudaStreamWaitEvent( s.stream, ev_seed[q], 0 ) );
            if( ( rc = run_frame( ctxs[r], s ) ) != PSIFT_OK ) { cleanup(); return rc; }
        }
#undef BX

    lap( "seeded passes queued" );
    /* 4. results: once every part's counters are known, each part's records and descriptors are downloaded straight to
     * their place in ONE page-locked block from the recycled pool (no merge copy); descriptor indices are re-based there */
    long long tf = 0, td = 0;
    const int n_parts = have_seed ? 2 * n_ctx : n_ctx;
    std::vector<long long> f0( n_parts + 1, 0 ), d0( n_parts + 1, 0 );
    std::vector<int> nf( n_parts, 0 ), nd( n_parts, 0 );
    auto part_job = [&]( int k ) { return k < n_ctx ? jobs[k] : seeded[k - n_ctx]; };
    for( int k = 0; k < n_parts; k++ ) {
        psift_job* j = part_job( k );
        Slot& s = slot_of( j );
        cudaSetDevice( j->ctx->device );
        if( ( rc = settle_frame( j->ctx, s ) ) != PSIFT_OK ) { cleanup(); return rc; }
        nf[k] = s.h_ct->ext_total; nd[k] = s.h_ct->ori_total;
        f0[k] = tf; d0[k] = td;
        tf += nf[k]; td += nd[k];
    }
    lap( "all parts finished" );
    psift_result_block* blk = result_pool().get( (int)std::max<long long>( tf, 1 ), (int)std::max<long long>( td, 1 ) );
    if( !blk ) { cleanup(); return fail( PSIFT_E_CUDA, "out of page-locked host memory for %lld keypoints / %lld descriptors", tf, td ); }
    cudaError_t ce = cudaSuccess;
    for( int k = 0; k < n_parts && ce == cudaSuccess; k++ ) {
        psift_job* j = part_job( k );
        Slot& s = slot_of( j );
        cudaSetDevice( j->ctx->device );
        if( nf[k] > 0 ) {
            ce = cudaMemcpyAsync( blk->feat + f0[k], s.feat, sizeof(psift_feature) * (size_t)nf[k], cudaMemcpyDeviceToHost, s.stream );
            if( ce == cudaSuccess ) ce = cudaMemcpyAsync( blk->ext + f0[k], s.ext, sizeof(psift_extremum) * (size_t)nf[k], cudaMemcpyDeviceToHost, s.stream );
        }
        if( ce == cudaSuccess && nd[k] > 0 )
            ce = cudaMemcpyAsync( blk->desc + (size_t)128 * d0[k], s.desc, sizeof(float) * 128 * (size_t)nd[k], cudaMemcpyDeviceToHost, s.stream );
        j->ctx->d2h_bytes += (long long)nf[k] * ( sizeof(psift_feature) + sizeof(psift_extremum) ) + (long long)nd[k] * 512;
    }
    for( int k = 0; k < n_parts && ce == cudaSuccess; k++ ) {
        psift_job* j = part_job( k );
        cudaSetDevice( j->ctx->device );
        ce = cudaStreamSynchronize( slot_of( j ).stream );
    }
    if( ce != cudaSuccess ) { cudaGetLastError(); result_pool().put( blk ); cleanup(); return fail( PSIFT_E_CUDA, "result download: %s", cudaGetErrorString( ce ) ); }
    lap( "downloaded" );
    for( int k = 1; k < n_parts; k++ ) {
        const int base = (int)d0[k];
        if( !base ) continue;
        psift_feature* f = blk->feat + f0[k]; psift_extremum* e = blk->ext + f0[k];
        for( int i = 0; i < nf[k]; i++ ) {
            for( int q = 0; q < PSIFT_MAX_ORI; q++ ) if( f[i].desc_idx[q] >= 0 ) f[i].desc_idx[q] += base;
            e[i].idx_ori += base;
        }
    }
    out->n_feat = (int)tf; out->n_desc = (int)td;
    out->feat = blk->feat; out->ext = blk->ext; out->desc = blk->desc; out->block = blk;
    lap( "re-based" );
    cleanup();
    return PSIFT_OK;
}

void psift_banded_free( psift_banded_result* r )
{
    if( !r ) return;
    result_pool().put( (psift_result_block*)r->block );       /* the arrays live in a page-locked block of the recycled pool */
    memset( r, 0, sizeof *r );
}

} // extern "C"
