/* k_extrema.cu -- DoG + 3x3x3 extrema search + sub-pixel refinement, fused, for sm_100a.
 *
 * The reference writes five DoG planes per octave (make_dog, s_pyramid_build.cu:74-92) and
 * scans them with one thread per voxel and point-texture reads (find_extrema_in_dog,
 * s_extrema.cu:300-562).  Here the DoG is never stored: one thread owns a pixel column
 * (x,y) of one octave, loads the Gaussian levels it needs once, forms D[l] = G[l+1]-G[l] on
 * the fly (the same single rounding as make_dog) and tests the `levels` inner DoG levels.
 * Only voxels that pass the first contrast test look at their 26 neighbours, and only strict
 * extrema run the Newton refinement, whose neighbour values are re-derived from the
 * L2-resident Gaussian planes.  One launch covers all octaves.
 *
 * The arithmetic of the refinement (gradient, Hessian, closed-form inverse, step rules,
 * contrast and edge tests) follows s_extrema.cu:122-504 and s_solve.h:24-85 operation by
 * operation; the fused-multiply-add placement was read off the SASS of the reference build
 * for sm_100 and is spelled out with intrinsics so that it cannot drift.
 */
#include "common.h"
#include "kernels.h"
#include <algorithm>

namespace psift {

struct DogView {
    const float* base;
    size_t       lstride;
    int          pitch, w, h, nlvl;      /* nlvl = number of Gaussian levels */
    int          y_lo, y_hi;             /* stored rows (the whole plane unless banded): reads stay inside them */
    /* DoG level z at (x,y), clamp-to-edge in x,y and clamp of the layer index like the
     * reference's layered point texture (sift_octave.cu:231-248, common/assist.h:66-75) */
    __device__ __forceinline__ float at( int x, int y, int z ) const
    {
        x = min( max( x, 0 ), w - 1 );
        y = min( max( y, y_lo ), y_hi - 1 );      /* y_lo = 0, y_hi = h: the reference's clamp */
        z = min( max( z, 0 ), nlvl - 2 );
        const float* p = base + (size_t)z * lstride + (size_t)y * pitch + x;
        return __fsub_rn( __ldg( p + lstride ), __ldg( p ) );
    }
};

/* strict maximum or strict minimum of the 26 neighbours (s_extrema.cu:56-120) */
__device__ __forceinline__ bool strict_extremum( const DogView& d, int x, int y, int z, float v )
{
    bool gt = true, lt = true;
    #pragma unroll
    for( int dz = 0; dz < 3; dz++ ) {
        /* own layer first: most candidates die there */
        const int zz = z + ( dz == 0 ? 0 : ( dz == 1 ? -1 : 1 ) );
        #pragma unroll
        for( int dy = -1; dy <= 1; dy++ )
            #pragma unroll
            for( int dx = -1; dx <= 1; dx++ ) {
                if( dz == 0 && dy == 0 && dx == 0 ) continue;
                const float f = d.at( x + dx, y + dy, zz );
                gt = gt && ( v > f );
                lt = lt && ( v < f );
            }
        if( !gt && !lt ) return false;
    }
    return gt || lt;
}

/* closed-form solve of the symmetric 3x3 system (s_solve.h:24-85); A = Dxx, B = Dyy, C = Dss,
 * a01 = Dxy, a02 = Dxs, a12 = Dys; b is overwritten by the solution.  false: singular. */
__device__ __forceinline__ bool solve3( float A, float B, float C, float a01, float a02, float a12, float3& b )
{
    const float det0 = __fmaf_rn( B,   C,   -__fmul_rn( a12, a12 ) );
    const float det1 = __fmaf_rn( a12, a02, -__fmul_rn( a01, C   ) );
    const float det2 = __fmaf_rn( a01, a12, -__fmul_rn( B,   a02 ) );
    const float det3 = __fmaf_rn( A,   C,   -__fmul_rn( a02, a02 ) );
    const float det4 = __fmaf_rn( a01, a02, -__fmul_rn( A,   a12 ) );
    const float det5 = __fmaf_rn( A,   B,   -__fmul_rn( a01, a01 ) );
    const float det  = __fmaf_rn( a02, det2, __fmaf_rn( A, det0, __fmul_rn( a01, det1 ) ) );
    if( det == 0.0f ) return false;
    const float r = __frcp_rn( det );
    const float i00 = __fmul_rn( det0, r ), i10 = __fmul_rn( det1, r ), i20 = __fmul_rn( det2, r );
    const float i11 = __fmul_rn( det3, r ), i12 = __fmul_rn( det4, r ), i22 = __fmul_rn( det5, r );
    const float bx = b.x, by = b.y, bz = b.z;
    b.x = __fmaf_rn( i20, bz, __fmaf_rn( i10, by, __fmaf_rn( i00, bx, 0.0f ) ) );
    b.y = __fmaf_rn( i12, bz, __fmaf_rn( i11, by, __fmaf_rn( i10, bx, 0.0f ) ) );
    b.z = __fmaf_rn( i22, bz, __fmaf_rn( i12, by, __fmaf_rn( i20, bx, 0.0f ) ) );
    return true;
}

template<int MODE>
__device__ __forceinline__ bool first_contrast_ok( float v, float thr )
{
    if( MODE == PSIFT_MODE_OPENCV ) return fabsf( v ) >= floorf( thr );                 /* s_extrema.cu:152 */
    if( MODE == PSIFT_MODE_VLFEAT ) return fabsf( v ) >= __fmul_rn( 0.8f * 2.0f, thr );  /* :203 */
    return fabsf( v ) >= __fmul_rn( 1.6f, thr );                                         /* :255 */
}

/* step rule of one Newton iteration: -1 reject, 0 iterate again, 1 converged
 * (s_extrema.cu:156-190 OpenCV, :207-233 VLFeat, :259-287 PopSift) */
template<int MODE>
__device__ __forceinline__ int step_rule( const float3& d, int3& n, int w, int h, int maxlevel, bool last )
{
    if( MODE == PSIFT_MODE_OPENCV ) {
        if( fabsf( d.x ) < 0.5f && fabsf( d.y ) < 0.5f && fabsf( d.z ) < 0.5f ) return 1;
        n.x = (int)( (float)n.x + roundf( d.x ) );
        n.y = (int)( (float)n.y + roundf( d.y ) );
        n.z = (int)( (float)n.z + roundf( d.z ) );
        return ( n.x < 5 || n.x >= w - 5 || n.y < 5 || n.y >= h - 5 || n.z < 1 || n.z > maxlevel - 2 ) ? -1 : 0;
    }
    if( last ) return 0;
    const int tx = ( ( d.x >= 0.6f && n.x < w - 2 ) ? 1 : 0 ) + ( ( d.x <= -0.6f && n.x > 1 ) ? -1 : 0 );
    const int ty = ( ( d.y >= 0.6f && n.y < h - 2 ) ? 1 : 0 ) + ( ( d.y <= -0.6f && n.y > 1 ) ? -1 : 0 );
    int tz = 0;
    if( MODE == PSIFT_MODE_POPSIFT )   /* VLFeat mode never changes the level, s_extrema.cu:221-229 */
        tz = ( ( d.z >= 0.6f && n.z < maxlevel - 1 ) ? 1 : 0 ) + ( ( d.z <= -0.6f && n.z > 1 ) ? -1 : 0 );
    if( tx == 0 && ty == 0 && tz == 0 ) return 1;
    n.x += tx; n.y += ty; n.z += tz;
    return 0;
}

template<int MODE, class View>
__device__ bool refine_extremum( const View& dv, const Consts& k, int x, int y, int level, float val,
                                 float w_div, float h_div, Cand& out )
{
    const int w = dv.w, h = dv.h, maxlevel = dv.nlvl - 1;
    float3 D, d;
    float  Dxx, Dyy, Dss, Dxy, Dxs, Dys;
    int3   n = make_int3( x, y, level );
    int    iter = 0;
    d = make_float3( 0.f, 0.f, 0.f );
    do {
        iter++;
        const float x2 = dv.at( n.x + 1, n.y, n.z ), x0 = dv.at( n.x - 1, n.y, n.z );
        const float y2 = dv.at( n.x, n.y + 1, n.z ), y0 = dv.at( n.x, n.y - 1, n.z );
        const float z2 = dv.at( n.x, n.y, n.z + 1 ), z0 = dv.at( n.x, n.y, n.z - 1 );
        const float c  = dv.at( n.x, n.y, n.z );
        D.x = scalbnf( __fsub_rn( x2, x0 ), -1 );
        D.y = scalbnf( __fsub_rn( y2, y0 ), -1 );
        D.z = scalbnf( __fsub_rn( z2, z0 ), -1 );
        const float c2 = scalbnf( c, 1 );
        Dxx = __fsub_rn( __fadd_rn( x2, x0 ), c2 );
        Dyy = __fsub_rn( __fadd_rn( y2, y0 ), c2 );
        Dss = __fsub_rn( __fadd_rn( z2, z0 ), c2 );
        /* mixed derivatives: ((++) + (--)) - (-+) - (+-), then /4  (s_extrema.cu:401-403) */
        Dxy = scalbnf( __fsub_rn( __fsub_rn( __fadd_rn( dv.at( n.x + 1, n.y + 1, n.z ), dv.at( n.x - 1, n.y - 1, n.z ) ),
                                             dv.at( n.x - 1, n.y + 1, n.z ) ), dv.at( n.x + 1, n.y - 1, n.z ) ), -2 );
        Dxs = scalbnf( __fsub_rn( __fsub_rn( __fadd_rn( dv.at( n.x + 1, n.y, n.z + 1 ), dv.at( n.x - 1, n.y, n.z - 1 ) ),
                                             dv.at( n.x - 1, n.y, n.z + 1 ) ), dv.at( n.x + 1, n.y, n.z - 1 ) ), -2 );
        Dys = scalbnf( __fsub_rn( __fsub_rn( __fadd_rn( dv.at( n.x, n.y + 1, n.z + 1 ), dv.at( n.x, n.y - 1, n.z - 1 ) ),
                                             dv.at( n.x, n.y + 1, n.z - 1 ) ), dv.at( n.x, n.y - 1, n.z + 1 ) ), -2 );
        float3 b = make_float3( -D.x, -D.y, -D.z );
        if( !solve3( Dxx, Dyy, Dss, Dxy, Dxs, Dys, b ) ) {      /* singular: keep the point, zero offset */
            d = make_float3( 0.f, 0.f, 0.f );
            break;
        }
        d = b;
        const int rv = step_rule<MODE>( d, n, w, h, maxlevel, iter == 5 );
        if( rv == -1 ) return false;
        if( rv == 1 ) break;
    } while( iter < 5 );

    if( MODE == PSIFT_MODE_OPENCV ) {
        if( iter >= 5 ) return false;                            /* s_extrema.cu:448-453 */
    } else {
        if( d.x >= 1.5f || d.y >= 1.5f || d.z >= 1.5f ) return false;   /* signed on purpose, :455-460 */
    }
    const float xn = __fadd_rn( (float)n.x, d.x );
    const float yn = __fadd_rn( (float)n.y, d.y );
    const float sn = __fadd_rn( (float)n.z, d.z );
    if( MODE != PSIFT_MODE_OPENCV ) {                             /* verify(), :235-246 / :289-298 */
        if( xn < 0.0f || xn > (float)w - 1.0f || yn < 0.0f || yn > (float)h - 1.0f || sn < 0.0f || sn > (float)maxlevel )
            return false;
    }
    const float dot   = __fmaf_rn( D.z, d.z, __fmaf_rn( D.y, d.y, __fmul_rn( D.x, d.x ) ) );
    const float contr = __fadd_rn( val, scalbnf( dot, -1 ) );
    const float tr    = __fadd_rn( Dxx, Dyy );
    const float det   = __fmaf_rn( Dxx, Dyy, -__fmul_rn( Dxy, Dxy ) );
    const float edge  = __fdiv_rn( __fmul_rn( tr, tr ), det );
    if( det <= 0.0f ) return false;
    if( fabsf( contr ) < scalbnf( k.threshold, 1 ) ) return false;
    const float el1 = __fadd_rn( k.edge_limit, 1.0f );
    if( edge >= __fdiv_rn( __fmul_rn( el1, el1 ), k.edge_limit ) ) return false;

    out.x     = xn;
    out.y     = yn;
    out.lpos  = (int)roundf( sn );
    out.sigma = __fmul_rn( k.sigma0, powf( k.sigma_k, sn ) );
    out.cell  = (int)__fadd_rn( __fmul_rn( floorf( __fdiv_rn( yn, h_div ) ), (float)k.grid_size ), floorf( __fdiv_rn( xn, w_div ) ) );
    return true;
}

/* refinement of one strict extremum and, if it survives, its append to the octave's list (s_extrema.cu:22-44) */
template<int MODE, class View>
__device__ __forceinline__ void refine_and_append( const View& dv, const Consts& K, int o, int x, int y, int level, float val,
                                                   Cand* cand, int cand_stride, Counters* ct )
{
    const float w_div = __fdiv_rn( (float)dv.w, (float)K.grid_size );
    const float h_div = __fdiv_rn( (float)dv.h, (float)K.grid_size );
    Cand c;
    if( refine_extremum<MODE>( dv, K, x, y, level, val, w_div, h_div, c ) ) {
        c.key_lo = (uint32_t)x | ( (uint32_t)( y & 0xf ) << 28 );
        c.key_hi = ( (uint32_t)y >> 4 ) | ( (uint32_t)level << 24 );
        const int idx = atomicAdd( &ct->ext_ct[o], 1 );
        if( idx < K.max_extrema ) cand[(size_t)o * cand_stride + idx] = c;
    }
}

/* one queued extremum: x | y << 16, level | octave << 8, DoG value */
template<int MODE>
__device__ __forceinline__ void refine_pending( const PyrLayout& L, const Consts& K, int n_lvl, unsigned xy, unsigned lo, float val,
                                                Cand* cand, int cand_stride, Counters* ct )
{
    const int o = (int)( lo >> 8 );
    DogView dv;
    dv.base = L.base[o]; dv.lstride = L.lstride[o]; dv.pitch = L.pitch[o]; dv.w = L.w[o]; dv.h = L.h[o]; dv.nlvl = n_lvl;
    dv.y_lo = L.row_lo[o]; dv.y_hi = L.row_hi[o];
    refine_and_append<MODE>( dv, K, o, (int)( xy & 0xffffu ), (int)( xy >> 16 ), (int)( lo & 0xffu ), val, cand, cand_stride, ct );
}

/* block -> (octave, tile) table */
struct ExtGrid {
    int first_block[MAX_OCT + 1];
    int tiles_x[MAX_OCT];
};

constexpr int EX_TX = 64, EX_TY = 16, EX_NT = 256;
constexpr int EX_SW = EX_TX + 8;           /* DoG tile row: tile column x sits at x+4, so that rows load and read as float4 */
constexpr int EX_SH = EX_TY + 2;
constexpr int EX_PEND = 96;                 /* strict extrema of a block waiting for their refinement */

static size_t extrema_smem( int n_lvl )
{
    /* Strict extrema of one kind cannot be neighbours, so a tile of 64 x 16 x nl voxels holds at most
     * 32 * 8 * ceil(nl / 2) maxima and as many minima: 256 * (nl + 1) entries bound the second queue. */
    return sizeof(float) * 2 * (size_t)n_lvl * EX_SH * EX_SW                      /* two raw tiles (the DoG replaces one in place) */
         + sizeof(float) * 256 * (size_t)( n_lvl - 2 )                            /* second queue: DoG values */
         + sizeof(unsigned short) * (size_t)( n_lvl - 3 ) * EX_TX * EX_TY         /* first queue */
         + sizeof(unsigned short) * 256 * (size_t)( n_lvl - 2 );                  /* second queue: voxels */
}

/* Persistent blocks; one step = one 64x16 tile of one octave, all levels.
 *   copy    : every Gaussian level of the NEXT tile (+1 halo) is requested with 16-byte asynchronous copies into the
 *             second buffer before the current tile is looked at, so a block always has a tile's worth of reads in
 *             flight and never waits for one (the one-tile-per-block version spent a memory latency per tile with
 *             nothing else to do: 0.50 of the HBM bandwidth, profiles/r02h_ncu_frame_summary.txt)
 *   DoG     : differences of neighbouring levels, in place
 *   phase 1 : first contrast test on each voxel's own value; survivors are queued (warp-aggregated)
 *   phase 2 : 26-neighbour test out of shared memory, balanced over the block; strict extrema queued
 *   phase 3 : Newton refinement for the few that are left (neighbours re-derived from the
 *             L2-resident Gaussian planes, since the iteration may walk out of the tile)
 * NLVL > 0 fixes the number of Gaussian levels at compile time; NLVL == 0 takes it from the layout. */
template<int MODE, int NLVL>
__global__ void __launch_bounds__( EX_NT )
k_extrema( const __grid_constant__ PyrLayout L, const __grid_constant__ ExtGrid G, const __grid_constant__ Consts K,
           const unsigned char* tmaps, Cand* cand, int cand_stride, Counters* ct )
{
    extern __shared__ __align__( 128 ) float ex_smem[];
    __shared__ int n1[2], n2[2];
    __shared__ __align__( 8 ) unsigned long long bars[2];
    __shared__ unsigned pend_xy[EX_PEND], pend_lo[EX_PEND];
    __shared__ float    pend_val[EX_PEND];
    int n_pend = 0;                                                             /* first warp only (uniform in it) */
    pdl_enter();
    const int n_lvl = NLVL ? NLVL : L.n_lvl;
    const int qcap  = ( n_lvl - 3 ) * EX_TX * EX_TY;
    const int tile_floats = n_lvl * EX_SH * EX_SW;
    float*          q2v = ex_smem + 2 * tile_floats;
    unsigned short* q1  = reinterpret_cast<unsigned short*>( q2v + 256 * ( n_lvl - 2 ) );
    unsigned short* q2  = q1 + qcap;
    const int tid = threadIdx.x;
    const int n_tiles = G.first_block[MAX_OCT];
    constexpr int LQ = EX_SW / 4;

    /* Tile column x is stored at x+4: shared column 3 is the left halo, EX_TX+4 the right one.  One thread asks the
     * TMA unit for the whole box (72 columns x 18 rows x all levels = one request); completion is counted in bytes on
     * the buffer's mbarrier. */
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared( bars );
    const unsigned buf0 = (unsigned)__cvta_generic_to_shared( ex_smem );
    const unsigned tile_bytes = 4u * tile_floats;
    auto request = [&]( int t, int kb ) {
        int o = 0;
        while( o + 1 < L.n_oct && t >= G.first_block[o + 1] ) o++;
        const int tile = t - G.first_block[o];
        const int tx0 = ( tile % G.tiles_x[o] ) * EX_TX;
        const int ty  = ( tile / G.tiles_x[o] ) * EX_TY;                         /* relative to the first stored row */
        asm volatile( "fence.proxy.async.shared::cta;" ::: "memory" );            /* the buffer was last written by ordinary stores (DoG) */
        asm volatile( "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"( bar0 + 8u * kb ), "r"( tile_bytes ) : "memory" );
        asm volatile( "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                      :: "r"( buf0 + kb * tile_bytes ), "l"( tmaps + 128 * o ), "r"( tx0 - 4 ), "r"( ty - 1 ), "r"( 0 ), "r"( bar0 + 8u * kb ) : "memory" );
    };
    if( tid == 0 ) {
        asm volatile( "mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"( bar0 ) : "memory" );
        asm volatile( "mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"( bar0 + 8u ) : "memory" );
        asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
        n1[0] = n2[0] = n1[1] = n2[1] = 0;
    }
    __syncthreads();

    /* Per tile: wait for the box / DoG + phase 1 / barrier / phase 2 / barrier / hand the survivors on.  Buffers, mbarriers
     * and counters alternate, so a step never resets or refills what a slow warp of the previous step still reads. */
    int t = blockIdx.x, k = 0;
    unsigned parity = 0;                                                          /* bit k: phase of buffer k's mbarrier */
    if( tid == 0 && t < n_tiles ) request( t, 0 );
    for( ; t < n_tiles; t += gridDim.x, k ^= 1 ) {
        float* D = ex_smem + k * tile_floats;                                  /* [n_lvl][EX_SH][EX_SW]; DoG in [0, n_lvl-1) */
        if( tid == 0 ) {
            n1[k ^ 1] = 0; n2[k ^ 1] = 0;                                      /* the next step's counters */
            if( t + (int)gridDim.x < n_tiles ) request( t + gridDim.x, k ^ 1 );
        }
        asm volatile( "{\n\t.reg .pred p;\n\tEXW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra EXD;\n\tbra EXW;\n\tEXD:\n\t}"
                      :: "r"( bar0 + 8u * k ), "r"( ( parity >> k ) & 1u ) : "memory" );
        parity ^= 1u << k;

        int o = 0;
        while( o + 1 < L.n_oct && t >= G.first_block[o + 1] ) o++;
        const int tile = t - G.first_block[o];
        const int w = L.w[o], h = L.h[o];
        const int tx0 = ( tile % G.tiles_x[o] ) * EX_TX;
        const int ty0 = L.row_lo[o] + ( tile / G.tiles_x[o] ) * EX_TY;

        /* ---- DoG in place, four columns at a time; phase 1 (first contrast test) on the thread's own interior chunk
         * straight from the registers.  Item i < 256 is the interior chunk (row y+1, chunk g+1) of thread i, the
         * remaining 68 items are the halo chunks. ---- */
        {
            const int g = tid % ( EX_TX / 4 ), y = tid / ( EX_TX / 4 );       /* 16 x 16 threads = 64 x 16 pixels */
            const int gy = ty0 + y;
            unsigned inside = 0;
            #pragma unroll
            for( int j = 0; j < 4; j++ ) {
                const int gx = tx0 + 4 * g + j;
                bool in = ( gx >= 1 ) && ( gy >= 1 ) && ( gx <= w - 2 ) && ( gy <= h - 2 );
                in = in && ( gy >= L.own_lo[o] ) && ( gy < L.own_hi[o] );       /* banded: each voxel belongs to one context */
                if( MODE == PSIFT_MODE_OPENCV ) in = in && !( gx < 5 || gy < 5 || gx >= w - 5 || gy >= h - 5 );   /* :336-340 */
                inside |= in ? ( 1u << j ) : 0u;
            }
            float4* d = reinterpret_cast<float4*>( D ) + ( y + 1 ) * LQ + g + 1;
            float4 prev = d[0];
            #pragma unroll
            for( int l = 1; l < n_lvl; l++ ) {
                const float4 cur = d[l * ( EX_SH * LQ )];
                const float4 v = make_float4( __fsub_rn( cur.x, prev.x ), __fsub_rn( cur.y, prev.y ), __fsub_rn( cur.z, prev.z ), __fsub_rn( cur.w, prev.w ) );
                d[( l - 1 ) * ( EX_SH * LQ )] = v;
                prev = cur;
                const int level = l - 1;
                if( inside && level >= 1 && level <= n_lvl - 3 ) {
                    unsigned m = ( first_contrast_ok<MODE>( v.x, K.threshold ) ? 1u : 0u ) | ( first_contrast_ok<MODE>( v.y, K.threshold ) ? 2u : 0u )
                               | ( first_contrast_ok<MODE>( v.z, K.threshold ) ? 4u : 0u ) | ( first_contrast_ok<MODE>( v.w, K.threshold ) ? 8u : 0u );
                    m &= inside;
                    if( m ) {
                        int pos = atomicAdd( &n1[k], __popc( m ) );
                        #pragma unroll
                        for( int j = 0; j < 4; j++ )
                            if( m & ( 1u << j ) ) q1[pos++] = (unsigned short)( ( 4 * g + j ) | ( y << 6 ) | ( level << 10 ) );
                    }
                }
            }
            if( tid < EX_SH * LQ - EX_NT ) {
                /* halo chunks: rows 0 and EX_SH-1 (LQ chunks each), then chunks 0 and LQ-1 of the rows between */
                int sy, c4;
                if( tid < 2 * LQ ) { sy = tid < LQ ? 0 : EX_SH - 1; c4 = tid % LQ; }
                else               { sy = 1 + ( tid - 2 * LQ ) / 2; c4 = ( ( tid - 2 * LQ ) & 1 ) ? LQ - 1 : 0; }
                float4* e = reinterpret_cast<float4*>( D ) + sy * LQ + c4;
                float4 pv = e[0];
                for( int l = 1; l < n_lvl; l++ ) {
                    const float4 cur = e[l * ( EX_SH * LQ )];
                    e[( l - 1 ) * ( EX_SH * LQ )] = make_float4( __fsub_rn( cur.x, pv.x ), __fsub_rn( cur.y, pv.y ), __fsub_rn( cur.z, pv.z ), __fsub_rn( cur.w, pv.w ) );
                    pv = cur;
                }
            }
        }
        __syncthreads();

        /* ---- phase 2: strict maximum or minimum of the 26 neighbours (s_extrema.cu:56-120) ---- */
        const int cnt1 = n1[k];
        for( int i = tid; i < cnt1; i += EX_NT ) {
            const unsigned e = q1[i];
            const int x = e & 63, y = ( e >> 6 ) & 15, level = e >> 10;
            const float* c = D + ( level * EX_SH + y + 1 ) * EX_SW + x + 4;
            const float v = *c;
            bool gt = true, lt = true;
            #pragma unroll
            for( int dy = -1; dy <= 1; dy++ )
                #pragma unroll
                for( int dx = -1; dx <= 1; dx++ ) {
                    if( dy == 0 && dx == 0 ) continue;
                    const float f = c[dy * EX_SW + dx];
                    gt = gt && ( v > f ); lt = lt && ( v < f );
                }
            if( !gt && !lt ) continue;
            #pragma unroll
            for( int dz = -1; dz <= 1; dz += 2 )
                #pragma unroll
                for( int dy = -1; dy <= 1; dy++ )
                    #pragma unroll
                    for( int dx = -1; dx <= 1; dx++ ) {
                        const float f = c[dz * EX_SH * EX_SW + dy * EX_SW + dx];
                        gt = gt && ( v > f ); lt = lt && ( v < f );
                    }
            if( gt || lt ) { const int at = atomicAdd( &n2[k], 1 ); q2[at] = (unsigned short)e; q2v[at] = v; }
        }
        __syncthreads();

        /* ---- phase 3: the few that are left wait in a list of the block; whenever 32 have come together the first
         * warp refines them, one per lane, while the others move on to the next tile (they meet again at its first
         * barrier).  Two in five tiles leave one or two survivors: refining them at once kept the block waiting for a
         * warp with one busy lane (1.3 us of dependent arithmetic) every other tile. ---- */
        const int cnt2 = n2[k];
        if( tid < 32 ) {
            for( int i = tid; i < cnt2; i += 32 ) {
                const unsigned e = q2[i];
                const int at = n_pend + i;
                if( at < EX_PEND ) {
                    pend_xy[at]  = (unsigned)( tx0 + (int)( e & 63 ) ) | ( (unsigned)( ty0 + (int)( ( e >> 6 ) & 15 ) ) << 16 );
                    pend_lo[at]  = (unsigned)( e >> 10 ) | ( (unsigned)o << 8 );
                    pend_val[at] = q2v[i];
                } else {                                                        /* more than the list holds in one tile: at once */
                    DogView dv;
                    dv.base = L.base[o]; dv.lstride = L.lstride[o]; dv.pitch = L.pitch[o]; dv.w = w; dv.h = h; dv.nlvl = n_lvl;
                    dv.y_lo = L.row_lo[o]; dv.y_hi = L.row_hi[o];
                    refine_and_append<MODE>( dv, K, o, tx0 + (int)( e & 63 ), ty0 + (int)( ( e >> 6 ) & 15 ), (int)( e >> 10 ), q2v[i], cand, cand_stride, ct );
                }
            }
            n_pend = min( n_pend + cnt2, EX_PEND );
            __syncwarp();
            while( n_pend >= 32 ) {
                n_pend -= 32;
                refine_pending<MODE>( L, K, n_lvl, pend_xy[n_pend + tid], pend_lo[n_pend + tid], pend_val[n_pend + tid], cand, cand_stride, ct );
                __syncwarp();
            }
        }
        /* no barrier here: the next step uses the other buffer and the other counters, and its own barriers come
         * before anything of this step is overwritten */
    }
    if( tid < 32 && tid < n_pend )
        refine_pending<MODE>( L, K, n_lvl, pend_xy[tid], pend_lo[tid], pend_val[tid], cand, cand_stride, ct );
}

/* Clamp the per-octave counters to max_extrema (s_extrema.cu:555-561), build the prefix sums,
 * and flag overflow.  One thread. */
__global__ void k_extrema_finish( Counters* ct, int n_oct, int max_extrema, int ext_capacity )
{
    pdl_enter();
    if( threadIdx.x != 0 || blockIdx.x != 0 ) return;
    int ps = 0, need = 0;
    for( int o = 0; o < MAX_OCT; o++ ) {
        int c = ( o < n_oct ) ? min( ct->ext_ct[o], max_extrema ) : 0;
        need += c;
        if( ps + c > ext_capacity ) { c = max( 0, ext_capacity - ps ); ct->overflow = 1; }
        ct->ext_ct[o] = c;
        ct->ext_ps[o] = ps;
        ps += c;
    }
    ct->ext_ps[MAX_OCT] = ps;
    ct->ext_total = ps;
    ct->need_ext = need;
}

/* Deterministic order: every candidate of an octave goes to ext_ps[o] + (its rank under the key (level, y, x) of its
 * detection voxel).  The reference's order is whatever its atomics produced (its own regression test sorts the output,
 * testOxfordDataset.sh.in:138-155).  Keys are unique, so this is a permutation.
 * Linear in the number of candidates (a giant image has > 100 000 per octave): counting sort by bucket = (level, band of
 * rows) - monotonic in the key -, then the exact rank inside the candidate's own bucket (a few dozen entries) by
 * comparison.  The scatter into bucket order uses an atomic cursor, so the order INSIDE a bucket of the intermediate
 * list is arbitrary; the rank step does not depend on it. */
constexpr int SORT_NT = 256;
__device__ __forceinline__ unsigned long long cand_key( const Cand& c ) { return ( (unsigned long long)c.key_hi << 32 ) | c.key_lo; }
__device__ __forceinline__ int cand_bucket( const Cand& c, int yshift )
{
    const int level = (int)( c.key_hi >> 24 );
    const int y     = (int)( ( ( c.key_hi & 0xffffffu ) << 4 ) | ( c.key_lo >> 28 ) );
    return min( level, MAX_LVL - 1 ) * SORT_NBY + min( y >> yshift, SORT_NBY - 1 );
}

__global__ void __launch_bounds__( SORT_NT )
k_sort_count( const Cand* cand, int cand_stride, const Counters* ct, SortScratch* ss, const __grid_constant__ SortShifts sh )
{
    pdl_enter();
    const int o = blockIdx.y, n = ct->ext_ct[o];
    const Cand* src = cand + (size_t)o * cand_stride;
    for( int i = blockIdx.x * SORT_NT + threadIdx.x; i < n; i += gridDim.x * SORT_NT )
        atomicAdd( &ss->start[o][cand_bucket( src[i], sh.y[o] )], 1 );
}

/* counts -> exclusive prefix, one block per octave */
__global__ void __launch_bounds__( 1024 )
k_sort_scan( const Counters* ct, SortScratch* ss )
{
    __shared__ int wsum[32];
    pdl_enter();
    const int o = blockIdx.x;
    if( ct->ext_ct[o] == 0 ) return;
    constexpr int PER = SORT_NB / 1024;
    static_assert( SORT_NB % 1024 == 0, "buckets per octave: a multiple of the block size" );
    int v[PER], tot = 0;
    #pragma unroll
    for( int k = 0; k < PER; k++ ) { v[k] = ss->start[o][threadIdx.x * PER + k]; tot += v[k]; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = tot;
    #pragma unroll
    for( int d = 1; d < 32; d <<= 1 ) { const int u = __shfl_up_sync( 0xffffffffu, inc, d ); if( lane >= d ) inc += u; }
    if( lane == 31 ) wsum[wid] = inc;
    __syncthreads();
    if( wid == 0 ) {
        int w = wsum[lane];
        #pragma unroll
        for( int d = 1; d < 32; d <<= 1 ) { const int u = __shfl_up_sync( 0xffffffffu, w, d ); if( lane >= d ) w += u; }
        wsum[lane] = w;
    }
    __syncthreads();
    int run = inc - tot + ( wid ? wsum[wid - 1] : 0 );
    #pragma unroll
    for( int k = 0; k < PER; k++ ) { ss->start[o][threadIdx.x * PER + k] = run; run += v[k]; }
}

__global__ void __launch_bounds__( SORT_NT )
k_sort_scatter( const Cand* cand, int cand_stride, const Counters* ct, SortScratch* ss, const __grid_constant__ SortShifts sh, Cand* grouped )
{
    pdl_enter();
    const int o = blockIdx.y, n = ct->ext_ct[o];
    const Cand* src = cand + (size_t)o * cand_stride;
    Cand* dst = grouped + ct->ext_ps[o];
    for( int i = blockIdx.x * SORT_NT + threadIdx.x; i < n; i += gridDim.x * SORT_NT ) {
        const Cand c = src[i];
        const int b = cand_bucket( c, sh.y[o] );
        dst[ss->start[o][b] + atomicAdd( &ss->fill[o][b], 1 )] = c;
    }
}

__global__ void __launch_bounds__( SORT_NT )
k_sort_rank( const Cand* grouped, const Counters* ct, const SortScratch* ss, const __grid_constant__ SortShifts sh, Cand* sorted )
{
    pdl_enter();
    const int o = blockIdx.y, n = ct->ext_ct[o];
    const Cand* src = grouped + ct->ext_ps[o];
    Cand* dst = sorted + ct->ext_ps[o];
    for( int i = blockIdx.x * SORT_NT + threadIdx.x; i < n; i += gridDim.x * SORT_NT ) {
        const Cand c = src[i];
        const int b = cand_bucket( c, sh.y[o] );
        const int lo = ss->start[o][b], hi = lo + ss->fill[o][b];
        const unsigned long long k = cand_key( c );
        int rank = 0;
        for( int j = lo; j < hi; j++ ) rank += ( cand_key( src[j] ) < k ) ? 1 : 0;
        dst[lo + rank] = c;
    }
}

/* deterministic_order = 0: the candidates of every octave, in the order their atomics arrived (what the reference returns) */
__global__ void __launch_bounds__( SORT_NT )
k_compact_unsorted( const Cand* cand, int cand_stride, const Counters* ct, Cand* sorted )
{
    pdl_enter();
    const int o = blockIdx.y, n = ct->ext_ct[o];
    const Cand* src = cand + (size_t)o * cand_stride;
    Cand* dst = sorted + ct->ext_ps[o];
    for( int i = blockIdx.x * SORT_NT + threadIdx.x; i < n; i += gridDim.x * SORT_NT ) dst[i] = src[i];
}

void extrema_box( int* columns, int* rows ) { *columns = EX_SW; *rows = EX_SH; }

cudaError_t launch_extrema( const PyrLayout& L, const Consts& K, const void* tensor_maps, Cand* cand, int cand_stride, Cand* sorted, Cand* sort_tmp,
                            int ext_capacity, Counters* ct, SortScratch* ss, cudaStream_t st, long long* launches )
{
    ExtGrid G;
    int nb = 0;
    for( int o = 0; o < L.n_oct; o++ ) {
        G.first_block[o] = nb;
        const int tx = ( L.w[o] + EX_TX - 1 ) / EX_TX;
        const int ty = ( L.row_hi[o] - L.row_lo[o] + EX_TY - 1 ) / EX_TY;
        G.tiles_x[o] = tx;
        if( o >= L.oct_begin && o < L.oct_end ) nb += tx * ty;          /* octaves outside the pass get no blocks */
    }
    for( int o = L.n_oct; o <= MAX_OCT; o++ ) G.first_block[o] = nb;
    if( nb == 0 ) return cudaErrorInvalidValue;
    const size_t smem = extrema_smem( L.n_lvl );
    if( L.n_lvl > 60 ) return cudaErrorInvalidValue;                      /* queue entries hold 6 bits of level */
#define PSIFT_EXT_LAUNCH( MODE )                                                                             \
    do {                                                                                                      \
        if( L.n_lvl == 6 ) { PSIFT_EXT_LAUNCH2( ( k_extrema<MODE, 6> ) ); }                                   \
        else               { PSIFT_EXT_LAUNCH2( ( k_extrema<MODE, 0> ) ); }                                   \
    } while( 0 )
#define PSIFT_EXT_LAUNCH2( KERNEL )                                                                           \
    do {                                                                                                      \
        cudaError_t e = cudaFuncSetAttribute( KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem ); \
        if( e != cudaSuccess ) return e;                                                                      \
        int per_sm = 0;                                                                                       \
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor( &per_sm, KERNEL, EX_NT, smem );                    \
        if( e != cudaSuccess || per_sm < 1 ) return e != cudaSuccess ? e : cudaErrorInvalidConfiguration;     \
        const int blocks = std::min( nb, per_sm * sm_count() );                                               \
        launch_pdl( KERNEL, dim3( blocks ), dim3( EX_NT ), smem, st, L, G, K, (const unsigned char*)tensor_maps, cand, cand_stride, ct );  \
    } while( 0 )
    switch( K.sift_mode ) {
    case PSIFT_MODE_OPENCV: PSIFT_EXT_LAUNCH( PSIFT_MODE_OPENCV ); break;
    case PSIFT_MODE_VLFEAT: PSIFT_EXT_LAUNCH( PSIFT_MODE_VLFEAT ); break;
    default:                PSIFT_EXT_LAUNCH( PSIFT_MODE_POPSIFT ); break;
    }
#undef PSIFT_EXT_LAUNCH2
    launch_pdl( k_extrema_finish, dim3( 1 ), dim3( 32 ), 0, st, ct, L.n_oct, K.max_extrema, ext_capacity );
    /* order the candidates (ss was zeroed together with the counters at the start of the frame) */
    SortShifts sh;
    for( int o = 0; o < MAX_OCT; o++ ) { int q = 0; while( o < L.n_oct && ( ( L.h[o] - 1 ) >> q ) >= SORT_NBY ) q++; sh.y[o] = q; }
    const int gx = std::max( 1, std::min( 4 * sm_count(), ( ext_capacity + SORT_NT - 1 ) / SORT_NT ) ) / std::max( 1, L.oct_end - L.oct_begin ) + 1;
    if( !K.deterministic ) {
        launch_pdl( k_compact_unsorted, dim3( gx, L.n_oct ), dim3( SORT_NT ), 0, st, cand, cand_stride, ct, sorted );
        if( launches ) *launches += 3;
        return cudaGetLastError();
    }
    launch_pdl( k_sort_count,   dim3( gx, L.n_oct ), dim3( SORT_NT ), 0, st, cand, cand_stride, ct, ss, sh );
    launch_pdl( k_sort_scan,    dim3( L.n_oct ),     dim3( 1024 ),    0, st, ct, ss );
    launch_pdl( k_sort_scatter, dim3( gx, L.n_oct ), dim3( SORT_NT ), 0, st, cand, cand_stride, ct, ss, sh, sort_tmp );
    launch_pdl( k_sort_rank,    dim3( gx, L.n_oct ), dim3( SORT_NT ), 0, st, sort_tmp, ct, ss, sh, sorted );
    if( launches ) *launches += 6;
    return cudaGetLastError();
}

} // namespace psift
