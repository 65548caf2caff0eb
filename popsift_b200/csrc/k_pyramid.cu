/* k_pyramid.cu -- Gaussian pyramid for sm_100a.
 *
 * One kernel builds one Gaussian level: it stages the source tile plus halo in shared
 * memory, runs the horizontal pass into a second shared buffer and the vertical pass
 * out of it, and writes the level once.  The horizontal-pass plane that the reference
 * stores and re-reads through textures (its _intm array, s_pyramid_build.cu:249-376) never
 * leaves the SM, and the "pick every second pixel" downscale that seeds the next octave
 * (s_pyramid_build.cu:50-71) is fused into the producer of level `levels`.
 *
 * Arithmetic follows the reference kernel by kernel so that planes come out bit-identical:
 *   level 0 of octave 0, horizontal (s_pyramid_build_ra.cu:40-54): taps from the outside in,
 *       fma((l+r), g, acc), centre last, then *255;
 *   other levels, horizontal (s_pyramid_build_aa.cu:29-49): centre first, then fma((l+r), g, acc)
 *       from the outside in;
 *   vertical (s_pyramid_build_aa.cu:71-86): from the outside in, one fma per side, centre last.
 * Taps beyond the span are zero, and fma(v, 0, acc) == acc, so iterating a padded, compile-time
 * number of taps (HALO = 4*HQ) changes nothing.
 *
 * The input image is sampled through the texture unit exactly like the reference
 * (normalized coordinates, bilinear, u8 -> unorm float, s_image.cu:140-169), because its
 * 16-bit fixed-point blend is not reproducible with fp32 arithmetic; but each upscaled pixel
 * is fetched ONCE into shared memory instead of once per tap (measured on B200: the tap
 * coordinate read_x -/+ k/W hits bit-identical texels to pixel x -/+ k, tools/texprobe.cu).
 */
#include "common.h"
#include "kernels.h"
#include <cstdlib>
#include <algorithm>

namespace psift {

bool pdl_enabled()
{
    static int v = -1;
    if( v < 0 ) v = getenv( "PSIFT_NO_PDL" ) ? 0 : 1;
    return v != 0;
}

int sm_count()
{
    static int n[16] = { 0 };
    int dev = 0; cudaGetDevice( &dev );
    if( dev < 0 || dev >= 16 ) { int v = 148; cudaDeviceGetAttribute( &v, cudaDevAttrMultiProcessorCount, dev ); return v; }
    if( !n[dev] ) { int v = 148; if( cudaDeviceGetAttribute( &v, cudaDevAttrMultiProcessorCount, dev ) != cudaSuccess ) { cudaGetLastError(); v = 148; } n[dev] = v; }
    return n[dev];
}

constexpr int TX = 128;      /* tile width  (outputs) */
constexpr int TY = 32;       /* tile height (outputs) */
constexpr int VY = 8;        /* consecutive rows per thread in the vertical pass */
constexpr int NT = 256;

struct BlurArgs {
    const float*        src;      /* level l-1 (ignored when FROM_TEX) */
    cudaTextureObject_t tex;      /* input image (FROM_TEX only) */
    float*              dst;
    float*              ds_dst;   /* level 0 of the next octave, or nullptr */
    int   w, h, pitch;            /* h: height of the whole plane (normalised texture coordinates) */
    int   r_lo, r_hi;             /* rows [r_lo, r_hi) of the source may be read: taps beyond are clamped to them (the plane's
                                     true border, or - inside a band - further than any tap reaches) */
    int   y_lo, y_hi;             /* rows [y_lo, y_hi) are computed here (the whole plane unless banded; within [r_lo, r_hi));
                                     src/dst/ds_dst point at the plane's row 0 even when that row is not stored */
    int   ds_pitch;
    float shift;                  /* FROM_TEX: texel shift, s_pyramid_build.cu:110-114 */
    Taps  th, tv;
};

template<int HQ, bool FROM_TEX>
__global__ void __launch_bounds__( NT )
k_blur_level( const __grid_constant__ BlurArgs a )
{
    constexpr int HALO = 4 * HQ;
    constexpr int SW   = TX + 2 * HALO;      /* source tile width  */
    constexpr int SH   = TY + 2 * HALO;      /* source tile height */
    extern __shared__ __align__( 16 ) float smem[];
    float* S = smem;                          /* [SH][SW]  source            */
    float* I = smem + SH * SW;                /* [SH][TX]  horizontal result */

    const int tid = threadIdx.x;
    const int tx0 = blockIdx.x * TX;
    const int ty0 = a.y_lo + blockIdx.y * TY;
    const int w = a.w, h = a.h;
    /* all SH rows are staged and filtered, including the <= 3 padding rows whose taps are zero:
     * 0 * (uninitialised shared memory) could be NaN */
    constexpr int r_lo = 0;
    constexpr int r_hi = SH;

    /* ---- stage the source tile (clamp to edge) ---- */
    if( FROM_TEX ) {
        const float fw = (float)w, fh = (float)h;
        for( int i = tid; i < ( r_hi - r_lo ) * SW; i += NT ) {
            const int r = r_lo + i / SW, c = i % SW;
            /* x is NOT clamped to the octave: the reference's tap coordinate read_x -/+ k/W leaves
             * [0,1) at the borders and is clamped by the texture unit on the SOURCE grid, which is
             * a different texel than "last octave pixel" when the input is downscaled.  y is
             * clamped to the octave: the reference's vertical pass clamps on its row buffer. */
            const int gx = tx0 - HALO + c;
            const int gy = min( max( ty0 - HALO + r, a.r_lo ), a.r_hi - 1 );
            const float rx = __fdiv_rn( (float)gx + a.shift, fw );
            const float ry = __fdiv_rn( (float)gy + a.shift, fh );
            S[r * SW + c] = tex2D<float>( a.tex, rx, ry );
        }
    } else {
        const bool interior = ( tx0 - HALO >= 0 ) && ( tx0 + TX + HALO <= w );
        if( interior ) {
            constexpr int SW4 = SW / 4;
            for( int i = tid; i < ( r_hi - r_lo ) * SW4; i += NT ) {
                const int r = r_lo + i / SW4, c4 = i % SW4;
                const int gy = min( max( ty0 - HALO + r, a.r_lo ), a.r_hi - 1 );
                const float4 v = __ldg( reinterpret_cast<const float4*>( a.src + (size_t)gy * a.pitch + ( tx0 - HALO ) ) + c4 );
                reinterpret_cast<float4*>( S + r * SW )[c4] = v;
            }
        } else {
            for( int i = tid; i < ( r_hi - r_lo ) * SW; i += NT ) {
                const int r = r_lo + i / SW, c = i % SW;
                const int gx = min( max( tx0 - HALO + c, 0 ), w - 1 );
                const int gy = min( max( ty0 - HALO + r, a.r_lo ), a.r_hi - 1 );
                S[r * SW + c] = __ldg( a.src + (size_t)gy * a.pitch + gx );
            }
        }
    }
    __syncthreads();

    /* ---- horizontal pass: 4 consecutive outputs per thread out of a register window ---- */
    for( int i = tid; i < ( r_hi - r_lo ) * ( TX / 4 ); i += NT ) {
        const int r  = r_lo + i / ( TX / 4 );
        const int x0 = ( i % ( TX / 4 ) ) * 4;
        float win[4 + 2 * HALO];
        const float4* p = reinterpret_cast<const float4*>( S + r * SW + x0 );
        #pragma unroll
        for( int k = 0; k < 1 + 2 * HQ; k++ ) {
            const float4 v = p[k];
            win[4 * k] = v.x; win[4 * k + 1] = v.y; win[4 * k + 2] = v.z; win[4 * k + 3] = v.w;
        }
        float o[4];
        #pragma unroll
        for( int j = 0; j < 4; j++ ) {
            float acc;
            if( FROM_TEX ) {
                acc = 0.0f;
                #pragma unroll
                for( int off = HALO; off > 0; off-- )
                    acc = __fmaf_rn( __fadd_rn( win[HALO + j - off], win[HALO + j + off] ), a.th.g[off], acc );
                acc = __fmaf_rn( win[HALO + j], a.th.g[0], acc );
                acc = __fmul_rn( acc, 255.0f );
            } else {
                acc = __fmaf_rn( win[HALO + j], a.th.g[0], 0.0f );
                #pragma unroll
                for( int off = HALO; off > 0; off-- )
                    acc = __fmaf_rn( __fadd_rn( win[HALO + j - off], win[HALO + j + off] ), a.th.g[off], acc );
            }
            o[j] = acc;
        }
        *reinterpret_cast<float4*>( I + r * TX + x0 ) = make_float4( o[0], o[1], o[2], o[3] );
    }
    __syncthreads();

    /* ---- vertical pass: VY consecutive rows per thread ---- */
    for( int i = tid; i < TX * ( TY / VY ); i += NT ) {
        const int c  = i % TX;
        const int y0 = ( i / TX ) * VY;
        const int gx = tx0 + c;
        float win[VY + 2 * HALO];
        #pragma unroll
        for( int k = 0; k < VY + 2 * HALO; k++ ) win[k] = I[( y0 + k ) * TX + c];
        #pragma unroll
        for( int j = 0; j < VY; j++ ) {
            float acc = 0.0f;
            #pragma unroll
            for( int off = HALO; off > 0; off-- ) {
                acc = __fmaf_rn( win[HALO + j - off], a.tv.g[off], acc );
                acc = __fmaf_rn( win[HALO + j + off], a.tv.g[off], acc );
            }
            acc = __fmaf_rn( win[HALO + j], a.tv.g[0], acc );
            const int gy = ty0 + y0 + j;
            if( gx < w && gy < a.y_hi ) {
                a.dst[(size_t)gy * a.pitch + gx] = acc;
                if( a.ds_dst && !( ( gx | gy ) & 1 ) )
                    a.ds_dst[(size_t)( gy >> 1 ) * a.ds_pitch + ( gx >> 1 )] = acc;
            }
        }
    }
}


/* ------------------------------------------------------------------------------------------
 * v2: exact tap count (template on H = span-1), packed fp32x2 arithmetic (FFMA2/FADD2, new on
 * sm_100) and a 128x64 tile.  The horizontal pass works on PAIRS OF ROWS: the staged source is
 * stored row-pair interleaved ((r,c),(r+1,c) adjacent), so one 64-bit register pair holds the same
 * column of two rows and every tap is one FADD2 + one FFMA2 for two outputs.  The vertical pass
 * works on PAIRS OF COLUMNS out of the row-major intermediate.  Per-element rounding of the packed
 * instructions is IEEE round-to-nearest, identical to the scalar FADD/FFMA sequence above, and the
 * tap order is unchanged, so the planes stay bit-identical (tests/test_parity_gpu.py).
 * ------------------------------------------------------------------------------------------ */
/* Two tile shapes: 128x32 for planes that fill the GPU several times over, 64x32 for the deeper
 * octaves, where more and shorter blocks cut the latency of each (dependent) level.  Measured on
 * B200 (profiles/): 128x64 tiles have less halo but only two blocks fit an SM, and the load /
 * horizontal / vertical phases of so few blocks do not overlap enough.  Two persistent variants
 * were built and measured as well - cp.async double buffering of a row-major tile (scalar
 * horizontal pass), and prefetching the next tile into registers - and both lost to plain
 * one-tile blocks (25.5 and 26.8 us against 23.3 us for the 11-tap level of a 3840x2160 plane):
 * the hardware block scheduler already overlaps the phases of three resident blocks. */
struct TileMed { static constexpr int TXv = 128, TYv = 32, VYv = 8, UNR = 2; static constexpr int minb( int ) { return 3; } };
struct TileSm  { static constexpr int TXv = 64,  TYv = 32, VYv = 4, UNR = 2; static constexpr int minb( int H ) { return H > 10 ? 3 : 4; } };

__device__ __forceinline__ float2 ffma2( float2 a, float b, float2 c )
{
    float2 d;
    asm( "{ .reg .b64 ra, rb, rc, rd;\n"
         "  mov.b64 ra, {%2, %3};\n  mov.b64 rb, {%4, %4};\n  mov.b64 rc, {%5, %6};\n"
         "  fma.rn.f32x2 rd, ra, rb, rc;\n  mov.b64 {%0, %1}, rd; }"
         : "=f"( d.x ), "=f"( d.y ) : "f"( a.x ), "f"( a.y ), "f"( b ), "f"( c.x ), "f"( c.y ) );
    return d;
}
__device__ __forceinline__ float2 fadd2( float2 a, float2 b )
{
    float2 d;
    asm( "{ .reg .b64 ra, rb, rd;\n"
         "  mov.b64 ra, {%2, %3};\n  mov.b64 rb, {%4, %5};\n"
         "  add.rn.f32x2 rd, ra, rb;\n  mov.b64 {%0, %1}, rd; }"
         : "=f"( d.x ), "=f"( d.y ) : "f"( a.x ), "f"( a.y ), "f"( b.x ), "f"( b.y ) );
    return d;
}
__device__ __forceinline__ float2 fmul2( float2 a, float b )
{
    float2 d;
    asm( "{ .reg .b64 ra, rb, rd;\n"
         "  mov.b64 ra, {%2, %3};\n  mov.b64 rb, {%4, %4};\n"
         "  mul.rn.f32x2 rd, ra, rb;\n  mov.b64 {%0, %1}, rd; }"
         : "=f"( d.x ), "=f"( d.y ) : "f"( a.x ), "f"( a.y ), "f"( b ) );
    return d;
}

template<int H, bool FROM_TEX, class T>
__global__ void __launch_bounds__( NT, T::minb( H ) )
k_blur2( const __grid_constant__ BlurArgs a )
{
    constexpr int TXv = T::TXv, TYv = T::TYv, VYv = T::VYv;
    constexpr int HX  = ( H + 3 ) / 4 * 4;          /* x halo, padded so that staged rows stay 16-byte aligned */
    constexpr int SW  = TXv + 2 * HX;                /* staged width */
    constexpr int SQ  = SW / 4;                      /* 16-byte chunks (2 columns x 2 rows) per half row */
    constexpr int RP  = TYv / 2 + H;                 /* staged row pairs */
    constexpr int D   = HX - H;
    constexpr int W0  = D & ~1;                      /* first window element (even: whole chunks) */
    constexpr int E0  = D & 1;
    constexpr int WN  = ( 4 + 2 * H + E0 + 1 ) & ~1; /* window length in float2, even */
    extern __shared__ __align__( 16 ) float smem[];
    /* S4[p][..]: row pair p as 16-byte chunks {(2p,2c),(2p+1,2c),(2p,2c+1),(2p+1,2c+1)}; the even
     * chunks of a row come first, then the odd ones, so that a warp whose lanes are two chunks
     * apart reads consecutive 16-byte words (no bank conflict). */
    float4* S4 = reinterpret_cast<float4*>( smem );  /* [RP][2*SQ] */
    float*  I  = smem + 2 * RP * SW;                  /* [2*RP][TXv]  horizontal result, row-major */

    const int tid = threadIdx.x;
    const int tx0 = blockIdx.x * TXv, ty0 = a.y_lo + blockIdx.y * TYv;
    const int w = a.w, h = a.h;
    /* programmatic dependent launch: let the next level's blocks be scheduled while this grid runs,
     * and do not touch the previous level's output before that grid has completed and flushed */
    pdl_enter();
    const int nrow = min( TYv, a.y_hi - ty0 );            /* output rows / columns this tile really has */
    const int ncol = min( TXv, w - tx0 );
    const int RPn  = ( nrow + 2 * H + 1 ) >> 1;
    const int ng   = ( ncol + 3 ) >> 2;              /* groups of 4 output columns */
    const int SWn  = 2 * HX + 4 * ng;

    /* ---- stage (clamp to edge) ---- */
    if( FROM_TEX ) {
        /* one fetch per staged pixel.  A warp owns row pairs, its lanes the columns: the normalised
         * x coordinates (one true division each, like the reference's (x + shift) / w) are computed
         * once per thread and reused for every row pair, the y coordinates once per row pair. */
        float2* S2 = reinterpret_cast<float2*>( smem );
        const float fw = (float)w, fh = (float)h;
        constexpr int NC = ( SW + 31 ) / 32;
        const int lane = tid & 31, wid = tid >> 5;
        float rx[NC];
        #pragma unroll
        for( int k = 0; k < NC; k++ ) rx[k] = __fdiv_rn( (float)( tx0 - HX + lane + 32 * k ) + a.shift, fw );   /* x unclamped: see k_blur_level */
        for( int p = wid; p < RPn; p += NT / 32 ) {
            const int g0 = min( max( ty0 - H + 2 * p, a.r_lo ), a.r_hi - 1 );
            const int g1 = min( max( ty0 - H + 2 * p + 1, a.r_lo ), a.r_hi - 1 );
            const float ry0 = __fdiv_rn( (float)g0 + a.shift, fh ), ry1 = __fdiv_rn( (float)g1 + a.shift, fh );
            #pragma unroll
            for( int k = 0; k < NC; k++ ) {
                const int c = lane + 32 * k, ch = c >> 1;
                if( c < SWn )
                    S2[( p * 2 * SQ + ( ch & 1 ) * SQ + ( ch >> 1 ) ) * 2 + ( c & 1 )] = make_float2( tex2D<float>( a.tex, rx[k], ry0 ), tex2D<float>( a.tex, rx[k], ry1 ) );
            }
        }
    } else if( tx0 - HX < 0 || tx0 + TXv + HX > w ) {
        float2* S2 = reinterpret_cast<float2*>( smem );
        /* four independent loads in flight per thread before the first store */
        for( int i0 = tid; i0 < RPn * SW; i0 += 4 * NT ) {
            float2 v[4];
            #pragma unroll
            for( int u = 0; u < 4; u++ ) {
                const int i = min( i0 + u * NT, RPn * SW - 1 );
                const int p = i / SW, c = i % SW;
                const int g0 = min( max( ty0 - H + 2 * p, a.r_lo ), a.r_hi - 1 );
                const int g1 = min( max( ty0 - H + 2 * p + 1, a.r_lo ), a.r_hi - 1 );
                const int gx = min( max( tx0 - HX + c, 0 ), w - 1 );
                v[u] = make_float2( __ldg( a.src + (size_t)g0 * a.pitch + gx ), __ldg( a.src + (size_t)g1 * a.pitch + gx ) );
            }
            #pragma unroll
            for( int u = 0; u < 4; u++ ) {
                const int i = i0 + u * NT;
                const int p = i / SW, c = i % SW, ch = c >> 1;
                if( i < RPn * SW && c < SWn )
                    S2[( p * 2 * SQ + ( ch & 1 ) * SQ + ( ch >> 1 ) ) * 2 + ( c & 1 )] = v[u];
            }
        }
    } else {
        #pragma unroll T::UNR
        for( int i = tid; i < RPn * SQ; i += NT ) {
            const int p = i / SQ, c4 = i % SQ;
            const int g0 = min( max( ty0 - H + 2 * p, a.r_lo ), a.r_hi - 1 );
            const int g1 = min( max( ty0 - H + 2 * p + 1, a.r_lo ), a.r_hi - 1 );
            const float4 A = __ldg( reinterpret_cast<const float4*>( a.src + (size_t)g0 * a.pitch + ( tx0 - HX ) ) + c4 );
            const float4 B = __ldg( reinterpret_cast<const float4*>( a.src + (size_t)g1 * a.pitch + ( tx0 - HX ) ) + c4 );
            S4[p * 2 * SQ + c4]      = make_float4( A.x, B.x, A.y, B.y );
            S4[p * 2 * SQ + SQ + c4] = make_float4( A.z, B.z, A.w, B.w );
        }
    }
    __syncthreads();

    /* ---- horizontal pass: 4 columns x 2 rows per thread ---- */
    for( int i = tid; i < RPn * ( TXv / 4 ); i += NT ) {
        const int p = i / ( TXv / 4 );
        const int g = i % ( TXv / 4 );
        if( g >= ng ) continue;
        float2 win[WN];
        const float4* src = S4 + p * 2 * SQ + g;
        #pragma unroll
        for( int k = 0; k < WN / 2; k++ ) {
            constexpr int C = W0 / 2;
            const float4 v = src[( ( C + k ) & 1 ) * SQ + ( ( C + k ) >> 1 )];
            win[2 * k] = make_float2( v.x, v.y ); win[2 * k + 1] = make_float2( v.z, v.w );
        }
        float2 o[4];
        #pragma unroll
        for( int j = 0; j < 4; j++ ) {
            constexpr int C0 = E0 + H;
            float2 acc;
            if( FROM_TEX ) {
                acc = make_float2( 0.0f, 0.0f );
                #pragma unroll
                for( int off = H; off > 0; off-- ) acc = ffma2( fadd2( win[C0 + j - off], win[C0 + j + off] ), a.th.g[off], acc );
                acc = ffma2( win[C0 + j], a.th.g[0], acc );
                acc = fmul2( acc, 255.0f );
            } else {
                acc = ffma2( win[C0 + j], a.th.g[0], make_float2( 0.0f, 0.0f ) );
                #pragma unroll
                for( int off = H; off > 0; off-- ) acc = ffma2( fadd2( win[C0 + j - off], win[C0 + j + off] ), a.th.g[off], acc );
            }
            o[j] = acc;
        }
        *reinterpret_cast<float4*>( I + ( 2 * p     ) * TXv + 4 * g ) = make_float4( o[0].x, o[1].x, o[2].x, o[3].x );
        *reinterpret_cast<float4*>( I + ( 2 * p + 1 ) * TXv + 4 * g ) = make_float4( o[0].y, o[1].y, o[2].y, o[3].y );
    }
    __syncthreads();

    /* ---- vertical pass: 2 columns x VYv rows per thread ---- */
    for( int i = tid; i < ( TXv / 2 ) * ( TYv / VYv ); i += NT ) {
        const int cp = i % ( TXv / 2 );
        const int y0 = ( i / ( TXv / 2 ) ) * VYv;
        const int gx = tx0 + 2 * cp;
        if( y0 >= nrow || gx >= w ) continue;
        float2 win[VYv + 2 * H];
        #pragma unroll
        for( int k = 0; k < VYv + 2 * H; k++ ) win[k] = *reinterpret_cast<const float2*>( I + ( y0 + k ) * TXv + 2 * cp );
        const int   nval = min( VYv, a.y_hi - ( ty0 + y0 ) );
        const bool  pair = gx + 1 < w;
        float*      row0 = a.dst + (size_t)( ty0 + y0 ) * a.pitch + gx;
        float*      ds0  = a.ds_dst ? a.ds_dst + (size_t)( ( ty0 + y0 ) >> 1 ) * a.ds_pitch + ( gx >> 1 ) : nullptr;
        #pragma unroll
        for( int j = 0; j < VYv; j++ ) {
            float2 acc = make_float2( 0.0f, 0.0f );
            #pragma unroll
            for( int off = H; off > 0; off-- ) {
                acc = ffma2( win[H + j - off], a.tv.g[off], acc );
                acc = ffma2( win[H + j + off], a.tv.g[off], acc );
            }
            acc = ffma2( win[H + j], a.tv.g[0], acc );
            /* ty0 and y0 are even, so row parity is j's: the downscale rows are known at compile time */
            if( j < nval ) {
                float* row = row0 + (size_t)j * a.pitch;
                if( pair ) *reinterpret_cast<float2*>( row ) = acc; else row[0] = acc.x;
                if( ( j & 1 ) == 0 && ds0 ) ds0[(size_t)( j >> 1 ) * a.ds_pitch] = acc.x;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * v3, "marching" blur for the large planes.  The tile kernels above recompute (and re-read) 2H halo
 * rows per 32-row tile - 1.3x (11 taps) to 1.8x (27 taps) of the horizontal pass - and pay two block
 * barriers plus a window of 2H+8 shared-memory reads per 8 outputs in the vertical pass.  Here ONE WARP
 * owns a strip of MW columns and walks down a segment of rows, RS rows per step:
 *   1. the RS source rows of the step two ahead are requested (4-byte cp.async straight into the row-pair
 *      interleaved layout of a ring of staging buffers: L2 latency hides behind two steps of arithmetic);
 *   2. horizontal pass, RS columns x 2 rows per lane out of a register window, result to shared memory;
 *   3. vertical pass: every lane keeps the last 2H+RS horizontal results of ITS two columns in
 *      registers (a sliding window), reads only the RS new ones, and emits RS output rows.
 * So every source value is read from global memory once per strip (halo columns aside), every
 * horizontal result is computed once and read from shared memory once, and there is no block barrier
 * at all (__syncwarp only; the warps of a block drift apart and overlap their phases).  Vertical
 * redundancy is 2H rows per SEGMENT (hundreds of rows), not per tile.  Same arithmetic, same order,
 * same packed instructions as k_blur2: planes stay bit-identical.
 *
 * Measured (B200, a 3840x2160 level, ncu with flushed caches): 19.8 ... 32.6 us for 11 ... 27 taps against 23.0 ...
 * 43.1 us for k_blur2.  What bounds it (profiles/r01d_ncu_blur_march.txt): a step of the 11-tap level is 176 packed
 * FFMA2/FADD2 (two issue cycles each) plus about 250 other instructions for 512 pixels, so the issue slots, not HBM
 * (66 MB in 20 us = 3.3 TB/s), set the time; more warps per SM do not help (8, 12 and 16 were measured equal, also with
 * the registers capped for 3 blocks per SM), and neither did taking 15 % of the non-FMA instructions out of the common
 * step (the unclamped staging / plain-store paths below).  Over the five tap counts the time fits
 * 12.5 us + 2.8 cycles per packed instruction and scheduler: the packed FMA pipe is the slope, the constant is next.
 * ------------------------------------------------------------------------------------------ */
constexpr int MW     = 64;       /* strip width (columns per warp) */
constexpr int MWARPS = 4;        /* warps per block; they do not cooperate */

struct MarchArgs {
    BlurArgs b;
    int nstrip, nseg, seg_rows;  /* seg_rows: even */
};

__device__ __forceinline__ void cp_async4( void* smem_dst, const void* gsrc )
{
    asm volatile( "cp.async.ca.shared.global [%0], [%1], 4;" :: "r"( (unsigned)__cvta_generic_to_shared( smem_dst ) ), "l"( gsrc ) : "memory" );
}
__device__ __forceinline__ void st_pred_v2( float* p, float2 v, bool ok )
{
    asm volatile( "{ .reg .pred q; setp.ne.b32 q, %0, 0; @q st.global.v2.f32 [%1], {%2, %3}; }" :: "r"( (int)ok ), "l"( p ), "f"( v.x ), "f"( v.y ) : "memory" );
}
__device__ __forceinline__ void st_pred( float* p, float v, bool ok )
{
    asm volatile( "{ .reg .pred q; setp.ne.b32 q, %0, 0; @q st.global.f32 [%1], %2; }" :: "r"( (int)ok ), "l"( p ), "f"( v ) : "memory" );
}
__device__ __forceinline__ void cp_async_commit() { asm volatile( "cp.async.commit_group;" ::: "memory" ); }
template<int N> __device__ __forceinline__ void cp_async_wait() { asm volatile( "cp.async.wait_group %0;" :: "n"( N ) : "memory" ); }

template<int H, int RS>
struct MarchGeo {
    static constexpr int HX  = ( H + 3 ) / 4 * 4;        /* x halo, padded to whole 16-byte words */
    static constexpr int SW  = MW + 2 * HX;               /* staged columns */
    static constexpr int NRP = RS / 2;                    /* row pairs per step */
    static constexpr int M   = NRP;                       /* 2-column chunks are dealt round-robin into M arrays: lanes that are
                                                             M chunks apart in the horizontal pass then read consecutive words */
    static constexpr int LPR = 32 / NRP;                  /* lanes per row pair in the horizontal pass */
    static constexpr int CPL = MW / LPR;                  /* columns per lane there (== RS) */
    static constexpr int ASm = ( SW / 2 + M - 1 ) / M;
    static constexpr int AS  = ASm + ( ( 8 / M - ASm % 8 ) + 8 ) % 8;   /* array stride in 16-byte words, == 8/M (mod 8): the M arrays
                                                             then land on disjoint banks for the staging writes of a half warp */
    static constexpr int RPB  = M * AS;                   /* 16-byte words per row pair */
    static constexpr int SBUF = NRP * RPB;                /* ... per staging buffer */
    static constexpr int PD   = RS >= 8 ? 2 : 3;          /* source rows are requested PD steps ahead ... */
    static constexpr int NB   = PD + 1;                   /* ... into a ring of NB staging buffers */
    static constexpr int NK   = ( SW + 15 ) / 16;         /* 16-column groups per staged row */
    /* horizontal results of a step: row-major for CPL == 4.  For CPL == 8 the two 4-column halves of every lane's 8
     * columns live 48 floats apart in rows of 80 floats: the float4 stores of a quarter warp and the float2 loads of a
     * half warp in the vertical pass then both touch every bank once */
    static constexpr int IW   = CPL == 8 ? 80 : MW;       /* row stride of the horizontal results (floats) */
    static constexpr int IH   = CPL == 8 ? 48 : 32;       /* offset of the second group of 4 columns */
    static constexpr int WARP_BYTES = NB * SBUF * 16 + RS * IW * 4;
};

template<int H, int RS, bool FROM_TEX>
__global__ void __launch_bounds__( 32 * MWARPS, 2 )
k_blur_march( const __grid_constant__ MarchArgs m )
{
    using G = MarchGeo<H, RS>;
    const BlurArgs& a = m.b;
    constexpr int HX = G::HX, SW = G::SW, NRP = G::NRP, M = G::M, LPR = G::LPR, CPL = G::CPL, AS = G::AS, RPB = G::RPB, SBUF = G::SBUF;
    constexpr int PD = G::PD, NB = G::NB, NK = G::NK;
    static_assert( CPL == 4 || CPL == 8, "horizontal pass: 4 or 8 columns per lane" );
    constexpr int D = HX - H, W0 = D & ~1, E0 = D & 1, C = W0 / 2;
    constexpr int WN = ( CPL + 2 * H + E0 + 1 ) & ~1;
    extern __shared__ __align__( 16 ) float smem[];

    pdl_enter();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int wid  = blockIdx.x * MWARPS + wib;
    if( wid >= m.nstrip * m.nseg ) return;
    const int strip = wid % m.nstrip, seg = wid / m.nstrip;
    const int x0 = strip * MW;
    const int r0 = a.y_lo + seg * m.seg_rows;
    const int r1 = min( r0 + m.seg_rows, a.y_hi );
    const int nsteps = ( r1 - r0 + 2 * H + RS - 1 ) / RS;
    const int w = a.w;
    float4* S4 = reinterpret_cast<float4*>( reinterpret_cast<char*>( smem ) + wib * G::WARP_BYTES );   /* [NB][NRP][M][AS]: staged source rows */
    float*  I  = reinterpret_cast<float*>( S4 + NB * SBUF );                                            /* [RS][MW]: horizontal results of this step */
    const bool edge = x0 - HX < 0 || x0 + MW + HX > w;

    /* staging: a half warp takes 16 consecutive columns of ONE row, the other half the same columns of the row below
     * (the two rows of a pair sit in alternate 4-byte words, so the two halves never share a bank) */
    const int hl = lane & 15, hp = lane >> 4;
    const int sbase = ( ( ( hl >> 1 ) % M ) * AS + ( hl >> 1 ) / M ) * 4 + ( hl & 1 ) * 2 + hp;   /* floats, within a row pair; +32/M per group */
    const int gx0 = x0 - HX + hl;
    float rx[FROM_TEX ? NK : 1];
    if( FROM_TEX ) {
        #pragma unroll
        for( int k = 0; k < NK; k++ ) rx[FROM_TEX ? k : 0] = __fdiv_rn( (float)( gx0 + 16 * k ) + a.shift, (float)w );   /* x unclamped: see k_blur_level */
    }
    /* request the RS source rows of step q into ring buffer q % NB.  Texture fetches return into registers:
     * they are parked in tq[][] and stored one step later (store_tex), so that no warp waits on the texture unit */
    float tq[FROM_TEX ? NRP : 1][FROM_TEX ? NK : 1];
    auto request = [&]( int q ) {
        float* sb = reinterpret_cast<float*>( S4 + ( q % NB ) * SBUF ) + sbase;
        const int row_first = r0 - H + q * RS;
        if( !FROM_TEX && !edge && row_first >= a.r_lo && row_first + RS <= a.r_hi ) {
            /* the common step: no clamping anywhere, one row pointer and constant offsets */
            const float* gr = a.src + (long long)( row_first + hp ) * a.pitch + gx0;
            #pragma unroll
            for( int rp = 0; rp < NRP; rp++ ) {
                float* sr = sb + rp * RPB * 4;
                #pragma unroll
                for( int k = 0; k < NK; k++ )
                    if( 16 * k + 15 < SW || hl + 16 * k < SW ) cp_async4( sr + ( 32 / M ) * k, gr + 16 * k );
                gr += 2 * a.pitch;
            }
            return;
        }
        #pragma unroll
        for( int rp = 0; rp < NRP; rp++ ) {
            const int g = min( max( r0 - H + q * RS + 2 * rp + hp, a.r_lo ), a.r_hi - 1 );
            if( FROM_TEX ) {
                const float ry = __fdiv_rn( (float)g + a.shift, (float)a.h );
                #pragma unroll
                for( int k = 0; k < NK; k++ ) tq[FROM_TEX ? rp : 0][FROM_TEX ? k : 0] = tex2D<float>( a.tex, rx[FROM_TEX ? k : 0], ry );
            } else {
                const float* gr = a.src + (long long)g * a.pitch;
                float* sr = sb + rp * RPB * 4;
                if( !edge ) {
                    #pragma unroll
                    for( int k = 0; k < NK; k++ )
                        if( 16 * k + 15 < SW || hl + 16 * k < SW ) cp_async4( sr + ( 32 / M ) * k, gr + gx0 + 16 * k );
                } else {
                    #pragma unroll
                    for( int k = 0; k < NK; k++ )
                        if( 16 * k + 15 < SW || hl + 16 * k < SW ) cp_async4( sr + ( 32 / M ) * k, gr + min( max( gx0 + 16 * k, 0 ), w - 1 ) );
                }
            }
        }
    };
    auto store_tex = [&]( int q ) {
        float* sb = reinterpret_cast<float*>( S4 + ( q % NB ) * SBUF ) + sbase;
        #pragma unroll
        for( int rp = 0; rp < NRP; rp++ ) {
            #pragma unroll
            for( int k = 0; k < NK; k++ )
                if( 16 * k + 15 < SW || hl + 16 * k < SW ) sb[rp * RPB * 4 + ( 32 / M ) * k] = tq[FROM_TEX ? rp : 0][FROM_TEX ? k : 0];
        }
    };

    float2 win[2 * H + RS];
    #pragma unroll
    for( int k = 0; k < 2 * H + RS; k++ ) win[k] = make_float2( 0.0f, 0.0f );
    const int gxo = x0 + 2 * lane;                   /* this lane's output columns gxo, gxo+1 (w is even) */
    /* horizontal pass: lane (p, g) filters columns CPL*g .. CPL*g+CPL-1 of row pair p */
    const int hp_p = lane / LPR, hp_g = lane % LPR;
    constexpr int IW = G::IW, IH = G::IH;
    float* const       i_st = I + 2 * hp_p * IW + 4 * hp_g;
    const float* const i_ld = I + ( CPL == 8 ? ( ( lane & 3 ) >> 1 ) * IH + ( lane >> 2 ) * 4 + ( lane & 1 ) * 2 : 2 * lane );

    #pragma unroll
    for( int q = 0; q < PD; q++ ) {
        if( q < nsteps ) { request( q ); if( FROM_TEX ) store_tex( q ); }
        cp_async_commit();
    }
    for( int s = 0; s < nsteps; s++ ) {
        /* ---- 1. request the rows of step s + PD (into the buffer step s - 1 has finished with), wait for step s ---- */
        if( FROM_TEX ) {
            if( s > 0 && s - 1 + PD < nsteps ) store_tex( s - 1 + PD );     /* fetched during step s - 1 */
            if( s + PD < nsteps ) request( s + PD );
        } else {
            if( s + PD < nsteps ) request( s + PD );
            cp_async_commit();
            cp_async_wait<PD>();
        }
        __syncwarp();

        /* ---- 2. horizontal pass: CPL columns x 2 rows per lane ---- */
        {
            float2 hw[WN];
            const float4* src = S4 + ( s % NB ) * SBUF + hp_p * RPB + hp_g;
            #pragma unroll
            for( int k = 0; k < WN / 2; k++ ) {
                const float4 v = src[( ( C + k ) % M ) * AS + ( C + k ) / M];
                hw[2 * k] = make_float2( v.x, v.y ); hw[2 * k + 1] = make_float2( v.z, v.w );
            }
            constexpr int C0 = E0 + H;
            float2 o[CPL];
            if( FROM_TEX ) {
                #pragma unroll
                for( int j = 0; j < CPL; j++ ) o[j] = make_float2( 0.0f, 0.0f );
                #pragma unroll
                for( int off = H; off > 0; off-- ) {
                    #pragma unroll
                    for( int j = 0; j < CPL; j++ ) o[j] = ffma2( fadd2( hw[C0 + j - off], hw[C0 + j + off] ), a.th.g[off], o[j] );
                }
                #pragma unroll
                for( int j = 0; j < CPL; j++ ) o[j] = fmul2( ffma2( hw[C0 + j], a.th.g[0], o[j] ), 255.0f );
            } else {
                #pragma unroll
                for( int j = 0; j < CPL; j++ ) o[j] = ffma2( hw[C0 + j], a.th.g[0], make_float2( 0.0f, 0.0f ) );
                #pragma unroll
                for( int off = H; off > 0; off-- ) {
                    #pragma unroll
                    for( int j = 0; j < CPL; j++ ) o[j] = ffma2( fadd2( hw[C0 + j - off], hw[C0 + j + off] ), a.th.g[off], o[j] );
                }
            }
            #pragma unroll
            for( int q = 0; q < CPL / 4; q++ ) {
                *reinterpret_cast<float4*>( i_st + IH * q )      = make_float4( o[4 * q].x, o[4 * q + 1].x, o[4 * q + 2].x, o[4 * q + 3].x );
                *reinterpret_cast<float4*>( i_st + IW + IH * q ) = make_float4( o[4 * q].y, o[4 * q + 1].y, o[4 * q + 2].y, o[4 * q + 3].y );
            }
        }
        __syncwarp();

        /* ---- 3. vertical pass: slide the register window down RS rows, emit RS output rows ---- */
        #pragma unroll
        for( int k = 0; k < 2 * H; k++ ) win[k] = win[k + RS];
        #pragma unroll
        for( int k = 0; k < RS; k++ ) win[2 * H + k] = *reinterpret_cast<const float2*>( i_ld + k * IW );
        if( ( s + 1 ) * RS > 2 * H ) {
            /* the RS accumulation chains are independent: written level by level so that they interleave, and the
             * stores are predicated instructions, not branches (a branch per row makes the compiler sink each chain
             * into its own block, where it runs as 2H+1 back-to-back dependent instructions) */
            float2 acc[RS];
            #pragma unroll
            for( int j = 0; j < RS; j++ ) acc[j] = make_float2( 0.0f, 0.0f );
            #pragma unroll
            for( int off = H; off > 0; off-- ) {
                #pragma unroll
                for( int j = 0; j < RS; j++ ) acc[j] = ffma2( win[H + j - off], a.tv.g[off], acc[j] );
                #pragma unroll
                for( int j = 0; j < RS; j++ ) acc[j] = ffma2( win[H + j + off], a.tv.g[off], acc[j] );
            }
            #pragma unroll
            for( int j = 0; j < RS; j++ ) acc[j] = ffma2( win[H + j], a.tv.g[0], acc[j] );
            const int rowb = r0 - 2 * H + s * RS;        /* output row of j = 0; even, like r0 */
            float* q  = a.dst + (long long)rowb * a.pitch + gxo;
            float* qd = a.ds_dst + (long long)( rowb >> 1 ) * a.ds_pitch + ( gxo >> 1 );
            if( rowb >= r0 && rowb + RS <= r1 && x0 + MW <= w ) {
                /* the common step: all RS rows and all 64 columns exist, plain stores off two running pointers */
                #pragma unroll
                for( int j = 0; j < RS; j++ ) {
                    *reinterpret_cast<float2*>( q ) = acc[j];
                    q += a.pitch;
                }
                if( a.ds_dst ) {
                    #pragma unroll
                    for( int j = 0; j < RS; j += 2 ) { *qd = acc[j].x; qd += a.ds_pitch; }
                }
            } else {
                const bool okx = gxo < w, okd = okx && a.ds_dst != nullptr;
                #pragma unroll
                for( int j = 0; j < RS; j++ ) {
                    const bool ok = (unsigned)( rowb + j - r0 ) < (unsigned)( r1 - r0 );
                    st_pred_v2( q + (long long)j * a.pitch, acc[j], ok && okx );
                    if( ( j & 1 ) == 0 ) st_pred( qd + (long long)( j >> 1 ) * a.ds_pitch, acc[j].x, ok && okd );
                }
            }
        }
    }
}

static int env_int( const char* name, int dflt )
{
    const char* e = getenv( name );
    return e ? atoi( e ) : dflt;
}
static long march_min_px()   { static long v = -1; if( v < 0 ) v = env_int( "PSIFT_MARCH", 1 ) ? (long)env_int( "PSIFT_MARCH_MIN", 2000000 ) : 0x7fffffffffffL; return v; }
static int  march_wps()      { static int v = -1; if( v < 0 ) v = env_int( "PSIFT_MARCH_WPS", 8 ); return v; }
static bool march_tex()      { static int v = -1; if( v < 0 ) v = env_int( "PSIFT_MARCH_TEX", 1 ); return v != 0; }


template<int H, int RS, bool FROM_TEX>
static cudaError_t launch_march_rs( const BlurArgs& a, cudaStream_t st )
{
    using G = MarchGeo<H, RS>;
    static int n_sm[16] = { 0 };
    int dev = 0; cudaGetDevice( &dev );
    int sms = dev < 16 ? n_sm[dev] : 0;
    const size_t smem = (size_t)MWARPS * G::WARP_BYTES;
    if( !sms ) {
        cudaError_t e = cudaFuncSetAttribute( k_blur_march<H, RS, FROM_TEX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem );
        if( e != cudaSuccess ) return e;
        cudaDeviceGetAttribute( &sms, cudaDevAttrMultiProcessorCount, dev );
        if( dev < 16 ) n_sm[dev] = sms;
    }
    MarchArgs m; m.b = a;
    const int rows = a.y_hi - a.y_lo;
    m.nstrip = ( a.w + MW - 1 ) / MW;
    /* segments per strip: about march_wps() warps per SM, picked among the neighbouring counts for the fewest steps of
     * the slowest SM (steps of a segment, RS rows each, 2H of warm-up) x (blocks on the fullest SM).  Segments start on
     * even rows (the fused downscale takes even rows). */
    /* planes under 3 Mpx (octave 1 of a 1080p frame) get fewer, longer segments: alone on the GPU that is no faster than the
     * tile kernel (too few warps), with other frames' kernels filling the SMs it is 2 % more frames per second */
    static int wps_small = -1; if( wps_small < 0 ) wps_small = env_int( "PSIFT_MARCH_WPS_SMALL", 6 );
    const int target = std::max( 1, sms * ( (long)a.w * rows < 3000000L ? wps_small : march_wps() ) / m.nstrip );
    long best = -1;
    for( int n = std::max( 1, target - 3 ); n <= target + 3; n++ ) {
        const int sr = std::max( 32, ( ( rows + n - 1 ) / n + 1 ) / 2 * 2 );
        const int ns = ( rows + sr - 1 ) / sr;
        const int bl = ( m.nstrip * ns + MWARPS - 1 ) / MWARPS;
        const long cost = (long)( ( sr + 2 * H + RS - 1 ) / RS ) * ( ( bl + sms - 1 ) / sms );
        if( best < 0 || cost < best ) { best = cost; m.seg_rows = sr; m.nseg = ns; }
    }
    const int blocks = ( m.nstrip * m.nseg + MWARPS - 1 ) / MWARPS;
    return launch_pdl( k_blur_march<H, RS, FROM_TEX>, dim3( blocks ), dim3( 32 * MWARPS ), smem, st, m );
}

template<int H, bool FROM_TEX>
static cudaError_t launch_march( const BlurArgs& a, cudaStream_t st )
{
    /* 8 rows per step: 8 columns x 2 rows per lane in the horizontal pass (4 rows / 4 columns was measured too: the
     * same time at twice the instructions per pixel for staging, barriers and addressing) */
    return launch_march_rs<H, 8, FROM_TEX>( a, st );
}

static long blur_small_max()
{
    static long v = -1;
    if( v < 0 ) { const char* e = getenv( "PSIFT_BLUR_SMALL_MAX" ); v = e ? atol( e ) : 3000000L; }
    return v;
}

template<int H, bool FROM_TEX, class T>
static cudaError_t launch_v2t( const BlurArgs& a, cudaStream_t st )
{
    constexpr int HX = ( H + 3 ) / 4 * 4;
    constexpr int RP = T::TYv / 2 + H;
    const size_t smem = sizeof(float) * ( (size_t)2 * RP * ( T::TXv + 2 * HX ) + (size_t)2 * RP * T::TXv );
    static bool attr_done[16] = { false };
    int dev = 0; cudaGetDevice( &dev );
    if( dev >= 16 || !attr_done[dev] ) {
        cudaError_t e = cudaFuncSetAttribute( k_blur2<H, FROM_TEX, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem );
        if( e != cudaSuccess ) return e;
        if( dev < 16 ) attr_done[dev] = true;
    }
    dim3 grid( ( a.w + T::TXv - 1 ) / T::TXv, ( a.y_hi - a.y_lo + T::TYv - 1 ) / T::TYv );
    return launch_pdl( k_blur2<H, FROM_TEX, T>, grid, dim3( NT ), smem, st, a );
}

template<int H, bool FROM_TEX>
static cudaError_t launch_v2( const BlurArgs& a, cudaStream_t st )
{
    if( (long)a.w * ( a.y_hi - a.y_lo ) >= march_min_px() && !( a.y_lo & 1 ) && !( a.w & 1 ) && ( !FROM_TEX || march_tex() ) ) return launch_march<H, FROM_TEX>( a, st );
    if( (long)a.w * ( a.y_hi - a.y_lo ) <= blur_small_max() ) return launch_v2t<H, FROM_TEX, TileSm>( a, st );
    return launch_v2t<H, FROM_TEX, TileMed>( a, st );
}

template<bool FROM_TEX>
static cudaError_t launch_blur_v2( const BlurArgs& a, cudaStream_t st, bool& handled )
{
    handled = true;
    /* one template parameter serves both passes; a config whose two tables differ in length
     * (never the case for the built-in Gauss modes) takes the generic kernel */
    if( a.th.span != a.tv.span ) { handled = false; return cudaSuccess; }
    switch( a.tv.span - 1 ) {
    case 1:  return launch_v2<1,  FROM_TEX>( a, st );
    case 2:  return launch_v2<2,  FROM_TEX>( a, st );
    case 3:  return launch_v2<3,  FROM_TEX>( a, st );
    case 4:  return launch_v2<4,  FROM_TEX>( a, st );
    case 5:  return launch_v2<5,  FROM_TEX>( a, st );
    case 6:  return launch_v2<6,  FROM_TEX>( a, st );
    case 7:  return launch_v2<7,  FROM_TEX>( a, st );
    case 8:  return launch_v2<8,  FROM_TEX>( a, st );
    case 9:  return launch_v2<9,  FROM_TEX>( a, st );
    case 10: return launch_v2<10, FROM_TEX>( a, st );
    case 11: return launch_v2<11, FROM_TEX>( a, st );
    case 12: return launch_v2<12, FROM_TEX>( a, st );
    case 13: return launch_v2<13, FROM_TEX>( a, st );
    case 14: return launch_v2<14, FROM_TEX>( a, st );
    case 15: return launch_v2<15, FROM_TEX>( a, st );
    case 16: return launch_v2<16, FROM_TEX>( a, st );
    default: handled = false; return cudaSuccess;
    }
}

template<int HQ, bool FROM_TEX>
static cudaError_t launch_hq( const BlurArgs& a, cudaStream_t st )
{
    constexpr int HALO = 4 * HQ;
    const size_t smem = sizeof(float) * ( (size_t)( TY + 2 * HALO ) * ( TX + 2 * HALO ) + (size_t)( TY + 2 * HALO ) * TX );
    static bool attr_done[8] = { false, false, false, false, false, false, false, false };
    int dev = 0; cudaGetDevice( &dev );
    if( dev < 8 && !attr_done[dev] ) {
        cudaError_t e = cudaFuncSetAttribute( k_blur_level<HQ, FROM_TEX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem );
        if( e != cudaSuccess ) return e;
        attr_done[dev] = true;
    } else if( dev >= 8 ) {
        cudaFuncSetAttribute( k_blur_level<HQ, FROM_TEX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem );
    }
    dim3 grid( ( a.w + TX - 1 ) / TX, ( a.y_hi - a.y_lo + TY - 1 ) / TY );
    k_blur_level<HQ, FROM_TEX><<<grid, NT, smem, st>>>( a );
    return cudaGetLastError();
}

template<bool FROM_TEX>
static cudaError_t launch_blur( const BlurArgs& a, cudaStream_t st )
{
    bool handled = false;
    cudaError_t e2 = launch_blur_v2<FROM_TEX>( a, st, handled );
    if( handled ) return e2;
    const int halo = max( a.th.span, a.tv.span ) - 1;
    const int hq   = max( 1, ( halo + 3 ) / 4 );
    switch( hq ) {
    case 1: return launch_hq<1, FROM_TEX>( a, st );
    case 2: return launch_hq<2, FROM_TEX>( a, st );
    case 3: return launch_hq<3, FROM_TEX>( a, st );
    case 4: return launch_hq<4, FROM_TEX>( a, st );
    case 5: return launch_hq<5, FROM_TEX>( a, st );
    case 6: return launch_hq<6, FROM_TEX>( a, st );
    case 7: return launch_hq<7, FROM_TEX>( a, st );
    default: return launch_hq<8, FROM_TEX>( a, st );
    }
}

/* ------------------------------------------------------------------------------------------
 * Gauss mode "relative" (VLFeat_Relative; reference s_pyramid_build_ai.cu:17-90 absoluteSourceInterpolated::horiz /
 * vert, K5): two neighbouring taps a, b are folded into ONE bilinear texture fetch at the fractional position
 * offset + (1 - u), u = a / (a + b), scaled by v = a + b (GaussTable::transformBlurTable, gauss_filter.cu:376-411).
 * The blend is the texture unit's (9-bit weights), so the fetches stay on the texture unit here as well: linear
 * textures over the pitch-linear planes, unnormalised coordinates + 0.5 (common/assist.h:66-81), clamp addressing.
 * The horizontal pass goes to a scratch plane, the vertical pass reads that through its own texture.  Level 0 of
 * octave 0 takes the ordinary horizontal pass over the input image (s_pyramid_build_ra.cu:17-55, one fetch per tap
 * at read_x -+ offset / w exactly as there) and the interpolated vertical pass.  Accumulation as compiled in the
 * reference build (SASS of oracle/_ref/obj/s_pyramid_build_ai.o): val = t(-off) + t(+off); out = fma(val, v, out);
 * centre last.  Not a fast path: one thread per pixel.
 * ------------------------------------------------------------------------------------------ */
struct RelArgs {
    cudaTextureObject_t src;
    float* dst; float* ds_dst;
    int   w, h, pitch, ds_pitch;
    int   span, vertical;
    float shift;
    float f[TAP_ALIGN + 1];
};

__global__ void __launch_bounds__( 256 )
k_rel_pass( const __grid_constant__ RelArgs a )
{
    pdl_enter();
    const int x = blockIdx.x * 64 + ( threadIdx.x & 63 ), y = blockIdx.y * 4 + ( threadIdx.x >> 6 );
    if( x >= a.w || y >= a.h ) return;
    const float xf = (float)x, yf = (float)y;
    float out = 0.0f;
    for( int offset = 1; offset <= a.span; offset += 2 ) {
        const float u   = a.f[offset];
        const float off = __fadd_rn( (float)offset, __fsub_rn( 1.0f, u ) );
        const float v   = a.f[offset + 1];
        float t0, t1;
        if( a.vertical ) {
            t0 = tex2D<float>( a.src, __fadd_rn( xf, 0.5f ), __fadd_rn( __fsub_rn( yf, off ), 0.5f ) );
            t1 = tex2D<float>( a.src, __fadd_rn( xf, 0.5f ), __fadd_rn( __fadd_rn( yf, off ), 0.5f ) );
        } else {
            t0 = tex2D<float>( a.src, __fadd_rn( __fsub_rn( xf, off ), 0.5f ), __fadd_rn( yf, 0.5f ) );
            t1 = tex2D<float>( a.src, __fadd_rn( __fadd_rn( xf, off ), 0.5f ), __fadd_rn( yf, 0.5f ) );
        }
        out = __fmaf_rn( __fadd_rn( t0, t1 ), v, out );
    }
    out = __fmaf_rn( tex2D<float>( a.src, __fadd_rn( xf, 0.5f ), __fadd_rn( yf, 0.5f ) ), a.f[0], out );
    a.dst[(size_t)y * a.pitch + x] = out;
    if( a.ds_dst && !( ( x | y ) & 1 ) ) a.ds_dst[(size_t)( y >> 1 ) * a.ds_pitch + ( x >> 1 )] = out;
}

/* horizontal pass over the input image, one texture fetch per tap (s_pyramid_build_ra.cu:17-55) */
__global__ void __launch_bounds__( 256 )
k_rel_h_input( const __grid_constant__ RelArgs a )
{
    pdl_enter();
    const int x = blockIdx.x * 64 + ( threadIdx.x & 63 ), y = blockIdx.y * 4 + ( threadIdx.x >> 6 );
    if( x >= a.w || y >= a.h ) return;
    const float read_x = __fdiv_rn( __fadd_rn( (float)x, a.shift ), (float)a.w );
    const float read_y = __fdiv_rn( __fadd_rn( (float)y, a.shift ), (float)a.h );
    float out = 0.0f;
    for( int offset = a.span; offset > 0; offset-- ) {
        const float offrel = __fdiv_rn( (float)offset, (float)a.w );
        const float v1 = tex2D<float>( a.src, __fsub_rn( read_x, offrel ), read_y );
        const float v2 = tex2D<float>( a.src, __fadd_rn( read_x, offrel ), read_y );
        out = __fmaf_rn( __fadd_rn( v1, v2 ), a.f[offset], out );
    }
    out = __fmaf_rn( tex2D<float>( a.src, read_x, read_y ), a.f[0], out );
    a.dst[(size_t)y * a.pitch + x] = __fmul_rn( out, 255.0f );
}

/* u / v table of one level (GaussTable::transformBlurTable, gauss_filter.cu:376-411) */
static void rel_table( const Taps& t, RelArgs& a )
{
    int spn = t.span;
    if( !( spn & 1 ) ) spn += 1;
    a.span = spn;
    for( int x = 0; x <= TAP_ALIGN; x++ ) a.f[x] = 0.0f;
    for( int x = 1; x < spn && x + 1 < TAP_ALIGN; x += 2 ) {
        const float ta = t.g[x], tb = t.g[x + 1];
        a.f[x]     = ta / ( ta + tb );
        a.f[x + 1] = ta + tb;
    }
    a.f[0] = t.g[0];
}

static cudaError_t launch_pyramid_relative( const PyrLayout& L, cudaTextureObject_t input_tex, float shift, bool direct_scaling, const GaussTables& T,
                                            const PyrStreams& S, long long* launches )
{
    if( !S.level_tex || !S.scratch_tex || !S.scratch ) return cudaErrorInvalidValue;
    const int ds_level = L.n_lvl - 3;
    cudaStream_t st = S.chain;
    cudaError_t e;
    for( int o = L.oct_begin; o < L.oct_end; o++ ) {
        const dim3 grid( ( L.w[o] + 63 ) / 64, ( L.h[o] + 3 ) / 4 );
        for( int l = ( o == 0 || direct_scaling ? 0 : 1 ); l < L.n_lvl; l++ ) {
            RelArgs a;
            a.w = L.w[o]; a.h = L.h[o]; a.pitch = L.pitch[o];
            a.shift = o == 0 ? shift : 0.5f;                          /* s_pyramid_build.cu:110-114: the scaled shift is octave 0's */
            /* horizontal pass -> scratch */
            a.dst = S.scratch; a.ds_dst = nullptr; a.ds_pitch = 0; a.vertical = 0;
            if( l == 0 ) {
                a.src = input_tex;
                a.span = T.dd[o].span;
                for( int x = 0; x < TAP_ALIGN; x++ ) a.f[x] = T.dd[o].g[x];
                a.f[TAP_ALIGN] = 0.0f;
                e = launch_pdl( k_rel_h_input, grid, dim3( 256 ), 0, st, a );
            } else {
                a.src = S.level_tex[(size_t)o * L.n_lvl + ( l - 1 )];
                rel_table( T.inc[l], a );
                e = launch_pdl( k_rel_pass, grid, dim3( 256 ), 0, st, a );
            }
            if( e != cudaSuccess ) return e;
            /* vertical pass scratch -> level l */
            a.src = S.scratch_tex[o];
            rel_table( T.inc[l], a );
            a.vertical = 1;
            a.dst = L.base[o] + (size_t)l * L.lstride[o];
            if( l == ds_level && o + 1 < L.alloc_end && !direct_scaling ) { a.ds_dst = L.base[o + 1]; a.ds_pitch = L.pitch[o + 1]; }
            if( ( e = launch_pdl( k_rel_pass, grid, dim3( 256 ), 0, st, a ) ) != cudaSuccess ) return e;
            if( launches ) *launches += 2;
        }
    }
    return cudaSuccess;
}

/* Build every Gaussian level of every octave (reference Pyramid::build_pyramid,
 * s_pyramid_build.cu:460-596, default branch :549-577). */

/* ------------------------------------------------------------------------------------------ */
/* Fixed9 / Fixed15 (reference s_pyramid_fixed.cu, K8): every level is filtered in ONE step from a fixed source
 * -- the input image for octave 0 (relativeTexAddress::octave_fixed, :127-186, tables abs_o0, result x255), level 0
 * of the same octave otherwise (absoluteTexAddress::octave_fixed, :66-119, tables abs_oN) -- with exactly S = 4 or 7
 * one-sided taps, VERTICAL pass first (:49-64, :131-147), then the horizontal one (the reference's warp shuffles,
 * :23-43, are just reads of the neighbouring columns' vertical results).  Accumulation order as compiled in the
 * reference build (SASS of oracle/_ref/obj/s_pyramid_fixed.o), both passes:
 *     acc = fma(centre, g0, pair1*g1);  acc = fma(pair_i, g_i, acc)  for i = 2..S;      pair_i = v(-i) + v(+i)
 * and the texture rows are fma(1/h, -+i, ypos).  Not a fast path: a 64x16 tile, vertical results of tile + 2S columns in
 * shared memory. */
struct FixedArgs {
    const float* src; cudaTextureObject_t tex; float* dst; float* ds_dst;
    int w, h, pitch, ds_pitch; float tshift;
    float g[8];
};
constexpr int FXW = 64, FXH = 16;

template<int S, bool FROM_TEX>
__global__ void __launch_bounds__( 256 )
k_fixed_level( const FixedArgs a )
{
    __shared__ float V[FXH][FXW + 2 * S];
    pdl_enter();
    const int x0 = blockIdx.x * FXW, y0 = blockIdx.y * FXH;
    const float mul_w = __frcp_rn( (float)a.w ), mul_h = __frcp_rn( (float)a.h );
    for( int q = threadIdx.x; q < FXH * ( FXW + 2 * S ); q += 256 ) {
        const int r = q / ( FXW + 2 * S ), cc = q - r * ( FXW + 2 * S );
        const int c = x0 - S + cc, y = y0 + r;
        float acc = 0.0f;
        if( y < a.h ) {
            float ctr, pr[S + 1];
            if( FROM_TEX ) {
                const float xpos = __fmul_rn( __fadd_rn( (float)c, a.tshift ), mul_w );
                const float ypos = __fmul_rn( __fadd_rn( (float)y, a.tshift ), mul_h );
                ctr = tex2D<float>( a.tex, xpos, ypos );
#pragma unroll
                for( int i = 1; i <= S; i++ )
                    pr[i] = __fadd_rn( tex2D<float>( a.tex, xpos, __fmaf_rn( mul_h, -(float)i, ypos ) ),
                                       tex2D<float>( a.tex, xpos, __fmaf_rn( mul_h,  (float)i, ypos ) ) );
            } else {
                const int cx = min( max( c, 0 ), a.w - 1 );
                ctr = a.src[(size_t)y * a.pitch + cx];
#pragma unroll
                for( int i = 1; i <= S; i++ )
                    pr[i] = __fadd_rn( a.src[(size_t)max( y - i, 0 ) * a.pitch + cx], a.src[(size_t)min( y + i, a.h - 1 ) * a.pitch + cx] );
            }
            acc = __fmaf_rn( ctr, a.g[0], __fmul_rn( pr[1], a.g[1] ) );
#pragma unroll
            for( int i = 2; i <= S; i++ ) acc = __fmaf_rn( pr[i], a.g[i], acc );
        }
        V[r][cc] = acc;
    }
    __syncthreads();
    const int lx = threadIdx.x & ( FXW - 1 ), x = x0 + lx;
    if( x >= a.w ) return;
#pragma unroll
    for( int k = 0; k < FXH / 4; k++ ) {
        const int r = ( threadIdx.x >> 6 ) + 4 * k, y = y0 + r;
        if( y >= a.h ) break;
        const float* v = &V[r][lx + S];
        float out = __fmaf_rn( v[0], a.g[0], __fmul_rn( __fadd_rn( v[-1], v[1] ), a.g[1] ) );
#pragma unroll
        for( int i = 2; i <= S; i++ ) out = __fmaf_rn( __fadd_rn( v[-i], v[i] ), a.g[i], out );
        if( FROM_TEX ) out = __fmul_rn( out, 255.0f );
        a.dst[(size_t)y * a.pitch + x] = out;
        if( a.ds_dst && !( x & 1 ) && !( y & 1 ) ) a.ds_dst[(size_t)( y >> 1 ) * a.ds_pitch + ( x >> 1 )] = out;
    }
}

template<bool FROM_TEX> static cudaError_t launch_blur( const BlurArgs& a, cudaStream_t st );

template<int S>
static cudaError_t launch_pyramid_fixed( const PyrLayout& L, cudaTextureObject_t input_tex, float up_pow, bool direct_scaling, const GaussTables& T,
                                         cudaStream_t st, long long* launches )
{
    const int ds_level = L.n_lvl - 3;
    for( int o = L.oct_begin; o < L.oct_end; o++ ) {
        if( o > 0 && direct_scaling ) {
            /* ScaleDirect with the fixed modes (s_pyramid_build.cu:478-486): level 0 of the octave comes from the input image
             * (horizontal dd[o], vertical inc[0]) instead of the previous octave; the levels above it as usual */
            BlurArgs b;
            b.w = L.w[o]; b.h = L.h[o]; b.pitch = L.pitch[o]; b.y_lo = L.cmp_lo[o]; b.y_hi = L.cmp_hi[o]; b.r_lo = L.row_lo[o]; b.r_hi = L.row_hi[o];
            b.dst = L.base[o]; b.ds_dst = nullptr; b.ds_pitch = 0; b.shift = 0.5f;
            b.src = nullptr; b.tex = input_tex; b.th = T.dd[o]; b.tv = T.inc[0];
            cudaError_t e = launch_blur<true>( b, st );
            if( e != cudaSuccess ) return e;
            if( launches ) ++*launches;
        }
        for( int l = ( o == 0 ? 0 : 1 ); l < L.n_lvl; l++ ) {
            FixedArgs a;
            a.w = L.w[o]; a.h = L.h[o]; a.pitch = L.pitch[o];
            a.dst = L.base[o] + (size_t)l * L.lstride[o];
            a.ds_dst = nullptr; a.ds_pitch = 0;
            if( l == ds_level && o + 1 < L.alloc_end && !direct_scaling ) { a.ds_dst = L.base[o + 1]; a.ds_pitch = L.pitch[o + 1]; }
            a.tshift = 0.5f * up_pow;                               /* s_pyramid_fixed.cu:219 */
            const Taps& t = ( o == 0 ) ? T.abs_o0[l] : T.abs_oN[l];
            for( int i = 0; i < 8; i++ ) a.g[i] = t.g[i];
            const dim3 grid( ( a.w + FXW - 1 ) / FXW, ( a.h + FXH - 1 ) / FXH );
            cudaError_t e;
            if( o == 0 ) { a.src = nullptr; a.tex = input_tex; e = launch_pdl( k_fixed_level<S, true>, grid, dim3( 256 ), 0, st, a ); }
            else         { a.src = L.base[o]; a.tex = 0;       e = launch_pdl( k_fixed_level<S, false>, grid, dim3( 256 ), 0, st, a ); }
            if( e != cudaSuccess ) return e;
            if( launches ) ++*launches;
        }
    }
    return cudaSuccess;
}

/* one Gaussian level of one octave (default Gauss modes: vlfeat, opencv, vlfeat-direct), onto stream st */
static cudaError_t launch_level( const PyrLayout& L, int o, int l, cudaTextureObject_t input_tex, float shift, bool direct_o0, bool direct_scaling,
                                 const GaussTables& T, cudaStream_t st )
{
    const int ds_level = L.n_lvl - 3;      /* _levels - PREV_LEVEL, s_pyramid_build.cu:229 */
    BlurArgs a;
    a.w = L.w[o]; a.h = L.h[o]; a.pitch = L.pitch[o];
    a.y_lo = L.cmp_lo[o]; a.y_hi = L.cmp_hi[o]; a.r_lo = L.row_lo[o]; a.r_hi = L.row_hi[o];
    a.dst = L.base[o] + (size_t)l * L.lstride[o];
    a.ds_dst = nullptr; a.ds_pitch = 0;
    if( l == ds_level && o + 1 < L.alloc_end && !direct_scaling ) { a.ds_dst = L.base[o + 1]; a.ds_pitch = L.pitch[o + 1]; }
    a.shift = o == 0 ? shift : 0.5f;                          /* ScaleDirect: only octave 0 has the scaled shift, s_pyramid_build.cu:110-114 */
    if( o == 0 && direct_o0 ) {
        a.src = nullptr; a.tex = input_tex;
        a.th = T.abs_o0[l]; a.tv = T.abs_o0[l];
        return launch_blur<true>( a, st );
    }
    if( l == 0 ) {
        /* ScaleDirect (s_pyramid_build.cu:499-516): horizontal pass with the octave's own table dd[o], vertical pass
         * with inc[0] (vert_from_interm( octave, 0, .. ) reads d_gauss.inc level 0 for every octave) */
        a.src = nullptr; a.tex = input_tex;
        a.th = T.dd[o]; a.tv = T.inc[0];
        return launch_blur<true>( a, st );
    }
    a.src = L.base[o] + (size_t)( l - 1 ) * L.lstride[o]; a.tex = 0;
    a.th = T.inc[l]; a.tv = T.inc[l];
    return launch_blur<false>( a, st );
}

/* the same for callers that drive the pyramid level by level (banded extraction with halo exchange, pipeline.cu) */
cudaError_t launch_pyramid_level( const PyrLayout& L, int o, int l, cudaTextureObject_t input_tex, float shift, int gauss_mode,
                                  const GaussTables& T, cudaStream_t st )
{
    if( gauss_mode != PSIFT_GAUSS_VLFEAT_COMPUTE && gauss_mode != PSIFT_GAUSS_OPENCV_COMPUTE && gauss_mode != PSIFT_GAUSS_VLFEAT_RELATIVE_ALL )
        return cudaErrorInvalidValue;
    return launch_level( L, o, l, input_tex, shift, gauss_mode == PSIFT_GAUSS_VLFEAT_RELATIVE_ALL, false, T, st );
}

cudaError_t launch_pyramid( const PyrLayout& L, cudaTextureObject_t input_tex, float shift, float up_pow, int gauss_mode, int scaling_mode,
                            const GaussTables& T, const PyrStreams& S, long long* launches )
{
    const bool direct_scaling = scaling_mode == 1;
    /* VLFeat_Relative_All ("vlfeat-direct"): every level of octave 0 is filtered straight from the
     * input image with the absolute tables (horiz_all + vert_all_abs0, s_pyramid_build.cu:545-548,
     * s_pyramid_build_ra.cu:93-129, s_pyramid_build_aa.cu:133-176); the other octaves are built as usual. */
    if( gauss_mode == PSIFT_GAUSS_VLFEAT_RELATIVE ) return launch_pyramid_relative( L, input_tex, shift, direct_scaling, T, S, launches );
    if( gauss_mode == PSIFT_GAUSS_FIXED9 )  return launch_pyramid_fixed<4>( L, input_tex, up_pow, direct_scaling, T, S.chain, launches );
    if( gauss_mode == PSIFT_GAUSS_FIXED15 ) return launch_pyramid_fixed<7>( L, input_tex, up_pow, direct_scaling, T, S.chain, launches );
    /* with ScaleDirect the reference never reaches its vlfeat-direct branch (s_pyramid_build.cu:499 comes first) */
    const bool direct_o0 = gauss_mode == PSIFT_GAUSS_VLFEAT_RELATIVE_ALL && !direct_scaling;
    const int  ds_level = L.n_lvl - 3;      /* _levels - PREV_LEVEL, s_pyramid_build.cu:229 */
    const bool fork = S.side[0] != nullptr;
    bool side_used[2] = { false, false };
    cudaError_t e;
    for( int o = L.oct_begin; o < L.oct_end; o++ ) {
        /* level 0 of octave 0 comes from the input image, of any other octave from the previous
         * octave's fused downscale (or, for the first octave of a seeded pass, from the caller) */
        for( int l = ( o == 0 || direct_scaling ? 0 : 1 ); l < L.n_lvl; l++ ) {
            cudaStream_t st = S.chain;
            if( fork && l > ds_level ) {
                st = S.side[o & 1];
                if( l == ds_level + 1 ) {
                    if( ( e = cudaEventRecord( S.seeded[o], S.chain ) ) != cudaSuccess ) return e;
                    if( ( e = cudaStreamWaitEvent( st, S.seeded[o], 0 ) ) != cudaSuccess ) return e;
                    side_used[o & 1] = true;
                }
            }
            const bool timed = o == 0 && S.lvl_ev != nullptr;
            if( timed && ( e = cudaEventRecord( S.lvl_ev[2 * l], st ) ) != cudaSuccess ) return e;
            e = launch_level( L, o, l, input_tex, shift, direct_o0, direct_scaling, T, st );
            if( e != cudaSuccess ) return e;
            if( timed && ( e = cudaEventRecord( S.lvl_ev[2 * l + 1], st ) ) != cudaSuccess ) return e;
            if( launches ) ++*launches;
        }
    }
    for( int k = 0; k < 2; k++ )
        if( side_used[k] ) {
            if( ( e = cudaEventRecord( S.side_done[k], S.side[k] ) ) != cudaSuccess ) return e;
            if( ( e = cudaStreamWaitEvent( S.chain, S.side_done[k], 0 ) ) != cudaSuccess ) return e;
        }
    return cudaSuccess;
}

} // namespace psift
