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

namespace psift {

constexpr int TX = 128;      /* tile width  (outputs) */
constexpr int TY = 32;       /* tile height (outputs) */
constexpr int VY = 8;        /* consecutive rows per thread in the vertical pass */
constexpr int NT = 256;

struct BlurArgs {
    const float*        src;      /* level l-1 (ignored when FROM_TEX) */
    cudaTextureObject_t tex;      /* input image (FROM_TEX only) */
    float*              dst;
    float*              ds_dst;   /* level 0 of the next octave, or nullptr */
    int   w, h, pitch;
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
    const int ty0 = blockIdx.y * TY;
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
            const int gy = min( max( ty0 - HALO + r, 0 ), h - 1 );
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
                const int gy = min( max( ty0 - HALO + r, 0 ), h - 1 );
                const float4 v = __ldg( reinterpret_cast<const float4*>( a.src + (size_t)gy * a.pitch + ( tx0 - HALO ) ) + c4 );
                reinterpret_cast<float4*>( S + r * SW )[c4] = v;
            }
        } else {
            for( int i = tid; i < ( r_hi - r_lo ) * SW; i += NT ) {
                const int r = r_lo + i / SW, c = i % SW;
                const int gx = min( max( tx0 - HALO + c, 0 ), w - 1 );
                const int gy = min( max( ty0 - HALO + r, 0 ), h - 1 );
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
            if( gx < w && gy < h ) {
                a.dst[(size_t)gy * a.pitch + gx] = acc;
                if( a.ds_dst && !( ( gx | gy ) & 1 ) )
                    a.ds_dst[(size_t)( gy >> 1 ) * a.ds_pitch + ( gx >> 1 )] = acc;
            }
        }
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
    dim3 grid( ( a.w + TX - 1 ) / TX, ( a.h + TY - 1 ) / TY );
    k_blur_level<HQ, FROM_TEX><<<grid, NT, smem, st>>>( a );
    return cudaGetLastError();
}

template<bool FROM_TEX>
static cudaError_t launch_blur( const BlurArgs& a, cudaStream_t st )
{
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

/* Build every Gaussian level of every octave (reference Pyramid::build_pyramid,
 * s_pyramid_build.cu:460-596, default branch :549-577). */
cudaError_t launch_pyramid( const PyrLayout& L, cudaTextureObject_t input_tex, float shift,
                            const GaussTables& T, cudaStream_t st, long long* launches )
{
    const int ds_level = L.n_lvl - 3;      /* _levels - PREV_LEVEL, s_pyramid_build.cu:229 */
    for( int o = 0; o < L.n_oct; o++ ) {
        for( int l = ( o == 0 ? 0 : 1 ); l < L.n_lvl; l++ ) {
            BlurArgs a;
            a.w = L.w[o]; a.h = L.h[o]; a.pitch = L.pitch[o];
            a.dst = L.base[o] + (size_t)l * L.lstride[o];
            a.ds_dst = nullptr; a.ds_pitch = 0;
            if( l == ds_level && o + 1 < L.n_oct ) { a.ds_dst = L.base[o + 1]; a.ds_pitch = L.pitch[o + 1]; }
            a.shift = shift;
            cudaError_t e;
            if( l == 0 ) {
                a.src = nullptr; a.tex = input_tex;
                a.th = T.dd[0]; a.tv = T.inc[0];
                e = launch_blur<true>( a, st );
            } else {
                a.src = L.base[o] + (size_t)( l - 1 ) * L.lstride[o]; a.tex = 0;
                a.th = T.inc[l]; a.tv = T.inc[l];
                e = launch_blur<false>( a, st );
            }
            if( e != cudaSuccess ) return e;
            if( launches ) ++*launches;
        }
    }
    return cudaSuccess;
}

} // namespace psift
