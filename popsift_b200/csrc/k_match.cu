/* Brute-force descriptor matching on device-resident descriptors.
 *
 * Replaces popsift::compute_distance (reference features.cu:184-225: one warp per left descriptor, serial over
 * the right ones, five shuffles per pair).  The result is defined by the reference's arithmetic, which is kept bit
 * for bit so that the integer outputs (best, second best, accept) are identical:
 *
 *   lane k of the reference warp handles elements 4k..4k+3:   p_k = fma(dw,dw, fma(dz,dz, fma(dx,dx, dy*dy)))
 *   (contraction read off the SASS of the reference build), then shuffle-down 16, 8, 4, 2, 1, i.e. lane 0 ends with
 *   the fixed tree  s16[i] = p_i + p_{i+16},  s8[i] = s16[i] + s16[i+8],  ...,  d = s2[0] + s2[1].
 *   Scan over the right descriptors in index order with strict `<` (features.cu:203-216)  ==  the two smallest
 *   under the order (distance, index).   accept = d1 / d2 < 0.8f with an IEEE division (features.cu:222).
 *
 * Here one thread owns a 4 x 2 block of pairs and evaluates the 32 partial sums in bit-reversed order with a
 * pairwise stack, which is that same tree without any shuffle; a 32 x 32 descriptor tile pair lives in shared
 * memory (the right tile padded to 132 floats per row so that the 16 lanes of a half warp read distinct banks).
 */
#include "common.h"
#include "kernels.h"
#include <climits>
#include <mutex>
#include <cstdio>
#include <math_constants.h>

namespace psift {

namespace {

constexpr int M_TL = 32, M_TR = 32, M_NT = 128;
constexpr int M_RS = 132;                       /* padded row of the right tile, floats */

struct Top2 { float v1, v2; int i1, i2; };

__device__ __forceinline__ bool lex_less( float v, int i, float w, int j ) { return v < w || ( v == w && i < j ); }

__device__ __forceinline__ void top2_insert( Top2& t, float v, int i )
{
    if( lex_less( v, i, t.v1, t.i1 ) ) { t.v2 = t.v1; t.i2 = t.i1; t.v1 = v; t.i1 = i; }
    else if( lex_less( v, i, t.v2, t.i2 ) ) { t.v2 = v; t.i2 = i; }
}

__device__ __forceinline__ float chunk_dist( const float4 l, const float4 r )
{
    const float dx = __fsub_rn( l.x, r.x ), dy = __fsub_rn( l.y, r.y ), dz = __fsub_rn( l.z, r.z ), dw = __fsub_rn( l.w, r.w );
    return __fmaf_rn( dw, dw, __fmaf_rn( dz, dz, __fmaf_rn( dx, dx, __fmul_rn( dy, dy ) ) ) );
}

__device__ __forceinline__ int bitrev5( int j )
{
    return ( ( j & 1 ) << 4 ) | ( ( j & 2 ) << 2 ) | ( j & 4 ) | ( ( j & 8 ) >> 2 ) | ( ( j & 16 ) >> 4 );
}

__global__ void __launch_bounds__( M_NT )
k_match( const float* __restrict__ l, int l_len, const float* __restrict__ r, int r_len, int3* __restrict__ out, float2* __restrict__ dist )
{
    __shared__ __align__( 16 ) float Ls[M_TL * 128];
    __shared__ __align__( 16 ) float Rs[M_TR * M_RS];
    const int tid = threadIdx.x;
    const int tr = tid & 15, tl = tid >> 4;                 /* 16 right lanes x 8 left groups */
    const int l0 = blockIdx.x * M_TL;

    for( int q = tid; q < M_TL * 32; q += M_NT ) {           /* left tile: rows past the end are zero, never stored */
        const int d = q >> 5, k = q & 31;
        float4 v = make_float4( 0.f, 0.f, 0.f, 0.f );
        if( l0 + d < l_len ) v = reinterpret_cast<const float4*>( l )[(size_t)( l0 + d ) * 32 + k];
        reinterpret_cast<float4*>( Ls )[q] = v;
    }

    Top2 best[4];
#pragma unroll
    for( int a = 0; a < 4; a++ ) { best[a].v1 = best[a].v2 = CUDART_INF_F; best[a].i1 = best[a].i2 = INT_MAX; }

    for( int r0 = 0; r0 < r_len; r0 += M_TR ) {
        __syncthreads();
        for( int q = tid; q < M_TR * 32; q += M_NT ) {
            const int d = q >> 5, k = q & 31;
            float4 v = make_float4( 0.f, 0.f, 0.f, 0.f );
            if( r0 + d < r_len ) v = reinterpret_cast<const float4*>( r )[(size_t)( r0 + d ) * 32 + k];
            *reinterpret_cast<float4*>( &Rs[d * M_RS + 4 * k] ) = v;
        }
        __syncthreads();

        float st[4][2][6];                                   /* pairwise-summation stacks (registers after unrolling) */
#pragma unroll
        for( int j = 0; j < 32; j++ ) {
            const int k = bitrev5( j );
            float4 lv[4], rv[2];
#pragma unroll
            for( int a = 0; a < 4; a++ ) lv[a] = *reinterpret_cast<const float4*>( &Ls[( tl * 4 + a ) * 128 + 4 * k] );
#pragma unroll
            for( int b = 0; b < 2; b++ ) rv[b] = *reinterpret_cast<const float4*>( &Rs[( tr + 16 * b ) * M_RS + 4 * k] );
            /* stack depth before the push = popcount(j); merges after it = trailing ones of j */
            int sp = __popc( j );
#pragma unroll
            for( int a = 0; a < 4; a++ )
#pragma unroll
                for( int b = 0; b < 2; b++ ) st[a][b][sp] = chunk_dist( lv[a], rv[b] );
            int t = j;
#pragma unroll
            for( int m = 0; m < 5; m++ ) {
                if( t & 1 ) {
#pragma unroll
                    for( int a = 0; a < 4; a++ )
#pragma unroll
                        for( int b = 0; b < 2; b++ ) st[a][b][sp - 1] = __fadd_rn( st[a][b][sp - 1], st[a][b][sp] );
                    sp--; t >>= 1;
                } else break;
            }
        }
#pragma unroll
        for( int b = 0; b < 2; b++ ) {                       /* ascending right index per thread: b, then tiles */
            const int ri = r0 + tr + 16 * b;
            if( ri < r_len ) {
#pragma unroll
                for( int a = 0; a < 4; a++ ) top2_insert( best[a], st[a][b][0], ri );
            }
        }
    }

    /* merge the 16 right lanes of each left group (they are 16 consecutive lanes of a warp) */
#pragma unroll
    for( int a = 0; a < 4; a++ ) {
#pragma unroll
        for( int s = 1; s < 16; s <<= 1 ) {
            const float ov1 = __shfl_xor_sync( 0xffffffffu, best[a].v1, s ), ov2 = __shfl_xor_sync( 0xffffffffu, best[a].v2, s );
            const int   oi1 = __shfl_xor_sync( 0xffffffffu, best[a].i1, s ), oi2 = __shfl_xor_sync( 0xffffffffu, best[a].i2, s );
            top2_insert( best[a], ov1, oi1 );
            top2_insert( best[a], ov2, oi2 );
        }
        const int li = l0 + tl * 4 + a;
        if( tr == 0 && li < l_len ) {
            const Top2 t = best[a];
            const int accept = ( __fdiv_rn( t.v1, t.v2 ) < 0.8f ) ? 1 : 0;        /* features.cu:222 */
            out[li] = make_int3( t.i1 == INT_MAX ? 0 : t.i1, t.i2 == INT_MAX ? 0 : t.i2, accept );
            if( dist ) dist[li] = make_float2( t.v1, t.v2 );
        }
    }
}

/* Feature records in the reference's layout (features.h:22-35: 36 bytes, padding, four Descriptor pointers) for a
 * device-resident result: prep_features with a device descriptor base (sift_pyramid.cu:323-331). */
struct RefFeature { int debug_octave; float xpos, ypos, sigma; int num_ori; float orientation[4]; int pad; float* desc[4]; };
static_assert( sizeof(RefFeature) == 72, "popsift::Feature is 72 bytes" );

__global__ void k_ref_features( const psift_feature* in, int n, float* desc_base, RefFeature* out )
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if( i >= n ) return;
    const psift_feature f = in[i];
    RefFeature o;
    o.debug_octave = f.debug_octave; o.xpos = f.xpos; o.ypos = f.ypos; o.sigma = f.sigma; o.num_ori = f.num_ori; o.pad = 0;
    for( int k = 0; k < 4; k++ ) { o.orientation[k] = f.orientation[k]; o.desc[k] = f.desc_idx[k] >= 0 ? desc_base + (size_t)128 * f.desc_idx[k] : nullptr; }
    out[i] = o;
}

} // namespace

cudaError_t launch_match( const float* l, int l_len, const float* r, int r_len, int* out3, float* dist2, cudaStream_t st )
{
    if( l_len <= 0 ) return cudaSuccess;
    k_match<<<dim3( ( l_len + M_TL - 1 ) / M_TL ), dim3( M_NT ), 0, st>>>( l, l_len, r, r_len, reinterpret_cast<int3*>( out3 ), reinterpret_cast<float2*>( dist2 ) );
    return cudaGetLastError();
}

cudaError_t launch_ref_features( const psift_feature* in, int n, float* desc_base, void* out72, cudaStream_t st )
{
    if( n <= 0 ) return cudaSuccess;
    k_ref_features<<<dim3( ( n + 127 ) / 128 ), dim3( 128 ), 0, st>>>( in, n, desc_base, reinterpret_cast<RefFeature*>( out72 ) );
    return cudaGetLastError();
}

} // namespace psift

/* ---------------------------------------------------------------------------------------------------------------
 * psift_match_dev_noalloc: the matcher for callers that must not allocate (code under CUDA graph capture, serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * psift_match_dev keeps its device-side result buffers (int3 + float2 per left descriptor) in memory it
 * cudaMallocs per call; here they live in ONE file-scope device array that every GPU which loads the library holds
 * until exit: PSIFT_NOALLOC_MATCH_MAX left descriptors x 20 bytes = 1.25 MiB.
 *   - the array's name is only a host-side stand-in in host code: its device address is taken once per device, with
 *     that device current, through cudaGetSymbolAddress, and that pointer is what the kernel gets;
 *   - all calls on a device share the array and it keeps the last call's contents: the rows a call uses are cleared
 *     first, and two calls that overlap on the same GPU (two host threads) race - serialise them;
 *   - nothing here calls cudaMalloc, cudaFree or any allocating routine.
 * ------------------------------------------------------------------------------------------------------------- */
namespace {
constexpr size_t NA_MATCH_BYTES = (size_t)PSIFT_NOALLOC_MATCH_MAX * sizeof(int3);     /* 786 432: a multiple of 256 */
constexpr size_t NA_TOTAL_BYTES = NA_MATCH_BYTES + (size_t)PSIFT_NOALLOC_MATCH_MAX * sizeof(float2);
static_assert( NA_MATCH_BYTES % 256 == 0, "the float2 part must stay 256-byte aligned" );
}
static __device__ __align__( 256 ) unsigned char g_noalloc_match_scratch[NA_TOTAL_BYTES];

namespace {
constexpr int NA_MAX_DEV = 64;
std::mutex     g_na_mu;
unsigned char* g_na_ptr[NA_MAX_DEV] = { nullptr };

/* device address of the scratch array on `device` (which must be current); looked up once per device */
cudaError_t noalloc_scratch( int device, unsigned char** out )
{
    std::lock_guard<std::mutex> lk( g_na_mu );
    if( device >= 0 && device < NA_MAX_DEV && g_na_ptr[device] ) { *out = g_na_ptr[device]; return cudaSuccess; }
    unsigned char* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress( (void**)&p, g_noalloc_match_scratch );
    if( e != cudaSuccess ) return e;
    if( device >= 0 && device < NA_MAX_DEV ) g_na_ptr[device] = p;
    *out = p;
    return cudaSuccess;
}
}

extern "C" int psift_match_dev_noalloc( int device, const float* l_desc_dev, int l_len, const float* r_desc_dev, int r_len,
                                        int* match3_host, float* dist2_host )
{
    using psift::set_error;
    if( l_len < 0 || r_len < 0 || ( l_len > 0 && ( !l_desc_dev || !match3_host ) ) || ( r_len > 0 && !r_desc_dev ) )
        return set_error( PSIFT_E_INVALID, "psift_match_dev_noalloc: bad argument" );
    if( l_len > PSIFT_NOALLOC_MATCH_MAX ) {
        char msg[200];
        snprintf( msg, sizeof msg, "psift_match_dev_noalloc: %d left descriptors exceed the static scratch (PSIFT_NOALLOC_MATCH_MAX = %d); use psift_match_dev",
                  l_len, PSIFT_NOALLOC_MATCH_MAX );
        return set_error( PSIFT_E_INVALID, msg );
    }
    if( l_len == 0 ) return PSIFT_OK;
    cudaError_t e = cudaSetDevice( device );
    unsigned char* base = nullptr;
    if( e == cudaSuccess ) e = noalloc_scratch( device, &base );
    if( e != cudaSuccess ) { cudaGetLastError(); return set_error( PSIFT_E_CUDA, cudaGetErrorString( e ) ); }
    int*   d_m = reinterpret_cast<int*>( base );
    float* d_d = reinterpret_cast<float*>( base + NA_MATCH_BYTES );
    /* the array keeps the previous call's rows: clear what this call uses (the kernel writes every row it owns) */
    e = cudaMemsetAsync( d_m, 0, sizeof(int3) * (size_t)l_len, 0 );
    if( e == cudaSuccess ) e = cudaMemsetAsync( d_d, 0, sizeof(float2) * (size_t)l_len, 0 );
    if( e == cudaSuccess ) e = psift::launch_match( l_desc_dev, l_len, r_desc_dev, r_len, d_m, d_d, 0 );
    if( e == cudaSuccess ) e = cudaMemcpy( match3_host, d_m, sizeof(int) * 3 * (size_t)l_len, cudaMemcpyDeviceToHost );
    if( e == cudaSuccess && dist2_host ) e = cudaMemcpy( dist2_host, d_d, sizeof(float) * 2 * (size_t)l_len, cudaMemcpyDeviceToHost );
    if( e != cudaSuccess ) { cudaGetLastError(); return set_error( PSIFT_E_CUDA, cudaGetErrorString( e ) ); }
    return PSIFT_OK;
}
