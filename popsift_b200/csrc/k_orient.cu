/* k_orient.cu -- orientation assignment and the orientation prefix sum, for sm_100a.
 *
 * Reference: ori_par (s_orientation.cu:60-242) runs one 32-thread block per keypoint and
 * accumulates the 36-bin histogram with shared-memory float atomics, so its sums depend on
 * the arrival order; ori_prefix_sum (s_orientation.cu:303-345) is a single 32x32 block.
 * Here a 4-WARP BLOCK owns a keypoint (taken from a device-side work queue: no host round
 * trip); the 128 threads stride over the window, every lane accumulates a private histogram
 * column col[warp][bin][lane] in shared memory in a fixed sample order, the columns are
 * summed in a fixed rotated lane order and the 4 warp partials in warp order, which makes
 * the result reproducible run to run.  Sample geometry, weights, bin rounding, the 6 box-filter
 * passes, the parabola fit and the 0.8-of-best rule follow the reference expression by
 * expression (quirks kept: integer-truncated squared distance, s_orientation.cu:123; all
 * four slots filled with bin -1 when the histogram has no peak at all, :215-230).
 */
#include "common.h"
#include "kernels.h"

namespace psift {

__device__ static const float PI_F  = 3.14159265358979323846f;          /* sift_constants.h:27-36 */
__device__ static const float PI2_F = 2.0f * 3.14159265358979323846f;

constexpr int ORI_WARPS = 4;

__global__ void __launch_bounds__( ORI_WARPS * 32 )
k_orientation( const __grid_constant__ PyrLayout L, const Cand* __restrict__ cand, Counters* __restrict__ ct,
               psift_extremum* __restrict__ ext )
{
    __shared__ float col[ORI_WARPS][ORI_NBINS][32];     /* lane-private histogram columns */
    __shared__ float hist[ORI_WARPS][ORI_NBINS];
    __shared__ float smh[ORI_WARPS][ORI_NBINS];
    __shared__ float yv[ORI_WARPS][64];
    __shared__ float ang[ORI_WARPS][64];

    __shared__ float part[ORI_WARPS][ORI_NBINS];
    __shared__ int   next_e;

    pdl_enter();
    const int tid  = threadIdx.x;
    const int lane = tid & 31;
    const int wid  = tid >> 5;
    const int total = ct->ext_total;

    /* a block per keypoint, handed out one at a time (windows differ in size by 5x and more): all
     * eight warps share the window, each lane summing into its own histogram column, and the
     * columns are added up in a fixed order - reproducible, unlike the reference's atomics */
    for( ;; ) {
        if( tid == 0 ) next_e = atomicAdd( &ct->ori_next, 1 );
        __syncthreads();
        const int e = next_e;
        if( e >= total ) break;
        int o = 0;
        while( o + 1 < L.n_oct && e >= ct->ext_ps[o + 1] ) o++;
        const Cand c = cand[e];
        const int  w = L.w[o], h = L.h[o];
        const float x = c.x, y = c.y, sig = c.sigma;
        const int   level = c.lpos;

        #pragma unroll
        for( int b = 0; b < ORI_NBINS; b++ ) col[wid][b][lane] = 0.0f;

        const float sigw   = __fmul_rn( 1.5f, sig );                       /* ORI_WINFACTOR */
        const int   rad    = (int)roundf( __fmul_rn( 3.0f, sigw ) );
        const float factor = __fdividef( -0.5f, __fmul_rn( sigw, sigw ) );
        const int   sq_thres = rad * rad;
        const int xmin = max( 1,     (int)roundf( x ) - rad );
        const int xmax = min( w - 2, (int)roundf( x ) + rad );
        const int ymin = max( 1,     (int)roundf( y ) - rad );
        const int ymax = min( h - 2, (int)roundf( y ) + rad );
        const int wx = xmax - xmin + 1;
        const int hy = ymax - ymin + 1;
        const int loops = ( wx > 0 && hy > 0 ) ? wx * hy : 0;
        const int pitch = L.pitch[o];
        const float* plane = L.base[o] + (size_t)min( max( level, 0 ), L.n_lvl - 1 ) * L.lstride[o];
        const int wxs = max( wx, 1 );
        constexpr int NTH = ORI_WARPS * 32;
        const int q = NTH / wxs, r = NTH - q * wxs;                         /* (row, column) advance per NTH pixels */
        int yy = ymin + tid / wxs, xx = xmin + tid % wxs;

        for( int i = tid; i < loops; i += NTH ) {
            /* 1 <= xx <= w-2 and 1 <= yy <= h-2: the four neighbours are inside the plane */
            const float* p = plane + (size_t)yy * pitch + xx;
            const float gx = __fsub_rn( __ldg( p + 1 ), __ldg( p - 1 ) );
            const float gy = __fsub_rn( __ldg( p + pitch ), __ldg( p - pitch ) );
            const float dx = __fsub_rn( (float)xx, x );
            const float dy = __fsub_rn( (float)yy, y );
            const int sq_dist = (int)__fmaf_rn( dy, dy, __fmul_rn( dx, dx ) );
            if( sq_dist <= sq_thres ) {
                const float grad  = hypotf( gx, gy );
                const float theta = atan2f( gy, gx );
                const float weight = __fmul_rn( grad, expf( __fmul_rn( (float)sq_dist, factor ) ) );
                int bidx = (int)roundf( __fdividef( float( ORI_NBINS ) * ( theta + PI_F ), PI2_F ) );
                bidx = ( bidx >= ORI_NBINS ) ? 0 : bidx;
                col[wid][bidx][lane] += weight;
            }
            yy += q; xx += r; if( xx > xmax ) { xx -= wx; yy++; }
        }
        __syncthreads();
        for( int i = tid; i < ORI_WARPS * ORI_NBINS; i += NTH ) {
            const int ww = i / ORI_NBINS, b = i - ww * ORI_NBINS;
            float s = 0.0f;
            #pragma unroll 8
            for( int l = 0; l < 32; l++ ) s += col[ww][b][( l + b ) & 31];       /* rotated start: no bank conflict, still a fixed order */
            part[ww][b] = s;
        }
        __syncthreads();
        if( wid != 0 ) continue;            /* the rest is a warp's worth of work; the others wait at the next hand-out */
        for( int b = lane; b < ORI_NBINS; b += 32 ) {
            float s = 0.0f;
            #pragma unroll
            for( int ww = 0; ww < ORI_WARPS; ww++ ) s += part[ww][b];
            hist[wid][b] = s;
        }
        __syncwarp();

        /* 3 x (two box passes), s_orientation.cu:143-160 */
        for( int it = 0; it < 3; it++ ) {
            for( int b = lane; b < ORI_NBINS; b += 32 ) {
                const int p = b == 0 ? ORI_NBINS - 1 : b - 1, n = b == ORI_NBINS - 1 ? 0 : b + 1;
                smh[wid][b] = __fdiv_rn( __fadd_rn( __fadd_rn( hist[wid][p], hist[wid][b] ), hist[wid][n] ), 3.0f );
            }
            __syncwarp();
            for( int b = lane; b < ORI_NBINS; b += 32 ) {
                const int p = b == 0 ? ORI_NBINS - 1 : b - 1, n = b == ORI_NBINS - 1 ? 0 : b + 1;
                hist[wid][b] = __fdiv_rn( __fadd_rn( __fadd_rn( smh[wid][p], smh[wid][b] ), smh[wid][n] ), 3.0f );
            }
            __syncwarp();
        }

        /* peaks and parabola fit, s_orientation.cu:180-205 */
        for( int b = lane; b < 64; b += 32 ) {
            const int p = b == 0 ? ORI_NBINS - 1 : b - 1, n = b == ORI_NBINS - 1 ? 0 : b + 1;
            bool pred = false;
            float hp = 0.f, hc = 0.f, hn = 0.f;
            if( b < ORI_NBINS ) {
                hp = hist[wid][p]; hc = hist[wid][b]; hn = hist[wid][n];
                pred = hc > fmaxf( hp, hn );
            }
            const float num  = pred ? 3.0f * hp - 4.0f * hc + 1.0f * hn : 0.0f;
            const float denB = pred ? 2.0f * ( hp - 2.0f * hc + hn ) : 1.0f;
            const float newbin = __fdividef( num, denB );
            pred = pred && newbin >= 0.0f && newbin <= 2.0f;
            ang[wid][b] = pred ? (float)p + newbin : -1.0f;
            yv[wid][b]  = pred ? -( num * num ) / ( 4.0f * denB ) + hp : -INFINITY;
        }
        __syncwarp();

        /* the up to four largest peaks in descending order (the reference sorts all 64 with a
         * bitonic network, common/warp_bitonic_sort.h; ties keep the lower bin first here) */
        float v0 = yv[wid][lane], v1 = yv[wid][lane + 32];
        float best_val = 0.f; int best_idx = 0; float first = 0.f;
        int n_ori = 0;
        float my_ori = 0.0f;
        for( int k = 0; k < MAX_ORI; k++ ) {
            float m = v0; int mi = lane;
            if( v1 > m ) { m = v1; mi = lane + 32; }
            #pragma unroll
            for( int s = 16; s > 0; s >>= 1 ) {
                const float om = __shfl_xor_sync( 0xffffffffu, m, s );
                const int   oi = __shfl_xor_sync( 0xffffffffu, mi, s );
                if( om > m || ( om == m && oi < mi ) ) { m = om; mi = oi; }
            }
            best_val = m; best_idx = mi;
            if( k == 0 ) first = m;
            const float yref = __fmul_rn( 0.8f, first );
            if( !( best_val >= yref ) ) break;
            float chosen = ang[wid][best_idx];
            if( chosen >= (float)ORI_NBINS ) chosen -= (float)ORI_NBINS;
            const float th = fmaf( PI2_F * chosen, 1.0f / ORI_NBINS, -PI_F );
            if( lane == k ) my_ori = th;
            n_ori++;
            if( mi == lane ) v0 = -INFINITY; else if( mi == lane + 32 ) v1 = -INFINITY;
            /* with no peak at all every value is -inf: -inf >= 0.8*-inf holds and the reference
             * writes four orientations from bin -1; removing one -inf leaves the rest, so the
             * loop naturally runs four times */
        }
        __syncwarp();

        psift_extremum* out = ext + e;
        if( lane < MAX_ORI ) out->orientation[lane] = ( lane < n_ori ) ? my_ori : 0.0f;
        if( lane == 0 ) {
            out->xpos = c.x; out->ypos = c.y; out->lpos = c.lpos; out->sigma = c.sigma;
            out->octave = o; out->num_ori = n_ori; out->idx_ori = 0;
        }
    }
}

/* Exclusive prefix sum of num_ori over all keypoints, per-octave descriptor counts, and the
 * descriptor -> keypoint map (s_orientation.cu:303-345; common/excl_blk_prefix_sum.h).  One block. */
constexpr int PS_NT = 1024;
__global__ void __launch_bounds__( PS_NT )
k_ori_prefix( psift_extremum* __restrict__ ext, Counters* __restrict__ ct, int* __restrict__ desc_map, int desc_capacity )
{
    __shared__ int warp_sum[32];
    __shared__ int carry;
    pdl_enter();
    const int total = ct->ext_total;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if( threadIdx.x == 0 ) carry = 0;
    __syncthreads();
    for( int base = 0; base < total; base += PS_NT ) {
        const int i = base + threadIdx.x;
        const int v = ( i < total ) ? ext[i].num_ori : 0;
        int inc = v;
        #pragma unroll
        for( int s = 1; s < 32; s <<= 1 ) {
            const int t = __shfl_up_sync( 0xffffffffu, inc, s );
            if( lane >= s ) inc += t;
        }
        if( lane == 31 ) warp_sum[wid] = inc;
        __syncthreads();
        if( wid == 0 ) {
            int ws = warp_sum[lane];
            #pragma unroll
            for( int s = 1; s < 32; s <<= 1 ) {
                const int t = __shfl_up_sync( 0xffffffffu, ws, s );
                if( lane >= s ) ws += t;
            }
            warp_sum[lane] = ws;
        }
        __syncthreads();
        const int excl = carry + ( wid ? warp_sum[wid - 1] : 0 ) + inc - v;
        if( i < total ) {
            int nv = v;
            if( excl + nv > desc_capacity ) {           /* clamp: keep only what fits */
                nv = max( 0, desc_capacity - excl );
                ext[i].num_ori = nv;
                ct->overflow = 1;
            }
            ext[i].idx_ori = min( excl, desc_capacity );
            for( int k = 0; k < nv; k++ ) desc_map[excl + k] = i;
        }
        __syncthreads();
        if( threadIdx.x == 0 ) carry += warp_sum[31];
        __syncthreads();
    }
    if( threadIdx.x == 0 ) {
        const int ori_total = min( carry, desc_capacity );
        ct->ori_total = ori_total;
        ct->need_desc = carry;
        for( int o = 0; o < MAX_OCT; o++ ) {
            const int fe = ct->ext_ps[o], n = ct->ext_ct[o];
            int lo = ori_total, hi = ori_total;
            if( n > 0 ) {
                lo = ext[fe].idx_ori;
                hi = ( fe + n < total ) ? ext[fe + n].idx_ori : ori_total;
            }
            ct->ori_ps[o] = lo;
            ct->ori_ct[o] = n > 0 ? hi - lo : 0;
        }
        ct->ori_ps[MAX_OCT] = ori_total;
    }
}

cudaError_t launch_orientation( const PyrLayout& L, const Cand* sorted, Counters* ct, psift_extremum* ext,
                                int* desc_map, int desc_capacity, int n_blocks, cudaStream_t st, long long* launches )
{
    launch_pdl( k_orientation, dim3( n_blocks ), dim3( ORI_WARPS * 32 ), 0, st, L, sorted, ct, ext );
    launch_pdl( k_ori_prefix, dim3( 1 ), dim3( PS_NT ), 0, st, ext, ct, desc_map, desc_capacity );
    if( launches ) *launches += 2;
    return cudaGetLastError();
}

} // namespace psift
