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
#include <cstdlib>

namespace psift {

__device__ static const float PI_F  = 3.14159265358979323846f;          /* sift_constants.h:27-36 */
__device__ static const float PI2_F = 2.0f * 3.14159265358979323846f;

#ifndef PSIFT_ORI_WARPS
#define PSIFT_ORI_WARPS 4
#endif
constexpr int ORI_WARPS = PSIFT_ORI_WARPS;
constexpr int ORI_CH    = ORI_WARPS >= 4 ? 1024 : 512;            /* samples per chunk: 32 steps of the reference's 32-thread loop */

/* Histogram accumulation in the REFERENCE'S ORDER.  ori_par (s_orientation.cu:104-138) is one warp: sample i is taken by
 * lane i % 32 in step i / 32 and added with a shared-memory float atomicAdd, which compiles to a compare-and-swap loop
 * (ATOMS.CAST.SPIN in the SASS of the reference build): the lanes of a step that hit the same bin are added one after
 * the other in lane order, the steps one after the other.  A float sum depends on its order, so the order is restated:
 * every sample's (bin, weight) goes to shared memory together with its rank among the samples of its bin in that order
 * (warp w takes a contiguous range of steps; __match_any_sync gives the lanes of a step that share a bin), the weights
 * are scattered into per-bin lists, and thread b adds up list b front to back, continuing from hist[b].  Windows with
 * more than ORI_CH samples take several chunks, in order.  Lane order within a step is ASCENDING: measured on B200
 * against the reference build - 1 180 of 1 180 orientations of eight golden cases bit-equal (1 049 with descending order),
 * and with them every descriptor. */
__global__ void __launch_bounds__( ORI_WARPS * 32 )
k_orientation( const __grid_constant__ PyrLayout L, const Cand* __restrict__ cand, Counters* __restrict__ ct,
               psift_extremum* __restrict__ ext )
{
    __shared__ float          sw[ORI_CH];                /* weight of sample (chunk-local index) */
    __shared__ float          lst[ORI_CH];               /* weights grouped by bin, in accumulation order */
    __shared__ unsigned short sp[ORI_CH];                /* rank of the sample among its warp's samples of the same bin */
    __shared__ unsigned char  sb[ORI_CH];                /* bin, or 63: outside the window's circle */
    __shared__ int   wc[ORI_WARPS][ORI_NBINS];           /* samples per bin in warp w's range of steps */
    __shared__ int   wbase[ORI_WARPS][ORI_NBINS];        /* where warp w's samples of bin b start in lst */
    __shared__ int   bstart[ORI_NBINS + 1];
    __shared__ float hist[ORI_WARPS][ORI_NBINS];
    __shared__ float smh[ORI_WARPS][ORI_NBINS];
    __shared__ float yv[ORI_WARPS][64];
    __shared__ float ang[ORI_WARPS][64];
    __shared__ int   next_e;

    pdl_enter();
    const int tid  = threadIdx.x;
    const int lane = tid & 31;
    const int wid  = tid >> 5;
    const int total = ct->ext_total;
    const unsigned lt_mask = ( 1u << lane ) - 1u;

    /* a block per keypoint, handed out one at a time (windows differ in size by 5x and more) */
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
        const int q32 = 32 / wxs, r32 = 32 - q32 * wxs;                     /* (row, column) advance per step */

        for( int b = tid; b < ORI_NBINS; b += ORI_WARPS * 32 ) hist[0][b] = 0.0f;
        for( int c0 = 0; c0 < loops; c0 += ORI_CH ) {
            const int n_it = min( ORI_CH / 32, ( loops - c0 + 31 ) >> 5 );   /* steps in this chunk */
            const int ipw  = ( n_it + ORI_WARPS - 1 ) / ORI_WARPS;          /* steps per warp: contiguous ranges */
            const int it0  = wid * ipw, it1 = min( n_it, it0 + ipw );
            for( int b = lane; b < ORI_NBINS; b += 32 ) wc[wid][b] = 0;
            __syncwarp();
            int i = c0 + 32 * it0 + lane;
            int yy = ymin + i / wxs, xx = xmin + i % wxs;
            for( int it = it0; it < it1; it++, i += 32 ) {
                int key = 63; float weight = 0.0f;
                if( i < loops ) {
                    /* 1 <= xx <= w-2 and 1 <= yy <= h-2: the four neighbours are inside the plane */
                    const float* p = plane + (size_t)yy * pitch + xx;
                    const float gx = __fsub_rn( __ldg( p + 1 ), __ldg( p - 1 ) );
                    const float gy = __fsub_rn( __ldg( p + pitch ), __ldg( p - pitch ) );
                    const float dx = __fsub_rn( (float)xx, x );
                    const float dy = __fsub_rn( (float)yy, y );
                    const int sq_dist = (int)__fmaf_rn( dy, dy, __fmul_rn( dx, dx ) );
                    if( sq_dist <= sq_thres ) {
                        const float grad  = hypot_as_lib( gx, gy );
                        const float theta = atan2f( gy, gx );
                        weight = __fmul_rn( grad, expf( __fmul_rn( (float)sq_dist, factor ) ) );
                        int bidx = (int)roundf( __fdividef( float( ORI_NBINS ) * ( theta + PI_F ), PI2_F ) );
                        key = ( bidx >= ORI_NBINS ) ? 0 : bidx;
                    }
                }
                const unsigned m = __match_any_sync( 0xffffffffu, key );
                int before = 0;
                if( key < ORI_NBINS ) before = wc[wid][key];
                __syncwarp();
                if( key < ORI_NBINS && ( m & lt_mask ) == 0 ) wc[wid][key] = before + __popc( m );
                __syncwarp();
                const int idx = 32 * it + lane;
                sw[idx] = weight; sb[idx] = (unsigned char)key;
                sp[idx] = (unsigned short)( before + __popc( m & lt_mask ) );
                yy += q32; xx += r32; if( xx > xmax ) { xx -= wx; yy++; }
            }
            __syncthreads();
            /* list layout: bins one after the other, within a bin warp 0's samples first */
            if( wid == 0 ) {
                int t0 = 0, t1 = 0;
                #pragma unroll
                for( int ww = 0; ww < ORI_WARPS; ww++ ) { t0 += wc[ww][lane]; if( lane < ORI_NBINS - 32 ) t1 += wc[ww][32 + lane]; }
                int s0 = t0;
                #pragma unroll
                for( int d = 1; d < 32; d <<= 1 ) { const int v = __shfl_up_sync( 0xffffffffu, s0, d ); if( lane >= d ) s0 += v; }
                const int all32 = __shfl_sync( 0xffffffffu, s0, 31 );
                int s1 = t1;
                #pragma unroll
                for( int d = 1; d < 4; d <<= 1 ) { const int v = __shfl_up_sync( 0xffffffffu, s1, d ); if( lane >= d ) s1 += v; }
                bstart[lane] = s0 - t0;
                if( lane < ORI_NBINS - 32 ) bstart[32 + lane] = all32 + s1 - t1;
                if( lane == ORI_NBINS - 32 - 1 ) bstart[ORI_NBINS] = all32 + s1;
                __syncwarp();
                for( int b = lane; b < ORI_NBINS; b += 32 ) {
                    int run = bstart[b];
                    #pragma unroll
                    for( int ww = 0; ww < ORI_WARPS; ww++ ) { wbase[ww][b] = run; run += wc[ww][b]; }
                }
            }
            __syncthreads();
            for( int idx = tid; idx < 32 * n_it; idx += ORI_WARPS * 32 ) {
                const int key = sb[idx];
                if( key < ORI_NBINS ) lst[wbase[( idx >> 5 ) / ipw][key] + sp[idx]] = sw[idx];
            }
            __syncthreads();
            for( int b = tid; b < ORI_NBINS; b += ORI_WARPS * 32 ) {
                float acc = hist[0][b];
                const int j1 = bstart[b + 1];
                for( int j = bstart[b]; j < j1; j++ ) acc = __fadd_rn( acc, lst[j] );
                hist[0][b] = acc;
            }
            __syncthreads();
        }
        __syncthreads();
        if( wid != 0 ) continue;            /* the rest is a warp's worth of work; the others wait at the next hand-out */
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
constexpr int DESC_CLASSES = 64;
__device__ __forceinline__ int desc_class( float sigma ) { return min( DESC_CLASSES - 1, max( 0, (int)( fabsf( sigma ) * 10.0f ) ) ); }
__global__ void __launch_bounds__( PS_NT )
k_ori_prefix( psift_extremum* __restrict__ ext, Counters* __restrict__ ct, int* __restrict__ desc_map, DescWork* __restrict__ work, int desc_capacity )
{
    __shared__ int warp_sum[32];
    __shared__ int carry;
    __shared__ int cls_ct[DESC_CLASSES], cls_at[DESC_CLASSES];
    pdl_enter();
    const int total = ct->ext_total;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if( threadIdx.x == 0 ) carry = 0;
    if( threadIdx.x < DESC_CLASSES ) cls_ct[threadIdx.x] = 0;
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
            if( nv > 0 ) atomicAdd( &cls_ct[desc_class( ext[i].sigma )], nv );
        }
        __syncthreads();
        if( threadIdx.x == 0 ) carry += warp_sum[31];
        __syncthreads();
    }
    /* work order of the descriptor kernel: largest footprints first (classes of sigma, descending; the order inside a
     * class is whatever the atomics give - it decides only WHEN a descriptor is computed, never where it is stored) */
    if( threadIdx.x == 0 ) {
        int at = 0;
        for( int c = DESC_CLASSES - 1; c >= 0; c-- ) { cls_at[c] = at; at += cls_ct[c]; }
    }
    __syncthreads();
    for( int i = threadIdx.x; i < total; i += PS_NT ) {
        const psift_extremum e = ext[i];
        if( e.num_ori > 0 ) {
            const int at = atomicAdd( &cls_at[desc_class( e.sigma )], e.num_ori );
            DescWork w;
            w.x = e.xpos; w.y = e.ypos; w.sigma = e.sigma; w.octave = e.octave; w.level = e.lpos; w.pad = 0;
            for( int k = 0; k < e.num_ori; k++ ) { w.ang = e.orientation[k]; w.d = e.idx_ori + k; work[at + k] = w; }
        }
    }
    /* one thread per octave (the two look-ups of an octave are dependent global loads: nine octaves in a row on one
     * thread were most of this kernel's time) */
    const int ori_total = min( carry, desc_capacity );
    if( threadIdx.x < MAX_OCT ) {
        const int o = threadIdx.x;
        const int fe = ct->ext_ps[o], n = ct->ext_ct[o];
        int lo = ori_total, hi = ori_total;
        if( n > 0 ) {
            lo = ext[fe].idx_ori;
            hi = ( fe + n < total ) ? ext[fe + n].idx_ori : ori_total;
        }
        ct->ori_ps[o] = lo;
        ct->ori_ct[o] = n > 0 ? hi - lo : 0;
    }
    if( threadIdx.x == 0 ) {
        ct->ori_total = ori_total;
        ct->need_desc = carry;
        ct->ori_ps[MAX_OCT] = ori_total;
    }
}

int orientation_blocks_per_sm() { return ORI_WARPS >= 4 ? 10 : ( ORI_WARPS == 2 ? 16 : 28 ); }

cudaError_t launch_orientation( const PyrLayout& L, const Cand* sorted, Counters* ct, psift_extremum* ext,
                                int* desc_map, DescWork* work, int desc_capacity, int n_blocks, cudaStream_t st, long long* launches )
{
    launch_pdl( k_orientation, dim3( n_blocks ), dim3( ORI_WARPS * 32 ), 0, st, L, sorted, ct, ext );
    launch_pdl( k_ori_prefix, dim3( 1 ), dim3( PS_NT ), 0, st, ext, ct, desc_map, work, desc_capacity );
    if( launches ) *launches += 2;
    return cudaGetLastError();
}

} // namespace psift
