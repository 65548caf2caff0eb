/* k_filter.cu -- grid filter: cap the number of keypoints per image cell, for sm_100a.
 *
 * Reference: Pyramid::extrema_filter_grid (s_filtergrid.cu:108-322), called from
 * Pyramid::orientation when filter_max_extrema > 0 and int(1.1 * limit) < total
 * (s_orientation.cu:362-367).  The reference runs ~20 Thrust launches, copies the cell histogram
 * to the host, computes the per-cell limit there and copies it back, with device-wide syncs in
 * between.  Here the whole decision stays on the device and in the frame's stream:
 *
 *   k_filter_plan    one block: cell histogram, the reference's limit rule (:203-263, on the
 *                    ascending-sorted cell counts), exclusive cell offsets
 *   k_filter_bucket  counting-sort scatter of (scale, index) into cell buckets
 *   k_filter_mark    rank of every keypoint inside its cell under the configured order
 *                    (largest / smallest scale first, :155-183; "random" = arrival order,
 *                    :185-193) and keep = rank < min(count, limit)  (:265-279)
 *   k_filter_compact one block: order-preserving compaction per octave, new counters (:281-316)
 *   k_filter_copy    compacted list back into the working list
 *
 * Every kernel returns at once when the filter does not apply, so an unfiltered configuration
 * only pays for five empty launches when filter_max_extrema > 0, and nothing otherwise.
 *
 * Order of equal keys: Thrust's merge sort is stable, so the reference breaks ties by arrival
 * order, which its atomics make nondeterministic.  Here arrival order is the deterministic
 * (octave, level, y, x) order of the rank sort, which is one of the orders the reference can
 * produce.
 */
#include "common.h"
#include "kernels.h"

namespace psift {

constexpr int FG_MAX_CELLS = 1024;      /* grid sizes up to 32 x 32 */
constexpr int FG_NT        = 1024;

struct FilterPlan {
    int active;
    int limit;
    int n_cells;
    int pad;
    int count[FG_MAX_CELLS];
    int offset[FG_MAX_CELLS + 1];
    int cursor[FG_MAX_CELLS];
};

__device__ __forceinline__ int octave_of( const Counters* ct, int n_oct, int i )
{
    int o = 0;
    while( o + 1 < n_oct && i >= ct->ext_ps[o + 1] ) o++;
    return o;
}

__global__ void __launch_bounds__( FG_NT )
k_filter_plan( const Cand* __restrict__ list, const Counters* __restrict__ ct, FilterPlan* plan,
               int grid_size, int filter_max )
{
    __shared__ int cnt[FG_MAX_CELLS];
    __shared__ int srt[FG_MAX_CELLS];
    const int tid = threadIdx.x;
    const int total = ct->ext_total;
    const int G = grid_size * grid_size;
    const bool active = filter_max > 0 && (int)( (double)filter_max * 1.1 ) < total;      /* s_orientation.cu:364 */
    if( tid == 0 ) { plan->active = active ? 1 : 0; plan->n_cells = G; }
    if( !active ) return;

    for( int c = tid; c < FG_MAX_CELLS; c += FG_NT ) cnt[c] = 0;
    __syncthreads();
    for( int i = tid; i < total; i += FG_NT ) atomicAdd( &cnt[min( max( list[i].cell, 0 ), G - 1 )], 1 );
    __syncthreads();

    /* ascending sort of the G counts (s_filtergrid.cu:221-222); empty cells count as zeros,
     * exactly like the zero-initialised tail of the reference's cell_counts vector */
    if( tid < G ) {
        const int mine = cnt[tid];
        int rank = 0;
        for( int j = 0; j < G; j++ ) rank += ( cnt[j] < mine || ( cnt[j] == mine && j < tid ) ) ? 1 : 0;
        srt[rank] = mine;
    }
    __syncthreads();

    if( tid == 0 ) {
        /* :226-252  sumup[i] = prefix[i] + count[i] * (cells after i); the cells whose sumup exceeds
         * the limit share the reduction */
        int prefix = 0, above = 0;
        for( int i = 0; i < G; i++ ) {
            prefix += srt[i];
            const int sumup = srt[i] * ( G - 1 - i ) + prefix;
            if( sumup > filter_max ) above++;
        }
        int tail = 0;
        for( int i = G - above; i < G; i++ ) tail += srt[i];
        int limit = 0x7fffffff;
        if( above > 0 ) {
            const float tailaverage = __fdiv_rn( (float)tail, (float)above );
            limit = (int)ceilf( __fsub_rn( tailaverage, (float)( ( total - filter_max ) / above ) ) );
        }
        plan->limit = limit;
        int off = 0;
        for( int c = 0; c < G; c++ ) { plan->count[c] = cnt[c]; plan->offset[c] = off; plan->cursor[c] = 0; off += cnt[c]; }
        plan->offset[G] = off;
    }
}

/* scale the reference sorts by: sigma * 2^octave (s_filtergrid.cu:66) */
__global__ void k_filter_bucket( const Cand* __restrict__ list, const Counters* __restrict__ ct, int n_oct,
                                 FilterPlan* plan, float* __restrict__ b_scale, int* __restrict__ b_index )
{
    if( !plan->active ) return;
    const int total = ct->ext_total, G = plan->n_cells;
    for( int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x ) {
        const int cell = min( max( list[i].cell, 0 ), G - 1 );
        const int pos  = plan->offset[cell] + atomicAdd( &plan->cursor[cell], 1 );
        b_scale[pos] = __fmul_rn( list[i].sigma, exp2f( (float)octave_of( ct, n_oct, i ) ) );
        b_index[pos] = i;
    }
}

__global__ void k_filter_mark( const Counters* __restrict__ ct, const FilterPlan* __restrict__ plan,
                               const float* __restrict__ b_scale, const int* __restrict__ b_index,
                               unsigned char* __restrict__ keep, int sorting )
{
    if( !plan->active ) return;
    const int total = ct->ext_total, G = plan->n_cells;
    for( int p = blockIdx.x * blockDim.x + threadIdx.x; p < total; p += gridDim.x * blockDim.x ) {
        int cell = 0;                                            /* bucket p belongs to: binary search in offset[] */
        { int lo = 0, hi = G; while( hi - lo > 1 ) { const int mid = ( lo + hi ) >> 1; if( plan->offset[mid] <= p ) lo = mid; else hi = mid; } cell = lo; }
        const int beg = plan->offset[cell], end = plan->offset[cell + 1];
        const float s = b_scale[p];
        const int   i = b_index[p];
        int rank = 0;
        for( int q = beg; q < end; q++ ) {
            const float sq = b_scale[q];
            const int   iq = b_index[q];
            bool before;
            if( sorting == PSIFT_FILTER_LARGEST_FIRST )       before = sq > s || ( sq == s && iq < i );
            else if( sorting == PSIFT_FILTER_SMALLEST_FIRST ) before = sq < s || ( sq == s && iq < i );
            else                                              before = iq < i;
            rank += before ? 1 : 0;
        }
        keep[i] = rank < min( end - beg, plan->limit ) ? 1 : 0;
    }
}

/* order-preserving compaction inside every octave; rewrites ext_ct / ext_ps / ext_total */
__global__ void __launch_bounds__( FG_NT )
k_filter_compact( const Cand* __restrict__ list, Cand* __restrict__ out, Counters* ct, int n_oct,
                  const FilterPlan* __restrict__ plan, const unsigned char* __restrict__ keep )
{
    if( !plan->active ) return;
    __shared__ int warp_sum[32];
    __shared__ int carry;
    __shared__ int new_ct[MAX_OCT];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int total = ct->ext_total;
    if( tid == 0 ) carry = 0;
    if( tid < MAX_OCT ) new_ct[tid] = 0;
    __syncthreads();
    for( int base = 0; base < total; base += FG_NT ) {
        const int i = base + tid;
        const int v = ( i < total ) ? keep[i] : 0;
        int inc = v;
        #pragma unroll
        for( int s = 1; s < 32; s <<= 1 ) { const int t = __shfl_up_sync( 0xffffffffu, inc, s ); if( lane >= s ) inc += t; }
        if( lane == 31 ) warp_sum[wid] = inc;
        __syncthreads();
        if( wid == 0 ) {
            int ws = warp_sum[lane];
            #pragma unroll
            for( int s = 1; s < 32; s <<= 1 ) { const int t = __shfl_up_sync( 0xffffffffu, ws, s ); if( lane >= s ) ws += t; }
            warp_sum[lane] = ws;
        }
        __syncthreads();
        const int excl = carry + ( wid ? warp_sum[wid - 1] : 0 ) + inc - v;
        if( v ) {
            out[excl] = list[i];
            atomicAdd( &new_ct[octave_of( ct, n_oct, i )], 1 );
        }
        __syncthreads();
        if( tid == FG_NT - 1 ) carry = excl + v;
        __syncthreads();
    }
    if( tid == 0 ) {
        int ps = 0;
        for( int o = 0; o < MAX_OCT; o++ ) { ct->ext_ct[o] = new_ct[o]; ct->ext_ps[o] = ps; ps += new_ct[o]; }
        ct->ext_ps[MAX_OCT] = ps;
        ct->ext_total = ps;
    }
}

__global__ void k_filter_copy( const Cand* __restrict__ src, Cand* __restrict__ dst, const Counters* __restrict__ ct,
                               const FilterPlan* __restrict__ plan )
{
    if( !plan->active ) return;
    const int total = ct->ext_total;
    for( int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x ) dst[i] = src[i];
}

size_t filter_scratch_bytes( int ext_capacity )
{
    return sizeof(FilterPlan) + (size_t)ext_capacity * ( sizeof(float) + sizeof(int) + 1 ) + 64;
}

cudaError_t launch_grid_filter( const PyrLayout& L, const psift_config& cfg, Cand* list, Cand* spare, int ext_capacity,
                                Counters* ct, void* scratch, cudaStream_t st, long long* launches )
{
    if( cfg.filter_max_extrema <= 0 ) return cudaSuccess;
    if( cfg.filter_grid_size < 1 || cfg.filter_grid_size * cfg.filter_grid_size > FG_MAX_CELLS ) return cudaErrorInvalidValue;
    char* p = static_cast<char*>( scratch );
    FilterPlan* plan = reinterpret_cast<FilterPlan*>( p );              p += sizeof(FilterPlan);
    float* b_scale   = reinterpret_cast<float*>( p );                   p += sizeof(float) * (size_t)ext_capacity;
    int*   b_index   = reinterpret_cast<int*>( p );                     p += sizeof(int) * (size_t)ext_capacity;
    unsigned char* keep = reinterpret_cast<unsigned char*>( p );
    k_filter_plan<<<1, FG_NT, 0, st>>>( list, ct, plan, cfg.filter_grid_size, cfg.filter_max_extrema );
    k_filter_bucket<<<sm_count(), 256, 0, st>>>( list, ct, L.n_oct, plan, b_scale, b_index );
    k_filter_mark<<<sm_count() * 4, 256, 0, st>>>( ct, plan, b_scale, b_index, keep, cfg.filter_sorting );
    k_filter_compact<<<1, FG_NT, 0, st>>>( list, spare, ct, L.n_oct, plan, keep );
    k_filter_copy<<<sm_count(), 256, 0, st>>>( spare, list, ct, plan );
    if( launches ) *launches += 5;
    return cudaGetLastError();
}

} // namespace psift
