/* kernels.h -- launchers of the sm_100a kernels (one translation unit per stage). */
#pragma once
#include "common.h"

namespace psift {

/* k_pyramid.cu */
/* Streams of one frame's pyramid build.  The dependency chain (levels 1..levels of every octave,
 * each octave seeded by the previous one's level `levels`) runs on `chain`; the two remaining
 * levels of each octave hang off that chain and run on side[octave & 1], so that the short,
 * latency-bound launches of the deep octaves overlap the large ones.  side[0] == nullptr: serial. */
struct PyrStreams {
    cudaStream_t chain;
    cudaStream_t side[2];
    cudaEvent_t  seeded[MAX_OCT];    /* level `levels` of octave o is complete */
    cudaEvent_t  side_done[2];
    /* optional (psift_debug_time_levels): event pairs around every launch of octave 0, recorded on the stream the
     * launch goes to; lvl_ev[2*l] before, lvl_ev[2*l+1] after level l.  nullptr: not timed */
    cudaEvent_t* lvl_ev;
    /* Gauss mode "relative" only (hardware-interpolated taps): bilinear textures over every Gaussian level
     * (level_tex[o * n_lvl + l]) and over one scratch plane that takes the horizontal pass (scratch_tex[o]: the same
     * memory seen with octave o's geometry) */
    const cudaTextureObject_t* level_tex;
    const cudaTextureObject_t* scratch_tex;
    float* scratch;
};
cudaError_t launch_pyramid( const PyrLayout& L, cudaTextureObject_t input_tex, float shift, float up_pow, int gauss_mode, int scaling_mode,
                            const GaussTables& T, const PyrStreams& S, long long* launches );
cudaError_t launch_pyramid_level( const PyrLayout& L, int o, int l, cudaTextureObject_t input_tex, float shift, int gauss_mode,
                                  const GaussTables& T, cudaStream_t st );
/* k_extrema.cu */
void extrema_box( int* columns, int* rows );      /* box of the octaves' tensor maps: one tile plus halo */
cudaError_t launch_extrema( const PyrLayout& L, const Consts& K, const void* tensor_maps, Cand* cand, int cand_stride, Cand* sorted, Cand* sort_tmp,
                            int ext_capacity, Counters* ct, SortScratch* ss, cudaStream_t st, long long* launches );
/* k_filter.cu */
size_t filter_scratch_bytes( int ext_capacity );
cudaError_t launch_grid_filter( const PyrLayout& L, const psift_config& cfg, Cand* list, Cand* spare, int ext_capacity,
                                Counters* ct, void* scratch, cudaStream_t st, long long* launches );
/* k_match.cu */
cudaError_t launch_match( const float* l, int l_len, const float* r, int r_len, int* out3, float* dist2, cudaStream_t st );
cudaError_t launch_ref_features( const psift_feature* in, int n, float* desc_base, void* out72, cudaStream_t st );
/* k_orient.cu */
int orientation_blocks_per_sm();
cudaError_t launch_orientation( const PyrLayout& L, const Cand* sorted, Counters* ct, psift_extremum* ext,
                                int* desc_map, DescWork* work, int desc_capacity, int n_blocks, cudaStream_t st, long long* launches );
/* k_desc.cu */
cudaError_t launch_descriptors( const PyrLayout& L, const Consts& K, const psift_extremum* ext, Counters* ct,
                                const DescWork* work, const cudaTextureObject_t* lintex, const DescTables* tab,
                                float* desc, psift_feature* feat, int n_blocks, cudaStream_t st, long long* launches );
/* pipeline.cu: sets the calling thread's psift_last_error() text and returns `code` (for entry points that live in other files) */
int set_error( int code, const char* msg );
/* pipeline.cu helper kernel: DoG plane for the parity hook */
cudaError_t launch_dog_plane( const float* lo, const float* hi, float* out, int w, int h, int pitch, cudaStream_t st );

} // namespace psift
