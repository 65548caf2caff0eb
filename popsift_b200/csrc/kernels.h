/* kernels.h -- launchers of the sm_100a kernels (one translation unit per stage). */
#pragma once
#include "common.h"

namespace psift {

/* k_pyramid.cu */
cudaError_t launch_pyramid( const PyrLayout& L, cudaTextureObject_t input_tex, float shift,
                            const GaussTables& T, cudaStream_t st, long long* launches );
/* k_extrema.cu */
cudaError_t launch_extrema( const PyrLayout& L, const Consts& K, Cand* cand, int cand_stride, Cand* sorted,
                            int ext_capacity, Counters* ct, cudaStream_t st, long long* launches );
/* k_orient.cu */
cudaError_t launch_orientation( const PyrLayout& L, const Cand* sorted, Counters* ct, psift_extremum* ext,
                                int* desc_map, int desc_capacity, int n_blocks, cudaStream_t st, long long* launches );
/* k_desc.cu */
cudaError_t launch_descriptors( const PyrLayout& L, const Consts& K, const psift_extremum* ext, Counters* ct,
                                const int* desc_map, const cudaTextureObject_t* lintex, const DescTables* tab,
                                float* desc, psift_feature* feat, int n_blocks, cudaStream_t st, long long* launches );
/* pipeline.cu helper kernel: DoG plane for the parity hook */
cudaError_t launch_dog_plane( const float* lo, const float* hi, float* out, int w, int h, int pitch, cudaStream_t st );

} // namespace psift
