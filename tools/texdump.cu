/* texdump.cu -- dump the 2x-upscaled view of a u8 image exactly as the reference's input
 * sampler produces it (normalized-coordinate bilinear u8 texture, s_image.cu:140-169, sampled at
 * (x+1)/W0,(y+1)/H0 as in s_pyramid_build_ra.cu:33-34).  Used once to fit the texture unit's
 * fixed-point blend for the CPU oracle.  TEST INFRASTRUCTURE.
 *   texdump in.u8 w h out.f32 */
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
__global__ void dump( cudaTextureObject_t tex, int W0, int H0, float shift, float* out )
{
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if( x < W0 ) out[(size_t)y * W0 + x] = tex2D<float>( tex, ( x + shift ) / W0, ( y + shift ) / H0 );
}
int main( int argc, char** argv )
{
    int w = atoi( argv[2] ), h = atoi( argv[3] );
    std::vector<unsigned char> img( (size_t)w * h );
    FILE* f = fopen( argv[1], "rb" ); if( !f || fread( img.data(), 1, img.size(), f ) != img.size() ) return 2; fclose( f );
    unsigned char* d; size_t pitch;
    cudaMallocPitch( (void**)&d, &pitch, w, h );
    cudaMemcpy2D( d, pitch, img.data(), w, w, h, cudaMemcpyHostToDevice );
    cudaTextureDesc td = {}; td.normalizedCoords = 1;
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
    td.readMode = cudaReadModeNormalizedFloat; td.filterMode = cudaFilterModeLinear;
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypePitch2D; rd.res.pitch2D.devPtr = d;
    rd.res.pitch2D.desc = cudaCreateChannelDesc( 8, 0, 0, 0, cudaChannelFormatKindUnsigned );
    rd.res.pitch2D.pitchInBytes = pitch; rd.res.pitch2D.width = w; rd.res.pitch2D.height = h;
    cudaTextureObject_t tex; cudaCreateTextureObject( &tex, &rd, &td, 0 );
    int W0 = 2 * w, H0 = 2 * h;
    float* o; cudaMalloc( &o, sizeof(float) * W0 * H0 );
    dump<<< dim3( ( W0 + 127 ) / 128, H0 ), 128 >>>( tex, W0, H0, 1.0f, o );
    std::vector<float> ho( (size_t)W0 * H0 );
    if( cudaMemcpy( ho.data(), o, sizeof(float) * W0 * H0, cudaMemcpyDeviceToHost ) != cudaSuccess ) return 3;
    f = fopen( argv[4], "wb" ); fwrite( ho.data(), 4, ho.size(), f ); fclose( f );
    return 0;
}
