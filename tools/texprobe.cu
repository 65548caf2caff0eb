/* texprobe.cu -- one-off hardware probe (test infrastructure, not product code).
 *
 * Question: the reference's level-0 kernel (s_pyramid_build_ra.cu:17-55) samples the
 * u8 input through a normalized-coordinate bilinear texture at coordinates that are,
 * for the 2x upscale, exact half-texel positions.  Can that be replaced by plain
 * loads + arithmetic without changing a single bit?
 *
 *  A: for every output x and tap offset, is tex(read_x - offset/W0) bit-identical to
 *     u[x-offset] where u[j] = tex((j + shift)/W0) with j clamped to [0,W0-1] ?
 *  B: dump tex value for every (a,b) u8 pair at fraction 0.5 and for fraction 0, so
 *     that the blend arithmetic can be fitted offline.
 *
 * build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o texprobe tools/texprobe.cu
 */
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if( e != cudaSuccess ) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while(0)

static cudaTextureObject_t make_tex( unsigned char* d, int w, int h, size_t pitch )
{
    cudaTextureDesc td = {};
    td.normalizedCoords = 1;
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.readMode = cudaReadModeNormalizedFloat;
    td.filterMode = cudaFilterModeLinear;
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypePitch2D;
    rd.res.pitch2D.devPtr = d;
    rd.res.pitch2D.desc = cudaCreateChannelDesc( 8, 0, 0, 0, cudaChannelFormatKindUnsigned );
    rd.res.pitch2D.pitchInBytes = pitch;
    rd.res.pitch2D.width = w;
    rd.res.pitch2D.height = h;
    cudaTextureObject_t t;
    CK( cudaCreateTextureObject( &t, &rd, &td, 0 ) );
    return t;
}

/* A: count mismatches between the reference's tap coordinate and the "row of upscaled
 * pixels" view, horizontally and vertically. */
__global__ void probeA( cudaTextureObject_t tex, int W0, int H0, float shift, int span, unsigned long long* bad, unsigned long long* badv )
{
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y;
    if( x >= W0 ) return;
    const float read_x = ( x + shift ) / W0;
    const float read_y = ( y + shift ) / H0;
    for( int off = -span; off <= span; off++ ) {
        const float offrel = float( abs( off ) ) / W0;
        const float v = tex2D<float>( tex, off < 0 ? read_x - offrel : read_x + offrel, read_y );
        int j = min( max( x + off, 0 ), W0 - 1 );
        const float u = tex2D<float>( tex, ( j + shift ) / W0, read_y );
        if( __float_as_uint( v ) != __float_as_uint( u ) ) atomicAdd( bad, 1ull );
    }
    /* vertical consistency: value at (x,y) equals a separable view ? compare tex(x,y)
     * against lerp of the two source rows sampled at exact texel centres */
    const float v  = tex2D<float>( tex, read_x, read_y );
    const int   Hin = H0 / 2, Win = W0 / 2;
    if( shift == 1.0f ) {
        /* source position y/2 - ... : texel space = read_y*Hin - 0.5 = y/2 */
        int   y0 = y >> 1, y1 = min( y0 + 1, Hin - 1 );
        float r0 = tex2D<float>( tex, read_x, ( y0 + 0.5f ) / Hin );
        float r1 = tex2D<float>( tex, read_x, ( y1 + 0.5f ) / Hin );
        float e  = ( y & 1 ) ? ( r0 + r1 ) * 0.5f : r0;
        if( __float_as_uint( v ) != __float_as_uint( e ) ) atomicAdd( badv, 1ull );
    }
}

/* B: pair table.  Texture is 256 rows x 512 cols; row r: col 2k = r, col 2k+1 = k. */
__global__ void probeB( cudaTextureObject_t tex, float* half, float* zero, float* quarter )
{
    int k = threadIdx.x, r = blockIdx.x;
    const float W = 512.0f, H = 256.0f;
    /* texel-space position 2k + 0.5 (half way between col 2k and 2k+1) -> normalized (2k+1)/W */
    half[r * 256 + k]    = tex2D<float>( tex, ( 2 * k + 1.0f ) / W, ( r + 0.5f ) / H );
    quarter[r * 256 + k] = tex2D<float>( tex, ( 2 * k + 0.75f ) / W, ( r + 0.5f ) / H );
    if( k == 0 ) zero[r] = tex2D<float>( tex, 0.5f / W, ( r + 0.5f ) / H );
}

int main( int argc, char** argv )
{
    const char* outdir = argc > 1 ? argv[1] : ".";
    unsigned long long *d_bad, *d_badv;
    CK( cudaMalloc( &d_bad, 8 ) ); CK( cudaMalloc( &d_badv, 8 ) );
    int sizes[][2] = { {64,48}, {640,480}, {1920,1080}, {3840,2160}, {333,217}, {1000,1000} };
    for( auto& s : sizes ) {
        int w = s[0], h = s[1];
        std::vector<unsigned char> img( (size_t)w * h );
        unsigned st = 12345u + w;
        for( auto& p : img ) { st = st * 1664525u + 1013904223u; p = st >> 24; }
        unsigned char* d; size_t pitch;
        CK( cudaMallocPitch( (void**)&d, &pitch, w, h ) );
        CK( cudaMemcpy2D( d, pitch, img.data(), w, w, h, cudaMemcpyHostToDevice ) );
        cudaTextureObject_t tex = make_tex( d, w, h, pitch );
        for( int up = 0; up < 2; up++ ) {
            int W0 = w << up, H0 = h << up;
            float shift = up ? 1.0f : 0.5f;
            CK( cudaMemset( d_bad, 0, 8 ) ); CK( cudaMemset( d_badv, 0, 8 ) );
            probeA<<< dim3( (W0 + 127) / 128, H0 ), 128 >>>( tex, W0, H0, shift, 8, d_bad, d_badv );
            CK( cudaDeviceSynchronize() );
            unsigned long long bad, badv;
            CK( cudaMemcpy( &bad, d_bad, 8, cudaMemcpyDeviceToHost ) );
            CK( cudaMemcpy( &badv, d_badv, 8, cudaMemcpyDeviceToHost ) );
            printf( "probeA %dx%d up=%d: tap-coordinate mismatches %llu of %llu ; vertical-lerp mismatches %llu of %llu\n",
                    w, h, up, bad, (unsigned long long)W0 * H0 * 17, badv, (unsigned long long)W0 * H0 );
        }
        cudaDestroyTextureObject( tex ); cudaFree( d );
    }
    {
        int w = 512, h = 256;
        std::vector<unsigned char> img( (size_t)w * h );
        for( int r = 0; r < h; r++ ) for( int k = 0; k < 256; k++ ) { img[r * w + 2 * k] = r; img[r * w + 2 * k + 1] = k; }
        unsigned char* d; size_t pitch;
        CK( cudaMallocPitch( (void**)&d, &pitch, w, h ) );
        CK( cudaMemcpy2D( d, pitch, img.data(), w, w, h, cudaMemcpyHostToDevice ) );
        cudaTextureObject_t tex = make_tex( d, w, h, pitch );
        float *dh, *dz, *dq;
        CK( cudaMalloc( &dh, 65536 * 4 ) ); CK( cudaMalloc( &dz, 256 * 4 ) ); CK( cudaMalloc( &dq, 65536 * 4 ) );
        probeB<<< 256, 256 >>>( tex, dh, dz, dq );
        CK( cudaDeviceSynchronize() );
        std::vector<float> hh( 65536 ), hz( 256 ), hq( 65536 );
        CK( cudaMemcpy( hh.data(), dh, 65536 * 4, cudaMemcpyDeviceToHost ) );
        CK( cudaMemcpy( hz.data(), dz, 256 * 4, cudaMemcpyDeviceToHost ) );
        CK( cudaMemcpy( hq.data(), dq, 65536 * 4, cudaMemcpyDeviceToHost ) );
        char fn[512];
        snprintf( fn, sizeof fn, "%s/texprobe_pairs.bin", outdir );
        FILE* f = fopen( fn, "wb" );
        fwrite( hz.data(), 4, 256, f ); fwrite( hh.data(), 4, 65536, f ); fwrite( hq.data(), 4, 65536, f );
        fclose( f );
        /* quick in-place candidates */
        int m1 = 0, m2 = 0, m3 = 0, z1 = 0, z2 = 0;
        for( int r = 0; r < 256; r++ ) {
            if( hz[r] != r / 255.0f ) z1++;
            if( hz[r] != r * ( 1.0f / 255.0f ) ) z2++;
            for( int k = 0; k < 256; k++ ) {
                float a = r / 255.0f, b = k / 255.0f, t = hh[r * 256 + k];
                if( t != ( a + b ) * 0.5f ) m1++;
                if( t != ( r + k ) / 510.0f ) m2++;
                if( t != a + ( b - a ) * 0.5f ) m3++;
            }
        }
        printf( "probeB: zero-frac mismatches: r/255.0f %d, r*(1/255) %d ; half-frac mismatches: (a+b)*0.5 %d, (r+k)/510 %d, a+(b-a)/2 %d\n", z1, z2, m1, m2, m3 );
    }
    return 0;
}
