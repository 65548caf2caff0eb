/*
 * ref_driver.cu -- TEST INFRASTRUCTURE, not product code.
 *
 * Boost-free driver around the UNMODIFIED reference sources under
 * /root/reference/src/popsift (compiled where they lie by oracle/build_ref.sh;
 * nothing is copied into this repo).  It does what the reference's worker
 * thread does per image (reference src/popsift/popsift.cpp:63-120 configure +
 * private_init, :182-213 extractDownloadLoop, :277-282 SiftJob::setImg) and
 * writes the results in a small binary format that tests/ and bench.py read.
 *
 * It needs a GPU: the reference has no CPU path.  Used for
 *   - golden outputs for parity tests (features, descriptors, Gaussian/DoG planes)
 *   - the timed "reference arm" of bench.py (--impl reference).
 */
/* test infrastructure only: reach the per-stage entry points (find_extrema, orientation,
 * descriptors are private in the reference's Pyramid, sift_pyramid.h:118-170) */
#define private public
#include "popsift/sift_pyramid.h"
#undef private
#include "popsift/sift_octave.h"
#include "popsift/s_image.h"
#include "popsift/gauss_filter.h"
#include "popsift/sift_constants.h"
#include "popsift/sift_conf.h"
#include "popsift/features.h"

#include <cuda_runtime.h>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>

using namespace popsift;

namespace popsift { __global__ void compute_distance( int3* match_matrix, Descriptor* l, int l_len, Descriptor* r, int r_len ); }
/* The reference DECLARES Config::setMaxExtrema (sift_conf.h:114) but never defines it; this is the missing one-line
 * definition, so that `--max-extrema` can lift the per-octave cap of 100 000 (sift_conf.cu:32) for the config-5 test
 * (which keypoints a clamp keeps is atomic arrival order: no parity to speak of once an octave is clamped). */
namespace popsift { void Config::setMaxExtrema( int extrema ) { _max_extrema = extrema; } }
namespace popsift { __global__ void ori_par( const int octave, const int ext_ct_prefix_sum, cudaTextureObject_t layer, const int w, const int h ); }

static void die( const char* msg ) { fprintf( stderr, "ref_driver: %s\n", msg ); exit( 2 ); }

static std::vector<unsigned char> read_file( const std::string& fn )
{
    FILE* f = fopen( fn.c_str(), "rb" );
    if( !f ) { fprintf( stderr, "ref_driver: cannot open %s\n", fn.c_str() ); exit( 2 ); }
    fseek( f, 0, SEEK_END ); long n = ftell( f ); fseek( f, 0, SEEK_SET );
    std::vector<unsigned char> b( n );
    if( fread( b.data(), 1, n, f ) != (size_t)n ) die( "short read" );
    fclose( f );
    return b;
}

struct FeatRec { int octave; float x, y, sigma; int num_ori; float ori[4]; int desc_idx[4]; };

static void write_features( const std::string& fn, Pyramid* pyr, FeaturesHost* fh, int w0, int h0 )
{
    FILE* f = fopen( fn.c_str(), "wb" );
    if( !f ) die( "cannot open feature output" );
    int hdr[6] = { 0x54465350, 1, pyr->getNumOctaves(), pyr->getNumLevels(), w0, h0 };
    fwrite( hdr, 4, 6, f );
    fwrite( hct.ext_ct, 4, MAX_OCTAVES, f );
    fwrite( hct.ori_ct, 4, MAX_OCTAVES, f );
    int tot[2] = { fh->getFeatureCount(), fh->getDescriptorCount() };
    fwrite( tot, 4, 2, f );
    Descriptor* base = fh->getDescriptors();
    for( int i = 0; i < tot[0]; i++ ) {
        const Feature& ft = fh->getFeatures()[i];
        FeatRec r;
        r.octave = ft.debug_octave; r.x = ft.xpos; r.y = ft.ypos; r.sigma = ft.sigma; r.num_ori = ft.num_ori;
        for( int k = 0; k < 4; k++ ) {
            r.ori[k] = ft.orientation[k];
            r.desc_idx[k] = ( k < ft.num_ori && ft.desc[k] ) ? int( ft.desc[k] - base ) : -1;
        }
        fwrite( &r, sizeof(r), 1, f );
    }
    if( tot[1] > 0 ) fwrite( base, sizeof(Descriptor), tot[1], f );
    /* internal Extremum records (octave-resolution coordinates, lpos): 44 B each */
    std::vector<Extremum> ext( std::max( tot[0], 1 ) );
    if( tot[0] > 0 )
        cudaMemcpy( ext.data(), dobuf_shadow.extrema, tot[0] * sizeof(Extremum), cudaMemcpyDeviceToHost );
    fwrite( ext.data(), sizeof(Extremum), tot[0], f );
    fclose( f );
}

/* Planes are read back through the reference's own public point-texture accessors
 * (sift_octave.h:110-131) with the reference's readTex addressing (common/assist.h:66-75). */
__global__ void read_layer( cudaTextureObject_t tex, int w, int h, int layer, float* out )
{
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y;
    if( x < w ) out[(size_t)y * w + x] = tex2DLayered<float>( tex, x + 0.5f, y + 0.5f, layer );
}

static void write_planes( const std::string& fn, Pyramid* pyr )
{
    FILE* f = fopen( fn.c_str(), "wb" );
    if( !f ) die( "cannot open plane output" );
    int hdr[3] = { 0x4e4c5050, pyr->getNumOctaves(), pyr->getNumLevels() };
    fwrite( hdr, 4, 3, f );
    for( int o = 0; o < pyr->getNumOctaves(); o++ ) {
        Octave& oc = pyr->getOctave( o );
        int w = oc.getWidth(), h = oc.getHeight(), L = oc.getLevels();
        int oh[2] = { w, h };
        fwrite( oh, 4, 2, f );
        std::vector<float> buf( (size_t)w * h );
        float* d_buf; cudaMalloc( &d_buf, buf.size() * sizeof(float) );
        cudaTextureObject_t texs[2] = { oc.getDataTexPoint(), oc.getDogTexturePoint() };
        int                 deps[2] = { L, L - 1 };
        for( int a = 0; a < 2; a++ ) {
            for( int l = 0; l < deps[a]; l++ ) {
                read_layer<<< dim3( (w + 127) / 128, h ), 128 >>>( texs[a], w, h, l, d_buf );
                if( cudaMemcpy( buf.data(), d_buf, buf.size() * sizeof(float), cudaMemcpyDeviceToHost ) != cudaSuccess )
                    die( "plane download failed" );
                fwrite( buf.data(), sizeof(float), buf.size(), f );
            }
        }
        cudaFree( d_buf );
    }
    fclose( f );
}

/* --match-desc L.f32 R.f32 --out X : the reference's matcher kernel (features.cu:184-225, what FeaturesDev::match
 * launches, features.cu:268-285) on two raw descriptor files; writes l_len x int3 {best, second, accept} to X. */
static int match_mode( const std::string& lf, const std::string& rf, const std::string& out )
{
    std::vector<unsigned char> lb = read_file( lf ), rb = read_file( rf );
    const int nl = (int)( lb.size() / sizeof(Descriptor) ), nr = (int)( rb.size() / sizeof(Descriptor) );
    FeaturesDev L( 1, std::max( nl, 1 ) ), R( 1, std::max( nr, 1 ) );
    cudaMemcpy( L.getDescriptors(), lb.data(), (size_t)nl * sizeof(Descriptor), cudaMemcpyHostToDevice );
    cudaMemcpy( R.getDescriptors(), rb.data(), (size_t)nr * sizeof(Descriptor), cudaMemcpyHostToDevice );
    int3* mm = 0; cudaMalloc( &mm, sizeof(int3) * std::max( nl, 1 ) );
    auto t0 = std::chrono::steady_clock::now();
    if( nl > 0 ) compute_distance<<< dim3( nl ), dim3( 32 ) >>>( mm, L.getDescriptors(), nl, R.getDescriptors(), nr );
    if( cudaDeviceSynchronize() != cudaSuccess ) die( "compute_distance failed" );
    const double ms = std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now() - t0 ).count();
    std::vector<int3> h( std::max( nl, 1 ) );
    cudaMemcpy( h.data(), mm, sizeof(int3) * nl, cudaMemcpyDeviceToHost );
    FILE* f = fopen( out.c_str(), "wb" );
    if( !f ) die( "cannot open match output" );
    fwrite( h.data(), sizeof(int3), nl, f );
    fclose( f );
    cudaFree( mm );
    printf( "{\"l_len\": %d, \"r_len\": %d, \"ms\": %.3f}\n", nl, nr, ms );
    return 0;
}

int main( int argc, char** argv )
{
    std::vector<std::string> inputs;
    std::string out_prefix;
    int w = 0, h = 0, iters = 1, warmup = 0, ori_o = -1, ori_i = 0;
    bool dump_planes = false, float_mode = false, tables = false, debug_stages = false;

    cudaSetDevice( 0 );
    Config conf;
    std::string match_l, match_r;

    for( int i = 1; i < argc; i++ ) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { if( i + 1 >= argc ) die( "missing value" ); return argv[++i]; };
        if(      a == "--in" )      inputs.push_back( next() );
        else if( a == "--match-desc" ) { match_l = next(); match_r = next(); }
        else if( a == "--out" )     out_prefix = next();
        else if( a == "--w" )       w = atoi( next() );
        else if( a == "--h" )       h = atoi( next() );
        else if( a == "--iters" )   iters = atoi( next() );
        else if( a == "--warmup" )  warmup = atoi( next() );
        else if( a == "--dump-planes" ) dump_planes = true;
        else if( a == "--float-mode" )  float_mode = true;
        else if( a == "--dump-tables" ) tables = true;
        else if( a == "--debug-stages" ) debug_stages = true;
        else if( a == "--ori-one" ) { debug_stages = true; ori_o = atoi( next() ); ori_i = atoi( next() ); }
        else if( a == "--mode" ) {
            std::string m = next();
            if( m == "popsift" ) conf.setMode( Config::PopSift );
            else if( m == "vlfeat" ) conf.setMode( Config::VLFeat );
            else if( m == "opencv" ) conf.setMode( Config::OpenCV );
            else die( "bad --mode" );
        }
        else if( a == "--gauss-mode" )   conf.setGaussMode( std::string( next() ) );
        else if( a == "--desc-mode" )    conf.setDescMode( std::string( next() ) );
        else if( a == "--norm-mode" )    conf.setNormMode( std::string( next() ) );
        else if( a == "--norm-multi" )   conf.setNormalizationMultiplier( atoi( next() ) );
        else if( a == "--downsampling" ) conf.setDownsampling( (float)atof( next() ) );
        else if( a == "--octaves" )      conf.setOctaves( atoi( next() ) );
        else if( a == "--levels" )       conf.setLevels( atoi( next() ) );
        else if( a == "--sigma" )        conf.setSigma( (float)atof( next() ) );
        else if( a == "--threshold" )    conf.setThreshold( (float)atof( next() ) );
        else if( a == "--edge-threshold" ) conf.setEdgeLimit( (float)atof( next() ) );
        else if( a == "--initial-blur" ) conf.setInitialBlur( (float)atof( next() ) );
        else if( a == "--max-extrema" )  conf.setMaxExtrema( atoi( next() ) );
        else if( a == "--filter-max-extrema" ) conf.setFilterMaxExtrema( atoi( next() ) );
        else if( a == "--filter-grid" )  conf.setFilterGridSize( atoi( next() ) );
        else if( a == "--filter-sort" )  conf.setFilterSorting( std::string( next() ) );
        else if( a == "--scaling" ) {
            std::string m = next();
            if( m == "direct" ) conf.setScalingMode( Config::ScaleDirect );
            else if( m == "default" ) conf.setScalingMode( Config::ScaleDefault );
            else die( "bad --scaling" );
        }
        else { fprintf( stderr, "ref_driver: unknown flag %s\n", a.c_str() ); return 2; }
    }
    if( !match_l.empty() ) return match_mode( match_l, match_r, out_prefix );
    if( inputs.empty() || w <= 0 || h <= 0 ) die( "need --in FILE --w W --h H" );

    /* PopSift::configure, reference popsift.cpp:63-87 */
    conf.levels = std::max( 2, conf.levels );
    init_filter( conf, conf.sigma, conf.levels );
    init_constants( conf.sigma, conf.levels, conf.getPeakThreshold(), conf._edge_limit,
                    conf.getMaxExtrema(), conf.getNormalizationMultiplier() );

    if( tables && !out_prefix.empty() ) {
        FILE* f = fopen( ( out_prefix + ".tables.bin" ).c_str(), "wb" );
        fwrite( &h_gauss, sizeof(GaussInfo), 1, f );
        fwrite( &h_consts, sizeof(ConstInfo), 1, f );
        fclose( f );
    }

    /* PopSift::private_init, reference popsift.cpp:89-120 */
    const float upscaleFactor = conf.getUpscaleFactor();
    const float scaleFactor   = 1.0f / powf( 2.0f, -upscaleFactor );
    if( conf.octaves < 0 ) {
        int oct = std::max( int( floor( logf( (float)std::min( w, h ) ) / logf( 2.0f ) ) - 3.0f + scaleFactor ), 1 );
        conf.octaves = oct;
    }
    const int w0 = ceilf( w * scaleFactor );
    const int h0 = ceilf( h * scaleFactor );

    std::vector<std::vector<unsigned char> > imgs;
    const size_t need = (size_t)w * h * ( float_mode ? 4 : 1 );
    for( auto& fn : inputs ) {
        imgs.push_back( read_file( fn ) );
        if( imgs.back().size() != need ) die( "input size mismatch" );
    }

    ImageBase* img = float_mode ? (ImageBase*)new ImageFloat : (ImageBase*)new Image;
    Pyramid*   pyr = new Pyramid( conf, w0, h0 );
    cudaDeviceSynchronize();

    FeaturesHost* last = 0;
    std::chrono::steady_clock::time_point t0;
    long total_feat = 0, total_desc = 0;
    for( int it = 0; it < warmup + iters; it++ ) {
        if( it == warmup ) { cudaDeviceSynchronize(); t0 = std::chrono::steady_clock::now(); }
        std::vector<unsigned char>& im = imgs[ it % imgs.size() ];
        img->resetDimensions( w, h );          /* SiftJob::setImg */
        img->load( im.data() );
        pyr->step1( conf, img );
        if( debug_stages ) {
            auto chk = [&]( const char* what ) {
                cudaError_t e = cudaDeviceSynchronize();
                if( e == cudaSuccess ) e = cudaGetLastError();
                fprintf( stderr, "stage %-12s : %s\n", what, cudaGetErrorString( e ) );
                if( e != cudaSuccess ) exit( 4 );
            };
            chk( "pyramid" );
            pyr->find_extrema( conf ); chk( "find_extrema" );
            ExtremaCounters c; cudaMemcpyFromSymbol( &c, dct, sizeof(c) );
            for( int o = 0; o < conf.octaves; o++ ) {
                int n = c.ext_ct[o];
                std::vector<int> off( std::max( n, 1 ) ); std::vector<InitialExtremum> ie( std::max( n, 1 ) );
                cudaMemcpy( off.data(), dobuf_shadow.i_ext_off[o], n * sizeof(int), cudaMemcpyDeviceToHost );
                cudaMemcpy( ie.data(), dobuf_shadow.i_ext_dat[o], n * sizeof(InitialExtremum), cudaMemcpyDeviceToHost );
                int bad = 0; for( int i = 0; i < n; i++ ) if( off[i] < 0 || off[i] >= n ) bad++;
                fprintf( stderr, "  octave %d: ext_ct %d, bad i_ext_off %d", o, n, bad );
                for( int i = 0; i < std::min( n, 3 ); i++ ) fprintf( stderr, " | off %d x %.2f y %.2f l %d s %.3f", off[i], ie[i].xpos, ie[i].ypos, ie[i].lpos, ie[i].sigma );
                fprintf( stderr, "\n" );
            }
            if( ori_o >= 0 ) {
                /* run the reference's ori_par on ONE extremum (fault bisection) */
                int one = ori_i;
                cudaMemcpy( dobuf_shadow.i_ext_off[ori_o], &one, sizeof(int), cudaMemcpyHostToDevice );
                ExtremaCounters c1 = c; c1.ext_ct[ori_o] = 1;
                cudaMemcpyToSymbol( dct, &c1, sizeof(c1) );
                Octave& oc = pyr->getOctave( ori_o );
                popsift::ori_par<<< 1, 32 >>>( ori_o, 0, oc.getDataTexPoint(), oc.getWidth(), oc.getHeight() );
                chk( "ori_par(one)" );
                Extremum e; cudaMemcpy( &e, dobuf_shadow.extrema, sizeof(e), cudaMemcpyDeviceToHost );
                fprintf( stderr, "  -> num_ori %d ori %.5f %.5f %.5f %.5f\n", e.num_ori, e.orientation[0], e.orientation[1], e.orientation[2], e.orientation[3] );
                exit( 0 );
            }
            pyr->orientation( conf );  chk( "orientation" );
            pyr->descriptors( conf );  chk( "descriptors" );
        } else
        pyr->step2( conf );
        FeaturesHost* fh = pyr->get_descriptors( conf );
        cudaDeviceSynchronize();
        if( it >= warmup ) { total_feat += fh->getFeatureCount(); total_desc += fh->getDescriptorCount(); }
        if( last ) delete last;
        last = fh;
    }
    cudaDeviceSynchronize();
    double sec = std::chrono::duration<double>( std::chrono::steady_clock::now() - t0 ).count();
    cudaError_t err = cudaGetLastError();
    if( err != cudaSuccess ) { fprintf( stderr, "ref_driver: CUDA error %s\n", cudaGetErrorString( err ) ); return 3; }

    if( !out_prefix.empty() ) {
        write_features( out_prefix + ".feat.bin", pyr, last, w0, h0 );
        if( dump_planes ) write_planes( out_prefix + ".planes.bin", pyr );
    }

    printf( "{\"impl\": \"reference-cuda\", \"w\": %d, \"h\": %d, \"octaves\": %d, \"iters\": %d, \"warmup\": %d, "
            "\"seconds\": %.6f, \"ms_per_frame\": %.6f, \"images_per_sec\": %.3f, "
            "\"features_last\": %d, \"descriptors_last\": %d, \"features_total\": %ld, \"descriptors_total\": %ld}\n",
            w, h, conf.octaves, iters, warmup, sec, 1e3 * sec / iters, iters / sec,
            last->getFeatureCount(), last->getDescriptorCount(), total_feat, total_desc );

    delete last;
    delete pyr;
    delete img;
    return 0;
}
