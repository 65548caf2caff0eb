/* api_demo.cpp -- exercises the drop-in C++ API the way the reference's demo does
 * (reference src/application/main.cpp:240-262, 300-324): Config setters, PopSift, enqueue(),
 * SiftJob::get(), Features iteration, Feature::print.  Compiled against include/popsift only.
 *
 *   api_demo <in.u8> <w> <h> <out.bin> [n_repeat] [flags as popsift-demo: --desc-mode X ...]
 *
 * out.bin: int n_feat, int n_desc, then per feature {octave,x,y,sigma,num_ori,ori[4],desc_idx[4]},
 * then n_desc*128 floats (the layout tests/refio.py reads as FEATREC).
 */
#include <popsift/popsift.h>
#include <popsift/features.h>
#include <popsift/sift_conf.h>
#include <popsift/common/device_prop.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <cmath>
#include <chrono>
#include <sstream>

int main( int argc, char** argv )
{
    if( argc < 5 ) { fprintf( stderr, "usage: api_demo in.u8 w h out.bin [repeat] [flags]\n" ); return 2; }
    const int w = atoi( argv[2] ), h = atoi( argv[3] );
    int repeat = 1, a = 5, tiled = 0, as_float = 0;
    if( argc > 5 && argv[5][0] != '-' ) { repeat = atoi( argv[5] ); a = 6; }

    popsift::Config config;
    for( ; a + 1 < argc; a += 2 ) {
        std::string k = argv[a], v = argv[a + 1];
        if( k == "--desc-mode" ) config.setDescMode( v );
        else if( k == "--gauss-mode" ) config.setGaussMode( v );
        else if( k == "--norm-mode" ) config.setNormMode( v );
        else if( k == "--norm-multi" ) config.setNormalizationMultiplier( atoi( v.c_str() ) );
        else if( k == "--downsampling" ) config.setDownsampling( (float)atof( v.c_str() ) );
        else if( k == "--levels" ) config.setLevels( atoi( v.c_str() ) );
        else if( k == "--tiled" ) tiled = atoi( v.c_str() );
        else if( k == "--float" ) as_float = atoi( v.c_str() );          /* in.u8 is a float32 file: PopSift::FloatImages */
        else if( k == "--log" ) config.setLogMode( v == "all" ? popsift::Config::All : popsift::Config::None );
        else if( k == "--mode" ) config.setMode( v == "vlfeat" ? popsift::Config::VLFeat : v == "opencv" ? popsift::Config::OpenCV : popsift::Config::PopSift );
        else { fprintf( stderr, "unknown flag %s\n", k.c_str() ); return 2; }
    }

    std::vector<unsigned char> img( (size_t)w * h * ( as_float ? sizeof(float) : 1 ) );
    FILE* f = fopen( argv[1], "rb" );
    if( !f || fread( img.data(), 1, img.size(), f ) != img.size() ) { fprintf( stderr, "cannot read %s\n", argv[1] ); return 2; }
    fclose( f );

    popsift::cuda::device_prop_t deviceInfo;
    deviceInfo.set( 0, false );

    PopSift sift( config, popsift::Config::ExtractingMode, as_float ? PopSift::FloatImages : PopSift::ByteImages );

    std::vector<SiftJob*> jobs;
    const auto t_start = std::chrono::steady_clock::now();
    for( int i = 0; i < repeat; i++ )
        jobs.push_back( as_float ? sift.enqueue( w, h, reinterpret_cast<const float*>( img.data() ) ) : sift.enqueue( w, h, img.data() ) );
    const std::vector<unsigned char> keep( img );
    std::fill( img.begin(), img.end(), 0 );          /* the job owns a private copy */

    popsift::Features* last = 0;
    int n_feat0 = -1;
    for( size_t i = 0; i < jobs.size(); i++ ) {
        popsift::Features* feats = jobs[i]->get();
        if( !feats ) { fprintf( stderr, "no features\n" ); return 3; }
        if( n_feat0 < 0 ) n_feat0 = feats->getFeatureCount();
        if( feats->getFeatureCount() != n_feat0 ) { fprintf( stderr, "run-to-run feature count differs\n" ); return 4; }
        delete jobs[i];
        if( last ) delete last;
        last = feats;
    }
    const double secs = std::chrono::duration<double>( std::chrono::steady_clock::now() - t_start ).count();
    if( repeat > 1 ) fprintf( stderr, "%d frames through enqueue()/get() in %.3f s: %.1f frames/s\n", repeat, secs, repeat / secs );
    fprintf( stderr, "Number of feature points: %d, number of feature descriptors: %d, GPUs: %d\n",
             last->getFeatureCount(), last->getDescriptorCount(), sift.deviceCount() );

    if( tiled > 0 ) {
        /* the same image cut into row bands (PopSift::extractTiled): the same keypoints, in another order */
        popsift::Features* t = sift.extractTiled( w, h, keep.data(), tiled );
        if( t->getFeatureCount() != last->getFeatureCount() || t->getDescriptorCount() != last->getDescriptorCount() ) {
            fprintf( stderr, "tiled: %d/%d features/descriptors, un-tiled %d/%d\n", t->getFeatureCount(), t->getDescriptorCount(), last->getFeatureCount(), last->getDescriptorCount() );
            return 5;
        }
        double sa[4] = { 0, 0, 0, 0 }, sb[4] = { 0, 0, 0, 0 };
        for( popsift::Features::F_const_iterator it = last->begin(); it != last->end(); ++it ) { sa[0] += it->xpos; sa[1] += it->ypos; sa[2] += it->sigma; for( int k = 0; k < it->num_ori; k++ ) for( int q = 0; q < 128; q++ ) sa[3] += it->desc[k]->features[q]; }
        for( popsift::Features::F_const_iterator it = t->begin(); it != t->end(); ++it )       { sb[0] += it->xpos; sb[1] += it->ypos; sb[2] += it->sigma; for( int k = 0; k < it->num_ori; k++ ) for( int q = 0; q < 128; q++ ) sb[3] += it->desc[k]->features[q]; }
        for( int k = 0; k < 4; k++ ) if( fabs( sa[k] - sb[k] ) > 1e-6 * ( 1.0 + fabs( sa[k] ) ) ) { fprintf( stderr, "tiled: checksum %d differs (%.9g vs %.9g)\n", k, sb[k], sa[k] ); return 6; }
        fprintf( stderr, "tiled in %d bands: identical keypoint set\n", tiled );
        delete t;
    }

    FILE* o = fopen( argv[4], "wb" );
    int hdr[2] = { last->getFeatureCount(), last->getDescriptorCount() };
    fwrite( hdr, 4, 2, o );
    popsift::Descriptor* base = last->getDescriptors();
    for( popsift::Features::F_const_iterator it = last->begin(); it != last->end(); ++it ) {
        const popsift::Feature& ft = *it;
        int   oi[2] = { ft.debug_octave, ft.num_ori };
        float xyz[3] = { ft.xpos, ft.ypos, ft.sigma };
        int   di[4];
        for( int k = 0; k < ORIENTATION_MAX_COUNT; k++ ) di[k] = ft.desc[k] ? int( ft.desc[k] - base ) : -1;
        fwrite( &oi[0], 4, 1, o ); fwrite( xyz, 4, 3, o ); fwrite( &oi[1], 4, 1, o );
        fwrite( ft.orientation, 4, 4, o ); fwrite( di, 4, 4, o );
    }
    fwrite( base, sizeof(popsift::Descriptor), last->getDescriptorCount(), o );
    fclose( o );

    std::ostringstream txt;                       /* exercise the text writer */
    txt << *last;
    fprintf( stderr, "text output: %zu bytes\n", txt.str().size() );
    delete last;
    sift.uninit();
    return 0;
}
