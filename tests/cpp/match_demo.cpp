/* match_demo.cpp -- the reference's matching mode through the drop-in C++ API (reference popsift.cpp:26-36, 215-236;
 * src/application/match.cpp: two images, PopSift(config, MatchingMode), SiftJob::getDev(), FeaturesDev::match).
 *
 *   match_demo <a.u8> <wa> <ha> <b.u8> <wb> <hb> <out.bin>
 *
 * stdout: the accept/reject lines of FeaturesDev::match.  out.bin: int l_len, int r_len, then l_len x {best, second,
 * accept} ints, then l_len x {d1, d2} floats.
 */
#include <popsift/popsift.h>
#include <popsift/features.h>
#include <popsift/sift_conf.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

static bool read_img( const char* fn, int w, int h, std::vector<unsigned char>& img )
{
    img.resize( (size_t)w * h );
    FILE* f = fopen( fn, "rb" );
    const bool ok = f && fread( img.data(), 1, img.size(), f ) == img.size();
    if( f ) fclose( f );
    return ok;
}

int main( int argc, char** argv )
{
    if( argc != 8 ) { fprintf( stderr, "usage: match_demo a.u8 wa ha b.u8 wb hb out.bin\n" ); return 2; }
    std::vector<unsigned char> a, b;
    const int wa = atoi( argv[2] ), ha = atoi( argv[3] ), wb = atoi( argv[5] ), hb = atoi( argv[6] );
    if( !read_img( argv[1], wa, ha, a ) || !read_img( argv[4], wb, hb, b ) ) { fprintf( stderr, "cannot read the images\n" ); return 2; }

    popsift::Config config;
    PopSift sift( config, popsift::Config::MatchingMode, PopSift::ByteImages );
    SiftJob* ja = sift.enqueue( wa, ha, a.data() );
    SiftJob* jb = sift.enqueue( wb, hb, b.data() );
    if( !ja || !jb ) return 3;
    popsift::FeaturesDev* fa = ja->getDev();
    popsift::FeaturesDev* fb = jb->getDev();
    if( !fa || !fb ) { fprintf( stderr, "no device features\n" ); return 3; }
    if( ja->getHost() != 0 ) { fprintf( stderr, "matching mode must not return host features\n" ); return 4; }
    fprintf( stderr, "left: %d features %d descriptors, right: %d features %d descriptors\n",
             fa->getFeatureCount(), fa->getDescriptorCount(), fb->getFeatureCount(), fb->getDescriptorCount() );

    fa->match( fb );                                              /* prints, like the reference */
    const std::vector<popsift::FeaturesDev::Match> m = fa->matchTo( fb );

    FILE* o = fopen( argv[7], "wb" );
    if( !o ) return 2;
    int hdr[2] = { fa->getDescriptorCount(), fb->getDescriptorCount() };
    fwrite( hdr, 4, 2, o );
    for( size_t i = 0; i < m.size(); i++ ) { int r[3] = { m[i].best, m[i].second, m[i].accept }; fwrite( r, 4, 3, o ); }
    for( size_t i = 0; i < m.size(); i++ ) { float d[2] = { m[i].dist_best, m[i].dist_second }; fwrite( d, 4, 2, o ); }
    fclose( o );

    delete fa; delete fb; delete ja; delete jb;
    sift.uninit();
    return 0;
}
