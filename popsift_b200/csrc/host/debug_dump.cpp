/* debug_dump.cpp -- Config::LogMode::All: the reference's debug dumps, in its on-disk formats.
 *
 * Reference: PopSift::extractDownloadLoop (popsift.cpp:201-209) calls Pyramid::download_and_save_array( "pyramid" )
 * (sift_pyramid.cu:79-83 -> Octave::download_and_save_array, sift_octave.cu:110-187) and Pyramid::save_descriptors
 * (sift_pyramid.cu:88-106, writeDescriptor :400-442); the plane writers are common/write_plane_2d.cu:48-176.
 * Files, relative to the working directory:
 *   dir-octave/pyramid-o-O-l-L.pgm          P2, every value truncated to int, no scaling
 *   dir-octave-dump/pyramid-o-O-l-L.dump    "floats\n<cols> <rows>\n" + raw little-endian fp32
 *   dir-dog/d-pyramid-o-O-l-L.pgm           P2, scaled to 0..255 by (v - min) * 255 / (max - min)
 *   dir-dog-txt/d-pyramid-o-O-l-L.txt       P2, truncated to int, +127
 *   dir-dog-dump/d-pyramid-o-O-l-L.dump     as above
 *   dir-desc/desc-pyramid.txt               x y sigma orientation[deg] d0..d127   per (keypoint, orientation)
 *   dir-fpt/desc-pyramid.txt                x y sigma orientation[deg]
 * The planes come from the context's arena through the parity hooks of the C ABI (psift_debug_read_level / _dog), which
 * are valid between psift_wait and the slot's next submit.  Two quirks of the reference are kept because a regression
 * script would see them: the scaled writer starts its running maximum at numeric_limits<float>::min() (the smallest
 * POSITIVE float), and writeDescriptor scales positions by 2^(octave - up_fac) although the Feature records it reads
 * are already in input-image coordinates.
 */
#include <popsift/sift_conf.h>
#include <popsift/features.h>
#include <psift.h>

#include <cmath>
#include <fstream>
#include <iomanip>
#include <limits>
#include <sstream>
#include <string>
#include <vector>
#include <sys/stat.h>

namespace popsift {

static void make_dir( const char* d )
{
    struct stat st;
    if( stat( d, &st ) == -1 ) mkdir( d, 0700 );
}

/* write_plane2Dunscaled, common/write_plane_2d.cu:113-140 */
static void write_unscaled( const std::string& fn, const float* p, int cols, int rows, int offset )
{
    std::ofstream of( fn.c_str(), std::ios::binary );
    of << "P2" << std::endl << cols << " " << rows << std::endl << "255" << std::endl;
    for( int y = 0; y < rows; y++ ) {
        for( int x = 0; x < cols; x++ ) of << (int)p[(size_t)y * cols + x] + offset << " ";
        of << std::endl;
    }
}

/* write_plane2D, common/write_plane_2d.cu:48-110 */
static void write_scaled( const std::string& fn, const float* p, int cols, int rows )
{
    float minval = std::numeric_limits<float>::max();
    float maxval = std::numeric_limits<float>::min();
    for( size_t i = 0; i < (size_t)cols * rows; i++ ) { minval = std::min( minval, p[i] ); maxval = std::max( maxval, p[i] ); }
    const float fmaxval = 255.0f / ( maxval - minval );
    std::ofstream of( fn.c_str(), std::ios::binary );
    of << "P2" << std::endl << cols << " " << rows << std::endl << "255" << std::endl;
    for( int y = 0; y < rows; y++ ) {
        for( int x = 0; x < cols; x++ ) {
            const float v = ( p[(size_t)y * cols + x] - minval ) * fmaxval;
            of << (int)(unsigned char)v << " ";
        }
        of << std::endl;
    }
}

/* dump_plane2Dfloat, common/write_plane_2d.cu:157-176 */
static void dump_float( const std::string& fn, const float* p, int cols, int rows )
{
    std::ofstream of( fn.c_str(), std::ios::binary );
    of << "floats" << std::endl << cols << " " << rows << std::endl;
    of.write( (const char*)p, (std::streamsize)( sizeof(float) * (size_t)cols * rows ) );
}

/* Pyramid::writeDescriptor, sift_pyramid.cu:400-442 */
static void write_descriptors( const Config& conf, std::ostream& ostr, FeaturesHost* features, bool really )
{
    if( features->getFeatureCount() == 0 ) return;
    const float up_fac = conf.getUpscaleFactor();
    const float pi2 = 2.0f * 3.14159265358979323846f;
    for( int i = 0; i < features->getFeatureCount(); i++ ) {
        const Feature& ext = features->getFeatures()[i];
        const int   octave = ext.debug_octave;
        const float xpos   = ext.xpos  * pow( 2.0f, octave - up_fac );
        const float ypos   = ext.ypos  * pow( 2.0f, octave - up_fac );
        const float sigma  = ext.sigma * pow( 2.0f, octave - up_fac );
        for( int ori = 0; ori < ext.num_ori; ori++ ) {
            float dom_ori = ext.orientation[ori];
            dom_ori = dom_ori / pi2 * 360;
            if( dom_ori < 0 ) dom_ori += 360;
            ostr << std::setprecision( 5 ) << xpos << " " << ypos << " " << sigma << " " << dom_ori << " ";
            if( really && ext.desc[ori] )
                for( int k = 0; k < 128; k++ ) ostr << ext.desc[ori]->features[k] << " ";
            ostr << std::endl;
        }
    }
}

/* called by the worker between psift_wait and psift_job_free of a frame */
void debug_dump_all( void* ctx_, const Config& conf, int n_octaves, int n_levels, FeaturesHost* features, const char* basename )
{
    psift_ctx* ctx = (psift_ctx*)ctx_;
    make_dir( "dir-octave" ); make_dir( "dir-octave-dump" ); make_dir( "dir-dog" ); make_dir( "dir-dog-txt" ); make_dir( "dir-dog-dump" );
    std::vector<float> buf;
    for( int o = 0; o < n_octaves; o++ ) {
        int w = 0, h = 0;
        if( psift_octave_size( ctx, o, &w, &h ) != PSIFT_OK ) break;
        buf.resize( (size_t)w * h );
        for( int l = 0; l < n_levels; l++ ) {
            if( psift_debug_read_level( ctx, o, l, buf.data() ) != PSIFT_OK ) return;
            std::ostringstream a, b;
            a << "dir-octave/" << basename << "-o-" << o << "-l-" << l << ".pgm";
            b << "dir-octave-dump/" << basename << "-o-" << o << "-l-" << l << ".dump";
            write_unscaled( a.str(), buf.data(), w, h, 0 );
            dump_float( b.str(), buf.data(), w, h );
        }
        for( int l = 0; l < n_levels - 1; l++ ) {
            if( psift_debug_read_dog( ctx, o, l, buf.data() ) != PSIFT_OK ) return;
            std::ostringstream a, b, c;
            a << "dir-dog/d-" << basename << "-o-" << o << "-l-" << l << ".pgm";
            b << "dir-dog-txt/d-" << basename << "-o-" << o << "-l-" << l << ".txt";
            c << "dir-dog-dump/d-" << basename << "-o-" << o << "-l-" << l << ".dump";
            write_scaled( a.str(), buf.data(), w, h );
            write_unscaled( b.str(), buf.data(), w, h, 127 );
            dump_float( c.str(), buf.data(), w, h );
        }
    }
    make_dir( "dir-desc" );
    { std::ostringstream fn; fn << "dir-desc/desc-" << basename << ".txt"; std::ofstream of( fn.str().c_str() ); write_descriptors( conf, of, features, true ); }
    make_dir( "dir-fpt" );
    { std::ostringstream fn; fn << "dir-fpt/desc-" << basename << ".txt"; std::ofstream of( fn.str().c_str() ); write_descriptors( conf, of, features, false ); }
}

} // namespace popsift
