/* popsift-demo -- command-line front end with the reference's flag names, Boost-free.
 *
 * Reference: src/application/main.cpp (option table :48-149, process_image :170-240, read_job
 * :242-262, main :264-334) and src/application/pgmread.cpp (P2/P3/P5/P6 reader, integer RGB->grey
 * with OpenCV's weights 4899/9617/1868 >> 14, :18-24).  The reference parses with
 * boost::program_options and loads through DevIL or pgmread; here both "--flag value" and
 * "--flag=value" are accepted by a small parser and images are read by the PGM/PPM reader below
 * (so --pgmread-loading is the only loader and is accepted as a no-op).  Output: the reference's
 * "output-features.txt" (Features::print, features.cu:308-328).
 *
 * Everything below the command line goes through the public C++ API (PopSift / SiftJob / Config).
 * 16-bit PGM/PPM samples are read most-significant byte first as the format defines; the
 * reference reads them in host byte order (pgmread.cpp:176, 213).
 */
#include <popsift/popsift.h>
#include <popsift/features.h>
#include <popsift/sift_conf.h>
#include <popsift/common/device_prop.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <queue>
#include <sstream>
#include <string>
#include <vector>
#include <dirent.h>
#include <sys/stat.h>

using namespace std;

static bool print_dev_info = false, print_time_info = false, write_as_uchar = false, dont_write = false, float_mode = false;

static const char* USAGE =
"Options:\n"
"  -h [ --help ]            Print usage\n"
"  -v [ --verbose ]\n"
"  -l [ --log ]             Write debugging files (the reference's dir-octave*/dir-dog*/dir-desc dumps)\n"
"  -i [ --input-file ] arg  Input file (PGM/PPM) or directory\n"
"Parameters:\n"
"  --octaves arg  --levels arg  --sigma arg  --threshold arg  --edge-threshold arg  --edge-limit arg\n"
"  --downsampling arg   Downscale width and height of input by 2^N\n"
"  --initial-blur arg   Assume initial blur, subtract when blurring first time\n"
"Modes:\n"
"  --gauss-mode arg     vlfeat (default), vlfeat-hw-interpolated, vlfeat-direct, opencv, fixed9, fixed15, relative\n"
"  --desc-mode arg      loop (default), iloop, grid, igrid, notile\n"
"  --popsift-mode  --vlfeat-mode  --opencv-mode  --direct-scaling\n"
"  --norm-multi arg  --norm-mode arg (RootSift, classic)  --root-sift\n"
"  --filter-max-extrema arg  --filter-grid arg  --filter-sort arg (random, up, down)\n"
"Informational:\n"
"  --print-gauss-tables  --print-dev-info  --print-time-info  --write-as-uchar  --dont-write\n"
"  --pgmread-loading  --float-mode\n";

[[noreturn]] static void usage_error( const string& what )
{
    cerr << "Error: " << what << endl << endl << "Usage:\n\n" << USAGE << endl;
    exit( EXIT_FAILURE );
}

static void parseargs( int argc, char** argv, popsift::Config& config, string& inputFile )
{
    for( int i = 1; i < argc; i++ ) {
        string a = argv[i], v;
        bool has_v = false;
        if( a.rfind( "--", 0 ) == 0 ) { const size_t eq = a.find( '=' ); if( eq != string::npos ) { v = a.substr( eq + 1 ); a = a.substr( 0, eq ); has_v = true; } }
        auto val = [&]() -> string { if( has_v ) return v; if( i + 1 >= argc ) usage_error( "the required argument for option '" + a + "' is missing" ); return string( argv[++i] ); };
        auto ival = [&]() { const string s = val(); char* e; long r = strtol( s.c_str(), &e, 10 ); if( *e ) usage_error( "the argument ('" + s + "') for option '" + a + "' is invalid" ); return (int)r; };
        auto fval = [&]() { const string s = val(); char* e; float r = strtof( s.c_str(), &e ); if( *e ) usage_error( "the argument ('" + s + "') for option '" + a + "' is invalid" ); return r; };
        if( a == "-h" || a == "--help" ) { cout << USAGE << '\n'; exit( 1 ); }
        else if( a == "-v" || a == "--verbose" ) config.setVerbose();
        else if( a == "-l" || a == "--log" ) config.setLogMode( popsift::Config::All );       /* dir-octave/, dir-dog/, dir-desc/ ... in the working directory */
        else if( a == "-i" || a == "--input-file" ) inputFile = val();
        else if( a == "--octaves" ) config.setOctaves( ival() );
        else if( a == "--levels" ) config.setLevels( ival() );
        else if( a == "--sigma" ) config.setSigma( fval() );
        else if( a == "--threshold" ) config.setThreshold( fval() );
        else if( a == "--edge-threshold" || a == "--edge-limit" ) config.setEdgeLimit( fval() );
        else if( a == "--downsampling" ) config.setDownsampling( fval() );
        else if( a == "--initial-blur" ) config.setInitialBlur( fval() );
        else if( a == "--gauss-mode" ) config.setGaussMode( val() );
        else if( a == "--desc-mode" ) config.setDescMode( val() );
        else if( a == "--popsift-mode" ) config.setMode( popsift::Config::PopSift );
        else if( a == "--vlfeat-mode" ) config.setMode( popsift::Config::VLFeat );
        else if( a == "--opencv-mode" ) config.setMode( popsift::Config::OpenCV );
        else if( a == "--direct-scaling" ) config.setScalingMode( popsift::Config::ScaleDirect );
        else if( a == "--norm-multi" ) config.setNormalizationMultiplier( ival() );
        else if( a == "--norm-mode" ) config.setNormMode( val() );
        else if( a == "--root-sift" ) config.setNormMode( popsift::Config::RootSift );
        else if( a == "--filter-max-extrema" ) config.setFilterMaxExtrema( ival() );
        else if( a == "--filter-grid" ) config.setFilterGridSize( ival() );
        else if( a == "--filter-sort" ) config.setFilterSorting( val() );
        else if( a == "--print-gauss-tables" ) config.setPrintGaussTables();
        else if( a == "--print-dev-info" ) print_dev_info = true;
        else if( a == "--print-time-info" ) print_time_info = true;
        else if( a == "--write-as-uchar" ) write_as_uchar = true;
        else if( a == "--dont-write" ) dont_write = true;
        else if( a == "--pgmread-loading" ) { }
        else if( a == "--float-mode" ) float_mode = true;
        else usage_error( "unrecognised option '" + a + "'" );
    }
    if( inputFile.empty() ) usage_error( "the option '--input-file' is required but missing" );
}

/* ---- PGM / PPM (pgmread.cpp) ---- */

static bool next_token( istream& in, string& tok )
{
    tok.clear();
    int c;
    for( ;; ) {
        c = in.get();
        if( c == EOF ) return false;
        if( c == '#' ) { string rest; getline( in, rest ); continue; }
        if( !isspace( c ) ) break;
    }
    do { tok.push_back( (char)c ); c = in.get(); } while( c != EOF && !isspace( c ) );
    return true;                                        /* exactly one whitespace byte consumed after the token */
}

static unsigned char* readPGMfile( const string& filename, int& w, int& h )
{
    ifstream f( filename.c_str(), ios::binary );
    if( !f.is_open() ) { cerr << "File " << filename << " could not be opened for reading" << endl; return 0; }
    string tok;
    if( !next_token( f, tok ) ) { cerr << "File " << filename << " is too short" << endl; return 0; }
    int type = 0;
    if( tok == "P2" ) type = 2; else if( tok == "P3" ) type = 3; else if( tok == "P5" ) type = 5; else if( tok == "P6" ) type = 6;
    else { cerr << "File " << filename << " can only contain P2, P3, P5 or P6 PGM images" << endl; return 0; }
    int maxval = 0;
    string tw, th, tm;
    if( !next_token( f, tw ) || !next_token( f, th ) || !next_token( f, tm ) ) { cerr << "File " << filename << " is too short" << endl; return 0; }
    w = atoi( tw.c_str() ); h = atoi( th.c_str() ); maxval = atoi( tm.c_str() );
    if( w <= 0 || h <= 0 ) { cerr << "File " << filename << " has meaningless image size" << endl; return 0; }
    if( maxval <= 0 || maxval > 65535 ) { cerr << "File " << filename << " has a meaningless max value" << endl; return 0; }
    const size_t n = (size_t)w * h, ch = ( type == 3 || type == 6 ) ? 3 : 1;
    vector<unsigned char> raw( n * ch );
    auto scale = [&]( unsigned v ) { return maxval == 255 ? (unsigned char)v : (unsigned char)( v * 255.0 / maxval ); };
    if( type == 2 || type == 3 ) {
        for( size_t i = 0; i < n * ch; i++ ) { int v; f >> v; if( f.fail() ) { cerr << "File " << filename << " file too short" << endl; return 0; } raw[i] = scale( (unsigned)v ); }
    } else if( maxval < 256 ) {
        f.read( (char*)raw.data(), (streamsize)( n * ch ) );
        if( f.gcount() != (streamsize)( n * ch ) ) { cerr << "File " << filename << " file too short" << endl; return 0; }
        if( maxval != 255 && ch == 1 ) { }              /* the reference passes sub-255 binary grey data through unscaled (pgmread.cpp:171) */
    } else {
        vector<unsigned char> two( n * ch * 2 );
        f.read( (char*)two.data(), (streamsize)two.size() );
        if( f.gcount() != (streamsize)two.size() ) { cerr << "File " << filename << " file too short" << endl; return 0; }
        for( size_t i = 0; i < n * ch; i++ ) raw[i] = (unsigned char)( ( ( two[2 * i] << 8 ) | two[2 * i + 1] ) * 255.0 / maxval );
    }
    unsigned char* out = new unsigned char[n];
    if( ch == 1 ) memcpy( out, raw.data(), n );
    else for( size_t i = 0; i < n; i++ ) {              /* RGB2GRAY_IN_INT, pgmread.cpp:18-24 */
        const unsigned r = raw[3 * i], g = raw[3 * i + 1], b = raw[3 * i + 2];
        out[i] = (unsigned char)( ( 4899u * r + 9617u * g + 1868u * b ) >> 14 );
    }
    return out;
}

static void collectFilenames( list<string>& files, const string& dir )
{
    DIR* d = opendir( dir.c_str() );
    if( !d ) return;
    vector<string> names;
    while( dirent* e = readdir( d ) ) { const string n = e->d_name; if( n != "." && n != ".." ) names.push_back( dir + "/" + n ); }
    closedir( d );
    for( auto& p : names ) {
        struct stat st;
        if( stat( p.c_str(), &st ) ) continue;
        if( S_ISREG( st.st_mode ) ) files.push_back( p ); else if( S_ISDIR( st.st_mode ) ) collectFilenames( files, p );
    }
}

static SiftJob* process_image( const string& inputFile, PopSift& sift )
{
    int w, h;
    unsigned char* image_data = readPGMfile( inputFile, w, h );
    if( image_data == 0 ) exit( -1 );
    cout << "Loading " << w << " x " << h << " image " << inputFile << endl;
    SiftJob* job;
    if( !float_mode ) job = sift.enqueue( w, h, image_data );
    else {
        vector<float> fdata( (size_t)w * h );
        for( size_t i = 0; i < fdata.size(); i++ ) fdata[i] = float( image_data[i] ) / 256.0f;       /* main.cpp:229-233 */
        job = sift.enqueue( w, h, fdata.data() );
    }
    delete [] image_data;
    return job;
}

static void read_job( SiftJob* job, bool really_write )
{
    popsift::Features* feature_list = job->get();
    cerr << "Number of feature points: " << feature_list->getFeatureCount()
         << " number of feature descriptors: " << feature_list->getDescriptorCount() << endl;
    if( really_write ) {
        ofstream of( "output-features.txt" );
        feature_list->print( of, write_as_uchar );
    }
    delete feature_list;
}

int main( int argc, char** argv )
{
    popsift::Config config;
    list<string> inputFiles;
    string inputFile;
    try { parseargs( argc, argv, config, inputFile ); cout << inputFile << endl; }
    catch( std::exception& e ) { cout << e.what() << endl; exit( 1 ); }

    struct stat st;
    if( stat( inputFile.c_str(), &st ) == 0 ) {
        if( S_ISDIR( st.st_mode ) ) {
            collectFilenames( inputFiles, inputFile );
            if( inputFiles.empty() ) { cerr << "No files in directory, nothing to do" << endl; exit( 0 ); }
        } else if( S_ISREG( st.st_mode ) ) inputFiles.push_back( inputFile );
        else { cout << "Input file is neither regular file nor directory, nothing to do" << endl; exit( -1 ); }
    } else { cerr << "File " << inputFile << " does not exist" << endl; exit( -1 ); }

    popsift::cuda::device_prop_t deviceInfo;
    deviceInfo.set( 0, print_dev_info );
    if( print_dev_info ) deviceInfo.print();

    PopSift sift( config, popsift::Config::ExtractingMode, float_mode ? PopSift::FloatImages : PopSift::ByteImages );
    if( sift.deviceCount() == 0 ) return EXIT_FAILURE;          /* no CUDA device, or the configuration was rejected: said on stderr */

    std::queue<SiftJob*> jobs;
    for( auto& f : inputFiles ) jobs.push( process_image( f, sift ) );
    while( !jobs.empty() ) {
        SiftJob* job = jobs.front(); jobs.pop();
        if( job ) { read_job( job, !dont_write ); delete job; }
    }
    sift.uninit();
    return 0;
}
