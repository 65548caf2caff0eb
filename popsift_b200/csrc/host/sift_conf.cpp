/* popsift::Config -- plain host code (reference src/popsift/sift_conf.cu; defaults :17-40,
 * string spellings :62-101, :131-141, :196-202; getPeakThreshold :275-278; equal :285-303). */
#include <iostream>
#include <cstdlib>
#include <string>
#include <popsift/sift_conf.h>
#include <psift.h>
#include <stdexcept>

namespace popsift {

/* A bad mode string: the reference prints the message and exits (POP_FATAL, sift_conf.cu:76,100,140,201).  Here the
 * message goes to stderr too - so that a caller without a try block still sees why it died - and then
 * std::invalid_argument is thrown; with POPSIFT_FATAL_EXIT=1 in the environment the process exits like the reference. */
[[noreturn]] static void bad_mode( const std::string& msg )
{
    std::cerr << "popsift: " << msg << std::endl;
    const char* e = getenv( "POPSIFT_FATAL_EXIT" );
    if( e && atoi( e ) != 0 ) exit( -1 );
    throw std::invalid_argument( msg );
}


Config::Config( )
    : octaves( -1 ), levels( 3 ), sigma( 1.6f ), _edge_limit( 10.0f )
    , _threshold( 0.04f ), _upscale_factor( 1.0f )
    , _log_mode( Config::None ), _scaling_mode( Config::ScaleDefault ), _desc_mode( Config::Loop )
    , _grid_filter_mode( Config::RandomScale ), verbose( false )
    , _max_extrema( 100000 ), _filter_max_extrema( -1 ), _filter_grid_size( 2 )
    , _gauss_mode( getGaussModeDefault() ), _sift_mode( Config::PopSift )
    , _assume_initial_blur( true ), _initial_blur( 0.5f )
    , _normalization_mode( getNormModeDefault() ), _normalization_multiplier( 0 )
    , _print_gauss_tables( false ), _gauss_group( 1 )
{ }

void Config::setMode( SiftMode m )          { _sift_mode = m; }
void Config::setGaussMode( GaussMode m )    { _gauss_mode = m; }
void Config::setDescMode( DescMode m )      { _desc_mode = m; }
void Config::setLogMode( LogMode m )        { _log_mode = m; }
void Config::setScalingMode( ScalingMode m ){ _scaling_mode = m; }
void Config::setVerbose( bool on )          { verbose = on; }
void Config::setGaussGroup( int g )         { _gauss_group = g; }
int  Config::getGaussGroup( ) const         { return _gauss_group; }

void Config::setDescMode( const std::string& text )
{
    if( text == "loop" )        setDescMode( Loop );
    else if( text == "iloop" )  setDescMode( ILoop );
    else if( text == "grid" )   setDescMode( Grid );
    else if( text == "igrid" )  setDescMode( IGrid );
    else if( text == "notile" ) setDescMode( NoTile );
    else bad_mode( "descriptor extraction mode must be one of loop, iloop, grid, igrid, notile" );
}

void Config::setGaussMode( const std::string& m )
{
    if( m == "vlfeat" )                      setGaussMode( VLFeat_Compute );
    else if( m == "vlfeat-hw-interpolated" ) setGaussMode( VLFeat_Relative );
    else if( m == "relative" )               setGaussMode( VLFeat_Relative );
    else if( m == "vlfeat-direct" )          setGaussMode( VLFeat_Relative_All );
    else if( m == "opencv" )                 setGaussMode( OpenCV_Compute );
    else if( m == "fixed9" )                 setGaussMode( Fixed9 );
    else if( m == "fixed15" )                setGaussMode( Fixed15 );
    else bad_mode( std::string( "Bad Gauss mode. " ) + getGaussModeUsage() );
}

Config::GaussMode Config::getGaussModeDefault( ) { return VLFeat_Compute; }

const char* Config::getGaussModeUsage( )
{
    return "Choice of Gauss filter method. Options are: vlfeat (default), vlfeat-hw-interpolated, "
           "vlfeat-direct, opencv, fixed9, fixed15, relative (synonym for vlfeat-hw-interpolated)";
}

bool Config::getCanFilterExtrema() const { return true; }

void Config::setFilterSorting( const std::string& text )
{
    if( text == "up" )          _grid_filter_mode = SmallestScaleFirst;
    else if( text == "down" )   _grid_filter_mode = LargestScaleFirst;
    else if( text == "random" ) _grid_filter_mode = RandomScale;
    else bad_mode( "filter sorting mode must be one of up, down or random" );
}
void Config::setFilterSorting( GridFilterMode m ) { _grid_filter_mode = m; }

void Config::setNormMode( NormMode m ) { _normalization_mode = m; }
void Config::setNormMode( const std::string& m )
{
    if( m == "RootSift" )     setNormMode( RootSift );
    else if( m == "classic" ) setNormMode( Classic );
    else bad_mode( std::string( "Bad Normalization mode. " ) + getNormModeUsage() );
}
void Config::setUseRootSift( bool on )         { _normalization_mode = on ? RootSift : Classic; }
bool Config::getUseRootSift( ) const           { return _normalization_mode == RootSift; }
Config::NormMode Config::getNormMode( NormMode ) const { return _normalization_mode; }
Config::NormMode Config::getNormModeDefault( ) { return RootSift; }
const char* Config::getNormModeUsage( )
{
    return "Choice of descriptor normalization modes. Options are: RootSift (default), classic";
}

void Config::setNormalizationMultiplier( int mul ) { _normalization_multiplier = mul; }
int  Config::getNormalizationMultiplier( ) const   { return _normalization_multiplier; }

void Config::setDownsampling( float v )     { _upscale_factor = -v; }
void Config::setOctaves( int v )            { octaves = v; }
void Config::setLevels( int v )             { levels = v; }
void Config::setSigma( float v )            { sigma = v; }
void Config::setEdgeLimit( float v )        { _edge_limit = v; }
void Config::setThreshold( float v )        { _threshold = v; }
void Config::setPrintGaussTables()          { _print_gauss_tables = true; }
void Config::setDPOrientation( bool )       { }
void Config::setMaxExtreme( int m )         { _max_extrema = m; }
void Config::setMaxExtrema( int m )         { _max_extrema = m; }
void Config::setFilterMaxExtrema( int ext ) { _filter_max_extrema = ext; }
void Config::setFilterGridSize( int sz )    { _filter_grid_size = sz; }

void Config::setInitialBlur( float blur )
{
    _assume_initial_blur = ( blur != 0.0f );
    _initial_blur        = blur;
}

Config::GaussMode Config::getGaussMode( ) const { return _gauss_mode; }
Config::SiftMode  Config::getSiftMode() const   { return _sift_mode; }
Config::LogMode   Config::getLogMode() const    { return _log_mode; }
bool  Config::hasInitialBlur( ) const           { return _assume_initial_blur; }
float Config::getInitialBlur( ) const           { return _initial_blur; }
float Config::getPeakThreshold() const          { return ( _threshold * 0.5f * 255.0f / levels ); }
bool  Config::ifPrintGaussTables() const        { return _print_gauss_tables; }

bool Config::equal( const Config& o ) const
{
    /* deliberately ignores descriptor mode, filter settings and log mode (sift_conf.cu:285-303) */
    return octaves == o.octaves && levels == o.levels && sigma == o.sigma && _edge_limit == o._edge_limit &&
           _threshold == o._threshold && _upscale_factor == o._upscale_factor && _scaling_mode == o._scaling_mode &&
           _max_extrema == o._max_extrema && _gauss_mode == o._gauss_mode && _sift_mode == o._sift_mode &&
           _assume_initial_blur == o._assume_initial_blur && _initial_blur == o._initial_blur &&
           _normalization_mode == o._normalization_mode && _normalization_multiplier == o._normalization_multiplier;
}

void Config::toC( psift_config* c ) const
{
    psift_config_default( c );
    c->octaves = octaves; c->levels = levels; c->sigma = sigma; c->edge_limit = _edge_limit;
    c->threshold = _threshold; c->upscale_factor = _upscale_factor;
    c->gauss_mode = (int)_gauss_mode; c->sift_mode = (int)_sift_mode; c->desc_mode = (int)_desc_mode;
    c->norm_mode = (int)_normalization_mode; c->norm_multi = _normalization_multiplier;
    c->max_extrema = _max_extrema; c->filter_max_extrema = _filter_max_extrema;
    c->filter_grid_size = _filter_grid_size; c->filter_sorting = (int)_grid_filter_mode;
    c->assume_initial_blur = _assume_initial_blur ? 1 : 0; c->initial_blur = _initial_blur;
    c->scaling_mode = _scaling_mode == ScaleDirect ? 1 : 0;
}

} // namespace popsift
