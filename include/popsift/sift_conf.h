/* <popsift/sift_conf.h> -- popsift::Config, source compatible with the reference
 * (reference src/popsift/sift_conf.h:28-320; defaults sift_conf.cu:17-40).
 * Differences: constructing a Config does not touch CUDA (the reference's constructor queries the
 * device and exits without one, sift_conf.cu:41-49), and bad mode strings throw
 * std::invalid_argument instead of calling exit() (sift_conf.cu:76,100). */
#pragma once
#include <string>

#define MAX_OCTAVES   20
#define MAX_LEVELS    10

#if defined(__GNUC__) || defined(__clang__)
#define POPSIFT_DEPRECATED(func) func __attribute__ ((deprecated))
#else
#define POPSIFT_DEPRECATED(func) func
#endif

struct psift_config;

namespace popsift
{

struct Config
{
    Config( );

    enum GaussMode      { VLFeat_Compute, VLFeat_Relative, VLFeat_Relative_All, OpenCV_Compute, Fixed9, Fixed15 };
    enum SiftMode       { PopSift, OpenCV, VLFeat, Default = PopSift };
    enum LogMode        { None, All };
    enum ScalingMode    { ScaleDirect, ScaleDefault };
    enum DescMode       { Loop, ILoop, Grid, IGrid, NoTile };
    enum NormMode       { RootSift, Classic };
    enum GridFilterMode { RandomScale, LargestScaleFirst, SmallestScaleFirst };
    enum ProcessingMode { ExtractingMode, MatchingMode };

    void setGaussMode( const std::string& m );
    void setGaussMode( GaussMode m );
    void setMode( SiftMode m );
    void setLogMode( LogMode mode = All );
    void setScalingMode( ScalingMode mode = ScaleDefault );
    void setVerbose( bool on = true );
    void setDescMode( const std::string& byname );
    void setDescMode( DescMode mode = Loop );
    void setGaussGroup( int groupsize );
    int  getGaussGroup( ) const;

    void setDownsampling( float v );
    void setOctaves( int v );
    void setLevels( int v );
    void setSigma( float v );
    void setEdgeLimit( float v );
    void setThreshold( float v );
    void setInitialBlur( float blur );
    void setMaxExtreme( int m );
    void setPrintGaussTables( );
    void setDPOrientation( bool on );
    void setMaxExtrema( int extrema );
    void setFilterMaxExtrema( int extrema );
    void setFilterGridSize( int sz );
    void setFilterSorting( const std::string& direction );
    void setFilterSorting( GridFilterMode m );

    bool  hasInitialBlur( ) const;
    float getInitialBlur( ) const;
    float getPeakThreshold() const;
    bool  ifPrintGaussTables() const;
    GaussMode getGaussMode( ) const;
    static GaussMode getGaussModeDefault( );
    static const char* getGaussModeUsage( );
    SiftMode getSiftMode() const;
    LogMode  getLogMode() const;

    int      octaves;
    int      levels;
    float    sigma;
    float    _edge_limit;

    void               setNormMode( NormMode m );
    void               setNormMode( const std::string& m );
    POPSIFT_DEPRECATED(void setUseRootSift( bool on ));
    bool               getUseRootSift( ) const;
    NormMode           getNormMode( NormMode m ) const;
    static NormMode    getNormModeDefault( );
    static const char* getNormModeUsage( );

    int  getNormalizationMultiplier( ) const;
    void setNormalizationMultiplier( int mul );

    inline float getUpscaleFactor( ) const    { return _upscale_factor; }
    int  getMaxExtrema( ) const               { return _max_extrema; }
    bool getCanFilterExtrema() const;
    int  getFilterMaxExtrema( ) const         { return _filter_max_extrema; }
    int  getFilterGridSize( ) const           { return _filter_grid_size; }
    GridFilterMode getFilterSorting() const   { return _grid_filter_mode; }
    inline ScalingMode getScalingMode() const { return _scaling_mode; }
    inline DescMode getDescMode() const       { return _desc_mode; }

    bool equal( const Config& other ) const;

    /* extension: the POD mirror that crosses the C ABI (include/psift.h) */
    void toC( psift_config* out ) const;

private:
    float          _threshold;
    float          _upscale_factor;
    LogMode        _log_mode;
    ScalingMode    _scaling_mode;
    DescMode       _desc_mode;
    GridFilterMode _grid_filter_mode;
public:
    bool           verbose;
private:
    int            _max_extrema;
    int            _filter_max_extrema;
    int            _filter_grid_size;
    GaussMode      _gauss_mode;
    SiftMode       _sift_mode;
    bool           _assume_initial_blur;
    float          _initial_blur;
    NormMode       _normalization_mode;
    int            _normalization_multiplier;
    bool           _print_gauss_tables;
    int            _gauss_group;
};

inline bool operator==( const Config& l, const Config& r ) { return l.equal( r ); }
inline bool operator!=( const Config& l, const Config& r ) { return !l.equal( r ); }

} // namespace popsift

/* BASELINE.json's north_star spells the namespace popart:: (an older release); both work. */
namespace popart = popsift;
