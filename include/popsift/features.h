/* <popsift/features.h> -- Feature / FeaturesHost containers, source compatible with the
 * reference (reference src/popsift/features.h:22-120).  Feature::desc[i] point into the same
 * FeaturesHost's descriptor block and die with it (reference sift_pyramid.cu:272). */
#pragma once
#include <iostream>
#include <vector>
#include "sift_constants.h"
#include "sift_extremum.h"

struct psift_dev_features;   /* psift.h: the C-ABI object behind FeaturesDev */

namespace popsift {

struct Feature
{
    int         debug_octave;
    float       xpos;
    float       ypos;
    float       sigma;
    int         num_ori;
    float       orientation[ORIENTATION_MAX_COUNT];
    Descriptor* desc[ORIENTATION_MAX_COUNT];

    void print( std::ostream& ostr, bool write_as_uchar ) const;
};

std::ostream& operator<<( std::ostream& ostr, const Feature& feature );

class FeaturesBase
{
    int _num_ext;
    int _num_ori;
public:
    FeaturesBase( );
    virtual ~FeaturesBase( );

    inline int  size() const                { return _num_ext; }
    inline int  getFeatureCount() const     { return _num_ext; }
    inline int  getDescriptorCount() const  { return _num_ori; }
    inline void setFeatureCount( int num_ext )    { _num_ext = num_ext; }
    inline void setDescriptorCount( int num_ori ) { _num_ori = num_ori; }
};

class FeaturesHost : public FeaturesBase
{
    Feature*    _ext;
    Descriptor* _ori;
    void*       _block;   /* page-locked result block that _ori points into (handed over by the extraction), or 0 */
public:
    FeaturesHost( );
    /* descriptors already sit in `block` (a psift_result_block, psift.h): adopt it instead of copying */
    FeaturesHost( int num_ext, int num_ori, Descriptor* descriptors_in_block, void* block );
    FeaturesHost( int num_ext, int num_ori );
    virtual ~FeaturesHost( );

    typedef Feature*       F_iterator;
    typedef const Feature* F_const_iterator;

    inline F_iterator       begin()       { return _ext; }
    inline F_const_iterator begin() const { return _ext; }
    inline F_iterator       end()         { return &_ext[size()]; }
    inline F_const_iterator end() const   { return &_ext[size()]; }

    void reset( int num_ext, int num_ori );
    void pin( );      /* kept for source compatibility: results arrive through the library's own */
    void unpin( );    /* pinned buffers, so these are no-ops (reference features.cu:84-108)       */

    inline Feature*    getFeatures()    { return _ext; }
    inline Descriptor* getDescriptors() { return _ori; }

    void print( std::ostream& ostr, bool write_as_uchar ) const;
};

typedef FeaturesHost Features;

std::ostream& operator<<( std::ostream& ostr, const FeaturesHost& feature );

/* Device-resident results of the matching mode (reference features.h:100-118, features.cu:126-302).  The three
 * pointers are DEVICE memory on device getDevice(); Feature::desc[] of the device records point into the device
 * descriptor block.  match() prints one accept/reject line per descriptor like the reference; matchTo() returns the
 * same result as data. */
class FeaturesDev : public FeaturesBase
{
    ::psift_dev_features* _handle;
    Feature*     _ext;
    Descriptor*  _ori;
    int*         _rev; // the reverse map from descriptors to extrema
    int          _device;

    FeaturesDev( const FeaturesDev& );
    FeaturesDev& operator=( const FeaturesDev& );
public:
    struct Match { int best; int second; int accept; float dist_best; float dist_second; };

    FeaturesDev( );
    FeaturesDev( int num_ext, int num_ori );
    explicit FeaturesDev( ::psift_dev_features* handle );   /* takes ownership (psift_wait_dev) */
    virtual ~FeaturesDev( );

    void reset( int num_ext, int num_ori );

    void match( FeaturesDev* other );
    std::vector<Match> matchTo( FeaturesDev* other );            /* extension */

    inline Feature*    getFeatures()    { return _ext; }
    inline Descriptor* getDescriptors() { return _ori; }
    inline int*        getReverseMap()  { return _rev; }
    inline int         getDevice() const { return _device; }     /* extension */
};

} // namespace popsift
