/* Feature / FeaturesHost / FeaturesDev (reference src/popsift/features.cu:20-110, 126-302, 308-336). */
#include <popsift/features.h>
#include <psift.h>
#include <cstdio>
#include <mutex>
#include <iostream>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iomanip>
#include <unistd.h>

namespace popsift {

FeaturesBase::FeaturesBase( ) : _num_ext( 0 ), _num_ori( 0 ) { }
FeaturesBase::~FeaturesBase( ) { }

FeaturesHost::FeaturesHost( ) : _ext( 0 ), _ori( 0 ), _block( 0 ) { }
FeaturesHost::FeaturesHost( int num_ext, int num_ori ) : _ext( 0 ), _ori( 0 ), _block( 0 ) { reset( num_ext, num_ori ); }
/* Page-aligned blocks (features.cu:58-82 uses posix_memalign) are recycled through a small process-wide cache: a
 * fresh 1.5 MB block costs the worker thread ~400 page faults per frame, a recycled one none.  Blocks are handed out by
 * size class (256 KB steps), at most 16 are kept, the cache is never torn down (results may outlive every PopSift). */
namespace {
struct BlockCache {
    std::mutex mu;
    struct E { size_t k; void* p; };
    E e[16]; int n = 0;
    static size_t klass( size_t b ) { const size_t g = 256 * 1024; return ( std::max<size_t>( b, 1 ) + g - 1 ) / g * g; }
    void* get( size_t bytes )
    {
        const size_t k = klass( bytes );
        {
            std::lock_guard<std::mutex> lk( mu );
            for( int i = 0; i < n; i++ ) if( e[i].k == k ) { void* p = e[i].p; e[i] = e[--n]; return p; }
        }
        void* p = 0;
        if( posix_memalign( &p, (size_t)sysconf( _SC_PAGESIZE ), k ) != 0 ) return 0;
        return p;
    }
    void put( void* p, size_t bytes )
    {
        if( !p ) return;
        {
            std::lock_guard<std::mutex> lk( mu );
            if( n < 16 ) { e[n].k = klass( bytes ); e[n].p = p; n++; return; }
        }
        free( p );
    }
};
BlockCache& block_cache() { static BlockCache* c = new BlockCache; return *c; }
}

FeaturesHost::~FeaturesHost( )
{
    block_cache().put( _ext, sizeof(Feature) * (size_t)getFeatureCount() );
    if( _block ) psift_result_block_free( (psift_result_block*)_block );
    else         block_cache().put( _ori, sizeof(Descriptor) * (size_t)getDescriptorCount() );
}

FeaturesHost::FeaturesHost( int num_ext, int num_ori, Descriptor* descriptors_in_block, void* block )
    : _ext( 0 ), _ori( descriptors_in_block ), _block( block )
{
    _ext = (Feature*)block_cache().get( sizeof(Feature) * (size_t)num_ext );
    if( !_ext ) {
        std::cerr << "popsift: failed to allocate memory for " << num_ext << " features" << std::endl;
        num_ext = num_ori = 0;
    }
    setFeatureCount( num_ext );
    setDescriptorCount( num_ori );
}

static void* page_alloc( size_t bytes ) { return block_cache().get( bytes ); }
static void  page_free( void* p, size_t bytes ) { block_cache().put( p, bytes ); }

void FeaturesHost::reset( int num_ext, int num_ori )
{
    page_free( _ext, sizeof(Feature) * (size_t)getFeatureCount() );
    if( _block ) { psift_result_block_free( (psift_result_block*)_block ); _block = 0; }
    else page_free( _ori, sizeof(Descriptor) * (size_t)getDescriptorCount() );
    _ext = 0; _ori = 0;
    _ext = (Feature*)page_alloc( sizeof(Feature) * (size_t)num_ext );
    _ori = (Descriptor*)page_alloc( sizeof(Descriptor) * (size_t)num_ori );
    if( !_ext || !_ori ) {
        std::cerr << "popsift: failed to allocate memory for " << num_ext << " features / " << num_ori << " descriptors" << std::endl;
        free( _ext ); free( _ori ); _ext = 0; _ori = 0;
        num_ext = num_ori = 0;
    }
    setFeatureCount( num_ext );
    setDescriptorCount( num_ori );
}

void FeaturesHost::pin( )   { }
void FeaturesHost::unpin( ) { }

void FeaturesHost::print( std::ostream& ostr, bool write_as_uchar ) const
{
    for( int i = 0; i < size(); i++ ) _ext[i].print( ostr, write_as_uchar );
}

std::ostream& operator<<( std::ostream& ostr, const FeaturesHost& feature )
{
    feature.print( ostr, false );
    return ostr;
}

/* ------------------------------------------------------------------ FeaturesDev */

FeaturesDev::FeaturesDev( ) : _handle( 0 ), _ext( 0 ), _ori( 0 ), _rev( 0 ), _device( 0 ) { }
FeaturesDev::FeaturesDev( int num_ext, int num_ori ) : _handle( 0 ), _ext( 0 ), _ori( 0 ), _rev( 0 ), _device( 0 ) { reset( num_ext, num_ori ); }

FeaturesDev::FeaturesDev( psift_dev_features* handle ) : _handle( handle ), _ext( 0 ), _ori( 0 ), _rev( 0 ), _device( 0 )
{
    psift_dev_info info;
    if( handle && psift_dev_features_info( handle, &info ) == PSIFT_OK ) {
        _ext = (Feature*)info.feat_ref; _ori = (Descriptor*)info.desc; _rev = info.rev; _device = info.device;
        setFeatureCount( info.n_feat );
        setDescriptorCount( info.n_desc );
    }
}

FeaturesDev::~FeaturesDev( ) { psift_dev_features_free( _handle ); }

void FeaturesDev::reset( int num_ext, int num_ori )
{
    psift_dev_features_free( _handle );
    _handle = 0; _ext = 0; _ori = 0; _rev = 0;
    setFeatureCount( 0 ); setDescriptorCount( 0 );
    psift_dev_info info;
    if( psift_dev_features_alloc( _device, num_ext, num_ori, &_handle ) != PSIFT_OK || psift_dev_features_info( _handle, &info ) != PSIFT_OK ) {
        std::cerr << "popsift: failed to allocate device memory for " << num_ext << " features / " << num_ori << " descriptors: "
                  << psift_last_error() << std::endl;
        return;
    }
    _ext = (Feature*)info.feat_ref; _ori = (Descriptor*)info.desc; _rev = info.rev;
    setFeatureCount( num_ext );
    setDescriptorCount( num_ori );
}

std::vector<FeaturesDev::Match> FeaturesDev::matchTo( FeaturesDev* other )
{
    std::vector<Match> out;
    const int n = getDescriptorCount();
    if( !other || !_handle || !other->_handle || n <= 0 ) return out;
    std::vector<int> m( 3 * (size_t)n ); std::vector<float> d( 2 * (size_t)n );
    if( psift_match( _handle, other->_handle, m.data(), d.data() ) != PSIFT_OK ) {
        std::cerr << "popsift: match failed: " << psift_last_error() << std::endl;
        return out;
    }
    out.resize( n );
    for( int i = 0; i < n; i++ ) { out[i].best = m[3 * i]; out[i].second = m[3 * i + 1]; out[i].accept = m[3 * i + 2]; out[i].dist_best = d[2 * i]; out[i].dist_second = d[2 * i + 1]; }
    return out;
}

/* the reference prints from a device kernel (show_distance, features.cu:227-266); same text, from the host */
void FeaturesDev::match( FeaturesDev* other )
{
    const std::vector<Match> m = matchTo( other );
    if( m.empty() ) return;
    std::vector<int> lf( getDescriptorCount() ), rf( std::max( other->getDescriptorCount(), 1 ) );
    psift_dev_download( _device, _rev, lf.data(), sizeof(int) * lf.size() );
    psift_dev_download( other->_device, other->_rev, rf.data(), sizeof(int) * (size_t)other->getDescriptorCount() );
    for( size_t i = 0; i < m.size(); i++ )
        printf( "%s feat %4d [%4d] matches feat %4d [%4d] ( 2nd feat %4d [%4d] ) dist %.3f vs %.3f\n",
                m[i].accept ? "accept" : "reject", lf[i], (int)i, rf[m[i].best], m[i].best, rf[m[i].second], m[i].second,
                m[i].dist_best, m[i].dist_second );
}

/* "x y 1/sigma^2 0 1/sigma^2 d0 ... d127", one line per orientation (features.cu:308-328) */
void Feature::print( std::ostream& ostr, bool write_as_uchar ) const
{
    const float sigval = 1.0f / ( sigma * sigma );
    for( int ori = 0; ori < num_ori; ori++ ) {
        ostr << xpos << " " << ypos << " " << sigval << " 0 " << sigval << " ";
        if( write_as_uchar ) {
            for( int i = 0; i < 128; i++ ) ostr << roundf( desc[ori]->features[i] ) << " ";
        } else {
            ostr << std::setprecision( 3 );
            for( int i = 0; i < 128; i++ ) ostr << desc[ori]->features[i] << " ";
            ostr << std::setprecision( 6 );
        }
        ostr << std::endl;
    }
}

std::ostream& operator<<( std::ostream& ostr, const Feature& feature )
{
    feature.print( ostr, false );
    return ostr;
}

} // namespace popsift
