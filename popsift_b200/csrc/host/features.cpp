/* Feature / FeaturesHost (reference src/popsift/features.cu:20-110, 308-336). */
#include <popsift/features.h>
#include <cmath>
#include <cstdlib>
#include <iomanip>
#include <unistd.h>

namespace popsift {

FeaturesBase::FeaturesBase( ) : _num_ext( 0 ), _num_ori( 0 ) { }
FeaturesBase::~FeaturesBase( ) { }

FeaturesHost::FeaturesHost( ) : _ext( 0 ), _ori( 0 ) { }
FeaturesHost::FeaturesHost( int num_ext, int num_ori ) : _ext( 0 ), _ori( 0 ) { reset( num_ext, num_ori ); }
FeaturesHost::~FeaturesHost( ) { free( _ext ); free( _ori ); }

static void* page_alloc( size_t bytes )
{
    void* p = 0;
    if( posix_memalign( &p, (size_t)sysconf( _SC_PAGESIZE ), bytes ? bytes : 1 ) != 0 ) return 0;
    return p;
}

void FeaturesHost::reset( int num_ext, int num_ori )
{
    free( _ext ); free( _ori ); _ext = 0; _ori = 0;
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
