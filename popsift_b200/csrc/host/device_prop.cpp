/* popsift::cuda::device_prop_t (reference src/popsift/common/device_prop.cu:24-90). */
#include <popsift/common/device_prop.h>
#include <cuda_runtime.h>
#include <iostream>

namespace popsift { namespace cuda {

device_prop_t::device_prop_t( ) : _num_devices( 0 )
{
    if( cudaGetDeviceCount( &_num_devices ) != cudaSuccess ) { cudaGetLastError(); _num_devices = 0; }
    for( int n = 0; n < _num_devices; n++ ) {
        cudaDeviceProp p;
        _names.push_back( cudaGetDeviceProperties( &p, n ) == cudaSuccess ? std::string( p.name ) : std::string( "?" ) );
    }
}
device_prop_t::~device_prop_t( ) { }

void device_prop_t::print( )
{
    for( int n = 0; n < _num_devices; n++ ) {
        cudaDeviceProp p;
        if( cudaGetDeviceProperties( &p, n ) != cudaSuccess ) continue;
        std::cout << "Device information:" << std::endl
                  << "    Name: " << p.name << std::endl
                  << "    Compute Capability:    " << p.major << "." << p.minor << std::endl
                  << "    Total device mem:      " << p.totalGlobalMem << " B" << std::endl
                  << "    Per-block shared mem:  " << p.sharedMemPerBlock << std::endl
                  << "    Number of SM(x)s:      " << p.multiProcessorCount << std::endl;
    }
}

void device_prop_t::set( int n, bool print_choice )
{
    if( n < 0 || n >= _num_devices ) { std::cerr << "popsift: no CUDA device " << n << std::endl; return; }
    cudaSetDevice( n );
    if( print_choice ) std::cout << "Choosing device " << n << ": " << _names[n] << std::endl;
}

}}
