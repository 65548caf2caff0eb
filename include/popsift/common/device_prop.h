/* <popsift/common/device_prop.h> -- device enumeration helper used by the demo application
 * (reference src/popsift/common/device_prop.h:14-28, device_prop.cu:24-90). */
#pragma once
#include <vector>
#include <string>

namespace popsift { namespace cuda {

class device_prop_t
{
    int _num_devices;
    std::vector<std::string> _names;
public:
    device_prop_t( );
    ~device_prop_t( );
    void print( );
    void set( int n, bool print_choice = false );
    int  count() const { return _num_devices; }     /* extension */
};

}}
