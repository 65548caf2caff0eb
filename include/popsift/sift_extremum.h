/* <popsift/sift_extremum.h> -- the SIFT descriptor record returned to callers
 * (reference src/popsift/sift_extremum.h:53-59).  The reference's internal InitialExtremum /
 * Extremum device structs are not part of this API; the C ABI exposes an equivalent
 * (psift_extremum, include/psift.h) for parity checks. */
#pragma once
#include "sift_constants.h"

namespace popsift {

struct Descriptor
{
    float features[128];
};

} // namespace popsift
