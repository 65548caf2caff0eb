/* <popsift/sift_constants.h> -- public constants of the drop-in API.
 * Mirrors the macros callers of the reference see (reference src/popsift/sift_constants.h:41-52). */
#pragma once

/* VLFeat keeps at most 4 orientations per keypoint (reference sift_constants.h:45-52) */
#define ORIENTATION_MAX_COUNT 4
