// Same include path as the reference (src/include/sens_loc/conversion/depth_to_max_curve.h).
#include "conversions.h"
