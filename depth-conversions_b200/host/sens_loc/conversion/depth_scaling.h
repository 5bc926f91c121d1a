// Same include path as the reference (src/include/sens_loc/conversion/depth_scaling.h).
#include "conversions.h"
