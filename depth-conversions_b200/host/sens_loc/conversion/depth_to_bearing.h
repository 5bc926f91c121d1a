// Same include path as the reference (src/include/sens_loc/conversion/depth_to_bearing.h).
#include "conversions.h"
