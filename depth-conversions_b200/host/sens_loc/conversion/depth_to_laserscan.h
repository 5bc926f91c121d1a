// Same include path as the reference (src/include/sens_loc/conversion/depth_to_laserscan.h).
#include "conversions.h"
