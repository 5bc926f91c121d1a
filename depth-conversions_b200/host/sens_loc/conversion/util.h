// Same include path as the reference (src/include/sens_loc/conversion/util.h).
#include "conversions.h"
