// TEST INFRASTRUCTURE ONLY (oracle/): see mat.hpp in this directory.
#ifndef ORACLE_REF_SHIM_OPENCV2_CORE_TYPES_HPP
#define ORACLE_REF_SHIM_OPENCV2_CORE_TYPES_HPP
#include <opencv2/core/mat.hpp>
#endif
