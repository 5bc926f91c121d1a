// TEST INFRASTRUCTURE ONLY (oracle/): see core/mat.hpp in this directory.
#ifndef ORACLE_REF_SHIM_OPENCV2_CORE_HPP
#define ORACLE_REF_SHIM_OPENCV2_CORE_HPP
#include <opencv2/core/mat.hpp>
#endif
