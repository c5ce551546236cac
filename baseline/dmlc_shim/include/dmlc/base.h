/*!
 * Minimal stand-in for dmlc-core's dmlc/base.h, only so that the UNMODIFIED reference
 * (bobjiang82/rabit, which expects dmlc-core next to it) compiles for the baseline arm of
 * bench.py. Written for this repository; it is not part of the product library.
 */
#ifndef DMLC_BASE_H_
#define DMLC_BASE_H_
#include <cstddef>
#include <cstdint>
#define DMLC_USE_CXX11 1
#define DMLC_ENABLE_STD_THREAD 1
#define DMLC_USE_GLOG 0
#define DMLC_LOG_FATAL_THROW 1
#endif  // DMLC_BASE_H_
