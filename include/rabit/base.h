/*!
 * rabit-b200: build-wide switches shared by every header.
 * Parity: reference include/rabit/base.h:1-19 (feature macros only).
 */
#ifndef RABIT_BASE_H_
#define RABIT_BASE_H_

#include <cstddef>
#include <cstdint>

/*! \brief library version reported by the C API and python module */
#define RABIT_B200_VERSION_MAJOR 0
#define RABIT_B200_VERSION_MINOR 1

/*! \brief C++11 is the floor for this code base; kept as a macro so client code that
 *  tests DMLC_USE_CXX11 (as xgboost does) keeps compiling */
#ifndef DMLC_USE_CXX11
#define DMLC_USE_CXX11 1
#endif

#if defined(__CUDACC__)
#define RABIT_HOST_DEVICE __host__ __device__ __forceinline__
#else
#define RABIT_HOST_DEVICE inline
#endif

#if defined(_WIN32)
#define RABIT_DLL extern "C" __declspec(dllexport)
#else
#define RABIT_DLL extern "C" __attribute__((visibility("default")))
#endif

#endif  // RABIT_BASE_H_
