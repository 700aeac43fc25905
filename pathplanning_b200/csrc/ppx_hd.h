// Host/device portability shim for the arithmetic headers of libppx.
//
// The PRODUCT is CUDA only: libppx.so is built by nvcc from the .cu files and has no CPU code
// path.  The same arithmetic headers can additionally be compiled by g++ (PPX_HOST_TWIN) into a
// test-only "twin" library (oracle/twin.cpp) so that the control logic of the kernels (state
// machines, tie orders, the reference's parity traps) can be debugged against the compiled
// reference in a container without a GPU.  Nothing in pathplanning_b200/ loads the twin.
#pragma once
#include <cstdint>
#include <climits>
#include <cmath>
#ifdef __CUDACC__
#include <cuda_runtime.h>
#define PPX_HD __device__ __forceinline__
#define PPX_CONST __device__ __constant__
#define PPX_LDG(p) __ldg(p)
#else
#ifndef PPX_HOST_TWIN
#error "ppx device headers are CUDA-only (define PPX_HOST_TWIN only for the test twin under oracle/)"
#endif
#include <cstring>
#define PPX_HD inline
#define PPX_CONST static const
#define PPX_LDG(p) (*(p))
#define __restrict__
static inline int __double2int_rz(double v) { return (int)v; }
static inline double __longlong_as_double(long long v) { double d; std::memcpy(&d, &v, 8); return d; }
static inline float __int_as_float(int v) { float f; std::memcpy(&f, &v, 4); return f; }
#endif
