/* Stub <optix.h>: just enough for the reference's src/Parameters.h, src/geometry/Primitives.h and the pure
 * math functions of shaders/tracer.cuh to compile as plain host C++.  OUR file, not the SDK's:
 * no OptiX entry point is declared, so nothing that traces can be built against it.  TEST INFRASTRUCTURE ONLY. */
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <math.h>
#include <cmath>
#include <algorithm>
#include <vector_types.h>
#include <vector_functions.h>
typedef unsigned long long OptixTraversableHandle;
typedef unsigned long long CUdeviceptr;
#define OPTIX_SBT_RECORD_ALIGNMENT 16ull
#define OPTIX_SBT_RECORD_HEADER_SIZE ((size_t)32)
#ifndef __align__
#define __align__(n) __attribute__((aligned(n)))
#endif
using std::min;
using std::max;
