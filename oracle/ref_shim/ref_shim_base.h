// ref_shim_base.h — TEST INFRASTRUCTURE (CPU oracle, "reference" build). Not part of the product path.
//
// Lets g++ compile the reference's device headers /root/reference/include/kernel/{rt,BSDFs,disneyBRDF,
// light_sample,cmj,math,Payload,color}.h UNMODIFIED on the host. It supplies only what the snapshot lacks
// (SURVEY §0.1, §0.3): CUDA qualifiers, the vector types/operators of the absent sutil/vec_math.h, the OptiX
// device intrinsics the headers call (optixTrace, optixGetPayload_*), tex2D<float4>, and a reconstructed
// HenjouRenderer/henjouRenderer.h (Params; field list from include/renderer/renderer.h:1175-1227).
// optixTrace / tex2D are routed to the oracle's shared scene code (hjo_common.h, hjo_trace.h).
#pragma once
// system headers first: kernel/cmj.h does `#define uint unsigned int`
#include <math.h>      // the C++ wrapper: brings the float overloads of cos/sin/acos/pow/... into the global
#include <cmath>       // namespace, which is what CUDA device code resolves `cos(float)` to
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>
#include <iostream>
#include <chrono>
#include <omp.h>
#include "../hjo_kernel.h"
#include "../hjo_trace.h"

#define __device__
#define __host__
#define __constant__
#define __forceinline__ inline
#define __forceinline inline

using std::min;
using std::max;

typedef unsigned long long OptixTraversableHandle;
typedef unsigned long long cudaTextureObject_t;
typedef unsigned int OptixVisibilityMask;
enum { OPTIX_RAY_FLAG_NONE = 0 };
