// TEST INFRASTRUCTURE (oracle reference build): stand-in for the absent <optix.h>; see ref_shim_base.h
#pragma once
#include "ref_shim_base.h"

// Device intrinsics used by kernel/rt.h:25,53 and kernel/Payload.h:61-66. Defined in ref_api.cpp.
void optixTrace(OptixTraversableHandle handle, float3 ray_origin, float3 ray_direction, float tmin, float tmax, float ray_time,
                OptixVisibilityMask mask, unsigned int ray_flags, unsigned int sbt_offset, unsigned int sbt_stride,
                unsigned int miss_index, unsigned int& p0, unsigned int& p1);
unsigned int optixGetPayload_0();
unsigned int optixGetPayload_1();
