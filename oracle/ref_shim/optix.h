/* Test-infrastructure shim (NOT product code): the reference's host headers
 * include <optix.h> only for two opaque handle types.  This lets the
 * reference's CPU sources (builder, Regions, loaders) compile in place from
 * /root/reference without the OptiX SDK. */
#pragma once
typedef unsigned long long OptixTraversableHandle;
typedef struct OptixDeviceContext_t *OptixDeviceContext;
