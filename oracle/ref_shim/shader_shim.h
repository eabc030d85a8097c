/* oracle/ref_shim/shader_shim.h — TEST INFRASTRUCTURE (our code, no reference source).
 *
 * Lets the reference's OWN device source programs/exabrick.cu compile as plain C++ with g++
 * ("shader source that compiles as plain C++"): the handful of OptiX / OWL-device / CUDA names
 * it uses are supplied here on the host.
 *
 *   __device__, __constant__             empty
 *   OPTIX_*_PROGRAM(name)                plain functions  raygen_name / intersect_name / ...
 *   optixGet*, optixReportIntersection   read / update a thread-local "current ray" record
 *   owl::getProgramData / getPRD / getLaunchIndex / getLaunchDims, owl::Ray
 *   owl::traceRay(handle, ray, prd)      closest hit by brute force over the primitives whose
 *                                        bounds program produced a non-empty box (what a BVH
 *                                        built from those boxes returns); the handle value says
 *                                        which geometry (set by the driver)
 *   tex1D<float4>(tex, u)                128-texel float4 table, normalised coordinates, clamp,
 *                                        linear filter (exa/Texture.h:143-147), full-precision weight
 */
#pragma once
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <ctime>
#include <algorithm>
#include <vector>

#define __device__
#define __constant__

#include "owl/owl.h"
#include "owl/common/math/vec.h"
#include "owl/common/math/box.h"
#include "owl/common/math/AffineSpace.h"

#define OPTIX_RAY_FLAG_DISABLE_ANYHIT 1u
#define OPTIX_RAYGEN_PROGRAM(n) void raygen_##n
#define OPTIX_MISS_PROGRAM(n) void miss_##n
#define OPTIX_CLOSEST_HIT_PROGRAM(n) void closesthit_##n
#define OPTIX_INTERSECT_PROGRAM(n) void intersect_##n
#define OPTIX_BOUNDS_PROGRAM(n) void bounds_##n

using std::min;
using std::max;

namespace shim {
  struct Ctx {
    owl::common::vec3f org, dir; float tmin = 0.f, tmax = 0.f;
    int primID = 0, hitPrim = -1; void *prd = nullptr; const void *programData = nullptr;
    owl::common::vec2i launchIndex, launchDims;
  };
  extern thread_local Ctx ctx;
  /* one geometry = its bounds boxes + programs; filled in by the driver */
  struct Geom {
    const void *programData = nullptr;
    void (*intersect)() = nullptr;
    void (*closestHit)() = nullptr;      /* may be null / empty */
    std::vector<owl::common::box3f> bounds;   /* from the bounds program; empty box = culled */
  };
  extern Geom geoms[8];                  /* indexed by the OptixTraversableHandle value */
}

inline int optixGetPrimitiveIndex() { return shim::ctx.primID; }
/* owl's vec_t <-> float3/float4 conversions exist only under nvcc: the shader's `float3` becomes this
 * type (macro at the end of the file), and tex1D returns a float4 stand-in that converts to vec4f */
struct shim_float3 {
  float x, y, z;
  shim_float3() : x(0), y(0), z(0) {}
  shim_float3(const owl::common::vec3f &v) : x(v.x), y(v.y), z(v.z) {}
  operator owl::common::vec3f() const { return owl::common::vec3f(x, y, z); }
};
struct shim_float4 {
  float x, y, z, w;
  operator owl::common::vec4f() const { return owl::common::vec4f(x, y, z, w); }
};
inline shim_float3 optixGetWorldRayOrigin() { return shim_float3(shim::ctx.org); }
inline shim_float3 optixGetWorldRayDirection() { return shim_float3(shim::ctx.dir); }
inline float optixGetRayTmin() { return shim::ctx.tmin; }
inline float optixGetRayTmax() { return shim::ctx.tmax; }
/* accepted iff tmin <= t < tmax; an accepted hit becomes the new tmax.  (OptiX accepts t == tmax too, but which of
 * two primitives with EQUAL t wins then depends on its BVH order; strict < makes the lowest index win.) */
inline bool optixReportIntersection(float t, unsigned)
{
  if (!(t >= shim::ctx.tmin && t < shim::ctx.tmax)) return false;
  shim::ctx.tmax = t;
  shim::ctx.hitPrim = shim::ctx.primID;
  return true;
}

namespace owl {
  template <typename T> inline const T &getProgramData() { return *(const T *)shim::ctx.programData; }
  template <typename T> inline T &getPRD() { return *(T *)shim::ctx.prd; }
  inline common::vec2i getLaunchIndex() { return shim::ctx.launchIndex; }
  inline common::vec2i getLaunchDims() { return shim::ctx.launchDims; }

  struct Ray {
    Ray() {}
    Ray(const common::vec3f &origin, const common::vec3f &direction, float tmin, float tmax)
      : origin(origin), direction(direction), tmin(tmin), tmax(tmax) {}
    common::vec3f origin, direction;
    float tmin = 0.f, tmax = 1e30f, time = 0.f;
  };

  template <typename PRD>
  inline void traceRay(OptixTraversableHandle handle, const Ray &ray, PRD &prd, uint32_t = 0u)
  {
    const shim::Geom &g = shim::geoms[handle];
    const shim::Ctx saved = shim::ctx;
    shim::ctx.org = ray.origin; shim::ctx.dir = ray.direction; shim::ctx.tmin = ray.tmin; shim::ctx.tmax = ray.tmax;
    shim::ctx.prd = &prd; shim::ctx.programData = g.programData; shim::ctx.hitPrim = -1;
    for (size_t i = 0; i < g.bounds.size(); i++) {
      const common::box3f &b = g.bounds[i];
      if (b.lower.x > b.upper.x || b.lower.y > b.upper.y || b.lower.z > b.upper.z) continue;   /* empty: never reported by a BVH */
      shim::ctx.primID = (int)i;
      g.intersect();
    }
    if (g.closestHit && shim::ctx.hitPrim >= 0) {            /* closest-hit program on the winner: tmax is its t */
      shim::ctx.primID = shim::ctx.hitPrim;
      g.closestHit();
    }
    shim::ctx = saved;
  }
}

/* CUDA texture fetch of the transfer function: cudaTextureObject_t holds a pointer to 128 float4 texels */
template <typename T> inline shim_float4 tex1D(cudaTextureObject_t tex, float u)
{
  const float4 *t = (const float4 *)(uintptr_t)tex;
  const float x = u * 128.f - 0.5f;
  const float fl = floorf(x), w = x - fl;
  const int i0 = std::max(0, std::min(127, (int)fl)), i1 = std::max(0, std::min(127, (int)fl + 1));
  const float4 a = t[i0], b = t[i1];
  shim_float4 r;
  r.x = a.x + w * (b.x - a.x); r.y = a.y + w * (b.y - a.y); r.z = a.z + w * (b.z - a.z); r.w = a.w + w * (b.w - a.w);
  return r;
}

/* from here on (the reference's programs/ headers and exabrick.cu) `float3` is the stand-in */
#define float3 shim_float3
