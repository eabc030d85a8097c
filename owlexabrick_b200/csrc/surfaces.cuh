// surfaces.cuh — geometry of the surface pass of the frame kernel (render.cu): streamline
// cones, contour planes and the ambient-occlusion sample frame.  Pure fp32 functions without
// memory traffic; host+device so that tests/test_surfaces_host.py can run the very same
// arithmetic under g++ against the oracle (the product only ever calls them from kernels).
//
//   rounded cone     programs/exabrick.cu:441-501   (intersectRoundedCone)
//   contour plane    programs/exabrick.cu:1267-1394 (intersectLinePlane, intersectBoxPlane,
//                    intersectRayTriangle, polygon part of traceContourRay)
//   AO sample frame  programs/exabrick.cu:78-94     (make_orthonormal_basis,
//                    cosine_sample_hemisphere)
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#ifdef __CUDACC__
#define EXA_HD __host__ __device__ __forceinline__
#else
#define EXA_HD inline
#endif

namespace exa { namespace surf {

EXA_HD float3 mk(float x, float y, float z) { float3 r; r.x = x; r.y = y; r.z = z; return r; }
EXA_HD float3 add(float3 a, float3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
EXA_HD float3 sub(float3 a, float3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
EXA_HD float3 mul(float3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
EXA_HD float dotp(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
EXA_HD float3 crossp(float3 a, float3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
EXA_HD float len(float3 a) { return sqrtf(dotp(a, a)); }
EXA_HD float3 unit(float3 a) { return mul(a, 1.f / sqrtf(dotp(a, a))); }      // owl: v * rsqrt(dot(v,v))

// Ray / capsule-with-two-radii test as the reference does it: the origin is first moved to within
// reach of the primitive (the quadratic is unstable for far origins) and the distance added back.
// Only the cap around pa is tested, as in the reference.  `n` is not normalised.
EXA_HD bool roundedCone(float3 pa, float3 pb, float ra, float rb, float3 ro, float3 rd, float &tHit, float3 &n)
{
  const float shift = fmaxf(0.f, fminf(len(sub(pa, ro)) - ra, len(sub(pb, ro)) - rb));
  ro = add(ro, mul(rd, shift));
  const float3 ba = sub(pb, pa), oa = sub(ro, pa);
  const float rr = ra - rb;
  const float m0 = dotp(ba, ba), m1 = dotp(ba, oa), m2 = dotp(ba, rd), m3 = dotp(rd, oa), m5 = dotp(oa, oa);
  const float d2 = m0 - rr * rr;
  const float k2 = d2 - m2 * m2;
  const float k1 = d2 * m3 - m1 * m2 + m2 * rr * ra;
  const float k0 = d2 * m5 - m1 * m1 + m1 * rr * ra * 2.0f - m0 * ra * ra;
  const float h = k1 * k1 - k0 * k2;
  if (h < 0.0f) return false;
  float t = (-sqrtf(h) - k1) / k2;
  const float y = m1 - ra * rr + t * m2;
  if (y > 0.0f && y < d2) {                                    // lateral surface
    tHit = shift + t;
    n = sub(mul(add(oa, mul(rd, t)), d2), mul(ba, y));
    return true;
  }
  const float h1 = m3 * m3 - m5 + ra * ra;                      // sphere around pa
  if (h1 > 0.0f) {
    t = -m3 - sqrtf(h1);
    tHit = shift + t;
    n = mul(add(oa, mul(rd, t)), 1.f / ra);
    return true;
  }
  return false;
}

// parameter along the edge p1->p2 at which it meets the plane dot(normal,x) = offset, or -1
EXA_HD float linePlane(float3 p1, float3 p2, float3 normal, float offset)
{
  const float3 e = sub(p2, p1);
  const float sdn = dotp(normal, unit(e));
  if (sdn == 0.f) return -1.f;
  const float t = (offset - dotp(normal, p1)) / sdn;
  if (t < 0.f || t > len(e)) return -1.f;
  return t;
}

// Polygon in which the plane cuts the UNIT cube, scaled to [wlo,whi] and ordered around
// pts[0] (the reference's selection sort on the sign of cross(.,.)·normal).  Edge enumeration:
// four corners of the cube that share no edge, three edges each.  Returns the point count (<= 6).
EXA_HD int planePolygon(float3 normal, float offset, const float wlo[3], const float whi[3], float3 pts[6])
{
  int cnt = 0;
  for (int i = 0; i < 4 && cnt < 6; ++i) {
    // key corners (0,0,0) (1,0,1) (1,1,0) (0,1,1)
    const float kx = (i == 1 || i == 2) ? 1.f : 0.f, ky = (i >= 2) ? 1.f : 0.f, kz = (i == 1 || i == 3) ? 1.f : 0.f;
    const float3 p2 = mk(kx, ky, kz);
    for (int j = 0; j < 3 && cnt < 6; ++j) {
      const float3 p1 = mk(j == 0 ? 1.f - kx : kx, j == 1 ? 1.f - ky : ky, j == 2 ? 1.f - kz : kz);
      const float t = linePlane(p1, p2, normal, offset);
      if (t >= 0.f) pts[cnt++] = add(p1, mul(unit(sub(p2, p1)), t));
    }
  }
  for (int i = 0; i < cnt; ++i) {
    pts[i] = mk(pts[i].x * (whi[0] - wlo[0]), pts[i].y * (whi[1] - wlo[1]), pts[i].z * (whi[2] - wlo[2]));
    pts[i] = add(pts[i], mk(wlo[0], wlo[1], wlo[2]));
  }
  for (int i = 0; i < cnt - 1; ++i) {
    int m = i;
    for (int j = i + 1; j < cnt; ++j)
      if (dotp(crossp(sub(pts[j], pts[0]), sub(pts[m], pts[0])), normal) < 0.f) m = j;
    const float3 tmp = pts[i]; pts[i] = pts[m]; pts[m] = tmp;
  }
  return cnt;
}

// ray parameter at which the ray meets triangle (v1, v1+e1, v1+e2); -1 if it misses.  The value
// may be negative for a hit behind the origin (the caller filters, as the reference does).
EXA_HD float rayTriangle(float3 ro, float3 rd, float3 v1, float3 e1, float3 e2)
{
  const float3 s1 = crossp(rd, e2);
  const float div = dotp(s1, e1);
  if (div == 0.f) return -1.f;
  const float inv = 1.f / div;
  const float3 d = sub(ro, v1);
  const float b1 = dotp(d, s1) * inv;
  if (b1 < 0.f || b1 > 1.f) return -1.f;
  const float3 s2 = crossp(d, e1);
  const float b2 = dotp(rd, s2) * inv;
  if (b2 < 0.f || b1 + b2 > 1.f) return -1.f;
  return dotp(e2, s2) * inv;
}

// nearest non-negative hit of the ray with the plane polygon (fan around pts[0]); -1: none
EXA_HD float contourPlaneT(float3 ro, float3 rd, float3 normal, float offset, const float wlo[3], const float whi[3])
{
  float3 pts[6];
  const int cnt = planePolygon(normal, offset, wlo, whi, pts);
  float t = -1.f;
  for (int i = 2; i < cnt; ++i) {
    const float tt = rayTriangle(ro, rd, pts[0], sub(pts[i - 1], pts[0]), sub(pts[i], pts[0]));
    if (tt >= 0.f && (tt < t || t < 0.f)) t = tt;
  }
  return t;
}

// frame (u, v, w) around the unit normal w, and a cosine-distributed direction in it
EXA_HD void orthoBasis(float3 &u, float3 &v, float3 w)
{
  v = fabsf(w.x) > fabsf(w.y) ? unit(mk(-w.z, 0.f, w.x)) : unit(mk(0.f, w.z, -w.y));
  u = crossp(v, w);
}
EXA_HD float3 cosineHemisphere(float u1, float u2)
{
  const float r = sqrtf(u1);
  const float theta = (float)(2.0 * 3.14159265358979323846 * (double)u2);     // 2.f*M_PI*u2 is a double product
  return mk(r * cosf(theta), r * sinf(theta), sqrtf(1.f - u1));
}
EXA_HD float3 aoDirection(float3 normal, float u1, float u2)
{
  float3 u, v; orthoBasis(u, v, normal);
  const float3 sp = cosineHemisphere(u1, u2);
  return unit(add(add(mul(u, sp.x), mul(v, sp.y)), mul(normal, sp.z)));
}

}} // namespace exa::surf
