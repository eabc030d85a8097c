// tests/harness/surfaces_host.cpp — TEST HARNESS ONLY.  Compiles the product's surface geometry
// (owlexabrick_b200/csrc/surfaces.cuh, host+device functions) with g++ so that the arithmetic the
// frame kernel runs can be compared with the oracle on the CPU.  Never linked into the library.
#include "../../owlexabrick_b200/csrc/surfaces.cuh"
using namespace exa::surf;
extern "C" {
int surf_rounded_cone(const float pa[3], const float pb[3], float ra, float rb, const float ro[3], const float rd[3], float *t, float n[3])
{
  float3 nn = mk(0, 0, 0); float tt = 0.f;
  const bool hit = roundedCone(mk(pa[0], pa[1], pa[2]), mk(pb[0], pb[1], pb[2]), ra, rb, mk(ro[0], ro[1], ro[2]), mk(rd[0], rd[1], rd[2]), tt, nn);
  *t = tt; n[0] = nn.x; n[1] = nn.y; n[2] = nn.z;
  return hit ? 1 : 0;
}
float surf_contour_plane_t(const float ro[3], const float rd[3], const float normal[3], float offset, const float wlo[3], const float whi[3],
                           float *pts18, int *cnt)
{
  float3 pts[6];
  const int c = planePolygon(mk(normal[0], normal[1], normal[2]), offset, wlo, whi, pts);
  if (cnt) *cnt = c;
  if (pts18) for (int i = 0; i < c; i++) { pts18[3 * i] = pts[i].x; pts18[3 * i + 1] = pts[i].y; pts18[3 * i + 2] = pts[i].z; }
  return contourPlaneT(mk(ro[0], ro[1], ro[2]), mk(rd[0], rd[1], rd[2]), mk(normal[0], normal[1], normal[2]), offset, wlo, whi);
}
void surf_ao_direction(const float normal[3], float u1, float u2, float out[3])
{
  const float3 d = aoDirection(mk(normal[0], normal[1], normal[2]), u1, u2);
  out[0] = d.x; out[1] = d.y; out[2] = d.z;
}
}
