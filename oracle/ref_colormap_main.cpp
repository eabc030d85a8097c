// oracle/ref_colormap_main.cpp — TEST INFRASTRUCTURE, not product code.
// Driver around the reference's own exa::ColorMapper (exa/ColorMapper.h:29-83, ColorMapper.cpp), compiled
// in place from /root/reference: control points (t, r, g, b) and sample positions in, colours out.
// usage: ref_colormap in.bin out.bin      in.bin = int32 numPoints; float cp[numPoints*4]; int32 n; float t[n]
#define STB_IMAGE_IMPLEMENTATION
#include "exa/ColorMapper.h"
#include <cstdio>
#include <vector>
using namespace exa;
int main(int ac, char **av)
{
  if (ac != 3) return 1;
  FILE *f = fopen(av[1], "rb"); if (!f) return 1;
  int32_t np = 0, n = 0;
  if (fread(&np, 4, 1, f) != 1) return 1;
  std::vector<float> cp(4 * (size_t)np); if (fread(cp.data(), 4, cp.size(), f) != cp.size()) return 1;
  if (fread(&n, 4, 1, f) != 1) return 1;
  std::vector<float> t(n); if (fread(t.data(), 4, t.size(), f) != t.size()) return 1;
  fclose(f);
  std::vector<ColorMapper::ControlPoint> pts;
  for (int i = 0; i < np; i++) pts.push_back({cp[4 * i], vec3f(cp[4 * i + 1], cp[4 * i + 2], cp[4 * i + 3])});
  ColorMapper cm({ColorMapper::ControlPoint{0.f, vec3f(0.f)}});
  cm.values_ = pts;
  std::vector<float> out(3 * (size_t)n);
  for (int i = 0; i < n; i++) { const vec3f c = cm(t[i]); out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z; }
  FILE *o = fopen(av[2], "wb"); fwrite(out.data(), 4, out.size(), o); fclose(o);
  return 0;
}
