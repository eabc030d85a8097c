// oracle/ref_colormap_main.cpp — TEST INFRASTRUCTURE, not product code.
// Driver around the reference's own exa::ColorMapper (exa/ColorMapper.h:29-83, ColorMapper.cpp), compiled
// in place from /root/reference: control points (t, r, g, b) and sample positions in, colours out.
// usage: ref_colormap in.bin out.bin      in.bin = int32 numPoints; float cp[numPoints*4]; int32 n; float t[n]
//        ref_colormap --embedded out.bin  the 15 embedded maps of exa/embedded_colormaps.h through the reference's PNG constructor
//                                         (stb_image): per map int32 numPoints; float cp[numPoints*4]
//        ref_colormap --string in.txt out.bin   ColorMapper(std::string) -> int32 numPoints; float cp[numPoints*4]
#define STB_IMAGE_IMPLEMENTATION
#include "exa/ColorMapper.h"
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <vector>
#include "exa/embedded_colormaps.h"
using namespace exa;
static void dump(FILE *o, const ColorMapper &cm)
{
  const int32_t n = (int32_t)cm.values_.size(); fwrite(&n, 4, 1, o);
  for (const auto &c : cm.values_) { const float v[4] = {c.first, c.second.x, c.second.y, c.second.z}; fwrite(v, 4, 4, o); }
}
int main(int ac, char **av)
{
  if (ac == 3 && !strcmp(av[1], "--embedded")) {
    FILE *o = fopen(av[2], "wb"); if (!o) return 1;
#define MAP(x) dump(o, ColorMapper(x, sizeof(x)));
    MAP(paraview_cool_warm) MAP(rainbow) MAP(matplotlib_plasma) MAP(matplotlib_virdis) MAP(samsel_linear_green)
    MAP(samsel_linear_ygb_1211g) MAP(cool_warm_extended) MAP(blackbody) MAP(jet) MAP(blue_gold) MAP(ice_fire) MAP(nic_edge)
    MAP(covise) MAP(jamie_draft) MAP(hsv)
#undef MAP
    fclose(o); return 0;
  }
  if (ac == 4 && !strcmp(av[1], "--string")) {
    std::ifstream in(av[2]); std::stringstream ss; ss << in.rdbuf();
    FILE *o = fopen(av[3], "wb"); if (!o) return 1;
    dump(o, ColorMapper(ss.str())); fclose(o); return 0;
  }
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
