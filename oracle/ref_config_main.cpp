// oracle/ref_config_main.cpp — TEST INFRASTRUCTURE, not product code.
//
// Driver around the reference's own config loader (exa/Config.cpp:57-180 parseConfigFile, with
// exa/ScalarField.cpp load / loadAndComputeMagnitude / createFromExpression), compiled from
// /root/reference in place by oracle/Makefile into oracle/_ref/ref_config.  Dumps every scalar field:
//   int32 numFields; per field: int32 nameLen; char name[]; int64 count; float range[2]; float values[count];
//   then float voxelSpaceTransform[12] (vx, vy, vz, p — Config::finalize) and float bounds[6] (Config::getBounds);
//   then the `triangles` meshes (TriangleMesh::load) concatenated: int32 numVerts; float xyz[]; int32 numTris; int32 idx[] (rebased)
// usage: ref_config in.exa out.bin
#include "exa/Config.h"
#include <cstdio>
using namespace exa;
int main(int ac, char **av)
{
  if (ac != 3) { fprintf(stderr, "usage: ref_config in.exa out.bin\n"); return 1; }
  Config::SP cfg = Config::parseConfigFile(av[1]);
  FILE *o = fopen(av[2], "wb");
  int32_t n = (int32_t)cfg->scalarFields.size();
  fwrite(&n, 4, 1, o);
  for (auto &sf : cfg->scalarFields) {
    int32_t nl = (int32_t)sf->name.size(); int64_t cnt = (int64_t)sf->value.size();
    float r[2] = {sf->valueRange.lower, sf->valueRange.upper};
    fwrite(&nl, 4, 1, o); fwrite(sf->name.data(), 1, nl, o); fwrite(&cnt, 8, 1, o); fwrite(r, 4, 2, o);
    fwrite(sf->value.data(), 4, sf->value.size(), o);
  }
  cfg->finalize();
  const affine3f &x = cfg->bricks.voxelSpaceTransform;
  const float m[12] = {x.l.vx.x, x.l.vx.y, x.l.vx.z, x.l.vy.x, x.l.vy.y, x.l.vy.z, x.l.vz.x, x.l.vz.y, x.l.vz.z, x.p.x, x.p.y, x.p.z};
  fwrite(m, 4, 12, o);
  const box3f b = cfg->getBounds();
  const float bb[6] = {b.lower.x, b.lower.y, b.lower.z, b.upper.x, b.upper.y, b.upper.z};
  fwrite(bb, 4, 6, o);
  std::vector<float> v; std::vector<int32_t> ix;
  for (auto &m : cfg->surfaces) {
    const int32_t base = (int32_t)(v.size() / 3);
    for (auto &p : m->vertex) { v.push_back(p.x); v.push_back(p.y); v.push_back(p.z); }
    for (auto &t : m->index) { ix.push_back(t.x + base); ix.push_back(t.y + base); ix.push_back(t.z + base); }
  }
  int32_t nv = (int32_t)(v.size() / 3), nt = (int32_t)(ix.size() / 3);
  fwrite(&nv, 4, 1, o); fwrite(v.data(), 4, v.size(), o); fwrite(&nt, 4, 1, o); fwrite(ix.data(), 4, ix.size(), o);
  fclose(o);
  return 0;
}
