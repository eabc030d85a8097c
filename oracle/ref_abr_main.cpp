// oracle/ref_abr_main.cpp — TEST INFRASTRUCTURE, not product code.
//
// Driver (our own code) around the *reference's own* ABR build
// (exa/Regions.cpp:242-320 ExaBrickRegions::buildFrom), compiled from
// /root/reference in place by oracle/Makefile into oracle/_ref/ref_abr.
// It reproduces the flatten + scalar gather of exa/OptixRenderer.cpp:71-141
// (needed as input to buildFrom) and dumps the result in a flat binary:
//
//   int32 numRegions; int32 leafListLen;
//   BrickRegion[numRegions] (44 B each, exa/Regions.h:31-41);
//   int32 leafList[leafListLen]
//
// usage: ref_abr in.bricks out.abr scalars0 [scalars1 ...]
#include "exa/Regions.h"
#include "exa/ScalarField.h"
#include <cstdio>
#include <cstring>

using namespace exa;

static std::vector<float> readFloats(const char *fn)
{
  FILE *f = fopen(fn, "rb");
  if (!f) { fprintf(stderr, "cannot open %s\n", fn); exit(1); }
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  std::vector<float> v(n / 4);
  if (fread(v.data(), 4, v.size(), f) != v.size()) exit(1);
  fclose(f);
  return v;
}

int main(int ac, char **av)
{
  if (ac < 4) { fprintf(stderr, "usage: ref_abr in.bricks out.abr scalars...\n"); return 1; }
  ExaBricks::SP input = ExaBricks::load(av[1]);
  std::vector<exa::Brick> bricks(input->bricks.size());
  std::vector<int> index(input->totalNumCells);
  size_t off = 0;
  for (size_t i = 0; i < bricks.size(); i++) {
    bricks[i].lower = input->bricks[i]->lower;
    bricks[i].size  = input->bricks[i]->size;
    bricks[i].level = input->bricks[i]->level;
    bricks[i].begin = (uint32_t)off;
    std::copy(input->bricks[i]->cellIDs.begin(), input->bricks[i]->cellIDs.end(), index.begin() + off);
    off += input->bricks[i]->cellIDs.size();
  }
  int nf = ac - 3;
  std::vector<float> scalars(index.size() * nf);
  std::vector<unsigned> offsets(nf);
  for (int f = 0; f < nf; f++) {
    std::vector<float> v = readFloats(av[3 + f]);
    offsets[f] = (unsigned)(f * index.size());
    for (size_t i = 0; i < index.size(); i++) scalars[offsets[f] + i] = v.at(index[i]);
  }
  ExaBrickRegions regions;
  regions.buildFrom(input, bricks.data(), bricks.size(), scalars.data(), offsets.data(), nf);
  static_assert(sizeof(ExaBrickRegions::BrickRegion) == 44, "region layout");
  FILE *o = fopen(av[2], "wb");
  int32_t hdr[2] = { (int32_t)regions.brickRegions.size(), (int32_t)regions.leafList.size() };
  fwrite(hdr, 4, 2, o);
  fwrite(regions.brickRegions.data(), 44, regions.brickRegions.size(), o);
  fwrite(regions.leafList.data(), 4, regions.leafList.size(), o);
  fclose(o);
  return 0;
}
