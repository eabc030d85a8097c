// oracle/oracle.cpp — TEST INFRASTRUCTURE ONLY.  See oracle/oracle.h for scope and the
// parity-pinning statement.  Every function cites the reference lines it restates
// (paths relative to /root/reference).  Compiled with -ffp-contract=off so that fp32
// expressions are evaluated as written.
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <climits>
#include <set>
#include <stdexcept>
#include <vector>
#include <omp.h>

// ---------------------------------------------------------------------------------
// small math helpers
// ---------------------------------------------------------------------------------
struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline float length(V3 a) { return sqrtf(dot(a, a)); }
// owl: normalize(v) = v * rsqrt(dot(v,v))  (owl/common/math/vec/functors.h)
static inline V3 normalize(V3 a) { return a * (1.f / sqrtf(dot(a, a))); }
static inline float get(V3 a, int d) { return d == 0 ? a.x : d == 1 ? a.y : a.z; }

extern "C" void orc_free(void *p) { free(p); }
extern "C" int orc_num_threads(void) { return omp_get_max_threads(); }

// =================================================================================
// 1. exaBuilder   (builder/builder.cpp)
// =================================================================================
namespace bld {
struct Cell { int x, y, z, level; };
struct Box4 {
  int lo[4] = {INT_MAX, INT_MAX, INT_MAX, INT_MAX};
  int hi[4] = {INT_MIN, INT_MIN, INT_MIN, INT_MIN};
  void extend(const Box4 &o) { for (int i = 0; i < 4; i++) { lo[i] = std::min(lo[i], o.lo[i]); hi[i] = std::max(hi[i], o.hi[i]); } }
  int size(int i) const { return hi[i] - lo[i]; }
};
// SingleCell::getBounds, builder.cpp:121-130
static inline Box4 cellBounds(const Cell &c)
{
  Box4 b; int cw = 1 << c.level;
  b.lo[0] = c.x; b.lo[1] = c.y; b.lo[2] = c.z; b.lo[3] = c.level;
  b.hi[0] = c.x + cw; b.hi[1] = c.y + cw; b.hi[2] = c.z + cw; b.hi[3] = c.level + 1;
  return b;
}
// builder.cpp:164-177 (note: the builder's own 'area' has no factor 2)
static inline size_t unitCellVolume(const Box4 &b) { return (size_t)b.size(0) * (size_t)b.size(1) * (size_t)b.size(2); }
static inline size_t areaOf(const Box4 &b)
{ return (size_t)b.size(0) * (size_t)b.size(1) + (size_t)b.size(1) * (size_t)b.size(2) + (size_t)b.size(2) * (size_t)b.size(0); }
// builder.cpp:53-67
static int divDown(int a, int b) { return a >= 0 ? a / b : (a - (b - 1)) / b; }
static int divUp(int a, int b) { return a >= 0 ? (a + b - 1) / b : a / b; }

struct Out { std::vector<uint8_t> bytes; };

struct Bricker {
  const Cell *cells; int maxLeafWidth; Out *out;

  // builder.cpp:301-350 — sort (cell,id) pairs; cell compared as two LE uint64
  // (builder.cpp:136-143) => key order (y,x) then (level,z) as unsigned 64-bit words.
  std::vector<int> dedup(int64_t n)
  {
    struct P { Cell c; int id; };
    std::vector<P> v(n);
    for (int64_t i = 0; i < n; i++) v[i] = P{cells[i], (int)i};
    auto key0 = [](const Cell &c) { return (uint64_t)(uint32_t)c.x | ((uint64_t)(uint32_t)c.y << 32); };
    auto key1 = [](const Cell &c) { return (uint64_t)(uint32_t)c.z | ((uint64_t)(uint32_t)c.level << 32); };
    std::sort(v.begin(), v.end(), [&](const P &a, const P &b) {
      uint64_t a0 = key0(a.c), b0 = key0(b.c);
      if (a0 != b0) return a0 < b0;
      uint64_t a1 = key1(a.c), b1 = key1(b.c);
      if (a1 != b1) return a1 < b1;
      return a.id < b.id;              // std::pair compares .second last
    });
    // builder.cpp:321-336 "overwrite inner nodes" pass, restated literally
    for (int64_t i = 1; i < n; i++)
      for (int64_t j = i - 1; j >= 0 && v[j].c.x == v[i].c.x && v[j].c.y == v[i].c.y && v[j].c.z == v[i].c.z; --j)
        if (v[j].c.level > v[i].c.level) v[j] = v[i];
    std::vector<int> r;
    if (n == 0) return r;
    r.push_back(v[0].id);
    for (int64_t i = 1; i < n; i++) {
      const Cell &a = v[i].c, &b = v[i - 1].c;
      if (a.x != b.x || a.y != b.y || a.z != b.z || a.level != b.level) r.push_back(v[i].id);
    }
    return r;
  }

  // builder.cpp:201-225
  Box4 coarseBounds(const std::vector<int> &ids)
  {
    Box4 b;
    for (int id : ids) b.extend(cellBounds(cells[id]));
    int cw = 1 << (b.hi[3] - 1);
    for (int d = 0; d < 3; d++) { b.lo[d] = cw * divDown(b.lo[d], cw); b.hi[d] = cw * divUp(b.hi[d], cw); }
    return b;
  }

  // builder.cpp:447-530
  bool tryMakeLeaf(const Box4 &b, const std::vector<int> &ids)
  {
    if (ids.empty()) throw std::runtime_error("empty ID vector in tryMakeLeaf!?");
    if (b.size(3) > 1) return false;
    const size_t cw = 1ULL << (b.hi[3] - 1);
    for (int d = 0; d < 3; d++) if ((size_t)b.size(d) / cw > (size_t)maxLeafWidth) return false;
    const size_t cellVol = cw * cw * cw;
    long long vol = (long long)b.size(0) * b.size(1) * b.size(2) * b.size(3);
    if ((size_t)vol != ids.size() * cellVol) return false;
    int sz[3] = {(int)(b.size(0) / (int)cw), (int)(b.size(1) / (int)cw), (int)(b.size(2) / (int)cw)};
    std::vector<int> cellIDs((size_t)sz[0] * sz[1] * sz[2], -1);
    for (int id : ids) {
      const Cell &c = cells[id];
      int ix = (c.x - b.lo[0]) / (int)cw, iy = (c.y - b.lo[1]) / (int)cw, iz = (c.z - b.lo[2]) / (int)cw;
      cellIDs[ix + (size_t)sz[0] * (iy + (size_t)sz[1] * iz)] = id;
    }
    // .bricks record, builder.cpp:895-902: size, lower, level, cellIDs
    int hdr[7] = {sz[0], sz[1], sz[2], b.lo[0], b.lo[1], b.lo[2], b.lo[3]};
    const uint8_t *p = (const uint8_t *)hdr;
    out->bytes.insert(out->bytes.end(), p, p + sizeof(hdr));
    p = (const uint8_t *)cellIDs.data();
    out->bytes.insert(out->bytes.end(), p, p + cellIDs.size() * 4);
    return true;
  }

  // builder.cpp:538-811, SAH_ALIKE_BUILDER (the default, builder.cpp:873-875)
  void buildRec(std::vector<int> &ids)
  {
    const Box4 cb = coarseBounds(ids);
    if (tryMakeLeaf(cb, ids)) return;
    const int ccw = 1 << (cb.hi[3] - 1);
    int dims[3] = {cb.size(0) / ccw, cb.size(1) / ccw, cb.size(2) / ccw};
    if (dims[0] == 1 && dims[1] == 1 && dims[2] == 1) throw std::runtime_error("coarse size 1 that's not a leaf!?");
    std::vector<size_t> volUsed[3]; std::vector<Box4> sb[3];
    for (int d = 0; d < 3; d++) { volUsed[d].assign(dims[d], 0); sb[d].assign(dims[d], Box4()); }
    for (int id : ids) {                                   // :575-590
      Box4 c = cellBounds(cells[id]);
      for (int d = 0; d < 3; d++) {
        int bin = (c.lo[d] - cb.lo[d]) / ccw;
        volUsed[d][bin] += unitCellVolume(c);
        sb[d][bin].extend(c);
      }
    }
    int bestDim = -1, bestPos = -1;
    double bestCost = std::numeric_limits<float>::infinity();
    for (int d = 0; d < 3; d++) {                          // :607-722
      if (dims[d] == 0) continue;
      const size_t expected = unitCellVolume(cb) / dims[d];
      for (int plane = 1; plane < dims[d]; plane++) {
        const int L = plane - 1, R = plane;
        const bool lFull = volUsed[d][L] == expected, rFull = volUsed[d][R] == expected;
        // :636-643 — an empty slice is never full, so its (overflowing) w-size never
        // decides anything; evaluate the w terms only when both are full.
        bool isBoundary = true;
        if (lFull && rFull)
          isBoundary = !(sb[d][L].lo[3] == sb[d][R].lo[3] && sb[d][L].size(3) == sb[d][R].size(3));
        if (!isBoundary) continue;
        Box4 lb, rb;
        for (int s = 0; s < plane; s++) lb.extend(sb[d][s]);
        for (int s = plane; s < dims[d]; s++) rb.extend(sb[d][s]);
        double cost = areaOf(lb) * (double)unitCellVolume(lb) * lb.size(3)       // :696-697
                    + areaOf(rb) * (double)unitCellVolume(rb) * rb.size(3);
        if (cost < bestCost) { bestCost = cost; bestDim = d; bestPos = cb.lo[d] + plane * ccw; }
      }
    }
    if (bestDim == -1) {                                   // :728-734
      bestDim = 0;
      for (int i = 1; i < 3; i++) if (dims[i] > dims[bestDim]) bestDim = i;
      bestPos = cb.lo[bestDim] + (dims[bestDim] / 2) * ccw;
    }
    std::vector<int> l, r;                                 // :753-770
    for (int id : ids) {
      Box4 c = cellBounds(cells[id]);
      if (c.lo[bestDim] >= bestPos) r.push_back(id);
      else if (c.hi[bestDim] <= bestPos) l.push_back(id);
      else throw std::runtime_error("cell straddles split plane!?");
    }
    if (l.empty() || r.empty()) throw std::runtime_error("invalid split...");
    std::vector<int>().swap(ids);
    buildRec(l);                                           // :808-810 left, then right
    buildRec(r);
  }
};
} // namespace bld

extern "C" int orc_build_bricks(const int32_t *cells, int64_t n, int maxLeafWidth,
                                uint8_t **outp, int64_t *outLen)
{
  try {
    bld::Out out;
    bld::Bricker b{(const bld::Cell *)cells, maxLeafWidth, &out};
    std::vector<int> ids = b.dedup(n);
    if (!ids.empty()) b.buildRec(ids);
    *outLen = (int64_t)out.bytes.size();
    *outp = (uint8_t *)malloc(out.bytes.size() ? out.bytes.size() : 1);
    memcpy(*outp, out.bytes.data(), out.bytes.size());
    return 0;
  } catch (std::exception &e) { fprintf(stderr, "orc_build_bricks: %s\n", e.what()); return 1; }
}

// =================================================================================
// 2. ABR build   (exa/Regions.cpp)
// =================================================================================
namespace abr {
struct Box3 { float lo[3], hi[3]; };
struct Prim { Box3 b; int id; };
struct Builder {
  std::vector<OrcRegion> regions; std::vector<int> leafList;

  // Regions.cpp:32-71
  void addLeaf(const std::vector<Prim> &prims, const Box3 &dom)
  {
    for (int d = 0; d < 3; d++) if (dom.lo[d] >= dom.hi[d]) return;
    std::set<int> ids;
    for (auto &p : prims) ids.insert(p.id);
    if (ids.empty()) return;
    OrcRegion r;
    for (int d = 0; d < 3; d++) { r.lo[d] = dom.lo[d]; r.hi[d] = dom.hi[d]; }
    r.vlo = INFINITY; r.vhi = -INFINITY;
    r.leafListBegin = (int)leafList.size(); r.leafListSize = (int)ids.size();
    r.finestLevelCellWidth = 0.f;
    for (int id : ids) leafList.push_back(id);
    regions.push_back(r);
  }
  // Regions.cpp:73-179
  void buildRec(std::vector<Prim> &prims, const Box3 &dom)
  {
    if (prims.empty()) return;
    for (int i = 0; i < 3; i++) if (dom.hi[i] == dom.lo[i]) return;
    float tgt[3], bestPos[3], bestDist[3];
    for (int d = 0; d < 3; d++) { tgt[d] = (dom.lo[d] + dom.hi[d]) / 2.f; bestPos[d] = dom.lo[d]; bestDist[d] = dom.hi[d] - dom.lo[d]; }
    for (auto &bp : prims)
      for (int d = 0; d < 3; d++)
        for (int side = 0; side < 2; side++) {
          float pos = side ? bp.b.lo[d] : bp.b.hi[d];
          if (pos <= dom.lo[d] || pos >= dom.hi[d]) continue;
          float dist = fabsf(tgt[d] - pos);
          if (dist >= bestDist[d]) continue;
          bestPos[d] = pos; bestDist[d] = dist;
        }
    int widest = 0;                                        // arg_max: first maximum
    for (int i = 1; i < 3; i++) if ((dom.hi[i] - dom.lo[i]) > (dom.hi[widest] - dom.lo[widest])) widest = i;
    int splitDim = -1; float splitPos = 0.f;
    for (int i = 0; i < 3; i++) {
      int d = (widest + i) % 3;
      if (bestPos[d] <= dom.lo[d] || bestPos[d] >= dom.hi[d]) continue;
      splitDim = d; splitPos = bestPos[d]; break;
    }
    if (splitDim < 0) { addLeaf(prims, dom); return; }
    Box3 dl = dom, dr = dom; dl.hi[splitDim] = splitPos; dr.lo[splitDim] = splitPos;
    std::vector<Prim> pl, pr;
    auto clip = [](const Box3 &a, const Box3 &b, Box3 &c) {
      bool ok = true;
      for (int d = 0; d < 3; d++) { c.lo[d] = std::max(a.lo[d], b.lo[d]); c.hi[d] = std::min(a.hi[d], b.hi[d]); ok &= c.lo[d] < c.hi[d]; }
      return ok;
    };
    for (auto &bp : prims) {
      Box3 c;
      if (clip(bp.b, dl, c)) pl.push_back(Prim{c, bp.id});
      if (clip(bp.b, dr, c)) pr.push_back(Prim{c, bp.id});
    }
    std::vector<Prim>().swap(prims);
    buildRec(pr, dr);                                      // :173-178 right first
    buildRec(pl, dl);
  }
};
} // namespace abr

// Brick::getDomain, exa/ExaBricks.h:55-60
static inline void brickDomain(const OrcBrick &b, float lo[3], float hi[3])
{
  const float cw = (float)(1 << b.level);
  for (int d = 0; d < 3; d++) { lo[d] = (float)b.lower[d] - 0.5f * cw; hi[d] = (float)b.lower[d] + ((float)b.size[d] + 0.5f) * cw; }
}

extern "C" int orc_build_abr(const OrcBrick *bricks, int64_t nb, const float *scalars,
                             int64_t numCells, int numFields,
                             OrcRegion **regionsOut, int64_t *nr, int32_t **leafOut, int64_t *nl)
{
  abr::Builder B;
  std::vector<abr::Prim> prims(nb);
  abr::Box3 bounds;
  for (int d = 0; d < 3; d++) { bounds.lo[d] = INFINITY; bounds.hi[d] = -INFINITY; }
  for (int64_t i = 0; i < nb; i++) {                       // Regions.cpp:264-278
    brickDomain(bricks[i], prims[i].b.lo, prims[i].b.hi); prims[i].id = (int)i;
    for (int d = 0; d < 3; d++) { bounds.lo[d] = std::min(bounds.lo[d], prims[i].b.lo[d]); bounds.hi[d] = std::max(bounds.hi[d], prims[i].b.hi[d]); }
  }
  B.buildRec(prims, bounds);
  const int64_t R = (int64_t)B.regions.size();
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t ri = 0; ri < R; ri++) {                     // Regions.cpp:290-300, 182-240
    OrcRegion &r = B.regions[ri];
    int finest = 1 << 30;
    for (int i = 0; i < r.leafListSize; i++) finest = std::min(finest, bricks[B.leafList[r.leafListBegin + i]].level);
    r.finestLevelCellWidth = (float)(1 << finest);
    float vlo = INFINITY, vhi = -INFINITY;
    for (int f = 0; f < numFields; f++)
      for (int i = 0; i < r.leafListSize; i++) {
        const OrcBrick &b = bricks[B.leafList[r.leafListBegin + i]];
        const float cw = (float)(1 << b.level);
        auto valid = [&](int d, int idx) {
          float pos = b.lower[d] + (idx + .5f) * cw;
          return (pos - cw <= r.hi[d]) && (pos + cw >= r.lo[d]);
        };
        for (int iz = 0; iz < b.size[2]; iz++) { if (!valid(2, iz)) continue;
          for (int iy = 0; iy < b.size[1]; iy++) { if (!valid(1, iy)) continue;
            for (int ix = 0; ix < b.size[0]; ix++) { if (!valid(0, ix)) continue;
              size_t idx = (size_t)f * numCells + b.begin + ix + (size_t)b.size[0] * iy + (size_t)b.size[0] * b.size[1] * iz;
              float s = scalars[idx];
              vlo = std::min(vlo, s); vhi = std::max(vhi, s);
            } } }
      }
    r.vlo = vlo; r.vhi = vhi;
  }
  *nr = R; *nl = (int64_t)B.leafList.size();
  *regionsOut = (OrcRegion *)malloc(sizeof(OrcRegion) * (R ? R : 1));
  memcpy(*regionsOut, B.regions.data(), sizeof(OrcRegion) * R);
  *leafOut = (int32_t *)malloc(4 * (B.leafList.size() ? B.leafList.size() : 1));
  memcpy(*leafOut, B.leafList.data(), 4 * B.leafList.size());
  return 0;
}

// =================================================================================
// 3. Scene + tent-basis reconstruction   (programs/exabrick.cu:581-942)
// =================================================================================
struct OrcScene {
  std::vector<OrcBrick> bricks; std::vector<float> scalars; int64_t numCells; int numFields;
  std::vector<OrcRegion> regions; std::vector<int32_t> leafList;
  // uniform grid for point location (oracle-private accelerator; regions are disjoint
  // half-open boxes, so the result equals a brute-force search)
  float glo[3], ghi[3]; int G; std::vector<int> cellStart, cellItems;
  // bounding-volume hierarchy over the region boxes for the closest-region ray query
  // (oracle-private accelerator standing in for the OptiX BVH; returns exactly what the
  // brute-force loop returns, see closestRegion); ORC_BRUTE_FORCE=1 disables it
  struct BvhNode { float lo[3], hi[3]; int left, right, first, count; };
  std::vector<BvhNode> bvh; std::vector<int> bvhItems; bool useBvh;
};

static int buildBvh(OrcScene &s, int first, int count)
{
  OrcScene::BvhNode n; n.left = n.right = -1; n.first = first; n.count = count;
  for (int d = 0; d < 3; d++) { n.lo[d] = INFINITY; n.hi[d] = -INFINITY; }
  for (int i = first; i < first + count; i++) {
    const OrcRegion &r = s.regions[s.bvhItems[i]];
    for (int d = 0; d < 3; d++) { n.lo[d] = std::min(n.lo[d], r.lo[d]); n.hi[d] = std::max(n.hi[d], r.hi[d]); }
  }
  const int id = (int)s.bvh.size(); s.bvh.push_back(n);
  if (count > 4) {
    int axis = 0;
    for (int d = 1; d < 3; d++) if (n.hi[d] - n.lo[d] > n.hi[axis] - n.lo[axis]) axis = d;
    auto key = [&](int ri) { return s.regions[ri].lo[axis] + s.regions[ri].hi[axis]; };
    std::nth_element(s.bvhItems.begin() + first, s.bvhItems.begin() + first + count / 2, s.bvhItems.begin() + first + count,
                     [&](int a, int b) { return key(a) < key(b); });
    const int l = buildBvh(s, first, count / 2), r = buildBvh(s, first + count / 2, count - count / 2);
    s.bvh[id].left = l; s.bvh[id].right = r; s.bvh[id].count = 0;
  }
  return id;
}

extern "C" OrcScene *orc_scene_create(const OrcBrick *bricks, int64_t nb, const float *scalars,
                                      int64_t numCells, int numFields,
                                      const OrcRegion *regions, int64_t nr,
                                      const int32_t *leafList, int64_t nl)
{
  OrcScene *s = new OrcScene;
  s->bricks.assign(bricks, bricks + nb);
  s->scalars.assign(scalars, scalars + numCells * numFields);
  s->numCells = numCells; s->numFields = numFields;
  s->regions.assign(regions, regions + nr);
  s->leafList.assign(leafList, leafList + nl);
  for (int d = 0; d < 3; d++) { s->glo[d] = INFINITY; s->ghi[d] = -INFINITY; }
  for (auto &r : s->regions) for (int d = 0; d < 3; d++) { s->glo[d] = std::min(s->glo[d], r.lo[d]); s->ghi[d] = std::max(s->ghi[d], r.hi[d]); }
  int G = (int)std::min(128.0, std::max(1.0, std::cbrt((double)nr))); s->G = G;
  std::vector<std::vector<int>> lists((size_t)G * G * G);
  for (int64_t ri = 0; ri < nr; ri++) {
    const OrcRegion &r = s->regions[ri]; int a[3], b[3];
    for (int d = 0; d < 3; d++) {
      float w = (s->ghi[d] - s->glo[d]) / G;
      a[d] = std::max(0, std::min(G - 1, (int)floorf((r.lo[d] - s->glo[d]) / w)));
      b[d] = std::max(0, std::min(G - 1, (int)floorf((r.hi[d] - s->glo[d]) / w)));
    }
    for (int z = a[2]; z <= b[2]; z++) for (int y = a[1]; y <= b[1]; y++) for (int x = a[0]; x <= b[0]; x++)
      lists[x + (size_t)G * (y + (size_t)G * z)].push_back((int)ri);
  }
  s->cellStart.resize(lists.size() + 1); s->cellStart[0] = 0;
  for (size_t i = 0; i < lists.size(); i++) s->cellStart[i + 1] = s->cellStart[i] + (int)lists[i].size();
  s->bvhItems.resize(nr); for (int64_t i = 0; i < nr; i++) s->bvhItems[i] = (int)i;
  s->useBvh = nr > 0 && !(getenv("ORC_BRUTE_FORCE") && atoi(getenv("ORC_BRUTE_FORCE")));
  if (s->useBvh) buildBvh(*s, 0, (int)nr);
  s->cellItems.resize(s->cellStart.back());
  for (size_t i = 0; i < lists.size(); i++) std::copy(lists[i].begin(), lists[i].end(), s->cellItems.begin() + s->cellStart[i]);
  return s;
}
extern "C" void orc_scene_destroy(OrcScene *s) { delete s; }

// Point query = traceVolumeRay with origin p, dir (1,1,1), [0,2e-10] (exabrick.cu:818-826,
// 949-954): hits the region with lo <= p < hi.  Miss -> -1 (the reference then indexes
// region -1: undefined behaviour we do not copy; a miss is an invalid sample).
static int locateRegion(const OrcScene &s, V3 p)
{
  int c[3]; float pp[3] = {p.x, p.y, p.z};
  for (int d = 0; d < 3; d++) {
    if (!(pp[d] >= s.glo[d] && pp[d] <= s.ghi[d])) return -1;
    float w = (s.ghi[d] - s.glo[d]) / s.G;
    c[d] = std::max(0, std::min(s.G - 1, (int)floorf((pp[d] - s.glo[d]) / w)));
  }
  // the grid cell of p may be off by one through rounding: search the 3^3 neighbourhood
  for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) {
    int x = c[0] + dx, y = c[1] + dy, z = c[2] + dz;
    if (x < 0 || y < 0 || z < 0 || x >= s.G || y >= s.G || z >= s.G) continue;
    size_t ci = x + (size_t)s.G * (y + (size_t)s.G * z);
    for (int k = s.cellStart[ci]; k < s.cellStart[ci + 1]; k++) {
      const OrcRegion &r = s.regions[s.cellItems[k]];
      if (pp[0] >= r.lo[0] && pp[0] < r.hi[0] && pp[1] >= r.lo[1] && pp[1] < r.hi[1] && pp[2] >= r.lo[2] && pp[2] < r.hi[2])
        return s.cellItems[k];
    }
  }
  return -1;
}

struct Acc { float sumWV = 0.f, sumW = 0.f; V3 sumD = {0, 0, 0}, sumDC = {0, 0, 0}; };

// addBasisFunctions<NEED_DERIVATIVE>, exabrick.cu:620-777 (corner order z-lo{y-lo{x-lo,x-hi},
// y-hi{..}}, z-hi{..}; weight = (wz*wy)*wx; derivative terms as written there).
template <bool DERIV>
static inline void addBasis(const OrcScene &s, Acc &a, int brickID, V3 pos, int channel, OrcStats *st)
{
  const OrcBrick &b = s.bricks[brickID];
  const float cw = (float)(1 << b.level);
  const float lp[3] = {(pos.x - (float)b.lower[0]) / cw - 0.5f, (pos.y - (float)b.lower[1]) / cw - 0.5f, (pos.z - (float)b.lower[2]) / cw - 0.5f};
  int lo[3], hi[3]; float fr[3], nf[3];
  for (int d = 0; d < 3; d++) {
    lo[d] = std::max(-1, (int)floorf(lp[d])); hi[d] = lo[d] + 1;
    fr[d] = lp[d] - (float)lo[d]; nf[d] = 1.f - fr[d];
  }
  const float *sc = s.scalars.data() + (size_t)channel * s.numCells + b.begin;   // getScalar :581-594
  if (st) st->brickVisits++;
  for (int cz = 0; cz < 2; cz++) {
    const int iz = cz ? hi[2] : lo[2]; const float wz = cz ? fr[2] : nf[2];
    if (cz ? !(iz < b.size[2]) : !(iz >= 0 && iz < b.size[2])) continue;
    for (int cy = 0; cy < 2; cy++) {
      const int iy = cy ? hi[1] : lo[1]; const float wy = cy ? fr[1] : nf[1];
      if (cy ? !(iy < b.size[1]) : !(iy >= 0 && iy < b.size[1])) continue;
      for (int cx = 0; cx < 2; cx++) {
        const int ix = cx ? hi[0] : lo[0]; const float wx = cx ? fr[0] : nf[0];
        if (cx ? !(ix < b.size[0]) : !(ix >= 0 && ix < b.size[0])) continue;
        const float scalar = sc[ix + iy * b.size[0] + iz * b.size[0] * b.size[1]];
        const float w = (wz * wy) * wx;
        if (st) st->cornerLoads++;
        if (DERIV) {
          const float dx = (wz * wy) * (cx ? 1.f : -1.f);
          const float dy = (wz * wx) * (cy ? 1.f : -1.f);
          const float dz = (wy * wx) * (cz ? 1.f : -1.f);
          a.sumDC = a.sumDC + v3(dx, dy, dz);
          a.sumD = a.sumD + v3(dx * scalar, dy * scalar, dz * scalar);
        }
        a.sumW += w; a.sumWV += w * scalar;
      }
    }
  }
}

// samplePoint, exabrick.cu:781-806
static inline bool samplePoint(const OrcScene &s, float &value, int leafID, V3 pos, int channel, OrcStats *st)
{
  if (leafID < 0) return false;
  const OrcRegion &r = s.regions[leafID];
  Acc a;
  for (int i = 0; i < r.leafListSize; i++) addBasis<false>(s, a, s.leafList[r.leafListBegin + i], pos, channel, st);
  if (st) st->samples++;
  if (a.sumW <= 1e-20f) return false;
  value = a.sumWV / a.sumW; return true;
}
// samplePointWithDerivative (ANALYTIC_GRADIENTS), exabrick.cu:883-927
static inline bool samplePointD(const OrcScene &s, float &value, V3 &g, int leafID, V3 pos, int channel, OrcStats *st)
{
  if (leafID < 0) return false;
  const OrcRegion &r = s.regions[leafID];
  Acc a;
  for (int i = 0; i < r.leafListSize; i++) addBasis<true>(s, a, s.leafList[r.leafListBegin + i], pos, channel, st);
  if (st) st->samples++;
  if (a.sumW <= 1e-20f) return false;
  value = a.sumWV / a.sumW;
  g = v3(a.sumW * a.sumD.x - a.sumWV * a.sumDC.x, a.sumW * a.sumD.y - a.sumWV * a.sumDC.y, a.sumW * a.sumD.z - a.sumWV * a.sumDC.z);
  return true;
}

static void addStats(OrcStats *dst, const OrcStats &s)
{
  if (!dst) return;
#pragma omp critical
  { dst->samples += s.samples; dst->brickVisits += s.brickVisits; dst->cornerLoads += s.cornerLoads; dst->segments += s.segments; }
}

extern "C" int orc_sample_points(const OrcScene *s, const float *xyz, int64_t n, int channel,
                                 float *value, uint8_t *valid, float *grad, OrcStats *stats)
{
  if (stats) memset(stats, 0, sizeof(*stats));
#pragma omp parallel
  {
    OrcStats st; memset(&st, 0, sizeof(st));
#pragma omp for schedule(static)
    for (int64_t i = 0; i < n; i++) {
      V3 p = v3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      int leaf = locateRegion(*s, p);
      float v = 0.f; V3 g = v3(0, 0, 0); bool ok;
      if (grad) ok = samplePointD(*s, v, g, leaf, p, channel, &st);
      else ok = samplePoint(*s, v, leaf, p, channel, &st);
      value[i] = ok ? v : 0.f; valid[i] = ok ? 1 : 0;
      if (grad) { grad[3 * i] = ok ? g.x : 0.f; grad[3 * i + 1] = ok ? g.y : 0.f; grad[3 * i + 2] = ok ? g.z : 0.f; }
    }
    addStats(stats, st);
  }
  return 0;
}

// =================================================================================
// 4. Frame: transfer function, region stepping, DVR, iso   (programs/exabrick.cu)
// =================================================================================
struct V4 { float x, y, z, w; };

// lookupTransferFunction, exabrick.cu:135-150; texture = 128 float4 texels, normalised
// coordinates, clamp addressing, linear filtering (exa/Texture.h:143-147): the unit
// samples at x = 128*u - 0.5 between texels floor(x) and floor(x)+1.
static inline V4 lookupXF(const OrcFrame &f, float in, int ch)
{
  float sc = 127.f * (in - f.xfDomain[ch][0]) / ((f.xfDomain[ch][1] - f.xfDomain[ch][0]) + 1e-20f);
  sc = fminf(fmaxf(sc + .5f, 0.f), 127.f);
  sc /= 127.f;
  float x = sc * 128.f - 0.5f;
  float fl = floorf(x); float w = x - fl;
  if (f.texFilter8bit) w = floorf(w * 256.f + 0.5f) / 256.f;
  int i0 = std::max(0, std::min(127, (int)fl)), i1 = std::max(0, std::min(127, (int)fl + 1));
  const float *a = f.xf[ch][i0], *b = f.xf[ch][i1];
  V4 r = {a[0] + w * (b[0] - a[0]), a[1] + w * (b[1] - a[1]), a[2] + w * (b[2] - a[2]), a[3] + w * (b[3] - a[3])};
  r.w *= f.xfOpacityScale;
  return r;
}

// activeForVolumeSampling, exabrick.cu:250-281
static bool activeForVolume(const OrcFrame &f, float vlo, float vhi, int ch)
{
  const float dlo = f.xfDomain[ch][0], dhi = f.xfDomain[ch][1];
  if (vlo > dhi) return false;
  if (vhi < dlo) return false;
  const float slo = (vlo - dlo) / ((dhi - dlo) + 1e-20f), shi = (vhi - dlo) / ((dhi - dlo) + 1e-20f);
  auto clampi = [](int v, int lo, int hi) { return std::min(hi, std::max(lo, v)); };
  const int ilo = clampi((int)(slo * 127), 0, 127), ihi = clampi((int)(shi * 127) + 1, 0, 127);
  for (int i = ilo; i <= ihi; i++) {
    float cv = (float)i / 127; cv *= dhi - dlo; cv += dlo;
    if (lookupXF(f, cv, ch).w > 0.f) return true;
  }
  return false;
}

struct Ray { V3 o, d; float tmin, tmax; };

// boxTest, exabrick.cu:197-210
static inline bool boxTest(const Ray &r, const float lo[3], const float hi[3], float &t0, float &t1)
{
  float tlx = (lo[0] - r.o.x) / r.d.x, thx = (hi[0] - r.o.x) / r.d.x;
  float tly = (lo[1] - r.o.y) / r.d.y, thy = (hi[1] - r.o.y) / r.d.y;
  float tlz = (lo[2] - r.o.z) / r.d.z, thz = (hi[2] - r.o.z) / r.d.z;
  float nr = fmaxf(fmaxf(fminf(tlx, thx), fminf(tly, thy)), fminf(tlz, thz));
  float fr = fminf(fminf(fmaxf(tlx, thx), fmaxf(tly, thy)), fmaxf(tlz, thz));
  t0 = fmaxf(r.tmin, nr); t1 = fminf(r.tmax, fr);
  return t0 < t1;
}

// Closest-hit over all active regions = what optixTrace + __intersection__VolumeBVH
// return (exabrick.cu:184-238); brute force.  Ties resolve to the lowest region ID.
static int closestRegion(const OrcScene &s, const std::vector<uint8_t> &active, const Ray &r, float &t0, float &t1)
{
  int best = -1; float bt0 = 0, bt1 = 0;
  if (s.useBvh) {
    // Same answer as the loop below: a region's [a,b) lies inside its ancestors' (fp subtraction
    // and division are monotonic), so a node whose own box test fails, or whose entry is
    // strictly behind the best entry found so far, cannot hold the winner; ties keep the
    // lowest region ID.
    int stack[64], sp = 0; stack[sp++] = 0;
    while (sp) {
      const OrcScene::BvhNode &n = s.bvh[stack[--sp]];
      float a, b;
      if (!boxTest(r, n.lo, n.hi, a, b)) continue;
      if (best >= 0 && a > bt0) continue;
      if (n.left < 0) {
        for (int i = n.first; i < n.first + n.count; i++) {
          const int ri = s.bvhItems[i];
          if (!active[ri]) continue;
          if (!boxTest(r, s.regions[ri].lo, s.regions[ri].hi, a, b)) continue;
          if (best < 0 || a < bt0 || (a == bt0 && ri < best)) { best = ri; bt0 = a; bt1 = b; }
        }
      } else { stack[sp++] = n.left; stack[sp++] = n.right; }
    }
    t0 = bt0; t1 = bt1; return best;
  }
  for (size_t i = 0; i < s.regions.size(); i++) {
    if (!active[i]) continue;
    float a, b;
    if (!boxTest(r, s.regions[i].lo, s.regions[i].hi, a, b)) continue;
    if (best < 0 || a < bt0) { best = (int)i; bt0 = a; bt1 = b; }
  }
  t0 = bt0; t1 = bt1; return best;
}

// gradientCD<ISO> (exabrick.cu:833-880, FAST_CENTRAL_DIFFERENCES off) behind samplePointWithDerivative built with
// ANALYTIC_GRADIENTS=0 (:928-931): six more point samples at pos -+ delta along the axes, each located by its own region
// query over the BVH of the ACTIVE regions (volume BVH, or the iso-surface BVH for ISO) with a ray of length 2e-10
// along (1,1,1); the result is f(pos - delta) - f(pos + delta) per axis (sic).  A position outside every active region
// samples 0.  Decided: a located position whose sample is invalid (sum of weights <= 1e-20) samples 0 too - the
// reference leaves that float uninitialised.
static V3 gradientCD(const OrcScene &s, const std::vector<uint8_t> &active, V3 pos, float delta, int channel)
{
  float sv[2][3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 2; j++) {
      const V3 dt = v3(i == 0 ? delta : 0.f, i == 1 ? delta : 0.f, i == 2 ? delta : 0.f);
      Ray ray; ray.o = j == 0 ? pos + dt : pos - dt; ray.d = v3(1.f, 1.f, 1.f); ray.tmin = 0.f; ray.tmax = 2e-10f;
      float t0, t1, v = 0.f;
      const int leaf = closestRegion(s, active, ray, t0, t1);
      sv[j][i] = (leaf >= 0 && samplePoint(s, v, leaf, ray.o, channel, nullptr)) ? v : 0.f;
    }
  return v3(sv[1][0] - sv[0][0], sv[1][1] - sv[0][1], sv[1][2] - sv[0][2]);
}
// samplePointWithDerivative<ISO> in either build: analytic (default) or central differences
static inline bool sampleWithGradient(const OrcScene &s, const OrcFrame &f, const std::vector<uint8_t> &active, float &value, V3 &g,
                                      int leafID, V3 pos, int channel, OrcStats *st)
{
  if (!f.centralDifferences) return samplePointD(s, value, g, leafID, pos, channel, st);
  if (!samplePoint(s, value, leafID, pos, channel, st)) return false;
  g = gradientCD(s, active, pos, (s.regions[leafID].finestLevelCellWidth + 1.f) * .5f, channel);
  return true;
}

struct IntResult { V4 *pc; float t_hit = -1.f; V3 grad = {0, 0, 0}; };

// integrateVolume, exabrick.cu:988-1016
static inline void integrateVolume(const OrcFrame &f, const Ray &ray, IntResult &res, float actual_dt,
                                   float cellValue, V3 g, int fcw, int ch)
{
  if (actual_dt == 0.f) return;
  V4 &pc = *res.pc;
  V4 smp = lookupXF(f, cellValue, ch);
  if (length(g) > fcw * 1e-6f) {
    const V3 l = -ray.d;
    const float scale = fabsf(dot(l, g)) / sqrtf(dot(g, g) * dot(l, l));
    smp.x *= scale; smp.y *= scale; smp.z *= scale;
  }
  smp.w = 1.f - powf(1.f - smp.w, actual_dt);
  const float k = (1.f - pc.w) * smp.w;
  pc.x += k * smp.x; pc.y += k * smp.y; pc.z += k * smp.z; pc.w += k * 1.f;
}

// IsoSurfaceIntegrationFunction, exabrick.cu:1018-1114
struct IsoFunc {
  float last_t = 0.f, lastCellValue = -1e36f;
  void operator()(const OrcScene &s, const OrcFrame &f, const std::vector<uint8_t> &isoActive, const Ray &ray, IntResult &res, float t_sample,
                  float cellValue, int leafID, int channel, OrcStats *st)
  {
    if (lastCellValue >= -1e35f) {
      for (int i = 0; i < 2; i++) {
        const float iso = f.isoValue[i];
        if (f.isoEnabled[i] && f.isoChannel[i] == channel &&
            ((lastCellValue <= iso && cellValue >= iso) || (lastCellValue >= iso && cellValue <= iso))) {
          float d1 = fabsf(lastCellValue - iso), d2 = fabsf(cellValue - iso);
          float w1 = 1.f - d1 / (d1 + d2), w2 = 1.f - d2 / (d1 + d2);
          float tavg = last_t * w1 + t_sample * w2;
          float cv = 0.f; V3 grad = v3(0, 0, 0); V4 smp = {1.f, 0.f, 0.f, 1.f};
          const V3 isopt = ray.o + tavg * ray.d;
          if (f.gradientShadingISO) {
            if (sampleWithGradient(s, f, isoActive, cv, grad, leafID, isopt, f.isoChannel[i], st)) {     // <ISO = true>
              smp = lookupXF(f, cv, f.isoChannel[i]);
              grad = normalize(grad);
              if (dot(grad, ray.d) > 0.f) grad = -grad;
            }
          } else {
            if (samplePoint(s, cv, leafID, isopt, f.isoChannel[i], st)) smp = lookupXF(f, cv, f.isoChannel[i]);
          }
          if (f.colormapChannel != 0) {
            cv = 0.f;
            if (samplePoint(s, cv, leafID, isopt, f.colormapChannel, st)) smp = lookupXF(f, cv, 0);
          }
          smp.w = 1.f;
          if (!std::isfinite(grad.x) || !std::isfinite(grad.y) || !std::isfinite(grad.z)) grad = v3(0, 0, 0);
          if (length(grad) > 0.f) {
            const V3 l = -ray.d;
            const float scale = .3f + .7f * fabsf(dot(l, grad)) / sqrtf(dot(grad, grad));
            smp.x *= scale; smp.y *= scale; smp.z *= scale;
          }
          V4 &pc = *res.pc; const float k = (1.f - pc.w) * smp.w;
          pc.x += k * smp.x; pc.y += k * smp.y; pc.z += k * smp.z; pc.w += k;
          res.t_hit = tavg; res.grad = grad;
        }
      }
    }
    last_t = t_sample; lastCellValue = cellValue;
  }
};

// common sample-lattice set-up of integrateBrick / isoIntegrateBrick, exabrick.cu:1141-1144
static inline float firstLatticeT(float t0, float dt, float offs)
{
  int i0 = (int)ceilf((t0 - dt * offs) / dt);
  float t_i = (offs + i0) * dt;
  while ((t_i - dt) >= t0) t_i = t_i - dt;
  while (t_i < t0) t_i += dt;
  return t_i;
}

// integrateBrick<GRADIENT_SHADING>, exabrick.cu:1116-1185
static thread_local bool g_dbg = false;
static void integrateBrick(const OrcScene &s, const OrcFrame &f, const std::vector<uint8_t> &volActive, IntResult &res, float offs, const Ray &ray,
                           int leafID, float t0, float t1, int numChannels, bool gradShading, OrcStats *st)
{
  const OrcRegion &r = s.regions[leafID];
  const float dt = f.dt * r.finestLevelCellWidth; const int fcw = (int)r.finestLevelCellWidth;
  float t_i = firstLatticeT(t0, dt, offs); float t_last = t0;
  if (st) st->segments++;
  if (g_dbg) printf("O seg rid=%d K=%d t0=%.9g t1=%.9g dt=%.9g tFirst=%.9g\n", leafID, r.leafListSize, t0, t1, dt, t_i);
  for (;; t_i += dt) {
    const float t_next = fminf(t_i, t1);
    const float t_sample = 0.5f * (fminf(t1, t_next) + t_last);
    const float actual_dt = t_next - t_last; t_last = t_next;
    const V3 pos = ray.o + t_sample * ray.d;
    float cv = 0.f; V3 g = v3(0, 0, 0);
    for (int c = 0; c < numChannels; c++) {
      if (g_dbg) { float v2; V3 g2; if (samplePointD(s, v2, g2, leafID, pos, c, nullptr)) printf("O   smp t=%.9g adt=%.9g v=%.9g g=(%.6g %.6g %.6g)\n", t_sample, actual_dt, v2, g2.x, g2.y, g2.z); }
      if (gradShading) { if (sampleWithGradient(s, f, volActive, cv, g, leafID, pos, c, st)) integrateVolume(f, ray, res, actual_dt, cv, g, fcw, c); }
      else             { if (samplePoint(s, cv, leafID, pos, c, st))     integrateVolume(f, ray, res, actual_dt, cv, g, fcw, c); }
    }
    if (res.pc->w >= f.ertThreshold) break;
    if (t_next >= t1) break;
  }
}

// isoIntegrateBrick, exabrick.cu:1187-1256
static void isoIntegrateBrick(const OrcScene &s, const OrcFrame &f, const std::vector<uint8_t> &isoActive, IsoFunc *funcs, IntResult &res, float offs,
                              const Ray &ray, int leafID, float t0, float t1, int numChannels, OrcStats *st)
{
  const OrcRegion &r = s.regions[leafID];
  const float dt = f.dt * r.finestLevelCellWidth;
  float t_i = firstLatticeT(t0, dt, offs); float t_last = t0;
  if (st) st->segments++;
  for (;; t_i += dt) {
    const float t_next = fminf(t_i, t1);
    const float t_sample = 0.5f * (fminf(t1, t_next) + t_last);
    t_last = t_next;
    const V3 pos = ray.o + t_sample * ray.d;
    for (int c = 0; c < numChannels; c++) {
      float cv = 0.f; V3 g = v3(0, 0, 0); bool ok;
      // the gradient of this sample is never used by the iso functor: the central-difference build's six extra samples
      // (samplePointWithDerivative<false>, :1237) have no effect and are not taken
      if (f.gradientShadingISO && !f.centralDifferences) ok = samplePointD(s, cv, g, leafID, pos, c, st);
      else ok = samplePoint(s, cv, leafID, pos, c, st);
      if (ok) {
        funcs[c](s, f, isoActive, ray, res, t_sample, cv, leafID, c, st);
        if (res.pc->w >= f.ertThreshold) break;
      }
    }
    if (t_next >= t1) break;
  }
}

// xfmPoint / xfmVector (owl/common/math/AffineSpace.h)
static inline V3 xfmVector(const OrcFrame &f, V3 v)
{ return v3(v.x * f.xf_vx[0] + v.y * f.xf_vy[0] + v.z * f.xf_vz[0], v.x * f.xf_vx[1] + v.y * f.xf_vy[1] + v.z * f.xf_vz[1], v.x * f.xf_vx[2] + v.y * f.xf_vy[2] + v.z * f.xf_vz[2]); }
static inline V3 xfmPoint(const OrcFrame &f, V3 v)
{ V3 r = xfmVector(f, v); return v3(r.x + f.xf_p[0], r.y + f.xf_p[1], r.z + f.xf_p[2]); }

struct Surface { int primID = -1; float t_hit = 0.f; V3 Ng = {0, 0, 0}; float ambient = 0.f; V3 base = {0, 0, 0}; };
enum { PRIMID_STREAMLINE = -25, PRIMID_PLANE = -24, PRIMID_ISOSURFACE = -23 };

// traceIsoRay, exabrick.cu:1408-1460
static Surface traceIsoRay(const OrcScene &s, const OrcFrame &f, const std::vector<uint8_t> &isoActive, Ray ray, OrcStats *st)
{
  ray.o = xfmPoint(f, ray.o); ray.d = xfmVector(f, ray.d);
  const float dt_scale = length(ray.d); ray.d = normalize(ray.d);
  float already = dt_scale * ray.tmin;
  IsoFunc funcs[10];
  Surface result; result.t_hit = ray.tmax;
  while (1) {
    ray.tmin = already; ray.tmax = ray.tmax * dt_scale;    // (sic) :1433-1434
    float t0, t1; int leaf = closestRegion(s, isoActive, ray, t0, t1);
    if (leaf < 0) break;
    V4 pc = {0, 0, 0, 0}; IntResult ir; ir.pc = &pc;
    isoIntegrateBrick(s, f, isoActive, funcs, ir, 0.f, ray, leaf, fmaxf(ray.tmin, t0), fminf(ray.tmax, t1), f.numPrimaryChannels, st);
    if (ir.t_hit >= 0.f) {
      result.primID = PRIMID_ISOSURFACE; result.t_hit = ir.t_hit / dt_scale;
      result.Ng = normalize(ir.grad); result.ambient = 0.f; result.base = v3(pc.x, pc.y, pc.z);
      return result;
    }
    already = t1 * 1.0000001f;
  }
  return result;
}

// intersectRoundedCone, exabrick.cu:441-501
static bool roundedCone(V3 pa, V3 pb, float ra, float rb, const Ray &ray, float &hit_t, V3 &n)
{
  V3 ro = ray.o; const V3 rd = ray.d;
  float minDist = fmaxf(0.f, fminf(length(pa - ro) - ra, length(pb - ro) - rb));
  ro = ro + minDist * rd;
  V3 ba = pb - pa, oa = ro - pa, ob = ro - pb; float rr = ra - rb;
  float m0 = dot(ba, ba), m1 = dot(ba, oa), m2 = dot(ba, rd), m3 = dot(rd, oa), m5 = dot(oa, oa);
  float d2 = m0 - rr * rr;
  float k2 = d2 - m2 * m2, k1 = d2 * m3 - m1 * m2 + m2 * rr * ra, k0 = d2 * m5 - m1 * m1 + m1 * rr * ra * 2.0f - m0 * ra * ra;
  float h = k1 * k1 - k0 * k2;
  if (h < 0.0f) return false;
  float t = (-sqrtf(h) - k1) / k2;
  float y = m1 - ra * rr + t * m2;
  if (y > 0.0f && y < d2) { hit_t = minDist + t; n = d2 * (oa + t * rd) - ba * y; return true; }
  float h1 = m3 * m3 - m5 + ra * ra;
  if (h1 > 0.0f) { t = -m3 - sqrtf(h1); hit_t = minDist + t; n = (oa + t * rd) * (1.f / ra); return true; }
  return false;
}

// Streamline closest hit (exabrick.cu:503-530 intersect, 532-570 bounds/visibility), brute force.
// prd comes in with primID -1 / t_hit 2e10 (:1505-1507); hits are accepted inside (ray.tmin, ray.tmax).
static void traceStreamlines(const OrcFrame &f, const Ray &ray, Surface &prd)
{
  if (!f.traces) return;
  float tmax = ray.tmax;
  for (int i = 0; i < f.numTraces; i++)
    for (int k = 0; k + 1 < f.numTimesteps; k++) {
      if (k >= f.currentTimestep - 1) break;
      const float *a = f.traces + 3 * ((size_t)i * f.numTimesteps + k), *b = a + 3;
      if (!(a[0] < 2e10f && b[0] < 2e10f)) continue;
      float t; V3 n;
      if (roundedCone(v3(a[0], a[1], a[2]), v3(b[0], b[1], b[2]), 2.f, 2.f, ray, t, n) && t > ray.tmin && t < tmax) {
        tmax = t; prd.primID = PRIMID_STREAMLINE; prd.t_hit = t; prd.Ng = normalize(n); prd.base = v3(.8f, .8f, .8f); prd.ambient = 0.f;
      }
    }
}

// intersectLinePlane, exabrick.cu:1267-1284
static float intersectLinePlane(V3 p1, V3 p2, V3 normal, float offset)
{
  const float sdn = dot(normal, normalize(p2 - p1));
  if (sdn == 0.f) return -1.f;
  const float t = (offset - dot(normal, p1)) / sdn;
  if (t < 0.f || t > length(p2 - p1)) return -1.f;
  return t;
}

// intersectBoxPlane, exabrick.cu:1287-1314: the 12 edges of the box as 4 key corners x 3 axes
static void intersectBoxPlane(const V3 corners[2], V3 normal, float offset, V3 *pts, int &cnt)
{
  static const int key[4][3] = {{0, 0, 0}, {1, 0, 1}, {1, 1, 0}, {0, 1, 1}};
  cnt = 0;
  for (int i = 0; i < 4 && cnt < 6; ++i)
    for (int j = 0; j < 3 && cnt < 6; ++j) {
      const V3 p1 = v3(j == 0 ? corners[1 - key[i][0]].x : corners[key[i][0]].x,
                       j == 1 ? corners[1 - key[i][1]].y : corners[key[i][1]].y,
                       j == 2 ? corners[1 - key[i][2]].z : corners[key[i][2]].z);
      const V3 p2 = v3(corners[key[i][0]].x, corners[key[i][1]].y, corners[key[i][2]].z);
      const float t = intersectLinePlane(p1, p2, normal, offset);
      if (t >= 0.f) pts[cnt++] = p1 + normalize(p2 - p1) * t;
    }
}

// intersectRayTriangle, exabrick.cu:1316-1344
static float intersectRayTriangle(const Ray &ray, V3 v1, V3 e1, V3 e2)
{
  const V3 s1 = cross(ray.d, e2);
  const float div = dot(s1, e1);
  if (div == 0.f) return -1.f;
  const float invDiv = 1.f / div;
  const V3 d = ray.o - v1;
  const float b1 = dot(d, s1) * invDiv;
  if (b1 < 0.f || b1 > 1.f) return -1.f;
  const V3 s2 = cross(d, e1);
  const float b2 = dot(ray.d, s2) * invDiv;
  if (b2 < 0.f || b1 + b2 > 1.f) return -1.f;
  return dot(e2, s2) * invDiv;
}

// traceContourRay, exabrick.cu:1346-1406.  The plane is given in the unit cube and scaled to the
// world-space bounds; the polygon is fanned from pts[0].  Quirks kept: the sample position is NOT
// taken through voxelSpaceTransform and always reads channel 0 (samplePointWithInfRay(pos, 0),
// :818-830, :1396), the XF is the plane's channel.  The reference leaves `value` uninitialised
// when the sample is invalid; here it is 0 (decision recorded in DESIGN.md).
static Surface traceContourRay(const OrcScene &s, const OrcFrame &f, const Ray &ray, V3 normal, float offset, int channel, OrcStats *st)
{
  Surface prd; prd.primID = -1; prd.t_hit = ray.tmax;
  V3 pts[6]; int cnt;
  const V3 unit[2] = {v3(0, 0, 0), v3(1, 1, 1)};
  intersectBoxPlane(unit, normal, offset, pts, cnt);
  for (int i = 0; i < cnt; ++i) {
    pts[i] = v3(pts[i].x * (f.worldHi[0] - f.worldLo[0]), pts[i].y * (f.worldHi[1] - f.worldLo[1]), pts[i].z * (f.worldHi[2] - f.worldLo[2]));
    pts[i] = pts[i] + v3(f.worldLo[0], f.worldLo[1], f.worldLo[2]);
  }
  float t = -1.f;
  for (int i = 0; i < cnt - 1; ++i) {                      // "cyclical selection-sort" :1367-1383
    int minIdx = i;
    for (int j = i + 1; j < cnt; ++j) {
      const V3 v = cross(pts[j] - pts[0], pts[minIdx] - pts[0]);
      if (dot(v, normal) < 0.f) minIdx = j;
    }
    std::swap(pts[i], pts[minIdx]);
  }
  for (int i = 2; i < cnt; ++i) {
    const V3 v1 = pts[0], e1 = pts[i - 1] - v1, e2 = pts[i] - v1;
    const float tt = intersectRayTriangle(ray, v1, e1, e2);
    if (tt >= 0.f && (tt < t || t < 0.f)) t = tt;
  }
  if (t < 0.f) return prd;
  const V3 pos = ray.o + ray.d * t;
  float value = 0.f;
  const int leaf = locateRegion(s, pos);                   // closest region of a ray of length 2e-10 from pos
  if (leaf >= 0) { float v2; if (samplePoint(s, v2, leaf, pos, 0, st)) value = v2; }
  const V4 smp = lookupXF(f, value, channel);
  prd.primID = PRIMID_PLANE; prd.t_hit = t; prd.Ng = normal; prd.ambient = 0.f; prd.base = v3(smp.x, smp.y, smp.z);
  return prd;
}

// make_orthonormal_basis / cosine_sample_hemisphere, exabrick.cu:78-94
static void makeOrthonormalBasis(V3 &u, V3 &v, V3 w)
{
  v = fabsf(w.x) > fabsf(w.y) ? normalize(v3(-w.z, 0, w.x)) : normalize(v3(0, w.z, -w.y));
  u = cross(v, w);
}
static V3 cosineSampleHemisphere(float u1, float u2)
{
  const float r = sqrtf(u1);
  const float theta = (float)(2.f * M_PI * u2);
  return v3(r * cosf(theta), r * sinf(theta), sqrtf(1.f - u1));
}

// traceSurfaces, exabrick.cu:1476-1529
static void traceSurfaces(const OrcScene &s, const OrcFrame &f, const std::vector<uint8_t> &isoActive, bool anyIso,
                          const Ray &ray, Surface &prd, bool withContourPlanes, OrcStats *st)
{
  prd.primID = -1; prd.t_hit = ray.tmax;
  // triangle geometry first (:1488-1490); closest hit program SurfaceBVH (:420-433).  OptiX's
  // built-in triangle test is replaced by a brute-force closest hit over all triangles (no
  // culling: the ray flags only disable any-hit programs)
  if (f.meshVerts && f.meshIndex) {
    float best = ray.tmax; int bestTri = -1;
    for (int t = 0; t < f.numMeshTris; t++) {
      const int32_t *ix = f.meshIndex + 3 * (size_t)t;
      const V3 A = v3(f.meshVerts[3 * ix[0]], f.meshVerts[3 * ix[0] + 1], f.meshVerts[3 * ix[0] + 2]);
      const V3 B = v3(f.meshVerts[3 * ix[1]], f.meshVerts[3 * ix[1] + 1], f.meshVerts[3 * ix[1] + 2]);
      const V3 Cc = v3(f.meshVerts[3 * ix[2]], f.meshVerts[3 * ix[2] + 1], f.meshVerts[3 * ix[2] + 2]);
      const float tt = intersectRayTriangle(ray, A, B - A, Cc - A);
      if (tt > ray.tmin && tt < best) { best = tt; bestTri = t; }
    }
    if (bestTri >= 0) {
      const int32_t *ix = f.meshIndex + 3 * (size_t)bestTri;
      const V3 A = v3(f.meshVerts[3 * ix[0]], f.meshVerts[3 * ix[0] + 1], f.meshVerts[3 * ix[0] + 2]);
      const V3 B = v3(f.meshVerts[3 * ix[1]], f.meshVerts[3 * ix[1] + 1], f.meshVerts[3 * ix[1] + 2]);
      const V3 Cc = v3(f.meshVerts[3 * ix[2]], f.meshVerts[3 * ix[2] + 1], f.meshVerts[3 * ix[2] + 2]);
      prd.primID = bestTri; prd.t_hit = best; prd.Ng = normalize(cross(B - A, Cc - A)); prd.ambient = .2f; prd.base = v3(.8f, .8f, .8f);
    }
  }
  if (withContourPlanes)
    for (int i = 0; i < 3; i++)
      if (f.cpEnabled[i]) {
        const Surface c = traceContourRay(s, f, ray, v3(f.cpNormal[i][0], f.cpNormal[i][1], f.cpNormal[i][2]), f.cpOffset[i], f.cpChannel[i], st);
        if (c.primID == PRIMID_PLANE && c.t_hit < prd.t_hit) prd = c;
      }
  {
    Surface sl; sl.primID = -1; sl.t_hit = 2e10f;
    traceStreamlines(f, ray, sl);
    if (sl.primID == PRIMID_STREAMLINE && sl.t_hit < prd.t_hit) prd = sl;
  }
  if (anyIso) {
    const Surface iso = traceIsoRay(s, f, isoActive, ray, st);
    if (iso.primID == PRIMID_ISOSURFACE && iso.t_hit < prd.t_hit) prd = iso;
  }
}

// test entry points for the surface geometry (tests/test_surfaces_host.py compares the product's
// csrc/surfaces.cuh, compiled for the host, with these)
extern "C" int orc_rounded_cone(const float pa[3], const float pb[3], float ra, float rb, const float ro[3], const float rd[3],
                                float *t, float n[3])
{
  Ray r; r.o = v3(ro[0], ro[1], ro[2]); r.d = v3(rd[0], rd[1], rd[2]); r.tmin = 0.f; r.tmax = 1e30f;
  V3 nn = v3(0, 0, 0); float tt = 0.f;
  const bool hit = roundedCone(v3(pa[0], pa[1], pa[2]), v3(pb[0], pb[1], pb[2]), ra, rb, r, tt, nn);
  *t = tt; n[0] = nn.x; n[1] = nn.y; n[2] = nn.z;
  return hit ? 1 : 0;
}
extern "C" float orc_contour_plane_t(const float ro[3], const float rd[3], const float normal[3], float offset,
                                     const float wlo[3], const float whi[3], float *pts18, int *cnt)
{
  // polygon part of traceContourRay (exabrick.cu:1355-1394), without the sample
  const V3 nrm = v3(normal[0], normal[1], normal[2]);
  V3 pts[6]; int c;
  const V3 unit[2] = {v3(0, 0, 0), v3(1, 1, 1)};
  intersectBoxPlane(unit, nrm, offset, pts, c);
  for (int i = 0; i < c; ++i) {
    pts[i] = v3(pts[i].x * (whi[0] - wlo[0]), pts[i].y * (whi[1] - wlo[1]), pts[i].z * (whi[2] - wlo[2]));
    pts[i] = pts[i] + v3(wlo[0], wlo[1], wlo[2]);
  }
  for (int i = 0; i < c - 1; ++i) {
    int minIdx = i;
    for (int j = i + 1; j < c; ++j) if (dot(cross(pts[j] - pts[0], pts[minIdx] - pts[0]), nrm) < 0.f) minIdx = j;
    std::swap(pts[i], pts[minIdx]);
  }
  Ray r; r.o = v3(ro[0], ro[1], ro[2]); r.d = v3(rd[0], rd[1], rd[2]); r.tmin = 0.f; r.tmax = 1e30f;
  float t = -1.f;
  for (int i = 2; i < c; ++i) {
    const float tt = intersectRayTriangle(r, pts[0], pts[i - 1] - pts[0], pts[i] - pts[0]);
    if (tt >= 0.f && (tt < t || t < 0.f)) t = tt;
  }
  if (cnt) *cnt = c;
  if (pts18) for (int i = 0; i < c; i++) { pts18[3 * i] = pts[i].x; pts18[3 * i + 1] = pts[i].y; pts18[3 * i + 2] = pts[i].z; }
  return t;
}
extern "C" void orc_ao_direction(const float normal[3], float u1, float u2, float out[3])
{
  V3 u, v; const V3 w = v3(normal[0], normal[1], normal[2]);
  makeOrthonormalBasis(u, v, w);
  const V3 sp = cosineSampleHemisphere(u1, u2);
  const V3 d = normalize(sp.x * u + sp.y * v + sp.z * w);
  out[0] = d.x; out[1] = d.y; out[2] = d.z;
}

// owl::LCG<16>, owl/common/math/random.h:56-96
struct LCG16 {
  uint32_t state;
  LCG16(uint32_t v0, uint32_t v1)
  {
    uint32_t s0 = 0;
    for (int n = 0; n < 16; n++) {
      s0 += 0x9e3779b9;
      v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4);
      v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761e);
    }
    state = v0;
  }
  float operator()() { state = 1664525u * state + 1013904223u; return ldexpf((float)state, -32); }
};

static inline float linear_to_srgb(float x) { return x <= 0.0031308f ? 12.92f * x : 1.055f * powf(x, 1.f / 2.4f) - 0.055f; }
static inline int make_8bit(float f) { return std::min(255, std::max(0, (int)(f * 256.f))); }

// __raygen__renderFrame, exabrick.cu:1576-1720
extern "C" int orc_render(const OrcScene *sp, const OrcFrame *fp, int x0, int y0, int x1, int y1,
                          float *accum, uint32_t *rgba8, OrcStats *stats)
{
  const OrcScene &s = *sp; const OrcFrame &f = *fp;
  const int W = f.width, H = f.height;
  if (stats) memset(stats, 0, sizeof(*stats));
  // bounds programs: exabrick.cu:285-312 (volume) and :373-402 (iso)
  std::vector<uint8_t> volActive(s.regions.size()), isoActive(s.regions.size());
  bool anyIso = false;
  for (int i = 0; i < 2; i++) anyIso |= f.isoEnabled[i] != 0;
  bool anyContour = false;
  for (int i = 0; i < 3; i++) anyContour |= f.cpEnabled[i] != 0;
  const bool rtl = getenv("ORC_RND_ARGS_RTL") && atoi(getenv("ORC_RND_ARGS_RTL"));
  for (size_t i = 0; i < s.regions.size(); i++) {
    const OrcRegion &r = s.regions[i];
    bool act = false;
    for (int c = 0; c < f.numPrimaryChannels && !act; c++) act |= activeForVolume(f, r.vlo, r.vhi, c);
    volActive[i] = (f.spaceSkipping && !anyContour) ? act : 1;   // no space skipping while a contour plane is on (OptixRenderer.cpp:518-530)
    bool ia = false;
    for (int k = 0; k < 2; k++) if (f.isoEnabled[k] && f.isoValue[k] >= r.vlo && f.isoValue[k] <= r.vhi) ia = true;
    isoActive[i] = ia;
  }
#pragma omp parallel
  {
    OrcStats st; memset(&st, 0, sizeof(st));
#pragma omp for schedule(dynamic, 8) collapse(2)
    for (int py = y0; py < y1; py++)
      for (int px = x0; px < x1; px++) {
        const int pixelIdx = px + W * py;
        { const char *e = getenv("ORC_DEBUG_PIXEL"); int dx = -1, dy = -1; g_dbg = e && sscanf(e, "%d,%d", &dx, &dy) == 2 && dx == px && dy == py; }
        LCG16 rnd((uint32_t)(f.frameID * W * H) + (uint32_t)px, (uint32_t)py);
        // vec2f(rnd(), rnd()) (:1594): the order of the two calls is unspecified in C++.  nvcc evaluates
        // left to right (x first), which is what the product and this oracle do; g++ evaluates right to
        // left, so the reference shader compiled with g++ (oracle/_ref/ref_shader) needs rtl = true
        float jx, jy;
        if (rtl) { jy = rnd(); jx = rnd(); } else { jx = rnd(); jy = rnd(); }
        const float sx = (float)px + jx, sy = (float)py + jy;
        Ray ray;                                           // Camera::generateRay, Camera.h:27-44
        ray.o = v3(f.cam_pos[0], f.cam_pos[1], f.cam_pos[2]);
        ray.d = normalize(v3(f.cam_dir00[0] + sx * f.cam_du[0] + sy * f.cam_dv[0],
                             f.cam_dir00[1] + sx * f.cam_du[1] + sy * f.cam_dv[1],
                             f.cam_dir00[2] + sx * f.cam_du[2] + sy * f.cam_dv[2]));
        ray.tmin = 1e-6f; ray.tmax = 1e8f;
        Surface surf;
        traceSurfaces(s, f, isoActive, anyIso, ray, surf, true, &st);        // :1598 ST_ALL_SURFACES
        V3 bg = v3(0, 0, 0);
        if (surf.primID != -1) {                           // :1604-1652
          const bool shade = surf.primID >= 0 || surf.primID == PRIMID_STREAMLINE || surf.primID == PRIMID_PLANE ||
                             (surf.primID == PRIMID_ISOSURFACE && f.gradientShadingISO);
          if (shade && length(surf.Ng) > 0.f) {
            const int aoSamples = f.aoEnabled ? 2 : 0;
            const V3 isect = ray.o + ray.d * surf.t_hit;
            V3 u, v; const V3 w = surf.Ng;
            makeOrthonormalBasis(u, v, w);
            int hitCnt = 0;
            for (int i = 0; i < aoSamples; ++i) {
              float u1, u2;                                  // cosine_sample_hemisphere(rnd(), rnd()) (:1624): same remark
              if (rtl) { u2 = rnd(); u1 = rnd(); } else { u1 = rnd(); u2 = rnd(); }
              const V3 sp = cosineSampleHemisphere(u1, u2);
              Ray ao; ao.o = isect; ao.d = normalize(sp.x * u + sp.y * v + sp.z * w); ao.tmin = 1e-4f; ao.tmax = f.aoLength;
              Surface hit;
              traceSurfaces(s, f, isoActive, anyIso, ao, hit, false, &st);   // ST_ALL & ~ST_CONTOUR_PLANES
              if (hit.primID != -1) hitCnt++;
            }
            const float shadow = f.aoEnabled ? (float)hitCnt / aoSamples : 0.f;
            const float k = fabsf(dot(ray.d, surf.Ng)) * (1.f - shadow);
            bg = v3(surf.ambient + surf.base.x * k, surf.ambient + surf.base.y * k, surf.ambient + surf.base.z * k);
          } else bg = surf.base;
        }
        V4 pc = {0, 0, 0, 0};
        const float offs = rnd();                          // :1655
        ray.tmax = surf.t_hit;
        if (f.clipEnabled) { float a, b; Ray r2 = ray; boxTest(r2, f.clipLo, f.clipHi, a, b); ray.tmin = a; ray.tmax = b; }   // clipRay :1258-1265
        surf.t_hit = ray.tmax;
        ray.o = xfmPoint(f, ray.o); ray.d = xfmVector(f, ray.d);
        const float dt_scale = length(ray.d); ray.d = normalize(ray.d);
        float already = dt_scale * ray.tmin;
        while (1) {                                        // :1675-1699
          ray.tmin = already; ray.tmax = surf.t_hit * dt_scale;
          float t0, t1; int leaf = closestRegion(s, volActive, ray, t0, t1);
          if (leaf < 0) break;
          IntResult ir; ir.pc = &pc;
          integrateBrick(s, f, volActive, ir, offs, ray, leaf, t0, t1, f.numPrimaryChannels, f.gradientShadingDVR != 0, &st);
          if (pc.w >= f.ertThreshold) { pc = V4{pc.x * pc.w, pc.y * pc.w, pc.z * pc.w, 1.f}; break; }
          already = t1 * 1.0000001f;
        }
        if (f.partial) {                                   // sort-last partial (not in reference)
          accum[4 * pixelIdx + 0] = pc.x; accum[4 * pixelIdx + 1] = pc.y; accum[4 * pixelIdx + 2] = pc.z; accum[4 * pixelIdx + 3] = pc.w;
          continue;
        }
        V3 color = v3(pc.w * pc.x + (1.f - pc.w) * bg.x, pc.w * pc.y + (1.f - pc.w) * bg.y, pc.w * pc.z + (1.f - pc.w) * bg.z);   // :1701 (sic)
        if (f.frameID > 0) color = color + v3(accum[4 * pixelIdx], accum[4 * pixelIdx + 1], accum[4 * pixelIdx + 2]);
        accum[4 * pixelIdx + 0] = color.x; accum[4 * pixelIdx + 1] = color.y; accum[4 * pixelIdx + 2] = color.z; accum[4 * pixelIdx + 3] = 1.f;
        const float inv = (float)f.frameID + 1.f;
        color = v3(linear_to_srgb(color.x / inv), linear_to_srgb(color.y / inv), linear_to_srgb(color.z / inv));
        rgba8[pixelIdx] = (uint32_t)(make_8bit(color.x) << 0) + (uint32_t)(make_8bit(color.y) << 8) + (uint32_t)(make_8bit(color.z) << 16) + (0xffu << 24);
      }
    addStats(stats, st);
  }
  return 0;
}

// =================================================================================
// 5. Streamlines: computeTraces (RK4), exabrick.cu:1531-1574; sampleDirection :945-963
// =================================================================================
static bool sampleDirection(const OrcScene &s, V3 pos, const int32_t ch[3], V3 &out)
{
  int leaf = locateRegion(s, pos);
  float v[3];
  for (int i = 0; i < 3; i++) if (!samplePoint(s, v[i], leaf, pos, ch[i], nullptr)) return false;
  out = v3(v[0], v[1], v[2]); return true;
}

extern "C" int orc_trace_streamlines(const OrcScene *sp, float *traces, int numTraces, int numTimesteps,
                                     const int32_t channels[3], float steplen,
                                     const float wlo[3], const float whi[3])
{
  const OrcScene &s = *sp;
#pragma omp parallel for schedule(dynamic, 4)
  for (int i = 0; i < numTraces; i++)
    for (int t = 1; t < numTimesteps; t++) {
      float *src = traces + 3 * ((size_t)i * numTimesteps + (t - 1)), *dst = src + 3;
      V3 p = v3(src[0], src[1], src[2]); const V3 pp = p;
      if (p.x < 2e10f) {
        bool valid = true; V3 k1 = v3(0, 0, 0), k2 = k1, k3 = k1, k4 = k1;
        valid &= sampleDirection(s, p, channels, k1); k1 = k1 * steplen; V3 p1 = p + k1 * .5f;
        valid &= sampleDirection(s, p1, channels, k2); k2 = k2 * steplen; V3 p2 = p + k2 * .5f;
        valid &= sampleDirection(s, p2, channels, k3); k3 = k3 * steplen; V3 p3 = p + k3;
        valid &= sampleDirection(s, p3, channels, k4); k4 = k4 * steplen;
        p = p + (1 / 6.f) * (k1 + 2.f * k2 + 2.f * k3 + k4);
        // box3f::contains: lower <= p <= upper (owl/common/math/box.h)
        bool inside = p.x >= wlo[0] && p.x <= whi[0] && p.y >= wlo[1] && p.y <= whi[1] && p.z >= wlo[2] && p.z <= whi[2];
        if (!valid || !inside || length(p - pp) < 1e-10f) p = v3(2e10f, 2e10f, 2e10f);
      }
      dst[0] = p.x; dst[1] = p.y; dst[2] = p.z;
    }
  return 0;
}
