// builder_common.h — the per-node DECISIONS of exaBuilder (builder/builder.cpp:447-530 tryMakeLeaf, 538-811 buildRec),
// shared by the host builder (builder.cpp) and the GPU builder (gpu_builder.cu).  Both reduce a node's cells to the same
// small statistics (4-D bounds; per axis and per coarse slice: used volume, cell count, 4-D bounds) - the host with OpenMP
// passes, the GPU with kernels - and both hand them to these functions, so the tree, and with it the brick order, is
// decided by one piece of code that the CPU tests pin against the reference's own exaBuilder bit for bit.
#pragma once
#include <climits>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <vector>

namespace exa {
namespace bld {

struct Box4 {
  int lo[4] = {INT_MAX, INT_MAX, INT_MAX, INT_MAX};
  int hi[4] = {INT_MIN, INT_MIN, INT_MIN, INT_MIN};
  inline void extendCell(int x, int y, int z, int level)
  {
    const int cw = 1 << level;                        // SingleCell::getBounds, builder.cpp:121-130
    lo[0] = lo[0] < x ? lo[0] : x; lo[1] = lo[1] < y ? lo[1] : y; lo[2] = lo[2] < z ? lo[2] : z; lo[3] = lo[3] < level ? lo[3] : level;
    hi[0] = hi[0] > x + cw ? hi[0] : x + cw; hi[1] = hi[1] > y + cw ? hi[1] : y + cw; hi[2] = hi[2] > z + cw ? hi[2] : z + cw;
    hi[3] = hi[3] > level + 1 ? hi[3] : level + 1;
  }
  inline void extend(const Box4 &o)
  { for (int i = 0; i < 4; i++) { lo[i] = lo[i] < o.lo[i] ? lo[i] : o.lo[i]; hi[i] = hi[i] > o.hi[i] ? hi[i] : o.hi[i]; } }
  inline int size(int i) const { return hi[i] - lo[i]; }
};
static_assert(sizeof(Box4) == 32, "Box4 is 8 ints (the GPU builder copies it as such)");

inline size_t unitCellVolume(const Box4 &b) { return (size_t)b.size(0) * (size_t)b.size(1) * (size_t)b.size(2); }   // builder.cpp:164-168
inline size_t areaOf(const Box4 &b)                                                                                    // :170-177 (no factor 2)
{ return (size_t)b.size(0) * (size_t)b.size(1) + (size_t)b.size(1) * (size_t)b.size(2) + (size_t)b.size(2) * (size_t)b.size(0); }
inline int divDown(int a, int b) { return a >= 0 ? a / b : (a - (b - 1)) / b; }   // :53-59
inline int divUp(int a, int b) { return a >= 0 ? (a + b - 1) / b : a / b; }       // :61-67

// computeCoarsestLevelBounds (:201-225): xyz snapped outward to multiples of the coarsest cell width; returns that width
inline int snapToCoarsest(Box4 &cb)
{
  const int ccw = 1 << (cb.hi[3] - 1);
  for (int d = 0; d < 3; d++) { cb.lo[d] = ccw * divDown(cb.lo[d], ccw); cb.hi[d] = ccw * divUp(cb.hi[d], ccw); }
  return ccw;
}

// the leaf test of tryMakeLeaf (:447-481): one level, at most maxLeafWidth cells per axis, completely full
inline bool isLeaf(const Box4 &cb, size_t n, int maxLeafWidth)
{
  if (cb.size(3) > 1) return false;
  const size_t cw = 1ULL << (cb.hi[3] - 1);
  for (int d = 0; d < 3; d++) if ((size_t)cb.size(d) / cw > (size_t)maxLeafWidth) return false;
  const long long vol = (long long)cb.size(0) * cb.size(1) * cb.size(2) * cb.size(3);
  return (size_t)vol == n * (cw * cw * cw);
}

// Split plane of an inner node (:603-734).  vol[d][i] / box[d][i]: used volume and 4-D bounds of coarse slice i along axis
// d (dims[d] slices).  SAH-alike: a plane between two slices is a candidate unless both are full and span the same levels
// (:636-643; an empty slice is never full); its cost is area * volume * level span of the union boxes left and right, in
// double (:695-697); strict < keeps the first minimum in order d = 0,1,2, plane ascending (:705-708).  Without a candidate
// (or for the spatial-median builder): the longest axis at dims/2 (:728-734).
inline void chooseSplit(const Box4 &cb, const int dims[3], int ccw, bool sahAlike, const size_t *const vol[3], const Box4 *const box[3],
                        int &bestDim, int &bestPos)
{
  bestDim = -1; bestPos = -1;
  if (sahAlike) {
    double bestCost = std::numeric_limits<float>::infinity();
    std::vector<Box4> suffix;
    for (int d = 0; d < 3; d++) {
      if (dims[d] == 0) continue;
      const size_t expected = unitCellVolume(cb) / dims[d];
      suffix.assign(dims[d] + 1, Box4());               // suffix[i] = union of slices [i, dims)
      for (int i = dims[d] - 1; i >= 0; i--) { suffix[i] = suffix[i + 1]; suffix[i].extend(box[d][i]); }
      Box4 prefix;                                      // union of slices [0, plane)
      for (int plane = 1; plane < dims[d]; plane++) {
        prefix.extend(box[d][plane - 1]);
        const int L = plane - 1, R = plane;
        const bool lFull = vol[d][L] == expected, rFull = vol[d][R] == expected;
        bool isBoundary = true;
        if (lFull && rFull)
          isBoundary = !(box[d][L].lo[3] == box[d][R].lo[3] && box[d][L].size(3) == box[d][R].size(3));
        if (!isBoundary) continue;
        const Box4 &lb = prefix, &rb = suffix[plane];
        const double cost = areaOf(lb) * (double)unitCellVolume(lb) * lb.size(3)
                          + areaOf(rb) * (double)unitCellVolume(rb) * rb.size(3);
        if (cost < bestCost) { bestCost = cost; bestDim = d; bestPos = cb.lo[d] + plane * ccw; }
      }
    }
  }
  if (bestDim == -1) {
    bestDim = 0;
    for (int i = 1; i < 3; i++) if (dims[i] > dims[bestDim]) bestDim = i;
    bestPos = cb.lo[bestDim] + (dims[bestDim] / 2) * ccw;
  }
}

} // namespace bld
} // namespace exa
