// builder.cpp — cells -> bricks kd builder, bit-exact with the reference's serial exaBuilder
// (builder/builder.cpp:301-350 dedup, 447-530 tryMakeLeaf, 538-811 buildRec, 895-907 writer).
//
// Same decisions, different machinery: cells travel with the recursion as compact
// {x,y,z,level,id} records (streaming passes instead of id -> cell gathers), duplicates are
// removed with a lock-free hash set instead of a 20-byte-pair std::sort (the reference's ID
// order inside a node never influences the output, only WHICH of two identical cells
// survives: the one with the smallest id, builder.cpp:338-342), per-axis slice statistics
// come from one pass, and the left/right union boxes of all candidate planes come from
// prefix/suffix unions instead of an O(planes x slices) rescan.  Passes over large nodes
// are OpenMP-parallel; the recursion itself stays serial (left, then right) because that is
// what defines the brick order.  The records live in two uninitialised arrays allocated once:
// a node partitions its span of one array into the same span of the other (left part first)
// and the children recurse with the roles swapped, so nothing is allocated, zero-filled or
// page-faulted serially on the way down (that, not the passes, dominated at 650 M cells).
#include "exa_internal.h"
#include "builder_common.h"
#include <algorithm>
#include <atomic>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <omp.h>

namespace exa {

struct BricksImpl;
BricksImpl *bricksFromParts(std::vector<ExaBrick> &&hdr, IdVector &&ids, std::vector<ExaKdTreeNode> &&kd);

namespace {

struct Rec { int x, y, z, level, id; };

using bld::Box4; using bld::unitCellVolume; using bld::areaOf; using bld::divDown; using bld::divUp;
inline void extendCell(Box4 &b, const Rec &c) { b.extendCell(c.x, c.y, c.z, c.level); }

static const size_t PAR_THRESHOLD = 1u << 16;

struct Builder {
  int builderType, maxLeafWidth;
  std::vector<ExaBrick> hdr; IdVector ids; std::vector<ExaKdTreeNode> tree;

  // ---- duplicate removal (builder.cpp:301-350) ----
  static inline uint64_t hashCell(const int32_t *c)
  {
    uint64_t h = (uint64_t)(uint32_t)c[0] * 0x9E3779B97F4A7C15ull;
    h ^= (uint64_t)(uint32_t)c[1] * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
    h ^= (uint64_t)(uint32_t)c[2] * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
    h ^= (uint64_t)(uint32_t)c[3] * 0x27D4EB2F165667C5ull + (h << 6) + (h >> 2);
    h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    return h;
  }
  // -> number of surviving records written to out[0..) (out has room for n)
  size_t dedup(const int32_t *cells, size_t n, Rec *out)
  {
    size_t cap = 1; while (cap < 2 * n + 16) cap <<= 1;
    // uninitialised storage, first touched by the parallel fill below
    std::atomic<int32_t> *table = reinterpret_cast<std::atomic<int32_t> *>(malloc(cap * sizeof(int32_t)));
    std::vector<uint8_t> keepStore; uint8_t *keep = nullptr;
    if (!table) throw std::runtime_error("out of memory in the duplicate-removal table");
    struct Free { void *p; ~Free() { free(p); } } freeTable{table};
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < cap; i++) table[i].store(-1, std::memory_order_relaxed);
    auto same = [&](int32_t a, const int32_t *c) { return memcmp(cells + 4 * (size_t)a, c, 16) == 0; };
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i++) {
      const int32_t *c = cells + 4 * i;
      size_t h = hashCell(c) & (cap - 1);
      for (;;) {
        int32_t cur = table[h].load(std::memory_order_acquire);
        if (cur == -1) {
          if (table[h].compare_exchange_strong(cur, (int32_t)i, std::memory_order_acq_rel)) break;
        }
        if (same(cur, c)) {           // slot owned by this cell: keep the smallest id
          while (cur > (int32_t)i && !table[h].compare_exchange_weak(cur, (int32_t)i, std::memory_order_acq_rel)) {}
          break;
        }
        h = (h + 1) & (cap - 1);
      }
    }
    keep = (uint8_t *)malloc(n ? n : 1);
    if (!keep) throw std::runtime_error("out of memory in duplicate removal");
    Free freeKeep{keep};
    const int T = omp_get_max_threads();
    std::vector<size_t> cnt(T + 1, 0);
#pragma omp parallel num_threads(T)
    {
      const int t = omp_get_thread_num();
      const size_t b = n * (size_t)t / T, e = n * (size_t)(t + 1) / T;
      size_t k = 0;
      for (size_t i = b; i < e; i++) {
        const int32_t *c = cells + 4 * i;
        size_t h = hashCell(c) & (cap - 1);
        for (;;) {
          const int32_t cur = table[h].load(std::memory_order_relaxed);
          if (same(cur, c)) { keep[i] = cur == (int32_t)i; k += keep[i]; break; }
          h = (h + 1) & (cap - 1);
        }
      }
      cnt[t + 1] = k;
#pragma omp barrier
#pragma omp single
      for (int q = 0; q < T; q++) cnt[q + 1] += cnt[q];
      size_t p = cnt[t];                                 // ascending id order, as the serial scan
      for (size_t i = b; i < e; i++)
        if (keep[i]) out[p++] = Rec{cells[4 * i], cells[4 * i + 1], cells[4 * i + 2], cells[4 * i + 3], (int)i};
    }
    return cnt[T];
  }

  Box4 boundsOf(const Rec *v, size_t n)
  {
    Box4 b;
    if (n < PAR_THRESHOLD) { for (size_t i = 0; i < n; i++) extendCell(b, v[i]); return b; }
#pragma omp parallel
    {
      Box4 lb;
#pragma omp for schedule(static) nowait
      for (size_t i = 0; i < n; i++) extendCell(lb, v[i]);
#pragma omp critical
      b.extend(lb);
    }
    return b;
  }

  // tryMakeLeaf, builder.cpp:447-530
  bool tryMakeLeaf(const Box4 &b, const Rec *v, size_t n, unsigned nodeID)
  {
    if (n == 0) throw std::runtime_error("empty ID vector in tryMakeLeaf!?");
    if (!bld::isLeaf(b, n, maxLeafWidth)) return false;
    const size_t cw = 1ULL << (b.hi[3] - 1);
    ExaBrick h;
    for (int d = 0; d < 3; d++) { h.lower[d] = b.lo[d]; h.size[d] = b.size(d) / (int)cw; }
    h.level = b.lo[3];
    if (ids.size() + n > 0xffffffffull) throw std::runtime_error("overflow in 32-bit brick offset");
    h.begin = (uint32_t)ids.size();
    const size_t base = ids.size();
    ids.resize(base + n, -1);
    int32_t *dst = ids.data() + base;
    const size_t sx = h.size[0], sxy = (size_t)h.size[0] * h.size[1];
#pragma omp parallel for schedule(static) if (n >= PAR_THRESHOLD)
    for (size_t i = 0; i < n; i++) {
      const Rec &c = v[i];
      dst[(c.x - b.lo[0]) / (int)cw + sx * ((c.y - b.lo[1]) / (int)cw) + sxy * ((c.z - b.lo[2]) / (int)cw)] = c.id;
    }
    // KdTreeNode::set_leaf(first_prim = brick index, count 1), exa/KdTree.h:102-107
    tree[nodeID].split_or_first_prim = (uint32_t)hdr.size();
    tree[nodeID].axis_bits = (3u << 30) | 1u;
    hdr.push_back(h);
    return true;
  }

  struct Slices { std::vector<size_t> vol; std::vector<Box4> box; };

  // Bricker::buildRec, builder.cpp:538-811
  // v = this node's records; o = the same span of the other array (scratch for the partition)
  void buildRec(Rec *v, Rec *o, size_t n, unsigned nodeID)
  {
    Box4 cb = boundsOf(v, n);                              // computeCoarsestLevelBounds :201-225
    const int ccw = bld::snapToCoarsest(cb);
    if (tryMakeLeaf(cb, v, n, nodeID)) return;
    const int dims[3] = {cb.size(0) / ccw, cb.size(1) / ccw, cb.size(2) / ccw};
    if (dims[0] == 1 && dims[1] == 1 && dims[2] == 1) throw std::runtime_error("coarse size 1 that's not a leaf!?");
    const int maxLevel = cb.hi[3] - 1;                   // == max over slice level lists, :592-598

    int bestDim = -1, bestPos = -1;
    if (builderType == EXA_BUILDER_SAH_ALIKE) {
      // ---- per-axis slice statistics, one pass (:557-590) ----
      Slices sl[3];
      for (int d = 0; d < 3; d++) { sl[d].vol.assign(dims[d], 0); sl[d].box.assign(dims[d], Box4()); }
      auto splat = [&](Slices *s, const Rec &c) {
        const int cw = 1 << c.level; const size_t cv = (size_t)cw * cw * cw;
        const int bin[3] = {(c.x - cb.lo[0]) / ccw, (c.y - cb.lo[1]) / ccw, (c.z - cb.lo[2]) / ccw};
        for (int d = 0; d < 3; d++) { s[d].vol[bin[d]] += cv; extendCell(s[d].box[bin[d]], c); }
      };
      if (n < PAR_THRESHOLD) { for (size_t i = 0; i < n; i++) splat(sl, v[i]); }
      else {
#pragma omp parallel
        {
          Slices loc[3];
          for (int d = 0; d < 3; d++) { loc[d].vol.assign(dims[d], 0); loc[d].box.assign(dims[d], Box4()); }
#pragma omp for schedule(static) nowait
          for (size_t i = 0; i < n; i++) splat(loc, v[i]);
#pragma omp critical
          for (int d = 0; d < 3; d++) for (int i = 0; i < dims[d]; i++) { sl[d].vol[i] += loc[d].vol[i]; sl[d].box[i].extend(loc[d].box[i]); }
        }
      }
      // ---- boundary planes and SAH-alike cost (:603-722): builder_common.h ----
      const size_t *const vol[3] = {sl[0].vol.data(), sl[1].vol.data(), sl[2].vol.data()};
      const Box4 *const box[3] = {sl[0].box.data(), sl[1].box.data(), sl[2].box.data()};
      bld::chooseSplit(cb, dims, ccw, true, vol, box, bestDim, bestPos);
    } else {
      const size_t *const vol[3] = {nullptr, nullptr, nullptr}; const Box4 *const box[3] = {nullptr, nullptr, nullptr};
      bld::chooseSplit(cb, dims, ccw, false, vol, box, bestDim, bestPos);   // spatial median, :728-734
    }

    // ---- partition (:753-770): v -> o, left part first, order kept ----
    auto coord = [&](const Rec &c) { return bestDim == 0 ? c.x : bestDim == 1 ? c.y : c.z; };
    size_t nl = 0;
    if (n < PAR_THRESHOLD) {
      for (size_t i = 0; i < n; i++) {
        const int lo = coord(v[i]);
        if (lo >= bestPos) continue;
        if (lo + (1 << v[i].level) > bestPos) throw std::runtime_error("cell straddles split plane!?");
        nl++;
      }
      size_t pl = 0, pr = nl;
      for (size_t i = 0; i < n; i++) { if (coord(v[i]) >= bestPos) o[pr++] = v[i]; else o[pl++] = v[i]; }
    } else {
      const int T = omp_get_max_threads();
      std::vector<size_t> cl(T + 1, 0);
      std::atomic<bool> straddle{false};
#pragma omp parallel num_threads(T)
      {
        const int t = omp_get_thread_num();
        const size_t b = n * (size_t)t / T, e = n * (size_t)(t + 1) / T;
        size_t k = 0;
        for (size_t i = b; i < e; i++) {
          const int lo = coord(v[i]);
          if (lo >= bestPos) continue;
          if (lo + (1 << v[i].level) <= bestPos) k++; else straddle.store(true, std::memory_order_relaxed);
        }
        cl[t + 1] = k;
#pragma omp barrier
#pragma omp single
        for (int q = 0; q < T; q++) cl[q + 1] += cl[q];
        size_t pl = cl[t], pr = cl[T] + (b - cl[t]);
        if (!straddle.load())
          for (size_t i = b; i < e; i++) { if (coord(v[i]) >= bestPos) o[pr++] = v[i]; else o[pl++] = v[i]; }
      }
      if (straddle.load()) throw std::runtime_error("cell straddles split plane!?");
      nl = cl[T];
    }
    if (nl == 0 || nl == n) throw std::runtime_error("invalid split...");
    const unsigned firstChild = (unsigned)tree.size();   // :794-800
    tree.emplace_back(ExaKdTreeNode{0, 0, 0}); tree.emplace_back(ExaKdTreeNode{0, 0, 0});
    tree[nodeID].axis_bits = ((unsigned)bestDim << 30) | firstChild;
    tree[nodeID].split_or_first_prim = (uint32_t)bestPos;
    tree[nodeID].max_level = maxLevel;
    buildRec(o, v, nl, firstChild);                      // :808-810 left, then right
    buildRec(o + nl, v + nl, n - nl, firstChild + 1);
  }
};

} // namespace

BricksImpl *buildBricks(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth)
{
  if (numCells > 0x7fffffffull) throw std::runtime_error("more than 2^31 cells: ids are 32-bit (builder.cpp:45)");
  Builder B; B.builderType = builderType; B.maxLeafWidth = maxLeafWidth;
  B.tree.emplace_back(ExaKdTreeNode{0, 0, 0});
  bool badLevel = false;
#pragma omp parallel for schedule(static) reduction(|| : badLevel)
  for (size_t i = 0; i < numCells; i++) badLevel = badLevel || cells[4 * i + 3] < 0 || cells[4 * i + 3] > 30;
  if (badLevel) throw std::runtime_error("cell level out of range");
  struct Buf { Rec *p = nullptr; ~Buf() { free(p); } } a, b;
  a.p = (Rec *)malloc(sizeof(Rec) * (numCells ? numCells : 1));
  if (!a.p) throw std::runtime_error("out of memory in the builder");
  const size_t n = B.dedup(cells, numCells, a.p);
  b.p = (Rec *)malloc(sizeof(Rec) * (n ? n : 1));
  if (!b.p) throw std::runtime_error("out of memory in the builder");
  if (n) B.buildRec(a.p, b.p, n, 0);
  return bricksFromParts(std::move(B.hdr), std::move(B.ids), std::move(B.tree));
}

} // namespace exa
