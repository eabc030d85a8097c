// abr.cu — GPU build of the active brick regions (ABRs), bit-exact with the reference's
// ExaBrickRegions::buildFrom (exa/Regions.cpp:242-320).
//
// The reference recursion (Regions.cpp:73-179) is stateless in the sense that the prims of
// a node are exactly the bricks whose basis domain (bounds +- 1/2 cell, exa/ExaBricks.h:55-60)
// overlaps the node box with positive volume, kept in brick order; the candidate planes
// are the faces of those (clipped) domains strictly inside the node box.  That lets the
// build run level-synchronously:
//
//   per level:  K_best     one thread per (node,prim) entry: 64-bit atomicMin of the key
//                          (dist-to-centre bits, prim index in node, side) per node and
//                          axis == "strictly closer wins, first seen wins ties" in the
//                          reference's prim-major / upper-before-lower scan order (:89-108)
//               K_decide   per node: widest-axis-first choice (:112-123) or leaf
//               K_flags    per entry: does the clipped prim survive in the left / right
//                          half (:147-158)
//               scans      stable partition offsets (order of prims is preserved)
//               K_scatter  write the two child entry lists
//
//   afterwards: subtree leaf / list counts bottom-up, then offsets top-down with the
//               RIGHT child first (:173-178) give every leaf its region index and leaf-list
//               offset in the reference's emission order; finest level and value range
//               (:182-240, 290-300) are computed by one CTA per (region, brick) pair.
//
// The kd-tree (which the reference discards) is kept: it is the traversal structure of the
// render kernels (render.cu).
#include "exa_internal.h"
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <algorithm>

namespace exa {

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    setError(std::string("CUDA error in abr build: ") + cudaGetErrorString(e_) + " at " #call); return EXA_ERR_CUDA; } } while (0)

namespace {

struct Box3 { float lo[3], hi[3]; };

template <typename T> struct DevBuf {
  T *p = nullptr; size_t cap = 0;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t reserve(size_t n, cudaStream_t s, bool keep = false, size_t used = 0)
  {
    if (n <= cap) return cudaSuccess;
    size_t nc = std::max(n, cap * 2 + 1024);
    T *q = nullptr;
    cudaError_t e = cudaMalloc(&q, nc * sizeof(T));
    if (e != cudaSuccess) return e;
    if (keep && p && used) cudaMemcpyAsync(q, p, used * sizeof(T), cudaMemcpyDeviceToDevice, s);
    if (p) { cudaStreamSynchronize(s); cudaFree(p); }
    p = q; cap = nc; return cudaSuccess;
  }
  T *release() { T *r = p; p = nullptr; cap = 0; return r; }
};

__device__ __forceinline__ void brickDomain(const ExaBrick &b, float lo[3], float hi[3])
{
  // Brick::getDomain, exa/ExaBricks.h:55-60
  const float cw = (float)(1 << b.level);
  lo[0] = (float)b.lower[0] - 0.5f * cw; hi[0] = (float)b.lower[0] + ((float)b.size[0] + 0.5f) * cw;
  lo[1] = (float)b.lower[1] - 0.5f * cw; hi[1] = (float)b.lower[1] + ((float)b.size[1] + 0.5f) * cw;
  lo[2] = (float)b.lower[2] - 0.5f * cw; hi[2] = (float)b.lower[2] + ((float)b.size[2] + 0.5f) * cw;
}

__global__ void k_rootBounds(const ExaBrick *bricks, int nb, unsigned *enc /*6*/)
{
  // order-preserving float->uint encoding; min of lows, max of highs
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nb) return;
  float lo[3], hi[3]; brickDomain(bricks[i], lo, hi);
  for (int d = 0; d < 3; d++) {
    unsigned a = __float_as_uint(lo[d]); a ^= (a >> 31) ? 0xffffffffu : 0x80000000u;
    unsigned b = __float_as_uint(hi[d]); b ^= (b >> 31) ? 0xffffffffu : 0x80000000u;
    atomicMin(&enc[d], a); atomicMax(&enc[3 + d], b);
  }
}

__global__ void k_initRoot(int nb, int *entBrick, int *entNode)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nb) { entBrick[i] = i; entNode[i] = 0; }
}

// one thread per entry: candidate planes -> per (node, axis) 64-bit atomicMin
__global__ void k_best(const ExaBrick *bricks, int m, const int *entBrick, const int *entNode,
                       const Box3 *nodeDom, const int *nodeBegin, unsigned long long *best)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  const int n = entNode[e];
  const Box3 dom = nodeDom[n];
  float lo[3], hi[3]; brickDomain(bricks[entBrick[e]], lo, hi);
  const unsigned local = (unsigned)(e - nodeBegin[n]);
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const float tgt = (dom.lo[d] + dom.hi[d]) / 2.f;      // box3f::center()
#pragma unroll
    for (int side = 0; side < 2; side++) {                // side 0 = upper face first (:91)
      const float pos = side ? fmaxf(lo[d], dom.lo[d]) : fminf(hi[d], dom.hi[d]);   // clipped prim face
      if (pos <= dom.lo[d] || pos >= dom.hi[d]) continue;
      const float dist = fabsf(tgt - pos);
      const unsigned long long key = ((unsigned long long)__float_as_uint(dist) << 32) | (unsigned long long)((local << 1) | (unsigned)side);
      atomicMin(&best[n * 3 + d], key);
    }
  }
}

// per node: pick split (widest axis first, Regions.cpp:110-123) or become a leaf
__global__ void k_decide(const ExaBrick *bricks, int n, const int *entBrick, const Box3 *nodeDom,
                         const int *nodeBegin, const int *nodeCount, const unsigned long long *best,
                         int *splitDim, float *splitPos, int *isSplit, int *isLeaf, int *leafEntries)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Box3 dom = nodeDom[i];
  int sd = -1; float sp = 0.f;
  if (nodeCount[i] > 0) {
    int widest = 0;
    for (int d = 1; d < 3; d++) if ((dom.hi[d] - dom.lo[d]) > (dom.hi[widest] - dom.lo[widest])) widest = d;
    for (int k = 0; k < 3 && sd < 0; k++) {
      const int d = (widest + k) % 3;
      const unsigned long long key = best[i * 3 + d];
      if (key == ~0ull) continue;
      const unsigned low = (unsigned)(key & 0xffffffffu);
      const int side = low & 1; const int local = low >> 1;
      float lo[3], hi[3]; brickDomain(bricks[entBrick[nodeBegin[i] + local]], lo, hi);
      sp = side ? fmaxf(lo[d], dom.lo[d]) : fminf(hi[d], dom.hi[d]);
      sd = d;
    }
  }
  splitDim[i] = sd; splitPos[i] = sp;
  isSplit[i] = sd >= 0;
  const int leaf = (sd < 0 && nodeCount[i] > 0);
  isLeaf[i] = leaf; leafEntries[i] = leaf ? nodeCount[i] : 0;
}

__global__ void k_flags(const ExaBrick *bricks, int m, const int *entBrick, const int *entNode,
                        const Box3 *nodeDom, const int *splitDim, const float *splitPos,
                        int *flagL, int *flagR)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e > m) return;
  if (e == m) { flagL[e] = 0; flagR[e] = 0; return; }     // sentinel so scan[m] = total
  const int n = entNode[e]; const int d = splitDim[n];
  int l = 0, r = 0;
  if (d >= 0) {
    float lo[3], hi[3]; brickDomain(bricks[entBrick[e]], lo, hi);
    const Box3 dom = nodeDom[n];
    const float clo = fmaxf(d == 0 ? lo[0] : d == 1 ? lo[1] : lo[2], dom.lo[d]);
    const float chi = fminf(d == 0 ? hi[0] : d == 1 ? hi[1] : hi[2], dom.hi[d]);
    const float sp = splitPos[n];
    l = clo < fminf(chi, sp);           // intersection with domain_l non-empty in the split axis
    r = fmaxf(clo, sp) < chi;           // ... with domain_r
  }
  flagL[e] = l; flagR[e] = r;
}

// per node: emit kd node, child records, leaf records
__global__ void k_children(int n, int levelBase, int nextBase, int leafBase, int poolBase,
                           const Box3 *nodeDom, const int *nodeBegin, const int *nodeCount,
                           const int *splitDim, const float *splitPos, const int *isSplit,
                           const int *splitRank, const int *leafRank, const int *leafEntryOfs,
                           const int *scanL, const int *scanR,
                           KdNode *kd, Box3 *nextDom, int *nextCount,
                           Box3 *leafDom, int *leafPoolBegin, int *leafCount, int *leafOfNode)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  KdNode node;
  const int b = nodeBegin[i], c = nodeCount[i];
  if (isSplit[i]) {
    const int r = splitRank[i], d = splitDim[i];
    node.split = splitPos[i];
    node.info = ((uint32_t)d << 30) | (uint32_t)(nextBase + 2 * r);
    Box3 dl = nodeDom[i], dr = dl;
    dl.hi[d] = splitPos[i]; dr.lo[d] = splitPos[i];
    nextDom[2 * r] = dl; nextDom[2 * r + 1] = dr;
    nextCount[2 * r] = scanL[b + c] - scanL[b];
    nextCount[2 * r + 1] = scanR[b + c] - scanR[b];
    leafOfNode[levelBase + i] = -1;
  } else if (c > 0) {
    const int lf = leafBase + leafRank[i];
    node.k = (uint32_t)c; node.info = (3u << 30) | (uint32_t)lf;      // temp leaf id, renumbered later
    leafDom[lf] = nodeDom[i]; leafPoolBegin[lf] = poolBase + leafEntryOfs[i]; leafCount[lf] = c;
    leafOfNode[levelBase + i] = lf;
  } else {
    node.k = 0; node.info = (3u << 30) | KD_EMPTY;
    leafOfNode[levelBase + i] = -1;
  }
  kd[levelBase + i] = node;
}

__global__ void k_scatter(int m, int poolBase, const int *entBrick, const int *entNode, const int *nodeBegin,
                          const int *isSplit, const int *isLeaf, const int *splitRank, const int *leafEntryOfs,
                          const int *flagL, const int *flagR, const int *scanL, const int *scanR,
                          const int *nextBegin, int *nextEntBrick, int *nextEntNode, int *pool)
{
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  const int n = entNode[e], b = nodeBegin[n], brick = entBrick[e];
  if (isSplit[n]) {
    const int r = splitRank[n];
    if (flagL[e]) { int dst = nextBegin[2 * r] + (scanL[e] - scanL[b]); nextEntBrick[dst] = brick; nextEntNode[dst] = 2 * r; }
    if (flagR[e]) { int dst = nextBegin[2 * r + 1] + (scanR[e] - scanR[b]); nextEntBrick[dst] = brick; nextEntNode[dst] = 2 * r + 1; }
  } else if (isLeaf[n]) {
    pool[poolBase + leafEntryOfs[n] + (e - b)] = brick;
  }
}

// bottom-up: leaves / list entries below every node
__global__ void k_countUp(int first, int n, const KdNode *kd, const int *leafOfNode, const int *leafCount,
                          int *cntLeaf, int *cntList)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int g = first + i; const KdNode nd = kd[g];
  if ((nd.info >> 30) != 3) { int c = nd.info & 0x3fffffff; cntLeaf[g] = cntLeaf[c] + cntLeaf[c + 1]; cntList[g] = cntList[c] + cntList[c + 1]; }
  else { int lf = leafOfNode[g]; cntLeaf[g] = lf >= 0 ? 1 : 0; cntList[g] = lf >= 0 ? leafCount[lf] : 0; }
}
// top-down: the reference recurses into the RIGHT half first (Regions.cpp:173-178)
__global__ void k_offsetDown(int first, int n, const KdNode *kd, const int *cntLeaf, const int *cntList,
                             int *offLeaf, int *offList)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int g = first + i; const KdNode nd = kd[g];
  if ((nd.info >> 30) == 3) return;
  const int c = nd.info & 0x3fffffff;
  offLeaf[c + 1] = offLeaf[g]; offList[c + 1] = offList[g];
  offLeaf[c] = offLeaf[g] + cntLeaf[c + 1]; offList[c] = offList[g] + cntList[c + 1];
}

__device__ __forceinline__ unsigned encF(float f) { unsigned a = __float_as_uint(f); return a ^ ((a >> 31) ? 0xffffffffu : 0x80000000u); }
__device__ __forceinline__ float decF(unsigned a) { a ^= (a >> 31) ? 0x80000000u : 0xffffffffu; return __uint_as_float(a); }

// per kd node that is a leaf: write region record, leaf list, blob; renumber the node
__global__ void k_finalize(int numNodes, KdNode *kd, const int *leafOfNode, const int *offLeaf, const int *offList,
                           const Box3 *leafDom, const int *leafPoolBegin, const int *leafCount, const int *pool,
                           const ExaBrick *bricks, ExaBrickRegion *regions, int *leafList, int *entryRegion, uint4 *blobs)
{
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= numNodes) return;
  const int lf = leafOfNode[g];
  if (lf < 0) return;
  const int rid = offLeaf[g], lb = offList[g], cnt = leafCount[lf], pb = leafPoolBegin[lf];
  const Box3 dom = leafDom[lf];
  // blob offset (16-B units): BLOB_HDR_UNITS header units per region + 2 per list entry
  const uint32_t blobOfs = (uint32_t)(BLOB_HDR_UNITS * rid + 2 * lb);
  int finest = 1 << 30;
  for (int k = 0; k < cnt; k++) {
    const int bid = pool[pb + k];
    leafList[lb + k] = bid; entryRegion[lb + k] = rid;
    const ExaBrick b = bricks[bid];
    finest = min(finest, b.level);
    BrickHdr h; h.lx = (float)b.lower[0]; h.ly = (float)b.lower[1]; h.lz = (float)b.lower[2];
    h.invCw = 1.f / (float)(1 << b.level); h.sx = b.size[0]; h.sy = b.size[1]; h.sz = b.size[2]; h.begin = 0;   // first record: k_blobBegins
    const uint4 *hp = reinterpret_cast<const uint4 *>(&h);
    blobs[blobOfs + BLOB_HDR_UNITS + 2 * k] = hp[0]; blobs[blobOfs + BLOB_HDR_UNITS + 2 * k + 1] = hp[1];
  }
  ExaBrickRegion r;
  for (int d = 0; d < 3; d++) { r.domain_lower[d] = dom.lo[d]; r.domain_upper[d] = dom.hi[d]; }
  r.valueRange_lower = 0.f; r.valueRange_upper = 0.f;            // filled by k_rangeDecode
  r.leafListBegin = lb; r.leafListSize = cnt; r.finestLevelCellWidth = (float)(1 << finest);
  regions[rid] = r;
  kd[g].k = (uint32_t)cnt; kd[g].info = (3u << 30) | blobOfs;
}

// computeValueRange, Regions.cpp:182-240: one CTA per (region, brick) leaf-list entry.
// regionOwned != null (sort-last data partition): only the regions of this rank are ranged, the
// bricks of the others are not resident
__global__ void k_valueRange(int L, const int *leafList, const int *entryRegion, const ExaBrick *bricks,
                             const ExaBrickRegion *regions, const float *scalars, unsigned long long numCells,
                             int numFields, const uint8_t *regionOwned, unsigned *rangeEnc)
{
  const int j = blockIdx.x;
  if (j >= L) return;
  const int rid = entryRegion[j];
  if (regionOwned && !regionOwned[rid]) return;
  const ExaBrick b = bricks[leafList[j]];
  const ExaBrickRegion r = regions[rid];
  const float cw = (float)(1 << b.level);
  int a[3], n[3];
  for (int d = 0; d < 3; d++) {
    // cell i is "valid" iff pos-cw <= region.hi and pos+cw >= region.lo with pos = lower+(i+.5)cw
    // (all quantities are multiples of 1/4 cell, exact in fp32)
    const float lo = (r.domain_lower[d] - (float)b.lower[d]) / cw - 1.5f;
    const float hi = (r.domain_upper[d] - (float)b.lower[d]) / cw + 0.5f;
    const int i0 = max(0, (int)ceilf(lo)), i1 = min(b.size[d] - 1, (int)floorf(hi));
    a[d] = i0; n[d] = max(0, i1 - i0 + 1);
  }
  const long long total = (long long)n[0] * n[1] * n[2];
  float vmin = INFINITY, vmax = -INFINITY;
  for (int f = 0; f < numFields; f++) {
    const float *sc = scalars + (size_t)f * numCells + b.begin;
    for (long long t = threadIdx.x; t < total; t += blockDim.x) {
      const int ix = a[0] + (int)(t % n[0]); const long long q = t / n[0];
      const int iy = a[1] + (int)(q % n[1]); const int iz = a[2] + (int)(q / n[1]);
      const float s = __ldg(sc + ix + (size_t)b.size[0] * iy + (size_t)b.size[0] * b.size[1] * iz);
      vmin = fminf(vmin, s); vmax = fmaxf(vmax, s);
    }
  }
  for (int o = 16; o > 0; o >>= 1) { vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o)); vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o)); }
  if ((threadIdx.x & 31) == 0 && total > 0) { atomicMin(&rangeEnc[2 * rid], encF(vmin)); atomicMax(&rangeEnc[2 * rid + 1], encF(vmax)); }
}

// first record of every brick header inside the region blobs: the brick's first dual-cell record
// or its first flat cell (both depend on which bricks are resident, so they are attached after
// the structure is built); one thread per leaf-list entry
__global__ void k_blobBegins(int L, const int *leafList, const int *entryRegion, const ExaBrick *bricks, const uint32_t *dualBegin, uint4 *blobs)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= L) return;
  const int bid = leafList[j];
  const size_t ofs = (size_t)BLOB_HDR_UNITS * (entryRegion[j] + 1) + (size_t)2 * j + 1;     // second 16-B unit of the header
  blobs[ofs].w = dualBegin ? dualBegin[bid] : bricks[bid].begin;
}

__global__ void k_rangeInit(int nr, unsigned *rangeEnc)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nr) { rangeEnc[2 * i] = 0xffffffffu; rangeEnc[2 * i + 1] = 0u; }
}

__global__ void k_rangeDecode(int nr, const unsigned *rangeEnc, ExaBrickRegion *regions, uint4 *blobs)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nr) return;
  const unsigned a = rangeEnc[2 * i], b = rangeEnc[2 * i + 1];
  ExaBrickRegion r = regions[i];
  r.valueRange_lower = (a == 0xffffffffu) ? INFINITY : decF(a);
  r.valueRange_upper = (b == 0u) ? -INFINITY : decF(b);
  regions[i] = r;
  RegionBlobHdr h;
  for (int d = 0; d < 3; d++) { h.lo[d] = r.domain_lower[d]; h.hi[d] = r.domain_upper[d]; }
  h.fcw = r.finestLevelCellWidth; h.k = r.leafListSize; h.regionID = i; h.vlo = r.valueRange_lower; h.vhi = r.valueRange_upper; h.pad = 0;
  const uint4 *hp = reinterpret_cast<const uint4 *>(&h);
  const size_t ofs = (size_t)BLOB_HDR_UNITS * i + (size_t)2 * r.leafListBegin;
  blobs[ofs] = hp[0]; blobs[ofs + 1] = hp[1]; blobs[ofs + 2] = hp[2]; blobs[ofs + 3] = make_uint4(0, 0, 0, 0);
}

struct Scanner {
  void *tmp = nullptr; size_t tmpBytes = 0;
  ~Scanner() { if (tmp) cudaFree(tmp); }
  cudaError_t run(const int *in, int *out, int n, cudaStream_t s)
  {
    size_t need = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, need, in, out, n, s);
    if (need > tmpBytes) { if (tmp) { cudaStreamSynchronize(s); cudaFree(tmp); } cudaError_t e = cudaMalloc(&tmp, need); if (e != cudaSuccess) return e; tmpBytes = need; }
    return cub::DeviceScan::ExclusiveSum(tmp, need, in, out, n, s);
  }
};

inline int blocks(long long n, int t = 256) { return (int)((n + t - 1) / t); }

} // namespace

void abrFree(AbrResult &r)
{
  cudaFree(r.regions); cudaFree(r.leafList); cudaFree(r.entryRegion); cudaFree(r.kd); cudaFree(r.blobs);
  r = AbrResult();
}

int abrBuild(const ExaBrick *d_bricks, int nb, void *streamV, AbrResult &out)
{
  cudaStream_t s = (cudaStream_t)streamV;
  out = AbrResult();
  if (nb <= 0) { setError("abrBuild: no bricks"); return EXA_ERR_INVALID; }
  uint64_t launches = 0;

  // ---- root bounds (Regions.cpp:264-278) ----
  DevBuf<unsigned> d_enc; CK(d_enc.reserve(6, s));
  unsigned encInit[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0, 0, 0};
  CK(cudaMemcpyAsync(d_enc.p, encInit, sizeof(encInit), cudaMemcpyHostToDevice, s));
  k_rootBounds<<<blocks(nb), 256, 0, s>>>(d_bricks, nb, d_enc.p); launches++;
  unsigned encOut[6]; CK(cudaMemcpyAsync(encOut, d_enc.p, sizeof(encOut), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  Box3 rootDom;
  for (int d = 0; d < 3; d++) {
    unsigned a = encOut[d], b = encOut[3 + d];
    a ^= (a >> 31) ? 0x80000000u : 0xffffffffu; b ^= (b >> 31) ? 0x80000000u : 0xffffffffu;
    memcpy(&rootDom.lo[d], &a, 4); memcpy(&rootDom.hi[d], &b, 4);
    out.boundsLo[d] = rootDom.lo[d]; out.boundsHi[d] = rootDom.hi[d];
  }

  // ---- level-synchronous build ----
  DevBuf<int> entBrick[2], entNode[2], nodeBegin[2], nodeCount[2];
  DevBuf<Box3> nodeDom[2];
  DevBuf<unsigned long long> best;
  DevBuf<int> splitDim, isSplit, isLeaf, leafEntries, splitRank, leafRank, leafEntryOfs, flagL, flagR, scanL, scanR;
  DevBuf<float> splitPos;
  DevBuf<KdNode> kd; DevBuf<int> leafOfNode;
  DevBuf<Box3> leafDom; DevBuf<int> leafPoolBegin, leafCount, pool;
  Scanner scanner;

  int cur = 0, n = 1, m = nb;
  CK(entBrick[0].reserve(m, s)); CK(entNode[0].reserve(m, s));
  CK(nodeBegin[0].reserve(2, s)); CK(nodeCount[0].reserve(2, s)); CK(nodeDom[0].reserve(1, s));
  k_initRoot<<<blocks(m), 256, 0, s>>>(nb, entBrick[0].p, entNode[0].p); launches++;
  { int z[2] = {0, nb}; CK(cudaMemcpyAsync(nodeBegin[0].p, z, 8, cudaMemcpyHostToDevice, s));
    int c[1] = {nb}; CK(cudaMemcpyAsync(nodeCount[0].p, c, 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(nodeDom[0].p, &rootDom, sizeof(Box3), cudaMemcpyHostToDevice, s)); }

  int levelBase = 0, numLeaves = 0, poolUsed = 0, depth = 0;
  std::vector<int> levelOffset; levelOffset.push_back(0);
  while (n > 0) {
    const int nxt = cur ^ 1;
    CK(best.reserve((size_t)n * 3, s)); CK(cudaMemsetAsync(best.p, 0xff, (size_t)n * 3 * 8, s));
    CK(splitDim.reserve(n, s)); CK(splitPos.reserve(n, s)); CK(isSplit.reserve(n + 1, s)); CK(isLeaf.reserve(n + 1, s));
    CK(leafEntries.reserve(n + 1, s)); CK(splitRank.reserve(n + 1, s)); CK(leafRank.reserve(n + 1, s)); CK(leafEntryOfs.reserve(n + 1, s));
    CK(flagL.reserve(m + 1, s)); CK(flagR.reserve(m + 1, s)); CK(scanL.reserve(m + 1, s)); CK(scanR.reserve(m + 1, s));
    if (m > 0) { k_best<<<blocks(m), 256, 0, s>>>(d_bricks, m, entBrick[cur].p, entNode[cur].p, nodeDom[cur].p, nodeBegin[cur].p, best.p); launches++; }
    k_decide<<<blocks(n), 256, 0, s>>>(d_bricks, n, entBrick[cur].p, nodeDom[cur].p, nodeBegin[cur].p, nodeCount[cur].p, best.p,
                                        splitDim.p, splitPos.p, isSplit.p, isLeaf.p, leafEntries.p); launches++;
    CK(cudaMemsetAsync(isSplit.p + n, 0, 4, s)); CK(cudaMemsetAsync(isLeaf.p + n, 0, 4, s)); CK(cudaMemsetAsync(leafEntries.p + n, 0, 4, s));
    CK(scanner.run(isSplit.p, splitRank.p, n + 1, s)); CK(scanner.run(isLeaf.p, leafRank.p, n + 1, s));
    CK(scanner.run(leafEntries.p, leafEntryOfs.p, n + 1, s)); launches += 3;
    k_flags<<<blocks(m + 1), 256, 0, s>>>(d_bricks, m, entBrick[cur].p, entNode[cur].p, nodeDom[cur].p, splitDim.p, splitPos.p, flagL.p, flagR.p); launches++;
    CK(scanner.run(flagL.p, scanL.p, m + 1, s)); CK(scanner.run(flagR.p, scanR.p, m + 1, s)); launches += 2;
    int h[3];
    CK(cudaMemcpyAsync(&h[0], splitRank.p + n, 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&h[1], leafRank.p + n, 4, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(&h[2], leafEntryOfs.p + n, 4, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    const int numSplit = h[0], newLeaves = h[1], newPool = h[2];
    const int nn = 2 * numSplit;
    CK(kd.reserve((size_t)levelBase + n, s, true, levelBase)); CK(leafOfNode.reserve((size_t)levelBase + n, s, true, levelBase));
    CK(leafDom.reserve((size_t)numLeaves + newLeaves, s, true, numLeaves)); CK(leafPoolBegin.reserve((size_t)numLeaves + newLeaves, s, true, numLeaves));
    CK(leafCount.reserve((size_t)numLeaves + newLeaves, s, true, numLeaves)); CK(pool.reserve((size_t)poolUsed + newPool, s, true, poolUsed));
    CK(nodeDom[nxt].reserve(nn + 1, s)); CK(nodeCount[nxt].reserve(nn + 1, s)); CK(nodeBegin[nxt].reserve(nn + 2, s));
    k_children<<<blocks(n), 256, 0, s>>>(n, levelBase, levelBase + n, numLeaves, poolUsed, nodeDom[cur].p, nodeBegin[cur].p, nodeCount[cur].p,
                                          splitDim.p, splitPos.p, isSplit.p, splitRank.p, leafRank.p, leafEntryOfs.p, scanL.p, scanR.p,
                                          kd.p, nodeDom[nxt].p, nodeCount[nxt].p, leafDom.p, leafPoolBegin.p, leafCount.p, leafOfNode.p); launches++;
    int mNext = 0;
    if (nn > 0) {
      CK(cudaMemsetAsync(nodeCount[nxt].p + nn, 0, 4, s));
      CK(scanner.run(nodeCount[nxt].p, nodeBegin[nxt].p, nn + 1, s)); launches++;
      CK(cudaMemcpyAsync(&mNext, nodeBegin[nxt].p + nn, 4, cudaMemcpyDeviceToHost, s));
      CK(cudaStreamSynchronize(s));
      CK(entBrick[nxt].reserve(mNext + 1, s)); CK(entNode[nxt].reserve(mNext + 1, s));
    }
    if (m > 0) {
      k_scatter<<<blocks(m), 256, 0, s>>>(m, poolUsed, entBrick[cur].p, entNode[cur].p, nodeBegin[cur].p, isSplit.p, isLeaf.p, splitRank.p, leafEntryOfs.p,
                                           flagL.p, flagR.p, scanL.p, scanR.p, nodeBegin[nxt].p, entBrick[nxt].p, entNode[nxt].p, pool.p); launches++;
    }
    levelBase += n; numLeaves += newLeaves; poolUsed += newPool; depth++;
    levelOffset.push_back(levelBase);
    n = nn; m = mNext; cur = nxt;
    if (depth > 4096) { setError("abrBuild: tree too deep"); return EXA_ERR_LIMIT; }
  }
  const int numNodes = levelBase;
  CK(cudaStreamSynchronize(s));

  // ---- right-first DFS numbering ----
  DevBuf<int> cntLeaf, cntList, offLeaf, offList;
  CK(cntLeaf.reserve((size_t)numNodes, s)); CK(cntList.reserve((size_t)numNodes, s));
  CK(offLeaf.reserve((size_t)numNodes, s)); CK(offList.reserve((size_t)numNodes, s));
  for (int lv = depth - 1; lv >= 0; lv--) {
    const int first = levelOffset[lv], cnt = levelOffset[lv + 1] - first;
    k_countUp<<<blocks(cnt), 256, 0, s>>>(first, cnt, kd.p, leafOfNode.p, leafCount.p, cntLeaf.p, cntList.p); launches++;
  }
  CK(cudaMemsetAsync(offLeaf.p, 0, 4, s)); CK(cudaMemsetAsync(offList.p, 0, 4, s));
  for (int lv = 0; lv < depth; lv++) {
    const int first = levelOffset[lv], cnt = levelOffset[lv + 1] - first;
    k_offsetDown<<<blocks(cnt), 256, 0, s>>>(first, cnt, kd.p, cntLeaf.p, cntList.p, offLeaf.p, offList.p); launches++;
  }
  const int nr = numLeaves, L = poolUsed;
  const size_t blobUnits = (size_t)BLOB_HDR_UNITS * nr + (size_t)2 * L;
  if (blobUnits >= KD_EMPTY) { setError("abrBuild: region blob exceeds 30-bit offset"); return EXA_ERR_LIMIT; }
  // the results go into `out` as soon as they exist: on an error below the caller's abrFree(out) releases them
  CK(cudaMalloc(&out.regions, sizeof(ExaBrickRegion) * (size_t)std::max(nr, 1))); CK(cudaMalloc(&out.leafList, 4 * (size_t)std::max(L, 1)));
  CK(cudaMalloc(&out.entryRegion, 4 * (size_t)std::max(L, 1)));
  CK(cudaMalloc(&out.blobs, 16 * std::max(blobUnits, (size_t)1)));
  k_finalize<<<blocks(numNodes), 256, 0, s>>>(numNodes, kd.p, leafOfNode.p, offLeaf.p, offList.p, leafDom.p, leafPoolBegin.p, leafCount.p, pool.p,
                                               d_bricks, out.regions, out.leafList, out.entryRegion, out.blobs); launches++;
  CK(cudaStreamSynchronize(s));
  CK(cudaGetLastError());
  out.numRegions = nr; out.leafListLen = L;
  out.kd = kd.release(); out.numKdNodes = numNodes; out.depth = depth; out.levelOffset = levelOffset;
  out.blobUnits = blobUnits; out.launches = launches;
  return EXA_OK;
}

// Second phase: attach the resident scalars.  Writes the first-record field of the brick headers in the
// region blobs, computes the value ranges (Regions.cpp:182-240) of the regions in d_regionOwned (null: all)
// and completes the region records / blob headers.
int abrAttachScalars(AbrResult &abr, const ExaBrick *d_bricks, const float *d_scalars, uint64_t numCells, int numFields,
                     const uint32_t *d_dualBegin, const uint8_t *d_regionOwned, void *streamV)
{
  cudaStream_t s = (cudaStream_t)streamV;
  const int nr = abr.numRegions, L = abr.leafListLen;
  DevBuf<unsigned> rangeEnc;
  CK(rangeEnc.reserve(2 * (size_t)std::max(nr, 1), s));
  // empty range: [+inf, -inf] after decode
  if (nr > 0) { k_rangeInit<<<blocks(nr), 256, 0, s>>>(nr, rangeEnc.p); abr.launches++; }
  if (L > 0) {
    k_blobBegins<<<blocks(L), 256, 0, s>>>(L, abr.leafList, abr.entryRegion, d_bricks, d_dualBegin, abr.blobs); abr.launches++;
    k_valueRange<<<L, 256, 0, s>>>(L, abr.leafList, abr.entryRegion, d_bricks, abr.regions, d_scalars, (unsigned long long)numCells, numFields,
                                   d_regionOwned, rangeEnc.p); abr.launches++;
  }
  if (nr > 0) { k_rangeDecode<<<blocks(nr), 256, 0, s>>>(nr, rangeEnc.p, abr.regions, abr.blobs); abr.launches++; }
  CK(cudaStreamSynchronize(s));
  CK(cudaGetLastError());
  return EXA_OK;
}

} // namespace exa
