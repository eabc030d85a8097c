// gpu_builder.cu — exaBuilder on the GPU: cells -> bricks, bit-exact with the reference's serial
// builder (builder/builder.cpp:301-350 dedup, 447-530 tryMakeLeaf, 538-811 buildRec, 895-907 writer).
//
// The reference recursion touches every cell of a node several times per level (bounds, three slice histograms, the
// partition); those passes are the cost (10.4 GB of cells on config C3) and they are what runs here as kernels, level by
// level over ALL nodes of the frontier at once:
//     k_dedup*      duplicate removal with a device hash set (smallest id of identical cells survives, :338-342)
//     k_bounds      4-D bounds of every frontier node                      (block reduce + 8 atomics per block)
//     k_slices      per node, axis and coarse slice: used volume, cell count, 4-D bounds
//                   (shared-memory histogram per block, run-length merged per thread, flushed with atomics)
//     k_partition   cells of every inner node to the left / right child span of the other record buffer
//                   (ballot ranks: coalesced stores; one atomic pair per block; order inside a node is free, :107)
//     k_leafScatter cell ids into the bricks (x fastest, :513-522)
// The DECISIONS (coarse snap, leaf test, boundary planes, SAH-alike cost in double, tie-breaks, fallback split) need only
// those small statistics and are taken on the host by the same functions the host builder uses (builder_common.h), which
// the CPU tests pin against the reference's own exaBuilder; brick order and kd numbering follow from a left-first walk of
// the finished tree (:794-810).  Records of a finished leaf simply stay where they are in the ping-pong buffers.
#include "exa_internal.h"
#include "builder_common.h"
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstring>
#include <stdexcept>
#include <omp.h>

namespace exa {

struct BricksImpl;
BricksImpl *bricksFromParts(std::vector<ExaBrick> &&hdr, IdVector &&ids, std::vector<ExaKdTreeNode> &&kd);

namespace {

using bld::Box4;

constexpr int BT = 256;              // threads per block
constexpr int RPT = 8;               // records per thread
constexpr int CHUNK = BT * RPT;      // records per block
constexpr int SMEM_SLICES = 1024;    // slices (all three axes) of a node that fit the shared-memory histogram

#define GCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    throw std::runtime_error(std::string("CUDA error in the GPU builder: ") + cudaGetErrorString(e_) + " at " #call); } while (0)

template <typename T> struct Dev {
  T *p = nullptr; size_t cap = 0;
  ~Dev() { if (p) cudaFree(p); }
  void reserve(size_t n) { if (n <= cap) return; if (p) { cudaFree(p); p = nullptr; } GCK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T))); cap = n; }
  void free() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct Work { unsigned long long start; unsigned node; unsigned pad; };            // one block: CHUNK records of one node from `start`
struct NodeInfo {                                                                   // one frontier node (inner nodes: all fields)
  unsigned long long begin, count, nl;                                              // span in the record buffers; cells left of the plane
  int cbLo[3], ccw, dims[3]; unsigned sliceBase;                                    // coarse bounds / slice layout: x slices, then y, then z
  int splitDim, splitPos; int pad[2];
};
struct LeafInfo { unsigned long long begin, count, outBegin; int lo[3], level, sx, sxy, parity, pad; };

__device__ __forceinline__ unsigned long long hashCell(int4 c)
{
  unsigned long long h = (unsigned long long)(unsigned)c.x * 0x9E3779B97F4A7C15ull;
  h ^= (unsigned long long)(unsigned)c.y * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
  h ^= (unsigned long long)(unsigned)c.z * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
  h ^= (unsigned long long)(unsigned)c.w * 0x27D4EB2F165667C5ull + (h << 6) + (h >> 2);
  h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
  return h;
}
__device__ __forceinline__ bool sameCell(int4 a, int4 b) { return a.x == b.x && a.y == b.y && a.z == b.z && a.w == b.w; }

// ---- duplicate removal (builder.cpp:301-350): every cell enters the hash set; a slot is owned by one (lower, level) key and
// holds the smallest index that carries it
__global__ void k_dedupInsert(const int4 *__restrict__ cells, unsigned long long first, unsigned long long n, int *table, unsigned long long mask, int *badLevel)
{
  const unsigned long long i = first + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4 c = cells[i];
  if (c.w < 0 || c.w > 30) { *badLevel = 1; return; }
  unsigned long long h = hashCell(c) & mask;
  for (;;) {
    int cur = table[h];
    if (cur == -1) { cur = atomicCAS(&table[h], -1, (int)i); if (cur == -1) return; }
    if (sameCell(cells[cur], c)) { atomicMin(&table[h], (int)i); return; }
    h = (h + 1) & mask;
  }
}
__global__ void k_dedupCompact(const int4 *__restrict__ cells, unsigned long long n, const int *__restrict__ table, unsigned long long mask,
                               int4 *outCell, int *outId, unsigned long long *counter)
{
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool keep = false; int4 c = make_int4(0, 0, 0, 0);
  if (i < n) {
    c = cells[i];
    unsigned long long h = hashCell(c) & mask;
    for (;;) {
      const int cur = table[h];
      if (cur < 0) break;                                         // not in the set (a cell k_dedupInsert rejected)
      if (sameCell(cells[cur], c)) { keep = cur == (int)i; break; }
      h = (h + 1) & mask;
    }
  }
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  if (!m) return;
  const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (keep) { const unsigned long long p = base + __popc(m & ((1u << lane) - 1u)); outCell[p] = c; outId[p] = (int)i; }
}

__global__ void k_initBoxes(int *box, unsigned long long n)      // n boxes of 8 ints: lo = INT_MAX, hi = INT_MIN
{
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n * 8ull) box[i] = (i & 7ull) < 4ull ? INT_MAX : INT_MIN;
}

// ---- 4-D bounds of every frontier node (SingleCell::getBounds, :121-130, united over the node)
__global__ void __launch_bounds__(BT) k_bounds(const Work *__restrict__ work, const NodeInfo *__restrict__ nodes, const int4 *__restrict__ cell, int *box)
{
  const Work w = work[blockIdx.x];
  const NodeInfo nd = nodes[w.node];
  const unsigned long long end = min(w.start + (unsigned long long)CHUNK, nd.begin + nd.count);
  int lo[4] = {INT_MAX, INT_MAX, INT_MAX, INT_MAX}, hi[4] = {INT_MIN, INT_MIN, INT_MIN, INT_MIN};
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const unsigned long long idx = w.start + (unsigned long long)i * BT + threadIdx.x;
    if (idx < end) {
      const int4 c = cell[idx]; const int cw = 1 << c.w;
      lo[0] = min(lo[0], c.x); lo[1] = min(lo[1], c.y); lo[2] = min(lo[2], c.z); lo[3] = min(lo[3], c.w);
      hi[0] = max(hi[0], c.x + cw); hi[1] = max(hi[1], c.y + cw); hi[2] = max(hi[2], c.z + cw); hi[3] = max(hi[3], c.w + 1);
    }
  }
#pragma unroll
  for (int k = 0; k < 4; k++)
    for (int o = 16; o > 0; o >>= 1) { lo[k] = min(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o)); hi[k] = max(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o)); }
  __shared__ int s[BT / 32][8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) for (int k = 0; k < 4; k++) { s[warp][k] = lo[k]; s[warp][4 + k] = hi[k]; }
  __syncthreads();
  if (threadIdx.x < 8) {
    int v = s[0][threadIdx.x];
    for (int q = 1; q < BT / 32; q++) v = threadIdx.x < 4 ? min(v, s[q][threadIdx.x]) : max(v, s[q][threadIdx.x]);
    if (threadIdx.x < 4) atomicMin(&box[8 * (size_t)w.node + threadIdx.x], v); else atomicMax(&box[8 * (size_t)w.node + threadIdx.x], v);
  }
}

// ---- per-axis slice statistics (:557-590), one pass: slot = sliceBase + [x slices | y slices | z slices]
struct SliceAcc {
  unsigned long long *vol; unsigned *cnt; int *box;      // box: 8 ints per slot (lo[4], hi[4])
  __device__ __forceinline__ void add(unsigned slot, unsigned long long v, unsigned c) { atomicAdd(&vol[slot], v); atomicAdd(&cnt[slot], c); }
  __device__ __forceinline__ void extend(unsigned slot, int4 c)
  {
    // min / max are monotone: a stale read can only cause a redundant atomic, never miss one
    volatile int *b = box + 8 * (size_t)slot; const int cw = 1 << c.w;
    if (c.x < b[0]) atomicMin(&box[8 * (size_t)slot + 0], c.x);
    if (c.y < b[1]) atomicMin(&box[8 * (size_t)slot + 1], c.y);
    if (c.z < b[2]) atomicMin(&box[8 * (size_t)slot + 2], c.z);
    if (c.w < b[3]) atomicMin(&box[8 * (size_t)slot + 3], c.w);
    if (c.x + cw > b[4]) atomicMax(&box[8 * (size_t)slot + 4], c.x + cw);
    if (c.y + cw > b[5]) atomicMax(&box[8 * (size_t)slot + 5], c.y + cw);
    if (c.z + cw > b[6]) atomicMax(&box[8 * (size_t)slot + 6], c.z + cw);
    if (c.w + 1 > b[7]) atomicMax(&box[8 * (size_t)slot + 7], c.w + 1);
  }
};

__global__ void __launch_bounds__(BT) k_slices(const Work *__restrict__ work, const NodeInfo *__restrict__ nodes, const int4 *__restrict__ cell,
                                               unsigned long long *gVol, unsigned *gCnt, int *gBox)
{
  extern __shared__ unsigned long long smemRaw[];
  const Work w = work[blockIdx.x];
  const NodeInfo nd = nodes[w.node];
  const unsigned S = (unsigned)(nd.dims[0] + nd.dims[1] + nd.dims[2]);
  const bool useSmem = S <= (unsigned)SMEM_SLICES;
  unsigned long long *sVol = smemRaw; int *sBox = reinterpret_cast<int *>(smemRaw + SMEM_SLICES); unsigned *sCnt = reinterpret_cast<unsigned *>(sBox + 8 * SMEM_SLICES);
  if (useSmem) {
    for (unsigned i = threadIdx.x; i < S; i += BT) { sVol[i] = 0; sCnt[i] = 0; }
    for (unsigned i = threadIdx.x; i < 8 * S; i += BT) sBox[i] = (i & 7u) < 4u ? INT_MAX : INT_MIN;
    __syncthreads();
  }
  SliceAcc A;
  unsigned base;
  if (useSmem) { A.vol = sVol; A.cnt = sCnt; A.box = sBox; base = 0; } else { A.vol = gVol; A.cnt = gCnt; A.box = gBox; base = nd.sliceBase; }
  const unsigned off[3] = {base, base + (unsigned)nd.dims[0], base + (unsigned)(nd.dims[0] + nd.dims[1])};
  // each thread takes RPT CONSECUTIVE records: neighbouring cells mostly fall into the same slices, so the volume / count
  // sums of a run are merged in registers and one pair of atomics is issued per run
  const unsigned long long end = min(w.start + (unsigned long long)CHUNK, nd.begin + nd.count);
  const unsigned long long first = w.start + (unsigned long long)threadIdx.x * RPT;
  int curBin[3] = {-1, -1, -1}; unsigned long long runVol[3] = {0, 0, 0}; unsigned runCnt[3] = {0, 0, 0};
  for (int i = 0; i < RPT; i++) {
    const unsigned long long idx = first + i;
    if (idx >= end) break;
    const int4 c = cell[idx];
    const unsigned long long cw = 1ull << c.w, cv = cw * cw * cw;
    const int bin[3] = {(c.x - nd.cbLo[0]) / nd.ccw, (c.y - nd.cbLo[1]) / nd.ccw, (c.z - nd.cbLo[2]) / nd.ccw};
#pragma unroll
    for (int d = 0; d < 3; d++) {
      if (bin[d] != curBin[d]) {
        if (runCnt[d]) A.add(off[d] + (unsigned)curBin[d], runVol[d], runCnt[d]);
        curBin[d] = bin[d]; runVol[d] = 0; runCnt[d] = 0;
      }
      runVol[d] += cv; runCnt[d]++;
      A.extend(off[d] + (unsigned)bin[d], c);
    }
  }
#pragma unroll
  for (int d = 0; d < 3; d++) if (runCnt[d]) A.add(off[d] + (unsigned)curBin[d], runVol[d], runCnt[d]);
  if (useSmem) {
    __syncthreads();
    SliceAcc G; G.vol = gVol; G.cnt = gCnt; G.box = gBox;
    for (unsigned i = threadIdx.x; i < S; i += BT) {
      if (!sCnt[i]) continue;
      const unsigned slot = nd.sliceBase + i;
      G.add(slot, sVol[i], sCnt[i]);
      volatile int *b = gBox + 8 * (size_t)slot;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        if (sBox[8 * i + k] < b[k]) atomicMin(&gBox[8 * (size_t)slot + k], sBox[8 * i + k]);
        if (sBox[8 * i + 4 + k] > b[4 + k]) atomicMax(&gBox[8 * (size_t)slot + 4 + k], sBox[8 * i + 4 + k]);
      }
    }
  }
}

// ---- partition (:753-770): lower >= pos -> right, upper <= pos -> left, anything else is the reference's
// "cell straddles split plane" error.  Left cells go to [begin, begin + nl), right cells behind them, in the other buffer.
__global__ void __launch_bounds__(BT) k_partition(const Work *__restrict__ work, const NodeInfo *__restrict__ nodes,
                                                  const int4 *__restrict__ cellIn, const int *__restrict__ idIn, int4 *cellOut, int *idOut,
                                                  unsigned long long *cursors, int *straddle)
{
  const Work w = work[blockIdx.x];
  const NodeInfo nd = nodes[w.node];
  const unsigned long long end = min(w.start + (unsigned long long)CHUNK, nd.begin + nd.count);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __shared__ unsigned cntL[RPT][BT / 32], cntR[RPT][BT / 32];
  __shared__ unsigned long long baseL, baseR;
  int4 c[RPT]; int id[RPT]; unsigned mL[RPT], mR[RPT];
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const unsigned long long idx = w.start + (unsigned long long)i * BT + threadIdx.x;
    bool left = false, right = false;
    c[i] = make_int4(0, 0, 0, 0); id[i] = 0;
    if (idx < end) {
      c[i] = cellIn[idx]; id[i] = idIn[idx];
      const int lo = nd.splitDim == 0 ? c[i].x : nd.splitDim == 1 ? c[i].y : c[i].z;
      right = lo >= nd.splitPos;
      left = !right;
      if (left && lo + (1 << c[i].w) > nd.splitPos) *straddle = 1;
    }
    mL[i] = __ballot_sync(0xffffffffu, left); mR[i] = __ballot_sync(0xffffffffu, right);
    if (lane == 0) { cntL[i][warp] = __popc(mL[i]); cntR[i][warp] = __popc(mR[i]); }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // exclusive prefix over (iteration, warp) in place, then one atomic pair reserves this block's slots in the child spans
    unsigned tl = 0, tr = 0;
    for (int i = 0; i < RPT; i++) for (int q = 0; q < BT / 32; q++) { const unsigned a = cntL[i][q], b = cntR[i][q]; cntL[i][q] = tl; cntR[i][q] = tr; tl += a; tr += b; }
    baseL = atomicAdd(&cursors[2 * (size_t)w.node], (unsigned long long)tl);
    baseR = atomicAdd(&cursors[2 * (size_t)w.node + 1], (unsigned long long)tr);
  }
  __syncthreads();
  const unsigned below = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const bool left = (mL[i] >> lane) & 1u, right = (mR[i] >> lane) & 1u;
    if (left) { const unsigned long long p = nd.begin + baseL + cntL[i][warp] + __popc(mL[i] & below); cellOut[p] = c[i]; idOut[p] = id[i]; }
    else if (right) { const unsigned long long p = nd.begin + nd.nl + baseR + cntR[i][warp] + __popc(mR[i] & below); cellOut[p] = c[i]; idOut[p] = id[i]; }
  }
}

// ---- brick writer (:513-522): cellIDs[(c.lower - lower) / cw] = ID, x fastest
__global__ void __launch_bounds__(BT) k_leafScatter(const Work *__restrict__ work, const LeafInfo *__restrict__ leaves,
                                                    const int4 *__restrict__ cell0, const int *__restrict__ id0, const int4 *__restrict__ cell1, const int *__restrict__ id1, int *ids)
{
  const Work w = work[blockIdx.x];
  const LeafInfo lf = leaves[w.node];
  const int4 *cell = lf.parity ? cell1 : cell0; const int *id = lf.parity ? id1 : id0;
  const unsigned long long end = min(w.start + (unsigned long long)CHUNK, lf.begin + lf.count);
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const unsigned long long idx = w.start + (unsigned long long)i * BT + threadIdx.x;
    if (idx < end) {
      const int4 c = cell[idx];
      const unsigned long long o = (unsigned long long)((c.x - lf.lo[0]) >> lf.level) + (unsigned long long)lf.sx * (unsigned)((c.y - lf.lo[1]) >> lf.level) +
                                   (unsigned long long)lf.sxy * (unsigned)((c.z - lf.lo[2]) >> lf.level);
      if (o < lf.count) ids[lf.outBegin + o] = id[idx];
    }
  }
}

// Host <-> device transfers of pageable arrays of gigabytes through two pinned staging buffers: the host-side memcpy of
// chunk c+1 (all cores, which is also the parallel first touch of a fresh destination) overlaps the DMA of chunk c.
struct Staging {
  static constexpr size_t CHUNK_BYTES = 64u << 20;
  char *buf[2] = {nullptr, nullptr}; cudaEvent_t ev[2] = {nullptr, nullptr};
  Staging() { for (int i = 0; i < 2; i++) { GCK(cudaMallocHost(&buf[i], CHUNK_BYTES)); GCK(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming)); } }
  ~Staging() { for (int i = 0; i < 2; i++) { if (buf[i]) cudaFreeHost(buf[i]); if (ev[i]) cudaEventDestroy(ev[i]); } }
  static void parallelCopy(char *dst, const char *src, size_t n)
  {
    const int T = std::max(1, std::min(omp_get_max_threads(), 16));
#pragma omp parallel for num_threads(T) schedule(static)
    for (int t = 0; t < T; t++) { const size_t b = n * t / T, e = n * (t + 1) / T; memcpy(dst + b, src + b, e - b); }
  }
  // afterChunk(firstByte, bytes) is called once the chunk's copy is enqueued (stream-ordered work on that range)
  template <typename F> void upload(char *dDst, const char *hSrc, size_t bytes, cudaStream_t s, F afterChunk)
  {
    size_t c = 0;
    for (size_t off = 0; off < bytes; off += CHUNK_BYTES, c++) {
      const size_t n = std::min(CHUNK_BYTES, bytes - off); const int b = (int)(c & 1);
      GCK(cudaEventSynchronize(ev[b]));                              // the DMA that last read this buffer is done
      parallelCopy(buf[b], hSrc + off, n);
      GCK(cudaMemcpyAsync(dDst + off, buf[b], n, cudaMemcpyHostToDevice, s));
      GCK(cudaEventRecord(ev[b], s));
      afterChunk(off, n);
    }
  }
  void download(char *hDst, const char *dSrc, size_t bytes, cudaStream_t s)
  {
    size_t c = 0, prevOff = 0, prevN = 0;
    for (size_t off = 0; off < bytes; off += CHUNK_BYTES, c++) {
      const size_t n = std::min(CHUNK_BYTES, bytes - off); const int b = (int)(c & 1);
      GCK(cudaMemcpyAsync(buf[b], dSrc + off, n, cudaMemcpyDeviceToHost, s));
      GCK(cudaEventRecord(ev[b], s));
      if (c) { GCK(cudaEventSynchronize(ev[b ^ 1])); parallelCopy(hDst + prevOff, buf[b ^ 1], prevN); }
      prevOff = off; prevN = n;
    }
    if (c) { GCK(cudaEventSynchronize(ev[(c - 1) & 1])); parallelCopy(hDst + prevOff, buf[(c - 1) & 1], prevN); }
  }
};

inline unsigned blocksFor(unsigned long long n, int t = 256) { return (unsigned)((n + t - 1) / t); }
inline double nowS() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct TNode { int dim = -1, pos = 0, maxLevel = 0, left = -1, right = -1, leaf = -1; };
struct Leaf { Box4 cb; unsigned long long begin, count; int parity; };

} // namespace

// timings[]: seconds of [0] upload of the cells, [1] duplicate removal, [2] level loop (bounds, slices, decisions,
// partition), [3] brick writer + download of the ids, [4] total; [5] levels, [6] frontier nodes visited
BricksImpl *buildBricksGpu(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth, int cudaDevice, double timings[8])
{
  if (numCells > 0x7fffffffull) throw std::runtime_error("more than 2^31 cells: ids are 32-bit (builder.cpp:45)");
  GCK(cudaSetDevice(cudaDevice));
  cudaStream_t s = nullptr;
  const double tStart = nowS();
  double tm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  std::vector<ExaBrick> hdr; IdVector ids; std::vector<ExaKdTreeNode> kd;
  kd.push_back(ExaKdTreeNode{0, 0, 0});
  if (numCells == 0) { if (timings) memcpy(timings, tm, sizeof(tm)); return bricksFromParts(std::move(hdr), std::move(ids), std::move(kd)); }

  // ---- upload + duplicate removal ----
  Staging staging;
  Dev<int4> recCell[2]; Dev<int> recId[2];
  unsigned long long n = 0;
  {
    Dev<int4> dCells; Dev<int> table, bad; Dev<unsigned long long> counter;
    dCells.reserve(numCells);
    unsigned long long cap = 1; while (cap < 2 * (unsigned long long)numCells + 16) cap <<= 1;
    table.reserve(cap); bad.reserve(1); counter.reserve(1);
    GCK(cudaMemsetAsync(table.p, 0xff, cap * 4, s)); GCK(cudaMemsetAsync(bad.p, 0, 4, s)); GCK(cudaMemsetAsync(counter.p, 0, 8, s));
    // the cells go up in chunks through pinned staging; every chunk enters the hash set as soon as it has arrived
    staging.upload(reinterpret_cast<char *>(dCells.p), reinterpret_cast<const char *>(cells), numCells * 16, s, [&](size_t off, size_t bytes) {
      const unsigned long long i0 = off / 16, i1 = i0 + bytes / 16;
      k_dedupInsert<<<blocksFor(i1 - i0), 256, 0, s>>>(dCells.p, i0, i1, table.p, cap - 1, bad.p);
    });
    GCK(cudaStreamSynchronize(s));
    tm[0] = nowS() - tStart;
    recCell[0].reserve(numCells); recId[0].reserve(numCells);
    k_dedupCompact<<<blocksFor(numCells), 256, 0, s>>>(dCells.p, numCells, table.p, cap - 1, recCell[0].p, recId[0].p, counter.p);
    int hBad = 0;
    GCK(cudaMemcpyAsync(&hBad, bad.p, 4, cudaMemcpyDeviceToHost, s)); GCK(cudaMemcpyAsync(&n, counter.p, 8, cudaMemcpyDeviceToHost, s));
    GCK(cudaStreamSynchronize(s)); GCK(cudaGetLastError());
    if (hBad) throw std::runtime_error("cell level out of range");
  }
  recCell[1].reserve(n); recId[1].reserve(n);
  tm[1] = nowS() - tStart - tm[0];

  // ---- level loop ----
  std::vector<TNode> tree(1);
  std::vector<Leaf> leaves;
  struct FNode { unsigned long long begin, count; int tnode; };
  std::vector<FNode> frontier; frontier.push_back(FNode{0, n, 0});
  Dev<Work> dWork; Dev<NodeInfo> dNodes; Dev<int> dBox, dSliceBox, dStraddle; Dev<unsigned long long> dVol, dCursors; Dev<unsigned> dCnt;
  dStraddle.reserve(1); GCK(cudaMemsetAsync(dStraddle.p, 0, 4, s));
  std::vector<Work> work; std::vector<NodeInfo> infos; std::vector<Box4> boxes;
  std::vector<unsigned long long> hVol; std::vector<unsigned> hCnt; std::vector<Box4> hSliceBox;
  int parity = 0, levels = 0; unsigned long long visited = 0;
  const double tLoop = nowS();
  auto makeWork = [&](const std::vector<NodeInfo> &nodes) {
    work.clear();
    for (size_t f = 0; f < nodes.size(); f++)
      for (unsigned long long o = 0; o < nodes[f].count; o += CHUNK) work.push_back(Work{nodes[f].begin + o, (unsigned)f, 0u});
    dWork.reserve(work.size());
    GCK(cudaMemcpyAsync(dWork.p, work.data(), work.size() * sizeof(Work), cudaMemcpyHostToDevice, s));
  };
  while (!frontier.empty()) {
    const size_t F = frontier.size(); levels++; visited += F;
    // bounds of every frontier node
    infos.assign(F, NodeInfo());
    for (size_t f = 0; f < F; f++) { infos[f].begin = frontier[f].begin; infos[f].count = frontier[f].count; }
    dNodes.reserve(F); GCK(cudaMemcpyAsync(dNodes.p, infos.data(), F * sizeof(NodeInfo), cudaMemcpyHostToDevice, s));
    makeWork(infos);
    dBox.reserve(8 * F);
    k_initBoxes<<<blocksFor(8ull * F), 256, 0, s>>>(dBox.p, (unsigned long long)F);
    k_bounds<<<(unsigned)work.size(), BT, 0, s>>>(dWork.p, dNodes.p, recCell[parity].p, dBox.p);
    boxes.resize(F);
    GCK(cudaMemcpyAsync(boxes.data(), dBox.p, F * sizeof(Box4), cudaMemcpyDeviceToHost, s));
    GCK(cudaStreamSynchronize(s)); GCK(cudaGetLastError());
    // leaf test / slice layout (host, builder_common.h)
    std::vector<NodeInfo> inner; std::vector<size_t> innerOf; std::vector<Box4> innerBox;
    unsigned long long totalSlices = 0;
    for (size_t f = 0; f < F; f++) {
      Box4 cb = boxes[f];
      const int ccw = bld::snapToCoarsest(cb);
      if (frontier[f].count == 0) throw std::runtime_error("empty ID vector in tryMakeLeaf!?");
      if (bld::isLeaf(cb, frontier[f].count, maxLeafWidth)) {
        tree[frontier[f].tnode].leaf = (int)leaves.size();
        leaves.push_back(Leaf{cb, frontier[f].begin, frontier[f].count, parity});
        continue;
      }
      NodeInfo ni = infos[f];
      for (int d = 0; d < 3; d++) { ni.cbLo[d] = cb.lo[d]; ni.dims[d] = cb.size(d) / ccw; }
      ni.ccw = ccw;
      if (ni.dims[0] == 1 && ni.dims[1] == 1 && ni.dims[2] == 1) throw std::runtime_error("coarse size 1 that's not a leaf!?");
      if (totalSlices + (unsigned long long)ni.dims[0] + ni.dims[1] + ni.dims[2] > 0x0fffffffull) throw std::runtime_error("GPU builder: too many slices in one level");
      ni.sliceBase = (unsigned)totalSlices; totalSlices += (unsigned long long)ni.dims[0] + ni.dims[1] + ni.dims[2];
      tree[frontier[f].tnode].maxLevel = cb.hi[3] - 1;       // :592-598
      inner.push_back(ni); innerOf.push_back(f); innerBox.push_back(cb);
    }
    if (inner.empty()) break;
    // slice statistics of every inner node
    const size_t I = inner.size();
    dNodes.reserve(I); GCK(cudaMemcpyAsync(dNodes.p, inner.data(), I * sizeof(NodeInfo), cudaMemcpyHostToDevice, s));
    makeWork(inner);
    dVol.reserve(totalSlices); dCnt.reserve(totalSlices); dSliceBox.reserve(8 * totalSlices);
    GCK(cudaMemsetAsync(dVol.p, 0, totalSlices * 8, s)); GCK(cudaMemsetAsync(dCnt.p, 0, totalSlices * 4, s));
    k_initBoxes<<<blocksFor(8ull * totalSlices), 256, 0, s>>>(dSliceBox.p, totalSlices);
    const size_t smem = (size_t)SMEM_SLICES * (8 + 32 + 4);
    k_slices<<<(unsigned)work.size(), BT, smem, s>>>(dWork.p, dNodes.p, recCell[parity].p, dVol.p, dCnt.p, dSliceBox.p);
    hVol.resize(totalSlices); hCnt.resize(totalSlices); hSliceBox.resize(totalSlices);
    GCK(cudaMemcpyAsync(hVol.data(), dVol.p, totalSlices * 8, cudaMemcpyDeviceToHost, s));
    GCK(cudaMemcpyAsync(hCnt.data(), dCnt.p, totalSlices * 4, cudaMemcpyDeviceToHost, s));
    GCK(cudaMemcpyAsync(hSliceBox.data(), dSliceBox.p, totalSlices * sizeof(Box4), cudaMemcpyDeviceToHost, s));
    GCK(cudaStreamSynchronize(s)); GCK(cudaGetLastError());
    // split decisions (host, builder_common.h: the same code as the host builder)
    static_assert(sizeof(size_t) == sizeof(unsigned long long), "size_t");
    bool invalid = false;
#pragma omp parallel for schedule(dynamic, 16) if (I > 64)
    for (long long k = 0; k < (long long)I; k++) {
      NodeInfo &ni = inner[k];
      const size_t b0 = ni.sliceBase, b1 = b0 + ni.dims[0], b2 = b1 + ni.dims[1];
      const size_t *const vol[3] = {reinterpret_cast<const size_t *>(hVol.data()) + b0, reinterpret_cast<const size_t *>(hVol.data()) + b1, reinterpret_cast<const size_t *>(hVol.data()) + b2};
      const Box4 *const box[3] = {hSliceBox.data() + b0, hSliceBox.data() + b1, hSliceBox.data() + b2};
      int dim, pos;
      bld::chooseSplit(innerBox[k], ni.dims, ni.ccw, builderType == EXA_BUILDER_SAH_ALIKE, vol, box, dim, pos);
      const int plane = (pos - innerBox[k].lo[dim]) / ni.ccw;
      const size_t bd = dim == 0 ? b0 : dim == 1 ? b1 : b2;
      unsigned long long nl = 0;
      for (int i = 0; i < plane; i++) nl += hCnt[bd + i];
      ni.splitDim = dim; ni.splitPos = pos; ni.nl = nl;
      if (nl == 0 || nl == ni.count) invalid = true;
    }
    if (invalid) {
      // nl counts cells by the slice of their LOWER corner; a cell that straddles the plane makes the reference throw first
      throw std::runtime_error("invalid split...");
    }
    std::vector<FNode> next; next.reserve(2 * I);
    for (size_t k = 0; k < I; k++) {
      const NodeInfo &ni = inner[k];
      const int tn = frontier[innerOf[k]].tnode;
      const int li = (int)tree.size();
      tree.push_back(TNode()); tree.push_back(TNode());
      tree[tn].dim = ni.splitDim; tree[tn].pos = ni.splitPos; tree[tn].left = li; tree[tn].right = li + 1;
      next.push_back(FNode{ni.begin, ni.nl, li}); next.push_back(FNode{ni.begin + ni.nl, ni.count - ni.nl, li + 1});
    }
    // partition into the other buffer
    GCK(cudaMemcpyAsync(dNodes.p, inner.data(), I * sizeof(NodeInfo), cudaMemcpyHostToDevice, s));
    dCursors.reserve(2 * I); GCK(cudaMemsetAsync(dCursors.p, 0, 2 * I * 8, s));
    k_partition<<<(unsigned)work.size(), BT, 0, s>>>(dWork.p, dNodes.p, recCell[parity].p, recId[parity].p, recCell[parity ^ 1].p, recId[parity ^ 1].p, dCursors.p, dStraddle.p);
    int hStraddle = 0;
    GCK(cudaMemcpyAsync(&hStraddle, dStraddle.p, 4, cudaMemcpyDeviceToHost, s));
    GCK(cudaStreamSynchronize(s)); GCK(cudaGetLastError());
    if (hStraddle) throw std::runtime_error("cell straddles split plane!?");
    parity ^= 1; frontier.swap(next);
    if (levels > 4096) throw std::runtime_error("GPU builder: tree too deep");
  }
  tm[2] = nowS() - tLoop; tm[5] = levels; tm[6] = (double)visited;

  // ---- left-first walk: brick order, kd numbering (:794-810), brick headers (:447-530) ----
  const double tOut = nowS();
  std::vector<LeafInfo> linfo(leaves.size());
  hdr.reserve(leaves.size());
  unsigned long long total = 0;
  {
    struct Item { int tnode; unsigned kdID; };
    std::vector<Item> st; st.push_back(Item{0, 0});
    while (!st.empty()) {
      const Item it = st.back(); st.pop_back();
      const TNode &t = tree[it.tnode];
      if (t.leaf >= 0) {
        const Leaf &lf = leaves[t.leaf];
        const int level = lf.cb.lo[3];
        ExaBrick h;
        for (int d = 0; d < 3; d++) { h.lower[d] = lf.cb.lo[d]; h.size[d] = lf.cb.size(d) >> level; }
        h.level = level;
        if (total + lf.count > 0xffffffffull) throw std::runtime_error("overflow in 32-bit brick offset");
        h.begin = (uint32_t)total;
        LeafInfo &li = linfo[hdr.size()];
        li.begin = lf.begin; li.count = lf.count; li.outBegin = total; li.level = level; li.parity = lf.parity;
        for (int d = 0; d < 3; d++) li.lo[d] = lf.cb.lo[d];
        li.sx = h.size[0]; li.sxy = h.size[0] * h.size[1]; li.pad = 0;
        kd[it.kdID].split_or_first_prim = (uint32_t)hdr.size();
        kd[it.kdID].axis_bits = (3u << 30) | 1u;            // KdTreeNode::set_leaf(first_prim, 1), exa/KdTree.h:102-107
        hdr.push_back(h); total += lf.count;
        continue;
      }
      const unsigned firstChild = (unsigned)kd.size();
      kd.push_back(ExaKdTreeNode{0, 0, 0}); kd.push_back(ExaKdTreeNode{0, 0, 0});
      kd[it.kdID].axis_bits = ((unsigned)t.dim << 30) | firstChild;
      kd[it.kdID].split_or_first_prim = (uint32_t)t.pos;
      kd[it.kdID].max_level = t.maxLevel;
      // NOTE the reference appends a node's children when the node is processed, i.e. in pre-order of the left-first
      // recursion: the left subtree's nodes are all numbered before the right child's children
      st.push_back(Item{t.right, firstChild + 1}); st.push_back(Item{t.left, firstChild});
    }
  }
  // ---- brick writer ----
  ids.resize(total);
  if (total) {
    Dev<int> dIds; Dev<LeafInfo> dLeaves;
    dIds.reserve(total); dLeaves.reserve(linfo.size());
    GCK(cudaMemsetAsync(dIds.p, 0xff, total * 4, s));
    GCK(cudaMemcpyAsync(dLeaves.p, linfo.data(), linfo.size() * sizeof(LeafInfo), cudaMemcpyHostToDevice, s));
    work.clear();
    for (size_t f = 0; f < linfo.size(); f++)
      for (unsigned long long o = 0; o < linfo[f].count; o += CHUNK) work.push_back(Work{linfo[f].begin + o, (unsigned)f, 0u});
    dWork.reserve(work.size());
    GCK(cudaMemcpyAsync(dWork.p, work.data(), work.size() * sizeof(Work), cudaMemcpyHostToDevice, s));
    k_leafScatter<<<(unsigned)work.size(), BT, 0, s>>>(dWork.p, dLeaves.p, recCell[0].p, recId[0].p, recCell[1].p, recId[1].p, dIds.p);
    GCK(cudaGetLastError());
    staging.download(reinterpret_cast<char *>(ids.data()), reinterpret_cast<const char *>(dIds.p), total * 4, s);
  }
  tm[3] = nowS() - tOut; tm[4] = nowS() - tStart;
  if (timings) memcpy(timings, tm, sizeof(tm));
  return bricksFromParts(std::move(hdr), std::move(ids), std::move(kd));
}

} // namespace exa
