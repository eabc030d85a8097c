// capi.cu — exa_renderer_* : the C ABI over the CUDA path (replaces exa::OptixRenderer,
// exa/OptixRenderer.{h,cpp}).  No CPU fallback anywhere: every entry point that needs the
// device fails with EXA_ERR_CUDA when CUDA is unusable.
#include "render.cuh"
#include <algorithm>
#include <vector>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <omp.h>

using namespace exa;

#define CKR(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    setError(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); return EXA_ERR_CUDA; } } while (0)

struct exa_renderer {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint64_t launches = 0;
  // scene
  ExaBrick *d_bricks = nullptr; int numBricks = 0;
  float *d_scalars = nullptr; uint64_t numCells = 0; int numFields = 0;
  unsigned *d_scalarOffsets = nullptr; int *d_primaryChannels = nullptr;
  uint2 *d_locGrid = nullptr; int locG[3] = {0, 0, 0}; float locScale[3] = {0, 0, 0};   // point-location grid
  float4 *d_dual = nullptr; uint64_t dualRecords = 0; uint32_t *d_dualBegin = nullptr;   // dual-cell records (render.cu:k_expandDual)
  AbrResult abr;
  std::vector<KdNode> h_kd;              // host copy (partitioning decisions)
  float voxelLo[3], voxelHi[3], worldLo[3], worldHi[3];
  // frame state
  ExaFrameState fs; ExaLaunchParams lp;
  float4 *d_xf = nullptr;                // [numFields][128]
  float dt = 0.5f; float ert = 0.98f;
  int spaceSkipping = 1, gradDVR = 1, gradISO = 1, gradCD = 0;
  bool masksDirty = true;
  uint8_t *d_volActive = nullptr, *d_isoActive = nullptr /* = d_volActive + numKdNodes */, *d_regionActive = nullptr /* = d_volActive + 2 numKdNodes */, *d_inPart = nullptr; int *d_kdParent = nullptr;
  // framebuffer
  int width = 0, height = 0;
  float4 *d_accum = nullptr; uint32_t *d_rgba8 = nullptr; bool ownRgba8 = false;
  uint32_t *strips[EXA_MAX_PARTS] = {}; int numStrips = 0; int stripRow[EXA_MAX_PARTS + 1] = {};   // exa_renderer_set_color_strips[_rows]
  uint32_t *d_tileList = nullptr; int tileListLen = 0; size_t tileListCap = 0;   // exa_renderer_set_tile_list (balanced image tiles)
  uint32_t *d_tileCost = nullptr; int tileCostLen = 0;
  // pipelined read-back (exa_renderer_read_rgba8_async): second renderer-owned RGBA8 buffer, copy stream
  uint32_t *d_rgba8Alt = nullptr; cudaStream_t copyStream = nullptr; cudaEvent_t evRendered = nullptr, evCopied[2] = {nullptr, nullptr}; int rbCur = 0;
  // pinned staging ring for updateXF (no stream synchronise per update)
  float4 *h_xfRing = nullptr; cudaEvent_t evXf[8] = {}; unsigned xfRingPos = 0;
  // partition
  int partMode = EXA_PARTITION_NONE, rank = 0, numRanks = 1; int partRoot = 0;
  // sort-last: partition tree over the top of the ABR kd-tree (buildPartitionTree); leaf: left < 0, owner = rank
  struct PartNode { int kd; int left, right; int rank; };
  std::vector<PartNode> partTree;
  int lastGrid[4] = {0, 0, 0, 0};        // tile rectangle of the last sort-last partial (a change restarts the cost-ordered tile schedule)
  bool dataPartitioned = false;          // exa_renderer_create_partitioned: only this rank's bricks are resident, the rest of the kd-tree is pruned
  int residentBricks = 0;
  double buildMs[4] = {0, 0, 0, 0};      // ABR geometry, scalar gather, dual records, value ranges (device time)
  // tracer
  int tracerEnabled = 0, numTraces = 1000, numTimesteps = 100, timestep = 0;   // numTraces x numTimesteps = the shape of d_traces
  int pendNumTraces = 1000, pendNumTimesteps = 100;                           // set_tracer_params; committed when d_traces is rebuilt
  float steplen = 1e-6f; int tracerChannels[3] = {0, 1, 2};
  float seedLo[3] = {.3f, .3f, .5f}, seedHi[3] = {.8f, .8f, .5f};
  float *d_traces = nullptr; size_t tracesCap = 0;
  float4 *d_meshNodes = nullptr; float *d_meshTris = nullptr; int numMeshTris = 0;          // triangle-mesh BVH (set_triangle_meshes)
  float *d_traceBounds = nullptr; int traceBoundsCap = 0; bool traceBoundsDirty = true;   // one box per trace (k_traceBounds) -> grid bounds
  int *d_coneStart = nullptr, *d_coneCursor = nullptr, *d_coneList = nullptr; size_t coneListCap = 0; bool coneGridValid = false;
  float coneLo[3] = {0, 0, 0}, coneCell[3] = {1, 1, 1};                                  // uniform grid over the visible cones (buildConeGrid)
  unsigned long long *d_stats = nullptr;
  // exa_sample_points_host: two sets of pinned / device staging buffers for STAGE_POINTS points each
  static constexpr size_t STAGE_POINTS = 262144;
  char *h_stage[2] = {nullptr, nullptr}, *d_stage[2] = {nullptr, nullptr}; cudaEvent_t evStage[2] = {nullptr, nullptr};
  // tile cost order, double buffered (render.cuh:TileOrder); valid once a frame has written it
  TileOrder *d_tileOrder[2] = {nullptr, nullptr}; int tileOrderCap = 0, tileOrderCur = 0; bool tileOrderValid = false;
};

static int cudaUsable()
{
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) { setError(std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0")); cudaGetLastError(); return 0; }
  return n;
}

static int g_scalarLayout = EXA_LAYOUT_AUTO;

static void identityXfm(float m[12]) { memset(m, 0, 48); m[0] = m[4] = m[8] = 1.f; }

// Pixel rectangle that contains every pixel whose primary ray can hit the scene bounds (the ABR root domain): the
// projections of its 8 corners, grown by 2 pixels (sub-pixel jitter, rounding).  Conservative: the whole frame when the
// camera is inside the box, a corner is not in front of it, or the camera basis is degenerate.  Tiles outside the
// rectangle skip ray set-up and traversal and write the background directly (render.cu).
static void boxScreenRect(const exa_renderer *r, const float boxLo[3], const float boxHi[3], int rect[4])
{
  rect[0] = 0; rect[1] = 0; rect[2] = r->width - 1; rect[3] = r->height - 1;
  const float *m = (const float *)&r->fs.voxelSpaceTransform;      // voxel = L * world + p  (columns vx, vy, vz)
  const double L[9] = {m[0], m[3], m[6], m[1], m[4], m[7], m[2], m[5], m[8]};   // row-major
  auto inv3 = [](const double A[9], double B[9]) {
    const double c0 = A[4] * A[8] - A[5] * A[7], c1 = A[5] * A[6] - A[3] * A[8], c2 = A[3] * A[7] - A[4] * A[6];
    const double det = A[0] * c0 + A[1] * c1 + A[2] * c2;
    if (!(std::fabs(det) > 1e-300)) return false;
    B[0] = c0 / det; B[1] = (A[2] * A[7] - A[1] * A[8]) / det; B[2] = (A[1] * A[5] - A[2] * A[4]) / det;
    B[3] = c1 / det; B[4] = (A[0] * A[8] - A[2] * A[6]) / det; B[5] = (A[2] * A[3] - A[0] * A[5]) / det;
    B[6] = c2 / det; B[7] = (A[1] * A[6] - A[0] * A[7]) / det; B[8] = (A[0] * A[4] - A[1] * A[3]) / det;
    return true;
  };
  double Li[9], Mi[9];
  const ExaFrameState &fs = r->fs;
  const double M[9] = {fs.camera.dir00[0], fs.camera.dirDu[0], fs.camera.dirDv[0], fs.camera.dir00[1], fs.camera.dirDu[1], fs.camera.dirDv[1],
                       fs.camera.dir00[2], fs.camera.dirDu[2], fs.camera.dirDv[2]};
  if (!inv3(L, Li) || !inv3(M, Mi)) return;
  double lo[2] = {1e300, 1e300}, hi[2] = {-1e300, -1e300};
  for (int c = 0; c < 8; c++) {
    const double v[3] = {(c & 1 ? boxHi[0] : boxLo[0]) - (double)m[9], (c & 2 ? boxHi[1] : boxLo[1]) - (double)m[10], (c & 4 ? boxHi[2] : boxLo[2]) - (double)m[11]};
    double q[3];
    for (int k = 0; k < 3; k++) q[k] = Li[3 * k] * v[0] + Li[3 * k + 1] * v[1] + Li[3 * k + 2] * v[2] - fs.camera.pos[k];
    const double a = Mi[0] * q[0] + Mi[1] * q[1] + Mi[2] * q[2], b = Mi[3] * q[0] + Mi[4] * q[1] + Mi[5] * q[2], d = Mi[6] * q[0] + Mi[7] * q[1] + Mi[8] * q[2];
    if (!(a > 1e-6 * (std::fabs(b) + std::fabs(d)) && a > 0.0)) return;       // corner beside / behind the camera: no bound
    lo[0] = std::min(lo[0], b / a); hi[0] = std::max(hi[0], b / a); lo[1] = std::min(lo[1], d / a); hi[1] = std::max(hi[1], d / a);
  }
  if (!(lo[0] <= hi[0]) || !std::isfinite(lo[0]) || !std::isfinite(hi[0]) || !std::isfinite(lo[1]) || !std::isfinite(hi[1])) return;
  auto clampi = [](double x, int a, int b) { return (int)std::min((double)b, std::max((double)a, x)); };
  rect[0] = clampi(std::floor(lo[0]) - 3.0, 0, r->width); rect[1] = clampi(std::floor(lo[1]) - 3.0, 0, r->height);
  rect[2] = clampi(std::ceil(hi[0]) + 2.0, -1, r->width - 1); rect[3] = clampi(std::ceil(hi[1]) + 2.0, -1, r->height - 1);
}

// Box of sort-last part `rank`: the scene bounds cut by the split planes on the way down the partition tree (a prefix of the
// ABR kd-tree, so every part is a box and all of the part's regions lie inside it).  false: that rank owns nothing.
static bool partBoxOf(const exa_renderer *r, int rank, float lo[3], float hi[3])
{
  struct Item { int node; float lo[3], hi[3]; };
  if (r->partTree.empty()) return false;
  std::vector<Item> st(1);
  st[0].node = 0; for (int d = 0; d < 3; d++) { st[0].lo[d] = r->abr.boundsLo[d]; st[0].hi[d] = r->abr.boundsHi[d]; }
  while (!st.empty()) {
    const Item it = st.back(); st.pop_back();
    const exa_renderer::PartNode &pn = r->partTree[it.node];
    if (pn.left < 0) {
      if (pn.rank != rank) continue;
      memcpy(lo, it.lo, 12); memcpy(hi, it.hi, 12); return true;
    }
    const KdNode nd = r->h_kd[pn.kd];
    const int axis = (int)(nd.info >> 30);
    Item l = it, rr = it;
    l.node = pn.left; l.hi[axis] = nd.split; rr.node = pn.right; rr.lo[axis] = nd.split;
    st.push_back(rr); st.push_back(l);
  }
  return false;
}
static inline bool sortLastParts(const exa_renderer *r) { return r->partMode == EXA_PARTITION_SORT_LAST && !r->partTree.empty(); }

static void fillParams(exa_renderer *r, RenderParams &P)
{
  memset(&P, 0, sizeof(P));
  DeviceScene &S = P.scene;
  S.bricks = r->d_bricks; S.numBricks = r->numBricks; S.scalars = r->d_scalars; S.numCells = r->numCells; S.numFields = r->numFields;
  S.regions = r->abr.regions; S.numRegions = r->abr.numRegions; S.leafList = r->abr.leafList; S.leafListLen = r->abr.leafListLen;
  S.dual = r->d_dual; S.dualRecords = r->dualRecords;
  S.locGrid = r->d_locGrid; for (int d = 0; d < 3; d++) { S.locG[d] = r->locG[d]; S.locScale[d] = r->locScale[d]; }
  S.kd = r->abr.kd; S.numKdNodes = r->abr.numKdNodes; S.kdDepth = r->abr.depth; S.blobs = r->abr.blobs;
  for (int d = 0; d < 3; d++) { S.boundsLo[d] = r->abr.boundsLo[d]; S.boundsHi[d] = r->abr.boundsHi[d]; }
  memcpy(P.camPos, r->fs.camera.pos, 12); memcpy(P.dir00, r->fs.camera.dir00, 12);
  memcpy(P.dirDu, r->fs.camera.dirDu, 12); memcpy(P.dirDv, r->fs.camera.dirDv, 12);
  memcpy(P.xfm, &r->fs.voxelSpaceTransform, 48);
  {
    // xfmPoint(xfm, camPos) exactly as the device evaluates it (render.cu: products and sums rounded separately, no FMA)
    const float *m = P.xfm, *v = P.camPos;
    for (int k = 0; k < 3; k++) {
      volatile float p0 = v[0] * m[k], p1 = v[1] * m[3 + k], p2 = v[2] * m[6 + k];
      volatile float s01 = p0 + p1; volatile float s012 = s01 + p2;
      volatile float q = s012 + m[9 + k];
      P.camVoxel[k] = q;
    }
  }
  P.width = r->width; P.height = r->height; P.frameID = r->fs.frameID; P.dt = r->dt;
  P.numChannels = r->numFields; P.colormapChannel = 0;     // multiFieldDvr (OptixRenderer.h:110, .cpp:276-283)
  P.gradDVR = r->gradDVR; P.gradISO = r->gradISO; P.gradCD = r->gradCD;
  P.xfOpacityScale = r->fs.xfOpacityScale; P.clockScale = r->fs.clockScale;
  memcpy(P.xfDomain, r->fs.xfDomain, sizeof(P.xfDomain));
  P.xf = r->d_xf;
  for (int i = 0; i < EXA_MAX_ISO_SURFACES; i++) {
    P.isoEnabled[i] = r->fs.isoSurface[i].enabled; P.isoValue[i] = r->fs.isoSurface[i].value; P.isoChannel[i] = r->fs.isoSurface[i].channel;
    P.anyIso |= P.isoEnabled[i];
  }
  P.clipEnabled = r->fs.clipBox.enabled; memcpy(P.clipLo, r->fs.clipBox.lower, 12); memcpy(P.clipHi, r->fs.clipBox.upper, 12);
  P.ertThreshold = r->ert;
  P.volActive = r->d_volActive; P.isoActive = r->d_isoActive; P.regionActive = r->d_regionActive;
  P.accum = r->d_accum; P.rgba8 = r->d_rgba8; P.partialOut = nullptr;
  // rays are clipped to the scene bounds - in sort-last mode to the box of this rank's part, outside of which it holds nothing
  for (int d = 0; d < 3; d++) { P.rayLo[d] = r->abr.boundsLo[d]; P.rayHi[d] = r->abr.boundsHi[d]; }
  bool ownsNothing = false;
  if (sortLastParts(r) && !partBoxOf(r, r->rank, P.rayLo, P.rayHi)) ownsNothing = true;
  boxScreenRect(r, P.rayLo, P.rayHi, P.sceneRect);
  if (ownsNothing) { P.sceneRect[0] = P.sceneRect[1] = 0; P.sceneRect[2] = P.sceneRect[3] = -1; }
  P.numStrips = r->numStrips; for (int k = 0; k < r->numStrips; k++) P.strip[k] = r->strips[k];
  for (int k = 0; k <= r->numStrips; k++) P.stripRow[k] = r->stripRow[k];
  P.tileMode = r->partMode == EXA_PARTITION_IMAGE_TILES; P.rank = r->rank; P.numRanks = r->numRanks;
  // streamline cones: segment k is drawn while k < currentTimestep-1 (exabrick.cu:532-545); the device
  // copy of the time step stops at numTimesteps (OptixRenderer.cpp:481-484)
  P.traces = nullptr; P.coneStart = nullptr; P.coneList = nullptr; P.coneG = 0; P.numTraces = 0; P.numTimesteps = 0; P.currentTimestep = 0;
  if (r->d_traces && r->numTraces > 0 && std::min(r->timestep, r->numTimesteps) >= 2) {
    P.traces = r->d_traces; P.numTraces = r->numTraces; P.numTimesteps = r->numTimesteps;
    P.currentTimestep = std::min(r->timestep, r->numTimesteps);
  }
  bool anyContour = false;
  for (int i = 0; i < EXA_MAX_CONTOUR_PLANES; i++) {
    P.cpEnabled[i] = r->fs.contourPlane[i].enabled; memcpy(P.cpNormal[i], r->fs.contourPlane[i].normal, 12);
    P.cpOffset[i] = r->fs.contourPlane[i].offset; P.cpChannel[i] = r->fs.contourPlane[i].channel;
    anyContour |= P.cpEnabled[i] != 0;
  }
  P.aoEnabled = r->fs.ao.enabled; P.aoLength = r->fs.ao.length;
  memcpy(P.worldLo, r->worldLo, 12); memcpy(P.worldHi, r->worldHi, 12);
  P.meshNodes = r->d_meshNodes; P.meshTris = r->d_meshTris; P.numMeshTris = r->numMeshTris;
  P.anySurf = (anyContour || P.traces || P.numMeshTris > 0 || (P.aoEnabled && P.anyIso)) ? 2 : P.anyIso ? 1 : 0;
  P.stats = nullptr;
  P.tileOrderIn = nullptr; P.tileOrderOut = nullptr; P.tileOrderCap = r->tileOrderCap;
  P.tileList = r->tileListLen > 0 ? r->d_tileList : nullptr; P.tileListLen = r->tileListLen; P.tileCost = r->d_tileCost;
  P.kdStack = r->abr.depth + 2;
  P.dbg = 0;
#ifdef EXA_DEBUG
  if (const char *e = getenv("EXA_DEBUG_PIXEL")) { int x = 0, y = 0; if (sscanf(e, "%d,%d", &x, &y) == 2) P.dbg = 1 + x + r->width * y; }
#endif
}

static void syncLaunchParams(exa_renderer *r)
{
  ExaLaunchParams &lp = r->lp;
  lp.deviceIndex = r->rank; lp.deviceCount = r->numRanks;
  lp.volumeBVH = (uint64_t)r->abr.kd; lp.surfaceModel = 0; lp.isoSurfaceBVH = (uint64_t)r->abr.kd; lp.streamlineBVH = 0;
  lp.fbSize[0] = r->width; lp.fbSize[1] = r->height;
  memcpy(lp.worldSpaceBounds_lo, r->worldLo, 12); memcpy(lp.worldSpaceBounds_hi, r->worldHi, 12);
  memcpy(lp.voxelSpaceBounds_lo, r->voxelLo, 12); memcpy(lp.voxelSpaceBounds_hi, r->voxelHi, 12);
  lp.colorBufferPtr = (uint64_t)r->d_rgba8; lp.accumBufferPtr = (uint64_t)r->d_accum; lp.frameStateBuffer = 0;
  lp.dt = r->dt;
  lp.sameBrickRegionsBuffer = (uint64_t)r->abr.regions; lp.sameBrickRegionsLeafList = (uint64_t)r->abr.leafList;
  lp.brickBuffer = (uint64_t)r->d_bricks; lp.scalarBuffers = (uint64_t)r->d_scalars; lp.scalarBufferOffsets = (uint64_t)r->d_scalarOffsets;
  lp.primaryChannels = (uint64_t)r->d_primaryChannels; lp.colormapChannel = 0; lp.numPrimaryChannels = r->numFields;
  lp.gradientShadingDVR = r->gradDVR; lp.gradientShadingISO = r->gradISO;
  lp.tracerEnabled = r->tracerEnabled; lp.traces = (uint64_t)r->d_traces;
  memcpy(lp.tracerChannels, r->tracerChannels, 12);
  lp.numTraces = r->numTraces; lp.numTimesteps = r->numTimesteps; lp.steplen = r->steplen; lp.currentTimestep = 0;
  for (int c = 0; c < EXA_MAX_CHANNELS; c++) r->fs.xfTexture[c] = c < r->numFields ? (uint64_t)(r->d_xf + 128 * c) : 0;
}

static void computeWorldBounds(exa_renderer *r)
{
  // OptixRenderer::setVoxelSpaceTransform (:330-332): world = rcp(xfm) * voxel bounds corners
  const float *m = (const float *)&r->fs.voxelSpaceTransform;
  const float *vx = m, *vy = m + 3, *vz = m + 6, *p = m + 9;
  const float cx[3] = {vy[1] * vz[2] - vy[2] * vz[1], vy[2] * vz[0] - vy[0] * vz[2], vy[0] * vz[1] - vy[1] * vz[0]};
  const float cy[3] = {vz[1] * vx[2] - vz[2] * vx[1], vz[2] * vx[0] - vz[0] * vx[2], vz[0] * vx[1] - vz[1] * vx[0]};
  const float cz[3] = {vx[1] * vy[2] - vx[2] * vy[1], vx[2] * vy[0] - vx[0] * vy[2], vx[0] * vy[1] - vx[1] * vy[0]};
  const float det = vx[0] * cx[0] + vx[1] * cx[1] + vx[2] * cx[2];
  float il[9] = {cx[0] / det, cy[0] / det, cz[0] / det, cx[1] / det, cy[1] / det, cz[1] / det, cx[2] / det, cy[2] / det, cz[2] / det};
  float ip[3];
  for (int k = 0; k < 3; k++) ip[k] = -(il[k] * p[0] + il[3 + k] * p[1] + il[6 + k] * p[2]);
  for (int k = 0; k < 3; k++) {
    r->worldLo[k] = il[k] * r->voxelLo[0] + il[3 + k] * r->voxelLo[1] + il[6 + k] * r->voxelLo[2] + ip[k];
    r->worldHi[k] = il[k] * r->voxelHi[0] + il[3 + k] * r->voxelHi[1] + il[6 + k] * r->voxelHi[2] + ip[k];
  }
}

extern "C" {

int exa_cuda_available(void)
{
  if (!cudaUsable()) return 0;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { setError("cudaGetDeviceProperties failed"); return 0; }
  if (prop.major != 10) { setError("device 0 is not compute capability 10.x (this library ships sm_100a code only)"); return 0; }
  return 1;
}

// ---- sort-last partition tree --------------------------------------------------------------------
// The partition is a prefix of the ABR kd-tree: starting from {root}, the subtree with the largest
// cost is replaced by its two children until there are numRanks subtrees (or only leaves are left).
// Rank k owns the k-th subtree in left-to-right order.  Every subtree is a convex box of disjoint
// regions, so for a pinhole camera the visibility order of the parts is the same for every pixel:
// at each inner node of the partition tree the side that holds the camera comes first.
static void nodeCosts(const std::vector<KdNode> &kd, const std::vector<ExaBrickRegion> &regions, const std::vector<int32_t> &leafList,
                      const ExaBrick *hb, bool byCells, std::vector<double> &cost)
{
  // region of a kd leaf: blob offset = BLOB_HDR_UNITS * rid + 2 * leafListBegin(rid), increasing in rid
  std::vector<uint32_t> blobOfs(regions.size());
  for (size_t i = 0; i < regions.size(); i++) blobOfs[i] = (uint32_t)(BLOB_HDR_UNITS * i + 2 * (size_t)regions[i].leafListBegin);
  cost.assign(kd.size(), 0.0);
  for (size_t g = kd.size(); g-- > 0;) {                       // children have larger indices than their parent (level order)
    const KdNode nd = kd[g];
    const uint32_t pl = nd.info & 0x3fffffffu;
    if ((nd.info >> 30) != 3) { cost[g] = cost[pl] + cost[pl + 1]; continue; }
    if (pl == KD_EMPTY) continue;
    const size_t rid = std::lower_bound(blobOfs.begin(), blobOfs.end(), pl) - blobOfs.begin();
    const ExaBrickRegion &r = regions[rid];
    const double vol = (double)(r.domain_upper[0] - r.domain_lower[0]) * (double)(r.domain_upper[1] - r.domain_lower[1]) * (double)(r.domain_upper[2] - r.domain_lower[2]);
    if (!byCells) cost[g] = vol / (double)r.finestLevelCellWidth;          // ~ lattice samples of a uniform ray bundle
    else for (int k = 0; k < r.leafListSize; k++) { const double cw = (double)(1 << hb[leafList[r.leafListBegin + k]].level); cost[g] += vol / (cw * cw * cw); }
  }
}

static void buildPartitionTree(exa_renderer *r, int numRanks, const std::vector<double> &cost)
{
  typedef exa_renderer::PartNode PN;
  r->partTree.clear();
  r->partTree.push_back(PN{0, -1, -1, 0});
  int parts = 1;
  while (parts < numRanks) {
    int best = -1; double bc = -1.0;
    for (size_t i = 0; i < r->partTree.size(); i++) {
      const PN &n = r->partTree[i];
      if (n.left >= 0 || (r->h_kd[n.kd].info >> 30) == 3) continue;     // already split / a kd leaf cannot be split
      if (cost[n.kd] > bc) { bc = cost[n.kd]; best = (int)i; }
    }
    if (best < 0) break;
    const int c = (int)(r->h_kd[r->partTree[best].kd].info & 0x3fffffffu);
    const int li = (int)r->partTree.size();
    r->partTree.push_back(PN{c, -1, -1, 0}); r->partTree.push_back(PN{c + 1, -1, -1, 0});
    r->partTree[best].left = li; r->partTree[best].right = li + 1;
    parts++;
  }
  // ranks in left-to-right order of the partition leaves
  int next = 0;
  std::vector<int> st; st.push_back(0);
  while (!st.empty()) {
    const int i = st.back(); st.pop_back();
    PN &n = r->partTree[i];
    if (n.left < 0) { n.rank = next++; continue; }
    st.push_back(n.right); st.push_back(n.left);
  }
}

static int partitionRootOf(const exa_renderer *r, int rank)
{
  for (const exa_renderer::PartNode &n : r->partTree) if (n.left < 0 && n.rank == rank) return n.kd;
  return -1;                                                   // fewer parts than ranks: this rank owns nothing
}

static int createImpl(const exa_bricks_t *bricks, const float *const *fields, const size_t *fieldCounts,
                      int numFields, int cudaDevice, int partRank, int partRanks, exa_renderer_t **out)
{
  if (!bricks || !out || !fields || !fieldCounts) { setError("exa_renderer_create: null argument"); return EXA_ERR_INVALID; }
  if (numFields < 1 || numFields > EXA_MAX_CHANNELS) { setError("exa_renderer_create: numFields must be 1.." + std::to_string(EXA_MAX_CHANNELS)); return EXA_ERR_INVALID; }
  if (partRanks < 1 || partRank < 0 || partRank >= partRanks) { setError("exa_renderer_create_partitioned: bad rank"); return EXA_ERR_INVALID; }
  const size_t nb = exa_bricks_count(bricks), ncAll = exa_bricks_num_cells(bricks);
  if (nb == 0 || ncAll == 0) { setError("exa_renderer_create: empty brick set"); return EXA_ERR_INVALID; }
  if (nb > 0x3fffffffull) { setError("too many bricks"); return EXA_ERR_LIMIT; }
  const int ndev = cudaUsable(); if (!ndev) return EXA_ERR_CUDA;
  if (cudaDevice < 0 || cudaDevice >= ndev) { setError("exa_renderer_create: bad device index"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(cudaDevice));
  if (const char *e = getenv("EXA_L2_FETCH")) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(e));   // experiment knob: 32/64/128
  exa_renderer *r = new exa_renderer;
  r->device = cudaDevice;
  memset(&r->fs, 0, sizeof(r->fs)); memset(&r->lp, 0, sizeof(r->lp));
  identityXfm((float *)&r->fs.voxelSpaceTransform);
  r->fs.xfOpacityScale = 1.f; r->fs.ao.length = 1e20f; r->fs.ao.enabled = 0;
  for (int i = 0; i < EXA_MAX_CONTOUR_PLANES; i++) { r->fs.contourPlane[i].normal[0] = 1.f; r->fs.contourPlane[i].offset = .5f; }
  for (int c = 0; c < EXA_MAX_CHANNELS; c++) { r->fs.xfDomain[c][0] = 0.f; r->fs.xfDomain[c][1] = 1.f; }
  r->numBricks = (int)nb; r->numFields = numFields;
  exa_bricks_get_bounds(bricks, r->voxelLo, r->voxelHi);
  memcpy(r->worldLo, r->voxelLo, 12); memcpy(r->worldHi, r->voxelHi, 12);
  cudaStream_t s = r->stream;
  int rc = EXA_OK;
  // temporaries of the set-up live in RAII holders: every error path below releases them
  struct Dev { void *p = nullptr; ~Dev() { if (p) cudaFree(p); } } d_ids, d_field, d_err, d_owned;
  struct Ev { cudaEvent_t e = nullptr; ~Ev() { if (e) cudaEventDestroy(e); } } ev[6];
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    setError(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " #call); exa_renderer_destroy(r); return EXA_ERR_CUDA; } } while (0)
  for (int i = 0; i < 6; i++) CKC(cudaEventCreate(&ev[i].e));
  const ExaBrick *hb = exa_bricks_headers(bricks);
  const int32_t *hids = exa_bricks_cell_ids(bricks);
  // ---- Brick[] (OptixRenderer.cpp:71-93) and the ABR geometry on the GPU (exa/Regions.cpp:242-287) ----
  CKC(cudaMalloc(&r->d_bricks, nb * sizeof(ExaBrick)));
  CKC(cudaMemcpyAsync(r->d_bricks, hb, nb * sizeof(ExaBrick), cudaMemcpyHostToDevice, s));
  CKC(cudaEventRecord(ev[0].e, s));
  rc = abrBuild(r->d_bricks, r->numBricks, s, r->abr);
  if (rc) { exa_renderer_destroy(r); return rc; }
  CKC(cudaEventRecord(ev[1].e, s));
  if (r->abr.depth + 2 > 160) { setError("ABR kd-tree deeper than the traversal stack limit"); exa_renderer_destroy(r); return EXA_ERR_LIMIT; }
  r->h_kd.resize(r->abr.numKdNodes);
  CKC(cudaMemcpy(r->h_kd.data(), r->abr.kd, sizeof(KdNode) * r->abr.numKdNodes, cudaMemcpyDeviceToHost));
  // ---- which bricks are resident: all of them, or (sort-last data partition) those of this rank's kd subtree ----
  std::vector<uint32_t> newBegin;                            // first flat cell of every resident brick in the compacted arrays
  std::vector<uint8_t> resident(nb, 1);
  size_t nc = ncAll;
  if (partRanks > 1) {
    std::vector<ExaBrickRegion> regions(r->abr.numRegions); std::vector<int32_t> leafList(r->abr.leafListLen);
    CKC(cudaMemcpy(regions.data(), r->abr.regions, sizeof(ExaBrickRegion) * regions.size(), cudaMemcpyDeviceToHost));
    CKC(cudaMemcpy(leafList.data(), r->abr.leafList, 4 * leafList.size(), cudaMemcpyDeviceToHost));
    std::vector<double> cost;
    const char *bal = getenv("EXA_PARTITION_BALANCE");
    nodeCosts(r->h_kd, regions, leafList, hb, bal && !strcmp(bal, "cells"), cost);
    buildPartitionTree(r, partRanks, cost);
    r->partMode = EXA_PARTITION_SORT_LAST; r->rank = partRank; r->numRanks = partRanks; r->dataPartitioned = true;
    r->partRoot = partitionRootOf(r, partRank);
    // owned leaves: DFS below the partition root; everything else becomes an empty leaf of THIS rank's kd-tree
    std::vector<uint8_t> ownedNode(r->h_kd.size(), 0), ownedRegion(std::max<size_t>(regions.size(), 1), 0);
    std::fill(resident.begin(), resident.end(), 0);
    std::vector<uint32_t> blobOfs(regions.size());
    for (size_t i = 0; i < regions.size(); i++) blobOfs[i] = (uint32_t)(BLOB_HDR_UNITS * i + 2 * (size_t)regions[i].leafListBegin);
    std::vector<int> st; if (r->partRoot >= 0) st.push_back(r->partRoot);
    while (!st.empty()) {
      const int g = st.back(); st.pop_back(); ownedNode[g] = 1;
      const KdNode nd = r->h_kd[g]; const uint32_t pl = nd.info & 0x3fffffffu;
      if ((nd.info >> 30) != 3) { st.push_back((int)pl); st.push_back((int)pl + 1); continue; }
      if (pl == KD_EMPTY) continue;
      const size_t rid = std::lower_bound(blobOfs.begin(), blobOfs.end(), pl) - blobOfs.begin();
      ownedRegion[rid] = 1;
      for (int k = 0; k < regions[rid].leafListSize; k++) resident[leafList[regions[rid].leafListBegin + k]] = 1;
    }
    for (size_t g = 0; g < r->h_kd.size(); g++)
      if (!ownedNode[g] && (r->h_kd[g].info >> 30) == 3) { r->h_kd[g].k = 0; r->h_kd[g].info = (3u << 30) | KD_EMPTY; }
    CKC(cudaMemcpy(r->abr.kd, r->h_kd.data(), sizeof(KdNode) * r->h_kd.size(), cudaMemcpyHostToDevice));
    CKC(cudaMalloc(&d_owned.p, ownedRegion.size()));
    CKC(cudaMemcpy(d_owned.p, ownedRegion.data(), ownedRegion.size(), cudaMemcpyHostToDevice));
    // compacted flat layout: resident bricks keep their order; the others get begin = 0 and are unreachable
    newBegin.assign(nb, 0); size_t tot = 0; int nres = 0;
    for (size_t i = 0; i < nb; i++) if (resident[i]) { newBegin[i] = (uint32_t)tot; tot += (size_t)hb[i].size[0] * hb[i].size[1] * hb[i].size[2]; nres++; }
    nc = std::max<size_t>(tot, 1); r->residentBricks = nres;
    std::vector<ExaBrick> hdr(hb, hb + nb);
    for (size_t i = 0; i < nb; i++) hdr[i].begin = newBegin[i];
    CKC(cudaMemcpy(r->d_bricks, hdr.data(), nb * sizeof(ExaBrick), cudaMemcpyHostToDevice));
  } else r->residentBricks = (int)nb;
  r->numCells = nc;
  // ---- flatten: scalar gather through the cell IDs (OptixRenderer.cpp:103-141) ----
  CKC(cudaEventRecord(ev[2].e, s));
  CKC(cudaMalloc(&r->d_scalars, sizeof(float) * nc * (size_t)numFields));
  {
    size_t fieldCap = 0;
    CKC(cudaMalloc(&d_ids.p, nc * 4)); CKC(cudaMalloc(&d_err.p, 4)); CKC(cudaMemsetAsync(d_err.p, 0, 4, s));
    if (partRanks > 1) {
      for (size_t i = 0; i < nb; i++) if (resident[i]) {
        const size_t n = (size_t)hb[i].size[0] * hb[i].size[1] * hb[i].size[2];
        CKC(cudaMemcpyAsync((int32_t *)d_ids.p + newBegin[i], hids + hb[i].begin, n * 4, cudaMemcpyHostToDevice, s));
      }
    } else CKC(cudaMemcpyAsync(d_ids.p, hids, nc * 4, cudaMemcpyHostToDevice, s));
    for (int f = 0; f < numFields; f++) {
      if (!fields[f] || fieldCounts[f] == 0) { setError("exa_renderer_create: empty scalar field"); rc = EXA_ERR_INVALID; break; }
      if (fieldCounts[f] > fieldCap) { if (d_field.p) { cudaStreamSynchronize(s); cudaFree(d_field.p); d_field.p = nullptr; } CKC(cudaMalloc(&d_field.p, fieldCounts[f] * 4)); fieldCap = fieldCounts[f]; }
      CKC(cudaMemcpyAsync(d_field.p, fields[f], fieldCounts[f] * 4, cudaMemcpyHostToDevice, s));
      if (launchGather((const float *)d_field.p, fieldCounts[f], (const int32_t *)d_ids.p, nc, r->d_scalars + (size_t)f * nc, (int *)d_err.p, s) < 0) { setError("gather launch failed"); rc = EXA_ERR_CUDA; break; }
      r->launches++;
      CKC(cudaStreamSynchronize(s));
    }
    int herr = 0;
    if (!rc) { CKC(cudaMemcpy(&herr, d_err.p, 4, cudaMemcpyDeviceToHost)); }
    cudaFree(d_ids.p); d_ids.p = nullptr; cudaFree(d_field.p); d_field.p = nullptr;
    if (!rc && herr) { setError(herr == 1 ? "negative cell ID" : "invalid cell ID"); rc = EXA_ERR_INVALID; }   // OptixRenderer.cpp:122-127
    if (rc) { exa_renderer_destroy(r); return rc; }
  }
  CKC(cudaEventRecord(ev[3].e, s));
  {
    std::vector<unsigned> offs(numFields); std::vector<int> prim(numFields);
    for (int f = 0; f < numFields; f++) { offs[f] = (unsigned)((size_t)f * nc); prim[f] = f; }   // 32-bit view valid while C*Nc < 2^32
    CKC(cudaMalloc(&r->d_scalarOffsets, 4 * numFields)); CKC(cudaMalloc(&r->d_primaryChannels, 4 * numFields));
    CKC(cudaMemcpy(r->d_scalarOffsets, offs.data(), 4 * numFields, cudaMemcpyHostToDevice));
    CKC(cudaMemcpy(r->d_primaryChannels, prim.data(), 4 * numFields, cudaMemcpyHostToDevice));
  }
  // ---- dual-cell records (8x the scalar memory, one DRAM sector per brick visit) ----
  {
    int mode = g_scalarLayout;
    if (const char *e = getenv("EXA_SCALAR_LAYOUT")) mode = atoi(e);
    std::vector<uint32_t> begin(nb, 0); uint64_t total = 0; bool fits32 = true;
    for (size_t i = 0; i < nb; i++) {
      if (total > 0xffffffffull) { fits32 = false; break; }
      begin[i] = (uint32_t)total;                              // non-resident bricks: empty span (keeps begin[] sorted for k_expandDual's search)
      if (!resident[i]) continue;
      total += (uint64_t)(hb[i].size[0] + 1) * (uint64_t)(hb[i].size[1] + 1) * (uint64_t)(hb[i].size[2] + 1);
    }
    size_t freeB = 0, totalB = 0; CKC(cudaMemGetInfo(&freeB, &totalB));
    const uint64_t need = total * 32ull * (uint64_t)numFields;
    // auto: build the records when they fit in 40% of the free HBM and 32-bit record offsets
    const bool want = mode == EXA_LAYOUT_DUAL || (mode == EXA_LAYOUT_AUTO && need <= (uint64_t)(0.4 * (double)freeB));
    if (want && total > 0 && fits32 && total <= 0xffffffffull && need + (64ull << 20) < freeB) {
      CKC(cudaMalloc(&r->d_dualBegin, nb * 4));
      CKC(cudaMemcpyAsync(r->d_dualBegin, begin.data(), nb * 4, cudaMemcpyHostToDevice, s));
      CKC(cudaMalloc(&r->d_dual, need));
      r->dualRecords = total;
      for (int f = 0; f < numFields; f++) {
        if (launchExpandDual(r->d_bricks, r->numBricks, r->d_dualBegin, total, r->d_scalars + (size_t)f * nc, r->d_dual + 2 * (size_t)f * total, s) < 0) {
          setError("dual-record expansion failed"); exa_renderer_destroy(r); return EXA_ERR_CUDA; }
        r->launches++;
      }
      CKC(cudaStreamSynchronize(s));                           // begin[] goes out of scope
    } else if (mode == EXA_LAYOUT_DUAL && total > 0) {
      setError("EXA_LAYOUT_DUAL requested but the dual-cell records do not fit (memory or 32-bit record offsets)");
      exa_renderer_destroy(r); return EXA_ERR_LIMIT;
    }
  }
  CKC(cudaEventRecord(ev[4].e, s));
  // ---- value ranges + scalar offsets of the region blobs (exa/Regions.cpp:182-240, 290-300) ----
  rc = abrAttachScalars(r->abr, r->d_bricks, r->d_scalars, r->numCells, numFields, r->d_dualBegin, (const uint8_t *)d_owned.p, s);
  if (rc) { exa_renderer_destroy(r); return rc; }
  CKC(cudaEventRecord(ev[5].e, s));
  CKC(cudaEventSynchronize(ev[5].e));
  {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev[0].e, ev[1].e); r->buildMs[0] = ms;      // ABR geometry (kd build, regions, leaf lists)
    cudaEventElapsedTime(&ms, ev[2].e, ev[3].e); r->buildMs[1] = ms;      // scalar gather incl. H2D of ids and fields
    cudaEventElapsedTime(&ms, ev[3].e, ev[4].e); r->buildMs[2] = ms;      // dual-cell records
    cudaEventElapsedTime(&ms, ev[4].e, ev[5].e); r->buildMs[3] = ms;      // value ranges
  }
  r->launches += r->abr.launches;
  {
    // point-location grid: roughly cubic cells (the domain need not be), ~32 per region, 4 K .. EXA_LOCGRID_CELLS (8 B each)
    double target = std::min(2097152.0, std::max(4096.0, 32.0 * r->abr.numRegions));
    if (const char *e = getenv("EXA_LOCGRID_CELLS")) target = std::max(4096.0, atof(e));        // experiment knob: exact cell budget
    double ext[3], vol = 1.0;
    for (int d = 0; d < 3; d++) { ext[d] = std::max(1e-20, (double)r->abr.boundsHi[d] - (double)r->abr.boundsLo[d]); vol *= ext[d]; }
    const double w = std::cbrt(vol / target);
    int G[3];
    for (int d = 0; d < 3; d++) G[d] = (int)std::min(2048.0, std::max(1.0, std::ceil(ext[d] / w)));
    if (!getenv("EXA_NO_LOCGRID")) {
      CKC(cudaMalloc(&r->d_locGrid, sizeof(uint2) * (size_t)G[0] * G[1] * G[2]));
      if (launchBuildLocGrid(r->abr.kd, G, r->abr.boundsLo, r->abr.boundsHi, r->d_locGrid, r->locScale, s) < 0) {
        setError("point-location grid kernel failed"); exa_renderer_destroy(r); return EXA_ERR_CUDA; }
      for (int d = 0; d < 3; d++) r->locG[d] = G[d];
      r->launches++;
    }
  }
  const int nn = std::max(r->abr.numKdNodes, 1), nr = std::max(r->abr.numRegions, 1);
  CKC(cudaMalloc(&r->d_volActive, 2 * (size_t)nn + nr)); r->d_isoActive = r->d_volActive + r->abr.numKdNodes; r->d_regionActive = r->d_volActive + 2 * (size_t)r->abr.numKdNodes; CKC(cudaMalloc(&r->d_inPart, nn));
  {
    // parent links for the activity pass (k_leafActiveUp climbs from the leaves)
    std::vector<int> parent(nn, -1);
    for (int g = 0; g < r->abr.numKdNodes; g++) {
      const KdNode nd = r->h_kd[g];
      if ((nd.info >> 30) == 3) continue;
      const uint32_t c = nd.info & 0x3fffffffu;
      if (c + 1 < (uint32_t)r->abr.numKdNodes) { parent[c] = g; parent[c + 1] = g; }
    }
    CKC(cudaMalloc(&r->d_kdParent, sizeof(int) * nn));
    CKC(cudaMemcpy(r->d_kdParent, parent.data(), sizeof(int) * nn, cudaMemcpyHostToDevice));
  }
  CKC(cudaMemset(r->d_inPart, 1, nn));
  CKC(cudaMalloc(&r->d_xf, sizeof(float4) * 128 * numFields));
  {
    // default XF: alpha ramp i/127 (viewer.cpp:558-566), grey colour ramp; callers normally updateXF
    std::vector<float4> lut(128 * numFields);
    for (int c = 0; c < numFields; c++) for (int i = 0; i < 128; i++) { float t = i / 127.f; lut[c * 128 + i] = make_float4(t, t, t, t); }
    CKC(cudaMemcpy(r->d_xf, lut.data(), sizeof(float4) * lut.size(), cudaMemcpyHostToDevice));
  }
  CKC(cudaMalloc(&r->d_stats, 8 * sizeof(unsigned long long)));
#undef CKC
  syncLaunchParams(r);
  *out = r;
  return EXA_OK;
}

int exa_renderer_create(const exa_bricks_t *bricks, const float *const *fields, const size_t *fieldCounts,
                        int numFields, int cudaDevice, exa_renderer_t **out)
{ return createImpl(bricks, fields, fieldCounts, numFields, cudaDevice, 0, 1, out); }

int exa_renderer_create_partitioned(const exa_bricks_t *bricks, const float *const *fields, const size_t *fieldCounts,
                                    int numFields, int cudaDevice, int rank, int numRanks, exa_renderer_t **out)
{ return createImpl(bricks, fields, fieldCounts, numFields, cudaDevice, rank, numRanks, out); }

int exa_renderer_get_build_times(const exa_renderer_t *r, double ms[4])
{
  if (!r || !ms) { setError("null argument"); return EXA_ERR_INVALID; }
  for (int i = 0; i < 4; i++) ms[i] = r->buildMs[i];
  return EXA_OK;
}

int exa_renderer_get_residency(const exa_renderer_t *r, int *residentBricks, uint64_t *residentCells)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (residentBricks) *residentBricks = r->residentBricks;
  if (residentCells) *residentCells = r->numCells;
  return EXA_OK;
}

void exa_renderer_destroy(exa_renderer_t *r)
{
  if (!r) return;
  cudaSetDevice(r->device);
  cudaFree(r->d_dual); cudaFree(r->d_dualBegin); cudaFree(r->d_locGrid);
  cudaFree(r->d_bricks); cudaFree(r->d_scalars); cudaFree(r->d_scalarOffsets); cudaFree(r->d_primaryChannels);
  abrFree(r->abr);
  cudaFree(r->d_xf); cudaFree(r->d_volActive); cudaFree(r->d_kdParent); cudaFree(r->d_inPart);
  cudaFree(r->d_accum); if (r->ownRgba8) cudaFree(r->d_rgba8);
  cudaFree(r->d_rgba8Alt); if (r->copyStream) cudaStreamDestroy(r->copyStream);
  if (r->evRendered) cudaEventDestroy(r->evRendered);
  for (int i = 0; i < 2; i++) if (r->evCopied[i]) cudaEventDestroy(r->evCopied[i]);
  for (int i = 0; i < 8; i++) if (r->evXf[i]) cudaEventDestroy(r->evXf[i]);
  if (r->h_xfRing) cudaFreeHost(r->h_xfRing);
  for (int i = 0; i < 2; i++) { if (r->h_stage[i]) cudaFreeHost(r->h_stage[i]); cudaFree(r->d_stage[i]); if (r->evStage[i]) cudaEventDestroy(r->evStage[i]); }
  cudaFree(r->d_traces); cudaFree(r->d_traceBounds); cudaFree(r->d_coneStart); cudaFree(r->d_coneCursor); cudaFree(r->d_coneList); cudaFree(r->d_meshNodes); cudaFree(r->d_meshTris); cudaFree(r->d_stats); cudaFree(r->d_tileOrder[0]); cudaFree(r->d_tileOrder[1]); cudaFree(r->d_tileList); cudaFree(r->d_tileCost);
  delete r;
}

int exa_device_malloc(size_t bytes, int cudaDevice, void **d_ptr)
{
  if (!d_ptr || bytes == 0) { setError("exa_device_malloc: bad argument"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(cudaDevice));
  CKR(cudaMalloc(d_ptr, bytes));                            // its own allocation: exportable with exa_ipc_export
  CKR(cudaMemset(*d_ptr, 0, bytes));
  return EXA_OK;
}
int exa_device_free(void *d_ptr) { if (d_ptr) CKR(cudaFree(d_ptr)); return EXA_OK; }
int exa_ipc_export(const void *d_ptr, uint8_t handle[64])
{
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t");
  if (!d_ptr || !handle) { setError("exa_ipc_export: null argument"); return EXA_ERR_INVALID; }
  cudaIpcMemHandle_t h;
  CKR(cudaIpcGetMemHandle(&h, const_cast<void *>(d_ptr)));
  memcpy(handle, &h, 64); return EXA_OK;
}
int exa_ipc_import(const uint8_t handle[64], int cudaDevice, void **d_ptr)
{
  if (!handle || !d_ptr) { setError("exa_ipc_import: null argument"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(cudaDevice));
  cudaIpcMemHandle_t h; memcpy(&h, handle, 64);
  CKR(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return EXA_OK;
}
int exa_ipc_close(void *d_ptr)
{
  if (!d_ptr) return EXA_OK;
  CKR(cudaIpcCloseMemHandle(d_ptr)); return EXA_OK;
}

int exa_set_scalar_layout(int mode)
{
  if (mode != EXA_LAYOUT_AUTO && mode != EXA_LAYOUT_FLAT && mode != EXA_LAYOUT_DUAL) { setError("exa_set_scalar_layout: bad mode"); return EXA_ERR_INVALID; }
  g_scalarLayout = mode; return EXA_OK;
}
int exa_renderer_get_scalar_layout(const exa_renderer_t *r, int *layout, uint64_t *bytes)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (layout) *layout = r->d_dual ? EXA_LAYOUT_DUAL : EXA_LAYOUT_FLAT;
  if (bytes) *bytes = r->d_dual ? r->dualRecords * 32ull * (uint64_t)r->numFields : r->numCells * 4ull * (uint64_t)r->numFields;
  return EXA_OK;
}

int exa_renderer_set_stream(exa_renderer_t *r, void *cudaStream)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  r->stream = (cudaStream_t)cudaStream; return EXA_OK;
}

int exa_renderer_get_abr(exa_renderer_t *r, ExaBrickRegion *regions, size_t *numRegions, int32_t *leafList, size_t *leafListLen)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (numRegions) *numRegions = r->abr.numRegions;
  if (leafListLen) *leafListLen = r->abr.leafListLen;
  if (regions && r->abr.numRegions) CKR(cudaMemcpy(regions, r->abr.regions, sizeof(ExaBrickRegion) * r->abr.numRegions, cudaMemcpyDeviceToHost));
  if (leafList && r->abr.leafListLen) CKR(cudaMemcpy(leafList, r->abr.leafList, 4 * (size_t)r->abr.leafListLen, cudaMemcpyDeviceToHost));
  return EXA_OK;
}

int exa_renderer_get_kd_info(exa_renderer_t *r, size_t *numNodes, int *depth)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (numNodes) *numNodes = r->abr.numKdNodes;
  if (depth) *depth = r->abr.depth;
  return EXA_OK;
}

int exa_renderer_get_launch_params(exa_renderer_t *r, ExaLaunchParams *lp, ExaFrameState *fs)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  syncLaunchParams(r);
  if (lp) *lp = r->lp;
  if (fs) *fs = r->fs;
  return EXA_OK;
}

int exa_renderer_set_voxel_space_transform(exa_renderer_t *r, const float xfm12[12])
{
  if (!r || !xfm12) { setError("null argument"); return EXA_ERR_INVALID; }
  memcpy(&r->fs.voxelSpaceTransform, xfm12, 48);
  computeWorldBounds(r);
  return EXA_OK;
}

int exa_renderer_resize_framebuffer(exa_renderer_t *r, void *fbPointer, int width, int height)
{
  if (!r || width <= 0 || height <= 0) { setError("resize_framebuffer: bad argument"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  CKR(cudaStreamSynchronize(r->stream));
  cudaFree(r->d_accum); r->d_accum = nullptr;
  if (r->copyStream) CKR(cudaStreamSynchronize(r->copyStream));
  if (r->ownRgba8) cudaFree(r->d_rgba8);
  r->d_rgba8 = nullptr; r->ownRgba8 = false;
  cudaFree(r->d_rgba8Alt); r->d_rgba8Alt = nullptr;
  const size_t n = (size_t)width * height;
  CKR(cudaMalloc(&r->d_accum, n * sizeof(float4)));
  CKR(cudaMemset(r->d_accum, 0, n * sizeof(float4)));
  if (fbPointer) r->d_rgba8 = (uint32_t *)fbPointer;
  else { CKR(cudaMalloc(&r->d_rgba8, n * 4)); CKR(cudaMemset(r->d_rgba8, 0, n * 4)); r->ownRgba8 = true; }
  r->width = width; r->height = height;
  for (int i = 0; i < 2; i++) { cudaFree(r->d_tileOrder[i]); r->d_tileOrder[i] = nullptr; }
  cudaFree(r->d_tileCost); r->d_tileCost = nullptr; r->tileCostLen = 0; r->tileListLen = 0;    // tile ids of the old size are meaningless
  r->numStrips = 0;                                           // so are strips laid out for the old height
  r->tileOrderCap = ((width + 15) / 16) * ((height + 7) / 8); r->tileOrderValid = false;
  if (!getenv("EXA_NO_TILE_ORDER"))
    for (int i = 0; i < 2; i++) CKR(cudaMalloc(&r->d_tileOrder[i], sizeof(TileOrder) + sizeof(uint32_t) * (size_t)TILE_BUCKETS * r->tileOrderCap));
  return EXA_OK;
}

int exa_renderer_set_color_buffer(exa_renderer_t *r, void *fbPointer)
{
  if (!r || !r->d_accum) { setError("set_color_buffer: resizeFrameBuffer was never called"); return EXA_ERR_INVALID; }
  if (fbPointer) { if (r->ownRgba8) { cudaFree(r->d_rgba8); r->ownRgba8 = false; } r->d_rgba8 = (uint32_t *)fbPointer; return EXA_OK; }
  if (!r->ownRgba8) {
    CKR(cudaSetDevice(r->device));
    CKR(cudaMalloc(&r->d_rgba8, (size_t)r->width * r->height * 4)); r->ownRgba8 = true;
  }
  return EXA_OK;
}

int exa_strip_rows(int k, int numStrips, int height, int *y0, int *y1)
{
  if (numStrips < 1 || k < 0 || k >= numStrips || height < 1 || !y0 || !y1) { setError("exa_strip_rows: bad argument"); return EXA_ERR_INVALID; }
  const int tileRows = (height + 7) / 8;
  *y0 = std::min(height, 8 * ((tileRows * k) / numStrips)); *y1 = std::min(height, 8 * ((tileRows * (k + 1)) / numStrips));
  return EXA_OK;
}

int exa_renderer_set_color_strips_rows(exa_renderer_t *r, void *const *stripPtrs, const int *firstRow, int numStrips)
{
  if (!r || !r->d_accum) { setError("set_color_strips: resizeFrameBuffer was never called"); return EXA_ERR_INVALID; }
  if (numStrips < 0 || numStrips > EXA_MAX_PARTS || (numStrips && (!stripPtrs || !firstRow))) { setError("set_color_strips: 0..16 strips"); return EXA_ERR_INVALID; }
  // strip k = pixel rows [firstRow[k], firstRow[k+1]): whole 8-row tile rows, ascending, covering the frame
  for (int k = 0; k < numStrips; k++) {
    const bool last = k == numStrips - 1;
    if ((firstRow[k] % 8 != 0 && firstRow[k] != r->height) || firstRow[k] > firstRow[k + 1] || (k == 0 && firstRow[0] != 0) || (last && firstRow[k + 1] != r->height) || (!last && firstRow[k + 1] > r->height)) {
      setError("set_color_strips: strip rows must start at multiples of 8, ascend from 0 and end at the frame height"); return EXA_ERR_INVALID;
    }
    if (!stripPtrs[k] && firstRow[k + 1] > firstRow[k]) { setError("set_color_strips: null strip"); return EXA_ERR_INVALID; }
  }
  for (int k = 0; k < numStrips; k++) { r->strips[k] = (uint32_t *)stripPtrs[k]; r->stripRow[k] = (firstRow[k] + 7) / 8; }
  r->stripRow[numStrips] = numStrips ? (r->height + 7) / 8 : 0;
  r->numStrips = numStrips;
  return EXA_OK;
}

int exa_renderer_set_color_strips(exa_renderer_t *r, void *const *stripPtrs, int numStrips)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (numStrips < 0 || numStrips > EXA_MAX_PARTS) { setError("set_color_strips: 0..16 strips"); return EXA_ERR_INVALID; }
  int rows[EXA_MAX_PARTS + 1] = {0};
  for (int k = 0; k < numStrips; k++) { int y0; exa_strip_rows(k, numStrips, std::max(r->height, 1), &y0, &rows[k + 1]); }
  return exa_renderer_set_color_strips_rows(r, stripPtrs, rows, numStrips);
}

int exa_renderer_update_camera(exa_renderer_t *r, const float pos[3], const float dir00[3], const float dirDu[3], const float dirDv[3])
{
  if (!r || !pos || !dir00 || !dirDu || !dirDv) { setError("null argument"); return EXA_ERR_INVALID; }
  memcpy(r->fs.camera.pos, pos, 12); memcpy(r->fs.camera.dir00, dir00, 12);
  memcpy(r->fs.camera.dirDu, dirDu, 12); memcpy(r->fs.camera.dirDv, dirDv, 12);
  return EXA_OK;
}

int exa_renderer_update_xf(exa_renderer_t *r, int channel, const float *opacities128, const float *rgb128x3,
                           float domainLower, float domainUpper, float opacityScale)
{
  if (!r || !opacities128 || !rgb128x3) { setError("null argument"); return EXA_ERR_INVALID; }
  if (channel < 0 || channel >= r->numFields) { setError("updateXF: channel out of range"); return EXA_ERR_INVALID; }
  r->fs.xfDomain[channel][0] = domainLower; r->fs.xfDomain[channel][1] = domainUpper;
  r->fs.xfOpacityScale = opacityScale;
  CKR(cudaSetDevice(r->device));
  if (!r->h_xfRing) {
    CKR(cudaMallocHost(&r->h_xfRing, sizeof(float4) * 128 * 8));
    for (int i = 0; i < 8; i++) CKR(cudaEventCreateWithFlags(&r->evXf[i], cudaEventDisableTiming));
  }
  // stream-ordered upload from a pinned ring slot: frames already enqueued keep the table they
  // were launched with, and the call does not wait for them
  const unsigned slot = r->xfRingPos++ % 8;
  CKR(cudaEventSynchronize(r->evXf[slot]));                  // only blocks when the GPU is 8 updates behind
  float4 *lut = r->h_xfRing + 128 * slot;
  for (int i = 0; i < 128; i++) lut[i] = make_float4(rgb128x3[3 * i], rgb128x3[3 * i + 1], rgb128x3[3 * i + 2], opacities128[i]);
  CKR(cudaMemcpyAsync(r->d_xf + 128 * channel, lut, sizeof(float4) * 128, cudaMemcpyHostToDevice, r->stream));
  CKR(cudaEventRecord(r->evXf[slot], r->stream));
  r->masksDirty = true;                                     // needVolumeBVHRebuild (:403)
  return EXA_OK;
}

int exa_renderer_update_iso_values(exa_renderer_t *r, const float *isoValues, const int *channels, const int *enabled)
{
  if (!r || !isoValues || !channels || !enabled) { setError("null argument"); return EXA_ERR_INVALID; }
  for (int i = 0; i < EXA_MAX_ISO_SURFACES; i++) {
    if (enabled[i] && (channels[i] < 0 || channels[i] >= r->numFields)) { setError("updateIsoValues: channel out of range"); return EXA_ERR_INVALID; }
    r->fs.isoSurface[i].value = isoValues[i]; r->fs.isoSurface[i].channel = channels[i]; r->fs.isoSurface[i].enabled = enabled[i] ? 1 : 0;
  }
  r->masksDirty = true;                                     // needIsoBVHRebuild (:501)
  return EXA_OK;
}

int exa_renderer_update_frame_id(exa_renderer_t *r, int frameID) { if (!r) { setError("null renderer"); return EXA_ERR_INVALID; } r->fs.frameID = frameID; return EXA_OK; }
int exa_renderer_update_dt(exa_renderer_t *r, float dt) { if (!r || !(dt > 0.f)) { setError("updateDt: dt must be > 0"); return EXA_ERR_INVALID; } r->dt = dt; return EXA_OK; }
int exa_renderer_set_space_skipping(exa_renderer_t *r, int enable) { if (!r) { setError("null renderer"); return EXA_ERR_INVALID; } r->spaceSkipping = enable != 0; r->masksDirty = true; return EXA_OK; }
int exa_renderer_set_gradient_shading_dvr(exa_renderer_t *r, int enable) { if (!r) { setError("null renderer"); return EXA_ERR_INVALID; } r->gradDVR = enable != 0; return EXA_OK; }
int exa_renderer_set_gradient_shading_iso(exa_renderer_t *r, int enable) { if (!r) { setError("null renderer"); return EXA_ERR_INVALID; } r->gradISO = enable != 0; return EXA_OK; }
int exa_renderer_set_gradient_mode(exa_renderer_t *r, int mode)
{
  if (!r || (mode != EXA_GRADIENTS_ANALYTIC && mode != EXA_GRADIENTS_CENTRAL_DIFFERENCES)) { setError("set_gradient_mode: bad argument"); return EXA_ERR_INVALID; }
  r->gradCD = mode == EXA_GRADIENTS_CENTRAL_DIFFERENCES; return EXA_OK;
}
int exa_renderer_set_ert_threshold(exa_renderer_t *r, float t) { if (!r) { setError("null renderer"); return EXA_ERR_INVALID; } r->ert = t; return EXA_OK; }
int exa_renderer_set_clip_box(exa_renderer_t *r, int enable, const float lo[3], const float hi[3])
{
  if (!r || (enable && (!lo || !hi))) { setError("null argument"); return EXA_ERR_INVALID; }
  r->fs.clipBox.enabled = enable ? 1 : 0;
  if (lo) memcpy(r->fs.clipBox.lower, lo, 12);
  if (hi) memcpy(r->fs.clipBox.upper, hi, 12);
  return EXA_OK;
}
int exa_renderer_update_contour_planes(exa_renderer_t *r, const float *normals3, const float *offsets, const int *channels, const int *enabled)
{
  if (!r || !normals3 || !offsets || !channels || !enabled) { setError("null argument"); return EXA_ERR_INVALID; }
  for (int i = 0; i < EXA_MAX_CONTOUR_PLANES; i++) {
    if (enabled[i] && (channels[i] < 0 || channels[i] >= r->numFields)) { setError("updateContourPlanes: channel out of range"); return EXA_ERR_INVALID; }
    // normalize(normals[i]) (OptixRenderer.cpp:511): v * (1 / sqrt(dot(v,v)))
    const float *n = normals3 + 3 * i;
    const float s = 1.f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    for (int d = 0; d < 3; d++) r->fs.contourPlane[i].normal[d] = n[d] * s;
    r->fs.contourPlane[i].offset = offsets[i]; r->fs.contourPlane[i].channel = channels[i]; r->fs.contourPlane[i].enabled = enabled[i] ? 1 : 0;
  }
  r->masksDirty = true;                                     // space skipping is off while a plane is enabled (:518-530)
  return EXA_OK;
}
// ---- triangle meshes (OptixRenderer ctor :556-608: one OWL triangle geom per mesh under an instance group) ----
namespace {
struct MeshBuild {
  const float *v; const int32_t *ix;
  std::vector<uint32_t> order;                                 // triangle ids in BVH order
  std::vector<float> cx[3];                                    // centroids
  std::vector<float4> nodes;
  void bounds(uint32_t first, uint32_t count, float lo[3], float hi[3]) const
  {
    for (int d = 0; d < 3; d++) { lo[d] = 3e38f; hi[d] = -3e38f; }
    for (uint32_t i = first; i < first + count; i++)
      for (int c = 0; c < 3; c++) { const float *p = v + 3 * (size_t)ix[3 * (size_t)order[i] + c];
        for (int d = 0; d < 3; d++) { lo[d] = std::min(lo[d], p[d]); hi[d] = std::max(hi[d], p[d]); } }
  }
  // node `ni` over order[first, first+count): median split on the centroid along the widest axis
  void build(uint32_t ni, uint32_t first, uint32_t count)
  {
    float lo[3], hi[3]; bounds(first, count, lo, hi);
    if (count <= 4) {
      nodes[2 * (size_t)ni] = make_float4(lo[0], lo[1], lo[2], __uint_as_float_host(first));
      nodes[2 * (size_t)ni + 1] = make_float4(hi[0], hi[1], hi[2], __uint_as_float_host(count));
      return;
    }
    int axis = 0; for (int d = 1; d < 3; d++) if (hi[d] - lo[d] > hi[axis] - lo[axis]) axis = d;
    const uint32_t half = count / 2;
    const std::vector<float> &c = cx[axis];
    std::nth_element(order.begin() + first, order.begin() + first + half, order.begin() + first + count,
                     [&](uint32_t a, uint32_t b) { return c[a] < c[b] || (c[a] == c[b] && a < b); });
    const uint32_t child = (uint32_t)(nodes.size() / 2);
    nodes.resize(nodes.size() + 4);
    nodes[2 * (size_t)ni] = make_float4(lo[0], lo[1], lo[2], __uint_as_float_host(child));
    nodes[2 * (size_t)ni + 1] = make_float4(hi[0], hi[1], hi[2], __uint_as_float_host(0u));
    build(child, first, half);
    build(child + 1, first + half, count - half);
  }
  static float __uint_as_float_host(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
};
}

int exa_renderer_set_triangle_meshes(exa_renderer_t *r, const float *vertices, size_t numVerts, const int32_t *indices, size_t numTris)
{
  if (!r || (numTris && (!vertices || !indices))) { setError("set_triangle_meshes: null argument"); return EXA_ERR_INVALID; }
  if (numTris > (size_t)1 << 30) { setError("set_triangle_meshes: too many triangles"); return EXA_ERR_INVALID; }
  for (size_t i = 0; i < 3 * numTris; i++)
    if (indices[i] < 0 || (size_t)indices[i] >= numVerts) { setError("broken triangle model"); return EXA_ERR_INVALID; }   // TriangleMesh.cpp:39-47
  CKR(cudaSetDevice(r->device));
  CKR(cudaStreamSynchronize(r->stream));
  cudaFree(r->d_meshNodes); cudaFree(r->d_meshTris); r->d_meshNodes = nullptr; r->d_meshTris = nullptr; r->numMeshTris = 0;
  if (!numTris) return EXA_OK;
  MeshBuild mb; mb.v = vertices; mb.ix = indices;
  mb.order.resize(numTris);
  for (int d = 0; d < 3; d++) mb.cx[d].resize(numTris);
  for (size_t t = 0; t < numTris; t++) {
    mb.order[t] = (uint32_t)t;
    for (int d = 0; d < 3; d++)
      mb.cx[d][t] = (vertices[3 * (size_t)indices[3 * t] + d] + vertices[3 * (size_t)indices[3 * t + 1] + d] + vertices[3 * (size_t)indices[3 * t + 2] + d]) * (1.f / 3.f);
  }
  mb.nodes.resize(2);
  mb.build(0, 0, (uint32_t)numTris);
  std::vector<float> tris(9 * numTris);
  for (size_t i = 0; i < numTris; i++)
    for (int c = 0; c < 3; c++) memcpy(&tris[9 * i + 3 * c], vertices + 3 * (size_t)indices[3 * (size_t)mb.order[i] + c], 12);
  CKR(cudaMalloc(&r->d_meshNodes, mb.nodes.size() * sizeof(float4)));
  CKR(cudaMalloc(&r->d_meshTris, tris.size() * sizeof(float)));
  CKR(cudaMemcpy(r->d_meshNodes, mb.nodes.data(), mb.nodes.size() * sizeof(float4), cudaMemcpyHostToDevice));
  CKR(cudaMemcpy(r->d_meshTris, tris.data(), tris.size() * sizeof(float), cudaMemcpyHostToDevice));
  r->numMeshTris = (int)numTris;
  return EXA_OK;
}
int exa_renderer_set_clock_scale(exa_renderer_t *r, float clockScale)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  r->fs.clockScale = clockScale;                                     // viewer.cpp:956 pokes frameState.clockScale
  return EXA_OK;
}
int exa_renderer_set_ao(exa_renderer_t *r, int enable, float length)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  r->fs.ao.enabled = enable ? 1 : 0; r->fs.ao.length = length;      // viewer.cpp:953-954 pokes frameState.ao
  return EXA_OK;
}
uint64_t exa_renderer_launch_count(const exa_renderer_t *r) { return r ? r->launches : 0; }

} // extern "C"

#include "capi_render.inl"
#include "capi_group.inl"

// ---- point query without any allocation -------------------------------------------------------------------------------
// exa_sample_points_host_noalloc: exa_sample_points_host (+ the counters of exa_sample_points_stats) for callers that must
// not reach cudaMalloc — CUDA-graph capture around them, tight serving loops, harnesses that forbid allocation.  No
// counterpart in the reference.  The device-side staging of the query (points
// up, values / gradients / flags down) and the kernel's counters live in ONE file-scope device array that every GPU which
// loads this library holds until exit; nothing is allocated, neither here nor lazily in the renderer.
//   layout of g_noallocScratch: [0,256) counters (cleared every call) | xyz 12 B/pt | value 4 B/pt | gradient 12 B/pt | valid 1 B/pt
// Every call shares the array and it keeps the last call's contents: the counters and the flags are cleared before the launch,
// and two calls that overlap on the GPU (two renderers / two host threads) race — callers serialise them.
#define EXA_NOALLOC_SCRATCH_BYTES (256 + (size_t)EXA_NOALLOC_MAX_POINTS * 29)
static __device__ __align__(256) unsigned char g_noallocScratch[EXA_NOALLOC_SCRATCH_BYTES];
static_assert(EXA_NOALLOC_MAX_POINTS % 256 == 0, "sections of the scratch array stay 256-byte aligned");

// In host code `g_noallocScratch` is only the host-side handle of the symbol; the device address is per device.
static unsigned char *noallocScratchOf(int device)
{
  static unsigned char *ptr[64] = {};
  static omp_lock_t lock; static bool lockInit = (omp_init_lock(&lock), true); (void)lockInit;
  if (device < 0 || device >= 64) { setError("sample_points_host_noalloc: device index out of range"); return nullptr; }
  omp_set_lock(&lock);
  unsigned char *p = ptr[device];
  if (!p) {                                                       // once per device, with that device current (the caller set it)
    void *a = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&a, g_noallocScratch);
    if (e != cudaSuccess || !a) setError(std::string("sample_points_host_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    else p = ptr[device] = (unsigned char *)a;
  }
  omp_unset_lock(&lock);
  return p;
}

extern "C" int exa_sample_points_host_noalloc(exa_renderer_t *r, const float *xyz, size_t n, int channel, float *values, uint8_t *valid, float *grad, ExaStats *stats)
{
  if (!r) { setError("null renderer"); return EXA_ERR_INVALID; }
  if (n > (size_t)EXA_NOALLOC_MAX_POINTS) {
    char msg[160]; snprintf(msg, sizeof msg, "sample_points_host_noalloc: %zu points exceed the static scratch (EXA_NOALLOC_MAX_POINTS = %d); split the query", n, (int)EXA_NOALLOC_MAX_POINTS);
    setError(msg); return EXA_ERR_INVALID;
  }
  if (n == 0) { if (stats) memset(stats, 0, sizeof(*stats)); return EXA_OK; }
  if (!xyz || !values || !valid) { setError("sample_points: null argument"); return EXA_ERR_INVALID; }
  if (channel < 0 || channel >= r->numFields) { setError("sample_points: channel out of range"); return EXA_ERR_INVALID; }
  CKR(cudaSetDevice(r->device));
  unsigned char *base = noallocScratchOf(r->device);
  if (!base) return EXA_ERR_CUDA;
  const size_t M = EXA_NOALLOC_MAX_POINTS;
  unsigned long long *d_cnt = (unsigned long long *)base;
  float *d_xyz = (float *)(base + 256), *d_val = (float *)(base + 256 + M * 12), *d_grad = (float *)(base + 256 + M * 16);
  uint8_t *d_ok = base + 256 + M * 28;
  CKR(cudaMemsetAsync(d_cnt, 0, 256, r->stream));                 // counters start at zero
  CKR(cudaMemsetAsync(d_ok, 0, n, r->stream));                    // a flag the kernel did not write reads "invalid", never the last call's
  CKR(cudaMemcpyAsync(d_xyz, xyz, n * 12, cudaMemcpyHostToDevice, r->stream));
  RenderParams RP; fillParams(r, RP);
  SampleParams P; P.scene = RP.scene; P.xyz = d_xyz; P.n = n; P.channel = channel; P.values = d_val; P.valid = d_ok; P.grad = grad ? d_grad : nullptr;
  P.stats = stats ? d_cnt : nullptr;
  const int k = launchSamplePoints(P, r->stream);
  if (k < 0) { setError(std::string("sample kernel launch failed: ") + cudaGetErrorString(cudaGetLastError())); return EXA_ERR_CUDA; }
  r->launches += k;
  unsigned long long h[4] = {0, 0, 0, 0};
  CKR(cudaMemcpyAsync(values, d_val, n * 4, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaMemcpyAsync(valid, d_ok, n, cudaMemcpyDeviceToHost, r->stream));
  if (grad) CKR(cudaMemcpyAsync(grad, d_grad, n * 12, cudaMemcpyDeviceToHost, r->stream));
  if (stats) CKR(cudaMemcpyAsync(h, d_cnt, 32, cudaMemcpyDeviceToHost, r->stream));
  CKR(cudaStreamSynchronize(r->stream));
  if (stats) { stats->samples = h[0]; stats->brickVisits = h[1]; stats->cornerLoads = h[2]; stats->segments = h[3]; }
  return EXA_OK;
}
