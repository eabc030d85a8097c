// render.cuh — parameters of the fused frame kernels (render.cu)
#pragma once
#include "exa_internal.h"
#include <cuda_runtime.h>

namespace exa {

// tile cost order of one frame: per log2(cycles) bucket a count and a list of tile ids
// (list[b * cap + i]); written by frame N, read by frame N+1 (render.cu:k_renderFrame)
static const int TILE_BUCKETS = 16;
// count[]/done of BOTH buffers are zeroed by the host when the order is invalidated; after that the
// last CTA of a frame clears the buffer it read (the next frame's output), so no per-frame memset
struct TileOrder { int count[TILE_BUCKETS]; int done; int pad[3]; uint32_t list[4]; };   // list: TILE_BUCKETS * cap entries

static const int EXA_MAX_PARTS = 16;     // ranks of one node (sort-last parts / frame strips)

struct RenderParams {
  DeviceScene scene;
  // FrameState (programs/FrameState.h) / LaunchParams (programs/LaunchParams.h) subset
  float camPos[3], dir00[3], dirDu[3], dirDv[3];
  int sceneRect[4];                    // pixel rectangle (x0, y0, x1, y1 inclusive) outside which no primary ray can hit the scene bounds (host: projected corners, grown)
  float rayLo[3], rayHi[3];            // box every primary / iso ray is clipped to before the traversal: the scene bounds, or in sort-last mode the box of this rank's part
  int gridX0, gridY0, gridW, gridH;    // gridW > 0 (sort-last partial): the grid covers only the tile rectangle that holds this part's projection; pixels outside it are not written
  int gradCD;                          // gradients by central differences (the reference built with ANALYTIC_GRADIENTS=0) instead of the analytic ones
  float clockScale;                    // > 0: the pixel's red channel becomes min(clockScale * cycles / 1e6, 1) (exabrick.cu:1703-1707, a profiling aid)
  float camVoxel[3];                   // voxelSpaceTransform applied to camPos (host, same rounding as the device's xfmPoint): origin of every primary ray in voxel space
  float xfm[12];                       // voxelSpaceTransform: vx, vy, vz, p
  int width, height, frameID;
  float dt;
  int numChannels, colormapChannel;
  int gradDVR, gradISO;
  float xfOpacityScale;
  float xfDomain[EXA_MAX_CHANNELS][2];
  const float4 *xf;                    // [numFields][128] rgba tables
  int isoEnabled[2]; float isoValue[2]; int isoChannel[2]; int anyIso;
  int clipEnabled; float clipLo[3], clipHi[3];
  float ertThreshold;
  const uint8_t *volActive;            // per kd node: subtree has an active region (space skipping x partition)
  const uint8_t *isoActive;            // per kd node: subtree has a region whose range holds an iso value
  const uint8_t *regionActive;         // per REGION: active for volume sampling (same classification as volActive's leaves); entry fast path
  float4 *accum; uint32_t *rgba8;
  // image-tile mode across GPUs: the RGBA8 frame as numStrips row strips (whole 8-row tile rows), strip k in the
  // memory of GPU k (peer-mapped); 0 = one buffer (rgba8).  Strip k holds tile rows [stripRow[k], stripRow[k+1]).
  uint32_t *strip[EXA_MAX_PARTS]; int numStrips; int stripRow[EXA_MAX_PARTS + 1];
  float4 *partialOut;                  // != null (sort-last): write the premultiplied partial, skip bg/accum/pack
  float4 *partialSurf;                 // with partialOut: nearest surface hit of this rank {shaded rgb, distance} (null: no surface pass)
  int tileMode, rank, numRanks;        // image-tile partition
  const float *traces; int numTraces, numTimesteps, currentTimestep;   // streamline cones (null / 0: none visible)
  // uniform grid over the visible cones (rebuilt after a tracer step, where the reference rebuilds an
  // OptiX BVH): cell c lists segment ids coneList[coneStart[c] .. coneStart[c+1]); id = trace * numTimesteps + k
  const int *coneStart; const int *coneList; int coneG; float coneLo[3], coneCell[3], coneInvCell[3];
  // contour planes (FrameState.h:45-52), ambient occlusion (:60-63), world bounds (LaunchParams.h:37)
  int cpEnabled[EXA_MAX_CONTOUR_PLANES]; float cpNormal[EXA_MAX_CONTOUR_PLANES][3]; float cpOffset[EXA_MAX_CONTOUR_PLANES]; int cpChannel[EXA_MAX_CONTOUR_PLANES];
  int aoEnabled; float aoLength;
  float worldLo[3], worldHi[3];
  // triangle meshes (surfaceModel, OptixRenderer.cpp:556-608): host-built BVH, 2 x float4 per node
  // {lo.xyz, a}{hi.xyz, b}: b == 0 inner node with children a, a+1; b > 0 leaf of triangles [a, a+b);
  // triangles in BVH order, 9 floats each (A, B, C), world space
  const float4 *meshNodes; const float *meshTris; int numMeshTris;
  int anySurf;                         // surface pass: 0 none, 1 iso-surfaces only, 2 + meshes / contour planes / visible streamlines / AO
  unsigned long long *stats;           // 4 counters (ExaStats) or null
  const TileOrder *tileOrderIn;        // previous frame's tile cost order (null: scan order)
  TileOrder *tileOrderOut; int tileOrderCap;   // where this frame records its own (null: off)
  // image tiles across GPUs with measured load balance: an explicit list of this rank's tiles (most expensive first;
  // replaces ownedTile() and the bucket order) and, per tile of the frame, the CTA cycles it took (input of the next deal)
  const uint32_t *tileList; int tileListLen; uint32_t *tileCost;
  int kdStack;                         // stack entries per warp
  int warpUnits;                       // set by launchRenderFrame: 16-B units of shared memory per warp
  int dbg;                             // EXA_DEBUG builds only: print segments/samples
};

// sort-last composite of one pixel span (render.cu:k_compositeFinish)
struct CompositeParams {
  const float4 *part[EXA_MAX_PARTS];   // partial images of all ranks, front-most first (whole-frame buffers, indexed by pixel)
  const float4 *surf[EXA_MAX_PARTS];   // their surface records (hasSurf)
  int rect[EXA_MAX_PARTS][4];          // pixel rectangle (x0, y0, x1, y1 inclusive) each part has written; outside it the part contributes nothing and is not read
  int n, hasSurf;
  size_t first, count;                 // pixel span [first, first + count) this rank composites
  uint32_t *out;                       // RGBA8 frame buffer (whole frame; only the span is written)
};

// stream-ordered barrier over peer memory (render.cu:k_peerBarrier); flags[k] = rank k's array of n counters
struct PeerBarrierParams { unsigned *flags[EXA_MAX_PARTS]; int rank, n; unsigned seq; long long timeoutCycles; unsigned *err; };
int launchPeerBarrier(const PeerBarrierParams &B, cudaStream_t s);

struct SampleParams {
  DeviceScene scene;
  const float *xyz; size_t n; int channel;
  float *values; uint8_t *valid; float *grad;
  unsigned long long *stats;
};

struct TraceParams {
  DeviceScene scene;
  float *traces; int numTraces, numTimesteps;
  int firstStep, lastStep;             // advance steps [firstStep, lastStep]
  int channels[3]; float steplen;
  float worldLo[3], worldHi[3];
};

// launchers (render.cu); each returns the number of kernels launched or <0 on error
int launchRenderFrame(const RenderParams &P, cudaStream_t s);
int launchSamplePoints(const SampleParams &P, cudaStream_t s);
int launchTrace(const TraceParams &P, cudaStream_t s);
int launchTraceBounds(const float *traces, int numTraces, int numTimesteps, int currentTimestep, float *bounds, cudaStream_t s);
// cone grid: fill = 0 counts the segments per cell (counts[] += 1), fill = 1 writes their ids at cursor[cell]++
int launchConeGrid(const float *traces, int numTraces, int numTimesteps, int currentTimestep, int G, const float lo[3], const float invCell[3],
                   int *countsOrCursor, int *list, int fill, cudaStream_t s);
int launchRegionActivity(const RenderParams &P, int numNodes, const int *kdParent, int spaceSkipping,
                         const uint8_t *nodeInPartition, uint8_t *volActive, uint8_t *isoActive /* = volActive + numNodes */,
                         uint8_t *regionActive /* = volActive + 2 * numNodes, numRegions bytes */, cudaStream_t s);
int launchPartitionMask(const KdNode *kd, const std::vector<int> &levelOffset, int partRoot, uint8_t *inPart, cudaStream_t s);
int launchGather(const float *d_field, size_t fieldCount, const int32_t *d_cellIDs, size_t numCells, float *d_out,
                 int *d_errFlag, cudaStream_t s);
int launchBuildLocGrid(const KdNode *kd, const int G[3], const float lo[3], const float hi[3], uint2 *d_grid, float scaleOut[3], cudaStream_t s);
int launchExpandDual(const ExaBrick *d_bricks, int numBricks, const uint32_t *d_dualBegin, uint64_t numRecords,
                     const float *d_scalars, float4 *d_dual, cudaStream_t s);
int launchCompositeFinish(const RenderParams &P, const CompositeParams &C, cudaStream_t s);

} // namespace exa
