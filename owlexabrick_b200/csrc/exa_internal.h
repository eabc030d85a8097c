// exa_internal.h — host/device structures shared by the kernels of libexabrick_b200.so
#pragma once
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <memory>
#include <utility>
#include "../../include/exabrick_b200.h"

static_assert(sizeof(ExaBrick) == 32, "programs/Brick.h layout");
static_assert(sizeof(ExaBrickRegion) == 44, "exa/Regions.h layout");
static_assert(sizeof(ExaKdTreeNode) == 12, "exa/KdTree.h layout");
static_assert(sizeof(ExaFrameState) == 408, "programs/FrameState.h layout");
static_assert(offsetof(ExaFrameState, isoSurface) == 48 && offsetof(ExaFrameState, contourPlane) == 72 &&
              offsetof(ExaFrameState, clipBox) == 144 && offsetof(ExaFrameState, ao) == 172 &&
              offsetof(ExaFrameState, clockScale) == 180 && offsetof(ExaFrameState, voxelSpaceTransform) == 184 &&
              offsetof(ExaFrameState, frameID) == 232 && offsetof(ExaFrameState, xfDomain) == 236 &&
              offsetof(ExaFrameState, xfTexture) == 320 && offsetof(ExaFrameState, xfOpacityScale) == 400,
              "programs/FrameState.h offsets");
static_assert(sizeof(ExaLaunchParams) == 240, "programs/LaunchParams.h layout");
static_assert(offsetof(ExaLaunchParams, volumeBVH) == 8 && offsetof(ExaLaunchParams, fbSize) == 40 &&
              offsetof(ExaLaunchParams, colorBufferPtr) == 96 && offsetof(ExaLaunchParams, accumBufferPtr) == 104 &&
              offsetof(ExaLaunchParams, frameStateBuffer) == 112 && offsetof(ExaLaunchParams, dt) == 120 &&
              offsetof(ExaLaunchParams, sameBrickRegionsBuffer) == 128 && offsetof(ExaLaunchParams, sameBrickRegionsLeafList) == 136 &&
              offsetof(ExaLaunchParams, brickBuffer) == 144 && offsetof(ExaLaunchParams, scalarBuffers) == 152 &&
              offsetof(ExaLaunchParams, scalarBufferOffsets) == 160 && offsetof(ExaLaunchParams, primaryChannels) == 168 &&
              offsetof(ExaLaunchParams, colormapChannel) == 176 && offsetof(ExaLaunchParams, gradientShadingDVR) == 184 &&
              offsetof(ExaLaunchParams, tracerEnabled) == 192 && offsetof(ExaLaunchParams, traces) == 200 &&
              offsetof(ExaLaunchParams, tracerChannels) == 208 && offsetof(ExaLaunchParams, numTraces) == 220 &&
              offsetof(ExaLaunchParams, steplen) == 228 && offsetof(ExaLaunchParams, currentTimestep) == 232,
              "programs/LaunchParams.h offsets");

namespace exa {

void setError(const std::string &msg);

// std::vector whose resize(n) leaves new elements uninitialised (cell-ID arrays of billions of entries are written
// right afterwards, in parallel; value-initialising them first is a serial page-fault pass over gigabytes)
template <typename T> struct DefaultInitAllocator : std::allocator<T> {
  template <typename U> struct rebind { typedef DefaultInitAllocator<U> other; };
  DefaultInitAllocator() = default;
  template <typename U> DefaultInitAllocator(const DefaultInitAllocator<U> &) {}
  template <typename U> void construct(U *p) { ::new (static_cast<void *>(p)) U; }
  template <typename U, typename... A> void construct(U *p, A &&...a) { ::new (static_cast<void *>(p)) U(std::forward<A>(a)...); }
};
typedef std::vector<int32_t, DefaultInitAllocator<int32_t>> IdVector;

// ---- traversal kd-tree over the ABRs (the tree Regions.cpp builds and throws away) -----
// inner: split plane + axis + index of the left child (right = left+1)
// leaf : axis==3, payload = offset of the region blob in 16-byte units (0x3fffffff = empty),
//        k = number of bricks in the region
struct KdNode {
  union { float split; uint32_t k; };
  uint32_t info;                        // axis<<30 | payload
};
static const uint32_t KD_EMPTY = 0x3fffffffu;

// Region blob, one per region, 32-byte aligned: a 64-B header (48 B used) followed by K brick
// headers of 32 B each, so that every brick header is ONE aligned 256-bit load
// (ld.global.nc.v8, LDG.E.ENL2.256 on sm_100a) and the region -> leafList -> Brick dependent
// chain of the reference is one contiguous read.
static const int BLOB_HDR_UNITS = 4;    // 16-B units
struct RegionBlobHdr {
  float lo[3], hi[3];
  float fcw;                            // finestLevelCellWidth
  int32_t k;                            // leafListSize
  int32_t regionID;
  float vlo, vhi;
  int32_t pad;
};
static_assert(sizeof(RegionBlobHdr) == 48, "blob header");
struct BrickHdr {                       // derived from ExaBrick: float lower, 1/cellWidth
  float lx, ly, lz, invCw;
  int32_t sx, sy, sz; uint32_t begin;
};
static_assert(sizeof(BrickHdr) == 32, "brick header");

struct DeviceScene {
  const ExaBrick *bricks; int numBricks;
  const float *scalars; uint64_t numCells; int numFields;   // channel c at scalars + c*numCells
  const ExaBrickRegion *regions; int numRegions;
  const int32_t *leafList; int leafListLen;
  const KdNode *kd; int numKdNodes; int kdDepth;
  const uint4 *blobs;                                       // region blobs, 16-B units
  const float4 *dual; uint64_t dualRecords;                 // dual-cell records (2 float4 each), channel c at dual + 2*c*dualRecords; null = flat layout
  float boundsLo[3], boundsHi[3];                           // root domain of the ABR tree
  // point-location grid: per cell the deepest kd node whose box contains the whole (slightly
  // grown) cell; a point query starts its descent there instead of at the root
  const uint2 *locGrid; int locG[3]; float locScale[3];       // locG[d] cells along axis d (cells are roughly cubic)
};

// result of the GPU ABR build (all device pointers owned by the caller afterwards)
struct AbrResult {
  ExaBrickRegion *regions = nullptr; int numRegions = 0;
  int32_t *leafList = nullptr; int leafListLen = 0;
  int32_t *entryRegion = nullptr;       // region of each leaf-list entry
  KdNode *kd = nullptr; int numKdNodes = 0; int depth = 0;
  std::vector<int> levelOffset;         // node index range of each tree level
  uint4 *blobs = nullptr; size_t blobUnits = 0;
  float boundsLo[3], boundsHi[3];
  uint64_t launches = 0;
};

// abr.cu
// Phase 1, geometry only: kd-tree, regions (domain, leaf list, finest level), blobs without scalar offsets.
int abrBuild(const ExaBrick *d_bricks, int numBricks, void *stream, AbrResult &out);
// Phase 2: attach the (resident) scalars: first-record field of the brick headers in the blobs
// (d_dualBegin != null: the brick's first dual-cell record, else its first flat cell = d_bricks[].begin),
// value ranges of the regions in d_regionOwned (null: all), blob headers.
int abrAttachScalars(AbrResult &abr, const ExaBrick *d_bricks, const float *d_scalars, uint64_t numCells, int numFields,
                     const uint32_t *d_dualBegin, const uint8_t *d_regionOwned, void *stream);
void abrFree(AbrResult &r);

} // namespace exa
