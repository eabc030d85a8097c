/* exabrick_b200.h — C ABI of libexabrick_b200.so
 *
 * Drop-in boundary for the ExaBricks rendering hot path of owl-project/owlExaBrick
 * (reference paths below are relative to the reference repo root).  The reference has no
 * FFI; its boundary is the C++ class exa::OptixRenderer (exa/OptixRenderer.h:32-188) plus
 * POD structs (programs/*.h) and file formats.  Every entry point here names the
 * reference interface it replaces.  The C++ mirror of the reference classes on top of
 * this ABI is include/exa/exa.hpp; INTEGRATION.md shows the reference-side binding.
 *
 * Conventions: every function returns EXA_OK (0) or an error code; the message of the last
 * error on the calling thread is exa_last_error().  No function falls back to the CPU:
 * device entry points fail with EXA_ERR_CUDA when no sm_100 device/driver is usable.
 * Handles are not thread-safe (same as the reference: one GLUT thread drives everything).
 */
#ifndef EXABRICK_B200_H
#define EXABRICK_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  EXA_OK = 0,
  EXA_ERR_INVALID = 1,   /* bad argument / malformed input (reference: std::runtime_error) */
  EXA_ERR_IO = 2,        /* file could not be opened / read */
  EXA_ERR_CUDA = 3,      /* CUDA runtime error or no usable device */
  EXA_ERR_LIMIT = 4      /* a size limit was exceeded */
};

#define EXA_NUM_XF_VALUES      128   /* exa/common.h:40 */
#define EXA_MAX_CHANNELS       10    /* exa/common.h:42 */
#define EXA_MAX_ISO_SURFACES   2     /* exa/common.h:43 */
#define EXA_MAX_CONTOUR_PLANES 3     /* exa/common.h:44 */

/* ---------------------------------------------------------------------------------------
 * POD contract (layouts probed from the reference with g++ 13 / x86-64)
 * ------------------------------------------------------------------------------------- */

/* programs/Brick.h:31-71 — 32 bytes */
typedef struct ExaBrick {
  int32_t  lower[3];
  int32_t  size[3];
  int32_t  level;
  uint32_t begin;      /* offset of the brick's first cell in the flattened scalar arrays */
} ExaBrick;

/* exa/Regions.h:31-41 (ExaBrickRegions::BrickRegion == SameBricksRegion) — 44 bytes */
typedef struct ExaBrickRegion {
  float   domain_lower[3], domain_upper[3];
  float   valueRange_lower, valueRange_upper;
  int32_t leafListBegin, leafListSize;
  float   finestLevelCellWidth;
} ExaBrickRegion;

/* exa/KdTree.h:25-135 — 12 bytes, as written by `exaBuilder -kd` */
typedef struct ExaKdTreeNode {
  uint32_t split_or_first_prim;
  uint32_t axis_bits;          /* axis<<30 | (first_child or num_prims) */
  int32_t  max_level;
} ExaKdTreeNode;

/* programs/FrameState.h:29-71 — 408 bytes */
typedef struct ExaFrameState {
  struct { float pos[3], dir00[3], dirDu[3], dirDv[3]; } camera;                 /*   0 */
  struct { uint8_t enabled; uint8_t _p[3]; float value; int32_t channel; }
      isoSurface[EXA_MAX_ISO_SURFACES];                                          /*  48 */
  struct { uint8_t enabled; uint8_t _p[3]; float normal[3]; int32_t channel; float offset; }
      contourPlane[EXA_MAX_CONTOUR_PLANES];                                      /*  72 */
  struct { float lower[3], upper[3]; uint8_t enabled; uint8_t _p[3]; } clipBox;  /* 144 */
  struct { float length; uint8_t enabled; uint8_t _p[3]; } ao;                   /* 172 */
  float    clockScale;                                                           /* 180 */
  struct { float vx[3], vy[3], vz[3], p[3]; } voxelSpaceTransform;               /* 184 */
  int32_t  frameID;                                                              /* 232 */
  float    xfDomain[EXA_MAX_CHANNELS][2];                                        /* 236 */
  uint32_t _pad0;                                                                /* 316 */
  uint64_t xfTexture[EXA_MAX_CHANNELS];  /* cudaTextureObject_t in the reference; here:
                                            device pointers to float4[128] tables   320 */
  float    xfOpacityScale;                                                       /* 400 */
  uint32_t _pad1;
} ExaFrameState;

/* programs/LaunchParams.h:26-80 — 240 bytes.  The four OptixTraversableHandle slots
 * become opaque 64-bit device pointers (kd-tree node arrays / activity masks). */
typedef struct ExaLaunchParams {
  int32_t  deviceIndex, deviceCount;                                             /*   0 */
  uint64_t volumeBVH, surfaceModel, isoSurfaceBVH, streamlineBVH;                /*   8 */
  uint32_t fbSize[2];                                                            /*  40 */
  float    worldSpaceBounds_lo[3], worldSpaceBounds_hi[3];                       /*  48 */
  float    voxelSpaceBounds_lo[3], voxelSpaceBounds_hi[3];                       /*  72 */
  uint64_t colorBufferPtr;                                                       /*  96 */
  uint64_t accumBufferPtr;                                                       /* 104 */
  uint64_t frameStateBuffer;                                                     /* 112 */
  float    dt; uint32_t _pad0;                                                   /* 120 */
  uint64_t sameBrickRegionsBuffer;                                               /* 128 */
  uint64_t sameBrickRegionsLeafList;                                             /* 136 */
  uint64_t brickBuffer;                                                          /* 144 */
  uint64_t scalarBuffers;                                                        /* 152 */
  uint64_t scalarBufferOffsets;                                                  /* 160 */
  uint64_t primaryChannels;                                                      /* 168 */
  int32_t  colormapChannel, numPrimaryChannels;                                  /* 176 */
  int32_t  gradientShadingDVR, gradientShadingISO;                               /* 184 */
  int32_t  tracerEnabled; uint32_t _pad1;                                        /* 192 */
  uint64_t traces;                                                               /* 200 */
  int32_t  tracerChannels[3];                                                    /* 208 */
  int32_t  numTraces, numTimesteps;                                              /* 220 */
  float    steplen;                                                              /* 228 */
  uint64_t currentTimestep;                                                      /* 232 */
} ExaLaunchParams;

/* counters behind the algorithmic-bytes model of SURVEY.md §8(d) / BASELINE.md §2 */
typedef struct ExaStats {
  uint64_t samples;       /* sample x channel reconstructions */
  uint64_t brickVisits;   /* sum over samples of bricks looped (4+32 B each) */
  uint64_t cornerLoads;   /* sum n_k: scalar loads issued (4 B each) */
  uint64_t segments;      /* region segments integrated (44 B each) */
} ExaStats;

const char *exa_last_error(void);
/* 1 if a CUDA device of compute capability 10.x is usable, else 0 (message set) */
int exa_cuda_available(void);
void exa_free(void *p);

/* ---------------------------------------------------------------------------------------
 * ExaBricks container — replaces exa::ExaBricks (exa/ExaBricks.h:23-45) and the builder
 * executable exaBuilder (builder/builder.cpp:836-915)
 * ------------------------------------------------------------------------------------- */
typedef struct exa_bricks exa_bricks_t;

enum { EXA_BUILDER_SPATIAL_MEDIAN = 0, EXA_BUILDER_SAH_ALIKE = 1 };   /* builder.cpp:31-33 */

/* exaBuilder in.cells -o out.bricks: cells = numCells x {x,y,z,level}
 * (loadExaJet builder.cpp:813-834; Bricker builder.cpp:406-811).  Output brick order and
 * cell IDs are bit-exact with the reference's serial builder.  maxLeafWidth: 127 default
 * (builder.cpp:50). */
int exa_bricks_build(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth,
                     exa_bricks_t **out);
/* The same builder on the GPU: duplicate removal, per-node bounds, per-slice statistics, the partition
 * and the brick writer run as CUDA kernels level by level over all nodes of the frontier; the split decisions are taken
 * by the same host code as exa_bricks_build from those statistics, so the output (brick order, cell IDs, kd file) is
 * bit-exact with it and with the reference.  timings (may be NULL): seconds of [0] cell upload, [1] duplicate removal,
 * [2] level loop, [3] brick writer + download, [4] total; [5] tree levels, [6] nodes visited.  Fails with
 * EXA_ERR_CUDA without a usable device (exa_bricks_build is the host builder; there is no silent fallback). */
int exa_bricks_build_gpu(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth,
                         int cudaDevice, exa_bricks_t **out, double timings[8]);
/* ExaBricks::load (exa/ExaBricks.cpp:21-55) */
int exa_bricks_load(const char *fileName, exa_bricks_t **out);
int exa_bricks_from_memory(const void *bytes, size_t numBytes, exa_bricks_t **out);
/* writer of builder.cpp:895-902 (+ optional kd file, :904-907; NULL to skip) */
int exa_bricks_save(const exa_bricks_t *b, const char *bricksFileName, const char *kdFileName);
void exa_bricks_destroy(exa_bricks_t *b);
size_t exa_bricks_count(const exa_bricks_t *b);
size_t exa_bricks_num_cells(const exa_bricks_t *b);          /* ExaBricks::totalNumCells */
/* flattened view = what OptixRenderer's ctor builds (exa/OptixRenderer.cpp:71-93) */
const ExaBrick *exa_bricks_headers(const exa_bricks_t *b);
const int32_t *exa_bricks_cell_ids(const exa_bricks_t *b);   /* indexVector */
/* ExaBricks::getBounds (exa/ExaBricks.cpp:57-63) */
int exa_bricks_get_bounds(const exa_bricks_t *b, float lower[3], float upper[3]);
size_t exa_bricks_kd_count(const exa_bricks_t *b);           /* 0 unless built here */
const ExaKdTreeNode *exa_bricks_kd_nodes(const exa_bricks_t *b);

/* exaRawToCells (tools/fromVolume/raw2cells.cpp): structured volume (x fastest) -> AMR cells +
 * scalars in the order the tool writes them; format = voxel type.  Both arrays malloc'd (exa_free). */
enum { EXA_RAW_BYTE = 0, EXA_RAW_FLOAT = 1, EXA_RAW_DOUBLE = 2 };
int exa_raw_to_cells(const void *voxels, int format, int nx, int ny, int nz, float threshold,
                     int maxLevel, int refinement, int32_t **cells, float **scalars, size_t *numCells);

/* TriangleMesh::load (exa/TriangleMesh.cpp:21-68): file = repeated {int32 numVerts; vec3f[];
 * int32 numTris; vec3i[]}.  All meshes of the file come back concatenated (indices rebased);
 * both arrays are malloc'd (exa_free).  Errors as the reference: "cannot open file",
 * "broken triangle model". */
int exa_triangle_mesh_load(const char *fileName, float **vertices, size_t *numVerts,
                           int32_t **indices, size_t *numTris);

/* ---------------------------------------------------------------------------------------
 * Loaders — replace exa::ScalarField (exa/ScalarField.cpp:22-96), exa::ColorMapper
 * (exa/ColorMapper.h:64-80) and exa::Config (exa/Config.cpp:23-180)
 * ------------------------------------------------------------------------------------- */
/* ScalarField::load. compat!=0 copies the reference quirk of sizing the vector by file
 * BYTES (ScalarField.cpp:27-34): *count = 4x the float count, zero tail, and valueRange
 * therefore includes 0.  values is malloc'd (exa_free). */
int exa_scalar_field_load(const char *fileName, int compat, float **values, size_t *count,
                          float valueRange[2]);
/* ScalarField::loadAndComputeMagnitude on three arrays already in memory */
/* ScalarField::createFromExpression (exa/ScalarField.cpp:96-230; `scalar <name> expr "<postfix>"` in a config):
 * postfix expression over `fields` per cell; tokens %k, numbers, + - * / ** == != < > <= >=, log abs sqrt, select */
int exa_scalar_field_expression(const float *const *fields, int numFields, size_t count,
                                const char *const *tokens, int numTokens, float *out, float valueRange[2]);
int exa_scalar_field_magnitude(const float *x, const float *y, const float *z, size_t count,
                               float *out, float valueRange[2]);
/* ColorMapper::operator()(t) over control points {t,r,g,b} sorted by t */
int exa_colormap_eval(const float *controlPoints4, int numPoints, float t, float rgb[3]);
/* ColorMapper(std::string) (exa/ColorMapper.cpp:92-131): lines "t,(r,g,b)" -> control points {t,r,g,b}, sorted by t when
 * they are not; malformed lines are skipped as in the reference.  controlPoints4 is malloc'd (exa_free). */
int exa_colormap_parse_string(const char *colorMapString, float **controlPoints4, int *numPoints);
/* ColorMapper(const uint8_t *buf, size_t size) (exa/ColorMapper.h:44-61): the first pixel row of a PNG image (8-bit grey,
 * RGB, palette, with or without alpha; non-interlaced) -> control points {i/(w-1), rgb/255}.  malloc'd (exa_free). */
int exa_colormap_from_png(const uint8_t *pngBytes, size_t numBytes, float **controlPoints4, int *numPoints);
/* the colour maps the viewer offers (exa/viewer.cpp:506-542): 0 "Paraview Cool Warm" (the default, :568) .. 14 "HSV",
 * 15 "Custom" (grey ramp); control points malloc'd (exa_free), name static.  include/exa/embedded_colormaps.h holds the
 * same maps as PNG arrays under the reference's symbol names for the PNG constructor. */
int exa_colormap_embedded_count(void);
int exa_colormap_embedded(int index, const char **name, float **controlPoints4, int *numPoints);

typedef struct exa_config exa_config_t;
/* Config::parseConfigFile + finalize (tokens remap_from remap_to scalar vector value_range
 * bricks triangles; 'scalar <name> expr "<postfix>"' evaluates ScalarField::createFromExpression) */
int exa_config_parse(const char *fileName, int scalarCompat, exa_config_t **out);
void exa_config_destroy(exa_config_t *c);
const exa_bricks_t *exa_config_bricks(const exa_config_t *c);
int exa_config_num_fields(const exa_config_t *c);
const float *exa_config_field(const exa_config_t *c, int i, size_t *count, float valueRange[2],
                              const char **name);
/* bricks.voxelSpaceTransform as 12 floats: vx, vy, vz, p (Config.cpp:23-44) */
int exa_config_voxel_space_transform(const exa_config_t *c, float xfm12[12]);
/* Config::getBounds — world-space bounds (Config.cpp:48-55) */
int exa_config_get_bounds(const exa_config_t *c, float lower[3], float upper[3]);
/* Config::surfaces (`triangles <file>`, Config.cpp:165-171), concatenated; owned by the config */
int exa_config_triangles(const exa_config_t *c, const float **vertices, size_t *numVerts,
                         const int32_t **indices, size_t *numTris);

/* ---------------------------------------------------------------------------------------
 * Renderer — replaces exa::OptixRenderer (exa/OptixRenderer.h:32-188)
 * ------------------------------------------------------------------------------------- */
typedef struct exa_renderer exa_renderer_t;

/* OptixRenderer::OptixRenderer(bricks, surfaces, scalarFields) (OptixRenderer.cpp:38-317):
 * flatten + scalar gather through cell IDs (:71-141, on the GPU), ABR build
 * (exa/Regions.cpp:242-320, on the GPU, bit-exact), kd-tree for traversal.
 * fields[i] has fieldCounts[i] floats indexed by cell ID.  multiFieldDvr semantics
 * (OptixRenderer.h:110): all fields are primary channels. */
int exa_renderer_create(const exa_bricks_t *bricks, const float *const *fields,
                        const size_t *fieldCounts, int numFields, int cudaDevice,
                        exa_renderer_t **out);
void exa_renderer_destroy(exa_renderer_t *r);
/* device time of the construction stages in ms: [0] ABR geometry (kd build, regions, leaf lists:
 * ExaBrickRegions::buildFrom, exa/Regions.cpp:242-287), [1] scalar gather incl. the upload of cell IDs and
 * fields (OptixRenderer.cpp:103-141), [2] dual-cell records, [3] value ranges (Regions.cpp:182-240, 290-300) */
int exa_renderer_get_build_times(const exa_renderer_t *r, double ms[4]);
/* Device layout of the scalars the sampling kernels read (new; the reference has only FLAT,
 * exa/OptixRenderer.cpp:103-132).  DUAL = one 32-byte record of the 8 corner scalars per dual
 * cell: 8x the memory, one DRAM sector per brick visit.  AUTO (default) = DUAL when the records
 * fit in 40% of the free device memory, else FLAT.  Applies to renderers created afterwards;
 * the environment variable EXA_SCALAR_LAYOUT=0|1|2 overrides it.  Results are identical. */
enum { EXA_LAYOUT_AUTO = 0, EXA_LAYOUT_FLAT = 1, EXA_LAYOUT_DUAL = 2 };
int exa_set_scalar_layout(int mode);
/* layout in use (EXA_LAYOUT_FLAT or EXA_LAYOUT_DUAL) and the bytes of scalar data the kernels read from */
int exa_renderer_get_scalar_layout(const exa_renderer_t *r, int *layout, uint64_t *bytes);
/* use an externally owned cudaStream_t for all launches (NULL = default stream) */
int exa_renderer_set_stream(exa_renderer_t *r, void *cudaStream);

/* ABR access (ExaBrickRegions::brickRegions / leafList, exa/Regions.h:62-64).
 * Pass NULL buffers to query the counts. */
int exa_renderer_get_abr(exa_renderer_t *r, ExaBrickRegion *regions, size_t *numRegions,
                         int32_t *leafList, size_t *leafListLen);
/* traversal kd-tree statistics: nodes, depth */
int exa_renderer_get_kd_info(exa_renderer_t *r, size_t *numNodes, int *depth);
/* host copies of the reference's launch contract (programs/LaunchParams.h, FrameState.h) */
int exa_renderer_get_launch_params(exa_renderer_t *r, ExaLaunchParams *lp, ExaFrameState *fs);

int exa_renderer_set_voxel_space_transform(exa_renderer_t *r, const float xfm12[12]); /* :322-339 */
/* resizeFrameBuffer(fbPointer, size) (:341-355). fbPointer: caller-owned RGBA8 buffer that
 * device code can write (device or managed memory), or NULL to let the renderer own one. */
int exa_renderer_resize_framebuffer(exa_renderer_t *r, void *fbPointer, int width, int height);
/* swap the RGBA8 target without touching the accumulation buffer (double-buffered read-back:
 * frame i+1 renders into one caller-owned buffer while frame i's is copied to the host).
 * NULL: back to a renderer-owned buffer. */
int exa_renderer_set_color_buffer(exa_renderer_t *r, void *fbPointer);
/* Multi-GPU image tiles: the RGBA8 frame as `numStrips` row strips, strip k = rows [y0, y1) of
 * exa_strip_rows (whole 8-row tile rows), stripPtrs[k] = device pointer of the strip's first pixel (local
 * or peer memory).  Every rank's frame kernel stores its tiles into the strip they fall in, so strip k
 * is complete on GPU k when all ranks have rendered and GPU k reads it back over its own PCIe link: the host
 * frame is filled by N concurrent copies.  numStrips = 0: back to the single colour buffer. */
int exa_renderer_set_color_strips(exa_renderer_t *r, void *const *stripPtrs, int numStrips);
int exa_strip_rows(int k, int numStrips, int height, int *y0, int *y1);
/* the same with caller-chosen strips: strip k = pixel rows [firstRow[k], firstRow[k+1]) (numStrips + 1 entries: multiples
 * of 8 ascending from 0, the last = the frame height; empty strips allowed).  For hosts whose GPUs do not reach host
 * memory equally fast (shared PCIe uplinks): strips sized by each GPU's measured copy bandwidth finish together. */
int exa_renderer_set_color_strips_rows(exa_renderer_t *r, void *const *stripPtrs, const int *firstRow, int numStrips);
int exa_renderer_update_camera(exa_renderer_t *r, const float pos[3], const float dir00[3],
                               const float dirDu[3], const float dirDv[3]);          /* :357-368 */
/* updateXF(chan, opacities[128], colorMap[128] rgb, domain, opacityScale) (:370-404) */
int exa_renderer_update_xf(exa_renderer_t *r, int channel, const float *opacities128,
                           const float *rgb128x3, float domainLower, float domainUpper,
                           float opacityScale);
int exa_renderer_update_iso_values(exa_renderer_t *r, const float *isoValues,
                                   const int *channels, const int *enabled);         /* :489-502 */
int exa_renderer_update_frame_id(exa_renderer_t *r, int frameID);                    /* :406-411 */
int exa_renderer_update_dt(exa_renderer_t *r, float dt);                             /* :413-416 */
int exa_renderer_set_space_skipping(exa_renderer_t *r, int enable);                  /* :418-432 */
int exa_renderer_set_gradient_shading_dvr(exa_renderer_t *r, int enable);            /* :434-437 */
int exa_renderer_set_gradient_shading_iso(exa_renderer_t *r, int enable);            /* :439-442 */
/* How the gradients of the two shading options above are computed.  ANALYTIC (default) is the reference's default build
 * (-DANALYTIC_GRADIENTS=1: quotient rule over the tent-basis sums, exabrick.cu:913-927); CENTRAL_DIFFERENCES is its
 * off-default build (-DANALYTIC_GRADIENTS=0: gradientCD, :833-880, six more point samples at +-(finestLevelCellWidth+1)/2,
 * each located among the ACTIVE regions).  A run-time switch here, a compile-time one in the reference.  Not available
 * together with the instrumented (stats) frame. */
enum { EXA_GRADIENTS_ANALYTIC = 0, EXA_GRADIENTS_CENTRAL_DIFFERENCES = 1 };
int exa_renderer_set_gradient_mode(exa_renderer_t *r, int mode);
int exa_renderer_set_clip_box(exa_renderer_t *r, int enable, const float lo[3], const float hi[3]);
/* updateContourPlanes(normals, offsets, channels, enabled) (OptixRenderer.cpp:504-531): arrays of
 * EXA_MAX_CONTOUR_PLANES; normals are normalised; offset is in the unit cube that is scaled to
 * the world-space bounds (exabrick.cu:1346-1364).  Forces space skipping off while a plane is on. */
int exa_renderer_update_contour_planes(exa_renderer_t *r, const float *normals3, const float *offsets,
                                       const int *channels, const int *enabled);
/* frameState.clockScale (programs/FrameState.h:57, set by the viewer at viewer.cpp:956): > 0 turns the red channel of
 * every pixel into min(clockScale * cycles / 1e6, 1), the SM cycles the pixel's thread ran (exabrick.cu:1703-1707; here
 * counted from the start of its 16x8-pixel CTA).  A profiling aid: values depend on the hardware, not on the reference. */
int exa_renderer_set_clock_scale(exa_renderer_t *r, float clockScale);
/* frameState.ao.{enabled,length} (programs/FrameState.h:60-63, poked by viewer.cpp:953-954): two
 * cosine-distributed occlusion rays of that length per shaded surface hit (exabrick.cu:1611-1646) */
int exa_renderer_set_ao(exa_renderer_t *r, int enable, float length);
/* the `surfaces` argument of the OptixRenderer constructor (OptixRenderer.cpp:556-608): world-space
 * triangle meshes, concatenated.  Builds a BVH on the host (set-up, like the reference's
 * owlGroupBuildAccel) and uploads it; numTris = 0 removes the meshes.  Hits shade as
 * ambient .2 + .8 |d.N| (closest hit SurfaceBVH, exabrick.cu:420-433). */
int exa_renderer_set_triangle_meshes(exa_renderer_t *r, const float *vertices, size_t numVerts,
                                     const int32_t *indices, size_t numTris);
/* early-ray-termination threshold (reference constant 0.98, exabrick.cu:49); >1 disables */
int exa_renderer_set_ert_threshold(exa_renderer_t *r, float threshold);
/* tracer (OptixRenderer.h:155-165, .cpp:444-487) */
int exa_renderer_set_tracer_enabled(exa_renderer_t *r, int enable);
int exa_renderer_set_tracer_params(exa_renderer_t *r, int numTraces, int numTimesteps,
                                   float steplen, const int channels[3],
                                   const float seedLo[3], const float seedHi[3]);
int exa_renderer_reset_tracer(exa_renderer_t *r);
/* replace the seeds generated by reset_tracer (libstdc++'s default_random_engine is what
 * the reference uses; this keeps tests independent of it) */
int exa_renderer_set_tracer_seeds(exa_renderer_t *r, const float *xyz, int numTraces);
int exa_renderer_advance_tracer(exa_renderer_t *r, int *rebuilt);
/* advance all remaining timesteps in one launch (the per-frame stepping is a viewer
 * animation artefact) */
int exa_renderer_trace_all(exa_renderer_t *r);
int exa_renderer_read_traces(exa_renderer_t *r, float *xyz /* numTraces*numTimesteps*3 */);
/* load trace vertices computed elsewhere (a saved tracer state): traces.buffer + timestepHost of
 * OptixRenderer.h:155-165; segments k < currentTimestep-1 are drawn */
int exa_renderer_set_traces(exa_renderer_t *r, const float *xyz, int numTraces, int numTimesteps, int currentTimestep);

/* render() (:531-552): refresh activity masks if dirty, launch the fused frame kernel,
 * synchronise */
int exa_renderer_render(exa_renderer_t *r);
int exa_renderer_render_async(exa_renderer_t *r);      /* no synchronise */
int exa_renderer_sync(exa_renderer_t *r);
/* instrumentation of the last exa_renderer_render_stats frame: out[0] = most kd nodes touched by
 * one ray, out[1] = most samples taken by one ray (frame kernel), out[2] = kd nodes touched in total */
int exa_renderer_debug_counters(exa_renderer_t *r, uint64_t out[4]);
/* after exa_sample_points_stats: out[0] = kd steps below the location grid summed over the slowest lane of every warp (what
 * the warps pay), out[1] = most steps by one point, out[2] = steps summed over all points */
/* histogram of the last frame's per-tile (16x8 pixels) cost: counts[b] = tiles whose CTA ran for
 * [2^(b+10), 2^(b+11)) SM cycles (b = 15: or more).  The frame kernel schedules the next frame's
 * tiles in descending order of these buckets. */
int exa_renderer_get_tile_cost_histogram(exa_renderer_t *r, int counts[16]);
/* headless read-back (replaces the GL blit of glutViewer/OWLViewer.cpp:140-197) */
int exa_renderer_read_rgba8(exa_renderer_t *r, uint32_t *hostOut);
/* pipelined read-back: enqueue the copy of the frame just rendered to PINNED host memory on the
 * renderer's copy stream and flip to a second frame buffer, so that the next render overlaps
 * the copy; exa_renderer_sync() waits for both.  Renderer-owned frame buffers only. */
int exa_renderer_read_rgba8_async(exa_renderer_t *r, uint32_t *pinnedHostOut);
int exa_renderer_read_accum(exa_renderer_t *r, float *hostOutRGBA);
/* restore the accumulation buffer of a progressive sequence (the next frame with frameID > 0 adds to it) */
int exa_renderer_write_accum(exa_renderer_t *r, const float *hostInRGBA);
/* device pointers of the current frame buffers (float4 accum, uint32 rgba8) */
int exa_renderer_device_buffers(exa_renderer_t *r, void **accum, void **rgba8);

/* tent-basis reconstruction at arbitrary points (samplePoint / samplePointWithDerivative,
 * programs/exabrick.cu:781-816, 883-942).  xyz in voxel space.  grad may be NULL.
 * _host copies in/out; _device takes device pointers and only enqueues on the stream. */
int exa_sample_points_host(exa_renderer_t *r, const float *xyz, size_t n, int channel,
                           float *values, uint8_t *valid, float *grad);
int exa_sample_points_device(exa_renderer_t *r, const float *d_xyz, size_t n, int channel,
                             float *d_values, uint8_t *d_valid, float *d_grad);

/* instrumented (untimed) variants: same kernels with counters on */
int exa_renderer_render_stats(exa_renderer_t *r, ExaStats *stats);
int exa_sample_points_stats(exa_renderer_t *r, const float *d_xyz, size_t n, int channel,
                            ExaStats *stats);
/* number of kernels this library has launched on this handle so far */
uint64_t exa_renderer_launch_count(const exa_renderer_t *r);

/* exa_sample_points_host for callers that must not reach cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid allocation).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels, same arguments and results as exa_sample_points_host (host arrays in and out; grad may be
 * NULL); `stats` may be NULL, else it receives the counters of exa_sample_points_stats for this query.  The device-side
 * staging and the counters live in one static device array of 256 + 29 * EXA_NOALLOC_MAX_POINTS bytes (1.8 MiB) that
 * every GPU holds from library load to exit; the call allocates nothing (no cudaMalloc / cudaFree / cudaMallocHost, now
 * or lazily).  n > EXA_NOALLOC_MAX_POINTS is refused with EXA_ERR_INVALID and a message; split larger queries.  All
 * calls, of all renderers on a device, share the array: calls that overlap on the GPU race, so serialise them (the
 * call itself returns only after its results are in the caller's arrays). */
#define EXA_NOALLOC_MAX_POINTS 65536
int exa_sample_points_host_noalloc(exa_renderer_t *r, const float *xyz, size_t n, int channel,
                                   float *values, uint8_t *valid, float *grad, ExaStats *stats);

/* ---------------------------------------------------------------------------------------
 * Multi-GPU (new; the reference only replicates)
 * ------------------------------------------------------------------------------------- */
enum { EXA_PARTITION_NONE = 0, EXA_PARTITION_SORT_LAST = 1, EXA_PARTITION_IMAGE_TILES = 2 };
/* Restrict this renderer to part `rank` of `numRanks`.
 * IMAGE_TILES: the scene is replicated, the rank renders its share of the 16x8-pixel tiles.
 * SORT_LAST on a renderer that holds the whole scene: the rank's kd subtree as a visibility mask
 * (same partition tree as exa_renderer_create_partitioned; used by tests and as the replicated variant). */
int exa_renderer_set_partition(exa_renderer_t *r, int mode, int rank, int numRanks);
/* Sort-last with the DATA partitioned (north_star's primary multi-GPU mode; the reference has none:
 * programs/LaunchParams.h:31-32 deviceIndex/deviceCount are never read).  The ABR kd-tree is built from
 * the brick headers (all ranks build the same one), cut into numRanks subtrees balanced by estimated
 * lattice samples (volume / finest cell width; EXA_PARTITION_BALANCE=cells: by cells), and only the
 * bricks listed by the regions of subtree `rank` are gathered, expanded and kept in device memory; the
 * rest of this rank's kd-tree is pruned to empty leaves.  Regions inside the part see every brick that
 * overlaps them (bricks straddling a cut are resident on both sides), so samples equal the single-GPU
 * renderer's.  exa_renderer_get_abr still returns all regions; value ranges exist for the owned ones. */
int exa_renderer_create_partitioned(const exa_bricks_t *bricks, const float *const *fields,
                                    const size_t *fieldCounts, int numFields, int cudaDevice,
                                    int rank, int numRanks, exa_renderer_t **out);
/* bricks / cells resident on this renderer's device (all of them unless created partitioned) */
int exa_renderer_get_residency(const exa_renderer_t *r, int *residentBricks, uint64_t *residentCells);
/* visibility order of the sort-last parts for the current camera: order[0] is the front-most rank.
 * All primary rays start at the camera position, so the order holds for every pixel. */
int exa_renderer_partition_order(exa_renderer_t *r, int *order /* numRanks */, int *numParts);
/* sort-last: render this rank's sub-volume to a premultiplied float4 partial image (device memory,
 * W*H*4 floats; a ray that met the early-termination threshold carries alpha + 2) and, when a surface
 * is enabled (iso / plane / mesh / cones), its nearest surface hit {shaded rgb, distance} per pixel
 * into d_partialSurf (W*H*4 floats; may be NULL when no surface is enabled). */
int exa_renderer_render_partial(exa_renderer_t *r, float *d_partialRGBA, float *d_partialSurf);
/* Fused composite + frame tail for the pixel span [first, first+count): reads the partials of ALL parts
 * front-most first (pointers may be peer memory of other GPUs - the exchange then happens inside this
 * kernel over NVLink - or local copies after an all-to-all), composites, picks the nearest surface,
 * then background blend, accumulation, sRGB, RGBA8 pack (exabrick.cu:1701-1719) into d_rgba8Out
 * (a whole-frame buffer; only the span is written). */
int exa_renderer_composite_partials(exa_renderer_t *r, const float *const *d_partials,
                                    const float *const *d_surfs /* or NULL */, int numParts,
                                    size_t first, size_t count, uint32_t *d_rgba8Out);

/* A device allocation of its own (cudaMalloc, zero-filled): unlike a sub-allocation of a caching
 * allocator it can be exported with exa_ipc_export.  Multi-GPU image-tile callers allocate the
 * shared frame buffers with it (double-buffered swap chain, bench.py). */
int exa_device_malloc(size_t bytes, int cudaDevice, void **d_ptr);
int exa_device_free(void *d_ptr);
/* Peer framebuffer (one process per GPU, one node): rank 0 exports its RGBA8 buffer, the other
 * ranks import it and pass the pointer to exa_renderer_resize_framebuffer, so that in
 * image-tile mode every rank's frame kernel stores its tiles straight into rank 0's frame
 * buffer over NVLink (no gather).  handle = the 64 bytes of a cudaIpcMemHandle_t. */
int exa_ipc_export(const void *d_ptr, uint8_t handle[64]);
int exa_ipc_import(const uint8_t handle[64], int cudaDevice, void **d_ptr);
int exa_ipc_close(void *d_ptr);

/* ---------------------------------------------------------------------------------------
 * Renderer group: N GPUs of one node driven by one host thread (the reference only replicates
 * its launch on every device, programs/LaunchParams.h:31-32 are never read).  One renderer per device;
 * mode EXA_PARTITION_IMAGE_TILES replicates the scene and splits the image, EXA_PARTITION_SORT_LAST partitions
 * the DATA along the ABR kd-tree (exa_renderer_create_partitioned) and composites the partial images through
 * NVLink peer memory.  exa_group_render renders one frame on all devices and returns when the whole RGBA8 frame
 * (scan-line layout) is in hostRgba8; every device copies its own strip / span over its own PCIe link.
 * Per-frame setters are forwarded to all devices; the others are called per device on exa_group_renderer(g, i).
 * A device may be listed more than once (several parts on one GPU: tests).
 * ------------------------------------------------------------------------------------- */
typedef struct exa_group exa_group_t;
int exa_group_create(const exa_bricks_t *bricks, const float *const *fields, const size_t *fieldCounts,
                     int numFields, const int *cudaDevices, int numDevices, int mode, exa_group_t **out);
void exa_group_destroy(exa_group_t *g);
int exa_group_size(const exa_group_t *g);
exa_renderer_t *exa_group_renderer(exa_group_t *g, int i);
int exa_group_resize_framebuffer(exa_group_t *g, int width, int height);
int exa_group_update_camera(exa_group_t *g, const float pos[3], const float dir00[3], const float dirDu[3], const float dirDv[3]);
int exa_group_update_xf(exa_group_t *g, int channel, const float *opacities128, const float *rgb128x3,
                        float domainLower, float domainUpper, float opacityScale);
int exa_group_update_frame_id(exa_group_t *g, int frameID);
int exa_group_set_voxel_space_transform(exa_group_t *g, const float xfm12[12]);
int exa_group_render(exa_group_t *g, uint32_t *hostRgba8);
/* image-tile mode: deal the tiles by measured cost (per-tile cycle counters of `rounds` frames, longest-processing-time-
 * first across the devices; exa_balance_tiles) instead of the static interleaved deal.  Call it after the camera or the
 * transfer function changed a lot - a viewer every few dozen frames.  maxOverMeanLoad may be NULL.  No-op for sort-last. */
int exa_group_rebalance(exa_group_t *g, int rounds, double *maxOverMeanLoad);

/* Image tiles balanced by measured cost (multi-GPU, data replicated).  With tile costs enabled the frame kernel
 * records, per 16x8-pixel tile it rendered, the SM cycles its CTA took.  The ranks combine their arrays (each tile
 * is rendered by one rank; max or sum), exa_balance_tiles deals the tiles longest-processing-time-first
 * (deterministic: every rank computes the same deal) and exa_renderer_set_tile_list installs this rank's tiles,
 * most expensive first, in place of the static round-robin deal.  numTiles = ceil(W/16) * ceil(H/8); a list of
 * length 0 restores the static deal.  The union of the ranks' tiles is the frame either way. */
int exa_renderer_enable_tile_costs(exa_renderer_t *r, int enable);
int exa_renderer_get_tile_costs(exa_renderer_t *r, uint32_t *costs, int numTiles);
int exa_renderer_set_tile_list(exa_renderer_t *r, const uint32_t *tiles, int numTiles);
int exa_balance_tiles(const uint32_t *costs, int numTiles, int numRanks, int rank, uint32_t *outTiles /* numTiles */, int *outCount);
/* Stream-ordered barrier between the ranks of one NVSwitch box over peer memory (one process per GPU).
 * flagPtrs[k] = rank k's flag array (exa_device_malloc of >= 17 uint32, zero-filled, IPC-mapped into every
 * rank).  Enqueues a one-warp kernel on `cudaStream` that publishes `seq` (increasing, first call 1) to every
 * rank and waits until every rank has published it: work enqueued after it on this stream sees what every
 * rank's stream did before its own call (frame strips, partial images).  Replaces a 4-byte NCCL all-reduce
 * (~64 us at 8 ranks) with ~2 us.  A rank that never arrives makes the wait time out after ~2 s and sets
 * word 16 of the local array to 1 (check it after synchronising). */
int exa_peer_barrier(void *const *flagPtrs, int rank, int numRanks, unsigned seq, void *cudaStream);
/* cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault) on a caller's stream: strip / span read-back into a
 * host frame shared by the ranks (each rank copies its part over its own PCIe link) */
int exa_copy_async(void *dst, const void *src, size_t bytes, void *cudaStream);
/* page-lock host memory the caller owns (e.g. a POSIX shared-memory frame mapped by all ranks) */
int exa_host_register(void *ptr, size_t bytes);
int exa_host_unregister(void *ptr);

#ifdef __cplusplus
}
#endif
#endif
