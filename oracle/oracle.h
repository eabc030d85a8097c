/* oracle/oracle.h — TEST INFRASTRUCTURE ONLY (never linked, imported or executed by
 * the product path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it).
 *
 * CPU restatement of the ExaBricks hot path of owl-project/owlExaBrick:
 *   - exaBuilder          builder/builder.cpp:301-350, 447-530, 538-811
 *   - ABR build           exa/Regions.cpp:32-320
 *   - tent-basis sampling programs/exabrick.cu:581-594, 620-777, 781-816, 883-942
 *   - DVR / iso frame     programs/exabrick.cu:135-150, 184-238, 250-331, 373-402,
 *                         988-1256, 1408-1460, 1576-1720; programs/Camera.h:27-44
 *   - streamlines (RK4)   programs/exabrick.cu:945-963, 1531-1574; rounded cones :441-573
 *   - contour planes, AO  programs/exabrick.cu:78-94, 1267-1406, 1476-1529, 1604-1652
 *
 * Parity status: PINNED by the reference's own code, compiled in place into oracle/_ref (see
 * oracle/Makefile, tests/golden/make_golden.py):
 *   - builder, ABR build, exaRawToCells: the reference's host sources (exaBuilder, ref_abr,
 *     exaRawToCells) — bit-exact (tests/test_oracle.py, tests/test_host.py);
 *   - reconstruction, transfer function, DVR / iso / contour-plane / streamline-cone / AO frames
 *     and the RK4 tracer: the reference's DEVICE source programs/exabrick.cu compiled as plain
 *     C++ (ref_shader_main.cpp + ref_shim/shader_shim.h supply the OptiX/OWL device API on the
 *     host) — the oracle reproduces its accumulation buffers, RGBA8 frames and trace vertices
 *     bit for bit on 24 frames + 1 tracer run (tests/test_oracle_pinned.py, golden vectors in
 *     tests/golden/ref_shader_frames.npz).  Two things in that comparison are not reference code:
 *     OptiX's closest-hit search (brute force in the shim) and the CUDA texture fetch of the
 *     transfer function (a linear-filtered 128-texel table with a full-precision weight).
 *     Triangle meshes: the closest-hit program, shading and AO are the reference's; OptiX's built-in
 *     triangle test (no source) is stood in for by the reference's own intersectRayTriangle.
 */
#pragma once
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrcBrick {        /* programs/Brick.h:31-71, 32 bytes */
  int32_t  lower[3];
  int32_t  size[3];
  int32_t  level;
  uint32_t begin;
} OrcBrick;

typedef struct OrcRegion {       /* exa/Regions.h:31-41, 44 bytes */
  float   lo[3], hi[3];
  float   vlo, vhi;
  int32_t leafListBegin, leafListSize;
  float   finestLevelCellWidth;
} OrcRegion;

typedef struct OrcFrame {
  /* camera (programs/FrameState.h:30-35) */
  float cam_pos[3], cam_dir00[3], cam_du[3], cam_dv[3];
  /* voxelSpaceTransform: linear part column vectors vx,vy,vz then translation p
     (owl AffineSpace: xfmPoint = p + x*vx + y*vy + z*vz) */
  float xf_vx[3], xf_vy[3], xf_vz[3], xf_p[3];
  int32_t width, height;
  int32_t frameID;
  float   dt;
  int32_t numPrimaryChannels;
  int32_t colormapChannel;
  int32_t gradientShadingDVR, gradientShadingISO;
  int32_t spaceSkipping;
  float   xfOpacityScale;
  float   xfDomain[10][2];
  float   xf[10][128][4];          /* rgba tables, one per channel */
  int32_t isoEnabled[2]; float isoValue[2]; int32_t isoChannel[2];
  int32_t clipEnabled; float clipLo[3], clipHi[3];
  float   ertThreshold;            /* reference: 0.98 (exabrick.cu:49) */
  int32_t texFilter8bit;           /* 1: emulate 1.8 fixed-point tex1D weights */
  int32_t partial;                 /* 1: write premultiplied partial rgba, no bg/accum */
  /* streamlines drawn as rounded cones (exabrick.cu:441-573) */
  int32_t numTraces, numTimesteps, currentTimestep;
  const float *traces;             /* numTraces*numTimesteps*3 or NULL */
  /* contour planes (exabrick.cu:1267-1406), ambient occlusion (:1611-1651) and the world-space
     bounds the plane polygon is scaled to (LaunchParams.h:37) */
  int32_t cpEnabled[3]; float cpNormal[3][3]; float cpOffset[3]; int32_t cpChannel[3];
  int32_t aoEnabled; float aoLength;
  float   worldLo[3], worldHi[3];
  /* triangle meshes (world space; exabrick.cu:420-433, OptixRenderer.cpp:556-608), all meshes
     concatenated: vertices 3*numVerts floats, indices 3*numTris ints */
  int32_t numMeshVerts, numMeshTris;
  const float *meshVerts; const int32_t *meshIndex;
  /* 1: gradients by central differences, the reference built with -DANALYTIC_GRADIENTS=0 (exabrick.cu:833-880, 928-931) */
  int32_t centralDifferences;
} OrcFrame;

typedef struct OrcStats {          /* counters for the algorithmic-bytes model */
  uint64_t samples;                /* sample x channel evaluations */
  uint64_t brickVisits;            /* sum over samples of bricks looped */
  uint64_t cornerLoads;            /* sum n_k */
  uint64_t segments;               /* region segments integrated */
} OrcStats;

void orc_free(void *p);
int  orc_num_threads(void);

/* exaBuilder restatement: cells = n x {x,y,z,level}; out = bytes of a .bricks file */
int orc_build_bricks(const int32_t *cells, int64_t n, int maxLeafWidth,
                     uint8_t **out, int64_t *outLen);

/* ABR build restatement; scalars is C*numCells floats channel-major (flattened order). */
int orc_build_abr(const OrcBrick *bricks, int64_t nb, const float *scalars,
                  int64_t numCells, int numFields,
                  OrcRegion **regions, int64_t *nr, int32_t **leafList, int64_t *nl);

typedef struct OrcScene OrcScene;
OrcScene *orc_scene_create(const OrcBrick *bricks, int64_t nb, const float *scalars,
                           int64_t numCells, int numFields,
                           const OrcRegion *regions, int64_t nr,
                           const int32_t *leafList, int64_t nl);
void orc_scene_destroy(OrcScene *);

/* tent-basis reconstruction at points (xyz, voxel space). grad may be NULL. */
int orc_sample_points(const OrcScene *, const float *xyz, int64_t n, int channel,
                      float *value, uint8_t *valid, float *grad, OrcStats *stats);

/* one frame (exabrick.cu:1576-1720) for pixels [x0,x1)x[y0,y1); accum is W*H*4 floats
   (read if frameID>0, written), rgba8 is W*H uint32 */
int orc_render(const OrcScene *, const OrcFrame *, int x0, int y0, int x1, int y1,
               float *accum, uint32_t *rgba8, OrcStats *stats);

/* RK4 advection (exabrick.cu:1531-1574): traces[numTraces*numTimesteps*3], seeds in
   slot 0 of each trace; fills steps 1..numTimesteps-1 */
int orc_trace_streamlines(const OrcScene *, float *traces, int numTraces, int numTimesteps,
                          const int32_t channels[3], float steplen,
                          const float worldLo[3], const float worldHi[3]);

/* surface geometry, exposed for tests (exabrick.cu:441-501, 1267-1394, 78-94) */
int   orc_rounded_cone(const float pa[3], const float pb[3], float ra, float rb, const float ro[3],
                       const float rd[3], float *t, float n[3]);
float orc_contour_plane_t(const float ro[3], const float rd[3], const float normal[3], float offset,
                          const float wlo[3], const float whi[3], float *pts18, int *cnt);
void  orc_ao_direction(const float normal[3], float u1, float u2, float out[3]);

#ifdef __cplusplus
}
#endif
