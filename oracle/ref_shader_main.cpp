// oracle/ref_shader_main.cpp — TEST INFRASTRUCTURE, not product code.
//
// Runs the reference's OWN device source, programs/exabrick.cu (raygen program, intersection and
// bounds programs, tent-basis reconstruction, transfer function, integration, iso-surfaces, contour
// planes, streamline cones, AO), compiled as plain C++ IN PLACE from /root/reference through
// oracle/ref_shim/shader_shim.h, on the CPU: one call of __raygen__renderFrame per pixel.  The
// scene is prepared by the reference's own host code (ExaBricks::load, ExaBrickRegions::buildFrom)
// with the flatten of exa/OptixRenderer.cpp:71-141 as in ref_abr_main.cpp.  What is NOT reference
// code: the shim (OptiX/OWL device API on the host: closest hit by brute force, tex1D as a 128-texel
// linear-filtered table) and this driver.  Triangle meshes use OptiX's built-in triangle test and
// are not run here; the tracer (computeTraces) is off, traces are an input.
//
// usage: ref_shader in.bricks frame.bin out.bin scalars0 [scalars1 ...]
//   frame.bin = OrcFrame (oracle/oracle.h) bytes, then numTraces*numTimesteps*3 floats
//   out.bin   = float accum[W*H*4], uint32 rgba8[W*H]
#include "shader_shim.h"
#include "programs/exabrick.cu"
#include "exa/ScalarField.h"
#include "oracle.h"
#include <cstdio>
#include <cstring>

thread_local shim::Ctx shim::ctx;
shim::Geom shim::geoms[8];
namespace exa { extern "C" { LaunchParams optixLaunchParams; } }

using namespace exa;

// stand-in for OptiX's built-in triangle intersection (not in the reference's source): the reference's own
// intersectRayTriangle (exabrick.cu:1316-1343) on triangle primID; hits strictly after tmin
static void intersect_Triangles()
{
  const SurfaceGeomData &self = owl::getProgramData<SurfaceGeomData>();
  const vec3i idx = self.indexBuffer[optixGetPrimitiveIndex()];
  const vec3f A = self.vertexBuffer[idx.x], B = self.vertexBuffer[idx.y], C = self.vertexBuffer[idx.z];
  owl::Ray ray(vec3f(optixGetWorldRayOrigin()), vec3f(optixGetWorldRayDirection()), optixGetRayTmin(), optixGetRayTmax());
  const float t = intersectRayTriangle(ray, A, B - A, C - A);
  if (t > optixGetRayTmin()) optixReportIntersection(t, 0);
}

static std::vector<float> readFloats(const char *fn)
{
  FILE *f = fopen(fn, "rb");
  if (!f) { fprintf(stderr, "cannot open %s\n", fn); exit(1); }
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  std::vector<float> v(n / 4);
  if (fread(v.data(), 4, v.size(), f) != v.size()) exit(1);
  fclose(f);
  return v;
}

int main(int ac, char **av)
{
  if (ac < 5) { fprintf(stderr, "usage: ref_shader in.bricks frame.bin out.bin scalars...\n"); return 1; }
  // ---- scene: the reference's loader + flatten + ABR build -------------------------------
  ExaBricks::SP input = ExaBricks::load(av[1]);
  std::vector<exa::Brick> bricks(input->bricks.size());
  std::vector<int> index(input->totalNumCells);
  size_t off = 0;
  for (size_t i = 0; i < bricks.size(); i++) {
    bricks[i].lower = input->bricks[i]->lower;
    bricks[i].size  = input->bricks[i]->size;
    bricks[i].level = input->bricks[i]->level;
    bricks[i].begin = (uint32_t)off;
    std::copy(input->bricks[i]->cellIDs.begin(), input->bricks[i]->cellIDs.end(), index.begin() + off);
    off += input->bricks[i]->cellIDs.size();
  }
  const int nf = ac - 4;
  std::vector<float> scalars(index.size() * nf);
  std::vector<unsigned> offsets(nf);
  for (int f = 0; f < nf; f++) {
    std::vector<float> v = readFloats(av[4 + f]);
    offsets[f] = (unsigned)(f * index.size());
    for (size_t i = 0; i < index.size(); i++) scalars[offsets[f] + i] = v.at(index[i]);
  }
  ExaBrickRegions regions;
  regions.buildFrom(input, bricks.data(), bricks.size(), scalars.data(), offsets.data(), nf);

  // ---- frame description ------------------------------------------------------------------
  OrcFrame of;
  std::vector<float> traces;
  std::vector<vec3f> meshVerts; std::vector<vec3i> meshIndex;
  {
    FILE *f = fopen(av[2], "rb");
    if (!f || fread(&of, sizeof(of), 1, f) != 1) { fprintf(stderr, "cannot read %s\n", av[2]); return 1; }
    if (of.numTraces > 0) {
      traces.resize((size_t)of.numTraces * of.numTimesteps * 3);
      if (fread(traces.data(), 4, traces.size(), f) != traces.size()) { fprintf(stderr, "short traces\n"); return 1; }
    }
    if (of.numMeshVerts > 0 && of.numMeshTris > 0) {         // then the triangle meshes: vec3f verts[], vec3i index[]
      meshVerts.resize(of.numMeshVerts); meshIndex.resize(of.numMeshTris);
      if (fread(meshVerts.data(), 12, meshVerts.size(), f) != meshVerts.size() || fread(meshIndex.data(), 12, meshIndex.size(), f) != meshIndex.size()) { fprintf(stderr, "short mesh\n"); return 1; }
    }
    fclose(f);
  }
  const int W = of.width, H = of.height;
  static FrameState fs;
  fs.camera.pos = vec3f(of.cam_pos[0], of.cam_pos[1], of.cam_pos[2]);
  fs.camera.dir00 = vec3f(of.cam_dir00[0], of.cam_dir00[1], of.cam_dir00[2]);
  fs.camera.dirDu = vec3f(of.cam_du[0], of.cam_du[1], of.cam_du[2]);
  fs.camera.dirDv = vec3f(of.cam_dv[0], of.cam_dv[1], of.cam_dv[2]);
  for (int i = 0; i < MAX_ISO_SURFACES; i++) { fs.isoSurface[i].enabled = of.isoEnabled[i] != 0; fs.isoSurface[i].value = of.isoValue[i]; fs.isoSurface[i].channel = of.isoChannel[i]; }
  for (int i = 0; i < MAX_CONTOUR_PLANES; i++) {
    fs.contourPlane[i].enabled = of.cpEnabled[i] != 0; fs.contourPlane[i].normal = vec3f(of.cpNormal[i][0], of.cpNormal[i][1], of.cpNormal[i][2]);
    fs.contourPlane[i].channel = of.cpChannel[i]; fs.contourPlane[i].offset = of.cpOffset[i];
  }
  fs.clipBox.enabled = of.clipEnabled != 0;
  fs.clipBox.coords = box3f(vec3f(of.clipLo[0], of.clipLo[1], of.clipLo[2]), vec3f(of.clipHi[0], of.clipHi[1], of.clipHi[2]));
  fs.ao.enabled = of.aoEnabled != 0; fs.ao.length = of.aoLength;
  fs.clockScale = 0.f;
  fs.voxelSpaceTransform.l.vx = vec3f(of.xf_vx[0], of.xf_vx[1], of.xf_vx[2]);
  fs.voxelSpaceTransform.l.vy = vec3f(of.xf_vy[0], of.xf_vy[1], of.xf_vy[2]);
  fs.voxelSpaceTransform.l.vz = vec3f(of.xf_vz[0], of.xf_vz[1], of.xf_vz[2]);
  fs.voxelSpaceTransform.p = vec3f(of.xf_p[0], of.xf_p[1], of.xf_p[2]);
  fs.frameID = of.frameID;
  static float4 xfTables[MAX_CHANNELS][128];
  for (int c = 0; c < nf; c++) {
    fs.xfDomain[c].lower = of.xfDomain[c][0]; fs.xfDomain[c].upper = of.xfDomain[c][1];
    for (int i = 0; i < 128; i++) xfTables[c][i] = make_float4(of.xf[c][i][0], of.xf[c][i][1], of.xf[c][i][2], of.xf[c][i][3]);
    fs.xfTexture[c] = (cudaTextureObject_t)(uintptr_t)&xfTables[c][0];
  }
  fs.xfOpacityScale = of.xfOpacityScale;

  std::vector<vec4f> accum((size_t)W * H, vec4f(0.f));
  std::vector<uint32_t> rgba((size_t)W * H, 0u);
  {                                                          // progressive frames read the previous sum
    FILE *f = fopen(av[3], "rb");
    if (f && of.frameID > 0) { if (fread(accum.data(), 16, accum.size(), f) != accum.size()) {} }
    if (f) fclose(f);
  }
  std::vector<int> primary(nf); for (int c = 0; c < nf; c++) primary[c] = c;
  static int currentTimestep; currentTimestep = of.currentTimestep;

  LaunchParams &lp = optixLaunchParams;
  memset(&lp, 0, sizeof(lp));
  lp.deviceIndex = 0; lp.deviceCount = 1;
  lp.volumeBVH = 1; lp.isoSurfaceBVH = 2; lp.streamlineBVH = 3; lp.surfaceModel = 4;   // indices into shim::geoms
  lp.fbSize = vec2ui(W, H);
  lp.worldSpaceBounds_lo = vec3f(of.worldLo[0], of.worldLo[1], of.worldLo[2]);
  lp.worldSpaceBounds_hi = vec3f(of.worldHi[0], of.worldHi[1], of.worldHi[2]);
  { const box3f vb = input->getBounds(); lp.voxelSpaceBounds_lo = vb.lower; lp.voxelSpaceBounds_hi = vb.upper; }
  lp.colorBufferPtr = rgba.data(); lp.accumBufferPtr = accum.data(); lp.frameStateBuffer = &fs;
  lp.dt = of.dt;
  lp.sameBrickRegionsBuffer = regions.brickRegions.data(); lp.sameBrickRegionsLeafList = regions.leafList.data();
  lp.brickBuffer = bricks.data(); lp.scalarBuffers = scalars.data(); lp.scalarBufferOffsets = offsets.data();
  lp.primaryChannels = primary.data(); lp.colormapChannel = of.colormapChannel; lp.numPrimaryChannels = of.numPrimaryChannels;
  lp.gradientShadingDVR = of.gradientShadingDVR; lp.gradientShadingISO = of.gradientShadingISO;
  lp.tracerEnabled = 0; lp.traces = traces.empty() ? nullptr : (vec3f *)traces.data();
  lp.tracerChannels = vec3i(0, 1, 2); lp.numTraces = of.numTraces; lp.numTimesteps = of.numTimesteps; lp.steplen = 0.f;
  lp.currentTimestep = &currentTimestep;

  // ---- geometries: run the reference's bounds programs, as owlGroupBuildAccel would ----------
  bool contour = false;
  for (int i = 0; i < MAX_CONTOUR_PLANES; i++) contour |= fs.contourPlane[i].enabled;
  static VolumeData vd; vd.frameStateBuffer = &fs; vd.sameBrickRegionsBuffer = lp.sameBrickRegionsBuffer; vd.brickBuffer = lp.brickBuffer;
  vd.valueRangePerBrickBuffer = nullptr; vd.numChannels = of.numPrimaryChannels;
  vd.spaceSkippingEnabled = (!contour && of.spaceSkipping) ? 1 : 0;                      // OptixRenderer.cpp:418-432, 518-530
  static IsoSurfaceData id; id.frameStateBuffer = &fs; id.sameBrickRegionsBuffer = lp.sameBrickRegionsBuffer; id.brickBuffer = lp.brickBuffer; id.valueRangePerBrickBuffer = nullptr;
  static StreamlineData sd; sd.traces = lp.traces; sd.numTraces = of.numTraces; sd.numTimesteps = of.numTimesteps; sd.currentTimestep = &currentTimestep;
  const int nr = (int)regions.brickRegions.size();
  shim::Geom &gv = shim::geoms[1]; gv.programData = &vd; gv.intersect = intersect_VolumeBVH; gv.bounds.resize(nr);
  for (int i = 0; i < nr; i++) bounds_VolumeBVH(&vd, gv.bounds[i], i);
  shim::Geom &gi = shim::geoms[2]; gi.programData = &id; gi.intersect = intersect_IsoSurface; gi.bounds.resize(nr);
  for (int i = 0; i < nr; i++) bounds_IsoSurface(&id, gi.bounds[i], i);
  shim::Geom &gs = shim::geoms[3]; gs.programData = &sd; gs.intersect = intersect_Streamline;
  if (lp.traces) {                                           // one prim per trace vertex (OptixRenderer.cpp:723-765)
    gs.bounds.resize((size_t)of.numTraces * of.numTimesteps);
    for (size_t i = 0; i + 1 < gs.bounds.size(); i++) bounds_Streamline(&sd, gs.bounds[i], (int)i);
    gs.bounds.back() = box3f();
  }
  // surfaceModel: the `surfaces` meshes (OptixRenderer.cpp:554-612).  OptiX's built-in triangle test has no source;
  // intersect_Triangles (above) stands in for it with the reference's own intersectRayTriangle, and the
  // reference's closest-hit program SurfaceBVH (:420-433) runs on the winner
  static SurfaceGeomData md; md.vertexBuffer = meshVerts.data(); md.indexBuffer = meshIndex.data();
  shim::Geom &gm = shim::geoms[4]; gm.programData = &md; gm.intersect = intersect_Triangles; gm.closestHit = closesthit_SurfaceBVH;
  gm.bounds.assign(meshIndex.size(), box3f(vec3f(0.f), vec3f(1.f)));      // non-empty: every triangle is tested

  // ---- point mode (ORC_POINTS=file of n x {x,y,z} floats, ORC_CHANNEL=c): the reference's reconstruction at
  // arbitrary points exactly as samplePointWithInfRay does it (exabrick.cu:818-830: region query with a ray of
  // length 2e-10, then samplePoint) plus samplePointWithDerivative (:883-942); out.bin = n x {value, valid, gx, gy, gz}
  if (const char *pf = getenv("ORC_POINTS")) {
    const int channel = getenv("ORC_CHANNEL") ? atoi(getenv("ORC_CHANNEL")) : 0;
    std::vector<float> pts = readFloats(pf);
    const size_t n = pts.size() / 3;
    std::vector<float> out(n * 5, 0.f);
#pragma omp parallel for schedule(dynamic, 256)
    for (size_t i = 0; i < n; i++) {
      const vec3f pos(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
      owl::Ray ray; ray.origin = pos; ray.direction = vec3f(1.f); ray.tmin = 0.f; ray.tmax = 2e-10f;
      VolumePRD prd = traceVolumeRay(ray);
      if (prd.leafID < 0) continue;                              // the reference would index region -1 here
      float v = 0.f, v2 = 0.f; vec3f g(0.f);
      const bool ok = samplePoint(v, prd.leafID, pos, channel);
      const bool ok2 = samplePointWithDerivative<false>(v2, g, prd.leafID, pos, channel);
      if (ok) { out[5 * i] = v; out[5 * i + 1] = (ok2 && v2 == v) ? 1.f : 2.f; out[5 * i + 2] = g.x; out[5 * i + 3] = g.y; out[5 * i + 4] = g.z; }
    }
    FILE *o = fopen(av[3], "wb");
    fwrite(out.data(), 4, out.size(), o);
    fclose(o);
    return 0;
  }

  // ---- tracer mode (ORC_TRACE="steplen,c0,c1,c2"): the reference's computeTraces (exabrick.cu:1531-1574),
  // one call per seed and time step as advanceTracer + render would do; out.bin = the trace vertices
  if (const char *e = getenv("ORC_TRACE")) {
    float steplen = 1.f; int c0 = 0, c1 = 1, c2 = 2;
    sscanf(e, "%f,%d,%d,%d", &steplen, &c0, &c1, &c2);
    lp.tracerEnabled = 1; lp.steplen = steplen; lp.tracerChannels = vec3i(c0, c1, c2);
    lp.fbSize = vec2ui(of.numTraces, 1);
    for (int t = 1; t < of.numTimesteps; t++) {
      currentTimestep = t;
#pragma omp parallel for schedule(dynamic, 4)
      for (int i = 0; i < of.numTraces; i++) {
        shim::ctx.launchIndex = vec2i(i, 0); shim::ctx.launchDims = vec2i(of.numTraces, 1);
        computeTraces();
      }
    }
    FILE *o = fopen(av[3], "wb");
    fwrite(traces.data(), 4, traces.size(), o);
    fclose(o);
    return 0;
  }

  // ---- launch: one raygen call per pixel (window given by ORC_WINDOW=x0,y0,x1,y1) ------------
  int x0 = 0, y0 = 0, x1 = W, y1 = H;
  if (const char *e = getenv("ORC_WINDOW")) sscanf(e, "%d,%d,%d,%d", &x0, &y0, &x1, &y1);
#pragma omp parallel for schedule(dynamic, 4) collapse(2)
  for (int y = y0; y < y1; y++)
    for (int x = x0; x < x1; x++) {
      shim::ctx.launchIndex = vec2i(x, y); shim::ctx.launchDims = vec2i(W, H);
      raygen_renderFrame();
    }
  FILE *o = fopen(av[3], "wb");
  fwrite(accum.data(), 16, accum.size(), o);
  fwrite(rgba.data(), 4, rgba.size(), o);
  fclose(o);
  return 0;
}
