// exa/exa.hpp — header-only C++ mirror of the reference's host classes for the rendering path,
// implemented on top of the C ABI in ../exabrick_b200.h (link with -lexabrick_b200).
//
// Same class and method names, argument meaning and error behaviour (std::runtime_error) as
//   exa::ExaBricks      exa/ExaBricks.h:22-45
//   exa::ScalarField    exa/ScalarField.h:24-40
//   exa::ColorMapper    exa/ColorMapper.h:29-83
//   exa::Config         exa/Config.h:26-49
//   exa::OptixRenderer  exa/OptixRenderer.h:32-188
// so that exa/viewer.cpp-style callers compile against it with the small math types below
// standing in for owl::vec3f / box3f / affine3f (same member names; no OWL dependency).
#pragma once
#include "../exabrick_b200.h"
#include <cstring>
#include <algorithm>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace exa {

struct vec2i { int x = 0, y = 0; };
struct vec3i { int x = 0, y = 0, z = 0; };
struct vec3f {
  float x = 0.f, y = 0.f, z = 0.f;
  vec3f() = default;
  explicit vec3f(float s) : x(s), y(s), z(s) {}
  vec3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
struct box3f { vec3f lower, upper; };
template <typename T> struct interval { T lower, upper; };
struct linear3f { vec3f vx{1, 0, 0}, vy{0, 1, 0}, vz{0, 0, 1}; };
struct affine3f { linear3f l; vec3f p; };

namespace detail {
inline void check(int rc) { if (rc != EXA_OK) throw std::runtime_error(exa_last_error()); }
}

// ---------------------------------------------------------------------------------------
struct ExaBricks {
  typedef std::shared_ptr<ExaBricks> SP;

  // ExaBricks::load (exa/ExaBricks.cpp:21-55)
  static SP load(const std::string &brickFileName)
  {
    exa_bricks_t *h = nullptr;
    detail::check(exa_bricks_load(brickFileName.c_str(), &h));
    return SP(new ExaBricks(h));
  }
  // exaBuilder (builder/builder.cpp:836-915) as a call: cells = n x {x,y,z,level}
  static SP build(const int32_t *cells, size_t numCells, int builderType = EXA_BUILDER_SAH_ALIKE, int maxLeafWidth = 127)
  {
    exa_bricks_t *h = nullptr;
    detail::check(exa_bricks_build(cells, numCells, builderType, maxLeafWidth, &h));
    return SP(new ExaBricks(h));
  }
  void save(const std::string &bricksFileName, const std::string &kdFileName = "") const
  { detail::check(exa_bricks_save(handle, bricksFileName.c_str(), kdFileName.empty() ? nullptr : kdFileName.c_str())); }

  box3f getBounds() const
  {
    box3f b; float lo[3], hi[3];
    detail::check(exa_bricks_get_bounds(handle, lo, hi));
    b.lower = vec3f(lo[0], lo[1], lo[2]); b.upper = vec3f(hi[0], hi[1], hi[2]);
    return b;
  }
  size_t numBricks() const { return exa_bricks_count(handle); }
  const ExaBrick *headers() const { return exa_bricks_headers(handle); }   // flattened Brick[] (programs/Brick.h)
  const int32_t *cellIDs() const { return exa_bricks_cell_ids(handle); }   // indexVector (OptixRenderer.h:100-104)

  size_t totalNumCells = 0;
  exa_bricks_t *handle = nullptr;
  ~ExaBricks() { if (owned) exa_bricks_destroy(handle); }
  ExaBricks(const ExaBricks &) = delete;
  ExaBricks &operator=(const ExaBricks &) = delete;
  explicit ExaBricks(exa_bricks_t *h, bool own = true) : totalNumCells(exa_bricks_num_cells(h)), handle(h), owned(own) {}
private:
  bool owned = true;
};

// ---------------------------------------------------------------------------------------
struct ScalarField {
  typedef std::shared_ptr<ScalarField> SP;

  // ScalarField::load (exa/ScalarField.cpp:22-45); `compat` keeps the reference's
  // bytes-as-count sizing quirk (value.size() == 4x the float count, zero tail)
  static SP load(const std::string &fieldName, const std::string &fileName, bool compat = true)
  {
    float *v = nullptr; size_t n = 0; float range[2];
    detail::check(exa_scalar_field_load(fileName.c_str(), compat ? 1 : 0, &v, &n, range));
    SP sf = std::make_shared<ScalarField>();
    sf->name = fieldName; sf->value.assign(v, v + n); sf->valueRange = {range[0], range[1]};
    exa_free(v);
    return sf;
  }
  // ScalarField::loadAndComputeMagnitude (exa/ScalarField.cpp:47-96)
  static SP loadAndComputeMagnitude(const std::string &fieldName, const std::string &fnx, const std::string &fny,
                                    const std::string &fnz, bool compat = true)
  {
    SP x = load("x", fnx, compat), y = load("y", fny, compat), z = load("z", fnz, compat);
    if (x->value.size() != y->value.size() || x->value.size() != z->value.size())
      throw std::runtime_error("vector field components differ in size");
    SP sf = std::make_shared<ScalarField>();
    sf->name = fieldName; sf->value.resize(x->value.size());
    float range[2];
    detail::check(exa_scalar_field_magnitude(x->value.data(), y->value.data(), z->value.data(), x->value.size(), sf->value.data(), range));
    sf->valueRange = {range[0], range[1]};
    return sf;
  }
  // ScalarField::createFromExpression (exa/ScalarField.cpp:96-230): postfix expression over `fields`
  static SP createFromExpression(const std::string &fieldName, const std::vector<SP> &fields, const std::vector<std::string> &tokens)
  {
    if (fields.empty()) throw std::runtime_error("createFromExpression: no input fields");
    std::vector<const float *> in; for (auto &f : fields) in.push_back(f->value.data());
    std::vector<const char *> tk; for (auto &t : tokens) tk.push_back(t.c_str());
    SP sf = std::make_shared<ScalarField>();
    sf->name = fieldName; sf->value.resize(fields[0]->value.size());
    float range[2];
    detail::check(exa_scalar_field_expression(in.data(), (int)in.size(), sf->value.size(), tk.data(), (int)tk.size(), sf->value.data(), range));
    sf->valueRange = {range[0], range[1]};
    return sf;
  }
  std::string name;
  interval<float> valueRange{0.f, 0.f};
  std::vector<float> value;
};

// ---------------------------------------------------------------------------------------
struct ColorMapper {
  typedef std::pair<float, vec3f> ControlPoint;
  ColorMapper(const vec3f *colors, std::size_t num)
  { for (std::size_t i = 0; i < num; i++) values_.push_back({num > 1 ? i / float(num - 1) : 0.f, colors[i]}); }
  ColorMapper(std::initializer_list<vec3f> colors) : ColorMapper(colors.begin(), colors.size()) {}
  ColorMapper(std::initializer_list<ControlPoint> colors) : values_(colors) {}
  // ColorMapper(std::string): "0.0,(0.0,0.1,0.2)\n1.0,(1.0,0.9,0.8)" (exa/ColorMapper.cpp:92-131)
  ColorMapper(std::string colorMapString)
  {
    float *cp = nullptr; int n = 0;
    detail::check(exa_colormap_parse_string(colorMapString.c_str(), &cp, &n));
    take(cp, n);
  }
  // ColorMapper(buf, size): first pixel row of a PNG (exa/ColorMapper.h:44-61); the viewer's embedded maps are in
  // include/exa/embedded_colormaps.h under the reference's names: ColorMapper(paraview_cool_warm, sizeof(paraview_cool_warm))
  ColorMapper(const uint8_t *buf, std::size_t buf_size)
  {
    float *cp = nullptr; int n = 0;
    detail::check(exa_colormap_from_png(buf, buf_size, &cp, &n));
    take(cp, n);
  }
  // ColorMapper::operator() (exa/ColorMapper.h:64-80): piecewise-linear over sorted control points
  vec3f operator()(float t) const
  {
    std::vector<float> cp; cp.reserve(values_.size() * 4);
    for (const ControlPoint &c : values_) { cp.push_back(c.first); cp.push_back(c.second.x); cp.push_back(c.second.y); cp.push_back(c.second.z); }
    float rgb[3];
    detail::check(exa_colormap_eval(cp.data(), (int)values_.size(), t, rgb));
    return vec3f(rgb[0], rgb[1], rgb[2]);
  }
  std::vector<ControlPoint> values_;
private:
  void take(float *cp, int n)
  {
    for (int i = 0; i < n; i++) values_.push_back({cp[4 * i], vec3f(cp[4 * i + 1], cp[4 * i + 2], cp[4 * i + 3])});
    exa_free(cp);
  }
};

// ---------------------------------------------------------------------------------------
// exa::TriangleMesh (exa/TriangleMesh.h:24-34): vertex / index arrays; load() reads every mesh of a file
struct TriangleMesh {
  typedef std::shared_ptr<TriangleMesh> SP;
  std::vector<vec3f> vertex;
  std::vector<vec3i> index;
  // the C ABI returns the meshes of a file concatenated; the mirror hands them back as one mesh
  static std::vector<SP> load(const std::string &fileName)
  {
    float *v = nullptr; int32_t *ix = nullptr; size_t nv = 0, nt = 0;
    detail::check(exa_triangle_mesh_load(fileName.c_str(), &v, &nv, &ix, &nt));
    SP m = std::make_shared<TriangleMesh>();
    m->vertex.resize(nv); m->index.resize(nt);
    if (nv) std::memcpy((void *)m->vertex.data(), v, nv * 12);
    if (nt) std::memcpy((void *)m->index.data(), ix, nt * 12);
    exa_free(v); exa_free(ix);
    std::vector<SP> out; if (nt) out.push_back(m);
    return out;
  }
};

// ---------------------------------------------------------------------------------------
struct Config {
  typedef std::shared_ptr<Config> SP;

  // Config::parseConfigFile + finalize (exa/Config.cpp:57-180)
  static SP parseConfigFile(const std::string &fileName, bool scalarCompat = true)
  {
    exa_config_t *c = nullptr;
    detail::check(exa_config_parse(fileName.c_str(), scalarCompat ? 1 : 0, &c));
    SP cfg(new Config(c));
    cfg->bricks.sp = ExaBricks::SP(new ExaBricks(const_cast<exa_bricks_t *>(exa_config_bricks(c)), /*own=*/false));
    const int n = exa_config_num_fields(c);
    for (int i = 0; i < n; i++) {
      size_t count = 0; float range[2]; const char *name = nullptr;
      const float *v = exa_config_field(c, i, &count, range, &name);
      ScalarField::SP sf = std::make_shared<ScalarField>();
      sf->name = name ? name : ""; sf->value.assign(v, v + count); sf->valueRange = {range[0], range[1]};
      cfg->scalarFields.push_back(sf);
    }
    {                                                      // config->surfaces (Config.cpp:165-171)
      const float *v = nullptr; const int32_t *ix = nullptr; size_t nv = 0, nt = 0;
      detail::check(exa_config_triangles(c, &v, &nv, &ix, &nt));
      if (nt) {
        TriangleMesh::SP mesh = std::make_shared<TriangleMesh>();
        mesh->vertex.resize(nv); mesh->index.resize(nt);
        std::memcpy((void *)mesh->vertex.data(), v, nv * 12); std::memcpy((void *)mesh->index.data(), ix, nt * 12);
        cfg->surfaces.push_back(mesh);
      }
    }
    float m[12];
    detail::check(exa_config_voxel_space_transform(c, m));
    affine3f &x = cfg->bricks.voxelSpaceTransform;
    x.l.vx = vec3f(m[0], m[1], m[2]); x.l.vy = vec3f(m[3], m[4], m[5]); x.l.vz = vec3f(m[6], m[7], m[8]); x.p = vec3f(m[9], m[10], m[11]);
    return cfg;
  }
  void finalize() {}   // done by parseConfigFile
  // Config::getBounds (exa/Config.cpp:48-55): world-space bounds
  box3f getBounds()
  {
    float lo[3], hi[3]; detail::check(exa_config_get_bounds(handle, lo, hi));
    box3f b; b.lower = vec3f(lo[0], lo[1], lo[2]); b.upper = vec3f(hi[0], hi[1], hi[2]); return b;
  }
  struct { ExaBricks::SP sp; affine3f voxelSpaceTransform; } bricks;
  std::vector<ScalarField::SP> scalarFields;
  std::vector<TriangleMesh::SP> surfaces;
  exa_config_t *handle = nullptr;
  ~Config() { bricks.sp.reset(); exa_config_destroy(handle); }
  Config(const Config &) = delete;
private:
  explicit Config(exa_config_t *c) : handle(c) {}
};

// ---------------------------------------------------------------------------------------
// exa::OptixRenderer (exa/OptixRenderer.h:32-188).  The name is kept so that exa/viewer.cpp
// compiles unchanged; nothing here uses OptiX.  `surfaces` are world-space triangle meshes.
struct OptixRenderer {
  typedef std::shared_ptr<OptixRenderer> SP;

  OptixRenderer(ExaBricks::SP input_, const std::vector<TriangleMesh::SP> &surfaces_, std::vector<ScalarField::SP> scalarField, int cudaDevice = 0)
      : scalarFields(std::move(scalarField)), input(std::move(input_))
  {
    if (!input) throw std::runtime_error("no bricks");
    std::vector<const float *> ptrs; std::vector<size_t> counts;
    for (auto &sf : scalarFields) { ptrs.push_back(sf->value.data()); counts.push_back(sf->value.size()); }
    detail::check(exa_renderer_create(input->handle, ptrs.data(), counts.data(), (int)ptrs.size(), cudaDevice, &handle));
    voxelSpaceBounds = input->getBounds(); worldSpaceBounds = voxelSpaceBounds;
    // one instance group over all meshes (OptixRenderer.cpp:556-608) = one concatenated mesh here
    std::vector<float> v; std::vector<int32_t> ix;
    for (auto &m : surfaces_) {
      if (!m) continue;
      const int32_t base = (int32_t)(v.size() / 3);
      for (auto &p : m->vertex) { v.push_back(p.x); v.push_back(p.y); v.push_back(p.z); }
      for (auto &t : m->index) { ix.push_back(t.x + base); ix.push_back(t.y + base); ix.push_back(t.z + base); }
    }
    if (!ix.empty()) detail::check(exa_renderer_set_triangle_meshes(handle, v.data(), v.size() / 3, ix.data(), ix.size() / 3));
  }
  ~OptixRenderer() { exa_renderer_destroy(handle); }
  OptixRenderer(const OptixRenderer &) = delete;

  void setVoxelSpaceTransform(const affine3f &x)
  {
    const float m[12] = {x.l.vx.x, x.l.vx.y, x.l.vx.z, x.l.vy.x, x.l.vy.y, x.l.vy.z, x.l.vz.x, x.l.vz.y, x.l.vz.z, x.p.x, x.p.y, x.p.z};
    detail::check(exa_renderer_set_voxel_space_transform(handle, m));
    ExaLaunchParams lp; detail::check(exa_renderer_get_launch_params(handle, &lp, nullptr));
    worldSpaceBounds.lower = vec3f(lp.worldSpaceBounds_lo[0], lp.worldSpaceBounds_lo[1], lp.worldSpaceBounds_lo[2]);
    worldSpaceBounds.upper = vec3f(lp.worldSpaceBounds_hi[0], lp.worldSpaceBounds_hi[1], lp.worldSpaceBounds_hi[2]);
  }
  // fbPointer: caller-owned device/managed RGBA8 buffer, or nullptr (renderer-owned; read_rgba8)
  void resizeFrameBuffer(void *fbPointer, const vec2i &size)
  { detail::check(exa_renderer_resize_framebuffer(handle, fbPointer, size.x, size.y)); fbSize = size; }
  void updateIsoValues(const float *isoValues, const int *channels, const int *enabled)
  { detail::check(exa_renderer_update_iso_values(handle, isoValues, channels, enabled)); }
  void updateContourPlanes(const vec3f *normals, const float *offsets, const int *channels, const int *enabled)
  { detail::check(exa_renderer_update_contour_planes(handle, &normals[0].x, offsets, channels, enabled)); }
  // the viewer pokes frameState.ao directly (viewer.cpp:953-954); here it is a call
  void setAO(bool enable, float length) { detail::check(exa_renderer_set_ao(handle, enable, length)); }
  void setCentralDifferenceGradients(bool on) { detail::check(exa_renderer_set_gradient_mode(handle, on ? EXA_GRADIENTS_CENTRAL_DIFFERENCES : EXA_GRADIENTS_ANALYTIC)); }   // the reference: -DANALYTIC_GRADIENTS=0 at build time
  void setClockScale(float s) { detail::check(exa_renderer_set_clock_scale(handle, s)); }   // the viewer writes frameState.clockScale (viewer.cpp:956)
  void updateCamera(const vec3f &pos, const vec3f &dir00, const vec3f &dirDu, const vec3f &dirDv)
  { detail::check(exa_renderer_update_camera(handle, &pos.x, &dir00.x, &dirDu.x, &dirDv.x)); }
  void updateXF(int chan, const float *opacities, const std::vector<vec3f> &colorMap, const interval<float> &xfDomain,
                float xfOpacityScale = .1f)
  {
    if (colorMap.size() != EXA_NUM_XF_VALUES) throw std::runtime_error("updateXF: colour map must have 128 entries");
    detail::check(exa_renderer_update_xf(handle, chan, opacities, &colorMap[0].x, xfDomain.lower, xfDomain.upper, xfOpacityScale));
  }
  void updateFrameID(int frameID) { detail::check(exa_renderer_update_frame_id(handle, frameID)); }
  void updateDt(float dt) { detail::check(exa_renderer_update_dt(handle, dt)); }
  void setSpaceSkipping(bool enable) { doSpaceSkipping = enable; detail::check(exa_renderer_set_space_skipping(handle, enable)); }
  void setGradientShadingDVR(bool enable) { gradientShadingDVR = enable; detail::check(exa_renderer_set_gradient_shading_dvr(handle, enable)); }
  void setGradientShadingISO(bool enable) { gradientShadingISO = enable; detail::check(exa_renderer_set_gradient_shading_iso(handle, enable)); }
  void setClipBox(bool enable, const box3f &b) { detail::check(exa_renderer_set_clip_box(handle, enable, &b.lower.x, &b.upper.x)); }
  void setTracerEnabled(bool enable) { traces.tracerEnabled = enable; detail::check(exa_renderer_set_tracer_enabled(handle, enable)); }
  void resetTracer()
  {
    const int ch[3] = {traces.tracerChannels.x, traces.tracerChannels.y, traces.tracerChannels.z};
    detail::check(exa_renderer_set_tracer_params(handle, traces.numTraces, traces.numTimesteps, traces.steplen, ch,
                                                 &traces.seedRegion.lower.x, &traces.seedRegion.upper.x));
    detail::check(exa_renderer_reset_tracer(handle));
  }
  bool advanceTracer() { int rebuilt = 0; detail::check(exa_renderer_advance_tracer(handle, &rebuilt)); return rebuilt != 0; }
  void render() { detail::check(exa_renderer_render(handle)); }

  // headless extras
  void readColorBuffer(uint32_t *hostRGBA8) { detail::check(exa_renderer_read_rgba8(handle, hostRGBA8)); }
  // pipelined variant: renderAsync(); readColorBufferAsync(pinned); ... sync() — frame i+1 renders while frame i is copied
  void renderAsync() { detail::check(exa_renderer_render_async(handle)); }
  void readColorBufferAsync(uint32_t *pinnedHostRGBA8) { detail::check(exa_renderer_read_rgba8_async(handle, pinnedHostRGBA8)); }
  void sync() { detail::check(exa_renderer_sync(handle)); }
  // swap the RGBA8 target without touching the accumulation buffer (swap chains, peer-mapped buffers); nullptr = renderer-owned
  void setColorBuffer(void *fbPointer) { detail::check(exa_renderer_set_color_buffer(handle, fbPointer)); }
  void readAccumBuffer(float *hostRGBA) { detail::check(exa_renderer_read_accum(handle, hostRGBA)); }
  void samplePoints(const float *xyz, size_t n, int channel, float *values, uint8_t *valid, float *grad = nullptr)
  { detail::check(exa_sample_points_host(handle, xyz, n, channel, values, valid, grad)); }

  std::vector<ScalarField::SP> scalarFields;
  ExaBricks::SP input;
  box3f voxelSpaceBounds, worldSpaceBounds;
  bool multiFieldDvr = true, gradientShadingDVR = true, gradientShadingISO = true, doSpaceSkipping = true;
  vec2i fbSize;
  struct {
    int tracerEnabled = false;
    vec3i tracerChannels{0, 1, 2};
    int numTraces = 1000, numTimesteps = 100;
    float steplen = 1e-6f;
    box3f seedRegion{{.3f, .3f, .5f}, {.8f, .8f, .5f}};
  } traces;
  exa_renderer_t *handle = nullptr;
};

} // namespace exa
