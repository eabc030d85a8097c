// host.cpp — host-side data model and loaders behind the C ABI (no CUDA here):
//   exa_bricks_*        exa::ExaBricks + .bricks reader/writer (exa/ExaBricks.cpp:21-63,
//                       builder/builder.cpp:895-907)
//   exa_scalar_field_*  exa::ScalarField (exa/ScalarField.cpp:22-96)
//   exa_colormap_eval   exa::ColorMapper::operator() (exa/ColorMapper.h:64-80)
//   exa_config_*        exa::Config (exa/Config.cpp:23-180)
#include "exa_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

namespace exa {

static thread_local std::string g_lastError;
void setError(const std::string &msg) { g_lastError = msg; }

struct BricksImpl {
  std::vector<ExaBrick> hdr;          // flattened headers, begin = running cell offset
  IdVector ids;                       // indexVector
  std::vector<ExaKdTreeNode> kd;      // builder kd-tree (empty when loaded from file)
};
BricksImpl *bricksFromParts(std::vector<ExaBrick> &&hdr, IdVector &&ids, std::vector<ExaKdTreeNode> &&kd)
{
  BricksImpl *b = new BricksImpl; b->hdr = std::move(hdr); b->ids = std::move(ids); b->kd = std::move(kd); return b;
}
BricksImpl *buildBricks(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth);   // builder.cpp
BricksImpl *buildBricksGpu(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth, int cudaDevice, double timings[8]);   // gpu_builder.cu

// .bricks stream: repeated {int32 size[3]; int32 lower[3]; int32 level; int32 cellID[sx*sy*sz]}
static BricksImpl *parseBricks(const uint8_t *p, size_t n)
{
  BricksImpl *b = new BricksImpl;
  size_t off = 0;
  while (off + 28 <= n) {                     // a short trailing header ends the stream, ExaBricks.cpp:27-34
    int32_t h[7]; memcpy(h, p + off, 28); off += 28;
    if (h[0] < 0 || h[1] < 0 || h[2] < 0) { delete b; throw std::runtime_error("negative brick size in .bricks stream"); }
    const size_t cells = (size_t)h[0] * (size_t)h[1] * (size_t)h[2];
    ExaBrick e; e.size[0] = h[0]; e.size[1] = h[1]; e.size[2] = h[2]; e.lower[0] = h[3]; e.lower[1] = h[4]; e.lower[2] = h[5]; e.level = h[6];
    if (b->ids.size() > 0x7fffffffull) { delete b; throw std::runtime_error("32-bit offset overflow"); }   // OptixRenderer.cpp:81-83
    e.begin = (uint32_t)b->ids.size();
    const size_t avail = std::min(cells, (n - off) / 4);
    const size_t base = b->ids.size();
    b->ids.resize(base + cells, 0);           // a truncated last brick reads as zeros (istream::read leaves the rest)
    memcpy(b->ids.data() + base, p + off, avail * 4);
    off += avail * 4;
    b->hdr.push_back(e);
    if (avail < cells) break;
  }
  return b;
}

} // namespace exa

using namespace exa;

struct exa_bricks { BricksImpl *impl; };

#define EXA_TRY try {
#define EXA_CATCH } catch (std::bad_alloc &) { setError("out of host memory"); return EXA_ERR_LIMIT; } \
                    catch (std::exception &e) { setError(e.what()); return EXA_ERR_INVALID; }

extern "C" {

const char *exa_last_error(void) { return g_lastError.c_str(); }
void exa_free(void *p) { free(p); }

int exa_bricks_build(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth, exa_bricks_t **out)
{
  if (!out || (!cells && numCells)) { setError("exa_bricks_build: null argument"); return EXA_ERR_INVALID; }
  if (builderType != EXA_BUILDER_SAH_ALIKE && builderType != EXA_BUILDER_SPATIAL_MEDIAN) { setError("exa_bricks_build: unknown builder type"); return EXA_ERR_INVALID; }
  if (maxLeafWidth <= 0) maxLeafWidth = 127;
  EXA_TRY
  *out = new exa_bricks{buildBricks(cells, numCells, builderType, maxLeafWidth)};
  return EXA_OK;
  EXA_CATCH
}

int exa_bricks_build_gpu(const int32_t *cells, size_t numCells, int builderType, int maxLeafWidth, int cudaDevice,
                         exa_bricks_t **out, double timings[8])
{
  if (!out || (!cells && numCells)) { setError("exa_bricks_build_gpu: null argument"); return EXA_ERR_INVALID; }
  if (builderType != EXA_BUILDER_SAH_ALIKE && builderType != EXA_BUILDER_SPATIAL_MEDIAN) { setError("exa_bricks_build_gpu: unknown builder type"); return EXA_ERR_INVALID; }
  if (maxLeafWidth <= 0) maxLeafWidth = 127;
  if (!exa_cuda_available()) return EXA_ERR_CUDA;            // no CPU fallback: exa_bricks_build is the host builder
  EXA_TRY
  *out = new exa_bricks{buildBricksGpu(cells, numCells, builderType, maxLeafWidth, cudaDevice, timings)};
  return EXA_OK;
  EXA_CATCH
}

static int readFile(const char *fn, std::vector<uint8_t> &buf)
{
  FILE *f = fn ? fopen(fn, "rb") : nullptr;
  if (!f) { setError(std::string("could not open ") + (fn ? fn : "(null)")); return EXA_ERR_IO; }
  fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
  buf.resize(n > 0 ? n : 0);
  size_t got = buf.empty() ? 0 : fread(buf.data(), 1, buf.size(), f);
  fclose(f);
  if (got != buf.size()) { setError(std::string("short read on ") + fn); return EXA_ERR_IO; }
  return EXA_OK;
}

int exa_bricks_from_memory(const void *bytes, size_t n, exa_bricks_t **out)
{
  if (!out || (!bytes && n)) { setError("exa_bricks_from_memory: null argument"); return EXA_ERR_INVALID; }
  EXA_TRY
  *out = new exa_bricks{parseBricks((const uint8_t *)bytes, n)};
  return EXA_OK;
  EXA_CATCH
}

int exa_bricks_load(const char *fileName, exa_bricks_t **out)
{
  std::vector<uint8_t> buf;
  int rc = readFile(fileName, buf); if (rc) return rc;
  return exa_bricks_from_memory(buf.data(), buf.size(), out);
}

int exa_bricks_save(const exa_bricks_t *b, const char *bricksFileName, const char *kdFileName)
{
  if (!b || !bricksFileName) { setError("exa_bricks_save: null argument"); return EXA_ERR_INVALID; }
  FILE *f = fopen(bricksFileName, "wb");
  if (!f) { setError(std::string("could not open ") + bricksFileName); return EXA_ERR_IO; }
  for (auto &h : b->impl->hdr) {
    int32_t rec[7] = {h.size[0], h.size[1], h.size[2], h.lower[0], h.lower[1], h.lower[2], h.level};
    fwrite(rec, 4, 7, f);
    fwrite(b->impl->ids.data() + h.begin, 4, (size_t)h.size[0] * h.size[1] * h.size[2], f);
  }
  fclose(f);
  if (kdFileName) {
    FILE *k = fopen(kdFileName, "wb");
    if (!k) { setError(std::string("could not open ") + kdFileName); return EXA_ERR_IO; }
    fwrite(b->impl->kd.data(), sizeof(ExaKdTreeNode), b->impl->kd.size(), k);
    fclose(k);
  }
  return EXA_OK;
}

void exa_bricks_destroy(exa_bricks_t *b) { if (b) { delete b->impl; delete b; } }
size_t exa_bricks_count(const exa_bricks_t *b) { return b ? b->impl->hdr.size() : 0; }
size_t exa_bricks_num_cells(const exa_bricks_t *b) { return b ? b->impl->ids.size() : 0; }
const ExaBrick *exa_bricks_headers(const exa_bricks_t *b) { return b ? b->impl->hdr.data() : nullptr; }
const int32_t *exa_bricks_cell_ids(const exa_bricks_t *b) { return b ? b->impl->ids.data() : nullptr; }
size_t exa_bricks_kd_count(const exa_bricks_t *b) { return b ? b->impl->kd.size() : 0; }
const ExaKdTreeNode *exa_bricks_kd_nodes(const exa_bricks_t *b) { return b ? b->impl->kd.data() : nullptr; }

int exa_bricks_get_bounds(const exa_bricks_t *b, float lower[3], float upper[3])
{
  if (!b || !lower || !upper) { setError("exa_bricks_get_bounds: null argument"); return EXA_ERR_INVALID; }
  for (int d = 0; d < 3; d++) { lower[d] = INFINITY; upper[d] = -INFINITY; }
  for (auto &h : b->impl->hdr)          // ExaBricks::getBounds over Brick::getBounds (ExaBricks.h:48-52)
    for (int d = 0; d < 3; d++) {
      lower[d] = std::min(lower[d], (float)h.lower[d]);
      upper[d] = std::max(upper[d], (float)(h.lower[d] + h.size[d] * (1 << h.level)));
    }
  return EXA_OK;
}

// ---------------------------------------------------------------------------------------
// exaRawToCells (tools/fromVolume/raw2cells.cpp:140-283): structured volume -> AMR cells.  Root
// cells of width refinement^maxLevel tile the volume (x fastest); a cell whose value range is
// <= threshold becomes one coarse cell carrying the average, otherwise its refinement^3 children
// are visited (x fastest), down to single voxels.  Arithmetic kept as written: the sum of
// `average` accumulates in the voxel type (bytes wrap), is divided by the UNCLIPPED cell volume
// as float and converted back to the voxel type; level = int(log2f(cellWidth)).
extern "C++" {
namespace {
template <typename T> struct RawVolume {
  const T *elt; int nx, ny, nz;
  T at(int x, int y, int z) const { return elt[x + (size_t)nx * (y + (size_t)ny * (size_t)z)]; }
};
template <typename T>
void rawBuildCell(const RawVolume<T> &v, int lx, int ly, int lz, int level, float threshold, int refinement,
                  std::vector<int32_t> &cells, std::vector<float> &scalars)
{
  if (lx >= v.nx || ly >= v.ny || lz >= v.nz) return;
  if (level == 0) {
    cells.insert(cells.end(), {lx, ly, lz, 0}); scalars.push_back((float)v.at(lx, ly, lz));
    return;
  }
  int childWidth = 1;
  for (int i = 0; i < level - 1; i++) childWidth *= refinement;
  const int cellWidth = childWidth * refinement;
  const int ux = std::min(v.nx, lx + cellWidth), uy = std::min(v.ny, ly + cellWidth), uz = std::min(v.nz, lz + cellWidth);
  T lo = v.at(lx, ly, lz), hi = lo;
  for (int z = lz; z < uz; z++) for (int y = ly; y < uy; y++) for (int x = lx; x < ux; x++) { const T t = v.at(x, y, z); lo = std::min(lo, t); hi = std::max(hi, t); }
  if ((hi - lo) <= threshold) {
    T sum = 0.f;
    for (int z = lz; z < uz; z++) for (int y = ly; y < uy; y++) for (int x = lx; x < ux; x++) sum += v.at(x, y, z);
    const T avg = sum / (float)(cellWidth * cellWidth * cellWidth);
    cells.insert(cells.end(), {lx, ly, lz, int(log2f((float)cellWidth))}); scalars.push_back((float)avg);
    return;
  }
  for (int iz = 0; iz < refinement; iz++) for (int iy = 0; iy < refinement; iy++) for (int ix = 0; ix < refinement; ix++)
    rawBuildCell(v, lx + ix * childWidth, ly + iy * childWidth, lz + iz * childWidth, level - 1, threshold, refinement, cells, scalars);
}
template <typename T>
void rawMakeCells(const void *voxels, int nx, int ny, int nz, float threshold, int maxLevel, int refinement,
                  std::vector<int32_t> &cells, std::vector<float> &scalars)
{
  RawVolume<T> v{(const T *)voxels, nx, ny, nz};
  int root = 1;
  for (int i = 0; i < maxLevel; i++) root *= refinement;
  const int rx = (nx + root - 1) / root, ry = (ny + root - 1) / root, rz = (nz + root - 1) / root;
  for (int z = 0; z < rz; z++) for (int y = 0; y < ry; y++) for (int x = 0; x < rx; x++)
    rawBuildCell(v, x * root, y * root, z * root, maxLevel, threshold, refinement, cells, scalars);
}
}
} // extern "C++"

int exa_raw_to_cells(const void *voxels, int format, int nx, int ny, int nz, float threshold, int maxLevel, int refinement,
                     int32_t **cells, float **scalars, size_t *numCells)
{
  if (!voxels || !cells || !scalars || !numCells || nx <= 0 || ny <= 0 || nz <= 0 || maxLevel < 0 || maxLevel > 20 || refinement < 2)
  { setError("exa_raw_to_cells: bad argument"); return EXA_ERR_INVALID; }
  std::vector<int32_t> c; std::vector<float> s;
  if (format == EXA_RAW_BYTE) rawMakeCells<uint8_t>(voxels, nx, ny, nz, threshold, maxLevel, refinement, c, s);
  else if (format == EXA_RAW_FLOAT) rawMakeCells<float>(voxels, nx, ny, nz, threshold, maxLevel, refinement, c, s);
  else if (format == EXA_RAW_DOUBLE) rawMakeCells<double>(voxels, nx, ny, nz, threshold, maxLevel, refinement, c, s);
  else { setError("unknown format!?"); return EXA_ERR_INVALID; }             // raw2cells.cpp:268
  int32_t *oc = (int32_t *)malloc(std::max<size_t>(1, c.size()) * 4);
  float *os = (float *)malloc(std::max<size_t>(1, s.size()) * 4);
  if (!oc || !os) { free(oc); free(os); setError("out of host memory"); return EXA_ERR_LIMIT; }
  memcpy(oc, c.data(), c.size() * 4); memcpy(os, s.data(), s.size() * 4);
  *cells = oc; *scalars = os; *numCells = s.size();
  return EXA_OK;
}

// ---------------------------------------------------------------------------------------
// TriangleMesh::load (exa/TriangleMesh.cpp:21-68): a file is a sequence of meshes, each
// {int32 numVerts; vec3f[numVerts]; int32 numTris; vec3i[numTris]}; here all meshes of the file are
// returned as ONE vertex / index array (indices rebased), which is what the renderer consumes.
int exa_triangle_mesh_load(const char *fileName, float **vertices, size_t *numVerts, int32_t **indices, size_t *numTris)
{
  if (!vertices || !numVerts || !indices || !numTris) { setError("exa_triangle_mesh_load: null argument"); return EXA_ERR_INVALID; }
  std::vector<uint8_t> buf;
  int rc = readFile(fileName, buf);
  if (rc) { setError("cannot open file"); return rc; }
  std::vector<float> v; std::vector<int32_t> ix;
  size_t off = 0;
  while (off + 4 <= buf.size()) {
    int32_t nv; memcpy(&nv, buf.data() + off, 4); off += 4;
    if (nv < 0 || off + (size_t)nv * 12 + 4 > buf.size()) { setError("broken triangle model"); return EXA_ERR_INVALID; }
    const size_t base = v.size() / 3;
    v.resize(v.size() + 3 * (size_t)nv);
    memcpy(v.data() + 3 * base, buf.data() + off, (size_t)nv * 12); off += (size_t)nv * 12;
    int32_t nt; memcpy(&nt, buf.data() + off, 4); off += 4;
    if (nt < 0 || off + (size_t)nt * 12 > buf.size()) { setError("broken triangle model"); return EXA_ERR_INVALID; }
    const size_t tbase = ix.size();
    ix.resize(ix.size() + 3 * (size_t)nt);
    memcpy(ix.data() + tbase, buf.data() + off, (size_t)nt * 12); off += (size_t)nt * 12;
    for (size_t i = tbase; i < ix.size(); i++) {
      if (ix[i] < 0 || ix[i] >= nv) { setError("broken triangle model"); return EXA_ERR_INVALID; }   // :39-47
      ix[i] += (int32_t)base;
    }
  }
  float *ov = (float *)malloc(std::max<size_t>(1, v.size()) * 4);
  int32_t *oi = (int32_t *)malloc(std::max<size_t>(1, ix.size()) * 4);
  if (!ov || !oi) { free(ov); free(oi); setError("out of host memory"); return EXA_ERR_LIMIT; }
  memcpy(ov, v.data(), v.size() * 4); memcpy(oi, ix.data(), ix.size() * 4);
  *vertices = ov; *numVerts = v.size() / 3; *indices = oi; *numTris = ix.size() / 3;
  return EXA_OK;
}

// ---------------------------------------------------------------------------------------
int exa_scalar_field_load(const char *fileName, int compat, float **values, size_t *count, float valueRange[2])
{
  if (!values || !count) { setError("exa_scalar_field_load: null argument"); return EXA_ERR_INVALID; }
  std::vector<uint8_t> buf;
  int rc = readFile(fileName, buf);
  if (rc) { setError(std::string("error in opening scalar file ") + (fileName ? fileName : "")); return rc; }
  const size_t nFloats = buf.size() / 4;
  // ScalarField.cpp:27-34 sizes the vector by file BYTES; compat keeps that (zero tail)
  const size_t n = compat ? buf.size() : nFloats;
  float *v = (float *)calloc(n ? n : 1, sizeof(float));
  if (!v) { setError("out of host memory"); return EXA_ERR_LIMIT; }
  memcpy(v, buf.data(), nFloats * 4);
  float lo = INFINITY, hi = -INFINITY;
  for (size_t i = 0; i < n; i++) { lo = std::min(lo, v[i]); hi = std::max(hi, v[i]); }
  if (valueRange) { valueRange[0] = lo; valueRange[1] = hi; }
  *values = v; *count = n;
  return EXA_OK;
}

// ScalarField::createFromExpression (exa/ScalarField.cpp:96-230): a postfix expression over the fields loaded
// so far, evaluated per cell on a float stack.  Tokens: %k (field k), numbers, + - * / ** == != < > <= >=,
// log abs sqrt, select (mask op1 op2 -> mask ? op1 : op2, mask truncated to int).  Tokens are trimmed of
// double quotes and blanks as the reference does.  Where the reference prints a complaint and carries on
// with an undefined value (bad token, stack underflow, more than one value left), this returns an error.
int exa_scalar_field_expression(const float *const *fields, int numFields, size_t count, const char *const *tokens, int numTokens,
                                float *out, float valueRange[2])
{
  if (!fields || numFields <= 0 || !tokens || numTokens <= 0 || !out) { setError("exa_scalar_field_expression: bad argument"); return EXA_ERR_INVALID; }
  enum Op { FIELD, CONST, ADD, SUB, MUL, DIV, POW, EQ, NE, LT, GT, LE, GE, LOG, ABS, SQRT, SELECT };
  struct Tok { Op op; int field; float value; };
  std::vector<Tok> prog;
  for (int i = 0; i < numTokens; i++) {
    std::string t = tokens[i] ? tokens[i] : "";
    auto strip = [&](char c) { while (!t.empty() && t.front() == c) t.erase(t.begin()); while (!t.empty() && t.back() == c) t.pop_back(); };
    strip('"'); strip(' ');
    if (t.empty()) { setError("expression: empty token"); return EXA_ERR_INVALID; }
    static const struct { const char *name; Op op; } table[] = {{"+", ADD}, {"-", SUB}, {"*", MUL}, {"/", DIV}, {"**", POW}, {"==", EQ}, {"!=", NE},
      {"<", LT}, {">", GT}, {"<=", LE}, {">=", GE}, {"log", LOG}, {"abs", ABS}, {"sqrt", SQRT}, {"select", SELECT}};
    Tok k; k.field = 0; k.value = 0.f; bool found = false;
    for (auto &e : table) if (t == e.name) { k.op = e.op; found = true; }
    if (!found) {
      try {
        if (t[0] == '%') { k.op = FIELD; k.field = std::stoi(t.substr(t.find_first_not_of('%'))); if (k.field < 0 || k.field >= numFields) { setError("expression: invalid placeholder token " + t); return EXA_ERR_INVALID; } }
        else { k.op = CONST; k.value = (float)std::stod(t); }
      } catch (...) { setError("expression: not a floating point token: " + t); return EXA_ERR_INVALID; }
    }
    prog.push_back(k);
  }
  int depth = 0;                                             // validate the stack discipline once
  for (auto &k : prog) {
    const int need = k.op <= CONST ? 0 : k.op == SELECT ? 3 : k.op >= LOG ? 1 : 2;
    if (depth < need) { setError("expression: insufficient operands"); return EXA_ERR_INVALID; }
    depth += 1 - need;
    if (depth > 62) { setError("expression: deeper than the 62-entry evaluation stack"); return EXA_ERR_LIMIT; }
  }
  if (depth != 1) { setError("Invalid expression"); return EXA_ERR_INVALID; }
  float lo = INFINITY, hi = -INFINITY;
#pragma omp parallel for reduction(min : lo) reduction(max : hi)
  for (long long i = 0; i < (long long)count; i++) {
    float st[64]; int sp = 0;
    for (auto &k : prog) {
      switch (k.op) {
        case FIELD: st[sp++] = fields[k.field][i]; break;
        case CONST: st[sp++] = k.value; break;
        case SELECT: { const float op2 = st[--sp], op1 = st[--sp]; const int mask = (int)st[--sp]; st[sp++] = mask ? op1 : op2; break; }
        case LOG: st[sp - 1] = logf(st[sp - 1]); break;
        case ABS: st[sp - 1] = fabsf(st[sp - 1]); break;
        case SQRT: st[sp - 1] = sqrtf(st[sp - 1]); break;
        default: {
          const float op2 = st[--sp], op1 = st[--sp]; float r = 0.f;
          switch (k.op) {
            case ADD: r = op1 + op2; break; case SUB: r = op1 - op2; break; case MUL: r = op1 * op2; break; case DIV: r = op1 / op2; break;
            case POW: r = powf(op1, op2); break; case EQ: r = op1 == op2; break; case NE: r = op1 != op2; break;
            case LT: r = op1 < op2; break; case GT: r = op1 > op2; break; case LE: r = op1 <= op2; break; case GE: r = op1 >= op2; break;
            default: break;
          }
          st[sp++] = r;
        }
      }
    }
    const float f = st[0];
    out[i] = f; lo = std::min(lo, f); hi = std::max(hi, f);
  }
  if (valueRange) { valueRange[0] = lo; valueRange[1] = hi; }
  return EXA_OK;
}

int exa_scalar_field_magnitude(const float *x, const float *y, const float *z, size_t count, float *out, float valueRange[2])
{
  if (!x || !y || !z || !out) { setError("exa_scalar_field_magnitude: null argument"); return EXA_ERR_INVALID; }
  float lo = INFINITY, hi = -INFINITY;
  for (size_t i = 0; i < count; i++) {
    const float f = sqrtf(x[i] * x[i] + y[i] * y[i] + z[i] * z[i]);   // owl length(vec3f)
    out[i] = f; lo = std::min(lo, f); hi = std::max(hi, f);
  }
  if (valueRange) { valueRange[0] = lo; valueRange[1] = hi; }
  return EXA_OK;
}

int exa_colormap_eval(const float *cp, int n, float t, float rgb[3])
{
  if (!cp || n <= 0 || !rgb) { setError("exa_colormap_eval: bad argument"); return EXA_ERR_INVALID; }
  // std::upper_bound on the control-point positions (ColorMapper.h:70-71)
  int it = 0;
  { int lo = 0, hi = n; while (lo < hi) { int mid = (lo + hi) / 2; if (t < cp[4 * mid]) hi = mid; else lo = mid + 1; } it = lo; }
  const int i1 = it > 0 ? it - 1 : it, i2 = it >= n ? it - 1 : it;
  const float t1 = cp[4 * i1], t2 = cp[4 * i2];
  const float x = t1 == t2 ? 1.f : (t - t1) / (t2 - t1);
  for (int c = 0; c < 3; c++) rgb[c] = (1.0f - x) * cp[4 * i1 + 1 + c] + x * cp[4 * i2 + 1 + c];
  return EXA_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------
// Config (exa/Config.cpp)
// ---------------------------------------------------------------------------------------
struct exa_config {
  exa_bricks_t *bricks = nullptr;
  struct Field { std::string name; float *values = nullptr; size_t count = 0; float range[2] = {INFINITY, -INFINITY}; };
  std::vector<Field> fields;
  float remapFrom[6] = {0, 0, 0, 1, 1, 1}, remapTo[6] = {0, 0, 0, 1, 1, 1};
  float xfm[12];
  float *meshVerts = nullptr; int32_t *meshIndex = nullptr; size_t numMeshVerts = 0, numMeshTris = 0;   // `triangles <file>`
};

// 3x3 (column vectors) + translation helpers restating owl AffineSpace3f rcp / operator*
static void affineInverse(const float m[12], float out[12])
{
  const float *vx = m, *vy = m + 3, *vz = m + 6, *p = m + 9;
  // adjoint / det, LinearSpace3::inverse (owl/common/math/LinearSpace.h)
  const float cx[3] = {vy[1] * vz[2] - vy[2] * vz[1], vy[2] * vz[0] - vy[0] * vz[2], vy[0] * vz[1] - vy[1] * vz[0]};   // cross(vy,vz)
  const float cy[3] = {vz[1] * vx[2] - vz[2] * vx[1], vz[2] * vx[0] - vz[0] * vx[2], vz[0] * vx[1] - vz[1] * vx[0]};   // cross(vz,vx)
  const float cz[3] = {vx[1] * vy[2] - vx[2] * vy[1], vx[2] * vy[0] - vx[0] * vy[2], vx[0] * vy[1] - vx[1] * vy[0]};   // cross(vx,vy)
  const float det = vx[0] * cx[0] + vx[1] * cx[1] + vx[2] * cx[2];
  // inverse has ROWS cx,cy,cz / det  => columns:
  float il[9] = {cx[0] / det, cy[0] / det, cz[0] / det, cx[1] / det, cy[1] / det, cz[1] / det, cx[2] / det, cy[2] / det, cz[2] / det};
  memcpy(out, il, sizeof(il));
  for (int r = 0; r < 3; r++) out[9 + r] = -(il[r] * p[0] + il[3 + r] * p[1] + il[6 + r] * p[2]);
}
static void affineMul(const float a[12], const float b[12], float out[12])
{
  float r[12];
  for (int c = 0; c < 4; c++)
    for (int row = 0; row < 3; row++) {
      const float *bc = b + 3 * c;
      r[3 * c + row] = a[row] * bc[0] + a[3 + row] * bc[1] + a[6 + row] * bc[2] + (c == 3 ? a[9 + row] : 0.f);
    }
  memcpy(out, r, sizeof(r));
}
static void translateScale(const float box[6], float out[12])
{
  memset(out, 0, 48);
  out[0] = box[3] - box[0]; out[4] = box[4] - box[1]; out[8] = box[5] - box[2];
  out[9] = box[0]; out[10] = box[1]; out[11] = box[2];
}

extern "C" {

int exa_config_parse(const char *fileName, int scalarCompat, exa_config_t **out)
{
  if (!fileName || !out) { setError("exa_config_parse: null argument"); return EXA_ERR_INVALID; }
  FILE *file = fopen(fileName, "r");
  if (!file) { setError(std::string("error in opening config file '") + fileName + "'"); return EXA_ERR_IO; }
  std::vector<std::string> tokens;
  static const int LINE_SZ = 10000; std::vector<char> line(LINE_SZ + 1);
  while (fgets(line.data(), LINE_SZ, file) && !feof(file)) {          // Config.cpp:72-78
    char *save = nullptr;
    char *tok = strtok_r(line.data(), " \t\n\r", &save);
    while (tok && tok[0] != '#') { tokens.push_back(tok); tok = strtok_r(nullptr, " \t\n\r", &save); }
  }
  fclose(file);
  std::string fn(fileName);
  const std::string basePath = fn.substr(0, fn.rfind('/')) + "/";
  exa_config *cfg = new exa_config;
  auto fail = [&](int rc, const std::string &msg) { if (!msg.empty()) setError(msg); exa_config_destroy(cfg); return rc; };
  try {
    size_t it = 0;
    auto need = [&](size_t k) { if (it + k >= tokens.size()) throw std::runtime_error("error in parsing config file: truncated '" + tokens[it] + "'"); };
    while (it < tokens.size()) {
      const std::string &t = tokens[it];
      if (t == "remap_from" || t == "remap_to") {
        need(6);
        float *dst = t == "remap_from" ? cfg->remapFrom : cfg->remapTo;
        for (int i = 0; i < 6; i++) dst[i] = std::stof(tokens[it + 1 + i]);
        it += 7; continue;
      }
      if (t == "scalar") {
        need(2);
        if (tokens[it + 2] == "expr") {                      // Config.cpp:113-123: tokens up to the one that ends with a quote
          if (cfg->fields.empty()) return fail(EXA_ERR_INVALID, "config: expression before any scalar field");
          std::vector<const char *> ex; size_t j = it + 3;
          for (;; j++) { if (j >= tokens.size()) return fail(EXA_ERR_INVALID, "config: unterminated expression"); ex.push_back(tokens[j].c_str()); if (tokens[j].back() == '"') break; }
          std::vector<const float *> in; for (auto &f : cfg->fields) in.push_back(f.values);
          exa_config::Field f; f.name = tokens[it + 1]; f.count = cfg->fields[0].count; f.values = (float *)malloc(4 * (f.count ? f.count : 1));
          int rc = exa_scalar_field_expression(in.data(), (int)in.size(), f.count, ex.data(), (int)ex.size(), f.values, f.range);
          if (rc) { free(f.values); return fail(rc, ""); }
          cfg->fields.push_back(f); it = j + 1; continue;
        }
        exa_config::Field f; f.name = tokens[it + 1];
        int rc = exa_scalar_field_load((basePath + tokens[it + 2]).c_str(), scalarCompat, &f.values, &f.count, f.range);
        if (rc) return fail(rc, "");
        cfg->fields.push_back(f); it += 3; continue;
      }
      if (t == "vector") {
        need(4);
        exa_config::Field f[3]; int rc = 0;
        for (int k = 0; k < 3 && !rc; k++) rc = exa_scalar_field_load((basePath + tokens[it + 2 + k]).c_str(), scalarCompat, &f[k].values, &f[k].count, f[k].range);
        if (!rc && (f[0].count != f[1].count || f[0].count != f[2].count)) { setError("vector components differ in size"); rc = EXA_ERR_INVALID; }
        if (rc) { for (auto &x : f) free(x.values); return fail(rc, ""); }
        exa_config::Field m; m.name = tokens[it + 1]; m.count = f[0].count; m.values = (float *)malloc(4 * (m.count ? m.count : 1));
        exa_scalar_field_magnitude(f[0].values, f[1].values, f[2].values, m.count, m.values, m.range);
        for (auto &x : f) free(x.values);
        cfg->fields.push_back(m); it += 5; continue;
      }
      if (t == "value_range") {
        need(2);
        if (cfg->fields.empty()) return fail(EXA_ERR_INVALID, "config: value_range before any scalar field");
        cfg->fields.back().range[0] = std::stof(tokens[it + 1]); cfg->fields.back().range[1] = std::stof(tokens[it + 2]);
        it += 3; continue;
      }
      if (t == "bricks") {
        need(1);
        if (cfg->bricks) { exa_bricks_destroy(cfg->bricks); cfg->bricks = nullptr; }
        int rc = exa_bricks_load((basePath + tokens[it + 1]).c_str(), &cfg->bricks);
        if (rc) return fail(rc, "");
        it += 2; continue;
      }
      if (t == "triangles") {                                 // Config.cpp:165-171
        need(1);
        free(cfg->meshVerts); free(cfg->meshIndex); cfg->meshVerts = nullptr; cfg->meshIndex = nullptr;
        int rc = exa_triangle_mesh_load((basePath + tokens[it + 1]).c_str(), &cfg->meshVerts, &cfg->numMeshVerts, &cfg->meshIndex, &cfg->numMeshTris);
        if (rc) return fail(rc, "");
        it += 2; continue;
      }
      return fail(EXA_ERR_INVALID, "error in parsing config file: unknown token '" + t + "'");
    }
  } catch (std::exception &e) { return fail(EXA_ERR_INVALID, e.what()); }
  // Config::finalize, Config.cpp:23-44: voxelSpaceTransform = voxelSys * rcp(worldSys)
  float vs[12], wsys[12], winv[12];
  translateScale(cfg->remapFrom, vs); translateScale(cfg->remapTo, wsys);
  affineInverse(wsys, winv); affineMul(vs, winv, cfg->xfm);
  *out = cfg;
  return EXA_OK;
}

void exa_config_destroy(exa_config_t *c)
{
  if (!c) return;
  exa_bricks_destroy(c->bricks);
  for (auto &f : c->fields) free(f.values);
  free(c->meshVerts); free(c->meshIndex);
  delete c;
}
int exa_config_triangles(const exa_config_t *c, const float **vertices, size_t *numVerts, const int32_t **indices, size_t *numTris)
{
  if (!c || !vertices || !numVerts || !indices || !numTris) { setError("exa_config_triangles: null argument"); return EXA_ERR_INVALID; }
  *vertices = c->meshVerts; *numVerts = c->numMeshVerts; *indices = c->meshIndex; *numTris = c->numMeshTris;
  return EXA_OK;
}
const exa_bricks_t *exa_config_bricks(const exa_config_t *c) { return c ? c->bricks : nullptr; }
int exa_config_num_fields(const exa_config_t *c) { return c ? (int)c->fields.size() : 0; }
const float *exa_config_field(const exa_config_t *c, int i, size_t *count, float valueRange[2], const char **name)
{
  if (!c || i < 0 || i >= (int)c->fields.size()) return nullptr;
  const auto &f = c->fields[i];
  if (count) *count = f.count;
  if (valueRange) { valueRange[0] = f.range[0]; valueRange[1] = f.range[1]; }
  if (name) *name = f.name.c_str();
  return f.values;
}
int exa_config_voxel_space_transform(const exa_config_t *c, float xfm12[12])
{
  if (!c || !xfm12) { setError("null argument"); return EXA_ERR_INVALID; }
  memcpy(xfm12, c->xfm, 48); return EXA_OK;
}
int exa_config_get_bounds(const exa_config_t *c, float lower[3], float upper[3])
{
  if (!c || !c->bricks) { setError("config has no bricks"); return EXA_ERR_INVALID; }
  float lo[3], hi[3], inv[12];
  exa_bricks_get_bounds(c->bricks, lo, hi);
  affineInverse(c->xfm, inv);                                // Config.cpp:48-55
  for (int r = 0; r < 3; r++) {
    lower[r] = inv[r] * lo[0] + inv[3 + r] * lo[1] + inv[6 + r] * lo[2] + inv[9 + r];
    upper[r] = inv[r] * hi[0] + inv[3 + r] * hi[1] + inv[6 + r] * hi[2] + inv[9 + r];
  }
  return EXA_OK;
}

} // extern "C"
