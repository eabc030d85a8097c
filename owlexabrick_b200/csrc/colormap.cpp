// colormap.cpp — the remaining constructors of exa::ColorMapper (exa/ColorMapper.h:29-61, exa/ColorMapper.cpp:92-131)
// behind the C ABI:
//   exa_colormap_parse_string   ColorMapper(std::string): "t,(r,g,b)" lines -> control points, sorted if needed
//   exa_colormap_from_png       ColorMapper(const uint8_t*, size_t): the first pixel row of a PNG -> control points i/(w-1)
//   exa_colormap_embedded       the 15 colour maps the viewer constructs (exa/viewer.cpp:526-541) + its grey "Custom" ramp
// The reference decodes its embedded PNGs with stb_image; here a small decoder of our own (zlib inflate: stored, fixed
// and dynamic Huffman blocks; PNG: 8-bit grey / RGB / palette / grey+alpha / RGBA, non-interlaced, all five filters)
// covers what a one-row colour-map strip needs, and the embedded maps are stored as decoded RGB8 rows
// (colormaps_data.inc, written by tests/golden/make_colormaps.py from the reference's own decode).
#include "exa_internal.h"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace exa {
namespace {

// ---- inflate (RFC 1951) -----------------------------------------------------------------
struct BitReader {
  const uint8_t *p; size_t n, pos = 0; uint32_t buf = 0; int cnt = 0; bool bad = false;
  int bits(int k)
  {
    while (cnt < k) { if (pos >= n) { bad = true; return 0; } buf |= (uint32_t)p[pos++] << cnt; cnt += 8; }
    const int v = (int)(buf & ((1u << k) - 1u)); buf >>= k; cnt -= k; return v;
  }
  void alignByte() { buf = 0; cnt = 0; }
};

struct Huffman {
  uint16_t count[16] = {}, symbol[320] = {};
  bool build(const uint8_t *len, int n)
  {
    memset(count, 0, sizeof(count));
    for (int i = 0; i < n; i++) count[len[i]]++;
    count[0] = 0;
    int left = 1;
    for (int l = 1; l < 16; l++) { left <<= 1; left -= count[l]; if (left < 0) return false; }
    uint16_t offs[16]; offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
    for (int i = 0; i < n; i++) if (len[i]) symbol[offs[len[i]]++] = (uint16_t)i;
    return true;
  }
  int decode(BitReader &br) const
  {
    int code = 0, first = 0, index = 0;
    for (int l = 1; l < 16; l++) {
      code |= br.bits(1);
      if (br.bad) return -1;
      const int c = count[l];
      if (code - c < first) return symbol[index + (code - first)];
      index += c; first += c; first <<= 1; code <<= 1;
    }
    return -1;
  }
};

// stops as soon as `want` bytes are out (the caller needs the first pixel row only: bounds the work and the memory a
// hostile stream can cost)
bool inflateZlib(const uint8_t *src, size_t n, std::vector<uint8_t> &out, size_t want)
{
  if (n < 6 || (src[0] & 0x0f) != 8 || ((src[0] << 8) | src[1]) % 31 != 0 || (src[1] & 0x20)) return false;
  BitReader br{src + 2, n - 2};
  static const uint16_t lenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
  static const uint8_t lenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
  static const uint16_t distBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
  static const uint8_t distExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
  for (;;) {
    const int last = br.bits(1), type = br.bits(2);
    if (br.bad) return false;
    if (type == 0) {
      br.alignByte();
      if (br.pos + 4 > br.n) return false;
      const unsigned len = br.p[br.pos] | (br.p[br.pos + 1] << 8), nlen = br.p[br.pos + 2] | (br.p[br.pos + 3] << 8);
      br.pos += 4;
      if ((len ^ 0xffffu) != nlen || br.pos + len > br.n) return false;
      out.insert(out.end(), br.p + br.pos, br.p + br.pos + len); br.pos += len;
      if (out.size() >= want) return true;
    } else if (type == 1 || type == 2) {
      Huffman lit, dist;
      uint8_t lens[320];
      if (type == 1) {
        for (int i = 0; i < 144; i++) lens[i] = 8;
        for (int i = 144; i < 256; i++) lens[i] = 9;
        for (int i = 256; i < 280; i++) lens[i] = 7;
        for (int i = 280; i < 288; i++) lens[i] = 8;
        lit.build(lens, 288);
        for (int i = 0; i < 30; i++) lens[i] = 5;
        dist.build(lens, 30);
      } else {
        const int nlen = br.bits(5) + 257, ndist = br.bits(5) + 1, ncode = br.bits(4) + 4;
        if (br.bad || nlen > 286 || ndist > 30) return false;
        static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
        uint8_t cl[19] = {};
        for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)br.bits(3);
        Huffman lc; if (!lc.build(cl, 19)) return false;
        int idx = 0;
        while (idx < nlen + ndist) {
          const int sym = lc.decode(br);
          if (sym < 0) return false;
          if (sym < 16) lens[idx++] = (uint8_t)sym;
          else {
            int prev = 0, rep;
            if (sym == 16) { if (!idx) return false; prev = lens[idx - 1]; rep = 3 + br.bits(2); }
            else if (sym == 17) rep = 3 + br.bits(3);
            else rep = 11 + br.bits(7);
            if (br.bad || idx + rep > nlen + ndist) return false;
            while (rep--) lens[idx++] = (uint8_t)prev;
          }
        }
        if (!lit.build(lens, nlen) || !dist.build(lens + nlen, ndist)) return false;
      }
      for (;;) {
        int sym = lit.decode(br);
        if (sym < 0) return false;
        if (sym < 256) { out.push_back((uint8_t)sym); if (out.size() >= want) return true; continue; }
        if (sym == 256) break;
        sym -= 257;
        if (sym >= 29) return false;
        const int len = lenBase[sym] + br.bits(lenExtra[sym]);
        const int ds = dist.decode(br);
        if (ds < 0 || ds >= 30) return false;
        const size_t d = distBase[ds] + (size_t)br.bits(distExtra[ds]);
        if (br.bad || d > out.size()) return false;
        for (int i = 0; i < len; i++) out.push_back(out[out.size() - d]);
        if (out.size() >= want) return true;
      }
    } else return false;
    if (last) break;
  }
  return true;
}

inline uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// first pixel row of a PNG as RGB8 (what ColorMapper's PNG constructor reads: stbi_load_from_memory(.., 3), row 0)
bool pngFirstRowRGB(const uint8_t *buf, size_t n, std::vector<uint8_t> &rgb, int &width, std::string &why)
{
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (n < 8 || memcmp(buf, sig, 8)) { why = "not a PNG"; return false; }
  size_t pos = 8; uint32_t w = 0, h = 0; int depth = 0, ctype = 0, interlace = 0;
  std::vector<uint8_t> idat, palette;
  while (pos + 12 <= n) {
    const uint32_t len = be32(buf + pos); const uint8_t *type = buf + pos + 4, *data = buf + pos + 8;
    if (pos + 12 + (size_t)len > n) { why = "truncated PNG chunk"; return false; }
    if (!memcmp(type, "IHDR", 4)) { if (len < 13) { why = "bad IHDR"; return false; } w = be32(data); h = be32(data + 4); depth = data[8]; ctype = data[9]; interlace = data[12]; }
    else if (!memcmp(type, "PLTE", 4)) palette.assign(data, data + len);
    else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
    else if (!memcmp(type, "IEND", 4)) break;
    pos += 12 + (size_t)len;
  }
  if (!w || !h || depth != 8 || interlace) { why = "unsupported PNG (need 8-bit, non-interlaced)"; return false; }
  const int chans = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
  if (!chans) { why = "unsupported PNG colour type"; return false; }
  if (w > (1u << 20)) { why = "PNG wider than 1 Mi pixels"; return false; }
  const size_t stride = (size_t)w * chans;
  std::vector<uint8_t> raw;
  if (!inflateZlib(idat.data(), idat.size(), raw, stride + 1)) { why = "corrupt PNG data stream"; return false; }
  if (raw.size() < stride + 1) { why = "short PNG data stream"; return false; }
  // un-filter row 0 (the row above it is all zeros)
  std::vector<uint8_t> row(raw.begin() + 1, raw.begin() + 1 + stride);
  const int ft = raw[0];
  for (size_t i = 0; i < stride; i++) {
    const int a = i >= (size_t)chans ? row[i - chans] : 0;
    int add = 0;
    switch (ft) { case 0: case 2: add = 0; break; case 1: add = a; break; case 3: add = a >> 1; break; case 4: add = a; break;   // Paeth with b = c = 0 picks a
                  default: why = "bad PNG filter"; return false; }
    row[i] = (uint8_t)(row[i] + add);
  }
  rgb.resize((size_t)w * 3); width = (int)w;
  for (uint32_t x = 0; x < w; x++) {
    const uint8_t *p = row.data() + (size_t)x * chans;
    uint8_t r, g, b;
    if (ctype == 0 || ctype == 4) r = g = b = p[0];
    else if (ctype == 3) { if ((size_t)p[0] * 3 + 2 >= palette.size()) { why = "PNG palette index out of range"; return false; } r = palette[p[0] * 3]; g = palette[p[0] * 3 + 1]; b = palette[p[0] * 3 + 2]; }
    else { r = p[0]; g = p[1]; b = p[2]; }
    rgb[3 * x] = r; rgb[3 * x + 1] = g; rgb[3 * x + 2] = b;
  }
  return true;
}

float *pointsFromRGB(const uint8_t *rgb, int w)
{
  float *cp = (float *)malloc(sizeof(float) * 4 * (size_t)std::max(w, 1));
  if (!cp) return nullptr;
  for (int i = 0; i < w; i++) {                              // ColorMapper.h:50-56
    cp[4 * i] = i / float(w - 1);
    cp[4 * i + 1] = rgb[3 * i] / 255.f; cp[4 * i + 2] = rgb[3 * i + 1] / 255.f; cp[4 * i + 3] = rgb[3 * i + 2] / 255.f;
  }
  return cp;
}

struct EmbeddedMap { const char *name; const char *symbol; int width; const uint8_t *rgb; };
#include "colormaps_data.inc"

std::string trimmed(const std::string &s, const char *ws)   // trim() of ColorMapper.cpp:22-51
{
  const size_t first = s.find_first_not_of(ws);
  if (first == std::string::npos) return "";
  size_t last = s.find_last_not_of(ws);
  if (last == std::string::npos) last = s.size() - 1;
  if (first > last) return "";
  return s.substr(first, last - first + 1);
}
std::vector<std::string> split(const std::string &s, char delim)   // string_split(): std::getline semantics
{
  std::vector<std::string> out; size_t pos = 0;
  if (s.empty()) return out;
  for (;;) {
    const size_t q = s.find(delim, pos);
    if (q == std::string::npos) { if (pos < s.size()) out.push_back(s.substr(pos)); break; }
    out.push_back(s.substr(pos, q - pos)); pos = q + 1;
  }
  return out;
}

} // namespace
} // namespace exa

using namespace exa;

extern "C" {

int exa_colormap_parse_string(const char *colorMapString, float **controlPoints4, int *numPoints)
{
  if (!colorMapString || !controlPoints4 || !numPoints) { setError("exa_colormap_parse_string: null argument"); return EXA_ERR_INVALID; }
  std::vector<float> cp;
  for (std::string s : split(colorMapString, '\n')) {        // ColorMapper.cpp:92-131
    if (s.empty()) continue;
    s = trimmed(s, " \t");
    std::vector<std::string> reals = split(s, ',');
    if (reals.size() != 4) continue;                          // "Not a valid control point"
    float v[4];
    try {
      for (int i = 0; i < 4; i++) { std::string t = trimmed(trimmed(reals[i], "("), ")"); v[i] = std::stof(t); }
    } catch (const std::exception &) { setError("ColorMapper: cannot parse control point '" + s + "'"); return EXA_ERR_INVALID; }   // the reference lets std::stof throw
    cp.insert(cp.end(), v, v + 4);
  }
  const int n = (int)(cp.size() / 4);
  // std::sort by position when not sorted (stable here: equal positions keep their order)
  std::vector<int> idx(n); for (int i = 0; i < n; i++) idx[i] = i;
  bool sorted = true; for (int i = 1; i < n; i++) if (cp[4 * i] < cp[4 * (i - 1)]) sorted = false;
  if (!sorted) std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return cp[4 * a] < cp[4 * b]; });
  float *out = (float *)malloc(sizeof(float) * 4 * (size_t)std::max(n, 1));
  if (!out) { setError("out of memory"); return EXA_ERR_LIMIT; }
  for (int i = 0; i < n; i++) memcpy(out + 4 * i, cp.data() + 4 * idx[i], 16);
  *controlPoints4 = out; *numPoints = n;
  return EXA_OK;
}

int exa_colormap_from_png(const uint8_t *buf, size_t numBytes, float **controlPoints4, int *numPoints)
{
  if (!buf || !controlPoints4 || !numPoints) { setError("exa_colormap_from_png: null argument"); return EXA_ERR_INVALID; }
  std::vector<uint8_t> rgb; int w = 0; std::string why;
  if (!pngFirstRowRGB(buf, numBytes, rgb, w, why)) { setError("exa_colormap_from_png: " + why); return EXA_ERR_INVALID; }
  float *cp = pointsFromRGB(rgb.data(), w);
  if (!cp) { setError("out of memory"); return EXA_ERR_LIMIT; }
  *controlPoints4 = cp; *numPoints = w;
  return EXA_OK;
}

int exa_colormap_embedded_count(void) { return (int)(sizeof(g_embeddedMaps) / sizeof(g_embeddedMaps[0])) + 1; }

int exa_colormap_embedded(int index, const char **name, float **controlPoints4, int *numPoints)
{
  const int n = exa_colormap_embedded_count();
  if (index < 0 || index >= n || !controlPoints4 || !numPoints) { setError("exa_colormap_embedded: bad argument"); return EXA_ERR_INVALID; }
  if (index == n - 1) {                                       // "Custom": ColorMapper({vec3f(0,0,0), vec3f(1,1,1)}) (viewer.cpp:542)
    float *cp = (float *)malloc(sizeof(float) * 8);
    if (!cp) { setError("out of memory"); return EXA_ERR_LIMIT; }
    const float v[8] = {0.f, 0.f, 0.f, 0.f, 1.f, 1.f, 1.f, 1.f};
    memcpy(cp, v, sizeof(v));
    if (name) *name = "Custom";
    *controlPoints4 = cp; *numPoints = 2;
    return EXA_OK;
  }
  const EmbeddedMap &m = g_embeddedMaps[index];
  float *cp = pointsFromRGB(m.rgb, m.width);
  if (!cp) { setError("out of memory"); return EXA_ERR_LIMIT; }
  if (name) *name = m.name;
  *controlPoints4 = cp; *numPoints = m.width;
  return EXA_OK;
}

} // extern "C"
