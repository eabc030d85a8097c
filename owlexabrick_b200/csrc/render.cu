// render.cu — fused sm_100a kernels of the ExaBricks rendering hot path.
//
//   k_renderFrame   one WARP per ray.  Ordered kd-tree traversal over the ABRs (replaces
//                   optixTrace + __intersection__VolumeBVH, programs/exabrick.cu:184-238,
//                   without per-region restarts), then per region segment the 32 lanes take
//                   32 CONSECUTIVE samples of the ray: tent-basis reconstruction
//                   (exabrick.cu:620-816, 883-942) from brick headers staged in shared
//                   memory, transfer function (:135-150) from a shared-memory table,
//                   opacity correction + gradient shading (:988-1016), and front-to-back
//                   compositing (:1116-1185) as a warp-wide prefix "over", with early ray
//                   termination located exactly (first sample whose running alpha crosses
//                   the threshold).  Implicit iso-surfaces (:1018-1114, 1187-1256,
//                   1408-1460) run as a first pass of the same kernel; accumulation, sRGB
//                   and RGBA8 pack (:1701-1719) as its tail.
//   k_samplePoints  one thread per point: kd point location + reconstruction.
//   k_trace         RK4 streamline advection (exabrick.cu:1531-1574), one thread per seed.
//   k_regionActive / k_nodeActiveUp   space-skipping masks (exabrick.cu:250-331, 373-402):
//                   an XF edit costs O(#regions), no BVH rebuild.
//   k_gather        scalar[i] = field[cellID[i]] (exa/OptixRenderer.cpp:103-132).
//   k_over / k_finish  sort-last compositing helpers.
#include "render.cuh"
#include <cstdio>

namespace exa {

#define FULL 0xffffffffu
static constexpr int WARPS_PER_CTA = 8;
static constexpr int TILE = 8;           // 8x8 pixel tile per CTA, one row per warp
static constexpr int MAXK = 16;          // brick headers staged per chunk
static constexpr int BLOB_UNITS = 3 + 2 * MAXK;

// -----------------------------------------------------------------------------------------
// small device helpers
// -----------------------------------------------------------------------------------------
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float len3(float3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ float3 normalize3(float3 a) { return a * (1.f / sqrtf(dot3(a, a))); }   // owl: v * rsqrt(dot)
// Per-ray set-up (camera ray, voxel-space transform, normalisation) is evaluated "as written"
// in the reference, products and sums rounded separately: it costs a few instructions per
// ray and makes the region entry/exit parameters t0,t1 of boxTest reproducible.
__device__ __forceinline__ float dot3nc(float3 a, float3 b)
{ return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z)); }
__device__ __forceinline__ float len3nc(float3 a) { return sqrtf(dot3nc(a, a)); }
__device__ __forceinline__ float3 normalize3nc(float3 a) { const float s = 1.f / sqrtf(dot3nc(a, a)); return f3(a.x * s, a.y * s, a.z * s); }
// ray.origin + t * ray.direction with the product and the sum rounded separately (no FMA
// contraction): single-sample segments put their sample at the midpoint of two region faces,
// which is exactly a dual-cell boundary of the coarser bricks, where the gradient jumps;
// evaluating the position as written in the reference keeps both sides on the same cell.
__device__ __forceinline__ float3 rayAt(float3 o, float3 d, float t)
{ return f3(__fadd_rn(o.x, __fmul_rn(t, d.x)), __fadd_rn(o.y, __fmul_rn(t, d.y)), __fadd_rn(o.z, __fmul_rn(t, d.z))); }
__device__ __forceinline__ float comp(float3 a, int d) { return d == 0 ? a.x : d == 1 ? a.y : a.z; }

struct Acc { float sw, swv; float3 sd, sdc; };
__device__ __forceinline__ void accZero(Acc &a) { a.sw = 0.f; a.swv = 0.f; a.sd = f3(0, 0, 0); a.sdc = f3(0, 0, 0); }

// addBasisFunctions (exabrick.cu:620-777) for k staged brick headers; `chan` = base of the
// channel's flattened scalars.  Invalid corners contribute weight 0 (== being skipped).
template <bool GRAD, bool STATS>
__device__ __forceinline__ void accumBricks(Acc &a, const uint4 *__restrict__ hdrs, int k, const float *__restrict__ chan,
                                            float px, float py, float pz, unsigned &nCorners)
{
  for (int i = 0; i < k; i++) {
    const uint4 u0 = hdrs[2 * i], u1 = hdrs[2 * i + 1];
    const float invCw = __uint_as_float(u0.w);
    const int sx = (int)u1.x, sy = (int)u1.y, sz = (int)u1.z;
    const float lx = (px - __uint_as_float(u0.x)) * invCw - 0.5f;
    const float ly = (py - __uint_as_float(u0.y)) * invCw - 0.5f;
    const float lz = (pz - __uint_as_float(u0.z)) * invCw - 0.5f;
    const int ix0 = max(-1, (int)floorf(lx)), iy0 = max(-1, (int)floorf(ly)), iz0 = max(-1, (int)floorf(lz));
    const float fx = lx - (float)ix0, fy = ly - (float)iy0, fz = lz - (float)iz0;
    const float gx = 1.f - fx, gy = 1.f - fy, gz = 1.f - fz;
    const bool x0 = ix0 >= 0 && ix0 < sx, x1 = ix0 + 1 < sx;
    const bool y0 = iy0 >= 0 && iy0 < sy, y1 = iy0 + 1 < sy;
    const bool z0 = iz0 >= 0 && iz0 < sz, z1 = iz0 + 1 < sz;
    const float *base = chan + u1.w + ix0 + iy0 * sx + iz0 * (sx * sy);
    const int dy = sx, dz = sx * sy;
    const bool v000 = z0 && y0 && x0, v001 = z0 && y0 && x1, v010 = z0 && y1 && x0, v011 = z0 && y1 && x1;
    const bool v100 = z1 && y0 && x0, v101 = z1 && y0 && x1, v110 = z1 && y1 && x0, v111 = z1 && y1 && x1;
    const float s000 = v000 ? __ldg(base) : 0.f;
    const float s001 = v001 ? __ldg(base + 1) : 0.f;
    const float s010 = v010 ? __ldg(base + dy) : 0.f;
    const float s011 = v011 ? __ldg(base + dy + 1) : 0.f;
    const float s100 = v100 ? __ldg(base + dz) : 0.f;
    const float s101 = v101 ? __ldg(base + dz + 1) : 0.f;
    const float s110 = v110 ? __ldg(base + dz + dy) : 0.f;
    const float s111 = v111 ? __ldg(base + dz + dy + 1) : 0.f;
    if (STATS) nCorners += v000 + v001 + v010 + v011 + v100 + v101 + v110 + v111;
    const float zy00 = gz * gy, zy01 = gz * fy, zy10 = fz * gy, zy11 = fz * fy;
#define EXA_CORNER(V, S, WZY, WX, WZ, WY, SXN, SYN, SZN)                                          \
    { const float w = (V) ? (WZY) * (WX) : 0.f; a.sw += w; a.swv += w * (S);                     \
      if (GRAD) { const float ddx = (V) ? (WZY) * (SXN) : 0.f, ddy = (V) ? ((WZ) * (WX)) * (SYN) : 0.f, \
                  ddz = (V) ? ((WY) * (WX)) * (SZN) : 0.f;                                       \
        a.sdc.x += ddx; a.sdc.y += ddy; a.sdc.z += ddz;                                          \
        a.sd.x += ddx * (S); a.sd.y += ddy * (S); a.sd.z += ddz * (S); } }
    EXA_CORNER(v000, s000, zy00, gx, gz, gy, -1.f, -1.f, -1.f)
    EXA_CORNER(v001, s001, zy00, fx, gz, gy, +1.f, -1.f, -1.f)
    EXA_CORNER(v010, s010, zy01, gx, gz, fy, -1.f, +1.f, -1.f)
    EXA_CORNER(v011, s011, zy01, fx, gz, fy, +1.f, +1.f, -1.f)
    EXA_CORNER(v100, s100, zy10, gx, fz, gy, -1.f, -1.f, +1.f)
    EXA_CORNER(v101, s101, zy10, fx, fz, gy, +1.f, -1.f, +1.f)
    EXA_CORNER(v110, s110, zy11, gx, fz, fy, -1.f, +1.f, +1.f)
    EXA_CORNER(v111, s111, zy11, fx, fz, fy, +1.f, +1.f, +1.f)
#undef EXA_CORNER
  }
}

// samplePoint / samplePointWithDerivative tail (exabrick.cu:800-806, 910-927)
template <bool GRAD>
__device__ __forceinline__ bool accFinish(const Acc &a, float &value, float3 &g)
{
  if (a.sw <= 1e-20f) return false;
  value = a.swv / a.sw;
  if (GRAD) g = f3(a.sw * a.sd.x - a.swv * a.sdc.x, a.sw * a.sd.y - a.swv * a.sdc.y, a.sw * a.sd.z - a.swv * a.sdc.z);
  return true;
}

// lookupTransferFunction (exabrick.cu:135-150).  The reference reads a 128-texel float4
// CUDA texture (normalised coords, clamp, linear filter, exa/Texture.h:143-147); here the
// table sits in shared memory and the lerp is done in fp32.
__device__ __forceinline__ float4 lookupXF(const float4 *__restrict__ table, float lo, float hi, float opacityScale, float in)
{
  float sc = 127.f * (in - lo) / ((hi - lo) + 1e-20f);
  sc = fminf(fmaxf(sc + .5f, 0.f), 127.f);
  sc /= 127.f;
  const float x = sc * 128.f - 0.5f;
  const float fl = floorf(x), w = x - fl;
  const int i0 = max(0, min(127, (int)fl)), i1 = max(0, min(127, (int)fl + 1));
  const float4 a = table[i0], b = table[i1];
  float4 r = make_float4(a.x + w * (b.x - a.x), a.y + w * (b.y - a.y), a.z + w * (b.z - a.z), a.w + w * (b.w - a.w));
  r.w *= opacityScale;
  return r;
}

// boxTest (exabrick.cu:197-210)
__device__ __forceinline__ bool boxTest(float3 o, float3 d, float tmin, float tmax, const float lo[3], const float hi[3], float &t0, float &t1)
{
  const float tlx = (lo[0] - o.x) / d.x, thx = (hi[0] - o.x) / d.x;
  const float tly = (lo[1] - o.y) / d.y, thy = (hi[1] - o.y) / d.y;
  const float tlz = (lo[2] - o.z) / d.z, thz = (hi[2] - o.z) / d.z;
  const float nr = fmaxf(fmaxf(fminf(tlx, thx), fminf(tly, thy)), fminf(tlz, thz));
  const float fr = fminf(fminf(fmaxf(tlx, thx), fmaxf(tly, thy)), fmaxf(tlz, thz));
  t0 = fmaxf(tmin, nr); t1 = fminf(tmax, fr);
  return t0 < t1;
}

// owl::LCG<16> (owl/common/math/random.h:56-96)
struct LCG16 {
  uint32_t state;
  __device__ LCG16(uint32_t v0, uint32_t v1)
  {
    uint32_t s0 = 0;
#pragma unroll
    for (int n = 0; n < 16; n++) {
      s0 += 0x9e3779b9;
      v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4);
      v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761e);
    }
    state = v0;
  }
  __device__ float next() { state = 1664525u * state + 1013904223u; return ldexpf((float)state, -32); }
};

__device__ __forceinline__ float linear_to_srgb(float x) { return x <= 0.0031308f ? 12.92f * x : 1.055f * powf(x, 1.f / 2.4f) - 0.055f; }
__device__ __forceinline__ int make_8bit(float f) { return min(255, max(0, (int)(f * 256.f))); }
__device__ __forceinline__ uint32_t make_rgba8(float3 c)
{ return (uint32_t)(make_8bit(c.x) << 0) + (uint32_t)(make_8bit(c.y) << 8) + (uint32_t)(make_8bit(c.z) << 16) + (0xffu << 24); }

// -----------------------------------------------------------------------------------------
// ordered kd traversal, one warp in lock-step; stack in shared memory
// -----------------------------------------------------------------------------------------
struct Trav {
  uint4 *stack;          // per warp: {node, tn bits, tf bits, -}
  int sp;
  int cur; float tn, tf; // node being descended (-1: pop)
  float3 o, d;
};

__device__ __forceinline__ void travInit(Trav &T, uint4 *stack, float3 o, float3 d, float tmin, float tmax, const DeviceScene &S)
{
  T.stack = stack; T.sp = 0; T.o = o; T.d = d; T.cur = -1;
  float t0, t1;
  if (boxTest(o, d, tmin, tmax, S.boundsLo, S.boundsHi, t0, t1)) {
    // widen a little: the root interval is only used for culling, leaves do the exact test
    T.cur = 0; T.tn = t0 - 1e-5f * fabsf(t0) - 1e-6f; T.tf = t1 + 1e-5f * fabsf(t1) + 1e-6f;
  }
}

// Finds the next region (front-to-back) with t0 = max(tmin, entry) < t1 = min(tmax, exit)
// (exabrick.cu:213-238); on success the region blob (header + up to MAXK brick headers) is
// staged in `blob` (shared memory) and true is returned.
__device__ __forceinline__ bool travNext(Trav &T, const DeviceScene &S, const uint8_t *__restrict__ active,
                                         float tmin, float tmax, uint4 *blob, int lane,
                                         uint32_t &blobOfs, int &K, float &t0, float &t1)
{
  for (;;) {
    if (T.cur < 0) {
      if (T.sp == 0) return false;
      T.sp--;
      const uint4 e = T.stack[T.sp];
      T.cur = (int)e.x; T.tn = __uint_as_float(e.y); T.tf = __uint_as_float(e.z);
    }
    if (T.tf < tmin - 1e-5f * fabsf(tmin) || T.tn > tmax || (active && !active[T.cur])) { T.cur = -1; continue; }
    const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(S.kd) + T.cur);
    const uint32_t axis = raw.y >> 30, payload = raw.y & 0x3fffffffu;
    if (axis == 3) {
      T.cur = -1;
      if (payload == KD_EMPTY) continue;
      K = (int)raw.x;
      const int units = 3 + 2 * min(K, MAXK);
      __syncwarp();
      for (int u = lane; u < units; u += 32) blob[u] = __ldg(S.blobs + payload + u);
      __syncwarp();
      const uint4 h0 = blob[0], h1 = blob[1];
      const float lo[3] = {__uint_as_float(h0.x), __uint_as_float(h0.y), __uint_as_float(h0.z)};
      const float hi[3] = {__uint_as_float(h0.w), __uint_as_float(h1.x), __uint_as_float(h1.y)};
      if (boxTest(T.o, T.d, tmin, tmax, lo, hi, t0, t1)) { blobOfs = payload; return true; }
      continue;
    }
    const float split = __uint_as_float(raw.x);
    const float oa = comp(T.o, (int)axis), da = comp(T.d, (int)axis);
    const float tp = (split - oa) / da;
    const bool belowFirst = (oa < split) || (oa == split && da <= 0.f);
    const int first = (int)payload + (belowFirst ? 0 : 1), second = (int)payload + (belowFirst ? 1 : 0);
    const float eps = 1e-5f * fabsf(tp) + 1e-6f;
    if (tp != tp) {                                  // ray inside the plane: visit both halves
      if (lane == 0) T.stack[T.sp] = make_uint4((uint32_t)second, __float_as_uint(T.tn), __float_as_uint(T.tf), 0);
      T.sp++; __syncwarp(); T.cur = first;
    } else if (tp > T.tf + eps || tp <= 0.f) {
      T.cur = first;
    } else if (tp < T.tn - eps) {
      T.cur = second;
    } else {
      if (lane == 0) T.stack[T.sp] = make_uint4((uint32_t)second, __float_as_uint(tp - eps), __float_as_uint(T.tf), 0);
      T.sp++; __syncwarp();
      T.cur = first; T.tf = tp + eps;
    }
  }
}

// common sample-lattice set-up of integrateBrick / isoIntegrateBrick (exabrick.cu:1141-1144)
__device__ __forceinline__ float firstLatticeT(float t0, float dt, float offs)
{
  const int i0 = (int)ceilf((t0 - dt * offs) / dt);
  float t_i = (offs + (float)i0) * dt;
  while ((t_i - dt) >= t0) t_i = t_i - dt;
  while (t_i < t0) t_i += dt;
  return t_i;
}

// premultiplied "over": a in front of b
__device__ __forceinline__ float4 over(float4 a, float4 b)
{
  const float k = 1.f - a.w;
  return make_float4(a.x + k * b.x, a.y + k * b.y, a.z + k * b.z, a.w + k * b.w);
}

// Reconstruct at `pos` for one channel, all lanes in lock-step over the same region.
template <bool GRAD, bool STATS>
__device__ __forceinline__ bool sampleRegion(const DeviceScene &S, uint4 *blob, uint32_t blobOfs, int K, int channel, int lane,
                                             float3 pos, float &value, float3 &g, unsigned &nCorners)
{
  Acc a; accZero(a);
  const float *chan = S.scalars + (size_t)channel * S.numCells;
  accumBricks<GRAD, STATS>(a, blob + 3, min(K, MAXK), chan, pos.x, pos.y, pos.z, nCorners);
  if (K > MAXK) {               // rare: more bricks than staged; stream the rest through the buffer
    for (int k0 = MAXK; k0 < K; k0 += MAXK) {
      const int kc = min(MAXK, K - k0);
      __syncwarp();
      for (int u = lane; u < 2 * kc; u += 32) blob[3 + u] = __ldg(S.blobs + blobOfs + 3 + 2 * k0 + u);
      __syncwarp();
      accumBricks<GRAD, STATS>(a, blob + 3, kc, chan, pos.x, pos.y, pos.z, nCorners);
    }
    __syncwarp();
    for (int u = lane; u < 2 * MAXK; u += 32) blob[3 + u] = __ldg(S.blobs + blobOfs + 3 + u);
    __syncwarp();
  }
  return accFinish<GRAD>(a, value, g);
}

struct WarpStats { unsigned samples, visits, corners, segments; };

// integrateBrick<GRADIENT_SHADING> (exabrick.cu:1116-1185) for one region segment; lanes take
// consecutive samples.  Returns true if the ray terminated (alpha >= threshold).
template <bool GRAD, bool STATS>
__device__ __forceinline__ bool integrateSegment(const RenderParams &P, const float4 *xfTables, uint4 *blob, uint32_t blobOfs, int K,
                                                 float3 o, float3 d, float t0, float t1, float offs, int lane,
                                                 float4 &pc, WarpStats &ws)
{
  const float fcw = __uint_as_float(blob[1].z);
  const float dt = P.dt * fcw;
  const float tFirst = firstLatticeT(t0, dt, offs);
  if (STATS) ws.segments++;
#ifdef EXA_DEBUG
  if (P.dbg && lane == 0) printf("G seg rid=%d K=%d t0=%.9g t1=%.9g dt=%.9g tFirst=%.9g\n", (int)blob[2].x, K, t0, t1, dt, tFirst);
#endif
  for (int base = 0;; base += 32) {
    const int j = base + lane;
    const float tj = fmaf((float)j, dt, tFirst);
    const float tjm1 = fmaf((float)(j - 1), dt, tFirst);
    const bool act = (j == 0) || (tjm1 < t1);
    const float tNext = fminf(tj, t1);
    const float tLast = (j == 0) ? t0 : tjm1;
    const float tSample = 0.5f * (tNext + tLast);
    const float actualDt = tNext - tLast;
    float4 L = make_float4(0.f, 0.f, 0.f, 0.f);
    if (act) {
      const float3 pos = rayAt(o, d, tSample);
      for (int c = 0; c < P.numChannels; c++) {
        float v = 0.f; float3 g = f3(0, 0, 0); unsigned nc = 0;
        const bool ok = sampleRegion<GRAD, STATS>(P.scene, blob, blobOfs, K, c, lane, pos, v, g, nc);
        if (STATS) { ws.samples++; ws.visits += K; ws.corners += nc; }
        if (ok && actualDt != 0.f) {                       // integrateVolume, exabrick.cu:988-1016
          float4 s = lookupXF(xfTables + c * 128, P.xfDomain[c][0], P.xfDomain[c][1], P.xfOpacityScale, v);
          if (GRAD && len3(g) > (float)(int)fcw * 1e-6f) {
            const float3 l = f3(-d.x, -d.y, -d.z);
            const float scale = fabsf(dot3(l, g)) / sqrtf(dot3(g, g) * dot3(l, l));
            s.x *= scale; s.y *= scale; s.z *= scale;
          }
          s.w = 1.f - powf(1.f - s.w, actualDt);
#ifdef EXA_DEBUG
          if (P.dbg) printf("G   smp j=%d t=%.9g adt=%.9g v=%.9g g=(%.6g %.6g %.6g) rgba=(%.6g %.6g %.6g %.6g)\n", j, tSample, actualDt, v, g.x, g.y, g.z, s.x, s.y, s.z, s.w);
#endif
          const float k = (1.f - L.w) * s.w;
          L.x += k * s.x; L.y += k * s.y; L.z += k * s.z; L.w += k;
        }
      }
    }
    // inclusive prefix "over" across lanes (samples are ordered by lane)
#pragma unroll
    for (int ofs = 1; ofs < 32; ofs <<= 1) {
      float4 prev;
      prev.x = __shfl_up_sync(FULL, L.x, ofs); prev.y = __shfl_up_sync(FULL, L.y, ofs);
      prev.z = __shfl_up_sync(FULL, L.z, ofs); prev.w = __shfl_up_sync(FULL, L.w, ofs);
      if (lane >= ofs) L = over(prev, L);
    }
    const float4 after = over(pc, L);                      // pixel colour after this lane's sample
    const unsigned term = __ballot_sync(FULL, act && after.w >= P.ertThreshold);
    const unsigned actMask = __ballot_sync(FULL, act);
    int src;
    bool done = false;
    if (term) { src = __ffs(term) - 1; done = true; }
    else src = 31 - __clz(actMask);                        // last active lane
    pc.x = __shfl_sync(FULL, after.x, src); pc.y = __shfl_sync(FULL, after.y, src);
    pc.z = __shfl_sync(FULL, after.z, src); pc.w = __shfl_sync(FULL, after.w, src);
    if (done) return true;
    // the reference loop ends after the first sample whose t_next reached t1
    const float t31 = fmaf((float)(base + 31), dt, tFirst);
    if (t31 >= t1) return false;
  }
}

// state of one IsoSurfaceIntegrationFunction (exabrick.cu:1018-1114), kept per channel
struct IsoState { float lastT, lastV; };

// Shade one iso crossing at ray parameter th (IsoSurfaceIntegrationFunction::operator(),
// exabrick.cu:1053-1101); executed by all lanes in lock-step with identical arguments.
template <bool STATS>
__device__ __forceinline__ void shadeIsoHit(const RenderParams &P, const float4 *xfTables, uint4 *blob, uint32_t blobOfs, int K,
                                            float3 o, float3 d, float th, int ii, int lane,
                                            float4 &pc, float &tHit, float3 &hitGrad, WarpStats &ws)
{
  const float3 isopt = rayAt(o, d, th);
  float cv = 0.f; float3 grad = f3(0, 0, 0); unsigned nc2 = 0;
  float4 smp = make_float4(1.f, 0.f, 0.f, 1.f);
  const int ich = P.isoChannel[ii];
  if (P.gradISO) {
    if (sampleRegion<true, STATS>(P.scene, blob, blobOfs, K, ich, lane, isopt, cv, grad, nc2)) {
      smp = lookupXF(xfTables + ich * 128, P.xfDomain[ich][0], P.xfDomain[ich][1], P.xfOpacityScale, cv);
      grad = normalize3(grad);
      if (dot3(grad, d) > 0.f) grad = f3(-grad.x, -grad.y, -grad.z);
    }
  } else {
    if (sampleRegion<false, STATS>(P.scene, blob, blobOfs, K, ich, lane, isopt, cv, grad, nc2))
      smp = lookupXF(xfTables + ich * 128, P.xfDomain[ich][0], P.xfDomain[ich][1], P.xfOpacityScale, cv);
  }
  if (STATS && lane == 0) { ws.samples++; ws.visits += K; ws.corners += nc2; }
  if (P.colormapChannel != 0) {
    cv = 0.f; float3 g2; unsigned nc3 = 0;
    if (sampleRegion<false, STATS>(P.scene, blob, blobOfs, K, P.colormapChannel, lane, isopt, cv, g2, nc3))
      smp = lookupXF(xfTables, P.xfDomain[0][0], P.xfDomain[0][1], P.xfOpacityScale, cv);
    if (STATS && lane == 0) { ws.samples++; ws.visits += K; ws.corners += nc3; }
  }
  smp.w = 1.f;
  if (!isfinite(grad.x) || !isfinite(grad.y) || !isfinite(grad.z)) grad = f3(0, 0, 0);
  if (len3(grad) > 0.f) {
    const float3 l = f3(-d.x, -d.y, -d.z);
    const float scale = .3f + .7f * fabsf(dot3(l, grad)) / sqrtf(dot3(grad, grad));
    smp.x *= scale; smp.y *= scale; smp.z *= scale;
  }
  const float k = (1.f - pc.w) * smp.w;
  pc.x += k * smp.x; pc.y += k * smp.y; pc.z += k * smp.z; pc.w += k;
  tHit = th; hitGrad = grad;
}

// crossing test of one functor call (exabrick.cu:1035-1052): the iso indices are visited in
// order, every match shades and overwrites t_hit; returns first and last matching index
__device__ __forceinline__ bool isoCross(const RenderParams &P, int c, float lastV, float lastT, float v, float tSample,
                                         float &tFirst, int &iFirst, float &tLast, int &iLast)
{
  bool crossed = false;
  if (lastV >= -1e35f) {
#pragma unroll
    for (int i = 0; i < EXA_MAX_ISO_SURFACES; i++) {
      const float isoV = P.isoValue[i];
      if (P.isoEnabled[i] && P.isoChannel[i] == c && ((lastV <= isoV && v >= isoV) || (lastV >= isoV && v <= isoV))) {
        const float d1 = fabsf(lastV - isoV), d2 = fabsf(v - isoV);
        const float w1 = 1.f - d1 / (d1 + d2), w2 = 1.f - d2 / (d1 + d2);
        const float tavg = lastT * w1 + tSample * w2;
        if (!crossed) { tFirst = tavg; iFirst = i; }
        tLast = tavg; iLast = i; crossed = true;
      }
    }
  }
  return crossed;
}

// isoIntegrateBrick (exabrick.cu:1187-1256) for one region segment; lanes take consecutive
// samples.  Colour comes from the first crossing; t_hit / gradient are overwritten by later
// crossings of the segment exactly as in the reference (its functor keeps running after a
// hit).  One primary channel: crossings are found with ballots.  Several channels: the values
// of a 32-sample batch are parked in shared memory and the reference's (sample, channel)
// loop, including its "break once opaque" rule (:1249), is replayed in lock-step.
template <bool STATS>
__device__ __forceinline__ void isoSegment(const RenderParams &P, const float4 *xfTables, uint4 *blob, uint32_t blobOfs, int K,
                                           float3 o, float3 d, float t0, float t1, int lane, IsoState *iso /*smem, per channel*/,
                                           float *scr /*smem scratch, multi-channel only*/,
                                           float4 &pc, float &tHit, float3 &hitGrad, WarpStats &ws)
{
  const float fcw = __uint_as_float(blob[1].z);
  const float dt = P.dt * fcw;
  const float tFirst = firstLatticeT(t0, dt, 0.f);
  const int C = P.numChannels;
  if (STATS) ws.segments++;
  for (int base = 0;; base += 32) {
    const int j = base + lane;
    const float tj = fmaf((float)j, dt, tFirst);
    const float tjm1 = fmaf((float)(j - 1), dt, tFirst);
    const bool act = (j == 0) || (tjm1 < t1);
    const float tNext = fminf(tj, t1);
    const float tLast = (j == 0) ? t0 : tjm1;
    const float tSample = 0.5f * (tNext + tLast);
    const float3 pos = rayAt(o, d, tSample);
    if (C == 1) {
      float v = 0.f; float3 g = f3(0, 0, 0); unsigned nc = 0; bool ok = false;
      if (act) {
        if (P.gradISO) ok = sampleRegion<true, STATS>(P.scene, blob, blobOfs, K, 0, lane, pos, v, g, nc);
        else ok = sampleRegion<false, STATS>(P.scene, blob, blobOfs, K, 0, lane, pos, v, g, nc);
        if (STATS) { ws.samples++; ws.visits += K; ws.corners += nc; }
      }
      // previous VALID sample: nearest lower valid lane, else the state carried in
      const unsigned okMask = __ballot_sync(FULL, ok);
      const unsigned below = okMask & ((1u << lane) - 1u);
      const int prevLane = below ? 31 - __clz(below) : -1;
      const float pv = __shfl_sync(FULL, v, prevLane < 0 ? 0 : prevLane);
      const float pt = __shfl_sync(FULL, tSample, prevLane < 0 ? 0 : prevLane);
      const float lastV = prevLane < 0 ? iso[0].lastV : pv;
      const float lastT = prevLane < 0 ? iso[0].lastT : pt;
      float ta = 0.f, tb = 0.f; int ia = 0, ib = 0;
      const bool crossed = ok && isoCross(P, 0, lastV, lastT, v, tSample, ta, ia, tb, ib);
      const unsigned crossMask = __ballot_sync(FULL, crossed);
      if (crossMask) {
        const int firstLane = __ffs(crossMask) - 1, lastLane = 31 - __clz(crossMask);
        const float tA = __shfl_sync(FULL, ta, firstLane); const int iA = __shfl_sync(FULL, ia, firstLane);
        const float tB = __shfl_sync(FULL, tb, lastLane); const int iB = __shfl_sync(FULL, ib, lastLane);
        shadeIsoHit<STATS>(P, xfTables, blob, blobOfs, K, o, d, tA, iA, lane, pc, tHit, hitGrad, ws);
        if (firstLane != lastLane || iA != iB)
          shadeIsoHit<STATS>(P, xfTables, blob, blobOfs, K, o, d, tB, iB, lane, pc, tHit, hitGrad, ws);
      }
      if (okMask) {
        const int lastOk = 31 - __clz(okMask);
        const float nv = __shfl_sync(FULL, v, lastOk), nt = __shfl_sync(FULL, tSample, lastOk);
        __syncwarp();
        if (lane == 0) { iso[0].lastV = nv; iso[0].lastT = nt; }
        __syncwarp();
      }
    } else {
      // scr: tS[32] | v[C][32] | okMask[C]
      unsigned *okMasks = reinterpret_cast<unsigned *>(scr + 32 + 32 * C);
      __syncwarp();
      scr[lane] = tSample;
      for (int c = 0; c < C; c++) {
        float v = 0.f; float3 g = f3(0, 0, 0); unsigned nc = 0; bool ok = false;
        if (act) {
          if (P.gradISO) ok = sampleRegion<true, STATS>(P.scene, blob, blobOfs, K, c, lane, pos, v, g, nc);
          else ok = sampleRegion<false, STATS>(P.scene, blob, blobOfs, K, c, lane, pos, v, g, nc);
          if (STATS) { ws.samples++; ws.visits += K; ws.corners += nc; }
        }
        scr[32 + 32 * c + lane] = v;
        const unsigned m = __ballot_sync(FULL, ok);
        if (lane == 0) okMasks[c] = m;
      }
      __syncwarp();
      const unsigned actMask = __ballot_sync(FULL, act);
      const int nAct = 32 - __clz(actMask);
      for (int jj = 0; jj < nAct; jj++) {
        const float ts = scr[jj];
        for (int c = 0; c < C; c++) {
          if (!((okMasks[c] >> jj) & 1u)) continue;
          const float v = scr[32 + 32 * c + jj];
          float ta = 0.f, tb = 0.f; int ia = 0, ib = 0;
          if (isoCross(P, c, iso[c].lastV, iso[c].lastT, v, ts, ta, ia, tb, ib)) {
            shadeIsoHit<STATS>(P, xfTables, blob, blobOfs, K, o, d, ta, ia, lane, pc, tHit, hitGrad, ws);
            if (ia != ib) shadeIsoHit<STATS>(P, xfTables, blob, blobOfs, K, o, d, tb, ib, lane, pc, tHit, hitGrad, ws);
          }
          __syncwarp();
          if (lane == 0) { iso[c].lastV = v; iso[c].lastT = ts; }
          __syncwarp();
          if (pc.w >= P.ertThreshold) break;
        }
      }
    }
    const float t31 = fmaf((float)(base + 31), dt, tFirst);
    if (t31 >= t1) return;
  }
}

__device__ __forceinline__ float3 xfmVector(const float *m, float3 v)
{ return f3(dot3nc(v, f3(m[0], m[3], m[6])), dot3nc(v, f3(m[1], m[4], m[7])), dot3nc(v, f3(m[2], m[5], m[8]))); }
__device__ __forceinline__ float3 xfmPoint(const float *m, float3 v)
{ const float3 r = xfmVector(m, v); return f3(__fadd_rn(r.x, m[9]), __fadd_rn(r.y, m[10]), __fadd_rn(r.z, m[11])); }

// -----------------------------------------------------------------------------------------
// __raygen__renderFrame (exabrick.cu:1576-1720): one warp per pixel
// -----------------------------------------------------------------------------------------
template <bool GRAD, bool STATS>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32) k_renderFrame(const __grid_constant__ RenderParams P)
{
  extern __shared__ uint4 smem[];
  // layout: [xf tables: numChannels*128 float4][per warp: stack kdStack][per warp: blob BLOB_UNITS][per warp: iso state 10*8B=5 units]
  float4 *xfTables = reinterpret_cast<float4 *>(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint4 *warpBase = smem + P.numChannels * 128 + warp * P.warpUnits;
  uint4 *stack = warpBase, *blob = warpBase + P.kdStack;
  IsoState *iso = reinterpret_cast<IsoState *>(warpBase + P.kdStack + BLOB_UNITS);
  float *isoScr = reinterpret_cast<float *>(warpBase + P.kdStack + BLOB_UNITS + 5);
  for (int i = threadIdx.x; i < P.numChannels * 128; i += blockDim.x) xfTables[i] = P.xf[i];
  __syncthreads();

  const int tilesX = (P.width + TILE - 1) / TILE;
  const int tileId = blockIdx.x;
  if (P.tileMode && (tileId % P.numRanks) != P.rank) return;
  const int tx = (tileId % tilesX) * TILE, ty = (tileId / tilesX) * TILE;
  const int py = ty + warp;
  WarpStats ws = {0, 0, 0, 0};
  if (py < P.height) {
    for (int px = tx; px < tx + TILE && px < P.width; px++) {
      const int pixelIdx = px + P.width * py;
#ifdef EXA_DEBUG
      if (P.dbg && pixelIdx != P.dbg - 1) continue;
#endif
      LCG16 rnd((uint32_t)(P.frameID * P.width * P.height) + (uint32_t)px, (uint32_t)py);
      const float jx = rnd.next(), jy = rnd.next();
      const float sx = (float)px + jx, sy = (float)py + jy;
      // Camera::generateRay (programs/Camera.h:27-44)
      float3 o = f3(P.camPos[0], P.camPos[1], P.camPos[2]);
      float3 d = normalize3nc(f3(__fadd_rn(__fadd_rn(P.dir00[0], __fmul_rn(sx, P.dirDu[0])), __fmul_rn(sy, P.dirDv[0])),
                                 __fadd_rn(__fadd_rn(P.dir00[1], __fmul_rn(sx, P.dirDu[1])), __fmul_rn(sy, P.dirDv[1])),
                                 __fadd_rn(__fadd_rn(P.dir00[2], __fmul_rn(sx, P.dirDu[2])), __fmul_rn(sy, P.dirDv[2]))));
      float tmin = 1e-6f, tmax = 1e8f;
      // ---- surfaces: implicit iso-surfaces (traceSurfaces :1514-1527, traceIsoRay :1408-1460)
      int surfPrim = -1; float surfT = tmax; float3 surfNg = f3(0, 0, 0), surfBase = f3(0, 0, 0); float surfAmbient = 0.f;
      if (P.anyIso) {
        const float3 vo = xfmPoint(P.xfm, o); float3 vd = xfmVector(P.xfm, d);
        const float dtScale = len3nc(vd); vd = normalize3nc(vd);
        float already = dtScale * tmin; float rtmax = tmax;
        __syncwarp();
        if (lane < EXA_MAX_CHANNELS) { iso[lane].lastV = -1e36f; iso[lane].lastT = 0.f; }
        __syncwarp();
        Trav T; rtmax = rtmax * dtScale;
        travInit(T, stack, vo, vd, already, rtmax, P.scene);
        for (;;) {
          uint32_t bo; int K; float t0, t1;
          if (!travNext(T, P.scene, P.isoActive, already, rtmax, blob, lane, bo, K, t0, t1)) break;
          float4 ipc = make_float4(0, 0, 0, 0); float th = -1.f; float3 hg = f3(0, 0, 0);
          isoSegment<STATS>(P, xfTables, blob, bo, K, vo, vd, fmaxf(already, t0), fminf(rtmax, t1), lane, iso, isoScr, ipc, th, hg, ws);
          if (th >= 0.f) {
            surfPrim = -23; surfT = th / dtScale; surfNg = normalize3(hg); surfAmbient = 0.f; surfBase = f3(ipc.x, ipc.y, ipc.z);
            break;
          }
          already = t1 * 1.0000001f;
          rtmax = rtmax * dtScale;                          // (sic) exabrick.cu:1434
        }
      }
      float3 bg = f3(0, 0, 0);
      if (surfPrim == -23) {                                // :1604-1652 with AO disabled
        const bool shade = P.gradISO != 0;
        if (shade && len3(surfNg) > 0.f) {
          const float k = fabsf(dot3(d, surfNg));
          bg = f3(surfAmbient + surfBase.x * k, surfAmbient + surfBase.y * k, surfAmbient + surfBase.z * k);
        } else bg = surfBase;
      }
      // ---- DVR (:1654-1699)
      float4 pc = make_float4(0, 0, 0, 0);
      const float offs = rnd.next();
      tmax = surfT;
      if (P.clipEnabled) { float a, b; boxTest(o, d, tmin, tmax, P.clipLo, P.clipHi, a, b); tmin = a; tmax = b; }
      const float tHitWorld = tmax;
      o = xfmPoint(P.xfm, o); d = xfmVector(P.xfm, d);
      const float dtScale = len3nc(d); d = normalize3nc(d);
      float already = dtScale * tmin;
      const float rayTmax = tHitWorld * dtScale;
      Trav T; travInit(T, stack, o, d, already, rayTmax, P.scene);
      for (;;) {
        uint32_t bo; int K; float t0, t1;
        if (!travNext(T, P.scene, P.volActive, already, rayTmax, blob, lane, bo, K, t0, t1)) break;
        const bool term = integrateSegment<GRAD, STATS>(P, xfTables, blob, bo, K, o, d, t0, t1, offs, lane, pc, ws);
        if (term || pc.w >= P.ertThreshold) { pc = make_float4(pc.x * pc.w, pc.y * pc.w, pc.z * pc.w, 1.f); break; }
        already = t1 * 1.0000001f;
      }
      if (lane == 0) {
        if (P.partialOut) P.partialOut[pixelIdx] = pc;
        else {
          float3 color = f3(pc.w * pc.x + (1.f - pc.w) * bg.x, pc.w * pc.y + (1.f - pc.w) * bg.y, pc.w * pc.z + (1.f - pc.w) * bg.z);   // :1701 (sic)
          if (P.frameID > 0) { const float4 acc = P.accum[pixelIdx]; color = color + f3(acc.x, acc.y, acc.z); }
          P.accum[pixelIdx] = make_float4(color.x, color.y, color.z, 1.f);
          const float inv = (float)P.frameID + 1.f;
          color = f3(linear_to_srgb(color.x / inv), linear_to_srgb(color.y / inv), linear_to_srgb(color.z / inv));
          P.rgba8[pixelIdx] = make_rgba8(color);
        }
      }
    }
  }
  if (STATS) {
    // per-lane counters: samples/visits/corners are per lane, segments per warp (lane 0)
    unsigned long long a = ws.samples, b = ws.visits, c = ws.corners;
    for (int o2 = 16; o2 > 0; o2 >>= 1) { a += __shfl_xor_sync(FULL, a, o2); b += __shfl_xor_sync(FULL, b, o2); c += __shfl_xor_sync(FULL, c, o2); }
    if (lane == 0 && P.stats) { atomicAdd(&P.stats[0], a); atomicAdd(&P.stats[1], b); atomicAdd(&P.stats[2], c); atomicAdd(&P.stats[3], (unsigned long long)ws.segments); }
  }
}

int launchRenderFrame(const RenderParams &Pin, cudaStream_t s)
{
  RenderParams P = Pin;
  // per-warp shared memory: kd stack | region blob | iso state | multi-channel iso scratch
  const int scrFloats = (P.anyIso && P.numChannels > 1) ? 32 + 33 * P.numChannels : 0;
  P.warpUnits = P.kdStack + BLOB_UNITS + 5 + (scrFloats + 3) / 4;
  const int tilesX = (P.width + TILE - 1) / TILE, tilesY = (P.height + TILE - 1) / TILE;
  const size_t smemBytes = 16 * ((size_t)P.numChannels * 128 + (size_t)WARPS_PER_CTA * P.warpUnits);
  const dim3 grid(tilesX * tilesY), block(WARPS_PER_CTA * 32);
  const bool stats = P.stats != nullptr;
#define LAUNCH(G, S_)                                                                                   \
  { cudaFuncSetAttribute(k_renderFrame<G, S_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes); \
    k_renderFrame<G, S_><<<grid, block, smemBytes, s>>>(P); }
  if (P.gradDVR) { if (stats) LAUNCH(true, true) else LAUNCH(true, false) }
  else { if (stats) LAUNCH(false, true) else LAUNCH(false, false) }
#undef LAUNCH
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// point location + reconstruction at arbitrary points (one thread per point)
// -----------------------------------------------------------------------------------------
__device__ __forceinline__ bool locate(const DeviceScene &S, float px, float py, float pz, uint32_t &blobOfs, int &K)
{
  if (!(px >= S.boundsLo[0] && px <= S.boundsHi[0] && py >= S.boundsLo[1] && py <= S.boundsHi[1] && pz >= S.boundsLo[2] && pz <= S.boundsHi[2]))
    return false;
  int node = 0;
  for (;;) {
    const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(S.kd) + node);
    const uint32_t axis = raw.y >> 30, payload = raw.y & 0x3fffffffu;
    if (axis == 3) {
      if (payload == KD_EMPTY) return false;
      blobOfs = payload; K = (int)raw.x;
      // regions are half-open [lo,hi): only the outer upper faces need the explicit check
      const uint4 h0 = __ldg(S.blobs + payload), h1 = __ldg(S.blobs + payload + 1);
      return px < __uint_as_float(h0.w) && py < __uint_as_float(h1.x) && pz < __uint_as_float(h1.y);
    }
    const float p = axis == 0 ? px : axis == 1 ? py : pz;
    node = (int)payload + (p < __uint_as_float(raw.x) ? 0 : 1);
  }
}

template <bool GRAD, bool STATS>
__device__ __forceinline__ bool samplePointGlobal(const DeviceScene &S, float px, float py, float pz, int channel,
                                                  float &value, float3 &g, unsigned &visits, unsigned &corners)
{
  uint32_t bo; int K;
  if (!locate(S, px, py, pz, bo, K)) return false;
  Acc a; accZero(a);
  accumBricks<GRAD, STATS>(a, S.blobs + bo + 3, K, S.scalars + (size_t)channel * S.numCells, px, py, pz, corners);
  if (STATS) visits += K;
  return accFinish<GRAD>(a, value, g);
}

template <bool GRAD, bool STATS>
__global__ void __launch_bounds__(256) k_samplePoints(const __grid_constant__ SampleParams P)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned visits = 0, corners = 0; bool counted = false;
  if (i < P.n) {
    const float px = P.xyz[3 * i], py = P.xyz[3 * i + 1], pz = P.xyz[3 * i + 2];
    float v = 0.f; float3 g = f3(0, 0, 0);
    const bool ok = samplePointGlobal<GRAD, STATS>(P.scene, px, py, pz, P.channel, v, g, visits, corners);
    counted = visits > 0;
    if (P.values) P.values[i] = ok ? v : 0.f;
    if (P.valid) P.valid[i] = ok ? 1 : 0;
    if (GRAD && P.grad) { P.grad[3 * i] = ok ? g.x : 0.f; P.grad[3 * i + 1] = ok ? g.y : 0.f; P.grad[3 * i + 2] = ok ? g.z : 0.f; }
  }
  if (STATS && P.stats) {
    unsigned long long a = counted ? 1 : 0, b = visits, c = corners;
    for (int o2 = 16; o2 > 0; o2 >>= 1) { a += __shfl_xor_sync(FULL, a, o2); b += __shfl_xor_sync(FULL, b, o2); c += __shfl_xor_sync(FULL, c, o2); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&P.stats[0], a); atomicAdd(&P.stats[1], b); atomicAdd(&P.stats[2], c); atomicAdd(&P.stats[3], a); }
  }
}

int launchSamplePoints(const SampleParams &P, cudaStream_t s)
{
  if (P.n == 0) return 0;
  const unsigned grid = (unsigned)((P.n + 255) / 256);
  const bool stats = P.stats != nullptr;
  if (P.grad) { if (stats) k_samplePoints<true, true><<<grid, 256, 0, s>>>(P); else k_samplePoints<true, false><<<grid, 256, 0, s>>>(P); }
  else { if (stats) k_samplePoints<false, true><<<grid, 256, 0, s>>>(P); else k_samplePoints<false, false><<<grid, 256, 0, s>>>(P); }
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// streamlines: computeTraces / sampleDirection (exabrick.cu:945-963, 1531-1574)
// -----------------------------------------------------------------------------------------
__device__ __forceinline__ bool sampleDirection(const DeviceScene &S, float3 p, const int ch[3], float3 &out)
{
  // a point query that misses every region is an invalid sample (the reference indexes
  // region -1 there: undefined behaviour that is not copied)
  float v[3]; float3 g; unsigned a = 0, b = 0;
  for (int i = 0; i < 3; i++) if (!samplePointGlobal<false, false>(S, p.x, p.y, p.z, ch[i], v[i], g, a, b)) return false;
  out = f3(v[0], v[1], v[2]); return true;
}

__global__ void k_trace(const __grid_constant__ TraceParams P)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.numTraces) return;
  for (int t = P.firstStep; t <= P.lastStep && t < P.numTimesteps; t++) {
    float *src = P.traces + 3 * ((size_t)i * P.numTimesteps + (t - 1)), *dst = src + 3;
    float3 p = f3(src[0], src[1], src[2]); const float3 pp = p;
    if (p.x < 2e10f) {
      bool valid = true; float3 k1 = f3(0, 0, 0), k2 = k1, k3 = k1, k4 = k1;
      valid &= sampleDirection(P.scene, p, P.channels, k1); k1 = k1 * P.steplen; const float3 p1 = p + k1 * .5f;
      valid &= sampleDirection(P.scene, p1, P.channels, k2); k2 = k2 * P.steplen; const float3 p2 = p + k2 * .5f;
      valid &= sampleDirection(P.scene, p2, P.channels, k3); k3 = k3 * P.steplen; const float3 p3 = p + k3;
      valid &= sampleDirection(P.scene, p3, P.channels, k4); k4 = k4 * P.steplen;
      p = p + (1 / 6.f) * (k1 + 2.f * k2 + 2.f * k3 + k4);
      const bool inside = p.x >= P.worldLo[0] && p.x <= P.worldHi[0] && p.y >= P.worldLo[1] && p.y <= P.worldHi[1] && p.z >= P.worldLo[2] && p.z <= P.worldHi[2];
      if (!valid || !inside || len3(p - pp) < 1e-10f) p = f3(2e10f, 2e10f, 2e10f);
    }
    dst[0] = p.x; dst[1] = p.y; dst[2] = p.z;
  }
}

int launchTrace(const TraceParams &P, cudaStream_t s)
{
  if (P.numTraces <= 0) return 0;
  k_trace<<<(P.numTraces + 63) / 64, 64, 0, s>>>(P);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// activity masks (space skipping): bounds programs exabrick.cu:285-331 (volume), 373-402 (iso)
// -----------------------------------------------------------------------------------------
__device__ bool activeForVolumeSampling(const RenderParams &P, float vlo, float vhi, int ch)
{
  const float dlo = P.xfDomain[ch][0], dhi = P.xfDomain[ch][1];
  if (vlo > dhi) return false;
  if (vhi < dlo) return false;
  const float slo = (vlo - dlo) / ((dhi - dlo) + 1e-20f), shi = (vhi - dlo) / ((dhi - dlo) + 1e-20f);
  const int ilo = min(127, max(0, (int)(slo * 127))), ihi = min(127, max(0, (int)(shi * 127) + 1));
  for (int i = ilo; i <= ihi; i++) {
    float cv = (float)i / 127; cv *= dhi - dlo; cv += dlo;
    if (lookupXF(P.xf + ch * 128, dlo, dhi, P.xfOpacityScale, cv).w > 0.f) return true;
  }
  return false;
}

__global__ void k_regionActive(const __grid_constant__ RenderParams P, int spaceSkipping, uint8_t *regionVol, uint8_t *regionIso)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.scene.numRegions) return;
  const ExaBrickRegion r = P.scene.regions[i];
  bool act = false;
  for (int c = 0; c < P.numChannels && !act; c++) act |= activeForVolumeSampling(P, r.valueRange_lower, r.valueRange_upper, c);
  regionVol[i] = spaceSkipping ? (act ? 1 : 0) : 1;
  bool ia = false;
  for (int k = 0; k < EXA_MAX_ISO_SURFACES; k++)
    if (P.isoEnabled[k] && P.isoValue[k] >= r.valueRange_lower && P.isoValue[k] <= r.valueRange_upper) ia = true;
  regionIso[i] = ia ? 1 : 0;
}

__global__ void k_nodeActiveUp(int first, int n, const KdNode *kd, const uint4 *blobs, const uint8_t *inPart,
                               const uint8_t *regionVol, const uint8_t *regionIso, uint8_t *volActive, uint8_t *isoActive)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int g = first + i; const KdNode nd = kd[g];
  const uint32_t payload = nd.info & 0x3fffffffu;
  if ((nd.info >> 30) == 3) {
    if (payload == KD_EMPTY) { volActive[g] = 0; isoActive[g] = 0; return; }
    const int rid = (int)blobs[payload + 2].x;
    const bool in = inPart ? inPart[g] != 0 : true;
    volActive[g] = in && regionVol[rid]; isoActive[g] = in && regionIso[rid];
  } else {
    volActive[g] = volActive[payload] | volActive[payload + 1];
    isoActive[g] = isoActive[payload] | isoActive[payload + 1];
  }
}

int launchRegionActivity(const RenderParams &P, const std::vector<int> &levelOffset, int spaceSkipping, const uint8_t *inPart,
                         uint8_t *regionVol, uint8_t *regionIso, uint8_t *volActive, uint8_t *isoActive, cudaStream_t s)
{
  int launches = 0;
  if (P.scene.numRegions > 0) { k_regionActive<<<(P.scene.numRegions + 127) / 128, 128, 0, s>>>(P, spaceSkipping, regionVol, regionIso); launches++; }
  for (int lv = (int)levelOffset.size() - 2; lv >= 0; lv--) {
    const int first = levelOffset[lv], n = levelOffset[lv + 1] - first;
    if (n <= 0) continue;
    k_nodeActiveUp<<<(n + 255) / 256, 256, 0, s>>>(first, n, P.scene.kd, P.scene.blobs, inPart, regionVol, regionIso, volActive, isoActive); launches++;
  }
  return cudaGetLastError() == cudaSuccess ? launches : -1;
}

__global__ void k_partDown(int first, int n, const KdNode *kd, uint8_t *inPart)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int g = first + i; const KdNode nd = kd[g];
  if ((nd.info >> 30) == 3) return;
  const uint32_t c = nd.info & 0x3fffffffu;
  if (inPart[g]) { inPart[c] = 1; inPart[c + 1] = 1; }
}

int launchPartitionMask(const KdNode *kd, const std::vector<int> &levelOffset, int partRoot, uint8_t *inPart, cudaStream_t s)
{
  const int numNodes = levelOffset.back();
  cudaMemsetAsync(inPart, 0, numNodes, s);
  if (partRoot < 0) return 0;
  cudaMemsetAsync(inPart + partRoot, 1, 1, s);
  int launches = 0;
  for (size_t lv = 0; lv + 1 < levelOffset.size(); lv++) {
    const int first = levelOffset[lv], n = levelOffset[lv + 1] - first;
    if (n <= 0) continue;
    k_partDown<<<(n + 255) / 256, 256, 0, s>>>(first, n, kd, inPart); launches++;
  }
  return cudaGetLastError() == cudaSuccess ? launches : -1;
}

// -----------------------------------------------------------------------------------------
// scalar flatten: scalar[i] = field[cellID[i]] (exa/OptixRenderer.cpp:103-132)
// -----------------------------------------------------------------------------------------
__global__ void k_gather(const float *field, size_t fieldCount, const int32_t *cellIDs, size_t n, float *out, int *err)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int id = cellIDs[i];
  if (id < 0) { atomicExch(err, 1); return; }                // "negative cell ID"
  if ((size_t)id >= fieldCount) { atomicExch(err, 2); return; }   // "invalid cell ID"
  out[i] = __ldg(field + id);
}

int launchGather(const float *d_field, size_t fieldCount, const int32_t *d_cellIDs, size_t numCells, float *d_out, int *d_err, cudaStream_t s)
{
  if (numCells == 0) return 0;
  k_gather<<<(unsigned)((numCells + 255) / 256), 256, 0, s>>>(d_field, fieldCount, d_cellIDs, numCells, d_out, d_err);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// sort-last helpers
// -----------------------------------------------------------------------------------------
__global__ void k_over(float4 *mine, const float4 *other, size_t n, int mineIsFront, float thr)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 a = mineIsFront ? mine[i] : other[i], b = mineIsFront ? other[i] : mine[i];
  // a partial that already hit the termination threshold is opaque (alpha was forced to 1)
  mine[i] = over(a, b);
  (void)thr;
}
int launchCompositeOver(float4 *mine, const float4 *other, size_t count, int mineIsFront, float thr, cudaStream_t s)
{
  if (!count) return 0;
  k_over<<<(unsigned)((count + 255) / 256), 256, 0, s>>>(mine, other, count, mineIsFront, thr);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

__global__ void k_finish(const __grid_constant__ RenderParams P, const float4 *comp, size_t first, size_t n, uint32_t *out)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 pc = comp[i];
  const size_t pixelIdx = first + i;
  float3 color = f3(pc.w * pc.x, pc.w * pc.y, pc.w * pc.z);   // :1701 with bg = 0
  if (P.frameID > 0) { const float4 acc = P.accum[pixelIdx]; color = color + f3(acc.x, acc.y, acc.z); }
  P.accum[pixelIdx] = make_float4(color.x, color.y, color.z, 1.f);
  const float inv = (float)P.frameID + 1.f;
  out[i] = make_rgba8(f3(linear_to_srgb(color.x / inv), linear_to_srgb(color.y / inv), linear_to_srgb(color.z / inv)));
}
int launchFinishPixels(const RenderParams &P, const float4 *composited, size_t first, size_t count, uint32_t *rgba8Out, cudaStream_t s)
{
  if (!count) return 0;
  k_finish<<<(unsigned)((count + 255) / 256), 256, 0, s>>>(P, composited, first, count, rgba8Out);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

} // namespace exa
