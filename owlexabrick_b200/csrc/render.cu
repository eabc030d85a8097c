// render.cu — fused sm_100a kernels of the ExaBricks rendering hot path.
//
//   k_renderFrame   one thread per ray, 8x4 pixel tile per warp.  Ordered kd-tree traversal
//                   over the ABRs with a per-thread stack in shared memory (replaces optixTrace
//                   + __intersection__VolumeBVH, programs/exabrick.cu:184-238, without the
//                   per-region restarts), then per region segment: tent-basis reconstruction
//                   (exabrick.cu:620-816, 883-942) straight from the region blob (header +
//                   brick headers, one contiguous 16-byte-aligned record per region),
//                   transfer function (:135-150) from a shared-memory table, opacity
//                   correction + gradient shading (:988-1016), front-to-back compositing with
//                   early ray termination (:1116-1185).  Implicit iso-surfaces (:1018-1114,
//                   1187-1256, 1408-1460) run as a first pass of the same kernel;
//                   accumulation, sRGB and RGBA8 pack (:1701-1719) are its tail.
//   k_samplePoints  one thread per point: kd point location + reconstruction.
//   k_trace         RK4 streamline advection (exabrick.cu:1531-1574), one thread per seed.
//   k_regionActive / k_nodeActiveUp   space-skipping masks (exabrick.cu:250-331, 373-402):
//                   an XF edit costs O(#regions), no BVH rebuild.
//   k_gather        scalar[i] = field[cellID[i]] (exa/OptixRenderer.cpp:103-132).
//   k_over / k_finish  sort-last compositing helpers.
#include "render.cuh"
#include "surfaces.cuh"
#include <algorithm>
#include <cstdio>

namespace exa {

#define FULL 0xffffffffu

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

// One aligned 256-bit read-only load (PTX ld.global.nc.v8.f32 -> SASS LDG.E.ENL2.256, new on
// sm_100): a 32-byte brick header or dual-cell record costs ONE L1 tag look-up per thread.
struct __align__(32) U8 { uint4 lo, hi; };
// L2N: with the .L2::64B qualifier (SASS LDG.E.ENL2.LTC64B.256).  A 256-bit load that misses makes the L2 fetch the whole
// 128-byte line from DRAM; for the random records of a point query three quarters of that line are never used, and
// .L2::64B halves the fetch (ncu: 5.36 M -> 3.28 M DRAM sectors per 1 M points, same 48.8 us).  The frame kernels keep
// the default: their neighbouring samples do use the rest of the line (0.5-1 % slower with the qualifier).
template <bool L2N = false>
__device__ __forceinline__ U8 ldg256(const void *p)
{
  U8 r;
  if (L2N)
    asm volatile("ld.global.nc.L2::64B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w) : "l"(p));
  else
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w) : "l"(p));
  return r;
}

struct Acc { float sw, swv; float3 sd, sdc; };
// base of channel c: flat scalars (reference layout) or dual-cell records (DeviceScene::dual)
template <bool DUAL>
__device__ __forceinline__ const void *chanBase(const DeviceScene &S, int c)
{
  if (DUAL) return S.dual + 2 * (size_t)c * S.dualRecords;
  return S.scalars + (size_t)c * S.numCells;
}
__device__ __forceinline__ void accZero(Acc &a) { a.sw = 0.f; a.swv = 0.f; a.sd = f3(0, 0, 0); a.sdc = f3(0, 0, 0); }

// addBasisFunctions (exabrick.cu:620-777) for k brick headers; `chan` = base of the channel's
// flattened scalars.  A corner is valid iff it is valid on each axis, so its tent weight
// (wz*wy)*wx factorises into per-axis weights that are zeroed on the invalid side: the eight
// corner terms of the reference collapse into three nested lerps (x, then y, then z), the
// weight sum into a product of three per-axis sums, and the derivative terms
//   sumDC.x = sum_valid (wz*wy)*sx      sumD.x = sum_valid (wz*wy)*sx*s      (sx = -1 low, +1 high)
// into differences of the same partial sums.  Same real-number result as the reference's
// corner loop, different fp32 rounding order (gate: 1e-5 relative).
template <bool GRAD, bool STATS, bool DUAL, bool L2N = false>
__device__ __forceinline__ void accumBricks(Acc &a, const U8 &hdr0, const uint4 *__restrict__ hdrs, int k, const void *__restrict__ chanV,
                                            float px, float py, float pz, unsigned &nCorners)
{
  for (int i = 0; i < k; i++) {
    // the first brick's header travels with the region segment (registers): for the common
    // one-brick region a sample has ONE dependent load (the record), not header -> record
    const U8 hdr = i == 0 ? hdr0 : ldg256<L2N>(hdrs + 2 * i);
    const uint4 u0 = hdr.lo, u1 = hdr.hi;
    const float invCw = __uint_as_float(u0.w);
    const int sx = (int)u1.x, sy = (int)u1.y, sz = (int)u1.z;
    const float lx = (px - __uint_as_float(u0.x)) * invCw - 0.5f;
    const float ly = (py - __uint_as_float(u0.y)) * invCw - 0.5f;
    const float lz = (pz - __uint_as_float(u0.z)) * invCw - 0.5f;
    const int ix0 = max(-1, (int)floorf(lx)), iy0 = max(-1, (int)floorf(ly)), iz0 = max(-1, (int)floorf(lz));
    const float fx = lx - (float)ix0, fy = ly - (float)iy0, fz = lz - (float)iz0;
    const bool x0 = ix0 >= 0 && ix0 < sx, x1 = ix0 + 1 < sx;
    const bool y0 = iy0 >= 0 && iy0 < sy, y1 = iy0 + 1 < sy;
    const bool z0 = iz0 >= 0 && iz0 < sz, z1 = iz0 + 1 < sz;
    float s000, s001, s010, s011, s100, s101, s110, s111;
    if (DUAL) {
      // dual-cell records: the 8 corner scalars of dual cell (ix0,iy0,iz0) lie in ONE 32-byte
      // sector (zeros where the corner is outside the brick); u1.w = first record of the brick
      float4 lo4 = make_float4(0, 0, 0, 0), hi4 = lo4;
      if (ix0 < sx && iy0 < sy && iz0 < sz) {
        const int rx = sx + 1, ry = sy + 1;
        const float4 *rec = reinterpret_cast<const float4 *>(chanV) +
                            2 * ((size_t)u1.w + (size_t)((ix0 + 1) + (iy0 + 1) * rx + (iz0 + 1) * (rx * ry)));
        const U8 q = ldg256<L2N>(rec);
        lo4 = make_float4(__uint_as_float(q.lo.x), __uint_as_float(q.lo.y), __uint_as_float(q.lo.z), __uint_as_float(q.lo.w));
        hi4 = make_float4(__uint_as_float(q.hi.x), __uint_as_float(q.hi.y), __uint_as_float(q.hi.z), __uint_as_float(q.hi.w));
      }
      s000 = lo4.x; s001 = lo4.y; s010 = lo4.z; s011 = lo4.w; s100 = hi4.x; s101 = hi4.y; s110 = hi4.z; s111 = hi4.w;
    } else {
      const int dy = sx, dz = sx * sy;
      const float *base = reinterpret_cast<const float *>(chanV) + u1.w + ix0 + iy0 * dy + iz0 * dz;
      const bool r00 = z0 && y0, r01 = z0 && y1, r10 = z1 && y0, r11 = z1 && y1;    // (z,y) rows
      s000 = (r00 && x0) ? __ldg(base) : 0.f;
      s001 = (r00 && x1) ? __ldg(base + 1) : 0.f;
      s010 = (r01 && x0) ? __ldg(base + dy) : 0.f;
      s011 = (r01 && x1) ? __ldg(base + dy + 1) : 0.f;
      s100 = (r10 && x0) ? __ldg(base + dz) : 0.f;
      s101 = (r10 && x1) ? __ldg(base + dz + 1) : 0.f;
      s110 = (r11 && x0) ? __ldg(base + dz + dy) : 0.f;
      s111 = (r11 && x1) ? __ldg(base + dz + dy + 1) : 0.f;
    }
    if (STATS) nCorners += ((int)z0 + (int)z1) * ((int)y0 + (int)y1) * ((int)x0 + (int)x1);
    const float wx0 = x0 ? 1.f - fx : 0.f, wx1 = x1 ? fx : 0.f;
    const float wy0 = y0 ? 1.f - fy : 0.f, wy1 = y1 ? fy : 0.f;
    const float wz0 = z0 ? 1.f - fz : 0.f, wz1 = z1 ? fz : 0.f;
    const float X00 = wx0 * s000 + wx1 * s001, X01 = wx0 * s010 + wx1 * s011;
    const float X10 = wx0 * s100 + wx1 * s101, X11 = wx0 * s110 + wx1 * s111;
    const float Y0 = wy0 * X00 + wy1 * X01, Y1 = wy0 * X10 + wy1 * X11;
    const float wsx = wx0 + wx1, wsy = wy0 + wy1, wsz = wz0 + wz1;
    a.swv += wz0 * Y0 + wz1 * Y1;
    a.sw += (wsz * wsy) * wsx;
    if (GRAD) {
      const float D00 = s001 - s000, D01 = s011 - s010, D10 = s101 - s100, D11 = s111 - s110;
      a.sd.x += wz0 * (wy0 * D00 + wy1 * D01) + wz1 * (wy0 * D10 + wy1 * D11);
      a.sd.y += wz0 * (X01 - X00) + wz1 * (X11 - X10);
      a.sd.z += Y1 - Y0;
      a.sdc.x += (wsz * wsy) * ((x1 ? 1.f : 0.f) - (x0 ? 1.f : 0.f));
      a.sdc.y += (wsz * wsx) * ((y1 ? 1.f : 0.f) - (y0 ? 1.f : 0.f));
      a.sdc.z += (wsy * wsx) * ((z1 ? 1.f : 0.f) - (z0 ? 1.f : 0.f));
    }
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
// Same lookup for the per-sample hot loop: `k127` = 127 / ((hi - lo) + 1e-20) is hoisted per
// channel, so the two IEEE divisions become multiplications (<= 2 ulp on the table coordinate).
__device__ __forceinline__ float4 lookupXFfast(const float4 *__restrict__ table, float lo, float k127, float opacityScale, float in)
{
  const float sc = fminf(fmaxf((in - lo) * k127 + .5f, 0.f), 127.f);
  const float x = sc * (128.f / 127.f) - 0.5f;
  const float fl = floorf(x), w = x - fl;
  const int i0 = max(0, (int)fl), i1 = min(127, (int)fl + 1);
  const float4 a = table[i0], b = table[i1];
  return make_float4(a.x + w * (b.x - a.x), a.y + w * (b.y - a.y), a.z + w * (b.z - a.z), (a.w + w * (b.w - a.w)) * opacityScale);
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

// powf as exp2(y * log2(x)): what the reference's --use_fast_math device build evaluates (exabrick.cu:53-60); the full-precision
// powf was 10 % of the opaque frame's instructions
__device__ __forceinline__ float linear_to_srgb(float x) { return x <= 0.0031308f ? 12.92f * x : 1.055f * __powf(x, 1.f / 2.4f) - 0.055f; }
__device__ __forceinline__ int make_8bit(float f) { return min(255, max(0, (int)(f * 256.f))); }
__device__ __forceinline__ uint32_t make_rgba8(float3 c)
{ return (uint32_t)(make_8bit(c.x) << 0) + (uint32_t)(make_8bit(c.y) << 8) + (uint32_t)(make_8bit(c.z) << 16) + (0xffu << 24); }

// -----------------------------------------------------------------------------------------
// ordered kd traversal, one thread per ray; per-thread stack in shared memory
// -----------------------------------------------------------------------------------------
static constexpr int FRAME_THREADS = 128;   // 16x8 pixel tile per CTA; also the stride of the per-thread shared-memory columns
struct Trav {
  uint2 *stack;          // this thread's column: entry e at stack[e * FRAME_THREADS]  {node, tf bits}
  int sp;
  int cur; float tn, tf; // node being descended (-1: pop)
  unsigned visits;       // kd nodes touched so far (instrumentation; only counted in the STATS instantiations)
  float3 invd;           // 1/d (IEEE): plane distances at inner nodes are (split - o) * invd; they only order and cull
};
// The ray (o, d) is passed to every call instead of living in Trav: for primary rays the voxel-space origin is the
// same for the whole grid (a kernel parameter, no registers), and d is the lane's own copy.

__device__ __forceinline__ void travInit(Trav &T, uint2 *stack, const float3 &o, const float3 &d, float tmin, float tmax, const float *boxLo, const float *boxHi)
{
  T.stack = stack; T.sp = 0; T.cur = -1; T.tn = 0.f; T.tf = 0.f; T.visits = 0;
  T.invd = f3(1.f / d.x, 1.f / d.y, 1.f / d.z);
  float t0, t1;
  if (boxTest(o, d, tmin, tmax, boxLo, boxHi, t0, t1)) {
    // widen a little: node intervals are only used for culling, leaves do the exact test
    T.cur = 0; T.tn = t0 - 1e-5f * fabsf(t0) - 1e-6f; T.tf = t1 + 1e-5f * fabsf(t1) + 1e-6f;
  }
}

struct Segment { uint32_t blobOfs; int K; float t0, t1, fcw; U8 hdr0; };   // hdr0: header of the region's first brick

// One step of the ordered kd traversal: pop if needed, then visit one node.  Returns 0 = keep
// going, 1 = `seg` holds the next region (front-to-back) with t0 = max(tmin, entry) < t1 =
// min(tmax, exit) (exabrick.cu:213-238), 2 = traversal exhausted.  Inner nodes only order and
// cull (plane distance by reciprocal, intervals widened by eps); leaves run the reference's
// boxTest verbatim, which alone decides t0/t1.  The node and its activity byte are fetched
// together, before the cull test needs either.
template <bool COUNT>
__device__ __forceinline__ int travStep(Trav &T, const float3 &o, const float3 &d, const DeviceScene &S, const uint8_t *__restrict__ active,
                                        float tmin, float tmax, Segment &seg)
{
  constexpr int stride = FRAME_THREADS;
  if (T.cur < 0) {
    if (T.sp == 0) return 2;
    T.sp--;
    const uint2 e = T.stack[T.sp * stride];
    T.cur = (int)e.x; T.tf = __uint_as_float(e.y);
    T.tn = tmin;                                     // conservative: everything before tmin is done
  }
  if (COUNT) T.visits++;
  const uint2 raw = __ldg(reinterpret_cast<const uint2 *>(S.kd) + T.cur);
  const unsigned act = __ldg(active + T.cur);
  if (T.tf < tmin - 1e-5f * fabsf(tmin) || T.tn > tmax || !act) { T.cur = -1; return 0; }
  const uint32_t axis = raw.y >> 30, payload = raw.y & 0x3fffffffu;
  if (axis == 3) {
    T.cur = -1;
    if (payload == KD_EMPTY) return 0;
    const U8 hb = ldg256(S.blobs + payload);
    const uint4 h0 = hb.lo, h1 = hb.hi;
    const float lo[3] = {__uint_as_float(h0.x), __uint_as_float(h0.y), __uint_as_float(h0.z)};
    const float hi[3] = {__uint_as_float(h0.w), __uint_as_float(h1.x), __uint_as_float(h1.y)};
    if (boxTest(o, d, tmin, tmax, lo, hi, seg.t0, seg.t1)) {
      seg.blobOfs = payload; seg.K = (int)raw.x; seg.fcw = __uint_as_float(h1.z);
      seg.hdr0 = ldg256(S.blobs + payload + BLOB_HDR_UNITS);
      return 1;
    }
    return 0;
  }
  const float split = __uint_as_float(raw.x);
  const float oa = comp(o, (int)axis), da = comp(d, (int)axis);
  const float tp = (split - oa) * comp(T.invd, (int)axis);
  const bool belowFirst = (oa < split) || (oa == split && da <= 0.f);
  const int first = (int)payload + (belowFirst ? 0 : 1), second = (int)payload + (belowFirst ? 1 : 0);
  const float eps = 1e-5f * fabsf(tp) + 1e-6f;
  if (tp != tp) {                                    // ray inside the plane: visit both halves
    T.stack[T.sp * stride] = make_uint2((uint32_t)second, __float_as_uint(T.tf));
    T.sp++; T.cur = first;
  } else if (tp > T.tf + eps || tp <= 0.f) {
    T.cur = first;
  } else if (tp < T.tn - eps) {
    T.cur = second;
  } else {
    T.stack[T.sp * stride] = make_uint2((uint32_t)second, __float_as_uint(T.tf));
    T.sp++;
    T.cur = first; T.tf = tp + eps;
  }
  return 0;
}

// next region of ONE ray; callable from divergent code (iso pass)
template <bool COUNT>
__device__ __forceinline__ bool travNext(Trav &T, const float3 &o, const float3 &d, const DeviceScene &S, const uint8_t *__restrict__ active,
                                         float tmin, float tmax, Segment &seg)
{
  for (;;) { const int r = travStep<COUNT>(T, o, d, S, active, tmin, tmax, seg); if (r) return r == 1; }
}

// next region for every lane with `need` set; must be called by ALL 32 lanes of the warp.  The
// lanes advance in lock step, one node visit per trip, until the last one has its answer: the
// cost is the LONGEST lane's node count.  (With per-lane loops and early returns the lanes of a
// warp ran their traversals one after the other: measured ~5000 cycles per call on C1/C2.)
template <bool COUNT>
__device__ __forceinline__ bool travNextWarp(Trav &T, const float3 &o, const float3 &d, const DeviceScene &S, const uint8_t *__restrict__ active,
                                             float tmin, float tmax, Segment &seg, bool need)
{
  bool found = false;
  while (__any_sync(FULL, need)) {
    if (need) { const int r = travStep<COUNT>(T, o, d, S, active, tmin, tmax, seg); if (r) { need = false; found = r == 1; } }
  }
  return found;
}

// common sample-lattice set-up of integrateBrick / isoIntegrateBrick (exabrick.cu:1141-1144)
__device__ __forceinline__ float firstLatticeT(float t0, float dt, float offs)
{
  const int i0 = (int)ceilf(__fsub_rn(t0, __fmul_rn(dt, offs)) / dt);
  float t_i = __fmul_rn(offs + (float)i0, dt);
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

struct RayStats { unsigned samples, visits, corners, segments; };

// samplePoint / samplePointWithDerivative (exabrick.cu:781-816, 883-942) inside a known region
template <bool GRAD, bool STATS, bool DUAL>
__device__ __forceinline__ bool sampleRegion(const DeviceScene &S, const Segment &seg, int channel, float3 pos,
                                             float &value, float3 &g, RayStats &rs)
{
  Acc a; accZero(a);
  unsigned nc = 0;
  accumBricks<GRAD, STATS, DUAL>(a, seg.hdr0, S.blobs + seg.blobOfs + BLOB_HDR_UNITS, seg.K, chanBase<DUAL>(S, channel), pos.x, pos.y, pos.z, nc);
  if (STATS) { rs.samples++; rs.visits += seg.K; rs.corners += nc; }
  return accFinish<GRAD>(a, value, g);
}

// One lattice sample of integrateBrick<GRADIENT_SHADING> + integrateVolume (exabrick.cu:1116-1185,
// 988-1016), as written: t_next = min(t_i, t1), sample at the midpoint of [t_last, t_next] with the
// actual segment length in the opacity correction.  The caller owns the loop (k_renderFrame runs
// it as a flat per-lane state machine so that the lanes of a warp stay converged).
template <bool DUAL> __device__ float3 gradientCD(const RenderParams &P, bool iso, float3 pos, float delta, int channel);

template <bool GRAD, bool STATS, bool DUAL, bool CD = false>
__device__ __forceinline__ void integrateSample(const RenderParams &P, const float4 *xfTables, const float *xfK, const Segment &seg,
                                                float3 o, float3 d, float t_sample, float actual_dt, float4 &pc, RayStats &rs)
{
  const float3 pos = rayAt(o, d, t_sample);
  for (int c = 0; c < P.numChannels; c++) {
    float v = 0.f; float3 g = f3(0, 0, 0);
    if (sampleRegion<GRAD && !CD, STATS, DUAL>(P.scene, seg, c, pos, v, g, rs) && actual_dt != 0.f) {
      if (GRAD && CD) g = gradientCD<DUAL>(P, false, pos, (seg.fcw + 1.f) * .5f, c);      // samplePointWithDerivative<false>, central differences (:928-931)
      float4 s = lookupXFfast(xfTables + c * 128, P.xfDomain[c][0], xfK[c], P.xfOpacityScale, v);
      if (GRAD) {
        const float gg = dot3(g, g), thr = (float)(int)seg.fcw * 1e-6f;
        if (gg > thr * thr) {                               // length(g) > fcw*1e-6 (:1004)
          const float3 l = f3(-d.x, -d.y, -d.z);
          const float scale = fabsf(dot3(l, g)) * rsqrtf(gg * dot3(l, l));
          s.x *= scale; s.y *= scale; s.z *= scale;
        }
      }
      // opacity correction 1 - (1-a)^dt (:1011); the reference's device build is
      // --use_fast_math, i.e. exactly this exp2(dt*log2(1-a)) form
      s.w = 1.f - __powf(1.f - s.w, actual_dt);
#ifdef EXA_DEBUG
      if (P.dbg) printf("G   smp t=%.9g adt=%.9g v=%.9g g=(%.6g %.6g %.6g) rgba=(%.6g %.6g %.6g %.6g)\n", t_sample, actual_dt, v, g.x, g.y, g.z, s.x, s.y, s.z, s.w);
#endif
      const float k = (1.f - pc.w) * s.w;
      pc.x += k * s.x; pc.y += k * s.y; pc.z += k * s.z; pc.w += k;
    }
  }
}

// IsoSurfaceIntegrationFunction::operator() (exabrick.cu:1024-1109) for channel `c`
template <bool STATS, bool DUAL, bool CD = false>
__device__ __forceinline__ void isoFunctor(const RenderParams &P, const float4 *xfTables, const Segment &seg, float3 o, float3 d,
                                           int c, float t_sample, float cellValue, float &lastT, float &lastV,
                                           float4 &pc, float &tHit, float3 &hitGrad, RayStats &rs)
{
  if (lastV >= -1e35f) {
    for (int i = 0; i < EXA_MAX_ISO_SURFACES; i++) {
      const float isoV = P.isoValue[i];
      if (P.isoEnabled[i] && P.isoChannel[i] == c &&
          ((lastV <= isoV && cellValue >= isoV) || (lastV >= isoV && cellValue <= isoV))) {
        const float d1 = fabsf(lastV - isoV), d2 = fabsf(cellValue - isoV);
        const float w1 = 1.f - d1 / (d1 + d2), w2 = 1.f - d2 / (d1 + d2);
        const float tavg = __fadd_rn(__fmul_rn(lastT, w1), __fmul_rn(t_sample, w2));
        const float3 isopt = rayAt(o, d, tavg);
        float cv = 0.f; float3 grad = f3(0, 0, 0);
        float4 smp = make_float4(1.f, 0.f, 0.f, 1.f);
        const int ich = P.isoChannel[i];
        if (P.gradISO) {
          if (sampleRegion<!CD, STATS, DUAL>(P.scene, seg, ich, isopt, cv, grad, rs)) {
            if (CD) grad = gradientCD<DUAL>(P, true, isopt, (seg.fcw + 1.f) * .5f, ich);     // samplePointWithDerivative<true>: the iso-surface BVH
            smp = lookupXF(xfTables + ich * 128, P.xfDomain[ich][0], P.xfDomain[ich][1], P.xfOpacityScale, cv);
            grad = normalize3(grad);
            if (dot3(grad, d) > 0.f) grad = f3(-grad.x, -grad.y, -grad.z);
          }
        } else {
          if (sampleRegion<false, STATS, DUAL>(P.scene, seg, ich, isopt, cv, grad, rs))
            smp = lookupXF(xfTables + ich * 128, P.xfDomain[ich][0], P.xfDomain[ich][1], P.xfOpacityScale, cv);
        }
        if (P.colormapChannel != 0) {
          cv = 0.f; float3 g2;
          if (sampleRegion<false, STATS, DUAL>(P.scene, seg, P.colormapChannel, isopt, cv, g2, rs))
            smp = lookupXF(xfTables, P.xfDomain[0][0], P.xfDomain[0][1], P.xfOpacityScale, cv);
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
        tHit = tavg; hitGrad = grad;
      }
    }
  }
  lastT = t_sample; lastV = cellValue;
}

// isoIntegrateBrick (exabrick.cu:1187-1256), as written; per-channel functor state lives in
// this thread's column of shared memory (isoT / isoV, entry c at [c * blockDim.x])
template <bool STATS, bool DUAL, bool CD = false>
__device__ __forceinline__ void isoSegment(const RenderParams &P, const float4 *xfTables, const Segment &seg, float3 o, float3 d,
                                           float t0, float t1, float *isoT, float *isoV,
                                           float4 &pc, float &tHit, float3 &hitGrad, RayStats &rs)
{
  const int stride = blockDim.x;
  const float dt = P.dt * seg.fcw;
  float t_i = firstLatticeT(t0, dt, 0.f);
  float t_last = t0;
  if (STATS) rs.segments++;
  for (;; t_i += dt) {
    const float t_next = fminf(t_i, t1);
    const float t_sample = 0.5f * (t_next + t_last);
    t_last = t_next;
    const float3 pos = rayAt(o, d, t_sample);
    for (int c = 0; c < P.numChannels; c++) {
      float v = 0.f; float3 g = f3(0, 0, 0); bool ok;
      // (the gradient of these samples is never used; the central-difference build's six extra samples are not taken)
      if (P.gradISO && !CD) ok = sampleRegion<true, STATS, DUAL>(P.scene, seg, c, pos, v, g, rs);
      else ok = sampleRegion<false, STATS, DUAL>(P.scene, seg, c, pos, v, g, rs);
      if (ok) {
        float lt = isoT[c * stride], lv = isoV[c * stride];
        isoFunctor<STATS, DUAL, CD>(P, xfTables, seg, o, d, c, t_sample, v, lt, lv, pc, tHit, hitGrad, rs);
        isoT[c * stride] = lt; isoV[c * stride] = lv;
        if (pc.w >= P.ertThreshold) break;
      }
    }
    if (t_next >= t1) return;
  }
}

__device__ __forceinline__ float3 xfmVector(const float *m, float3 v)
{ return f3(dot3nc(v, f3(m[0], m[3], m[6])), dot3nc(v, f3(m[1], m[4], m[7])), dot3nc(v, f3(m[2], m[5], m[8]))); }
__device__ __forceinline__ float3 xfmPoint(const float *m, float3 v)
{ const float3 r = xfmVector(m, v); return f3(__fadd_rn(r.x, m[9]), __fadd_rn(r.y, m[10]), __fadd_rn(r.z, m[11])); }

// -----------------------------------------------------------------------------------------
// surface pass (traceSurfaces, exabrick.cu:1476-1529): contour planes, streamline cones and
// implicit iso-surfaces; primary rays and the two ambient-occlusion rays of a hit share it.
// -----------------------------------------------------------------------------------------
enum { PRIM_NONE = -1, PRIM_ISOSURFACE = -23, PRIM_PLANE = -24, PRIM_STREAMLINE = -25 };
struct SurfHit { int prim; float t; float3 Ng, base; float ambient; };
static constexpr int SURF_THREADS = FRAME_THREADS;            // stride of the per-thread smem columns

template <bool GRAD, bool STATS, bool DUAL>
__device__ __forceinline__ bool samplePointGlobal(const DeviceScene &S, float px, float py, float pz, int channel,
                                                  float &value, float3 &g, unsigned &visits, unsigned &corners, unsigned *kdSteps = nullptr);
__device__ __forceinline__ bool locate(const DeviceScene &S, float px, float py, float pz, uint32_t &blobOfs, int &K, unsigned *steps);

// traceIsoRay (exabrick.cu:1408-1460): march the iso-active regions front to back with lattice
// offset 0 until two consecutive valid samples bracket an enabled iso value.
// (not __noinline__: ptxas 12.9 segfaults on an out-of-line copy of this function)
template <bool STATS, bool DUAL, bool CD = false>
__device__ void traceIsoRay(const RenderParams &P, const float4 *xfTables, uint2 *stack, float *isoT, float *isoV,
                                         float3 o, float3 d, float tmin, float tmax, SurfHit &out, RayStats &rs)
{
  out.prim = PRIM_NONE; out.t = tmax; out.Ng = f3(0, 0, 0); out.base = f3(0, 0, 0); out.ambient = 0.f;
  const float3 vo = xfmPoint(P.xfm, o); float3 vd = xfmVector(P.xfm, d);
  const float dtScale = len3nc(vd); vd = normalize3nc(vd);
  float already = dtScale * tmin; float rtmax = tmax;
  for (int c = 0; c < P.numChannels; c++) { isoV[c * SURF_THREADS] = -1e36f; isoT[c * SURF_THREADS] = 0.f; }
  rtmax = rtmax * dtScale;
  Trav T; travInit(T, stack, vo, vd, already, rtmax, P.rayLo, P.rayHi);
  for (;;) {
    Segment seg;
    if (!travNext<STATS>(T, vo, vd, P.scene, P.isoActive, already, rtmax, seg)) break;
    float4 ipc = make_float4(0, 0, 0, 0); float th = -1.f; float3 hg = f3(0, 0, 0);
    isoSegment<STATS, DUAL, CD>(P, xfTables, seg, vo, vd, fmaxf(already, seg.t0), fminf(rtmax, seg.t1), isoT, isoV, ipc, th, hg, rs);
    if (th >= 0.f) {
      out.prim = PRIM_ISOSURFACE; out.t = th / dtScale; out.Ng = normalize3(hg); out.ambient = 0.f; out.base = f3(ipc.x, ipc.y, ipc.z);
      return;
    }
    already = seg.t1 * 1.0000001f;
    rtmax = rtmax * dtScale;                          // (sic) exabrick.cu:1434
  }
}

// traceContourRay for every enabled plane (exabrick.cu:1346-1406, 1491-1501).  Quirks kept: the
// sample position is world-space (no voxelSpaceTransform) and reads channel 0
// (samplePointWithInfRay(pos, 0), :818-830, :1396); the XF is the plane's channel.  An invalid
// sample classifies the value 0 (the reference reads an uninitialised float there).
template <bool STATS, bool DUAL>
__device__ void traceContourPlanes(const RenderParams &P, const float4 *xfTables, float3 o, float3 d, SurfHit &prd, RayStats &rs)
{
  for (int i = 0; i < EXA_MAX_CONTOUR_PLANES; i++) {
    if (!P.cpEnabled[i]) continue;
    const float3 n = f3(P.cpNormal[i][0], P.cpNormal[i][1], P.cpNormal[i][2]);
    const float t = surf::contourPlaneT(o, d, n, P.cpOffset[i], P.worldLo, P.worldHi);
    if (t < 0.f) continue;
    const float3 pos = rayAt(o, d, t);
    float value = 0.f, v2 = 0.f; float3 g; unsigned vis = 0, cor = 0;
    if (samplePointGlobal<false, STATS, DUAL>(P.scene, pos.x, pos.y, pos.z, 0, v2, g, vis, cor)) value = v2;
    if (STATS && vis) { rs.samples++; rs.visits += vis; rs.corners += cor; }
    const int ch = P.cpChannel[i];
    const float4 smp = lookupXF(xfTables + ch * 128, P.xfDomain[ch][0], P.xfDomain[ch][1], P.xfOpacityScale, value);
    if (t < prd.t) { prd.prim = PRIM_PLANE; prd.t = t; prd.Ng = n; prd.ambient = 0.f; prd.base = f3(smp.x, smp.y, smp.z); }
  }
}

// Streamline closest hit (bounds :532-570, intersect :503-530): segment k of a trace is visible
// while k < currentTimestep-1 and both ends are alive; radius 2 at both ends.  The reference
// rebuilds an OptiX BVH over all segments after every tracer step; here the segments are binned
// into a uniform grid (k_coneGrid, same moment) and a ray walks the grid cells it crosses front to
// back (3D DDA), testing the segments listed in each cell, and stops at the first cell whose exit
// lies behind the closest hit found.  Ties go to the lower segment id (the order of a serial loop).
__device__ __noinline__ void traceCones(const RenderParams &P, float3 o, float3 d, float tmin, float tmax, SurfHit &sl)
{
  sl.prim = PRIM_NONE; sl.t = 2e10f; sl.Ng = f3(0, 0, 0); sl.base = f3(.8f, .8f, .8f); sl.ambient = 0.f;
  if (!P.traces || !P.coneStart) return;
  const int G = P.coneG;
  const float glo[3] = {P.coneLo[0], P.coneLo[1], P.coneLo[2]};
  const float ghi[3] = {glo[0] + P.coneCell[0] * G, glo[1] + P.coneCell[1] * G, glo[2] + P.coneCell[2] * G};
  float t0, t1;
  if (!boxTest(o, d, tmin, tmax, glo, ghi, t0, t1)) return;
  const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
  int cell[3], step[3]; float tNext[3], tDelta[3];
  const float tStart = t0 + 1e-5f * fabsf(t0);
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const float pa = oo[a] + tStart * dd[a];
    cell[a] = min(G - 1, max(0, (int)floorf((pa - glo[a]) * P.coneInvCell[a])));
    if (dd[a] > 0.f) { step[a] = 1; tNext[a] = (glo[a] + (cell[a] + 1) * P.coneCell[a] - oo[a]) / dd[a]; tDelta[a] = P.coneCell[a] / dd[a]; }
    else if (dd[a] < 0.f) { step[a] = -1; tNext[a] = (glo[a] + cell[a] * P.coneCell[a] - oo[a]) / dd[a]; tDelta[a] = -P.coneCell[a] / dd[a]; }
    else { step[a] = 0; tNext[a] = 3e38f; tDelta[a] = 3e38f; }
  }
  float best = tmax; int bestId = 0x7fffffff;
  for (;;) {
    const int c = cell[0] + G * (cell[1] + G * cell[2]);
    const int e0 = __ldg(P.coneStart + c), e1 = __ldg(P.coneStart + c + 1);
    for (int e = e0; e < e1; e++) {
      const int id = __ldg(P.coneList + e);
      const float *q = P.traces + 3 * (size_t)id;
      const float3 pa = f3(__ldg(q), __ldg(q + 1), __ldg(q + 2)), pb = f3(__ldg(q + 3), __ldg(q + 4), __ldg(q + 5));
      float t; float3 n;
      if (surf::roundedCone(pa, pb, 2.f, 2.f, o, d, t, n) && t > tmin && (t < best || (t == best && id < bestId))) {
        best = t; bestId = id; sl.prim = PRIM_STREAMLINE; sl.t = t; sl.Ng = normalize3(n);
      }
    }
    const int a = tNext[0] <= tNext[1] ? (tNext[0] <= tNext[2] ? 0 : 2) : (tNext[1] <= tNext[2] ? 1 : 2);
    const float tExit = tNext[a];
    if (best <= tExit || tExit > t1) break;                    // the hit lies in a cell already visited / ray leaves the grid or its range
    cell[a] += step[a];
    if (cell[a] < 0 || cell[a] >= G) break;
    tNext[a] += tDelta[a];
  }
}

// Triangle meshes (owl::traceRay(surfaceModel) + closest hit SurfaceBVH, exabrick.cu:420-433,
// 1488-1490): stack traversal of the host-built BVH, closest hit inside (tmin, tmax); no culling.
// The ray/triangle test is the one the reference itself uses for contour planes (:1316-1344);
// OptiX's built-in triangle test differs from it only on shared edges.
__device__ __noinline__ void traceMesh(const RenderParams &P, float3 o, float3 d, float tmin, float tmax, SurfHit &prd)
{
  if (!P.meshNodes || P.numMeshTris <= 0) return;
  float best = tmax; int bestTri = -1;
  const float3 invd = f3(1.f / d.x, 1.f / d.y, 1.f / d.z);
  uint32_t stack[64]; int sp = 0;
  stack[sp++] = 0;
  while (sp) {
    const uint32_t ni = stack[--sp];
    const float4 n0 = __ldg(P.meshNodes + 2 * (size_t)ni), n1 = __ldg(P.meshNodes + 2 * (size_t)ni + 1);
    const float px = 1e-4f * (n1.x - n0.x) + 1e-6f * (fabsf(n0.x) + fabsf(n1.x)), py = 1e-4f * (n1.y - n0.y) + 1e-6f * (fabsf(n0.y) + fabsf(n1.y)),
                pz = 1e-4f * (n1.z - n0.z) + 1e-6f * (fabsf(n0.z) + fabsf(n1.z));
    const float ax = (n0.x - px - o.x) * invd.x, bx = (n1.x + px - o.x) * invd.x;
    const float ay = (n0.y - py - o.y) * invd.y, by = (n1.y + py - o.y) * invd.y;
    const float az = (n0.z - pz - o.z) * invd.z, bz = (n1.z + pz - o.z) * invd.z;
    const float t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), tmin));
    const float t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), best));
    if (t0 > t1) continue;
    const uint32_t a = __float_as_uint(n0.w), b = __float_as_uint(n1.w);
    if (b) {
      for (uint32_t k = 0; k < b; k++) {
        const float *t = P.meshTris + 9 * (size_t)(a + k);
        const float3 A = f3(__ldg(t), __ldg(t + 1), __ldg(t + 2));
        const float3 e1 = f3(__ldg(t + 3) - A.x, __ldg(t + 4) - A.y, __ldg(t + 5) - A.z);
        const float3 e2 = f3(__ldg(t + 6) - A.x, __ldg(t + 7) - A.y, __ldg(t + 8) - A.z);
        const float tt = surf::rayTriangle(o, d, A, e1, e2);
        if (tt > tmin && tt < best) { best = tt; bestTri = (int)(a + k); }
      }
    } else if (sp < 62) { stack[sp++] = a; stack[sp++] = a + 1; }   // median-split BVH over < 2^30 triangles: depth <= 30, at most depth + 1 entries pending
  }
  if (bestTri >= 0) {
    const float *t = P.meshTris + 9 * (size_t)bestTri;
    const float3 A = f3(__ldg(t), __ldg(t + 1), __ldg(t + 2));
    const float3 e1 = f3(__ldg(t + 3) - A.x, __ldg(t + 4) - A.y, __ldg(t + 5) - A.z);
    const float3 e2 = f3(__ldg(t + 6) - A.x, __ldg(t + 7) - A.y, __ldg(t + 8) - A.z);
    prd.prim = bestTri; prd.t = best; prd.Ng = normalize3(surf::crossp(e1, e2)); prd.ambient = .2f; prd.base = f3(.8f, .8f, .8f);
  }
}

// ALL = false: iso-surfaces only (the instantiation the iso + DVR workload C4 runs; meshes, planes,
// cones and AO cost it registers and 10 % of its frame time when compiled in)
template <bool STATS, bool DUAL, bool ALL, bool CD = false>
__device__ __forceinline__ void traceSurfaces(const RenderParams &P, const float4 *xfTables, uint2 *stack, float *isoT, float *isoV,
                                              float3 o, float3 d, float tmin, float tmax, bool withContourPlanes, SurfHit &prd, RayStats &rs)
{
  prd.prim = PRIM_NONE; prd.t = tmax; prd.Ng = f3(0, 0, 0); prd.base = f3(0, 0, 0); prd.ambient = 0.f;
  if (ALL) {
    traceMesh(P, o, d, tmin, tmax, prd);
    if (withContourPlanes) traceContourPlanes<STATS, DUAL>(P, xfTables, o, d, prd, rs);
    if (P.traces) {
      SurfHit sl; traceCones(P, o, d, tmin, tmax, sl);
      if (sl.prim == PRIM_STREAMLINE && sl.t < prd.t) prd = sl;
    }
  }
  if (P.anyIso) {
    SurfHit iso; traceIsoRay<STATS, DUAL, CD>(P, xfTables, stack, isoT, isoV, o, d, tmin, tmax, iso, rs);
    if (iso.prim == PRIM_ISOSURFACE && iso.t < prd.t) prd = iso;
  }
}

// -----------------------------------------------------------------------------------------
// __raygen__renderFrame (exabrick.cu:1576-1720): one thread per pixel, 8x4 pixel tile per warp
// -----------------------------------------------------------------------------------------
// final colour of a pixel (exabrick.cu:1701-1719): background blend (sic: premultiplied colour
// times alpha again), progressive accumulation, sRGB, RGBA8 pack; or the raw premultiplied
// partial in sort-last mode
// Sort-last: the premultiplied partial of this rank's sub-volume, NOT yet scaled by the early-termination
// rule (a terminated ray is marked by alpha + 2), and - when a surface pass ran - the shaded colour and
// world-space distance of this rank's nearest surface hit (w = 1e30: none); k_compositeFinish puts them together.
__device__ __forceinline__ void writePixel(const RenderParams &P, int pixelIdx, float4 pc, float3 bg, long long tStart, float surfT = 1e30f)
{
  if (P.partialOut) {
    P.partialOut[pixelIdx] = pc;
    if (P.partialSurf) P.partialSurf[pixelIdx] = make_float4(bg.x, bg.y, bg.z, surfT);
    return;
  }
  float3 color = f3(pc.w * pc.x + (1.f - pc.w) * bg.x, pc.w * pc.y + (1.f - pc.w) * bg.y, pc.w * pc.z + (1.f - pc.w) * bg.z);   // :1701 (sic)
  // clockScale heat-map (:1703-1707): the cycles this pixel took (here: since its CTA started), scaled, into the red channel
  if (P.clockScale > 0.f) color.x = fminf(P.clockScale * (float)(clock64() - tStart) / 1000000.f, 1.f);
  if (P.frameID > 0) { const float4 acc = P.accum[pixelIdx]; color = color + f3(acc.x, acc.y, acc.z); }
  P.accum[pixelIdx] = make_float4(color.x, color.y, color.z, 1.f);
  const float inv = (float)P.frameID + 1.f;
  color = f3(linear_to_srgb(color.x / inv), linear_to_srgb(color.y / inv), linear_to_srgb(color.z / inv));
  if (P.numStrips > 0) {
    // frame distributed as row strips over the GPUs: store into the strip that holds this pixel's tile row
    const int tr = (pixelIdx / P.width) >> 3;
    int k = 0;
    while (tr >= P.stripRow[k + 1]) k++;                         // warp-uniform: a warp's 4 pixel rows lie in one tile row
    P.strip[k][pixelIdx - (P.stripRow[k] << 3) * P.width] = make_rgba8(color);
  } else P.rgba8[pixelIdx] = make_rgba8(color);
}

#ifndef EXA_FRAME_MINBLOCKS
#define EXA_FRAME_MINBLOCKS 6     // 80 registers, 24 warps / SM: C3 opaque 0.718 -> 0.660 ms, translucent 7.29 -> 6.92 ms against 5 (96 registers); 8 (64 registers, spills) is slower (round 2, GPU calls A-C)
#endif
static constexpr int TILE_W = 16, TILE_H = 8;

// SURF: the surface pass is compiled in only when a surface is enabled (RenderParams::anySurf):
// 1 = implicit iso-surfaces only, 2 = + triangle meshes, contour planes, streamline cones, AO; the plain
// DVR instantiation is less than half the code, which matters because every CTA runs through
// the kernel once, front to back (instruction fetch, not data, is what a lone CTA waits for).
// Image-tile mode: tiles are dealt in blocks of numRanks consecutive tiles and the deal rotates by
// one from block to block, so that a rank does not get the same tile columns in every tile row
// (a plain t mod N does when N divides the tiles per row: 3840/16 = 240 = 30 * 8).
__host__ __device__ inline int ownedTile(int j, int rank, int numRanks) { return j * numRanks + (rank + j) % numRanks; }
__host__ __device__ inline int numOwnedTiles(int numTiles, int rank, int numRanks)
{
  const int full = numTiles / numRanks, rem = numTiles % numRanks;
  return full + ((rank + full) % numRanks < rem ? 1 : 0);
}

// Measured and dropped in round 1 (profiles/perf_quick_r01_*; the code is in the history before round 2): a
// one-node traversal look-ahead, the first region by point location, next-record L1 prefetch, and a second
// kernel that finishes long rays one warp per ray.  All parity-green, all slower on an issue-bound kernel.
template <bool GRAD, bool STATS, bool DUAL, int SURF, bool CD = false>
__global__ void __launch_bounds__(FRAME_THREADS, EXA_FRAME_MINBLOCKS) k_renderFrame(const __grid_constant__ RenderParams P)
{
  extern __shared__ uint4 smem[];
  // layout: [xf tables: numChannels*128 float4][kd stacks: kdStack x 128 uint2][iso state: 2 x numChannels x 128 float]
  float4 *xfTables = reinterpret_cast<float4 *>(smem);
  uint2 *stackBase = reinterpret_cast<uint2 *>(smem + P.numChannels * 128);
  float *isoBase = reinterpret_cast<float *>(stackBase + (size_t)P.kdStack * FRAME_THREADS);
  __shared__ float xfK[EXA_MAX_CHANNELS];      // 127 / ((hi - lo) + 1e-20) per channel (lookupXFfast)
  // the previous frame's bucket counts: 4 independent 128-bit loads issued before anything else
  // (a loop of 16 dependent loads with an early exit cost the cheap tiles ~4 000 cycles each)
  int4 tc[TILE_BUCKETS / 4];
  if (P.tileOrderIn) {
    const int4 *cp = reinterpret_cast<const int4 *>(P.tileOrderIn->count);
#pragma unroll
    for (int i = 0; i < TILE_BUCKETS / 4; i++) tc[i] = cp[i];
  }
  for (int i = threadIdx.x; i < P.numChannels * 128; i += FRAME_THREADS) xfTables[i] = P.xf[i];
  if (threadIdx.x < P.numChannels) xfK[threadIdx.x] = 127.f / ((P.xfDomain[threadIdx.x][1] - P.xfDomain[threadIdx.x][0]) + 1e-20f);
  __syncthreads();

  const int tilesX = (P.width + TILE_W - 1) / TILE_W;
  // Tile schedule: CTA q renders the q-th tile of the PREVIOUS frame's cost order (most expensive
  // first: longest-processing-time-first keeps the SMs busy to the end of the frame; without it
  // a few tiles of long rays finish alone).  Costs are CTA cycle counts, bucketed by log2; each
  // frame writes the lists the next one reads.  Every thread runs the 15-bucket search itself: doing it in
  // warp 0 only (prefix sum over the counts, tile id through shared memory) saves 3 % of the opaque frame's
  // instructions and is 0.6 % SLOWER (the CTA then waits for one warp's dependent loads; profiles/perf_quick_r02_n_*).
  // Image-tile mode (multi-GPU, data replicated): the grid holds only the owned tiles (ownedTile()) or an explicit list.
  const long long t_start = clock64();
  int tileId;
  if (P.tileList) tileId = (int)P.tileList[blockIdx.x];
  else if (P.tileOrderIn) {
    int q = (int)blockIdx.x, b = 0; bool found = false;
#pragma unroll
    for (int i = TILE_BUCKETS - 1; i > 0; i--) {            // most expensive bucket first
      const int4 v = tc[i >> 2]; const int n = (i & 3) == 0 ? v.x : (i & 3) == 1 ? v.y : (i & 3) == 2 ? v.z : v.w;
      if (!found) { if (q < n) { b = i; found = true; } else q -= n; }
    }
    tileId = (int)P.tileOrderIn->list[(size_t)b * P.tileOrderCap + q];
  } else if (P.gridW > 0) tileId = (P.gridY0 + (int)blockIdx.x / P.gridW) * tilesX + P.gridX0 + (int)blockIdx.x % P.gridW;
  else tileId = P.tileMode ? ownedTile((int)blockIdx.x, P.rank, P.numRanks) : (int)blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // warp w covers the 8x4 sub-tile (w&1, w>>1) of the 16x8 CTA tile
  const int px = (tileId % tilesX) * TILE_W + (warp & 1) * 8 + (lane & 7);
  const int py = (tileId / tilesX) * TILE_H + (warp >> 1) * 4 + (lane >> 3);
  RayStats rs = {0, 0, 0, 0};
  bool inside = px < P.width && py < P.height;
  // A tile that lies outside the screen rectangle of the scene bounds: none of its rays can hit the volume, and without a
  // surface pass the pixel is the (black) background whatever the jitter: no random numbers, no ray, no traversal.
  if (SURF == 0) {
    const int tx0 = (tileId % tilesX) * TILE_W, ty0 = (tileId / tilesX) * TILE_H;
    if (tx0 > P.sceneRect[2] || tx0 + TILE_W - 1 < P.sceneRect[0] || ty0 > P.sceneRect[3] || ty0 + TILE_H - 1 < P.sceneRect[1]) {
      if (inside) writePixel(P, px + P.width * py, make_float4(0, 0, 0, 0), f3(0, 0, 0), t_start);
      inside = false;
    }
  }
#ifdef EXA_DEBUG
  if (P.dbg && (px + P.width * py) != P.dbg - 1) inside = false;
#endif
  // per-lane DVR state; the integration loop below is shared by all 32 lanes of the warp.  The voxel-space
  // ray origin is the transformed camera position for EVERY ray of the frame: it comes from the kernel
  // parameters (computed once on the host with the same separately rounded products and sums as xfmPoint),
  // costs no registers and makes the plane-side tests of the traversal warp-uniform.
  const int pixelIdx = px + P.width * py;
  bool live = false;                                         // lane still has volume to integrate
  const float3 o = f3(P.camVoxel[0], P.camVoxel[1], P.camVoxel[2]);
  float3 d = f3(0, 0, 1), bg = f3(0, 0, 0);
  float4 pc = make_float4(0, 0, 0, 0);
  float offs = 0.f, already = 0.f, rayTmax = 0.f;
  Trav T; T.stack = stackBase + threadIdx.x; T.sp = 0; T.cur = -1; T.tn = 0.f; T.tf = 0.f; T.visits = 0; T.invd = f3(0, 0, 0);
  bool haveSeg = false; Segment seg; seg.blobOfs = 0; seg.K = 0; seg.t0 = seg.t1 = 0.f; seg.fcw = 1.f;   // current region segment of this lane
  float dtS = 0.f, t_i = 0.f, t_last = 0.f;
  float surfHitT = 1e30f;                                    // sort-last: distance of this rank's nearest surface hit
  if (inside) {
    uint2 *stack = stackBase + threadIdx.x;
    LCG16 rnd((uint32_t)(P.frameID * P.width * P.height) + (uint32_t)px, (uint32_t)py);
    const float jx = rnd.next(), jy = rnd.next();
    const float sx = (float)px + jx, sy = (float)py + jy;
    // Camera::generateRay (programs/Camera.h:27-44)
    d = normalize3nc(f3(__fadd_rn(__fadd_rn(P.dir00[0], __fmul_rn(sx, P.dirDu[0])), __fmul_rn(sy, P.dirDv[0])),
                        __fadd_rn(__fadd_rn(P.dir00[1], __fmul_rn(sx, P.dirDu[1])), __fmul_rn(sy, P.dirDv[1])),
                        __fadd_rn(__fadd_rn(P.dir00[2], __fmul_rn(sx, P.dirDu[2])), __fmul_rn(sy, P.dirDv[2]))));
    float tmin = 1e-6f, tmax = 1e8f;
    // ---- surfaces (traceSurfaces :1476-1529) and their shading with ambient occlusion (:1604-1652)
    float surfT = tmax;
    if (SURF && P.anySurf) {
      float *isoT = isoBase + threadIdx.x, *isoV = isoBase + (size_t)P.numChannels * FRAME_THREADS + threadIdx.x;
      const float3 ow = f3(P.camPos[0], P.camPos[1], P.camPos[2]);
      SurfHit sh;
      traceSurfaces<STATS, DUAL, SURF == 2, CD>(P, xfTables, stack, isoT, isoV, ow, d, tmin, tmax, true, sh, rs);
      surfT = sh.t;
      if (sh.prim != PRIM_NONE) {
        // sort-last: an iso-surface hit lies inside THIS rank's part (everything behind the part is hidden); meshes, planes
        // and cones are replicated, every rank finds the same hit and has clipped its own samples at it: marked by -t
        surfHitT = sh.prim == PRIM_ISOSURFACE ? sh.t : -sh.t;
        const bool shade = sh.prim >= 0 || sh.prim == PRIM_STREAMLINE || sh.prim == PRIM_PLANE || (sh.prim == PRIM_ISOSURFACE && P.gradISO != 0);
        if (shade && len3(sh.Ng) > 0.f) {
          const int aoSamples = (SURF == 2 && P.aoEnabled) ? 2 : 0;
          const float3 isect = rayAt(ow, d, sh.t);
          int hitCnt = 0;
          for (int i = 0; i < aoSamples; i++) {
            const float u1 = rnd.next(), u2 = rnd.next();
            const float3 adir = surf::aoDirection(sh.Ng, u1, u2);
            SurfHit ah;                                       // ST_ALL_SURFACES & ~ST_CONTOUR_PLANES (:1638-1640)
            traceSurfaces<STATS, DUAL, SURF == 2, CD>(P, xfTables, stack, isoT, isoV, isect, adir, 1e-4f, P.aoLength, false, ah, rs);
            if (ah.prim != PRIM_NONE) hitCnt++;
          }
          const float shadow = aoSamples ? (float)hitCnt / (float)aoSamples : 0.f;
          const float k = fabsf(dot3(d, sh.Ng)) * (1.f - shadow);
          bg = f3(sh.ambient + sh.base.x * k, sh.ambient + sh.base.y * k, sh.ambient + sh.base.z * k);
        } else bg = sh.base;
      }
    }
    // ---- DVR set-up (:1654-1674)
    offs = rnd.next();
    tmax = surfT;
    if (P.clipEnabled) { float a, b; boxTest(f3(P.camPos[0], P.camPos[1], P.camPos[2]), d, tmin, tmax, P.clipLo, P.clipHi, a, b); tmin = a; tmax = b; }
    const float tHitWorld = tmax;
    d = xfmVector(P.xfm, d);
    const float dtScale = len3nc(d); d = normalize3nc(d);
    already = dtScale * tmin;
    rayTmax = tHitWorld * dtScale;
    travInit(T, stack, o, d, already, rayTmax, P.rayLo, P.rayHi);
    live = true;
  }
  // ---- DVR region loop (:1675-1699) + integrateBrick's sample loop (:1141-1184), flattened into
  // one per-lane state machine: each trip, lanes without a current region segment advance the kd
  // traversal to their next one, then every live lane takes exactly ONE lattice sample.  With the
  // nested loops as written, a warp pays sum-over-segments of the longest lane; flat, it pays the
  // longest lane's total.  __all_sync at the head keeps the warp converged.  Arithmetic and
  // per-ray order of operations are unchanged.
  for (;;) {
    if (__all_sync(FULL, !live)) break;
    const bool need = live && !haveSeg;
    const bool got = travNextWarp<STATS>(T, o, d, P.scene, P.volActive, already, rayTmax, seg, need);
    if (need) {
      if (!got) live = false;
      else {
        dtS = P.dt * seg.fcw; t_i = firstLatticeT(seg.t0, dtS, offs); t_last = seg.t0; haveSeg = true;
        if (STATS) rs.segments++;
#ifdef EXA_DEBUG
        if (P.dbg) printf("G seg rid=%d K=%d t0=%.9g t1=%.9g dt=%.9g tFirst=%.9g\n", (int)P.scene.blobs[seg.blobOfs + 2].x, seg.K, seg.t0, seg.t1, dtS, t_i);
#endif
      }
    }
    if (live && haveSeg) {
      const float t_next = fminf(t_i, seg.t1);
      const float t_sample = 0.5f * (t_next + t_last);
      const float actual_dt = t_next - t_last;
      t_last = t_next;
      integrateSample<GRAD, STATS, DUAL, CD>(P, xfTables, xfK, seg, o, d, t_sample, actual_dt, pc, rs);
      if (pc.w >= P.ertThreshold) {                                                                               // :1180, :1694-1697
        if (P.partialOut) pc.w += 2.f; else pc = make_float4(pc.x * pc.w, pc.y * pc.w, pc.z * pc.w, 1.f);
        live = false;
      }
      else if (t_next >= seg.t1) { haveSeg = false; already = seg.t1 * 1.0000001f; }                              // :1698
      else t_i += dtS;
    }
  }
  if (inside) writePixel(P, pixelIdx, pc, bg, t_start, surfHitT);
  if (STATS && P.stats && inside) {                          // instrumentation: per-ray maxima
    atomicMax(&P.stats[4], (unsigned long long)T.visits); atomicAdd(&P.stats[6], (unsigned long long)T.visits);
    atomicMax(&P.stats[5], (unsigned long long)rs.samples);
  }
  if (STATS) {
    unsigned long long a = rs.samples, b = rs.visits, c = rs.corners, e = rs.segments;
    for (int o2 = 16; o2 > 0; o2 >>= 1) {
      a += __shfl_xor_sync(FULL, a, o2); b += __shfl_xor_sync(FULL, b, o2); c += __shfl_xor_sync(FULL, c, o2); e += __shfl_xor_sync(FULL, e, o2);
    }
    if (lane == 0 && P.stats) { atomicAdd(&P.stats[0], a); atomicAdd(&P.stats[1], b); atomicAdd(&P.stats[2], c); atomicAdd(&P.stats[3], e); }
  }
  if (P.tileCost) {                                          // cross-rank balancing: this tile's cycles, for the next deal
    __syncthreads();
    if (threadIdx.x == 0) P.tileCost[tileId] = (uint32_t)min((unsigned long long)(clock64() - t_start), 0xffffffffull);
  }
  if (P.tileOrderOut) {                                      // record this tile's cost for the next frame's schedule
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned long long cyc = (unsigned long long)(clock64() - t_start);
      const int b = min(TILE_BUCKETS - 1, max(0, 63 - __clzll(cyc | 1ull) - 10));      // log2(cycles / 1024)
      const int idx = atomicAdd(&P.tileOrderOut->count[b], 1);
      P.tileOrderOut->list[(size_t)b * P.tileOrderCap + idx] = (uint32_t)tileId;
      // last CTA of the frame: every CTA has read its slot of tileOrderIn (its tile id came from
      // there, long before this point), which is the next frame's output -> clear it here
      // instead of a memset in front of every launch.  No fence: nothing is published.
      if (atomicAdd(&P.tileOrderOut->done, 1) == (int)gridDim.x - 1) {
        P.tileOrderOut->done = 0;
        if (P.tileOrderIn) for (int i = 0; i < TILE_BUCKETS; i++) const_cast<TileOrder *>(P.tileOrderIn)->count[i] = 0;
      }
    }
  }
}

int launchRenderFrame(const RenderParams &P, cudaStream_t s)
{
  const int tilesX = (P.width + TILE_W - 1) / TILE_W, tilesY = (P.height + TILE_H - 1) / TILE_H;
  size_t smemBytes = 16 * (size_t)P.numChannels * 128 + 8 * (size_t)P.kdStack * FRAME_THREADS +
                           (P.anyIso ? 8 * (size_t)P.numChannels * FRAME_THREADS : 0);
  const int numTiles = tilesX * tilesY;
  // experiment knob: at most EXA_FRAME_MAX_CTAS CTAs per SM (dynamic shared memory padded so that no more fit)
  static const int maxCtas = getenv("EXA_FRAME_MAX_CTAS") ? atoi(getenv("EXA_FRAME_MAX_CTAS")) : 0;
  if (maxCtas > 0 && maxCtas < EXA_FRAME_MINBLOCKS) smemBytes = std::max(smemBytes, (size_t)(227 * 1024) / maxCtas - 2048);
  const int numMine = P.tileList ? P.tileListLen : P.gridW > 0 ? P.gridW * P.gridH : P.tileMode ? numOwnedTiles(numTiles, P.rank, P.numRanks) : numTiles;
  if (numMine <= 0) return 0;
  const dim3 grid(numMine), block(FRAME_THREADS);
  const bool stats = P.stats != nullptr;
  // the shared-memory attribute is per device (a process may hold renderers on several devices); cudaFuncSetAttribute
  // is a slow driver call: raise each instantiation's dynamic shared memory limit only when it has to grow
  int dev = 0; cudaGetDevice(&dev); dev &= 63;
#define LAUNCH(G, S_, D_)                                                                                   \
  { static size_t capFd[64] = {}; size_t &capF = capFd[dev];                                                \
    if (smemBytes > capF) { cudaFuncSetAttribute(k_renderFrame<G, S_, D_, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);            \
                            cudaFuncSetAttribute(k_renderFrame<G, S_, D_, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);            \
                            cudaFuncSetAttribute(k_renderFrame<G, S_, D_, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes); capF = smemBytes; } \
    if (P.anySurf == 2) k_renderFrame<G, S_, D_, 2><<<grid, block, smemBytes, s>>>(P);                      \
    else if (P.anySurf) k_renderFrame<G, S_, D_, 1><<<grid, block, smemBytes, s>>>(P);                      \
    else k_renderFrame<G, S_, D_, 0><<<grid, block, smemBytes, s>>>(P); }
#define LAUNCH_D(G, S_) { if (P.scene.dual) LAUNCH(G, S_, true) else LAUNCH(G, S_, false) }
  // central-difference gradients (the reference's ANALYTIC_GRADIENTS=0 build): separate instantiations, so that the default
  // kernels are the code they were; only where a gradient is used at all (DVR shading, or shaded iso-surfaces)
  const bool cd = P.gradCD && !stats && (P.gradDVR || (P.anySurf && P.gradISO));
#define LAUNCH_CD(G, D_, SF)                                                                                \
  { static size_t capCd[64] = {}; size_t &capC = capCd[dev];                                                \
    if (smemBytes > capC) { cudaFuncSetAttribute(k_renderFrame<G, false, D_, SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes); capC = smemBytes; } \
    k_renderFrame<G, false, D_, SF, true><<<grid, block, smemBytes, s>>>(P); }
#define LAUNCH_CD_D(G, SF) { if (P.scene.dual) LAUNCH_CD(G, true, SF) else LAUNCH_CD(G, false, SF) }
  if (cd) {
    if (P.gradDVR) { if (P.anySurf == 2) LAUNCH_CD_D(true, 2) else if (P.anySurf) LAUNCH_CD_D(true, 1) else LAUNCH_CD_D(true, 0) }
    else { if (P.anySurf == 2) LAUNCH_CD_D(false, 2) else LAUNCH_CD_D(false, 1) }
  }
  else if (P.gradDVR) { if (stats) LAUNCH_D(true, true) else LAUNCH_D(true, false) }
  else { if (stats) LAUNCH_D(false, true) else LAUNCH_D(false, false) }
#undef LAUNCH_CD_D
#undef LAUNCH_CD
#undef LAUNCH_D
#undef LAUNCH
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// point location + reconstruction at arbitrary points (one thread per point)
// -----------------------------------------------------------------------------------------
__device__ __forceinline__ bool locate(const DeviceScene &S, float px, float py, float pz, uint32_t &blobOfs, int &K, unsigned *steps)
{
  // Regions are the half-open leaf boxes [lo,hi) of the kd-tree and tile its root box, so
  // "lo <= p < hi of the root" + a descent with (p < split -> left) finds the region with
  // lo <= p < hi; empty leaves are holes.  (NaN fails the first test.)
  if (!(px >= S.boundsLo[0] && px < S.boundsHi[0] && py >= S.boundsLo[1] && py < S.boundsHi[1] && pz >= S.boundsLo[2] && pz < S.boundsHi[2]))
    return false;
  uint2 raw;
  if (S.locGrid) {
    // the grid cell holds a COPY of the deepest kd node that contains the whole cell: for most
    // cells that is already the leaf, and the query is one 8-byte load
    const int cx = min(S.locG[0] - 1, (int)((px - S.boundsLo[0]) * S.locScale[0]));
    const int cy = min(S.locG[1] - 1, (int)((py - S.boundsLo[1]) * S.locScale[1]));
    const int cz = min(S.locG[2] - 1, (int)((pz - S.boundsLo[2]) * S.locScale[2]));
    raw = __ldg(S.locGrid + cx + (size_t)S.locG[0] * (cy + (size_t)S.locG[1] * cz));
  } else raw = __ldg(reinterpret_cast<const uint2 *>(S.kd));
  for (;;) {
    const uint32_t axis = raw.y >> 30, payload = raw.y & 0x3fffffffu;
    if (axis == 3) {
      if (payload == KD_EMPTY) return false;
      blobOfs = payload; K = (int)raw.x;
      return true;
    }
    const float p = axis == 0 ? px : axis == 1 ? py : pz;
    if (steps) ++*steps;
    raw = __ldg(reinterpret_cast<const uint2 *>(S.kd) + payload + (p < __uint_as_float(raw.x) ? 0 : 1));
  }
}

template <bool GRAD, bool STATS, bool DUAL>
__device__ __forceinline__ bool samplePointGlobal(const DeviceScene &S, float px, float py, float pz, int channel,
                                                  float &value, float3 &g, unsigned &visits, unsigned &corners, unsigned *kdSteps)
{
  uint32_t bo; int K;
  if (!locate(S, px, py, pz, bo, K, STATS ? kdSteps : nullptr)) return false;
  Acc a; accZero(a);
  accumBricks<GRAD, STATS, DUAL, true>(a, ldg256<true>(S.blobs + bo + BLOB_HDR_UNITS), S.blobs + bo + BLOB_HDR_UNITS, K, chanBase<DUAL>(S, channel), px, py, pz, corners);
  if (STATS) visits += K;
  return accFinish<GRAD>(a, value, g);
}

// gradientCD<ISO> (exabrick.cu:833-880, FAST_CENTRAL_DIFFERENCES off) for samplePointWithDerivative of the reference built
// with ANALYTIC_GRADIENTS=0 (:928-931): six more samples at pos -+ delta along the axes.  The reference locates each with a
// ray of length 2e-10 through the BVH of the ACTIVE regions (volume BVH, iso-surface BVH for ISO): here the point location
// (regions are half-open boxes, as that ray sees them) followed by the activity test of the region found - the per-region
// volume mask, or the iso values against the region's value range.  Outside every active region, and (decided; the
// reference reads an uninitialised float) where the sample is invalid, the sample is 0.  Result: f(pos - delta) -
// f(pos + delta) per axis (sic).  Only the CD instantiations of the frame kernel call it (not __noinline__: ptxas 12.9
// segfaults on an out-of-line copy, as for traceIsoRay).
template <bool DUAL>
__device__ float3 gradientCD(const RenderParams &P, bool iso, float3 pos, float delta, int channel)
{
  const DeviceScene &S = P.scene;
  float sv[6];
#pragma unroll 1
  for (int k = 0; k < 6; k++) {
    const float dd = (k & 1) ? -delta : delta;                 // j = 0: pos + dt[i], j = 1: pos - dt[i]
    const int i = k >> 1;
    const float qx = i == 0 ? __fadd_rn(pos.x, dd) : pos.x, qy = i == 1 ? __fadd_rn(pos.y, dd) : pos.y, qz = i == 2 ? __fadd_rn(pos.z, dd) : pos.z;
    float v = 0.f;
    uint32_t bo; int K;
    if (locate(S, qx, qy, qz, bo, K, nullptr)) {
      const uint4 h2 = __ldg(S.blobs + bo + 2);                // {regionID, value range lower, upper, -}
      bool act = false;
      if (iso) {
        const float vlo = __uint_as_float(h2.y), vhi = __uint_as_float(h2.z);
        for (int q = 0; q < EXA_MAX_ISO_SURFACES; q++) act |= P.isoEnabled[q] && P.isoValue[q] >= vlo && P.isoValue[q] <= vhi;
      } else act = P.regionActive[(int)h2.x] != 0;
      if (act) {
        Acc a; accZero(a); unsigned nc = 0; float val; float3 g;
        accumBricks<false, false, DUAL>(a, ldg256(S.blobs + bo + BLOB_HDR_UNITS), S.blobs + bo + BLOB_HDR_UNITS, K, chanBase<DUAL>(S, channel), qx, qy, qz, nc);
        if (accFinish<false>(a, val, g)) v = val;
      }
    }
    sv[k] = v;
  }
  return f3(sv[1] - sv[0], sv[3] - sv[2], sv[5] - sv[4]);
}

// 128 threads x 16 CTAs per SM = 2048 resident threads at 32 registers (12 bytes of spills): the kernel is a
// chain of dependent random loads, so resident threads are what hides the latency.  1 M points on C2
// (tools/perf_quick.py): 256 x default registers (40 -> 1536 threads) 53.5 us, 256 x 8 53.3 us, 128 x 16 50.7 us
#ifndef EXA_POINTS_THREADS
#define EXA_POINTS_THREADS 128
#endif
#ifndef EXA_POINTS_MINBLOCKS
#define EXA_POINTS_MINBLOCKS 16
#endif
// One point per thread.  More chains in flight per SM make it SLOWER, not faster (round 2, profiles/perf_quick_r02_s_*): 2 / 3 / 4
// points per thread with the descents advanced together: 63-69 / 88 / 94 us against 50.6 us; so do persistent lanes (93-109 us), a
// two-level location grid (53.7 us although the kd steps a warp pays drop from 11.9 to 8.1), and halving the DRAM reads with
// .L2::64B on the record loads (the default fetches 128 B per miss; 5.36 M -> 3.28 M sectors) changes nothing (48.8 us): the
// query is bound neither by DRAM bandwidth nor by the chain length but by how many random sector requests the L1 / L2 / fabric
// path turns around, and 2048 single-chain threads per SM sit at that optimum.

template <bool GRAD, bool STATS, bool DUAL>
__global__ void __launch_bounds__(EXA_POINTS_THREADS, EXA_POINTS_MINBLOCKS) k_samplePoints(const __grid_constant__ SampleParams P)
{
  unsigned visits = 0, corners = 0, counted = 0, kdSteps = 0;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P.n) {
    const float px = P.xyz[3 * i], py = P.xyz[3 * i + 1], pz = P.xyz[3 * i + 2];
    float v = 0.f; float3 g = f3(0, 0, 0);
    const bool ok = samplePointGlobal<GRAD, STATS, DUAL>(P.scene, px, py, pz, P.channel, v, g, visits, corners, &kdSteps);
    counted = visits > 0;
    if (P.values) P.values[i] = ok ? v : 0.f;
    if (P.valid) P.valid[i] = ok ? 1 : 0;
    if (GRAD && P.grad) { P.grad[3 * i] = ok ? g.x : 0.f; P.grad[3 * i + 1] = ok ? g.y : 0.f; P.grad[3 * i + 2] = ok ? g.z : 0.f; }
  }
  if (STATS && P.stats) {
    unsigned long long a = counted, b = visits, c = corners;
    for (int o2 = 16; o2 > 0; o2 >>= 1) { a += __shfl_xor_sync(FULL, a, o2); b += __shfl_xor_sync(FULL, b, o2); c += __shfl_xor_sync(FULL, c, o2); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&P.stats[0], a); atomicAdd(&P.stats[1], b); atomicAdd(&P.stats[2], c); atomicAdd(&P.stats[3], a); }
    // point location: kd steps below the location grid - total, most by one point, and the total a warp PAYS (its slowest lane)
    unsigned wmax = kdSteps;
    for (int o2 = 16; o2 > 0; o2 >>= 1) wmax = max(wmax, __shfl_xor_sync(FULL, wmax, o2));
    atomicAdd(&P.stats[6], (unsigned long long)kdSteps); atomicMax(&P.stats[5], (unsigned long long)kdSteps);
    if ((threadIdx.x & 31) == 0) atomicAdd(&P.stats[4], (unsigned long long)wmax);
  }
}

// Tried and removed (round 2, profiles/perf_quick_r02_{f,g}_points_*.json): a persistent-lane variant (a warp owns a chunk of
// points in shared memory, every lane walks its own grid -> kd -> header -> record chain one dependent load per trip and takes
// the next point when done).  It cut the kd steps a warp pays from 11.9 to 3.5 per point but ran 1 M random points on C3 in
// 93-109 us (1 / 2 / 4 / 8 points per lane) against 50 us here: the query is bound by random 32-B DRAM sectors, not by the
// descent's divergence.

int launchSamplePoints(const SampleParams &P, cudaStream_t s)
{
  if (P.n == 0) return 0;
  const bool stats = P.stats != nullptr;
  const size_t perBlock = (size_t)EXA_POINTS_THREADS;
  const unsigned grid = (unsigned)((P.n + perBlock - 1) / perBlock);
#define LAUNCH_P(G, S_) { if (P.scene.dual) k_samplePoints<G, S_, true><<<grid, EXA_POINTS_THREADS, 0, s>>>(P); else k_samplePoints<G, S_, false><<<grid, EXA_POINTS_THREADS, 0, s>>>(P); }
  if (P.grad) { if (stats) LAUNCH_P(true, true) else LAUNCH_P(true, false) }
  else { if (stats) LAUNCH_P(false, true) else LAUNCH_P(false, false) }
#undef LAUNCH_P
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// streamlines: computeTraces / sampleDirection (exabrick.cu:945-963, 1531-1574)
// -----------------------------------------------------------------------------------------
template <bool DUAL>
__device__ __forceinline__ bool sampleDirection(const DeviceScene &S, float3 p, const int ch[3], float3 &out)
{
  // a point query that misses every region is an invalid sample (the reference indexes
  // region -1 there: undefined behaviour that is not copied)
  float v[3]; float3 g; unsigned a = 0, b = 0;
  for (int i = 0; i < 3; i++) if (!samplePointGlobal<false, false, DUAL>(S, p.x, p.y, p.z, ch[i], v[i], g, a, b)) return false;
  out = f3(v[0], v[1], v[2]); return true;
}

template <bool DUAL>
__global__ void k_trace(const __grid_constant__ TraceParams P)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.numTraces) return;
  for (int t = P.firstStep; t <= P.lastStep && t < P.numTimesteps; t++) {
    float *src = P.traces + 3 * ((size_t)i * P.numTimesteps + (t - 1)), *dst = src + 3;
    float3 p = f3(src[0], src[1], src[2]); const float3 pp = p;
    if (p.x < 2e10f) {
      bool valid = true; float3 k1 = f3(0, 0, 0), k2 = k1, k3 = k1, k4 = k1;
      valid &= sampleDirection<DUAL>(P.scene, p, P.channels, k1); k1 = k1 * P.steplen; const float3 p1 = p + k1 * .5f;
      valid &= sampleDirection<DUAL>(P.scene, p1, P.channels, k2); k2 = k2 * P.steplen; const float3 p2 = p + k2 * .5f;
      valid &= sampleDirection<DUAL>(P.scene, p2, P.channels, k3); k3 = k3 * P.steplen; const float3 p3 = p + k3;
      valid &= sampleDirection<DUAL>(P.scene, p3, P.channels, k4); k4 = k4 * P.steplen;
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
  if (P.scene.dual) k_trace<true><<<(P.numTraces + 63) / 64, 64, 0, s>>>(P); else k_trace<false><<<(P.numTraces + 63) / 64, 64, 0, s>>>(P);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// One box per trace around its visible cones (OPTIX_BOUNDS_PROGRAM(Streamline), exabrick.cu:532-570:
// segment k counts while k < currentTimestep-1 and both end points are alive; radius 2).  One
// thread per trace; lo > hi marks a trace without a visible cone.  Re-run after every tracer step
// (the reference rebuilds its streamline BVH there, OptixRenderer.cpp:539-543).
__global__ void k_traceBounds(const float *__restrict__ traces, int numTraces, int numTimesteps, int currentTimestep, float *__restrict__ bounds)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= numTraces) return;
  const int nSeg = min(numTimesteps - 1, currentTimestep - 1);
  float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
  const float *tr = traces + 3 * (size_t)i * numTimesteps;
  for (int k = 0; k < nSeg; k++) {
    const float *a = tr + 3 * k, *b = a + 3;
    if (!(a[0] < 2e10f && b[0] < 2e10f)) continue;
    for (int d = 0; d < 3; d++) {
      lo[d] = fminf(lo[d], fminf(a[d], b[d]) - 2.f);
      hi[d] = fmaxf(hi[d], fmaxf(a[d], b[d]) + 2.f);
    }
  }
  for (int d = 0; d < 3; d++) { bounds[6 * (size_t)i + d] = lo[d]; bounds[6 * (size_t)i + 3 + d] = hi[d]; }
}

int launchTraceBounds(const float *traces, int numTraces, int numTimesteps, int currentTimestep, float *bounds, cudaStream_t s)
{
  if (numTraces <= 0) return 0;
  k_traceBounds<<<(numTraces + 127) / 128, 128, 0, s>>>(traces, numTraces, numTimesteps, currentTimestep, bounds);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// Bin the visible cone segments into the uniform grid: one thread per (trace, segment).  fill = 0:
// counts[cell] += 1 for every cell the segment's box (end points -+ radius 2, grown a little)
// overlaps; fill = 1: list[cursor[cell]++] = segment id.  The host turns the counts into cell starts.
__global__ void k_coneGrid(const float *__restrict__ traces, int numTraces, int numTimesteps, int currentTimestep, int G,
                           float lox, float loy, float loz, float icx, float icy, float icz, int *countsOrCursor, int *list, int fill)
{
  const int nSeg = min(numTimesteps - 1, currentTimestep - 1);
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (nSeg <= 0 || g >= numTraces * nSeg) return;
  const int i = g / nSeg, k = g % nSeg;
  const int id = i * numTimesteps + k;
  const float *a = traces + 3 * (size_t)id, *b = a + 3;
  if (!(a[0] < 2e10f && b[0] < 2e10f)) return;
  const float lo[3] = {lox, loy, loz}, ic[3] = {icx, icy, icz};
  int c0[3], c1[3];
  for (int d = 0; d < 3; d++) {
    const float pad = 2.f + 1e-3f * (2.f + fabsf(a[d]) * 1e-3f) + 0.01f / ic[d];
    c0[d] = min(G - 1, max(0, (int)floorf((fminf(a[d], b[d]) - pad - lo[d]) * ic[d])));
    c1[d] = min(G - 1, max(0, (int)floorf((fmaxf(a[d], b[d]) + pad - lo[d]) * ic[d])));
  }
  for (int z = c0[2]; z <= c1[2]; z++)
    for (int y = c0[1]; y <= c1[1]; y++)
      for (int x = c0[0]; x <= c1[0]; x++) {
        const int c = x + G * (y + G * z);
        const int slot = atomicAdd(countsOrCursor + c, 1);
        if (fill) list[slot] = id;
      }
}

int launchConeGrid(const float *traces, int numTraces, int numTimesteps, int currentTimestep, int G, const float lo[3], const float invCell[3],
                   int *countsOrCursor, int *list, int fill, cudaStream_t s)
{
  const int nSeg = std::min(numTimesteps - 1, currentTimestep - 1);
  if (numTraces <= 0 || nSeg <= 0) return 0;
  const long long n = (long long)numTraces * nSeg;
  k_coneGrid<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(traces, numTraces, numTimesteps, currentTimestep, G, lo[0], lo[1], lo[2],
                                                        invCell[0], invCell[1], invCell[2], countsOrCursor, list, fill);
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

// One launch per XF / iso / partition change: every kd LEAF classifies its region (value range
// against the XF, iso values against the range) and, if active, marks itself and its
// ancestors.  A thread stops climbing at the first ancestor that is already marked: whoever
// marked it had found it unmarked (or raced with another marker) and keeps climbing, so the
// chain always reaches the root.  Both masks were cleared by one memset (they are adjacent).
__global__ void k_leafActiveUp(const __grid_constant__ RenderParams P, int numNodes, const int *kdParent, const uint8_t *inPart,
                               int spaceSkipping, uint8_t *volActive, uint8_t *isoActive, uint8_t *regionActive)
{
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= numNodes) return;
  const KdNode nd = P.scene.kd[g];
  const uint32_t payload = nd.info & 0x3fffffffu;
  if ((nd.info >> 30) != 3 || payload == KD_EMPTY) return;
  if (inPart && !inPart[g]) return;
  const int rid = (int)P.scene.blobs[payload + 2].x;
  const ExaBrickRegion r = P.scene.regions[rid];
  bool act = !spaceSkipping;
  for (int c = 0; c < P.numChannels && !act; c++) act |= activeForVolumeSampling(P, r.valueRange_lower, r.valueRange_upper, c);
  bool ia = false;
  for (int k = 0; k < EXA_MAX_ISO_SURFACES; k++)
    if (P.isoEnabled[k] && P.isoValue[k] >= r.valueRange_lower && P.isoValue[k] <= r.valueRange_upper) ia = true;
  volatile uint8_t *va = volActive, *ia8 = isoActive;
  if (act) { regionActive[rid] = 1; for (int n = g; n >= 0 && !va[n]; n = kdParent[n]) va[n] = 1; }
  if (ia) for (int n = g; n >= 0 && !ia8[n]; n = kdParent[n]) ia8[n] = 1;
}

int launchRegionActivity(const RenderParams &P, int numNodes, const int *kdParent, int spaceSkipping, const uint8_t *inPart,
                         uint8_t *volActive, uint8_t *isoActive, uint8_t *regionActive, cudaStream_t s)
{
  if (isoActive != volActive + numNodes || regionActive != volActive + 2 * (size_t)numNodes) return -1;   // one allocation, one memset
  cudaMemsetAsync(volActive, 0, 2 * (size_t)numNodes + (size_t)std::max(P.scene.numRegions, 0), s);
  if (numNodes <= 0 || P.scene.numRegions <= 0) return 0;
  k_leafActiveUp<<<(numNodes + 127) / 128, 128, 0, s>>>(P, numNodes, kdParent, inPart, spaceSkipping, volActive, isoActive, regionActive);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
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
// point-location grid (see DeviceScene::locGrid).  The cell box is grown by 1% of a cell on
// every side before it is tested against a split plane, which is orders of magnitude more
// than the rounding of the cell-index computation in locate(): every point that maps to a
// cell lies inside the grown box, so starting at the stored node gives exactly the leaf the
// root descent would reach.
// -----------------------------------------------------------------------------------------
__global__ void k_buildLocGrid(const KdNode *kd, int gx, int gy, int gz, float lox, float loy, float loz, float wx, float wy, float wz, uint2 *grid)
{
  const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= (size_t)gx * gy * gz) return;
  const int cx = (int)(c % gx), cy = (int)((c / gx) % gy), cz = (int)(c / ((size_t)gx * gy));
  const float lo[3] = {lox + ((float)cx - 0.01f) * wx, loy + ((float)cy - 0.01f) * wy, loz + ((float)cz - 0.01f) * wz};
  const float hi[3] = {lox + ((float)cx + 1.01f) * wx, loy + ((float)cy + 1.01f) * wy, loz + ((float)cz + 1.01f) * wz};
  int node = 0;
  for (;;) {
    const KdNode nd = kd[node];
    const uint32_t axis = nd.info >> 30;
    if (axis == 3) break;
    const int child = (int)(nd.info & 0x3fffffffu);
    if (hi[axis] < nd.split) node = child;                  // p < split for every p in the cell
    else if (lo[axis] >= nd.split) node = child + 1;        // p >= split for every p in the cell
    else break;
  }
  grid[c] = reinterpret_cast<const uint2 *>(kd)[node];
}

int launchBuildLocGrid(const KdNode *kd, const int G[3], const float lo[3], const float hi[3], uint2 *d_grid, float scaleOut[3], cudaStream_t s)
{
  float w[3];
  for (int d = 0; d < 3; d++) { w[d] = (hi[d] - lo[d]) / (float)G[d]; scaleOut[d] = (float)G[d] / (hi[d] - lo[d]); }
  const size_t n = (size_t)G[0] * G[1] * G[2];
  k_buildLocGrid<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(kd, G[0], G[1], G[2], lo[0], lo[1], lo[2], w[0], w[1], w[2], d_grid);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// dual-cell records: for every brick and every dual cell (i,j,k) in [-1,sx-1]x[-1,sy-1]x[-1,sz-1]
// the 8 corner scalars {s000,s001,s010,s011,s100,s101,s110,s111} (x fastest, 0 outside the
// brick) as one 32-byte record, so that one tent-basis brick visit reads ONE DRAM sector
// instead of >= 4.  8x the scalar memory: a capacity-for-bandwidth trade that 180 GB of HBM3e
// affords; the flat reference layout stays (ABI, value-range pass, fallback).
// -----------------------------------------------------------------------------------------
__global__ void k_expandDual(const ExaBrick *bricks, int numBricks, const uint32_t *dualBegin, uint64_t numRecords,
                             const float *scalars, float4 *dual)
{
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= numRecords) return;
  int lo = 0, hi = numBricks - 1;                            // last brick with dualBegin <= r
  while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if ((uint64_t)__ldg(dualBegin + mid) <= r) lo = mid; else hi = mid - 1; }
  const ExaBrick b = bricks[lo];
  const int sx = b.size[0], sy = b.size[1], sz = b.size[2], rx = sx + 1, ry = sy + 1;
  uint32_t q = (uint32_t)(r - dualBegin[lo]);
  const int i = (int)(q % rx) - 1; q /= rx;
  const int j = (int)(q % ry) - 1, k = (int)(q / ry) - 1;
  const float *base = scalars + b.begin;
  float v[8];
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const int x = i + (c & 1), y = j + ((c >> 1) & 1), z = k + (c >> 2);
    v[c] = (x >= 0 && x < sx && y >= 0 && y < sy && z >= 0 && z < sz) ? __ldg(base + x + (size_t)sx * y + (size_t)sx * sy * z) : 0.f;
  }
  dual[2 * r] = make_float4(v[0], v[1], v[2], v[3]);
  dual[2 * r + 1] = make_float4(v[4], v[5], v[6], v[7]);
}

int launchExpandDual(const ExaBrick *d_bricks, int numBricks, const uint32_t *d_dualBegin, uint64_t numRecords,
                     const float *d_scalars, float4 *d_dual, cudaStream_t s)
{
  if (numRecords == 0) return 0;
  k_expandDual<<<(unsigned)((numRecords + 255) / 256), 256, 0, s>>>(d_bricks, numBricks, d_dualBegin, numRecords, d_scalars, d_dual);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// sort-last compositing, fused with the frame's tail: for every pixel of this rank's span, read the
// partials of ALL ranks in visibility order (part[0] is the front-most; the pointers are peer
// memory of the other GPUs, read over NVLink - or local copies after an all-to-all), composite
// front to back, apply the early-termination rule (exabrick.cu:1180, 1694-1697) where the
// single-GPU ray would have met it at the latest, pick the nearest surface hit of all ranks as
// background, then background blend, accumulation, sRGB, RGBA8 pack (:1701-1719).
// -----------------------------------------------------------------------------------------
__global__ void k_compositeFinish(const __grid_constant__ RenderParams P, const __grid_constant__ CompositeParams C)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C.count) return;
  const size_t pixelIdx = C.first + i;
  const int px = (int)(pixelIdx % (size_t)P.width), py = (int)(pixelIdx / (size_t)P.width);
  float4 q[EXA_MAX_PARTS];
  unsigned covered = 0;                                       // parts whose rectangle holds this pixel: the others wrote nothing here and are not read
#pragma unroll
  for (int j = 0; j < EXA_MAX_PARTS; j++)                     // all loads in flight before the dependent chain
    if (j < C.n) {
      const bool in = px >= C.rect[j][0] && px <= C.rect[j][2] && py >= C.rect[j][1] && py <= C.rect[j][3];
      q[j] = in ? __ldcs(C.part[j] + pixelIdx) : make_float4(0.f, 0.f, 0.f, 0.f);
      covered |= in ? 1u << j : 0u;
    }
  float4 pc = make_float4(0, 0, 0, 0);
  int jStop = C.n - 1;                                        // last part that contributes: the first one with a surface hit of its own
  if (C.hasSurf)
    for (int j = 0; j < C.n; j++) { if (!(covered >> j & 1u)) continue; const float w = __ldcs(&C.surf[j][pixelIdx].w); if (w > 0.f && w < 1e30f) { jStop = j; break; } }
#pragma unroll
  for (int j = 0; j < EXA_MAX_PARTS; j++) {
    if (j >= C.n || j > jStop) break;
    float4 a = q[j];
    const bool term = a.w >= 2.f;
    if (term) a.w -= 2.f;
    const float k = 1.f - pc.w;
    pc = make_float4(pc.x + k * a.x, pc.y + k * a.y, pc.z + k * a.z, pc.w + k * a.w);
    if (term || pc.w >= P.ertThreshold) { pc = make_float4(pc.x * pc.w, pc.y * pc.w, pc.z * pc.w, 1.f); break; }
  }
  float3 bg = f3(0, 0, 0);
  if (C.hasSurf) {
    // background = the nearest surface.  Parts are convex and ordered along the ray, so the first part (front to back) with
    // a hit of its OWN (iso-surface, w = t > 0) holds the nearest one and hides every part behind it; a replicated surface
    // (w = -t) is reported identically by all parts, which have clipped their samples at it themselves.
    bool have = false;
    for (int j = 0; j < C.n && j <= jStop; j++) {
      if (!(covered >> j & 1u)) continue;
      const float4 sv = __ldcs(C.surf[j] + pixelIdx);
      if (sv.w >= 1e30f) continue;
      if (sv.w > 0.f) { bg = f3(sv.x, sv.y, sv.z); break; }
      if (!have) { bg = f3(sv.x, sv.y, sv.z); have = true; }
    }
  }
  float3 color = f3(pc.w * pc.x + (1.f - pc.w) * bg.x, pc.w * pc.y + (1.f - pc.w) * bg.y, pc.w * pc.z + (1.f - pc.w) * bg.z);   // :1701 (sic)
  if (P.frameID > 0) { const float4 acc = P.accum[pixelIdx]; color = color + f3(acc.x, acc.y, acc.z); }
  P.accum[pixelIdx] = make_float4(color.x, color.y, color.z, 1.f);
  const float inv = (float)P.frameID + 1.f;
  C.out[pixelIdx] = make_rgba8(f3(linear_to_srgb(color.x / inv), linear_to_srgb(color.y / inv), linear_to_srgb(color.z / inv)));
}

int launchCompositeFinish(const RenderParams &P, const CompositeParams &C, cudaStream_t s)
{
  if (!C.count) return 0;
  k_compositeFinish<<<(unsigned)((C.count + 255) / 256), 256, 0, s>>>(P, C);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// -----------------------------------------------------------------------------------------
// Barrier between the ranks of one NVSwitch box over peer memory, stream-ordered: lane k stores this rank's
// sequence number into slot `rank` of rank k's flag array (a peer pointer), then spins until slot k of the
// local array has reached it.  Everything the stream did before (pixels stored into other GPUs' frame
// strips, a partial image) completed with its kernel, so a rank that sees the flag sees that data.  ~2 us
// instead of a 4-byte NCCL all-reduce (64 us at 8 ranks); the spin is bounded, a missing rank sets *err.
// -----------------------------------------------------------------------------------------
__global__ void k_peerBarrier(PeerBarrierParams B)
{
  const int k = threadIdx.x;
  if (k >= B.n) return;
  __threadfence_system();
  volatile unsigned *remote = B.flags[k] + B.rank;
  *remote = B.seq;
  __threadfence_system();
  volatile unsigned *mine = B.flags[B.rank] + k;
  const long long t0 = clock64();
  while ((int)(*mine - B.seq) < 0) {
    if (clock64() - t0 > B.timeoutCycles) { atomicExch(B.err, 1u); break; }
    __nanosleep(100);
  }
  __threadfence_system();
}

int launchPeerBarrier(const PeerBarrierParams &B, cudaStream_t s)
{
  k_peerBarrier<<<1, 32, 0, s>>>(B);
  return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

} // namespace exa
