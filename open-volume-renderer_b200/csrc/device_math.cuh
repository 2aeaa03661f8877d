// Device-side math shared by the B200 volume kernels (sm_100a). Semantics follow the reference's gdt helpers so
// that results stay pixel-matched: NaN-suppressing min/max clamps, normalize = v*1/sqrt(dot),
// xfmPoint as nested madd. Parity-critical expressions use __f*_rn intrinsics to stop nvcc from contracting them
// into FMAs the reference arithmetic does not have.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>

namespace b200k {

struct f3 { float x, y, z; };
struct i3 { int x, y, z; };

__host__ __device__ __forceinline__ f3 mk3(float a, float b, float c) { f3 r; r.x = a; r.y = b; r.z = c; return r; }
__host__ __device__ __forceinline__ f3 mk3(float a) { return mk3(a, a, a); }
__host__ __device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
__host__ __device__ __forceinline__ f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__host__ __device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
__host__ __device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
__host__ __device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ f3 cross3(f3 a, f3 b)
{
  return mk3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
__host__ __device__ __forceinline__ f3 normalize3(f3 v) { return (v * 1.f) / sqrtf(dot3(v, v)); }
__host__ __device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(hi, fmaxf(lo, v)); }
__host__ __device__ __forceinline__ float clamp01(float v) { return fminf(1.f, fmaxf(0.f, v)); }
__host__ __device__ __forceinline__ f3 clamp01(f3 v) { return mk3(clamp01(v.x), clamp01(v.y), clamp01(v.z)); }

struct lin3 { f3 vx, vy, vz; };          // columns
struct aff3 { lin3 l; f3 p; };

__host__ __device__ __forceinline__ f3 xfm_vector(const lin3& l, f3 v) { return v.x * l.vx + (v.y * l.vy + v.z * l.vz); } // gdt madd order
__host__ __device__ __forceinline__ f3 xfm_point(const aff3& m, f3 p)
{
  return mk3(p.x) * m.l.vx + (mk3(p.y) * m.l.vy + (mk3(p.z) * m.l.vz + m.p));
}
__host__ __device__ __forceinline__ lin3 transposed(const lin3& m)
{
  lin3 r;
  r.vx = mk3(m.vx.x, m.vy.x, m.vz.x);
  r.vy = mk3(m.vx.y, m.vy.y, m.vz.y);
  r.vz = mk3(m.vx.z, m.vy.z, m.vz.z);
  return r;
}
__host__ __device__ __forceinline__ lin3 inverse(const lin3& m)
{
  const float det = dot3(m.vx, cross3(m.vy, m.vz));
  lin3 a;
  a.vx = cross3(m.vy, m.vz);
  a.vy = cross3(m.vz, m.vx);
  a.vz = cross3(m.vx, m.vy);
  a = transposed(a);
  lin3 r;
  r.vx = a.vx / det;
  r.vy = a.vy / det;
  r.vz = a.vz / det;
  return r;
}
__host__ __device__ __forceinline__ aff3 inverse(const aff3& m)
{
  aff3 r;
  r.l = inverse(m.l);
  r.p = -((m.p.x * r.l.vx + m.p.y * r.l.vy) + m.p.z * r.l.vz); // rcp(AffineSpace): -(il * a.p)
  return r;
}
__host__ __device__ __forceinline__ aff3 from_3x4(const float* m)
{
  aff3 a;
  a.l.vx = mk3(m[0], m[4], m[8]);
  a.l.vy = mk3(m[1], m[5], m[9]);
  a.l.vz = mk3(m[2], m[6], m[10]);
  a.p = mk3(m[3], m[7], m[11]);
  return a;
}

// RandomTEA (reference: ovr/common/random/random.h:146-188): 16 rounds, state advanced in place, two floats per draw
struct Tea {
  uint32_t v0, v1;
  __device__ __forceinline__ Tea(uint32_t idx, uint32_t seed) : v0(idx), v1(seed) {}
  __device__ __forceinline__ float2 draw()
  {
    uint32_t s0 = v0, s1 = v1, sum = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
      sum += 0x9e3779b9u;
      s0 += ((s1 << 4) + 0xa341316cu) ^ (s1 + sum) ^ ((s1 >> 5) + 0xc8013ea4u);
      s1 += ((s0 << 4) + 0xad90777du) ^ (s0 + sum) ^ ((s0 >> 5) + 0x7e95761eu);
    }
    v0 = s0;
    v1 = s1;
    const float tofloat = 2.3283064365386962890625e-10f;
    return make_float2(__uint2float_rn(v0) * tofloat, __uint2float_rn(v1) * tofloat);
  }
};

// intersect_box (reference: ovr/devices/optix7/shaders_common.h:156-184): slab test with __frcp_rn and FLT_MIN guards
__device__ __forceinline__ bool intersect_box(float& t0, float& t1, f3 org, f3 dir, f3 lower, f3 upper)
{
  const bool sx = fabsf(dir.x) < FLT_MIN, sy = fabsf(dir.y) < FLT_MIN, sz = fabsf(dir.z) < FLT_MIN;
  const f3 rcp = mk3(__frcp_rn(dir.x), __frcp_rn(dir.y), __frcp_rn(dir.z));
  const f3 lo = mk3(sx ? FLT_MAX : __fmul_rn(lower.x - org.x, rcp.x), sy ? FLT_MAX : __fmul_rn(lower.y - org.y, rcp.y),
                    sz ? FLT_MAX : __fmul_rn(lower.z - org.z, rcp.z));
  const f3 hi = mk3(sx ? -FLT_MAX : __fmul_rn(upper.x - org.x, rcp.x), sy ? -FLT_MAX : __fmul_rn(upper.y - org.y, rcp.y),
                    sz ? -FLT_MAX : __fmul_rn(upper.z - org.z, rcp.z));
  const float nr = fmaxf(fmaxf(fminf(lo.x, hi.x), fminf(lo.y, hi.y)), fminf(lo.z, hi.z));
  const float fr = fminf(fminf(fmaxf(lo.x, hi.x), fmaxf(lo.y, hi.y)), fmaxf(lo.z, hi.z));
  t0 = fmaxf(t0, nr);
  t1 = fminf(t1, fr);
  return t1 > t0;
}

} // namespace b200k
