// TEST INFRASTRUCTURE ONLY -- CPU oracle for the B200 volume device.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// Small fp32 vector / affine helpers restating the gdt semantics the reference shaders rely on
// (all citations relative to the reference tree):
//   * device min/max are CUDA fminf/fmaxf (NaN-suppressing)      extern/gdt/gdt/gdt.h:118-137
//   * clamp(v,lo,hi) = min(hi, max(lo, v))                       extern/gdt/gdt/math/vec/functors.h:55-57
//   * normalize(v)   = v * 1.f / sqrt(dot(v,v))                  extern/gdt/gdt/math/vec.h:443-448
//   * xfmPoint  = nested madd                                    extern/gdt/gdt/math/mat/AffineSpace.h:133
//   * xfmNormal = transpose(inverse(l)) * n, inverse=adjoint/det extern/gdt/gdt/math/mat/LinearSpace.h:215-221,321
#pragma once
#include <cmath>
#include <cstdint>
#include <cfloat>

namespace ovr_oracle {

struct vec3 { float x, y, z; };
struct vec3i_t { int x, y, z; };

static inline vec3 v3(float a, float b, float c) { return vec3{ a, b, c }; }
static inline vec3 v3(float a) { return vec3{ a, a, a }; }
static inline vec3 operator+(vec3 a, vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline vec3 operator-(vec3 a, vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline vec3 operator*(vec3 a, vec3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline vec3 operator/(vec3 a, vec3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline vec3 operator*(float s, vec3 a) { return v3(s * a.x, s * a.y, s * a.z); }
static inline vec3 operator*(vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline vec3 operator/(vec3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
static inline vec3 operator-(vec3 a) { return v3(-a.x, -a.y, -a.z); }
static inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline vec3 cross(vec3 a, vec3 b)
{
  return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
static inline vec3 normalize(vec3 v) { return (v * 1.f) / std::sqrt(dot(v, v)); }

// CUDA fminf/fmaxf: if one operand is NaN the other is returned.
static inline float cmin(float a, float b) { return std::fmin(a, b); }
static inline float cmax(float a, float b) { return std::fmax(a, b); }
static inline float cclamp(float v, float lo, float hi) { return cmin(hi, cmax(lo, v)); }
static inline float clamp01(float v) { return cclamp(v, 0.f, 1.f); }
static inline vec3 clamp01(vec3 v) { return v3(clamp01(v.x), clamp01(v.y), clamp01(v.z)); }

struct linear3 { vec3 vx, vy, vz; };      // columns
struct affine3 { linear3 l; vec3 p; };

static inline vec3 xfmVector(const linear3& l, vec3 v)
{
  // gdt xfmVector (LinearSpace.h:320): madd(a.x, vx, madd(a.y, vy, a.z*vz))
  return v.x * l.vx + (v.y * l.vy + v.z * l.vz);
}
static inline vec3 madd(vec3 a, vec3 b, vec3 c) { return a * b + c; }
static inline vec3 xfmPoint(const affine3& m, vec3 p)
{
  return madd(v3(p.x), m.l.vx, madd(v3(p.y), m.l.vy, madd(v3(p.z), m.l.vz, m.p)));
}
static inline linear3 transposed(const linear3& m)
{
  return linear3{ v3(m.vx.x, m.vy.x, m.vz.x), v3(m.vx.y, m.vy.y, m.vz.y), v3(m.vx.z, m.vy.z, m.vz.z) };
}
static inline linear3 inverse(const linear3& m)
{
  const float det = dot(m.vx, cross(m.vy, m.vz));
  const linear3 adj = transposed(linear3{ cross(m.vy, m.vz), cross(m.vz, m.vx), cross(m.vx, m.vy) });
  return linear3{ adj.vx / det, adj.vy / det, adj.vz / det };
}
static inline vec3 xfmNormal(const linear3& l, vec3 n) { return xfmVector(transposed(inverse(l)), n); }
static inline affine3 inverse(const affine3& m)
{
  const linear3 inv = inverse(m.l);
  affine3 r;
  r.l = inv;
  r.p = -((m.p.x * inv.vx + m.p.y * inv.vy) + m.p.z * inv.vz); // rcp(AffineSpace): -(il * a.p), LinearSpace operator*
  return r;
}
// 3x4 row-major (the OptiX instance layout, ovr/devices/optix7/volume.cpp:25-40) -> affine3
static inline affine3 from_3x4(const float* m)
{
  affine3 a;
  a.l.vx = v3(m[0], m[4], m[8]);
  a.l.vy = v3(m[1], m[5], m[9]);
  a.l.vz = v3(m[2], m[6], m[10]);
  a.p = v3(m[3], m[7], m[11]);
  return a;
}

} // namespace ovr_oracle
