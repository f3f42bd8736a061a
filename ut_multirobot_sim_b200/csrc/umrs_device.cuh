// Device-side primitives of the scan core (sm_100a).  Citations are file:line
// in the reference tree (ut-amrl/ut_multirobot_sim).
//
// Exactness discipline: this translation unit is compiled with -fmad=false and
// WITHOUT --use_fast_math, so every float * + - / sqrtf below is one IEEE
// round-to-nearest operation, exactly like the reference built with
// -ffp-contract=off.  The places where the reference's expression promotes to
// double (atan2, cos/sin in vector_map.cpp, AngleMod's rint) use double here too.
// Where a cheaper evaluation is used it is explicit (__fmaf_rn, __fdividef,
// atan2f) and comes with the argument for why every DECISION the reference
// takes is unchanged.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/umrs.h"

namespace umrs {

constexpr double kPiD = 3.14159265358979323846;
constexpr double k2PiD = 2.0 * kPiD;                 // M_2PI, math_util.h:28-30
constexpr float k2PiF = 6.28318530717958647692f;     // T(M_2PI) for T = float
constexpr float kPiF = 3.14159265358979323846f;

// math_util::AngleMod<float>, math_util.h:55-59 (rint is the double C function
// in vector_map.cpp's translation unit; the product and subtraction promote).
__device__ __forceinline__ float angle_mod_f(float angle) {
  const float q = angle / k2PiF;
  const double k = rint((double)q);
  return (float)((double)angle - (double)k2PiF * k);
}
// math_util::AngleMod<double>
__device__ __forceinline__ double angle_mod_d(double angle) {
  return angle - k2PiD * rint(angle / k2PiD);
}

// Ray angle i of a fan: vector_map.cpp:465,467.
__device__ __forceinline__ float ray_angle(float angle_min, float da, int i) {
  return angle_mod_f(angle_min + (float)i * da);
}

// cos/sin of a ray angle as the reference evaluates them (vector_map.cpp:472:
// the double functions on a float angle, narrowed to float).  |a| <= pi after
// AngleMod, so the general-argument machinery of sincos() is not needed:
// Cody-Waite reduction by k*pi/2 (two-part constant, exact for |k| <= 2) and the
// fdlibm kernel polynomials on [-pi/4, pi/4] give both values to ~1 ulp in
// double, i.e. the same float after rounding as any other faithful double
// implementation (the GPU/glibc pair already in use agreed on every ray tested).
__device__ __forceinline__ void ray_dir(float a, float& rx, float& ry) {
  const float kf = rintf(a * 0.636619772367581343f);   // nearest multiple of pi/2
  const double k = (double)kf;
  double r = fma(-k, 1.57079632673412561417e+00, (double)a);
  r = fma(-k, 6.07710050650619224932e-11, r);
  const double z = r * r;
  double ps = fma(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08);
  ps = fma(z, ps, 2.75573137070700676789e-06);
  ps = fma(z, ps, -1.98412698298579493134e-04);
  ps = fma(z, ps, 8.33333333332248946124e-03);
  ps = fma(z, ps, -1.66666666666666324348e-01);
  const double sn = fma(z * r, ps, r);
  double pc = fma(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09);
  pc = fma(z, pc, -2.75573143513906633035e-07);
  pc = fma(z, pc, 2.48015872894767294178e-05);
  pc = fma(z, pc, -1.38888888888741095749e-03);
  pc = fma(z, pc, 4.16666666666666019037e-02);
  const double cs = 1.0 - fma(z, 0.5, -(z * z) * pc);
  const int q = (int)kf & 3;
  const double c = (q & 1) ? sn : cs, s = (q & 1) ? cs : sn;
  // q: 0 (s, c)  1 (c, -s)  2 (-s, -c)  3 (-c, s)   [as (sin, cos)]
  rx = (float)((q == 1 || q == 2) ? -c : c);
  ry = (float)((q == 2 || q == 3) ? -s : s);
}

// ---- per-fan ray lattice ----------------------------------------------------------
// da = span / num_rays (vector_map.cpp:465).  cull_ok: the fan is a proper one
// (positive step, at most one turn) so that a candidate's angular interval maps
// to a ray-index interval; otherwise every candidate is tested against every ray
// (still exact).  lazy_ok: endpoint angles may be taken from atan2f unless a ray
// lies within `th_idx` index units of them (see make_cast).
struct Lattice {
  float amin, da, inv_da;
  float period;      // rays per full turn
  int p_lo, p_hi;    // integer bracket of period, for the wrapped image of a span
  int num_rays;
  bool cull_ok, lazy_ok;
  float th_idx;
};

// Angular slack (rad) inside which an approximate endpoint angle could order
// differently against a ray angle than the reference's float(atan2(double)):
//   atan2f error <= 2 ulp + final rounding of the reference value 0.5 ulp,
//   ulp(pi) = 2.4e-7                                           -> 6.0e-7
//   ray angle vs its ideal lattice point (rounding of i*da, of the sum with
//   angle_min, of AngleMod)                                    -> 8.4e-7
//   float evaluation of (a - amin) mod 2pi and of u * inv_da at index <= 4096
//                                                              -> 2.2e-6
// sum 3.6e-6; 6e-6 is used.
constexpr float kAngleSlack = 6e-6f;

__device__ __forceinline__ Lattice make_lattice(float amin, float amax, int num_rays) {
  Lattice L;
  const float span = amax - amin;
  L.amin = amin;
  L.num_rays = num_rays;
  L.da = span / (float)num_rays;
  L.cull_ok = (L.da > 1e-7f) && (span <= 6.2832f);
  L.inv_da = L.cull_ok ? 1.0f / L.da : 0.f;
  L.period = L.cull_ok ? k2PiF * L.inv_da : 0.f;
  L.p_hi = (int)ceilf(L.period) + 1;
  L.p_lo = (int)floorf(L.period) - 1;
  L.th_idx = kAngleSlack * L.inv_da;
  L.lazy_ok = L.cull_ok && fabsf(amin) <= 8.0f && L.th_idx < 0.25f && num_rays <= 4096;
  return L;
}

// True when some ray of the lattice may lie within kAngleSlack of angle a.
__device__ __forceinline__ bool near_a_ray(const Lattice& L, float a) {
  float u = a - L.amin;
  u -= k2PiF * floorf(u * (1.0f / k2PiF));   // ~[0, 2pi]
  const float t = u * L.inv_da;
  const float d = fabsf(t - rintf(t));
  // ray 0 also sits at u == 2pi, i.e. t == period (not an integer in general)
  return d <= L.th_idx || fabsf(t - L.period) <= L.th_idx;
}

// float(atan2(double)) of vector_map.cpp:449-450; out of line (rare path, large body).
static __device__ __noinline__ float atan2_ref(float y, float x) {
  return (float)atan2((double)y, (double)x);
}

// Per-candidate constants of the ray fill (vector_map.cpp:444-459,471,473).
// wraps_around is not stored: after the reference's swap, a wrapping interval
// has a0 > a1 and a plain one a0 < a1 (|a0-a1| >= 1e-4 is guaranteed), so
// wraps == (a0 > a1).
struct SegCast {
  float a0, a1, nx, ny;  // one LDS.128 in the ray loop
  float c;               // n . p0'
  bool valid;
  float arc;             // CCW extent a0 -> a1 in [0, pi], for chunk culling only
};

// Input: the segment already translated by -loc (vector_map.cpp:447-448).
//
// Lazily exact angles.  The reference computes a0/a1 as float(atan2(double))
// (:449-450) and uses them ONLY in comparisons: |a0-a1| < 1e-4 (:451),
// |a1-a0| > pi (:452), a0 < a1 (:453) and a0 <= a <= a1 against ray angles
// (:469-470).  atan2f is within 6e-7 rad of that value; if neither threshold
// comparison nor any ray comparison can come out differently (no ray within
// kAngleSlack of either endpoint angle, both differences further than 4e-6
// from their thresholds) the cheap angles are stored -- every decision, hence
// every output bit, is the same.  Otherwise the double atan2 is evaluated.
// The range itself never depends on a0/a1 beyond the swap decision.
//
// Each endpoint is decided on its own: a threshold comparison within 4e-6 of
// its threshold makes BOTH exact; otherwise only an endpoint with a ray within
// kAngleSlack of it needs the exact value (the thresholds keep their margin when
// one angle moves by <= 6e-7, and the ray comparisons of the two endpoints are
// independent).
struct CastAngles {
  float a0, a1;        // atan2f values
  bool need0, need1;   // the reference's float(atan2(double)) must replace them
};

__device__ __forceinline__ CastAngles cast_angles(float p0x, float p0y, float p1x, float p1y,
                                                  const Lattice& L) {
  CastAngles r;
  r.a0 = atan2f(p0y, p0x);
  r.a1 = atan2f(p1y, p1x);
  const float dab = fabsf(r.a0 - r.a1);
  const bool both = !L.lazy_ok || fabsf(dab - 1e-4f) <= 4e-6f || fabsf(dab - kPiF) <= 4e-6f;
  r.need0 = both || near_a_ray(L, r.a0);
  r.need1 = both || near_a_ray(L, r.a1);
  return r;
}

// Replaces the flagged angles by float(atan2(double)): one call site, executed a
// second time only by lanes that need both (a diverged warp pays per call).
__device__ __forceinline__ void exact_angles(CastAngles& r, float p0x, float p0y, float p1x,
                                             float p1y) {
  for (int it = 0; it < 2; ++it) {
    if (!(r.need0 || r.need1)) break;
    const bool first = r.need0;
    const float v = atan2_ref(first ? p0y : p1y, first ? p0x : p1x);  // :449-450 (double atan2)
    if (first) { r.a0 = v; r.need0 = false; } else { r.a1 = v; r.need1 = false; }
  }
}

// vector_map.cpp:451-459,471,473 on settled angles.
__device__ __forceinline__ SegCast finish_cast(float a0, float a1, float p0x, float p0y, float p1x,
                                               float p1y) {
  SegCast k;
  k.valid = false;
  k.a0 = 0.f; k.a1 = 0.f; k.nx = 0.f; k.ny = 0.f; k.c = 0.f; k.arc = 0.f;
  if (fabs((double)(a0 - a1)) < 0.0001) return k;     // :451
  const bool wraps = fabs((double)(a1 - a0)) > kPiD;  // :452
  if ((wraps && a0 < a1) || (!wraps && a0 > a1)) {    // :453-457
    float t;
    t = a0; a0 = a1; a1 = t;
    t = p0x; p0x = p1x; p1x = t;
    t = p0y; p0y = p1y; p1y = t;
  }
  // UnitNormal = Perp(Dir()), line2d.h:62-69; Eigen normalized() divides.
  float dx = p1x - p0x, dy = p1y - p0y;
  const float z = dx * dx + dy * dy;
  if (z > 0.0f) {
    const float len = sqrtf(z);
    dx = dx / len;
    dy = dy / len;
  }
  k.a0 = a0;
  k.a1 = a1;
  k.nx = -dy;
  k.ny = dx;
  k.c = k.nx * p0x + k.ny * p0y;   // :473 n.dot(p0)
  k.valid = true;
  k.arc = wraps ? (a1 + k2PiF) - a0 : a1 - a0;
  return k;
}

__device__ __forceinline__ SegCast make_cast(float p0x, float p0y, float p1x, float p1y,
                                             const Lattice& L) {
  CastAngles r = cast_angles(p0x, p0y, p1x, p1y, L);
  exact_angles(r, p0x, p0y, p1x, p1y);
  return finish_cast(r.a0, r.a1, p0x, p0y, p1x, p1y);
}

// vector_map.cpp:469-470, inclusive at both ends.
__device__ __forceinline__ bool cast_covers(float a0, float a1, float a) {
  const bool p = a0 <= a, q = a1 >= a;
  return (a0 > a1) ? (p || q) : (p && q);
}

// GetSceneLines AABB cull, vector_map.cpp:195-204.
__device__ __forceinline__ bool aabb_keep(float4 s, float lx, float ly, float max_range) {
  const float x_min = lx - max_range, y_min = ly - max_range;
  const float x_max = lx + max_range, y_max = ly + max_range;
  if (s.x < x_min && s.z < x_min) return false;
  if (s.y < y_min && s.w < y_min) return false;
  if (s.x > x_max && s.z > x_max) return false;
  if (s.y > y_max && s.w > y_max) return false;
  return true;
}

// 1 / x to 1 ulp, one MUFU.RCP (no denormal-scaling wrapper around it).
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// atan2 to 2e-5 rad (Abramowitz & Stegun 4.4.49 on the octant-reduced ratio).
__device__ __forceinline__ float atan2_coarse(float y, float x) {
  const float ax = fabsf(x), ay = fabsf(y);
  const float z = fminf(ax, ay) * rcp_approx(fmaxf(ax, ay));
  const float z2 = z * z;
  float r = __fmaf_rn(z2, 0.0208351f, -0.0851330f);
  r = __fmaf_rn(z2, r, 0.1801410f);
  r = __fmaf_rn(z2, r, -0.3302995f);
  r = __fmaf_rn(z2, r, 0.9998660f);
  r *= z;
  if (ay > ax) r = 1.57079633f - r;
  if (x < 0.f) r = 3.14159265f - r;
  return y < 0.f ? -r : r;
}

// ---- Philox4x32-10 (counter-based; replaces std::default_random_engine,
// simulator.h:99-100, ackermann_model.h:17-18) ------------------------------
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// Box-Muller on two words -> two N(0,1).  The angle is 2*pi*y - pi in (-pi, pi)
// (same distribution as 2*pi*y; keeps the fast sincos in its accurate range).
// The noise is checked distributionally, not bitwise, so the hardware
// log/sin/cos approximations are used (absolute error ~4e-7 each).
__device__ __forceinline__ float2 box_muller(uint32_t u0, uint32_t u1) {
  const float x = ((float)(u0 >> 8) + 0.5f) * (1.0f / 16777216.0f);
  const float y = ((float)(u1 >> 8) + 0.5f) * (1.0f / 16777216.0f);
  const float r = sqrtf(-2.0f * __logf(x));
  float s, c;
  __sincosf(k2PiF * y - kPiF, &s, &c);
  return make_float2(r * c, r * s);
}

// Noise streams (counter.w): which consumer draws.
constexpr uint32_t kStreamLaser = 0x4C415352u;   // per-ray range noise
constexpr uint32_t kStreamMotion = 0x4D4F5445u;  // Ackermann angular error

}  // namespace umrs
