// The reference's analytic scene render, restated for host and device ("tier A"):
//   TrimOcclusion            vector_map.cpp:55-189
//   VectorMap::SceneRender   vector_map.cpp:217-258
//   Line<float> primitives   line2d.h:54-60,123-137,180-200, geometry::Cross geometry.h:56-59
// Used by the opt-in reference-render scan (umrs_scan_batch_reference_render), whose output
// is bit-identical to the reference's own GetPredictedScan -- including the occlusion-trimming
// defects the default (exact nearest hit) path does not have.
//
// Every float expression below is the reference's, operation for operation; the file must be
// compiled without FMA contraction (nvcc -fmad=false, g++ -ffp-contract=off).  No libm here.
#pragma once

#ifdef __CUDACC__
#define UMRS_HD __host__ __device__ __forceinline__
#else
#define UMRS_HD inline
#endif

namespace umrs {
namespace render {

struct V2 { float x, y; };
struct Seg { V2 p0, p1; };   // same 16 bytes as a float4 {x0, y0, x1, y1}

UMRS_HD V2 sub(V2 a, V2 b) { V2 r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }
UMRS_HD float cross2(V2 a, V2 b) { return a.x * b.y - b.x * a.y; }   // geometry.h:58
UMRS_HD float dot2(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
UMRS_HD float sq_norm(V2 a) { return a.x * a.x + a.y * a.y; }
UMRS_HD float sq_length(const Seg& s) { return sq_norm(sub(s.p0, s.p1)); }   // line2d.h:58-60
UMRS_HD float fmin2(float a, float b) { return b < a ? b : a; }   // std::min
UMRS_HD float fmax2(float a, float b) { return a < b ? b : a; }   // std::max

// Line::RayIntersects, line2d.h:180-187
UMRS_HD bool ray_intersects(const Seg& s, V2 p, V2 dir) {
  V2 v0 = sub(s.p0, p), v1 = sub(s.p1, p);
  if (cross2(v0, v1) < 0.0f) { const V2 t = v0; v0 = v1; v1 = t; }
  return cross2(v0, dir) >= 0.0f && cross2(v1, dir) <= 0.0f;
}

// Line::Touches, line2d.h:189-193
UMRS_HD bool touches(const Seg& s, V2 p) {
  const V2 v0 = sub(s.p0, p), v1 = sub(s.p1, p);
  return cross2(v0, v1) == 0.0f && dot2(v0, v1) < 0.0f;
}

// Line::Intersects(p2, p3), line2d.h:123-137
UMRS_HD bool intersects(const Seg& s, V2 p2, V2 p3) {
  if (fmin2(s.p0.x, s.p1.x) > fmax2(p2.x, p3.x)) return false;
  if (fmax2(s.p0.x, s.p1.x) < fmin2(p2.x, p3.x)) return false;
  if (fmin2(s.p0.y, s.p1.y) > fmax2(p2.y, p3.y)) return false;
  if (fmax2(s.p0.y, s.p1.y) < fmin2(p2.y, p3.y)) return false;
  const V2 d1 = sub(s.p1, s.p0), d2 = sub(p3, p2);
  if (cross2(d1, sub(p3, s.p0)) * cross2(d1, sub(p2, s.p0)) > 0.0f) return false;
  if (cross2(d2, sub(s.p1, p2)) * cross2(d2, sub(s.p0, p2)) > 0.0f) return false;
  return true;
}

// Line::RayIntersection, line2d.h:195-200
UMRS_HD V2 ray_intersection(const Seg& s, V2 p, V2 dir) {
  const V2 p0p = sub(s.p0, p), p1p = sub(s.p1, p);
  const float alpha = cross2(p1p, p0p) / cross2(dir, sub(s.p0, s.p1));
  V2 r;
  r.x = p.x + alpha * dir.x;
  r.y = p.y + alpha * dir.y;
  return r;
}

// TrimOcclusion, vector_map.cpp:55-189: trims `trim` by what `test` hides of it as seen from
// loc.  Returns true when a piece was split off (the reference push_backs it onto the list
// it is iterating, :174); the piece is then in *pushed.
UMRS_HD bool trim_occlusion(V2 loc, const Seg& test, Seg* trim, Seg* pushed) {
  const float sqeps = 1e-8f;                      // :62
  // end points (ta, tb of `test`; ua, ub of `trim`) and the same relative to loc (ra, rb; sa, sb)
  V2 ta = test.p0, tb = test.p1;                  // :65-67
  V2 ra = sub(ta, loc), rb = sub(tb, loc);
  V2 ua = trim->p0, ub = trim->p1;                // :73-75
  V2 sa = sub(ua, loc), sb = sub(ub, loc);
  // each pair in right-handed order as seen from loc, :82-89
  if (cross2(ra, rb) < 0.0f) {
    V2 t = ra; ra = rb; rb = t;
    t = ta; ta = tb; tb = t;
  }
  if (cross2(sa, sb) < 0.0f) {
    V2 t = sa; sa = sb; sb = t;
    t = ua; ua = ub; ub = t;
  }
  // no interaction, :98-105
  if ((cross2(ra, sa) >= 0.0f && cross2(rb, sa) >= 0.0f) ||
      (cross2(sa, ra) >= 0.0f && cross2(sb, ra) >= 0.0f)) {
    return false;
  }
  const bool ray_occlusion0 = ray_intersects(*trim, loc, ra);   // :109
  const bool ray_occlusion1 = ray_intersects(*trim, loc, rb);   // :112
  const bool complete = intersects(test, loc, ua) && intersects(test, loc, ub);   // :121-122
  const bool occlusion0 = ray_occlusion0 &&                        // :125-127
      (touches(*trim, loc) || touches(*trim, ta) || !intersects(*trim, loc, ta));
  const bool occlusion1 = ray_occlusion1 &&                        // :130-132
      (touches(*trim, loc) || touches(*trim, tb) || !intersects(*trim, loc, tb));
  if (complete) {                                                  // :143-149
    const float z = (float)0.1;
    trim->p0.x = z; trim->p0.y = z; trim->p1.x = z; trim->p1.y = z;
    return false;
  }
  if (occlusion0 && occlusion1) {                                  // :150-170
    const V2 right_end = ray_intersection(*trim, loc, ra);
    const V2 left_end = ray_intersection(*trim, loc, rb);
    trim->p0 = ua;
    trim->p1 = right_end;
    if (sq_norm(sub(left_end, ub)) > sqeps) {
      pushed->p0 = left_end;
      pushed->p1 = ub;
      return true;
    }
    return false;
  }
  if (occlusion0) {                                                // :171-175
    const V2 right_end = ray_intersection(*trim, loc, ra);
    trim->p0 = ua;
    trim->p1 = right_end;
  } else if (occlusion1) {                                         // :176-180
    const V2 left_end = ray_intersection(*trim, loc, rb);
    trim->p0 = left_end;
    trim->p1 = ub;
  }
  return false;
}

constexpr float kRenderEps = 0.0001f;     // SceneRender's eps on squared lengths, :220
constexpr int kRenderMaxLines = 2000;     // :221; entries past it are never read

// SceneRender, :217-258, sequentially (host tests; the device kernel runs the same steps
// with a warp, see k_scan_reference_render).  `lines` holds n_lines candidates from
// GetSceneLines and room for kRenderMaxLines; `scene` room for kRenderMaxLines.
// Returns the scene size (entries with sq_length <= eps are dead and must be skipped).
UMRS_HD int scene_render_sequential(V2 loc, Seg* lines, int n_lines, Seg* scene) {
  int n_scene = 0;
  for (int i = 0; i < n_lines && i < kRenderMaxLines; ++i) {
    Seg cur = lines[i];
    Seg piece;
    for (int j = 0; j < n_scene && sq_length(cur) >= kRenderEps; ++j) {
      if (sq_length(scene[j]) < kRenderEps) continue;
      if (trim_occlusion(loc, scene[j], &cur, &piece)) {
        if (n_lines < kRenderMaxLines) lines[n_lines] = piece;
        ++n_lines;
      }
    }
    if (sq_length(cur) > kRenderEps) {
      for (int j = 0; j < n_scene; ++j) {
        if (sq_length(scene[j]) < kRenderEps) continue;
        if (trim_occlusion(loc, cur, &scene[j], &piece)) {
          if (n_lines < kRenderMaxLines) lines[n_lines] = piece;
          ++n_lines;
        }
      }
      scene[n_scene++] = cur;
    }
  }
  return n_scene;
}

}  // namespace render
}  // namespace umrs
