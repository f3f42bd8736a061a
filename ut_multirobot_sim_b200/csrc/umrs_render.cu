// Opt-in reference render on the device: VectorMap::GetPredictedScan exactly as the
// reference computes it (vector_map.cpp:420-478) -- GetSceneLines, the O(L^2) SceneRender /
// TrimOcclusion visibility pass, then "first rendered piece that covers the ray" -- so that
// the output is bit-identical to the reference's, occlusion-trimming defects included
// (SURVEY.md 8c tier A).  The default scan (umrs_kernels.cu) is the exact nearest hit instead.
//
// One warp per fan for big batches, one 4-warp CTA per fan for small ones (a third of the
// latency).  The render is sequential by nature (every candidate is trimmed by the
// scene built so far, pieces split off are appended to the list being iterated), so the group
// reproduces the reference's order of events:
//   * trimming the current line by the scene (:231-234) changes the line, so the 32 lanes test
//     32 consecutive scene entries against the CURRENT line, the lowest lane whose entry
//     changes anything is applied, and the scan resumes behind it with the new line --
//     entries that change nothing are no-ops in the reference's loop as well;
//   * trimming the scene by the current line (:237-240) is independent per entry; pieces
//     split off are appended in entry order (ballot prefix);
//   * GetSceneLines and the LineCast list (:439-456) are ordered compactions.
// Work.  TrimOcclusion's first test (vector_map.cpp:98-105) returns "no interaction" -- both lines
// untouched -- whenever the two segments' angular intervals as seen from loc are disjoint, and
// that is the outcome of >90 % of the (line, scene entry) pairs.  Every scene entry therefore
// keeps a coarse angular interval {lo, span} of the vectors the reference itself crosses (end
// point - loc, right-handed); per candidate line the group first compacts, IN SCENE ORDER, the
// entries whose interval comes within 2e-3 rad of the line's, and runs the reference's two
// loops over that list only.  Skipped pairs are exactly those the reference's own test would
// dismiss (proof at seg_interval / intervals_may_touch), so order of events and every bit are kept.
//
// Storage.  The scene is what the O(L^2) pass reads over and over: its first kSceneSmem
// entries per warp (256; a 1000-segment floor plan renders to 50-150) live in shared
// memory, the rest -- and the candidate list, which is read once, in order, one entry ahead --
// in per-fan global scratch.  The cast list of the ray fill is built in place over the scene.
#include "umrs_device.cuh"
#include "umrs_internal.h"
#include "umrs_render.h"

namespace umrs {

using render::Seg;
using render::V2;

constexpr int kRenderWarpsPerCta = 4;
constexpr int kSceneSmem = 256;                 // scene / cast entries per warp kept in shared memory
constexpr int kRenderCtaPerFanUpTo = 148 * 6;   // one resident wave of 128-thread CTAs (80 registers)

__device__ __forceinline__ Seg load_seg(const float4* p) {
  const float4 v = *p;
  Seg s;
  s.p0.x = v.x; s.p0.y = v.y; s.p1.x = v.z; s.p1.y = v.w;
  return s;
}
__device__ __forceinline__ void store_seg(float4* p, const Seg& s) {
  *p = make_float4(s.p0.x, s.p0.y, s.p1.x, s.p1.y);
}
__device__ __forceinline__ Seg shfl_seg(const Seg& s, int src) {
  Seg r;
  r.p0.x = __shfl_sync(0xffffffffu, s.p0.x, src);
  r.p0.y = __shfl_sync(0xffffffffu, s.p0.y, src);
  r.p1.x = __shfl_sync(0xffffffffu, s.p1.x, src);
  r.p1.y = __shfl_sync(0xffffffffu, s.p1.y, src);
  return r;
}
__device__ __forceinline__ bool same_bits(const Seg& a, const Seg& b) {
  return __float_as_uint(a.p0.x) == __float_as_uint(b.p0.x) && __float_as_uint(a.p0.y) == __float_as_uint(b.p0.y) &&
         __float_as_uint(a.p1.x) == __float_as_uint(b.p1.x) && __float_as_uint(a.p1.y) == __float_as_uint(b.p1.y);
}

// A fan is rendered by a GROUP = one CTA of kW warps: kW = 1 (best throughput for big batches)
// or kW = 4 (a third of the latency for a single scan or a small batch).  The helpers below
// are the group-wide forms of ballot / prefix.
struct GroupShared {
  int first[4];     // per warp: lowest acting lane (32: none)
  int count[4];     // per warp: lanes that append
  float4 seg;       // the line a group adopts after an acting entry
  int flag;
};

template <int kW>
__device__ __forceinline__ void group_sync() {
  if (kW == 1) __syncwarp(); else __syncthreads();
}

// Ordered append: group thread order (GetSceneLines :199-214, the pieces of :174, the casts).
// Returns this thread's position; *total = entries appended by the group.
template <int kW>
__device__ __forceinline__ int group_append_pos(GroupShared* gs, bool keep, int base, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  int before = __popc(m & ((1u << lane) - 1u));
  if (kW == 1) {
    *total = __popc(m);
    return base + before;
  }
  if (lane == 0) gs->count[warp] = __popc(m);
  __syncthreads();
  int t = 0;
  for (int w = 0; w < kW; ++w) {
    const int c = gs->count[w];
    if (w < warp) before += c;
    t += c;
  }
  __syncthreads();          // count[] is reused by the next call
  *total = t;
  return base + before;
}

// Lowest group thread whose `acts` is set (-1: none).
template <int kW>
__device__ __forceinline__ int group_first(GroupShared* gs, bool acts) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned m = __ballot_sync(0xffffffffu, acts);
  if (kW == 1) return m ? __ffs(m) - 1 : -1;
  if (lane == 0) gs->first[warp] = m ? __ffs(m) - 1 : 32;
  __syncthreads();
  int r = -1;
  for (int w = kW - 1; w >= 0; --w) {
    const int f = gs->first[w];
    if (f < 32) r = w * 32 + f;
  }
  __syncthreads();
  return r;
}

// ---- angular intervals -------------------------------------------------------------------
// {lo, span} in units of 2^-32 turn: direction of the right-hand end point as seen from loc and
// the CCW extent to the other one (TrimOcclusion orders its end points the same way, :82-89).
// Computed from the very float vectors the reference crosses (x - loc), with a 2e-5 rad polynomial
// atan2.  span == kSpanUnknown (zero vector, NaN, a line through loc): "may touch" everything.
// Unsigned arithmetic makes the wrap-around free: lo2 - lo1 IS the CCW distance from lo1 to lo2.
constexpr uint32_t kSpanUnknown = 0xffffffffu;
constexpr float kTurnToU32 = 4294967296.0f / k2PiF;
constexpr uint32_t kIvMargin = 1400000u;   // 2.05e-3 rad: 100x the angle errors (4 x 2e-5) + cross-product rounding
__device__ __forceinline__ uint2 seg_interval(V2 loc, const Seg& s) {
  V2 a = render::sub(s.p0, loc), b = render::sub(s.p1, loc);
  if (render::cross2(a, b) < 0.0f) { const V2 t = a; a = b; b = t; }
  const float lo = atan2_coarse(a.y, a.x), hi = atan2_coarse(b.y, b.x);
  float span = hi - lo;
  if (span < 0.f) span += k2PiF;
  if (!(span <= 3.15f)) return make_uint2(0u, kSpanUnknown);     // also NaN
  // angle -> [0, 2^32): lo in (-pi, pi] -> lo + 2 pi if negative; float -> uint32 saturates, never traps
  const float lo_pos = lo < 0.f ? lo + k2PiF : lo;
  return make_uint2((uint32_t)(lo_pos * kTurnToU32), (uint32_t)(span * kTurnToU32));
}
// False only when the intervals are certainly disjoint with kIvMargin to spare.  Then, going CCW
// round loc, one meets A, a gap g1, S, a gap g2: |A| + g1 + |S| + g2 = one turn, so one of
// d12 = |A| + g1 and d21 = |S| + g2 is at most half a turn -- which is the reference's first (or
// second) "no interaction" clause: cross(ra, sa) >= 0 && cross(rb, sa) >= 0, i.e. sa lies within
// half a turn CCW of both ra and rb (:98-105).  The clause that holds must hold robustly: its two
// angles (d12 and g1, or d21 and g2) stay kIvMargin clear of 0 and of half a turn, where the sign
// of a cross product is rounding; so the exact float test takes the same branch.
__device__ __forceinline__ bool intervals_may_touch(uint2 a, uint2 b) {
  if (a.y == kSpanUnknown || b.y == kSpanUnknown) return true;
  const uint32_t d12 = b.x - a.x, d21 = a.x - b.x;                 // CCW distances lo1 -> lo2, lo2 -> lo1
  const uint32_t off_half = d12 > 0x80000000u ? d12 - 0x80000000u : 0x80000000u - d12;
  const bool disjoint = d12 > a.y + kIvMargin && d21 > b.y + kIvMargin && off_half > kIvMargin;
  return !disjoint;
}

// The scene (then the cast list) of one fan: entries below `cap` in shared memory, the rest in
// the fan's global scratch.
struct SceneStore {
  float4* sm;
  float* sm_c;
  uint2* iv;       // [cap] angular interval of entry j as it was appended (entries only shrink)
  uint16_t* list;  // [cap] scene entries that may touch the current line, in scene order
  float4* gl;
  float* gl_c;
  int cap;
  uint2* gl_iv;
  uint16_t* gl_list;
  int list_cap;    // list entries kept in shared memory
  __device__ __forceinline__ uint2 ld_iv(int j) const { return j < cap ? iv[j] : gl_iv[j]; }
  __device__ __forceinline__ void st_iv(int j, uint2 v) const { if (j < cap) iv[j] = v; else gl_iv[j] = v; }
  __device__ __forceinline__ int ld_list(int k) const { return (int)(k < list_cap ? list[k] : gl_list[k]); }
  __device__ __forceinline__ void st_list(int k, int j) const { if (k < list_cap) list[k] = (uint16_t)j; else gl_list[k] = (uint16_t)j; }
  __device__ __forceinline__ Seg ld(int j) const { return load_seg(j < cap ? sm + j : gl + j); }
  __device__ __forceinline__ void st(int j, const Seg& t) const { store_seg(j < cap ? sm + j : gl + j, t); }
  __device__ __forceinline__ float4 ld_cast(int j) const { return j < cap ? sm[j] : gl[j]; }
  __device__ __forceinline__ float ld_c(int j) const { return j < cap ? sm_c[j] : gl_c[j]; }
  __device__ __forceinline__ void st_cast(int j, float4 k, float c) const {
    if (j < cap) { sm[j] = k; sm_c[j] = c; } else { gl[j] = k; gl_c[j] = c; }
  }
};

// One fan, rendered by a group of kW warps (threads gt = 0 .. 32 kW - 1 of the caller's CTA).
// `slot` picks the fan's global scratch; sm / sm_c / iv / list are the group's kSceneSmem * kW
// shared-memory entries.
// Speculative form (kSpec, kW = 4): most candidate lines -- 85 % on a floor plan -- end up completely
// occluded and leave the scene exactly as they found it, so the NEXT candidates' trims by the scene
// do not depend on them.  Each of the four warps therefore takes one of four consecutive candidates
// and trims it by the same scene snapshot (warp-level loops, own candidate list); the results are
// then committed in candidate order: the pieces a candidate split off are appended to the list
// (warp 0 appends directly -- its snapshot is the true scene -- the others stage theirs), and the
// first candidate that survives stops the round: the whole CTA trims the scene by it (:237-240),
// appends it, and the candidates behind it are redone against the new scene.  Same events, same
// order as the sequential loop, hence the same bits (tests/test_render.py).
constexpr int kStageMax = 32;      // pieces a speculative candidate may split off before it is redone in order
struct SpecShared {
  float4 stage[kRenderWarpsPerCta][kStageMax];
  float4 cur[kRenderWarpsPerCta];
  int alive[kRenderWarpsPerCta], npush[kRenderWarpsPerCta], valid[kRenderWarpsPerCta], nlist[kRenderWarpsPerCta];
};

template <int kW, bool kSpec = false>
__device__ void render_fan(const RenderScanParams& p, int f, int slot, GroupShared* gs, float4* sm,
                           float* sm_c, uint2* iv, uint16_t* list, SpecShared* ss = nullptr) {
  static_assert(!kSpec || kW == kRenderWarpsPerCta, "the speculative form is one 4-warp CTA per fan");
  constexpr int kGroup = 32 * kW;                        // threads per fan
  const int gt = (int)threadIdx.x % kGroup;              // thread index inside the group
  float4* lines = p.scratch_lines + (size_t)slot * render::kRenderMaxLines;
  SceneStore scene;
  scene.cap = kSceneSmem * kW;
  scene.sm = sm;
  scene.sm_c = sm_c;
  scene.iv = iv;
  scene.list = list;
  scene.gl = p.scratch_scene + (size_t)slot * render::kRenderMaxLines;
  // per-fan block of scratch_c: [2000 cast constants | 2000 x {lo, span} | 2000 x u16 list]
  float* blk = p.scratch_c + (size_t)slot * kRenderScratchCWords;
  scene.gl_c = blk;
  scene.gl_iv = reinterpret_cast<uint2*>(blk + render::kRenderMaxLines);
  scene.gl_list = reinterpret_cast<uint16_t*>(blk + 3 * render::kRenderMaxLines);
  scene.list_cap = scene.cap;
  const float4 fp = p.fans[f];
  V2 loc;
  loc.x = fp.x; loc.y = fp.y;
  const float amin = fp.z, amax = fp.w;
  // batch mode: one shared object-line set, one laser.  World mode: the object lines of the
  // fan's world minus the robot's own body, the robot's own laser (simulator.cpp:398-411).
  float range_max = p.range_max;
  int num_rays = p.num_rays;
  const float4* obj = p.obj_segs;
  int n_obj = p.n_obj, skip_lo = 0, skip_hi = 0;
  float* out = p.ranges + (size_t)f * (size_t)p.num_rays;
  int wr = 0, ww = 0;
  if (p.world_R > 0) {
    ww = f / p.world_R;
    wr = f % p.world_R;
    range_max = p.robots[wr].laser.range_max;
    num_rays = p.num_rays_r[wr];
    obj = p.world_obj_lines + (size_t)ww * (size_t)p.L_obj;
    n_obj = p.L_obj;
    skip_lo = p.ent_off[p.first_robot_ent + wr];
    skip_hi = p.ent_off[p.first_robot_ent + wr + 1];
    out = p.ranges + (size_t)ww * (size_t)p.rays_per_world + (size_t)p.ray_off[wr];
  }

  // ---- GetSceneLines: map lines, then object lines, in order --------------------------
  int n_lines = 0;
  for (int b = 0; b < p.n_map + n_obj; b += kGroup) {
    // one index space: [map | objects]; a trip may straddle the boundary, order is kept
    const int k = b + gt;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    bool keep = false;
    if (k < p.n_map) {
      s = __ldg(p.segs + k);
      keep = aabb_keep(s, loc.x, loc.y, range_max);
    } else if (k - p.n_map < n_obj) {
      const int o = k - p.n_map;
      if (!(o >= skip_lo && o < skip_hi)) {
        s = obj[o];
        keep = aabb_keep(s, loc.x, loc.y, range_max);
      }
    }
    int total;
    const int pos = group_append_pos<kW>(gs, keep, n_lines, &total);
    if (keep && pos < render::kRenderMaxLines) lines[pos] = s;
    n_lines += total;
  }
  group_sync<kW>();

  // ---- SceneRender, vector_map.cpp:226-246 ---------------------------------------------
  int n_scene = 0;
  if constexpr (kSpec) {
    const int warp = gt >> 5, lane = gt & 31;
    uint16_t* gl_lists = reinterpret_cast<uint16_t*>(blk + 3 * render::kRenderMaxLines);   // 4 x u16[2000]
    SceneStore sw = scene;             // this warp's view: its own candidate list
    sw.list = list + warp * kSceneSmem;
    sw.list_cap = kSceneSmem;
    sw.gl_list = gl_lists + warp * render::kRenderMaxLines;
    int i = 0;
    while (i < n_lines && i < render::kRenderMaxLines) {
      const int lim = n_lines < render::kRenderMaxLines ? n_lines : render::kRenderMaxLines;
      const int n_avail = lim - i < kRenderWarpsPerCta ? lim - i : kRenderWarpsPerCta;
      if (warp < n_avail) {
        Seg cur = load_seg(lines + i + warp);
        const uint2 iv_cur = seg_interval(loc, cur);
        int n_list = 0;
        for (int b = 0; b < n_scene; b += 32) {
          const int j = b + lane;
          const bool keep = j < n_scene && intervals_may_touch(iv_cur, scene.ld_iv(j));
          int total;
          const int pos = group_append_pos<1>(gs, keep, n_list, &total);
          if (keep) sw.st_list(pos, j);
          n_list += total;
        }
        __syncwarp();
        int npush = 0, k0 = 0;
        bool valid = true;
        while (k0 < n_list && render::sq_length(cur) >= render::kRenderEps) {     // :231
          const int k = k0 + lane;
          Seg mine = cur, piece;
          piece.p0.x = piece.p0.y = piece.p1.x = piece.p1.y = 0.f;
          bool pushed = false, acts = false;
          if (k < n_list) {
            const Seg t = scene.ld(sw.ld_list(k));
            if (!(render::sq_length(t) < render::kRenderEps)) {                 // :232
              pushed = render::trim_occlusion(loc, t, &mine, &piece);
              acts = pushed || !same_bits(mine, cur);
            }
          }
          const int first = group_first<1>(gs, acts);
          if (first < 0) { k0 += 32; continue; }
          cur = shfl_seg(mine, first);
          if (__shfl_sync(0xffffffffu, pushed ? 1 : 0, first) != 0) {
            if (warp == 0) {     // the first candidate of a round sees the true scene: append directly (:174)
              if (lane == first && n_lines + npush < render::kRenderMaxLines) store_seg(lines + n_lines + npush, piece);
            } else {
              if (npush >= kStageMax) { valid = false; break; }   // redone in order, as warp 0 of a later round
              if (lane == first) ss->stage[warp][npush] = make_float4(piece.p0.x, piece.p0.y, piece.p1.x, piece.p1.y);
            }
            ++npush;
          }
          k0 += first + 1;
        }
        if (lane == 0) {
          ss->cur[warp] = make_float4(cur.p0.x, cur.p0.y, cur.p1.x, cur.p1.y);
          ss->alive[warp] = render::sq_length(cur) > render::kRenderEps ? 1 : 0;   // :236
          ss->npush[warp] = npush;
          ss->valid[warp] = valid ? 1 : 0;
          ss->nlist[warp] = n_list;
        }
      }
      __syncthreads();
      // ---- commit, in candidate order (block-uniform decisions) ----
      int adv = 0;
      for (int w = 0; w < n_avail; ++w) {
        if (w > 0 && !ss->valid[w]) break;
        const int np = ss->npush[w];
        if (w > 0)
          for (int k = gt; k < np; k += kGroup)
            if (n_lines + k < render::kRenderMaxLines) lines[n_lines + k] = ss->stage[w][k];
        n_lines += np;
        adv = w + 1;
        if (ss->alive[w]) {
          // the survivor trims the scene (:237-240) -- the whole CTA, over the survivor's own list -- and joins it
          SceneStore sb = scene;
          sb.list = list + w * kSceneSmem;
          sb.list_cap = kSceneSmem;
          sb.gl_list = gl_lists + w * render::kRenderMaxLines;
          const Seg cur = load_seg(&ss->cur[w]);
          const int n_it = ss->nlist[w];
          for (int b = 0; b < n_it; b += kGroup) {
            const int k = b + gt;
            Seg piece;
            piece.p0.x = piece.p0.y = piece.p1.x = piece.p1.y = 0.f;
            bool pushed = false;
            if (k < n_it) {
              const int j = sb.ld_list(k);
              Seg t = scene.ld(j);
              if (!(render::sq_length(t) < render::kRenderEps)) {               // :238
                pushed = render::trim_occlusion(loc, cur, &t, &piece);
                scene.st(j, t);
              }
            }
            int total;
            const int pos = group_append_pos<kW>(gs, pushed, n_lines, &total);
            if (pushed && pos < render::kRenderMaxLines) store_seg(lines + pos, piece);
            n_lines += total;
          }
          if (gt == 0) {
            scene.st(n_scene, cur);
            scene.st_iv(n_scene, seg_interval(loc, cur));
          }
          ++n_scene;
          break;
        }
      }
      i += adv;
      __syncthreads();
    }
  } else {
  // the candidate list is read one entry ahead: entries below n_lines never change (pieces are
  // appended at n_lines and beyond), so a prefetched entry is valid iff it existed when it was read
  Seg ahead;
  ahead.p0.x = ahead.p0.y = ahead.p1.x = ahead.p1.y = 0.f;
  bool ahead_ok = false;
  if (n_lines > 0) { ahead = load_seg(lines); ahead_ok = true; }
  for (int i = 0; i < n_lines && i < render::kRenderMaxLines; ++i) {
    Seg cur = ahead_ok ? ahead : load_seg(lines + i);
    ahead_ok = i + 1 < n_lines && i + 1 < render::kRenderMaxLines;
    if (ahead_ok) ahead = load_seg(lines + i + 1);
    // ---- the scene entries this line can interact with, in scene order ----------------------
    // (intervals and list beyond the shared-memory part live in the fan's global scratch)
    const uint2 iv_cur = seg_interval(loc, cur);
    int n_list = 0;
    for (int b = 0; b < n_scene; b += kGroup) {
      const int j = b + gt;
      const bool keep = j < n_scene && intervals_may_touch(iv_cur, scene.ld_iv(j));
      int total;
      const int pos = group_append_pos<kW>(gs, keep, n_list, &total);
      if (keep) scene.st_list(pos, j);
      n_list += total;
    }
    const int n_it = n_list;
    group_sync<kW>();
    int k0 = 0;
    while (k0 < n_it && render::sq_length(cur) >= render::kRenderEps) {     // :231
      const int k = k0 + gt;
      Seg mine = cur, piece;
      piece.p0.x = piece.p0.y = piece.p1.x = piece.p1.y = 0.f;
      bool pushed = false, acts = false;
      if (k < n_it) {
        const int j = scene.ld_list(k);
        const Seg t = scene.ld(j);
        if (!(render::sq_length(t) < render::kRenderEps)) {                 // :232
          pushed = render::trim_occlusion(loc, t, &mine, &piece);
          acts = pushed || !same_bits(mine, cur);
        }
      }
      const int first = group_first<kW>(gs, acts);
      if (first < 0) { k0 += kGroup; continue; }
      // the lowest acting entry is what the reference's loop meets first: adopt its result
      if (kW == 1) {
        cur = shfl_seg(mine, first);
        const bool fp_ = __shfl_sync(0xffffffffu, pushed ? 1 : 0, first) != 0;
        if (fp_) {
          if (gt == first && n_lines < render::kRenderMaxLines) store_seg(lines + n_lines, piece);
          ++n_lines;
        }
      } else {
        if (gt == first) {
          gs->seg = make_float4(mine.p0.x, mine.p0.y, mine.p1.x, mine.p1.y);
          gs->flag = pushed ? 1 : 0;
          if (pushed && n_lines < render::kRenderMaxLines) store_seg(lines + n_lines, piece);
        }
        __syncthreads();
        cur = load_seg(&gs->seg);
        if (gs->flag) ++n_lines;
        __syncthreads();
      }
      k0 += first + 1;
    }
    group_sync<kW>();
    if (render::sq_length(cur) > render::kRenderEps) {                       // :236
      for (int b = 0; b < n_it; b += kGroup) {
        const int k = b + gt;
        Seg piece;
        piece.p0.x = piece.p0.y = piece.p1.x = piece.p1.y = 0.f;
        bool pushed = false;
        if (k < n_it) {
          const int j = scene.ld_list(k);
          Seg t = scene.ld(j);
          if (!(render::sq_length(t) < render::kRenderEps)) {               // :238
            pushed = render::trim_occlusion(loc, cur, &t, &piece);
            scene.st(j, t);
          }
        }
        int total;
        const int pos = group_append_pos<kW>(gs, pushed, n_lines, &total);
        if (pushed && pos < render::kRenderMaxLines) store_seg(lines + pos, piece);
        n_lines += total;
      }
      if (gt == 0) {
        scene.st(n_scene, cur);
        scene.st_iv(n_scene, seg_interval(loc, cur));
      }
      ++n_scene;
    }
    group_sync<kW>();
  }
  }   // !kSpec

  // ---- LineCast list, :439-456 (ordered; built in place over the scene: a trip's entries are
  // all read before any of its casts is written, and cast positions never run ahead of the trip) ----
  const Lattice L = make_lattice(amin, amax, num_rays);
  int n_cast = 0;
  for (int b = 0; b < n_scene; b += kGroup) {
    const int j = b + gt;
    bool keep = false;
    SegCast k;
    k.a0 = k.a1 = k.nx = k.ny = k.c = k.arc = 0.f;
    k.valid = false;
    if (j < n_scene) {
      const Seg t = scene.ld(j);
      if (render::sq_length(t) > render::kRenderEps) {                       // :255-257
        const float p0x = t.p0.x - loc.x, p0y = t.p0.y - loc.y;             // :443-444
        const float p1x = t.p1.x - loc.x, p1y = t.p1.y - loc.y;
        k = finish_cast(atan2_ref(p0y, p0x), atan2_ref(p1y, p1x), p0x, p0y, p1x, p1y);
        keep = k.valid;                                                      // :447
      }
    }
    int total;
    const int pos = group_append_pos<kW>(gs, keep, n_cast, &total);
    if (keep) scene.st_cast(pos, make_float4(k.a0, k.a1, k.nx, k.ny), k.c);
    n_cast += total;
  }
  group_sync<kW>();

  // ---- ray fill, :458-476: the FIRST cast that covers the ray; in world mode followed by
  // publishLaser's dropout + noise (simulator.cpp:412-418) on the step kernel's Philox stream
  // (ray i = 128 s + 32 q + l takes normal q of block 32 s + l, see finalise in umrs_kernels.cu)
  const bool noisy = p.world_R > 0 && p.laser_noise != 0;
  const float stddev = noisy ? p.robots[wr].laser.noise_stddev : 0.f;
  const uint32_t fan_global = (uint32_t)(p.world_base + ww) * (uint32_t)(p.world_R > 0 ? p.world_R : 1) + (uint32_t)wr;
  for (int i = gt; i < num_rays; i += kGroup) {
    const float a = ray_angle(L.amin, L.da, i);
    float r = range_max;                                                     // :433
    for (int c = 0; c < n_cast; ++c) {
      const float4 k = scene.ld_cast(c);
      if (cast_covers(k.x, k.y, a)) {                                        // :469-470
        float rx, ry;
        ray_dir(a, rx, ry);
        r = scene.ld_c(c) / (k.z * rx + k.w * ry);                           // :473
        break;
      }
    }
    if (noisy) {
      const uint32_t blk = (uint32_t)(((i >> 7) << 5) + (i & 31));
      const uint4 w4 = philox4x32(make_uint4(blk, fan_global, p.step_lo, p.step_hi ^ kStreamLaser),
                                  make_uint2(p.key0, p.key1));
      const int q = (i >> 5) & 3;
      const float2 g2 = (q < 2) ? box_muller(w4.x, w4.y) : box_muller(w4.z, w4.w);
      const float g = (q & 1) ? g2.y : g2.x;
      if ((double)r > (double)range_max - 0.1) r = 0.f;
      else r = fmaxf(0.f, r + stddev * g);
    }
    out[i] = r;
  }
}

// ---- small batches: one 4-warp CTA per fan ---------------------------------------------------
__global__ void __launch_bounds__(32 * kRenderWarpsPerCta, 6)
k_scan_reference_render(const RenderScanParams p) {
  __shared__ GroupShared gs;
  __shared__ float4 scene_sm[kRenderWarpsPerCta * kSceneSmem];
  __shared__ float scene_sm_c[kRenderWarpsPerCta * kSceneSmem];
  __shared__ uint2 scene_iv[kRenderWarpsPerCta * kSceneSmem];
  __shared__ uint16_t scene_list[kRenderWarpsPerCta * kSceneSmem];
  __shared__ SpecShared ss;
  const int slot = (int)blockIdx.x;
  if (slot >= p.n_fans) return;
  render_fan<kRenderWarpsPerCta, true>(p, p.fan_begin + slot, slot, &gs, scene_sm, scene_sm_c, scene_iv, scene_list, &ss);
}

// ---- big batches: one 1-warp CTA per fan; 28 CTAs per SM = 4144 resident fans, so that a 4096-fan
// launch is ONE wave and a finished fan frees its slot at once -----------------------------------
__global__ void __launch_bounds__(32, 28)
k_scan_reference_render_warp(const RenderScanParams p) {
  __shared__ GroupShared gs;
  __shared__ float4 scene_sm[kSceneSmem];
  __shared__ float scene_sm_c[kSceneSmem];
  __shared__ uint2 scene_iv[kSceneSmem];
  __shared__ uint16_t scene_list[kSceneSmem];
  const int slot = (int)blockIdx.x;
  if (slot >= p.n_fans) return;
  render_fan<1>(p, p.fan_begin + slot, slot, &gs, scene_sm, scene_sm_c, scene_iv, scene_list);
}

cudaError_t launch_reference_render(const RenderScanParams& p, int fan_begin, int fan_end,
                                    cudaStream_t stream) {
  if (fan_end <= fan_begin) return cudaSuccess;
  RenderScanParams q = p;
  q.fan_begin = fan_begin;            // the launch covers fans [fan_begin, fan_end)
  q.n_fans = fan_end - fan_begin;
  if (q.n_fans <= (g_tuning.render_cta_fans > 0 ? g_tuning.render_cta_fans : kRenderCtaPerFanUpTo)) {
    // few fans: a whole CTA per fan (latency)
    k_scan_reference_render<<<q.n_fans, 32 * kRenderWarpsPerCta, 0, stream>>>(q);
    note_launch();
    return cudaGetLastError();
  }
  // many fans: a warp per fan, one wave.  (Also tried: the heaviest eighth of the fans -- by AABB
  // candidate count -- on 4-warp CTAs in the same launch, light fans four to a CTA behind them:
  // 7.3 ms against 5.7 ms for 4096 fans; the launch ends on a handful of fans whose cost is the
  // serial trim chain of ~700 occluded candidates, which four warps do not shorten.)
  k_scan_reference_render_warp<<<q.n_fans, 32, 0, stream>>>(q);
  note_launch();
  return cudaGetLastError();
}

}  // namespace umrs
