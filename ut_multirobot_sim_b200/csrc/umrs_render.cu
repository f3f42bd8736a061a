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
// Per fan: three scratch arrays in global memory (candidate list, scene, cast constants).
#include "umrs_device.cuh"
#include "umrs_internal.h"
#include "umrs_render.h"

namespace umrs {

using render::Seg;
using render::V2;

constexpr int kRenderWarpsPerCta = 4;
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

// A fan is rendered by a GROUP of kW warps of a 4-warp CTA: kW = 1 (four fans per CTA: best
// throughput for big batches) or kW = 4 (one fan per CTA: a third of the latency for a single
// scan or a small batch).  The helpers below are the group-wide forms of ballot / prefix.
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

template <int kW>
__global__ void __launch_bounds__(32 * kRenderWarpsPerCta)
k_scan_reference_render(const RenderScanParams p) {
  __shared__ GroupShared gs_all[kRenderWarpsPerCta / kW];
  constexpr int kGroup = 32 * kW;                        // threads per fan
  const int gt = (int)threadIdx.x % kGroup;              // thread index inside the group
  const int slot = (int)blockIdx.x * (kRenderWarpsPerCta / kW) + (int)threadIdx.x / kGroup;
  GroupShared* gs = gs_all + (int)threadIdx.x / kGroup;
  // (kW == 4: the whole CTA is one group and takes the __syncthreads below together;
  //  kW == 1: no CTA-wide barrier is ever used, so a warp may leave early)
  if (slot >= p.n_fans) return;
  const int f = p.fan_begin + slot;           // scratch is sized for one launch
  float4* lines = p.scratch_lines + (size_t)slot * render::kRenderMaxLines;
  float4* scene = p.scratch_scene + (size_t)slot * render::kRenderMaxLines;
  float* cast_c = p.scratch_c + (size_t)slot * render::kRenderMaxLines;
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
  for (int i = 0; i < n_lines && i < render::kRenderMaxLines; ++i) {
    Seg cur = load_seg(lines + i);
    int j0 = 0;
    while (j0 < n_scene && render::sq_length(cur) >= render::kRenderEps) {   // :231
      const int j = j0 + gt;
      Seg mine = cur, piece;
      piece.p0.x = piece.p0.y = piece.p1.x = piece.p1.y = 0.f;
      bool pushed = false, acts = false;
      if (j < n_scene) {
        const Seg t = load_seg(scene + j);
        if (!(render::sq_length(t) < render::kRenderEps)) {                 // :232
          pushed = render::trim_occlusion(loc, t, &mine, &piece);
          acts = pushed || !same_bits(mine, cur);
        }
      }
      const int first = group_first<kW>(gs, acts);
      if (first < 0) { j0 += kGroup; continue; }
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
      j0 += first + 1;
    }
    group_sync<kW>();
    if (render::sq_length(cur) > render::kRenderEps) {                       // :236
      for (int b = 0; b < n_scene; b += kGroup) {
        const int j = b + gt;
        Seg piece;
        piece.p0.x = piece.p0.y = piece.p1.x = piece.p1.y = 0.f;
        bool pushed = false;
        if (j < n_scene) {
          Seg t = load_seg(scene + j);
          if (!(render::sq_length(t) < render::kRenderEps)) {               // :238
            pushed = render::trim_occlusion(loc, cur, &t, &piece);
            store_seg(scene + j, t);
          }
        }
        int total;
        const int pos = group_append_pos<kW>(gs, pushed, n_lines, &total);
        if (pushed && pos < render::kRenderMaxLines) store_seg(lines + pos, piece);
        n_lines += total;
      }
      if (gt == 0) store_seg(scene + n_scene, cur);
      ++n_scene;
    }
    group_sync<kW>();
  }

  // ---- LineCast list, :439-456 (ordered; the candidate list is free now and takes it) ----
  const Lattice L = make_lattice(amin, amax, num_rays);
  int n_cast = 0;
  for (int b = 0; b < n_scene; b += kGroup) {
    const int j = b + gt;
    bool keep = false;
    SegCast k;
    k.a0 = k.a1 = k.nx = k.ny = k.c = k.arc = 0.f;
    k.valid = false;
    if (j < n_scene) {
      const Seg t = load_seg(scene + j);
      if (render::sq_length(t) > render::kRenderEps) {                       // :255-257
        const float p0x = t.p0.x - loc.x, p0y = t.p0.y - loc.y;             // :443-444
        const float p1x = t.p1.x - loc.x, p1y = t.p1.y - loc.y;
        k = finish_cast(atan2_ref(p0y, p0x), atan2_ref(p1y, p1x), p0x, p0y, p1x, p1y);
        keep = k.valid;                                                      // :447
      }
    }
    int total;
    const int pos = group_append_pos<kW>(gs, keep, n_cast, &total);
    if (keep) {
      lines[pos] = make_float4(k.a0, k.a1, k.nx, k.ny);
      cast_c[pos] = k.c;
    }
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
      const float4 k = lines[c];
      if (cast_covers(k.x, k.y, a)) {                                        // :469-470
        float rx, ry;
        ray_dir(a, rx, ry);
        r = cast_c[c] / (k.z * rx + k.w * ry);                               // :473
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

cudaError_t launch_reference_render(const RenderScanParams& p, int fan_begin, int fan_end,
                                    cudaStream_t stream) {
  if (fan_end <= fan_begin) return cudaSuccess;
  RenderScanParams q = p;
  q.fan_begin = fan_begin;            // the launch covers fans [fan_begin, fan_end)
  q.n_fans = fan_end - fan_begin;
  // few fans: a whole CTA per fan (latency); many: a warp per fan (throughput)
  if (q.n_fans <= kRenderCtaPerFanUpTo) {
    k_scan_reference_render<kRenderWarpsPerCta><<<q.n_fans, 32 * kRenderWarpsPerCta, 0, stream>>>(q);
  } else {
    const int ctas = (q.n_fans + kRenderWarpsPerCta - 1) / kRenderWarpsPerCta;
    k_scan_reference_render<1><<<ctas, 32 * kRenderWarpsPerCta, 0, stream>>>(q);
  }
  note_launch();
  return cudaGetLastError();
}

}  // namespace umrs
