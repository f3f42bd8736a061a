// Opt-in reference render on the device: VectorMap::GetPredictedScan exactly as the
// reference computes it (vector_map.cpp:420-478) -- GetSceneLines, the O(L^2) SceneRender /
// TrimOcclusion visibility pass, then "first rendered piece that covers the ray" -- so that
// the output is bit-identical to the reference's, occlusion-trimming defects included
// (SURVEY.md 8c tier A).  The default scan (umrs_kernels.cu) is the exact nearest hit instead.
//
// One WARP per fan.  The render is sequential by nature (every candidate is trimmed by the
// scene built so far, pieces split off are appended to the list being iterated), so the warp
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

// Ordered append of the lanes' candidates (GetSceneLines, vector_map.cpp:199-214).
__device__ __forceinline__ void append_ordered(float4* lines, int& n_lines, bool keep, const float4& s,
                                               int lane) {
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  if (keep) {
    const int pos = n_lines + __popc(m & ((1u << lane) - 1u));
    if (pos < render::kRenderMaxLines) lines[pos] = s;
  }
  n_lines += __popc(m);
}

__global__ void __launch_bounds__(32 * kRenderWarpsPerCta)
k_scan_reference_render(const RenderScanParams p) {
  const int lane = threadIdx.x & 31;
  const int slot = (int)blockIdx.x * kRenderWarpsPerCta + (int)(threadIdx.x >> 5);
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
  for (int b = 0; b < p.n_map; b += 32) {
    const int k = b + lane;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    bool keep = false;
    if (k < p.n_map) { s = __ldg(p.segs + k); keep = aabb_keep(s, loc.x, loc.y, range_max); }
    append_ordered(lines, n_lines, keep, s, lane);
  }
  for (int b = 0; b < n_obj; b += 32) {
    const int k = b + lane;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    bool keep = false;
    if (k < n_obj && !(k >= skip_lo && k < skip_hi)) {
      s = obj[k];
      keep = aabb_keep(s, loc.x, loc.y, range_max);
    }
    append_ordered(lines, n_lines, keep, s, lane);
  }
  __syncwarp();

  // ---- SceneRender, vector_map.cpp:226-246 ---------------------------------------------
  int n_scene = 0;
  for (int i = 0; i < n_lines && i < render::kRenderMaxLines; ++i) {
    Seg cur = load_seg(lines + i);
    int j0 = 0;
    while (j0 < n_scene && render::sq_length(cur) >= render::kRenderEps) {   // :231
      const int j = j0 + lane;
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
      const unsigned m = __ballot_sync(0xffffffffu, acts);
      if (m == 0u) { j0 += 32; continue; }
      const int first = __ffs(m) - 1;
      cur = shfl_seg(mine, first);
      const bool first_pushed = __shfl_sync(0xffffffffu, pushed ? 1 : 0, first) != 0;
      if (first_pushed) {
        if (lane == first && n_lines < render::kRenderMaxLines) store_seg(lines + n_lines, piece);
        ++n_lines;
      }
      j0 += first + 1;
    }
    __syncwarp();
    if (render::sq_length(cur) > render::kRenderEps) {                       // :236
      for (int b = 0; b < n_scene; b += 32) {
        const int j = b + lane;
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
        const unsigned m = __ballot_sync(0xffffffffu, pushed);
        if (pushed) {
          const int pos = n_lines + __popc(m & ((1u << lane) - 1u));
          if (pos < render::kRenderMaxLines) store_seg(lines + pos, piece);
        }
        n_lines += __popc(m);
      }
      if (lane == 0) store_seg(scene + n_scene, cur);
      ++n_scene;
    }
    __syncwarp();
  }

  // ---- LineCast list, :439-456 (ordered; the candidate list is free now and takes it) ----
  const Lattice L = make_lattice(amin, amax, num_rays);
  int n_cast = 0;
  for (int b = 0; b < n_scene; b += 32) {
    const int j = b + lane;
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
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const int pos = n_cast + __popc(m & ((1u << lane) - 1u));
      lines[pos] = make_float4(k.a0, k.a1, k.nx, k.ny);
      cast_c[pos] = k.c;
    }
    n_cast += __popc(m);
  }
  __syncwarp();

  // ---- ray fill, :458-476: the FIRST cast that covers the ray; in world mode followed by
  // publishLaser's dropout + noise (simulator.cpp:412-418) on the step kernel's Philox stream
  // (ray i = 128 s + 32 q + l takes normal q of block 32 s + l, see finalise in umrs_kernels.cu)
  const bool noisy = p.world_R > 0 && p.laser_noise != 0;
  const float stddev = noisy ? p.robots[wr].laser.noise_stddev : 0.f;
  const uint32_t fan_global = (uint32_t)(p.world_base + ww) * (uint32_t)(p.world_R > 0 ? p.world_R : 1) + (uint32_t)wr;
  for (int i = lane; i < num_rays; i += 32) {
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
  const int ctas = (q.n_fans + kRenderWarpsPerCta - 1) / kRenderWarpsPerCta;
  k_scan_reference_render<<<ctas, 32 * kRenderWarpsPerCta, 0, stream>>>(q);
  note_launch();
  return cudaGetLastError();
}

}  // namespace umrs
