// Scan kernels (K1 shared-memory scan, K2 uniform-grid walk) and the fused world
// step, sm_100a.  One CTA = one ray fan (one robot's LaserScan of one world), in one of
// seven CTA shapes (threads x resident CTAs per SM) picked per launch: see the launchers.
//
// Replaces
//   VectorMap::GetPredictedScan   vector_map.cpp:420-478   (per-ray fill)
//   VectorMap::SceneRender        vector_map.cpp:217-258   (NOT reproduced: the
//       O(L^2) occlusion trimming is replaced by a per-ray minimum over all
//       candidates -- the same visible surface; SURVEY.md 8c "tier B")
//   VectorMap::GetSceneLines      vector_map.cpp:192-215   (AABB cull, order)
// and, in the world-step form, Simulator::update + publishLaser
//   (simulator.cpp:474-509, 393-426) with the model Step()s.
//
// Phases of a CTA (scan_fan):
//   A  (world step only) integrate every entity of the world, build the
//      world's object lines in shared memory -- or, for large batches, read the
//      fan and the object lines k_world_entities left in global memory (the
//      two-launch tick: the serial entity step leaves every fan's critical path)
//   then, per PASS over at most `cap` source segments (one pass unless the
//   fan's source set is larger than the shared-memory slots):
//   S1 gather: threads stream the fan's source rows (the whole map for K1; the
//      grid rows overlapping [loc +- range_max] for K2; then the object lines),
//      apply the reference's AABB cull and compact the survivors' translated
//      end points + hit index into shared memory (warp ballot + one atomic)
//   S1b pre-cull (fans with more than 256 candidates): coarse angles and
//      distances drop every candidate that lies wholly behind a wall covering
//      its chunks, before any exact work is spent on it
//   S2 records: one thread per survivor: lazily exact end-point angles, swap,
//      unit normal -> {a0,a1,nx,ny | c}; conservative ray-index span; the span
//      is turned into per-32-ray-chunk BUCKETS (count, scan, fill) so that the
//      ray loop only ever touches candidates that can cover its rays
//   C  ray loop: a warp owns 32 consecutive rays (lane = ray), chunks handed out
//      dynamically; every bucket entry is broadcast from shared memory and
//      tested by the 32 rays with the reference's inclusive angular predicate;
//      the running winner is selected on a 2-ulp quotient, the winner's range
//      is then evaluated with the reference's exact expression
//   F  finalise: publishLaser post-process (dropout + Philox noise), coalesced
//      stores of ranges / hit indices
#include <stdlib.h>

#include <mutex>

#include "umrs_device.cuh"
#include "umrs_internal.h"

#ifndef UMRS_GATHER_DEPTH
#define UMRS_GATHER_DEPTH 2
#endif


// Dev build (-DUMRS_PHASE_TIMING): thread 0 of every fan CTA adds the clock64 cycles between
// phase boundaries to counters[8 + k] (k: 0 world phase A, 1 setup, 2 gather, 3 records,
// 4 buckets, 5 ray loop, 6 finalise) -- where a CTA's critical path goes.  Off: nothing.
#ifdef UMRS_PHASE_TIMING
#define UMRS_STAMP(ctrs, k)                                                        \
  do {                                                                             \
    if (threadIdx.x == 0 && (ctrs) != nullptr) {                                   \
      const long long n_ = clock64();                                              \
      atomicAdd((ctrs) + 8 + (k), (unsigned long long)(n_ - umrs_tlast));          \
      umrs_tlast = n_;                                                             \
    }                                                                              \
  } while (0)
#define UMRS_STAMP_INIT long long umrs_tlast = clock64()
#else
#define UMRS_STAMP(ctrs, k) do { } while (0)
#define UMRS_STAMP_INIT do { } while (0)
#endif

namespace umrs {

// Bulk-copy (TMA) staging of the source rows: variant builds only, see stage_* below.
#ifdef UMRS_TMA_STAGE
constexpr bool kTmaBuilt = true;
#else
constexpr bool kTmaBuilt = false;
#endif

// ---- shared-memory carve-up --------------------------------------------------
// Checked build (-DUMRS_CHECKED; compute-sanitizer is closed on this pool): every array of the
// carve-up carries its allocated length and every index / pointer offset is asserted against it
// (a device assert prints file:line and fails the launch), as are the raw shared-address loads of
// the ray loop.  The whole GPU test suite runs against that build (scripts/checked_build.sh).
// Shipped build: plain pointers, nothing.
#ifdef UMRS_CHECKED
#include <assert.h>
#define UMRS_ASSERT(c) assert(c)
template <typename T>
struct Arr {
  T* p;
  int n;
  __device__ __forceinline__ T& operator[](int i) const { assert(i >= 0 && i < n); return p[i]; }
  __device__ __forceinline__ T* operator+(int i) const { assert(i >= 0 && i <= n); return p + i; }
  __device__ __forceinline__ operator T*() const { return p; }
};
template <typename T>
__device__ __forceinline__ Arr<T> make_arr(unsigned char* base, const uint32_t* off, int k) {
  Arr<T> a;
  a.p = reinterpret_cast<T*>(base + off[k]);
  a.n = (int)((off[k + 1] - off[k]) / sizeof(T));   // what smem_layout allocated (16-byte granules)
  return a;
}
#else
#define UMRS_ASSERT(c) do { } while (0)
template <typename T>
using Arr = T*;
template <typename T>
__device__ __forceinline__ Arr<T> make_arr(unsigned char* base, const uint32_t* off, int k) {
  return reinterpret_cast<T*>(base + off[k]);
}
#endif

struct SmemView {
  Arr<float4> rec;       // [2*cap] slot s = rec[2s], rec[2s+1]:
                         //   S1: {p0x,p0y,p1x,p1y} translated | {-, hit index, -, -}
                         //   S2: {a0, a1, nx, ny}             | {c = n . p0', hit index, -, -}
  Arr<uint16_t> pool;    // [3*cap] bucket entries (slots): chunk c owns [c*K, (c+1)*K)
  Arr<uint16_t> wide;    // [cap/2] slots of candidates that met a full bucket: span-checked per chunk
  Arr<uint16_t> surv;    // [cap] slots that survived the coarse occlusion pre-cull (S1b)
  Arr<float> best_r;     // [max_rays] range of the running winner (the reference's exact expression)
  Arr<float> best_q;     // [max_rays] the quotient the winner was selected on (multi-pass plans only)
  Arr<int32_t> best_i;   // [max_rays] its hit index (-1: none)
  Arr<int32_t> cnt;      // [n_chunks_max] (wrapping entries << 16) | plain entries of the bucket
  Arr<uint32_t> zmax2;   // [n_chunks_max] float bits: squared distance beyond which nothing in the chunk can win
  Arr<int32_t> row_v0;   // [max_rows + 1] first virtual index of source row r
  Arr<int32_t> row_s0;   // [max_rows] first source index of row r (object row: 0)
  Arr<int32_t> ctl;      // [16] n_cand, wide overflow, n_wide, next_chunk, n_rows, wedge flag, n_surv, -, wedge[4]
  float4* stage;         // tma_stage: [cap] raw source segments, over pad + pool + wide + best_r + best_i
  Arr<float4> obj;       // [n_obj_lines] world object lines (world step only)
  Arr<float> ent;        // [n_ent][8] {x, y, theta, sin, cos (AngleMod theta), sin, cos (theta), -} (world step only)
};
enum { kCtlNCand = 0, kCtlWideOverflow = 1, kCtlNWide = 2, kCtlNextChunk = 3, kCtlNRows = 4,
       kCtlWedgeOn = 5, kCtlNSurv = 6, kCtlWedge = 8 /* 4 floats: e1x, e1y, e2x, e2y */,
       kCtlBar = 12 /* 8 bytes: the staging mbarrier (tma_stage) */ };

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

// Bucket pool: 3 entries per slot, but never less than 48 per chunk of the widest fan.
__host__ __device__ inline size_t pool_entries(int cap, int n_chunks_max) {
  const size_t a = 3 * (size_t)cap, b = 48 * (size_t)n_chunks_max;
  return a > b ? a : b;
}

struct SmemOffsets { size_t o[17]; };

// Staging area of the bulk-copy variant (tma_stage): the raw source rows, 16 B per segment, land
// in the arrays that are dead until the bucket fill / ray loop -- pool, wide, best_r, best_i, which
// are contiguous -- preceded by whatever padding makes the area hold `cap` segments.
__host__ __device__ inline size_t stage_pad(int cap, int max_rays, int n_chunks, int tma_stage) {
  if (!tma_stage) return 0;
  const size_t have = align16(sizeof(uint16_t) * pool_entries(cap, n_chunks)) + align16(sizeof(uint16_t) * (size_t)(cap / 2)) +
                      align16(sizeof(float) * (size_t)max_rays) + align16(sizeof(int32_t) * (size_t)max_rays);
  const size_t need = sizeof(float4) * (size_t)cap;
  return need > have ? need - have : 0;
}

__host__ __device__ inline SmemOffsets smem_layout(int cap, int max_rows, int max_rays,
                                                   int n_obj_lines, int n_ent, int multipass, int tma_stage) {
  const size_t n_chunks = (size_t)((max_rays + 31) >> 5);
  const size_t sizes[16] = {
      sizeof(float4) * 2 * (size_t)cap,        // rec
      sizeof(uint16_t) * (size_t)cap,          // surv
      0,                                       // (spare)
      stage_pad(cap, max_rays, (int)n_chunks, tma_stage),   // stage = this pad + pool + wide + best_r + best_i
      sizeof(uint16_t) * pool_entries(cap, (int)n_chunks),  // pool
      sizeof(uint16_t) * (size_t)(cap / 2),    // wide
      sizeof(float) * (size_t)max_rays,        // best_r
      sizeof(int32_t) * (size_t)max_rays,      // best_i
      sizeof(int32_t) * n_chunks,              // cnt
      sizeof(uint32_t) * n_chunks,             // zmax2
      sizeof(int32_t) * (size_t)(max_rows + 1),  // row_v0
      sizeof(int32_t) * (size_t)(max_rows + 1),  // row_s0
      sizeof(int32_t) * 16,                    // ctl
      sizeof(float4) * (size_t)n_obj_lines,    // obj
      sizeof(float) * 8 * (size_t)n_ent,       // ent
      multipass ? sizeof(float) * (size_t)max_rays : 0};  // best_q
  SmemOffsets L;
  size_t o = 0;
  for (int k = 0; k < 16; ++k) {
    L.o[k] = o;
    o += align16(sizes[k]);
  }
  L.o[16] = o;
  return L;
}

bool tma_stage_built() { return kTmaBuilt; }

size_t scan_smem_bytes(int cap, int max_rows, int max_rays, int n_obj_lines, int n_ent,
                       int multipass, int tma_stage) {
  return smem_layout(cap, max_rows, max_rays, n_obj_lines, n_ent, multipass, tma_stage).o[16];
}

// The offsets come with the launch parameters (constant bank): see fill_smem_offsets.
__device__ __forceinline__ SmemView carve(unsigned char* base, const ScanCommon& c) {
  SmemView v;
  v.rec = make_arr<float4>(base, c.smem_off, 0);
  v.surv = make_arr<uint16_t>(base, c.smem_off, 1);
  v.pool = make_arr<uint16_t>(base, c.smem_off, 4);
  v.wide = make_arr<uint16_t>(base, c.smem_off, 5);
  v.best_r = make_arr<float>(base, c.smem_off, 6);
  v.best_i = make_arr<int32_t>(base, c.smem_off, 7);
  v.cnt = make_arr<int32_t>(base, c.smem_off, 8);
  v.zmax2 = make_arr<uint32_t>(base, c.smem_off, 9);
  v.row_v0 = make_arr<int32_t>(base, c.smem_off, 10);
  v.row_s0 = make_arr<int32_t>(base, c.smem_off, 11);
  v.ctl = make_arr<int32_t>(base, c.smem_off, 12);
  v.obj = make_arr<float4>(base, c.smem_off, 13);
  v.ent = make_arr<float>(base, c.smem_off, 14);
  v.best_q = make_arr<float>(base, c.smem_off, 15);
  v.stage = reinterpret_cast<float4*>(base + c.smem_off[3]);
  return v;
}

// Host side of carve(): returns the CTA's dynamic shared-memory size.
static size_t fill_smem_offsets(ScanCommon& c, int n_obj_lines, int n_ent) {
  const SmemOffsets L = smem_layout(c.cap, c.max_rows, c.max_rays, n_obj_lines, n_ent, c.multipass, c.tma_stage);
  for (int k = 0; k <= 16; ++k) c.smem_off[k] = (uint32_t)L.o[k];
  return L.o[16];
}

struct Fan {
  float lx, ly;        // laser origin
  float amin, amax;    // world-frame fan limits (angle_min + theta, ...)
  int num_rays;
  float range_max;
};

// publishLaser post-process (simulator.cpp:412-418) with Philox noise.
struct PostProcess {
  bool enabled;
  float stddev;
  uint32_t key0, key1, step_lo, step_hi, fan_global;
};

// ---- bulk-copy staging (TMA engine: cp.async.bulk + mbarrier) ---------------------------
// VARIANT BUILD ONLY (-DUMRS_TMA_STAGE): measured on B200 and rejected for every plan
// (profiles/r2_tma_ab.txt: -1 % on K1 maps, -25 % on the grid walk; the gather is not
// latency-bound, and bulk copies bypass the L1 the co-resident fans share).  In the shipped
// build kTmaBuilt is false and every use below folds away.

// One thread arms the mbarrier with the byte count and issues the copies; the copy engine
// streams global -> shared memory on its own and completes the barrier's transaction count;
// every thread then waits on the barrier's phase.  SASS: UBLKCP.S.G, SYNCS.ARRIVE.TRANS64,
// SYNCS.PHASECHK.TRANS64.TRYWAIT.
__device__ __forceinline__ void stage_bar_init(int32_t* bar_words) {
  const unsigned bar = (unsigned)__cvta_generic_to_shared(bar_words);
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// Arms the barrier for `bytes` in total (one arrival: the issuing thread's).
__device__ __forceinline__ void stage_expect(int32_t* bar_words, unsigned bytes) {
  const unsigned bar = (unsigned)__cvta_generic_to_shared(bar_words);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// bytes: a multiple of 16; src and dst 16-byte aligned.
__device__ __forceinline__ void stage_copy(void* dst_smem, const void* src_gmem, unsigned bytes, int32_t* bar_words) {
  const unsigned bar = (unsigned)__cvta_generic_to_shared(bar_words);
  const unsigned dst = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src_gmem), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void stage_wait(int32_t* bar_words, unsigned parity) {
  const unsigned bar = (unsigned)__cvta_generic_to_shared(bar_words);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// The staging area was last written through the generic proxy (buckets / winners of the previous
// fan): order those writes before the async proxy's.
__device__ __forceinline__ void stage_fence_generic_to_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// K1: the whole map in ONE bulk copy; issued by one thread as early as possible (it does not
// depend on the pose, so it runs under the entity steps).
__device__ __forceinline__ void stage_issue_map(const SmemView& sm, const MapSource& src) {
  const unsigned bytes = (unsigned)src.n * 16u;
  stage_expect(sm.ctl + kCtlBar, bytes);
  if (bytes) stage_copy(sm.stage, src.segs, bytes, sm.ctl + kCtlBar);
}

// ---- source rows of a fan -----------------------------------------------------------
// Virtual index space = concatenation of the rows; row r covers virtual
// [row_v0[r], row_v0[r+1]) = source [row_s0[r], ...).  The LAST row is always the
// object-line row (possibly empty).  Written by thread 0.
__device__ void build_rows(const SmemView& sm, const MapSource& src, const Fan& fan, int n_obj,
                           int max_rows, uint32_t* flags) {
  int r = 0, v = 0;
  if (src.gx == 0) {
    sm.row_v0[r] = v; sm.row_s0[r] = 0; v += src.n; ++r;
  } else {
    // cells whose binned segments (by midpoint) can pass the AABB cull: widen the
    // query box by the largest midpoint-to-endpoint offset, plus a margin that
    // dominates every float rounding in this computation and in the host binning
    const float mx = grid_margin(src.hmax_x, src.cell, fabsf(fan.lx), fan.range_max);
    const float my = grid_margin(src.hmax_y, src.cell, fabsf(fan.ly), fan.range_max);
    const float fx0 = ((fan.lx - fan.range_max - mx) - src.x0) * src.inv_cell;
    const float fx1 = ((fan.lx + fan.range_max + mx) - src.x0) * src.inv_cell;
    const float fy0 = ((fan.ly - fan.range_max - my) - src.y0) * src.inv_cell;
    const float fy1 = ((fan.ly + fan.range_max + my) - src.y0) * src.inv_cell;
    // NaN-safe clamps (a NaN pose yields an empty set)
    int cx0 = fx0 > 0.f ? (fx0 < (float)src.gx ? (int)fx0 : src.gx) : 0;
    int cx1 = fx1 >= 0.f ? (fx1 < (float)src.gx ? (int)fx1 : src.gx - 1) : -1;
    int cy0 = fy0 > 0.f ? (fy0 < (float)src.gy ? (int)fy0 : src.gy) : 0;
    int cy1 = fy1 >= 0.f ? (fy1 < (float)src.gy ? (int)fy1 : src.gy - 1) : -1;
    if (!(fx0 == fx0) || !(fx1 == fx1) || !(fy0 == fy0) || !(fy1 == fy1)) cx1 = -1;
    if (cx0 <= cx1) {
      for (int cy = cy0; cy <= cy1; ++cy) {
        const int s0 = __ldg(src.cell_start + cy * src.gx + cx0);
        const int s1 = __ldg(src.cell_start + cy * src.gx + cx1 + 1);
        if (s1 <= s0) continue;
        if (r >= max_rows - 2) {   // cannot happen while the host bound (capi_scan_plan) holds; never silent
          if (flags != nullptr) atomicOr(flags, kFlagRowsTruncated);
          break;
        }
        sm.row_v0[r] = v; sm.row_s0[r] = s0; v += s1 - s0; ++r;
      }
    }
    // always-visited tail: segments longer than two cells
    if (src.n > src.n_binned) {
      sm.row_v0[r] = v; sm.row_s0[r] = src.n_binned; v += src.n - src.n_binned; ++r;
    }
  }
  sm.row_v0[r] = v; sm.row_s0[r] = 0; v += n_obj; ++r;   // object row
  sm.row_v0[r] = v;
  sm.ctl[kCtlNRows] = r;
}

// sinf / cosf of the reference (float overloads), kept out of line: each inlined
// copy drags a Payne-Hanek slow path along, and the step kernel's instruction
// footprint is what its no_instruction stalls pay for.
__device__ __noinline__ float2 sincos_ref(float a) { return make_float2(sinf(a), cosf(a)); }

constexpr float kWedgeMargin = 1e-3f;   // rad; dominates every rounding in the wedge test by 1000x

constexpr int kGatherDepth = UMRS_GATHER_DEPTH;
// ---- S1: gather + AABB cull + compaction ------------------------------------------
// Virtual indices [v_begin, v_end) of this pass.  Object lines [skip_lo, skip_hi)
// are the fan's own body.
__device__ void gather(const SmemView& sm, const MapSource& src, const Fan& fan,
                       const float4* obj_segs, int skip_lo, int skip_hi, int v_begin, int v_end,
                       bool staged) {
  const int n_rows = sm.ctl[kCtlNRows];
  const int lane = threadIdx.x & 31;
  // Back-wedge cull (fans wider than half a turn): the directions no ray of the fan
  // takes form a convex wedge; a segment with both (translated) end points inside
  // the wedge, contracted by 1e-3 rad at both edges, lies inside it entirely, so
  // no ray angle can satisfy the reference's a0 <= a <= a1 (:469-470) for it.
  const bool wedge_on = sm.ctl[kCtlWedgeOn] != 0;
  const float* wf = reinterpret_cast<const float*>(sm.ctl + kCtlWedge);
  const float e1x = wf[0], e1y = wf[1], e2x = wf[2], e2y = wf[3];
  // Warp-uniform trip counts (the ballots are full-warp), kGatherDepth trips at a time: their
  // loads are issued back to back, so a fan pays one L2 round trip per kGatherDepth trips (the
  // fan's critical path is what the launch's time follows -- DESIGN 4.2).
  const int nt = (int)blockDim.x;
  for (int vb0 = v_begin + (int)(threadIdx.x & ~31u); vb0 < v_end; vb0 += kGatherDepth * nt) {
    float4 s[kGatherDepth];
    int hit[kGatherDepth];
    bool keep[kGatherDepth];
#pragma unroll
    for (int u = 0; u < kGatherDepth; ++u) {
      const int v = vb0 + u * nt + lane;
      keep[u] = false;
      s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      hit[u] = -1;
      if (v < v_end) {
        // row of v: the table is short (2 rows for K1, a few dozen for K2)
        int lo = 0, hi = n_rows - 1;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (sm.row_v0[mid] <= v) lo = mid; else hi = mid - 1;
        }
        const int k = v - sm.row_v0[lo];
        if (lo == n_rows - 1) {
          s[u] = obj_segs[k];
          hit[u] = src.n_map_total + k;
          keep[u] = !(k >= skip_lo && k < skip_hi);
        } else {
          const int si = sm.row_s0[lo] + k;
          s[u] = staged ? sm.stage[v] : __ldg(src.segs + si);   // K1: v == si; K2: rows are staged back to back
          hit[u] = src.orig ? __ldg(src.orig + si) : si;
          keep[u] = true;
        }
      }
    }
    // cull every trip of the group, then ONE slot reservation (a returning shared atomic) for all of them
    unsigned m[kGatherDepth];
    int total = 0;
#pragma unroll
    for (int u = 0; u < kGatherDepth; ++u) {
      bool kp = keep[u] && aabb_keep(s[u], fan.lx, fan.ly, fan.range_max);
      // translate by -loc (vector_map.cpp:447-448)
      s[u] = make_float4(s[u].x - fan.lx, s[u].y - fan.ly, s[u].z - fan.lx, s[u].w - fan.ly);
      if (wedge_on) {
        // cross(e1, p) > 0 && cross(p, e2) > 0 for p = a, b (NaN compares false: kept)
        const bool in_a = __fmaf_rn(e1x, s[u].y, -(e1y * s[u].x)) > 0.f && __fmaf_rn(s[u].x, e2y, -(s[u].y * e2x)) > 0.f;
        const bool in_b = __fmaf_rn(e1x, s[u].w, -(e1y * s[u].z)) > 0.f && __fmaf_rn(s[u].z, e2y, -(s[u].w * e2x)) > 0.f;
        kp = kp && !(in_a && in_b);
      }
      keep[u] = kp;
      m[u] = __ballot_sync(0xffffffffu, kp);   // (lanes beyond v_end hold keep = false)
      total += __popc(m[u]);
    }
    int base = 0;
    if (lane == 0 && total) base = atomicAdd(&sm.ctl[kCtlNCand], total);
    base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
    for (int u = 0; u < kGatherDepth; ++u) {
      if (keep[u]) {
        const int pos = base + __popc(m[u] & ((1u << lane) - 1u));
        sm.rec[2 * pos] = s[u];
        sm.rec[2 * pos + 1] = make_float4(0.f, __int_as_float(hit[u]), 0.f, 0.f);
      }
      base += __popc(m[u]);
    }
  }
}

// ---- span -> chunk ranges ------------------------------------------------------------
struct ChunkRanges {
  int xa, xb;   // chunks of the span itself        (empty: xa > xb)
  int ya, yb;   // chunks of its image one turn down (empty: ya > yb)
};

__device__ __forceinline__ ChunkRanges chunk_ranges(int lo, int hi, const Lattice& L) {
  ChunkRanges c;
  const int last = L.num_rays - 1;
  int l1 = lo < 0 ? 0 : lo, h1 = hi > last ? last : hi;
  c.xa = l1 >> 5; c.xb = h1 >> 5;
  if (l1 > h1) { c.xa = 1; c.xb = 0; }
  int l2 = lo - L.p_hi, h2 = hi - L.p_lo;
  l2 = l2 < 0 ? 0 : l2; h2 = h2 > last ? last : h2;
  c.ya = l2 >> 5; c.yb = h2 >> 5;
  if (l2 > h2 || !L.cull_ok) { c.ya = 1; c.yb = 0; }
  if (c.xa <= c.xb && c.ya <= c.yb && c.yb >= c.xa) {   // overlapping images: one range
    c.xa = c.ya < c.xa ? c.ya : c.xa;
    c.xb = c.yb > c.xb ? c.yb : c.xb;
    c.ya = 1; c.yb = 0;
  }
  return c;
}

// Bucket of chunk c = pool[c*K, (c+1)*K): plain intervals (a0 < a1) fill it from
// the front, wrapping ones (a0 > a1) from the back, so that the ray loop tests
// each kind with a two-compare predicate.  cnt[c] = (n_wrap << 16) | n_plain.
// A candidate that meets a full bucket goes to the `wide` list instead, which
// the ray loop span-checks per chunk (it may then be tested twice in chunks that
// had room -- harmless for a minimum).
__device__ __forceinline__ int bucket_capacity(int cap, int n_chunks_max, int n_chunks) {
  const int k = (int)((unsigned)pool_entries(cap, n_chunks_max) / (unsigned)n_chunks) & ~7;
  return k > 248 ? 248 : k;   // >= 48 by construction of the pool
}

// Conservative ray-index interval of a record (a0 -> a1 counter-clockwise); the
// exact decision is the float predicate in the ray loop.
__device__ __forceinline__ int2 span_of(float a0, float a1, const Lattice& L) {
  if (!L.cull_ok) return make_int2(0, L.num_rays - 1);
  const float arc = (a0 > a1) ? (a1 + k2PiF) - a0 : a1 - a0;
  float u0 = a0 - L.amin;
  u0 -= k2PiF * floorf(u0 / k2PiF);              // [0, 2pi)
  return make_int2((int)floorf(u0 * L.inv_da) - 2, (int)ceilf((u0 + arc) * L.inv_da) + 2);
}

// ---- S2: records + buckets ---------------------------------------------------------------
// Two sub-phases with a barrier between them:
//   S2a make_records: per survivor, the record {a0,a1,nx,ny | c, idx, chunk ranges,
//       mind2}; a well-conditioned candidate lowers zmax2[c] of every chunk whose 32
//       rays it certainly covers to its farthest end point (squared);
//   S2b fill_buckets: the candidate enters the bucket of chunk c unless its nearest
//       point is farther than zmax2[c] -- it then loses on every ray of the chunk.
// A slot that is not a candidate for this fan (dropped by :451, or outside the
// fan) gets a0 = a1 = NaN: no ray angle compares true against it.
//
// Occlusion cull, exactness.  For a ray covered by candidate Y (:469-470) the
// range c/(n.r) is the distance to a point of segment Y, so <= maxd(Y); for a ray
// covered by X it is >= mind(X).  Both hold for the FLOAT quotients to 3e-5
// relative when |n.r| >= 0.01, which |c| > 0.01 maxd guarantees ("well
// conditioned": |n.r| = |c| / range >= |c| / maxd); others neither occlude nor
// get culled.  X is dropped from chunk c only if mind2(X) > 1.001 zmax2[c], i.e.
// mind(X) > 1.0005 maxd(Y) for a Y covering all rays of the chunk: X's quotient
// is strictly larger than Y's on every ray, so neither the minimum nor a tie can
// involve X.
constexpr float kOcclWell = 1e-4f;     // c^2 > kOcclWell * maxd^2
constexpr float kOcclMargin = 1.001f;  // on squared distances

// chunk ranges -> 30 bits: xa | xb << 10 | (yb + 1) << 20 (ya == 0 whenever the
// image range exists: lo < period; 0 = no image); n_chunks <= 512.
__device__ __forceinline__ int pack_ranges(const ChunkRanges& c) {
  const int y = (c.ya <= c.yb) ? c.yb + 1 : 0;
  return c.xa | (c.xb << 10) | (y << 20);
}

// chunks whose rays all lie in the index interval [li, hi] (certain coverage)
__device__ __forceinline__ void lower_zmax(const SmemView& sm, int li, int hi, int last, int n_chunks,
                                           uint32_t bits) {
  const int ca = li > 0 ? (li + 31) >> 5 : 0;
  const int cb = hi >= last ? n_chunks - 1 : ((hi + 1) >> 5) - 1;
  for (int c = ca; c <= cb; ++c) atomicMin(&sm.zmax2[c], bits);
}

// ---- S1b: coarse occlusion pre-cull ------------------------------------------------------
// Most gathered candidates are walls behind nearer walls.  Before the expensive S2
// work (exact end-point angles, normals, buckets) a cheap pass drops those that
// cannot win on ANY ray, with the same argument as the per-chunk cull of S2 (below)
// but on coarse quantities only:
//   * approximate end-point angles (polynomial atan, |error| < 2e-5 rad; every use
//     carries kPreSlack = 1e-3 rad of margin on top of the S2 index margins);
//   * mind2 = squared distance origin -> segment (a lower bound of every hit range
//     on the segment), maxd2 = farthest end point (an upper bound);
//   * only WELL-CONDITIONED segments take part, as occluder or as occluded:
//     c^2 > 1.2e-4 maxd2 on the coarse c implies the c^2 > 1e-4 maxd2 of the S2
//     argument for the exact c (the coarse c^2 is within 1e-4 relative: the cross
//     product is taken against the short edge vector, so it does not cancel).  For
//     these the reference's a0 -> a1 order is the geometric CCW order (their arc is
//     in [1e-4, pi - 0.02]; an arc the coarse angles cannot order comes out near 2 pi
//     and disqualifies the segment), so the exact coverage interval lies inside the
//     widened coarse span, and contains the narrowed one.
// Pass 1 lowers zmax2[c] for every chunk an occluder certainly covers; after a barrier
// pass 2 keeps X iff it is not well-conditioned or some chunk its widened span touches
// has mind2(X) <= 1.002 zmax2[c].  A dropped X loses, by > 0.05 %, on every ray it
// covers, to a candidate that stays (an occluder Y of chunk c has maxd2(Y) >= zmax2[c],
// so Y itself is kept in c), hence neither the minimum nor a tie can involve it.
// zmax2 is kept for S2, whose own lowering can only tighten it.  Survivors are
// compacted into sm.surv; a dropped slot gets a0 = a1 = NaN (for the sweep-all path).
constexpr float kPreWell = 1.2e-4f;
constexpr float kPreMargin = 1.002f;
constexpr float kPreSlack = 1e-3f;
#ifndef UMRS_PRECULL_MIN   // dev builds: 1 << 30 switches S1b off
#define UMRS_PRECULL_MIN 256
#endif
constexpr int kPreMinCand = UMRS_PRECULL_MIN;   // fewer candidates than the widest CTA has threads: S2 is one trip anyway

__device__ void precull_mark(const SmemView& sm, const Lattice& L) {
  const int n_cand = sm.ctl[kCtlNCand];
  const int last = L.num_rays - 1, n_chunks = (L.num_rays + 31) >> 5;
  const int m = (int)(kPreSlack * L.inv_da) + 2;
  for (int slot = threadIdx.x; slot < n_cand; slot += blockDim.x) {
    const float4 p = sm.rec[2 * slot];
    const float ex = p.z - p.x, ey = p.w - p.y;
    const float d0 = __fmaf_rn(p.x, p.x, p.y * p.y), d1 = __fmaf_rn(p.z, p.z, p.w * p.w);
    const float maxd2 = fmaxf(d0, d1);
    const float len2 = __fmaf_rn(ex, ex, ey * ey);
    const float cr = __fmaf_rn(p.x, ey, -(p.y * ex));       // cross(p0, p1 - p0) = |c| * len
    const float c2 = (cr * cr) * rcp_approx(len2);
    int packed = -1;   // -1: takes no part (always survives)
    float mind2 = 0.f;
    if (c2 > kPreWell * maxd2 && c2 <= maxd2) {   // NaN / inf: false
      float aa = atan2_coarse(p.y, p.x), ab = atan2_coarse(p.w, p.z);
      if (cr < 0.f) { const float t = aa; aa = ab; ab = t; }
      float arc = ab - aa;
      if (arc < 0.f) arc += k2PiF;
      if (arc < 3.2f) {
        float u0 = aa - L.amin;
        u0 -= k2PiF * floorf(u0 * (1.0f / k2PiF));
        const int lo = (int)floorf(u0 * L.inv_da) - m, hi = (int)ceilf((u0 + arc) * L.inv_da) + m;
        const float dot0 = -__fmaf_rn(p.x, ex, p.y * ey);
        mind2 = dot0 <= 0.f ? d0 : (dot0 >= len2 ? d1 : c2);
        packed = pack_ranges(chunk_ranges(lo, hi, L));
        const uint32_t bits = __float_as_uint(maxd2);
        const int li = lo + 2 * m + 2, ui = hi - 2 * m - 2;
#pragma unroll 1
        for (int img = 0; img < 2; ++img)   // the span and its image one turn down
          lower_zmax(sm, li - (img ? L.p_lo : 0), ui - (img ? L.p_hi : 0), last, n_chunks, bits);
      }
    }
    float4* r2 = sm.rec + 2 * slot + 1;
    r2->z = __int_as_float(packed);
    r2->w = mind2;
  }
}

__device__ void precull_compact(const SmemView& sm) {
  const int n_cand = sm.ctl[kCtlNCand];
  const int lane = threadIdx.x & 31;
  const float nanf_ = __int_as_float(0x7fc00000);
  for (int sb = (int)(threadIdx.x & ~31u); sb < n_cand; sb += blockDim.x) {
    const int slot = sb + lane;
    bool keep = false;
    if (slot < n_cand) {
      const float4 r2 = sm.rec[2 * slot + 1];
      const int packed = __float_as_int(r2.z);
      if (packed < 0) {
        keep = true;
      } else {
#pragma unroll 1
        for (int part = 0; part < 2 && !keep; ++part) {
          const int ca = part ? 0 : (packed & 1023);
          const int cb = part ? ((packed >> 20) & 1023) - 1 : ((packed >> 10) & 1023);
          for (int c = ca; c <= cb; ++c)
            if (!(r2.w > kPreMargin * __uint_as_float(sm.zmax2[c]))) { keep = true; break; }
        }
        if (!keep) sm.rec[2 * slot] = make_float4(nanf_, nanf_, 0.f, 0.f);
      }
    }
    const unsigned mk = __ballot_sync(0xffffffffu, keep);
    int base = 0;
    if (lane == 0 && mk) base = atomicAdd(&sm.ctl[kCtlNSurv], __popc(mk));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (keep) sm.surv[base + __popc(mk & ((1u << lane) - 1u))] = (uint16_t)slot;
  }
}

__device__ void make_records(const SmemView& sm, const Lattice& L, bool listed,
                             unsigned long long* counters) {
  const int n_cand = sm.ctl[kCtlNCand];
  const float nanf_ = __int_as_float(0x7fc00000);
  const int last = L.num_rays - 1, n_chunks = (L.num_rays + 31) >> 5;
  int n_valid = 0;
  const int n_list = listed ? sm.ctl[kCtlNSurv] : n_cand;
  for (int j = threadIdx.x; j < n_list; j += blockDim.x) {
    const int slot = listed ? (int)sm.surv[j] : j;
    const float4 p = sm.rec[2 * slot];
    const SegCast k = make_cast(p.x, p.y, p.z, p.w, L);
    float4 rec = make_float4(nanf_, nanf_, 0.f, 0.f);
    int packed = 0;
    float mind2 = 0.f;
    if (k.valid) {
      ++n_valid;
      const int2 sp = span_of(k.a0, k.a1, L);
      const ChunkRanges cr = chunk_ranges(sp.x, sp.y, L);
      if (cr.xa <= cr.xb || cr.ya <= cr.yb) {
        rec = make_float4(k.a0, k.a1, k.nx, k.ny);
        packed = pack_ranges(cr);
        if (L.lazy_ok) {
          const float maxd2 = fmaxf(__fmaf_rn(p.x, p.x, p.y * p.y), __fmaf_rn(p.z, p.z, p.w * p.w));
          const float c2 = k.c * k.c;
          if (c2 > kOcclWell * maxd2) {   // well conditioned (NaN: false)
            // along-line coordinates of the end points; direction d = (ny, -nx)
            const float t0 = __fmaf_rn(p.x, k.ny, -(p.y * k.nx));
            const float t1 = __fmaf_rn(p.z, k.ny, -(p.w * k.nx));
            mind2 = c2 + ((t0 * t1 > 0.f) ? fminf(t0 * t0, t1 * t1) : 0.f);
            const uint32_t bits = __float_as_uint(maxd2);
#pragma unroll 1
            for (int img = 0; img < 2; ++img)   // the span and its image one turn down
              lower_zmax(sm, sp.x + 4 - (img ? L.p_lo : 0), sp.y - 4 - (img ? L.p_hi : 0), last, n_chunks, bits);
          }
        }
      }
    }
    sm.rec[2 * slot] = rec;
    float4* r2 = sm.rec + 2 * slot + 1;
    r2->x = k.c;
    r2->z = __int_as_float(packed);
    r2->w = mind2;
  }
  if (counters != nullptr) {
    for (int o = 16; o > 0; o >>= 1) n_valid += __shfl_down_sync(0xffffffffu, n_valid, o);
    if ((threadIdx.x & 31) == 0 && n_valid) atomicAdd(counters + 1, (unsigned long long)n_valid);
  }
}

__device__ void fill_buckets(const SmemView& sm, int K, int wide_cap, bool listed) {
  const int n_list = listed ? sm.ctl[kCtlNSurv] : sm.ctl[kCtlNCand];
  for (int j = threadIdx.x; j < n_list; j += blockDim.x) {
    const int slot = listed ? (int)sm.surv[j] : j;
    const float4 r1 = sm.rec[2 * slot];
    if (!(r1.x == r1.x)) continue;   // not a candidate
    const float4 r2 = sm.rec[2 * slot + 1];
    const int packed = __float_as_int(r2.z);
    const float mind2 = r2.w;
    const bool wrap = r1.x > r1.y;
    const int inc = wrap ? 0x10000 : 1;
    const uint16_t entry = (uint16_t)slot;
    bool spilled = false;
#pragma unroll 1   // one copy of the bucket insertion: the kernel is instruction-fetch sensitive
    for (int part = 0; part < 2; ++part) {
      const int ca = part ? 0 : (packed & 1023);
      const int cb = part ? ((packed >> 20) & 1023) - 1 : ((packed >> 10) & 1023);
      for (int c = ca; c <= cb; ++c) {
        if (mind2 > kOcclMargin * __uint_as_float(sm.zmax2[c])) continue;   // occluded in this chunk
        const int old = atomicAdd(&sm.cnt[c], inc);
        const int na = old & 0xffff, nb = old >> 16;
        if (na + nb < K) {
          sm.pool[c * K + (wrap ? K - 1 - nb : na)] = entry;
        } else {
          atomicSub(&sm.cnt[c], inc);   // bucket full: stays full from here on
          spilled = true;
        }
      }
    }
    if (spilled) {
      const int w = atomicAdd(&sm.ctl[kCtlNWide], 1);
      if (w < wide_cap) sm.wide[w] = entry;
      else sm.ctl[kCtlWideOverflow] = 1;   // ray loop falls back to sweeping every slot
    }
  }
}

// Shared-memory loads of the hot loop on 32-bit shared addresses (register +
// immediate), so that no address arithmetic lands on the integer pipe.
__device__ __forceinline__ float4 lds_f4(unsigned addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f1_off16(unsigned addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+16];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ int lds_i1_off20(unsigned addr) {
  int v;
  asm volatile("ld.shared.b32 %0, [%1+20];" : "=r"(v) : "r"(addr));
  return v;
}
template <int kOff>
__device__ __forceinline__ unsigned lds_u16(unsigned addr) {
  unsigned v;
  asm volatile("ld.shared.u16 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(kOff));
  return v;
}
// slot -> shared address of its 32-byte record: one IMAD (float/IMAD pipe).
__device__ __forceinline__ unsigned rec_addr(unsigned slot, unsigned rec_s) {
  unsigned a;
  asm("mad.lo.u32 %0, %1, 32, %2;" : "=r"(a) : "r"(slot), "r"(rec_s));
  return a;
}

// The rarely needed parts of the ray loop, out of line (the kernel is instruction-fetch
// sensitive: they stay out of the hot loop's cache lines) and on shared-memory addresses only.
struct RayBest { float q; int slot; int tie; };

// One candidate against this lane's ray, any interval kind.
__device__ __forceinline__ void test_general(unsigned rec_s, int s, float a, float rx, float ry, RayBest& b) {
  const unsigned ra = rec_addr((unsigned)s, rec_s);
  const float4 r4 = lds_f4(ra);
  const float q = lds_f1_off16(ra) * rcp_approx(__fmaf_rn(r4.z, rx, r4.w * ry));
  const bool cov = cast_covers(r4.x, r4.y, a);
  if (cov && q == b.q) b.tie = 1;
  if (cov && q < b.q) { b.q = q; b.slot = s; }
}

// The candidates that met a full bucket (list_s = shared address of the wide list), or every slot
// (list_s = 0), against the chunk's rays, continuing from the running winner b.
__device__ __noinline__ RayBest slow_candidates(unsigned rec_s, unsigned list_s, int n, float a, float rx,
                                                float ry, RayBest b) {
  for (int e = 0; e < n; ++e)
    test_general(rec_s, list_s ? (int)lds_u16<0>(list_s + 2u * (unsigned)e) : e, a, rx, ry, b);
  return b;
}

// Careful sweep after an equal quotient: among the covering candidates with q == best the lowest
// hit index wins (the reference's scan order with a strict <).  Returns the winner's slot.
__device__ __noinline__ int tie_sweep(unsigned rec_s, unsigned bucket_s, int K, int n_plain, int n_wrap,
                                      unsigned list_s, int n_list, float a, float rx, float ry, float best,
                                      int best_s, int prev_i) {
  int best_idx = best_s >= 0 ? lds_i1_off20(rec_addr((unsigned)best_s, rec_s)) : prev_i;
  const int n_b = n_plain + n_wrap;
  for (int e = 0; e < n_b + n_list; ++e) {
    int slot;
    if (e < n_plain) slot = (int)lds_u16<0>(bucket_s + 2u * (unsigned)e);
    else if (e < n_b) slot = (int)lds_u16<0>(bucket_s + 2u * (unsigned)(K - 1 - (e - n_plain)));
    else slot = list_s ? (int)lds_u16<0>(list_s + 2u * (unsigned)(e - n_b)) : e - n_b;
    const unsigned ra = rec_addr((unsigned)slot, rec_s);
    const float4 r4 = lds_f4(ra);
    if (!cast_covers(r4.x, r4.y, a)) continue;
    const float q = lds_f1_off16(ra) * rcp_approx(__fmaf_rn(r4.z, rx, r4.w * ry));
    const int id = lds_i1_off20(ra);
    if (q == best && id < best_idx) { best_s = slot; best_idx = id; }
  }
  return best_s;
}

// ---- C: ray loop ----------------------------------------------------------------------
__device__ void ray_loop(const SmemView& sm, const Fan& fan, const Lattice& L, int K, int wide_cap,
                         int pass, bool multipass, unsigned long long* counters) {
  const int lane = threadIdx.x & 31;
  const int n_chunks = (fan.num_rays + 31) >> 5;
  const int n_cand = sm.ctl[kCtlNCand];
  // not a proper fan, or more spilled candidates than the wide list holds: sweep every slot
  const bool sweep_all = !L.cull_ok || sm.ctl[kCtlWideOverflow] != 0;
  const int n_wide = sweep_all ? 0 : sm.ctl[kCtlNWide];
  unsigned long long n_pairs = 0;
  const float inf = __int_as_float(0x7f800000);
  const unsigned rec_s = (unsigned)__cvta_generic_to_shared(sm.rec);
  const unsigned pool_s = (unsigned)__cvta_generic_to_shared(sm.pool);
  const unsigned wide_s = (unsigned)__cvta_generic_to_shared(sm.wide);

  for (;;) {
    int chunk = 0;
    if (lane == 0) chunk = atomicAdd(&sm.ctl[kCtlNextChunk], 1);
    chunk = __shfl_sync(0xffffffffu, chunk, 0);
    if (chunk >= n_chunks) break;
    const int i = (chunk << 5) + lane;
    const bool live = i < fan.num_rays;
    float a = __int_as_float(0x7fc00000);  // NaN: covers nothing
    float rx = 0.f, ry = 0.f;
    if (live) {
      a = ray_angle(L.amin, L.da, i);                    // vector_map.cpp:467
      ray_dir(a, rx, ry);                                // :472
    }
    const int prev_i = (pass > 0 && live) ? sm.best_i[i] : -1;
    float best = prev_i >= 0 ? sm.best_q[i] : inf;
    int best_s = -1;          // slot of the running winner of this pass
    int tie = 0;   // set when a covering candidate equals the running best (collinear wall pieces
                   // share n and c exactly, so the equality must be qualified by coverage)

    // One candidate (slot S) against the 32 rays of the chunk.
    //  * The winner is selected on q = c / (n . r) evaluated with one FMA and a
    //    1-ulp reciprocal; the reference's exact expression is evaluated once, for
    //    the winner, below.  An equal q only raises `tie` (careful sweep below).
    //  * Coverage (:469-470, inclusive) as the sign of one product, so that it
    //    costs float-pipe adds instead of integer-pipe compares:
    //      plain    a0 <= a && a <= a1   <=>   (a - a0) * (a1 - a) >= 0
    //      wrapping a0 <= a || a <= a1   <=>   (a - a1) * (a0 - a) <= 0
    //    (a float difference is zero only for equal operands and has the exact
    //    sign; the product cannot underflow because |a1 - a0| >= 1e-4 and nonzero
    //    angle differences are >= 1e-9; NaN -- a dead lane -- compares false.)
#define UMRS_TEST(S, COVER_EXPR)                                                          \
    {                                                                                     \
      const int s_ = (int)(S);                                                            \
      UMRS_ASSERT(s_ >= 0 && s_ < n_cand);                                                \
      const unsigned ra_ = rec_addr((unsigned)s_, rec_s);                                 \
      const float4 r4 = lds_f4(ra_);                          /* broadcast LDS.128 */     \
      const float cc = lds_f1_off16(ra_);                                                 \
      const float q = cc * rcp_approx(__fmaf_rn(r4.z, rx, r4.w * ry));                    \
      const bool cov = (COVER_EXPR);                                                      \
      if (cov && q == best) tie = 1;                                                      \
      if (cov && q < best) { best = q; best_s = s_; }                                     \
    }
#define UMRS_COV_PLAIN ((a - r4.x) * (r4.y - a) >= 0.f)
#define UMRS_COV_WRAP ((a - r4.y) * (r4.x - a) <= 0.f)

    int n_plain = 0, n_wrap = 0;
    const unsigned bucket_s = pool_s + 2u * (unsigned)(chunk * K);
    if (!sweep_all) {
      const int packed = sm.cnt[chunk];
      n_plain = packed & 0xffff;
      n_wrap = packed >> 16;
#ifdef UMRS_CHECKED
      assert(n_plain + n_wrap <= K && (chunk + 1) * K <= sm.pool.n);
#endif
      unsigned bp = bucket_s;
      int n = n_plain;
      for (; n >= 4; n -= 4, bp += 8) {               // four candidates in flight
        const unsigned e0 = lds_u16<0>(bp), e1 = lds_u16<2>(bp), e2 = lds_u16<4>(bp), e3 = lds_u16<6>(bp);
        UMRS_TEST(e0, UMRS_COV_PLAIN)
        UMRS_TEST(e1, UMRS_COV_PLAIN)
        UMRS_TEST(e2, UMRS_COV_PLAIN)
        UMRS_TEST(e3, UMRS_COV_PLAIN)
      }
      for (; n > 0; --n, bp += 2) UMRS_TEST(lds_u16<0>(bp), UMRS_COV_PLAIN)
      bp = bucket_s + 2u * (unsigned)(K - 1);
      for (int e = 0; e < n_wrap; ++e, bp -= 2) UMRS_TEST(lds_u16<0>(bp), UMRS_COV_WRAP)
      n_pairs += n_plain + n_wrap;
    }
#undef UMRS_TEST
#undef UMRS_COV_PLAIN
#undef UMRS_COV_WRAP
    // spilled candidates (no span check: the coverage predicate is the exact test anyway), or every slot
    const int n_list = sweep_all ? n_cand : n_wide;
    const unsigned list_s = sweep_all ? 0u : wide_s;
    if (n_list > 0) {
      RayBest b;
      b.q = best; b.slot = best_s; b.tie = tie;
      b = slow_candidates(rec_s, list_s, n_list, a, rx, ry, b);
      best = b.q; best_s = b.slot; tie = b.tie;
      n_pairs += n_list;
    }
    if (__any_sync(0xffffffffu, tie != 0)) {
      if (counters != nullptr && lane == 0) atomicAdd(counters + 4, 1ull);
      best_s = tie_sweep(rec_s, bucket_s, K, n_plain, n_wrap, list_s, n_list, a, rx, ry, best, best_s, prev_i);
    }

    if (live) {
      if (best_s >= 0) {
        const unsigned ra = rec_addr((unsigned)best_s, rec_s);
        const float4 r4 = lds_f4(ra);
        const float den = r4.z * rx + r4.w * ry;         // n.dot(r), vector_map.cpp:473
        sm.best_r[i] = lds_f1_off16(ra) / den;
        if (multipass) sm.best_q[i] = best;
        sm.best_i[i] = lds_i1_off20(ra);
      } else if (pass == 0) {
        sm.best_r[i] = inf;
        sm.best_i[i] = -1;
      }
    }
  }
  if (counters != nullptr && threadIdx.x == 0) {
    if (n_wide) atomicAdd(counters + 5, (unsigned long long)n_wide);
    if (sweep_all) atomicAdd(counters + 6, 1ull);
  }
  if (counters != nullptr && lane == 0 && n_pairs) {
    atomicAdd(counters + 2, n_pairs);
    atomicAdd(counters + 0, n_pairs * 32ull);  // tests = pairs x 32 lanes
  }
}

// ---- F: finalise ------------------------------------------------------------------------
// Noise counter layout: ray i = 128*s + 32*q + l (l < 32, q < 4) takes normal q of
// Philox block {32*s + l, fan_global, step_lo, step_hi ^ 'LASR'} -- one Philox call
// and two Box-Muller transforms serve four rays, and a warp's stores stay coalesced.
__device__ void finalise(const SmemView& sm, const Fan& fan, float* __restrict__ ranges,
                         int32_t* __restrict__ hit_idx, const PostProcess& post) {
  const int n_blocks = ((fan.num_rays + 127) >> 7) << 5;
  for (int b = threadIdx.x; b < n_blocks; b += blockDim.x) {
    float g[4] = {0.f, 0.f, 0.f, 0.f};
    if (post.enabled) {
      const uint4 w = philox4x32(make_uint4((uint32_t)b, post.fan_global, post.step_lo,
                                            post.step_hi ^ kStreamLaser),
                                 make_uint2(post.key0, post.key1));
      const float2 g01 = box_muller(w.x, w.y), g23 = box_muller(w.z, w.w);
      g[0] = g01.x; g[1] = g01.y; g[2] = g23.x; g[3] = g23.y;
    }
    const int i0 = ((b >> 5) << 7) + (b & 31);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i = i0 + (q << 5);
      if (i >= fan.num_rays) break;
      const int bi = sm.best_i[i];
      float out = bi >= 0 ? sm.best_r[i] : fan.range_max;   // :433 fill value
      if (post.enabled) {
        // simulator.cpp:412-418; range_max - 0.1 is a double expression
        if ((double)out > (double)fan.range_max - 0.1) out = 0.f;
        else out = fmaxf(0.f, out + post.stddev * g[q]);
      }
      ranges[i] = out;
      if (hit_idx != nullptr) hit_idx[i] = bi;
    }
  }
}

// ---- one fan, all passes --------------------------------------------------------------
__device__ void scan_fan(const SmemView& sm, const ScanCommon& c, const Fan& fan,
                         const float4* obj_segs, int n_obj, int skip_lo, int skip_hi,
                         float* ranges, int32_t* hit_idx, const PostProcess& post, unsigned stage_parity,
                         unsigned long long* counters) {
  // `counters`: the launch's work-accounting array, or nullptr -- a compile-time nullptr in the
  // instances ordinary launches use (kCounters = false), so that none of the accounting code is in them
  UMRS_STAMP_INIT;
  const Lattice L = make_lattice(fan.amin, fan.amax, fan.num_rays);
  const int n_chunks = (fan.num_rays + 31) >> 5;
  const int K = bucket_capacity(c.cap, (c.max_rays + 31) >> 5, n_chunks);
  if (threadIdx.x == 0) build_rows(sm, c.src, fan, n_obj, c.max_rows, c.flags);
  if (threadIdx.x >= 32 && threadIdx.x < 36) sm.ctl[threadIdx.x - 32] = 0;
  if (threadIdx.x == 36) sm.ctl[kCtlNSurv] = 0;
  if (threadIdx.x == 64) {
    // wedge of directions behind the fan: CCW from e1 (last ray + margin) to e2
    // (first ray - margin); used only when it is narrower than half a turn
    const float last = (float)(fan.num_rays - 1) * L.da;
    const float width = k2PiF - last - 2.f * kWedgeMargin;
    const bool on = L.cull_ok && fabsf(fan.amin) <= 8.0f && width > 1e-2f && width < kPiF - 1e-2f;
    float* wf = reinterpret_cast<float*>(sm.ctl + kCtlWedge);
    float s1 = 0.f, c1 = 0.f, s2 = 0.f, c2 = 0.f;
    if (on) {
      const float2 w1 = sincos_ref(fan.amin + last + kWedgeMargin);
      const float2 w2 = sincos_ref(fan.amin - kWedgeMargin);
      s1 = w1.x; c1 = w1.y; s2 = w2.x; c2 = w2.y;
    }
    wf[0] = c1; wf[1] = s1; wf[2] = c2; wf[3] = s2;
    sm.ctl[kCtlWedgeOn] = on ? 1 : 0;
  }
  for (int k = threadIdx.x; k < n_chunks; k += blockDim.x) {
    sm.cnt[k] = 0;
    sm.zmax2[k] = 0x7f800000u;   // +inf; valid across passes (a bound from any candidate holds)
  }
  __syncthreads();
  UMRS_STAMP(counters, 1);
  const int v_total = sm.row_v0[sm.ctl[kCtlNRows]];
  // Bulk-copy staging: single-pass fans only (the area aliases the per-ray winners, which a
  // multi-pass fan carries from pass to pass).
  const bool staged = kTmaBuilt && c.tma_stage != 0 && v_total <= c.cap;
  if (kTmaBuilt && c.tma_stage == 2) {
    // K2: the rows are known now -- one bulk copy per grid row, back to back in virtual order
    if (threadIdx.x == 0) {
      const int n_rows = sm.ctl[kCtlNRows];
      const int n_src = staged ? sm.row_v0[n_rows - 1] : 0;   // all rows but the object row
      stage_expect(sm.ctl + kCtlBar, (unsigned)n_src * 16u);
      for (int r = 0; staged && r < n_rows - 1; ++r) {
        const int len = sm.row_v0[r + 1] - sm.row_v0[r];
        if (len > 0) stage_copy(sm.stage + sm.row_v0[r], c.src.segs + sm.row_s0[r], (unsigned)len * 16u, sm.ctl + kCtlBar);
      }
    }
  }
  if (kTmaBuilt && c.tma_stage != 0) stage_wait(sm.ctl + kCtlBar, stage_parity);   // also when not used: keeps the phases in step
  int pass = 0;
  for (int v0 = 0; v0 < v_total || pass == 0; v0 += c.cap, ++pass) {
    const int v1 = v0 + c.cap < v_total ? v0 + c.cap : v_total;
    gather(sm, c.src, fan, obj_segs, skip_lo, skip_hi, v0, v1, staged);
    __syncthreads();
    UMRS_STAMP(counters, 2);
    if (pass == 0 || sm.ctl[kCtlNCand] > 0) {   // block-uniform: a later pass may be empty
      // S1b pays for itself once S2 would need more than one trip of the CTA
      // (block-uniform: n_cand is settled behind the barrier above)
      const bool pre = L.lazy_ok && sm.ctl[kCtlNCand] > kPreMinCand;
      if (pre) {
        precull_mark(sm, L);
        __syncthreads();
        precull_compact(sm);
        __syncthreads();
      }
#ifdef UMRS_HIST   // dev build: histogram of the candidates S2 prepares per fan -> counters[8..15]
      if (threadIdx.x == 0 && counters != nullptr) {
        const int ns = pre ? sm.ctl[kCtlNSurv] : sm.ctl[kCtlNCand];
        const int edges[7] = {192, 256, 320, 384, 448, 512, 608};
        int b = 0;
        while (b < 7 && ns > edges[b]) ++b;
        atomicAdd(counters + 8 + b, 1ull);
      }
#endif
      make_records(sm, L, pre, counters);
      __syncthreads();
      UMRS_STAMP(counters, 3);
      fill_buckets(sm, K, c.cap / 2, pre);
      __syncthreads();
      UMRS_STAMP(counters, 4);
      ray_loop(sm, fan, L, K, c.cap / 2, pass, c.multipass != 0, counters);
    }
    __syncthreads();
    UMRS_STAMP(counters, 5);
    if (v0 + c.cap < v_total) {   // another pass: reset the per-pass state
      if (threadIdx.x < 4) sm.ctl[threadIdx.x] = 0;
      if (threadIdx.x == 4) sm.ctl[kCtlNSurv] = 0;
      for (int k = threadIdx.x; k < n_chunks; k += blockDim.x) sm.cnt[k] = 0;
      __syncthreads();
    }
  }
  if (counters != nullptr && threadIdx.x == 0) {
    atomicAdd(counters + 3, (unsigned long long)pass);
    // source segments streamed by the gather (16 B each) | source rows, packed: rows << 40
    atomicAdd(counters + 7, (unsigned long long)v_total | ((unsigned long long)sm.ctl[kCtlNRows] << 40));
  }
  finalise(sm, fan, ranges, hit_idx, post);
#ifdef UMRS_PHASE_TIMING
  __syncthreads();
  UMRS_STAMP(counters, 6);
#endif
}

// ---- scan-only kernel (umrs_predicted_scan / umrs_scan_batch) -------------------
// kThreads x kCtas = the CTA shape and the resident fan CTAs per SM the register budget is
// held to (see the shape table at the launchers).
template <int kThreads, int kCtas, bool kCounters>
__global__ void __launch_bounds__(kThreads, kCtas)
k_scan_batch(const __grid_constant__ ScanBatchParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const SmemView sm = carve(smem_raw, p.c);
  if (kTmaBuilt && p.c.tma_stage != 0 && threadIdx.x == 0) stage_bar_init(sm.ctl + kCtlBar);
  unsigned stage_parity = 0;
  for (int f = blockIdx.x; f < p.n_fans; f += gridDim.x, stage_parity ^= 1u) {
    if (kTmaBuilt && p.c.tma_stage != 0) {
      __syncthreads();   // the previous fan is done with the arrays the staging area covers (and the barrier is initialised)
      if (threadIdx.x == 0) {
        stage_fence_generic_to_async();
        if (p.c.tma_stage == 1) stage_issue_map(sm, p.c.src);
      }
    }
    const float4 fp = __ldg(p.fans + f);
    Fan fan;
    fan.lx = fp.x; fan.ly = fp.y; fan.amin = fp.z; fan.amax = fp.w;
    fan.num_rays = p.num_rays;
    fan.range_max = p.range_max;
    PostProcess post;
    post.enabled = false;
    post.stddev = 0.f; post.key0 = post.key1 = post.step_lo = post.step_hi = post.fan_global = 0;
    const size_t off = (size_t)f * (size_t)p.num_rays;
    __syncthreads();  // previous fan's finalise is done with the best arrays
    scan_fan(sm, p.c, fan, p.obj_segs, p.n_obj, 0, 0, p.c.ranges + off,
             p.c.hit_idx ? p.c.hit_idx + off : nullptr, post, stage_parity, kCounters ? p.c.counters : nullptr);
  }
}

// ---- phase A: kinematic steps (per entity, one thread each) ------------------
__device__ __forceinline__ void boundf(float lo, float hi, float& x) {  // math_util.h:188-192
  if (x > hi) x = hi;
  if (x < lo) x = lo;
}
__device__ __forceinline__ int signf_(float v) { return (0.0f < v) - (v < 0.0f); }

// DriveCallback of each model: ackermann_model.cpp:41-50,
// diff_drive_model.cpp:170-197, omnidirectional_model.cpp:49-58.
// kCarsOnly: the roster holds Ackermann cars only (the launcher checked), so the other models'
// code is not even compiled into this instance -- the kernel's instruction footprint is what its
// no_instruction stalls pay for (same-box A/B: +1.5 % on the bench workload).
template <bool kCarsOnly>
__device__ void robot_command(RobotState& s, const umrs_robot_params& p, const float* cmd) {
  if (kCarsOnly || p.model_type == UMRS_ACKERMANN_DRIVE) {
    if (!isfinite(cmd[0]) || !isfinite(cmd[1])) return;
    s.cmd[0] = cmd[0]; s.cmd[1] = cmd[1]; s.cmd[2] = 0.f;
    s.cmd_age = 0;
  } else if (p.model_type == UMRS_DIFF_DRIVE) {
    double x = (double)cmd[0], z = (double)cmd[1];
    if (p.invert_linear_vel_cmds) x *= -1;
    if (p.invert_angular_vel_cmds) z *= -1;
    if ((double)p.max_linear_vel != 0.0) {
      if (fabs(x) > (double)p.max_linear_vel)
        x = (x > 0) ? (double)p.max_linear_vel : (double)-p.max_linear_vel;
    }
    if ((double)p.max_angular != 0.0) {
      if (fabs(z) > (double)p.max_angular)
        z = (z > 0) ? (double)p.max_angular : (double)-p.max_angular;
    }
    s.cmd[0] = (float)x; s.cmd[1] = (float)z; s.cmd[2] = 0.f;
    s.cmd_age = 0;
  } else {
    s.cmd[0] = cmd[0]; s.cmd[1] = cmd[1]; s.cmd[2] = cmd[2];
    s.cmd_age = 0;
  }
}

__device__ void accel_limit(double& v, float target, float pos_lim, float neg_lim) {
  // diff_drive_model.cpp:91-130 (limits per tick)
  if (v < (double)target) {
    if ((double)pos_lim == 0.0 || (double)target - v < (double)pos_lim) v = (double)target;
    else v += (double)pos_lim;
  } else if (v > (double)target) {
    if ((double)neg_lim == 0.0 || v - (double)target < (double)neg_lim) v = (double)target;
    else v -= (double)neg_lim;
  }
}

template <bool kCarsOnly>
__device__ void robot_step(RobotState& s, const umrs_robot_params& p, double dt, float n1, float n2) {
  const bool stale = (double)s.cmd_age * dt > 0.1;  // kMaxCommandAge in sim time
  if (kCarsOnly || p.model_type == UMRS_ACKERMANN_DRIVE) {
    // ackermann_model.cpp:52-96
    if (stale) s.cmd[0] = 0.f;
    const float vel = s.vel[0];
    const float kEpsilonCurvature = (float)(1.0 / 1E3);
    float desired_vel = s.cmd[0];
    boundf(-p.ak_max_speed, p.ak_max_speed, desired_vel);
    const float max_curvature = (float)(1.0 / (double)p.ak_min_turn_radius);
    float desired_curvature = s.cmd[1];
    boundf(-max_curvature, max_curvature, desired_curvature);
    const bool linear_motion = fabsf(desired_curvature) < kEpsilonCurvature;
    const float dv_max = (float)(dt * (double)p.ak_max_accel);
    float bounded_dv = desired_vel - vel;
    boundf(-dv_max, dv_max, bounded_dv);
    s.vel[0] = vel + bounded_dv;
    const float dist = (float)((double)s.vel[0] * dt);
    float dx = 0.f, dy = 0.f, dtheta = 0.f;
    if (linear_motion) {
      dx = dist;
      dtheta = (float)(dt * (double)p.ak_angular_error_bias);
    } else {
      const float r = (float)(1.0 / (double)desired_curvature);
      const float arc = dist * desired_curvature;
      dtheta = (float)((double)arc + ((double)n1 * dt) * (double)p.ak_angular_error_bias +
                       (double)((n2 * p.ak_angular_error_rate) * fabsf(arc)));
      const float2 sc = sincos_ref(dtheta);
      dx = r * sc.x;
      dy = (float)((double)r * (1.0 - (double)sc.y));
    }
    const float2 scp = sincos_ref(s.pose[2]);
    const float sn = scp.x, cs = scp.y;
    s.pose[0] = s.pose[0] + (cs * dx - sn * dy);
    s.pose[1] = s.pose[1] + (sn * dx + cs * dy);
    s.pose[2] = angle_mod_f(s.pose[2] + dtheta);
  } else if (p.model_type == UMRS_DIFF_DRIVE) {
    // diff_drive_model.cpp:83-148
    if (stale) { s.cmd[0] = 0.f; s.cmd[1] = 0.f; }
    accel_limit(s.linear_vel, s.cmd[0], p.linear_pos_accel_limit, p.linear_neg_accel_limit);
    accel_limit(s.angular_vel, s.cmd[1], p.angular_pos_accel_limit, p.angular_neg_accel_limit);
    const float acc_times_dt = (float)(s.linear_vel - (double)s.vel[0]);
    const float dx = (float)((double)s.vel[0] * dt + (0.5 * (double)acc_times_dt) * dt);
    s.vel[0] = (float)s.linear_vel;
    s.vel[2] = (float)s.angular_vel;
    const float2 scp = sincos_ref(s.pose[2]);
    s.pose[0] = s.pose[0] + scp.y * dx;
    s.pose[1] = s.pose[1] + scp.x * dx;
    s.pose[2] = (float)angle_mod_d((double)s.pose[2] + (double)s.vel[2] * dt);
  } else {
    // omnidirectional_model.cpp:82-123
    if (stale) { s.cmd[0] = 0.f; s.cmd[1] = 0.f; s.cmd[2] = 0.f; }
    float dvx = s.cmd[0], dvy = s.cmd[1];
    float z = dvx * dvx + dvy * dvy;
    if (sqrtf(z) > p.co_max_speed) {
      if (z > 0.f) { const float n = sqrtf(z); dvx = dvx / n; dvy = dvy / n; }
      dvx = dvx * p.co_max_speed;
      dvy = dvy * p.co_max_speed;
    }
    float ax = dvx - s.vel[0], ay = dvy - s.vel[1];
    const float max_accel = (float)((double)p.co_max_accel * dt);
    z = ax * ax + ay * ay;
    if (sqrtf(z) > max_accel) {
      if (z > 0.f) { const float n = sqrtf(z); ax = ax / n; ay = ay / n; }
      ax = ax * max_accel;
      ay = ay * max_accel;
    }
    s.vel[0] = s.vel[0] + ax;
    s.vel[1] = s.vel[1] + ay;
    float desired_ang = s.cmd[2];
    if (fabsf(desired_ang) > p.co_max_angle_vel)
      desired_ang = (float)signf_(desired_ang) * p.co_max_angle_vel;
    const float max_angle_accel = (float)((double)p.co_max_angle_accel * dt);
    float delta_ang = desired_ang - s.vel[2];
    if (fabsf(desired_ang) > max_angle_accel)  // :115 quirk kept
      delta_ang = (float)signf_(delta_ang) * max_angle_accel;
    s.vel[2] = s.vel[2] + delta_ang;
    const float2 scp = sincos_ref(s.pose[2]);
    const float sn = scp.x, cs = scp.y;
    const float rx = cs * s.vel[0] - sn * s.vel[1];
    const float ry = sn * s.vel[0] + cs * s.vel[1];
    const float dtf = (float)dt;
    s.pose[0] = s.pose[0] + rx * dtf;
    s.pose[1] = s.pose[1] + ry * dtf;
    s.pose[2] = (float)angle_mod_d((double)s.pose[2] + (double)s.vel[2] * dt);
  }
  if (s.cmd_age < kAgeNever) s.cmd_age += 1;
}

// HumanObject::Step + CheckReachGoal, human_object.cpp:147-161,178-192.
__device__ void human_step(HumanState& s, const umrs_human_params& p, double dt) {
  const double avg_speed = (double)p.avg_speed, max_speed = (double)p.max_speed;
  float gx = s.goal[0] - s.pose[0], gy = s.goal[1] - s.pose[1];
  float z = gx * gx + gy * gy;
  if (z > 0.f) { const float n = sqrtf(z); gx = gx / n; gy = gy / n; }
  float vx = gx * (float)avg_speed, vy = gy * (float)avg_speed;
  z = vx * vx + vy * vy;
  if ((double)sqrtf(z) > max_speed) {
    if (z > 0.f) { const float n = sqrtf(z); vx = vx / n; vy = vy / n; }
    vx = vx * (float)max_speed;
    vy = vy * (float)max_speed;
  }
  s.trans_vel[0] = vx;
  s.trans_vel[1] = vy;
  s.pose[2] = (float)((double)s.pose[2] + (double)s.rot_vel * dt);
  s.pose[0] = s.pose[0] + vx * (float)dt;
  s.pose[1] = s.pose[1] + vy * (float)dt;
  const float ex = s.pose[0] - s.goal[0], ey = s.pose[1] - s.goal[1];
  if ((double)sqrtf(ex * ex + ey * ey) < (double)p.reach_goal_threshold) {
    if (p.mode == 0) {
      s.trans_vel[0] = 0.f; s.trans_vel[1] = 0.f; s.rot_vel = 0.f;
    } else if (p.mode == 1) {
      for (int k = 0; k < 3; ++k) { const float t = s.goal[k]; s.goal[k] = s.start[k]; s.start[k] = t; }
    }
  }
}

// ---- phase A for one entity --------------------------------------------------------
// Entity e of world w: static objects first, then humans, then robots.  Steps it from the
// read-only "in" state, stores the "out" state when `write`, returns the post-step pose.
template <bool kCarsOnly>
__device__ __forceinline__ void step_entity(const WorldStepParams& p, int w, int e, bool write, double dt,
                                            float& px, float& py, float& pth) {
  if (e < p.O) {
    const float* sp = p.static_pose + ((size_t)w * p.O + e) * 3;
    px = sp[0]; py = sp[1]; pth = sp[2];
  } else if (!kCarsOnly && e < p.O + p.H) {
    const int h = e - p.O;
    HumanState s = p.hs_in[(size_t)w * p.H + h];
    human_step(s, p.humans[h], dt);
    if (write) p.hs_out[(size_t)w * p.H + h] = s;
    px = s.pose[0]; py = s.pose[1]; pth = s.pose[2];
  } else {
    const int rr = e - p.O - p.H;
    const size_t gi = (size_t)w * p.R + rr;
    RobotState s = p.rs_in[gi];
    const umrs_robot_params& rp = p.robots[rr];
    if (p.cmds != nullptr && (p.cmd_mask == nullptr || p.cmd_mask[gi] != 0)) {
      const float cmd[3] = {p.cmds[gi * 3 + 0], p.cmds[gi * 3 + 1], p.cmds[gi * 3 + 2]};
      robot_command<kCarsOnly>(s, rp, cmd);
    }
    float n1 = 0.f, n2 = 0.f;
    if (p.motion_noise && (kCarsOnly || rp.model_type == UMRS_ACKERMANN_DRIVE)) {
      const uint4 wd = philox4x32(
          make_uint4(0u, (uint32_t)(p.world_base + w) * (uint32_t)p.R + (uint32_t)rr,
                     p.step_lo, p.step_hi ^ kStreamMotion),
          make_uint2(p.key0, p.key1));
      const float2 g = box_muller(wd.x, wd.y);
      n1 = g.x; n2 = g.y;
    }
    robot_step<kCarsOnly>(s, rp, dt, n1, n2);
    if (write) {
      p.rs_out[gi] = s;
      if (p.robot_pose_out != nullptr) {
        p.robot_pose_out[gi * 3 + 0] = s.pose[0];
        p.robot_pose_out[gi * 3 + 1] = s.pose[1];
        p.robot_pose_out[gi * 3 + 2] = s.pose[2];
      }
    }
    px = s.pose[0]; py = s.pose[1]; pth = s.pose[2];
  }
}

// laserLoc = t + Rotation2Df(angle) * (laser_x, laser_y) with sin/cos of the raw pose angle
// (float overloads inside Eigen), simulator.cpp:398-401
__device__ __forceinline__ float4 fan_of(const umrs_robot_params& me, float px, float py, float pth,
                                         float sn, float cs) {
  return make_float4(px + (cs * me.laser.loc_x - sn * me.laser.loc_y),
                     py + (sn * me.laser.loc_x + cs * me.laser.loc_y),
                     me.laser.angle_min + pth, me.laser.angle_max + pth);
}

// R * template + T (human_object.cpp:135-138)
__device__ __forceinline__ float4 place_line(const float4 tm, float px, float py, float sn, float cs) {
  float4 o;
  o.x = (cs * tm.x - sn * tm.y) + px;
  o.y = (sn * tm.x + cs * tm.y) + py;
  o.z = (cs * tm.z - sn * tm.w) + px;
  o.w = (sn * tm.z + cs * tm.w) + py;
  return o;
}

// ---- entity kernel of the two-launch step --------------------------------------------
// One thread per (world, entity): the step, the entity's object lines and -- for a robot -- its
// fan, all to global memory; k_world_step<..., kPreStepped> then runs the scans from there.
// Splitting the tick takes the serial entity step (one thread, ~5 k cycles of double-precision
// math) off the critical path of every fan CTA; the fused form stays for small batches, where
// one launch is the shorter path.
template <bool kCarsOnly>
__global__ void __launch_bounds__(128)
k_world_entities(const __grid_constant__ WorldStepParams p) {
  const int n_ent = p.O + p.H + p.R;
  const int g = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (g >= p.world_count * n_ent) return;
  const int w = p.world_first + g / n_ent, e = g % n_ent;
  float px, py, pth;
  step_entity<kCarsOnly>(p, w, e, true, (double)p.dt, px, py, pth);
  const float2 sca = sincos_ref(angle_mod_f(pth));   // HumanObject::Transform rotation, human_object.cpp:132
  const int l0 = p.ent_off[e], l1 = p.ent_off[e + 1];
  for (int l = l0; l < l1; ++l)
    p.obj_lines_out[(size_t)w * p.L_obj + l] = place_line(__ldg(p.tmpl + l), px, py, sca.x, sca.y);
  if (e >= p.O + p.H) {
    const int r = e - p.O - p.H;
    const float2 scl = sincos_ref(pth);
    p.fan_params[(size_t)w * p.R + r] = fan_of(p.robots[r], px, py, pth, scl.x, scl.y);
  }
}

// ---- fused world-step kernel -------------------------------------------------
// grid = world_count * R fans; CTA (w, r).  Every CTA of world w integrates ALL
// entities of that world (R + H tiny steps, redundantly across the R CTAs) from
// the read-only "in" state so that no inter-CTA exchange is needed; the r == 0
// CTA alone writes the "out" state (ping-pong buffers).  Order as in
// Simulator::Run (simulator.cpp:531-541): update() first, scan from the
// post-step pose against the post-step object lines.
// kPreStepped: k_world_entities has run -- the fan and the object lines come from global memory.
template <int kThreads, int kCtas, bool kCarsOnly, bool kCounters, bool kPreStepped>
__global__ void __launch_bounds__(kThreads, kCtas)
k_world_step(const __grid_constant__ WorldStepParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n_ent = p.O + p.H + p.R;
  const SmemView sm = carve(smem_raw, p.c);
  const int w = p.world_first + (int)(blockIdx.x / (unsigned)p.R);
  const int r = (int)(blockIdx.x % (unsigned)p.R);
  const bool owner = (r == 0);
  const double dt = (double)p.dt;  // `const double& dt` bound to float CONFIG_DT
  const int t = threadIdx.x;
  UMRS_STAMP_INIT;
  const umrs_robot_params& me = p.robots[r];
  Fan fan;
  const float4* obj = sm.obj;
  if (kPreStepped) {
    const float4 fp = p.fan_params[(size_t)w * p.R + r];
    fan.lx = fp.x; fan.ly = fp.y; fan.amin = fp.z; fan.amax = fp.w;
    obj = p.obj_lines_out + (size_t)w * p.L_obj;
  } else {
    if (kTmaBuilt && p.c.tma_stage != 0 && t == 32 % blockDim.x) {
      // the map streams into shared memory (copy engine) while thread 0 steps the robot
      stage_bar_init(sm.ctl + kCtlBar);
      if (p.c.tma_stage == 1) stage_issue_map(sm, p.c.src);
    }

    // ---- phase A: entity steps -> sm.ent[e] = {x, y, theta, sin, cos} ----------
    for (int e = t; e < n_ent; e += blockDim.x) {
      float px, py, pth;
      step_entity<kCarsOnly>(p, w, e, owner, dt, px, py, pth);
      // HumanObject::Transform rotation, human_object.cpp:132: R(AngleMod(angle))
      const float am = angle_mod_f(pth);
      float* en = sm.ent + 8 * e;
      en[0] = px; en[1] = py; en[2] = pth;
      const float2 sca = sincos_ref(am);
      en[3] = sca.x; en[4] = sca.y;
      if (e >= p.O + p.H) {   // laser pose of a robot: sin/cos of the raw angle (simulator.cpp:398-401)
        const float2 scl = sincos_ref(pth);
        en[5] = scl.x; en[6] = scl.y;
      }
    }
    __syncthreads();

    // ---- object lines: R * template + T (human_object.cpp:135-138) --------------
    for (int l = t; l < p.L_obj; l += blockDim.x) {
      const float* en = sm.ent + 8 * p.line_ent[l];
      const float4 o = place_line(__ldg(p.tmpl + l), en[0], en[1], en[3], en[4]);
      sm.obj[l] = o;
      if (owner && p.obj_lines_out != nullptr) p.obj_lines_out[(size_t)w * p.L_obj + l] = o;
    }

    // ---- fan of robot r (simulator.cpp:398-411) ----------------------------------
    const float* en = sm.ent + 8 * (p.O + p.H + r);
    const float4 fp = fan_of(me, en[0], en[1], en[2], en[5], en[6]);
    fan.lx = fp.x; fan.ly = fp.y; fan.amin = fp.z; fan.amax = fp.w;
    if (t == 0 && p.fan_params != nullptr) p.fan_params[(size_t)w * p.R + r] = fp;
    __syncthreads();
    if (p.skip_scan) return;   // reference-render worlds: the scan is k_scan_reference_render's
  }
  fan.num_rays = p.num_rays[r];
  fan.range_max = me.laser.range_max;

  const int skip_lo = p.ent_off[p.O + p.H + r], skip_hi = p.ent_off[p.O + p.H + r + 1];
  PostProcess post;
  post.enabled = p.laser_noise != 0;
  post.stddev = me.laser.noise_stddev;
  post.key0 = p.key0; post.key1 = p.key1;
  post.step_lo = p.step_lo; post.step_hi = p.step_hi;
  post.fan_global = (uint32_t)(p.world_base + w) * (uint32_t)p.R + (uint32_t)r;
  const size_t off = (size_t)w * (size_t)p.rays_per_world + (size_t)p.ray_off[r];
  UMRS_STAMP(p.c.counters, 0);
  scan_fan(sm, p.c, fan, obj, p.L_obj, skip_lo, skip_hi, p.c.ranges + off,
           p.c.hit_idx ? p.c.hit_idx + off : nullptr, post, 0u, kCounters ? p.c.counters : nullptr);
}

// ---- K3: statistics of a ranges buffer -------------------------------------------
// out4 = {count of finite positive ranges, min, max, sum}; one atomic set per warp.
__global__ void k3_stats(const float* __restrict__ ranges, size_t n, double* out4) {
  double cnt = 0.0, sum = 0.0;
  float mn = __int_as_float(0x7f800000), mx = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    const float v = ranges[i];
    if (v > 0.f && v < __int_as_float(0x7f800000)) {
      cnt += 1.0; sum += (double)v;
      mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    sum += __shfl_down_sync(0xffffffffu, sum, o);
    mn = fminf(mn, __shfl_down_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(out4 + 0, cnt);
    atomicAdd(out4 + 3, sum);
    // positive floats order like their bit patterns
    atomicMin(reinterpret_cast<unsigned long long*>(out4 + 1),
              (unsigned long long)__float_as_uint(mn));
    atomicMax(reinterpret_cast<unsigned long long*>(out4 + 2),
              (unsigned long long)__float_as_uint(mx));
  }
}

// ---- launchers ----------------------------------------------------------------
// CTA shapes.  The fan code needs 64 registers to run without spills, so an SM holds 1024
// fan threads whatever the shape; what differs is how many independent fans those threads
// serve.  A fan's phases are separated by barriers, and it is the other resident fans that
// fill an SM while one waits, so more and smaller CTAs win as long as shared memory seats
// them (same-box A/B on 8192 worlds):
//     shape            regs  64-seg map  256-seg map  needs shared memory <=
//     256 thr x 4 CTAs  64   0.207 ms    0.229 ms     56 KB  (any plan)
//     192 thr x 5 CTAs  64   0.191 ms    0.217 ms     44.6 KB
//     192 thr x 6 CTAs  56   0.180 ms    0.210 ms     37 KB
//     128 thr x 8 CTAs  64   0.159 ms    0.204 ms     27.5 KB
//      96 thr x 10 CTAs 64   0.148 ms    0.209 ms     21.8 KB (used up to 18 KB)
// (256 threads at 48 / 40 registers for 5 / 6 CTAs: 0.198 / 0.190 ms, spills; more threads
// per CTA, 320 / 384 at 48 / 40 registers: slower than 256 on every map.)
// The shape is picked per launch: 96 x 10, 128 x 8 or 192 x 6 if the plan's shared memory seats it
// (asked of the occupancy calculator), else 256 x 4; a grid too small to fill the SMs at
// that count keeps the widest CTA (a single scan runs fastest on 256 threads).  The
// 1024-segment bench map (52.4 KB) stays at 256 x 4: squeezed to 44 KB (per-ray winners
// kept in dead record fields) and run as 192 x 5 it was 2.5 % slower, so that was dropped.
enum { kShape256x4 = 0, kShape192x5 = 1, kShape192x6 = 2, kShape128x8 = 3, kShape96x10 = 4, kShape512x2 = 5, kShape320x3 = 6, kNumShapes = 7 };
struct DevLimits {
  int n_sm; int smem_per_sm; bool known;
  size_t configured[kNumShapes];
  size_t pick_smem; int pick; bool pick_known;
};
static DevLimits g_dev[64];
static std::mutex g_dev_mutex;

static cudaError_t dev_limits(DevLimits** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  DevLimits& d = g_dev[dev & 63];
  if (!d.known) {
    if ((e = cudaDeviceGetAttribute(&d.n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&d.smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev)) != cudaSuccess) return e;
    d.known = true;
  }
  *out = &d;
  return cudaSuccess;
}

template <int kThreads, int kCtas>
static bool seats(size_t smem) {
  int n = 0, m = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_world_step<kThreads, kCtas, false, false, false>, kThreads, smem) != cudaSuccess ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&m, k_scan_batch<kThreads, kCtas, false>, kThreads, smem) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return n >= kCtas && m >= kCtas;
}

constexpr size_t kWideCtaFrom = 76 * 1024;   // above this only two CTAs fit the 227 KB of an SM

static int shape_ctas(int shape) {
  return shape == kShape96x10 ? 10 : shape == kShape128x8 ? 8 : shape == kShape192x6 ? 6 : shape == kShape192x5 ? 5 :
         shape == kShape512x2 ? 2 : shape == kShape320x3 ? 3 : 4;
}

static int pick_shape(DevLimits& d, size_t smem, int fans) {
  // umrs_tuning_set("cta_shape"): experiments and tests.  Every shape runs any plan that fits the
  // opt-in shared memory at all (a plan too big to seat kCtas CTAs just runs at lower occupancy).
  const int forced = g_tuning.cta_shape;
  if (forced >= 0 && forced < kNumShapes) return forced;
  if (!d.pick_known || d.pick_smem != smem) {
    int c = kShape256x4;
    if (smem <= 48 * 1024) {   // above 48 KB not even 5 CTAs fit 228 KB: no need to ask
      // 96 x 10 seats up to 21.8 KB but loses to 128 x 8 from ~19 KB (192 segments) on
      if (smem <= 18 * 1024 && seats<96, 10>(smem)) c = kShape96x10;
      else if (seats<128, 8>(smem)) c = kShape128x8;
      else if (seats<192, 6>(smem)) c = kShape192x6;
      // 192 x 5 (37-44.6 KB) is not picked: the gain of a smaller CTA shrinks as the
      // candidate count grows, and at 1024 segments 192 x 5 lost 2.5 % to 256 x 4
    } else if (smem > kWideCtaFrom) {
      // plans so large that only two fans fit an SM (3601-ray fans on a 1500-segment map: 115 KB): 256
      // threads a fan would leave the SM at 16 warps -- twice the threads per fan instead
      c = kShape512x2;
    } else if (4 * (smem + 1024) > (size_t)d.smem_per_sm) {   // four CTAs (+1 KB each, reserved) do not fit
      // three fans per SM (1100-1536-segment maps, the 100k-segment grid walk): 320 threads a fan put
      // 30 warps on the SM instead of 24 (+4-8 %; 512 x 2 loses 6-11 % here)
      c = kShape320x3;
    }
    d.pick_smem = smem;
    d.pick = c;
    d.pick_known = true;
  }
  // a grid that is resident all at once even at two CTAs per SM is a latency question: the widest CTA
  // (one 1081-ray scan 47 -> 38 us through the host-buffer call, a 148-world tick 29 -> 24 us)
  if ((long long)fans <= 2LL * d.n_sm) return kShape512x2;
  // the grid must fill every SM at (almost) the shape's CTA count for the smaller CTA to pay
  if (d.pick != kShape512x2 && (long long)fans < (long long)(shape_ctas(d.pick) - 1) * d.n_sm) return kShape256x4;
  return d.pick;
}

template <int kThreads, int kCtas>
static cudaError_t configure_shape(DevLimits& d, int shape, size_t smem_bytes) {
  if (smem_bytes <= d.configured[shape]) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(k_scan_batch<kThreads, kCtas, false>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_world_step<kThreads, kCtas, false, false, false>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_world_step<kThreads, kCtas, true, false, false>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_world_step<kThreads, kCtas, false, false, true>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  d.configured[shape] = smem_bytes;
  return cudaSuccess;
}

template <int kThreads, int kCtas>
static cudaError_t launch_scan_shape(DevLimits& d, int shape, const ScanBatchParams& p, size_t smem,
                                     cudaStream_t stream) {
  cudaError_t e = configure_shape<kThreads, kCtas>(d, shape, smem);
  if (e != cudaSuccess) return e;
  k_scan_batch<kThreads, kCtas, false><<<p.n_fans, kThreads, smem, stream>>>(p);
  return cudaSuccess;
}

template <int kThreads, int kCtas>
static cudaError_t launch_step_shape(DevLimits& d, int shape, const WorldStepParams& p, int fans,
                                     size_t smem, cudaStream_t stream) {
  cudaError_t e = configure_shape<kThreads, kCtas>(d, shape, smem);
  if (e != cudaSuccess) return e;
  // a roster of Ackermann cars only (and no humans) runs the instance compiled without the other models
  if (p.pre_stepped) k_world_step<kThreads, kCtas, false, false, true><<<fans, kThreads, smem, stream>>>(p);
  else if (p.cars_only) k_world_step<kThreads, kCtas, true, false, false><<<fans, kThreads, smem, stream>>>(p);
  else k_world_step<kThreads, kCtas, false, false, false><<<fans, kThreads, smem, stream>>>(p);
  return cudaSuccess;
}

// Work accounting on (umrs_*_enable_counters): ONE generic instance per kernel carries the accounting
// code (256 x 4, every model); the counts do not depend on the CTA shape.
static size_t g_counted_smem[2] = {0, 0};
template <typename Kernel, typename Params>
static cudaError_t launch_counted(Kernel kernel, int which, const Params& p, int ctas, size_t smem, cudaStream_t stream) {
  if (smem > g_counted_smem[which]) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    g_counted_smem[which] = smem;
  }
  kernel<<<ctas, 256, smem, stream>>>(p);
  note_launch();
  return cudaGetLastError();
}

cudaError_t launch_scan_batch(const ScanBatchParams& p_in, cudaStream_t stream) {
  if (p_in.n_fans <= 0) return cudaSuccess;
  ScanBatchParams p = p_in;
  const size_t smem = fill_smem_offsets(p.c, 0, 0);
  std::lock_guard<std::mutex> lock(g_dev_mutex);
  if (p.c.counters != nullptr) return launch_counted(k_scan_batch<256, 4, true>, 0, p, p.n_fans, smem, stream);
  DevLimits* d = nullptr;
  cudaError_t e = dev_limits(&d);
  if (e != cudaSuccess) return e;
  const int shape = pick_shape(*d, smem, p.n_fans);
  switch (shape) {
    case kShape96x10: e = launch_scan_shape<96, 10>(*d, shape, p, smem, stream); break;
    case kShape128x8: e = launch_scan_shape<128, 8>(*d, shape, p, smem, stream); break;
    case kShape192x6: e = launch_scan_shape<192, 6>(*d, shape, p, smem, stream); break;
    case kShape192x5: e = launch_scan_shape<192, 5>(*d, shape, p, smem, stream); break;
    case kShape512x2: e = launch_scan_shape<512, 2>(*d, shape, p, smem, stream); break;
    case kShape320x3: e = launch_scan_shape<320, 3>(*d, shape, p, smem, stream); break;
    default: e = launch_scan_shape<256, 4>(*d, shape, p, smem, stream); break;
  }
  if (e != cudaSuccess) return e;
  note_launch();
  return cudaGetLastError();
}

cudaError_t launch_world_step(const WorldStepParams& p_in, cudaStream_t stream) {
  const int fans = p_in.world_count * p_in.R;
  if (fans <= 0) return cudaSuccess;
  WorldStepParams p = p_in;
  const size_t smem = fill_smem_offsets(p.c, p.L_obj, p.O + p.H + p.R);
  std::lock_guard<std::mutex> lock(g_dev_mutex);   // the per-device shape cache is shared by all handles
  if (p.c.counters != nullptr) return launch_counted(k_world_step<256, 4, false, true, false>, 1, p, fans, smem, stream);
  DevLimits* d = nullptr;
  cudaError_t e = dev_limits(&d);
  if (e != cudaSuccess) return e;
  const int shape = pick_shape(*d, smem, fans);
  // Two-launch step (umrs_tuning "split_step": -1 auto, 0 never, 1 always): worth a second launch
  // once the fans fill the SMs more than twice over (measured break-even: two resident waves, at
  // 256 x 4 as at 96 x 10); reference-render worlds are split already.
  const bool can_split = !p.skip_scan && p.fan_params != nullptr && p.obj_lines_out != nullptr;
  const bool split = can_split && (g_tuning.split_step > 0 ||
                                   (g_tuning.split_step < 0 && (long long)fans > 2LL * shape_ctas(shape) * d->n_sm));
  if (split) {
    const int n_thr = p.world_count * (p.O + p.H + p.R);
    if (p.cars_only) k_world_entities<true><<<(n_thr + 127) / 128, 128, 0, stream>>>(p);
    else k_world_entities<false><<<(n_thr + 127) / 128, 128, 0, stream>>>(p);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    note_launch();
    p.pre_stepped = 1;
  }
  switch (shape) {
    case kShape96x10: e = launch_step_shape<96, 10>(*d, shape, p, fans, smem, stream); break;
    case kShape128x8: e = launch_step_shape<128, 8>(*d, shape, p, fans, smem, stream); break;
    case kShape192x6: e = launch_step_shape<192, 6>(*d, shape, p, fans, smem, stream); break;
    case kShape192x5: e = launch_step_shape<192, 5>(*d, shape, p, fans, smem, stream); break;
    case kShape512x2: e = launch_step_shape<512, 2>(*d, shape, p, fans, smem, stream); break;
    case kShape320x3: e = launch_step_shape<320, 3>(*d, shape, p, fans, smem, stream); break;
    default: e = launch_step_shape<256, 4>(*d, shape, p, fans, smem, stream); break;
  }
  if (e != cudaSuccess) return e;
  note_launch();
  return cudaGetLastError();
}

cudaError_t launch_stats(const float* ranges, size_t n, double* out4_dev, cudaStream_t stream) {
  // out4_dev must be pre-initialised by the caller: {0, bits(+inf), 0, 0}
  k3_stats<<<148 * 4, 256, 0, stream>>>(ranges, n, out4_dev);
  note_launch();
  return cudaGetLastError();
}

}  // namespace umrs
