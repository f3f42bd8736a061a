// Test harness (CPU): the host build of csrc/umrs_render.h (the restated TrimOcclusion /
// SceneRender, vector_map.cpp:55-258) run with PROVENANCE -- every candidate, split-off piece
// and rendered scene entry carries the index of the GetSceneLines candidate it was cut from,
// and every piece the render drops for being <= 1 cm (eps on the squared length,
// vector_map.cpp:222,238,256) is recorded.  tests/parity_tools.py uses it to say WHY a ray of
// the reference's GetPredictedScan differs from the exact nearest hit (classes D1-D6 of
// SURVEY.md 8c).  The per-ray fill (vector_map.cpp:443-477) is the oracle port's
// orc_candidate_range / orc_ray_angle; tests pin the result to the compiled reference's
// GetPredictedScan bit for bit before trusting the provenance.
#include <math.h>
#include <string.h>

#include <vector>

#include "../../oracle/umrs_oracle.h"
#include "../../ut_multirobot_sim_b200/csrc/umrs_render.h"

using umrs::render::Seg;
using umrs::render::V2;
using umrs::render::kRenderEps;
using umrs::render::kRenderMaxLines;
using umrs::render::sq_length;
using umrs::render::trim_occlusion;

namespace {
bool is_sentinel(const Seg& s) {   // the "completely occluded" marker, vector_map.cpp:143-149
  return s.p0.x == 0.1f && s.p0.y == 0.1f && s.p1.x == 0.1f && s.p1.y == 0.1f;
}
}  // namespace

// cand_xyxy: the GetSceneLines candidates (AABB-culled, map then object lines), n_cand of them.
// Outputs:
//   ranges[num_rays], ray_piece[num_rays] (index into pieces, -1: none covered the ray)
//   pieces[cap_p][4] / piece_parent[cap_p]: the rendered scene (entries with len^2 > eps), in order
//   dead[cap_d][4] / dead_parent[cap_d]: pieces dropped with 0 < len^2 <= eps
//   counts[4] = {n_pieces, n_dead, cap hit (candidate list reached 2000), complete-occlusion sentinels}
extern "C" void umrs_test_render_provenance(const float* cand_xyxy, int n_cand, float lx, float ly,
                                            float range_max, float angle_min, float angle_max,
                                            int num_rays, float* ranges, int* ray_piece,
                                            float* pieces, int* piece_parent, int cap_p,
                                            float* dead, int* dead_parent, int cap_d, int* counts) {
  std::vector<Seg> lines(kRenderMaxLines), scene(kRenderMaxLines);
  std::vector<int> lpar(kRenderMaxLines), spar(kRenderMaxLines);
  int n_lines = n_cand;
  const int stored = n_cand < kRenderMaxLines ? n_cand : kRenderMaxLines;
  if (stored > 0) memcpy(lines.data(), cand_xyxy, sizeof(Seg) * (size_t)stored);
  for (int i = 0; i < stored; ++i) lpar[i] = i;
  V2 loc;
  loc.x = lx;
  loc.y = ly;
  int n_scene = 0, n_dead = 0, n_sentinel = 0;
  auto note_dead = [&](const Seg& s, int parent) {
    if (is_sentinel(s)) { ++n_sentinel; return; }
    if (!(sq_length(s) > 0.0f)) return;
    if (n_dead < cap_d) { memcpy(dead + 4 * n_dead, &s, sizeof(Seg)); dead_parent[n_dead] = parent; }
    ++n_dead;
  };
  // same statements, same order as umrs::render::scene_render_sequential (tested equal below
  // through the ranges), plus the bookkeeping
  for (int i = 0; i < n_lines && i < kRenderMaxLines; ++i) {
    Seg cur = lines[i];
    const int cp = lpar[i];
    Seg piece;
    for (int j = 0; j < n_scene && sq_length(cur) >= kRenderEps; ++j) {
      if (sq_length(scene[j]) < kRenderEps) continue;
      if (trim_occlusion(loc, scene[j], &cur, &piece)) {
        if (n_lines < kRenderMaxLines) { lines[n_lines] = piece; lpar[n_lines] = cp; }
        ++n_lines;
      }
    }
    if (sq_length(cur) > kRenderEps) {
      for (int j = 0; j < n_scene; ++j) {
        if (sq_length(scene[j]) < kRenderEps) continue;
        const bool pushed = trim_occlusion(loc, cur, &scene[j], &piece);
        if (pushed) {
          if (n_lines < kRenderMaxLines) { lines[n_lines] = piece; lpar[n_lines] = spar[j]; }
          ++n_lines;
        }
        if (!(sq_length(scene[j]) > kRenderEps)) note_dead(scene[j], spar[j]);
      }
      scene[n_scene] = cur;
      spar[n_scene] = cp;
      ++n_scene;
    } else {
      note_dead(cur, cp);
    }
  }
  int n_pieces = 0;
  std::vector<Seg> live;
  for (int j = 0; j < n_scene; ++j) {
    if (!(sq_length(scene[j]) > kRenderEps)) continue;   // :255-257
    if (n_pieces < cap_p) { memcpy(pieces + 4 * n_pieces, &scene[j], sizeof(Seg)); piece_parent[n_pieces] = spar[j]; }
    live.push_back(scene[j]);
    ++n_pieces;
  }
  counts[0] = n_pieces;
  counts[1] = n_dead;
  counts[2] = n_lines >= kRenderMaxLines ? 1 : 0;
  counts[3] = n_sentinel;
  // per-ray fill: first rendered piece whose interval contains the ray (:466-477)
  for (int i = 0; i < num_rays; ++i) {
    ranges[i] = range_max;
    ray_piece[i] = -1;
    for (size_t k = 0; k < live.size(); ++k) {
      const float r = orc_candidate_range(&live[k].p0.x, lx, ly, angle_min, angle_max, num_rays, i);
      if (r == r) { ranges[i] = r; ray_piece[i] = (int)k; break; }
    }
  }
}
