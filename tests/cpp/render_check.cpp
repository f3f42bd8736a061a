// Test harness: runs the host build of csrc/umrs_render.h (the restated TrimOcclusion /
// SceneRender that the device kernel k_scan_reference_render executes) so that
// tests/test_render.py can compare it, without a GPU, with the compiled reference's own
// SceneRender.  Built by the test with g++ -ffp-contract=off.
#include <string.h>

#include <vector>

#include "../../ut_multirobot_sim_b200/csrc/umrs_render.h"

using umrs::render::Seg;
using umrs::render::V2;

extern "C" int umrs_test_scene_render(const float* lines_xyxy, int n, float lx, float ly,
                                      float* out_xyxy, int cap) {
  std::vector<Seg> lines(umrs::render::kRenderMaxLines), scene(umrs::render::kRenderMaxLines);
  const int stored = n < umrs::render::kRenderMaxLines ? n : umrs::render::kRenderMaxLines;
  if (stored > 0) memcpy(lines.data(), lines_xyxy, sizeof(Seg) * (size_t)stored);
  V2 loc;
  loc.x = lx;
  loc.y = ly;
  const int n_scene = umrs::render::scene_render_sequential(loc, lines.data(), n, scene.data());
  int k = 0;
  for (int j = 0; j < n_scene; ++j) {
    if (!(umrs::render::sq_length(scene[j]) > umrs::render::kRenderEps)) continue;   // :255-257
    if (k < cap) memcpy(out_xyxy + 4 * k, &scene[j], sizeof(Seg));
    ++k;
  }
  return k;
}
