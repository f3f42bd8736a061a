// Host-only helpers of libumrs (no CUDA types).
#pragma once
#include <vector>
#include "../../include/umrs.h"

namespace umrs {
// VectorMap::Cleanup + ShrinkLine (vector_map.cpp:359-395).  false: the split queue ran away
// (inputs on which the reference's own loop never ends).
bool CleanupSegments(const float* in_xyxy, int n, std::vector<float>* out_xyxy);
// The same output by the reference's quadratic loop / through a uniform grid / chosen by size.
bool CleanupSegmentsQuadratic(const float* in_xyxy, int n, std::vector<float>* out_xyxy);
bool CleanupSegmentsGrid(const float* in_xyxy, int n, std::vector<float>* out_xyxy);
bool CleanupSegmentsWith(const float* in_xyxy, int n, int algorithm, std::vector<float>* out_xyxy);
// VectorMap::Load's parse loop (vector_map.cpp:404-407); false if unreadable.
bool LoadMapFile(const char* path, std::vector<float>* raw_xyxy);
// Footprint templates at pose (0,0,0) (entity_base.h:37-38).
void HumanTemplate(const umrs_human_params& p, std::vector<float>* tmpl_xyxy);
void SquareTemplate(float* tmpl_xyxy16);
void RobotTemplate(const umrs_robot_params& p, float* tmpl_xyxy16);
}  // namespace umrs
