/* TEST INFRASTRUCTURE ONLY -- never linked into or called by the product.
 *
 * Plain-C CPU restatement ("port") of the reference algorithm for the hot path
 * (ut-amrl/ut_multirobot_sim; citations are file:line in the reference tree).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this.
 *
 * Parity status: the scan part is PINNED against the reference itself run in
 * the build container (oracle/_ref/libumrs_ref.so = the unmodified
 * vector_map.cpp; fixtures in tests/golden/ made by tests/golden/make_golden.py).
 * The kinematic steps (ackermann/diff-drive/omni/human) are PINNED as well: the
 * reference's own model sources compile unmodified into oracle/_ref against
 * data-only ROS stand-ins (oracle/shim, oracle/ref_models_bridge.cpp), and
 * tests/test_ref_models.py steps them side by side with the restatement below,
 * bit for bit after every step.  Tier B is additionally pinned as the exact
 * nearest hit by an independent float64 ray cast (tests/parity_tools.py).
 */
#ifndef UMRS_ORACLE_H_
#define UMRS_ORACLE_H_

#include <stdint.h>
#include "../include/umrs.h" /* POD parameter structs only */

#ifdef __cplusplus
extern "C" {
#endif

/* ---- scan, tier B ------------------------------------------- */
/* Per-ray minimum over ALL candidates of the reference's own range formula,
 * candidates in GetSceneLines order (map then objects, AABB-culled,
 * vector_map.cpp:192-215).  Objects with index in [skip_lo, skip_hi) are left
 * out (a robot's own footprint).  idx: index into map||objects, -1 = none. */
void orc_scan_tierb(const float* map_xyxy, int32_t n_map, const float* obj_xyxy,
                    int32_t n_obj, int32_t skip_lo, int32_t skip_hi, float loc_x,
                    float loc_y, float range_max, float angle_min, float angle_max,
                    int32_t num_rays, float* ranges, int32_t* idx);

/* Ray angle i of a fan: AngleMod(angle_min + i*da), da = span/num_rays
 * (vector_map.cpp:465-467). */
float orc_ray_angle(float angle_min, float angle_max, int32_t num_rays, int32_t i);

/* Range of ONE candidate at ray i by the same formula, NaN when the ray is
 * outside the candidate's angular interval or the candidate is dropped
 * (|a0-a1| < 1e-4).  Used by tests to certify ties. */
float orc_candidate_range(const float seg_xyxy[4], float loc_x, float loc_y,
                          float angle_min, float angle_max, int32_t num_rays,
                          int32_t i);

/* Laser post-process, simulator.cpp:412-418.  normals: one N(0,1) per ray
 * (NULL = noise off: the r > range_max - 0.1 -> 0 rule and max(0, r) still apply
 * only when apply_dropout != 0). */
void orc_laser_postprocess(float* ranges, int32_t n, float range_max, float stddev,
                           const float* normals);

/* Laser origin, simulator.cpp:398-400. */
void orc_laser_origin(const float pose_xyt[3], float laser_x, float laser_y,
                      float out_xy[2]);

/* ---- map cleanup (vector_map.cpp:359-395) --------------------------------- */
int32_t orc_map_cleanup(const float* in_xyxy, int32_t n, float* out_xyxy, int32_t cap);

/* Line<float>::Intersection (line2d.h:139-160), the predicate Cleanup splits
 * with: returns 1 and the crossing point of segments a and b, else 0. */
int32_t orc_line_intersection(const float a_xyxy[4], const float b_xyxy[4], float out_xy[2]);

/* ---- kinematic steps ---------------------------------------------------------- */
typedef struct orc_robot_state {
  float pose[3];      /* pose_  {x, y, angle}         entity_base.h:36 */
  float vel[3];       /* vel_   {x, y, angle}         robot_model.h:31 */
  double linear_vel;  /* DiffDriveModel::linear_vel_  diff_drive_model.h:30 */
  double angular_vel; /* DiffDriveModel::angular_vel_ */
  float cmd[3];       /* last_cmd_ / target_*_vel_ */
  int32_t cmd_age;    /* ticks since the last accepted command */
} orc_robot_state;

void orc_robot_init(orc_robot_state* s, const float pose_xyt[3]);
/* DriveCallback of the robot's model: validation/clamping + t_last_cmd_ reset. */
void orc_robot_command(orc_robot_state* s, const umrs_robot_params* p, const float cmd[3]);
/* Step(dt) of the robot's model; n1, n2 = the two angular_error_ draws
 * (ackermann_model.cpp:87-89), ignored by the other models. */
void orc_robot_step(orc_robot_state* s, const umrs_robot_params* p, float dt,
                    float n1, float n2);

/* n robots of one kind, one tick each (command then Step), and the fan each scan
 * then uses (simulator.cpp:398-411): fans_out[n][4] = {loc_x, loc_y, angle_min +
 * theta, angle_max + theta}.  Noise draws are zero.  Batch form of the calls
 * above for the CPU baseline (no Python in the timed loop). */
void orc_robots_tick(orc_robot_state* s, int32_t n, const umrs_robot_params* p,
                     const float* cmds_n3, float dt, float* fans_out_n4);

typedef struct orc_human_state {
  float pose[3], start[3], goal[3];
  float trans_vel[2];
  double rot_vel;
} orc_human_state;

void orc_human_init(orc_human_state* s, const umrs_human_params* p);
void orc_human_step(orc_human_state* s, const umrs_human_params* p, float dt);
/* Template + pose lines: human_object.cpp:95-112,131-139. n = num_segments. */
void orc_human_template(const umrs_human_params* p, float* tmpl_xyxy);
void orc_transform_lines(const float* tmpl_xyxy, int32_t n, const float pose_xyt[3],
                         float* out_xyxy);
/* ShortTermObject square, short_term_object.cpp:53-58 (double r, eps). */
void orc_square_template(float* tmpl_xyxy /* 4 segs */);
/* Robot body rectangle (NEW, no reference behaviour):
 * car_length x car_width centred at -rear_axle_offset along the heading,
 * corners gapped 1 mm like the square above. */
void orc_robot_template(const umrs_robot_params* p, float* tmpl_xyxy /* 4 segs */);

/* ---- Philox4x32-10 + the normal transform used by the product ----------------- */
void orc_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
/* Box-Muller on two 32-bit words -> two N(0,1) (same formula as the kernel). */
void orc_box_muller(uint32_t u0, uint32_t u1, float out[2]);

#ifdef __cplusplus
}
#endif
#endif
