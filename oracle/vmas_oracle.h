/*
 * vmas_oracle.h - TEST INFRASTRUCTURE: scalar CPU restatement of the reference's batched physics step.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * It is the checker, never the product: nothing under vectorizedmultiagentsimulator_b200/ links or calls it.
 *
 * Parity pin: the reference holds no golden vectors for this path; this restatement is pinned
 * against outputs of the reference itself (torch 2.11.0 CPU fp32, oracle/_ref) stored in tests/golden/*.npz
 * by tests/golden/make_golden.py (tests/test_oracle_golden.py).
 *
 * State layout here is the fixture layout, NOT the product layout: state[E][N][6] = px py vx vy rot ang_vel.
 */
#ifndef VMAS_ORACLE_H_
#define VMAS_ORACLE_H_
#include <stdint.h>
#include "../include/vmasb200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One World.step (core.py:1971-2014).
 *  state  [E][N][6] in/out
 *  force  [A][N][2] in/out (agent.state.force; written back clamped, core.py:2017-2027), A = #agents in entity order
 *  torque [A][N]    in/out
 *  jrot   [J][N]    per-env fixed rotation of every joint (used when fixed_rotation_per_env), may be NULL
 *  masks  [S][E][E] out, World.collides verdicts (core.py:2787-2802) per substep, may be NULL
 *  n_threads: OpenMP threads over envs (<=0: all)
 */
int vmas_oracle_step(const vmasb200_world_desc* desc, int64_t n, float* state, float* force, float* torque,
                     const float* jrot, uint8_t* masks, int n_threads);

/* Lidar.measure(vectorized=True) (sensors.py:100-122 -> core.py:1661-1785) for desc->lidars[lidar_index];
 * out [N][n_rays]. */
int vmas_oracle_lidar(const vmasb200_world_desc* desc, int64_t n, const float* state, int lidar_index, float* out,
                      int n_threads);

/* World.cast_rays with explicit world-frame angles [N][n_rays] and explicit target list. */
int vmas_oracle_cast_rays(const vmasb200_world_desc* desc, int64_t n, const float* state, int entity,
                          const float* angles, int n_rays, float max_range, const int32_t* targets, int n_targets,
                          float* out, int n_threads);

/* World.get_distance (core.py:1821-1907), World.is_overlapping (core.py:1909-1966) and
 * World.get_distance_from_point (core.py:1787-1819).  dims = {radius, length, width} as the shape's Python floats;
 * state_x = the entity's [N][6] block of the fixture state; points [N][2]; out [N]. */
int vmas_oracle_get_distance(int shape_a, const double* dims_a, const float* state_a, int shape_b, const double* dims_b,
                             const float* state_b, int64_t n, float* out);
int vmas_oracle_is_overlapping(int shape_a, const double* dims_a, const float* state_a, int shape_b,
                               const double* dims_b, const float* state_b, int64_t n, uint8_t* out);
int vmas_oracle_distance_from_point(int shape_a, const double* dims_a, const float* state_a, const float* points,
                                    int64_t n, float* out);

#ifdef __cplusplus
}
#endif
#endif
