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

/* ONE substep `substep` of World.step (the body of the loop core.py:1974-2009; drag only when substep == 0,
 * core.py:2868, 2893) from the given state, with what the reference's loop holds in between:
 *  masks        [S][E][E] out (row `substep` written), may be NULL
 *  tot_force    [E][N][2] out: forces_dict as _integrate_state reads it; tot_torque [E][N]: torques_dict
 *  joint_force  [J][2][N][2] out: force on entity_a / entity_b of every joint as handed to update_env_forces
 *                                 (core.py:2282-2291); joint_torque [J][2][N].  Any of the four may be NULL.
 *  trig         [E][N][4] in, may be NULL: cos(rot) sin(rot) cos(rot + pi/2) sin(rot + pi/2) of every entity at the
 *               substep's entry, used instead of libm's cosf / sinf - a test passes what torch.cos / torch.sin return for
 *               the same fp32 angles (MKL VML on the CPU; third-party, closed source, not restatable);
 *  jexp         [J][N] in, may be NULL: torch.exp(|rot_a - (rot_b + fixed_rotation)|) of core.py:2852, likewise.
 *  vec_width    0, or W = 2 x Vec::size() of the ATen build that made a golden fixture (32 with AVX-512, 16 with AVX2):
 *               ATen's logaddexp (core.py:2820) runs whole groups of W elements of its [N, P] operand through SLEEF and
 *               the tail through libm; with W the oracle takes the same function for the same element. */
int vmas_oracle_substep(const vmasb200_world_desc* desc, int64_t n, float* state, float* force, float* torque,
                        const float* jrot, uint8_t* masks, int substep, float* tot_force, float* tot_torque,
                        float* joint_force, float* joint_torque, const float* trig, const float* jexp, int vec_width,
                        int n_threads);

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
