/*
 * vmasb200.h - C ABI of libvmasb200.so: the B200-native (sm_100a) batched 2-D physics step of VMAS.
 *
 * This is the drop-in boundary for ONE hot path of the reference
 * (robj0nes/VectorizedMultiAgentSimulator, all citations relative to its tree):
 *
 *   vmas/simulator/core.py:1971  World.step()        -> vmasb200_step / vmasb200_step_host
 *   vmas/simulator/core.py:1661  World.cast_rays()   -> vmasb200_cast_rays (and the LIDAR epilogue of
 *                                                        vmasb200_step, called from sensors.py:100 Lidar.measure)
 *
 * The reference has no FFI layer (it is pure Python/PyTorch); the binding a maintainer would add is the
 * ctypes stub shown in INTEGRATION.md, and the host-side mirror that monkey-patches those two bound methods
 * lives in vectorizedmultiagentsimulator_b200/dropin.py.
 *
 * Conventions
 *   - plain C types only; every pointer is either HOST ("h:") or DEVICE ("d:") memory as annotated;
 *   - every entry point returns 0 on success or a negative vmasb200_status; a thread-local, human readable
 *     message for the last failure is returned by vmasb200_last_error();
 *   - all device work is enqueued on the cudaStream_t passed as `stream` (NULL = legacy default stream);
 *     no entry point synchronises the host unless its comment says so;
 *   - a plan owns small device scratch (flag words) and must be used by one stream at a time;
 *   - there is NO CPU fallback anywhere in this library: without a CUDA device every compute entry point
 *     fails with VMASB200_ERR_CUDA.
 *
 * Data layout in HBM: structure of arrays, fp32, one ROW per scalar field and `ld` floats between rows
 * (ld >= n_envs, ld % 4 == 0, every buffer 16-byte aligned, so that any tile of a row is a legal cp.async.bulk /
 * TMA 1-D source or destination and a warp's 32 lanes touch 32 consecutive environments):
 *   dyn   [dyn_rows][ld]    6 rows per dynamic entity (movable or rotatable), in entity order:
 *                            px py vx vy rot ang_vel                           (core.py:205-315 EntityState)
 *   stat  [stat_rows][ld]   3 rows (px py rot) per static entity that can collide / is jointed / is a LIDAR
 *                            target; then one row per joint whose fixed_rotation is per-env (joints.py:139-143)
 *   actout[act_rows][ld]    3 rows (clamped fx fy torque) per agent that has max_f/f_range/max_t/t_range
 *                            (the write-back of core.py:2017-2040)
 *   lidar [lidar_rows][ld]  one row per ray, sensors concatenated in desc order (sensors.py:100-122)
 * vmasb200_plan_layout() / _slots() report all row numbers.  A torch view of one field of all environments is
 * `buf[row]` (contiguous [n_envs]); of an entity's position `buf[row:row+2].T` ([n_envs, 2], strides (1, ld)).
 */
#ifndef VMASB200_H_
#define VMASB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VMASB200_ABI_VERSION 3

typedef enum vmasb200_status {
  VMASB200_OK = 0,
  VMASB200_ERR_INVALID = -1,      /* bad argument / inconsistent description            */
  VMASB200_ERR_UNSUPPORTED = -2,  /* description uses something the kernel does not do  */
  VMASB200_ERR_CUDA = -3,         /* a CUDA runtime call failed (no device, launch...)  */
  VMASB200_ERR_NOMEM = -4
} vmasb200_status;

/* shape classes: core.py:102 Box, :140 Sphere, :171 Line */
enum { VMASB200_SPHERE = 0, VMASB200_BOX = 1, VMASB200_LINE = 2 };

/* entity flag bits (which optional attributes of core.py:537 Entity / :829 Agent are set) */
enum {
  VMASB200_E_MOVABLE = 1u << 0,
  VMASB200_E_ROTATABLE = 1u << 1,
  VMASB200_E_AGENT = 1u << 2,
  VMASB200_E_HOLLOW = 1u << 3,        /* Box(hollow=True), core.py:103                      */
  VMASB200_E_MAX_SPEED = 1u << 4,
  VMASB200_E_V_RANGE = 1u << 5,
  VMASB200_E_DRAG = 1u << 6,          /* entity.drag overrides world drag, core.py:2865     */
  VMASB200_E_LIN_FRICTION = 1u << 7,  /* entity.linear_friction is not None, core.py:2075   */
  VMASB200_E_ANG_FRICTION = 1u << 8,
  VMASB200_E_GRAVITY = 1u << 9,       /* entity.gravity is not None, core.py:2048           */
  VMASB200_E_MAX_F = 1u << 10,
  VMASB200_E_F_RANGE = 1u << 11,
  VMASB200_E_MAX_T = 1u << 12,
  VMASB200_E_T_RANGE = 1u << 13,
  VMASB200_E_COLLIDE = 1u << 14       /* entity.collide, core.py:620                        */
};

/* One row per entry of World.entities (landmarks first, then agents: core.py:1220). */
typedef struct vmasb200_entity {
  int32_t shape;              /* VMASB200_SPHERE / BOX / LINE                                         */
  uint32_t flags;             /* VMASB200_E_*                                                         */
  float radius;               /* Sphere.radius                                                        */
  float length;               /* Box.length / Line.length                                             */
  float width;                /* Box.width                                                            */
  float mass;
  float moment_of_inertia;    /* Shape.moment_of_inertia(mass), core.py:122,159,186 (double -> fp32)  */
  float max_speed;
  float v_range;
  double drag;                /* entity.drag as the Python float it is: the reference forms 1 - drag in double before
                                 the product with an fp32 tensor (core.py:2866-2869)                      */
  float linear_friction;
  float angular_friction;
  float gravity[2];
  float max_f, f_range, max_t, t_range;
  double circumscribed_radius;/* core.py:125,162,189; kept in double: the broad phase adds two of them
                                 in double before the fp32 compare (core.py:2796-2800)                */
} vmasb200_entity;

/* One row per JointConstraint (joints.py:147), in the order World._apply_vectorized_enviornment_force
 * meets them (pair enumeration a<b, core.py:2111-2121). */
typedef struct vmasb200_joint {
  int32_t entity_a, entity_b;    /* indices into entities[] of constraint.entity_a / entity_b          */
  float delta_a[2], delta_b[2];  /* shape.get_delta_from_anchor(anchor), joints.py:185-201             */
  float dist;                    /* constraint.dist                                                     */
  int32_t rotate;                /* constraint.rotate                                                   */
  float fixed_rotation;          /* scalar fixed rotation (ignored when fixed_rotation_per_env != 0)   */
  int32_t fixed_rotation_per_env;/* 1: value is read from the stat record (joints.py:139-143)          */
} vmasb200_joint;

/* One row per Lidar (sensors.py:46) whose measurement is produced by the step epilogue. */
typedef struct vmasb200_lidar {
  int32_t entity;          /* the sensing agent                                                       */
  int32_t n_rays;
  const float* angles;     /* h: [n_rays] sensor-frame ray angles (row 0 of Lidar._angles)            */
  float max_range;
  int32_t n_targets;
  const int32_t* targets;  /* h: [n_targets] entity indices passing entity_filter, in entity order    */
} vmasb200_lidar;

typedef struct vmasb200_world_desc {
  int32_t abi_version;     /* VMASB200_ABI_VERSION */
  int32_t n_entities;
  const vmasb200_entity* entities;  /* h: [n_entities]                                               */
  /* h: [n_entities*n_entities] row-major; collide_mask[a*E+b] (a<b) != 0 iff the pair passes the STATIC
   * part of World.collides (collide flags, both collision_filters, not both static, core.py:2787-2795)
   * and is not excluded by a dist-0 joint (core.py:2118-2121).  The batch-wide distance test of
   * core.py:2796-2800 is evaluated on the device every substep. */
  const uint8_t* collide_mask;
  int32_t n_joints;
  const vmasb200_joint* joints;     /* h: [n_joints]                                                 */
  int32_t n_lidars;
  const vmasb200_lidar* lidars;     /* h: [n_lidars]                                                 */
  /* World constants, core.py:1090-1143 */
  double dt;               /* Python floats: sub_dt = dt / substeps (core.py:1124) and 1 - drag (core.py:2866) are  */
  int32_t substeps;        /* formed in double and only then meet an fp32 tensor                                    */
  double drag;
  float linear_friction, angular_friction;
  int32_t has_x_semidim, has_y_semidim;
  float x_semidim, y_semidim;
  float collision_force, joint_force, torque_constraint_force, contact_margin;
  float gravity[2];
  /* Kernel shape override: environments per CTA, 32, 64 or 128 (a CTA has 128 threads: 4, 2 or 1 of them share an
   * environment's work); 0 = chosen per step from the batch size (vmasb200_plan_tile). */
  int32_t envs_per_cta;
} vmasb200_world_desc;

enum { VMASB200_PAIR_SS = 0, VMASB200_PAIR_LS, VMASB200_PAIR_LL, VMASB200_PAIR_BS, VMASB200_PAIR_BL, VMASB200_PAIR_BB };

typedef struct vmasb200_pair {
  int32_t type;   /* VMASB200_PAIR_*                                                                   */
  int32_t a, b;   /* entity indices in the reference's per-type order: (line,sphere) (box,sphere) (box,line) */
} vmasb200_pair;

typedef struct vmasb200_layout {
  int32_t dyn_rows, stat_rows, act_rows, lidar_rows; /* rows of the four SoA buffers                     */
  int32_t n_pairs;        /* candidate pairs, in flag-bit order (ss, ls, ll, bs, bl, bb; a<b inside)     */
  int32_t pair_words;     /* ceil(n_pairs/32): words per substep in pair_active_out                      */
  int32_t envs_per_cta;   /* the description's override, or 32 (smallest tile) when 0; see vmasb200_plan_tile */
  int32_t n_dyn, n_stat, n_agents, n_act_out;
  int64_t alg_bytes_per_env; /* algorithmic HBM bytes per env-step (SURVEY 8d formula)                    */
} vmasb200_layout;

typedef struct vmasb200_plan vmasb200_plan;

int vmasb200_version(void);
const char* vmasb200_last_error(void);

/* Builds the static plan (entity / pair / joint / sensor tables, record layout) and uploads it.
 * Host only reads `desc` during the call.  Replaces the per-substep Python pair enumeration of core.py:2103-2173. */
int vmasb200_plan_create(const vmasb200_world_desc* desc, vmasb200_plan** out);
int vmasb200_plan_destroy(vmasb200_plan* plan);
int vmasb200_plan_layout(const vmasb200_plan* plan, vmasb200_layout* out);
/* Per-entity rows: dyn_slot[e] / stat_slot[e] = first row of the entity inside the dyn / stat buffer, or -1;
 * act_slot[e] = first row of the clamped (fx fy torque) triple in the actout buffer, or -1.
 * h: each array [n_entities] (any may be NULL). */
int vmasb200_plan_slots(const vmasb200_plan* plan, int32_t* dyn_slot, int32_t* stat_slot, int32_t* act_slot);
/* h: pairs[n_pairs] */
int vmasb200_plan_pairs(const vmasb200_plan* plan, vmasb200_pair* pairs);
/* row inside the stat buffer of the per-env fixed rotation of joint j, or -1. h: [n_joints] */
int vmasb200_plan_joint_slots(const vmasb200_plan* plan, int32_t* joint_slot);
/* first row inside the lidar buffer of sensor i. h: [n_lidars] */
int vmasb200_plan_lidar_slots(const vmasb200_plan* plan, int32_t* lidar_slot);
/* entity index of every cos/sin cache slot (the row order of vmasb200_step_io.trig_in); returns their number.
 * h: trig_entities [n_entities] or NULL. */
int vmasb200_plan_trig_entities(const vmasb200_plan* plan, int32_t* trig_entities);
/* environments per CTA a step over n_envs environments runs with (32, 64 or 128): the description's envs_per_cta,
 * or, when that is 0, chosen from the batch size.  128 / tile threads share each environment's work. */
int vmasb200_plan_tile(const vmasb200_plan* plan, int64_t n_envs);

/* World-compiled step kernels.  For small worlds (simple_spread, navigation, transport, most MPE scenarios) the library prints the plan as C++ with the world's constants, entity attributes
 * and pair list as literals, compiles it with NVRTC for sm_100a (same parity flags as the library build) and launches
 * that instead of the generic table-walking kernel: an environment's state then stays in registers for all substeps.
 * Results are bit-identical to the generic kernel.  Cubins are cached in `_jit/` beside the library (or
 * $VMASB200_JIT_CACHE).  Without libnvrtc / libcuda, for other worlds, with VMASB200_JIT=0 in the environment or after
 * vmasb200_jit_enable(0) the generic kernel runs.
 *  vmasb200_jit_enable: process-wide switch, returns the previous setting.  0 off; 1 (default) worlds the compiled
 *  kernel is measured to win on: all pairs sphere-sphere, or at most 32 candidate pairs and 2 joints; 2 (also
 *  VMASB200_JIT=2) every world the generator covers, whatever its size (tests).
 *  vmasb200_plan_jit: compiles (no GPU needed) the kernel a step over n_envs would take.  *state: 1 compiled / cached,
 *  0 the world is not covered or the switch is off, -1 compilation failed (vmasb200_last_error says why).
 *  *launches: launches of this plan that took a compiled kernel so far; *tail_launches: launches of the 1-CTA
 *  verification kernel the library puts behind each of them when the world can have thin contacts (any pair that is
 *  not sphere-sphere).  Every out pointer may be NULL. */
int vmasb200_jit_enable(int on);
int vmasb200_plan_jit(vmasb200_plan* plan, int64_t n_envs, int32_t* state, int64_t* launches, int64_t* tail_launches);

typedef struct vmasb200_step_io {
  const float* dyn_in;             /* d: [dyn_rows][ld]                                                 */
  float* dyn_out;                  /* d: [dyn_rows][ld]; may alias dyn_in                               */
  const float* stat;               /* d: [stat_rows][ld] (NULL iff stat_rows == 0)                      */
  /* agent.state.force / agent.state.torque of every agent, in entity order of the agents (core.py:2017-2040).
   * h: arrays of n_agents device pointers + row strides in floats; a NULL pointer reads as zeros.        */
  const float* const* force_ptrs;
  const int64_t* force_strides;
  const float* const* torque_ptrs;
  const int64_t* torque_strides;
  float* act_out;                  /* d: [act_rows][ld] (NULL iff act_rows == 0)                        */
  float* lidar_out;                /* d: [lidar_rows][ld]; NULL skips the LIDAR epilogue                */
  uint32_t* pair_active_out;       /* d: [substeps][pair_words] batch-wide broad-phase results of
                                      core.py:2796-2800 per (substep, pair), or NULL                     */
  int64_t n_envs;
  int64_t ld;                      /* floats between rows of every SoA buffer: >= n_envs, multiple of 4 */
  /* 0: one fused speculative launch + a device-side verified exact fix-up (no host sync);
   * 1: always take the exact per-substep path (flag reduction before every substep).                    */
  int32_t exact;
  void* stream;                    /* cudaStream_t                                                       */
  /* ---- probe fields (parity tests, tests/test_gpu_substeps.py); all zero / NULL in normal use; they need exact != 0.
   * substep_begin / substep_end: run only substeps [begin, end) of the World.step loop (core.py:1974; drag is applied
   *   when substep 0 is among them, core.py:2868); both 0 = all substeps.  The clamp of core.py:2017-2040 is replayed
   *   for the substeps before `begin` unless force_ptrs already hold the write-back of substep begin - 1
   *   (forces_are_substep_inputs != 0).
   * trig_in  d: [4 * n_trig][ld] cos(rot) sin(rot) cos(rot + pi/2) sin(rot + pi/2) of every entity with a cos/sin
   *   cache, in plan order (vmasb200_plan_trig_entities): used INSTEAD of computing them at the start of the launch
   *   (a test hands in the reference's own torch.cos / torch.sin values: MKL's on the CPU), or NULL.
   * jexp_in  d: [n_joints][ld] exp(|rot_a - (rot_b + fixed_rotation)|) of every joint (core.py:2848-2852: torch.exp, MKL's on
 *   the CPU; exp(a) - 1 cancels for small a), used instead of expf for the joints with rotate == 0; needs a one-substep
 *   range; or NULL.
 * force_out d: [3 * n_dyn][ld] forces_dict / torques_dict (core.py:1975-1992) of every dynamic entity (Fx Fy T, plan
   *   order of the dynamic entities) as _integrate_state of the LAST substep run read them, or NULL.
   * joint_out d: [4 * n_joints][ld] per joint (fx fy on entity_a; torque on a; torque on b) of the last substep run -
   *   what core.py:2282-2291 hands to update_env_forces (the force on b is the negation) -, or NULL. */
  int32_t substep_begin, substep_end;
  int32_t forces_are_substep_inputs;
  const float* trig_in;
  float* force_out;
  float* joint_out;
  const float* jexp_in;
} vmasb200_step_io;

/* One World.step(): all substeps of {action/friction/gravity forces, joints, six collision pair types,
 * semi-implicit Euler integration} plus the LIDAR epilogue, for every env.  Asynchronous. */
int vmasb200_step(vmasb200_plan* plan, const vmasb200_step_io* io);

/* Same step with HOST buffers (same SoA layout, row stride ld): H2D of inputs, kernel(s), D2H of results, then a
 * sync.  The batch is cut into chunks of environments that are pipelined over several streams so that the H2D of
 * one chunk, the kernel of another and the D2H of a third overlap (pinned host buffers make the copies truly
 * asynchronous; pageable ones still work).  h_force [n_envs][n_agents][2], h_torque [n_envs][n_agents] (either may
 * be NULL = zeros).  Output pointers may be NULL to skip that copy.  h_dyn_in / h_stat may be NULL from the second call on:
 * the step then starts from the state (static rows) the plan's device staging holds from the previous call of the same
 * batch size - a rollout through this entry point only moves the actions in and the results out.  Pinned (page-locked,
 * mapped) h_force / h_torque of up to 131 072 envs are not copied: the kernel loads them over PCIe itself (they must stay
 * unchanged until the call returns, which it does only after the step has completed).  h_pair_active / exact != 0 need the batch-wide
 * verdicts and therefore run the batch as ONE chunk.  Device staging is owned by the plan and grows on demand. */
int vmasb200_step_host(vmasb200_plan* plan, const float* h_dyn_in, const float* h_stat, const float* h_force,
                       const float* h_torque, float* h_dyn_out, float* h_act_out, float* h_lidar_out,
                       uint32_t* h_pair_active, int64_t n_envs, int64_t ld, int32_t exact);

/* World.cast_rays(entity, angles, max_range, entity_filter) (core.py:1661-1785) on the current SoA buffers.
 * angles d: [n_envs][n_rays] WORLD-frame ray angles with row stride `angles_stride` floats (the torch tensor
 * Lidar.measure builds, sensors.py:115); targets h: [n_targets] entity indices (already filtered);
 * out d: [n_envs][n_rays] contiguous (the tensor World.cast_rays returns). */
int vmasb200_cast_rays(vmasb200_plan* plan, const float* dyn, const float* stat, int64_t ld, int32_t entity,
                       const float* angles, int64_t angles_stride, int32_t n_rays, float max_range,
                       const int32_t* targets, int32_t n_targets, float* out, int64_t n_envs, void* stream);

/* Synchronises `stream` and reports any deferred CUDA error; also returns, in *n_fixups (may be NULL), how many
 * steps of this plan needed the exact fix-up path since creation. */
int vmasb200_sync_check(vmasb200_plan* plan, void* stream, int64_t* n_fixups);

/* ---- distance queries of reward code: World.get_distance_from_point (core.py:1787-1819),
 * World.get_distance (core.py:1821-1907) and World.is_overlapping (core.py:1909-1966).  They take an entity's shape
 * and pointers to its pose tensors as they are (any strides, device memory), so they work for every entity of a World,
 * staged in a plan's SoA buffers (then pos is a strided view into them) or not; no plan is needed. */
typedef struct vmasb200_pose {
  int32_t shape;                /* VMASB200_SPHERE / BOX / LINE                                          */
  double radius, length, width; /* the shape's Python floats; the library rounds them the way torch does   */
  const float* pos;             /* d: x of env i at pos[i * pos_stride], y at pos[i * pos_stride + pos_stride_xy] */
  int64_t pos_stride, pos_stride_xy;
  const float* rot;             /* d: rot of env i at rot[i * rot_stride]; may be NULL for a sphere         */
  int64_t rot_stride;
} vmasb200_pose;

/* out d: [n_envs] float, World.get_distance(a, b).  Asynchronous on `stream`. */
int vmasb200_get_distance(const vmasb200_pose* a, const vmasb200_pose* b, float* out, int64_t n_envs, void* stream);
/* out d: [n_envs] uint8 (0 / 1), World.is_overlapping(a, b). */
int vmasb200_is_overlapping(const vmasb200_pose* a, const vmasb200_pose* b, uint8_t* out, int64_t n_envs, void* stream);
/* points d: test point of env i at points[i * points_stride] (x) and [.. + points_stride_xy] (y);
 * out d: [n_envs] float, World.get_distance_from_point(a, points). */
int vmasb200_get_distance_from_point(const vmasb200_pose* a, const float* points, int64_t points_stride,
                                     int64_t points_stride_xy, float* out, int64_t n_envs, void* stream);

/* ---- action pre-processing of Environment._set_action for continuous actions without communication / noise
 * (environment.py:615-668): u = clamp(action, -u_range, +u_range) (when clamp != 0) * u_multiplier, one
 * launch per agent instead of ~10 and no host synchronisation.
 *  action d: [n_envs][action_size] fp32, element (i, k) at action[i * stride_env + k * stride_k]
 *  u_range, u_multiplier h: [action_size] (agent.action.u_range_tensor / u_multiplier_tensor as fp32), action_size <= 8
 *  u_out d: [n_envs][action_size] contiguous (the new agent.action.u)
 *  flags d: one int32 the launch ORs into: bit 0 = a NaN action (environment.py:622), bit 1 = |action| > u_range after
 *  the optional clamp (environment.py:647-649).  The caller reads it when it wants the reference's assertions. */
int vmasb200_set_action(const float* action, int64_t stride_env, int64_t stride_k, int32_t action_size,
                        const float* u_range, const float* u_multiplier, int32_t clamp, float* u_out, int32_t* flags,
                        int64_t n_envs, void* stream);

/* ---- the same with COMMUNICATION actions and / or action NOISE (environment.py:633-645, 711-754), continuous actions:
 *  action d: [n_envs][>= action_size + n_comm] fp32; physical columns 0 .. action_size - 1, communication columns
 *            comm_col0 .. comm_col0 + n_comm - 1 (comm_col0 = action_size when env.clamp_action, else world.dim_p: the
 *            reference slices `action[:, dim_p:]` there, environment.py:742-743)
 *  u_randn d: [n_envs][action_size] a torch.randn draw or NULL: u += u_randn * u_noise (environment.py:711-720; the
 *            caller draws it with the reference's own call, so both consume the generator alike)
 *  c_randn d: [n_envs][n_comm] or NULL: c += c_randn * c_noise (environment.py:745-754)
 *  c_out  d: [n_envs][n_comm] the new agent.action.c (never clamped: with clamp_action the reference clamps a copy it then
 *            does not use, environment.py:640-645 vs 737-743)
 *  flags: bit 0 NaN, bit 1 |u| > u_range, bit 3 a communication action outside [0, 1] (environment.py:738-740) */
typedef struct vmasb200_action {
  const float* action;
  int64_t stride_env, stride_k;
  int32_t action_size, comm_col0, n_comm, clamp;
  const float* u_range;       /* h: [action_size] */
  const float* u_multiplier;  /* h: [action_size] */
  const float* u_randn;
  const float* c_randn;
  double u_noise, c_noise;    /* agent.u_noise / agent.c_noise (Python floats) */
  float* u_out;
  float* c_out;
  int32_t* flags;
} vmasb200_action;
int vmasb200_set_action_ex(const vmasb200_action* a, int64_t n_envs, void* stream);

/* ---- the same for DISCRETE physical actions without communication / noise (environment.py:660-710):
 * a flat action (the index into the cartesian product of the agent's action spaces, one column) or a multi-discrete
 * one (one column per component) becomes u[k] = (a_k / (nvec[k] - 1)) * 2 u_range[k] - u_range[k], with the
 * reference's re-ordering for odd nvec[k] (action 0 maps to u = 0), times u_multiplier[k].
 *  action d: [n_envs][1 or action_size] integers of `action_dtype` (VMASB200_I64 / VMASB200_I32), element (i, k) at
 *            action[i * stride_env + k * stride_k] (strides in elements)
 *  nvec h: [action_size] agent.discrete_action_nvec (every entry > 1), action_size <= 8
 *  flags: bit 2 = a component outside [0, nvec[k]) (environment.py:608-612, "Discrete physical actions are out of bounds") */
enum { VMASB200_I64 = 0, VMASB200_I32 = 1 };
int vmasb200_set_action_discrete(const void* action, int32_t action_dtype, int64_t stride_env, int64_t stride_k,
                                 int32_t action_size, int32_t multidiscrete, const int32_t* nvec, const float* u_range,
                                 const float* u_multiplier, float* u_out, int32_t* flags, int64_t n_envs, void* stream);

/* ---- reset path
 * vmasb200_reset_rows: World.reset(env_index) (core.py:1178-1180 -> EntityState._reset, core.py:285-295) IN PLACE on a
 * plan's SoA buffer: rows [row0, row0 + n_rows) of buf d: [rows][ld] are zeroed for env `env_index`, for every env
 * (env_index < 0, mask NULL) or for the envs whose byte of mask d: [n_envs] is non-zero (env_index < 0).  The state
 * tensors of the reference stay the views they were: no re-pack at the next step.
 * vmasb200_spawn_select: one iteration of the rejection loop of ScenarioUtils.find_random_pos_for_entity
 * (utils.py:285-319): overlaps = (torch.cdist(occupied [B][K][2], pos [B][1][2]) < min_dist).any(K);
 * pos[overlaps] = proposed[overlaps]; *flag |= overlaps.any() (the loop's exit test, read by the caller: the reference
 * synchronises on it too).  The proposals are the caller's torch `uniform_` draws, so placement is bit-identical. */
int vmasb200_reset_rows(float* buf, int32_t row0, int32_t n_rows, int64_t ld, int64_t n_envs, int64_t env_index,
                        const uint8_t* mask, void* stream);
int vmasb200_spawn_select(float* pos, const float* proposed, const float* occupied, int32_t n_occupied, float min_dist,
                          int32_t* flag, int64_t batch, void* stream);

/* ---- scenario hooks: what Environment._get_from_scenario (environment.py:253-308) asks of a scenario
 * after every World.step - observation / reward / done / info - as one or two launches per step instead of tens to
 * hundreds of small torch ops with host synchronisations (World.collides, boolean-mask indexing).  All tensors are read
 * and written in place through (pointer, strides); nothing is allocated and nothing synchronises. */
typedef struct vmasb200_vec {     /* a [n_envs][k] device tensor: element (i, k) at p[i * stride_env + k * stride_k] */
  const void* p;
  int64_t stride_env, stride_k;
} vmasb200_vec;

enum {
  VMASB200_OBS_COPY = 0, /* a            (fp32)                                              */
  VMASB200_OBS_SUB = 1,  /* a - b        (fp32; e.g. agent.state.pos - goal.state.pos)         */
  VMASB200_OBS_CSUB = 2, /* c - a        (fp32; e.g. sensor._max_range - sensor.measure())     */
  VMASB200_OBS_BOOL = 3, /* a != 0 ? 1 : 0 (a is a torch.bool / uint8 tensor; cat promotes it) */
  VMASB200_OBS_COS = 4,  /* cosf(a)      (joint_passage.py:41-42 angle_to_vector)                  */
  VMASB200_OBS_SIN = 5   /* sinf(a)                                                              */
};
/* Flags of vmasb200_obs_seg.flags.  NEG0: column 0 of BOTH operands is negated before the op (football.py:1376-1397:
 * a red agent sees every x mirrored, `tensor[..., X] = -tensor[..., X]` on clones of the operands). */
#define VMASB200_OBS_NEG0 1
/* One block of columns of one observation tensor: out[i][col0 + k] = op(a[i][k], b[i][k]), k < width
 * (what one operand of the scenario's torch.cat([...], dim=-1) contributes). */
typedef struct vmasb200_obs_seg {
  vmasb200_vec a, b;
  int32_t op;
  float c;
  int32_t width;
  int32_t out;    /* index into outs[] */
  int32_t col0;
  int32_t flags;  /* VMASB200_OBS_NEG0 */
  vmasb200_vec noise; /* p NULL, or a [n_envs][width] fp32 tensor ADDED to the op's result (joint_passage.py:531-541:
                       * `obs + noise`, the noise drawn by the caller from torch's generator) */
} vmasb200_obs_seg;
#define VMASB200_OBS_MAX_SEGS 128
#define VMASB200_OBS_MAX_OUTS 32
#define VMASB200_OBS_MAX_WIDTH 352
/* The observations of ALL agents in one launch (SoA -> [n_envs][width] transposition through shared memory).
 * segs h: [n_segs]; outs h: [n_outs] device pointers, outs[o] d: [n_envs][out_widths[o]] contiguous fp32.
 * Every column of every out must be covered by exactly one segment (checked). */
int vmasb200_gather_obs(const vmasb200_obs_seg* segs, int32_t n_segs, float* const* outs, const int32_t* out_widths,
                        int32_t n_outs, int64_t n_envs, void* stream);

/* navigation.py:199-232 (reward, agent_reward), :262-273 (done), :275-280 (info): everything `Scenario.reward` computes
 * when called for the first agent, for all agents, plus the per-agent return values and the done flags. */
#define VMASB200_NAV_MAX_AGENTS 16
typedef struct vmasb200_nav_agent {
  vmasb200_vec pos, goal_pos;   /* agent.state.pos, agent.goal.state.pos: [n_envs][2]                              */
  double radius, goal_radius;   /* agent.shape.radius (collisions, done), agent.goal.shape.radius (on_goal)         */
  const float* pos_shaping;     /* d in : [n_envs] agent.pos_shaping of the previous step                          */
  float* pos_shaping_out;       /* d out: [n_envs] the new agent.pos_shaping (may alias pos_shaping)               */
  float* distance_to_goal;      /* d out: [n_envs]                                                                 */
  uint8_t* on_goal;             /* d out: [n_envs] 0 / 1 (torch.bool storage)                                      */
  float* pos_rew;               /* d out: [n_envs] agent.pos_rew                                                   */
  float* agent_collision_rew;   /* d out: [n_envs] (the reference zeroes it and adds the penalties, :209, 222-228)  */
  float* reward;                /* d out: [n_envs] what Scenario.reward(agent) returns (:230-231)                   */
} vmasb200_nav_agent;
typedef struct vmasb200_navigation {
  int32_t n_agents;
  int32_t shared_rew;
  double pos_shaping_factor, final_reward, agent_collision_penalty, min_collision_distance;
  vmasb200_nav_agent agent[VMASB200_NAV_MAX_AGENTS];
  /* collidable[i][j], j < i: the static part of World.collides(agent_i, agent_j) (core.py:2788-2795); the batch-wide
   * distance test (core.py:2796-2800) is evaluated by the call itself on `pair_flags`. */
  uint8_t collidable[VMASB200_NAV_MAX_AGENTS][VMASB200_NAV_MAX_AGENTS];
  float* pos_rew;               /* d out: [n_envs] Scenario.pos_rew (sum over agents in order, :205-208)            */
  float* final_rew;             /* d out: [n_envs] Scenario.final_rew                                               */
  uint8_t* all_goal_reached;    /* d out: [n_envs]                                                                  */
  uint8_t* done;                /* d out: [n_envs] Scenario.done() or NULL                                           */
  uint32_t* pair_flags;         /* d scratch: 4 words, zeroed and filled by the call                                 */
} vmasb200_navigation;
int vmasb200_navigation_reward(const vmasb200_navigation* nv, int64_t n_envs, void* stream);

/* Allocation-free variant of vmasb200_navigation_reward.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid allocation).  Same inputs, outputs, launches and results as vmasb200_navigation_reward,
 * but `nv->pair_flags` is IGNORED (may be NULL): the scratch of the flag reduction lives in a file-scope
 * `static __device__ __align__(256)` array of VMASB200_NAV_STATIC_WORDS words (256 bytes, held by every GPU that loads
 * the library until the process exits).  Limit: n_agents <= VMASB200_NAV_MAX_AGENTS (16 agents = 120 pairs = 4 flag
 * words); more is rejected with VMASB200_ERR_INVALID and a message.  The call makes no cudaMalloc / cudaFree and calls no
 * allocating library routine; the array's device address is looked up once per device (cudaGetSymbolAddress, with that
 * device current) on the first call there.  Every call on a device shares the one array: the words in use are zeroed on
 * `stream` at the start of each call, and two calls that overlap on the GPU (different streams, no ordering) RACE on it -
 * order them, or use vmasb200_navigation_reward with a scratch of their own. */
#define VMASB200_NAV_STATIC_WORDS 64
int vmasb200_navigation_reward_static(const vmasb200_navigation* nv, int64_t n_envs, void* stream);

/* transport.py:130-164 (reward), :186-193 (done): package.dist_to_goal, package.on_goal (World.is_overlapping with its
 * goal), package.color, the shared reward and the new package.global_shaping, for every package in one launch. */
#define VMASB200_TRANSPORT_MAX_PACKAGES 8
typedef struct vmasb200_transport_package {
  vmasb200_pose package, goal;
  const float* global_shaping;  /* d in : [n_envs]                                                                 */
  float* global_shaping_out;    /* d out: [n_envs] (may alias global_shaping)                                      */
  float* dist_to_goal;          /* d out: [n_envs]                                                                 */
  uint8_t* on_goal;             /* d out: [n_envs]                                                                 */
  float* color;                 /* d out: [n_envs][3] contiguous, or NULL                                          */
} vmasb200_transport_package;
typedef struct vmasb200_transport {
  int32_t n_packages;
  double shaping_factor;
  float color_off_goal[3], color_on_goal[3]; /* Color.RED.value, Color.GREEN.value */
  vmasb200_transport_package package[VMASB200_TRANSPORT_MAX_PACKAGES];
  float* rew;                   /* d out: [n_envs] Scenario.rew                                                    */
  uint8_t* done;                /* d out: [n_envs] Scenario.done() or NULL                                          */
} vmasb200_transport;
int vmasb200_transport_reward(const vmasb200_transport* tr, int64_t n_envs, void* stream);

/* mpe/simple_spread.py:79-110 (reward): for every agent in turn, minus the distance of the closest agent to every
 * landmark, then minus 1 per other agent overlapping it (World.is_overlapping, spheres), all in the reference's
 * order of fp32 subtractions. */
#define VMASB200_SPREAD_MAX 16
typedef struct vmasb200_simple_spread {
  int32_t n_agents, n_landmarks;
  vmasb200_vec agent_pos[VMASB200_SPREAD_MAX];     /* agent.state.pos    [n_envs][2] */
  vmasb200_vec landmark_pos[VMASB200_SPREAD_MAX];  /* landmark.state.pos [n_envs][2] */
  double agent_radius[VMASB200_SPREAD_MAX];
  uint8_t collide[VMASB200_SPREAD_MAX];            /* agent.collide                  */
  float* rew;                                      /* d out: [n_envs] Scenario.rew   */
} vmasb200_simple_spread;
int vmasb200_simple_spread_reward(const vmasb200_simple_spread* sp, int64_t n_envs, void* stream);

/* joint_passage.py:387-490 (reward for the first agent: pos / rot shaping of the joint before and after the passage,
 * collision and energy terms), :547-563 (done).  `done` couples environments in the reference: the [n] position test
 * is multiplied with the [n, 1] rotation test, which broadcasts to [n, n] before `all(dim=1)`, so
 * done[i] = rot_ok[i] AND (pos_ok of EVERY env); the call reduces `all pos_ok` over the batch first (done_flag). */
#define VMASB200_JP_MAX_BODIES 4
#define VMASB200_JP_MAX_PASSAGES 24
#define VMASB200_JP_MAX_WALLS 8
typedef struct vmasb200_joint_passage {
  int32_t n_agents;                 /* world.agents: all_passed (:401-404), energy (:476-484)                      */
  int32_t n_bodies;                 /* agents (+ the mass when asym_package): the bodies of the collision loop :454 */
  int32_t n_passages, n_walls;
  int32_t all_passed_rot;           /* :432                                                                         */
  int32_t dim_p;                    /* world.dim_p (:480)                                                           */
  vmasb200_pose body[VMASB200_JP_MAX_BODIES];      /* agents first, then the mass                                   */
  vmasb200_vec action_u[VMASB200_JP_MAX_BODIES];   /* agent.action.u [n_envs][2], agents only                       */
  double f_range[VMASB200_JP_MAX_BODIES];          /* agent.f_range                                                 */
  vmasb200_pose passage[VMASB200_JP_MAX_PASSAGES]; /* self.passages in order                                        */
  uint8_t passage_collide[VMASB200_JP_MAX_PASSAGES];
  vmasb200_pose wall[VMASB200_JP_MAX_WALLS];
  vmasb200_pose joint;              /* self.joint.landmark (a line)                                                 */
  vmasb200_vec goal_pos, goal_rot;  /* self.goal.state.pos [n][2], .rot [n][1]                                      */
  double passage_width, pos_shaping_factor, rot_shaping_factor, collision_reward, energy_reward_coeff,
      min_collision_distance, middle_angle;
  const float* shaping_in[4];       /* d in : [n_envs] joint.pos_shaping_pre, pos_shaping_post, rot_shaping_pre, rot_shaping_post */
  float* shaping_out[4];            /* d out: the new ones (may alias shaping_in)                                   */
  float* pos_rew;                   /* d out: [n_envs] Scenario.pos_rew (written in place by the reference)         */
  float* rot_rew;
  float* collision_rew;
  float* energy_expenditure;
  float* energy_rew;
  float* rew;                       /* d out: [n_envs] what Scenario.reward returns                                 */
  uint8_t* all_passed;              /* d out: [n_envs]                                                              */
  uint8_t* done;                    /* d out: [n_envs] Scenario.done(), or NULL                                      */
  uint32_t* done_flag;              /* d scratch: one word, zeroed and filled by the call (needed when done != NULL) */
} vmasb200_joint_passage;
int vmasb200_joint_passage_reward(const vmasb200_joint_passage* jp, int64_t n_envs, void* stream);

/* football.py:1118-1226 (reward for the first agent: sparse scoring reward, done, dense shaping of ball-to-goal and
 * closest-agent-to-ball for both teams). */
#define VMASB200_FB_MAX_TEAM 16
typedef struct vmasb200_football_team {
  int32_t n_agents;
  int32_t dense;                    /* the dense terms of this team are computed (:1143-1151)                       */
  vmasb200_vec agent_pos[VMASB200_FB_MAX_TEAM];
  float goal_pos[2];                /* right_goal_pos (blue) / left_goal_pos (red), fp32 as the reference holds them */
  const float* pos_shaping;         /* d in : [n_envs] ball.pos_shaping_blue / _red                                 */
  float* pos_shaping_out;
  const float* pos_shaping_agent;   /* d in : [n_envs] ball.pos_shaping_agent_blue / _red                           */
  float* pos_shaping_agent_out;
  float* distance_to_goal;          /* d out: [n_envs] ball.distance_to_goal_blue / _red                            */
  float* pos_rew;                   /* d out: ball.pos_rew_blue / _red                                              */
  float* min_agent_dist_to_ball;    /* d out                                                                        */
  float* pos_rew_agent;             /* d out: ball.pos_rew_agent_blue / _red                                        */
  float* dense_reward;              /* d out: _dense_reward_blue / _red                                             */
  float* sparse_reward;             /* d out: _sparse_reward_blue / _red                                            */
  float* reward;                    /* d out: sparse + dense, what Scenario.reward returns for this team's agents   */
} vmasb200_football_team;
typedef struct vmasb200_football {
  vmasb200_vec ball_pos, ball_vel;
  double pitch_length, ball_size, goal_size, scoring_reward, pos_shaping_factor_ball_goal,
      pos_shaping_factor_agent_ball, distance_to_ball_trigger;
  vmasb200_football_team team[2];   /* blue, red                                                                    */
  uint8_t* done;                    /* d out: [n_envs] Scenario._done                                               */
} vmasb200_football;
int vmasb200_football_reward(const vmasb200_football* fb, int64_t n_envs, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VMASB200_H_ */
