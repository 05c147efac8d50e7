// tables.h - plain-data plan tables shared by the host plan builder (plan.cpp) and the kernels.
//
// The plan is the static half of the reference's per-substep Python work: entity attributes (core.py:537-1085),
// the pair enumeration + bucketing of core.py:2103-2173, the joint list (joints.py:147-215), the accumulation order
// of update_env_forces (core.py:2190-2198) and the Lidar target lists (core.py:1677-1690).
//
// Execution model the tables are laid out for: ONE THREAD PER ENVIRONMENT.  Every loop over entities / pairs /
// joints / sensors is uniform across the warp, so every table read below is a warp-uniform (broadcast) load and
// all index arithmetic runs once per warp; only the env's own state differs between lanes.  That state lives in
// shared memory as structure of arrays, "rows" of TE floats (TE = envs per CTA): word `row * TE + tid`.
//   env rows  [0, dyn_rows)                  staged dyn buffer rows   (6 per dynamic entity: px py vx vy rot w)
//             [dyn_rows, dyn_rows+stat_rows) staged stat buffer rows  (3 per static entity: px py rot; joint rots)
//   work rows fin  3 per dynamic entity (Fx Fy T of the running substep, core.py:1975-1992)
//             trig 4 per entity that needs cos/sin (c s c2 s2)
//             frc  3 per agent (state.force x y, state.torque; clamped in place, core.py:2017-2040)
//             lmask one per (sensor, 32-target word)
//             strig 2 per sensor (cos/sin of the sensing entity's rotation, for the ray cull)
//             jres 4 per joint, pres 5 per result slot (fx fy ta tb of an evaluated pair, pair index), sv / bp one per 32-pair
//             block (survivor / verdict bits of the running substep), taint 1 (record index + 1 of a tainted env),
//             th 1 + one per 32-pair block (thin events of the running substep)
// A TEAM of Q = NT / TE warps-groups shares the TE columns: thread t works on column t % TE as member q = t / TE and
// takes the entities / joints / pair blocks / sensors with index = q (mod Q); order-sensitive sums stay with member 0.
#pragma once
#include <stdint.h>

#include "../../include/vmasb200.h"

namespace vb {

// extra (resolved) entity flag bits on top of VMASB200_E_* (bits 0..14)
enum : uint32_t {
  DE_LINF_ON = 1u << 16,   // a linear friction coefficient applies (entity's, else world's > 0; core.py:2075-2087)
  DE_ANGF_ON = 1u << 17,   // same for angular friction (core.py:2088-2101)
  DE_TRIG = 1u << 18,      // cos/sin of the rotation are needed by some pair / joint / ray
  DE_DYNAMIC = 1u << 19    // has dyn rows (movable or rotatable)
};

struct DEnt {               // 32 words
  int32_t shape;
  uint32_t flags;
  float radius, length, width;
  float mass, moi;
  float max_speed, v_range;
  float one_minus_drag;     // 1 - (entity.drag ?? world.drag) evaluated in double, then fp32 (core.py:2866)
  float lin_fric, ang_fric; // resolved coefficient
  float gx, gy;             // entity gravity
  float max_f, f_range, max_t, t_range;
  int32_t dyn_off;          // first row inside the dyn buffer (px py vx vy rot w) or -1
  int32_t stat_off;         // first row inside the stat buffer (px py rot) or -1
  int32_t agent;            // agent slot (index into the force / torque pointer tables) or -1
  int32_t act_off;          // first row of (fx fy tq) inside the actout buffer or -1
  int32_t trig_row;         // work row of the cos/sin cache (c s c2 s2) or -1
  int32_t fin_row;          // work row of (Fx Fy T) (dynamic entities only) or -1
  int32_t pos_row;          // env row of px (py = +1): dyn_off, or dyn_rows + stat_off; -1 = not staged
  int32_t rot_row;          // env row of rot: pos_row + 4 (dynamic) / pos_row + 2 (static)
  int32_t pad[6];
};
static_assert(sizeof(DEnt) == 32 * 4, "DEnt must be 32 words");

enum { IT_SS = 0, IT_LS, IT_LL, IT_BS, IT_BL, IT_BB, IT_JOINT };

// accumulation switches of update_env_forces (core.py:2190-2198)
enum : int32_t { ACC_A_MOV = 1, ACC_A_ROT = 2, ACC_B_MOV = 4, ACC_B_ROT = 8 };

struct DItem {              // 16 words; joints first (in reference order), then candidate pairs in type order
  int32_t type;
  int32_t a, b;             // entity indices, already in the reference's (line,sphere)/(box,sphere)/(box,line) order
  float thr;                // broad-phase threshold (float)(Rc_a + Rc_b in double), core.py:2796-2800
  float dax, day, dbx, dby; // joint anchor deltas (joints.py:185-201)
  float dist;               // joint distance
  int32_t rotate;           // joint.rotate
  float fixed_rot;
  int32_t fr_row;           // env row of the per-env fixed rotation or -1
  int32_t fin_a, fin_b;     // work rows of the (Fx Fy T) accumulators of a / b, or -1 (static entity)
  int32_t acc;              // ACC_* bits
  uint32_t own;             // dynamic-entity index (position in dyn_list) of a | of b << 16: who of the team sums it
};
static_assert(sizeof(DItem) == 16 * 4, "DItem must be 16 words");

// Broad-phase row of candidate pair p (same order as the pair items): everything the cull of a substep reads, as two
// 16-byte loads.  The tests are sqrt-free but EXACT:
//   verdict of World.collides (core.py:2796-2800):  sqrtf(d2) <= thr   <=>  d2 <= t2_bp
//   (t2_bp = the largest float whose correctly rounded square root is <= thr; d2 = fma(dy,dy,dx*dx) is the very
//   argument torch.linalg.vector_norm takes the root of)
// and the cull only drops pairs whose force is provably zero (dist > dist_min, core.py:2835):
//   circle cull: keep iff !(d2 > t2_cull)   (sphere-sphere: exact image of dist > r_a + r_b; other types:
//                (Rc_a + Rc_b + LINE_MIN_DIST) with a safety margin); such rows carry hx = hy = +inf
//   frame cull (a = line or box, b = sphere): b's centre in a's frame, keep iff |x| <= hx and |y| <= hy (half
//                extents + radius + LINE_MIN_DIST with a safety margin); trig_row = work row of a's cos/sin;
//                such rows carry t2_cull = +inf.
// Every row is evaluated with BOTH tests (branch-free); NaN passes both.
struct DBroad {             // 8 words
  int32_t pa_row, pb_row;   // env rows of the two positions
  int32_t trig_row;
  int32_t pad;
  float t2_bp, t2_cull, hx, hy;
};
static_assert(sizeof(DBroad) == 8 * 4, "DBroad must be 8 words");

struct DSensor {            // 8 words
  int32_t entity;
  int32_t n_rays;
  float max_range;
  int32_t n_targets;
  int32_t tg_begin;         // index into the target list
  int32_t out_off;          // first row inside the lidar buffer
  int32_t tw_off;           // first lmask work row of this sensor
  int32_t n_tw;             // words of that mask = ceil(n_targets / 32)
};

constexpr int kSlots = 8;        // narrow-phase results a team holds per env and round (more survivors = more rounds)
constexpr int kMaxAgents = 96;   // by-value kernel parameter tables

struct AgentIO {            // kernel parameter (by value)
  const float* force[kMaxAgents];
  const float* torque[kMaxAgents];
  int32_t fstride[kMaxAgents];
  int32_t tstride[kMaxAgents];
};

struct KParams {            // kernel parameter (by value)
  // sizes
  int32_t E, n_dyn, J, P, n_items, S, A, n_trig, n_sensors, n_rays, pair_words, max_rays;
  int32_t dyn_rows, stat_rows, act_rows, lidar_rows;
  int32_t TE;               // envs per CTA (= threads that run the physics)
  int32_t NT;               // threads per CTA: TE, or a multiple of it when helper warps share the LIDAR epilogue
  // work rows: [fin 3*n_dyn][trig 4*n_trig][frc 3*A][lmask][strig 2*n_sensors]
  int32_t work_rows, row_fin, row_trig, row_frc, row_lmask, n_lmask_words, row_strig;
  // team rows: [jres 4*J][sv n_blk][bp n_blk][pres 5*kSlots][taint 1][th 1 + n_blk]; n_blk = ceil(P / 32)
  int32_t row_jres, row_sv, row_bp, row_pres, row_taint, n_blk;
  int32_t row_th;           // [1 + n_blk]: marker + thin-event bits of the running substep (speculative launch)
  // device tables
  const DEnt* ents;
  const DItem* items;
  const DBroad* broad;      // [P]
  const int32_t* dyn_list;  // entity index of every dynamic entity, in entity order
  const int32_t* trig_list; // entity index per trig slot
  const int32_t* agent_ents;// entity index per agent slot
  const float* ray_angles;  // [n_rays] sensor-frame angles, sensors concatenated (row 0 of Lidar._angles)
  const DSensor* sensors;
  const int32_t* targets;
  // [n targets] LIDAR culls per (sensor, target): x = squared reach (max_range + Rc, with margin) of the sensor-level
  // cull; y = lateral reach of the ray-level cull (Rc with margin); z = how far behind the origin the centre may be;
  // w = slack per radian of ray angle (the cull uses an approximate ray direction)
  const float4* tcull;
  const float2* ray_cs;     // [n_rays] (cos, sin) of the sensor-frame ray angles (ray cull only)
  // world constants (core.py:1090-1143)
  float sub_dt, Kc, Kj, Kt, margin;
  float xs, ys;
  int32_t has_xs, has_ys;
  float wgx, wgy;
  int32_t has_wgrav;
};

}  // namespace vb
