// tables.h - plain-data plan tables shared by the host plan builder (plan.cpp part of vmasb200.cu) and the kernels.
//
// The plan is the static half of the reference's per-substep Python work: entity attributes (core.py:537-1085),
// the pair enumeration + bucketing of core.py:2103-2173, the joint list (joints.py:147-215), the per-entity
// accumulation order of update_env_forces (core.py:2190-2198) and the Lidar target lists (core.py:1677-1690).
// Everything is 32-bit words so the tables can be read with ld.global.nc by every lane.
#pragma once
#include <stdint.h>

#include "../../include/vmasb200.h"

namespace vb {

// extra (resolved) entity flag bits on top of VMASB200_E_* (bits 0..14)
enum : uint32_t {
  DE_LINF_ON = 1u << 16,   // a linear friction coefficient applies (entity's, else world's > 0; core.py:2075-2087)
  DE_ANGF_ON = 1u << 17,   // same for angular friction (core.py:2088-2101)
  DE_TRIG = 1u << 18,      // cos/sin of the rotation are needed by some pair / joint / ray
  DE_DYNAMIC = 1u << 19    // has a dyn record (movable or rotatable)
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
  int32_t dyn_off;          // float offset inside the dyn record (px py vx vy rot w) or -1
  int32_t stat_off;         // float offset inside the stat record (px py rot) or -1
  int32_t agent;            // agent slot (index into the force / torque pointer tables) or -1
  int32_t act_off;          // float offset of (fx fy tq) inside the actout record or -1
  int32_t rot_slot;         // slot of the cos/sin cache or -1
  int32_t fin_off;          // float offset of (Fx Fy T) in the per-env work area (dynamic entities only)
  int32_t pad[8];
};
static_assert(sizeof(DEnt) == 32 * 4, "DEnt must be 32 words");

enum { IT_SS = 0, IT_LS, IT_LL, IT_BS, IT_BL, IT_BB, IT_JOINT };

struct DItem {              // 12 words; joints first (in reference order), then candidate pairs in type order
  int32_t type;
  int32_t a, b;             // entity indices, already in the reference's (line,sphere)/(box,sphere)/(box,line) order
  float thr;                // broad-phase threshold (float)(Rc_a + Rc_b in double), core.py:2796-2800
  float dax, day, dbx, dby; // joint anchor deltas (joints.py:185-201)
  float dist;               // joint distance
  int32_t rotate;           // joint.rotate
  float fixed_rot;
  int32_t fr_off;           // float offset inside the stat record of the per-env fixed rotation or -1
};
static_assert(sizeof(DItem) == 12 * 4, "DItem must be 12 words");

// Broad-phase row of candidate pair p (same order as the pair items): everything phase A of a substep reads, as two
// 16-byte loads.  Position offsets are encoded: >= 0 float offset inside the dyn record, < 0: ~offset inside the stat
// record.  The tests are sqrt-free but EXACT:
//   verdict of World.collides (core.py:2796-2800):  sqrtf(d2) <= thr   <=>  d2 <= t2_bp
//   (t2_bp = the largest float whose correctly rounded square root is <= thr; d2 = fma(dy,dy,dx*dx) is the very
//   argument torch.linalg.vector_norm takes the root of)
// and the cull only drops pairs whose force is provably zero (dist > dist_min, core.py:2835):
//   rot_off < 0   circle cull: keep iff !(d2 > t2_cull)   (sphere-sphere: exact image of dist > r_a + r_b;
//                 other types: (Rc_a + Rc_b + LINE_MIN_DIST) with a safety margin)
//   rot_off >= 0  frame cull (a = line or box, b = sphere): b's centre in a's frame, keep iff |x| <= hx and |y| <= hy
//                 (half extents + radius + LINE_MIN_DIST with a safety margin); rot_off = float offset of a's
//                 cos/sin cache inside the env work area.
struct DBroad {             // 8 words
  int32_t pa_off, pb_off;
  int32_t rot_off;
  int32_t entry;            // list entry of the pair: item index | a << 12 | b << 21  (bit 30 = verdict, set at run time)
  float t2_bp, t2_cull, hx, hy;
};
static_assert(sizeof(DBroad) == 8 * 4, "DBroad must be 8 words");
constexpr int kEntryItemBits = 12, kEntryEntBits = 9;   // limits: 4096 items, 512 entities

struct DRay {               // one per (sensor, ray): 4 words
  int32_t sensor;
  float angle;              // sensor-frame angle (row 0 of Lidar._angles)
  int32_t out_off;          // float offset inside the lidar record
  int32_t pad;
};

struct DSensor {            // 8 words
  int32_t entity;
  int32_t n_rays;
  float max_range;
  int32_t n_targets;
  int32_t tg_begin;         // index into the target list
  int32_t out_off;
  int32_t tw_off;           // first word of this sensor's target survival mask inside the env's lmask area
  int32_t n_tw;             // words of that mask = ceil(n_targets / 32)
};

struct DSWord {             // one per (sensor, 32-target word) of the LIDAR cull: 4 words
  int32_t sensor;
  int32_t tg_begin;         // index into targets[] / tcull2[] of bit 0
  int32_t n;                // targets in this word (<= 32)
  int32_t pad;
};

constexpr int kMaxAgents = 96;   // by-value kernel parameter tables

struct AgentIO {            // kernel parameter (by value)
  const float* force[kMaxAgents];
  const float* torque[kMaxAgents];
  int32_t fstride[kMaxAgents];
  int32_t tstride[kMaxAgents];
};

struct KParams {            // kernel parameter (by value)
  // sizes
  int32_t E, n_dyn, J, P, n_items, S, A, n_rot, n_sensors, n_rays, pair_words;
  int32_t dyn_stride, stat_stride, act_stride, lidar_stride;
  int32_t G, TE;            // lanes per env, envs per CTA (threads per CTA = G * TE)
  // per-env work area (4-byte words): [fin 3*n_dyn][rot 4*n_rot][res 4*n_items][frc 3*A][list n_items][lmask]
  int32_t work_stride, off_fin, off_rot, off_res, off_frc, off_list, off_lmask, n_lmask_words;
  // device tables
  const DEnt* ents;
  const DItem* items;
  const DBroad* broad;      // [P]
  const int32_t* dyn_list;  // entity index of every dynamic entity, in entity order
  const int32_t* rot_list;  // entity index per rot slot
  const DRay* rays;
  const DSensor* sensors;
  const int32_t* targets;
  const float* tcull2;      // [n targets] squared LIDAR cull distance per (sensor, target): (max_range + Rc) with margin
  const DSWord* swords;     // [n_lmask_words]
  // world constants (core.py:1090-1143)
  float sub_dt, Kc, Kj, Kt, margin;
  float xs, ys;
  int32_t has_xs, has_ys;
  float wgx, wgy;
  int32_t has_wgrav;
};

}  // namespace vb
