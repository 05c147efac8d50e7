// plan.cpp - builds the static plan from a vmasb200_world_desc (host only, no CUDA).
//
// Replaces the per-substep Python work of World._apply_vectorized_enviornment_force (core.py:2103-2173): pair
// enumeration a<b, joint lookup, bucketing into s_s / l_s / l_l / b_s / b_l / b_b lists, and records the order in
// which update_env_forces (core.py:2190-2198) adds each item to each entity.
#include "plan.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace vb {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
const char* last_error() { return g_err.c_str(); }

static inline int align4(int x) { return (x + 3) & ~3; }

// Largest float x with sqrtf(x) <= t (sqrtf is correctly rounded and monotonic), so that for every float d2
//   sqrtf(d2) <= t  <=>  d2 <= sq_le(t)   and   sqrtf(d2) > t  <=>  d2 > sq_le(t).
static float sq_le(float t) {
  if (!(t >= 0.f)) return -1.f;  // nothing satisfies sqrtf(x) <= negative (or NaN) threshold
  if (std::isinf(t)) return t;
  float x = (float)((double)t * (double)t);
  while (std::sqrt(x) <= t) x = std::nextafter(x, INFINITY);
  while (std::sqrt(x) > t) x = std::nextafter(x, -INFINITY);
  return x;
}
// conservative reach used by the culls: anything farther than this is provably force-free / unseen
static double with_margin(double reach) { return reach * (1.0 + 1e-3) + 1e-5; }
static float sq_up(double reach) { return std::nextafter((float)(reach * reach), INFINITY); }
static const double kLMDd = 4.0 / 6e2;  // utils.py:27 LINE_MIN_DIST

static int pair_type(int sa, int sb, bool& swap) {
  swap = false;
  if (sa == VMASB200_SPHERE && sb == VMASB200_SPHERE) return IT_SS;
  if ((sa == VMASB200_LINE && sb == VMASB200_SPHERE) || (sb == VMASB200_LINE && sa == VMASB200_SPHERE)) {
    swap = sb != VMASB200_SPHERE;  // (line, sphere), core.py:2134-2141
    return IT_LS;
  }
  if (sa == VMASB200_LINE && sb == VMASB200_LINE) return IT_LL;
  if ((sa == VMASB200_BOX && sb == VMASB200_SPHERE) || (sb == VMASB200_BOX && sa == VMASB200_SPHERE)) {
    swap = sb != VMASB200_SPHERE;  // (box, sphere), core.py:2145-2152
    return IT_BS;
  }
  if ((sa == VMASB200_BOX && sb == VMASB200_LINE) || (sb == VMASB200_BOX && sa == VMASB200_LINE)) {
    swap = sb != VMASB200_LINE;  // (box, line), core.py:2156-2163
    return IT_BL;
  }
  return IT_BB;
}

int build_plan(const vmasb200_world_desc* d, vmasb200_plan* pl) {
  if (!d || d->abi_version != VMASB200_ABI_VERSION) { set_error("world desc: bad abi_version"); return VMASB200_ERR_INVALID; }
  const int E = d->n_entities;
  if (E <= 0 || !d->entities || !d->collide_mask) { set_error("world desc: no entities"); return VMASB200_ERR_INVALID; }
  if (d->substeps < 1 || !(d->dt > 0.0)) { set_error("world desc: substeps >= 1 and dt > 0 required"); return VMASB200_ERR_INVALID; }
  if (d->n_joints < 0 || d->n_lidars < 0 || (d->n_joints && !d->joints) || (d->n_lidars && !d->lidars)) {
    set_error("world desc: bad joint / lidar tables"); return VMASB200_ERR_INVALID;
  }
  for (int e = 0; e < E; ++e) {
    int sh = d->entities[e].shape;
    if (sh != VMASB200_SPHERE && sh != VMASB200_BOX && sh != VMASB200_LINE) { set_error("world desc: unknown shape"); return VMASB200_ERR_INVALID; }
  }
  KParams& kp = pl->kp;
  std::memset(&kp, 0, sizeof(kp));
  kp.E = E; kp.S = d->substeps; kp.J = d->n_joints;

  // ---- work items: joints first (core.py:2175), then candidate pairs bucketed by type (core.py:2124-2173)
  pl->items.clear(); pl->pairs.clear();
  for (int j = 0; j < d->n_joints; ++j) {
    const vmasb200_joint& jt = d->joints[j];
    if (jt.entity_a < 0 || jt.entity_a >= E || jt.entity_b < 0 || jt.entity_b >= E || jt.entity_a == jt.entity_b) {
      set_error("world desc: joint entity index out of range"); return VMASB200_ERR_INVALID;
    }
    DItem it{};
    it.type = IT_JOINT; it.a = jt.entity_a; it.b = jt.entity_b; it.thr = 0.f;
    it.dax = jt.delta_a[0]; it.day = jt.delta_a[1]; it.dbx = jt.delta_b[0]; it.dby = jt.delta_b[1];
    it.dist = jt.dist; it.rotate = jt.rotate; it.fixed_rot = jt.fixed_rotation; it.fr_row = -1;
    pl->items.push_back(it);
  }
  for (int t = 0; t < 6; ++t)
    for (int a = 0; a < E; ++a)
      for (int b = a + 1; b < E; ++b) {
        if (!d->collide_mask[a * E + b]) continue;
        bool sw;
        if (pair_type(d->entities[a].shape, d->entities[b].shape, sw) != t) continue;
        DItem it{};
        it.type = t; it.a = sw ? b : a; it.b = sw ? a : b; it.fr_row = -1;
        it.thr = (float)(d->entities[a].circumscribed_radius + d->entities[b].circumscribed_radius);
        pl->items.push_back(it);
        pl->pairs.push_back(vmasb200_pair{t, it.a, it.b});
      }
  kp.n_items = (int)pl->items.size();
  kp.P = kp.n_items - kp.J;
  kp.pair_words = (kp.P + 31) / 32;

  // ---- sensors (sensors.py:46-69, core.py:1677-1690)
  pl->sensors.clear(); pl->ray_angles.clear(); pl->targets.clear(); pl->lidar_slot.assign(d->n_lidars, 0);
  pl->tcull.clear(); pl->ray_cs.clear();
  int ray_off = 0, n_lmask = 0, max_rays = 0;
  for (int k = 0; k < d->n_lidars; ++k) {
    const vmasb200_lidar& ld = d->lidars[k];
    if (ld.entity < 0 || ld.entity >= E || ld.n_rays < 0 || ld.n_targets < 0 || (ld.n_rays && !ld.angles) || (ld.n_targets && !ld.targets)) {
      set_error("world desc: bad lidar row"); return VMASB200_ERR_INVALID;
    }
    DSensor sn{};
    sn.entity = ld.entity; sn.n_rays = ld.n_rays; sn.max_range = ld.max_range; sn.n_targets = ld.n_targets;
    sn.tg_begin = (int)pl->targets.size(); sn.out_off = ray_off;
    for (int t = 0; t < ld.n_targets; ++t) {
      if (ld.targets[t] < 0 || ld.targets[t] >= E) { set_error("world desc: lidar target out of range"); return VMASB200_ERR_INVALID; }
      pl->targets.push_back(ld.targets[t]);
      // a ray of length max_range cannot reach a shape whose centre is farther than max_range + circumscribed radius,
      // nor one whose centre is farther than Rc from the ray's line, nor one that lies behind the origin
      const vmasb200_entity& te = d->entities[ld.targets[t]];
      const double reach = with_margin((double)ld.max_range + te.circumscribed_radius);
      const double lat = with_margin(te.circumscribed_radius);
      pl->tcull.push_back(make_float4(sq_up(reach), (float)lat, te.shape == VMASB200_SPHERE ? 1e-5f : (float)lat,
                                      (float)(reach * 4.8e-7)));
    }
    sn.tw_off = n_lmask;  // relative to row_lmask
    sn.n_tw = (ld.n_targets + 31) / 32;
    n_lmask += sn.n_tw;
    for (int r = 0; r < ld.n_rays; ++r) {
      pl->ray_angles.push_back(ld.angles[r]);
      pl->ray_cs.push_back(make_float2((float)std::cos((double)ld.angles[r]), (float)std::sin((double)ld.angles[r])));
    }
    pl->lidar_slot[k] = ray_off;
    ray_off += ld.n_rays;
    max_rays = std::max(max_rays, ld.n_rays);
    pl->sensors.push_back(sn);
  }
  if (d->n_lidars > 255 || max_rays > 32767) { set_error("more than 255 sensors / 32767 rays per sensor are not supported"); return VMASB200_ERR_UNSUPPORTED; }
  kp.n_sensors = d->n_lidars; kp.n_rays = ray_off; kp.max_rays = max_rays;
  kp.lidar_rows = ray_off;
  kp.n_lmask_words = n_lmask;

  // ---- which entities are read by the step at all, and which need cos/sin
  std::vector<char> used(E, 0), trig(E, 0);
  for (const DItem& it : pl->items) {
    used[it.a] = used[it.b] = 1;
    switch (it.type) {
      case IT_SS: break;
      case IT_LS: case IT_BS: trig[it.a] = 1; break;
      default: trig[it.a] = trig[it.b] = 1; break;  // LL, BL, BB, joints
    }
  }
  for (const DSensor& sn : pl->sensors) {
    used[sn.entity] = 1;
    for (int t = 0; t < sn.n_targets; ++t) {
      int e = pl->targets[sn.tg_begin + t];
      used[e] = 1;
      if (d->entities[e].shape != VMASB200_SPHERE) trig[e] = 1;
    }
  }

  // ---- SoA row layout + entity table
  pl->ents.assign(E, DEnt{});
  pl->dyn_slot.assign(E, -1); pl->stat_slot.assign(E, -1); pl->act_slot.assign(E, -1);
  pl->dyn_list.clear(); pl->trig_list.clear(); pl->agent_entities.clear();
  int dyn_off = 0, stat_off = 0, act_off = 0, n_act = 0, n_stat = 0;
  for (int e = 0; e < E; ++e) {
    const vmasb200_entity& s = d->entities[e];
    DEnt& o = pl->ents[e];
    o.shape = s.shape; o.flags = s.flags & 0x7fffu;
    o.radius = s.radius; o.length = s.length; o.width = s.width;
    o.mass = s.mass; o.moi = s.moment_of_inertia;
    o.max_speed = s.max_speed; o.v_range = s.v_range;
    const double drag = (s.flags & VMASB200_E_DRAG) ? s.drag : d->drag;  // Python floats
    o.one_minus_drag = (float)(1.0 - drag);
    if (s.flags & VMASB200_E_LIN_FRICTION) { o.flags |= DE_LINF_ON; o.lin_fric = s.linear_friction; }
    else if (d->linear_friction > 0.f) { o.flags |= DE_LINF_ON; o.lin_fric = d->linear_friction; }
    if (s.flags & VMASB200_E_ANG_FRICTION) { o.flags |= DE_ANGF_ON; o.ang_fric = s.angular_friction; }
    else if (d->angular_friction > 0.f) { o.flags |= DE_ANGF_ON; o.ang_fric = d->angular_friction; }
    o.gx = s.gravity[0]; o.gy = s.gravity[1];
    o.max_f = s.max_f; o.f_range = s.f_range; o.max_t = s.max_t; o.t_range = s.t_range;
    o.dyn_off = o.stat_off = o.agent = o.act_off = o.trig_row = o.fin_row = o.pos_row = o.rot_row = -1;
    const bool dynamic = s.flags & (VMASB200_E_MOVABLE | VMASB200_E_ROTATABLE);
    if (s.flags & VMASB200_E_AGENT) {
      o.agent = (int)pl->agent_entities.size();
      pl->agent_entities.push_back(e);
    }
    if (dynamic) {
      o.flags |= DE_DYNAMIC;
      o.dyn_off = dyn_off; pl->dyn_slot[e] = dyn_off; dyn_off += 6;
      pl->dyn_list.push_back(e);
      if ((s.flags & VMASB200_E_AGENT) &&
          (s.flags & (VMASB200_E_MAX_F | VMASB200_E_F_RANGE | VMASB200_E_MAX_T | VMASB200_E_T_RANGE))) {
        o.act_off = act_off; pl->act_slot[e] = act_off; act_off += 3; ++n_act;
      }
    } else if (used[e]) {
      o.stat_off = stat_off; pl->stat_slot[e] = stat_off; stat_off += 3; ++n_stat;
    }
    if (trig[e]) {
      o.flags |= DE_TRIG;
      pl->trig_list.push_back(e);
    }
  }
  kp.A = (int)pl->agent_entities.size();
  if (kp.A > kMaxAgents) { set_error("more than 96 agents are not supported by the CUDA step"); return VMASB200_ERR_UNSUPPORTED; }
  pl->joint_slot.assign(d->n_joints, -1);
  for (int j = 0; j < d->n_joints; ++j)
    if (d->joints[j].fixed_rotation_per_env && !d->joints[j].rotate) { pl->joint_slot[j] = stat_off; stat_off += 1; }
  kp.n_dyn = (int)pl->dyn_list.size();
  kp.n_trig = (int)pl->trig_list.size();
  kp.dyn_rows = dyn_off; kp.stat_rows = stat_off; kp.act_rows = act_off;

  // work rows
  kp.row_fin = 0;
  kp.row_trig = kp.row_fin + 3 * kp.n_dyn;
  kp.row_frc = kp.row_trig + 4 * kp.n_trig;
  kp.row_lmask = kp.row_frc + 3 * kp.A;
  kp.row_strig = kp.row_lmask + kp.n_lmask_words;
  kp.n_blk = (kp.P + 31) / 32;
  kp.row_jres = kp.row_strig + 2 * kp.n_sensors;
  kp.row_sv = kp.row_jres + 4 * kp.J;
  kp.row_bp = kp.row_sv + kp.n_blk;
  kp.row_pres = kp.row_bp + kp.n_blk;
  kp.row_taint = kp.row_pres + 5 * kSlots;
  kp.row_th = kp.row_taint + 1;
  kp.work_rows = kp.row_th + 1 + kp.n_blk;
  for (int i = 0; i < kp.n_dyn; ++i) pl->ents[pl->dyn_list[i]].fin_row = kp.row_fin + 3 * i;
  for (int i = 0; i < kp.n_trig; ++i) pl->ents[pl->trig_list[i]].trig_row = kp.row_trig + 4 * i;
  for (int e = 0; e < E; ++e) {
    DEnt& o = pl->ents[e];
    if (o.dyn_off >= 0) { o.pos_row = o.dyn_off; o.rot_row = o.dyn_off + 4; }
    else if (o.stat_off >= 0) { o.pos_row = kp.dyn_rows + o.stat_off; o.rot_row = o.pos_row + 2; }
  }
  for (DSensor& sn : pl->sensors) sn.tw_off += kp.row_lmask;
  for (int i = 0; i < kp.n_items; ++i) {
    DItem& it = pl->items[i];
    const DEnt &a = pl->ents[it.a], &b = pl->ents[it.b];
    it.fin_a = a.fin_row; it.fin_b = b.fin_row;
    it.own = (uint32_t)(a.fin_row >= 0 ? (a.fin_row - kp.row_fin) / 3 : 0) | (uint32_t)(b.fin_row >= 0 ? (b.fin_row - kp.row_fin) / 3 : 0) << 16;
    it.acc = ((a.flags & VMASB200_E_MOVABLE) ? ACC_A_MOV : 0) | ((a.flags & VMASB200_E_ROTATABLE) ? ACC_A_ROT : 0) |
             ((b.flags & VMASB200_E_MOVABLE) ? ACC_B_MOV : 0) | ((b.flags & VMASB200_E_ROTATABLE) ? ACC_B_ROT : 0);
    if (i < kp.J && pl->joint_slot[i] >= 0) it.fr_row = kp.dyn_rows + pl->joint_slot[i];
  }

  // ---- broad-phase table (one row per candidate pair, same order as the pair items)
  pl->broad.clear();
  for (int p = 0; p < kp.P; ++p) {
    const DItem& it = pl->items[kp.J + p];
    const vmasb200_entity &sa = d->entities[it.a], &sb = d->entities[it.b];
    DBroad b{};
    b.pa_row = pl->ents[it.a].pos_row; b.pb_row = pl->ents[it.b].pos_row;
    b.trig_row = 0; b.hx = b.hy = INFINITY;  // circle cull only (any valid work row; the frame test always passes)
    b.t2_bp = sq_le(it.thr);
    if (it.type == IT_SS) {
      const float dmin = sa.radius + sb.radius;  // fp32 sum = the kernel's dist_min (core.py:2312)
      b.t2_cull = sq_le(dmin);
    } else {
      b.t2_cull = sq_up(with_margin(sa.circumscribed_radius + sb.circumscribed_radius + kLMDd));
      if (it.type == IT_LS || it.type == IT_BS) {  // a = line / box, b = sphere: cull in a's frame
        const double reach = with_margin((double)sb.radius + kLMDd);
        b.hx = (float)((double)sa.length / 2 + reach);
        b.hy = (float)((it.type == IT_BS ? (double)sa.width / 2 : 0.0) + reach);
        b.trig_row = pl->ents[it.a].trig_row;
        b.t2_cull = INFINITY;
      }
    }
    pl->broad.push_back(b);
  }
  while (pl->broad.size() % 32) {  // padding rows of the unrolled broad phase: d2 <= -1 never holds
    DBroad b{};
    b.trig_row = 0; b.t2_bp = -1.f; b.t2_cull = -1.f; b.hx = b.hy = INFINITY;
    pl->broad.push_back(b);
  }

  // ---- kernel shape: one thread per env, TE envs per CTA; keep >= 2 CTAs per SM resident when the world allows it
  // per env: staged rows, work rows, lidar rows + ray queue (one int per ray), sensor queue; per warp: 32 narrow-phase
  // tasks (int + float4)
  const int per_env = 4 * (kp.dyn_rows + kp.stat_rows + kp.work_rows + 2 * kp.lidar_rows + kp.n_sensors) + 20;
  const int fixed = 64 + 4 * kp.S * kp.pair_words;
  int TE = d->envs_per_cta;
  if (TE != 0 && TE != 32 && TE != 64 && TE != 128) { set_error("envs_per_cta must be 0, 32, 64 or 128"); return VMASB200_ERR_INVALID; }
  pl->te_fixed = TE;
  if (TE == 0) TE = 32;  // the launch picks 32 or 64 by batch size (vmasb200.cu tile_for)
  if (fixed + TE * per_env > 226 * 1024) { set_error("world too large for one CTA's shared memory"); return VMASB200_ERR_UNSUPPORTED; }
  kp.TE = TE;
  kp.NT = 128;  // Q = 128 / TE team members per env

  // ---- world constants
  kp.sub_dt = (float)(d->dt / (double)d->substeps);  // core.py:1124 (Python doubles)
  kp.Kc = d->collision_force; kp.Kj = d->joint_force; kp.Kt = d->torque_constraint_force; kp.margin = d->contact_margin;
  kp.xs = d->x_semidim; kp.ys = d->y_semidim; kp.has_xs = d->has_x_semidim; kp.has_ys = d->has_y_semidim;
  kp.wgx = d->gravity[0]; kp.wgy = d->gravity[1];
  kp.has_wgrav = !(d->gravity[0] == 0.f && d->gravity[1] == 0.f);  // core.py:2044

  vmasb200_layout& L = pl->layout;
  L.dyn_rows = kp.dyn_rows; L.stat_rows = kp.stat_rows; L.act_rows = kp.act_rows; L.lidar_rows = kp.lidar_rows;
  L.n_pairs = kp.P; L.pair_words = kp.pair_words; L.envs_per_cta = kp.TE;
  L.n_dyn = kp.n_dyn; L.n_stat = n_stat; L.n_agents = kp.A; L.n_act_out = n_act;
  // state read + write, static pose read, action read, clamped action write, lidar write
  L.alg_bytes_per_env = 4ll * (12 * kp.n_dyn + 3 * n_stat + 3 * kp.A + 3 * n_act + kp.n_rays);
  return VMASB200_OK;
}

}  // namespace vb
