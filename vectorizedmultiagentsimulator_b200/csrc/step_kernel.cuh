// step_kernel.cuh - device code of the fused World.step kernel (sm_100a): one thread per environment.
//
// One CTA owns a tile of TE consecutive environments.  The tile's rows of the SoA buffers (dyn + stat) are staged into
// shared memory by one cp.async.bulk (TMA 1-D) per row, stay resident across all substeps, and are written back by
// one bulk store per output row.  Thread `tid` owns column `tid` of every row, so every shared-memory access of a
// warp is 32 consecutive words (conflict-free) and every loop below - entities, joints, pairs, sensors - has a
// warp-uniform trip count and warp-uniform table reads.  Per substep (core.py:1974-2009):
//   1  joints (core.py:2200-2291), accumulated straight into the (Fx Fy T) rows;
//   2  pairs in the reference's type order (core.py:2124-2173): a sqrt-free EXACT broad phase (tables.h DBroad) takes
//      the reference's own verdict (core.py:2796-2800) and drops every pair whose force is provably zero; the verdict
//      bits of 32 pairs are OR-reduced over the warp with one REDUX; only surviving pairs run the narrow phase
//      (closest points, penalty force, torques) and add into the accumulators in the reference's order
//      (core.py:2175-2198; a dropped pair would have added exact zeros);
//   3  integrate every dynamic entity (core.py:2861-2907), refresh cos/sin of a rotated entity, open the next
//      substep's accumulators (action / friction / gravity, core.py:1994-2005).
// After the last substep the LIDAR epilogue (core.py:1661-1785): per (env, sensor) a mask of the targets within reach;
// the (env, sensor) pairs that can see anything are compacted into a CTA queue and the CTA's threads then stride over
// (queue entry, ray) so that ray casting runs on full warps.
#pragma once
#include "geom.cuh"
#include "tables.h"

namespace vb {

// eval_item has three call sites (joints, pair survivors, repair); out of line keeps the hot path's code compact
#ifdef VB_EVAL_INLINE
#define VB_EVAL VB_DI
#else
#define VB_EVAL __device__ __noinline__
#endif

// ---------------------------------------------------------------------------------------- TMA / mbarrier PTX
VB_DI uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
VB_DI void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
VB_DI void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
VB_DI void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
VB_DI void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
VB_DI void mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(phase)
      : "memory");
}
// global -> shared bulk copy, completion counted on the mbarrier (bytes % 16 == 0, both addresses 16B aligned)
VB_DI void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
VB_DI void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
               "r"(bytes)
               : "memory");
}
VB_DI void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
VB_DI void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// the sources in shared memory have been read (enough to leave the CTA; the writes complete by the end of the kernel)
VB_DI void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// x / c for a divisor that is normally a positive finite constant (mass, moment of inertia).  +-0 / c is +-0 itself, and
// a zero numerator sends the IEEE division to its slow path (ncu: ~230 instructions per warp and step in navigation,
// whose spheres never receive a torque): branch around it.  Same bits as x / c in every case.
VB_DI float div_pos(float x, float c) {
#ifndef VB_AB_PLAIN_DIV
  if (x == 0.f && c > 0.f && c < 3.0e38f) return x;
  asm volatile("" : "+f"(x));  // keeps the division on this side of the branch (the compiler would otherwise divide, slow path and all, and select)
#endif
  return x / c;
}

// Plan tables every substep walks, in the constant bank: the walks are warp-uniform, so these become uniform-datapath
// constant loads (LDCU) with uniform address arithmetic instead of per-thread global loads.  One copy per device; the
// host re-uploads it (stream-ordered) when a launch belongs to another plan (vmasb200.cu bind_const).  Worlds that do
// not fit take the same kernel with the tables in global memory.
constexpr int kCPairs = 1024, kCEnts = 160, kCSensors = 64;
struct CTab {
  DBroad broad[kCPairs];
  DEnt ents[kCEnts];
  int32_t dyn_list[kCEnts], trig_list[kCEnts], agent_ents[kMaxAgents];
  DSensor sensors[kCSensors];
};
__constant__ CTab c_tab;
// parity probe (vmasb200_step_io.jexp_in): slot 0 of a joint's result rows holds the reference's exp(|delta rot|) for the
// constraint torque of the substep; set around a probe launch only
__constant__ int c_probe_jexp;

// table accessors: CT = the walk tables are in the constant bank (c_tab), else in global memory behind kp
template <bool CT> VB_DI const DEnt& T_ent(const KParams& kp, int e) { if (CT) return c_tab.ents[e]; else return kp.ents[e]; }
template <bool CT> VB_DI const DBroad& T_broad(const KParams& kp, int p) { if (CT) return c_tab.broad[p]; else return kp.broad[p]; }
template <bool CT> VB_DI const DSensor& T_sensor(const KParams& kp, int k) { if (CT) return c_tab.sensors[k]; else return kp.sensors[k]; }
template <bool CT> VB_DI int T_dyn(const KParams& kp, int d) { if (CT) return c_tab.dyn_list[d]; else return kp.dyn_list[d]; }
template <bool CT> VB_DI int T_trig(const KParams& kp, int r) { if (CT) return c_tab.trig_list[r]; else return kp.trig_list[r]; }
template <bool CT> VB_DI int T_agent(const KParams& kp, int a) { if (CT) return c_tab.agent_ents[a]; else return kp.agent_ents[a]; }

// ---------------------------------------------------------------------------------------- per-env view of the tile
// TE (envs per CTA = floats per smem row) is a compile-time constant: row addressing is one shift, row + 1 an immediate
template <int TE_>
struct EnvT {
  static constexpr int TE = TE_;
  float* env;   // &envS[column]: row r of this env's dyn / stat state at env[r * TE]
  float* work;  // &workS[column]
};
template <class EV> VB_DI float& ER(const EV& E, int row) { return E.env[row * EV::TE]; }
template <class EV> VB_DI float& WR(const EV& E, int row) { return E.work[row * EV::TE]; }
template <class EV> VB_DI V2 pos_at(const EV& E, int row) { return mk(ER(E, row), ER(E, row + 1)); }

template <class EV>
VB_DI Rot trig_of(const EV& E, const DEnt* __restrict__ en) {
  const int r = en->trig_row;
  Rot o;
  o.c = WR(E, r); o.s = WR(E, r + 1); o.c2 = WR(E, r + 2); o.s2 = WR(E, r + 3);
  return o;
}

// cos/sin cache of one entity (physics.py:299-301, 412; utils.py:183-184)
template <class EV>
VB_DI void store_trig(const DEnt* __restrict__ en, const EV& E, float rot) {
  const int r = en->trig_row;
  float s, c;
  sincosf(rot, &s, &c);
  WR(E, r) = c; WR(E, r + 1) = s;
  if (en->shape == VMASB200_BOX) {
    float rot2 = rot + (float)(3.14159265358979323846 / 2);
    sincosf(rot2, &s, &c);
    WR(E, r + 2) = c; WR(E, r + 3) = s;
  }
}

// Start of a substep for one dynamic entity: F = T = 0, then action force/torque (core.py:2017-2040, with the
// clamped value written back to the env's frc rows), friction (core.py:2053-2101), gravity (core.py:2042-2051).
template <class EV>
VB_DI void begin_forces(const KParams& kp, const DEnt* __restrict__ en, const EV& E) {
  const uint32_t fl = en->flags;
  V2 F = mk(0.f, 0.f);
  float T = 0.f;
  const int d = en->dyn_off;
  if (fl & VMASB200_E_AGENT) {
    const int fr = kp.row_frc + 3 * en->agent;
    if (fl & VMASB200_E_MOVABLE) {
      V2 f = mk(WR(E, fr), WR(E, fr + 1));
      if (fl & VMASB200_E_MAX_F) f = clamp_norm2(f, en->max_f);
      if (fl & VMASB200_E_F_RANGE) {
        float r = en->f_range;
        f = mk(clampf(f.x, -r, r), clampf(f.y, -r, r));
      }
      WR(E, fr) = f.x; WR(E, fr + 1) = f.y;
      F = F + f;
    }
    if (fl & VMASB200_E_ROTATABLE) {
      float t = WR(E, fr + 2);
      if (fl & VMASB200_E_MAX_T) t = clamp_norm1(t, en->max_t);
      if (fl & VMASB200_E_T_RANGE) t = clampf(t, -en->t_range, en->t_range);
      WR(E, fr + 2) = t;
      T = T + t;
    }
  }
  if (fl & DE_LINF_ON) F = F + friction2(mk(ER(E, d + 2), ER(E, d + 3)), en->lin_fric, en->mass, kp.sub_dt);
  if (fl & DE_ANGF_ON) T = T + friction1(ER(E, d + 5), en->ang_fric, en->moi, kp.sub_dt);
  if (fl & VMASB200_E_MOVABLE) {
    if (kp.has_wgrav) F = F + mk(en->mass * kp.wgx, en->mass * kp.wgy);
    if (fl & VMASB200_E_GRAVITY) F = F + mk(en->mass * en->gx, en->mass * en->gy);
  }
  const int fin = en->fin_row;
  WR(E, fin) = F.x; WR(E, fin + 1) = F.y; WR(E, fin + 2) = T;
}

// One work item (joint constraint or candidate pair): force on a (the force on b is its exact negation) and the
// two torques.  `type` is warp-uniform, so the switch does not diverge.
template <bool CT, class EV>
VB_EVAL void eval_item(const KParams& kp, const DItem* __restrict__ it, const EV& E, V2& fa, float& ta, float& tb) {
  const int type = it->type;
  const DEnt* __restrict__ ea = &T_ent<CT>(kp, it->a);
  const DEnt* __restrict__ eb = &T_ent<CT>(kp, it->b);
  const V2 pa = pos_at(E, ea->pos_row), pb = pos_at(E, eb->pos_row);
  const float Kc = kp.Kc, k = kp.margin;
  ta = 0.f; tb = 0.f;
  switch (type) {
    case IT_JOINT: {  // core.py:2200-2291, joints.py:206-215
      Rot qa = trig_of(E, ea), qb = trig_of(E, eb);
      V2 ja = pa + rotv(mk(it->dax, it->day), qa.c, qa.s);
      V2 jb = pb + rotv(mk(it->dbx, it->dby), qb.c, qb.s);
      V2 att, rep;
      constraint_force_pair(ja, jb, it->dist, kp.Kj, k, att, rep);
      fa = att + rep;
      V2 fb = (-att) + (-rep);
      ta = cross(ja - pa, fa);
      tb = cross(jb - pb, fb);
      if (!it->rotate) {
        const float ra = ER(E, ea->rot_row), rb = ER(E, eb->rot_row);
        float fr = it->fr_row >= 0 ? ER(E, it->fr_row) : it->fixed_rot;
        float tq;
        if (c_probe_jexp) {
          const float d = ra - (rb + fr);
          tq = constraint_torque_exp(d, sqrtf(d * d), WR(E, kp.row_jres + 4 * (int)(it - kp.items)), kp.Kt);
        } else {
          tq = constraint_torque(ra, rb + fr, kp.Kt);
        }
        ta = ta + (-tq);
        tb = tb + tq;
      }
    } break;
    case IT_SS: {  // core.py:2293
      fa = constraint_force(pa, pb, ea->radius + eb->radius, Kc, k, false);
    } break;
    case IT_LS: {  // core.py:2340; a = line, b = sphere
      Rot qa = trig_of(E, ea);
      V2 cp = closest_point_line(pa, qa.c, qa.s, ea->length, pb);
      V2 fs = constraint_force(pb, cp, eb->radius + kLMD, Kc, k, false);
      fa = -fs;
      ta = cross(cp - pa, fa);
    } break;
    case IT_LL: {  // core.py:2393
      Rot qa = trig_of(E, ea), qb = trig_of(E, eb);
      V2 ca, cb;
      closest_points_line_line(pa, qa.c, qa.s, ea->length, pb, qb.c, qb.s, eb->length, ca, cb);
      fa = constraint_force(ca, cb, kLMD, Kc, k, false);
      ta = cross(ca - pa, fa);
      tb = cross(cb - pb, -fa);
    } break;
    case IT_BS: {  // core.py:2458; a = box, b = sphere
      Rot qa = trig_of(E, ea);
      V2 cp = closest_point_box(pa, qa, ea->width, ea->length, pb);
      V2 ip = cp;
      float dd = 0.f;
      if (!(ea->flags & VMASB200_E_HOLLOW)) ip = inner_point_box(pb, cp, pa, dd);
      V2 fs = constraint_force(pb, ip, (eb->radius + kLMD) + dd, Kc, k, false);
      fa = -fs;
      ta = cross(cp - pa, fa);
    } break;
    case IT_BL: {  // core.py:2553; a = box, b = line
      Rot qa = trig_of(E, ea), qb = trig_of(E, eb);
      V2 cbx, cl;
      closest_line_box(pa, qa, ea->width, ea->length, pb, qb.c, qb.s, eb->length, cbx, cl);
      V2 ip = cbx;
      float dd = 0.f;
      if (!(ea->flags & VMASB200_E_HOLLOW)) ip = inner_point_box(cl, cbx, pa, dd);
      fa = constraint_force(ip, cl, kLMD + dd, Kc, k, false);
      ta = cross(cbx - pa, fa);
      tb = cross(cl - pb, -fa);
    } break;
    default: {  // core.py:2654 box-box
      Rot qa = trig_of(E, ea), qb = trig_of(E, eb);
      V2 ca, cb;
      closest_box_box(pa, qa, ea->width, ea->length, pb, qb, eb->width, eb->length, ca, cb);
      V2 ia = ca, ib = cb;
      float da = 0.f, db = 0.f;
      if (!(ea->flags & VMASB200_E_HOLLOW)) ia = inner_point_box(cb, ca, pa, da);
      if (!(eb->flags & VMASB200_E_HOLLOW)) ib = inner_point_box(ca, cb, pb, db);
      fa = constraint_force(ia, ib, (da + db) + kLMD, Kc, k, false);
      ta = cross(ca - pa, fa);
      tb = cross(cb - pb, -fa);
    } break;
  }
}

// update_env_forces (core.py:2190-2198): F_a += f_a if a.movable, T_a += t_a if a.rotatable, then b with -f_a.
template <int Q, class EV>
VB_DI void accumulate(const DItem* __restrict__ it, const EV& E, V2 fa, float ta, float tb, int q) {
  // member q of the env's team sums the entities it owns (dyn index = q mod Q): every accumulator still receives
  // its addends in the reference's order
  const int acc = it->acc;
  const uint32_t own = it->own;
  if ((acc & (ACC_A_MOV | ACC_A_ROT)) && (int)(own & (Q - 1)) == q) {
    const int r = it->fin_a;
    if (acc & ACC_A_MOV) { WR(E, r) = WR(E, r) + fa.x; WR(E, r + 1) = WR(E, r + 1) + fa.y; }
    if (acc & ACC_A_ROT) WR(E, r + 2) = WR(E, r + 2) + ta;
  }
  if ((acc & (ACC_B_MOV | ACC_B_ROT)) && (int)((own >> 16) & (Q - 1)) == q) {
    const int r = it->fin_b;
    if (acc & ACC_B_MOV) { WR(E, r) = WR(E, r) + (-fa.x); WR(E, r + 1) = WR(E, r + 1) + (-fa.y); }
    if (acc & ACC_B_ROT) WR(E, r + 2) = WR(E, r + 2) + tb;
  }
}

// _integrate_state of one dynamic entity (core.py:2861-2907) from its accumulated (Fx Fy T).
template <class EV>
VB_DI void integrate(const KParams& kp, const DEnt* __restrict__ en, const EV& E, int substep) {
  const uint32_t fl = en->flags;
  const int fin = en->fin_row, d = en->dyn_off;
  const float sub_dt = kp.sub_dt;
  if (fl & VMASB200_E_MOVABLE) {
    V2 F = mk(WR(E, fin), WR(E, fin + 1));
    V2 vel = mk(ER(E, d + 2), ER(E, d + 3));
    if (substep == 0) vel = vel * en->one_minus_drag;
    const float mass = en->mass;
    V2 acc = mass == 1.f ? F : mk(div_pos(F.x, mass), div_pos(F.y, mass));  // x / 1 == x exactly
    vel = vel + acc * sub_dt;
    if (fl & VMASB200_E_MAX_SPEED) vel = clamp_norm2(vel, en->max_speed);
    if (fl & VMASB200_E_V_RANGE) {
      float r = en->v_range;
      vel = mk(clampf(vel.x, -r, r), clampf(vel.y, -r, r));
    }
    V2 np = mk(ER(E, d), ER(E, d + 1)) + vel * sub_dt;
    if (kp.has_xs) np.x = clampf(np.x, -kp.xs, kp.xs);
    if (kp.has_ys) np.y = clampf(np.y, -kp.ys, kp.ys);
    ER(E, d + 2) = vel.x; ER(E, d + 3) = vel.y; ER(E, d) = np.x; ER(E, d + 1) = np.y;
  }
  if (fl & VMASB200_E_ROTATABLE) {
    float w = ER(E, d + 5);
    if (substep == 0) w = w * en->one_minus_drag;
    w = w + div_pos(WR(E, fin + 2), en->moi) * sub_dt;
    ER(E, d + 5) = w;
    float nr = ER(E, d + 4) + w * sub_dt;
    ER(E, d + 4) = nr;
    if (en->trig_row >= 0) store_trig(en, E, nr);
  }
}

// One LIDAR ray (sensors.py:115-118 -> core.py:1661-1785) of env view E: min over the sensor's targets within reach
// (the sensor's survival mask words are in the env's lmask rows; an unreachable target would have given max_range).
template <bool CT, class EV>
VB_DI float cast_one_ray(const KParams& kp, const DSensor* __restrict__ sn, const EV& E, V2 o, float ang) {
  const float max_range = sn->max_range;
  float best = max_range;
  float ds, dc;
  sincosf(ang, &ds, &dc);
  for (int w = 0; w < sn->n_tw; ++w) {
    uint32_t m = __float_as_uint(WR(E, sn->tw_off + w));
    const int tb = sn->tg_begin + 32 * w;
    while (m) {
      const int t = __ffs(m) - 1;
      m &= m - 1;
      const DEnt* __restrict__ en = &T_ent<CT>(kp, __ldg(kp.targets + tb + t));
      V2 p = pos_at(E, en->pos_row);
      float d;
      if (en->shape == VMASB200_SPHERE) {
        d = ray_sphere(p, en->radius, o, dc, ds, max_range);
      } else {
        Rot q = trig_of(E, en);
        if (en->shape == VMASB200_BOX) d = ray_box(p, q.c, q.s, en->length, en->width, o, dc, ds, max_range);
        else d = ray_line(p, q.c, q.s, en->length, o, dc, ds, max_range);
      }
      best = pmin(best, d);
    }
  }
  return best;
}

}  // namespace vb
