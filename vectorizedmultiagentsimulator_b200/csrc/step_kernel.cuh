// step_kernel.cuh - the fused World.step kernel (sm_100a).
//
// One CTA owns a tile of TE consecutive environments.  The tile's env records (dyn + stat) are staged into
// shared memory by ONE cp.async.bulk (TMA 1-D) each, stay resident across all substeps, and are written back by
// one bulk store per output record type.  G lanes (a sub-warp group, G in {1,2,4,8,16,32}) cooperate on one
// environment; per substep they run three phases separated by warp barriers:
//   A  broad phase: lanes stride over the candidate pairs (tables.h DBroad), take the reference's own verdict
//      (core.py:2796-2800) sqrt-free, drop every pair whose force is provably zero, and compact the survivors - by
//      warp ballot, in item order - into the env's active list (behind the joint constraints, always active);
//   B  narrow phase: lanes stride over the ACTIVE LIST only and write each item's (force on a, torque a, torque b);
//   C  lanes stride over the dynamic entities, add the active items up IN THE REFERENCE'S ORDER (core.py:2175-2198;
//      a dropped item would have added exact zeros), integrate (core.py:2861-2907), refresh cos/sin of a rotated
//      entity and start the next substep's force accumulator (action / friction / gravity, core.py:1994-2005);
// and after the last substep the LIDAR epilogue (core.py:1661-1785): per sensor a mask of the targets within
// reach, then lanes stride over the rays and cast only against those.
#pragma once
#include "geom.cuh"
#include "tables.h"

namespace vb {

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

// ---------------------------------------------------------------------------------------- per-env accessors
struct EnvPtrs {
  float* dyn;         // this env's dyn record in smem
  const float* stat;  // this env's stat record in smem
  float* work;        // this env's work area in smem
};

VB_DI void pose(const DEnt* __restrict__ en, const EnvPtrs& ep, V2& p, float& rot) {
  int d = en->dyn_off;
  if (d >= 0) {
    p = mk(ep.dyn[d], ep.dyn[d + 1]);
    rot = ep.dyn[d + 4];
  } else {
    int s = en->stat_off;
    p = mk(ep.stat[s], ep.stat[s + 1]);
    rot = ep.stat[s + 2];
  }
}
// encoded position offset of the broad-phase table: >= 0 dyn record, < 0 ~offset into the stat record
VB_DI V2 pos_enc(const EnvPtrs& ep, int off) {
  const float* q = off >= 0 ? ep.dyn + off : ep.stat + ~off;
  return mk(q[0], q[1]);
}
VB_DI V2 posof(const DEnt* __restrict__ en, const EnvPtrs& ep) {
  int d = en->dyn_off;
  if (d >= 0) return mk(ep.dyn[d], ep.dyn[d + 1]);
  int s = en->stat_off;
  return mk(ep.stat[s], ep.stat[s + 1]);
}
VB_DI Rot rotof(const KParams& kp, const DEnt* __restrict__ en, const EnvPtrs& ep) {
  const float* r = ep.work + kp.off_rot + 4 * en->rot_slot;
  Rot o;
  o.c = r[0]; o.s = r[1]; o.c2 = r[2]; o.s2 = r[3];
  return o;
}

// cos/sin cache of one entity (physics.py:299-301, 412; utils.py:183-184)
VB_DI void store_trig(const KParams& kp, const DEnt* __restrict__ en, const EnvPtrs& ep, float rot) {
  float* r = ep.work + kp.off_rot + 4 * en->rot_slot;
  float s, c;
  sincosf(rot, &s, &c);
  r[0] = c; r[1] = s;
  if (en->shape == VMASB200_BOX) {
    float rot2 = rot + (float)(3.14159265358979323846 / 2);
    sincosf(rot2, &s, &c);
    r[2] = c; r[3] = s;
  }
}

// Start of a substep for one dynamic entity: F = T = 0, then action force/torque (core.py:2017-2040, with the
// clamped value written back to the env's force slots), friction (core.py:2053-2101), gravity (core.py:2042-2051).
VB_DI void begin_forces(const KParams& kp, const DEnt* __restrict__ en, const EnvPtrs& ep) {
  const uint32_t fl = en->flags;
  V2 F = mk(0.f, 0.f);
  float T = 0.f;
  const int d = en->dyn_off;
  if (fl & VMASB200_E_AGENT) {
    float* fr = ep.work + kp.off_frc + 3 * en->agent;
    if (fl & VMASB200_E_MOVABLE) {
      V2 f = mk(fr[0], fr[1]);
      if (fl & VMASB200_E_MAX_F) f = clamp_norm2(f, en->max_f);
      if (fl & VMASB200_E_F_RANGE) {
        float r = en->f_range;
        f = mk(clampf(f.x, -r, r), clampf(f.y, -r, r));
      }
      fr[0] = f.x; fr[1] = f.y;
      F = F + f;
    }
    if (fl & VMASB200_E_ROTATABLE) {
      float t = fr[2];
      if (fl & VMASB200_E_MAX_T) t = clamp_norm1(t, en->max_t);
      if (fl & VMASB200_E_T_RANGE) t = clampf(t, -en->t_range, en->t_range);
      fr[2] = t;
      T = T + t;
    }
  }
  if (fl & DE_LINF_ON) F = F + friction2(mk(ep.dyn[d + 2], ep.dyn[d + 3]), en->lin_fric, en->mass, kp.sub_dt);
  if (fl & DE_ANGF_ON) T = T + friction1(ep.dyn[d + 5], en->ang_fric, en->moi, kp.sub_dt);
  if (fl & VMASB200_E_MOVABLE) {
    if (kp.has_wgrav) F = F + mk(en->mass * kp.wgx, en->mass * kp.wgy);
    if (fl & VMASB200_E_GRAVITY) F = F + mk(en->mass * en->gx, en->mass * en->gy);
  }
  float* fin = ep.work + kp.off_fin + en->fin_off;
  fin[0] = F.x; fin[1] = F.y; fin[2] = T;
}

// One work item (joint constraint or candidate pair): force on a (the force on b is its exact negation) and the
// two torques.
VB_DI void eval_item(const KParams& kp, const DItem* __restrict__ it, const EnvPtrs& ep, V2& fa, float& ta, float& tb) {
  const int type = it->type;
  const DEnt* __restrict__ ea = kp.ents + it->a;
  const DEnt* __restrict__ eb = kp.ents + it->b;
  V2 pa, pb;
  float ra, rb;
  pose(ea, ep, pa, ra);
  pose(eb, ep, pb, rb);
  const float Kc = kp.Kc, k = kp.margin;
  ta = 0.f; tb = 0.f;
  if (type == IT_JOINT) {  // core.py:2200-2291, joints.py:206-215
    Rot qa = rotof(kp, ea, ep), qb = rotof(kp, eb, ep);
    V2 ja = pa + rotv(mk(it->dax, it->day), qa.c, qa.s);
    V2 jb = pb + rotv(mk(it->dbx, it->dby), qb.c, qb.s);
    V2 att = constraint_force(ja, jb, it->dist, kp.Kj, k, true);
    V2 rep = constraint_force(ja, jb, it->dist, kp.Kj, k, false);
    fa = att + rep;
    V2 fb = (-att) + (-rep);
    ta = cross(ja - pa, fa);
    tb = cross(jb - pb, fb);
    if (!it->rotate) {
      float fr = it->fr_off >= 0 ? ep.stat[it->fr_off] : it->fixed_rot;
      float tq = constraint_torque(ra, rb + fr, kp.Kt);
      ta = ta + (-tq);
      tb = tb + tq;
    }
    return;
  }
  switch (type) {
    case IT_SS: {  // core.py:2293
      fa = constraint_force(pa, pb, ea->radius + eb->radius, Kc, k, false);
    } break;
    case IT_LS: {  // core.py:2340; a = line, b = sphere
      Rot qa = rotof(kp, ea, ep);
      V2 cp = closest_point_line(pa, qa.c, qa.s, ea->length, pb);
      V2 fs = constraint_force(pb, cp, eb->radius + kLMD, Kc, k, false);
      fa = -fs;
      ta = cross(cp - pa, fa);
    } break;
    case IT_LL: {  // core.py:2393
      Rot qa = rotof(kp, ea, ep), qb = rotof(kp, eb, ep);
      V2 ca, cb;
      closest_points_line_line(pa, qa.c, qa.s, ea->length, pb, qb.c, qb.s, eb->length, ca, cb);
      fa = constraint_force(ca, cb, kLMD, Kc, k, false);
      ta = cross(ca - pa, fa);
      tb = cross(cb - pb, -fa);
    } break;
    case IT_BS: {  // core.py:2458; a = box, b = sphere
      Rot qa = rotof(kp, ea, ep);
      V2 cp = closest_point_box(pa, qa, ea->width, ea->length, pb);
      V2 ip = cp;
      float dd = 0.f;
      if (!(ea->flags & VMASB200_E_HOLLOW)) ip = inner_point_box(pb, cp, pa, dd);
      V2 fs = constraint_force(pb, ip, (eb->radius + kLMD) + dd, Kc, k, false);
      fa = -fs;
      ta = cross(cp - pa, fa);
    } break;
    case IT_BL: {  // core.py:2553; a = box, b = line
      Rot qa = rotof(kp, ea, ep), qb = rotof(kp, eb, ep);
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
      Rot qa = rotof(kp, ea, ep), qb = rotof(kp, eb, ep);
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

// Accumulate the active items of one dynamic entity in reference order (core.py:2190-2198; the list is in item
// order) and integrate (core.py:2861-2907).
VB_DI void finish_entity(const KParams& kp, const DEnt* __restrict__ en, int e, const EnvPtrs& ep, int substep,
                         const int* __restrict__ list, int count) {
  const uint32_t fl = en->flags;
  const float* fin = ep.work + kp.off_fin + en->fin_off;
  V2 F = mk(fin[0], fin[1]);
  float T = fin[2];
  const float4* res = reinterpret_cast<const float4*>(ep.work + kp.off_res);
  const bool mov = fl & VMASB200_E_MOVABLE, rotb = fl & VMASB200_E_ROTATABLE;
  constexpr int em = (1 << kEntryEntBits) - 1;
  for (int k = 0; k < count; ++k) {
    const int entry = list[k];
    const int a = (entry >> kEntryItemBits) & em, b = (entry >> (kEntryItemBits + kEntryEntBits)) & em;
    if (a == e) {
      float4 r = res[k];
      if (mov) F = F + mk(r.x, r.y);
      if (rotb) T = T + r.z;
    } else if (b == e) {
      float4 r = res[k];
      if (mov) F = F + mk(-r.x, -r.y);
      if (rotb) T = T + r.w;
    }
  }
  float* st = ep.dyn + en->dyn_off;
  const float sub_dt = kp.sub_dt;
  if (mov) {
    V2 vel = mk(st[2], st[3]);
    if (substep == 0) vel = vel * en->one_minus_drag;
    V2 acc = F / en->mass;
    vel = vel + acc * sub_dt;
    if (fl & VMASB200_E_MAX_SPEED) vel = clamp_norm2(vel, en->max_speed);
    if (fl & VMASB200_E_V_RANGE) {
      float r = en->v_range;
      vel = mk(clampf(vel.x, -r, r), clampf(vel.y, -r, r));
    }
    V2 np = mk(st[0], st[1]) + vel * sub_dt;
    if (kp.has_xs) np.x = clampf(np.x, -kp.xs, kp.xs);
    if (kp.has_ys) np.y = clampf(np.y, -kp.ys, kp.ys);
    st[2] = vel.x; st[3] = vel.y; st[0] = np.x; st[1] = np.y;
  }
  if (rotb) {
    float w = st[5];
    if (substep == 0) w = w * en->one_minus_drag;
    w = w + (T / en->moi) * sub_dt;
    st[5] = w;
    float nr = st[4] + w * sub_dt;
    st[4] = nr;
    if (en->rot_slot >= 0) store_trig(kp, en, ep, nr);
  }
}

// One LIDAR ray (sensors.py:115-118 -> core.py:1661-1785): min over the sensor's targets within reach (`lm`: the
// sensor's survival mask words; an unreachable target would have contributed max_range).
VB_DI float cast_one_ray(const KParams& kp, const DSensor* __restrict__ sn, const EnvPtrs& ep, V2 o, float ang,
                         const uint32_t* __restrict__ lm) {
  const float max_range = sn->max_range;
  float best = max_range;
  const int ntw = sn->n_tw;
  uint32_t any = 0u;
  for (int w = 0; w < ntw; ++w) any |= lm[w];
  if (!any) return best;
  float ds, dc;
  sincosf(ang, &ds, &dc);
  for (int w = 0; w < ntw; ++w) {
    uint32_t m = lm[w];
    const int tb = sn->tg_begin + 32 * w;
    while (m) {
      const int t = __ffs(m) - 1;
      m &= m - 1;
      const DEnt* __restrict__ en = kp.ents + __ldg(kp.targets + tb + t);
      V2 p = posof(en, ep);
      float d;
      if (en->shape == VMASB200_SPHERE) {
        d = ray_sphere(p, en->radius, o, dc, ds, max_range);
      } else {
        Rot q = rotof(kp, en, ep);
        if (en->shape == VMASB200_BOX) d = ray_box(p, q.c, q.s, en->length, en->width, o, dc, ds, max_range);
        else d = ray_line(p, q.c, q.s, en->length, o, dc, ds, max_range);
      }
      best = pmin(best, d);
    }
  }
  return best;
}

}  // namespace vb
