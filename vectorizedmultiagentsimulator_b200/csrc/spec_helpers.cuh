// spec_helpers.cuh - register-resident twins of begin_forces / integrate / the sphere-sphere item of step_kernel.cuh
// for the WORLD-COMPILED step kernel (jit.cpp).  This file is not part of libvmasb200.so's own translation unit: it is
// embedded in the library as text and compiled at plan time by NVRTC, between a generated prelude that defines the
// world as compile-time constants (namespace vb::gen: ND, NST, NA, S, PW, SUB_DT, KC ...) and generated straight-line
// calls of the functions below with literal entity attributes.  After inlining every `en.flags & ...` test, every
// index into the state arrays and every table read of the generic kernel is a constant: the state lives in registers
// and only the reference's arithmetic is left.
//
// Every function performs the SAME fp32 operations in the SAME order as its generic twin (cited beside it), so the
// compiled kernel is bit-identical to the generic one (tests/test_gpu_jit.py), which is the one pinned to the oracle.
#pragma once

namespace vb {
namespace gen {

struct EntC {  // the attributes of tables.h DEnt the physics reads, by value (constant-folded)
  uint32_t flags;
  float radius, mass, moi, max_speed, v_range, one_minus_drag, lin_fric, ang_fric, gx, gy, max_f, f_range, max_t, t_range;
};

constexpr int kND = ND > 0 ? ND : 1, kNST = NST > 0 ? NST : 1, kNA = NA > 0 ? NA : 1, kPW = PW > 0 ? PW : 1;

struct St {
  float px[kND], py[kND], vx[kND], vy[kND], rot[kND], w[kND];  // dynamic entities (dyn buffer rows)
  float sx[kNST], sy[kNST], srot[kNST];                       // static entities (stat buffer rows)
  float fx[kNA], fy[kNA], tq[kNA];                            // agent.state.force / torque, clamped in place
  float Fx[kND], Fy[kND], T[kND];                             // accumulators of the running substep
  uint32_t verd[kPW];                                         // World.collides verdict bits of the running substep
};

// step_kernel.cuh begin_forces (core.py:2017-2101): d = dynamic slot, a = agent slot
template <int d, int a>
VB_DI void begin_r(const EntC en, St& st) {
  const uint32_t fl = en.flags;
  V2 F = mk(0.f, 0.f);
  float T = 0.f;
  if (fl & VMASB200_E_AGENT) {
    constexpr int aa = a >= 0 ? a : 0;
    if (fl & VMASB200_E_MOVABLE) {
      V2 f = mk(st.fx[aa], st.fy[aa]);
      if (fl & VMASB200_E_MAX_F) f = clamp_norm2(f, en.max_f);
      if (fl & VMASB200_E_F_RANGE) {
        float r = en.f_range;
        f = mk(clampf(f.x, -r, r), clampf(f.y, -r, r));
      }
      st.fx[aa] = f.x; st.fy[aa] = f.y;
      F = F + f;
    }
    if (fl & VMASB200_E_ROTATABLE) {
      float t = st.tq[aa];
      if (fl & VMASB200_E_MAX_T) t = clamp_norm1(t, en.max_t);
      if (fl & VMASB200_E_T_RANGE) t = clampf(t, -en.t_range, en.t_range);
      st.tq[aa] = t;
      T = T + t;
    }
  }
  if (fl & DE_LINF_ON) F = F + friction2(mk(st.vx[d], st.vy[d]), en.lin_fric, en.mass, SUB_DT);
  if (fl & DE_ANGF_ON) T = T + friction1(st.w[d], en.ang_fric, en.moi, SUB_DT);
  if (fl & VMASB200_E_MOVABLE) {
    if (HAS_WGRAV) F = F + mk(en.mass * WGX, en.mass * WGY);
    if (fl & VMASB200_E_GRAVITY) F = F + mk(en.mass * en.gx, en.mass * en.gy);
  }
  st.Fx[d] = F.x; st.Fy[d] = F.y; st.T[d] = T;
}

// step_kernel.cuh integrate (core.py:2861-2907)
template <int d>
VB_DI void integrate_r(const EntC en, St& st, int substep) {
  const uint32_t fl = en.flags;
  if (fl & VMASB200_E_MOVABLE) {
    V2 F = mk(st.Fx[d], st.Fy[d]);
    V2 vel = mk(st.vx[d], st.vy[d]);
    if (substep == 0) vel = vel * en.one_minus_drag;
    const float mass = en.mass;
    V2 acc = mass == 1.f ? F : F / mass;  // x / 1 == x exactly
    vel = vel + acc * SUB_DT;
    if (fl & VMASB200_E_MAX_SPEED) vel = clamp_norm2(vel, en.max_speed);
    if (fl & VMASB200_E_V_RANGE) {
      float r = en.v_range;
      vel = mk(clampf(vel.x, -r, r), clampf(vel.y, -r, r));
    }
    V2 np = mk(st.px[d], st.py[d]) + vel * SUB_DT;
    if (HAS_XS) np.x = clampf(np.x, -XS, XS);
    if (HAS_YS) np.y = clampf(np.y, -YS, YS);
    st.vx[d] = vel.x; st.vy[d] = vel.y; st.px[d] = np.x; st.py[d] = np.y;
  }
  if (fl & VMASB200_E_ROTATABLE) {
    float w = st.w[d];
    if (substep == 0) w = w * en.one_minus_drag;
    w = w + (st.T[d] / en.moi) * SUB_DT;
    st.w[d] = w;
    st.rot[d] = st.rot[d] + w * SUB_DT;
  }
}

// One sphere-sphere candidate pair of a substep: the broad-phase row (kernels.cuh run_substeps: verdict of
// World.collides, core.py:2796-2800, and the exact cull dist > r_a + r_b), eval_item IT_SS (core.py:2293-2338) and
// accumulate (core.py:2190-2198).  da / db: dynamic slot of a / b or -1 (static: no accumulator); ACC: tables.h ACC_*.
template <int ACC, int da, int db, int word, int bit>
VB_DI void pair_ss_r(St& st, V2 pa, V2 pb, float t2_bp, float t2_cull, float ra, float rb) {
  const float dx = pa.x - pb.x, dy = pa.y - pb.y;
  const float d2 = fmaf(dy, dy, dx * dx);
  if (d2 <= t2_bp) st.verd[word] |= 1u << bit;
  if (!(d2 > t2_cull)) {  // NaN passes, as in the generic broad phase
    const V2 fa = constraint_force(pa, pb, ra + rb, KC, MARGIN, false);
    constexpr int ia = da >= 0 ? da : 0, ib = db >= 0 ? db : 0;
    if (ACC & ACC_A_MOV) { st.Fx[ia] = st.Fx[ia] + fa.x; st.Fy[ia] = st.Fy[ia] + fa.y; }
    if (ACC & ACC_A_ROT) st.T[ia] = st.T[ia] + 0.f;
    if (ACC & ACC_B_MOV) { st.Fx[ib] = st.Fx[ib] + (-fa.x); st.Fy[ib] = st.Fy[ib] + (-fa.y); }
    if (ACC & ACC_B_ROT) st.T[ib] = st.T[ib] + 0.f;
  }
}

}  // namespace gen
}  // namespace vb
