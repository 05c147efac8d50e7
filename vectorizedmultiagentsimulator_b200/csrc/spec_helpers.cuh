// spec_helpers.cuh - register-resident twins of begin_forces / integrate / eval_item / accumulate of step_kernel.cuh
// for the WORLD-COMPILED step kernel (jit.h).  This file is not part of libvmasb200.so's own translation unit: it is
// embedded in the library as text and compiled at plan time by NVRTC, between a generated prelude that defines the
// world as compile-time constants (namespace vb::gen: ND, NST, NA, NTR, NJF, S, PW, SUB_DT, KC ...) and generated
// straight-line calls of the functions below with literal entity attributes.  After inlining, every `en.flags & ...`
// test, every index into the state arrays and every table read of the generic kernel is a constant: the state lives
// in registers and only the reference's arithmetic is left.
//
// Every function performs the SAME fp32 operations in the SAME order as its generic twin (cited beside it), so the
// compiled kernel is bit-identical to the generic one (tests/test_gpu_jit.py), which is the one pinned to the oracle.
#pragma once

namespace vb {
namespace gen {

struct EntC {  // the attributes of tables.h DEnt the physics reads, by value (constant-folded)
  int shape;
  uint32_t flags;
  float radius, length, width, mass, moi, max_speed, v_range, one_minus_drag, lin_fric, ang_fric, gx, gy, max_f, f_range,
      max_t, t_range;
};
struct JointC {  // tables.h DItem of a joint
  float dax, day, dbx, dby, dist, fixed_rot;
  int rotate;
};

constexpr int kND = ND > 0 ? ND : 1, kNST = NST > 0 ? NST : 1, kNA = NA > 0 ? NA : 1, kPW = PW > 0 ? PW : 1;
constexpr int kNTR = NTR > 0 ? NTR : 1, kNJF = NJF > 0 ? NJF : 1;

// PARK (BIG worlds): what the pair phase of a substep does not read - velocities, rotations, the agents' forces - and the
// (Fx Fy T) accumulators (which the table-driven narrow phase adds to in the tile anyway) live in the thread's column of
// the shared-memory tile instead of registers: football's kernel needed 197 registers (two warps per scheduler), and the
// scheduler's register file holds three at 168.  Same syntax as the register arrays (st.vx[d]), literal rows after inlining.
template <int STRIDE>
struct TileArr {
  float* p;  // row of element 0 in this thread's column
  VB_DI float& operator[](int i) const { return p[i * (STRIDE * NT)]; }
};
template <bool B, class X, class Y> struct SelT { using type = X; };
template <class X, class Y> struct SelT<false, X, Y> { using type = Y; };
template <int STRIDE> VB_DI void park(TileArr<STRIDE>& a, float* col, int row) { a.p = col + row * NT; }
template <int N> VB_DI void park(float (&)[N], float*, int) {}
using DynArr = SelT<PARK, TileArr<1>, float[kND]>::type;
using AgtArr = SelT<PARK, TileArr<1>, float[kNA]>::type;
using AccArr = SelT<PARK, TileArr<3>, float[kND]>::type;

struct St {
  float px[kND], py[kND];                                     // dynamic entities (dyn buffer rows)
  DynArr vx, vy, rot, w;
  float sx[kNST], sy[kNST], srot[kNST];                       // static entities (stat buffer rows)
  float jfr[kNJF];                                            // per-env fixed rotations of joints (stat buffer rows)
  AgtArr fx, fy, tq;                                          // agent.state.force / torque, clamped in place
  AccArr Fx, Fy, T;                                           // accumulators of the running substep
  float tc[kNTR], ts[kNTR], tc2[kNTR], ts2[kNTR];             // cos/sin cache (step_kernel.cuh store_trig)
  uint32_t verd[kPW];                                         // World.collides verdict bits of the running substep
  uint32_t thin[kPW];                                         // thin events of the running substep (kernels.cuh)
  uint32_t sv[kPW];                                           // BIG worlds: survivors of the broad phase of the running substep
  float* ss;                                                  // BIG worlds: this thread's column of the static tile (row r at ss[r * NT])
  float* coop;                                                // BIG worlds: this WARP's scratch of the cooperative narrow phase (narrow_coop)
};

// BIG worlds (jit.h jit_big_world): cos/sin cache of a STATIC entity in the static tile, rows r .. r + 1 (+ 2, 3 for a box)
template <int r, int BOX>
VB_DI Rot rot_s(const St& st) {
  Rot o;
  o.c = st.ss[r * NT]; o.s = st.ss[(r + 1) * NT];
  if (BOX) { o.c2 = st.ss[(r + 2) * NT]; o.s2 = st.ss[(r + 3) * NT]; }
  else { o.c2 = 0.f; o.s2 = 0.f; }
  return o;
}
template <int r, int BOX>
VB_DI void trig_s(St& st, float rot) {
  float s, c;
  sincosf(rot, &s, &c);
  st.ss[r * NT] = c; st.ss[(r + 1) * NT] = s;
  if (BOX) {
    float rot2 = rot + (float)(3.14159265358979323846 / 2);
    sincosf(rot2, &s, &c);
    st.ss[(r + 2) * NT] = c; st.ss[(r + 3) * NT] = s;
  }
}

// ... the same with the tile address at run time (looped runs of static entities, jit.h): p = the entity's cos row in this
// thread's column
template <int BOX>
VB_DI void trig_sp(float* p, float rot) {
  float s, c;
  sincosf(rot, &s, &c);
  p[0] = c; p[NT] = s;
  if (BOX) {
    float rot2 = rot + (float)(3.14159265358979323846 / 2);
    sincosf(rot2, &s, &c);
    p[2 * NT] = c; p[3 * NT] = s;
  }
}

template <int t>
VB_DI Rot rot_r(const St& st) {
  Rot o;
  constexpr int k = t >= 0 ? t : 0;
  o.c = st.tc[k]; o.s = st.ts[k]; o.c2 = st.tc2[k]; o.s2 = st.ts2[k];
  return o;
}
// step_kernel.cuh store_trig (physics.py:299-301, 412; utils.py:183-184)
template <int t, int SHAPE>
VB_DI void trig_r(St& st, float rot) {
  if (t < 0) return;
  constexpr int k = t >= 0 ? t : 0;
  float s, c;
  sincosf(rot, &s, &c);
  st.tc[k] = c; st.ts[k] = s;
  if (SHAPE == VMASB200_BOX) {
    float rot2 = rot + (float)(3.14159265358979323846 / 2);
    sincosf(rot2, &s, &c);
    st.tc2[k] = c; st.ts2[k] = s;
  }
}

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

// step_kernel.cuh integrate (core.py:2861-2907); t = trig slot of the entity or -1
template <int d, int t>
VB_DI void integrate_r(const EntC en, St& st, int substep) {
  const uint32_t fl = en.flags;
  if (fl & VMASB200_E_MOVABLE) {
    V2 F = mk(st.Fx[d], st.Fy[d]);
    V2 vel = mk(st.vx[d], st.vy[d]);
    if (substep == 0) vel = vel * en.one_minus_drag;
    const float mass = en.mass;
    V2 acc = mass == 1.f ? F : mk(div_pos(F.x, mass), div_pos(F.y, mass));  // x / 1 == x exactly
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
    w = w + div_pos(st.T[d], en.moi) * SUB_DT;
    st.w[d] = w;
    const float nr = st.rot[d] + w * SUB_DT;
    st.rot[d] = nr;
    trig_r<t, VMASB200_SPHERE>(st, nr);  // the box branch below needs the shape at compile time
  }
}
template <int d, int t>
VB_DI void integrate_box_r(const EntC en, St& st, int substep) {  // a rotatable box also refreshes cos/sin(rot + pi/2)
  integrate_r<d, -1>(en, st, substep);
  if (en.flags & VMASB200_E_ROTATABLE) trig_r<t, VMASB200_BOX>(st, st.rot[d >= 0 ? d : 0]);
}

// step_kernel.cuh accumulate = update_env_forces (core.py:2190-2198)
template <int ACC, int da, int db>
VB_DI void accumulate_r(St& st, V2 fa, float ta, float tb) {
  constexpr int ia = da >= 0 ? da : 0, ib = db >= 0 ? db : 0;
  if (ACC & ACC_A_MOV) { st.Fx[ia] = st.Fx[ia] + fa.x; st.Fy[ia] = st.Fy[ia] + fa.y; }
  if (ACC & ACC_A_ROT) st.T[ia] = st.T[ia] + ta;
  if (ACC & ACC_B_MOV) { st.Fx[ib] = st.Fx[ib] + (-fa.x); st.Fy[ib] = st.Fy[ib] + (-fa.y); }
  if (ACC & ACC_B_ROT) st.T[ib] = st.T[ib] + tb;
}

// step_kernel.cuh eval_item for the six pair types (core.py:2293-2785)
template <int TYPE>
VB_DI void eval_pair_r(const EntC ea, const EntC eb, V2 pa, V2 pb, const Rot qa, const Rot qb, V2& fa, float& ta, float& tb) {
  ta = 0.f; tb = 0.f;
  if (TYPE == IT_SS) {
    fa = constraint_force(pa, pb, ea.radius + eb.radius, KC, MARGIN, false);
  } else if (TYPE == IT_LS) {
    V2 cp = closest_point_line(pa, qa.c, qa.s, ea.length, pb);
    V2 fs = constraint_force(pb, cp, eb.radius + kLMD, KC, MARGIN, false);
    fa = -fs;
    ta = cross(cp - pa, fa);
  } else if (TYPE == IT_LL) {
    V2 ca, cb;
    closest_points_line_line(pa, qa.c, qa.s, ea.length, pb, qb.c, qb.s, eb.length, ca, cb);
    fa = constraint_force(ca, cb, kLMD, KC, MARGIN, false);
    ta = cross(ca - pa, fa);
    tb = cross(cb - pb, -fa);
  } else if (TYPE == IT_BS) {
    V2 cp = closest_point_box(pa, qa, ea.width, ea.length, pb);
    V2 ip = cp;
    float dd = 0.f;
    if (!(ea.flags & VMASB200_E_HOLLOW)) ip = inner_point_box(pb, cp, pa, dd);
    V2 fs = constraint_force(pb, ip, (eb.radius + kLMD) + dd, KC, MARGIN, false);
    fa = -fs;
    ta = cross(cp - pa, fa);
  } else if (TYPE == IT_BL) {
    V2 cbx, cl;
    closest_line_box(pa, qa, ea.width, ea.length, pb, qb.c, qb.s, eb.length, cbx, cl);
    V2 ip = cbx;
    float dd = 0.f;
    if (!(ea.flags & VMASB200_E_HOLLOW)) ip = inner_point_box(cl, cbx, pa, dd);
    fa = constraint_force(ip, cl, kLMD + dd, KC, MARGIN, false);
    ta = cross(cbx - pa, fa);
    tb = cross(cl - pb, -fa);
  } else {
    V2 ca, cb;
    closest_box_box(pa, qa, ea.width, ea.length, pb, qb, eb.width, eb.length, ca, cb);
    V2 ia = ca, ib = cb;
    float da = 0.f, db = 0.f;
    if (!(ea.flags & VMASB200_E_HOLLOW)) ia = inner_point_box(cb, ca, pa, da);
    if (!(eb.flags & VMASB200_E_HOLLOW)) ib = inner_point_box(ca, cb, pb, db);
    fa = constraint_force(ia, ib, (da + db) + kLMD, KC, MARGIN, false);
    ta = cross(ca - pa, fa);
    tb = cross(cb - pb, -fa);
  }
}
// box-sphere narrow phase out of line: one copy of the closest-point code per kernel keeps the unrolled pair list
// compact (joint_passage: 39 such pairs); only the varying operands travel
__device__ __noinline__ void eval_bs_r(float width, float length, bool hollow, float rb, V2 pa, V2 pb, const Rot qa, V2& fa,
                                       float& ta) {
  V2 cp = closest_point_box(pa, qa, width, length, pb);
  V2 ip = cp;
  float dd = 0.f;
  if (!hollow) ip = inner_point_box(pb, cp, pa, dd);
  V2 fs = constraint_force(pb, ip, (rb + kLMD) + dd, KC, MARGIN, false);
  fa = -fs;
  ta = cross(cp - pa, fa);
}

// the same for sphere-sphere and line-sphere in BIG worlds (football: 21 + 70 such pairs unrolled)
__device__ __noinline__ void eval_ss_r(float rsum, V2 pa, V2 pb, V2& fa) { fa = constraint_force(pa, pb, rsum, KC, MARGIN, false); }
__device__ __noinline__ void eval_ls_r(float length, float rb, V2 pa, V2 pb, float c, float s, V2& fa, float& ta) {
  V2 cp = closest_point_line(pa, c, s, length, pb);
  V2 fs = constraint_force(pb, cp, rb + kLMD, KC, MARGIN, false);
  fa = -fs;
  ta = cross(cp - pa, fa);
}

// One candidate pair of a substep: the broad-phase row (kernels.cuh run_substeps: verdict of World.collides,
// core.py:2796-2800, and the cull that only drops pairs whose force is provably zero), the narrow phase, the thin-event
// test of the speculative launch and update_env_forces.  da / db: dynamic slot of a / b or -1; FRAME: the row carries
// a frame cull (b's centre in a's frame against hx, hy) instead of the circle cull.
template <int TYPE, int ACC, int da, int db, int word, int bit, bool FRAME, bool THIN>
VB_DI void pair_r(St& st, const EntC ea, const EntC eb, V2 pa, V2 pb, const Rot qa, const Rot qb, float t2_bp, float t2_cull,
                  float hx, float hy) {
  const float dx = pa.x - pb.x, dy = pa.y - pb.y;
  const float d2 = fmaf(dy, dy, dx * dx);
  const bool verdict = d2 <= t2_bp;
  if (verdict) st.verd[word] |= 1u << bit;
  bool sv = !(d2 > t2_cull);  // NaN passes, as in the generic broad phase
  if (FRAME) {
    const float lx = fmaf(dx, qa.c, dy * qa.s), ly = fmaf(dx, qa.s, -(dy * qa.c));
    sv = sv && !(fabsf(lx) > hx) && !(fabsf(ly) > hy);
  }
  if (sv) {
    V2 fa;
    float ta, tb;
    if (TYPE == IT_BS) { tb = 0.f; eval_bs_r(ea.width, ea.length, (ea.flags & VMASB200_E_HOLLOW) != 0, eb.radius, pa, pb, qa, fa, ta); }
    else if (BIG && TYPE == IT_SS) { ta = 0.f; tb = 0.f; eval_ss_r(ea.radius + eb.radius, pa, pb, fa); }
    else if (BIG && TYPE == IT_LS) { tb = 0.f; eval_ls_r(ea.length, eb.radius, pa, pb, qa.c, qa.s, fa, ta); }
    else eval_pair_r<TYPE>(ea, eb, pa, pb, qa, qb, fa, ta, tb);
    // THIN = false: a sphere-sphere pair whose fp32 r_a + r_b does not exceed the verdict's threshold cannot be thin
    if (THIN && (fa.x != 0.f || fa.y != 0.f) && !verdict) st.thin[word] |= 1u << bit;
    accumulate_r<ACC, da, db>(st, fa, ta, tb);
  }
}

// the bit of a row as ONE predicated OR behind its compare(s) (the compiler's own choice is a select per bit plus adds to
// merge them: 48 of the 100 instructions of football's loop body)
template <int i>
VB_DI void bit_le(uint32_t& v, float d2, float t) {  // v |= (d2 <= t) << i
  asm("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p or.b32 %0, %0, %3;\n\t}" : "+r"(v) : "f"(d2), "f"(t), "n"(1u << i));
}
template <int i>
VB_DI void bit_ngt(uint32_t& s, float d2, float t) {  // s |= !(d2 > t) << i   (NaN passes, as in the generic broad phase)
  asm("{\n\t.reg .pred p;\n\tsetp.gt.f32 p, %1, %2;\n\t@!p or.b32 %0, %0, %3;\n\t}" : "+r"(s) : "f"(d2), "f"(t), "n"(1u << i));
}
template <int i>
VB_DI void bit_frame(uint32_t& s, float lx, float ly, float hx, float hy) {  // s |= !(|lx| > hx || |ly| > hy) << i
  asm("{\n\t.reg .pred p;\n\t.reg .f32 a, b;\n\tabs.f32 a, %1;\n\tabs.f32 b, %2;\n\tsetp.gt.f32 p, a, %3;\n\tsetp.gt.or.f32 p, b, %4, p;\n\t"
      "@!p or.b32 %0, %0, %5;\n\t}"
      : "+r"(s) : "f"(lx), "f"(ly), "f"(hx), "f"(hy), "n"(1u << i));
}
// BIG worlds: the broad-phase row alone (kernels.cuh run_substeps: the verdict of World.collides and the conservative cull),
// branch-free; the survivors are evaluated afterwards by narrow_big.
template <int word, int bit, bool FRAME>
VB_DI void pair_b(St& st, V2 pa, V2 pb, const Rot qa, float t2_bp, float t2_cull, float hx, float hy) {
  const float dx = pa.x - pb.x, dy = pa.y - pb.y;
  const float d2 = fmaf(dy, dy, dx * dx);
  bit_le<bit>(st.verd[word], d2, t2_bp);
  if (FRAME && t2_cull == CUDART_INF_F) {  // plan.cpp: rows with a frame cull have no circle cull
    const float lx = fmaf(dx, qa.c, dy * qa.s), ly = fmaf(dx, qa.s, -(dy * qa.c));
    bit_frame<bit>(st.sv[word], lx, ly, hx, hy);
  } else {
    bool sv = !(d2 > t2_cull);  // NaN passes, as in the generic broad phase
    if (FRAME) {
      const float lx = fmaf(dx, qa.c, dy * qa.s), ly = fmaf(dx, qa.s, -(dy * qa.c));
      sv = sv && !(fabsf(lx) > hx) && !(fabsf(ly) > hy);
    }
    st.sv[word] |= (sv ? 1u : 0u) << bit;
  }
}
// The same row inside a LOOPED run (jit.h: consecutive pairs of one list of dynamic entities against static entities whose
// tile rows are a fixed stride apart run as a loop over the static entities - the straight-line list of football's 91 pairs
// is 20 KB of code that every warp streams through once per substep, and 42 % of that kernel's stall samples were
// instruction fetches): the static partner's rows and the thresholds are run-time values, the pair's bit is bit i of the
// block's masks.  FRAME rows have no circle cull (plan.cpp: t2_cull = inf), the others no frame cull (hx = hy = inf).
template <int i, bool FRAME>
VB_DI void pair_l(uint32_t& v, uint32_t& s, V2 pa, V2 pb, float qc, float qs, float t2_bp, float t2_cull, float hx, float hy) {
  const float dx = pa.x - pb.x, dy = pa.y - pb.y;
  const float d2 = fmaf(dy, dy, dx * dx);
  bit_le<i>(v, d2, t2_bp);
  if (FRAME) {
    const float lx = fmaf(dx, qc, dy * qs), ly = fmaf(dx, qs, -(dy * qc));
    bit_frame<i>(s, lx, ly, hx, hy);
  } else {
    bit_ngt<i>(s, d2, t2_cull);
  }
}
// Two rows of a block at once (bits i and i + 1; the same static partner `ps`, dynamic partners b0 / b1) on the packed fp32
// pipe of sm_100 (FADD2 / FMUL2 / FFMA2: two IEEE fp32 operations per instruction, each rounded like its scalar twin, so
// the bits are pair_l's).  STAT_A: the static entity is the pair's a (dx = ps - b), else its b (dx = b - ps).
template <int i, bool FRAME, bool STAT_A>
VB_DI void pair_l2(uint32_t& v, uint32_t& s, V2 ps, V2 b0, V2 b1, float qc, float qs, float t2_bp0, float t2_cull0, float hx0, float hy0,
                   float t2_bp1, float t2_cull1, float hx1, float hy1) {
  const float2 sx = make_float2(ps.x, ps.x), sy = make_float2(ps.y, ps.y);
  const float2 bx = make_float2(b0.x, b1.x), by = make_float2(b0.y, b1.y);
  const float2 dx = STAT_A ? __fadd2_rn(sx, make_float2(-bx.x, -bx.y)) : __fadd2_rn(bx, make_float2(-sx.x, -sx.y));
  const float2 dy = STAT_A ? __fadd2_rn(sy, make_float2(-by.x, -by.y)) : __fadd2_rn(by, make_float2(-sy.x, -sy.y));
  const float2 d2 = __ffma2_rn(dy, dy, __fmul2_rn(dx, dx));
  bit_le<i>(v, d2.x, t2_bp0);
  bit_le<i + 1>(v, d2.y, t2_bp1);
  if (FRAME) {
    const float2 c2 = make_float2(qc, qc), s2 = make_float2(qs, qs);
    const float2 lx = __ffma2_rn(dx, c2, __fmul2_rn(dy, s2));
    const float2 ly = __ffma2_rn(dx, s2, __fmul2_rn(make_float2(-dy.x, -dy.y), c2));  // (-dy) * c == -(dy * c)
    bit_frame<i>(s, lx.x, ly.x, hx0, hy0);
    bit_frame<i + 1>(s, lx.y, ly.y, hx1, hy1);
  } else {
    bit_ngt<i>(s, d2.x, t2_cull0);
    bit_ngt<i + 1>(s, d2.y, t2_cull1);
  }
}
// shift a block's k bits into a run's W-word accumulator (blocks are visited last to first: block 0 ends at bit 0)
template <int W, int k>
VB_DI void run_push(uint32_t* a, uint32_t bits) {
#pragma unroll
  for (int w = W - 1; w > 0; --w) a[w] = __funnelshift_l(a[w - 1], a[w], k);
  a[0] = (a[0] << k) | bits;
}
// the accumulator's bits [32 * tw - p0, +32) : what lands in word tw of the pair masks when the run starts at pair p0
template <int W, int p0, int tw>
VB_DI uint32_t run_word(const uint32_t* a) {
  uint32_t o = 0u;
#pragma unroll
  for (int q = 0; q < W; ++q) {
    const int rel = p0 + 32 * q - 32 * tw;
    if (rel >= 0 && rel < 32) o |= a[q] << rel;
    else if (rel < 0 && rel > -32) o |= a[q] >> (-rel);
  }
  return o;
}

// ... and their narrow phase: every lane walks its own survivors in pair order through the GENERIC eval_item / accumulate
// (step_kernel.cuh: same code, hence same bits) on its column of the tile, with the tables of jit.h (tile rows, global
// memory).  Lanes with contacts on different pairs run together; only the pair TYPE can diverge.
//
// narrow_coop: the same, shared out over the WARP.  A lane's survivors cost max-over-lanes iterations of the walk above
// while most lanes have none (joint_passage: 1.7 of 32 lanes active in the narrow phase, 40 % of the kernel's
// instructions).  So the warp's survivors are numbered (prefix sum of the lanes' counts), written to a queue in shared
// memory as (lane, pair), evaluated one per lane on the OWNER's column of the tile (the state of every env of the warp is
// there), and the results (fa, ta, tb) go back through shared memory; every lane then applies ITS OWN results in its pair
// order - the accumulation order per entity is the reference's (and narrow_big's), so the bits do not change.
constexpr int kCoopQ = 64;                               // queue entries per pass (more survivors: several passes)
constexpr int kCoopWords = kCoopQ / 2 + 4 * kCoopQ;      // 4-byte words per warp: u16 queue, then float4 results
VB_DI void narrow_coop(const KParams& kp, St& st, const int cnt) {
  const int lane = threadIdx.x & 31;
  unsigned short* q = reinterpret_cast<unsigned short*>(st.coop);
  float4* res = reinterpret_cast<float4*>(st.coop + kCoopQ / 2);
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += up;
  }
  const int off = incl - cnt, T = __shfl_sync(0xffffffffu, incl, 31);
  if (T == 0) return;
  float* col0 = st.ss - lane;  // column of lane 0 of this warp
  EnvT<NT> Eo;
  Eo.env = st.ss; Eo.work = st.ss;
  for (int base = 0; base < T; base += kCoopQ) {
    {
      int g = off - base;
#pragma unroll
      for (int w = 0; w < PW; ++w) {
        uint32_t m = st.sv[w];
        while (m) {
          const int b = __ffs(m) - 1;
          m &= m - 1;
          if (g >= 0 && g < kCoopQ) q[g] = (unsigned short)(lane | (w * 32 + b) << 5);
          ++g;
        }
      }
    }
    __syncwarp();
    const int n = min(kCoopQ, T - base);
    for (int e = lane; e < n; e += 32) {
      const int ent = q[e];
      const DItem* it = kp.items + (NJ + (ent >> 5));
      if (COOP_EDGES && it->type == IT_BL) continue;  // below, four lanes per entry
      EnvT<NT> E;
      E.env = col0 + (ent & 31); E.work = E.env;
      V2 fa;
      float ta, tb;
      eval_item<false>(kp, it, E, fa, ta, tb);
      res[e] = make_float4(fa.x, fa.y, ta, tb);
    }
    if (COOP_EDGES) {
      // box-line entries (eval_item IT_BL: closest_line_box = 4 box edges x a segment-segment closest-point search, ~1500
      // instructions on ONE lane while the warp's other lanes idle: joint_passage's bar against the wall boxes): one lane per
      // box EDGE, the four candidates folded in edge order exactly like closest_line_box's loop, the rest of the item by all
      // four lanes alike.  Same functions, same operations per result: the bits do not change.
      for (int e0 = 0; e0 < n; e0 += 32) {
        const int e = e0 + lane;
        const bool isbl = e < n && kp.items[NJ + (q[e < n ? e : 0] >> 5)].type == IT_BL;
        uint32_t bm = __ballot_sync(0xffffffffu, isbl);
        while (bm) {  // eight entries per round (warp-uniform loop)
          const uint32_t sel = __fns(bm, 0, (lane >> 2) + 1);
          const bool have = sel != 0xffffffffu;
          const int ee = e0 + (have ? (int)sel : 0);
          const int ent = q[have ? ee : 0];
          const DItem* it = kp.items + (NJ + (ent >> 5));
          EnvT<NT> E;
          E.env = col0 + (ent & 31); E.work = E.env;
          const DEnt* ea = &kp.ents[it->a];
          const DEnt* eb = &kp.ents[it->b];
          const V2 pa = pos_at(E, ea->pos_row), pb = pos_at(E, eb->pos_row);
          const Rot qa = trig_of(E, ea), qb = trig_of(E, eb);
          V2 p_box = vinf(), p_line = vinf();
          float d = CUDART_INF_F;
          if (have) {
            const Edge ed = box_edge(lane & 3, pa, qa, ea->width, ea->length);
            closest_points_line_line(ed.p, ed.c, ed.s, ed.len, pb, qb.c, qb.s, eb->length, p_box, p_line);
            d = norm(p_box - p_line);
          }
          V2 b1 = vinf(), b2 = vinf();
          float dist = CUDART_INF_F;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int src = (lane & ~3) + j;
            const float dj = __shfl_sync(0xffffffffu, d, src);
            const float bx = __shfl_sync(0xffffffffu, p_box.x, src), by = __shfl_sync(0xffffffffu, p_box.y, src);
            const float lx = __shfl_sync(0xffffffffu, p_line.x, src), ly = __shfl_sync(0xffffffffu, p_line.y, src);
            if (dj < dist) { b1 = mk(bx, by); b2 = mk(lx, ly); dist = dj; }
          }
          if (have) {
            const V2 cbx = b1, cl = b2;
            V2 ip = cbx;
            float dd = 0.f;
            if (!(ea->flags & VMASB200_E_HOLLOW)) ip = inner_point_box(cl, cbx, pa, dd);
            const V2 fa = constraint_force(ip, cl, kLMD + dd, kp.Kc, kp.margin, false);
            const float ta = cross(cbx - pa, fa), tb = cross(cl - pb, -fa);
            if ((lane & 3) == 0) res[ee] = make_float4(fa.x, fa.y, ta, tb);
          }
#pragma unroll
          for (int k8 = 0; k8 < 8; ++k8) bm &= bm - 1;
        }
      }
    }
    __syncwarp();
    {
      int g = off - base;
#pragma unroll
      for (int w = 0; w < PW; ++w) {
        uint32_t m = st.sv[w];
        while (m) {
          const int b = __ffs(m) - 1;
          m &= m - 1;
          if (g >= 0 && g < kCoopQ) {
            const float4 r = res[g];
            if (THIN_POSSIBLE && (r.x != 0.f || r.y != 0.f) && !((st.verd[w] >> b) & 1u)) st.thin[w] |= 1u << b;
            accumulate<1>(kp.items + (NJ + w * 32 + b), Eo, mk(r.x, r.y), r.z, r.w, 0);
          }
          ++g;
        }
      }
    }
    __syncwarp();
  }
}
VB_DI void narrow_big(const KParams& kp, St& st) {
  if (COOP) {
    // (Worlds with box-box pairs do not share out - jit.h: when every lane has many survivors (waterfall's piled boxes, ~10
    // each) 32 consecutive queue entries mix all pair types and every round runs the ~5000-flop box-box code, while lanes
    // walking their OWN lists in pair order stay on one type together: 337 us own lists, 617 us shared out, 478 us generic.)
    int cnt = 0;
#pragma unroll
    for (int w = 0; w < PW; ++w) cnt += __popc(st.sv[w]);
    narrow_coop(kp, st, cnt);
    return;
  }
  EnvT<NT> E;
  E.env = st.ss; E.work = st.ss;
#pragma unroll
  for (int w = 0; w < PW; ++w) {
    uint32_t m = st.sv[w];
    while (m) {
      const int b = __ffs(m) - 1;
      m &= m - 1;
      const DItem* it = kp.items + (NJ + w * 32 + b);
      V2 fa;
      float ta, tb;
      eval_item<false>(kp, it, E, fa, ta, tb);
      if (THIN_POSSIBLE && (fa.x != 0.f || fa.y != 0.f) && !((st.verd[w] >> b) & 1u)) st.thin[w] |= 1u << b;
      accumulate<1>(it, E, fa, ta, tb, 0);
    }
  }
}

// step_kernel.cuh eval_item IT_JOINT (core.py:2200-2291, joints.py:206-215); jf = slot of the per-env fixed rotation
template <int ACC, int da, int db, int jf>
VB_DI void joint_r(St& st, const JointC jc, V2 pa, V2 pb, const Rot qa, const Rot qb, float ra, float rb) {
  V2 ja = pa + rotv(mk(jc.dax, jc.day), qa.c, qa.s);
  V2 jb = pb + rotv(mk(jc.dbx, jc.dby), qb.c, qb.s);
  V2 att, rep;
  constraint_force_pair(ja, jb, jc.dist, KJ, MARGIN, att, rep);
  V2 fa = att + rep;
  V2 fb = (-att) + (-rep);
  float ta = cross(ja - pa, fa);
  float tb = cross(jb - pb, fb);
  if (!jc.rotate) {
    float fr = jf >= 0 ? st.jfr[jf >= 0 ? jf : 0] : jc.fixed_rot;
    float tq = constraint_torque(ra, rb + fr, KT);
    ta = ta + (-tq);
    tb = tb + tq;
  }
  accumulate_r<ACC, da, db>(st, fa, ta, tb);
}

}  // namespace gen
}  // namespace vb
