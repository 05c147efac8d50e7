/*
 * vmas_oracle.c - TEST INFRASTRUCTURE (see vmas_oracle.h).  Scalar fp32 restatement of
 *   vmas/simulator/core.py:1971-2014 (World.step), :2017-2101 (action/friction/gravity),
 *   :2103-2198 (pair enumeration, update_env_forces), :2200-2291 (joints), :2293-2785 (six pair types),
 *   :2787-2857 (collides, constraint forces/torques), :2861-2907 (integrate),
 *   vmas/simulator/physics.py (closest-point geometry), vmas/simulator/utils.py:168-201 (TorchUtils),
 *   core.py:1280-1371,1413-1489,1543-1625,1661-1785 + sensors.py:60-69,100-122 (LIDAR).
 *
 * Every torch op of the reference is one rounded fp32 operation here (compile with -ffp-contract=off);
 * the only fused operation is torch.linalg.vector_norm over a contiguous last dim of size 2, which ATen
 * evaluates as sqrt(fma(y, y, x*x)) (measured against torch 2.11.0 CPU).
 *
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC)
 */
#include "vmas_oracle.h"
#include "sleef_f32.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXE 512
#define LMD ((float)(4.0 / 6e2)) /* utils.py:27 LINE_MIN_DIST, cast to fp32 where it meets a tensor */

typedef struct { float x, y; } v2;

static inline v2 V(float x, float y) { v2 r = {x, y}; return r; }
static inline v2 vsub(v2 a, v2 b) { return V(a.x - b.x, a.y - b.y); }
static inline v2 vadd(v2 a, v2 b) { return V(a.x + b.x, a.y + b.y); }
static inline v2 vscale(v2 a, float s) { return V(a.x * s, a.y * s); }
static inline v2 vdivs(v2 a, float s) { return V(a.x / s, a.y / s); }
/* torch.linalg.vector_norm(dim=-1) on [..., 2] contiguous */
static inline float vnorm(v2 a) { return sqrtf(fmaf(a.y, a.y, a.x * a.x)); }
/* (a * b).sum(-1) */
static inline float vdot(v2 a, v2 b) { return a.x * b.x + a.y * b.y; }
/* utils.py:194-197 TorchUtils.cross */
static inline float vcross(v2 a, v2 b) { return a.x * b.y - a.y * b.x; }
/* torch.sign: (0 < x) - (x < 0) */
static inline float fsign(float x) { return (float)((0.f < x) - (x < 0.f)); }
static inline float clampf(float x, float lo, float hi) { float m = x < lo ? lo : x; return m > hi ? hi : m; }
static const v2 VINF = {INFINITY, INFINITY};

/* cos / sin of an entity's rotation as the reference's call sites take them: (c, s) = torch.cos / torch.sin of rot
 * (utils.py:183-184, physics.py:299, 412), (c2, s2) of rot + pi/2 (physics.py:300-301, an fp32 sum).  By default they come
 * from libm; a substep may instead be handed the values torch computed for the same angles (vmas_oracle_substep `trig`):
 * torch's CPU cos / sin / exp are Intel MKL VML's (vsCos / vsSin / vsExp, high-accuracy mode; closed source, not
 * correctly rounded: they differ from libm's correctly rounded results by 1 ulp on ~5 % of arguments), and a 1-ulp
 * difference in an anchor's cos / sin is what a stiff joint (joint_force 900 on a moment of inertia of 5e-4) amplifies. */
typedef struct { float c, s, c2, s2; } tg;
static inline tg tg_of(float rot) {
  tg t;
  float rot2 = rot + (float)(M_PI / 2);
  t.c = cosf(rot); t.s = sinf(rot); t.c2 = cosf(rot2); t.s2 = sinf(rot2);
  return t;
}
static inline tg tg_cs(float c, float s) { tg t = {c, s, 0.f, 0.f}; return t; }

/* utils.py:176-191 rotate_vector */
static inline v2 vrot_cs(v2 v, float c, float s) { return V(v.x * c - v.y * s, v.x * s + v.y * c); }
static inline v2 vrot(v2 v, float ang) { return vrot_cs(v, cosf(ang), sinf(ang)); }

/* utils.py:168-173 clamp_with_norm on a 2-vector */
static inline v2 clamp_norm2(v2 t, float maxn) {
  float n = vnorm(t);
  v2 nt = vscale(vdivs(t, n), maxn);
  return (n > maxn) ? nt : t;
}
/* same on a [N,1] tensor: the norm over a size-1 dim is sqrt(x*x) */
static inline float clamp_norm1(float t, float maxn) {
  float n = sqrtf(t * t);
  float nt = (t / n) * maxn;
  return (n > maxn) ? nt : t;
}

/* torch.logaddexp(0, x) (ATen BinaryOpsKernel logaddexp: max + log1p(exp(-|a-b|)), inf==inf shortcut irrelevant for
 * a=0).  Its vectorised CPU path takes exp and log1p from Vectorized<float>, i.e. SLEEF's expf_u10 / log1pf_u10
 * (sleef_f32.h restates them; checked bit for bit against the Sleef_* symbols of torch 2.11.0's libtorch_cpu.so). */
static inline float softplus0v(float x, int vectorised) {
  float m = 0.f > x ? 0.f : x;
  /* ATen's cpu_kernel_vec runs whole groups of 2 x Vec::size() elements through the Vectorized (SLEEF) lambda and the
   * remaining tail of the tensor through the scalar one (std::exp, std::log1p: libm), Loops.h vectorized_loop */
  return vectorised ? m + sleef_log1pf_u10(sleef_expf_u10(-fabsf(0.f - x))) : m + log1pf(expf(-fabsf(0.f - x)));
}

/* physics.py:399-428 _get_closest_point_line */
static inline v2 closest_point_line(v2 lp, tg rot, float length, v2 q, int limit) {
  v2 u = V(rot.c, rot.s);
  v2 delta = vsub(lp, q);
  float dot = vdot(delta, u);
  float sg = fsign(dot);
  float half = length / 2.f;
  float a = fabsf(dot);
  float m = limit ? (a < half ? a : half) : a;
  float sm = sg * m;
  return vsub(lp, V(sm * u.x, sm * u.y));
}

typedef struct { v2 p[4]; tg rot[4]; float len[4]; } boxlines;

/* physics.py:297-324 _get_all_lines_box */
static inline void all_lines_box(v2 pos, tg rot, float width, float length, boxlines* o) {
  v2 rv = V(rot.c, rot.s);
  v2 rv2 = V(rot.c2, rot.s2);
  tg rot2 = tg_cs(rot.c2, rot.s2);
  float hl = length / 2.f, hw = width / 2.f;
  o->p[0] = vadd(pos, vscale(rv, hl));
  o->p[1] = vsub(pos, vscale(rv, hl));
  o->p[2] = vadd(pos, vscale(rv2, hw));
  o->p[3] = vsub(pos, vscale(rv2, hw));
  o->rot[0] = o->rot[1] = rot2;
  o->rot[2] = o->rot[3] = rot;
  o->len[0] = o->len[1] = width;
  o->len[2] = o->len[3] = length;
}

/* physics.py:262-294 _get_closest_point_box */
static inline v2 closest_point_box(v2 pos, tg rot, float width, float length, v2 q) {
  boxlines bl;
  all_lines_box(pos, rot, width, length, &bl);
  v2 best = VINF;
  float dist = INFINITY;
  for (int i = 0; i < 4; ++i) {
    v2 p = closest_point_line(bl.p[i], bl.rot[i], bl.len[i], q, 1);
    float d = vnorm(vsub(q, p));
    if (d < dist) { best = p; dist = d; }
  }
  return best;
}

/* physics.py:12-22 _get_inner_point_box */
static inline v2 inner_point_box(v2 outside, v2 surface, v2 box_pos, float* d_out) {
  v2 v = vsub(surface, outside);
  v2 u = vsub(box_pos, surface);
  float vn = vnorm(v);
  float xm = vdot(v, u) / vn;
  v2 x = vscale(vdivs(v, vn), xm);
  if (vn == 0.f) { x = surface; xm = 0.f; }
  *d_out = fabsf(xm);
  return vadd(surface, x);
}

/* physics.py:131-259 _get_line_extrema + _get_intersection_point_line_line + _get_closest_points_line_line */
static inline void closest_points_line_line(v2 pa, tg ra, float la, v2 pb, tg rb, float lb, v2* o1, v2* o2) {
  v2 xya = V((la / 2.f) * ra.c, (la / 2.f) * ra.s);
  v2 xyb = V((lb / 2.f) * rb.c, (lb / 2.f) * rb.s);
  v2 a1 = vadd(pa, xya), a2 = vsub(pa, xya);
  v2 b1 = vadd(pb, xyb), b2 = vsub(pb, xyb);
  /* intersection */
  v2 r = vsub(a2, a1), s = vsub(b2, b1);
  v2 qp = vsub(b1, a1);
  float cqpr = vcross(qp, r), cqps = vcross(qp, s), crs = vcross(r, s);
  float u = cqpr / crs, t = cqps / crs;
  int hit = (crs != 0.f) && (0.f <= u) && (u <= 1.f) && (0.f <= t) && (t <= 1.f);
  v2 pi = vadd(a1, vscale(r, t)); /* p + t * r */
  v2 c1[4], c2[4];
  c1[0] = a1; c2[0] = closest_point_line(pb, rb, lb, a1, 1);
  c1[1] = a2; c2[1] = closest_point_line(pb, rb, lb, a2, 1);
  c1[2] = closest_point_line(pa, ra, la, b1, 1); c2[2] = b1;
  c1[3] = closest_point_line(pa, ra, la, b2, 1); c2[3] = b2;
  v2 best1 = VINF, best2 = VINF;
  float md = INFINITY;
  for (int i = 0; i < 4; ++i) {
    float d = vnorm(vsub(c1[i], c2[i]));
    if (d < md) { best1 = c1[i]; best2 = c2[i]; md = d; }
  }
  if (hit) { best1 = pi; best2 = pi; }
  *o1 = best1; *o2 = best2;
}

/* physics.py:327-381 _get_closest_line_box -> (point on box, point on line) */
static inline void closest_line_box(v2 bpos, tg brot, float bw, float bl_, v2 lpos, tg lrot, float llen,
                                    v2* pbox, v2* pline) {
  boxlines bl;
  all_lines_box(bpos, brot, bw, bl_, &bl);
  v2 b1 = VINF, b2 = VINF;
  float dist = INFINITY;
  for (int i = 0; i < 4; ++i) {
    v2 p_box, p_line;
    closest_points_line_line(bl.p[i], bl.rot[i], bl.len[i], lpos, lrot, llen, &p_box, &p_line);
    float d = vnorm(vsub(p_box, p_line));
    if (d < dist) { b1 = p_box; b2 = p_line; dist = d; }
  }
  *pbox = b1; *pline = b2;
}

/* physics.py:25-128 _get_closest_box_box -> (point on box1, point on box2) */
static inline void closest_box_box(v2 p1, tg r1, float w1, float l1, v2 p2, tg r2, float w2, float l2, v2* o1,
                                   v2* o2) {
  boxlines la, lb;
  all_lines_box(p1, r1, w1, l1, &la);
  all_lines_box(p2, r2, w2, l2, &lb);
  v2 c1[8], c2[8];
  for (int i = 0; i < 4; ++i) { /* box2 vs edges of box1: (on box2, on edge of box1) */
    v2 on_box2, on_line1;
    closest_line_box(p2, r2, w2, l2, la.p[i], la.rot[i], la.len[i], &on_box2, &on_line1);
    c1[i] = on_line1; c2[i] = on_box2;
  }
  for (int i = 0; i < 4; ++i) { /* box1 vs edges of box2: (on box1, on edge of box2) */
    v2 on_box1, on_line2;
    closest_line_box(p1, r1, w1, l1, lb.p[i], lb.rot[i], lb.len[i], &on_box1, &on_line2);
    c1[4 + i] = on_box1; c2[4 + i] = on_line2;
  }
  v2 b1 = VINF, b2 = VINF;
  float dist = INFINITY;
  for (int i = 0; i < 8; ++i) {
    float d = vnorm(vsub(c1[i], c2[i]));
    if (d < dist) { b1 = c1[i]; b2 = c2[i]; dist = d; }
  }
  *o1 = b1; *o2 = b2;
}

/* ------------------------------------------------------------------------------------------------------------
 * core.py:2804-2838 _get_constraint_forces; returns force on a (force on b is its negation).
 * K = force_multiplier, k = contact_margin. */
static inline v2 constraint_force(v2 pa, v2 pb, float dist_min, float K, float k, int attractive, int vectorised) {
  v2 delta = vsub(pa, pb);
  float dist = vnorm(delta);
  float sg = attractive ? -1.f : 1.f;
  float pen = softplus0v((dist_min - dist) * sg / k, vectorised) * k;
  float sk = attractive ? -K : K; /* python: sign * force_multiplier, then one fp32 scalar */
  float den = dist > 0.f ? dist : 1e-8f;
  v2 f = V(sk * delta.x / den * pen, sk * delta.y / den * pen);
  if (dist < 1e-6f) f = V(0.f, 0.f);
  if (!attractive) { if (dist > dist_min) f = V(0.f, 0.f); }
  else { if (dist < dist_min) f = V(0.f, 0.f); }
  return f;
}

/* core.py:2840-2857 _get_constraint_torques; returns torque on b (torque on a is its negation). */
static inline float constraint_torque(float rot_a, float rot_b, float Kt, const float* exp_a) {
  float d = rot_a - rot_b;
  float a = sqrtf(d * d);
  float pen = (exp_a ? *exp_a : expf(a)) - 1.f; /* torch.exp: MKL vsExp on the CPU (see tg above) */
  float tq = Kt * fsign(d) * pen;
  if (a < 1e-9f) tq = 0.f;
  return tq;
}

typedef struct { int type, a, b; } opair;

typedef struct {
  const vmasb200_world_desc* d;
  int E, A, S, J;
  int agent_of[MAXE];   /* entity -> agent slot or -1 */
  int npair;
  opair* pairs;         /* candidate pairs in per-type order ss, ls, ll, bs, bl, bb */
  int type_begin[7];
  float sub_dt;
} octx;

static int build_ctx(const vmasb200_world_desc* d, octx* c) {
  memset(c, 0, sizeof(*c));
  c->d = d; c->E = d->n_entities; c->S = d->substeps; c->J = d->n_joints;
  if (c->E > MAXE) return -1;
  int A = 0;
  for (int e = 0; e < c->E; ++e) c->agent_of[e] = (d->entities[e].flags & VMASB200_E_AGENT) ? A++ : -1;
  c->A = A;
  c->sub_dt = (float)(d->dt / (double)d->substeps); /* core.py:1124 in Python doubles */
  int E = c->E;
  c->pairs = (opair*)malloc(sizeof(opair) * (size_t)(E * E / 2 + 1));
  int n = 0;
  for (int t = 0; t < 6; ++t) { /* core.py:2124-2173 bucketing, lists kept in a<b enumeration order */
    c->type_begin[t] = n;
    for (int a = 0; a < E; ++a)
      for (int b = a + 1; b < E; ++b) {
        if (!d->collide_mask[a * E + b]) continue;
        int sa = d->entities[a].shape, sb = d->entities[b].shape;
        int ty, x = a, y = b;
        if (sa == VMASB200_SPHERE && sb == VMASB200_SPHERE) ty = VMASB200_PAIR_SS;
        else if ((sa == VMASB200_LINE && sb == VMASB200_SPHERE) || (sb == VMASB200_LINE && sa == VMASB200_SPHERE)) {
          ty = VMASB200_PAIR_LS; if (sb != VMASB200_SPHERE) { x = b; y = a; }
        } else if (sa == VMASB200_LINE && sb == VMASB200_LINE) ty = VMASB200_PAIR_LL;
        else if ((sa == VMASB200_BOX && sb == VMASB200_SPHERE) || (sb == VMASB200_BOX && sa == VMASB200_SPHERE)) {
          ty = VMASB200_PAIR_BS; if (sb != VMASB200_SPHERE) { x = b; y = a; }
        } else if ((sa == VMASB200_BOX && sb == VMASB200_LINE) || (sb == VMASB200_BOX && sa == VMASB200_LINE)) {
          ty = VMASB200_PAIR_BL; if (sb != VMASB200_LINE) { x = b; y = a; }
        } else ty = VMASB200_PAIR_BB;
        if (ty != t) continue;
        c->pairs[n].type = ty; c->pairs[n].a = x; c->pairs[n].b = y; ++n;
      }
  }
  c->type_begin[6] = n;
  c->npair = n;
  return 0;
}

#define ST(e, i, k) state[((size_t)(e) * (size_t)n + (size_t)(i)) * 6 + (k)]

static inline void add_env_forces(const vmasb200_entity* ents, v2* F, float* T, int a, v2 fa, float ta, int b, v2 fb,
                                  float tb) { /* core.py:2190-2198 */
  if (ents[a].flags & VMASB200_E_MOVABLE) F[a] = vadd(F[a], fa);
  if (ents[a].flags & VMASB200_E_ROTATABLE) T[a] = T[a] + ta;
  if (ents[b].flags & VMASB200_E_MOVABLE) F[b] = vadd(F[b], fb);
  if (ents[b].flags & VMASB200_E_ROTATABLE) T[b] = T[b] + tb;
}

static inline v2 vneg(v2 a) { return V(-a.x, -a.y); }

/* get_friction_force, core.py:2054-2073, for a 2-vector (linear) */
static inline v2 friction2(v2 vel, float coeff, float mass, float sub_dt) {
  float speed = vnorm(vel);
  int st = speed == 0.f;
  float c = coeff * mass;
  float den = st ? 1e-8f : speed;
  float mx = (fabsf(vel.x) / sub_dt) * mass, my = (fabsf(vel.y) / sub_dt) * mass;
  v2 f = V(-(vel.x / den) * (c < mx ? c : mx), -(vel.y / den) * (c < my ? c : my));
  if (st) f = V(0.f, 0.f);
  return f;
}
static inline float friction1(float w, float coeff, float inertia, float sub_dt) {
  float speed = sqrtf(w * w);
  int st = speed == 0.f;
  float c = coeff * inertia;
  float den = st ? 1e-8f : speed;
  float m = (fabsf(w) / sub_dt) * inertia;
  float f = -(w / den) * (c < m ? c : m);
  if (st) f = 0.f;
  return f;
}

/* ------------------------------------------------------------------------------------------------------------
 * One substep for one env (core.py:1975-2009) given the batch-wide pair verdicts `active`. */
typedef struct { /* optional per-substep inputs / outputs of vmas_oracle_substep (any may be NULL) */
  const float* trig;  /* [E][N][4] in: cos(rot) sin(rot) cos(rot + pi/2) sin(rot + pi/2) to use instead of libm's    */
  int vec_width;      /* 0: every logaddexp element takes the Vectorized path; W > 0: elements of a [N, P] logaddexp call
                         beyond its last whole group of W take the scalar path, as ATen does with W = 2 x Vec::size() */
  const float* jexp;  /* [J][N] in: exp(|rot_a - (rot_b + fixed_rotation)|) of core.py:2852 to use instead of libm's */
  float* tot_force;   /* [E][N][2] forces_dict as _integrate_state reads it (core.py:2876)   */
  float* tot_torque;  /* [E][N]    torques_dict (core.py:2899)                                */
  float* joint_force; /* [J][2][N][2] what core.py:2282-2291 hands to update_env_forces: a, b */
  float* joint_torque;/* [J][2][N]                                                            */
} odebug;

static void substep_env(const octx* c, int64_t n, int64_t i, float* state, float* force, float* torque,
                        const float* jrot, const uint8_t* active, const int* rank, const int* type_count, int substep,
                        const odebug* dbg) {
  const vmasb200_world_desc* d = c->d;
  const vmasb200_entity* ents = d->entities;
  const int E = c->E;
  const float sub_dt = c->sub_dt;
  const float Kc = d->collision_force, Kj = d->joint_force, k = d->contact_margin;
  v2 F[MAXE];
  float T[MAXE];
  for (int e = 0; e < E; ++e) { F[e] = V(0.f, 0.f); T[e] = 0.f; }
  /* cos / sin of an entity's rotation at the substep's entry: computed when first needed, or the caller's */
  /* position of this env's element inside the reference's logaddexp call over a [N, P] tensor (env-major) */
  const int W = dbg ? dbg->vec_width : 0;
#define VEC(idx, numel) (W <= 0 || (int64_t)(idx) < ((int64_t)(numel) / W) * W)
  tg TG[MAXE];
  uint8_t have_tg[MAXE];
  memset(have_tg, 0, (size_t)E);
#define TGE(e) (have_tg[e] ? TG[e] : (have_tg[e] = 1, TG[e] = (dbg && dbg->trig) ? *(const tg*)(dbg->trig + ((size_t)(e) * (size_t)n + (size_t)i) * 4) : tg_of(ST(e, i, 4))))

  for (int e = 0; e < E; ++e) {
    const vmasb200_entity* en = &ents[e];
    v2 vel = V(ST(e, i, 2), ST(e, i, 3));
    float w = ST(e, i, 5);
    if (en->flags & VMASB200_E_AGENT) {
      int a = c->agent_of[e];
      if (en->flags & VMASB200_E_MOVABLE) { /* core.py:2017-2027 */
        float* fp = &force[((size_t)a * (size_t)n + (size_t)i) * 2];
        v2 f = V(fp[0], fp[1]);
        if (en->flags & VMASB200_E_MAX_F) f = clamp_norm2(f, en->max_f);
        if (en->flags & VMASB200_E_F_RANGE) f = V(clampf(f.x, -en->f_range, en->f_range), clampf(f.y, -en->f_range, en->f_range));
        fp[0] = f.x; fp[1] = f.y;
        F[e] = vadd(F[e], f);
      }
      if (en->flags & VMASB200_E_ROTATABLE) { /* core.py:2029-2040 */
        float* tp = &torque[(size_t)a * (size_t)n + (size_t)i];
        float t = *tp;
        if (en->flags & VMASB200_E_MAX_T) t = clamp_norm1(t, en->max_t);
        if (en->flags & VMASB200_E_T_RANGE) t = clampf(t, -en->t_range, en->t_range);
        *tp = t;
        T[e] = T[e] + t;
      }
    }
    /* friction, core.py:2075-2101 */
    if (en->flags & VMASB200_E_LIN_FRICTION) F[e] = vadd(F[e], friction2(vel, en->linear_friction, en->mass, sub_dt));
    else if (d->linear_friction > 0.f) F[e] = vadd(F[e], friction2(vel, d->linear_friction, en->mass, sub_dt));
    if (en->flags & VMASB200_E_ANG_FRICTION) T[e] = T[e] + friction1(w, en->angular_friction, en->moment_of_inertia, sub_dt);
    else if (d->angular_friction > 0.f) T[e] = T[e] + friction1(w, d->angular_friction, en->moment_of_inertia, sub_dt);
    /* gravity, core.py:2042-2051 */
    if (en->flags & VMASB200_E_MOVABLE) {
      if (!(d->gravity[0] == 0.f && d->gravity[1] == 0.f))
        F[e] = vadd(F[e], V(en->mass * d->gravity[0], en->mass * d->gravity[1]));
      if (en->flags & VMASB200_E_GRAVITY) F[e] = vadd(F[e], V(en->mass * en->gravity[0], en->mass * en->gravity[1]));
    }
  }

  /* joints, core.py:2200-2291 */
  for (int j = 0; j < c->J; ++j) {
    const vmasb200_joint* jt = &d->joints[j];
    int a = jt->entity_a, b = jt->entity_b;
    v2 pa = V(ST(a, i, 0), ST(a, i, 1)), pb = V(ST(b, i, 0), ST(b, i, 1));
    float ra = ST(a, i, 4), rb = ST(b, i, 4);
    tg qa = TGE(a), qb = TGE(b);
    v2 ja = vadd(pa, vrot_cs(V(jt->delta_a[0], jt->delta_a[1]), qa.c, qa.s)); /* joints.py:206-215 */
    v2 jb = vadd(pb, vrot_cs(V(jt->delta_b[0], jt->delta_b[1]), qb.c, qb.s));
    const int vj = VEC(i * c->J + j, n * c->J);
    v2 fa_att = constraint_force(ja, jb, jt->dist, Kj, k, 1, vj);
    v2 fa_rep = constraint_force(ja, jb, jt->dist, Kj, k, 0, vj);
    v2 fa = vadd(fa_att, fa_rep);
    v2 fb = vadd(vneg(fa_att), vneg(fa_rep));
    float ta = vcross(vsub(ja, pa), fa), tb = vcross(vsub(jb, pb), fb);
    if (!jt->rotate) {
      float fr = jt->fixed_rotation_per_env ? jrot[(size_t)j * (size_t)n + (size_t)i] : jt->fixed_rotation;
      float tq = constraint_torque(ra, rb + fr, d->torque_constraint_force,
                                   (dbg && dbg->jexp) ? dbg->jexp + (size_t)j * (size_t)n + (size_t)i : NULL);
      ta = ta + (-tq);
      tb = tb + tq;
    }
    add_env_forces(ents, F, T, a, fa, ta, b, fb, tb);
    if (dbg && dbg->joint_force) {
      float* o = dbg->joint_force + ((size_t)j * 2 * (size_t)n + (size_t)i) * 2;
      o[0] = fa.x; o[1] = fa.y; o[(size_t)n * 2] = fb.x; o[(size_t)n * 2 + 1] = fb.y;
    }
    if (dbg && dbg->joint_torque) {
      dbg->joint_torque[(size_t)j * 2 * (size_t)n + (size_t)i] = ta;
      dbg->joint_torque[((size_t)j * 2 + 1) * (size_t)n + (size_t)i] = tb;
    }
  }

  /* collisions, core.py:2293-2785, lists in type order, each in a<b enumeration order */
  for (int p = 0; p < c->npair; ++p) {
    if (!active[p]) continue;
    int a = c->pairs[p].a, b = c->pairs[p].b;
    const vmasb200_entity *ea = &ents[a], *eb = &ents[b];
    v2 pa = V(ST(a, i, 0), ST(a, i, 1)), pb = V(ST(b, i, 0), ST(b, i, 1));
    const int pt = type_count[c->pairs[p].type];
    const int vp = VEC(i * pt + rank[p], n * pt);
    switch (c->pairs[p].type) {
      case VMASB200_PAIR_SS: { /* :2293 */
        v2 fa = constraint_force(pa, pb, ea->radius + eb->radius, Kc, k, 0, vp);
        add_env_forces(ents, F, T, a, fa, 0.f, b, vneg(fa), 0.f);
      } break;
      case VMASB200_PAIR_LS: { /* :2340 a=line b=sphere */
        v2 cp = closest_point_line(pa, TGE(a), ea->length, pb, 1);
        v2 fs = constraint_force(pb, cp, eb->radius + LMD, Kc, k, 0, vp);
        v2 fl = vneg(fs);
        float tl = vcross(vsub(cp, pa), fl);
        add_env_forces(ents, F, T, a, fl, tl, b, fs, 0.f);
      } break;
      case VMASB200_PAIR_LL: { /* :2393 */
        v2 ca, cb;
        closest_points_line_line(pa, TGE(a), ea->length, pb, TGE(b), eb->length, &ca, &cb);
        v2 fa = constraint_force(ca, cb, LMD, Kc, k, 0, vp);
        v2 fb = vneg(fa);
        add_env_forces(ents, F, T, a, fa, vcross(vsub(ca, pa), fa), b, fb, vcross(vsub(cb, pb), fb));
      } break;
      case VMASB200_PAIR_BS: { /* :2458 a=box b=sphere */
        v2 cp = closest_point_box(pa, TGE(a), ea->width, ea->length, pb);
        v2 ip = cp; float dd = 0.f;
        if (!(ea->flags & VMASB200_E_HOLLOW)) ip = inner_point_box(pb, cp, pa, &dd);
        v2 fs = constraint_force(pb, ip, (eb->radius + LMD) + dd, Kc, k, 0, vp);
        v2 fbx = vneg(fs);
        add_env_forces(ents, F, T, a, fbx, vcross(vsub(cp, pa), fbx), b, fs, 0.f);
      } break;
      case VMASB200_PAIR_BL: { /* :2553 a=box b=line */
        v2 cbx, cl;
        closest_line_box(pa, TGE(a), ea->width, ea->length, pb, TGE(b), eb->length, &cbx, &cl);
        v2 ip = cbx; float dd = 0.f;
        if (!(ea->flags & VMASB200_E_HOLLOW)) ip = inner_point_box(cl, cbx, pa, &dd);
        v2 fbx = constraint_force(ip, cl, LMD + dd, Kc, k, 0, vp);
        v2 fl = vneg(fbx);
        add_env_forces(ents, F, T, a, fbx, vcross(vsub(cbx, pa), fbx), b, fl, vcross(vsub(cl, pb), fl));
      } break;
      default: { /* :2654 box-box */
        v2 ca, cb;
        closest_box_box(pa, TGE(a), ea->width, ea->length, pb, TGE(b), eb->width, eb->length, &ca, &cb);
        v2 ia = ca, ib = cb; float da = 0.f, db = 0.f;
        if (!(ea->flags & VMASB200_E_HOLLOW)) ia = inner_point_box(cb, ca, pa, &da);
        if (!(eb->flags & VMASB200_E_HOLLOW)) ib = inner_point_box(ca, cb, pb, &db);
        v2 fa = constraint_force(ia, ib, (da + db) + LMD, Kc, k, 0, vp);
        v2 fb = vneg(fa);
        add_env_forces(ents, F, T, a, fa, vcross(vsub(ca, pa), fa), b, fb, vcross(vsub(cb, pb), fb));
      } break;
    }
  }

  if (dbg && dbg->tot_force)
    for (int e = 0; e < E; ++e) { float* o = dbg->tot_force + ((size_t)e * (size_t)n + (size_t)i) * 2; o[0] = F[e].x; o[1] = F[e].y; }
  if (dbg && dbg->tot_torque)
    for (int e = 0; e < E; ++e) dbg->tot_torque[(size_t)e * (size_t)n + (size_t)i] = T[e];

  /* integrate, core.py:2861-2907 */
  for (int e = 0; e < E; ++e) {
    const vmasb200_entity* en = &ents[e];
    double drag = (en->flags & VMASB200_E_DRAG) ? en->drag : d->drag; /* Python floats */
    float one_minus_drag = (float)(1.0 - drag);
    if (en->flags & VMASB200_E_MOVABLE) {
      v2 vel = V(ST(e, i, 2), ST(e, i, 3));
      if (substep == 0) vel = vscale(vel, one_minus_drag);
      v2 acc = vdivs(F[e], en->mass);
      vel = vadd(vel, vscale(acc, sub_dt));
      if (en->flags & VMASB200_E_MAX_SPEED) vel = clamp_norm2(vel, en->max_speed);
      if (en->flags & VMASB200_E_V_RANGE) vel = V(clampf(vel.x, -en->v_range, en->v_range), clampf(vel.y, -en->v_range, en->v_range));
      v2 np_ = vadd(V(ST(e, i, 0), ST(e, i, 1)), vscale(vel, sub_dt));
      if (d->has_x_semidim) np_.x = clampf(np_.x, -d->x_semidim, d->x_semidim);
      if (d->has_y_semidim) np_.y = clampf(np_.y, -d->y_semidim, d->y_semidim);
      ST(e, i, 2) = vel.x; ST(e, i, 3) = vel.y; ST(e, i, 0) = np_.x; ST(e, i, 1) = np_.y;
    }
    if (en->flags & VMASB200_E_ROTATABLE) {
      float w = ST(e, i, 5);
      if (substep == 0) w = w * one_minus_drag;
      w = w + (T[e] / en->moment_of_inertia) * sub_dt;
      ST(e, i, 5) = w;
      ST(e, i, 4) = ST(e, i, 4) + w * sub_dt;
    }
  }
}

static int run_substeps(const vmasb200_world_desc* desc, int64_t n, float* state, float* force, float* torque,
                        const float* jrot, uint8_t* masks, int n_threads, int s_begin, int s_end, const odebug* dbg) {
  octx c;
  if (build_ctx(desc, &c)) return -1;
  if (s_begin < 0 || s_end > c.S || s_begin > s_end) { free(c.pairs); return -2; }
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#else
  n_threads = 1;
#endif
  const int E = c.E;
  uint8_t* active = (uint8_t*)calloc((size_t)c.npair + 1, 1);
  int* rank = (int*)calloc((size_t)c.npair + 1, sizeof(int));
  int type_count[6];
  for (int s = s_begin; s < s_end; ++s) {
    /* batch-wide broad phase, core.py:2796-2800 */
    for (int p = 0; p < c.npair; ++p) {
      int a = c.pairs[p].a, b = c.pairs[p].b;
      float thr = (float)(desc->entities[a].circumscribed_radius + desc->entities[b].circumscribed_radius);
      int any = 0;
#pragma omp parallel for reduction(| : any) num_threads(n_threads) if (n > 4096)
      for (int64_t i = 0; i < n; ++i) {
        v2 dlt = V(ST(a, i, 0) - ST(b, i, 0), ST(a, i, 1) - ST(b, i, 1));
        any |= (vnorm(dlt) <= thr);
      }
      active[p] = (uint8_t)any;
      if (masks) {
        int lo = a < b ? a : b, hi = a < b ? b : a;
        masks[((size_t)s * E + lo) * E + hi] = (uint8_t)any;
      }
    }
    memset(type_count, 0, sizeof(type_count));
    for (int p = 0; p < c.npair; ++p)
      if (active[p]) rank[p] = type_count[c.pairs[p].type]++;
#pragma omp parallel for num_threads(n_threads) schedule(static) if (n > 64)
    for (int64_t i = 0; i < n; ++i) substep_env(&c, n, i, state, force, torque, jrot, active, rank, type_count, s, dbg);
  }
  free(active);
  free(rank);
  free(c.pairs);
  return 0;
}

int vmas_oracle_step(const vmasb200_world_desc* desc, int64_t n, float* state, float* force, float* torque,
                     const float* jrot, uint8_t* masks, int n_threads) {
  return run_substeps(desc, n, state, force, torque, jrot, masks, n_threads, 0, desc->substeps, NULL);
}

int vmas_oracle_substep(const vmasb200_world_desc* desc, int64_t n, float* state, float* force, float* torque,
                        const float* jrot, uint8_t* masks, int substep, float* tot_force, float* tot_torque,
                        float* joint_force, float* joint_torque, const float* trig, const float* jexp, int vec_width,
                        int n_threads) {
  odebug dbg = {trig, vec_width, jexp, tot_force, tot_torque, joint_force, joint_torque};
  return run_substeps(desc, n, state, force, torque, jrot, masks, n_threads, substep, substep + 1, &dbg);
}

/* ------------------------------------------------------------------------------------------------------------
 * LIDAR: core.py:1280-1371 (box), :1413-1489 (sphere), :1543-1625 (line), :1661-1785 (min over targets). */
/* torch.min / torch.max over a stacked dim propagate NaN */
static inline float pmin(float a, float b) { return (a != a || b != b) ? NAN : (a < b ? a : b); }
static inline float pmax(float a, float b) { return (a != a || b != b) ? NAN : (a > b ? a : b); }

static inline float ray_box(v2 bpos, float brot, float blen, float bwid, v2 o, float ang, float max_range) {
  v2 pos_origin = vsub(o, bpos);
  v2 pos_aabb = vrot(pos_origin, -brot);
  v2 dir_w = V(cosf(ang), sinf(ang));
  v2 dir = vrot(dir_w, -brot);
  float tx1 = (-blen / 2.f - pos_aabb.x) / dir.x;
  float tx2 = (blen / 2.f - pos_aabb.x) / dir.x;
  float tmin = pmin(tx1, tx2), tmax = pmax(tx1, tx2);
  float ty1 = (-bwid / 2.f - pos_aabb.y) / dir.y;
  float ty2 = (bwid / 2.f - pos_aabb.y) / dir.y;
  float tymin = pmin(ty1, ty2), tymax = pmax(ty1, ty2);
  tmin = pmax(tmin, tymin);
  tmax = pmin(tmax, tymax);
  v2 ia = vadd(vscale(dir, tmin), pos_aabb); /* tmin * ray_dir_aabb + pos_aabb */
  v2 iw = vadd(vrot(ia, brot), bpos);
  int collision = (tmax >= tmin) && (tmin > 0.f);
  float dist = vnorm(vsub(o, iw));
  return collision ? dist : max_range;
}

static inline float ray_sphere(v2 spos, float rad, v2 o, float ang, float max_range) {
  v2 dir = V(cosf(ang), sinf(ang));
  v2 line_pos = vadd(o, vscale(dir, max_range / 2.f));
  v2 cp = closest_point_line(line_pos, tg_cs(dir.x, dir.y), max_range, spos, 0);
  v2 dd = vsub(spos, cp);
  float dn = vnorm(dd);
  int hit = dn < rad;
  float a = rad * rad - dn * dn; /* radius**2 - d_norm**2 */
  float m = sqrtf(a > 0.f ? a : 1e-8f);
  v2 u = vsub(spos, o);
  v2 u1 = vsub(cp, o);
  int front = vdot(u, dir) > 0.f;
  float dist = vnorm(u1) - m;
  return (hit && front) ? dist : max_range;
}

static inline float ray_line(v2 lpos, float lrot, float llen, v2 o, float ang, float max_range) {
  v2 r = V(cosf(lrot) * llen, sinf(lrot) * llen);
  v2 s = V(cosf(ang), sinf(ang));
  float rxs = vcross(r, s);
  v2 qp = vsub(o, lpos);
  float t = vcross(qp, vdivs(s, rxs));
  float u = vcross(qp, vdivs(r, rxs));
  float d = vnorm(V(u * s.x, u * s.y));
  if (rxs == 0.f) d = max_range;
  if (t > 0.5f) d = max_range;
  if (t < -0.5f) d = max_range;
  if (u < 0.f) d = max_range;
  return d;
}

int vmas_oracle_cast_rays(const vmasb200_world_desc* desc, int64_t n, const float* state_c, int entity,
                          const float* angles, int n_rays, float max_range, const int32_t* targets, int n_targets,
                          float* out, int n_threads) {
  float* state = (float*)state_c; /* ST() macro reads only */
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#endif
#pragma omp parallel for num_threads(n_threads) schedule(static) if (n > 64)
  for (int64_t i = 0; i < n; ++i) {
    v2 o = V(ST(entity, i, 0), ST(entity, i, 1));
    for (int r = 0; r < n_rays; ++r) {
      float ang = angles[(size_t)i * n_rays + r];
      float best = max_range;
      for (int k = 0; k < n_targets; ++k) {
        int e = targets[k];
        const vmasb200_entity* en = &desc->entities[e];
        v2 p = V(ST(e, i, 0), ST(e, i, 1));
        float rot = ST(e, i, 4);
        float dd;
        if (en->shape == VMASB200_BOX) dd = ray_box(p, rot, en->length, en->width, o, ang, max_range);
        else if (en->shape == VMASB200_SPHERE) dd = ray_sphere(p, en->radius, o, ang, max_range);
        else dd = ray_line(p, rot, en->length, o, ang, max_range);
        best = pmin(best, dd);
      }
      out[(size_t)i * n_rays + r] = best;
    }
  }
  return 0;
}

int vmas_oracle_lidar(const vmasb200_world_desc* desc, int64_t n, const float* state_c, int lidar_index, float* out,
                      int n_threads) {
  if (lidar_index < 0 || lidar_index >= desc->n_lidars) return -1;
  const vmasb200_lidar* ld = &desc->lidars[lidar_index];
  float* state = (float*)state_c;
  float* ang = (float*)malloc(sizeof(float) * (size_t)n * (size_t)ld->n_rays);
  for (int64_t i = 0; i < n; ++i)
    for (int r = 0; r < ld->n_rays; ++r) /* sensors.py:117: self._angles + agent.state.rot */
      ang[(size_t)i * ld->n_rays + r] = ld->angles[r] + ST(ld->entity, i, 4);
  int rc = vmas_oracle_cast_rays(desc, n, state_c, ld->entity, ang, ld->n_rays, ld->max_range, ld->targets,
                                 ld->n_targets, out, n_threads);
  free(ang);
  return rc;
}

/* ---- distance queries.  A shape is (kind, radius, length, width) as Python floats (doubles); a pose is
 * the entity's [N][6] block of the fixture state. */
typedef struct { int shape; double radius, length, width; const float* st; } qent;
#define QP(e, i) V((e)->st[(size_t)(i) * 6 + 0], (e)->st[(size_t)(i) * 6 + 1])
#define QR(e, i) ((e)->st[(size_t)(i) * 6 + 4])

/* core.py:1787-1819 get_distance_from_point */
static float q_from_point(const qent* e, int64_t i, v2 q) {
  if (e->shape == VMASB200_SPHERE) return vnorm(vsub(QP(e, i), q)) - (float)e->radius;
  v2 cp = e->shape == VMASB200_BOX ? closest_point_box(QP(e, i), tg_of(QR(e, i)), (float)e->width, (float)e->length, q)
                                   : closest_point_line(QP(e, i), tg_of(QR(e, i)), (float)e->length, q, 1);
  return vnorm(vsub(q, cp)) - LMD;
}

/* core.py:1936-1964 is_overlapping, box / sphere branch */
static int q_overlap_box_sphere(const qent* box, const qent* sph, int64_t i) {
  v2 bp = QP(box, i), sp = QP(sph, i);
  v2 cp = closest_point_box(bp, tg_of(QR(box, i)), (float)box->width, (float)box->length, sp);
  float d_sphere_cp = vnorm(vsub(sp, cp));
  float d_sphere_box = vnorm(vsub(sp, bp));
  float d_cp_box = vnorm(vsub(bp, cp));
  float dist_min = (float)(sph->radius + 4.0 / 6e2); /* Python double sum, cast where it meets the tensor */
  return (d_sphere_box < d_cp_box) || (d_sphere_cp < dist_min);
}

/* core.py:1821-1907 get_distance */
static float q_distance(const qent* a, const qent* b, int64_t i) {
  int sa = a->shape, sb = b->shape;
  if (sa == VMASB200_SPHERE && sb == VMASB200_SPHERE) return q_from_point(a, i, QP(b, i)) - (float)b->radius;
  if ((sa == VMASB200_BOX && sb == VMASB200_SPHERE) || (sb == VMASB200_BOX && sa == VMASB200_SPHERE)) {
    const qent* box = sb == VMASB200_SPHERE ? a : b;
    const qent* sph = sb == VMASB200_SPHERE ? b : a;
    float d = q_from_point(box, i, QP(sph, i)) - (float)sph->radius;
    return q_overlap_box_sphere(box, sph, i) ? -1.f : d;
  }
  if ((sa == VMASB200_LINE && sb == VMASB200_SPHERE) || (sb == VMASB200_LINE && sa == VMASB200_SPHERE)) {
    const qent* line = sb == VMASB200_SPHERE ? a : b;
    const qent* sph = sb == VMASB200_SPHERE ? b : a;
    return q_from_point(line, i, QP(sph, i)) - (float)sph->radius;
  }
  v2 p1, p2;
  if (sa == VMASB200_LINE && sb == VMASB200_LINE) {
    closest_points_line_line(QP(a, i), tg_of(QR(a, i)), (float)a->length, QP(b, i), tg_of(QR(b, i)), (float)b->length, &p1, &p2);
  } else if (sa == VMASB200_BOX && sb == VMASB200_BOX) {
    closest_box_box(QP(a, i), tg_of(QR(a, i)), (float)a->width, (float)a->length, QP(b, i), tg_of(QR(b, i)), (float)b->width,
                    (float)b->length, &p1, &p2);
  } else {
    const qent* box = sa == VMASB200_BOX ? a : b;
    const qent* line = sa == VMASB200_BOX ? b : a;
    closest_line_box(QP(box, i), tg_of(QR(box, i)), (float)box->width, (float)box->length, QP(line, i), tg_of(QR(line, i)),
                     (float)line->length, &p1, &p2);
  }
  return vnorm(vsub(p1, p2)) - LMD;
}

int vmas_oracle_get_distance(int shape_a, const double* dims_a, const float* state_a, int shape_b, const double* dims_b,
                             const float* state_b, int64_t n, float* out) {
  qent a = {shape_a, dims_a[0], dims_a[1], dims_a[2], state_a}, b = {shape_b, dims_b[0], dims_b[1], dims_b[2], state_b};
  for (int64_t i = 0; i < n; ++i) out[i] = q_distance(&a, &b, i);
  return 0;
}

int vmas_oracle_is_overlapping(int shape_a, const double* dims_a, const float* state_a, int shape_b,
                               const double* dims_b, const float* state_b, int64_t n, uint8_t* out) {
  qent a = {shape_a, dims_a[0], dims_a[1], dims_a[2], state_a}, b = {shape_b, dims_b[0], dims_b[1], dims_b[2], state_b};
  int bs = (shape_a == VMASB200_BOX && shape_b == VMASB200_SPHERE) || (shape_b == VMASB200_BOX && shape_a == VMASB200_SPHERE);
  for (int64_t i = 0; i < n; ++i) {
    if (bs) out[i] = (uint8_t)q_overlap_box_sphere(shape_b == VMASB200_SPHERE ? &a : &b, shape_b == VMASB200_SPHERE ? &b : &a, i);
    else out[i] = q_distance(&a, &b, i) < 0.f; /* core.py:1934 */
  }
  return 0;
}

int vmas_oracle_distance_from_point(int shape_a, const double* dims_a, const float* state_a, const float* points,
                                    int64_t n, float* out) {
  qent a = {shape_a, dims_a[0], dims_a[1], dims_a[2], state_a};
  for (int64_t i = 0; i < n; ++i) out[i] = q_from_point(&a, i, V(points[2 * i], points[2 * i + 1]));
  return 0;
}
