// geom.cuh - device-side fp32 math of the VMAS physics step (sm_100a).
//
// Every function restates ONE reference routine with the reference's operation order, so that each torch op
// becomes exactly one rounded fp32 instruction (the translation unit is compiled with -fmad=false; the only
// fused multiply-add is the one ATen itself performs inside linalg.vector_norm).  Reference citations are
// relative to the reference tree (vmas/simulator/...).
//
// Difference to the reference's data flow: cos/sin of an entity's rotation are computed ONCE per entity per
// substep (struct Rot) and shared by all pairs that touch the entity, instead of once per pair; the values are
// identical because every pair evaluates cos/sin of the same fp32 angle.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

namespace vb {

#define VB_DI __device__ __forceinline__

constexpr float kLMD = (float)(4.0 / 6e2);  // utils.py:27 LINE_MIN_DIST

struct V2 { float x, y; };
VB_DI V2 mk(float x, float y) { V2 r; r.x = x; r.y = y; return r; }
VB_DI V2 operator+(V2 a, V2 b) { return mk(a.x + b.x, a.y + b.y); }
VB_DI V2 operator-(V2 a, V2 b) { return mk(a.x - b.x, a.y - b.y); }
VB_DI V2 operator-(V2 a) { return mk(-a.x, -a.y); }
VB_DI V2 operator*(V2 a, float s) { return mk(a.x * s, a.y * s); }
VB_DI V2 operator/(V2 a, float s) { return mk(a.x / s, a.y / s); }
// torch.linalg.vector_norm(dim=-1) over a contiguous size-2 dim: sqrt(fma(y, y, x*x))
VB_DI float norm(V2 a) { return sqrtf(fmaf(a.y, a.y, a.x * a.x)); }
VB_DI float dot(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
VB_DI float cross(V2 a, V2 b) { return a.x * b.y - a.y * b.x; }  // utils.py:194-197
VB_DI float sgn(float x) { return (float)((0.f < x) - (x < 0.f)); }  // torch.sign
VB_DI float clampf(float x, float lo, float hi) { float m = x < lo ? lo : x; return m > hi ? hi : m; }
VB_DI V2 vinf() { return mk(CUDART_INF_F, CUDART_INF_F); }

// cos/sin of an entity rotation; c2/s2 are cos/sin of (rot + pi/2) (box edges, physics.py:299-301)
struct Rot { float c, s, c2, s2; };

// utils.py:176-191 rotate_vector with precomputed cos/sin
VB_DI V2 rotv(V2 v, float c, float s) { return mk(v.x * c - v.y * s, v.x * s + v.y * c); }

// utils.py:168-173 clamp_with_norm
VB_DI V2 clamp_norm2(V2 t, float maxn) {
  float n = norm(t);
  V2 nt = (t / n) * maxn;
  return (n > maxn) ? nt : t;
}
VB_DI float clamp_norm1(float t, float maxn) {
  float n = sqrtf(t * t);
  float nt = (t / n) * maxn;
  return (n > maxn) ? nt : t;
}

// torch.logaddexp(0, x): max(0,x) + log1p(exp(-|0-x|))
VB_DI float softplus0(float x) {
  float m = 0.f > x ? 0.f : x;
  return m + log1pf(expf(-fabsf(0.f - x)));
}

// physics.py:399-428 _get_closest_point_line; (c,s) = cos/sin(line_rot)
VB_DI V2 closest_point_line(V2 lp, float c, float s, float length, V2 q, bool limit = true) {
  V2 u = mk(c, s);
  V2 delta = lp - q;
  float dp = dot(delta, u);
  float sg = sgn(dp);
  float half = length / 2.f;
  float a = fabsf(dp);
  float m = limit ? (a < half ? a : half) : a;
  float sm = sg * m;
  return lp - mk(sm * u.x, sm * u.y);
}

// the 4 edges of a box as segments: physics.py:297-324 _get_all_lines_box
struct Edge { V2 p; float c, s, len; };
VB_DI Edge box_edge(int i, V2 pos, const Rot& r, float width, float length) {
  Edge e;
  float hl = length / 2.f, hw = width / 2.f;
  if (i == 0) { e.p = pos + mk(r.c, r.s) * hl; }
  else if (i == 1) { e.p = pos - mk(r.c, r.s) * hl; }
  else if (i == 2) { e.p = pos + mk(r.c2, r.s2) * hw; }
  else { e.p = pos - mk(r.c2, r.s2) * hw; }
  if (i <= 1) { e.c = r.c2; e.s = r.s2; e.len = width; }
  else { e.c = r.c; e.s = r.s; e.len = length; }
  return e;
}

// physics.py:262-294 _get_closest_point_box
VB_DI V2 closest_point_box(V2 pos, const Rot& r, float width, float length, V2 q) {
  V2 best = vinf();
  float dist = CUDART_INF_F;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    Edge e = box_edge(i, pos, r, width, length);
    V2 p = closest_point_line(e.p, e.c, e.s, e.len, q);
    float d = norm(q - p);
    if (d < dist) { best = p; dist = d; }
  }
  return best;
}

// physics.py:12-22 _get_inner_point_box
VB_DI V2 inner_point_box(V2 outside, V2 surface, V2 box_pos, float& d_out) {
  V2 v = surface - outside;
  V2 u = box_pos - surface;
  float vn = norm(v);
  float xm = dot(v, u) / vn;
  V2 x = (v / vn) * xm;
  if (vn == 0.f) { x = surface; xm = 0.f; }
  d_out = fabsf(xm);
  return surface + x;
}

// physics.py:131-259 segment-segment closest points (+ intersection override)
__device__ __noinline__ void closest_points_line_line(V2 pa, float ca, float sa, float la, V2 pb, float cb, float sb,
                                                      float lb, V2& o1, V2& o2) {
  V2 xya = mk((la / 2.f) * ca, (la / 2.f) * sa);
  V2 xyb = mk((lb / 2.f) * cb, (lb / 2.f) * sb);
  V2 a1 = pa + xya, a2 = pa - xya;
  V2 b1 = pb + xyb, b2 = pb - xyb;
  V2 r = a2 - a1, s = b2 - b1;
  V2 qp = b1 - a1;
  float cqpr = cross(qp, r), cqps = cross(qp, s), crs = cross(r, s);
  float u = cqpr / crs, t = cqps / crs;
  bool hit = (crs != 0.f) && (0.f <= u) && (u <= 1.f) && (0.f <= t) && (t <= 1.f);
  V2 pi = a1 + r * t;
  V2 best1 = vinf(), best2 = vinf();
  float md = CUDART_INF_F;
  {
    V2 c2 = closest_point_line(pb, cb, sb, lb, a1);
    float d = norm(a1 - c2);
    if (d < md) { best1 = a1; best2 = c2; md = d; }
  }
  {
    V2 c2 = closest_point_line(pb, cb, sb, lb, a2);
    float d = norm(a2 - c2);
    if (d < md) { best1 = a2; best2 = c2; md = d; }
  }
  {
    V2 c1 = closest_point_line(pa, ca, sa, la, b1);
    float d = norm(c1 - b1);
    if (d < md) { best1 = c1; best2 = b1; md = d; }
  }
  {
    V2 c1 = closest_point_line(pa, ca, sa, la, b2);
    float d = norm(c1 - b2);
    if (d < md) { best1 = c1; best2 = b2; md = d; }
  }
  if (hit) { best1 = pi; best2 = pi; }
  o1 = best1; o2 = best2;
}

// physics.py:327-381 _get_closest_line_box -> (point on box, point on line)
__device__ __noinline__ void closest_line_box(V2 bpos, const Rot& br, float bw, float bl, V2 lpos, float lc, float ls,
                                              float llen, V2& pbox, V2& pline) {
  V2 b1 = vinf(), b2 = vinf();
  float dist = CUDART_INF_F;
  for (int i = 0; i < 4; ++i) {
    Edge e = box_edge(i, bpos, br, bw, bl);
    V2 p_box, p_line;
    closest_points_line_line(e.p, e.c, e.s, e.len, lpos, lc, ls, llen, p_box, p_line);
    float d = norm(p_box - p_line);
    if (d < dist) { b1 = p_box; b2 = p_line; dist = d; }
  }
  pbox = b1; pline = b2;
}

// physics.py:25-128 _get_closest_box_box -> (point on box1, point on box2)
__device__ __noinline__ void closest_box_box(V2 p1, const Rot& r1, float w1, float l1, V2 p2, const Rot& r2, float w2,
                                             float l2, V2& o1, V2& o2) {
  V2 b1 = vinf(), b2 = vinf();
  float dist = CUDART_INF_F;
  for (int i = 0; i < 4; ++i) {  // box2 against the edges of box1
    Edge e = box_edge(i, p1, r1, w1, l1);
    V2 on_box2, on_edge1;
    closest_line_box(p2, r2, w2, l2, e.p, e.c, e.s, e.len, on_box2, on_edge1);
    float d = norm(on_edge1 - on_box2);
    if (d < dist) { b1 = on_edge1; b2 = on_box2; dist = d; }
  }
  for (int i = 0; i < 4; ++i) {  // box1 against the edges of box2
    Edge e = box_edge(i, p2, r2, w2, l2);
    V2 on_box1, on_edge2;
    closest_line_box(p1, r1, w1, l1, e.p, e.c, e.s, e.len, on_box1, on_edge2);
    float d = norm(on_box1 - on_edge2);
    if (d < dist) { b1 = on_box1; b2 = on_edge2; dist = d; }
  }
  o1 = b1; o2 = b2;
}

// core.py:2804-2838 _get_constraint_forces: force on a; the force on b is its exact negation.
VB_DI V2 constraint_force(V2 pa, V2 pb, float dist_min, float K, float k, bool attractive) {
  V2 delta = pa - pb;
  float dist = norm(delta);
  float sg = attractive ? -1.f : 1.f;
  float pen = softplus0((dist_min - dist) * sg / k) * k;
  float sk = attractive ? -K : K;
  float den = dist > 0.f ? dist : 1e-8f;
  V2 f = mk(sk * delta.x / den * pen, sk * delta.y / den * pen);
  if (dist < 1e-6f) f = mk(0.f, 0.f);
  if (!attractive) { if (dist > dist_min) f = mk(0.f, 0.f); }
  else { if (dist < dist_min) f = mk(0.f, 0.f); }
  return f;
}

// A joint's pair of calls (core.py:2236-2257: the attractive and the repulsive constraint force between the same two
// anchors) in one go: att = constraint_force(.., true), rep = constraint_force(.., false), bit for bit.  Their softplus
// arguments are exact negations of each other ((a * -1) / k == -(a / k) in IEEE arithmetic), and
// softplus0(x) = max(0, x) + log1p(exp(-|x|)): the transcendental term is the SAME number for both and is evaluated once
// (a third of a joint's instructions; joints are 18 % of joint_passage's).
VB_DI void constraint_force_pair(V2 pa, V2 pb, float dist_min, float K, float k, V2& att, V2& rep) {
  V2 delta = pa - pb;
  float dist = norm(delta);
  float xr = (dist_min - dist) * 1.f / k, xa = (dist_min - dist) * -1.f / k;
  float L = log1pf(expf(-fabsf(0.f - xr)));  // == log1pf(expf(-fabsf(0.f - xa)))
  float pen_r = ((0.f > xr ? 0.f : xr) + L) * k, pen_a = ((0.f > xa ? 0.f : xa) + L) * k;
  float den = dist > 0.f ? dist : 1e-8f;
  rep = mk(K * delta.x / den * pen_r, K * delta.y / den * pen_r);
  att = mk(-K * delta.x / den * pen_a, -K * delta.y / den * pen_a);
  if (dist < 1e-6f) { rep = mk(0.f, 0.f); att = mk(0.f, 0.f); }
  if (dist > dist_min) rep = mk(0.f, 0.f);
  if (dist < dist_min) att = mk(0.f, 0.f);
}

// core.py:2840-2857 _get_constraint_torques: torque on b; torque on a is its negation.
// ea = torch.exp(a), a = |rot_a - rot_b| (core.py:2848-2852).  exp(a) - 1 cancels for the small a of a held joint: one ulp
// of exp(a) is a relative 6e-8 / a of the torque, which is why a parity probe may hand in the reference's own value.
VB_DI float constraint_torque_exp(float d, float a, float ea, float Kt) {
  float pen = ea - 1.f;
  float tq = Kt * sgn(d) * pen;
  if (a < 1e-9f) tq = 0.f;
  return tq;
}
VB_DI float constraint_torque(float rot_a, float rot_b, float Kt) {
  float d = rot_a - rot_b;
  float a = sqrtf(d * d);
  return constraint_torque_exp(d, a, expf(a), Kt);
}

// get_friction_force, core.py:2054-2073
VB_DI V2 friction2(V2 vel, float coeff, float mass, float sub_dt) {
  float speed = norm(vel);
  bool st = speed == 0.f;
  float c = coeff * mass;
  float den = st ? 1e-8f : speed;
  float mx = (fabsf(vel.x) / sub_dt) * mass, my = (fabsf(vel.y) / sub_dt) * mass;
  V2 f = mk(-(vel.x / den) * (c < mx ? c : mx), -(vel.y / den) * (c < my ? c : my));
  if (st) f = mk(0.f, 0.f);
  return f;
}
VB_DI float friction1(float w, float coeff, float inertia, float sub_dt) {
  float speed = sqrtf(w * w);
  bool st = speed == 0.f;
  float c = coeff * inertia;
  float den = st ? 1e-8f : speed;
  float m = (fabsf(w) / sub_dt) * inertia;
  float f = -(w / den) * (c < m ? c : m);
  if (st) f = 0.f;
  return f;
}

// ---- LIDAR (core.py:1280-1371 box, :1413-1489 sphere, :1543-1625 line) ---------------------------------------
// torch.min/max over a stacked dim propagate NaN
VB_DI float pmin(float a, float b) { return (a != a || b != b) ? CUDART_NAN_F : (a < b ? a : b); }
VB_DI float pmax(float a, float b) { return (a != a || b != b) ? CUDART_NAN_F : (a > b ? a : b); }

// (dc, ds) = cos/sin of the world-frame ray angle; (bc, bs) = cos/sin of the box rotation
VB_DI float ray_box(V2 bpos, float bc, float bs, float blen, float bwid, V2 o, float dc, float ds, float max_range) {
  V2 pos_origin = o - bpos;
  float nbs = -bs;  // sin(-rot) == -sin(rot); cos(-rot) == cos(rot)
  V2 pos_aabb = rotv(pos_origin, bc, nbs);
  V2 dir = rotv(mk(dc, ds), bc, nbs);
  float tx1 = (-blen / 2.f - pos_aabb.x) / dir.x;
  float tx2 = (blen / 2.f - pos_aabb.x) / dir.x;
  float tmin = pmin(tx1, tx2), tmax = pmax(tx1, tx2);
  float ty1 = (-bwid / 2.f - pos_aabb.y) / dir.y;
  float ty2 = (bwid / 2.f - pos_aabb.y) / dir.y;
  float tymin = pmin(ty1, ty2), tymax = pmax(ty1, ty2);
  tmin = pmax(tmin, tymin);
  tmax = pmin(tmax, tymax);
  V2 ia = dir * tmin + pos_aabb;
  V2 iw = rotv(ia, bc, bs) + bpos;
  bool collision = (tmax >= tmin) && (tmin > 0.f);
  float dist = norm(o - iw);
  return collision ? dist : max_range;
}

VB_DI float ray_sphere(V2 spos, float rad, V2 o, float dc, float ds, float max_range) {
  V2 dir = mk(dc, ds);
  V2 line_pos = o + dir * (max_range / 2.f);
  V2 cp = closest_point_line(line_pos, dc, ds, max_range, spos, false);
  float dn = norm(spos - cp);
  bool hit = dn < rad;
  float a = rad * rad - dn * dn;
  float m = sqrtf(a > 0.f ? a : 1e-8f);
  bool front = dot(spos - o, dir) > 0.f;
  float dist = norm(cp - o) - m;
  return (hit && front) ? dist : max_range;
}

VB_DI float ray_line(V2 lpos, float lc, float ls, float llen, V2 o, float dc, float ds, float max_range) {
  V2 r = mk(lc * llen, ls * llen);
  V2 s = mk(dc, ds);
  float rxs = cross(r, s);
  V2 qp = o - lpos;
  float t = cross(qp, s / rxs);
  float u = cross(qp, r / rxs);
  float d = norm(mk(u * s.x, u * s.y));
  if (rxs == 0.f) d = max_range;
  if (t > 0.5f) d = max_range;
  if (t < -0.5f) d = max_range;
  if (u < 0.f) d = max_range;
  return d;
}

}  // namespace vb
