// queries.cuh - the distance queries reward code calls between steps: World.get_distance_from_point
// (core.py:1787-1819), World.get_distance (core.py:1821-1907), World.is_overlapping (core.py:1909-1966).
//
// One thread per environment; the two shapes are uniform over the launch, so the shape switch does not diverge.
// Poses are read through the tensors' own strides: an entity staged in a plan's SoA rows is read coalesced
// (pos_stride 1), any other torch tensor works too.  Bytes per env: 4 * (3 + 3 + 1) at most: HBM / launch bound.
#pragma once
#include "geom.cuh"

#include "../../include/vmasb200.h"

namespace vb {

struct QPose {  // vmasb200_pose by value, dimensions rounded the way torch rounds a Python scalar operand
  int shape;
  float radius, length, width;
  float overlap_min;  // (float)(radius + LINE_MIN_DIST), the sum taken in double (core.py:1960)
  const float* pos; long long ps, pxy;
  const float* rot; long long rs;
};

VB_DI V2 q_pos(const QPose& p, long long i) { return mk(p.pos[i * p.ps], p.pos[i * p.ps + p.pxy]); }
VB_DI Rot q_rot(const QPose& p, long long i) {
  Rot r;
  r.c = 1.f; r.s = 0.f; r.c2 = 0.f; r.s2 = 1.f;
  if (p.shape == VMASB200_SPHERE) return r;
  const float rot = p.rot[i * p.rs];
  sincosf(rot, &r.s, &r.c);
  if (p.shape == VMASB200_BOX) sincosf(rot + (float)(3.14159265358979323846 / 2), &r.s2, &r.c2);  // physics.py:299-301
  return r;
}

// core.py:1787-1819
VB_DI float dist_from_point(const QPose& e, V2 pos, const Rot& r, V2 q) {
  if (e.shape == VMASB200_SPHERE) return norm(pos - q) - e.radius;
  const V2 cp = e.shape == VMASB200_BOX ? closest_point_box(pos, r, e.width, e.length, q)
                                        : closest_point_line(pos, r.c, r.s, e.length, q);
  return norm(q - cp) - kLMD;
}

// core.py:1944-1966: box / sphere overlap
VB_DI bool overlap_box_sphere(const QPose& box, V2 bp, const Rot& br, const QPose& sph, V2 sp) {
  const V2 cp = closest_point_box(bp, br, box.width, box.length, sp);
  const float d_sphere_cp = norm(sp - cp);
  const float d_sphere_box = norm(sp - bp);
  const float d_cp_box = norm(bp - cp);
  return (d_sphere_box < d_cp_box) || (d_sphere_cp < sph.overlap_min);
}

// core.py:1821-1907
VB_DI float pair_distance(const QPose& a, const QPose& b, long long i) {
  const V2 pa = q_pos(a, i), pb = q_pos(b, i);
  const Rot ra = q_rot(a, i), rb = q_rot(b, i);
  const int sa = a.shape, sb = b.shape;
  if (sa == VMASB200_SPHERE && sb == VMASB200_SPHERE) return dist_from_point(a, pa, ra, pb) - b.radius;
  if ((sa == VMASB200_BOX && sb == VMASB200_SPHERE) || (sb == VMASB200_BOX && sa == VMASB200_SPHERE)) {
    const bool a_box = sb == VMASB200_SPHERE;
    const QPose& box = a_box ? a : b;
    const QPose& sph = a_box ? b : a;
    const V2 bp = a_box ? pa : pb, sp = a_box ? pb : pa;
    const Rot& br = a_box ? ra : rb;
    const float d = dist_from_point(box, bp, br, sp) - sph.radius;
    return overlap_box_sphere(box, bp, br, sph, sp) ? -1.f : d;
  }
  if ((sa == VMASB200_LINE && sb == VMASB200_SPHERE) || (sb == VMASB200_LINE && sa == VMASB200_SPHERE)) {
    const bool a_line = sb == VMASB200_SPHERE;
    const QPose& line = a_line ? a : b;
    const QPose& sph = a_line ? b : a;
    return dist_from_point(line, a_line ? pa : pb, a_line ? ra : rb, a_line ? pb : pa) - sph.radius;
  }
  V2 p1, p2;
  if (sa == VMASB200_LINE && sb == VMASB200_LINE) {
    closest_points_line_line(pa, ra.c, ra.s, a.length, pb, rb.c, rb.s, b.length, p1, p2);
  } else if (sa == VMASB200_BOX && sb == VMASB200_BOX) {
    closest_box_box(pa, ra, a.width, a.length, pb, rb, b.width, b.length, p1, p2);
  } else {  // box / line
    const bool a_box = sa == VMASB200_BOX;
    const QPose& box = a_box ? a : b;
    const QPose& line = a_box ? b : a;
    const Rot& lr = a_box ? rb : ra;
    closest_line_box(a_box ? pa : pb, a_box ? ra : rb, box.width, box.length, a_box ? pb : pa, lr.c, lr.s, line.length, p1, p2);
  }
  return norm(p1 - p2) - kLMD;
}

// core.py:1909-1966
VB_DI bool pair_overlap(const QPose& a, const QPose& b, long long i) {
  const bool bs = (a.shape == VMASB200_BOX && b.shape == VMASB200_SPHERE) || (b.shape == VMASB200_BOX && a.shape == VMASB200_SPHERE);
  if (bs) {  // core.py:1936-1964
    const bool a_box = b.shape == VMASB200_SPHERE;
    const QPose& box = a_box ? a : b;
    const QPose& sph = a_box ? b : a;
    return overlap_box_sphere(box, q_pos(box, i), q_rot(box, i), sph, q_pos(sph, i));
  }
  return pair_distance(a, b, i) < 0.f;  // core.py:1934
}

enum { QM_DISTANCE = 0, QM_OVERLAP = 1 };

template <int MODE>
__global__ void __launch_bounds__(256) pair_query_kernel(const __grid_constant__ QPose a, const __grid_constant__ QPose b,
                                                         float* __restrict__ out_f, uint8_t* __restrict__ out_b, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (MODE == QM_DISTANCE) {
    out_f[i] = pair_distance(a, b, i);
  } else {
    const bool r = pair_overlap(a, b, i);
    out_b[i] = r ? 1 : 0;
  }
}

__global__ void __launch_bounds__(256) point_query_kernel(const __grid_constant__ QPose a, const float* __restrict__ pts,
                                                          long long ps, long long pxy, float* __restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = dist_from_point(a, q_pos(a, i), q_rot(a, i), mk(pts[i * ps], pts[i * ps + pxy]));
}

// Environment._set_action, continuous physical actions (environment.py:615-668): one thread per (env, component).
struct ActParams { float range[8], mult[8]; };
__global__ void __launch_bounds__(256) set_action_kernel(const float* __restrict__ action, long long se, long long sk, int k_n,
                                                         const __grid_constant__ ActParams ap, int clamp,
                                                         float* __restrict__ u_out, int* __restrict__ flags, long long n) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int bad = 0;
  if (gid < n * k_n) {
    const long long i = gid / k_n;
    const int k = (int)(gid - i * k_n);
    float x = action[i * se + k * sk];
    const float r = ap.range[k];
    if (x != x) bad |= 1;                                                   // assert not action.isnan().any()
    if (clamp && x == x) x = fminf(fmaxf(x, -r), r);                         // torch.clamp keeps NaN
    if (fabsf(x) > r) bad |= 2;                                             // assert not any(|u| > u_range)
    u_out[gid] = x * ap.mult[k];                                            // agent.action.u *= u_multiplier_tensor
  }
  bad = __reduce_or_sync(0xffffffffu, bad);
  if (bad && (threadIdx.x & 31) == 0) atomicOr(flags, bad);
}

// Environment._set_action, continuous actions WITH communication and / or action noise (environment.py:615-668, 711-754):
// one thread per (env, column) over the physical columns and the communication columns.
//   physical column k: NaN check, clamp to +-u_range (clamp_action), range assertion, u = x * u_multiplier, then
//     u += randn * u_noise (the caller's torch.randn tensor: the same draw the reference makes, environment.py:712-720);
//   communication column c (source column comm_col0 + c: action_size when the reference has clamped and re-concatenated
//     the action, world.dim_p otherwise - environment.py:633-645, 742-743): NaN check, assertion 0 <= c <= 1 (flag bit 3;
//     never clamped, see below), c += randn * c_noise.
struct ActExParams { float range[8], mult[8]; int k_n, comm_col0, n_comm, clamp; float u_noise, c_noise; };
__global__ void __launch_bounds__(256) set_action_ex_kernel(const float* __restrict__ action, long long se, long long sk,
                                                            const __grid_constant__ ActExParams ap,
                                                            const float* __restrict__ u_randn, const float* __restrict__ c_randn,
                                                            float* __restrict__ u_out, float* __restrict__ c_out,
                                                            int* __restrict__ flags, long long n) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cols = ap.k_n + ap.n_comm;
  int bad = 0;
  if (gid < n * cols) {
    const long long i = gid / cols;
    const int k = (int)(gid - i * cols);
    if (k < ap.k_n) {
      float x = action[i * se + k * sk];
      const float r = ap.range[k];
      if (x != x) bad |= 1;
      if (ap.clamp && x == x) x = fminf(fmaxf(x, -r), r);
      if (fabsf(x) > r) bad |= 2;
      float u = x * ap.mult[k];
      if (u_randn) u = u + u_randn[i * ap.k_n + k] * ap.u_noise;
      u_out[i * ap.k_n + k] = u;
    } else {
      const int c = k - ap.k_n;
      float x = action[i * se + (ap.comm_col0 + c) * sk];
      if (x != x) bad |= 1;
      // (with clamp_action the reference concatenates comm.clamp(0, 1) into `action`, but keeps asserting on and storing
      //  the UNCLAMPED slice it took before, environment.py:640-645, 737-743: communication actions are never clamped)
      if (x > 1.f || x < 0.f) bad |= 8;                                       // "Comm actions are out of range [0,1]"
      if (c_randn) x = x + c_randn[i * ap.n_comm + c] * ap.c_noise;
      c_out[i * ap.n_comm + c] = x;
    }
  }
  bad = __reduce_or_sync(0xffffffffu, bad);
  if (bad && (threadIdx.x & 31) == 0) atomicOr(flags, bad);
}

// ---- reset path
// World.reset(env_index) (core.py:1178-1180 -> EntityState._reset, core.py:285-295) on the SoA buffers, IN PLACE: the rows
// [row0, row0 + n_rows) of one buffer are zeroed for env `env_index`, for every env (env_index < 0), or for the envs whose
// mask byte is set.  One thread per (row, env), env fastest (coalesced).
__global__ void __launch_bounds__(256) reset_rows_kernel(float* __restrict__ buf, int n_rows, long long ld, long long n,
                                                         long long env_index, const uint8_t* __restrict__ mask) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (env_index >= 0) {  // one column
    if (gid < n_rows) buf[gid * ld + env_index] = 0.f;
    return;
  }
  if (gid >= (long long)n_rows * n) return;
  const long long r = gid / n, i = gid - r * n;
  if (!mask || mask[i]) buf[r * ld + i] = 0.f;
}

// One iteration of the rejection loop of ScenarioUtils.find_random_pos_for_entity (utils.py:301-306):
//   dist = torch.cdist(occupied [B,K,2], pos [B,1,2]); overlaps = (dist < min_dist).any(K)
//   pos[overlaps] = proposed[overlaps]; and the loop's own `overlaps.any()` into *flag.
// cdist's p = 2 kernel accumulates diff * diff per coordinate and takes the root: sqrt(fma(dy, dy, dx * dx)).
__global__ void __launch_bounds__(256) spawn_select_kernel(float* __restrict__ pos, const float* __restrict__ proposed,
                                                           const float* __restrict__ occupied, int K, float min_dist,
                                                           int* __restrict__ flag, long long B) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int over = 0;
  if (b < B) {
    const float px = pos[2 * b], py = pos[2 * b + 1];
    for (int k = 0; k < K; ++k) {
      const float dx = occupied[(b * K + k) * 2] - px, dy = occupied[(b * K + k) * 2 + 1] - py;
      over |= sqrtf(fmaf(dy, dy, dx * dx)) < min_dist;
    }
    if (over) { pos[2 * b] = proposed[2 * b]; pos[2 * b + 1] = proposed[2 * b + 1]; }
  }
  over = __reduce_or_sync(0xffffffffu, over);
  if (over && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// Environment._set_action, discrete physical actions (environment.py:660-707): one thread per (env, component).
//   flat action (one column): component k = (flat mod prod(nvec[k:])) div prod(nvec[k+1:]) with Python's floor
//   semantics (component 0 is not reduced: an action beyond the product fails the range check, as in the reference);
//   multi-discrete: component k is column k.
//   range check [0, n) on the raw component (environment.py:608-612) -> flag bit 2;
//   odd n: 0 -> n div 2, 1..n div 2 -> one less (the first action maps to u = 0);
//   u = (float(a) / (n - 1)) * (2 u_max) - u_max, then u *= u_multiplier (fp32, the reference's op order).
struct DiscParams { int nvec[8]; long long below[8]; float range[8], mult[8]; };
VB_DI long long floor_div(long long a, long long b) { long long q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }
template <class AT>
__global__ void __launch_bounds__(256) set_action_discrete_kernel(const AT* __restrict__ action, long long se, long long sk, int k_n,
                                                                  int multi, const __grid_constant__ DiscParams dp,
                                                                  float* __restrict__ u_out, int* __restrict__ flags, long long n) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int bad = 0;
  if (gid < n * k_n) {
    const long long i = gid / k_n;
    const int k = (int)(gid - i * k_n);
    long long a;
    if (multi) {
      a = (long long)action[i * se + k * sk];
    } else {
      long long flat = (long long)action[i * se];
      if (k > 0) {  // flat % prod(nvec[k:]) (sign of the divisor), then // prod(nvec[k+1:])
        const long long m = dp.below[k] * dp.nvec[k];
        flat -= floor_div(flat, m) * m;
      }
      a = floor_div(flat, dp.below[k]);
    }
    const int nk = dp.nvec[k];
    if (a < 0 || a >= nk) bad |= 4;
    if (nk & 1) {
      if (a == 0) a = nk / 2;
      else if (a > 0 && a <= nk / 2) a -= 1;
    }
    const float um = dp.range[k];
    const float u = ((float)a / (float)(nk - 1)) * (2.f * um) - um;
    u_out[gid] = u * dp.mult[k];
  }
  bad = __reduce_or_sync(0xffffffffu, bad);
  if (bad && (threadIdx.x & 31) == 0) atomicOr(flags, bad);
}

}  // namespace vb
