// kernels.cuh - __global__ entry points: fused step, batch-wide broad-phase flags, standalone cast_rays.
#pragma once
#include "step_kernel.cuh"

namespace vb {

struct StepArgs {
  const float* dyn_in;
  float* dyn_out;
  const float* stat;
  float* act_out;
  float* lidar_out;
  long long n_envs;
  int s_begin, s_end;        // substeps [s_begin, s_end) of this launch
  const uint32_t* flags_in;  // [S][pair_words] batch-wide World.collides verdicts, or NULL = treat every pair as active
  uint32_t* any_out;         // [S][pair_words] OR over envs of the env's own broad-phase verdict (or NULL)
  uint32_t* thin_out;        // [S][pair_words] OR over envs of (own verdict false AND force applied != 0) (or NULL)
  int do_lidar;
};

VB_DI int align4(int x) { return (x + 3) & ~3; }

// dynamic shared memory carve-up (floats unless noted); every tile array starts 16B aligned
struct SmemMap {
  int bar, dyn, stat, act, lid, work, bits, total_bytes;
};
__host__ __device__ inline SmemMap smem_map(const KParams& kp) {
  SmemMap m;
  int o = 0;
  m.bar = o; o += 4;  // 16 bytes for the mbarrier
  m.dyn = o; o += kp.TE * kp.dyn_stride;
  m.stat = o; o += kp.TE * kp.stat_stride;
  m.act = o; o += kp.TE * kp.act_stride;
  m.lid = o; o += kp.TE * kp.lidar_stride;
  m.work = o; o += ((kp.TE * kp.work_stride + 3) & ~3);
  m.bits = o; o += 2 * kp.S * kp.pair_words;
  m.total_bytes = o * 4;
  return m;
}

__global__ void __launch_bounds__(256) step_kernel(const __grid_constant__ KParams kp, const __grid_constant__ AgentIO aio,
                                                   const __grid_constant__ StepArgs sa) {
  extern __shared__ __align__(16) float smem[];
  const SmemMap m = smem_map(kp);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + m.bar);
  float* dynS = smem + m.dyn;
  float* statS = smem + m.stat;
  float* actS = smem + m.act;
  float* lidS = smem + m.lid;
  float* workS = smem + m.work;
  uint32_t* anyS = reinterpret_cast<uint32_t*>(smem + m.bits);
  uint32_t* thinS = anyS + kp.S * kp.pair_words;

  const int tid = threadIdx.x, nthr = blockDim.x;
  const int G = kp.G, TE = kp.TE;
  const long long env0 = (long long)blockIdx.x * TE;
  const int n_valid = (int)min((long long)TE, sa.n_envs - env0);
  const int el = tid / G, lg = tid % G;  // env inside the tile, lane inside the env group
  const bool env_ok = el < n_valid;

  // ---- stage the tile: TMA bulk copies for the env records, plain loads for the per-agent force tensors
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  __syncthreads();
  if (tid == 0) {
    uint32_t bd = (uint32_t)n_valid * kp.dyn_stride * 4u, bs = (uint32_t)n_valid * kp.stat_stride * 4u;
    mbar_expect_tx(bar, bd + bs);
    bulk_g2s(dynS, sa.dyn_in + env0 * kp.dyn_stride, bd, bar);
    if (bs) bulk_g2s(statS, sa.stat + env0 * kp.stat_stride, bs, bar);
  }
  for (int w = tid; w < 2 * kp.S * kp.pair_words; w += nthr) anyS[w] = 0u;
  for (int idx = tid; idx < TE * kp.A; idx += nthr) {
    int e = idx % TE, a = idx / TE;
    float fx = 0.f, fy = 0.f, tq = 0.f;
    if (e < n_valid) {
      const float* fp = aio.force[a];
      const float* tp = aio.torque[a];
      if (fp) { fx = fp[(env0 + e) * aio.fstride[a]]; fy = fp[(env0 + e) * aio.fstride[a] + 1]; }
      if (tp) tq = tp[(env0 + e) * aio.tstride[a]];
    }
    float* fr = workS + e * kp.work_stride + kp.off_frc + 3 * a;
    fr[0] = fx; fr[1] = fy; fr[2] = tq;
  }
  mbar_wait(bar, 0);
  __syncthreads();

  EnvPtrs ep;
  ep.dyn = dynS + el * kp.dyn_stride;
  ep.stat = statS + el * kp.stat_stride;
  ep.work = workS + el * kp.work_stride;

  // ---- prologue: cos/sin cache of every entity that needs it, first force accumulators
  for (int r = lg; r < kp.n_rot; r += G) {
    const DEnt* en = kp.ents + __ldg(kp.rot_list + r);
    V2 p; float rot;
    pose(en, ep, p, rot);
    store_trig(kp, en, ep, rot);
  }
  // the clamp of core.py:2017-2040 is re-applied to its own result every substep: replay it for the substeps
  // that earlier launches of a split step already did
  for (int rep = 0; rep <= sa.s_begin; ++rep)
    for (int d = lg; d < kp.n_dyn; d += G) begin_forces(kp, kp.ents + __ldg(kp.dyn_list + d), ep);
  __syncwarp();

  float4* res = reinterpret_cast<float4*>(ep.work + kp.off_res);
  for (int s = sa.s_begin; s < sa.s_end; ++s) {
    // ---- phase B: joints and pairs
    for (int i = lg; i < kp.n_items; i += G) {
      const DItem* it = kp.items + i;
      const int p = i - kp.J;  // pair index (flag bit)
      bool active = true;
      if (p >= 0 && sa.flags_in) active = (sa.flags_in[s * kp.pair_words + (p >> 5)] >> (p & 31)) & 1u;
      V2 fa = mk(0.f, 0.f);
      float ta = 0.f, tb = 0.f;
      if (active) {
        bool bp = eval_item(kp, it, ep, fa, ta, tb);
        if (p >= 0 && env_ok) {
          const uint32_t bit = 1u << (p & 31);
          const int w = s * kp.pair_words + (p >> 5);
          if (bp) {
            if (!(anyS[w] & bit)) atomicOr(&anyS[w], bit);
          } else if (fa.x != 0.f || fa.y != 0.f) {
            if (!(thinS[w] & bit)) atomicOr(&thinS[w], bit);
          }
        }
      }
      res[i] = make_float4(fa.x, fa.y, ta, tb);
    }
    __syncwarp();
    // ---- phase C: accumulate + integrate, then open the next substep's accumulators
    for (int d = lg; d < kp.n_dyn; d += G) {
      const DEnt* en = kp.ents + __ldg(kp.dyn_list + d);
      finish_entity(kp, en, ep, s);
      if (s + 1 < sa.s_end) begin_forces(kp, en, ep);
    }
    __syncwarp();
  }

  // ---- epilogue: LIDAR, clamped action write-back
  if (sa.do_lidar && sa.lidar_out) {
    float* lid = lidS + el * kp.lidar_stride;
    for (int r = lg; r < kp.n_rays; r += G) {
      const DRay* ry = kp.rays + r;
      const DSensor* sn = kp.sensors + ry->sensor;
      V2 o; float rot;
      pose(kp.ents + sn->entity, ep, o, rot);
      lid[ry->out_off] = cast_one_ray(kp, sn, ep, o, ry->angle + rot);  // sensors.py:117
    }
  }
  if (sa.act_out) {
    float* act = actS + el * kp.act_stride;
    for (int d = lg; d < kp.n_dyn; d += G) {
      const DEnt* en = kp.ents + __ldg(kp.dyn_list + d);
      if (en->act_off >= 0) {
        const float* fr = ep.work + kp.off_frc + 3 * en->agent;
        act[en->act_off] = fr[0]; act[en->act_off + 1] = fr[1]; act[en->act_off + 2] = fr[2];
      }
    }
  }
  fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    bulk_s2g(sa.dyn_out + env0 * kp.dyn_stride, dynS, (uint32_t)n_valid * kp.dyn_stride * 4u);
    if (sa.act_out && kp.act_stride) bulk_s2g(sa.act_out + env0 * kp.act_stride, actS, (uint32_t)n_valid * kp.act_stride * 4u);
    if (sa.do_lidar && sa.lidar_out && kp.lidar_stride)
      bulk_s2g(sa.lidar_out + env0 * kp.lidar_stride, lidS, (uint32_t)n_valid * kp.lidar_stride * 4u);
    bulk_commit();
  }
  // batch-wide flags: CTA-level OR already done in smem, one atomic per word that still adds a bit
  const int nw = kp.S * kp.pair_words;
  if (sa.any_out)
    for (int w = tid; w < nw; w += nthr) {
      uint32_t v = anyS[w];
      if (v & ~__ldcg(sa.any_out + w)) atomicOr(sa.any_out + w, v);
    }
  if (sa.thin_out)
    for (int w = tid; w < nw; w += nthr) {
      uint32_t v = thinS[w];
      if (v & ~__ldcg(sa.thin_out + w)) atomicOr(sa.thin_out + w, v);
    }
  if (tid == 0) bulk_wait_all();
}

// World.collides' batch-wide distance test (core.py:2796-2800) for substep s on the records in HBM:
// flags[s][p] = any over envs of ||p_a - p_b|| <= Rc_a + Rc_b.  One thread per env, ballot per pair.
__global__ void __launch_bounds__(256) flags_kernel(const __grid_constant__ KParams kp, const float* __restrict__ dyn,
                                                    const float* __restrict__ stat, long long n_envs, int s,
                                                    uint32_t* __restrict__ flags) {
  const long long env = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = env < n_envs;
  const float* dr = dyn + (ok ? env : 0) * kp.dyn_stride;
  const float* sr = stat + (ok ? env : 0) * kp.stat_stride;
  for (int p = 0; p < kp.P; ++p) {
    const DItem* it = kp.items + kp.J + p;
    const DEnt* ea = kp.ents + it->a;
    const DEnt* eb = kp.ents + it->b;
    V2 pa = ea->dyn_off >= 0 ? mk(dr[ea->dyn_off], dr[ea->dyn_off + 1]) : mk(sr[ea->stat_off], sr[ea->stat_off + 1]);
    V2 pb = eb->dyn_off >= 0 ? mk(dr[eb->dyn_off], dr[eb->dyn_off + 1]) : mk(sr[eb->stat_off], sr[eb->stat_off + 1]);
    bool bp = ok && (norm(pa - pb) <= it->thr);
    unsigned any = __ballot_sync(0xffffffffu, bp);
    if (any && (threadIdx.x & 31) == 0) {
      uint32_t* w = flags + s * kp.pair_words + (p >> 5);
      uint32_t bit = 1u << (p & 31);
      if (!(__ldcg(w) & bit)) atomicOr(w, bit);
    }
  }
}

struct RayTargets {  // by-value kernel parameter: the (already filtered) target entities of one cast_rays call
  int n;
  int e[64];
};

// World.cast_rays (core.py:1661-1785) with caller-provided world-frame angles and target list.
// One thread per (env, ray); records are read straight from HBM (L1/L2 serve the 1-to-n_rays reuse).
__global__ void __launch_bounds__(256) cast_rays_kernel(const __grid_constant__ KParams kp, const float* __restrict__ dyn,
                                                        const float* __restrict__ stat, int entity,
                                                        const float* __restrict__ angles, long long angles_stride,
                                                        int n_rays, float max_range, const __grid_constant__ RayTargets rt,
                                                        float* __restrict__ out, long long n_envs) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= n_envs * n_rays) return;
  const long long env = gid / n_rays;
  const int r = (int)(gid % n_rays);
  const float* dr = dyn + env * kp.dyn_stride;
  const float* sr = stat + env * kp.stat_stride;
  auto getpose = [&](const DEnt* en, V2& p, float& rot) {
    if (en->dyn_off >= 0) { p = mk(dr[en->dyn_off], dr[en->dyn_off + 1]); rot = dr[en->dyn_off + 4]; }
    else { p = mk(sr[en->stat_off], sr[en->stat_off + 1]); rot = sr[en->stat_off + 2]; }
  };
  V2 o; float orot;
  getpose(kp.ents + entity, o, orot);
  float ds, dc;
  sincosf(angles[env * angles_stride + r], &ds, &dc);
  float best = max_range;
  for (int t = 0; t < rt.n; ++t) {
    const DEnt* en = kp.ents + rt.e[t];
    V2 p; float rot;
    getpose(en, p, rot);
    float d;
    if (en->shape == VMASB200_SPHERE) d = ray_sphere(p, en->radius, o, dc, ds, max_range);
    else {
      float s, c;
      sincosf(rot, &s, &c);
      if (en->shape == VMASB200_BOX) d = ray_box(p, c, s, en->length, en->width, o, dc, ds, max_range);
      else d = ray_line(p, c, s, en->length, o, dc, ds, max_range);
    }
    best = pmin(best, d);
  }
  out[env * n_rays + r] = best;
}

}  // namespace vb
