// kernels.cuh - __global__ entry points: fused step (+ verified repair), batch-wide broad-phase flags, cast_rays.
//
// Exactness of the fused (speculative) launch.  The reference decides per (substep, pair) for the WHOLE batch whether
// the pair is evaluated: World.collides ends in (||p_a - p_b|| <= Rc_a + Rc_b).any() over envs (core.py:2796-2800).
// The fused kernel evaluates every pair in every env.  That differs from the reference only for an env whose own
// verdict is false while its force is non-zero ("thin" event: contact through the LINE_MIN_DIST shell at a box
// corner / line end) AND no env of the batch has a true verdict.  Each env therefore tracks a taint bit (set after
// its first thin event); untainted envs OR their verdicts into any_exact, tainted ones into any_spec.  The last CTA
// to finish checks thin & ~any_exact == 0, which proves (by induction over substeps) that every env equals the
// reference; otherwise it re-simulates only the tainted envs from their saved input records, substep by substep,
// with flags = any_exact | OR over the re-simulated envs' verdicts - which is exactly the reference's flag.
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
  const uint32_t* flags_in;  // [S][pair_words] batch-wide World.collides verdicts (exact path), or NULL = speculative
  // speculative mode scratch (all owned by the plan, zero between launches)
  uint32_t* flags_out;       // [S][pair_words] result: the reference's verdicts
  uint32_t* any_exact;       // [S][pair_words]
  uint32_t* any_spec;        // [S][pair_words]
  uint32_t* thin;            // [S][pair_words]
  int* counters;             // [0] finished CTAs, [1] tainted envs, [2] repairs done since plan creation
  int* u_list;               // [n_envs] env index of every tainted env
  float* u_buf;              // [n_envs][dyn_stride] their input records (then their evolving state)
  int do_lidar;
};

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
  m.bits = o; o += 3 * kp.S * kp.pair_words;
  m.total_bytes = o * 4;
  return m;
}

struct Tile {
  float *dynS, *statS, *actS, *lidS, *workS;
  uint32_t *bitsE, *bitsS, *bitsT;  // any_exact / any_spec / thin of this CTA
};

VB_DI void set_bit(uint32_t* words, int w, uint32_t bit) {
  if (!(words[w] & bit)) atomicOr(&words[w], bit);
}

// agent.state.force / torque of env `env` into the env's work area
VB_DI void load_forces(const KParams& kp, const AgentIO& aio, float* workE, long long env, int a) {
  float fx = 0.f, fy = 0.f, tq = 0.f;
  const float* fp = aio.force[a];
  const float* tp = aio.torque[a];
  if (fp) { fx = fp[env * aio.fstride[a]]; fy = fp[env * aio.fstride[a] + 1]; }
  if (tp) tq = tp[env * aio.tstride[a]];
  float* fr = workE + kp.off_frc + 3 * a;
  fr[0] = fx; fr[1] = fy; fr[2] = tq;
}

// Substeps [s_begin, s_end) of one env on the staged tile.  flags: batch-wide verdicts to obey (smem or global), or
// NULL = speculative (every pair considered, verdict / thin bits collected into the tile's bit arrays).
// Returns the env's taint (speculative mode).
VB_DI bool run_env(const KParams& kp, const Tile& t, int el, int lg, bool env_ok, int s_begin, int s_end,
                   const uint32_t* flags, bool do_lidar) {
  const int G = kp.G;
  EnvPtrs ep;
  ep.dyn = t.dynS + el * kp.dyn_stride;
  ep.stat = t.statS + el * kp.stat_stride;
  ep.work = t.workS + el * kp.work_stride;
  const int gshift = (threadIdx.x & 31) & ~(G - 1);
  const unsigned gbits = G == 32 ? 0xffffffffu : ((1u << G) - 1u);
  const unsigned gmask = gbits << gshift;
  const unsigned below = (1u << lg) - 1u;

  // prologue: cos/sin cache of every entity that needs it, first force accumulators, joints at the list head
  for (int r = lg; r < kp.n_rot; r += G) {
    const DEnt* en = kp.ents + __ldg(kp.rot_list + r);
    V2 p; float rot;
    pose(en, ep, p, rot);
    store_trig(kp, en, ep, rot);
  }
  // the clamp of core.py:2017-2040 is re-applied to its own result every substep: replay it for the substeps
  // that earlier launches of a split step already did
  for (int rep = 0; rep <= s_begin; ++rep)
    for (int d = lg; d < kp.n_dyn; d += G) begin_forces(kp, kp.ents + __ldg(kp.dyn_list + d), ep);
  int* list = reinterpret_cast<int*>(ep.work + kp.off_list);
  for (int j = lg; j < kp.J; j += G) {
    const DItem* it = kp.items + j;
    list[j] = j | (it->a << kEntryItemBits) | (it->b << (kEntryItemBits + kEntryEntBits));
  }
  __syncwarp();

  bool tainted = false;
  float4* res = reinterpret_cast<float4*>(ep.work + kp.off_res);
  for (int s = s_begin; s < s_end; ++s) {
    // ---- phase A: broad phase + compaction of the surviving pairs (all lanes of the warp run the same trip count)
    int count = kp.J;
    for (int base = 0; base < kp.P; base += G) {
      const int p = base + lg;
      bool surv = false;
      int entry = 0;
      if (p < kp.P) {
        const int4 bi = __ldg(reinterpret_cast<const int4*>(kp.broad + p));
        const float4 bf = __ldg(reinterpret_cast<const float4*>(kp.broad + p) + 1);
        const V2 pa = pos_enc(ep, bi.x), pb = pos_enc(ep, bi.y);
        const float dx = pa.x - pb.x, dy = pa.y - pb.y;
        const float d2 = fmaf(dy, dy, dx * dx);
        const bool bp = d2 <= bf.x;  // == (norm(pa - pb) <= Rc_a + Rc_b), core.py:2796-2800
        const int w = s * kp.pair_words + (p >> 5);
        const uint32_t bit = 1u << (p & 31);
        bool active = true;
        if (flags) active = (flags[w] & bit) != 0u;
        else if (bp && env_ok) set_bit(tainted ? t.bitsS : t.bitsE, w, bit);
        if (bi.z < 0) {
          surv = !(d2 > bf.y);
        } else {  // sphere centre in the line / box frame
          const float c = ep.work[bi.z], sn = ep.work[bi.z + 1];
          const float lx = -(dx * c + dy * sn), ly = dx * sn - dy * c;
          surv = !(fabsf(lx) > bf.z) && !(fabsf(ly) > bf.w);
        }
        surv = surv && active;
        entry = bi.w | (bp ? (1 << 30) : 0);
      }
      const unsigned votes = (__ballot_sync(0xffffffffu, surv) >> gshift) & gbits;
      if (surv) list[count + __popc(votes & below)] = entry;
      count += __popc(votes);
    }
    __syncwarp();
    // ---- phase B: joints and surviving pairs
    bool thin_now = false;
    for (int k = lg; k < count; k += G) {
      const int entry = list[k];
      const int i = entry & ((1 << kEntryItemBits) - 1);
      V2 fa = mk(0.f, 0.f);
      float ta = 0.f, tb = 0.f;
      eval_item(kp, kp.items + i, ep, fa, ta, tb);
      if (!flags && i >= kp.J && env_ok && !(entry & (1 << 30)) && (fa.x != 0.f || fa.y != 0.f)) {
        const int p = i - kp.J;
        set_bit(t.bitsT, s * kp.pair_words + (p >> 5), 1u << (p & 31));
        thin_now = true;
      }
      res[k] = make_float4(fa.x, fa.y, ta, tb);
    }
    // warp-wide vote doubles as the phase barrier; the taint applies from the NEXT substep on
    const unsigned vote = __ballot_sync(0xffffffffu, thin_now);
    tainted = tainted || (vote & gmask);
    // ---- phase C: accumulate + integrate, then open the next substep's accumulators
    for (int d = lg; d < kp.n_dyn; d += G) {
      const int e = __ldg(kp.dyn_list + d);
      const DEnt* en = kp.ents + e;
      finish_entity(kp, en, e, ep, s, list, count);
      if (s + 1 < s_end) begin_forces(kp, en, ep);
    }
    __syncwarp();
  }

  if (do_lidar) {  // LIDAR epilogue
    uint32_t* lm = reinterpret_cast<uint32_t*>(ep.work + kp.off_lmask);
    for (int sw = lg; sw < kp.n_lmask_words; sw += G) {  // which targets can a ray of this sensor reach at all
      const DSWord* q = kp.swords + sw;
      const V2 o = posof(kp.ents + kp.sensors[q->sensor].entity, ep);
      uint32_t m = 0u;
      for (int k = 0; k < q->n; ++k) {
        const V2 p = posof(kp.ents + __ldg(kp.targets + q->tg_begin + k), ep);
        const float dx = o.x - p.x, dy = o.y - p.y;
        if (!(fmaf(dy, dy, dx * dx) > __ldg(kp.tcull2 + q->tg_begin + k))) m |= 1u << k;
      }
      lm[sw] = m;
    }
    __syncwarp();
    float* lid = t.lidS + el * kp.lidar_stride;
    for (int r = lg; r < kp.n_rays; r += G) {
      const DRay* ry = kp.rays + r;
      const DSensor* sn = kp.sensors + ry->sensor;
      V2 o; float rot;
      pose(kp.ents + sn->entity, ep, o, rot);
      lid[ry->out_off] = cast_one_ray(kp, sn, ep, o, ry->angle + rot, lm + sn->tw_off);  // sensors.py:117
    }
  }
  return tainted;
}

// Re-simulation of the tainted envs by the last CTA (rare path; see the header comment).  flagS (smem, [S][pair_words])
// starts as any_exact and ends as the reference's verdicts.
__device__ __noinline__ void repair_tainted(const KParams& kp, const AgentIO& aio, const StepArgs& sa, const Tile& t,
                                            uint32_t* flagS, int nu) {
  const int tid = threadIdx.x, nthr = blockDim.x, G = kp.G, TE = kp.TE;
  const int el = tid / G, lg = tid % G;
  const int nw = kp.S * kp.pair_words;
  for (int w = tid; w < nw; w += nthr) flagS[w] = __ldcg(sa.any_exact + w);
  __syncthreads();
  for (int s = 0; s <= kp.S; ++s) {  // s == S: LIDAR + write-back pass
    if (s < kp.S) {
      // pass 1: verdicts of the tainted envs at their true state
      for (int base = 0; base < nu; base += TE) {
        const int idx = base + el;
        if (idx < nu) {
          EnvPtrs ep;
          ep.dyn = sa.u_buf + (size_t)idx * kp.dyn_stride;
          ep.stat = sa.stat + (size_t)sa.u_list[idx] * kp.stat_stride;
          ep.work = nullptr;
          for (int p = lg; p < kp.P; p += G) {
            const DBroad* bq = kp.broad + p;
            V2 pa = pos_enc(ep, bq->pa_off), pb = pos_enc(ep, bq->pb_off);
            const float dx = pa.x - pb.x, dy = pa.y - pb.y;
            if (fmaf(dy, dy, dx * dx) <= bq->t2_bp) set_bit(flagS, s * kp.pair_words + (p >> 5), 1u << (p & 31));
          }
        }
      }
      __syncthreads();
    }
    // pass 2: one substep with the now exact flags (or the epilogue)
    for (int base = 0; base < nu; base += TE) {
      const int nv = min(TE, nu - base);
      for (int i = tid; i < nv * kp.dyn_stride; i += nthr) t.dynS[i] = sa.u_buf[(size_t)base * kp.dyn_stride + i];
      for (int i = tid; i < nv * kp.stat_stride; i += nthr) {
        int e = i / kp.stat_stride, k = i % kp.stat_stride;
        t.statS[i] = sa.stat[(size_t)sa.u_list[base + e] * kp.stat_stride + k];
      }
      for (int i = tid; i < nv * kp.A; i += nthr) {
        int e = i % nv, a = i / nv;
        load_forces(kp, aio, t.workS + e * kp.work_stride, sa.u_list[base + e], a);
      }
      __syncthreads();
      const bool last = s == kp.S;
      run_env(kp, t, el, lg, el < nv, s, last ? s : s + 1, flagS, last && sa.lidar_out);
      __syncthreads();
      if (!last) {
        for (int i = tid; i < nv * kp.dyn_stride; i += nthr) sa.u_buf[(size_t)base * kp.dyn_stride + i] = t.dynS[i];
      } else {
        for (int i = tid; i < nv * kp.dyn_stride; i += nthr) {
          int e = i / kp.dyn_stride, k = i % kp.dyn_stride;
          sa.dyn_out[(size_t)sa.u_list[base + e] * kp.dyn_stride + k] = t.dynS[i];
        }
        if (sa.lidar_out)
          for (int i = tid; i < nv * kp.lidar_stride; i += nthr) {
            int e = i / kp.lidar_stride, k = i % kp.lidar_stride;
            sa.lidar_out[(size_t)sa.u_list[base + e] * kp.lidar_stride + k] = t.lidS[i];
          }
      }
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(256) step_kernel(const __grid_constant__ KParams kp, const __grid_constant__ AgentIO aio,
                                                   const __grid_constant__ StepArgs sa) {
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_last;
  const SmemMap m = smem_map(kp);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + m.bar);
  Tile t;
  t.dynS = smem + m.dyn; t.statS = smem + m.stat; t.actS = smem + m.act; t.lidS = smem + m.lid; t.workS = smem + m.work;
  const int nw = kp.S * kp.pair_words;
  t.bitsE = reinterpret_cast<uint32_t*>(smem + m.bits); t.bitsS = t.bitsE + nw; t.bitsT = t.bitsS + nw;

  const int tid = threadIdx.x, nthr = blockDim.x;
  const int G = kp.G, TE = kp.TE;
  const long long env0 = (long long)blockIdx.x * TE;
  const int n_valid = (int)min((long long)TE, sa.n_envs - env0);
  const int el = tid / G, lg = tid % G;  // env inside the tile, lane inside the env group
  const bool env_ok = el < n_valid;
  const bool spec = sa.flags_in == nullptr;

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
    bulk_g2s(t.dynS, sa.dyn_in + env0 * kp.dyn_stride, bd, bar);
    if (bs) bulk_g2s(t.statS, sa.stat + env0 * kp.stat_stride, bs, bar);
  }
  for (int w = tid; w < 3 * nw; w += nthr) t.bitsE[w] = 0u;
  for (int idx = tid; idx < n_valid * kp.A; idx += nthr) {
    int e = idx % n_valid, a = idx / n_valid;
    load_forces(kp, aio, t.workS + e * kp.work_stride, env0 + e, a);
  }
  mbar_wait(bar, 0);
  __syncthreads();

  const bool tainted = run_env(kp, t, el, lg, env_ok, sa.s_begin, sa.s_end, sa.flags_in, sa.do_lidar && sa.lidar_out);

  if (sa.act_out) {  // clamped action write-back (core.py:2019-2026, 2031-2038)
    float* act = t.actS + el * kp.act_stride;
    const float* work = t.workS + el * kp.work_stride;
    for (int d = lg; d < kp.n_dyn; d += G) {
      const DEnt* en = kp.ents + __ldg(kp.dyn_list + d);
      if (en->act_off >= 0) {
        const float* fr = work + kp.off_frc + 3 * en->agent;
        act[en->act_off] = fr[0]; act[en->act_off + 1] = fr[1]; act[en->act_off + 2] = fr[2];
      }
    }
  }
  if (spec && tainted && env_ok) {  // keep the INPUT record of a tainted env (dyn_in is intact until our own store)
    int idx = 0;
    if (lg == 0) {
      idx = atomicAdd(sa.counters + 1, 1);
      sa.u_list[idx] = (int)(env0 + el);
    }
    idx = __shfl_sync(__activemask(), idx, (threadIdx.x & 31) & ~(G - 1));
    const float* src = sa.dyn_in + (env0 + el) * kp.dyn_stride;
    float* dst = sa.u_buf + (size_t)idx * kp.dyn_stride;
    for (int k = lg; k < kp.dyn_stride; k += G) dst[k] = src[k];
  }
  fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    bulk_s2g(sa.dyn_out + env0 * kp.dyn_stride, t.dynS, (uint32_t)n_valid * kp.dyn_stride * 4u);
    if (sa.act_out && kp.act_stride) bulk_s2g(sa.act_out + env0 * kp.act_stride, t.actS, (uint32_t)n_valid * kp.act_stride * 4u);
    if (sa.do_lidar && sa.lidar_out && kp.lidar_stride)
      bulk_s2g(sa.lidar_out + env0 * kp.lidar_stride, t.lidS, (uint32_t)n_valid * kp.lidar_stride * 4u);
    bulk_commit();
  }
  if (!spec) {
    if (tid == 0) bulk_wait_all();
    return;
  }
  // ---- batch-wide bits: CTA-level OR is in smem, one global atomic per word that still adds a bit
  for (int w = tid; w < nw; w += nthr) {
    uint32_t v = t.bitsE[w];
    if (v & ~__ldcg(sa.any_exact + w)) atomicOr(sa.any_exact + w, v);
    v = t.bitsS[w];
    if (v & ~__ldcg(sa.any_spec + w)) atomicOr(sa.any_spec + w, v);
    v = t.bitsT[w];
    if (v & ~__ldcg(sa.thin + w)) atomicOr(sa.thin + w, v);
  }
  if (tid == 0) bulk_wait_all();
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(sa.counters, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  // ---- last CTA: verify, repair if needed, publish the verdicts, reset the scratch for the next launch
  __threadfence();
  int bad = 0;
  for (int w = tid; w < nw; w += nthr) bad |= (__ldcg(sa.thin + w) & ~__ldcg(sa.any_exact + w)) != 0u;
  bad = __syncthreads_or(bad);
  const int nu = *reinterpret_cast<volatile int*>(sa.counters + 1);
  if (bad) {
    repair_tainted(kp, aio, sa, t, t.bitsE, nu);
    for (int w = tid; w < nw; w += nthr) sa.flags_out[w] = t.bitsE[w];
  } else {
    for (int w = tid; w < nw; w += nthr) sa.flags_out[w] = __ldcg(sa.any_exact + w) | __ldcg(sa.any_spec + w);
  }
  for (int w = tid; w < nw; w += nthr) { sa.any_exact[w] = 0u; sa.any_spec[w] = 0u; sa.thin[w] = 0u; }
  if (tid == 0) {
    sa.counters[0] = 0; sa.counters[1] = 0;
    if (bad) sa.counters[2] += 1;
  }
}

// World.collides' batch-wide distance test (core.py:2796-2800) for substep s on the records in HBM:
// flags[s][p] = any over envs of ||p_a - p_b|| <= Rc_a + Rc_b.  One thread per env, ballot per pair.
__global__ void __launch_bounds__(256) flags_kernel(const __grid_constant__ KParams kp, const float* __restrict__ dyn,
                                                    const float* __restrict__ stat, long long n_envs, int s,
                                                    uint32_t* __restrict__ flags) {
  const long long env = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = env < n_envs;
  EnvPtrs ep;
  ep.dyn = const_cast<float*>(dyn) + (ok ? env : 0) * kp.dyn_stride;
  ep.stat = stat + (ok ? env : 0) * kp.stat_stride;
  ep.work = nullptr;
  for (int p = 0; p < kp.P; ++p) {
    const DBroad* bq = kp.broad + p;
    V2 pa = pos_enc(ep, bq->pa_off), pb = pos_enc(ep, bq->pb_off);
    const float dx = pa.x - pb.x, dy = pa.y - pb.y;
    bool bp = ok && (fmaf(dy, dy, dx * dx) <= bq->t2_bp);
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
  EnvPtrs ep;
  ep.dyn = const_cast<float*>(dyn) + env * kp.dyn_stride;
  ep.stat = stat + env * kp.stat_stride;
  ep.work = nullptr;
  V2 o = posof(kp.ents + entity, ep);
  float ds, dc;
  sincosf(angles[env * angles_stride + r], &ds, &dc);
  float best = max_range;
  for (int t = 0; t < rt.n; ++t) {
    const DEnt* en = kp.ents + rt.e[t];
    V2 p; float rot;
    pose(en, ep, p, rot);
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
