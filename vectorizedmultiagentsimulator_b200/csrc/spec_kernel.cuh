// spec_kernel.cuh - the WORLD-COMPILED step kernel (NVRTC, jit.h).  Included after kernels.cuh, the generated prelude
// (vb::gen constants), spec_helpers.cuh and the generated world functions:
//   world_load(st, aio, sa, env)   coalesced row loads of the env's state and agent forces into registers
//   world_begin(st)                cos/sin cache + begin_r of every dynamic entity
//   world_items(st)                joint_r of every joint, then pair_r of every candidate pair, in the reference's order
//   world_integrate(st, s, more)   integrate_r (+ begin_r of the next substep) of every dynamic entity
//   world_store(st, sa, env)       dyn_out / act_out rows
//   world_stage_lidar(st, env, work) px py rot (and cos/sin of non-sphere targets) of the entities the epilogue reads
//   world_lidar_init / _pass1, the p_* / k_* tables of the (sensor, target) pairs and sensors: see lidar_world below
//
// One thread per environment, state in registers for all substeps, no shared memory and no barrier on the physics
// path.  The speculation protocol is the generic kernel's (kernels.cuh header): verdicts of untainted envs go to
// any_exact, an env is tainted after its first thin event and from then on files its verdicts and thin events in its
// record; the CTA that finishes last verifies thin & ~any_exact == 0 and otherwise raises counters[3] for
// kernels.cuh repair_kernel (generic validation + repair, compiled into the library), which the host launches right
// behind every launch of a world that can have thin events.  THIN_POSSIBLE = false (sphere-sphere worlds) compiles all
// of that away.
// LIDAR: lidar_world below (sensors, targets and rays as constants) on a compact tile, or the generic three-pass
// epilogue (kernels.cuh lidar_epilogue) for the worlds jit.h does not cover; such a CTA keeps 128 threads so that the
// passes run on full warps even when only TE = 32 of them carry an environment.
#pragma once

namespace vb {

// The LIDAR epilogue with the world's sensors as constants (the culls of kernels.cuh lidar_epilogue, the same exact
// arithmetic for the rays that can hit):
//   1  thread per env, straight from the registers: per sensor the mask of targets within reach (world_lidar_pass1,
//      generated: unrolled over sensors and targets, thresholds as literals); every (sensor, target) within reach is
//      one entry of the pair queue (one shared atomic per env);
//   2  thread per queued pair: the target is turned into the sensor's frame once, then every ray of the sensor is
//      tested against it with the plan's cos / sin table of the ray angles (6 flops per ray, no memory access when all
//      sensors share their rays); a ray that may hit is queued once (its output cell is the flag);
//   3  thread per queued ray: the reference's exact arithmetic (step_kernel.cuh cast_one_ray) over the sensor's mask.
// A culled ray keeps max_range, which is what the exact arithmetic gives a ray that cannot reach: the culls are
// conservative (plan.cpp tcull: reach, lateral and behind-the-origin bounds with a slack for the approximate
// direction), so the result does not depend on them.
// Every thread of the CTA calls it (CTA barriers inside); s_qn[0..1] must be zero on entry (set at kernel start).
// It also counts the CTA as finished (sa.counters[0]) and leaves in *s_last whether it was the step's last one.
namespace gen {
constexpr int kRaySentinel = 0x7fc0dead;  // a NaN no range is equal to: "already queued"
VB_DI void lidar_world(const KParams& kp, const StepArgs& sa, const Tile& t, const St& st, bool phys, bool ok, int* s_qn, int* s_last,
                       long long env0) {
  const int tid = threadIdx.x;
  unsigned short* pq = reinterpret_cast<unsigned short*>(t.queue);  // [NPT * TE]
  unsigned short* rq = pq + NPT * TE;                                // [RAYS_TOTAL * TE]
  if (!DIRECT_LID) world_lidar_init(t.lidS, tid);  // every ray starts at max_range
  if (phys) {
    if (DIRECT_LID)  // "queued" bits of the env's sensors (one word each)
      for (int k = 0; k < (L_Q - L_LID); ++k) t.lidS[k * TE + tid] = 0.f;
    world_stage_lidar(st, t.envS + tid, t.workS + tid);
    world_lidar_pass1(st, ok, tid, t.workS + tid, pq, s_qn);
  }
  __syncthreads();
  // the physics of the whole CTA is behind this barrier: count the CTA now (see the kernel's tail), by a thread of the
  // last warp (a helper warp when NT > TE), so that the fence and the atomic's round trip overlap the passes below
#ifndef VB_AB_LATE_COUNT
  if (tid == NT - 1) {
    __threadfence();
    *s_last = (atomicAdd(sa.counters, 1) == sa.total_ctas - 1);
  }
#endif
  const int np = s_qn[0];
  for (int task = tid; task < np; task += NT) {
    const int e = pq[task], col = e & (TE - 1), p = e >> CB;
    const float* ec = t.envS + col;
    const float* wc = t.workS + col;
    const int es = p_es(p), ts = p_ts(p), cr = p_cr(p);
    const float ox = ec[es * TE], oy = ec[(es + 1) * TE];
    const float srot = ec[(es + 2) * TE];
    const float slack = p_slk(p) * (fabsf(srot) + 8.f);
#ifndef VB_AB_SINCOS_P2
    const float c0 = wc[cr * TE], s0 = wc[(cr + 1) * TE];  // pass 1 (measured: 146 us against 154 us at 1 M envs when it is done here)
#else
    float c0, s0;
    sincosf(srot, &s0, &c0);
#endif
    const float rx = ec[ts * TE] - ox, ry = ec[(ts + 1) * TE] - oy;
    const float u = rx * c0 + ry * s0, v = ry * c0 - rx * s0;  // the target in the sensor's frame
    const float lat_max = p_lat(p) + slack, along_min = -(p_along(p) + slack);
    const int out = p_out(p);
    const int tag = col | (p_k(p) << CB);
    int* cell = reinterpret_cast<int*>(t.lidS) + out * TE + col;
    auto may_hit = [&](float2 cs) {
      const float lat = fabsf(u * cs.y - v * cs.x), along = u * cs.x + v * cs.y;
      return !(lat > lat_max) && !(along < along_min);
    };
    uint32_t* queued = reinterpret_cast<uint32_t*>(t.lidS) + p_k(p) * TE + col;  // DIRECT_LID
    auto push = [&](int r) {  // once per ray: its output cell (its bit of the sensor's word) is the flag
      if (DIRECT_LID) {
        if (!(atomicOr(queued, 1u << r) >> r & 1u)) rq[atomicAdd(s_qn + 1, 1)] = (unsigned short)(tag | (r << (CB + KB)));
      } else {
        if (atomicExch(cell + r * TE, kRaySentinel) != kRaySentinel) rq[atomicAdd(s_qn + 1, 1)] = (unsigned short)(tag | (r << (CB + KB)));
      }
    };
#ifdef VB_AB_DIRECT_PUSH
    if constexpr (SHARED_RAYS > 0) {
#pragma unroll
      for (int r = 0; r < SHARED_RAYS; ++r)
        if (may_hit(T_raycs[r])) push(r);
    } else
#endif
    if constexpr (SHARED_RAYS > 0 && SHARED_RAYS <= 32) {  // branch-free sweep into a mask, then only the rays that passed
      uint32_t cm = 0u;
#pragma unroll
      for (int r = 0; r < SHARED_RAYS; ++r) cm |= (may_hit(T_raycs[r]) ? 1u : 0u) << r;
      while (cm) { push(__ffs(cm) - 1); cm &= cm - 1; }
    } else if constexpr (SHARED_RAYS > 32) {
      unsigned long long cm = 0ull;
#pragma unroll
      for (int r = 0; r < SHARED_RAYS; ++r) cm |= (may_hit(T_raycs[r]) ? 1ull : 0ull) << r;
      while (cm) { push(__ffsll((long long)cm) - 1); cm &= cm - 1; }
    } else {
      const int nr = p_nr(p);
      for (int r = 0; r < nr; ++r)
        if (may_hit(__ldg(kp.ray_cs + out + r))) push(r);
    }
  }
  __syncthreads();
  const int nq = s_qn[1];
  for (int qi = tid; qi < nq; qi += NT) {
    const int item = rq[qi], col = item & (TE - 1), k = (item >> CB) & ((1 << KB) - 1), r = item >> (CB + KB);
    const float* ec = t.envS + col;
    const float* wc = t.workS + col;
    const int es = k_es(k), out = k_out(k), tg0 = k_tg0(k);
    const V2 o = mk(ec[es * TE], ec[(es + 1) * TE]);
    const float ang = __ldg(kp.ray_angles + out + r) + ec[(es + 2) * TE];  // sensors.py:117
    const float max_range = k_range(k);
    float best = max_range, ds, dc;
    sincosf(ang, &ds, &dc);
    uint32_t m = __float_as_uint(wc[k_mrow(k) * TE]);
    while (m) {  // step_kernel.cuh cast_one_ray with the targets' dimensions as constants
      const int p = tg0 + __ffs(m) - 1;
      m &= m - 1;
      const int ts = p_ts(p), shape = p_shape(p);
      const V2 tp = mk(ec[ts * TE], ec[(ts + 1) * TE]);
      float d;
      if (shape == VMASB200_SPHERE) {
        d = ray_sphere(tp, p_radius(p), o, dc, ds, max_range);
      } else {
        const int tr = p_trig(p);
        const float tc = wc[tr * TE], tsn = wc[(tr + 1) * TE];
        d = shape == VMASB200_BOX ? ray_box(tp, tc, tsn, p_length(p), p_width(p), o, dc, ds, max_range)
                                  : ray_line(tp, tc, tsn, p_length(p), o, dc, ds, max_range);
      }
      best = pmin(best, d);
    }
    if (DIRECT_LID) sa.lidar_out[(size_t)(out + r) * sa.ld + env0 + col] = best;  // over the fill of the kernel's start
    else t.lidS[(out + r) * TE + col] = best;
  }
  __syncthreads();
}
}  // namespace gen

extern "C" __global__ void __launch_bounds__(gen::NT, gen::MINB)
vmasb200_step_world(const __grid_constant__ KParams kp, const __grid_constant__ AgentIO aio, const __grid_constant__ StepArgs sa) {
  using namespace gen;
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_last;
  __shared__ int s_qn[2];  // LIDAR: queued pairs, queued rays
  const int tid = threadIdx.x;
  // programmatic dependent launch (vmasb200.cu): nothing of the previous grid on the stream may be read or written before
  // this returns (at once when the launch carries no such dependency); the next grid may be set up from here on
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  bool counted = false;  // the CTA is in sa.counters[0] and s_last is valid
  if (LIDAR && LIDAR_COMPILED) {  // before the physics, so that the epilogue needs no barrier of its own for it
    if (tid < 2) s_qn[tid] = 0;
    __syncthreads();
  }
  const long long env0 = sa.env_begin + (long long)blockIdx.x * TE;
  const int n_valid = (int)min((long long)TE, sa.env_end - env0);
  const bool phys = tid < TE;
  const bool ok = phys && tid < n_valid;
  const long long env = env0 + (ok ? tid : 0);
  // LIDAR rows of the tile in global memory; the batch's last tile also carries the padding columns up to ld
  const int n_cols = (int)min((long long)TE, (sa.env_end == sa.n_envs ? sa.ld : sa.env_end) - env0);
  if (LIDAR && LIDAR_COMPILED && DIRECT_LID && sa.do_lidar && sa.lidar_out) {
    // every ray starts at max_range: filled here by the helper warps (all threads when there are none), so that the stores
    // drain while the physics runs; pass 3 of the epilogue stores the hits over them (two CTA barriers later)
    const bool vec = ((n_cols | (int)(env0 & 3)) & 3) == 0;
    if (NT > TE) { if (!phys) world_lidar_fill(sa.lidar_out + env0, (size_t)sa.ld, n_cols, tid - TE, NT - TE, vec); }
    else world_lidar_fill(sa.lidar_out + env0, (size_t)sa.ld, n_cols, tid, NT, vec);
  }
  St st;
  st.ss = BIG ? smem + tid : nullptr;  // static tile of a pair-heavy world (non-LIDAR): TILE_ROWS rows of NT columns
  st.coop = BIG && COOP ? smem + TILE_ROWS * NT + (tid >> 5) * kCoopWords : nullptr;  // ... then every warp's scratch (narrow_coop)
  park(st.vx, st.ss, PARK_ROW); park(st.vy, st.ss, PARK_ROW + ND); park(st.rot, st.ss, PARK_ROW + 2 * ND); park(st.w, st.ss, PARK_ROW + 3 * ND);
  park(st.fx, st.ss, PARK_ROW + 4 * ND); park(st.fy, st.ss, PARK_ROW + 4 * ND + NA); park(st.tq, st.ss, PARK_ROW + 4 * ND + 2 * NA);
  park(st.Fx, st.ss, ACC_ROW); park(st.Fy, st.ss, ACC_ROW + 1); park(st.T, st.ss, ACC_ROW + 2);
  int taint = 0;  // 0, or 1 + the env's index in sa.u_list (kernels.cuh run_substeps)
  if (phys) {
    world_load(st, aio, sa, env);
    world_begin(st);
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int w = 0; w < kPW; ++w) { st.verd[w] = 0u; st.thin[w] = 0u; st.sv[w] = 0u; }
      world_items(st, kp);
      const bool tainted = THIN_POSSIBLE && taint != 0;
      {
        // the warp's bits of this substep into the batch-wide words: all reductions first, then lane 0 reads every word
        // it may have to add to (independent L2 loads in flight together; one after the other they were a chain of L2
        // round trips per substep, 5.6 % of the samples of the football kernel), then the atomics that add a bit
        uint32_t e[kPW], sp[kPW], th[kPW];
#pragma unroll
        for (int w = 0; w < PW; ++w) {
          e[w] = __reduce_or_sync(0xffffffffu, (ok && !tainted) ? st.verd[w] : 0u);
          sp[w] = THIN_POSSIBLE ? __reduce_or_sync(0xffffffffu, (ok && tainted) ? st.verd[w] : 0u) : 0u;
          th[w] = THIN_POSSIBLE ? __reduce_or_sync(0xffffffffu, ok ? st.thin[w] : 0u) : 0u;
        }
        if ((tid & 31) == 0) {
          uint32_t ce[kPW], cs[kPW], ct[kPW];
#pragma unroll
          for (int w = 0; w < PW; ++w) {
            ce[w] = e[w] ? __ldcg(sa.any_exact + s * PW + w) : 0xffffffffu;
            cs[w] = sp[w] ? __ldcg(sa.any_spec + s * PW + w) : 0xffffffffu;
            ct[w] = th[w] ? __ldcg(sa.thin + s * PW + w) : 0xffffffffu;
          }
#pragma unroll
          for (int w = 0; w < PW; ++w) {
            if (e[w] & ~ce[w]) atomicOr(sa.any_exact + s * PW + w, e[w]);
            if (sp[w] & ~cs[w]) atomicOr(sa.any_spec + s * PW + w, sp[w]);
            if (th[w] & ~ct[w]) atomicOr(sa.thin + s * PW + w, th[w]);
          }
        }
        if (THIN_POSSIBLE && ok && tainted && taint <= sa.rec_cap)
#pragma unroll
          for (int w = 0; w < PW; ++w) sa.u_verd[((size_t)(taint - 1) * S + s) * PW + w] = st.verd[w];
      }
      world_integrate(st, s, s + 1 < S);
      if (THIN_POSSIBLE && ok) {  // thin events of this substep: tainted from the NEXT substep on, with a record
        uint32_t any = 0u;
#pragma unroll
        for (int w = 0; w < PW; ++w) any |= st.thin[w];
        if (any) {
          const size_t rw = (size_t)S * PW;
          if (taint == 0) {
            taint = atomicAdd(sa.counters + 1, 1) + 1;
            sa.u_list[taint - 1] = (int)env;
            if (taint <= sa.rec_cap)
              for (size_t i = 0; i < rw; ++i) { sa.u_verd[(taint - 1) * rw + i] = 0u; sa.u_thin[(taint - 1) * rw + i] = 0u; }
          }
          if (taint <= sa.rec_cap)
#pragma unroll
            for (int w = 0; w < PW; ++w)
              if (st.thin[w]) sa.u_thin[(taint - 1) * rw + (size_t)s * PW + w] = st.thin[w];
        }
      }
    }
  }
  // Everything the step's last CTA reads (verdict words, taint records) is written by now; the state and LIDAR rows
  // are not part of that hand-shake (a repair runs in a kernel of its own).  So the CTA is counted HERE by one thread
  // - barrier, fence, atomic; the result stays in a register - and the round trip overlaps the stores below.  (Worlds
  // with the compiled LIDAR epilogue count behind its first barrier instead, which saves this one.)
  int count_old = -1;
#ifndef VB_AB_LATE_COUNT2
  if (!(LIDAR && LIDAR_COMPILED && sa.do_lidar && sa.lidar_out)) {
    __syncthreads();
    if (tid == NT - 1) {
      __threadfence();
      count_old = atomicAdd(sa.counters, 1);
    }
  }
#endif
  if (phys) {
    if (ok) {
      if (THIN_POSSIBLE && taint)  // keep the INPUT state of a tainted env (dyn_in is intact until our own store)
        for (int r = 0; r < DYN_ROWS; ++r) sa.u_buf[(size_t)(taint - 1) * DYN_ROWS + r] = sa.dyn_in[(size_t)r * sa.ld + env];
      world_store(st, sa, env);
    }
  }
  if (LIDAR && sa.do_lidar && sa.lidar_out) {
    Tile t;
    if (LIDAR_COMPILED) {  // the generator's own carve-up (jit.h): no mbarrier, 16-bit queues
      t.envS = smem + L_ENV * TE; t.workS = smem + L_WORK * TE; t.lidS = smem + L_LID * TE;
      t.queue = reinterpret_cast<int*>(smem + L_Q * TE); t.rayq = nullptr; t.qcount = nullptr;
    } else {
      const SmemMap m = smem_map(kp);
      t.envS = smem + m.env; t.workS = smem + m.work; t.lidS = smem + m.lid;
      t.queue = reinterpret_cast<int*>(smem + m.queue); t.rayq = reinterpret_cast<int*>(smem + m.rayq);
      t.qcount = reinterpret_cast<int*>(smem + m.bar) + 2;
    }
    // LIDAR rows: 16-byte stores straight from the tile (rows are 16-byte aligned: ld % 4 == 0, env0 % TE == 0; a TMA
    // bulk store per row costs the issuing warp ~14 instructions per row, one lane at a time, and a wait before the CTA
    // may leave).
    auto store_rows = [&]() {
      constexpr int LPR = TE / 4, RSTEP = NT / LPR;  // lanes per row, rows per sweep of the CTA
      const int rows = LIDAR_COMPILED ? RAYS_TOTAL : kp.lidar_rows;
      if (((n_cols | (int)(env0 & 3)) & 3) == 0) {
        const int c = (tid % LPR) * 4;
        int r = tid / LPR;
        if (c < n_cols) {
          float* g = sa.lidar_out + (size_t)r * sa.ld + env0 + c;
          const float* l = t.lidS + r * TE + c;
          for (; r < rows; r += RSTEP, g += (size_t)RSTEP * sa.ld, l += RSTEP * TE)
            *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(l);
        }
      } else {
        for (int i = tid; i < rows * TE; i += NT) {
          const int r = i / TE, c = i - r * TE;
          if (c < n_cols) sa.lidar_out[(size_t)r * sa.ld + env0 + c] = t.lidS[r * TE + c];
        }
      }
    };
    if (LIDAR_COMPILED) {
      lidar_world(kp, sa, t, st, phys, ok, s_qn, &s_last, env0);
      if (!DIRECT_LID) {
#ifdef VB_AB_TMA_STORE
      fence_proxy_async();
      __syncthreads();
      if (tid < 32) {
        for (int r = tid; r < kp.lidar_rows; r += 32) bulk_s2g(sa.lidar_out + (size_t)r * sa.ld + env0, t.lidS + r * TE, (uint32_t)n_cols * 4u);
        bulk_commit();
        bulk_wait_read();
      }
#else
      store_rows();
#endif
      }
#ifndef VB_AB_LATE_COUNT
      counted = true;
#endif
    } else {
      if (phys) world_stage_lidar(st, t.envS + tid, t.workS + tid);
      __syncthreads();
      lidar_epilogue<true, TE>(kp, t, n_valid);
      __syncthreads();
      store_rows();
    }
  }
  // ---- last CTA of the step: verify (repair if needed), publish the verdicts, reset the scratch for the next launch
  // release by one thread: the barrier orders the CTA's writes before that thread's fence (cumulativity), which orders
  // them before its atomic - no fence per thread (ncu: stall_membar was the top stall of the small worlds).  What the
  // last CTA reads (verdict words, records) is written by the physics, so lidar_world counts the CTA right behind the
  // physics and the atomic's latency hides under the LIDAR passes; its closing barrier publishes s_last.
  if (!counted) {
#ifndef VB_AB_LATE_COUNT2
    if (tid == NT - 1) s_last = (count_old == sa.total_ctas - 1);
#else
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      s_last = (atomicAdd(sa.counters, 1) == sa.total_ctas - 1);
    }
#endif
    __syncthreads();
  }
  if (!s_last) return;
  __threadfence();
  constexpr int nw = S * PW;
  if (!THIN_POSSIBLE) {
    for (int w = tid; w < nw; w += blockDim.x) {
      sa.flags_out[w] = __ldcg(sa.any_exact + w);
      sa.any_exact[w] = 0u;
    }
    if (tid == 0) sa.counters[0] = 0;
    return;
  }
  int bad = 0;
  for (int w = tid; w < nw; w += blockDim.x) bad |= (__ldcg(sa.thin + w) & ~__ldcg(sa.any_exact + w)) != 0u;
  bad = __syncthreads_or(bad);
  if (bad) {
    // hand over to kernels.cuh repair_kernel, which the host launches right behind this kernel: it validates the
    // speculation, re-simulates the envs that left the reference's trajectory, publishes and resets the scratch
    if (tid == 0) {
      sa.counters[0] = 0;
      __threadfence();
      atomicExch(sa.counters + 3, 1);
    }
    return;
  }
  for (int w = tid; w < nw; w += blockDim.x) {
    sa.flags_out[w] = __ldcg(sa.any_exact + w) | __ldcg(sa.any_spec + w);
    sa.any_exact[w] = 0u; sa.any_spec[w] = 0u; sa.thin[w] = 0u;
  }
  if (tid == 0) { sa.counters[0] = 0; sa.counters[1] = 0; }
}

}  // namespace vb
