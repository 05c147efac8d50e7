// spec_kernel.cuh - the WORLD-COMPILED step kernel (NVRTC, jit.h).  Included after kernels.cuh, the generated prelude
// (vb::gen constants), spec_helpers.cuh and the generated world functions:
//   world_load(st, aio, sa, env)   coalesced row loads of the env's state and agent forces into registers
//   world_begin(st)                begin_r of every dynamic entity
//   world_pairs(st)                pair_ss_r of every candidate pair, in the reference's pair order
//   world_integrate(st, s, more)   integrate_r (+ begin_r of the next substep) of every dynamic entity
//   world_store(st, sa, env)       dyn_out / act_out rows
//   world_stage_lidar(st, envS)    px py rot of the sensing entities and their targets into the epilogue's tile
//
// One thread per environment, state in registers for all substeps, no shared memory and no barrier on the physics
// path; worlds this kernel is generated for (sphere-sphere pairs only, no joints) cannot produce a "thin" contact, so
// every env equals the reference without the speculation records of the generic kernel: only the verdict bits are
// OR-ed into any_exact and published by the CTA that finishes last (same protocol and scratch as kernels.cuh, so
// chunked steps and vmasb200_step_io.pair_active_out work unchanged).
// LIDAR: the generic three-pass epilogue (kernels.cuh lidar_epilogue) on a compact tile; the CTA keeps 128 threads so
// that its passes run on full warps even when only TE = 32 of them carry an environment.
#pragma once

namespace vb {

extern "C" __global__ void __launch_bounds__(gen::NT)
vmasb200_step_world(const __grid_constant__ KParams kp, const __grid_constant__ AgentIO aio, const __grid_constant__ StepArgs sa) {
  using namespace gen;
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_last;
  const int tid = threadIdx.x;
  const long long env0 = sa.env_begin + (long long)blockIdx.x * TE;
  const int n_valid = (int)min((long long)TE, sa.env_end - env0);
  const bool phys = tid < TE;
  const bool ok = phys && tid < n_valid;
  const long long env = env0 + (ok ? tid : 0);
  St st;
  if (phys) {
    world_load(st, aio, sa, env);
    world_begin(st);
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int w = 0; w < kPW; ++w) st.verd[w] = 0u;
      world_pairs(st);
#pragma unroll
      for (int w = 0; w < PW; ++w) {
        const uint32_t e = __reduce_or_sync(0xffffffffu, ok ? st.verd[w] : 0u);
        if ((tid & 31) == 0 && e) or_global(sa.any_exact + s * PW + w, e);
      }
      world_integrate(st, s, s + 1 < S);
    }
    if (ok) world_store(st, sa, env);
  }
  if (LIDAR && sa.do_lidar && sa.lidar_out) {
    const SmemMap m = smem_map(kp);
    Tile t;
    t.envS = smem + m.env; t.workS = smem + m.work; t.lidS = smem + m.lid;
    t.queue = reinterpret_cast<int*>(smem + m.queue); t.rayq = reinterpret_cast<int*>(smem + m.rayq);
    t.qcount = reinterpret_cast<int*>(smem + m.bar) + 2;
    if (phys) world_stage_lidar(st, t.envS + tid);
    __syncthreads();
    lidar_epilogue<true, TE>(kp, t, n_valid);
    // one TMA bulk store per LIDAR row; the batch's last tile also carries the padding columns up to ld
    const int n_cols = (int)min((long long)TE, (sa.env_end == sa.n_envs ? sa.ld : sa.env_end) - env0);
    const uint32_t row_bytes = (uint32_t)n_cols * 4u;
    fence_proxy_async();
    __syncthreads();
    if (tid < 32) {
      for (int r = tid; r < kp.lidar_rows; r += 32) bulk_s2g(sa.lidar_out + (size_t)r * sa.ld + env0, t.lidS + r * TE, row_bytes);
      bulk_commit();
      bulk_wait_all();
    }
  }
  // ---- last CTA of the step: publish the verdicts, reset the scratch for the next launch
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(sa.counters, 1) == sa.total_ctas - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int w = tid; w < S * PW; w += blockDim.x) {
    sa.flags_out[w] = __ldcg(sa.any_exact + w);
    sa.any_exact[w] = 0u;
  }
  if (tid == 0) sa.counters[0] = 0;
}

}  // namespace vb
