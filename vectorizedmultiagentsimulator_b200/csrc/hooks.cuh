// hooks.cuh - scenario hooks after World.step: the observation / reward / done / info tensors
// Environment._get_from_scenario (environment.py:253-308) collects, as one or two launches per step.
//
//   gather_obs_kernel       every agent's observation = torch.cat([...], dim=-1) of state fields, differences of state
//                           fields and sensor readings: SoA rows (env fastest) are read coalesced, transposed through
//                           shared memory and written as [n_envs][width] rows (column fastest), all agents in one launch
//   nav_flags_kernel +      navigation.py:199-232, 262-280: the batch-wide World.collides verdict per agent pair
//   nav_reward_kernel       (core.py:2796-2800), then per env distances to goal, shaping rewards, collision penalties,
//                           final reward, done
//   transport_reward_kernel transport.py:130-164, 186-193
// Bytes per env: inputs + outputs once (navigation 4 agents: 48 B in, ~100 B out; obs 4 x 18 floats): HBM / launch bound.
#pragma once
#include "queries.cuh"

namespace vb {

struct DVec {
  const void* p;
  long long se, sk;
};
VB_DI float vget(const DVec& v, long long i, int k) { return static_cast<const float*>(v.p)[i * v.se + k * v.sk]; }

// ------------------------------------------------------------------------------------------------ observations
struct ObsSeg {
  DVec a, b;
  int op;
  float c;
  int width, out, col0;
};
struct ObsTable {
  int n_segs, n_outs;
  float* outs[VMASB200_OBS_MAX_OUTS];
  int width[VMASB200_OBS_MAX_OUTS];
  ObsSeg seg[VMASB200_OBS_MAX_SEGS];
};
constexpr int kObsTile = 32;  // envs per CTA
// grid (tiles of 32 envs, outs); 256 threads.  smem: width x 33 floats.
__global__ void __launch_bounds__(256) gather_obs_kernel(const __grid_constant__ ObsTable tb, long long n) {
  extern __shared__ float tile[];
  const int o = blockIdx.y;
  const int W = tb.width[o];
  const long long env0 = (long long)blockIdx.x * kObsTile;
  const int n_valid = (int)min((long long)kObsTile, n - env0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long i = env0 + (lane < n_valid ? lane : 0);
  for (int s = 0; s < tb.n_segs; ++s) {
    const ObsSeg& sg = tb.seg[s];
    if (sg.out != o) continue;
    for (int k = warp; k < sg.width; k += 8) {
      float v;
      switch (sg.op) {
        case VMASB200_OBS_COPY: v = vget(sg.a, i, k); break;
        case VMASB200_OBS_SUB: v = vget(sg.a, i, k) - vget(sg.b, i, k); break;
        case VMASB200_OBS_CSUB: v = sg.c - vget(sg.a, i, k); break;
        default: v = static_cast<const uint8_t*>(sg.a.p)[i * sg.a.se + k * sg.a.sk] ? 1.f : 0.f; break;
      }
      tile[(sg.col0 + k) * (kObsTile + 1) + lane] = v;
    }
  }
  __syncthreads();
  float* __restrict__ out = tb.outs[o] + env0 * W;
  const int total = n_valid * W;
  for (int idx = threadIdx.x; idx < total; idx += 256) {
    const int e = idx / W, c = idx - e * W;
    out[idx] = tile[c * (kObsTile + 1) + e];
  }
}

// ------------------------------------------------------------------------------------------------ navigation
struct NavAgent {
  DVec pos, goal;
  float radius, goal_radius;
  const float* shaping_in;
  float* shaping_out;
  float* dist;
  uint8_t* on_goal;
  float* pos_rew;
  float* coll_rew;
  float* reward;
};
struct NavParams {
  int n_agents, shared_rew;
  float factor, final_reward, penalty, min_coll_dist;
  NavAgent ag[VMASB200_NAV_MAX_AGENTS];
  uint8_t coll[VMASB200_NAV_MAX_AGENTS][VMASB200_NAV_MAX_AGENTS];
  float rc_sum[VMASB200_NAV_MAX_AGENTS][VMASB200_NAV_MAX_AGENTS];  // (float)(Rc_i + Rc_j), the sum taken in double
  float* pos_rew;
  float* final_rew;
  uint8_t* all_goal;
  uint8_t* done;
  uint32_t* flags;
};
VB_DI int nav_pair(int i, int j) { return i * (i - 1) / 2 + j; }  // j < i

// World.collides' batch-wide verdict (core.py:2796-2800) of every agent pair the reward loop asks about
// (navigation.py:217-221): one thread per env, verdicts OR-reduced over the warp, one atomicOr per warp and word.
template <int MAXA>
__global__ void __launch_bounds__(256) nav_flags_kernel(const __grid_constant__ NavParams np, long long n) {
  const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = gi < n;
  const long long i = ok ? gi : 0;
  V2 p[MAXA];
#pragma unroll
  for (int a = 0; a < MAXA; ++a)
    if (a < np.n_agents) p[a] = mk(vget(np.ag[a].pos, i, 0), vget(np.ag[a].pos, i, 1));
  uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
  for (int a = 1; a < MAXA; ++a)
#pragma unroll
    for (int b = 0; b < a; ++b)
      if (a < np.n_agents && np.coll[a][b]) {
        const int q = nav_pair(a, b);
        if (ok && norm(p[a] - p[b]) <= np.rc_sum[a][b]) w[q >> 5] |= 1u << (q & 31);
      }
  const int n_words = (nav_pair(np.n_agents - 1, np.n_agents - 2) >> 5) + 1;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (k < n_words) {
      const uint32_t any = __reduce_or_sync(0xffffffffu, w[k]);
      if (any && (threadIdx.x & 31) == 0 && (any & ~__ldcg(np.flags + k))) atomicOr(np.flags + k, any);
    }
}

template <int MAXA>
__global__ void __launch_bounds__(256) nav_reward_kernel(const __grid_constant__ NavParams np, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  V2 p[MAXA];
  float coll[MAXA], prew[MAXA];
  float sum = 0.f;          // self.pos_rew[:] = 0; += agent_reward(a) in agent order (navigation.py:203-208)
  bool all_goal = true, done = true;
#pragma unroll
  for (int a = 0; a < MAXA; ++a)
    if (a < np.n_agents) {
      const NavAgent& ag = np.ag[a];
      p[a] = mk(vget(ag.pos, i, 0), vget(ag.pos, i, 1));
      const V2 g = mk(vget(ag.goal, i, 0), vget(ag.goal, i, 1));
      const float d = norm(p[a] - g);                 // agent.distance_to_goal (:234-237)
      const bool og = d < ag.goal_radius;             // agent.on_goal (:238)
      const float sh = d * np.factor;                 // pos_shaping (:240)
      prew[a] = ag.shaping_in[i] - sh;                // agent.pos_rew (:241)
      ag.shaping_out[i] = sh;
      ag.dist[i] = d;
      ag.on_goal[i] = og ? 1 : 0;
      ag.pos_rew[i] = prew[a];
      sum = sum + prew[a];
      all_goal = all_goal && og;
      done = done && (d < ag.radius);                 // done(): < agent.shape.radius (:262-273)
      coll[a] = 0.f;                                  // a.agent_collision_rew[:] = 0 (:209)
    }
  const float fin = all_goal ? np.final_reward : 0.f; // self.final_rew[self.all_goal_reached] = final_reward (:211-215)
#pragma unroll
  for (int a = 1; a < MAXA; ++a)                      // i > j pairs in loop order (:217-228)
#pragma unroll
    for (int b = 0; b < a; ++b)
      if (a < np.n_agents && np.coll[a][b]) {
        const int q = nav_pair(a, b);
        if ((np.flags[q >> 5] >> (q & 31)) & 1u) {    // if self.world.collides(a, b)
          const float d = (norm(p[a] - p[b]) - np.ag[a].radius) - np.ag[b].radius;  // World.get_distance, core.py:1830-1836
          if (d <= np.min_coll_dist) { coll[a] = coll[a] + np.penalty; coll[b] = coll[b] + np.penalty; }
        }
      }
  np.pos_rew[i] = sum;
  np.final_rew[i] = fin;
  np.all_goal[i] = all_goal ? 1 : 0;
  if (np.done) np.done[i] = done ? 1 : 0;
#pragma unroll
  for (int a = 0; a < MAXA; ++a)
    if (a < np.n_agents) {
      np.ag[a].coll_rew[i] = coll[a];
      np.ag[a].reward[i] = ((np.shared_rew ? sum : prew[a]) + fin) + coll[a];  // :230-231
    }
}

// ------------------------------------------------------------------------------------------------ transport
struct TrPackage {
  QPose pkg, goal;
  const float* shaping_in;
  float* shaping_out;
  float* dist;
  uint8_t* on_goal;
  float* color;
};
struct TrParams {
  int n_packages;
  float factor;
  float off[3], on[3];
  TrPackage pk[VMASB200_TRANSPORT_MAX_PACKAGES];
  float* rew;
  uint8_t* done;
};
__global__ void __launch_bounds__(256) transport_reward_kernel(const __grid_constant__ TrParams tp, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float rew = 0.f;  // self.rew = zeros (transport.py:134-138)
  bool done = true;
  for (int k = 0; k < tp.n_packages; ++k) {
    const TrPackage& pk = tp.pk[k];
    const float d = norm(q_pos(pk.pkg, i) - q_pos(pk.goal, i));  // package.dist_to_goal (:141-143)
    const bool og = pair_overlap(pk.pkg, pk.goal, i);             // world.is_overlapping(package, package.goal) (:144)
    const float sh = d * tp.factor;                               // package_shaping (:156)
    if (!og) rew = rew + (pk.shaping_in[i] - sh);                 // self.rew[~on_goal] += global_shaping - shaping (:157-160)
    pk.shaping_out[i] = sh;                                       // package.global_shaping = package_shaping (:161)
    pk.dist[i] = d;
    pk.on_goal[i] = og ? 1 : 0;
    if (pk.color) {                                               // :145-154
      const float* c = og ? tp.on : tp.off;
      pk.color[3 * i] = c[0]; pk.color[3 * i + 1] = c[1]; pk.color[3 * i + 2] = c[2];
    }
    done = done && og;
  }
  tp.rew[i] = rew;
  if (tp.done) tp.done[i] = done ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------ simple_spread
struct SpreadParams {
  int n_agents, n_landmarks;
  DVec apos[VMASB200_SPREAD_MAX], lpos[VMASB200_SPREAD_MAX];
  float radius[VMASB200_SPREAD_MAX];
  uint8_t collide[VMASB200_SPREAD_MAX];
  float* rew;
};
__global__ void __launch_bounds__(256) simple_spread_reward_kernel(const __grid_constant__ SpreadParams sp, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  V2 p[VMASB200_SPREAD_MAX];
  float closest[VMASB200_SPREAD_MAX];
  for (int a = 0; a < sp.n_agents; ++a) p[a] = mk(vget(sp.apos[a], i, 0), vget(sp.apos[a], i, 1));
  for (int l = 0; l < sp.n_landmarks; ++l) {  // torch.min over agents of ||a.pos - landmark.pos|| (simple_spread.py:90-101)
    const V2 q = mk(vget(sp.lpos[l], i, 0), vget(sp.lpos[l], i, 1));
    float m = norm(p[0] - q);
    for (int a = 1; a < sp.n_agents; ++a) m = pmin(m, norm(p[a] - q));
    closest[l] = m;
  }
  float rew = 0.f;
  for (int s = 0; s < sp.n_agents; ++s) {     // for single_agent in self.world.agents (:88)
    for (int l = 0; l < sp.n_landmarks; ++l) rew = rew - closest[l];                       // self.rew -= closest (:102)
    if (sp.collide[s])
      for (int a = 0; a < sp.n_agents; ++a)
        if (a != s && (norm(p[a] - p[s]) - sp.radius[a]) - sp.radius[s] < 0.f) rew = rew - 1.f;  // is_overlapping(a, single_agent) (:104-107)
  }
  sp.rew[i] = rew;
}

}  // namespace vb
