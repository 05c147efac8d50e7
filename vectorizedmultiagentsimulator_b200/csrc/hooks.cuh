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
  DVec a, b, noise;
  int op;
  float c;
  int width, out, col0, flags;
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
      const bool neg = (sg.flags & VMASB200_OBS_NEG0) && k == 0;  // football.py:1376-1397: x mirrored on the operands
      switch (sg.op) {
        case VMASB200_OBS_COPY: v = vget(sg.a, i, k); if (neg) v = -v; break;
        case VMASB200_OBS_SUB: {
          float x = vget(sg.a, i, k), y = vget(sg.b, i, k);
          if (neg) { x = -x; y = -y; }
          v = x - y;
        } break;
        case VMASB200_OBS_CSUB: v = sg.c - vget(sg.a, i, k); break;
        case VMASB200_OBS_COS: v = cosf(vget(sg.a, i, k)); break;
        case VMASB200_OBS_SIN: v = sinf(vget(sg.a, i, k)); break;
        default: v = static_cast<const uint8_t*>(sg.a.p)[i * sg.a.se + k * sg.a.sk] ? 1.f : 0.f; break;
      }
      if (sg.noise.p) v = v + vget(sg.noise, i, k);  // joint_passage.py:531-541 obs + noise
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

// ------------------------------------------------------------------------------------------------ joint_passage
// torch.remainder for fp32 (ATen BinaryOpsKernel remainder_kernel): fmod, then the divisor's sign
VB_DI float remainder_f(float a, float b) {
  float m = fmodf(a, b);
  if (m != 0.f && ((b < 0.f) != (m < 0.f))) m += b;
  return m;
}
// joint_passage.py:24-38 get_line_angle_dist_0_180(angle, goal); g, g_minus_pi: the goal taken modulo pi and that
// minus pi, fp32 (a tensor goal: fp32 arithmetic; a Python float goal: double arithmetic rounded once, by the caller)
VB_DI float line_angle_dist(float rot, float g, float g_minus_pi, float pi_f) {
  const float a = remainder_f(rot, pi_f);
  const float d0 = fabsf(a - g), d1 = fabsf(a - g_minus_pi), d2 = fabsf((a - pi_f) - g);
  return pmin(d0, pmin(d1, d2));
}
struct JpParams {
  int n_agents, n_bodies, n_passages, n_walls, all_passed_rot;
  QPose body[VMASB200_JP_MAX_BODIES];
  DVec action_u[VMASB200_JP_MAX_BODIES];
  float e_div[VMASB200_JP_MAX_BODIES];     // (float)sqrt(dim_p * f_range^2)
  QPose passage[VMASB200_JP_MAX_PASSAGES];
  uint8_t pcollide[VMASB200_JP_MAX_PASSAGES];
  QPose wall[VMASB200_JP_MAX_WALLS];
  QPose joint;
  DVec goal_pos, goal_rot;
  float half_width, pos_factor, rot_factor, coll_rew, energy_coeff, min_coll, pi_f, mid_g, mid_g_minus_pi;
  const float* sh_in[4];
  float* sh_out[4];
  float *pos_rew, *rot_rew, *collision_rew, *energy_exp, *energy_rew, *rew;
  uint8_t *all_passed, *done;
  uint32_t* done_flag;
};
// joint_passage.py:387-490 reward of the first agent; also done()'s two tests (:547-563): rot_ok goes to done[], a
// failed position test raises done_flag (the reference's [n] * [n, 1] product couples every env with all of them)
__global__ void __launch_bounds__(128) jp_reward_kernel(const __grid_constant__ JpParams jp, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool pos_bad = false;
  if (i < n) {
    const V2 jpos = q_pos(jp.joint, i);
    const float jrot = jp.joint.rot[i * jp.joint.rs];
    const bool joint_passed = jpos.y > 0.f;                                              // :399
    bool all_passed = true;                                                              // :400-404
    for (int a = 0; a < jp.n_agents; ++a) all_passed = all_passed && (q_pos(jp.body[a], i).y > jp.half_width);
    float closest = 0.f;                                                                 // :407-416
    bool first = true;
    for (int p = 0; p < jp.n_passages; ++p)
      if (!jp.pcollide[p]) {
        const float d = norm(jpos - q_pos(jp.passage[p], i));
        closest = first ? d : pmin(closest, d);
        first = false;
      }
    float pos_rew = 0.f, rot_rew = 0.f;                                                   // :396-397
    const float sh_pre = closest * jp.pos_factor;                                        // :417
    if (!joint_passed) pos_rew = pos_rew + (jp.sh_in[0][i] - sh_pre);                    // :418-420
    const V2 gpos = mk(vget(jp.goal_pos, i, 0), vget(jp.goal_pos, i, 1));
    const float dist_goal = norm(jpos - gpos);                                           // :423-426
    const float sh_post = dist_goal * jp.pos_factor;
    if (joint_passed) pos_rew = pos_rew + (jp.sh_in[1][i] - sh_post);                    // :428-430
    const bool rot_passed = jp.all_passed_rot ? all_passed : joint_passed;               // :434
    const float rsh_pre = line_angle_dist(jrot, jp.mid_g, jp.mid_g_minus_pi, jp.pi_f) * jp.rot_factor;   // :435-438
    if (!rot_passed) rot_rew = rot_rew + (jp.sh_in[2][i] - rsh_pre);                     // :439-441
    const float g = remainder_f(vget(jp.goal_rot, i, 0), jp.pi_f);
    const float rot_dist_goal = line_angle_dist(jrot, g, g - jp.pi_f, jp.pi_f);          // :444-446
    const float rsh_post = rot_dist_goal * jp.rot_factor;
    if (rot_passed) rot_rew = rot_rew + (jp.sh_in[3][i] - rsh_post);                     // :448-450
    jp.sh_out[0][i] = sh_pre; jp.sh_out[1][i] = sh_post; jp.sh_out[2][i] = rsh_pre; jp.sh_out[3][i] = rsh_post;
    float coll = 0.f;                                                                    // :398, 453-473
    if (jp.coll_rew != 0.f) {  // k additions of the same constant; adding 0 any number of times leaves 0
      int k = 0;
      for (int a = 0; a < jp.n_bodies; ++a) {
        int walls = 0;
        for (int w = 0; w < jp.n_walls; ++w) walls += pair_distance(jp.body[a], jp.wall[w], i) <= jp.min_coll;
        k += walls * jp.n_passages;                                                      // the wall loop sits INSIDE the passage loop (:461)
        for (int p = 0; p < jp.n_passages; ++p)
          if (jp.pcollide[p]) k += pair_distance(jp.body[a], jp.passage[p], i) <= jp.min_coll;
      }
      for (int p = 0; p < jp.n_passages; ++p)
        if (jp.pcollide[p]) k += pair_distance(jp.passage[p], jp.joint, i) <= jp.min_coll;
      for (int q = 0; q < k; ++q) coll = coll + jp.coll_rew;
    }
    float energy = 0.f;                                                                  // :476-484 (sum over <= 2 agents)
    for (int a = 0; a < jp.n_agents; ++a) {
      const float e = norm(mk(vget(jp.action_u[a], i, 0), vget(jp.action_u[a], i, 1))) / jp.e_div[a];
      energy = a == 0 ? e : energy + e;
    }
    const float energy_rew = (-energy) * jp.energy_coeff;                                // :485
    jp.pos_rew[i] = pos_rew; jp.rot_rew[i] = rot_rew; jp.collision_rew[i] = coll;
    jp.energy_exp[i] = energy; jp.energy_rew[i] = energy_rew;
    jp.rew[i] = ((pos_rew + rot_rew) + coll) + energy_rew;                               // :487-489
    jp.all_passed[i] = all_passed ? 1 : 0;
    if (jp.done) {
      jp.done[i] = rot_dist_goal <= 0.01f ? 1 : 0;                                       // :555-560
      pos_bad = !(dist_goal <= 0.01f);                                                   // :549-554
    }
  }
  if (jp.done) {
    const unsigned any = __ballot_sync(0xffffffffu, pos_bad);
    if (any && (threadIdx.x & 31) == 0 && !__ldcg(jp.done_flag)) atomicOr(jp.done_flag, 1u);
  }
}
__global__ void __launch_bounds__(256) jp_done_kernel(uint8_t* __restrict__ done, const uint32_t* __restrict__ flag, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && *flag) done[i] = 0;
}

// ------------------------------------------------------------------------------------------------ football
struct FbTeam {
  int n_agents, dense;
  DVec apos[VMASB200_FB_MAX_TEAM];
  float gx, gy;
  const float* sh_in;
  float* sh_out;
  const float* sha_in;
  float* sha_out;
  float *dist_goal, *pos_rew, *min_dist, *pos_rew_agent, *dense_rew, *sparse_rew, *reward;
};
struct FbParams {
  DVec ball_pos, ball_vel;
  float x_right, x_left, y_hi, y_lo, scoring, f_ball_goal, f_agent_ball, trigger;
  FbTeam team[2];
  uint8_t* done;
};
// football.py:1118-1226 (reward of the first agent; reward_ball_to_goal; reward_all_agent_to_ball with
// get_closest_agent_to_ball :585-599, whose torch.cdist on these operands is ATen's scalar loop agg += diff * diff,
// contracted by the compiler: sqrt(fma(dy, dy, dx * dx)), the form of linalg.vector_norm; pinned by the golden vectors)
__global__ void __launch_bounds__(256) fb_reward_kernel(const __grid_constant__ FbParams fb, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const V2 b = mk(vget(fb.ball_pos, i, 0), vget(fb.ball_pos, i, 1));
  const bool over_right = b.x > fb.x_right, over_left = b.x < fb.x_left;                 // :1122-1127
  const bool goal_mask = (b.y <= fb.y_hi) && (b.y >= fb.y_lo);                           // :1128-1130
  const bool blue_score = over_right && goal_mask, red_score = over_left && goal_mask;
  const float sparse_blue = fb.scoring * (blue_score ? 1.f : 0.f) - fb.scoring * (red_score ? 1.f : 0.f);  // :1133-1135
  const float sparse[2] = {sparse_blue, -sparse_blue};                                   // :1136
  fb.done[i] = (blue_score || red_score) ? 1 : 0;                                        // :1138
  const V2 bv = mk(vget(fb.ball_vel, i, 0), vget(fb.ball_vel, i, 1));
  const bool ball_moving = norm(bv) > 1e-6f;                                             // :1197
  for (int t = 0; t < 2; ++t) {
    const FbTeam& tm = fb.team[t];
    float dense = 0.f;                                                                   // :1140-1141
    if (tm.dense) {
      const float dg = norm(b - mk(tm.gx, tm.gy));                                       // :1160-1171
      const float sh = dg * fb.f_ball_goal;                                              // :1173
      const float pr = tm.sh_in[i] - sh;                                                 // :1176, 1180
      float md = 0.f;
      for (int a = 0; a < tm.n_agents; ++a) {                                            // :585-599
        const float d = norm(mk(vget(tm.apos[a], i, 0), vget(tm.apos[a], i, 1)) - b);
        md = a == 0 ? d : pmin(md, d);
      }
      const float sha = md * fb.f_agent_ball;                                            // :1195
      const bool close = md < fb.trigger;                                                // :1198
      const float pra = (close || ball_moving) ? 0.f : (tm.sha_in[i] - sha);             // :1201-1215
      tm.dist_goal[i] = dg; tm.sh_out[i] = sh; tm.pos_rew[i] = pr;
      tm.min_dist[i] = md; tm.sha_out[i] = sha; tm.pos_rew_agent[i] = pra;
      dense = pr + pra;                                                                  // :1144-1151
      tm.dense_rew[i] = dense;
    }
    tm.sparse_rew[i] = sparse[t];
    tm.reward[i] = sparse[t] + dense;                                                    // :1153-1157
  }
}

}  // namespace vb
