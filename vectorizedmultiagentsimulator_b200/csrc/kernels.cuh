// kernels.cuh - __global__ entry points: fused step (+ verified repair), batch-wide broad-phase flags, cast_rays.
//
// Exactness of the fused (speculative) launch.  The reference decides per (substep, pair) for the WHOLE batch whether
// the pair is evaluated: World.collides ends in (||p_a - p_b|| <= Rc_a + Rc_b).any() over envs (core.py:2796-2800).
// The fused kernel considers every pair in every env.  That differs from the reference only for an env whose own
// verdict is false while its force is non-zero ("thin" event: contact through the LINE_MIN_DIST shell at a box
// corner / line end) AND no env of the batch has a true verdict.  Each env therefore tracks a taint (set after its
// first thin event); untainted envs OR their verdicts into any_exact, tainted ones into any_spec and, per env, into a
// record (u_verd) next to the record of their thin events (u_thin).  The last CTA to finish checks
// thin & ~any_exact == 0, which proves (by induction over substeps) that every env equals the reference.  Otherwise
// it walks the substeps once (validate_speculation): an env whose thin events so far were all on pairs that exact
// sources activate is exact itself, so its verdicts join the exact flags of the next substep; the envs left over -
// typically the one or two whose thin force the reference would really have skipped - are re-simulated from their
// saved input state, substep by substep, with flags = exact flags | the re-simulated envs' verdicts, which is exactly
// the reference's flag.
#pragma once
#include "step_kernel.cuh"

namespace vb {

struct StepArgs {
  const float* dyn_in;       // [dyn_rows][ld]
  float* dyn_out;            // [dyn_rows][ld]
  const float* stat;         // [stat_rows][ld]
  float* act_out;            // [act_rows][ld]
  float* lidar_out;          // [lidar_rows][ld]
  long long n_envs, ld;
  long long env_begin, env_end;  // envs of this launch (a step may be cut into several launches on several streams)
  int total_ctas;            // CTAs of ALL launches of the step: the one that finishes last verifies the speculation
  int s_begin, s_end;        // substeps [s_begin, s_end) of this launch
  const uint32_t* flags_in;  // [S][pair_words] batch-wide World.collides verdicts (exact path), or NULL = speculative
  // speculative mode scratch (all owned by the plan, zero between launches)
  uint32_t* flags_out;       // [S][pair_words] result: the reference's verdicts
  uint32_t* any_exact;       // [S][pair_words]
  uint32_t* any_spec;        // [S][pair_words]
  uint32_t* thin;            // [S][pair_words]
  int* counters;             // [0] finished CTAs, [1] tainted envs, [2] repairs done since plan creation, [3] repair_kernel wanted
  int* u_list;               // [n_envs] env index of every tainted env (bit 31: to be re-simulated)
  float* u_buf;              // [n_envs][dyn_rows] their input state (then their evolving state)
  int* u_rep;                // [n_envs] indices into u_list of the envs to re-simulate
  uint32_t* u_verd;          // [rec_cap][S][pair_words] verdicts of tainted env i from its taint on
  uint32_t* u_thin;          // [rec_cap][S][pair_words] its thin events
  int rec_cap;               // tainted envs beyond this have no record and are re-simulated when anything is
  int* repairs_host;         // mapped host copy of counters[2] (read by the host-buffer entry point without a copy)
  int do_lidar;
  // probe fields of the exact path (vmasb200_step_io: parity tests), all 0 / NULL otherwise
  int forces_fresh;          // the frc inputs are this launch's substep inputs: no replay of the earlier substeps' clamps
  const float* trig_in;      // [4 * n_trig][ld] cos/sin cache to use instead of computing it
  const float* jexp_in;      // [J][ld] exp(|delta rot|) of every joint's constraint torque (one-substep launches, c_probe_jexp)
  float* dbg_force;          // [3 * n_dyn][ld] (Fx Fy T) of the launch's last substep
  float* dbg_joint;          // [4 * J][ld] (fx fy ta tb) per joint of the launch's last substep
};

// dynamic shared memory carve-up (4-byte words); every row starts 16B aligned because TE % 4 == 0
struct SmemMap {
  int bar, env, work, lid, queue, rayq, flags, total_bytes;
};
constexpr int kThreads = 128;   // threads per CTA; a team of Q = kThreads / TE members works on every env of the tile
__host__ __device__ inline SmemMap smem_map(const KParams& kp) {
  SmemMap m;
  int o = 0;
  m.bar = o; o += 4;  // 16 bytes: mbarrier (8) + the two LIDAR queue counters (4 + 4)
  m.env = o; o += kp.TE * (kp.dyn_rows + kp.stat_rows);
  m.work = o; o += kp.TE * kp.work_rows;
  m.lid = o; o += kp.TE * kp.lidar_rows;
  m.queue = o; o += kp.TE * kp.n_sensors;
  m.rayq = o; o += kp.TE * kp.lidar_rows;
  m.flags = o; o += kp.S * kp.pair_words;
  m.total_bytes = o * 4;
  return m;
}

struct Tile {
  float *envS, *workS, *lidS;
  int *queue, *rayq, *qcount;  // qcount[0]: sensor queue length, qcount[1]: ray queue length
};
template <int TE>
VB_DI EnvT<TE> env_view(const Tile& t, int col) {
  EnvT<TE> E;
  E.env = t.envS + col; E.work = t.workS + col;
  return E;
}

VB_DI void or_global(uint32_t* word, uint32_t v) {
  if (v & ~__ldcg(word)) atomicOr(word, v);
}

// agent.state.force / torque of env `env` into the env's frc rows (agents q, q + Q, ...)
template <class EV>
VB_DI void load_forces(const KParams& kp, const AgentIO& aio, const EV& E, long long env, int q, int Q) {
  for (int a = q; a < kp.A; a += Q) {
    float fx = 0.f, fy = 0.f, tq = 0.f;
    const float* fp = aio.force[a];
    const float* tp = aio.torque[a];
    if (fp) { fx = fp[env * aio.fstride[a]]; fy = fp[env * aio.fstride[a] + 1]; }
    if (tp) tq = tp[env * aio.tstride[a]];
    const int fr = kp.row_frc + 3 * a;
    WR(E, fr) = fx; WR(E, fr + 1) = fy; WR(E, fr + 2) = tq;
  }
}

// Substeps [s_begin, s_end) of the tile.  EVERY thread of the CTA calls this (CTA barriers inside): thread t is
// member q = t / TE of the team of column t % TE.  flags: batch-wide verdicts to obey (smem or global), or NULL =
// speculative (every pair considered; verdict / thin bits OR-ed into sa.any_exact / any_spec / thin).  collect (= flags,
// shared memory): the tile's own verdicts are OR-ed into the flags before they are obeyed (a tile that holds every env
// whose verdicts are still missing from the flags).
// Work split of a substep (tables.h): joints, 8-pair slices of the broad phase, narrow-phase survivors (in the env's
// pair order, survivor i to member i mod Q), then per dynamic entity the ordered sum of its forces + integration.
// env: index of the column's environment.  Returns the column's taint: 0, or 1 + its index in sa.u_list.
template <bool CT, int TE>
VB_DI int run_substeps(const KParams& kp, const StepArgs& sa, const Tile& t, bool ok, long long env, int s_begin,
                       int s_end, const uint32_t* flags, uint32_t* collect = nullptr) {
  constexpr int Q = kThreads / TE;
  constexpr int R = kSlots / Q;     // survivors a member evaluates per round
  const int col = threadIdx.x % TE, q = threadIdx.x / TE;
  const int lane = threadIdx.x & 31;
  const EnvT<TE> E = env_view<TE>(t, col);
  __syncthreads();  // the caller's tile loads
  // prologue: cos/sin cache of every entity that needs it, first force accumulators
  for (int r = q; r < kp.n_trig; r += Q) {
    const DEnt* en = &T_ent<CT>(kp, T_trig<CT>(kp, r));
    if (sa.trig_in) {  // probe: the caller's cos/sin (the reference's own torch.cos / torch.sin values)
      if (ok)
        for (int k = 0; k < 4; ++k) WR(E, en->trig_row + k) = sa.trig_in[(size_t)(4 * r + k) * sa.ld + env];
    } else {
      store_trig(en, E, ER(E, en->rot_row));
    }
  }
  // the clamp of core.py:2017-2040 is re-applied to its own result every substep: replay it for the substeps
  // that earlier launches of a split step already did
  for (int d = q; d < kp.n_dyn; d += Q)
    for (int rep = sa.forces_fresh ? s_begin : 0; rep <= s_begin; ++rep) begin_forces(kp, &T_ent<CT>(kp, T_dyn<CT>(kp, d)), E);
  if (sa.jexp_in && ok)
    for (int j = q; j < kp.J; j += Q) WR(E, kp.row_jres + 4 * j) = sa.jexp_in[(size_t)j * sa.ld + env];
  if (q == 0) {
    WR(E, kp.row_taint) = 0.f;
    for (int b = 0; b <= kp.n_blk; ++b) WR(E, kp.row_th + b) = 0.f;
  }
  __syncthreads();

  const int n_oct = kp.n_blk * 4, n_real = (kp.P + 7) >> 3;  // octets of the sv / bp words; octets that hold pairs
  for (int s = s_begin; s < s_end; ++s) {
    // ---- joints (core.py:2175, always evaluated), from the last member down (the broad phase starts at member 0)
    for (int j = Q - 1 - q; j < kp.J; j += Q) {
      V2 fa; float ta, tb;
      eval_item<CT>(kp, kp.items + j, E, fa, ta, tb);
      const int r = kp.row_jres + 4 * j;
      WR(E, r) = fa.x; WR(E, r + 1) = fa.y; WR(E, r + 2) = ta; WR(E, r + 3) = tb;
    }
    // ---- broad phase, 8 candidate pairs at a time
    const int tix = __float_as_int(WR(E, kp.row_taint));
    const bool tainted = tix != 0;
    for (int u = q; u < n_oct; u += Q) {
      uint32_t bpw = 0u, sv = 0u;
      // The table is padded to a multiple of 32 rows with rows that never pass (plan.cpp).
      if (u < n_real)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        // branch-free, so that the bodies interleave: a circle row carries hx = hy = +inf (and any valid trig
        // row), a frame row t2_cull = +inf; NaN passes every test and is left to the narrow phase
        const DBroad& bq = T_broad<CT>(kp, u * 8 + k);
        const int ra = bq.pa_row, rb = bq.pb_row, rt = bq.trig_row;
        const float dx = ER(E, ra) - ER(E, rb), dy = ER(E, ra + 1) - ER(E, rb + 1);
        const float d2 = fmaf(dy, dy, dx * dx);
        const float c = WR(E, rt), sn = WR(E, rt + 1);
        const float lx = fmaf(dx, c, dy * sn), ly = fmaf(dx, sn, -(dy * c));  // b's centre in a's frame
        bpw |= (d2 <= bq.t2_bp ? 1u : 0u) << k;  // == (norm(pa - pb) <= Rc_a + Rc_b), core.py:2796-2800
        sv |= ((!(d2 > bq.t2_cull) && !(fabsf(lx) > bq.hx) && !(fabsf(ly) > bq.hy)) ? 1u : 0u) << k;
      }
      const int n = kp.P - u * 8;  // the mask only matters for NaN state in the padding rows
      const uint32_t valid = n >= 8 ? 0xffu : (n > 0 ? (1u << n) - 1u : 0u);
      bpw &= valid; sv &= valid;
      const int w = s * kp.pair_words + (u >> 2), sh = 8 * (u & 3);
      if (flags) {  // obeyed when the survivors are taken up below
        if (collect) {
          const uint32_t e = __reduce_or_sync(0xffffffffu, ok ? bpw : 0u) << sh;
          if (lane == 0 && (e & ~collect[w])) atomicOr(collect + w, e);
        }
      } else {
        const uint32_t e = __reduce_or_sync(0xffffffffu, (ok && !tainted) ? bpw : 0u);
        const uint32_t sp = __reduce_or_sync(0xffffffffu, (ok && tainted) ? bpw : 0u);
        if (lane == 0) {
          if (e) or_global(sa.any_exact + w, e << sh);
          if (sp) or_global(sa.any_spec + w, sp << sh);
        }
        if (ok && tainted && tix <= sa.rec_cap)
          reinterpret_cast<uint8_t*>(sa.u_verd + ((size_t)(tix - 1) * kp.S + s) * kp.pair_words + (u >> 2))[u & 3] = (uint8_t)bpw;
      }
      if (!ok) sv = 0u;
      reinterpret_cast<uint8_t*>(&WR(E, kp.row_sv + (u >> 2)))[u & 3] = (uint8_t)sv;
      reinterpret_cast<uint8_t*>(&WR(E, kp.row_bp + (u >> 2)))[u & 3] = (uint8_t)bpw;
    }
    __syncthreads();
    // ---- narrow phase of the survivors in rounds of kSlots per env, then the ordered sums
    for (int lo = 0;; lo += kSlots) {
      int cnt = 0;
      int task[R];
#pragma unroll
      for (int r = 0; r < R; ++r) task[r] = -1;
      for (int b = 0; b < kp.n_blk; ++b) {
        uint32_t m = __float_as_uint(WR(E, kp.row_sv + b));
        if (flags) m &= flags[s * kp.pair_words + b];
        while (m) {
          const int k = __ffs(m) - 1;
          m &= m - 1u;
#pragma unroll
          for (int r = 0; r < R; ++r)
            if (cnt == lo + q + r * Q) task[r] = b * 32 + k;
          ++cnt;
        }
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int p = task[r];
        if (p >= 0) {
          V2 fa; float ta, tb;
          eval_item<CT>(kp, kp.items + kp.J + p, E, fa, ta, tb);
          const int pr = kp.row_pres + 5 * (q + r * Q);
          WR(E, pr) = fa.x; WR(E, pr + 1) = fa.y; WR(E, pr + 2) = ta; WR(E, pr + 3) = tb;
          WR(E, pr + 4) = __int_as_float(p);
          if (!flags && (fa.x != 0.f || fa.y != 0.f) &&
              !((__float_as_uint(WR(E, kp.row_bp + (p >> 5))) >> (p & 31)) & 1u)) {
            or_global(sa.thin + s * kp.pair_words + (p >> 5), 1u << (p & 31));
            atomicOr(reinterpret_cast<uint32_t*>(&WR(E, kp.row_th + 1 + (p >> 5))), 1u << (p & 31));
            WR(E, kp.row_th) = 1.f;  // member 0 files the substep's thin events below
          }
        }
      }
      const int more = __syncthreads_or(cnt > lo + kSlots);
      if (lo == 0)
        for (int j = 0; j < kp.J; ++j) {
          const int r = kp.row_jres + 4 * j;
          accumulate<Q>(kp.items + j, E, mk(WR(E, r), WR(E, r + 1)), WR(E, r + 2), WR(E, r + 3), q);
        }
      const int n = min(cnt - lo, kSlots);
      for (int i = 0; i < n; ++i) {
        const int pr = kp.row_pres + 5 * i;
        accumulate<Q>(kp.items + kp.J + __float_as_int(WR(E, pr + 4)), E, mk(WR(E, pr), WR(E, pr + 1)), WR(E, pr + 2),
                      WR(E, pr + 3), q);
      }
      if (!more) break;
      __syncthreads();  // the result slots are reused
    }
    // ---- integrate the entities this member sums, then open their next accumulators
    for (int d = q; d < kp.n_dyn; d += Q) {
      const DEnt* en = &T_ent<CT>(kp, T_dyn<CT>(kp, d));
      integrate(kp, en, E, s);
      if (s + 1 < s_end) begin_forces(kp, en, E);
    }
    // ---- thin events of this substep: the env is tainted from the NEXT substep on and keeps a record
    if (!flags && q == 0 && WR(E, kp.row_th) != 0.f) {
      int tx = __float_as_int(WR(E, kp.row_taint));
      const size_t rw = (size_t)kp.S * kp.pair_words;
      if (tx == 0) {
        tx = atomicAdd(sa.counters + 1, 1) + 1;
        sa.u_list[tx - 1] = (int)env;
        WR(E, kp.row_taint) = __int_as_float(tx);
        if (tx <= sa.rec_cap)
          for (size_t i = 0; i < rw; ++i) { sa.u_verd[(tx - 1) * rw + i] = 0u; sa.u_thin[(tx - 1) * rw + i] = 0u; }
      }
      for (int b = 0; b < kp.n_blk; ++b) {
        const uint32_t w = __float_as_uint(WR(E, kp.row_th + 1 + b));
        if (w) {
          if (tx <= sa.rec_cap) sa.u_thin[(tx - 1) * rw + (size_t)s * kp.pair_words + b] = w;
          WR(E, kp.row_th + 1 + b) = 0.f;
        }
      }
      WR(E, kp.row_th) = 0.f;
    }
    __syncthreads();
  }
  return __float_as_int(WR(E, kp.row_taint));
}

// LIDAR epilogue for the whole tile (every thread of the CTA calls it, helper warps included; contains CTA barriers).
// n_cols = columns of the tile that hold real environments.  Three passes, each on full warps:
//   1  thread per (env, sensor mod Q): per sensor the mask of targets a ray could reach at all; sensors that see something go to
//      the sensor queue (and get cos/sin of their rotation cached); every ray starts at max_range;
//   2  thread per (queued sensor, ray): ray-level cull with an APPROXIMATE direction (angle-sum formula) and a slack
//      that covers the approximation: a target farther than Rc from the ray's line, or behind the origin, cannot be
//      hit (core.py:1467-1489 sphere, :1344-1371 box, :1598-1625 line); rays that may hit go to the ray queue;
//   3  thread per queued ray: the reference's exact arithmetic (sincosf of the fp32 angle sum, sensors.py:117).
template <bool CT, int TE>
VB_DI void lidar_epilogue(const KParams& kp, const Tile& t, int n_cols) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  if (tid < 2) t.qcount[tid] = 0;
  for (int k = 0; k < kp.n_sensors; ++k) {  // every ray starts at max_range
    const DSensor* sn = &T_sensor<CT>(kp, k);
    const float mr = sn->max_range;
    for (int i = tid; i < sn->n_rays * TE; i += nthr) t.lidS[sn->out_off * TE + i] = mr;
  }
  __syncthreads();
  {
    constexpr int Q = kThreads / TE;
    const int col = tid % TE;
    const EnvT<TE> E = env_view<TE>(t, col);
    for (int k = tid / TE; k < kp.n_sensors; k += Q) {
      const DSensor* sn = &T_sensor<CT>(kp, k);
      const DEnt* se = &T_ent<CT>(kp, sn->entity);
      const V2 o = pos_at(E, se->pos_row);
      uint32_t any = 0u;
      for (int w = 0; w < sn->n_tw; ++w) {
        const int tb = sn->tg_begin + 32 * w, nt = min(32, sn->n_targets - 32 * w);
        uint32_t m = 0u;
        for (int q = 0; q < nt; ++q) {
          const V2 p = pos_at(E, T_ent<CT>(kp, __ldg(kp.targets + tb + q)).pos_row);
          const float dx = o.x - p.x, dy = o.y - p.y;
          if (!(fmaf(dy, dy, dx * dx) > __ldg(&kp.tcull[tb + q].x))) m |= 1u << q;
        }
        WR(E, sn->tw_off + w) = __uint_as_float(m);
        any |= m;
      }
      if (any && col < n_cols) {
        float s, c;
        sincosf(ER(E, se->rot_row), &s, &c);
        WR(E, kp.row_strig + 2 * k) = c; WR(E, kp.row_strig + 2 * k + 1) = s;
        t.queue[atomicAdd(t.qcount, 1)] = col | (k << 16);
      }
    }
  }
  __syncthreads();
  const int total = t.qcount[0] * kp.max_rays;
  for (int task = tid; task < total; task += nthr) {
    const int qi = task / kp.max_rays, r = task - qi * kp.max_rays;
    const int item = t.queue[qi], k = item >> 16, col = item & 0xffff;
    const DSensor* sn = &T_sensor<CT>(kp, k);
    if (r >= sn->n_rays) continue;
    const EnvT<TE> E = env_view<TE>(t, col);
    const DEnt* se = &T_ent<CT>(kp, sn->entity);
    const V2 o = pos_at(E, se->pos_row);
    const float c0 = WR(E, kp.row_strig + 2 * k), s0 = WR(E, kp.row_strig + 2 * k + 1);
    const float2 cs = __ldg(kp.ray_cs + sn->out_off + r);
    const float dc = c0 * cs.x - s0 * cs.y, ds = s0 * cs.x + c0 * cs.y;  // ~ cos/sin(angle_r + rot)
    const float aslack = fabsf(ER(E, se->rot_row)) + 8.f;                // radians the angle sum may have rounded over
    bool cand = false;
    for (int w = 0; w < sn->n_tw; ++w) {
      uint32_t m = __float_as_uint(WR(E, sn->tw_off + w));
      const int tb = sn->tg_begin + 32 * w;
      while (m) {
        const int q = __ffs(m) - 1;
        m &= m - 1u;
        const float4 tc = __ldg(kp.tcull + tb + q);
        const V2 p = pos_at(E, T_ent<CT>(kp, __ldg(kp.targets + tb + q)).pos_row);
        const float rx = p.x - o.x, ry = p.y - o.y;
        const float slack = tc.w * aslack;
        const float lat = fabsf(rx * ds - ry * dc), along = rx * dc + ry * ds;
        cand = cand || (!(lat > tc.y + slack) && !(along < -(tc.z + slack)));
      }
    }
    if (cand) t.rayq[atomicAdd(t.qcount + 1, 1)] = col | (k << 8) | (r << 16);
  }
  __syncthreads();
  const int nq = t.qcount[1];
  for (int qi = tid; qi < nq; qi += nthr) {
    const int item = t.rayq[qi], col = item & 0xff, k = (item >> 8) & 0xff, r = item >> 16;
    const DSensor* sn = &T_sensor<CT>(kp, k);
    const EnvT<TE> E = env_view<TE>(t, col);
    const DEnt* se = &T_ent<CT>(kp, sn->entity);
    const float ang = __ldg(kp.ray_angles + sn->out_off + r) + ER(E, se->rot_row);  // sensors.py:117
    t.lidS[(sn->out_off + r) * TE + col] = cast_one_ray<CT>(kp, sn, E, pos_at(E, se->pos_row), ang);
  }
  __syncthreads();
}

// One pass over the substeps by the last CTA (rare path; see the header comment): which tainted envs are exact anyway.
// flagS (smem, [S][pair_words]) ends as any_exact | the verdicts of those envs; the others are listed in sa.u_rep.
// Returns how many those are.
__device__ __noinline__ int validate_speculation(const KParams& kp, const StepArgs& sa, uint32_t* flagS, int nu) {
  __shared__ int s_nd;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int pw = kp.pair_words, nw = kp.S * pw;
  for (int w = tid; w < nw; w += nthr) flagS[w] = __ldcg(sa.any_exact + w);
  if (tid == 0) s_nd = 0;
  const int nrec = min(nu, sa.rec_cap);
  for (int i = nrec + tid; i < nu; i += nthr) sa.u_list[i] |= 0x80000000;  // no record: cannot be cleared
  __syncthreads();
  for (int s = 0; s < kp.S; ++s) {
    // verdicts of the envs that are exact up to here are reference verdicts
    for (int i = tid; i < nrec; i += nthr)
      if (sa.u_list[i] >= 0)
        for (int w = 0; w < pw; ++w) {
          const uint32_t v = __ldcg(sa.u_verd + ((size_t)i * kp.S + s) * pw + w);
          if (v & ~flagS[s * pw + w]) atomicOr(flagS + s * pw + w, v);
        }
    __syncthreads();
    // a thin event on a pair nobody exact activates: the env leaves the reference's trajectory after this substep
    for (int i = tid; i < nrec; i += nthr)
      if (sa.u_list[i] >= 0) {
        bool off = false;
        for (int w = 0; w < pw; ++w) off = off || (__ldcg(sa.u_thin + ((size_t)i * kp.S + s) * pw + w) & ~flagS[s * pw + w]);
        if (off) sa.u_list[i] |= 0x80000000;
      }
    __syncthreads();
  }
  for (int i = tid; i < nu; i += nthr)
    if (sa.u_list[i] < 0) {
      sa.u_list[i] &= 0x7fffffff;
      sa.u_rep[atomicAdd(&s_nd, 1)] = i;
    }
  __syncthreads();
  return s_nd;
}

// Re-simulation of the nd envs listed in sa.u_rep by the last CTA.  flagS comes from validate_speculation and ends as
// the reference's verdicts.  u_buf holds the evolving state of tainted env `idx` as dyn_rows consecutive floats; each
// team works on one of them at a time in its own tile column.
template <bool CT, int TE>
__device__ __noinline__ void repair_tainted(const KParams& kp, const AgentIO& aio, const StepArgs& sa, const Tile& t,
                                            uint32_t* flagS, int nd) {
  constexpr int Q = kThreads / TE;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int col = tid % TE, q = tid / TE;
  const EnvT<TE> E = env_view<TE>(t, col);
  if (nd <= TE) {
    // the usual case, one tile: it stays in shared memory for all substeps and adds its own verdicts to the flags
    const bool ok = col < nd;
    const int idx = ok ? sa.u_rep[col] : 0;
    const long long env = ok ? sa.u_list[idx] : 0;
    if (ok) {
      for (int r = q; r < kp.dyn_rows; r += Q) ER(E, r) = sa.u_buf[(size_t)idx * kp.dyn_rows + r];
      for (int r = q; r < kp.stat_rows; r += Q) ER(E, kp.dyn_rows + r) = sa.stat[(size_t)r * sa.ld + env];
      load_forces(kp, aio, E, env, q, Q);
    }
    run_substeps<CT, TE>(kp, sa, t, ok, env, 0, kp.S, flagS, flagS);
    if (sa.lidar_out && kp.n_rays) lidar_epilogue<CT, TE>(kp, t, nd);
    if (ok) {
      for (int r = q; r < kp.dyn_rows; r += Q) sa.dyn_out[(size_t)r * sa.ld + env] = ER(E, r);
      if (sa.lidar_out)
        for (int r = q; r < kp.lidar_rows; r += Q) sa.lidar_out[(size_t)r * sa.ld + env] = t.lidS[r * TE + col];
    }
    __syncthreads();
    return;
  }
  for (int s = 0; s <= kp.S; ++s) {  // s == S: LIDAR + write-back pass
    if (s < kp.S) {
      // pass 1: verdicts of the tainted envs at their true state
      for (int i = tid; i < nd; i += nthr) {
        const int idx = sa.u_rep[i];
        const float* dyn = sa.u_buf + (size_t)idx * kp.dyn_rows;
        const long long env = sa.u_list[idx];
        auto px = [&](int row, int k) {
          return row < kp.dyn_rows ? dyn[row + k] : sa.stat[(size_t)(row - kp.dyn_rows + k) * sa.ld + env];
        };
        for (int p = 0; p < kp.P; ++p) {
          const DBroad* bq = &T_broad<CT>(kp, p);
          const float dx = px(bq->pa_row, 0) - px(bq->pb_row, 0), dy = px(bq->pa_row, 1) - px(bq->pb_row, 1);
          if (fmaf(dy, dy, dx * dx) <= bq->t2_bp) {
            uint32_t* w = flagS + s * kp.pair_words + (p >> 5);
            const uint32_t bit = 1u << (p & 31);
            if (!(*w & bit)) atomicOr(w, bit);
          }
        }
      }
      __syncthreads();
    }
    // pass 2: one substep with the now exact flags (or the epilogue), a tile of TE tainted envs at a time
    for (int base = 0; base < nd; base += TE) {
      const bool ok = base + col < nd;
      const int idx = ok ? sa.u_rep[base + col] : 0;
      const long long env = ok ? sa.u_list[idx] : 0;
      if (ok) {
        for (int r = q; r < kp.dyn_rows; r += Q) ER(E, r) = sa.u_buf[(size_t)idx * kp.dyn_rows + r];
        for (int r = q; r < kp.stat_rows; r += Q) ER(E, kp.dyn_rows + r) = sa.stat[(size_t)r * sa.ld + env];
        load_forces(kp, aio, E, env, q, Q);
      }
      const bool last = s == kp.S;
      // s == S: prologue only (cos/sin cache, clamps)
      run_substeps<CT, TE>(kp, sa, t, ok, env, s, last ? s : s + 1, flagS);
      if (!last) {
        if (ok) for (int r = q; r < kp.dyn_rows; r += Q) sa.u_buf[(size_t)idx * kp.dyn_rows + r] = ER(E, r);
      } else {
        if (sa.lidar_out && kp.n_rays) lidar_epilogue<CT, TE>(kp, t, min(TE, nd - base));
        if (ok) {
          for (int r = q; r < kp.dyn_rows; r += Q) sa.dyn_out[(size_t)r * sa.ld + env] = ER(E, r);
          if (sa.lidar_out)
            for (int r = q; r < kp.lidar_rows; r += Q) sa.lidar_out[(size_t)r * sa.ld + env] = t.lidS[r * TE + col];
        }
      }
      __syncthreads();
    }
  }
}

// Fused step of one tile of TE environments.  Registers: kThreads * 7 CTAs of TE = 32 must fit the register file for
// one wave at 32768 envs, hence the tighter bound for the 4-member teams.
template <bool CT, int TE>
__global__ void __launch_bounds__(kThreads, TE == 32 ? 7 : 4)
step_kernel(const __grid_constant__ KParams kp, const __grid_constant__ AgentIO aio, const __grid_constant__ StepArgs sa) {
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_last;
  constexpr int Q = kThreads / TE;
  const SmemMap m = smem_map(kp);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + m.bar);
  Tile t;
  t.envS = smem + m.env; t.workS = smem + m.work; t.lidS = smem + m.lid;
  t.queue = reinterpret_cast<int*>(smem + m.queue); t.rayq = reinterpret_cast<int*>(smem + m.rayq);
  t.qcount = reinterpret_cast<int*>(smem + m.bar) + 2;
  uint32_t* flagS = reinterpret_cast<uint32_t*>(smem + m.flags);
  const int nw = kp.S * kp.pair_words;

  const int tid = threadIdx.x;
  const int col = tid % TE, q = tid / TE;
  const long long env0 = sa.env_begin + (long long)blockIdx.x * TE;
  const int n_valid = (int)min((long long)TE, sa.env_end - env0);  // real environments of the tile
  // columns moved by TMA (multiple of 4): the batch's last tile also carries the padding columns up to ld
  const int n_cols = (int)min((long long)TE, (sa.env_end == sa.n_envs ? sa.ld : sa.env_end) - env0);
  const bool ok = col < n_valid;
  const bool spec = sa.flags_in == nullptr;
  const uint32_t row_bytes = (uint32_t)n_cols * 4u;
  const int n_in = kp.dyn_rows + kp.stat_rows;

  // ---- stage the tile: one TMA bulk copy per SoA row, plain loads for the per-agent force tensors
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  __syncthreads();
  if (tid < 32) {
    if (tid == 0) mbar_expect_tx(bar, row_bytes * (uint32_t)n_in);
    __syncwarp();
    for (int r = tid; r < n_in; r += 32) {
      const float* src = r < kp.dyn_rows ? sa.dyn_in + (size_t)r * sa.ld + env0
                                         : sa.stat + (size_t)(r - kp.dyn_rows) * sa.ld + env0;
      bulk_g2s(t.envS + r * TE, src, row_bytes, bar);
    }
  }
  const EnvT<TE> E = env_view<TE>(t, col);
  if (ok) load_forces(kp, aio, E, env0 + col, q, Q);
  mbar_wait(bar, 0);
  const int taint = run_substeps<CT, TE>(kp, sa, t, ok, env0 + col, sa.s_begin, sa.s_end, sa.flags_in);
  if (sa.dbg_force && ok)
    for (int r = q; r < 3 * kp.n_dyn; r += Q) sa.dbg_force[(size_t)r * sa.ld + env0 + col] = WR(E, kp.row_fin + r);
  if (sa.dbg_joint && ok)
    for (int r = q; r < 4 * kp.J; r += Q) sa.dbg_joint[(size_t)r * sa.ld + env0 + col] = WR(E, kp.row_jres + r);
  const bool lidar = sa.do_lidar && sa.lidar_out && kp.n_rays;
  if (lidar) lidar_epilogue<CT, TE>(kp, t, n_valid);

  if (spec && taint && ok && q == 0) {  // keep the INPUT state of a tainted env (dyn_in is intact until our own store)
    const int idx = taint - 1;
    for (int r = 0; r < kp.dyn_rows; ++r) sa.u_buf[(size_t)idx * kp.dyn_rows + r] = sa.dyn_in[(size_t)r * sa.ld + env0 + col];
  }
  // ---- write back: one TMA bulk store per output row
  fence_proxy_async();
  __syncthreads();
  if (tid < 32) {
    for (int r = tid; r < kp.dyn_rows; r += 32) bulk_s2g(sa.dyn_out + (size_t)r * sa.ld + env0, t.envS + r * TE, row_bytes);
    if (sa.act_out)  // clamped action write-back (core.py:2019-2026, 2031-2038): the frc rows of the clamped agents
      for (int a = tid; a < kp.A; a += 32) {
        const int ao = T_ent<CT>(kp, T_agent<CT>(kp, a)).act_off;
        if (ao >= 0)
          for (int k = 0; k < 3; ++k)
            bulk_s2g(sa.act_out + (size_t)(ao + k) * sa.ld + env0, t.workS + (kp.row_frc + 3 * a + k) * TE, row_bytes);
      }
    if (lidar)
      for (int r = tid; r < kp.lidar_rows; r += 32) bulk_s2g(sa.lidar_out + (size_t)r * sa.ld + env0, t.lidS + r * TE, row_bytes);
    bulk_commit();
    bulk_wait_all();
  }
  if (!spec) return;
  // ---- last CTA: verify, repair if needed, publish the verdicts, reset the scratch for the next launch
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(sa.counters, 1) == sa.total_ctas - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  int bad = 0;
  uint32_t keep = 0u;  // the published word of this thread when one pass covers them all
  for (int w = tid; w < nw; w += blockDim.x) {
    const uint32_t th = __ldcg(sa.thin + w), ex = __ldcg(sa.any_exact + w), sp = __ldcg(sa.any_spec + w);
    bad |= (th & ~ex) != 0u;
    keep = ex | sp;
  }
  bad = __syncthreads_or(bad);
  int nd = 0;
  if (bad) {
    const int nu = *reinterpret_cast<volatile int*>(sa.counters + 1);
    nd = validate_speculation(kp, sa, flagS, nu);
    if (nd) repair_tainted<CT, TE>(kp, aio, sa, t, flagS, nd);
    for (int w = tid; w < nw; w += blockDim.x) sa.flags_out[w] = flagS[w];
  } else if (nw <= (int)blockDim.x) {
    if (tid < nw) sa.flags_out[tid] = keep;
  } else {
    for (int w = tid; w < nw; w += blockDim.x) sa.flags_out[w] = __ldcg(sa.any_exact + w) | __ldcg(sa.any_spec + w);
  }
  for (int w = tid; w < nw; w += blockDim.x) { sa.any_exact[w] = 0u; sa.any_spec[w] = 0u; sa.thin[w] = 0u; }
  if (tid == 0) {
    sa.counters[0] = 0; sa.counters[1] = 0;
    if (nd) {
      sa.counters[2] += 1;
      *sa.repairs_host = sa.counters[2];
      __threadfence_system();
    }
  }
}

// Verification tail of a world-compiled launch (spec_kernel.cuh) whose last CTA found a thin event on a pair no exact
// env activates (counters[3] raised): the same validation, repair, publication and scratch reset as the tail of
// step_kernel.  One CTA, launched behind every such launch; returns at once when there is nothing to do.  A chunked
// step launches one per chunk stream: the one that claims the flag does the work.
template <bool CT, int TE>
__global__ void __launch_bounds__(kThreads)
repair_kernel(const __grid_constant__ KParams kp, const __grid_constant__ AgentIO aio, const __grid_constant__ StepArgs sa) {
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_claim;
  const int tid = threadIdx.x;
  // programmatic dependent launch (vmasb200.cu): set up while the step's grid still runs, reads nothing before that grid has
  // completed; the next step's grid may be set up behind this one in turn (it waits for THIS grid, hence for both)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (tid == 0) s_claim = atomicExch(sa.counters + 3, 0);
  __syncthreads();
  if (!s_claim) return;
  __threadfence();
  const SmemMap m = smem_map(kp);
  Tile t;
  t.envS = smem + m.env; t.workS = smem + m.work; t.lidS = smem + m.lid;
  t.queue = reinterpret_cast<int*>(smem + m.queue); t.rayq = reinterpret_cast<int*>(smem + m.rayq);
  t.qcount = reinterpret_cast<int*>(smem + m.bar) + 2;
  uint32_t* flagS = reinterpret_cast<uint32_t*>(smem + m.flags);
  const int nw = kp.S * kp.pair_words;
  const int nu = *reinterpret_cast<volatile int*>(sa.counters + 1);
  const int nd = validate_speculation(kp, sa, flagS, nu);
  if (nd) repair_tainted<CT, TE>(kp, aio, sa, t, flagS, nd);
  for (int w = tid; w < nw; w += blockDim.x) sa.flags_out[w] = flagS[w];
  for (int w = tid; w < nw; w += blockDim.x) { sa.any_exact[w] = 0u; sa.any_spec[w] = 0u; sa.thin[w] = 0u; }
  if (tid == 0) {
    sa.counters[1] = 0;
    if (nd) {
      sa.counters[2] += 1;
      *sa.repairs_host = sa.counters[2];
      __threadfence_system();
    }
  }
}

// World.collides' batch-wide distance test (core.py:2796-2800) for substep s on the SoA buffers in HBM:
// flags[s][p] = any over envs of ||p_a - p_b|| <= Rc_a + Rc_b.  One thread per env (coalesced row reads), the verdicts
// of 32 pairs are OR-reduced over the warp with one REDUX.
__global__ void __launch_bounds__(256) flags_kernel(const __grid_constant__ KParams kp, const float* __restrict__ dyn,
                                                    const float* __restrict__ stat, long long n_envs, long long ld, int s,
                                                    uint32_t* __restrict__ flags) {
  const long long env = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = env < n_envs;
  const long long col = ok ? env : 0;
  auto px = [&](int row, int k) {
    return row < kp.dyn_rows ? dyn[(size_t)(row + k) * ld + col] : stat[(size_t)(row - kp.dyn_rows + k) * ld + col];
  };
  for (int base = 0; base < kp.P; base += 32) {
    const int n = min(32, kp.P - base);
    uint32_t bpw = 0u;
    for (int k = 0; k < n; ++k) {
      const DBroad* bq = kp.broad + base + k;
      const float dx = px(bq->pa_row, 0) - px(bq->pb_row, 0), dy = px(bq->pa_row, 1) - px(bq->pb_row, 1);
      bpw |= (fmaf(dy, dy, dx * dx) <= bq->t2_bp ? 1u : 0u) << k;
    }
    const uint32_t any = __reduce_or_sync(0xffffffffu, ok ? bpw : 0u);
    if (any && (threadIdx.x & 31) == 0) or_global(flags + s * kp.pair_words + (base >> 5), any);
  }
}

struct RayTargets {  // by-value kernel parameter: the (already filtered) target entities of one cast_rays call
  int n;
  int e[64];
};

// World.cast_rays (core.py:1661-1785) with caller-provided world-frame angles and target list.
// One thread per (env, ray) with the env index fastest, so that the SoA rows are read coalesced.
__global__ void __launch_bounds__(256) cast_rays_kernel(const __grid_constant__ KParams kp, const float* __restrict__ dyn,
                                                        const float* __restrict__ stat, long long ld, int entity,
                                                        const float* __restrict__ angles, long long angles_stride,
                                                        int n_rays, float max_range, const __grid_constant__ RayTargets rt,
                                                        float* __restrict__ out, long long n_envs) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= n_envs * n_rays) return;
  const int r = (int)(gid / n_envs);
  const long long env = gid - (long long)r * n_envs;
  auto field = [&](const DEnt* en, int k) {  // k: 0 px, 1 py, 2 rot
    return en->dyn_off >= 0 ? dyn[(size_t)(en->dyn_off + (k == 2 ? 4 : k)) * ld + env]
                            : stat[(size_t)(en->stat_off + k) * ld + env];
  };
  const DEnt* se = kp.ents + entity;
  const V2 o = mk(field(se, 0), field(se, 1));
  float ds, dc;
  sincosf(angles[env * angles_stride + r], &ds, &dc);
  float best = max_range;
  for (int t = 0; t < rt.n; ++t) {
    const DEnt* en = kp.ents + rt.e[t];
    const V2 p = mk(field(en, 0), field(en, 1));
    float d;
    if (en->shape == VMASB200_SPHERE) d = ray_sphere(p, en->radius, o, dc, ds, max_range);
    else {
      float s, c;
      sincosf(field(en, 2), &s, &c);
      if (en->shape == VMASB200_BOX) d = ray_box(p, c, s, en->length, en->width, o, dc, ds, max_range);
      else d = ray_line(p, c, s, en->length, o, dc, ds, max_range);
    }
    best = pmin(best, d);
  }
  out[env * n_rays + r] = best;
}

}  // namespace vb
