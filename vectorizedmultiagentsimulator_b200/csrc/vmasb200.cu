// vmasb200.cu - C ABI of libvmasb200.so (include/vmasb200.h): plan upload, launches, host-buffer entry point.
// Compiled with -fmad=false: every reference torch op is one rounded fp32 instruction (see geom.cuh).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <cuda.h>
#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: a range costs one pointer test unless a tool (nsys, ncu --nvtx) is attached  // types of cuLaunchKernelEx only: the driver entry points are resolved with dlsym (jit.h)

#include "kernels.cuh"
#include "queries.cuh"
#include "hooks.cuh"
#include "plan.h"
#include "jit.h"

namespace vb {

// One NVTX range per C-ABI call that launches work (the reference has no tracing hooks at all): `nsys` / `ncu --nvtx`
// show the step, its host-buffer pipeline, the queries, the action / reward / reset launches by name.
static bool nvtx_on() {  // VMASB200_NVTX=0 switches the ranges off
  static const bool on = [] { const char* e = getenv("VMASB200_NVTX"); return !(e && e[0] == '0'); }();
  return on;
}
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) : on(nvtx_on()) { if (on) nvtxRangePushA(name); }
  ~NvtxRange() { if (on) nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
const char* last_error();

#define VB_CUDA(call)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess) {                                                                   \
      set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                           \
      return VMASB200_ERR_CUDA;                                                                \
    }                                                                                          \
  } while (0)

template <class T>
static size_t put(std::vector<char>& blob, const std::vector<T>& v) {
  size_t off = (blob.size() + 15) & ~size_t(15);
  blob.resize(off + std::max<size_t>(v.size() * sizeof(T), 16));
  if (!v.empty()) std::memcpy(blob.data() + off, v.data(), v.size() * sizeof(T));
  return off;
}

// the step kernel instantiations: tables in the constant bank or in global memory x CTA tile (envs per CTA)
using StepFn = void (*)(const KParams, const AgentIO, const StepArgs);
static StepFn step_fn(bool ct, int te) {
  if (ct) return te == 32 ? step_kernel<true, 32> : te == 64 ? step_kernel<true, 64> : step_kernel<true, 128>;
  return te == 32 ? step_kernel<false, 32> : te == 64 ? step_kernel<false, 64> : step_kernel<false, 128>;
}

static size_t bits_words(const KParams& kp) { return (size_t)kp.S * std::max(kp.pair_words, 1); }

int upload_plan(vmasb200_plan* pl) {
  if (pl->uploaded) return VMASB200_OK;
  int ndev = 0;
  VB_CUDA(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) { set_error("no CUDA device: libvmasb200 has no CPU fallback"); return VMASB200_ERR_CUDA; }
  VB_CUDA(cudaGetDevice(&pl->device));
  std::vector<char> blob;
  size_t o_ent = put(blob, pl->ents), o_item = put(blob, pl->items), o_broad = put(blob, pl->broad);
  size_t o_dyn = put(blob, pl->dyn_list), o_trig = put(blob, pl->trig_list), o_ag = put(blob, pl->agent_entities);
  size_t o_ray = put(blob, pl->ray_angles), o_sen = put(blob, pl->sensors), o_tg = put(blob, pl->targets);
  size_t o_tc = put(blob, pl->tcull), o_cs = put(blob, pl->ray_cs);
  VB_CUDA(cudaMalloc(&pl->d_blob, blob.size()));
  VB_CUDA(cudaMemcpy(pl->d_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice));
  char* b = static_cast<char*>(pl->d_blob);
  KParams& kp = pl->kp;
  kp.ents = reinterpret_cast<const DEnt*>(b + o_ent);
  kp.items = reinterpret_cast<const DItem*>(b + o_item);
  kp.broad = reinterpret_cast<const DBroad*>(b + o_broad);
  kp.dyn_list = reinterpret_cast<const int32_t*>(b + o_dyn);
  kp.trig_list = reinterpret_cast<const int32_t*>(b + o_trig);
  kp.agent_ents = reinterpret_cast<const int32_t*>(b + o_ag);
  kp.ray_angles = reinterpret_cast<const float*>(b + o_ray);
  kp.sensors = reinterpret_cast<const DSensor*>(b + o_sen);
  kp.targets = reinterpret_cast<const int32_t*>(b + o_tg);
  kp.tcull = reinterpret_cast<const float4*>(b + o_tc);
  kp.ray_cs = reinterpret_cast<const float2*>(b + o_cs);
  // 4 bit arrays (flags, any_exact, any_spec, thin) + 4 counters
  const size_t nbits = (4 * bits_words(kp) + 4) * sizeof(uint32_t);
  VB_CUDA(cudaMalloc(&pl->d_bits, nbits));
  VB_CUDA(cudaMemset(pl->d_bits, 0, nbits));
  VB_CUDA(cudaHostAlloc(&pl->h_repairs, sizeof(int), cudaHostAllocMapped));
  *pl->h_repairs = 0;
  for (int ct = 0; ct < 2; ++ct)
    for (int te : {32, 64, 128})
      VB_CUDA(cudaFuncSetAttribute(step_fn(ct != 0, te), cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  VB_CUDA(cudaFuncSetAttribute(repair_kernel<false, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
  pl->uploaded = true;
  return VMASB200_OK;
}

static int check_io(const vmasb200_plan* pl, const vmasb200_step_io* io) {
  const KParams& kp = pl->kp;
  if (!io || io->n_envs < 0) { set_error("step: bad io"); return VMASB200_ERR_INVALID; }
  if (io->ld < io->n_envs || (io->ld & 3)) { set_error("step: ld must be >= n_envs and a multiple of 4"); return VMASB200_ERR_INVALID; }
  if (!io->dyn_in || !io->dyn_out) { set_error("step: dyn_in / dyn_out are required"); return VMASB200_ERR_INVALID; }
  const bool probe = io->substep_begin || io->substep_end || io->forces_are_substep_inputs || io->trig_in || io->jexp_in || io->force_out || io->joint_out;
  if (io->jexp_in && io->substep_end != io->substep_begin + 1) { set_error("step: jexp_in needs a one-substep range"); return VMASB200_ERR_INVALID; }
  if (probe && !io->exact) { set_error("step: the probe fields (substep range, trig_in, force_out, joint_out) need exact != 0"); return VMASB200_ERR_INVALID; }
  if ((io->substep_begin || io->substep_end) &&
      (io->substep_begin < 0 || io->substep_end > kp.S || io->substep_begin >= io->substep_end)) {
    set_error("step: substep range must satisfy 0 <= begin < end <= substeps"); return VMASB200_ERR_INVALID;
  }
  if (kp.stat_rows && !io->stat) { set_error("step: stat buffer missing"); return VMASB200_ERR_INVALID; }
  if (kp.act_rows && !io->act_out) { set_error("step: act_out missing"); return VMASB200_ERR_INVALID; }
  if (kp.A && (!io->force_ptrs || !io->torque_ptrs || !io->force_strides || !io->torque_strides)) {
    set_error("step: force / torque pointer tables missing"); return VMASB200_ERR_INVALID;
  }
  auto mis = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(io->dyn_in) || mis(io->dyn_out) || mis(io->stat) || mis(io->act_out) || mis(io->lidar_out)) {
    set_error("step: SoA buffers must be 16-byte aligned (TMA bulk copies)"); return VMASB200_ERR_INVALID;
  }
  return VMASB200_OK;
}

// The constant-bank copy of the walk tables (kernels.cuh CTab): one per device, owned by the plan that launched last.
struct ConstOwner {
  const vmasb200_plan* plan = nullptr;
  uint64_t serial = 0;
  cudaEvent_t last_use = nullptr;  // recorded after every (eager) launch that reads the tables
  bool graphs = false;             // a launch was captured into a CUDA graph on this device: see bind_const
};
static ConstOwner g_const[64];
static std::mutex g_const_mu;
static uint64_t g_plan_serial = 0;

static bool stream_capturing(cudaStream_t st) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); return false; }
  return cs == cudaStreamCaptureStatusActive;
}

// Makes c_tab hold this plan's tables for the launch that follows on `st`.
// Eager launches: the copy is skipped while the plan still owns the bank; another plan's upload is ordered after the
// previous owner's launches through `last_use`.  A launch that is being CAPTURED into a CUDA graph brings its own
// device-to-device upload node, so that every replay finds its tables whoever used the bank in between; the library
// cannot see replays, so from then on eager launches on this device re-upload every time (a 50 KB device-to-device
// copy), and replays / eager steps of different plans must be ordered by the caller's streams (they are when
// everything runs on one stream, as under torch).
static int bind_const(vmasb200_plan* pl, cudaStream_t st, bool* ct, bool* capturing) {
  const KParams& kp = pl->kp;
  *capturing = stream_capturing(st);
  *ct = (int)pl->broad.size() <= kCPairs && kp.E <= kCEnts && kp.n_sensors <= kCSensors && pl->device >= 0 && pl->device < 64;
  if (!*ct) return VMASB200_OK;
  std::lock_guard<std::mutex> lock(g_const_mu);
  ConstOwner& o = g_const[pl->device];
  if (pl->serial == 0) pl->serial = ++g_plan_serial;
  if (!pl->d_ctab) {
    if (*capturing) {
      set_error("step: the plan's first launch cannot be captured into a CUDA graph (run one step eagerly first)");
      return VMASB200_ERR_INVALID;
    }
    pl->h_ctab.reset(new CTab());
    CTab& t = *pl->h_ctab;
    std::memset(&t, 0, sizeof(t));
    std::copy(pl->broad.begin(), pl->broad.end(), t.broad);
    std::copy(pl->ents.begin(), pl->ents.end(), t.ents);
    std::copy(pl->dyn_list.begin(), pl->dyn_list.end(), t.dyn_list);
    std::copy(pl->trig_list.begin(), pl->trig_list.end(), t.trig_list);
    std::copy(pl->agent_entities.begin(), pl->agent_entities.end(), t.agent_ents);
    std::copy(pl->sensors.begin(), pl->sensors.end(), t.sensors);
    VB_CUDA(cudaMalloc(&pl->d_ctab, sizeof(CTab)));
    VB_CUDA(cudaMemcpy(pl->d_ctab, pl->h_ctab.get(), sizeof(CTab), cudaMemcpyHostToDevice));
  }
  if (*capturing) {
    VB_CUDA(cudaMemcpyToSymbolAsync(c_tab, pl->d_ctab, sizeof(CTab), 0, cudaMemcpyDeviceToDevice, st));
    o.graphs = true;
    o.plan = nullptr;
    return VMASB200_OK;
  }
  if (!o.last_use) VB_CUDA(cudaEventCreateWithFlags(&o.last_use, cudaEventDisableTiming));
  if (!o.graphs && o.plan == pl && o.serial == pl->serial) return VMASB200_OK;
  // another plan's launches may still be reading the tables: order the upload after them
  if (o.plan) VB_CUDA(cudaStreamWaitEvent(st, o.last_use, 0));
  VB_CUDA(cudaMemcpyToSymbolAsync(c_tab, pl->d_ctab, sizeof(CTab), 0, cudaMemcpyDeviceToDevice, st));
  VB_CUDA(cudaEventRecord(o.last_use, st));
  o.plan = pl; o.serial = pl->serial;
  return VMASB200_OK;
}

// CTA tile of a step over n environments (measured on B200, profiles/README.md): 4-member teams on 32 envs while that
// fills the GPU in about one wave - the shortest critical path; beyond, throughput counts: whole-env threads (128 envs
// per CTA, no work shared, fewest barriers) for small worlds - at most 32 candidate pairs, no LIDAR epilogue, two such
// CTAs fit an SM (transport, simple_spread: +40 %) - else 2-member teams on 64 envs (football: +35 % over 128).
static int tile_for(const vmasb200_plan* pl, long long n) {
  if (pl->te_fixed) return pl->te_fixed;
  auto smem = [&](int te) { KParams k = pl->kp; k.TE = te; return smem_map(k).total_bytes; };
  if (n <= 48 * 1024) return 32;
  if (!pl->kp.n_rays && pl->kp.n_blk <= 1 && smem(128) <= 113 * 1024) return 128;
  return smem(64) <= 226 * 1024 ? 64 : 32;
}

// Envs [env_begin, env_end) of one World.step on stream io->stream.  A step may be issued as several such launches
// (on different streams); together they must cover the whole batch: the CTA that finishes last - of ALL launches -
// verifies the speculation over the whole batch.  env_begin must be a multiple of the CTA tile.
static int launch_step(vmasb200_plan* pl, const vmasb200_step_io* io, long long env_begin, long long env_end) {
  KParams kp = pl->kp;
  kp.TE = tile_for(pl, io->n_envs);
  // every launch boundary of a chunked step is a multiple of the tile
  const long long total_ctas = (io->n_envs + kp.TE - 1) / kp.TE;
  const int smem_bytes = smem_map(kp).total_bytes;
  cudaStream_t st = static_cast<cudaStream_t>(io->stream);
  AgentIO aio;
  std::memset(&aio, 0, sizeof(aio));
  for (int a = 0; a < kp.A; ++a) {
    aio.force[a] = io->force_ptrs[a]; aio.torque[a] = io->torque_ptrs[a];
    aio.fstride[a] = (int32_t)io->force_strides[a]; aio.tstride[a] = (int32_t)io->torque_strides[a];
  }
  const int nw = kp.S * kp.pair_words;
  const size_t bw = bits_words(kp);
  uint32_t* flags = pl->d_bits;
  StepArgs sa{};
  sa.dyn_in = io->dyn_in; sa.dyn_out = io->dyn_out; sa.stat = io->stat; sa.act_out = io->act_out;
  sa.lidar_out = io->lidar_out; sa.n_envs = io->n_envs; sa.ld = io->ld;
  sa.env_begin = env_begin; sa.env_end = env_end; sa.total_ctas = (int)total_ctas;
  sa.flags_out = flags; sa.any_exact = flags + bw; sa.any_spec = flags + 2 * bw; sa.thin = flags + 3 * bw;
  sa.counters = reinterpret_cast<int*>(flags + 4 * bw);
  // world-compiled kernel (jit.h) when the plan has one: same scratch protocol, same bits, no constant-bank upload
  const int jte = jit_tile(pl, io->n_envs);
  JitKernel* jk = io->exact ? nullptr : jit_get(pl, jte, false, nullptr);
  const unsigned grid = (unsigned)((env_end - env_begin + kp.TE - 1) / kp.TE);
  const unsigned block = (unsigned)kp.NT;
  bool ct = false;
  bool capturing = false;
  int rc_ct = jk ? VMASB200_OK : bind_const(pl, st, &ct, &capturing);
  if (rc_ct) return rc_ct;
  if (io->exact) {
    // exact path: the batch-wide World.collides verdicts are reduced before every substep (core.py:2796-2800)
    if (nw) VB_CUDA(cudaMemsetAsync(flags, 0, nw * sizeof(uint32_t), st));
    const bool ranged = io->substep_begin || io->substep_end;
    const int s0 = ranged ? io->substep_begin : 0, s1 = ranged ? io->substep_end : kp.S;
    sa.trig_in = io->trig_in;  // only the first launch starts from the caller's state
    sa.jexp_in = io->jexp_in;
    const int probe_on = 1, probe_off = 0;
    if (io->jexp_in) VB_CUDA(cudaMemcpyToSymbolAsync(c_probe_jexp, &probe_on, sizeof(int), 0, cudaMemcpyHostToDevice, st));
    sa.forces_fresh = io->forces_are_substep_inputs;
    for (int s = s0; s < s1; ++s) {
      if (kp.P) {
        unsigned fg = (unsigned)((io->n_envs + 255) / 256);
        flags_kernel<<<fg, 256, 0, st>>>(kp, sa.dyn_in, sa.stat, io->n_envs, io->ld, s, flags);
      }
      sa.s_begin = s; sa.s_end = s + 1; sa.flags_in = flags;
      sa.do_lidar = (s == kp.S - 1);
      sa.dbg_force = s == s1 - 1 ? io->force_out : nullptr;
      sa.dbg_joint = s == s1 - 1 ? io->joint_out : nullptr;
      step_fn(ct, kp.TE)<<<grid, block, smem_bytes, st>>>(kp, aio, sa);
      sa.dyn_in = io->dyn_out;
      sa.trig_in = nullptr;
    }
    if (io->jexp_in) VB_CUDA(cudaMemcpyToSymbolAsync(c_probe_jexp, &probe_off, sizeof(int), 0, cudaMemcpyHostToDevice, st));
  } else {
    // fused speculative launch; the last CTA verifies it and repairs the (rare) tainted envs in place
    if (io->n_envs > pl->u_cap) {
      cudaFree(pl->d_ulist); cudaFree(pl->d_ubuf); cudaFree(pl->d_urec);
      pl->d_ulist = nullptr; pl->d_ubuf = nullptr; pl->d_urec = nullptr; pl->u_cap = 0;
      const size_t rec = (size_t)std::min<int64_t>(io->n_envs, kRecCap) * bw;  // words of one record array
      VB_CUDA(cudaMalloc(&pl->d_ulist, 2 * (size_t)io->n_envs * sizeof(int)));
      VB_CUDA(cudaMalloc(&pl->d_ubuf, (size_t)io->n_envs * std::max(kp.dyn_rows, 1) * sizeof(float)));
      VB_CUDA(cudaMalloc(&pl->d_urec, 2 * rec * sizeof(uint32_t)));
      pl->u_cap = io->n_envs;
    }
    sa.u_list = pl->d_ulist; sa.u_rep = pl->d_ulist + pl->u_cap; sa.u_buf = pl->d_ubuf;
    sa.rec_cap = (int)std::min<int64_t>(pl->u_cap, kRecCap);
    sa.u_verd = pl->d_urec; sa.u_thin = pl->d_urec + (size_t)sa.rec_cap * bw;
    VB_CUDA(cudaHostGetDevicePointer(&sa.repairs_host, pl->h_repairs, 0));
    sa.s_begin = 0; sa.s_end = kp.S; sa.flags_in = nullptr; sa.do_lidar = 1;
    if (jk) {
      if (env_begin % jte) { set_error("step: chunk boundary is not a multiple of the compiled kernel's tile"); return VMASB200_ERR_INVALID; }
      sa.total_ctas = (int)((io->n_envs + jte - 1) / jte);
      KParams jkp = jk->kp;
      void* args[3] = {&jkp, &aio, &sa};
      const unsigned jgrid = (unsigned)((env_end - env_begin + jte - 1) / jte);
      // Programmatic dependent launch: the kernel starts with griddepcontrol.wait + launch_dependents (spec_kernel.cuh), so a
      // step launched right behind another one on the stream is set up on the SMs while its predecessor drains and only its
      // first instruction waits for that grid to complete (with its memory visible).  Any other predecessor (a torch kernel,
      // a copy) simply completes first, as without the attribute.  Not inside a stream capture.
      static const bool pdl = [] { const char* e = getenv("VMASB200_PDL"); return !(e && e[0] == '0'); }();
      int rc;
      if (pdl && !capturing && g_drv.LaunchKernelEx) {
        CUlaunchAttribute attr{};
        attr.id = CU_LAUNCH_ATTRIBUTE_PROGRAMMATIC_STREAM_SERIALIZATION;
        attr.value.programmaticStreamSerializationAllowed = 1;
        CUlaunchConfig cfg{};
        cfg.gridDimX = jgrid; cfg.gridDimY = 1; cfg.gridDimZ = 1;
        cfg.blockDimX = (unsigned)jk->nt; cfg.blockDimY = 1; cfg.blockDimZ = 1;
        cfg.sharedMemBytes = (unsigned)jk->smem_bytes;
        cfg.hStream = reinterpret_cast<CUstream>(st);
        cfg.attrs = &attr; cfg.numAttrs = 1;
        rc = g_drv.LaunchKernelEx(&cfg, jk->func, args, nullptr);
      } else {
        rc = g_drv.LaunchKernel(jk->func, jgrid, 1, 1, (unsigned)jk->nt, 1, 1, (unsigned)jk->smem_bytes, st, args, nullptr);
      }
      if (rc) { set_error("cuLaunchKernel(vmasb200_step_world) failed: CUresult " + std::to_string(rc)); return VMASB200_ERR_CUDA; }
      pl->jit_launches += 1;
      if (jk->thin_possible) {  // verification tail of the compiled launch (returns at once unless its last CTA asked)
        KParams gk = jk->kpg;
        if (pdl && !capturing) {  // same linkage as the step's own launch: 6 us per step of launch gaps otherwise (transport)
          cudaLaunchAttribute ra{};
          ra.id = cudaLaunchAttributeProgrammaticStreamSerialization;
          ra.val.programmaticStreamSerializationAllowed = 1;
          cudaLaunchConfig_t rcfg{};
          rcfg.gridDim = dim3(1); rcfg.blockDim = dim3(kThreads);
          rcfg.dynamicSmemBytes = (size_t)smem_map(gk).total_bytes; rcfg.stream = st;
          rcfg.attrs = &ra; rcfg.numAttrs = 1;
          VB_CUDA(cudaLaunchKernelEx(&rcfg, repair_kernel<false, 32>, gk, aio, sa));
        } else {
          repair_kernel<false, 32><<<1, kThreads, smem_map(gk).total_bytes, st>>>(gk, aio, sa);
        }
        pl->jit_tail_launches += 1;
      }
    } else {
      step_fn(ct, kp.TE)<<<grid, block, smem_bytes, st>>>(kp, aio, sa);
    }
  }
  if (ct && !capturing) VB_CUDA(cudaEventRecord(g_const[pl->device].last_use, st));
  if (io->pair_active_out && nw) VB_CUDA(cudaMemcpyAsync(io->pair_active_out, flags, nw * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

}  // namespace vb

using namespace vb;

extern "C" {

int vmasb200_version(void) { return VMASB200_ABI_VERSION; }
const char* vmasb200_last_error(void) { return vb::last_error(); }

int vmasb200_plan_create(const vmasb200_world_desc* desc, vmasb200_plan** out) {
  NvtxRange nvtx_("vmasb200_plan_create");
  if (!out) { set_error("plan_create: out is NULL"); return VMASB200_ERR_INVALID; }
  *out = nullptr;
  vmasb200_plan* pl = new (std::nothrow) vmasb200_plan();
  if (!pl) { set_error("out of host memory"); return VMASB200_ERR_NOMEM; }
  int rc = build_plan(desc, pl);
  if (rc != VMASB200_OK) { delete pl; return rc; }
  *out = pl;
  return VMASB200_OK;
}

int vmasb200_plan_destroy(vmasb200_plan* pl) {
  if (!pl) return VMASB200_OK;
  jit_forget(pl);
  {
    std::lock_guard<std::mutex> lock(g_const_mu);
    for (ConstOwner& o : g_const) if (o.plan == pl) o.plan = nullptr;
  }
  if (pl->uploaded) {
    cudaFree(pl->d_blob); cudaFree(pl->d_ctab); cudaFree(pl->d_bits); cudaFreeHost(pl->h_repairs); cudaFree(pl->d_ulist); cudaFree(pl->d_ubuf); cudaFree(pl->d_urec);
    cudaFree(pl->d_dyn); cudaFree(pl->d_stat); cudaFree(pl->d_force); cudaFree(pl->d_torque); cudaFree(pl->d_act); cudaFree(pl->d_lidar);
    for (cudaStream_t s : pl->streams) if (s) cudaStreamDestroy(s);
  }
  delete pl;
  return VMASB200_OK;
}

int vmasb200_plan_layout(const vmasb200_plan* pl, vmasb200_layout* out) {
  if (!pl || !out) { set_error("plan_layout: NULL argument"); return VMASB200_ERR_INVALID; }
  *out = pl->layout;
  return VMASB200_OK;
}

int vmasb200_plan_slots(const vmasb200_plan* pl, int32_t* dyn_slot, int32_t* stat_slot, int32_t* act_slot) {
  if (!pl) { set_error("plan_slots: NULL plan"); return VMASB200_ERR_INVALID; }
  const size_t n = pl->dyn_slot.size() * sizeof(int32_t);
  if (dyn_slot) std::memcpy(dyn_slot, pl->dyn_slot.data(), n);
  if (stat_slot) std::memcpy(stat_slot, pl->stat_slot.data(), n);
  if (act_slot) std::memcpy(act_slot, pl->act_slot.data(), n);
  return VMASB200_OK;
}

int vmasb200_plan_pairs(const vmasb200_plan* pl, vmasb200_pair* pairs) {
  if (!pl || (!pairs && !pl->pairs.empty())) { set_error("plan_pairs: NULL argument"); return VMASB200_ERR_INVALID; }
  if (!pl->pairs.empty()) std::memcpy(pairs, pl->pairs.data(), pl->pairs.size() * sizeof(vmasb200_pair));
  return VMASB200_OK;
}

int vmasb200_plan_joint_slots(const vmasb200_plan* pl, int32_t* joint_slot) {
  if (!pl) { set_error("plan_joint_slots: NULL plan"); return VMASB200_ERR_INVALID; }
  if (joint_slot && !pl->joint_slot.empty()) std::memcpy(joint_slot, pl->joint_slot.data(), pl->joint_slot.size() * sizeof(int32_t));
  return VMASB200_OK;
}

int vmasb200_plan_lidar_slots(const vmasb200_plan* pl, int32_t* lidar_slot) {
  if (!pl) { set_error("plan_lidar_slots: NULL plan"); return VMASB200_ERR_INVALID; }
  if (lidar_slot && !pl->lidar_slot.empty()) std::memcpy(lidar_slot, pl->lidar_slot.data(), pl->lidar_slot.size() * sizeof(int32_t));
  return VMASB200_OK;
}

int vmasb200_plan_trig_entities(const vmasb200_plan* pl, int32_t* trig_entities) {
  if (!pl) { set_error("plan_trig_entities: NULL plan"); return VMASB200_ERR_INVALID; }
  if (trig_entities && !pl->trig_list.empty()) std::memcpy(trig_entities, pl->trig_list.data(), pl->trig_list.size() * sizeof(int32_t));
  return (int)pl->trig_list.size();
}

int vmasb200_plan_tile(const vmasb200_plan* pl, int64_t n_envs) {
  if (!pl) return 0;
  {
    std::lock_guard<std::mutex> lock(g_jit_mu);
    auto it = g_jit_plans.find(pl);
    if (jit_on() && it != g_jit_plans.end()) {
      auto k = it->second.by_te.find(jit_tile(pl, n_envs));
      if (k != it->second.by_te.end() && k->second.func) return k->first;
    }
  }
  return tile_for(pl, n_envs);
}

int vmasb200_jit_enable(int on) {
  std::lock_guard<std::mutex> lock(g_jit_mu);
  const int prev = jit_mode();
  g_jit_on = on < 0 ? 0 : on > 2 ? 2 : on;
  return prev;
}

int vmasb200_plan_jit(vmasb200_plan* pl, int64_t n_envs, int32_t* state, int64_t* launches, int64_t* tail_launches) {
  NvtxRange nvtx_("vmasb200_plan_jit");
  if (!pl) { set_error("plan_jit: NULL plan"); return VMASB200_ERR_INVALID; }
  std::string why;
  int st = 0;
  if (!jit_on() || !jit_eligible(pl, jit_tile(pl, n_envs))) st = 0;
  else {
    jit_get(pl, jit_tile(pl, n_envs), true, &why);
    st = why == "compiled" ? 1 : -1;
    if (st < 0) set_error("plan_jit: " + why);
  }
  if (state) *state = st;
  if (launches) *launches = pl->jit_launches;
  if (tail_launches) *tail_launches = pl->jit_tail_launches;
  return VMASB200_OK;
}

int vmasb200_step(vmasb200_plan* pl, const vmasb200_step_io* io) {
  NvtxRange nvtx_("vmasb200_step");
  if (!pl) { set_error("step: NULL plan"); return VMASB200_ERR_INVALID; }
  int rc = check_io(pl, io);
  if (rc) return rc;
  if ((rc = upload_plan(pl))) return rc;
  if (io->n_envs == 0 || pl->kp.n_dyn == 0) return VMASB200_OK;
  return launch_step(pl, io, 0, io->n_envs);
}

static int grow_host_staging(vmasb200_plan* pl, int64_t ld) {
  if (ld <= pl->host_cap) return VMASB200_OK;
  pl->res_n = -1; pl->res_stat = false;
  const KParams& kp = pl->kp;
  cudaFree(pl->d_dyn); cudaFree(pl->d_stat); cudaFree(pl->d_force); cudaFree(pl->d_torque); cudaFree(pl->d_act); cudaFree(pl->d_lidar);
  pl->d_dyn = pl->d_stat = pl->d_force = pl->d_torque = pl->d_act = pl->d_lidar = nullptr;
  pl->host_cap = 0;
  auto bytes = [&](int rows) { return std::max<size_t>((size_t)ld * rows * sizeof(float), 16); };
  VB_CUDA(cudaMalloc(&pl->d_dyn, bytes(kp.dyn_rows)));
  VB_CUDA(cudaMalloc(&pl->d_stat, bytes(kp.stat_rows)));
  VB_CUDA(cudaMalloc(&pl->d_force, bytes(2 * kp.A)));
  VB_CUDA(cudaMalloc(&pl->d_torque, bytes(kp.A)));
  VB_CUDA(cudaMalloc(&pl->d_act, bytes(kp.act_rows)));
  VB_CUDA(cudaMalloc(&pl->d_lidar, bytes(kp.lidar_rows)));
  pl->host_cap = ld;
  return VMASB200_OK;
}

// device alias of a pinned (mapped) host buffer; false when `h` is pageable (NULL stays NULL)
static bool host_dev_ptr(const void* h, void** d) {
  *d = nullptr;
  if (!h) return true;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, h) != cudaSuccess) { cudaGetLastError(); return false; }
  if (at.type != cudaMemoryTypeHost || !at.devicePointer) return false;
  *d = at.devicePointer;
  return true;
}

int vmasb200_step_host(vmasb200_plan* pl, const float* h_dyn_in, const float* h_stat, const float* h_force,
                       const float* h_torque, float* h_dyn_out, float* h_act_out, float* h_lidar_out,
                       uint32_t* h_pair_active, int64_t n, int64_t ld, int32_t exact) {
  NvtxRange nvtx_("vmasb200_step_host");
  if (!pl || n < 0 || ld < n || (ld & 3)) { set_error("step_host: bad argument (ld >= n_envs, ld % 4 == 0)"); return VMASB200_ERR_INVALID; }
  int rc = upload_plan(pl);
  if (rc) return rc;
  const KParams& kp = pl->kp;
  if (n == 0) return VMASB200_OK;
  if ((rc = grow_host_staging(pl, ld))) return rc;
  // h_dyn_in == NULL / h_stat == NULL: the rows this plan's device staging holds from the previous call (a rollout through
  // the host API only moves the actions in: the state it would send is the state it just received)
  if (!h_dyn_in && pl->res_n != n) { set_error("step_host: h_dyn_in is NULL but the plan holds no state of this batch size from a previous call"); return VMASB200_ERR_INVALID; }
  if (kp.stat_rows && !h_stat && !(pl->res_stat && pl->res_n == n)) { set_error("step_host: stat buffer missing (NULL = the one uploaded by the previous call of this batch size)"); return VMASB200_ERR_INVALID; }
  if (h_dyn_in && pl->res_n != n) pl->res_stat = false;
  for (cudaStream_t& s : pl->streams)
    if (!s) VB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  const int64_t dld = pl->host_cap;  // row stride of the device staging (>= ld)

  // chunks of envs, one stream each: H2D(k+1) overlaps kernel(k) overlaps D2H(k-1).  Boundaries are multiples of
  // 256 envs (every CTA tile size divides 256); the exact path needs batch-wide flags between substeps: one chunk.
  // chunk size: 16384 envs measured best at 32 K - 256 K envs (profiles/README.md); VMASB200_HOST_CHUNK_ENVS overrides
  static const int64_t chunk_envs = [] { const char* e = std::getenv("VMASB200_HOST_CHUNK_ENVS"); return e ? std::max<int64_t>(256, std::atoll(e)) : int64_t(16384); }();
  int64_t nchunk = exact ? 1 : std::min<int64_t>(kHostStreams, (n + chunk_envs - 1) / chunk_envs);
  int64_t csize = ((n + nchunk - 1) / nchunk + 255) & ~int64_t(255);
  nchunk = (n + csize - 1) / csize;
  const float* fptr[kMaxAgents]; const float* tptr[kMaxAgents];
  int64_t fstr[kMaxAgents], tstr[kMaxAgents];
  for (int a = 0; a < kp.A; ++a) {  // h_force [n][A][2], h_torque [n][A]
    fptr[a] = h_force ? pl->d_force + 2 * a : nullptr; fstr[a] = 2 * kp.A;
    tptr[a] = h_torque ? pl->d_torque + a : nullptr; tstr[a] = kp.A;
  }
  const bool want_lidar = kp.lidar_rows && h_lidar_out;
  auto rows2d = [&](void* dst, size_t dpitch, const void* src, size_t spitch, int64_t c0, int64_t ncol, int rows,
                    cudaMemcpyKind kind, cudaStream_t st) {
    return cudaMemcpy2DAsync(static_cast<char*>(dst) + c0 * 4, dpitch * 4, static_cast<const char*>(src) + c0 * 4,
                             spitch * 4, (size_t)ncol * 4, (size_t)rows, kind, st);
  };
  auto d2h = [&](int64_t c0, int64_t ncol, cudaStream_t st) -> int {
    if (h_dyn_out && kp.dyn_rows) VB_CUDA(rows2d(h_dyn_out, ld, pl->d_dyn, dld, c0, ncol, kp.dyn_rows, cudaMemcpyDeviceToHost, st));
    if (h_act_out && kp.act_rows) VB_CUDA(rows2d(h_act_out, ld, pl->d_act, dld, c0, ncol, kp.act_rows, cudaMemcpyDeviceToHost, st));
    if (want_lidar) VB_CUDA(rows2d(h_lidar_out, ld, pl->d_lidar, dld, c0, ncol, kp.lidar_rows, cudaMemcpyDeviceToHost, st));
    return VMASB200_OK;
  };
  // Zero-copy path: when every buffer is pinned (page-locked, mapped) host memory, ONE launch over the whole batch
  // reads its tiles from and writes its results to host memory directly - the TMA bulk copies and the force loads go
  // over PCIe themselves, in both directions at once and overlapped with the arithmetic of the other CTAs - instead of
  // H2D copies, kernels and D2H copies chained per chunk.  Opt-in (VMASB200_HOST_DIRECT=1): measured on B200 it is
  // +5 % (navigation) / +17 % (transport) at 32 768 envs, equal at 262 144 and -8 % for football (CTAs that wait
  // for PCIe hold the SM's shared memory), see profiles/README.md; same bits either way (tests/test_gpu_parity.py).
  const char* env_direct = std::getenv("VMASB200_HOST_DIRECT");
  const bool direct_ok = env_direct && std::atoi(env_direct) != 0;
  if (direct_ok && !exact && h_dyn_in && (h_stat || !kp.stat_rows) && h_dyn_out && kp.n_dyn && (!kp.act_rows || h_act_out)) {
    void *d_in, *d_out, *d_st, *d_f, *d_t, *d_act, *d_lid;
    if (host_dev_ptr(h_dyn_in, &d_in) && host_dev_ptr(h_dyn_out, &d_out) && host_dev_ptr(h_stat, &d_st) &&
        host_dev_ptr(h_force, &d_f) && host_dev_ptr(h_torque, &d_t) && host_dev_ptr(h_act_out, &d_act) &&
        host_dev_ptr(want_lidar ? h_lidar_out : nullptr, &d_lid)) {
      const float* f2[kMaxAgents]; const float* t2[kMaxAgents];
      for (int a = 0; a < kp.A; ++a) {
        f2[a] = d_f ? static_cast<const float*>(d_f) + 2 * a : nullptr;
        t2[a] = d_t ? static_cast<const float*>(d_t) + a : nullptr;
      }
      cudaStream_t st = pl->streams[0];
      vmasb200_step_io io{};
      io.dyn_in = static_cast<const float*>(d_in); io.dyn_out = static_cast<float*>(d_out);
      io.stat = static_cast<const float*>(d_st);
      io.force_ptrs = f2; io.force_strides = fstr; io.torque_ptrs = t2; io.torque_strides = tstr;
      io.act_out = kp.act_rows ? static_cast<float*>(d_act) : nullptr;
      io.lidar_out = static_cast<float*>(d_lid);
      io.pair_active_out = nullptr; io.n_envs = n; io.ld = ld; io.exact = 0; io.stream = st;
      if ((rc = check_io(pl, &io)) == VMASB200_OK) {
        if ((rc = launch_step(pl, &io, 0, n))) return rc;
        VB_CUDA(cudaStreamSynchronize(st));
        pl->res_n = -1;  // the state went host -> host: the staging holds nothing
        if (h_pair_active && kp.S * kp.pair_words)
          VB_CUDA(cudaMemcpy(h_pair_active, pl->d_bits, (size_t)kp.S * kp.pair_words * 4, cudaMemcpyDeviceToHost));
        return VMASB200_OK;
      }
    }
  }
  // VMASB200_HOST_TRACE=1: where the host time of a call goes (enqueue of copies + launches vs waiting), every 100 calls
  static const bool trace = std::getenv("VMASB200_HOST_TRACE") != nullptr;
  static double tr_enq = 0, tr_wait = 0;
  static long tr_n = 0;
  const auto tr0 = std::chrono::steady_clock::now();
  // Actions straight from PINNED host memory (batches up to 131 072 envs; VMASB200_HOST_ACTIONS_DIRECT=0 switches it off):
  // the kernel's force loads go over PCIe themselves instead of an H2D copy in front of the launch, and with the state
  // resident the batch is ONE chunk - the per-step chain is launch -> D2H instead of two chunks of H2D -> launch -> D2H.
  // Measured at 32 768 envs on one box: navigation 0.251 -> 0.229 ms per step, transport 0.170 -> 0.140, simple_spread
  // 0.168 -> 0.099, football 0.274 -> 0.235.  Pageable buffers keep the copies.
  static const bool act_direct = [] { const char* e = std::getenv("VMASB200_HOST_ACTIONS_DIRECT"); return !(e && std::atoi(e) == 0); }();
  bool acts_mapped = false;
  if (act_direct && !exact && n <= 131072 && (h_force || h_torque)) {
    void *d_f = nullptr, *d_t = nullptr;
    if (host_dev_ptr(h_force, &d_f) && host_dev_ptr(h_torque, &d_t)) {
      acts_mapped = true;
      if (!h_dyn_in) { nchunk = 1; csize = (n + 255) & ~int64_t(255); }
      for (int a = 0; a < kp.A; ++a) {
        fptr[a] = d_f ? static_cast<const float*>(d_f) + 2 * a : nullptr;
        tptr[a] = d_t ? static_cast<const float*>(d_t) + a : nullptr;
      }
    }
  }
  // repairs so far (mapped host word the verifying CTA writes): every earlier step of this plan has been waited for
  const int repairs_before = *static_cast<volatile int*>(pl->h_repairs);
  for (int64_t k = 0; k < nchunk; ++k) {
    cudaStream_t st = pl->streams[k % kHostStreams];
    const int64_t c0 = k * csize, c1 = std::min(n, c0 + csize), ncol = c1 - c0;
    if (kp.dyn_rows && h_dyn_in) VB_CUDA(rows2d(pl->d_dyn, dld, h_dyn_in, ld, c0, ncol, kp.dyn_rows, cudaMemcpyHostToDevice, st));
    if (kp.stat_rows && h_stat) VB_CUDA(rows2d(pl->d_stat, dld, h_stat, ld, c0, ncol, kp.stat_rows, cudaMemcpyHostToDevice, st));
    if (kp.A && h_force && !acts_mapped) VB_CUDA(cudaMemcpyAsync(pl->d_force + c0 * 2 * kp.A, h_force + c0 * 2 * kp.A, (size_t)ncol * kp.A * 8, cudaMemcpyHostToDevice, st));
    if (kp.A && h_torque && !acts_mapped) VB_CUDA(cudaMemcpyAsync(pl->d_torque + c0 * kp.A, h_torque + c0 * kp.A, (size_t)ncol * kp.A * 4, cudaMemcpyHostToDevice, st));
    vmasb200_step_io io{};
    io.dyn_in = pl->d_dyn; io.dyn_out = pl->d_dyn; io.stat = pl->d_stat;
    io.force_ptrs = fptr; io.force_strides = fstr; io.torque_ptrs = tptr; io.torque_strides = tstr;
    io.act_out = kp.act_rows ? pl->d_act : nullptr;
    io.lidar_out = want_lidar ? pl->d_lidar : nullptr;
    io.pair_active_out = nullptr; io.n_envs = n; io.ld = dld; io.exact = exact; io.stream = st;
    if (kp.n_dyn && (rc = launch_step(pl, &io, c0, c1))) return rc;
    if ((rc = d2h(c0, ncol, st))) return rc;
  }
  const auto tr1 = std::chrono::steady_clock::now();
  for (int64_t k = 0; k < std::min<int64_t>(nchunk, kHostStreams); ++k) VB_CUDA(cudaStreamSynchronize(pl->streams[k]));
  if (trace) {
    const auto tr2 = std::chrono::steady_clock::now();
    tr_enq += std::chrono::duration<double, std::micro>(tr1 - tr0).count();
    tr_wait += std::chrono::duration<double, std::micro>(tr2 - tr1).count();
    if (++tr_n % 100 == 0) {
      fprintf(stderr, "libvmasb200 step_host: %ld calls, %lld chunks: enqueue %.1f us, wait %.1f us per call\n", tr_n, (long long)nchunk, tr_enq / 100, tr_wait / 100);
      tr_enq = tr_wait = 0;
    }
  }
  if (nchunk > 1) {
    // a repair by the step's last CTA may have rewritten envs of a chunk whose results were already copied out
    if (*static_cast<volatile int*>(pl->h_repairs) != repairs_before) {
      if ((rc = d2h(0, n, pl->streams[0]))) return rc;
      VB_CUDA(cudaStreamSynchronize(pl->streams[0]));
    }
  }
  if (h_pair_active && kp.S * kp.pair_words)
    VB_CUDA(cudaMemcpy(h_pair_active, pl->d_bits, (size_t)kp.S * kp.pair_words * 4, cudaMemcpyDeviceToHost));
  pl->res_n = n;  // the kernel stepped the staging in place: it now holds the new state
  if (h_stat) pl->res_stat = true;
  return VMASB200_OK;
}

int vmasb200_cast_rays(vmasb200_plan* pl, const float* dyn, const float* stat, int64_t ld, int32_t entity,
                       const float* angles, int64_t angles_stride, int32_t n_rays, float max_range,
                       const int32_t* targets, int32_t n_targets, float* out, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_cast_rays");
  if (!pl || !dyn || !angles || !out || n < 0 || ld < n || n_rays < 0 || n_targets < 0 || (n_targets && !targets)) {
    set_error("cast_rays: bad argument"); return VMASB200_ERR_INVALID;
  }
  int rc = upload_plan(pl);
  if (rc) return rc;
  const KParams& kp = pl->kp;
  if (entity < 0 || entity >= kp.E) { set_error("cast_rays: entity out of range"); return VMASB200_ERR_INVALID; }
  auto has_rec = [&](int e) { return pl->ents[e].dyn_off >= 0 || pl->ents[e].stat_off >= 0; };
  if (!has_rec(entity)) { set_error("cast_rays: the casting entity has no rows in this plan"); return VMASB200_ERR_INVALID; }
  if (n_targets > 64) { set_error("cast_rays: more than 64 targets"); return VMASB200_ERR_UNSUPPORTED; }
  for (int t = 0; t < n_targets; ++t)
    if (targets[t] < 0 || targets[t] >= kp.E || !has_rec(targets[t])) {
      set_error("cast_rays: a target has no rows in this plan"); return VMASB200_ERR_INVALID;
    }
  if (n == 0 || n_rays == 0) return VMASB200_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  RayTargets rt;
  rt.n = n_targets;
  for (int t = 0; t < n_targets; ++t) rt.e[t] = targets[t];
  const long long total = (long long)n * n_rays;
  cast_rays_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(kp, dyn, stat ? stat : dyn, ld, entity, angles,
                                                                   angles_stride, n_rays, max_range, rt, out, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

static int to_qpose(const vmasb200_pose* p, QPose* q, const char* who) {
  if (!p || !p->pos || (p->shape != VMASB200_SPHERE && p->shape != VMASB200_BOX && p->shape != VMASB200_LINE) ||
      (p->shape != VMASB200_SPHERE && !p->rot)) {
    set_error(std::string(who) + ": bad pose (shape, pos, rot of a box / line)");
    return VMASB200_ERR_INVALID;
  }
  q->shape = p->shape;
  q->radius = (float)p->radius; q->length = (float)p->length; q->width = (float)p->width;
  q->overlap_min = (float)(p->radius + 4.0 / 6e2);
  q->pos = p->pos; q->ps = p->pos_stride; q->pxy = p->pos_stride_xy;
  q->rot = p->rot; q->rs = p->rot_stride;
  return VMASB200_OK;
}

int vmasb200_get_distance(const vmasb200_pose* a, const vmasb200_pose* b, float* out, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_get_distance");
  QPose qa, qb;
  int rc;
  if ((rc = to_qpose(a, &qa, "get_distance")) || (rc = to_qpose(b, &qb, "get_distance"))) return rc;
  if (!out || n < 0) { set_error("get_distance: bad argument"); return VMASB200_ERR_INVALID; }
  if (n == 0) return VMASB200_OK;
  pair_query_kernel<QM_DISTANCE><<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(qa, qb, out, nullptr, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_is_overlapping(const vmasb200_pose* a, const vmasb200_pose* b, uint8_t* out, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_is_overlapping");
  QPose qa, qb;
  int rc;
  if ((rc = to_qpose(a, &qa, "is_overlapping")) || (rc = to_qpose(b, &qb, "is_overlapping"))) return rc;
  if (!out || n < 0) { set_error("is_overlapping: bad argument"); return VMASB200_ERR_INVALID; }
  if (n == 0) return VMASB200_OK;
  pair_query_kernel<QM_OVERLAP><<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(qa, qb, nullptr, out, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_get_distance_from_point(const vmasb200_pose* a, const float* points, int64_t points_stride,
                                     int64_t points_stride_xy, float* out, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_get_distance_from_point");
  QPose qa;
  int rc;
  if ((rc = to_qpose(a, &qa, "get_distance_from_point"))) return rc;
  if (!points || !out || n < 0) { set_error("get_distance_from_point: bad argument"); return VMASB200_ERR_INVALID; }
  if (n == 0) return VMASB200_OK;
  point_query_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(qa, points, points_stride, points_stride_xy, out, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_set_action(const float* action, int64_t stride_env, int64_t stride_k, int32_t action_size,
                        const float* u_range, const float* u_multiplier, int32_t clamp, float* u_out, int32_t* flags,
                        int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_set_action");
  if (!action || !u_range || !u_multiplier || !u_out || !flags || n < 0 || action_size < 1 || action_size > 8) {
    set_error("set_action: bad argument (1 <= action_size <= 8)"); return VMASB200_ERR_INVALID;
  }
  if (n == 0) return VMASB200_OK;
  ActParams ap{};
  for (int k = 0; k < action_size; ++k) { ap.range[k] = u_range[k]; ap.mult[k] = u_multiplier[k]; }
  const long long total = (long long)n * action_size;
  set_action_kernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      action, stride_env, stride_k, action_size, ap, clamp, u_out, flags, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_reset_rows(float* buf, int32_t row0, int32_t n_rows, int64_t ld, int64_t n, int64_t env_index, const uint8_t* mask,
                        void* stream) {
  NvtxRange nvtx_("vmasb200_reset_rows");
  if (!buf || row0 < 0 || n_rows < 0 || n < 0 || ld < n || env_index >= n) { set_error("reset_rows: bad argument"); return VMASB200_ERR_INVALID; }
  if (n_rows == 0 || n == 0) return VMASB200_OK;
  const long long total = env_index >= 0 ? n_rows : (long long)n_rows * n;
  reset_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      buf + (size_t)row0 * ld, n_rows, ld, n, env_index, mask);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_spawn_select(float* pos, const float* proposed, const float* occupied, int32_t n_occupied, float min_dist,
                          int32_t* flag, int64_t batch, void* stream) {
  NvtxRange nvtx_("vmasb200_spawn_select");
  if (!pos || !proposed || !flag || batch < 0 || n_occupied < 0 || (n_occupied && !occupied)) {
    set_error("spawn_select: bad argument"); return VMASB200_ERR_INVALID;
  }
  if (batch == 0) return VMASB200_OK;
  spawn_select_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      pos, proposed, occupied, n_occupied, min_dist, flag, batch);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_set_action_ex(const vmasb200_action* a, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_set_action_ex");
  if (!a || !a->action || !a->u_range || !a->u_multiplier || !a->u_out || !a->flags || n < 0 || a->action_size < 1 ||
      a->action_size > 8 || a->n_comm < 0 || a->n_comm > 64 || a->comm_col0 < 0 || (a->n_comm && !a->c_out)) {
    set_error("set_action_ex: bad argument (1 <= action_size <= 8, 0 <= n_comm <= 64, c_out when n_comm > 0)"); return VMASB200_ERR_INVALID;
  }
  if (n == 0) return VMASB200_OK;
  ActExParams ap{};
  for (int k = 0; k < a->action_size; ++k) { ap.range[k] = a->u_range[k]; ap.mult[k] = a->u_multiplier[k]; }
  ap.k_n = a->action_size; ap.comm_col0 = a->comm_col0; ap.n_comm = a->n_comm; ap.clamp = a->clamp;
  ap.u_noise = (float)a->u_noise; ap.c_noise = (float)a->c_noise;
  const long long total = (long long)n * (a->action_size + a->n_comm);
  set_action_ex_kernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      a->action, a->stride_env, a->stride_k, ap, a->u_randn, a->n_comm ? a->c_randn : nullptr, a->u_out, a->c_out, a->flags, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_set_action_discrete(const void* action, int32_t action_dtype, int64_t stride_env, int64_t stride_k,
                                 int32_t action_size, int32_t multidiscrete, const int32_t* nvec, const float* u_range,
                                 const float* u_multiplier, float* u_out, int32_t* flags, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_set_action_discrete");
  if (!action || !nvec || !u_range || !u_multiplier || !u_out || !flags || n < 0 || action_size < 1 || action_size > 8 ||
      (action_dtype != VMASB200_I64 && action_dtype != VMASB200_I32)) {
    set_error("set_action_discrete: bad argument (1 <= action_size <= 8, int64 / int32 actions)"); return VMASB200_ERR_INVALID;
  }
  DiscParams dp{};
  for (int k = 0; k < action_size; ++k) {
    if (nvec[k] < 2) { set_error("set_action_discrete: nvec entries must be > 1"); return VMASB200_ERR_INVALID; }
    dp.nvec[k] = nvec[k]; dp.range[k] = u_range[k]; dp.mult[k] = u_multiplier[k];
  }
  long long below = 1;
  for (int k = action_size - 1; k >= 0; --k) { dp.below[k] = below; below *= nvec[k]; }
  if (n == 0) return VMASB200_OK;
  const long long total = (long long)n * action_size;
  const unsigned grid = (unsigned)((total + 255) / 256);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (action_dtype == VMASB200_I64)
    set_action_discrete_kernel<long long><<<grid, 256, 0, st>>>(static_cast<const long long*>(action), stride_env, stride_k,
                                                                  action_size, multidiscrete, dp, u_out, flags, n);
  else
    set_action_discrete_kernel<int><<<grid, 256, 0, st>>>(static_cast<const int*>(action), stride_env, stride_k, action_size,
                                                          multidiscrete, dp, u_out, flags, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

// ---------------------------------------------------------------------------------------------- scenario hooks (8 f1)
static DVec to_dvec(const vmasb200_vec& v) { DVec d; d.p = v.p; d.se = v.stride_env; d.sk = v.stride_k; return d; }

int vmasb200_gather_obs(const vmasb200_obs_seg* segs, int32_t n_segs, float* const* outs, const int32_t* out_widths,
                        int32_t n_outs, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_gather_obs");
  if (!segs || !outs || !out_widths || n < 0 || n_segs < 1 || n_segs > VMASB200_OBS_MAX_SEGS || n_outs < 1 ||
      n_outs > VMASB200_OBS_MAX_OUTS) {
    set_error("gather_obs: bad argument (1 <= n_segs <= 128, 1 <= n_outs <= 32)"); return VMASB200_ERR_INVALID;
  }
  // by value in the kernel's parameter space (grid constant): pointers change every step, nothing is uploaded
  static thread_local ObsTable tb;
  tb.n_segs = n_segs; tb.n_outs = n_outs;
  int max_w = 0;
  std::vector<std::vector<uint8_t>> cover((size_t)n_outs);
  for (int o = 0; o < n_outs; ++o) {
    if (!outs[o] || out_widths[o] < 1 || out_widths[o] > VMASB200_OBS_MAX_WIDTH) {
      set_error("gather_obs: out pointer NULL or width outside 1..352"); return VMASB200_ERR_INVALID;
    }
    tb.outs[o] = outs[o]; tb.width[o] = out_widths[o];
    max_w = std::max(max_w, (int)out_widths[o]);
    cover[o].assign((size_t)out_widths[o], 0);
  }
  for (int s = 0; s < n_segs; ++s) {
    const vmasb200_obs_seg& g = segs[s];
    if (g.out < 0 || g.out >= n_outs || g.width < 1 || g.col0 < 0 || g.col0 + g.width > out_widths[g.out] || !g.a.p ||
        g.op < VMASB200_OBS_COPY || g.op > VMASB200_OBS_SIN || (g.op == VMASB200_OBS_SUB && !g.b.p)) {
      set_error("gather_obs: bad segment (out, columns, operands or op)"); return VMASB200_ERR_INVALID;
    }
    for (int k = 0; k < g.width; ++k)
      if (cover[g.out][(size_t)(g.col0 + k)]++) { set_error("gather_obs: a column is written by two segments"); return VMASB200_ERR_INVALID; }
    ObsSeg& d = tb.seg[s];
    d.a = to_dvec(g.a); d.b = to_dvec(g.b); d.noise = to_dvec(g.noise); d.op = g.op; d.c = g.c; d.width = g.width; d.out = g.out;
    d.col0 = g.col0; d.flags = g.flags;
  }
  for (int o = 0; o < n_outs; ++o)
    for (uint8_t c : cover[o])
      if (!c) { set_error("gather_obs: a column of an observation is not covered by any segment"); return VMASB200_ERR_INVALID; }
  if (n == 0) return VMASB200_OK;
  const dim3 grid((unsigned)((n + kObsTile - 1) / kObsTile), (unsigned)n_outs);
  gather_obs_kernel<<<grid, 256, (size_t)max_w * (kObsTile + 1) * sizeof(float), static_cast<cudaStream_t>(stream)>>>(tb, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_navigation_reward(const vmasb200_navigation* nv, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_navigation_reward");
  if (!nv || n < 0 || nv->n_agents < 1 || nv->n_agents > VMASB200_NAV_MAX_AGENTS || !nv->pos_rew || !nv->final_rew ||
      !nv->all_goal_reached || !nv->pair_flags) {
    set_error("navigation_reward: bad argument (1 <= n_agents <= 16, output pointers)"); return VMASB200_ERR_INVALID;
  }
  static thread_local NavParams np;
  np.n_agents = nv->n_agents; np.shared_rew = nv->shared_rew;
  np.factor = (float)nv->pos_shaping_factor; np.final_reward = (float)nv->final_reward;
  np.penalty = (float)nv->agent_collision_penalty; np.min_coll_dist = (float)nv->min_collision_distance;
  bool any_pair = false;
  for (int a = 0; a < nv->n_agents; ++a) {
    const vmasb200_nav_agent& g = nv->agent[a];
    if (!g.pos.p || !g.goal_pos.p || !g.pos_shaping || !g.pos_shaping_out || !g.distance_to_goal || !g.on_goal ||
        !g.pos_rew || !g.agent_collision_rew || !g.reward) {
      set_error("navigation_reward: an agent has a NULL tensor"); return VMASB200_ERR_INVALID;
    }
    NavAgent& d = np.ag[a];
    d.pos = to_dvec(g.pos); d.goal = to_dvec(g.goal_pos);
    d.radius = (float)g.radius; d.goal_radius = (float)g.goal_radius;
    d.shaping_in = g.pos_shaping; d.shaping_out = g.pos_shaping_out; d.dist = g.distance_to_goal; d.on_goal = g.on_goal;
    d.pos_rew = g.pos_rew; d.coll_rew = g.agent_collision_rew; d.reward = g.reward;
    for (int b = 0; b < a; ++b) {
      np.coll[a][b] = nv->collidable[a][b] ? 1 : 0;
      any_pair = any_pair || np.coll[a][b];
      np.rc_sum[a][b] = (float)(g.radius + nv->agent[b].radius);  // circumscribed radius of a sphere = its radius
    }
  }
  np.pos_rew = nv->pos_rew; np.final_rew = nv->final_rew; np.all_goal = nv->all_goal_reached; np.done = nv->done;
  np.flags = nv->pair_flags;
  if (n == 0) return VMASB200_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)((n + 255) / 256);
  VB_CUDA(cudaMemsetAsync(nv->pair_flags, 0, 4 * sizeof(uint32_t), st));
  const int A = nv->n_agents;
  if (any_pair) {
    if (A <= 4) nav_flags_kernel<4><<<grid, 256, 0, st>>>(np, n);
    else if (A <= 8) nav_flags_kernel<8><<<grid, 256, 0, st>>>(np, n);
    else nav_flags_kernel<16><<<grid, 256, 0, st>>>(np, n);
  }
  if (A <= 4) nav_reward_kernel<4><<<grid, 256, 0, st>>>(np, n);
  else if (A <= 8) nav_reward_kernel<8><<<grid, 256, 0, st>>>(np, n);
  else nav_reward_kernel<16><<<grid, 256, 0, st>>>(np, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

// ---- allocation-free variant of vmasb200_navigation_reward (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops, harnesses that
// forbid allocation).  The scratch of the flag reduction (nav_flags_kernel -> nav_reward_kernel: one bit per agent pair,
// 120 pairs = 4 words at the maximum of 16 agents) lives in a file-scope device array that every GPU loading the library
// holds until exit.  g_nav_static_flags is a HOST-side stand-in for that array in host code, not a device address: the
// device address is asked for once per device, with that device current, by cudaGetSymbolAddress.
static __device__ __align__(256) uint32_t g_nav_static_flags[VMASB200_NAV_STATIC_WORDS];
static uint32_t* g_nav_static_ptr[64];  // per device, filled on the first call there
static std::mutex g_nav_static_mu;

int vmasb200_navigation_reward_static(const vmasb200_navigation* nv, int64_t n, void* stream) {
  if (!nv) { set_error("navigation_reward_static: NULL argument"); return VMASB200_ERR_INVALID; }
  if (nv->n_agents < 1 || nv->n_agents > VMASB200_NAV_MAX_AGENTS) {
    set_error("navigation_reward_static: n_agents = " + std::to_string(nv->n_agents) + " is outside 1.." +
              std::to_string(VMASB200_NAV_MAX_AGENTS) + ": the static scratch holds the pair flags of at most " +
              std::to_string(VMASB200_NAV_MAX_AGENTS) + " agents (120 pairs, 4 of its " +
              std::to_string(VMASB200_NAV_STATIC_WORDS) + " words)");
    return VMASB200_ERR_INVALID;
  }
  static_assert((VMASB200_NAV_MAX_AGENTS * (VMASB200_NAV_MAX_AGENTS - 1) / 2 + 31) / 32 <= VMASB200_NAV_STATIC_WORDS, "scratch too small");
  int dev = -1;
  VB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_error("navigation_reward_static: device index outside 0..63"); return VMASB200_ERR_UNSUPPORTED; }
  uint32_t* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_nav_static_mu);
    if (!g_nav_static_ptr[dev]) {
      void* p = nullptr;
      VB_CUDA(cudaGetSymbolAddress(&p, g_nav_static_flags));
      g_nav_static_ptr[dev] = static_cast<uint32_t*>(p);
    }
    scratch = g_nav_static_ptr[dev];
  }
  // same launches as vmasb200_navigation_reward, which zeroes the words it uses (cudaMemsetAsync on the stream) before
  // nav_flags_kernel ORs into them: the array keeps the previous call's bits otherwise
  vmasb200_navigation with_scratch = *nv;
  with_scratch.pair_flags = scratch;
  return vmasb200_navigation_reward(&with_scratch, n, stream);
}

int vmasb200_transport_reward(const vmasb200_transport* tr, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_transport_reward");
  if (!tr || n < 0 || tr->n_packages < 1 || tr->n_packages > VMASB200_TRANSPORT_MAX_PACKAGES || !tr->rew) {
    set_error("transport_reward: bad argument (1 <= n_packages <= 8, rew)"); return VMASB200_ERR_INVALID;
  }
  static thread_local TrParams tp;
  tp.n_packages = tr->n_packages; tp.factor = (float)tr->shaping_factor;
  for (int k = 0; k < 3; ++k) { tp.off[k] = tr->color_off_goal[k]; tp.on[k] = tr->color_on_goal[k]; }
  for (int k = 0; k < tr->n_packages; ++k) {
    const vmasb200_transport_package& g = tr->package[k];
    int rc;
    if ((rc = to_qpose(&g.package, &tp.pk[k].pkg, "transport_reward")) || (rc = to_qpose(&g.goal, &tp.pk[k].goal, "transport_reward"))) return rc;
    if (!g.global_shaping || !g.global_shaping_out || !g.dist_to_goal || !g.on_goal) {
      set_error("transport_reward: a package has a NULL tensor"); return VMASB200_ERR_INVALID;
    }
    tp.pk[k].shaping_in = g.global_shaping; tp.pk[k].shaping_out = g.global_shaping_out; tp.pk[k].dist = g.dist_to_goal;
    tp.pk[k].on_goal = g.on_goal; tp.pk[k].color = g.color;
  }
  tp.rew = tr->rew; tp.done = tr->done;
  if (n == 0) return VMASB200_OK;
  transport_reward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(tp, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_simple_spread_reward(const vmasb200_simple_spread* sp, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_simple_spread_reward");
  if (!sp || n < 0 || sp->n_agents < 1 || sp->n_agents > VMASB200_SPREAD_MAX || sp->n_landmarks < 0 ||
      sp->n_landmarks > VMASB200_SPREAD_MAX || !sp->rew) {
    set_error("simple_spread_reward: bad argument (1 <= n_agents <= 16, n_landmarks <= 16, rew)"); return VMASB200_ERR_INVALID;
  }
  static thread_local SpreadParams p;
  p.n_agents = sp->n_agents; p.n_landmarks = sp->n_landmarks; p.rew = sp->rew;
  for (int a = 0; a < sp->n_agents; ++a) {
    if (!sp->agent_pos[a].p) { set_error("simple_spread_reward: an agent position is NULL"); return VMASB200_ERR_INVALID; }
    p.apos[a] = to_dvec(sp->agent_pos[a]); p.radius[a] = (float)sp->agent_radius[a]; p.collide[a] = sp->collide[a] ? 1 : 0;
  }
  for (int l = 0; l < sp->n_landmarks; ++l) {
    if (!sp->landmark_pos[l].p) { set_error("simple_spread_reward: a landmark position is NULL"); return VMASB200_ERR_INVALID; }
    p.lpos[l] = to_dvec(sp->landmark_pos[l]);
  }
  if (n == 0) return VMASB200_OK;
  simple_spread_reward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(p, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_joint_passage_reward(const vmasb200_joint_passage* g, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_joint_passage_reward");
  if (!g || n < 0 || g->n_agents < 1 || g->n_agents > 2 || g->n_bodies < g->n_agents || g->n_bodies > VMASB200_JP_MAX_BODIES ||
      g->n_passages < 1 || g->n_passages > VMASB200_JP_MAX_PASSAGES || g->n_walls < 0 || g->n_walls > VMASB200_JP_MAX_WALLS) {
    set_error("joint_passage_reward: bad argument (1..2 agents, <= 4 bodies, 1..24 passages, <= 8 walls)"); return VMASB200_ERR_INVALID;
  }
  if (!g->pos_rew || !g->rot_rew || !g->collision_rew || !g->energy_expenditure || !g->energy_rew || !g->rew || !g->all_passed ||
      !g->goal_pos.p || !g->goal_rot.p || (g->done && !g->done_flag)) {
    set_error("joint_passage_reward: a NULL tensor"); return VMASB200_ERR_INVALID;
  }
  static thread_local JpParams jp;
  int rc;
  jp.n_agents = g->n_agents; jp.n_bodies = g->n_bodies; jp.n_passages = g->n_passages; jp.n_walls = g->n_walls;
  jp.all_passed_rot = g->all_passed_rot;
  bool any_open = false;
  for (int a = 0; a < g->n_bodies; ++a) {
    if ((rc = to_qpose(&g->body[a], &jp.body[a], "joint_passage_reward"))) return rc;
    if (a < g->n_agents) {
      if (!g->action_u[a].p) { set_error("joint_passage_reward: agent.action.u is NULL"); return VMASB200_ERR_INVALID; }
      jp.action_u[a] = to_dvec(g->action_u[a]);
      jp.e_div[a] = (float)std::sqrt((double)g->dim_p * (g->f_range[a] * g->f_range[a]));  // math.sqrt(dim_p * f_range**2)
    }
  }
  for (int p = 0; p < g->n_passages; ++p) {
    if ((rc = to_qpose(&g->passage[p], &jp.passage[p], "joint_passage_reward"))) return rc;
    jp.pcollide[p] = g->passage_collide[p] ? 1 : 0;
    any_open = any_open || !jp.pcollide[p];
  }
  if (!any_open) { set_error("joint_passage_reward: no passage without collision (the gap the joint is shaped towards)"); return VMASB200_ERR_INVALID; }
  for (int w = 0; w < g->n_walls; ++w)
    if ((rc = to_qpose(&g->wall[w], &jp.wall[w], "joint_passage_reward"))) return rc;
  if ((rc = to_qpose(&g->joint, &jp.joint, "joint_passage_reward"))) return rc;
  if (!jp.joint.rot) { set_error("joint_passage_reward: the joint landmark needs a rotation"); return VMASB200_ERR_INVALID; }
  jp.goal_pos = to_dvec(g->goal_pos); jp.goal_rot = to_dvec(g->goal_rot);
  const double pi = 3.141592653589793;
  jp.half_width = (float)(g->passage_width / 2); jp.pos_factor = (float)g->pos_shaping_factor; jp.rot_factor = (float)g->rot_shaping_factor;
  jp.coll_rew = (float)g->collision_reward; jp.energy_coeff = (float)g->energy_reward_coeff; jp.min_coll = (float)g->min_collision_distance;
  jp.pi_f = (float)pi;
  double mg = std::fmod(g->middle_angle, pi);  // Python float % float
  if (mg != 0.0 && mg < 0.0) mg += pi;
  jp.mid_g = (float)mg; jp.mid_g_minus_pi = (float)(mg - pi);
  for (int k = 0; k < 4; ++k) {
    if (!g->shaping_in[k] || !g->shaping_out[k]) { set_error("joint_passage_reward: a shaping tensor is NULL"); return VMASB200_ERR_INVALID; }
    jp.sh_in[k] = g->shaping_in[k]; jp.sh_out[k] = g->shaping_out[k];
  }
  jp.pos_rew = g->pos_rew; jp.rot_rew = g->rot_rew; jp.collision_rew = g->collision_rew; jp.energy_exp = g->energy_expenditure;
  jp.energy_rew = g->energy_rew; jp.rew = g->rew; jp.all_passed = g->all_passed; jp.done = g->done; jp.done_flag = g->done_flag;
  if (n == 0) return VMASB200_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g->done) VB_CUDA(cudaMemsetAsync(g->done_flag, 0, sizeof(uint32_t), st));
  jp_reward_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(jp, n);
  if (g->done) jp_done_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g->done, g->done_flag, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_football_reward(const vmasb200_football* g, int64_t n, void* stream) {
  NvtxRange nvtx_("vmasb200_football_reward");
  if (!g || n < 0 || !g->ball_pos.p || !g->ball_vel.p || !g->done) { set_error("football_reward: bad argument"); return VMASB200_ERR_INVALID; }
  static thread_local FbParams fb;
  fb.ball_pos = to_dvec(g->ball_pos); fb.ball_vel = to_dvec(g->ball_vel);
  fb.x_right = (float)(g->pitch_length / 2 + g->ball_size / 2); fb.x_left = (float)(-g->pitch_length / 2 - g->ball_size / 2);
  fb.y_hi = (float)(g->goal_size / 2); fb.y_lo = (float)(-g->goal_size / 2);
  fb.scoring = (float)g->scoring_reward; fb.f_ball_goal = (float)g->pos_shaping_factor_ball_goal;
  fb.f_agent_ball = (float)g->pos_shaping_factor_agent_ball; fb.trigger = (float)g->distance_to_ball_trigger;
  for (int t = 0; t < 2; ++t) {
    const vmasb200_football_team& m = g->team[t];
    FbTeam& d = fb.team[t];
    if (m.n_agents < 0 || m.n_agents > VMASB200_FB_MAX_TEAM || !m.sparse_reward || !m.reward) {
      set_error("football_reward: bad team (<= 16 agents, sparse_reward, reward)"); return VMASB200_ERR_INVALID;
    }
    d.n_agents = m.n_agents; d.dense = m.dense ? 1 : 0;
    if (d.dense) {
      if (m.n_agents < 1 || !m.pos_shaping || !m.pos_shaping_out || !m.pos_shaping_agent || !m.pos_shaping_agent_out ||
          !m.distance_to_goal || !m.pos_rew || !m.min_agent_dist_to_ball || !m.pos_rew_agent || !m.dense_reward) {
        set_error("football_reward: a dense-reward tensor is NULL"); return VMASB200_ERR_INVALID;
      }
      for (int a = 0; a < m.n_agents; ++a) {
        if (!m.agent_pos[a].p) { set_error("football_reward: an agent position is NULL"); return VMASB200_ERR_INVALID; }
        d.apos[a] = to_dvec(m.agent_pos[a]);
      }
    }
    d.gx = m.goal_pos[0]; d.gy = m.goal_pos[1];
    d.sh_in = m.pos_shaping; d.sh_out = m.pos_shaping_out; d.sha_in = m.pos_shaping_agent; d.sha_out = m.pos_shaping_agent_out;
    d.dist_goal = m.distance_to_goal; d.pos_rew = m.pos_rew; d.min_dist = m.min_agent_dist_to_ball; d.pos_rew_agent = m.pos_rew_agent;
    d.dense_rew = m.dense_reward; d.sparse_rew = m.sparse_reward; d.reward = m.reward;
  }
  fb.done = g->done;
  if (n == 0) return VMASB200_OK;
  fb_reward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(fb, n);
  VB_CUDA(cudaGetLastError());
  return VMASB200_OK;
}

int vmasb200_sync_check(vmasb200_plan* pl, void* stream, int64_t* n_fixups) {
  NvtxRange nvtx_("vmasb200_sync_check");
  VB_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  VB_CUDA(cudaGetLastError());
  if (n_fixups) {
    *n_fixups = 0;
    if (pl && pl->uploaded) {
      int c = 0;
      VB_CUDA(cudaMemcpy(&c, reinterpret_cast<int*>(pl->d_bits + 4 * bits_words(pl->kp)) + 2, sizeof(int), cudaMemcpyDeviceToHost));
      *n_fixups = c;
    }
  }
  return VMASB200_OK;
}

}  // extern "C"
