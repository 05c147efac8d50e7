// jit.h - world-compiled step kernels: host side (included by vmasb200.cu after kernels.cuh; host code only).
//
// The generic step kernel (kernels.cuh) INTERPRETS the plan tables: per entity and per pair it reads flags, row
// indices and constants and branches on them, which ncu showed to be ~10x the instructions of the physics itself for
// small worlds (profiles/README.md: simple_spread 2200 instructions per env for 3 agents and 3 pairs).  For worlds
// whose candidate pairs are all sphere-sphere and that have no joints (simple_spread, navigation and most MPE
// scenarios) the plan is instead COMPILED: `generate()` prints the world as C++ - constants, one literal EntC per
// entity, straight-line calls per pair in the reference's order - around csrc/spec_helpers.cuh / spec_kernel.cuh,
// NVRTC compiles it for sm_100a with the parity flags of the library build (--fmad=false), and the cubin is loaded
// with the driver API and launched on the caller's stream.  The state of an environment then lives in registers for
// all substeps.  The library's own sources are embedded as text (jit_sources.inc, made by the Makefile), libnvrtc and
// libcuda are dlopen-ed on first use, cubins are cached on disk next to the library (`_jit/`, or $VMASB200_JIT_CACHE).
// No NVRTC, an ineligible world, VMASB200_JIT=0 or vmasb200_jit_enable(0): the generic kernel runs (same bits).
#pragma once
#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdarg>
#include <fstream>
#include <map>
#include <sstream>

#include "jit_sources.inc"

namespace vb {

// ------------------------------------------------------------------------------------------------ dynamic APIs
struct NvrtcApi {
  void* h = nullptr;
  int (*CreateProgram)(void**, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
  int (*CompileProgram)(void*, int, const char* const*) = nullptr;
  int (*GetCUBINSize)(void*, size_t*) = nullptr;
  int (*GetCUBIN)(void*, char*) = nullptr;
  int (*GetProgramLogSize)(void*, size_t*) = nullptr;
  int (*GetProgramLog)(void*, char*) = nullptr;
  int (*DestroyProgram)(void**) = nullptr;
  int (*Version)(int*, int*) = nullptr;
  bool ok = false;
};
struct DriverApi {
  void* h = nullptr;
  int (*ModuleLoadData)(void**, const void*) = nullptr;
  int (*ModuleUnload)(void*) = nullptr;
  int (*ModuleGetFunction)(void**, void*, const char*) = nullptr;
  int (*ModuleGetGlobal)(unsigned long long*, size_t*, void*, const char*) = nullptr;
  int (*MemcpyHtoD)(unsigned long long, const void*, size_t) = nullptr;
  int (*FuncSetAttribute)(void*, int, int) = nullptr;
  int (*LaunchKernel)(void*, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void*, void**, void**) = nullptr;
  int (*LaunchKernelEx)(const void*, void*, void**, void**) = nullptr;  // optional (CUDA >= 11.8): programmatic dependent launch
  bool ok = false;
};
static std::mutex g_jit_mu;
static NvrtcApi g_nvrtc;
static DriverApi g_drv;
static int g_jit_on = -1;  // -1: read VMASB200_JIT on first use

template <class F> static bool sym(void* h, const char* name, F& f) {
  f = reinterpret_cast<F>(dlsym(h, name));
  return f != nullptr;
}
static bool load_nvrtc() {
  if (g_nvrtc.h) return g_nvrtc.ok;
  // the toolkit's NVRTC first: same release as the nvcc that built the library (same libdevice math)
  for (const char* name : {"/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so.12", "libnvrtc.so"}) {
    g_nvrtc.h = dlopen(name, RTLD_NOW | RTLD_LOCAL);
    if (g_nvrtc.h) break;
  }
  if (!g_nvrtc.h) { g_nvrtc.h = reinterpret_cast<void*>(1); return false; }
  NvrtcApi& a = g_nvrtc;
  a.ok = sym(a.h, "nvrtcCreateProgram", a.CreateProgram) && sym(a.h, "nvrtcCompileProgram", a.CompileProgram) &&
         sym(a.h, "nvrtcGetCUBINSize", a.GetCUBINSize) && sym(a.h, "nvrtcGetCUBIN", a.GetCUBIN) &&
         sym(a.h, "nvrtcGetProgramLogSize", a.GetProgramLogSize) && sym(a.h, "nvrtcGetProgramLog", a.GetProgramLog) &&
         sym(a.h, "nvrtcDestroyProgram", a.DestroyProgram) && sym(a.h, "nvrtcVersion", a.Version);
  return a.ok;
}
static bool load_driver() {
  if (g_drv.h) return g_drv.ok;
  g_drv.h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_LOCAL);
  if (!g_drv.h) { g_drv.h = reinterpret_cast<void*>(1); return false; }
  DriverApi& d = g_drv;
  d.ok = sym(d.h, "cuModuleLoadData", d.ModuleLoadData) && sym(d.h, "cuModuleUnload", d.ModuleUnload) &&
         sym(d.h, "cuModuleGetFunction", d.ModuleGetFunction) && sym(d.h, "cuModuleGetGlobal_v2", d.ModuleGetGlobal) &&
         sym(d.h, "cuMemcpyHtoD_v2", d.MemcpyHtoD) && sym(d.h, "cuFuncSetAttribute", d.FuncSetAttribute) &&
         sym(d.h, "cuLaunchKernel", d.LaunchKernel);
  sym(d.h, "cuLaunchKernelEx", d.LaunchKernelEx);
  return d.ok;
}
static int jit_mode() {  // 0 off, 1 worlds the compiled kernel is measured to win on, 2 every world the generator covers
  if (g_jit_on < 0) {
    const char* e = getenv("VMASB200_JIT");
    g_jit_on = (e && e[0] == '0') ? 0 : (e && e[0] == '2') ? 2 : 1;
  }
  return g_jit_on;
}
static bool jit_on() { return jit_mode() != 0; }

// ------------------------------------------------------------------------------------------------ eligibility
// Worlds the generator covers: every item type, within sizes that keep the unrolled code and the register state small.
static bool jit_big_world(const vmasb200_plan* pl);
static bool jit_eligible(const vmasb200_plan* pl, int te = 0) {  // te: the CTA tile of the launch in question (32 = small batch), 0 = any
  const KParams& kp = pl->kp;
  if (kp.n_dyn < 1 || kp.n_dyn > 16 || kp.P > 128 || kp.J > 16 || kp.S > 64 || kp.A > 16 || kp.n_sensors > 32 || kp.n_trig > 48)
    return false;
  int n_stat_used = 0;
  for (const DEnt& e : pl->ents) n_stat_used += e.stat_off >= 0;
  if (n_stat_used > 48) return false;
  if (jit_mode() >= 2) return true;
  // Measured on B200 (profiles/README.md): everything in registers wins while the pair list is short or all
  // sphere-sphere (simple_spread 3.4x, transport 2.7x, navigation 2.3x at 1 M envs).  Pair-heavy worlds without LIDAR take
  // the BIG variant (jit_big_world: static tile, looped broad phase, warp-cooperative table-driven narrow phase): football
  // 2.5x at 1 M envs, 1.6x at 32 K; joint_passage 2x.  Pair-heavy worlds WITH a LIDAR epilogue keep the generic kernel.
  bool all_ss = kp.J == 0;
  for (int p = 0; p < kp.P && all_ss; ++p) all_ss = pl->items[kp.J + p].type == IT_SS;
  if (all_ss || (kp.P <= 32 && kp.J <= 2)) return true;
  if (!jit_big_world(pl)) return false;
  // ... where it was measured to win (tools/ab_worlds.sh): small batches (32-env tiles, one wave) need more than one substep
  // to amortise the register load of the whole world (passage, ball_passage: S = 1, 0.9x); large batches lose with box-box
  // pairs, whose narrow phase is ~5000 flops on a 168-register kernel (waterfall 0.8x at 262 144 envs, 1.4x at 32 768)
  bool has_bb = false;
  for (int p = 0; p < kp.P; ++p) has_bb = has_bb || pl->items[kp.J + p].type == IT_BB;
  if (te == 32) return kp.S > 1;
  if (te != 0) return !has_bb;
  return true;
}

// ------------------------------------------------------------------------------------------------ generator
static std::string fl(float v) {  // exact fp32 literal
  char b[64];
  if (std::isinf(v)) return v > 0 ? "CUDART_INF_F" : "(-CUDART_INF_F)";
  if (std::isnan(v)) return "CUDART_NAN_F";
  snprintf(b, sizeof b, "%af", (double)v);
  return b;
}
static void outf(std::string& s, const char* fmt, ...) {
  char b[2048];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(b, sizeof b, fmt, ap);
  va_end(ap);
  s += b;
}

struct JitLayout {            // the compact LIDAR tile of the compiled kernel (rows of TE floats)
  std::vector<int> lid_ents;  // entities staged for the epilogue: 3 env rows each (px py rot)
  std::vector<int> lid_trig;  // of those, the non-sphere targets: 4 work rows each (cos/sin cache)
  int work_rows = 0;
  int smem_rows = 0;          // > 0: the compiled epilogue's own carve-up (rows of TE words), else kernels.cuh smem_map
  int stat_tile_rows = 0;     // > 0: pair-heavy world: static entities (pose + cos/sin) live in a shared-memory tile, one column per thread
  // ... and, behind them, per dynamic entity px py [c s [c2 s2]] and its (Fx Fy T) accumulators: what the table-driven
  // narrow phase (spec_helpers.cuh narrow_big -> step_kernel.cuh eval_item / accumulate) reads and adds to.  Tile rows:
  std::vector<int> tile_pos, tile_trig, tile_rot, tile_fin;  // per entity: row of px / of cos / of rot / of Fx, or -1
  bool thin_possible = false;
  bool coop = false;          // narrow_coop scratch behind the tile: (nt / 32) * (64 / 2 + 4 * 64) words
};

// Pair-heavy worlds (football: 91 pairs, joint_passage: 54 pairs + 3 joints over 10 substeps): with everything in registers
// the straight-line kernel needs 213 - 255 of them and issues at ~10 %.  For those the generator keeps the STATIC entities
// (pose and cos / sin, which never change inside a step) in a shared-memory tile with one private column per thread
// (literal row offsets: one LDS per use, no barrier) and calls the sphere-sphere / line-sphere narrow phases out of line,
// like box-sphere: ~50 registers and most of the code size less.
// The narrow phase itself is TABLE-DRIVEN there (narrow_big): the unrolled list only does the broad phase (verdict and
// survivor bits, branch-free); then every lane walks ITS OWN survivors through the generic eval_item / accumulate on the
// tile, so that lanes whose contacts are on different pairs still run together (straight-line narrow phases ran once per
// pair that ANY lane of the warp had, which is what made the all-in-registers kernel lose on these worlds).
static bool jit_big_world(const vmasb200_plan* pl) {
  const KParams& kp = pl->kp;
  if (kp.n_rays > 0) return false;
  if (const char* e = getenv("VMASB200_JIT_BIG")) return atoi(e) != 0;
  bool all_ss = kp.J == 0;
  for (int p = 0; p < kp.P && all_ss; ++p) all_ss = pl->items[kp.J + p].type == IT_SS;
  return !(all_ss || (kp.P <= 32 && kp.J <= 2));
}

// The world of `pl` as C++: prelude constants, then (after spec_helpers.cuh) the world_* functions.
static std::string generate(const vmasb200_plan* pl, int te, int nt, JitLayout* lay) {
  const KParams& kp = pl->kp;
  std::string s;
  s += "#include \"kernels.cuh\"\nnamespace vb { namespace gen {\n";
  int n_stat = 0, n_jf = 0;
  const size_t E = pl->ents.size();
  std::vector<int> stat_slot(E, -1), dyn_slot(E, -1), trig_slot(E, -1);
  for (size_t e = 0; e < E; ++e)
    if (pl->ents[e].stat_off >= 0) stat_slot[e] = n_stat++;
  for (int d = 0; d < kp.n_dyn; ++d) dyn_slot[pl->dyn_list[d]] = d;
  for (int t = 0; t < kp.n_trig; ++t) trig_slot[pl->trig_list[t]] = t;
  std::vector<int> jf_slot((size_t)kp.J, -1);
  for (int j = 0; j < kp.J; ++j)
    if (pl->items[j].fr_row >= 0) jf_slot[j] = n_jf++;
  const bool lidar = kp.n_rays > 0;
  const bool big = jit_big_world(pl);
  // static tile rows of entity e (big worlds): px py rot [c s [c2 s2]]
  std::vector<int> srow(E, -1);
  int n_srows = 0;
  if (big)
    for (size_t e = 0; e < E; ++e)
      if (stat_slot[e] >= 0) {
        srow[e] = n_srows;
        n_srows += 3 + (trig_slot[e] >= 0 ? (pl->ents[e].shape == VMASB200_BOX ? 4 : 2) : 0);
      }
  // behind the static rows: dynamic entities' px py [c s [c2 s2]], then every dynamic entity's Fx Fy T
  std::vector<int> drow(E, -1), frow(E, -1);
  if (big) {
    for (int d = 0; d < kp.n_dyn; ++d) {
      const int e = pl->dyn_list[d];
      drow[e] = n_srows;
      n_srows += 2 + (trig_slot[e] >= 0 ? (pl->ents[e].shape == VMASB200_BOX ? 4 : 2) : 0);
    }
    for (int d = 0; d < kp.n_dyn; ++d) { frow[pl->dyn_list[d]] = n_srows; n_srows += 3; }
  }
  // ... then (spec_helpers.cuh PARK) vx vy rot w of every dynamic entity and fx fy tq of every agent
  // worth it when it frees enough registers for another warp per scheduler (football: 49 words, 197 -> 167 registers,
  // 511 -> 417 us at 1 M envs); joint_passage's 22 words only grow its tile (3 -> 2 CTAs per SM: 3300 -> 3980 us)
  bool park = big && 4 * kp.n_dyn + 3 * kp.A >= 32;
  if (const char* e = getenv("VMASB200_JIT_PARK")) park = big && atoi(e) != 0;  // experiments
  const int acc_row = kp.n_dyn > 0 && big ? frow[pl->dyn_list[0]] : 0, park_row = n_srows;
  if (park) n_srows += 4 * kp.n_dyn + 3 * kp.A;
  lay->stat_tile_rows = n_srows;
  lay->tile_pos.assign(E, -1); lay->tile_trig.assign(E, -1); lay->tile_rot.assign(E, -1); lay->tile_fin.assign(E, -1);
  for (size_t e = 0; e < E && big; ++e) {
    if (srow[e] >= 0) { lay->tile_pos[e] = srow[e]; lay->tile_rot[e] = srow[e] + 2; if (trig_slot[e] >= 0) lay->tile_trig[e] = srow[e] + 3; }
    if (drow[e] >= 0) { lay->tile_pos[e] = drow[e]; if (trig_slot[e] >= 0) lay->tile_trig[e] = drow[e] + 2; lay->tile_fin[e] = frow[e]; }
  }
  bool thin_possible = false;
  std::vector<char> thin((size_t)kp.P, 1);
  for (int p = 0; p < kp.P; ++p) {
    const DItem& it = pl->items[kp.J + p];
    // a sphere-sphere force is non-zero only within fp32(r_a + r_b); no thin event when the verdict's threshold covers it
    if (it.type == IT_SS && pl->ents[it.a].radius + pl->ents[it.b].radius <= it.thr) thin[p] = 0;
    thin_possible = thin_possible || thin[p];
  }
  lay->thin_possible = thin_possible;
  outf(s, "constexpr int TE = %d, NT = %d, ND = %d, NST = %d, NA = %d, NTR = %d, NJF = %d, S = %d, P = %d, PW = %d, DYN_ROWS = %d, NJ = %d;\n",
       te, nt, kp.n_dyn, big ? 0 : n_stat, kp.A, kp.n_trig, n_jf, kp.S, kp.P, kp.pair_words, kp.dyn_rows, kp.J);
  outf(s, "constexpr bool LIDAR = %s, HAS_XS = %s, HAS_YS = %s, HAS_WGRAV = %s, THIN_POSSIBLE = %s, BIG = %s;\n", lidar ? "true" : "false",
       kp.has_xs ? "true" : "false", kp.has_ys ? "true" : "false", kp.has_wgrav ? "true" : "false", thin_possible ? "true" : "false",
       big ? "true" : "false");
  {
    // resident CTAs per SM the register allocator must leave room for (VMASB200_JIT_MINB, experiments only)
    // 32-env tiles of a LIDAR world (128 threads) serve batches up to 48 K envs: 32 768 envs are 1024 CTAs, which fit the 148
    // SMs in ONE wave only at 7 CTAs per SM, i.e. at most 72 registers (navigation compiled to 80: 888 resident CTAs and a
    // second wave of 136)
    // 64-env tiles (96 threads, large batches): 10 CTAs per SM = 64 registers (navigation compiles to 80 uncapped; 16 more
    // bytes of spill; 125.8 -> 117.9 us at 1 M envs with the DIRECT epilogue's smaller tile)
    // pair-heavy worlds: three warps per scheduler = at most 168 registers (joint_passage compiles to 172 uncapped)
    int minb = (lidar && te == 32 && nt == 128) ? 7 : (lidar && te == 64 && nt == 96) ? 10 : big ? 384 / nt : 1;  // otherwise measured (tools/ab_big2.sh): capping the registers of the BIG variant below its natural 168 trades the occupancy for spills one to one
    if (const char* e = getenv("VMASB200_JIT_MINB")) minb = std::max(1, atoi(e));
    outf(s, "constexpr int MINB = %d;\n", minb);
  }
  {
    // BIG worlds: narrow phase shared out over the warp (spec_helpers.cuh narrow_coop); needs the warp's envs in one tile
    // (all threads carry an env: no LIDAR) and pair numbers that fit the queue's 11 bits
    bool has_bbox = false;
    for (int p = 0; p < kp.P; ++p) has_bbox = has_bbox || pl->items[kp.J + p].type == IT_BB;
    const bool coop = big && kp.P < 2048 && !has_bbox && !getenv("VMASB200_JIT_NOCOOP");  // box-box worlds: spec_helpers.cuh narrow_big
    bool has_bl = false;
    for (int p = 0; p < kp.P; ++p) has_bl = has_bl || pl->items[kp.J + p].type == IT_BL;
    outf(s, "constexpr bool COOP_EDGES = %s;\n", coop && has_bl && !getenv("VMASB200_JIT_NOEDGES") ? "true" : "false");
    outf(s, "constexpr bool COOP = %s, PARK = %s;\nconstexpr int TILE_ROWS = %d, PARK_ROW = %d, ACC_ROW = %d;\n", coop ? "true" : "false",
         park ? "true" : "false", n_srows, park_row, acc_row);
    lay->coop = coop;
  }
  outf(s, "constexpr float SUB_DT = %s, KC = %s, KJ = %s, KT = %s, MARGIN = %s, XS = %s, YS = %s, WGX = %s, WGY = %s;\n",
       fl(kp.sub_dt).c_str(), fl(kp.Kc).c_str(), fl(kp.Kj).c_str(), fl(kp.Kt).c_str(), fl(kp.margin).c_str(), fl(kp.xs).c_str(),
       fl(kp.ys).c_str(), fl(kp.wgx).c_str(), fl(kp.wgy).c_str());
  s += "} }\n#include \"spec_helpers.cuh\"\nnamespace vb { namespace gen {\n";
  for (size_t e = 0; e < E; ++e) {  // one literal per entity (constant-folded after inlining)
    const DEnt& d = pl->ents[e];
    outf(s, "#define VBE%d EntC{%d, 0x%xu, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s, %s}\n", (int)e, d.shape, d.flags,
         fl(d.radius).c_str(), fl(d.length).c_str(), fl(d.width).c_str(), fl(d.mass).c_str(), fl(d.moi).c_str(),
         fl(d.max_speed).c_str(), fl(d.v_range).c_str(), fl(d.one_minus_drag).c_str(), fl(d.lin_fric).c_str(),
         fl(d.ang_fric).c_str(), fl(d.gx).c_str(), fl(d.gy).c_str(), fl(d.max_f).c_str(), fl(d.f_range).c_str(),
         fl(d.max_t).c_str(), fl(d.t_range).c_str());
  }
  auto posx = [&](int e) {
    std::string r;
    if (dyn_slot[e] >= 0) outf(r, "mk(st.px[%d], st.py[%d])", dyn_slot[e], dyn_slot[e]);
    else if (big) outf(r, "mk(st.ss[%d * NT], st.ss[%d * NT])", srow[e], srow[e] + 1);
    else outf(r, "mk(st.sx[%d], st.sy[%d])", stat_slot[e], stat_slot[e]);
    return r;
  };
  auto rotx = [&](int e) {
    std::string r;
    if (dyn_slot[e] >= 0) outf(r, "st.rot[%d]", dyn_slot[e]);
    else if (stat_slot[e] >= 0 && big) outf(r, "st.ss[%d * NT]", srow[e] + 2);
    else if (stat_slot[e] >= 0) outf(r, "st.srot[%d]", stat_slot[e]);
    else r = "0.f";
    return r;
  };
  auto trigx = [&](int e) {
    std::string r;
    if (big && e >= 0 && stat_slot[e] >= 0 && trig_slot[e] >= 0)
      outf(r, "rot_s<%d, %d>(st)", srow[e] + 3, pl->ents[e].shape == VMASB200_BOX ? 1 : 0);
    else outf(r, "rot_r<%d>(st)", trig_slot[e]);
    return r;
  };
  // ---- looped runs of the broad phase (big worlds; spec_helpers.cuh pair_l): pairs p0 + j * k + i, j < n, i < k, where block
  // j is one static entity (tile rows a fixed stride apart, same shape, same role) against the same k dynamic entities
  struct Run { int p0, n, k, s0, stride; bool stat_is_a; std::vector<int> cls; int kc; };
  std::vector<Run> runs;
  std::vector<int> run_of((size_t)kp.P, -1);
  if (big && !getenv("VMASB200_JIT_NOLOOP")) {
    auto stat_of = [&](int p, bool* is_a) {  // the pair's static entity when it has exactly one, else -1
      const DItem& it = pl->items[kp.J + p];
      const bool sa_ = stat_slot[it.a] >= 0, sb_ = stat_slot[it.b] >= 0;
      if (sa_ == sb_) return -1;
      *is_a = sa_;
      return sa_ ? (int)it.a : (int)it.b;
    };
    auto frame_of = [&](int p) { return !std::isinf(pl->broad[p].hx) || !std::isinf(pl->broad[p].hy); };
    for (int p = 0; p < kp.P;) {
      bool is_a = false;
      const int S0 = stat_of(p, &is_a);
      if (S0 < 0) { ++p; continue; }
      int k = 1;
      for (bool a2; p + k < kp.P && stat_of(p + k, &a2) == S0 && a2 == is_a; ++k) {}
      auto dyn_of = [&](int q) { const DItem& it = pl->items[kp.J + q]; return is_a ? (int)it.b : (int)it.a; };
      int n = 1, stride = 0;
      for (;; ++n) {
        const int q0 = p + n * k;
        if (q0 + k > kp.P) break;
        bool a2 = false;
        const int Sj = stat_of(q0, &a2);
        if (Sj < 0 || a2 != is_a || pl->ents[Sj].shape != pl->ents[S0].shape || (trig_slot[Sj] >= 0) != (trig_slot[S0] >= 0)) break;
        const int st_j = srow[Sj] - srow[S0];
        if (n == 1) stride = st_j;
        if (st_j != n * stride || stride <= 0) break;
        bool same = true;
        for (int i = 0; i < k && same; ++i) {
          bool a3 = false;
          same = stat_of(q0 + i, &a3) == Sj && a3 == is_a && dyn_of(q0 + i) == dyn_of(p + i) && frame_of(q0 + i) == frame_of(p + i) &&
                 (frame_of(p + i) ? std::isinf(pl->broad[q0 + i].t2_cull) : true);
        }
        // the block must END here: the next pair may not belong to the same static entity (else k would differ)
        bool a4 = false;
        if (same && q0 + k < kp.P && stat_of(q0 + k, &a4) == Sj && a4 == is_a) same = false;
        if (!same) break;
      }
      bool ok_run = n >= 2 && n * k >= 6 && k <= 32;
      for (int i = 0; i < k && ok_run; ++i)
        if (frame_of(p + i) && !std::isinf(pl->broad[p + i].t2_cull)) ok_run = false;
      if (!ok_run) { p += 1; continue; }
      Run r{p, n, k, srow[S0], stride, is_a, std::vector<int>((size_t)k, 0), 0};
      for (int i = 0; i < k; ++i) {  // columns with the same thresholds in every block share their table column
        int c = -1;
        for (int i2 = 0; i2 < i && c < 0; ++i2) {
          bool eq = true;
          for (int j = 0; j < n && eq; ++j) {
            const DBroad &x = pl->broad[p + j * k + i], &y = pl->broad[p + j * k + i2];
            eq = x.t2_bp == y.t2_bp && x.t2_cull == y.t2_cull && x.hx == y.hx && x.hy == y.hy;
          }
          if (eq) c = r.cls[(size_t)i2];
        }
        r.cls[(size_t)i] = c >= 0 ? c : r.kc++;
      }
      for (int q = p; q < p + n * k; ++q) run_of[(size_t)q] = (int)runs.size();
      runs.push_back(r);
      p += n * k;
    }
    for (size_t ri = 0; ri < runs.size(); ++ri) {  // thresholds [block][class][t2_bp t2_cull hx hy]
      const Run& r = runs[ri];
      outf(s, "__constant__ uint32_t RT%d[%d] = {", (int)ri, r.n * r.kc * 4);  // bit patterns (inf has no literal)
      for (int j = 0; j < r.n; ++j)
        for (int c = 0; c < r.kc; ++c) {
          int i = 0;
          while (r.cls[(size_t)i] != c) ++i;
          const DBroad& b = pl->broad[r.p0 + j * r.k + i];
          const float v4[4] = {b.t2_bp, b.t2_cull, b.hx, b.hy};
          for (float f : v4) {
            uint32_t u;
            memcpy(&u, &f, 4);
            outf(s, "0x%08xu, ", u);
          }
        }
      s += "};\n";
    }
  }
  // ---- load
  s += "VB_DI void world_load(St& st, const AgentIO& aio, const StepArgs& sa, long long env) {\n";
  // rows through a running pointer (one 64-bit add per row instead of a 64-bit multiply-add: world_load was 12 % of the
  // football kernel's instructions)
  s += "  const size_t ld_ = (size_t)sa.ld;\n";
  {
    int at = -1;
    for (int d = 0; d < kp.n_dyn; ++d) {
      const DEnt& e = pl->ents[pl->dyn_list[d]];
      const char* f[6] = {"px", "py", "vx", "vy", "rot", "w"};
      for (int k = 0; k < 6; ++k) {
        const int row = e.dyn_off + k;
        if (at < 0) outf(s, "  const float* pd = sa.dyn_in + (size_t)%d * ld_ + env;\n", row);
        else if (row == at + 1) s += "  pd += ld_;\n";
        else outf(s, "  pd = sa.dyn_in + (size_t)%d * ld_ + env;\n", row);
        at = row;
        outf(s, "  st.%s[%d] = *pd;\n", f[k], d);
      }
    }
  }
  int at_s = -1;
  for (size_t e = 0; e < E; ++e)
    if (stat_slot[e] >= 0) {
      const char* f[3] = {"sx", "sy", "srot"};
      for (int k = 0; k < 3; ++k) {
        const int row = pl->ents[e].stat_off + k;
        if (at_s < 0) outf(s, "  const float* ps = sa.stat + (size_t)%d * ld_ + env;\n", row);
        else if (row == at_s + 1) s += "  ps += ld_;\n";
        else outf(s, "  ps = sa.stat + (size_t)%d * ld_ + env;\n", row);
        at_s = row;
        if (big) outf(s, "  st.ss[%d * NT] = *ps;\n", srow[e] + k);
        else outf(s, "  st.%s[%d] = *ps;\n", f[k], stat_slot[e]);
      }
    }
  for (int j = 0; j < kp.J; ++j)
    if (jf_slot[j] >= 0) outf(s, "  st.jfr[%d] = sa.stat[(size_t)%d * sa.ld + env];\n", jf_slot[j], pl->items[j].fr_row - kp.dyn_rows);
  for (int a = 0; a < kp.A; ++a) {  // kernels.cuh load_forces
    outf(s, "  { const float* fp = aio.force[%d]; const float* tp = aio.torque[%d]; float fx = 0.f, fy = 0.f, tq = 0.f;\n", a, a);
    outf(s, "    if (fp) { fx = fp[env * aio.fstride[%d]]; fy = fp[env * aio.fstride[%d] + 1]; }\n", a, a);
    outf(s, "    if (tp) tq = tp[env * aio.tstride[%d]];\n    st.fx[%d] = fx; st.fy[%d] = fy; st.tq[%d] = tq; }\n", a, a, a, a);
  }
  s += "}\n";
  // ---- prologue (cos/sin cache, first accumulators) / integrate
  s += "VB_DI void world_begin(St& st) {\n";
  for (int t = 0; t < kp.n_trig; ++t) {
    const int e = pl->trig_list[t];
    if (big && stat_slot[e] >= 0 && !getenv("VMASB200_JIT_NOLOOP")) {
      // static entities of one shape whose tile rows are a fixed stride apart: one loop (10 inlined sincosf are 7 KB of code)
      int n = 1, stride = 0;
      for (; t + n < kp.n_trig; ++n) {
        const int e2 = pl->trig_list[t + n];
        if (stat_slot[e2] < 0 || pl->ents[e2].shape != pl->ents[e].shape) break;
        if (n == 1) stride = srow[e2] - srow[e];
        if (stride <= 0 || srow[e2] - srow[e] != n * stride) break;
      }
      if (n >= 3) {
        outf(s, "#pragma unroll 1\n  for (int j = 0; j < %d; ++j) { float* sj = st.ss + j * (%d * NT); trig_sp<%d>(sj + %d * NT, sj[%d * NT]); }\n",
             n, stride, pl->ents[e].shape == VMASB200_BOX ? 1 : 0, srow[e] + 3, srow[e] + 2);
        t += n - 1;
        continue;
      }
    }
    if (big && stat_slot[e] >= 0)
      outf(s, "  trig_s<%d, %d>(st, %s);\n", srow[e] + 3, pl->ents[e].shape == VMASB200_BOX ? 1 : 0, rotx(e).c_str());
    else
      outf(s, "  trig_r<%d, %d>(st, %s);\n", t, pl->ents[e].shape == VMASB200_BOX ? VMASB200_BOX : VMASB200_SPHERE, rotx(e).c_str());
  }
  for (int d = 0; d < kp.n_dyn; ++d) {
    const int e = pl->dyn_list[d];
    outf(s, "  begin_r<%d, %d>(VBE%d, st);\n", d, pl->ents[e].agent, e);
  }
  s += "}\nVB_DI void world_integrate(St& st, int s, bool more) {\n";
  for (int d = 0; d < kp.n_dyn; ++d) {
    const int e = pl->dyn_list[d];
    const bool box = pl->ents[e].shape == VMASB200_BOX && trig_slot[e] >= 0;
    outf(s, "  %s<%d, %d>(VBE%d, st, s);\n  if (more) begin_r<%d, %d>(VBE%d, st);\n", box ? "integrate_box_r" : "integrate_r", d,
         trig_slot[e], e, d, pl->ents[e].agent, e);
  }
  s += "}\n";
  // ---- joints, then pairs
  std::vector<char> run_done(runs.size(), 0);
  s += "VB_DI void world_items(St& st, const KParams& kp) {\n";
  for (int j = 0; j < kp.J; ++j) {
    const DItem& it = pl->items[j];
    outf(s, "  joint_r<%d, %d, %d, %d>(st, JointC{%s, %s, %s, %s, %s, %s, %d}, %s, %s, %s, %s, %s, %s);\n", it.acc, dyn_slot[it.a],
         dyn_slot[it.b], jf_slot[j], fl(it.dax).c_str(), fl(it.day).c_str(), fl(it.dbx).c_str(), fl(it.dby).c_str(),
         fl(it.dist).c_str(), fl(it.fixed_rot).c_str(), it.rotate, posx(it.a).c_str(), posx(it.b).c_str(), trigx(it.a).c_str(),
         trigx(it.b).c_str(), rotx(it.a).c_str(), rotx(it.b).c_str());
  }
  for (int p = 0; p < kp.P; ++p) {
    const DItem& it = pl->items[kp.J + p];
    const DBroad& b = pl->broad[p];
    const bool frame = !std::isinf(b.hx) || !std::isinf(b.hy);
    if (big && run_of[(size_t)p] >= 0) {
      const int ri0 = run_of[(size_t)p];
      const Run& r = runs[(size_t)ri0];
      if (p != r.p0 || run_done[(size_t)ri0]) continue;  // the whole run is emitted at its first pair
      // runs over the SAME static entities share one loop (joint_passage: the 13 wall boxes against the bar, then against the
      // two agents): the static rows are loaded once per block, one loop overhead; every run keeps its own bit accumulators
      std::vector<int> group{ri0};
      for (size_t rj = (size_t)ri0 + 1; rj < runs.size() && !getenv("VMASB200_JIT_NOFUSE"); ++rj) {
        const Run& o = runs[rj];
        if (!run_done[rj] && o.n == r.n && o.s0 == r.s0 && o.stride == r.stride && o.stat_is_a == r.stat_is_a) group.push_back((int)rj);
      }
      s += "  {\n";
      for (int ri : group) {
        const Run& g = runs[(size_t)ri];
        const int W = (g.n * g.k + 31) / 32;
        outf(s, "    uint32_t av%d[%d] = {}, as%d[%d] = {};  // pairs %d .. %d: %d static entities x %d dynamic ones\n", ri, W, ri, W, g.p0,
             g.p0 + g.n * g.k - 1, g.n, g.k);
      }
      outf(s, "#pragma unroll 1\n    for (int j = %d; j >= 0; --j) {\n      const float* sj = st.ss + j * (%d * NT);\n", r.n - 1, r.stride);
      outf(s, "      const V2 ps = mk(sj[%d * NT], sj[%d * NT]);\n", r.s0, r.s0 + 1);
      const bool packed = !getenv("VMASB200_JIT_NOF2");
      for (int ri : group) {
        const Run& g = runs[(size_t)ri];
        run_done[(size_t)ri] = 1;
        const int W = (g.n * g.k + 31) / 32, q0 = g.p0;
        outf(s, "      const uint32_t* th%d = RT%d + j * %d;\n      uint32_t v%d = 0u, sv%d = 0u;\n", ri, ri, g.kc * 4, ri, ri);
        for (int i = 0; i < g.k; ++i) {
          const DItem& iti = pl->items[kp.J + q0 + i];
          const bool fr = !std::isinf(pl->broad[q0 + i].hx) || !std::isinf(pl->broad[q0 + i].hy);
          if (packed && i + 1 < g.k) {  // rows i and i + 1 on the packed fp32 pipe when they share the frame (pair_l2)
            const DItem& itn = pl->items[kp.J + q0 + i + 1];
            const bool fr1 = !std::isinf(pl->broad[q0 + i + 1].hx) || !std::isinf(pl->broad[q0 + i + 1].hy);
            const bool same_frame = fr == fr1 && (!fr || g.stat_is_a || iti.a == itn.a);
            if (same_frame) {
              const int d0 = dyn_slot[g.stat_is_a ? iti.b : iti.a], d1 = dyn_slot[g.stat_is_a ? itn.b : itn.a];
              std::string qc = "0.f", qs = "0.f";
              if (fr) {
                qc.clear(); qs.clear();
                if (g.stat_is_a) { outf(qc, "sj[%d * NT]", g.s0 + 3); outf(qs, "sj[%d * NT]", g.s0 + 4); }
                else { outf(qc, "st.tc[%d]", trig_slot[iti.a]); outf(qs, "st.ts[%d]", trig_slot[iti.a]); }
              }
              const int c0 = g.cls[(size_t)i] * 4, c1 = g.cls[(size_t)i + 1] * 4;
              outf(s, "      pair_l2<%d, %s, %s>(v%d, sv%d, ps, mk(st.px[%d], st.py[%d]), mk(st.px[%d], st.py[%d]), %s, %s", i, fr ? "true" : "false",
                   g.stat_is_a ? "true" : "false", ri, ri, d0, d0, d1, d1, qc.c_str(), qs.c_str());
              for (int c : {c0, c0 + 1, c0 + 2, c0 + 3, c1, c1 + 1, c1 + 2, c1 + 3}) outf(s, ", __uint_as_float(th%d[%d])", ri, c);
              s += ");\n";
              ++i;
              continue;
            }
          }
          const int dynE = g.stat_is_a ? (int)iti.b : (int)iti.a;
          std::string pd;
          outf(pd, "mk(st.px[%d], st.py[%d])", dyn_slot[dynE], dyn_slot[dynE]);
          std::string qc = "0.f", qs = "0.f";
          if (fr) {  // the frame is a's (plan.cpp: a = line / box)
            if (g.stat_is_a) { qc.clear(); qs.clear(); outf(qc, "sj[%d * NT]", g.s0 + 3); outf(qs, "sj[%d * NT]", g.s0 + 4); }
            else { qc.clear(); qs.clear(); outf(qc, "st.tc[%d]", trig_slot[iti.a]); outf(qs, "st.ts[%d]", trig_slot[iti.a]); }
          }
          const int c4 = g.cls[(size_t)i] * 4;
          outf(s, "      pair_l<%d, %s>(v%d, sv%d, %s, %s, %s, %s, __uint_as_float(th%d[%d]), __uint_as_float(th%d[%d]), __uint_as_float(th%d[%d]), __uint_as_float(th%d[%d]));\n",
               i, fr ? "true" : "false", ri, ri, g.stat_is_a ? "ps" : pd.c_str(), g.stat_is_a ? pd.c_str() : "ps", qc.c_str(), qs.c_str(), ri, c4, ri,
               c4 + 1, ri, c4 + 2, ri, c4 + 3);
        }
        outf(s, "      run_push<%d, %d>(av%d, v%d); run_push<%d, %d>(as%d, sv%d);\n", W, g.k, ri, ri, W, g.k, ri, ri);
      }
      s += "    }\n";
      for (int ri : group) {
        const Run& g = runs[(size_t)ri];
        const int W = (g.n * g.k + 31) / 32;
        for (int tw = g.p0 >> 5; tw <= (g.p0 + g.n * g.k - 1) >> 5; ++tw)
          outf(s, "    st.verd[%d] |= run_word<%d, %d, %d>(av%d); st.sv[%d] |= run_word<%d, %d, %d>(as%d);\n", tw, W, g.p0, tw, ri, tw, W, g.p0, tw, ri);
      }
      s += "  }\n";
      continue;
    }
    if (big) {
      outf(s, "  pair_b<%d, %d, %s>(st, %s, %s, %s, %s, %s, %s, %s);\n", p >> 5, p & 31, frame ? "true" : "false", posx(it.a).c_str(),
           posx(it.b).c_str(), frame ? trigx(it.a).c_str() : "Rot{}", fl(b.t2_bp).c_str(), fl(b.t2_cull).c_str(), fl(b.hx).c_str(),
           fl(b.hy).c_str());
      continue;
    }
    outf(s, "  pair_r<%d, %d, %d, %d, %d, %d, %s, %s>(st, VBE%d, VBE%d, %s, %s, %s, %s, %s, %s, %s, %s);\n", it.type, it.acc,
         dyn_slot[it.a], dyn_slot[it.b], p >> 5, p & 31, frame ? "true" : "false", thin[p] ? "true" : "false", it.a, it.b,
         posx(it.a).c_str(), posx(it.b).c_str(), trigx(it.a).c_str(), trigx(it.b).c_str(), fl(b.t2_bp).c_str(),
         fl(b.t2_cull).c_str(), fl(b.hx).c_str(), fl(b.hy).c_str());
  }
  if (big && kp.P > 0) {
    // what the table-driven narrow phase reads and adds to goes to the tile, the sums come back (spec_helpers.cuh narrow_big)
    for (int d = 0; d < kp.n_dyn; ++d) {
      const int e = pl->dyn_list[d];
      outf(s, "  st.ss[%d * NT] = st.px[%d]; st.ss[%d * NT] = st.py[%d];\n", drow[e], d, drow[e] + 1, d);
      if (trig_slot[e] >= 0) {
        const int t = trig_slot[e];
        outf(s, "  st.ss[%d * NT] = st.tc[%d]; st.ss[%d * NT] = st.ts[%d];\n", drow[e] + 2, t, drow[e] + 3, t);
        if (pl->ents[e].shape == VMASB200_BOX) outf(s, "  st.ss[%d * NT] = st.tc2[%d]; st.ss[%d * NT] = st.ts2[%d];\n", drow[e] + 4, t, drow[e] + 5, t);
      }
      if (!park) outf(s, "  st.ss[%d * NT] = st.Fx[%d]; st.ss[%d * NT] = st.Fy[%d]; st.ss[%d * NT] = st.T[%d];\n", frow[e], d, frow[e] + 1, d, frow[e] + 2, d);
    }
    s += "  narrow_big(kp, st);\n";
    for (int d = 0; d < kp.n_dyn && !park; ++d) {
      const int e = pl->dyn_list[d];
      outf(s, "  st.Fx[%d] = st.ss[%d * NT]; st.Fy[%d] = st.ss[%d * NT]; st.T[%d] = st.ss[%d * NT];\n", d, frow[e], d, frow[e] + 1, d, frow[e] + 2);
    }
  }
  s += "}\n";
  // ---- store
  s += "VB_DI void world_store(const St& st, const StepArgs& sa, long long env) {\n";
  s += "  const size_t ld_ = (size_t)sa.ld;\n";
  int at_o = -1;
  for (int d = 0; d < kp.n_dyn; ++d) {
    const DEnt& e = pl->ents[pl->dyn_list[d]];
    const char* f[6] = {"px", "py", "vx", "vy", "rot", "w"};
    for (int k = 0; k < 6; ++k) {
      const int row = e.dyn_off + k;
      if (at_o < 0) outf(s, "  float* po = sa.dyn_out + (size_t)%d * ld_ + env;\n", row);
      else if (row == at_o + 1) s += "  po += ld_;\n";
      else outf(s, "  po = sa.dyn_out + (size_t)%d * ld_ + env;\n", row);
      at_o = row;
      outf(s, "  *po = st.%s[%d];\n", f[k], d);
    }
    if (e.act_off >= 0 && e.agent >= 0) {
      const char* g[3] = {"fx", "fy", "tq"};
      for (int k = 0; k < 3; ++k)
        outf(s, "  if (sa.act_out) sa.act_out[(size_t)%d * sa.ld + env] = st.%s[%d];\n", e.act_off + k, g[k], e.agent);
    }
  }
  s += "}\n";
  // ---- LIDAR staging: compact tile, 3 env rows (px py rot) per entity the epilogue reads; work rows =
  // [target masks][2 per sensor: cos/sin of its rotation][4 per non-sphere target: its cos/sin cache]
  lay->lid_ents.clear(); lay->lid_trig.clear();
  std::vector<int> lid_slot(E, -1);
  auto need = [&](int e) {
    if (lid_slot[e] < 0) { lid_slot[e] = (int)lay->lid_ents.size(); lay->lid_ents.push_back(e); }
  };
  for (const DSensor& sn : pl->sensors) {
    need(sn.entity);
    for (int q = 0; q < sn.n_targets; ++q) need(pl->targets[sn.tg_begin + q]);
  }
  for (int e : lay->lid_ents)
    if (pl->ents[e].shape != VMASB200_SPHERE && trig_slot[e] >= 0) lay->lid_trig.push_back(e);
  const int trig0 = kp.n_lmask_words + 2 * kp.n_sensors;
  lay->work_rows = trig0 + 4 * (int)lay->lid_trig.size();
  s += "VB_DI void world_stage_lidar(const St& st, float* col, float* wcol) {\n";
  for (size_t k = 0; k < lay->lid_ents.size(); ++k) {
    const int e = lay->lid_ents[k];
    outf(s, "  { const V2 p = %s; col[%d * TE] = p.x; col[%d * TE] = p.y; col[%d * TE] = %s; }\n", posx(e).c_str(), (int)(3 * k),
         (int)(3 * k + 1), (int)(3 * k + 2), rotx(e).c_str());
  }
  for (size_t k = 0; k < lay->lid_trig.size(); ++k) {
    const int t = trig_slot[lay->lid_trig[k]], r = trig0 + 4 * (int)k;
    outf(s, "  wcol[%d * TE] = st.tc[%d]; wcol[%d * TE] = st.ts[%d]; wcol[%d * TE] = st.tc2[%d]; wcol[%d * TE] = st.ts2[%d];\n", r, t,
         r + 1, t, r + 2, t, r + 3, t);
  }
  s += "}\n";
  // ---- LIDAR epilogue with the sensors as constants (spec_kernel.cuh lidar_world) when every sensor's target mask is
  // one word and the queue entries fit 16 bits; else the generic epilogue runs on the compact tile.
  //   pair entry  col | p << CB             p: index of (sensor, target) in the plan's target list
  //   ray entry   col | k << CB | r << (CB + KB)
  // both queues live in the tile's queue + rayq regions ((NS + R) * TE words): NPT * TE + R * TE 16-bit entries.
  auto bits = [](int n) { int b = 0; while ((1 << b) < n) ++b; return b; };
  const int NPT = (int)pl->targets.size(), R = kp.lidar_rows, NS = kp.n_sensors;
  const int CB = bits(te), KB = bits(std::max(NS, 1)), RB = bits(std::max(kp.max_rays, 1)), PB = bits(std::max(NPT, 1));
  bool lc = lidar;
  for (const DSensor& sn : pl->sensors) lc = lc && sn.n_tw == 1 && sn.n_targets >= 1 && sn.n_rays >= 1;
  lc = lc && (te & (te - 1)) == 0 && CB + KB + RB <= 16 && CB + PB <= 16;
  outf(s, "constexpr bool LIDAR_COMPILED = %s;\n", lc ? "true" : "false");
  lay->smem_rows = 0;
  if (lc) {
    // shared-memory rows of the compiled epilogue: [env 3 per staged entity][work][lid R][queues: (NPT + R) 16-bit entries per env]
    // DIRECT: no readings tile.  The rows are filled with max_range in global memory at the START of the kernel (by the
    // helper warps while the physics runs) and pass 3 stores its few hits over them; what is left of the tile's R rows is
    // one word per sensor ("this ray is queued" bits).  Needs at most 32 rays per sensor.
    // Measured (navigation): 128.6 -> 125.8 us at 1 M envs (-> 117.9 us with the occupancy the smaller tile allows), but
    // 10.8 -> 11.1 us at 32 768 envs: the one-wave 32-env tiles keep the readings tile.
    bool direct = te >= 64;
    if (const char* e = getenv("VMASB200_JIT_LIDDIRECT")) direct = atoi(e) != 0;  // experiments
    for (const DSensor& sn : pl->sensors) direct = direct && sn.n_rays <= 32;
    const int l_env = 0, l_work = l_env + 3 * (int)lay->lid_ents.size(), l_lid = l_work + lay->work_rows, l_q = l_lid + (direct ? NS : R);
    lay->smem_rows = l_q + (NPT + R + 1) / 2;
    outf(s, "constexpr bool DIRECT_LID = %s;\n", direct ? "true" : "false");
    outf(s, "constexpr int L_ENV = %d, L_WORK = %d, L_LID = %d, L_Q = %d;\n", l_env, l_work, l_lid, l_q);
    outf(s, "constexpr int NPT = %d, CB = %d, KB = %d, RAYS_TOTAL = %d;\n", NPT, CB, KB, R);
    // per (sensor, target) tables; a column whose entries are all equal becomes a literal
    struct Col { const char* name; const char* type; std::vector<std::string> v; };
    std::vector<Col> cols = {{"p_es", "int", {}}, {"p_cr", "int", {}}, {"p_ts", "int", {}}, {"p_out", "int", {}}, {"p_nr", "int", {}},
                             {"p_k", "int", {}}, {"p_lat", "float", {}}, {"p_along", "float", {}}, {"p_slk", "float", {}}};
    for (int k = 0; k < NS; ++k) {
      const DSensor& sn = pl->sensors[k];
      for (int q = 0; q < sn.n_targets; ++q) {
        const float4 tc = pl->tcull[sn.tg_begin + q];
        const int vals[6] = {3 * lid_slot[sn.entity], kp.n_lmask_words + 2 * k, 3 * lid_slot[pl->targets[sn.tg_begin + q]], sn.out_off, sn.n_rays, k};
        for (int c = 0; c < 6; ++c) cols[c].v.push_back(std::to_string(vals[c]));
        cols[6].v.push_back(fl(tc.y)); cols[7].v.push_back(fl(tc.z)); cols[8].v.push_back(fl(tc.w));
      }
    }
    std::vector<Col> scols = {{"k_es", "int", {}}, {"k_mrow", "int", {}}, {"k_tg0", "int", {}}, {"k_out", "int", {}}, {"k_range", "float", {}}};
    std::vector<Col> tcols = {{"p_shape", "int", {}}, {"p_trig", "int", {}}, {"p_radius", "float", {}}, {"p_length", "float", {}}, {"p_width", "float", {}}};
    for (int k = 0; k < NS; ++k) {
      const DSensor& sn = pl->sensors[k];
      const int vals[4] = {3 * lid_slot[sn.entity], sn.tw_off - kp.row_lmask, sn.tg_begin, sn.out_off};
      for (int c = 0; c < 4; ++c) scols[c].v.push_back(std::to_string(vals[c]));
      scols[4].v.push_back(fl(sn.max_range));
      for (int q = 0; q < sn.n_targets; ++q) {
        const int tg = pl->targets[sn.tg_begin + q];
        const DEnt& en = pl->ents[tg];
        int trow = 0;
        for (size_t i = 0; i < lay->lid_trig.size(); ++i)
          if (lay->lid_trig[i] == tg) trow = trig0 + 4 * (int)i;
        tcols[0].v.push_back(std::to_string(en.shape)); tcols[1].v.push_back(std::to_string(trow));
        tcols[2].v.push_back(fl(en.radius)); tcols[3].v.push_back(fl(en.length)); tcols[4].v.push_back(fl(en.width));
      }
    }
    auto emit_cols = [&](const std::vector<Col>& cs) {
      for (const Col& c : cs) {
        bool same = true;
        for (const std::string& v : c.v) same = same && v == c.v[0];
        if (same) {
          outf(s, "VB_DI %s %s(int) { return %s; }\n", c.type, c.name, c.v[0].c_str());
        } else {
          const bool f = c.type[0] == 'f';
          outf(s, "__constant__ %s T_%s[%d] = {", f ? "float" : "unsigned short", c.name, (int)c.v.size());
          for (size_t i = 0; i < c.v.size(); ++i) outf(s, "%s%s", i ? ", " : "", c.v[i].c_str());
          outf(s, "};\nVB_DI %s %s(int p) { return T_%s[p]; }\n", c.type, c.name, c.name);
        }
      }
    };
    emit_cols(cols); emit_cols(scols); emit_cols(tcols);
    // ray directions: when every sensor has the same rays the sweep is unrolled over a constant-bank table
    bool shared_rays = true;
    const DSensor& s0 = pl->sensors[0];
    for (const DSensor& sn : pl->sensors) {
      shared_rays = shared_rays && sn.n_rays == s0.n_rays;
      for (int r = 0; shared_rays && r < sn.n_rays; ++r)
        shared_rays = pl->ray_cs[sn.out_off + r].x == pl->ray_cs[s0.out_off + r].x && pl->ray_cs[sn.out_off + r].y == pl->ray_cs[s0.out_off + r].y;
    }
    shared_rays = shared_rays && s0.n_rays <= 64;
    outf(s, "constexpr int SHARED_RAYS = %d;\n__constant__ float2 T_raycs[%d] = {", shared_rays ? s0.n_rays : 0, shared_rays ? s0.n_rays : 1);
    for (int r = 0; r < (shared_rays ? s0.n_rays : 1); ++r)
      outf(s, "%s{%s, %s}", r ? ", " : "", fl(pl->ray_cs[s0.out_off + r].x).c_str(), fl(pl->ray_cs[s0.out_off + r].y).c_str());
    s += "};\n";
    // init: rows of one sensor share its max_range
    s += "VB_DI void world_lidar_init(float* lidS, int tid) {\n";
    for (size_t k = 0; k < pl->sensors.size();) {  // consecutive sensors of equal range share one fill
      size_t k2 = k;
      int rows = 0;
      while (k2 < pl->sensors.size() && pl->sensors[k2].max_range == pl->sensors[k].max_range) rows += pl->sensors[k2++].n_rays;
      const std::string v = fl(pl->sensors[k].max_range);
      outf(s, "  { const float4 v = make_float4(%s, %s, %s, %s); float4* p = reinterpret_cast<float4*>(lidS + %d * TE);\n"
              "#pragma unroll\n    for (int i = 0; i < (%d * TE / 4 + NT - 1) / NT; ++i) if (i * NT + tid < %d * TE / 4) p[i * NT + tid] = v; }\n",
           v.c_str(), v.c_str(), v.c_str(), v.c_str(), pl->sensors[k].out_off, rows, rows);
      k = k2;
    }
    s += "}\n";
    // DIRECT: the same fill in global memory (g = lidar_out + env0), by threads t of nth; vec: 16-byte stores are aligned
    s += "VB_DI void world_lidar_fill(float* g, size_t ld, int n_cols, int t, int nth, bool vec) {\n";
    for (size_t k = 0; k < pl->sensors.size();) {
      size_t k2 = k;
      int rows = 0;
      while (k2 < pl->sensors.size() && pl->sensors[k2].max_range == pl->sensors[k].max_range) rows += pl->sensors[k2++].n_rays;
      const std::string v = fl(pl->sensors[k].max_range);
      outf(s, "  if (vec) { const float4 v = make_float4(%s, %s, %s, %s);\n"
              "    for (int i = t; i < %d * (TE / 4); i += nth) { const int r = i / (TE / 4), c = (i %% (TE / 4)) * 4;\n"
              "      if (c < n_cols) *reinterpret_cast<float4*>(g + (size_t)(%d + r) * ld + c) = v; }\n"
              "  } else {\n    for (int i = t; i < %d * TE; i += nth) { const int r = i / TE, c = i %% TE; if (c < n_cols) g[(size_t)(%d + r) * ld + c] = %s; } }\n",
           v.c_str(), v.c_str(), v.c_str(), v.c_str(), rows, pl->sensors[k].out_off, rows, pl->sensors[k].out_off, v.c_str());
      k = k2;
    }
    s += "}\n";
    // pass 1: reach masks from the registers; every (sensor, target) within reach becomes one entry of the pair queue
    s += "VB_DI void world_lidar_pass1(const St& st, bool ok, int col, float* wcol, unsigned short* pq, int* qn) {\n  int n = 0;\n";
    for (int k = 0; k < NS; ++k) {
      const DSensor& sn = pl->sensors[k];
      const int mrow = sn.tw_off - kp.row_lmask;
      outf(s, "  uint32_t m%d = 0u;\n  { const V2 o = %s;\n", k, posx(sn.entity).c_str());
      for (int q = 0; q < sn.n_targets; ++q) {
        const int tg = pl->targets[sn.tg_begin + q];
        outf(s, "    { const V2 p = %s; const float dx = o.x - p.x, dy = o.y - p.y; if (!(fmaf(dy, dy, dx * dx) > %s)) m%d |= %uu; }\n",
             posx(tg).c_str(), fl(pl->tcull[sn.tg_begin + q].x).c_str(), k, 1u << q);
      }
      outf(s, "    wcol[%d * TE] = __uint_as_float(m%d);\n    if (!ok) m%d = 0u;\n", mrow, k, k);
      outf(s, "#ifndef VB_AB_SINCOS_P2\n    if (m%d) { float sn_, cs_; sincosf(%s, &sn_, &cs_); wcol[%d * TE] = cs_; wcol[%d * TE] = sn_; }\n#endif\n", k,
           rotx(sn.entity).c_str(), kp.n_lmask_words + 2 * k, kp.n_lmask_words + 2 * k + 1);
      outf(s, "    n += __popc(m%d); }\n", k);
    }
    s += "  if (n) {\n    int b = atomicAdd(qn, n);\n";
    for (int k = 0; k < NS; ++k) {
      const DSensor& sn = pl->sensors[k];
      for (int q = 0; q < sn.n_targets; ++q)
        outf(s, "    if (m%d & %uu) pq[b++] = (unsigned short)(col | (%d << CB));\n", k, 1u << q, sn.tg_begin + q);
    }
    s += "  }\n}\n";
  } else {
    s += "constexpr bool DIRECT_LID = false;\nVB_DI void world_lidar_fill(float*, size_t, int, int, int, bool) {}\n";
    s += "constexpr int L_ENV = 0, L_WORK = 0, L_LID = 0, L_Q = 0;\nconstexpr int NPT = 1, CB = 0, KB = 0, RAYS_TOTAL = 1, SHARED_RAYS = 0;\n__constant__ float2 T_raycs[1] = {{0.f, 0.f}};\n"
         "VB_DI int p_es(int) { return 0; }\nVB_DI int p_cr(int) { return 0; }\nVB_DI int p_ts(int) { return 0; }\n"
         "VB_DI int p_out(int) { return 0; }\nVB_DI int p_nr(int) { return 0; }\nVB_DI int p_k(int) { return 0; }\n"
         "VB_DI float p_lat(int) { return 0.f; }\nVB_DI float p_along(int) { return 0.f; }\nVB_DI float p_slk(int) { return 0.f; }\n"
         "VB_DI int k_es(int) { return 0; }\nVB_DI int k_mrow(int) { return 0; }\nVB_DI int k_tg0(int) { return 0; }\n"
         "VB_DI int k_out(int) { return 0; }\nVB_DI float k_range(int) { return 0.f; }\nVB_DI int p_shape(int) { return 0; }\n"
         "VB_DI int p_trig(int) { return 0; }\nVB_DI float p_radius(int) { return 0.f; }\nVB_DI float p_length(int) { return 0.f; }\n"
         "VB_DI float p_width(int) { return 0.f; }\n"
         "VB_DI void world_lidar_init(float*, int) {}\n"
         "VB_DI void world_lidar_pass1(const St&, bool, int, float*, unsigned short*, int*) {}\n";
  }
  s += "} }\n#include \"spec_kernel.cuh\"\n";
  return s;
}

// ------------------------------------------------------------------------------------------------ compile + cache
struct JitKernel {
  void* module = nullptr;
  void* func = nullptr;
  int te = 0, nt = 0, smem_bytes = 0;
  KParams kp{};   // the epilogue's view: compact tile rows, this plan's device tables
  KParams kpg{};  // the generic view (kernels.cuh repair_kernel: 32-env tile, tables in global memory)
  bool thin_possible = false;
  bool failed = false;
  void* d_tables = nullptr;  // BIG worlds: DEnt[E] + DItem[n_items] with the rows of the compiled kernel's tile (narrow_big)
};
struct JitPlanState {
  std::map<int, JitKernel> by_te;
  std::map<int, std::string> cubin;  // compiled (or read from the cache) but not yet loaded
};
static std::map<const vmasb200_plan*, JitPlanState> g_jit_plans;

static uint64_t fnv1a(const std::string& s, uint64_t h = 1469598103934665603ull) {
  for (unsigned char c : s) { h ^= c; h *= 1099511628211ull; }
  return h;
}
static std::string cache_dir() {
  if (const char* e = getenv("VMASB200_JIT_CACHE")) return e;
  Dl_info info;
  if (dladdr(reinterpret_cast<void*>(&cache_dir), &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    const size_t k = p.rfind('/');
    return (k == std::string::npos ? std::string(".") : p.substr(0, k)) + "/_jit";
  }
  return "_jit";
}

static const char* kShimStdint =
    "#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;\n"
    "typedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;\n"
    "typedef unsigned long uintptr_t;\n";
static const char* kShimMathConst =
    "#pragma once\n#define CUDART_INF_F __int_as_float(0x7f800000)\n#define CUDART_NAN_F __int_as_float(0x7fffffff)\n";

// cubin of the world-compiled kernel of (pl, te): from the disk cache, else NVRTC.  Empty string + message on failure.
static std::string jit_compile(const vmasb200_plan* pl, int te, int nt, JitLayout* lay, std::string* err) {
  const std::string src = generate(pl, te, nt, lay);
  if (const char* dump = getenv("VMASB200_JIT_DUMP")) {  // the generated source, for profiles (ncu line numbers) and docs
    std::ofstream f(dump);
    f << src;
  }
  // -default-device: the C prototypes of include/vmasb200.h carry no execution space
  std::vector<const char*> opts = {"--gpu-architecture=sm_100a", "--std=c++17", "--fmad=false", "-lineinfo", "-default-device"};
  // VMASB200_JIT_DEFINES="-DVB_X=1 -DVB_Y": extra options for A/B measurements of the kernel sources (part of the cache key)
  std::vector<std::string> extra;
  if (const char* d = getenv("VMASB200_JIT_DEFINES")) {
    std::stringstream ss(d);
    std::string w;
    while (ss >> w) extra.push_back(w);
  }
  for (const std::string& w : extra) opts.push_back(w.c_str());
  uint64_t h = fnv1a(src);
  for (int i = 0; i < kJitNumSources; ++i) h = fnv1a(kJitSrcText[i], h);
  for (const char* o : opts) h = fnv1a(o, h);
  char name[64];
  snprintf(name, sizeof name, "/world_%016llx.cubin", (unsigned long long)h);
  const std::string dir = cache_dir(), path = dir + name;
  {
    std::ifstream f(path, std::ios::binary);
    if (f) {
      std::stringstream ss;
      ss << f.rdbuf();
      if (ss.str().size() > 64) return ss.str();
    }
  }
  if (!load_nvrtc()) { *err = "libnvrtc.so.12 not found"; return ""; }
  std::vector<const char*> hn, hs;
  for (int i = 0; i < kJitNumSources; ++i) { hn.push_back(kJitSrcNames[i]); hs.push_back(kJitSrcText[i]); }
  hn.push_back("stdint.h"); hs.push_back(kShimStdint);
  hn.push_back("math_constants.h"); hs.push_back(kShimMathConst);
  hn.push_back("cuda_runtime.h"); hs.push_back("#pragma once\n");
  void* prog = nullptr;
  if (g_nvrtc.CreateProgram(&prog, src.c_str(), "vmasb200_world.cu", (int)hn.size(), hs.data(), hn.data()) != 0) {
    *err = "nvrtcCreateProgram failed";
    return "";
  }
  const int rc = g_nvrtc.CompileProgram(prog, (int)opts.size(), opts.data());
  std::string out;
  if (rc != 0) {
    size_t n = 0;
    g_nvrtc.GetProgramLogSize(prog, &n);
    std::string log(n, '\0');
    if (n) g_nvrtc.GetProgramLog(prog, &log[0]);
    *err = "NVRTC: " + log.substr(0, 1200);
  } else {
    size_t n = 0;
    g_nvrtc.GetCUBINSize(prog, &n);
    out.resize(n);
    if (n) g_nvrtc.GetCUBIN(prog, &out[0]);
  }
  g_nvrtc.DestroyProgram(&prog);
  if (!out.empty()) {  // best effort: a read-only tree just recompiles next time
    mkdir(dir.c_str(), 0755);
    const std::string tmp = path + ".tmp" + std::to_string((long long)getpid());
    std::ofstream f(tmp, std::ios::binary);
    if (f) {
      f.write(out.data(), (std::streamsize)out.size());
      f.close();
      if (rename(tmp.c_str(), path.c_str()) != 0) remove(tmp.c_str());
    }
  }
  return out;
}

// CTA shape of the compiled kernel for a batch of n envs: one warp of envs per CTA while that fills the GPU in about
// one wave (shortest critical path), else 128; with LIDAR the CTA always has 128 threads (kernels.cuh kThreads).
static int jit_tile(const vmasb200_plan* pl, long long n) {
  if (pl->te_fixed) return pl->te_fixed;
  if (const char* e = getenv("VMASB200_JIT_TE")) {  // experiments (tools/ab_big.sh)
    const int v = atoi(e);
    if (v == 32 || v == 64 || v == 128) return v;
  }
  if (n <= 48 * 1024) return 32;
  return pl->kp.n_rays ? 64 : 128;
}

// threads of a LIDAR world's CTA: the TE physics threads plus helper warps for the epilogue's passes
static int jit_lidar_threads(int te) {
  if (const char* e = getenv("VMASB200_JIT_NT")) {  // experiments (tools/ab_jit.sh): a multiple of 32, >= te
    const int v = atoi(e);
    if (v >= te && v % 32 == 0 && v <= 1024) return v;
  }
  // measured (navigation, tools/ab_jit.sh): 32-env tiles 128 threads (64: 14.4 us against 12.7 us at 32 768 envs);
  // 64-env tiles 96 (141.5 us at 1 M envs; 128: 145.0, 64: 148.1, 256: 250.7)
  return te == 64 ? 96 : std::max(te, kThreads);
}
static void jit_kparams(const vmasb200_plan* pl, int te, const JitLayout& lay, KParams* out) {
  KParams k = pl->kp;
  k.TE = te; k.NT = kThreads;
  k.dyn_rows = 3 * (int)lay.lid_ents.size(); k.stat_rows = 0;
  k.work_rows = lay.work_rows;
  k.row_lmask = 0; k.row_strig = pl->kp.n_lmask_words;
  *out = k;
}
// the generic kernel's view for the last CTA's repair: its 32-env tile with 4-member teams, tables in global memory
static void jit_kparams_generic(const vmasb200_plan* pl, KParams* out) {
  KParams k = pl->kp;
  k.TE = 32; k.NT = kThreads;
  *out = k;
}

// The compiled kernel of (pl, tile) ready to launch, or nullptr: not eligible / switched off / no NVRTC / no driver API
// (reason in *why when given).  compile_only: stop after the cubin exists (no CUDA context needed).
static JitKernel* jit_get(vmasb200_plan* pl, int te, bool compile_only, std::string* why) {
  std::lock_guard<std::mutex> lock(g_jit_mu);
  auto say = [&](const std::string& m) { if (why) *why = m; return (JitKernel*)nullptr; };
  if (!jit_on()) return say("switched off (VMASB200_JIT=0 / vmasb200_jit_enable(0))");
  if (!jit_eligible(pl, te)) return say("world not covered by the generator, or the generic kernel is the measured choice for it at this batch size");
  JitPlanState& ps = g_jit_plans[pl];
  JitKernel& jk = ps.by_te[te];
  if (jk.failed) return say("an earlier attempt failed");
  if (jk.func) return &jk;
  const bool lidar = pl->kp.n_rays > 0;
  JitLayout lay;
  const int nt = lidar ? jit_lidar_threads(te) : te;
  std::string err;
  std::string& cubin = ps.cubin[te];
  if (cubin.empty()) cubin = jit_compile(pl, te, nt, &lay, &err);
  else generate(pl, te, nt, &lay);
  jk.thin_possible = lay.thin_possible;
  if (cubin.empty()) {
    jk.failed = true;
    fprintf(stderr, "libvmasb200: world-compiled kernel unavailable, using the generic kernel (%s)\n", err.c_str());
    return say(err);
  }
  jk.te = te; jk.nt = nt;
  if (compile_only) return say("compiled");
  if (!load_driver()) { jk.failed = true; return say("libcuda.so.1 not found"); }
  jit_kparams(pl, te, lay, &jk.kp);
  jit_kparams_generic(pl, &jk.kpg);
  jk.smem_bytes = !lidar ? lay.stat_tile_rows * nt * 4 + (lay.coop ? (nt / 32) * (64 / 2 + 4 * 64) * 4 : 0) : lay.smem_rows ? lay.smem_rows * te * 4 : smem_map(jk.kp).total_bytes;
  auto fail = [&](const char* what, int rc) {
    jk.failed = true;
    if (jk.module) { g_drv.ModuleUnload(jk.module); jk.module = nullptr; }
    fprintf(stderr, "libvmasb200: world-compiled kernel unavailable, using the generic kernel (%s: CUresult %d)\n", what, rc);
    return say(what);
  };
  int rc;
  if ((rc = g_drv.ModuleLoadData(&jk.module, cubin.data()))) return fail("cuModuleLoadData", rc);
  void* fn = nullptr;
  if ((rc = g_drv.ModuleGetFunction(&fn, jk.module, "vmasb200_step_world"))) return fail("cuModuleGetFunction", rc);
  if (lidar) {
    // this module's own constant bank: the epilogue's tables with the compact tile's rows; uploaded once, never shared
    std::unique_ptr<CTab> t(new CTab());
    std::memset(t.get(), 0, sizeof(CTab));
    std::copy(pl->ents.begin(), pl->ents.end(), t->ents);
    std::copy(pl->sensors.begin(), pl->sensors.end(), t->sensors);
    for (size_t e = 0; e < pl->ents.size(); ++e) { t->ents[e].pos_row = -1; t->ents[e].rot_row = -1; t->ents[e].trig_row = -1; }
    for (size_t k = 0; k < lay.lid_ents.size(); ++k) {
      t->ents[lay.lid_ents[k]].pos_row = 3 * (int)k;
      t->ents[lay.lid_ents[k]].rot_row = 3 * (int)k + 2;
    }
    for (size_t k = 0; k < lay.lid_trig.size(); ++k)
      t->ents[lay.lid_trig[k]].trig_row = pl->kp.n_lmask_words + 2 * pl->kp.n_sensors + 4 * (int)k;
    for (size_t k = 0; k < pl->sensors.size(); ++k) t->sensors[k].tw_off = pl->sensors[k].tw_off - pl->kp.row_lmask;
    unsigned long long dptr = 0;
    size_t bytes = 0;
    if ((rc = g_drv.ModuleGetGlobal(&dptr, &bytes, jk.module, "_ZN2vb5c_tabE"))) return fail("cuModuleGetGlobal(c_tab)", rc);
    if (bytes != sizeof(CTab)) return fail("c_tab size mismatch", (int)bytes);
    if ((rc = g_drv.MemcpyHtoD(dptr, t.get(), sizeof(CTab)))) return fail("cuMemcpyHtoD(c_tab)", rc);
  }
  if (!lidar && lay.stat_tile_rows > 0) {
    // the generic narrow phase runs on the compiled kernel's tile: entity / item tables with ITS rows, in global memory
    std::vector<DEnt> ents = pl->ents;
    std::vector<DItem> items = pl->items;
    for (size_t e = 0; e < ents.size(); ++e) {
      ents[e].pos_row = lay.tile_pos[e]; ents[e].trig_row = lay.tile_trig[e]; ents[e].rot_row = lay.tile_rot[e];
      ents[e].fin_row = lay.tile_fin[e];
    }
    for (DItem& it : items) { it.fin_a = lay.tile_fin[it.a]; it.fin_b = lay.tile_fin[it.b]; it.own = 0u; }
    const size_t be = ents.size() * sizeof(DEnt), bi = items.size() * sizeof(DItem);
    if (cudaMalloc(&jk.d_tables, be + bi) != cudaSuccess) return fail("cudaMalloc(tile tables)", -1);
    cudaMemcpy(jk.d_tables, ents.data(), be, cudaMemcpyHostToDevice);
    cudaMemcpy(static_cast<char*>(jk.d_tables) + be, items.data(), bi, cudaMemcpyHostToDevice);
    jk.kp = pl->kp;
    jk.kp.TE = te; jk.kp.NT = nt;
    jk.kp.ents = static_cast<const DEnt*>(jk.d_tables);
    jk.kp.items = reinterpret_cast<const DItem*>(static_cast<char*>(jk.d_tables) + be);
  }
  if (jk.smem_bytes > 48 * 1024 &&
      (rc = g_drv.FuncSetAttribute(fn, 8 /* CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES */, jk.smem_bytes)))
    return fail("cuFuncSetAttribute", rc);
  jk.func = fn;
  cubin.clear();
  cubin.shrink_to_fit();
  return &jk;
}

static void jit_forget(const vmasb200_plan* pl) {
  std::lock_guard<std::mutex> lock(g_jit_mu);
  auto it = g_jit_plans.find(pl);
  if (it == g_jit_plans.end()) return;
  for (auto& kv : it->second.by_te) {
    if (kv.second.module && g_drv.ok) g_drv.ModuleUnload(kv.second.module);
    if (kv.second.d_tables) cudaFree(kv.second.d_tables);
  }
  g_jit_plans.erase(it);
}

}  // namespace vb
