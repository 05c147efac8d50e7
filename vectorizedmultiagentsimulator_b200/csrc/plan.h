// plan.h - host-side plan object behind the opaque vmasb200_plan of include/vmasb200.h.
#pragma once
#include <cuda_runtime.h>

#include <memory>
#include <string>
#include <vector>

#include "../../include/vmasb200.h"
#include "tables.h"

namespace vb {
struct CTab;
constexpr int kHostStreams = 4;      // chunk streams of the host-buffer entry point
constexpr int64_t kRecCap = 65536;  // tainted envs that get a verdict / thin-event record
}

struct vmasb200_plan {
  vb::KParams kp{};  // device pointers are filled by upload()
  vmasb200_layout layout{};
  // host copies of the tables
  std::vector<vb::DEnt> ents;
  std::vector<vb::DItem> items;
  std::vector<vb::DBroad> broad;
  std::vector<int32_t> dyn_list, trig_list, targets;
  std::vector<float> ray_angles;
  std::vector<float4> tcull;
  std::vector<float2> ray_cs;
  std::vector<vb::DSensor> sensors;
  std::vector<vmasb200_pair> pairs;
  std::vector<int32_t> dyn_slot, stat_slot, act_slot, joint_slot, lidar_slot, agent_entities;
  int64_t jit_launches = 0;  // launches that took the world-compiled kernel (jit.h)
  int64_t jit_tail_launches = 0;  // 1-CTA repair_kernel launches behind them (worlds that can have thin contacts)
  int te_fixed = 0;          // envs per CTA asked for by the world description (0 = by batch size, see tile_for)
  uint64_t serial = 0;                      // identity for the constant-bank ownership check
  std::shared_ptr<vb::CTab> h_ctab;         // host image of the constant-bank tables (built on first use)
  void* d_ctab = nullptr;                   // its device copy: uploads to the constant bank are device-to-device
  // device side (lazy: nothing CUDA happens before the first compute call)
  bool uploaded = false;
  int device = -1;
  void* d_blob = nullptr;
  // [4][S][pair_words] bit arrays (flags, any_exact, any_spec, thin) + 4 counters, and the tainted-env scratch of
  // the speculative launch.  One step at a time per plan.
  uint32_t* d_bits = nullptr;
  int64_t u_cap = 0;
  int* d_ulist = nullptr;
  float* d_ubuf = nullptr;
  int* h_repairs = nullptr;    // mapped pinned host word: repairs done since plan creation (written by the kernel)
  uint32_t* d_urec = nullptr;  // per tainted env: verdict record + thin-event record ([S][pair_words] each)
  // staging of vmasb200_step_host
  int64_t host_cap = 0;        // row stride (floats) of the staging buffers
  int64_t res_n = -1;          // envs of the state / static rows the staging holds from the last step_host call (-1: none)
  bool res_stat = false;       // ... and whether the static rows were ever uploaded for that batch
  float *d_dyn = nullptr, *d_stat = nullptr, *d_force = nullptr, *d_torque = nullptr, *d_act = nullptr, *d_lidar = nullptr;
  cudaStream_t streams[vb::kHostStreams] = {};
};

namespace vb {
void set_error(const std::string& msg);
const char* last_error();
int build_plan(const vmasb200_world_desc* desc, vmasb200_plan* pl);  // pure host
int upload_plan(vmasb200_plan* pl);                                  // first CUDA touch
}  // namespace vb
