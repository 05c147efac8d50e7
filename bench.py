#!/usr/bin/env python
"""bench.py - env-steps/s of the batched physics step (World.step + LIDAR epilogue) on N B200s.

Contract: `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0 (see DESIGN.md "Measurement").
A "step" is one World.step (all substeps + the LIDAR epilogue) over the whole env batch.
`--impl reference` times the reference's own torch CPU path (oracle/_ref, the patched copy of the reference made by
oracle/make_ref.sh) when it is present, else the C oracle port, on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "env_steps_per_sec"
L2_BYTES = 126 * 1024 * 1024


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "reference-cuda", "dropin", "dropin-hooks", "dropin-graph", "queries"],
                    help="b200: the CUDA step through the C ABI (the driver's arm); reference: the reference's CPU path; "
                         "reference-cuda / dropin: the reference Environment on this GPU with its own torch World.step, "
                         "or with vectorizedmultiagentsimulator_b200.dropin enabled (reported comparison, needs oracle/_ref)")
    ap.add_argument("--scenario", default="navigation")
    ap.add_argument("--num-envs", type=int, default=32768, help="environments PER GPU (weak scaling)")
    ap.add_argument("--envs-per-cta", type=int, default=0, help="CTA tile override (0 = plan default)")
    ap.add_argument("--exact", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--static-io", type=int, default=0, help="dropin arms: step in place on persistent buffers")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_capture(scenario, n, compiled=False):
    """What the committed `ncu --set full` capture of this workload says about one launch of the step kernel
    (profiles/r1z_traffic.json, else the earlier r1w / r1 sets; written by tools/profile_collect.py):
    dram_bytes = dram__bytes_read.sum + dram__bytes_write.sum, warp_inst = smsp__inst_executed.sum.  Only a capture of
    the SAME kernel counts (world-compiled vmasb200_step_world or generic vb::step_kernel); {} when there is none."""
    for name in ("r1z_traffic.json", "r1w_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                cap = json.load(f).get(f"{scenario}@{n}", {})
        except (OSError, ValueError):
            continue
        if cap and ("vmasb200_step_world" in cap.get("kernel", "step_kernel")) == bool(compiled):
            return cap
    return {}


def alg_flops_per_env(plan, desc):
    """SURVEY 8(d) estimate of the step's algorithmic FP32 work per env-step (op lists of the reference; the kernel's
    culls execute far fewer): per substep 35 / 55 / 190 / 180 / 730 / 5000 per S-S / L-S / L-L / B-S / B-L / B-B pair,
    110 per joint, 35 per dynamic entity; plus 45 per LIDAR (ray, target)."""
    per_type = [35, 55, 190, 180, 730, 5000]
    pairs = sum(per_type[t] for t, _, _ in plan.pairs)
    rays = sum(len(l.angles) * len(l.targets) for l in desc.lidars)
    return plan.S * (pairs + 110 * len(desc.joints) + 35 * plan.layout.n_dyn) + 45 * rays


def cpu_port_rate(wl, seconds=8.0, n_sample=None, threads=0):
    """The C oracle (oracle/vmas_oracle.c: scalar port of the reference step, OpenMP over envs) timed on the host."""
    from oracle.oracle import Oracle

    n = min(wl.n_envs, n_sample or wl.n_envs)
    orc = Oracle(wl.desc, n_threads=threads)
    st = np.ascontiguousarray(wl.state0[:, :n].copy())
    jr = np.ascontiguousarray(wl.jrot0[:, :n]) if wl.jrot0.size else None
    f, t = wl.random_actions(4)
    f = [np.ascontiguousarray(x[:n].transpose(1, 0, 2)) for x in f]
    t = [np.ascontiguousarray(x[:n].transpose(1, 0)) for x in t]
    steps, t0 = 0, time.perf_counter()
    while True:
        orc.step(st, f[steps % 4].copy(), t[steps % 4].copy(), jr, want_masks=False)
        for k in range(len(wl.desc.lidars)):
            orc.lidar(st, k)
        steps += 1
        el = time.perf_counter() - t0
        if el >= seconds and steps >= 3:
            break
    cores = threads or (os.cpu_count() or 1)
    return n * steps / el, cores, f"{steps} steps x {n} envs of {wl.scenario} ({el:.1f} s)"


def run_reference(args, rank, world):
    """Reference arm: the reference's own CPU implementation of the path, on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from vectorizedmultiagentsimulator_b200.workload import Workload

    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    line = {"impl": "reference", "metric": METRIC, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic"}
    n = args.num_envs
    value = kind = sample = None
    cores = os.cpu_count() or 1
    if os.path.isdir(os.path.join(ref_dir, "vmas")):
        try:
            value, sample = _time_ref_torch(args, ref_dir, n)
            kind = "reference"
        except Exception as e:  # the copy did not import on this box: fall back to the port, say so
            sample = f"oracle/_ref failed ({type(e).__name__}: {e}); "
    if value is None:
        wl = Workload(args.scenario, n, seed=0)
        value, cores, s2 = cpu_port_rate(wl, seconds=max(2.0, 0.05 * args.steps))
        kind, sample = "port", (sample or "") + s2
    line.update({"value": value, "ms_per_step": 1e3 * n / value,
                 "config": {"workload": f"{args.scenario} num_envs={n} (host CPU, {kind})"},
                 "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": kind, "sample": sample},
                 "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(line), flush=True)


def _time_ref_torch(args, ref_dir, n):
    """World.step + Lidar.measure of the UNMODIFIED reference physics (oracle/_ref = reference copy with the one-line
    import fix of oracle/make_ref.sh) under torch CPU with all host threads."""
    import torch

    sys.path.insert(0, ref_dir)
    sys.path.insert(0, os.path.join(ROOT, "oracle", "gym_stub"))
    import vmas  # noqa

    from vectorizedmultiagentsimulator_b200.workload import Workload

    torch.set_num_threads(os.cpu_count() or 1)
    wl = Workload(args.scenario, 1, seed=0)  # fixture meta: reference scenario name, kwargs, substeps, action scale
    env = vmas.make_env(wl.meta["scenario"], num_envs=n, device="cpu", seed=0, continuous_actions=True, **wl.meta["kwargs"])
    world = env.world
    if world._substeps != wl.desc.substeps:  # "transport, substeps>1": imposed on the built world (note i)
        world._substeps, world._sub_dt = wl.desc.substeps, world._dt / wl.desc.substeps
    agents = list(world.agents)
    g = torch.Generator().manual_seed(1234)
    scale = [float(v) for v in wl.f_scale]

    def one():
        for a, sc in zip(agents, scale):
            a.state.force = (torch.rand(n, 2, generator=g) * 2 - 1) * sc
        world.step()
        for a in agents:
            for s in a.sensors or []:
                s.measure()

    for _ in range(args.warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one()
    el = time.perf_counter() - t0
    return n * args.steps / el, f"{args.steps} steps x {n} envs, reference torch {torch.__version__} CPU World.step + Lidar.measure, {torch.get_num_threads()} threads ({el:.1f} s)"


def run_b200(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    from vectorizedmultiagentsimulator_b200.plan import StepPlan
    from vectorizedmultiagentsimulator_b200.workload import Workload

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the CUDA step has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.num_envs
    wl = Workload(args.scenario, n, seed=rank)
    wl.desc.envs_per_cta = args.envs_per_cta
    plan = StepPlan(wl.desc)
    A, K, W = plan.A, args.steps, args.warmup

    # R independent env batches visited round-robin so that a step's inputs come from HBM, not from L2
    ld = plan.ld_for(n)
    per_batch = 4 * ld * (2 * plan.dyn_rows + plan.stat_rows + plan.act_rows + plan.lidar_rows + 3 * A)
    R = max(2, min(64, -(-2 * L2_BYTES // max(per_batch, 1))))
    dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
    f_all, t_all = wl.random_actions(R)
    stream = torch.cuda.current_stream().cuda_stream
    ios, keep = [], []
    for r in range(R):
        dyn = torch.from_numpy(dyn0).to(dev)
        stat = torch.from_numpy(stat0).to(dev) if plan.stat_rows else None
        fr = torch.from_numpy(f_all[r]).to(dev)  # [N, A, 2]
        tr = torch.from_numpy(t_all[r]).to(dev)  # [N, A]
        forces = [fr[:, a, :] for a in range(A)]
        torques = [tr[:, a:a + 1] for a in range(A)]
        act = torch.zeros(plan.act_rows, ld, device=dev) if plan.act_rows else None
        lid = torch.zeros(plan.lidar_rows, ld, device=dev) if plan.lidar_rows else None
        ios.append(plan.make_io(dyn, dyn, stat, forces, torques, act, lid, None, n, ld, bool(args.exact), stream))
        keep.append((dyn, stat, fr, tr, act, lid))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(W, 3) * R):  # every batch gets >= 3 warm-up steps
        plan.step_io(ios[i % R])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local_rank)
    barrier()
    e0.record()
    for i in range(K):
        plan.step_io(ios[i % R])
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    repairs = plan.sync_check(stream)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * n * K / (ms * 1e-3)

    # L2-resident variant (one batch stepped in place, what a rollout does): reported beside, not as `value`
    for _ in range(3):
        plan.step_io(ios[0])
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        plan.step_io(ios[0])
    e1.record()
    torch.cuda.synchronize()
    ms_hot = e0.elapsed_time(e1)

    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(plan, wl, n, K, world, dev)
    if rank == 0:
        peak, peak_src = peaks()
        jit_state, jit_launches = plan.jit(n)
        compiled = jit_launches > 0   # the launches took the world-compiled kernel (csrc/jit.h), not the generic one
        cap = ncu_capture(args.scenario, n, compiled)
        alg = plan.alg_bytes_per_env * n
        achieved = alg / (ms / K * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.scenario} num_envs={n} per GPU, substeps={plan.S}, {len(wl.desc.lidars)} lidars",
                       "l2": f"{R} env batches visited round-robin, working set {R * per_batch / 2**20:.0f} MiB > L2 126 MiB",
                       "envs_per_cta": plan.tile_for(n),
                       "layout": ("SoA rows [field][env]; world-compiled kernel (NVRTC): one thread per env, state in registers"
                                  if compiled else
                                  "SoA rows [field][env]; 128 threads per CTA, 128 / envs_per_cta of them share an env"),
                       "exact": int(args.exact), "speculation_repairs": repairs,
                       "l2_resident_ms_per_step": ms_hot / K},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": cap.get("dram_bytes"), "peak_source": peak_src, "alg_bytes_per_env": plan.alg_bytes_per_env,
                         "kernel": "vmasb200_step_world (world-compiled, NVRTC)" if compiled else "vb::step_kernel"},
            # one launch per step; a compiled world that can have thin contacts is followed by its 1-CTA repair_kernel
            "gpu_launches": K * (2 * plan.S if args.exact else (2 if compiled and plan.jit_tail_launches else 1)),
            "clocks": clocks,
        }
        # the step's roofline is the slower of its HBM bytes at peak bandwidth and its FP32 FLOPs at the
        # non-tensor FP32 peak (148 SMs x 128 lanes x 2 x max SM clock); both fractions are reported, `binding` names it
        flops = alg_flops_per_env(plan, wl.desc)
        fp32_peak = 148 * 128 * 2 * ((clocks or {}).get("sm_max_mhz") or 1965.0) * 1e6
        fp32_frac = flops * value / world / fp32_peak
        line["roofline"]["fp32"] = {"alg_flops_per_env": flops, "achieved_tflops": flops * value / world / 1e12,
                                    "peak_tflops": fp32_peak / 1e12, "frac": fp32_frac}
        line["roofline"]["binding"] = "fp32" if flops / fp32_peak > plan.alg_bytes_per_env / (peak * 1e9) else "hbm"
        if cap.get("warp_inst"):
            # the bound that actually binds: warp instructions of one launch (ncu) over the live launch
            # time, against 4 issue slots x 148 SMs at the SM clock sampled during the timed region
            mhz = (clocks or {}).get("sm_mhz") or 1965.0
            ips = cap["warp_inst"] / (ms / K * 1e-3)
            line["roofline"]["issue"] = {"warp_inst_per_launch": cap["warp_inst"], "achieved_ginst_s": ips / 1e9,
                                         "peak_ginst_s": 4 * 148 * mhz * 1e-3, "frac": ips / (4 * 148 * mhz * 1e6)}
        if e2e is not None:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_port_rate(wl, seconds=10.0)
            line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(plan, wl, n, K, world, dev):
    """Same metric through the C-ABI HOST-buffer call (vmasb200_step_host): pinned host records in, H2D, kernel,
    D2H of state / clamped actions / LIDAR, stream sync - every step."""
    import torch
    import torch.distributed as dist

    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()

    dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
    ld = dyn0.shape[1]
    f, t = wl.random_actions(1)
    keep = [pinned(dyn0), pinned(stat0), pinned(f[0]), pinned(t[0])]
    dyn, stat, fr, tq = (k[1] for k in keep)
    outs = [pinned(np.zeros((plan.dyn_rows, ld), np.float32)), pinned(np.zeros((max(plan.act_rows, 1), ld), np.float32)),
            pinned(np.zeros((max(plan.lidar_rows, 1), ld), np.float32))]
    out = (outs[0][1], outs[1][1][:plan.act_rows] if plan.act_rows else None,
           outs[2][1][:plan.lidar_rows] if plan.lidar_rows else None)
    for _ in range(3):
        plan.step_host(dyn, stat, fr, tq, exact=False, want_masks=False, out=out)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
        # the step's results land in pinned host memory; the next step starts from them (a rollout through the host API)
        plan.step_host(dyn, stat, fr, tq, exact=False, want_masks=False, out=out)
        dyn, out = out[0], (dyn, out[1], out[2])
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([el], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        el = float(tt.item())
    h2d = 4 * n * (plan.dyn_rows + plan.stat_rows + 3 * plan.A)
    d2h = 4 * n * (plan.dyn_rows + plan.act_rows + plan.lidar_rows)
    return {"value": world * n * K / el, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "api": "vmasb200_step_host (C ABI, pinned host SoA buffers, chunked over 4 streams)"}


def run_env_cuda(args, rank):
    """The reference's Environment on cuda:0, stepped with pre-generated random actions; World.step is timed on its own
    (CUDA events around the bound method) and Environment.step as a whole.  --impl dropin patches the world first."""
    if rank != 0:
        return
    import torch

    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "vmas")):
        print(json.dumps({"impl": args.impl, "unavailable": "oracle/_ref (reference install) not present"}))
        return
    sys.path.insert(0, ref_dir)
    sys.path.insert(0, os.path.join(ROOT, "oracle", "gym_stub"))
    import vmas

    from vectorizedmultiagentsimulator_b200.workload import Workload

    wl = Workload(args.scenario, 1, seed=0)
    n, K, W = args.num_envs, args.steps, max(args.warmup, 3)
    env = vmas.make_env(wl.meta["scenario"], num_envs=n, device="cuda", seed=0, continuous_actions=True, **wl.meta["kwargs"])
    world = env.world
    if world._substeps != wl.desc.substeps:
        world._substeps, world._sub_dt = wl.desc.substeps, world._dt / wl.desc.substeps
    hooks = None
    if args.impl in ("dropin", "dropin-hooks"):
        from vectorizedmultiagentsimulator_b200 import dropin

        # dropin-hooks: also the fused reward / observation / done launches of SURVEY 8 f1 (navigation, transport)
        hooks = dropin.enable(env, hooks=args.impl == "dropin-hooks", static_io=bool(args.static_io)).hooks
    torch.manual_seed(1234)
    acts = [[torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in env.agents] for _ in range(8)]
    inner = world.step
    ev = []

    def timed_step():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        inner()
        e1.record()
        ev.append((e0, e1))

    world.step = timed_step
    for i in range(W):
        env.step([a.clone() for a in acts[i % 8]])
    torch.cuda.synchronize()
    ev.clear()
    t0 = time.perf_counter()
    for i in range(K):
        env.step([a.clone() for a in acts[i % 8]])
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    ws_gpu = sum(a.elapsed_time(b) for a, b in ev) / K
    # host-side cost of World.step alone (launch-bound torch code is limited by it, not by the GPU)
    t1 = time.perf_counter()
    for i in range(K):
        inner()
    torch.cuda.synchronize()
    ws_wall = (time.perf_counter() - t1) / K * 1e3
    line = {"impl": args.impl, "metric": METRIC, "unit": "env-steps/s", "n_gpus": 1, "steps": K, "warmup": W,
            "config": {"workload": f"{args.scenario} num_envs={n}, reference Environment on cuda:0 (torch {torch.__version__})"},
            "world_step_ms_gpu_events": ws_gpu, "world_step_ms_wall": ws_wall, "env_step_ms_wall": el / K * 1e3,
            "world_step_env_steps_per_s": n / (ws_wall * 1e-3), "value": n * K / el, "data": "synthetic"}
    if hooks is not None:
        line["fused_hook_calls"], line["fallback_hook_calls"] = hooks.n_fused, hooks.n_fallback
    print(json.dumps(line), flush=True)


def run_env_graph(args, rank):
    """The reference Environment.step captured ONCE as a CUDA graph (vectorizedmultiagentsimulator_b200.graph) and
    replayed: actions copied in, obs / rewards / dones / infos left in static tensors.  Timed with CUDA events around K
    replays (the step is GPU-bound now) and by wall clock including the per-step host work of GraphedEnvironment.step."""
    if rank != 0:
        return
    import torch

    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    sys.path[:0] = [ref_dir, os.path.join(ROOT, "oracle", "gym_stub")]
    import vmas

    from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment
    from vectorizedmultiagentsimulator_b200.workload import Workload

    wl = Workload(args.scenario, 1, seed=0)
    n, K, W = args.num_envs, args.steps, max(args.warmup, 3)
    env = vmas.make_env(wl.meta["scenario"], num_envs=n, device="cuda", seed=0, continuous_actions=True, **wl.meta["kwargs"])
    if env.world._substeps != wl.desc.substeps:
        env.world._substeps, env.world._sub_dt = wl.desc.substeps, env.world._dt / wl.desc.substeps
    genv = GraphedEnvironment(env, check_actions=False)
    torch.manual_seed(1234)
    acts = [[torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in env.agents] for _ in range(8)]
    for i in range(W):
        genv.step(acts[i % 8])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(K):
        genv.step(acts[i % 8])
    e1.record()
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    line = {"impl": args.impl, "metric": METRIC, "unit": "env-steps/s", "n_gpus": 1, "steps": K, "warmup": W,
            "config": {"workload": f"{args.scenario} num_envs={n}, reference Environment.step on cuda:0 as one CUDA graph "
                                   f"(torch {torch.__version__}), actions copied in every step"},
            "env_step_ms_gpu_events": e0.elapsed_time(e1) / K, "env_step_ms_wall": el / K * 1e3, "value": n * K / el,
            "graph_captures": genv.n_captures, "data": "synthetic"}
    print(json.dumps(line), flush=True)


def run_queries(args, rank):
    """World.get_distance on device-resident poses, one launch per call, per shape combination; the
    reference's torch implementation on the same tensors beside it when oracle/_ref travelled.  One JSON line."""
    if rank != 0:
        return
    import torch

    from vectorizedmultiagentsimulator_b200 import queries
    from vectorizedmultiagentsimulator_b200.worlddesc import BOX, LINE, SPHERE

    dev = torch.device("cuda", 0)
    n, K, W = args.num_envs, args.steps, max(args.warmup, 3)
    R = max(2, -(-2 * L2_BYTES // (n * 4 * 7)))  # pose sets visited round-robin: inputs come from HBM
    g = torch.Generator().manual_seed(1)
    sets = [[((torch.rand(n, 2, generator=g) * 2 - 1).to(dev), (torch.rand(n, 1, generator=g) * 6.28).to(dev)) for _ in range(2)]
            for _ in range(R)]
    dims = {SPHERE: (0.1, 0.0, 0.0), BOX: (0.0, 0.4, 0.2), LINE: (0.0, 0.5, 0.0)}
    names = {SPHERE: "sphere", BOX: "box", LINE: "line"}
    ref_world = None
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if os.path.isdir(os.path.join(ref_dir, "vmas")):
        sys.path[:0] = [ref_dir, os.path.join(ROOT, "oracle", "gym_stub")]
        from vmas.simulator import core as rcore

        ref_world = rcore.World(batch_dim=n, device=dev)
        mk = {SPHERE: lambda: rcore.Sphere(0.1), BOX: lambda: rcore.Box(length=0.4, width=0.2), LINE: lambda: rcore.Line(length=0.5)}
    peak, peak_src = peaks()
    res = {}
    for sa, sb in ((SPHERE, SPHERE), (BOX, SPHERE), (LINE, SPHERE), (LINE, LINE), (BOX, LINE), (BOX, BOX)):
        poses = [(queries.make_pose(sa, *dims[sa], s[0][0], s[0][1], torch), queries.make_pose(sb, *dims[sb], s[1][0], s[1][1], torch))
                 for s in sets]
        for i in range(W * R):
            queries.get_distance(poses[i % R][0][0], poses[i % R][1][0], n, sets[0][0][0], torch)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(K):
            queries.get_distance(poses[i % R][0][0], poses[i % R][1][0], n, sets[0][0][0], torch)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        alg = 4 * n * (2 + (sa != SPHERE) + 2 + (sb != SPHERE) + 1)
        r = {"ms": ms, "queries_per_s": n / (ms * 1e-3), "hbm_frac": alg / (ms * 1e-3) / 1e9 / peak}
        if ref_world is not None:
            ea = rcore.Landmark(f"a{sa}{sb}", shape=mk[sa]())
            eb = rcore.Landmark(f"b{sa}{sb}", shape=mk[sb]())
            for e, s in ((ea, sets[0][0]), (eb, sets[0][1])):
                ref_world.add_landmark(e)  # batch_dim + device (core.py:1178)
                e.state.pos, e.state.rot = s[0], s[1]
            for _ in range(3):
                ref_world.get_distance(ea, eb)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(max(3, K // 10)):
                ref_world.get_distance(ea, eb)
            e1.record()
            torch.cuda.synchronize()
            r["reference_torch_cuda_ms"] = e0.elapsed_time(e1) / max(3, K // 10)
        res[f"{names[sa]}-{names[sb]}"] = r
    ss = res["sphere-sphere"]
    print(json.dumps({"impl": "queries", "metric": "distance_queries_per_sec", "value": ss["queries_per_s"], "unit": "queries/s",
                      "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ss["ms"], "higher_is_better": True, "dtype": "f32",
                      "data": "synthetic", "config": {"workload": f"World.get_distance, num_envs={n}, one launch per call",
                                                     "l2": f"{R} pose sets visited round-robin"},
                      "roofline": {"bound": "hbm", "achieved": ss["hbm_frac"] * peak, "peak": peak, "unit": "GB/s",
                                   "frac": ss["hbm_frac"], "traffic": None, "peak_source": peak_src, "kernel": "vb::pair_query_kernel"},
                      "per_shape_pair": res, "gpu_launches": K}), flush=True)


def main():
    args = parse()
    rank, local_rank, world = dist_env()
    if args.impl == "dropin-graph":
        run_env_graph(args, rank)
    elif args.impl in ("reference-cuda", "dropin", "dropin-hooks"):
        run_env_cuda(args, rank)
    elif args.impl == "queries":
        run_queries(args, rank)
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
