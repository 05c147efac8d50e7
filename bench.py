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
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "reference-cuda", "dropin", "dropin-hooks", "dropin-graph", "queries", "reset"],
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
    ap.add_argument("--fixture-inputs", action="store_true", help="tiled fixture reset states instead of the reference's own make_env reset")
    ap.add_argument("--no-torch-cuda", action="store_true", help="skip the torch_cuda_baseline block (reference World.step on this GPU)")
    ap.add_argument("--no-1m", action="store_true", help="skip the roofline_1m block (the same kernel at 1 048 576 envs)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --num-envs per GPU; strong: --num-envs in total, split evenly over the GPUs (SURVEY 8d football sweep)")
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

    def mark(self):
        """the timed region starts here: only samples from now on count"""
        self.first = len(self.rows)

    def stop(self, busy=None):
        """busy: a callable that keeps the GPU under the timed region's own load (the same launches, untimed).  A region of
        a few milliseconds ends before nvidia-smi (one sample per 100 ms, slower to start with 8 ranks) has seen it, so the
        same load is kept up until three samples exist (at most 3 s): the clocks reported are clocks UNDER this load."""
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        first = getattr(self, "first", 0)
        t0 = time.time()
        while busy is not None and len(self.rows) < first + 3 and time.time() - t0 < 3.0:
            busy()
        if busy is None:
            time.sleep(0.15)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows[first:]:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm),
                "window": "from the start of the timed region; the same launches are kept up (untimed) until 3 samples exist"
                          if busy is not None else "timed region + 0.15 s"}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_capture(scenario, n, compiled=False):
    """What the committed `ncu --set full` capture of this workload says about one launch of the step kernel
    (profiles/r2_traffic.json, else the earlier r1z / r1w / r1 sets; written by tools/profile_collect.py):
    dram_bytes = dram__bytes_read.sum + dram__bytes_write.sum, warp_inst = smsp__inst_executed.sum.  Only a capture of
    the SAME kernel counts (world-compiled vmasb200_step_world or generic vb::step_kernel); {} when there is none."""
    for name in ("r2_traffic.json", "r1z_traffic.json", "r1w_traffic.json", "r1_traffic.json"):
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


def ref_available():
    return os.path.isdir(os.path.join(ROOT, "oracle", "_ref", "vmas"))


def import_vmas():
    for p in (os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import vmas

    return vmas


class RefInputs:
    """The synthetic inputs SURVEY 8(d) specifies, made by the reference itself (oracle/_ref, INPUT GENERATION only, before
    any timed region): `make_env(scenario, num_envs=N, seed=seed, continuous_actions=True, **kw)`, its reset state, and
    for n_sets steps the forces / torques its own pre-step path (`get_random_action` under torch.manual_seed(1234) ->
    `_set_action` -> `env_process_action`, environment.py:386-390, 535-561) hands to World.step.  Same attributes as
    vectorizedmultiagentsimulator_b200.workload.Workload, which tiles 32 recorded reset states instead and is the
    fallback where the reference install did not travel."""

    def __init__(self, fixture, n, seed, n_sets, device="cpu"):
        import torch

        from vectorizedmultiagentsimulator_b200.dropin import _joints_in_order
        from vectorizedmultiagentsimulator_b200.workload import Workload
        from vectorizedmultiagentsimulator_b200.worlddesc import WorldDesc

        vmas = import_vmas()
        fx = Workload(fixture, 1, seed=0)  # fixture meta: reference scenario name, kwargs, substeps
        self.scenario, self.meta, self.n_envs = fixture, fx.meta, n
        env = vmas.make_env(fx.meta["scenario"], num_envs=n, device=device, seed=seed, continuous_actions=True, **fx.meta["kwargs"])
        world = env.world
        if world._substeps != fx.desc.substeps:  # "transport, substeps>1": imposed on the built world (note i)
            world._substeps, world._sub_dt = fx.desc.substeps, world._dt / fx.desc.substeps
        self.env = env
        self.desc = WorldDesc.from_world(world)
        ents = list(world.entities)
        self.state0 = torch.stack([torch.cat([e.state.pos, e.state.vel, e.state.rot, e.state.ang_vel], -1) for e in ents]).cpu().numpy()
        jr = []
        for j in (_joints_in_order(world, ents) if getattr(world, "_joints", None) else []):
            fr = j.fixed_rotation
            jr.append(fr.reshape(-1).expand(n).cpu().numpy() if hasattr(fr, "reshape") else np.full(n, float(fr), np.float32))
        self.jrot0 = np.stack(jr).astype(np.float32) if jr else np.zeros((0, n), np.float32)
        agents = list(world.agents)
        self.A = len(agents)
        torch.manual_seed(1234)
        f = np.zeros((n_sets, n, self.A, 2), np.float32)
        t = np.zeros((n_sets, n, self.A), np.float32)
        for k in range(n_sets):
            acts = [env.get_random_action(a) for a in env.agents]
            for i, a in enumerate(env.agents):
                env._set_action(acts[i], a)
            for a in world.agents:
                env.scenario.env_process_action(a)
            for i, a in enumerate(agents):
                if a.state.force is not None:
                    f[k, :, i] = a.state.force.cpu().numpy()
                if a.state.torque is not None:
                    t[k, :, i] = a.state.torque.cpu().numpy()[:, 0]
        self._f, self._t = f, t
        self.source = (f"reference make_env('{fx.meta['scenario']}', num_envs={n}, seed={seed}) reset state; forces of "
                       f"get_random_action under torch.manual_seed(1234), pre-generated ({n_sets} sets)")

    def random_actions(self, n_sets):
        assert n_sets <= self._f.shape[0]
        return self._f[:n_sets], self._t[:n_sets]


def make_inputs(args, n, seed, n_sets, need_env=False):
    """SURVEY 8(d) inputs from the reference when its install travelled (oracle/_ref), else the tiled fixture states."""
    from vectorizedmultiagentsimulator_b200.workload import Workload

    if ref_available() and not args.fixture_inputs:
        try:
            import torch

            dev = "cuda" if (n > 65536 and torch.cuda.is_available() and not need_env) else "cpu"
            return RefInputs(args.scenario, n, seed, n_sets, device=dev)
        except Exception as e:  # the copy did not import / build this scenario on this box: say so in the line
            print(f"bench.py: reference inputs unavailable ({type(e).__name__}: {e}); using the tiled fixture states", file=sys.stderr)
    wl = Workload(args.scenario, n, seed=seed)
    wl.source = "32 recorded reset states of the reference (tests/golden) tiled with +-0.02 jitter; forces U(-a, a), a = largest recorded force"
    wl.env = None
    return wl


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
    """Reference arm: the reference's own CPU implementation of the path, on the host cores (rank 0 only), on the same
    inputs as the b200 arm (make_inputs), forces pre-generated outside the timed region."""
    if rank != 0:
        return
    line = {"impl": "reference", "metric": METRIC, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic"}
    n = args.num_envs
    value = kind = sample = None
    cores = os.cpu_count() or 1
    wl = make_inputs(args, n, 0, 8, need_env=True)
    if getattr(wl, "env", None) is not None:
        try:
            value, sample = _time_ref_torch(args, wl, n)
            kind = "reference"
        except Exception as e:  # the copy did not run on this box: fall back to the port, say so
            sample = f"oracle/_ref failed ({type(e).__name__}: {e}); "
    if value is None:
        value, cores, s2 = cpu_port_rate(wl, seconds=max(2.0, 0.05 * args.steps))
        kind, sample = "port", (sample or "") + s2
    line.update({"value": value, "ms_per_step": 1e3 * n / value,
                 "config": {"workload": f"{args.scenario} num_envs={n} per GPU, substeps={wl.desc.substeps}, {len(wl.desc.lidars)} lidars",
                            "arm": f"host CPU, {kind}", "inputs": wl.source},
                 "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": kind, "sample": sample},
                 "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(line), flush=True)


def _time_ref_torch(args, wl, n):
    """World.step + Lidar.measure of the UNMODIFIED reference physics (oracle/_ref = reference copy with the one-line
    import fix of oracle/make_ref.sh) under torch CPU with all host threads, from the reset state and with the
    pre-generated forces of `wl` (RefInputs)."""
    import torch

    torch.set_num_threads(os.cpu_count() or 1)
    world = wl.env.world
    agents = list(world.agents)
    f, t = wl.random_actions(8)
    fs = [[torch.from_numpy(np.ascontiguousarray(f[k][:, a])) for a in range(len(agents))] for k in range(8)]
    ts = [[torch.from_numpy(np.ascontiguousarray(t[k][:, a:a + 1])) for a in range(len(agents))] for k in range(8)]

    def one(k):
        for i, a in enumerate(agents):
            if a.state.force is not None:
                a.state.force = fs[k % 8][i]
            if a.state.torque is not None:
                a.state.torque = ts[k % 8][i]
        world.step()
        for a in agents:
            for s in a.sensors or []:
                s.measure()

    for k in range(args.warmup):
        one(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        one(k)
    el = time.perf_counter() - t0
    return n * args.steps / el, f"{args.steps} steps x {n} envs, reference torch {torch.__version__} CPU World.step + Lidar.measure, {torch.get_num_threads()} threads ({el:.1f} s)"


def torch_cuda_baseline(args, n, value_per_gpu, ms_kernel):
    """The reference on THIS GPU (what north_star's ">= 10x the reference's torch-CUDA env-steps/s" refers to):
    its Environment on cuda:0 at the same N, the pre-generated actions of the protocol, World.step + Lidar.measure alone
    (host clock with a synchronize on both sides: the torch path is launch-bound, the GPU idles between its kernels) and
    the whole Environment.step; then the same Environment with the drop-in enabled, and as one CUDA graph."""
    import torch

    from vectorizedmultiagentsimulator_b200.workload import Workload

    vmas = import_vmas()
    fx = Workload(args.scenario, 1, seed=0)
    K, W = max(5, min(args.steps, 30)), 3

    def make():
        env = vmas.make_env(fx.meta["scenario"], num_envs=n, device="cuda", seed=0, continuous_actions=True, **fx.meta["kwargs"])
        if env.world._substeps != fx.desc.substeps:
            env.world._substeps, env.world._sub_dt = fx.desc.substeps, env.world._dt / fx.desc.substeps
        return env

    def timed(fn, k):
        for i in range(W):
            fn(i)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(k):
            fn(i)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / k

    def arms(env):
        world = env.world
        torch.manual_seed(1234)
        acts = [[torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in env.agents] for _ in range(8)]
        agents = list(world.agents)

        def world_step(i):
            world.step()
            for a in agents:
                for s_ in a.sensors or []:
                    s_.measure()

        t_env = timed(lambda i: env.step([a.clone() for a in acts[i % 8]]), K)
        t_ws = timed(world_step, K)
        return t_ws, t_env

    out = {"device": f"cuda:0 ({torch.cuda.get_device_name(0)}), torch {torch.__version__}", "num_envs": n, "steps": K, "warmup": W,
           "unit": "env-steps/s", "timing": "host clock, torch.cuda.synchronize() on both sides"}
    ref = make()
    t_ws, t_env = arms(ref)
    out["reference"] = {"world_step": n / t_ws, "world_step_ms": 1e3 * t_ws, "environment_step": n / t_env, "environment_step_ms": 1e3 * t_env}
    del ref
    from vectorizedmultiagentsimulator_b200 import dropin

    env = make()
    dropin.enable(env, hooks=True)
    t_ws, t_env = arms(env)
    out["dropin"] = {"world_step": n / t_ws, "world_step_ms": 1e3 * t_ws, "environment_step": n / t_env, "environment_step_ms": 1e3 * t_env,
                     "what": "the same Environment with World.step / cast_rays / scenario hooks routed through libvmasb200 (dropin.enable(env, hooks=True))"}
    try:
        from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment

        genv = GraphedEnvironment(make(), check_actions=False)
        torch.manual_seed(1234)
        acts = [[torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in genv.env.agents] for _ in range(8)]
        t_g = timed(lambda i: genv.step(acts[i % 8]), max(K, 50))
        out["dropin_graph"] = {"environment_step": n / t_g, "environment_step_ms": 1e3 * t_g,
                               "what": "the reference's Environment.step captured once as a CUDA graph and replayed (graph.GraphedEnvironment)"}
    except Exception as e:
        out["dropin_graph"] = {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    r = out["reference"]
    out["speedup"] = {"kernel_vs_reference_world_step": value_per_gpu / r["world_step"],
                      "dropin_world_step": out["dropin"]["world_step"] / r["world_step"],
                      "dropin_environment_step": out["dropin"]["environment_step"] / r["environment_step"]}
    if "environment_step" in out["dropin_graph"]:
        out["speedup"]["graph_environment_step"] = out["dropin_graph"]["environment_step"] / r["environment_step"]
    return out


def pin_rank(local_rank, world):
    """One rank per GPU: give each rank its own slice of the host cores (all GPUs of these boxes report the same CPU
    affinity, so the slices are equal shares), before any pinned host buffer is allocated."""
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // world)
        mine = cores[local_rank * per:(local_rank + 1) * per] or cores
        os.sched_setaffinity(0, mine)
        import torch

        torch.set_num_threads(max(1, len(mine)))
    except (AttributeError, OSError):
        pass


def fp32_peak_tflops(clocks):
    """Non-tensor FP32 peak: the FFMA micro-benchmark committed under profiles/ (tools/fma_peak.cu) when there is one, else
    148 SMs x 128 lanes x 2 x the max SM clock."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_fma_peak.json")) as f:
            m = json.load(f)
        return float(m["ffma_tflops"]), f"measured FFMA micro-benchmark (profiles/r2_fma_peak.json: {m['ffma_tflops']:.1f} TFLOP/s at {m.get('sm_mhz', '?')} MHz)"
    except (OSError, ValueError, KeyError):
        mhz = (clocks or {}).get("sm_max_mhz") or 1965.0
        return 148 * 128 * 2 * mhz * 1e6 / 1e12, "148 SMs x 128 lanes x 2 x max SM clock"


def time_steps(plan, ios, K, W, barrier, torch):
    """K launches over the batches of `ios` round-robin, CUDA events on the launching stream, >= 3 warm-up steps per batch."""
    R = len(ios)
    for i in range(max(W, 3) * R):
        plan.step_io(ios[i % R])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        plan.step_io(ios[i % R])
    e1.record()
    barrier()
    return e0.elapsed_time(e1)


def device_batches(plan, wl, n, dev, exact, torch, max_sets=64):
    """R independent env batches (same initial state, different forces) so that a step's inputs come from HBM, not L2."""
    A = plan.A
    ld = plan.ld_for(n)
    per_batch = 4 * ld * (2 * plan.dyn_rows + plan.stat_rows + plan.act_rows + plan.lidar_rows + 3 * A)
    R = max(2, min(max_sets, -(-2 * L2_BYTES // max(per_batch, 1))))
    dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
    f_all, t_all = wl.random_actions(R)
    stream = torch.cuda.current_stream().cuda_stream
    ios, keep = [], []
    for r in range(R):
        dyn = torch.from_numpy(dyn0).to(dev)
        stat = torch.from_numpy(stat0).to(dev) if plan.stat_rows else None
        fr = torch.from_numpy(np.ascontiguousarray(f_all[r])).to(dev)  # [N, A, 2]
        tr = torch.from_numpy(np.ascontiguousarray(t_all[r])).to(dev)  # [N, A]
        forces = [fr[:, a, :] for a in range(A)]
        torques = [tr[:, a:a + 1] for a in range(A)]
        act = torch.zeros(plan.act_rows, ld, device=dev) if plan.act_rows else None
        lid = torch.zeros(plan.lidar_rows, ld, device=dev) if plan.lidar_rows else None
        ios.append(plan.make_io(dyn, dyn, stat, forces, torques, act, lid, None, n, ld, bool(exact), stream))
        keep.append((dyn, stat, fr, tr, act, lid))
    return ios, keep, R, per_batch


def roofline_block(args, plan, desc, n, ms_per_step, clocks, compiled, peak, peak_src):
    """Both roofs of SURVEY 8(d) for one launch over n envs: HBM (algorithmic bytes at the measured copy bandwidth) and FP32
    (algorithmic FLOPs at the non-tensor FP32 peak); `bound` names the one that binds, i.e. whose ideal time is longer."""
    cap = ncu_capture(args.scenario, n, compiled)
    alg = plan.alg_bytes_per_env * n
    hbm_achieved = alg / (ms_per_step * 1e-3) / 1e9
    flops = alg_flops_per_env(plan, desc)
    fpeak, fsrc = fp32_peak_tflops(clocks)
    fp32_achieved = flops * n / (ms_per_step * 1e-3) / 1e12
    binding = "fp32" if flops / (fpeak * 1e12) > plan.alg_bytes_per_env / (peak * 1e9) else "hbm"
    hbm = {"achieved": hbm_achieved, "peak": peak, "unit": "GB/s", "frac": hbm_achieved / peak, "peak_source": peak_src,
           "alg_bytes_per_env": plan.alg_bytes_per_env}
    fp32 = {"achieved": fp32_achieved, "peak": fpeak, "unit": "TFLOP/s", "frac": fp32_achieved / fpeak, "peak_source": fsrc,
            "alg_flops_per_env": flops, "flops_source": "SURVEY 8(d) per-primitive op counts of the reference (the kernel's culls execute fewer)"}
    fcap = cap
    if not fcap.get("fp32_flops"):  # no capture at this batch size: the per-env count of the same kernel's capture at another one
        for other in (32768, 1 << 20):
            c2 = ncu_capture(args.scenario, other, compiled)
            if c2.get("fp32_flops"):
                fcap = c2
                break
    if fcap.get("fp32_flops"):
        # FP32 operations the kernel really executed (ncu: thread-level FADD + FMUL + 2 FFMA of one launch).  The algorithmic
        # count above is the REFERENCE's (every narrow phase and every ray for every env); the kernel's exact
        # verdicts and conservative culls skip what is provably zero, so `executed_frac` is the real load of the FP32 pipe
        per_env = fcap["fp32_flops"] / fcap.get("n_envs", n)
        fp32["executed_flops_per_env"] = per_env
        fp32["executed_tflops"] = per_env * n / (ms_per_step * 1e-3) / 1e12
        fp32["executed_frac"] = fp32["executed_tflops"] / fpeak
        fp32["executed_source"] = f"ncu capture {args.scenario}@{fcap.get('n_envs', n)} (profiles/r2_traffic.json)"
    b = hbm if binding == "hbm" else fp32
    r = {"bound": binding, "achieved": b["achieved"], "peak": b["peak"], "unit": b["unit"], "frac": b["frac"],
         "traffic": cap.get("dram_bytes"), "hbm": hbm, "fp32": fp32, "num_envs": n, "ms_per_launch": ms_per_step,
         "kernel": "vmasb200_step_world (world-compiled, NVRTC)" if compiled else "vb::step_kernel"}
    r["bound_rule"] = ("the roof whose ideal time for the ALGORITHMIC work of one launch is longer (SURVEY 8d bytes and the reference's "
                       "op counts); fp32.executed_frac and issue.frac say how busy the FP32 pipe and the issue slots really are")
    if cap.get("warp_inst"):
        mhz = (clocks or {}).get("sm_mhz") or 1965.0
        ips = cap["warp_inst"] / (ms_per_step * 1e-3)
        r["issue"] = {"warp_inst_per_launch": cap["warp_inst"], "achieved_ginst_s": ips / 1e9,
                      "peak_ginst_s": 4 * 148 * mhz * 1e-3, "frac": ips / (4 * 148 * mhz * 1e6)}
    return r


def run_b200(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    from vectorizedmultiagentsimulator_b200.plan import StepPlan

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the CUDA step has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        pin_rank(local_rank, world)
    # weak: --num-envs per GPU; strong: --num-envs in total, split evenly ("N split evenly" for the football sweep)
    n = args.num_envs if args.scaling == "weak" else -(-args.num_envs // world)
    A_guess_sets = 64
    wl = make_inputs(args, n, rank, A_guess_sets if n <= 65536 else 4)
    wl.desc.envs_per_cta = args.envs_per_cta
    plan = StepPlan(wl.desc)
    K, W = args.steps, args.warmup
    ios, keep, R, per_batch = device_batches(plan, wl, n, dev, args.exact, torch)
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up pass (also triggers the NVRTC build / cubin load), then the timed region with the clock sampler running
    sampler = ClockSampler(local_rank)  # started early: nvidia-smi takes ~0.1 s to deliver its first sample
    time_steps(plan, ios, 3, W, barrier, torch)
    sampler.mark()
    ms = time_steps(plan, ios, K, W, barrier, torch)
    clocks = sampler.stop(busy=lambda: time_steps(plan, ios, max(K, 200), 0, lambda: torch.cuda.synchronize(), torch))
    repairs = plan.sync_check(stream)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * n * K / (ms * 1e-3)

    # L2-resident variant (one batch stepped in place, what a rollout does): reported beside, not as `value`
    ms_hot = time_steps(plan, ios[:1], K, 3, lambda: torch.cuda.synchronize(), torch)

    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(plan, wl, n, K, world, dev)
    if rank == 0:
        peak, peak_src = peaks()
        jit_state, jit_launches = plan.jit(n)
        compiled = jit_launches > 0   # the launches took the world-compiled kernel (csrc/jit.h), not the generic one
        line = {
            "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.scenario} num_envs={n} per GPU, substeps={plan.S}, {len(wl.desc.lidars)} lidars",
                       "inputs": wl.source,
                       "l2": f"{R} env batches visited round-robin, working set {R * per_batch / 2**20:.0f} MiB > L2 126 MiB",
                       "envs_per_cta": plan.tile_for(n),
                       "layout": ("SoA rows [field][env]; world-compiled kernel (NVRTC): one thread per env, state in registers"
                                  if compiled else
                                  "SoA rows [field][env]; 128 threads per CTA, 128 / envs_per_cta of them share an env"),
                       "exact": int(args.exact), "speculation_repairs": repairs,
                       "launch": ("consecutive steps are launched back to back on one stream with programmatic dependent launch "
                                  "(griddepcontrol: a step's CTAs are set up while its predecessor drains, its first instruction waits "
                                  "for that grid); ms_per_step is the period of that pipeline" if compiled and os.environ.get("VMASB200_PDL", "1") != "0"
                                  else "consecutive steps serialised on one stream"),
                       "l2_resident_ms_per_step": ms_hot / K},
            "roofline": roofline_block(args, plan, wl.desc, n, ms / K, clocks, compiled, peak, peak_src),
            # one launch per step; a compiled world that can have thin contacts is followed by its 1-CTA repair_kernel
            "gpu_launches": K * (2 * plan.S if args.exact else (2 if compiled and plan.jit_tail_launches else 1)),
            "clocks": clocks,
        }
        if n < 262144:
            line["roofline"]["note"] = (f"one launch over {n} envs moves {plan.alg_bytes_per_env * n / 1e6:.1f} MB in a single wave of "
                                        f"{-(-n // plan.tile_for(n))} CTAs: {1e3 * plan.alg_bytes_per_env * n / (peak * 1e9):.4f} ms at the HBM peak, below one "
                                        "kernel launch - latency-bound; the roofline fraction of this path is meaningful from ~256 K envs "
                                        "(SURVEY 8d), see roofline_1m (same kernel, same run)")
        if e2e is not None:
            line["e2e"] = e2e
    # ---- the same kernel at 1 048 576 envs, measured in this run: what the north_star's ">= 50 % of the HBM roofline" is about
    if world == 1 and n < 262144 and not args.no_1m:
        try:
            from vectorizedmultiagentsimulator_b200.workload import Workload

            n1 = 1 << 20
            del ios, keep
            torch.cuda.empty_cache()
            wl1 = Workload(args.scenario, n1, seed=0)
            wl1.desc.envs_per_cta = args.envs_per_cta
            plan1 = StepPlan(wl1.desc)
            ios1, keep1, R1, pb1 = device_batches(plan1, wl1, n1, dev, args.exact, torch, max_sets=2)
            time_steps(plan1, ios1, 3, 3, lambda: torch.cuda.synchronize(), torch)
            s1 = ClockSampler(local_rank)
            K1 = max(10, min(K, 50))
            s1.mark()
            ms1 = time_steps(plan1, ios1, K1, 3, lambda: torch.cuda.synchronize(), torch)
            ck1 = s1.stop(busy=lambda: time_steps(plan1, ios1, K1, 0, lambda: torch.cuda.synchronize(), torch))
            _, jl1 = plan1.jit(n1)
            r1 = roofline_block(args, plan1, wl1.desc, n1, ms1 / K1, ck1, jl1 > 0, peak, peak_src)
            r1.update({"value": n1 * K1 / (ms1 * 1e-3), "unit_value": "env-steps/s", "steps": K1, "clocks": ck1,
                       "inputs": "tiled fixture reset states (workload.Workload)", "l2": f"{R1} batches of {pb1 / 2**20:.0f} MiB visited round-robin"})
            line["roofline_1m"] = r1
            del ios1, keep1
            torch.cuda.empty_cache()
        except Exception as e:
            line["roofline_1m"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    if rank == 0:
        if world == 1 and not args.no_torch_cuda and ref_available():
            try:
                line["torch_cuda_baseline"] = torch_cuda_baseline(args, n, value, ms / K)
            except Exception as e:
                line["torch_cuda_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_port_rate(wl, seconds=10.0)
            line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(plan, wl, n, K, world, dev):
    """Same metric through the C-ABI HOST-buffer call (vmasb200_step_host) as a rollout drives it: every step the actions
    come from pinned host memory (H2D), the kernel runs, and the new state + LIDAR readings (what observations are made
    of) go back to pinned host memory (D2H), then a stream sync.  The state itself stays on the device between the calls
    (h_dyn_in = NULL from the second call on: it is what the previous call just returned)."""
    import torch
    import torch.distributed as dist

    def pinned(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()

    dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
    ld = dyn0.shape[1]
    f, t = wl.random_actions(2)
    keep = [pinned(dyn0), pinned(stat0), pinned(f[0]), pinned(t[0]), pinned(f[1]), pinned(t[1])]
    dyn, stat = keep[0][1], keep[1][1]
    acts = [(keep[2][1], keep[3][1]), (keep[4][1], keep[5][1])]
    outs = [pinned(np.zeros((plan.dyn_rows, ld), np.float32)), pinned(np.zeros((max(plan.lidar_rows, 1), ld), np.float32))]
    out = (outs[0][1], None, outs[1][1][:plan.lidar_rows] if plan.lidar_rows else None)
    plan.step_host(dyn, stat, acts[0][0], acts[0][1], exact=False, want_masks=False, out=out)  # uploads state + static rows
    # one prepared C-ABI call per action buffer (plan.host_stepper: the ctypes call and nothing else)
    calls = [plan.host_stepper(n, ld, a[0], a[1], out[0], None, out[2]) for a in acts]
    for i in range(20):  # the first calls pay for lazy driver work (pinned-buffer mappings, copy-engine warm-up): 0.28 -> 0.25 ms
        calls[i % 2]()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(K):
        calls[i % 2]()
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([el], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        el = float(tt.item())
    h2d = 4 * n * 3 * plan.A
    d2h = 4 * n * (plan.dyn_rows + plan.lidar_rows)
    # this box's PCIe link, measured now: one flat device-to-host copy of the step's output bytes into pinned memory
    src = torch.empty(d2h // 4, device=dev)
    dst = torch.empty(d2h // 4).pin_memory()
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    for _ in range(20):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    copy_ms = (time.perf_counter() - t1) / 20 * 1e3
    return {"value": world * n * K / el, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "ms_per_step": 1e3 * el / K,
            "pcie_floor_ms": copy_ms, "pcie_d2h_gbs": d2h / copy_ms / 1e6,
            "api": "vmasb200_step_host (C ABI, pinned host buffers): actions in, new state + LIDAR out every step; the state stays "
                   "device-resident between calls (h_dyn_in = NULL)"}


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


def run_reset(args, rank):
    """Environment.reset() and reset_at(i) followed by one step (which pays for the re-packing of the state
    tensors a reset rebinds), wall clock with a synchronize on both sides: the reference, the drop-in with the reference's
    reset path, and the drop-in with fused_reset=True.  One JSON line."""
    if rank != 0:
        return
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin
    from vectorizedmultiagentsimulator_b200.workload import Workload

    vmas = import_vmas()
    fx = Workload(args.scenario, 1, seed=0)
    n, K = args.num_envs, max(3, min(args.steps, 20))

    def arm(mode):
        env = vmas.make_env(fx.meta["scenario"], num_envs=n, device="cuda", seed=0, continuous_actions=True, **fx.meta["kwargs"])
        b = None
        if mode != "reference":
            b = dropin.enable(env, fused_reset=(mode == "fused"))
        acts = [torch.zeros(n, a.action_size, device="cuda") for a in env.agents]
        env.step(acts)

        def timed(fn):
            fn(); fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(K):
                fn()
            torch.cuda.synchronize()
            return (time.perf_counter() - t0) / K * 1e3

        full = timed(lambda: (env.reset(), env.step(acts)))
        one = timed(lambda: (env.reset_at(7), env.step(acts)))
        step = timed(lambda: env.step(acts))
        out = {"reset_plus_step_ms": full, "reset_at_plus_step_ms": one, "step_ms": step,
               "reset_ms": full - step, "reset_at_ms": one - step}
        if b is not None:
            out["state_fields_repacked"] = b.n_copies
        return out

    line = {"impl": "reset", "metric": "reset_ms", "unit": "ms", "n_gpus": 1, "steps": K, "higher_is_better": False, "data": "synthetic",
            "config": {"workload": f"{args.scenario} num_envs={n}: Environment.reset() / reset_at(7), each followed by one step"},
            "reference": arm("reference"), "dropin": arm("dropin"), "dropin_fused_reset": arm("fused")}
    line["value"] = line["dropin_fused_reset"]["reset_ms"]
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
    elif args.impl == "reset":
        run_reset(args, rank)
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
