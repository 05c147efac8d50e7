#!/usr/bin/env python
"""Where the time of one vmasb200_step_host call goes (navigation, 32768 envs): the same rollout with outputs switched
off one by one, for the chunk size given by VMASB200_HOST_CHUNK_ENVS.  Prints ms per step per variant.
    gpurun -- 'for c in 8192 16384 32768; do VMASB200_HOST_CHUNK_ENVS=$c python tools/e2e_probe.py; done'"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vectorizedmultiagentsimulator_b200.plan import StepPlan  # noqa: E402
from vectorizedmultiagentsimulator_b200.workload import Workload  # noqa: E402

scenario = sys.argv[1] if len(sys.argv) > 1 else "navigation"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
wl = Workload(scenario, n, seed=0)
plan = StepPlan(wl.desc)


def pinned(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t, t.numpy()


dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
ld = dyn0.shape[1]
f, t = wl.random_actions(1)
keep = [pinned(dyn0), pinned(stat0), pinned(f[0]), pinned(t[0]), pinned(np.zeros((plan.dyn_rows, ld), np.float32)),
        pinned(np.zeros((max(plan.act_rows, 1), ld), np.float32)), pinned(np.zeros((max(plan.lidar_rows, 1), ld), np.float32))]
dyn, stat, fr, tq, o_dyn, o_act, o_lid = (k[1] for k in keep)
o_act = o_act[:plan.act_rows] if plan.act_rows else None
o_lid = o_lid[:plan.lidar_rows] if plan.lidar_rows else None
L = __import__("vectorizedmultiagentsimulator_b200._lib", fromlist=["lib"]).lib()
P = lambda a: None if a is None else a.ctypes.data


def run(name, d_in, s_in, force, torque, od, oa, ol, K=200):
    args = (plan._h, P(d_in), P(s_in), P(force), P(torque), P(od), P(oa), P(ol), None, n, ld, 0)
    for _ in range(5):
        assert L.vmasb200_step_host(*args) == 0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
        L.vmasb200_step_host(*args)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / K * 1e3
    print(f"  {name:<46s} {ms:7.3f} ms  {n / ms / 1e3:7.1f} M env-steps/s")


print(f"{scenario} n={n} chunk={os.environ.get('VMASB200_HOST_CHUNK_ENVS', 'default')} direct={os.environ.get('VMASB200_HOST_DIRECT', '0')}")
run("state in, all out", dyn, stat, fr, tq, o_dyn, o_act, o_lid)
run("resident, state + act + lidar out", None, None, fr, tq, o_dyn, o_act, o_lid)
run("resident, state + lidar out", None, None, fr, tq, o_dyn, None, o_lid)
run("resident, state out", None, None, fr, tq, o_dyn, None, None)
run("resident, lidar out", None, None, fr, tq, None, None, o_lid)
run("resident, nothing out", None, None, fr, tq, None, None, None)
run("resident, no actions in, nothing out", None, None, None, None, None, None, None)
