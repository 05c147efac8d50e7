"""Synthetic workloads for bench.py / smoke(): a world description + N seeded environments + random actions.

The world descriptions are the ones extracted from the reference scenarios by tests/golden/make_golden.py
(``desc_json`` inside tests/golden/<scenario>.npz); nothing here needs the reference at run time.  Initial states
are the reference's own reset states (the fixture's ``state0``, e.g. navigation.py:140-197 spawns) tiled to N
environments with a small seeded jitter on the dynamic entities so that all environments differ; actions are
U(-a, a) per agent with ``a`` the largest force the reference produced for that agent in the recorded rollout
(``Environment.get_random_actions`` x ``u_multiplier``, environment.py:535-561, :709).
"""
import json
import os

import numpy as np

from .worlddesc import WorldDesc

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


class Workload:
    def __init__(self, scenario: str, n_envs: int, seed: int = 0):
        z = np.load(os.path.join(GOLDEN, scenario + ".npz"))
        self.scenario = scenario
        self.desc = WorldDesc.from_json(str(z["desc_json"]))
        self.meta = json.loads(str(z["meta_json"]))
        self.n_envs = n_envs
        rng = np.random.default_rng(seed)
        s0 = z["state0"]  # [E, n0, 6]
        n0 = s0.shape[1]
        idx = np.arange(n_envs) % n0
        state = np.ascontiguousarray(s0[:, idx, :])
        for e, en in enumerate(self.desc.entities):
            if en.movable:
                state[e, :, 0:2] += rng.uniform(-0.02, 0.02, (n_envs, 2)).astype(np.float32)
        self.state0 = state
        j0 = z["jrot0"]
        self.jrot0 = np.ascontiguousarray(j0[:, idx]) if j0.size else j0
        f = z["forces"]  # [T, A, n0, 2]
        t = z["torques"]
        self.A = f.shape[1]
        self.f_scale = np.abs(f).max(axis=(0, 2, 3)) if f.size else np.zeros(0)
        self.t_scale = np.abs(t).max(axis=(0, 2)) if t.size else np.zeros(0)
        self._rng = rng

    def random_actions(self, n_sets: int):
        """force [n_sets, N, A, 2], torque [n_sets, N, A] (host API layout)."""
        f = self._rng.uniform(-1, 1, (n_sets, self.n_envs, self.A, 2)).astype(np.float32)
        t = self._rng.uniform(-1, 1, (n_sets, self.n_envs, self.A)).astype(np.float32)
        f *= self.f_scale.astype(np.float32)[None, None, :, None]
        t *= self.t_scale.astype(np.float32)[None, None, :]
        return f, t
