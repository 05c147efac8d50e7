#!/usr/bin/env python
"""Golden vectors of Environment._set_action for DISCRETE actions (environment.py:660-710) from the
UNMODIFIED reference: agent.action.u after _set_action for flat and multi-discrete actions, odd and even action
counts, non-unit u_range / u_multiplier.

TEST INFRASTRUCTURE.  Run in the build container (torch CPU):  python tests/golden/make_golden_actions.py
Output: tests/golden/actions/discrete.npz (case_<i>_{action,nvec,u_range,u_mult,multi,u}).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub"), ROOT, HERE]

import vmas  # noqa: E402

N = 64
CASES = [  # (multidiscrete, nvec, u_range, u_multiplier)
    (False, [3, 3], [1.0, 1.0], [1.0, 1.0]),
    (True, [3, 3], [1.0, 1.0], [1.0, 1.0]),
    (False, [5, 4], [0.7, 1.3], [0.5, 2.0]),
    (True, [5, 4], [0.7, 1.3], [0.5, 2.0]),
    (False, [2, 7], [1.0, 0.3], [3.0, 1.0]),
    (True, [9, 2], [0.1, 2.5], [1.0, 0.25]),
]


def main():
    out = {}
    g = torch.Generator().manual_seed(7)
    for ci, (multi, nvec, ur, um) in enumerate(CASES):
        env = vmas.make_env("navigation", num_envs=N, device="cpu", seed=0, continuous_actions=False,
                            multidiscrete_actions=multi, n_agents=2)
        agent = env.world.agents[0]
        agent.discrete_action_nvec = list(nvec)
        agent.action._u_range = list(ur)
        agent.action._u_multiplier = list(um)
        agent.action._u_range_tensor = None
        agent.action._u_multiplier_tensor = None
        if multi:
            act = torch.stack([torch.randint(0, n, (N,), generator=g) for n in nvec], dim=-1)
        else:
            act = torch.randint(0, int(np.prod(nvec)), (N, 1), generator=g)
            act[: int(np.prod(nvec)), 0] = torch.arange(int(np.prod(nvec)))[:N]  # every flat action at least once
        env._set_action(act.clone(), agent)
        assert torch.equal(agent.action.u_range_tensor, torch.tensor(ur, dtype=torch.float32))
        out[f"case_{ci}_action"] = act.numpy()
        out[f"case_{ci}_nvec"] = np.asarray(nvec, np.int32)
        out[f"case_{ci}_u_range"] = np.asarray(ur, np.float32)
        out[f"case_{ci}_u_mult"] = np.asarray(um, np.float32)
        out[f"case_{ci}_multi"] = np.asarray(int(multi))
        out[f"case_{ci}_u"] = agent.action.u.detach().numpy().astype(np.float32)
    out["n_cases"] = np.asarray(len(CASES))
    np.savez_compressed(os.path.join(HERE, "actions", "discrete.npz"), **out)
    print("wrote tests/golden/actions/discrete.npz:", len(CASES), "cases")


if __name__ == "__main__":
    main()
