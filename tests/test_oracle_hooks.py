"""CPU: pins the numpy restatement of the scenario hooks (oracle/hooks_oracle.py) against outputs of the
reference's own navigation / transport hooks (tests/golden/hooks/*.npz from tests/golden/make_golden_hooks.py), bit for
bit; and checks the batch-wide World.collides semantics of the collision penalty on a synthetic thin case."""
import os

import numpy as np
import pytest

from oracle import hooks_oracle as ho

HERE = os.path.dirname(os.path.abspath(__file__))
NAV = ["navigation", "navigation_noshare", "navigation_nocoll"]


def load(name):
    return dict(np.load(os.path.join(HERE, "golden", "hooks", name + ".npz")))


def nav_inputs(g, t):
    A = g["pos"].shape[1]
    factor, final, pen, mcd, shared, max_range, all_goals = [float(v) for v in g["params"]]
    collidable = np.zeros((A, A), bool)
    if g["lidar"].shape[-1]:  # collisions=True: agents collide with each other (navigation.py:100-101)
        collidable[np.tril_indices(A, -1)] = True
    return dict(pos=g["pos"][t], goal=g["goal"][t], radius=g["radius"], goal_radius=g["goal_radius"], shaping_in=g["shaping_in"][t],
                collidable=collidable, factor=factor, final_reward=final, penalty=pen, min_coll=mcd, shared_rew=bool(shared))


def nav_obs_parts(g, t, a):
    A = g["pos"].shape[1]
    max_range, all_goals = float(g["params"][5]), bool(g["params"][6])
    parts = [(0, g["pos"][t, a], None, 0.0), (0, g["vel"][t, a], None, 0.0)]
    for b in (range(A) if all_goals else [a]):
        parts.append((1, g["pos"][t, a], g["goal"][t, b], 0.0))
    if g["lidar"].shape[-1]:
        parts.append((2, g["lidar"][t, a], None, max_range))
    return parts


@pytest.mark.parametrize("name", NAV)
def test_navigation_reward_done_match_reference(name):
    g = load(name)
    for t in range(g["pos"].shape[0]):
        r = ho.navigation_reward(**nav_inputs(g, t))
        for k in ("dist", "on_goal", "shaping_out", "pos_rew_agent", "pos_rew", "all_goal", "final_rew", "coll_rew", "reward", "done"):
            assert np.array_equal(r[k], g[k][t]), f"{name} step {t}: {k} differs"
        assert np.array_equal(r["flags"], g["flags"][t]), f"{name} step {t}: World.collides verdicts differ"


@pytest.mark.parametrize("name", NAV)
def test_navigation_observation_matches_reference(name):
    g = load(name)
    for t in range(g["pos"].shape[0]):
        for a in range(g["pos"].shape[1]):
            assert np.array_equal(ho.gather_obs(nav_obs_parts(g, t, a)), g["obs"][t, a]), f"{name} step {t} agent {a}"


def test_navigation_penalty_obeys_the_batch_wide_verdict():
    """distance <= min_collision_distance with centres farther than r_a + r_b: a penalty only if SOME env of the batch
    has the pair's centres within r_a + r_b (World.collides, core.py:2796-2800, navigation.py:220)."""
    n = 8
    pos = np.zeros((2, n, 2), np.float32)
    pos[1, :, 0] = 1.0
    pos[1, 3, 0] = 0.203  # 0.003 outside contact
    kw = dict(goal=np.ones((2, n, 2), np.float32) * 5, radius=[0.1, 0.1], goal_radius=[0.1, 0.1],
              shaping_in=np.zeros((2, n), np.float32), collidable=np.array([[0, 0], [1, 0]], bool), factor=1.0,
              final_reward=0.01, penalty=-1.0, min_coll=0.005, shared_rew=True)
    r = ho.navigation_reward(pos=pos, **kw)
    assert not r["flags"].any() and not r["coll_rew"].any()
    pos[1, 6, 0] = 0.19  # another env in contact: now env 3 is penalised too
    r = ho.navigation_reward(pos=pos, **kw)
    assert r["flags"][1, 0] and r["coll_rew"][0, 3] == -1 and r["coll_rew"][1, 3] == -1 and r["coll_rew"][0, 6] == -1
    assert r["coll_rew"][:, [0, 1, 2, 4, 5, 7]].sum() == 0


def tr_obs_parts(g, t, a):
    parts = [(0, g["agent_pos"][t, a], None, 0.0), (0, g["agent_vel"][t, a], None, 0.0)]
    for k in range(g["pkg_state"].shape[1]):
        p = g["pkg_state"][t, k]
        parts += [(1, p[:, 0:2], g["goal_pos"][t, k], 0.0), (1, p[:, 0:2], g["agent_pos"][t, a], 0.0), (0, p[:, 2:4], None, 0.0),
                  (3, g["on_goal"][t, k], None, 0.0)]
    return parts


def test_transport_hooks_match_reference():
    g = load("transport")
    for t in range(g["rew"].shape[0]):
        r = ho.transport_reward(g["pkg_state"][t], g["pkg_dims"], g["goal_pos"][t], g["goal_radius"], g["shaping_in"][t], float(g["params"][0]))
        for k in ("dist", "on_goal", "shaping_out", "color", "rew", "done"):
            assert np.array_equal(r[k], g[k][t]), f"transport step {t}: {k} differs"
        for a in range(g["agent_pos"].shape[1]):
            assert np.array_equal(ho.gather_obs(tr_obs_parts(g, t, a)), g["obs"][t, a]), f"transport step {t} agent {a}: obs"


def spread_obs_parts(g, t, a):
    A = g["agent_pos"].shape[1]
    parts = [(0, g["agent_pos"][t, a], None, 0.0), (0, g["agent_vel"][t, a], None, 0.0)]
    parts += [(1, lp, g["agent_pos"][t, a], 0.0) for lp in g["landmark_pos"][t]]
    if g["params"][0]:
        parts += [(1, g["agent_pos"][t, o], g["agent_pos"][t, a], 0.0) for o in range(A) if o != a]
    return parts


def test_simple_spread_hooks_match_reference():
    g = load("simple_spread")
    hit = 0
    for t in range(g["rew"].shape[0]):
        r = ho.simple_spread_reward(g["agent_pos"][t], g["landmark_pos"][t], g["radius"], g["collide"])
        assert np.array_equal(r, g["rew"][t]), f"simple_spread step {t}: rew differs by {np.abs(r - g['rew'][t]).max():.3e}"
        hit += int((np.abs(r) > 3).sum())
        for a in range(g["agent_pos"].shape[1]):
            assert np.array_equal(ho.gather_obs(spread_obs_parts(g, t, a)), g["obs"][t, a]), f"simple_spread step {t} agent {a}: obs"
    assert not g["done"].any()


@pytest.mark.parametrize("name", ["joint_passage", "joint_passage_terms"])
def test_joint_passage_hooks_match_reference(name):
    g = load(name)
    A = int(g["params"][10])
    seen = dict(all_passed=0, done=0, coll=0)
    for t in range(g["joint"].shape[0]):
        r = ho.joint_passage_reward(g, t)
        for k in ("pos_rew", "rot_rew", "collision_rew", "energy_expenditure", "energy_rew", "rew", "all_passed", "done", "shaping_out"):
            assert np.array_equal(r[k], g[k][t]), f"{name} sample {t}: {k}"
        seen["all_passed"] += int(r["all_passed"].sum()); seen["done"] += int(r["done"].sum()); seen["coll"] += int((r["collision_rew"] != 0).sum())
        for a in range(A):
            obs = ho.gather_obs(ho.joint_passage_obs_parts(g, t, a))
            exp = g["obs"][t, a]
            w = obs.shape[1] - (4 if g["params"][9] else 2)  # the trailing cos / sin columns: libm vs torch's Sleef
            assert np.array_equal(obs[:, :w], exp[:, :w]), f"{name} sample {t}: observation of agent {a}"
            assert np.allclose(obs[:, w:], exp[:, w:], rtol=0, atol=2.5e-7), f"{name} sample {t}: angle vector of agent {a}"
    assert seen["all_passed"] > 0 and seen["done"] > 0 and (seen["coll"] > 0) == (name == "joint_passage_terms")


@pytest.mark.parametrize("name", ["football", "football_sparse"])
def test_football_hooks_match_reference(name):
    g = load(name)
    dense = bool(g["params"][7])
    n_done = 0
    for t in range(g["ball"].shape[0]):
        r = ho.football_reward(g, t)
        assert np.array_equal(r["done"], g["done"][t]) and np.array_equal(r["sparse"], g["sparse"][t])
        assert np.array_equal(r["rew"][0], g["rew_blue"][t]) and np.array_equal(r["rew"][1], g["rew_red"][t]), f"{name} sample {t}: reward"
        n_done += int(r["done"].sum())
        if dense:
            for k in ("dist_goal", "pos_rew", "min_dist", "pos_rew_agent"):
                assert np.array_equal(np.stack(r[k]), g[k][t]), f"{name} sample {t}: {k}"
            assert np.array_equal(np.stack(r["shaping_out"]), g["shaping_out"][t])
        nb = g["blue"].shape[1]
        for j in range(g["obs"].shape[1]):
            team, i = ("blue", j) if j < nb else ("red", j - nb)
            obs = ho.gather_obs(ho.football_obs_parts(g, t, team, i))
            assert obs.shape == g["obs"][t, j].shape and np.array_equal(obs.view(np.uint32), g["obs"][t, j].view(np.uint32)), \
                f"{name} sample {t}: observation of {team} agent {i}"
    assert n_done > 0
