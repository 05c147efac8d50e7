"""GPU: a whole Environment.step of the reference captured as ONE CUDA graph (vectorizedmultiagentsimulator_b200.graph).

The graphed environment and an eager twin (same drop-in, same fused hooks, same in-place stepping) get the same
actions; observations, rewards, dones, infos and the full world state must be BIT-identical at every step - across a
full reset (state tensors and goal positions are rebound by the scenario: the graph's inputs must be repaired) and a
partial reset_at (in-place writes) - and the graph must have been captured exactly once."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


def _vmas():
    if not os.path.isdir(os.path.join(REF, "vmas")):
        pytest.skip("oracle/_ref (the reference install) did not travel to this box")
    for p in (REF, os.path.join(ROOT, "oracle", "gym_stub")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import vmas

    return vmas


def _leaves(x):
    if isinstance(x, dict):
        return [t for k in sorted(x) for t in _leaves(x[k])]
    if isinstance(x, (list, tuple)):
        return [t for v in x for t in _leaves(v)]
    return [x]


def _state(env):
    import torch

    return torch.stack([torch.cat([e.state.pos, e.state.vel, e.state.rot, e.state.ang_vel], -1) for e in env.world.entities])


@pytest.mark.parametrize("scenario,kw", [("navigation", dict(n_agents=4)), ("transport", dict(n_agents=4)),
                                         ("navigation", dict(n_agents=3, shared_rew=False)), ("simple_spread", dict(n_agents=3)),
                                         ("joint_passage", dict()), ("joint_passage", dict(fixed_passage=False))])
def test_graphed_step_equals_eager_step_bitwise(scenario, kw):
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin
    from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment

    vmas = _vmas()
    n, steps = 160, 30
    mk = lambda: vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True, max_steps=25, **kw)
    e1, e2 = mk(), mk()
    dropin.enable(e1, static_io=True, hooks=True)
    before = (_state(e2).clone(), e2.steps.clone(), type(e2).vmas_random_state[0].clone())
    g = GraphedEnvironment(e2, warmup=2)
    # the warm-up steps of the capture leave no trace: state, step counter and the environment's random state are restored
    assert torch.equal(_state(e2), before[0]) and torch.equal(e2.steps, before[1])
    assert torch.equal(type(e2).vmas_random_state[0], before[2])
    rs = type(e1).vmas_random_state
    saved = list(rs)
    e1.reset(seed=1)
    rs[:] = saved
    e2.reset(seed=1)
    torch.manual_seed(7)
    for t in range(steps):
        if t == 10:
            saved = list(rs)
            o1 = e1.reset(seed=3)
            rs[:] = saved
            o2 = g.reset(seed=3)
            assert all(torch.equal(a, c) for a, c in zip(_leaves(o1), _leaves(o2)))
        if t == 20:
            saved = list(rs)
            torch.manual_seed(11)
            e1.reset_at(5)
            rs[:] = saved
            torch.manual_seed(11)
            g.reset_at(5)
            torch.manual_seed(12)
        acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in e1.agents]
        if t % 7 == 3:  # agents onto their goals / packages onto theirs in a few envs (in-place writes)
            with torch.no_grad():
                for env in (e1, e2):
                    if scenario == "navigation":
                        for a in env.world.agents:
                            a.state.pos[:6] = a.goal.state.pos[:6]
                    elif scenario == "simple_spread":
                        env.world.agents[1].state.pos[:6] = env.world.agents[0].state.pos[:6] + 0.1
                    elif scenario == "transport":
                        for p in env.scenario.packages:
                            p.state.pos[:6] = p.goal.state.pos[:6]
                    elif scenario == "joint_passage":
                        env.scenario.joint.landmark.state.pos[:6, 1] = 0.3
                    else:
                        env.scenario.ball.state.pos[:6, 0] = 5.0  # beyond the goal line: a score
        r1 = e1.step([a.clone() for a in acts])
        r2 = g.step(acts)
        l1, l2 = _leaves(r1), _leaves(r2)
        assert len(l1) == len(l2)
        for i, (a, c) in enumerate(zip(l1, l2)):
            assert a.dtype == c.dtype and torch.equal(a.nan_to_num(), c.nan_to_num()), f"{scenario}: output {i} differs at step {t}"
        assert torch.equal(_state(e1), _state(e2)), f"{scenario}: state differs at step {t}"
        assert torch.equal(e1.steps, e2.steps)
    # joint_passage's resets may change collision filters (joint_passage.py:615-622: with fixed_passage=False the gap moves):
    # a new plan, hence a new capture - whose warm-up must leave no trace either (the per-step equality above covers it)
    assert (g.n_captures == 1 or scenario == "joint_passage") and g.n_replays == steps and g.n_input_fixes > 0


def test_recapture_after_a_plan_change_leaves_no_trace():
    """A reset that changes the world's static description (here: an agent gets heavier) makes GraphedEnvironment.step
    re-capture; the warm-up steps of that re-capture run on the freshly reset env and must not move it: the graphed env
    stays bit-identical to the eager twin afterwards."""
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin
    from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment

    vmas = _vmas()
    n = 96
    mk = lambda: vmas.make_env("navigation", num_envs=n, device="cuda", seed=0, continuous_actions=True, n_agents=4)
    e1, e2 = mk(), mk()
    dropin.enable(e1, static_io=True, hooks=True)
    g = GraphedEnvironment(e2, warmup=3)
    rs = type(e1).vmas_random_state
    torch.manual_seed(3)
    for t in range(12):
        if t == 5:
            for env in (e1, e2):
                env.world.agents[1]._mass = 2.5  # entity attributes are part of the plan (like the collision filters a reset may change)
            saved = list(rs)
            e1.reset(seed=9)
            rs[:] = saved
            g.reset(seed=9)
        acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in e1.agents]
        r1 = e1.step([a.clone() for a in acts])
        r2 = g.step(acts)
        for i, (a, c) in enumerate(zip(_leaves(r1), _leaves(r2))):
            if not torch.equal(a.nan_to_num(), c.nan_to_num()):
                d = (a.float() - c.float()).abs()
                cols = d.reshape(d.shape[0], -1).amax(0)
                raise AssertionError(f"output {i} differs at step {t}: columns {cols.nonzero().flatten().tolist()} max {float(d.max()):.3e}; "
                                     f"state equal {torch.equal(_state(e1), _state(e2))}")
        assert torch.equal(_state(e1), _state(e2)) and torch.equal(e1.steps, e2.steps), f"state differs at step {t}"
    assert g.n_captures == 2, "the plan change did not lead to a re-capture"


def test_graphed_step_keeps_the_action_assertions():
    import torch

    from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment

    vmas = _vmas()
    n = 64
    env = vmas.make_env("navigation", num_envs=n, device="cuda", seed=0, continuous_actions=True, n_agents=4)
    g = GraphedEnvironment(env)
    acts = [torch.zeros(n, a.action_size, device="cuda") for a in env.agents]
    g.step(acts)
    acts[2][9, 1] = float("nan")
    with pytest.raises(AssertionError):
        g.step(acts)
    acts[2][9, 1] = 0.0
    g.step(acts)


def test_graph_needs_fused_hooks():
    from vectorizedmultiagentsimulator_b200 import _lib
    from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment

    vmas = _vmas()
    env = vmas.make_env("football", num_envs=8, device="cuda", seed=0, continuous_actions=True, ai_red_agents=False)
    with pytest.raises(_lib.VmasB200Error):
        GraphedEnvironment(env)


def test_graph_refuses_what_the_scenario_itself_makes_uncapturable():
    """football's ball_action_script (reference code, football.py:1617-1677) copies Python floats to the device every
    step: the capture must fail with the library's error, and the eager fused hooks keep working."""
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin
    from vectorizedmultiagentsimulator_b200._lib import VmasB200Error
    from vectorizedmultiagentsimulator_b200.graph import GraphedEnvironment

    vmas = _vmas()
    env = vmas.make_env("football", num_envs=64, device="cuda", seed=0, continuous_actions=True, ai_red_agents=False)
    with pytest.raises(VmasB200Error, match="cannot be captured"):
        GraphedEnvironment(env, warmup=1)
    torch.cuda.synchronize()
    b = dropin.enable(env, hooks=True)
    acts = [torch.zeros(64, a.action_size, device="cuda") for a in env.agents]
    for _ in range(3):
        env.step(acts)
    assert b.hooks.n_fused["reward"] >= 3 and b.hooks.n_fallback["reward"] == 0
