"""GPU: the fused reset path (dropin.enable(env, fused_reset=True)) against the reference's own
Environment.reset / reset_at on the SAME random state.

The reference rebinds a fresh tensor for every state field of every entity on a reset - also on reset_at(i)
(core.py:285-295: torch.where over the whole batch) - and its spawn sampler (utils.py:274-319) runs ~10 small launches and
two host synchronisations per rejection round.  Fused: World.reset zeroes the SoA buffers in place (vmasb200_reset_rows),
Entity.set_* write through the bound views, a rejection round is one launch (vmasb200_spawn_select) on the reference's own
uniform_ draws.  Required: entity state, observations and everything a later step produces BIT-identical to the
unpatched reference on the same seed; no state field re-packed at the step after a reset (n_copies unchanged); the
kernels really ran."""
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


def _state(env):
    import torch

    return torch.stack([torch.cat([e.state.pos, e.state.vel, e.state.rot, e.state.ang_vel], -1) for e in env.world.entities])


def _flat(x):
    import torch

    if isinstance(x, dict):
        return torch.cat([_flat(v) for _, v in sorted(x.items())], -1)
    if isinstance(x, (list, tuple)):
        return torch.cat([_flat(v) for v in x], -1)
    return x.reshape(x.shape[0], -1).float()


CASES = [("navigation", dict(n_agents=4)), ("transport", dict(n_agents=4)), ("simple_spread", dict(n_agents=3)),
         ("joint_passage", dict()), ("football", dict(ai_red_agents=False)), ("flocking", dict()), ("balance", dict())]


@pytest.mark.parametrize("scenario,kw", CASES)
def test_fused_reset_equals_reference_reset_bitwise(scenario, kw):
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n = 512
    mk = lambda: vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True, **kw)
    ref, twin, env = mk(), mk(), mk()
    dropin.enable(twin)                       # the drop-in WITHOUT the fused reset: same physics, reference reset path
    b = dropin.enable(env, fused_reset=True)
    rs = type(ref).vmas_random_state          # the Environment's random state is class-level (environment.py:58-64)

    def on_all(fn):
        """fn on the three envs from the same random state (environment + CUDA generator)."""
        saved, cuda = list(rs), torch.cuda.get_rng_state()
        outs = []
        for e in (ref, twin, env):
            rs[:] = saved
            torch.cuda.set_rng_state(cuda)
            outs.append(fn(e))
        return outs

    def same(what):
        s_ref, s_twin, s_env = _state(ref), _state(twin), _state(env)
        assert torch.equal(s_env, s_twin), f"{scenario}: {what}: fused reset vs reference reset (same physics)"
        return s_ref

    torch.manual_seed(3)
    o = on_all(lambda e: e.reset(seed=11))
    s_ref = same("state after reset(seed=11)")
    assert torch.equal(_state(env), s_ref), f"{scenario}: state after reset(seed=11) vs the unpatched reference"
    assert torch.equal(_flat(o[2]), _flat(o[0])) or torch.allclose(_flat(o[2]), _flat(o[0]), rtol=1e-5, atol=1e-5)
    copies = None
    for t in range(12):
        acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in ref.agents]
        for e in (twin, env):
            e.step([a.clone() for a in acts])
        same(f"state at step {t}")
        if t == 0:
            copies = b.n_copies  # the first step after enable packed the state once
        if t == 3:
            on_all(lambda e: [e.reset_at(i) for i in (5, 0, n - 1, 77)])
            same("state after reset_at")
        if t == 7:
            on_all(lambda e: e.reset(seed=4))
            same("state after reset(seed=4)")
    # no state field was re-packed after any of the resets, unless a reset changed the plan itself (new buffers)
    plan_changes = b.n_plan_builds - 1
    assert b.n_copies == copies or plan_changes > 0, f"{scenario}: {b.n_copies - copies} state fields re-packed after a reset"
    assert b.n_reset_launches >= 6, "World.reset did not take vmasb200_reset_rows"
    if scenario in ("navigation", "transport", "flocking"):
        assert b.n_spawn_launches > 0, "the spawn sampler did not take vmasb200_spawn_select"
    twin_b = twin.world._vmasb200
    assert twin_b.n_copies > copies, "the reference reset path no longer rebinds state tensors: the comparison is moot"
    dropin.disable(env)
    dropin.disable(twin)
    import importlib

    su = importlib.import_module("vmas.simulator.utils").ScenarioUtils
    assert su.find_random_pos_for_entity.__name__ == "find_random_pos_for_entity" and "fused" not in su.find_random_pos_for_entity.__qualname__
