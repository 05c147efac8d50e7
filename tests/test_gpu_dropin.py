"""GPU: the drop-in (vectorizedmultiagentsimulator_b200.dropin.enable) inside the reference's own Environment.step.

Two environments are built from the UNMODIFIED reference (oracle/_ref, see oracle/make_ref.sh) on the same CUDA device
with the same seed; one keeps the reference's torch World.step / cast_rays, the other is patched.  Both are driven with
the same random actions; observations, rewards, dones and the full entity state must agree (fp32 rtol 1e-5 / atol 1e-6
after one step; the stated drift bound of tests/fixtures.drift_bounds over the rollout), including across a full
reset and a partial reset_at."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")

CASES = [
    ("simple_spread", dict(n_agents=3), 1),
    ("navigation", dict(n_agents=4), 1),
    ("transport", dict(n_agents=4), 1),
    ("transport", dict(n_agents=4), 5),
    ("joint_passage", dict(), 1),
    ("football", dict(ai_red_agents=False), 1),
    ("flocking", dict(), 1),
    ("balance", dict(), 1),
    ("give_way", dict(), 1),
    ("simple_reference", dict(), 1),  # world.dim_c = 10: communication actions (environment.py:721-754) and state (core.py:2909-2912)
]


def _vmas():
    if not os.path.isdir(os.path.join(REF, "vmas")):
        pytest.skip("oracle/_ref (the reference install) did not travel to this box")
    for p in (REF, os.path.join(ROOT, "oracle", "gym_stub")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import vmas

    return vmas


def _make(vmas, scenario, kw, n, substeps):
    env = vmas.make_env(scenario, num_envs=n, device="cuda", seed=0, continuous_actions=True, **kw)
    if substeps > 1:
        env.world._substeps = substeps
        env.world._sub_dt = env.world._dt / substeps
    return env


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


@pytest.mark.parametrize("scenario,kw,substeps", CASES)
def test_environment_step_matches_reference(scenario, kw, substeps):
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n, steps = 96, 30
    ref = _make(vmas, scenario, kw, n, substeps)
    env = _make(vmas, scenario, kw, n, substeps)
    b = dropin.enable(env)
    jointed = bool(getattr(env.world, "_joints", {}))
    pos_tol, vel_tol = (1e-2, 2e-1) if jointed else (1e-4, 1e-3)
    # Stiff jointed worlds (joint_force 900, 10 substeps, inertia ~5e-4) are chaotic: one input ulp grows to O(0.1)
    # within ~15 steps.  Their bound is therefore calibrated on the reference itself: a twin reference env whose
    # positions are moved by ONE fp32 ulp after every (re)synchronisation shows how many envs leave the tolerance by
    # sensitivity alone; the drop-in may not lose more than twice that many (+2).
    twin = _make(vmas, scenario, kw, n, substeps) if jointed else None

    def nudge():
        if twin is not None:
            for e in twin.world.entities:
                if e.movable:
                    e.state.pos = torch.nextafter(e.state.pos, torch.full_like(e.state.pos, float("inf")))

    nudge()
    torch.manual_seed(7)
    for t in range(steps):
        if t == 12:
            o1, o2 = ref.reset(seed=3), env.reset(seed=3)
            assert torch.allclose(_flat(o1), _flat(o2), rtol=1e-5, atol=1e-6)
            if twin is not None:
                twin.reset(seed=3)
                nudge()
        if t == 20:  # partial reset: in-place writes through the bound views (the spawn draws from the global RNG)
            torch.manual_seed(11)
            ref.reset_at(5)
            torch.manual_seed(11)
            env.reset_at(5)
            if twin is not None:
                torch.manual_seed(11)
                twin.reset_at(5)
            torch.manual_seed(12)
        acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in ref.agents]
        if ref.world.dim_c > 0:  # communication columns in [0, 1] behind the physical ones (environment.py:545-557)
            acts = [torch.cat([x, torch.rand(n, ref.world.dim_c, device="cuda")], -1) if not a.silent else x for x, a in zip(acts, ref.agents)]
        o1, r1, d1, _ = ref.step([a.clone() for a in acts])
        o2, r2, d2, _ = env.step([a.clone() for a in acts])
        s1, s2 = _state(ref), _state(env)
        if twin is not None:
            twin.step([a.clone() for a in acts])
            if t not in (0, 12):
                s3 = _state(twin)
                lost = lambda a, b: int(((a[..., 0:2] - b[..., 0:2]).abs().amax(dim=(0, 2)) > pos_tol).sum())
                assert lost(s1, s2) <= 2 * lost(s1, s3) + 2, f"{scenario}: {lost(s1, s2)} envs diverged at step {t}, ulp twin {lost(s1, s3)}"
                assert torch.equal(d1, d2)
                continue
        if t == 0:
            assert torch.allclose(s1, s2, rtol=1e-5, atol=1e-6), f"{scenario}: state after one step {float((s1 - s2).abs().max()):.3e}"
            assert torch.allclose(_flat(o1), _flat(o2), rtol=1e-5, atol=1e-5), f"{scenario}: obs after one step"
        assert float((s1[..., 0:2] - s2[..., 0:2]).abs().max()) <= pos_tol, f"{scenario}: pos drift at step {t}"
        assert float((s1[..., 2:4] - s2[..., 2:4]).abs().max()) <= vel_tol, f"{scenario}: vel drift at step {t}"
        assert torch.equal(d1, d2), f"{scenario}: done flags differ at step {t}"
        assert torch.allclose(_flat(r1), _flat(r2), rtol=1e-3, atol=10 * vel_tol), f"{scenario}: reward at step {t}"
        if ref.world.dim_c > 0:  # _update_comm_state (core.py:2909-2912): state.c = action.c of every non-silent agent
            for a1, a2 in zip(ref.world.agents, env.world.agents):
                assert a1.state.c is not None and torch.equal(a1.state.c, a2.state.c), f"{scenario}: comm state of {a1.name} at step {t}"
    assert b.n_steps == steps
    assert b.n_fallback["set_action"] == 0, "an action took the reference's torch _set_action"
    if ref.world.dim_c > 0:
        assert b.n_fused_actions >= steps * len(env.agents)
    dropin.disable(env)


def test_lidar_measure_uses_epilogue_and_falls_back():
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n = 64
    ref = _make(vmas, "navigation", dict(n_agents=4), n, 1)
    env = _make(vmas, "navigation", dict(n_agents=4), n, 1)
    b = dropin.enable(env)
    acts = [torch.zeros(n, 2, device="cuda") for _ in ref.agents]
    ref.step(acts)
    env.step(acts)
    for a1, a2 in zip(ref.agents, env.agents):
        m1, m2 = a1.sensors[0].measure(), a2.sensors[0].measure()
        assert torch.allclose(m1, m2, rtol=1e-5, atol=1e-6)
    assert b._state_untouched()
    # moving an agent invalidates the epilogue hand-off: measure must see the new position (standalone cast_rays)
    p = env.agents[0].state.pos.clone()
    p[:, 0] += 0.05
    env.agents[0].set_pos(p, batch_index=None)
    ref.agents[0].set_pos(p.clone(), batch_index=None)
    for a1, a2 in zip(ref.agents, env.agents):
        assert torch.allclose(a1.sensors[0].measure(), a2.sensors[0].measure(), rtol=1e-5, atol=1e-6)
    # in-place write through the bound view
    env.agents[1].state.pos[:, 1] += 0.03
    ref.agents[1].state.pos[:, 1] += 0.03
    for a1, a2 in zip(ref.agents, env.agents):
        assert torch.allclose(a1.sensors[0].measure(), a2.sensors[0].measure(), rtol=1e-5, atol=1e-6)


def test_cpu_world_is_rejected():
    from vectorizedmultiagentsimulator_b200 import _lib, dropin

    vmas = _vmas()
    env = vmas.make_env("simple_spread", num_envs=4, device="cpu", seed=0, continuous_actions=True)
    with pytest.raises(_lib.VmasB200Error):
        dropin.enable(env)


@pytest.mark.parametrize("scenario,kw", [("navigation", dict(n_agents=4)), ("transport", dict(n_agents=4)),
                                         ("joint_passage", dict()), ("football", dict(ai_red_agents=False)),
                                         ("pollock", dict(lidar=True))])
def test_distance_queries_match_reference(scenario, kw):
    """world.get_distance / is_overlapping / get_distance_from_point of a patched world (what the scenario's reward
    code calls) against the reference's torch implementation on the very same state tensors, for every entity pair."""
    import itertools

    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n = 96
    env = _make(vmas, scenario, kw, n, 1)
    b = dropin.enable(env)
    torch.manual_seed(3)
    for _ in range(5):
        env.step([env.get_random_action(a) for a in env.agents])
    world = env.world
    ents = list(world.entities)
    flips = 0
    for a, c in itertools.permutations(ents, 2):
        d_ref = b.orig_get_distance(a, c)
        d = world.get_distance(a, c)
        assert d.shape == d_ref.shape and torch.allclose(d, d_ref, rtol=1e-5, atol=1e-6), (a.name, c.name)
        o_ref, o = b.orig_is_overlapping(a, c), world.is_overlapping(a, c)
        assert o.dtype == torch.bool and o.shape == o_ref.shape
        diff = o != o_ref
        flips += int(diff.sum())
        assert not (diff & (d_ref.abs() > 1e-6)).any(), (a.name, c.name)  # only a distance within rounding of 0 may flip
    assert flips <= 2
    pts = torch.rand(n, 2, device="cuda") * 2 - 1
    for e in ents:
        assert torch.allclose(world.get_distance_from_point(e, pts), b.orig_get_distance_from_point(e, pts), rtol=1e-5, atol=1e-6)
    # env_index and autograd keep the reference's own path
    a0, a1 = env.agents[0], env.agents[1]  # spheres (the reference's env_index handling breaks for box / sphere)
    assert torch.equal(world.get_distance(a0, a1, env_index=3), b.orig_get_distance(a0, a1, 3))
    dropin.disable(env)
    assert "get_distance" not in world.__dict__



@pytest.mark.parametrize("scenario,kw", [("navigation", dict(n_agents=4)), ("transport", dict(n_agents=4)),
                                         ("joint_passage", dict()), ("football", dict(ai_red_agents=False))])
def test_static_io_equals_fresh_buffers_bitwise(scenario, kw):
    """enable(env, static_io=True) steps in place on persistent buffers; against a drop-in env that binds fresh output
    buffers every step, everything Environment.step returns and the whole world state must be bit-identical at every
    step, across a full reset and a partial reset_at; the state tensors keep their addresses."""
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n, steps = 96, 30
    e1, e2 = _make(vmas, scenario, kw, n, 1), _make(vmas, scenario, kw, n, 1)
    dropin.enable(e1)
    b2 = dropin.enable(e2, static_io=True)

    def leaves(x):
        if isinstance(x, dict):
            return [t for k in sorted(x) for t in leaves(x[k])]
        if isinstance(x, (list, tuple)):
            return [t for v in x for t in leaves(v)]
        return [x]

    torch.manual_seed(7)
    ptrs = None
    for t in range(steps):
        if t == 10:
            o1, o2 = e1.reset(seed=3), e2.reset(seed=3)
            assert all(torch.equal(a, c) for a, c in zip(leaves(o1), leaves(o2)))
            ptrs = None
        if t == 20:
            # the reference keeps ONE host RNG state for all Environments (class attribute vmas_random_state,
            # environment.py:58): both resets must start from the same one (joint_passage draws a randperm from it)
            rs = type(e1).vmas_random_state
            saved = list(rs)
            torch.manual_seed(11)
            e1.reset_at(5)
            rs[:] = saved
            torch.manual_seed(11)
            e2.reset_at(5)
            torch.manual_seed(12)
        acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in e1.agents]
        r1 = e1.step([a.clone() for a in acts])
        r2 = e2.step([a.clone() for a in acts])
        for i, (a, c) in enumerate(zip(leaves(r1), leaves(r2))):
            assert torch.equal(a.nan_to_num(), c.nan_to_num()), f"{scenario}: output {i} differs at step {t}"
        assert torch.equal(_state(e1), _state(e2)), f"{scenario}: state differs at step {t}"
        now = [a.state.pos.data_ptr() for a in e2.agents]
        assert ptrs is None or now == ptrs, "state views moved"
        ptrs = now
    assert b2.n_steps == steps


def test_set_action_matches_reference_and_keeps_its_assertions():
    """Environment._set_action of a patched env (one launch per agent, vmasb200_set_action) against the reference's:
    agent.action.u bit-identical for in-range and clamped actions; a NaN action and - with clamp_action off - an action
    out of range still raise AssertionError inside Environment.step, before the world moves."""
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n = 200
    for scenario, kw in (("navigation", dict(n_agents=4)), ("transport", dict(n_agents=4)), ("football", dict(ai_red_agents=False))):
        ref = _make(vmas, scenario, kw, n, 1)
        env = _make(vmas, scenario, kw, n, 1)
        b = dropin.enable(env)
        assert "_set_action" in env.__dict__
        torch.manual_seed(5)
        for scale in (1.0, 1.7):
            for a1, a2 in zip(ref.agents, env.agents):
                act = (torch.rand(n, a1.action_size, device="cuda") * 2 - 1) * a1.u_range * scale
                if not ref.clamp_action and scale > 1:
                    continue
                ref._set_action(act.clone(), a1)
                env._set_action(act.clone(), a2)
                assert torch.equal(a1.action.u, a2.action.u), (scenario, scale)
                # a strided view of a wider tensor is read in place
                wide = torch.zeros(n, a1.action_size + 3, device="cuda")
                wide[:, 1:1 + a1.action_size] = act
                env._set_action(wide[:, 1:1 + a1.action_size], a2)
                assert torch.equal(a1.action.u, a2.action.u)
        b.check_actions()
        before = _state(env).clone()
        acts = [torch.zeros(n, a.action_size, device="cuda") for a in env.agents]
        acts[1][7, 0] = float("nan")
        with pytest.raises(AssertionError):
            env.step(acts)
        assert torch.equal(_state(env), before), "the world moved although an action was NaN"
        acts[1][7, 0] = 0.0
        env.step(acts)  # and it recovers
        dropin.disable(env)
        assert "_set_action" not in env.__dict__


@pytest.mark.parametrize("scenario,kw", [("flocking", dict()), ("navigation", dict(n_agents=4)), ("pollock", dict(lidar=True))])
def test_scalar_cast_ray_matches_reference(scenario, kw):
    """World.cast_ray (core.py:1627-1659, the scalar twin Lidar.measure(vectorized=False) uses) through the same launch
    with one ray per env, against the reference's own scalar implementation on the same CUDA state; and
    Lidar.measure(vectorized=False) of the patched world against the reference's."""
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n = 200
    ref, env = _make(vmas, scenario, kw, n, 1), _make(vmas, scenario, kw, n, 1)
    b = dropin.enable(env, strict=True)
    torch.manual_seed(3)
    for _ in range(4):
        acts = [torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range for a in ref.agents]
        ref.step([a.clone() for a in acts]); env.step([a.clone() for a in acts])
    for er, ee in zip(ref.world.entities, env.world.entities):  # identical state on both sides
        for f in ("pos", "vel", "rot", "ang_vel"):
            setattr(ee.state, f, getattr(er.state, f).clone())
    checked = 0
    for ar, ae in zip(ref.agents, env.agents):
        for sr, se in zip(ar.sensors or [], ae.sensors or []):
            for k in range(3):
                ang = torch.rand(n, device="cuda") * 6.2831853
                d_ref = ref.world.cast_ray(ar, ang, sr._max_range, entity_filter=sr.entity_filter)
                d_got = env.world.cast_ray(ae, ang.clone(), se._max_range, entity_filter=se.entity_filter)
                assert d_got.shape == d_ref.shape == (n,)
                bad = (d_got - d_ref).abs() > 1e-6 + 1e-5 * d_ref.abs()
                assert int(bad.sum()) == 0 and float((d_got - d_ref).abs().max()) <= 1e-5, f"{scenario} {ar.name}: {float((d_got - d_ref).abs().max()):.3e}"
                checked += 1
            m_ref, m_got = sr.measure(vectorized=False), se.measure(vectorized=False)
            assert torch.allclose(m_got, m_ref, rtol=1e-5, atol=1e-5), f"{scenario} {ar.name}: Lidar.measure(vectorized=False)"
    assert checked > 0 and b.n_fallback["cast_rays"] == 0
    dropin.disable(env)
    assert "cast_ray" not in env.world.__dict__


def test_fallbacks_warn_once_and_strict_raises():
    """A secondary call the library does not cover takes the reference's torch code with ONE warning and a count;
    under strict=True it raises.  (World.step itself never falls back.)"""
    import warnings

    import torch

    from vectorizedmultiagentsimulator_b200 import dropin
    from vectorizedmultiagentsimulator_b200._lib import VmasB200Error

    vmas = _vmas()
    n = 32
    env = _make(vmas, "navigation", dict(n_agents=2), n, 1)
    b = dropin.enable(env)
    acts = [torch.zeros(n, a.action_size, device="cuda", dtype=torch.float64) for a in env.agents]  # fp64 actions: not covered
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        env.step(acts)
        env.step(acts)
    msgs = [str(x.message) for x in w if "vmas-b200" in str(x.message)]
    assert len(msgs) == 1 and "set_action" in msgs[0], msgs
    assert b.n_fallback["set_action"] == 4 and b.n_steps == 2
    b.strict = True
    with pytest.raises(VmasB200Error, match="strict"):
        env.step(acts)


def test_invalidate_picks_up_attribute_changes():
    """The reference re-reads entity attributes every step; the binding reads them when the plan is built.  After
    binding.invalidate() a changed mass / max_speed is what the next step uses (and without it the old plan keeps running)."""
    import torch

    from vectorizedmultiagentsimulator_b200 import dropin

    vmas = _vmas()
    n = 64
    ref = _make(vmas, "transport", dict(n_agents=4), n, 1)
    env = _make(vmas, "transport", dict(n_agents=4), n, 1)
    b = dropin.enable(env)
    torch.manual_seed(11)
    acts = [[torch.rand(n, 2, device="cuda") * 2 - 1 for _ in env.agents] for _ in range(8)]
    for t in range(8):
        if t == 4:  # heavier first agent, slower package - the same change on both worlds
            for w in (ref.world, env.world):
                w.agents[0]._mass = 5.0
                w.agents[1]._max_speed = 0.05
            builds = b.n_plan_builds
            b.invalidate()
        ref.step([a.clone() for a in acts[t]])
        env.step([a.clone() for a in acts[t]])
        assert torch.allclose(_state(ref), _state(env), rtol=1e-5, atol=1e-6), f"step {t}"
    assert b.n_plan_builds == builds + 1
