"""GPU: vmasb200_set_action_discrete (environment.py:660-710) through the C ABI against the reference's
golden outputs, bit for bit, and inside the reference Environment (drop-in) against the unpatched Environment on the
same device, including the out-of-bounds assertion."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from tests.test_oracle_actions import cases
from vectorizedmultiagentsimulator_b200 import _lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("dtype", ["int64", "int32"])
@pytest.mark.parametrize("c", list(cases()), ids=lambda c: f"multi{int(c['multi'])}_nvec{'x'.join(map(str, c['nvec']))}")
def test_c_abi_reproduces_the_reference(c, dtype):
    import torch

    L = _lib.lib()
    dev = torch.device("cuda:0")
    a = torch.from_numpy(c["action"].astype(dtype)).to(dev)
    k = len(c["nvec"])
    n = a.shape[0]
    u = torch.empty(n, k, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    nv = (C.c_int32 * k)(*[int(v) for v in c["nvec"]])
    rng = (C.c_float * k)(*[float(v) for v in c["u_range"]])
    mul = (C.c_float * k)(*[float(v) for v in c["u_mult"]])
    _lib.check(L.vmasb200_set_action_discrete(a.data_ptr(), 0 if dtype == "int64" else 1, a.stride(0), a.stride(1), k,
                                              int(c["multi"]), nv, rng, mul, u.data_ptr(), flags.data_ptr(), n,
                                              _lib.raw_stream(torch, dev)), "vmasb200_set_action_discrete")
    torch.cuda.synchronize()
    assert int(flags.item()) == 0
    assert np.array_equal(u.cpu().numpy().view(np.uint32), c["u"].view(np.uint32))
    # one component out of bounds -> flag bit 2, the other envs unchanged
    a2 = a.clone()
    a2[3, 0] = int(np.prod(c["nvec"])) if not c["multi"] else int(c["nvec"][0])
    _lib.check(L.vmasb200_set_action_discrete(a2.data_ptr(), 0 if dtype == "int64" else 1, a2.stride(0), a2.stride(1), k,
                                              int(c["multi"]), nv, rng, mul, u.data_ptr(), flags.data_ptr(), n,
                                              _lib.raw_stream(torch, dev)), "vmasb200_set_action_discrete")
    torch.cuda.synchronize()
    assert int(flags.item()) == 4
    keep = np.arange(n) != 3
    assert np.array_equal(u.cpu().numpy()[keep].view(np.uint32), c["u"][keep].view(np.uint32))


@pytest.mark.parametrize("multi", [False, True])
@pytest.mark.parametrize("scenario", ["navigation", "transport"])
def test_dropin_discrete_actions_in_the_reference_environment(scenario, multi):
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "vmas")):
        pytest.skip("oracle/_ref (the reference, installed by oracle/make_ref.sh) is not present")
    sys.path[:0] = [p for p in (ref_dir, os.path.join(ROOT, "oracle", "gym_stub")) if p not in sys.path]
    import torch
    import vmas

    from vectorizedmultiagentsimulator_b200 import dropin

    n = 512
    mk = lambda: vmas.make_env(scenario, num_envs=n, device="cuda:0", seed=3, continuous_actions=False,  # noqa: E731
                               multidiscrete_actions=multi)
    ref, env = mk(), mk()
    b = dropin.enable(env)
    try:
        g = torch.Generator(device="cuda:0").manual_seed(5)
        for _ in range(4):
            acts = []
            for a in ref.world.agents:
                nvec = a.discrete_action_nvec
                if multi:
                    acts.append(torch.stack([torch.randint(0, m, (n,), generator=g, device="cuda:0") for m in nvec], dim=-1))
                else:
                    acts.append(torch.randint(0, int(np.prod(nvec)), (n, 1), generator=g, device="cuda:0"))
            o1, r1, d1, _ = ref.step([a.clone() for a in acts])
            o2, r2, d2, _ = env.step([a.clone() for a in acts])
            for a1, a2 in zip(ref.world.agents, env.world.agents):
                assert torch.equal(a1.action.u, a2.action.u), "agent.action.u differs"
            for x, y in zip(o1, o2):
                assert torch.allclose(x, y, rtol=1e-5, atol=1e-6)
            assert torch.equal(d1, d2)
        assert b._act_flags is not None, "the discrete actions did not take vmasb200_set_action_discrete"
        bad = [a.clone() for a in acts]
        bad[0][7, 0] = 99
        with pytest.raises(AssertionError):
            env.step(bad)
    finally:
        dropin.disable(env)


@pytest.mark.parametrize("clamp", [True, False])
def test_comm_actions_and_action_noise_in_the_environment(clamp):
    """Environment._set_action with communication columns and u / c noise (environment.py:633-645, 711-754) through
    vmasb200_set_action_ex: the patched env draws the noise with the reference's own torch.randn calls, so with the same
    generator state action.u, action.c and - after World.step - state.c (core.py:2909-2912) equal the reference's BIT for
    bit; and the reference's assertions (NaN, physical range, comm range) still fire."""
    import os
    import sys

    import torch

    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "vmas")):
        pytest.skip("oracle/_ref (the reference install) did not travel to this box")
    for p in (ref_dir, os.path.join(ROOT, "oracle", "gym_stub")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import vmas

    from vectorizedmultiagentsimulator_b200 import dropin

    n = 300
    mk = lambda: vmas.make_env("simple_reference", num_envs=n, device="cuda", seed=0, continuous_actions=True, clamp_actions=clamp)
    ref, env = mk(), mk()
    for e in (ref, env):
        for a in e.agents:
            a.action._u_noise = 0.05
            a._c_noise = 0.02
    b = dropin.enable(env, strict=True)  # strict: a call that is not covered would raise, not fall back
    torch.manual_seed(5)
    for t in range(6):
        acts = []
        for a in ref.agents:
            phys = torch.rand(n, a.action_size, device="cuda") * 2 * a.u_range - a.u_range
            comm = torch.rand(n, ref.world.dim_c, device="cuda")
            if clamp and t == 2:  # out-of-range physical values the clamp must bring back (comm actions are never clamped)
                phys[:7] *= 3.0
            acts.append(torch.cat([phys, comm], -1))
        st = torch.cuda.get_rng_state()
        torch.manual_seed(100 + t)
        ref.step([x.clone() for x in acts])
        torch.cuda.set_rng_state(st)
        torch.manual_seed(100 + t)
        env.step([x.clone() for x in acts])
        for a1, a2 in zip(ref.agents, env.agents):
            assert torch.equal(a1.action.u, a2.action.u), f"action.u of {a1.name} at step {t}"
            assert torch.equal(a1.action.c, a2.action.c), f"action.c of {a1.name} at step {t}"
            assert torch.equal(a1.state.c, a2.state.c), f"state.c of {a1.name} at step {t}"
    assert b.n_fused_actions == 6 * len(env.agents) and b.n_fallback["set_action"] == 0
    if True:
        bad = [torch.cat([torch.zeros(n, a.action_size, device="cuda"), torch.full((n, ref.world.dim_c), 0.5, device="cuda")], -1) for a in env.agents]
        bad[1][3, 5] = 1.5
        with pytest.raises(AssertionError, match="Comm actions"):
            env.step(bad)
