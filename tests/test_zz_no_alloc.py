"""GPU: vmasb200_navigation_reward_static (include/vmasb200.h) - the allocation-free variant of vmasb200_navigation_reward
whose flag-reduction scratch is a file-scope `static __device__` array of the library.  It has no counterpart in the
reference.  Through the C ABI, on the golden inputs of the reference's
navigation scenario: two calls in a row must give the bits of the existing entry point (which gets a scratch tensor of
its own) - the hook tests' bar is bit equality - and the stated limit must be rejected with a message.  Collected last
(file name), runs alone in a fresh process as well as after the other tests."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def _soa(torch, x, pad=5):
    n, k = x.shape
    buf = torch.zeros(k, n + pad, device="cuda")
    buf[:, :n] = torch.from_numpy(np.ascontiguousarray(x.T)).cuda()
    return buf[:, :n].t()


def _call(torch, L, H, g, t, static):
    """One reward call on step t of the golden rollout; returns (float outputs, byte outputs) as numpy."""
    from vectorizedmultiagentsimulator_b200 import _lib

    A, n = g["pos"].shape[1:3]
    factor, final, pen, mcd, shared = [float(v) for v in g["params"][:5]]
    pos = [_soa(torch, g["pos"][t, a]) for a in range(A)]
    goal = [torch.from_numpy(g["goal"][t, a]).cuda() for a in range(A)]
    shaping = torch.from_numpy(g["shaping_in"][t]).cuda()
    fout = torch.full((5 * A + 2, n), float("nan"), device="cuda")
    bout = torch.full((A + 2, n), 7, dtype=torch.uint8, device="cuda")
    flags = torch.full((4,), -1, dtype=torch.int32, device="cuda")
    cn = H.CNavigation()
    cn.n_agents, cn.shared_rew = A, int(shared)
    cn.pos_shaping_factor, cn.final_reward, cn.agent_collision_penalty, cn.min_collision_distance = factor, final, pen, mcd
    for a in range(A):
        ag = cn.agent[a]
        H._vec(ag.pos, pos[a]); H._vec(ag.goal_pos, goal[a])
        ag.radius, ag.goal_radius = float(g["radius"][a]), float(g["goal_radius"][a])
        ag.pos_shaping = shaping[a].data_ptr()
        for j, fld in enumerate(("pos_shaping_out", "distance_to_goal", "pos_rew", "agent_collision_rew", "reward")):
            setattr(ag, fld, fout[5 * a + j].data_ptr())
        ag.on_goal = bout[a].data_ptr()
        for b in range(a):
            cn.collidable[a][b] = 1
    cn.pos_rew, cn.final_rew = fout[5 * A].data_ptr(), fout[5 * A + 1].data_ptr()
    cn.all_goal_reached, cn.done = bout[A].data_ptr(), bout[A + 1].data_ptr()
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        cn.pair_flags = None  # ignored: the scratch is the library's static device array
        _lib.check(L.vmasb200_navigation_reward_static(C.byref(cn), n, stream), "vmasb200_navigation_reward_static")
    else:
        cn.pair_flags = flags.data_ptr()
        _lib.check(L.vmasb200_navigation_reward(C.byref(cn), n, stream), "vmasb200_navigation_reward")
    torch.cuda.synchronize()
    return fout.cpu().numpy(), bout.cpu().numpy()


def test_static_scratch_equals_existing_entry_point_twice_in_a_row():
    import torch

    from vectorizedmultiagentsimulator_b200 import _lib, hooks as H

    g = dict(np.load(os.path.join(HERE, "golden", "hooks", "navigation.npz")))
    T, A = g["pos"].shape[:2]
    assert T >= 5
    L = _lib.lib()
    any_collision_penalty = False
    for t in (0, 3, 4, 0, T - 1):  # in a row on the shared array: steps with and without pair flags set, one repeated
        f_ref, b_ref = _call(torch, L, H, g, t, static=False)
        f_new, b_new = _call(torch, L, H, g, t, static=True)
        assert not np.isnan(f_new).any() and (b_new <= 1).all(), f"step {t}: an output was not written"
        assert np.array_equal(f_new.view(np.uint32), f_ref.view(np.uint32)), f"step {t}: float outputs differ"
        assert np.array_equal(b_new, b_ref), f"step {t}: flag outputs differ"
        # and both are the reference's own outputs (tests/golden/hooks/navigation.npz)
        assert np.array_equal(f_new[4:5 * A:5], g["reward"][t]) and np.array_equal(b_new[A + 1].astype(bool), g["done"][t])
        any_collision_penalty = any_collision_penalty or bool((g["coll_rew"][t] != 0).any())
    assert any_collision_penalty, "the fixture no longer exercises the flag scratch"


def test_static_scratch_rejects_more_agents_than_the_array_holds():
    import torch  # noqa: F401  (CUDA context of the process)

    from vectorizedmultiagentsimulator_b200 import _lib, hooks as H

    L = _lib.lib()
    cn = H.CNavigation()
    cn.n_agents = 17
    rc = L.vmasb200_navigation_reward_static(C.byref(cn), 8, None)
    assert rc == -1
    msg = L.vmasb200_last_error().decode()
    assert "17" in msg and "static scratch" in msg
