"""CPU: pins the C oracle (oracle/vmas_oracle.c) against outputs of the reference itself (tests/golden)."""
import numpy as np
import pytest

from oracle.oracle import Oracle
from tests.fixtures import Fixture, close, drift_bounds, conditioned_close, max_err, names, ulp_perturb

ALL = names()


@pytest.mark.parametrize("name", ALL)
def test_one_step_state_and_masks(name):
    fx = Fixture(name)
    orc = Oracle(fx.desc)
    for state_in, t, k in fx.one_step_cases():
        st = np.ascontiguousarray(state_in.copy())
        f = np.ascontiguousarray(fx.forces[t].copy())
        tq = np.ascontiguousarray(fx.torques[t].copy())
        masks = orc.step(st, f, tq, fx.jrot_at(t))
        exp = fx.state_after[k]
        # conditioning of this step: output change under a 1-ulp perturbation of the input state
        sp = np.ascontiguousarray(ulp_perturb(state_in))
        orc.step(sp, fx.forces[t].copy(), fx.torques[t].copy(), fx.jrot_at(t), want_masks=False)
        sens = [max_err(sp[..., i], st[..., i]) for i in range(6)]
        assert conditioned_close(st, exp, sens), f"{name} step {t}: max err {max_err(st, exp):.3e} sens {sens}"
        # collision masks (World.collides verdicts per substep) are bit-exact
        assert (masks == fx.masks[t]).all(), f"{name} step {t}: mask mismatch {np.argwhere(masks != fx.masks[t])[:5]}"
        # clamped force / torque write-back
        assert close(f, fx.force_after[k]) and close(tq, fx.torque_after[k])


@pytest.mark.parametrize("name", [n for n in ALL if Fixture(n).lidar])
def test_lidar(name):
    fx = Fixture(name)
    orc = Oracle(fx.desc)
    for k, step in enumerate(fx.snaps):
        st = np.ascontiguousarray(fx.state_after[k])
        for li in range(len(fx.desc.lidars)):
            got = orc.lidar(st, li)
            assert close(got, fx.lidar[li][k]), f"{name} lidar {li} step {step}: {max_err(got, fx.lidar[li][k]):.3e}"


@pytest.mark.parametrize("name", ALL)
def test_rollout_drift(name):
    """Free-running replay of the recorded forces; the stated drift bound is checked at every snapshot
    for envs whose collision masks never diverged from the reference's."""
    fx = Fixture(name)
    orc = Oracle(fx.desc)
    st = np.ascontiguousarray(fx.state0.copy())
    diverged = False
    pos_tol, vel_tol = drift_bounds(fx.desc)
    for t in range(fx.steps):
        fx.apply_overrides(t, st)
        f = np.ascontiguousarray(fx.forces[t].copy())
        tq = np.ascontiguousarray(fx.torques[t].copy())
        masks = orc.step(st, f, tq, fx.jrot_at(t))
        diverged = diverged or bool((masks != fx.masks[t]).any())
        if (t + 1) in fx.snaps and not diverged:
            exp = fx.state_after[fx.snap_index(t + 1)]
            assert max_err(st[..., 0:2], exp[..., 0:2]) <= pos_tol, f"{name} pos drift at step {t+1}"
            assert max_err(st[..., 2:4], exp[..., 2:4]) <= vel_tol, f"{name} vel drift at step {t+1}"
    assert np.isfinite(st).all()
