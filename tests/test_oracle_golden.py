"""CPU: pins the C oracle (oracle/vmas_oracle.c) against outputs of the reference itself (tests/golden).

Two modes of the oracle:
* reference math (`step_torch_math`): cos / sin / exp taken from torch's own CPU kernels (MKL VML: third party, closed
  source), exp / log1p of logaddexp from the SLEEF restatement with ATen's vector / tail split.  Then the oracle IS the
  reference, bit for bit: every one-step case of all 18 fixtures and whole 100-step rollouts, stiff jointed worlds included.
* libm math (`step`, what the timed cpu_baseline runs): plain fp32 bar (rtol 1e-5 / atol 1e-6) on every world but the three
  stiff jointed ones, where the bar is widened ELEMENT by element by what a 1-ulp jitter of the cos / sin values does."""
import numpy as np
import pytest

from oracle.oracle import Oracle
from tests.fixtures import (ATEN_VEC_WIDTH, Fixture, close, conditioned_close, drift_bounds, max_err, names,
                            reference_math_available, trig_sensitivity)

ALL = names()
STIFF = {"joint_passage", "asym_joint", "waterfall"}


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def refmath():
    if not reference_math_available():
        pytest.skip("torch.cos / torch.sin on this CPU do not return the fixture machine's values (other MKL code path)")


@pytest.mark.parametrize("name", ALL)
def test_one_step_bit_exact_with_reference_math(name, refmath):
    fx = Fixture(name)
    orc = Oracle(fx.desc)
    for state_in, t, k in fx.one_step_cases():
        st = np.ascontiguousarray(state_in.copy())
        f, tq = np.ascontiguousarray(fx.forces[t].copy()), np.ascontiguousarray(fx.torques[t].copy())
        masks = orc.step_torch_math(st, f, tq, fx.jrot_at(t), vec_width=ATEN_VEC_WIDTH)
        assert np.array_equal(bits(st), bits(fx.state_after[k])), f"{name} step {t}: {max_err(st, fx.state_after[k]):.3e}"
        assert (masks == fx.masks[t]).all(), f"{name} step {t}: mask mismatch"
        assert np.array_equal(bits(f), bits(fx.force_after[k])) and np.array_equal(bits(tq), bits(fx.torque_after[k]))


@pytest.mark.parametrize("name", ALL)
def test_rollout_bit_exact_with_reference_math(name, refmath):
    """Free-running replay of the recorded forces over the whole recorded rollout: state at every snapshot and the
    World.collides verdicts of every substep of every step equal the reference's, bit for bit."""
    fx = Fixture(name)
    orc = Oracle(fx.desc)
    st = np.ascontiguousarray(fx.state0.copy())
    for t in range(fx.steps):
        fx.apply_overrides(t, st)
        f, tq = np.ascontiguousarray(fx.forces[t].copy()), np.ascontiguousarray(fx.torques[t].copy())
        masks = orc.step_torch_math(st, f, tq, fx.jrot_at(t), vec_width=ATEN_VEC_WIDTH)
        assert (masks == fx.masks[t]).all(), f"{name} step {t}: mask mismatch"
        if (t + 1) in fx.snaps:
            exp = fx.state_after[fx.snap_index(t + 1)]
            assert np.array_equal(bits(st), bits(exp)), f"{name} after step {t + 1}: {max_err(st, exp):.3e}"


@pytest.mark.parametrize("name", ALL)
def test_one_step_libm_math(name):
    fx = Fixture(name)
    orc = Oracle(fx.desc)
    for state_in, t, k in fx.one_step_cases():
        st = np.ascontiguousarray(state_in.copy())
        f = np.ascontiguousarray(fx.forces[t].copy())
        tq = np.ascontiguousarray(fx.torques[t].copy())
        masks = orc.step(st, f, tq, fx.jrot_at(t))
        exp = fx.state_after[k]
        if name in STIFF:
            sens = trig_sensitivity(orc, state_in, fx.forces[t], fx.torques[t], fx.jrot_at(t))
            assert conditioned_close(st, exp, sens), f"{name} step {t}: max err {max_err(st, exp):.3e}"
        else:
            assert close(st, exp), f"{name} step {t}: max err {max_err(st, exp):.3e}"
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
def test_rollout_drift_libm_math(name):
    """Free-running replay with the oracle's own libm math; the stated drift bound is checked at every
    snapshot for the envs whose collision masks have not diverged from the reference's, and worlds without stiff
    joints must never diverge."""
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
    assert not diverged or name in STIFF, f"{name}: collision masks diverged from the reference's"
