"""GPU: the world-compiled step kernel (csrc/jit.h, NVRTC) against the generic table-walking kernel, BIT for bit, on
every golden world the generator covers - state, clamped actions, LIDAR, collision masks - over the recorded rollout,
and at a batch size that takes the large-batch CTA shape.  The generic kernel is the one pinned to the oracle and the
reference's golden outputs (tests/test_gpu_parity.py, which with exact=False also runs the compiled kernel against
them when the world has one)."""
import numpy as np
import pytest

from tests.fixtures import Fixture, names
from vectorizedmultiagentsimulator_b200 import _lib
from vectorizedmultiagentsimulator_b200.plan import StepPlan

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _every_covered_world():
    """mode 2: also the worlds the library would leave to the generic kernel for speed (football, joint_passage)"""
    L = _lib.lib()
    prev = L.vmasb200_jit_enable(2)
    yield
    L.vmasb200_jit_enable(prev)

COVERED = [n for n in names() if n != "pollock"]  # pollock: 990 pairs, beyond what the generator unrolls


def _same(a, b):
    if a is None or b is None:
        return a is None and b is None
    return np.array_equal(np.asarray(a).view(np.uint32), np.asarray(b).view(np.uint32))


def _forces(fx, t):
    return np.ascontiguousarray(fx.forces[t].transpose(1, 0, 2)), np.ascontiguousarray(fx.torques[t].transpose(1, 0))


@pytest.mark.parametrize("name", COVERED)
def test_compiled_equals_generic_bitwise(name):
    L = _lib.lib()
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    assert pl.jit(fx.n_envs)[0] == 1, L.vmasb200_last_error()
    n, st = fx.n_envs, fx.state0.copy()
    launches = 0
    for t in range(min(fx.steps, 30)):
        fx.apply_overrides(t, st)
        dyn, stat = pl.pack(st, fx.jrot_at(t))
        f, tq = _forces(fx, t)
        a = pl.step_host(dyn, stat, f, tq, exact=False)
        now = pl.jit(n)[1]
        assert now > launches, "the step did not take the world-compiled kernel"
        launches = now
        L.vmasb200_jit_enable(0)
        try:
            b = pl.step_host(dyn, stat, f, tq, exact=False)
        finally:
            L.vmasb200_jit_enable(2)
        assert pl.jit(n)[1] == launches, "the generic step took the compiled kernel"
        rows = pl.dyn_rows
        assert _same(a[0][:rows, :n], b[0][:rows, :n]), f"{name} t={t}: state differs"
        for k, what in ((1, "clamped actions"), (2, "lidar")):
            if a[k] is not None or b[k] is not None:
                assert _same(a[k][:, :n], b[k][:, :n]), f"{name} t={t}: {what} differ"
        assert _same(a[3], b[3]), f"{name} t={t}: collision masks differ"
        st = pl.unpack(a[0], st)


@pytest.mark.parametrize("name", ["navigation", "simple_spread", "flocking", "transport", "football", "joint_passage", "wheel"])
@pytest.mark.parametrize("n", [1000, 70001])
def test_compiled_equals_generic_large_batch(name, n):
    """ragged batch sizes (last tile partly empty, ld padding) and, at 70001 envs, the large-batch CTA shape"""
    L = _lib.lib()
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    rng = np.random.default_rng(7)
    reps = -(-n // fx.n_envs)
    st = np.tile(fx.state0, (1, reps, 1))[:, :n].copy()
    st[..., 0:2] += rng.uniform(-0.05, 0.05, st[..., 0:2].shape).astype(np.float32)   # contacts come and go
    f = np.tile(fx.forces[0], (1, reps, 1))[:, :n]
    tq = np.tile(fx.torques[0], (1, reps))[:, :n]
    f, tq = np.ascontiguousarray(f.transpose(1, 0, 2)), np.ascontiguousarray(tq.transpose(1, 0))
    jr = fx.jrot_at(0)
    dyn, stat = pl.pack(st, np.tile(jr, (1, reps))[:, :n] if jr is not None and jr.size else None)
    for _ in range(3):
        a = pl.step_host(dyn, stat, f, tq, exact=False)
        assert pl.jit(n)[1] > 0
        L.vmasb200_jit_enable(0)
        try:
            b = pl.step_host(dyn, stat, f, tq, exact=False)
        finally:
            L.vmasb200_jit_enable(2)
        assert _same(a[0][:pl.dyn_rows, :n], b[0][:pl.dyn_rows, :n]), "state differs"
        if a[2] is not None:
            assert _same(a[2][:, :n], b[2][:, :n]), "lidar differs"
        assert _same(a[3], b[3]), "collision masks differ"
        dyn = a[0]


@pytest.mark.parametrize("name,rays", [("transport", (12, 12)), ("joint_passage", (24, 30)), ("wheel", (7, 5)), ("football", (20, 20)), ("football", (8, 8))])
def test_compiled_lidar_on_boxes_and_lines(name, rays):
    """the golden worlds only point their LIDARs at spheres: give the agents of worlds with boxes and lines sensors that
    see every other entity (equal ray counts: the unrolled constant-bank sweep; different ones: the table-driven sweep)
    and require the compiled epilogue (pair queue, frame sweep, exact rays with literal dimensions) to equal the
    generic one bit for bit"""
    import copy
    import math
    from vectorizedmultiagentsimulator_b200.worlddesc import LidarDesc
    L = _lib.lib()
    fx = Fixture(name)
    desc = copy.deepcopy(fx.desc)
    E = len(desc.entities)
    desc.lidars = []
    for i, a in enumerate(desc.agent_indices[:4]):
        nr = rays[i % len(rays)]
        desc.lidars.append(LidarDesc(entity=a, angles=[2 * math.pi * r / nr for r in range(nr)], max_range=0.4 + 0.1 * (i % 2),
                                     targets=[e for e in range(E) if e != a][:32]))
    pl = StepPlan(desc)
    n = 5000
    rng = np.random.default_rng(11)
    reps = -(-n // fx.n_envs)
    st = np.tile(fx.state0, (1, reps, 1))[:, :n].copy()
    st[..., 0:2] += rng.uniform(-0.2, 0.2, st[..., 0:2].shape).astype(np.float32)
    st[..., 4] += rng.uniform(-3.0, 3.0, st[..., 4].shape).astype(np.float32)
    f = np.tile(fx.forces[0], (1, reps, 1))[:, :n]
    tq = np.tile(fx.torques[0], (1, reps))[:, :n]
    f, tq = np.ascontiguousarray(f.transpose(1, 0, 2)), np.ascontiguousarray(tq.transpose(1, 0))
    jr = fx.jrot_at(0)
    dyn, stat = pl.pack(st, np.tile(jr, (1, reps))[:, :n] if jr is not None and jr.size else None)
    assert pl.jit(n)[0] == 1, L.vmasb200_last_error()
    for _ in range(2):
        a = pl.step_host(dyn, stat, f, tq, exact=False)
        assert pl.jit(n)[1] > 0
        L.vmasb200_jit_enable(0)
        try:
            b = pl.step_host(dyn, stat, f, tq, exact=False)
        finally:
            L.vmasb200_jit_enable(2)
        assert _same(a[0][:pl.dyn_rows, :n], b[0][:pl.dyn_rows, :n]), "state differs"
        assert a[2] is not None and _same(a[2][:, :n], b[2][:, :n]), "lidar differs"
        assert (a[2][:, :n] < np.float32(0.39)).any(), "no ray hit anything: the test does not exercise the exact rays"
        dyn = a[0]


def test_compiled_launch_reaches_the_repair():
    """joint_passage at a few thousand envs has thin contacts at the ends of the passage's wall segments whose pair no
    exact env activates: the compiled launch must hand over to kernels.cuh repair_kernel, and the result must still be
    the generic kernel's, bit for bit."""
    L = _lib.lib()
    fx = Fixture("joint_passage")
    pl = StepPlan(fx.desc)
    n = 8192
    reps = -(-n // fx.n_envs)
    rng = np.random.default_rng(3)
    st = np.tile(fx.state0, (1, reps, 1))[:, :n].copy()
    dynamic = [e for e in range(st.shape[0]) if pl.dyn_slot[e] >= 0]
    st[dynamic, :, 0:2] += rng.uniform(-0.02, 0.02, st[dynamic, :, 0:2].shape).astype(np.float32)
    jr = fx.jrot_at(0)
    dyn, stat = pl.pack(st, np.tile(jr, (1, reps))[:, :n] if jr is not None and jr.size else None)
    A = len(pl.agents)
    repairs0 = pl.sync_check()
    for t in range(40):
        f = rng.uniform(-1, 1, (n, A, 2)).astype(np.float32)
        tq = np.zeros((n, A), np.float32)
        a = pl.step_host(dyn, stat, f, tq, exact=False)
        L.vmasb200_jit_enable(0)
        try:
            b = pl.step_host(dyn, stat, f, tq, exact=False)
        finally:
            L.vmasb200_jit_enable(2)
        assert _same(a[0][:pl.dyn_rows, :n], b[0][:pl.dyn_rows, :n]), f"t={t}: state differs"
        assert _same(a[3], b[3]), f"t={t}: collision masks differ"
        dyn = a[0]
    assert pl.jit(n)[1] > 0
    assert pl.sync_check() > repairs0, "no step needed a repair: the test does not exercise the hand-over"
