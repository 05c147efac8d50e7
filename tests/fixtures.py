"""Loader / replay helpers for tests/golden/*.npz (written by tests/golden/make_golden.py)."""
import glob
import json
import os

import numpy as np

from vectorizedmultiagentsimulator_b200.worlddesc import WorldDesc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# parity bar of BASELINE.json north_star: fp32 rtol 1e-5 / atol 1e-6 after one step
RTOL, ATOL = 1e-5, 1e-6


def names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


class Fixture:
    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.name = name
        self.desc = WorldDesc.from_json(str(z["desc_json"]))
        self.meta = json.loads(str(z["meta_json"]))
        self.state0 = z["state0"]
        self.jrot0 = z["jrot0"]
        self.forces = z["forces"]
        self.torques = z["torques"]
        self.snaps = [int(s) for s in z["snaps"]]
        self.state_after = z["state_after"]
        self.force_after = z["force_after"]
        self.torque_after = z["torque_after"]
        self.done = z["done"]
        shp = tuple(int(v) for v in z["masks_shape"])
        self.masks = np.unpackbits(z["masks_packed"])[: int(np.prod(shp))].reshape(shp).astype(bool)
        self.override_steps = z["override_steps"]
        self.override_entities = z["override_entities"]
        self.override_states = z["override_states"]
        self.jrot_override_steps = z["jrot_override_steps"]
        self.jrot_override_values = z["jrot_override_values"]
        self.lidar = [z[f"lidar_{k}"] for k in range(len(self.desc.lidars))]
        self.n_envs = self.state0.shape[1]
        self.steps = self.forces.shape[0]

    def snap_index(self, step):
        return self.snaps.index(step)

    def one_step_cases(self):
        """(input state, step index t, expected state index k): every t whose input state is stored."""
        cases = [(self.state0, 0, self.snap_index(1))]
        for s in self.snaps:
            if s + 1 in self.snaps and not (self.override_steps == s).any() and not (self.jrot_override_steps == s).any():
                cases.append((self.state_after[self.snap_index(s)], s, self.snap_index(s + 1)))
        return cases

    def jrot_at(self, t):
        j = self.jrot0
        for s, v in zip(self.jrot_override_steps, self.jrot_override_values):
            if s <= t:
                j = v
        return j

    def apply_overrides(self, t, state):
        for s, e, v in zip(self.override_steps, self.override_entities, self.override_states):
            if s == t:
                state[e] = v


class QueryFixture:
    """tests/golden/queries/<name>.npz (tests/golden/make_golden_queries.py): reference outputs of World.get_distance /
    is_overlapping / get_distance_from_point on recorded states of fixture `name`."""

    def __init__(self, name):
        self.fx = Fixture(name)
        z = np.load(os.path.join(GOLDEN, "queries", name + ".npz"))
        self.snaps = [int(k) for k in z["snaps"]]
        self.states = [self.fx.state0 if k < 0 else self.fx.state_after[k] for k in self.snaps]
        self.pairs = [(int(a), int(b)) for a, b in z["pairs"]]
        self.dist = z["dist"]
        shp = tuple(int(v) for v in z["overlap_shape"])
        self.overlap = np.unpackbits(z["overlap"])[: int(np.prod(shp))].reshape(shp).astype(bool)
        self.pt_entities = [int(e) for e in z["pt_entities"]]
        self.points = z["points"]
        self.pt_dist = z["pt_dist"]
        self.dims = z["dims"]          # [E, 3] radius, length, width as Python floats
        self.shapes = [int(s) for s in z["shapes"]]


def close(a, b, rtol=RTOL, atol=ATOL):
    return np.allclose(a, b, rtol=rtol, atol=atol, equal_nan=True)


def max_err(a, b):
    return float(np.nanmax(np.abs(a.astype(np.float64) - b.astype(np.float64)))) if a.size else 0.0


# 2 x Vectorized<float>::size() of the ATen build that generated tests/golden/*.npz (AVX-512 container: 2 x 16): ATen's
# logaddexp runs whole groups of that many elements through SLEEF and the tail through libm (oracle/vmas_oracle.h)
ATEN_VEC_WIDTH = 32


def reference_math_available():
    """The oracle's bit-exact mode evaluates torch.cos / torch.sin / torch.exp (MKL VML) on the test machine.  True when
    this machine's torch returns the values the fixture machine's did (checked on the stored angles of a stiff fixture)."""
    import torch

    z = np.load(os.path.join(GOLDEN, "substeps", "asym_joint.npz"))
    rot = torch.from_numpy(np.ascontiguousarray(z["states"][:, :-1, :, :, 4]))
    return bool(np.array_equal(rot.cos().numpy(), z["trig"][..., 0]) and np.array_equal(rot.sin().numpy(), z["trig"][..., 1]))


def trig_sensitivity(orc, state_in, force, torque, jrot, trials=6, seed=0):
    """Per ELEMENT of the output state: how far one World.step moves when every cos / sin value (and the exp of a
    rotation-holding joint's torque, core.py:2852) of every substep is off by -1 / 0 / +1 ulp at random (what libm or CUDA do to the reference's MKL values): max over `trials` draws of
    |jittered - exact|.  Zero for entities no stiff joint reaches; ~1e-3 in ang_vel of a line held by joint_force 900."""
    base = np.ascontiguousarray(state_in.copy())
    orc.step_torch_math(base, np.ascontiguousarray(force.copy()), np.ascontiguousarray(torque.copy()), jrot, vec_width=ATEN_VEC_WIDTH)
    rng = np.random.default_rng(seed)
    sens = np.zeros(base.shape, np.float64)
    for _ in range(trials):
        st = np.ascontiguousarray(state_in.copy())
        orc.step_torch_math(st, np.ascontiguousarray(force.copy()), np.ascontiguousarray(torque.copy()), jrot,
                            vec_width=ATEN_VEC_WIDTH, trig_jitter=rng)
        sens = np.maximum(sens, np.abs(st.astype(np.float64) - base.astype(np.float64)))
    return sens


def conditioned_close(got, exp, sens, factor=8.0):
    """|got - exp| <= atol + rtol |exp| + factor * sens, ELEMENT by element (`sens` has the state's shape, from
    trig_sensitivity): only the elements a stiff joint really drives get slack, everything else keeps the plain fp32 bar."""
    err = np.abs(got.astype(np.float64) - exp.astype(np.float64))
    tol = ATOL + RTOL * np.abs(exp.astype(np.float64)) + factor * np.asarray(sens, np.float64)
    return bool((err <= tol).all())


def drift_bounds(desc):
    """Stated 100-step drift bound (pos, vel) for envs whose collision masks never diverged.
    Worlds without joints reproduce the reference to a few ulp (measured: <= 1e-6 / 3e-6 over 100 steps);
    stiff jointed worlds (joint_force >= 130, substeps >= 5) jitter chaotically at the 1e-2 velocity level, so
    the bound there only says the trajectories stay together (measured: 2e-3 / 6e-2)."""
    return (1e-2, 2e-1) if desc.joints else (1e-4, 1e-3)
