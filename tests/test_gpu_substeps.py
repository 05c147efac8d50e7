"""GPU: ONE substep of the CUDA step (vmasb200_step's probe fields, exact path) against the reference's World.step
inside its substep loop (tests/golden/substeps/*.npz) - every substep starts from the reference's own substep input.

1. With the reference's cos / sin values (`trig_in`) and, for joints that hold a rotation, its exp value (`jexp_in`)
   handed in (the reference's are MKL VML's, which no other library reproduces bit for bit): state, forces_dict / torques_dict and the per-joint constraint force and torque (SURVEY 8
   a7 / a8 / a9) at the PLAIN fp32 bar, rtol 1e-5 / atol 1e-6, no conditioning - stiff jointed worlds included - and the
   World.collides verdicts and clamped actions bit-exact.
2. With the kernel's own sincosf: the same plain bar on every world without stiff joints; on the three stiff ones the
   miss of a substep is bounded and - by 1. - is the 1-ulp cos / sin difference and nothing else."""
import glob
import json
import os

import numpy as np
import pytest

from tests.fixtures import GOLDEN, Fixture, close, max_err
from vectorizedmultiagentsimulator_b200.plan import StepPlan

pytestmark = pytest.mark.gpu
NAMES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "substeps", "*.npz")))
STIFF = {"joint_passage", "asym_joint", "waterfall"}


def load(name):
    z = dict(np.load(os.path.join(GOLDEN, "substeps", name + ".npz")))
    z["meta"] = json.loads(str(z["meta_json"]))
    return z


def _joint_check(fx, z, k, s, out, where):
    """core.py:2282-2291: (force_a, torque_a, force_b, torque_b) of every joint; the kernel keeps fx fy ta tb (fb = -fa)."""
    for j, jt in enumerate(fx.desc.joints):
        got = out["joint"][j]
        assert close(got[:, 0:2], z["joint_force"][k, s, j, 0]), where + f": joint {j} force on a {max_err(got[:, 0:2], z['joint_force'][k, s, j, 0]):.3e}"
        assert close(-got[:, 0:2], z["joint_force"][k, s, j, 1]), where + f": joint {j} force on b"
        assert close(got[:, 2], z["joint_torque"][k, s, j, 0]), where + f": joint {j} torque on a {max_err(got[:, 2], z['joint_torque'][k, s, j, 0]):.3e}"
        assert close(got[:, 3], z["joint_torque"][k, s, j, 1]), where + f": joint {j} torque on b {max_err(got[:, 3], z['joint_torque'][k, s, j, 1]):.3e}"


@pytest.mark.parametrize("name", NAMES)
def test_substep_with_reference_trig(name):
    import torch

    fx, z = Fixture(name), load(name)
    pl = StepPlan(fx.desc)
    n_bit = n_all = 0
    for k, t in enumerate(z["steps"]):
        t = int(t)
        for s in range(fx.desc.substeps):
            where = f"{name} step {t} substep {s}"
            out = pl.substep_probe(torch, z["states"][k, s], z["act_force_in"][k, s], z["act_torque_in"][k, s], fx.jrot_at(t), s,
                                   trig=z["trig"][k, s], jexp=z["jexp"][k, s])
            exp = z["states"][k, s + 1]
            assert close(out["state"], exp), where + f": state {max_err(out['state'], exp):.3e}"
            assert (out["masks"] == fx.masks[t][s]).all(), where + ": World.collides verdicts"
            assert np.array_equal(out["force"], z["act_force_out"][k, s]) and np.array_equal(out["torque"], z["act_torque_out"][k, s])
            de = out["dyn_entities"]
            mov = [e for e in de if fx.desc.entities[e].movable]
            rot = [e for e in de if fx.desc.entities[e].rotatable]
            assert close(out["total_force"][mov], z["total_force"][k, s][mov]), where + ": forces_dict"
            assert close(out["total_torque"][rot], z["total_torque"][k, s][rot]), where + ": torques_dict"
            _joint_check(fx, z, k, s, out, where)
            n_all += 1
            n_bit += int(np.array_equal(out["state"].view(np.uint32), exp.view(np.uint32)))
    print(f"{name}: {n_bit}/{n_all} substeps bit-identical to the reference (CUDA expf / log1pf vs SLEEF / libm account for the rest)")


@pytest.mark.parametrize("name", NAMES)
def test_substep_with_own_sincosf(name):
    import torch

    fx, z = Fixture(name), load(name)
    pl = StepPlan(fx.desc)
    worst = 0.0
    for k, t in enumerate(z["steps"]):
        t = int(t)
        for s in range(fx.desc.substeps):
            out = pl.substep_probe(torch, z["states"][k, s], z["act_force_in"][k, s], z["act_torque_in"][k, s], fx.jrot_at(t), s)
            exp = z["states"][k, s + 1]
            assert (out["masks"] == fx.masks[t][s]).all()
            if name not in STIFF:
                assert close(out["state"], exp), f"{name} step {t} substep {s}: {max_err(out['state'], exp):.3e}"
            worst = max(worst, max_err(out["state"], exp))
    if name in STIFF:
        # the oracle with libm's cos / sin measures 2.4e-5 / 2.1e-5 / 7.2e-6 per substep (tests/test_oracle_substeps.py)
        assert worst <= 1e-4, f"{name}: per-substep deviation {worst:.3e} with CUDA sincosf"
    print(f"{name}: worst per-substep deviation with CUDA sincosf {worst:.3e}")
