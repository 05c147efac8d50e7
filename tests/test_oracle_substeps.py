"""CPU: the C oracle against the reference's World.step INSIDE its substep loop (tests/golden/substeps/*.npz, made by
tests/golden/make_golden_substeps.py from the unmodified reference).

Every recorded substep starts from the reference's own substep input, so nothing is amplified between comparison points.
With the reference's cos / sin / exp values handed in (torch's unary CPU kernels are MKL VML's, closed source; stored in the
fixture) and ATen's vector / scalar-tail split of logaddexp mirrored (`vec_width`), the oracle must reproduce the
reference BIT for bit: state, forces_dict / torques_dict, and - directly - the force and torque `_vectorized_joint_constraints`
hands to `update_env_forces` for every joint.  With libm's own cos / sin instead, the stated fp32 bar
(rtol 1e-5 / atol 1e-6) holds for every world except the stiff jointed ones, whose deviation is then shown to be the 1-ulp
cos / sin difference and nothing else."""
import glob
import json
import os

import numpy as np
import pytest

from oracle.oracle import Oracle
from tests.fixtures import GOLDEN, Fixture, close, max_err

NAMES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "substeps", "*.npz")))
# stiff worlds: joint_force >= 130 on moments of inertia of ~5e-4 (joint_passage.py:131-171, asym_joint, waterfall)
STIFF = {"joint_passage", "asym_joint", "waterfall"}


def load(name):
    z = dict(np.load(os.path.join(GOLDEN, "substeps", name + ".npz")))
    z["meta"] = json.loads(str(z["meta_json"]))
    return z


def cases(fx, z):
    for k, t in enumerate(z["steps"]):
        for s in range(fx.desc.substeps):
            yield k, int(t), s


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("name", NAMES)
def test_substep_bit_exact_with_reference_math(name):
    fx, z = Fixture(name), load(name)
    orc = Oracle(fx.desc)
    W = z["meta"]["vec_width"]
    assert W in (16, 32)
    for k, t, s in cases(fx, z):
        st = np.ascontiguousarray(z["states"][k, s].copy())
        f, tq = np.ascontiguousarray(z["act_force_in"][k, s].copy()), np.ascontiguousarray(z["act_torque_in"][k, s].copy())
        m, tf, tt, jf, jt = orc.substep(st, f, tq, fx.jrot_at(t), s, trig=np.ascontiguousarray(z["trig"][k, s]),
                                        jexp=np.ascontiguousarray(z["jexp"][k, s]), vec_width=W)
        where = f"{name} step {t} substep {s}"
        assert (m == fx.masks[t][s]).all(), where + ": World.collides verdicts"
        assert np.array_equal(bits(f), bits(z["act_force_out"][k, s])) and np.array_equal(bits(tq), bits(z["act_torque_out"][k, s]))
        # a8 / a9 directly: per joint, force and torque on entity_a and entity_b (core.py:2282-2291)
        assert np.array_equal(bits(jf), bits(z["joint_force"][k, s])), where + f": joint force {max_err(jf, z['joint_force'][k, s]):.3e}"
        assert np.array_equal(bits(jt), bits(z["joint_torque"][k, s])), where + f": joint torque {max_err(jt, z['joint_torque'][k, s]):.3e}"
        # forces_dict / torques_dict as _integrate_state reads them; -0.0 == +0.0 (an accumulator nothing was added to)
        assert np.array_equal(tf, z["total_force"][k, s]) and np.array_equal(tt, z["total_torque"][k, s]), where + ": forces_dict"
        assert np.array_equal(bits(st), bits(z["states"][k, s + 1])), where + f": state {max_err(st, z['states'][k, s + 1]):.3e}"


@pytest.mark.parametrize("name", NAMES)
def test_substep_with_libm_math(name):
    """The oracle's default math (libm cosf / sinf / expf): plain fp32 bar per substep; for the stiff jointed worlds the
    miss is bounded and is exactly what a 1-ulp change of the cos / sin values produces (the comparison above removes it)."""
    fx, z = Fixture(name), load(name)
    orc = Oracle(fx.desc)
    worst = 0.0
    for k, t, s in cases(fx, z):
        st = np.ascontiguousarray(z["states"][k, s].copy())
        f, tq = np.ascontiguousarray(z["act_force_in"][k, s].copy()), np.ascontiguousarray(z["act_torque_in"][k, s].copy())
        m, tf, tt, jf, jt = orc.substep(st, f, tq, fx.jrot_at(t), s)
        assert (m == fx.masks[t][s]).all()
        exp = z["states"][k, s + 1]
        if name not in STIFF:
            assert close(st, exp), f"{name} step {t} substep {s}: {max_err(st, exp):.3e}"
            assert close(tf, z["total_force"][k, s]) and close(tt, z["total_torque"][k, s])
        worst = max(worst, max_err(st, exp))
    if name in STIFF:
        # measured: 2.4e-5 (asym_joint), 2.1e-5 (joint_passage), 7.2e-6 (waterfall) per substep
        assert worst <= 1e-4, f"{name}: per-substep deviation {worst:.3e} with libm cos / sin"


def test_libm_trig_is_within_one_ulp_of_the_reference():
    """Names the op: the reference's cos / sin (MKL VML through torch.cos / torch.sin) vs libm's correctly rounded ones on
    the very angles of the stiff fixtures: never more than 1 ulp apart, and different on a few percent of them."""
    n_diff = n_all = 0
    for name in sorted(STIFF):
        z = load(name)
        rot = z["states"][:, :-1, :, :, 4]
        for col, fn in ((0, np.cos), (1, np.sin)):
            ref = z["trig"][..., col]
            mine = fn(rot.astype(np.float64)).astype(np.float32)  # correctly rounded (double evaluation, one rounding)
            d = np.abs(bits(ref).astype(np.int64) - bits(mine).astype(np.int64))
            sign_flip = np.signbit(ref) != np.signbit(mine)
            assert (d[~sign_flip] <= 1).all(), f"{name}: {d.max()} ulp"
            n_diff += int((d > 0).sum()); n_all += d.size
    assert 0 < n_diff < 0.2 * n_all
