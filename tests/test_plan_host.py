"""CPU: the C-ABI library loads, exports every symbol of include/vmasb200.h, and its host-side plan builder
(pair enumeration / bucketing of core.py:2103-2173, record layout) agrees with the oracle's own enumeration."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tests.fixtures import Fixture, names
from vectorizedmultiagentsimulator_b200 import _lib
from vectorizedmultiagentsimulator_b200.plan import StepPlan
from vectorizedmultiagentsimulator_b200.worlddesc import BOX, LINE, SPHERE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    _lib.build()
    L = _lib.lib()
    hdr = open(os.path.join(ROOT, "include", "vmasb200.h")).read()
    declared = set(re.findall(r"\b(vmasb200_[a-z_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.vmasb200_version() == 3


def _expected_pairs(desc):
    E = len(desc.entities)
    out = []
    order = {(SPHERE, SPHERE): 0, (LINE, SPHERE): 1, (LINE, LINE): 2, (BOX, SPHERE): 3, (BOX, LINE): 4, (BOX, BOX): 5}
    for t in range(6):
        for a in range(E):
            for b in range(a + 1, E):
                if not desc.collide_mask[a][b]:
                    continue
                sa, sb = desc.entities[a].shape, desc.entities[b].shape
                if (sa, sb) in order:
                    ty, x, y = order[(sa, sb)], a, b
                else:
                    ty, x, y = order[(sb, sa)], b, a
                if ty == t:
                    out.append((ty, x, y))
    return out


@pytest.mark.parametrize("name", names())
def test_plan_layout(name):
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    assert pl.pairs == _expected_pairs(fx.desc)
    ents = fx.desc.entities
    n_dyn = sum(e.dynamic for e in ents)
    assert pl.n_dyn == n_dyn and pl.dyn_rows == 6 * n_dyn and pl.envs_per_cta % 32 == 0
    assert all((pl.dyn_slot[i] >= 0) == e.dynamic for i, e in enumerate(ents))
    assert all(not (pl.dyn_slot[i] >= 0 and pl.stat_slot[i] >= 0) for i in range(len(ents)))
    # pack / unpack round trip
    dyn, stat = pl.pack(fx.state0, fx.jrot0)
    assert dyn.shape[1] % 4 == 0 and dyn.shape[1] >= fx.n_envs
    back = pl.unpack(dyn, np.zeros_like(fx.state0))
    for i, e in enumerate(ents):
        if e.dynamic:
            assert (back[i] == fx.state0[i]).all()
    assert pl.alg_bytes_per_env > 0


def test_bad_desc_is_rejected():
    fx = Fixture("simple_spread")
    d, keep = fx.desc.to_ctypes()
    d.abi_version = 99
    h = C.c_void_p()
    rc = _lib.lib().vmasb200_plan_create(C.byref(d), C.byref(h))
    assert rc == -1 and b"abi" in _lib.lib().vmasb200_last_error()


def test_no_gpu_fails_loudly():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    fx = Fixture("simple_spread")
    pl = StepPlan(fx.desc)
    dyn, stat = pl.pack(fx.state0, fx.jrot0)
    with pytest.raises(_lib.VmasB200Error):
        pl.step_host(dyn, stat, np.ascontiguousarray(fx.forces[0].transpose(1, 0, 2)), None)
