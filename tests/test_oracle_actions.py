"""CPU: the numpy restatement of Environment._set_action for discrete actions (oracle/hooks_oracle.py) against outputs
of the unmodified reference (tests/golden/actions/discrete.npz, tests/golden/make_golden_actions.py), bit for bit."""
import os

import numpy as np
import pytest

from oracle import hooks_oracle as ho

GOLD = os.path.join(os.path.dirname(__file__), "golden", "actions", "discrete.npz")


def cases():
    z = np.load(GOLD)
    for i in range(int(z["n_cases"])):
        yield {k: z[f"case_{i}_{k}"] for k in ("action", "nvec", "u_range", "u_mult", "multi", "u")}


@pytest.mark.parametrize("c", list(cases()), ids=lambda c: f"multi{int(c['multi'])}_nvec{'x'.join(map(str, c['nvec']))}")
def test_oracle_reproduces_the_reference(c):
    u, bad = ho.set_action_discrete(c["action"], [int(v) for v in c["nvec"]], c["u_range"], c["u_mult"], bool(c["multi"]))
    assert not bad
    assert np.array_equal(u.view(np.uint32), c["u"].view(np.uint32))


def test_out_of_bounds_is_reported():
    u, bad = ho.set_action_discrete(np.array([[9]]), [3, 3], [1.0, 1.0], [1.0, 1.0], False)
    assert bad
    u, bad = ho.set_action_discrete(np.array([[-1]]), [3, 3], [1.0, 1.0], [1.0, 1.0], False)
    assert bad
    u, bad = ho.set_action_discrete(np.array([[1, 3]]), [3, 3], [1.0, 1.0], [1.0, 1.0], True)
    assert bad
