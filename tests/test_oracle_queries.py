"""CPU: pins the oracle's distance queries (oracle/vmas_oracle.c) against outputs of the reference's own
World.get_distance / is_overlapping / get_distance_from_point (tests/golden/queries)."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.fixtures import QueryFixture, close, max_err, names

ALL = names()


@pytest.mark.parametrize("name", ALL)
def test_pair_queries(name):
    q = QueryFixture(name)
    for k, st in enumerate(q.states):
        for m, (a, b) in enumerate(q.pairs):
            got = orc.get_distance(q.shapes[a], q.dims[a], st[a], q.shapes[b], q.dims[b], st[b])
            assert close(got, q.dist[k, m]), f"{name} state {k} pair {(a, b)}: {max_err(got, q.dist[k, m]):.3e}"
            ov = orc.is_overlapping(q.shapes[a], q.dims[a], st[a], q.shapes[b], q.dims[b], st[b])
            assert np.array_equal(ov, q.overlap[k, m]), f"{name} state {k} pair {(a, b)}: overlap differs"


@pytest.mark.parametrize("name", ALL)
def test_point_queries(name):
    q = QueryFixture(name)
    for k, st in enumerate(q.states):
        for p, e in enumerate(q.pt_entities):
            got = orc.distance_from_point(q.shapes[e], q.dims[e], st[e], q.points[k, p])
            assert close(got, q.pt_dist[k, p]), f"{name} state {k} entity {e}: {max_err(got, q.pt_dist[k, p]):.3e}"
