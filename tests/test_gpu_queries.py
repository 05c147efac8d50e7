"""GPU: the CUDA distance queries (C ABI: vmasb200_get_distance / vmasb200_is_overlapping /
vmasb200_get_distance_from_point) against the CPU oracle and the golden outputs of the reference's own
World.get_distance / is_overlapping / get_distance_from_point (tests/golden/queries), on the recorded states.

Bar: distances fp32 rtol 1e-5 / atol 1e-6 against the golden vectors and BIT-EXACT against the oracle (same operation
order, both built without FMA contraction); overlap flags bit-exact against both."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.fixtures import QueryFixture, close, max_err, names

pytestmark = pytest.mark.gpu
ALL = names()


def _pose(q, st_dev, e, torch):
    from vectorizedmultiagentsimulator_b200 import queries

    r, l, w = (float(v) for v in q.dims[e])
    return queries.make_pose(q.shapes[e], r, l, w, st_dev[e, :, 0:2], st_dev[e, :, 4:5], torch)


@pytest.mark.parametrize("name", ALL)
def test_pair_queries(name):
    import torch

    from vectorizedmultiagentsimulator_b200 import queries

    q = QueryFixture(name)
    dev = torch.device("cuda", 0)
    n = q.fx.n_envs
    for k, st in enumerate(q.states):
        st_dev = torch.from_numpy(np.ascontiguousarray(st)).to(dev)
        for m, (a, b) in enumerate(q.pairs):
            pa, ka = _pose(q, st_dev, a, torch)
            pb, kb = _pose(q, st_dev, b, torch)
            got = queries.get_distance(pa, pb, n, st_dev, torch).cpu().numpy()
            ov = queries.is_overlapping(pa, pb, n, st_dev, torch).cpu().numpy()
            assert close(got, q.dist[k, m]), f"{name} state {k} pair {(a, b)}: {max_err(got, q.dist[k, m]):.3e}"
            assert np.array_equal(ov, q.overlap[k, m]), f"{name} state {k} pair {(a, b)}: overlap differs from the reference"
            ref = orc.get_distance(q.shapes[a], q.dims[a], st[a], q.shapes[b], q.dims[b], st[b])
            assert close(got, ref, rtol=1e-6, atol=1e-7), f"{name} state {k} pair {(a, b)} vs oracle: {max_err(got, ref):.3e}"
            assert np.array_equal(ov, orc.is_overlapping(q.shapes[a], q.dims[a], st[a], q.shapes[b], q.dims[b], st[b]))


@pytest.mark.parametrize("name", ALL)
def test_point_queries(name):
    import torch

    from vectorizedmultiagentsimulator_b200 import queries

    q = QueryFixture(name)
    dev = torch.device("cuda", 0)
    n = q.fx.n_envs
    for k, st in enumerate(q.states):
        st_dev = torch.from_numpy(np.ascontiguousarray(st)).to(dev)
        for p, e in enumerate(q.pt_entities):
            pose, keep = _pose(q, st_dev, e, torch)
            pts = torch.from_numpy(q.points[k, p]).to(dev)
            got = queries.get_distance_from_point(pose, pts, n, torch).cpu().numpy()
            assert close(got, q.pt_dist[k, p]), f"{name} state {k} entity {e}: {max_err(got, q.pt_dist[k, p]):.3e}"


def test_strided_views_and_large_batch():
    """Poses read through SoA-style strided views (pos_stride 1, xy stride ld) give the same bits as contiguous ones, at
    a batch of many CTAs; is_overlapping == (get_distance < 0) where the reference defines it so (core.py:1934)."""
    import torch

    from vectorizedmultiagentsimulator_b200 import queries
    from vectorizedmultiagentsimulator_b200.worlddesc import BOX, LINE, SPHERE

    dev = torch.device("cuda", 0)
    n = 100_003
    g = torch.Generator(device="cpu").manual_seed(5)
    pos = [(torch.rand(n, 2, generator=g) * 2 - 1).to(dev) for _ in range(2)]
    rot = [(torch.rand(n, 1, generator=g) * 6.28).to(dev) for _ in range(2)]
    soa = torch.empty(6, n + 5, device=dev)
    soa[0:2, :n], soa[2, :n], soa[3:5, :n], soa[5, :n] = pos[0].t(), rot[0][:, 0], pos[1].t(), rot[1][:, 0]
    shapes = {SPHERE: (0.1, 0.0, 0.0), BOX: (0.0, 0.4, 0.2), LINE: (0.0, 0.5, 0.0)}
    for sa, da in shapes.items():
        for sb, db in shapes.items():
            a, _ = queries.make_pose(sa, *da, pos[0], rot[0], torch)
            b, _ = queries.make_pose(sb, *db, pos[1], rot[1], torch)
            a2, k1 = queries.make_pose(sa, *da, soa[0:2, :n].t(), soa[2:3, :n].t(), torch)
            b2, k2 = queries.make_pose(sb, *db, soa[3:5, :n].t(), soa[5:6, :n].t(), torch)
            d1, d2 = queries.get_distance(a, b, n, pos[0], torch), queries.get_distance(a2, b2, n, pos[0], torch)
            assert torch.equal(d1.view(torch.int32), d2.view(torch.int32))
            ov = queries.is_overlapping(a2, b2, n, pos[0], torch)
            if {sa, sb} != {BOX, SPHERE}:
                assert torch.equal(ov, d1 < 0)
            else:
                assert torch.equal(ov, d1 == -1)  # core.py:1842-1843
