#!/usr/bin/env python
"""Golden vectors of the distance queries from the UNMODIFIED reference: World.get_distance,
World.is_overlapping, World.get_distance_from_point (core.py:1787-1966).

TEST INFRASTRUCTURE.  Run in the build container after tests/golden/make_golden.py:

    python tests/golden/make_golden_queries.py [fixture ...]

For every step fixture the reference scenario is rebuilt, its entities are put into recorded states (state0, a middle
and the last snapshot) and the three functions are called for entity pairs of every shape combination the world
holds - the pairs whose centres come closest, in both argument orders - and for random test points around the
entities.  Output: tests/golden/queries/<fixture>.npz.
"""
import itertools
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_ref"), os.path.join(ROOT, "oracle", "gym_stub"), ROOT, HERE]

import vmas  # noqa: E402
from make_golden import CONFIGS  # noqa: E402

from tests.fixtures import Fixture  # noqa: E402

PER_COMBO = 5


def run(name):
    cfg = CONFIGS[name]
    fx = Fixture(name)
    n = fx.n_envs
    torch.manual_seed(0)
    env = vmas.make_env(cfg["scenario"], num_envs=n, device="cpu", seed=0, continuous_actions=True, **cfg["kw"])
    world = env.world
    env.reset(seed=0)
    ents = list(world.entities)
    assert [e.name for e in ents] == fx.meta["names"]
    E = len(ents)
    shapes = [fx.desc.entities[e].shape for e in range(E)]
    snaps = sorted({-1, len(fx.snaps) // 2, len(fx.snaps) - 1})
    states = [fx.state0 if k < 0 else fx.state_after[k] for k in snaps]

    # pairs: per shape combination the PER_COMBO pairs whose centres come closest in any env of any chosen state
    best = {}
    for a, b in itertools.combinations(range(E), 2):
        d = min(float(np.linalg.norm(st[a, :, 0:2] - st[b, :, 0:2], axis=-1).min()) for st in states)
        best.setdefault(tuple(sorted((shapes[a], shapes[b]))), []).append((d, a, b))
    pairs = []
    for combo, lst in sorted(best.items()):
        for d, a, b in sorted(lst)[:PER_COMBO]:
            pairs += [(a, b), (b, a)]
    pt_entities = []
    for sh in sorted(set(shapes)):
        pt_entities += [e for e in range(E) if shapes[e] == sh][:3]
    rng = np.random.default_rng(7)
    dist = np.zeros((len(snaps), len(pairs), n), np.float32)
    over = np.zeros((len(snaps), len(pairs), n), bool)
    points = np.zeros((len(snaps), len(pt_entities), n, 2), np.float32)
    pdist = np.zeros((len(snaps), len(pt_entities), n), np.float32)
    for k, st in enumerate(states):
        for e, ent in enumerate(ents):
            ent.state.pos = torch.from_numpy(st[e, :, 0:2].copy())
            ent.state.rot = torch.from_numpy(st[e, :, 4:5].copy())
        for m, (a, b) in enumerate(pairs):
            dist[k, m] = world.get_distance(ents[a], ents[b]).numpy()
            over[k, m] = world.is_overlapping(ents[a], ents[b]).numpy()
        for p, e in enumerate(pt_entities):
            scale = 0.05 if p % 2 else 0.4  # close to the surface / farther out
            points[k, p] = st[e, :, 0:2] + rng.uniform(-scale, scale, (n, 2)).astype(np.float32)
            pdist[k, p] = world.get_distance_from_point(ents[e], torch.from_numpy(points[k, p])).numpy()
    dims = np.array([[float(getattr(e.shape, "radius", 0.0)), float(getattr(e.shape, "length", 0.0)),
                      float(getattr(e.shape, "width", 0.0))] for e in ents], np.float64)
    path = os.path.join(HERE, "queries", f"{name}.npz")
    np.savez_compressed(path, snaps=np.array(snaps, np.int32), pairs=np.array(pairs, np.int32).reshape(-1, 2),
                        dist=dist, overlap=np.packbits(over.reshape(-1)), overlap_shape=np.array(over.shape, np.int32),
                        pt_entities=np.array(pt_entities, np.int32), points=points, pt_dist=pdist, dims=dims,
                        shapes=np.array(shapes, np.int32), torch=np.array(torch.__version__))
    combos = sorted({tuple(sorted((shapes[a], shapes[b]))) for a, b in pairs})
    print(f"{name}: {len(pairs)} ordered pairs, combos {combos}, overlapping {over.mean():.3f}, "
          f"{len(pt_entities)} point entities -> {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    for nm in sys.argv[1:] or list(CONFIGS):
        run(nm)
