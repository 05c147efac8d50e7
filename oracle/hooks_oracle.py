"""CPU restatement (numpy, fp32) of the scenario hooks of SURVEY 8 f1.  TEST INFRASTRUCTURE: only tests/ import this.

Pinned by tests/test_oracle_hooks.py against outputs of the UNMODIFIED reference (tests/golden/hooks/*.npz, made by
tests/golden/make_golden_hooks.py under torch 2.11.0 CPU), bit for bit.

Every operation is a single correctly rounded fp32 operation in the reference's order.  The 2-vector norm is
sqrtf(fmaf(y, y, x * x)), what torch.linalg.vector_norm computes for a last dimension of 2 (same as vnorm in
oracle/vmas_oracle.c); numpy has no fma, so `_fmaf` evaluates it in float64 and repairs the double-rounding ties."""
import numpy as np

from . import oracle as _orc
from vectorizedmultiagentsimulator_b200.worlddesc import BOX, SPHERE

F = np.float32


def _fmaf(a, b, c):
    """Correctly rounded fp32 fma(a, b, c), elementwise.  a * b is exact in float64; p + c rounds once to 53 bits, and
    the second rounding to 24 bits differs from a single rounding only when the float64 sum is inexact AND lands exactly
    halfway between two fp32 values: those (rare) elements are decided with exact rational arithmetic."""
    a64, b64, c64 = a.astype(np.float64), b.astype(np.float64), c.astype(np.float64)
    p = a64 * b64
    s = p + c64
    r = s.astype(F)
    bb = s - p
    err = (p - (s - bb)) + (c64 - bb)  # TwoSum: the part of p + c the float64 sum lost
    half = (s.view(np.uint64) & np.uint64((1 << 29) - 1)) == np.uint64(1 << 28)
    idx = np.nonzero((err != 0) & half & np.isfinite(s))
    if idx[0].size:
        from fractions import Fraction

        for i in zip(*idx):
            exact = Fraction(float(a64[i])) * Fraction(float(b64[i])) + Fraction(float(c64[i]))
            cands = (np.nextafter(r[i], F(-np.inf)), r[i], np.nextafter(r[i], F(np.inf)))
            r[i] = min((abs(Fraction(float(v)) - exact), int(F(v).view(np.uint32)) & 1, float(v)) for v in cands)[2]
    return r


def _fma_norm(d):
    """sqrtf(fmaf(y, y, x * x)) over the last axis (size 2)."""
    x, y = np.ascontiguousarray(d[..., 0], F), np.ascontiguousarray(d[..., 1], F)
    return np.sqrt(_fmaf(y, y, (x * x).astype(F)), dtype=F)


def gather_obs(parts):
    """torch.cat([...], dim=-1) of one agent: parts = list of (op, a, b, c) like vmasb200_obs_seg."""
    cols = []
    for op, a, b, c in parts:
        a = np.asarray(a)
        if a.ndim == 1:
            a = a[:, None]
        if op == 0:
            cols.append(a.astype(F))
        elif op == 1:
            cols.append((a.astype(F) - np.asarray(b, F).reshape(a.shape)).astype(F))
        elif op == 2:
            cols.append((F(c) - a.astype(F)).astype(F))
        else:
            cols.append((a != 0).astype(F))
    return np.concatenate(cols, axis=-1)


def navigation_reward(pos, goal, radius, goal_radius, shaping_in, collidable, factor, final_reward, penalty, min_coll,
                      shared_rew):
    """navigation.py:199-245, 262-273.  pos, goal [A, N, 2]; shaping_in [A, N]; collidable [A, A] (i > j): the static
    part of World.collides.  Returns a dict named like the reference's attributes."""
    A, n = pos.shape[0], pos.shape[1]
    dist = np.stack([_fma_norm(pos[a] - goal[a]) for a in range(A)])                    # :234-237
    on_goal = np.stack([dist[a] < F(goal_radius[a]) for a in range(A)])                  # :238
    shaping = (dist * F(factor)).astype(F)                                               # :240
    pos_rew_agent = (shaping_in.astype(F) - shaping).astype(F)                           # :241
    pos_rew = np.zeros(n, F)
    for a in range(A):
        pos_rew = (pos_rew + pos_rew_agent[a]).astype(F)                                 # :205-208
    all_goal = on_goal.all(axis=0)                                                       # :211-214
    final_rew = np.where(all_goal, F(final_reward), F(0)).astype(F)                      # :215
    coll = np.zeros((A, n), F)
    flags = np.zeros((A, A), bool)
    for i in range(A):
        for j in range(i):
            if not collidable[i][j]:
                continue
            nrm = _fma_norm(pos[i] - pos[j])
            flags[i, j] = bool((nrm <= F(float(radius[i]) + float(radius[j]))).any())   # core.py:2796-2800
            if flags[i, j]:
                d = ((nrm - F(radius[i])).astype(F) - F(radius[j])).astype(F)            # core.py:1830-1836
                m = d <= F(min_coll)
                coll[i] = np.where(m, coll[i] + F(penalty), coll[i]).astype(F)           # :222-228
                coll[j] = np.where(m, coll[j] + F(penalty), coll[j]).astype(F)
    reward = np.stack([(((pos_rew if shared_rew else pos_rew_agent[a]) + final_rew).astype(F) + coll[a]).astype(F) for a in range(A)])
    done = np.stack([dist[a] < F(radius[a]) for a in range(A)]).all(axis=0)              # :262-273
    return dict(dist=dist, on_goal=on_goal, shaping_out=shaping, pos_rew_agent=pos_rew_agent, pos_rew=pos_rew,
                all_goal=all_goal, final_rew=final_rew, coll_rew=coll, reward=reward, done=done, flags=flags)


def transport_reward(pkg_state, pkg_dims, goal_pos, goal_radius, shaping_in, factor, red=(0.75, 0.25, 0.25), green=(0.25, 0.75, 0.25)):
    """transport.py:130-164, 186-193.  pkg_state [P, N, 6], pkg_dims [P, 3] (radius, length, width), goal_pos [P, N, 2]."""
    P, n = pkg_state.shape[0], pkg_state.shape[1]
    rew = np.zeros(n, F)
    dist, on_goal, shaping, color = [], [], [], []
    for k in range(P):
        d = _fma_norm(pkg_state[k, :, 0:2] - goal_pos[k])                                # :141-143
        gst = np.zeros((n, 6), F)
        gst[:, 0:2] = goal_pos[k]
        og = _orc.is_overlapping(BOX, pkg_dims[k], pkg_state[k], SPHERE, (float(goal_radius[k]), 0.0, 0.0), gst)  # :144
        sh = (d * F(factor)).astype(F)                                                   # :156
        rew = np.where(~og, (rew + (shaping_in[k].astype(F) - sh).astype(F)).astype(F), rew)  # :157-160
        dist.append(d); on_goal.append(og); shaping.append(sh)
        color.append(np.where(og[:, None], np.asarray(green, F)[None], np.asarray(red, F)[None]).astype(F))  # :145-154
    on_goal = np.stack(on_goal)
    return dict(dist=np.stack(dist), on_goal=on_goal, shaping_out=np.stack(shaping), color=np.stack(color), rew=rew,
                done=on_goal.all(axis=0))                                                # :186-193


def simple_spread_reward(agent_pos, landmark_pos, radius, collide):
    """mpe/simple_spread.py:79-110.  agent_pos [A, N, 2], landmark_pos [L, N, 2]."""
    A, n = agent_pos.shape[0], agent_pos.shape[1]
    closest = [np.stack([_fma_norm(agent_pos[a] - lp) for a in range(A)], -1).min(-1) for lp in landmark_pos]  # :90-101
    rew = np.zeros(n, F)
    for s in range(A):                                                                   # :88
        for c in closest:
            rew = (rew - c).astype(F)                                                    # :102
        if collide[s]:
            for a in range(A):
                if a != s:                                                               # :104-107, core.py:1934
                    d = ((_fma_norm(agent_pos[a] - agent_pos[s]) - F(radius[a])).astype(F) - F(radius[s])).astype(F)
                    rew = np.where(d < 0, (rew - F(1)).astype(F), rew)
    return rew
