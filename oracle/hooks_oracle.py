"""CPU restatement (numpy, fp32) of the scenario hooks of SURVEY 8 f1.  TEST INFRASTRUCTURE: only tests/ import this.

Pinned by tests/test_oracle_hooks.py against outputs of the UNMODIFIED reference (tests/golden/hooks/*.npz, made by
tests/golden/make_golden_hooks.py under torch 2.11.0 CPU), bit for bit.

Every operation is a single correctly rounded fp32 operation in the reference's order.  The 2-vector norm is
sqrtf(fmaf(y, y, x * x)), what torch.linalg.vector_norm computes for a last dimension of 2 (same as vnorm in
oracle/vmas_oracle.c); numpy has no fma, so `_fmaf` evaluates it in float64 and repairs the double-rounding ties."""
import numpy as np

from . import oracle as _orc
from vectorizedmultiagentsimulator_b200.worlddesc import BOX, LINE, SPHERE

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
    for part in parts:
        op, a, b, c = part[:4]
        neg0 = len(part) > 4 and part[4] & 1
        noise = part[5] if len(part) > 5 else None
        a = np.asarray(a)
        if a.ndim == 1:
            a = a[:, None]
        if op in (0, 1) and neg0:  # football.py:1376-1397: column 0 of both operands negated before the op
            a = a.astype(F).copy()
            a[:, 0] = -a[:, 0]
            if b is not None:
                b = np.asarray(b, F).reshape(a.shape).copy()
                b[:, 0] = -b[:, 0]
        if op == 0:
            v = a.astype(F)
        elif op == 1:
            v = (a.astype(F) - np.asarray(b, F).reshape(a.shape)).astype(F)
        elif op == 2:
            v = (F(c) - a.astype(F)).astype(F)
        elif op == 4:
            v = np.cos(a.astype(F), dtype=F)  # libm, not torch's Sleef: compared within an ulp or two, not bitwise
        elif op == 5:
            v = np.sin(a.astype(F), dtype=F)
        else:
            v = (a != 0).astype(F)
        if noise is not None:
            v = (v + np.asarray(noise, F).reshape(v.shape)).astype(F)
        cols.append(v)
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


def _remainder(a, b):
    """torch.remainder for fp32 (ATen remainder_kernel): fmod, then the divisor's sign."""
    m = np.fmod(a.astype(F), F(b)).astype(F)
    fix = (m != 0) & ((F(b) < 0) != (m < 0))
    return np.where(fix, (m + F(b)).astype(F), m)


def _line_angle_dist(rot, g, g_minus_pi):
    """joint_passage.py:24-38; g / g_minus_pi: fp32 arrays or scalars (see csrc/hooks.cuh line_angle_dist)."""
    pi = F(np.pi)
    a = _remainder(rot, pi)
    d0 = np.abs((a - g).astype(F))
    d1 = np.abs((a - g_minus_pi).astype(F))
    d2 = np.abs(((a - pi).astype(F) - g).astype(F))
    return np.minimum(d0, np.minimum(d1, d2))


def joint_passage_reward(g, t):
    """joint_passage.py:387-490, 547-563 on sample t of tests/golden/hooks/joint_passage*.npz."""
    (width, pos_f, rot_f, coll_rew, e_coeff, min_coll, middle, all_rot, dim_p, _, n_agents) = [float(v) for v in g["params"]]
    A = int(n_agents)
    bodies, joint, goal, passages, walls = g["bodies"][t], g["joint"][t], g["goal"][t], g["passages"][t], g["walls"][t]
    n = joint.shape[0]
    collide = g["passage_collide"]
    joint_passed = joint[:, 1] > 0                                                            # :399
    all_passed = np.all(np.stack([bodies[a][:, 1] > F(width / 2) for a in range(A)], 1), 1)   # :400-404
    closest = np.stack([_fma_norm(joint[:, 0:2] - passages[p][:, 0:2]) for p in range(len(passages)) if not collide[p]], 1).min(1)
    sh = g["shaping_in"][t].astype(F)
    sh_pre = (closest * F(pos_f)).astype(F)
    pos_rew = np.where(~joint_passed, (F(0) + (sh[0] - sh_pre).astype(F)).astype(F), F(0)).astype(F)   # :418-420
    dist_goal = _fma_norm(joint[:, 0:2] - goal[:, 0:2])
    sh_post = (dist_goal * F(pos_f)).astype(F)
    pos_rew = np.where(joint_passed, (pos_rew + (sh[1] - sh_post).astype(F)).astype(F), pos_rew)        # :428-430
    rot_passed = all_passed if all_rot else joint_passed
    mg = float(middle) % float(np.pi)
    rsh_pre = (_line_angle_dist(joint[:, 4], F(mg), F(mg - float(np.pi))) * F(rot_f)).astype(F)
    rot_rew = np.where(~rot_passed, (F(0) + (sh[2] - rsh_pre).astype(F)).astype(F), F(0)).astype(F)
    gm = _remainder(goal[:, 4], F(np.pi))
    rot_dist_goal = _line_angle_dist(joint[:, 4], gm, (gm - F(np.pi)).astype(F))
    rsh_post = (rot_dist_goal * F(rot_f)).astype(F)
    rot_rew = np.where(rot_passed, (rot_rew + (sh[3] - rsh_post).astype(F)).astype(F), rot_rew)
    coll = np.zeros(n, F)
    if coll_rew != 0.0:
        def hit(sa, da, xa, sb, db, xb):
            return _orc.get_distance(sa, da, xa, sb, db, xb) <= F(min_coll)
        def add(m):
            nonlocal coll
            coll = np.where(m, (coll + F(coll_rew)).astype(F), coll)
        for a in range(bodies.shape[0]):                                                    # :454-466
            da = (float(g["body_radius"][a]), 0.0, 0.0)
            for p in range(len(passages)):
                if collide[p]:
                    add(hit(SPHERE, da, bodies[a], BOX, (0.0, *[float(v) for v in g["passage_dims"][p]]), passages[p]))
                for w in range(len(walls)):
                    add(hit(SPHERE, da, bodies[a], LINE, (0.0, float(g["wall_length"][w]), 0.0), walls[w]))
        for p in range(len(passages)):                                                      # :468-473
            if collide[p]:
                add(hit(BOX, (0.0, *[float(v) for v in g["passage_dims"][p]]), passages[p], LINE, (0.0, float(g["joint_length"][0]), 0.0), joint))
    energy = None
    for a in range(A):                                                                      # :476-484
        e = (_fma_norm(g["action_u"][t][a]) / F(np.sqrt(dim_p * float(g["f_range"][a]) ** 2))).astype(F)
        energy = e if energy is None else (energy + e).astype(F)
    energy_rew = ((-energy) * F(e_coeff)).astype(F)
    rew = (((pos_rew + rot_rew).astype(F) + coll).astype(F) + energy_rew).astype(F)
    done = (rot_dist_goal <= F(0.01)) & bool((dist_goal <= F(0.01)).all())                   # :547-563 ([n] * [n, 1] -> [n, n])
    return dict(pos_rew=pos_rew, rot_rew=rot_rew, collision_rew=coll, energy_expenditure=energy, energy_rew=energy_rew, rew=rew,
                all_passed=all_passed, done=done, shaping_out=np.stack([sh_pre, sh_post, rsh_pre, rsh_post]))


def joint_passage_obs_parts(g, t, a):
    """joint_passage.py:500-545 with obs_noise = 0 (the noise then is an exact zero)."""
    bodies, joint, goal, passages = g["bodies"][t], g["joint"][t], g["goal"][t], g["passages"][t]
    me = bodies[a]
    parts = [(0, me[:, 0:2], None, 0.0), (0, me[:, 2:4], None, 0.0), (1, me[:, 0:2], goal[:, 0:2], 0.0)]
    parts += [(1, me[:, 0:2], passages[p][:, 0:2], 0.0) for p in range(len(passages)) if not g["passage_collide"][p]]
    parts += [(4, goal[:, 4:5], None, 0.0), (5, goal[:, 4:5], None, 0.0)]
    if g["params"][9]:
        parts += [(4, joint[:, 4:5], None, 0.0), (5, joint[:, 4:5], None, 0.0)]
    return parts


def football_reward(g, t):
    """football.py:1118-1226 on sample t of tests/golden/hooks/football*.npz.  torch.cdist (football.py:593) on these
    small operands runs ATen's scalar loop `agg = agg + diff * diff`, which the compiler contracts: sqrt(fma(dy, dy,
    dx * dx)), the same form as linalg.vector_norm (pinned by the golden min_agent_dist_to_ball)."""
    L, bsz, gsz, scoring, f_bg, f_ab, trigger, dense, _, _ = [float(v) for v in g["params"]]
    ball = g["ball"][t]
    bx, by = ball[:, 0], ball[:, 1]
    over_right, over_left = bx > F(L / 2 + bsz / 2), bx < F(-L / 2 - bsz / 2)
    goal_mask = (by <= F(gsz / 2)) & (by >= F(-gsz / 2))
    blue_score, red_score = over_right & goal_mask, over_left & goal_mask
    sparse_blue = ((F(scoring) * blue_score.astype(F)).astype(F) - (F(scoring) * red_score.astype(F)).astype(F)).astype(F)
    sparse = [sparse_blue, (-sparse_blue).astype(F)]
    moving = _fma_norm(ball[:, 2:4]) > F(1e-6)
    out = dict(done=blue_score | red_score, sparse=np.stack(sparse), rew=[], dist_goal=[], pos_rew=[], min_dist=[], pos_rew_agent=[], shaping_out=[])
    for k, name in enumerate(("blue", "red")):
        d = np.zeros_like(bx)
        if dense:
            dg = _fma_norm(ball[:, 0:2] - g["goal_pos"][k][None])
            sh = (dg * F(f_bg)).astype(F)
            pr = (g["shaping_in"][t][2 * k] - sh).astype(F)
            md = np.stack([_fma_norm(a[:, 0:2] - ball[:, 0:2]) for a in g[name][t]], 1).min(1)
            sha = (md * F(f_ab)).astype(F)
            pra = np.where((md < F(trigger)) | moving, F(0), (g["shaping_in"][t][2 * k + 1] - sha).astype(F)).astype(F)
            d = (pr + pra).astype(F)
            out["dist_goal"].append(dg); out["pos_rew"].append(pr); out["min_dist"].append(md); out["pos_rew_agent"].append(pra)
            out["shaping_out"] += [sh, sha]
        out["rew"].append((sparse[k] + d).astype(F))
    return out


def football_obs_parts(g, t, team, i):
    """football.py:1228-1462 for agent i of `team` ("blue" / "red"), flat observation, no shooting."""
    blue = team == "blue"
    mine, other = (g["blue"][t], g["red"][t]) if blue else (g["red"][t], g["blue"][t])
    mine_f, other_f = (g["blue_force"][t], g["red_force"][t]) if blue else (g["red_force"][t], g["blue_force"][t])
    ball, bf = g["ball"][t], g["ball_force"][t]
    n = ball.shape[0]
    goal = np.broadcast_to(g["goal_pos"][0 if blue else 1][None], (n, 2))
    f = 0 if blue else 1
    me, mf = mine[i], mine_f[i]
    parts = [(0, mf, None, 0.0, f), (1, me[:, 0:2], ball[:, 0:2], 0.0, f), (1, me[:, 2:4], ball[:, 2:4], 0.0, f),
             (1, ball[:, 0:2], goal, 0.0, f), (0, ball[:, 2:4], None, 0.0, f), (0, bf, None, 0.0, f),
             (1, me[:, 0:2], goal, 0.0, f), (0, me[:, 2:4], None, 0.0, f)]
    groups = []
    if g["params"][9] and len(other):
        groups.append((other, other_f, range(len(other))))
    if g["params"][8]:
        groups.append((mine, mine_f, [j for j in range(len(mine)) if j != i]))
    for st, fo, idx in groups:
        for j in idx:
            parts += [(1, me[:, 0:2], st[j][:, 0:2], 0.0, f), (1, me[:, 2:4], st[j][:, 2:4], 0.0, f), (0, st[j][:, 2:4], None, 0.0, f),
                      (0, fo[j], None, 0.0, f)]
    return parts


def set_action_discrete(action, nvec, u_range, u_mult, multidiscrete):
    """Environment._set_action for discrete physical actions (environment.py:660-710): integer `action` [n, 1] (flat index
    into the cartesian product of the action spaces) or [n, k] (multi-discrete) -> (u [n, k] fp32, out_of_bounds).
    Floor division / modulo as torch's on integers; fp32 arithmetic in the reference's order."""
    a = np.asarray(action).astype(np.int64)
    k = len(nvec)
    if multidiscrete:
        comp = [a[:, i].copy() for i in range(k)]
    else:
        flat, comp = a[:, 0].copy(), []
        for i in range(k):
            n = int(np.prod(nvec[i + 1:], dtype=np.int64))
            comp.append(flat // n)
            flat = flat % n
    u = np.zeros((a.shape[0], k), np.float32)
    bad = False
    for i, n in enumerate(nvec):
        p = comp[i]
        bad = bad or bool(((p < 0) | (p >= n)).any())
        if n % 2:
            stay, dec = p == 0, (p > 0) & (p <= n // 2)
            p = np.where(stay, n // 2, np.where(dec, p - 1, p))
        um = np.float32(u_range[i])
        u[:, i] = (p.astype(np.float32) / np.float32(n - 1)) * (np.float32(2) * um) - um
        u[:, i] = u[:, i] * np.float32(u_mult[i])
    return u, bad
