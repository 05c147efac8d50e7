"""Fused scenario hooks: what ``Environment._get_from_scenario`` (environment.py:253-308) asks of a
scenario after every ``World.step`` - ``reward`` / ``observation`` / ``done`` / ``info`` - as one or two launches per
step (include/vmasb200.h: vmasb200_gather_obs, vmasb200_navigation_reward, vmasb200_transport_reward) instead of tens
to hundreds of small torch ops with host synchronisations (``World.collides``, boolean-mask indexing).

Opt-in and keyed by scenario (``dropin.enable(env, hooks=True)``): the hook API stays the reference's - the same bound
methods are patched on ONE scenario instance, they return tensors of the same shape / dtype, and every attribute the
reference's hooks leave behind (``agent.pos_shaping``, ``agent.on_goal``, ``scenario.pos_rew``, ``package.on_goal``,
``package.color`` ...) is left behind here too, written in place where the reference writes in place and rebound where
it rebinds.  A configuration a fused hook does not cover (autograd in play, shapes it does not know, non-CUDA tensors)
takes the saved original method: the reference's torch code is the fallback of the HOOKS, never of the physics.

Scenarios: ``navigation`` (navigation.py:199-280), ``transport`` (transport.py:130-193), ``simple_spread``
(mpe/simple_spread.py:79-129), ``joint_passage`` (joint_passage.py:387-563) and ``football`` (football.py:1118-1467)."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

from . import _lib
from .queries import CPose, make_pose, shape_of
from .worlddesc import SPHERE

OBS_COPY, OBS_SUB, OBS_CSUB, OBS_BOOL, OBS_COS, OBS_SIN = 0, 1, 2, 3, 4, 5
OBS_NEG0 = 1
NAV_MAX_AGENTS, TRANSPORT_MAX_PACKAGES, OBS_MAX_SEGS, OBS_MAX_OUTS, OBS_MAX_WIDTH, SPREAD_MAX = 16, 8, 128, 32, 352, 16


class CVec(C.Structure):
    _fields_ = [("p", C.c_void_p), ("stride_env", C.c_int64), ("stride_k", C.c_int64)]


class CObsSeg(C.Structure):
    _fields_ = [("a", CVec), ("b", CVec), ("op", C.c_int32), ("c", C.c_float), ("width", C.c_int32), ("out", C.c_int32),
                ("col0", C.c_int32), ("flags", C.c_int32), ("noise", CVec)]


class CNavAgent(C.Structure):
    _fields_ = [("pos", CVec), ("goal_pos", CVec), ("radius", C.c_double), ("goal_radius", C.c_double),
                ("pos_shaping", C.c_void_p), ("pos_shaping_out", C.c_void_p), ("distance_to_goal", C.c_void_p),
                ("on_goal", C.c_void_p), ("pos_rew", C.c_void_p), ("agent_collision_rew", C.c_void_p),
                ("reward", C.c_void_p)]


class CNavigation(C.Structure):
    _fields_ = [("n_agents", C.c_int32), ("shared_rew", C.c_int32), ("pos_shaping_factor", C.c_double),
                ("final_reward", C.c_double), ("agent_collision_penalty", C.c_double),
                ("min_collision_distance", C.c_double), ("agent", CNavAgent * NAV_MAX_AGENTS),
                ("collidable", (C.c_uint8 * NAV_MAX_AGENTS) * NAV_MAX_AGENTS), ("pos_rew", C.c_void_p),
                ("final_rew", C.c_void_p), ("all_goal_reached", C.c_void_p), ("done", C.c_void_p),
                ("pair_flags", C.c_void_p)]


class CTransportPackage(C.Structure):
    _fields_ = [("package", CPose), ("goal", CPose), ("global_shaping", C.c_void_p), ("global_shaping_out", C.c_void_p),
                ("dist_to_goal", C.c_void_p), ("on_goal", C.c_void_p), ("color", C.c_void_p)]


class CTransport(C.Structure):
    _fields_ = [("n_packages", C.c_int32), ("shaping_factor", C.c_double), ("color_off_goal", C.c_float * 3),
                ("color_on_goal", C.c_float * 3), ("package", CTransportPackage * TRANSPORT_MAX_PACKAGES),
                ("rew", C.c_void_p), ("done", C.c_void_p)]


class CSimpleSpread(C.Structure):
    _fields_ = [("n_agents", C.c_int32), ("n_landmarks", C.c_int32), ("agent_pos", CVec * SPREAD_MAX),
                ("landmark_pos", CVec * SPREAD_MAX), ("agent_radius", C.c_double * SPREAD_MAX),
                ("collide", C.c_uint8 * SPREAD_MAX), ("rew", C.c_void_p)]


JP_MAX_BODIES, JP_MAX_PASSAGES, JP_MAX_WALLS, FB_MAX_TEAM = 4, 24, 8, 16


class CJointPassage(C.Structure):
    _fields_ = [("n_agents", C.c_int32), ("n_bodies", C.c_int32), ("n_passages", C.c_int32), ("n_walls", C.c_int32),
                ("all_passed_rot", C.c_int32), ("dim_p", C.c_int32), ("body", CPose * JP_MAX_BODIES),
                ("action_u", CVec * JP_MAX_BODIES), ("f_range", C.c_double * JP_MAX_BODIES),
                ("passage", CPose * JP_MAX_PASSAGES), ("passage_collide", C.c_uint8 * JP_MAX_PASSAGES),
                ("wall", CPose * JP_MAX_WALLS), ("joint", CPose), ("goal_pos", CVec), ("goal_rot", CVec),
                ("passage_width", C.c_double), ("pos_shaping_factor", C.c_double), ("rot_shaping_factor", C.c_double),
                ("collision_reward", C.c_double), ("energy_reward_coeff", C.c_double), ("min_collision_distance", C.c_double),
                ("middle_angle", C.c_double), ("shaping_in", C.c_void_p * 4), ("shaping_out", C.c_void_p * 4),
                ("pos_rew", C.c_void_p), ("rot_rew", C.c_void_p), ("collision_rew", C.c_void_p),
                ("energy_expenditure", C.c_void_p), ("energy_rew", C.c_void_p), ("rew", C.c_void_p),
                ("all_passed", C.c_void_p), ("done", C.c_void_p), ("done_flag", C.c_void_p)]


class CFootballTeam(C.Structure):
    _fields_ = [("n_agents", C.c_int32), ("dense", C.c_int32), ("agent_pos", CVec * FB_MAX_TEAM), ("goal_pos", C.c_float * 2),
                ("pos_shaping", C.c_void_p), ("pos_shaping_out", C.c_void_p), ("pos_shaping_agent", C.c_void_p),
                ("pos_shaping_agent_out", C.c_void_p), ("distance_to_goal", C.c_void_p), ("pos_rew", C.c_void_p),
                ("min_agent_dist_to_ball", C.c_void_p), ("pos_rew_agent", C.c_void_p), ("dense_reward", C.c_void_p),
                ("sparse_reward", C.c_void_p), ("reward", C.c_void_p)]


class CFootball(C.Structure):
    _fields_ = [("ball_pos", CVec), ("ball_vel", CVec), ("pitch_length", C.c_double), ("ball_size", C.c_double),
                ("goal_size", C.c_double), ("scoring_reward", C.c_double), ("pos_shaping_factor_ball_goal", C.c_double),
                ("pos_shaping_factor_agent_ball", C.c_double), ("distance_to_ball_trigger", C.c_double),
                ("team", CFootballTeam * 2), ("done", C.c_void_p)]


def _vec(v: CVec, t):
    v.p, v.stride_env = t.data_ptr(), t.stride(0)
    v.stride_k = t.stride(1) if t.dim() > 1 else 0


class ObsGather:
    """torch.cat([...], dim=-1) of every agent in one launch.  ``spec`` = per agent a list of (op, a, b, c) with a / b
    tensors [N, w] (or [N] -> one column), optionally followed by flags (OBS_NEG0) and a noise tensor [N, w] that is
    added to the op's result."""

    def __init__(self, torch, n: int, device):
        self.torch, self.n, self.device = torch, n, device
        self.segs = (CObsSeg * OBS_MAX_SEGS)()
        self.outs = (C.c_void_p * OBS_MAX_OUTS)()
        self.widths = (C.c_int32 * OBS_MAX_OUTS)()

    @staticmethod
    def usable(torch, t, boolean=False) -> bool:
        return (torch.is_tensor(t) and t.is_cuda and t.dim() in (1, 2) and
                (t.dtype in (torch.bool, torch.uint8) if boolean else t.dtype == torch.float32) and
                not (torch.is_grad_enabled() and t.requires_grad))

    def run(self, spec) -> Optional[List]:
        """One launch per group of observations that fits a table (OBS_MAX_SEGS segments, OBS_MAX_OUTS outputs):
        one launch for the small scenarios, two for football 3 v 3 (36 segments per agent)."""
        outs, group = [], []
        n_segs = 0
        for parts in spec:
            if len(parts) > OBS_MAX_SEGS:
                return None
            if group and (n_segs + len(parts) > OBS_MAX_SEGS or len(group) == OBS_MAX_OUTS):
                r = self._run(group)
                if r is None:
                    return None
                outs += r
                group, n_segs = [], 0
            group.append(parts)
            n_segs += len(parts)
        if group:
            r = self._run(group)
            if r is None:
                return None
            outs += r
        return outs

    def _run(self, spec) -> Optional[List]:
        torch, n = self.torch, self.n
        s, outs = 0, []
        for o, parts in enumerate(spec):
            col = 0
            for part in parts:
                op, a, b, c = part[:4]
                flags = part[4] if len(part) > 4 else 0
                noise = part[5] if len(part) > 5 else None
                if not self.usable(torch, a, op == OBS_BOOL) or a.shape[0] != n:
                    return None
                w = a.shape[1] if a.dim() == 2 else 1
                g = self.segs[s]
                _vec(g.a, a)
                if op == OBS_SUB:
                    if not self.usable(torch, b) or tuple(b.shape) != tuple(a.shape):
                        return None
                    _vec(g.b, b)
                else:
                    g.b.p = None
                if noise is not None:
                    if not self.usable(torch, noise) or noise.shape[0] != n or (noise.shape[1] if noise.dim() == 2 else 1) != w:
                        return None
                    _vec(g.noise, noise)
                else:
                    g.noise.p = None
                g.op, g.c, g.width, g.out, g.col0, g.flags = op, float(c), w, o, col, int(flags)
                col += w
                s += 1
            if not 1 <= col <= OBS_MAX_WIDTH:
                return None
            out = torch.empty(n, col, device=self.device)
            outs.append(out)
            self.outs[o], self.widths[o] = out.data_ptr(), col
        _lib.check(_lib.lib().vmasb200_gather_obs(self.segs, s, self.outs, self.widths, len(spec), n,
                                                  _lib.raw_stream(torch, self.device)), "vmasb200_gather_obs")
        return outs


def scenario_file(scenario) -> str:
    """'navigation' for a Scenario defined in .../scenarios/navigation.py.  The reference loads scenario modules from
    their file path under an EMPTY module name (vmas/scenarios/__init__.py), so the class's code object is asked."""
    import os

    code = getattr(getattr(type(scenario), "reward", None), "__code__", None)
    return os.path.splitext(os.path.basename(code.co_filename))[0] if code is not None else ""


class ScenarioHooks:
    """Base: patches reward / observation / done of one scenario instance; ``remove`` restores them."""

    NAMES = ("reward", "observation", "done")

    def __init__(self, env, binding):
        self.env, self.b = env, binding
        self.torch, self.scenario, self.world = binding.torch, env.scenario, binding.world
        self.n, self.device = binding.n, binding.device
        self.gather = ObsGather(self.torch, self.n, self.device)
        self.n_fused = {"reward": 0, "observation": 0, "done": 0}
        self.n_fallback = {"reward": 0, "observation": 0, "done": 0}
        self._obs = None
        self._obs_key = None
        self._done = None
        self._done_key = None
        self._rew_ok = False
        self.orig = {k: getattr(self.scenario, k) for k in self.NAMES}
        self.attach()

    def _not_covered(self, kind):
        """The fused hook cannot take this call (tensor layouts / shapes it does not know, autograd): the reference's
        torch hook answers; counted, a one-time warning, an error under binding.strict (dropin.WorldBinding._fallback)."""
        self.n_fallback[kind] += 1
        self.b._fallback("hook " + kind, f"{type(self.scenario).__module__}: operand tensors are not CUDA fp32 of the expected "
                                          "shapes, or autograd is in play")

    def attach(self):
        sc = self.scenario
        sc.reward = lambda agent: self.reward(agent)
        sc.observation = lambda agent: self.observation(agent)
        sc.done = lambda: self.done()

    def remove(self):
        for k in self.NAMES:
            self.scenario.__dict__.pop(k, None)

    def state_key(self):
        """Changes whenever the world's state may have: a step, an in-place write through a state view (version
        counter of the SoA buffers), a rebound state tensor, a reset (binding.dirty)."""
        b = self.b
        if b.dirty or b.plan is None:
            return None
        for e, views in b._dyn_views:
            st = e.state
            if st._pos is not views[0] or st._vel is not views[1] or st._rot is not views[2] or st._ang_vel is not views[3]:
                return None
        return (b.n_steps, id(b.dyn), b.dyn._version) + tuple((id(t), t._version) for t in self.extra_state())

    def extra_state(self):
        return ()

    def _f32_vec(self, t, k=None) -> bool:
        torch = self.torch
        return (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32 and t.shape[0] == self.n and
                (k is None or (t.dim() == 2 and t.shape[1] == k)) and not (torch.is_grad_enabled() and t.requires_grad))

    def _flat(self, t) -> bool:
        return self._f32_vec(t) and t.dim() == 1 and t.is_contiguous()

    # -- observation / done share the cache discipline
    def observation(self, agent):
        agents = self.world.agents
        key = self.state_key()
        if agent is agents[0] or self._obs is None or key is None or key != self._obs_key:
            spec = self.obs_spec() if key is not None else None
            self._obs = self.gather.run(spec) if spec is not None else None
            self._obs_key = key
            if self._obs is not None:
                self.n_fused["observation"] += 1
        if self._obs is None:
            if key is None:
                self.n_fallback["observation"] += 1  # between a reset and the next step: the reference's code answers, by design
            else:
                self._not_covered("observation")
            return self.orig["observation"](agent)
        return self._obs[self._agent_index(agent)]

    def _agent_index(self, agent) -> int:
        for i, a in enumerate(self.world.agents):
            if a is agent:
                return i
        raise ValueError("agent is not in the world")

    def done(self):
        key = self.state_key()
        if self._done is not None and key is not None and key == self._done_key:
            self.n_fused["done"] += 1
            return self._done
        self.n_fallback["done"] += 1
        return self.orig["done"]()


class NavigationHooks(ScenarioHooks):
    """navigation.py:199-280.  ``reward`` for the first agent launches the whole computation (as the reference's
    ``is_first`` branch does); the other agents read what it left."""

    @staticmethod
    def matches(scenario) -> bool:
        return (scenario_file(scenario) == "navigation" and
                all(hasattr(scenario, k) for k in ("pos_shaping_factor", "final_reward", "agent_collision_penalty",
                                                   "min_collision_distance", "shared_rew", "pos_rew", "final_rew")))

    def __init__(self, env, binding):
        super().__init__(env, binding)
        self.cn = CNavigation()
        self.flags = self.torch.zeros(4, dtype=self.torch.int32, device=self.device)
        self._rewards = {}

    def extra_state(self):
        return tuple(a.goal.state.pos for a in self.world.agents)

    def graph_inputs(self):
        """(object, attribute) of every tensor the fused hooks READ that a reset may rebind (graph.py)."""
        out = [(self.scenario, "pos_rew"), (self.scenario, "final_rew")]
        for a in self.world.agents:
            out += [(a.goal.state, "pos"), (a, "pos_shaping"), (a, "agent_collision_rew")]
        return out

    def _static_collides(self, a, b) -> bool:
        """World.collides without its batch-wide distance test (core.py:2788-2795)."""
        w = self.world
        if (not a.collides(b)) or (not b.collides(a)) or a is b:
            return False
        if not a.movable and not a.rotatable and not b.movable and not b.rotatable:
            return False
        return {a.shape.__class__, b.shape.__class__} in w._collidable_pairs

    def _reward_usable(self) -> bool:
        sc, agents = self.scenario, self.world.agents
        if not 1 <= len(agents) <= NAV_MAX_AGENTS or not self._flat(sc.pos_rew) or not self._flat(sc.final_rew):
            return False
        for a in agents:
            sh, gsh = shape_of(a), shape_of(a.goal)
            if sh is None or sh[0] != SPHERE or gsh is None or gsh[0] != SPHERE:
                return False
            if not (self._f32_vec(a.state.pos, 2) and self._f32_vec(a.goal.state.pos, 2) and
                    self._flat(getattr(a, "pos_shaping", None)) and self._flat(getattr(a, "agent_collision_rew", None))):
                return False
        return True

    def reward(self, agent):
        sc, agents, torch = self.scenario, self.world.agents, self.torch
        if agent is not agents[0]:
            if self._rew_ok:
                return self._rewards[id(agent)]
            return self.orig["reward"](agent)
        self._rew_ok = self._reward_usable()
        if not self._rew_ok:
            self._not_covered("reward")
            self._done = None
            return self.orig["reward"](agent)
        n, A, cn, inplace = self.n, len(agents), self.cn, self.b.static_io
        fbuf = torch.empty(4 * A, n, device=self.device)          # per agent: pos_shaping, distance, pos_rew, reward
        bbuf = torch.empty(A + 2, n, dtype=torch.uint8, device=self.device)  # per agent on_goal; all_goal_reached; done
        cn.n_agents, cn.shared_rew = A, int(bool(sc.shared_rew))
        cn.pos_shaping_factor, cn.final_reward = float(sc.pos_shaping_factor), float(sc.final_reward)
        cn.agent_collision_penalty, cn.min_collision_distance = float(sc.agent_collision_penalty), float(sc.min_collision_distance)
        fp, bp, row = fbuf.data_ptr(), bbuf.data_ptr(), 4 * n
        for i, a in enumerate(agents):
            g = cn.agent[i]
            _vec(g.pos, a.state.pos)
            _vec(g.goal_pos, a.goal.state.pos)
            g.radius, g.goal_radius = float(a.shape.radius), float(a.goal.shape.radius)
            g.pos_shaping = a.pos_shaping.data_ptr()
            # static_io: the shaping term is carried IN PLACE (a replayed CUDA graph must find it where it left it)
            g.pos_shaping_out = g.pos_shaping if inplace else fp + (4 * i) * row
            g.distance_to_goal = fp + (4 * i + 1) * row
            g.pos_rew = fp + (4 * i + 2) * row
            g.reward = fp + (4 * i + 3) * row
            g.on_goal = bp + i * n
            g.agent_collision_rew = a.agent_collision_rew.data_ptr()  # written in place (navigation.py:209, 222-228)
            for j in range(i):
                cn.collidable[i][j] = int(self._static_collides(a, agents[j]))
        cn.pos_rew, cn.final_rew = sc.pos_rew.data_ptr(), sc.final_rew.data_ptr()  # in place (navigation.py:203-204)
        cn.all_goal_reached, cn.done = bp + A * n, bp + (A + 1) * n
        cn.pair_flags = self.flags.data_ptr()
        _lib.check(_lib.lib().vmasb200_navigation_reward(C.byref(cn), n, _lib.raw_stream(torch, self.device)),
                   "vmasb200_navigation_reward")
        bools = bbuf.view(torch.bool)
        for i, a in enumerate(agents):
            if not inplace:
                a.pos_shaping = fbuf[4 * i]
            a.distance_to_goal, a.pos_rew = fbuf[4 * i + 1], fbuf[4 * i + 2]
            a.on_goal = bools[i]
            self._rewards[id(a)] = fbuf[4 * i + 3]
        sc.all_goal_reached = bools[A]
        self._done, self._done_key = bools[A + 1], self.state_key()
        self.n_fused["reward"] += 1
        return self._rewards[id(agent)]

    def obs_spec(self):
        sc, agents = self.scenario, self.world.agents
        spec = []
        for a in agents:
            parts = [(OBS_COPY, a.state.pos, None, 0.0), (OBS_COPY, a.state.vel, None, 0.0)]
            for g in (agents if sc.observe_all_goals else [a]):
                parts.append((OBS_SUB, a.state.pos, g.goal.state.pos, 0.0))
            if sc.collisions:
                s0 = a.sensors[0]
                parts.append((OBS_CSUB, s0.measure(), None, float(s0._max_range)))
            spec.append(parts)
        return spec


class TransportHooks(ScenarioHooks):
    """transport.py:130-193."""

    @staticmethod
    def matches(scenario) -> bool:
        return (scenario_file(scenario) == "transport" and
                all(hasattr(scenario, k) for k in ("packages", "shaping_factor")))

    def __init__(self, env, binding):
        super().__init__(env, binding)
        self.ct = CTransport()
        color = getattr(type(self.scenario).reward, "__globals__", {}).get("Color")  # vmas.simulator.utils.Color
        red = color.RED.value if color is not None else (0.75, 0.25, 0.25)
        green = color.GREEN.value if color is not None else (0.25, 0.75, 0.25)
        for k in range(3):
            self.ct.color_off_goal[k], self.ct.color_on_goal[k] = float(red[k]), float(green[k])

    def extra_state(self):
        out = []
        for p in self.scenario.packages:
            out += [p.goal.state.pos, p.on_goal]
        return tuple(out)

    def graph_inputs(self):
        out = []
        for p in self.scenario.packages:
            out += [(p.goal.state, "pos"), (p.goal.state, "rot"), (p, "global_shaping"), (p, "on_goal")]
        return out

    def _reward_usable(self) -> bool:
        pk = self.scenario.packages
        if not 1 <= len(pk) <= TRANSPORT_MAX_PACKAGES:
            return False
        for p in pk:
            if shape_of(p) is None or shape_of(p.goal) is None:
                return False
            for e in (p, p.goal):
                if not self._f32_vec(e.state.pos, 2) or (shape_of(e)[0] != SPHERE and not self._f32_vec(e.state.rot)):
                    return False
            if not self._flat(getattr(p, "global_shaping", None)):
                return False
        return True

    def reward(self, agent):
        sc, torch = self.scenario, self.torch
        if agent is not self.world.agents[0]:
            return sc.rew if self._rew_ok else self.orig["reward"](agent)
        self._rew_ok = self._reward_usable()
        if not self._rew_ok:
            self._not_covered("reward")
            self._done = None
            return self.orig["reward"](agent)
        n, ct, pk, inplace = self.n, self.ct, sc.packages, self.b.static_io
        P = len(pk)
        fbuf = torch.empty(2 * P + 1, n, device=self.device)   # per package: global_shaping, dist_to_goal; rew
        bbuf = torch.empty(P + 1, n, dtype=torch.uint8, device=self.device)
        colors = [torch.empty(n, 3, device=self.device) for _ in pk]
        ct.n_packages, ct.shaping_factor = P, float(sc.shaping_factor)
        keep = []
        fp, bp, row = fbuf.data_ptr(), bbuf.data_ptr(), 4 * n
        for k, p in enumerate(pk):
            g = ct.package[k]
            pa, ka = make_pose(*shape_of(p), p.state.pos, p.state.rot, torch)
            pg, kg = make_pose(*shape_of(p.goal), p.goal.state.pos, p.goal.state.rot, torch)
            keep += ka + kg
            g.package, g.goal = pa, pg
            g.global_shaping = p.global_shaping.data_ptr()
            g.global_shaping_out = g.global_shaping if inplace else fp + (2 * k) * row  # see NavigationHooks.reward
            g.dist_to_goal = fp + (2 * k + 1) * row
            g.on_goal, g.color = bp + k * n, colors[k].data_ptr()
        ct.rew, ct.done = fp + (2 * P) * row, bp + P * n
        _lib.check(_lib.lib().vmasb200_transport_reward(C.byref(ct), n, _lib.raw_stream(torch, self.device)),
                   "vmasb200_transport_reward")
        bools = bbuf.view(torch.bool)
        for k, p in enumerate(pk):
            if not inplace:
                p.global_shaping = fbuf[2 * k]
            p.dist_to_goal, p.on_goal, p.color = fbuf[2 * k + 1], bools[k], colors[k]
        sc.rew = fbuf[2 * P]
        self._done, self._done_key = bools[P], self.state_key()
        self.n_fused["reward"] += 1
        return sc.rew

    def obs_spec(self):
        spec = []
        for a in self.world.agents:
            parts = [(OBS_COPY, a.state.pos, None, 0.0), (OBS_COPY, a.state.vel, None, 0.0)]
            for p in self.scenario.packages:
                parts += [(OBS_SUB, p.state.pos, p.goal.state.pos, 0.0), (OBS_SUB, p.state.pos, a.state.pos, 0.0),
                          (OBS_COPY, p.state.vel, None, 0.0), (OBS_BOOL, p.on_goal, None, 0.0)]
            spec.append(parts)
        return spec


class SimpleSpreadHooks(ScenarioHooks):
    """mpe/simple_spread.py:79-129.  The scenario keeps BaseScenario.done (scenario.py:326: all False, built from a
    host list every call - a host-to-device copy no CUDA graph can hold); here the same values come from a device fill."""

    @staticmethod
    def matches(scenario) -> bool:
        return scenario_file(scenario) == "simple_spread" and hasattr(scenario, "obs_agents")

    def __init__(self, env, binding):
        super().__init__(env, binding)
        self.cs = CSimpleSpread()

    def extra_state(self):
        return tuple(l.state.pos for l in self.world.landmarks)

    def graph_inputs(self):
        return [(l.state, "pos") for l in self.world.landmarks]

    def _reward_usable(self) -> bool:
        agents, lms = self.world.agents, self.world.landmarks
        if not (1 <= len(agents) <= SPREAD_MAX and len(lms) <= SPREAD_MAX):
            return False
        for a in agents:
            sh = shape_of(a)
            if sh is None or sh[0] != SPHERE or not self._f32_vec(a.state.pos, 2):
                return False
        return all(self._f32_vec(l.state.pos, 2) for l in lms)

    def reward(self, agent):
        sc, torch, agents = self.scenario, self.torch, self.world.agents
        if agent is not agents[0]:
            return sc.rew if self._rew_ok else self.orig["reward"](agent)
        self._rew_ok = self._reward_usable()
        if not self._rew_ok:
            self._not_covered("reward")
            return self.orig["reward"](agent)
        cs, lms = self.cs, self.world.landmarks
        cs.n_agents, cs.n_landmarks = len(agents), len(lms)
        for i, a in enumerate(agents):
            _vec(cs.agent_pos[i], a.state.pos)
            cs.agent_radius[i], cs.collide[i] = float(a.shape.radius), int(bool(a.collide))
        for i, l in enumerate(lms):
            _vec(cs.landmark_pos[i], l.state.pos)
        rew = torch.empty(self.n, device=self.device)  # the reference binds a fresh tensor too (simple_spread.py:83)
        cs.rew = rew.data_ptr()
        _lib.check(_lib.lib().vmasb200_simple_spread_reward(C.byref(cs), self.n, _lib.raw_stream(torch, self.device)),
                   "vmasb200_simple_spread_reward")
        sc.rew = rew
        self.n_fused["reward"] += 1
        return rew

    def done(self):
        self.n_fused["done"] += 1
        return self.torch.zeros(self.n, dtype=self.torch.bool, device=self.device)

    def obs_spec(self):
        sc, agents = self.scenario, self.world.agents
        spec = []
        for a in agents:
            parts = [(OBS_COPY, a.state.pos, None, 0.0), (OBS_COPY, a.state.vel, None, 0.0)]
            parts += [(OBS_SUB, l.state.pos, a.state.pos, 0.0) for l in self.world.landmarks]
            if sc.obs_agents:
                parts += [(OBS_SUB, o.state.pos, a.state.pos, 0.0) for o in agents if o is not a]
            spec.append(parts)
        return spec


class JointPassageHooks(ScenarioHooks):
    """joint_passage.py:387-563.  The observation draws its noise from torch's generator exactly as the reference does
    (one ``uniform_`` per observation operand, even when ``obs_noise`` is 0), so the random streams stay in step."""

    SHAPING = ("pos_shaping_pre", "pos_shaping_post", "rot_shaping_pre", "rot_shaping_post")

    @staticmethod
    def matches(scenario) -> bool:
        return (scenario_file(scenario) == "joint_passage" and
                all(hasattr(scenario, k) for k in ("joint", "passages", "walls", "goal", "pos_rew", "rot_rew", "collision_rew",
                                                   "all_passed_rot", "min_collision_distance", "middle_angle")))

    def __init__(self, env, binding):
        super().__init__(env, binding)
        self.cj = CJointPassage()
        self.flag = self.torch.zeros(1, dtype=self.torch.int32, device=self.device)

    def extra_state(self):
        sc = self.scenario
        return (sc.goal.state.pos, sc.goal.state.rot) + tuple(a.action.u for a in self.world.agents if a.action.u is not None)

    def graph_inputs(self):
        sc = self.scenario
        out = [(sc.goal.state, "pos"), (sc.goal.state, "rot")] + [(sc.joint, k) for k in self.SHAPING]
        for e in list(sc.passages) + list(sc.walls):
            out += [(e.state, "pos"), (e.state, "rot")]
        return out

    def _bodies(self):
        sc = self.scenario
        return list(self.world.agents) + ([sc.mass] if getattr(sc, "asym_package", False) else [])

    def _reward_usable(self) -> bool:
        sc, agents = self.scenario, self.world.agents
        bodies = self._bodies()
        if not (1 <= len(agents) <= 2 and len(bodies) <= JP_MAX_BODIES and 1 <= len(sc.passages) <= JP_MAX_PASSAGES and
                len(sc.walls) <= JP_MAX_WALLS and any(not p.collide for p in sc.passages)):
            return False
        for e in bodies + list(sc.passages) + list(sc.walls) + [sc.joint.landmark]:
            sh = shape_of(e)
            if sh is None or not self._f32_vec(e.state.pos, 2) or (sh[0] != SPHERE and not self._f32_vec(e.state.rot)):
                return False
        if shape_of(sc.joint.landmark)[0] == SPHERE:
            return False
        for a in agents:
            if not self._f32_vec(a.action.u, 2) or a.f_range is None:
                return False
        if not (self._f32_vec(sc.goal.state.pos, 2) and self._f32_vec(sc.goal.state.rot)):
            return False
        if not all(self._flat(getattr(sc.joint, k, None)) for k in self.SHAPING):
            return False
        return all(self._flat(getattr(sc, k, None)) for k in ("pos_rew", "rot_rew", "collision_rew"))

    def reward(self, agent):
        sc, torch, agents = self.scenario, self.torch, self.world.agents
        if agent is not agents[0]:
            return sc.rew if self._rew_ok else self.orig["reward"](agent)
        self._rew_ok = self._reward_usable()
        if not self._rew_ok:
            self._not_covered("reward")
            self._done = None
            return self.orig["reward"](agent)
        n, cj, inplace = self.n, self.cj, self.b.static_io
        bodies = self._bodies()
        keep = []

        def pose(dst, e):
            p, k = make_pose(*shape_of(e), e.state.pos, e.state.rot, torch)
            keep.extend(k)
            C.memmove(C.addressof(dst), C.addressof(p), C.sizeof(CPose))

        cj.n_agents, cj.n_bodies, cj.n_passages, cj.n_walls = len(agents), len(bodies), len(sc.passages), len(sc.walls)
        cj.all_passed_rot, cj.dim_p = int(bool(sc.all_passed_rot)), int(self.world.dim_p)
        for i, e in enumerate(bodies):
            pose(cj.body[i], e)
        for i, a in enumerate(agents):
            _vec(cj.action_u[i], a.action.u)
            cj.f_range[i] = float(a.f_range)
        for i, p in enumerate(sc.passages):
            pose(cj.passage[i], p)
            cj.passage_collide[i] = int(bool(p.collide))
        for i, w in enumerate(sc.walls):
            pose(cj.wall[i], w)
        pose(cj.joint, sc.joint.landmark)
        _vec(cj.goal_pos, sc.goal.state.pos)
        _vec(cj.goal_rot, sc.goal.state.rot)
        cj.passage_width, cj.pos_shaping_factor = float(sc.passage_width), float(sc.pos_shaping_factor)
        cj.rot_shaping_factor, cj.collision_reward = float(sc.rot_shaping_factor), float(sc.collision_reward)
        cj.energy_reward_coeff, cj.min_collision_distance = float(sc.energy_reward_coeff), float(sc.min_collision_distance)
        cj.middle_angle = float(sc.middle_angle)
        fbuf = torch.empty(7, n, device=self.device)  # 4 new shaping terms, energy_expenditure, energy_rew, rew
        bbuf = torch.empty(2, n, dtype=torch.uint8, device=self.device)  # all_passed, done
        fp, bp, row = fbuf.data_ptr(), bbuf.data_ptr(), 4 * n
        for k, name in enumerate(self.SHAPING):
            cur = getattr(sc.joint, name)
            cj.shaping_in[k] = cur.data_ptr()
            cj.shaping_out[k] = cur.data_ptr() if inplace else fp + k * row  # see NavigationHooks.reward
        cj.pos_rew, cj.rot_rew, cj.collision_rew = sc.pos_rew.data_ptr(), sc.rot_rew.data_ptr(), sc.collision_rew.data_ptr()
        cj.energy_expenditure, cj.energy_rew, cj.rew = fp + 4 * row, fp + 5 * row, fp + 6 * row
        cj.all_passed, cj.done, cj.done_flag = bp, bp + n, self.flag.data_ptr()
        _lib.check(_lib.lib().vmasb200_joint_passage_reward(C.byref(cj), n, _lib.raw_stream(torch, self.device)),
                   "vmasb200_joint_passage_reward")
        bools = bbuf.view(torch.bool)
        if not inplace:
            for k, name in enumerate(self.SHAPING):
                setattr(sc.joint, name, fbuf[k])
        sc.energy_expenditure, sc.energy_rew, sc.rew, sc.all_passed = fbuf[4], fbuf[5], fbuf[6], bools[0]
        self._done, self._done_key = bools[1], self.state_key()
        self.n_fused["reward"] += 1
        return sc.rew

    def obs_spec(self):
        sc, torch = self.scenario, self.torch
        if sc.observe_joint_angle and sc.joint_angle_obs_noise:
            return None  # the reference then adds noise to the landmark's rotation IN PLACE: its own code does that
        lo = float(sc.obs_noise)
        spec = []
        for a in self.world.agents:
            parts = [(OBS_COPY, a.state.pos, None), (OBS_COPY, a.state.vel, None), (OBS_SUB, a.state.pos, sc.goal.state.pos)]
            parts += [(OBS_SUB, a.state.pos, p.state.pos) for p in sc.passages if not p.collide]
            angles = [sc.goal.state.rot] + ([sc.joint.landmark.state.rot] if sc.observe_joint_angle else [])
            out = []
            for op, x, y in parts:  # one draw per operand of the cat, in the reference's order (joint_passage.py:531-541)
                noise = torch.zeros(*x.shape, device=self.device).uniform_(-lo, lo)
                out.append((op, x, y, 0.0, 0, noise))
            for ang in angles:      # angle_to_vector(angle) = cat([cos, sin], dim=1): one [n, 2] draw
                noise = torch.zeros(ang.shape[0], 2, device=self.device).uniform_(-lo, lo)
                out += [(OBS_COS, ang, None, 0.0, 0, noise[:, 0:1]), (OBS_SIN, ang, None, 0.0, 0, noise[:, 1:2])]
            spec.append(out)
        return spec


class FootballHooks(ScenarioHooks):
    """football.py:1118-1226 (reward), :1228-1462 (observation), :1464-1467 (done) for teams of policy agents
    (``ai_red_agents`` / ``ai_blue_agents`` teams keep their scripted controller and get no observation)."""

    @staticmethod
    def matches(scenario) -> bool:
        return (scenario_file(scenario) == "football" and
                all(hasattr(scenario, k) for k in ("ball", "blue_agents", "red_agents", "pitch_length", "goal_size",
                                                   "scoring_reward", "left_goal_pos", "right_goal_pos")))

    def __init__(self, env, binding):
        super().__init__(env, binding)
        self.cf = CFootball()
        sc = self.scenario
        self.goal_xy = [[float(v) for v in sc.right_goal_pos.tolist()], [float(v) for v in sc.left_goal_pos.tolist()]]
        self._rewards = None

    def extra_state(self):
        return tuple(a.state.force for a in self.world.agents if self.torch.is_tensor(a.state.force))

    def graph_inputs(self):
        b = self.scenario.ball
        names = ("pos_shaping_blue", "pos_shaping_red", "pos_shaping_agent_blue", "pos_shaping_agent_red")
        return [(b, k) for k in names if self.torch.is_tensor(getattr(b, k, None))]

    def _teams(self):
        sc = self.scenario
        return ((sc.blue_agents, "blue", bool(sc.ai_blue_agents)), (sc.red_agents, "red", bool(sc.ai_red_agents)))

    def _reward_usable(self) -> bool:
        sc = self.scenario
        ball = sc.ball
        if not (self._f32_vec(ball.state.pos, 2) and self._f32_vec(ball.state.vel, 2)):
            return False
        for team, name, ai in self._teams():
            if len(team) > FB_MAX_TEAM or not all(self._f32_vec(a.state.pos, 2) for a in team):
                return False
            if sc.dense_reward and not ai:
                if not team or not self._flat(getattr(ball, "pos_shaping_" + name, None)) or \
                        not self._flat(getattr(ball, "pos_shaping_agent_" + name, None)):
                    return False
        return True

    def reward(self, agent):
        sc, torch = self.scenario, self.torch
        first = agent is None or agent is self.world.agents[0]
        if not first:
            if self._rewards is not None:
                return self._rewards[0 if any(agent is a for a in sc.blue_agents) else 1]
            return self.orig["reward"](agent)
        self._rew_ok = agent is not None and self._reward_usable()
        if not self._rew_ok:
            self._not_covered("reward")
            self._rewards = None
            self._done = None
            return self.orig["reward"](agent)
        n, cf, ball, inplace = self.n, self.cf, sc.ball, self.b.static_io
        _vec(cf.ball_pos, ball.state.pos)
        _vec(cf.ball_vel, ball.state.vel)
        cf.pitch_length, cf.ball_size, cf.goal_size = float(sc.pitch_length), float(sc.ball_size), float(sc.goal_size)
        cf.scoring_reward = float(sc.scoring_reward)
        cf.pos_shaping_factor_ball_goal = float(sc.pos_shaping_factor_ball_goal)
        cf.pos_shaping_factor_agent_ball = float(sc.pos_shaping_factor_agent_ball)
        cf.distance_to_ball_trigger = float(sc.distance_to_ball_trigger)
        fbuf = torch.empty(18, n, device=self.device)  # per team 9 rows
        dbuf = torch.empty(n, dtype=torch.uint8, device=self.device)
        fp, row = fbuf.data_ptr(), 4 * n
        dense = []
        for t, (team, name, ai) in enumerate(self._teams()):
            m = cf.team[t]
            m.n_agents, m.dense = len(team), int(bool(sc.dense_reward) and not ai)
            dense.append(bool(m.dense))
            m.goal_pos[0], m.goal_pos[1] = self.goal_xy[t]
            base = fp + 9 * t * row
            if m.dense:
                for i, a in enumerate(team):
                    _vec(m.agent_pos[i], a.state.pos)
                sh, sha = getattr(ball, "pos_shaping_" + name), getattr(ball, "pos_shaping_agent_" + name)
                m.pos_shaping, m.pos_shaping_agent = sh.data_ptr(), sha.data_ptr()
                m.pos_shaping_out = sh.data_ptr() if inplace else base  # see NavigationHooks.reward
                m.pos_shaping_agent_out = sha.data_ptr() if inplace else base + row
            m.distance_to_goal, m.pos_rew, m.min_agent_dist_to_ball = base + 2 * row, base + 3 * row, base + 4 * row
            m.pos_rew_agent, m.dense_reward, m.sparse_reward, m.reward = base + 5 * row, base + 6 * row, base + 7 * row, base + 8 * row
        cf.done = dbuf.data_ptr()
        _lib.check(_lib.lib().vmasb200_football_reward(C.byref(cf), n, _lib.raw_stream(torch, self.device)),
                   "vmasb200_football_reward")
        for t, (team, name, ai) in enumerate(self._teams()):
            r = fbuf[9 * t: 9 * t + 9]
            if dense[t]:
                if not inplace:
                    setattr(ball, "pos_shaping_" + name, r[0])
                    setattr(ball, "pos_shaping_agent_" + name, r[1])
                setattr(ball, "distance_to_goal_" + name, r[2])
                setattr(ball, "pos_rew_" + name, r[3])
                setattr(sc, "min_agent_dist_to_ball_" + name, r[4])
                setattr(ball, "pos_rew_agent_" + name, r[5])
            setattr(sc, "_dense_reward_" + name, r[6] if dense[t] else 0)
            setattr(sc, "_sparse_reward_" + name, r[7])
        sc._done = dbuf.view(torch.bool)
        self._rewards = (fbuf[8], fbuf[17])
        self._done, self._done_key = sc._done, self.state_key()
        self.n_fused["reward"] += 1
        return self._rewards[0 if any(agent is a for a in sc.blue_agents) else 1]

    def done(self):
        sc = self.scenario
        if sc.ai_blue_agents and sc.ai_red_agents:
            return self.orig["done"]()
        return super().done()

    def observation(self, agent, *args, **kwargs):
        if args or kwargs:  # the scripted controllers call it with explicit tensors (football.py:1228-1246)
            return self.orig["observation"](agent, *args, **kwargs)
        return super().observation(agent)

    def _agent_index(self, agent) -> int:
        for i, a in enumerate(self._obs_agents):
            if a is agent:
                return i
        raise ValueError("agent has no fused observation")

    def obs_spec(self):
        sc, n = self.scenario, self.n
        if sc.dict_obs or sc.enable_shooting:
            return None
        ball = sc.ball
        self._obs_agents = [a for a in self.world.agents if a is not ball and a.action_script is None]
        if not self._obs_agents or self._obs_agents[0] is not self.world.agents[0]:
            return None  # the cache is refreshed when the first agent of the world asks
        spec = []
        for a in self._obs_agents:
            blue = any(a is x for x in sc.blue_agents)
            mine, other = (sc.blue_agents, sc.red_agents) if blue else (sc.red_agents, sc.blue_agents)
            goal = (sc.right_goal_pos if blue else sc.left_goal_pos).unsqueeze(0).expand(n, 2)
            f = 0 if blue else OBS_NEG0  # a red agent sees every x mirrored (football.py:1376-1397)
            st, bs = a.state, ball.state
            parts = [(OBS_COPY, st.force, None, 0.0, f), (OBS_SUB, st.pos, bs.pos, 0.0, f), (OBS_SUB, st.vel, bs.vel, 0.0, f),
                     (OBS_SUB, bs.pos, goal, 0.0, f), (OBS_COPY, bs.vel, None, 0.0, f), (OBS_COPY, bs.force, None, 0.0, f),
                     (OBS_SUB, st.pos, goal, 0.0, f), (OBS_COPY, st.vel, None, 0.0, f)]
            groups = ([other] if sc.observe_adversaries and len(other) else []) + \
                     ([[x for x in mine if x is not a]] if sc.observe_teammates else [])
            for group in groups:
                for o in group:
                    parts += [(OBS_SUB, st.pos, o.state.pos, 0.0, f), (OBS_SUB, st.vel, o.state.vel, 0.0, f),
                              (OBS_COPY, o.state.vel, None, 0.0, f), (OBS_COPY, o.state.force, None, 0.0, f)]
            spec.append(parts)
        return spec


REGISTRY = (NavigationHooks, TransportHooks, SimpleSpreadHooks, JointPassageHooks, FootballHooks)


def install(env, binding) -> Optional[ScenarioHooks]:
    """Fused hooks for ``env.scenario`` when its scenario has them, else None (the scenario keeps its own)."""
    sc = getattr(env, "scenario", None)
    if sc is None:
        return None
    for cls in REGISTRY:
        if cls.matches(sc):
            return cls(env, binding)
    return None
