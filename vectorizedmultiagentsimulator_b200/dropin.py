"""Drop-in for the two bound methods that make up the reference's physics hot path:

* ``vmas.simulator.core.World.step``       (core.py:1971)  -> one fused CUDA launch (``vmasb200_step``)
* ``vmas.simulator.core.World.cast_rays``  (core.py:1661)  -> ``vmasb200_cast_rays``; ``Lidar.measure``
  (sensors.py:100) additionally returns the measurement the step's LIDAR epilogue already produced when nothing
  touched the state in between.
* ``World.get_distance / is_overlapping / get_distance_from_point`` (core.py:1787-1966, the queries reward code makes
  between steps) -> one launch each (``vmasb200_get_distance`` ...), reading the entities' state tensors in place.

``enable(world_or_env)`` patches ONE world instance (and its Lidar sensors); ``make_env``, ``Environment.step``, the
scenario hooks and the returned obs / reward / done / info tensors stay untouched.  Nothing from ``vmas`` is imported:
the world is duck-typed, so the same code serves any fork that keeps the reference's attribute names.

State ownership.  The kernel works on structure-of-arrays buffers ``[row][ld]`` (include/vmasb200.h).  After a step
every ``entity.state.pos / vel / rot / ang_vel`` of a dynamic entity is bound to a strided VIEW of a freshly allocated
output buffer (the reference also rebinds fresh tensors every step, core.py:2866-2907, so stale references a scenario
kept stay valid), and agents whose action is clamped get ``state.force / torque`` bound to views of the clamped
write-back (core.py:2019-2038).  Before a step each field is compared with the view it should be (data_ptr + strides);
only fields a scenario or ``reset`` rebound are copied in.  In-place writes through the views land in the buffer directly.

What is NOT done here, on purpose: no CPU fallback (a world that is not on a CUDA device is an error), and no autograd
(if grad mode is on and a state / force tensor requires grad, the saved original ``World.step`` runs: the reference's
torch path remains the gradient path).  ``world.forces_dict / torques_dict`` are not materialised (no scenario reads
them).
"""
from __future__ import annotations

import ctypes as C
import types
from typing import Dict, List, Optional

import numpy as np

from . import _lib, hooks as _hooks, queries
from .plan import StepPlan
from .worlddesc import CStepIO, UnsupportedWorld, WorldDesc  # noqa: F401

_FIELDS = ("pos", "vel", "rot", "ang_vel")


def _same_view(t, v) -> bool:
    return t is v or (t.data_ptr() == v.data_ptr() and t.shape == v.shape and t.stride() == v.stride()
                      and t.dtype == v.dtype)


class CAction(C.Structure):  # vmasb200_action (include/vmasb200.h)
    _fields_ = [("action", C.c_void_p), ("stride_env", C.c_int64), ("stride_k", C.c_int64), ("action_size", C.c_int32),
                ("comm_col0", C.c_int32), ("n_comm", C.c_int32), ("clamp", C.c_int32), ("u_range", C.c_void_p),
                ("u_multiplier", C.c_void_p), ("u_randn", C.c_void_p), ("c_randn", C.c_void_p), ("u_noise", C.c_double),
                ("c_noise", C.c_double), ("u_out", C.c_void_p), ("c_out", C.c_void_p), ("flags", C.c_void_p)]


class WorldBinding:
    """Plan + device buffers of one enabled World."""

    def __init__(self, world, exact: bool = False):
        import torch

        self.torch = torch
        self.world = world
        self.exact = exact
        dev = torch.device(world.device)
        if dev.type != "cuda":
            raise _lib.VmasB200Error(
                f"vmas-b200: world.device is {dev}; the CUDA step has no CPU fallback (create the env with device='cuda')")
        self.device = dev
        self.n = int(world.batch_dim)
        self.ld = StepPlan.ld_for(self.n)
        self.orig_step = world.step
        self.orig_cast_rays = world.cast_rays
        self.dirty = True
        self._act_consts: Dict[int, tuple] = {}
        self._act_flags = None
        self._act_pending = False
        self.static_io = False  # persistent buffers stepped in place (enable(..., static_io=True))
        self.defer_action_check = False  # graph.py: no host read of the flag word while a CUDA graph is captured
        self.hooks = None
        self.desc_json = None
        self.plan: Optional[StepPlan] = None
        self.lidar_valid = False
        self.n_steps = 0
        self.n_copies = 0
        self.n_fallback = {"set_action": 0, "cast_rays": 0}  # calls that took the reference's torch code instead (see `strict`)
        self.n_fused_actions = 0
        self.strict = False  # True: such a call raises instead
        self.fused_reset = False
        self.n_plan_builds = 0
        self.record_masks = False  # True: every step also leaves World.collides' verdicts per (substep, pair) in self.masks
        self.masks = None

    def _fallback(self, kind: str, why: str):
        """A secondary call (never World.step itself) that this library does not cover takes the reference's own torch
        code.  Counted in n_fallback; a one-time warning says so; with strict=True it is an error instead."""
        if self.strict:
            raise _lib.VmasB200Error(f"vmas-b200 (strict): {kind} is not covered here: {why}")
        if self.n_fallback.get(kind, 0) == 0:
            import warnings

            warnings.warn(f"vmas-b200: {kind} takes the reference's torch code: {why} (binding.strict = True makes this an error)",
                          RuntimeWarning, stacklevel=3)
        self.n_fallback[kind] = self.n_fallback.get(kind, 0) + 1

    # ------------------------------------------------------------------ plan (re)build
    def invalidate(self):
        """The static description of the world changed outside a reset (a mass, max_speed, a force range, gravity, a collide flag
        or collision filter, a joint, a LIDAR's targets): re-extract it before the next step.  The reference re-reads these
        attributes every step; this binding reads them when the plan is built - at enable(), after every World.reset /
        reset_at, and after invalidate().  Cheap when nothing changed (one description compare)."""
        self.dirty = True

    def refresh(self):
        """Re-extract the static description (entity table, collision filters, joints, sensors); rebuild the plan and
        the buffers only when it changed (collision filters may depend on state set at reset: joint_passage.py:615-622)."""
        torch = self.torch
        desc = WorldDesc.from_world(self.world)
        js = desc.to_json()
        self.dirty = False
        if js == self.desc_json:
            return
        self.desc_json = js
        self.desc = desc
        self.n_plan_builds += 1
        self.plan = pl = StepPlan(desc)
        n, ld, dev = self.n, self.ld, self.device
        self.ents = list(self.world.entities)
        self.agents = [self.ents[i] for i in pl.agents]
        self.dyn = torch.zeros(pl.dyn_rows, ld, device=dev)
        self.stat = torch.zeros(max(pl.stat_rows, 1), ld, device=dev)
        self.act = None
        self.lid = None
        self._act_views = []
        self.dyn_ents = [(e, int(pl.dyn_slot[i])) for i, e in enumerate(self.ents) if pl.dyn_slot[i] >= 0]
        self.stat_ents = [(e, int(pl.stat_slot[i])) for i, e in enumerate(self.ents) if pl.stat_slot[i] >= 0]
        self.act_agents = [(e, int(pl.act_slot[i])) for i, e in enumerate(self.ents) if pl.act_slot[i] >= 0]
        self.joint_rows = []
        if len(pl.joint_slot) and (pl.joint_slot >= 0).any():
            order = _joints_in_order(self.world, self.ents)
            self.joint_rows = [(order[j], int(o)) for j, o in enumerate(pl.joint_slot) if o >= 0]
        self.sensors = {}
        for k, ld_ in enumerate(desc.lidars):
            sensor = self.ents[ld_.entity].sensors[ld_.agent_sensor_index]
            self.sensors[id(sensor)] = k
        self._bind_static()
        self._dyn_views = self._views_of(self.dyn)
        self.lidar_valid = False
        # reusable C structures
        A = pl.A
        self._io = io = CStepIO()
        self._fp = (C.c_void_p * max(A, 1))()
        self._tp = (C.c_void_p * max(A, 1))()
        self._fs = (C.c_int64 * max(A, 1))()
        self._ts = (C.c_int64 * max(A, 1))()
        io.force_ptrs, io.force_strides = C.cast(self._fp, C.POINTER(C.c_void_p)), C.cast(self._fs, C.POINTER(C.c_int64))
        io.torque_ptrs, io.torque_strides = C.cast(self._tp, C.POINTER(C.c_void_p)), C.cast(self._ts, C.POINTER(C.c_int64))
        io.n_envs, io.ld, io.exact = n, ld, int(self.exact)

    def _views_of(self, dyn):
        # [n, k] views with strides (1, ld) of rows o .. o + k - 1 (= dyn[o:o + k, :n].t(), one call instead of three)
        n, ld = self.n, self.ld
        out = []
        for e, o in self.dyn_ents:
            out.append((e, (dyn.as_strided((n, 2), (1, ld), o * ld), dyn.as_strided((n, 2), (1, ld), (o + 2) * ld),
                            dyn.as_strided((n, 1), (1, ld), (o + 4) * ld), dyn.as_strided((n, 1), (1, ld), (o + 5) * ld))))
        return out

    def _bind_static(self):
        n = self.n
        self._stat_views = [(e, (self.stat[o:o + 2, :n].t(), self.stat[o + 2:o + 3, :n].t())) for e, o in self.stat_ents]

    # ------------------------------------------------------------------ state hand-over
    def sync_in(self):
        """Copies into the buffers every state field that is not (any more) the view the last step bound."""
        for e, views in self._dyn_views:
            st = e.state
            for name, v in zip(_FIELDS, views):
                t = getattr(st, name)
                if not _same_view(t, v):
                    v.copy_(t)
                    setattr(st, "_" + name, v)
                    self.n_copies += 1
                    self.lidar_valid = False
        for e, (vp, vr) in self._stat_views:
            st = e.state
            if not _same_view(st.pos, vp):
                vp.copy_(st.pos)
                st._pos = vp
                self.n_copies += 1
                self.lidar_valid = False
            if not _same_view(st.rot, vr):
                vr.copy_(st.rot)
                st._rot = vr
                self.n_copies += 1
                self.lidar_valid = False
        for joint, row in self.joint_rows:
            fr = joint.fixed_rotation
            self.stat[row, :self.n].copy_(fr.reshape(-1).expand(self.n) if hasattr(fr, "reshape") else
                                          self.torch.full((self.n,), float(fr), device=self.device))

    def _needs_grad(self) -> bool:
        if not self.torch.is_grad_enabled():
            return False
        for e in self.ents:
            for name in _FIELDS:
                t = getattr(e.state, name)
                if t is not None and t.requires_grad:
                    return True
        for a in self.agents:
            for t in (a.state.force, a.state.torque):
                if t is not None and t.requires_grad:
                    return True
        return False

    # ------------------------------------------------------------------ World.step
    def step(self):
        world, torch = self.world, self.torch
        if not self.defer_action_check:
            self.check_actions()
        if self.dirty or len(world.entities) != len(getattr(self, "ents", ())):
            self.refresh()
        if self._needs_grad():
            self.lidar_valid = False
            return self.orig_step()  # the reference's torch path is the gradient path
        pl, n, ld = self.plan, self.n, self.ld
        world.entity_index_map = {e: i for i, e in enumerate(world.entities)}  # core.py:1972
        self.sync_in()
        io = self._io
        keep = []
        for a, ag in enumerate(self.agents):
            f, t = ag.state.force, ag.state.torque
            if f is not None:
                if f.dtype != torch.float32 or f.stride(1) != 1:
                    f = f.to(torch.float32).contiguous()
                keep.append(f)
                self._fp[a], self._fs[a] = f.data_ptr(), f.stride(0)
            else:
                self._fp[a], self._fs[a] = None, 0
            if t is not None:
                if t.dtype != torch.float32:
                    t = t.to(torch.float32)
                keep.append(t)
                self._tp[a], self._ts[a] = t.data_ptr(), t.stride(0)
            else:
                self._tp[a], self._ts[a] = None, 0
        if self.static_io:
            # persistent buffers, stepped in place (the kernel stages a tile completely before it stores it): state views
            # keep their addresses and nothing is allocated
            dyn_out = self.dyn
            act = self.act if self.act is not None else (torch.empty(pl.act_rows, ld, device=self.device) if pl.act_rows else None)
            lid = self.lid if self.lid is not None else (torch.empty(pl.lidar_rows, ld, device=self.device) if pl.lidar_rows else None)
        else:
            # fresh output buffers every step: the reference rebinds new tensors too (core.py:2866-2907)
            dyn_out = torch.empty_like(self.dyn)
            act = torch.empty(pl.act_rows, ld, device=self.device) if pl.act_rows else None
            lid = torch.empty(pl.lidar_rows, ld, device=self.device) if pl.lidar_rows else None
        io.dyn_in, io.dyn_out = self.dyn.data_ptr(), dyn_out.data_ptr()
        io.stat = self.stat.data_ptr() if pl.stat_rows else None
        io.act_out = act.data_ptr() if act is not None else None
        io.lidar_out = lid.data_ptr() if lid is not None else None
        if self.record_masks and pl.pair_words:
            if self.masks is None or self.masks.shape != (pl.S, pl.pair_words):
                self.masks = torch.zeros(pl.S, pl.pair_words, dtype=torch.int32, device=self.device)
            io.pair_active_out = self.masks.data_ptr()
        else:
            io.pair_active_out = None
        io.stream = _lib.raw_stream(torch, self.device)
        pl.step_io(io)
        # bind the results
        if dyn_out is not self.dyn:
            self.dyn = dyn_out
            self._dyn_views = self._views_of(dyn_out)
        for e, views in self._dyn_views:  # (sync_in left every field bound to its view: nothing to do when in place)
            st = e.state
            if st._pos is not views[0]:
                st._pos, st._vel, st._rot, st._ang_vel = views
        if act is not None:
            if act is not self.act:
                self.act = act
                self._act_views = [(ag, act.as_strided((n, 2), (1, ld), o * ld), act.as_strided((n, 1), (1, ld), (o + 2) * ld))
                                   for ag, o in self.act_agents]
            for ag, vf, vt in self._act_views:
                ag.state._force, ag.state._torque = vf, vt
        if lid is not None:
            self.lid = lid
            self.lidar_valid = True
            self._versions = (dyn_out._version, self.stat._version)
        if world._dim_c > 0:  # core.py:2909-2912
            for ag in world._agents:
                if not ag.silent:
                    ag.state.c = ag.action.c
        self.n_steps += 1

    # ------------------------------------------------------------------ World.cast_rays / Lidar.measure
    def _state_untouched(self) -> bool:
        if not self.lidar_valid or self._versions != (self.dyn._version, self.stat._version):
            return False
        for e, views in self._dyn_views:
            st = e.state
            if st._pos is not views[0] or st._rot is not views[2]:
                return False
        for e, (vp, vr) in self._stat_views:
            if e.state._pos is not vp or e.state._rot is not vr:
                return False
        return True

    def measure(self, sensor, orig_measure, vectorized: bool = True):
        k = self.sensors.get(id(sensor))
        if vectorized and k is not None and not self.dirty and self._state_untouched():
            o, r = int(self.plan.lidar_slot[k]), len(self.desc.lidars[k].angles)
            m = self.lid.as_strided((self.n, r), (1, self.ld), o * self.ld)
            sensor._last_measurement = m
            return m
        return orig_measure(vectorized)

    def cast_rays(self, entity, angles, max_range: float, entity_filter=lambda _: False):
        torch = self.torch
        if self.dirty or self.plan is None:
            self.refresh()
        if self.torch.is_grad_enabled() and angles.requires_grad:
            return self.orig_cast_rays(entity, angles, max_range, entity_filter)
        ents = self.ents
        index = {id(e): i for i, e in enumerate(ents)}
        targets = []
        for e in ents:
            if e is entity or not entity_filter(e):
                continue
            assert e.collides(entity) and entity.collides(e), "Rays are only casted among collidables"  # core.py:1680
            targets.append(index[id(e)])
        pl = self.plan
        me = index[id(entity)]
        if any(pl.dyn_slot[i] < 0 and pl.stat_slot[i] < 0 for i in targets + [me]) or len(targets) > 64:
            self._fallback("cast_rays", "more than 64 targets, or an entity the step never reads (no rows in the plan's buffers)")
            return self.orig_cast_rays(entity, angles, max_range, entity_filter)
        self.sync_in()
        ang = angles if (angles.dtype == torch.float32 and angles.stride(1) == 1) else angles.to(torch.float32).contiguous()
        out = torch.empty(ang.shape[0], ang.shape[1], device=self.device)
        pl.cast_rays(self.dyn, self.stat, self.ld, me, ang, float(max_range), np.asarray(targets, np.int32), out,
                     _lib.raw_stream(torch, self.device))
        return out


    # ------------------------------------------------------------------ Environment._set_action
    def set_action(self, env, action, agent):
        """Continuous physical actions without communication / noise: one launch (vmasb200_set_action) instead of the
        reference's ~10 ops and 3 host synchronisations per agent (environment.py:615-668).  The reference's assertions
        (NaN action, action out of range) are kept: the launches OR into one flag word that ``step`` reads - a single
        synchronisation per Environment.step - before it touches the world.  Everything else takes the original."""
        torch = self.torch
        act = getattr(agent, "action", None)
        noise = getattr(act, "u_noise", 0)
        noisy = any(v > 0 for v in noise) if isinstance(noise, (list, tuple)) else noise > 0
        comm = self.world.dim_c > 0 and not agent.silent
        base = (getattr(env, "continuous_actions", False) and torch.is_tensor(action) and action.is_cuda
                and action.dtype == torch.float32 and action.dim() == 2 and action.shape[0] == self.n
                and act is not None and 1 <= agent.action_size <= 8
                and not (getattr(env, "grad_enabled", False) and action.requires_grad))
        if base and (comm or noisy) and not isinstance(noise, (list, tuple)):
            if self._set_action_ex(env, action, agent, act, comm):
                return None
        ok = base and not comm and not noisy and action.shape[1] == agent.action_size
        if not ok:
            if self._set_action_discrete(env, action, agent, act, noisy):
                return None
            self._fallback("set_action", f"agent {getattr(agent, 'name', '?')}: not a CUDA fp32 / int action tensor of the agent's "
                                         "size, autograd in play, per-component noise, or discrete actions with communication")
            return self.orig_set_action(action, agent)
        key = (id(agent), repr(act.u_range), repr(act.u_multiplier))
        c = self._act_consts.get(id(agent))
        if c is None or c[0] != key:
            k = agent.action_size
            rng = (C.c_float * k)(*[float(v) for v in act.u_range_tensor.tolist()])
            mul = (C.c_float * k)(*[float(v) for v in act.u_multiplier_tensor.tolist()])
            c = self._act_consts[id(agent)] = (key, rng, mul)
        if self._act_flags is None:
            self._act_flags = torch.zeros(1, dtype=torch.int32, device=self.device)
        u = torch.empty(self.n, agent.action_size, device=self.device)
        _lib.check(_lib.lib().vmasb200_set_action(
            action.data_ptr(), action.stride(0), action.stride(1), agent.action_size, c[1], c[2],
            int(bool(getattr(env, "clamp_action", False))), u.data_ptr(), self._act_flags.data_ptr(), self.n,
            _lib.raw_stream(torch, self.device)), "vmasb200_set_action")
        act.u = u
        self._act_pending = True

    def _set_action_ex(self, env, action, agent, act, comm):
        """Continuous actions with communication columns and / or action noise (environment.py:633-645, 711-754): one
        launch (vmasb200_set_action_ex).  The noise is drawn here with the reference's own torch.randn calls, in its
        order (u first, then c), so the generator advances exactly as in the reference."""
        torch = self.torch
        k, dim_c = agent.action_size, (self.world.dim_c if comm else 0)
        clamp = bool(getattr(env, "clamp_action", False))
        col0 = k if clamp else self.world.dim_p
        n_comm = (dim_c if clamp else action.shape[1] - col0) if comm else 0
        if action.shape[1] != k + dim_c or n_comm < 0 or n_comm > 64:
            return False
        key = (id(agent), repr(act.u_range), repr(act.u_multiplier))
        c = self._act_consts.get(id(agent))
        if c is None or c[0] != key:
            rng = (C.c_float * k)(*[float(v) for v in act.u_range_tensor.tolist()])
            mul = (C.c_float * k)(*[float(v) for v in act.u_multiplier_tensor.tolist()])
            c = self._act_consts[id(agent)] = (key, rng, mul)
        if self._act_flags is None:
            self._act_flags = torch.zeros(1, dtype=torch.int32, device=self.device)
        u_noise, c_noise = float(getattr(act, "u_noise", 0) or 0), float(getattr(agent, "c_noise", 0) or 0)
        u = torch.empty(self.n, k, device=self.device)
        u_randn = torch.randn(self.n, k, device=self.device, dtype=torch.float32) if u_noise > 0 else None
        cc = torch.empty(self.n, n_comm, device=self.device) if n_comm else None
        c_randn = torch.randn(self.n, n_comm, device=self.device, dtype=torch.float32) if (n_comm and c_noise > 0) else None
        a = CAction()
        a.action, a.stride_env, a.stride_k = action.data_ptr(), action.stride(0), action.stride(1)
        a.action_size, a.comm_col0, a.n_comm, a.clamp = k, col0, n_comm, int(clamp)
        a.u_range, a.u_multiplier = C.cast(c[1], C.c_void_p), C.cast(c[2], C.c_void_p)
        a.u_randn = u_randn.data_ptr() if u_randn is not None else None
        a.c_randn = c_randn.data_ptr() if c_randn is not None else None
        a.u_noise, a.c_noise = u_noise, c_noise
        a.u_out, a.c_out, a.flags = u.data_ptr(), (cc.data_ptr() if cc is not None else None), self._act_flags.data_ptr()
        _lib.check(_lib.lib().vmasb200_set_action_ex(C.byref(a), self.n, _lib.raw_stream(torch, self.device)), "vmasb200_set_action_ex")
        act.u = u
        if cc is not None:
            act.c = cc
        self._act_pending = True
        self.n_fused_actions += 1
        return True

    def _set_action_discrete(self, env, action, agent, act, noisy):
        """Discrete physical actions (flat or multi-discrete, int64 / int32) without communication / noise
        (environment.py:660-710): one launch (vmasb200_set_action_discrete) instead of ~25 ops and a host
        synchronisation per action component.  False = not this case, the caller takes the original."""
        torch = self.torch
        if getattr(env, "continuous_actions", True) or act is None or noisy or not torch.is_tensor(action):
            return False
        multi = bool(getattr(env, "multidiscrete_actions", False))
        nvec = list(getattr(agent, "discrete_action_nvec", []) or [])
        k = agent.action_size
        ok = (action.is_cuda and action.dtype in (torch.int64, torch.int32) and action.dim() == 2 and action.shape[0] == self.n
              and (self.world.dim_c == 0 or agent.silent) and 1 <= k <= 8 and len(nvec) == k
              and action.shape[1] == (k if multi else 1) and all(int(v) > 1 for v in nvec))
        if not ok:
            return False
        key = (id(agent), "d", repr(act.u_range), repr(act.u_multiplier), tuple(int(v) for v in nvec))
        c = self._act_consts.get((id(agent), "d"))
        if c is None or c[0] != key:
            rng = (C.c_float * k)(*[float(v) for v in act.u_range_tensor.tolist()])
            mul = (C.c_float * k)(*[float(v) for v in act.u_multiplier_tensor.tolist()])
            nv = (C.c_int32 * k)(*[int(v) for v in nvec])
            c = self._act_consts[(id(agent), "d")] = (key, rng, mul, nv)
        if self._act_flags is None:
            self._act_flags = torch.zeros(1, dtype=torch.int32, device=self.device)
        u = torch.empty(self.n, k, device=self.device)
        L = _lib.lib()
        _lib.check(L.vmasb200_set_action_discrete(
            action.data_ptr(), 0 if action.dtype == torch.int64 else 1, action.stride(0), action.stride(1), k, int(multi),
            c[3], c[1], c[2], u.data_ptr(), self._act_flags.data_ptr(), self.n, _lib.raw_stream(torch, self.device)),
            "vmasb200_set_action_discrete")
        act.u = u
        self._act_pending = True
        return True

    def check_actions(self):
        """The reference's assertions on the actions set since the last check (environment.py:622, 647-649)."""
        if not self._act_pending:
            return
        self._act_pending = False
        bad = int(self._act_flags.item())
        if bad:
            self._act_flags.zero_()
            assert not bad & 1, "an action holds NaN"
            assert not bad & 2, "Physical actions of an agent are out of its range"
            assert not bad & 4, "Discrete physical actions are out of bounds"
            assert not bad & 8, "Comm actions are out of range [0,1]"

    # ------------------------------------------------------------------ distance queries of reward code
    def _poses(self, *entities):
        """CPose per entity from its live state tensors, or None when the reference's torch path must answer
        (autograd in play, or a shape the library does not know: the reference raises for those itself)."""
        torch = self.torch
        out, keep = [], []
        for e in entities:
            sh = queries.shape_of(e)
            pos, rot = e.state.pos, e.state.rot
            if sh is None or pos is None or not pos.is_cuda:
                return None
            if torch.is_grad_enabled() and (pos.requires_grad or (rot is not None and rot.requires_grad)):
                return None
            p, k = queries.make_pose(*sh, pos, rot, torch)
            out.append(p)
            keep.append(k)
        return out, keep

    def get_distance(self, entity_a, entity_b, env_index=None):
        ps = self._poses(entity_a, entity_b) if env_index is None else None  # env_index: the reference's own quirks
        if ps is None:
            return self.orig_get_distance(entity_a, entity_b, env_index)
        (pa, pb), keep = ps
        return queries.get_distance(pa, pb, self.n, entity_a.state.pos, self.torch)

    def is_overlapping(self, entity_a, entity_b, env_index=None):
        ps = self._poses(entity_a, entity_b) if env_index is None else None
        if ps is None:
            return self.orig_is_overlapping(entity_a, entity_b, env_index)
        (pa, pb), keep = ps
        return queries.is_overlapping(pa, pb, self.n, entity_a.state.pos, self.torch)

    def get_distance_from_point(self, entity, test_point_pos, env_index=None):
        torch = self.torch
        ok = (env_index is None and torch.is_tensor(test_point_pos) and test_point_pos.is_cuda
              and tuple(test_point_pos.shape) == (self.n, 2)
              and not (torch.is_grad_enabled() and test_point_pos.requires_grad))
        ps = self._poses(entity) if ok else None
        if ps is None:
            return self.orig_get_distance_from_point(entity, test_point_pos, env_index)
        (pa,), keep = ps
        return queries.get_distance_from_point(pa, test_point_pos, self.n, torch)


# ---------------------------------------------------------------------------------------------- reset path
_SPAWN_PATCH = {"count": 0, "orig": None, "cls": None}


def _fused_find_random_pos(occupied_positions, env_index, world, min_dist_between_entities, x_bounds, y_bounds,
                           disable_warn: bool = False):
    """ScenarioUtils.find_random_pos_for_entity (utils.py:274-319) for a world with the fused reset enabled: the same
    torch `uniform_` draws in the same order (so the generator advances exactly as in the reference and the placement is
    bit-identical), the cdist / compare / any / masked assignment of an iteration as ONE launch (vmasb200_spawn_select),
    and the loop's exit test read from a flag word (the reference synchronises on `overlaps.any()` there too)."""
    b = getattr(world, "_vmasb200", None)
    if b is None or not getattr(b, "fused_reset", False) or not occupied_positions.is_cuda:
        return _SPAWN_PATCH["orig"](occupied_positions, env_index, world, min_dist_between_entities, x_bounds, y_bounds, disable_warn)
    torch = b.torch
    batch = world.batch_dim if env_index is None else 1
    occ = occupied_positions if occupied_positions.is_contiguous() else occupied_positions.contiguous()
    K = int(occ.shape[1])
    flag = b._spawn_flag
    if flag is None:
        flag = b._spawn_flag = torch.zeros(1, dtype=torch.int32, device=b.device)
    L, stream = _lib.lib(), _lib.raw_stream(torch, b.device)
    pos, tries = None, 0
    while True:
        proposed = torch.cat([torch.empty((batch, 1, 1), device=world.device, dtype=torch.float32).uniform_(*x_bounds),
                              torch.empty((batch, 1, 1), device=world.device, dtype=torch.float32).uniform_(*y_bounds)], dim=2)
        if pos is None:
            pos = proposed
        if K == 0:
            break
        flag.zero_()
        _lib.check(L.vmasb200_spawn_select(pos.data_ptr(), proposed.data_ptr(), occ.data_ptr(), K, float(min_dist_between_entities),
                                           flag.data_ptr(), batch, stream), "vmasb200_spawn_select")
        b.n_spawn_launches += 1
        if not int(flag.item()):
            break
        tries += 1
        if tries > 50_000 and not disable_warn:
            import warnings

            warnings.warn("It is taking many iterations to spawn the entity, make sure the bounds or "
                          "the min_dist_between_entities are not too tight to fit all entities."
                          "You can disable this warning by setting disable_warn=True")
    return pos


def _install_fused_reset(world, b):
    """World.reset / Entity.set_pos|vel|rot|ang_vel / ScenarioUtils.find_random_pos_for_entity for ONE world: state is
    zeroed and written IN PLACE through the bound views of the SoA buffers (the reference rebinds fresh tensors,
    core.py:285-295, 742-755, which the next step would have to re-pack field by field)."""
    torch = b.torch
    b.fused_reset, b._spawn_flag, b.n_spawn_launches, b.n_reset_launches = True, None, 0, 0
    b.sync_in()  # pack the current state now: from here on every state field is a view of the SoA buffers

    def reset(self, env_index=None):
        b.dirty = True  # collision filters / joint rotations may depend on what the scenario sets next
        if not (env_index is None or isinstance(env_index, int)) or b.plan is None:
            return b._orig_reset(env_index)
        L, stream, pl = _lib.lib(), _lib.raw_stream(torch, b.device), b.plan
        idx = -1 if env_index is None else int(env_index)
        in_dyn = {id(e) for e, views in b._dyn_views if all(_same_view(getattr(e.state, f), v) for f, v in zip(_FIELDS, views))}
        in_stat = {id(e) for e, (vp, vr) in b._stat_views if _same_view(e.state.pos, vp) and _same_view(e.state.rot, vr)}
        if len(in_dyn) == len(b._dyn_views) and pl.dyn_rows:
            _lib.check(L.vmasb200_reset_rows(b.dyn.data_ptr(), 0, pl.dyn_rows, b.ld, b.n, idx, None, stream), "vmasb200_reset_rows")
            b.n_reset_launches += 1
        else:
            in_dyn = set()
        if len(in_stat) == len(b._stat_views) and b._stat_views:
            rows = 3 * len(b._stat_views)  # entity rows only: the per-env joint rotations behind them are not state
            _lib.check(L.vmasb200_reset_rows(b.stat.data_ptr(), 0, rows, b.ld, b.n, idx, None, stream), "vmasb200_reset_rows")
            b.n_reset_launches += 1
        else:
            in_stat = set()
        b.lidar_valid = False
        for e in self.entities:
            # everything else an entity resets (Agent._reset: action, dynamics, AgentState's c / force / torque,
            # core.py:372-387, 1046-1049; fields outside the buffers) stays the reference's own code: the fields already
            # zeroed in place are hidden from EntityState._reset (it skips None, core.py:285-288) and put back
            st = e.state
            hide = _FIELDS if id(e) in in_dyn else (("pos", "rot") if id(e) in in_stat else ())
            saved = [(k, getattr(st, "_" + k)) for k in hide]
            for k, _ in saved:
                setattr(st, "_" + k, None)
            try:
                e._reset(env_index)
            finally:
                for k, t in saved:
                    setattr(st, "_" + k, t)

    def set_prop(self, prop, state, new, batch_index):
        assert self.batch_dim is not None, f"Tried to set property of {self.name} without adding it to the world"
        self._check_batch_index(batch_index)
        new = new.to(self.device)
        cur = prop.fget(state)
        if batch_index is None:
            src = new if (len(new.shape) > 1 and new.shape[0] == self.batch_dim) else new.repeat(self.batch_dim, 1)
            if (cur is not None and cur.shape == src.shape and cur.dtype == src.dtype
                    and not (torch.is_grad_enabled() and (src.requires_grad or cur.requires_grad))):
                cur.copy_(src)  # in place: the field stays the view of the SoA buffer it was
            else:
                prop.fset(state, src)
        else:
            cur[batch_index] = new
        self.notify_observers()

    world.reset = types.MethodType(reset, world)
    b._set_prop_patched = []
    for e in world.entities:
        if hasattr(e, "_set_state_property"):
            e._set_state_property = types.MethodType(set_prop, e)
            b._set_prop_patched.append(e)
    try:
        import importlib

        cls = importlib.import_module(type(world).__module__.rsplit(".", 1)[0] + ".utils").ScenarioUtils
    except Exception:  # a World that is not the reference's: nothing to patch
        cls = None
    if cls is not None:
        if _SPAWN_PATCH["count"] == 0:
            _SPAWN_PATCH["orig"], _SPAWN_PATCH["cls"] = cls.find_random_pos_for_entity, cls
            cls.find_random_pos_for_entity = staticmethod(_fused_find_random_pos)
        _SPAWN_PATCH["count"] += 1
        b._spawn_patched = True


def _remove_fused_reset(world, b):
    for e in getattr(b, "_set_prop_patched", []):
        e.__dict__.pop("_set_state_property", None)
    if getattr(b, "_spawn_patched", False):
        b._spawn_patched = False
        _SPAWN_PATCH["count"] -= 1
        if _SPAWN_PATCH["count"] == 0 and _SPAWN_PATCH["cls"] is not None:
            _SPAWN_PATCH["cls"].find_random_pos_for_entity = staticmethod(_SPAWN_PATCH["orig"])
            _SPAWN_PATCH["orig"] = _SPAWN_PATCH["cls"] = None
    b.fused_reset = False


def _joints_in_order(world, ents):
    """JointConstraints in the order _apply_vectorized_enviornment_force meets them (core.py:2111-2121)."""
    out = []
    wj = getattr(world, "_joints", {})
    for a in range(len(ents)):
        for b in range(a + 1, len(ents)):
            j = wj.get(frozenset({ents[a].name, ents[b].name}), None)
            if j is not None:
                out.append(j)
    return out


def enable(world_or_env, exact: bool = False, static_io: bool = False, hooks: bool = False, strict: bool = False,
           fused_reset: bool = False) -> WorldBinding:
    """Routes ``world.step`` / ``world.cast_rays`` / its Lidars' ``measure`` / the distance queries of ONE world through
    libvmasb200.  ``exact=True`` takes the per-substep exact path instead of the verified speculative single launch.
    ``static_io=True`` (opt-in): the step works IN PLACE on persistent buffers instead of binding fresh output tensors
    every step - no allocations and no new views per step, state tensors keep their addresses (what CUDA-graph capture
    of user code around the step needs) - at the price that a state tensor a caller kept from an earlier step changes
    with the world instead of staying a snapshot, unlike the reference (core.py:2866-2907 rebinds fresh tensors).
    ``hooks=True`` (opt-in, needs an Environment): the scenario's reward / observation / done take the fused launches of
    ``hooks.py`` when the scenario has them (navigation, transport); ``binding.hooks`` is None otherwise.
    ``strict=True``: a secondary call this library does not cover (an action layout, a cast_rays over entities outside
    the plan, a hook on shapes it does not know) raises instead of taking the reference's torch code with a one-time
    warning.  ``World.step`` itself never falls back: an unsupported world is an error at plan build.
    ``fused_reset=True`` (opt-in): ``World.reset`` zeroes the SoA buffers in place (one launch per buffer,
    for one env or all), ``Entity.set_pos / set_vel / set_rot / set_ang_vel`` write through the bound views, and
    ``ScenarioUtils.find_random_pos_for_entity`` runs one launch per rejection round on the reference's own ``uniform_``
    draws - placement bit-identical to the reference, and no re-packing of rebound state tensors at the next step.  The
    price: a state tensor the caller passed to ``set_pos(..., batch_index=None)`` is copied, not adopted."""
    world = getattr(world_or_env, "world", world_or_env)
    if getattr(world, "_vmasb200", None) is not None:
        b = world._vmasb200  # already enabled: opt-ins can still be switched on
        b.static_io = b.static_io or bool(static_io)
        b.strict = b.strict or bool(strict)
        if hooks and b.hooks is None and getattr(b, "_env", None) is not None and b._env is world_or_env:
            b.hooks = _hooks.install(world_or_env, b)
        return b
    _lib.lib()  # fail loudly when the extension is missing
    b = WorldBinding(world, exact=exact)
    b.strict = bool(strict)
    b.refresh()
    world._vmasb200 = b
    world.step = types.MethodType(lambda self: b.step(), world)
    world.cast_rays = types.MethodType(
        lambda self, entity, angles, max_range, entity_filter=lambda _: False: b.cast_rays(entity, angles, max_range, entity_filter),
        world)
    # the scalar twin World.cast_ray (core.py:1627-1659; Lidar.measure(vectorized=False), sensors.py:104-113): same arithmetic,
    # one ray per env - the same launch with n_rays = 1
    b.orig_cast_ray = world.cast_ray

    def cast_ray(self, entity, angles, max_range, entity_filter=lambda _: False):
        pos = entity.state.pos
        assert pos.ndim == 2 and angles.ndim == 1  # core.py:1636-1637
        assert pos.shape[0] == angles.shape[0]
        return b.cast_rays(entity, angles.unsqueeze(-1), max_range, entity_filter)[:, 0]

    world.cast_ray = types.MethodType(cast_ray, world)
    b.orig_get_distance, b.orig_is_overlapping = world.get_distance, world.is_overlapping
    b.orig_get_distance_from_point = world.get_distance_from_point
    world.get_distance = types.MethodType(lambda self, a, c, env_index=None: b.get_distance(a, c, env_index), world)
    world.is_overlapping = types.MethodType(lambda self, a, c, env_index=None: b.is_overlapping(a, c, env_index), world)
    world.get_distance_from_point = types.MethodType(
        lambda self, entity, test_point_pos, env_index=None: b.get_distance_from_point(entity, test_point_pos, env_index), world)
    orig_reset = world.reset

    def reset(self, env_index=None):
        b.dirty = True  # collision filters / joint rotations may depend on what reset sets
        return orig_reset(env_index)

    world.reset = types.MethodType(reset, world)
    b._orig_reset = orig_reset
    if fused_reset:
        _install_fused_reset(world, b)
    b._orig_measures = []
    for ag in world.agents:
        for sensor in getattr(ag, "sensors", None) or []:
            if hasattr(sensor, "measure") and hasattr(sensor, "_angles"):
                om = sensor.measure
                b._orig_measures.append((sensor, om))
                sensor.measure = (lambda s, o: (lambda vectorized=True: b.measure(s, o, vectorized)))(sensor, om)
    b.static_io = bool(static_io)
    b._env = None
    if world is not world_or_env and hasattr(world_or_env, "_set_action"):  # an Environment: its action pre-processing too
        env = world_or_env
        b._env, b.orig_set_action = env, env._set_action
        env._set_action = types.MethodType(lambda self, action, agent: b.set_action(self, action, agent), env)
    b.hooks = _hooks.install(world_or_env, b) if (hooks and b._env is not None) else None
    return b


def disable(world_or_env) -> None:
    world = getattr(world_or_env, "world", world_or_env)
    b = getattr(world, "_vmasb200", None)
    if b is None:
        return
    if getattr(b, "fused_reset", False):
        _remove_fused_reset(world, b)
    for name in ("step", "cast_rays", "cast_ray", "reset", "get_distance", "is_overlapping", "get_distance_from_point"):
        if name in world.__dict__:
            del world.__dict__[name]
    for sensor, _ in b._orig_measures:
        if "measure" in sensor.__dict__:
            del sensor.__dict__["measure"]
    if getattr(b, "_env", None) is not None:
        b._env.__dict__.pop("_set_action", None)
    if getattr(b, "hooks", None) is not None:
        b.hooks.remove()
        b.hooks = None
    world._vmasb200 = None
