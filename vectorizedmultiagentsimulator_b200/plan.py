"""StepPlan: Python handle on a ``vmasb200_plan`` (include/vmasb200.h) built from a :class:`WorldDesc`.

The plan replaces the Python half of ``World.step`` (core.py:1971-2014): pair enumeration and bucketing
(core.py:2103-2173), joint lookup, accumulation order.  Compute entry points take raw pointers: torch CUDA tensors
for :meth:`step` / :meth:`cast_rays`, numpy host arrays for :meth:`step_host`.
"""
import ctypes as C
from typing import List, Optional

import numpy as np

from . import _lib
from .worlddesc import CLayout, CPair, CStepIO, WorldDesc


def _ptr(t) -> Optional[int]:
    if t is None:
        return None
    return t.data_ptr() if hasattr(t, "data_ptr") else t.ctypes.data


class StepPlan:
    def __init__(self, desc: WorldDesc):
        self.desc = desc
        L = _lib.lib()
        self._cdesc, self._keep = desc.to_ctypes()
        h = C.c_void_p()
        _lib.check(L.vmasb200_plan_create(C.byref(self._cdesc), C.byref(h)), "vmasb200_plan_create")
        self._h = h
        lay = CLayout()
        _lib.check(L.vmasb200_plan_layout(h, C.byref(lay)), "vmasb200_plan_layout")
        self.layout = lay
        for f, _ in CLayout._fields_:
            setattr(self, f, int(getattr(lay, f)))
        E = len(desc.entities)
        ds, ss, as_ = (np.zeros(max(E, 1), np.int32) for _ in range(3))
        i32p = C.POINTER(C.c_int32)
        _lib.check(L.vmasb200_plan_slots(h, ds.ctypes.data_as(i32p), ss.ctypes.data_as(i32p), as_.ctypes.data_as(i32p)),
                   "vmasb200_plan_slots")
        self.dyn_slot, self.stat_slot, self.act_slot = ds[:E], ss[:E], as_[:E]
        pairs = (CPair * max(self.n_pairs, 1))()
        _lib.check(L.vmasb200_plan_pairs(h, pairs), "vmasb200_plan_pairs")
        self.pairs = [(int(p.type), int(p.a), int(p.b)) for p in pairs[: self.n_pairs]]
        js = np.zeros(max(len(desc.joints), 1), np.int32)
        _lib.check(L.vmasb200_plan_joint_slots(h, js.ctypes.data_as(i32p)), "vmasb200_plan_joint_slots")
        self.joint_slot = js[: len(desc.joints)]
        ls = np.zeros(max(len(desc.lidars), 1), np.int32)
        _lib.check(L.vmasb200_plan_lidar_slots(h, ls.ctypes.data_as(i32p)), "vmasb200_plan_lidar_slots")
        self.lidar_slot = ls[: len(desc.lidars)]
        self.E, self.S = E, desc.substeps
        self.agents = desc.agent_indices
        self.A = len(self.agents)

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().vmasb200_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ record packing (numpy; tests / bench / host API)
    def pack(self, state: np.ndarray, jrot: Optional[np.ndarray] = None):
        """fixture layout state[E,N,6] (+ jrot[J,N]) -> env records dyn[N,dyn_stride], stat[N,stat_stride]."""
        n = state.shape[1]
        dyn = np.zeros((n, self.dyn_stride), np.float32)
        stat = np.zeros((n, self.stat_stride), np.float32)
        for e in range(self.E):
            if self.dyn_slot[e] >= 0:
                dyn[:, self.dyn_slot[e]: self.dyn_slot[e] + 6] = state[e]
            elif self.stat_slot[e] >= 0:
                o = self.stat_slot[e]
                stat[:, o: o + 2] = state[e][:, 0:2]
                stat[:, o + 2] = state[e][:, 4]
        for j, o in enumerate(self.joint_slot):
            if o >= 0 and jrot is not None and jrot.size:
                stat[:, o] = jrot[j]
        return dyn, stat

    def unpack(self, dyn: np.ndarray, state: np.ndarray) -> np.ndarray:
        """Writes the dyn records back into a fixture-layout state (static entities untouched)."""
        for e in range(self.E):
            if self.dyn_slot[e] >= 0:
                state[e] = dyn[:, self.dyn_slot[e]: self.dyn_slot[e] + 6]
        return state

    def masks_matrix(self, words: np.ndarray) -> np.ndarray:
        """pair_active words [S, pair_words] -> bool [S, E, E] (a<b), the shape the golden fixtures store."""
        m = np.zeros((self.S, self.E, self.E), bool)
        if not self.pair_words:
            return m
        words = np.asarray(words, np.uint32).reshape(self.S, self.pair_words)
        for p, (_, a, b) in enumerate(self.pairs):
            lo, hi = min(a, b), max(a, b)
            m[:, lo, hi] = (words[:, p >> 5] >> np.uint32(p & 31)) & np.uint32(1)
        return m

    # ------------------------------------------------------------------ compute: host buffers (H2D + kernel + D2H inside)
    def step_host(self, dyn, stat, force=None, torque=None, exact=False, want_lidar=True, want_masks=True):
        """dyn [N,dyn_stride], stat [N,stat_stride], force [N,A,2], torque [N,A] numpy fp32.
        Returns (dyn_out, act_out, lidar_out, pair_active_words)."""
        n = dyn.shape[0]
        assert dyn.dtype == np.float32 and dyn.flags.c_contiguous and dyn.shape == (n, self.dyn_stride)
        for a, shp in ((stat, (n, self.stat_stride)), (force, (n, self.A, 2)), (torque, (n, self.A))):
            assert a is None or (a.dtype == np.float32 and a.flags.c_contiguous and a.shape == shp), shp
        dyn_out = np.empty_like(dyn)
        act = np.zeros((n, self.act_stride), np.float32) if self.act_stride else None
        lid = np.zeros((n, self.lidar_stride), np.float32) if (self.lidar_stride and want_lidar) else None
        pa = np.zeros((self.S, self.pair_words), np.uint32) if (self.pair_words and want_masks) else None
        rc = _lib.lib().vmasb200_step_host(self._h, _ptr(dyn), _ptr(stat), _ptr(force), _ptr(torque), _ptr(dyn_out),
                                           _ptr(act), _ptr(lid), _ptr(pa), n, int(exact))
        _lib.check(rc, "vmasb200_step_host")
        return dyn_out, act, lid, pa

    # ------------------------------------------------------------------ compute: device buffers (asynchronous)
    def make_io(self, dyn_in, dyn_out, stat, forces: List, torques: List, act_out, lidar_out, pair_active_out,
                n_envs: int, exact: bool, stream: int):
        """Builds a reusable vmasb200_step_io; forces/torques are per-agent CUDA tensors ([N,2] / [N,1]) or None."""
        io = CStepIO()
        fp = (C.c_void_p * max(self.A, 1))()
        tp = (C.c_void_p * max(self.A, 1))()
        fs = (C.c_int64 * max(self.A, 1))()
        ts = (C.c_int64 * max(self.A, 1))()
        for a in range(self.A):
            f, t = forces[a], torques[a]
            fp[a] = _ptr(f)
            tp[a] = _ptr(t)
            fs[a] = f.stride(0) if f is not None else 0
            ts[a] = t.stride(0) if t is not None else 0
        io.dyn_in, io.dyn_out, io.stat = _ptr(dyn_in), _ptr(dyn_out), _ptr(stat)
        io.force_ptrs, io.force_strides = C.cast(fp, C.POINTER(C.c_void_p)), C.cast(fs, C.POINTER(C.c_int64))
        io.torque_ptrs, io.torque_strides = C.cast(tp, C.POINTER(C.c_void_p)), C.cast(ts, C.POINTER(C.c_int64))
        io.act_out, io.lidar_out, io.pair_active_out = _ptr(act_out), _ptr(lidar_out), _ptr(pair_active_out)
        io.n_envs, io.exact, io.stream = n_envs, int(exact), stream
        io._keep = (fp, tp, fs, ts, dyn_in, dyn_out, stat, forces, torques, act_out, lidar_out, pair_active_out)
        return io

    def step_io(self, io):
        _lib.check(_lib.lib().vmasb200_step(self._h, C.byref(io)), "vmasb200_step")

    def cast_rays(self, dyn, stat, entity: int, angles, max_range: float, targets, out, stream: int):
        n, r = angles.shape
        tg = np.ascontiguousarray(targets, np.int32)
        rc = _lib.lib().vmasb200_cast_rays(self._h, _ptr(dyn), _ptr(stat), entity, _ptr(angles), angles.stride(0), r,
                                           float(max_range), tg.ctypes.data_as(C.POINTER(C.c_int32)), len(tg),
                                           _ptr(out), n, stream)
        _lib.check(rc, "vmasb200_cast_rays")
        return out

    def sync_check(self, stream: int = 0) -> int:
        nf = C.c_int64(0)
        _lib.check(_lib.lib().vmasb200_sync_check(self._h, stream, C.byref(nf)), "vmasb200_sync_check")
        return int(nf.value)
