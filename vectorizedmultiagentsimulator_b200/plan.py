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

    # ------------------------------------------------------------------ SoA packing (numpy; tests / bench / host API)
    @staticmethod
    def ld_for(n: int) -> int:
        """Row stride of the SoA buffers for n environments: a multiple of 4 floats (16-byte rows for TMA)."""
        return max(4, (n + 3) & ~3)

    def pack(self, state: np.ndarray, jrot: Optional[np.ndarray] = None):
        """fixture layout state[E,N,6] (+ jrot[J,N]) -> SoA buffers dyn[dyn_rows, ld], stat[stat_rows, ld]."""
        n = state.shape[1]
        ld = self.ld_for(n)
        dyn = np.zeros((self.dyn_rows, ld), np.float32)
        stat = np.zeros((self.stat_rows, ld), np.float32)
        for e in range(self.E):
            if self.dyn_slot[e] >= 0:
                dyn[self.dyn_slot[e]: self.dyn_slot[e] + 6, :n] = state[e].T
            elif self.stat_slot[e] >= 0:
                o = self.stat_slot[e]
                stat[o: o + 2, :n] = state[e][:, 0:2].T
                stat[o + 2, :n] = state[e][:, 4]
        for j, o in enumerate(self.joint_slot):
            if o >= 0 and jrot is not None and jrot.size:
                stat[o, :n] = jrot[j]
        return dyn, stat

    def unpack(self, dyn: np.ndarray, state: np.ndarray) -> np.ndarray:
        """Writes the dyn rows back into a fixture-layout state (static entities untouched)."""
        n = state.shape[1]
        for e in range(self.E):
            if self.dyn_slot[e] >= 0:
                state[e] = dyn[self.dyn_slot[e]: self.dyn_slot[e] + 6, :n].T
        return state

    def lidar_of(self, lid: np.ndarray, k: int, n: int) -> np.ndarray:
        """[n, n_rays] measurement of sensor k out of the lidar SoA buffer."""
        o, r = self.lidar_slot[k], len(self.desc.lidars[k].angles)
        return lid[o: o + r, :n].T

    def act_of(self, act: np.ndarray, e: int, n: int):
        """clamped (force [n,2], torque [n]) of entity e out of the actout SoA buffer."""
        o = self.act_slot[e]
        return act[o: o + 2, :n].T, act[o + 2, :n]

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
    def step_host(self, dyn, stat, force=None, torque=None, exact=False, want_lidar=True, want_masks=True, out=None):
        """dyn [dyn_rows, ld], stat [stat_rows, ld], force [N,A,2], torque [N,A] numpy fp32 (N = force/torque rows, else
        ld).  out = optional (dyn_out, act, lidar) buffers to fill (e.g. pinned).  dyn / stat None (from the second call of a
        batch on): start from the state / static rows the plan's device staging holds from the previous call; `out` then
        gives the row stride.  Returns (dyn_out, act_out, lidar_out, pair_active_words)."""
        ld = dyn.shape[1] if dyn is not None else out[0].shape[1]
        n = force.shape[0] if force is not None else (torque.shape[0] if torque is not None else ld)
        assert dyn is None or (dyn.dtype == np.float32 and dyn.flags.c_contiguous and dyn.shape == (self.dyn_rows, ld))
        for a, shp in ((stat, (self.stat_rows, ld)), (force, (n, self.A, 2)), (torque, (n, self.A))):
            assert a is None or (a.dtype == np.float32 and a.flags.c_contiguous and a.shape == shp), shp
        if out is not None:
            dyn_out, act, lid = out
        else:
            dyn_out = np.zeros((self.dyn_rows, ld), np.float32)
            act = np.zeros((self.act_rows, ld), np.float32) if self.act_rows else None
            lid = np.zeros((self.lidar_rows, ld), np.float32) if (self.lidar_rows and want_lidar) else None
        pa = np.zeros((self.S, self.pair_words), np.uint32) if (self.pair_words and want_masks) else None
        rc = _lib.lib().vmasb200_step_host(self._h, _ptr(dyn), _ptr(stat) if (self.stat_rows and stat is not None) else None, _ptr(force),
                                           _ptr(torque), _ptr(dyn_out), _ptr(act), _ptr(lid), _ptr(pa), n, ld, int(exact))
        _lib.check(rc, "vmasb200_step_host")
        return dyn_out, act, lid, pa

    def host_stepper(self, n, ld, force, torque, dyn_out, act_out, lidar_out, dyn_in=None, stat=None):
        """A closure over ONE prepared vmasb200_step_host call (numpy host buffers, ideally pinned): what a rollout loop
        calls every step without re-validating its buffers - the C ABI call itself, nothing else.  dyn_in / stat None =
        continue from the state the plan's device staging holds (include/vmasb200.h)."""
        fn = _lib.lib().vmasb200_step_host
        args = (self._h, _ptr(dyn_in), _ptr(stat), _ptr(force), _ptr(torque), _ptr(dyn_out), _ptr(act_out), _ptr(lidar_out),
                None, int(n), int(ld), 0)
        keep = (force, torque, dyn_out, act_out, lidar_out, dyn_in, stat)

        def step(_keep=keep):
            rc = fn(*args)
            if rc:
                _lib.check(rc, "vmasb200_step_host")

        return step

    # ------------------------------------------------------------------ compute: device buffers (asynchronous)
    def make_io(self, dyn_in, dyn_out, stat, forces: List, torques: List, act_out, lidar_out, pair_active_out,
                n_envs: int, ld: int, exact: bool, stream: int):
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
        io.n_envs, io.ld, io.exact, io.stream = n_envs, ld, int(exact), stream
        io._keep = (fp, tp, fs, ts, dyn_in, dyn_out, stat, forces, torques, act_out, lidar_out, pair_active_out)
        return io

    # ------------------------------------------------------------------ probe: ONE substep on the GPU (parity tests)
    def trig_entities(self):
        """Entity index per cos/sin cache slot (row order of vmasb200_step_io.trig_in)."""
        buf = np.zeros(max(self.E, 1), np.int32)
        n = _lib.lib().vmasb200_plan_trig_entities(self._h, buf.ctypes.data_as(C.POINTER(C.c_int32)))
        return [int(e) for e in buf[:n]]

    def substep_probe(self, torch, state, force, torque, jrot, substep, trig=None, jexp=None):
        """Substep `substep` of World.step (the loop body of core.py:1974-2009) through vmasb200_step's probe fields, on
        fixture-layout numpy inputs: state [E,N,6], force [A,N,2] / torque [A,N] AS THEY ENTER THE SUBSTEP, jrot [J,N],
        trig [E,N,4] (cos sin cos2 sin2 per entity) to use instead of the kernel's own sincosf, or None; jexp [J,N]
        exp(|delta rot|) of the joints' constraint torque to use instead of the kernel's expf, or None.
        Returns dict(state [E,N,6], force [A,N,2], torque [A,N] (clamped write-back, where the plan has act rows),
        masks [E,E] bool, total_force [E,N,2], total_torque [E,N], joint [J,N,4] = fx fy (on entity_a) ta tb)."""
        n = state.shape[1]
        ld = self.ld_for(n)
        dev = torch.device("cuda", torch.cuda.current_device())
        dyn_h, stat_h = self.pack(state, jrot)
        dyn = torch.from_numpy(dyn_h).to(dev)
        stat = torch.from_numpy(stat_h).to(dev) if self.stat_rows else None
        dyn_out = torch.zeros_like(dyn)
        forces = [torch.from_numpy(np.ascontiguousarray(force[a])).to(dev) for a in range(self.A)]
        torques = [torch.from_numpy(np.ascontiguousarray(torque[a])).to(dev).unsqueeze(-1).contiguous() for a in range(self.A)]
        act = torch.zeros(self.act_rows, ld, device=dev) if self.act_rows else None
        pa = torch.zeros(self.S, max(self.pair_words, 1), dtype=torch.int32, device=dev) if self.pair_words else None
        n_dyn, J = self.n_dyn, len(self.desc.joints)
        fout = torch.full((3 * max(n_dyn, 1), ld), float("nan"), device=dev)
        jout = torch.full((4 * max(J, 1), ld), float("nan"), device=dev)
        tin = None
        if trig is not None:
            te = self.trig_entities()
            tin = torch.zeros(4 * max(len(te), 1), ld, device=dev)
            for r, e in enumerate(te):
                tin[4 * r: 4 * r + 4, :n] = torch.from_numpy(np.ascontiguousarray(trig[e].T)).to(dev)
        io = self.make_io(dyn, dyn_out, stat, forces, torques, act, None, pa, n, ld, True,
                          torch.cuda.current_stream().cuda_stream)
        io.substep_begin, io.substep_end, io.forces_are_substep_inputs = substep, substep + 1, 1
        jin = None
        if jexp is not None and J:
            jin = torch.zeros(J, ld, device=dev)
            jin[:, :n] = torch.from_numpy(np.ascontiguousarray(jexp)).to(dev)
        io.trig_in, io.force_out, io.joint_out, io.jexp_in = _ptr(tin), _ptr(fout), _ptr(jout) if J else None, _ptr(jin)
        self.step_io(io)
        torch.cuda.synchronize()
        out = dict(state=self.unpack(dyn_out.cpu().numpy(), state.copy()))
        f_out, t_out = force.copy(), torque.copy()
        if act is not None:
            a_h = act.cpu().numpy()
            for slot, e in enumerate(self.agents):
                if self.act_slot[e] >= 0:
                    f_out[slot], t_out[slot] = self.act_of(a_h, e, n)
        out["force"], out["torque"] = f_out, t_out
        out["masks"] = self.masks_matrix(pa.cpu().numpy().view(np.uint32))[substep] if pa is not None else np.zeros((self.E, self.E), bool)
        tf, tt = np.zeros((self.E, n, 2), np.float32), np.zeros((self.E, n), np.float32)
        fo = fout.cpu().numpy()
        dyn_ents = [e for e in range(self.E) if self.dyn_slot[e] >= 0]
        for i, e in enumerate(dyn_ents):
            tf[e] = fo[3 * i: 3 * i + 2, :n].T
            tt[e] = fo[3 * i + 2, :n]
        out["total_force"], out["total_torque"], out["dyn_entities"] = tf, tt, dyn_ents
        out["joint"] = jout.cpu().numpy()[: 4 * J, :n].reshape(J, 4, n).transpose(0, 2, 1) if J else np.zeros((0, n, 4), np.float32)
        return out

    def step_io(self, io):
        _lib.check(_lib.lib().vmasb200_step(self._h, C.byref(io)), "vmasb200_step")

    def cast_rays(self, dyn, stat, ld: int, entity: int, angles, max_range: float, targets, out, stream: int):
        n, r = angles.shape
        tg = np.ascontiguousarray(targets, np.int32)
        rc = _lib.lib().vmasb200_cast_rays(self._h, _ptr(dyn), _ptr(stat), ld, entity, _ptr(angles), angles.stride(0), r,
                                           float(max_range), tg.ctypes.data_as(C.POINTER(C.c_int32)), len(tg),
                                           _ptr(out), n, stream)
        _lib.check(rc, "vmasb200_cast_rays")
        return out

    def tile_for(self, n_envs: int) -> int:
        """Environments per CTA a step over n_envs environments runs with (vmasb200_plan_tile)."""
        return int(_lib.lib().vmasb200_plan_tile(self._h, int(n_envs)))

    def jit(self, n_envs: int):
        """(state, launches) of the world-compiled kernel a step over n_envs takes (include/vmasb200.h
        vmasb200_plan_jit): state 1 compiled (NVRTC, no GPU needed) / 0 world not covered or switched off / -1 failed."""
        st, ln, tl = C.c_int32(0), C.c_int64(0), C.c_int64(0)
        _lib.check(_lib.lib().vmasb200_plan_jit(self._h, int(n_envs), C.byref(st), C.byref(ln), C.byref(tl)), "vmasb200_plan_jit")
        self.jit_tail_launches = tl.value
        return st.value, ln.value

    def sync_check(self, stream: int = 0) -> int:
        nf = C.c_int64(0)
        _lib.check(_lib.lib().vmasb200_sync_check(self._h, stream, C.byref(nf)), "vmasb200_sync_check")
        return int(nf.value)
