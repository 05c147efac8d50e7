"""ctypes front-end of oracle/libvmas_oracle.so - TEST INFRASTRUCTURE, never on the product path.

State layout is the fixture layout: ``state[E, N, 6]`` = px py vx vy rot ang_vel (fp32, C-contiguous),
``force[A, N, 2]``, ``torque[A, N]``, ``jrot[J, N]``; masks ``[S, E, E]`` uint8.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(HERE, "libvmas_oracle.so")
    srcs = [os.path.join(HERE, f) for f in ("vmas_oracle.c", "vmas_oracle.h", "sleef_f32.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(so) < os.path.getmtime(x) for x in srcs):
        subprocess.run(["make", "-C", HERE, "-B"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        fp = C.POINTER(C.c_float)
        _LIB.vmas_oracle_step.restype = C.c_int
        _LIB.vmas_oracle_step.argtypes = [C.c_void_p, C.c_int64, fp, fp, fp, fp, C.POINTER(C.c_uint8), C.c_int]
        _LIB.vmas_oracle_substep.restype = C.c_int
        _LIB.vmas_oracle_substep.argtypes = [C.c_void_p, C.c_int64, fp, fp, fp, fp, C.POINTER(C.c_uint8), C.c_int, fp, fp, fp, fp, fp, fp, C.c_int, C.c_int]
        _LIB.vmas_oracle_lidar.restype = C.c_int
        _LIB.vmas_oracle_lidar.argtypes = [C.c_void_p, C.c_int64, fp, C.c_int, fp, C.c_int]
        _LIB.vmas_oracle_cast_rays.restype = C.c_int
        _LIB.vmas_oracle_cast_rays.argtypes = [
            C.c_void_p, C.c_int64, fp, C.c_int, fp, C.c_int, C.c_float, C.POINTER(C.c_int32), C.c_int, fp, C.c_int,
        ]
        dp = C.POINTER(C.c_double)
        _LIB.vmas_oracle_get_distance.restype = C.c_int
        _LIB.vmas_oracle_get_distance.argtypes = [C.c_int, dp, fp, C.c_int, dp, fp, C.c_int64, fp]
        _LIB.vmas_oracle_is_overlapping.restype = C.c_int
        _LIB.vmas_oracle_is_overlapping.argtypes = [C.c_int, dp, fp, C.c_int, dp, fp, C.c_int64, C.POINTER(C.c_uint8)]
        _LIB.vmas_oracle_distance_from_point.restype = C.c_int
        _LIB.vmas_oracle_distance_from_point.argtypes = [C.c_int, dp, fp, fp, C.c_int64, fp]
    return _LIB


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class Oracle:
    """Holds the ctypes world description (and its backing buffers) for repeated calls."""

    def __init__(self, desc, n_threads: int = 1):
        self.desc = desc
        self.cdesc, self._keep = desc.to_ctypes()
        self.n_threads = n_threads
        self.E = len(desc.entities)
        self.A = len(desc.agent_indices)
        self.S = desc.substeps
        self.J = len(desc.joints)

    def step(self, state, force, torque, jrot=None, want_masks=True):
        """In-place World.step on (state, force, torque). Returns masks [S,E,E] (bool) or None."""
        assert state.dtype == np.float32 and state.flags.c_contiguous and state.shape[0] == self.E
        n = state.shape[1]
        if force is None:
            force = np.zeros((self.A, n, 2), np.float32)
        if torque is None:
            torque = np.zeros((self.A, n), np.float32)
        assert force.dtype == np.float32 and force.flags.c_contiguous and force.shape == (self.A, n, 2)
        assert torque.dtype == np.float32 and torque.flags.c_contiguous and torque.shape == (self.A, n)
        masks = np.zeros((self.S, self.E, self.E), np.uint8) if want_masks else None
        if jrot is not None:
            jrot = np.ascontiguousarray(jrot, np.float32)
            assert jrot.shape == (self.J, n)
        rc = lib().vmas_oracle_step(
            C.byref(self.cdesc), n, _fp(state), _fp(force), _fp(torque),
            _fp(jrot) if jrot is not None and jrot.size else None,
            masks.ctypes.data_as(C.POINTER(C.c_uint8)) if masks is not None else None, self.n_threads,
        )
        if rc:
            raise RuntimeError(f"vmas_oracle_step failed: {rc}")
        return masks.astype(bool) if masks is not None else None

    def torch_math(self, state, jrot):
        """(trig [E,N,4], jexp [J,N]): what the reference's torch CPU ops return for the cos / sin / exp calls of one
        substep from `state` (utils.py:183-184, physics.py:299-301, 412; core.py:2852).  torch's unary CPU kernels are
        MKL VML's, which cannot be restated (closed source); they are pure per-element functions (tests/test_oracle_sleef.py),
        so evaluating them here on the same fp32 arguments gives the reference's values."""
        import torch

        rot = torch.from_numpy(np.ascontiguousarray(state[..., 4]))
        rot2 = rot + torch.pi / 2  # physics.py:300
        trig = torch.stack([rot.cos(), rot.sin(), rot2.cos(), rot2.sin()], -1).numpy()
        n = state.shape[1]
        jexp = np.zeros((self.J, n), np.float32)
        for j, jt in enumerate(self.desc.joints):
            if not jt.rotate:
                fr = jrot[j] if (jt.fixed_rotation_per_env and jrot is not None and len(jrot)) else np.float32(jt.fixed_rotation)
                ra, rb = torch.from_numpy(state[jt.entity_a, :, 4].copy()), torch.from_numpy(state[jt.entity_b, :, 4].copy())
                d = ra - (rb + torch.as_tensor(fr, dtype=torch.float32))
                jexp[j] = torch.exp(torch.linalg.vector_norm(d.unsqueeze(-1), dim=-1)).numpy()  # core.py:2848-2852
        return np.ascontiguousarray(trig, np.float32), jexp

    def step_torch_math(self, state, force, torque, jrot=None, vec_width=0, trig_jitter=None):
        """World.step like `step`, substep by substep with the reference's own cos / sin / exp values (torch_math).
        trig_jitter: a numpy Generator; every cos / sin / exp value is then moved by -1 / 0 / +1 ulp at random - what another
        math library (libm, CUDA) does to them - to measure how far that alone moves the result.  Returns masks [S,E,E]."""
        masks = np.zeros((self.S, self.E, self.E), bool)
        for s in range(self.S):
            trig, jexp = self.torch_math(state, jrot)
            if trig_jitter is not None:
                def jit(x):
                    k = trig_jitter.integers(-1, 2, x.shape)
                    up = np.nextafter(x, np.float32(np.inf)).astype(np.float32)
                    dn = np.nextafter(x, np.float32(-np.inf)).astype(np.float32)
                    return np.ascontiguousarray(np.where(k > 0, up, np.where(k < 0, dn, x)), np.float32)
                trig, jexp = jit(trig), jit(jexp)
            masks[s] = self.substep(state, force, torque, jrot, s, trig=trig, jexp=jexp, vec_width=vec_width)[0]
        return masks

    def substep(self, state, force, torque, jrot, substep, trig=None, jexp=None, vec_width=0):
        """In place: ONE substep of World.step (core.py:1974-2009) on (state, force, torque).  Returns
        (masks [E,E] bool, total_force [E,N,2], total_torque [E,N], joint_force [J,2,N,2], joint_torque [J,2,N]).
        trig [E,N,4] / jexp [J,N]: cos / sin / exp values to use instead of libm's (see torch_math)."""
        n = state.shape[1]
        assert state.dtype == np.float32 and state.flags.c_contiguous and state.shape[0] == self.E
        assert force.dtype == np.float32 and force.flags.c_contiguous and force.shape == (self.A, n, 2)
        assert torque.dtype == np.float32 and torque.flags.c_contiguous and torque.shape == (self.A, n)
        masks = np.zeros((self.S, self.E, self.E), np.uint8)
        tf, tt = np.zeros((self.E, n, 2), np.float32), np.zeros((self.E, n), np.float32)
        jf, jt = np.zeros((self.J, 2, n, 2), np.float32), np.zeros((self.J, 2, n), np.float32)
        if jrot is not None:
            jrot = np.ascontiguousarray(jrot, np.float32)
        if trig is not None:
            assert trig.dtype == np.float32 and trig.flags.c_contiguous and trig.shape == (self.E, n, 4)
        if jexp is not None:
            assert jexp.dtype == np.float32 and jexp.flags.c_contiguous and jexp.shape == (self.J, n)
        rc = lib().vmas_oracle_substep(
            C.byref(self.cdesc), n, _fp(state), _fp(force), _fp(torque),
            _fp(jrot) if jrot is not None and jrot.size else None, masks.ctypes.data_as(C.POINTER(C.c_uint8)), substep,
            _fp(tf), _fp(tt), _fp(jf) if self.J else None, _fp(jt) if self.J else None,
            _fp(trig) if trig is not None else None, _fp(jexp) if (jexp is not None and self.J) else None, int(vec_width),
            self.n_threads)
        if rc:
            raise RuntimeError(f"vmas_oracle_substep failed: {rc}")
        return masks[substep].astype(bool), tf, tt, jf, jt

    def lidar(self, state, k):
        n = state.shape[1]
        out = np.empty((n, len(self.desc.lidars[k].angles)), np.float32)
        rc = lib().vmas_oracle_lidar(C.byref(self.cdesc), n, _fp(state), k, _fp(out), self.n_threads)
        if rc:
            raise RuntimeError(f"vmas_oracle_lidar failed: {rc}")
        return out

    def cast_rays(self, state, entity, angles, max_range, targets):
        n = state.shape[1]
        angles = np.ascontiguousarray(angles, np.float32)
        tg = np.ascontiguousarray(targets, np.int32)
        out = np.empty_like(angles)
        rc = lib().vmas_oracle_cast_rays(
            C.byref(self.cdesc), n, _fp(state), entity, _fp(angles), angles.shape[1], max_range,
            tg.ctypes.data_as(C.POINTER(C.c_int32)), len(tg), _fp(out), self.n_threads,
        )
        if rc:
            raise RuntimeError(f"vmas_oracle_cast_rays failed: {rc}")
        return out


# ---- distance queries: shape = VMASB200_SPHERE/BOX/LINE, dims = (radius, length, width) Python floats,
# state = the entity's [N, 6] block of the fixture state
def _dims(d):
    a = np.ascontiguousarray(d, np.float64)
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


def get_distance(shape_a, dims_a, state_a, shape_b, dims_b, state_b):
    sa, sb = np.ascontiguousarray(state_a, np.float32), np.ascontiguousarray(state_b, np.float32)
    (ka, da), (kb, db) = _dims(dims_a), _dims(dims_b)
    out = np.empty(sa.shape[0], np.float32)
    rc = lib().vmas_oracle_get_distance(int(shape_a), da, _fp(sa), int(shape_b), db, _fp(sb), sa.shape[0], _fp(out))
    assert rc == 0
    return out


def is_overlapping(shape_a, dims_a, state_a, shape_b, dims_b, state_b):
    sa, sb = np.ascontiguousarray(state_a, np.float32), np.ascontiguousarray(state_b, np.float32)
    (ka, da), (kb, db) = _dims(dims_a), _dims(dims_b)
    out = np.empty(sa.shape[0], np.uint8)
    rc = lib().vmas_oracle_is_overlapping(int(shape_a), da, _fp(sa), int(shape_b), db, _fp(sb), sa.shape[0],
                                          out.ctypes.data_as(C.POINTER(C.c_uint8)))
    assert rc == 0
    return out.astype(bool)


def distance_from_point(shape_a, dims_a, state_a, points):
    sa, pts = np.ascontiguousarray(state_a, np.float32), np.ascontiguousarray(points, np.float32)
    ka, da = _dims(dims_a)
    out = np.empty(sa.shape[0], np.float32)
    rc = lib().vmas_oracle_distance_from_point(int(shape_a), da, _fp(sa), _fp(pts), sa.shape[0], _fp(out))
    assert rc == 0
    return out
