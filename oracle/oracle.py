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
    src = os.path.join(HERE, "vmas_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-B"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        fp = C.POINTER(C.c_float)
        _LIB.vmas_oracle_step.restype = C.c_int
        _LIB.vmas_oracle_step.argtypes = [C.c_void_p, C.c_int64, fp, fp, fp, fp, C.POINTER(C.c_uint8), C.c_int]
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
