"""Host side of the distance queries (include/vmasb200.h: vmasb200_get_distance / vmasb200_is_overlapping /
vmasb200_get_distance_from_point): the reference's World.get_distance (core.py:1821-1907), World.is_overlapping
(core.py:1909-1966) and World.get_distance_from_point (core.py:1787-1819) on CUDA tensors, one launch per call.

Same argument meaning and error behaviour as the reference: shapes other than Sphere / Box / Line raise RuntimeError.
There is no fallback: tensors must live on a CUDA device."""
import ctypes as C

from . import _lib
from .worlddesc import BOX, LINE, SPHERE, UnsupportedWorld, _shape_kind


class CPose(C.Structure):
    _fields_ = [
        ("shape", C.c_int32),
        ("radius", C.c_double), ("length", C.c_double), ("width", C.c_double),
        ("pos", C.c_void_p), ("pos_stride", C.c_int64), ("pos_stride_xy", C.c_int64),
        ("rot", C.c_void_p), ("rot_stride", C.c_int64),
    ]


def shape_of(entity):
    """(VMASB200_* shape id, radius, length, width) of a reference entity (duck-typed on the shape's class name)."""
    sh = entity.shape
    try:
        kind = _shape_kind(sh)
    except UnsupportedWorld:
        return None
    if kind == SPHERE:
        return SPHERE, float(sh.radius), 0.0, 0.0
    if kind == BOX:
        return BOX, 0.0, float(sh.length), float(sh.width)
    return LINE, 0.0, float(sh.length), 0.0


def _f32(t, torch):
    if t.dtype != torch.float32:
        t = t.to(torch.float32)
    if not t.is_cuda:
        raise _lib.VmasB200Error("distance queries need CUDA tensors (libvmasb200 has no CPU path)")
    return t


def make_pose(shape, radius, length, width, pos, rot, torch):
    """pos [N, 2], rot [N, 1] or [N] (any strides).  Returns (CPose, tensors to keep alive)."""
    pos = _f32(pos, torch)
    p = CPose()
    p.shape, p.radius, p.length, p.width = shape, radius, length, width
    p.pos, p.pos_stride, p.pos_stride_xy = pos.data_ptr(), pos.stride(0), pos.stride(1)
    keep = [pos]
    if shape != SPHERE:
        rot = _f32(rot, torch)
        p.rot, p.rot_stride = rot.data_ptr(), rot.stride(0)
        keep.append(rot)
    else:
        p.rot, p.rot_stride = None, 0
    return p, keep


def _stream(t, torch):
    return _lib.raw_stream(torch, t.device)


def get_distance(pose_a, pose_b, n, like, torch):
    out = torch.empty(n, dtype=torch.float32, device=like.device)
    with torch.cuda.device(like.device):
        _lib.check(_lib.lib().vmasb200_get_distance(C.byref(pose_a), C.byref(pose_b), out.data_ptr(), n, _stream(like, torch)),
                   "vmasb200_get_distance")
    return out


def is_overlapping(pose_a, pose_b, n, like, torch):
    out = torch.empty(n, dtype=torch.uint8, device=like.device)
    with torch.cuda.device(like.device):
        _lib.check(_lib.lib().vmasb200_is_overlapping(C.byref(pose_a), C.byref(pose_b), out.data_ptr(), n, _stream(like, torch)),
                   "vmasb200_is_overlapping")
    return out.view(torch.bool)


def get_distance_from_point(pose_a, points, n, torch):
    points = _f32(points, torch)
    out = torch.empty(n, dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        _lib.check(_lib.lib().vmasb200_get_distance_from_point(C.byref(pose_a), points.data_ptr(), points.stride(0),
                                                               points.stride(1), out.data_ptr(), n, _stream(points, torch)),
                   "vmasb200_get_distance_from_point")
    return out
