"""ctypes binding of libvmasb200.so (include/vmasb200.h).  There is no fallback: a missing library is an error."""
import ctypes as C
import os
import subprocess

from .worlddesc import CLayout, CPair, CStepIO, CWorldDesc

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libvmasb200.so")
CSRC = os.path.join(HERE, "csrc")

EXPORTS = [
    "vmasb200_version", "vmasb200_last_error", "vmasb200_plan_create", "vmasb200_plan_destroy",
    "vmasb200_plan_layout", "vmasb200_plan_slots", "vmasb200_plan_pairs", "vmasb200_plan_joint_slots",
    "vmasb200_plan_lidar_slots", "vmasb200_step", "vmasb200_step_host", "vmasb200_cast_rays",
    "vmasb200_sync_check", "vmasb200_plan_tile", "vmasb200_get_distance", "vmasb200_is_overlapping",
    "vmasb200_get_distance_from_point", "vmasb200_set_action", "vmasb200_set_action_discrete", "vmasb200_gather_obs",
    "vmasb200_navigation_reward",
    "vmasb200_transport_reward", "vmasb200_simple_spread_reward", "vmasb200_jit_enable", "vmasb200_plan_jit",
    "vmasb200_joint_passage_reward", "vmasb200_football_reward", "vmasb200_navigation_reward_static", "vmasb200_plan_trig_entities", "vmasb200_set_action_ex", "vmasb200_reset_rows", "vmasb200_spawn_select",
]

_LIB = None


class VmasB200Error(RuntimeError):
    pass


def build(force: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a (csrc/Makefile); cross-compiles without a GPU."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "vmasb200.h")]
    stale = not os.path.exists(SO_PATH) or any(os.path.getmtime(s) > os.path.getmtime(SO_PATH) for s in srcs)
    if force or stale:
        r = subprocess.run(["make", "-C", CSRC] + (["-B"] if force else []), capture_output=True, text=True)
        if r.returncode != 0:
            raise VmasB200Error("building libvmasb200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return SO_PATH


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(SO_PATH):
        raise VmasB200Error(
            f"{SO_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(libvmasb200 is the only compute path; there is no CPU or torch fallback)"
        )
    L = C.CDLL(SO_PATH)
    i32p, vp = C.POINTER(C.c_int32), C.c_void_p
    L.vmasb200_version.restype = C.c_int
    L.vmasb200_last_error.restype = C.c_char_p
    L.vmasb200_plan_create.argtypes = [C.POINTER(CWorldDesc), C.POINTER(vp)]
    L.vmasb200_plan_destroy.argtypes = [vp]
    L.vmasb200_plan_layout.argtypes = [vp, C.POINTER(CLayout)]
    L.vmasb200_plan_slots.argtypes = [vp, i32p, i32p, i32p]
    L.vmasb200_plan_pairs.argtypes = [vp, C.POINTER(CPair)]
    L.vmasb200_plan_joint_slots.argtypes = [vp, i32p]
    L.vmasb200_plan_lidar_slots.argtypes = [vp, i32p]
    L.vmasb200_plan_trig_entities.argtypes = [vp, i32p]
    L.vmasb200_step.argtypes = [vp, C.POINTER(CStepIO)]
    L.vmasb200_step_host.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int64, C.c_int64, C.c_int32]
    L.vmasb200_cast_rays.argtypes = [
        vp, vp, vp, C.c_int64, C.c_int32, vp, C.c_int64, C.c_int32, C.c_float, i32p, C.c_int32, vp, C.c_int64, vp,
    ]
    L.vmasb200_sync_check.argtypes = [vp, vp, C.POINTER(C.c_int64)]
    L.vmasb200_plan_tile.argtypes = [vp, C.c_int64]
    L.vmasb200_get_distance.argtypes = [vp, vp, vp, C.c_int64, vp]
    L.vmasb200_is_overlapping.argtypes = [vp, vp, vp, C.c_int64, vp]
    L.vmasb200_get_distance_from_point.argtypes = [vp, vp, C.c_int64, C.c_int64, vp, C.c_int64, vp]
    L.vmasb200_set_action.argtypes = [vp, C.c_int64, C.c_int64, C.c_int32, vp, vp, C.c_int32, vp, vp, C.c_int64, vp]
    L.vmasb200_set_action_discrete.argtypes = [vp, C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32, vp, vp, vp, vp, vp,
                                               C.c_int64, vp]
    L.vmasb200_set_action_ex.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_reset_rows.argtypes = [vp, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_int64, vp, vp]
    L.vmasb200_spawn_select.argtypes = [vp, vp, vp, C.c_int32, C.c_float, vp, C.c_int64, vp]
    L.vmasb200_gather_obs.argtypes = [vp, C.c_int32, vp, vp, C.c_int32, C.c_int64, vp]
    L.vmasb200_navigation_reward.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_navigation_reward_static.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_transport_reward.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_simple_spread_reward.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_joint_passage_reward.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_football_reward.argtypes = [vp, C.c_int64, vp]
    L.vmasb200_jit_enable.argtypes = [C.c_int]
    L.vmasb200_plan_jit.argtypes = [vp, C.c_int64, i32p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    for name in EXPORTS:
        if name != "vmasb200_last_error":
            getattr(L, name).restype = C.c_int
    _LIB = L
    return L


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().vmasb200_last_error()
        raise VmasB200Error(f"{what} failed ({rc}): {msg.decode() if msg else ''}")


def raw_stream(torch, device) -> int:
    """cudaStream_t of torch's CURRENT stream on `device` (the stream every launch of this package goes to).
    torch._C._cuda_getCurrentRawStream is the cheap accessor (what torch's own Triton launcher uses); the public
    torch.cuda.current_stream(device).cuda_stream costs ~10 us per call, several times per step."""
    idx = device.index
    if idx is None:
        idx = torch.cuda.current_device()
    get = getattr(torch._C, "_cuda_getCurrentRawStream", None)
    return get(idx) if get is not None else torch.cuda.current_stream(device).cuda_stream
