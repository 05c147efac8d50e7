"""CPU: the world-compiled kernels (csrc/jit.h) compile without a GPU - NVRTC is a compiler, the cubin is only loaded at
the first launch - for exactly the worlds the generator covers, and the generic kernel stays the path of the others."""
import os

import pytest

from tests.fixtures import Fixture
from vectorizedmultiagentsimulator_b200 import _lib
from vectorizedmultiagentsimulator_b200.plan import StepPlan

HAVE_NVRTC = os.path.exists("/usr/local/cuda/lib64/libnvrtc.so.12")
COVERED = ["simple_spread", "navigation", "transport", "flocking", "wheel", "dropout",  # short or all-sphere pair lists
           "football", "joint_passage"]  # pair-heavy without LIDAR: the BIG variant (static tile, table-driven narrow phase)
NOT_COVERED = ["pollock"]                # 990 pairs and a LIDAR epilogue: the generic kernel's teams
FORCED = ["waterfall", "passage"]        # covered by the generator but not chosen by default: mode 2 (used by the GPU tests)


@pytest.mark.skipif(not HAVE_NVRTC, reason="no libnvrtc in this image")
@pytest.mark.parametrize("name", COVERED)
def test_covered_worlds_compile(name, tmp_path, monkeypatch):
    monkeypatch.setenv("VMASB200_JIT_CACHE", str(tmp_path))
    fx = Fixture(name)
    pl = StepPlan(fx.desc)
    for n in (fx.n_envs, 1 << 20):  # the two CTA shapes (jit.h jit_tile)
        state, launches = pl.jit(n)
        assert state == 1, _lib.lib().vmasb200_last_error()
        assert launches == 0
    assert any(f.endswith(".cubin") for f in os.listdir(tmp_path)), "the cubin cache was not written"
    # a second plan of the same world takes the cached cubin (same source, same hash)
    assert StepPlan(fx.desc).jit(fx.n_envs)[0] == 1


@pytest.mark.parametrize("name", NOT_COVERED)
def test_other_worlds_keep_the_generic_kernel(name):
    fx = Fixture(name)
    assert StepPlan(fx.desc).jit(fx.n_envs) == (0, 0)


def test_switch():
    L = _lib.lib()
    fx = Fixture("simple_spread")
    prev = L.vmasb200_jit_enable(0)
    try:
        assert StepPlan(fx.desc).jit(fx.n_envs) == (0, 0)
    finally:
        L.vmasb200_jit_enable(prev)


@pytest.mark.skipif(not HAVE_NVRTC, reason="no libnvrtc in this image")
@pytest.mark.parametrize("name", FORCED)
def test_mode_2_compiles_every_covered_world(name, tmp_path, monkeypatch):
    monkeypatch.setenv("VMASB200_JIT_CACHE", str(tmp_path))
    L = _lib.lib()
    fx = Fixture(name)
    prev = L.vmasb200_jit_enable(2)
    try:
        assert StepPlan(fx.desc).jit(fx.n_envs)[0] == 1, L.vmasb200_last_error()
    finally:
        L.vmasb200_jit_enable(prev)


@pytest.mark.skipif(not HAVE_NVRTC, reason="no libnvrtc in this image")
def test_experiment_switches_are_part_of_the_cache_key(tmp_path, monkeypatch):
    """VMASB200_JIT_DEFINES (extra NVRTC options) and VMASB200_JIT_NT (threads of a LIDAR world's CTA), the switches
    of tools/ab_jit.sh, must give cubins of their own: a variant may never be served from another variant's cache"""
    monkeypatch.setenv("VMASB200_JIT_CACHE", str(tmp_path))
    fx = Fixture("navigation")
    seen = set()
    for defs, nt in (("", ""), ("-DVB_AB_SINCOS_P2", ""), ("", "64")):
        monkeypatch.setenv("VMASB200_JIT_DEFINES", defs)
        if nt:
            monkeypatch.setenv("VMASB200_JIT_NT", nt)
        else:
            monkeypatch.delenv("VMASB200_JIT_NT", raising=False)
        assert StepPlan(fx.desc).jit(fx.n_envs)[0] == 1, _lib.lib().vmasb200_last_error()
        now = {f for f in os.listdir(tmp_path) if f.endswith(".cubin")}
        assert len(now - seen) == 1, (defs, nt, now)
        seen = now


@pytest.mark.skipif(not HAVE_NVRTC, reason="no libnvrtc in this image")
def test_pair_heavy_worlds_get_the_looped_broad_phase(tmp_path, monkeypatch):
    """The generator's shape for pair-heavy worlds: football's 70 line-sphere pairs are ONE loop over its 10
    static lines (7 rows per block, packed in pairs), joint_passage's two runs over the 13 wall boxes share one loop, the
    narrow phase is shared out over the warp, and football parks its cold state in the tile."""
    import re

    monkeypatch.setenv("VMASB200_JIT_CACHE", str(tmp_path))
    src = {}
    for name in ("football", "joint_passage", "waterfall"):
        dump = tmp_path / f"{name}.cu"
        monkeypatch.setenv("VMASB200_JIT_DUMP", str(dump))
        L = _lib.lib()
        prev = L.vmasb200_jit_enable(2)
        try:
            assert StepPlan(Fixture(name).desc).jit(1 << 20)[0] == 1, L.vmasb200_last_error()
        finally:
            L.vmasb200_jit_enable(prev)
        src[name] = dump.read_text()
    fb, jp, wf = src["football"], src["joint_passage"], src["waterfall"]
    assert "BIG = true" in fb and "COOP = true, PARK = true" in fb
    assert len(re.findall(r"for \(int j = 9; j >= 0; --j\)", fb)) == 1          # one loop over the 10 lines
    assert fb.count("pair_l2<") == 3 and fb.count("pair_l<") == 1               # 7 rows per block: 3 packed pairs + 1
    assert fb.count("pair_b<") == 21                                            # the sphere-sphere pairs stay unrolled
    assert "COOP = true, PARK = false" in jp and "COOP_EDGES = true" in jp      # box-line pairs over four lanes
    loops = re.findall(r"for \(int j = 12; j >= 0; --j\)", jp)
    assert len(loops) == 1 and jp.count("uint32_t av") == 3                     # two runs over the 13 boxes in one loop, + the 4 lines
    assert "COOP = false" in wf                                                 # box-box pairs: lanes keep their own lists
