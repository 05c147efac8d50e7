"""CPU: oracle/sleef_f32.h (scalar restatement of SLEEF's expf_u10 / log1pf_u10 / sinf_u10 / cosf_u10, the third-party
math behind ATen's Vectorized<float>) against torch itself, bit for bit, and the facts the oracle's `torch_math` mode
relies on: torch.logaddexp's vector body IS exp_u10 + log1p_u10, its tail is libm, and torch.cos / sin / exp are pure
per-element functions (so evaluating them on the same fp32 arguments gives the reference's values)."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SHIM = r"""
#include "sleef_f32.h"
void v_exp(const float* x, float* o, long n) { for (long i = 0; i < n; ++i) o[i] = sleef_expf_u10(x[i]); }
void v_log1p(const float* x, float* o, long n) { for (long i = 0; i < n; ++i) o[i] = sleef_log1pf_u10(x[i]); }
void v_softplus(const float* x, float* o, long n) {
  for (long i = 0; i < n; ++i) { float m = 0.f > x[i] ? 0.f : x[i]; o[i] = m + sleef_log1pf_u10(sleef_expf_u10(-fabsf(0.f - x[i]))); }
}
void g_softplus(const float* x, float* o, long n) {
  for (long i = 0; i < n; ++i) { float m = 0.f > x[i] ? 0.f : x[i]; o[i] = m + log1pf(expf(-fabsf(0.f - x[i]))); }
}
"""


@pytest.fixture(scope="module")
def shim():
    d = tempfile.mkdtemp()
    src, so = os.path.join(d, "shim.c"), os.path.join(d, "shim.so")
    open(src, "w").write(SHIM)
    subprocess.run(["gcc", "-O2", "-mfma", "-ffp-contract=off", "-fPIC", "-shared", "-I", os.path.join(ROOT, "oracle"),
                    "-o", so, src, "-lm"], check=True)
    return C.CDLL(so)


def run(lib, fn, x):
    o = np.empty_like(x)
    fp = C.POINTER(C.c_float)
    getattr(lib, fn)(x.ctypes.data_as(fp), o.ctypes.data_as(fp), C.c_long(x.size))
    return o


def bits(a):
    return a.view(np.uint32)


def test_log1p_is_sleef_u10(shim):
    rng = np.random.default_rng(0)
    for x in (rng.uniform(0, 1, 1 << 20), np.exp(rng.uniform(-30, 0, 1 << 20)), rng.uniform(-0.99, 5, 1 << 20)):
        x = np.ascontiguousarray(x, np.float32)
        assert np.array_equal(bits(run(shim, "v_log1p", x)), bits(torch.log1p(torch.from_numpy(x)).numpy()))


def test_logaddexp_body_is_sleef_and_tail_is_libm(shim):
    """The arguments of core.py:2820: logaddexp(0, (dist_min - dist) * sign / k), k = 1e-3."""
    rng = np.random.default_rng(1)
    cap = torch.backends.cpu.get_cpu_capability()
    W = {"AVX512": 32, "AVX2": 16}.get(cap)
    if W is None:
        pytest.skip(f"ATen CPU capability {cap}: no FMA vector path to compare with")
    torch.set_num_threads(1)
    n = 64 * W + W - 3  # whole groups + a tail
    for x in (rng.uniform(-30, 30, n), rng.normal(0, 2, n), rng.uniform(-100, 5, n)):
        x = np.ascontiguousarray(x, np.float32)
        ref = torch.logaddexp(torch.tensor(0.0), torch.from_numpy(x)).numpy()
        body = (n // W) * W
        assert np.array_equal(bits(run(shim, "v_softplus", x))[:body], bits(ref)[:body])
        assert np.array_equal(bits(run(shim, "g_softplus", x))[body:], bits(ref)[body:])


@pytest.mark.parametrize("fn", [torch.cos, torch.sin, torch.exp])
def test_unary_kernels_are_per_element_functions(fn):
    torch.set_num_threads(1)
    rng = np.random.default_rng(2)
    x = torch.from_numpy(rng.uniform(-7, 7, 1 << 14).astype(np.float32))
    full = fn(x)
    for n in (1, 2, 3, 7, 8, 9, 15, 16, 17, 31, 33, 48, 100):
        for off in (0, 1, 3, 8, 13):
            assert torch.equal(fn(x[off:off + n].clone()), full[off:off + n])
    assert torch.equal(fn(x[:48].clone().reshape(16, 3, 1)).reshape(-1), full[:48])
    assert torch.equal(fn(x[:96].clone().reshape(48, 2)[:, 0]), full[:96:2])
