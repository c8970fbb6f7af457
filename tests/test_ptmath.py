"""include/ptmath.h: accuracy against double precision and a known-answer hash (the functions define the parity
arithmetic, so their bits must never drift)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from helpers import ROOT, sha

SRC = r'''
#include "ptmath.h"
void v_sin(const float*x,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_sinf(x[i]);}
void v_cos(const float*x,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_cosf(x[i]);}
void v_asin(const float*x,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_asinf(x[i]);}
void v_log(const float*x,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_logf(x[i]);}
void v_exp(const float*x,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_expf(x[i]);}
void v_atan2(const float*a,const float*b,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_atan2f(a[i],b[i]);}
void v_pow(const float*a,const float*b,float*y,int n){for(int i=0;i<n;i++)y[i]=pt_powf(a[i],b[i]);}
'''


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("ptm")
    (d / "t.c").write_text(SRC)
    so = str(d / "t.so")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", os.path.join(ROOT, "include"),
                           str(d / "t.c"), "-o", so, "-lm"])
    return C.CDLL(so)


def _run1(L, name, x):
    fp = np.ctypeslib.ndpointer(np.float32)
    f = getattr(L, name); f.argtypes = [fp, fp, C.c_int]
    y = np.empty_like(x); f(x, y, len(x)); return y


def _run2(L, name, a, b):
    fp = np.ctypeslib.ndpointer(np.float32)
    f = getattr(L, name); f.argtypes = [fp, fp, fp, C.c_int]
    y = np.empty_like(a); f(a, b, y, len(a)); return y


def _ulp(y, ref):
    sp = np.spacing(np.abs(ref.astype(np.float32))).astype(np.float64)
    return float((np.abs(y.astype(np.float64) - ref) / sp).max())


def test_accuracy(lib):
    rng = np.random.default_rng(0)
    n = 400_000
    x = rng.uniform(-7, 7, n).astype(np.float32)
    assert np.abs(_run1(lib, "v_sin", x) - np.sin(x.astype(np.float64))).max() < 1.2e-7
    assert np.abs(_run1(lib, "v_cos", x) - np.cos(x.astype(np.float64))).max() < 1.2e-7
    x = rng.uniform(-1, 1, n).astype(np.float32)
    assert _ulp(_run1(lib, "v_asin", x), np.arcsin(x.astype(np.float64))) < 3.0
    x = np.exp(rng.uniform(-20, 5, n)).astype(np.float32)
    assert _ulp(_run1(lib, "v_log", x), np.log(x.astype(np.float64))) < 1.5
    x = rng.uniform(-20, 20, n).astype(np.float32)
    assert _ulp(_run1(lib, "v_exp", x), np.exp(x.astype(np.float64))) < 1.5
    a, b = rng.normal(size=n).astype(np.float32), rng.normal(size=n).astype(np.float32)
    assert _ulp(_run2(lib, "v_atan2", a, b), np.arctan2(a.astype(np.float64), b.astype(np.float64))) < 4.0
    a, b = rng.uniform(1e-6, 1e-2, n).astype(np.float32), rng.uniform(0, 1, n).astype(np.float32)
    assert _ulp(_run2(lib, "v_pow", a, b), np.power(a.astype(np.float64), b.astype(np.float64))) < 20.0


def test_special_values(lib):
    x = np.array([0.0, -0.0, 1.0, -1.0, 0.5, 1.0000001], np.float32)
    y = _run1(lib, "v_asin", x)
    assert y[0] == 0.0 and abs(y[2] - np.pi / 2) < 2e-7 and abs(y[3] + np.pi / 2) < 2e-7 and np.isfinite(y).all()
    a = np.array([0.0, 1.0, -1.0, 0.0, 0.0], np.float32); b = np.array([1.0, 0.0, 0.0, -1.0, 0.0], np.float32)
    y = _run2(lib, "v_atan2", a, b)
    assert y[0] == 0.0 and abs(y[1] - np.pi / 2) < 2e-7 and abs(y[2] + np.pi / 2) < 2e-7 and abs(y[3] - np.pi) < 3e-7 and y[4] == 0.0


def test_known_answer_hash(lib):
    """The exact bits of the parity arithmetic's transcendental functions on a fixed grid."""
    x = np.linspace(-6.5, 6.5, 4001, dtype=np.float32)
    u = np.linspace(-1, 1, 2001, dtype=np.float32)
    got = sha(np.concatenate([_run1(lib, "v_sin", x), _run1(lib, "v_cos", x), _run1(lib, "v_asin", u),
                              _run1(lib, "v_log", np.linspace(1e-6, 2.0, 2001, dtype=np.float32)),
                              _run2(lib, "v_atan2", u, u[::-1].copy() + np.float32(0.25))]))
    path = os.path.join(ROOT, "tests", "golden", "ptmath_sha256.txt")
    if not os.path.exists(path):
        pytest.skip("tests/golden/ptmath_sha256.txt missing; run tests/golden/make_golden.py")
    assert got == open(path).read().strip()
