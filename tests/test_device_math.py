"""Accuracy of the E-step's scalar math (open-network-insight_b200/csrc/device_math.cuh).
The header compiles for the host too, so its algorithms are checked here on the
CPU against libm and against the oracle's digamma; the GPU tests then cover the
device build through the parity of whole E-steps."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


@pytest.fixture(scope="module")
def dm(tmp_path_factory):
    out = tmp_path_factory.mktemp("dm") / "libdm.so"
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                    "-I", str(ROOT / "open-network-insight_b200" / "csrc"),
                    "-I", str(ROOT / "include"),
                    str(ROOT / "tests" / "device_math_host.cpp"),
                    str(ROOT / "open-network-insight_b200" / "csrc" / "hostmath.c"), "-o", str(out)], check=True)
    L = C.CDLL(str(out))
    for f in (L.dm_rcp, L.dm_log, L.dm_log_tab, L.dm_exp, L.dm_digamma, L.dm_exp_tab, L.dm_exp_tab_shallow,
              L.dm_psi_tab, L.dm_psi_tab_shallow, L.dm_epsi_tab):
        f.argtypes = [_f64p, _f64p, C.c_long]

    def call(fn, x):
        x = np.ascontiguousarray(x, np.float64)
        y = np.empty_like(x)
        fn(x, y, len(x))
        return y
    L.call = call
    return L


def ulps(a, b):
    return np.abs(a - b) / np.spacing(np.abs(b))


def test_rcp(dm):
    rng = np.random.default_rng(0)
    x = np.exp(rng.uniform(-600, 600, 200000))
    assert ulps(dm.call(dm.dm_rcp, x), 1.0 / x).max() <= 1.0


def test_log(dm):
    rng = np.random.default_rng(1)
    x = np.concatenate([np.exp(rng.uniform(-700, 700, 200000)), rng.uniform(0.5, 2.0, 200000),
                        6 + np.exp(rng.uniform(-30, 15, 200000)), [1.0, 2.0, 0.5, np.sqrt(2), 1 / np.sqrt(2), 6.0]])
    got, want = dm.call(dm.dm_log, x), np.log(x)
    err = np.abs(got - want)
    assert np.all(err <= 2.5 * np.spacing(np.abs(want)) + 2.3e-16)       # <= 2 ulp, or ~1 ulp of 1 near log = 0
    assert dm.call(dm.dm_log, np.array([1.0]))[0] == 0.0


def test_log_by_table(dm):
    rng = np.random.default_rng(4)
    x = np.concatenate([np.exp(rng.uniform(-700, 700, 200000)), rng.uniform(0.5, 2.0, 200000),
                        6 + np.exp(rng.uniform(-30, 15, 200000)), 1 + np.arange(129) / 128.0,
                        np.nextafter(1 + np.arange(1, 129) / 128.0, 0), [1.0, 2.0, 0.5, 6.0, 1e-300, 1e300]])
    got, want = dm.call(dm.dm_log_tab, x), np.log(x)
    err = np.abs(got - want)
    assert np.all(err <= 2.5 * np.spacing(np.abs(want)) + 4.5e-16)       # <= 2 ulp, or ~2 ulp of 1 near log = 0


def test_exp(dm):
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.uniform(-700, 700, 300000), rng.uniform(-1, 1, 100000), [0.0, -707.0, 1e-300]])
    assert ulps(dm.call(dm.dm_exp, x), np.exp(x)).max() <= 2.0
    tiny = dm.call(dm.dm_exp, np.array([-1000.0, -1e6, -708.0]))
    assert np.all(tiny >= 0) and np.all(tiny < 1e-306)                   # far below anything that matters, never NaN


def test_digamma_against_reference_series(dm, oracle):
    rng = np.random.default_rng(3)
    x = np.concatenate([np.exp(rng.uniform(-25, 14, 300000)), [0.05, 0.1, 1.0, 2.5, 100.0, 1e-12]])
    want = np.array([oracle.lib().oracle_digamma(v) for v in x[:20000]])
    got = dm.call(dm.dm_digamma, x[:20000])
    assert np.max(np.abs(got - want) / np.maximum(np.abs(want), 1.0)) < 4e-15


def test_exp_by_table(dm):
    """fast_exp_tab (both polynomial forms): the E-step's exp(psi(gamma))."""
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(-700, 700, 300000), rng.uniform(-1, 1, 100000), rng.uniform(-40, 14, 300000),
                        [0.0, -707.0, 1e-300]])
    for fn in (dm.dm_exp_tab, dm.dm_exp_tab_shallow):
        # 2 ulp plus the one-constant range reduction: a relative error of |x| * 1.2e-16, i.e. half an ulp of x
        got, want = dm.call(fn, x), np.exp(x)
        assert np.all(np.abs(got - want) <= want * (4.5e-16 + 1.2e-16 * np.abs(x)))
        tiny = dm.call(fn, np.array([-1000.0, -1e6, -708.0, -0.99e9]))
        assert np.all(tiny >= 0) and np.all(tiny < 1e-306)


def test_psi_by_table_against_reference_series(dm, oracle):
    """psi = R(gamma) - 1/gamma with R from the piecewise degree-5 rows (hostmath.c) against the
    reference's series (utils.c:55-64) over the table's whole range, cell edges included."""
    rng = np.random.default_rng(6)
    edges = 1.0 + np.arange(1, 64 * 6) / 64.0 - 1.0            # gamma with gamma + 1 on cell edges of the first octaves
    x = np.concatenate([np.exp(rng.uniform(np.log(1e-8), np.log(2.0 ** 20 - 2), 30000)), rng.uniform(0, 2, 10000),
                        edges[edges > 0], np.nextafter(edges[edges > 0], 0), [1e-8, 0.05, 0.1, 1.0, 2.5, 100.0, 2.0 ** 20 - 1.5]])
    want = np.array([oracle.lib().oracle_digamma(v) for v in x])
    for fn in (dm.dm_psi_tab, dm.dm_psi_tab_shallow):
        got = dm.call(fn, x)
        # the reference forms its pole term from the rounded shifted argument, 1/((x+6)-6) (utils.c:62); the
        # table form has the exact 1/x: compare with that one term exchanged
        adj = want + 1.0 / ((x + 6.0) - 6.0) - 1.0 / x
        assert np.max(np.abs(got - adj) / np.maximum(np.abs(adj), 1.0)) < 1e-14
    e = dm.call(dm.dm_epsi_tab, x)
    ok = want > -600
    adj = want + 1.0 / ((x + 6.0) - 6.0) - 1.0 / x
    assert np.max(np.abs(e[ok] - np.exp(adj[ok])) / np.exp(adj[ok])) < 2e-13
