"""Shared helpers of the parity tests: build the same scene in the CUDA
library (through the C ABI) and in the CPU oracle."""
import numpy as np

import xrtailor_b200 as X
from oracle import oracle as O
from xrtailor_b200 import scene


def params(**kw):
    p = X.default_params()
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def pair(sc, bindings=None, **kw):
    """(cuda solver, oracle scene) on identical inputs and colours."""
    kw.setdefault("particle_diameter", sc.particle_diameter())
    kw.setdefault("morton_fast_div", 0)  # IEEE division: bit parity with the CPU oracle
    cs, _ = X.greedy_color(sc.stretch_idx, len(sc.cloth_pos))
    cb, _ = X.greedy_color(sc.bend_idx, len(sc.cloth_pos))
    s = scene.make_solver(sc, colors=(cs, cb), bindings=bindings, **kw)
    o = O.Scene(sc, s.params, (cs, cb))
    if bindings is not None:
        o.set_binding(*bindings)
    return s, o


def rel_rms(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.sqrt(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-300)))


def rms(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.sqrt(((a - b) ** 2).sum(-1).mean()))


def assert_bvh_equal(g, o):
    np.testing.assert_array_equal(g["morton"], o["morton"])
    np.testing.assert_array_equal(g["object"], o["object"])
    np.testing.assert_array_equal(g["nodes"], o["nodes"])
    np.testing.assert_array_equal(g["esc_int"], o["esc_int"])
    np.testing.assert_array_equal(g["esc_ext"], o["esc_ext"])
    np.testing.assert_array_equal(g["aabb"], o["aabb"])  # min/max only: bit exact


def stretch_stats(sc, x):
    """RMS / max of (|xi-xj|-rest)/rest, the statistic xrt_stats reports."""
    x = np.asarray(x, np.float64)
    d = np.linalg.norm(x[sc.stretch_idx[:, 0]] - x[sc.stretch_idx[:, 1]], axis=1)
    e = (d - sc.stretch_rest) / sc.stretch_rest
    return float(np.sqrt((e * e).mean())), float(np.abs(e).max())
