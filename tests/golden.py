"""Loader for the fixtures in tests/golden/ (made by oracle/gen_golden.py from the reference)."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ALL_D = (3, 4, 5, 6, 7, 9, 15, 25)


def load(name):
    return np.load(os.path.join(GOLDEN_DIR, name))


def unpack_views(bits, total, d):
    """packed bits -> (total, 2, d, d) uint8"""
    n = int(total) * 2 * d * d
    return np.unpackbits(np.asarray(bits))[:n].reshape(int(total), 2, d, d)


def per_d(d):
    g = load("golden_d%d.npz" % d)
    out = {k: g[k] for k in g.files}
    out["views"] = unpack_views(g["persp_bits"], g["persp_total"], d)
    out["views_last"] = unpack_views(g["persp_bits_last"], g["persp_total_last"], d)
    return out


def exhaustive(d):
    g = load("golden_exhaustive.npz")
    return {"qubits": g["q%d" % d], "state": g["s%d" % d], "gs": g["gs%d" % d],
            "counts": g["counts%d" % d], "positions": g["pos%d" % d],
            "views": unpack_views(g["bits%d" % d], g["total%d" % d], d)}


def batch_d7():
    g = load("golden_batch_d7.npz")
    out = {k: g[k] for k in g.files}
    out["views"] = unpack_views(g["persp_bits"], g["persp_total"], 7)
    return out
