"""Shared helpers for the parity tests (test infrastructure)."""
import glob
import os

import numpy as np

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases(prefix):
    return sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def system_and_grid(d):
    s = O.System.from_dump(d)
    g = O.grid([0, 0, 0], d["box"], float(d["rc"][0]), int(d["meta"][1]))
    return s, g


def vec_rel_err(a, b):
    """max_i |a_i - b_i| / max(|b_i|, 1e-3*max_j|b_j|): relative error of per-molecule vectors, as the
    reference's LinkedCellsTest.cpp:620-633 does (relative on the vector norm, not per component)."""
    a = np.asarray(a, dtype=np.float64).reshape(len(b), -1)
    b = np.asarray(b, dtype=np.float64).reshape(len(b), -1)
    nb = np.linalg.norm(b, axis=1)
    scale = nb.max() if len(nb) else 0.0
    if scale == 0.0:
        return float(np.abs(a).max()) if a.size else 0.0
    return float((np.linalg.norm(a - b, axis=1) / np.maximum(nb, 1e-3 * scale)).max())


def initial_state(d):
    """owned molecules after the t=0 force pass, sorted by id: the state the time loop starts from"""
    own = d["f_owned"].astype(bool)
    ids = d["f_id"][own]
    order = np.argsort(ids)
    inorder = np.argsort(d["in_id"])
    assert np.array_equal(ids[order], d["in_id"][inorder])
    return dict(id=ids[order].astype(np.uint64), cid=d["in_cid"][inorder].astype(np.uint32),
                r=d["f_r"][own][order], v=d["in_v"][inorder], q=d["f_q"][own][order], L=d["in_L"][inorder],
                F=d["f_F"][own][order], M=d["f_M"][own][order])
