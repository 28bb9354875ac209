"""
Cross-check of the golden fixtures with outputs of the REAL reference (TensorFlow markovflow), when somebody has
produced them with baseline/run_tf_reference.py on a box that has the reference's environment (tests/golden_tf/
*_tf.npz).  Not runnable in the build image (no TensorFlow): skipped while the directory is empty.  When the files
exist this is what turns "parity pinned to the oracle" into "parity pinned to the reference": the oracle's fixtures
(tests/golden/*.npz, which the CUDA path reproduces to 1e-9 in tests/test_golden.py) must equal the reference's.
"""
import glob
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
TF_FILES = sorted(glob.glob(os.path.join(HERE, "golden_tf", "*_tf.npz")))
TOL = 1e-9

PAIRS = [("cvi_elbo", None), ("cvi_m", None), ("cvi_S", None), ("cvi_d1", None), ("cvi_d2", None),
         ("cvi_g1", 0), ("cvi_gd", 1), ("cvi_gs", 2),
         ("vgp_elbo", None), ("vgp_A", None), ("vgp_b", None), ("vgp_lam", None), ("vgp_psi", None), ("vgp_m", None),
         ("vgp_S", None), ("vgp_q_mean", None), ("vgp_q_chol", None)]


@pytest.mark.skipif(not TF_FILES, reason="no reference dumps (baseline/run_tf_reference.py needs TensorFlow 2.4)")
@pytest.mark.parametrize("path", TF_FILES, ids=[os.path.basename(p)[:-7] for p in TF_FILES])
def test_oracle_fixtures_equal_the_reference_dump(path):
    tf = np.load(path)
    g = np.load(os.path.join(HERE, "golden", os.path.basename(path).replace("_tf.npz", ".npz")))
    floor = g["site_floor_f64_vs_f80"]
    for key, site in PAIRS:
        want, got = np.asarray(tf["tf_" + key]).reshape(g[key].shape), g[key]
        err = np.abs(got - want).max() / max(np.abs(want).max(), 1e-300)
        bar = TOL if site is None else max(TOL, 3.0 * (floor[site] if floor[site] < 1e-6 else 0.0))
        if site is not None and np.abs(want).max() < 1e-6:      # OU prior: the exact sites are 0
            continue
        assert err < bar, (key, err)
