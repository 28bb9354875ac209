"""On-disk formats: the data-set schema of generate_data.py:128-140 / exp_dp_utils.py:108-124 and
the checkpoint schemas of cvi_dp.py:140-149 / vi_markov_gp.py:171-174."""
import io
import os
import zipfile

import numpy as np
import pytest

from vi_diffusion_processes_b200 import io as vio

REF_ZIP = "/root/reference/docs/diffusion_processes/data.zip"


def test_dataset_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    grid = np.round(np.arange(0, 1.01, 0.01), 2)
    path = tmp_path / "7.npz"
    vio.save_exp_data(path, "dw", Q=np.ones((1, 1)), x0=np.ones((1, 1)), sigma=0.1,
                      latent_process=rng.normal(size=(101, 1)), observations=rng.normal(size=(8, 1)),
                      observation_grid=grid[::13][:8], time_grid=grid, test_observations=rng.normal(size=(2, 1)),
                      test_grid=grid[5::40][:2])
    assert set(np.load(path).keys()) == set(vio.DATASET_KEYS)
    Q, x0, noise, latent, obs, tg, test = vio.load_exp_data(str(path))
    assert noise.shape == (1, 1) and noise[0, 0] == 0.1
    assert latent.shape == (101, 1) and obs[0].shape == (8,) and obs[1].shape == (8, 1)
    assert test[0].shape == (2,) and tg.shape == (101,) and Q.shape == (1, 1)


@pytest.mark.skipif(not os.path.exists(REF_ZIP), reason="the reference's shipped data sets are not on this machine")
@pytest.mark.parametrize("member", ["data/linear/15/0.npz", "data/dw/3/0.npz", "data/vanderpol/33/0.npz"])
def test_reads_the_reference_data_files(member, tmp_path):
    with zipfile.ZipFile(REF_ZIP) as z:
        names = [n for n in z.namelist() if n.endswith(member)]
        assert names, member
        raw = z.read(names[0])
    assert set(np.load(io.BytesIO(raw)).keys()) == set(vio.DATASET_KEYS)
    p = tmp_path / "d.npz"
    p.write_bytes(raw)
    Q, x0, noise, latent, obs, tg, test = vio.load_exp_data(str(p))
    D = latent.shape[-1]          # the 1-D files store Q and x0 as scalars
    assert np.size(Q) == D * D and obs[1].shape == (obs[0].shape[0], D) and test[1].shape[0] == test[0].shape[0]
    # every observation sits on a grid point (variational_cvi_sde.py:117)
    assert np.isin(obs[0], tg).all() and np.isin(test[0], tg).all()


def test_checkpoint_key_sets():
    assert vio.CVI_MODEL_KEYS == ("data_sites_nat1", "data_sites_nat2", "girsanov_sites_nat1",
                                  "girsanov_sites_nat2_diag", "girsanov_sites_nat2_subdiag")
    assert vio.VI_GP_MODEL_KEYS == ("A", "b", "lambda_lagrange", "psi_lagrange", "x0_m", "x0_S")
