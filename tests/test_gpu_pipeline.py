"""GPU: the file contract end to end -- prepare_from_matrix -> factorize (replicate-batched, all K at once)
-> combine -> consensus, through the cNMF object and through the CLI verbs, in the reference's default
online mode and in batch mode.  Checks the files the reference writes (names, index/columns, dtypes) and
that a rerun with skip_completed_runs only fills the gaps."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def counts():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(400, 150, k_true=4, seed=14, libsize=500.0)


@pytest.mark.parametrize("mode", ["online", "batch"])
def test_factorize_combine_consensus_files(tmp_path, counts, mode):
    from cnmf_torch_b200 import cnmf as host
    genes = [f"gene{i}" for i in range(counts.shape[1])]
    obj = host.cNMF(output_dir=str(tmp_path), name="run")
    rp, kw = obj.prepare_from_matrix(counts, [3, 4, 5], n_iter=6, seed=14, beta_loss="frobenius", var_names=genes)
    assert kw["mode"] == "online" and kw["algo"] == "mu" and kw["online_chunk_size"] == 5000     # cnmf.py:757-771
    override = None if mode == "online" else dict(mode="batch", batch_max_iter=80)
    obj.factorize(nmf_kwargs_override=override)
    for k in (3, 4, 5):
        for it in range(6):
            df = host.load_df_from_npz(obj.paths["iter_spectra"] % (k, it))
            assert df.shape == (k, len(genes)) and list(df.index) == list(range(1, k + 1))
            assert list(df.columns) == genes and df.values.dtype == np.float32 and np.isfinite(df.values).all()
            assert (df.values >= 0).all() and df.values.sum() > 0
    obj.combine()
    m = host.load_df_from_npz(obj.paths["merged_spectra"] % 4)
    assert m.shape == (24, len(genes)) and m.index[0] == "iter0_topic1" and m.index[-1] == "iter5_topic4"
    med, labels, dens = obj.consensus(4, density_threshold=2.0, local_neighborhood_size=0.3)
    assert med.shape == (4, len(genes)) and np.allclose(med.sum(axis=1), 1.0, atol=1e-5)
    assert set(labels.unique()) <= {1, 2, 3, 4} and len(labels) == 24 and dens.shape == (24, 1)
    for key in ("local_density_cache",):
        assert os.path.exists(obj.paths[key] % 4)
    assert os.path.exists(obj.paths["consensus_spectra"] % (4, "2_0"))          # dt string rule, cnmf.py:1049-1052
    assert os.path.exists(obj.paths["consensus_spectra__txt"] % (4, "2_0"))
    # resume: delete two replicate files, mark the ledger, rerun only the gaps
    os.remove(obj.paths["iter_spectra"] % (3, 1)); os.remove(obj.paths["iter_spectra"] % (5, 4))
    stamp = os.path.getmtime(obj.paths["iter_spectra"] % (4, 0))
    obj.update_nmf_iter_params()
    obj.factorize(skip_completed_runs=True, nmf_kwargs_override=override)
    assert os.path.exists(obj.paths["iter_spectra"] % (3, 1)) and os.path.exists(obj.paths["iter_spectra"] % (5, 4))
    assert os.path.getmtime(obj.paths["iter_spectra"] % (4, 0)) == stamp          # untouched


def test_cli_verbs(tmp_path, counts):
    from cnmf_torch_b200 import cnmf as host
    obj = host.cNMF(output_dir=str(tmp_path), name="cli")
    obj.prepare_from_matrix(counts, [3], n_iter=4, seed=1, beta_loss="frobenius")
    env = dict(os.environ, PYTHONPATH=ROOT)
    base = [sys.executable, "-m", "cnmf_torch_b200.cnmf"]
    common = ["--output-dir", str(tmp_path), "--name", "cli"]
    for w in (0, 1):   # two workers, the reference's stride rule
        subprocess.run(base + ["factorize"] + common + ["--worker-index", str(w), "--total-workers", "2"], check=True, env=env)
    subprocess.run(base + ["combine"] + common, check=True, env=env)
    subprocess.run(base + ["consensus"] + common + ["-k", "3", "--local-density-threshold", "2.0"], check=True, env=env)
    assert os.path.exists(obj.paths["merged_spectra"] % 3)
    assert os.path.exists(obj.paths["consensus_spectra__txt"] % (3, "2_0"))


@pytest.mark.parametrize("beta_loss", ["frobenius", "kullback-leibler"])
def test_factorize_writes_merged_spectra_identical_to_combine(tmp_path, counts, beta_loss):
    """a single worker that ran every replicate writes each K's merged file itself; it must be the
    file combine_nmf builds from the per-replicate files (cnmf.py:895-920), and combine must rebuild it as soon as a
    replicate file is newer."""
    from cnmf_torch_b200 import cnmf as host
    obj = host.cNMF(output_dir=str(tmp_path), name="m")
    obj.prepare_from_matrix(counts, [3, 4], n_iter=5, seed=14, beta_loss=beta_loss)
    obj.factorize(nmf_kwargs_override=dict(mode="batch", batch_max_iter=30))
    direct = {k: host.load_df_from_npz(obj.paths["merged_spectra"] % k) for k in (3, 4)}
    stamp = os.path.getmtime(obj.paths["merged_spectra"] % 3)
    obj.combine()                                                  # up to date: untouched
    assert os.path.getmtime(obj.paths["merged_spectra"] % 3) == stamp
    for k in (3, 4):
        os.remove(obj.paths["merged_spectra"] % k)
    obj.combine()                                                  # rebuilt from the replicate files
    for k in (3, 4):
        rebuilt = host.load_df_from_npz(obj.paths["merged_spectra"] % k)
        assert list(rebuilt.index) == list(direct[k].index) and list(rebuilt.columns) == list(direct[k].columns)
        assert np.array_equal(rebuilt.values, direct[k].values) and rebuilt.values.dtype == direct[k].values.dtype
    # a replicate re-run later makes the merged file stale
    os.remove(obj.paths["iter_spectra"] % (3, 2))
    obj.update_nmf_iter_params()
    obj.factorize(skip_completed_runs=True, nmf_kwargs_override=dict(mode="batch", batch_max_iter=30))
    assert not obj._merged_is_current(3, host.load_df_from_npz(obj.paths["nmf_replicate_parameters"]).query("n_components == 3"))
    obj.combine(components=3)
    assert obj._merged_is_current(3, host.load_df_from_npz(obj.paths["nmf_replicate_parameters"]).query("n_components == 3"))
