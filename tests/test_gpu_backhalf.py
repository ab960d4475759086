"""GPU: the consensus back half through the cNMF mirror (usage refit, reordering by usage, TPM-unit spectra, z-score
spectra, final usage refit; cnmf.py:1092-1160) against files written by the REFERENCE's own cNMF.consensus
(tests/golden/consensus_backhalf.npz, made by tests/golden/make_golden.py with torch seeded so that the reference's
unseeded torch.rand inits are reproducible)."""
import os

import numpy as np
import pandas as pd
import pytest

pytestmark = pytest.mark.gpu


def _pearson(a, b):
    return float(np.corrcoef(np.asarray(a, dtype=np.float64).ravel(), np.asarray(b, dtype=np.float64).ravel())[0, 1])


@pytest.mark.parametrize("tag,flags", [("refit", dict(refit_usage=True)),
                                       ("norefit", dict(refit_usage=False, normalize_tpm_spectra=True))])
def test_consensus_back_half_matches_reference_files(tmp_path, golden_dir, tag, flags):
    from cnmf_torch_b200 import cnmf as host
    g = np.load(os.path.join(golden_dir, "consensus_backhalf.npz"), allow_pickle=True)
    k = int(g["k"])
    hvgs, genes_all, cells = list(g["hvgs"]), list(g["genes_all"]), list(g["cells"])
    obj = host.cNMF(output_dir=str(tmp_path), name="back")
    obj.prepare_from_matrix(g["norm"], [k], n_iter=g["spectra"].shape[0] // k, seed=14, beta_loss="frobenius",
                            obs_names=cells, var_names=hvgs, batch_size=int(g["batch_size"]), tpm=g["tpm"],
                            tpm_var_names=genes_all)
    idx = [f"iter{i // k}_topic{i % k + 1}" for i in range(g["spectra"].shape[0])]
    host.save_df_to_npz(pd.DataFrame(g["spectra"], index=idx, columns=hvgs), obj.paths["merged_spectra"] % k)
    obj.consensus(k, density_threshold=0.5, local_neighborhood_size=0.30, seed=int(g["torch_seed"]), **flags)
    for key, rtol in (("consensus_spectra", 1e-5), ("consensus_usages", 2e-3), ("gene_spectra_tpm", 2e-3),
                      ("gene_spectra_score", 2e-3)):
        df = host.load_df_from_npz(obj.paths[key] % (k, "0_5"))
        ref = g[f"{tag}_{key}"]
        assert df.shape == ref.shape, key
        assert [str(i) for i in df.index] == list(g[f"{tag}_{key}_index"]), key          # same program order, same labels
        assert [str(c) for c in df.columns] == list(g[f"{tag}_{key}_columns"]), key
        scale = np.abs(ref).max()
        np.testing.assert_allclose(df.values, ref, rtol=rtol, atol=rtol * scale, err_msg=key)
        assert _pearson(df.values, ref) >= 0.9999, key
        assert os.path.exists(obj.paths[key + "__txt"] % (k, "0_5"))
    usage, scores, tpm_sp, top = obj.load_results(k, 0.5, n_top_genes=10)          # cnmf.py:1335-1385
    assert usage.shape == (len(cells), k) and np.allclose(usage.sum(axis=1), 1.0) and list(usage.columns) == [1, 2, 3, 4]
    assert scores.shape == (len(genes_all), k) and tpm_sp.shape == (len(genes_all), k) and top.shape == (10, k)
    assert top.iloc[0, 0] == scores.iloc[:, 0].idxmax()


def test_consensus_without_tpm_writes_spectra_and_usages_only(tmp_path, golden_dir):
    from cnmf_torch_b200 import cnmf as host
    g = np.load(os.path.join(golden_dir, "consensus_backhalf.npz"), allow_pickle=True)
    k = int(g["k"])
    obj = host.cNMF(output_dir=str(tmp_path), name="notpm")
    obj.prepare_from_matrix(g["norm"], [k], n_iter=g["spectra"].shape[0] // k, seed=14, batch_size=100)
    host.save_df_to_npz(pd.DataFrame(g["spectra"]), obj.paths["merged_spectra"] % k)
    med, labels, dens = obj.consensus(k, density_threshold=0.5, seed=1)
    us = host.load_df_from_npz(obj.paths["consensus_usages"] % (k, "0_5"))
    assert us.shape == (g["norm"].shape[0], k) and list(us.columns) == list(range(1, k + 1))
    tot = us.div(us.sum(axis=1), axis=0).sum(axis=0).values
    assert np.all(np.diff(tot) <= 1e-6)                                   # programs ordered by total usage
    assert not os.path.exists(obj.paths["gene_spectra_tpm"] % (k, "0_5"))
    a = obj.consensus(k, density_threshold=0.5, seed=1)[0]
    assert np.array_equal(a.values, med.values)                           # seeded: reproducible
