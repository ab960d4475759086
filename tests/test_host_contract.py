"""CPU: the host-side mirror of the reference interface against goldens produced by the reference's
own functions (ledger seeds, worker striding, .df.npz format, combine)."""
import io
import os

import numpy as np
import pandas as pd
import pytest

from cnmf_torch_b200 import cnmf as host


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=True)


@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_ledger_matches_reference(golden_dir, tmp_path, tag):
    g = _load(golden_dir, "ledger.npz")
    n_iter, seed = [int(v) for v in g[tag + "_args"]]
    ks = [int(v) for v in g[tag + "_ks"]]
    obj = host.cNMF(output_dir=str(tmp_path), name="t")
    rp, kw = obj.get_nmf_iter_params(ks if tag != "e" else ks[0], n_iter=n_iter, random_state_seed=seed,
                                     beta_loss="frobenius")
    np.testing.assert_array_equal(rp[["n_components", "iter", "nmf_seed"]].values.astype(np.int64),
                                  g[tag + "_values"])
    np.testing.assert_array_equal(rp["completed"].values.astype(bool), g[tag + "_completed"])
    assert sorted(kw.keys()) == list(g["kwargs_keys"])
    assert repr(sorted(kw.items())) == str(g["kwargs_repr"])


def test_ledger_known_answers():
    """KATs quoted in SURVEY.md 8(a1), computed with the reference's code, seed 14."""
    from cnmf_torch_b200.synth import ledger_seeds
    s = ledger_seeds(45, 14)
    assert list(s[:3]) == [59886188, 1812018521, 1173234957] and s[-1] == 1288885239
    assert ledger_seeds(900, 14)[-1] == 1393000


def test_worker_filter_matches_reference(golden_dir):
    g = _load(golden_dir, "worker_filter.npz")
    for key in g.files:
        n, w, t = [int(v) for v in key.split("_")]
        np.testing.assert_array_equal(np.array(list(host.worker_filter(range(n), w, t)), dtype=np.int64), g[key])


def test_df_npz_format_is_interchangeable(golden_dir, tmp_path):
    g = _load(golden_dir, "combine.npz")
    # a file written by the reference's save_df_to_npz loads with our loader ...
    ref_file = tmp_path / "ref.df.npz"
    ref_file.write_bytes(g["iter_file_bytes"].tobytes())
    df = host.load_df_from_npz(str(ref_file))
    np.testing.assert_array_equal(df.values, g["k3_i0"])
    assert list(df.index) == [1, 2, 3] and list(df.columns) == [f"g{i}" for i in range(12)]
    # ... and a file we write has the same members, dtypes and contents
    ours = tmp_path / "ours.df.npz"
    host.save_df_to_npz(df, str(ours))
    with np.load(str(ref_file), allow_pickle=True) as a, np.load(str(ours), allow_pickle=True) as b:
        assert sorted(a.files) == sorted(b.files) == ["columns", "data", "index"]
        for k in a.files:
            assert a[k].dtype == b[k].dtype
            np.testing.assert_array_equal(a[k], b[k])


def test_combine_matches_reference(golden_dir, tmp_path):
    g = _load(golden_dir, "combine.npz")
    obj = host.cNMF(output_dir=str(tmp_path), name="comb")
    rp, kw = obj.get_nmf_iter_params([3, 4], n_iter=3, random_state_seed=14, beta_loss="frobenius")
    obj.save_nmf_iter_params(rp, kw)
    genes = [f"g{i}" for i in range(12)]
    for kk in (3, 4):
        for it in range(3):
            df = pd.DataFrame(g[f"k{kk}_i{it}"], index=np.arange(1, kk + 1), columns=genes)
            host.save_df_to_npz(df, obj.paths["iter_spectra"] % (kk, it))
    for kk in (3, 4):
        m = obj.combine_nmf(kk)
        np.testing.assert_array_equal(m.values, g[f"merged_k{kk}_data"])
        assert list(m.index) == list(g[f"merged_k{kk}_index"])
        assert list(m.columns) == list(g[f"merged_k{kk}_columns"])
        m2 = host.load_df_from_npz(obj.paths["merged_spectra"] % kk)
        np.testing.assert_array_equal(m2.values, m.values)
    # missing file raises unless skip_missing_files (cnmf.py:904-909)
    os.remove(obj.paths["iter_spectra"] % (3, 1))
    with pytest.raises(FileNotFoundError):
        obj.combine_nmf(3)
    assert obj.combine_nmf(3, skip_missing_files=True).shape[0] == 6
    # update_nmf_iter_params marks completion by file existence (cnmf.py:780-795)
    obj.update_nmf_iter_params()
    rp2 = host.load_df_from_npz(obj.paths["nmf_replicate_parameters"])
    done = rp2[(rp2.n_components == 3)].sort_values("iter")["completed"].tolist()
    assert done == [True, False, True]


def test_prepare_from_matrix_rejects_zero_count_cells(tmp_path):
    obj = host.cNMF(output_dir=str(tmp_path), name="z")
    x = np.ones((5, 4), dtype=np.float32)
    x[2] = 0
    with pytest.raises(Exception, match="cells have zero counts of overdispersed genes"):
        obj.prepare_from_matrix(x, [2], n_iter=2, seed=1)


def test_norm_counts_side_channel_roundtrip(tmp_path):
    obj = host.cNMF(output_dir=str(tmp_path), name="n")
    rng = np.random.default_rng(0)
    x = (rng.random((7, 5)) * (rng.random((7, 5)) < 0.5)).astype(np.float32)
    x[:, 0] += 1
    obj.save_norm_counts(x, var_names=[f"g{i}" for i in range(5)])
    X, obs, var = obj.load_norm_counts()
    np.testing.assert_array_equal(X.toarray(), x)
    assert list(var) == [f"g{i}" for i in range(5)]


def test_product_does_not_import_the_oracle():
    """No file of the product package may reference oracle/ (the oracle is the checker only)."""
    import cnmf_torch_b200
    root = os.path.dirname(cnmf_torch_b200.__file__)
    for dirpath, _, files in os.walk(root):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "import oracle" not in text and "from oracle" not in text, fn


def test_tpm_side_channel_and_stats(tmp_path):
    """save_tpm writes what prepare leaves for the consensus back half: the TPM matrix and per-gene mean / population
    std (cnmf.py:570-580: StandardScaler var ** .5 for sparse input, std(ddof=0) for dense)."""
    import scipy.sparse as sp
    from sklearn.preprocessing import StandardScaler
    from cnmf_torch_b200 import cnmf as host
    rng = np.random.default_rng(3)
    tpm = (rng.gamma(0.5, 50.0, size=(40, 17)) * (rng.random((40, 17)) < 0.4))
    tpm[:, 5] = 0.0
    obj = host.cNMF(output_dir=str(tmp_path), name="t")
    obj.save_tpm(tpm, obs_names=[f"c{i}" for i in range(40)], var_names=[f"g{i}" for i in range(17)])
    T, obs, var = obj.load_tpm()
    assert np.array_equal(T.toarray(), tpm) and list(var) == [f"g{i}" for i in range(17)] and obs[3] == "c3"
    stats = host.load_df_from_npz(obj.paths["tpm_stats"])
    sc = StandardScaler(with_mean=False).fit(sp.csr_matrix(tpm))
    assert list(stats.columns) == ["__mean", "__std"] and list(stats.index) == list(var)
    np.testing.assert_allclose(stats["__mean"].values.astype(float), sc.mean_, rtol=1e-12)
    np.testing.assert_allclose(stats["__std"].values.astype(float), sc.var_ ** .5, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(stats["__std"].values.astype(float), tpm.std(axis=0, ddof=0), rtol=1e-9, atol=1e-12)
    assert host.cNMF(output_dir=str(tmp_path), name="none").load_tpm() is None


def test_cli_accepts_every_flag_the_reference_declares(tmp_path, capsys):
    """cnmf.py:1410-1435: a command line written for the reference must parse here.  ``prepare`` is outside this
    path and says so; ``combine`` runs on the CPU and is exercised end to end."""
    obj = host.cNMF(output_dir=str(tmp_path), name="cli")
    rp, kw = obj.get_nmf_iter_params([2], n_iter=2, random_state_seed=1, beta_loss="frobenius")
    obj.save_nmf_iter_params(rp, kw)
    for it in range(2):
        host.save_df_to_npz(pd.DataFrame(np.full((2, 3), it + 1.0, dtype=np.float32), index=[1, 2],
                                         columns=["a", "b", "c"]), obj.paths["iter_spectra"] % (2, it))
    every_flag = ["--name", "cli", "--output-dir", str(tmp_path), "-c", "counts.h5ad", "-k", "2", "-n", "2",
                  "--total-workers", "1", "--use_gpu", "--seed", "3", "--genes-file", "g.txt", "--numgenes", "500",
                  "--tpm", "tpm.df.npz", "--max-nmf-iter", "200", "--beta-loss", "itakura-saito", "--init", "nndsvd",
                  "--densify", "--batch_size", "1000", "--skip-completed-runs", "--local-density-threshold", "0.3",
                  "--local-neighborhood-size", "0.2", "--show-clustering", "--build-reference"]
    host.main(["combine"] + every_flag)
    assert host.load_df_from_npz(obj.paths["merged_spectra"] % 2).shape == (4, 3)
    with pytest.raises(SystemExit) as e:
        host.main(["prepare"] + every_flag)
    assert e.value.code == 2 and "prepare_from_matrix" in capsys.readouterr().err
    with pytest.raises(SystemExit):
        host.main(["plot_everything", "--name", "cli"])          # not one of the reference's verbs


def test_prepare_from_matrix_writes_the_reference_default_kwargs(tmp_path):
    """use_gpu=False unless asked, like ``cnmf prepare`` without --use_gpu (cnmf.py:461, 770, 1420)."""
    import yaml
    obj = host.cNMF(output_dir=str(tmp_path), name="d")
    obj.prepare_from_matrix(np.ones((4, 3), dtype=np.float32), [2], n_iter=2, seed=1)
    kw = yaml.load(open(obj.paths["nmf_run_parameters"]), Loader=yaml.FullLoader)
    assert kw == dict(alpha_W=0.0, alpha_H=0.0, l1_ratio_H=0.0, l1_ratio_W=0.0, beta_loss="frobenius", algo="mu",
                      tol=1e-4, mode="online", online_chunk_max_iter=1000, online_chunk_size=5000, init="random",
                      n_jobs=-1, use_gpu=False)
    with pytest.raises(NotImplementedError, match="zero"):
        obj.prepare_from_matrix(np.ones((4, 3), dtype=np.float32), [2], n_iter=2, beta_loss="itakura-saito")


def test_kl_kernel_family_by_padded_k():
    """SparseKLEngine picks the kernel family per stack (kl_engine.kl_impl_for): slab where its line layout fills the
    half-warp (Kp <= 10, 14), gather elsewhere; an explicit impl or CNMF_B200_KL_IMPL overrides."""
    from cnmf_torch_b200 import _lib
    from cnmf_torch_b200.kl_engine import kl_impl_for, pad2
    assert [kl_impl_for(pad2(k)) for k in (5, 6, 7, 8, 9, 10)] == [_lib.KL_IMPL_SLAB] * 6
    assert kl_impl_for(pad2(11)) == kl_impl_for(12) == _lib.KL_IMPL_GATHER
    assert kl_impl_for(pad2(13)) == _lib.KL_IMPL_SLAB
    assert kl_impl_for(12, 8) == _lib.KL_IMPL_SLAB and kl_impl_for(12, 13) == _lib.KL_IMPL_GATHER   # a short stack
    assert all(kl_impl_for(kp) == _lib.KL_IMPL_GATHER for kp in range(16, 34, 2))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_text_tables_are_byte_identical_to_pandas(tmp_path, dtype):
    """save_df_to_text (reference cnmf.py:44-45) through the library's host-side writer cnmf_b200_write_tsv: the
    bytes pandas' to_csv(sep='\\t') writes -- ordinary values, the notation cut-overs, signed zeros, NaN / inf,
    subnormals and 180 000 random bit patterns; a frame the fast path does not take (small, or labels that need
    quoting) still goes through pandas."""
    import pandas as pd
    from cnmf_torch_b200.cnmf import save_df_to_text
    rng = np.random.default_rng(5)
    n = 30000
    a = rng.random((n, 9)).astype(dtype)
    edge = [1e-7, 3.5e10, 0.0, np.nan, 1e-45 if dtype == np.float32 else 5e-324, -2.5, 1e16, 123456789.0, 1e-4,
            9.9999e-5, 1e15, 9.999999e15, -0.0, np.inf, -np.inf, 1.0, 100.0, 999999.0, 1e6, 1234567.0, 16777216.0,
            0.00012345, 1e-5, 65504.0, 3.4e38, 1.17549435e-38]
    a[:len(edge), 0] = np.array(edge, dtype=dtype)
    bits = rng.integers(0, 2 ** 32 if dtype == np.float32 else 2 ** 63, size=(20000, 9), dtype=np.uint64)
    a[5000:25000] = bits.astype(np.uint32).view(np.float32) if dtype == np.float32 else bits.view(np.float64)
    for name in (None, "cell"):
        df = pd.DataFrame(a, index=["cell_%d" % i for i in range(n)], columns=np.arange(1, 10))
        df.index.name = name
        save_df_to_text(df, tmp_path / "fast.txt")
        df.to_csv(tmp_path / "pandas.txt", sep="\t")
        assert (tmp_path / "fast.txt").read_bytes() == (tmp_path / "pandas.txt").read_bytes()
    wide = pd.DataFrame(a[:9 * 2000].reshape(9, -1)[:, :2000], index=np.arange(1, 10),
                        columns=["gene_%d" % i for i in range(2000)])          # a spectra table: few rows, many columns
    save_df_to_text(wide, tmp_path / "wide.txt")
    wide.to_csv(tmp_path / "wide_pandas.txt", sep="\t")
    assert (tmp_path / "wide.txt").read_bytes() == (tmp_path / "wide_pandas.txt").read_bytes()
    odd = pd.DataFrame(a[:1500], index=['c"%d' % i for i in range(1500)], columns=list("abcdefghi"))
    save_df_to_text(odd, tmp_path / "odd.txt")
    odd.to_csv(tmp_path / "odd_pandas.txt", sep="\t")
    assert (tmp_path / "odd.txt").read_bytes() == (tmp_path / "odd_pandas.txt").read_bytes()


def test_large_frames_round_trip_through_the_fast_npz_writer(tmp_path):
    """save_df_to_npz (reference cnmf.py:32-34) takes the zip writer of the replicate files for a large frame; the
    reference's loader (np.load + DataFrame(**f), cnmf.py:37-41) reads it back unchanged, suffix rule included."""
    import pandas as pd
    from cnmf_torch_b200.cnmf import load_df_from_npz, save_df_to_npz
    rng = np.random.default_rng(3)
    df = pd.DataFrame(rng.random((30000, 7)).astype(np.float32), index=["cell_%d" % i for i in range(30000)],
                      columns=np.arange(1, 8))
    save_df_to_npz(df, str(tmp_path / "usages.df.npz"))
    save_df_to_npz(df, str(tmp_path / "nosuffix"))
    for name in ("usages.df.npz", "nosuffix.npz"):
        back = load_df_from_npz(tmp_path / name)
        pd.testing.assert_frame_equal(back, df, check_names=False)
        with np.load(tmp_path / name, allow_pickle=True) as f:
            assert list(f.keys()) == ["data", "index", "columns"]


def test_native_replicate_file_writer_matches_the_python_writer(tmp_path):
    """cnmf_b200_write_npz_batch (the per-replicate spectra files of factorize, reference cnmf.py:889-892 through
    save_df_to_npz :32-34): the same bytes as the Python zip writer, and the reference's loader reads them back."""
    from cnmf_torch_b200 import cnmf as c
    rng = np.random.default_rng(2)
    genes = np.array(["gene_%d" % i for i in range(300)], dtype=object)
    cols = c._deflated_member("columns.npy", c._npy_bytes(genes))
    W = rng.random((40, 9, 300)).astype(np.float32)
    jobs = [(str(tmp_path / ("n_%d.df.npz" % j)), np.ascontiguousarray(W[j, :5 + j % 5]), np.arange(1, 6 + j % 5))
            for j in range(40)]
    jobs.append((str(tmp_path / "f64.df.npz"), W[0].astype(np.float64), np.arange(1, 10)))     # takes the Python writer
    c.save_replicate_files(jobs, cols)
    for fn, d, ix in jobs:
        df = c.load_df_from_npz(fn)
        assert np.array_equal(df.values, d) and np.array_equal(df.index.values, ix) and list(df.columns) == list(genes)
        c.save_arrays_to_df_npz(fn + ".ref.npz", d, ix, cols)
        assert open(fn, "rb").read() == open(fn + ".ref.npz", "rb").read()
