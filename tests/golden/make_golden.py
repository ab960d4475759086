"""Generate the committed golden fixtures by running the REFERENCE's own code.

Runs only in the build container (needs /root/reference).  The reference module is imported
from where it lies, with empty stand-ins for the packages that are absent from this image
(`nmf`, `scanpy`, `matplotlib`); nothing is copied.  Outputs (small .npz files beside this
script) are committed; tests never read /root/reference.

    python tests/golden/make_golden.py
"""
import importlib.util
import os
import sys
import tempfile
import types

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src/cnmf/cnmf.py"


def load_reference():
    for name in ("nmf", "scanpy", "matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["nmf"].run_nmf = None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    spec = importlib.util.spec_from_file_location("ref_cnmf", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def synth_spectra(k, n_rep, g, seed, outlier_frac=0.05):
    """Replicate spectra: jittered copies of k base programs + a few outlier rows."""
    rng = np.random.default_rng(seed)
    base = rng.gamma(0.3, 1.0, size=(k, g)) + 1e-3
    rows = []
    for r in range(n_rep):
        perm = rng.permutation(k)
        rows.append(base[perm] * rng.lognormal(0.0, 0.1, size=(k, g)))
    s = np.concatenate(rows, 0)
    n_out = max(1, int(outlier_frac * s.shape[0]))
    idx = rng.choice(s.shape[0], n_out, replace=False)
    s[idx] = rng.gamma(0.3, 1.0, size=(n_out, g)) + 1e-3
    return s.astype(np.float32)


class _Stop(Exception):
    pass


def make_ols(ref):
    """efficient_ols_all_cols (cnmf.py:56-126) and get_mean_var on a sparse and a dense target matrix."""
    import scipy.sparse as sps
    rng = np.random.default_rng(5)
    n, k, t = 700, 6, 45
    X = rng.gamma(1.0, 1.0, size=(n, k)).astype(np.float32)
    Yd = (rng.gamma(0.5, 2.0, size=(n, t)) * (rng.random((n, t)) < 0.3)).astype(np.float32)
    Yd[:, 7] = 0.0                      # a constant column exercises the variance floor
    out = dict(X=X, Y=Yd)
    out["beta_dense_raw"] = ref.efficient_ols_all_cols(X, Yd, batch_size=128, normalize_y=False)
    out["beta_dense_norm"] = ref.efficient_ols_all_cols(X, Yd, batch_size=128, normalize_y=True)
    out["beta_sparse_norm"] = ref.efficient_ols_all_cols(X, sps.csr_matrix(Yd), batch_size=256, normalize_y=True)
    m, v = ref.get_mean_var(sps.csr_matrix(Yd))
    out["mean"], out["var"] = m, v
    np.savez_compressed(os.path.join(HERE, "ols.npz"), **out)
    print("ols golden:", {k_: getattr(v_, "shape", None) for k_, v_ in out.items()})


class _FakeVarObs:
    def __init__(self, index):
        self.index = pd.Index(index)


class _FakeAnnData:
    """The attributes cNMF.consensus touches on sc.read(...) results: X, obs.index, var.index, [:, genes]."""
    def __init__(self, X, obs, var):
        self.X, self.obs, self.var = X, _FakeVarObs(obs), _FakeVarObs(var)

    def __getitem__(self, key):
        rows, cols = key
        assert rows == slice(None)
        pos = self.var.index.get_indexer(list(cols))
        return _FakeAnnData(np.array(self.X[:, pos], copy=True), self.obs.index, list(cols))


def make_backhalf(ref, tmp):
    """The reference's whole cNMF.consensus (front half + usage refit, reordering, TPM spectra, z-score spectra, final
    usage refit; cnmf.py:997-1160) on a small dense problem, with torch seeded so that the three unseeded
    torch.rand inits of fit_H_online are reproducible."""
    import torch
    rng = np.random.default_rng(21)
    n, g_all, k, n_rep = 260, 140, 4, 18
    hv_pos = np.sort(rng.choice(g_all, 70, replace=False))
    genes_all = np.array([f"g{i}" for i in range(g_all)])
    hvgs = list(genes_all[hv_pos])
    cells = np.array([f"c{i}" for i in range(n)])
    progs = rng.gamma(0.4, 1.0, size=(k, g_all)) + 1e-2
    usage = rng.dirichlet(np.full(k, 0.4), size=n)
    counts = rng.poisson(usage @ progs * 30.0).astype(np.float64)
    counts[counts.sum(1) == 0, 0] = 1.0
    tpm = counts / counts.sum(1, keepdims=True) * 1e6
    norm = counts[:, hv_pos] / counts[:, hv_pos].std(axis=0, ddof=1)
    spectra = synth_spectra(k, n_rep, len(hvgs), seed=8)
    obj = ref.cNMF(output_dir=tmp, name="back")
    rp, kw = obj.get_nmf_iter_params([k], n_iter=n_rep, random_state_seed=14, beta_loss="frobenius", batch_size=100)
    obj.save_nmf_iter_params(rp, kw)
    idx = [f"iter{i // k}_topic{i % k + 1}" for i in range(spectra.shape[0])]
    ref.save_df_to_npz(pd.DataFrame(spectra, index=idx, columns=hvgs), obj.paths["merged_spectra"] % k)
    with open(obj.paths["nmf_genes_list"], "w") as F:
        F.write("\n".join(hvgs))
    stats = pd.DataFrame([tpm.mean(axis=0), tpm.std(axis=0, ddof=0)], index=["__mean", "__std"], columns=genes_all).T
    ref.save_df_to_npz(stats, obj.paths["tpm_stats"])
    fake = {obj.paths["normalized_counts"]: _FakeAnnData(norm, cells, hvgs),
            obj.paths["tpm"]: _FakeAnnData(tpm, cells, genes_all)}
    ref.sc.read = lambda path, *_a, **_k: fake[path]
    out = {}
    for tag, flags in {"refit": dict(refit_usage=True), "norefit": dict(refit_usage=False, normalize_tpm_spectra=True)}.items():
        torch.manual_seed(77)
        obj.consensus(k, density_threshold=0.5, local_neighborhood_size=0.30, show_clustering=False, build_ref=False,
                      **flags)
        dt = "0_5"
        for key in ("consensus_spectra", "consensus_usages", "gene_spectra_tpm", "gene_spectra_score"):
            df = ref.load_df_from_npz(obj.paths[key] % (k, dt))
            out[f"{tag}_{key}"] = df.values
            out[f"{tag}_{key}_index"] = np.array(list(df.index), dtype=str)
            out[f"{tag}_{key}_columns"] = np.array(list(df.columns), dtype=str)
        os.remove(obj.paths["local_density_cache"] % k)
    np.savez_compressed(os.path.join(HERE, "consensus_backhalf.npz"), norm=norm.astype(np.float32), tpm=tpm,
                        spectra=spectra, k=k, hvgs=np.array(hvgs), genes_all=genes_all, cells=cells, torch_seed=77,
                        batch_size=100, **out)
    print("backhalf golden:", {k_: v.shape for k_, v in out.items() if not k_.endswith(("index", "columns"))})


def main():
    ref = load_reference()
    if "--only-ols" in sys.argv:
        return make_ols(ref)
    if "--only-backhalf" in sys.argv:
        return make_backhalf(ref, tempfile.mkdtemp())
    tmp = tempfile.mkdtemp()
    make_ols(ref)
    make_backhalf(ref, tempfile.mkdtemp())

    # ---- ledger (cnmf.py:701-777) ------------------------------------------------------
    led = {}
    for tag, (ks, n_iter, seed) in {
        "a": ([5, 6, 7], 15, 14), "b": ([7, 8, 9], 15, 14), "c": (list(range(5, 14)), 100, 14),
        "d": ([9, 5, 5, 7], 4, 3), "e": (6, 3, 0),
    }.items():
        obj = ref.cNMF(output_dir=tmp, name="led_" + tag)
        rp, kw = obj.get_nmf_iter_params(ks, n_iter=n_iter, random_state_seed=seed,
                                         beta_loss="frobenius")
        led[tag + "_values"] = rp[["n_components", "iter", "nmf_seed"]].values.astype(np.int64)
        led[tag + "_completed"] = rp["completed"].values.astype(bool)
        led[tag + "_args"] = np.array([n_iter, seed], dtype=np.int64)
        led[tag + "_ks"] = np.atleast_1d(np.array(ks, dtype=np.int64))
    led["kwargs_keys"] = np.array(sorted(kw.keys()))
    led["kwargs_repr"] = np.array(repr(sorted(kw.items())))
    np.savez_compressed(os.path.join(HERE, "ledger.npz"), **led)

    # ---- worker_filter (cnmf.py:53-54) -------------------------------------------------
    wf = {}
    for n, w, t in [(10, 1, 4), (10, 0, 1), (7, 2, 3), (5, 4, 8), (0, 0, 2)]:
        wf[f"{n}_{w}_{t}"] = np.array(list(ref.worker_filter(range(n), w, t)), dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "worker_filter.npz"), **wf)

    # ---- fit_H_online (cnmf.py:260-388): pins the MU pattern, eps rule, stop rule -------
    rng = np.random.default_rng(14)
    n, g, k = 230, 60, 5
    w = rng.gamma(0.5, 1.0, size=(k, g)).astype(np.float32)
    h_true = rng.dirichlet(np.full(k, 0.3), size=n).astype(np.float32) * 20
    x = rng.poisson(h_true @ w).astype(np.float32)
    h_init = np.abs(rng.normal(size=(n, k))).astype(np.float32)
    h_init[3, 2] = 0.0
    w2 = w.copy(); w2[4] = 0.0  # a dead component exercises denom < eps
    fh = dict(x=x, w=w, w_dead=w2, h_init=h_init)
    for tag, kw in {
        "default": dict(chunk_size=100, chunk_max_iter=200, h_tol=0.05),
        "tight": dict(chunk_size=64, chunk_max_iter=50, h_tol=1e-3),
        "reg": dict(chunk_size=100, chunk_max_iter=30, h_tol=1e-3, l1_reg_H=0.1, l2_reg_H=0.2),
        "one": dict(chunk_size=5000, chunk_max_iter=1, h_tol=0.0),
    }.items():
        fh["out_" + tag] = ref.fit_H_online(x, w, H_init=h_init.copy(), **kw)
        fh["kw_" + tag] = np.array(repr(sorted(kw.items())))
    fh["out_dead"] = ref.fit_H_online(x, w2, H_init=h_init.copy(), chunk_size=100,
                                      chunk_max_iter=20, h_tol=1e-3)
    np.savez_compressed(os.path.join(HERE, "fit_h_online.npz"), **fh)

    # ---- .df.npz helpers + combine_nmf (cnmf.py:32-41, 895-920) -------------------------
    obj = ref.cNMF(output_dir=tmp, name="comb")
    rp, kw = obj.get_nmf_iter_params([3, 4], n_iter=3, random_state_seed=14, beta_loss="frobenius")
    obj.save_nmf_iter_params(rp, kw)
    genes = np.array([f"g{i}" for i in range(12)], dtype=object)
    per = {}
    for kk in (3, 4):
        for it in range(3):
            sp_ = pd.DataFrame(rng.random((kk, 12)).astype(np.float32),
                               index=np.arange(1, kk + 1), columns=genes)
            ref.save_df_to_npz(sp_, obj.paths["iter_spectra"] % (kk, it))
            per[f"k{kk}_i{it}"] = sp_.values
    comb = {}
    for kk in (3, 4):
        m = obj.combine_nmf(kk)
        comb[f"merged_k{kk}_data"] = m.values
        comb[f"merged_k{kk}_index"] = np.array(list(m.index), dtype=str)
        comb[f"merged_k{kk}_columns"] = np.array(list(m.columns), dtype=str)
    comb.update(per)
    with open(obj.paths["iter_spectra"] % (3, 0), "rb") as f:
        comb["iter_file_bytes"] = np.frombuffer(f.read(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "combine.npz"), **comb)

    # ---- consensus front half (cnmf.py:997-1093) run through the reference's own method -
    captured = {}
    real_kmeans = ref.KMeans

    class RecordingKMeans(real_kmeans):
        def fit(self, X, *a, **kw):
            out = super().fit(X, *a, **kw)
            captured["labels"] = self.labels_.copy()
            captured["inertia"] = float(self.inertia_)
            captured["centers"] = self.cluster_centers_.copy()
            return out

    ref.KMeans = RecordingKMeans
    fake_adata = types.SimpleNamespace(X=np.zeros((2, 2)), obs=types.SimpleNamespace(index=[0, 1]),
                                       var=types.SimpleNamespace(index=[0, 1]))
    ref.sc.read = lambda *_a, **_k: fake_adata

    cases = {
        "small": dict(k=5, n_rep=30, g=80, seed=1, dt=0.5, lns=0.30),
        "nofilter": dict(k=4, n_rep=25, g=64, seed=2, dt=2.0, lns=0.30),
        "tightfilter": dict(k=6, n_rep=40, g=120, seed=3, dt=0.3, lns=0.25),
        "mid": dict(k=8, n_rep=50, g=300, seed=4, dt=0.5, lns=0.30),
    }
    for tag, c in cases.items():
        spectra = synth_spectra(c["k"], c["n_rep"], c["g"], c["seed"])
        obj = ref.cNMF(output_dir=tmp, name="cons_" + tag)
        idx = [f"iter{i // c['k']}_topic{i % c['k'] + 1}" for i in range(spectra.shape[0])]
        cols = [f"g{i}" for i in range(c["g"])]
        ref.save_df_to_npz(pd.DataFrame(spectra, index=idx, columns=cols),
                           obj.paths["merged_spectra"] % c["k"])

        def grab(self_x, median_spectra, _cap=captured):
            _cap["median"] = median_spectra.values.copy()
            _cap["median_index"] = np.array(median_spectra.index)
            raise _Stop()

        obj.refit_usage = grab
        captured.clear()
        try:
            obj.consensus(c["k"], density_threshold=c["dt"], local_neighborhood_size=c["lns"],
                          show_clustering=False)
        except _Stop:
            pass
        dens = ref.load_df_from_npz(obj.paths["local_density_cache"] % c["k"])
        np.savez_compressed(
            os.path.join(HERE, f"consensus_{tag}.npz"),
            spectra=spectra, k=c["k"], density_threshold=c["dt"],
            local_neighborhood_size=c["lns"],
            local_density=dens.values[:, 0], labels=captured["labels"] + 1,
            median_spectra=captured["median"], median_index=captured["median_index"],
            inertia=captured["inertia"], centers=captured["centers"])
        print(tag, spectra.shape, "kept", len(captured["labels"]), "dens dtype", dens.values.dtype,
              "median dtype", captured["median"].dtype)

    # density filter that removes everything must raise (cnmf.py:1079-1080)
    obj = ref.cNMF(output_dir=tmp, name="cons_empty")
    spectra = synth_spectra(3, 10, 20, 9)
    ref.save_df_to_npz(pd.DataFrame(spectra), obj.paths["merged_spectra"] % 3)
    try:
        obj.consensus(3, density_threshold=1e-6, show_clustering=False)
        msg = ""
    except RuntimeError as e:
        msg = str(e)
    np.savez_compressed(os.path.join(HERE, "consensus_empty.npz"), spectra=spectra, message=np.array(msg))
    print("empty-filter message:", msg)


if __name__ == "__main__":
    main()
