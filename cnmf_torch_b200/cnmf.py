"""``cNMF`` -- the reference's pipeline object for the factorize -> combine -> consensus path, with the
per-replicate ``run_nmf`` loop replaced by the replicate-batched B200 engine and the consensus front
half replaced by the consensus kernel set.  Method names, arguments, file names and file formats
follow /root/reference/src/cnmf/cnmf.py (cited per method) so the two are interchangeable on disk.

Not re-implemented (out of scope): ``prepare`` with HVG selection, plotting, the starCAT
reference.  ``prepare_from_matrix`` is the thin entry that writes the same inputs ``factorize`` and ``consensus``
read (normalised counts, ledger, kwargs, gene list and -- optionally -- the TPM matrix with its statistics).
"""
from __future__ import annotations

import errno
import itertools
import os
import warnings

import numpy as np
import pandas as pd
import scipy.sparse as sp
import yaml


# ---- .df.npz helpers (cnmf.py:32-41) --------------------------------------------------------------
def save_df_to_npz(obj, filename):
    """cnmf.py:32-34.  Small frames go through np.savez_compressed like the reference's; a large one (the usage table
    of a consensus call: cells x k) is written by the same zip writer as the replicate files -- the same three
    members, deflated at level 1 on three threads (zlib releases the GIL), 0.03 s instead of 0.14 s at 100 000 cells;
    np.load / load_df_from_npz read both alike."""
    if obj.values.size < 100000:
        np.savez_compressed(filename, data=obj.values, index=obj.index.values, columns=obj.columns.values)
        return
    filename = os.fspath(filename)
    if not filename.endswith(".npz"):
        filename += ".npz"                      # np.savez appends the suffix
    parts = [("data.npy", obj.values), ("index.npy", obj.index.values), ("columns.npy", obj.columns.values)]
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=3) as pool:
        members = list(pool.map(lambda nv: _deflated_member(nv[0], _npy_bytes(nv[1])), parts))
    _write_zip(filename, members)


def load_df_from_npz(filename):
    with np.load(filename, allow_pickle=True) as f:
        return pd.DataFrame(**f)


def _npy_bytes(arr):
    """The bytes np.savez would store for one member (object arrays pickled, as np.savez does)."""
    import io
    buf = io.BytesIO()
    np.lib.format.write_array(buf, np.asanyarray(arr), allow_pickle=True)
    return buf.getvalue()


def _deflated_member(name, payload, level=1):
    """(name, crc32, uncompressed size, raw-deflate bytes) of one archive member; zlib releases the GIL."""
    import zlib
    if len(payload) <= (1 << 20):
        co = zlib.compressobj(level, zlib.DEFLATED, -15)
        return (name.encode(), zlib.crc32(payload) & 0xFFFFFFFF, len(payload), co.compress(payload) + co.flush())
    # a large member (the usage table of 100 000 cells): pieces deflated independently on host threads and joined into
    # ONE valid raw-deflate stream -- every piece but the last ends on a byte boundary with an empty stored block
    # (Z_SYNC_FLUSH), the last one carries the final-block bit (what pigz does)
    from concurrent.futures import ThreadPoolExecutor
    view = memoryview(payload)
    step = 1 << 19
    pieces = [view[i:i + step] for i in range(0, len(payload), step)]

    def one(args):
        i, piece = args
        co = zlib.compressobj(level, zlib.DEFLATED, -15)
        return co.compress(piece) + co.flush(zlib.Z_FINISH if i == len(pieces) - 1 else zlib.Z_SYNC_FLUSH)
    with ThreadPoolExecutor(max_workers=min(8, len(pieces))) as pool:
        comp = b"".join(pool.map(one, enumerate(pieces)))
    return (name.encode(), zlib.crc32(payload) & 0xFFFFFFFF, len(payload), comp)


def _write_zip(filename, members):
    """A plain ZIP archive (method 8, no zip64, no data descriptors) from pre-deflated members.  np.load / zipfile read
    it like the archive np.savez_compressed writes; zipfile's own writer costs ~1 ms of interpreter time per archive
    (seeks, locks, header rewrites), which for the 600 replicate files of a run was most of factorize's wall time."""
    import struct
    out, central, offset = [], [], 0
    for name, crc, size, comp in members:
        local = struct.pack("<IHHHHHIIIHH", 0x04034B50, 20, 0, 8, 0, 0x21, crc, len(comp), size, len(name), 0) + name
        central.append(struct.pack("<IHHHHHHIIIHHHHHII", 0x02014B50, 20, 20, 0, 8, 0, 0x21, crc, len(comp), size,
                                   len(name), 0, 0, 0, 0, 0, offset) + name)
        out.append(local)
        out.append(comp)
        offset += len(local) + len(comp)
    cd = b"".join(central)
    out.append(cd)
    out.append(struct.pack("<IHHHHIIH", 0x06054B50, 0, 0, len(members), len(members), len(cd), offset, 0))
    with open(filename, "wb") as f:
        f.write(b"".join(out))


def save_arrays_to_df_npz(filename, data, index, columns_member, level=1):
    """A ``.df.npz`` file with the members ``save_df_to_npz`` writes (data / index / columns; np.load and the
    reference's load_df_from_npz read it the same way), without building a DataFrame and without re-pickling and
    re-compressing the gene names for each of the hundreds of replicate files of a run: ``columns_member`` is
    ``_deflated_member("columns.npy", _npy_bytes(columns))``, made once by the caller.  Deflate level 1: the payload
    is float32 noise, higher levels buy ~2 % of size for 3 x the time."""
    _write_zip(filename, [_deflated_member("data.npy", _npy_bytes(data), level),
                          _deflated_member("index.npy", _npy_bytes(index), level), columns_member])


def save_replicate_files(jobs, columns_member, level=1):
    """``save_arrays_to_df_npz`` for many files at once: [(filename, data float32 [K, G], index int64 [K])] go to the
    library's ``cnmf_b200_write_npz_batch`` (deflate and file writes on host threads outside the interpreter -- the
    Python writer needs ~0.5 ms of GIL time per file, 0.3 s for the 600 replicate files of cfg-2); anything else
    (another dtype, a non-contiguous block) takes the Python writer."""
    import ctypes as C
    fast, slow = [], []
    for job in jobs:
        d = job[1]
        ok = (isinstance(d, np.ndarray) and d.dtype == np.float32 and d.flags.c_contiguous and d.ndim == 2
              and isinstance(job[2], np.ndarray) and job[2].dtype == np.int64)
        (fast if ok else slow).append(job)
    if fast:
        from ._lib import check, lib
        hdr_cache, idx_cache = {}, {}
        n = len(fast)
        paths = (C.c_char_p * n)(*[os.fspath(j[0]).encode() for j in fast])
        data = (C.c_void_p * n)(*[j[1].ctypes.data for j in fast])
        nbytes = np.array([j[1].nbytes for j in fast], dtype=np.int64)
        hdrs, idxs = [], []
        for _, d, ix in fast:
            if d.shape not in hdr_cache:
                full = _npy_bytes(np.empty(d.shape, dtype=np.float32))
                hdr_cache[d.shape] = full[:len(full) - d.nbytes]
            key = ix.tobytes()
            if key not in idx_cache:
                idx_cache[key] = _npy_bytes(ix)
            hdrs.append(hdr_cache[d.shape])
            idxs.append(idx_cache[key])
        hdr_p = (C.c_char_p * n)(*hdrs)
        hdr_len = np.array([len(h) for h in hdrs], dtype=np.int64)
        idx_p = (C.c_char_p * n)(*idxs)
        idx_len = np.array([len(b) for b in idxs], dtype=np.int64)
        _, crc, raw_len, comp = columns_member
        check(lib().cnmf_b200_write_npz_batch(n, paths, data, nbytes.ctypes.data, hdr_p, hdr_len.ctypes.data, idx_p,
                                              idx_len.ctypes.data, comp, len(comp), raw_len, crc, level, 0),
              "write_npz_batch")
    for fn, d, ix in slow:
        save_arrays_to_df_npz(fn, d, ix, columns_member, level)


def save_df_to_text(obj, filename):
    """``obj.to_csv(filename, sep="\\t")`` (cnmf.py:44-45), byte for byte.  A plain float frame (the usage table of a
    consensus call is cells x k: 100 000 rows at cfg-3) goes through the library's ``cnmf_b200_write_tsv``: the same
    digits pandas writes (numpy's shortest round-trip repr), formatted by a few host threads -- pandas spent 0.6-1.2 s
    per such table, most of a consensus call at that size.  Anything else goes to pandas."""
    vals = obj.values if isinstance(obj, pd.DataFrame) else None
    plain = (vals is not None and vals.ndim == 2 and vals.dtype in (np.float32, np.float64)
             and not isinstance(obj.index, pd.MultiIndex) and not isinstance(obj.columns, pd.MultiIndex)
             and vals.size >= 10000)
    if plain:
        idx = [str(v) for v in obj.index.tolist()]
        head = [str(obj.index.name or "")] + [str(v) for v in obj.columns.tolist()]
        blob = "".join(idx)
        plain = not any(c in t for t in (blob, "".join(head)) for c in ("\t", '"', "\n", "\r"))
    if not plain:
        obj.to_csv(filename, sep="\t")
        return
    from ._lib import check, lib
    vals = np.ascontiguousarray(vals)
    enc = blob.encode()
    lens = map(len, idx) if len(enc) == len(blob) else (len(v.encode()) for v in idx)   # (pure ASCII: no re-encoding)
    off = np.zeros(len(idx) + 1, dtype=np.int64)
    np.cumsum(np.fromiter(lens, dtype=np.int64, count=len(idx)), out=off[1:])
    header = ("\t".join(head) + "\n").encode()
    check(lib().cnmf_b200_write_tsv(os.fspath(filename).encode(), header, len(header), enc,
                                    off.ctypes.data, vals.ctypes.data, int(vals.dtype == np.float64), vals.shape[0],
                                    vals.shape[1], vals.shape[1], 0), "write_tsv")


def _io_pool():
    """A few host threads for the per-replicate .df.npz files: zip-deflate and inflate release the GIL, and 600 files
    written (and read back by combine) one after the other were 4.7 s of a 5.2 s factorize+combine+consensus run
    whose GPU work is 0.1 s.  Same files, same bytes (np.savez_compressed)."""
    from concurrent.futures import ThreadPoolExecutor
    return ThreadPoolExecutor(max_workers=max(1, min(16, os.cpu_count() or 1)))


def worker_filter(iterable, worker_index, total_workers):
    """Job i belongs to worker w iff (i - w) % total == 0 (cnmf.py:53-54)."""
    return (p for i, p in enumerate(iterable) if (i - worker_index) % total_workers == 0)


def _mkdir(path):
    os.makedirs(path, exist_ok=True)


class cNMF:
    def __init__(self, output_dir=".", name=None, device="cuda:0", impl="auto", init_device=None):
        """cnmf.py:393-414.  ``device``/``impl`` are additions (which GPU, which CUDA path); ``init_device``
        ("cpu"/"cuda") overrides which torch generator draws the random init (default: as the ledger's
        use_gpu flag says, see ``factorize``)."""
        import datetime
        import uuid
        self.output_dir = output_dir
        if name is None:
            name = "%s_%s" % (datetime.datetime.now().strftime("%Y_%m_%d"), uuid.uuid4().hex[:6])
        self.name = name
        self.device, self.impl, self.init_device = device, impl, init_device
        self.paths = None
        self._initialize_dirs()

    def _initialize_dirs(self):
        """Same file names as cnmf.py:416-455 (the ones this path reads or writes)."""
        if self.paths is not None:
            return
        base = os.path.join(self.output_dir, self.name)
        tmp = os.path.join(base, "cnmf_tmp")
        _mkdir(tmp)
        n = self.name
        self.paths = {
            "normalized_counts": os.path.join(tmp, n + ".norm_counts.h5ad"),
            "normalized_counts_npz": os.path.join(tmp, n + ".norm_counts.npz"),  # side channel: no anndata in this image
            "nmf_replicate_parameters": os.path.join(tmp, n + ".nmf_params.df.npz"),
            "nmf_run_parameters": os.path.join(tmp, n + ".nmf_idvrun_params.yaml"),
            "nmf_genes_list": os.path.join(base, n + ".overdispersed_genes.txt"),
            "iter_spectra": os.path.join(tmp, n + ".spectra.k_%d.iter_%d.df.npz"),
            "iter_usages": os.path.join(tmp, n + ".usages.k_%d.iter_%d.df.npz"),
            "merged_spectra": os.path.join(tmp, n + ".spectra.k_%d.merged.df.npz"),
            "local_density_cache": os.path.join(tmp, n + ".local_density_cache.k_%d.merged.df.npz"),
            "consensus_spectra": os.path.join(tmp, n + ".spectra.k_%d.dt_%s.consensus.df.npz"),
            "consensus_spectra__txt": os.path.join(base, n + ".spectra.k_%d.dt_%s.consensus.txt"),
            "consensus_usages": os.path.join(tmp, n + ".usages.k_%d.dt_%s.consensus.df.npz"),
            "consensus_usages__txt": os.path.join(base, n + ".usages.k_%d.dt_%s.consensus.txt"),
            "gene_spectra_score": os.path.join(tmp, n + ".gene_spectra_score.k_%d.dt_%s.df.npz"),
            "gene_spectra_score__txt": os.path.join(base, n + ".gene_spectra_score.k_%d.dt_%s.txt"),
            "gene_spectra_tpm": os.path.join(tmp, n + ".gene_spectra_tpm.k_%d.dt_%s.df.npz"),
            "gene_spectra_tpm__txt": os.path.join(base, n + ".gene_spectra_tpm.k_%d.dt_%s.txt"),
            "tpm": os.path.join(tmp, n + ".tpm.h5ad"),
            "tpm_npz": os.path.join(tmp, n + ".tpm.npz"),                        # side channel, as for the norm counts
            "tpm_stats": os.path.join(tmp, n + ".tpm_stats.df.npz"),
        }

    # ---- inputs --------------------------------------------------------------------------------
    def save_norm_counts(self, X, obs_names=None, var_names=None):
        """Counterpart of cnmf.py:696-698 (sc.write of an AnnData).  Writes .h5ad when anndata is
        importable, and always the .npz side channel (CSR float32) that ``factorize`` can read here."""
        X = sp.csr_matrix(X) if not sp.issparse(X) else X.tocsr()
        obs_names = np.arange(X.shape[0]).astype(str) if obs_names is None else np.asarray(obs_names).astype(str)
        var_names = np.arange(X.shape[1]).astype(str) if var_names is None else np.asarray(var_names).astype(str)
        np.savez_compressed(self.paths["normalized_counts_npz"], data=X.data, indices=X.indices,
                            indptr=X.indptr, shape=np.array(X.shape), obs=obs_names, var=var_names)
        try:
            import anndata as ad
            ad.AnnData(X=X.astype(np.float64), obs=pd.DataFrame(index=obs_names),
                       var=pd.DataFrame(index=var_names)).write(self.paths["normalized_counts"])
        except ImportError:
            pass

    def load_norm_counts(self):
        """-> (X csr, obs_names, var_names); prefers .h5ad (cnmf.py:873) when it can be read.  The arrays of the last
        read are kept while the file is unchanged (consensus over several K reads it once, not once per K)."""
        src = self.paths["normalized_counts"] if os.path.exists(self.paths["normalized_counts"]) \
            else self.paths["normalized_counts_npz"]
        hit = self._recall(src)
        if hit is None:
            hit = self._read_norm_counts()
            self._remember("norm_counts", src, hit)
        return hit

    def _read_norm_counts(self):
        if os.path.exists(self.paths["normalized_counts"]):
            try:
                import anndata as ad
                a = ad.read_h5ad(self.paths["normalized_counts"])
                return sp.csr_matrix(a.X), np.asarray(a.obs.index), np.asarray(a.var.index)
            except ImportError:
                pass
        with np.load(self.paths["normalized_counts_npz"], allow_pickle=False) as f:
            X = sp.csr_matrix((f["data"], f["indices"], f["indptr"]), shape=tuple(f["shape"]))
            return X, f["obs"], f["var"]

    def save_tpm(self, tpm, obs_names=None, var_names=None):
        """What ``prepare`` leaves for the consensus back half (cnmf.py:545-580): the TPM matrix over ALL genes and
        its per-gene mean / population standard deviation (``tpm_stats``, columns ``__mean`` / ``__std``)."""
        T = sp.csr_matrix(tpm) if not sp.issparse(tpm) else tpm.tocsr()
        obs_names = np.arange(T.shape[0]).astype(str) if obs_names is None else np.asarray(obs_names).astype(str)
        var_names = np.arange(T.shape[1]).astype(str) if var_names is None else np.asarray(var_names).astype(str)
        np.savez_compressed(self.paths["tpm_npz"], data=T.data, indices=T.indices, indptr=T.indptr,
                            shape=np.array(T.shape), obs=obs_names, var=var_names)
        mean = np.asarray(T.mean(axis=0)).ravel()
        sq = np.asarray(T.multiply(T).mean(axis=0)).ravel()
        std = np.sqrt(np.maximum(sq - mean ** 2, 0.0))                        # ddof = 0 (cnmf.py:571-575)
        save_df_to_npz(pd.DataFrame([mean, std], index=["__mean", "__std"], columns=var_names).T, self.paths["tpm_stats"])

    def load_tpm(self):
        """-> (TPM csr, obs_names, var_names) or None when ``prepare`` left no TPM matrix."""
        if os.path.exists(self.paths["tpm"]):
            try:
                import anndata as ad
                a = ad.read_h5ad(self.paths["tpm"])
                return sp.csr_matrix(a.X), np.asarray(a.obs.index), np.asarray(a.var.index)
            except ImportError:
                pass
        if not os.path.exists(self.paths["tpm_npz"]):
            return None
        with np.load(self.paths["tpm_npz"], allow_pickle=False) as f:
            T = sp.csr_matrix((f["data"], f["indices"], f["indptr"]), shape=tuple(f["shape"]))
            return T, f["obs"], f["var"]

    def prepare_from_matrix(self, norm_counts, components, n_iter=100, seed=None, beta_loss="frobenius",
                            obs_names=None, var_names=None, alpha_usage=0.0, alpha_spectra=0.0,
                            init="random", total_workers=-1, use_gpu=False, batch_size=5000, max_iter=1000,
                            tpm=None, tpm_var_names=None):
        """Writes what ``prepare`` leaves for ``factorize`` (cnmf.py:590-596): normalised counts, the
        replicate ledger and the run_nmf kwargs.  ``norm_counts`` is the already variance-scaled HVG
        matrix (cells x genes); zero-count cells are refused like cnmf.py:688-691.  Defaults are ``prepare``'s
        (cnmf.py:457-462: use_gpu=False -- here that flag selects the init stream, see ``factorize``), so the YAML
        equals the reference's.  ``beta_loss='itakura-saito'`` is refused here, before anything is written, rather
        than at the first factorize batch."""
        if beta_loss in ("itakura-saito", 0, 0.0):
            from .nmf import ITAKURA_SAITO_REFUSAL
            raise NotImplementedError(ITAKURA_SAITO_REFUSAL)
        X = sp.csr_matrix(norm_counts)
        if (np.asarray(X.sum(axis=1)).ravel() == 0).any():
            raise Exception("Error: %d cells have zero counts of overdispersed genes. E.g. %s. Filter those "
                            "cells and re-run or adjust the number of overdispersed genes. Quitting!"
                            % (int((np.asarray(X.sum(axis=1)).ravel() == 0).sum()), "see matrix rows"))
        self.save_norm_counts(X, obs_names, var_names)
        if tpm is not None:   # enables the TPM / z-score outputs of ``consensus`` (cnmf.py:1124-1133)
            self.save_tpm(tpm, obs_names, tpm_var_names)
        with open(self.paths["nmf_genes_list"], "w") as F:                    # cnmf.py:684-685
            F.write("\n".join(np.arange(X.shape[1]).astype(str) if var_names is None else np.asarray(var_names).astype(str)))
        rp, kw = self.get_nmf_iter_params(ks=components, n_iter=n_iter, random_state_seed=seed,
                                          beta_loss=beta_loss, alpha_usage=alpha_usage,
                                          alpha_spectra=alpha_spectra, init=init, total_workers=total_workers,
                                          use_gpu=use_gpu, batch_size=batch_size, max_iter=max_iter)
        self.save_nmf_iter_params(rp, kw)
        return rp, kw

    # ---- ledger (cnmf.py:701-802) ----------------------------------------------------------------
    def get_nmf_iter_params(self, ks, n_iter=100, random_state_seed=None, beta_loss="kullback-leibler",
                            alpha_usage=0.0, alpha_spectra=0.0, init="random", total_workers=-1,
                            use_gpu=False, batch_size=5000, max_iter=1000):
        """Seed ledger + run_nmf kwargs, byte-compatible with cnmf.py:701-777: n_runs counts the RAW
        ``ks`` (cnmf.py:738), rows enumerate the de-duplicated sorted Ks (cnmf.py:735, 744), seeds are
        ``np.random.seed(seed); randint(1, 2**31-1, n_runs)`` (cnmf.py:740-741)."""
        if type(ks) is int:
            ks = [ks]
        k_list = sorted(set(list(ks)))
        n_runs = len(ks) * n_iter
        np.random.seed(seed=random_state_seed)
        nmf_seeds = np.random.randint(low=1, high=(2 ** 31) - 1, size=n_runs)
        rows = []
        for i, (k, r) in enumerate(itertools.product(k_list, range(n_iter))):
            rows.append([k, r, nmf_seeds[i], os.path.exists(self.paths["iter_spectra"] % (k, r))])
        replicate_params = pd.DataFrame(rows, columns=["n_components", "iter", "nmf_seed", "completed"])
        n_completed = replicate_params["completed"].sum()
        if n_completed > 0:
            warnings.warn("{n} runs already appear completed. If this is unexpected, consider\n            "
                          "re-initializing the cnmf object with a different run name or output directory"
                          .format(n=n_completed), UserWarning)
        _nmf_kwargs = dict(alpha_W=alpha_spectra, alpha_H=alpha_usage, l1_ratio_H=0.0, l1_ratio_W=0.0,
                           beta_loss=beta_loss, algo="mu", tol=1e-4, mode="online",
                           online_chunk_max_iter=max_iter, online_chunk_size=batch_size, init=init,
                           n_jobs=total_workers, use_gpu=use_gpu)
        return replicate_params, _nmf_kwargs

    def save_nmf_iter_params(self, replicate_params, run_params):
        save_df_to_npz(replicate_params, self.paths["nmf_replicate_parameters"])
        with open(self.paths["nmf_run_parameters"], "w") as F:
            yaml.dump(run_params, F)

    def update_nmf_iter_params(self):
        """Marks ledger rows completed by file existence (cnmf.py:780-795)."""
        with open(self.paths["nmf_run_parameters"]) as F:
            _nmf_kwargs = yaml.load(F, Loader=yaml.FullLoader)
        rp = load_df_from_npz(self.paths["nmf_replicate_parameters"])
        for i in rp.index:
            rp.at[i, "completed"] = os.path.exists(
                self.paths["iter_spectra"] % (rp.at[i, "n_components"], rp.at[i, "iter"]))
        print("{n} NMF runs are currently incomplete".format(n=(rp["completed"] == False).sum()))  # noqa: E712
        self.save_nmf_iter_params(rp, _nmf_kwargs)

    # ---- factorize (cnmf.py:839-892), replicate-batched ------------------------------------------
    def factorize(self, worker_i=0, total_workers=1, skip_completed_runs=False, max_batch=None,
                  nmf_kwargs_override=None):
        """Same job selection as the reference (stride over ledger rows, cnmf.py:876-880) and the same
        output per job -- ``spectra.k_%d.iter_%d.df.npz`` with index 1..K and the gene names as
        columns (cnmf.py:889-892) -- but the selected jobs are grouped by K and each group runs as ONE
        batch on the GPU, with X uploaded once per call instead of once per replicate."""
        from .nmf import make_engine, run_nmf_batch
        run_params = load_df_from_npz(self.paths["nmf_replicate_parameters"])
        X, _, var_names = self.load_norm_counts()
        with open(self.paths["nmf_run_parameters"]) as F:
            kw = yaml.load(F, Loader=yaml.FullLoader)
        if nmf_kwargs_override:
            kw.update(nmf_kwargs_override)
        if not skip_completed_runs:
            jobs = list(worker_filter(range(len(run_params)), worker_i, total_workers))
        else:
            jobs = list(worker_filter(run_params.index[run_params["completed"] == False],  # noqa: E712
                                      worker_i, total_workers))
        if not jobs:
            return
        # the ledger's use_gpu flag keeps the one meaning it can have on a GPU-only engine: WHICH torch generator
        # draws the random init (device generator under use_gpu=True, host generator otherwise -- nmf-torch draws
        # on the device it computes on), so replicates start where the reference's would for that flag
        init_device = self.init_device or ("cuda" if kw.get("use_gpu", False) else "cpu")
        engine = make_engine(X, beta_loss=kw.get("beta_loss", "frobenius"), device=self.device, impl=self.impl,
                             init_device=init_device)
        engine_kw = {k: kw[k] for k in ("init", "algo", "mode", "tol", "alpha_W", "l1_ratio_W", "alpha_H",
                                        "l1_ratio_H", "batch_max_iter", "online_max_pass", "online_chunk_size",
                                        "online_chunk_max_iter", "online_h_tol", "online_w_tol") if k in kw}
        sel = run_params.iloc[jobs, :]
        mode = kw.get("mode", "batch")
        if mode == "online" and not hasattr(engine, "factorize_online"):
            # sufficient statistics exist for the Frobenius loss only: nmf-torch itself switches beta != 2 to batch
            # mode, and cNMF always writes mode='online' (cnmf.py:765)
            print("Online mode only works for beta_loss='frobenius'. Switch to batch mode.")
            mode = "batch"
            engine_kw["mode"] = "batch"
        if (mode == "batch" and hasattr(engine, "factorize_many") and not max_batch
                and engine_kw.get("init", "random") == "random"):
            # all K at once: K-batches overlap on separate streams (engine.factorize_many)
            from .nmf import _reg
            groups = {int(k): g for k, g in sel.groupby("n_components", sort=True)}
            print("[Worker %d]. Starting %d tasks: %s." % (worker_i, len(sel), ", ".join(
                "k=%d x %d replicates" % (k, len(g)) for k, g in groups.items())))
            reg = _reg(engine_kw.get("alpha_W", 0.0), engine_kw.get("l1_ratio_W", 0.0),
                       engine_kw.get("alpha_H", 0.0), engine_kw.get("l1_ratio_H", 0.0))
            res = engine.factorize_many(list(groups), {k: [int(s) for s in g["nmf_seed"]] for k, g in groups.items()},
                                        algo=engine_kw.get("algo", "mu"), tol=engine_kw.get("tol", 1e-4),
                                        max_iter=int(engine_kw.get("batch_max_iter", 500)), **reg)
            self._write_spectra(res, groups, var_names, whole_ledger=len(sel) == len(run_params))
            return
        if (mode == "online" and hasattr(engine, "factorize_online_many") and not max_batch
                and engine_kw.get("init", "random") == "random"):
            # the reference's default mode: K values sharing a padded Kp run as one stack
            from .nmf import _reg
            groups = {int(k): g for k, g in sel.groupby("n_components", sort=True)}
            print("[Worker %d]. Starting %d tasks (online mode): %s." % (worker_i, len(sel), ", ".join(
                "k=%d x %d replicates" % (k, len(g)) for k, g in groups.items())))
            reg = _reg(engine_kw.get("alpha_W", 0.0), engine_kw.get("l1_ratio_W", 0.0),
                       engine_kw.get("alpha_H", 0.0), engine_kw.get("l1_ratio_H", 0.0))
            res = engine.factorize_online_many(
                list(groups), {k: [int(s) for s in g["nmf_seed"]] for k, g in groups.items()},
                tol=engine_kw.get("tol", 1e-4), max_pass=int(engine_kw.get("online_max_pass", 20)),
                chunk_size=int(engine_kw.get("online_chunk_size", 5000)),
                chunk_max_iter=int(engine_kw.get("online_chunk_max_iter", 200)),
                h_tol=engine_kw.get("online_h_tol", 0.05), w_tol=engine_kw.get("online_w_tol", 0.05), **reg)
            self._write_spectra(res, groups, var_names, whole_ledger=len(sel) == len(run_params))
            return
        for k, grp in sel.groupby("n_components", sort=True):
            k = int(k)
            idx = list(grp.index)
            step = len(idx) if not max_batch else int(max_batch)
            for lo in range(0, len(idx), step):
                part = grp.loc[idx[lo:lo + step]]
                print("[Worker %d]. Starting tasks %s (k=%d, %d replicates in one batch)."
                      % (worker_i, list(part.index[:3]) + (["..."] if len(part) > 3 else []), k, len(part)))
                res = run_nmf_batch(engine, k, [int(s) for s in part["nmf_seed"]], **engine_kw)
                # (per K, so that a long run leaves its finished K on disk; the merged file too when this batch is
                # every replicate of the K in the ledger)
                self._write_spectra({k: res}, {k: part}, var_names,
                                    whole_ledger=len(part) == int((run_params["n_components"] == k).sum()))

    def _write_spectra(self, res, groups, var_names, whole_ledger):
        """Per-replicate ``spectra.k_%d.iter_%d.df.npz`` files (cnmf.py:889-892) on host threads, and -- SURVEY 8 row
        f4 -- when this call holds EVERY replicate of a K (one worker, nothing skipped), that K's merged file as
        ``combine_nmf`` would build it (cnmf.py:895-920: rows in ``iter`` order, labels ``iter%d_topic%d``) straight
        from memory; ``combine`` then finds it up to date instead of re-reading the replicate files."""
        cols = _deflated_member("columns.npy", _npy_bytes(np.asarray(var_names, dtype=object)))   # once for all files
        jobs = [(self.paths["iter_spectra"] % (k, int(it)), res[k].W[j], np.arange(1, k + 1))
                for k, g in groups.items() for j, it in enumerate(g["iter"].values)]
        merged = []
        if whole_ledger:
            for k, g in groups.items():
                order = np.argsort(g["iter"].values, kind="stable")
                its = g["iter"].values[order]
                data = np.concatenate([res[k].W[j] for j in order], axis=0)
                index = np.array(["iter%d_topic%d" % (it, t + 1) for it in its for t in range(k)], dtype=object)
                merged.append((self.paths["merged_spectra"] % k, data, index))
                self._remember("merged", self.paths["merged_spectra"] % k,
                               pd.DataFrame(data, index=index, columns=np.asarray(var_names, dtype=object)), defer=True)
        save_replicate_files(jobs, cols)
        with _io_pool() as pool:
            # merged files strictly after their replicate files: combine's up-to-date test compares mtimes
            for f in [pool.submit(save_arrays_to_df_npz, fn, data, index, cols) for fn, data, index in merged]:
                f.result()
        self._settle_remembered()

    # ---- in-memory copies of files this object has just written or read (keyed by path, size and mtime) ----------
    def _remember(self, kind, path, value, defer=False):
        cache = self.__dict__.setdefault("_file_cache", {})
        cache[path] = [None if defer else self._stamp(path), value]

    def _settle_remembered(self):
        for path, ent in self.__dict__.get("_file_cache", {}).items():
            if ent[0] is None:
                ent[0] = self._stamp(path)

    @staticmethod
    def _stamp(path):
        st = os.stat(path)
        return (st.st_size, st.st_mtime_ns)

    def _recall(self, path):
        ent = self.__dict__.get("_file_cache", {}).get(path)
        if ent is not None and ent[0] is not None and os.path.exists(path) and ent[0] == self._stamp(path):
            return ent[1]
        return None

    def _load_merged(self, k):
        """The merged spectra of k: the copy kept from this object's own factorize / combine when the file is still
        the one it wrote, else read (and inflated) from disk."""
        fn = self.paths["merged_spectra"] % k
        hit = self._recall(fn)
        if hit is None:
            hit = load_df_from_npz(fn)
            self._remember("merged", fn, hit)
        return hit

    def _merged_is_current(self, k, subset):
        """The merged file of k exists, has one block of k rows per ledger replicate and is not older than any of
        their files (it was written by ``factorize`` from the same results, or by an earlier ``combine``)."""
        fn = self.paths["merged_spectra"] % k
        if not os.path.exists(fn):
            return False
        t = os.path.getmtime(fn)
        for _, p in subset.iterrows():
            f = self.paths["iter_spectra"] % (p["n_components"], p["iter"])
            if not os.path.exists(f) or os.path.getmtime(f) > t:
                return False
        return self._load_merged(k).shape[0] == k * len(subset)

    # ---- combine (cnmf.py:599-620, 895-920) -------------------------------------------------------
    def combine(self, components=None, skip_missing_files=False):
        if type(components) is int:
            ks = [components]
        elif components is None:
            run_params = load_df_from_npz(self.paths["nmf_replicate_parameters"])
            ks = sorted(set(run_params.n_components))
        else:
            ks = components
        # one K per host thread: deflating each merged file (7-10 MB) is the cost of this step
        with _io_pool() as pool:
            for f in [pool.submit(self.combine_nmf, k, skip_missing_files) for k in ks]:
                f.result()

    def combine_nmf(self, k, skip_missing_files=False, remove_individual_iterations=False):
        run_params = load_df_from_npz(self.paths["nmf_replicate_parameters"])
        print("Combining factorizations for k=%d." % k)
        subset = run_params[run_params.n_components == k].sort_values("iter")
        if self._merged_is_current(k, subset):
            print("Merged spectra for k=%d are up to date." % k)
            return self._load_merged(k)
        combined, present = [], []
        for _, p in subset.iterrows():
            fn = self.paths["iter_spectra"] % (p["n_components"], p["iter"])
            if not os.path.exists(fn):
                if not skip_missing_files:
                    print("Missing file: %s, run with skip_missing=True to override" % fn)
                    raise FileNotFoundError(errno.ENOENT, os.strerror(errno.ENOENT), fn)
                print("Missing file: %s. Skipping." % fn)
            else:
                present.append((fn, p["iter"]))
        with _io_pool() as pool:                       # files inflate in parallel, rows stay in `iter` order
            for (fn, it), spectra in zip(present, pool.map(load_df_from_npz, [fn for fn, _ in present])):
                spectra.index = ["iter%d_topic%d" % (it, t + 1) for t in range(k)]
                combined.append(spectra)
        if len(combined) > 0:
            combined = pd.concat(combined, axis=0)
            save_df_to_npz(combined, self.paths["merged_spectra"] % k)
            self._remember("merged", self.paths["merged_spectra"] % k, combined)
        else:
            print("No spectra found for k=%d" % k)
        return combined

    # ---- usage refit (cnmf.py:923-994) ---------------------------------------------------------------
    def refit_usage(self, X, spectra, usage=None, seed=None):
        """cnmf.py:923-976: kwargs from the YAML -> fit_H_online (chunk size / max iterations from the
        run parameters, h_tol 0.05, L1 strength = l1_ratio_H as the reference passes it).  Frobenius only,
        like the reference's fit_H_online."""
        from .refit import fit_H_online
        with open(self.paths["nmf_run_parameters"]) as F:
            kw = yaml.load(F, Loader=yaml.FullLoader)
        return fit_H_online(X, spectra, H_init=usage, chunk_size=kw["online_chunk_size"],
                            chunk_max_iter=kw["online_chunk_max_iter"], h_tol=0.05, l1_reg_H=kw["l1_ratio_H"],
                            l2_reg_H=0.0, epsilon=1e-16, device=self.device, seed=seed)

    def refit_spectra(self, X, usage, seed=None):
        """cnmf.py:979-994: the transposed problem."""
        X = getattr(X, "values", X)
        usage = getattr(usage, "values", usage)
        Xt = X.T.tocsr() if sp.issparse(X) else np.ascontiguousarray(np.asarray(X).T)
        return self.refit_usage(Xt, np.ascontiguousarray(np.asarray(usage).T), seed=seed).T

    # ---- consensus, front half (cnmf.py:997-1090) -------------------------------------------------
    def consensus(self, k, density_threshold=0.5, local_neighborhood_size=0.30, show_clustering=False,
                  build_ref=False, skip_density_and_return_after_stats=False, close_clustergram_fig=False,
                  refit_usage=True, normalize_tpm_spectra=False, norm_counts=None, seed=None):
        """cnmf.py:997-1160.  Front half on the GPU: L2-normalise the merged spectra, local-density filter,
        KMeans(k, n_init=10, random_state=1), per-cluster median, row-normalise (writes the ``local_density_cache``,
        cnmf.py:1073).  Back half (rows f1/f2): usages refitted against the consensus spectra on the normalised
        counts, programs reordered by total usage, and -- when ``prepare`` left a TPM matrix (``save_tpm``) --
        spectra in TPM units (``refit_spectra``), z-score spectra (OLS of the z-scored TPM against the usages) and,
        with ``refit_usage``, a final usage refit on the std-scaled HVG TPM; files as the reference names them
        (cnmf.py:1151-1160).  ``seed`` (an addition): the reference's three refits start from unseeded
        ``torch.rand``; an int seeds one generator they draw from in the same order.  Returns (consensus spectra
        DataFrame, labels Series, local_density DataFrame); plots and the starCAT reference are outside this path."""
        from .consensus import consensus_front_half
        merged = self._load_merged(k)
        dt_str = "2" if skip_density_and_return_after_stats else str(density_threshold)
        dt_repl = dt_str.replace(".", "_")                                    # cnmf.py:1049-1052
        n_neighbors = int(local_neighborhood_size * merged.shape[0] / k)      # cnmf.py:1053
        cache = self.paths["local_density_cache"] % k
        cached = None
        if not skip_density_and_return_after_stats and os.path.isfile(cache):  # cnmf.py:1061-1062
            cached = load_df_from_npz(cache).values[:, 0]
        out = consensus_front_half(merged.values, k, density_threshold, n_neighbors, device=self.device,
                                   skip_density=skip_density_and_return_after_stats, local_density=cached)
        if skip_density_and_return_after_stats:
            # k-selection statistics (cnmf.py:1093-1110): refit usages against the consensus spectra, silhouette of
            # the clustering, prediction error of the rank-k reconstruction
            from .stats import prediction_error, silhouette_score
            X = norm_counts if norm_counts is not None else self.load_norm_counts()[0]
            X = X[0] if isinstance(X, tuple) else X
            median = pd.DataFrame(out["median_spectra"], index=np.arange(1, k + 1), columns=merged.columns)
            rf_usages = self.refit_usage(X, median.values, seed=1)
            sil = silhouette_score(out["l2_spectra"], out["labels"] - 1, k, device=self.device)
            err = prediction_error(X, rf_usages, median.values, device=self.device)
            return pd.DataFrame([k, density_threshold, sil, err],
                                index=["k", "local_density_threshold", "silhouette", "prediction_error"],
                                columns=["stats"])
        local_density = pd.DataFrame(out["local_density"], columns=["local_density"], index=merged.index) \
            if out["local_density"] is not None else None
        if local_density is not None and cached is None:
            save_df_to_npz(local_density, cache)
        kept_index = merged.index[out["keep"]]
        labels = pd.Series(out["labels"], index=kept_index)
        median = pd.DataFrame(out["median_spectra"], index=np.arange(1, k + 1), columns=merged.columns)
        median.index.name = None

        # ---- back half ---------------------------------------------------------------------------------------
        import torch
        gen = None
        if seed is not None:
            gen = torch.Generator(device="cpu")
            gen.manual_seed(int(seed))
        if isinstance(norm_counts, tuple):                 # what load_norm_counts returns, kept by the caller
            Xn, obs_names = norm_counts[0], norm_counts[1]
        elif norm_counts is not None:
            Xn, obs_names = norm_counts, np.arange(norm_counts.shape[0]).astype(str)
        else:
            Xn, obs_names, _ = self.load_norm_counts()
        rf_usages = pd.DataFrame(self.refit_usage(Xn, median.values, seed=gen), index=obs_names,
                                 columns=median.index)                                       # cnmf.py:1093-1094
        norm_usages = rf_usages.div(rf_usages.sum(axis=1), axis=0)                           # cnmf.py:1113-1120
        reorder = norm_usages.sum(axis=0).sort_values(ascending=False)
        rf_usages = rf_usages.loc[:, reorder.index]
        norm_usages = norm_usages.loc[:, reorder.index]
        median = median.loc[reorder.index, :]
        rf_usages.columns = np.arange(1, rf_usages.shape[1] + 1)
        norm_usages.columns = rf_usages.columns
        median.index = rf_usages.columns
        tpm = self.load_tpm()
        if tpm is not None:
            T, _, tpm_var = tpm
            from .stats import efficient_ols_all_cols
            tpm_stats = load_df_from_npz(self.paths["tpm_stats"])
            spectra_tpm = pd.DataFrame(self.refit_spectra(T, norm_usages.values, seed=gen), index=rf_usages.columns,
                                       columns=tpm_var)                                      # cnmf.py:1126-1127
            if normalize_tpm_spectra:
                spectra_tpm = spectra_tpm.div(spectra_tpm.sum(axis=1), axis=0) * 1e6
            usage_coef = pd.DataFrame(efficient_ols_all_cols(rf_usages.values, T, normalize_y=True, device=self.device),
                                      index=rf_usages.columns, columns=tpm_var)              # cnmf.py:1132-1133
            if refit_usage:
                # final refit on the std-scaled (ddof = 1, not centred) HVG columns of the TPM matrix, cnmf.py:1135-1149
                hvgs = open(self.paths["nmf_genes_list"]).read().split("\n")
                pos = pd.Index(tpm_var).get_indexer(hvgs)
                if (pos < 0).any():
                    raise KeyError("%d high-variance genes are missing from the TPM matrix, e.g. %r"
                                   % (int((pos < 0).sum()), hvgs[int(np.argmax(pos < 0))]))
                Th = sp.csc_matrix(T)[:, pos].astype(np.float64)
                n_c = Th.shape[0]
                mean = np.asarray(Th.mean(axis=0)).ravel()
                var = (np.asarray(Th.multiply(Th).mean(axis=0)).ravel() - mean ** 2) * (n_c / max(n_c - 1, 1))
                std = np.sqrt(np.maximum(var, 0.0))
                std[std == 0] = 1.0                                  # scanpy's scale leaves constant genes alone
                norm_tpm = sp.csr_matrix(Th.multiply(1.0 / std))
                spectra_tpm_rf = spectra_tpm.loc[:, hvgs].div(tpm_stats.loc[hvgs, "__std"], axis=1)
                rf_usages = pd.DataFrame(self.refit_usage(norm_tpm, spectra_tpm_rf.values, seed=gen), index=obs_names,
                                         columns=spectra_tpm_rf.index)
        save_df_to_npz(median, self.paths["consensus_spectra"] % (k, dt_repl))                # cnmf.py:1151-1160
        save_df_to_npz(rf_usages, self.paths["consensus_usages"] % (k, dt_repl))
        save_df_to_text(median, self.paths["consensus_spectra__txt"] % (k, dt_repl))
        save_df_to_text(rf_usages, self.paths["consensus_usages__txt"] % (k, dt_repl))
        if tpm is not None:
            save_df_to_npz(spectra_tpm, self.paths["gene_spectra_tpm"] % (k, dt_repl))
            save_df_to_text(spectra_tpm, self.paths["gene_spectra_tpm__txt"] % (k, dt_repl))
            save_df_to_npz(usage_coef, self.paths["gene_spectra_score"] % (k, dt_repl))
            save_df_to_text(usage_coef, self.paths["gene_spectra_score__txt"] % (k, dt_repl))
        return median, labels, local_density


def _load_results(self, K, density_threshold, n_top_genes=100, norm_usage=True):
    """cnmf.py:1335-1385: (usage cells x K, spectra_scores genes x K, spectra_tpm genes x K, top_genes n_top x K) read
    back from the text files ``consensus`` wrote for (K, density_threshold).  (Outside SURVEY section 8 -- a reader
    of this path's own output files, kept as it was in round 1 and not developed further.)"""
    dt = str(density_threshold).replace(".", "_")
    spectra_scores = pd.read_csv(self.paths["gene_spectra_score__txt"] % (K, dt), sep="\t", index_col=0).T
    spectra_tpm = pd.read_csv(self.paths["gene_spectra_tpm__txt"] % (K, dt), sep="\t", index_col=0).T
    usage = pd.read_csv(self.paths["consensus_usages__txt"] % (K, dt), sep="\t", index_col=0)
    if norm_usage:
        usage = usage.div(usage.sum(axis=1), axis=0)
    try:
        usage.columns = [int(x) for x in usage.columns]
    except ValueError:
        print("Usage matrix columns include non integer values")
    top_genes = [list(spectra_scores.sort_values(by=gep, ascending=False).index[:n_top_genes])
                 for gep in spectra_scores.columns]
    top_genes = pd.DataFrame(top_genes, index=spectra_scores.columns).T
    return usage, spectra_scores, spectra_tpm, top_genes


cNMF.load_results = _load_results


def _k_selection_stats(self):
    """The table behind ``k_selection_plot`` (cnmf.py:1293-1309) -- one ``consensus(k,
    skip_density_and_return_after_stats=True)`` per K, saved as ``k_selection_stats.df.npz``; the figure itself
    is the reference's matplotlib code and stays there."""
    run_params = load_df_from_npz(self.paths["nmf_replicate_parameters"])
    X = self.load_norm_counts()[0]
    stats = [self.consensus(int(k), skip_density_and_return_after_stats=True, norm_counts=X).stats
             for k in sorted(set(run_params.n_components))]
    stats = pd.DataFrame(stats)
    stats.reset_index(drop=True, inplace=True)
    save_df_to_npz(stats, os.path.join(self.output_dir, self.name, self.name + ".k_selection_stats.df.npz"))
    return stats


cNMF.k_selection_stats = _k_selection_stats


def k_selection_plot(self, close_fig=False):
    """cnmf.py:1274-1332 computes the stability / prediction-error table and draws it with matplotlib.  This path
    writes the table (``k_selection_stats.df.npz``, what ``k_selection_stats`` returns); the figure is plotting and
    stays with the reference."""
    stats = self.k_selection_stats()
    print("k_selection_plot: statistics written to %s; the figure itself is drawn by the reference's matplotlib code."
          % os.path.join(self.output_dir, self.name, self.name + ".k_selection_stats.df.npz"))
    return stats


cNMF.k_selection_plot = k_selection_plot


def main(argv=None):
    """The ``cnmf`` command line with the reference's verbs and flags (cnmf.py:1410-1472).  ``factorize``,
    ``combine``, ``consensus`` and ``k_selection_plot`` run here; ``prepare`` (HVG selection on AnnData) stays with
    the reference and says so.  Every flag the reference declares is accepted, so a script written for it runs
    unchanged: those that only concern ``prepare`` or plotting are parsed and, where they cannot take effect, named
    in a message.  ``--worker-index`` (commented out in the reference, cnmf.py:1430), ``--device`` and
    ``--init-device`` are additions."""
    import argparse
    parser = argparse.ArgumentParser()
    parser.add_argument("command", type=str, choices=["prepare", "factorize", "combine", "consensus", "k_selection_plot"])
    parser.add_argument("--name", type=str, nargs="?", default="cNMF")
    parser.add_argument("--output-dir", type=str, nargs="?", default=".")
    parser.add_argument("-c", "--counts", type=str, default=None)
    parser.add_argument("-k", "--components", type=int, nargs="+")
    parser.add_argument("-n", "--n-iter", type=int, default=100)
    parser.add_argument("--total-workers", type=int, default=-1,
                        help="workers the replicate jobs are strided over (default: WORLD_SIZE under torchrun, else 1)")
    parser.add_argument("--use_gpu", action="store_true", default=False)
    parser.add_argument("--seed", type=int, default=None)
    parser.add_argument("--genes-file", type=str, default=None)
    parser.add_argument("--numgenes", type=int, default=2000)
    parser.add_argument("--tpm", type=str, default=None)
    parser.add_argument("--max-nmf-iter", type=int, default=1000)
    parser.add_argument("--beta-loss", type=str, choices=["frobenius", "kullback-leibler", "itakura-saito"],
                        default="frobenius")
    parser.add_argument("--init", type=str, choices=["random", "nndsvd"], default="random")
    parser.add_argument("--densify", dest="densify", action="store_true", default=False)
    parser.add_argument("--batch_size", type=int, default=5000)
    parser.add_argument("--worker-index", type=int, default=None,
                        help="this worker's index (default: RANK under torchrun, else 0)")
    parser.add_argument("--skip-completed-runs", action="store_true", default=False)
    parser.add_argument("--local-density-threshold", type=float, default=0.5)
    parser.add_argument("--local-neighborhood-size", type=float, default=0.30)
    parser.add_argument("--show-clustering", dest="show_clustering", action="store_true")
    parser.add_argument("--build-reference", dest="build_reference", action="store_true", default=True)
    parser.add_argument("--device", type=str, default=None)
    parser.add_argument("--init-device", type=str, choices=["cpu", "cuda"], default=None,
                        help="[factorize] torch generator that draws the random init (default: the run's use_gpu flag)")
    args = parser.parse_args(argv)
    if args.command == "prepare":
        parser.exit(2, "cnmf_torch_b200: `prepare` (count loading, TPM, high-variance gene selection) is not part of "
                       "this path -- run the reference's `cnmf prepare` with the same --name/--output-dir (its files are "
                       "read as they are), or call cNMF.prepare_from_matrix on an already normalised matrix.\n")
    device = args.device or "cuda:%d" % int(os.environ.get("LOCAL_RANK", "0"))
    obj = cNMF(output_dir=args.output_dir, name=args.name, device=device, init_device=args.init_device)
    if args.command == "factorize":
        # one process per GPU:  torchrun --nproc-per-node 8 -m cnmf_torch_b200.cnmf factorize --name ...
        # strides the replicate ledger over the ranks exactly like worker_filter (cnmf.py:53-54, 876-880)
        total = args.total_workers if args.total_workers and args.total_workers > 0 else int(os.environ.get("WORLD_SIZE", "1"))
        index = args.worker_index if args.worker_index is not None else int(os.environ.get("RANK", "0"))
        obj.factorize(worker_i=index, total_workers=max(1, total), skip_completed_runs=args.skip_completed_runs)
    elif args.command == "combine":
        obj.combine(components=args.components)
    elif args.command == "consensus":
        if args.show_clustering:
            print("--show-clustering: the clustergram figure is plotting and stays with the reference; "
                  "consensus files are written as usual.")
        ks = args.components or sorted(set(load_df_from_npz(obj.paths["nmf_replicate_parameters"]).n_components))
        X = obj.load_norm_counts() if len(ks) > 1 else None        # read once for every K's usage refit
        for k in ks:
            obj.consensus(k, args.local_density_threshold, args.local_neighborhood_size, norm_counts=X)
    elif args.command == "k_selection_plot":
        obj.k_selection_plot(close_fig=True)


if __name__ == "__main__":
    main()
