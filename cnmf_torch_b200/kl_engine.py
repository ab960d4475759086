"""Replicate-batched sparse KL engine (host side): CSR X and CSR X^T resident on one GPU, R
replicates interleaved in Hr [N x R*Kp] / Wt [G x R*Kp].  Replaces the per-replicate
``run_nmf(beta_loss='kullback-leibler')`` of /root/reference/src/cnmf/cnmf.py:805-821."""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np
import scipy.sparse as sp
import torch

from . import _lib
from ._lib import KlProblem, check, lib, ptr
from .engine import EPS_DEFAULT, BatchResult, SeedInit, draw_init

_KL_IMPL = {"slab": _lib.KL_IMPL_SLAB, "gather": _lib.KL_IMPL_GATHER}


def kl_impl_for(Kp: int, R: int = 1 << 30) -> int:
    """The faster kernel family for a stack of R replicates of padded K = Kp, measured at 100 000 x 2000 cells (ms per
    batched iteration, R = 200 / R = 26; profiles/NOTES.md): Kp=6 slab 8.6 / 1.45 vs gather 9.8 / 1.70; 8: 9.6 / 1.67 vs
    11.4 / 1.92; 10: 13.5 / 2.16 vs 13.6 / 2.40; 12: 19.0 / 2.83 vs 15.6 / 2.68; 14: 19.1 / 2.85 vs 19.0 / 3.34.
    A slab line holds floor(16 / (Kp/2)) replicate pairs: at Kp = 12 only 12 of a half-warp's 16 lanes have work and a
    group is 4 replicates, so the gather family wins there -- except for a stack of a few replicates (a widely sharded
    ledger), which fills a gather warp (32 consecutive replicates) badly: R = 4 / 8 / 12 slab 0.56 / 0.95 / 1.34 vs
    gather 0.79 / 1.03 / 1.38 ms.  (A batch compacted during a run keeps the family it started with.)  From Kp = 16 on the slab kernels' accumulators no longer fit the register file."""
    if Kp <= 10 or Kp == 14 or (Kp == 12 and R <= 12):
        return _lib.KL_IMPL_SLAB
    return _lib.KL_IMPL_GATHER


def pad2(k: int) -> int:
    """The KL kernels take any even Kp: every padded component is a gathered value and two FMAs per stored entry."""
    m = 4 if os.environ.get("CNMF_B200_KL_PAD4") else 2      # (dev knob: the multiple-of-4 padding of the dense path)
    return (int(k) + m - 1) // m * m


class SparseKLEngine:
    def __init__(self, X, device="cuda:0", init_device="cpu", impl=None, normalise=True):
        """``impl``: "slab" (shared-memory slab kernels) or "gather" (coalesced L2-gather kernels); default (None, or
        env CNMF_B200_KL_IMPL): per stack, by its padded K -- see ``kl_impl_for``.
        ``normalise``: the kernels gather fp16 copies of the factors, so the engine works on X * c^2 with c the power
        of two that brings mean(X * c^2) to about 1 -- KL multiplicative updates are exactly covariant (H, W scale
        by c, a power of two: the same bits), and factor entries then sit in the middle of the fp16 range whatever
        the units of X.  Results and objectives are returned in the units of X."""
        self.lib = lib()
        self.init_device = init_device   # which torch generator draws the 'random' init (engine.draw_init)
        self._impl_arg = impl or os.environ.get("CNMF_B200_KL_IMPL")
        if not torch.cuda.is_available():
            raise _lib.Cnmfb200Error("CUDA device required: cnmf_torch_b200 has no CPU path")
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        if torch.is_tensor(X):
            X = X.cpu().numpy()
        csr = sp.csr_matrix(X, dtype=np.float32, copy=True)   # never touch the caller's buffers
        csr.eliminate_zeros()
        csr.sort_indices()
        self._csr_host = csr
        if csr.nnz and float(csr.data.min()) < 0.0:
            raise ValueError("X must be non-negative")   # the kernels rely on stored entries being > 0
        self.N, self.G = csr.shape
        self.impl = _KL_IMPL[self._impl_arg] if self._impl_arg else None      # None: per stack (kl_impl_for)
        self.nnz = int(csr.nnz)
        if self.nnz >= 2 ** 31:
            raise ValueError("nnz must fit 32-bit offsets")
        sum_x = float(csr.data.sum(dtype=np.float64))
        self.x_mean = sum_x / (float(self.N) * float(self.G))         # of the caller's X: the init is the reference's
        self.c = 1.0
        if normalise and self.x_mean > 0:
            self.c = 2.0 ** int(round(-0.5 * math.log2(self.x_mean)))
        dev = self.device
        c2 = np.float32(self.c * self.c)
        up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a.astype(dt))).to(dev)
        self.row_ptr, self.col_idx, self.val = up(csr.indptr, np.int32), up(csr.indices, np.int32), up(csr.data * c2, np.float32)
        # CSR of X^T on the device: a stable sort of the entries by column keeps the rows of every column ascending
        # (scipy's csr.T.tocsr() + sort_indices is a second of host time at 15 M entries -- more than the iterations of
        # a small worker share)
        counts = torch.bincount(self.col_idx.long(), minlength=self.G)
        self.t_row_ptr = torch.zeros(self.G + 1, device=dev, dtype=torch.int32)
        self.t_row_ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
        rows = torch.repeat_interleave(torch.arange(self.N, device=dev, dtype=torch.int32),
                                       (self.row_ptr[1:] - self.row_ptr[:-1]).long())
        perm = torch.sort(self.col_idx, stable=True).indices
        self.t_col_idx = rows[perm].contiguous()
        self.t_val = self.val[perm].contiguous()
        del rows, perm, counts
        out = torch.zeros(2, device=dev, dtype=torch.float64)
        check(self.lib.cnmf_b200_kl_constants(ptr(self.val), self.nnz, ptr(out), self._stream()), "kl_constants")
        o = out.cpu().numpy()
        self.sumX, self.xlogx = float(o[0]), float(o[1])               # of the scaled matrix the kernels see

    @staticmethod
    def _stream():
        return torch.cuda.current_stream().cuda_stream

    def host_matrix(self) -> np.ndarray:
        return np.asarray(self._csr_host.toarray(), dtype=np.float64)

    def make_init(self, K, seeds, init_device=None):
        return draw_init(self.x_mean, self.N, self.N, 0, self.G, int(K), seeds, self.device,
                         init_device or self.init_device)

    def _init_for(self, K, seeds):
        """(H0, W0) host arrays from the host generator, or a SeedInit marker: the draw then happens on the device
        (cnmf_b200_init_random, bit-identical to torch's CUDA generator) straight into the batch."""
        if self.init_device == "cuda" and self.N * int(K) < 2 ** 31 and int(K) * self.G < 2 ** 31:
            return SeedInit([int(s) for s in seeds]), None
        return self.make_init(K, seeds)

    def new_batch(self, K, seeds=None, H0=None, W0=None, l1_H=0.0, l2_H=0.0, l1_W=0.0, l2_W=0.0, eps=EPS_DEFAULT):
        if H0 is None or W0 is None:
            H0, W0 = self._init_for(K, seeds)
        return KlBatch(self, K, H0, W0, l1_H, l2_H, l1_W, l2_W, eps)

    def factorize(self, K, seeds=None, H0=None, W0=None, algo="mu", tol=1e-4, max_iter=500, l1_H=0.0, l2_H=0.0,
                  l1_W=0.0, l2_W=0.0, eps=EPS_DEFAULT, check_every=10, return_H=False, compact=0.8) -> BatchResult:
        """Batch-mode KL MU of R replicates; same stopping rule as the Frobenius engine."""
        if algo != "mu":
            raise NotImplementedError("only multiplicative updates exist for beta_loss='kullback-leibler'")
        b = self.new_batch(K, seeds, H0, W0, l1_H, l2_H, l1_W, l2_W, eps)
        r = _run_kl_batch(b, tol, max_iter, check_every, compact)
        W, H = b.results(return_H)
        return BatchResult(W=W, H=H, launches=int(b.launches), **r)

    def factorize_many(self, ks, seeds_by_k, algo="mu", tol=1e-4, max_iter=500, l1_H=0.0, l2_H=0.0, l1_W=0.0,
                       l2_W=0.0, eps=EPS_DEFAULT, check_every=10, return_H=False, compact=0.8):
        """Several K in one call (the shape of a cNMF run: K = 5..13 x 100 replicates).  K values that round up to
        the same even Kp are stacked into ONE batch (K = 9, 10 -> Kp 10): the kernels only see Kp, a taller stack
        fills the replicate groups of the slab kernel better (13 replicates of one K on one of 8 GPUs are 2 groups
        and a third; 26 are 4 and a third), and X is streamed once for both.  A slab kernel launch occupies every SM,
        so stacks run one after the other (no side streams as on the dense path).  Returns {K: BatchResult}."""
        if algo != "mu":
            raise NotImplementedError("only multiplicative updates exist for beta_loss='kullback-leibler'")
        stacks = {}
        for k in [int(k) for k in ks]:
            stacks.setdefault(pad2(k), []).append(k)
        out = {}
        for group in stacks.values():
            groups = [(k,) + tuple(self._init_for(k, seeds_by_k[k])) for k in group]
            b = KlBatch(self, None, None, None, l1_H, l2_H, l1_W, l2_W, eps, groups=groups)
            r = _run_kl_batch(b, tol, max_iter, check_every, compact)
            for k, sl, W, H in b.results_by_group(return_H):
                out[k] = BatchResult(W=W, H=H, err=r["err"][sl], n_iter=r["n_iter"][sl],
                                     trace=[(it, v[sl]) for it, v in r["trace"]], launches=int(b.launches))
            del b
        return out


def _run_kl_batch(b: "KlBatch", tol, max_iter, check_every, compact=0.8):
    """run_nmf's batch loop on a resident KL batch: objective every ``check_every`` iterations, replicates with
    prev <= cur or |prev - cur| / init < tol are frozen.

    Frozen replicates still occupy lanes (a slab group of 6 is skipped only when all of them are frozen; a gather warp
    covers 32 consecutive replicates), and the replicates of a stack converge at very different iteration counts, so
    once no more than ``compact`` of the working batch is alive the live replicates are copied into a smaller batch
    (``KlBatch.subset``) and the loop continues there; their state is copied back at the end (or at the next
    compaction).  Replicates are independent and every reduction is per replicate in a fixed order, so the results
    are bit-identical to the uncompacted run (``tests/test_gpu_kl.py::test_kl_compaction_is_bit_identical``).
    ``compact=0`` turns it off."""
    R = b.R
    init = b.objective()
    prev, cur = init.copy(), init.copy()
    trace = [(0, init.copy())]
    n_iter = np.full(R, max_iter, dtype=np.int64)
    alive = np.ones(R, dtype=bool)
    wb, widx = b, np.arange(R)          # the working batch and its replicates' positions in b
    it = 0
    while it < max_iter and alive.any():
        step = min(check_every, max_iter - it)
        wb.iterate(step)
        it += step
        cur[widx] = wb.read_loss()
        trace.append((it, cur.copy()))
        if it % check_every == 0:
            conv = alive & ((prev <= cur) | (np.abs((prev - cur) / init) < tol))
            if conv.any():
                n_iter[conv] = it
                alive &= ~conv
                wb.set_active(alive[widx])
            prev = np.where(alive, cur, prev)
            n_alive = int(alive.sum())
            if compact > 0 and 0 < n_alive <= compact * wb.R and wb.R - n_alive >= 6 and it < max_iter:
                if wb is not b:
                    b.absorb(wb, widx)
                else:
                    b.release_workspace()
                widx = np.flatnonzero(alive)
                wb = b.subset(widx)
    if wb is not b:
        b.absorb(wb, widx)
    return dict(err=np.sqrt(np.maximum(2.0 * cur, 0.0)), n_iter=n_iter, trace=trace)


class KlBatch:
    """Replicates resident on the device.  ``groups`` = [(K, H0, W0) | (K, SeedInit, None), ...] stacks replicates of
    different K that share the same even Kp (padding components are zero and stay zero)."""

    def __init__(self, eng: SparseKLEngine, K, H0, W0, l1_H, l2_H, l1_W, l2_W, eps, groups=None):
        if groups is None:
            groups = [(int(K), H0, W0)]
        self.eng = eng
        self.Ks = [int(g[0]) for g in groups]
        self.Rs = [len(g[1].seeds) if isinstance(g[1], SeedInit) else int(g[2].shape[0]) for g in groups]
        self.K = max(self.Ks)
        self.R, self.Kp = sum(self.Rs), pad2(self.K)
        if self.Kp > _lib.MAX_KP:
            raise ValueError(f"n_components={self.K} exceeds the engine limit {_lib.MAX_KP}")
        if any(pad2(k) != self.Kp for k in self.Ks):
            raise ValueError("all K of one stack must round up to the same even number")
        dev = eng.device
        R, Kp = self.R, self.Kp
        self.Hr = torch.zeros((eng.N, R, Kp), device=dev, dtype=torch.float32)
        self.Wt = torch.zeros((eng.G, R, Kp), device=dev, dtype=torch.float32)
        lo = 0
        for (k, h0, w0), r in zip(groups, self.Rs):
            if isinstance(h0, SeedInit):
                # device draw (one launch for the r replicates of this K) into the dense engine's K-major layout,
                # then one permuted copy into the replicate-interleaved layout of this path
                ldh, ldw = (eng.N + 3) // 4 * 4, (eng.G + 3) // 4 * 4
                Ht = torch.zeros((r, Kp, ldh), device=dev, dtype=torch.float32)
                Wall = torch.zeros((r, Kp, ldw), device=dev, dtype=torch.float32)
                seeds = torch.tensor(h0.seeds, dtype=torch.int64, device=dev)
                check(eng.lib.cnmf_b200_init_random(ptr(seeds), r, k, Kp, eng.N, eng.N, 0, eng.G,
                                                    math.sqrt(float(eng.x_mean) / k), ptr(Ht), ldh, ptr(Wall), ldw,
                                                    eng._stream()), "init_random")
                self.Hr[:, lo:lo + r, :] = Ht[:, :, :eng.N].permute(2, 0, 1)
                self.Wt[:, lo:lo + r, :] = Wall[:, :, :eng.G].permute(2, 0, 1)
                del Ht, Wall
            else:
                h0 = torch.as_tensor(h0, dtype=torch.float32).to(dev)   # [r, N, k]
                w0 = torch.as_tensor(w0, dtype=torch.float32).to(dev)   # [r, k, G]
                self.Hr[:, lo:lo + r, :k] = h0.permute(1, 0, 2)
                self.Wt[:, lo:lo + r, :k] = w0.permute(2, 0, 1)
            lo += r
        if eng.c != 1.0:               # the engine's units: X * c^2 = (cH)(cW); a power of two, exact
            self.Hr.mul_(eng.c)
            self.Wt.mul_(eng.c)
        self._regs = (l1_H, l2_H, l1_W, l2_W, eps)
        self._setup_problem()

    def _setup_problem(self, impl=None):
        """Masks, outputs, the C-ABI problem struct and its workspace for the factors this batch holds."""
        eng, R, Kp = self.eng, self.R, self.Kp
        dev = eng.device
        l1_H, l2_H, l1_W, l2_W, eps = self._regs
        self.active = torch.ones(R, device=dev, dtype=torch.uint8)
        self.err2 = torch.zeros(R, device=dev, dtype=torch.float64)
        self.p = KlProblem(
            n_cells=eng.N, n_genes=eng.G, nnz=eng.nnz, row_ptr=ptr(eng.row_ptr), col_idx=ptr(eng.col_idx),
            val=ptr(eng.val), t_row_ptr=ptr(eng.t_row_ptr), t_col_idx=ptr(eng.t_col_idx), t_val=ptr(eng.t_val),
            R=R, K=self.K, Kp=Kp, impl=impl if impl is not None else (eng.impl if eng.impl is not None else kl_impl_for(Kp, R)),
            Hr=ptr(self.Hr), Wt=ptr(self.Wt), active=ptr(self.active),
            # an L1 strength is compared with a numerator, which scales like the factors; an L2 strength multiplies a
            # factor entry against a denominator that scales the same way
            l1_H=l1_H * eng.c, l2_H=l2_H, l1_W=l1_W * eng.c, l2_W=l2_W, eps=eps, sumX=eng.sumX, xlogx=eng.xlogx,
            err2=ptr(self.err2))
        nbytes = check(eng.lib.cnmf_b200_kl_workspace_bytes(C.byref(self.p)), "kl_workspace_bytes")
        self.ws = torch.empty(max(int(nbytes), 256), device=dev, dtype=torch.uint8)
        self.p.workspace, self.p.workspace_bytes = ptr(self.ws), self.ws.numel()
        check(eng.lib.cnmf_b200_kl_range_flag_clear(C.byref(self.p), eng._stream()), "kl_range_flag_clear")
        self.launches = 0

    def subset(self, idx) -> "KlBatch":
        """A new batch holding copies of the replicates ``idx`` of this one (current state, same strengths): what
        ``_run_kl_batch`` continues on once enough replicates of a stack are frozen.  ``absorb`` copies them back."""
        sub = object.__new__(KlBatch)
        sub.eng, sub.Ks, sub.Rs, sub.K, sub.Kp, sub.R = self.eng, None, None, self.K, self.Kp, len(idx)
        ii = torch.as_tensor(np.asarray(idx), device=self.eng.device, dtype=torch.long)
        sub.Hr = self.Hr.index_select(1, ii)
        sub.Wt = self.Wt.index_select(1, ii)
        sub._regs = self._regs
        sub._setup_problem(impl=self.p.impl)   # same kernel family: a replicate's bits do not depend on its neighbours
        return sub

    def absorb(self, sub: "KlBatch", idx):
        """Replicates ``idx`` of this batch take the state of ``sub`` (made by ``subset(idx)``)."""
        ii = torch.as_tensor(np.asarray(idx), device=self.eng.device, dtype=torch.long)
        self.Hr.index_copy_(1, ii, sub.Hr)
        self.Wt.index_copy_(1, ii, sub.Wt)
        self.launches += sub.launches
        self.invalidate()

    def release_workspace(self):
        """Give the workspace back (a stack whose live replicates moved to a subset keeps only its factors)."""
        self.ws = None
        self.p.workspace, self.p.workspace_bytes = None, 0

    def objective(self):
        check(self.eng.lib.cnmf_b200_kl_objective(C.byref(self.p), self.eng._stream()), "kl_objective")
        self.launches += self.eng.lib.cnmf_b200_last_launch_count()
        return self.read_loss()

    def iterate(self, n):
        check(self.eng.lib.cnmf_b200_kl_iterate(C.byref(self.p), int(n), self.eng._stream()), "kl_iterate")
        self.launches += self.eng.lib.cnmf_b200_last_launch_count()
        if n > 0:
            self.p.warm = 1    # the next call reuses the tables and fp16 copies in the workspace (see invalidate)

    def invalidate(self):
        """Call after writing ``Hr`` / ``Wt`` directly: the next iterate rebuilds the fp16 copies from them."""
        self.p.warm = 0

    def check_range(self):
        """Raise if a factor entry (or, on the slab path, a ratio x / (WH)) left the fp16 range of the gathered
        copies since the batch was created -- the affected entries are inf/NaN, nothing was clamped silently."""
        flag = C.c_int32(0)
        check(self.eng.lib.cnmf_b200_kl_range_flag(C.byref(self.p), C.byref(flag), self.eng._stream()), "kl_range_flag")
        if flag.value != 0:
            raise _lib.Cnmfb200Error(
                "sparse KL: a factor entry exceeded 65504 in the engine's units (X scaled to mean ~1): the fp16 copies "
                "the kernels gather cannot hold it. This happens with an initial H0/W0 many orders of magnitude away "
                "from sqrt(mean(X)/K); rescale the initial factors (H*a, W/a leave the product unchanged).")

    def read_loss(self):
        self.check_range()
        return 0.5 * self.err2.cpu().numpy() / (self.eng.c * self.eng.c)

    def set_active(self, alive):
        self.active.copy_(torch.from_numpy(alive.astype(np.uint8)))

    def results(self, return_H=False):
        """(W [R, K, G], H [R, N, K] or None) in the units of the caller's X -- for a mixed-K stack use
        ``results_by_group``."""
        inv = 1.0 / self.eng.c
        W = (self.Wt[:, :, :self.K].permute(1, 2, 0) * inv).contiguous().cpu().numpy()
        H = (self.Hr[:, :, :self.K].permute(1, 0, 2) * inv).contiguous().cpu().numpy() if return_H else None
        return W, H

    def results_by_group(self, return_H=False):
        """[(K, slice of replicates, W [R_k, K, G], H or None)] for every group of the stack."""
        inv, out, lo = 1.0 / self.eng.c, [], 0
        for k, r in zip(self.Ks, self.Rs):
            W = (self.Wt[:, lo:lo + r, :k].permute(1, 2, 0) * inv).contiguous().cpu().numpy()
            H = (self.Hr[:, lo:lo + r, :k].permute(1, 0, 2) * inv).contiguous().cpu().numpy() if return_H else None
            out.append((k, slice(lo, lo + r), W, H))
            lo += r
        return out
