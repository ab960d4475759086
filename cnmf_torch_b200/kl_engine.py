"""Replicate-batched sparse KL engine (host side): CSR X and CSR X^T resident on one GPU, R
replicates of one K interleaved in Hr [N x R*Kp] / Wt [G x R*Kp].  Replaces the per-replicate
``run_nmf(beta_loss='kullback-leibler')`` of /root/reference/src/cnmf/cnmf.py:805-821."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import scipy.sparse as sp
import torch

from . import _lib
from ._lib import KlProblem, check, lib, ptr
from .engine import EPS_DEFAULT, BatchResult, draw_init


def pad2(k: int) -> int:
    """The KL kernels take any even Kp: every padded component is a gathered value and two FMAs per stored entry."""
    m = 4 if os.environ.get("CNMF_B200_KL_PAD4") else 2      # (dev knob: the multiple-of-4 padding of the dense path)
    return (int(k) + m - 1) // m * m


class SparseKLEngine:
    def __init__(self, X, device="cuda:0", init_device="cpu"):
        self.lib = lib()
        self.init_device = init_device   # which torch generator draws the 'random' init (engine.draw_init)
        if not torch.cuda.is_available():
            raise _lib.Cnmfb200Error("CUDA device required: cnmf_torch_b200 has no CPU path")
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        if torch.is_tensor(X):
            X = X.cpu().numpy()
        csr = sp.csr_matrix(X, dtype=np.float32)
        self._csr_host = csr
        csr.eliminate_zeros()
        csr.sort_indices()
        if csr.nnz and float(csr.data.min()) < 0.0:
            raise ValueError("X must be non-negative")   # the kernels rely on stored entries being > 0
        csc = csr.T.tocsr()
        csc.sort_indices()
        self.N, self.G = csr.shape
        self.nnz = int(csr.nnz)
        if self.nnz >= 2 ** 31:
            raise ValueError("nnz must fit 32-bit offsets")
        dev = self.device
        up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a.astype(dt))).to(dev)
        self.row_ptr, self.col_idx, self.val = up(csr.indptr, np.int32), up(csr.indices, np.int32), up(csr.data, np.float32)
        self.t_row_ptr, self.t_col_idx, self.t_val = up(csc.indptr, np.int32), up(csc.indices, np.int32), up(csc.data, np.float32)
        out = torch.zeros(2, device=dev, dtype=torch.float64)
        check(self.lib.cnmf_b200_kl_constants(ptr(self.val), self.nnz, ptr(out), self._stream()), "kl_constants")
        o = out.cpu().numpy()
        self.sumX, self.xlogx = float(o[0]), float(o[1])
        self.x_mean = self.sumX / (float(self.N) * float(self.G))

    @staticmethod
    def _stream():
        return torch.cuda.current_stream().cuda_stream

    def host_matrix(self) -> np.ndarray:
        return np.asarray(self._csr_host.toarray(), dtype=np.float64)

    def make_init(self, K, seeds, init_device=None):
        return draw_init(self.x_mean, self.N, self.N, 0, self.G, int(K), seeds, self.device,
                         init_device or self.init_device)

    def new_batch(self, K, seeds=None, H0=None, W0=None, l1_H=0.0, l2_H=0.0, l1_W=0.0, l2_W=0.0, eps=EPS_DEFAULT):
        if H0 is None or W0 is None:
            H0, W0 = self.make_init(K, seeds)
        return KlBatch(self, K, H0, W0, l1_H, l2_H, l1_W, l2_W, eps)

    def factorize(self, K, seeds=None, H0=None, W0=None, algo="mu", tol=1e-4, max_iter=500, l1_H=0.0, l2_H=0.0,
                  l1_W=0.0, l2_W=0.0, eps=EPS_DEFAULT, check_every=10, return_H=False) -> BatchResult:
        """Batch-mode KL MU of R replicates; same stopping rule as the Frobenius engine."""
        if algo != "mu":
            raise NotImplementedError("only multiplicative updates exist for beta_loss='kullback-leibler'")
        b = self.new_batch(K, seeds, H0, W0, l1_H, l2_H, l1_W, l2_W, eps)
        R = b.R
        init = b.objective()
        prev, cur = init.copy(), init
        trace = [(0, init.copy())]
        n_iter = np.full(R, max_iter, dtype=np.int64)
        alive = np.ones(R, dtype=bool)
        it = 0
        while it < max_iter and alive.any():
            step = min(check_every, max_iter - it)
            b.iterate(step)
            it += step
            cur = b.read_loss()
            trace.append((it, cur.copy()))
            if it % check_every == 0:
                conv = alive & ((prev <= cur) | (np.abs((prev - cur) / init) < tol))
                if conv.any():
                    n_iter[conv] = it
                    alive &= ~conv
                    b.set_active(alive)
                prev = np.where(alive, cur, prev)
        W, H = b.results(return_H)
        return BatchResult(W=W, H=H, err=np.sqrt(np.maximum(2.0 * cur, 0.0)), n_iter=n_iter, trace=trace,
                           launches=int(b.launches))


class KlBatch:
    def __init__(self, eng: SparseKLEngine, K, H0, W0, l1_H, l2_H, l1_W, l2_W, eps):
        self.eng, self.K = eng, int(K)
        self.R, self.Kp = int(W0.shape[0]), pad2(K)
        if self.Kp > _lib.MAX_KP:
            raise ValueError(f"n_components={K} exceeds the engine limit {_lib.MAX_KP}")
        dev = eng.device
        R, Kp = self.R, self.Kp
        H0 = torch.as_tensor(H0, dtype=torch.float32).to(dev)   # [R, N, K]
        W0 = torch.as_tensor(W0, dtype=torch.float32).to(dev)   # [R, K, G]
        self.Hr = torch.zeros((eng.N, R, Kp), device=dev, dtype=torch.float32)
        self.Wt = torch.zeros((eng.G, R, Kp), device=dev, dtype=torch.float32)
        self.Hr[:, :, :K] = H0.permute(1, 0, 2)
        self.Wt[:, :, :K] = W0.permute(2, 0, 1)
        self.active = torch.ones(R, device=dev, dtype=torch.uint8)
        self.err2 = torch.zeros(R, device=dev, dtype=torch.float64)
        self.p = KlProblem(
            n_cells=eng.N, n_genes=eng.G, nnz=eng.nnz, row_ptr=ptr(eng.row_ptr), col_idx=ptr(eng.col_idx),
            val=ptr(eng.val), t_row_ptr=ptr(eng.t_row_ptr), t_col_idx=ptr(eng.t_col_idx), t_val=ptr(eng.t_val),
            R=R, K=self.K, Kp=Kp, Hr=ptr(self.Hr), Wt=ptr(self.Wt), active=ptr(self.active), l1_H=l1_H, l2_H=l2_H,
            l1_W=l1_W, l2_W=l2_W, eps=eps, sumX=eng.sumX, xlogx=eng.xlogx, err2=ptr(self.err2))
        nbytes = check(eng.lib.cnmf_b200_kl_workspace_bytes(C.byref(self.p)), "kl_workspace_bytes")
        self.ws = torch.empty(max(int(nbytes), 256), device=dev, dtype=torch.uint8)
        self.p.workspace, self.p.workspace_bytes = ptr(self.ws), self.ws.numel()
        self.launches = 0

    def objective(self):
        check(self.eng.lib.cnmf_b200_kl_objective(C.byref(self.p), self.eng._stream()), "kl_objective")
        self.launches += self.eng.lib.cnmf_b200_last_launch_count()
        return self.read_loss()

    def iterate(self, n):
        check(self.eng.lib.cnmf_b200_kl_iterate(C.byref(self.p), int(n), self.eng._stream()), "kl_iterate")
        self.launches += self.eng.lib.cnmf_b200_last_launch_count()

    def read_loss(self):
        return 0.5 * self.err2.cpu().numpy()

    def set_active(self, alive):
        self.active.copy_(torch.from_numpy(alive.astype(np.uint8)))

    def results(self, return_H=False):
        W = self.Wt[:, :, :self.K].permute(1, 2, 0).contiguous().cpu().numpy()
        H = self.Hr[:, :, :self.K].permute(1, 0, 2).contiguous().cpu().numpy() if return_H else None
        return W, H
