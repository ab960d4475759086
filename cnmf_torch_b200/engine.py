"""Replicate-batched NMF engine (host side).

Holds the count matrix resident on one GPU and runs R replicates of one K at a time through
libcnmf_b200.so.  Replaces what ``cNMF._nmf`` -> ``run_nmf`` does once per replicate
(/root/reference/src/cnmf/cnmf.py:805-821): the reference re-densifies, re-casts and
re-uploads X for every replicate; here X is ingested once and every replicate of a K shares
each pass over it.

torch is used for device memory, streams and host<->device copies only.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field

import numpy as np
import torch

from . import _lib
from ._lib import FrobProblem, check, lib, ptr

EPS_DEFAULT = 1e-16  # cnmf.py:269


def pad4(n: int) -> int:
    return (int(n) + 3) // 4 * 4


def pad_ld(n: int) -> int:
    """Leading dimension: rows start on 128-byte lines (a TMA box row is 128 B; a misaligned row
    pitch costs a fifth more L2 sectors per box, measured on the W half-step)."""
    return (int(n) + 31) // 32 * 32


def init_random(x_mean: float, n: int, g: int, k: int, seed: int):
    """'random' init of run_nmf: avg = sqrt(mean(X)/K), |avg * N(0,1)|, H then W, torch CPU
    generator seeded with ``random_state`` (sklearn _nmf.py:296-309 formula; the seed is the
    ledger's nmf_seed, cnmf.py:885)."""
    gen = torch.Generator(device="cpu")
    gen.manual_seed(int(seed))
    avg = math.sqrt(float(x_mean) / k)
    h = torch.randn((n, k), generator=gen, dtype=torch.float32).mul_(avg).abs_()
    w = torch.randn((k, g), generator=gen, dtype=torch.float32).mul_(avg).abs_()
    return h, w


def draw_init(x_mean, n_total, n, row_offset, g, k, seeds, device, init_device="cpu"):
    """H0 [R, n, k], W0 [R, k, g] for R = len(seeds) replicates: |sqrt(mean(X)/k) * N(0,1)|, H then W from one
    generator seeded with the replicate's ``random_state`` (see ``init_random``).

    ``init_device`` picks WHICH torch generator draws, i.e. which of nmf-torch's two streams the replicates
    start from: "cpu" = the host generator (what the reference draws with use_gpu=False, cnmf.py:461/770, and
    what the oracle draws), written into pinned staging buffers; "cuda" = the device generator (what
    ``torch.manual_seed(seed); torch.randn(..., device='cuda')`` yields, i.e. the reference with use_gpu=True),
    drawn straight into device memory -- no host RNG time and no H2D copy of the initial state.
    A row-sharded engine (n_total != n) draws the GLOBAL H0 and keeps its own rows, so a sharded run starts from
    exactly the state a single-GPU run would."""
    if init_device not in ("cpu", "cuda"):
        raise ValueError(f"init_device={init_device!r}: expected 'cpu' or 'cuda'")
    seeds = [int(s) for s in seeds]
    R = len(seeds)
    avg = math.sqrt(float(x_mean) / k)
    sharded = n_total != n
    if init_device == "cuda":
        f32 = dict(device=device, dtype=torch.float32)
        H0 = torch.empty((R, n, k), **f32)
        W0 = torch.empty((R, k, g), **f32)
        gen = torch.Generator(device=device)
        tmp = torch.empty((n_total, k), **f32) if sharded else None
        for i in range(R):
            gen.manual_seed(seeds[i])
            if sharded:
                torch.randn((n_total, k), generator=gen, out=tmp)
                H0[i].copy_(tmp[row_offset:row_offset + n])
            else:
                torch.randn((n, k), generator=gen, out=H0[i])
            torch.randn((k, g), generator=gen, out=W0[i])
        H0.mul_(avg).abs_()
        W0.mul_(avg).abs_()
        return H0, W0
    pin = torch.cuda.is_available()
    H0 = torch.empty((R, n, k), dtype=torch.float32, pin_memory=pin)
    W0 = torch.empty((R, k, g), dtype=torch.float32, pin_memory=pin)
    for i in range(R):
        if sharded:
            h, w = init_random(x_mean, n_total, g, k, seeds[i])
            H0[i].copy_(h[row_offset:row_offset + n])
            W0[i].copy_(w)
        else:
            # the same draws as init_random, written straight into the pinned staging buffers (about 0.1 ms
            # per replicate at 2700 x 2000, K=10; a thread pool over replicates measured slower than this loop)
            gen = torch.Generator(device="cpu")
            gen.manual_seed(seeds[i])
            torch.randn((n, k), generator=gen, dtype=torch.float32, out=H0[i]).mul_(avg).abs_()
            torch.randn((k, g), generator=gen, dtype=torch.float32, out=W0[i]).mul_(avg).abs_()
    return H0, W0


@dataclass
class SeedInit:
    """Stands in for (H0, W0) in a FrobBatch group: 'draw the random init of these seeds on the device, straight
    into the stacked layout' (cnmf_b200_init_random; bit-identical to draw_init(..., init_device="cuda"))."""
    seeds: list


@dataclass
class BatchResult:
    W: np.ndarray                  # [R, K, G] spectra
    H: np.ndarray | None           # [R, N, K] usages (None unless requested)
    err: np.ndarray                # [R] sqrt(2 * objective)  (run_nmf's third return value)
    n_iter: np.ndarray             # [R] iterations (batch) or passes (online) run
    trace: list = field(default_factory=list)   # [(iteration, loss[R] float64)]
    launches: int = 0              # kernels enqueued by libcnmf_b200 for this batch


class DenseFrobeniusEngine:
    """X resident as dense fp32 [N x G] and its transpose [G x N]; Frobenius MU / HALS."""

    def __init__(self, X, device="cuda:0", impl=_lib.IMPL_AUTO, row_group=None, row_offset=0, init_device="cpu"):
        """``row_group``: torch.distributed group over which the CELLS are sharded (online mode for very
        large N); X is then this rank's block of rows and ||X||^2 / mean(X) are made global.
        ``init_device``: which torch generator draws the 'random' init (see ``draw_init``)."""
        self.lib = lib()  # raises if the CUDA library is missing -- no fallback
        self.impl = impl
        self.init_device = init_device
        self.row_group = row_group
        self.row_offset = int(row_offset)  # first global row of this rank's block
        if not torch.cuda.is_available():
            raise _lib.Cnmfb200Error("CUDA device required: cnmf_torch_b200 has no CPU path")
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        if hasattr(X, "toarray"):
            X = X.toarray()
        xh = torch.as_tensor(np.ascontiguousarray(X)) if not torch.is_tensor(X) else X
        self._x_host = xh   # kept only for host-side inits that need the matrix (NNDSVD)
        self.N, self.G = int(xh.shape[0]), int(xh.shape[1])
        self.ldx, self.ldxt = pad_ld(self.G), pad_ld(self.N)
        xd = xh.to(self.device, dtype=torch.float32, non_blocking=True)
        self.X = torch.zeros((self.N, self.ldx), device=self.device, dtype=torch.float32)
        self.X[:, :self.G] = xd
        self.Xt = torch.zeros((self.G, self.ldxt), device=self.device, dtype=torch.float32)
        self.Xt[:, :self.N] = xd.t()
        del xd
        ws = torch.empty(self.lib.cnmf_b200_sumsq_workspace_bytes(), device=self.device, dtype=torch.uint8)
        out = torch.zeros(2, device=self.device, dtype=torch.float64)
        st = torch.cuda.current_stream().cuda_stream
        # mean(X) for the random init: from the fp32 matrix, fp64 on device
        x_sum = float(self.X[:, :self.G].sum(dtype=torch.float64).item())
        if impl != _lib.IMPL_SIMT:
            # static operands of the tcgen05 path: round once to tf32 (nearest) so that the tensor core's
            # mantissa truncation is unbiased.  The objective is then evaluated for the matrix the tensor
            # cores actually see, ||tf32(X)||^2 (relative difference to ||X||^2 ~1e-7), which keeps the
            # trace identity self-consistent.
            check(self.lib.cnmf_b200_round_tf32(ptr(self.X), self.X.numel(), st), "round_tf32")
            check(self.lib.cnmf_b200_round_tf32(ptr(self.Xt), self.Xt.numel(), st), "round_tf32")
        check(self.lib.cnmf_b200_sumsq(ptr(self.X), self.N, self.G, self.ldx, ptr(out), ptr(ws), st), "sumsq")
        self.normX2 = float(out[0].item())
        self.N_total = self.N
        if row_group is not None:
            from . import dist as _d
            tot = _d.allreduce_scalars([self.normX2, x_sum, float(self.N)], self.device, row_group)
            self.normX2, x_sum, self.N_total = float(tot[0]), float(tot[1]), int(round(tot[2]))
        self.x_mean = x_sum / (float(self.N_total) * self.G)

    def _chunked_xt(self, chunk_size):
        """X^T stored chunk by chunk, [n_chunks, G, ldc] with chunk ci = (rows ci*chunk .. of X)^T: the A operand of the
        chunk's B = h^T x product.  Taken from the global X^T its 128-row TMA boxes would stride over N floats per row
        -- 4 MB at a million cells, every box row on its own page.  Built once per chunk size."""
        cached = getattr(self, "_xtc", None)
        if cached is not None and cached[0] == int(chunk_size):
            return cached[1], cached[2]
        cs = int(chunk_size)
        n_chunks = max(1, (self.N + cs - 1) // cs)
        ldc = pad_ld(min(cs, self.N))
        xtc = torch.zeros((n_chunks, self.G, ldc), device=self.device, dtype=torch.float32)
        for ci in range(n_chunks):
            lo = ci * cs
            c = min(cs, self.N - lo)
            xtc[ci, :, :c] = self.X[lo:lo + c, :self.G].t()
        self._xtc = (cs, xtc, ldc)
        return xtc, ldc

    def host_matrix(self) -> np.ndarray:
        return np.asarray(self._x_host.cpu().numpy() if torch.is_tensor(self._x_host) else self._x_host, dtype=np.float64)

    # ------------------------------------------------------------------------------------
    def make_init(self, K: int, seeds, init_device=None):
        """H0 [R, N, K], W0 [R, K, G] of the 'random' init, one generator per replicate seed (``draw_init``)."""
        return draw_init(self.x_mean, self.N_total, self.N, self.row_offset, self.G, int(K), seeds, self.device,
                         init_device or self.init_device)

    def _init_for(self, K, seeds):
        """(H0, W0) for a FrobBatch: host arrays from the host generator, or a SeedInit marker when the init is
        drawn by the device generator (then it never exists outside the batch's own layout)."""
        if self.init_device == "cuda" and self.N_total * int(K) < 2 ** 31 and int(K) * self.G < 2 ** 31:
            return SeedInit([int(s) for s in seeds]), None
        return self.make_init(K, seeds)

    def _upload_state(self, H0, W0, K, Kp):
        R = int(W0.shape[0])
        H0 = torch.as_tensor(H0, dtype=torch.float32)
        W0 = torch.as_tensor(W0, dtype=torch.float32)
        ldw, ldh = pad_ld(self.G), pad_ld(self.N)
        Wall = torch.zeros((R, Kp, ldw), device=self.device, dtype=torch.float32)
        Ht = torch.zeros((R, Kp, ldh), device=self.device, dtype=torch.float32)
        Wall[:, :K, :self.G] = W0.to(self.device, non_blocking=True)
        Ht[:, :K, :self.N] = H0.to(self.device, non_blocking=True).permute(0, 2, 1)
        return Wall, Ht, ldw, ldh

    def new_batch(self, K: int, seeds=None, H0=None, W0=None, algo="mu", l1_H=0.0, l2_H=0.0, l1_W=0.0,
                  l2_W=0.0, eps=EPS_DEFAULT) -> "FrobBatch":
        """Device-resident state of R replicates of one K, ready to iterate."""
        if H0 is None or W0 is None:
            H0, W0 = self._init_for(K, seeds)
        return FrobBatch(self, K, H0, W0, algo, l1_H, l2_H, l1_W, l2_W, eps)

    def factorize(self, K: int, seeds=None, H0=None, W0=None, algo="mu", tol=1e-4, max_iter=500,
                  l1_H=0.0, l2_H=0.0, l1_W=0.0, l2_W=0.0, eps=EPS_DEFAULT, check_every=10,
                  return_H=False, compact=0.5) -> BatchResult:
        """Batch-mode Frobenius NMF of R replicates (R = len(seeds) or W0.shape[0]).
        Convergence per replicate as run_nmf: every ``check_every`` iterations,
        stop when prev <= cur or |prev - cur| / init < tol; converged replicates are frozen."""
        b = self.new_batch(K, seeds, H0, W0, algo, l1_H, l2_H, l1_W, l2_W, eps)
        return run_batches([b], tol, max_iter, check_every, return_H, compact=compact)[0]

    def factorize_many(self, ks, seeds_by_k, algo="mu", tol=1e-4, max_iter=500, l1_H=0.0, l2_H=0.0, l1_W=0.0,
                       l2_W=0.0, eps=EPS_DEFAULT, check_every=10, return_H=False, compact=0.5):
        """Several K at once (the shape of a cNMF run: K = 5..13 x 100 replicates).  The K-batches are
        independent, so their kernels run on separate streams and overlap, their inits are drawn by a
        thread pool while earlier batches are already uploading, and convergence is checked for all of
        them in lock-step (one small D2H per batch per check period).  Returns {K: BatchResult}."""
        from concurrent.futures import ThreadPoolExecutor
        ks = [int(k) for k in ks]
        stacks = {}
        for k in ks:                       # K values with the same padded Kp share one stack
            stacks.setdefault(pad4(k), []).append(k)
        # host draws: a few threads, so that earlier batches upload while later ones are drawn; device draws are
        # just kernel launches on the current stream
        with ThreadPoolExecutor(max_workers=1 if self.init_device == "cuda" else max(1, min(len(ks), 3))) as pool:
            futs = {k: pool.submit(self._init_for, k, seeds_by_k[k]) for k in ks}
            batches = []
            for kp, group in stacks.items():
                groups = [(k,) + tuple(futs[k].result()) for k in group]
                batches.append(FrobBatch(self, None, None, None, algo, l1_H, l2_H, l1_W, l2_W, eps, groups=groups))
        streams = [torch.cuda.Stream(device=self.device) for _ in batches]
        res = run_batches(batches, tol, max_iter, check_every, return_H, streams, compact)
        out = {}
        for b, rb in zip(batches, res):
            for k, sl, W, H in b.results_by_group(return_H):
                out[k] = BatchResult(W=W, H=H, err=rb.err[sl], n_iter=rb.n_iter[sl],
                                     trace=[(it, v[sl]) for it, v in rb.trace], launches=rb.launches)
        return out

    # ------------------------------------------------------------------------------------
    def factorize_online_many(self, ks, seeds_by_k, **kw):
        """Online mode for several K: K values that share a padded Kp run as ONE stack (as ``factorize_many``
        does in batch mode).  Returns {K: BatchResult}."""
        stacks = {}
        for k in [int(k) for k in ks]:
            stacks.setdefault(pad4(k), []).append(k)
        out = {}
        for group in stacks.values():
            groups = [(k,) + tuple(self._init_for(k, seeds_by_k[k])) for k in group]
            out.update(self.factorize_online(None, groups=groups, **kw))
        return out

    def factorize_online(self, K: int, seeds=None, H0=None, W0=None, tol=1e-4, max_pass=20, chunk_size=5000,
                         chunk_max_iter=200, h_tol=0.05, w_tol=0.05, l1_H=0.0, l2_H=0.0, l1_W=0.0, l2_W=0.0,
                         eps=EPS_DEFAULT, return_H=False, inner_batch=4, groups=None, profile=None):
        """Online-mode Frobenius MU of R replicates (run_nmf(mode='online'), what cNMF configures at
        cnmf.py:757-771).  Per pass and per chunk of ``chunk_size`` cells: inner MU loop on h against the
        cached x W^T (stopping rule of cnmf.py:365-378), A += h^T h, B += h^T x, inner MU loop on W against
        (A, B); then H is refreshed on all cells and the objective decides convergence per replicate.
        With ``row_group`` the cells are sharded over ranks: A, B and the objective terms are summed with
        one all-reduce per chunk; W stays replicated (every rank runs the same W loop on the same sums).
        ``profile`` (a dict): device time spent in the all-reduces of [A | B] is added to profile["allreduce_ms"]
        (CUDA events around every collective; read after the last pass)."""
        from . import dist as _d
        if self.impl == _lib.IMPL_SIMT:
            raise NotImplementedError("online mode runs on the tcgen05 path (impl auto/tcgen05)")
        if chunk_size % 4 != 0:
            raise ValueError("online_chunk_size must be a multiple of 4 (16-byte aligned chunk operands)")
        L, dev, grp = self.lib, self.device, self.row_group
        if groups is None and (H0 is None or W0 is None):
            H0, W0 = self._init_for(K, seeds)
        b = FrobBatch(self, K, H0, W0, "mu", l1_H, l2_H, l1_W, l2_W, eps, groups=groups)
        R, Kp, N, G = b.R, b.Kp, self.N, self.G
        rk = R * Kp
        ldw, ldh = b.p.ldw, b.p.ldh
        _stream_now = FrobBatch._stream()     # one lookup per call (a torch.cuda.current_stream() costs ~14 us; a small run
        st = lambda: _stream_now              # made 18 of them): the call runs on the stream current at entry
        cmax = max(1, min(chunk_size, N))
        ldp = pad_ld(cmax)
        f32 = dict(device=dev, dtype=torch.float32)
        Pt = torch.zeros((rk, ldp), **f32)
        ABst = torch.zeros(R * Kp * Kp + rk * ldw, **f32)     # the pass's statistics [A | B], same layout as [Ac | Bc]
        A = ABst[:R * Kp * Kp].view(R, Kp, Kp)
        Bst = ABst[R * Kp * Kp:].view(rk, ldw)
        # the chunk's statistics [Ac | Bc] in ONE buffer: the message of the all-reduce, no packing copies
        AcBc = torch.zeros(R * Kp * Kp + rk * ldw, **f32)
        Ac = AcBc[:R * Kp * Kp].view(R, Kp, Kp)
        Bc = AcBc[R * Kp * Kp:].view(rk, ldw)
        Wop = torch.zeros((rk, ldw), **f32)
        Wlo = torch.zeros((rk, ldw), **f32)        # tf32(W - tf32 W): second GEMM pass of the refresh / objective
        Hs = torch.zeros((rk, ldp), **f32)         # the chunk's usages, compact (fp32 / tf32 copy)
        Hopc = torch.zeros((rk, ldp), **f32)
        cross = torch.zeros(R, device=dev, dtype=torch.float64)
        ws_gram = torch.empty(max(256, int(L.cnmf_b200_gram_workspace_bytes(R, Kp, max(cmax, G)))), device=dev,
                              dtype=torch.uint8)
        Wall, Ht, active = b.Wall, b.Ht, b.active
        n_chunks = _d.allreduce_max_int((N + chunk_size - 1) // chunk_size, dev, grp)

        iters_hw = torch.zeros((2, R), device=dev, dtype=torch.int32)     # inner iterations per replicate: h loops, W loops
        ws_loop = torch.empty(max(256, int(L.cnmf_b200_online_loop_workspace_bytes(R))), device=dev, dtype=torch.uint8)
        xtc, ldc = self._chunked_xt(chunk_size)      # X^T chunk by chunk: [n_local_chunks, G, ldc]
        # one library call per chunk phase (cnmf_b200_online_chunk): the dozen kernels of a chunk are enqueued from C,
        # and the inner loops stop on the device -- nothing in a pass waits for the host
        ca = _lib.OnlineChunkArgs(
            ldx=self.ldx, ldxt=ldc, n_genes=G, R=R, Kp=Kp, Wall=ptr(Wall), Wop=ptr(Wop), ldw=ldw, ldh=ldh, Hs=ptr(Hs),
            Hopc=ptr(Hopc), Pt=ptr(Pt), ldp=ldp, Wlo=ptr(Wlo), cross=ptr(cross), gramH_acc=ptr(b.gramH),
            gramW=ptr(b.gramW), A=ptr(A), Ac=ptr(Ac), Bst=ptr(Bst), Bc=ptr(Bc), active=ptr(active), l1_H=l1_H,
            l2_H=l2_H, l1_W=l1_W, l2_W=l2_W, eps=eps, h_tol=h_tol, w_tol=w_tol, max_iter=int(chunk_max_iter),
            iters_h=ptr(iters_hw[0]), iters_w=ptr(iters_hw[1]), ws_loop=ptr(ws_loop), ws_gram=ptr(ws_gram))
        x_ptr, ht_ptr, xtc_ptr = ptr(self.X), ptr(Ht), ptr(xtc)
        xtc_stride = 4 * G * ldc

        def chunk_phase(ci, lo, c, phase):
            ca.x, ca.xt, ca.c = x_ptr + 4 * lo * self.ldx, xtc_ptr + ci * xtc_stride, c
            ca.Ht = ht_ptr + 4 * lo
            check(L.cnmf_b200_online_chunk(C.byref(ca), phase, st()), "online_chunk")
            b.launches += L.cnmf_b200_last_launch_count()

        def refresh_and_objective(iterate: bool):
            """H on every cell against the current W (``iterate``; else evaluation only) and the objective by the trace
            identity ||X||^2 - 2 <H, X W^T> + <W W^T, H^T H>, its data-dependent terms gathered chunk by chunk on the
            way: no separate pass over X (at a million cells the frozen half-step the batch path uses costs 60 ms)."""
            cross.zero_()
            b.gramH.zero_()
            ca.max_iter = int(chunk_max_iter) if iterate else 0
            for ci in range(n_chunks):
                lo = ci * chunk_size
                c = max(0, min(chunk_size, N - lo))
                if c > 0:
                    chunk_phase(ci, lo, c, _lib.ONLINE_REFRESH)
            ca.max_iter = int(chunk_max_iter)
            sums = torch.zeros((R, 3), device=dev, dtype=torch.float64)
            sums[:, 0] = cross
            if l1_W != 0.0:
                sums[:, 1] = Wall[:, :, :G].sum(dim=(1, 2), dtype=torch.float64)
            if l1_H != 0.0:
                sums[:, 2] = Ht[:, :, :N].sum(dim=(1, 2), dtype=torch.float64)
            if grp is not None:
                sw = sums[:, 1].clone()
                _d.allreduce_sum_([sums, b.gramH], grp)
                sums[:, 1] = sw                   # W is replicated: its sum must not be multiplied by P
            check(L.cnmf_b200_frob_objective_combine(self.normX2, ptr(sums), ptr(b.gramW), ptr(b.gramH), R, Kp,
                                                     l1_H, l2_H, l1_W, l2_W, ptr(b.err2), st()), "objective_combine")
            return b.read_loss()

        ar_events = []
        ca.max_iter = 0                                   # W's tf32 copy, its residual and Gram; W itself untouched
        chunk_phase(0, 0, 1, _lib.ONLINE_SPECTRA)
        ca.max_iter = int(chunk_max_iter)
        init = refresh_and_objective(False)
        prev, cur = init.copy(), init
        trace = [(0, init.copy())]
        n_pass = np.full(R, max_pass, dtype=np.int64)
        alive = np.ones(R, dtype=bool)
        for ps in range(1, max_pass + 1):
            ABst.zero_()
            for ci in range(n_chunks):
                lo = ci * chunk_size
                c = max(0, min(chunk_size, N - lo))
                if c > 0:
                    chunk_phase(ci, lo, c, _lib.ONLINE_USAGE)
                else:                                     # a rank that has run out of cells contributes zeros
                    Ac.zero_()
                    Bc.zero_()
                if profile is not None and grp is not None:
                    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                    ev[0].record()
                    _d.allreduce_flat_(AcBc, grp)
                    ev[1].record()
                    ar_events.append(ev)
                else:
                    _d.allreduce_flat_(AcBc, grp)         # the only data-path collective
                chunk_phase(ci, lo, max(c, 1), _lib.ONLINE_SPECTRA)
            # end of pass: H on every cell against the final W.  (A frozen replicate is skipped by every kernel: its
            # objective is the one it was frozen at.)
            cur = np.where(alive, refresh_and_objective(True), cur)
            trace.append((ps, cur.copy()))
            conv = alive & ((prev <= cur) | (np.abs((prev - cur) / init) < tol))
            if conv.any():
                n_pass[conv] = ps
                alive &= ~conv
                b.set_active(alive)
            prev = np.where(alive, cur, prev)
            if not alive.any():
                break
        err = np.sqrt(np.maximum(2.0 * cur, 0.0))
        if profile is not None and ar_events:
            torch.cuda.synchronize()
            profile["allreduce_ms"] = profile.get("allreduce_ms", 0.0) + float(sum(a.elapsed_time(b) for a, b in ar_events))
        it_host = iters_hw.cpu().numpy()
        inner = {"h": int(it_host[0].max()), "w": int(it_host[1].max())}   # of the replicate that iterated most
        if groups is not None:   # a stack of several K: one result per K
            out = {}
            for k, sl, W, H in b.results_by_group(return_H):
                out[k] = BatchResult(W=W, H=H, err=err[sl], n_iter=n_pass[sl], trace=[(it, v[sl]) for it, v in trace],
                                     launches=int(b.launches))
                out[k].inner_iterations = inner
            return out
        W, H = b.results(return_H)
        res = BatchResult(W=W, H=H, err=err, n_iter=n_pass, trace=trace, launches=int(b.launches))
        res.inner_iterations = inner
        return res


def run_batches(batches, tol, max_iter, check_every=10, return_H=False, streams=None, compact=0.5):
    """run_nmf's batch loop for a list of resident batches in lock-step: every ``check_every`` iterations the
    objective of every replicate is read back (R doubles per batch) and replicates with
    prev <= cur or |prev - cur| / init < tol are frozen (their state stops changing on the device).

    A frozen replicate still rides through the GEMM of its tile, and the replicates of a stack stop at very different
    checks; once no more than ``compact`` of a working batch is alive (half: a step of this path is a few ms, so a
    compaction -- new buffers, Grams, tensor maps -- must buy a halved GEMM) and at least 16 replicates are frozen the
    live ones are copied into a smaller batch (``FrobBatch.subset``) and the loop continues there; their factors are
    copied back at the end.  A replicate's arithmetic depends on its neighbours in the stack only through the tile
    shape (fp32 summation order, tests/test_gpu_properties.py), so the results are those of the uncompacted run to
    that rounding (``test_cfg2_compaction_matches_uncompacted_run``).  ``compact=0`` turns it off."""
    st = [dict(init=b.objective()) for b in batches]
    for s, b in zip(st, batches):
        s.update(prev=s["init"].copy(), cur=s["init"].copy(), trace=[(0, s["init"].copy())],
                 n_iter=np.full(b.R, max_iter, dtype=np.int64), alive=np.ones(b.R, dtype=bool),
                 wb=b, widx=np.arange(b.R))      # the working batch and its replicates' positions in b
    it = 0
    while it < max_iter and any(s["alive"].any() for s in st):
        step = min(check_every, max_iter - it)
        live = [i for i, s in enumerate(st) if s["alive"].any()]
        iterate_concurrently([st[i]["wb"] for i in live], step, [streams[i] for i in live] if streams else None)
        it += step
        for i in live:
            s, b, wb = st[i], batches[i], st[i]["wb"]
            s["cur"][s["widx"]] = wb.read_loss()
            cur = s["cur"]
            s["trace"].append((it, cur.copy()))
            if it % check_every == 0:
                conv = s["alive"] & ((s["prev"] <= cur) | (np.abs((s["prev"] - cur) / s["init"]) < tol))
                if conv.any():
                    s["n_iter"][conv] = it
                    s["alive"] &= ~conv
                    wb.set_active(s["alive"][s["widx"]])
                s["prev"] = np.where(s["alive"], cur, s["prev"])
                n_alive = int(s["alive"].sum())
                if compact > 0 and 0 < n_alive <= compact * wb.R and wb.R - n_alive >= 16 and it < max_iter:
                    if wb is not b:
                        b.absorb(wb, s["widx"])
                    else:
                        b.release_workspace()
                    s["widx"] = np.flatnonzero(s["alive"])
                    s["wb"] = b.subset(s["widx"])
    out = []
    for s, b in zip(st, batches):
        if s["wb"] is not b:
            b.absorb(s["wb"], s["widx"])
            s["wb"] = None
        W, H = b.results(return_H) if len(b.Ks) == 1 else (None, None)   # mixed-K stacks: results_by_group
        out.append(BatchResult(W=W, H=H, err=np.sqrt(np.maximum(2.0 * s["cur"], 0.0)), n_iter=s["n_iter"],
                               trace=s["trace"], launches=int(b.launches)))
    return out


def iterate_concurrently(batches, n: int, streams=None):
    """Enqueue ``n`` iterations of several independent batches (different K) on side streams that fork
    from / join into the current stream, so that their kernels overlap: one K-batch alone leaves SMs idle
    (its persistent GEMM has fewer tiles than SMs and its small Gram kernels are latency-bound)."""
    cur = torch.cuda.current_stream()
    if streams is None or len(batches) == 1:
        for b in batches:
            b.iterate(n)
        return
    fork = torch.cuda.Event()
    fork.record(cur)
    for b, s in zip(batches, streams):
        s.wait_event(fork)
        with torch.cuda.stream(s):
            b.iterate(n)
        done = torch.cuda.Event()
        done.record(s)
        cur.wait_event(done)


class FrobBatch:
    """Replicates resident on the device (stacked Wall / Ht, Grams, workspace).  Usually R replicates of one
    K; ``groups`` = [(K, H0, W0), ...] stacks replicates of DIFFERENT K that share the same padded Kp (e.g.
    K = 9, 10, 11, 12 -> Kp = 12): the kernels only see Kp (padding rows are zero and stay zero), and a
    taller stack means more tiles per launch -- better SM fill, and the epilogue of one tile overlaps the
    main loop of the next."""

    def __init__(self, eng: DenseFrobeniusEngine, K, H0, W0, algo, l1_H, l2_H, l1_W, l2_W, eps, groups=None):
        if groups is None:
            groups = [(int(K), H0, W0)]
        self.eng = eng
        self.Ks = [int(g[0]) for g in groups]
        self.Rs = [len(g[1].seeds) if isinstance(g[1], SeedInit) else int(g[2].shape[0]) for g in groups]
        self.K = max(self.Ks)
        self.R = sum(self.Rs)
        self.Kp = pad4(self.K)
        if self.Kp > _lib.MAX_KP:
            raise ValueError(f"n_components={self.K} exceeds the engine limit {_lib.MAX_KP}")
        if any(pad4(k) != self.Kp for k in self.Ks):
            raise ValueError("all K of one stack must round up to the same multiple of 4")
        algo_id = {"mu": _lib.ALGO_MU, "hals": _lib.ALGO_HALS}[algo]
        dev = eng.device
        R, Kp = self.R, self.Kp
        if all(isinstance(g[1], SeedInit) for g in groups):
            # drawn on the device straight into the stacked layout: one launch per K of the stack
            ldw, ldh = pad_ld(eng.G), pad_ld(eng.N)
            self.Wall = torch.zeros((R, Kp, ldw), device=dev, dtype=torch.float32)
            self.Ht = torch.zeros((R, Kp, ldh), device=dev, dtype=torch.float32)
            lo = 0
            for (k, si, _), r in zip(groups, self.Rs):
                seeds = torch.tensor(si.seeds, dtype=torch.int64, device=dev)
                check(eng.lib.cnmf_b200_init_random(ptr(seeds), r, k, Kp, eng.N_total, eng.N, eng.row_offset, eng.G,
                                                    math.sqrt(float(eng.x_mean) / k), ptr(self.Ht[lo:]), ldh,
                                                    ptr(self.Wall[lo:]), ldw, self._stream()), "init_random")
                lo += r
        else:
            parts = [eng._upload_state(h0, w0, k, self.Kp) for k, h0, w0 in groups]
            ldw, ldh = parts[0][2], parts[0][3]
            self.Wall = parts[0][0] if len(parts) == 1 else torch.cat([q[0] for q in parts], 0)
            self.Ht = parts[0][1] if len(parts) == 1 else torch.cat([q[1] for q in parts], 0)
            del parts
        # the true K of every replicate of a mixed-K stack: the tensor-core epilogue works on those components only
        self.rep_k = None
        if any(k != Kp for k in self.Ks):
            self.rep_k = torch.from_numpy(np.repeat(np.asarray(self.Ks, dtype=np.uint8), self.Rs)).to(dev)
        self._regs = (algo_id, l1_H, l2_H, l1_W, l2_W, eps, ldw, ldh)
        self._setup_problem()

    def _setup_problem(self):
        """Grams, masks, outputs, the C-ABI problem struct and its workspace for the factors this batch holds."""
        eng, R, Kp = self.eng, self.R, self.Kp
        dev = eng.device
        algo_id, l1_H, l2_H, l1_W, l2_W, eps, ldw, ldh = self._regs
        self.gramW = torch.zeros((R, Kp, Kp), device=dev, dtype=torch.float32)
        self.gramH = torch.zeros((R, Kp, Kp), device=dev, dtype=torch.float32)
        self.active = torch.ones(R, device=dev, dtype=torch.uint8)
        self.err2 = torch.zeros(R, device=dev, dtype=torch.float64)
        self.p = FrobProblem(
            n_cells=eng.N, n_genes=eng.G, X=ptr(eng.X), ldx=eng.ldx, Xt=ptr(eng.Xt), ldxt=eng.ldxt,
            R=R, K=self.K, Kp=Kp, algo=algo_id, impl=eng.impl, Wall=ptr(self.Wall), ldw=ldw,
            Ht=ptr(self.Ht), ldh=ldh, gramW=ptr(self.gramW), gramH=ptr(self.gramH),
            active=ptr(self.active), l1_H=l1_H, l2_H=l2_H, l1_W=l1_W, l2_W=l2_W, eps=eps,
            normX2=eng.normX2, err2=ptr(self.err2), rep_k=ptr(self.rep_k))
        nbytes = check(eng.lib.cnmf_b200_frob_workspace_bytes(C.byref(self.p)), "frob_workspace_bytes")
        self.ws = torch.empty(max(int(nbytes), 256), device=dev, dtype=torch.uint8)
        self.p.workspace, self.p.workspace_bytes = ptr(self.ws), self.ws.numel()
        self.launches = 0
        check(eng.lib.cnmf_b200_frob_prepare(C.byref(self.p), self._stream()), "frob_prepare")

    def subset(self, idx) -> "FrobBatch":
        """A new batch holding copies of the replicates ``idx`` of this one (current state, same settings): what
        ``run_batches`` continues on once enough replicates of a stack are frozen.  ``absorb`` copies them back."""
        sub = object.__new__(FrobBatch)
        sub.eng, sub.Ks, sub.Rs, sub.K, sub.Kp, sub.R = self.eng, None, None, self.K, self.Kp, len(idx)
        ii = torch.as_tensor(np.asarray(idx), device=self.eng.device, dtype=torch.long)
        sub.Wall = self.Wall.index_select(0, ii)
        sub.Ht = self.Ht.index_select(0, ii)
        sub.rep_k = None if self.rep_k is None else self.rep_k.index_select(0, ii)
        sub._regs = self._regs
        sub._setup_problem()
        return sub

    def absorb(self, sub: "FrobBatch", idx) -> None:
        """Replicates ``idx`` of this batch take the factors of ``sub`` (made by ``subset(idx)``)."""
        ii = torch.as_tensor(np.asarray(idx), device=self.eng.device, dtype=torch.long)
        self.Wall.index_copy_(0, ii, sub.Wall)
        self.Ht.index_copy_(0, ii, sub.Ht)
        self.launches += sub.launches

    def release_workspace(self) -> None:
        """Give the workspace back (a stack whose live replicates moved to a subset keeps only its factors)."""
        self.ws = None
        self.p.workspace, self.p.workspace_bytes = None, 0

    @staticmethod
    def _stream():
        return torch.cuda.current_stream().cuda_stream

    def objective(self) -> np.ndarray:
        """loss[R] of the current state (factors untouched)."""
        check(self.eng.lib.cnmf_b200_frob_objective(C.byref(self.p), self._stream()), "frob_objective")
        self.launches += self.eng.lib.cnmf_b200_last_launch_count()
        return self.read_loss()

    def iterate(self, n: int) -> None:
        """Enqueue n iterations (+ objective) on the current stream; asynchronous."""
        check(self.eng.lib.cnmf_b200_frob_iterate(C.byref(self.p), int(n), self._stream()), "frob_iterate")
        self.launches += self.eng.lib.cnmf_b200_last_launch_count()

    def read_loss(self) -> np.ndarray:
        return 0.5 * self.err2.cpu().numpy()

    def set_active(self, alive: np.ndarray) -> None:
        self.active.copy_(torch.from_numpy(alive.astype(np.uint8)))

    @staticmethod
    def _to_host(views):
        """Device views -> host arrays through pinned staging buffers (torch's caching host allocator), all copies
        enqueued before the one synchronisation; pageable destinations cost ~7x more at 45 MB of spectra."""
        bufs = []
        for v in views:
            if v is None:
                bufs.append(None)
                continue
            b = torch.empty(v.shape, dtype=v.dtype, pin_memory=True)
            b.copy_(v, non_blocking=True)
            bufs.append(b)
        torch.cuda.current_stream().synchronize()
        return [None if b is None else b.numpy() for b in bufs]

    def results(self, return_H=False):
        """(W [R, K, G], H [R, N, K] or None) -- for a mixed-K stack use ``results_by_group``."""
        eng = self.eng
        W, H = self._to_host([self.Wall[:, :self.K, :eng.G],
                              self.Ht[:, :self.K, :eng.N].permute(0, 2, 1) if return_H else None])
        return W, H

    def results_by_group(self, return_H=False):
        """[(K, slice of replicates, W [R_k, K, G], H or None)] for every group of the stack."""
        eng, views, lo = self.eng, [], 0
        for k, r in zip(self.Ks, self.Rs):
            views.append(self.Wall[lo:lo + r, :k, :eng.G])
            views.append(self.Ht[lo:lo + r, :k, :eng.N].permute(0, 2, 1) if return_H else None)
            lo += r
        host, out, lo = self._to_host(views), [], 0
        for i, (k, r) in enumerate(zip(self.Ks, self.Rs)):
            out.append((k, slice(lo, lo + r), host[2 * i], host[2 * i + 1]))
            lo += r
        return out
