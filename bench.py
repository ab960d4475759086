"""bench.py -- NMF replicate-iterations/s of the cNMF factorize hot path on B200.

Workload at every N: BASELINE.json configs[1] ("PBMC-3k shape 2.7k x 2k HVG, K=7-12, 100 replicates,
frobenius HALS"), synthetic negative-binomial counts, variance-scaled as the reference does
(SURVEY.md 8d).  One STEP = one convergence-check period of run_nmf's batch loop -- 10 iterations
(H update + W update each) plus the objective -- for all 100 replicates of every K in 7..12, i.e.
6000 replicate-iterations per GPU per step.  With N GPUs every rank runs the whole workload on its own
replicate seeds (replicates are independent units; no data-path collective) => weak scaling.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config cfg2|cfg3|cfg5]

--config (default cfg2, the line the driver records):
  cfg3  BASELINE.json configs[2]: 100 000 x 2000 CSR, K=5..13 x 100 replicates, KL MU.  ONE ledger of 900 jobs split over
        the ranks with the reference's stride rule (worker_filter, cnmf.py:53-54) => strong scaling; a step = 10
        iterations + objective of every job of the rank.
  cfg5  BASELINE.json configs[4]: 1 000 000 x 2000 dense, online mode, chunk 5000, K=15 x 50 replicates, cells
        row-sharded over the ranks with one NCCL all-reduce of the chunk statistics [A | B] per step; a step = one
        pass over all cells (nmf-torch counts passes as the iterations of online mode).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_CELLS, N_GENES = 2700, 2000
KS = list(range(7, 13))
REPLICATES = 100
ITERS_PER_STEP = 10
ALGO = "hals"
E2E_ITERS = 100  # iterations per replicate in one end-to-end public-API call
KL_PROFILE_CSV = "r01_kl_half_step_full_raw.csv"   # the gather kernels are round 1's; same code
KL_SLAB_PROFILE_CSV = "r02_kl_slab_d_raw.csv"
TC_PROFILE_CSV = "r02_tc_half_step_full_raw.csv"   # ncu --set full raw page of the dominant kernel, this round
METRIC = "nmf_replicate_iterations_per_second"
UNIT = "replicate-iters/s"
WORKLOAD = ("cfg2: 2700x2000 NB counts (variance-scaled), K=7..12 x 100 replicates, frobenius HALS, "
            "batch mode, 10 iterations + objective per step")


CFG3 = dict(cells=100000, genes=2000, ks=list(range(5, 14)), replicates=100, libsize=170.0, e2e_iters=20)
CFG5 = dict(cells=1000000, genes=2000, k=15, replicates=50, chunk=5000)
WORKLOAD3 = ("cfg3: 100000x2000 CSR NB counts (~8 % stored, variance-scaled), K=5..13 x 100 replicates = ONE ledger of "
             "900 jobs strided over the ranks (worker_filter), beta_loss=kullback-leibler MU, batch mode, 10 iterations "
             "+ objective per step")
WORKLOAD5 = ("cfg5: 1000000x2000 dense NB counts, online mode, chunk 5000, K=15 x 50 replicates, cells row-sharded over "
             "the ranks, one all-reduce of the chunk statistics [A | B] per chunk step; one step = one pass over all cells "
             "(inner h / W loops to tol 0.05, H refresh, objective)")


def _config(name, path="auto", streams=1):
    """The config object of a bench line -- the SAME for the b200 arm and the reference arm of a config, sampling rule
    of the CPU arm included."""
    if name == "cfg2":
        return {"workload": WORKLOAD, "replicates_per_gpu": len(KS) * REPLICATES, "l2_flush_between_steps": True,
                "path": path, "k_batches_on_separate_streams": bool(streams),
                "e2e_step": f"DenseFrobeniusEngine(X_pinned).factorize_many(K=7..12, seeds, hals, tol=0, "
                            f"max_iter={E2E_ITERS}) per step: H2D of X from pinned host memory, random inits drawn by "
                            f"the device generator (init_device='cuda', nmf-torch's use_gpu=True stream), D2H of the "
                            f"objectives every 10 iterations and of all spectra",
                "reference_arm_sampling": f"the CPU arm (--impl reference) runs 1 replicate per K (K=7..12) x "
                                          f"{ITERS_PER_STEP} HALS iterations + objective per step: replicates are "
                                          f"independent and the CPU path runs them one after the other, so its "
                                          f"replicate-iterations/s does not depend on how many are run"}
    if name == "cfg3":
        return {"workload": WORKLOAD3, "jobs": len(CFG3["ks"]) * CFG3["replicates"], "inputs_larger_than_l2": True,
                "e2e_step": f"SparseKLEngine(X csr on the host).factorize_many(this rank's jobs, tol=0, "
                            f"max_iter={CFG3['e2e_iters']}) per step: CSR upload, device-drawn inits, D2H of objectives "
                            f"and spectra",
                "reference_arm_sampling": "the CPU arm runs 1 replicate of K=9 x 2 KL MU iterations + objective (dense, "
                                          "torch CPU) per step on the same matrix; replicates are independent and run "
                                          "one after the other on the CPU"}
    if name == "cfg5":
        return {"workload": WORKLOAD5, "inputs_larger_than_l2": True,
                "e2e_step": "DenseFrobeniusEngine(this rank's cells, pinned host memory).factorize_online(max_pass=1) "
                            "per step: H2D of X, device-drawn inits, D2H of spectra",
                "reference_arm_sampling": "the CPU arm runs 1 replicate x 1 pass over the first 100 000 cells (20 chunks) "
                                          "per step and divides by 10: the chunk loop is linear in the number of cells"}
    raise ValueError(name)


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def _traffic_from_profile(csv_name, kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum of the named kernel's launch in a committed `ncu --set full`
    raw page (profiles/<csv_name>), per launch; None when the capture is absent (then the line says so)."""
    import csv
    path = os.path.join(ROOT, "profiles", csv_name)
    if not os.path.exists(path):
        return None
    try:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        best = None
        for r in rows[2:]:
            if kernel_substr in r[ki]:
                v = float(r[ri].replace(",", "")) * mult.get(units[ri], 1.0) + float(r[wi].replace(",", "")) * mult.get(units[wi], 1.0)
                best = v if best is None else max(best, v)
        return best
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU while the timed region runs (NVML every 10 ms; falls
    back to the nvidia-smi query of the profiling recipe every 200 ms if NVML cannot be loaded)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.sm, self.max_sm, self.reasons, self._stop_evt = index, [], [], set(), threading.Event()
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        self.max_sm.append(float(n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)))
        r = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                          ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
            c = [v.strip() for v in out.split(",")]
            self.sm.append(float(c[0])); self.max_sm.append(float(c[1]))
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], c[3:7]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)

    def run(self):
        while not self._stop_evt.is_set():
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop_evt.wait(0.01 if self.nvml is not None else 0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                "sm_max_mhz": max(self.max_sm) if self.max_sm else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def _use_all_host_threads(torch):
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host thread it can."""
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def _cpu_baseline(x, seconds_target=12.0):
    """The oracle port of run_nmf (torch CPU, all host threads) on a bounded sample of the same workload: K=10, HALS,
    a few replicates x 10 iterations + objective (trace identity, as a competent CPU implementation evaluates it),
    repeated until ~seconds_target.  scikit-learn's coordinate-descent solver from the same init is timed beside it
    (``sklearn_cd``): an independent, compiled CPU implementation of the same HALS sweep, so that the ratio does not
    rest on the restatement alone."""
    import torch
    from oracle import nmf_oracle as no
    cores = _use_all_host_threads(torch)
    k = 10
    done, t0, reps = 0, time.perf_counter(), 0
    while True:
        h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), x.shape[0], x.shape[1], k, 1000 + reps)
        no.nmf_batch(x, h0, w0, algo=ALGO, tol=0.0, max_iter=ITERS_PER_STEP, objective="trace")
        done += ITERS_PER_STEP
        reps += 1
        el = time.perf_counter() - t0
        if el > seconds_target or reps >= 200:
            break
    out = {"value": done / el, "unit": UNIT, "cores": int(cores), "kind": "port",
           "sample": f"{reps} replicates x {ITERS_PER_STEP} HALS iterations + objective (trace identity), K={k}, same X, "
                     f"{el:.1f}s of torch-CPU work (oracle/nmf_oracle.py; nmf-torch itself is not installable here)"}
    try:
        from sklearn.decomposition import non_negative_factorization
        x64 = x.astype(np.float64)
        h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), x.shape[0], x.shape[1], k, 999)
        n_it, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < 4.0:
            non_negative_factorization(x64, W=h0.astype(np.float64).copy(), H=w0.astype(np.float64).copy(),
                                       n_components=k, init="custom", solver="cd", max_iter=ITERS_PER_STEP, tol=0)
            n_it += ITERS_PER_STEP
        out["sklearn_cd"] = {"value": n_it / (time.perf_counter() - t0), "unit": UNIT,
                             "note": "sklearn non_negative_factorization(solver='cd', init='custom') from the same "
                                     "kind of init, fp64, same X, K=10; its own threading"}
    except Exception as e:   # the comparison is optional
        out["sklearn_cd"] = {"error": str(e)[:100]}
    return out


def _extras(dev):
    """Side measurements on the other BASELINE.json configs (not the bench metric): wall-times of the
    device paths, with the host library path of the reference beside the consensus number."""
    import torch
    from cnmf_torch_b200 import consensus as cs
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    from cnmf_torch_b200.synth import nb_counts, replicate_spectra
    out = {}
    # cfg-4: consensus front half, 200 replicates x K=20 over 2000 genes
    s = replicate_spectra(20, 200, 2000, seed=14)
    cs.consensus_front_half(s, 20, 0.5, 60, device=dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        r = cs.consensus_front_half(s, 20, 0.5, 60, device=dev)
    torch.cuda.synchronize()
    out["consensus_cfg4_ms"] = (time.perf_counter() - t0) / 3 * 1e3
    try:
        from oracle import consensus_oracle as co
        t0 = time.perf_counter()
        ref = co.consensus_front_half(s, 20, 0.5, 0.30)
        out["consensus_cfg4_host_sklearn_ms"] = (time.perf_counter() - t0) * 1e3
        out["consensus_cfg4_labels_identical"] = bool(np.array_equal(ref["labels"], r["labels"]))
    except Exception as e:  # the oracle is optional here
        out["consensus_cfg4_host_sklearn_ms"] = None
    # cfg-3: sparse KL MU, K=9, R=100, one K-batch alone -- at 20 000 cells (round 1's extra) and at the config's own
    # 100 000 cells (`--config cfg3` runs the whole K=5..13 ledger)
    for cells, key in ((20000, "kl_sparse_20k_cells"), (CFG3["cells"], "kl_sparse_cfg3_100k_cells")):
        xk = nb_counts(cells, 2000, seed=14, libsize=170.0)
        eng = SparseKLEngine(xk, device=dev, init_device="cuda")
        b = eng.new_batch(9, seeds=list(range(100)))
        b.iterate(2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        b.iterate(10)
        e1.record()
        torch.cuda.synchronize()
        dt = e0.elapsed_time(e1) / 10 * 1e-3
        out[key] = {"density": float(eng.nnz) / (cells * 2000.0), "nnz": int(eng.nnz), "ms_per_batched_iter": dt * 1e3,
                    "replicate_iters_per_s": 100 / dt,
                    "algorithmic_hbm_gb_s": (2 * 8 * eng.nnz + 100 * 8 * 9 * (cells + 2000)) / dt / 1e9,
                    "fp32_fma_tflops_algorithmic": 2.0 * 2 * eng.nnz * 100 * 2 * 9 / dt / 1e12,
                    "kernels": ("slab" if b.p.impl == 1 else "gather") + " family (the engine's choice for this Kp); "
                               "includes 1/10 of an objective evaluation"}
        del eng, b, xk
        torch.cuda.empty_cache()
    # cfg-1/2 in the reference's default mode: online MU, chunk 5000, K=9, R=100, 3 passes
    xo = nb_counts(3000, 2000, seed=14)
    e2 = DenseFrobeniusEngine(xo, device=dev, init_device="cuda")   # host-generator draws would be 3/4 of the time
    e2.factorize_online(9, seeds=list(range(100)), tol=0.0, max_pass=1, chunk_max_iter=1000)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = e2.factorize_online(9, seeds=list(range(100)), tol=0.0, max_pass=3, chunk_max_iter=1000)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out["online_mu_3000x2000_K9_R100"] = {"s_per_3_passes": dt, "inner_iterations": res.inner_iterations,
                                          "replicate_passes_per_s": 300 / dt}
    del e2
    out["pipeline_cfg2"] = _pipeline_wall(dev)
    return out


def _pipeline_wall(dev):
    """BASELINE.json's second metric on cfg-2: wall time of factorize + combine + consensus through the cNMF mirror
    (files and all): K = 7..12 x 100 replicates, frobenius HALS batch mode with the reference's tol (1e-4) and
    stopping rule, then the consensus front half for every K."""
    import shutil
    import tempfile
    import torch
    from cnmf_torch_b200 import cnmf as host
    from cnmf_torch_b200.synth import nb_counts
    x = nb_counts(N_CELLS, N_GENES, seed=14)
    tmp = tempfile.mkdtemp(prefix="cnmf_b200_bench_")
    try:
        # warm-up on a small matrix: module loading, thread pools and pandas' first calls are not the pipeline's time
        warm = host.cNMF(output_dir=tmp, name="warm", device=str(dev))
        warm.prepare_from_matrix(x[:300, :200] + 0.01, [3], n_iter=4, seed=1, beta_loss="frobenius", use_gpu=True)
        warm.factorize(nmf_kwargs_override=dict(mode="batch", algo=ALGO, batch_max_iter=20))
        warm.combine()
        warm.consensus(3, density_threshold=2.0)
        obj = host.cNMF(output_dir=tmp, name="cfg2", device=str(dev))
        obj.prepare_from_matrix(x, KS, n_iter=REPLICATES, seed=14, beta_loss="frobenius", use_gpu=True)  # `cnmf prepare --use_gpu`
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        obj.factorize(nmf_kwargs_override=dict(mode="batch", algo=ALGO, batch_max_iter=500))
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        obj.combine()
        t2 = time.perf_counter()
        for k in KS:
            obj.consensus(k, density_threshold=0.5, local_neighborhood_size=0.30)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        return {"factorize_s": t1 - t0, "combine_s": t2 - t1, "consensus_s": t3 - t2, "total_s": t3 - t0,
                "what": f"cNMF mirror on {N_CELLS}x{N_GENES}: factorize K={KS[0]}..{KS[-1]} x {REPLICATES} replicates "
                        f"({ALGO}, batch, tol 1e-4, up to 500 iterations, {len(KS) * REPLICATES} spectra files written), "
                        f"combine, consensus(k, 0.5, 0.30) for every K"}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def _cfg3_matrix():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(CFG3["cells"], CFG3["genes"], seed=14, libsize=CFG3["libsize"])


def _cfg5_block_device(n_rows, seed, dev):
    """One rank's block of the cfg-5 matrix, drawn on the device (a 1 M x 2000 host draw takes minutes): NB counts
    from 10 gene programs, every gene scaled to unit variance like the reference's normalisation.  fp32 [n_rows x G]."""
    import torch
    g = torch.Generator(device=dev)
    g.manual_seed(14)                                        # the programs are the same on every rank
    G, kt = CFG5["genes"], 10
    prog = torch._standard_gamma(torch.full((kt, G), 0.3, device=dev), generator=g) + 1e-6
    prog /= prog.sum(1, keepdim=True)
    g.manual_seed(1000 + seed)
    out = torch.empty((n_rows, G), device=dev, dtype=torch.float32)
    for lo in range(0, n_rows, 50000):
        hi = min(n_rows, lo + 50000)
        usage = torch._standard_gamma(torch.full((hi - lo, kt), 0.2, device=dev), generator=g) + 1e-8
        usage /= usage.sum(1, keepdim=True)
        lib = torch.exp(torch.randn((hi - lo, 1), device=dev, generator=g) * 0.35 + math.log(600.0))
        mu = lib * (usage @ prog)
        lam = torch._standard_gamma(torch.full_like(mu, 2.0), generator=g) * (mu / 2.0)
        out[lo:hi] = torch.poisson(lam, generator=g)
    out[:, :] += (out.sum(1, keepdim=True) == 0) * (1.0 / G)      # no all-zero cell
    std = out.std(dim=0, unbiased=True)
    std[std == 0] = 1.0
    out /= std
    return out


def run_reference(args):
    """--impl reference: the reference's CPU path for this metric.  nmf-torch is absent from the image
    (SURVEY.md 0.2), so this times the oracle port with all host threads; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from cnmf_torch_b200.synth import nb_counts
    from oracle import nmf_oracle as no
    cores = _use_all_host_threads(torch)
    cfg = args.config
    if cfg == "cfg2":
        x = nb_counts(N_CELLS, N_GENES, seed=14)

        def step(seed0):   # every K, 1 replicate per K x 10 iterations + objective
            n = 0
            for k in KS:
                h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), N_CELLS, N_GENES, k, seed0 + k)
                no.nmf_batch(x, h0, w0, algo=ALGO, tol=0.0, max_iter=ITERS_PER_STEP, objective="trace")
                n += ITERS_PER_STEP
            return n
        scaling, dtype = "weak", "f32"
    elif cfg == "cfg3":
        x = _cfg3_matrix()

        def step(seed0):   # 1 replicate of K=9 x 2 KL iterations + objective on the dense matrix
            h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), x.shape[0], x.shape[1], 9, seed0)
            no.nmf_batch(x, h0, w0, beta_loss="kullback-leibler", tol=0.0, max_iter=2, check_every=2)
            return 2
        scaling, dtype = "strong", "f32"
    else:
        from cnmf_torch_b200.synth import nb_counts as nbc
        x = nbc(100000, CFG5["genes"], seed=14, libsize=600.0)

        def step(seed0):   # 1 replicate x 1 pass over 100 000 cells = 1/10 of a pass over the config's 1 M cells
            h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), x.shape[0], x.shape[1], CFG5["k"], seed0)
            no.nmf_online(x, h0, w0, tol=0.0, max_pass=1, chunk_size=CFG5["chunk"], chunk_max_iter=200)
            return 0.1
        scaling, dtype = "strong", "f32"
    for w in range(args.warmup):
        step(w)
    t0 = time.perf_counter()
    total = 0
    for s in range(args.steps):
        total += step(100 + s)
    el = time.perf_counter() - t0
    val = total / el
    config = _config(cfg, args.path, args.streams)
    sample = config["reference_arm_sampling"]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * el / max(1, args.steps),
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": dtype, "data": "synthetic",
        "config": config,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": int(cores), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def _dist_setup():
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return rank, world, local, dev


def _timed_steps(args, one_step, world, local, flush=None):
    """W warm-up steps, then K steps timed with CUDA events on the current stream, barrier + synchronize on both sides;
    returns (sum of the step times in ms on this rank, clocks sampled meanwhile)."""
    import torch
    import torch.distributed as dist
    for _ in range(args.warmup):
        one_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for s_ in range(args.steps):
        if flush is not None:
            flush.zero_()
        ev[s_][0].record()
        one_step()
        ev[s_][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    return float(sum(a.elapsed_time(b) for a, b in ev)), clocks


def _max_over_ranks(values, dev, world):
    import torch
    import torch.distributed as dist
    if world == 1:
        return [float(v) for v in values]
    t = torch.tensor(values, dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t]


def run_cfg3(args):
    """BASELINE.json configs[2]: one ledger of 900 KL jobs, strided over the ranks (strong scaling)."""
    import scipy.sparse as sp
    import torch
    import torch.distributed as dist
    from cnmf_torch_b200 import cnmf as host
    from cnmf_torch_b200.kl_engine import KlBatch, SparseKLEngine, pad2
    from cnmf_torch_b200.synth import ledger_seeds
    rank, world, local, dev = _dist_setup()
    ks, R = CFG3["ks"], CFG3["replicates"]
    x = _cfg3_matrix()
    xs = sp.csr_matrix(x)
    seeds = ledger_seeds(len(ks) * R, seed=14)
    ledger = [(k, int(seeds[i * R + r])) for i, k in enumerate(ks) for r in range(R)]      # cnmf.py:744 order
    mine = list(host.worker_filter(ledger, rank, world))
    by_k = {}
    for k, sd in mine:
        by_k.setdefault(k, []).append(sd)
    eng = SparseKLEngine(xs, device=dev, init_device="cuda")
    stacks = {}
    for k in by_k:
        stacks.setdefault(pad2(k), []).append(k)
    batches = [KlBatch(eng, None, None, None, 0.0, 0.0, 0.0, 0.0, 1e-16,
                       groups=[(k,) + tuple(eng._init_for(k, by_k[k])) for k in group]) for group in stacks.values()]
    n_mine = len(mine)

    def one_step():
        for b in batches:
            b.iterate(ITERS_PER_STEP)

    # dominant kernel: the H half-step (slab kernel) of the tallest stack, timed alone through one iteration's events
    big = max(batches, key=lambda b: (b.R, b.Kp))
    big_kp = big.Kp
    big_slab = big.p.impl == 1
    big.iterate(1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    big.iterate(4)
    e1.record()
    torch.cuda.synchronize()
    it_ms = e0.elapsed_time(e1) / 4
    launches0 = sum(b.launches for b in batches)
    ms_total, clocks = _timed_steps(args, one_step, world, local)
    launches = sum(b.launches for b in batches) - launches0
    losses = [b.read_loss() for b in batches]
    assert all(np.isfinite(l).all() for l in losses)
    del batches, big

    def e2e_step():
        e = SparseKLEngine(xs, device=dev, init_device="cuda")
        res = e.factorize_many(list(by_k), by_k, tol=0.0, max_iter=CFG3["e2e_iters"])
        h2d = xs.data.nbytes * 2 + xs.indices.nbytes * 2 + xs.indptr.nbytes + (xs.shape[1] + 1) * 4
        d2h = sum(r.W.nbytes + (CFG3["e2e_iters"] // 10 + 1) * len(by_k[k]) * 8 for k, r in res.items())
        return h2d, d2h
    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    h2d, d2h = e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    ms_total, e2e_s, it_ms_max = _max_over_ranks([ms_total, e2e_s, it_ms], dev, world)
    pipeline = _cfg3_pipeline(xs, rank, world, dev) if args.pipeline else None
    jobs = len(ledger)
    value = jobs * ITERS_PER_STEP * args.steps / (ms_total * 1e-3)
    if rank == 0:
        peaks, src = _peaks()
        nnz = int(xs.nnz)
        rk = sum(len(v) * k for k, v in by_k.items() if pad2(k) == big_kp)       # K x replicates of the big stack
        r_big = sum(len(v) for k, v in by_k.items() if pad2(k) == big_kp)
        # algorithmic bytes of one batched iteration of that stack: X's CSR twice + factors read and written
        abytes = 2 * (8 * nnz + 4 * (x.shape[0] + 1)) + 8 * rk * (x.shape[0] + x.shape[1])
        aflops = 2.0 * 2 * nnz * 2 * rk                 # 2 half-steps x (s and numerator) x K FMAs per entry
        fma_peak = 148 * 128 * 2 * (clocks["sm_mhz"] or 1900.0) * 1e6 / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f16 operands, f32 accumulation", "data": "synthetic", "config": _config("cfg3"),
            "e2e": {"value": jobs * CFG3["e2e_iters"] / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": abytes / (it_ms_max * 1e-3) / 1e9, "peak": peaks["hbm_gbs"],
                         "unit": "GB/s", "frac": abytes / (it_ms_max * 1e-3) / 1e9 / peaks["hbm_gbs"],
                         "traffic": _traffic_from_profile(KL_SLAB_PROFILE_CSV, "kl_slab_kernel") if big_slab
                         else _traffic_from_profile(KL_PROFILE_CSV, "kl_half_step_kernel"),
                         "kernel": f"one batched KL iteration of the Kp={big_kp} stack on rank 0 ({r_big} replicates): "
                                   + ("kl_slab_kernel x2 (slab family)" if big_slab else "kl_half_step_kernel x2 (gather family)")
                                   + " + finish passes + column sums; stacks of Kp <= 10 and 14 run the slab family",
                         "kernel_ms": it_ms_max,
                         "peak_source": src,
                         "note": "the algorithmic bytes are tiny against the arithmetic: this path is bound by the fp32 FMA "
                                 "pipe (instruction issue), see roofline_fma"},
            "roofline_fma": {"bound": "fp32 FMA pipe", "achieved": aflops / (it_ms_max * 1e-3) / 1e12, "peak": fma_peak,
                             "unit": "TFLOP/s", "frac": aflops / (it_ms_max * 1e-3) / 1e12 / fma_peak,
                             "peak_source": "148 SMs x 128 lanes x 2 x measured SM clock (FFMA / FHFMA issue rate "
                                            "measured at 122-124 of 128 lanes/clk/SM: profiles/r02_simt_rate.txt)"},
            "jobs_per_rank": n_mine,
        }
        if pipeline is not None:
            line["pipeline_cfg3"] = pipeline
        if not args.no_cpu_baseline and world == 1:
            import torch as _t
            from oracle import nmf_oracle as no
            cores = _use_all_host_threads(_t)
            h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), x.shape[0], x.shape[1], 9, 1)
            t0 = time.perf_counter()
            no.nmf_batch(x, h0, w0, beta_loss="kullback-leibler", tol=0.0, max_iter=4, check_every=4)
            el = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": 4 / el, "unit": UNIT, "cores": int(cores), "kind": "port",
                                    "sample": f"1 replicate x 4 KL MU iterations + objective, K=9, dense on the host, {el:.1f}s"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _cfg3_pipeline(xs, rank, world, dev):
    """BASELINE.json's second metric at cfg-3: wall time of factorize (the ledger strided over the ranks, replicate files
    written) + combine (rank 0) + consensus of every K (the K values dealt over the ranks), tol and iteration limits as
    `cnmf prepare` writes them.  Max over ranks per stage."""
    import shutil
    import tempfile
    import torch
    import torch.distributed as dist
    from cnmf_torch_b200 import cnmf as host
    box = [tempfile.mkdtemp(prefix="cnmf_b200_cfg3_") if rank == 0 else None]
    if world > 1:
        dist.broadcast_object_list(box, 0)
    obj = host.cNMF(output_dir=box[0], name="cfg3", device=str(dev))
    try:
        if rank == 0:
            obj.prepare_from_matrix(xs, CFG3["ks"], n_iter=CFG3["replicates"], seed=14, beta_loss="kullback-leibler",
                                    use_gpu=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        obj.factorize(worker_i=rank, total_workers=world)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if world > 1:
            dist.barrier()
        t1b = time.perf_counter()
        mine_k = CFG3["ks"][rank::world]
        if mine_k:
            obj.combine(components=mine_k)            # `cnmf combine -k ...`: one K's merged file per call, dealt over the ranks
        if world > 1:
            dist.barrier()
        t2 = time.perf_counter()
        for k in CFG3["ks"][rank::world]:
            obj.consensus(k, density_threshold=0.5, local_neighborhood_size=0.30)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        if world > 1:
            dist.barrier()
        t4 = time.perf_counter()
        st = _max_over_ranks([t1 - t0, t2 - t1b, t3 - t2, t4 - t0], dev, world)
        return {"factorize_s": st[0], "combine_s": st[1], "consensus_s": st[2], "total_s": st[3],
                "what": "cNMF mirror: factorize(worker_i=rank, total_workers=N) of the 900-job KL ledger (tol 1e-4, up to 500 "
                        "iterations, replicate files written), combine and consensus(k, 0.5, 0.30) incl. the usage "
                        "refit over all cells, K values dealt over the ranks; max over ranks per stage"}
    finally:
        if world > 1:
            dist.barrier()
        if rank == 0:
            shutil.rmtree(box[0], ignore_errors=True)


def run_cfg5(args):
    """BASELINE.json configs[4]: online mode over row-sharded cells with the NCCL all-reduce of [A | B]."""
    import torch
    import torch.distributed as dist
    from cnmf_torch_b200 import dist as d
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    rank, world, local, dev = _dist_setup()
    n_total, K, R, chunk = CFG5["cells"], CFG5["k"], CFG5["replicates"], CFG5["chunk"]
    lo, hi = d.shard_rows(n_total, rank, world)
    xd = _cfg5_block_device(hi - lo, rank, dev)
    eng = DenseFrobeniusEngine(xd, device=dev, row_group=dist.group.WORLD if world > 1 else None, row_offset=lo,
                               init_device="cuda")
    seeds = list(range(R))
    prof = {}

    def one_step():
        res = eng.factorize_online(K, seeds=seeds, tol=0.0, max_pass=1, chunk_size=chunk, chunk_max_iter=200,
                                   profile=prof)
        prof["last"] = res
    ms_total, clocks = _timed_steps(args, one_step, world, local)
    res = prof["last"]
    ar_ms = prof.get("allreduce_ms", 0.0)
    # end to end: this rank's cells in pinned host memory -> engine (H2D, transposes, tf32 rounding) -> one pass -> spectra
    x_host = xd.cpu().pin_memory()
    del eng
    torch.cuda.empty_cache()

    def e2e_step():
        e = DenseFrobeniusEngine(x_host, device=dev, row_group=dist.group.WORLD if world > 1 else None, row_offset=lo,
                                 init_device="cuda")
        r = e.factorize_online(K, seeds=seeds, tol=0.0, max_pass=1, chunk_size=chunk, chunk_max_iter=200)
        return x_host.numel() * 4, r.W.nbytes + 2 * R * 8
    del xd
    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    h2d, d2h = e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    ms_total, ar_ms, e2e_s = _max_over_ranks([ms_total, ar_ms, e2e_s], dev, world)
    value = R * args.steps / (ms_total * 1e-3)
    if rank == 0:
        peaks, src = _peaks()
        n_steps = args.steps + args.warmup
        flops = 2.0 * 2 * n_total * CFG5["genes"] * K * R * 2      # x W^T twice (chunk pass + refresh) and h^T x
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "tf32", "data": "synthetic (drawn on the device)", "config": _config("cfg5"),
            "e2e": {"value": R / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "note": "per rank; one pass per step, engine construction (upload, transposes) included"},
            "gpu_launches": int(res.launches), "clocks": clocks,
            "online": {"inner_iterations_per_pass": res.inner_iterations, "chunks_per_rank": (hi - lo + chunk - 1) // chunk,
                       "allreduce_ms_per_pass": ar_ms / max(1, n_steps),
                       "allreduce_share_of_pass": (ar_ms / max(1, n_steps)) / (ms_total / args.steps) if world > 1 else 0.0,
                       "allreduce_bytes_per_chunk_step": int(R * 16 * (16 + CFG5["genes"]) * 4)},
            "roofline": {"bound": "tensor", "achieved": flops / (ms_total / args.steps * 1e-3) / 1e12 * 0.75,
                         "peak": peaks["bf16_tflops"] / 2.0 * world, "unit": "TFLOP/s",
                         "frac": flops * 0.75 / (ms_total / args.steps * 1e-3) / 1e12 / (peaks["bf16_tflops"] / 2.0 * world),
                         "traffic": None, "kernel": "whole pass: 3 GEMM sweeps over X (x W^T, h^T x, x W^T of the refresh) "
                                                    "against the pass time, all ranks", "peak_source": f"{src} bf16 burst / 2 (tf32)"},
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg2", choices=["cfg2", "cfg3", "cfg5"])
    ap.add_argument("--pipeline", action="store_true",
                    help="[cfg3] also time factorize + combine + consensus through the cNMF mirror, jobs strided over the ranks")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--path", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--no-extras", action="store_true", help="skip the consensus / KL / online side measurements")
    ap.add_argument("--streams", type=int, default=1, help="1: one stream per K-batch (default); 0: single stream")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    if args.config == "cfg3":
        return run_cfg3(args)
    if args.config == "cfg5":
        return run_cfg5(args)

    import torch
    import torch.distributed as dist
    from cnmf_torch_b200 import _lib
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.synth import ledger_seeds, nb_counts

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.lib()
    impl = {"auto": _lib.IMPL_AUTO, "simt": _lib.IMPL_SIMT, "tcgen05": _lib.IMPL_TCGEN05}[args.path]

    x = nb_counts(N_CELLS, N_GENES, seed=14)
    x_pinned = torch.from_numpy(x).pin_memory()
    seeds_all = ledger_seeds(len(KS) * REPLICATES * world, seed=14).reshape(world, len(KS), REPLICATES)
    seeds = seeds_all[rank]

    # ---------------- device-resident arm: `value` ------------------------------------------
    eng = DenseFrobeniusEngine(x_pinned, device=dev, impl=impl)
    # replicates of K values that share a padded Kp are stacked into one batch (K=7,8 -> Kp 8; K=9..12 -> Kp 12),
    # exactly as the public factorize_many does
    from cnmf_torch_b200.engine import FrobBatch, pad4
    stacks = {}
    for i, k in enumerate(KS):
        stacks.setdefault(pad4(k), []).append((k,) + tuple(eng.make_init(k, seeds[i])))
    batches = [FrobBatch(eng, None, None, None, ALGO, 0.0, 0.0, 0.0, 0.0, 1e-16, groups=g) for g in stacks.values()]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    from cnmf_torch_b200.engine import iterate_concurrently
    streams = [torch.cuda.Stream(device=dev) for _ in batches] if args.streams else None

    def one_step():
        iterate_concurrently(batches, ITERS_PER_STEP, streams)

    # ---------------- dominant kernel alone: roofline (before the long timed region: burst clocks,
    # compared with the burst peak; after a sustained tensor load the 1 kW cap lowers the SM clock) ----
    # the H half-step of the Kp=12 stack (K=9..12, 400 replicates), exactly as the timed steps launch it
    # (live HALS update, tf32 operand copy written, no partial sums), on a scratch copy of the state
    big = max(batches, key=lambda b: b.R * b.Kp)
    p = big.p
    st = torch.cuda.current_stream().cuda_stream
    n_rep = 30
    Wop = big.Wall.clone()
    _lib.check(lib.cnmf_b200_round_tf32(Wop.data_ptr(), Wop.numel(), st))
    Hs, Hop = big.Ht.clone(), big.Ht.clone()

    def dominant():
        _lib.check(lib.cnmf_b200_dense_half_step_k(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, Hs.data_ptr(),
                                                   Hop.data_ptr(), p.ldh, p.gramW, None, p.rep_k, p.R, p.Kp, p.algo,
                                                   0.0, 0.0, p.eps, None, st))
    roof = None
    if impl != _lib.IMPL_SIMT:
        for _ in range(5):
            dominant()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_rep):
            dominant()
        e1.record()
        torch.cuda.synchronize()
        k_ms = e0.elapsed_time(e1) / n_rep
        cols = sum(k * r for k, r in zip(big.Ks, big.Rs))  # algorithmic factor columns (unpadded K)
        flops = 2.0 * eng.N * eng.G * cols               # 2*N*G*K per replicate and half-step
        peaks, src = _peaks()
        peak = peaks["bf16_tflops"] / 2.0                # tf32 tensor rate = half the bf16 rate
        ach = flops / (k_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of this launch, read from the committed `ncu --set full`
                # capture of the same launch (null when no capture of this round's kernel is committed); algorithmic
                # bytes (X + W copy + H read + H and H-copy write) = 216 MB
                "traffic": _traffic_from_profile(TC_PROFILE_CSV, "tc_half_step_kernel"),
                "traffic_source": "profiles/" + TC_PROFILE_CSV,
                "kernel": f"tc_half_step_kernel<{big.Kp},HALS> (H half-step, K={big.Ks}, R={big.R}, live update)",
                "kernel_ms": k_ms, "padded_frac": ach / peak * big.Kp * big.R / cols,
                "peak_source": f"{src} bf16 burst / 2 (tf32)"}
    del Wop, Hs, Hop

    for _ in range(args.warmup):
        one_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = sum(b.launches for b in batches)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for s in range(args.steps):
        flush.zero_()  # evict L2 between timed steps (inputs are smaller than L2)
        ev[s][0].record()
        one_step()
        ev[s][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms_total = float(sum(a.elapsed_time(b) for a, b in ev))
    launches = sum(b.launches for b in batches) - launches0
    losses = [b.read_loss() for b in batches]
    assert all(np.isfinite(l).all() for l in losses)

    # ---------------- end-to-end arm: public API with host buffers -----------------------------
    def e2e_step(init_device):
        e = DenseFrobeniusEngine(x_pinned, device=dev, impl=impl, init_device=init_device)   # H2D of X (pinned)
        h2d, d2h = x_pinned.numel() * 4, 0
        res = e.factorize_many(KS, {k: seeds[i] for i, k in enumerate(KS)}, algo=ALGO, tol=0.0,
                               max_iter=E2E_ITERS)                        # init draws, all K, D2H of spectra
        for k in KS:
            if init_device == "cpu":                                      # host-drawn inits cross PCIe too
                h2d += REPLICATES * k * (N_CELLS + N_GENES) * 4
            d2h += res[k].W.nbytes + (E2E_ITERS // 10 + 1) * REPLICATES * 8
        return h2d, d2h

    def time_e2e(init_device, n=3):
        e2e_step(init_device)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            h2d_, d2h_ = e2e_step(init_device)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n, h2d_, d2h_
    # headline: the init drawn by the device generator (nmf-torch's use_gpu=True stream, cNMF.factorize's default
    # here); the host-generator stream (use_gpu=False, the oracle's) is timed beside it
    e2e_s, h2d, d2h = time_e2e("cuda")
    e2e_host_s, _, _ = time_e2e("cpu", n=2)

    # ---------------- other configs of BASELINE.json, reported as extras (rank 0, N=1 only) -----
    extras = None
    if world == 1 and not args.no_extras:
        extras = _extras(dev)

    # ---------------- reduce over ranks ------------------------------------------------------
    if world > 1:
        t = torch.tensor([ms_total, e2e_s, e2e_host_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, e2e_s, e2e_host_s = float(t[0]), float(t[1]), float(t[2])
        ln = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(ln)
        launches = int(ln[0])
    units_per_step = len(KS) * REPLICATES * ITERS_PER_STEP * world
    value = units_per_step * args.steps / (ms_total * 1e-3)
    e2e_val = len(KS) * REPLICATES * E2E_ITERS * world / e2e_s

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "tf32" if impl != _lib.IMPL_SIMT else "f32",
            "data": "synthetic",
            "config": _config("cfg2", args.path, args.streams),
            "e2e_host_generator_init": {
                "value": len(KS) * REPLICATES * E2E_ITERS * world / e2e_host_s, "unit": UNIT,
                "note": "the e2e step with init_device='cpu': inits drawn by torch's host generator (the reference's "
                        "use_gpu=False stream, the oracle's) and uploaded"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
        }
        if extras is not None:
            line["extras"] = extras
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = _cpu_baseline(x)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
