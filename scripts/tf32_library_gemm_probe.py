"""Calibration probe (test-only, not a product path): what a LIBRARY tf32 GEMM (torch.matmul -> cuBLAS, allow_tf32)
reaches on this box for the shapes of the dense half-step, timed with the bench's protocol (warm-up, CUDA events,
L2 flush between launches optional).  Answers VERDICT r01 item 6a: is the shared-memory-port model of DESIGN.md a
property of the machine (a library kernel with its own tile shape lands in the same place) or of our kernel?
    python scripts/tf32_library_gemm_probe.py
"""
import json
import torch

torch.backends.cuda.matmul.allow_tf32 = True
torch.backends.cudnn.allow_tf32 = True
dev = "cuda"


def timed(f, n=30, warm=5):
    for _ in range(warm):
        f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out = {}
for name, (M, K, N) in {
    "H half-step cfg-2 stack [2700x2000]x[2000x4800]": (2700, 2000, 4800),
    "W half-step cfg-2 stack [2000x2700]x[2700x4800]": (2000, 2700, 4800),
    "long-K probe shape [4736x8000]x[8000x1920]": (4736, 8000, 1920),
    "square 8192": (8192, 8192, 8192),
}.items():
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev)            # NT form, as the half-step reads its operands
    C = torch.empty(M, N, device=dev)
    ms = timed(lambda: torch.matmul(A, B.t(), out=C))
    tf = 2.0 * M * K * N / (ms * 1e-3) / 1e12
    Ab, Bb = A.bfloat16(), B.bfloat16()
    Cb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    msb = timed(lambda: torch.matmul(Ab, Bb.t(), out=Cb))
    out[name] = {"tf32_ms": ms, "tf32_tflops": tf, "bf16_ms": msb, "bf16_tflops": 2.0 * M * K * N / (msb * 1e-3) / 1e12}
    print(f"{name}: tf32 {ms * 1e3:.1f} us = {tf:.0f} TFLOP/s; bf16 {msb * 1e3:.1f} us = {out[name]['bf16_tflops']:.0f} TFLOP/s")
print(json.dumps(out))
