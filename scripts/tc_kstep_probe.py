"""Dev probe: clocks per k-block of the tcgen05 main loop (pure GEMM, store epilogue) on a shape where every SM
gets the same number of tiles and the K loop is long, for the tile plan forced through CNMF_B200_TC_T / _CG /
_STAGES.  Prints ns per k-block per SM."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnmf_torch_b200 import _lib
lib = _lib.lib()
KP = int(os.environ.get("KP", "12")); T = int(os.environ.get("CNMF_B200_TC_T", "20")); CG = int(os.environ.get("CNMF_B200_TC_CG", "1"))
TILES_PER_SM = 2
M = 128 * 37 * (2 if CG == 2 else 1)          # 37 m-tiles (or m-groups)
NT = 4 * TILES_PER_SM                           # n tiles -> 37*8 = 296 tiles = 2 per SM (or per pair: 148 pair tiles... )
if CG == 2: NT = 4                              # 37 groups * 4 = 148 pair-tiles over 74 pairs = 2 per pair
R = T * NT
KD = int(os.environ.get("KD", "8000"))
ld = lambda n: (n + 31) // 32 * 32
A = torch.randn(M, ld(KD), device="cuda"); F = torch.randn(R * KP, ld(KD), device="cuda"); P = torch.zeros(R * KP, ld(M), device="cuda")
st = torch.cuda.current_stream().cuda_stream
def run(): _lib.check(lib.cnmf_b200_dense_gemm_tf32(A.data_ptr(), ld(KD), M, KD, F.data_ptr(), ld(KD), P.data_ptr(), ld(M), None, R, KP, 0, st))
for _ in range(3): run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): run()
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 10 * 1e3
nkb = (KD + 31) // 32
rows = 128 + T * KP // CG
print(f"CG={CG} T={T} BN={T*KP} stages={os.environ.get('CNMF_B200_TC_STAGES','auto')} M={M} R={R}: {us:.1f} us, "
      f"{us*1e3/(TILES_PER_SM*nkb):.0f} ns/k-block/SM  ({rows} box rows, {rows*128/1024:.1f} KB per CTA per k-block; MMA {T*KP*2*4/4:.0f} clk)")
