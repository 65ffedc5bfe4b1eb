"""What one tcgen05 kind::tf32 MMA (M = 64, K = 8) costs as a function of N: every SM issues 20 000 of them back to back.
python profiles/microbench/tc_mma_rate.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, hfrep
from hfrep._lib import lib, check
torch.zeros(1, device="cuda")
it = 20000
for N in (8, 16, 32, 64, 144, 256):
    best = 1e9
    for rep in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); check(lib.hfrep_tc_mma_rate(N, it, None)); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    ns = best * 1e6 / it
    print(f"N={N:3d}: {ns:6.1f} ns per MMA ({ns * 1.965:6.0f} cycles at 1965 MHz), {64 * N * 8 * 2 / ns / 1e3 * 148:7.1f} TFLOP/s over 148 SMs")
