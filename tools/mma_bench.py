"""clocks per tcgen05.mma by operand layout (b200_tc_mma_bench).  python tools/mma_bench.py"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import masters_thesis_b200 as B
from masters_thesis_b200._lib import check, load, ptr
lib = load(); out = torch.zeros(1, device="cuda:0")
print("variant 0 = A,B K-major; 1 = A K-major, B MN-major; 2 = A,B MN-major; clocks per MMA (M=128, K=16)")
for variant in (0, 1, 2):
    for n in ((64, 128, 256) if variant == 0 else (64, 128)):
        row = []
        for layout in (0, 1):
            check(lib.b200_tc_mma_bench(layout, variant, n, 4096, ptr(out), None), "bench"); torch.cuda.synchronize()
            row.append(float(out))
        print(f"variant {variant} N={n:3d}: interleaved {row[0]:6.1f}  swizzle128 {row[1]:6.1f}  (math floor {n / 2:.0f})")
