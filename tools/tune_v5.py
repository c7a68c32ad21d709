import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
CASES = [(256, 56, 64, 64, 3, 1, 1), (256, 56, 64, 64, 1, 1, 0), (256, 56, 256, 64, 1, 1, 0), (256, 28, 128, 128, 3, 1, 1), (256, 28, 512, 128, 1, 1, 0), (256, 56, 256, 128, 1, 2, 0),
         (128, 149, 32, 32, 3, 1, 0), (128, 147, 32, 64, 3, 1, 1), (128, 35, 48, 64, 5, 1, 2), (128, 35, 64, 96, 3, 1, 1), (128, 35, 96, 96, 3, 1, 1), (128, 17, 128, 128, 7, 1, 3), (128, 73, 64, 80, 1, 1, 0)]
for (n, hi, cin, cout, k, stride, pad) in CASES:
    for impl in (2, 5, 6):
        try:
            ms, tf, gbs = time_conv(impl, n, hi, hi, cin, cout, k, k, stride, pad, pad, False, 0, 0)
            print(f"{hi}x{hi} {cin}->{cout} k{k} s{stride} impl{impl}: {ms*1e3:.1f} us {tf:.0f} TF {gbs:.0f} GB/s", flush=True)
        except Exception as e:
            print("ERR", hi, cin, cout, k, impl, str(e)[:80], flush=True)
