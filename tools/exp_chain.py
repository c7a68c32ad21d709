import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
for (n, hi, cin, cout, k, pad, bn) in [(256, 56, 64, 64, 3, 1, 64), (128, 149, 32, 32, 3, 0, 32), (256, 28, 128, 128, 3, 1, 128), (256, 56, 256, 64, 1, 0, 64)]:
    ms, tf, gbs = time_conv(2, n, hi, hi, cin, cout, k, k, 1, pad, pad, False, bn, 0)
    print(f"{hi} {cin}->{cout} k{k} bn{bn}: {ms*1e3:.1f} us {tf:.0f} TF", flush=True)
