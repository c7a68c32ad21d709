import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
for (n, hi, cin, cout, k, pad) in [(256, 56, 64, 64, 3, 1), (256, 28, 128, 128, 3, 1), (256, 14, 256, 256, 3, 1)]:
    for st in (0, 4, 6):
        ms, tf, gbs = time_conv(4, n, hi, hi, cin, cout, k, k, 1, pad, pad, False, 0, st)
        print(f"{hi} {cin}->{cout} k{k} bstages={st}: {ms*1e3:.1f} us {tf:.0f} TF", flush=True)
