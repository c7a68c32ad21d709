import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
# (n, hi, cin, cout, k, pad, bn): bytes per k-block = 16 KB (A) + bn*128 (B)
for (n, hi, cin, cout, k, pad, bn) in [(256, 56, 64, 64, 3, 1, 64), (256, 56, 64, 32, 3, 1, 32), (128, 149, 32, 32, 3, 0, 32), (256, 56, 256, 64, 1, 0, 64),
                                       (256, 56, 256, 256, 1, 0, 256), (256, 28, 512, 128, 1, 0, 128), (256, 14, 256, 256, 3, 1, 256), (256, 14, 1024, 256, 1, 0, 256)]:
    ms, tf, gbs = time_conv(2, n, hi, hi, cin, cout, k, k, 1, pad, pad, False, bn, 0)
    ho = hi + 2 * pad - k + 1
    tiles = (n * ho * ho + 127) // 128 * (cout // bn)
    kb = k * k * ((cin + 63) // 64)
    byts = tiles * kb * (16384 + bn * 128)
    print(f"{hi} {cin}->{cout} k{k} bn{bn}: {ms*1e3:.1f} us  smem fill {byts/ms/1e9:.2f} TB/s  = {byts/ms/1e6/148/1.965:.1f} B/cyc/SM  ({tf:.0f} TF if real)", flush=True)
