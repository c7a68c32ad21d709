import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
CASES = [(256, 56, 64, 64, 3, 3, 1, 1), (256, 28, 128, 128, 3, 3, 1, 1), (256, 14, 256, 256, 3, 3, 1, 1), (128, 35, 64, 96, 3, 3, 1, 1), (128, 35, 64, 64, 5, 5, 2, 2),
         (128, 17, 128, 128, 1, 7, 0, 3), (128, 17, 192, 192, 7, 1, 3, 0), (128, 147, 64, 64, 3, 3, 1, 1)]
for (n, hi, cin, cout, kh, kw, ph, pw) in CASES:
    for impl, bn in ((2, 0), (4, 0), (4, 64), (4, 128), (4, 256)):
        if bn and (cout % bn or bn > cout):
            continue
        try:
            ms, tf, gbs = time_conv(impl, n, hi, hi, cin, cout, kh, kw, 1, ph, pw, False, bn, 0)
            print(f"{hi}x{hi} {cin}->{cout} {kh}x{kw} impl{impl} bn={bn}: {ms*1e3:.1f} us {tf:.0f} TF {gbs:.0f} GB/s", flush=True)
        except Exception as e:
            print("ERR", hi, cin, cout, kh, kw, impl, bn, e, flush=True)
