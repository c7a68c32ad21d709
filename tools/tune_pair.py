import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv, RESNET
extra = [(35, 64, 96, 3, 1, 1, False), (17, 128, 192, 7, 1, 3, False), (8, 448, 384, 3, 1, 1, False), (17, 768, 192, 1, 1, 0, False), (35, 288, 384, 3, 2, 0, False)]
n = 256
for (hi, cin, cout, k, stride, pad, res) in RESNET + extra:
    nn = n if (hi, cin, cout, k, stride, pad, res) in RESNET else 128
    for impl, bn in ((2, 0), (5, 0), (5, 128), (5, 256)):
        if bn and (cout % bn or bn > cout):
            continue
        try:
            ms, tf, gbs = time_conv(impl, nn, hi, hi, cin, cout, k, k, stride, pad, pad, res, bn, 0)
            print(f"{hi}x{hi} {cin}->{cout} k{k} s{stride}{' +res' if res else ''} | impl{impl} bn={bn} | {ms*1e3:.1f} us | {tf:.0f} TF | {gbs:.0f} GB/s", flush=True)
        except Exception as e:
            print("ERR", hi, cin, cout, k, impl, bn, str(e)[:100], flush=True)
