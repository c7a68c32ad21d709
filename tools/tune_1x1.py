import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
for (hi, cin, cout, res) in [(56, 64, 64, False), (56, 64, 256, True), (56, 256, 64, False), (28, 128, 512, True), (28, 512, 128, False),
                             (14, 256, 1024, True), (14, 1024, 256, False), (7, 512, 2048, True), (7, 2048, 512, False)]:
    for bn in (0, 128, 256):
        if bn and (cout % bn or bn > cout):
            continue
        ms, tf, gbs = time_conv(2, 256, hi, hi, cin, cout, 1, 1, 1, 0, 0, res, bn, 0)
        print(f"{hi}x{hi} {cin}->{cout}{' +res' if res else ''} bn={bn}: {ms*1e3:.1f} us {tf:.0f} TF {gbs:.0f} GB/s", flush=True)
