import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
# 14x14, Cin 1024 (16 k-blocks), 1x1: MMAs per SM = tiles/SM * 64
n, hi, cin = 256, 14, 1024
for cout, bn in [(32, 32), (64, 64), (128, 128), (256, 256), (512, 256), (192, 192), (96, 96)]:
    ms, tf, gbs = time_conv(2, n, hi, hi, cin, cout, 1, 1, 1, 0, 0, False, bn, 0)
    tiles = (n * hi * hi + 127) // 128 * (cout // bn)
    mmas_per_sm = -(-tiles // 148) * 16 * 4
    cyc = ms * 1e-3 * 1.965e9
    print(f"N={bn}: {ms*1e3:.1f} us, {cyc / mmas_per_sm:.0f} cycles per MMA (max tiles/SM {-(-tiles // 148)}) -> {128*bn*16/(cyc / mmas_per_sm):.0f} MAC/cycle/SM", flush=True)
