"""Runs a few single-layer shapes through the persistent conv kernel (for ncu)."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.tune_conv import time_conv
SHAPES = [  # n, hi, cin, cout, k, stride, pad, res
    (128, 149, 32, 32, 3, 1, 0, False),   # InceptionV3 conv2d_1
    (256, 56, 64, 64, 3, 1, 1, False),    # ResNet50 conv2 3x3
    (256, 56, 64, 256, 1, 1, 0, True),    # ResNet50 conv2 expand + residual
    (256, 14, 256, 256, 3, 1, 1, False),  # ResNet50 conv4 3x3
]
for (n, hi, cin, cout, k, s, p, res) in SHAPES:
    ms, tf, gbs = time_conv(2, n, hi, hi, cin, cout, k, k, s, p, p, res, iters=3, nbuf=2)
    print(f"{hi} {cin}->{cout} k{k} res={res}: {ms*1e3:.1f} us {tf:.0f} TF {gbs:.0f} GB/s", flush=True)
