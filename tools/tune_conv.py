"""Times the convolution kernels on single layer shapes through dml_conv2d_bf16
(CUDA events, L2-flushing rotation of buffers).  Usage on the GPU box:
   python tools/tune_conv.py [resnet|inception|custom] > gpurun_out/tune.txt
"""
import importlib
import sys
import os

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lib = importlib.import_module("distributed-machine-learning_b200._lib")
L = lib.lib()


def time_conv(impl, n, hi, wi, cin, cout, kh, kw, stride, pad_h, pad_w, res=False, tile_n=0, stages=0, iters=20, nbuf=3):
    dev = "cuda"
    cin_pad = (cin + 63) // 64 * 64
    ho, wo = (hi + 2 * pad_h - kh) // stride + 1, (wi + 2 * pad_w - kw) // stride + 1
    xs = [torch.randn((n, hi, wi, cin), device=dev).bfloat16() for _ in range(nbuf)]
    w = (torch.randn((cout, kh * kw, cin_pad), device=dev) / (kh * kw * cin) ** 0.5).bfloat16()
    b = torch.randn(cout, device=dev)
    outs = [torch.empty((n, ho, wo, cout), dtype=torch.bfloat16, device=dev) for _ in range(nbuf)]
    rs = [torch.randn((n, ho, wo, cout), device=dev).bfloat16() for _ in range(nbuf)] if res else None

    def run(i):
        k = i % nbuf
        rc = L.dml_conv2d_bf16(impl, xs[k].data_ptr(), w.data_ptr(), b.data_ptr(), rs[k].data_ptr() if res else None,
                               outs[k].data_ptr(), n, hi, wi, cin, cin, cout, cout, cout if res else 0, kh, kw, stride,
                               pad_h, pad_w, cin_pad, 1, tile_n, stages, None)
        if rc != 0:
            raise RuntimeError(L.dml_last_error().decode())

    for i in range(3):
        run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    flop = 2.0 * n * ho * wo * cout * kh * kw * cin
    act = (n * hi * wi * cin + n * ho * wo * cout * (2 if res else 1)) * 2
    return ms, flop / ms / 1e9, act / ms / 1e6


RESNET = [  # hi, cin, cout, k, stride, pad, res
    (56, 64, 64, 1, 1, 0, False), (56, 64, 64, 3, 1, 1, False), (56, 64, 256, 1, 1, 0, True), (56, 256, 64, 1, 1, 0, False),
    (56, 256, 512, 1, 2, 0, False), (28, 128, 128, 3, 1, 1, False), (28, 128, 512, 1, 1, 0, True), (28, 512, 128, 1, 1, 0, False),
    (14, 256, 256, 3, 1, 1, False), (14, 256, 1024, 1, 1, 0, True), (14, 1024, 256, 1, 1, 0, False),
    (7, 512, 512, 3, 1, 1, False), (7, 512, 2048, 1, 1, 0, True), (7, 2048, 512, 1, 1, 0, False),
]

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "resnet"
    batches = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [32, 256]
    print("shape | batch | impl bn stages | us | TFLOP/s | act GB/s")
    for (hi, cin, cout, k, stride, pad, res) in RESNET:
        for n in batches:
            if n * hi * hi * max(cin, cout) * 2 > 1.5e9:
                continue
            cands = [(1, 0, 0), (2, 0, 0)]
            for bn in (64, 128, 256):
                if cout % bn == 0 and bn <= cout:
                    cands.append((2, bn, 0))
            for impl, bn, st in cands:
                try:
                    ms, tf, gbs = time_conv(impl, n, hi, hi, cin, cout, k, k, stride, pad, pad, res, bn, st)
                    print(f"{hi}x{hi} {cin}->{cout} k{k} s{stride}{' +res' if res else ''} | {n} | v{impl} bn={bn} st={st} | "
                          f"{ms * 1e3:.1f} | {tf:.0f} | {gbs:.0f}", flush=True)
                except Exception as exc:
                    print("ERR", hi, cin, cout, k, n, impl, bn, exc, flush=True)
