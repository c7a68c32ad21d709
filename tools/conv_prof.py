"""Time one convolution through dml_conv2d_bf16 for a list of kernel variants / tiles / k-blocks per stage.

    DML_TC_PROF=1 python tools/conv_prof.py n hi wi cin cout k stride pad  impl:tile_n:stages:group [...]

With DML_TC_PROF set, the persistent kernel (impl 2) also prints its per-role cycle counters.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    lib = importlib.import_module("distributed-machine-learning_b200._lib")
    n, hi, wi, cin, cout, k, stride, pad = [int(x) for x in sys.argv[1:9]]
    cin_pad = (cin + 63) // 64 * 64
    ho = (hi + 2 * pad - k) // stride + 1
    rng = np.random.default_rng(0)
    dev = torch.device("cuda")
    x = torch.from_numpy(rng.standard_normal((n, hi, wi, cin)).astype(np.float32)).to(dev).bfloat16()
    wp = torch.zeros((cout, k * k, cin_pad), dtype=torch.bfloat16, device=dev)
    wp[:, :, :cin] = torch.from_numpy((rng.standard_normal((cout, k * k, cin)) / np.sqrt(k * k * cin)).astype(np.float32)).to(dev).bfloat16()
    bias = torch.zeros(cout, dtype=torch.float32, device=dev)
    out = torch.empty((n, ho, ho, cout), dtype=torch.bfloat16, device=dev)
    with_res = os.environ.get("DML_PROF_RES") is not None   # add a residual input (the 1x1 expand layers of ResNet50)
    res = torch.from_numpy(rng.standard_normal((n, ho, ho, cout)).astype(np.float32)).to(dev).bfloat16() if with_res else None
    geom = (n, hi, wi, cin, cin, cout, cout, cout if with_res else 0, k, k, stride, pad, pad, cin_pad, 1)
    flop = 2.0 * n * ho * ho * cout * k * k * cin
    first = None
    for spec in sys.argv[9:]:
        impl, tile_n, stages, group = [int(v) for v in spec.split(":")]

        def run():
            lib.check(lib.lib().dml_conv2d_bf16(impl, x.data_ptr(), wp.data_ptr(), bias.data_ptr(), res.data_ptr() if res is not None else None, out.data_ptr(), *geom,
                                                tile_n, stages | (group << 8) | (int(os.environ.get("DML_PROF_RESMODE", "0")) << 16), None))
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(3):
            e0.record()
            for _ in range(5):
                run()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 5)
        same = ""
        if first is None:
            first = out.clone()
        else:
            same = "bit-identical" if torch.equal(first, out) else "DIFFERS"
        print(f"impl {impl} tile_n {tile_n} stages {stages} group {group}: {best * 1e3:.1f} us  {flop / best / 1e9:.0f} TFLOP/s {same}", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
