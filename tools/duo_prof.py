"""Time the expand + next-reduce kernel (dml_duo_bf16) against the two separate tcgen05 kernels on one shape.

    [DML_CHAIN_PROF=1] [DML_DUO_PF=n] python tools/duo_prof.py n h w kmid

With DML_CHAIN_PROF set the kernel prints its per-role wait-cycle counters.  Builder tool, not a bench record.
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
    n, h, w, kmid = [int(x) for x in sys.argv[1:5]]
    nexp, nred, m = 4 * kmid, kmid, n * h * w
    rng = np.random.default_rng(0)
    dev = torch.device("cuda")

    def t(shape, scale=1.0, dtype=torch.bfloat16):
        return torch.from_numpy((rng.standard_normal(shape) * scale).astype(np.float32)).to(dev).to(dtype)

    y2, res = t((m, kmid)).relu(), t((m, nexp)).relu()
    we, be = t((nexp, kmid), 1 / np.sqrt(kmid)), t((nexp,), 0.1, torch.float32)
    wr, br = t((nred, nexp), 1 / np.sqrt(nexp)), t((nred,), 0.1, torch.float32)
    out = torch.empty((m, nexp), dtype=torch.bfloat16, device=dev)
    y1n = torch.empty((m, nred), dtype=torch.bfloat16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def fused():
        lib.check(lib.lib().dml_duo_bf16(y2.data_ptr(), we.data_ptr(), be.data_ptr(), res.data_ptr(), None, out.data_ptr(), wr.data_ptr(),
                                         br.data_ptr(), y1n.data_ptr(), n, h, w, kmid, 0, 1, nexp, nred, None))

    def conv(x, wp, b, r, o, cin, cout):
        lib.check(lib.lib().dml_conv2d_bf16(5, x.data_ptr(), wp.data_ptr(), b.data_ptr(), r.data_ptr() if r is not None else None, o.data_ptr(),
                                            n, h, w, cin, cin, cout, cout, cout if r is not None else 0, 1, 1, 1, 0, 0, cin, 1, 0, 0, None))

    def separate():
        conv(y2, we, be, res, out, kmid, nexp)
        conv(out, wr, br, None, y1n, nexp, nred)

    mb = (m * (kmid + 2 * nexp + nred) * 2) / 1e6
    for name, fn in (("fused", fused), ("separate", separate)):
        if name == "separate" and os.environ.get("DML_CHAIN_PROF"):
            continue
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            flush.zero_()  # cold L2, as inside a forward pass over 256 images
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print(f"{name}: {best * 1e3:.1f} us  ({mb:.0f} MB of activations without the re-read -> {mb / best / 1e3:.2f} TB/s)", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
