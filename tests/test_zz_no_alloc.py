"""dml_head_static: the allocation-free dense + softmax + top-5 entry point.

Some users of this library cannot allocate device memory at the point where they
need the classifier head: they are inside a stream capture, on a latency-critical
serving thread that must not hit cudaMalloc's device-wide synchronisation, or inside
a host process whose GPU memory is owned by another framework's pool.  For them the
library keeps the n x 1000 logits in a fixed array of its own (at most 1024 images per
call).  This test drives it through the C ABI twice in a row and checks both results
against the regular head (dml_infer), which uses the caller-provided workspace.
The file name sorts last on purpose: a -x run has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def test_head_static_matches_regular_head_twice(images):
    import torch
    lib = pkg("_lib")
    L = lib.lib()
    kw = pkg("keras_weights")
    model = "ResNet50"
    W = kw.generate(model)
    names = [n for n, _ in lib.weight_spec(model)]
    dev = torch.device("cuda", 0)
    dense_w = torch.from_numpy(W[names.index("predictions/kernel")]).to(dev).contiguous()
    dense_b = torch.from_numpy(W[names.index("predictions/bias")]).to(dev).contiguous()
    eng = pkg("engine").Engine(model, max_batch=32, weights=W)
    try:
        for first, count in ((0, 32), (40, 9)):  # two calls in a row; the second reuses the static array
            sel = images[first:first + count]
            idx_ref, prob_ref, probs_ref = eng.predict_images(sel, want_probs=True)
            _, _, _, cin, _, eb, ptr = eng.tensors()["avg_pool"]
            assert eb == 4 and cin == 2048
            # the engine's pooled features, copied out so that the inputs are plainly the caller's
            off = ptr - eng.workspace.data_ptr()
            pooled = eng.workspace[off:off + count * cin * 4].view(torch.float32).view(count, cin).clone()
            probs = torch.full((count, 1000), -1.0, dtype=torch.float32, device=dev)
            idx = torch.full((count, 5), -1, dtype=torch.int32, device=dev)
            prob = torch.full((count, 5), -1.0, dtype=torch.float32, device=dev)
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            rc = L.dml_head_static(pooled.data_ptr(), dense_w.data_ptr(), dense_b.data_ptr(), count, cin, probs.data_ptr(),
                                   idx.data_ptr(), prob.data_ptr(), None)
            assert rc == 0, L.dml_last_error()
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            # same kernels, same inputs: the tolerance of the other head tests is 0 here too
            assert np.array_equal(idx.cpu().numpy(), idx_ref)
            assert np.abs(prob.cpu().numpy() - prob_ref).max() <= 1e-6
            assert np.abs(probs.cpu().numpy() - probs_ref).max() <= 1e-6
        # sizes beyond the static scratch are refused with a clear message, nothing is launched
        rc = L.dml_head_static(pooled.data_ptr(), dense_w.data_ptr(), dense_b.data_ptr(), 1025, cin, None, idx.data_ptr(),
                               prob.data_ptr(), None)
        assert rc == -4 and b"1024" in L.dml_last_error()
        assert L.dml_head_static(None, None, None, 0, cin, None, None, None, None) == 0  # n = 0 is a no-op
    finally:
        eng.close()
