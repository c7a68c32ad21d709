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
    # The first call on a device does the documented one-time set-up (it looks up the static array's device address;
    # with CUDA's lazy module loading that is also when the driver materialises the 4,096,000-byte array itself).
    # Make it here, so that the two measured calls below must leave the device's free memory exactly unchanged.
    warm = torch.zeros((1, 2048), dtype=torch.float32, device=dev)
    wi = torch.zeros((1, 5), dtype=torch.int32, device=dev)
    wp = torch.zeros((1, 5), dtype=torch.float32, device=dev)
    assert L.dml_head_static(warm.data_ptr(), dense_w.data_ptr(), dense_b.data_ptr(), 1, 2048, None, wi.data_ptr(),
                             wp.data_ptr(), None) == 0, L.dml_last_error()
    torch.cuda.synchronize()
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
            # cudaMemGetInfo sees every allocation on the device, the library's own included (torch's
            # memory_allocated() only counts torch's caching allocator)
            free_before, _ = torch.cuda.mem_get_info()
            rc = L.dml_head_static(pooled.data_ptr(), dense_w.data_ptr(), dense_b.data_ptr(), count, cin, probs.data_ptr(),
                                   idx.data_ptr(), prob.data_ptr(), None)
            assert rc == 0, L.dml_last_error()
            torch.cuda.synchronize()
            free_after, _ = torch.cuda.mem_get_info()
            assert free_after == free_before
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


@pytest.mark.parametrize("cin", [32, 520, 2048, 3072])
@pytest.mark.parametrize("n", [1, 37, 1024])
def test_head_static_sizes_against_fp32_reference(cin, n):
    """Edge sizes of the static head: feature counts that take the TF32 GEMM (multiples of 32) and one that
    takes the exact fp32 kernel (520: a multiple of 8 only, and small enough that the kernel's partial-sum
    area is larger than its activation area), the smallest and the largest image counts."""
    import torch
    L = pkg("_lib").lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(1000 * cin + n)
    pooled = torch.rand((n, cin), generator=g, dtype=torch.float32)
    w = torch.randn((cin, 1000), generator=g, dtype=torch.float32) * 0.05
    b = torch.randn((1000,), generator=g, dtype=torch.float32) * 0.1
    ref = torch.softmax(pooled.double() @ w.double() + b.double(), dim=1).float().numpy()
    d_pooled, d_w, d_b = pooled.to(dev), w.to(dev), b.to(dev)
    probs = torch.full((n, 1000), -1.0, dtype=torch.float32, device=dev)
    idx = torch.full((n, 5), -1, dtype=torch.int32, device=dev)
    prob = torch.full((n, 5), -1.0, dtype=torch.float32, device=dev)
    rc = L.dml_head_static(d_pooled.data_ptr(), d_w.data_ptr(), d_b.data_ptr(), n, cin, probs.data_ptr(), idx.data_ptr(),
                           prob.data_ptr(), None)
    assert rc == 0, L.dml_last_error()
    torch.cuda.synchronize()
    got = probs.cpu().numpy()
    # TF32 operands (10-bit mantissa) for cin % 32 == 0, exact fp32 otherwise
    tol = 2e-3 if cin % 32 == 0 else 1e-5
    assert np.abs(got - ref).max() <= tol
    assert np.abs(got.sum(1) - 1.0).max() <= 1e-4
    # top-5 of the kernel's own probabilities: descending, ties to the larger index
    order = np.lexsort((-np.arange(1000)[None, :].repeat(n, 0), -got), axis=1)[:, :5]
    assert np.array_equal(idx.cpu().numpy(), order.astype(np.int32))
    assert np.array_equal(prob.cpu().numpy(), np.take_along_axis(got, order, axis=1))


def test_head_static_rejects_bad_feature_counts():
    L = pkg("_lib").lib()
    for cin in (0, 4, 12, 3080, 4096):
        assert L.dml_head_static(None, None, None, 1, cin, None, None, None, None) == -4
        assert b"cin" in L.dml_last_error()
