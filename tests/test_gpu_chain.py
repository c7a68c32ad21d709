"""The bottleneck-tail chain kernel (csrc/conv_chain.inc) against the three convolutions it fuses.

ResNet50's conv2 blocks end in a 3x3 (64 -> 64), a 1x1 expand (64 -> 256) with the residual add (or, in the first
block, the folded projection shortcut as a second GEMM source), and are followed by the next block's 1x1 reduce
(256 -> 64) -- /root/reference/models.py:65 runs them as separate Keras layers.  The fused kernel keeps the 3x3's
output in shared memory and feeds the block output to the reduce from shared memory too; it must not change one bit
against the separate tcgen05 kernels, and it must agree with a plain torch fp32 reference of the same three ops.
"""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def _case(n, h, w, variant, seed):
    import torch
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda")

    def t(shape, scale=1.0, dtype=torch.bfloat16):
        return torch.from_numpy((rng.standard_normal(shape) * scale).astype(np.float32)).to(dev).to(dtype)

    c = dict(n=n, h=h, w=w, variant=variant)
    c["y1"] = t((n, h, w, 64)).relu()
    c["w3"] = t((64, 9, 64), 1 / np.sqrt(576))
    c["b3"] = t((64,), 0.1, torch.float32)
    kexp = 128 if variant == "src2" else 64
    c["we"] = t((256, kexp), 1 / np.sqrt(kexp))
    c["be"] = t((256,), 0.1, torch.float32)
    c["res"] = t((n, h, w, 256)).relu() if variant != "src2" else None
    c["src2"] = t((n, h, w, 64)).relu() if variant == "src2" else None
    if variant != "last":
        c["wr"] = t((64, 256), 1 / np.sqrt(256))
        c["br"] = t((64,), 0.1, torch.float32)
    else:
        c["wr"] = c["br"] = None
    return c


def _ptr(t):
    return t.data_ptr() if t is not None else None


def _fused(c):
    import torch
    lib = pkg("_lib")
    n, h, w = c["n"], c["h"], c["w"]
    out = torch.full((n, h, w, 256), 7.0, dtype=torch.bfloat16, device="cuda")
    y1n = torch.full((n, h, w, 64), 7.0, dtype=torch.bfloat16, device="cuda") if c["wr"] is not None else None
    lib.check(lib.lib().dml_chain_bf16(_ptr(c["y1"]), _ptr(c["w3"]), _ptr(c["b3"]), _ptr(c["we"]), _ptr(c["be"]), _ptr(c["res"]),
                                       _ptr(c["src2"]), _ptr(out), _ptr(c["wr"]), _ptr(c["br"]), _ptr(y1n), n, h, w, None))
    torch.cuda.synchronize()
    return out, y1n


def _conv(impl, x, wp, bias, res, n, h, w, cin, cout, k, pad, tile_n=0):
    import torch
    lib = pkg("_lib")
    out = torch.empty((n, h, w, cout), dtype=torch.bfloat16, device="cuda")
    lib.check(lib.lib().dml_conv2d_bf16(impl, x.data_ptr(), wp.data_ptr(), bias.data_ptr(), _ptr(res), out.data_ptr(), n, h, w, cin, cin,
                                        cout, cout, cout if res is not None else 0, k, k, 1, pad, pad, cin, 1, tile_n, 0, None))
    torch.cuda.synchronize()
    return out


def _separate(c):
    """The same three layers through the engine's separate tcgen05 kernels (halo kernel on CTA pairs, pair GEMM kernel)."""
    import torch
    n, h, w = c["n"], c["h"], c["w"]
    y2 = _conv(8, c["y1"], c["w3"], c["b3"], None, n, h, w, 64, 64, 3, 1)
    if c["src2"] is not None:
        # the folded projection shortcut = a second K block: the same GEMM over the concatenated channels
        x = torch.cat([y2, c["src2"]], dim=-1).contiguous()
        out = _conv(5, x, c["we"], c["be"], None, n, h, w, 128, 256, 1, 0)
    else:
        out = _conv(5, y2, c["we"], c["be"], c["res"], n, h, w, 64, 256, 1, 0)
    y1n = _conv(5, out, c["wr"], c["br"], None, n, h, w, 256, 64, 1, 0) if c["wr"] is not None else None
    return y2, out, y1n


def _torch_reference(c):
    import torch
    import torch.nn.functional as F
    x = c["y1"].float().permute(0, 3, 1, 2)
    w3 = c["w3"].float().reshape(64, 3, 3, 64).permute(0, 3, 1, 2)
    y2 = F.conv2d(x, w3, c["b3"], padding=1).relu().bfloat16().float()  # the kernel keeps y2 as bf16, like the separate layers do
    we = c["we"].float()
    o = torch.einsum("nchw,kc->nkhw", y2, we[:, :64]) + c["be"].view(1, -1, 1, 1)
    if c["src2"] is not None:
        o = o + torch.einsum("nhwc,kc->nkhw", c["src2"].float(), we[:, 64:])
    else:
        o = o + c["res"].float().permute(0, 3, 1, 2)
    out = o.relu().bfloat16().float()
    y1n = None
    if c["wr"] is not None:
        y1n = (torch.einsum("nchw,kc->nkhw", out, c["wr"].float()) + c["br"].view(1, -1, 1, 1)).relu().permute(0, 2, 3, 1)
    return out.permute(0, 2, 3, 1), y1n


@pytest.mark.parametrize("variant", ["res", "src2", "last"])
@pytest.mark.parametrize("n", [1, 3, 40])
def test_chain_matches_separate_kernels_bit_for_bit(variant, n):
    import torch
    c = _case(n, 56, 56, variant, seed=100 * n + len(variant))
    out, y1n = _fused(c)
    _, out_s, y1n_s = _separate(c)
    assert torch.equal(out, out_s)
    if variant != "last":
        assert torch.equal(y1n, y1n_s)
    ref_out, ref_y1n = _torch_reference(c)
    assert (out.float() - ref_out).abs().max().item() <= 2e-2 * max(1.0, ref_out.abs().max().item())
    if variant != "last":
        assert (y1n.float() - ref_y1n).abs().max().item() <= 2e-2 * max(1.0, ref_y1n.abs().max().item())


def test_chain_other_geometries_and_refusals():
    """24 x 24 images (4 rows per CTA tile, 96-pixel tiles) work too; shapes the kernel cannot tile are refused with a message."""
    import torch
    lib = pkg("_lib")
    c = _case(5, 24, 24, "res", seed=7)
    out, y1n = _fused(c)
    _, out_s, y1n_s = _separate(c)
    assert torch.equal(out, out_s) and torch.equal(y1n, y1n_s)
    bad = _case(1, 30, 30, "res", seed=8)   # 4 rows of 30 = 120 pixels per tile: not a multiple of the 16-row store boxes
    o = torch.empty((1, 30, 30, 256), dtype=torch.bfloat16, device="cuda")
    y = torch.empty((1, 30, 30, 64), dtype=torch.bfloat16, device="cuda")
    rc = lib.lib().dml_chain_bf16(_ptr(bad["y1"]), _ptr(bad["w3"]), _ptr(bad["b3"]), _ptr(bad["we"]), _ptr(bad["be"]), _ptr(bad["res"]),
                                  None, _ptr(o), _ptr(bad["wr"]), _ptr(bad["br"]), _ptr(y), 1, 30, 30, None)
    assert rc == -1 and b"not eligible" in lib.lib().dml_last_error()


def test_engine_with_and_without_the_chain_gives_the_same_bits(images, monkeypatch):
    """Whole ResNet50 through the engine: fused chains (forced, whatever the autotuner would pick) against DML_NO_CHAIN."""
    kw = pkg("keras_weights")
    W = kw.generate("ResNet50", "folding_stress", 9)
    sel = images[10:47]  # 37 images: the last pair tile of the grid is not the last of an image
    outs = []
    for chain in (True, False):
        monkeypatch.delenv("DML_NO_CHAIN", raising=False)
        monkeypatch.delenv("DML_FORCE_CHAIN", raising=False)
        monkeypatch.setenv("DML_FORCE_CHAIN" if chain else "DML_NO_CHAIN", "1")
        monkeypatch.setenv("DML_NO_DUO", "1")  # (the conv3 / conv4 duos have their own test)
        eng = pkg("engine").Engine("ResNet50", max_batch=40, weights=W)
        try:
            names = [o[0] for o in eng.ops() if o[1] == 0 and o[3] == 0]  # convs that report no work of their own: chain members
            assert (len(names) == 5) == chain, names  # 3 expands + 2 next-block reduces
            idx, prob, probs = eng.predict_images(sel, want_probs=True)
            outs.append((idx, prob, probs, eng.logits(len(sel)).cpu().numpy()))
        finally:
            eng.close()
    for a, b in zip(outs[0], outs[1]):
        assert np.array_equal(a, b)
