"""The expand + next-reduce kernel (csrc/conv_duo.inc) against the two convolutions it fuses.

ResNet50's conv3 / conv4 identity blocks end in a 1x1 expand (128 -> 512 at 28 x 28, 256 -> 1024 at 14 x 14) with the residual
add, followed by the next block's 1x1 reduce -- /root/reference/models.py:65 runs them as separate Keras layers.  The fused
kernel feeds the block output to the reduce from shared memory; it must not change one bit against the separate tcgen05
kernels, and it must agree with a plain torch fp32 reference of the same two ops.
"""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def _case(n, h, w, kmid, seed, k2=0, s2=1):
    import torch
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda")
    nexp, nred = 4 * kmid, kmid

    def t(shape, scale=1.0, dtype=torch.bfloat16):
        return torch.from_numpy((rng.standard_normal(shape) * scale).astype(np.float32)).to(dev).to(dtype)

    c = dict(n=n, h=h, w=w, kmid=kmid, nexp=nexp, nred=nred, k2=k2, s2=s2)
    c["y2"] = t((n, h, w, kmid)).relu()
    c["we"] = t((nexp, kmid + k2), 1 / np.sqrt(kmid + k2))
    c["be"] = t((nexp,), 0.1, torch.float32)
    c["res"] = t((n, h, w, nexp)).relu() if k2 == 0 else None      # identity shortcut ...
    c["src2"] = t((n, h * s2, w * s2, k2)).relu() if k2 else None    # ... or the projection shortcut's input (read at stride s2) as a second GEMM source
    c["wr"] = t((nred, nexp), 1 / np.sqrt(nexp))
    c["br"] = t((nred,), 0.1, torch.float32)
    return c


def _ptr(t):
    return t.data_ptr() if t is not None else None


def _fused(c):
    import torch
    lib = pkg("_lib")
    n, h, w = c["n"], c["h"], c["w"]
    out = torch.full((n, h, w, c["nexp"]), 7.0, dtype=torch.bfloat16, device="cuda")
    y1n = torch.full((n, h, w, c["nred"]), 7.0, dtype=torch.bfloat16, device="cuda")
    lib.check(lib.lib().dml_duo_bf16(c["y2"].data_ptr(), c["we"].data_ptr(), c["be"].data_ptr(), _ptr(c["res"]), _ptr(c["src2"]), out.data_ptr(),
                                     c["wr"].data_ptr(), c["br"].data_ptr(), y1n.data_ptr(), n, h, w, c["kmid"], c["k2"], c["s2"], c["nexp"], c["nred"], None))
    torch.cuda.synchronize()
    return out, y1n


def _conv(x, wp, bias, res, n, h, w, cin, cout):
    import torch
    lib = pkg("_lib")
    out = torch.empty((n, h, w, cout), dtype=torch.bfloat16, device="cuda")
    lib.check(lib.lib().dml_conv2d_bf16(5, x.data_ptr(), wp.data_ptr(), bias.data_ptr(), res.data_ptr() if res is not None else None,
                                        out.data_ptr(), n, h, w, cin, cin, cout, cout, cout if res is not None else 0, 1, 1, 1, 0, 0, cin,
                                        1, 0, 0, None))
    torch.cuda.synchronize()
    return out


def _separate(c):
    import torch
    n, h, w = c["n"], c["h"], c["w"]
    if c["src2"] is not None:
        # the folded projection shortcut = more K blocks: the same GEMM over the concatenated channels
        x = torch.cat([c["y2"], c["src2"][:, ::c["s2"], ::c["s2"], :]], dim=-1).contiguous()
        out = _conv(x, c["we"], c["be"], None, n, h, w, c["kmid"] + c["k2"], c["nexp"])
    else:
        out = _conv(c["y2"], c["we"], c["be"], c["res"], n, h, w, c["kmid"], c["nexp"])
    y1n = _conv(out, c["wr"], c["br"], None, n, h, w, c["nexp"], c["nred"])
    return out, y1n


def torch_cat(a, b):
    import torch
    return torch.cat([a, b], dim=-1)


def _torch_reference(c):
    if c["src2"] is not None:
        o = torch_cat(c["y2"], c["src2"][:, ::c["s2"], ::c["s2"], :]).float() @ c["we"].float().t() + c["be"]
    else:
        o = c["y2"].float() @ c["we"].float().t() + c["be"] + c["res"].float()
    out = o.relu().bfloat16().float()
    y1n = (out @ c["wr"].float().t() + c["br"]).relu()
    return out, y1n


@pytest.mark.parametrize("n,h,w,kmid,k2,s2", [(1, 14, 14, 256, 0, 1), (3, 28, 28, 128, 0, 1), (40, 28, 28, 128, 0, 1), (37, 14, 14, 256, 0, 1),
                                              (256, 14, 14, 256, 0, 1), (70, 28, 28, 128, 0, 1), (2, 16, 16, 64, 0, 1), (5, 12, 12, 192, 0, 1),
                                              (1, 56, 56, 64, 64, 1), (37, 56, 56, 64, 64, 1), (3, 20, 20, 128, 64, 1),
                                              (1, 28, 28, 128, 256, 2), (37, 28, 28, 128, 256, 2), (5, 10, 10, 64, 128, 2)])
def test_duo_matches_separate_kernels_bit_for_bit(n, h, w, kmid, k2, s2):
    import torch
    c = _case(n, h, w, kmid, seed=n + kmid, k2=k2, s2=s2)
    out, y1n = _fused(c)
    out_s, y1n_s = _separate(c)
    assert torch.equal(out, out_s)
    assert torch.equal(y1n, y1n_s)
    ref_out, ref_y1n = _torch_reference(c)
    assert (out.float() - ref_out).abs().max().item() <= 2e-2 * max(1.0, ref_out.abs().max().item())
    assert (y1n.float() - ref_y1n).abs().max().item() <= 2e-2 * max(1.0, ref_y1n.abs().max().item())


def test_duo_refuses_what_it_cannot_tile():
    import torch
    lib = pkg("_lib")
    c = _case(1, 8, 8, 128, seed=3)
    out = torch.empty((64, 512), dtype=torch.bfloat16, device="cuda")
    y = torch.empty((64, 128), dtype=torch.bfloat16, device="cuda")
    rc = lib.lib().dml_duo_bf16(c["y2"].data_ptr(), c["we"].data_ptr(), c["be"].data_ptr(), c["res"].data_ptr(), None, out.data_ptr(),
                                c["wr"].data_ptr(), c["br"].data_ptr(), y.data_ptr(), 1, 8, 8, 96, 0, 1, 512, 128, None)  # kmid not a multiple of 64
    assert rc == -1 and b"not eligible" in lib.lib().dml_last_error()


def test_engine_with_and_without_the_duo_gives_the_same_bits(images, monkeypatch):
    """Whole ResNet50 through the engine: fused duos (forced, whatever the autotuner would pick) against DML_NO_DUO."""
    kw = pkg("keras_weights")
    W = kw.generate("ResNet50", "folding_stress", 9)
    sel = images[10:47]  # 37 images: partial last tiles at 28 x 28 and 14 x 14
    outs = []
    for duo in (True, False):
        monkeypatch.delenv("DML_NO_DUO", raising=False)
        monkeypatch.delenv("DML_FORCE_DUO", raising=False)
        monkeypatch.setenv("DML_NO_CHAIN", "1")
        monkeypatch.setenv("DML_FORCE_DUO" if duo else "DML_NO_DUO", "1")
        eng = pkg("engine").Engine("ResNet50", max_batch=40, weights=W)
        try:
            names = [o[0] for o in eng.ops() if o[1] == 0 and o[3] == 0]  # convs that report no work of their own: duo members
            assert (len(names) == 8) == duo, names  # next-block reduces: conv2_block2's and conv3_block2's (behind folded shortcuts), 2 more in conv3, 4 in conv4
            idx, prob, probs = eng.predict_images(sel, want_probs=True)
            outs.append((idx, prob, probs, eng.logits(len(sel)).cpu().numpy()))
        finally:
            eng.close()
    for a, b in zip(outs[0], outs[1]):
        assert np.array_equal(a, b)
