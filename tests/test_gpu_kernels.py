"""GPU parity of the individual kernels through the C ABI: K1 (bit-exact against
Pillow + numpy), and the convolution kernels against a torch fp32 reference on
bf16-rounded operands (the tcgen05 kernel additionally against the CUDA-core one)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

from conftest import pkg
from oracle import frontend as ofe

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def engines():
    E = pkg("engine").Engine
    out = {}
    for m in ("ResNet50", "InceptionV3"):
        out[m, "fp32"] = E(m, max_batch=200, micro_batch=8, fp32_check=True, load=False)
        out[m, "bf16"] = E(m, max_batch=200, micro_batch=8, load=False)
    yield out
    for e in out.values():
        e.close()


@pytest.mark.parametrize("model", ["ResNet50", "InceptionV3"])
def test_preprocess_bit_exact_all_200(model, engines, images, jpeg_bytes, frontend_golden):
    """K1 == load_img (Pillow NEAREST) + preprocess_input, bit for bit, on every bundled image."""
    import torch
    fe = pkg("frontend")
    size, mode = ofe.MODEL_INPUT[model]
    g = frontend_golden
    for kind in ("fp32", "bf16"):
        eng = engines[model, kind]
        hb = fe.HostBatch(images, size)
        eng.upload(hb)
        eng.preprocess(hb.n)
        eng.sync()
        got = eng.input_tensor(hb.n)
        if kind == "fp32":
            got = got.cpu().numpy()
            for i in range(hb.n):
                assert sha(got[i]) == g[f"sha_pre_{size}_{mode}"][i], (model, i)
        else:
            # bf16 mode stores the round-to-nearest-even bf16 of the same fp32 value
            for i in range(0, hb.n, 7):
                want = ofe.load_and_preprocess(jpeg_bytes[str(g["names"][i])], model)
                w16 = torch.from_numpy(want).to(torch.bfloat16)
                assert torch.equal(got[i].cpu(), w16), (model, i)


def test_preprocess_ragged_and_empty(engines):
    fe = pkg("frontend")
    eng = engines["InceptionV3", "fp32"]
    rng = np.random.default_rng(1)
    imgs = [rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8) for h, w in [(1, 1), (299, 299), (300, 169), (114, 300), (7, 513)]]
    hb = fe.HostBatch(imgs, 299)
    eng.upload(hb)
    eng.preprocess(hb.n)
    eng.sync()
    got = eng.input_tensor(hb.n).cpu().numpy()
    for i, im in enumerate(imgs):
        want = ofe.preprocess(ofe.resize_nearest_tables(im, 299), "tf")
        assert np.array_equal(got[i].view(np.uint32), want.view(np.uint32)), i
    eng.preprocess(0)  # n = 0 is a no-op, not an error


def _conv_case(rng, n, hi, wi, cin, cout, kh, kw, stride, pad_h, pad_w, relu, with_res, in_extra=0, out_extra=0, cin_pad=None):
    import torch
    import torch.nn.functional as F
    dev = torch.device("cuda")
    in_pitch = cin + in_extra
    cin_pad = cin_pad or (cin + 63) // 64 * 64
    ho, wo = (hi + 2 * pad_h - kh) // stride + 1, (wi + 2 * pad_w - kw) // stride + 1
    out_pitch = cout + out_extra
    x = torch.from_numpy(rng.standard_normal((n, hi, wi, in_pitch)).astype(np.float32)).to(dev).bfloat16()
    w = torch.from_numpy((rng.standard_normal((cout, kh, kw, cin)) / np.sqrt(kh * kw * cin)).astype(np.float32)).to(dev).bfloat16()
    wp = torch.zeros((cout, kh * kw, cin_pad), dtype=torch.bfloat16, device=dev)
    wp[:, :, :cin] = w.reshape(cout, kh * kw, cin)
    bias = torch.from_numpy(rng.standard_normal(cout).astype(np.float32)).to(dev)
    res = torch.from_numpy(rng.standard_normal((n, ho, wo, cout)).astype(np.float32)).to(dev).bfloat16() if with_res else None
    ref = F.conv2d(x[..., :cin].float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), bias, stride=stride,
                   padding=(pad_h, pad_w)).permute(0, 2, 3, 1)
    if with_res:
        ref = ref + res.float()
    if relu:
        ref = ref.relu()
    return dict(x=x, wp=wp, bias=bias, res=res, ref=ref, geom=(n, hi, wi, cin, in_pitch, cout, out_pitch, cout if with_res else 0,
                kh, kw, stride, pad_h, pad_w, cin_pad, int(relu)), ho=ho, wo=wo)


def _run_conv(impl, case, tile_n=0, stages=0):
    import torch
    lib = pkg("_lib")
    n, hi, wi, cin, in_pitch, cout, out_pitch = case["geom"][:7]
    out = torch.full((n, case["ho"], case["wo"], out_pitch), 7.0, dtype=torch.bfloat16, device="cuda")
    res = case["res"]
    lib.check(lib.lib().dml_conv2d_bf16(impl, case["x"].data_ptr(), case["wp"].data_ptr(), case["bias"].data_ptr(),
                                        res.data_ptr() if res is not None else None, out.data_ptr(), *case["geom"],
                                        tile_n, stages, None))
    torch.cuda.synchronize()
    return out


CONV_SHAPES = [
    # n, hi, wi, cin, cout, kh, kw, stride, pad_h, pad_w, relu, res, in_extra, out_extra
    (2, 56, 56, 64, 64, 1, 1, 1, 0, 0, True, False, 0, 0),
    (2, 56, 56, 64, 256, 1, 1, 1, 0, 0, True, True, 0, 0),
    (3, 56, 56, 256, 128, 1, 1, 2, 0, 0, True, False, 0, 0),
    (2, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0),
    (5, 7, 7, 512, 512, 3, 3, 1, 1, 1, True, False, 0, 0),
    (3, 14, 14, 1024, 2048, 1, 1, 2, 0, 0, False, False, 0, 0),
    (2, 35, 35, 48, 64, 5, 5, 1, 2, 2, True, False, 0, 32),      # ragged Cin, concat-slice store
    (2, 35, 35, 288, 384, 3, 3, 2, 0, 0, True, False, 0, 384),   # 3x3 s2 valid into a concat slice
    (2, 17, 17, 160, 160, 1, 7, 1, 0, 3, True, False, 0, 0),
    (2, 17, 17, 160, 192, 7, 1, 1, 3, 0, True, False, 0, 576),
    (3, 8, 8, 448, 384, 3, 3, 1, 1, 1, True, False, 0, 0),
    (2, 8, 8, 384, 384, 1, 3, 1, 0, 1, True, False, 0, 1664),
    (2, 8, 8, 384, 384, 3, 1, 1, 1, 0, True, False, 0, 1664),
    (2, 73, 73, 80, 192, 3, 3, 1, 0, 0, True, False, 0, 0),
    (1, 147, 147, 32, 64, 3, 3, 1, 1, 1, True, False, 0, 0),
    (2, 17, 17, 192, 320, 3, 3, 2, 0, 0, True, False, 576, 0),   # input is a channel slice of a wider buffer
    (1, 5, 5, 64, 32, 1, 1, 1, 0, 0, False, False, 0, 0),        # tiny: one partial M tile
]


@pytest.mark.parametrize("shape", CONV_SHAPES, ids=lambda s: "x".join(map(str, s[:10])))
def test_conv_cuda_core_vs_torch(shape):
    rng = np.random.default_rng(hash(shape) % 2**32)
    case = _conv_case(rng, *shape)
    out = _run_conv(0, case)
    cout = shape[4]
    got = out[..., :cout].float()
    tol = 2e-2 * max(1.0, case["ref"].abs().max().item())  # one bf16 rounding of the output
    assert (got - case["ref"]).abs().max().item() <= tol
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()  # neighbours of the channel slice untouched


def _need_legacy():
    """impl 1 / 3 / 4 are superseded A/B twins: only in libraries built with DML_B200_LEGACY_VARIANTS=1 (csrc: -DDML_LEGACY_VARIANTS)."""
    if not pkg("_lib").has_legacy_variants():
        pytest.skip("superseded kernel variant: not in the product build (DML_B200_LEGACY_VARIANTS=1 builds it)")


@pytest.mark.parametrize("impl", [1, 2], ids=["tile_per_cta", "persistent"])
@pytest.mark.parametrize("shape", CONV_SHAPES, ids=lambda s: "x".join(map(str, s[:10])))
def test_conv_tcgen05_vs_cuda_core_and_torch(shape, impl):
    if impl == 1:
        _need_legacy()
    rng = np.random.default_rng(hash(shape) % 2**32)
    case = _conv_case(rng, *shape)
    base = _run_conv(0, case)
    out = _run_conv(impl, case)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    tol = 2e-2 * max(1.0, ref.abs().max().item())
    assert (got - ref).abs().max().item() <= tol
    # same bf16 operands, fp32 accumulation in both kernels: equal up to summation order
    d = (got - base[..., :cout].float()).abs().max().item()
    assert d <= 1.6e-2 * max(1.0, ref.abs().max().item()), d
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()


PERSISTENT_CASES = [
    # shape, tile_n, stages: many tiles per CTA, both accumulators, ring wrap-around, all store widths
    ((8, 56, 56, 64, 256, 1, 1, 1, 0, 0, True, True, 0, 0), 0, 0),
    ((8, 56, 56, 64, 256, 1, 1, 1, 0, 0, True, True, 0, 0), 64, 2),
    ((8, 56, 56, 64, 256, 1, 1, 1, 0, 0, False, False, 0, 0), 256, 0),
    ((16, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0), 128, 3),
    ((16, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0), 32, 0),
    ((9, 35, 35, 64, 96, 3, 3, 1, 1, 1, True, False, 0, 160), 0, 0),    # bn 96 -> 32-column boxes
    ((9, 35, 35, 48, 64, 5, 5, 1, 2, 2, True, False, 0, 32), 16, 2),     # 16-column boxes
    ((6, 73, 73, 64, 80, 1, 1, 1, 0, 0, True, False, 0, 0), 0, 0),       # bn 80 -> 16-column boxes
    ((32, 14, 14, 1024, 256, 1, 1, 1, 0, 0, True, False, 0, 0), 0, 0),
]


@pytest.mark.parametrize("case_id", range(len(PERSISTENT_CASES)))
def test_conv_tcgen05_persistent_tiles(case_id):
    shape, tile_n, stages = PERSISTENT_CASES[case_id]
    rng = np.random.default_rng(1234 + case_id)
    case = _conv_case(rng, *shape)
    out = _run_conv(2, case, tile_n, stages)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    assert (got - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()


@pytest.mark.parametrize("shape", [(3, 224, 224, 3, 64, 7, 7, 2, 3, 3, True, False), (2, 299, 299, 3, 32, 3, 3, 2, 0, 0, True, False),
                                   (1, 37, 53, 3, 64, 7, 7, 2, 3, 3, True, False)], ids=["resnet", "inception", "ragged"])
def test_stem_conv_tcgen05(shape):
    _need_legacy()
    rng = np.random.default_rng(99)
    case = _conv_case(rng, *shape, cin_pad=3)
    ref = case["ref"]
    tol = 2e-2 * max(1.0, ref.abs().max().item())
    base = _run_conv(0, case)
    assert (base.float() - ref).abs().max().item() <= tol
    out = _run_conv(3, case)
    assert (out.float() - ref).abs().max().item() <= tol


HALO_CASES = [
    # n, hi, wi, cin, cout, kh, kw, stride, pad_h, pad_w, relu, res, in_extra, out_extra ; tile_n, stages
    ((2, 56, 56, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0), 0, 0),
    ((5, 56, 56, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0), 32, 3),
    ((3, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0), 0, 0),
    ((9, 28, 28, 128, 128, 3, 3, 1, 1, 1, False, False, 0, 0), 64, 0),
    ((4, 14, 14, 256, 256, 3, 3, 1, 1, 1, True, False, 0, 0), 0, 0),
    ((2, 35, 35, 64, 96, 3, 3, 1, 0, 0, True, False, 0, 160), 0, 0),      # valid conv, concat-slice store, 32-column chunks
    ((2, 35, 35, 64, 64, 5, 5, 1, 2, 2, True, False, 0, 32), 0, 0),       # 5x5 same
    ((3, 17, 17, 128, 192, 1, 7, 1, 0, 3, True, False, 0, 0), 0, 0),      # 1x7
    ((3, 17, 17, 192, 192, 7, 1, 1, 3, 0, True, False, 576, 0), 0, 0),    # 7x1, input is a channel slice
    ((1, 147, 147, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0), 0, 0),      # rows wider than one tile
    ((2, 8, 8, 384, 384, 1, 3, 1, 0, 1, True, False, 0, 1664), 0, 0),
]


@pytest.mark.parametrize("case_id", range(len(HALO_CASES)))
def test_conv_tcgen05_halo_kernel(case_id):
    _need_legacy()
    shape, tile_n, stages = HALO_CASES[case_id]
    rng = np.random.default_rng(4321 + case_id)
    case = _conv_case(rng, *shape)
    out = _run_conv(4, case, tile_n, stages)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    assert (got - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()


PAIR_CASES = [
    ((8, 56, 56, 64, 256, 1, 1, 1, 0, 0, True, True, 0, 0), 0, 0),      # 1x1 + residual, tiled A
    ((8, 56, 56, 64, 256, 1, 1, 1, 0, 0, True, False, 0, 0), 256, 0),
    ((3, 56, 56, 256, 128, 1, 1, 2, 0, 0, True, False, 0, 0), 0, 0),     # stride-2 1x1 (im2col)
    ((16, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0), 0, 0),
    ((5, 7, 7, 512, 512, 3, 3, 1, 1, 1, True, False, 0, 0), 0, 2),       # odd tile count, 2 stages
    ((32, 14, 14, 1024, 256, 1, 1, 1, 0, 0, True, False, 0, 0), 0, 0),
    ((9, 35, 35, 64, 96, 3, 3, 1, 1, 1, True, False, 0, 160), 0, 0),     # bn 96, concat slice
    ((2, 17, 17, 160, 192, 7, 1, 1, 3, 0, True, False, 0, 576), 0, 0),
    ((1, 5, 5, 64, 32, 1, 1, 1, 0, 0, False, False, 0, 0), 0, 0),        # one partial tile: peer CTA has no rows
    ((2, 73, 73, 80, 192, 3, 3, 1, 0, 0, True, False, 0, 0), 0, 0),      # Cin 80: ragged last channel block
    ((2, 73, 73, 80, 192, 3, 3, 1, 0, 0, True, False, 0, 0), 64, 2 | (3 << 8)),  # the same, 3 k-blocks per stage
    ((3, 35, 35, 288, 64, 1, 1, 1, 0, 0, True, False, 0, 224), 0, 0),    # Cin 288, 1x1 tiled A into a concat slice
    ((2, 17, 17, 96, 96, 3, 3, 2, 0, 0, True, False, 32, 0), 0, 0),      # Cin 96 stride 2, input is a channel slice
]


@pytest.mark.parametrize("case_id", range(len(PAIR_CASES)))
def test_conv_tcgen05_cta_pair(case_id):
    shape, tile_n, stages = PAIR_CASES[case_id]
    rng = np.random.default_rng(777 + case_id)
    case = _conv_case(rng, *shape)
    out = _run_conv(5, case, tile_n, stages)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    assert (got - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()


RESIDENT_HALO_CASES = [
    # n, hi, wi, cin, cout, kh, kw, stride, pad_h, pad_w, relu, res, in_extra, out_extra
    (2, 56, 56, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0),         # ResNet50 conv2 3x3
    (300, 14, 14, 64, 64, 3, 3, 1, 1, 1, False, False, 0, 0),      # more tiles than SMs: ring and accumulators wrap
    (2, 149, 149, 32, 32, 3, 3, 1, 0, 0, True, False, 0, 0),       # InceptionV3 conv2d_1: 64-byte rows, valid conv
    (1, 147, 147, 32, 64, 3, 3, 1, 1, 1, True, False, 0, 0),       # conv2d_2: 64-byte rows, same padding
    (3, 35, 35, 64, 96, 3, 3, 1, 1, 1, True, False, 0, 160),       # Cout 96 into a concat slice
    (2, 35, 35, 48, 64, 3, 3, 1, 1, 1, True, False, 16, 0),        # ragged Cin, input is a channel slice
    (2, 20, 20, 16, 32, 5, 5, 1, 2, 2, True, False, 0, 0),         # 5x5, one K step
    (2, 17, 17, 24, 48, 1, 7, 1, 0, 3, False, False, 8, 16),       # 1x7, ragged 64-byte rows
    (2, 17, 17, 96, 48, 7, 1, 1, 3, 0, True, False, 0, 0),         # 7x1, two channel blocks, Cout 48
    (2, 30, 30, 128, 32, 3, 3, 1, 1, 1, True, False, 0, 0),        # 3x3 with two channel blocks (compile-time path)
    (3, 112, 112, 16, 64, 4, 4, 1, 2, 2, True, False, 0, 0),       # 32-byte rows: the space-to-depth ResNet50 stem's shape
    (2, 150, 150, 16, 32, 2, 2, 1, 0, 0, True, False, 0, 0),       # 32-byte rows, 2x2 valid: the InceptionV3 stem's shape
    (2, 21, 21, 8, 48, 3, 3, 1, 1, 1, False, False, 8, 0),         # 32-byte rows, generic path, ragged channels
    (3, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0),       # filter too big for one CTA: Cout split in 2 slices
    (2, 35, 35, 48, 64, 5, 5, 1, 2, 2, True, False, 0, 32),        # 5x5, Cout split, concat slice
    (2, 17, 17, 128, 128, 1, 7, 1, 0, 3, True, False, 0, 0),       # 1x7, two channel blocks, Cout split
    (2, 17, 17, 128, 192, 7, 1, 1, 3, 0, False, False, 0, 576),    # 7x1, Cout 192 in 3 or 4 slices
]


@pytest.mark.parametrize("shape", RESIDENT_HALO_CASES, ids=lambda s: "x".join(map(str, s[:10])))
def test_conv_tcgen05_resident_filter_halo(shape):
    """impl 7: halo kernel with the whole filter resident in shared memory (both row widths)."""
    import torch
    rng = np.random.default_rng(hash(shape) % 2**32)
    case = _conv_case(rng, *shape)
    out = _run_conv(7, case)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    assert (got - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()
    assert torch.equal(out, _run_conv(2, case))   # same bits as the im2col kernel


PAIR_HALO_CASES = [
    (3, 28, 28, 128, 128, 3, 3, 1, 1, 1, True, False, 0, 0),       # ResNet50 conv3 3x3
    (5, 28, 28, 128, 128, 3, 3, 1, 1, 1, False, False, 0, 0),      # odd image count: tiles wrap over the clusters
    (2, 35, 35, 48, 64, 5, 5, 1, 2, 2, True, False, 0, 32),        # InceptionV3 5x5, ragged Cin, concat slice
    (2, 35, 35, 96, 96, 3, 3, 1, 1, 1, True, False, 0, 160),       # 3x3 96 -> 96 (two channel blocks, second one ragged)
    (2, 17, 17, 128, 128, 1, 7, 1, 0, 3, True, False, 0, 0),       # 1x7
    (2, 17, 17, 160, 192, 7, 1, 1, 3, 0, True, False, 0, 576),     # 7x1, three channel blocks, Cout 192
    (300, 14, 14, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0),       # many tiles per cluster: rings and accumulators wrap
    (2, 56, 56, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0),         # two rows per tile
    (3, 9, 9, 32, 32, 3, 3, 1, 0, 0, False, False, 0, 0),          # 64-byte rows, valid conv, generic loop
]


@pytest.mark.parametrize("shape", PAIR_HALO_CASES, ids=lambda s: "x".join(map(str, s[:10])))
def test_conv_tcgen05_pair_resident_filter_halo(shape):
    """impl 8: resident-filter halo kernel on CTA pairs (row-aligned tiles)."""
    import torch
    rng = np.random.default_rng(hash(shape) % 2**32)
    case = _conv_case(rng, *shape)
    out = _run_conv(8, case)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    assert (got - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()
    assert torch.equal(out, _run_conv(2, case))   # same bits as the im2col kernel


NARROW_ROW_CASES = [
    # Cin that 32-channel K blocks (64-byte shared-memory rows, bit 18 of `stages`) pad less than 64-channel ones
    (2, 73, 73, 80, 192, 3, 3, 1, 0, 0, True, False, 0, 0),         # InceptionV3 conv2d_4: 80 -> 96 instead of 128
    (3, 35, 35, 96, 96, 3, 3, 1, 1, 1, True, False, 0, 160),        # mixed0..2 3x3: 96 instead of 128, concat-slice store
    (3, 17, 17, 160, 160, 1, 7, 1, 0, 3, True, False, 0, 0),        # mixed5/6 1x7: 160 instead of 192
    (5, 17, 17, 160, 192, 7, 1, 1, 3, 0, True, False, 32, 0),       # 7x1, input is a channel slice
    (2, 17, 17, 128, 128, 1, 7, 1, 0, 3, True, False, 0, 0),        # no padding either way: still the same bits
]


@pytest.mark.parametrize("impl", [7, 8], ids=["one_cta", "cta_pair"])
@pytest.mark.parametrize("shape", NARROW_ROW_CASES, ids=lambda s: "x".join(map(str, s[:10])))
def test_conv_tcgen05_halo_kernels_with_64_byte_rows(shape, impl):
    import torch
    rng = np.random.default_rng(hash(shape) % 2**32)
    case = _conv_case(rng, *shape)
    try:
        out = _run_conv(impl, case, 0, 1 << 18)
    except pkg("_lib").DmlError as e:
        if "does not fit" in str(e):
            pytest.skip("the whole filter of this layer does not fit next to two halo stages in one CTA's shared memory")
        raise
    cout = shape[4]
    ref = case["ref"]
    assert (out[..., :cout].float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()
    assert torch.equal(out, _run_conv(2, case))      # same bits as the im2col kernel
    try:
        assert torch.equal(out, _run_conv(impl, case))   # and as the same kernel with 128-byte rows
    except pkg("_lib").DmlError as e:
        assert "does not fit" in str(e)  # (with 128-byte rows this layer's filter needs more shared memory than one CTA has)


def test_conv_tcgen05_variants_are_bit_identical():
    """Tile shape and kernel variant must not change a single bit (the engine may pick any of them)."""
    import torch
    for shape in [(4, 28, 28, 128, 256, 3, 3, 1, 1, 1, True, False, 0, 0), (6, 56, 56, 64, 256, 1, 1, 1, 0, 0, True, True, 0, 0),
                  (3, 17, 17, 160, 192, 1, 7, 1, 0, 3, True, False, 0, 0)]:
        rng = np.random.default_rng(2024)
        case = _conv_case(rng, *shape)
        big = 128 if shape[4] % 128 == 0 else 96
        g = lambda k, stages=0: (k << 8) | stages   # k-blocks per pipeline stage ride in bits 8..15 of `stages`
        outs = [_run_conv(2, case, 0, 0), _run_conv(2, case, 64, 2), _run_conv(2, case, big, 0), _run_conv(5, case, 0, 0),
                _run_conv(5, case, 64, 3), _run_conv(2, case, 64, g(1)), _run_conv(2, case, 64, g(4)),
                _run_conv(2, case, 32, g(5, 2)), _run_conv(5, case, 64, g(1)), _run_conv(5, case, 64, g(3)),
                _run_conv(5, case, big, g(2, 2)), _run_conv(5, case, 32, g(7)),
                _run_conv(5, case, 0, 1 << 16), _run_conv(5, case, 0, 2 << 16), _run_conv(5, case, 64, (2 << 16) | g(2))]  # residual paths
        if pkg("_lib").has_legacy_variants():
            outs.append(_run_conv(1, case, 0, 0))
        for o in outs[1:]:
            assert torch.equal(o, outs[0])


WEIGHTS_AS_A_CASES = [
    (2, 56, 56, 64, 64, 3, 3, 1, 1, 1, True, False, 0, 0),
    (5, 56, 56, 256, 64, 1, 1, 1, 0, 0, True, False, 0, 0),       # tiled pixel operand
    (9, 28, 28, 128, 128, 3, 3, 1, 1, 1, False, False, 0, 0),
    (3, 56, 56, 256, 128, 1, 1, 2, 0, 0, True, False, 0, 0),      # stride-2 1x1
    (2, 35, 35, 48, 64, 5, 5, 1, 2, 2, True, False, 0, 32),       # ragged Cin, concat slice
    (2, 35, 35, 64, 96, 3, 3, 1, 1, 1, True, False, 0, 160),      # Cout 96
    (1, 147, 147, 32, 32, 3, 3, 1, 0, 0, True, False, 0, 0),      # Cout 32: three idle epilogue warps
    (3, 17, 17, 128, 128, 1, 7, 1, 0, 3, True, False, 0, 0),
    (2, 17, 17, 160, 80, 7, 1, 1, 3, 0, True, False, 576, 0),
    (1, 5, 5, 64, 32, 1, 1, 1, 0, 0, False, False, 0, 0),
]


@pytest.mark.parametrize("shape", WEIGHTS_AS_A_CASES, ids=lambda s: "x".join(map(str, s[:10])))
def test_conv_tcgen05_weights_as_a(shape):
    import torch
    rng = np.random.default_rng(hash(shape) % 2**32)
    case = _conv_case(rng, *shape)
    out = _run_conv(6, case)
    cout = shape[4]
    ref = case["ref"]
    got = out[..., :cout].float()
    assert (got - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    if shape[13]:
        assert (out[..., cout:].float() == 7.0).all()
    assert torch.equal(out, _run_conv(2, case))  # same bits as the pixels-as-A kernels
