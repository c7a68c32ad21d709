"""CPU tests of the C-ABI boundary: the library loads without a GPU, exports every
symbol include/dml_b200.h declares, describes the models, and fails loudly
(no CPU fallback) when asked to compute without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, pkg
from oracle import frontend as ofe, keras_nets


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "dml_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dml_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = pkg("_lib")
    l = lib.lib()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(l, s), f"{s} declared in include/dml_b200.h but not exported"
    assert set(lib.EXPORTS) <= set(syms)


def test_weight_spec_matches_oracle():
    lib = pkg("_lib")
    for m in ("ResNet50", "InceptionV3"):
        assert lib.weight_spec(m) == keras_nets.weight_spec(m)
    assert lib.model_info("ResNet50") == (224, "caffe")
    assert lib.model_info("InceptionV3") == (299, "tf")


def test_preprocess_lut_is_bit_exact():
    lib = pkg("_lib")
    v = np.arange(256, dtype=np.uint8).reshape(-1, 1).repeat(3, 1)
    for m, (_, mode) in ofe.MODEL_INPUT.items():
        ref = ofe.preprocess(v, mode)  # [256, 3]: output channel c for source byte v
        assert np.array_equal(lib.preprocess_lut(m).T.view(np.uint32), ref.view(np.uint32))


def test_errors_are_reported_not_swallowed():
    lib = pkg("_lib")
    l = lib.lib()
    n, t = C.c_int(), C.c_int64()
    assert l.dml_model_num_weights(7, C.byref(n), C.byref(t)) == -1
    assert b"unknown model" in l.dml_last_error()
    with pytest.raises(lib.DmlError):
        lib.model_id("VGG16")
    import torch
    if not torch.cuda.is_available():
        h = C.c_void_p()
        rc = l.dml_create(0, 0, 8, 0, 0, C.byref(h))
        assert rc == -3 and b"no CPU fallback" in l.dml_last_error()
        with pytest.raises(lib.DmlError):
            pkg("engine").Engine("ResNet50")


def test_default_weights_are_deterministic():
    kw = pkg("keras_weights")
    a = kw.to_blob(kw.generate("InceptionV3"))
    b = kw.to_blob(kw.generate("InceptionV3"))
    assert a.size == 23_851_784 and np.array_equal(a, b)
    w = kw.generate("ResNet50", "folding_stress", 1)
    names = [n for n, _ in pkg("_lib").weight_spec("ResNet50")]
    var = w[names.index("conv1_bn/moving_variance")]
    assert var.min() >= 0.5 and var.max() <= 1.5
