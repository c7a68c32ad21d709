import importlib
import io
import os
import sys
import tarfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
PKG = "distributed-machine-learning_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pkg(sub=None):
    return importlib.import_module(PKG + ("." + sub if sub else ""))


@pytest.fixture(scope="session")
def jpeg_bytes():
    """{name: bytes} of the 200 bundled test images (tests/golden/testfiles.tar)."""
    out = {}
    with tarfile.open(os.path.join(GOLDEN, "testfiles.tar")) as tf:
        for m in tf.getmembers():
            out[m.name] = tf.extractfile(m).read()
    return out


@pytest.fixture(scope="session")
def frontend_golden():
    return np.load(os.path.join(GOLDEN, "frontend.npz"))


@pytest.fixture(scope="session")
def images(jpeg_bytes, frontend_golden):
    """Decoded uint8 RGB arrays in golden order (1..200)."""
    fe = pkg("frontend")
    return [fe.decode_rgb(jpeg_bytes[str(n)]) for n in frontend_golden["names"]]


@pytest.fixture(scope="session")
def oracle_golden():
    return {m: np.load(os.path.join(GOLDEN, f"{m}_oracle.npz")) for m in ("ResNet50", "InceptionV3")}


@pytest.fixture(scope="session")
def image_dir(tmp_path_factory, jpeg_bytes):
    d = tmp_path_factory.mktemp("testfiles")
    for name, data in jpeg_bytes.items():
        (d / name).write_bytes(data)
    return str(d)
