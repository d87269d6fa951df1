"""Test configuration.  `-m "not gpu"` runs on the CPU-only dev container; `-m gpu` needs a B200."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def ctx():
    from speech_enhancement_b200 import api
    c = api.Context(0)          # raises loudly if the library or the GPU is missing: no fallback
    yield c
    c.close()


def snr_db(ref, test):
    ref = np.asarray(ref, np.float64)
    test = np.asarray(test, np.float64)
    err = np.sum((ref - test) ** 2)
    sig = np.sum(ref ** 2)
    if err == 0:
        return 300.0
    if sig == 0:
        return -300.0
    return 10.0 * np.log10(sig / err)


def rel_err(ref, test, floor):
    ref = np.asarray(ref, np.float64)
    test = np.asarray(test, np.float64)
    return np.abs(ref - test) / np.maximum(np.maximum(np.abs(ref), np.abs(test)), floor)


def ceps_err(ref, test):
    """Cepstral parity metric: |a-b| / max(|a|, |b|, s) with s = the frame's largest |c1..c12| (for c1..c12)
    or 1e-3 (for c0, logE).  c1..c12 are DCT sums of 23 log-band energies with cancellation; element-wise
    relative error of a coefficient that happens to be near zero says nothing about parity (the
    reference's own float FFT carries the same rounding noise), so the scale is the frame's vector."""
    ref = np.asarray(ref, np.float64).reshape(-1, 14)
    test = np.asarray(test, np.float64).reshape(-1, 14)
    floor = np.full(ref.shape, 1e-3)
    if len(ref):
        floor[:, :12] = np.maximum(np.max(np.abs(ref[:, :12]), axis=1, keepdims=True), 1e-3)
    return np.abs(ref - test) / np.maximum(np.maximum(np.abs(ref), np.abs(test)), floor)
