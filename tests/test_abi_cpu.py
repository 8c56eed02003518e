"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every
symbol include/taco_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

from tacotron_prosody_multi_gpu_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    return build.build()


def test_library_exports_every_declared_symbol(lib_path):
    hdr = open(os.path.join(ROOT, "include", "taco_b200.h")).read()
    declared = set(re.findall(r"\b(taco_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    L = ctypes.CDLL(lib_path)
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in ctypes.cast(L.taco_version, ctypes.CFUNCTYPE(ctypes.c_char_p))()


def test_hparams_struct_matches_header():
    hdr = open(os.path.join(ROOT, "include", "taco_b200.h")).read()
    body = hdr[hdr.index("typedef struct TacoHParams {"):hdr.index("} TacoHParams;")]
    fields = re.findall(r"^\s*(?:int32_t|float)\s+([a-z_0-9]+)(?:\[\d+\])?;", body, flags=re.M)
    assert fields == [f[0] for f in _lib.TacoHParams._fields_]


def test_engine_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    with pytest.raises(_lib.TacoError):
        TacoEngine(default_hparams())


def test_library_exports_every_training_primitive(lib_path):
    from tacotron_prosody_multi_gpu_b200.train import ops
    hdr = open(os.path.join(ROOT, "include", "taco_b200_train.h")).read()
    declared = set(re.findall(r"\b(taco_op_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(ops.TRAIN_SYMBOLS)
    L = ctypes.CDLL(lib_path)
    for s in declared:
        assert hasattr(L, s), s


def test_training_ops_fail_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from tacotron_prosody_multi_gpu_b200.train.ops import CudaOps
    with pytest.raises(_lib.TacoError):
        CudaOps(0)
