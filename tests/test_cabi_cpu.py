"""CPU-only checks of the drop-in boundary: the shared library loads without a GPU driver and
exports every symbol include/*.h declares; with no device the product fails loudly."""
import ctypes
import os
import re

import numpy as np
import pytest

from qre_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    inc = os.path.join(ROOT, "include")
    text = "".join(open(os.path.join(inc, f)).read() for f in sorted(os.listdir(inc)) if f.endswith(".h"))
    return sorted(set(re.findall(r"QRE_API\s+(?:const\s+char\s*\*|int)\s+(qre_\w+)\s*\(", text)))


def test_header_symbols_all_exported():
    lib = _cabi.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_cabi.EXPORTS) == names
    assert lib.qre_abi_version() == 2


def test_struct_layout_matches_header():
    assert _cabi.CAND_DTYPE.itemsize == 32
    assert ctypes.sizeof(_cabi.Params) == 8 * 3 + 4 * 4 + 8 * 6 + 4 * 4
    assert ctypes.sizeof(_cabi.Result) == 8 * len(_cabi.RESULT_FIELDS)


def test_encoder_structs_and_arguments():
    assert ctypes.sizeof(_cabi.EncoderConfig) == 40
    assert ctypes.sizeof(_cabi.EncoderLayer) == 16 * 8
    lib = _cabi.load()
    h = ctypes.c_void_p()
    bad = _cabi.EncoderConfig(hidden=100)
    assert lib.qre_encoder_create(0, ctypes.byref(bad), ctypes.c_int64(128), 4, ctypes.byref(h)) == 1
    assert b"multiple of 64" in lib.qre_last_error()
    bad = _cabi.EncoderConfig(heads=5)
    assert lib.qre_encoder_create(0, ctypes.byref(bad), ctypes.c_int64(128), 4, ctypes.byref(h)) == 1
    assert lib.qre_encoder_destroy(None) == 0
    assert lib.qre_encoder_debug_gemm(0, 4, 30, 64, None, None, None, None, 0, None, None, ctypes.c_float(0), None) == 1


def test_encoder_without_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_cabi.QreError, match="no CUDA device|cudaGetDeviceCount"):
        _cabi.Encoder(_cabi.EncoderConfig(), {}, max_tokens=128, max_seqs=4)


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_cabi.QreError, match="no CUDA device|CPU fallback|cudaGetDeviceCount"):
        _cabi.Shard(dim=384, capacity=1024, max_batch=8)


def test_bad_arguments_rejected_before_any_cuda_call():
    lib = _cabi.load()
    h = ctypes.c_void_p()
    assert lib.qre_cache_create(0, 100, ctypes.c_int64(10), 4, ctypes.byref(h)) == 1   # dim % 8
    assert b"multiple of 8" in lib.qre_last_error()
    assert lib.qre_cache_create(0, 384, ctypes.c_int64(0), 4, ctypes.byref(h)) == 1
    assert lib.qre_cache_destroy(None) == 0


def test_every_encoder_kernel_joins_the_dependent_launch_chain():
    """Source check: each __global__ kernel of the embedder is launched with programmatic stream
    serialization (qre_encoder.cuh), which is only safe if EVERY kernel of the chain executes
    griddepcontrol.wait before touching global memory.  A kernel added without pdl_wait() would break the
    transitive completion guarantee silently, so fail here instead."""
    csrc = os.path.join(ROOT, "distributed-llm_b200", "csrc")
    seen = 0
    for fn in ("qre_encoder.cu", "qre_enc_gemm.cu"):
        text = open(os.path.join(csrc, fn)).read()
        for m in re.finditer(r"__global__[^{;]*?\b(\w+)\s*\([^;{]*?\)\s*\{", text, re.S):
            body = text[m.end(): m.end() + 6000]
            first_wait = body.find("pdl_wait();")
            assert first_wait >= 0, (fn, m.group(1))
            # nothing that dereferences a kernel argument may precede the wait
            head = body[:first_wait]
            assert "offsets[" not in head and "__ldg" not in head and "tma_load" not in head, (fn, m.group(1))
            seen += 1
    assert seen >= 9, seen
    # and no kernel of the chain is launched with the <<<>>> syntax (which cannot carry the attribute)
    for fn in ("qre_encoder.cu", "qre_enc_gemm.cu"):
        assert "<<<" not in open(os.path.join(csrc, fn)).read(), fn
