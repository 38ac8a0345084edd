"""Build libqre_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

Freshness is decided by CONTENT, not by file times: every object file carries a stamp with the
SHA-256 of its source, all headers and the compiler flags, and the library carries the hash of its
objects' stamps.  A checkout with reset mtimes (or a prebuilt library that travelled with an older
tree) is therefore rebuilt exactly when a source actually differs.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(PKG, "csrc")
LIB_DIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIB_DIR, "libqre_b200.so")
SOURCES = ["qre_api.cu", "qre_kernels.cu", "qre_gemm_topk.cu", "qre_gemm_topk2.cu", "qre_noalloc.cu",
           "qre_host_features.cu", "qre_enc_gemm.cu", "qre_encoder.cu", "qre_wordpiece.cu"]
HEADERS = ["qre_common.cuh", "qre_ptx.cuh", "qre_scan_epilogue.cuh", "qre_encoder.cuh", "qre_unicode_tables.inc",
           os.path.join("..", "..", "include", "qre_b200.h"), os.path.join("..", "..", "include", "qre_encoder_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
]


def _sha(*parts: bytes) -> str:
    h = hashlib.sha256()
    for p in parts:
        h.update(p)
        h.update(b"\0")
    return h.hexdigest()


def _read(path: str) -> bytes:
    with open(path, "rb") as f:
        return f.read()


def _headers_digest() -> bytes:
    return _sha(*[_read(os.path.join(CSRC, h)) for h in HEADERS]).encode()


def _object_digest(src: str, hdr: bytes) -> str:
    return _sha(_read(os.path.join(CSRC, src)), hdr, " ".join(NVCC_FLAGS).encode())


def _stamp(path: str) -> str:
    try:
        with open(path + ".sha256") as f:
            return f.read().strip()
    except OSError:
        return ""


def _write_stamp(path: str, digest: str) -> None:
    with open(path + ".sha256", "w") as f:
        f.write(digest + "\n")


def source_digest() -> str:
    """Digest of everything the library is built from (what `lib/libqre_b200.so.sha256` must hold)."""
    hdr = _headers_digest()
    return _sha(*[_object_digest(s, hdr).encode() for s in SOURCES])


def _stale() -> bool:
    return not os.path.isfile(LIB) or _stamp(LIB) != source_digest()


def build_lib(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(LIB_DIR, exist_ok=True)
    hdr = _headers_digest()

    def compile_one(src: str) -> str:
        obj = os.path.join(LIB_DIR, src.replace(".cu", ".o"))
        digest = _object_digest(src, hdr)
        if not force and os.path.isfile(obj) and _stamp(obj) == digest:
            return obj
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True)
        _write_stamp(obj, digest)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    subprocess.run([nvcc, "-shared", "-o", LIB, *objs, "-cudart", "static"], check=True)
    _write_stamp(LIB, source_digest())
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
