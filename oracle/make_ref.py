#!/usr/bin/env python
"""Recipe for `oracle/_ref/`: the reference's OWN cache module, compiled where it lies.

TEST / BASELINE INFRASTRUCTURE ONLY.  The reference is Python (`/root/reference/src/cache.py`,
numpy-only, 553 lines); its "build" is byte-compilation.  This script byte-compiles the
unmodified file straight from `/root/reference/src` into `oracle/_ref/cache.bin` (CPython bytecode;
the `.pyc` suffix is avoided because snapshot tools skip `*.pyc`) — a binary, no
source is copied into the repository; `oracle/_ref/` is git-ignored and travels to the GPU box
with the snapshot like the repo's own built `.so`.  `bench.py --impl reference` and the
`cpu_baseline` leg load it with a sourceless loader and time `QueryCache.lookup`
(cache.py:236-326) itself; nothing under `distributed-llm_b200/` ever touches it.

    python oracle/make_ref.py          # in the build container (needs /root/reference)
"""
from __future__ import annotations

import importlib.machinery
import importlib.util
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src"
OUT = os.path.join(HERE, "_ref")
MODULES = ("cache",)


def build() -> list:
    if not os.path.isfile(os.path.join(REF_SRC, "cache.py")):
        raise RuntimeError(f"{REF_SRC} not present: oracle/_ref can only be built in the build container")
    os.makedirs(OUT, exist_ok=True)
    made = []
    for m in MODULES:
        made.append(py_compile.compile(os.path.join(REF_SRC, m + ".py"), cfile=os.path.join(OUT, m + ".bin"),
                                       doraise=True, optimize=0))
    return made


def available() -> bool:
    return os.path.isfile(os.path.join(OUT, "cache.bin"))


def load_cache_module():
    """The reference's `cache` module from the compiled file (same interpreter version needed:
    the GPU box runs this image).  Imported under a private name so it can never shadow the
    drop-in `cache` module of the product."""
    path = os.path.join(OUT, "cache.bin")
    loader = importlib.machinery.SourcelessFileLoader("_qre_reference_cache", path)
    spec = importlib.util.spec_from_loader("_qre_reference_cache", loader)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["_qre_reference_cache"] = mod      # dataclasses look their module up by name
    loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    for p in build():
        print(p)
