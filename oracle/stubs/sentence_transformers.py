"""Stub of the third-party `sentence_transformers` package (TEST INFRASTRUCTURE ONLY).

The reference imports `SentenceTransformer` at qre:37 and calls `.encode(list[str])`
at qre:126/152/176/181/511/571 and cache.py:412.  The real package and the MiniLM
weights are absent here (no network), so the oracle injects deterministic synthetic
embeddings instead.  A test may install an explicit text -> vector table with
`set_table`; any other text gets a vector seeded by crc32(text).
"""
import zlib

import numpy as np

DIM = 384
_TABLE = {}


def set_table(table, dim=None):
    global DIM
    _TABLE.clear()
    _TABLE.update(table)
    if dim is not None:
        DIM = int(dim)


def embed_text(text, dim=None):
    dim = DIM if dim is None else dim
    v = _TABLE.get(text)
    if v is not None:
        return np.asarray(v, dtype=np.float32)
    rng = np.random.Generator(np.random.PCG64(zlib.crc32(text.encode("utf-8"))))
    return rng.standard_normal(dim).astype(np.float32)


class SentenceTransformer:
    def __init__(self, model_name="all-MiniLM-L6-v2", *a, **k):
        self.model_name = model_name

    def encode(self, texts, *a, **k):
        if isinstance(texts, str):
            return embed_text(texts)
        return np.stack([embed_text(t) for t in texts]).astype(np.float32)
