"""The C++ host WordPiece (qre_wordpiece_encode) against the oracle restatement, which is itself pinned
against transformers' BertTokenizer (tests/test_minilm_oracle_cpu.py).  No GPU involved."""
import os
import random

import numpy as np

from oracle import minilm_oracle as MO
from qre_b200 import _cabi
from tests.wordpiece_cases import TEXTS, VOCAB


def _vocab_file(tmp_path):
    vp = os.path.join(tmp_path, "vocab.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(VOCAB) + "\n")
    return vp, {t: i for i, t in enumerate(VOCAB)}


def _split(ids, lens):
    out, o = [], 0
    for n in lens:
        if n < 0:
            out.append(None)
        else:
            out.append(ids[o:o + n].tolist())
            o += n
    assert o == ids.size
    return out


def test_fixed_texts(tmp_path):
    vp, vocab = _vocab_file(tmp_path)
    wp = _cabi.WordPiece(vp)
    for max_len in (256, 16, 3, 2):
        ids, lens = wp.encode(TEXTS, max_len=max_len)
        got = _split(ids, lens)
        for text, g in zip(TEXTS, got):          # non-ASCII texts ("Café ÉE 中文abc") included
            assert g == MO.tokenize_ids(text, vocab, max_len).tolist(), (text, max_len)


def test_random_ascii_texts_threaded(tmp_path):
    vp, vocab = _vocab_file(tmp_path)
    wp = _cabi.WordPiece(vp)
    rnd = random.Random(7)
    words = [v for v in VOCAB if v.isascii() and not v.startswith("[") and not v.startswith("##")]
    glue = [" ", "  ", "\t", "", "-", "'", ", ", ". ", "\n", "\x01", "_"]
    texts = []
    for _ in range(3000):
        n = rnd.randint(0, 30)
        s = "".join(rnd.choice(words).upper() if rnd.random() < 0.2 else rnd.choice(words) + rnd.choice(glue)
                    for _ in range(n))
        if rnd.random() < 0.1:
            s += rnd.choice(["zzz", "q" * 150, "helloworlds", "abcabc", "123", "x" * 99])
        texts.append(s)
    for threads in (1, 4):
        ids, lens = wp.encode(texts, max_len=24, n_threads=threads)
        got = _split(ids, lens)
        for text, g in zip(texts, got):
            assert g == MO.tokenize_ids(text, vocab, 24).tolist(), repr(text)
    assert (lens >= 2).all()


def test_errors(tmp_path):
    import ctypes
    import pytest
    lib = _cabi.load()
    with pytest.raises(_cabi.QreError, match="cannot open"):
        _cabi.WordPiece(os.path.join(tmp_path, "missing.txt"))
    bad = os.path.join(tmp_path, "bad.txt")
    open(bad, "w").write("a\nb\n")
    with pytest.raises(_cabi.QreError, match="lacks"):
        _cabi.WordPiece(bad)
    vp, _ = _vocab_file(tmp_path)
    wp = _cabi.WordPiece(vp)
    texts = (ctypes.c_char_p * 1)(b"hello world hello world")
    lens = np.zeros(1, np.int32)
    ids = np.zeros(2, np.int32)
    rc = lib.qre_wordpiece_encode(wp._h, ctypes.c_int64(1), texts, None, 256, ids.ctypes.data_as(ctypes.c_void_p),
                                  ctypes.c_int64(2), lens.ctypes.data_as(ctypes.c_void_p), 1)
    assert rc == 3 and b"capacity" in lib.qre_last_error()


def test_duplicate_vocab_lines_take_the_last_id(tmp_path):
    """A piece listed twice gets the id of its last line — what loading vocab.txt into a dict does."""
    vp = os.path.join(tmp_path, "dup.txt")
    pieces = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "go", "##ing", "x", "##ing", "go"]
    with open(vp, "w") as f:
        f.write("\n".join(pieces) + "\n")
    vocab = {t: i for i, t in enumerate(pieces)}
    wp = _cabi.WordPiece(vp)
    ids, lens = wp.encode(["going x"], max_len=16)
    assert ids.tolist() == MO.tokenize_ids("going x", vocab, 16).tolist() == [2, 8, 7, 6, 3]


def test_product_unicode_fallback_equals_oracle():
    """qre_b200.embedder.wordpiece_unicode (the Python path for texts the C++ tokenizer declines) against the
    restatement, on generated texts with accents, CJK, emoji, control characters and odd whitespace."""
    from qre_b200.embedder import wordpiece_unicode
    vocab = {t: i for i, t in enumerate(VOCAB)}
    rnd = random.Random(5)
    words = [v for v in VOCAB if not v.startswith("[") and not v.startswith("##")]
    extra = ["É", "ü", "ñ", "ß", "中", "文", "日本", "😀", " ", " ", "​", "\x07", "…", "—", "İ", "ﬁ", "½"]
    glue = [" ", "\t", "\n", "", "-", "'", ",", ".", "!", "(", ")", "$", "_", "#"]
    for _ in range(1500):
        text = "".join(rnd.choice(words + extra) + rnd.choice(glue) for _ in range(rnd.randint(0, 12)))
        for max_len in (256, 7):
            assert wordpiece_unicode(text, vocab, max_len) == MO.tokenize_ids(text, vocab, max_len).tolist(), repr(text)


def test_packed_entry_equals_pointer_entry(tmp_path):
    """qre_wordpiece_encode_packed (one blob + offsets) against qre_wordpiece_encode (one pointer per
    text) through ctypes: same ids, same lengths, non-ASCII texts declined by both, empty texts and an
    empty batch included."""
    import ctypes as C
    from qre_b200 import _cabi
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + list("abcdefghijklmnopqrstuvwxyz0123456789!?,.'") + \
        ["##" + c for c in "abcdefghijklmnopqrstuvwxyz"] + ["hello", "world", "token", "##ing", "##s", "cache", "route"]
    vp = tmp_path / "vocab.txt"
    vp.write_text("\n".join(vocab) + "\n")
    wp = _cabi.WordPiece(str(vp))
    rs = np.random.Generator(np.random.PCG64(5))
    base = ["hello", "world!", "tokens", "routing", "cache,", "café", "ΣΑΣ", "", "  ", "Hello  WORLD", "x" * 300]
    texts = [" ".join(base[int(j)] for j in rs.integers(0, len(base), int(rs.integers(0, 7)))) for _ in range(700)]
    ids_p, lens_p = wp.encode(texts, max_len=64, n_threads=3)            # the packed entry
    raw = [t.encode("utf-8") for t in texts]
    arr = (C.c_char_p * len(raw))(*raw)
    tl = np.array([len(b) for b in raw], np.int64)
    cap = int(np.minimum(tl + 2, 64).sum())
    ids = np.empty(max(cap, 1), np.int32)
    lens = np.empty(len(raw), np.int32)
    rc = wp._lib.qre_wordpiece_encode(wp._h, C.c_int64(len(raw)), arr, _cabi._ptr(tl), C.c_int(64), _cabi._ptr(ids),
                                      C.c_int64(cap), _cabi._ptr(lens), C.c_int(2))
    assert rc == 0
    assert np.array_equal(lens, lens_p) and np.array_equal(ids[: int(lens[lens > 0].sum())], ids_p)
    assert (lens == -1).sum() > 0 and (lens == 2).sum() > 0               # declined texts; empty texts = [CLS][SEP]
    e_ids, e_lens = wp.encode([], max_len=64)
    assert e_ids.size == 0 and e_lens.size == 0


UNI_VOCAB = VOCAB + ["é", "##é", "日", "本", "ᄒ", "##ᅡ", "##ᆫ", "한", "ﬁ", "##ﬁ", "…", "—", "ss", "##ss", "i", "##i",
                     "й", "##й", "и", "##и", "α", "##α", "σ", "##σ", "ς", "😀", "½", "ß", "##ß", "ü", "u", "##u", "n", "##n"]


def unicode_texts(seed=11, n_mixed=2500, n_random=1500):
    """Texts for the general tokenizer path: words glued with accents, CJK incl. compatibility
    ideographs, Hangul, emoji, zero-width / control characters, odd spaces, ligatures, over-long words;
    then texts built from arbitrary ASSIGNED code points (three plane ranges)."""
    import unicodedata
    rnd = random.Random(seed)
    words = [v for v in UNI_VOCAB if not v.startswith("[") and not v.startswith("##")]
    extra = ["É", "ü", "Ü", "ñ", "ß", "中", "文", "日本", "한국", "한", "豈", "😀", "\u00a0", "\u2003", "\u200b", "\u2028", "\x07",
             "\x85", "\ufeff", "…", "—", "«", "»", "¿", "İ", "ﬁ", "½", "e\u0301", "Й", "й", "ά", "Ω", "ς", "\ufffd",
             "\U0001f600", "\u0e01\u0e34", "x" * 101, "é" * 101, "é" * 100, "\U00020000", "\u3400", "\u202f", "\u1680"]
    glue = [" ", "\t", "\n", "", "-", "'", ",", ".", "!", "(", ")", "$", "_", "#", "\r\n", "\x0b", "\x1c"]
    texts = ["".join(rnd.choice(words + extra) + rnd.choice(glue) for _ in range(rnd.randint(0, 12)))
             for _ in range(n_mixed)]
    pools = [[c for c in range(lo, hi) if unicodedata.category(chr(c)) not in ("Cn", "Cs", "Co")]
             for lo, hi in ((0x80, 0x3000), (0x3000, 0x10000), (0x10000, 0x30000))]
    for _ in range(n_random):
        cps = [rnd.choice(rnd.choice(pools)) for _ in range(rnd.randint(1, 6))]
        texts.append(" hello".join(chr(c) for c in cps) + rnd.choice(glue) + "World")
    return texts, n_mixed


def test_unicode_texts_equal_the_restatement(tmp_path):
    """The C++ general path (UTF-8 decode, cleaning, CJK spacing, NFD + strip accents + lower from the
    generated tables, Unicode whitespace / punctuation, pieces on character boundaries) against the
    restatement, both casings.  A text is handed back (-1) only when it holds a character the tables
    list as neighbour-dependent or as handled differently by the two reference tokenizers; that must
    stay rare on ordinary text.  (tests/test_minilm_oracle_cpu.py runs the same texts against
    transformers' BertTokenizer itself.)"""
    vp = os.path.join(tmp_path, "vocab_u.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(UNI_VOCAB) + "\n")
    vocab = {t: i for i, t in enumerate(UNI_VOCAB)}
    texts, n_mixed = unicode_texts()
    for lower in (True, False):
        wp = _cabi.WordPiece(vp, do_lower_case=lower)
        for max_len, threads in ((256, 1), (7, 3)):
            ids, lens = wp.encode(texts, max_len=max_len, n_threads=threads)
            got = _split(ids, lens)
            for text, g in zip(texts, got):
                if g is not None:
                    assert g == MO.tokenize_ids(text, vocab, max_len, lower).tolist(), (repr(text), lower, max_len)
            declined_mixed = sum(g is None for g in got[:n_mixed])
            declined_random = sum(g is None for g in got[n_mixed:])
            assert declined_mixed <= 0.01 * n_mixed, declined_mixed            # only the texts with a capital sigma etc.
            assert declined_random <= 0.10 * (len(texts) - n_mixed), declined_random
    # capital sigma (final-sigma rule of str.lower), unassigned code points and malformed UTF-8 are handed
    # back, not guessed
    wp = _cabi.WordPiece(vp)
    _, lens = wp.encode(["ΟΔΥΣΣΕΥΣ", "fine é", "unassigned \u0378 here"], max_len=16)
    assert lens.tolist()[0] == -1 and lens.tolist()[1] > 2 and lens.tolist()[2] == -1
