"""Generates csrc/qre_unicode_tables.inc: the Unicode data the C++ WordPiece needs to tokenise
non-ASCII texts exactly as BERT's BasicTokenizer does (transformers' BertTokenizer, which
SentenceTransformer.encode runs: reference src/query_router_engine.py:126, 181).

    python distributed-llm_b200/qre_b200/gen_unicode_tables.py

The tables are taken from THIS interpreter's `unicodedata` (the module BertTokenizer itself uses), so
the C++ path and the Python tokenizers agree for the Unicode version recorded in the file's header.
Per code point the tokenizer needs:

  pre-class   what `_clean_text` / `_tokenize_chinese_chars` do with the ORIGINAL character:
              dropped (U+0000, U+FFFD, categories C* other than tab / LF / CR), turned into a space
              (tab, LF, CR, category Zs), or kept; CJK ranges are coded in the C++ source.
  mapping     for uncased models the kept character then goes through NFD, loses its Mn marks and is
              lowered: the table holds the resulting sequence for every character where that is not
              the identity (Hangul syllables are decomposed arithmetically in C++, not listed).
  post-class  of every MAPPED character: whitespace (`str.split`), punctuation (ASCII ranges or
              category P*), or part of a word.
  declined    characters whose handling depends on their neighbours — capital sigma (final-sigma
              rule of `str.lower`) and combining marks with a non-zero combining class that are not
              Mn (canonical reordering would move them) — and every code point that the Rust
              `tokenizers` backend of BertTokenizerFast (what sentence-transformers loads) treats
              differently from the Python algorithm above because its Unicode tables are of another
              version: unassigned code points, a few hundred recent marks / format characters, one
              CJK extension range.  Texts containing one go to the caller's tokenizer, so whatever
              the C++ path accepts is tokenised identically by both implementations.
"""
from __future__ import annotations

import os
import sys
import unicodedata as U

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "csrc", "qre_unicode_tables.inc")

SKIP, SPACE = 1, 2
HANGUL_LO, HANGUL_HI = 0xAC00, 0xD7A3


def ranges(pred):
    out, start = [], None
    for cp in range(0x80, 0x110000):
        if pred(cp):
            if start is None:
                start = cp
        elif start is not None:
            out.append((start, cp - 1))
            start = None
    if start is not None:
        out.append((start, 0x10FFFF))
    return out


def pre_class(cp: int) -> int:
    ch = chr(cp)
    if cp == 0 or cp == 0xFFFD:
        return SKIP
    cat = U.category(ch)
    if ch in "\t\n\r" or cat == "Zs":
        return SPACE
    if cat.startswith("C"):
        return SKIP
    return 0


def mapped(cp: int) -> str:
    return "".join(c for c in U.normalize("NFD", chr(cp)) if U.category(c) != "Mn").lower()


CJK = ((0x4E00, 0x9FFF), (0x3400, 0x4DBF), (0x20000, 0x2A6DF), (0x2A700, 0x2B73F), (0x2B740, 0x2B81F),
       (0x2B820, 0x2CEAF), (0xF900, 0xFAFF), (0x2F800, 0x2FA1F))


def python_view(cp: int, lower: bool):
    """(normalised text, class of the result inside a word) as the Python algorithm sees one character."""
    pc = pre_class(cp)
    if pc == SKIP:
        return "", "w"
    if pc == SPACE:
        return " ", "w"
    m = mapped(cp) if lower else chr(cp)
    if any(lo <= cp <= hi for lo, hi in CJK):
        m = " " + m + " "
    return m, None


def rust_disagreements():
    """Code points that tokenizers' BertNormalizer / BertPreTokenizer handle differently from the Python
    algorithm (empty set, with a note, when the package is not installed)."""
    try:
        import tokenizers
        from tokenizers.normalizers import BertNormalizer
        from tokenizers.pre_tokenizers import BertPreTokenizer
    except ImportError:
        return set(), "tokenizers not installed: no cross-check"
    norm = {lw: BertNormalizer(clean_text=True, handle_chinese_chars=True, strip_accents=None, lowercase=lw)
            for lw in (True, False)}
    pre = BertPreTokenizer()
    bad = set()
    for cp in range(0x80, 0x110000):
        if 0xD800 <= cp <= 0xDFFF:
            continue
        for lw in (True, False):
            want = python_view(cp, lw)[0]
            got = norm[lw].normalize_str(chr(cp))
            if got != want and not (got.strip() == "" and want.strip() == "" and (got == "") == (want == "")):
                # (U+2028 / U+2029: ' ' from one, the character itself from the other; both split there)
                if not (cp in (0x2028, 0x2029)):
                    bad.add(cp)
        if cp in bad or pre_class(cp) != 0:
            continue
        # inside a word: separator (whitespace), token of its own (punctuation), or word character
        for c in (mapped(cp) or ""):
            o = ord(c)
            if o < 0x80:
                continue
            py = "s" if c.isspace() else ("p" if U.category(c).startswith("P") else "w")
            parts = [t for t, _ in pre.pre_tokenize_str("a" + c + "b")]
            ru = "s" if parts == ["a", "b"] else ("p" if parts == ["a", c, "b"] else ("w" if parts == ["a" + c + "b"] else "?"))
            if py != ru:
                bad.add(cp)
    return bad, f"cross-checked against tokenizers {tokenizers.__version__}"


def main() -> None:
    rust_bad, rust_note = rust_disagreements()
    skip = ranges(lambda cp: pre_class(cp) == SKIP and cp not in rust_bad)
    space = ranges(lambda cp: pre_class(cp) == SPACE)
    punct = ranges(lambda cp: U.category(chr(cp)).startswith("P"))
    white = ranges(lambda cp: chr(cp).isspace())
    decline = ranges(lambda cp: cp in rust_bad or (pre_class(cp) == 0 and (
        cp == 0x03A3 or (U.combining(chr(cp)) != 0 and U.category(chr(cp)) != "Mn"))))
    # a mapped sequence must not itself contain a character that would be declined or re-mapped
    entries, pool = [], []
    for cp in range(0x80, 0x110000):
        if pre_class(cp) != 0 or HANGUL_LO <= cp <= HANGUL_HI or cp in rust_bad:
            continue
        m = mapped(cp)
        if m == chr(cp):
            continue
        for c in m:
            assert mapped(ord(c)) == c or ord(c) < 0x80, (hex(cp), m)
        entries.append((cp, len(pool), len(m)))
        pool.extend(ord(c) for c in m)
    # Hangul: the arithmetic decomposition in C++ must equal unicodedata's, and Jamo map to themselves
    for cp in (HANGUL_LO, HANGUL_LO + 1, 0xB098, HANGUL_HI):
        s = cp - HANGUL_LO
        want = [0x1100 + s // 588, 0x1161 + (s % 588) // 28] + ([0x11A7 + s % 28] if s % 28 else [])
        assert [ord(c) for c in mapped(cp)] == want, hex(cp)
    lines = [
        "// GENERATED by qre_b200/gen_unicode_tables.py - do not edit.",
        f"// unicodedata.unidata_version = {U.unidata_version} (Python {sys.version_info.major}.{sys.version_info.minor}); {rust_note}:",
        f"// {len(rust_bad)} code points on which the two disagree are in kUniDecline.",
        f"static const char kUnicodeVersion[] = \"{U.unidata_version}\";",
    ]

    def emit_ranges(name, rs):
        lines.append(f"static const uint32_t {name}[][2] = {{")
        for i in range(0, len(rs), 6):
            lines.append("  " + " ".join(f"{{0x{a:X}, 0x{b:X}}}," for a, b in rs[i:i + 6]))
        lines.append("};")
        lines.append(f"static const int {name}N = {len(rs)};")

    lines.append("struct UniMap { uint32_t cp; uint32_t off; uint32_t len; };")
    emit_ranges("kUniSkip", skip)
    emit_ranges("kUniSpace", space)
    emit_ranges("kUniPunct", punct)
    emit_ranges("kUniWhite", white)
    emit_ranges("kUniDecline", decline)
    # ---- for the heuristic router's regular expressions (qre_host_features.cu): Python `re` classes on str
    # patterns (\w = isalnum or '_', \d = isdecimal, \s = isspace = kUniWhite) and the full str.lower()
    # mapping (context-free part; capital sigma is declined there as well)
    emit_ranges("kUniAlnum", ranges(lambda cp: chr(cp).isalnum()))
    emit_ranges("kUniDecimal", ranges(lambda cp: chr(cp).isdecimal()))
    # final-sigma rule of str.lower(): capital sigma becomes the final form iff a CASED character precedes it
    # and none follows, case-IGNORABLE characters skipped on both sides.  CPython does not expose the two
    # properties; they are read off lower() itself.
    emit_ranges("kUniCased", ranges(lambda cp: not 0xD800 <= cp <= 0xDFFF and (chr(cp) + "\u03a3").lower()[-1] == "\u03c2"))
    emit_ranges("kUniCaseIgnorable", ranges(lambda cp: not 0xD800 <= cp <= 0xDFFF
                                            and (chr(cp) + "\u03a3").lower()[-1] != "\u03c2"
                                            and ("a" + chr(cp) + "\u03a3").lower()[-1] == "\u03c2"))
    low_entries, low_pool = [], []
    for cp in range(0x80, 0x110000):
        if 0xD800 <= cp <= 0xDFFF:
            continue
        m = chr(cp).lower()
        if m != chr(cp):
            low_entries.append((cp, len(low_pool), len(m)))
            low_pool.extend(ord(c) for c in m)
    lines.append("static const UniMap kUniLower[] = {")
    for i in range(0, len(low_entries), 5):
        lines.append("  " + " ".join(f"{{0x{c:X}, {o}, {n}}}," for c, o, n in low_entries[i:i + 5]))
    lines.append("};")
    lines.append(f"static const int kUniLowerN = {len(low_entries)};")
    lines.append("static const uint32_t kUniLowerPool[] = {")
    for i in range(0, len(low_pool), 12):
        lines.append("  " + " ".join(f"0x{c:X}," for c in low_pool[i:i + 12]))
    lines.append("  0};")
    lines.append("static const UniMap kUniMap[] = {")
    for i in range(0, len(entries), 5):
        lines.append("  " + " ".join(f"{{0x{c:X}, {o}, {n}}}," for c, o, n in entries[i:i + 5]))
    lines.append("};")
    lines.append(f"static const int kUniMapN = {len(entries)};")
    lines.append("static const uint32_t kUniPool[] = {")
    for i in range(0, len(pool), 12):
        lines.append("  " + " ".join(f"0x{c:X}," for c in pool[i:i + 12]))
    lines.append("  0};")
    with open(OUT, "w", encoding="ascii") as f:
        f.write("\n".join(lines) + "\n")
    print(f"{OUT}: skip {len(skip)} space {len(space)} punct {len(punct)} white {len(white)} decline {len(decline)} "
          f"ranges, {len(entries)} mappings, {len(pool)} pooled code points")


if __name__ == "__main__":
    main()
