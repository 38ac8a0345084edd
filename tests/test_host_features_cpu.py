"""CPU: qre_host_features (C++ threads) against HeuristicRouter.classify (Python `re`) on a large
generated corpus that stresses every pattern, word-boundary quirk and rule order — ASCII and, for
the general path, Unicode text (accented neighbours of the keywords, non-ASCII spaces and digits,
characters whose lower() changes length); and, when the reference tree is mounted, against the
reference's own HeuristicRouter."""
import numpy as np
import pytest

from oracle import ref_harness
from qre_b200 import _cabi
from tests import replay_common as R
from tests.fake_engine import OracleEngine

VOCAB = """write implement code program script build refactor debug fix traceback exception error segfault timeout
hanging api flask fastapi docker kubernetes ssh tunnel nginx system design architecture distributed scalability
load balanc balancing prove lemma theorem corollary derivative integral gradient time space complexity big-o big o
dynamic programming dp graph dijkstra bfs dfs a* a*b compare contrast difference between pros and cons vs vs. versus
evaluate assess critique analyze essay report proposal research paper literature review methodology comprehensive
in-depth step-by-step step by step detailed summarize synthesis everything all so far entire transcript debate
dialogue format as json markdown table etl pipeline spark hadoop presto sql csv excel dataframe dataset deduplicate
normalize clean transform parse extract symptom diagnosis treatment therapy prognosis chronic severe pain migraine
dizziness fatigue nausea inflammation anxiety depression dietary meal training exercise recovery plan schedule
regimen mental health psycholog psychology counseling therapist physician hospital using the context history above
based on our conversation continue expand build on follow up previous earlier hi hello hey yo sup good morning
afternoon evening thanks thank you what is who is where is when did how many capital of tell me a joke fun fact
random how to how do i can you tell me benefit benefits of tip tips for advice on daily intake how often how much
healthy habit routine lifestyle define meaning of definition what does mean 2+2 12 * 3 ? what is 7/8? the of a an
cat dog blue apple_pie under_score x9 versus. fixes prefix codex apis hi5 thanksgiving report's""".split()
FILLERS = ["the", "of", "quick", "brown", "fox", "zzz", "lorem", "ipsum", "42", "ok", "maybe", "really"]
PUNCT = [" ", " ", " ", ", ", ". ", "? ", "! ", "\n", "\t", " - ", "; ", " {", "} ", "==", "->", "::", "```", "."]
MARKERS = ["def ", "class ", "import ", "Traceback", "Exception", "SELECT ", "FROM ", "JOIN ", "WITH ", "!="]


UNI_WORDCH = ["é", "ß", "ñ", "Ü", "İ", "ı", "中", "文", "й", "Ж", "α", "٣", "K", "ǅ", "ﬁ", "²", "한", "e\u0301"]
UNI_PUNCT = ["…", "—", "«", "»", "¿", "·", "。", "😀", "\u0301", "€"]
UNI_SPACE = ["\u00a0", "\u2003", "\u3000", "\u2028", "\x85", "\u1680"]
UNI_MISC = ["İstanbul", "straße", "naïve", "DÉBUG", "fıx", "Kelvin", "日本語", "привет", "API’s", "ＳＱＬ", "\u200bhi\u200b", "ſql", "ſpark", "eſſay", "hı", "Kubernetes", "dıjkstra"]
UNI_DIGITS = ["٣", "12", "７", "४२", "9", "²", "½"]


def _corpus(n, seed, unicode_mix=False):
    rs = np.random.Generator(np.random.PCG64(seed))
    out = []
    for i in range(n):
        k = int(rs.choice([1, 2, 3, 5, 8, 14, 18, 30, 120]))
        parts = []
        plain = rs.random() < 0.2              # fillers only: reaches the length / word-count / fallback rules
        for _ in range(k):
            u = 0.5 + 0.4 * rs.random() if plain else rs.random()
            w = str(rs.choice(VOCAB)) if u < 0.45 else (str(rs.choice(FILLERS)) if u < 0.9 else str(rs.choice(MARKERS)))
            if rs.random() < 0.15:
                w = w.upper() if rs.random() < 0.5 else w.capitalize()
            parts.append(w)
            parts.append(" " if plain else str(rs.choice(PUNCT)))
        q = "".join(parts)
        if rs.random() < 0.2:
            q = " \t" + q + "  \n"
        if rs.random() < 0.03:
            q = f"{rs.integers(0, 99)} {rs.choice(list('+-*/'))} {rs.integers(0, 99)}{rs.choice(['', '?', ' ?'])}"
        if rs.random() < 0.02:
            q = "what is " + f"{rs.integers(0, 99)}{rs.choice(list('+-*/'))}{rs.integers(0, 99)}?"
        if rs.random() < 0.02:
            q = ""
        if rs.random() < 0.03:
            q += " café naïve"            # non-ASCII: the general path
        if unicode_mix and rs.random() < 0.5:
            # glue Unicode word characters, spaces, digits and punctuation onto / between the words
            bits = q.split(" ")
            for j in range(len(bits)):
                u = rs.random()
                if u < 0.15:
                    bits[j] = str(rs.choice(UNI_WORDCH)) + bits[j]            # \b must NOT see a boundary here
                elif u < 0.30:
                    bits[j] = bits[j] + str(rs.choice(UNI_WORDCH))
                elif u < 0.40:
                    bits[j] = bits[j] + str(rs.choice(UNI_PUNCT))             # still a boundary
                elif u < 0.45:
                    bits[j] = str(rs.choice(UNI_MISC))
            q = str(rs.choice(UNI_SPACE + [" "] * 6)).join(bits)
            if rs.random() < 0.1:
                q = str(rs.choice(UNI_SPACE)) + q + str(rs.choice(UNI_SPACE))
        if unicode_mix and rs.random() < 0.03:
            q = f"{rs.choice(UNI_DIGITS)}{rs.choice(['', ' ', chr(0x3000)])}{rs.choice(list('+-*/'))} {rs.choice(UNI_DIGITS)}{rs.choice(['', '?'])}"
        if unicode_mix and rs.random() < 0.05:
            # capital sigma: str.lower() picks the final form by looking at the neighbours
            q += str(rs.choice([" ΣΟΦΙΑ", " ΟΔΥΣΣΕΥΣ", "Σ", " ΑΣ.", " ΑΣ'Α", " Σ Σ", "aΣ", " ΑΣ\u0301", "ΑΣ:", " ΑΣ)Σ", "ΑΣ "]))
        ctx_len = int(rs.choice([0, 0, 5, 799, 800, 3199, 3200, 9000]))
        out.append((q, None if ctx_len == 0 else "c" * ctx_len))
    return out


CFGS = [dict(), dict(heuristic_long_chars=800, heuristic_multi_qmarks=2, heuristic_code_markers_needed=2,
                     heuristic_context_chars=3200, token_threshold=1000),
        dict(heuristic_long_chars=40, heuristic_multi_qmarks=1, heuristic_code_markers_needed=1,
             heuristic_context_chars=5, token_threshold=3)]


@pytest.mark.parametrize("ci", range(len(CFGS)))
def test_cpp_features_equal_python_classifier(ci):
    _, qre_mod, _ = R.load_modules(None, OracleEngine)
    hr = qre_mod.HeuristicRouter(CFGS[ci])
    corpus = _corpus(6000, 10 + ci)
    qs, cs = [q for q, _ in corpus], [c for _, c in corpus]
    feats, toks = hr.classify_batch(qs, cs)
    prefix = {"complex pattern": 0, "long query chars": 1, "multi-question count": 2,
              "code/debug markers detected": 3, "large context chars": 4, "simple pattern": 5,
              "short everyday query": 6, "no heuristic match": 7}
    exp_rule = []
    for (q, c), f, t in zip(corpus, feats, toks):
        exp = hr.classify(q, c)
        assert f == exp, (q, c and len(c), f, exp)
        assert t == hr._token_fallback.count(q, c)
        exp_rule.append(prefix[exp[3].split("=")[0].split(" ->")[0]])
    hp = _cabi.HeurParams(hr.long_text_threshold, hr.multi_question_threshold, hr.code_markers_needed,
                          hr.context_chars_threshold, hr._token_fallback.threshold, 2)
    r = _cabi.host_features(qs, [len(c) if c else 0 for c in cs], hp, n_threads=4)
    assert np.array_equal(r["rule"], np.array(exp_rule, np.int32))
    assert np.array_equal(r["heur_dev"], np.array([f[0] for f in feats], np.uint8))
    assert np.array_equal(r["heur_conf"], np.array([f[1] for f in feats], np.float64))
    assert np.array_equal(r["tok"], np.array(toks, np.int32))
    if ci == 0:      # with the default thresholds the corpus reaches every rule
        assert set(exp_rule) == {0, 1, 2, 3, 4, 5, 6, 7}


@pytest.mark.parametrize("ci", range(len(CFGS)))
def test_cpp_features_on_unicode_text(ci):
    """The general path: every keyword with Unicode word characters, punctuation and spaces around it,
    non-ASCII digits in the tiny-math rule, characters whose lower() changes length (İ), Unicode
    whitespace in strip() / split() / len(), capital sigmas in and out of the final position.  Rule,
    device, confidence, token estimate and the cache's exact-match key must equal Python's for every
    text; only malformed UTF-8 (not expressible as a str) is handed back."""
    import hashlib
    _, qre_mod, _ = R.load_modules(None, OracleEngine)
    hr = qre_mod.HeuristicRouter(CFGS[ci])
    corpus = _corpus(5000, 40 + ci, unicode_mix=True)
    qs, cs = [q for q, _ in corpus], [c for _, c in corpus]
    prefix = {"complex pattern": 0, "long query chars": 1, "multi-question count": 2,
              "code/debug markers detected": 3, "large context chars": 4, "simple pattern": 5,
              "short everyday query": 6, "no heuristic match": 7}
    hp = _cabi.HeurParams(hr.long_text_threshold, hr.multi_question_threshold, hr.code_markers_needed,
                          hr.context_chars_threshold, hr._token_fallback.threshold, 2)
    cks = [("default", "ctx-ü")[i % 2] for i in range(len(qs))]
    r = _cabi.host_features_packed(qs, [len(c) if c else 0 for c in cs], hp, context_keys=cks, n_threads=3)
    n_uni = n_sigma = 0
    for i, (q, c) in enumerate(corpus):
        assert r["rule"][i] >= 0, repr(q)
        n_sigma += "Σ" in q
        exp = hr.classify(q, c)
        got = (int(r["heur_dev"][i]), float(r["heur_conf"][i]), int(r["rule"][i]))
        assert got == (exp[0], exp[1], prefix[exp[3].split("=")[0].split(" ->")[0]]), (repr(q), got, exp)
        assert int(r["tok"][i]) == hr._token_fallback.count(q, c), repr(q)
        assert r["keys"][i] == hashlib.md5(f"{cks[i]}||{q.lower().strip()}".encode("utf-8")).hexdigest(), repr(q)
        n_uni += not q.isascii()
    assert n_uni > 1500 and n_sigma > 100
    # malformed UTF-8 through the raw entry point: handed back
    import ctypes as C
    raw = [b"ok text", b"bad \xff\xfe bytes", b"truncated \xe4\xb8", b"overlong \xc0\xaf"]
    arr = (C.c_char_p * len(raw))(*raw)
    ql = np.array([len(x) for x in raw], np.int64)
    out = {k: np.zeros(len(raw), t) for k, t in (("tok", np.int32), ("dev", np.uint8), ("conf", np.float64),
                                                   ("rule", np.int32), ("cat", np.int32))}
    rc = _cabi.load().qre_host_features(C.c_int64(len(raw)), arr, _cabi._ptr(ql), None, C.byref(hp), C.c_int(1),
                                        _cabi._ptr(out["tok"]), _cabi._ptr(out["dev"]), _cabi._ptr(out["conf"]),
                                        _cabi._ptr(out["rule"]), _cabi._ptr(out["cat"]))
    assert rc == 0 and out["rule"][0] >= 0 and out["rule"][1:].tolist() == [-1, -1, -1]


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")
def test_cpp_features_equal_live_reference():
    _, ref_qre, _ = ref_harness.load_reference()
    _, qre_mod, _ = R.load_modules(None, OracleEngine)
    cfg = dict(ref_qre.PRODUCTION_CFG)
    a, b = ref_qre.HeuristicRouter(cfg), qre_mod.HeuristicRouter(cfg)
    corpus = _corpus(4000, 99)
    feats, _ = b.classify_batch([q for q, _ in corpus], [c for _, c in corpus])
    for (q, c), f in zip(corpus, feats):
        d = a.route(q, c)
        assert (["nano", "orin"].index(d.device), d.confidence, d.method, d.reasoning, d.complexity_score) == f, (q, f, d)
