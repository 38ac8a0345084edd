// qre_host_features.cu — host-side per-query features for a whole batch, in C++ threads
// (SURVEY §8f-2).  No device code: it lives in libqre_b200.so so that one binding serves both.
//
// Restates, for ASCII input, the reference's HeuristicRouter.route (src/query_router_engine.py:
// 323-364) — its ordered rule list and the regular expressions of qre:231-294 — and the
// no-tokenizer token estimate of TokenBasedRouter (qre:91,96: max(1, len(text)//4)).  The
// Python `re` patterns are compiled by hand into "segments": sets of literal alternatives with
// \b conditions at either end, joined by `.*` (which, as in Python, does not cross a newline).
// ASCII queries are matched byte-wise.  Other queries are decoded, stripped and lowered with the
// full str.lower() table, and every character gets Python `re`'s class on str patterns (\w = isalnum
// or '_', \s = isspace, \d = isdecimal; tables generated from unicodedata, qre_unicode_tables.inc), so
// \b, \s, \d, len() and split() mean what they mean in the reference (str.lower()'s final-sigma rule
// included).  Only malformed UTF-8 is handed back (rule = -1).
#include <stdint.h>
#include <string.h>

#include <string>
#include <cmath>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/qre_b200.h"

namespace {

inline bool is_word(unsigned char c) {
  return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_';
}
inline bool is_space(unsigned char c) {          // str.isspace() / \s for code points < 128
  return c == 0x20 || (c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x1F);
}
inline bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }

#include "qre_unicode_tables.inc"

inline bool uni_in(const uint32_t (*r)[2], int n, uint32_t cp) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (cp < r[mid][0]) hi = mid - 1;
    else if (cp > r[mid][1]) lo = mid + 1;
    else return true;
  }
  return false;
}
constexpr uint8_t kWord = 1, kSpace = 2, kDigit = 4, kCont = 0x80;   // per-byte classes of a Unicode text
inline uint8_t class_of(uint32_t cp) {
  if (cp < 0x80) {
    const unsigned char c = static_cast<unsigned char>(cp);
    return static_cast<uint8_t>((is_word(c) ? kWord : 0) | (is_space(c) ? kSpace : 0) | (is_digit(c) ? kDigit : 0));
  }
  return static_cast<uint8_t>((uni_in(kUniAlnum, kUniAlnumN, cp) ? kWord : 0) | (uni_in(kUniWhite, kUniWhiteN, cp) ? kSpace : 0) |
                              (uni_in(kUniDecimal, kUniDecimalN, cp) ? kDigit : 0));
}
inline bool utf8_next(const char* s, int64_t n, int64_t& i, uint32_t& cp) {      // strict decode, as in qre_wordpiece.cu
  const unsigned char c0 = static_cast<unsigned char>(s[i]);
  int extra;
  uint32_t v, min_v;
  if (c0 < 0x80) { cp = c0; ++i; return true; }
  if ((c0 & 0xE0) == 0xC0) { extra = 1; v = c0 & 0x1F; min_v = 0x80; }
  else if ((c0 & 0xF0) == 0xE0) { extra = 2; v = c0 & 0x0F; min_v = 0x800; }
  else if ((c0 & 0xF8) == 0xF0) { extra = 3; v = c0 & 0x07; min_v = 0x10000; }
  else return false;
  if (i + extra >= n) return false;
  for (int k = 1; k <= extra; ++k) {
    const unsigned char c = static_cast<unsigned char>(s[i + k]);
    if ((c & 0xC0) != 0x80) return false;
    v = (v << 6) | (c & 0x3F);
  }
  if (v < min_v || v > 0x10FFFF || (v >= 0xD800 && v <= 0xDFFF)) return false;
  cp = v;
  i += extra + 1;
  return true;
}
inline void utf8_put(std::string& w, uint32_t cp) {
  if (cp < 0x80) {
    w.push_back(static_cast<char>(cp));
  } else if (cp < 0x800) {
    w.push_back(static_cast<char>(0xC0 | (cp >> 6)));
    w.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
  } else if (cp < 0x10000) {
    w.push_back(static_cast<char>(0xE0 | (cp >> 12)));
    w.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
    w.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
  } else {
    w.push_back(static_cast<char>(0xF0 | (cp >> 18)));
    w.push_back(static_cast<char>(0x80 | ((cp >> 12) & 0x3F)));
    w.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
    w.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
  }
}

struct View {
  const char* p;
  int n;
  const uint8_t* cls = nullptr;                  // Unicode text only: class bits of the character starting at each byte
  int nchars = 0;                                // characters (== n for ASCII text)
  uint64_t has[4] = {0, 0, 0, 0};                // which byte values occur (prefilter for the literal search)
  void index() {
    for (int i = 0; i < n; ++i) {
      const unsigned char c = static_cast<unsigned char>(p[i]);
      has[c >> 6] |= 1ull << (c & 63);
    }
  }
  bool may_contain(const std::string& lit) const {   // every literal here is non-empty
    const unsigned char a = static_cast<unsigned char>(lit[0]), b = static_cast<unsigned char>(lit.back());
    return ((has[a >> 6] >> (a & 63)) & (has[b >> 6] >> (b & 63)) & 1ull) != 0 && static_cast<int>(lit.size()) <= n;
  }
  unsigned char at(int i) const { return (i >= 0 && i < n) ? static_cast<unsigned char>(p[i]) : 0; }
  // class tests of the character that STARTS at byte i (false outside the text)
  bool word_at(int i) const { return i >= 0 && i < n && (cls ? (cls[i] & kWord) != 0 : is_word(at(i))); }
  bool space_at(int i) const { return i >= 0 && i < n && (cls ? (cls[i] & kSpace) != 0 : is_space(at(i))); }
  bool digit_at(int i) const { return i >= 0 && i < n && (cls ? (cls[i] & kDigit) != 0 : is_digit(at(i))); }
  int next(int i) const {                        // start of the following character
    ++i;
    if (cls) while (i < n && (cls[i] & kCont)) ++i;
    return i;
  }
  bool word_before(int i) const {                // the character that ends at byte i
    if (i <= 0) return false;
    int j = i - 1;
    if (cls) while (j > 0 && (cls[j] & kCont)) --j;
    return word_at(j);
  }
  bool boundary(int i) const { return word_before(i) != word_at(i); }      // \b between positions i-1 and i
};

struct Seg {
  bool wb_start, wb_end;
  std::vector<std::string> alts;
};
struct Pat {
  std::vector<Seg> segs;      // joined by .*
  int special = 0;            // 1: a* ; 2: tiny math ; 3: "what is" tiny math
};

// all (start, end) matches of one segment
void seg_matches(const View& s, const Seg& g, std::vector<std::pair<int, int>>& out) {
  out.clear();
  for (const std::string& a : g.alts) {
    if (!s.may_contain(a)) continue;
    const int L = static_cast<int>(a.size());
    const char* end = s.p + (s.n - L + 1);         // one past the last possible start
    for (const char* h = s.p; h < end; ++h) {
      h = static_cast<const char*>(memchr(h, a[0], static_cast<size_t>(end - h)));
      if (h == nullptr) break;
      if (memcmp(h, a.data(), L) != 0) continue;
      const int i = static_cast<int>(h - s.p);
      if (g.wb_start && !s.boundary(i)) continue;
      if (g.wb_end && !s.boundary(i + L)) continue;
      out.emplace_back(i, i + L);
    }
  }
}

bool match_math(const View& s, bool with_what_is) {
  // ^\s*(what is\s+)?\d+\s*[\+\-\*/]\s*\d+\s*\??\s*$   (\s and \d in Python's Unicode sense)
  int i = 0;
  while (s.space_at(i)) i = s.next(i);
  if (with_what_is) {
    if (i + 7 > s.n || memcmp(s.p + i, "what is", 7) != 0) return false;
    i += 7;
    if (!s.space_at(i)) return false;
    while (s.space_at(i)) i = s.next(i);
  }
  int d0 = i;
  while (s.digit_at(i)) i = s.next(i);
  if (i == d0) return false;
  while (s.space_at(i)) i = s.next(i);
  const unsigned char op = s.at(i);
  if (!(op == '+' || op == '-' || op == '*' || op == '/')) return false;
  ++i;
  while (s.space_at(i)) i = s.next(i);
  d0 = i;
  while (s.digit_at(i)) i = s.next(i);
  if (i == d0) return false;
  while (s.space_at(i)) i = s.next(i);
  if (i < s.n && s.at(i) == '?') ++i;
  while (s.space_at(i)) i = s.next(i);
  return i == s.n;
}

bool pat_search(const View& s, const Pat& p) {
  if (p.special == 1) {                          // (?:\b|^)a\*(?:\s|$|\W)
    for (int i = 0; i + 2 <= s.n; ++i)
      if (s.at(i) == 'a' && s.at(i + 1) == '*' && !s.word_before(i) && !s.word_at(i + 2))
        return true;
    return false;
  }
  if (p.special == 2) return match_math(s, false);
  if (p.special == 3) return match_math(s, true);
  std::vector<std::pair<int, int>> cur, nxt, reach;
  seg_matches(s, p.segs[0], cur);
  if (cur.empty()) return false;
  for (size_t k = 1; k < p.segs.size(); ++k) {
    seg_matches(s, p.segs[k], nxt);
    reach.clear();
    for (const auto& m2 : nxt) {
      bool ok = false;
      for (const auto& m1 : cur) {
        if (m2.first < m1.second) continue;
        bool nl = false;                          // `.` does not match '\n'
        for (int j = m1.second; j < m2.first && !nl; ++j) nl = s.at(j) == '\n';
        if (!nl) { ok = true; break; }
      }
      if (ok) reach.push_back(m2);
    }
    cur.swap(reach);
    if (cur.empty()) return false;
  }
  return true;
}

Pat words(std::vector<std::string> alts, bool wb_start = true, bool wb_end = true) {
  Pat p;
  p.segs.push_back(Seg{wb_start, wb_end, std::move(alts)});
  return p;
}
Pat gap(std::vector<std::string> a, std::vector<std::string> b) {
  Pat p;
  p.segs.push_back(Seg{true, true, std::move(a)});
  p.segs.push_back(Seg{true, true, std::move(b)});
  return p;
}
Pat special(int k) {
  Pat p;
  p.special = k;
  return p;
}

struct Tables {
  std::vector<std::vector<Pat>> complex_cats, simple_cats;   // in the reference's dict order
  std::vector<std::string> code_markers;
  Tables() {
    complex_cats = {
        // code_build_debug
        {words({"write", "implement", "code", "program", "script", "build", "refactor", "debug", "fix"}),
         words({"traceback", "exception", "error", "segfault", "timeout", "hanging"}),
         words({"api", "flask", "fastapi", "docker", "kubernetes", "ssh", "tunnel", "nginx"}),
         words({"system design", "architecture", "distributed", "scalability", "load balanc"})},
        // math_cs_theory
        {words({"prove", "lemma", "theorem", "corollary", "derivative", "integral", "gradient"}),
         words({"time complexity", "space complexity", "big-o", "big o"}),
         words({"dynamic programming", "dp", "graph", "dijkstra", "bfs", "dfs"}), special(1)},
        // reasoning_comparison
        {words({"compare", "contrast", "difference between", "pros and cons", "vs.", "vs", "versus"}),
         words({"evaluate", "assess", "critique", "analyze"})},
        // long_form_generation
        {words({"essay", "report", "proposal", "research paper", "literature review", "methodology"}),
         words({"comprehensive", "in-depth", "step-by-step", "step by step", "step-by step", "step by-step",
                "detailed"}),
         gap({"summarize", "synthesis"}, {"everything", "all", "so far", "entire"}),
         words({"transcript", "debate", "dialogue", "format as json", "markdown table"})},
        // data_engineering
        {words({"etl", "pipeline", "spark", "hadoop", "presto", "sql", "csv", "excel", "dataframe", "dataset"}),
         words({"deduplicate", "normalize", "clean", "transform", "parse", "extract"})},
        // medical_analysis
        {words({"symptom", "diagnosis", "treatment", "therapy", "prognosis", "chronic", "severe"}),
         words({"pain", "migraine", "dizziness", "fatigue", "nausea", "inflammation", "anxiety", "depression"}),
         gap({"dietary", "meal", "training", "exercise", "recovery"}, {"plan", "schedule", "regimen"}),
         words({"mental health", "psycholog", "counseling", "therapist", "physician", "hospital"})},
        // context_heavy
        {words({"using all context", "using all history", "using all above", "using the context",
                "using the history", "using the above", "based on the conversation", "based on the context",
                "based on our conversation", "based on our context"}),
         gap({"continue", "expand", "build on", "follow up"}, {"previous", "earlier", "above"})},
    };
    simple_cats = {
        // greeting
        {words({"hi", "hello", "hey", "yo", "sup"}), words({"good morning", "good afternoon", "good evening"}),
         words({"thanks", "thank you"})},
        // general_knowledge
        {words({"what is", "who is", "where is", "when is", "when did", "how many", "capital of"}),
         words({"tell me a joke", "fun fact", "random fact"}), words({"how to", "how do i", "can you tell me"})},
        // wellness_tips
        {words({"benefit of", "benefits of", "tip for", "tips for", "advice on"}),
         words({"daily intake", "how often", "how much"}),
         gap({"healthy", "good"}, {"habit", "routine", "lifestyle"})},
        // short_definition
        {words({"define", "meaning of", "definition of"}), gap({"what does"}, {"mean"})},
        // tiny_math
        {special(2), special(3)},
    };
    code_markers = {"```", "def ", "class ", "import ", "Traceback", "Exception", "ModuleNotFoundError",
                    "SELECT ", "WITH ", "FROM ", "JOIN ", ";", "{", "}", "->", "::", "==", "!="};
  }
};

const Tables& tables() {
  static const Tables t;
  return t;
}

bool contains(const View& s, const std::string& m) {
  if (!s.may_contain(m)) return false;
  const int L = static_cast<int>(m.size());
  const char* end = s.p + (s.n - L + 1);
  for (const char* h = s.p; h < end; ++h) {
    h = static_cast<const char*>(memchr(h, m[0], static_cast<size_t>(end - h)));
    if (h == nullptr) return false;
    if (memcmp(h, m.data(), L) == 0) return true;
  }
  return false;
}

// The ordered rule list (qre:323-364) on the stripped query `q` and its lowered form `ql`.
void apply_rules(const View& q, const View& ql, bool has_ctx, int64_t ctx_len, const qre_heur_params_t& hp, int64_t t,
                 uint8_t* dev, double* conf, int32_t* rule, int32_t* cat) {
  const Tables& T = tables();
  const uint8_t heavy = static_cast<uint8_t>(hp.n_devices - 1), light = 0;
  *cat = -1;
  for (size_t c = 0; c < T.complex_cats.size(); ++c)
    for (const Pat& p : T.complex_cats[c])
      if (pat_search(ql, p)) { *rule = 0; *cat = static_cast<int32_t>(c); *dev = heavy; *conf = 0.92; return; }
  if (q.nchars >= hp.long_chars) { *rule = 1; *dev = heavy; *conf = 0.80; return; }
  int qm = 0;
  for (int i = 0; i < q.n; ++i) qm += q.p[i] == '?';
  if (qm >= hp.multi_qmarks) { *rule = 2; *cat = qm; *dev = heavy; *conf = 0.80; return; }
  int marks = 0;
  for (const std::string& m : T.code_markers) marks += contains(q, m) ? 1 : 0;
  if (marks >= hp.code_markers_needed) { *rule = 3; *dev = heavy; *conf = 0.88; return; }
  if (has_ctx && ctx_len >= hp.context_chars) { *rule = 4; *dev = heavy; *conf = 0.75; return; }
  for (size_t c = 0; c < T.simple_cats.size(); ++c)
    for (const Pat& p : T.simple_cats[c])
      if (pat_search(ql, p)) { *rule = 5; *cat = static_cast<int32_t>(c); *dev = light; *conf = 0.90; return; }
  int words_n = 0;                               // len(low.split())
  for (int i = 0; i < ql.n;) {
    while (ql.space_at(i)) i = ql.next(i);
    if (i < ql.n) ++words_n;
    while (i < ql.n && !ql.space_at(i)) i = ql.next(i);
  }
  if (words_n <= 15 && q.nchars <= 100) { *rule = 6; *dev = light; *conf = 0.75; return; }
  // no rule fired: token router's device, half its confidence (qre:357-364)
  const int thr = hp.token_threshold;
  const int64_t diff = t > thr ? t - thr : thr - t;
  double c = static_cast<double>(diff) / static_cast<double>(thr > 1 ? thr : 1);
  if (c > 1.0) c = 1.0;
  *rule = 7;
  *dev = t > thr ? 1 : 0;
  *conf = c * 0.5;
}

inline const UniMap* lower_of(uint32_t cp) {
  int lo = 0, hi = kUniLowerN - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (cp < kUniLower[mid].cp) hi = mid - 1;
    else if (cp > kUniLower[mid].cp) lo = mid + 1;
    else return &kUniLower[mid];
  }
  return nullptr;
}

// `low_out` (optional) receives query.lower().strip() as UTF-8 — what the cache's exact-match key hashes.
void classify_one(const char* q_raw, int64_t q_len, int64_t ctx_len, const qre_heur_params_t& hp,
                  int32_t* tok, uint8_t* dev, double* conf, int32_t* rule, int32_t* cat, std::string* low_out = nullptr) {
  const bool has_ctx = ctx_len > 0;               // `if context` — an empty string is falsy
  bool ascii = true;
  for (int64_t i = 0; i < q_len && ascii; ++i) ascii = static_cast<unsigned char>(q_raw[i]) < 0x80;
  auto declined = [&]() { *rule = -1; *cat = -1; *dev = 0; *conf = 0.0; *tok = 1; };
  std::string low_local;
  std::string& low = low_out ? *low_out : low_local;
  low.clear();
  if (ascii) {
    // token estimate over f"{context}\n{query}" (qre:91,96); the ORIGINAL query, not the stripped one
    const int64_t full_len = has_ctx ? ctx_len + 1 + q_len : q_len;
    const int64_t t = full_len / 4 > 1 ? full_len / 4 : 1;
    *tok = static_cast<int32_t>(t > 0x7fffffff ? 0x7fffffff : t);
    // q = query.strip(); ql = q.lower()
    int64_t b = 0, e = q_len;
    while (b < e && is_space(static_cast<unsigned char>(q_raw[b]))) ++b;
    while (e > b && is_space(static_cast<unsigned char>(q_raw[e - 1]))) --e;
    View q{q_raw + b, static_cast<int>(e - b)};
    q.nchars = q.n;
    q.index();
    low.assign(q.p, static_cast<size_t>(q.n));
    for (char& c : low)
      if (c >= 'A' && c <= 'Z') c = static_cast<char>(c - 'A' + 'a');
    View ql{low.data(), q.n};
    ql.nchars = q.n;
    ql.index();
    apply_rules(q, ql, has_ctx, ctx_len, hp, t, dev, conf, rule, cat);
    return;
  }
  // ---- general Unicode text ----
  std::vector<uint32_t> cps;
  std::vector<int64_t> at;                       // byte offset of each character (+ the end)
  for (int64_t i = 0; i < q_len;) {
    uint32_t cp;
    at.push_back(i);
    if (!utf8_next(q_raw, q_len, i, cp)) { declined(); return; }                   // malformed UTF-8
    cps.push_back(cp);
  }
  at.push_back(q_len);
  const int64_t n_all = static_cast<int64_t>(cps.size());
  const int64_t full_len = has_ctx ? ctx_len + 1 + n_all : n_all;                  // len() counts characters
  const int64_t t = full_len / 4 > 1 ? full_len / 4 : 1;
  *tok = static_cast<int32_t>(t > 0x7fffffff ? 0x7fffffff : t);
  int64_t b = 0, e = n_all;
  while (b < e && (class_of(cps[b]) & kSpace)) ++b;
  while (e > b && (class_of(cps[e - 1]) & kSpace)) --e;
  View q{q_raw + at[b], static_cast<int>(at[e] - at[b])};
  q.nchars = static_cast<int>(e - b);
  q.index();
  // `fold` is what the patterns are matched on: the lowered text with the two characters that Python's
  // re.IGNORECASE treats as equal to an ASCII letter replaced by it (dotless i ~ i, long s ~ s; both are
  // word characters like the letters they stand for).  `low` stays the true lower() for the cache key.
  std::string fold;
  std::vector<uint8_t> cls;
  int low_chars = 0;
  auto put = [&](uint32_t cp) {
    if (low_out) utf8_put(low, cp);
    const uint32_t m = cp == 0x0131 ? 'i' : (cp == 0x017F ? 's' : cp);
    const size_t start = fold.size();
    utf8_put(fold, m);
    cls.push_back(class_of(m));
    for (size_t k = start + 1; k < fold.size(); ++k) cls.push_back(kCont);
    ++low_chars;
  };
  // str.lower()'s one context rule: capital sigma becomes the FINAL form iff a cased character precedes it
  // and none follows, case-ignorable characters skipped on both sides (looked up over the whole query:
  // the stripped-off whitespace is neither)
  auto cased = [](uint32_t c) {
    return c < 0x80 ? ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) : uni_in(kUniCased, kUniCasedN, c);
  };
  auto ignorable = [](uint32_t c) {
    return c < 0x80 ? (c == '\'' || c == '.' || c == ':' || c == '^' || c == '`') : uni_in(kUniCaseIgnorable, kUniCaseIgnorableN, c);
  };
  auto final_sigma = [&](int64_t i) {
    int64_t j = i - 1;
    while (j >= 0 && ignorable(cps[j])) --j;
    if (j < 0 || !cased(cps[j])) return false;
    int64_t k = i + 1;
    while (k < n_all && ignorable(cps[k])) ++k;
    return k == n_all || !cased(cps[k]);
  };
  for (int64_t i = b; i < e; ++i) {
    const uint32_t cp = cps[i];
    if (cp < 0x80) {
      put(cp >= 'A' && cp <= 'Z' ? cp - 'A' + 'a' : cp);
    } else if (cp == 0x03A3) {
      put(final_sigma(i) ? 0x03C2 : 0x03C3);
    } else if (const UniMap* m = lower_of(cp)) {
      for (uint32_t k = 0; k < m->len; ++k) put(kUniLowerPool[m->off + k]);
    } else {
      put(cp);
    }
  }
  View ql{fold.data(), static_cast<int>(fold.size())};
  ql.cls = cls.data();
  ql.nchars = low_chars;
  ql.index();
  apply_rules(q, ql, has_ctx, ctx_len, hp, t, dev, conf, rule, cat);
}

// ---- MD5 (RFC 1321), for the cache's exact-match key: md5("<ctx>||<query lowered, stripped>") ----
struct Md5 {
  uint32_t K[64];
  Md5() {
    for (int i = 0; i < 64; ++i) K[i] = static_cast<uint32_t>(std::floor(std::fabs(std::sin(static_cast<double>(i + 1))) * 4294967296.0));
  }
  static uint32_t rotl(uint32_t x, int c) { return (x << c) | (x >> (32 - c)); }
  void block(uint32_t st[4], const unsigned char* p) const {
    static const int R[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9,  14, 20, 5, 9,
                              14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                              4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
    uint32_t M[16];
    for (int i = 0; i < 16; ++i)
      M[i] = static_cast<uint32_t>(p[4 * i]) | (static_cast<uint32_t>(p[4 * i + 1]) << 8) |
             (static_cast<uint32_t>(p[4 * i + 2]) << 16) | (static_cast<uint32_t>(p[4 * i + 3]) << 24);
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    for (int i = 0; i < 64; ++i) {
      uint32_t f;
      int g;
      if (i < 16) { f = (b & c) | (~b & d); g = i; }
      else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
      else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
      else { f = c ^ (b | ~d); g = (7 * i) & 15; }
      const uint32_t t = d;
      d = c;
      c = b;
      b = b + rotl(a + f + K[i] + M[g], R[i]);
      a = t;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
  }
  // hex digest of the concatenation of up to three byte ranges
  void hex(const char* const* parts, const int64_t* lens, int n_parts, char* out32) const {
    uint32_t st[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    unsigned char buf[64];
    int fill = 0;
    uint64_t total = 0;
    for (int k = 0; k < n_parts; ++k) {
      const unsigned char* p = reinterpret_cast<const unsigned char*>(parts[k]);
      int64_t n = lens[k];
      total += static_cast<uint64_t>(n);
      while (n > 0) {
        const int take = static_cast<int>(n < 64 - fill ? n : 64 - fill);
        memcpy(buf + fill, p, take);
        fill += take; p += take; n -= take;
        if (fill == 64) { block(st, buf); fill = 0; }
      }
    }
    buf[fill++] = 0x80;
    if (fill > 56) { memset(buf + fill, 0, 64 - fill); block(st, buf); fill = 0; }
    memset(buf + fill, 0, 56 - fill);
    const uint64_t bits = total * 8;
    for (int i = 0; i < 8; ++i) buf[56 + i] = static_cast<unsigned char>(bits >> (8 * i));
    block(st, buf);
    static const char* H = "0123456789abcdef";
    for (int i = 0; i < 16; ++i) {
      const unsigned char v = static_cast<unsigned char>(st[i >> 2] >> (8 * (i & 3)));
      out32[2 * i] = H[v >> 4];
      out32[2 * i + 1] = H[v & 15];
    }
  }
};
const Md5& md5() {
  static const Md5 m;
  return m;
}

}  // namespace

extern "C" QRE_API int qre_host_features(int64_t n, const char* const* queries, const int64_t* query_len,
                                         const int64_t* context_len, const qre_heur_params_t* hp, int n_threads,
                                         int32_t* tok, uint8_t* heur_dev, double* heur_conf, int32_t* heur_rule,
                                         int32_t* heur_cat) {
  if (n < 0 || hp == nullptr || (n > 0 && (!queries || !query_len || !tok || !heur_dev || !heur_conf ||
                                           !heur_rule || !heur_cat)))
    return QRE_ERR_INVALID;
  if (hp->n_devices < 1) return QRE_ERR_INVALID;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  if (n < 64) n_threads = 1;
  tables();                                       // build the tables before any thread needs them
  auto work = [&](int64_t i0, int64_t i1) {
    for (int64_t i = i0; i < i1; ++i)
      classify_one(queries[i] ? queries[i] : "", queries[i] ? query_len[i] : 0, context_len ? context_len[i] : 0,
                   *hp, &tok[i], &heur_dev[i], &heur_conf[i], &heur_rule[i], &heur_cat[i]);
  };
  if (n_threads == 1) {
    work(0, n);
    return QRE_OK;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t) th.emplace_back(work, n * t / n_threads, n * (t + 1) / n_threads);
  for (auto& x : th) x.join();
  return QRE_OK;
}

// The same over ONE byte blob (queries[i] = blob[offsets[i] .. offsets[i+1])), which spares the
// caller an array of n pointers, plus — when key_hex is given — the cache's exact-match key of every
// ASCII query: md5(ctx_keys[ctx_index[i]] + "||" + lower(strip(query))) as 32 hex characters
// (src/cache.py:518-521).  Rows with heur_rule -1 (non-ASCII query) are left for the caller.
extern "C" QRE_API int qre_host_features_packed(int64_t n, const char* blob, const int64_t* offsets,
                                                const int64_t* context_len, const qre_heur_params_t* hp,
                                                int n_threads, int32_t* tok, uint8_t* heur_dev, double* heur_conf,
                                                int32_t* heur_rule, int32_t* heur_cat, const char* ctx_blob,
                                                const int64_t* ctx_offsets, const int32_t* ctx_index,
                                                int32_t n_ctx, char* key_hex) {
  if (n < 0 || hp == nullptr ||
      (n > 0 && (!blob || !offsets || !tok || !heur_dev || !heur_conf || !heur_rule || !heur_cat)))
    return QRE_ERR_INVALID;
  if (hp->n_devices < 1) return QRE_ERR_INVALID;
  if (key_hex != nullptr && (!ctx_blob || !ctx_offsets || !ctx_index || n_ctx < 1)) return QRE_ERR_INVALID;
  for (int64_t i = 0; i < n; ++i)
    if (offsets[i + 1] < offsets[i]) return QRE_ERR_INVALID;
  if (key_hex != nullptr)
    for (int64_t i = 0; i < n; ++i)
      if (ctx_index[i] < 0 || ctx_index[i] >= n_ctx) return QRE_ERR_INVALID;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  if (n < 64) n_threads = 1;
  tables();
  const Md5& H = md5();
  auto work = [&](int64_t i0, int64_t i1) {
    std::string low;
    for (int64_t i = i0; i < i1; ++i) {
      const char* q = blob + offsets[i];
      const int64_t len = offsets[i + 1] - offsets[i];
      classify_one(q, len, context_len ? context_len[i] : 0, *hp, &tok[i], &heur_dev[i], &heur_conf[i],
                   &heur_rule[i], &heur_cat[i], &low);
      if (key_hex == nullptr) continue;
      char* out = key_hex + 32 * i;
      if (heur_rule[i] < 0) { memset(out, 0, 32); continue; }
      // `low` = query.lower().strip(), left there by the classifier
      const int32_t ci = ctx_index[i];
      const char* parts[3] = {ctx_blob + ctx_offsets[ci], "||", low.data()};
      const int64_t lens[3] = {ctx_offsets[ci + 1] - ctx_offsets[ci], 2, static_cast<int64_t>(low.size())};
      H.hex(parts, lens, 3, out);
    }
  };
  if (n_threads == 1) {
    work(0, n);
    return QRE_OK;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t) th.emplace_back(work, n * t / n_threads, n * (t + 1) / n_threads);
  for (auto& x : th) x.join();
  return QRE_OK;
}
