// qre_wordpiece.cu — host-side WordPiece for the on-GPU embedder (include/qre_encoder_b200.h).
//
// SentenceTransformer.encode (src/query_router_engine.py:181, 571) tokenises in Python before
// the model runs; at the batch sizes the fused lookup is built for, that step has to keep up.
// This is BERT's uncased scheme for ASCII text on C++ threads: clean (drop NUL / control
// characters, fold \t \n \r to space), lower-case, split on whitespace, make every ASCII
// punctuation character its own token, then greedy longest-match-first pieces with the "##"
// continuation prefix (a word with no full cover, or longer than 100 characters, is [UNK]).
// Output is `[CLS] pieces [SEP]`, truncated to max_len.  Restated in oracle/minilm_oracle.py and
// pinned against transformers' BertTokenizer there.  Texts with non-ASCII characters take a second
// path here (encode_one_unicode): UTF-8 decode, BasicTokenizer's cleaning / CJK spacing on the
// original characters, NFD + strip-Mn + lower through a table generated from Python's unicodedata
// (qre_unicode_tables.inc, qre_b200/gen_unicode_tables.py), Unicode whitespace and punctuation
// classes, pieces matched on character boundaries.  Only texts holding a character whose handling
// depends on its neighbours (capital sigma, reordering combining marks) or invalid UTF-8 are left
// to the caller (length -1).
// No CUDA in this file; it is a .cu only so that one compiler invocation builds the library.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "qre_encoder.cuh"

using namespace qre;

struct qre_wordpiece {
  std::unordered_map<std::string, int32_t> vocab;
  int32_t cls = -1, sep = -1, unk = -1;
  size_t max_piece = 0;      // longest vocab entry in bytes: bounds the longest-match search
  bool lower = true;
};

namespace {

inline bool is_punct(unsigned char c) {
  return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
}

#include "qre_unicode_tables.inc"

inline bool in_ranges(const uint32_t (*r)[2], int n, uint32_t cp) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (cp < r[mid][0]) hi = mid - 1;
    else if (cp > r[mid][1]) lo = mid + 1;
    else return true;
  }
  return false;
}
inline const UniMap* find_map(uint32_t cp) {
  int lo = 0, hi = kUniMapN - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (cp < kUniMap[mid].cp) hi = mid - 1;
    else if (cp > kUniMap[mid].cp) lo = mid + 1;
    else return &kUniMap[mid];
  }
  return nullptr;
}
inline bool is_cjk(uint32_t cp) {            // BasicTokenizer._is_chinese_char
  return (cp >= 0x4E00 && cp <= 0x9FFF) || (cp >= 0x3400 && cp <= 0x4DBF) || (cp >= 0x20000 && cp <= 0x2A6DF) ||
         (cp >= 0x2A700 && cp <= 0x2B73F) || (cp >= 0x2B740 && cp <= 0x2B81F) || (cp >= 0x2B820 && cp <= 0x2CEAF) ||
         (cp >= 0xF900 && cp <= 0xFAFF) || (cp >= 0x2F800 && cp <= 0x2FA1F);
}
// Strict UTF-8 decode of the character at s[i]; false on a malformed, overlong or surrogate sequence.
inline bool decode_utf8(const char* s, int64_t n, int64_t& i, uint32_t& cp) {
  const unsigned char c0 = static_cast<unsigned char>(s[i]);
  int extra;
  uint32_t v, min_v;
  if (c0 < 0x80) { cp = c0; ++i; return true; }
  if ((c0 & 0xE0) == 0xC0) { extra = 1; v = c0 & 0x1F; min_v = 0x80; }
  else if ((c0 & 0xF0) == 0xE0) { extra = 2; v = c0 & 0x0F; min_v = 0x800; }
  else if ((c0 & 0xF8) == 0xF0) { extra = 3; v = c0 & 0x07; min_v = 0x10000; }
  else return false;
  if (i + extra >= n) return false;                 // truncated sequence
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
inline void append_utf8(std::string& w, uint32_t cp) {
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

// word_to_ids for a UTF-8 word of `nchars` characters: pieces start and end on character boundaries.
void word_to_ids_u(const qre_wordpiece& t, const std::string& w, int nchars, std::vector<int32_t>& out,
                   std::string& probe, std::vector<uint32_t>& b) {
  if (nchars > 100) {
    out.push_back(t.unk);
    return;
  }
  b.clear();
  for (size_t i = 0; i < w.size(); ++i)
    if ((static_cast<unsigned char>(w[i]) & 0xC0) != 0x80) b.push_back(static_cast<uint32_t>(i));
  b.push_back(static_cast<uint32_t>(w.size()));
  const size_t n = b.size() - 1, mark = out.size();
  size_t start = 0;
  while (start < n) {
    size_t end = n;
    while (end > start + 1 && b[end] - b[start] > t.max_piece) --end;     // longer than any vocab entry
    int32_t found = -1;
    for (; end > start; --end) {
      probe.clear();
      if (start > 0) probe.append("##");
      probe.append(w, b[start], b[end] - b[start]);
      auto it = t.vocab.find(probe);
      if (it != t.vocab.end()) {
        found = it->second;
        break;
      }
    }
    if (found < 0) {
      out.resize(mark);
      out.push_back(t.unk);
      return;
    }
    out.push_back(found);
    start = end;
  }
}

// Appends the pieces of one word; returns false when the word has no cover ([UNK]).
void word_to_ids(const qre_wordpiece& t, const std::string& w, std::vector<int32_t>& out, std::string& probe) {
  if (w.size() > 100) {
    out.push_back(t.unk);
    return;
  }
  const size_t mark = out.size();
  size_t start = 0;
  while (start < w.size()) {
    size_t end = std::min(w.size(), start + t.max_piece);
    int32_t found = -1;
    for (; end > start; --end) {
      probe.clear();
      if (start > 0) probe.append("##");
      probe.append(w, start, end - start);
      auto it = t.vocab.find(probe);
      if (it != t.vocab.end()) {
        found = it->second;
        break;
      }
    }
    if (found < 0) {
      out.resize(mark);
      out.push_back(t.unk);
      return;
    }
    out.push_back(found);
    start = end;
  }
}

// The general path (any Unicode text).  Returns false — nothing appended — for a text that must go to
// the caller's tokenizer: malformed UTF-8, or a character whose handling depends on its neighbours.
bool encode_one_unicode(const qre_wordpiece& t, const char* s, int64_t n, int max_len, std::vector<int32_t>& out,
                        std::string& word, std::string& probe) {
  const size_t mark = out.size();
  std::vector<uint32_t> bounds;
  out.push_back(t.cls);
  word.clear();
  int nchars = 0;
  auto flush = [&]() {
    if (!word.empty()) {
      word_to_ids_u(t, word, nchars, out, probe, bounds);
      word.clear();
      nchars = 0;
    }
  };
  auto put = [&](uint32_t m) {                   // one character of the normalised text
    if (m < 0x80) {
      const unsigned char c = static_cast<unsigned char>(m);
      if (c == ' ' || (c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x1F)) {
        flush();
      } else if (is_punct(c)) {
        flush();
        word.push_back(static_cast<char>(c));
        nchars = 1;
        flush();
      } else {
        word.push_back(static_cast<char>(c));
        ++nchars;
      }
    } else if (in_ranges(kUniWhite, kUniWhiteN, m)) {
      flush();
    } else if (in_ranges(kUniPunct, kUniPunctN, m)) {
      flush();
      append_utf8(word, m);
      nchars = 1;
      flush();
    } else {
      append_utf8(word, m);
      ++nchars;
    }
  };
  int64_t i = 0;
  while (i < n) {
    uint32_t cp;
    if (!decode_utf8(s, n, i, cp)) {
      out.resize(mark);
      return false;
    }
    if (cp < 0x80) {
      unsigned char c = static_cast<unsigned char>(cp);
      if (c == ' ' || c == '\t' || c == '\n' || c == '\r') { flush(); continue; }
      if (c < 32 || c == 127) continue;          // NUL and the other control characters vanish
      if (t.lower && c >= 'A' && c <= 'Z') c = static_cast<unsigned char>(c - 'A' + 'a');
      put(c);
      continue;
    }
    if (in_ranges(kUniDecline, kUniDeclineN, cp)) {
      out.resize(mark);
      return false;
    }
    if (cp == 0xFFFD || in_ranges(kUniSkip, kUniSkipN, cp)) continue;
    if (in_ranges(kUniSpace, kUniSpaceN, cp)) { flush(); continue; }
    const bool cjk = is_cjk(cp);                 // spaced out BEFORE normalisation
    if (cjk) flush();
    if (!t.lower) {
      put(cp);
    } else if (cp >= 0xAC00 && cp <= 0xD7A3) {   // Hangul syllable: NFD is arithmetic
      const uint32_t si = cp - 0xAC00;
      put(0x1100 + si / 588);
      put(0x1161 + (si % 588) / 28);
      if (si % 28) put(0x11A7 + si % 28);
    } else if (const UniMap* m = find_map(cp)) {
      for (uint32_t k = 0; k < m->len; ++k) put(kUniPool[m->off + k]);
    } else {
      put(cp);
    }
    if (cjk) flush();
  }
  flush();
  const size_t keep = static_cast<size_t>(max_len > 2 ? max_len - 2 : 0);
  if (out.size() - mark - 1 > keep) out.resize(mark + 1 + keep);
  out.push_back(t.sep);
  return true;
}

// ASCII texts take the byte-wise path below; anything else the general one.
bool encode_one(const qre_wordpiece& t, const char* s, int64_t n, int max_len, std::vector<int32_t>& out,
                std::string& word, std::string& probe) {
  for (int64_t i = 0; i < n; ++i)
    if (static_cast<unsigned char>(s[i]) >= 0x80) return encode_one_unicode(t, s, n, max_len, out, word, probe);
  const size_t mark = out.size();
  out.push_back(t.cls);
  word.clear();
  auto flush = [&]() {
    if (!word.empty()) {
      word_to_ids(t, word, out, probe);
      word.clear();
    }
  };
  for (int64_t i = 0; i < n; ++i) {
    unsigned char c = static_cast<unsigned char>(s[i]);
    if (c == ' ' || c == '\t' || c == '\n' || c == '\r') {
      flush();
    } else if (c == 0 || c < 32 || c == 127) {
      continue;                                  // control characters vanish without splitting the word
    } else if (is_punct(c)) {
      flush();
      word.push_back(static_cast<char>(c));
      flush();
    } else {
      if (t.lower && c >= 'A' && c <= 'Z') c = static_cast<unsigned char>(c - 'A' + 'a');
      word.push_back(static_cast<char>(c));
    }
  }
  flush();
  const size_t keep = static_cast<size_t>(max_len > 2 ? max_len - 2 : 0);
  if (out.size() - mark - 1 > keep) out.resize(mark + 1 + keep);
  out.push_back(t.sep);
  return true;
}

}  // namespace

// Shared body: text i is the bytes [text_at(i), text_at(i) + len_at(i)).
template <typename TextAt, typename LenAt>
static int wordpiece_encode_impl(const qre_wordpiece_t* t, int64_t n, TextAt text_at, LenAt len_at, int max_len,
                                 int32_t* ids_out, int64_t cap, int32_t* lens_out, int n_threads);

extern "C" {

int qre_wordpiece_create(const char* vocab_path, int do_lower_case, qre_wordpiece_t** out) {
  if (vocab_path == nullptr || out == nullptr) return set_error(QRE_ERR_INVALID, "null argument");
  *out = nullptr;
  FILE* f = fopen(vocab_path, "rb");
  if (f == nullptr) return set_error(QRE_ERR_INVALID, "wordpiece: cannot open %s", vocab_path);
  qre_wordpiece* t = new qre_wordpiece();
  t->lower = do_lower_case != 0;
  std::string line;
  int32_t id = 0;
  int ch;
  auto commit = [&]() {
    while (!line.empty() && (line.back() == '\r' || line.back() == '\n')) line.pop_back();
    // vocab.txt lines are pieces verbatim; a duplicated piece takes the id of its LAST line, as the dict the
    // Python tokenizers load the file into does
    t->vocab[line] = id;
    t->max_piece = std::max(t->max_piece, line.size());
    ++id;
    line.clear();
  };
  while ((ch = fgetc(f)) != EOF) {
    if (ch == '\n') commit();
    else line.push_back(static_cast<char>(ch));
  }
  if (!line.empty()) commit();
  fclose(f);
  auto find = [&](const char* k) {
    auto it = t->vocab.find(k);
    return it == t->vocab.end() ? -1 : it->second;
  };
  t->cls = find("[CLS]");
  t->sep = find("[SEP]");
  t->unk = find("[UNK]");
  if (t->cls < 0 || t->sep < 0 || t->unk < 0) {
    delete t;
    return set_error(QRE_ERR_INVALID, "wordpiece: vocab lacks [CLS], [SEP] or [UNK]");
  }
  *out = t;
  return QRE_OK;
}

int qre_wordpiece_destroy(qre_wordpiece_t* t) {
  delete t;
  return QRE_OK;
}

int qre_wordpiece_encode(const qre_wordpiece_t* t, int64_t n, const char* const* texts, const int64_t* text_len,
                         int max_len, int32_t* ids_out, int64_t cap, int32_t* lens_out, int n_threads) {
  if (t == nullptr || (n > 0 && (texts == nullptr || lens_out == nullptr))) return set_error(QRE_ERR_INVALID, "null argument");
  return wordpiece_encode_impl(
      t, n, [&](int64_t i) { return texts[i]; },
      [&](int64_t i) {
        return texts[i] == nullptr ? 0 : (text_len != nullptr ? text_len[i] : static_cast<int64_t>(strlen(texts[i])));
      },
      max_len, ids_out, cap, lens_out, n_threads);
}

// The same with the texts in ONE byte blob (text i = blob[offsets[i] .. offsets[i+1]), n + 1 offsets):
// a Python caller builds that with one join + encode instead of n bytes objects and n pointers, which
// at 16 host threads cost more than the tokenisation itself.
int qre_wordpiece_encode_packed(const qre_wordpiece_t* t, int64_t n, const char* blob, const int64_t* offsets,
                                int max_len, int32_t* ids_out, int64_t cap, int32_t* lens_out, int n_threads) {
  if (t == nullptr || (n > 0 && (blob == nullptr || offsets == nullptr || lens_out == nullptr)))
    return set_error(QRE_ERR_INVALID, "null argument");
  for (int64_t i = 0; i < n; ++i)
    if (offsets[i + 1] < offsets[i]) return set_error(QRE_ERR_INVALID, "wordpiece: offsets must not decrease");
  return wordpiece_encode_impl(
      t, n, [&](int64_t i) { return blob + offsets[i]; }, [&](int64_t i) { return offsets[i + 1] - offsets[i]; },
      max_len, ids_out, cap, lens_out, n_threads);
}

}  // extern "C"

template <typename TextAt, typename LenAt>
static int wordpiece_encode_impl(const qre_wordpiece_t* t, int64_t n, TextAt text_at, LenAt len_at, int max_len,
                                 int32_t* ids_out, int64_t cap, int32_t* lens_out, int n_threads) {
  if (n < 0 || max_len < 2 || cap < 0) return set_error(QRE_ERR_INVALID, "wordpiece: need n >= 0, max_len >= 2");
  if (n == 0) return QRE_OK;
  int nt = n_threads > 0 ? n_threads : static_cast<int>(std::thread::hardware_concurrency());
  nt = std::max(1, std::min<int>(nt, static_cast<int>(std::min<int64_t>(n, 256))));
  if (n < 64) nt = 1;
  std::vector<std::vector<int32_t>> part(nt);
  const int64_t per = (n + nt - 1) / nt;
  auto work = [&](int w) {
    const int64_t i0 = w * per, i1 = std::min<int64_t>(n, i0 + per);
    std::string word, probe;
    for (int64_t i = i0; i < i1; ++i) {
      const char* s = text_at(i);
      const int64_t len = len_at(i);
      const size_t before = part[w].size();
      lens_out[i] = encode_one(*t, s, len, max_len, part[w], word, probe) ? static_cast<int32_t>(part[w].size() - before) : -1;
    }
  };
  if (nt == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int w = 0; w < nt; ++w) th.emplace_back(work, w);
    for (auto& x : th) x.join();
  }
  int64_t total = 0;
  for (auto& p : part) total += static_cast<int64_t>(p.size());
  if (total > cap) return set_error(QRE_ERR_CAPACITY, "wordpiece: %lld ids, output capacity %lld", (long long)total, (long long)cap);
  if (total > 0 && ids_out == nullptr) return set_error(QRE_ERR_INVALID, "null ids_out");
  int64_t o = 0;
  for (auto& p : part) {
    if (!p.empty()) memcpy(ids_out + o, p.data(), sizeof(int32_t) * p.size());
    o += static_cast<int64_t>(p.size());
  }
  return QRE_OK;
}
