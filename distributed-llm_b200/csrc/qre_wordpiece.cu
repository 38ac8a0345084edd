// qre_wordpiece.cu — host-side WordPiece for the on-GPU embedder (include/qre_encoder_b200.h).
//
// SentenceTransformer.encode (src/query_router_engine.py:181, 571) tokenises in Python before
// the model runs; at the batch sizes the fused lookup is built for, that step has to keep up.
// This is BERT's uncased scheme for ASCII text on C++ threads: clean (drop NUL / control
// characters, fold \t \n \r to space), lower-case, split on whitespace, make every ASCII
// punctuation character its own token, then greedy longest-match-first pieces with the "##"
// continuation prefix (a word with no full cover, or longer than 100 characters, is [UNK]).
// Output is `[CLS] pieces [SEP]`, truncated to max_len.  Restated with full Unicode handling in
// oracle/minilm_oracle.py and pinned against transformers' BertTokenizer there; texts with any
// non-ASCII byte are left to the caller (length -1), as qre_host_features does.
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

// Returns false for a text with a non-ASCII byte.
bool encode_one(const qre_wordpiece& t, const char* s, int64_t n, int max_len, std::vector<int32_t>& out,
                std::string& word, std::string& probe) {
  for (int64_t i = 0; i < n; ++i)
    if (static_cast<unsigned char>(s[i]) >= 0x80) return false;
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
