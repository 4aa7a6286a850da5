// tokenizer.hpp -- text <-> token ids for the server.
//
// Two back ends behind one interface:
//   * Bpe: byte-level BPE read from an HF `tokenizer.json` (the file the reference's model download
//     puts next to the weights, llm-d-deploy.yaml:121-132,184): GPT-2 byte-to-unicode alphabet, the
//     Qwen2 or the GPT-2 pre-tokenizer split, merges by rank, added/special tokens matched first.
//     Not implemented: Unicode normalisation (Qwen2 declares NFC; it only matters for text that
//     contains combining sequences) and SentencePiece-style models (Llama/Phi-3 "Metaspace").
//   * byte fallback (no tokenizer file, e.g. `--load-format dummy`): one UTF-8 byte = one id in
//     [0, 256); other ids render as "<|id|>".  Token-id prompts bypass tokenisation entirely, which is
//     what the benchmark uses.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../json.hpp"
#include "unicode_tables.hpp"

namespace b2tok {

// ---- UTF-8 helpers -------------------------------------------------------------------------
inline void utf8_append(std::string& o, uint32_t c) {
  if (c < 0x80) o += (char)c;
  else if (c < 0x800) { o += (char)(0xC0 | (c >> 6)); o += (char)(0x80 | (c & 0x3F)); }
  else if (c < 0x10000) { o += (char)(0xE0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
  else { o += (char)(0xF0 | (c >> 18)); o += (char)(0x80 | ((c >> 12) & 0x3F)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
}
// Decodes `s` into code points with the byte offset of each (invalid bytes become U+FFFD, one per byte).
inline void utf8_decode(const std::string& s, std::vector<uint32_t>& cps, std::vector<uint32_t>& offs) {
  size_t i = 0;
  while (i < s.size()) {
    unsigned char c = (unsigned char)s[i];
    int n = c < 0x80 ? 1 : (c >> 5) == 0x6 ? 2 : (c >> 4) == 0xE ? 3 : (c >> 3) == 0x1E ? 4 : 0;
    uint32_t cp = 0xFFFD;
    if (n == 1) cp = c;
    else if (n >= 2 && i + n <= s.size()) {
      cp = c & (0xFF >> (n + 1));
      bool good = true;
      for (int k = 1; k < n; ++k) {
        unsigned char d = (unsigned char)s[i + k];
        if ((d & 0xC0) != 0x80) { good = false; break; }
        cp = (cp << 6) | (d & 0x3F);
      }
      if (!good) { cp = 0xFFFD; n = 1; }
    } else {
      n = 1;
    }
    cps.push_back(cp);
    offs.push_back((uint32_t)i);
    i += (size_t)n;
  }
  offs.push_back((uint32_t)s.size());
}

// ---- pre-tokenizer splits (hand-written scanners for the two regexes byte-level BPE models ship) --------
// Both return the end (code-point index) of the piece that starts at i.  Ordered alternation, greedy
// quantifiers with the backtracking the regex engine would do.
struct CharClass {
  static bool L(uint32_t c) { return c < 0x80 ? ((c | 0x20) >= 'a' && (c | 0x20) <= 'z') : b2uni::is_letter(c); }
  static bool N(uint32_t c) { return c < 0x80 ? (c >= '0' && c <= '9') : b2uni::is_number(c); }
  static bool S(uint32_t c) { return c < 0x80 ? (c == ' ' || (c >= 9 && c <= 13)) : b2uni::is_space(c); }
  static bool other(uint32_t c) { return !S(c) && !L(c) && !N(c); }  // [^\s\p{L}\p{N}]
};

inline size_t contraction_len(const std::vector<uint32_t>& c, size_t i, bool ignore_case) {
  if (c[i] != '\'' || i + 1 >= c.size()) return 0;
  auto lower = [&](uint32_t x) { return (ignore_case && x >= 'A' && x <= 'Z') ? x + 32 : x; };
  const uint32_t a = lower(c[i + 1]);
  const uint32_t b = i + 2 < c.size() ? lower(c[i + 2]) : 0;
  // order of the alternatives: 's 't 're 've 'm 'll 'd
  if (a == 's' || a == 't') return 2;
  if ((a == 'r' && b == 'e') || (a == 'v' && b == 'e')) return 3;
  if (a == 'm') return 2;
  if (a == 'l' && b == 'l') return 3;
  if (a == 'd') return 2;
  return 0;
}
inline size_t space_tail(const std::vector<uint32_t>& c, size_t i) {  // `\s+(?!\S)` then `\s+`
  size_t j = i;
  while (j < c.size() && CharClass::S(c[j])) ++j;
  if (j == i) return i;
  if (j == c.size()) return j;   // \s+(?!\S) at the end of the text
  if (j - i >= 2) return j - 1;  // leave the last space for the word that follows
  return j;                      // \s+
}
// Qwen2 / Qwen3:
// (?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+
inline size_t next_piece_qwen2(const std::vector<uint32_t>& c, size_t i) {
  const size_t n = c.size();
  if (size_t k = contraction_len(c, i, true)) return i + k;
  {
    size_t j = i;
    if (c[j] != '\r' && c[j] != '\n' && !CharClass::L(c[j]) && !CharClass::N(c[j]) && j + 1 < n && CharClass::L(c[j + 1])) ++j;
    if (CharClass::L(c[j])) {
      while (j < n && CharClass::L(c[j])) ++j;
      return j;
    }
  }
  if (CharClass::N(c[i])) return i + 1;
  {
    size_t j = i;
    if (c[j] == ' ' && j + 1 < n && CharClass::other(c[j + 1])) ++j;
    if (CharClass::other(c[j])) {
      while (j < n && CharClass::other(c[j])) ++j;
      while (j < n && (c[j] == '\r' || c[j] == '\n')) ++j;
      return j;
    }
  }
  {  // \s*[\r\n]+ : up to and including the last CR/LF of the whitespace run
    size_t j = i, last = (size_t)-1;
    while (j < n && CharClass::S(c[j])) {
      if (c[j] == '\r' || c[j] == '\n') last = j;
      ++j;
    }
    if (last != (size_t)-1) return last + 1;
  }
  return space_tail(c, i);
}
// GPT-2 (OPT and other ByteLevel(use_regex=true) models):
// 's|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+
inline size_t next_piece_gpt2(const std::vector<uint32_t>& c, size_t i) {
  const size_t n = c.size();
  if (size_t k = contraction_len(c, i, false)) return i + k;
  size_t j = i;
  if (c[j] == ' ' && j + 1 < n && !CharClass::S(c[j + 1])) ++j;
  if (!CharClass::S(c[j])) {
    if (CharClass::L(c[j])) { while (j < n && CharClass::L(c[j])) ++j; return j; }
    if (CharClass::N(c[j])) { while (j < n && CharClass::N(c[j])) ++j; return j; }
    while (j < n && CharClass::other(c[j])) ++j;
    return j;
  }
  return space_tail(c, i);
}

// ---- byte-level BPE -------------------------------------------------------------------------------------
class Bpe {
 public:
  // Throws std::runtime_error with a message naming what is unsupported.
  explicit Bpe(const std::string& tokenizer_json) {
    b2json::Value j = b2json::parse(tokenizer_json);
    const b2json::Value* model = j.find("model");
    if (!model || !model->find("vocab") || !model->find("merges")) throw std::runtime_error("tokenizer.json: no BPE model");
    const b2json::Value* type = model->find("type");
    if (type && type->is_string() && type->as_string() != "BPE") throw std::runtime_error("tokenizer.json: model type " + type->as_string() + " not supported");
    build_alphabet();
    const b2json::Value* vocab = model->find("vocab");
    id_to_tok_.resize(vocab->size());
    for (auto& kv : *vocab->obj) {
      const int id = (int)kv.second.as_int(-1);
      if (id < 0) continue;
      tok_to_id_[kv.first] = id;
      if ((size_t)id >= id_to_tok_.size()) id_to_tok_.resize((size_t)id + 1);
      id_to_tok_[(size_t)id] = kv.first;
    }
    const b2json::Value* merges = model->find("merges");
    int rank = 0;
    for (size_t i = 0; i < merges->size(); ++i) {
      const b2json::Value& m = merges->at(i);
      std::string a, b;
      if (m.is_string()) {  // "a b"
        const size_t sp = m.str.find(' ');
        if (sp == std::string::npos) continue;
        a = m.str.substr(0, sp);
        b = m.str.substr(sp + 1);
      } else if (m.is_array() && m.size() == 2) {
        a = m.at(0).as_string();
        b = m.at(1).as_string();
      } else {
        continue;
      }
      ranks_[a + '\x01' + b] = rank++;
    }
    // added tokens (special markers such as <|im_start|>): matched in the raw text before anything else
    if (const b2json::Value* added = j.find("added_tokens"))
      for (size_t i = 0; i < added->size(); ++i) {
        const b2json::Value& t = added->at(i);
        const b2json::Value* c = t.find("content");
        const b2json::Value* id = t.find("id");
        if (!c || !id) continue;
        Added a;
        a.content = c->as_string();
        a.id = (int)id->as_int();
        const b2json::Value* sp = t.find("special");
        a.special = sp && sp->as_bool();
        if (a.content.empty()) continue;
        added_.push_back(a);
        if ((size_t)a.id >= id_to_tok_.size()) id_to_tok_.resize((size_t)a.id + 1);
        id_to_added_[a.id] = added_.size() - 1;
      }
    // pre-tokenizer: which split pattern, and ByteLevel's add_prefix_space
    std::string pattern;
    bool bytelevel_regex = false;
    scan_pretok(j.find("pre_tokenizer"), pattern, bytelevel_regex);
    if (!pattern.empty()) {
      if (pattern.find("(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\\r\\n\\p{L}\\p{N}]?\\p{L}+|\\p{N}|") == 0) mode_ = QWEN2;
      else throw std::runtime_error("tokenizer.json: pre-tokenizer pattern not supported: " + pattern);
    } else if (bytelevel_regex) {
      mode_ = GPT2;
    } else {
      throw std::runtime_error("tokenizer.json: no supported pre-tokenizer (need the Qwen2 Split pattern or ByteLevel with use_regex)");
    }
  }

  int vocab_size() const { return (int)id_to_tok_.size(); }

  std::vector<int32_t> encode(const std::string& text) const {
    std::vector<int32_t> out;
    size_t pos = 0;
    while (pos < text.size()) {
      // earliest (then longest) added token at or after pos
      size_t best = std::string::npos, best_k = 0;
      for (size_t k = 0; k < added_.size(); ++k) {
        const size_t f = text.find(added_[k].content, pos);
        if (f == std::string::npos) continue;
        if (f < best || (f == best && added_[k].content.size() > added_[best_k].content.size())) { best = f; best_k = k; }
      }
      const size_t end = best == std::string::npos ? text.size() : best;
      if (end > pos) encode_plain(text.substr(pos, end - pos), out);
      if (best == std::string::npos) break;
      out.push_back(added_[best_k].id);
      pos = best + added_[best_k].content.size();
    }
    return out;
  }

  // Bytes of one token (empty for special tokens when skip_special is set; U+FFFD never appears here:
  // the caller's incremental UTF-8 assembler decides what is printable).
  void token_bytes(int32_t id, bool skip_special, std::string& bytes) const {
    auto it = id_to_added_.find(id);
    if (it != id_to_added_.end()) {
      if (!(skip_special && added_[it->second].special)) bytes += added_[it->second].content;
      return;
    }
    if (id < 0 || (size_t)id >= id_to_tok_.size()) return;
    std::vector<uint32_t> cps, offs;
    utf8_decode(id_to_tok_[(size_t)id], cps, offs);
    for (uint32_t c : cps) {
      auto b = cp_to_byte_.find(c);
      if (b != cp_to_byte_.end()) bytes += (char)b->second;
    }
  }

 private:
  enum Mode { QWEN2, GPT2 } mode_ = QWEN2;
  struct Added { std::string content; int id = 0; bool special = false; };
  std::unordered_map<std::string, int> tok_to_id_;
  std::vector<std::string> id_to_tok_;
  std::unordered_map<std::string, int> ranks_;
  std::vector<Added> added_;
  std::unordered_map<int, size_t> id_to_added_;
  std::string byte_to_str_[256];
  std::unordered_map<uint32_t, unsigned char> cp_to_byte_;
  bool add_prefix_space_ = false;
  mutable std::unordered_map<std::string, std::vector<int32_t>> cache_;  // piece -> ids, shared by the HTTP threads
  mutable std::mutex cache_mu_;

  // GPT-2's bytes_to_unicode: printable Latin-1 bytes stand for themselves, the rest map to U+0100...
  void build_alphabet() {
    int n = 0;
    for (int b = 0; b < 256; ++b) {
      const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
      const uint32_t cp = keep ? (uint32_t)b : (uint32_t)(256 + n++);
      utf8_append(byte_to_str_[b], cp);
      cp_to_byte_[cp] = (unsigned char)b;
    }
  }
  void scan_pretok(const b2json::Value* p, std::string& pattern, bool& bytelevel_regex) {
    if (!p || !p->is_object()) return;
    const std::string type = p->find("type") ? p->find("type")->as_string() : "";
    if (type == "Sequence") {
      if (const b2json::Value* l = p->find("pretokenizers"))
        for (size_t i = 0; i < l->size(); ++i) scan_pretok(&l->at(i), pattern, bytelevel_regex);
    } else if (type == "Split") {
      const b2json::Value* pat = p->find("pattern");
      if (pat && pat->find("Regex")) pattern = pat->find("Regex")->as_string();
    } else if (type == "ByteLevel") {
      const b2json::Value* ur = p->find("use_regex");
      if (!ur || ur->as_bool(true)) bytelevel_regex = true;
      const b2json::Value* ps = p->find("add_prefix_space");
      if (ps && ps->as_bool()) add_prefix_space_ = true;
    }
  }

  void encode_plain(const std::string& raw, std::vector<int32_t>& out) const {
    std::string text = raw;
    if (add_prefix_space_ && !text.empty() && text[0] != ' ') text.insert(text.begin(), ' ');
    std::vector<uint32_t> cps, offs;
    utf8_decode(text, cps, offs);
    size_t i = 0;
    while (i < cps.size()) {
      size_t j = mode_ == QWEN2 ? next_piece_qwen2(cps, i) : next_piece_gpt2(cps, i);
      if (j <= i) j = i + 1;  // cannot happen for a total pattern; never loop forever
      bpe_piece(text.substr(offs[i], offs[j] - offs[i]), out);
      i = j;
    }
  }

  void bpe_piece(const std::string& piece, std::vector<int32_t>& out) const {
    {
      std::lock_guard<std::mutex> l(cache_mu_);
      auto hit = cache_.find(piece);
      if (hit != cache_.end()) {
        out.insert(out.end(), hit->second.begin(), hit->second.end());
        return;
      }
    }
    std::vector<std::string> sym;
    sym.reserve(piece.size());
    for (unsigned char b : piece) sym.push_back(byte_to_str_[b]);
    // merge the lowest-ranked adjacent pair (leftmost on ties) until none is left
    while (sym.size() > 1) {
      int best_rank = INT32_MAX;
      size_t best = 0;
      for (size_t k = 0; k + 1 < sym.size(); ++k) {
        auto r = ranks_.find(sym[k] + '\x01' + sym[k + 1]);
        if (r != ranks_.end() && r->second < best_rank) { best_rank = r->second; best = k; }
      }
      if (best_rank == INT32_MAX) break;
      sym[best] += sym[best + 1];
      sym.erase(sym.begin() + (long)best + 1);
    }
    std::vector<int32_t> ids;
    for (auto& s : sym) {
      auto t = tok_to_id_.find(s);
      if (t != tok_to_id_.end()) { ids.push_back(t->second); continue; }
      // a symbol the vocabulary lacks (cannot happen for a complete byte alphabet): fall back to its bytes
      std::vector<uint32_t> cps, offs;
      utf8_decode(s, cps, offs);
      for (uint32_t c : cps) {
        std::string one;
        utf8_append(one, c);
        auto t1 = tok_to_id_.find(one);
        if (t1 != tok_to_id_.end()) ids.push_back(t1->second);
      }
    }
    {
      std::lock_guard<std::mutex> l(cache_mu_);
      if (cache_.size() < 200000) cache_[piece] = ids;
    }
    out.insert(out.end(), ids.begin(), ids.end());
  }
};

// ---- the interface the server uses ----------------------------------------------------------------------
// One process-wide tokenizer: set once at start-up (before the HTTP threads exist), read-only afterwards
// except for Bpe's piece cache, which has its own lock.
inline std::shared_ptr<Bpe>& global_bpe() {
  static std::shared_ptr<Bpe> g;
  return g;
}

inline std::vector<int32_t> encode(const std::string& text) {
  if (auto& b = global_bpe()) return b->encode(text);
  std::vector<int32_t> ids;
  ids.reserve(text.size());
  for (unsigned char c : text) ids.push_back((int32_t)c);
  return ids;
}

// Incremental detokeniser: holds back an incomplete UTF-8 sequence until its last byte arrives so
// every streamed delta is valid UTF-8.
class Detok {
 public:
  std::string push(int32_t id) {
    if (auto& b = global_bpe()) {
      b->token_bytes(id, /*skip_special=*/true, pending_);
      return flush(false);
    }
    if (id >= 0 && id < 256) {
      pending_ += (char)(unsigned char)id;
      return flush(false);
    }
    std::string out = flush(true);
    out += "<|" + std::to_string(id) + "|>";
    return out;
  }
  std::string finish() { return flush(true); }

 private:
  std::string pending_;
  static int need(unsigned char c) { return c < 0x80 ? 1 : (c >> 5) == 0x6 ? 2 : (c >> 4) == 0xE ? 3 : (c >> 3) == 0x1E ? 4 : 0; }
  std::string flush(bool force) {
    std::string out;
    size_t i = 0;
    while (i < pending_.size()) {
      unsigned char c = (unsigned char)pending_[i];
      int n = need(c);
      if (n == 0) { out += "\xEF\xBF\xBD"; ++i; continue; }  // stray continuation byte
      if (i + n > pending_.size()) {
        if (!force) break;
        out += "\xEF\xBF\xBD";
        i = pending_.size();
        break;
      }
      bool good = true;
      for (int k = 1; k < n; ++k) if (((unsigned char)pending_[i + k] & 0xC0) != 0x80) good = false;
      if (good) { out.append(pending_, i, (size_t)n); i += (size_t)n; }
      else { out += "\xEF\xBF\xBD"; ++i; }
    }
    pending_.erase(0, i);
    return out;
  }
};

}  // namespace b2tok
