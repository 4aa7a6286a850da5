// tokenizer.hpp -- byte-level fallback tokenizer.  No tokenizer files exist offline (the reference
// downloads Qwen/Qwen3-0.6B with an HF token, llm-d-deploy.yaml:121-132,184), so text prompts map one
// UTF-8 byte to one id in [0, 256) and generated ids map back the same way; ids outside the byte range
// (any id a random-init model emits) render as "<|id|>".  Token-id prompts (`"prompt": [1,2,3]`) bypass
// this entirely, which is what the benchmark uses.
#pragma once
#include <string>
#include <vector>

namespace b2tok {

inline std::vector<int32_t> encode(const std::string& text) {
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
