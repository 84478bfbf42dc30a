// Incremental output text of one request: detokenise token by token, match stop strings, hold back the tail
// that could still become a stop string.  Follows vllm:v1/engine/detokenizer.py:95-164 (update,
// get_next_output_text) and :309-344 (check_stop_strings): lengths there are in characters, so everything
// here that counts text counts UTF-8 code points, not bytes.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "tokenizer.h"

namespace kge {

inline size_t utf8_chars(const std::string& s) {
  size_t n = 0;
  for (unsigned char c : s) n += (c & 0xC0) != 0x80;
  return n;
}
// byte offset reached by walking `n` code points back from byte offset `from`
inline size_t utf8_back(const std::string& s, size_t from, size_t n) {
  size_t i = from;
  while (n > 0 && i > 0) {
    --i;
    if ((static_cast<unsigned char>(s[i]) & 0xC0) != 0x80) --n;
  }
  return i;
}

struct TextStream {
  std::vector<std::string> stop;
  bool include_stop = false;
  int min_tokens = 0;

  std::string text;          // output text so far (truncated at a matched stop string)
  std::string pending;       // bytes of an unfinished UTF-8 sequence
  size_t emitted = 0;        // bytes of `text` already handed out by next_delta()
  int num_tokens = 0;
  int matched = -1;          // index into `stop` once a stop string was found

  // One new token.  `stop_terminated`: the engine finished the request on this token because it is EOS or a stop
  // token id; its text is then left out unless include_stop (detokenizer.py:107-113).  `last`: no token follows.
  // Returns true when a stop string matched (the caller aborts the request).
  bool push(const Tokenizer& tok, int32_t id, bool stop_terminated, bool last) {
    if (matched >= 0) return true;
    size_t check_from = text.size();
    ++num_tokens;
    if (!(stop_terminated && !include_stop)) text += tok.decode_stream(id, &pending);
    if (last) text += Tokenizer::flush_stream(&pending);
    if (min_tokens && num_tokens <= min_tokens) check_from = text.size();
    if (stop.empty() || num_tokens <= min_tokens || text.size() == check_from) return false;
    for (size_t i = 0; i < stop.size(); ++i) {              // first stop string IN LIST ORDER that occurs
      const std::string& st = stop[i];
      if (st.empty()) continue;
      const size_t n = utf8_chars(st);
      // str.find(stop, 1 - new_chars - len(stop)): start len(stop)-1 characters before the new text
      const size_t start = utf8_back(text, check_from, n - 1);
      const size_t pos = text.find(st, start);
      if (pos == std::string::npos) continue;
      matched = (int)i;
      text.resize(include_stop ? pos + st.size() : pos);
      if (emitted > text.size()) emitted = text.size();   // cannot happen with the hold-back; kept as a guard
      return true;
    }
    return false;
  }

  // Text not yet handed out.  While the request runs, the last (longest stop string - 1) characters stay back.
  std::string next_delta(bool finished) {
    size_t hold = 0;
    if (!finished)
      for (auto& s : stop) { const size_t n = utf8_chars(s); if (n > 0 && n - 1 > hold) hold = n - 1; }
    const size_t end = utf8_back(text, text.size(), hold);
    if (end <= emitted) return "";
    std::string d = text.substr(emitted, end - emitted);
    emitted = end;
    return d;
  }
};

}  // namespace kge
