// C view of the byte-level BPE tokenizer (host/tokenizer.h) so that front ends other than the C++ server - the Python
// mirror of vllm.LLM first of all - can take text prompts and return text, as the engine they replace does.  Host only.
#include <string.h>

#include <string>
#include <vector>

#include "../host/tokenizer.h"
#include "common.cuh"
#include "kge.h"

struct kge_tokenizer { kge::Tokenizer tok; };

extern "C" kge_tokenizer* kge_tokenizer_open(const char* model_dir) {
  if (!model_dir) { kge::set_error("kge_tokenizer_open: null directory"); return nullptr; }
  auto* t = new kge_tokenizer();
  std::string err;
  if (!t->tok.load(model_dir, &err)) { kge::set_error("kge_tokenizer_open: %s", err.c_str()); delete t; return nullptr; }
  return t;
}
extern "C" void kge_tokenizer_close(kge_tokenizer* t) { delete t; }
extern "C" int kge_tokenizer_vocab_size(const kge_tokenizer* t) { return t ? t->tok.vocab_size() : 0; }

extern "C" int64_t kge_tokenizer_encode(const kge_tokenizer* t, const char* text, int64_t text_bytes, int add_special_tokens,
                                        int32_t* out_ids, int64_t cap) {
  if (!t || !text || text_bytes < 0) { kge::set_error("kge_tokenizer_encode: bad arguments"); return -1; }
  const std::vector<int32_t> ids = t->tok.encode(std::string(text, (size_t)text_bytes), add_special_tokens != 0);
  if (out_ids && (int64_t)ids.size() <= cap) memcpy(out_ids, ids.data(), ids.size() * sizeof(int32_t));
  return (int64_t)ids.size();          // call with cap = 0 to size the buffer
}

extern "C" int64_t kge_tokenizer_decode(const kge_tokenizer* t, const int32_t* ids, int64_t n, char* out, int64_t cap) {
  if (!t || (!ids && n > 0) || n < 0) { kge::set_error("kge_tokenizer_decode: bad arguments"); return -1; }
  const std::string s = t->tok.decode(std::vector<int32_t>(ids, ids + n));
  if (out && (int64_t)s.size() <= cap) memcpy(out, s.data(), s.size());
  return (int64_t)s.size();            // bytes of UTF-8 (no terminator); call with cap = 0 to size the buffer
}
