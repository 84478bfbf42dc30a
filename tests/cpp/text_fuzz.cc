// Tokenizer, NFC and stop-string code under sanitizers on hostile input: random bytes (invalid UTF-8 included), random
// token ids, long runs of combining marks.  Built by tests/test_sanitizers_cpu.py with -fsanitize=address,undefined.
// usage: text_fuzz TOKENIZER_DIR
#include <stdio.h>

#include <random>
#include <string>
#include <vector>

#include "nfc.h"
#include "stop_strings.h"
#include "tokenizer.h"

using namespace kge;

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  Tokenizer tok;
  std::string err;
  if (!tok.load(argv[1], &err)) { fprintf(stderr, "%s\n", err.c_str()); return 2; }
  std::mt19937 rng(1);
  size_t total = 0;
  for (int it = 0; it < 3000; ++it) {
    std::string s;
    const int n = (int)(rng() % 80);
    for (int i = 0; i < n; ++i) {
      const unsigned r = rng() % 10;
      if (r < 5) s += (char)(32 + rng() % 95);
      else if (r < 7) s += (char)(rng() % 256);                                     // arbitrary byte
      else if (r < 8) { s += (char)0xCC; s += (char)(0x80 + rng() % 0x40); }        // combining marks U+0300..U+033F
      else if (r < 9) { s += (char)0xE1; s += (char)0x84; s += (char)(0x80 + rng() % 0x13); }   // Hangul jamo
      else s += "<|im_end|>";
    }
    const std::string nf = nfc_utf8(s);
    const std::vector<int32_t> ids = tok.encode(s, it % 2 == 0);
    total += nf.size() + ids.size();
    const std::string back = tok.decode(ids);
    std::string pending, streamed;
    for (int32_t id : ids) streamed += tok.decode_stream(id, &pending);
    streamed += Tokenizer::flush_stream(&pending);
    if (streamed != back) { fprintf(stderr, "stream/decode mismatch at %d\n", it); return 1; }
    TextStream ts;
    ts.stop = {"ab", "\xCC\x81", "<|", std::string(1, (char)(32 + rng() % 95))};
    ts.include_stop = it % 3 == 0;
    ts.min_tokens = it % 5;
    std::string out;
    for (size_t i = 0; i < ids.size(); ++i) {
      const bool hit = ts.push(tok, (int32_t)(rng() % 4 ? ids[i] : (int32_t)(rng() % (tok.vocab_size() + 50)) - 10), false, i + 1 == ids.size());
      out += ts.next_delta(hit || i + 1 == ids.size());
      if (hit) break;
    }
    if (out != ts.text && !ids.empty()) { fprintf(stderr, "deltas do not add up to the text at %d\n", it); return 1; }
  }
  // JSON bodies as a client might send them, mutated: the parser must reject or accept, never crash; numbers convert
  // to integers without undefined behaviour
  const std::string sample = "{\"model\":\"m\",\"prompt\":[1,2,3,1e30,-1e30,0.5],\"max_tokens\":1e99,\"stop\":[\"a\",\"\\u00e9\\ud83d\\ude00\"],"
                             "\"stream_options\":{\"include_usage\":true},\"nested\":[[[[[[[[1]]]]]]]],\"seed\":-0,\"x\":null}";
  for (int it = 0; it < 20000; ++it) {
    std::string b = sample;
    const int muts = 1 + (int)(rng() % 4);
    for (int m = 0; m < muts; ++m) {
      const size_t pos = rng() % b.size();
      switch (rng() % 4) {
        case 0: b[pos] = (char)(rng() % 256); break;
        case 1: b.erase(pos, 1 + rng() % 3); break;
        case 2: b.insert(pos, std::string(1 + rng() % 200, "[{\"\\"[rng() % 4])); break;
        default: b.insert(pos, std::to_string((double)rng() * 1e300)); break;
      }
      if (b.empty()) b = "x";
    }
    Json j;
    std::string e;
    if (json_parse(b, &j, &e) && j.is_obj()) {
      total += (size_t)(j.int_or("max_tokens", 1) & 1) + (size_t)(j.int_or("seed", 0) & 1);
      if (const Json* p = j.get("prompt")) if (p->is_arr()) for (auto& t : p->arr) total += (size_t)(t.as_i64() & 1);
      total += json_escape(j.str_or("model", "")).size();
    }
  }
  printf("ok %zu\n", total);
  return 0;
}
