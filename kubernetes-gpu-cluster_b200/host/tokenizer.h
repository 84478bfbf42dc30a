// Byte-level BPE tokenizer reading HF `tokenizer.json` (the format of every model the reference's
// values files name: OPT = GPT-2 BPE, Llama-3, Qwen2.5 / Qwen3), so the server can take string
// prompts and chat messages like the engine it replaces (vllm:entrypoints/openai/completion/
// protocol.py:46-52 accepts str | [str] | [int] | [[int]]).
//
// Pre-tokenisation: the three regular expressions those tokenizers use are matched by hand-written
// scanners over code points (std::regex has no \p{L}); which one applies is read from the file.
// Encoding: added (special) tokens are split out first, each pre-token is mapped through GPT-2's
// byte<->unicode table and merged by rank; `ignore_merges` (Llama-3) is honoured.  Decoding maps back to
// bytes and only releases complete UTF-8 sequences (streaming).
#pragma once
#include <stdint.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include "json.h"
#include "nfc.h"
#include "unicode_tables.h"

namespace kge {

struct ChatMessage { std::string role, content; };

class Tokenizer {
 public:
  enum Pattern { GPT2, LLAMA3, QWEN };

  bool load(const std::string& dir, std::string* err) {
    std::ifstream f(dir + "/tokenizer.json", std::ios::binary);
    if (!f) return load_vocab_merges(dir, err);      // GPT-2 layout (facebook/opt-125m ships vocab.json + merges.txt only)
    std::stringstream ss;
    ss << f.rdbuf();
    Json j;
    std::string perr;
    if (!json_parse(ss.str(), &j, &perr)) { *err = "tokenizer.json: " + perr; return false; }
    const Json* model = j.get("model");
    if (!model || model->str_or("type", "BPE") != "BPE") { *err = "tokenizer.json: only BPE models are supported"; return false; }
    const Json* vocab = model->get("vocab");
    const Json* merges = model->get("merges");
    if (!vocab || !vocab->is_obj() || !merges || !merges->is_arr()) { *err = "tokenizer.json: vocab/merges missing"; return false; }
    ignore_merges_ = model->bool_or("ignore_merges", false);
    for (auto& kv : vocab->obj) {
      const int id = (int)kv.second.num;
      vocab_[kv.first] = id;
      if (id >= (int)id_to_tok_.size()) id_to_tok_.resize(id + 1);
      id_to_tok_[id] = kv.first;
    }
    int rank = 0;
    for (auto& m : merges->arr) {
      std::string a, b;
      if (m.is_arr() && m.arr.size() == 2) { a = m.arr[0].str; b = m.arr[1].str; }
      else if (m.is_str()) { size_t sp = m.str.find(' '); if (sp == std::string::npos) continue; a = m.str.substr(0, sp); b = m.str.substr(sp + 1); }
      else continue;
      ranks_[a + " " + b] = rank++;
    }
    if (const Json* at = j.get("added_tokens"))
      for (auto& t : at->arr) {
        const int id = (int)t.int_or("id", -1);
        const std::string c = t.str_or("content", "");
        if (id < 0 || c.empty()) continue;
        added_.push_back({c, id});
        if (t.bool_or("special", false)) special_ids_.push_back(id);
        if (id >= (int)id_to_tok_.size()) id_to_tok_.resize(id + 1);
        id_to_tok_[id] = c;
        added_ids_.push_back(id);
      }
    std::sort(added_.begin(), added_.end(), [](auto& a, auto& b) { return a.first.size() > b.first.size(); });
    // normalizer: NFC alone or inside a Sequence (other normalizers are not used by the supported families)
    if (const Json* nz = j.get("normalizer")) {
      if (nz->is_obj()) {
        if (nz->str_or("type", "") == "NFC") nfc_ = true;
        if (const Json* seq = nz->get("normalizers"))
          if (seq->is_arr()) for (auto& x : seq->arr) if (x.is_obj() && x.str_or("type", "") == "NFC") nfc_ = true;
      }
    }
    // which split pattern?
    pattern_ = GPT2;
    std::string dump = ss.str();
    if (dump.find("\\\\p{N}{1,3}") != std::string::npos || dump.find("\\p{N}{1,3}") != std::string::npos) pattern_ = LLAMA3;
    else if (dump.find("(?i:'s|'t|'re|'ve|'m|'ll|'d)") != std::string::npos) pattern_ = QWEN;
    // tokens a TemplateProcessing post-processor puts in front of a single sequence (BOS)
    if (const Json* pp = j.get("post_processor")) collect_prefix(*pp);
    build_byte_maps();
    // chat format
    if (vocab_.count("<|start_header_id|>") || has_added("<|start_header_id|>")) chat_ = "llama3";
    else if (has_added("<|im_start|>")) chat_ = "chatml";
    return true;
  }

  // vocab.json + merges.txt (+ tokenizer_config.json / special_tokens_map.json): the slow-tokenizer layout of GPT-2 style
  // models.  Byte-level BPE with the GPT-2 split pattern; bos/eos/pad/unk named in the config become special tokens and,
  // with add_bos_token (OPT), the bos id is put in front of every prompt - what GPT2TokenizerFast does with these files.
  bool load_vocab_merges(const std::string& dir, std::string* err) {
    std::ifstream fv(dir + "/vocab.json", std::ios::binary), fm(dir + "/merges.txt", std::ios::binary);
    if (!fv || !fm) { *err = "no tokenizer.json (or vocab.json + merges.txt) in " + dir; return false; }
    std::stringstream ss;
    ss << fv.rdbuf();
    Json v;
    std::string perr;
    if (!json_parse(ss.str(), &v, &perr) || !v.is_obj()) { *err = "vocab.json: " + perr; return false; }
    for (auto& kv : v.obj) {
      if (!kv.second.is_num()) continue;
      const int id = (int)kv.second.as_i64();
      if (id < 0) continue;
      vocab_[kv.first] = id;
      if (id >= (int)id_to_tok_.size()) id_to_tok_.resize(id + 1);
      id_to_tok_[id] = kv.first;
    }
    std::string line;
    int rank = 0;
    while (std::getline(fm, line)) {
      if (!line.empty() && line.back() == '\r') line.pop_back();
      if (line.empty() || line.compare(0, 9, "#version:") == 0) continue;
      const size_t sp = line.find(' ');
      if (sp == std::string::npos) continue;
      ranks_[line] = rank++;                          // "a b"
    }
    pattern_ = GPT2;
    auto special = [&](const Json& cfg, const char* key) -> int {
      const Json* t = cfg.get(key);
      std::string name;
      if (t && t->is_str()) name = t->str;
      else if (t && t->is_obj()) name = t->str_or("content", "");
      auto it = vocab_.find(name);
      if (name.empty() || it == vocab_.end()) return -1;
      if (std::find(added_ids_.begin(), added_ids_.end(), it->second) == added_ids_.end()) {
        added_.push_back({name, it->second});
        added_ids_.push_back(it->second);
        special_ids_.push_back(it->second);
      }
      return it->second;
    };
    for (const char* file : {"/special_tokens_map.json", "/tokenizer_config.json"}) {
      std::ifstream fc(dir + file, std::ios::binary);
      if (!fc) continue;
      std::stringstream sc;
      sc << fc.rdbuf();
      Json cfg;
      if (!json_parse(sc.str(), &cfg, &perr) || !cfg.is_obj()) continue;
      const int bos = special(cfg, "bos_token");
      special(cfg, "eos_token"); special(cfg, "pad_token"); special(cfg, "unk_token");
      if (cfg.bool_or("add_bos_token", false) && bos >= 0 && prefix_.empty()) prefix_.push_back(bos);
    }
    std::sort(added_.begin(), added_.end(), [](auto& a, auto& b) { return a.first.size() > b.first.size(); });
    build_byte_maps();
    return true;
  }

  std::vector<int32_t> encode(const std::string& text, bool add_special = true) const {
    std::vector<int32_t> out;
    if (add_special) out.insert(out.end(), prefix_.begin(), prefix_.end());
    size_t i = 0, plain0 = 0;
    while (i < text.size()) {
      int hit = -1;
      for (size_t a = 0; a < added_.size(); ++a)
        if (text.compare(i, added_[a].first.size(), added_[a].first) == 0) { hit = (int)a; break; }
      if (hit < 0) { ++i; continue; }
      encode_plain(text.substr(plain0, i - plain0), &out);
      out.push_back(added_[hit].second);
      i += added_[hit].first.size();
      plain0 = i;
    }
    encode_plain(text.substr(plain0), &out);
    return out;
  }

  // bytes of one token; special tokens decode to nothing (skip_special_tokens=True, vLLM's default)
  std::string token_bytes(int32_t id) const {
    if (id < 0 || id >= (int)id_to_tok_.size()) return "";
    if (std::find(special_ids_.begin(), special_ids_.end(), id) != special_ids_.end()) return "";
    if (std::find(added_ids_.begin(), added_ids_.end(), id) != added_ids_.end()) return id_to_tok_[id];
    std::string out;
    const std::string& t = id_to_tok_[id];
    for (size_t i = 0; i < t.size();) {
      uint32_t cp;
      i += decode_utf8(t, i, &cp);
      auto it = cp_to_byte_.find(cp);
      if (it != cp_to_byte_.end()) out += (char)it->second;
    }
    return out;
  }
  // Bytes -> valid UTF-8, replacing every maximal invalid subpart with U+FFFD exactly as Rust's
  // String::from_utf8_lossy does (what HF's ByteLevel decoder, and so vLLM's detokenizer, produce).
  // With final == false a sequence that is still a valid PREFIX at the end of the buffer is left in `*rest`.
  static std::string utf8_lossy(const std::string& b, bool final, std::string* rest) {
    std::string out;
    size_t i = 0;
    const size_t n = b.size();
    auto at = [&](size_t k) { return (unsigned char)b[k]; };
    while (i < n) {
      const unsigned char c = at(i);
      if (c < 0x80) { out += (char)c; ++i; continue; }
      size_t need = 0;
      unsigned char lo = 0x80, hi = 0xBF;                      // allowed range of the SECOND byte
      if (c >= 0xC2 && c <= 0xDF) need = 1;
      else if (c == 0xE0) { need = 2; lo = 0xA0; }
      else if ((c >= 0xE1 && c <= 0xEC) || c == 0xEE || c == 0xEF) need = 2;
      else if (c == 0xED) { need = 2; hi = 0x9F; }
      else if (c == 0xF0) { need = 3; lo = 0x90; }
      else if (c >= 0xF1 && c <= 0xF3) need = 3;
      else if (c == 0xF4) { need = 3; hi = 0x8F; }
      if (need == 0) { out += "\xEF\xBF\xBD"; ++i; continue; }   // not a lead byte
      size_t k = 1;
      bool ok = true, cut = false;
      for (; k <= need; ++k) {
        if (i + k >= n) { cut = true; break; }
        const unsigned char d = at(i + k);
        const unsigned char l2 = k == 1 ? lo : 0x80, h2 = k == 1 ? hi : 0xBF;
        if (d < l2 || d > h2) { ok = false; break; }
      }
      if (cut) {
        if (!final) { if (rest) *rest = b.substr(i); return out; }
        out += "\xEF\xBF\xBD";                                // truncated at the very end: one replacement
        i = n;
      } else if (ok) {
        out.append(b, i, need + 1);
        i += need + 1;
      } else {
        out += "\xEF\xBF\xBD";                                // lead + the valid continuation bytes seen so far
        i += k;
      }
    }
    if (rest) rest->clear();
    return out;
  }
  std::string decode(const std::vector<int32_t>& ids) const {
    std::string b;
    for (int32_t id : ids) b += token_bytes(id);
    return utf8_lossy(b, true, nullptr);
  }
  // Streaming: append one token's bytes to `pending`, return the text that is final so far (invalid bytes
  // already replaced), keep a possibly-incomplete sequence pending.  flush_stream() ends the stream.
  std::string decode_stream(int32_t id, std::string* pending) const {
    const std::string b = *pending + token_bytes(id);
    return utf8_lossy(b, false, pending);
  }
  static std::string flush_stream(std::string* pending) {
    std::string out = utf8_lossy(*pending, true, nullptr);
    pending->clear();
    return out;
  }

  bool has_chat_template() const { return !chat_.empty(); }
  // Llama-3 and ChatML (Qwen) conversation formats with the generation prompt appended.
  std::string render_chat(const std::vector<ChatMessage>& msgs) const {
    std::string s;
    if (chat_ == "llama3") {
      s = "<|begin_of_text|>";
      for (auto& m : msgs) s += "<|start_header_id|>" + m.role + "<|end_header_id|>\n\n" + m.content + "<|eot_id|>";
      s += "<|start_header_id|>assistant<|end_header_id|>\n\n";
    } else {
      for (auto& m : msgs) s += "<|im_start|>" + m.role + "\n" + m.content + "<|im_end|>\n";
      s += "<|im_start|>assistant\n";
    }
    return s;
  }
  int vocab_size() const { return (int)id_to_tok_.size(); }
  Pattern pattern() const { return pattern_; }

  // ---- pre-tokenisation (public for tests)
  std::vector<std::string> split(const std::string& text) const {
    std::vector<uint32_t> cp;
    std::vector<size_t> off;
    for (size_t i = 0; i < text.size();) { uint32_t c; off.push_back(i); i += decode_utf8(text, i, &c); cp.push_back(c); }
    off.push_back(text.size());
    std::vector<std::string> out;
    size_t i = 0;
    const size_t n = cp.size();
    while (i < n) {
      size_t e = match(cp, i);
      if (e <= i) e = i + 1;
      out.push_back(text.substr(off[i], off[e] - off[i]));
      i = e;
    }
    return out;
  }

 private:
  static size_t decode_utf8(const std::string& s, size_t i, uint32_t* cp) {
    const unsigned char c = (unsigned char)s[i];
    auto cont = [&](size_t k) { return i + k < s.size() && ((unsigned char)s[i + k] & 0xC0) == 0x80; };
    if (c < 0x80) { *cp = c; return 1; }
    if (c >= 0xF0 && cont(1) && cont(2) && cont(3)) { *cp = ((c & 7) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F); return 4; }
    if (c >= 0xE0 && c < 0xF0 && cont(1) && cont(2)) { *cp = ((c & 15) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F); return 3; }
    if (c >= 0xC0 && c < 0xE0 && cont(1)) { *cp = ((c & 31) << 6) | (s[i + 1] & 0x3F); return 2; }
    *cp = 0xFFFD;   // invalid byte: one replacement character per byte
    return 1;
  }
  static void append_utf8(std::string& o, uint32_t c) {
    if (c < 0x80) o += (char)c;
    else if (c < 0x800) { o += (char)(0xC0 | (c >> 6)); o += (char)(0x80 | (c & 0x3F)); }
    else if (c < 0x10000) { o += (char)(0xE0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
    else { o += (char)(0xF0 | (c >> 18)); o += (char)(0x80 | ((c >> 12) & 0x3F)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
  }
  template <size_t N>
  static bool in_ranges(const uint32_t (&r)[N][2], uint32_t c) {
    size_t lo = 0, hi = N;
    while (lo < hi) {
      const size_t mid = (lo + hi) / 2;
      if (c < r[mid][0]) hi = mid;
      else if (c > r[mid][1]) lo = mid + 1;
      else return true;
    }
    return false;
  }
  static bool is_L(uint32_t c) { return c < 0x80 ? ((c | 32) >= 'a' && (c | 32) <= 'z') : in_ranges(kLetterRanges, c); }
  static bool is_N(uint32_t c) { return c < 0x80 ? (c >= '0' && c <= '9') : in_ranges(kNumberRanges, c); }
  static bool is_S(uint32_t c) {
    for (uint32_t w : kWhiteSpace) if (w == c) return true;
    return false;
  }
  static bool is_nl(uint32_t c) { return c == '\r' || c == '\n'; }
  static bool is_other(uint32_t c) { return !is_S(c) && !is_L(c) && !is_N(c); }   // [^\s\p{L}\p{N}]

  // end of the match starting at i (ordered alternation, as in the tokenizers' regexes)
  size_t match(const std::vector<uint32_t>& t, size_t i) const {
    const size_t n = t.size();
    auto at = [&](size_t k) -> uint32_t { return k < n ? t[k] : 0; };
    // 1. contractions
    if (t[i] == '\'' && i + 1 < n) {
      const bool ci = pattern_ != GPT2;
      auto lc = [&](size_t k) -> uint32_t { uint32_t c = at(k); return (ci && c >= 'A' && c <= 'Z') ? c + 32 : c; };
      const uint32_t a = lc(i + 1), b2 = lc(i + 2);
      if (a == 's' || a == 't' || a == 'm' || a == 'd') return i + 2;
      if ((a == 'r' && b2 == 'e') || (a == 'v' && b2 == 'e') || (a == 'l' && b2 == 'l')) return i + 3;
    }
    if (pattern_ == GPT2) {
      size_t j = i + (t[i] == ' ' ? 1 : 0);
      if (j < n && is_L(t[j])) { while (j < n && is_L(t[j])) ++j; return j; }            //  ?\p{L}+
      if (j < n && is_N(t[j])) { while (j < n && is_N(t[j])) ++j; return j; }            //  ?\p{N}+
      if (j < n && is_other(t[j])) { while (j < n && is_other(t[j])) ++j; return j; }    //  ?[^\s\p{L}\p{N}]+
    } else {
      // [^\r\n\p{L}\p{N}]?\p{L}+
      if (!is_nl(t[i]) && !is_L(t[i]) && !is_N(t[i]) && i + 1 < n && is_L(t[i + 1])) { size_t j = i + 1; while (j < n && is_L(t[j])) ++j; return j; }
      if (is_L(t[i])) { size_t j = i; while (j < n && is_L(t[j])) ++j; return j; }
      // \p{N}{1,3} | \p{N}
      if (is_N(t[i])) { size_t j = i, lim = pattern_ == LLAMA3 ? 3 : 1; while (j < n && j - i < lim && is_N(t[j])) ++j; return j; }
      //  ?[^\s\p{L}\p{N}]+[\r\n]*
      {
        size_t j = i + (t[i] == ' ' ? 1 : 0);
        if (j < n && is_other(t[j])) { while (j < n && is_other(t[j])) ++j; while (j < n && is_nl(t[j])) ++j; return j; }
      }
      // \s*[\r\n]+ : through the last newline of the whitespace run
      if (is_S(t[i])) {
        size_t j = i, last_nl = 0; bool any = false;
        while (j < n && is_S(t[j])) { if (is_nl(t[j])) { last_nl = j; any = true; } ++j; }
        if (any) return last_nl + 1;
      }
    }
    // \s+(?!\S) | \s+
    if (is_S(t[i])) {
      size_t j = i;
      while (j < n && is_S(t[j])) ++j;
      if (j == n) return j;               // run to the end of the text
      if (j - i > 1) return j - 1;        // leave the last space for the following word
      return j;                           // single space before something that did not take it
    }
    return i + 1;
  }

  void build_byte_maps() {
    // GPT-2's bytes_to_unicode: printable bytes map to themselves, the rest to 256, 257, ...
    std::vector<int> bs;
    for (int b = '!'; b <= '~'; ++b) bs.push_back(b);
    for (int b = 0xA1; b <= 0xAC; ++b) bs.push_back(b);
    for (int b = 0xAE; b <= 0xFF; ++b) bs.push_back(b);
    std::vector<bool> have(256, false);
    for (int b : bs) have[b] = true;
    int nxt = 0;
    for (int b = 0; b < 256; ++b) {
      uint32_t cp = have[b] ? (uint32_t)b : (uint32_t)(256 + nxt++);
      std::string u;
      append_utf8(u, cp);
      byte_to_str_[b] = u;
      cp_to_byte_[cp] = (uint8_t)b;
    }
  }
  bool has_added(const std::string& c) const {
    for (auto& a : added_) if (a.first == c) return true;
    return false;
  }
  void collect_prefix(const Json& pp) {
    const std::string type = pp.str_or("type", "");
    if (type == "Sequence") {
      if (const Json* ps = pp.get("processors")) for (auto& p : ps->arr) collect_prefix(p);
    } else if (type == "TemplateProcessing") {
      if (const Json* single = pp.get("single"))
        for (auto& item : single->arr) {
          if (item.get("Sequence")) break;                 // only what precedes the text
          if (const Json* sp = item.get("SpecialToken")) {
            const std::string name = sp->str_or("id", "");
            auto it = vocab_.find(name);
            if (it != vocab_.end()) prefix_.push_back(it->second);
            else for (auto& a : added_) if (a.first == name) prefix_.push_back(a.second);
          }
        }
    } else if (type == "RobertaProcessing" || type == "BertProcessing") {
      if (const Json* cls = pp.get("cls")) if (cls->is_arr() && cls->arr.size() == 2) prefix_.push_back((int)cls->arr[1].num);
    }
  }
  void encode_plain(const std::string& raw, std::vector<int32_t>* out) const {
    if (raw.empty()) return;
    const std::string text = nfc_ ? nfc_utf8(raw) : raw;      // "normalizer": NFC (Qwen2.5 / Qwen3)
    for (const std::string& piece : split(text)) {
      std::vector<std::string> sym;
      std::string whole;
      for (unsigned char b : piece) { sym.push_back(byte_to_str_[b]); whole += byte_to_str_[b]; }
      if (ignore_merges_) {
        auto it = vocab_.find(whole);
        if (it != vocab_.end()) { out->push_back(it->second); continue; }
      }
      while (sym.size() > 1) {
        int best = INT32_MAX; size_t bi = 0;
        for (size_t i = 0; i + 1 < sym.size(); ++i) {
          auto it = ranks_.find(sym[i] + " " + sym[i + 1]);
          if (it != ranks_.end() && it->second < best) { best = it->second; bi = i; }
        }
        if (best == INT32_MAX) break;
        const std::string a = sym[bi], b2 = sym[bi + 1];
        std::vector<std::string> merged;
        for (size_t i = 0; i < sym.size();) {
          if (i + 1 < sym.size() && sym[i] == a && sym[i + 1] == b2) { merged.push_back(a + b2); i += 2; }
          else merged.push_back(sym[i++]);
        }
        sym.swap(merged);
      }
      for (auto& s : sym) {
        auto it = vocab_.find(s);
        if (it != vocab_.end()) out->push_back(it->second);
      }
    }
  }

  std::unordered_map<std::string, int> vocab_, ranks_;
  std::vector<std::string> id_to_tok_;
  std::vector<std::pair<std::string, int>> added_;
  std::vector<int> special_ids_, added_ids_, prefix_;
  std::string byte_to_str_[256];
  std::unordered_map<uint32_t, uint8_t> cp_to_byte_;
  bool ignore_merges_ = false;
  bool nfc_ = false;
  Pattern pattern_ = GPT2;
  std::string chat_;
};

}  // namespace kge
