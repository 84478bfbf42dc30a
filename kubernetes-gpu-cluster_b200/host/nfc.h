// Unicode Normalization Form C for the tokenizer's "normalizer": {"type": "NFC"} (Qwen2.5 / Qwen3 tokenizer.json):
// canonical decomposition, canonical ordering, canonical composition (UAX #15), Hangul handled arithmetically.
// Tables come from tools/gen_unicode.py (host/unicode_nfc.h).
#pragma once
#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

#include "unicode_nfc.h"

namespace kge {

inline int nfc_ccc(uint32_t cp) {
  if (cp < 0x300) return 0;
  size_t lo = 0, hi = sizeof(kNfcCcc) / sizeof(kNfcCcc[0]);
  while (lo < hi) {
    const size_t mid = (lo + hi) / 2;
    if (cp < kNfcCcc[mid][0]) hi = mid;
    else if (cp > kNfcCcc[mid][1]) lo = mid + 1;
    else return (int)kNfcCcc[mid][2];
  }
  return 0;
}

inline void nfc_decompose(uint32_t cp, std::vector<uint32_t>* out) {
  if (cp >= 0xAC00 && cp <= 0xD7A3) {                       // Hangul syllable -> L V (T)
    const uint32_t s = cp - 0xAC00;
    out->push_back(0x1100 + s / 588);
    out->push_back(0x1161 + (s % 588) / 28);
    if (s % 28) out->push_back(0x11A7 + s % 28);
    return;
  }
  if (cp >= 0xC0) {
    size_t lo = 0, hi = sizeof(kNfcDecomp) / sizeof(kNfcDecomp[0]);
    while (lo < hi) {
      const size_t mid = (lo + hi) / 2;
      if (kNfcDecomp[mid][0] < cp) lo = mid + 1; else hi = mid;
    }
    if (lo < sizeof(kNfcDecomp) / sizeof(kNfcDecomp[0]) && kNfcDecomp[lo][0] == cp) {
      for (int k = 1; k <= 4 && kNfcDecomp[lo][k]; ++k) out->push_back(kNfcDecomp[lo][k]);
      return;
    }
  }
  out->push_back(cp);
}

inline uint32_t nfc_compose_pair(uint32_t a, uint32_t b) {
  if (a >= 0x1100 && a < 0x1113 && b >= 0x1161 && b < 0x1176) return 0xAC00 + ((a - 0x1100) * 21 + (b - 0x1161)) * 28;   // L + V
  if (a >= 0xAC00 && a <= 0xD7A3 && (a - 0xAC00) % 28 == 0 && b > 0x11A7 && b < 0x11C3) return a + (b - 0x11A7);        // LV + T
  size_t lo = 0, hi = sizeof(kNfcCompose) / sizeof(kNfcCompose[0]);
  while (lo < hi) {
    const size_t mid = (lo + hi) / 2;
    if (kNfcCompose[mid][0] < a || (kNfcCompose[mid][0] == a && kNfcCompose[mid][1] < b)) lo = mid + 1; else hi = mid;
  }
  if (lo < sizeof(kNfcCompose) / sizeof(kNfcCompose[0]) && kNfcCompose[lo][0] == a && kNfcCompose[lo][1] == b) return kNfcCompose[lo][2];
  return 0;
}

// Valid UTF-8 in, NFC UTF-8 out.  Invalid bytes are passed through untouched (each as its own unit).
inline std::string nfc_utf8(const std::string& in) {
  bool plain = true;                                         // nothing below U+0300 changes under NFC
  for (unsigned char c : in) if (c >= 0xCC) { plain = false; break; }
  if (plain) return in;
  std::vector<uint32_t> cps;                                 // invalid bytes travel as 0x110000 + byte
  for (size_t i = 0; i < in.size();) {
    const unsigned char c = (unsigned char)in[i];
    int n = c < 0x80 ? 1 : (c >> 5) == 6 ? 2 : (c >> 4) == 14 ? 3 : (c >> 3) == 30 ? 4 : 0;
    uint32_t cp = 0;
    bool ok = n > 0 && i + n <= in.size();
    if (ok) {
      cp = n == 1 ? c : c & (0xFF >> (n + 1));
      for (int k = 1; k < n; ++k) {
        const unsigned char d = (unsigned char)in[i + k];
        if ((d & 0xC0) != 0x80) { ok = false; break; }
        cp = (cp << 6) | (d & 0x3F);
      }
    }
    if (!ok) { cps.push_back(0x110000 + c); ++i; continue; }
    nfc_decompose(cp, &cps);
    i += n;
  }
  // canonical ordering: stable sort of each run of non-starters by combining class
  for (size_t i = 0; i < cps.size();) {
    if (cps[i] >= 0x110000 || nfc_ccc(cps[i]) == 0) { ++i; continue; }
    size_t j = i;
    while (j < cps.size() && cps[j] < 0x110000 && nfc_ccc(cps[j]) != 0) ++j;
    std::stable_sort(cps.begin() + i, cps.begin() + j, [](uint32_t a, uint32_t b) { return nfc_ccc(a) < nfc_ccc(b); });
    i = j;
  }
  // canonical composition: a character combines with the last starter unless something between them has a
  // combining class that is zero or not lower than its own (UAX #15, "blocked")
  std::vector<uint32_t> out;
  int starter = -1;
  for (uint32_t cp : cps) {
    const bool raw = cp >= 0x110000;
    const int cc = raw ? 0 : nfc_ccc(cp);
    if (starter >= 0 && !raw) {
      const bool blocked = (int)out.size() - 1 > starter && nfc_ccc(out.back()) >= cc;
      if (!blocked) {
        const uint32_t c2 = nfc_compose_pair(out[starter], cp);
        if (c2) { out[starter] = c2; continue; }
      }
    }
    out.push_back(cp);
    if (raw) starter = -1;
    else if (cc == 0) starter = (int)out.size() - 1;
  }
  std::string s;
  for (uint32_t cp : out) {
    if (cp >= 0x110000) s += (char)(cp - 0x110000);
    else if (cp < 0x80) s += (char)cp;
    else if (cp < 0x800) { s += (char)(0xC0 | (cp >> 6)); s += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { s += (char)(0xE0 | (cp >> 12)); s += (char)(0x80 | ((cp >> 6) & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
    else { s += (char)(0xF0 | (cp >> 18)); s += (char)(0x80 | ((cp >> 12) & 0x3F)); s += (char)(0x80 | ((cp >> 6) & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
  }
  return s;
}

}  // namespace kge
