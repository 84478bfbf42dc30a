// Error channel, version/device queries and the host-side RoPE cache builder of the C ABI.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <vector>

#include "common.cuh"
#include "half_bits.h"
#include "kge.h"

namespace kge {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static std::atomic<int> g_pdl{-1};   // -1: not decided yet
bool pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) { v = getenv("KGE_DISABLE_PDL") ? 0 : 1; g_pdl.store(v, std::memory_order_relaxed); }
  return v != 0;
}
void pdl_set_enabled(bool on) { g_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }
static thread_local bool g_pdl_break = false;
void pdl_break_next() { g_pdl_break = true; }
bool pdl_for_this_launch() {
  const bool brk = g_pdl_break;
  g_pdl_break = false;
  return pdl_enabled() && !brk;
}

static std::atomic<int> g_tun[TUN_COUNT] = {{1}, {0}, {0}, {0}, {1}, {0}, {1}, {0}, {1}, {1}};   // even grid measured +6.5 % on the C3 gate_up GEMM (profiles/r01_knob_sweep.md)
static const char* const kTunNames[TUN_COUNT] = {"gemm_even_grid", "attn_decode_splits", "attn_split_balance", "attn_cluster_merge", "attn_tma", "gemm_ws", "smem_carveout", "gemm_l2_hints", "attn_persistent", "gemm_wide"};
int tunable(Tunable t) { return g_tun[t].load(std::memory_order_relaxed); }
}  // namespace kge

extern "C" int kge_set_tunable(const char* name, int value) {
  KGE_REQUIRE(name, "kge_set_tunable: null name");
  if (!strcmp(name, "pdl")) { kge::pdl_set_enabled(value != 0); return 0; }
  for (int i = 0; i < kge::TUN_COUNT; ++i)
    if (!strcmp(name, kge::kTunNames[i])) { kge::g_tun[i].store(value, std::memory_order_relaxed); return 0; }
  kge::set_error("kge_set_tunable: unknown knob '%s' (pdl, gemm_even_grid, attn_decode_splits, attn_split_balance, attn_cluster_merge, attn_tma, gemm_ws, smem_carveout, gemm_l2_hints, attn_persistent, gemm_wide)", name);
  return -1;
}

extern "C" const char* kge_last_error(void) { return kge::g_err; }
extern "C" int kge_abi_version(void) { return KGE_ABI_VERSION; }

extern "C" int kge_device_sm_count(void) {
  int n = 0, dev = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { (void)cudaGetLastError(); return 0; }
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  return sms;
}

static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

// vllm:model_executor/layers/rotary_embedding/base.py:69-93 (fp32 inv_freq, fp32 outer product,
// fp32 cos/sin, cast to the model dtype) and llama3_rope.py:33-54 (frequency rescale).
static int rope_cache_host(uint16_t* out_host, int max_pos, int hd, double theta, double factor,
                           double low_freq_factor, double high_freq_factor, int orig_max_pos, bool f16) {
  KGE_REQUIRE(out_host && max_pos > 0 && hd > 0 && hd % 2 == 0, "kge_rope_cache_host: bad arguments");
  const int half = hd / 2;
  auto to16 = [f16](float v) { return f16 ? kge::f32_to_f16_bits(v) : f32_to_bf16_rne(v); };
  std::vector<float> inv(half);
  for (int i = 0; i < half; ++i) {
    const float e = (float)(2 * i) / (float)hd;               // arange(0,rd,2,float)/rd
    const float pw = powf((float)theta, e);                   // base ** e in fp32
    float f = 1.0f / pw;
    if (factor != 0.0) {
      const float low_wl = (float)orig_max_pos / (float)low_freq_factor;
      const float high_wl = (float)orig_max_pos / (float)high_freq_factor;
      const float wl = (float)(2.0 * M_PI) / f;               // python float * fp32 tensor -> fp32
      if (wl < high_wl) {
        // keep
      } else if (wl > low_wl) {
        f = f / (float)factor;
      } else {
        float smooth = 0.f;
        if (low_freq_factor != high_freq_factor)
          smooth = ((float)orig_max_pos / wl - (float)low_freq_factor) /
                   (float)(high_freq_factor - low_freq_factor);
        f = (1.0f - smooth) * f / (float)factor + smooth * f;
      }
    }
    inv[i] = f;
  }
  for (int p = 0; p < max_pos; ++p) {
    for (int i = 0; i < half; ++i) {
      const float fr = (float)p * inv[i];
      // correctly rounded fp32 cos/sin of the fp32 angle
      out_host[(size_t)p * hd + i] = to16((float)cos((double)fr));
      out_host[(size_t)p * hd + half + i] = to16((float)sin((double)fr));
    }
  }
  return 0;
}

extern "C" int kge_rope_cache_host(uint16_t* out_host, int max_pos, int hd, double theta, double factor,
                                   double low_freq_factor, double high_freq_factor, int orig_max_pos) {
  return rope_cache_host(out_host, max_pos, hd, theta, factor, low_freq_factor, high_freq_factor, orig_max_pos, false);
}
extern "C" int kge_rope_cache_host_f16(uint16_t* out_host, int max_pos, int hd, double theta, double factor,
                                       double low_freq_factor, double high_freq_factor, int orig_max_pos) {
  return rope_cache_host(out_host, max_pos, hd, theta, factor, low_freq_factor, high_freq_factor, orig_max_pos, true);
}
