// Replays the recorded vLLM scheduler traces (tests/golden/sched_vllm.json, sched_prefix_vllm.json) through the C++
// Scheduler directly.  Built by tests/test_sanitizers_cpu.py with -fsanitize=address,undefined: the point is memory
// safety of the block pool (intrusive free list, hash index) and the request bookkeeping, on top of the parity the
// Python-driven tests already check.  usage: sched_replay FILE prefix_caching(0|1)
#include <stdio.h>
#include <stdlib.h>

#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "json.h"
#include "scheduler.h"

using namespace kge;

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  std::ifstream f(argv[1]);
  std::stringstream ss;
  ss << f.rdbuf();
  Json root;
  std::string err;
  if (!json_parse(ss.str(), &root, &err) || !root.is_arr()) { fprintf(stderr, "bad json: %s\n", err.c_str()); return 2; }
  const bool pc = atoi(argv[2]) != 0;
  long steps = 0;
  for (auto& c : root.arr) {
    const Json& k = *c.get("config");
    SchedConfig sc;
    sc.max_num_seqs = (int)k.int_or("max_num_seqs", 8);
    sc.max_num_batched_tokens = (int)k.int_or("max_num_batched_tokens", 64);
    sc.max_model_len = (int)k.int_or("max_model_len", 128);
    sc.block_size = (int)k.int_or("block_size", 16);
    sc.num_blocks = (int)k.int_or("num_blocks", 64);
    sc.reserve_full_isl = k.bool_or("reserve_full_isl", true);
    sc.enable_prefix_caching = pc;
    sc.eos_token_id = (int)c.int_or("eos", -1);
    Scheduler s(sc);
    const Json& reqs = *c.get("requests");
    auto add = [&](const Json& r) {
      std::vector<int32_t> p;
      for (auto& t : r.get("prompt")->arr) p.push_back((int32_t)t.num);
      SamplingParams sp;
      sp.max_tokens = (int)r.int_or("max_tokens", 16);
      sp.min_tokens = (int)r.int_or("min_tokens", 0);
      sp.ignore_eos = r.bool_or("ignore_eos", false);
      for (auto& t : r.get("stop")->arr) sp.stop_token_ids.push_back((int32_t)t.num);
      const char* why = nullptr;
      if (!s.add((int64_t)r.int_or("rid", 0), p.data(), (int32_t)p.size(), sp, &why)) { fprintf(stderr, "add: %s\n", why); exit(1); }
    };
    const bool has_arrivals = c.get("steps")->arr.size() && c.get("steps")->arr[0].get("arrivals");
    if (!has_arrivals) for (auto& r : reqs.arr) add(r);
    int si = 0;
    for (auto& st : c.get("steps")->arr) {
      if (has_arrivals)
        for (auto& a : st.get("arrivals")->arr)
          for (auto& r : reqs.arr) if ((int64_t)r.int_or("rid", -1) == (int64_t)a.num) add(r);
      const auto& sched = s.schedule();
      const auto& want = st.get("sched")->arr;
      if (sched.size() != want.size()) { fprintf(stderr, "seed %lld step %d: %zu scheduled, want %zu\n", (long long)c.int_or("seed", -1), si, sched.size(), want.size()); return 1; }
      for (size_t i = 0; i < sched.size(); ++i)
        if (sched[i].req->id != (int64_t)want[i].arr[0].num || sched[i].num_tokens != (int32_t)want[i].arr[1].num) {
          fprintf(stderr, "seed %lld step %d entry %zu differs\n", (long long)c.int_or("seed", -1), si, i);
          return 1;
        }
      std::vector<int32_t> toks;
      for (auto& t : st.get("tokens")->arr) toks.push_back((int32_t)t.arr[1].num);
      std::vector<StepResult> res;
      s.update(toks, &res);
      if (const Json* fb = st.get("free_blocks"))
        if (s.free_blocks() != (int32_t)fb->num) { fprintf(stderr, "seed %lld step %d: free %d want %d\n", (long long)c.int_or("seed", -1), si, s.free_blocks(), (int)fb->num); return 1; }
      ++si; ++steps;
    }
    if (s.has_unfinished() || s.free_blocks() != sc.num_blocks - 1) { fprintf(stderr, "seed %lld: not drained\n", (long long)c.int_or("seed", -1)); return 1; }
    // abort path: fill the scheduler again and abort everything mid-flight
    for (auto& r : reqs.arr) add(r);
    for (int i = 0; i < 3 && s.has_unfinished(); ++i) { s.schedule(); std::vector<int32_t> t(4096, 7); std::vector<StepResult> rr; s.update(t, &rr); }
    s.invalidate_prefix_cache();                 // blocks written behind the scheduler's back: forget every hash, keep running
    for (int i = 0; i < 3 && s.has_unfinished(); ++i) { s.schedule(); std::vector<int32_t> t(4096, 7); std::vector<StepResult> rr; s.update(t, &rr); }
    for (auto& r : reqs.arr) s.abort((int64_t)r.int_or("rid", 0));
    if (s.has_unfinished() || s.free_blocks() != sc.num_blocks - 1) { fprintf(stderr, "abort left state behind\n"); return 1; }
  }
  printf("ok %ld steps\n", steps);
  return 0;
}
