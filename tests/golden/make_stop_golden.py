"""Generate tests/golden/stop_vllm.json: what vLLM's REAL incremental detokenizer (vllm/v1/engine/detokenizer.py of the
installed 0.22.0: FastIncrementalDetokenizer.update / get_next_output_text, check_stop_strings) does with stop strings,
min_tokens, include_stop_str_in_output and stop-token termination, token by token, on a small byte-level BPE tokenizer
(its tokenizer.json is stored in the fixture so the C++ side reads the same file).  Runs only in the build container.

  python tests/golden/make_stop_golden.py
"""
import json
import os
import random
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
import test_tokenizer_cpu as tt  # noqa: E402  (trains the tokenizer used by the tokenizer tests)

TEXTS = [
    "The quick brown fox jumps over the lazy dog. It's 12345 o'clock, isn't it? They're here; we'll see.",
    "def main(argc, argv):\n    for i in range(10):\n        print(i)\n\n\treturn 0\n",
    "Ünïcödé: café naïve Ελληνικά Русский 日本語 中文 — emoji 🙂🚀 and math ∑∫√ done.",
    "User: hello\nAssistant: hi there\nUser: bye\nAssistant: goodbye\n",
    "aaaaabaaaaabaaaaab END of it END",
]
STOPS = [["."], ["\n"], ["\n\n", "return"], ["User:"], ["é"], ["日本"], ["🙂"], ["END"], ["ab", "aab"], ["o'clock", "12345"], ["zzz"],
         ["lazy dog. It"], [" "], ["Assistant: goodbye\n"]]


def main():
    from transformers import PreTrainedTokenizerFast
    from vllm.sampling_params import SamplingParams
    from vllm.v1.engine import EngineCoreRequest
    from vllm.v1.engine.detokenizer import FastIncrementalDetokenizer
    d = tempfile.mkdtemp()
    tok = tt.build("qwen", d)
    tokenizer_json = open(os.path.join(d, "tokenizer.json")).read()
    hf = PreTrainedTokenizerFast(tokenizer_file=os.path.join(d, "tokenizer.json"))
    rng = random.Random(0)
    cases = []
    for text in TEXTS:
        ids = tok.encode(text, add_special_tokens=False).ids
        for stop in STOPS:
            if not any(s in text for s in stop) and rng.random() < 0.7:
                continue
            for include in (False, True):
                for min_tokens in (0, rng.choice([3, 8, 20])):
                    for stop_terminated in (False, True):
                        sp = SamplingParams(max_tokens=len(ids) + 32, min_tokens=min_tokens, stop=list(stop),
                                            include_stop_str_in_output=include, skip_special_tokens=True, temperature=0.0)
                        req = EngineCoreRequest(request_id="r", prompt_token_ids=[1], mm_features=None, sampling_params=sp,
                                                pooling_params=None, arrival_time=0.0, lora_request=None, cache_salt=None,
                                                data_parallel_rank=None)
                        det = FastIncrementalDetokenizer(hf, req)
                        deltas, matched, used = [], None, 0
                        for i, t in enumerate(ids):
                            last = i + 1 == len(ids)
                            m = det.update([t], last and stop_terminated)
                            used += 1
                            deltas.append(det.get_next_output_text(finished=(m is not None) or last, delta=True))
                            if m is not None:
                                matched = m
                                break
                        cases.append(dict(ids=ids, stop=stop, include_stop_str_in_output=include, min_tokens=min_tokens,
                                          stop_terminated=stop_terminated, text=det.output_text, deltas=deltas,
                                          matched=-1 if matched is None else stop.index(matched), tokens_used=used))
    with open(os.path.join(HERE, "stop_vllm.json"), "w") as f:
        json.dump(dict(tokenizer_json=tokenizer_json, cases=cases), f, separators=(",", ":"), ensure_ascii=True)
    print("stop_vllm.json:", len(cases), "cases,", sum(c["matched"] >= 0 for c in cases), "with a match")


if __name__ == "__main__":
    main()
