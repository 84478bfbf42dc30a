"""Python mirror of vllm.LLM, host side only (the engine calls are stubbed): prompt forms (token ids, text, vLLM's dict
forms, a single prompt), the tokenizer behind the C ABI (kge_tokenizer_*) against HF `tokenizers`, decoded text in the
outputs, and the error when text arrives without a tokenizer.  CPU only."""
import pytest

tokenizers = pytest.importorskip("tokenizers")


class StubLLM:
    """LLM with the engine replaced by a script: every request emits token ids 5, 6, 7 and finishes by length."""

    def __new__(cls, kge, model):
        class _Stub(kge.LLM):
            def __init__(self, model):            # no engine
                self._model, self._tok, self._h, self._next_id, self._added, self._step = model, None, None, 0, [], 0

            def add_request(self, prompt_token_ids, sp):
                rid = self._next_id
                self._next_id += 1
                self._added.append(list(prompt_token_ids))
                return rid

            def has_unfinished(self):
                return self._step < 3

            def step(self):
                self._step += 1
                return [(rid, 4 + self._step, 2 if self._step == 3 else 0, 1) for rid in range(self._next_id)]

            def step_logprobs(self, rid):
                return None

            def close(self):
                pass
        return _Stub(model)


def test_tokenizer_c_abi_matches_hf(kge, tmp_path):
    from test_tokenizer_cpu import TEXTS, build
    for kind in ("gpt2", "llama3", "qwen"):
        d = tmp_path / kind
        d.mkdir()
        hf = build(kind, str(d))
        t = kge.Tokenizer(str(d))
        assert t.vocab_size == hf.get_vocab_size()
        for s in TEXTS:
            ids = t.encode(s)
            assert ids == hf.encode(s).ids, (kind, s)
            assert t.encode(s, add_special_tokens=False) == hf.encode(s, add_special_tokens=False).ids
            assert t.decode(ids) == hf.decode(ids, skip_special_tokens=True)
        assert t.encode("") == hf.encode("").ids and t.decode([]) == ""
        t.close()
    with pytest.raises(kge.KgeError, match="tokenizer"):
        kge.Tokenizer(str(tmp_path / "missing"))


def test_generate_accepts_vllm_prompt_forms(kge, tmp_path):
    from test_tokenizer_cpu import build
    hf = build("llama3", str(tmp_path))
    sp = kge.SamplingParams(max_tokens=3)
    on_step = lambda outs: None  # noqa: E731  (the stepwise path; the one-call path needs an engine)
    llm = StubLLM(kge, str(tmp_path))
    outs = llm.generate(["Hello, world!", [9, 8, 7], {"prompt": "hi"}, {"prompt_token_ids": [1, 2]}], sp, on_step=on_step)
    assert llm._added == [hf.encode("Hello, world!").ids, [9, 8, 7], hf.encode("hi").ids, [1, 2]]
    assert [o.outputs[0].token_ids for o in outs] == [[5, 6, 7]] * 4 and all(o.outputs[0].finish_reason == "length" for o in outs)
    assert outs[0].outputs[0].text == hf.decode([5, 6, 7], skip_special_tokens=True)
    assert outs[1].prompt_token_ids == [9, 8, 7]
    one = StubLLM(kge, str(tmp_path))
    assert len(one.generate("Hello", sp, on_step=on_step)) == 1 and one._added == [hf.encode("Hello").ids]
    ids = StubLLM(kge, str(tmp_path))
    assert len(ids.generate([3, 4, 5], sp, on_step=on_step)) == 1 and ids._added == [[3, 4, 5]]       # one prompt given as ids
    bare = StubLLM(kge, str(tmp_path / "nothing-here"))
    assert bare.generate([[3, 4]], sp, on_step=on_step)[0].outputs[0].text == ""                        # ids need no tokenizer
    with pytest.raises(kge.KgeError, match="tokenizer"):
        bare.generate(["text"], sp, on_step=on_step)
