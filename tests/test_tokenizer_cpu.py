"""The server's C++ tokenizer (host/tokenizer.h, driven through `vllm tokenize DIR`) against the HF
`tokenizers` library on tokenizer.json files of the three families the reference's values files name:
GPT-2 byte-level BPE (OPT), Llama-3 and Qwen2.5/Qwen3 (their published pre-tokenisation regexes).
The tokenizers are trained here on a small corpus (no tokenizer files exist offline); what is under test
is the tokenizer.json reader, the pre-tokenisation scanners, the BPE merge loop, added/special tokens, the
BOS template, decoding and streaming decode.  CPU only."""
import json
import os
import subprocess

import pytest

tokenizers = pytest.importorskip("tokenizers")
from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, processors, trainers  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kubernetes-gpu-cluster_b200", "bin", "vllm")
pytestmark = pytest.mark.skipif(not os.path.exists(BIN), reason="server binary not built")

LLAMA3_RE = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")
QWEN_RE = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")

CORPUS = [
    "The quick brown fox jumps over the lazy dog. It's 12345 o'clock, isn't it? They're here; we'll see, I'd say.",
    "def main(argc, argv):\n    for i in range(10):\n        print(f\"{i:03d} -> {i*i}\")\n\n\treturn 0\n",
    "Kubernetes GPU cluster: kubeadm init --control-plane-endpoint 192.168.1.10:6443 --upload-certs",
    "Prices: $1,234.56 + 7% = ?  Multiple   spaces\tand\ttabs.  Trailing space \n\n\nNew paragraph!",
    "Ünïcödé: café naïve Ελληνικά Русский 日本語 中文 한국어 — emoji 🙂🚀 and math ∑∫√",
    "I'M SHOUTING AND YOU'LL HEAR IT. we've they'd he's",
] * 4

TEXTS = [
    "Hello, world!", "  leading spaces", "trailing spaces   ", "a  b   c    d", "it's, IT'S, they'RE and o'clock",
    "line1\nline2\r\n\r\nline3 \n \n end", "tabs\t\tand\nnewlines\n", "1234567 12 3.14159 1,000,000", "x=1;y=2;//comment!!!\n",
    "café Ελληνικά Русский 日本語 🙂", "", " ", "\n", "a", "...", "snake_case and camelCase and kebab-case", "100%  sure?!  (maybe)",
    "The GPU cluster's nodes aren't ready yet.\n\nkubectl get nodes", "<|special|> in text", "mixed123abc456 7up",
]


def build(kind, d):
    tok = Tokenizer(models.BPE(ignore_merges=(kind == "llama3")))
    if kind == "gpt2":
        tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=True)
        specials = ["</s>", "<pad>"]
    else:
        tok.pre_tokenizer = pre_tokenizers.Sequence([
            pre_tokenizers.Split(Regex(LLAMA3_RE if kind == "llama3" else QWEN_RE), behavior="isolated", invert=False),
            pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
        specials = (["<|begin_of_text|>", "<|end_of_text|>", "<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"]
                    if kind == "llama3" else ["<|endoftext|>", "<|im_start|>", "<|im_end|>"])
    tok.decoder = decoders.ByteLevel()
    tr = trainers.BpeTrainer(vocab_size=700, special_tokens=specials, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                             show_progress=False)
    tok.train_from_iterator(CORPUS, tr)
    if kind in ("gpt2", "llama3"):      # BOS in front of a single sequence, as OPT's and Llama-3's tokenizers do
        bos = specials[0]
        tok.post_processor = processors.TemplateProcessing(single=f"{bos} $A", special_tokens=[(bos, tok.token_to_id(bos))])
    tok.save(os.path.join(d, "tokenizer.json"))
    return tok


def run_cli(d, objs):
    inp = "".join(json.dumps(o) + "\n" for o in objs)
    p = subprocess.run([BIN, "tokenize", d], input=inp, capture_output=True, text=True, timeout=60)
    assert p.returncode == 0, p.stderr
    return [json.loads(line) for line in p.stdout.splitlines()]


@pytest.mark.parametrize("kind", ["gpt2", "llama3", "qwen"])
def test_encode_decode_match_hf_tokenizers(kind, tmp_path):
    tok = build(kind, str(tmp_path))
    outs = run_cli(str(tmp_path), [{"text": t} for t in TEXTS])
    for t, o in zip(TEXTS, outs):
        ref = tok.encode(t)
        assert o["ids"] == ref.ids, (kind, t, o["pieces"], [tok.id_to_token(i) for i in ref.ids])
    # without the BOS template
    outs = run_cli(str(tmp_path), [{"text": t, "add_special_tokens": False} for t in TEXTS[:6]])
    for t, o in zip(TEXTS, outs):
        assert o["ids"] == tok.encode(t, add_special_tokens=False).ids
    # decode (special tokens skipped) and streaming decode give the text back
    id_lists = [tok.encode(t).ids for t in TEXTS]
    outs = run_cli(str(tmp_path), [{"ids": ids} for ids in id_lists])
    for t, ids, o in zip(TEXTS, id_lists, outs):
        assert o["text"] == tok.decode(ids, skip_special_tokens=True) == o["streamed"], (kind, t)
        if "<|special|>" not in t:
            assert o["text"] == t


def test_lossy_decoding_of_arbitrary_id_sequences(tmp_path):
    """Random ids decode to arbitrary bytes; invalid UTF-8 must become U+FFFD exactly as in HF's (Rust's lossy)
    decoding, and streamed pieces must concatenate to the same text."""
    import random
    tok = build("qwen", str(tmp_path))
    rng = random.Random(0)
    seqs = [[rng.randrange(3, tok.get_vocab_size()) for _ in range(rng.randrange(1, 40))] for _ in range(200)]
    seqs += [tok.encode("日本語 🙂 café").ids[:k] for k in range(1, 12)]        # truncated multi-byte sequences
    outs = run_cli(str(tmp_path), [{"ids": s} for s in seqs])
    for s, o in zip(seqs, outs):
        ref = tok.decode(s, skip_special_tokens=True)
        assert o["text"] == ref, s
        assert o["streamed"] == ref, s


def test_added_tokens_are_matched_in_text(tmp_path):
    tok = build("qwen", str(tmp_path))
    t = "<|im_start|>user\nhi<|im_end|>\n"
    o = run_cli(str(tmp_path), [{"text": t}])[0]
    assert o["ids"] == tok.encode(t).ids
    assert tok.token_to_id("<|im_start|>") in o["ids"] and tok.token_to_id("<|im_end|>") in o["ids"]


def test_chat_formats(tmp_path):
    msgs = [{"role": "system", "content": "Be brief."}, {"role": "user", "content": "Hi there"}]
    d1 = tmp_path / "l3"
    d1.mkdir()
    tok = build("llama3", str(d1))
    o = run_cli(str(d1), [{"messages": msgs}])[0]
    assert o["text"] == ("<|begin_of_text|><|start_header_id|>system<|end_header_id|>\n\nBe brief.<|eot_id|>"
                         "<|start_header_id|>user<|end_header_id|>\n\nHi there<|eot_id|>"
                         "<|start_header_id|>assistant<|end_header_id|>\n\n")
    assert o["ids"] == tok.encode(o["text"], add_special_tokens=False).ids
    d2 = tmp_path / "qw"
    d2.mkdir()
    tok = build("qwen", str(d2))
    o = run_cli(str(d2), [{"messages": msgs}])[0]
    assert o["text"] == "<|im_start|>system\nBe brief.<|im_end|>\n<|im_start|>user\nHi there<|im_end|>\n<|im_start|>assistant\n"
    assert o["ids"] == tok.encode(o["text"], add_special_tokens=False).ids
