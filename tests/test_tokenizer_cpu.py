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
from tokenizers import Regex, Tokenizer, decoders, models, normalizers, pre_tokenizers, processors, trainers  # noqa: E402

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


def build(kind, d, nfc=False):
    tok = Tokenizer(models.BPE(ignore_merges=(kind == "llama3")))
    if nfc:
        tok.normalizer = normalizers.NFC()          # what Qwen2.5 / Qwen3 tokenizer.json carries
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
    return [json.loads(line) for line in p.stdout.split("\n") if line]


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


def test_nfc_normalizer_like_qwen(tmp_path):
    """Qwen's tokenizer.json has "normalizer": NFC.  The C++ normaliser must agree with unicodedata on every code point,
    on decomposed / reordered mark sequences and Hangul, and the tokenizer must apply it before pre-tokenisation."""
    import random
    import unicodedata
    tok = build("qwen", str(tmp_path), nfc=True)
    rng = random.Random(1)
    singles = ["".join(chr(c) for c in range(a, min(a + 64, 0x110000)) if not 0xD800 <= c <= 0xDFFF) for a in range(0, 0x110000, 64)]
    pool = [chr(c) for c in list(range(0x20, 0x7F)) + list(range(0xC0, 0x250)) + list(range(0x300, 0x370)) + list(range(0x1E00, 0x1F00))
            + list(range(0x1100, 0x1200)) + list(range(0xAC00, 0xAC80)) + [0x212B, 0x2126, 0xFB1D, 0x958, 0xF43, 0x1D15E, 0x2ADC, 0x344, 0x3099,
                                                                            0x304B, 0xB47, 0xB3E, 0xB56, 0xDD9, 0xDCA, 0xDCF, 0x1B05, 0x1B35, 0x11099, 0x110BA]]
    mixes = ["".join(rng.choice(pool) for _ in range(rng.randrange(1, 24))) for _ in range(3000)]
    texts = [t for t in singles if t] + mixes + [unicodedata.normalize("NFD", t) for t in mixes[:1000]]
    outs = run_cli(str(tmp_path), [{"nfc": t} for t in texts])
    for t, o in zip(texts, outs):
        assert o["text"] == unicodedata.normalize("NFC", t), [hex(ord(c)) for c in t][:12]
    samples = ["cafe\u0301 nai\u0308ve", unicodedata.normalize("NFD", "Ünïcödé Ελληνικά 한국어 が ぱ"), "A\u030a \u212b \u1100\u1161\u11a8", "plain ascii"]
    outs = run_cli(str(tmp_path), [{"text": t} for t in samples])
    for t, o in zip(samples, outs):
        assert o["ids"] == tok.encode(t).ids, t
    plain = build("qwen", str(tmp_path / "x") if (tmp_path / "x").mkdir() is None else "", nfc=False)
    assert plain.encode(samples[0]).ids != tok.encode(samples[0]).ids      # the normaliser matters for this input


def test_gpt2_layout_vocab_json_and_merges_txt(tmp_path):
    """facebook/opt-125m (the reference's first values file) ships vocab.json + merges.txt, no tokenizer.json: the server
    must read that layout the way transformers' GPT-2 tokenizer does, OPT's leading </s> included."""
    transformers = pytest.importorskip("transformers")
    src = tmp_path / "src"
    src.mkdir()
    tok = build("gpt2", str(src))
    d = tmp_path / "opt"
    d.mkdir()
    tok.model.save(str(d))                                   # vocab.json + merges.txt
    (d / "tokenizer_config.json").write_text(json.dumps({"bos_token": "</s>", "eos_token": "</s>", "pad_token": "<pad>", "unk_token": "</s>",
                                                         "add_bos_token": True, "add_prefix_space": False, "tokenizer_class": "GPT2Tokenizer"}))
    (d / "special_tokens_map.json").write_text(json.dumps({"bos_token": "</s>", "eos_token": "</s>", "pad_token": "<pad>", "unk_token": "</s>"}))
    hf = transformers.AutoTokenizer.from_pretrained(str(d))
    if (d / "tokenizer.json").exists():                      # from_pretrained does not write files; keep the layout pure
        os.remove(d / "tokenizer.json")
    texts = [t for t in TEXTS if "<|special|>" not in t]
    outs = run_cli(str(d), [{"text": t} for t in texts])
    bos = hf.convert_tokens_to_ids("</s>")
    for t, o in zip(texts, outs):
        ref = hf(t)["input_ids"]
        assert ref[0] == bos and o["ids"] == ref, (t, o["ids"], ref)
    outs = run_cli(str(d), [{"text": "a</s>b<pad>", "add_special_tokens": False}])
    assert outs[0]["ids"] == hf("a</s>b<pad>", add_special_tokens=False)["input_ids"]
    ids = [hf(t)["input_ids"] for t in texts]
    outs = run_cli(str(d), [{"ids": i} for i in ids])
    for t, i, o in zip(texts, ids, outs):
        assert o["text"] == hf.decode(i, skip_special_tokens=True) == t
