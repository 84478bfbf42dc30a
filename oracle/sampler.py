"""Sampler oracle.  TEST INFRASTRUCTURE (see oracle/__init__.py).

Restates vllm:v1/sample/sampler.py:67-145,235-296 and
vllm:v1/sample/ops/topk_topp_sampler.py:159-186,328-417.
"""
from __future__ import annotations

import torch

GREEDY_EPS = 1e-5  # vllm:sampling_params.py:25 - temperature below this means greedy


def greedy(logits: torch.Tensor) -> torch.Tensor:
    """sampler.py:91,235-236: fp32 cast, argmax, first maximal index wins."""
    return logits.to(torch.float32).argmax(dim=-1)


def apply_top_k_top_p(logits: torch.Tensor, k: torch.Tensor | None,
                      p: torch.Tensor | None) -> torch.Tensor:
    """topk_topp_sampler.py:328-369.  Returns masked logits (-inf where dropped).
    k==V or p==1 disables that filter for the row."""
    logits = logits.to(torch.float32).clone()
    if k is None and p is None:
        return logits
    srt, idx = logits.sort(dim=-1, descending=False)
    if k is not None:
        kth = srt.size(1) - k.to(torch.long)
        kth = srt.gather(1, kth.unsqueeze(1))
        srt.masked_fill_(srt < kth, float("-inf"))
    if p is not None:
        ps = srt.softmax(dim=-1)
        cum = torch.cumsum(ps, dim=-1)
        mask = cum <= 1 - p.unsqueeze(1)
        mask[:, -1] = False
        srt.masked_fill_(mask, float("-inf"))
    return logits.scatter_(dim=-1, index=idx, src=srt)


def sample(logits: torch.Tensor, temperature: torch.Tensor, k: torch.Tensor | None,
           p: torch.Tensor | None, q: torch.Tensor) -> torch.Tensor:
    """sampler.py:254-296 + topk_topp_sampler.py:395-417 with the Exp(1) noise q given by the
    caller ([B,V], > 0) so the draw is reproducible: rows with temperature < GREEDY_EPS take
    argmax(logits); others argmax(softmax(masked(logits / T)) / q)."""
    lg = logits.to(torch.float32)
    g = greedy(lg)
    t = torch.where(temperature < GREEDY_EPS, torch.ones_like(temperature), temperature)
    masked = apply_top_k_top_p(lg / t.unsqueeze(1), k, p)
    probs = masked.softmax(dim=-1, dtype=torch.float32)
    r = probs.div(q).argmax(dim=-1)
    return torch.where(temperature < GREEDY_EPS, g, r)
