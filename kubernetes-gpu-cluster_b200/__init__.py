"""kubernetes-gpu-cluster_b200: B200-native serving engine behind the container contract of the
reference's Helm values (ref:values-01-minimal-example*.yaml -> `vllm serve <model> ...`).

This package is a thin ctypes layer over libkge.so (CUDA kernels + C++ engine, include/kge.h).
It mirrors the deployed engine's Python surface for the hot path - `LLM.generate` with
`SamplingParams` (vllm:entrypoints/llm.py, vllm:sampling_params.py) - so parity tests read like
tests of the engine the values files deploy.  There is NO CPU path: every compute call raises
if libkge.so is missing or no CUDA device is present.

The directory name is not a Python identifier; load it with kge_loader.load() (repo root).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkge.so")


class KgeError(RuntimeError):
    pass


class EngineConfig(C.Structure):
    _fields_ = [("max_num_seqs", C.c_int32), ("max_num_batched_tokens", C.c_int32),
                ("max_model_len", C.c_int32), ("block_size", C.c_int32), ("num_blocks", C.c_int32),
                ("gpu_memory_utilization", C.c_float), ("tp_rank", C.c_int32), ("tp_size", C.c_int32),
                ("enable_chunked_prefill", C.c_int32), ("device", C.c_int32), ("dtype", C.c_int32)]


class StepBatch(C.Structure):
    _fields_ = [("T", C.c_int32), ("B", C.c_int32), ("token_ids", C.c_void_p), ("positions", C.c_void_p),
                ("query_start_loc", C.c_void_p), ("seq_lens", C.c_void_p), ("block_table", C.c_void_p),
                ("max_blocks", C.c_int32), ("slot_mapping", C.c_void_p), ("logits_rows", C.c_void_p),
                ("R", C.c_int32)]


class CSamplingParams(C.Structure):
    _fields_ = [("max_tokens", C.c_int32), ("min_tokens", C.c_int32), ("ignore_eos", C.c_int32),
                ("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float), ("seed", C.c_uint64),
                ("stop_token_ids", C.c_void_p), ("num_stop_token_ids", C.c_int32), ("has_seed", C.c_int32)]


class StepOutput(C.Structure):
    _fields_ = [("request_id", C.c_int64), ("token", C.c_int32), ("finished", C.c_int32),
                ("num_scheduled", C.c_int32)]


def _event_dtype():
    import numpy as np
    return np.dtype({"names": ["request_id", "token", "finished", "num_scheduled"], "formats": ["<i8", "<i4", "<i4", "<i4"],
                     "offsets": [0, 8, 12, 16], "itemsize": C.sizeof(StepOutput)})


class TpPeerInfo(C.Structure):
    """kge_tp_peer_info: what a tensor-parallel rank publishes so that its peers can map its exchange window."""
    _fields_ = [("ipc_handle", C.c_char * 64), ("ptr", C.c_uint64), ("pid", C.c_int64), ("bytes", C.c_int64),
                ("device", C.c_int32), ("rank", C.c_int32)]


class KernelStats(C.Structure):
    _fields_ = [("launches", C.c_int64), ("ms", C.c_double), ("bytes", C.c_double), ("flops", C.c_double)]


NUM_KERNEL_CATS = 10
KERNEL_CATS = ["gemm_decode", "gemm_prefill", "attn_decode", "attn_general", "norm", "rope_kvwrite",
               "activation", "embed_gather", "sampling", "all_reduce"]

# name -> (restype, argtypes); every symbol include/kge.h declares
_P, _I, _L, _F = C.c_void_p, C.c_int, C.c_int64, C.c_float
SIGNATURES = {
    "kge_last_error": (C.c_char_p, []),
    "kge_abi_version": (_I, []),
    "kge_device_sm_count": (_I, []),
    "kge_embed": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "kge_rmsnorm": (_I, [_P, _P, _P, _I, _I, _F, _P]),
    "kge_add_rmsnorm": (_I, [_P, _P, _P, _P, _I, _I, _F, _P]),
    "kge_head_rmsnorm": (_I, [_P, _P, _P, _I, _I, _I, _I, _F, _P]),
    "kge_layernorm": (_I, [_P, _P, _P, _P, _P, _I, _I, _F, _P]),
    "kge_silu_mul": (_I, [_P, _P, _I, _I, _P]),
    "kge_gemm_bf16": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _P, _L, _P]),
    "kge_gemm_workspace_bytes": (_L, [_I, _I, _I]),
    "kge_gemm_silu_gate_bf16": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "kge_gemm_silu_gate_ws_bf16": (_I, [_P, _P, _P, _I, _I, _I, _P, _L, _P]),
    "kge_rope_kvwrite": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "kge_rope_cache_host": (_I, [_P, _I, _I, C.c_double, C.c_double, C.c_double, C.c_double, _I]),
    "kge_paged_attn": (_I, [_P, _P, _I, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _F, _P, _L, _P]),
    "kge_attn_workspace_bytes": (_L, [_I, _I, _I, _I]),
    "kge_paged_attn_decode_noalloc": (_I, [_P, _P, _I, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _I, _I, _F, _P]),
    "kge_paged_attn_decode_noalloc_max_bytes": (_L, []),
    "kge_argmax": (_I, [_P, _P, _I, _I, _I, _P]),
    "kge_sample": (_I, [_P, _P, _I, _I, _I, _P, _P, _P, _P, _P, C.c_uint64, _P, _L, _P]),
    "kge_sample_workspace_bytes": (_L, [_I, _I]),
    "kge_engine_create": (_I, [C.POINTER(_P), C.c_char_p, C.POINTER(EngineConfig), _L]),
    "kge_engine_destroy": (None, [_P]),
    "kge_engine_num_blocks": (_I, [_P]),
    "kge_engine_dtype": (_I, [_P]),
    "kge_tokenizer_open": (_P, [C.c_char_p]),
    "kge_tokenizer_close": (None, [_P]),
    "kge_tokenizer_vocab_size": (_I, [_P]),
    "kge_tokenizer_encode": (C.c_int64, [_P, C.c_char_p, C.c_int64, _I, _P, C.c_int64]),
    "kge_tokenizer_decode": (C.c_int64, [_P, _P, C.c_int64, _P, C.c_int64]),
    "kge_model_open": (_P, [C.c_char_p]),
    "kge_model_close": (None, [_P]),
    "kge_model_spec_json": (_I, [_P, C.c_char_p, _I]),
    "kge_model_read_bf16": (C.c_int64, [_P, C.c_char_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, _P, C.c_int64]),
    "kge_model_read_f16": (C.c_int64, [_P, C.c_char_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, _P, C.c_int64]),
    "kge_engine_vocab": (_I, [_P]),
    "kge_engine_eos_token_id": (_I, [_P]),
    "kge_engine_max_model_len": (_I, [_P]),
    "kge_engine_weight_bytes": (_L, [_P]),
    "kge_engine_launches": (_L, [_P]),
    "kge_step": (_I, [_P, C.POINTER(StepBatch), _P, _P]),
    "kge_engine_set_timing": (_I, [_P, _I]),
    "kge_engine_get_timing": (_I, [_P, _P]),
    "kge_engine_io_bytes": (_I, [_P, _P, _P]),
    "kge_engine_run_resident": (_I, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, _P, _P]),
    "kge_engine_record_steps": (_I, [_P, _I]),
    "kge_engine_replay_recorded": (_I, [_P, _P, _P, _L, C.POINTER(_L), C.POINTER(C.c_int32), _P, _P]),
    "kge_nccl_unique_id": (_I, [_P]),
    "kge_engine_init_tp": (_I, [_P, _P]),
    "kge_tp_comm_create": (_I, [C.POINTER(_P), _I, _I, _I, _L]),
    "kge_tp_comm_destroy": (None, [_P]),
    "kge_tp_comm_info": (_I, [_P, C.POINTER(TpPeerInfo)]),
    "kge_tp_comm_connect": (_I, [_P, C.POINTER(TpPeerInfo)]),
    "kge_tp_comm_connected": (_I, [_P]),
    "kge_tp_comm_data": (_P, [_P]),
    "kge_tp_reduce_add_norm": (_I, [_P, _L, _P, _L, _P, _P, _I, _I, _F, _I, _P]),
    "kge_tp_all_gather_cols": (_I, [_P, _L, _P, _I, _I, _I, _I, _P]),
    "kge_engine_tp_info": (_I, [_P, C.POINTER(TpPeerInfo)]),
    "kge_engine_tp_connect": (_I, [_P, C.POINTER(TpPeerInfo)]),
    "kge_add_request": (_I, [_P, _L, _P, C.c_int32, C.POINTER(CSamplingParams)]),
    "kge_request_set_logprobs": (_I, [_P, C.c_int64, C.c_int32]),
    "kge_engine_step_logprobs": (_I, [_P, C.c_int64, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int32]),
    "kge_logprobs": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "kge_abort_request": (_I, [_P, _L]),
    "kge_has_unfinished": (_I, [_P]),
    "kge_engine_run_to_completion": (_I, [_P, _P, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "kge_engine_step": (_I, [_P, C.POINTER(StepOutput), C.c_int32]),
    "kge_engine_stats": (_I, [_P, _P, _P, _P, _P]),
    "kge_sched_create": (_P, [C.c_int32] * 7),
    "kge_set_tunable": (C.c_int, [C.c_char_p, C.c_int]),
    "kge_gemm_set_trace": (None, [C.c_void_p]),
    "kge_sched_destroy": (None, [_P]),
    "kge_sched_set_reserve_full_isl": (None, [_P, C.c_int32]),
    "kge_engine_set_reserve_full_isl": (None, [_P, C.c_int32]),
    "kge_sched_set_prefix_caching": (None, [_P, C.c_int32]),
    "kge_engine_set_prefix_caching": (None, [_P, C.c_int32]),
    "kge_sched_cached_tokens": (C.c_int32, [_P, C.c_int64]),
    "kge_engine_prefix_cache_stats": (_I, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "kge_sched_add": (_I, [_P, _L, _P, C.c_int32, C.POINTER(CSamplingParams)]),
    "kge_sched_schedule": (_I, [_P, _P, _P, C.c_int32]),
    "kge_sched_update": (_I, [_P, _P, _P, C.c_int32, C.POINTER(StepOutput), C.c_int32]),
    "kge_sched_counts": (_I, [_P, _P, _P, _P]),
}

# float16 twins of the element-type dependent ops: same signatures (include/kge.h)
for _n, _t in {"kge_embed": "kge_embed_f16", "kge_rmsnorm": "kge_rmsnorm_f16", "kge_add_rmsnorm": "kge_add_rmsnorm_f16",
               "kge_head_rmsnorm": "kge_head_rmsnorm_f16", "kge_layernorm": "kge_layernorm_f16", "kge_silu_mul": "kge_silu_mul_f16",
               "kge_gemm_bf16": "kge_gemm_f16", "kge_gemm_silu_gate_bf16": "kge_gemm_silu_gate_f16",
               "kge_gemm_silu_gate_ws_bf16": "kge_gemm_silu_gate_ws_f16",
               "kge_rope_kvwrite": "kge_rope_kvwrite_f16", "kge_rope_cache_host": "kge_rope_cache_host_f16",
               "kge_paged_attn": "kge_paged_attn_f16", "kge_argmax": "kge_argmax_f16", "kge_sample": "kge_sample_f16",
               "kge_logprobs": "kge_logprobs_f16", "kge_tp_reduce_add_norm": "kge_tp_reduce_add_norm_f16"}.items():
    SIGNATURES[_t] = SIGNATURES[_n]

DTYPES = {"auto": 0, "bfloat16": 1, "float16": 2, "half": 2}
_lib = None


def lib() -> C.CDLL:
    """Load libkge.so (built in-tree by __graft_entry__.build() / make).  Raises if missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise KgeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`"
                           " - there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc: int, what: str = "") -> int:
    if rc < 0:
        raise KgeError(f"{what}: {lib().kge_last_error().decode()} (rc={rc})")
    return rc


def require_cuda() -> int:
    n = lib().kge_device_sm_count()
    if n <= 0:
        raise KgeError("no CUDA device: this engine has no CPU path")
    return n


# ------------------------------------------------------------------------- public API


@dataclass
class SamplingParams:
    """Subset of vllm:sampling_params.py used on the hot path (same names and defaults)."""
    max_tokens: int = 16
    min_tokens: int = 0
    temperature: float = 1.0
    top_k: int = 0
    top_p: float = 1.0
    seed: int | None = None
    ignore_eos: bool = False
    stop_token_ids: list = field(default_factory=list)
    logprobs: int | None = None         # top-N log-probabilities per sampled token (0: the sampled token's only)

    def to_c(self):
        arr = (C.c_int32 * max(1, len(self.stop_token_ids)))(*self.stop_token_ids)
        c = CSamplingParams(self.max_tokens, self.min_tokens, int(self.ignore_eos), self.temperature,
                            self.top_k, self.top_p, self.seed or 0,
                            C.cast(arr, C.c_void_p), len(self.stop_token_ids), int(self.seed is not None))
        c._keep = arr
        return c


@dataclass
class Logprob:
    """vllm:logprobs.py Logprob (decoded_token is left to the caller's tokenizer)."""
    logprob: float
    rank: int | None = None


@dataclass
class CompletionOutput:
    token_ids: list
    finish_reason: str | None = None
    logprobs: list | None = None        # per generated token: {token_id: Logprob}, sampled token first
    text: str = ""                      # decoded token_ids when the model directory has a tokenizer


class Tokenizer:
    """The model directory's tokenizer through the C ABI (kge_tokenizer_*): what lets `LLM.generate` take text."""

    def __init__(self, model_dir: str):
        self._h = lib().kge_tokenizer_open(str(model_dir).encode())
        if not self._h:
            raise KgeError("kge_tokenizer_open: " + lib().kge_last_error().decode())

    def close(self):
        if getattr(self, "_h", None):
            lib().kge_tokenizer_close(self._h)
            self._h = None

    __del__ = close

    @property
    def vocab_size(self) -> int:
        return lib().kge_tokenizer_vocab_size(self._h)

    def encode(self, text: str, add_special_tokens: bool = True) -> list:
        import numpy as np
        b = text.encode("utf-8")
        n = lib().kge_tokenizer_encode(self._h, b, len(b), int(add_special_tokens), None, 0)
        if n < 0:
            raise KgeError("kge_tokenizer_encode: " + lib().kge_last_error().decode())
        out = np.zeros(max(1, n), dtype=np.int32)
        lib().kge_tokenizer_encode(self._h, b, len(b), int(add_special_tokens), out.ctypes.data, n)
        return out[:n].tolist()

    def decode(self, ids) -> str:
        import numpy as np
        a = np.ascontiguousarray(ids, dtype=np.int32)
        n = lib().kge_tokenizer_decode(self._h, a.ctypes.data, int(a.size), None, 0)
        if n < 0:
            raise KgeError("kge_tokenizer_decode: " + lib().kge_last_error().decode())
        buf = C.create_string_buffer(max(1, n))
        lib().kge_tokenizer_decode(self._h, a.ctypes.data, int(a.size), buf, n)
        return buf.raw[:n].decode("utf-8")


@dataclass
class RequestOutput:
    request_id: int
    prompt_token_ids: list
    outputs: list


class LLM:
    """`vllm.LLM`-shaped front end over the C++ engine (vllm:entrypoints/llm.py):
    LLM(model=<HF directory>, ...).generate(prompt_token_ids, sampling_params)."""

    def __init__(self, model: str, *, max_model_len: int = 0, max_num_seqs: int = 0,
                 max_num_batched_tokens: int = 0, gpu_memory_utilization: float = 0.9,
                 block_size: int = 16, num_gpu_blocks_override: int = 0, tensor_parallel_size: int = 1,
                 tp_rank: int = 0, enable_chunked_prefill: bool = True, load_format: str = "auto",
                 seed: int = 0, device: int = 0, scheduler_reserve_full_isl: bool = True,
                 enable_prefix_caching: bool = True, dtype: str = "auto"):
        require_cuda()
        if dtype not in DTYPES:
            raise KgeError(f"dtype {dtype!r}: this engine computes in bfloat16 or float16 (auto, bfloat16, float16)")
        self.cfg = EngineConfig(max_num_seqs, max_num_batched_tokens, max_model_len, block_size,
                                num_gpu_blocks_override, gpu_memory_utilization, tp_rank,
                                tensor_parallel_size, int(enable_chunked_prefill), device, DTYPES[dtype])
        h = C.c_void_p()
        dummy = seed if load_format == "dummy" else -1
        check(lib().kge_engine_create(C.byref(h), model.encode(), C.byref(self.cfg), dummy), "kge_engine_create")
        self._h = h
        self._model = model
        self._tok = None                # opened on first use; False = the directory has no tokenizer
        lib().kge_engine_set_reserve_full_isl(h, int(scheduler_reserve_full_isl))
        lib().kge_engine_set_prefix_caching(h, int(enable_prefix_caching))      # on by default, as in vLLM V1
        self._next_id = 0
        self._outs = (StepOutput * 4096)()

    def close(self):
        if getattr(self, "_h", None):
            lib().kge_engine_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def dtype(self) -> str:
        """What the engine computes in: "bfloat16" or "float16" (what `dtype="auto"` settled on)."""
        return {1: "bfloat16", 2: "float16"}.get(lib().kge_engine_dtype(self._h), "?")

    @property
    def num_gpu_blocks(self) -> int:
        return lib().kge_engine_num_blocks(self._h)

    @property
    def vocab_size(self) -> int:
        return lib().kge_engine_vocab(self._h)

    @property
    def weight_bytes(self) -> int:
        return lib().kge_engine_weight_bytes(self._h)

    @property
    def kernel_launches(self) -> int:
        return lib().kge_engine_launches(self._h)

    def set_timing(self, enable: bool):
        check(lib().kge_engine_set_timing(self._h, int(enable)), "kge_engine_set_timing")

    def kernel_stats(self) -> dict:
        """Per-category device time and algorithmic work since set_timing(True)."""
        arr = (KernelStats * NUM_KERNEL_CATS)()
        check(lib().kge_engine_get_timing(self._h, arr), "kge_engine_get_timing")
        return {KERNEL_CATS[i]: dict(launches=arr[i].launches, ms=arr[i].ms, bytes=arr[i].bytes, flops=arr[i].flops)
                for i in range(NUM_KERNEL_CATS) if arr[i].launches}

    def prefix_cache_stats(self):
        """(tokens looked up, tokens served from the prefix cache) since the engine was created."""
        q, h = C.c_int64(), C.c_int64()
        check(lib().kge_engine_prefix_cache_stats(self._h, C.byref(q), C.byref(h)), "kge_engine_prefix_cache_stats")
        return q.value, h.value

    def io_bytes(self):
        a, b = C.c_int64(), C.c_int64()
        check(lib().kge_engine_io_bytes(self._h, C.byref(a), C.byref(b)), "kge_engine_io_bytes")
        return a.value, b.value

    def run_resident(self, prompts, out_len: int, prefill_batch: int):
        """Device-resident replay of a fixed-shape greedy job (see include/kge.h).
        Returns (tokens [B][out_len], total_ms, decode_ms)."""
        import numpy as np
        pr = np.ascontiguousarray(np.asarray(prompts, dtype=np.int32))
        B, P = pr.shape
        out = np.zeros((B, out_len), dtype=np.int32)
        ms, dms = C.c_float(), C.c_float()
        check(lib().kge_engine_run_resident(self._h, pr.ctypes.data, B, P, out_len, prefill_batch, out.ctypes.data,
                                            C.byref(ms), C.byref(dms)), "kge_engine_run_resident")
        return out.tolist(), ms.value, dms.value

    def record_steps(self, enable: bool):
        """While on, every scheduler-driven step of generate() is kept for replay_recorded()."""
        check(lib().kge_engine_record_steps(self._h, int(enable)), "kge_engine_record_steps")

    def replay_recorded(self, max_samples: int):
        """Device-resident replay of the recorded job.  Returns ({request_id: [tokens]}, steps, total_ms, trailing_decode_ms)."""
        import numpy as np
        toks = np.zeros(max_samples, dtype=np.int32)
        rids = np.zeros(max_samples, dtype=np.int64)
        n, steps = C.c_int64(), C.c_int32()
        ms, dms = C.c_float(), C.c_float()
        check(lib().kge_engine_replay_recorded(self._h, toks.ctypes.data, rids.ctypes.data, max_samples, C.byref(n), C.byref(steps),
                                               C.byref(ms), C.byref(dms)), "kge_engine_replay_recorded")
        out = {}
        for r, t in zip(rids[:n.value].tolist(), toks[:n.value].tolist()):
            out.setdefault(r, []).append(t)
        return out, steps.value, ms.value, dms.value

    def init_tp(self, unique_id: bytes):
        """NCCL exchange (what --disable-custom-all-reduce selects): every rank passes rank 0's kge_nccl_unique_id."""
        check(lib().kge_engine_init_tp(self._h, unique_id), "kge_engine_init_tp")

    def tp_info(self) -> bytes:
        """This rank's exchange-window record (96 bytes) to hand to every peer, by any transport."""
        info = TpPeerInfo()
        check(lib().kge_engine_tp_info(self._h, C.byref(info)), "kge_engine_tp_info")
        return bytes(info)

    def tp_connect(self, infos):
        """infos: every rank's tp_info(), indexed by rank.  Switches the forward pass to the peer-memory exchange
        (fused all-reduce + residual + norm, logits all-gather), programmatic dependent launch and decode graphs."""
        arr = (TpPeerInfo * len(infos))()
        for i, b in enumerate(infos):
            C.memmove(C.addressof(arr[i]), bytes(b), C.sizeof(TpPeerInfo))
        check(lib().kge_engine_tp_connect(self._h, arr), "kge_engine_tp_connect")

    def tp_connect_torch(self, group=None):
        """tp_info / tp_connect over an initialised torch.distributed process group (rank order = TP rank order)."""
        import torch
        import torch.distributed as dist
        dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        mine = torch.frombuffer(bytearray(self.tp_info()), dtype=torch.uint8).to(dev)
        outs = [torch.empty_like(mine) for _ in range(dist.get_world_size(group))]
        dist.all_gather(outs, mine, group=group)
        self.tp_connect([bytes(o.cpu().numpy().tobytes()) for o in outs])
        dist.barrier(group=group)        # nobody starts a step before every rank has mapped every window

    def add_request(self, prompt_token_ids, sp: SamplingParams) -> int:
        rid = self._next_id
        self._next_id += 1
        import numpy as np
        arr = np.ascontiguousarray(prompt_token_ids, dtype=np.int32)      # 3x cheaper than a ctypes array for long prompts
        c = sp.to_c()
        check(lib().kge_add_request(self._h, rid, arr.ctypes.data, int(arr.size), C.byref(c)), "kge_add_request")
        if sp.logprobs is not None:
            check(lib().kge_request_set_logprobs(self._h, rid, int(sp.logprobs)), "kge_request_set_logprobs")
        return rid

    def step_logprobs(self, rid: int):
        """{token_id: Logprob} of the token request `rid` sampled in the last step (sampled token first), or None."""
        vals, ids, rank = (C.c_float * 21)(), (C.c_int32 * 21)(), C.c_int32()
        n = check(lib().kge_engine_step_logprobs(self._h, rid, vals, ids, C.byref(rank), 21), "kge_engine_step_logprobs")
        if n <= 0:
            return None
        d = {ids[0]: Logprob(vals[0], rank.value)}
        for k in range(1, n):
            d.setdefault(ids[k], Logprob(vals[k], k))
        return d

    def step(self):
        """One engine step (schedule + forward + sample + update).  Returns a list of
        (request_id, token or -1, finished 0|1|2, num_scheduled)."""
        n = check(lib().kge_engine_step(self._h, self._outs, len(self._outs)), "kge_engine_step")
        return [(o.request_id, o.token, o.finished, o.num_scheduled) for o in self._outs[:n]]

    def has_unfinished(self) -> bool:
        return bool(lib().kge_has_unfinished(self._h))

    def stats(self):
        r, w, p = C.c_int32(), C.c_int32(), C.c_int64()
        u = C.c_float()
        check(lib().kge_engine_stats(self._h, C.byref(r), C.byref(w), C.byref(u), C.byref(p)), "kge_engine_stats")
        return {"num_running": r.value, "num_waiting": w.value, "kv_usage": u.value, "num_preemptions": p.value}

    def get_tokenizer(self):
        """The model directory's tokenizer, or None when it has none (token-id prompts still work)."""
        if self._tok is None:
            try:
                self._tok = Tokenizer(self._model)
            except KgeError:
                self._tok = False
        return self._tok or None

    def _prompt_ids(self, p):
        """vLLM's prompt forms: token ids, text, {"prompt_token_ids": [...]} or {"prompt": "..."}."""
        if isinstance(p, dict):
            p = p["prompt_token_ids"] if "prompt_token_ids" in p else p["prompt"]
        if isinstance(p, str):
            tok = self.get_tokenizer()
            if tok is None:
                raise KgeError("text prompts need a tokenizer.json (or vocab.json + merges.txt) in the model directory; pass token ids")
            return tok.encode(p)
        return p

    def generate(self, prompt_token_ids, sampling_params: SamplingParams | list | None = None,
                 on_step=None):
        """prompt_token_ids: one prompt or a list of prompts, each a list of token ids, a string, or vLLM's dict forms.
        Returns a RequestOutput per prompt, in order."""
        if isinstance(prompt_token_ids, (str, dict)) or (isinstance(prompt_token_ids, (list, tuple)) and prompt_token_ids
                                                          and isinstance(prompt_token_ids[0], int)):
            prompt_token_ids = [prompt_token_ids]
        prompt_token_ids = [self._prompt_ids(p) for p in prompt_token_ids]
        sps = sampling_params or SamplingParams()
        if not isinstance(sps, list):
            sps = [sps] * len(prompt_token_ids)
        ids = [self.add_request(p, sp) for p, sp in zip(prompt_token_ids, sps)]
        toks = {i: [] for i in ids}
        fin = {}
        want_lp = {i for i, sp in zip(ids, sps) if sp.logprobs is not None}
        lps = {i: [] for i in want_lp}
        if on_step is None and not want_lp:
            # the whole job in one C call: no Python between engine steps
            import numpy as np
            cap = sum(sp.max_tokens for sp in sps) + len(ids) + 16
            ev = np.zeros(cap, dtype=_event_dtype())
            n, steps = C.c_int64(), C.c_int32()
            check(lib().kge_engine_run_to_completion(self._h, ev.ctypes.data, cap, C.byref(n), C.byref(steps)),
                  "kge_engine_run_to_completion")
            ev = ev[:n.value]
            rid, tok, f = ev["request_id"].tolist(), ev["token"].tolist(), ev["finished"].tolist()
            for r, t, ff in zip(rid, tok, f):
                if t >= 0:
                    toks[r].append(t)
                if ff:
                    fin[r] = "stop" if ff == 1 else "length"
            return self._outputs(ids, prompt_token_ids, toks, fin, {})
        while self.has_unfinished():
            outs = self.step()
            if not outs:
                raise KgeError("engine made no progress (KV pool too small for the waiting request?)")
            for rid, tok, f, _ in outs:
                if tok >= 0:
                    toks[rid].append(tok)
                    if rid in want_lp:
                        lps[rid].append(self.step_logprobs(rid))
                if f:
                    fin[rid] = "stop" if f == 1 else "length"
            if on_step:
                on_step(outs)
        return self._outputs(ids, prompt_token_ids, toks, fin, lps)

    def _outputs(self, ids, prompts, toks, fin, lps):
        tok = self.get_tokenizer()
        return [RequestOutput(i, list(p), [CompletionOutput(toks[i], fin.get(i), lps.get(i), tok.decode(toks[i]) if tok else "")])
                for i, p in zip(ids, prompts)]

    # raw step with caller-built metadata (tests / bench device-resident leg)
    def raw_step(self, token_ids, positions, query_start_loc, seq_lens, block_table, max_blocks,
                 slot_mapping, logits_rows, want_logits=False):
        import numpy as np
        a = lambda x, dt: np.ascontiguousarray(np.asarray(x, dtype=dt))
        tid, pos, qsl = a(token_ids, np.int32), a(positions, np.int32), a(query_start_loc, np.int32)
        sl, bt, sm, rows = a(seq_lens, np.int32), a(block_table, np.int32), a(slot_mapping, np.int64), a(logits_rows, np.int32)
        b = StepBatch(len(tid), len(sl), tid.ctypes.data, pos.ctypes.data, qsl.ctypes.data, sl.ctypes.data,
                      bt.ctypes.data, max_blocks, sm.ctypes.data, rows.ctypes.data, len(rows))
        out = np.zeros(max(1, len(rows)), dtype=np.int32)
        lg = np.zeros((len(rows), self.vocab_size), dtype=np.float32) if want_logits else None
        check(lib().kge_step(self._h, C.byref(b), out.ctypes.data, lg.ctypes.data if want_logits else None), "kge_step")
        return out[:len(rows)], lg
