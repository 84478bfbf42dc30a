"""Same-box table: this repo's kernels next to the sm_100 libraries installed in the image, on the C3 shapes.
  GEMM    gemm_tc_kernel (through kge_gemm_bf16) vs cuBLAS (torch.matmul), prefill and decode shapes
  prefill attn_tc_kernel (through kge_paged_attn) vs FlashInfer's trtllm-gen fmha (trtllm_batch_context_with_kv_cache)
  decode  attn_decode_kernel vs FlashInfer's trtllm-gen decode (trtllm_batch_decode_with_kv_cache)
CUDA events on the launching (default) stream, buffers rotated past L2 for the HBM-bound cases.  The library calls are
comparison points only (tools/, never the product).  Writes one JSON document to --out.
  python tools/bench_vs_libs.py --out gpurun_out/r02_vs_libs.json"""
import argparse
import json
import os
import sys
import traceback

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kge_loader import load  # noqa: E402

kge = load()
L = kge.lib()
BF = torch.bfloat16
_FI_WS = None


def fi_workspace():
    """One workspace for every FlashInfer call of the process (its trtllm-gen wrappers keep state keyed to the buffer)."""
    global _FI_WS
    if _FI_WS is None:
        _FI_WS = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
    return _FI_WS


def time_ms(fn, iters, warm=3):
    """Per-launch device time of fn(i, stream): `iters` launches are captured into ONE CUDA graph on a side stream and the
    graph is replayed, so host launch cost (which exceeds the run time of the decode-sized kernels here) is out of the
    measurement for both sides; falls back to eager back-to-back launches if the callee cannot be captured."""
    st = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        for i in range(warm):
            fn(i, st.cuda_stream)
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for i in range(iters):
                    fn(i, st.cuda_stream)
            g.replay()
            torch.cuda.synchronize()
            best = None
            for _ in range(3):
                e0.record(st)
                g.replay()
                e1.record(st)
                torch.cuda.synchronize()
                t = e0.elapsed_time(e1) / iters
                best = t if best is None else min(best, t)
            return best
        except Exception as ex:
            print(f"(graph capture failed, eager timing: {type(ex).__name__}: {str(ex)[:120]})", flush=True)
            torch.cuda.synchronize()
            e0.record(st)
            for i in range(iters):
                fn(i, st.cuda_stream)
            e1.record(st)
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / iters


def gemm_rows():
    rows = []
    shapes = [(8192, 6144, 4096), (8192, 4096, 4096), (8192, 28672, 4096), (8192, 4096, 14336),
              (64, 6144, 4096), (64, 4096, 4096), (64, 28672, 4096), (64, 4096, 14336), (64, 128256, 4096)]
    for (T, N, K) in shapes:
        ncopies = min(12, max(2, int(400e6 // (N * K * 2)) + 1))
        ws = [torch.randn(N, K, device="cuda", dtype=BF) * 0.02 for _ in range(ncopies)]
        x = torch.randn(T, K, device="cuda", dtype=BF)
        y = torch.empty(T, N, device="cuda", dtype=BF)

        def ours(i, st):
            kge.check(L.kge_gemm_bf16(y.data_ptr(), x.data_ptr(), ws[i % ncopies].data_ptr(), None, T, N, K, 0, None, 0, st), "gemm")
        it = 20 if T > 256 else 60
        ms = time_ms(ours, it)
        ms_sk = None
        if T <= 128:          # the stream-K variant (knob gemm_ws) next to the default
            L.kge_set_tunable(b"gemm_ws", 1)
            nws = L.kge_gemm_workspace_bytes(T, N, K)
            if nws:
                wsb = torch.zeros(nws, dtype=torch.uint8, device="cuda")

                def ours_sk(i, st):
                    kge.check(L.kge_gemm_bf16(y.data_ptr(), x.data_ptr(), ws[i % ncopies].data_ptr(), None, T, N, K, 0, wsb.data_ptr(), nws, st), "gemm")
                ms_sk = time_ms(ours_sk, it)
            L.kge_set_tunable(b"gemm_ws", 0)
        ms2 = time_ms(lambda i, st: torch.matmul(x, ws[i % ncopies].t(), out=y), it)
        by, fl = 2.0 * (N * K + T * K + T * N), 2.0 * T * N * K
        rows.append({"T": T, "N": N, "K": K, "kge_ms": round(ms, 4), "cublas_ms": round(ms2, 4),
                     "kge_TFLOP/s": round(fl / ms / 1e9, 1), "cublas_TFLOP/s": round(fl / ms2 / 1e9, 1),
                     "kge_GB/s": round(by / ms / 1e6, 1), "cublas_GB/s": round(by / ms2 / 1e6, 1),
                     "winner": "kge" if ms <= ms2 else "cublas"})
        if ms_sk is not None:
            rows[-1]["kge_stream_k_ms"] = round(ms_sk, 4)
        print(json.dumps(rows[-1]), flush=True)
        del ws
    return rows


def paged_setup(B, ctx, KV, hd, copies, bs=16):
    nb_seq = (ctx + bs - 1) // bs
    nblk = B * nb_seq + 1
    kv = [torch.randn(nblk, 2, KV, bs, hd, device="cuda", dtype=BF) for _ in range(copies)]   # FlashInfer HND: [page, 2, KV, page_size, hd]
    bt = (torch.arange(B * nb_seq, dtype=torch.int32).view(B, nb_seq) + 1).cuda()
    return bs, nb_seq, nblk, kv, bt


def prefill_rows():
    rows = []
    for (B, qlen, H, KV, hd, bsz) in [(4, 2048, 32, 8, 128, 64), (1, 8192, 32, 8, 128, 64), (4, 2048, 32, 8, 128, 16)]:
        bs, nb_seq, nblk, kvs, bt = paged_setup(B, qlen, KV, hd, 1, bsz)
        kv = kvs[0]
        kc, vc = kv[:, 0].contiguous(), kv[:, 1].contiguous()
        sl = torch.full((B,), qlen, dtype=torch.int32, device="cuda")
        qsl = (torch.arange(B + 1, dtype=torch.int32) * qlen).cuda()
        qw = (H + 2 * KV) * hd
        qkv = torch.randn(B * qlen, qw, device="cuda", dtype=BF)
        out = torch.empty(B * qlen, H * hd, device="cuda", dtype=BF)

        def ours(i, st):
            kge.check(L.kge_paged_attn(out.data_ptr(), qkv.data_ptr(), qw, kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), nb_seq,
                                       sl.data_ptr(), qsl.data_ptr(), B, B * qlen, qlen, qlen, H, KV, hd, bs, hd ** -0.5, None, 0, st), "attn")
        ms = time_ms(ours, 10)
        L.kge_set_tunable(b"attn_tma", 0)
        ms_old = time_ms(ours, 10)
        L.kge_set_tunable(b"attn_tma", 1)
        flops = 4.0 * H * hd * B * qlen * (qlen + 1) / 2
        row = {"B": B, "q_len": qlen, "H": H, "KV": KV, "hd": hd, "kv_block": bs, "kge_ms": round(ms, 4), "kge_TFLOP/s": round(flops / ms / 1e9, 1),
               "kge_cp_async_variant_ms": round(ms_old, 4), "kge_cp_async_variant_TFLOP/s": round(flops / ms_old / 1e9, 1)}
        try:
            import flashinfer
            q = qkv[:, :H * hd].reshape(B * qlen, H, hd).contiguous()
            wsb = fi_workspace()
            o2 = torch.empty(B * qlen, H, hd, device="cuda", dtype=BF)

            def fi(i, st):
                flashinfer.prefill.trtllm_batch_context_with_kv_cache(
                    q, kv, wsb, bt, sl, qlen, qlen, hd ** -0.5, 1.0, B, qsl, qsl, out=o2, kv_layout="HND")
            ms2 = time_ms(fi, 10)
            err = (o2.reshape(B * qlen, H * hd).float() - out.float()).abs().max().item()
            row.update({"flashinfer_trtllm_gen_ms": round(ms2, 4), "flashinfer_TFLOP/s": round(flops / ms2 / 1e9, 1),
                        "max_abs_diff_between_the_two": err, "winner": "kge" if ms <= ms2 else "flashinfer"})
        except Exception as ex:      # comparison point only: record why it is missing
            row["flashinfer_error"] = f"{type(ex).__name__}: {ex}"[:400]
            traceback.print_exc()
        rows.append(row)
        print(json.dumps(row), flush=True)
    return rows


def decode_rows():
    rows = []
    for (B, ctx, H, KV, hd, bsz) in [(64, 2176, 32, 8, 128, 64), (64, 2176, 32, 8, 128, 16), (32, 192, 32, 8, 64, 64)]:
        bs, nb_seq, nblk, kvs, bt = paged_setup(B, ctx, KV, hd, 3, bsz)
        kcs = [kv[:, 0].contiguous() for kv in kvs]
        vcs = [kv[:, 1].contiguous() for kv in kvs]
        sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
        qsl = torch.arange(B + 1, dtype=torch.int32, device="cuda")
        qw = (H + 2 * KV) * hd
        qkv = torch.randn(B, qw, device="cuda", dtype=BF)
        out = torch.empty(B, H * hd, device="cuda", dtype=BF)
        nws = L.kge_attn_workspace_bytes(B, H, hd, ctx)
        ws = torch.empty(nws, dtype=torch.uint8, device="cuda")

        def ours(i, st):
            kge.check(L.kge_paged_attn(out.data_ptr(), qkv.data_ptr(), qw, kcs[i % 3].data_ptr(), vcs[i % 3].data_ptr(), bt.data_ptr(),
                                       nb_seq, sl.data_ptr(), qsl.data_ptr(), B, B, 1, ctx, H, KV, hd, bs, hd ** -0.5, ws.data_ptr(), nws, st), "attn")
        ms = time_ms(ours, 60)
        L.kge_set_tunable(b"attn_tma", 0)
        ms_old = time_ms(ours, 60)
        L.kge_set_tunable(b"attn_tma", 1)
        by = B * ctx * KV * hd * 4.0
        row = {"B": B, "ctx": ctx, "H": H, "KV": KV, "hd": hd, "kv_block": bs, "kge_ms": round(ms, 4), "kge_GB/s": round(by / ms / 1e6, 1),
               "kge_cp_async_variant_ms": round(ms_old, 4), "kge_cp_async_variant_GB/s": round(by / ms_old / 1e6, 1)}
        try:
            import flashinfer
            q = qkv[:, :H * hd].reshape(B, H, hd).contiguous()
            wsb = fi_workspace()
            o2 = torch.empty(B, H, hd, device="cuda", dtype=BF)

            def fi(i, st):
                flashinfer.decode.trtllm_batch_decode_with_kv_cache(q, kvs[i % 3], wsb, bt, sl, ctx, hd ** -0.5, 1.0, out=o2, kv_layout="HND")
            ms2 = time_ms(fi, 60)
            row.update({"flashinfer_trtllm_gen_ms": round(ms2, 4), "flashinfer_GB/s": round(by / ms2 / 1e6, 1),
                        "winner": "kge" if ms <= ms2 else "flashinfer"})
        except Exception as ex:
            row["flashinfer_error"] = f"{type(ex).__name__}: {ex}"[:400]
            traceback.print_exc()
        rows.append(row)
        print(json.dumps(row), flush=True)
    return rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/r02_vs_libs.json")
    ap.add_argument("--skip", default="")
    a = ap.parse_args()
    kge.require_cuda()
    doc = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__}
    if "gemm" not in a.skip:
        doc["gemm_vs_cublas"] = gemm_rows()
    if "prefill" not in a.skip:
        doc["prefill_attention_vs_flashinfer"] = prefill_rows()
    if "decode" not in a.skip:
        doc["decode_attention_vs_flashinfer"] = decode_rows()
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    json.dump(doc, open(a.out, "w"), indent=1)
