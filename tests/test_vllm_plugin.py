"""integration/vllm_kge_ops.py: libkge.so registered as provider "kge" of vLLM's own IR ops (vllm/ir/op.py).
CPU: the provider registers, declares what it covers, and vLLM falls through to `native` where it does not.
GPU: vLLM's dispatcher, told to prefer "kge", runs the sm_100a kernels and the results equal `native` within 1 bf16 ulp."""
import pytest
import torch

pytest.importorskip("vllm")


@pytest.fixture(scope="module")
def ops():
    import integration.vllm_kge_ops as plug
    from vllm.ir.ops.layernorm import fused_add_rms_norm, rms_norm
    return plug, rms_norm, fused_add_rms_norm


def test_provider_registers_and_falls_through_on_cpu(ops):
    plug, rms_norm, fused = ops
    assert "kge" in rms_norm.impls and "kge" in fused.impls
    assert rms_norm.impls["kge"].supported == torch.cuda.is_available()
    x = torch.randn(3, 64).to(torch.bfloat16)
    w = torch.ones(64, dtype=torch.bfloat16)
    assert not plug._rms_ok(x, w, 1e-5)                                # CPU tensor: not covered
    assert not plug._rms_ok(x, None, 1e-5) and not plug._rms_ok(x, w, 1e-5, 32)
    assert not plug._fused_ok(x, x[:2], w, 1e-5)
    with rms_norm.set_priority(["kge", "native"]):                      # dispatch still works: native takes the call
        y = rms_norm(x, w, 1e-5)
    with rms_norm.set_priority(["native"]):
        assert torch.equal(y, rms_norm(x, w, 1e-5))


def test_silu_and_mul_layer_is_replaced(ops):
    from vllm.model_executor.custom_op import op_registry_oot
    plug = ops[0]
    assert op_registry_oot["SiluAndMul"] is plug.KgeSiluAndMul
    assert not plug._silu_covered(torch.zeros(2, 32, dtype=torch.bfloat16))          # CPU tensor
    assert not plug._silu_covered(torch.zeros(2, 20, dtype=torch.bfloat16))          # half width not a multiple of 8


@pytest.mark.gpu
def test_kge_silu_and_mul_matches_vllm_native(ops):
    from vllm.model_executor.layers.activation import SiluAndMul
    plug = ops[0]
    g = torch.Generator().manual_seed(3)
    for shape in [(64, 2 * 14336), (5, 2 * 512), (2, 3, 2 * 64)]:
        x = (torch.randn(*shape, generator=g) * 2).to(torch.bfloat16).cuda()
        assert plug._silu_covered(x)
        got = plug.KgeSiluAndMul.forward_cuda(None, x)
        ref = SiluAndMul.forward_native(x)
        diff = (got.view(torch.int16).int() - ref.view(torch.int16).int()).abs()
        assert got.shape == ref.shape and int(diff.max()) <= 1 and float((diff > 0).float().mean()) < 5e-3


@pytest.mark.gpu
@pytest.mark.parametrize("T,d", [(1, 4096), (64, 4096), (7, 768), (300, 8192)])
def test_vllm_dispatches_to_kge_kernels(ops, T, d):
    plug, rms_norm, fused = ops
    g = torch.Generator().manual_seed(T + d)
    x = torch.randn(T, d, generator=g).to(torch.bfloat16).cuda()
    r = torch.randn(T, d, generator=g).to(torch.bfloat16).cuda()
    w = (1 + 0.1 * torch.randn(d, generator=g)).to(torch.bfloat16).cuda()
    assert plug._rms_ok(x, w, 1e-5) and plug._fused_ok(x, r, w, 1e-5)

    def ulp_ok(a, b):
        ai, bi = a.view(torch.int16).int(), b.view(torch.int16).int()
        diff = (ai - bi).abs()
        return int(diff.max()) <= 1 and float((diff > 0).float().mean()) < 5e-3

    with rms_norm.set_priority(["native"]):
        ref = rms_norm(x, w, 1e-5)
    with rms_norm.set_priority(["kge", "native"]):
        got = rms_norm(x, w, 1e-5)
    assert ulp_ok(got, ref)
    with fused.set_priority(["native"]):
        ref_y, ref_r = fused(x.clone(), r.clone(), w, 1e-5)
    with fused.set_priority(["kge", "native"]):
        got_y, got_r = fused(x.clone(), r.clone(), w, 1e-5)
    assert torch.equal(got_r, ref_r) and ulp_ok(got_y, ref_y)
    x32 = x.float()                                                     # uncovered dtype: vLLM falls through to native
    with rms_norm.set_priority(["kge", "native"]):
        assert rms_norm(x32, w.float(), 1e-5).dtype == torch.float32
