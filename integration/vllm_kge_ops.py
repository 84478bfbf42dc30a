"""vLLM op provider "kge": routes vLLM's own IR ops through libkge.so.

This is the reference-side binding for the engine the reference deploys: vLLM 0.22 resolves `vllm.ir.ops.rms_norm` and
`fused_add_rms_norm` through a provider registry (vllm/ir/op.py: IrOp.register_impl, set_default / set_priority).
Importing this module registers the sm_100a kernels of libkge.so as provider "kge"; a deployment then puts it in front:

    import integration.vllm_kge_ops                      # e.g. from a vLLM general plugin entry point
    from vllm.ir.ops.layernorm import rms_norm, fused_add_rms_norm
    rms_norm.set_default(["kge", "native"])
    fused_add_rms_norm.set_default(["kge", "native"])

Arguments the kernels do not cover (non-bf16, non-contiguous, variance_size, no weight, hidden size not a multiple of 8
or above 16384) fall through to vLLM's native implementation via `supports_args`.  The library path is
$KGE_LIB or the in-tree build.  The kernels follow vLLM's cast points (weight multiply in the weight dtype, residual =
rounded sum), so the outputs are bit-equal to `native` except where fp32 reduction order differs (<= 1 bf16 ulp).
"""
import ctypes as C
import os

import torch
from torch import Tensor
from vllm.ir.ops.layernorm import fused_add_rms_norm, rms_norm

_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        here = os.path.dirname(os.path.abspath(__file__))
        path = os.environ.get("KGE_LIB") or os.path.join(here, "..", "kubernetes-gpu-cluster_b200", "libkge.so")
        lib = C.CDLL(path)
        lib.kge_last_error.restype = C.c_char_p
        P, I, F = C.c_void_p, C.c_int, C.c_float
        lib.kge_rmsnorm.argtypes = [P, P, P, I, I, F, P]
        lib.kge_add_rmsnorm.argtypes = [P, P, P, P, I, I, F, P]
        lib.kge_rmsnorm.restype = lib.kge_add_rmsnorm.restype = I
        _LIB = lib
    return _LIB


def _check(rc, what):
    if rc:
        raise RuntimeError(f"{what}: {_lib().kge_last_error().decode()}")


def _covered(x, weight, variance_size):
    d = x.shape[-1] if x.dim() else 0
    return (variance_size is None and weight is not None and x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
            and x.is_contiguous() and weight.is_contiguous() and x.dim() >= 1 and d % 8 == 0 and 0 < d <= 16384 and weight.numel() == d)


def _rms_ok(x, weight, epsilon, variance_size=None):
    return _covered(x, weight, variance_size)


def _fused_ok(x, x_residual, weight, epsilon, variance_size=None):
    return (_covered(x, weight, variance_size) and x_residual.shape == x.shape and x_residual.dtype == x.dtype
            and x_residual.is_cuda and x_residual.is_contiguous())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@rms_norm.register_impl("kge", supported=torch.cuda.is_available(), supports_args=_rms_ok)
def rms_norm_kge(x: Tensor, weight: Tensor | None, epsilon: float, variance_size: int | None = None) -> Tensor:
    d = x.shape[-1]
    y = torch.empty_like(x)
    if x.numel():
        _check(_lib().kge_rmsnorm(y.data_ptr(), x.data_ptr(), weight.data_ptr(), x.numel() // d, d, float(epsilon), _stream()), "kge_rmsnorm")
    return y


@fused_add_rms_norm.register_impl("kge", supported=torch.cuda.is_available(), supports_args=_fused_ok)
def fused_add_rms_norm_kge(x: Tensor, x_residual: Tensor, weight: Tensor | None, epsilon: float,
                           variance_size: int | None = None) -> tuple[Tensor, Tensor]:
    d = x.shape[-1]
    y = torch.empty_like(x)
    res = x_residual.clone()              # kge_add_rmsnorm updates the residual in place; vLLM's op returns a new one
    if x.numel():
        _check(_lib().kge_add_rmsnorm(y.data_ptr(), res.data_ptr(), x.data_ptr(), weight.data_ptr(), x.numel() // d, d, float(epsilon),
                                      _stream()), "kge_add_rmsnorm")
    return y, res


# ---- SwiGLU activation: vLLM's out-of-tree layer registry (vllm/model_executor/custom_op.py: CustomOp.register_oot).
# Every `SiluAndMul()` a model builds after this import is a KgeSiluAndMul; inputs the kernel does not cover use vLLM's own
# forward.
from vllm.model_executor.custom_op import CustomOp  # noqa: E402
from vllm.model_executor.layers.activation import SiluAndMul  # noqa: E402


def _silu_covered(x):
    f = x.shape[-1] // 2 if x.dim() else 0
    return x.is_cuda and x.dtype == torch.bfloat16 and x.is_contiguous() and x.shape[-1] % 2 == 0 and f % 8 == 0 and f > 0


@CustomOp.register_oot(name="SiluAndMul")
class KgeSiluAndMul(SiluAndMul):
    def forward_cuda(self, x: Tensor) -> Tensor:
        if not _silu_covered(x):
            return SiluAndMul.forward_cuda(self, x)
        f = x.shape[-1] // 2
        out = torch.empty(x.shape[:-1] + (f,), dtype=x.dtype, device=x.device)
        if x.numel():
            lib = _lib()
            lib.kge_silu_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
            lib.kge_silu_mul.restype = C.c_int
            _check(lib.kge_silu_mul(out.data_ptr(), x.data_ptr(), x.numel() // (2 * f), f, _stream()), "kge_silu_mul")
        return out

    forward_oot = forward_cuda
