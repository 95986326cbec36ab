"""Drop-in for the reference's `asy_tools` pybind module (/root/reference/src/dagr/asynchronous/asy_tools/main.cu:239-244;
call sites asynchronous/conv.py:236-238, linear.py:46-48, batch_norm.py:40, base/utils.py:42-43): same four function names
and argument lists, running the sm_100a kernels of csrc/asy.cu through the C ABI.  CUDA tensors only (the reference
asserts the same through CHECK_CUDA, main.cu:8-11)."""
import torch

from .. import _lib


def _chk(t, name, dtype):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}")
    return _lib.ptr(t)


def masked_lin(indices, x_in, x_out, weight, bias, add):
    L = _lib.lib()
    _lib.check(L.evad_asy_masked_lin(_chk(indices, "indices", torch.int64), indices.shape[0], _chk(x_in, "x_in", torch.float32),
                                     _chk(x_out, "x_out", torch.float32), _chk(weight.detach(), "weight", torch.float32),
                                     _chk(bias.detach(), "bias", torch.float32), weight.shape[1], weight.shape[0], int(bool(add)),
                                     _lib.stream()), "evad_asy_masked_lin")


def masked_lin_no_bias(indices, x_in, x_out, weight, add):
    L = _lib.lib()
    _lib.check(L.evad_asy_masked_lin(_chk(indices, "indices", torch.int64), indices.shape[0], _chk(x_in, "x_in", torch.float32),
                                     _chk(x_out, "x_out", torch.float32), _chk(weight.detach(), "weight", torch.float32), None,
                                     weight.shape[1], weight.shape[0], int(bool(add)), _lib.stream()), "evad_asy_masked_lin")


def masked_inplace_BN(indices, x, x_out, running_mean, running_var, weight, bias, eps):
    L = _lib.lib()
    _lib.check(L.evad_asy_masked_bn(_chk(indices, "indices", torch.int64), indices.shape[0], _chk(x, "x", torch.float32),
                                    _chk(x_out, "x_out", torch.float32), _chk(running_mean, "running_mean", torch.float32),
                                    _chk(running_var, "running_var", torch.float32), _chk(weight.detach(), "weight", torch.float32),
                                    _chk(bias.detach(), "bias", torch.float32), x.shape[1], float(eps), _lib.stream()),
               "evad_asy_masked_bn")


def masked_isdiff(indices, x_old, x_new, atol, rtol):
    """Returns the entries of `indices` whose rows differ; like the reference, `indices` is overwritten (-1 = unchanged)."""
    L = _lib.lib()
    _lib.check(L.evad_asy_masked_isdiff(_chk(indices, "indices", torch.int64), indices.shape[0], _chk(x_old, "x_old", torch.float32),
                                        _chk(x_new, "x_new", torch.float32), x_old.shape[1], float(atol), float(rtol),
                                        _lib.stream()), "evad_asy_masked_isdiff")
    return indices[indices > -1]
