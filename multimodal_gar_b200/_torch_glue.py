"""Small helpers shared by the extension-module shims: tensor checks, stream, device guard."""
import torch

from . import _lib


def need_cuda(**tensors):
    """The reference prints and exit(-1)s on a bad tensor (e.g. ball_query.cpp:14-26); we raise."""
    for name, t in tensors.items():
        if t is None:
            continue
        if not isinstance(t, torch.Tensor):
            raise TypeError("%s must be a torch.Tensor" % name)
        if not t.is_cuda:
            raise _lib.MgarError("%s must be a CUDA tensor (libmgar_lidar has no CPU fallback)" % name)
        if not t.is_contiguous():
            raise _lib.MgarError("%s must be contiguous" % name)


def need_dtype(dtype, **tensors):
    for name, t in tensors.items():
        if t is not None and t.dtype != dtype:
            raise _lib.MgarError("%s must be %s, got %s" % (name, dtype, t.dtype))


def ptr(t):
    return 0 if t is None else t.data_ptr()


def stream_of(t):
    return torch.cuda.current_stream(t.device).cuda_stream


class on_device:
    """CUDAGuard: make the tensor's device current for the launch (the reference has none)."""

    def __init__(self, t):
        self.guard = torch.cuda.device(t.device)

    def __enter__(self):
        self.guard.__enter__()

    def __exit__(self, *a):
        return self.guard.__exit__(*a)
