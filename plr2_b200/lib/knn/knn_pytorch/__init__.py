"""Stand-in for the reference's compiled pybind module ``lib.knn.knn_pytorch.knn_pytorch``
(lib/knn/src/vision.cpp:3-5): ``knn_pytorch.knn(ref, query, idx)`` fills the caller-allocated
int64 ``idx`` (B, k, n_query) in place and returns 1 (lib/knn/src/knn.h:12-42).  With this module
registered as ``lib.knn.knn_pytorch`` the UNMODIFIED reference wrapper (lib/knn/__init__.py:7,19)
runs on the B200 kernel."""
import sys

import torch

from .... import _cabi


def knn(ref, query, idx, mode=0):
    if not (ref.is_cuda and query.is_cuda and idx.is_cuda):
        raise RuntimeError("Not compiled with CPU support: plr2_b200 knn needs CUDA tensors")
    if ref.dtype != torch.float32 or query.dtype != torch.float32:
        raise RuntimeError("expected scalar type Float")  # data_ptr<float>() in knn.h:22-23
    if idx.dtype != torch.int64:
        raise RuntimeError("expected scalar type Long")   # data_ptr<long>() in knn.h:24
    b, d, n_ref = ref.shape
    k, n_query = idx.shape[1], query.shape[2]
    if not idx.is_contiguous():
        raise RuntimeError("idx must be contiguous (written in place, knn.h:24)")
    # bind the contiguous copies to locals: a temporary would go back to the caching allocator (and could be handed
    # to the next .contiguous()) before the kernel that reads it has been enqueued
    r, q = ref.contiguous(), query.contiguous()
    with torch.cuda.device(ref.device):
        st = _cabi.lib().plr2_knn_f32(_cabi.ptr(r), _cabi.ptr(q), _cabi.ptr(idx), None, b, d, n_ref, n_query, k,
                                      mode, None, 0, _cabi.stream_ptr(ref.device))
    del r, q
    _cabi.check(st, "plr2_knn_f32")
    return 1


knn_pytorch = sys.modules[__name__]


def install_as_reference_module():
    """Register under the names the reference imports (``from lib.knn.knn_pytorch import knn_pytorch``)."""
    sys.modules.setdefault("lib.knn.knn_pytorch", sys.modules[__name__])
