"""Drop-in for the reference's ``lib.knn`` (lib/knn/__init__.py:9-21).

``KNearestNeighbor.apply(ref, query, k)`` -> (B, k, n_query) int64, 1-BASED reference indices.
ref (B, D, n_ref), query (B, D, n_query).  Like the reference, inputs are converted to float32 on
the GPU unconditionally (lib/knn/__init__.py:14-15) and there is no backward.  Unlike
``torch.autograd.Function.apply`` it also accepts ``k=`` as a keyword, which the reference's own
refiner loss uses (lib/loss_refiner.py:43) and which fails on torch >= 2 with the original class.
"""
import torch

from ... import _cabi
from .knn_pytorch import knn_pytorch  # noqa: F401  (same import the reference performs, :7)

KNN_EXACT = 0
KNN_FAST = 1


def knn_indices(ref, query, k=1, mode=KNN_EXACT, return_dist=False):
    """Raw entry: contiguous float32 CUDA tensors in the reference layout."""
    if not ref.is_cuda:
        raise _cabi.Plr2Error("plr2_b200 KNN needs CUDA tensors (there is no CPU fallback)")
    assert ref.dim() == 3 and query.dim() == 3 and ref.shape[:2] == query.shape[:2]
    b, d, n_ref = ref.shape
    n_query = query.shape[2]
    inds = torch.empty((b, k, n_query), dtype=torch.int64, device=ref.device)
    dist = torch.empty((b, k, n_query), dtype=torch.float32, device=ref.device) if return_dist else None
    with torch.cuda.device(ref.device):
        st = _cabi.lib().plr2_knn_f32(_cabi.ptr(ref), _cabi.ptr(query), _cabi.ptr(inds), _cabi.ptr(dist),
                                      b, d, n_ref, n_query, k, mode, None, 0,
                                      _cabi.stream_ptr(ref.device))
    _cabi.check(st, "plr2_knn_f32")
    return (inds, dist) if return_dist else inds


class _KNearestNeighborFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ref, query, k):
        ref = ref.float().cuda().contiguous()
        query = query.float().cuda().contiguous()
        inds = knn_indices(ref, query, int(k), KNearestNeighbor.mode)
        ctx.mark_non_differentiable(inds)
        return inds

    @staticmethod
    def backward(ctx, *grads):  # the reference defines none; indices carry no gradient
        return None, None, None


class KNearestNeighbor:
    """Compute k nearest neighbors for each query point (same name and call as the reference)."""
    mode = KNN_EXACT  # class-wide arithmetic mode (KNN_EXACT | KNN_FAST)

    @staticmethod
    def apply(ref, query, k=1):
        return _KNearestNeighborFn.apply(ref, query, k)

    # the reference's (broken) unit test constructs the class directly; keep that working too
    def __new__(cls, ref, query, k=1):
        return cls.apply(ref, query, k)

    forward = staticmethod(lambda ctx, ref, query, k: _KNearestNeighborFn.forward(ctx, ref, query, k))
