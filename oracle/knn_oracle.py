"""ORACLE — test infrastructure only.  ctypes wrappers of oracle/knn_oracle.c (our C restatement
of lib/knn/src/cuda/knn.cu:88-91,127-173 and lib/knn/src/cpu/knn_cpu.cpp:8-43) and, when built,
of the reference's own unmodified knn_cpu (oracle/_ref/libknn_cpu_ref.so)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(_HERE, "libknn_oracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libknn_cpu_ref.so")


def build(quiet=True):
    """Compile the C restatement (and oracle/_ref when /root/reference is present)."""
    out = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


def _load(path):
    if not os.path.exists(path):
        build()
    return C.CDLL(path)


_oracle = None
_ref = None


def knn_oracle(ref, query, k=1, contract=True, want_second=False):
    """ref (D, n_ref) f32, query (D, n_query) f32 -> idx (k, n_query) i64 1-based, dist (k, n_query)
    [, second (n_query,)].  contract=True follows the reference CUDA kernel's fma arithmetic,
    contract=False the reference CPU build's separate multiply/add."""
    global _oracle
    if _oracle is None:
        _oracle = _load(_ORACLE_SO)
        _oracle.knn_oracle.restype = C.c_int
        _oracle.knn_oracle.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    ref = np.ascontiguousarray(ref, dtype=np.float32)
    query = np.ascontiguousarray(query, dtype=np.float32)
    d, n_ref = ref.shape
    n_query = query.shape[1]
    idx = np.zeros((k, n_query), dtype=np.int64)
    dist = np.zeros((k, n_query), dtype=np.float32)
    second = np.zeros((n_query,), dtype=np.float32) if want_second else None
    rc = _oracle.knn_oracle(ref.ctypes.data, n_ref, query.ctypes.data, n_query, d, k,
                            int(bool(contract)), idx.ctypes.data, dist.ctypes.data,
                            second.ctypes.data if want_second else None)
    if rc != 0:
        raise ValueError("knn_oracle: bad arguments")
    return (idx, dist, second) if want_second else (idx, dist)


def have_reference_knn():
    return os.path.exists(_REF_SO)


def knn_reference_cpu(ref, query, k=1):
    """The reference's own knn_cpu (O(n_ref^2) per query — keep sizes small)."""
    global _ref
    if _ref is None:
        if not os.path.exists(_REF_SO):
            raise FileNotFoundError(_REF_SO + " (build with `make -C oracle` where /root/reference exists)")
        _ref = C.CDLL(_REF_SO)
        _ref.ref_knn_cpu.restype = C.c_int
        _ref.ref_knn_cpu.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    ref = np.ascontiguousarray(ref, dtype=np.float32)
    query = np.ascontiguousarray(query, dtype=np.float32)
    d, n_ref = ref.shape
    n_query = query.shape[1]
    idx = np.zeros((k, n_query), dtype=np.int64)
    rc = _ref.ref_knn_cpu(ref.ctypes.data, n_ref, query.ctypes.data, n_query, d, k, idx.ctypes.data)
    assert rc == 0
    return idx


_REF_CUDA_SO = os.path.join(_HERE, "_ref", "libknn_cuda_ref.so")
_ref_cuda = None


def have_reference_cuda_knn():
    return os.path.exists(_REF_CUDA_SO)


def _ref_cuda_lib():
    global _ref_cuda
    if _ref_cuda is None:
        if not os.path.exists(_REF_CUDA_SO):
            raise FileNotFoundError(_REF_CUDA_SO + " (build with `make -C oracle` where /root/reference exists)")
        _ref_cuda = C.CDLL(_REF_CUDA_SO)
        _ref_cuda.ref_knn_cuda.restype = C.c_int
        _ref_cuda.ref_knn_cuda.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p, C.c_void_p]
        _ref_cuda.ref_knn_cuda_scratch_bytes.restype = C.c_size_t
        _ref_cuda.ref_knn_cuda_scratch_bytes.argtypes = [C.c_int, C.c_int]
    return _ref_cuda


def knn_reference_cuda(ref, query, k=1, scratch=None, return_dist=False):
    """The reference's OWN CUDA kernels (lib/knn/src/cuda/knn.cu: cuComputeDistanceGlobal + cuInsertionSort through
    knn_device, compiled unmodified for sm_100a) on torch CUDA tensors ref (B,D,n_ref), query (B,D,n_query), driven
    like lib/knn/src/knn.h:29-41.  Returns idx (B,k,n_query) i64 1-based [, the first k rows of the distance matrix
    of the LAST batch element].  `scratch`: optional float32 CUDA tensor of >= n_ref*n_query elements (re-used
    across calls the way torch's caching allocator re-uses the reference's THCudaMalloc block)."""
    import torch
    L = _ref_cuda_lib()
    assert ref.is_cuda and query.is_cuda and ref.dtype == torch.float32 and query.dtype == torch.float32
    ref, query = ref.contiguous(), query.contiguous()
    B, D, n_ref = ref.shape
    n_query = query.shape[2]
    if scratch is None:
        scratch = torch.empty(n_ref * n_query, dtype=torch.float32, device=ref.device)
    assert scratch.numel() >= n_ref * n_query
    idx = torch.empty((B, k, n_query), dtype=torch.int64, device=ref.device)
    with torch.cuda.device(ref.device):
        rc = L.ref_knn_cuda(ref.data_ptr(), query.data_ptr(), idx.data_ptr(), B, D, n_ref, n_query, k,
                            scratch.data_ptr(), torch.cuda.current_stream(ref.device).cuda_stream)
    if rc != 0:
        raise RuntimeError(f"reference knn.cu failed with cudaError {rc}")
    if return_dist:
        return idx, scratch[: k * n_query].view(k, n_query)
    return idx
