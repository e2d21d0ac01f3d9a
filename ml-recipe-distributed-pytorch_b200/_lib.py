"""ctypes binding of libbqa.so (C ABI declared in include/bqa.h).

The library is the product: there is no Python or CPU fallback behind these calls.  If the
shared object is missing the import fails with instructions to build it; if a call fails the
C side's message is raised as ``RuntimeError``.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbqa.so")

EPI_LINEAR, EPI_GELU_FWD, EPI_GELU_BWD, EPI_F32_ADD = 0, 1, 2, 3


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("M", ctypes.c_int32), ("N", ctypes.c_int32), ("K", ctypes.c_int32),
        ("A", ctypes.c_void_p), ("lda", ctypes.c_int32), ("a_mn_major", ctypes.c_int32),
        ("B", ctypes.c_void_p), ("ldb", ctypes.c_int32), ("b_mn_major", ctypes.c_int32),
        ("C", ctypes.c_void_p), ("ldc", ctypes.c_int32),
        ("C2", ctypes.c_void_p), ("epilogue", ctypes.c_int32),
        ("bias", ctypes.c_void_p),
        ("residual", ctypes.c_void_p), ("ld_res", ctypes.c_int32),
        ("aux", ctypes.c_void_p), ("ld_aux", ctypes.c_int32),
        ("k_splits", ctypes.c_int32),
        ("dropout_p", ctypes.c_float),
        ("drop_seed", ctypes.c_uint64), ("drop_site", ctypes.c_uint32),
    ]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(needs nvcc; there is no fallback implementation)")
    lib = ctypes.CDLL(LIB_PATH)
    lib.bqa_last_error.restype = ctypes.c_char_p
    lib.bqa_last_error.argtypes = []
    lib.bqa_abi_version.restype = ctypes.c_int
    lib.bqa_gemm_bf16.restype = ctypes.c_int
    lib.bqa_gemm_bf16.argtypes = [ctypes.POINTER(GemmDesc), ctypes.c_void_p]
    return lib


lib = _load()


def check(rc, what):
    if rc != 0:
        msg = lib.bqa_last_error()
        raise RuntimeError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def current_stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def gemm_bf16(A, B, C, *, a_mn_major=False, b_mn_major=False, epilogue=EPI_LINEAR, C2=None, bias=None,
              residual=None, aux=None, k_splits=1, dropout_p=0.0, drop_seed=0, drop_site=0, M=None, N=None,
              K=None):
    """C = epilogue(op(A) . op(B)) on the tcgen05 kernel.  Shapes follow include/bqa.h:
    A is [M,K] (K-major) or [K,M] (MN-major); B is [N,K] or [K,N]; C is [M,N]."""
    if M is None:
        M = A.shape[1] if a_mn_major else A.shape[0]
    if K is None:
        K = A.shape[0] if a_mn_major else A.shape[1]
    if N is None:
        N = B.shape[1] if b_mn_major else B.shape[0]
    d = GemmDesc()
    d.M, d.N, d.K = M, N, K
    d.A, d.lda, d.a_mn_major = A.data_ptr(), A.stride(0), int(a_mn_major)
    d.B, d.ldb, d.b_mn_major = B.data_ptr(), B.stride(0), int(b_mn_major)
    d.C, d.ldc = C.data_ptr(), C.stride(0)
    d.C2 = C2.data_ptr() if C2 is not None else None
    d.epilogue = epilogue
    d.bias = bias.data_ptr() if bias is not None else None
    d.residual = residual.data_ptr() if residual is not None else None
    d.ld_res = residual.stride(0) if residual is not None else 0
    d.aux = aux.data_ptr() if aux is not None else None
    d.ld_aux = aux.stride(0) if aux is not None else 0
    d.k_splits = k_splits
    d.dropout_p = dropout_p
    d.drop_seed = drop_seed
    d.drop_site = drop_site
    check(lib.bqa_gemm_bf16(ctypes.byref(d), current_stream_ptr()), "bqa_gemm_bf16")
    return C
