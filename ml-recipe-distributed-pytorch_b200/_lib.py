"""ctypes binding of libbqa.so (C ABI declared in include/bqa.h).

The library is the product: there is no Python or CPU fallback behind these calls.  If the
shared object is missing the import fails with instructions to build it; if a call fails the
C side's message is raised as ``RuntimeError``.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbqa.so")
ABI_VERSION = 4

EPI_LINEAR, EPI_GELU_FWD, EPI_GELU_BWD, EPI_F32_ADD = 0, 1, 2, 3
CLS_SMOOTH, CLS_CE, CLS_FOCAL = 0, 1, 2

c_i32, c_i64, c_u32, c_u64, c_f32 = ctypes.c_int32, ctypes.c_int64, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_float
vp = ctypes.c_void_p


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("M", c_i32), ("N", c_i32), ("K", c_i32),
        ("A", vp), ("lda", c_i32), ("a_mn_major", c_i32),
        ("B", vp), ("ldb", c_i32), ("b_mn_major", c_i32),
        ("C", vp), ("ldc", c_i32),
        ("C2", vp), ("epilogue", c_i32),
        ("bias", vp),
        ("residual", vp), ("ld_res", c_i32),
        ("aux", vp), ("ld_aux", c_i32),
        ("k_splits", c_i32),
        ("dropout_p", c_f32),
        ("drop_seed", c_u64), ("drop_site", c_u32),
        ("colsum_out", vp), ("dot_with", vp), ("ld_dot", c_i32), ("dot_out", vp), ("dot_rows", c_i32),
    ]


class Config(ctypes.Structure):
    _fields_ = [("num_layers", c_i32), ("hidden", c_i32), ("heads", c_i32), ("intermediate", c_i32),
                ("vocab", c_i32), ("max_pos", c_i32), ("type_vocab", c_i32), ("num_labels", c_i32),
                ("ln_eps", c_f32), ("hidden_dropout", c_f32), ("attn_dropout", c_f32),
                ("pad_token_id", c_i32), ("position_mode", c_i32)]


LAYOUT_FIELDS = ["word_emb", "pos_emb", "type_emb", "emb_ln_g", "emb_ln_b", "layer0", "layer_stride",
                 "qkv_w", "qkv_b", "ao_w", "ao_b", "ln1_g", "ln1_b", "i_w", "i_b", "o_w", "o_b", "ln2_g", "ln2_b",
                 "pool_w", "pool_b", "span_w", "span_b", "cls_w", "cls_b", "regs_w", "regs_b", "rege_w", "rege_b",
                 "total"]


class Layout(ctypes.Structure):
    _fields_ = [(n, c_i64) for n in LAYOUT_FIELDS]


class LossConfig(ctypes.Structure):
    _fields_ = [("w_start", c_f32), ("w_end", c_f32), ("w_start_reg", c_f32), ("w_end_reg", c_f32),
                ("w_cls", c_f32), ("cls_loss", c_i32), ("smooth_alpha", c_f32), ("cls_weight", vp),
                ("loss_scale", c_f32), ("focal_alpha", c_f32), ("focal_gamma", c_f32)]


class AdamWChunk(ctypes.Structure):
    _fields_ = [("begin", c_u64), ("end", c_u64), ("weight_decay", c_f32), ("group", c_i32)]


# name -> (restype, argtypes); the CPU test-suite checks that every symbol of include/bqa.h is here
SIGNATURES = {
    "bqa_last_error": (ctypes.c_char_p, []),
    "bqa_abi_version": (ctypes.c_int, []),
    "bqa_gemm_bf16": (ctypes.c_int, [ctypes.POINTER(GemmDesc), vp]),
    "bqa_attention_keep_words": (ctypes.c_int, [c_i32]),
    "bqa_attention_forward": (ctypes.c_int, [vp, vp, vp, vp, vp, c_i32, c_i32, c_i32, c_f32, c_u64, c_u32, vp]),
    "bqa_attention_backward": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, c_i32, c_i32, c_i32, c_f32, c_u64,
                                              c_u32, vp, vp]),
    "bqa_layernorm_forward": (ctypes.c_int, [vp, vp, vp, vp, c_i32, c_i32, c_f32, vp]),
    "bqa_layernorm_backward": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, c_i32, c_i32, c_f32, c_f32, c_u32,
                                              c_f32, c_u32, c_u64, vp]),
    "bqa_colsum_bf16": (ctypes.c_int, [vp, c_i32, vp, c_i32, c_i32, vp]),
    "bqa_sumsq_scratch_floats": (ctypes.c_int, []),
    "bqa_grad_sumsq": (ctypes.c_int, [vp, c_u64, vp, vp, vp]),
    "bqa_grad_sumsq_noalloc": (ctypes.c_int, [vp, c_u64, vp, vp]),
    "bqa_clip_adamw": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, c_i32, vp, c_f32, ctypes.POINTER(c_f32), c_i32,
                                      c_f32, c_f32, c_f32, c_i32, vp]),
    "bqa_clip_adamod": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, c_i32, vp, c_f32, ctypes.POINTER(c_f32), c_i32,
                                       c_f32, c_f32, c_f32, c_f32, c_i64, c_i32, vp]),
    "bqa_f32_to_bf16": (ctypes.c_int, [vp, vp, c_u64, vp]),
    "bqa_decode": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, c_i32, c_i32, c_i32, vp]),
    "bqa_prepare_inputs": (ctypes.c_int, [vp, c_i32, c_i32, c_i32, c_i32, vp, vp, vp, vp, vp]),
    "bqa_param_layout": (ctypes.c_int, [ctypes.POINTER(Config), ctypes.POINTER(Layout)]),
    "bqa_workspace_bytes": (ctypes.c_int, [ctypes.POINTER(Config), c_i32, c_i32, c_i32, ctypes.POINTER(c_u64)]),
    "bqa_create": (ctypes.c_int, [ctypes.POINTER(Config), ctypes.POINTER(vp)]),
    "bqa_destroy": (ctypes.c_int, [vp]),
    "bqa_bind": (ctypes.c_int, [vp, vp, vp, vp, vp, c_u64]),
    "bqa_forward": (ctypes.c_int, [vp, vp, vp, vp, c_i32, c_i32, c_i32, c_u64, c_u64, vp, vp, vp, vp, vp, vp]),
    "bqa_loss": (ctypes.c_int, [vp] + [vp] * 10 + [ctypes.POINTER(LossConfig), c_i32, vp, vp]),
    "bqa_qa_loss": (ctypes.c_int, [vp] * 17 + [c_i32, c_i32, c_i32, ctypes.POINTER(LossConfig), c_i32, vp]),
    "bqa_input_status": (ctypes.c_int, [vp, ctypes.POINTER(c_i32), vp]),
    "bqa_set_output_grads": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp]),
    "bqa_backward": (ctypes.c_int, [vp, c_i32, c_i32, vp]),
    "bqa_num_backward_phases": (ctypes.c_int, [vp]),
    "bqa_phase_grad_begin": (c_i64, [vp, c_i32]),
    "bqa_launch_count": (c_i64, [vp]),
    "bqa_profile_begin": (ctypes.c_int, [vp]),
    "bqa_profile_end": (ctypes.c_int, [vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                       ctypes.POINTER(c_i64)]),
}


def _load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(needs nvcc; there is no fallback implementation)")
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = library and binding out of sync
        fn.restype = res
        fn.argtypes = args
    if lib.bqa_abi_version() != ABI_VERSION:
        raise ImportError(f"{path}: ABI version {lib.bqa_abi_version()} != expected {ABI_VERSION}; rebuild")
    return lib


lib = _load()


def check(rc, what):
    if rc != 0:
        msg = lib.bqa_last_error()
        raise RuntimeError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def current_stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def gemm_bf16(A, B, C, *, a_mn_major=False, b_mn_major=False, epilogue=EPI_LINEAR, C2=None, bias=None,
              residual=None, aux=None, k_splits=1, dropout_p=0.0, drop_seed=0, drop_site=0, M=None, N=None,
              K=None, colsum_out=None, dot_with=None, dot_out=None, dot_rows=0):
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
    d.colsum_out = colsum_out.data_ptr() if colsum_out is not None else None
    d.dot_with = dot_with.data_ptr() if dot_with is not None else None
    d.ld_dot = dot_with.stride(0) if dot_with is not None else 0
    d.dot_out = dot_out.data_ptr() if dot_out is not None else None
    d.dot_rows = dot_rows
    check(lib.bqa_gemm_bf16(ctypes.byref(d), current_stream_ptr()), "bqa_gemm_bf16")
    return C
