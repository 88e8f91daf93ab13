"""ctypes binding of librmem_b200.so (the C ABI in include/rmem_b200.h).

There is no fallback: if the shared library is missing or an entry point is absent, importing the
ops raises.  ``build()`` compiles the library in-tree with nvcc for sm_100a (no GPU needed).
"""
import ctypes
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "librmem_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "rmem_b200.h")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
]


def sources():
    return sorted(os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith(".cu"))


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC)] + [HEADER_PATH]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """nvcc -> rmem_ocu_b200/librmem_b200.so (cross-compiles without a GPU)."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + os.environ.get("RMEM_NVCC_EXTRA", "").split() + ["-o", LIB_PATH] + sources()
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr, file=sys.stderr)
    return LIB_PATH


_lib = None

_c_void_p = ctypes.c_void_p
_c_int = ctypes.c_int
_c_float = ctypes.c_float

# name -> argtypes; every function returns int (status) unless listed in _RESTYPES
_c_float_p = ctypes.POINTER(ctypes.c_float)
_SIGNATURES = {
    "rmem_abi_version": [],
    "rmem_last_error": [],
    "rmem_lt_attn_deaot_partial": [_c_void_p] * 4 + [_c_int] * 4 + [_c_void_p] * 4,
    "rmem_lt_attn_deaot_combine": [_c_void_p] * 5 + [_c_int] * 3 + [_c_void_p] * 4,
    "rmem_lt_attn_deaot_single": [_c_void_p] * 3 + [_c_int] + [_c_void_p] * 3,
    "rmem_lt_attn_deaot_sk_workspace_bytes": [_c_int, _c_int, _c_int, _c_int],
    "rmem_lt_attn_deaot_sk_partial": [_c_void_p] * 4 + [_c_int] * 5 + [_c_void_p, _c_void_p, _c_int, _c_int, _c_void_p,
                                                                     ctypes.c_int64, _c_void_p],
    "rmem_lt_attn_deaot_sk_combine": [_c_int, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_int, _c_void_p,
                                      ctypes.c_int64, _c_void_p],
    "rmem_bank_tick_batched": [_c_void_p, _c_int, _c_void_p],
    "rmem_logits_head_batched": [_c_void_p, _c_int] + [_c_void_p] * 3 + [_c_int] * 4 + [_c_void_p] * 2 +
                                [_c_int, _c_void_p, _c_int] + [_c_void_p] * 2 + [_c_int] * 2 + [_c_void_p] * 3,
    "rmem_layernorm_copy": [_c_void_p, _c_int] + [_c_void_p] * 3 + [_c_int] * 2 + [_c_void_p] + [_c_int] * 2 + [_c_void_p],
    "rmem_gn_upsample_add_nhwc": [_c_void_p] * 4 + [_c_int] * 4 + [_c_void_p] * 3 + [_c_int] * 4 + [_c_void_p] * 2,
    "rmem_groupnorm_nhwc_ex": [_c_void_p] * 5 + [_c_int] * 6 + [_c_void_p] * 2,
    "rmem_upsample_add_nhwc_ex": [_c_void_p, _c_int, _c_int] + [_c_void_p] * 3 + [_c_int] * 5 + [_c_void_p],
    "rmem_id_embed_soft": [_c_void_p, _c_int, _c_void_p, _c_int, _c_int] + [_c_void_p] * 4 + [_c_int] + [_c_void_p] * 3 +
                          [_c_int] * 4 + [_c_void_p],
    "rmem_id_embed_batched": [_c_void_p, _c_int, _c_int] + [_c_void_p] * 4 + [_c_int] + [_c_void_p] * 3 + [_c_int] * 4 +
                             [_c_void_p],
    "rmem_groupnorm_nhwc_batched": [_c_void_p] * 5 + [_c_int] * 5 + [_c_void_p] * 2,
    "rmem_upsample_add_nhwc_batched": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_int, _c_int,
                                       _c_int, _c_int, _c_void_p],
    "rmem_finish_logits_batched": [_c_void_p] + [ctypes.c_int64] * 4 + [_c_void_p, _c_void_p] + [_c_int] * 3 +
                                  [_c_void_p, _c_int, _c_void_p, _c_int, _c_int, _c_void_p],
    "rmem_pack_image_s2d_batched": [_c_void_p, _c_int, _c_float_p, _c_float_p, _c_int, _c_int, _c_void_p, _c_int,
                                    _c_int, _c_void_p],
    "rmem_maxpool3s2_nhwc_batched": [_c_void_p, _c_int, _c_int, _c_int, _c_void_p, _c_int, _c_void_p, _c_int, _c_int,
                                     _c_void_p],
    "rmem_argmax_label_batched": [_c_void_p, _c_int, _c_int, _c_void_p, _c_int, _c_void_p],
    "rmem_dwconv5_batched": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p, _c_int, _c_int, _c_int, _c_int,
                             _c_int, _c_int, _c_void_p],
    "rmem_groupnorm_tokens_batched": [_c_void_p, _c_int] + [_c_void_p] * 4 + [_c_int] * 5 + [_c_void_p] * 2,
    "rmem_local_attn_scores_batched": [_c_void_p] * 3 + [_c_int, _c_void_p, _c_void_p, _c_int, _c_int, _c_int, _c_void_p],
    "rmem_local_attn_aggregate_batched": [_c_void_p] * 4 + [_c_int, _c_int, _c_int, _c_void_p],
    "rmem_mha_partial": [_c_void_p, _c_int, _c_void_p, _c_int, ctypes.c_int64, _c_void_p, _c_int, ctypes.c_int64,
                         _c_void_p] + [_c_int] * 4 + [_c_void_p] * 4,
    "rmem_mha_combine": [_c_void_p] * 5 + [_c_int] * 3 + [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p],
    "rmem_tpe_prepare_mha": [_c_void_p, _c_int] + [_c_void_p] * 4 + [_c_int, _c_void_p],
    "rmem_layernorm_ex": [_c_void_p, _c_int, _c_void_p, _c_int, _c_int, _c_void_p, _c_void_p,
                          _c_void_p, _c_int, _c_int, _c_void_p, _c_int, _c_void_p, _c_int, _c_int,
                          _c_void_p, _c_int, _c_int, _c_int, _c_int, _c_void_p],
    "rmem_groupnorm_gelu_tokens": [_c_void_p] * 4 + [_c_int] * 3 + [_c_void_p] * 2,
    "rmem_linear": [_c_void_p, _c_int, _c_void_p, _c_int, _c_void_p, _c_int, _c_int, _c_int,
                    _c_void_p, _c_int, _c_void_p],
    "rmem_linear_ex": [_c_void_p, _c_int, _c_void_p, _c_int, _c_void_p, _c_int, _c_int, _c_int,
                       _c_void_p, _c_int, _c_int, _c_void_p],
    "rmem_layernorm": [_c_void_p] * 4 + [_c_int] * 2 + [_c_void_p] * 4 + [_c_int] * 5 + [_c_void_p],
    "rmem_dwconv5": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p, _c_int, _c_int, _c_int, _c_int,
                     _c_int, _c_void_p],
    "rmem_groupnorm_tokens": [_c_void_p, _c_int] + [_c_void_p] * 4 + [_c_int] * 4 + [_c_void_p] * 2,
    "rmem_groupnorm_tokens_static": [_c_void_p, _c_int] + [_c_void_p] * 4 + [_c_int] * 4 + [_c_void_p],
    "rmem_tpe_prepare": [_c_void_p] * 5 + [_c_int] * 2 + [_c_void_p],
    "rmem_local_banded_elems": [_c_int, _c_int],
    "rmem_local_attn_scores": [_c_void_p] * 3 + [_c_int, _c_void_p, _c_void_p, _c_int, _c_int, _c_void_p],
    "rmem_local_attn_aggregate": [_c_void_p] * 4 + [_c_int, _c_int, _c_void_p],
    "rmem_local_attn": [_c_void_p] * 4 + [_c_int, _c_void_p, _c_void_p, _c_int, _c_int, _c_void_p],
    "rmem_groupnorm_nhwc": [_c_void_p] * 5 + [_c_int] * 4 + [_c_void_p] * 2,
    "rmem_upsample_add_nhwc": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_int, _c_int, _c_int,
                               _c_void_p],
    "rmem_finish_logits": [_c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, _c_void_p, _c_void_p] +
                          [_c_int] * 4 + [_c_void_p, _c_int, _c_int, _c_void_p],
    "rmem_pack_image": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p],
    "rmem_pack_image_s2d": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p],
    "rmem_pack_image_s2d_ex": [_c_void_p, _c_int, _c_float_p, _c_float_p, _c_int, _c_int, _c_void_p, _c_int, _c_void_p],
    "rmem_maxpool3s2_nhwc_ex": [_c_void_p, _c_int, _c_int, _c_int, _c_void_p, _c_int, _c_void_p, _c_int, _c_void_p],
    "rmem_maxpool3s2_nhwc": [_c_void_p, _c_int, _c_int, _c_int, _c_void_p, _c_int, _c_void_p, _c_void_p],
    "rmem_argmax_label": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p],
    "rmem_bank_init": [_c_void_p, _c_int, _c_void_p],
    "rmem_bank_tick": [_c_void_p, _c_void_p],
    "rmem_id_embed": [_c_void_p, _c_int, _c_int] + [_c_void_p] * 4 + [_c_int] + [_c_void_p] * 3 + [_c_int] * 3 + [_c_void_p],
    "rmem_bank_append": [_c_void_p, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_int, _c_void_p],
    "rmem_bank_score_evict": [_c_void_p] * 3 + [_c_int] * 3 + [_c_void_p],
    "rmem_foreground_prob": [_c_void_p] + [_c_int] * 3 + [_c_void_p, _c_int, _c_int, _c_void_p],
}


class GemmSegment(ctypes.Structure):
    """rmem_gemm_segment (include/rmem_b200.h)."""
    _fields_ = [("col_begin", ctypes.c_int32), ("col_end", ctypes.c_int32), ("act", ctypes.c_int32),
                ("mode", ctypes.c_int32), ("dst", ctypes.c_void_p), ("ld", ctypes.c_int32),
                ("dst_col0", ctypes.c_int32)]


ABI_VERSION = 11          # must equal RMEM_ABI_VERSION of the library that gets loaded (include/rmem_b200.h)
MAX_BANK = 16
MAX_CHUNKS_PER_FRAME = 4
BANK_STATE_WORDS = 5 * MAX_BANK + 8
ACT_NONE, ACT_SILU = 0, 1
OUT_BF16, OUT_F32, OUT_F32_ACCUM = 0, 1, 2
_RESTYPES = {"rmem_last_error": ctypes.c_char_p, "rmem_local_banded_elems": ctypes.c_int64,
             "rmem_lt_attn_deaot_sk_workspace_bytes": ctypes.c_int64}


def declared_symbols():
    """Entry points declared in include/rmem_b200.h (parsed, so tests can check the export list)."""
    import re
    with open(HEADER_PATH) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rmem_[a-z0-9_]+)\s*\(", text)))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "librmem_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; "
            "g.build()'`. There is no CPU or PyTorch fallback for the RMem hot path." % LIB_PATH)
    handle = ctypes.CDLL(LIB_PATH)
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(handle, name)  # AttributeError if the symbol is missing: fail loudly
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, _c_int)
    got = handle.rmem_abi_version()
    if got != ABI_VERSION:
        # a stale binary (the .so is git-ignored and shipped out of band) would be called with mismatched
        # signatures / struct layouts and corrupt memory silently
        raise RuntimeError("librmem_b200.so has ABI version %d, the Python binding expects %d: rebuild it "
                           "(python -c 'import __graft_entry__ as g; g.build()')" % (got, ABI_VERSION))
    _lib = handle
    return _lib


class RmemError(RuntimeError):
    pass


def check(status, what):
    if status != 0:
        msg = lib().rmem_last_error()
        raise RmemError("%s failed (%d): %s" % (what, status, msg.decode() if msg else "?"))
