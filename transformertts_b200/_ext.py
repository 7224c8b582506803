"""ctypes binding of the C ABI in include/tts_b200.h (libtts_b200.so, built in-tree by build.py).

There is deliberately no fallback: if the CUDA library is missing or a call fails, this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtts_b200.so")

vp, fp, i32, f32 = C.c_void_p, C.c_void_p, C.c_int32, C.c_float   # device pointers travel as void*


class AttnW(C.Structure):
    _fields_ = [("w_in", vp), ("b_in", fp), ("w_out", vp), ("b_out", fp)]


class FfnW(C.Structure):
    _fields_ = [("w1", vp), ("b1", fp), ("w2", vp), ("b2", fp)]


class LnW(C.Structure):
    _fields_ = [("g", fp), ("b", fp)]


class EncLayerW(C.Structure):
    _fields_ = [("sa", AttnW), ("ffn", FfnW), ("ln1", LnW), ("ln2", LnW)]


class DecLayerW(C.Structure):
    _fields_ = [("sa", AttnW), ("ca", AttnW), ("ffn", FfnW), ("ln1", LnW), ("ln2", LnW), ("ln3", LnW)]


class Weights(C.Structure):
    _fields_ = [
        ("emb", vp), ("n_phon", i32), ("reserved0", i32),
        ("enc_conv_w", vp * 3), ("enc_conv_b", fp * 3),
        ("enc_lin_w", vp), ("enc_lin_b", fp),
        ("enc", EncLayerW * 3),
        ("pre1_w", vp), ("pre1_b", fp),
        ("pre2_w", vp), ("pre2_b", fp),
        ("dec", DecLayerW * 3),
        ("head_w", vp), ("head_b", fp),
        ("post_w", vp * 5), ("post_b", fp * 5),
        ("pe", fp), ("alpha", f32), ("reserved1", i32),
        ("ckv_w", vp), ("ckv_b", fp),
        ("dec_kt", (vp * 3) * 3), ("pre2_kt", vp),
    ]


# name -> (restype, argtypes); mirrors include/tts_b200.h one to one
SIGNATURES = {
    "tts_abi_version": (C.c_int, []),
    "tts_last_error": (C.c_char_p, []),
    "tts_launch_count": (C.c_uint64, []),
    "tts_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "tts_forward": (C.c_int, [C.POINTER(Weights), vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                              vp, vp, vp, vp, vp, C.c_size_t, vp]),
    "tts_encode": (C.c_int, [C.POINTER(Weights), vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_size_t, vp]),
    "tts_decode": (C.c_int, [C.POINTER(Weights), vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp, vp, vp, vp, vp,
                             C.c_int, C.c_int, vp, vp]),
    "tts_decode_max_clusters": (C.c_int, [C.c_int, C.c_int]),
    "tts_decode_capacity": (C.c_int, []),
    "tts_postnet": (C.c_int, [C.POINTER(Weights), vp, C.c_int, C.c_int, vp, vp, C.c_size_t, vp]),
    "tts_teacher_forcing_mix": (C.c_int, [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp]),
    "tts_loss_scratch_bytes": (C.c_size_t, []),
    "tts_loss": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, vp, vp, C.c_size_t, vp]),
    "tts_loss_static": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_float, vp, vp]),
    "tts_pack_outputs": (C.c_int, [vp, vp, vp, C.c_longlong, vp, vp]),
    "tts_op_linear": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_float, C.c_int, vp, vp, vp,
                                vp, vp, vp]),
    "tts_op_linear_chain": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_int, vp, C.c_float, C.c_int,
                                      vp, vp, vp]),
    "tts_op_conv1d": (C.c_int, [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]),
    "tts_op_ffn": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, C.c_int, vp, vp]),
    "tts_op_attention": (C.c_int, [vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int,
                                   vp, vp, vp]),
}

_lib = None


def load():
    """Load libtts_b200.so (raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build the CUDA extension first (python -m transformertts_b200.build). "
                "There is no CPU or PyTorch fallback for this path.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        if lib.tts_abi_version() != 3:
            raise RuntimeError("libtts_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def check(status: int, what: str):
    if status != 0:
        raise RuntimeError(f"{what}: {load().tts_last_error().decode()}")


def ptr(t):
    """Device (or host) address of a torch tensor, or NULL."""
    return None if t is None else C.c_void_p(t.data_ptr())
