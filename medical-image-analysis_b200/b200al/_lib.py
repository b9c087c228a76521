"""ctypes binding of libb200al.so (include/b200al.h).

The shared library is built in-tree by `b200al.build.build()` (nvcc, sm_100a).
There is NO fallback: if the library is missing or a call fails, a
RuntimeError is raised - nothing on the product path computes on the CPU.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200al.so")

NUM_BLOCKS = 18
NUM_UPSAMPLES = 4
MASK_CHANNELS = 2944
EMBED_DIM = 512
TOPK_STATIC_MAX_N = 262144  # B200AL_TOPK_STATIC_MAX_N

ENTROPY, MARGIN, CONFIDENCE, DICE = 0, 1, 2, 3
L2, COSINE = 0, 1
BF16X3, BF16, MIXED, MIXED_DEEP = 0, 1, 2, 3
U8, F32 = 0, 1

MODES = {"entropy": ENTROPY, "margin": MARGIN, "confidence": CONFIDENCE, "dice": DICE}
METRICS = {"l2": L2, "cosine": COSINE}
CRIT_MIN, CRIT_MEAN = 0, 1
CRITERIA = {"min": CRIT_MIN, "mean": CRIT_MEAN}
PRECISIONS = {"bf16x3": BF16X3, "bf16": BF16, "mixed": MIXED, "mixed-deep": MIXED_DEEP}


class BlockWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("d_conv_weight", "d_conv_bias", "d_bn_gamma", "d_bn_beta", "d_bn_mean", "d_bn_var")]


class Weights(C.Structure):
    _fields_ = [
        ("blocks", BlockWeights * NUM_BLOCKS),
        ("d_up_weight", C.c_void_p * NUM_UPSAMPLES),
        ("d_up_bias", C.c_void_p * NUM_UPSAMPLES),
        ("d_head_weight", C.c_void_p),
        ("d_head_bias", C.c_void_p),
        ("bn_eps", C.c_float),
        ("leaky_slope", C.c_float),
    ]


# name -> (restype, argtypes); must list every symbol include/b200al.h declares
PROTOTYPES = {
    "b200al_ctx_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "b200al_ctx_destroy": (None, [C.c_void_p]),
    "b200al_last_error": (C.c_char_p, [C.c_void_p]),
    "b200al_load_weights": (C.c_int, [C.c_void_p, C.POINTER(Weights), C.c_void_p]),
    "b200al_prep_images": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_kcenter_fused": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_standardize_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "b200al_kmeans_pp_scratch_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "b200al_kmeans_pp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                   C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_badge_embed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p,
                                     C.c_int, C.c_void_p]),
    "b200al_unet_scores": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float,
                                     C.c_uint64, C.c_int64, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_unet_embed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200al_uncertainty_from_logits": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                 C.c_float, C.c_void_p, C.c_void_p]),
    "b200al_workspace_bytes_topk": (C.c_size_t, [C.c_int]),
    "b200al_workspace_bytes_kcenter_fused": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "b200al_workspace_reserve": (C.c_int, [C.c_void_p, C.c_size_t]),
    "b200al_topk": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_topk_static": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_kcenter_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                      C.c_void_p, C.c_int, C.c_void_p]),
    "b200al_kcenter_run": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200al_kcenter_mailbox_bytes": (C.c_size_t, [C.c_int]),
    "b200al_kcenter_run_sharded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                             C.c_int, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                             C.POINTER(C.c_int64), C.c_int, C.c_void_p]),
    "b200al_normalize_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "b200al_kernel_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "b200al_conv_time_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_double)]),
    "b200al_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "b200al_debug_layer": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200al_debug_layer_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "b200al_debug_layer_times": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_int]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library and bind every prototype.  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  b200al has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, ctx=None) -> None:
    if rc != 0:
        msg = load().b200al_last_error(ctx)
        raise RuntimeError(f"libb200al error {rc}: {msg.decode() if msg else '?'}")
