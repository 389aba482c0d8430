"""ctypes binding of the C-ABI library (include/cbx.h).  Fails loudly: there is no fallback path."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CBX_LIBCBX") or os.path.join(HERE, "libcbx.so")     # CBX_LIBCBX: A/B another build of the library


class CbxError(RuntimeError):
    pass


class CbxT3Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("n_layers", C.c_int32), ("max_pos", C.c_int32), ("kv_fp32", C.c_int32),
                ("speech_pos_rows", C.c_int32)]


class CbxT3Weights(C.Structure):
    _fields_ = [("wstream", C.c_void_p), ("stride_units", C.c_int64), ("ln1", C.c_void_p), ("ln2", C.c_void_p),
                ("final_norm", C.c_void_p), ("speech_emb", C.c_void_p), ("speech_pos_emb", C.c_void_p),
                ("rope_cos", C.c_void_p), ("rope_sin", C.c_void_p), ("wplain", C.c_void_p)]


class CbxSampleParams(C.Structure):
    _fields_ = [("cfg_weight", C.c_float), ("temperature", C.c_float), ("repetition_penalty", C.c_float),
                ("min_p", C.c_float), ("top_p", C.c_float), ("greedy", C.c_int32), ("stop_on_eos", C.c_int32),
                ("seed", C.c_uint64)]


class CbxT3FrontWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("spkr_w", "spkr_b", "query", "norm_g", "norm_b", "wq", "bq", "wk", "bk", "wv", "bv",
                                          "wo", "bo", "emotion_w", "text_emb", "text_pos_emb", "speech_emb", "speech_pos_emb")]


class CbxT3BatchConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("n_layers", C.c_int32), ("n_slots", C.c_int32), ("max_pos", C.c_int32),
                ("kv_fp32", C.c_int32), ("max_new_cap", C.c_int32), ("keep_logits", C.c_int32),
                ("speech_pos_rows", C.c_int32)]


class CbxCfmWeights(C.Structure):
    _fields_ = [("bf16_blob", C.c_void_p), ("f32_blob", C.c_void_p), ("bf16_off", C.POINTER(C.c_int64)),
                ("f32_off", C.POINTER(C.c_int64))]


class CbxFlowFrontWeights(C.Structure):
    _fields_ = [("bf16_blob", C.c_void_p), ("f32_blob", C.c_void_p), ("bf16_off", C.POINTER(C.c_int64)),
                ("f32_off", C.POINTER(C.c_int64))]


class CbxHiftWeights(C.Structure):
    _fields_ = [("bf16_blob", C.c_void_p), ("f32_blob", C.c_void_p), ("bf16_off", C.POINTER(C.c_int64)),
                ("f32_off", C.POINTER(C.c_int64))]


# every symbol include/cbx.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "cbx_last_error": (C.c_char_p, []),
    "cbx_version": (C.c_int, []),
    "cbx_t3_grid": (C.c_int, [C.c_int]),
    "cbx_t3_kwarps": (C.c_int, []),
    "cbx_t3_plan_count": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "cbx_t3_plan_unit": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "cbx_t3_create": (C.c_int, [C.POINTER(CbxT3Config), C.POINTER(CbxT3Weights), C.POINTER(C.c_void_p)]),
    "cbx_t3_destroy": (C.c_int, [C.c_void_p]),
    "cbx_t3_prefill": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "cbx_t3_decode": (C.c_int, [C.c_void_p, C.POINTER(CbxSampleParams), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_void_p]),
    "cbx_t3_set_profile": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "cbx_t3_set_alignment": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cbx_t3_front_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(CbxT3FrontWeights), C.POINTER(C.c_void_p)]),
    "cbx_t3_front_destroy": (C.c_int, [C.c_void_p]),
    "cbx_t3_front": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                               C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "cbx_t3_front_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "cbx_t3_batch_create": (C.c_int, [C.POINTER(CbxT3BatchConfig), C.POINTER(CbxT3Weights), C.POINTER(C.c_void_p)]),
    "cbx_t3_batch_destroy": (C.c_int, [C.c_void_p]),
    "cbx_t3_batch_prefill": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(CbxSampleParams), C.c_int,
                                       C.c_void_p, C.c_void_p]),
    "cbx_t3_batch_prefill_many": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(CbxSampleParams), C.c_int,
                                            C.c_void_p, C.c_void_p]),
    "cbx_t3_decode_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_void_p]),
    "cbx_t3_batch_poll": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                    C.c_void_p]),
    "cbx_t3_batch_tokens": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "cbx_t3_batch_logits": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "cbx_t3_batch_release": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "cbx_t3_batch_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "cbx_cfm_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(CbxCfmWeights), C.POINTER(C.c_void_p)]),
    "cbx_cfm_destroy": (C.c_int, [C.c_void_p]),
    "cbx_cfm_estimator": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int,
                                    C.c_void_p, C.c_void_p]),
    "cbx_cfm_solve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float, C.c_void_p, C.c_void_p]),
    "cbx_cfm_create_batched": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(CbxCfmWeights), C.POINTER(C.c_void_p)]),
    "cbx_cfm_solve_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int32),
                                      C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float, C.c_void_p,
                                      C.c_void_p]),
    "cbx_cfm_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "cbx_cfm_cache_sizes": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "cbx_flowfe_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(CbxFlowFrontWeights), C.POINTER(C.c_void_p)]),
    "cbx_flowfe_destroy": (C.c_int, [C.c_void_p]),
    "cbx_flowfe_run": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p]),
    "cbx_flowfe_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "cbx_hift_create": (C.c_int, [C.c_int, C.c_int, C.POINTER(CbxHiftWeights), C.POINTER(C.c_void_p)]),
    "cbx_hift_destroy": (C.c_int, [C.c_void_p]),
    "cbx_hift_f0": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "cbx_hift_source": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                  C.c_void_p]),
    "cbx_hift_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "cbx_hift_inference": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "cbx_hift_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "cbx_hift_source_static_max_frames": (C.c_int, []),
    "cbx_hift_source_static": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_uint64, C.c_void_p, C.c_void_p]),
    "cbx_t3_state": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                               C.POINTER(C.c_int64)]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libcbx.so (built in-tree by chatterbox_b200/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CbxError(f"{LIB_PATH} is missing: run `python -m chatterbox_b200.build` "
                           "(chatterbox_b200 has no CPU or PyTorch fallback)")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise CbxError(lib().cbx_last_error().decode())
