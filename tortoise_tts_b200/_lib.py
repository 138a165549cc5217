"""ctypes binding of include/tts_ar.h (libtts_ar.so, built in-tree by __graft_entry__.build()).

There is no CPU fallback: if the shared library is missing or the device is not an sm_100 GPU the
import / create call raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libtts_ar.so")

SYMBOLS = [
    "tts_ar_abi_version", "tts_ar_create", "tts_ar_destroy", "tts_ar_last_error", "tts_ar_load_weights",
    "tts_ar_prefill", "tts_ar_generate", "tts_ar_step", "tts_ar_logits_debug", "tts_ar_forward_latents",
    "tts_ar_inference_speech_host", "tts_ar_sync_result", "tts_ar_debug_copy", "tts_ar_fixup_codes",
    "tts_ar_debug_gemm", "tts_ar_set_option", "tts_ar_stream_begin", "tts_ar_stream_poll", "tts_ar_stream_end",
    "tts_ar_gemm_f32_noalloc",
    "tts_ar_clvp_create", "tts_ar_clvp_destroy", "tts_ar_clvp_last_error", "tts_ar_clvp_load_weights", "tts_ar_clvp_score",
    "tts_ar_clvp_sync", "tts_ar_clvp_topk", "tts_ar_load_conditioning_weights", "tts_ar_get_conditioning", "tts_ar_mel_spectrogram",
]

ABI_VERSION = 2
STATUS = {0: "OK", -1: "BAD_ARG", -2: "CAPACITY", -3: "CUDA", -4: "STATE", -5: "DEVICE", -6: "WATCHDOG"}


class TtsArConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "struct_size", "device", "n_layers", "model_dim", "n_heads", "n_mel_codes", "n_text_rows", "n_mel_pos",
        "n_text_pos", "start_text_token", "stop_text_token", "start_mel_token", "stop_mel_token", "max_utts",
        "max_seqs", "max_text_len", "max_new_tokens")]


class TtsArLayerWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "ln_1_w", "ln_1_b", "c_attn_w", "c_attn_b", "attn_proj_w", "attn_proj_b", "ln_2_w", "ln_2_b",
        "c_fc_w", "c_fc_b", "mlp_proj_w", "mlp_proj_b")]


class TtsArWeights(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("dtype", C.c_int32), ("n_layers", C.c_int32),
                ("layers", C.POINTER(TtsArLayerWeights))] + [(n, C.c_void_p) for n in (
                    "ln_f_w", "ln_f_b", "final_norm_w", "final_norm_b", "mel_head_w", "mel_head_b",
                    "mel_embedding", "mel_pos_embedding", "text_embedding", "text_pos_embedding")]


class TtsArSampling(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("do_sample", C.c_int32), ("temperature", C.c_float),
                ("top_k", C.c_int32), ("top_p", C.c_float), ("repetition_penalty", C.c_float),
                ("seed", C.c_uint64), ("seq_index_base", C.c_int64), ("ignore_eos", C.c_int32),
                ("typical_mass", C.c_float), ("seq_indices", C.c_void_p)]


class TtsArStats(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("steps_run", C.c_int32), ("all_finished", C.c_int32),
                ("watchdog", C.c_int32), ("kernel_launches", C.c_int64), ("last_algorithmic_bytes", C.c_double),
                ("last_device_ms", C.c_float), ("last_prefill_ms", C.c_float),
                ("kv_pages", C.c_int32), ("ctas", C.c_int32)]


class TtsArCondBlockWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("norm_w", "norm_b", "qkv_w", "qkv_b", "proj_w", "proj_b")]


class TtsArCondWeights(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("dtype", C.c_int32), ("n_blocks", C.c_int32), ("spec_dim", C.c_int32),
                ("init_w", C.c_void_p), ("init_b", C.c_void_p), ("blocks", C.POINTER(TtsArCondBlockWeights))]


class TtsClvpConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "struct_size", "device", "dim", "heads", "depth", "n_text_tokens", "n_speech_tokens", "max_seqs", "max_text_len",
        "max_speech_len")]


class TtsClvpLayerWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "attn_norm_g", "to_q_w", "to_k_w", "to_v_w", "to_out_w", "to_out_b", "ff_norm_g", "ff_proj_w", "ff_proj_b",
        "ff_out_w", "ff_out_b")]


class TtsClvpEncoderWeights(C.Structure):
    _fields_ = [("emb", C.c_void_p), ("layers", C.POINTER(TtsClvpLayerWeights)), ("norm_w", C.c_void_p),
                ("norm_b", C.c_void_p), ("to_latent_w", C.c_void_p)]


class TtsClvpWeights(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("dtype", C.c_int32), ("depth", C.c_int32),
                ("text", TtsClvpEncoderWeights), ("speech", TtsClvpEncoderWeights), ("temperature", C.c_void_p)]


_lib = None


def load():
    """Loads libtts_ar.so (raises if it has not been built -- there is no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a). tortoise_tts_b200 has no CPU / PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64p, f32p, i32p = C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p
    lib.tts_ar_abi_version.restype = C.c_int
    lib.tts_ar_create.argtypes = [C.POINTER(TtsArConfig), C.POINTER(vp)]
    lib.tts_ar_destroy.argtypes = [vp]
    lib.tts_ar_destroy.restype = None
    lib.tts_ar_last_error.argtypes = [vp]
    lib.tts_ar_last_error.restype = C.c_char_p
    lib.tts_ar_load_weights.argtypes = [vp, C.POINTER(TtsArWeights), vp]
    lib.tts_ar_prefill.argtypes = [vp, f32p, i32p, i32, C.POINTER(C.c_int32), i32, i32, vp]
    lib.tts_ar_generate.argtypes = [vp, C.POINTER(TtsArSampling), i32, i64p, f32p, i32p, i32p, vp]
    lib.tts_ar_step.argtypes = [vp, C.POINTER(TtsArSampling), i32, i64p, f32p, i32p, i32p, vp]
    lib.tts_ar_logits_debug.argtypes = [vp, f32p, C.POINTER(C.c_int32), vp]
    lib.tts_ar_forward_latents.argtypes = [vp, f32p, i32p, i32, C.POINTER(C.c_int32), i32p, i32,
                                           C.POINTER(C.c_int32), i32, f32p, vp]
    lib.tts_ar_inference_speech_host.argtypes = [vp, f32p, i32p, i32, C.POINTER(C.c_int32), i32, i32,
                                                 C.POINTER(TtsArSampling), i32, i64p, C.POINTER(C.c_int32), vp]
    lib.tts_ar_debug_gemm.argtypes = [vp, vp, vp, f32p, f32p, i32, i32, i32, i32, vp]
    lib.tts_ar_gemm_f32_noalloc.argtypes = [f32p, vp, f32p, f32p, i32, i32, i32, i32, i32p, vp]
    lib.tts_ar_load_conditioning_weights.argtypes = [vp, C.POINTER(TtsArCondWeights), vp]
    lib.tts_ar_get_conditioning.argtypes = [vp, f32p, i32, i32, f32p, vp]
    lib.tts_ar_mel_spectrogram.argtypes = [vp, f32p, i32, i32, f32p, f32p, vp]
    lib.tts_ar_clvp_create.argtypes = [C.POINTER(TtsClvpConfig), C.POINTER(vp)]
    lib.tts_ar_clvp_destroy.argtypes = [vp]
    lib.tts_ar_clvp_destroy.restype = None
    lib.tts_ar_clvp_last_error.argtypes = [vp]
    lib.tts_ar_clvp_last_error.restype = C.c_char_p
    lib.tts_ar_clvp_load_weights.argtypes = [vp, C.POINTER(TtsClvpWeights), vp]
    lib.tts_ar_clvp_score.argtypes = [vp, i32p, i32, i32, i32, i64p, i32, i32, i32, f32p, f32p, vp]
    lib.tts_ar_clvp_topk.argtypes = [vp, f32p, i32, i32, i64p, vp]
    lib.tts_ar_clvp_sync.argtypes = [vp, C.POINTER(C.c_int64), vp]
    lib.tts_ar_set_option.argtypes = [vp, C.c_char_p, i32]
    lib.tts_ar_stream_begin.argtypes = [vp, C.POINTER(TtsArSampling), i32, i64p, f32p, vp]
    lib.tts_ar_stream_poll.argtypes = [vp, i32, i32]
    lib.tts_ar_stream_end.argtypes = [vp, C.POINTER(TtsArStats), vp]
    lib.tts_ar_sync_result.argtypes = [vp, C.POINTER(TtsArStats), vp]
    lib.tts_ar_debug_copy.argtypes = [vp, C.c_char_p, vp, C.c_uint64, vp]
    lib.tts_ar_fixup_codes.argtypes = [vp, i64p, i32, i32, i32, i32, i32p, i32p, vp]
    for s in SYMBOLS:
        getattr(lib, s)
        if s not in ("tts_ar_destroy", "tts_ar_last_error", "tts_ar_clvp_destroy", "tts_ar_clvp_last_error"):
            getattr(lib, s).restype = C.c_int
    assert lib.tts_ar_abi_version() == ABI_VERSION
    _lib = lib
    return lib


def check(lib, handle, rc, what):
    """C status -> the exception the reference would raise (AssertionError/ValueError become ValueError)."""
    if rc == 0:
        return
    msg = lib.tts_ar_last_error(handle)
    msg = msg.decode() if msg else ""
    text = f"{what}: {STATUS.get(rc, rc)}: {msg}"
    if rc in (-1, -2):
        raise ValueError(text)
    raise RuntimeError(text)
