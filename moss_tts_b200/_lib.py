"""ctypes binding of libmosstts_sm100.so (declarations mirror include/mosstts.h).

The library is the product: there is no Python/CPU fallback.  If the shared object is
missing or a call fails, a `MossTTSNativeError` is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmosstts_sm100.so")
CSRC = os.path.join(_HERE, "csrc")

MT_OK, MT_ERR_ARG, MT_ERR_CUDA, MT_ERR_UNSUPPORTED = 0, -1, -2, -3
MT_DTYPE_BF16, MT_DTYPE_F32, MT_DTYPE_F16 = 0, 1, 2
MT_GEMM_AUTO, MT_GEMM_SIMT, MT_GEMM_W_STATIC, MT_GEMM_W_TILED = 0, 1, 2, 4
MT_CODEC_MAX_STAGES = 8
MT_GEMM_STATIC_SCRATCH_BYTES = 8 * 1024 * 1024

EXPORTS = [
    "mt_version", "mt_last_error", "mt_launch_count", "mt_debug_force_simt", "mt_debug_set_pdl", "mt_debug_set_attn_cluster", "mt_debug_set_trace", "mt_codec_debug_set_absmax", "mt_codec_debug_set_attention", "mt_gemm_debug_set_pair", "mt_gemm", "mt_gemm_static_scratch",
    "mt_lm_query_sizes", "mt_lm_create", "mt_lm_destroy", "mt_lm_forward", "mt_lm_embed", "mt_lm_generate",
    "mt_lm_stream_begin", "mt_lm_stream_steps", "mt_lm_stream_end", "mt_stream_gather_frames", "mt_crossfade",
    "mt_sample_tokens", "mt_sample_tokens_full", "mt_sample_full_scratch_bytes", "mt_find_last_equal", "mt_apply_delay_pattern", "mt_apply_de_delay_pattern",
    "mt_codec_create", "mt_codec_destroy", "mt_codec_workspace_bytes", "mt_rvq_fold_tables",
    "mt_rvq_dequant", "mt_codec_decode", "mt_rlfq_quantize", "mt_codec_encode",
    "mt_codec_stream_bytes", "mt_codec_stream_begin", "mt_codec_stream_push", "mt_codec_stream_push_encode", "mt_codec_stream_end",
    "mt_loudness_normalize", "mt_pack_unified_codes", "mt_pad_left",
    "mt_local_query_sizes", "mt_local_create", "mt_local_destroy", "mt_local_forward", "mt_local_generate",
    "mt_rt_query_sizes", "mt_rt_create", "mt_rt_destroy", "mt_rt_prefill", "mt_rt_step", "mt_rt_finish", "mt_rt_state",
    "mt_rt_local_generate", "mt_rt_last_hidden",
]


class MossTTSNativeError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [("dtype_in", C.c_int), ("dtype_out", C.c_int), ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
                ("x", C.c_void_p), ("ldx", C.c_longlong), ("w", C.c_void_p), ("ldw", C.c_longlong),
                ("out", C.c_void_p), ("ldo", C.c_longlong), ("bias", C.c_void_p), ("act", C.c_int),
                ("residual", C.c_void_p), ("ldr", C.c_longlong), ("scale", C.c_void_p), ("round_bf16", C.c_int),
                ("workspace", C.c_void_p), ("workspace_bytes", C.c_longlong), ("impl", C.c_int),
                ("out_splits", C.POINTER(C.c_int))]


class LMConfigC(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("hidden_size", "intermediate_size", "num_layers", "num_heads", "num_kv_heads",
                                       "head_dim", "vocab_size", "n_vq", "audio_head_size")] + \
               [("rms_norm_eps", C.c_float)] + \
               [(n, C.c_int) for n in ("pad_token_id", "im_start_token_id", "im_end_token_id", "audio_start_token_id",
                                       "audio_end_token_id", "audio_assistant_gen_slot_token_id",
                                       "audio_assistant_delay_slot_token_id", "audio_pad_code", "max_batch", "max_seq",
                                       "max_tokens", "weights_tiled")]


class LMWeightsC(C.Structure):
    _fields_ = [("embed_tokens", C.c_void_p), ("emb_ext", C.c_void_p),
                ("w_qkv", C.POINTER(C.c_void_p)), ("w_o", C.POINTER(C.c_void_p)),
                ("w_gate_up", C.POINTER(C.c_void_p)), ("w_down", C.POINTER(C.c_void_p)),
                ("ln_in", C.POINTER(C.c_void_p)), ("ln_post", C.POINTER(C.c_void_p)),
                ("q_norm", C.POINTER(C.c_void_p)), ("k_norm", C.POINTER(C.c_void_p)),
                ("final_norm", C.c_void_p), ("text_head", C.c_void_p), ("audio_heads", C.c_void_p),
                ("rope_cos", C.c_void_p), ("rope_sin", C.c_void_p)]


class GenParamsC(C.Structure):
    _fields_ = [("max_new_tokens", C.c_int), ("text_temperature", C.c_float), ("text_top_p", C.c_float),
                ("text_top_k", C.c_int), ("audio_temperature", C.c_float), ("audio_top_p", C.c_float),
                ("audio_top_k", C.c_int), ("audio_repetition_penalty", C.c_float), ("seed", C.c_uint),
                ("use_cuda_graph", C.c_int), ("check_stop_every", C.c_int)]


class LocalConfigC(C.Structure):
    _fields_ = [("lm", LMConfigC)] + [(n, C.c_int) for n in ("local_hidden_size", "local_ffn_hidden_size",
                                                            "local_num_layers", "additional_mlp_ffn_hidden_size")]


class LocalWeightsC(C.Structure):
    _fields_ = [("lm", LMWeightsC)] + \
               [(n, C.POINTER(C.c_void_p)) for n in ("l_w_qkv", "l_w_o", "l_w_gate_up", "l_w_down", "l_ln_in", "l_ln_post",
                                                     "l_q_norm", "l_k_norm")] + \
               [("l_final_norm", C.c_void_p), ("in_gate_up", C.c_void_p), ("in_down", C.c_void_p)] + \
               [(n, C.POINTER(C.c_void_p)) for n in ("out_gate_up", "out_down", "head_norm", "heads")]


class LocalGenParamsC(C.Structure):
    _fields_ = [("max_new_tokens", C.c_int), ("n_vq_for_inference", C.c_int), ("eos_token_id", C.c_int),
                ("do_sample", C.POINTER(C.c_int)), ("temperature", C.POINTER(C.c_float)), ("top_p", C.POINTER(C.c_float)),
                ("top_k", C.POINTER(C.c_int)), ("repetition_penalty", C.POINTER(C.c_float)), ("seed", C.c_uint),
                ("use_cuda_graph", C.c_int), ("check_stop_every", C.c_int), ("attention_mask", C.c_void_p)]


class RtConfigC(C.Structure):
    _fields_ = [("lm", LMConfigC)] + [(n, C.c_int) for n in ("local_num_layers", "local_intermediate_size",
                                                            "local_num_heads", "local_num_kv_heads", "local_head_dim")] + \
               [("local_rms_norm_eps", C.c_float), ("max_frames", C.c_int), ("audio_eos_token", C.c_int),
                ("text_pad_id", C.c_longlong)]


class RtWeightsC(C.Structure):
    _fields_ = [("lm", LMWeightsC)] + \
               [(n, C.POINTER(C.c_void_p)) for n in ("l_w_qkv", "l_w_o", "l_w_gate_up", "l_w_down", "l_ln_in", "l_ln_post",
                                                     "l_q_norm", "l_k_norm")] + \
               [("l_final_norm", C.c_void_p), ("l_embed", C.c_void_p), ("l_heads", C.POINTER(C.c_void_p)),
                ("l_rope_cos", C.c_void_p), ("l_rope_sin", C.c_void_p)]


class RtSampleParamsC(C.Structure):
    _fields_ = [("do_sample", C.c_int), ("temperature", C.c_float), ("top_p", C.c_float), ("top_k", C.c_int),
                ("repetition_penalty", C.c_float), ("repetition_window", C.c_int), ("seed", C.c_uint)]


_IntArr = C.c_int * MT_CODEC_MAX_STAGES


class CodecConfigC(C.Structure):
    _fields_ = [("num_stages", C.c_int)] + \
               [(n, _IntArr) for n in ("input_dim", "d_model", "num_heads", "num_layers", "dim_ff", "output_dim",
                                       "patch", "context")] + \
               [(n, C.c_int) for n in ("code_dim", "rvq_dim", "num_quantizers", "codebook_size", "codebook_dim")] + \
               [("max_period", C.c_float), ("encoder", C.c_int), ("layer_gemm_f16", C.c_int)]


class CodecQuantWeightsC(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("in_w", "in_b", "q_in_w", "q_in_b", "codebook", "codebook_n", "codebook_nsq",
                                          "q_out_w", "q_out_b")]


class CodecLayerC(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("norm1_w", "norm1_b", "norm2_w", "norm2_b", "in_proj", "out_proj",
                                          "linear1", "linear2", "scale1", "scale2")]


class CodecWeightsC(C.Structure):
    _fields_ = [("rvq_table", C.c_void_p), ("rvq_out_bias", C.c_void_p),
                ("input_proj", C.c_void_p * MT_CODEC_MAX_STAGES), ("output_proj", C.c_void_p * MT_CODEC_MAX_STAGES),
                ("layers", C.POINTER(CodecLayerC) * MT_CODEC_MAX_STAGES),
                ("quant", C.POINTER(CodecQuantWeightsC))]


_lib = None


def build(force: bool = False) -> str:
    """Compile the shared library in-tree with nvcc for sm_100a (works without a GPU)."""
    if force or not os.path.exists(LIB_PATH) or _stale():
        subprocess.run(["make", "-C", CSRC, "-j8"], check=True, capture_output=not bool(os.environ.get("MT_VERBOSE")))
    return LIB_PATH


def _stale() -> bool:
    try:
        t = os.path.getmtime(LIB_PATH)
        srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))]
        srcs.append(os.path.join(os.path.dirname(_HERE), "include", "mosstts.h"))
        return any(os.path.getmtime(s) > t for s in srcs)
    except OSError:
        return True


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MossTTSNativeError(
                f"{LIB_PATH} is missing: build it with `make -C moss_tts_b200/csrc` "
                "(or __graft_entry__.build()); there is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.mt_last_error.restype = C.c_char_p
        L.mt_launch_count.restype = C.c_longlong
        L.mt_sample_full_scratch_bytes.restype = C.c_longlong
        if os.environ.get("MT_FORCE_SIMT") == "1":      # test-only: cross-check everything on the SIMT GEMM
            L.mt_debug_force_simt(1)
        _lib = L
    return _lib


def check(status: int, what: str = "") -> None:
    if status != MT_OK:
        msg = lib().mt_last_error().decode("utf-8", "replace")
        if status == MT_ERR_ARG:
            raise ValueError(f"{what}: {msg}")
        if status == MT_ERR_UNSUPPORTED:
            raise NotImplementedError(f"{what}: {msg}")
        raise MossTTSNativeError(f"{what}: {msg}")


def ptr(t) -> C.c_void_p:
    """Device (or host) address of a torch tensor, None -> NULL."""
    return C.c_void_p(None if t is None else t.data_ptr())


def stream_ptr(device=None) -> C.c_void_p:
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr_array(tensors):
    arr = (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    return arr


def launch_count() -> int:
    return int(lib().mt_launch_count())
