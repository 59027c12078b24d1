"""Loader and build recipe for the UNMODIFIED reference code (TEST / BASELINE INFRASTRUCTURE ONLY).

The reference is a pure-Python package: "building" it means copying the modules the hot path
needs, byte for byte, from /root/reference into the git-ignored directory `oracle/_ref/` so that they travel to
the GPU box with the snapshot (the box has no /root/reference).  Nothing under `oracle/_ref/` is ever committed
and nothing in `moss_tts_b200/` imports it; `bench.py --impl reference` / `cpu_baseline` time it on the host
cores and `oracle/make_golden*.py` execute it to mint the fixtures.

    python -m oracle.ref            # (re)build oracle/_ref from /root/reference

Modules copied (reference @ fe397ee):
    moss_tts_delay/{modeling_moss_tts,inference_utils,configuration_moss_tts,processing_moss_tts}.py
    moss_audio_tokenizer/{__init__,modeling_moss_audio_tokenizer,configuration_moss_audio_tokenizer}.py + config.json
    moss_tts_realtime/mossttsrealtime/{__init__,configuration_mossttsrealtime,modeling_mossttsrealtime,
                                       modeling_mossttsrealtime_local,streaming_mossttsrealtime}.py
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("MOSS_REFERENCE", "/root/reference")
DST = os.path.join(HERE, "_ref")

FILES = [
    "LICENSE",
    "moss_tts_delay/modeling_moss_tts.py",
    "moss_tts_delay/inference_utils.py",
    "moss_tts_delay/configuration_moss_tts.py",
    "moss_tts_delay/processing_moss_tts.py",
    "moss_audio_tokenizer/__init__.py",
    "moss_audio_tokenizer/modeling_moss_audio_tokenizer.py",
    "moss_audio_tokenizer/configuration_moss_audio_tokenizer.py",
    "moss_audio_tokenizer/config.json",
    "moss_tts_realtime/mossttsrealtime/__init__.py",
    "moss_tts_realtime/mossttsrealtime/configuration_mossttsrealtime.py",
    "moss_tts_realtime/mossttsrealtime/modeling_mossttsrealtime.py",
    "moss_tts_realtime/mossttsrealtime/modeling_mossttsrealtime_local.py",
    "moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py",
]


def build(verbose: bool = False) -> bool:
    """Copy the reference modules into oracle/_ref (only where /root/reference exists: the build container).
    Returns True when oracle/_ref is complete afterwards."""
    if os.path.isdir(SRC):
        manifest = {}
        for rel in FILES:
            src = os.path.join(SRC, rel)
            if not os.path.exists(src):
                continue
            dst = os.path.join(DST, rel)
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            shutil.copyfile(src, dst)
            manifest[rel] = hashlib.sha256(open(src, "rb").read()).hexdigest()
        with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
            json.dump({"source": "xiami2019/MOSS-TTS@fe397ee (copied unmodified by oracle/ref.py)", "sha256": manifest}, f, indent=1)
        if verbose:
            print(f"oracle/_ref: {len(manifest)} files copied from {SRC}")
    return available()


def available() -> bool:
    return ref_root() is not None


def ref_root():
    """Directory holding the reference packages: /root/reference in the build container, else oracle/_ref."""
    for root in (SRC, DST):
        if os.path.exists(os.path.join(root, "moss_tts_delay", "modeling_moss_tts.py")):
            return root
    return None


def _ensure_path():
    root = ref_root()
    if root is None:
        raise RuntimeError("the reference is not available: neither /root/reference nor oracle/_ref exists "
                           "(run `python -m oracle.ref` in the build container)")
    if root not in sys.path:
        sys.path.insert(0, root)
    sys.dont_write_bytecode = True          # never leave __pycache__ next to the reference sources
    return root


def import_delay():
    """(MossTTSDelayConfig, MossTTSDelayModel, inference_utils module, MossTTSDelayProcessor)"""
    _ensure_path()
    from moss_tts_delay.configuration_moss_tts import MossTTSDelayConfig
    from moss_tts_delay.modeling_moss_tts import MossTTSDelayModel
    from moss_tts_delay import inference_utils
    from moss_tts_delay.processing_moss_tts import MossTTSDelayProcessor
    return MossTTSDelayConfig, MossTTSDelayModel, inference_utils, MossTTSDelayProcessor


def import_codec():
    """(MossAudioTokenizerConfig, MossAudioTokenizerModel).  transformers >= 5.5 dataclass-wraps PreTrainedConfig
    subclasses, which the reference's bare class annotations do not survive: the wrapper is
    switched off while that one module is imported."""
    _ensure_path()
    import transformers.configuration_utils as cu
    orig = cu.PreTrainedConfig.__dict__.get("__init_subclass__")
    cu.PreTrainedConfig.__init_subclass__ = classmethod(lambda cls, **kw: None)
    try:
        from moss_audio_tokenizer.configuration_moss_audio_tokenizer import MossAudioTokenizerConfig
    finally:
        if orig is not None:
            cu.PreTrainedConfig.__init_subclass__ = orig
        else:
            del cu.PreTrainedConfig.__init_subclass__
    from moss_audio_tokenizer.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    return MossAudioTokenizerConfig, MossAudioTokenizerModel


def qwen3_config(cfg, **over):
    from transformers import Qwen3Config
    kw = dict(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size, num_hidden_layers=cfg.num_layers,
              num_attention_heads=cfg.num_heads, num_key_value_heads=cfg.num_kv_heads, head_dim=cfg.head_dim,
              vocab_size=cfg.vocab_size, max_position_embeddings=8192, rope_theta=cfg.rope_theta,
              rms_norm_eps=cfg.rms_norm_eps, attn_implementation="eager")
    kw.update(over)
    return Qwen3Config(**kw)


def build_lm(cfg, weights, dtype):
    """The reference's MossTTSDelayModel with `weights` (oracle/weights.py names) loaded, eager attention."""
    import torch
    MossTTSDelayConfig, MossTTSDelayModel, _, _ = import_delay()
    mc = MossTTSDelayConfig(language_config=qwen3_config(cfg), n_vq=cfg.n_vq)
    model = MossTTSDelayModel(mc).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected and all("rotary" in m for m in missing), (missing, unexpected)
    inv_freq = model.language_model.rotary_emb.inv_freq.clone()
    model = model.to(dtype)
    # from_pretrained(dtype=bf16) keeps the (non-persistent, explicitly fp32) inv_freq buffer in fp32;
    # a blanket .to(bf16) would round it, which no deployment of the reference does.
    model.language_model.rotary_emb.inv_freq = inv_freq
    model.language_model.rotary_emb.original_inv_freq = inv_freq.clone()
    model.config._attn_implementation = "eager"
    model.language_model.config._attn_implementation = "eager"
    del torch
    return model


def codec_config(cfg, layer_scale: float = 0.01):
    """MossAudioTokenizerConfig with the decoder (and, when cfg has them, encoder) stages of a CodecConfig."""
    MossAudioTokenizerConfig, _ = import_codec()

    def tr(st):
        return dict(causal=True, conv_layout=True, d_model=st.d_model, dim_feedforward=st.dim_feedforward,
                    gating="none", input_dimension=st.input_dim, layer_scale=layer_scale, max_period=cfg.max_period,
                    module_type="Transformer", norm="layer_norm", num_heads=st.num_heads,
                    num_layers=st.num_layers, output_dimension=st.output_dim, positional_embedding="rope")
    dec = []
    for st in cfg.stages:
        dec.append(tr(st))
        dec.append(dict(module_type="PatchedPretransform", patch_size=st.patch_size))
    enc = [dict(module_type="PatchedPretransform", patch_size=cfg.downsample_rate)]      # decode-only stand-in
    q = dict(codebook_dim=cfg.codebook_dim, codebook_size=cfg.codebook_size, input_dim=cfg.code_dim,
             num_quantizers=cfg.num_quantizers, output_dim=cfg.code_dim, quantizer_type="rlfq", rvq_dim=cfg.rvq_dim)
    return MossAudioTokenizerConfig(sampling_rate=cfg.sampling_rate, downsample_rate=cfg.downsample_rate,
                                    causal_transformer_context_duration=cfg.context_duration,
                                    encoder_kwargs=enc, decoder_kwargs=dec, quantizer_type="rlfq", quantizer_kwargs=q)


def build_codec(cfg, weights):
    """The reference's MossAudioTokenizerModel (decode side) with `weights` (oracle/weights.py names) loaded."""
    _, MossAudioTokenizerModel = import_codec()
    model = MossAudioTokenizerModel(codec_config(cfg)).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected, unexpected
    # only encode-side parameters (quantizer input projections, encoder) may be absent
    assert all(("in_proj." in m or "input_proj." in m or "encoder" in m) and "decoder" not in m for m in missing), missing
    return model


def bare_processor(audio_tokenizer, audio_pad_code: int = 1024):
    """A MossTTSDelayProcessor without tokenizer (a hub asset): enough for _parse_audio_codes / decode_audio_codes."""
    _, _, _, MossTTSDelayProcessor = import_delay()

    class _MC:
        pass
    mc = _MC()
    mc.audio_pad_code = audio_pad_code
    proc = object.__new__(MossTTSDelayProcessor)
    proc.model_config = mc
    proc.audio_tokenizer = audio_tokenizer
    return proc


# ------------------------------------------------------------------------------------------------------------
# bench-sized reference models (random init, no checkpoint): bench.py's CPU arm
# ------------------------------------------------------------------------------------------------------------
def build_lm_bench(cfg, dtype=None, alias_layers: bool = True, never_stop: bool = True):
    """The reference's MossTTSDelayModel at `cfg` (e.g. the assumed released 8B shape) with its own random init.
    alias_layers: layers 1..L-1 share layer 0's parameter storage (772 MB at the 8B shape, far beyond any LLC, so a
    layer still streams its weights from DRAM) -- 7 GB instead of 34 GB of fp32 and no 2-minute init.  Stated in the
    bench line's config."""
    import torch
    from transformers.models.qwen3.modeling_qwen3 import Qwen3DecoderLayer
    MossTTSDelayConfig, MossTTSDelayModel, _, _ = import_delay()
    L = cfg.num_layers
    lc = qwen3_config(cfg, num_hidden_layers=1 if alias_layers else L)
    mc = MossTTSDelayConfig(language_config=lc, n_vq=cfg.n_vq)
    model = MossTTSDelayModel(mc).eval()
    if alias_layers and L > 1:
        lm = model.language_model
        lc2 = lm.config
        lc2.num_hidden_layers = L
        if getattr(lc2, "layer_types", None) is not None:
            lc2.layer_types = ["full_attention"] * L
        sd0 = lm.layers[0].state_dict()
        for i in range(1, L):
            with torch.device("meta"):
                li = Qwen3DecoderLayer(lc2, i)
            li.load_state_dict(sd0, assign=True)
            lm.layers.append(li.eval())
        model.config.language_config.num_hidden_layers = L
    if never_stop:      # exact tie -> argmax keeps choosing the gen slot
        with torch.no_grad():
            w = model.lm_heads[0].weight
            w[mc.audio_assistant_delay_slot_token_id] = w[mc.audio_assistant_gen_slot_token_id]
    if dtype is not None and dtype != torch.float32:
        inv_freq = model.language_model.rotary_emb.inv_freq.clone()
        model = model.to(dtype)
        model.language_model.rotary_emb.inv_freq = inv_freq
        model.language_model.rotary_emb.original_inv_freq = inv_freq.clone()
    return model


def build_codec_bench():
    """The reference's MossAudioTokenizerModel built from ITS OWN config.json (the released architecture), own init."""
    MossAudioTokenizerConfig, MossAudioTokenizerModel = import_codec()
    with open(os.path.join(ref_root(), "moss_audio_tokenizer", "config.json")) as f:
        d = json.load(f)
    for k in ("architectures", "auto_map", "transformers_version", "dtype", "torch_dtype", "model_type"):
        d.pop(k, None)
    return MossAudioTokenizerModel(MossAudioTokenizerConfig(**d)).eval()


if __name__ == "__main__":
    ok = build(verbose=True)
    print("oracle/_ref available:", ok)
