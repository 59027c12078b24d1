"""Shape/constant containers for the MOSS-TTS synthesis hot path.

These carry exactly the hyper-parameters the kernels need; they mirror
`MossTTSDelayConfig` (reference moss_tts_delay/configuration_moss_tts.py:62-106),
the Qwen3 backbone config it wraps, and the codec's `config.json`
(reference moss_audio_tokenizer/config.json:9-186).
"""
from __future__ import annotations

from dataclasses import dataclass, field, asdict
from typing import List, Optional


@dataclass
class LMConfig:
    hidden_size: int = 4096
    intermediate_size: int = 12288
    num_layers: int = 36
    num_heads: int = 32
    num_kv_heads: int = 8
    head_dim: int = 128
    vocab_size: int = 151936
    n_vq: int = 32
    audio_vocab_size: int = 1024
    rope_theta: float = 1_000_000.0
    rms_norm_eps: float = 1e-6
    # special ids, reference configuration_moss_tts.py:67-78
    pad_token_id: int = 151643
    im_start_token_id: int = 151644
    im_end_token_id: int = 151645
    audio_user_slot_token_id: int = 151654
    audio_assistant_gen_slot_token_id: int = 151656
    audio_assistant_delay_slot_token_id: int = 151662
    audio_start_token_id: int = 151652
    audio_end_token_id: int = 151653
    audio_pad_code: int = 1024
    sampling_rate: int = 24000

    @property
    def audio_head_size(self) -> int:
        return self.audio_vocab_size + 1

    @property
    def q_size(self) -> int:
        return self.num_heads * self.head_dim

    @property
    def kv_size(self) -> int:
        return self.num_kv_heads * self.head_dim

    def to_dict(self):
        return asdict(self)

    # ---- named shapes ------------------------------------
    @staticmethod
    def delay_8b(n_vq: int = 32) -> "LMConfig":
        """Assumed released MossTTSDelay-8B shape = public Qwen3-8B backbone."""
        return LMConfig(n_vq=n_vq)

    @staticmethod
    def tiny(n_vq: int = 4) -> "LMConfig":
        return LMConfig(hidden_size=128, intermediate_size=256, num_layers=2, num_heads=4,
                        num_kv_heads=2, head_dim=64, n_vq=n_vq)

    @staticmethod
    def mid(n_vq: int = 32) -> "LMConfig":
        return LMConfig(hidden_size=512, intermediate_size=1536, num_layers=3, num_heads=8,
                        num_kv_heads=2, head_dim=128, n_vq=n_vq)


@dataclass
class LocalConfig(LMConfig):
    """MossTTSLocal (reference moss_tts_local/configuration_moss_tts.py:62-106): a Qwen3 "global" backbone
    (the LMConfig fields) plus a small "local"/depth transformer that emits the 1+n_vq tokens of one frame
    sequentially.  The local transformer copies the backbone's head layout (num_heads / num_kv_heads /
    head_dim, modeling_moss_tts.py:585-590) with its own hidden / ffn sizes and no rotary embedding."""
    local_hidden_size: int = 1536
    local_ffn_hidden_size: int = 8960
    local_num_layers: int = 4
    additional_mlp_ffn_hidden_size: int = 2048

    @staticmethod
    def local_1p7b(n_vq: int = 32) -> "LocalConfig":
        """Assumed released MossTTSLocal shape = public Qwen3-1.7B backbone (SURVEY.md §8 d)."""
        return LocalConfig(hidden_size=2048, intermediate_size=6144, num_layers=28, num_heads=16, num_kv_heads=8,
                           head_dim=128, n_vq=n_vq)

    @staticmethod
    def tiny(n_vq: int = 4) -> "LocalConfig":
        return LocalConfig(hidden_size=128, intermediate_size=256, num_layers=2, num_heads=4, num_kv_heads=2,
                           head_dim=64, n_vq=n_vq, local_hidden_size=192, local_ffn_hidden_size=320,
                           local_num_layers=2, additional_mlp_ffn_hidden_size=256)

    @staticmethod
    def mid(n_vq: int = 8) -> "LocalConfig":
        return LocalConfig(hidden_size=512, intermediate_size=1024, num_layers=2, num_heads=8, num_kv_heads=2,
                           head_dim=128, n_vq=n_vq, local_hidden_size=384, local_ffn_hidden_size=768,
                           local_num_layers=2, additional_mlp_ffn_hidden_size=512)


@dataclass
class RealtimeConfig(LMConfig):
    """MOSS-TTS-Realtime (reference moss_tts_realtime/mossttsrealtime/configuration_mossttsrealtime.py:35-112): a Qwen3
    backbone whose input is the sum of one text and `rvq` audio embeddings (audio vocabulary 1027 = 1024 codes + pad 1024,
    bos 1025, eos 1026) and a small Qwen3-shaped "local transformer" WITH rotary embedding that emits the `rvq` codes of
    one frame sequentially from the backbone's last hidden state (streaming_mossttsrealtime.py:301-357).  The LMConfig
    fields describe the backbone (`n_vq` = rvq); the local transformer shares its hidden size."""
    n_vq: int = 16
    audio_vocab_size: int = 1026          # audio_head_size = 1027 (configuration_mossttsrealtime.py:76-77)
    local_num_layers: int = 4
    local_intermediate_size: int = 6144
    local_num_heads: int = 16
    local_num_kv_heads: int = 8
    local_head_dim: int = 128
    local_rope_theta: float = 1_000_000.0
    local_rms_norm_eps: float = 1e-6
    audio_bos_token: int = 1025           # streaming_mossttsrealtime.py:47-51
    audio_eos_token: int = 1026
    text_pad_id: int = 151655
    reference_audio_pad: int = 151654

    @property
    def rvq(self) -> int:
        return self.n_vq

    @property
    def local_q_size(self) -> int:
        return self.local_num_heads * self.local_head_dim

    @property
    def local_kv_size(self) -> int:
        return self.local_num_kv_heads * self.local_head_dim

    @staticmethod
    def realtime_1p7b() -> "RealtimeConfig":
        """Assumed released shape: public Qwen3-1.7B backbone + the local-transformer defaults of the reference config."""
        return RealtimeConfig(hidden_size=2048, intermediate_size=6144, num_layers=28, num_heads=16, num_kv_heads=8,
                              head_dim=128)

    @staticmethod
    def wide(num_layers: int = 2) -> "RealtimeConfig":
        """Every per-layer shape of the assumed 1.7 B model (backbone and the full 4-layer local transformer), with
        `num_layers` of the backbone's 28 layers: the shape of the reference-executed golden rt_wide_bf16."""
        c = RealtimeConfig.realtime_1p7b()
        c.num_layers = num_layers
        return c

    @staticmethod
    def tiny(n_vq: int = 4) -> "RealtimeConfig":
        return RealtimeConfig(hidden_size=128, intermediate_size=256, num_layers=2, num_heads=4, num_kv_heads=2,
                              head_dim=64, n_vq=n_vq, local_num_layers=2, local_intermediate_size=320,
                              local_num_heads=4, local_num_kv_heads=2, local_head_dim=64)

    @staticmethod
    def mid(n_vq: int = 16) -> "RealtimeConfig":
        return RealtimeConfig(hidden_size=512, intermediate_size=1024, num_layers=2, num_heads=8, num_kv_heads=2,
                              head_dim=128, n_vq=n_vq, local_num_layers=4, local_intermediate_size=1536,
                              local_num_heads=4, local_num_kv_heads=2, local_head_dim=128)


@dataclass
class CodecStage:
    """One ProjectedTransformer stage followed by a patch upsample (reference
    modeling_moss_audio_tokenizer.py:888-953)."""
    input_dim: int
    d_model: int
    num_heads: int
    num_layers: int
    dim_feedforward: int
    output_dim: int
    patch_size: int      # upsample factor applied after this stage
    context: int         # attention window in tokens of this stage


@dataclass
class CodecConfig:
    sampling_rate: int = 24000
    downsample_rate: int = 1920
    context_duration: float = 10.0
    code_dim: int = 768          # quantizer input/output dim
    rvq_dim: int = 512
    num_quantizers: int = 32
    codebook_size: int = 1024
    codebook_dim: int = 8
    max_period: float = 10000.0
    stages: List[CodecStage] = field(default_factory=list)
    # encoder side: same CodecStage record, but `patch_size` is the DOWNsample applied BEFORE the
    # stage (PatchedPretransform.encode :936-944) and `input_dim` the width after it.  Empty = decode-only codec.
    enc_stages: List[CodecStage] = field(default_factory=list)

    @property
    def frame_rate(self) -> float:
        return self.sampling_rate / self.downsample_rate

    def to_dict(self):
        return asdict(self)

    @staticmethod
    def from_dict(d: dict) -> "CodecConfig":
        d = dict(d)
        stages = [CodecStage(**s) for s in d.pop("stages")]
        enc = [CodecStage(**s) for s in d.pop("enc_stages", [])]
        return CodecConfig(stages=stages, enc_stages=enc, **d)

    def with_encoder(self, dims, patch_sizes) -> "CodecConfig":
        """Adds encoder stages; dims = (input_dim after the patch, d_model, heads, layers, ff, output_dim) per stage."""
        rate = float(self.sampling_rate)
        self.enc_stages = []
        for (inp, d, h, L, ff, out), p in zip(dims, patch_sizes):
            rate /= p
            self.enc_stages.append(CodecStage(inp, d, h, L, ff, out, p, int(rate * self.context_duration)))
        return self

    @staticmethod
    def _build(dims, sampling_rate, context_duration, code_dim, rvq_dim, num_quantizers, patch_sizes):
        downsample = 1
        for p in patch_sizes:
            downsample *= p
        cfg = CodecConfig(sampling_rate=sampling_rate, downsample_rate=downsample,
                          context_duration=context_duration, code_dim=code_dim, rvq_dim=rvq_dim,
                          num_quantizers=num_quantizers)
        rate = sampling_rate / downsample
        for (inp, d, h, L, ff, out), p in zip(dims, patch_sizes):
            # reference modeling_moss_audio_tokenizer.py:1368: context=int(frame_rate * duration)
            cfg.stages.append(CodecStage(inp, d, h, L, ff, out, p, int(rate * context_duration)))
            rate *= p
        return cfg

    @staticmethod
    def released() -> "CodecConfig":
        """decoder_kwargs of reference moss_audio_tokenizer/config.json:11-92."""
        dims = [(768, 1280, 20, 32, 5120, 1280),
                (640, 768, 12, 12, 3072, 768),
                (384, 768, 12, 12, 3072, 768),
                (384, 768, 12, 12, 3072, 240)]
        enc = [(240, 768, 12, 12, 3072, 384),        # encoder_kwargs of config.json:93-176
               (768, 768, 12, 12, 3072, 384),
               (768, 768, 12, 12, 3072, 640),
               (1280, 1280, 20, 32, 5120, 768)]
        return CodecConfig._build(dims, 24000, 10.0, 768, 512, 32, [2, 2, 2, 240]).with_encoder(enc, [240, 2, 2, 2])

    @staticmethod
    def mini(context_duration: float = 1.0, num_quantizers: int = 8) -> "CodecConfig":
        """Same topology, small widths; windows of 12/25/50/100 tokens so a 2 s clip
        exercises the sliding window in every stage."""
        dims = [(96, 128, 2, 2, 256, 128),
                (64, 128, 2, 1, 256, 128),
                (64, 64, 1, 1, 128, 64),
                (32, 64, 1, 1, 128, 240)]
        enc = [(240, 64, 1, 1, 128, 32),
               (64, 64, 1, 1, 128, 64),
               (128, 128, 2, 1, 256, 64),
               (128, 128, 2, 2, 256, 96)]
        return CodecConfig._build(dims, 24000, context_duration, 96, 64, num_quantizers,
                                  [2, 2, 2, 240]).with_encoder(enc, [240, 2, 2, 2])
