"""Host-side mirror of `MossAudioTokenizerModel` (reference
moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:1312-1755): `decode`, `batch_decode`,
`_decode_frame`, `quantizer.decode_codes`, and the encode side `encode`, `batch_encode`, `_encode_frame`,
`quantizer(hidden, lengths, n)` (:1403-1429, :1478-1511, :1536-1648, :1238-1274) when the
config carries encoder stages.  Arithmetic runs in libmosstts_sm100.

Differences from the reference, all output-preserving:
  * decode(chunk_duration=...) does not loop over chunks: the streaming ring cache's visibility
    rule is applied as a mask in one pass (see oracle/codec.py), so the waveform equals the
    reference's chunked output.  The reference's B==1 restriction is kept as the same ValueError
    unless `allow_batched_chunks=True` is passed (used by the processor to batch requests).
  * encode(chunk_duration=...) likewise applies the ring rule in one pass and returns what the reference's chunk loop
    returns: floor(len_chunk / downsample_rate) frames per chunk (the ragged tail frame of the last chunk is dropped).
  * the quantiser returns codes (and lengths); the straight-through `quantized_out` tensor of ResidualLFQ.forward
    equals `decode_codes(codes)` on the valid frames and is not materialised.
"""
from __future__ import annotations

import ctypes as C
from contextlib import contextmanager
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from . import _lib as L
from .config import CodecConfig


@dataclass
class MossAudioTokenizerDecoderOutput:
    audio: Optional[torch.Tensor] = None
    audio_lengths: Optional[torch.Tensor] = None


@dataclass
class MossAudioTokenizerEncoderOutput:
    audio_codes: Optional[torch.Tensor] = None
    audio_codes_lengths: Optional[torch.Tensor] = None
    encoder_hidden_states: Optional[torch.Tensor] = None


class _Quantizer:
    """`model.quantizer` facade: decode_codes (reference :1276-1283) and the inference forward (:1238-1274)."""

    @torch.inference_mode()
    def __call__(self, z: torch.Tensor, input_length: torch.Tensor, n_quantizers: Optional[int] = None):
        """z [B, code_dim, T] fp32 -> (None, codes [nq, B, T], input_length)."""
        o = self._o
        if o._enc_engine is None:
            raise L.MossTTSNativeError("this codec was built without encoder / quantiser weights")
        B, D, T = z.shape
        nq = n_quantizers or self.num_quantizers
        hidden = z.to(device=o.device, dtype=torch.float32).transpose(1, 2).contiguous()      # token-major
        lens = input_length.to(device=o.device, dtype=torch.int64).contiguous()
        codes = torch.empty(nq, B, T, dtype=torch.int64, device=o.device)
        if B * T > 0:
            ws = torch.empty(B * T * o.config.rvq_dim, dtype=torch.float32, device=o.device)
            with torch.cuda.device(o.device):
                L.check(L.lib().mt_rlfq_quantize(o._enc_engine, L.ptr(hidden), B, T, L.ptr(lens), nq, L.ptr(codes), L.ptr(ws),
                                                 C.c_longlong(ws.numel() * 4), L.stream_ptr(o.device)), "mt_rlfq_quantize")
        return None, codes, input_length

    def __init__(self, owner: "MossAudioTokenizerModel"):
        self._o = owner
        self.num_quantizers = owner.config.num_quantizers

    @torch.inference_mode()
    def decode_codes(self, codes: torch.Tensor) -> torch.Tensor:
        o = self._o
        if codes.dim() != 3:
            raise ValueError("codes must be (num_quantizers, batch, time)")
        nq, B, T = codes.shape
        codes = codes.to(device=o.device, dtype=torch.int64).contiguous()
        out = torch.empty(B, o.config.code_dim, T, dtype=torch.float32, device=o.device)
        if B * T == 0:
            return out
        with torch.cuda.device(o.device):
            L.check(L.lib().mt_rvq_dequant(o._engine, L.ptr(codes), nq, B, T, 1, L.ptr(out), L.stream_ptr(o.device)),
                    "mt_rvq_dequant")
        return out


class MossAudioTokenizerModel:
    def __init__(self, config: CodecConfig, state_dict: Dict[str, torch.Tensor], device="cuda:0",
                 workspace_bytes: int = 8 << 30, layer_gemm: Optional[str] = None):
        """layer_gemm: "fp16" (default; MT_CODEC_GEMM overrides) runs the four Linear layers of every transformer block
        as fp16-operand / fp32-accumulate tensor-core GEMMs (TF32's mantissa at half the operand bytes);
        "tf32" keeps fp32 storage with TF32 math (fp32 exponent range for weights outside fp16's +-65504)."""
        if not torch.cuda.is_available():
            raise L.MossTTSNativeError("MossAudioTokenizerModel needs a CUDA (sm_100) device; there is no CPU path")
        self.config = config
        self.device = torch.device(device)
        self.sampling_rate = config.sampling_rate
        self.downsample_rate = config.downsample_rate
        self.causal_transformer_context_duration = config.context_duration
        self._ws_cap = workspace_bytes
        self._ws = None
        self._enc_engine = None
        import os
        layer_gemm = layer_gemm or os.environ.get("MT_CODEC_GEMM", "fp16")
        if layer_gemm not in ("fp16", "tf32"):
            raise ValueError(f"layer_gemm must be 'fp16' or 'tf32', got {layer_gemm!r}")
        self.layer_gemm = layer_gemm
        self._build(state_dict)
        if config.enc_stages and "quantizer.input_proj.bias" in state_dict:
            self._build_encoder(state_dict)
        self.quantizer = _Quantizer(self)

    @classmethod
    def random_init(cls, config: CodecConfig, device="cuda:0", seed: int = 0, layer_scale: float = 0.01, **kw):
        """Random decoder weights generated on the device (bench)."""
        dev = torch.device(device)
        gen = torch.Generator(device=dev)
        gen.manual_seed(seed)

        def rn(*shape, s=1.0, mean=0.0):
            return torch.randn(*shape, device=dev, generator=gen) * s + mean
        sd = {}
        c = config
        for q in range(c.num_quantizers):
            p = f"quantizer.quantizers.{q}."
            sd[p + "codebook.weight"] = rn(c.codebook_size, c.codebook_dim)
            sd[p + "out_proj.parametrizations.weight.original0"] = rn(c.rvq_dim, 1, 1, s=0.1, mean=1.0)
            sd[p + "out_proj.parametrizations.weight.original1"] = rn(c.rvq_dim, c.codebook_dim, 1, s=0.3)
            sd[p + "out_proj.bias"] = rn(c.rvq_dim, s=0.05)
        p = "quantizer.output_proj."
        sd[p + "parametrizations.weight.original0"] = rn(c.code_dim, 1, 1, s=0.1, mean=1.0)
        sd[p + "parametrizations.weight.original1"] = rn(c.code_dim, c.rvq_dim, 1, s=0.05)
        sd[p + "bias"] = rn(c.code_dim, s=0.05)
        if kw.pop("encoder", False) and c.enc_stages:
            p = "quantizer.input_proj."
            sd[p + "parametrizations.weight.original0"] = rn(c.rvq_dim, 1, 1, s=0.1, mean=1.0)
            sd[p + "parametrizations.weight.original1"] = rn(c.rvq_dim, c.code_dim, 1, s=0.05)
            sd[p + "bias"] = rn(c.rvq_dim, s=0.05)
            for q in range(c.num_quantizers):
                p = f"quantizer.quantizers.{q}.in_proj."
                sd[p + "parametrizations.weight.original0"] = rn(c.codebook_dim, 1, 1, s=0.1, mean=1.0)
                sd[p + "parametrizations.weight.original1"] = rn(c.codebook_dim, c.rvq_dim, 1, s=0.1)
                sd[p + "bias"] = rn(c.codebook_dim, s=0.05)
            enc = [(f"encoder.{2 * s_ + 1}.", st) for s_, st in enumerate(c.enc_stages)]
        else:
            enc = []
        for base, st in [(f"decoder.{2 * s_}.", st) for s_, st in enumerate(c.stages)] + enc:
            d = st.d_model
            if st.input_dim != d:
                sd[base + "input_proj.weight"] = rn(d, st.input_dim, s=st.input_dim ** -0.5)
            if st.output_dim != d:
                sd[base + "output_proj.weight"] = rn(st.output_dim, d, s=d ** -0.5)
            for l in range(st.num_layers):
                p = base + f"transformer.layers.{l}."
                sd[p + "self_attn.in_projs.0.weight"] = rn(3 * d, d, s=d ** -0.5)
                sd[p + "self_attn.out_projs.0.weight"] = rn(d, d, s=d ** -0.5)
                sd[p + "norm1.weight"] = rn(d, s=0.05, mean=1.0)
                sd[p + "norm1.bias"] = rn(d, s=0.05)
                sd[p + "norm2.weight"] = rn(d, s=0.05, mean=1.0)
                sd[p + "norm2.bias"] = rn(d, s=0.05)
                sd[p + "linear1.weight"] = rn(st.dim_feedforward, d, s=d ** -0.5)
                sd[p + "linear2.weight"] = rn(d, st.dim_feedforward, s=st.dim_feedforward ** -0.5)
                sd[p + "layer_scale_1.scale"] = torch.full((d,), layer_scale, device=dev)
                sd[p + "layer_scale_2.scale"] = torch.full((d,), layer_scale, device=dev)
        return cls(config, sd, device=device, **kw)

    # ---- weights ------------------------------------------------------------------------------
    def _build(self, sd: Dict[str, torch.Tensor]):
        c, dev = self.config, self.device

        def g(name):
            return sd[name].to(device=dev, dtype=torch.float32).contiguous()

        nq = c.num_quantizers
        cb = torch.stack([g(f"quantizer.quantizers.{q}.codebook.weight") for q in range(nq)]).contiguous()
        qg = torch.stack([g(f"quantizer.quantizers.{q}.out_proj.parametrizations.weight.original0").reshape(-1)
                          for q in range(nq)]).contiguous()
        qv = torch.stack([g(f"quantizer.quantizers.{q}.out_proj.parametrizations.weight.original1").reshape(c.rvq_dim, c.codebook_dim)
                          for q in range(nq)]).contiguous()
        qb = torch.stack([g(f"quantizer.quantizers.{q}.out_proj.bias") for q in range(nq)]).contiguous()
        og = g("quantizer.output_proj.parametrizations.weight.original0").reshape(-1).contiguous()
        ov = g("quantizer.output_proj.parametrizations.weight.original1").reshape(c.code_dim, c.rvq_dim).contiguous()
        self._rvq_bias = g("quantizer.output_proj.bias")
        self._rvq_table = torch.empty(nq, c.codebook_size, c.code_dim, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            L.check(L.lib().mt_rvq_fold_tables(L.ptr(cb), L.ptr(qg), L.ptr(qv), L.ptr(qb), L.ptr(og), L.ptr(ov), nq,
                                               c.codebook_size, c.codebook_dim, c.rvq_dim, c.code_dim,
                                               L.ptr(self._rvq_table), L.stream_ptr(dev)), "mt_rvq_fold_tables")
            torch.cuda.current_stream(dev).synchronize()    # cb/qv/... may be freed after this scope
        self._keep = []
        cc = L.CodecConfigC()
        cc.num_stages = len(c.stages)
        cc.layer_gemm_f16 = 1 if self.layer_gemm == "fp16" else 0
        wc = L.CodecWeightsC()
        wc.rvq_table = self._rvq_table.data_ptr()
        wc.rvq_out_bias = self._rvq_bias.data_ptr()
        self._layer_arrays = []
        self._fill_stages(sd, c.stages, "decoder.{}.", 0, cc, wc)
        cc.code_dim, cc.rvq_dim, cc.num_quantizers = c.code_dim, c.rvq_dim, c.num_quantizers
        cc.codebook_size, cc.codebook_dim, cc.max_period = c.codebook_size, c.codebook_dim, c.max_period
        h = C.c_void_p()
        L.check(L.lib().mt_codec_create(C.byref(cc), C.byref(wc), C.byref(h)), "mt_codec_create")
        self._engine = h

    def _fill_stages(self, sd, stages, base_fmt: str, index_offset: int, cc, wc):
        """Stage shapes -> cc, device pointers -> wc (tensors kept alive in self._keep)."""
        dev = self.device

        def g(name):
            return sd[name].to(device=dev, dtype=torch.float32).contiguous()
        for s, st in enumerate(stages):
            cc.input_dim[s], cc.d_model[s], cc.num_heads[s] = st.input_dim, st.d_model, st.num_heads
            cc.num_layers[s], cc.dim_ff[s], cc.output_dim[s] = st.num_layers, st.dim_feedforward, st.output_dim
            cc.patch[s], cc.context[s] = st.patch_size, st.context
            base = base_fmt.format(2 * s + index_offset)
            ip = g(base + "input_proj.weight") if st.input_dim != st.d_model else None
            op = g(base + "output_proj.weight") if st.output_dim != st.d_model else None
            self._keep += [ip, op]
            wc.input_proj[s] = None if ip is None else ip.data_ptr()
            wc.output_proj[s] = None if op is None else op.data_ptr()
            arr = (L.CodecLayerC * st.num_layers)()
            for l in range(st.num_layers):
                p = base + f"transformer.layers.{l}."
                names = ["norm1.weight", "norm1.bias", "norm2.weight", "norm2.bias", "self_attn.in_projs.0.weight",
                         "self_attn.out_projs.0.weight", "linear1.weight", "linear2.weight", "layer_scale_1.scale",
                         "layer_scale_2.scale"]
                ts = [g(p + n) for n in names]
                if self.layer_gemm == "fp16":                 # in_proj, out_proj, linear1, linear2 -> half, once
                    for k in range(4, 8):
                        if ts[k].abs().max().item() > 6.0e4:
                            raise ValueError(f"{p + names[k]} exceeds the fp16 range; build the codec with layer_gemm='tf32'")
                        ts[k] = ts[k].half().contiguous()
                self._keep += ts
                for field, t in zip(("norm1_w", "norm1_b", "norm2_w", "norm2_b", "in_proj", "out_proj", "linear1",
                                     "linear2", "scale1", "scale2"), ts):
                    setattr(arr[l], field, t.data_ptr())
            self._layer_arrays.append(arr)
            wc.layers[s] = C.cast(arr, C.POINTER(L.CodecLayerC))

    def _build_encoder(self, sd: Dict[str, torch.Tensor]):
        """Encoder handle: encoder_kwargs stages + the quantiser's encode-side parameters (weight-norm applied once
        here, W = g * v / ||v||, reference WNConv1d :967-969; codebook rows normalised as decode_latents :1082 does on
        every call)."""
        c, dev = self.config, self.device

        def g(name):
            return sd[name].to(device=dev, dtype=torch.float32).contiguous()

        def wn(prefix):
            gg = g(prefix + "parametrizations.weight.original0")
            v = g(prefix + "parametrizations.weight.original1")
            n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, 1, 1)
            return (gg * v / n).squeeze(-1).contiguous()
        nq = c.num_quantizers
        cb = torch.stack([g(f"quantizer.quantizers.{q}.codebook.weight") for q in range(nq)]).contiguous()
        cbn = torch.nn.functional.normalize(cb, dim=-1).contiguous()
        qt = {
            "in_w": wn("quantizer.input_proj.") if c.code_dim != c.rvq_dim else None,
            "in_b": g("quantizer.input_proj.bias") if c.code_dim != c.rvq_dim else None,
            "q_in_w": torch.stack([wn(f"quantizer.quantizers.{q}.in_proj.") for q in range(nq)]).contiguous(),
            "q_in_b": torch.stack([g(f"quantizer.quantizers.{q}.in_proj.bias") for q in range(nq)]).contiguous(),
            "codebook": cb, "codebook_n": cbn, "codebook_nsq": cbn.pow(2).sum(-1).contiguous(),
            "q_out_w": torch.stack([wn(f"quantizer.quantizers.{q}.out_proj.") for q in range(nq)]).contiguous(),
            "q_out_b": torch.stack([g(f"quantizer.quantizers.{q}.out_proj.bias") for q in range(nq)]).contiguous(),
        }
        self._quant_tensors = qt
        qc = L.CodecQuantWeightsC()
        for k, t in qt.items():
            setattr(qc, k, None if t is None else t.data_ptr())
        cc = L.CodecConfigC()
        cc.num_stages = len(c.enc_stages)
        cc.encoder = 1
        cc.layer_gemm_f16 = 1 if self.layer_gemm == "fp16" else 0
        wc = L.CodecWeightsC()
        wc.quant = C.pointer(qc)
        self._fill_stages(sd, c.enc_stages, "encoder.{}.", 1, cc, wc)
        cc.code_dim, cc.rvq_dim, cc.num_quantizers = c.code_dim, c.rvq_dim, c.num_quantizers
        cc.codebook_size, cc.codebook_dim, cc.max_period = c.codebook_size, c.codebook_dim, c.max_period
        h = C.c_void_p()
        L.check(L.lib().mt_codec_create(C.byref(cc), C.byref(wc), C.byref(h)), "mt_codec_create (encoder)")
        self._enc_engine = h

    def __del__(self):
        try:
            if getattr(self, "_enc_engine", None) is not None:
                L.lib().mt_codec_destroy(self._enc_engine)
                self._enc_engine = None
            if getattr(self, "_engine", None) is not None:
                L.lib().mt_codec_destroy(self._engine)
                self._engine = None
        except Exception:
            pass

    def _workspace(self, B: int, T: int, engine=None) -> torch.Tensor:
        engine = self._engine if engine is None else engine
        need = C.c_longlong(0)
        L.check(L.lib().mt_codec_workspace_bytes(engine, B, T, C.byref(need)), "mt_codec_workspace_bytes")
        one = C.c_longlong(0)
        L.check(L.lib().mt_codec_workspace_bytes(engine, 1, T, C.byref(one)), "mt_codec_workspace_bytes")
        want = max(min(need.value, self._ws_cap), one.value)
        if self._ws is None or self._ws.numel() < want:
            self._ws = None
            self._ws = torch.empty(want, dtype=torch.uint8, device=self.device)
        return self._ws

    # ---- streaming (reference :1393-1400, StreamingModule :121-177, RingKVCache :413-480) ------------------
    @contextmanager
    def streaming(self, batch_size: int = 1):
        """`with model.streaming(batch_size): model._decode_frame(chunk)` / `model._encode_frame(chunk)` per push: every
        decoder and encoder layer keeps a ring cache of `context` tokens on the device between pushes (mt_codec_stream_*;
        the reference's _start_streaming reaches every StreamingModule of both halves, :1375-1391)."""
        if getattr(self, "_streaming", False):
            raise RuntimeError("MossAudioTokenizerModel is already streaming!")
        self._streaming = True
        opened = []
        try:
            for engine in (self._engine, self._enc_engine):
                if engine is None:
                    opened.append((None, None))
                    continue
                need = C.c_longlong(0)
                L.check(L.lib().mt_codec_stream_bytes(engine, int(batch_size), C.byref(need)), "mt_codec_stream_bytes")
                cache = torch.empty(need.value, dtype=torch.uint8, device=self.device)
                h = C.c_void_p()
                L.check(L.lib().mt_codec_stream_begin(engine, int(batch_size), L.ptr(cache), C.c_longlong(cache.numel()),
                                                      C.byref(h)), "mt_codec_stream_begin")
                opened.append((h, cache))
            (self._stream, self._stream_cache), (self._enc_stream, self._enc_stream_cache) = opened
            self._stream_batch = int(batch_size)
            yield
        finally:
            torch.cuda.current_stream(self.device).synchronize()
            for h, _ in opened:
                if h is not None:
                    L.lib().mt_codec_stream_end(h)
            self._stream = self._stream_cache = self._enc_stream = self._enc_stream_cache = None
            self._streaming = False

    def _stream_push(self, codes: torch.Tensor, codes_lengths: torch.Tensor) -> MossAudioTokenizerDecoderOutput:
        nq, B, T = codes.shape
        if B != self._stream_batch:
            raise ValueError(f"streaming batch size is {self._stream_batch}, got {B}")
        wav = torch.empty(B, 1, T * self.downsample_rate, dtype=torch.float32, device=self.device)
        if T > 0:
            ws = self._workspace(B, T)
            need = C.c_longlong(0)
            L.check(L.lib().mt_codec_workspace_bytes(self._engine, B, T, C.byref(need)), "mt_codec_workspace_bytes")
            if ws.numel() < need.value:
                raise L.MossTTSNativeError("codec workspace too small for a streaming push of the whole batch")
            with torch.cuda.device(self.device):
                L.check(L.lib().mt_codec_stream_push(self._stream, L.ptr(codes), nq, T, L.ptr(wav), L.ptr(ws),
                                                     C.c_longlong(ws.numel()), L.stream_ptr(self.device)), "mt_codec_stream_push")
        return MossAudioTokenizerDecoderOutput(audio=wav, audio_lengths=codes_lengths.to(self.device) * self.downsample_rate)

    # ---- reference surface ----------------------------------------------------------------------
    @torch.inference_mode()
    def _decode_frame(self, codes: torch.Tensor, codes_lengths: Optional[torch.Tensor] = None,
                      chunk_frames: int = 0) -> MossAudioTokenizerDecoderOutput:
        nq, B, T = codes.shape
        codes = codes.to(device=self.device, dtype=torch.int64).contiguous()
        if codes_lengths is None:
            codes_lengths = torch.full((B,), T, device=self.device, dtype=torch.long)
        if getattr(self, "_stream", None) is not None:
            if chunk_frames:
                # the reference would enter decoder_module.streaming() a second time here (:1732)
                raise RuntimeError("decoder is already streaming!")
            return self._stream_push(codes, codes_lengths)
        wav = torch.empty(B, 1, T * self.downsample_rate, dtype=torch.float32, device=self.device)
        if B * T > 0:
            ws = self._workspace(B, T)
            with torch.cuda.device(self.device):
                L.check(L.lib().mt_codec_decode(self._engine, L.ptr(codes), nq, B, T, int(chunk_frames), L.ptr(wav),
                                                L.ptr(ws), C.c_longlong(ws.numel()), L.stream_ptr(self.device)),
                        "mt_codec_decode")
        return MossAudioTokenizerDecoderOutput(audio=wav, audio_lengths=codes_lengths.to(self.device) * self.downsample_rate)

    # ---- encode side ---------------------------------------------------------
    @torch.inference_mode()
    def _encode_frame(self, input_values: torch.Tensor, input_lengths: Optional[torch.Tensor] = None,
                      n_quantizers: Optional[int] = None, chunk_frames: int = 0) -> MossAudioTokenizerEncoderOutput:
        """reference :1478-1511.  input_values [B,1,T] or [B,T] fp32; lengths in samples."""
        if self._enc_engine is None:
            raise L.MossTTSNativeError("this codec was built without encoder / quantiser weights")
        if input_values.dim() == 2:
            input_values = input_values.unsqueeze(1)
        B, _, T = input_values.shape
        D = self.downsample_rate
        if input_lengths is None:
            input_lengths = torch.full((B,), T, device=self.device, dtype=torch.long)
        Tf = (T + D - 1) // D
        nq = n_quantizers or self.config.num_quantizers
        wav = torch.zeros(B, Tf * D, dtype=torch.float32, device=self.device)          # pad to whole frames (:1494-1496)
        wav[:, :T] = input_values[:, 0].to(device=self.device, dtype=torch.float32)
        code_lens = input_lengths.to(device=self.device, dtype=torch.int64)
        for st in self.config.enc_stages:                                              # floor at every stage (:943)
            code_lens = code_lens // st.patch_size
        code_lens = code_lens.contiguous()
        codes = torch.empty(nq, B, Tf, dtype=torch.int64, device=self.device)
        hidden = torch.empty(B, Tf, self.config.code_dim, dtype=torch.float32, device=self.device)
        if getattr(self, "_enc_stream", None) is not None:
            if chunk_frames:
                raise RuntimeError("encoder is already streaming!")    # the reference would enter encoder_module.streaming() again (:1612)
            if B != self._stream_batch:
                raise ValueError(f"streaming batch size is {self._stream_batch}, got {B}")
            if Tf > 0:
                ws = self._workspace(B, Tf, self._enc_engine)
                need = C.c_longlong(0)
                L.check(L.lib().mt_codec_workspace_bytes(self._enc_engine, B, Tf, C.byref(need)), "mt_codec_workspace_bytes")
                if ws.numel() < need.value:
                    raise L.MossTTSNativeError("codec workspace too small for a streaming push of the whole batch")
                with torch.cuda.device(self.device):
                    L.check(L.lib().mt_codec_stream_push_encode(self._enc_stream, L.ptr(wav), Tf, L.ptr(code_lens), nq,
                                                                L.ptr(codes), L.ptr(hidden), L.ptr(ws),
                                                                C.c_longlong(ws.numel()), L.stream_ptr(self.device)),
                            "mt_codec_stream_push_encode")
        elif B * Tf > 0:
            ws = self._workspace(B, Tf, self._enc_engine)
            with torch.cuda.device(self.device):
                L.check(L.lib().mt_codec_encode(self._enc_engine, L.ptr(wav), B, Tf, L.ptr(code_lens), nq, int(chunk_frames),
                                                L.ptr(codes), L.ptr(hidden), L.ptr(ws), C.c_longlong(ws.numel()),
                                                L.stream_ptr(self.device)), "mt_codec_encode")
        return MossAudioTokenizerEncoderOutput(audio_codes=codes, audio_codes_lengths=code_lens,
                                               encoder_hidden_states=hidden.transpose(1, 2))

    @torch.inference_mode()
    def encode(self, input_values: torch.Tensor, padding_mask: Optional[torch.Tensor] = None,
               num_quantizers: Optional[int] = None, return_dict: Optional[bool] = None,
               chunk_duration: Optional[float] = None):
        """reference :1536-1648 (same argument checks and messages)."""
        return_dict = True if return_dict is None else return_dict
        if input_values.dim() == 2:
            input_values = input_values.unsqueeze(1)
        B, _, T = input_values.shape
        if padding_mask is not None:
            input_lengths = padding_mask.sum(dim=-1).long().to(self.device)
        else:
            input_lengths = torch.full((B,), T, device=self.device, dtype=torch.long)
        if chunk_duration is None:
            out = self._encode_frame(input_values, input_lengths, num_quantizers)
        else:
            if chunk_duration <= 0:
                raise ValueError("`chunk_duration` must be > 0 when provided.")
            if chunk_duration > self.causal_transformer_context_duration:
                raise ValueError("`chunk_duration` must be <= `config.causal_transformer_context_duration` "
                                 f"({self.causal_transformer_context_duration}), got {chunk_duration}.")
            if B != 1:
                raise ValueError("Streaming encode via `chunk_duration` currently only supports batch_size=1.")
            chunk_length = int(round(chunk_duration * self.sampling_rate))
            if chunk_length <= 0:
                raise ValueError("`chunk_duration` is too small and results in chunk_length <= 0.")
            if chunk_length % self.downsample_rate != 0:
                raise ValueError("`chunk_duration * config.sampling_rate` must be divisible by `config.downsample_rate`. "
                                 f"Got chunk_length={chunk_length}, downsample_rate={self.downsample_rate}.")
            n = int(input_lengths[0].item())
            if n <= chunk_length:
                out = self._encode_frame(input_values[..., :n], input_lengths, num_quantizers)
            else:
                cf = chunk_length // self.downsample_rate
                full = self._encode_frame(input_values[..., :n], input_lengths, num_quantizers, chunk_frames=cf)
                # the reference's chunk loop keeps floor(len_i / downsample_rate) frames of every chunk (:1624-1626)
                keep = (n // chunk_length) * cf + (n % chunk_length) // self.downsample_rate
                out = MossAudioTokenizerEncoderOutput(
                    audio_codes=full.audio_codes[:, :, :keep],
                    audio_codes_lengths=torch.tensor([keep], device=self.device, dtype=torch.long),
                    encoder_hidden_states=full.encoder_hidden_states[:, :, :keep])
        if not return_dict:
            return out.audio_codes, out.audio_codes_lengths
        return out

    @torch.inference_mode()
    def batch_encode(self, wav_list: List[torch.Tensor], num_quantizers: Optional[int] = None):
        """reference :1403-1429."""
        if len(wav_list) == 0:
            raise ValueError("`wav_list` must contain at least one waveform.")
        B = len(wav_list)
        max_length = max(int(w.shape[-1]) for w in wav_list)
        x = torch.zeros(B, 1, max_length, device=self.device)
        lens = torch.zeros(B, device=self.device, dtype=torch.long)
        for i, w in enumerate(wav_list):
            x[i, 0, : w.shape[-1]] = w.to(self.device)
            lens[i] = w.shape[-1]
        return self._encode_frame(x, lens, n_quantizers=num_quantizers)

    @torch.inference_mode()
    def decode(self, audio_codes: torch.Tensor, padding_mask: Optional[torch.Tensor] = None,
               return_dict: Optional[bool] = None, chunk_duration: Optional[float] = None,
               num_quantizers: Optional[int] = None, allow_batched_chunks: bool = False):
        return_dict = True if return_dict is None else return_dict
        if audio_codes.dim() == 2:
            audio_codes = audio_codes.unsqueeze(1)
        if num_quantizers is not None:
            if num_quantizers > audio_codes.shape[0]:
                raise ValueError(f"`num_quantizers` ({num_quantizers}) must be <= audio_codes.shape[0] ({audio_codes.shape[0]}).")
            audio_codes = audio_codes[:num_quantizers]
        _, B, T = audio_codes.shape
        if padding_mask is not None:
            codes_lengths = padding_mask.sum(dim=-1).long().to(self.device)
        else:
            codes_lengths = torch.full((B,), T, device=self.device, dtype=torch.long)
        if chunk_duration is None:
            out = self._decode_frame(audio_codes, codes_lengths)
        else:
            if chunk_duration <= 0:
                raise ValueError("`chunk_duration` must be > 0 when provided.")
            if chunk_duration > self.causal_transformer_context_duration:
                raise ValueError("`chunk_duration` must be <= `config.causal_transformer_context_duration` "
                                 f"({self.causal_transformer_context_duration}), got {chunk_duration}.")
            if B != 1 and not allow_batched_chunks:
                raise ValueError("Streaming decode via `chunk_duration` currently only supports batch_size=1.")
            chunk_length = int(round(chunk_duration * self.sampling_rate))
            if chunk_length <= 0:
                raise ValueError("`chunk_duration` is too small and results in chunk_length <= 0.")
            if chunk_length % self.downsample_rate != 0:
                raise ValueError("`chunk_duration * config.sampling_rate` must be divisible by `config.downsample_rate`. "
                                 f"Got chunk_length={chunk_length}, downsample_rate={self.downsample_rate}.")
            cf = chunk_length // self.downsample_rate
            if B == 1:
                n = int(codes_lengths[0].item())
                out = self._decode_frame(audio_codes[..., :n], codes_lengths, chunk_frames=cf if n > cf else 0)
                out.audio_lengths = torch.tensor([out.audio.shape[-1]], device=self.device, dtype=torch.long)
            else:
                # batched extension: every row must have the same length for the ring rule to be exact
                lens = codes_lengths.tolist()
                if any(n != lens[0] for n in lens):
                    raise ValueError("batched chunked decode needs equal code lengths")
                n = lens[0]
                out = self._decode_frame(audio_codes[..., :n], codes_lengths, chunk_frames=cf if n > cf else 0)
        if not return_dict:
            return (out.audio,)
        return out

    @torch.inference_mode()
    def batch_decode(self, codes_list: List[torch.Tensor], num_quantizers: Optional[int] = None):
        if len(codes_list) == 0:
            raise ValueError("`codes_list` must contain at least one code tensor.")
        nqs = [int(c.shape[0]) for c in codes_list]
        if num_quantizers is None:
            num_quantizers = nqs[0]
            if any(nq != num_quantizers for nq in nqs):
                raise ValueError("All elements in `codes_list` must have the same number of quantizers when "
                                 "`num_quantizers` is None. Pass `num_quantizers=...` to decode a common prefix.")
        elif min(nqs) < num_quantizers:
            raise ValueError("`num_quantizers` must be <= the number of quantizers for every element in `codes_list`. "
                             f"Got num_quantizers={num_quantizers}, min(codes.shape[0])={min(nqs)}.")
        max_len = max(int(c.shape[-1]) for c in codes_list)
        B = len(codes_list)
        codes = torch.zeros(num_quantizers, B, max_len, device=self.device, dtype=torch.long)
        lens = torch.zeros(B, device=self.device, dtype=torch.long)
        for i, c in enumerate(codes_list):
            c = c[:num_quantizers].to(self.device)
            codes[:, i, : c.shape[-1]] = c
            lens[i] = c.shape[-1]
        return self._decode_frame(codes, lens)
