"""Oracle for MOSS-Audio-Tokenizer (decode side; encode side = SURVEY.md §8f rank 1) and the processor's
de-delay glue (TEST INFRASTRUCTURE ONLY).

Restates reference moss_audio_tokenizer/modeling_moss_audio_tokenizer.py
  ResidualLFQ.decode_codes :1276-1283, LFQ.decode_code :1065-1074, WNConv1d :967-969
  ProjectedTransformer.forward :916-920, Transformer.forward :860-885,
  TransformerLayer :757-789, MultiheadAttention.forward :627-666, apply_rope :255-300,
  LayerScale :230-234, PatchedPretransform.decode :947-953,
  _decode_frame :1513-1534, decode :1650-1755 (incl. the chunked path's RingKVCache
  :413-480 semantics), batch_decode :1431-1475
  encode side: PatchedPretransform.encode :936-944, _encode_frame :1478-1511, ResidualLFQ.forward :1238-1274,
  LFQ.forward / decode_latents :1055-1092, encode :1536-1648, batch_encode :1403-1429
and reference moss_tts_delay/processing_moss_tts.py
  apply_delay_pattern :516-525, apply_de_delay_pattern :528-537,
  _parse_audio_codes :668-709, decode_audio_codes :880-930.

Chunked decode is restated WITHOUT a chunk loop or ring buffer: writing a chunk of
`len` new keys into a ring of `capacity` slots leaves only keys with position
>= chunk_start + len - capacity alive, so for a query at position p in that chunk
the visible keys are  max(p-context+1, chunk_start+len-capacity, 0) <= k <= p.
`tests/test_oracle_golden.py` checks this closed form against the reference's
actual streaming implementation.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

from moss_tts_b200.config import CodecConfig


# ---- delay pattern (integer work, numpy) ---------------------------------------------

def apply_delay_pattern(codes: np.ndarray, pad_code: int) -> np.ndarray:
    T, n = codes.shape
    out = np.full((T + n - 1, n), pad_code, dtype=codes.dtype)
    for t in range(T):
        for i in range(n):
            out[t + i, i] = codes[t, i]
    return out


def apply_de_delay_pattern(delayed: np.ndarray) -> np.ndarray:
    Td, n = delayed.shape
    T = Td - n + 1
    out = np.zeros((max(T, 0), n), dtype=delayed.dtype)
    for t in range(T):
        for i in range(n):
            out[t, i] = delayed[t + i, i]
    return out


def split_audio_segments(codes: np.ndarray, pad_code: int) -> List[np.ndarray]:
    """Runs of rows that are not all-pad (reference _parse_audio_codes :672-685)."""
    segs, cur = [], []
    for t in range(codes.shape[0]):
        if np.all(codes[t] == pad_code):
            if cur:
                segs.append(np.stack(cur))
                cur = []
        else:
            cur.append(codes[t])
    if cur:
        segs.append(np.stack(cur))
    return segs


def trim_first_segment(wavs: List[torch.Tensor], seg_lens: List[int], start_length: int):
    """Waveform-level trim by start_length ratio (reference _parse_audio_codes :692-707)."""
    if start_length > 0 and wavs and seg_lens and seg_lens[0] > 0:
        ratio = max(0.0, min(float(start_length) / float(seg_lens[0]), 1.0))
        if ratio >= 1.0:
            return wavs[1:]
        if ratio > 0.0:
            n = int(wavs[0].shape[-1] * ratio)
            wavs = [wavs[0][..., n:]] + wavs[1:]
    return wavs


# ---- codec ---------------------------------------------------------------------------

class OracleCodec:
    def __init__(self, cfg: CodecConfig, weights: Dict[str, torch.Tensor]):
        self.cfg = cfg
        self.w = {k: v.float() for k, v in weights.items()}

    @staticmethod
    def _wn(g, v):
        """weight_norm: W = g * v / ||v|| with the norm over all dims but 0."""
        n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
        return (g * v / n).squeeze(-1)

    def decode_codes(self, codes: torch.Tensor) -> torch.Tensor:
        """codes [nq,B,T] -> [B, code_dim, T] fp32."""
        nq, B, T = codes.shape
        emb = torch.zeros(B, self.cfg.rvq_dim, T)
        for q in range(nq):
            p = f"quantizer.quantizers.{q}."
            W = self._wn(self.w[p + "out_proj.parametrizations.weight.original0"],
                         self.w[p + "out_proj.parametrizations.weight.original1"])      # [rvq, cb_dim]
            e = self.w[p + "codebook.weight"][codes[q]]                                  # [B,T,cb_dim]
            emb += (e @ W.T + self.w[p + "out_proj.bias"]).transpose(1, 2)
        p = "quantizer.output_proj."
        Wo = self._wn(self.w[p + "parametrizations.weight.original0"], self.w[p + "parametrizations.weight.original1"])
        return torch.einsum("oc,bct->bot", Wo, emb) + self.w[p + "bias"][None, :, None]

    def _rope(self, q, k):
        """Adjacent-pair rotation, fp32 (reference apply_rope :255-300); q,k [B,H,T,D]."""
        B, H, T, D = q.shape
        j = torch.arange(D // 2, dtype=torch.float32)
        freqs = torch.exp(j * (-math.log(self.cfg.max_period) * 2 / D))
        ang = torch.arange(T, dtype=torch.float32)[:, None] * freqs[None, :]
        cr, sr = torch.cos(ang), torch.sin(ang)

        def rot(x):
            xr, xi = x[..., 0::2], x[..., 1::2]
            out = torch.empty_like(x)
            out[..., 0::2] = xr * cr - xi * sr
            out[..., 1::2] = xr * sr + xi * cr
            return out
        return rot(q), rot(k)

    def _visible(self, T: int, context: int, chunk, up: int = 1) -> torch.Tensor:
        """Boolean [T,T] (query, key) visibility; see module docstring for the chunked form.
        chunk: None, a uniform chunk length in tokens, or a list of per-push lengths in FRAMES (streaming pushes of
        uneven size, `with model.streaming(B): _decode_frame(...)` repeatedly; scaled by `up` tokens per frame)."""
        qp = torch.arange(T)[:, None]
        kp = torch.arange(T)[None, :]
        vis = (kp <= qp) & (qp - kp < context)
        if isinstance(chunk, (list, tuple)):
            end = torch.zeros(T, dtype=torch.long)
            pos = 0
            for n in chunk:
                end[pos: pos + n * up] = pos + n * up
                pos += n * up
            assert pos == T, (pos, T)
            return vis & (kp >= end[:, None] - context)
        if chunk is not None and T > chunk:
            start = (qp // chunk) * chunk
            ln = torch.clamp(T - start, max=chunk)
            vis = vis & (kp >= start + ln - context)      # ring capacity == context (:597-600)
        return vis

    def _stage(self, s: int, x: torch.Tensor, chunk_frames: Optional[int], up: int, enc: bool = False) -> torch.Tensor:
        """x [B, input_dim, T] -> [B, output_dim, T]."""
        st = self.cfg.enc_stages[s] if enc else self.cfg.stages[s]
        base = f"encoder.{2 * s + 1}." if enc else f"decoder.{2 * s}."
        x = x.transpose(1, 2)
        if st.input_dim != st.d_model:
            x = x @ self.w[base + "input_proj.weight"].T
        B, T, d = x.shape
        hd = d // st.num_heads
        if isinstance(chunk_frames, (list, tuple)):
            vis = self._visible(T, st.context, chunk_frames, up)
        else:
            vis = self._visible(T, st.context, None if chunk_frames is None else chunk_frames * up)
        for l in range(st.num_layers):
            p = base + f"transformer.layers.{l}."
            h = F.layer_norm(x, (d,), self.w[p + "norm1.weight"], self.w[p + "norm1.bias"], 1e-5)
            qkv = (h @ self.w[p + "self_attn.in_projs.0.weight"].T).reshape(B, T, 3, st.num_heads, hd)
            q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
            q, k = self._rope(q, k)
            sc = (q @ k.transpose(2, 3)) / math.sqrt(hd)
            sc = sc.masked_fill(~vis[None, None], float("-inf"))
            a = (torch.softmax(sc, dim=-1) @ v).transpose(1, 2).reshape(B, T, d)
            a = a @ self.w[p + "self_attn.out_projs.0.weight"].T
            x = x + self.w[p + "layer_scale_1.scale"] * a
            h = F.layer_norm(x, (d,), self.w[p + "norm2.weight"], self.w[p + "norm2.bias"], 1e-5)
            f = F.gelu(h @ self.w[p + "linear1.weight"].T) @ self.w[p + "linear2.weight"].T
            x = x + self.w[p + "layer_scale_2.scale"] * f
        if st.output_dim != d:
            x = x @ self.w[base + "output_proj.weight"].T
        return x.transpose(1, 2)

    @staticmethod
    def _unpatch(x: torch.Tensor, h: int) -> torch.Tensor:
        """out[b, d, l*h + j] = in[b, d*h + j, l]."""
        B, dh, L = x.shape
        d = dh // h
        out = torch.empty(B, d, L * h)
        for j in range(h):
            out[:, :, j::h] = x[:, j::h, :]
        return out

    @staticmethod
    def _patch(x: torch.Tensor, h: int) -> torch.Tensor:
        """PatchedPretransform.encode :936-944: out[b, d*h + j, l] = in[b, d, l*h + j]."""
        B, d, Lh = x.shape
        out = torch.empty(B, d * h, Lh // h)
        for j in range(h):
            out[:, j::h, :] = x[:, :, j::h]
        return out

    @torch.no_grad()
    def quantize(self, e: torch.Tensor, lengths: torch.Tensor, nq: Optional[int] = None, return_margins: bool = False):
        """ResidualLFQ.forward :1238-1274 with LFQ.forward / decode_latents :1055-1092.
        e [B, code_dim, T] fp32, lengths [B] (frames) -> codes [nq, B, T] int64 (+ per-code margin between the best and
        the second-best negative distance, for near-tie classification in the GPU tests)."""
        c = self.cfg
        nq = nq or c.num_quantizers
        p = "quantizer.input_proj."
        Wi = self._wn(self.w[p + "parametrizations.weight.original0"], self.w[p + "parametrizations.weight.original1"])
        z = torch.einsum("oc,bct->bot", Wi, e.float()) + self.w[p + "bias"][None, :, None]
        B, _, T = z.shape
        mask = (torch.arange(T)[None, :] < lengths[:, None]).unsqueeze(1).float()
        residual = z.clone()
        codes, margins = [], []
        for q in range(nq):
            p = f"quantizer.quantizers.{q}."
            Win = self._wn(self.w[p + "in_proj.parametrizations.weight.original0"],
                           self.w[p + "in_proj.parametrizations.weight.original1"])     # [cb_dim, rvq]
            Wout = self._wn(self.w[p + "out_proj.parametrizations.weight.original0"],
                            self.w[p + "out_proj.parametrizations.weight.original1"])   # [rvq, cb_dim]
            z_e = torch.einsum("oc,bct->bot", Win, residual * mask) + self.w[p + "in_proj.bias"][None, :, None]
            enc = F.normalize(z_e.transpose(1, 2).reshape(-1, c.codebook_dim))
            cb = self.w[p + "codebook.weight"]
            cbn = F.normalize(cb)
            dist = enc.pow(2).sum(1, keepdim=True) - 2 * enc @ cbn.t() + cbn.pow(2).sum(1, keepdim=True).t()
            neg = -dist
            idx = neg.max(1)[1]
            if return_margins:
                top2 = neg.topk(2, dim=1).values
                margins.append((top2[:, 0] - top2[:, 1]).reshape(B, T))
            idx = idx.reshape(B, T)
            z_q = cb[idx].transpose(1, 2)                                                 # [B, cb_dim, T]
            z_q = torch.einsum("oc,bct->bot", Wout, z_q) + self.w[p + "out_proj.bias"][None, :, None]
            residual = residual - z_q * mask
            codes.append(idx)
        codes = torch.stack(codes) if codes else torch.empty(0, B, T, dtype=torch.long)
        return (codes, torch.stack(margins)) if return_margins else codes

    @torch.no_grad()
    def encode_hidden(self, wav: torch.Tensor, chunk_frames: Optional[int] = None) -> torch.Tensor:
        """Encoder stack of _encode_frame :1490-1500.  wav [B, 1, T] with T a multiple of downsample_rate
        -> encoder_hidden_states [B, code_dim, T / downsample_rate]."""
        c = self.cfg
        x = wav.float()
        down = c.downsample_rate
        for s, st in enumerate(c.enc_stages):
            x = self._patch(x, st.patch_size)
            down //= st.patch_size
            x = self._stage(s, x, chunk_frames, down, enc=True)
        return x

    @torch.no_grad()
    def encode_frame(self, wav: torch.Tensor, lengths: Optional[torch.Tensor] = None, nq: Optional[int] = None,
                     chunk_frames: Optional[int] = None, return_margins: bool = False):
        """_encode_frame :1478-1511.  wav [B,1,T] or [B,T]; lengths in samples.
        Returns (codes [nq,B,Tf], code_lengths [B], hidden [B,code_dim,Tf][, margins])."""
        c = self.cfg
        if wav.dim() == 2:
            wav = wav[:, None, :]
        B, _, T = wav.shape
        if lengths is None:
            lengths = torch.full((B,), T, dtype=torch.long)
        if T % c.downsample_rate:
            wav = F.pad(wav, (0, c.downsample_rate - T % c.downsample_rate))
        e = self.encode_hidden(wav, chunk_frames)
        el = lengths.clone()
        for st in c.enc_stages:
            el = el // st.patch_size                    # floor at every stage (:943)
        q = self.quantize(e, el, nq, return_margins)
        return (q[0], el, e, q[1]) if return_margins else (q, el, e)

    @torch.no_grad()
    def decode_frames(self, codes: torch.Tensor, chunk_frames: Optional[int] = None,
                      return_stages: bool = False):
        """codes [nq,B,T] -> audio [B,1,T*downsample_rate].  `chunk_frames` reproduces the
        mask of decode(chunk_duration=...) (only meaningful when T > chunk_frames)."""
        x = self.decode_codes(codes)
        outs = [x]
        up = 1
        for s, st in enumerate(self.cfg.stages):
            x = self._stage(s, x, chunk_frames, up)
            x = self._unpatch(x, st.patch_size)
            up *= st.patch_size
            outs.append(x)
        return (x, outs) if return_stages else x

    @torch.no_grad()
    def decode(self, audio_codes: torch.Tensor, padding_mask=None, chunk_duration=None, num_quantizers=None):
        """Argument handling of reference decode :1650-1755.  Returns (audio [B,1,L], audio_lengths [B])."""
        c = self.cfg
        if audio_codes.dim() == 2:
            audio_codes = audio_codes[:, None, :]
        if num_quantizers is not None:
            if num_quantizers > audio_codes.shape[0]:
                raise ValueError("`num_quantizers` must be <= audio_codes.shape[0]")
            audio_codes = audio_codes[:num_quantizers]
        _, B, T = audio_codes.shape
        lens = padding_mask.sum(-1).long() if padding_mask is not None else torch.full((B,), T, dtype=torch.long)
        if chunk_duration is None:
            return self.decode_frames(audio_codes), lens * c.downsample_rate
        if chunk_duration <= 0 or chunk_duration > c.context_duration:
            raise ValueError("bad chunk_duration")
        if B != 1:
            raise ValueError("Streaming decode via `chunk_duration` currently only supports batch_size=1.")
        chunk_len = int(round(chunk_duration * c.sampling_rate))
        if chunk_len <= 0 or chunk_len % c.downsample_rate != 0:
            raise ValueError("bad chunk_duration")
        cf = chunk_len // c.downsample_rate
        n = int(lens[0])
        audio = self.decode_frames(audio_codes[..., :n], chunk_frames=cf if n > cf else None)
        return audio, torch.tensor([audio.shape[-1]])

    @torch.no_grad()
    def decode_audio_codes(self, segments: List[torch.Tensor], chunk_duration=8) -> List[torch.Tensor]:
        """processor.decode_audio_codes :880-930 for a list of [T,NQ] segments (the reference
        only works for one segment because of the B==1 check; the oracle loops)."""
        out = []
        for seg in segments:
            audio, lens = self.decode(seg.T.contiguous().long()[:, None, :], chunk_duration=chunk_duration)
            out.append(audio[0, 0, : int(lens[0])].float())
        return out
