"""Decode half of `MossTTSDelayProcessor` (reference moss_tts_delay/processing_moss_tts.py):
  apply_delay_pattern :516-525, apply_de_delay_pattern :528-537, _parse_audio_codes :668-709,
  decode :711-733, decode_audio_codes :880-930.
The prompt-building half (chat template, tokenizer, audio encode) is out of scope (SURVEY.md §8 f).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Tuple, Union

import torch

from . import _lib as L
from .config import LMConfig

AUDIO_PLACEHOLDER = "<|audio|>"


@dataclass
class AssistantMessage:
    content: str = AUDIO_PLACEHOLDER
    audio_codes_list: List[Union[str, torch.Tensor]] = field(default_factory=list)


class MossTTSDelayProcessor:
    def __init__(self, model_config: LMConfig, audio_tokenizer=None, tokenizer=None):
        self.model_config = model_config
        self.audio_tokenizer = audio_tokenizer
        self.tokenizer = tokenizer

    # ---- delay pattern (integer kernels) ---------------------------------------------------------
    @staticmethod
    def apply_delay_pattern(codes: torch.Tensor, pad_code: int) -> torch.Tensor:
        if not codes.is_cuda:
            raise L.MossTTSNativeError("apply_delay_pattern: codes must live on a CUDA device (no CPU path)")
        T, n = codes.shape
        src = codes.to(torch.int64).contiguous()
        out = torch.empty(T + n - 1, n, dtype=torch.int64, device=codes.device)
        if out.numel():
            with torch.cuda.device(codes.device):
                L.check(L.lib().mt_apply_delay_pattern(L.ptr(src), T, n, C.c_longlong(int(pad_code)), L.ptr(out),
                                                       L.stream_ptr(codes.device)), "mt_apply_delay_pattern")
        return out.to(codes.dtype)

    @staticmethod
    def apply_de_delay_pattern(delay_codes: torch.Tensor) -> torch.Tensor:
        if not delay_codes.is_cuda:
            raise L.MossTTSNativeError("apply_de_delay_pattern: codes must live on a CUDA device (no CPU path)")
        Td, n = delay_codes.shape
        src = delay_codes.to(torch.int64).contiguous()
        out = torch.empty(max(Td - n + 1, 0), n, dtype=torch.int64, device=delay_codes.device)
        if out.numel():
            with torch.cuda.device(delay_codes.device):
                L.check(L.lib().mt_apply_de_delay_pattern(L.ptr(src), Td, n, L.ptr(out), L.stream_ptr(delay_codes.device)),
                        "mt_apply_de_delay_pattern")
        return out.to(delay_codes.dtype)

    # ---- segments ----------------------------------------------------------------------------------
    def _split_segments(self, audio_codes: torch.Tensor) -> List[torch.Tensor]:
        """Runs of rows that are not all-pad (:672-685); audio_codes is the de-delayed [T', n_vq] matrix."""
        is_pad = (audio_codes == self.model_config.audio_pad_code).all(dim=1)
        non_pad = ~is_pad
        if not bool(non_pad.any()):
            return []
        idx = torch.nonzero(non_pad).squeeze(1)
        breaks = (torch.where(idx[1:] != idx[:-1] + 1)[0] + 1).tolist()
        pieces = torch.tensor_split(idx, breaks) if breaks else [idx]
        return [audio_codes[s] for s in pieces]

    def _split_segments_many(self, mats: List[torch.Tensor]) -> List[List[torch.Tensor]]:
        """_split_segments for a list of de-delayed matrices with ONE device->host transfer (the all-pad row masks of
        every utterance); the runs are then plain row slices."""
        if not mats:
            return []
        pad = self.model_config.audio_pad_code
        masks = [(m == pad).all(dim=1) for m in mats]
        flat = torch.cat(masks).cpu().numpy()
        out, off = [], 0
        for m in mats:
            n = int(m.shape[0])
            is_pad = flat[off: off + n]
            off += n
            segs, start = [], None
            for r in range(n):
                if not is_pad[r] and start is None:
                    start = r
                elif is_pad[r] and start is not None:
                    segs.append(m[start:r]); start = None
            if start is not None:
                segs.append(m[start:n])
            out.append(segs)
        return out

    @staticmethod
    def _trim_first(wavs: List[torch.Tensor], first_len: int, start_length: int) -> List[torch.Tensor]:
        """:692-707"""
        if start_length > 0 and wavs and first_len > 0:
            ratio = max(0.0, min(float(start_length) / float(first_len), 1.0))
            if ratio >= 1.0:
                return wavs[1:]
            if ratio > 0.0:
                n = int(wavs[0].shape[-1] * ratio)
                wavs = [wavs[0][..., n:]] + wavs[1:]
        return wavs

    def decode_audio_codes(self, audio_tokens_list: Union[torch.Tensor, List[torch.Tensor]]) -> List[torch.Tensor]:
        """List of [T, NQ] code matrices -> list of 1-D fp32 CPU waveforms (:880-930).  Segments are
        decoded one codec call per distinct length (the reference handles exactly one segment)."""
        if self.audio_tokenizer is None:
            raise RuntimeError("audio_tokenizer is not set on processor.")
        if isinstance(audio_tokens_list, torch.Tensor):
            audio_tokens_list = [audio_tokens_list]
        if len(audio_tokens_list) == 0:
            return []
        wavs = self._decode_segments_device(audio_tokens_list)
        return [w.float().cpu() for w in wavs]

    def _decode_segments_device(self, segs: List[torch.Tensor]) -> List[torch.Tensor]:
        tok = self.audio_tokenizer
        out: List[Optional[torch.Tensor]] = [None] * len(segs)
        by_len = {}
        for i, s in enumerate(segs):
            by_len.setdefault(int(s.shape[0]), []).append(i)
        for T, idxs in by_len.items():
            codes = torch.stack([segs[i].to(tok.device).long().T for i in idxs], dim=1).contiguous()   # [NQ, B, T]
            dec = tok.decode(codes, return_dict=True, chunk_duration=8, allow_batched_chunks=True)
            for j, i in enumerate(idxs):
                out[i] = dec.audio[j, 0, : T * tok.downsample_rate]
        return out    # type: ignore[return-value]

    def _parse_audio_codes(self, start_length: int, audio_codes: torch.Tensor) -> List[torch.Tensor]:
        audio_codes = self.apply_de_delay_pattern(audio_codes)
        segs = self._split_segments(audio_codes)
        if not segs:
            return []
        wavs = self.decode_audio_codes(segs)
        return self._trim_first(wavs, int(segs[0].shape[0]), int(start_length))

    def _parse_text_codes(self, start_length, text_codes) -> str:
        if self.tokenizer is None:
            return AUDIO_PLACEHOLDER          # no tokenizer offline: content cannot be rendered
        text = self.tokenizer.decode(text_codes)
        prefix = self.tokenizer.decode(text_codes[:start_length])
        return text[len(prefix):]

    def decode_device(self, output: List[Tuple[int, torch.Tensor]]) -> List[List[torch.Tensor]]:
        """generate() output -> per-utterance lists of trimmed waveforms that stay on the device.
        All utterances' segments go through the codec in as few batched calls as possible; results
        equal the per-utterance reference path (:668-709)."""
        per_utt = []
        all_segs = []
        dds = [self.apply_de_delay_pattern(gen[:, 1:]) for _, gen in output]
        for (start_length, _), segs in zip(output, self._split_segments_many(dds)):
            per_utt.append((int(start_length), len(all_segs), [int(s.shape[0]) for s in segs]))
            all_segs += segs
        wavs = self._decode_segments_device(all_segs) if all_segs else []
        return [self._trim_first(wavs[off: off + len(lens)], lens[0] if lens else 0, sl) for sl, off, lens in per_utt]

    def decode(self, output: List[Tuple[int, torch.Tensor]]) -> List[Optional[AssistantMessage]]:
        """generate() output -> messages whose audio_codes_list holds 1-D fp32 CPU waveforms (:711-733)."""
        dev = self.decode_device(output)
        messages = []
        for wavs, (start_length, gen) in zip(dev, output):
            content = self._parse_text_codes(start_length, gen[:, 0])
            host = [w.float().cpu() for w in wavs]
            messages.append(None if content == "" else AssistantMessage(content=content, audio_codes_list=host))
        return messages
