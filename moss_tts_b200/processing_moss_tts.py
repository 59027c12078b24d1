"""`MossTTSDelayProcessor` (reference moss_tts_delay/processing_moss_tts.py).
Decode half: apply_delay_pattern :516-525, apply_de_delay_pattern :528-537, _parse_audio_codes :668-709,
  decode :711-733, decode_audio_codes :880-930.
Prompt half: loudness_normalize :736-748, encode_audios_from_wav :778-852,
  _replace_audio_placeholders :434-468, _merge_consecutive_audio_placeholders :471-513, _get_unified_codes :539-641,
  _pad :410-431.  String handling stays on the host as in the reference; every tensor operation is a kernel of
  libmosstts_sm100 (mt_loudness_normalize, mt_codec_encode, mt_pack_unified_codes, mt_pad_left) and the result stays on the
  device, ready for `generate`.  The chat template / tokenizer themselves are hub assets and are passed in by the caller
  (any object with `encode` and `convert_ids_to_tokens`); `__call__` is not mirrored.
"""
from __future__ import annotations

import ctypes as C
import re
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

    # ---- prompt side ----------------------------------------------------------------------------------
    def _device(self) -> torch.device:
        dev = getattr(self.audio_tokenizer, "device", None)
        return torch.device(dev) if dev is not None else torch.device("cuda", torch.cuda.current_device())

    def _token(self, token_id: int) -> str:
        tok = self.tokenizer.convert_ids_to_tokens(int(token_id))
        if isinstance(tok, list):
            return tok[0] if len(tok) > 0 else ""
        return tok

    @staticmethod
    def loudness_normalize(wav: torch.Tensor, target_dbfs: float = -20, gain_range: Tuple[float, float] = (-3.0, 3.0)):
        """reference :736-748 (a new fp32 tensor; the input is left untouched)."""
        if not wav.is_cuda:
            raise L.MossTTSNativeError("loudness_normalize: wav must live on a CUDA device (no CPU path)")
        out = wav.to(torch.float32).contiguous().clone()
        if out.numel() == 0:
            return out
        n = torch.tensor([out.numel()], dtype=torch.int64, device=out.device)
        with torch.cuda.device(out.device):
            L.check(L.lib().mt_loudness_normalize(L.ptr(out), C.c_longlong(out.numel()), L.ptr(n), 1, C.c_float(target_dbfs),
                                                  C.c_float(gain_range[0]), C.c_float(gain_range[1]),
                                                  L.stream_ptr(out.device)), "mt_loudness_normalize")
        return out

    def encode_audios_from_wav(self, wav_list: Union[torch.Tensor, List[torch.Tensor]], sampling_rate: int,
                               n_vq: Optional[int] = None) -> List[torch.Tensor]:
        """reference :778-852: mono mix-down, (resample), loudness normalisation, codec encode -> list of [T_i, NQ]
        int64 CPU tensors.  The batch is normalised by ONE kernel launch and encoded by ONE mt_codec_encode call."""
        if self.audio_tokenizer is None:
            raise RuntimeError("audio_tokenizer is not set on processor.")
        if isinstance(wav_list, torch.Tensor):
            wav_list = [wav_list]
        dev = self._device()
        mono = []
        for wav in wav_list:
            if wav.dim() == 1:
                wav = wav.unsqueeze(0)
            if wav.shape[0] > 1:
                wav = torch.mean(wav, dim=0, keepdim=True)
            if sampling_rate != self.model_config.sampling_rate:
                try:
                    import torchaudio
                except Exception as e:      # the reference depends on torchaudio for this step as well
                    raise NotImplementedError("resampling needs torchaudio; pass audio at the model's sampling rate") from e
                wav = torchaudio.functional.resample(waveform=wav, orig_freq=sampling_rate,
                                                     new_freq=self.model_config.sampling_rate)
            mono.append(wav.squeeze(0))
        B = len(mono)
        if B == 0:
            return []
        max_len = max(int(w.shape[-1]) for w in mono)
        x = torch.zeros(B, max_len, dtype=torch.float32, device=dev)
        lens_h = torch.tensor([int(w.shape[-1]) for w in mono], dtype=torch.int64)
        for i, w in enumerate(mono):
            x[i, : w.shape[-1]] = w.to(device=dev, dtype=torch.float32)
        lens = lens_h.to(dev)
        with torch.cuda.device(dev):
            L.check(L.lib().mt_loudness_normalize(L.ptr(x), C.c_longlong(max_len), L.ptr(lens), B, C.c_float(-20.0),
                                                  C.c_float(-3.0), C.c_float(3.0), L.stream_ptr(dev)), "mt_loudness_normalize")
        enc = self.audio_tokenizer._encode_frame(x.unsqueeze(1), lens, n_quantizers=n_vq)
        codes = enc.audio_codes.cpu()                       # the reference's contract: (T, NQ) on the CPU
        out_lens = enc.audio_codes_lengths.cpu().tolist()
        return [codes[:, i, : int(out_lens[i])].transpose(0, 1).contiguous().to(torch.long) for i in range(B)]

    @staticmethod
    def _replace_audio_placeholders(content: str, lengths: List[int], n_vq: int, gen_slot_token: str,
                                    delay_slot_token: str, audio_start_token: str, audio_end_token: str) -> str:
        """reference :434-468 (host string work)."""
        if n_vq < 1:
            raise ValueError(f"n_vq must be >= 1, got {n_vq}")
        num_placeholders = content.count(AUDIO_PLACEHOLDER)
        if num_placeholders != len(lengths):
            raise ValueError(f"Number of {AUDIO_PLACEHOLDER} ({num_placeholders}) does not match lengths ({len(lengths)})")
        blocks = []
        for length in lengths:
            if length < 0:
                raise ValueError(f"length must be >= 0, got {length}")
            if length == 0:
                blocks.append(f"{audio_start_token}{audio_end_token}")
            else:
                blocks.append(f"{audio_start_token}{gen_slot_token * length}{delay_slot_token * (n_vq - 1)}{audio_end_token}")
        it = iter(blocks)
        return re.sub(re.escape(AUDIO_PLACEHOLDER), lambda m: next(it), content)

    @staticmethod
    def _merge_consecutive_audio_placeholders(content: str, audio_codes_list: List[torch.Tensor]):
        """reference :471-513: placeholders separated only by whitespace become one audio block."""
        matches = list(re.finditer(re.escape(AUDIO_PLACEHOLDER), content))
        if len(matches) <= 1:
            return content, audio_codes_list
        if len(matches) != len(audio_codes_list):
            raise ValueError("Audio placeholders do not match the provided audio codes list.")
        merged, parts, last, i = [], [], 0, 0
        while i < len(matches):
            j = i
            while j + 1 < len(matches) and content[matches[j].end(): matches[j + 1].start()].strip() == "":
                j += 1
            parts += [content[last: matches[i].start()], AUDIO_PLACEHOLDER]
            last = matches[j].end()
            merged.append(audio_codes_list[i] if j == i else torch.cat(audio_codes_list[i: j + 1], dim=0))
            i = j + 1
        parts.append(content[last:])
        return "".join(parts), merged

    def _get_unified_codes(self, role: str, content: str, audio_codes_list: List[torch.Tensor], truncation: bool):
        """reference :539-641 -> [rows, 1 + n_vq] int64 on the device.  The segment table is derived from the token list
        on the host (the ids come from the tokenizer as a Python list anyway); one kernel writes every channel."""
        mc = self.model_config
        if role == "user":
            gen_tok = delay_tok = self._token(mc.audio_user_slot_token_id)
            truncation = False
        else:
            gen_tok = self._token(mc.audio_assistant_gen_slot_token_id)
            delay_tok = self._token(mc.audio_assistant_delay_slot_token_id)
        n_vq = int(audio_codes_list[0].shape[1]) if len(audio_codes_list) else mc.n_vq
        if len(audio_codes_list) > 1 and AUDIO_PLACEHOLDER in content:
            content, audio_codes_list = self._merge_consecutive_audio_placeholders(content, audio_codes_list)
        content = self._replace_audio_placeholders(content, [len(c) for c in audio_codes_list], n_vq, gen_tok, delay_tok,
                                                   self._token(mc.audio_start_token_id), self._token(mc.audio_end_token_id))
        ids = list(self.tokenizer.encode(content))
        starts = [i for i, t in enumerate(ids) if t == mc.audio_start_token_id]
        ends = [i for i, t in enumerate(ids) if t == mc.audio_end_token_id]
        if len(starts) != len(audio_codes_list) or len(ends) != len(audio_codes_list):
            raise ValueError("Audio placeholders do not match the provided audio codes list.")
        dev = self._device()
        blocks, row, prefix, off = [], 0, 0, 0
        for a, e, codes in zip(starts, ends, audio_codes_list):
            T = int(codes.shape[0])
            blocks.append((row, a - prefix + 1, -1, 0))
            row += a - prefix + 1
            blocks.append((row, T + n_vq - 1, T, off))
            row += T + n_vq - 1
            off += T
            prefix = e
        if audio_codes_list:
            if truncation:
                r0, n, T, o = blocks[-1]
                kept = n - (n_vq - 1) if n_vq > 1 else 0          # `[: -(n_vq - 1)]` is empty for n_vq == 1 (:626-629)
                blocks[-1] = (r0, kept, T, o)
                row = r0 + kept
            else:
                blocks.append((row, len(ids) - ends[-1], -1, 0))
                row += len(ids) - ends[-1]
        else:
            row = len(ids)
        if len(ids) < row:           # the reference's torch.cat would fail on mismatched lengths
            raise ValueError("Audio placeholders do not match the provided audio codes list.")
        out = torch.empty(row, 1 + n_vq, dtype=torch.int64, device=dev)
        if row == 0:
            return out
        text = torch.tensor(ids[:row], dtype=torch.int64).to(dev, non_blocking=True)
        bt = torch.tensor(blocks, dtype=torch.int32).reshape(-1, 4).to(dev, non_blocking=True) if blocks else None
        codes = torch.cat([c.to(device=dev, dtype=torch.int64) for c in audio_codes_list]).contiguous() \
            if audio_codes_list else None
        with torch.cuda.device(dev):
            L.check(L.lib().mt_pack_unified_codes(L.ptr(text), L.ptr(bt), len(blocks), L.ptr(codes), n_vq,
                                                  C.c_longlong(mc.audio_pad_code), row, L.ptr(out), L.stream_ptr(dev)),
                    "mt_pack_unified_codes")
        return out

    def _pad(self, input_ids_list: List[torch.Tensor]):
        """reference :410-431: left padding (pad_token_id in channel 0, audio_pad_code elsewhere) + bool attention mask."""
        dev = input_ids_list[0].device
        if dev.type != "cuda":
            raise L.MossTTSNativeError("_pad: input ids must live on a CUDA device (no CPU path)")
        rows = [t.to(torch.int64).contiguous() for t in input_ids_list]
        B, Cn = len(rows), int(rows[0].shape[1])
        Lmax = max(int(t.shape[0]) for t in rows)
        ptrs = torch.tensor([t.data_ptr() for t in rows], dtype=torch.int64).to(dev)
        lens = torch.tensor([int(t.shape[0]) for t in rows], dtype=torch.int32).to(dev)
        out = torch.empty(B, Lmax, Cn, dtype=torch.int64, device=dev)
        mask = torch.empty(B, Lmax, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            L.check(L.lib().mt_pad_left(L.ptr(ptrs), L.ptr(lens), B, Lmax, Cn, C.c_longlong(self.model_config.pad_token_id),
                                        C.c_longlong(self.model_config.audio_pad_code), L.ptr(out), L.ptr(mask),
                                        L.stream_ptr(dev)), "mt_pad_left")
        return {"input_ids": out, "attention_mask": mask.bool()}

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
