"""CPU restatement of the reference's chunked stream decoder, moss_tts_realtime/mossttsrealtime/
streaming_mossttsrealtime.py:682-799 (AudioStreamDecoder: frame buffer :700-735, per-chunk decode :737-774,
crossfade :776-799, overlap rule :801-804), on top of the oracle codec.  TEST INFRASTRUCTURE ONLY: imported by
tests/ (and nothing in the product path).  Pinned against tests/golden/stream_decoder_mini.npz, which
oracle/make_golden_asd.py mints by executing the reference class with the reference codec.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch


def linspace_fade_out(n: int) -> np.ndarray:
    """torch.linspace(1.0, 0.0, n) in fp32: start + i*step for i < n//2, end - (n-1-i)*step above."""
    if n == 1:
        return np.ones(1, dtype=np.float32)
    step = np.float64(np.float32(np.float32(0.0 - 1.0) / np.float32(n - 1)))
    i = np.arange(n, dtype=np.float64)
    # ATen evaluates base + step * i with a fused multiply-add: the f32 x f32 product is exact in f64, one rounding at the end
    lo = (1.0 + step * i).astype(np.float32)
    hi = (0.0 - step * (np.float64(n - 1) - i)).astype(np.float32)
    return np.where(np.arange(n) < n // 2, lo, hi).astype(np.float32)


def crossfade(prev_tail: np.ndarray, head: np.ndarray) -> np.ndarray:
    """:794-796: prev_tail * fade_out + head * (1 - fade_out), both [overlap] fp32."""
    fo = linspace_fade_out(prev_tail.shape[0])
    return (prev_tail * fo + head * (np.float32(1.0) - fo)).astype(np.float32)


class OracleStreamDecoder:
    def __init__(self, codec, chunk_frames: int = 40, overlap_frames: int = 4):
        self.codec, self.chunk_frames, self.overlap_frames = codec, chunk_frames, overlap_frames
        self.buf = np.zeros((0, 0), dtype=np.int64)
        self.prev_tail: Optional[np.ndarray] = None

    def push_tokens(self, tokens: np.ndarray):              # [T, C]
        self.buf = tokens.copy() if self.buf.size == 0 else np.concatenate([self.buf, tokens], 0)

    def _decode(self, tokens: np.ndarray) -> np.ndarray:    # :737-774 with chunk_duration = 0.32
        audio, lens = self.codec.decode(torch.from_numpy(tokens.T.copy()).long(), chunk_duration=0.32)
        return audio[0, 0, : int(lens[0])].float().numpy()

    def _overlap(self, wav: np.ndarray) -> int:             # :801-804
        return 0 if self.chunk_frames <= 0 else int(wav.size * (self.overlap_frames / self.chunk_frames))

    def _xfade(self, wav: np.ndarray, final: bool = False) -> np.ndarray:
        if self.overlap_frames <= 0:
            return wav
        if self.prev_tail is None:
            self.prev_tail = None if final else wav[wav.size - self._overlap(wav):].copy()
            return wav
        ov = min(self._overlap(wav), self.prev_tail.size)
        if ov == 0:
            return wav
        pt = self.prev_tail
        out = np.concatenate([pt[: pt.size - ov], crossfade(pt[pt.size - ov:], wav[:ov]), wav[ov:]])
        self.prev_tail = None if final else wav[wav.size - ov:].copy()
        return out

    def audio_chunks(self) -> List[np.ndarray]:
        out = []
        while self.buf.shape[0] >= self.chunk_frames:
            chunk, self.buf = self.buf[: self.chunk_frames], self.buf[self.chunk_frames:]
            out.append(self._xfade(self._decode(chunk)))
        return out

    def flush(self) -> Optional[np.ndarray]:
        if self.buf.shape[0] == 0:
            return None
        chunk, self.buf = self.buf, self.buf[:0]
        return self._xfade(self._decode(chunk), final=True)
