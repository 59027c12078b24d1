"""Streaming synthesis: chunked codec decode overlapped with LM decode.

Two pieces, both on the native path (libmosstts_sm100.so; no CPU fallback):

* ``AudioStreamDecoder`` mirrors moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py:682-799 (same
  constructor, ``push_tokens`` / ``audio_chunks`` / ``flush``, same crossfade rule): frames are buffered, every
  ``chunk_frames`` of them are decoded as an independent utterance with ``codec.decode(chunk_duration=0.32)`` and the
  chunk heads are cross-faded against the previous chunk's tail (``mt_crossfade``).
* ``generate_stream`` drives ``mt_lm_stream_begin / _steps`` (the sliced form of ``generate``) on the caller's stream
  and, on a second stream, de-delays the frames whose ``n_vq`` codes are final and pushes them through the codec's
  incremental decoder (``mt_codec_stream_push``, ring KV caches carried between pushes), so the first audio leaves
  after ``chunk_frames + n_vq - 1`` steps instead of after the last one.  The concatenated chunks equal one
  ``codec.streaming()`` decode of the finished codes, and the token ids equal ``generate``'s.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Iterator, List, Optional, Tuple

import numpy as np
import torch

from . import _lib as L


class AudioStreamDecoder:
    """Reference-shaped chunk decoder with crossfade (streaming_mossttsrealtime.py:682-799)."""

    def __init__(self, codec, chunk_frames: int = 40, overlap_frames: int = 4, decode_kwargs: Optional[dict] = None,
                 device: Optional[torch.device] = None):
        self.codec = codec
        self.chunk_frames = chunk_frames
        self.overlap_frames = overlap_frames
        self.decode_kwargs = decode_kwargs or {}
        self.device = device
        self._buffer: List[torch.Tensor] = []
        self._buffer_len = 0
        self._prev_tail: Optional[torch.Tensor] = None

    def push_tokens(self, audio_tokens):
        if isinstance(audio_tokens, np.ndarray):
            audio_tokens = torch.from_numpy(audio_tokens)
        if audio_tokens.dim() != 2:
            raise ValueError(f"Expected [T, C] audio tokens, got {tuple(audio_tokens.shape)}")
        self._buffer.append(audio_tokens)
        self._buffer_len += audio_tokens.shape[0]

    def audio_chunks(self) -> Iterable[torch.Tensor]:
        while self._buffer_len >= self.chunk_frames:
            wav = self._decode(self._consume_frames(self.chunk_frames), chunk_duration=0.32)
            yield self._apply_crossfade(wav)

    def flush(self) -> Optional[torch.Tensor]:
        if self._buffer_len == 0:
            return None
        wav = self._decode(self._consume_frames(self._buffer_len))
        return self._apply_crossfade(wav, final_chunk=True)

    def _consume_frames(self, num_frames: int) -> torch.Tensor:
        taken, left = [], num_frames
        while left > 0 and self._buffer:
            head = self._buffer[0]
            if head.shape[0] <= left:
                taken.append(self._buffer.pop(0))
                left -= head.shape[0]
            else:
                taken.append(head[:left])
                self._buffer[0] = head[left:]
                left = 0
        self._buffer_len -= num_frames - left
        return torch.cat(taken, dim=0)

    def _decode(self, tokens: torch.Tensor, chunk_duration: Optional[float] = 0.32) -> torch.Tensor:
        device = self.device if self.device is not None else getattr(self.codec, "device", None)
        if device is not None:
            tokens = tokens.to(device)
        kwargs = dict(self.decode_kwargs)
        if "chunk_duration" in kwargs:                       # caller override; None / <= 0 switches chunking off (:744-756)
            override = kwargs.pop("chunk_duration")
            try:
                override = None if override is None else float(override)
            except Exception:
                override = None
            chunk_duration = None if override is None or override <= 0 else override
        decoded = self.codec.decode(tokens.permute(1, 0), chunk_duration=chunk_duration, **kwargs)
        wav = decoded["audio"][0] if isinstance(decoded, dict) else getattr(decoded, "audio", decoded)
        if isinstance(wav, np.ndarray):
            wav = torch.from_numpy(wav)
        return wav.reshape(-1) if wav.dim() > 1 else wav

    def _overlap_samples(self, wav: torch.Tensor) -> int:
        if self.chunk_frames <= 0:
            return 0
        return int(wav.numel() * (self.overlap_frames / self.chunk_frames))

    def _apply_crossfade(self, wav: torch.Tensor, final_chunk: bool = False) -> torch.Tensor:
        if self.overlap_frames <= 0:
            return wav
        if self._prev_tail is None:
            self._prev_tail = wav[-self._overlap_samples(wav):].clone() if not final_chunk else None
            return wav
        overlap = min(self._overlap_samples(wav), self._prev_tail.numel())
        if overlap == 0:
            return wav
        if not wav.is_cuda:
            raise L.MossTTSNativeError("AudioStreamDecoder: waveforms must live on a CUDA device (no CPU path)")
        prev_tail = self._prev_tail.contiguous()
        wav = wav.contiguous()
        keep = prev_tail.numel() - overlap
        merged = torch.empty(keep + wav.numel(), dtype=torch.float32, device=wav.device)
        merged[:keep] = prev_tail[:keep]
        merged[keep + overlap:] = wav[overlap:]
        with torch.cuda.device(wav.device):
            L.check(L.lib().mt_crossfade(L.ptr(prev_tail[keep:]), L.ptr(wav), overlap, L.ptr(merged[keep:]),
                                         L.stream_ptr(wav.device)), "mt_crossfade")
        self._prev_tail = None if final_chunk else wav[-overlap:].clone()
        return merged


def scan_frames(codes: np.ndarray, next_row: List[int], n_rows: int, frame_rows: List[List[int]], pad_code: int) -> None:
    """Frames exactly as the processor defines them (processing_moss_tts.py:668-685): a frame is a de-delayed row
    `[codes[r + q, q] for q < n_vq]` that is NOT all pad; all-pad rows separate audio spans.  `codes` [B, >= n_rows, n_vq]
    holds the audio channels of every row known so far (prompt + generated, absolute row index), `next_row[b]` the first
    opening row not yet judged.  A row r can be judged once row r + n_vq - 1 exists; the opening rows of the frames
    found are appended to frame_rows[b] (ascending).  This covers what a gen-slot-only rule misses: the FIRST delay-slot
    row of a generated span still has every channel sampled (modeling_moss_tts.py:475-481 builds the mask from the
    pre-update delayed_lengths) and is a complete frame, and frames opened in the prompt whose diagonal reaches into the
    generated rows."""
    n_vq = codes.shape[2]
    q = np.arange(n_vq)
    for b in range(codes.shape[0]):
        hi = n_rows - n_vq + 1
        if hi <= next_row[b]:
            continue
        cand = np.arange(next_row[b], hi)
        diag = codes[b, cand[:, None] + q[None, :], q[None, :]]
        frame_rows[b].extend(int(r) for r in cand[(diag != pad_code).any(axis=1)])
        next_row[b] = hi


def scan_finished(new_tokens: np.ndarray, finished: List[bool], im_end_id: int) -> None:
    """A row that has emitted <|im_end|> produces no more frames (it is fed pad codes from then on)."""
    for b in range(new_tokens.shape[0]):
        if (new_tokens[b] == im_end_id).any():
            finished[b] = True


def plan_push(frame_rows: List[List[int]], finished: List[bool], pushed: int, chunk_frames: int,
              max_push: int) -> Optional[Tuple[int, List[int]]]:
    """Next lockstep push, or None.  frame_rows holds complete frames only (scan_frames).  Unfinished rows must all have
    `chunk_frames` frames beyond the `pushed` already sent; once every row has finished, whatever is left goes out.
    Returns (T, frames of each row inside the push)."""
    avail = [len(fr) - pushed for fr in frame_rows]
    live = [a for a, f in zip(avail, finished) if not f]
    T = min(live) if live else max(avail + [0])
    if T <= 0 or (live and T < chunk_frames):
        return None
    T = min(T, max_push)
    return T, [max(0, min(a, T)) for a in avail]


@torch.inference_mode()
def generate_stream(model, codec, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, *,
                    chunk_frames: int = 8, first_chunk_frames: Optional[int] = None, lm_slice: Optional[int] = None,
                    max_new_tokens: int = 1000,
                    text_temperature: float = 1.5, text_top_p: float = 1.0, text_top_k: int = 50,
                    audio_temperature: float = 1.7, audio_top_p: float = 0.8, audio_top_k: int = 25,
                    audio_repetition_penalty: float = 1.0, seed: int = 0,
                    use_cuda_graph: bool = True) -> Iterator[Tuple[torch.Tensor, List[int]]]:
    """Streaming form of ``model.generate`` + codec decode for a batch of B rows.

    Yields ``(wav [B, T * downsample_rate] float32 on the device, valid_samples [B])`` per chunk: the next T frames of
    every row (rows that have already finished are padded, ``valid_samples[b]`` says how much of the chunk is row b's).
    The frames of a row are exactly the ones ``processor.decode`` decodes for it: the de-delayed rows of the returned
    ids that are not all pad (every audio span in order, including the frame on the first delay-slot row of a span and
    frames of the returned prompt rows), so the streamed samples of a row total the samples of its decoded segments
    BEFORE the processor's ``start_length`` waveform trim (processing_moss_tts.py:692-707 needs the finished first
    segment's length, which a stream does not have; it is not applied here).
    A frame is pushed once all of its n_vq delayed codes exist, rows advance in lockstep, and a chunk leaves as soon
    as ``chunk_frames`` frames are final for every unfinished row (everything left at the end);
    ``first_chunk_frames`` (< ``chunk_frames``) lets the very first chunk leave earlier - short time to first audio with
    the throughput of large chunks afterwards.  The LM keeps one slice of ``lm_slice`` (default: the smaller of the
    two chunk sizes) steps queued on the caller's stream while the previous slice is scanned / gathered / decoded on a
    side stream.  Afterwards ``model.last_stream_output`` holds what ``generate``
    would have returned, ``[(start_length, ids)]``.  All audio spans of a row feed one codec context.
    """
    c = model.config
    if input_ids.dim() != 3 or input_ids.shape[-1] != c.n_vq + 1:
        raise ValueError("`Input_ids`'s shape should be exactly (batch_size, sequence_length, 1 + n_vq).")
    if chunk_frames < 1 or (first_chunk_frames is not None and first_chunk_frames < 1):
        raise ValueError("chunk_frames / first_chunk_frames must be >= 1")
    first_chunk = min(int(first_chunk_frames), chunk_frames) if first_chunk_frames else chunk_frames
    dev = model.device
    B, S0, Cc = input_ids.shape
    nvq = c.n_vq
    ids = input_ids.to(device=dev, dtype=torch.int64).contiguous()
    if attention_mask is None:
        attention_mask = torch.ones(B, S0, dtype=torch.bool)
    mask = attention_mask.to(device=dev).to(torch.uint8).contiguous()
    max_len = S0 + max_new_tokens
    model._ensure_engine(B, max_len, B * S0)
    gen = torch.empty(B, max_len, Cc, dtype=torch.int64, device=dev)
    starts = torch.empty(B, dtype=torch.int64, device=dev)
    status_dev = torch.zeros(2, 2, dtype=torch.int32, device=dev)
    text_host = torch.empty(B, max_new_tokens, dtype=torch.int64).pin_memory()
    codes_host = torch.empty(B, max_len, nvq, dtype=torch.int16).pin_memory()      # audio channels of every row, for scan_frames
    codes_host[:, :S0].copy_(ids[:, :, 1:].to(torch.int16))
    starts_host = torch.zeros(B, dtype=torch.int64).pin_memory()
    status_host = torch.zeros(2, dtype=torch.int32).pin_memory()
    gp = L.GenParamsC(max_new_tokens, float(text_temperature), float(text_top_p), int(text_top_k or 0),
                      float(audio_temperature), float(audio_top_p), int(audio_top_k or 0),
                      float(audio_repetition_penalty), seed & 0xFFFFFFFF, int(use_cuda_graph), 0)
    lib = L.lib()
    lm_stream = torch.cuda.current_stream(dev)
    lm_ptr = C.c_void_p(lm_stream.cuda_stream)
    side = torch.cuda.Stream(dev)
    slice_steps = int(lm_slice or first_chunk)
    max_push = max(chunk_frames, int(round(codec.causal_transformer_context_duration * codec.sampling_rate))
                   // codec.downsample_rate)
    frame_rows: List[List[int]] = [[] for _ in range(B)]      # absolute opening rows of the complete frames of each row
    next_row: Optional[List[int]] = None                      # first row not yet judged (starts at the returned slice)
    finished = [False] * B
    scanned = pushed = n_slices = 0
    launched = C.c_int(1)
    pending: List[Tuple[torch.cuda.Event, int]] = []
    all_done = False

    def enqueue(n: int):
        nonlocal n_slices
        slot = n_slices & 1
        with torch.cuda.device(dev):
            L.check(lib.mt_lm_stream_steps(model._engine, n, L.ptr(status_dev[slot]), C.byref(launched), lm_ptr),
                    "mt_lm_stream_steps")
        ev = torch.cuda.Event()
        ev.record(lm_stream)
        pending.append((ev, slot))
        n_slices += 1

    with torch.cuda.device(dev):
        L.check(lib.mt_lm_stream_begin(model._engine, L.ptr(ids), L.ptr(mask), B, S0, C.byref(gp), L.ptr(gen),
                                       L.ptr(starts), lm_ptr), "mt_lm_stream_begin")
    try:
        with codec.streaming(B):
            enqueue(0)                            # time_step 0 (prefill + first sample) is its own slice
            while True:
                if launched.value < max_new_tokens and not all_done and len(pending) < 2:
                    enqueue(slice_steps)          # one slice runs while the one before it is consumed below
                    continue
                if not pending:
                    break
                ev, slot = pending.pop(0)
                chunks = []
                with torch.cuda.stream(side):
                    side.wait_event(ev)
                    status_host.copy_(status_dev[slot], non_blocking=True)
                    side.synchronize()
                    steps_exec, done = int(status_host[0]), int(status_host[1])
                    if next_row is None:          # time_step 0 has run: gen_init's start indices are final
                        starts_host.copy_(starts, non_blocking=True)
                        side.synchronize()
                        next_row = [int(v) for v in starts_host]
                    if steps_exec > scanned:
                        text_host[:, scanned:steps_exec].copy_(gen[:, S0 + scanned:S0 + steps_exec, 0], non_blocking=True)
                        codes_host[:, S0 + scanned:S0 + steps_exec].copy_(
                            gen[:, S0 + scanned:S0 + steps_exec, 1:].to(torch.int16), non_blocking=True)
                        side.synchronize()
                        scan_finished(text_host[:, scanned:steps_exec].numpy(), finished, c.im_end_token_id)
                        scanned = steps_exec
                        scan_frames(codes_host.numpy(), next_row, S0 + scanned, frame_rows, c.audio_pad_code)
                    if done:
                        all_done = True
                        pending.clear()           # slices queued behind an all-stopped state change nothing
                    last = all_done or (not pending and launched.value >= max_new_tokens)
                    if last:
                        finished = [True] * B
                    while True:
                        first = pushed == 0 and first_chunk < chunk_frames
                        plan = plan_push(frame_rows, finished, pushed, first_chunk if first else chunk_frames,
                                         first_chunk if first else max_push)
                        if plan is None:
                            break
                        T, counts = plan
                        rows = torch.full((B, T), -1, dtype=torch.int32)
                        valid = [n * codec.downsample_rate for n in counts]
                        for b, n in enumerate(counts):
                            if n:
                                rows[b, :n] = torch.tensor(frame_rows[b][pushed:pushed + n], dtype=torch.int32)
                        rows_dev = rows.to(dev)
                        codes = torch.empty(nvq, B, T, dtype=torch.int64, device=dev)
                        L.check(lib.mt_stream_gather_frames(L.ptr(gen), B, max_len, Cc, 0, L.ptr(rows_dev), T, L.ptr(codes),
                                                            C.c_void_p(side.cuda_stream)), "mt_stream_gather_frames")
                        chunks.append((codec._decode_frame(codes).audio[:, 0], valid))
                        pushed += T
                    side.synchronize()
                for ch in chunks:                 # outside the side-stream context: the caller keeps its own stream
                    yield ch
                if last:
                    break
            lm_stream.synchronize()
    finally:
        lib.mt_lm_stream_end(model._engine)
    model.last_n_steps = scanned
    starts_h = starts.cpu().tolist()
    model.last_stream_output = [(S0 - starts_h[b], gen[b, starts_h[b]:S0 + scanned]) for b in range(B)]
