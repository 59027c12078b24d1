"""Oracle for the prompt half of MossTTSDelayProcessor (TEST INFRASTRUCTURE ONLY), numpy restatement of reference
moss_tts_delay/processing_moss_tts.py: loudness_normalize :736-748, _get_unified_codes :588-641 (tensor part),
_pad :410-431.  Pinned against reference-executed goldens (tests/golden/prompt.npz, oracle/make_golden_prompt.py)."""
from __future__ import annotations

import re
from typing import List

import numpy as np

from oracle.codec import apply_delay_pattern

SPECIAL = {"<|audio_start|>": 151652, "<|audio_end|>": 151653, "<|audio_user_slot|>": 151654,
           "<|audio_gen_slot|>": 151656, "<|audio_delay_slot|>": 151662, "<|im_start|>": 151644, "<|im_end|>": 151645}


class StubTokenizer:
    """Deterministic stand-in for the hub tokenizer: special-token strings map to their ids, every other character to
    1 + ord(c) % 1000.  Only `encode` / `convert_ids_to_tokens` are used by the code under test."""
    _pat = re.compile("(" + "|".join(re.escape(k) for k in SPECIAL) + ")")

    def encode(self, content: str) -> List[int]:
        ids: List[int] = []
        for part in self._pat.split(content):
            if part in SPECIAL:
                ids.append(SPECIAL[part])
            else:
                ids.extend(1 + ord(c) % 1000 for c in part)
        return ids

    def convert_ids_to_tokens(self, token_id: int) -> str:
        for k, v in SPECIAL.items():
            if v == token_id:
                return k
        raise KeyError(token_id)


def loudness_normalize(wav: np.ndarray, target_dbfs: float = -20.0, gain_range=(-3.0, 3.0)) -> np.ndarray:
    wav = wav.astype(np.float32)
    if wav.size == 0:
        return wav
    current = np.float32(10.0) * np.log10(np.mean(wav * wav, dtype=np.float32) + np.float32(1e-9))
    gain = float(target_dbfs - float(current))
    gain = max(gain_range[0], min(gain, gain_range[1]))
    return (wav * np.float32(10.0 ** (gain / 20.0))).astype(np.float32)


def unified_codes(ids: List[int], audio_codes_list: List[np.ndarray], n_vq: int, truncation: bool, pad_code: int,
                  audio_start: int, audio_end: int) -> np.ndarray:
    """Tensor part of _get_unified_codes :588-641 for an already tokenised, placeholder-expanded message."""
    text = np.asarray(ids, dtype=np.int64)
    starts = [i for i, t in enumerate(ids) if t == audio_start]
    ends = [i for i, t in enumerate(ids) if t == audio_end]
    assert len(starts) == len(ends) == len(audio_codes_list)
    if not audio_codes_list:
        audio = np.full((len(ids), n_vq), pad_code, dtype=np.int64)
    else:
        parts, prefix = [], 0
        for a, e, codes in zip(starts, ends, audio_codes_list):
            parts.append(np.full((a - prefix + 1, n_vq), pad_code, dtype=np.int64))
            parts.append(apply_delay_pattern(codes.astype(np.int64), pad_code))
            prefix = e
        if truncation:
            parts[-1] = parts[-1][: -(n_vq - 1)]
        else:
            parts.append(np.full((len(ids) - ends[-1], n_vq), pad_code, dtype=np.int64))
        audio = np.concatenate(parts)
    text = text[: audio.shape[0]]
    return np.concatenate([text[:, None], audio], axis=1)


def pad_left(rows: List[np.ndarray], pad_token: int, pad_code: int):
    L = max(r.shape[0] for r in rows)
    out = np.full((len(rows), L, rows[0].shape[1]), pad_code, dtype=np.int64)
    mask = np.zeros((len(rows), L), dtype=bool)
    for b, r in enumerate(rows):
        out[b, : L - r.shape[0], 0] = pad_token
        out[b, L - r.shape[0]:] = r
        mask[b, L - r.shape[0]:] = True
    return out, mask
