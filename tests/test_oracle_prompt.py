"""Pins oracle/prompt.py (prompt half of the processor) against reference-executed goldens (tests/golden/prompt.npz)."""
import os

import numpy as np

from conftest import GOLDEN
from moss_tts_b200.config import LMConfig
from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor
from oracle.make_golden_prompt import CASES
from oracle.prompt import StubTokenizer, loudness_normalize, unified_codes, pad_left


def _expanded_ids(proc, lm, role, content, codes):
    """placeholder expansion exactly as the product's host code does it (string work, no device)."""
    tok = proc.tokenizer
    if role == "user":
        gen = delay = tok.convert_ids_to_tokens(lm.audio_user_slot_token_id)
    else:
        gen = tok.convert_ids_to_tokens(lm.audio_assistant_gen_slot_token_id)
        delay = tok.convert_ids_to_tokens(lm.audio_assistant_delay_slot_token_id)
    if len(codes) > 1:
        import torch
        content, merged = proc._merge_consecutive_audio_placeholders(content, [torch.from_numpy(c) for c in codes])
        codes = [m.numpy() for m in merged]
    content = proc._replace_audio_placeholders(content, [len(c) for c in codes], lm.n_vq, gen, delay,
                                               tok.convert_ids_to_tokens(lm.audio_start_token_id),
                                               tok.convert_ids_to_tokens(lm.audio_end_token_id))
    return tok.encode(content), codes


def test_unified_codes_and_pad_match_reference():
    g = np.load(os.path.join(GOLDEN, "prompt.npz"), allow_pickle=False)
    lm = LMConfig.tiny(n_vq=8)
    proc = MossTTSDelayProcessor(lm, tokenizer=StubTokenizer())
    outs = []
    for i, (role, content, lens, trunc) in enumerate(CASES):
        codes = [g[f"case{i}_codes{j}"] for j in range(len(lens))]
        ids, codes = _expanded_ids(proc, lm, role, content, codes)
        u = unified_codes(ids, codes, lm.n_vq, trunc and role != "user", lm.audio_pad_code, lm.audio_start_token_id,
                          lm.audio_end_token_id)
        assert np.array_equal(u, g[f"case{i}_out"]), i
        outs.append(u)
    ids, mask = pad_left(outs[:3], lm.pad_token_id, lm.audio_pad_code)
    assert np.array_equal(ids, g["pad_ids"]) and np.array_equal(mask, g["pad_mask"])


def test_loudness_normalize_matches_reference():
    g = np.load(os.path.join(GOLDEN, "prompt.npz"), allow_pickle=False)
    for i in range(3):
        out = loudness_normalize(g[f"loud{i}_in"])
        assert np.abs(out - g[f"loud{i}_out"]).max() <= 1e-6 * max(1.0, np.abs(g[f"loud{i}_out"]).max())
