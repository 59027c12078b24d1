"""Mints tests/golden/prompt.npz by EXECUTING THE REFERENCE's prompt-side processor code
(/root/reference/moss_tts_delay/processing_moss_tts.py: loudness_normalize, _get_unified_codes, _pad,
encode_audios_from_wav) with a stub tokenizer (the real one is a hub asset) and the reference codec loaded with the
synthetic weights of oracle/weights.py.  Build container only.  TEST INFRASTRUCTURE ONLY.

  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden_prompt.py
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")

from moss_tts_b200.config import CodecConfig, LMConfig       # noqa: E402
from oracle.weights import make_codec_weights, _gen          # noqa: E402
from oracle.make_golden import _versions                     # noqa: E402
from oracle.make_golden_enc import build_ref_codec_full      # noqa: E402
from oracle.prompt import StubTokenizer                      # noqa: E402

CASES = [  # (role, content, segment lengths, truncation)
    ("user", "<|im_start|>user\nclone <|audio|> and <|audio|> please<|im_end|>\n", [5, 9], False),
    ("assistant", "<|im_start|>assistant\n<|audio|><|im_end|>\n", [7], False),
    ("assistant", "<|im_start|>assistant\n<|audio|>", [6], True),
    ("user", "<|im_start|>user\nno audio here<|im_end|>\n", [], False),
    ("assistant", "<|im_start|>assistant\n<|audio|> \n<|audio|>tail<|audio|><|im_end|>", [3, 4, 2], False),
]


def main():
    sys.path.insert(0, REF)
    from moss_tts_delay.processing_moss_tts import MossTTSDelayProcessor
    lm = LMConfig.tiny(n_vq=8)
    cfg = CodecConfig.mini()
    model = build_ref_codec_full(cfg, make_codec_weights(cfg, seed=11))
    tok = StubTokenizer()
    proc = object.__new__(MossTTSDelayProcessor)

    class _MC:
        pass
    mc = _MC()
    for k in ("pad_token_id", "audio_user_slot_token_id", "audio_assistant_gen_slot_token_id",
              "audio_assistant_delay_slot_token_id", "audio_start_token_id", "audio_end_token_id", "audio_pad_code",
              "n_vq", "sampling_rate"):
        setattr(mc, k, getattr(lm, k))
    proc.model_config = mc
    proc.tokenizer = tok
    proc.audio_tokenizer = model
    proc.audio_user_slot_token = tok.convert_ids_to_tokens(lm.audio_user_slot_token_id)
    proc.audio_assistant_gen_slot_token = tok.convert_ids_to_tokens(lm.audio_assistant_gen_slot_token_id)
    proc.audio_assistant_delay_slot_token = tok.convert_ids_to_tokens(lm.audio_assistant_delay_slot_token_id)
    proc.audio_start_token = tok.convert_ids_to_tokens(lm.audio_start_token_id)
    proc.audio_end_token = tok.convert_ids_to_tokens(lm.audio_end_token_id)

    g = _gen(21, "prompt")
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": 11, "n_cases": len(CASES)}
    unified = []
    for i, (role, content, lens, trunc) in enumerate(CASES):
        codes = [torch.randint(0, 1024, (n, lm.n_vq), generator=g) for n in lens]
        u = proc._get_unified_codes(role, content, codes, trunc)
        rec[f"case{i}_out"] = u.numpy()
        for j, c in enumerate(codes):
            rec[f"case{i}_codes{j}"] = c.numpy()
        unified.append(u)
    padded = proc._pad(unified[:3])
    rec["pad_ids"] = padded["input_ids"].numpy()
    rec["pad_mask"] = padded["attention_mask"].numpy()
    # loudness: quiet (gain clamps at +3 dB), loud (clamps at -3 dB), in range
    for i, amp in enumerate((0.01, 0.9, 0.1)):
        w = torch.randn(5000 + 37 * i, generator=g) * amp
        rec[f"loud{i}_in"] = w.numpy()
        rec[f"loud{i}_out"] = MossTTSDelayProcessor.loudness_normalize(w).numpy()
    # encode_audios_from_wav: stereo + mono, ragged lengths, at the model's sampling rate
    D = cfg.downsample_rate
    wavs = [torch.randn(2, 11 * D + 123, generator=g) * 0.05, torch.randn(1, 6 * D, generator=g) * 0.4]
    with torch.no_grad():
        codes_list = proc.encode_audios_from_wav(wavs, sampling_rate=lm.sampling_rate, n_vq=5)
    for i, (w, c) in enumerate(zip(wavs, codes_list)):
        rec[f"wav{i}"] = w.numpy()
        rec[f"wav{i}_codes"] = c.numpy()
    np.savez_compressed(os.path.join(OUT, "prompt.npz"), **rec)
    print("prompt golden ok", [rec[f"case{i}_out"].shape for i in range(len(CASES))], rec["pad_ids"].shape,
          [c.shape for c in codes_list])


if __name__ == "__main__":
    main()
