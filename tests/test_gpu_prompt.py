"""Prompt half of MossTTSDelayProcessor on the device vs reference-executed goldens."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200.config import CodecConfig, LMConfig
from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor
from oracle.make_golden_prompt import CASES
from oracle.prompt import StubTokenizer
from oracle.weights import make_codec_weights
from gpu_util import DEV

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup():
    g = np.load(os.path.join(GOLDEN, "prompt.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    codec = MossAudioTokenizerModel(cfg, make_codec_weights(cfg, seed=int(g["seed"])), device=DEV)
    lm = LMConfig.tiny(n_vq=8)
    return g, lm, MossTTSDelayProcessor(lm, audio_tokenizer=codec, tokenizer=StubTokenizer())


def test_unified_codes_bit_exact(setup):
    g, lm, proc = setup
    outs = []
    for i, (role, content, lens, trunc) in enumerate(CASES):
        codes = [torch.from_numpy(g[f"case{i}_codes{j}"]).to(DEV) for j in range(len(lens))]
        u = proc._get_unified_codes(role, content, codes, trunc)
        assert u.is_cuda and u.dtype == torch.int64
        assert np.array_equal(u.cpu().numpy(), g[f"case{i}_out"]), i
        outs.append(u)
    padded = proc._pad(outs[:3])
    assert np.array_equal(padded["input_ids"].cpu().numpy(), g["pad_ids"])
    assert padded["attention_mask"].dtype == torch.bool
    assert np.array_equal(padded["attention_mask"].cpu().numpy(), g["pad_mask"])
    with pytest.raises(ValueError):
        proc._get_unified_codes("assistant", "<|audio|><|audio|>x<|audio|>", [torch.zeros(2, 8, dtype=torch.long, device=DEV)], False)


def test_unified_codes_single_codebook_truncation_quirk(setup):
    """n_vq == 1: `[: -(n_vq - 1)]` is `[:-0]`, i.e. the whole delayed block is dropped (reference :626-629)."""
    g, lm, proc = setup
    codes = torch.arange(4, device=DEV).reshape(4, 1)
    u = proc._get_unified_codes("assistant", "ab<|audio|>", [codes], True)
    assert u.shape == (3, 2) and u[:, 1].tolist() == [1024, 1024, 1024]


def test_loudness_normalize(setup):
    g, lm, proc = setup
    for i in range(3):
        x = torch.from_numpy(g[f"loud{i}_in"]).to(DEV)
        out = proc.loudness_normalize(x).cpu().numpy()
        assert np.abs(out - g[f"loud{i}_out"]).max() <= 2e-6 * max(1.0, np.abs(g[f"loud{i}_out"]).max())
        assert torch.equal(x.cpu(), torch.from_numpy(g[f"loud{i}_in"]))          # input untouched
    assert proc.loudness_normalize(torch.zeros(0, device=DEV)).numel() == 0


def test_encode_audios_from_wav_matches_reference(setup):
    g, lm, proc = setup
    wavs = [torch.from_numpy(g["wav0"]).to(DEV), torch.from_numpy(g["wav1"]).to(DEV)]
    out = proc.encode_audios_from_wav(wavs, sampling_rate=lm.sampling_rate, n_vq=5)
    assert len(out) == 2
    agree = []
    for i, c in enumerate(out):
        ref = g[f"wav{i}_codes"]
        assert c.device.type == "cpu" and c.dtype == torch.long and tuple(c.shape) == ref.shape
        agree.append(float((c.numpy()[:, 0] == ref[:, 0]).mean()))
    print(f"[encode_audios_from_wav] first-quantiser agreement with the fp32 reference: {agree}")
    assert min(agree) >= 0.8            # TF32 encoder vs fp32 reference: codes may flip at near-ties only
