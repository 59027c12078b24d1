"""CPU-only: the C-ABI library builds/loads and exports every symbol include/mosstts.h declares."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from moss_tts_b200 import _lib as L


def _declared():
    hdr = open(os.path.join(ROOT, "include", "mosstts.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(mt_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    L.build()
    lib = ctypes.CDLL(L.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in mosstts.h but not exported"
    assert sorted(L.EXPORTS) == names
    lib.mt_version.restype = ctypes.c_int
    assert lib.mt_version() == 100        # no device needed


def test_no_cpu_fallback_in_product_path():
    """The product package must not import the oracle, and must refuse to run without CUDA."""
    import torch
    pkg = os.path.join(ROOT, "moss_tts_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert "import oracle" not in src and "from oracle" not in src, f
    if not torch.cuda.is_available():
        from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
        from moss_tts_b200.config import LMConfig
        with pytest.raises(L.MossTTSNativeError):
            MossTTSDelayModel(LMConfig.tiny(), {}, device="cuda:0")
        from moss_tts_b200.inference_utils import sample_token, find_last_equal_C
        with pytest.raises(L.MossTTSNativeError):
            sample_token(torch.zeros(2, 8, dtype=torch.bfloat16), do_sample=False)
        with pytest.raises(L.MossTTSNativeError):
            find_last_equal_C(torch.zeros(2, 8, dtype=torch.int64), 1)
