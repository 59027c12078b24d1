"""Helpers shared by the GPU parity tests."""
import ctypes as C
import json
import os

import numpy as np
import torch

from moss_tts_b200 import _lib as L

DEV = "cuda:0"


def gemm(x, w, *, out_dtype=None, bias=None, act=0, residual=None, scale=None, round_bf16=0, impl=L.MT_GEMM_AUTO,
         workspace_mb=64, N=None):
    M, K = x.shape
    N = w.shape[0] if N is None else N       # N must be given for a flat tile-major weight buffer
    code = {torch.bfloat16: L.MT_DTYPE_BF16, torch.float32: L.MT_DTYPE_F32, torch.float16: L.MT_DTYPE_F16}
    din = code[x.dtype]
    out_dtype = out_dtype or x.dtype
    dout = code[out_dtype]
    out = torch.empty(M, N, dtype=out_dtype, device=x.device)
    ws = torch.empty(workspace_mb << 20, dtype=torch.uint8, device=x.device) if workspace_mb else None
    a = L.GemmArgs(din, dout, M, N, K, x.data_ptr(), x.stride(0), w.data_ptr(), w.stride(0) if w.dim() == 2 else K, out.data_ptr(), N,
                   None if bias is None else bias.data_ptr(), act,
                   None if residual is None else residual.data_ptr(), 0 if residual is None else residual.stride(0),
                   None if scale is None else scale.data_ptr(), round_bf16,
                   None if ws is None else ws.data_ptr(), 0 if ws is None else ws.numel(), impl)
    L.check(L.lib().mt_gemm(C.byref(a), L.stream_ptr(x.device)), "mt_gemm")
    torch.cuda.synchronize()
    return out


def snr_db(ref, got):
    ref = ref.double().flatten()
    got = got.double().flatten()
    noise = (ref - got).pow(2).sum().item()
    sig = ref.pow(2).sum().item()
    return float("inf") if noise == 0 else 10.0 * np.log10(sig / noise)


def bf16_ulp(x: float) -> float:
    x = abs(float(x))
    if x == 0.0:
        return 2.0 ** -133
    return 2.0 ** (np.floor(np.log2(x)) - 7)


def classify_token_mismatch(mine, ref, trace, start_length, cfg):
    """First differing (row, channel) between two generated id matrices and the oracle's margin there,
    in bf16 ulps of the winning logit.  trace[step][channel][b] are the oracle's (already
    temperature-scaled) logits of that step."""
    bad = np.argwhere(mine != ref)
    if len(bad) == 0:
        return None
    r, c = bad[0]
    step = int(r) - int(start_length)
    return {"row": int(r), "channel": int(c), "step": step, "mine": int(mine[r, c]), "ref": int(ref[r, c])}
