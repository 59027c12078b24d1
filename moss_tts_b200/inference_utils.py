"""Mirror of reference moss_tts_delay/inference_utils.py (:111-165) on top of the C ABI.

`sample_token` keeps the reference's argument names.  Differences that are part of the
contract: logits must be bf16 CUDA tensors; the random draw is an
inverse-CDF over (prob desc, index asc) driven by `uniforms` (or a seeded generator) rather
than torch.multinomial's Philox stream.  Any top_k is accepted: 1..1024 runs on the shared-memory candidate list,
`top_k=None` / `<= 0` (no filter, reference :135) and `top_k > 1024` on the full-vocabulary counting sort
(csrc/sampling_full.cuh, mt_sample_tokens_full).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L


def _bitmap_from_tokens(prev_tokens: torch.Tensor, vocab: int) -> torch.Tensor:
    """Union of every id in prev_tokens as a device bitmap (what torch.unique feeds the penalty,
    reference :78-86)."""
    words = (vocab + 31) // 32
    flat = prev_tokens.reshape(-1).to(torch.int64)
    flat = flat[(flat >= 0) & (flat < vocab)]
    present = torch.zeros(words * 32, dtype=torch.bool, device=flat.device)
    present[flat] = True
    weights = (1 << torch.arange(32, device=flat.device, dtype=torch.int64))
    packed = (present.view(words, 32).to(torch.int64) * weights).sum(dim=1)
    return packed.to(torch.int32)        # two's-complement wrap keeps the low 32 bits


def _launch(x: torch.Tensor, bitmap, repetition_penalty, top_p, top_k, do_sample, uniforms, out: torch.Tensor) -> None:
    """x [N, V] bf16 (row stride x.stride(0), unit column stride) -> out [N] int64 (contiguous)."""
    N, V = x.shape
    k = int(top_k or 0)
    scratch = None
    if do_sample and (k > 1024 or (k <= 0 and V > 1024)):
        scratch = torch.empty(L.lib().mt_sample_full_scratch_bytes(N), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.lib().mt_sample_tokens_full(L.ptr(x), C.c_longlong(x.stride(0)), N, V, C.c_float(1.0), L.ptr(bitmap),
                                              C.c_float(repetition_penalty if bitmap is not None else 1.0),
                                              C.c_float(1.0 if top_p is None else top_p), k, int(do_sample),
                                              L.ptr(uniforms if do_sample else None), L.ptr(out), L.ptr(scratch),
                                              C.c_longlong(0 if scratch is None else scratch.numel()),
                                              L.stream_ptr(x.device)), "mt_sample_tokens_full")


def sample_token(logits: torch.Tensor, prev_tokens: Optional[torch.Tensor] = None, repetition_penalty: float = 1.0,
                 top_p=None, top_k=None, do_sample: bool = True, *, uniforms: Optional[torch.Tensor] = None,
                 generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """reference sample_token (:111-145): penalty -> (greedy: argmax) -> top-k -> top-p -> softmax -> draw.
    Repetition penalty (:62-108): 2-D logits [N, V] take the union of ALL prev_tokens ("case 1"); 3-D logits [B, H, V]
    take, for head h, the ids of prev_tokens[..., h] only ("case 2"), applied to every batch row of that head."""
    if not logits.is_cuda:
        raise L.MossTTSNativeError("sample_token: logits must live on a CUDA device (no CPU path)")
    if logits.dtype != torch.bfloat16:
        raise NotImplementedError("sample_token: the CUDA sampler takes bf16 logits (the model dtype; its rounding points are "
                                  "the reference's bf16 ones)")
    shape = logits.shape[:-1]
    V = logits.shape[-1]
    N = 1
    for d in shape:
        N *= int(d)
    out = torch.empty(N, dtype=torch.int64, device=logits.device)
    if N == 0:
        return out.view(shape)
    penalise = prev_tokens is not None and repetition_penalty != 1.0
    if penalise and logits.dim() not in (2, 3):
        raise AssertionError("Delay Pattern audio logits must be [B, H, V]")        # the reference's assert (:90)
    if do_sample:
        if top_k is None or top_k <= 0:
            top_k = 0                                # no top-k filter (:135)
        if uniforms is None:
            uniforms = torch.rand(N, device=logits.device, generator=generator, dtype=torch.float32)
        uniforms = uniforms.to(device=logits.device, dtype=torch.float32).reshape(-1).contiguous()
    if penalise and logits.dim() == 3:
        Bh, H, _ = logits.shape
        prev = prev_tokens.to(logits.device)
        x3 = logits if logits.stride(-1) == 1 else logits.contiguous()
        out2 = out.view(Bh, H)
        for h in range(H):                       # one id set per head, shared by the batch rows of that head
            bitmap = _bitmap_from_tokens(prev[..., h], V).contiguous()
            oh = torch.empty(Bh, dtype=torch.int64, device=logits.device)
            uh = uniforms.view(Bh, H)[:, h].contiguous() if do_sample else None
            _launch(x3[:, h, :], bitmap, repetition_penalty, top_p, top_k, do_sample, uh, oh)
            out2[:, h] = oh
        return out.view(shape)
    x = logits.reshape(-1, V)
    if x.stride(-1) != 1:
        x = x.contiguous()
    bitmap = _bitmap_from_tokens(prev_tokens.to(x.device), V).contiguous() if penalise else None
    _launch(x, bitmap, repetition_penalty, top_p, top_k, do_sample, uniforms, out)
    return out.view(shape)


def find_last_equal_C(tensor: torch.Tensor, C_value: int) -> torch.Tensor:
    """tensor [B,S] int64 (CUDA) -> [B] last index equal to C, -1 if absent."""
    if not tensor.is_cuda:
        raise L.MossTTSNativeError("find_last_equal_C: tensor must live on a CUDA device (no CPU path)")
    t = tensor.to(torch.int64)
    if t.stride(-1) != 1:
        t = t.contiguous()
    B, S = t.shape
    out = torch.empty(B, dtype=torch.int64, device=t.device)
    if B == 0:
        return out
    if S == 0:
        return out.fill_(-1)
    with torch.cuda.device(t.device):
        L.check(L.lib().mt_find_last_equal(L.ptr(t), B, S, C.c_longlong(t.stride(0)), C.c_longlong(int(C_value)),
                                           L.ptr(out), L.stream_ptr(t.device)), "mt_find_last_equal")
    return out
