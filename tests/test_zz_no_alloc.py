"""mt_gemm_static_scratch (the allocation-free variant of mt_gemm; no counterpart in the reference)
against mt_gemm, through the C ABI.  Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

from moss_tts_b200 import _lib as L
from gpu_util import DEV

pytestmark = pytest.mark.gpu

TOL_F32 = 2e-3        # the fp32-output tolerance of tests/test_gpu_gemm.py
_CODE = {torch.bfloat16: L.MT_DTYPE_BF16, torch.float32: L.MT_DTYPE_F32, torch.float16: L.MT_DTYPE_F16}


def _args(x, w, out, workspace=None):
    M, K = x.shape
    N = w.shape[0]
    return L.GemmArgs(_CODE[x.dtype], _CODE[out.dtype], M, N, K, x.data_ptr(), x.stride(0), w.data_ptr(), w.stride(0),
                      out.data_ptr(), N, None, 0, None, 0, None, 0,
                      None if workspace is None else workspace.data_ptr(), 0 if workspace is None else workspace.numel(),
                      L.MT_GEMM_AUTO, None)


def _static(x, w, out):
    a = _args(x, w, out)
    L.check(L.lib().mt_gemm_static_scratch(C.byref(a), L.stream_ptr(x.device)), "mt_gemm_static_scratch")


def _existing(x, w, out, workspace):
    a = _args(x, w, out, workspace)
    L.check(L.lib().mt_gemm(C.byref(a), L.stream_ptr(x.device)), "mt_gemm")


@pytest.mark.parametrize("M,N,K,dtype", [(64, 4096, 4096, torch.bfloat16), (8, 1024, 12288, torch.bfloat16),
                                         (32, 768, 3072, torch.float16)])
def test_static_scratch_gemm_matches_mt_gemm_twice(M, N, K, dtype):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    # everything is allocated BEFORE the calls under test; the existing entry point gets a caller workspace of the same size
    # as the library's static scratch, so both pick the same split count and run the same kernels
    ws = torch.empty(L.MT_GEMM_STATIC_SCRATCH_BYTES, dtype=torch.uint8, device=DEV)
    xs = [torch.randn(M, K, device=DEV, generator=g).to(dtype) for _ in range(2)]
    ws_ = [(torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(dtype) for _ in range(2)]
    want = [torch.empty(M, N, dtype=torch.float32, device=DEV) for _ in range(2)]
    got = [torch.full((M, N), float("nan"), dtype=torch.float32, device=DEV) for _ in range(2)]
    for i in range(2):
        _existing(xs[i], ws_[i], want[i], ws)
    torch.cuda.synchronize()
    _static(xs[0], ws_[0], got[0])          # first call on this device: resolves the symbol address, loads the kernels
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    _static(xs[1], ws_[1], got[1])          # second call in a row, different operands: the scratch holds call 1's partials
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    for i in range(2):
        ref = xs[i].float() @ ws_[i].float().T
        assert bool(torch.isfinite(got[i]).all())
        err = (got[i] - want[i]).abs().max().item()
        assert err <= TOL_F32, f"call {i + 1}: mt_gemm_static_scratch differs from mt_gemm by {err}"
        assert (got[i] - ref).abs().max().item() <= TOL_F32
        assert torch.equal(got[i], want[i]), f"call {i + 1}: same kernels and split count must give the same bits"
    assert free_after >= free_before, "the second call took device memory"
    # and once more with the first operands: the result does not depend on what the scratch held
    again = torch.empty_like(got[0])
    _static(xs[0], ws_[0], again)
    torch.cuda.synchronize()
    assert torch.equal(again, got[0])


def test_static_scratch_gemm_rejects_what_does_not_fit():
    x = torch.zeros(256, 128, dtype=torch.bfloat16, device=DEV)
    w = torch.zeros(8192, 128, dtype=torch.bfloat16, device=DEV)
    out = torch.empty(256, 8192, dtype=torch.float32, device=DEV)
    with pytest.raises(ValueError, match="MT_GEMM_STATIC_SCRATCH_BYTES"):      # 2 * 256 * 8192 * 4 = 16 MiB > 8 MiB
        _static(x, w, out)
    x = torch.zeros(300, 128, dtype=torch.bfloat16, device=DEV)
    w = torch.zeros(256, 128, dtype=torch.bfloat16, device=DEV)
    out = torch.empty(300, 256, dtype=torch.float32, device=DEV)
    with pytest.raises(ValueError, match="exceeds 256"):
        _static(x, w, out)
    ws = torch.empty(1 << 20, dtype=torch.uint8, device=DEV)
    a = _args(x[:64], w, out[:64], ws)
    with pytest.raises(ValueError, match="must be NULL"):
        L.check(L.lib().mt_gemm_static_scratch(C.byref(a), L.stream_ptr(x.device)), "mt_gemm_static_scratch")
