"""tcgen05/TMA GEMM vs a plain torch fp32 reference (and the SIMT cross-check kernel), through mt_gemm."""
import pytest
import torch

from moss_tts_b200 import _lib as L
from gpu_util import DEV, gemm

pytestmark = pytest.mark.gpu


def _ref(x, w, bias=None, act=0, residual=None, scale=None, round_bf16=0, out_dtype=None):
    v = x.float() @ w.float().T
    if bias is not None:
        v = v + bias
    if act == 1:
        v = torch.nn.functional.gelu(v)
    if round_bf16:
        v = v.to(torch.bfloat16).float()
    if residual is not None:
        v = residual.float() + (scale * v if scale is not None else v)
    elif scale is not None:
        v = scale * v
    return v


SHAPES = [(1, 256, 128), (16, 384, 256), (64, 1024, 512), (64, 1025 * 4, 256), (200, 512, 192), (256, 640, 1280),
          (300, 512, 256), (1000, 1536, 512), (130, 240, 768), (4096, 1024, 1024)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("impl", [L.MT_GEMM_SIMT, L.MT_GEMM_AUTO])
def test_gemm_bf16(M, N, K, impl):
    g = torch.Generator(device=DEV).manual_seed(M * 7 + N)
    x = torch.randn(M, K, device=DEV, generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(torch.bfloat16)
    out = gemm(x, w, impl=impl)
    ref = _ref(x, w)
    err = (out.float() - ref).abs().max().item()
    assert err <= 2e-2 * max(1.0, ref.abs().max().item()), err
    # fp32 output exposes accumulation differences without the bf16 rounding
    out32 = gemm(x, w, out_dtype=torch.float32, impl=impl)
    assert (out32 - ref).abs().max().item() <= 2e-3, (out32 - ref).abs().max().item()


@pytest.mark.parametrize("M,N,K", [(64, 512, 4096), (8, 4096, 12288)])
def test_gemm_bf16_splitk_matches_unsplit(M, N, K):
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(M, K, device=DEV, generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(torch.bfloat16)
    a = gemm(x, w, out_dtype=torch.float32, workspace_mb=64)
    b = gemm(x, w, out_dtype=torch.float32, workspace_mb=0)
    ref = _ref(x, w)
    assert (a - ref).abs().max().item() <= 2e-3
    assert (b - ref).abs().max().item() <= 2e-3


@pytest.mark.parametrize("M,N,K", [(32, 768, 512), (250, 1280, 768), (1000, 3072, 768), (6000, 2304, 768), (700, 768, 3072)])
@pytest.mark.parametrize("impl", [L.MT_GEMM_SIMT, L.MT_GEMM_AUTO])
def test_gemm_fp16_operands(M, N, K, impl):
    """MT_DTYPE_F16 inputs (the codec's layer GEMMs): fp32 / fp16 outputs, GELU, LayerScale + fp32 residual."""
    g = torch.Generator(device=DEV).manual_seed(M + N)
    x = torch.randn(M, K, device=DEV, generator=g).half()
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).half()
    res = torch.randn(M, N, device=DEV, generator=g)
    scale = torch.rand(N, device=DEV, generator=g)
    ref = x.float() @ w.float().T
    out = gemm(x, w, out_dtype=torch.float32, impl=impl)
    assert (out - ref).abs().max().item() <= 2e-3
    out = gemm(x, w, out_dtype=torch.float32, residual=res, scale=scale, impl=impl)
    assert (out - (res + scale * ref)).abs().max().item() <= 2e-3
    out = gemm(x, w, out_dtype=torch.float16, act=1, impl=impl)
    gref = torch.nn.functional.gelu(ref)
    assert out.dtype == torch.float16 and (out.float() - gref).abs().max().item() <= 4e-3


@pytest.mark.parametrize("M,N,K", [(64, 24576, 4096), (8, 512, 12288), (64, 384, 512)])
def test_gemm_partials_for_fused_consumer(M, N, K):
    """mt_gemm_args.out_splits: the engines' calling form - fp32 slabs [splits][M][N] in the workspace, no reduce kernel."""
    import ctypes as C
    g = torch.Generator(device=DEV).manual_seed(5)
    x = torch.randn(M, K, device=DEV, generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(torch.bfloat16)
    ws = torch.zeros(96 << 20, dtype=torch.uint8, device=DEV)
    out = torch.full((M, N), 7.0, dtype=torch.bfloat16, device=DEV)
    splits = C.c_int(0)
    n0 = L.launch_count()
    a = L.GemmArgs(L.MT_DTYPE_BF16, L.MT_DTYPE_BF16, M, N, K, x.data_ptr(), K, w.data_ptr(), K, out.data_ptr(), N,
                   None, 0, None, 0, None, 0, ws.data_ptr(), ws.numel(), L.MT_GEMM_AUTO, C.pointer(splits))
    L.check(L.lib().mt_gemm(C.byref(a), L.stream_ptr(torch.device(DEV))), "mt_gemm")
    torch.cuda.synchronize()
    assert L.launch_count() - n0 == 1 and splits.value >= 1
    part = ws.view(torch.float32)[: splits.value * M * N].view(splits.value, M, N)
    assert (part.sum(0) - _ref(x, w)).abs().max().item() <= 2e-3
    assert (out == 7.0).all()                      # `out` is left to the consumer


@pytest.mark.parametrize("M,N,K", [(32, 768, 512), (250, 1280, 768), (1000, 3072, 768), (700, 240, 768), (700, 768, 240), (40, 64, 240)])
@pytest.mark.parametrize("impl", [L.MT_GEMM_SIMT, L.MT_GEMM_AUTO])
def test_gemm_tf32_epilogues(M, N, K, impl):
    g = torch.Generator(device=DEV).manual_seed(N)
    x = torch.randn(M, K, device=DEV, generator=g)
    w = torch.randn(N, K, device=DEV, generator=g) * K ** -0.5
    bias = torch.randn(N, device=DEV, generator=g)
    scale = torch.randn(N, device=DEV, generator=g)
    res = torch.randn(M, N, device=DEV, generator=g)
    tol = 1e-4 if impl == L.MT_GEMM_SIMT else 1e-2      # TF32 inputs keep 10 mantissa bits
    for kw in (dict(), dict(bias=bias), dict(act=1), dict(residual=res, scale=scale), dict(bias=bias, act=1, residual=res)):
        out = gemm(x, w, impl=impl, **kw)
        ref = _ref(x, w, **kw)
        err = (out - ref).abs().max().item()
        assert err <= tol * max(1.0, ref.abs().max().item()), (kw.keys(), err)


def test_gemm_bf16_lm_epilogue():
    g = torch.Generator(device=DEV).manual_seed(11)
    M, N, K = 48, 512, 256
    x = torch.randn(M, K, device=DEV, generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(torch.bfloat16)
    res = torch.randn(M, N, device=DEV, generator=g).to(torch.bfloat16)
    out = gemm(x, w, residual=res, round_bf16=1)
    ref = _ref(x, w, residual=res, round_bf16=1).to(torch.bfloat16)
    # identical rounding points: at most 1 bf16 ulp apart, and almost everywhere identical
    diff = (out.float() - ref.float()).abs()
    assert diff.max().item() <= 2e-2
    assert (diff > 0).float().mean().item() < 0.02


@pytest.mark.skipif(__import__("os").environ.get("MT_FORCE_SIMT") == "1", reason="TMA alignment only applies to the tcgen05 path")
def test_gemm_rejects_bad_args():
    x = torch.zeros(4, 100, device=DEV, dtype=torch.bfloat16)     # K*2 bytes not a multiple of 16
    w = torch.zeros(8, 100, device=DEV, dtype=torch.bfloat16)
    with pytest.raises(ValueError):
        gemm(x, w)


@pytest.mark.parametrize("M,N,K,dt", [(6000, 2304, 768, torch.float32), (6144, 6144, 1024, torch.bfloat16),
                                      (333, 4000, 320, torch.bfloat16)])
def test_gemm_persistent_many_tiles(M, N, K, dt):
    """More 128x256 tiles than SMs: the persistent loop, TMEM double buffering and ring wrap-around."""
    g = torch.Generator(device=DEV).manual_seed(5)
    x = torch.randn(M, K, device=DEV, generator=g).to(dt)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(dt)
    out = gemm(x, w, out_dtype=torch.float32)
    ref = x.float() @ w.float().T
    tol = 2e-3 if dt == torch.bfloat16 else 1e-2
    assert (out - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())
    again = gemm(x, w, out_dtype=torch.float32)
    assert torch.equal(out, again)           # deterministic


@pytest.mark.parametrize("M,N,K,dt", [(64, 1025 * 4, 256, torch.bfloat16), (17, 384, 512, torch.bfloat16),
                                      (700, 1100, 512, torch.bfloat16), (900, 768, 256, torch.float32)])
@pytest.mark.parametrize("impl", [L.MT_GEMM_SIMT, L.MT_GEMM_AUTO])
def test_gemm_tile_major_weights(M, N, K, dt, impl):
    """MT_GEMM_W_TILED: same results from the tile-major weight layout (bit-identical to row-major)."""
    from moss_tts_b200.modeling_moss_tts import tile_major
    g = torch.Generator(device=DEV).manual_seed(9)
    x = torch.randn(M, K, device=DEV, generator=g).to(dt)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(dt)
    a = gemm(x, w, out_dtype=torch.float32, impl=impl, workspace_mb=0)
    b = gemm(x, tile_major(w), out_dtype=torch.float32, impl=impl | L.MT_GEMM_W_TILED | L.MT_GEMM_W_STATIC,
             workspace_mb=0, N=N)
    assert torch.equal(a, b)


@pytest.mark.parametrize("M,N,K,dt", [(256, 256, 64, torch.bfloat16), (257, 300, 192, torch.bfloat16), (6144, 6144, 1024, torch.bfloat16),
                                      (6144, 4096, 4096, torch.bfloat16), (1000, 3072, 768, torch.float16), (8000, 1280, 5120, torch.float16),
                                      (16000, 768, 768, torch.float16), (700, 240, 768, torch.float32), (6000, 2304, 768, torch.float32),
                                      (40000, 768, 256, torch.float16)])
def test_gemm_cta_pair_kernel_matches_one_sm_kernel(M, N, K, dt):
    """gemm_pair.cu (CTA pairs, 256 x 256 tiles, tcgen05.mma.cta_group::2: the product path for K >= 2048) against the one-SM
    kernel with both of its epilogues (warp-transposed coalesced stores / per-lane row stores) and a torch fp32 reference:
    same K order inside the tensor core and the same arithmetic per element, so all outputs are bit-identical; ragged M / N
    tiles, ring wrap-around, more pair tiles than pairs, every epilogue."""
    g = torch.Generator(device=DEV).manual_seed(M + 3 * N + K)
    x = torch.randn(M, K, device=DEV, generator=g).to(dt)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(dt)
    res = torch.randn(M, N, device=DEV, generator=g)
    scale = torch.rand(N, device=DEV, generator=g)
    bias = torch.randn(N, device=DEV, generator=g)
    half_out = torch.float16 if dt == torch.float16 else (torch.bfloat16 if dt == torch.bfloat16 else torch.float32)
    cases = [dict(out_dtype=torch.float32), dict(out_dtype=torch.float32, residual=res, scale=scale),
             dict(out_dtype=half_out, act=1, bias=bias)]
    if dt == torch.bfloat16:
        from moss_tts_b200.modeling_moss_tts import tile_major
        if K % 64 == 0:
            cases.append(dict(out_dtype=torch.float32, impl=L.MT_GEMM_AUTO | L.MT_GEMM_W_TILED | L.MT_GEMM_W_STATIC, N=N, tiled=True))
    ref = x.float() @ w.float().T
    tol = (2e-3 if dt != torch.float32 else 1e-2) * max(1.0, ref.abs().max().item())
    for kw in cases:
        kw = dict(kw)
        ww = tile_major(w) if kw.pop("tiled", False) else w
        try:
            L.lib().mt_gemm_debug_set_pair(4 + 2)            # one-SM kernel, per-lane row stores (the round-1 form)
            old = gemm(x, ww, **kw)
            L.lib().mt_gemm_debug_set_pair(4)                # one-SM kernel, warp-transposed epilogue
            mid = gemm(x, ww, **kw)
            L.lib().mt_gemm_debug_set_pair(1)                # CTA pairs
            new = gemm(x, ww, **kw)
            again = gemm(x, ww, **kw)
        finally:
            L.lib().mt_gemm_debug_set_pair(0)
        assert torch.equal(old, mid), (kw.keys(), (old.float() - mid.float()).abs().max().item())
        assert torch.equal(new, gemm(x, ww, **kw))           # the product's choice for this shape
        assert torch.isfinite(new.float()).all()
        assert torch.equal(old, new), (kw.keys(), (old.float() - new.float()).abs().max().item())
        if set(kw) <= {"out_dtype", "impl", "N"} and kw["out_dtype"] == torch.float32:
            assert (new - ref).abs().max().item() <= tol
        assert torch.equal(new, again)           # deterministic
