"""A/B of the token-major GEMM kernels on one box: CTA pairs (gemm_pair.cu) vs the one-SM kernel (gemm_big.cu) on the codec's
and the prefill's shapes through mt_gemm, and the whole codec decode of the bench batch under both."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from moss_tts_b200 import _lib as L  # noqa: E402
from gpu_util import DEV  # noqa: E402

code = {torch.bfloat16: L.MT_DTYPE_BF16, torch.float32: L.MT_DTYPE_F32, torch.float16: L.MT_DTYPE_F16}


def time_gemm(M, N, K, dt, out_dt, act=0, residual=False, iters=20):
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(M, K, device=DEV, generator=g).to(dt)
    w = (torch.randn(N, K, device=DEV, generator=g) * K ** -0.5).to(dt)
    out = torch.empty(M, N, dtype=out_dt, device=DEV)
    res = torch.randn(M, N, device=DEV, generator=g) if residual else None
    scale = torch.rand(N, device=DEV, generator=g) if residual else None
    a = L.GemmArgs(code[dt], code[out_dt], M, N, K, x.data_ptr(), K, w.data_ptr(), K, out.data_ptr(), N, None, act,
                   None if res is None else res.data_ptr(), 0 if res is None else N, None if scale is None else scale.data_ptr(), 0,
                   None, 0, L.MT_GEMM_AUTO | L.MT_GEMM_W_STATIC)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    s = L.stream_ptr(DEV)
    out_ms = {}
    for mode in (1, 4, 6, 36):
        L.lib().mt_gemm_debug_set_pair(mode)
        for _ in range(3):
            L.check(L.lib().mt_gemm(C.byref(a), s), "mt_gemm")
        ts = []
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); L.check(L.lib().mt_gemm(C.byref(a), s), "mt_gemm"); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        out_ms[{4: "one_sm", 1: "pair", 6: "direct", 36: "ew8"}[mode]] = ts[len(ts) // 2]
    L.lib().mt_gemm_debug_set_pair(0)
    fl = 2.0 * M * N * K
    return {"M": M, "N": N, "K": K, "dtype": str(dt).split(".")[-1], "epilogue": "gelu->half" if act else ("scale+residual fp32" if residual else "plain"),
            "one_sm_us": round(out_ms["one_sm"] * 1e3, 1), "pair_us": round(out_ms["pair"] * 1e3, 1), "one_sm_direct_epilogue_us": round(out_ms["direct"] * 1e3, 1), "one_sm_8_epilogue_warps_us": round(out_ms["ew8"] * 1e3, 1),
            "one_sm_tflops": round(fl / out_ms["one_sm"] / 1e9, 1), "pair_tflops": round(fl / out_ms["pair"] / 1e9, 1)}


def main():
    rows = []
    # codec layer GEMMs at the bench batch (64 x 125 frames): stage widths 1280 (x1 rate), 768 (x2, x4, x8)
    for M, d, ff in [(8000, 1280, 5120), (16000, 768, 3072), (32000, 768, 3072), (64000, 768, 3072)]:
        rows.append(time_gemm(M, 3 * d, d, torch.float16, torch.float16))                       # in_proj
        rows.append(time_gemm(M, d, d, torch.float16, torch.float32, residual=True))            # out_proj
        rows.append(time_gemm(M, ff, d, torch.float16, torch.float16, act=1))                   # linear1
        rows.append(time_gemm(M, d, ff, torch.float16, torch.float32, residual=True))           # linear2
    # prefill (M = 64 x 96 tokens), Qwen3-8B widths
    for N, K in [(6144, 4096), (4096, 4096), (24576, 4096), (4096, 12288)]:
        rows.append(time_gemm(6144, N, K, torch.bfloat16, torch.bfloat16))
    for r in rows:
        print(json.dumps(r))
    # whole codec decode of the bench batch
    from moss_tts_b200.config import CodecConfig
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    codec = MossAudioTokenizerModel.random_init(CodecConfig.released(), device=DEV, seed=99, workspace_bytes=24 << 30)
    codes = torch.randint(0, 1024, (32, 64, 125), device=DEV)
    res = {}
    for mode in (6, 4, 1, 0, 16, 6, 4, 1, 0, 16):
        L.lib().mt_gemm_debug_set_pair(mode)
        for _ in range(2):
            codec.decode(codes)
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); codec.decode(codes); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        res.setdefault({4: "one_sm", 1: "pair", 6: "one_sm_direct_epilogue", 0: "product_choice", 16: "product_with_rotary_prepass"}[mode], []).append(round(ts[2], 2))
    L.lib().mt_gemm_debug_set_pair(0)
    print(json.dumps({"codec_decode_ms_64x125": res}))


if __name__ == "__main__":
    main()
