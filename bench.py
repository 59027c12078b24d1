"""bench.py — MOSS-TTS synthesis hot path on B200 (contract: see DESIGN.md §Measurement).

Workload (BASELINE.json configs[1]): MossTTSDelay, assumed released 8B shape (Qwen3-8B backbone,
n_vq=32, random-init bf16 weights, never-stop head tie), batch 64 utterances per GPU, 96-token
prompt, 156 greedy steps = 125 codec frames = 10 s of audio per utterance, then
MOSS-Audio-Tokenizer (released config.json shape; fp32 residual stream and accumulation, fp16-operand
tensor-core GEMMs / attention in the transformer blocks) RVQ dequantise + decoder to a 24 kHz waveform.  One "step" = that whole pass over one batch.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
For N > 1 launch with torchrun (one rank per GPU); ranks shard the utterances (weak scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch                                                          # noqa: E402

FRAME_RATE = 12.5
PROMPT_LEN, NEW_TOKENS, FRAMES = 96, 156, 125
BATCH_PER_GPU = 64
METRIC = "generated audio-sec/sec"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), "measured"
    return 6650.0, 1590.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_prompts(cfg, batch, seed):
    g = torch.Generator().manual_seed(seed)
    ids = torch.full((batch, PROMPT_LEN, 1 + cfg.n_vq), cfg.audio_pad_code, dtype=torch.int64)
    ids[:, :, 0] = torch.randint(0, cfg.pad_token_id, (batch, PROMPT_LEN), generator=g)
    ids[:, PROMPT_LEN - 4, 0] = cfg.im_start_token_id
    ids[:, PROMPT_LEN - 1, 0] = cfg.audio_start_token_id
    return ids, torch.ones(batch, PROMPT_LEN, dtype=torch.bool)


# ------------------------------------------------------------------------------------------------
# CPU baseline: the oracle (a port of the reference's PyTorch path) on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_baseline(decode_steps: int = 3):
    """B=1 sample of the same workload at the full shapes: prefill(96) + `decode_steps` decode steps of
    the 8B-shape LM (fp32, eager attention: the reference's CPU choice, README.md:218-219) + a full
    10 s codec decode, extrapolated to 156 steps.  All 36 layers alias ONE layer's weight tensors
    (772 MB > LLC, so timing equals distinct layers) to avoid a 34 GB / 140 s init."""
    from moss_tts_b200.config import LMConfig, CodecConfig
    from oracle.lm import OracleLM
    from oracle.codec import OracleCodec
    from oracle.weights import make_codec_weights
    torch.manual_seed(0)
    cfg = LMConfig.delay_8b()
    H, I = cfg.hidden_size, cfg.intermediate_size

    def rn(*s):
        return torch.randn(*s) * 0.02
    layer = {"self_attn.q_proj.weight": rn(cfg.q_size, H), "self_attn.k_proj.weight": rn(cfg.kv_size, H),
             "self_attn.v_proj.weight": rn(cfg.kv_size, H), "self_attn.o_proj.weight": rn(H, cfg.q_size),
             "self_attn.q_norm.weight": torch.ones(cfg.head_dim), "self_attn.k_norm.weight": torch.ones(cfg.head_dim),
             "mlp.gate_proj.weight": rn(I, H), "mlp.up_proj.weight": rn(I, H), "mlp.down_proj.weight": rn(H, I),
             "input_layernorm.weight": torch.ones(H), "post_attention_layernorm.weight": torch.ones(H)}
    w = {"language_model.embed_tokens.weight": rn(cfg.vocab_size, H), "language_model.norm.weight": torch.ones(H)}
    for l in range(cfg.num_layers):
        for k, v in layer.items():
            w[f"language_model.layers.{l}.{k}"] = v
    emb = rn(cfg.audio_head_size, H)
    head = rn(cfg.audio_head_size, H)
    for i in range(cfg.n_vq):
        w[f"emb_ext.{i}.weight"] = emb
        w[f"lm_heads.{i + 1}.weight"] = head
    w["lm_heads.0.weight"] = rn(cfg.vocab_size, H)
    w["lm_heads.0.weight"][cfg.audio_assistant_delay_slot_token_id] = w["lm_heads.0.weight"][cfg.audio_assistant_gen_slot_token_id]
    lm = OracleLM.__new__(OracleLM)
    lm.cfg, lm.dtype, lm.w = cfg, torch.float32, w                       # no per-tensor copies
    half = torch.arange(0, cfg.head_dim, 2, dtype=torch.int64).float() / cfg.head_dim
    lm.inv_freq = 1.0 / (cfg.rope_theta ** half)
    ids, mask = make_prompts(cfg, 1, 0)
    cache = [None] * cfg.num_layers
    t0 = time.perf_counter()
    lm.forward(ids, mask, cache, 0, last_only=True)
    t_prefill = time.perf_counter() - t0
    step_ids = torch.full((1, 1, 1 + cfg.n_vq), 7, dtype=torch.int64)
    ts = []
    for s in range(decode_steps):
        m = torch.ones(1, PROMPT_LEN + s + 1, dtype=torch.bool)
        t0 = time.perf_counter()
        lm.forward(step_ids, m, cache, PROMPT_LEN + s, last_only=True)
        ts.append(time.perf_counter() - t0)
    t_step = statistics.median(ts)
    del lm, w, layer, cache
    ccfg = CodecConfig.released()
    oc = OracleCodec(ccfg, make_codec_weights(ccfg, seed=0, layer_scale=0.01))
    codes = torch.randint(0, 1024, (32, 1, FRAMES))
    t0 = time.perf_counter()
    oc.decode(codes, chunk_duration=8)
    t_codec = time.perf_counter() - t0
    total = t_prefill + (NEW_TOKENS - 1) * t_step + t_codec
    return {"value": FRAMES / FRAME_RATE / total, "unit": "audio-s/s", "cores": torch.get_num_threads(),
            "kind": "port",
            "sample": f"B=1 at full 8B/codec shapes, fp32: prefill(96)={t_prefill:.2f}s + median of {decode_steps} decode "
                      f"steps={t_step:.3f}s extrapolated to {NEW_TOKENS - 1} + codec 10s={t_codec:.2f}s",
            "prefill_s": t_prefill, "s_per_decode_step": t_step, "codec_s_per_10s": t_codec}


# ------------------------------------------------------------------------------------------------
# dominant kernel: the decode-orientation tcgen05 GEMM streaming one layer's gate/up weights
# ------------------------------------------------------------------------------------------------
def time_dominant_kernel(model, batch, iters=3):
    """Times gemm_dec_kernel on the largest per-layer decode GEMM (gate|up: [2I, H] weights, M = batch) the way the
    engine launches it (mt_gemm with out_splits: fp32 split-K partials left for the consumer kernel, ONE kernel per
    call), cycling over all layers' weights so every launch streams from HBM (36 x 201 MB >> 126 MB L2).
    Returns (avg ms per launch, algorithmic bytes per launch, splits, avg ms incl. the stand-alone reduce kernel)."""
    import ctypes as C
    from moss_tts_b200 import _lib as L
    cfg = model.config
    H, N = cfg.hidden_size, 2 * cfg.intermediate_size
    dev = model.device
    x = torch.randn(batch, H, device=dev).to(torch.bfloat16)
    out = torch.empty(batch, N, dtype=torch.bfloat16, device=dev)
    ws = torch.empty(96 << 20, dtype=torch.uint8, device=dev)
    sp = L.stream_ptr(dev)
    splits = C.c_int(0)

    def launch(w, partial):
        a = L.GemmArgs(L.MT_DTYPE_BF16, L.MT_DTYPE_BF16, batch, N, H, x.data_ptr(), H, w.data_ptr(), H, out.data_ptr(), N,
                       None, 0, None, 0, None, 0, ws.data_ptr(), ws.numel(),
                       L.MT_GEMM_W_STATIC | (L.MT_GEMM_W_TILED if model.weights_tiled else 0),
                       C.pointer(splits) if partial else None)
        L.check(L.lib().mt_gemm(C.byref(a), sp), "mt_gemm")

    def run(partial):
        for w in model._w["gu"]:
            launch(w, partial)                         # warm-up pass
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 0
        for _ in range(iters):
            for w in model._w["gu"]:
                launch(w, partial)
                n += 1
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / n
    ms_full = run(False)                               # finished bf16 tensor: GEMM + split-K reduce kernel
    ms = run(True)                                     # the kernel alone, as inside a decode step
    alg = N * H * 2 + batch * H * 2 + splits.value * batch * N * 4
    return ms, alg, splits.value, ms_full


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel's launch, from the committed `ncu --set full`
    capture of this round (profiles/ncu_traffic.json, written by scripts/gpu_job_profile.sh); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return float(json.load(f)["dram_bytes_per_launch"])
    except Exception:
        return None


def run_ours(args):
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.config import LMConfig, CodecConfig
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback for the product arm)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=dev)
    L.build()
    B = args.batch
    cfg = LMConfig.delay_8b() if not args.small else LMConfig.mid(32)
    ccfg = CodecConfig.released() if not args.small else CodecConfig.mini(context_duration=10.0, num_quantizers=32)
    model = MossTTSDelayModel.random_init(cfg, device=dev, seed=1234 + rank, never_stop=True)
    codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=24 << 30)
    proc = MossTTSDelayProcessor(cfg, audio_tokenizer=codec)
    ids_h, mask_h = make_prompts(cfg, B, seed=rank)
    ids_h, mask_h = ids_h.pin_memory(), mask_h.pin_memory()
    ids_d, mask_d = ids_h.to(dev), mask_h.to(dev)
    gen_kw = dict(max_new_tokens=NEW_TOKENS, text_temperature=0, audio_temperature=0)
    wav_h = torch.empty(B, FRAMES * ccfg.downsample_rate, dtype=torch.float32).pin_memory()

    def step_device():
        out = model.generate(ids_d, mask_d, **gen_kw)
        return proc.decode_device(out)

    def step_e2e():
        out = model.generate(ids_h, mask_h, **gen_kw)          # H2D of the prompt inside
        wavs = proc.decode_device(out)
        n = 0
        for b, ws in enumerate(wavs):
            w = ws[0]
            wav_h[b, : w.numel()].copy_(w, non_blocking=True)   # D2H of the result
            n += w.numel()
        if dist is not None:                                     # NCCL only gathers outputs (lengths here)
            lens = torch.tensor([n], device=dev)
            allv = [torch.zeros_like(lens) for _ in range(world)]
            dist.all_gather(allv, lens)
        torch.cuda.current_stream(dev).synchronize()
        return n

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, r

    for _ in range(args.warmup):
        wavs = step_device()
    step_e2e()
    frames = sum(w[0].numel() for w in wavs) / ccfg.downsample_rate
    audio_s_per_step = B * FRAMES / FRAME_RATE              # codec-decoded frames / 12.5
    sampler = ClockSampler(local)
    sampler.start()
    l0 = L.launch_count()
    ms_dev, _ = timed(step_device, args.steps)
    launches = L.launch_count() - l0
    ms_e2e, n_samples = timed(step_e2e, args.steps)
    clocks = sampler.stop()
    # per-step decode latency from one instrumented run (events between steps)
    out = model.generate(ids_d, mask_d, record_step_times=True, **gen_kw)
    step_ms = model.last_step_ms or []
    dec = sorted(step_ms[2:]) if len(step_ms) > 3 else [float("nan")]
    dec_med = dec[len(dec) // 2]
    prefill_ms = step_ms[0] if step_ms else float("nan")
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record(); proc.decode_device(out); t1.record(); torch.cuda.synchronize(dev)
    codec_ms = t0.elapsed_time(t1)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    hbm, tflops, src = measured_peaks()
    k_ms, k_bytes, k_splits, k_ms_full = time_dominant_kernel(model, B)
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    weight_bytes = sum(t.numel() for k in ("qkv", "o", "gu", "down") for t in model._w[k]) * 2 + \
        (model._w["text_head"].numel() + model._w["audio_heads"].numel()) * 2
    line = {
        "metric": METRIC, "value": world * audio_s_per_step * args.steps / (ms_dev * 1e-3), "unit": "audio-s/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "configs[1]: MossTTSDelay batch-64 x 10 s utterances bf16 (greedy, 96-token prompt, 156 steps) "
                               "+ MOSS-Audio-Tokenizer decode to 24 kHz" + (" [SMALL DEBUG SHAPES]" if args.small else ""),
                   "lm_shape": "assumed released 8B = Qwen3-8B backbone H4096 L36 32q/8kv x128 I12288 n_vq32 (hub-only config)"
                   if not args.small else "mid debug shape",
                   "codec": "config.json decoder (32+12+12+12 layers): fp32 residual stream / accumulation / softmax / LayerNorm, "
                            "fp16-operand tensor-core GEMMs and attention in the transformer blocks (TF32's mantissa), TF32 stage "
                            "projections; waveform SNR vs the fp32 reference 48.9-50.9 dB (gate: >= 40 dB)",
                   "batch_per_gpu": B, "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "frames": FRAMES,
                   "sharding": f"request-sharded x{world}, no in-step collective",
                   "l2_policy": "inputs larger than L2: 15.4 GB of weights streamed every decode step"},
        "e2e": {"value": world * audio_s_per_step * args.steps / (ms_e2e * 1e-3), "unit": "audio-s/s",
                "h2d_bytes_per_step": ids_h.numel() * 8 + mask_h.numel(), "d2h_bytes_per_step": int(n_samples) * 4,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "decode_ms_per_step": dec_med, "prefill_ms": prefill_ms, "codec_ms": codec_ms,
        "decode_weight_stream_gbs": weight_bytes / (dec_med * 1e-3) / 1e9 if dec_med == dec_med else None,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                     "traffic": ncu_traffic(), "kernel": "gemm_dec_kernel<BN=64,bf16> gate|up decode GEMM (M=%d,N=%d,K=%d) through mt_gemm as the engine "
                               "launches it: %d split-K fp32 partial slabs left for the fused consumer" % (B, 2 * cfg.intermediate_size, cfg.hidden_size, k_splits),
                     "ms_per_launch": k_ms, "algorithmic_bytes": k_bytes, "peak_source": src,
                     "ms_per_call_with_own_reduce_kernel": k_ms_full},
        "clocks": clocks,
    }
    if not args.no_cpu:
        try:
            line["cpu_baseline"] = cpu_baseline() if not args.small else {"value": None, "kind": "port", "sample": "skipped (--small)"}
        except Exception as e:           # the baseline must not take the product number down with it
            line["cpu_baseline"] = {"value": None, "kind": "port", "error": repr(e)[:200]}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path (the oracle port; the reference is
    pure Python/PyTorch and cannot be shipped to the box), all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    res = []
    for _ in range(max(1, min(args.steps, 2))):
        res.append(cpu_baseline(decode_steps=2))
    best = max(res, key=lambda r: r["value"])
    v = best["value"]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "audio-s/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": FRAMES / FRAME_RATE / v * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1]: MossTTSDelay batch-64 x 10 s utterances bf16 (greedy, 96-token prompt, 156 steps) "
                                   "+ MOSS-Audio-Tokenizer decode to 24 kHz",
                       "sample": "bounded sample of that workload on the host cores: B=1 at the full 8B / codec shapes, fp32 "
                                 "(the reference's CPU dtype), prefill + timed decode steps extrapolated to 156 + codec decode of 10 s",
                       "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "frames": FRAMES},
            "cpu_baseline": best,
            "e2e": {"value": v, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH_PER_GPU)
    ap.add_argument("--small", action="store_true", help="debug shapes (NOT a valid bench number)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
