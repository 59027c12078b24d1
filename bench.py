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
# CPU baseline: the reference's own code (oracle/_ref, built by oracle/ref.py) on the host cores;
# the oracle port only when the reference is not in the snapshot
# ------------------------------------------------------------------------------------------------
_REF_MODELS = None


def host_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1: the CPU arm must use the box's cores at every N."""
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(n)
    return torch.get_num_threads()


def reference_models():
    """(reference MossTTSDelayModel at the assumed 8B shape, reference processor over the reference codec), fp32,
    the reference's own random init; built once per process."""
    global _REF_MODELS
    if _REF_MODELS is None:
        from moss_tts_b200.config import LMConfig
        from oracle import ref
        t0 = time.perf_counter()
        torch.manual_seed(0)
        lm = ref.build_lm_bench(LMConfig.delay_8b(), alias_layers=True, never_stop=True)
        codec = ref.build_codec_bench()
        _REF_MODELS = (lm, ref.bare_processor(codec), time.perf_counter() - t0)
    return _REF_MODELS


def reference_utterance(new_tokens: int = NEW_TOKENS):
    """ONE utterance of the bench workload through the reference's own public path, unmodified:
    MossTTSDelayModel.generate (modeling_moss_tts.py:392-525, greedy) -> MossTTSDelayProcessor._parse_audio_codes
    (processing_moss_tts.py:668-709: de-delay, split, codec.decode(chunk_duration=8), trim).  Returns
    (seconds, codec-decoded frames, seconds inside generate)."""
    import contextlib
    import io
    from moss_tts_b200.config import LMConfig
    lm, proc, _ = reference_models()
    ids, mask = make_prompts(LMConfig.delay_8b(), 1, 0)
    t0 = time.perf_counter()
    with contextlib.redirect_stderr(io.StringIO()):              # the reference prints a tqdm bar per step
        out = lm.generate(ids, mask, max_new_tokens=new_tokens, text_temperature=0, audio_temperature=0)
    t1 = time.perf_counter()
    samples = 0
    for start_length, g in out:
        for w in proc._parse_audio_codes(int(start_length), g[:, 1:]):
            samples += w.numel()
    t2 = time.perf_counter()
    # the reference trims start_length/first_codes_length of the first segment (1 frame here): count decoded frames
    frames = new_tokens + 1 - 32 if new_tokens + 1 > 32 else 0
    return t2 - t0, frames, t1 - t0, samples


def cpu_baseline():
    """The bench workload for ONE utterance (B = 1 of the 64) at the full shapes on the host cores."""
    threads = host_threads()
    from oracle import ref
    if ref.available():
        _, _, t_build = reference_models()
        secs, frames, t_gen, samples = reference_utterance()
        return {"value": frames / FRAME_RATE / secs, "unit": "audio-s/s", "cores": threads, "kind": "reference",
                "sample": f"B=1 (one of the {BATCH_PER_GPU} utterances), the reference's own generate() + processor decode, fp32, "
                          f"eager attention, assumed 8B shape with layers 1..35 aliasing layer 0's weights (772 MB/layer, DRAM-streamed), "
                          f"released codec config.json: {NEW_TOKENS} steps in {t_gen:.1f}s + codec {secs - t_gen:.2f}s = {secs:.1f}s "
                          f"for {frames} frames ({samples} samples); model build {t_build:.0f}s not timed",
                "seconds": secs, "generate_s": t_gen, "frames": frames}
    return cpu_baseline_port()


def cpu_baseline_port(decode_steps: int = 3):
    """Fallback when oracle/_ref is not in the snapshot: the oracle PORT (kind "port"), B=1 at the full shapes:
    prefill(96) + `decode_steps` decode steps + a 10 s codec decode, extrapolated to 156 steps."""
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
    Returns (avg ms per launch, algorithmic bytes per launch, splits, avg ms incl. the stand-alone reduce kernel,
    bytes of fp32 split-K slabs the launch writes)."""
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
    alg = N * H * 2 + batch * H * 2 + batch * N * 2      # weights + x + the bf16 output, nothing else
    slab = splits.value * batch * N * 4                   # implementation traffic (fp32 split-K slabs), reported apart
    return ms, alg, splits.value, ms_full, slab


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel's launch, from the committed `ncu --set full`
    capture of this round (profiles/ncu_traffic.json, written by scripts/gpu_job_profile.sh); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return float(json.load(f)["dram_bytes_per_launch"])
    except Exception:
        return None


def refuse_mt_env():
    """The product library ignores the environment (A/B switches exist only in an EXPERIMENTS=1 build); a bench run
    with any MT_* variable set is refused anyway, so that no number can come from a non-default configuration."""
    bad = sorted(k for k in os.environ if k.startswith("MT_"))
    if bad:
        print(json.dumps({"metric": METRIC, "value": None, "invalid": "MT_* environment variables are set: " + ", ".join(bad)}), flush=True)
        sys.exit(2)


def run_ours(args):
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.config import LMConfig, CodecConfig
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    refuse_mt_env()
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
    wav_all = torch.empty(B * world, FRAMES * ccfg.downsample_rate, dtype=torch.float32).pin_memory() if (world > 1 and rank == 0) else None
    if world > 1:
        from moss_tts_b200.sharding import gather_variable

    def step_device():
        out = model.generate(ids_d, mask_d, **gen_kw)
        return proc.decode_device(out)

    gathered_bytes = [0]

    def step_e2e():
        out = model.generate(ids_h, mask_h, **gen_kw)          # H2D of the prompt inside
        wavs = proc.decode_device(out)
        n = 0
        if dist is not None:
            # north_star: "NCCL used only to gather outputs": all_gather of the lengths, then ncclGather of the padded
            # waveforms to rank 0 (moss_tts_b200/sharding.py), device tensors over NVLink, then ONE D2H on rank 0
            rows = gather_variable([ws[0] for ws in wavs], device=dev)
            n = sum(ws[0].numel() for ws in wavs)
            if rank == 0:
                for b, w in enumerate(rows):
                    wav_all[b, : w.numel()].copy_(w, non_blocking=True)
                gathered_bytes[0] = sum(w.numel() for w in rows) * 4
        else:
            for b, ws in enumerate(wavs):
                w = ws[0]
                wav_h[b, : w.numel()].copy_(w, non_blocking=True)   # D2H of the result
                n += w.numel()
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
    # codec-decoded frames: the waveform of a row is its 125 frames minus the processor's start_length trim of 1/125
    samples = sum(w[0].numel() for w in wavs)
    frames = B * FRAMES
    assert len(wavs) == B and all(len(w) == 1 for w in wavs), "every utterance must decode to exactly one audio segment"
    assert abs(samples - B * (FRAMES - 1) * ccfg.downsample_rate) <= B, \
        f"decoded {samples} samples, expected {B} x {FRAMES} frames (minus the 1-frame start_length trim)"
    assert model.last_n_steps == NEW_TOKENS, f"generate ran {model.last_n_steps} steps, expected {NEW_TOKENS}"
    audio_s_per_step = frames / FRAME_RATE                  # codec-decoded frames / 12.5
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
    k_ms, k_bytes, k_splits, k_ms_full, k_slab = time_dominant_kernel(model, B)
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    # bytes a decode step of THIS run streams from HBM: all layer weights + the audio heads + the text-head row tile(s)
    # that hold the gen/delay-slot rows (every row of the bench batch is inside an audio span at every step, so the
    # text-head filter is on throughout and the other 1.24 GB of the text head are never read)
    H = cfg.hidden_size
    lo, hi = sorted((cfg.audio_assistant_gen_slot_token_id, cfg.audio_assistant_delay_slot_token_id))
    text_tiles = hi // 128 - lo // 128 + 1
    layer_bytes = sum(t.numel() for k in ("qkv", "o", "gu", "down") for t in model._w[k]) * 2
    streamed_bytes = layer_bytes + model._w["audio_heads"].numel() * 2 + text_tiles * 128 * H * 2
    full_head_bytes = layer_bytes + (model._w["text_head"].numel() + model._w["audio_heads"].numel()) * 2
    kv_bytes_mid = B * (PROMPT_LEN + NEW_TOKENS // 2) * cfg.num_layers * 2 * cfg.kv_size * 2
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
                            "projections; waveform SNR vs the reference decoder at the full 32+12+12+12-layer depth: 49.1 dB (fp16 operands) / 48.8 dB (TF32), "
                            "unchunked and chunk_duration=8 (tests/test_gpu_baseline_shapes.py, golden codec_full.npz; gate: >= 40 dB)",
                   "batch_per_gpu": B, "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "frames": FRAMES,
                   "sharding": f"request-sharded x{world}, no in-step collective",
                   "l2_policy": "inputs larger than L2: %.2f GB of weights streamed every decode step (text head filtered to %d "
                                "row tile; %.2f GB with the full text head)" % (streamed_bytes / 1e9, text_tiles, full_head_bytes / 1e9),
                   "nccl": ("all_gather of lengths + ncclGather of padded fp32 waveforms to rank 0 inside the e2e step: %d bytes per step"
                            % gathered_bytes[0]) if world > 1 else "none (N=1)"},
        "e2e": {"value": world * audio_s_per_step * args.steps / (ms_e2e * 1e-3), "unit": "audio-s/s",
                "h2d_bytes_per_step": ids_h.numel() * 8 + mask_h.numel(), "d2h_bytes_per_step": int(gathered_bytes[0]) if world > 1 else int(n_samples) * 4,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "decode_ms_per_step": dec_med, "prefill_ms": prefill_ms, "codec_ms": codec_ms,
        "decode_weight_stream_gbs": streamed_bytes / (dec_med * 1e-3) / 1e9 if dec_med == dec_med else None,
        "roofline_step": {"bound": "hbm", "streamed_bytes": streamed_bytes, "kv_bytes_mid_generation": kv_bytes_mid,
                          "achieved": (streamed_bytes + kv_bytes_mid) / (dec_med * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                          "frac": (streamed_bytes + kv_bytes_mid) / (dec_med * 1e-3) / 1e9 / hbm,
                          "what": "whole decode step in situ: weight + KV bytes actually streamed / median decode-step time"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                     "traffic": ncu_traffic(), "kernel": "gemm_dec_kernel<BN=64,bf16> gate|up decode GEMM (M=%d,N=%d,K=%d) through mt_gemm as the engine "
                               "launches it: %d split-K fp32 partial slabs left for the fused consumer" % (B, 2 * cfg.intermediate_size, cfg.hidden_size, k_splits),
                     "ms_per_launch": k_ms, "algorithmic_bytes": k_bytes, "algorithmic_bytes_def": "weights N*K*2 + x M*K*2 + bf16 output M*N*2",
                     "splitk_slab_bytes_written": k_slab, "peak_source": src,
                     "ms_per_call_with_own_reduce_kernel": k_ms_full},
        "clocks": clocks,
    }
    if not args.no_cpu and world == 1:          # cpu_baseline: rank 0 at N = 1 only
        try:
            line["cpu_baseline"] = cpu_baseline() if not args.small else {"value": None, "kind": "port", "sample": "skipped (--small)"}
        except Exception as e:           # the baseline must not take the product number down with it
            line["cpu_baseline"] = {"value": None, "kind": "reference", "error": repr(e)[:300]}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_reference(args):
    """Reference arm: the UNMODIFIED reference (oracle/_ref) on the host cores, all host threads, through its own public
    path.  A "step" is a bounded sample of the workload: ONE of the 64 utterances, start to finish (96-token prompt, 156
    greedy steps, processor decode to 24 kHz) -- timed as it ran, no extrapolation.  The step count is clamped so the run
    ends within a few minutes.  Falls back to the oracle port (kind "port") only when oracle/_ref is not in the snapshot."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    threads = host_threads()
    from oracle import ref
    if not ref.available():
        best = cpu_baseline_port(decode_steps=2)
        v, ms, steps = best["value"], FRAMES / FRAME_RATE / best["value"] * 1e3, 1
        sample = best["sample"]
    else:
        _, _, t_build = reference_models()
        reference_utterance(new_tokens=36)                             # warm-up: allocator, mkldnn primitives
        steps = max(1, min(args.steps, 2))
        runs = [reference_utterance()]
        while len(runs) < steps and sum(r[0] for r in runs) < 60.0:      # keep the whole arm within a few minutes
            runs.append(reference_utterance())
        steps = len(runs)
        secs = sum(r[0] for r in runs)
        frames = sum(r[1] for r in runs)
        v, ms = frames / FRAME_RATE / secs, secs / steps * 1e3
        best = {"value": v, "unit": "audio-s/s", "cores": threads, "kind": "reference",
                "sample": f"{steps} x one utterance (B=1) through the reference's generate() + processor decode at the full shapes",
                "generate_s": [r[2] for r in runs], "seconds": [r[0] for r in runs], "model_build_s_not_timed": t_build}
        sample = ("bounded sample: B=1 (one of the 64 utterances of a step), fp32 (the reference's CPU dtype), eager attention; "
                  "assumed 8B shape with layers 1..35 aliasing layer 0's parameter storage (772 MB per layer, far beyond the LLC), "
                  "released codec config.json with the reference's own random init")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "audio-s/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": 1, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[1]: MossTTSDelay batch-64 x 10 s utterances bf16 (greedy, 96-token prompt, 156 steps) "
                                   "+ MOSS-Audio-Tokenizer decode to 24 kHz",
                       "sample": sample, "batch_per_step": 1,
                       "prompt_len": PROMPT_LEN, "new_tokens": NEW_TOKENS, "frames": FRAMES, "host_threads": threads},
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
