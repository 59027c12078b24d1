import sys, time, torch
sys.path.insert(0, '/root/repo')
import bench
from moss_tts_b200.config import LMConfig, CodecConfig
from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor
dev = torch.device('cuda', 0)
cfg = LMConfig.delay_8b(); ccfg = CodecConfig.released()
model = MossTTSDelayModel.random_init(cfg, device=dev, seed=1234, never_stop=True)
codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=24 << 30)
proc = MossTTSDelayProcessor(cfg, audio_tokenizer=codec)
ids, mask = bench.make_prompts(cfg, 64, seed=0); ids, mask = ids.to(dev), mask.to(dev)
kw = dict(max_new_tokens=bench.NEW_TOKENS, text_temperature=0, audio_temperature=0)
for _ in range(3):
    out = model.generate(ids, mask, **kw); proc.decode_device(out)
torch.cuda.synchronize()
for it in range(3):
    t0 = time.perf_counter(); out = model.generate(ids, mask, **kw); torch.cuda.synchronize(); t1 = time.perf_counter()
    w = proc.decode_device(out); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"generate {1e3*(t1-t0):.1f} ms  decode_device {1e3*(t2-t1):.1f} ms")
out = model.generate(ids, mask, record_step_times=True, **kw)
ms = model.last_step_ms
print("sum of step events", sum(ms), "prefill", ms[0], "step1 (eager)", ms[1], "step2 (capture)", ms[2], "median", sorted(ms[3:])[len(ms[3:])//2])
import cProfile, pstats
pr = cProfile.Profile(); pr.enable(); w = proc.decode_device(out); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(12)
