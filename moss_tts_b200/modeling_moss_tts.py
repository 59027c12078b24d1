"""Host-side mirror of `MossTTSDelayModel` (reference moss_tts_delay/modeling_moss_tts.py:159-525).

Same method names and argument meaning as the reference (`generate`, `forward`,
`get_input_embeddings`), but the tensors are owned here and all arithmetic runs in
libmosstts_sm100 (hand-written sm_100a CUDA) through the C ABI of include/mosstts.h.
No CPU fallback: constructing the model without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib as L
from .config import LMConfig


@dataclass
class MossTTSDelayOutputWithPast:
    """Subset of the reference output class (modeling_moss_tts.py:52-86) that inference fills."""
    logits: Optional[List[torch.Tensor]] = None
    past_key_values: Optional[object] = None
    loss: Optional[torch.Tensor] = None


def _rope_tables(cfg: LMConfig, max_seq: int):
    """cos/sin exactly as Qwen3RotaryEmbedding computes them (modeling_qwen3.py:128-158): fp32 on the
    host, then one cast to bf16.  Only head_dim/2 distinct columns exist (emb = cat(freqs, freqs))."""
    half = torch.arange(0, cfg.head_dim, 2, dtype=torch.int64).float() / cfg.head_dim
    inv_freq = 1.0 / (cfg.rope_theta ** half)
    ang = torch.arange(max_seq, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return ang.cos().to(torch.bfloat16), ang.sin().to(torch.bfloat16)


def tile_major(w: torch.Tensor) -> torch.Tensor:
    """[N, K] row-major Linear weight -> the tile-major HBM layout the GEMM kernels stream
    ([ceil(N/128)][K/ke][128][ke], ke = 128 bytes of K; include/mosstts.h MT_GEMM_W_TILED): every TMA box
    becomes one contiguous 16 KB read.  Rows past N are zero padding.  Returned flat."""
    N, K = w.shape
    ke = 128 // w.element_size()
    if K % ke:
        raise ValueError(f"tile-major weights need K % {ke} == 0, got {K}")
    Np = (N + 127) // 128 * 128
    if Np != N:
        w = torch.cat([w, w.new_zeros(Np - N, K)], dim=0)
    return w.view(Np // 128, 128, K // ke, ke).permute(0, 2, 1, 3).contiguous().view(-1)


_TILED_KEYS = ("qkv", "o", "gu", "down", "text_head", "audio_heads")


class MossTTSDelayModel:
    weights_tiled = True    # HBM layout of the Linear weights (False = plain nn.Linear row-major)

    def __init__(self, config: LMConfig, state_dict: Dict[str, torch.Tensor], device="cuda:0"):
        if not torch.cuda.is_available():
            raise L.MossTTSNativeError("MossTTSDelayModel needs a CUDA (sm_100) device; there is no CPU path")
        self.config = config
        self.device = torch.device(device)
        self.dtype = torch.bfloat16
        self._pack(state_dict)
        self._engine = None
        self._cap = (0, 0, 0)
        self._bufs = {}
        self.last_step_ms: Optional[List[float]] = None
        self.last_n_steps = 0

    # ---- weights ------------------------------------------------------------------------------
    def _pack(self, sd: Dict[str, torch.Tensor]):
        c, dev = self.config, self.device

        def g(name):
            return sd[name].to(device=dev, dtype=torch.bfloat16).contiguous()

        w = {}
        w["embed"] = g("language_model.embed_tokens.weight")
        w["emb_ext"] = torch.stack([g(f"emb_ext.{i}.weight") for i in range(c.n_vq)]).contiguous()
        for key in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn"):
            w[key] = []
        for l in range(c.num_layers):
            p = f"language_model.layers.{l}."
            w["qkv"].append(torch.cat([g(p + "self_attn.q_proj.weight"), g(p + "self_attn.k_proj.weight"),
                                       g(p + "self_attn.v_proj.weight")], dim=0).contiguous())
            w["o"].append(g(p + "self_attn.o_proj.weight"))
            w["gu"].append(torch.cat([g(p + "mlp.gate_proj.weight"), g(p + "mlp.up_proj.weight")], dim=0).contiguous())
            w["down"].append(g(p + "mlp.down_proj.weight"))
            w["ln_in"].append(g(p + "input_layernorm.weight"))
            w["ln_post"].append(g(p + "post_attention_layernorm.weight"))
            w["qn"].append(g(p + "self_attn.q_norm.weight"))
            w["kn"].append(g(p + "self_attn.k_norm.weight"))
        w["norm"] = g("language_model.norm.weight")
        w["text_head"] = g("lm_heads.0.weight")
        w["audio_heads"] = torch.cat([g(f"lm_heads.{i + 1}.weight") for i in range(c.n_vq)], dim=0).contiguous()
        self._w = w
        self._retile()

    def _retile(self):
        """Record logical shapes and convert every Linear weight to the tile-major layout."""
        w = self._w
        self._shapes = {k: (tuple(w[k][0].shape) if isinstance(w[k], list) else tuple(w[k].shape)) for k in _TILED_KEYS}
        if not self.weights_tiled:
            return
        for k in _TILED_KEYS:
            if isinstance(w[k], list):
                for i in range(len(w[k])):
                    w[k][i] = tile_major(w[k][i])
            else:
                w[k] = tile_major(w[k])

    @classmethod
    def random_init(cls, config: LMConfig, device="cuda:0", seed: int = 0, never_stop: bool = True, std: float = 0.02):
        """Random weights generated directly on the device (bench: no checkpoint, no network)."""
        self = cls.__new__(cls)
        self.config, self.device, self.dtype = config, torch.device(device), torch.bfloat16
        c = config
        gen = torch.Generator(device=self.device)
        gen.manual_seed(seed)

        def rn(*shape, s=std, mean=0.0):
            return (torch.randn(*shape, device=self.device, dtype=torch.float32, generator=gen) * s + mean).to(torch.bfloat16)

        H, I = c.hidden_size, c.intermediate_size
        w = {"embed": rn(c.vocab_size, H), "emb_ext": rn(c.n_vq, c.audio_head_size, H)}
        for key in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn"):
            w[key] = []
        for _ in range(c.num_layers):
            w["qkv"].append(rn(c.q_size + 2 * c.kv_size, H))
            w["o"].append(rn(H, c.q_size))
            w["gu"].append(rn(2 * I, H))
            w["down"].append(rn(H, I))
            w["ln_in"].append(rn(H, s=0.05, mean=1.0))
            w["ln_post"].append(rn(H, s=0.05, mean=1.0))
            w["qn"].append(rn(c.head_dim, s=0.05, mean=1.0))
            w["kn"].append(rn(c.head_dim, s=0.05, mean=1.0))
        w["norm"] = rn(H, s=0.05, mean=1.0)
        w["text_head"] = rn(c.vocab_size, H, s=4 * std)
        w["audio_heads"] = rn(c.n_vq * c.audio_head_size, H, s=4 * std)
        if never_stop:   # exact tie -> argmax keeps choosing the gen slot
            w["text_head"][c.audio_assistant_delay_slot_token_id] = w["text_head"][c.audio_assistant_gen_slot_token_id]
        self._w = w
        self._retile()
        self._engine, self._cap, self._bufs = None, (0, 0, 0), {}
        self.last_step_ms, self.last_n_steps = None, 0
        return self

    def weight_bytes(self) -> int:
        n = 0
        for v in self._w.values():
            n += sum(t.numel() for t in v) * 2 if isinstance(v, list) else v.numel() * 2
        return n

    # ---- engine -------------------------------------------------------------------------------
    def _ensure_engine(self, batch: int, seq: int, tokens: int):
        mb, ms, mt_ = self._cap
        if self._engine is not None and batch <= mb and seq <= ms and tokens <= mt_:
            return
        self._destroy()
        c = self.config
        mb, ms, mt_ = max(mb, batch), max(ms, seq), max(mt_, tokens, batch)
        cc = L.LMConfigC(c.hidden_size, c.intermediate_size, c.num_layers, c.num_heads, c.num_kv_heads, c.head_dim,
                         c.vocab_size, c.n_vq, c.audio_head_size, c.rms_norm_eps, c.pad_token_id, c.im_start_token_id,
                         c.im_end_token_id, c.audio_start_token_id, c.audio_end_token_id,
                         c.audio_assistant_gen_slot_token_id, c.audio_assistant_delay_slot_token_id, c.audio_pad_code,
                         mb, ms, mt_, int(self.weights_tiled))
        kvb, wsb = C.c_longlong(0), C.c_longlong(0)
        L.check(L.lib().mt_lm_query_sizes(C.byref(cc), C.byref(kvb), C.byref(wsb)), "mt_lm_query_sizes")
        self._kv = torch.empty(kvb.value, dtype=torch.uint8, device=self.device)
        self._ws = torch.empty(wsb.value, dtype=torch.uint8, device=self.device)
        cos, sin = _rope_tables(c, ms)
        self._rope = (cos.to(self.device).contiguous(), sin.to(self.device).contiguous())
        w = self._w
        self._ptr_arrays = {k: L.ptr_array(w[k]) for k in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn")}
        pa = self._ptr_arrays
        wc = L.LMWeightsC(w["embed"].data_ptr(), w["emb_ext"].data_ptr(), pa["qkv"], pa["o"], pa["gu"], pa["down"],
                          pa["ln_in"], pa["ln_post"], pa["qn"], pa["kn"], w["norm"].data_ptr(),
                          w["text_head"].data_ptr(), w["audio_heads"].data_ptr(), self._rope[0].data_ptr(),
                          self._rope[1].data_ptr())
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_lm_create(C.byref(cc), C.byref(wc), L.ptr(self._kv), L.ptr(self._ws), C.byref(h)),
                    "mt_lm_create")
        self._engine, self._cap = h, (mb, ms, mt_)

    def _destroy(self):
        if getattr(self, "_engine", None) is not None:
            L.lib().mt_lm_destroy(self._engine)
            self._engine = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    # ---- reference surface ----------------------------------------------------------------------
    @torch.inference_mode()
    def get_input_embeddings(self, input_ids: torch.Tensor) -> torch.Tensor:
        """[B, T, 1+n_vq] ids -> [B, T, H] bf16 summed codebook embedding (reference :196-213; embed_sum_kernel)."""
        c = self.config
        if input_ids.dim() != 3 or input_ids.shape[-1] != c.n_vq + 1:
            raise ValueError("`Input_ids`'s shape should be exactly (batch_size, sequence_length, 1 + n_vq).")
        B, T, _ = input_ids.shape
        ids = input_ids.to(device=self.device, dtype=torch.int64).contiguous()
        out = torch.empty(B, T, c.hidden_size, dtype=torch.bfloat16, device=self.device)
        if B * T == 0:
            return out
        self._ensure_engine(1, 1, 1)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_lm_embed(self._engine, L.ptr(ids), C.c_longlong(B * T), L.ptr(out), L.stream_ptr(self.device)),
                    "mt_lm_embed")
        return out

    @torch.inference_mode()
    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, **unused):
        """Teacher-forced logits for all positions: list of 1+n_vq tensors [B,S,V_i] (bf16)."""
        c = self.config
        if input_ids.dim() != 3 or input_ids.shape[-1] != c.n_vq + 1:
            raise ValueError("`Input_ids`'s shape should be exactly (batch_size, sequence_length, 1 + n_vq).")
        B, S, _ = input_ids.shape
        ids = input_ids.to(device=self.device, dtype=torch.int64).contiguous()
        if attention_mask is None:
            attention_mask = torch.ones(B, S, dtype=torch.bool)
        mask = attention_mask.to(device=self.device).to(torch.uint8).contiguous()
        self._ensure_engine(B, S, B * S)
        text = torch.empty(B, S, c.vocab_size, dtype=torch.bfloat16, device=self.device)
        audio = torch.empty(B, S, c.n_vq, c.audio_head_size, dtype=torch.bfloat16, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_lm_forward(self._engine, L.ptr(ids), L.ptr(mask), C.c_longlong(S), B, S, 0, 0,
                                          L.ptr(text), L.ptr(audio), L.stream_ptr(self.device)), "mt_lm_forward")
        logits = [text] + [audio[:, :, i, :] for i in range(c.n_vq)]
        return MossTTSDelayOutputWithPast(logits=logits)

    __call__ = forward

    @torch.inference_mode()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                 max_new_tokens: int = 1000, text_temperature: float = 1.5, text_top_p: float = 1.0,
                 text_top_k: int = 50, audio_temperature: float = 1.7, audio_top_p: float = 0.8,
                 audio_top_k: int = 25, audio_repetition_penalty: float = 1.0, *, seed: int = 0,
                 use_cuda_graph: bool = True, record_step_times: bool = False,
                 check_stop_every: int = 8) -> List[Tuple[int, torch.Tensor]]:
        """Same contract as the reference generate(): returns [(start_length, ids[T_i, 1+n_vq])]."""
        c = self.config
        if input_ids.dim() != 3 or input_ids.shape[-1] != c.n_vq + 1:
            raise ValueError("`Input_ids`'s shape should be exactly (batch_size, sequence_length, 1 + n_vq).")
        B, S0, Cc = input_ids.shape
        ids = input_ids.to(device=self.device, dtype=torch.int64, non_blocking=True).contiguous()
        if attention_mask is None:
            attention_mask = torch.ones(B, S0, dtype=torch.bool)
        mask = attention_mask.to(device=self.device, non_blocking=True).to(torch.uint8).contiguous()
        max_len = S0 + max_new_tokens
        self._ensure_engine(B, max_len, B * S0)
        gen = torch.empty(B, max_len, Cc, dtype=torch.int64, device=self.device)
        starts = torch.empty(B, dtype=torch.int64, device=self.device)
        gp = L.GenParamsC(max_new_tokens, float(text_temperature), float(text_top_p), int(text_top_k or 0),
                          float(audio_temperature), float(audio_top_p), int(audio_top_k or 0),
                          float(audio_repetition_penalty), seed & 0xFFFFFFFF, int(use_cuda_graph), check_stop_every)
        n_steps = C.c_int(0)
        step_ms = (C.c_float * max(1, max_new_tokens))() if record_step_times else None
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_lm_generate(self._engine, L.ptr(ids), L.ptr(mask), B, S0, C.byref(gp), L.ptr(gen),
                                           L.ptr(starts), C.byref(n_steps), step_ms, L.stream_ptr(self.device)),
                    "mt_lm_generate")
        self.last_n_steps = n_steps.value
        self.last_step_ms = [step_ms[i] for i in range(n_steps.value)] if record_step_times else None
        total = S0 + n_steps.value
        starts_h = starts.cpu().tolist()
        out = []
        for b in range(B):
            out.append((S0 - starts_h[b], gen[b, starts_h[b]:total]))
        return out
