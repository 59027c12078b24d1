"""Host-side mirror of MOSS-TTS-Realtime's step-wise decode:
`MossTTSRealtimeInference` (reference moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py:35-421) over a
`MossTTSRealtimeModel` that owns the weights (reference modeling_mossttsrealtime.py:80-99: Qwen3 backbone + 4-layer
RoPE local transformer, modeling_mossttsrealtime_local.py:258-448).

Same calls as the reference: `prefill(input_ids, text_prefix_ids, ...)`, `step(text_token, ...)`, `finish(max_steps, ...)`,
`generate_local_transformer(hidden_states, ...)`, `is_finished`, `reset_generation_state(keep_cache)`, same sampling
arguments and defaults, same `ValueError`s.  All arithmetic runs in libmosstts_sm100 (include/mosstts.h, mt_rt_*); the
session state (KV cache, attention mask, history, stopping flags) lives on the device.  No CPU fallback.

Differences that are stated rather than hidden
  * the random draw is the library's inverse CDF over counter-hash uniforms (seed, frame, row, channel), not
    torch.multinomial's RNG stream; sampling needs 1 <= top_k <= 1024 (the reference default is 30); top-k ties: the lower
    index wins (torch.topk's threshold form keeps every tie of the k-th value).
  * `past_key_values` objects are not exchanged: the cache belongs to the engine; `reset_generation_state(keep_cache=True)`
    followed by `prefill` appends the next turn behind it exactly like the reference's mask concatenation (:177-179).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, List, Optional

import numpy as np
import torch

from . import _lib as L
from .config import RealtimeConfig
from .modeling_moss_tts import tile_major


def _rope(theta: float, head_dim: int, n: int):
    half = torch.arange(0, head_dim, 2, dtype=torch.int64).float() / head_dim
    inv_freq = 1.0 / (theta ** half)
    ang = torch.arange(n, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return ang.cos().to(torch.bfloat16), ang.sin().to(torch.bfloat16)


class MossTTSRealtimeModel:
    """Weights of `MossTTSRealtime` (state_dict keys of the reference) packed for the engine."""
    weights_tiled = True

    def __init__(self, config: RealtimeConfig, state_dict: Dict[str, torch.Tensor], device="cuda:0"):
        if not torch.cuda.is_available():
            raise L.MossTTSNativeError("MossTTSRealtimeModel needs a CUDA (sm_100) device; there is no CPU path")
        self.config = config
        self.device = torch.device(device)
        self.dtype = torch.bfloat16
        self._pack(state_dict)
        self._engine = None
        self._cap = None

    def _lin(self, t: torch.Tensor) -> torch.Tensor:
        t = t.to(device=self.device, dtype=torch.bfloat16).contiguous()
        return tile_major(t) if self.weights_tiled else t

    def _pack(self, sd):
        c, dev = self.config, self.device

        def g(name):
            return sd[name].to(device=dev, dtype=torch.bfloat16).contiguous()

        def layer(p, w, sfx):
            w["qkv" + sfx].append(self._lin(torch.cat([sd[p + "self_attn.q_proj.weight"], sd[p + "self_attn.k_proj.weight"],
                                                       sd[p + "self_attn.v_proj.weight"]], dim=0)))
            w["o" + sfx].append(self._lin(sd[p + "self_attn.o_proj.weight"]))
            w["gu" + sfx].append(self._lin(torch.cat([sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]], dim=0)))
            w["down" + sfx].append(self._lin(sd[p + "mlp.down_proj.weight"]))
            w["ln_in" + sfx].append(g(p + "input_layernorm.weight"))
            w["ln_post" + sfx].append(g(p + "post_attention_layernorm.weight"))
            w["qn" + sfx].append(g(p + "self_attn.q_norm.weight"))
            w["kn" + sfx].append(g(p + "self_attn.k_norm.weight"))

        w: Dict[str, object] = {}
        for sfx in ("", "_l"):
            for key in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn"):
                w[key + sfx] = []
        w["embed"] = g("embed_tokens.0.weight")
        w["emb_ext"] = torch.stack([g(f"embed_tokens.{i}.weight") for i in range(1, c.rvq + 1)]).contiguous()
        for l in range(c.num_layers):
            layer(f"language_model.layers.{l}.", w, "")
        w["norm"] = g("language_model.norm.weight")
        for l in range(c.local_num_layers):
            layer(f"local_transformer.model.layers.{l}.", w, "_l")
        w["norm_l"] = g("local_transformer.model.norm.weight")
        if c.rvq > 1:
            w["embed_l"] = torch.stack([g(f"local_transformer.model.embed_tokens.{i}.weight") for i in range(c.rvq - 1)]).contiguous()
        else:
            w["embed_l"] = torch.zeros(1, c.audio_head_size, c.hidden_size, dtype=torch.bfloat16, device=dev)
        w["heads_l"] = [self._lin(sd[f"local_transformer.local_lm_heads.{i}.weight"]) for i in range(c.rvq)]
        self._w = w

    @classmethod
    def random_init(cls, config: RealtimeConfig, device="cuda:0", seed: int = 0, std: float = 0.02):
        """Random weights generated directly on the device (benchmarks: no checkpoint, no network)."""
        self = cls.__new__(cls)
        self.config, self.device, self.dtype = config, torch.device(device), torch.bfloat16
        c = config
        gen = torch.Generator(device=self.device)
        gen.manual_seed(seed)

        def rn(*shape, s=std, mean=0.0):
            return (torch.randn(*shape, device=self.device, dtype=torch.float32, generator=gen) * s + mean).to(torch.bfloat16)

        def lin(n, k, s=std):
            return self._lin(rn(n, k, s=s))

        H = c.hidden_size
        w: Dict[str, object] = {"embed": rn(c.vocab_size, H), "emb_ext": rn(c.rvq, c.audio_head_size, H)}
        shapes = (("", c.num_layers, c.intermediate_size, c.q_size, c.kv_size, c.head_dim),
                  ("_l", c.local_num_layers, c.local_intermediate_size, c.local_q_size, c.local_kv_size, c.local_head_dim))
        for sfx, n_layers, ffn, qs, kvs, D in shapes:
            for key in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn"):
                w[key + sfx] = []
            for _ in range(n_layers):
                w["qkv" + sfx].append(lin(qs + 2 * kvs, H))
                w["o" + sfx].append(lin(H, qs))
                w["gu" + sfx].append(lin(2 * ffn, H))
                w["down" + sfx].append(lin(H, ffn))
                w["ln_in" + sfx].append(rn(H, s=0.05, mean=1.0))
                w["ln_post" + sfx].append(rn(H, s=0.05, mean=1.0))
                w["qn" + sfx].append(rn(D, s=0.05, mean=1.0))
                w["kn" + sfx].append(rn(D, s=0.05, mean=1.0))
        w["norm"] = rn(H, s=0.05, mean=1.0)
        w["norm_l"] = rn(H, s=0.05, mean=1.0)
        w["embed_l"] = rn(max(c.rvq - 1, 1), c.audio_head_size, H, s=1.0)
        w["heads_l"] = [lin(c.audio_head_size, H, s=4 * std) for _ in range(c.rvq)]
        self._w = w
        self._engine, self._cap = None, None
        return self

    def weight_bytes(self) -> int:
        n = 0
        for v in self._w.values():
            n += sum(t.numel() for t in v) * 2 if isinstance(v, list) else v.numel() * 2
        return n

    # ---- engine -------------------------------------------------------------------------------
    def _ensure_engine(self, batch: int, seq: int, tokens: int, frames: int, eos: int, text_pad: int):
        cap = self._cap
        if self._engine is not None and batch <= cap[0] and seq <= cap[1] and tokens <= cap[2] and frames <= cap[3] \
                and (eos, text_pad) == cap[4:]:
            return False
        self._destroy()
        c = self.config
        if cap is not None:
            batch, seq, tokens, frames = max(batch, cap[0]), max(seq, cap[1]), max(tokens, cap[2]), max(frames, cap[3])
        tokens = max(tokens, batch)
        lm = L.LMConfigC(c.hidden_size, c.intermediate_size, c.num_layers, c.num_heads, c.num_kv_heads, c.head_dim,
                         c.vocab_size, c.rvq, c.audio_head_size, c.rms_norm_eps, c.pad_token_id, c.im_start_token_id,
                         c.im_end_token_id, c.audio_start_token_id, c.audio_end_token_id,
                         c.audio_assistant_gen_slot_token_id, c.audio_assistant_delay_slot_token_id, c.audio_pad_code,
                         batch, seq, tokens, int(self.weights_tiled))
        cc = L.RtConfigC(lm, c.local_num_layers, c.local_intermediate_size, c.local_num_heads, c.local_num_kv_heads,
                         c.local_head_dim, c.local_rms_norm_eps, frames, int(eos), int(text_pad))
        kvb, wsb = C.c_longlong(0), C.c_longlong(0)
        L.check(L.lib().mt_rt_query_sizes(C.byref(cc), C.byref(kvb), C.byref(wsb)), "mt_rt_query_sizes")
        self._kv = torch.empty(kvb.value, dtype=torch.uint8, device=self.device)
        self._ws = torch.empty(wsb.value, dtype=torch.uint8, device=self.device)
        cos, sin = _rope(c.rope_theta, c.head_dim, seq)
        lcos, lsin = _rope(c.local_rope_theta, c.local_head_dim, c.rvq)
        self._rope_t = [t.to(self.device).contiguous() for t in (cos, sin, lcos, lsin)]
        w = self._w
        keys = [k + sfx for sfx in ("", "_l") for k in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn")] + ["heads_l"]
        self._ptr_arrays = pa = {k: L.ptr_array(w[k]) for k in keys}
        lw = L.LMWeightsC(w["embed"].data_ptr(), w["emb_ext"].data_ptr(), pa["qkv"], pa["o"], pa["gu"], pa["down"],
                          pa["ln_in"], pa["ln_post"], pa["qn"], pa["kn"], w["norm"].data_ptr(), None, None,
                          self._rope_t[0].data_ptr(), self._rope_t[1].data_ptr())
        wc = L.RtWeightsC(lw, pa["qkv_l"], pa["o_l"], pa["gu_l"], pa["down_l"], pa["ln_in_l"], pa["ln_post_l"], pa["qn_l"],
                          pa["kn_l"], w["norm_l"].data_ptr(), w["embed_l"].data_ptr(), pa["heads_l"],
                          self._rope_t[2].data_ptr(), self._rope_t[3].data_ptr())
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_create(C.byref(cc), C.byref(wc), L.ptr(self._kv), L.ptr(self._ws), C.byref(h)), "mt_rt_create")
        self._engine, self._cap = h, (batch, seq, tokens, frames, int(eos), int(text_pad))
        return True

    def _destroy(self):
        if getattr(self, "_engine", None) is not None:
            L.lib().mt_rt_destroy(self._engine)
            self._engine = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass


class MossTTSRealtimeInference:
    """Step-wise inference wrapper (reference streaming_mossttsrealtime.py:35-357): prefill / step / finish."""

    def __init__(self, model: MossTTSRealtimeModel, tokenizer=None, max_length: int = 1000, channels: int = 16,
                 audio_channel_pad: int = 1024, audio_bos_token: int = 1025, audio_eos_token: int = 1026,
                 text_pad_id: int = 151655, aud_pad_id: int = 151654, max_seq: Optional[int] = None,
                 max_batch: int = 1, seed: int = 0, use_cuda_graph: bool = True):
        if channels != model.config.rvq:
            raise ValueError(f"channels ({channels}) must equal the model's rvq ({model.config.rvq})")
        self.model = model
        self.tokenizer = tokenizer
        self.max_length = max_length
        self.channels = channels
        self.audio_channel_pad = audio_channel_pad
        self.audio_bos_token = audio_bos_token
        self.audio_eos_token = audio_eos_token
        self.text_pad_id = text_pad_id
        self.aud_pad_id = aud_pad_id
        self.seed = seed
        self.use_cuda_graph = use_cuda_graph
        self._max_seq = max_seq
        self._max_batch = max_batch
        self._batch = 0
        self._has_cache = False
        self._last_audio_tokens: Optional[torch.Tensor] = None
        self._generated_tokens: List[torch.Tensor] = []

    # ---- state -----------------------------------------------------------------------------------
    @property
    def device(self):
        return self.model.device

    def _state(self, want_mask=False):
        B = self._batch
        stopping = torch.empty(B, dtype=torch.uint8, device=self.device)
        ln, st, fin = C.c_int(0), C.c_int(0), C.c_int(0)
        mask = None
        cap = self.model._cap[1]
        if want_mask:
            mask = torch.zeros(B, cap, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_state(self.model._engine, L.ptr(stopping), L.ptr(mask), C.c_longlong(cap), C.byref(ln),
                                        C.byref(st), C.byref(fin), L.stream_ptr(self.device)), "mt_rt_state")
        return stopping.bool(), (mask[:, :ln.value].bool() if want_mask else None), ln.value, st.value, bool(fin.value)

    @property
    def is_finished(self) -> bool:
        return self._last_audio_tokens is not None and self._state()[4]

    @property
    def _is_stopping(self):
        return None if self._last_audio_tokens is None else self._state()[0]

    @property
    def attention_mask(self):
        return self._state(True)[1] if self._has_cache else None

    @property
    def _step_idx(self) -> int:
        return 0 if self._last_audio_tokens is None else self._state()[3]

    def reset_generation_state(self, keep_cache: bool = True):
        if not keep_cache:
            self._has_cache = False
        self._generated_tokens = []
        self._last_audio_tokens = None

    # ---- argument normalisation (:88-119) ------------------------------------------------------------
    def _normalize_input_ids(self, input_ids):
        if isinstance(input_ids, torch.Tensor):
            input_ids = input_ids.detach().cpu().numpy()
        if isinstance(input_ids, np.ndarray):
            if input_ids.ndim == 2:
                return [input_ids]
            if input_ids.ndim == 3:
                return [input_ids[i] for i in range(input_ids.shape[0])]
        if isinstance(input_ids, (list, tuple)):
            return [np.array(item) for item in input_ids]
        raise ValueError("input_ids must be a list/array/tensor of shape [T, C] or [B, T, C].")

    def _normalize_text_prefix(self, text_prefix_ids, batch_size: int):
        if text_prefix_ids is None:
            raise ValueError("text_prefix_ids must be provided for prefill.")
        if isinstance(text_prefix_ids, torch.Tensor):
            text_prefix_ids = text_prefix_ids.detach().cpu().tolist()
        if isinstance(text_prefix_ids, np.ndarray):
            text_prefix_ids = text_prefix_ids.tolist()
        if isinstance(text_prefix_ids, list):
            if len(text_prefix_ids) == 0:
                return [[] for _ in range(batch_size)]
            if isinstance(text_prefix_ids[0], (int, np.integer)):
                return [list(text_prefix_ids)]
            if len(text_prefix_ids) == 1 and batch_size > 1:
                return [list(text_prefix_ids[0]) for _ in range(batch_size)]
            if len(text_prefix_ids) != batch_size:
                raise ValueError(f"text_prefix_ids batch size mismatch: got {len(text_prefix_ids)}, expected {batch_size}.")
            return [list(item) for item in text_prefix_ids]
        raise ValueError("text_prefix_ids must be list-like or tensor-like.")

    def _params(self, temperature, top_p, top_k, do_sample, repetition_penalty, repetition_window):
        sampling = bool(do_sample) and temperature != 0
        if sampling:
            if top_k is None:
                top_k = min(1024, self.model.config.audio_head_size)       # no top-k filter == keep the k largest of <= 1027
                if self.model.config.audio_head_size > 1024:
                    raise NotImplementedError("sampling without top_k over more than 1024 codes is not supported")
            if not isinstance(top_k, int) or top_k <= 0:
                raise ValueError(f"`top_k` has to be a strictly positive integer, but is {top_k}")
            if top_p is None:
                top_p = 1.0
            top_p = float(top_p)
            if top_p < 0 or top_p > 1.0:
                raise ValueError(f"`top_p` has to be a float > 0 and < 1, but is {top_p}")
        return L.RtSampleParamsC(int(sampling), float(temperature), float(top_p if top_p is not None else 1.0),
                                 int(top_k if top_k is not None else 0), float(repetition_penalty or 1.0),
                                 int(repetition_window or 0), self.seed & 0xFFFFFFFF)

    # ---- prefill / step / finish --------------------------------------------------------------------------
    def build_prefill_batch(self, input_ids, text_prefix_ids, max_prefill_len: Optional[int] = None):
        """The padded batch prefill() feeds the backbone (:139-176): rows = prompt | text prefix with audio_bos under its
        last token, left-padded with (pad text id, audio_channel_pad); mask False on the padding."""
        input_ids_list = self._normalize_input_ids(input_ids)
        batch_size = len(input_ids_list)
        text_prefix_list = self._normalize_text_prefix(text_prefix_ids, batch_size)
        rows = []
        for i in range(batch_size):
            prefix = text_prefix_list[i]
            if max_prefill_len is not None:
                prefix = prefix[:max_prefill_len]
            if len(prefix) == 0:
                raise ValueError("Prefill requires at least one text token.")
            seg = np.full((len(prefix), self.channels + 1), self.audio_channel_pad, dtype=np.int64)
            seg[:, 0] = np.array(prefix, dtype=np.int64)
            seg[len(prefix) - 1, 1] = self.audio_bos_token
            rows.append(np.concatenate([input_ids_list[i], seg], axis=0))
        max_len = max(r.shape[0] for r in rows)
        pad_id = getattr(self.tokenizer, "pad_token_id", None)
        pad_id = pad_id if pad_id is not None else self.text_pad_id
        ids = np.full((batch_size, max_len, self.channels + 1), self.audio_channel_pad, dtype=np.int64)
        ids[:, :, 0] = pad_id
        mask = np.zeros((batch_size, max_len), dtype=np.uint8)
        for b, r in enumerate(rows):
            ids[b, max_len - r.shape[0]:] = r
            mask[b, max_len - r.shape[0]:] = 1
        return torch.from_numpy(ids), torch.from_numpy(mask)

    @torch.inference_mode()
    def prefill(self, input_ids, text_prefix_ids, max_prefill_len: Optional[int] = None, past_key_values=None, device=None,
                temperature: float = 0.8, top_p: float = 0.6, top_k: int = 30, do_sample: bool = True,
                repetition_penalty: Optional[float] = 1.1, repetition_window: Optional[int] = 50) -> torch.Tensor:
        if past_key_values is not None:
            raise NotImplementedError("the KV cache belongs to the engine; use reset_generation_state(keep_cache=True)")
        ids, mask = self.build_prefill_batch(input_ids, text_prefix_ids, max_prefill_len)
        B, S0, _ = ids.shape
        keep = self._has_cache and B == self._batch
        past = self._state()[2] if keep else 0
        need_seq = past + S0 + (1 if keep else self.max_length + 1)     # a continued turn must fit what was reserved
        seq = max(need_seq, self._max_seq or 0)
        created = self.model._ensure_engine(max(B, self._max_batch), seq, B * S0, self.max_length + 1,
                                            self.audio_eos_token, self.text_pad_id)
        if created and keep:
            raise ValueError("the engine had to grow; the cached conversation is lost — construct the inferencer with a "
                             "larger max_seq / max_length before the first turn")
        p = self._params(temperature, top_p, top_k, do_sample, repetition_penalty, repetition_window)
        ids_d = ids.to(self.device, non_blocking=True).contiguous()
        mask_d = mask.to(self.device, non_blocking=True).contiguous()
        out = torch.empty(B, self.channels, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_prefill(self.model._engine, L.ptr(ids_d), L.ptr(mask_d), B, S0, int(keep), C.byref(p),
                                          L.ptr(out), L.stream_ptr(self.device)), "mt_rt_prefill")
        self._batch, self._has_cache = B, True
        self._generated_tokens = [out]
        self._last_audio_tokens = out
        return out

    @torch.inference_mode()
    def step(self, text_token, temperature: float = 0.8, top_p: float = 0.6, top_k: int = 30, do_sample: bool = True,
             repetition_penalty: Optional[float] = 1.1, repetition_window: Optional[int] = 50) -> torch.Tensor:
        if self._last_audio_tokens is None or not self._has_cache:
            raise ValueError("You must call prefill() before step().")
        B = self._batch
        text_d = None
        if text_token is not None:
            if isinstance(text_token, torch.Tensor):
                text_d = text_token.detach().to(device=self.device, dtype=torch.int64).reshape(-1).contiguous()
            elif isinstance(text_token, (list, tuple, np.ndarray)):
                text_d = torch.as_tensor(list(text_token), dtype=torch.int64).to(self.device)
            else:
                text_d = torch.tensor([int(text_token)], dtype=torch.int64, device=self.device)
            if text_d.numel() != B:
                raise ValueError(f"text_token batch size mismatch: got {text_d.numel()}, expected {B}.")
        p = self._params(temperature, top_p, top_k, do_sample, repetition_penalty, repetition_window)
        out = torch.empty(B, self.channels, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_step(self.model._engine, L.ptr(text_d), C.byref(p), int(self.use_cuda_graph), L.ptr(out),
                                       L.stream_ptr(self.device)), "mt_rt_step")
        self._generated_tokens.append(out)
        self._last_audio_tokens = out
        return out

    @torch.inference_mode()
    def finish(self, max_steps: Optional[int] = None, temperature: float = 0.8, top_p: float = 0.6, top_k: int = 30,
               do_sample: bool = True, repetition_penalty: Optional[float] = 1.1,
               repetition_window: Optional[int] = 50) -> List[torch.Tensor]:
        if self._last_audio_tokens is None or not self._has_cache:
            raise ValueError("You must call prefill() before step().")
        steps = max_steps if max_steps is not None else self.max_length
        _, _, ln, st, fin = self._state()
        steps = min(steps, self.model._cap[3] - st, self.model._cap[1] - ln)
        if steps <= 0 or fin:
            return []
        p = self._params(temperature, top_p, top_k, do_sample, repetition_penalty, repetition_window)
        frames = torch.empty(steps, self._batch, self.channels, dtype=torch.int64, device=self.device)
        n = C.c_int(0)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_finish(self.model._engine, steps, C.byref(p), int(self.use_cuda_graph), L.ptr(frames),
                                         C.byref(n), L.stream_ptr(self.device)), "mt_rt_finish")
        outs = [frames[i] for i in range(n.value)]
        self._generated_tokens.extend(outs)
        if outs:
            self._last_audio_tokens = outs[-1]
        return outs

    def last_hidden_state(self) -> torch.Tensor:
        """Backbone last_hidden_state [B, H] the newest frame was generated from (outputs.last_hidden_state[:, -1])."""
        out = torch.empty(self._batch, self.model.config.hidden_size, dtype=torch.bfloat16, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_last_hidden(self.model._engine, self._batch, L.ptr(out), L.stream_ptr(self.device)),
                    "mt_rt_last_hidden")
        return out

    @torch.inference_mode()
    def generate_local_transformer(self, hidden_states: torch.Tensor, temperature: float = 0.8, top_p: float = 0.6,
                                   top_k: int = 30, do_sample: bool = True, repetition_penalty: Optional[float] = 1.1,
                                   repetition_window: Optional[int] = 50, generated_tokens: Optional[torch.Tensor] = None,
                                   gen_step: int = 0, forced_tokens: Optional[torch.Tensor] = None,
                                   return_logits: bool = False):
        """:301-357.  hidden_states [B,1,H] or [B,H]; generated_tokens [B,T>=gen_step,rvq].  `forced_tokens` /
        `return_logits` (not in the reference) teacher-force the depth loop and return every channel's raw logits
        [rvq, B, V] for parity tests."""
        c = self.model.config
        hs = hidden_states.reshape(-1, c.hidden_size).to(device=self.device, dtype=torch.bfloat16).contiguous()
        B = hs.shape[0]
        if self.model._engine is None:
            self.model._ensure_engine(max(B, self._max_batch), self._max_seq or 64, B, self.max_length + 1,
                                      self.audio_eos_token, self.text_pad_id)
        hist = None
        if generated_tokens is not None and gen_step > 0:
            hist = generated_tokens[:, :gen_step].to(device=self.device, dtype=torch.int64).contiguous()
        forced = None if forced_tokens is None else forced_tokens.to(device=self.device, dtype=torch.int64).contiguous()
        p = self._params(temperature, top_p, top_k, do_sample, repetition_penalty if hist is not None else 1.0, repetition_window)
        out = torch.empty(B, self.channels, dtype=torch.int64, device=self.device)
        logits = torch.empty(self.channels, B, c.audio_head_size, dtype=torch.bfloat16, device=self.device) if return_logits else None
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_rt_local_generate(self.model._engine, L.ptr(hs), B, L.ptr(hist), int(gen_step),
                                                 C.byref(p), L.ptr(forced), L.ptr(logits), L.ptr(out),
                                                 L.stream_ptr(self.device)), "mt_rt_local_generate")
        if forced is not None:
            out = forced
        return (out, logits) if return_logits else out
