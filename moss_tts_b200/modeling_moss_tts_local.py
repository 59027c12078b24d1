"""Host-side mirror of MossTTSLocal's `MossTTSDelayModel` (reference moss_tts_local/modeling_moss_tts.py:573-611,
generate() -> CustomMixin._sample :316-477).

Same call as the reference: `generate(input_ids, attention_mask=None, generation_config=..., **overrides)` with the
reference's generation_config fields (`max_new_tokens`, `n_vq_for_inference`, `do_samples`, `layers`,
`eos_token_id`) and the same return value `[(start_length, ids[start_idx:])]`.  All arithmetic runs in
libmosstts_sm100 (include/mosstts.h, mt_local_*).  No CPU fallback.

Differences that are stated rather than hidden
  * left-padded prompts: the padding is masked out of the attention keys and every generated frame is visible (HF's
    generation loop extends the mask it is given with ones).  Positions are the column index where HF derives them
    from the mask's cumulative sum — a per-row constant shift rotary attention does not see (fp32 goldens with HF's
    position_ids are reproduced exactly, tests/test_oracle_local.py).  Masks with holes other than left padding are
    rejected (`ValueError`).
  * the local transformer is causal (what `attn_implementation="sdpa"` gives, the README's setting).
  * `top_k=None` (no TopKLogitsWarper): audio channels keep their 1024 real codes (the pad code is -inf); the text
    channel, or any `top_k > 1024`, runs the full-vocabulary counting-sort sampler (csrc/sampling_full.cuh).
  * top-k ties: lower index wins (HF keeps every tie of the k-th value).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib as L
from .config import LocalConfig
from .modeling_moss_tts import _rope_tables, tile_major


class MossTTSLocalModel:
    weights_tiled = True

    def __init__(self, config: LocalConfig, state_dict: Dict[str, torch.Tensor], device="cuda:0"):
        if not torch.cuda.is_available():
            raise L.MossTTSNativeError("MossTTSLocalModel needs a CUDA (sm_100) device; there is no CPU path")
        self.config = config
        self.device = torch.device(device)
        self.dtype = torch.bfloat16
        self._pack(state_dict)
        self._engine = None
        self._cap = (0, 0, 0)
        self.last_step_ms: Optional[List[float]] = None
        self.last_n_steps = 0

    # ---- weights ------------------------------------------------------------------------------
    def _lin(self, t: torch.Tensor) -> torch.Tensor:
        t = t.to(device=self.device, dtype=torch.bfloat16).contiguous()
        return tile_major(t) if self.weights_tiled else t

    def _pack(self, sd: Dict[str, torch.Tensor]):
        c, dev = self.config, self.device
        C_ = c.n_vq + 1

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
        w["embed"] = g("model.embedding_list.0.weight")
        w["emb_ext"] = torch.stack([g(f"model.embedding_list.{i}.weight") for i in range(1, C_)]).contiguous()
        for l in range(c.num_layers):
            layer(f"model.language_model.layers.{l}.", w, "")
        w["norm"] = g("model.language_model.norm.weight")
        for l in range(c.local_num_layers):
            layer(f"local_transformer.layers.{l}.", w, "_l")
        w["norm_l"] = g("local_transformer.norm.weight")
        p = "speech_embedding_to_local_mlp."
        w["in_gu"] = self._lin(torch.cat([sd[p + "gate_proj.weight"], sd[p + "up_proj.weight"]], dim=0))
        w["in_down"] = self._lin(sd[p + "down_proj.weight"])
        w["out_gu"], w["out_down"], w["head_norm"], w["heads"] = [], [], [], []
        for ch in range(C_):
            p = f"local_to_speech_embedding_mlps.{ch}."
            w["out_gu"].append(self._lin(torch.cat([sd[p + "gate_proj.weight"], sd[p + "up_proj.weight"]], dim=0)))
            w["out_down"].append(self._lin(sd[p + "down_proj.weight"]))
            w["head_norm"].append(g(f"layer_norm_before_lm_heads.{ch}.weight"))
            w["heads"].append(self._lin(sd[f"lm_heads.{ch}.weight"]))
        self._w = w

    @classmethod
    def random_init(cls, config: LocalConfig, device="cuda:0", seed: int = 0, std: float = 0.02):
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

        H, I, C_ = c.hidden_size, c.intermediate_size, c.n_vq + 1
        LH, LF, AF = c.local_hidden_size, c.local_ffn_hidden_size, c.additional_mlp_ffn_hidden_size
        w: Dict[str, object] = {"embed": rn(c.vocab_size, H), "emb_ext": rn(c.n_vq, c.audio_head_size, H)}
        for sfx, (n_layers, hid, ffn) in (("", (c.num_layers, H, I)), ("_l", (c.local_num_layers, LH, LF))):
            for key in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn"):
                w[key + sfx] = []
            for _ in range(n_layers):
                w["qkv" + sfx].append(lin(c.q_size + 2 * c.kv_size, hid))
                w["o" + sfx].append(lin(hid, c.q_size))
                w["gu" + sfx].append(lin(2 * ffn, hid))
                w["down" + sfx].append(lin(hid, ffn))
                w["ln_in" + sfx].append(rn(hid, s=0.05, mean=1.0))
                w["ln_post" + sfx].append(rn(hid, s=0.05, mean=1.0))
                w["qn" + sfx].append(rn(c.head_dim, s=0.05, mean=1.0))
                w["kn" + sfx].append(rn(c.head_dim, s=0.05, mean=1.0))
        w["norm"] = rn(H, s=0.05, mean=1.0)
        w["norm_l"] = rn(LH, s=0.05, mean=1.0)
        w["in_gu"], w["in_down"] = lin(2 * AF, H), lin(LH, AF)
        w["out_gu"] = [lin(2 * AF, LH) for _ in range(C_)]
        w["out_down"] = [lin(H, AF) for _ in range(C_)]
        w["head_norm"] = [rn(H, s=0.05, mean=1.0) for _ in range(C_)]
        w["heads"] = [lin(c.vocab_size if ch == 0 else c.audio_head_size, H, s=4 * std) for ch in range(C_)]
        self._w = w
        self._engine, self._cap = None, (0, 0, 0)
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
        lm = L.LMConfigC(c.hidden_size, c.intermediate_size, c.num_layers, c.num_heads, c.num_kv_heads, c.head_dim,
                         c.vocab_size, c.n_vq, c.audio_head_size, c.rms_norm_eps, c.pad_token_id, c.im_start_token_id,
                         c.im_end_token_id, c.audio_start_token_id, c.audio_end_token_id,
                         c.audio_assistant_gen_slot_token_id, c.audio_assistant_delay_slot_token_id, c.audio_pad_code,
                         mb, ms, mt_, int(self.weights_tiled))
        cc = L.LocalConfigC(lm, c.local_hidden_size, c.local_ffn_hidden_size, c.local_num_layers,
                            c.additional_mlp_ffn_hidden_size)
        kvb, wsb = C.c_longlong(0), C.c_longlong(0)
        L.check(L.lib().mt_local_query_sizes(C.byref(cc), C.byref(kvb), C.byref(wsb)), "mt_local_query_sizes")
        self._kv = torch.empty(kvb.value, dtype=torch.uint8, device=self.device)
        self._ws = torch.empty(wsb.value, dtype=torch.uint8, device=self.device)
        cos, sin = _rope_tables(c, ms)
        self._rope = (cos.to(self.device).contiguous(), sin.to(self.device).contiguous())
        w = self._w
        keys = [k + sfx for sfx in ("", "_l") for k in ("qkv", "o", "gu", "down", "ln_in", "ln_post", "qn", "kn")]
        keys += ["out_gu", "out_down", "head_norm", "heads"]
        self._ptr_arrays = pa = {k: L.ptr_array(w[k]) for k in keys}
        lw = L.LMWeightsC(w["embed"].data_ptr(), w["emb_ext"].data_ptr(), pa["qkv"], pa["o"], pa["gu"], pa["down"],
                          pa["ln_in"], pa["ln_post"], pa["qn"], pa["kn"], w["norm"].data_ptr(), None, None,
                          self._rope[0].data_ptr(), self._rope[1].data_ptr())
        wc = L.LocalWeightsC(lw, pa["qkv_l"], pa["o_l"], pa["gu_l"], pa["down_l"], pa["ln_in_l"], pa["ln_post_l"],
                             pa["qn_l"], pa["kn_l"], w["norm_l"].data_ptr(), w["in_gu"].data_ptr(),
                             w["in_down"].data_ptr(), pa["out_gu"], pa["out_down"], pa["head_norm"], pa["heads"])
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_local_create(C.byref(cc), C.byref(wc), L.ptr(self._kv), L.ptr(self._ws), C.byref(h)),
                    "mt_local_create")
        self._engine, self._cap = h, (mb, ms, mt_)

    def _destroy(self):
        if getattr(self, "_engine", None) is not None:
            L.lib().mt_local_destroy(self._engine)
            self._engine = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    # ---- reference surface ----------------------------------------------------------------------
    def _mask(self, attention_mask, B, S0):
        if attention_mask is None:
            return None
        m = attention_mask.to(torch.bool)
        if tuple(m.shape) != (B, S0):
            raise ValueError(f"attention_mask must be (batch_size, sequence_length) = ({B}, {S0})")
        if bool(m.all()):
            return None
        if not bool((m[:, 1:] >= m[:, :-1]).all()) or not bool(m[:, -1].all()):
            raise ValueError("attention_mask must be left padding (zeros only before the first real token)")
        return m.to(device=self.device, dtype=torch.uint8).contiguous()

    @torch.inference_mode()
    def frame_logits(self, input_ids: torch.Tensor, frame_ids: torch.Tensor,
                     n_vq_for_inference: Optional[int] = None, attention_mask=None) -> List[torch.Tensor]:
        """Teacher-forced logits of the frame that follows `input_ids` [B,S,1+n_vq] when the depth loop is fed
        `frame_ids` [B,1+n_vq] (one iteration of the reference loop :379-450): list of [B, V_c] bf16 tensors."""
        c = self.config
        C_ = c.n_vq + 1
        if input_ids.dim() != 3 or input_ids.shape[-1] != C_ or tuple(frame_ids.shape) != (input_ids.shape[0], C_):
            raise ValueError("input_ids must be (batch, seq, 1 + n_vq) and frame_ids (batch, 1 + n_vq)")
        n_inf = c.n_vq if n_vq_for_inference is None else int(n_vq_for_inference)
        B, S0, _ = input_ids.shape
        ids = input_ids.to(device=self.device, dtype=torch.int64).contiguous()
        fr = frame_ids.to(device=self.device, dtype=torch.int64).contiguous()
        self._ensure_engine(B, S0, B * S0)
        n_ch = min(C_, 1 + n_inf)
        ldv = max(c.vocab_size, c.audio_head_size)
        out = torch.empty(n_ch, B, ldv, dtype=torch.bfloat16, device=self.device)
        am = self._mask(attention_mask, B, S0)
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_local_forward(self._engine, L.ptr(ids), L.ptr(am), B, S0, n_inf, L.ptr(fr), L.ptr(out),
                                             C.c_longlong(ldv), L.stream_ptr(self.device)), "mt_local_forward")
        return [out[ch, :, :(c.vocab_size if ch == 0 else c.audio_head_size)] for ch in range(n_ch)]

    @torch.inference_mode()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                 generation_config=None, *, max_new_tokens: Optional[int] = None,
                 n_vq_for_inference: Optional[int] = None, do_samples: Optional[Sequence[bool]] = None,
                 layers: Optional[Sequence[dict]] = None, eos_token_id: Optional[int] = "config", seed: int = 0,
                 use_cuda_graph: bool = True, record_step_times: bool = False,
                 check_stop_every: int = 4) -> List[Tuple[int, torch.Tensor]]:
        """`generation_config` may be any object carrying the reference's fields; keyword arguments override it."""
        c = self.config
        C_ = c.n_vq + 1
        if input_ids.dim() != 3 or input_ids.shape[-1] != C_:
            raise ValueError("`Input_ids`'s shape should be exactly (batch_size, sequence_length, 1 + n_vq).")

        def pick(name, given, default):
            if given is not None:
                return given
            v = getattr(generation_config, name, None) if generation_config is not None else None
            return default if v is None else v

        max_new_tokens = int(pick("max_new_tokens", max_new_tokens, 1000))
        n_inf = int(pick("n_vq_for_inference", n_vq_for_inference, c.n_vq))
        do_samples = list(pick("do_samples", do_samples, [False] * C_))
        layers = list(pick("layers", layers, [{} for _ in range(C_)]))
        if eos_token_id == "config":
            eos_token_id = getattr(generation_config, "eos_token_id", None) if generation_config is not None else None
            if eos_token_id is None:
                eos_token_id = c.audio_end_token_id
        if len(do_samples) != C_ or len(layers) != C_:
            raise ValueError(f"do_samples / layers need {C_} entries")
        ds = (C.c_int * C_)(); tk = (C.c_int * C_)()
        tm = (C.c_float * C_)(); tp = (C.c_float * C_)(); rp = (C.c_float * C_)()
        for ch in range(C_):
            lc = layers[ch] or {}
            ds[ch] = int(bool(do_samples[ch]))
            tm[ch] = float(lc["temperature"]) if lc.get("temperature") is not None else 1.0
            tp[ch] = float(lc["top_p"]) if lc.get("top_p") is not None else 1.0
            rp[ch] = float(lc["repetition_penalty"]) if lc.get("repetition_penalty") is not None else 1.0
            if lc.get("top_k") is not None:
                tk[ch] = int(lc["top_k"])
            else:
                tk[ch] = 1024 if ch > 0 and c.audio_head_size <= 1025 else 0
        B, S0, _ = input_ids.shape
        ids = input_ids.to(device=self.device, dtype=torch.int64, non_blocking=True).contiguous()
        max_len = S0 + max_new_tokens
        self._ensure_engine(B, max_len, B * S0)
        gen = torch.empty(B, max_len, C_, dtype=torch.int64, device=self.device)
        am = self._mask(attention_mask, B, S0)
        gp = L.LocalGenParamsC(max_new_tokens, n_inf, -1 if eos_token_id is None else int(eos_token_id), ds, tm, tp, tk, rp,
                               seed & 0xFFFFFFFF, int(use_cuda_graph), check_stop_every, None if am is None else am.data_ptr())
        n_steps = C.c_int(0)
        step_ms = (C.c_float * max(1, max_new_tokens))() if record_step_times else None
        with torch.cuda.device(self.device):
            L.check(L.lib().mt_local_generate(self._engine, L.ptr(ids), B, S0, C.byref(gp), L.ptr(gen), C.byref(n_steps),
                                              step_ms, L.stream_ptr(self.device)), "mt_local_generate")
        self.last_n_steps = n_steps.value
        self.last_step_ms = [step_ms[i] for i in range(n_steps.value)] if record_step_times else None
        total = S0 + n_steps.value
        # :466-475  start index = last audio_start token of the prompt row
        from .inference_utils import find_last_equal_C
        starts = find_last_equal_C(ids[..., 0], c.audio_start_token_id).cpu().tolist()
        return [(S0 - starts[b] - 1, gen[b, starts[b]:total]) for b in range(B)]
