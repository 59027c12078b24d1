"""Oracle for MossTTSLocal (TEST INFRASTRUCTURE ONLY) — SURVEY.md §8 row a15.

Restates, in plain torch-CPU, the per-frame algorithm of
  reference moss_tts_local/modeling_moss_tts.py
      CustomMixin._sample :316-477 (global forward, depth loop, EOS handling, output slicing)
      MosiTTSModel._prepare_multi_modal_inputs :515-530
      MossTTSLocalTransformer :178-292, MossTTSAttentionWithoutPositionalEmbedding :126-176
      MossTTSMLP :47-95, MossTTSRMSNorm :34-44, module wiring :573-611
  third-party transformers models/qwen3/modeling_qwen3.py (decoder layer, RMSNorm, MLP, rotary)
  HF logits processors in the order the reference builds them (:359-370):
      RepetitionPenalty (per row, over that row's history of the channel) -> Temperature -> TopK -> TopP

Notes that matter for parity
  * the reference `generate()` itself cannot run under the installed transformers, so the
    golden fixtures are minted by `oracle/make_golden_local.py`, which drives the reference's own
    sub-modules (global model with a DynamicCache, local transformer, MLPs, norms, heads) through the same
    loop; this file is the independent restatement that is checked against them.
  * the local transformer is called with `is_causal=True` and no mask (:161-172): under the README's
    `attn_implementation="sdpa"` that is causal attention, which is what is restated here.  The prefix is
    recomputed every depth step like the reference (`use_cache=False`, :205); a causal stack makes that
    equal to keeping the prefix K/V, which is what the CUDA path does.
  * left-padded prompts: HF's generation loop honours the attention mask it is given (padding keys are masked),
    extends it with ones for every new frame and derives position_ids from its cumulative sum; the restatement
    masks the padding and keeps column-index positions (a per-row constant shift, invisible to rotary attention
    up to the rounding of the cos/sin tables).
  * sampling: greedy channels use no processor at all (:361-362).  Sampled channels apply penalty ->
    temperature, then the same top-k / top-p / inverse-CDF rules as oracle/sampling.py (lower index wins
    top-k ties, where HF's TopKLogitsWarper keeps every tie of the k-th value).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from moss_tts_b200.config import LocalConfig
from . import sampling as S


class OracleLocal:
    def __init__(self, cfg: LocalConfig, weights: Dict[str, torch.Tensor], dtype=torch.float32):
        self.cfg, self.dtype = cfg, dtype
        self.w = {k: v.to(dtype) for k, v in weights.items()}
        half = torch.arange(0, cfg.head_dim, 2, dtype=torch.int64).float() / cfg.head_dim
        self.inv_freq = 1.0 / (cfg.rope_theta ** half)

    # ---- building blocks ------------------------------------------------------------
    def qwen_rmsnorm(self, x, gain):
        """Qwen3RMSNorm (modeling_qwen3.py:59-64): fp32 statistics, cast, then gain."""
        xf = x.float()
        xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.cfg.rms_norm_eps)
        return gain * xf.to(x.dtype)

    @staticmethod
    def moss_rmsnorm(x, gain, eps=1e-6):
        """MossTTSRMSNorm (:34-44): every op in the working dtype."""
        norm = x.pow(2).mean(dim=-1, keepdim=True)
        x = x * torch.rsqrt(norm + eps)
        return x * gain

    def swiglu(self, x, p):
        g = F.linear(x, self.w[p + "gate_proj.weight"])
        u = F.linear(x, self.w[p + "up_proj.weight"])
        return F.linear(F.silu(g) * u, self.w[p + "down_proj.weight"])

    @staticmethod
    def _rot_half(x):
        h = x.shape[-1] // 2
        return torch.cat([-x[..., h:], x[..., :h]], dim=-1)

    def attention(self, p, x, allowed, rope=None, cache=None, l=0):
        """Qwen3Attention with q/k norm; `rope` = (cos, sin) or None (local transformer)."""
        c = self.cfg
        B, T, _ = x.shape
        q = F.linear(x, self.w[p + "q_proj.weight"]).view(B, T, c.num_heads, c.head_dim)
        k = F.linear(x, self.w[p + "k_proj.weight"]).view(B, T, c.num_kv_heads, c.head_dim)
        v = F.linear(x, self.w[p + "v_proj.weight"]).view(B, T, c.num_kv_heads, c.head_dim)
        q = self.qwen_rmsnorm(q, self.w[p + "q_norm.weight"]).transpose(1, 2)
        k = self.qwen_rmsnorm(k, self.w[p + "k_norm.weight"]).transpose(1, 2)
        v = v.transpose(1, 2)
        if rope is not None:
            cs, sn = rope[0][None, None], rope[1][None, None]
            q = (q * cs) + (self._rot_half(q) * sn)
            k = (k * cs) + (self._rot_half(k) * sn)
        if cache is not None:
            if cache[l] is not None:
                k = torch.cat([cache[l][0], k], dim=2)
                v = torch.cat([cache[l][1], v], dim=2)
            cache[l] = (k, v)
        g = c.num_heads // c.num_kv_heads
        kk, vv = k.repeat_interleave(g, dim=1), v.repeat_interleave(g, dim=1)
        scores = (q @ kk.transpose(2, 3)) * (c.head_dim ** -0.5)
        neg = torch.finfo(self.dtype).min
        scores = scores + torch.where(allowed, 0.0, neg).to(self.dtype)
        probs = torch.softmax(scores, dim=-1, dtype=torch.float32).to(q.dtype)
        o = (probs @ vv).transpose(1, 2).reshape(B, T, c.q_size)
        return F.linear(o, self.w[p + "o_proj.weight"])

    def stack(self, prefix, n_layers, x, allowed, rope, cache, final_norm):
        for l in range(n_layers):
            p = f"{prefix}layers.{l}."
            h = self.qwen_rmsnorm(x, self.w[p + "input_layernorm.weight"])
            x = x + self.attention(p + "self_attn.", h, allowed, rope, cache, l)
            h = self.qwen_rmsnorm(x, self.w[p + "post_attention_layernorm.weight"])
            x = x + self.swiglu(h, p + "mlp.")
        return self.qwen_rmsnorm(x, self.w[final_norm])

    # ---- the two transformers ---------------------------------------------------------
    def embed(self, ids, n_vq_inf):
        """:515-530: zeros + embedding_list[0] + ... (left to right, rounding after each add)."""
        x = torch.zeros(*ids.shape[:2], self.cfg.hidden_size, dtype=self.dtype)
        for i in range(min(ids.shape[-1], 1 + n_vq_inf)):
            x = x + self.w[f"model.embedding_list.{i}.weight"][ids[..., i]]
        return x

    def global_forward(self, ids, cache, past_len, n_vq_inf, attention_mask=None):
        """attention_mask [B, past_len + T] bool or None: left padding is masked out of the keys (HF extends the mask it
        is given with ones, GenerationMixin._update_model_kwargs_for_generation).  Positions stay the column index: HF
        derives them from the mask's cumulative sum, a per-row constant shift that rotary attention does not see."""
        B, T, _ = ids.shape
        qpos = torch.arange(T) + past_len
        ang = qpos.float()[:, None] * self.inv_freq[None, :]
        ang = torch.cat([ang, ang], dim=-1)
        rope = (ang.cos().to(self.dtype), ang.sin().to(self.dtype))
        allowed = (torch.arange(past_len + T)[None, :] <= qpos[:, None])[None, None]
        if attention_mask is not None:
            allowed = allowed & attention_mask[:, None, None, :past_len + T].bool()
        x = self.embed(ids, n_vq_inf)
        h = self.stack("model.language_model.", self.cfg.num_layers, x, allowed, rope, cache,
                       "model.language_model.norm.weight")
        return h[:, -1, :]

    def local_forward(self, xs):
        """xs [B, t, LH] -> last position of the causal local transformer, [B, LH]."""
        t = xs.shape[1]
        allowed = (torch.arange(t)[None, :] <= torch.arange(t)[:, None])[None, None]
        h = self.stack("local_transformer.", self.cfg.local_num_layers, xs, allowed, None, None,
                       "local_transformer.norm.weight")
        return h[:, -1, :]

    def channel_logits(self, h_local, c):
        h = self.swiglu(h_local, f"local_to_speech_embedding_mlps.{c}.")
        h = self.moss_rmsnorm(h, self.w[f"layer_norm_before_lm_heads.{c}.weight"])
        lg = F.linear(h, self.w[f"lm_heads.{c}.weight"])
        if c != 0:
            lg[:, self.cfg.audio_pad_code] = float("-inf")
        return lg

    @torch.no_grad()
    def frame_logits(self, input_ids, frame_ids, n_vq_for_inference: Optional[int] = None):
        """One iteration of the :379-450 loop with the chosen tokens replaced by `frame_ids` [B, C]."""
        cfg = self.cfg
        B, _, C = input_ids.shape
        n_inf = cfg.n_vq if n_vq_for_inference is None else n_vq_for_inference
        h_glob = self.global_forward(input_ids, [None] * cfg.num_layers, 0, n_inf)
        cur_in = self.swiglu(h_glob, "speech_embedding_to_local_mlp.")
        local_in = torch.zeros(B, 0, cfg.local_hidden_size, dtype=self.dtype)
        out = []
        for c in range(min(C, 1 + n_inf)):
            local_in = torch.cat([local_in, cur_in[:, None, :]], dim=1)
            out.append(self.channel_logits(self.local_forward(local_in), c))
            cur_in = self.swiglu(self.w[f"model.embedding_list.{c}.weight"][frame_ids[:, c]], "speech_embedding_to_local_mlp.")
        return out

    # ---- generate -----------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, input_ids, max_new_tokens: int, *, n_vq_for_inference: Optional[int] = None,
                 do_samples: Optional[Sequence[bool]] = None, layers: Optional[Sequence[dict]] = None,
                 eos_token_id: Optional[int] = 151653, seed: int = 0, return_trace: bool = False,
                 attention_mask: Optional[torch.Tensor] = None):
        """One frame per iteration (:379-450).  Returns [(start_length, ids[start_idx:])] (:466-475)."""
        cfg = self.cfg
        B, S0, C = input_ids.shape
        n_inf = cfg.n_vq if n_vq_for_inference is None else n_vq_for_inference
        do_samples = [False] * C if do_samples is None else list(do_samples)
        layers = [{} for _ in range(C)] if layers is None else list(layers)
        gen = input_ids.clone()
        unfinished = [True] * B
        cache: List = [None] * cfg.num_layers
        cur = input_ids
        trace = []
        for step in range(max_new_tokens):
            past_len = gen.shape[1] - cur.shape[1]
            am = None
            if attention_mask is not None:
                am = torch.cat([attention_mask.bool(), torch.ones(B, gen.shape[1] - S0, dtype=torch.bool)], dim=1)
            h_glob = self.global_forward(cur, cache, past_len, n_inf, am)
            cur_in = self.swiglu(h_glob, "speech_embedding_to_local_mlp.")
            local_in = torch.zeros(B, 0, cfg.local_hidden_size, dtype=self.dtype)
            nxt = torch.zeros(B, C, dtype=torch.int64)
            step_logits = []
            for c in range(min(C, 1 + n_inf)):
                local_in = torch.cat([local_in, cur_in[:, None, :]], dim=1)
                lg = self.channel_logits(self.local_forward(local_in), c)
                if return_trace:
                    step_logits.append(lg.clone())
                toks = []
                for b in range(B):
                    row = lg[b]
                    if do_samples[c]:
                        lc = layers[c]
                        pen = lc.get("repetition_penalty")
                        if pen is not None and c != 0:
                            row = S.apply_penalty_row(row, np.unique(gen[b, :, c].numpy()), pen)
                        if lc.get("temperature") is not None:
                            row = row / lc["temperature"]
                        tok = S.sample_row(row, top_p=lc.get("top_p"), top_k=lc.get("top_k"), do_sample=True,
                                           u=S.uniform_u(seed, step, b, c))
                    else:
                        tok = int(torch.argmax(row.float()).item())
                    toks.append(tok)
                tk = torch.tensor(toks, dtype=torch.int64)
                nxt[:, c] = tk
                cur_in = self.swiglu(self.w[f"model.embedding_list.{c}.weight"][tk], "speech_embedding_to_local_mlp.")
            if return_trace:
                trace.append(step_logits)
            if eos_token_id is not None:                        # :430-433
                for b in range(B):
                    if not unfinished[b]:
                        nxt[b, 0] = eos_token_id
                        nxt[b, 1:] = cfg.audio_pad_code
            cur = nxt[:, None, :]
            gen = torch.cat([gen, cur], dim=1)
            if eos_token_id is not None:
                for b in range(B):
                    if int(nxt[b, 0]) == eos_token_id:
                        unfinished[b] = False
            if not any(unfinished):
                break
        starts = S.find_last_equal(input_ids[..., 0], cfg.audio_start_token_id)
        out = [(int(S0 - starts[b] - 1), gen[b, int(starts[b]):]) for b in range(B)]
        return (out, trace) if return_trace else out
