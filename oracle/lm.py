"""Oracle for the MossTTSDelay language model (TEST INFRASTRUCTURE ONLY).

Restates, in plain torch-CPU with explicit per-row Python state, the algorithm of
  reference moss_tts_delay/modeling_moss_tts.py
      get_input_embeddings :196-213     forward :225-300     generate :392-525
  third-party transformers (pinned 5.0.0; 5.5.0 installed) models/qwen3/modeling_qwen3.py
      Qwen3RMSNorm :59-64   Qwen3MLP :81-83   rotary :137-181
      eager_attention_forward :196-217   Qwen3Attention :252-291
      Qwen3DecoderLayer :305-335   Qwen3Model.forward :378-439
  cache_utils.DynamicLayer.update :102-121 (append along the time axis)

Every op is written in the working dtype (`torch.float32` or `torch.bfloat16`)
so that bf16 rounding happens where the reference's PyTorch code rounds:
after every Linear, after every elementwise op, softmax in fp32 then cast.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from moss_tts_b200.config import LMConfig
from . import sampling as S

INT64_MAX = np.iinfo(np.int64).max


class OracleLM:
    def __init__(self, cfg: LMConfig, weights: Dict[str, torch.Tensor], dtype=torch.float32):
        self.cfg = cfg
        self.dtype = dtype
        self.w = {k: v.to(dtype) for k, v in weights.items()}
        half = torch.arange(0, cfg.head_dim, 2, dtype=torch.int64).float() / cfg.head_dim
        self.inv_freq = 1.0 / (cfg.rope_theta ** half)            # modeling_qwen3.py:128-131

    # ---- pieces ---------------------------------------------------------------------
    def embed(self, ids: torch.Tensor) -> torch.Tensor:
        """E_text[id0] + sum_i emb_ext[i][id_{i+1}], left-to-right, rounding after each add."""
        x = self.w["language_model.embed_tokens.weight"][ids[..., 0]]
        for i in range(self.cfg.n_vq):
            x = x + self.w[f"emb_ext.{i}.weight"][ids[..., i + 1]]
        return x

    def rmsnorm(self, x, gain, eps=None):
        eps = self.cfg.rms_norm_eps if eps is None else eps
        xf = x.float()
        xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
        return gain * xf.to(x.dtype)

    def rope_tables(self, positions: torch.Tensor):
        ang = positions.float()[:, None] * self.inv_freq[None, :]
        ang = torch.cat([ang, ang], dim=-1)
        return ang.cos().to(self.dtype), ang.sin().to(self.dtype)   # [T, head_dim]

    @staticmethod
    def _rot_half(x):
        h = x.shape[-1] // 2
        return torch.cat([-x[..., h:], x[..., :h]], dim=-1)

    def attention_layer(self, l, x, cos, sin, add_mask, cache):
        c, p = self.cfg, f"language_model.layers.{l}.self_attn."
        B, T, _ = x.shape
        q = F.linear(x, self.w[p + "q_proj.weight"]).view(B, T, c.num_heads, c.head_dim)
        k = F.linear(x, self.w[p + "k_proj.weight"]).view(B, T, c.num_kv_heads, c.head_dim)
        v = F.linear(x, self.w[p + "v_proj.weight"]).view(B, T, c.num_kv_heads, c.head_dim)
        q = self.rmsnorm(q, self.w[p + "q_norm.weight"]).transpose(1, 2)
        k = self.rmsnorm(k, self.w[p + "k_norm.weight"]).transpose(1, 2)
        v = v.transpose(1, 2)
        cs, sn = cos[None, None], sin[None, None]
        q = (q * cs) + (self._rot_half(q) * sn)
        k = (k * cs) + (self._rot_half(k) * sn)
        if cache is not None:
            if cache[l] is not None:
                k = torch.cat([cache[l][0], k], dim=2)
                v = torch.cat([cache[l][1], v], dim=2)
            cache[l] = (k, v)
        g = c.num_heads // c.num_kv_heads
        kk = k.repeat_interleave(g, dim=1)
        vv = v.repeat_interleave(g, dim=1)
        scores = (q @ kk.transpose(2, 3)) * (c.head_dim ** -0.5)
        scores = scores + add_mask
        probs = torch.softmax(scores, dim=-1, dtype=torch.float32).to(q.dtype)
        o = (probs @ vv).transpose(1, 2).reshape(B, T, c.q_size)
        return F.linear(o, self.w[p + "o_proj.weight"])

    def mlp(self, l, x):
        p = f"language_model.layers.{l}.mlp."
        g = F.linear(x, self.w[p + "gate_proj.weight"])
        u = F.linear(x, self.w[p + "up_proj.weight"])
        return F.linear(F.silu(g) * u, self.w[p + "down_proj.weight"])

    def backbone(self, x, attention_mask, cache, past_len):
        """x [B,T,H]; attention_mask [B,past_len+T] bool (True = attend)."""
        B, T, _ = x.shape
        S = past_len + T
        qpos = torch.arange(T) + past_len          # padded-column positions, shared by all rows
        cos, sin = self.rope_tables(qpos)
        allowed = (torch.arange(S)[None, :] <= qpos[:, None])[None] & attention_mask[:, None, :S].bool()
        neg = torch.finfo(self.dtype).min
        add_mask = torch.where(allowed, 0.0, neg).to(self.dtype)[:, None]   # [B,1,T,S]
        for l in range(self.cfg.num_layers):
            p = f"language_model.layers.{l}."
            h = self.rmsnorm(x, self.w[p + "input_layernorm.weight"])
            x = x + self.attention_layer(l, h, cos, sin, add_mask, cache)
            h = self.rmsnorm(x, self.w[p + "post_attention_layernorm.weight"])
            x = x + self.mlp(l, h)
        return self.rmsnorm(x, self.w["language_model.norm.weight"])

    def heads(self, h) -> List[torch.Tensor]:
        out = []
        for i in range(self.cfg.n_vq + 1):
            lg = F.linear(h, self.w[f"lm_heads.{i}.weight"])
            if i > 0:
                lg[..., -1] = float("-inf")
            out.append(lg)
        return out

    # ---- public ---------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, ids, attention_mask=None, cache=None, past_len=0, last_only=False):
        """Teacher-forced logits: list of 1+n_vq tensors [B,T,V_i] (reference forward :225-300)."""
        B, T, C = ids.shape
        if C != self.cfg.n_vq + 1:
            raise ValueError("`Input_ids`'s shape should be exactly (batch_size, sequence_length, 1 + n_vq).")
        if attention_mask is None:
            attention_mask = torch.ones(B, past_len + T, dtype=torch.bool)
        h = self.backbone(self.embed(ids), attention_mask, cache, past_len)
        if last_only:
            h = h[:, -1:, :]
        return self.heads(h)

    @torch.no_grad()
    def generate(self, input_ids, attention_mask=None, max_new_tokens=1000,
                 text_temperature=1.5, text_top_p=1.0, text_top_k=50,
                 audio_temperature=1.7, audio_top_p=0.8, audio_top_k=25,
                 audio_repetition_penalty=1.0, seed: int = 0, return_trace: bool = False):
        """Delay-pattern decode loop (reference generate :392-525), one explicit state
        machine per batch row."""
        c = self.cfg
        text_sample, audio_sample = text_temperature > 0, audio_temperature > 0
        if not text_sample:
            text_temperature = 1
        if not audio_sample:
            audio_temperature = 1
        B, S0, C = input_ids.shape
        n_vq = C - 1
        if attention_mask is None:
            attention_mask = torch.ones(B, S0, dtype=torch.bool)
        gen = input_ids.clone()
        mask = attention_mask.clone().bool()
        stopping = [False] * B
        audio_len = [0] * B
        delayed = [INT64_MAX] * B
        last_start = S.find_last_equal(input_ids[..., 0], c.audio_start_token_id)
        is_audio = [False] * B
        for b in range(B):
            last_tok = int(input_ids[b, -1, 0])
            cont = last_tok in (c.audio_start_token_id, c.audio_assistant_gen_slot_token_id)
            if cont and last_start[b] != -1:
                audio_len[b] = S0 - int(last_start[b])
                is_audio[b] = True
        cache: List[Optional[Tuple[torch.Tensor, torch.Tensor]]] = [None] * c.num_layers
        cur = input_ids
        trace = []
        for step in range(max_new_tokens):
            past_len = gen.shape[1] - cur.shape[1]
            logits = self.forward(cur, mask, cache, past_len, last_only=True)
            text_logits = (logits[0][:, -1, :] / text_temperature).clone()
            audio_logits = [lg[:, -1, :] / audio_temperature for lg in logits[1:]]
            if return_trace:
                trace.append([text_logits.clone()] + [a.clone() for a in audio_logits])
            # repetition-penalty id sets: union over the batch and the whole history (a5)
            pen_on = audio_repetition_penalty != 1.0
            set_ch0 = S.penalty_set(gen[:, :, 1]) if pen_on else None
            set_rest = S.penalty_set(gen[:, :, 2:]) if (pen_on and n_vq > 1) else None
            nxt = torch.full((B, C), c.audio_pad_code, dtype=torch.int64)
            for b in range(B):
                # ---- text channel :453-473
                if stopping[b]:
                    tok = c.pad_token_id
                elif delayed[b] < n_vq:
                    tok = c.audio_assistant_delay_slot_token_id
                elif delayed[b] == n_vq:
                    tok = c.audio_end_token_id
                    is_audio[b] = False
                else:
                    row = text_logits[b].clone()
                    if is_audio[b]:
                        keep = [c.audio_assistant_gen_slot_token_id, c.audio_assistant_delay_slot_token_id]
                        saved = row[keep].clone()
                        row[:] = float("-inf")
                        row[keep] = saved
                    else:
                        row[[c.pad_token_id, c.audio_assistant_gen_slot_token_id,
                             c.audio_assistant_delay_slot_token_id, c.audio_end_token_id]] = float("-inf")
                    if step == 0:
                        row[151662] = float("-inf")            # hard-coded in the reference :462
                    if step <= n_vq:
                        row[c.im_end_token_id] = float("-inf")
                    tok = S.sample_row(row, top_p=text_top_p, top_k=text_top_k, do_sample=text_sample,
                                       u=S.uniform_u(seed, step, b, 0))
                nxt[b, 0] = tok
                # ---- audio channels :475-503 (uses the lengths from *before* this step's update)
                for i in range(n_vq):
                    active = audio_len[b] > i and (delayed[b] == INT64_MAX or i > delayed[b] - 1)
                    if not active:
                        continue
                    row = audio_logits[i][b].clone()
                    row[c.audio_pad_code] = float("-inf")
                    prev = (set_ch0 if i == 0 else set_rest)
                    nxt[b, i + 1] = S.sample_row(row, prev_ids=prev, repetition_penalty=audio_repetition_penalty,
                                                 top_p=audio_top_p, top_k=audio_top_k, do_sample=audio_sample,
                                                 u=S.uniform_u(seed, step, b, i + 1))
            for b in range(B):
                tok = int(nxt[b, 0])
                # state update :472-473,505-509
                if tok == c.audio_start_token_id:
                    is_audio[b] = True
                if tok == c.im_end_token_id:
                    stopping[b] = True
                if tok in (c.audio_start_token_id, c.audio_assistant_gen_slot_token_id,
                           c.audio_assistant_delay_slot_token_id):
                    audio_len[b] += 1
                if tok == c.audio_end_token_id:
                    audio_len[b] = 0
                if delayed[b] == INT64_MAX and tok == c.audio_assistant_delay_slot_token_id:
                    delayed[b] = 0
                if delayed[b] != INT64_MAX:
                    delayed[b] += 1
                if delayed[b] != INT64_MAX and delayed[b] > n_vq:
                    delayed[b] = INT64_MAX
            cur = nxt[:, None, :]
            mask = torch.cat([mask, torch.tensor([[not s] for s in stopping])], dim=1)
            gen = torch.cat([gen, cur], dim=1)
            if all(stopping):
                break
        starts = S.find_last_equal(input_ids[..., 0], c.im_start_token_id) + 3
        out = [(int(S0 - starts[b]), gen[b, int(starts[b]):]) for b in range(B)]
        if return_trace:
            return out, trace
        return out
