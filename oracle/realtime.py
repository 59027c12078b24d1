"""Oracle for MOSS-TTS-Realtime's step-wise decode (TEST INFRASTRUCTURE ONLY) — SURVEY.md §8 row f4.

Restates, in plain torch-CPU with explicit per-row loops, the algorithm of
  reference moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py
      MossTTSRealtimeInference.prefill :121-218     step :220-272      finish :274-299
      generate_local_transformer :301-357           apply_repetition_penalty :359-375
      sample_token :377-397  apply_top_k :399-406  apply_top_p :408-421
  reference moss_tts_realtime/mossttsrealtime/modeling_mossttsrealtime.py
      MossTTSRealtime.get_input_embeddings :101-109 (text + rvq audio embeddings, summed left to right)
      forward :111-142 (Qwen3Model backbone, last_hidden_state AFTER the final norm)
  reference moss_tts_realtime/mossttsrealtime/modeling_mossttsrealtime_local.py
      RMSNorm :40-52, MLP :58-72, rotary :75-88,230-255, attention :123-186 (q/k norm, RoPE, StaticCache of `rvq`
      slots written at cache_position = channel index), decoder layer :189-227, model forward :270-378
      (channel 0 input = backbone hidden state; channel i > 0 input = embed_tokens[i-1](token i-1)),
      heads :381-448 (local_lm_heads[codebook_idx] on the final-norm state)

Notes that matter for parity
  * the StaticCache holds `rvq` slots; slots past the current channel are masked by the causal mask, so a depth step at
    channel i attends to slots 0..i — restated as a growing K/V list.
  * repetition penalty is applied to the RAW logits of every channel (also when decoding greedily), per row, over the last
    `repetition_window` frames of that row's history of the channel (`generated_tokens[:, :gen_step, i]`); duplicates in the
    window gather the same unpenalised score, so a token is penalised once.  `scores * penalty` / `scores / penalty` are
    computed in the working dtype.
  * greedy = argmax of the penalised logits (`not do_sample or temperature == 0`); sampling = logits / temperature -> top-k
    -> top-p -> softmax -> draw.  The draw is this test suite's explicit inverse CDF with counter-hash uniforms
    (oracle/sampling.py) in place of torch.multinomial's RNG stream; top-k ties: lower index wins.
  * a stopped row (channel-0 token == audio_eos) keeps being computed; its later positions are masked out of the backbone's
    attention mask for the steps that follow (`(~is_stopping)` appended, :247).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

from moss_tts_b200.config import RealtimeConfig
from . import sampling as S


class OracleRealtime:
    def __init__(self, cfg: RealtimeConfig, weights: Dict[str, torch.Tensor], dtype=torch.float32):
        self.cfg, self.dtype = cfg, dtype
        self.w = {k: v.to(dtype) for k, v in weights.items()}
        half = torch.arange(0, cfg.head_dim, 2, dtype=torch.int64).float() / cfg.head_dim
        self.inv_freq = 1.0 / (cfg.rope_theta ** half)
        half = torch.arange(0, cfg.local_head_dim, 2, dtype=torch.int64).float() / cfg.local_head_dim
        self.local_inv_freq = 1.0 / (cfg.local_rope_theta ** half)
        self.reset()

    def reset(self):
        self.cache: List = [None] * self.cfg.num_layers
        self.mask: Optional[torch.Tensor] = None        # [B, filled positions] bool
        self.history: List[torch.Tensor] = []           # frames [B, rvq]
        self.stopping: Optional[torch.Tensor] = None
        self.step_idx = 0

    # ---- building blocks ----------------------------------------------------------------------------------
    def rmsnorm(self, x, gain, eps):
        xf = x.float()
        xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
        return gain * xf.to(x.dtype)

    @staticmethod
    def _rot_half(x):
        h = x.shape[-1] // 2
        return torch.cat([-x[..., h:], x[..., :h]], dim=-1)

    def _rope(self, positions, inv_freq):
        ang = positions.float()[:, None] * inv_freq[None, :]
        ang = torch.cat([ang, ang], dim=-1)
        return ang.cos().to(self.dtype), ang.sin().to(self.dtype)

    def _attention(self, p, x, allowed, rope, cache, l, n_heads, n_kv, D, eps):
        B, T, _ = x.shape
        q = F.linear(x, self.w[p + "q_proj.weight"]).view(B, T, n_heads, D)
        k = F.linear(x, self.w[p + "k_proj.weight"]).view(B, T, n_kv, D)
        v = F.linear(x, self.w[p + "v_proj.weight"]).view(B, T, n_kv, D)
        q = self.rmsnorm(q, self.w[p + "q_norm.weight"], eps).transpose(1, 2)
        k = self.rmsnorm(k, self.w[p + "k_norm.weight"], eps).transpose(1, 2)
        v = v.transpose(1, 2)
        cs, sn = rope[0][None, None], rope[1][None, None]
        q = (q * cs) + (self._rot_half(q) * sn)
        k = (k * cs) + (self._rot_half(k) * sn)
        if cache[l] is not None:
            k = torch.cat([cache[l][0], k], dim=2)
            v = torch.cat([cache[l][1], v], dim=2)
        cache[l] = (k, v)
        g = n_heads // n_kv
        kk, vv = k.repeat_interleave(g, dim=1), v.repeat_interleave(g, dim=1)
        scores = (q @ kk.transpose(2, 3)) * (D ** -0.5)
        neg = torch.finfo(self.dtype).min
        scores = scores + torch.where(allowed, 0.0, neg).to(self.dtype)
        probs = torch.softmax(scores, dim=-1, dtype=torch.float32).to(q.dtype)
        o = (probs @ vv).transpose(1, 2).reshape(B, T, n_heads * D)
        return F.linear(o, self.w[p + "o_proj.weight"])

    def _mlp(self, p, x):
        g = F.linear(x, self.w[p + "gate_proj.weight"])
        u = F.linear(x, self.w[p + "up_proj.weight"])
        return F.linear(F.silu(g) * u, self.w[p + "down_proj.weight"])

    def _stack(self, prefix, n_layers, x, allowed, rope, cache, final_norm, n_heads, n_kv, D, eps):
        for l in range(n_layers):
            p = f"{prefix}layers.{l}."
            h = self.rmsnorm(x, self.w[p + "input_layernorm.weight"], eps)
            x = x + self._attention(p + "self_attn.", h, allowed, rope, cache, l, n_heads, n_kv, D, eps)
            h = self.rmsnorm(x, self.w[p + "post_attention_layernorm.weight"], eps)
            x = x + self._mlp(p + "mlp.", h)
        return self.rmsnorm(x, self.w[final_norm], eps)

    # ---- backbone -----------------------------------------------------------------------------------------
    def embed(self, ids):
        """modeling_mossttsrealtime.py:101-109."""
        x = self.w["embed_tokens.0.weight"][ids[..., 0]]
        for i in range(1, self.cfg.rvq + 1):
            x = x + self.w[f"embed_tokens.{i}.weight"][ids[..., i]]
        return x

    def backbone(self, ids, mask):
        """ids [B,T,1+rvq] appended to the cache; mask [B, past+T] bool.  Returns the final-norm state of the last position."""
        c = self.cfg
        B, T, _ = ids.shape
        S_ = mask.shape[1]
        past = S_ - T
        qpos = torch.arange(T) + past
        rope = self._rope(qpos, self.inv_freq)
        allowed = ((torch.arange(S_)[None, :] <= qpos[:, None])[None] & mask[:, None, :].bool())[:, None]
        h = self._stack("language_model.", c.num_layers, self.embed(ids), allowed, rope, self.cache,
                        "language_model.norm.weight", c.num_heads, c.num_kv_heads, c.head_dim, c.rms_norm_eps)
        return h[:, -1, :]

    # ---- local transformer (streaming_mossttsrealtime.py:301-357) --------------------------------------------
    def local_logits(self, hidden, forced: Optional[torch.Tensor] = None, chooser=None):
        """Depth loop over the rvq channels.  `chooser(i, logits[B,V]) -> tokens[B]` picks each channel's token (or `forced`
        [B,rvq] teacher-forces them).  Returns (tokens [B,rvq], list of rvq raw-logit tensors [B,V])."""
        c = self.cfg
        B = hidden.shape[0]
        cache: List = [None] * c.local_num_layers
        x = hidden[:, None, :]
        toks = torch.zeros(B, c.rvq, dtype=torch.int64)
        all_logits = []
        for i in range(c.rvq):
            rope = self._rope(torch.tensor([i]), self.local_inv_freq)
            allowed = torch.ones(1, 1, 1, i + 1, dtype=torch.bool)
            h = self._stack("local_transformer.model.", c.local_num_layers, x, allowed, rope, cache,
                            "local_transformer.model.norm.weight", c.local_num_heads, c.local_num_kv_heads,
                            c.local_head_dim, c.local_rms_norm_eps)
            lg = F.linear(h[:, 0, :], self.w[f"local_transformer.local_lm_heads.{i}.weight"])
            all_logits.append(lg)
            tk = forced[:, i] if forced is not None else chooser(i, lg)
            toks[:, i] = tk
            if i + 1 < c.rvq:
                x = self.w[f"local_transformer.model.embed_tokens.{i}.weight"][tk][:, None, :]
        return toks, all_logits

    def penalise(self, lg, history, i, gen_step, penalty, window):
        """apply_repetition_penalty :359-375 on channel i: lg [B,V], history [B,T,rvq] or None."""
        if not penalty or penalty == 1.0 or history is None:
            return lg
        out = lg.clone()
        for b in range(lg.shape[0]):
            ht = history[b, :gen_step, i]
            if window is not None and window > 0:
                ht = ht[-window:]
            if ht.numel():
                out[b] = S.apply_penalty_row(lg[b], np.unique(ht.numpy()), penalty)
        return out

    def generate_local_transformer(self, hidden, *, temperature=0.8, top_p=0.6, top_k=30, do_sample=True,
                                   repetition_penalty=1.1, repetition_window=50, generated_tokens=None, gen_step=0,
                                   seed=0, return_logits=False):
        def chooser(i, lg):
            lg = self.penalise(lg, generated_tokens, i, gen_step, repetition_penalty, repetition_window)
            out = []
            for b in range(lg.shape[0]):
                if not do_sample or temperature == 0:
                    out.append(int(torch.argmax(lg[b].float()).item()))
                else:
                    row = lg[b] / temperature
                    out.append(S.sample_row(row, top_p=top_p, top_k=top_k, do_sample=True,
                                            u=S.uniform_u(seed, gen_step, b, i)))
            return torch.tensor(out, dtype=torch.int64)
        toks, logits = self.local_logits(hidden, chooser=chooser)
        return (toks, logits) if return_logits else toks

    # ---- session (prefill / step) ----------------------------------------------------------------------------
    @torch.no_grad()
    def prefill(self, input_ids, attention_mask, **kw):
        """input_ids [B,S0,1+rvq] already concatenated with the text prefix and left-padded (:139-176)."""
        mask = attention_mask.bool()
        if self.mask is not None:
            mask = torch.cat([self.mask, mask], dim=1)
        h = self.backbone(input_ids, mask)
        self.mask = mask
        self.last_hidden = h
        toks = self.generate_local_transformer(h, generated_tokens=None, gen_step=0, **kw)
        self.history = [toks]
        self.stopping = toks[:, 0] == self.cfg.audio_eos_token
        self.step_idx = 1
        return toks

    @torch.no_grad()
    def step(self, text_token=None, **kw):
        c = self.cfg
        last = self.history[-1]
        B = last.shape[0]
        if bool(self.stopping.all()):
            return last
        if text_token is None:
            text_token = [c.text_pad_id] * B
        text = torch.as_tensor(text_token, dtype=torch.int64).reshape(B)
        ids = torch.cat([text[:, None, None], last[:, None, :]], dim=2)
        self.mask = torch.cat([self.mask, (~self.stopping)[:, None]], dim=1)
        h = self.backbone(ids, self.mask)
        self.last_hidden = h
        hist = torch.stack(self.history, dim=1)
        toks = self.generate_local_transformer(h, generated_tokens=hist, gen_step=self.step_idx, **kw)
        self.history.append(toks)
        self.stopping = self.stopping | (toks[:, 0] == c.audio_eos_token)
        self.step_idx += 1
        return toks
