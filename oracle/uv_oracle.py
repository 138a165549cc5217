"""CPU restatement of the reference AR hot path (TEST INFRASTRUCTURE, see oracle/__init__.py).

Plain torch fp32, no dependency on /root/reference or on `transformers`, so it runs on the
GPU box.  Every function cites what it restates.  "ref" = /root/reference/tortoise_tts,
"HF" = transformers 5.5.0 (the un-vendored dependency that owns the GPT-2 block math,
setup.py:56).  Pinned against the imported reference classes by oracle/make_golden.py ->
tests/golden/*.npz and tests/test_oracle_golden.py.
"""
import math
import numpy as np
import torch
import torch.nn.functional as F

from .weights import ARConfig
from . import philox


def gelu_new(x):
    """HF activations.NewGELUActivation (GPT2Config.activation_function default "gelu_new")."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


class OracleAR:
    """UnifiedVoice AR sampler, restated.  Weights: {name: fp32 tensor} as in oracle/weights.py."""

    def __init__(self, sd, cfg: ARConfig = ARConfig(), dtype=torch.float32, device="cpu"):
        self.cfg = cfg
        self.dtype = dtype
        self.device = device
        self.sd = {k: v.to(device=device, dtype=dtype) for k, v in sd.items()}
        self.hd = cfg.model_dim // cfg.heads

    # ---- embeddings -------------------------------------------------------------------
    def build_prefix(self, latent, text):
        """ref models/unified_voice.py:639-644 (inference_speech) == :615-619 (compute_embeddings):
        text -> pad stop(0) right, start(255) left; text_embedding + text_pos_embedding(arange);
        prefix = cat([latent[:,None], text_emb]) -> [U, P=T+3, d]."""
        cfg = self.cfg
        text = F.pad(text, (0, 1), value=cfg.stop_text_token)
        text = F.pad(text, (1, 0), value=cfg.start_text_token)
        te = self.sd["text_embedding.weight"][text]
        tp = self.sd["text_pos_embedding.emb.weight"][torch.arange(text.shape[1], device=text.device)]
        emb = te + tp[None]
        return torch.cat([latent.to(self.dtype)[:, None, :], emb], dim=1)

    def mel_emb(self, tok, pos_index):
        """ref unified_voice.py:205-206 (prefill, pos via LearnedPositionEmbeddings.forward = arange)
        and :213-214 (decode, get_fixed_embedding(attention_mask.shape[1]-mel_len))."""
        return self.sd["mel_embedding.weight"][tok] + self.sd["mel_pos_embedding.emb.weight"][pos_index]

    # ---- GPT-2 block (HF modeling_gpt2.GPT2Block / GPT2Attention / GPT2MLP / Conv1D) ---
    def _ln(self, x, name):
        return F.layer_norm(x, (self.cfg.model_dim,), self.sd[name + ".weight"], self.sd[name + ".bias"], 1e-5)

    def _block(self, x, i, past):
        """x: [B,S,d]; past: None or (k,v) each [B,H,L,hd].  Returns (x, (k,v)).
        HF GPT2Block.forward: x += c_proj(attn(ln_1 x)); x += c_proj(gelu_new(c_fc(ln_2 x))).
        Conv1D = addmm(bias, x, W[in,out]).  Attention: softmax(q k^T / sqrt(hd) + causal) v,
        softmax in fp32 (HF eager_attention_forward / SDPA)."""
        cfg, sd, p = self.cfg, self.sd, f"gpt.h.{i}."
        B, S, d = x.shape
        H, hd = cfg.heads, self.hd
        h = self._ln(x, p + "ln_1")
        qkv = h @ sd[p + "attn.c_attn.weight"] + sd[p + "attn.c_attn.bias"]
        q, k, v = qkv.split(d, dim=2)
        q = q.view(B, S, H, hd).transpose(1, 2)
        k = k.view(B, S, H, hd).transpose(1, 2)
        v = v.view(B, S, H, hd).transpose(1, 2)
        if past is not None:                      # HF DynamicCache.update: cat on dim=-2
            k = torch.cat([past[0], k], dim=-2)
            v = torch.cat([past[1], v], dim=-2)
        L = k.shape[-2]
        att = (q.float() @ k.float().transpose(-1, -2)) / math.sqrt(hd)
        if S > 1:                                 # causal mask over the last S positions
            qpos = torch.arange(L - S, L, device=x.device)[:, None]
            kpos = torch.arange(L, device=x.device)[None, :]
            att = att.masked_fill(kpos > qpos, float("-inf"))
        att = torch.softmax(att, dim=-1).to(self.dtype)
        a = (att @ v).transpose(1, 2).reshape(B, S, d)
        x = x + (a @ sd[p + "attn.c_proj.weight"] + sd[p + "attn.c_proj.bias"])
        h = self._ln(x, p + "ln_2")
        h = gelu_new(h @ sd[p + "mlp.c_fc.weight"] + sd[p + "mlp.c_fc.bias"])
        x = x + (h @ sd[p + "mlp.c_proj.weight"] + sd[p + "mlp.c_proj.bias"])
        return x, (k, v)

    def transformer(self, emb, past=None):
        """HF GPT2Model.forward with inputs_embeds; wpe == zeros (ref unified_voice.py:77-78,
        425-426); 30 blocks; ln_f.  Returns (last_hidden_state, new_past)."""
        x = emb
        new_past = []
        for i in range(self.cfg.layers):
            x, kv = self._block(x, i, None if past is None else past[i])
            new_past.append(kv)
        return self._ln(x, "gpt.ln_f"), new_past

    def head(self, hidden):
        """ref unified_voice.py:106,239: lm_head = Sequential(final_norm, mel_head) applied on the
        ln_f output (quirk Q1: two LayerNorms).  Returns (logits, latent=final_norm(hidden))."""
        lat = self._ln(hidden, "final_norm")
        return lat @ self.sd["mel_head.weight"].t() + self.sd["mel_head.bias"], lat

    # ---- prefill / step -----------------------------------------------------------------
    def prefill(self, latent, text, n_cand=1):
        """GPT2InferenceModel.forward prefill branch (ref unified_voice.py:203-211) for
        input_ids = [1]*P + [start_mel]: emb = cat([prefix repeat_interleave(N), mel_emb(start)+pos[0]]).
        Returns (logits_last [B,V], latent_last [B,d], past, P)."""
        prefix = self.build_prefix(latent, text)
        U, P, d = prefix.shape
        start = torch.full((U, 1), self.cfg.start_mel_token, dtype=torch.long, device=prefix.device)
        emb = torch.cat([prefix, self.mel_emb(start, torch.arange(1, device=prefix.device))], dim=1)
        emb = emb.repeat_interleave(n_cand, 0)
        hidden, past = self.transformer(emb)
        logits, lat = self.head(hidden[:, -1])
        return logits, lat, past, P

    def step(self, tok, k_generated, past):
        """Cached decode branch (ref unified_voice.py:212-214).  `tok` [B] is the k-th generated
        token (k>=1); attention_mask.shape[1] - mel_len == (P+1+k) - P == k+1 -> mel_pos[k+1]
        (quirk Q2: index 1 is never used)."""
        emb = self.mel_emb(tok[:, None], torch.tensor([k_generated + 1], device=tok.device))
        hidden, past = self.transformer(emb, past)
        logits, lat = self.head(hidden[:, -1])
        return logits, lat, past

    # ---- sample loop ----------------------------------------------------------------------
    def generate(self, latent, text, n_cand=1, max_new=None, do_sample=False, temperature=1.0,
                 top_k=0, top_p=1.0, repetition_penalty=1.0, seed=1234, seq_index_base=0,
                 ignore_eos=False, forced=None, want_logits=0, sampler="philox", typical_mass=None):
        """ref stream_generator.py:1106-1190 (sample_stream), driven by the ref's
        inference_speech bookkeeping (unified_voice.py:647-668).  Per step: forward -> logits[:, -1]
        -> processors (RepetitionPenalty over ALL input_ids incl. the dummy prompt ids 1 and 8192,
        quirk Q4) -> warpers (Temperature, TopK, TopP; order of _get_logits_warper :56-101) ->
        softmax/multinomial (or argmax when do_sample=False) -> pad finished rows with 8193 (Q7) ->
        append -> unfinished &= tok != eos -> stop when all finished or max_length reached.

        sampler="philox": multinomial = inverse CDF in vocabulary order with a Philox4x32-10 uniform
        keyed by (seed, global sequence index, step) -- the CUDA path's definition (torch's global
        generator cannot be reproduced on device, quirk Q5).  sampler="torch": torch.multinomial.
        `forced` [B, max_new]: teacher forcing (the token appended is forced[:, s]; the model's own
        choice is still recorded in out["choice"]).
        Returns dict(codes [B, n_steps] int64, choice, logits [want_logits, B, V], latents, pos_trace)."""
        cfg = self.cfg
        eos = pad = cfg.stop_mel_token
        max_new = (cfg.max_mel_tokens - 1) if max_new is None else max_new
        logits, lat, past, P = self.prefill(latent, text, n_cand)
        B = logits.shape[0]
        dev = logits.device
        # fake_inputs: [1]*P + [start_mel]  (ref unified_voice.py:647-649), expanded N times
        ids = torch.full((B, P + 1), 1, dtype=torch.long, device=dev)
        ids[:, -1] = cfg.start_mel_token
        unfinished = torch.ones(B, dtype=torch.long, device=dev)
        codes, choice, keep_logits, lats, pos_trace = [], [], [], [], [0]
        for s in range(max_new):
            if len(keep_logits) < want_logits:
                keep_logits.append(logits.float().clone())
            lats.append(lat.float().clone())
            scores = logits.float()
            scores = apply_repetition_penalty(scores, ids, repetition_penalty)
            scores = apply_typical(scores, typical_mass)      # custom processor, merged after the default ones (:659)
            if do_sample:
                scores = apply_temperature(scores, temperature)
                scores = apply_top_k(scores, top_k)
                scores = apply_top_p(scores, top_p)
                if sampler == "torch":
                    nxt = torch.multinomial(torch.softmax(scores, -1), 1).squeeze(1)
                else:
                    nxt = torch.as_tensor(philox.multinomial_rows(
                        scores.cpu().numpy(), seed, seq_index_base + np.arange(B), s), device=dev)
            else:
                nxt = scores.argmax(-1)
            choice.append(nxt.clone())
            if forced is not None:
                nxt = forced[:, s].to(dev)
            nxt = nxt * unfinished + pad * (1 - unfinished)
            codes.append(nxt)
            ids = torch.cat([ids, nxt[:, None]], dim=1)
            if not ignore_eos:
                unfinished = unfinished * (nxt != eos).long()
            if unfinished.max() == 0 or ids.shape[1] >= P + 1 + max_new:
                break
            pos_trace.append(s + 2)
            logits, lat, past = self.step(nxt, s + 1, past)
        return dict(codes=torch.stack(codes, 1), choice=torch.stack(choice, 1),
                    logits=torch.stack(keep_logits) if keep_logits else None,
                    latents=torch.stack(lats, 1), pos_trace=pos_trace, prefix_len=P)

    # ---- latent re-forward ------------------------------------------------------------------
    def forward_latents(self, latent, text, codes):
        """ref unified_voice.py:544-600 with return_latent=True, clip_inputs=False, text_first=True,
        and wav_lengths = M*1024 (set_mel_padding is then a no-op, :494-506): one causal pass over
        [cond, start_text+text+stop_text+stop_text?, ...].  Exactly: text -> pad stop, then
        build_aligned -> [start, text, stop] (T+2); mel -> pad stop, build_aligned -> [start, codes,
        stop] (M+2); positions contiguous 0..; enc = final_norm(ln_f(h)[:, 1:]); returns the last
        M+2 rows minus 2 -> [B, M, d]."""
        cfg, sd = self.cfg, self.sd
        t = F.pad(text, (0, 1), value=cfg.stop_text_token)
        t = F.pad(t, (1, 0), value=cfg.start_text_token)
        te = sd["text_embedding.weight"][t] + sd["text_pos_embedding.emb.weight"][torch.arange(t.shape[1])][None]
        m = F.pad(codes, (0, 1), value=cfg.stop_mel_token)
        m = F.pad(m, (1, 0), value=cfg.start_mel_token)
        me = sd["mel_embedding.weight"][m] + sd["mel_pos_embedding.emb.weight"][torch.arange(m.shape[1])][None]
        emb = torch.cat([latent.to(self.dtype)[:, None, :], te, me], dim=1)
        hidden, _ = self.transformer(emb)
        enc = self._ln(hidden[:, 1:], "final_norm")
        return enc[:, -m.shape[1]:][:, :-2]


# ---- HF logits processors / warpers, restated (transformers/generation/logits_process.py) ----
def apply_repetition_penalty(scores, ids, penalty):
    """RepetitionPenaltyLogitsProcessor: s = gather(scores, ids); s = where(s<0, s*p, s/p); scatter."""
    if penalty == 1.0:
        return scores
    s = torch.gather(scores, 1, ids)
    s = torch.where(s < 0, s * penalty, s / penalty)
    return scores.scatter(1, ids, s)


def apply_temperature(scores, t):
    """TemperatureLogitsWarper: scores / temperature (only appended when != 1.0)."""
    return scores if t == 1.0 else scores / t


def apply_top_k(scores, k):
    """TopKLogitsWarper: remove scores < (k-th largest); filter value -inf."""
    if not k:
        return scores
    k = min(int(k), scores.shape[-1])
    kth = torch.topk(scores, k)[0][..., -1, None]
    return scores.masked_fill(scores < kth, float("-inf"))


def apply_top_p(scores, p):
    """TopPLogitsWarper: ascending sort, softmax, cumsum; remove cum <= 1-p; keep >= 1 token."""
    if p >= 1.0:
        return scores
    srt, idx = torch.sort(scores, descending=False)
    cum = srt.softmax(dim=-1).cumsum(dim=-1)
    rem = cum <= (1 - p)
    rem[..., -1:] = False
    rem = rem.scatter(1, idx, rem)
    return scores.masked_fill(rem, float("-inf"))


def apply_typical(scores, mass):
    """Reference TypicalLogitsWarper (unified_voice.py:47-75; handed to generate() as a logits PROCESSOR by
    inference_speech :659, so it runs after the repetition penalty and before temperature / top-k / top-p, in greedy
    mode too): entropy of softmax(scores); shifted = |-log p - H|; ascending sort of shifted; cumulative
    probability in that order; last_ind = #{cum < mass}; remove shifted > sorted_shifted[last_ind]."""
    if mass is None or not (0.0 < mass < 1.0):
        return scores
    normalized = torch.log_softmax(scores, dim=-1)
    p = torch.exp(normalized)
    ent = -(normalized * p).nansum(-1, keepdim=True)
    shifted = torch.abs((-normalized) - ent)
    srt, idx = torch.sort(shifted, descending=False)
    cum = scores.gather(-1, idx).softmax(dim=-1).cumsum(dim=-1)
    last_ind = (cum < mass).sum(dim=1)
    last_ind[last_ind < 0] = 0
    rem = srt > srt.gather(1, last_ind.view(-1, 1))
    rem = rem.scatter(1, idx, rem)
    return scores.masked_fill(rem, float("-inf"))
