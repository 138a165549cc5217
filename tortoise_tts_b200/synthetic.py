"""Synthetic UnifiedVoice AR weights and inputs (data generation only -- no model arithmetic): used by
bench.py for the random-init workload BASELINE.json names, and re-exported by oracle/weights.py for the tests.

Tensor names and shapes are the reference's `UnifiedVoice.state_dict()` entries that
the AR hot path touches (reference: tortoise_tts/models/unified_voice.py:396-451 and the
HF GPT2Model it builds at :408-418; Conv1D weights are stored [in, out]).  Values come
from one seeded CPU torch.Generator in a fixed order, so the same (cfg, seed) gives the
same weights in the build container (where they are loaded into the imported reference
classes to make golden vectors) and on the GPU box (where /root/reference is absent).
"""
from dataclasses import dataclass
import torch


@dataclass(frozen=True)
class ARConfig:
    layers: int = 30
    model_dim: int = 1024
    heads: int = 16
    max_text_tokens: int = 402
    max_mel_tokens: int = 604
    max_conditioning_inputs: int = 2
    number_text_tokens: int = 255
    start_text_token: int = 255
    stop_text_token: int = 0
    number_mel_codes: int = 8194
    start_mel_token: int = 8192
    stop_mel_token: int = 8193

    @property
    def max_mel_seq_len(self):          # unified_voice.py:403
        return self.max_mel_tokens + 2 + self.max_conditioning_inputs

    @property
    def max_text_seq_len(self):         # unified_voice.py:404
        return self.max_text_tokens + 2


def make_state_dict(cfg: ARConfig = ARConfig(), seed: int = 0, w_std: float = 0.02,
                    attn_std: float = None, dtype=torch.float32):
    """Returns {name: tensor} for every AR-path parameter.

    w_std is the HF GPT-2 init scale (normal(0, 0.02)).  `attn_std` (default = w_std)
    lets tests sharpen the attention logits so a wrong KV index cannot hide behind a
    near-uniform softmax.  LayerNorm gains/biases and linear biases are perturbed away
    from the ones/zeros default so that a dropped bias or gain is visible.
    """
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    d = cfg.model_dim
    attn_std = w_std if attn_std is None else attn_std

    def n(*shape, std):
        return (torch.randn(*shape, generator=g, dtype=torch.float32) * std).to(dtype)

    sd = {}
    sd["text_embedding.weight"] = n(cfg.number_text_tokens + 1, d, std=w_std)
    sd["mel_embedding.weight"] = n(cfg.number_mel_codes, d, std=w_std)
    sd["mel_pos_embedding.emb.weight"] = n(cfg.max_mel_seq_len, d, std=w_std)
    sd["text_pos_embedding.emb.weight"] = n(cfg.max_text_seq_len, d, std=w_std)
    for i in range(cfg.layers):
        p = f"gpt.h.{i}."
        sd[p + "ln_1.weight"] = 1.0 + n(d, std=0.1)
        sd[p + "ln_1.bias"] = n(d, std=0.05)
        sd[p + "attn.c_attn.weight"] = n(d, 3 * d, std=attn_std)
        sd[p + "attn.c_attn.bias"] = n(3 * d, std=w_std)
        sd[p + "attn.c_proj.weight"] = n(d, d, std=w_std)
        sd[p + "attn.c_proj.bias"] = n(d, std=w_std)
        sd[p + "ln_2.weight"] = 1.0 + n(d, std=0.1)
        sd[p + "ln_2.bias"] = n(d, std=0.05)
        sd[p + "mlp.c_fc.weight"] = n(d, 4 * d, std=w_std)
        sd[p + "mlp.c_fc.bias"] = n(4 * d, std=w_std)
        sd[p + "mlp.c_proj.weight"] = n(4 * d, d, std=w_std)
        sd[p + "mlp.c_proj.bias"] = n(d, std=w_std)
    sd["gpt.ln_f.weight"] = 1.0 + n(d, std=0.1)
    sd["gpt.ln_f.bias"] = n(d, std=0.05)
    sd["final_norm.weight"] = 1.0 + n(d, std=0.1)
    sd["final_norm.bias"] = n(d, std=0.05)
    sd["mel_head.weight"] = n(cfg.number_mel_codes, d, std=w_std)
    sd["mel_head.bias"] = n(cfg.number_mel_codes, std=w_std)
    return sd


def make_inputs(cfg: ARConfig, n_utt: int, text_lens, seed: int = 1):
    """Synthetic conditioning latents [U, d] and text ids (list of LongTensor[T_u], ids in
    [1, 254]) -- SURVEY.md section 8(d)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    latent = torch.randn(n_utt, cfg.model_dim, generator=g, dtype=torch.float32)
    texts = [torch.randint(1, cfg.number_text_tokens, (int(t),), generator=g, dtype=torch.long)
             for t in text_lens]
    return latent, texts
