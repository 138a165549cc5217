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


# ---- CLVP (candidate scoring): reference tortoise_tts/models/clvp.py state_dict keys ----------------
@dataclass
class ClvpConfig:
    dim: int = 768
    heads: int = 12
    depth: int = 20                     # both encoders (clvp.py:33,38)
    num_text_tokens: int = 256
    num_speech_tokens: int = 8192
    dim_head: int = 64
    rotary_dim: int = 32
    ff_inner: int = 1536                # dim * ff_mult (2)


def make_clvp_state_dict(cfg: ClvpConfig, seed=0, bf16=True):
    """Synthetic weights under the reference's state_dict keys, drawn like the reference's default initialisers
    (N(0,1) embeddings, U(+-1/sqrt(fan_in)) linears) so that activations have realistic scales; `bf16` rounds every
    tensor to a bf16-representable value (the fp32 reference and the bf16 CUDA path then hold identical weights)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def lin(out_f, in_f):
        b = in_f ** -0.5
        return (torch.rand(out_f, in_f, generator=g) * 2 - 1) * b

    def vec(n, scale):
        return torch.randn(n, generator=g) * scale

    sd["temperature"] = torch.tensor(1.0)
    for side, n_tok in (("text", cfg.num_text_tokens), ("speech", cfg.num_speech_tokens)):
        sd[f"{side}_emb.weight"] = torch.randn(n_tok, cfg.dim, generator=g)
        sd[f"to_{side}_latent.weight"] = lin(cfg.dim, cfg.dim)
        p = f"{side}_transformer.transformer."
        for i in range(cfg.depth):
            a, f = f"{p}attn_layers.layers.{2 * i}.", f"{p}attn_layers.layers.{2 * i + 1}."
            sd[a + "0.0.g"] = 1 + vec(cfg.dim, 0.1)
            for n in ("to_q", "to_k", "to_v"):
                sd[a + f"1.wrap.{n}.weight"] = lin(cfg.heads * cfg.dim_head, cfg.dim)
            sd[a + "1.wrap.to_out.weight"] = lin(cfg.dim, cfg.heads * cfg.dim_head)
            sd[a + "1.wrap.to_out.bias"] = vec(cfg.dim, 0.02)
            sd[f + "0.0.g"] = 1 + vec(cfg.dim, 0.1)
            sd[f + "1.wrap.net.0.proj.weight"] = lin(2 * cfg.ff_inner, cfg.dim)
            sd[f + "1.wrap.net.0.proj.bias"] = vec(2 * cfg.ff_inner, 0.02)
            sd[f + "1.wrap.net.3.weight"] = lin(cfg.dim, cfg.ff_inner)
            sd[f + "1.wrap.net.3.bias"] = vec(cfg.dim, 0.02)
        sd[p + "attn_layers.rotary_pos_emb.inv_freq"] = 1.0 / (10000 ** (torch.arange(0, cfg.rotary_dim, 2).float() / cfg.rotary_dim))
        sd[p + "norm.weight"] = 1 + vec(cfg.dim, 0.1)
        sd[p + "norm.bias"] = vec(cfg.dim, 0.02)
    if bf16:
        sd = {k: (v if k.endswith("inv_freq") else v.bfloat16().float()) for k, v in sd.items()}
    return sd


def make_clvp_inputs(cfg: ClvpConfig, n, text_len, speech_len, seed=0):
    g = torch.Generator().manual_seed(seed)
    text = torch.randint(1, cfg.num_text_tokens, (1, text_len), generator=g)
    codes = torch.randint(0, cfg.num_speech_tokens, (n, speech_len), generator=g)
    return text, codes


# ---- conditioning-latent encoder: reference unified_voice.py:269-293 state_dict keys ------------------
def make_cond_state_dict(seed: int = 0, dim: int = 1024, spec_dim: int = 80, blocks: int = 6, bf16: bool = True):
    """`conditioning_encoder.*` entries of UnifiedVoice.state_dict(): init 1x1 conv, then per AttentionBlock GroupNorm
    gain/bias, qkv 1x1 conv, proj_out 1x1 conv.  proj_out is zero-initialised in the reference (zero_module, arch_utils.py:171):
    with that default every block is the identity, so the synthetic values are drawn non-zero.  `bf16` rounds the GEMM weights
    (qkv, proj_out) to bf16-representable values so that the fp32 reference and the CUDA path hold identical weights."""
    g = torch.Generator().manual_seed(seed)
    p = "conditioning_encoder."
    sd = {p + "init.weight": (torch.rand(dim, spec_dim, 1, generator=g) * 2 - 1) * spec_dim ** -0.5,
          p + "init.bias": torch.randn(dim, generator=g) * 0.05}
    for i in range(blocks):
        a = f"{p}attn.{i}."
        sd[a + "norm.weight"] = 1 + torch.randn(dim, generator=g) * 0.1
        sd[a + "norm.bias"] = torch.randn(dim, generator=g) * 0.05
        sd[a + "qkv.weight"] = torch.randn(3 * dim, dim, 1, generator=g) * 0.06
        sd[a + "qkv.bias"] = torch.randn(3 * dim, generator=g) * 0.02
        sd[a + "proj_out.weight"] = torch.randn(dim, dim, 1, generator=g) * 0.03
        sd[a + "proj_out.bias"] = torch.randn(dim, generator=g) * 0.02
    if bf16:
        sd = {k: (v.bfloat16().float() if k.endswith(("qkv.weight", "proj_out.weight")) else v) for k, v in sd.items()}
    return sd


def make_cond_mel(n_batch: int, n_clips: int, frames: int, seed: int = 0, spec_dim: int = 80):
    """A mel-spectrogram-shaped input [n_batch, n_clips, 80, frames]: log-mel values are roughly in [-11.5, 2]."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n_batch, n_clips, spec_dim, frames, generator=g) * 2.0 - 5.0
