"""A stand-in for the reference's `UnifiedVoice` nn.Module (TEST INFRASTRUCTURE): the attribute names and the
`state_dict()` keys that `UnifiedVoiceB200.from_module` / `patch_reference_module` read (reference
tortoise_tts/models/unified_voice.py:335-354, 396-451; HF GPT2Model under `.gpt`), no arithmetic.  /root/reference
(and its `transformers` GPT-2) cannot travel to the GPU box, so the boundary adapter is exercised against this."""
import torch
import torch.nn as nn


class Conv1D(nn.Module):
    """HF pytorch_utils.Conv1D: weight [in, out], bias [out]."""

    def __init__(self, nf, nx):
        super().__init__()
        self.nf = nf
        self.weight = nn.Parameter(torch.empty(nx, nf))
        self.bias = nn.Parameter(torch.zeros(nf))


class _Attn(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.c_attn = Conv1D(3 * d, d)
        self.c_proj = Conv1D(d, d)


class _MLP(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.c_fc = Conv1D(4 * d, d)
        self.c_proj = Conv1D(d, 4 * d)


class _Block(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.ln_1 = nn.LayerNorm(d)
        self.attn = _Attn(d)
        self.ln_2 = nn.LayerNorm(d)
        self.mlp = _MLP(d)


class _GPT(nn.Module):
    def __init__(self, layers, d):
        super().__init__()
        self.h = nn.ModuleList([_Block(d) for _ in range(layers)])
        self.ln_f = nn.LayerNorm(d)


class _Pos(nn.Module):
    def __init__(self, n, d):
        super().__init__()
        self.emb = nn.Embedding(n, d)


class StandInUnifiedVoice(nn.Module):
    def __init__(self, cfg, state_dict):
        super().__init__()
        d = cfg.model_dim
        self.layers, self.model_dim, self.heads = cfg.layers, d, cfg.heads
        self.max_text_tokens, self.max_mel_tokens = cfg.max_text_tokens, cfg.max_mel_tokens
        self.max_conditioning_inputs = cfg.max_conditioning_inputs
        self.number_text_tokens, self.start_text_token, self.stop_text_token = cfg.number_text_tokens, cfg.start_text_token, cfg.stop_text_token
        self.number_mel_codes, self.start_mel_token, self.stop_mel_token = cfg.number_mel_codes, cfg.start_mel_token, cfg.stop_mel_token
        self.mel_length_compression = 1024
        self.text_embedding = nn.Embedding(cfg.number_text_tokens + 1, d)
        self.mel_embedding = nn.Embedding(cfg.number_mel_codes, d)
        self.mel_pos_embedding = _Pos(cfg.max_mel_seq_len, d)
        self.text_pos_embedding = _Pos(cfg.max_text_seq_len, d)
        self.gpt = _GPT(cfg.layers, d)
        self.final_norm = nn.LayerNorm(d)
        self.mel_head = nn.Linear(d, cfg.number_mel_codes)
        missing, unexpected = self.load_state_dict(state_dict, strict=True)
        assert not missing and not unexpected

    def forward(self, *a, **k):     # the training-loss forward of the reference: not on the AR inference path
        return "reference forward"


class LoRADelta(nn.Module):
    """Same arithmetic as the reference's ParameterizedLoRA.forward (lora.py:121-124): x + (B @ A) * scaling."""

    def __init__(self, shape, rank=4, alpha=8, seed=0):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.lora_B = nn.Parameter(torch.randn(shape[0], rank, generator=g) * 0.05)
        self.lora_A = nn.Parameter(torch.randn(rank, shape[1], generator=g) * 0.05)
        self.scaling = alpha / rank
        self.enabled = True

    def forward(self, x):
        if self.enabled:
            return x + torch.matmul(self.lora_B, self.lora_A).view(x.shape) * self.scaling
        return x
