"""CPU oracle of CLVP candidate scoring -- TEST INFRASTRUCTURE, never on a product path.

fp32 restatement of what the reference runs for `clvp(text_tokens.repeat(n, 1), codes, return_loss=False)`
(reference tortoise_tts/inference.py:392-396) in eval mode:

  CLVP.forward                         tortoise_tts/models/clvp.py:100-133
    two x-transformers encoders        clvp.py:55-84 (dim 768, depth 20, heads 12, ff_mult 2, ff_glu, use_rmsnorm, rotary)
      AttentionLayers.forward          tortoise_tts/models/xtransformers.py:912-1016 (pre-norm, residual, layer types a,f,...)
      RMSNorm                          xtransformers.py:337-346  (x / max(||x|| * dim^-0.5, 1e-8) * g)
      Attention.forward                xtransformers.py:583-731  (no-bias q/k/v, 64-dim heads, scale 64^-0.5, softmax, to_out with bias;
                                        eval masks are all-true (clvp.py:110-112) so nothing is masked)
      rotary embedding                 xtransformers.py:266-290, 786-787, 660-664: rotary dim = max(64 // 2, 32) = 32, applied to the
                                        first 32 dims of q, k AND v; rotate_half pairs dim i with dim i + 16
      FeedForward with GLU             xtransformers.py:431-479  (proj 768 -> 3072, x * gelu(gate) with exact erf GELU, 1536 -> 768)
      final LayerNorm                  xtransformers.py:1217, 1238 (ContinuousTransformerWrapper.norm)
    masked_mean (all-true mask)        clvp.py:15-17, 120-121
    to_*_latent, L2 normalise          clvp.py:123-126
    sim = <text, speech> * exp(temp)   clvp.py:128-132

Pinned by tests/test_oracle_golden.py against tests/golden/clvp.npz, which oracle/make_golden_clvp.py produced from the
reference's own CLVP class.
"""
import torch
import torch.nn.functional as F


from tortoise_tts_b200.synthetic import ClvpConfig, make_clvp_inputs, make_clvp_state_dict  # noqa: F401,E402  (pure data generation)


class OracleCLVP:
    def __init__(self, sd, cfg: ClvpConfig):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.cfg = cfg

    def rotary(self, t, freqs):
        """apply_rotary_pos_emb on the first rotary_dim dims (xtransformers.py:279-290, 660-664); t [b, h, n, 64]."""
        l = freqs.shape[-1]
        tl, tr = t[..., :l], t[..., l:]
        x1, x2 = tl[..., : l // 2], tl[..., l // 2:]
        rot = torch.cat((-x2, x1), dim=-1)
        return torch.cat((tl * freqs.cos() + rot * freqs.sin(), tr), dim=-1)

    def encode(self, side, tokens):
        """[b, n] ids -> [b, n, dim] encoder output after the final LayerNorm."""
        c, sd = self.cfg, self.sd
        p = f"{side}_transformer.transformer."
        x = sd[f"{side}_emb.weight"][tokens]
        b, n, _ = x.shape
        inv = sd[p + "attn_layers.rotary_pos_emb.inv_freq"]
        fr = torch.arange(n, device=inv.device).float()[:, None] * inv[None, :]
        freqs = torch.cat((fr, fr), dim=-1)[None, None]                    # [1, 1, n, 32]

        def rms(v, g):
            norm = torch.norm(v, dim=-1, keepdim=True) * c.dim ** -0.5
            return v / norm.clamp(min=1e-8) * g

        for i in range(c.depth):
            a, f = f"{p}attn_layers.layers.{2 * i}.", f"{p}attn_layers.layers.{2 * i + 1}."
            h = rms(x, sd[a + "0.0.g"])
            q, k, v = (F.linear(h, sd[a + f"1.wrap.{nm}.weight"]).view(b, n, c.heads, c.dim_head).transpose(1, 2)
                       for nm in ("to_q", "to_k", "to_v"))
            q, k, v = (self.rotary(t, freqs) for t in (q, k, v))
            dots = torch.einsum("bhid,bhjd->bhij", q, k) * c.dim_head ** -0.5
            o = torch.einsum("bhij,bhjd->bhid", dots.softmax(dim=-1), v).transpose(1, 2).reshape(b, n, c.heads * c.dim_head)
            x = x + F.linear(o, sd[a + "1.wrap.to_out.weight"], sd[a + "1.wrap.to_out.bias"])
            h = rms(x, sd[f + "0.0.g"])
            u, gate = F.linear(h, sd[f + "1.wrap.net.0.proj.weight"], sd[f + "1.wrap.net.0.proj.bias"]).chunk(2, dim=-1)
            x = x + F.linear(u * F.gelu(gate), sd[f + "1.wrap.net.3.weight"], sd[f + "1.wrap.net.3.bias"])
        return F.layer_norm(x, (c.dim,), sd[p + "norm.weight"], sd[p + "norm.bias"])

    def latents(self, side, tokens):
        enc = self.encode(side, tokens)
        lat = F.linear(enc.mean(dim=1), self.sd[f"to_{side}_latent.weight"])   # masked_mean with an all-true mask
        return F.normalize(lat, p=2, dim=-1)

    def score(self, text, codes):
        """text [1 or n, T], codes [n, L] -> sim [n] (clvp.py:128-132, return_loss=False)."""
        with torch.inference_mode():
            tl = self.latents("text", text)
            sl = self.latents("speech", codes)
            return (tl * sl).sum(-1) * self.sd["temperature"].exp()
