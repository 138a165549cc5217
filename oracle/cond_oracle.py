"""CPU oracle of the conditioning-latent encoder -- TEST INFRASTRUCTURE, never on a product path.

fp32 restatement of UnifiedVoice.get_conditioning (reference tortoise_tts/models/unified_voice.py:535-542):
  ConditioningEncoder.forward      unified_voice.py:269-293   init 1x1 conv (80 -> 1024), 6 AttentionBlocks, h[:, :, 0]
  AttentionBlock._forward          tortoise_tts/models/arch_utils.py:183-190  x + proj_out(attention(qkv(norm(x))))
  normalization -> GroupNorm32     arch_utils.py:24-44         32 groups, eps 1e-5, statistics over (32 channels x T), fp32
  QKVAttentionLegacy.forward       arch_utils.py:59-95         channels = [head][q | k | v][64]; q, k each scaled by 64^-0.25;
                                                               softmax in fp32 over all T positions; no mask, no rel_pos
  mean over the clips              unified_voice.py:538-541
Pinned by tests/test_oracle_golden.py against tests/golden/cond.npz (oracle/make_golden_cond.py: the reference's own classes).
"""
import torch
import torch.nn.functional as F

from tortoise_tts_b200.synthetic import make_cond_mel, make_cond_state_dict  # noqa: F401  (pure data generation)


class OracleConditioning:
    def __init__(self, sd, heads=16, blocks=6):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.heads, self.blocks = heads, blocks

    def encode(self, mel):
        """mel [b, 80, T] -> [b, 1024] (ConditioningEncoder.forward with mean=False)."""
        sd, p = self.sd, "conditioning_encoder."
        h = F.conv1d(mel.float(), sd[p + "init.weight"], sd[p + "init.bias"])
        b, c, T = h.shape
        ch = c // self.heads
        for i in range(self.blocks):
            a = f"{p}attn.{i}."
            qkv = F.conv1d(F.group_norm(h, 32, sd[a + "norm.weight"], sd[a + "norm.bias"], eps=1e-5), sd[a + "qkv.weight"], sd[a + "qkv.bias"])
            q, k, v = qkv.reshape(b * self.heads, ch * 3, T).split(ch, dim=1)
            scale = ch ** -0.25
            w = torch.einsum("bct,bcs->bts", q * scale, k * scale).softmax(dim=-1)
            att = torch.einsum("bts,bcs->bct", w, v).reshape(b, -1, T)
            h = h + F.conv1d(att, sd[a + "proj_out.weight"], sd[a + "proj_out.bias"])
        return h[:, :, 0]

    def get_conditioning(self, x):
        """[b, n_clips, 80, T] or [b, 80, T] -> [b, 1024]."""
        with torch.inference_mode():
            x = x.unsqueeze(1) if x.dim() == 3 else x
            return torch.stack([self.encode(x[:, j]) for j in range(x.shape[1])], dim=1).mean(dim=1)
