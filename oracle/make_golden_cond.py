"""Generates tests/golden/cond.npz from the REFERENCE's own UnifiedVoice.get_conditioning (TEST INFRASTRUCTURE; build
container only):  python -m oracle.make_golden_cond
Synthetic `conditioning_encoder.*` weights (tortoise_tts_b200/synthetic.py make_cond_state_dict) loaded into the reference's
UnifiedVoice; inputs shaped like emb/mel.py:92-97 builds them ([1, n_clips, 80, frames])."""
import os

import numpy as np
import torch

from . import ref_shim
from .weights import ARConfig
from .cond_oracle import make_cond_mel, make_cond_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    uv = ref_shim.load_reference()
    cfg = ARConfig(layers=1)
    m = uv.UnifiedVoice(layers=1, model_dim=cfg.model_dim, heads=cfg.heads, max_text_tokens=cfg.max_text_tokens,
                        max_mel_tokens=cfg.max_mel_tokens, max_conditioning_inputs=cfg.max_conditioning_inputs,
                        number_text_tokens=cfg.number_text_tokens, start_text_token=cfg.start_text_token,
                        number_mel_codes=cfg.number_mel_codes, start_mel_token=cfg.start_mel_token,
                        stop_mel_token=cfg.stop_mel_token, checkpointing=False)
    out = {}
    for name, w_seed, in_seed, b, clips, frames in (("one_clip", 31, 32, 1, 1, 150), ("two_clips", 33, 34, 2, 2, 77)):
        sd = make_cond_state_dict(seed=w_seed)
        missing, unexpected = m.load_state_dict(sd, strict=False)
        assert not unexpected and not any(k.startswith("conditioning_encoder.") for k in missing)
        m.eval()
        mel = make_cond_mel(b, clips, frames, seed=in_seed)
        with torch.inference_mode():
            lat = m.get_conditioning(mel)
        out[f"{name}_spec"] = str(dict(w_seed=w_seed, in_seed=in_seed, batch=b, clips=clips, frames=frames))
        out[f"{name}_latent"] = lat.numpy()
        print(name, lat.shape, lat.abs().max().item(), lat[0, :4].tolist())
    out["versions"] = str(dict(torch=torch.__version__, numpy=np.__version__))
    np.savez_compressed(os.path.join(OUT, "cond.npz"), **out)


if __name__ == "__main__":
    main()
