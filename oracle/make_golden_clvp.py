"""Generates tests/golden/clvp.npz from the REFERENCE's own CLVP class (TEST INFRASTRUCTURE; build container only):
    python -m oracle.make_golden_clvp
The call is the one of reference inference.py:392-396: clvp(text_tokens.repeat(n, 1), codes, return_loss=False), then top-k.
Two cases on synthetic bf16-representable weights (oracle/clvp_oracle.py make_clvp_state_dict):
  d2    depth 2, 5 candidates x 37 codes, 11 text tokens    (layer-level pin: also stores the encoder outputs)
  d20   depth 20 (the reference's size), 16 candidates x 120 codes, 40 text tokens
"""
import os

import numpy as np
import torch

from . import ref_shim
from .clvp_oracle import ClvpConfig, make_clvp_inputs, make_clvp_state_dict

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    out = {}
    for name, depth, n, T, L, w_seed, in_seed in (("d2", 2, 5, 11, 37, 11, 12), ("d20", 20, 16, 40, 120, 13, 14)):
        cfg = ClvpConfig(depth=depth)
        sd = make_clvp_state_dict(cfg, seed=w_seed)
        text, codes = make_clvp_inputs(cfg, n, T, L, seed=in_seed)
        m = ref_shim.build_reference_clvp(sd, cfg)
        with torch.inference_mode():
            scores = m(text.repeat(n, 1), codes, return_loss=False)
            enc_speech = m.speech_transformer(m.speech_emb(codes), mask=torch.ones_like(codes).bool())
            enc_text = m.text_transformer(m.text_emb(text), mask=torch.ones_like(text).bool())
        out[f"{name}_spec"] = str(dict(depth=depth, n=n, text_len=T, speech_len=L, w_seed=w_seed, in_seed=in_seed))
        out[f"{name}_scores"] = scores.numpy()
        out[f"{name}_topk"] = torch.topk(scores, k=min(n, 4)).indices.numpy()
        out[f"{name}_enc_speech_row0"] = enc_speech[0].numpy() if depth == 2 else enc_speech[0, :4].numpy()
        out[f"{name}_enc_text"] = enc_text[0].numpy() if depth == 2 else enc_text[0, :4].numpy()
        print(name, scores.tolist())
    out["versions"] = str(dict(torch=torch.__version__, numpy=np.__version__))
    np.savez_compressed(os.path.join(OUT, "clvp.npz"), **out)


if __name__ == "__main__":
    main()
