"""Golden vectors for the typical-sampling warper: the REFERENCE'S OWN class
(`TypicalLogitsWarper`, /root/reference/tortoise_tts/models/unified_voice.py:47-75, imported by path through
oracle/ref_shim.py) run in this container on fixed random logits.  TEST INFRASTRUCTURE; writes
tests/golden/typical.npz.  Rows: flat (std 1), normal (std 2.5), peaked (std 6), one with a dominant token, one
with -inf entries (what a preceding filter leaves)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_shim import load_reference  # noqa: E402


def main():
    uv = load_reference()
    g = torch.Generator().manual_seed(21)
    rows = [torch.randn(8194, generator=g) * s for s in (1.0, 2.5, 2.5, 6.0, 6.0)]
    dom = torch.randn(8194, generator=g)
    dom[4000] = 14.0
    rows.append(dom)
    holes = torch.randn(8194, generator=g) * 2.5
    holes[torch.rand(8194, generator=g) < 0.3] = float("-inf")
    rows.append(holes)
    logits = torch.stack(rows)
    ids = torch.zeros(len(rows), 1, dtype=torch.long)
    out = dict(logits=logits.numpy(), torch=torch.__version__)
    for mass in (0.9, 0.5, 0.2, 0.99):
        w = uv.TypicalLogitsWarper(mass=mass)
        out[f"keep{int(mass * 100)}"] = torch.isfinite(w(ids, logits.clone())).numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "typical.npz"), **out)
    print({k: (v.sum(1).tolist() if k.startswith("keep") else None) for k, v in out.items() if k.startswith("keep")})


if __name__ == "__main__":
    main()
