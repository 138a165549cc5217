"""Generates tests/golden/mel.npz from the REFERENCE's own TorchMelSpectrogram (TEST INFRASTRUCTURE; build container only):
    python -m oracle.make_golden_mel
The module loads the reference's data/models/mel_norms.pth (80 values, stored in the fixture: data, not code)."""
import importlib
import os

import numpy as np
import torch

from . import ref_shim
from .mel_oracle import make_wav

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    ref_shim.load_reference()
    au = importlib.import_module("tortoise_tts.models.arch_utils")
    m = au.TorchMelSpectrogram()
    assert m.mel_norms is not None
    out = {"mel_norms": m.mel_norms.numpy(), "versions": str(dict(torch=torch.__version__))}
    import torchaudio
    out["versions"] = str(dict(torch=torch.__version__, torchaudio=torchaudio.__version__))
    for name, b, n, seed in (("short", 2, 9000, 51), ("clip", 1, 132300, 52)):       # 132300 = cond_length of emb/mel.py:50
        wav = make_wav(b, n, seed=seed)
        with torch.inference_mode():
            mel = m(wav.unsqueeze(1))
        out[f"{name}_spec"] = str(dict(batch=b, samples=n, seed=seed))
        out[f"{name}_mel"] = mel.numpy().astype(np.float32) if name == "short" else mel[:, :, ::8].numpy().astype(np.float32)
        print(name, mel.shape, float(mel.min()), float(mel.max()))
    np.savez_compressed(os.path.join(OUT, "mel.npz"), **out)


if __name__ == "__main__":
    main()
