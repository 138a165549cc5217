"""Writes tests/golden/fixup.npz by EXECUTING the reference's own source lines (run in the build container, where
/root/reference is mounted):  python -m oracle.make_golden_fixup

The epilogue is inline in TTS.inference (tortoise_tts/inference.py), which cannot be imported here (h5py, soundfile,
librosa are missing: SURVEY.md 8c), so the two loops are sliced out of the source file by their line anchors,
dedented and exec'd unchanged on torch tensors with a stand-in for `autoregressive` that carries stop_mel_token."""
import os
import re
import textwrap
import types

import numpy as np
import torch

REF = "/root/reference/tortoise_tts/inference.py"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _slice(lines, start_pat, end_pat):
    a = next(i for i, l in enumerate(lines) if re.search(start_pat, l))
    b = next(i for i in range(a + 1, len(lines)) if re.search(end_pat, lines[i]))
    return textwrap.dedent("".join(lines[a:b]))


def main():
    lines = open(REF).readlines()
    fix_src = _slice(lines, r"for i, code in enumerate\( codes \):", r"wav_lengths = torch\.tensor")
    trim_src = _slice(lines, r"^\s*calm_tokens = 0\s*$", r"# clvp pass")
    rng = np.random.default_rng(11)
    cases, outs, trims = [], [], []
    specs = [(4, 40, [5, 17, 39, 36]), (3, 12, [0, 3, 11]), (16, 500, None), (2, 64, [40, 50]), (1, 30, [12])]
    for rows, n, stops in specs:
        codes = rng.integers(0, 8192, size=(rows, n)).astype(np.int64)
        codes[:, rng.integers(0, n, size=3)] = 83                       # some calm tokens in the body
        if stops is None:
            stops = rng.integers(20, n, size=rows).tolist()
        for r, s in enumerate(stops):
            codes[r, s:] = 8193                                          # Q7: rows are padded with the stop token
        t = torch.from_numpy(codes.copy())
        env = {"codes": t, "autoregressive": types.SimpleNamespace(stop_mel_token=8193), "torch": torch}
        exec(fix_src, env)
        fixed = env["codes"].numpy().copy()
        # the trim loop slices `latents`; give it a [rows, len, 1] tensor and read the cut off its shape
        env2 = {"codes": torch.from_numpy(fixed.copy()), "calm_token": 83, "latents": torch.zeros(rows, n, 1)}
        exec("for _ in [0]:\n" + textwrap.indent(trim_src, "    "), env2)
        cases.append(codes)
        outs.append(fixed)
        trims.append(env2["latents"].shape[1])
    out = {"n_cases": np.array(len(cases)), "trims": np.array(trims, dtype=np.int64)}
    for i, (c, o) in enumerate(zip(cases, outs)):
        out[f"in_{i}"] = c
        out[f"out_{i}"] = o
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fixup.npz"), **out)
    print("wrote tests/golden/fixup.npz:", [c.shape for c in cases], "trims", trims)


if __name__ == "__main__":
    main()
