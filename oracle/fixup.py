"""CPU restatement (test infrastructure only -- nothing under tortoise_tts_b200/ may import this) of the epilogue the
reference runs on the sampled mel codes between the AR sampler and its consumers:

  * stop-token fix-up, tortoise_tts/inference.py:353-366 (identical loop in train.py:135-147)
  * calm-token trim of the latents, inference.py:381-389 (train.py:161-169)

Pinned by tests/golden/fixup.npz, which oracle/make_golden_fixup.py produced by EXECUTING those reference source
lines in the build container (tests/test_oracle_golden.py)."""
import numpy as np


def fixup_codes(codes, stop_token=8193, calm_token=83, strict=True):
    """codes: int64 [rows, len] (modified copy returned).  Follows inference.py:353-366 line by line."""
    codes = np.array(codes, dtype=np.int64, copy=True)
    first = np.full(codes.shape[0], -1, dtype=np.int64)
    for i in range(codes.shape[0]):                                  # :353
        idx = np.nonzero(codes[i] == stop_token)[0]                  # :354
        if len(idx) == 0:                                            # :355 .min() of an empty tensor raises (quirk Q10)
            if strict:
                raise RuntimeError("row without a stop token")
            continue                                                 # :357-358 what the reference meant to do
        stm = int(idx.min())                                         # :355
        first[i] = stm
        codes[i][idx] = calm_token                                   # :360
        codes[i][stm:] = calm_token                                  # :361
        if stm - 3 < codes[i].shape[0]:                              # :363 (always true)
            codes[i][-3] = 45                                        # :364
            codes[i][-2] = 45                                        # :365
            codes[i][-1] = 248                                       # :366
    return codes, first


def calm_trim(codes, calm_token=83):
    """Index at which inference.py:381-389 cuts the latents (uses row 0 only); codes.shape[-1] if it never does."""
    calm_tokens = 0                                                  # :381
    for k in range(codes.shape[-1]):                                 # :382
        if codes[0, k] == calm_token:                                # :383
            calm_tokens += 1
        else:
            calm_tokens = 0                                          # :386
        if calm_tokens > 8:                                          # :387
            return k                                                 # :388 latents[:, :k]
    return codes.shape[-1]
