"""CPU: the oracle restatement (oracle/uv_oracle.py) against golden vectors produced by the
reference's own classes (oracle/make_golden.py).  fp32 tolerance 1e-4 relative (BASELINE.json)."""
import ast
import os

import numpy as np
import pytest
import torch

from oracle import philox
from oracle.uv_oracle import (OracleAR, apply_repetition_penalty, apply_temperature, apply_top_k,
                              apply_top_p)
from oracle.weights import ARConfig, make_inputs, make_state_dict

FP32_REL = 1e-4


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    assert philox.philox4x32_10((0, 0, 0, 0), (0, 0)) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert philox.philox4x32_10((0xffffffff,) * 4, (0xffffffff,) * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert philox.philox4x32_10((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    u = philox.uniform(1234, 5, 7)
    assert 0.0 <= float(u) < 1.0


def test_oracle_matches_reference_l2(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_l2.npz"))
    spec = ast.literal_eval(str(g["spec"]))
    cfg = ARConfig(layers=spec["layers"])
    sd = make_state_dict(cfg, seed=spec["w_seed"], attn_std=spec["attn_std"])
    lat, texts = make_inputs(cfg, spec["n_utt"], [spec["text_len"]] * spec["n_utt"], seed=spec["in_seed"])
    text = torch.stack(texts)
    o = OracleAR(sd, cfg)
    r = o.generate(lat, text, n_cand=spec["n_cand"], max_new=spec["max_new"], want_logits=3,
                   repetition_penalty=spec["repetition_penalty"])
    assert np.array_equal(r["codes"].numpy(), g["codes"])                  # greedy ids, bit-exact
    assert r["pos_trace"] == g["pos_trace"].tolist() == [0] + list(range(2, spec["max_new"] + 1))  # Q2
    assert r["prefix_len"] == int(g["prefix_len"]) == spec["text_len"] + 3                     # Q3/Q9
    assert rel(r["logits"].numpy(), g["logits"]) < FP32_REL
    assert rel(r["latents"].numpy(), g["latents"]) < FP32_REL
    codes = torch.from_numpy(g["reforward_codes"])
    lat2, text2 = lat.repeat_interleave(2, 0), text.repeat_interleave(2, 0)
    assert rel(o.forward_latents(lat2, text2, codes).numpy(), g["reforward_latents"]) < FP32_REL


def test_oracle_matches_reference_full_c1_prefix(golden_dir):
    """Config C1 at full depth: the first 12 greedy steps (the whole 200 run in the GPU tests and bench)."""
    g = np.load(os.path.join(golden_dir, "ref_full_c1.npz"))
    cfg = ARConfig()
    sd = make_state_dict(cfg, seed=0)
    lat, texts = make_inputs(cfg, 1, [64], seed=1)
    torch.set_num_threads(os.cpu_count() or 1)
    r = OracleAR(sd, cfg).generate(lat, torch.stack(texts), max_new=12, ignore_eos=True, want_logits=2)
    assert np.array_equal(r["codes"].numpy(), g["codes"][:, :12])
    assert rel(r["logits"].numpy(), g["logits_first"]) < FP32_REL
    assert rel(r["latents"][:, :2].numpy(), g["latents_sel"][:, :2]) < FP32_REL
    assert int(g["prefix_len"]) == 67


def test_warpers_match_hf(golden_dir):
    g = np.load(os.path.join(golden_dir, "hf_warpers.npz"))
    logits, ids = torch.from_numpy(g["logits"]), torch.from_numpy(g["ids"])
    assert torch.equal(apply_repetition_penalty(logits, ids, 2.0), torch.from_numpy(g["rep2"]))
    assert torch.allclose(apply_temperature(logits, 0.8), torch.from_numpy(g["temp08"]), rtol=0, atol=0)
    assert np.array_equal(torch.isfinite(apply_top_k(logits, 16)).numpy(), g["topk16"])
    for p in (0.8, 0.5, 0.95):
        assert np.array_equal(torch.isfinite(apply_top_p(logits, p)).numpy(), g[f"topp{int(p * 100)}"])
    chain = apply_top_p(apply_top_k(apply_temperature(apply_repetition_penalty(logits, ids, 2.0), 0.8), 50), 0.8)
    assert np.array_equal(torch.isfinite(chain).numpy(), np.isfinite(g["chain"]))
    fin = np.isfinite(g["chain"])
    assert np.allclose(chain.numpy()[fin], g["chain"][fin], rtol=1e-6)


def test_top_p_mass_rule_equals_hf_support(golden_dir):
    """The device sampler keeps token i iff (probability mass of strictly larger scores) < top_p.
    That is HF's `cumsum(ascending) <= 1-p -> remove` rule without the sort; check the two support
    sets are identical on the HF golden rows."""
    g = np.load(os.path.join(golden_dir, "hf_warpers.npz"))
    logits = g["logits"].astype(np.float64)
    for p in (0.8, 0.5, 0.95):
        want = g[f"topp{int(p * 100)}"]
        for r in range(logits.shape[0]):
            e = np.exp(logits[r] - logits[r].max())
            pr = e / e.sum()
            order = np.argsort(-logits[r], kind="stable")
            above = np.zeros_like(pr)
            above[order] = np.concatenate([[0.0], np.cumsum(pr[order])[:-1]])
            assert np.array_equal(above < p, want[r])


def test_multinomial_row_is_inverse_cdf():
    rng = np.random.default_rng(0)
    s = rng.standard_normal(50).astype(np.float32)
    s[rng.random(50) < 0.5] = -np.inf
    s[7] = 3.0
    counts = np.zeros(50)
    for i in range(4000):
        tok, _ = philox.multinomial_row(s, philox.uniform(99, i, 0))
        counts[tok] += 1
    assert counts[~np.isfinite(s)].sum() == 0
    e = np.where(np.isfinite(s), np.exp(s - s.max()), 0)
    pr = e / e.sum()
    chi2 = (((counts - 4000 * pr) ** 2) / np.maximum(4000 * pr, 1e-9))[pr > 0].sum()
    assert chi2 < 2.0 * (pr > 0).sum() + 30


def test_fixup_oracle_matches_the_executed_reference_lines(golden_dir):
    """oracle/fixup.py vs tests/golden/fixup.npz, which oracle/make_golden_fixup.py produced by exec'ing the reference's
    own stop-token fix-up and calm-token trim loops (inference.py:353-366, 381-389).  Bit-exact (integer work)."""
    from oracle.fixup import fixup_codes, calm_trim
    g = np.load(os.path.join(golden_dir, "fixup.npz"))
    for i in range(int(g["n_cases"])):
        fixed, first = fixup_codes(g[f"in_{i}"])
        assert np.array_equal(fixed, g[f"out_{i}"]), i
        assert calm_trim(fixed) == int(g["trims"][i]), i
        assert np.array_equal(first, np.argmax(g[f"in_{i}"] == 8193, axis=1)), i
    # a row without a stop token: the reference raises (Q10); non-strict leaves it alone
    row = np.arange(20, dtype=np.int64)[None]
    with pytest.raises(RuntimeError):
        fixup_codes(row)
    fixed, first = fixup_codes(row, strict=False)
    assert np.array_equal(fixed, row) and first[0] == -1 and calm_trim(fixed) == 20


def test_typical_warper_matches_the_reference_class(golden_dir):
    """oracle.apply_typical vs tests/golden/typical.npz = the reference's own TypicalLogitsWarper
    (unified_voice.py:47-75) run by oracle/make_golden_typical.py.  Support sets are bit-exact (same torch ops)."""
    from oracle.uv_oracle import apply_typical
    g = np.load(os.path.join(golden_dir, "typical.npz"))
    logits = torch.from_numpy(g["logits"])
    for mass in (0.9, 0.5, 0.2, 0.99):
        got = torch.isfinite(apply_typical(logits.clone(), mass)).numpy()
        assert np.array_equal(got, g[f"keep{int(mass * 100)}"]), mass
    assert torch.equal(apply_typical(logits, None), logits) and torch.equal(apply_typical(logits, 1.0), logits)


def test_typical_mass_rule_equals_reference_support(golden_dir):
    """The device sampler keeps token i iff (probability mass of tokens with a strictly smaller
    |s_i - E_p[s]|) < mass, with E_p[s] = smax + sum e (s - smax) / sum e: the reference's sort + cumsum rule without
    the sort (csrc/ar_sample.cuh, pass -1).  Identical support sets on the reference's golden rows, evaluated in
    float64; a token may differ only if its shifted value ties the threshold to 1e-6 (fp32 rounding of the
    reference's log_softmax)."""
    g = np.load(os.path.join(golden_dir, "typical.npz"))
    logits = g["logits"].astype(np.float64)
    for mass in (0.9, 0.5, 0.2, 0.99):
        want = g[f"keep{int(mass * 100)}"]
        for r in range(logits.shape[0]):
            s = logits[r]
            fin = np.isfinite(s)
            e = np.where(fin, np.exp(s - s[fin].max()), 0.0)
            cT = s[fin].max() + (e[fin] * (s[fin] - s[fin].max())).sum() / e.sum()
            shifted = np.where(fin, np.abs(cT - np.where(fin, s, 0.0)), np.inf)
            pr = e / e.sum()
            order = np.argsort(shifted, kind="stable")
            below = np.zeros_like(pr)
            cs = np.cumsum(pr[order])
            # mass of strictly smaller shifted values (ties share the same "below")
            sh_sorted = shifted[order]
            first_of_group = np.searchsorted(sh_sorted, sh_sorted, side="left")
            below[order] = np.where(first_of_group > 0, cs[np.maximum(first_of_group - 1, 0)], 0.0)
            got = (below < mass) & fin
            diff = np.nonzero(got != want[r])[0]
            if diff.size:
                thr = shifted[want[r]].max()
                assert np.all(np.abs(shifted[diff] - thr) < 1e-6 * max(1.0, thr)), (mass, r, diff[:5])
                assert diff.size <= 2


# ---- CLVP candidate scoring: oracle/clvp_oracle.py vs the reference's CLVP class ----------------
@pytest.mark.parametrize("case", ["d2", "d20"])
def test_clvp_oracle_matches_the_reference_class(golden_dir, case):
    """tests/golden/clvp.npz holds what the REFERENCE's CLVP.forward(text.repeat(n, 1), codes, return_loss=False)
    (inference.py:392-396 call) returned on synthetic weights; the restatement must reproduce scores, top-k order and the
    encoder outputs to fp32 round-off."""
    import ast
    from oracle.clvp_oracle import ClvpConfig, OracleCLVP, make_clvp_inputs, make_clvp_state_dict
    g = np.load(os.path.join(golden_dir, "clvp.npz"))
    spec = ast.literal_eval(str(g[f"{case}_spec"]))
    cfg = ClvpConfig(depth=spec["depth"])
    sd = make_clvp_state_dict(cfg, seed=spec["w_seed"])
    text, codes = make_clvp_inputs(cfg, spec["n"], spec["text_len"], spec["speech_len"], seed=spec["in_seed"])
    o = OracleCLVP(sd, cfg)
    scores = o.score(text, codes)
    want = torch.from_numpy(g[f"{case}_scores"])
    assert (scores - want).abs().max() < 2e-5, (scores - want).abs().max()
    assert torch.topk(scores, k=min(spec["n"], 4)).indices.tolist() == g[f"{case}_topk"].tolist()
    with torch.inference_mode():
        enc = o.encode("speech", codes[:1])[0]
        enc_t = o.encode("text", text)[0]
    ws, wt = torch.from_numpy(g[f"{case}_enc_speech_row0"]), torch.from_numpy(g[f"{case}_enc_text"])
    assert (enc[: ws.shape[0]] - ws).abs().max() < 2e-4
    assert (enc_t[: wt.shape[0]] - wt).abs().max() < 2e-4


# ---- conditioning-latent encoder: oracle/cond_oracle.py vs the reference's get_conditioning --------
@pytest.mark.parametrize("case", ["one_clip", "two_clips"])
def test_conditioning_oracle_matches_the_reference_class(golden_dir, case):
    import ast
    from oracle.cond_oracle import OracleConditioning, make_cond_mel, make_cond_state_dict
    g = np.load(os.path.join(golden_dir, "cond.npz"))
    spec = ast.literal_eval(str(g[f"{case}_spec"]))
    o = OracleConditioning(make_cond_state_dict(seed=spec["w_seed"]))
    lat = o.get_conditioning(make_cond_mel(spec["batch"], spec["clips"], spec["frames"], seed=spec["in_seed"]))
    want = torch.from_numpy(g[f"{case}_latent"])
    assert lat.shape == want.shape == (spec["batch"], 1024)
    assert rel(lat.numpy(), want.numpy()) < FP32_REL


# ---- mel-spectrogram front end: oracle/mel_oracle.py vs the reference's TorchMelSpectrogram ---------
@pytest.mark.parametrize("case", ["short", "clip"])
def test_mel_oracle_matches_the_reference_class(golden_dir, case):
    import ast
    from oracle.mel_oracle import make_wav, oracle_mel
    g = np.load(os.path.join(golden_dir, "mel.npz"))
    spec = ast.literal_eval(str(g[f"{case}_spec"]))
    mel = oracle_mel(make_wav(spec["batch"], spec["samples"], seed=spec["seed"]), torch.from_numpy(g["mel_norms"]))
    want = torch.from_numpy(g[f"{case}_mel"])
    if case == "clip":
        assert mel.shape[-1] == 517                      # 132300 samples = the reference's conditioning clip (emb/mel.py:50)
        mel = mel[:, :, ::8]
    assert (mel - want).abs().max() < 1e-5
