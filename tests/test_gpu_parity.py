"""GPU (-m gpu): the CUDA path, called through the C ABI (ctypes, tortoise_tts_b200/_lib.py), against
 (a) golden vectors produced by the REFERENCE's own classes (tests/golden/*.npz, oracle/make_golden.py) and
 (b) the CPU oracle (oracle/uv_oracle.py, pinned to (a) by tests/test_oracle_golden.py) on the same seeded
     inputs.
Tolerances (BASELINE.json north_star): logits within 2e-2 relative (max |diff| / max |ref|) in bf16;
KV-cache indexing, stop-token handling and every integer output bit-exact.
"""
import ast
import ctypes as C
import os

import numpy as np
import pytest
import torch

from oracle import philox
from oracle.uv_oracle import (OracleAR, apply_repetition_penalty, apply_temperature, apply_top_k, apply_top_p,
                              apply_typical)
from oracle.weights import ARConfig, make_inputs, make_state_dict

pytestmark = pytest.mark.gpu

BF16_REL = 2e-2          # north_star: "prefill and step logits within 2e-2 relative error in bf16"
STRESS_REL = 4e-2        # ref_l2.npz only: the reference holds UNROUNDED fp32 weights there, so the bf16 snapshot of sharpened (attn_std 0.08)
                         # weights is part of the error; ref_l2_bf16w.npz (identical weights) is held to BF16_REL


PARITY = {}              # measured errors / agreements of this run -> parity_r02.json (gpurun_out/ on the GPU box)


def record(name, **kw):
    PARITY.setdefault(name, {}).update({k: (float(v) if isinstance(v, (int, float, np.floating)) else v) for k, v in kw.items()})


@pytest.fixture(scope="module", autouse=True)
def _write_parity_record():
    yield
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out_dir = os.path.join(root, "gpurun_out") if os.path.isdir(os.path.join(root, "gpurun_out")) else os.path.join(root, "profiles")
    try:
        with open(os.path.join(out_dir, "parity_r02.json"), "w") as f:
            json.dump({"bars": {"bf16_rel": BF16_REL, "greedy_identical": 0.99}, "measured": PARITY}, f, indent=1, sort_keys=True)
    except OSError:
        pass


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def b200():
    import __graft_entry__ as g
    g.build()
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    return UnifiedVoiceB200, ARGeometry


def bf16_weights(sd):
    """bf16-representable weights: the fp32 oracle and the bf16 CUDA path then hold IDENTICAL weights."""
    return {k: v.bfloat16().float() for k, v in sd.items()}


@pytest.fixture(scope="module", params=["ref_l2_bf16w.npz", "ref_l2.npz"])
def l2_case(golden_dir, request):
    g = np.load(os.path.join(golden_dir, request.param))
    spec = ast.literal_eval(str(g["spec"]))
    cfg = ARConfig(layers=spec["layers"])
    sd = make_state_dict(cfg, seed=spec["w_seed"], attn_std=spec["attn_std"])
    tol = STRESS_REL
    if spec.get("bf16_weights"):
        sd = bf16_weights(sd)
        tol = BF16_REL
    lat, texts = make_inputs(cfg, spec["n_utt"], [spec["text_len"]] * spec["n_utt"], seed=spec["in_seed"])
    return g, spec, cfg, sd, lat, torch.stack(texts), tol, request.param


# ---- against the reference's golden vectors ------------------------------------------------------------
def test_l2_reference_golden_teacher_forced(b200, l2_case):
    """2 layers, sharpened attention, 2 utterances x 2 candidates, repetition_penalty 2.0: prefill and
    step logits, per-step latents, greedy choices under teacher forcing with the reference's codes."""
    UV, Geo = b200
    g, spec, cfg, sd, lat, text, tol, gname = l2_case
    m = UV(sd, Geo(layers=2), max_utts=2, max_seqs=4, max_text_len=16, max_new_tokens=16)
    ref_codes = torch.from_numpy(g["codes"])
    codes, lats, choice = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=10,
                                             repetition_penalty=2.0, forced_tokens=ref_codes, return_latents=True,
                                             return_choice=True)
    assert torch.equal(codes.cpu(), ref_codes)                       # forced ids are written back bit-exact
    e_lat = rel(lats, g["latents"])
    agree = (choice.cpu().long() == ref_codes).float().mean().item()
    record("l2_golden_teacher_forced/" + gname, latents_rel=e_lat, choice_agreement=agree, tol=tol)
    assert e_lat < tol, e_lat
    assert agree >= 0.9, agree
    # step logits through the streaming entry (tts_ar_step), still teacher-forced
    m._prefill(lat.cuda(), text.cuda(), 2)
    s = m._sampling(dict(do_sample=False, repetition_penalty=2.0))
    errs = [rel(m.logits_debug(), g["logits"][0][::2])]               # prefill: one row per utterance
    for k in range(2):
        out = torch.zeros(4, 1, dtype=torch.int64, device="cuda")
        forced = ref_codes[:, k:k + 1].to("cuda", torch.int32).contiguous()
        m._check(m.lib.tts_ar_step(m.handle, C.byref(s), 1, C.c_void_p(out.data_ptr()), C.c_void_p(0),
                                   C.c_void_p(forced.data_ptr()), C.c_void_p(0), m._stream()), "step")
        m.sync_stats()
        errs.append(rel(m.logits_debug(), g["logits"][k + 1]))
    record("l2_golden_teacher_forced/" + gname, step_logits_rel=max(errs))
    assert max(errs) < tol, errs
    m.close()


def test_l2_reference_golden_latent_reforward(b200, l2_case):
    """UnifiedVoice.forward(return_latent=True) (unified_voice.py:544-600) vs the reference's output."""
    UV, Geo = b200
    g, spec, cfg, sd, lat, text, tol, gname = l2_case
    m = UV(sd, Geo(layers=2), max_utts=2, max_seqs=4, max_text_len=16, max_new_tokens=16)
    codes = torch.from_numpy(g["reforward_codes"])
    B, M = codes.shape
    l2, t2 = lat.repeat_interleave(2, 0), text.repeat_interleave(2, 0)
    out = m.forward(l2.cuda(), t2.cuda(), torch.full((B,), t2.shape[1]), codes.cuda(), torch.full((B,), M * 1024),
                    return_latent=True, clip_inputs=False)
    assert out.shape == (B, M, 1024)
    e = rel(out, g["reforward_latents"])
    record("l2_golden_latent_reforward/" + gname, latents_rel=e, tol=tol)
    assert e < tol, e
    # the reference caller's own shapes: lengths as [n, 1] (inference.py:371-379 wav_lengths.repeat(candidates, 1))
    out2 = m.forward(l2.cuda(), t2.cuda(), torch.full((B, 1), t2.shape[1]), codes.cuda(), torch.full((B, 1), M * 1024),
                     return_latent=True, clip_inputs=False)
    assert torch.equal(out2, out)
    m.close()


@pytest.fixture(scope="module")
def full_model(b200):
    UV, Geo = b200
    cfg = ARConfig()
    sd = bf16_weights(make_state_dict(cfg, seed=0))
    lat, texts = make_inputs(cfg, 1, [64], seed=1)
    m = UV(sd, Geo(), max_utts=1, max_seqs=16, max_text_len=64, max_new_tokens=520)
    yield m, cfg, lat, torch.stack(texts)
    m.close()


def test_full_c1_reference_golden(full_model, golden_dir):
    """BASELINE config C1 (30 layers, T=64, 1 candidate, greedy, 200 tokens) on IDENTICAL (bf16-representable)
    weights: prefill logits within 2e-2; greedy choices under teacher forcing vs the reference's 200 tokens.
    With random-init weights 20% of the reference's own top-2 gaps are below 1e-2 of the logit scale, so the
    argmax criterion is applied where the reference's gap exceeds the logits tolerance actually achieved."""
    m, cfg, lat, text = full_model
    g = np.load(os.path.join(golden_dir, "ref_full_c1_bf16w.npz"))
    ref = torch.from_numpy(g["codes"][:, :200])
    m._prefill(lat.cuda(), text.cuda(), 1)
    e0 = rel(m.logits_debug(), g["logits_first"][0])
    assert e0 < BF16_REL, e0
    codes, choice, lats = None, None, None
    codes, lats, choice = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=1, max_generate_length=200,
                                             forced_tokens=ref, return_latents=True, return_choice=True, ignore_eos=True)
    assert torch.equal(codes.cpu(), ref)
    same = (choice.cpu().long() == ref)[0].numpy()
    gap = g["top2_vals"][:, 0] - g["top2_vals"][:, 1]
    scale = float(np.abs(g["logits_first"]).max())
    clear = gap > 2 * e0 * scale                      # the reference's own decision is outside our error band
    print(f"C1 teacher-forced agreement {same.mean():.4f} overall, {same[clear].mean():.4f} on {clear.sum()} clear steps; "
          f"prefill logits rel err {e0:.2e}")
    e_lat = rel(lats[:, [0, 1, 199]], g["latents_sel"])
    record("c1_full_golden", prefill_logits_rel=e0, greedy_identical=float(same.mean()), greedy_identical_clear=float(same[clear].mean()),
           latents_rel=e_lat)
    # The same 200 teacher-forced steps through the oracle in bf16 on this GPU (torch eager: what the reference itself
    # would decide in its low-precision mode): the yardstick for "identical up to bf16 near-ties".
    ob = OracleAR(bf16_weights(make_state_dict(cfg, seed=0)), cfg, dtype=torch.bfloat16, device="cuda")
    with torch.inference_mode():
        rb = ob.generate(lat.cuda(), text.cuda(), n_cand=1, max_new=200, forced=ref.cuda(), ignore_eos=True)
    same_bf16 = (rb["choice"].cpu() == ref)[0].numpy()
    print(f"   torch-bf16 oracle on the same GPU agrees with the fp32 reference on {same_bf16.mean():.4f} of the steps")
    record("c1_full_golden", greedy_identical_torch_bf16_oracle=float(same_bf16.mean()))
    # north star: greedy sequences >= 99 % identical.  Every step whose reference top-2 gap exceeds the measured logits
    # error band must agree (asserted at 100 % below); inside the band (15 % of the steps with random-init weights) a
    # flip is a bf16 near-tie that the reference's own bf16 mode breaks differently too, so overall agreement is bounded
    # by rounding luck: it must reach 99 % up to those ties and never fall below the torch-bf16 oracle's own agreement.
    assert same[clear].mean() >= 0.99 and bool(same[clear].all())
    assert same.mean() >= min(0.99, same_bf16.mean()), (same.mean(), same_bf16.mean())
    assert same.mean() >= 0.98
    assert e_lat < BF16_REL


def test_full_c1_original_fp32_weights(b200, golden_dir):
    """Same, with the reference holding unrounded fp32 weights (ref_full_c1.npz): the bf16 snapshot of the weights
    is now part of the error."""
    UV, Geo = b200
    cfg = ARConfig()
    sd = make_state_dict(cfg, seed=0)
    lat, texts = make_inputs(cfg, 1, [64], seed=1)
    text = torch.stack(texts)
    g = np.load(os.path.join(golden_dir, "ref_full_c1.npz"))
    m = UV(sd, Geo(), max_utts=1, max_seqs=1, max_text_len=64, max_new_tokens=8)
    m._prefill(lat.cuda(), text.cuda(), 1)
    assert rel(m.logits_debug(), g["logits_first"][0]) < BF16_REL
    m.close()


# ---- against the oracle on the same seeded inputs -------------------------------------------------------------
def make_small(b200, layers=2, U=2, T=12, attn_std=0.04, **kw):
    UV, Geo = b200
    cfg = ARConfig(layers=layers)
    sd = bf16_weights(make_state_dict(cfg, seed=7, attn_std=attn_std))
    lens = T if isinstance(T, (list, tuple)) else [T] * U
    lat, texts = make_inputs(cfg, U, lens, seed=9)
    m = UV(sd, Geo(layers=layers), max_text_len=max(max(lens), 8), **kw)
    return m, OracleAR(sd, cfg), cfg, lat, texts


def kv_rows(m, layer, page, kv, head, n_pages):
    """Python restatement of layout.h kv_row_offset: pool[layer][head][page][kv][slot 32][64] bf16."""
    pool = m.debug_buffer("kv", ((layer + 1) * n_pages * 2 * 16 * 32 * 64,), dtype=torch.bfloat16)
    off = ((((layer * 16 + head) * n_pages + page) * 2 + kv) * 32) * 64
    return pool[off: off + 32 * 64].view(32, 64).float().cpu()


def test_kv_cache_indexing_and_prefix_sharing(b200):
    """Prompt K rows land in the utterance's shared prefix pages at slot = position % 32; decode K rows of
    candidate c land in ITS private pages; candidates of one utterance share the prompt pages (values vs the
    oracle's DynamicCache-equivalent past, bf16 tolerance; the mapping itself is exact or the values are garbage)."""
    T = 40                                               # P+1 = 44 -> 2 prefix pages per utterance
    m, o, cfg, lat, texts = make_small(b200, U=2, T=T, max_utts=2, max_seqs=6, max_new_tokens=40)
    text = torch.stack(texts)
    N, steps = 3, 36                                     # private segment crosses a page boundary (32)
    ref = o.generate(lat, text, n_cand=N, max_new=steps, ignore_eos=True)
    m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=N, max_generate_length=steps,
                       forced_tokens=ref["codes"], ignore_eos=True)
    # oracle cache for sequence (u=1, c=2): replay to get K of layer 1
    logits, _, past, P = o.prefill(lat, text, N)
    for k in range(steps - 1):
        logits, _, past = o.step(ref["codes"][:, k], k + 1, past)
    K = past[1][0]                                       # layer 1 keys [B, H, L, 64]
    V = past[1][1]
    nqp = (m.cfg.max_new_tokens + 31) // 32              # private pages reserved per sequence by create()
    n_pages = m.last_stats.kv_pages
    plen = T + 4
    pp = (plen + 31) // 32                               # prefix pages actually used per utterance
    seq, head = 1 * N + 2, 5
    # prompt page 1 of utterance 1 = global page pp*1 + 1
    got = kv_rows(m, 1, pp * 1 + 1, 0, head, n_pages)[: plen - 32]
    assert rel(got, K[seq, head, 32:plen]) < 1e-2
    got = kv_rows(m, 1, pp * 1 + 0, 1, head, n_pages)
    assert rel(got, V[seq, head, :32]) < 1e-2
    # private pages start after all prompt pages: sequence s owns [2*pp + s*nqp, ...)
    base = 2 * pp + seq * nqp
    got = kv_rows(m, 1, base, 0, head, n_pages)
    assert rel(got, K[seq, head, plen:plen + 32]) < 1e-2
    got = kv_rows(m, 1, base + 1, 0, head, n_pages)[: steps - 1 - 32]
    assert rel(got, K[seq, head, plen + 32: plen + steps - 1]) < 1e-2
    m.close()


def test_stop_token_handling_bit_exact(b200):
    """Quirk Q7 (stream_generator.py:1169-1190): a row that emits 8193 keeps emitting 8193, the other rows go
    on, the loop ends when every row has finished (rectangular output, early exit) or at max_new."""
    m, o, cfg, lat, texts = make_small(b200, U=2, T=10, max_utts=2, max_seqs=4, max_new_tokens=32)
    text = torch.stack(texts)
    stop = cfg.stop_mel_token
    forced = torch.randint(0, 8192, (4, 20), generator=torch.Generator().manual_seed(3))
    forced[0, 3] = stop
    forced[1, 0] = stop
    forced[2, 11] = stop
    forced[3, 7] = stop
    ref = o.generate(lat, text, n_cand=2, max_new=20, forced=forced)
    got, choice = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=20,
                                     forced_tokens=forced, return_choice=True)
    assert got.shape == ref["codes"].shape == (4, 12)                 # stopped after the last row finished
    assert torch.equal(got.cpu(), ref["codes"])
    assert m.last_stats.all_finished == 1
    for r, k in ((0, 3), (1, 0), (2, 11), (3, 7)):
        assert (got[r, k:] == stop).all()
    # no stop: runs to max_new exactly; ignore_eos: a forced stop does not finish the row
    got2 = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=7,
                              forced_tokens=forced[:, 4:11].contiguous())
    assert got2.shape == (4, 7) and m.last_stats.all_finished == 0
    got3 = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=20,
                              forced_tokens=forced, ignore_eos=True)
    assert got3.shape == (4, 20) and torch.equal(got3.cpu(), forced)
    m.close()


def test_greedy_free_running_matches_oracle(b200):
    """Free-running greedy with repetition_penalty 2.0 (train.py:127): every row is identical to the oracle's up to
    its first divergence, and a divergence may only happen where the oracle's own top-2 gap is inside the bf16
    logits tolerance (a tie the reference itself would break differently in bf16)."""
    UV, Geo = b200
    cfg = ARConfig(layers=2)
    sd = make_state_dict(cfg, seed=7, attn_std=0.04)
    sd["mel_head.weight"] = sd["mel_head.weight"] * 8.0
    sd = bf16_weights(sd)
    lat, texts = make_inputs(cfg, 2, [12, 12], seed=9)
    text = torch.stack(texts)
    o = OracleAR(sd, cfg)
    steps = 48
    ref = o.generate(lat, text, n_cand=2, max_new=steps, repetition_penalty=2.0, want_logits=steps)
    m = UV(sd, Geo(layers=2), max_utts=2, max_seqs=4, max_text_len=16, max_new_tokens=64)
    got = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=steps,
                             repetition_penalty=2.0).cpu()
    ids = torch.full((4, 16), 1, dtype=torch.long)
    ids[:, -1] = cfg.start_mel_token
    n_same = 0
    for r in range(4):
        diff = (got[r] != ref["codes"][r]).nonzero()
        k = int(diff[0]) if len(diff) else steps
        n_same += k
        if k < steps:        # the oracle's processed scores at step k (history identical up to k)
            hist = torch.cat([ids[r], ref["codes"][r, :k]])[None]
            sc = apply_repetition_penalty(ref["logits"][k][r][None].clone(), hist, 2.0)[0]
            top2 = torch.topk(sc, 2).values
            gap = (top2[0] - top2[1]).item() / ref["logits"][k][r].abs().max().item()
            assert gap < BF16_REL, (r, k, gap)
            assert int(got[r, k]) in torch.topk(sc, 2).indices.tolist()
    print("free-running greedy: identical prefix fraction", n_same / (4 * steps))
    record("greedy_free_running_l2", identical_prefix_fraction=n_same / (4 * steps))
    assert n_same / (4 * steps) >= 0.5     # every divergence is asserted above to sit at a reference near-tie
    m.close()


def test_ragged_batch_equals_single_utterance_runs(b200):
    """Quirk Q8 contract: a ragged batch (each utterance at its true text length, paged KV) gives, per
    utterance, what the oracle gives for that utterance alone."""
    lens = [5, 33, 18]
    m, o, cfg, lat, texts = make_small(b200, U=3, T=lens, max_utts=3, max_seqs=6, max_new_tokens=16)
    Tm = max(lens)
    text = torch.zeros(3, Tm, dtype=torch.long)
    for u, t in enumerate(texts):
        text[u, : len(t)] = t
    forced = torch.randint(0, 8192, (6, 6), generator=torch.Generator().manual_seed(5))
    got, lats = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=6,
                                   text_lengths=lens, forced_tokens=forced, return_latents=True)
    for u in range(3):
        ref = o.generate(lat[u:u + 1], texts[u][None], n_cand=2, max_new=6, forced=forced[2 * u: 2 * u + 2])
        assert rel(lats[2 * u: 2 * u + 2], ref["latents"]) < BF16_REL, u
    m.close()


@pytest.mark.parametrize("kw", [dict(top_p=0.8, temperature=0.8, top_k=0), dict(top_k=16, temperature=1.3, top_p=1.0),
                                dict(top_k=50, top_p=0.5, temperature=0.7, repetition_penalty=2.0),
                                # edges of the count-terminated threshold search: a one-token support, a nucleus of a
                                # few tokens on a peaked row (first probe round ends it), an almost complete nucleus on
                                # a flat row (most rounds), k close to the vocabulary size
                                dict(top_k=1, temperature=1.0, top_p=1.0), dict(top_k=0, top_p=0.02, temperature=1.0),
                                dict(top_k=50, top_p=0.8, temperature=0.02), dict(top_k=0, top_p=0.999, temperature=3.0, min_checked=0.5),
                                dict(top_k=8000, top_p=0.9, temperature=1.0, repetition_penalty=1.5, min_checked=0.5)])
def test_device_sampler_equals_oracle_on_device_logits(b200, kw):
    """The fused warpers + Philox inverse-CDF multinomial vs the oracle's HF-equivalent processors
    (pinned to HF's own classes by tests/test_oracle_golden.py) applied to the DEVICE's logits: identical
    tokens except where u*Z is within fp32 summation-order distance of a CDF edge."""
    kw = dict(kw)
    # a flat row that keeps ~8000 tokens has CDF steps of 1.2e-4: a third of the draws fall within the 1e-5 skip margin
    min_checked = kw.pop("min_checked", 0.85)
    m, o, cfg, lat, texts = make_small(b200, U=1, T=12, max_utts=1, max_seqs=24, max_new_tokens=16)
    text = torch.stack(texts)
    N, seed, base = 24, 4321, 1000
    m._prefill(lat.cuda(), text.cuda(), N)
    s = m._sampling(dict(do_sample=True, seed=seed, seq_index_base=base, **kw))
    ids = torch.full((N, text.shape[1] + 4), 1, dtype=torch.long)
    ids[:, -1] = cfg.start_mel_token
    n_checked = 0
    for step in range(6):
        lg = m.logits_debug().cpu()
        if step == 0:
            lg = lg.repeat_interleave(N, 0)
        sc = apply_repetition_penalty(lg.clone(), ids, kw.get("repetition_penalty", 1.0))
        sc = apply_top_p(apply_top_k(apply_temperature(sc, kw["temperature"]), kw["top_k"]), kw["top_p"])
        out = torch.zeros(N, 1, dtype=torch.int64, device="cuda")
        m._check(m.lib.tts_ar_step(m.handle, C.byref(s), 1, C.c_void_p(out.data_ptr()), C.c_void_p(0), C.c_void_p(0),
                                   C.c_void_p(0), m._stream()), "step")
        m.sync_stats()
        out = out.cpu()[:, 0]
        for r in range(N):
            tok, margin = philox.multinomial_row(sc[r].numpy(), philox.uniform(seed, base + r, step))
            assert torch.isfinite(sc[r, out[r]]), "device sampled a token outside the HF support set"
            if margin > 1e-5:
                assert int(out[r]) == tok, (step, r, int(out[r]), tok, margin)
                n_checked += 1
        ids = torch.cat([ids, out[:, None]], 1)
    assert n_checked > min_checked * 6 * N
    m.close()


def test_sampling_is_invariant_to_sharding(b200):
    """Philox key = (seed, GLOBAL sequence index, step): 8 candidates in one call == two calls of 4 with
    seq_index_base 0 and 4 (what two ranks would run)."""
    m, o, cfg, lat, texts = make_small(b200, U=1, T=12, max_utts=1, max_seqs=8, max_new_tokens=16)
    text = torch.stack(texts)
    kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, seed=77, max_generate_length=12, ignore_eos=True)
    whole = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=8, **kw).cpu()
    a = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=4, seq_index_base=0, **kw).cpu()
    b = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=4, seq_index_base=4, **kw).cpu()
    assert torch.equal(whole, torch.cat([a, b]))
    again = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=8, **kw).cpu()
    assert torch.equal(whole, again)                    # deterministic
    m.close()


def test_head_cluster_kernel_agrees_with_interpreter_kernel(b200):
    """<= 16 sequences decode in ar_decode_kernel (32 clusters of 4 CTAs); more, or TTS_AR_DECODE=interp, in the
    interpreter kernel.  Same weights, same teacher-forced tokens: per-step latents agree to bf16-path noise, the
    sampler picks are the same wherever the oracle margin is not a near-tie, and the pad/stop bookkeeping is identical.
    Also covers a batch of 7 (partly filled 16-row tile, second row-half cluster half empty)."""
    kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, seed=11, max_generate_length=14, ignore_eos=True)
    outs = {}
    for mode in ("cluster", "interp"):
        m, o, cfg, lat, texts = make_small(b200, U=1, T=12, max_utts=1, max_seqs=7, max_new_tokens=16)
        m.set_option("decode_kernel", 1 if mode == "interp" else 0)
        text = torch.stack(texts)
        free = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=7, **kw).cpu()
        forced = outs["cluster"][0] if mode == "interp" else free
        codes, lats, choice = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=7, return_latents=True,
                                                 return_choice=True, forced_tokens=forced, **kw)
        outs[mode] = (free, codes.cpu(), lats.cpu(), choice.cpu())
        m.close()
    a, b = outs["cluster"], outs["interp"]
    assert torch.equal(a[1], b[1])                                   # forced tokens echo back identically
    assert rel(a[2], b[2]) < 5e-3                                    # same math, different summation trees
    assert torch.equal(a[3].long(), a[0])                            # teacher forcing with its own picks reproduces them
    # the sampler's own picks under identical forced history: they differ only where u*Z is within the two kernels'
    # logit difference (~1e-3) of a CDF edge -- frequent on this near-flat 8194-way toy distribution (p ~ 1e-4 each)
    assert (a[3] == b[3]).float().mean() > 0.6


def test_streaming_generator_matches_batch_path(b200):
    m, o, cfg, lat, texts = make_small(b200, U=1, T=12, max_utts=1, max_seqs=2, max_new_tokens=40)
    text = torch.stack(texts)
    ids = m.compute_embeddings(lat.cuda(), text.cuda())
    assert ids.shape == (1, 16) and int(ids[0, -1]) == cfg.start_mel_token and (ids[0, :-1] == 1).all()
    kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, seed=5)
    toks, lts = [], []
    for t, l in m.get_generator(ids, max_length=16 + 20, num_return_sequences=2, chunk=8, **kw):   # Q9: total length
        toks.append(t.cpu())
        lts.append(l.cpu())
    toks = torch.stack(toks, 1)
    assert toks.shape == (2, 20)
    whole, lats = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=2, max_generate_length=20,
                                     return_latents=True, **kw)
    assert torch.equal(toks, whole.cpu())
    assert torch.equal(torch.stack(lts, 1), lats.cpu())
    # stop tokens end the stream early (rows pad with the stop token until the last one has finished, Q7), an abandoned
    # generator releases the stream, and a second stream on the same handle works
    g = m.get_generator(ids, max_length=16 + 20, num_return_sequences=2, **kw)
    first = next(g)
    assert torch.equal(first[0].cpu(), whole[:, 0].cpu())
    g.close()
    toks2 = torch.stack([t.cpu() for t, _ in m.get_generator(ids, max_length=16 + 6, num_return_sequences=2, **kw)], 1)
    assert torch.equal(toks2, whole[:, :6].cpu())
    m.close()


def test_host_buffer_entry_and_errors(b200):
    m, o, cfg, lat, texts = make_small(b200, U=1, T=12, max_utts=1, max_seqs=4, max_new_tokens=16)
    text = torch.stack(texts)
    kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, seed=5, ignore_eos=True)
    a = m.inference_speech_host(lat, text, num_return_sequences=4, max_generate_length=10, **kw)
    b = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=4, max_generate_length=10, **kw)
    assert a.device.type == "cpu" and torch.equal(a, b.cpu())
    with pytest.raises(ValueError):
        m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=5, max_generate_length=4)    # capacity
    with pytest.raises(ValueError):
        m.inference_speech(lat.cuda(), text.cuda(), do_sample=True, temperature=0.0)
    with pytest.raises(NotImplementedError):
        m.inference_speech(lat.cuda(), text.cuda(), num_beams=4)
    m.close()


def test_c2_shape_properties(full_model):
    """BASELINE config C2 at full size (16 candidates, top-p 0.8, 500 tokens): size-independent properties --
    determinism, candidates differ, every sampled id is a valid mel code, a sequence prefix does not depend on how
    many steps are requested, and the watchdog never fires."""
    m, cfg, lat, text = full_model
    kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, seed=1234, ignore_eos=True)
    a = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=16, max_generate_length=500, **kw).cpu()
    st = m.last_stats
    assert a.shape == (16, 500) and st.watchdog == 0 and st.steps_run == 500
    assert int(a.min()) >= 0 and int(a.max()) < cfg.number_mel_codes
    assert len({tuple(r.tolist()) for r in a}) == 16
    b = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=16, max_generate_length=120, **kw).cpu()
    assert torch.equal(a[:, :120], b)


def test_fixup_codes_epilogue_bit_exact(b200, golden_dir):
    """the stop-token fix-up + calm-token trim as a device epilogue (tts_ar_fixup_codes) against the
    vectors produced by exec'ing the reference's own loops (tests/golden/fixup.npz), against the oracle on random rows
    at BASELINE sizes (16 x 500, 256 x 604), on a strided view, and through its size-independent properties."""
    from oracle.fixup import fixup_codes, calm_trim
    m, o, cfg, lat, texts = make_small(b200, U=1, T=8, max_utts=1, max_seqs=2, max_new_tokens=8)
    g = np.load(os.path.join(golden_dir, "fixup.npz"))
    for i in range(int(g["n_cases"])):
        codes = torch.from_numpy(g[f"in_{i}"]).cuda()
        out, trim = m.fix_autoregressive_codes(codes)
        assert out.data_ptr() == codes.data_ptr()                        # in place, like the reference
        assert torch.equal(out.cpu(), torch.from_numpy(g[f"out_{i}"])), i
        assert trim == int(g["trims"][i]), i
    rng = np.random.default_rng(5)
    for rows, n in [(16, 500), (256, 604), (1, 3), (7, 33)]:
        codes = rng.integers(0, 8192, size=(rows, n)).astype(np.int64)
        codes[rng.random(size=codes.shape) < 0.05] = 83
        for r in range(rows):
            codes[r, int(rng.integers(0, n)):] = 8193
        ref, first = fixup_codes(codes)
        wide = torch.full((rows, n + 13), -7, dtype=torch.int64, device="cuda")   # strided view: stride > len
        view = wide[:, :n]
        view.copy_(torch.from_numpy(codes))
        out, trim = m.fix_autoregressive_codes(view)
        assert torch.equal(out.cpu(), torch.from_numpy(ref))
        assert trim == calm_trim(ref)
        assert bool((wide[:, n:] == -7).all())                           # nothing written past the row
        # properties: no stop token survives, so a second pass finds nothing to fix (strict would raise)
        assert not bool((out == 8193).any())
        again, trim2 = m.fix_autoregressive_codes(out.clone(), strict=False)
        assert torch.equal(again, out) and trim2 == trim
        with pytest.raises(RuntimeError):
            m.fix_autoregressive_codes(out.clone())
    with pytest.raises(ValueError):
        m.fix_autoregressive_codes(torch.zeros(2, 8, dtype=torch.int32, device="cuda"))
    with pytest.raises(ValueError):                                      # len < 3: BAD_ARG from the C ABI
        m.fix_autoregressive_codes(torch.full((1, 2), 8193, dtype=torch.int64, device="cuda"))
    m.close()


def typical_row_is_stable(scores, mass, eps=3e-6):
    """True when no token sits within fp32 distance of the typical threshold (shifted value or mass), so that the
    reference's support set does not depend on how log_softmax was rounded.  float64."""
    s = scores.double().numpy()
    e = np.exp(s - s.max())
    pr = e / e.sum()
    cT = (pr * s).sum()
    sh = np.abs(cT - s)
    order = np.argsort(sh, kind="stable")
    cs = np.cumsum(pr[order])
    j = int((cs < mass).sum())                    # index of the threshold token in sorted order
    if j >= len(cs) - 1:
        return False
    lo = cs[j - 1] if j > 0 else 0.0
    gap_next = sh[order[j + 1]] - sh[order[j]]
    gap_prev = sh[order[j]] - sh[order[j - 1]] if j > 0 else 1.0
    return min(mass - lo, cs[j] - mass) > eps and min(gap_next, gap_prev) > eps * max(1.0, sh[order[j]])


@pytest.mark.parametrize("kw", [dict(typical_mass=0.9, temperature=1.0, top_k=0, top_p=1.0),
                                dict(typical_mass=0.5, temperature=0.8, top_k=0, top_p=0.8, repetition_penalty=2.0),
                                dict(typical_mass=0.2, temperature=1.3, top_k=40, top_p=1.0),
                                dict(typical_mass=0.99, temperature=0.7, top_k=50, top_p=0.9)])
def test_typical_sampling_equals_reference_warper_on_device_logits(b200, kw):
    """typical_sampling=True (reference unified_voice.py:47-75, :659): repetition penalty -> TypicalLogitsWarper ->
    temperature -> top-k -> top-p -> multinomial on the DEVICE's logits vs the oracle chain (apply_typical is pinned
    to the reference's own class by tests/test_oracle_golden.py).  Rows with a token inside fp32 distance of the
    typical threshold are skipped (the reference's own answer there depends on log_softmax rounding)."""
    kw = dict(kw)
    mass = kw["typical_mass"]
    m, o, cfg, lat, texts = make_small(b200, U=1, T=12, max_utts=1, max_seqs=24, max_new_tokens=16)
    text = torch.stack(texts)
    N, seed, base = 24, 99, 500
    m._prefill(lat.cuda(), text.cuda(), N)
    s = m._sampling(dict(do_sample=True, seed=seed, seq_index_base=base, **kw))
    ids = torch.full((N, text.shape[1] + 4), 1, dtype=torch.long)
    ids[:, -1] = cfg.start_mel_token
    n_checked = n_rows = 0
    for step in range(6):
        lg = m.logits_debug().cpu()
        if step == 0:
            lg = lg.repeat_interleave(N, 0)
        pre = apply_repetition_penalty(lg.clone(), ids, kw.get("repetition_penalty", 1.0))
        sc = apply_typical(pre.clone(), mass)
        sc = apply_top_p(apply_top_k(apply_temperature(sc, kw["temperature"]), kw["top_k"]), kw["top_p"])
        out = torch.zeros(N, 1, dtype=torch.int64, device="cuda")
        m._check(m.lib.tts_ar_step(m.handle, C.byref(s), 1, C.c_void_p(out.data_ptr()), C.c_void_p(0), C.c_void_p(0),
                                   C.c_void_p(0), m._stream()), "step")
        m.sync_stats()
        out = out.cpu()[:, 0]
        for r in range(N):
            n_rows += 1
            if not typical_row_is_stable(pre[r], mass):
                continue
            tok, margin = philox.multinomial_row(sc[r].numpy(), philox.uniform(seed, base + r, step))
            assert torch.isfinite(sc[r, out[r]]), "device sampled a token outside the reference's typical support set"
            if margin > 1e-5:
                assert int(out[r]) == tok, (step, r, int(out[r]), tok, margin)
                n_checked += 1
        ids = torch.cat([ids, out[:, None]], 1)
    print("typical: rows checked", n_checked, "of", n_rows)
    assert n_checked > 0.5 * n_rows
    m.close()


def test_typical_sampling_greedy_and_cluster_kernel(b200):
    """The typical warper is a logits processor, so it also filters the greedy argmax (the most likely token is NOT
    always in the typical set).  Teacher-forced, 16 sequences -> the head-cluster kernel; the recorded choices equal
    the oracle's argmax over the typical-filtered scores wherever the oracle's own top-2 gap exceeds the bf16 logits
    tolerance."""
    UV, Geo = b200
    cfg = ARConfig(layers=2)
    sd = make_state_dict(cfg, seed=7, attn_std=0.04)
    sd["mel_head.weight"] = sd["mel_head.weight"] * 8.0
    sd = bf16_weights(sd)
    lat, texts = make_inputs(cfg, 1, [12], seed=9)
    text = torch.stack(texts)
    o = OracleAR(sd, cfg)
    steps, N, mass = 12, 16, 0.3
    ref = o.generate(lat, text, n_cand=N, max_new=steps, typical_mass=mass, repetition_penalty=2.0, want_logits=steps,
                     ignore_eos=True)
    m = UV(sd, Geo(layers=2), max_utts=1, max_seqs=N, max_text_len=16, max_new_tokens=32)
    got, choice = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=N, max_generate_length=steps,
                                     typical_sampling=True, typical_mass=mass, repetition_penalty=2.0,
                                     forced_tokens=ref["codes"], return_choice=True, ignore_eos=True)
    assert torch.equal(got.cpu(), ref["codes"])
    ids = torch.full((N, 16), 1, dtype=torch.long)
    ids[:, -1] = cfg.start_mel_token
    n_ok = n_all = n_not_mode = 0
    for k in range(steps):
        hist = torch.cat([ids, ref["codes"][:, :k]], 1)
        pre = apply_repetition_penalty(ref["logits"][k].clone(), hist, 2.0)
        sc = apply_typical(pre.clone(), mass)
        for r in range(N):
            n_all += 1
            if not typical_row_is_stable(pre[r], mass, eps=2e-2):      # bf16 logits move the threshold too
                continue
            top2 = torch.topk(sc[r], 2).values
            if (top2[0] - top2[1]).item() / ref["logits"][k][r].abs().max().item() < BF16_REL:
                continue
            assert int(choice[r, k]) == int(sc[r].argmax()), (k, r)
            n_ok += 1
            n_not_mode += int(sc[r].argmax() != pre[r].argmax())
    print("typical greedy: checked", n_ok, "of", n_all, "; argmax outside the typical set in", n_not_mode)
    assert n_ok >= 0.3 * n_all
    m.close()


@pytest.mark.parametrize("N", [24, 32, 40, 64, 100])
def test_wide_launches_equal_16_sequence_launches_bit_exact(b200, N):
    """17..32 / 33..64 sequences run in ONE launch of the 32- / 64-row head-cluster kernel (csrc/ar_decode.cu MODE 2 / 4:
    every streamed weight chunk is multiplied with all rows), more than 64 as consecutive launches over groups of 64.  A
    row's arithmetic (k-split, accumulator chains, attention quarters, reduction orders) does not depend on the MODE, the
    sequences are independent and the Philox key is the global sequence index, so one N-candidate call must equal, bit
    for bit, the 16-candidate calls (MODE 1, the kernel every oracle test above pins) that cover the same global
    indices -- codes, per-step latents and early termination included (stop tokens honoured: launches finish at
    different steps, the rectangle is padded with the stop token, Q7)."""
    UV, Geo = b200
    cfg = ARConfig(layers=2)
    sd = make_state_dict(cfg, seed=7, attn_std=0.04)
    sd["mel_head.weight"] = sd["mel_head.weight"] * 8.0
    sd["mel_head.bias"] = sd["mel_head.bias"].clone()
    sd["mel_head.bias"][cfg.stop_mel_token] = 14.0           # stop tokens do occur
    sd = bf16_weights(sd)
    lat, texts = make_inputs(cfg, 1, [12], seed=9)
    text = torch.stack(texts)
    m = UV(sd, Geo(layers=2), max_utts=1, max_seqs=N, max_text_len=16, max_new_tokens=48)
    kw = dict(do_sample=True, top_p=0.9, temperature=1.0, top_k=0, seed=5, max_generate_length=40, return_latents=True)
    whole, wl = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=N, **kw)
    whole, wl = whole.cpu(), wl.cpu()
    n = whole.shape[1]
    stop = cfg.stop_mel_token
    steps = []
    for s0 in range(0, N, 16):
        cnt = min(16, N - s0)
        a, al = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=cnt, seq_index_base=s0, **kw)
        a, al = a.cpu(), al.cpu()
        k = a.shape[1]
        steps.append(k)
        assert k <= n
        assert torch.equal(whole[s0:s0 + cnt, :k], a), s0
        assert bool((whole[s0:s0 + cnt, k:] == stop).all())
        assert torch.equal(wl[s0:s0 + cnt, :k], al), s0
    assert n == max(steps)
    print(f"N={N}: steps {n}, 16-sequence launches ran {steps}, stops {int((whole == stop).any(1).sum())}")
    if N <= 32:   # the host-buffer entry pads the columns an early launch never wrote
        h = m.inference_speech_host(lat, text, num_return_sequences=N, do_sample=True, top_p=0.9, temperature=1.0, top_k=0, seed=5,
                                    max_generate_length=40)
        assert torch.equal(h, whole)
    m.close()


def test_dense_pass_tiled_attention_matches_oracle_and_row_path(b200):
    """Prompt-shaped passes with >= 1184 rows run their causal attention as 16-row mma tiles (csrc/ar_attn.cuh
    attn_tile16; below that one decode-shaped unit per (row, head)).  The latent re-forward (unified_voice.py:544-600)
    of 4 sequences x 317 positions crosses the threshold: vs the fp32 oracle at the bf16 tolerance, and vs the row path
    (TTS_AR_DBG=2) of the same library.  317 is not a multiple of 16, so tiles that straddle two sequences (the per-row
    fallback inside the tiled path) are exercised as well."""
    UV, Geo = b200
    cfg = ARConfig(layers=2)
    sd = bf16_weights(make_state_dict(cfg, seed=7, attn_std=0.04))
    lat, texts = make_inputs(cfg, 4, [12] * 4, seed=9)
    text = torch.stack(texts)
    M = 300
    g = torch.Generator().manual_seed(5)
    codes = torch.randint(0, 8192, (4, M), generator=g)
    o = OracleAR(sd, cfg)
    want = o.forward_latents(lat, text, codes)
    m = UV(sd, Geo(layers=2), max_utts=4, max_seqs=4, max_text_len=16, max_new_tokens=M + 4)
    args = (lat.cuda(), text.cuda(), torch.full((4,), 12), codes.cuda(), torch.full((4,), M * 1024))
    m.set_option("attn_tiles", 2)
    tiled = m.forward(*args, return_latent=True, clip_inputs=False).cpu()
    m.set_option("attn_tiles", 1)
    rows = m.forward(*args, return_latent=True, clip_inputs=False).cpu()
    m.set_option("attn_tiles", 0)
    m.set_option("dense_kernel", 1)          # the interpreter kernel in prompt mode (fallback path) agrees with the tcgen05 pipeline
    interp = m.forward(*args, return_latent=True, clip_inputs=False).cpu()
    m.set_option("dense_kernel", 0)
    e_i = rel(interp, want)
    record("dense_pass_4x317", tiled_vs_oracle=rel(tiled, want), rows_vs_oracle=rel(rows, want), interpreter_vs_oracle=e_i)
    assert e_i < BF16_REL
    assert tiled.shape == (4, M, 1024)
    e_t, e_r, e_tr = rel(tiled, want), rel(rows, want), rel(tiled, rows)
    print(f"dense attention: tiled vs oracle {e_t:.3e}, row path vs oracle {e_r:.3e}, tiled vs row path {e_tr:.3e}")
    assert e_t < BF16_REL and e_r < BF16_REL and e_tr < BF16_REL
    assert not torch.equal(tiled, rows)          # the two paths really are different code
    m.close()


# ---- the tcgen05 / TMEM GEMM of the dense passes, alone ------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (300, 384, 1024, 128), (1000, 3072, 1024, 256),
                                      (257, 8194, 1024, 128), (130, 1024, 4096, 256), (68, 4096, 1024, 128)])
def test_tc_gemm_matches_torch(b200, M, N, K, bn):
    """csrc/tc_gemm.cuh (tcgen05.mma, TMA-staged 128-byte-swizzled operands, TMEM accumulators) through the C-ABI parity
    hook: out = A . Wt^T + bias with bf16 operands and fp32 accumulation vs torch fp32 on the same bf16 values.  Covers
    ragged M (TMA zero fill, masked rows), ragged N (mel head: 8194 columns), both tile widths, K = 64 .. 4096."""
    m, o, cfg, lat, texts = make_small(b200, U=1, T=8, max_utts=1, max_seqs=2, max_new_tokens=8)
    g = torch.Generator().manual_seed(M * 7 + N)
    a = torch.randn(M, K, generator=g).bfloat16().cuda()
    wt = (torch.randn(N, K, generator=g) * 0.05).bfloat16().cuda()
    bias = torch.randn(N, generator=g).cuda()
    out = torch.full((M, N), float("nan"), device="cuda")
    m._check(m.lib.tts_ar_debug_gemm(m.handle, C.c_void_p(a.data_ptr()), C.c_void_p(wt.data_ptr()), C.c_void_p(bias.data_ptr()),
                                     C.c_void_p(out.data_ptr()), M, N, K, bn, m._stream()), "tts_ar_debug_gemm")
    m.sync_stats()
    want = a.float() @ wt.float().t() + bias
    e = rel(out, want)
    record(f"tc_gemm/{M}x{N}x{K}/bn{bn}", rel=e)
    assert torch.isfinite(out).all()
    assert e < 1e-4, e                       # fp32 accumulation of exact bf16 products: summation-order noise only
    m.close()


# ---- BASELINE shapes against the oracle ------------------------------------------------------
def test_c2_full_size_teacher_forced_vs_oracle(full_model):
    """BASELINE config C2 at FULL size (30 layers, 16 candidates of one 64-token utterance), teacher-forced for 8 steps
    with a fixed token matrix: per-step latents and step logits within 2e-2 of the fp32 oracle on identical
    (bf16-representable) weights, the sampler's greedy choice equal to the oracle's wherever the oracle's top-2 gap is
    outside the measured error band."""
    m, cfg, lat, text = full_model
    steps, N = 8, 16
    forced = torch.randint(0, 8192, (N, steps), generator=torch.Generator().manual_seed(21))
    sd = bf16_weights(make_state_dict(cfg, seed=0))
    o = OracleAR(sd, cfg)
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.inference_mode():
        ref = o.generate(lat, text, n_cand=N, max_new=steps, forced=forced, want_logits=steps, ignore_eos=True)
    codes, lats, choice = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=N, max_generate_length=steps,
                                             forced_tokens=forced, return_latents=True, return_choice=True, ignore_eos=True)
    assert torch.equal(codes.cpu(), forced)
    e_lat = rel(lats, ref["latents"])
    # step logits of the last step through the parity hook (state after `steps - 1` fed tokens)
    m._prefill(lat.cuda(), text.cuda(), N)
    s = m._sampling(dict(do_sample=False, ignore_eos=True))
    errs = [rel(m.logits_debug(), ref["logits"][0][:1])]
    for k in range(3):
        out = torch.zeros(N, 1, dtype=torch.int64, device="cuda")
        f = forced[:, k:k + 1].to("cuda", torch.int32).contiguous()
        m._check(m.lib.tts_ar_step(m.handle, C.byref(s), 1, C.c_void_p(out.data_ptr()), C.c_void_p(0), C.c_void_p(f.data_ptr()),
                                   C.c_void_p(0), m._stream()), "step")
        m.sync_stats()
        errs.append(rel(m.logits_debug(), ref["logits"][k + 1]))
    e_log = max(errs)
    scale = ref["logits"].abs().amax(dim=(1, 2))                       # per step
    top2 = torch.topk(ref["logits"], 2, dim=-1).values
    gap = (top2[..., 0] - top2[..., 1]) / scale[:, None]               # [steps, N]
    clear = (gap > 2 * e_log).t()
    same = choice.cpu().long() == ref["choice"]
    record("c2_full_size_teacher_forced", latents_rel=e_lat, step_logits_rel=e_log, choice_agreement=float(same.float().mean()),
           choice_agreement_clear=float(same[clear].float().mean()), clear_fraction=float(clear.float().mean()))
    print(f"C2 full size: latents {e_lat:.2e}, step logits {e_log:.2e}, choices {same.float().mean():.3f} ({same[clear].float().mean():.3f} clear)")
    assert e_lat < BF16_REL and e_log < BF16_REL
    assert bool(same[clear].all())


@pytest.mark.parametrize("U,N", [(3, 16), (4, 16), (8, 16), (1, 24), (1, 32), (2, 16)])
def test_many_sequences_ragged_vs_oracle(b200, U, N):
    """More than 16 sequences (the shapes the C3 / C4 / C5 per-GPU shares run: 24, 32, 48, 64, 128), ragged text lengths,
    44 teacher-forced steps so private KV pages cross a page boundary and long rows are split over several attention
    units; rows finish at different steps (forced stop tokens).  Quirk Q8 contract: per utterance, the latents equal the
    oracle's run on that utterance ALONE (<= 2e-2), codes / padding / termination bit-exact."""
    steps = 44
    rng = np.random.default_rng(U * 100 + N)
    lens = [int(x) for x in rng.integers(5, 41, size=U)]
    B = U * N
    m, o, cfg, lat, texts = make_small(b200, U=U, T=lens, max_utts=U, max_seqs=B, max_new_tokens=48)
    Tm = max(lens)
    text = torch.zeros(U, Tm, dtype=torch.long)
    for u, t in enumerate(texts):
        text[u, : len(t)] = t
    forced = torch.from_numpy(rng.integers(0, 8192, size=(B, steps))).long()
    stop = cfg.stop_mel_token
    stops = {}
    for r in range(0, B, 3):                                         # every third row stops somewhere in [4, steps)
        k = int(rng.integers(4, steps))
        forced[r, k] = stop
        stops[r] = k
    got, lats = m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=N, max_generate_length=steps,
                                   text_lengths=lens, forced_tokens=forced, return_latents=True)
    got, lats = got.cpu(), lats.cpu()
    worst = 0.0
    for u in range(U):
        ref = o.generate(lat[u:u + 1], texts[u][None], n_cand=N, max_new=steps, forced=forced[N * u: N * u + N])
        n = ref["codes"].shape[1]                                    # the oracle stops when ITS rows have all finished (never: some run on)
        assert torch.equal(got[N * u: N * u + N, :n], ref["codes"]), u
        for r in range(N):
            k = stops.get(N * u + r, steps)
            valid = min(k + 1, n)                                    # latents of steps up to and including the stop step are defined
            e = rel(lats[N * u + r, :valid], ref["latents"][r, :valid])
            worst = max(worst, e)
    for r, k in stops.items():
        assert (got[r, k:] == stop).all()                            # pad after stop (Q7)
    record(f"ragged_vs_oracle/U{U}xN{N}", latents_rel=worst, sequences=B, steps=steps)
    print(f"U={U} N={N}: worst per-row latents rel err {worst:.2e}")
    assert worst < BF16_REL
    m.close()


# ---- the drop-in boundary: patch_reference_module on a reference-shaped module ----------------------------------------
def test_patch_reference_module_with_reference_call_sites(b200):
    """`patch_reference_module` on a module with the reference's attribute names / state_dict keys (tests/standin.py),
    called with exactly the kwargs of the reference's callers: inference.py:334-346 (inference_speech), :266-285
    (compute_embeddings + get_generator), :371-379 (forward(return_latent=True) with repeated [n, 1] lengths) and
    train.py:119-129; results equal a UnifiedVoiceB200 built directly from the same weights.  A LoRA parametrization
    (reference lora.py:163-195) on c_attn must change the result exactly as merged weights do."""
    import torch.nn.utils.parametrize as parametrize
    from tests.standin import LoRADelta, StandInUnifiedVoice
    from tortoise_tts_b200 import patch_reference_module
    UV, Geo = b200
    cfg = ARConfig(layers=2)
    sd = bf16_weights(make_state_dict(cfg, seed=7, attn_std=0.04))
    lat, texts = make_inputs(cfg, 1, [12], seed=9)
    text = torch.stack(texts)
    mod = StandInUnifiedVoice(cfg, sd).cuda()
    patch_reference_module(mod, max_utts=1, max_seqs=4, max_text_len=16, max_new_tokens=40)
    direct = UV(sd, Geo(layers=2), max_utts=1, max_seqs=4, max_text_len=16, max_new_tokens=40)
    candidates = 2
    kw = dict(do_sample=True, top_k=50, top_p=0.8, temperature=0.8, num_return_sequences=candidates, num_beams=1,
              length_penalty=1.0, repetition_penalty=2.0, max_generate_length=24)
    codes = mod.inference_speech(lat.cuda(), text.cuda(), **kw)                                   # inference.py:334-346
    want = direct.inference_speech(lat.cuda(), text.cuda(), **kw)
    assert codes.dtype == torch.int64 and codes.is_cuda and torch.equal(codes, want)
    codes[0, -1] = 83                                                                             # the caller mutates the result in place (inference.py:360-366)
    # forward(return_latent=True) with the caller's repeated [n, 1] lengths                       # inference.py:371-379
    text_lengths = torch.Tensor([text.shape[1]]).to(dtype=torch.int32)
    wav_lengths = torch.tensor([codes.shape[-1] * mod.mel_length_compression], device="cuda")
    l1 = mod.forward(lat.cuda().repeat(candidates, 1), text.cuda().repeat(candidates, 1), text_lengths.repeat(candidates, 1), codes,
                     wav_lengths.repeat(candidates, 1), return_latent=True, clip_inputs=False)
    l2 = direct.forward(lat.cuda().repeat(candidates, 1), text.cuda().repeat(candidates, 1), torch.full((candidates,), text.shape[1]),
                        codes, torch.full((candidates,), codes.shape[-1] * 1024), return_latent=True, clip_inputs=False)
    assert l1.shape == (candidates, codes.shape[1], 1024) and torch.equal(l1, l2)
    assert mod.forward(1, 2) == "reference forward"                                               # every other forward() stays the reference's
    # streaming call site                                                                         # inference.py:266-285
    inputs = mod.compute_embeddings(lat.cuda(), text.cuda())
    gen = mod.get_generator(inputs=inputs, top_k=50, top_p=0.8, temperature=0.8, do_sample=True, num_beams=1, num_return_sequences=1,
                            length_penalty=1.0, repetition_penalty=2.0, output_attentions=False, output_hidden_states=True)
    toks = []
    for c, latent in gen:
        assert c.shape == (1,) and latent.shape == (1, 1024)
        toks.append(int(c[0]))
        if len(toks) == 12:
            break
    one = direct.inference_speech(lat.cuda(), text.cuda(), do_sample=True, top_k=50, top_p=0.8, temperature=0.8, repetition_penalty=2.0,
                                  max_generate_length=12)
    assert toks == one[0, : len(toks)].tolist()
    # LoRA: parametrized c_attn of layer 1 -> re-patch -> equals a model built from the merged weights
    layer = mod.gpt.h[1].attn.c_attn
    parametrize.register_parametrization(layer, "weight", LoRADelta(tuple(layer.weight.shape), seed=3).cuda())
    mod._b200.load_state_dict(UV.module_weights(mod))                                             # "refresh after LoRA toggles" (lora.py:197-205)
    sd2 = dict(sd)
    lo = layer.parametrizations.weight[0]
    sd2["gpt.h.1.attn.c_attn.weight"] = sd["gpt.h.1.attn.c_attn.weight"] + ((lo.lora_B @ lo.lora_A).detach() * lo.scaling).cpu()
    merged = UV(sd2, Geo(layers=2), max_utts=1, max_seqs=4, max_text_len=16, max_new_tokens=40)
    kw2 = dict(do_sample=False, num_return_sequences=1, repetition_penalty=2.0, max_generate_length=16)   # train.py:119-129 (greedy eval)
    a = mod.inference_speech(lat.cuda(), text.cuda(), **kw2)
    b = merged.inference_speech(lat.cuda(), text.cuda(), **kw2)
    c = direct.inference_speech(lat.cuda(), text.cuda(), **kw2)
    assert torch.equal(a, b)
    assert not torch.equal(a, c)                      # the delta is really in: the base model decodes something else
    for x in (mod._b200, direct, merged):
        x.close()
