"""GPU (-m gpu): CLVP candidate scoring through the C ABI (tts_ar_clvp_*, tortoise_tts_b200/clvp.py)
against (a) tests/golden/clvp.npz -- outputs of the REFERENCE's own CLVP class (oracle/make_golden_clvp.py) -- and (b) the CPU
oracle (oracle/clvp_oracle.py, pinned to (a) by tests/test_oracle_golden.py) on the same seeded inputs.

Tolerances: the GEMMs run in bf16 with fp32 accumulation (the bar BASELINE.json's north_star sets for the AR path: 2e-2
relative in bf16): L2-normalised speech latents within 2e-2 (max |diff| / max |ref|); scores are cosine * exp(temperature),
compared at 2e-2 * exp(temperature) absolute, and the top-k order must agree wherever the reference's gaps exceed twice
the measured error.  Integer outputs (none here) would be exact.
"""
import ast
import json
import os

import numpy as np
import pytest
import torch

from oracle.clvp_oracle import ClvpConfig, OracleCLVP, make_clvp_inputs, make_clvp_state_dict

pytestmark = pytest.mark.gpu

BF16_REL = 2e-2
PARITY = {}


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module", autouse=True)
def _write_parity_record():
    yield
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out_dir = os.path.join(root, "gpurun_out") if os.path.isdir(os.path.join(root, "gpurun_out")) else os.path.join(root, "profiles")
    with open(os.path.join(out_dir, "parity_r02_clvp.json"), "w") as f:
        json.dump(PARITY, f, indent=1, sort_keys=True)


@pytest.fixture(scope="module")
def clvp_cls():
    import __graft_entry__ as g
    g.build()
    from tortoise_tts_b200.clvp import CLVPB200
    return CLVPB200


def order_agrees(got, want, band):
    """top-k order of `got` equals that of `want` wherever neighbouring reference scores differ by more than `band`."""
    idx = torch.argsort(want, descending=True)
    g = got[idx]
    w = want[idx]
    for i in range(len(idx) - 1):
        if (w[i] - w[i + 1]) > band and not g[i] > g[i + 1]:
            return False
    return True


@pytest.mark.parametrize("case", ["d2", "d20"])
def test_clvp_scores_match_the_reference_goldens(clvp_cls, golden_dir, case):
    g = np.load(os.path.join(golden_dir, "clvp.npz"))
    spec = ast.literal_eval(str(g[f"{case}_spec"]))
    cfg = ClvpConfig(depth=spec["depth"])
    sd = make_clvp_state_dict(cfg, seed=spec["w_seed"])
    text, codes = make_clvp_inputs(cfg, spec["n"], spec["text_len"], spec["speech_len"], seed=spec["in_seed"])
    m = clvp_cls(sd, depth=cfg.depth, max_seqs=16, max_text_len=64, max_speech_len=128)
    # the reference's call (inference.py:394): text repeated per candidate
    scores = m(text.repeat(spec["n"], 1).cuda(), codes.cuda(), return_loss=False).cpu()
    want = torch.from_numpy(g[f"{case}_scores"])
    temp = float(sd["temperature"].exp())
    err = (scores - want).abs().max().item()
    PARITY[f"golden/{case}"] = dict(score_abs_err=err, score_abs_max=want.abs().max().item(), tol=BF16_REL * temp)
    print(f"clvp {case}: max |score - reference| = {err:.3e} (scores up to {want.abs().max():.3f}, tolerance {BF16_REL * temp:.3e})")
    assert torch.isfinite(scores).all()
    assert err < BF16_REL * temp
    assert order_agrees(scores, want, 2 * err + 1e-6)
    # one text row for every candidate: the same latent, computed once
    s1 = m(text.cuda(), codes.cuda()).cpu()
    assert torch.equal(s1, scores)
    m.close()


def test_clvp_vs_oracle_ragged_shapes_and_distinct_texts(clvp_cls):
    """Lengths that are not multiples of the attention tiles (16 / 64) or the GEMM tiles (128 rows), a different text per
    candidate (n_text == n), against the fp32 oracle: speech latents and scores."""
    cfg = ClvpConfig(depth=3)
    sd = make_clvp_state_dict(cfg, seed=21)
    o = OracleCLVP(sd, cfg)
    m = clvp_cls(sd, depth=3, max_seqs=8, max_text_len=80, max_speech_len=200)
    temp = float(sd["temperature"].exp())
    for n, T, L, seed in ((3, 7, 50, 1), (5, 33, 131, 2), (1, 1, 1, 3), (8, 80, 200, 4)):
        g = torch.Generator().manual_seed(seed)
        text = torch.randint(0, cfg.num_text_tokens, (n, T), generator=g)
        codes = torch.randint(0, cfg.num_speech_tokens, (n, L), generator=g)
        scores, lat = m.forward(text.cuda(), codes.cuda(), return_latents=True)
        want = o.score(text, codes)
        with torch.inference_mode():
            want_lat = o.latents("speech", codes)
        e_s, e_l = (scores.cpu() - want).abs().max().item(), rel(lat, want_lat)
        PARITY[f"oracle/n{n}_T{T}_L{L}"] = dict(score_abs_err=e_s, latent_rel=e_l)
        print(f"clvp n={n} T={T} L={L}: score err {e_s:.3e}, speech latent rel {e_l:.3e}")
        assert e_l < BF16_REL and e_s < BF16_REL * temp
    m.close()


def test_clvp_full_size_candidates_are_scored_independently(clvp_cls):
    """BASELINE config C2's scoring call at full size (depth 20, 16 candidates x 500 codes, 64 text tokens).  Size-independent
    property: a candidate's score does not depend on its position in the batch -- scoring a permutation of the candidates
    returns the permuted scores BIT-exactly -- and two candidates checked against the fp32 oracle."""
    cfg = ClvpConfig()
    sd = make_clvp_state_dict(cfg, seed=5)
    text, codes = make_clvp_inputs(cfg, 16, 64, 500, seed=6)
    m = clvp_cls(sd, depth=cfg.depth, max_seqs=16, max_text_len=64, max_speech_len=500)
    scores = m(text.cuda(), codes.cuda())
    perm = torch.randperm(16, generator=torch.Generator().manual_seed(7))
    s_perm = m(text.cuda(), codes[perm].cuda())
    assert torch.isfinite(scores).all()
    assert torch.equal(s_perm, scores[perm.cuda()])
    for k in (1, 4, 16):                              # inference.py:395: indices = torch.topk(scores, k=candidates).indices
        assert m.topk(scores, k).tolist() == torch.topk(scores, k).indices.tolist()
    with pytest.raises(ValueError):
        m.topk(scores, 17)
    sub = m(text.cuda(), codes[[3, 12]].cuda())
    assert torch.equal(sub, scores[[3, 12]])
    want = OracleCLVP(sd, cfg).score(text, codes[[3, 12]])
    err = (sub.cpu() - want).abs().max().item()
    PARITY["oracle/full_size_2_of_16"] = dict(score_abs_err=err, launches=m.kernel_launches)
    print(f"clvp full size: score err vs oracle {err:.3e}; top-4 {torch.topk(scores, 4).indices.tolist()}")
    assert err < BF16_REL * float(sd["temperature"].exp())
    m.close()


def test_clvp_errors_and_from_module(clvp_cls):
    cfg = ClvpConfig(depth=1)
    sd = make_clvp_state_dict(cfg, seed=2)

    class StandIn:                                       # what CLVPB200.from_module reads of the reference module
        xformers = True

        def state_dict(self):
            return sd

    m = clvp_cls.from_module(StandIn(), max_seqs=4, max_text_len=16, max_speech_len=32)
    assert (m.depth, m.dim, m.heads) == (1, 768, 12)
    text, codes = make_clvp_inputs(cfg, 4, 16, 32, seed=3)
    ok = m(text.cuda(), codes.cuda())
    assert ok.shape == (4,) and torch.isfinite(ok).all()
    with pytest.raises(ValueError):                      # capacity
        m(text.cuda(), torch.zeros(5, 32, dtype=torch.long).cuda())
    with pytest.raises(ValueError):
        m(text.cuda(), torch.zeros(4, 33, dtype=torch.long).cuda())
    bad = codes.clone()
    bad[2, 5] = 8192                                     # the reference's nn.Embedding raises IndexError
    with pytest.raises(ValueError):
        m(text.cuda(), bad.cuda())
    with pytest.raises(NotImplementedError):
        m(text.cuda(), codes.cuda(), return_loss=True)
    again = m(text.cuda(), codes.cuda())                 # the error was reported once and cleared
    assert torch.equal(again, ok)
    m.close()


def test_clvp_bf16_source_weights(clvp_cls):
    """TtsClvpWeights.dtype = 1: a bfloat16 state_dict (the packing kernels' other instantiation).  The GEMM weights are
    bf16-representable already; gains / biases / embeddings are rounded once more, so the scores move by bf16 noise only."""
    cfg = ClvpConfig(depth=2)
    sd = make_clvp_state_dict(cfg, seed=8)
    text, codes = make_clvp_inputs(cfg, 4, 12, 40, seed=9)
    m32 = clvp_cls(sd, depth=2, max_seqs=4, max_text_len=16, max_speech_len=64)
    m16 = clvp_cls({k: v.bfloat16() for k, v in sd.items()}, depth=2, max_seqs=4, max_text_len=16, max_speech_len=64)
    a, b = m32(text.cuda(), codes.cuda()), m16(text.cuda(), codes.cuda())
    assert torch.isfinite(b).all() and (a - b).abs().max().item() < BF16_REL * float(sd["temperature"].exp())
    m32.close()
    m16.close()
