"""Generates tests/golden/*.npz from the REFERENCE's own classes (TEST INFRASTRUCTURE; runs only in
the build container where /root/reference is mounted):  python -m oracle.make_golden

The reference has no tests or golden vectors for the AR path, so these
fixtures ARE the pin: reference UnifiedVoice + GPT2InferenceModel (imported by oracle/ref_shim.py)
with HF transformers' GPT-2 blocks and logits processors, fed the synthetic weights / inputs of
oracle/weights.py.  Versions are recorded in each file.  tests/test_oracle_golden.py checks the
restatement (oracle/uv_oracle.py) against them on CPU; tests/test_gpu_parity.py checks the CUDA
path against them on the B200.
"""
import os
import numpy as np
import torch

from .weights import ARConfig, make_state_dict, make_inputs
from . import ref_shim

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def versions():
    import transformers
    return dict(torch=torch.__version__, transformers=transformers.__version__, numpy=np.__version__)


def case_l2(bf16_weights=False):
    """2 layers at full width, sharpened attention, 2 utterances x 2 candidates, greedy with
    repetition_penalty 2.0 (train.py:127), plus the latent re-forward on the generated codes.
    bf16_weights=True -> ref_l2_bf16w.npz: every weight rounded to a bf16-representable value first (the reference
    still computes in fp32), so the reference and the bf16 CUDA path hold IDENTICAL weights and the north-star
    tolerance (2e-2) applies undiluted; ref_l2.npz keeps the unrounded fp32 weights (the bf16 snapshot of the
    weights is then part of the error)."""
    cfg = ARConfig(layers=2)
    spec = dict(layers=2, w_seed=3, attn_std=0.08, in_seed=5, n_utt=2, text_len=12, n_cand=2,
                max_new=10, repetition_penalty=2.0, bf16_weights=bf16_weights)
    sd = make_state_dict(cfg, seed=spec["w_seed"], attn_std=spec["attn_std"])
    if bf16_weights:
        sd = {k: v.bfloat16().float() for k, v in sd.items()}
    lat, texts = make_inputs(cfg, spec["n_utt"], [spec["text_len"]] * spec["n_utt"], seed=spec["in_seed"])
    text = torch.stack(texts)
    m = ref_shim.build_reference_model(sd, cfg)
    r = ref_shim.ref_generate(m, lat, text, n_cand=2, max_new=10, want_logits=3, repetition_penalty=2.0)
    codes = r["codes"][:, :8].clone()
    B = codes.shape[0]
    l2, t2 = lat.repeat_interleave(2, 0), text.repeat_interleave(2, 0)
    with torch.inference_mode():
        rl = m(l2, t2, torch.full((B,), t2.shape[1]), codes.clone(),
               torch.full((B,), codes.shape[1] * 1024), return_latent=True, clip_inputs=False)
    np.savez_compressed(os.path.join(OUT, "ref_l2_bf16w.npz" if bf16_weights else "ref_l2.npz"), spec=str(spec), versions=str(versions()),
                        codes=r["codes"].numpy(), logits=r["logits"].numpy(),
                        latents=r["latents"].numpy(), pos_trace=np.array(r["pos_trace"]),
                        prefix_len=r["prefix_len"], reforward_codes=codes.numpy(),
                        reforward_latents=rl.float().numpy())
    print("ref_l2 done", r["codes"][0].tolist())


def case_full(bf16_weights=False):
    """BASELINE config C1: 30 layers fp32, one utterance T=64, 1 candidate, greedy, 200 tokens,
    stop token ignored so exactly 200 steps run.  bf16_weights=True: the same run with every weight
    rounded to a bf16-representable value first (still fp32 arithmetic in the reference), i.e. the
    reference and the bf16 CUDA path hold IDENTICAL weights -> ref_full_c1_bf16w.npz."""
    cfg = ARConfig()
    spec = dict(layers=30, w_seed=0, w_std=0.02, in_seed=1, n_utt=1, text_len=64, n_cand=1, max_new=200,
                ignore_eos=True, bf16_weights=bf16_weights)
    sd = make_state_dict(cfg, seed=0)
    if bf16_weights:
        sd = {k: v.bfloat16().float() for k, v in sd.items()}
    lat, texts = make_inputs(cfg, 1, [64], seed=1)
    text = torch.stack(texts)
    m = ref_shim.build_reference_model(sd, cfg)
    r = ref_shim.ref_generate(m, lat, text, n_cand=1, max_new=200, want_logits=200, ignore_eos=True)
    lg = r["logits"]                                   # [200, 1, V]
    top2 = torch.topk(lg[:, 0], 2, dim=-1)
    name = "ref_full_c1_bf16w.npz" if bf16_weights else "ref_full_c1.npz"
    np.savez_compressed(os.path.join(OUT, name), spec=str(spec), versions=str(versions()),
                        codes=r["codes"].numpy(), logits_first=lg[:2].numpy(), logits_last=lg[-1:].numpy(),
                        top2_vals=top2.values.numpy(), top2_idx=top2.indices.numpy(),
                        latents_sel=r["latents"][:, [0, 1, 199]].numpy(),
                        pos_trace=np.array(r["pos_trace"]), prefix_len=r["prefix_len"])
    print(name, "done", r["codes"][0, :12].tolist())


def case_warpers():
    """HF's own processor classes on fixed random logits: the support sets the sampler must match."""
    from transformers.generation.logits_process import (
        RepetitionPenaltyLogitsProcessor, TemperatureLogitsWarper, TopKLogitsWarper, TopPLogitsWarper)
    g = torch.Generator().manual_seed(11)
    logits = torch.randn(6, 8194, generator=g) * 2.5
    ids = torch.randint(0, 8194, (6, 40), generator=g)
    ids[:, 0] = 1
    ids[:, 1] = 8192
    out = dict(logits=logits.numpy(), ids=ids.numpy(), versions=str(versions()))
    out["rep2"] = RepetitionPenaltyLogitsProcessor(penalty=2.0)(ids, logits.clone()).numpy()
    out["temp08"] = TemperatureLogitsWarper(0.8)(ids, logits.clone()).numpy()
    out["topk16"] = TopKLogitsWarper(top_k=16, min_tokens_to_keep=1)(ids, logits.clone()).numpy()
    for p in (0.8, 0.5, 0.95):
        out[f"topp{int(p * 100)}"] = TopPLogitsWarper(top_p=p, min_tokens_to_keep=1)(ids, logits.clone()).numpy()
    chain = TemperatureLogitsWarper(0.8)(ids, RepetitionPenaltyLogitsProcessor(penalty=2.0)(ids, logits.clone()))
    chain = TopPLogitsWarper(top_p=0.8, min_tokens_to_keep=1)(ids, TopKLogitsWarper(top_k=50, min_tokens_to_keep=1)(ids, chain))
    out["chain"] = chain.numpy()
    out = {k: (np.isfinite(v) if k.startswith(("topk", "topp", "chain")) and k != "chain" else v) for k, v in out.items()}
    np.savez_compressed(os.path.join(OUT, "hf_warpers.npz"), **out)
    print("hf_warpers done")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    case_warpers()
    case_l2()
    case_l2(bf16_weights=True)
    case_full()
    case_full(bf16_weights=True)
