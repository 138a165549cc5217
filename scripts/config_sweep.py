"""Per-GPU shares of BASELINE.json's other configs (C3, C4, C5) and the latent re-forward, device-timed on ONE B200.
Not a bench line: evidence for which kernel serves which shape and at what speed.

  C2  1 utterance x 16 candidates, T=64, 500 tokens                 (the bench workload, head-cluster kernel)
  C3  1 utterance x 32 candidates, T=64, 500 tokens                 (256 candidates over 8 GPUs -> 32 per GPU)
  C4  8 utterances x 16 candidates, T ~ U{16..402}, 500 tokens      (64 utterances over 8 GPUs -> 8 per GPU)
  C5  4 utterances x 16 candidates, T ~ U{32..200}, 603 tokens      (32 sentences over 8 GPUs) + latent re-forward
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import __graft_entry__ as ge
    ge.build()
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    from tortoise_tts_b200.synthetic import ARConfig, make_inputs, make_state_dict
    dev = torch.device("cuda", 0)
    cfg = ARConfig()
    sd = make_state_dict(cfg, seed=0)
    only = sys.argv[1:]
    cases = [
        ("C2", 1, 16, [64], 500),
        ("C3_share", 1, 32, [64], 500),
        ("C4_share", 8, 16, None, 500),
        ("C5_share", 4, 16, None, 603),
    ]
    out = []
    for name, U, N, lens, max_new in cases:
        if only and name not in only:
            continue
        g = torch.Generator().manual_seed(100 if name.startswith("C4") else 300)
        if lens is None:
            lo, hi = (16, 402) if name.startswith("C4") else (32, 200)
            lens = torch.randint(lo, hi + 1, (U,), generator=g).tolist()
        Tmax = max(lens)
        lat, texts = make_inputs(cfg, U, lens, seed=1)
        text = torch.zeros(U, Tmax, dtype=torch.long)
        for u, t in enumerate(texts):
            text[u, : len(t)] = t
        m = UnifiedVoiceB200(sd, ARGeometry(), device=dev, max_utts=U, max_seqs=U * N, max_text_len=Tmax,
                             max_new_tokens=min(max_new + 4, 606))
        kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, repetition_penalty=1.0, seed=1234, ignore_eos=True)
        lat_d, text_d = lat.to(dev), text.to(dev)
        rec = {"case": name, "utterances": U, "candidates": N, "text_lens": lens, "max_new": max_new}
        for it in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            codes = m.inference_speech(lat_d, text_d, num_return_sequences=N, max_generate_length=max_new,
                                       text_lengths=lens, **kw)
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
            st = m.last_stats
        assert codes.shape == (U * N, max_new), codes.shape
        dec_ms = st.last_device_ms - st.last_prefill_ms
        rec.update(device_ms=round(st.last_device_ms, 2), prefill_ms=round(st.last_prefill_ms, 2),
                   us_per_step=round(dec_ms / (max_new - 1) * 1e3, 1), tokens_per_s=round(U * N * max_new / (st.last_device_ms * 1e-3), 1),
                   wall_ms=round(wall * 1e3, 2), algorithmic_gbs=round(st.last_algorithmic_bytes / (dec_ms * 1e-3) / 1e9, 1))
        if name.startswith("C5") or name == "C2":
            # latent re-forward over the sampled codes (reference inference.py:371-379): dense, tensor-bound
            M = max_new
            wav = torch.full((U * N,), M * 1024, dtype=torch.long)
            lat_rep = lat_d.repeat_interleave(N, 0)
            text_rep = text_d.repeat_interleave(N, 0)
            tl = torch.tensor(lens).repeat_interleave(N)
            for it in range(2):
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                o = m.forward(lat_rep, text_rep, tl, codes.clamp(max=8191), wav, return_latent=True, clip_inputs=False)
                e1.record()
                torch.cuda.synchronize()
            rows = sum(text.shape[1] + 3 + M + 2 for _ in range(U * N))
            flops = 2.0 * 377_886_720 * rows + 4.0 * 30 * 1024 * sum((text.shape[1] + 3 + M + 2) ** 2 / 2 for _ in range(U * N))
            ms = e0.elapsed_time(e1)
            rec.update(reforward_ms=round(ms, 2), reforward_rows=rows, reforward_tflops=round(flops / (ms * 1e-3) / 1e12, 2))
            assert o.shape == (U * N, M, 1024)
        m.close()
        del m
        torch.cuda.empty_cache()
        print(json.dumps(rec), flush=True)
        out.append(rec)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "config_sweep.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
