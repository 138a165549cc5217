#!/usr/bin/env python
"""bench.py -- AR mel tokens/s of the UnifiedVoice sampler (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  (N > 1: python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...)

A "step" is one whole `inference_speech` job on one batch of synthetic input: prompt pass + 500 sampled mel
tokens for 16 candidates per GPU (BASELINE config C2: bf16, 16 candidates, top-p 0.8, temperature 0.8, one 64-token
sentence, stop token ignored so exactly 500 steps run -- SURVEY.md 8(d)).  At N GPUs every rank runs 16 more
candidates of the same utterance (C3's sharding: candidates split across GPUs, weak scaling) and the timed region
ends with the NCCL all-gather of the codes.

value : tokens/s, inputs resident in HBM, CUDA events, max over ranks.
e2e   : the same job through the host-buffer C-ABI entry (tts_ar_inference_speech_host): pinned host inputs,
        H2D + prompt pass + sample loop + D2H of the codes inside the timed region.
roofline : the persistent decode kernel (ar_decode_kernel): SURVEY.md 8(d) algorithmic bytes / its device time.
configs  : sub-records for the other BASELINE configs, run after the headline in the same process(es), STRONG scaling
        (the global job is fixed, each rank runs its share): C3 = 256 candidates of one utterance; C4 = 64 ragged
        utterances x 16 candidates, length-balanced utterance shards; C5 = 32 utterances x 16 candidates x 603 tokens,
        then the stop-token fix-up, the latent re-forward (tcgen05 dense pass) and the NCCL all-gather of codes + latents.
reference_same_gpu : the oracle (the reference's math in torch eager, torch.cat KV cache) on THIS GPU in fp32 and bf16,
        bounded sample of the C2 job -- "the reference path on the same GPU" (BASELINE.md section 4).
cpu_baseline : the oracle (CPU restatement of the reference path) on config C1, bounded sample, host cores.
--impl reference : the reference's CPU implementation of the path (oracle port; the reference's own
        generate() does not run on the installed transformers) on the C2 workload, bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_CAND = int(os.environ.get("TTS_BENCH_CANDIDATES", "16"))   # 32 = the per-GPU share of C3 (256 candidates over 8 GPUs); default = C2
TEXT_LEN = 64
MAX_NEW = int(os.environ.get("TTS_BENCH_MAX_NEW", "500"))   # override ONLY for short ncu captures (not a bench value)
SAMPLING = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, repetition_penalty=1.0, seed=1234)
REF_SAMPLE_TOKENS = 24          # --impl reference: decode steps per timed step (bounded sample)
METRIC = "AR mel tokens/sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the oracle legs (cpu_baseline, reference_same_gpu)")
    ap.add_argument("--no-sub-configs", action="store_true", help="headline C2 record only (no C3 / C4 / C5 sub-records)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def oracle_cpu_c1(budget_s=20.0):
    """cpu_baseline leg: the oracle on BASELINE config C1 (fp32, 1 utterance T=64, 1 candidate, greedy), bounded."""
    import torch
    from oracle.uv_oracle import OracleAR
    from tortoise_tts_b200.synthetic import ARConfig, make_inputs, make_state_dict
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = ARConfig()
    sd = make_state_dict(cfg, seed=0)
    lat, texts = make_inputs(cfg, 1, [TEXT_LEN], seed=1)
    o = OracleAR(sd, cfg)
    text = torch.stack(texts)
    with torch.inference_mode():
        o.prefill(lat, text)                                   # warm-up (thread pool, allocator)
        t0 = time.perf_counter()
        logits, _, past, _ = o.prefill(lat, text)
        t_pf = time.perf_counter() - t0
        n = 0
        t1 = time.perf_counter()
        while n < 200 and (time.perf_counter() - t1) < budget_s:
            tok = logits.argmax(-1)
            logits, _, past = o.step(tok, n + 1, past)
            n += 1
        t_dec = time.perf_counter() - t1
    return {"value": round(n / (t_pf + t_dec), 3), "unit": "tokens/s", "cores": cores, "kind": "port",
            "sample": f"config C1 (fp32, U=1, N=1, T=64, greedy): prompt pass {t_pf:.3f}s + {n} decode steps {t_dec:.3f}s, "
                      f"torch {torch.__version__} CPU, {cores} threads"}


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port) on the C2 workload; each step is a bounded sample:
    prompt pass + REF_SAMPLE_TOKENS sampled tokens for 16 candidates."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    from oracle.uv_oracle import OracleAR
    from tortoise_tts_b200.synthetic import ARConfig, make_inputs, make_state_dict
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = ARConfig()
    sd = make_state_dict(cfg, seed=0)
    lat, texts = make_inputs(cfg, 1, [TEXT_LEN], seed=1)
    text = torch.stack(texts)
    o = OracleAR(sd, cfg)
    kw = dict(SAMPLING)
    kw.pop("seed")

    def one():
        with torch.inference_mode():
            r = o.generate(lat, text, n_cand=N_CAND, max_new=REF_SAMPLE_TOKENS, ignore_eos=True, seed=1234, **kw)
        return r["codes"].numel()

    for _ in range(min(args.warmup, 1)):
        one()
    t0 = time.perf_counter()
    toks = 0
    for _ in range(args.steps):
        toks += one()
    dt = time.perf_counter() - t0
    val = toks / dt
    sample = (f"per step: prompt pass + {REF_SAMPLE_TOKENS} sampled tokens x {N_CAND} candidates (of the 500-token C2 job), "
              f"fp32, torch CPU {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": round(dt / args.steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": round(val, 3), "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(val, 3), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def workload_config(n_gpus):
    return {"workload": ("C2" if N_CAND == 16 else f"C3 share ({N_CAND} candidates per GPU)") + ": UnifiedVoice AR (30 layers, d=1024, 16 heads, 8194 mel codes) bf16, 1 utterance, "
                        f"{TEXT_LEN} text tokens, {N_CAND} candidates per GPU, top_p=0.8 temperature=0.8, "
                        f"{MAX_NEW} mel tokens (stop ignored), random-init weights seed 0",
            "candidates_per_gpu": N_CAND, "global_candidates": N_CAND * n_gpus, "max_new_tokens": MAX_NEW,
            "text_tokens": TEXT_LEN, "parallelism": f"candidate-sharded x{n_gpus}, weight replica per GPU",
            "l2": "inputs larger than L2: 773 MB of weights streamed every decode step vs 126 MB L2"}



def _dist_max(ms, dev, world):
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def _timed(fn, dev, world, reps=1):
    """reps calls of fn bracketed by barrier + synchronize, CUDA events on the current stream, max over ranks -> ms per call."""
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = None
    for _ in range(reps):
        out = fn()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    return _dist_max(e0.elapsed_time(e1), dev, world) / reps, out


def _ragged(cfg, n_utt, lo, hi, seed):
    import torch
    from tortoise_tts_b200.synthetic import make_inputs
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(lo, hi + 1, (n_utt,), generator=g).tolist()
    lat, texts = make_inputs(cfg, n_utt, lens, seed=1)
    text = torch.zeros(n_utt, max(lens), dtype=torch.long)
    for u, t in enumerate(texts):
        text[u, : len(t)] = t
    return lens, lat, text


def sub_configs(cfg, sd, dev, world, rank, peaks):
    """BASELINE configs C3, C4, C5 (strong scaling: fixed global job, each rank its share).  One warm-up + one timed job each."""
    import torch
    import torch.distributed as dist
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    from tortoise_tts_b200.shard import (allgather_inplace, gather_codes, gather_rows, global_sequence_indices,
                                         plan_balanced_utterance_shards, plan_shards)
    from tortoise_tts_b200.synthetic import make_inputs
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    tens = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
    kw = dict(SAMPLING, ignore_eos=True)
    stop = cfg.stop_mel_token
    out = {}

    def decode_stats(m, n_seq_local, max_new):
        st = m.last_stats
        dec_ms = st.last_device_ms - st.last_prefill_ms
        gbs = st.last_algorithmic_bytes / (dec_ms * 1e-3) / 1e9
        return {"sequences_per_gpu": n_seq_local, "prefill_ms": round(st.last_prefill_ms, 2),
                "us_per_decode_step": round(dec_ms / (max_new - 1) * 1e3, 1), "algorithmic_gbs_per_gpu": round(gbs, 1),
                "roofline_frac": round(gbs / hbm, 4)}

    # ---- C3: one utterance, 256 candidates ----
    n_total, max_new = 256, 500
    shards = plan_shards(1, n_total, world)
    me = shards[rank]
    lat, texts = make_inputs(cfg, 1, [TEXT_LEN], seed=1)
    text = torch.stack(texts)
    m = UnifiedVoiceB200(sd, ARGeometry(), device=dev, max_utts=1, max_seqs=me.n_cand, max_text_len=TEXT_LEN, max_new_tokens=max_new + 4)
    lat_d, text_d = lat.to(dev), text.to(dev)

    def job_c3():
        codes = m.inference_speech(lat_d, text_d, num_return_sequences=me.n_cand, max_generate_length=max_new,
                                   seq_index_base=me.seq_index_base, **kw)
        return gather_codes(codes, stop, shards) if world > 1 else codes
    job_c3()
    ms, codes = _timed(job_c3, dev, world)
    assert codes.shape == (n_total, max_new)
    rec = {"workload": "C3: 1 utterance x 256 candidates, T=64, 500 tokens, candidates split over the GPUs", "scaling": "strong",
           "tokens_per_s": round(n_total * max_new / (ms * 1e-3), 1), "ms": round(ms, 2)}
    rec.update(decode_stats(m, me.n_cand, max_new))
    out["C3"] = rec
    m.close()
    del m

    # ---- C4: 64 ragged utterances x 16 candidates, length-balanced utterance shards ----
    n_utt, n_cand, max_new = 64, 16, 500
    lens, lat, text = _ragged(cfg, n_utt, 16, 402, 100)
    plan = plan_balanced_utterance_shards(lens, world)
    mine = plan[rank]
    idx_lists = [global_sequence_indices(u, n_cand) for u in plan]
    my_lens = [lens[u] for u in mine]
    m = UnifiedVoiceB200(sd, ARGeometry(), device=dev, max_utts=len(mine), max_seqs=len(mine) * n_cand, max_text_len=max(my_lens),
                         max_new_tokens=max_new + 4)
    lat_d = lat[mine].to(dev)
    text_d = text[mine][:, : max(my_lens)].contiguous().to(dev)

    def job_c4():
        codes = m.inference_speech(lat_d, text_d, num_return_sequences=n_cand, max_generate_length=max_new, text_lengths=my_lens,
                                   seq_indices=idx_lists[rank], **kw)
        return gather_rows(codes, idx_lists, stop) if world > 1 else codes
    job_c4()
    ms, codes = _timed(job_c4, dev, world)
    assert codes.shape == (n_utt * n_cand, max_new)
    rec = {"workload": "C4: 64 utterances x 16 candidates, text lengths U{16..402} (seed 100), 500 tokens, paged KV, "
                       "length-balanced utterance shards", "scaling": "strong",
           "tokens_per_s": round(n_utt * n_cand * max_new / (ms * 1e-3), 1), "ms": round(ms, 2),
           "prompt_tokens_per_gpu": [sum(lens[u] for u in p) for p in plan]}
    rec.update(decode_stats(m, len(mine) * n_cand, max_new))
    out["C4"] = rec
    m.close()
    del m

    # ---- C5: 32 utterances x 16 candidates x 603 tokens -> fix-up -> latent re-forward -> all-gather of codes + latents ----
    n_utt, n_cand, max_new = 32, 16, 603
    lens, lat, text = _ragged(cfg, n_utt, 32, 200, 300)
    shards = plan_shards(n_utt, n_cand, world)
    me = shards[rank]
    mine = list(range(me.utt_lo, me.utt_hi))
    my_lens = [lens[u] for u in mine]
    n_local = len(mine) * n_cand
    assert n_utt % world == 0
    m = UnifiedVoiceB200(sd, ARGeometry(), device=dev, max_utts=len(mine), max_seqs=n_local, max_text_len=max(my_lens),
                         max_new_tokens=max_new + 3)
    lat_d = lat[mine].to(dev)
    text_d = text[mine][:, : max(my_lens)].contiguous().to(dev)
    lat_rep = lat_d.repeat_interleave(n_cand, 0)
    text_rep = text_d.repeat_interleave(n_cand, 0)
    tl_rep = torch.tensor(my_lens).repeat_interleave(n_cand)
    wav = torch.full((n_local,), max_new * 1024, dtype=torch.long)
    codes_all = torch.empty((world, n_local, max_new), dtype=torch.int64, device=dev)
    lat_all = torch.empty((world, n_local, max_new, cfg.model_dim), dtype=torch.float32, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    rf_rows = sum(t + 3 + max_new + 2 for t in my_lens) * n_cand
    rf_flops = 2.0 * 377_886_720 * rf_rows + 4.0 * 30 * 1024 * sum((t + 3 + max_new + 2) ** 2 / 2 for t in my_lens) * n_cand

    def job_c5():
        codes = m.inference_speech(lat_d, text_d, num_return_sequences=n_cand, max_generate_length=max_new, text_lengths=my_lens,
                                   seq_index_base=me.seq_index_base, **kw)
        ev[0].record()
        codes, _ = m.fix_autoregressive_codes(codes, strict=False)      # inference.py:353-366 (no stop token under ignore_eos: rows kept)
        codes_all[rank].copy_(codes)
        # the re-forward writes straight into this rank's block of the all-gather buffer (no pack / pad kernel before NCCL)
        m.forward(lat_rep, text_rep, tl_rep, codes, wav, return_latent=True, clip_inputs=False,
                  out=lat_all[rank], ragged_text=True)
        ev[1].record()
        if world > 1:
            allgather_inplace(codes_all)
            allgather_inplace(lat_all)
        ev[2].record()
        return codes_all
    job_c5()
    ms, _ = _timed(job_c5, dev, world)
    rf_ms = _dist_max(ev[0].elapsed_time(ev[1]), dev, world)
    ga_ms = _dist_max(ev[1].elapsed_time(ev[2]), dev, world)      # includes waiting for the slowest rank to reach the collective
    ga_pure_ms = 0.0
    if world > 1:                                                  # the transfer alone: all ranks start together
        ga_pure_ms, _ = _timed(lambda: (allgather_inplace(codes_all), allgather_inplace(lat_all)), dev, world)
    rec = {"workload": "C5: 32 utterances x 16 candidates, text lengths U{32..200} (seed 300), 603 tokens, then stop-token fix-up, "
                       "latent re-forward and NCCL all-gather of codes + latents (utterance-major shards)", "scaling": "strong",
           "tokens_per_s_incl_reforward_and_gather": round(n_utt * n_cand * max_new / (ms * 1e-3), 1), "ms": round(ms, 2),
           "reforward_ms": round(rf_ms, 2), "reforward_rows_per_gpu": rf_rows,
           "reforward_tflops_per_gpu": round(rf_flops / (rf_ms * 1e-3) / 1e12, 1),
           "reforward_frac_of_bf16_sustained_peak": round(rf_flops / (rf_ms * 1e-3) / 1e12 / tens, 4),
           "allgather_ms": round(ga_ms, 3), "allgather_transfer_only_ms": round(ga_pure_ms, 3), "allgather_bytes_per_gpu": int(lat_all[0].numel() * 4 + codes_all[0].numel() * 8),
           "limiting_collective": "ncclAllGather of the fp32 latents (NVLink 5 / NVSwitch)" if world > 1 else "none (1 GPU)"}
    rec.update(decode_stats(m, n_local, max_new))
    out["C5"] = rec
    m.close()
    del m, lat_all, codes_all
    torch.cuda.empty_cache()
    return out


def streaming_latency(cfg, sd, dev):
    """The reference's only streaming use (inference.py:266-285: 1 candidate, HiFiGAN branch): compute_embeddings + get_generator,
    host-side arrival time of every (token, latent) pair.  One persistent launch, flag ring in mapped pinned memory."""
    import torch
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    from tortoise_tts_b200.synthetic import make_inputs
    lat, texts = make_inputs(cfg, 1, [TEXT_LEN], seed=1)
    text = torch.stack(texts).to(dev)
    lat = lat.to(dev)
    m = UnifiedVoiceB200(sd, ARGeometry(), device=dev, max_utts=1, max_seqs=1, max_text_len=TEXT_LEN, max_new_tokens=256)
    n_tok = 200
    res = None
    for it in range(2):                                        # warm-up pass, then the measured one
        inputs = m.compute_embeddings(lat, text)
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        stamps = []
        for tok, latent in m.get_generator(inputs, max_length=inputs.shape[1] + n_tok, do_sample=True, top_p=0.8, temperature=0.8,
                                           top_k=0, num_return_sequences=1, ignore_eos=True, seed=1234):
            stamps.append(time.perf_counter())
        gaps = sorted((b - a) * 1e6 for a, b in zip(stamps, stamps[1:]))
        res = {"candidates": 1, "tokens": len(stamps), "first_token_ms": round((stamps[0] - t0) * 1e3, 3),
               "per_token_us_median": round(gaps[len(gaps) // 2], 1), "per_token_us_p99": round(gaps[int(len(gaps) * 0.99)], 1),
               "tokens_per_s": round(len(stamps) / (stamps[-1] - t0), 1),
               "how": "host perf_counter at the arrival of each yielded (token, latent); first token includes the prompt pass"}
    m.close()
    return res


def clvp_scoring(dev, peaks, with_reference):
    """the CLVP scoring call that follows the C2 job (reference inference.py:392-396): 16 candidates x 500
    codes against one 64-token text, 20-layer encoders, random-init weights.  Device-timed (CUDA events), plus -- when the
    baseline legs are on -- the reference math (oracle, torch eager fp32) on the same GPU."""
    import torch
    from tortoise_tts_b200.clvp import CLVPB200
    from tortoise_tts_b200.synthetic import ClvpConfig, make_clvp_inputs, make_clvp_state_dict
    ccfg = ClvpConfig()
    csd = make_clvp_state_dict(ccfg, seed=5)
    text, codes = make_clvp_inputs(ccfg, N_CAND, TEXT_LEN, MAX_NEW, seed=6)
    text, codes = text.to(dev), codes.to(dev)
    m = CLVPB200(csd, depth=ccfg.depth, device=dev, max_seqs=N_CAND, max_text_len=TEXT_LEN, max_speech_len=MAX_NEW)
    m(text, codes)                                            # warm-up
    l0 = m.kernel_launches
    reps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(reps):
        scores = m(text, codes)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    rows = N_CAND * MAX_NEW + TEXT_LEN
    flops = ccfg.depth * (rows * 2.0 * (768 * 6144 + 1536 * 768) + 4.0 * 64 * 12 * (N_CAND * MAX_NEW ** 2 + TEXT_LEN ** 2))
    tens = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
    rec = {"workload": f"{N_CAND} candidates x {MAX_NEW} codes, {TEXT_LEN} text tokens, 2 x {ccfg.depth} layers, d=768",
           "ms": round(ms, 3), "rows": rows, "tflops": round(flops / (ms * 1e-3) / 1e12, 1),
           "frac_of_bf16_sustained_peak": round(flops / (ms * 1e-3) / 1e12 / tens, 4),
           "kernel_launches_per_call": int((m.kernel_launches - l0) / reps), "top4": torch.topk(scores, 4).indices.tolist(),
           "how": "CUDA events around 5 calls of CLVPB200.forward (each ends with a stream sync), device-resident inputs"}
    m.close()
    if with_reference:
        try:
            from oracle.clvp_oracle import OracleCLVP
            o = OracleCLVP({k: v.to(dev) for k, v in csd.items()}, ccfg)
            o.score(text, codes)
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            want = o.score(text, codes)
            torch.cuda.synchronize(dev)
            rec["reference_same_gpu_fp32_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
            rec["score_abs_err_vs_reference_math"] = round(float((scores - want).abs().max()), 6)
        except Exception as e:
            rec["reference_same_gpu_fp32_ms"] = None
            rec["reference_unavailable"] = repr(e)[:160]
    return rec


def reference_same_gpu(cfg, sd, dev):
    """The oracle (reference math in torch eager; HF-style torch.cat KV cache, one host sync per step for the sampler, like the
    reference's sample_stream) on this GPU, fp32 and bf16: prompt pass + REF_SAMPLE_TOKENS sampled tokens x 16 candidates."""
    import torch
    from oracle.uv_oracle import OracleAR
    from tortoise_tts_b200.synthetic import make_inputs
    lat, texts = make_inputs(cfg, 1, [TEXT_LEN], seed=1)
    text = torch.stack(texts).to(dev)
    lat = lat.to(dev)
    kw = dict(SAMPLING)
    kw.pop("seed")
    res = {"sample": f"prompt pass + {REF_SAMPLE_TOKENS} sampled tokens x {N_CAND} candidates of the C2 job (contexts 68..{68 + REF_SAMPLE_TOKENS}), "
                     f"torch {torch.__version__} eager on the same GPU"}
    for name, dt in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
        o = OracleAR(sd, cfg, dtype=dt, device=dev)
        with torch.inference_mode():
            o.generate(lat, text, n_cand=N_CAND, max_new=4, ignore_eos=True, seed=1234, **kw)      # warm-up
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            r = o.generate(lat, text, n_cand=N_CAND, max_new=REF_SAMPLE_TOKENS, ignore_eos=True, seed=1234, **kw)
            torch.cuda.synchronize(dev)
            dt_s = time.perf_counter() - t0
        res[name] = {"value": round(r["codes"].numel() / dt_s, 1), "unit": "tokens/s", "ms": round(dt_s * 1e3, 1)}
        del o
        torch.cuda.empty_cache()
    return res


def run_ours(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from tortoise_tts_b200 import ARGeometry, UnifiedVoiceB200
    from tortoise_tts_b200.shard import gather_codes, plan_shards
    from tortoise_tts_b200.synthetic import ARConfig, make_inputs, make_state_dict

    cfg = ARConfig()
    sd = make_state_dict(cfg, seed=0)
    lat, texts = make_inputs(cfg, 1, [TEXT_LEN], seed=1)
    text = torch.stack(texts)
    shards = plan_shards(1, N_CAND * world, world)
    me = shards[rank]
    m = UnifiedVoiceB200(sd, ARGeometry(), device=dev, max_utts=1, max_seqs=N_CAND, max_text_len=TEXT_LEN,
                         max_new_tokens=MAX_NEW + 4)
    kw = dict(SAMPLING, ignore_eos=True, seq_index_base=me.seq_index_base)
    lat_d, text_d = lat.to(dev), text.to(dev)
    lat_h, text_h = lat.pin_memory(), text.to(torch.int32).pin_memory()

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def job_device():
        codes = m.inference_speech(lat_d, text_d, num_return_sequences=me.n_cand, max_generate_length=MAX_NEW, **kw)
        if world > 1:
            codes = gather_codes(codes, cfg.stop_mel_token, shards)
        return codes

    def job_host():
        codes = m.inference_speech_host(lat_h, text_h, num_return_sequences=me.n_cand, max_generate_length=MAX_NEW, **kw)
        if world > 1:
            codes = gather_codes(codes.to(dev), cfg.stop_mel_token, shards).cpu()
        return codes

    def timed(job, steps):
        """K jobs bracketed by barrier + synchronize; CUDA events on the launching (current) stream; max over ranks."""
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        out = None
        for _ in range(steps):
            out = job()
        e1.record()
        sync_all()
        wall = time.perf_counter() - t0
        ms = torch.tensor([e0.elapsed_time(e1), wall * 1e3], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms[0].item(), ms[1].item(), out

    for _ in range(args.warmup):
        out = job_device()
    assert out.shape == (N_CAND * world, MAX_NEW), out.shape
    launches0 = m.sync_stats().kernel_launches
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    dec_ms, dec_bytes = [], []
    sync_all()
    # timed region 1: device-resident inputs
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = job_device()
        st = m.last_stats
        dec_ms.append(st.last_device_ms - st.last_prefill_ms)
        dec_bytes.append(st.last_algorithmic_bytes)
    e1.record()
    sync_all()
    ms_dev = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_dev, op=dist.ReduceOp.MAX)
    ms_dev = ms_dev.item()
    launches = m.sync_stats().kernel_launches - launches0
    # timed region 2: host buffers through the C-ABI host entry (wall clock == device timeline here: the call is SYNC)
    job_host()
    ms_e2e_ev, ms_e2e_wall, out_h = timed(job_host, args.steps)
    clk = clocks.stop() if rank == 0 else None
    assert out_h.shape == (N_CAND * world, MAX_NEW)
    tokens_per_step = N_CAND * world * MAX_NEW
    ms_per_step = ms_dev / args.steps
    value = tokens_per_step / (ms_per_step * 1e-3)
    e2e_ms = max(ms_e2e_ev, ms_e2e_wall) / args.steps
    line = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = (sum(dec_bytes) / len(dec_bytes)) / (sum(dec_ms) / len(dec_ms) * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("ar_decode_kernel_c2_launch_bytes")
        except Exception:
            pass
        line = {"metric": METRIC, "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(world),
                "e2e": {"value": round(tokens_per_step / (e2e_ms * 1e-3), 1), "unit": "tokens/s",
                        "h2d_bytes_per_step": int(lat_h.numel() * 4 + text_h.numel() * 4),
                        "d2h_bytes_per_step": int(me.n_cand * MAX_NEW * 8), "ms_per_step": round(e2e_ms, 3)},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                             "frac": round(achieved / peak, 4), "traffic": traffic,
                             "traffic_source": "static: dram__bytes_read+write of one ncu --set full capture of this kernel on this workload "
                                               "(profiles/traffic.json), not re-measured by this run",
                             "kernel": "ar_decode_kernel (persistent head-cluster decode loop, one launch per job)",
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s",
                             "algorithmic_bytes_per_launch": sum(dec_bytes) / len(dec_bytes),
                             "launch_ms": round(sum(dec_ms) / len(dec_ms), 3),
                             "us_per_decode_step": round(sum(dec_ms) / len(dec_ms) / (MAX_NEW - 1) * 1e3, 2)},
                "clocks": clk}
    m.close()
    del m
    torch.cuda.empty_cache()
    peaks_all = {}
    try:
        peaks_all = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    subs = None
    if not args.no_sub_configs:
        subs = sub_configs(cfg, sd, dev, world, rank, peaks_all)
    if rank == 0:
        if subs is not None:
            line["configs"] = subs
        if not args.no_sub_configs:
            try:
                line["streaming"] = streaming_latency(cfg, sd, dev)
            except Exception as e:
                line["streaming"] = {"unavailable": repr(e)[:200]}
        if not args.no_sub_configs:
            try:
                line["clvp_scoring"] = clvp_scoring(dev, peaks_all, not args.no_cpu_baseline)
            except Exception as e:
                line["clvp_scoring"] = {"unavailable": repr(e)[:200]}
        if not args.no_cpu_baseline:
            try:
                line["reference_same_gpu"] = reference_same_gpu(cfg, sd, dev)
            except Exception as e:                                   # a baseline leg must never take the bench line down
                line["reference_same_gpu"] = {"unavailable": repr(e)[:200]}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = oracle_cpu_c1()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    a = parse()
    sys.exit(run_reference(a) if a.impl == "reference" else run_ours(a))
