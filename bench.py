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
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    del sd
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
                             "kernel": "ar_decode_kernel (persistent head-cluster decode loop, one launch per job)",
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s",
                             "algorithmic_bytes_per_launch": sum(dec_bytes) / len(dec_bytes),
                             "launch_ms": round(sum(dec_ms) / len(dec_ms), 3),
                             "us_per_decode_step": round(sum(dec_ms) / len(dec_ms) / (MAX_NEW - 1) * 1e3, 2)},
                "clocks": clk}
    m.close()
    if rank == 0:
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
