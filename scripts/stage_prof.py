"""Bring-up tool: per-stage timestamps (CTA 0) of one decode step of the C2 workload.
usage: TTS_AR_PROF_STEP=200 python scripts/stage_prof.py [n_cand] [steps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tortoise_tts_b200 import UnifiedVoiceB200, ARGeometry  # noqa: E402
from tortoise_tts_b200.synthetic import ARConfig, make_state_dict, make_inputs  # noqa: E402

OPS = ["LN_QKV", "ATTN", "PROJ", "LN_FC", "PROJ2", "HEAD", "SAMPLE", "LATENT"]
n_cand = int(sys.argv[1]) if len(sys.argv) > 1 else 16
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 300
layers = int(os.environ.get("LAYERS", "30"))
cfg = ARConfig(layers=layers)
sd = make_state_dict(cfg, seed=0)
lat, texts = make_inputs(cfg, 1, [64], seed=1)
text = torch.stack(texts)
m = UnifiedVoiceB200(sd, ARGeometry(layers=layers), max_utts=1, max_seqs=n_cand, max_text_len=64, max_new_tokens=520)
if "TTS_AR_PROF_STEP" in os.environ:
    m.set_option("prof_step", int(os.environ["TTS_AR_PROF_STEP"]))
kw = dict(do_sample=True, top_p=0.8, temperature=0.8, top_k=0, seed=1234, ignore_eos=True)
for _ in range(2):
    m.inference_speech(lat.cuda(), text.cuda(), num_return_sequences=n_cand, max_generate_length=steps, **kw)
s = m.last_stats
dec = s.last_device_ms - s.last_prefill_ms
print(f"N={n_cand} steps={steps}: device {s.last_device_ms:.2f} ms, prefill {s.last_prefill_ms:.2f} ms, decode {dec / (steps - 1) * 1e3:.1f} us/step, "
      f"alg {s.last_algorithmic_bytes / (dec * 1e-3) / 1e9:.0f} GB/s")
if "TTS_AR_PROF_STEP" in os.environ:
    n_instr = 1 + layers * 5 + 1
    prog = np.array([[6]] + [[o] for _ in range(layers) for o in (0, 1, 2, 3, 4)] + [[5]])   # ProgIter order
    raw = m.debug_buffer("prof", (16384,), dtype=torch.int64).cpu().numpy()
    bar = raw[8192:8192 + 128 * 8].reshape(128, 8)
    if bar[:, 0].any():
        for b, name in enumerate(["C->A (before ln_1)", "A->B (before ln_2)", "B->C (before mlp.c_proj)"]):
            arr, rel = bar[:, b], bar[:, b + 3]
            t0 = arr.min()
            order = np.argsort(arr)
            print(f"barrier {name}: arrivals span {(arr.max() - t0) / 1e3:.2f} us (median +{(np.median(arr) - t0) / 1e3:.2f}), "
                  f"last arrival -> first release {(rel.min() - arr.max()) / 1e3:.2f} us, -> last release {(rel.max() - arr.max()) / 1e3:.2f} us; "
                  f"last 4 CTAs {order[-4:].tolist()}, first 4 {order[:4].tolist()}")
            if os.environ.get("BAR_DETAIL"):
                cl = (arr - t0).reshape(32, 4)
                print("   per-cluster arrival offset ns (max over the 4 CTAs), cluster id ascending:")
                print("   " + " ".join(f"{int(x):4d}" for x in cl.max(axis=1)))
                print("   per-cluster release offset ns after last arrival (max over the 4 CTAs):")
                print("   " + " ".join(f"{int(x):4d}" for x in (rel - arr.max()).reshape(32, 4).max(axis=1)))
    aw = raw[10240:10240 + 128 * 16].reshape(128, 16)
    if aw.any():
        dur = (aw[:, :8] - aw[:, 8:16]) / 1e3          # per (CTA, warp): q arrived -> granule loop done, us
        start = aw[:, 8:16].min()
        end = (aw[:, :8] - start) / 1e3
        print("attention of layer 3, per warp (row = w >> 2, quarter = w & 3 at 16 rows): granule-loop us, mean / max over the 128 CTAs")
        print("   " + "  ".join(f"w{w}: {dur[:, w].mean():.2f}/{dur[:, w].max():.2f}" for w in range(8)))
        print(f"   CTA-level (slowest warp): mean {dur.max(axis=1).mean():.2f} us, max {dur.max():.2f}; last warp of the grid ends {end.max():.2f} us after the first q")
    raw = raw[:1024 + 4096]
    t = raw[: 2 * n_instr].reshape(n_instr, 2)
    marks = raw[1024:].reshape(-1, 2)
    marks = marks[marks[:, 1] > 0]
    names = {1: "A-ready", 2: "mma-done", 3: "epi-done", 11: "attn-plan", 4: "reduce-synced", 5: "a-frags+bias issued", 6: "chunks probed", 12: "attn-partials", 13: "attn-merged", 21: "ln-loaded", 22: "ln-normalised", 23: "ln-dsmem-stored", 31: "smp-scores+max", 32: "smp-exp+Z", 33: "smp-top-p", 34: "smp-multinomial", 35: "smp-bookkeeping"}
    names.update({100 + i: "pre-bar " + o for i, o in enumerate(OPS)})
    names.update({200 + i: "post-bar " + o for i, o in enumerate(OPS)})
    ids = marks[:, 0].tolist()
    if 106 in ids and 100 in ids:
        i0 = ids.index(106); i1 = ids.index(206); i2 = ids.index(100)
        print(f"SAMPLE: barrier wait {(marks[i1,1]-marks[i0,1])/1e3:.2f} us, body {(marks[i2,1]-marks[i1,1])/1e3:.2f} us")
        if 105 in ids:
            i3 = ids.index(105); i4 = ids.index(205)
            print(f"HEAD: barrier wait {(marks[i4,1]-marks[i3,1])/1e3:.2f} us; step (SAMPLE pre-bar -> HEAD post-bar) {(marks[i4,1]-marks[i0,1])/1e3:.1f} us")
    if 206 in ids:
        i1 = ids.index(206)
        for i in range(i1 + 1, min(i1 + 8, len(marks))):  # sampler marks
            print(f"   {names.get(int(marks[i, 0]), marks[i, 0]):18s} +{int(marks[i, 1] - marks[i - 1, 1])}")
    if 205 in ids:
        i4 = ids.index(205)
        print("marks of the HEAD phase (delta ns from previous mark):")
        for i in range(i4, len(marks)):
            print(f"   {names.get(int(marks[i, 0]), marks[i, 0]):18s} +{int(marks[i, 1] - marks[i - 1, 1])}")
    sel = [i for i in range(len(marks)) if marks[i, 0] == 100]       # pre-barrier LN_QKV marks = layer starts
    if len(sel) > 4:
        a, b = sel[3], sel[4]
        print("marks of layer 3 (delta ns from previous mark):")
        for i in range(a, b + 1):
            print(f"   {names.get(int(marks[i, 0]), marks[i, 0]):18s} +{int(marks[i, 1] - marks[i - 1, 1])}")
    if not t.any():
        m.close(); sys.exit(0)
    before, after = t[:, 0], t[:, 1]
    wait = after - before                       # barrier wait seen by CTA 0
    work = np.append(before[1:], before[-1]) - after   # stage body on CTA 0 (until it reaches the next barrier)
    tot = {}
    for i in range(n_instr):
        op = OPS[prog[i, 0]]
        a = tot.setdefault(op, [0, 0.0, 0.0])
        a[0] += 1; a[1] += wait[i]; a[2] += work[i]
    print("step span (SAMPLE barrier -> HEAD start):", (before[-1] - before[0]) / 1e3, "us")
    for op, (n, w, k) in tot.items():
        print(f"  {op:8s} n={n:3d}  barrier-wait avg {w / n / 1e3:7.2f} us   body(CTA0) avg {k / n / 1e3:7.2f} us")
    L = 3
    print("layer", L, [(OPS[prog[i, 0]], int(wait[i]), int(work[i])) for i in range(1 + 5 * L, 6 + 5 * L)])
m.close()
