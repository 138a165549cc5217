"""Times the stop-token fix-up + calm-token trim epilogue (tts_ar_fixup_codes) on the device with
CUDA events (kernel only, inputs resident), end to end through the adapter (launch + one D2H read of the two int32
results), and the oracle restatement of the reference's Python loop on the host cores.  One JSON line per size."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tortoise_tts_b200 import UnifiedVoiceB200, ARGeometry  # noqa: E402
from tortoise_tts_b200.synthetic import ARConfig, make_state_dict  # noqa: E402
from oracle.fixup import fixup_codes, calm_trim  # noqa: E402  (cpu_baseline leg only)

cfg = ARConfig(layers=1)
m = UnifiedVoiceB200(make_state_dict(cfg, seed=0), ARGeometry(layers=1), max_utts=1, max_seqs=2, max_text_len=8, max_new_tokens=8)
rng = np.random.default_rng(0)
for rows, n in [(16, 500), (256, 604)]:
    codes = rng.integers(0, 8192, size=(rows, n)).astype(np.int64)
    for r in range(rows):
        codes[r, int(rng.integers(n // 2, n)):] = 8193
    src = torch.from_numpy(codes).cuda()
    work = src.clone()
    first = torch.empty(rows, dtype=torch.int32, device="cuda")
    trim = torch.empty(rows, dtype=torch.int32, device="cuda")
    import ctypes as C
    args = lambda: (m.handle, C.c_void_p(work.data_ptr()), rows, n, n, 83, C.c_void_p(first.data_ptr()), C.c_void_p(trim.data_ptr()), m._stream())
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for it in range(25):
        work.copy_(src)
        torch.cuda.synchronize()
        ev[0].record()
        m._check(m.lib.tts_ar_fixup_codes(*args()), "fixup")
        ev[1].record()
        torch.cuda.synchronize()
        if it >= 5:
            ts.append(ev[0].elapsed_time(ev[1]) * 1e3)
    e2e = []
    for it in range(25):
        work.copy_(src)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m.fix_autoregressive_codes(work)
        e2e.append((time.perf_counter() - t0) * 1e6)
    t0 = time.perf_counter()
    reps = 20
    for _ in range(reps):
        fixed, _f = fixup_codes(codes)
        calm_trim(fixed)
    cpu_us = (time.perf_counter() - t0) / reps * 1e6
    byts = rows * n * 8
    print(json.dumps({"op": "fixup_codes", "rows": rows, "len": n, "kernel_us": round(float(np.median(ts)), 2),
                      "adapter_e2e_us": round(float(np.median(e2e[5:])), 1), "cpu_port_us": round(cpu_us, 1),
                      "algorithmic_bytes": 2 * byts, "note": "launch-latency bound: the array is %d KB" % (byts // 1024)}))
m.close()
