"""CPU: the C-ABI library loads and exports what include/tts_ar.h declares, the layout arithmetic is
consistent, and the host-side sharding logic (world_size 2, gloo) reassembles the global code matrix."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    g.build()
    return g


def test_c_abi_exports_every_declared_symbol(built):
    hdr = open(os.path.join(ROOT, "include", "tts_ar.h")).read()
    declared = set(re.findall(r"\b(tts_ar_[a-z_0-9]+)\s*\(", hdr))
    assert len(declared) >= 13
    lib = C.CDLL(os.path.join(ROOT, "tortoise_tts_b200", "lib", "libtts_ar.so"))
    for name in declared:
        assert hasattr(lib, name), name
    lib.tts_ar_abi_version.restype = C.c_int
    assert lib.tts_ar_abi_version() == 2
    from tortoise_tts_b200 import _lib
    assert set(_lib.SYMBOLS) == declared


def test_product_fails_loudly_without_a_gpu(built):
    """No CPU fallback: creating a handle on a box without an sm_100 device must raise."""
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from tortoise_tts_b200 import UnifiedVoiceB200
    with pytest.raises(RuntimeError):
        UnifiedVoiceB200({}, device="cpu")
    with pytest.raises(RuntimeError):
        UnifiedVoiceB200({}, device="cuda:0")
    from tortoise_tts_b200.clvp import CLVPB200
    with pytest.raises(RuntimeError):
        CLVPB200({}, depth=1)


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "tortoise_tts_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f


@pytest.fixture(scope="module")
def lc(built):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "liblayout_check.so"))
    lib.lc_afrag_elem_offset.restype = C.c_longlong
    lib.lc_kv_row_offset.restype = C.c_longlong
    lib.lc_float_key.restype = C.c_uint
    lib.lc_float_key.argtypes = [C.c_uint]
    return lib


def test_weight_chunk_map_is_a_bijection(lc):
    seen = np.zeros(1024 * 8, dtype=np.int32)
    for k in range(1024):
        for n in range(8):
            seen[lc.lc_chunk_elem_offset(k, n)] += 1
    assert (seen == 1).all()
    assert lc.lc_chunk_bytes() == 16384
    # the two k of a 32-bit word are adjacent; the four words of a lane's uint4 are b0/b1 of two k16-blocks
    assert lc.lc_chunk_elem_offset(1, 0) == lc.lc_chunk_elem_offset(0, 0) + 1
    assert lc.lc_chunk_elem_offset(8, 0) == lc.lc_chunk_elem_offset(0, 0) + 2
    assert lc.lc_chunk_elem_offset(16, 0) == lc.lc_chunk_elem_offset(0, 0) + 4


def test_afrag_map_is_a_bijection(lc):
    for K in (1024, 4096):
        offs = {lc.lc_afrag_elem_offset(r, c, K) for r in range(32) for c in range(0, K, 1)}
        assert len(offs) == 32 * K and max(offs) == 32 * K - 1


def test_kv_page_rows_do_not_overlap(lc):
    n_pages = 5
    offs = set()
    for layer in range(2):
        for page in range(n_pages):
            for kv in range(2):
                for head in range(16):
                    for slot in range(32):
                        offs.add(lc.lc_kv_row_offset(layer, n_pages, page, kv, head, slot))
    assert len(offs) == 2 * n_pages * 2 * 16 * 32
    assert all(o % 64 == 0 for o in offs) and max(offs) == (len(offs) - 1) * 64


@pytest.mark.parametrize("G", [148, 132, 7])
def test_tile_schedule_covers_every_tile_once(lc, G):
    rot = 0
    for n_tiles in (384, 128, 512, 128, 1025):
        got = []
        t0, cnt = C.c_int(), C.c_int()
        for cta in range(G):
            lc.lc_tile_range(n_tiles, G, rot, cta, C.byref(t0), C.byref(cnt))
            got += list(range(t0.value, t0.value + cnt.value))
        assert sorted(got) == list(range(n_tiles))
        rot = lc.lc_next_rot(n_tiles, G, rot)
        assert 0 <= rot < G


def test_float_key_is_monotone(lc):
    vals = np.array([-np.inf, -3.5, -1e-30, -0.0, 0.0, 1e-30, 2.0, 7.25, np.inf], dtype=np.float32)
    keys = [lc.lc_float_key(int(v.view(np.uint32))) for v in vals]
    assert all(a <= b for a, b in zip(keys, keys[1:]))
    assert keys[0] < keys[1] and keys[-2] < keys[-1]


def test_plan_shards():
    from tortoise_tts_b200.shard import plan_shards
    s = plan_shards(1, 256, 8)                       # C3: one utterance, 256 candidates, 8 GPUs
    assert [x.n_seqs for x in s] == [32] * 8 and [x.seq_index_base for x in s] == list(range(0, 256, 32))
    s = plan_shards(64, 16, 8)                       # C4: utterance-major
    assert [x.n_utt for x in s] == [8] * 8 and s[3].seq_index_base == 24 * 16
    s = plan_shards(3, 5, 2)
    assert sorted(i for x in s for i in x.global_indices()) == list(range(15))
    s = plan_shards(2, 6, 8)
    assert sorted(i for x in s for i in x.global_indices()) == list(range(12))
    s = plan_shards(3, 4, 8)                         # 2 ranks per utterance, 2 idle
    assert sorted(i for x in s for i in x.global_indices()) == list(range(12))


_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from tortoise_tts_b200.shard import plan_shards, gather_codes
backend = sys.argv[4] if len(sys.argv) > 4 else "gloo"
dev = torch.device("cuda", int(sys.argv[3])) if backend == "nccl" else torch.device("cpu")
if backend == "nccl":
    torch.cuda.set_device(dev)
dist.init_process_group(backend, init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
for (U, N) in ((1, 6), (3, 2)):
    shards = plan_shards(U, N, 2)
    me = shards[rank]
    steps = 4 + rank                                  # ranks stop at different lengths
    g = torch.tensor(me.global_indices())
    codes = (g[:, None] * 100 + torch.arange(steps)[None]).long().to(dev)
    lat = codes[..., None].float().repeat(1, 1, 3)
    out, out_lat = gather_codes(codes, 8193, shards, latents=lat)
    assert out.shape == (U * N, 5), out.shape
    for i in range(U * N):
        r = [k for k, s in enumerate(shards) if i in s.global_indices()][0]
        n = 4 + r
        assert out[i, :n].tolist() == [i * 100 + k for k in range(n)], (i, out[i])
        assert (out[i, n:] == 8193).all()
        assert out_lat[i, :n, 0].tolist() == [float(i * 100 + k) for k in range(n)]
dist.barrier()
dist.destroy_process_group()
print("ok", rank)
'''


def run_world2(tmp_path, backend):
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, str(port), str(r), backend], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


def test_gather_codes_world2_gloo(tmp_path):
    run_world2(tmp_path, "gloo")


@pytest.mark.gpu
def test_gather_codes_world2_nccl(tmp_path):
    """The same two-rank gather over NCCL on two GPUs (skipped on a one-GPU box; run with `gpurun --gpus 2`)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    run_world2(tmp_path, "nccl")


# ---- the device sampler's threshold search, restated in numpy (csrc/ar_sample.cuh) ---------------------------------
def _keys(a):
    b = a.astype(np.float32).view(np.uint32).astype(np.uint64)
    return np.where(b & 0x80000000, (~b) & 0xFFFFFFFF, b | 0x80000000).astype(np.uint64)


def _key_to_float(k):
    b = (k & 0x7FFFFFFF) if (k & 0x80000000) else ((~k) & 0xFFFFFFFF)
    return np.uint32(b).view(np.float32)


def _threshold_search(keys, wgt, lim, lo, n_lo, smin, smax):
    """Smallest key T in [lo, kmax] with weight{key > T} < lim: 4-way rounds over the score interval that also count
    the tokens left inside; <= 32 of them end the search with a descending sort + prefix sum.  Returns (T, rounds)."""
    f32 = np.float32
    kmax = int(keys.max())
    hi, n_hi, m_hi, first, rounds = kmax, 0, f32(0), True, 0
    while lo < hi:
        if n_lo - n_hi <= 32:
            sel = (keys >= lo) & (keys <= hi)
            order = np.argsort(-keys[sel].astype(np.int64), kind="stable")
            ck, cw = keys[sel][order], wgt[sel][order]
            incl = np.cumsum(cw, dtype=np.float32)
            brk = np.nonzero(m_hi + incl >= lim)[0]
            if len(brk):
                lo = int(ck[brk[0]])
            break
        fl, fh = max(_key_to_float(lo), f32(smin)), _key_to_float(hi)
        t, ok = [], True
        for j in range(3):
            with np.errstate(all="ignore"):
                pj = f32(fl + (fh - fl) * f32(0.25 * (j + 1)))
                if first:
                    pj = max(pj, f32(smax - (16.0, 4.0, 1.0)[j])) if np.isfinite(pj) else f32(smax - (16.0, 4.0, 1.0)[j])
            ok = ok and bool(np.abs(pj) <= 3.0e38)
            t.append(int(_keys(np.array([pj], np.float32))[0]))
        ok = ok and lo < t[0] < t[1] < t[2] < hi
        if not ok:
            q = (hi - lo) >> 2
            t = [lo + q * (j + 1) - 1 if q else min(lo + j, hi) for j in range(3)]
        first = False
        rounds += 1
        ms = [wgt[keys > tj].sum(dtype=np.float32) for tj in t]
        cn = [int((keys > tj).sum()) for tj in t]
        if ms[0] < lim:
            hi, m_hi, n_hi = t[0], ms[0], cn[0]
        elif ms[1] < lim:
            lo, n_lo, hi, m_hi, n_hi = t[0] + 1, cn[0], t[1], ms[1], cn[1]
        elif ms[2] < lim:
            lo, n_lo, hi, m_hi, n_hi = t[1] + 1, cn[1], t[2], ms[2], cn[2]
        else:
            lo, n_lo = t[2] + 1, cn[2]
    return lo, rounds


@pytest.mark.parametrize("std,top_k,top_p", [(0.02, 0, 0.8), (1.0, 50, 0.8), (6.0, 0, 0.5), (0.5, 16, 1.0),
                                             (3.0, 8000, 0.9), (1.0, 1, 1.0), (0.1, 0, 0.02)])
def test_threshold_search_keeps_the_hf_support_set(std, top_k, top_p):
    """The search the device sampler runs for top-k and top-p (one routine, ar_sample.cuh) selects exactly HF's
    TopKLogitsWarper / TopPLogitsWarper support (oracle.uv_oracle, pinned to HF's classes by test_oracle_golden.py),
    ties included, in a handful of rounds instead of 16 key-space bisection steps."""
    from oracle.uv_oracle import apply_top_k, apply_top_p
    rng = np.random.default_rng(int(std * 100) + top_k)
    V = 8194
    for trial in range(3):
        sc = rng.normal(0, std, V).astype(np.float32)
        sc[rng.integers(0, V, 40)] = sc[rng.integers(0, V, 40)]            # ties
        keys = _keys(sc)
        smin, smax = sc.min(), sc.max()
        kmin, n_keep, rounds = int(_keys(np.array([-np.inf], np.float32))[0]), V, 0
        if 0 < top_k < V:
            kmin, r = _threshold_search(keys, np.ones(V, np.float32), np.float32(top_k), kmin, n_keep, smin, smax)
            n_keep, rounds = int((keys >= kmin).sum()), rounds + r
        e = np.where(keys >= kmin, np.exp(sc - smax), 0).astype(np.float32)
        if top_p < 1.0:
            lim = np.float32(top_p) * e.sum(dtype=np.float32)
            kmin, r = _threshold_search(keys, e, lim, kmin, n_keep, smin, smax)
            rounds += r
        ref = apply_top_p(apply_top_k(torch.from_numpy(sc)[None], top_k), top_p)[0]
        ref_keep = torch.isfinite(ref).numpy()
        mine = keys >= kmin
        # the nucleus edge may differ by the tokens whose cumulative mass is within fp32 summation error of p
        diff = np.nonzero(mine != ref_keep)[0]
        if len(diff):
            p = torch.softmax(apply_top_k(torch.from_numpy(sc)[None], top_k)[0].double(), -1).numpy()
            order = np.argsort(-sc, kind="stable")
            cum_excl = np.empty(V)
            cum_excl[order] = np.cumsum(p[order]) - p[order]
            assert np.all(np.abs(cum_excl[diff] - top_p) < 1e-4), (diff[:5], cum_excl[diff][:5])
        assert rounds <= 16, rounds


@pytest.mark.parametrize("std,mass", [(1.0, 0.9), (2.5, 0.5), (6.0, 0.2), (0.3, 0.99), (2.5, 0.9)])
def test_threshold_search_keeps_the_typical_support_set(std, mass):
    """Pass -1 of the device sampler (csrc/ar_sample.cuh): the same search over the key -|s - E_p[s]| with weight
    exp(s - max) and limit mass * Z selects the support of the reference's TypicalLogitsWarper
    (oracle.apply_typical, pinned to the reference's own class by tests/test_oracle_golden.py).  fp32 throughout, as
    on the device; tokens may differ only where |shifted - threshold| or the cumulative mass is within fp32 rounding."""
    from oracle.uv_oracle import apply_typical
    f32 = np.float32
    rng = np.random.default_rng(int(std * 10) + int(mass * 100))
    V = 8194
    for trial in range(3):
        sc = rng.normal(0, std, V).astype(np.float32)
        smax, smin = sc.max(), sc.min()
        d = (sc - smax).astype(np.float32)
        e = np.exp(d).astype(np.float32)
        Z = e.sum(dtype=np.float32)
        cT = f32(smax + f32((e * d).sum(dtype=np.float32) / Z))
        tk = _keys((-np.abs(cT - sc)).astype(np.float32))
        lo0 = int(_keys(np.array([-np.inf], np.float32))[0])
        pmn = f32(-max(abs(cT - smax), abs(cT - smin)))
        T, rounds = _threshold_search(tk, e, f32(mass) * Z, lo0, V, pmn, f32(0.0))
        mine = tk >= T
        ref = torch.isfinite(apply_typical(torch.from_numpy(sc)[None], mass)[0]).numpy()
        diff = np.nonzero(mine != ref)[0]
        if len(diff):
            s64 = sc.astype(np.float64)
            p = np.exp(s64 - s64.max()); p /= p.sum()
            sh = np.abs((p * s64).sum() - s64)
            thr = sh[ref].max()
            order = np.argsort(sh, kind="stable")
            cum = np.cumsum(p[order])
            j = int((cum < mass).sum())
            near_mass = min(abs(cum[j] - mass), abs((cum[j - 1] if j else 0.0) - mass)) < 1e-5
            assert len(diff) <= 2 and (near_mass or np.all(np.abs(sh[diff] - thr) < 1e-5 * max(1.0, thr))), (diff[:5], sh[diff][:5], thr)
        assert mine.sum() >= 1 and rounds <= 16, rounds


def test_module_weights_reads_lora_merged_values():
    """Boundary: `from_module` collects weights by reading `.weight` (the LoRA-merged value under
    torch.nn.utils.parametrize, reference lora.py:163-195), not `state_dict()` (which holds
    `parametrizations.weight.original`, the un-merged base)."""
    import torch
    import torch.nn.utils.parametrize as parametrize
    from tests.standin import LoRADelta, StandInUnifiedVoice
    from tortoise_tts_b200.synthetic import ARConfig, make_state_dict
    from tortoise_tts_b200.unified_voice import UnifiedVoiceB200
    cfg = ARConfig(layers=2, model_dim=64, heads=2, number_mel_codes=40, start_mel_token=38, stop_mel_token=39,
                   max_text_tokens=10, max_mel_tokens=12, number_text_tokens=15, start_text_token=15)
    sd = make_state_dict(cfg, seed=1)
    mod = StandInUnifiedVoice(cfg, sd)
    plain = UnifiedVoiceB200.module_weights(mod)
    assert set(sd) <= set(plain)
    for k, v in sd.items():
        assert torch.equal(plain[k], v), k
    layer = mod.gpt.h[1].attn.c_attn
    parametrize.register_parametrization(layer, "weight", LoRADelta(tuple(layer.weight.shape), seed=3))
    assert "gpt.h.1.attn.c_attn.weight" not in mod.state_dict()                    # what state_dict() would have missed
    assert "gpt.h.1.attn.c_attn.parametrizations.weight.original" in mod.state_dict()
    merged = UnifiedVoiceB200.module_weights(mod)
    lo = layer.parametrizations.weight[0]
    want = sd["gpt.h.1.attn.c_attn.weight"] + (lo.lora_B @ lo.lora_A).detach() * lo.scaling
    assert torch.allclose(merged["gpt.h.1.attn.c_attn.weight"], want)
    assert not torch.equal(merged["gpt.h.1.attn.c_attn.weight"], sd["gpt.h.1.attn.c_attn.weight"])
    assert torch.equal(merged["gpt.h.0.attn.c_attn.weight"], sd["gpt.h.0.attn.c_attn.weight"])
    lo.enabled = False                                                             # disable_lora (lora.py:204-205)
    assert torch.equal(UnifiedVoiceB200.module_weights(mod)["gpt.h.1.attn.c_attn.weight"], sd["gpt.h.1.attn.c_attn.weight"])
    # a swapped-in LoRALinear-style module (lora.py:21-92): base .weight + un-merged lora_A / lora_B
    head = mod.mel_head
    g = torch.Generator().manual_seed(9)
    head.lora_B = torch.nn.Parameter(torch.randn(head.out_features, 4, generator=g) * 0.05)
    head.lora_A = torch.nn.Parameter(torch.randn(4, head.in_features, generator=g) * 0.05)
    head.scaling, head.merged, head.enabled = 0.5, False, True
    w = UnifiedVoiceB200.module_weights(mod)["mel_head.weight"]
    assert torch.allclose(w, sd["mel_head.weight"] + (head.lora_B @ head.lora_A).detach() * 0.5)
    head.merged = True                                                             # eval() with merge_weights: already in .weight
    assert torch.equal(UnifiedVoiceB200.module_weights(mod)["mel_head.weight"], sd["mel_head.weight"])


def test_set_mel_padding_per_row_lengths():
    """forward(): lengths arrive as [n] or [n, 1] (reference inference.py:371-379) and are applied per row like the
    reference's set_mel_padding (unified_voice.py:494-506) -- host-side shape logic only."""
    import torch
    n, M, stop = 3, 6, 99
    codes = torch.arange(n * M).reshape(n, M)
    for wav in (torch.tensor([2048, 4096, 1024]), torch.tensor([[2048], [4096], [1024]])):
        w = torch.as_tensor(wav).reshape(-1)
        mel_lengths = torch.div(w, 1024, rounding_mode="trunc")
        idx = torch.arange(M)[None, :]
        out = torch.where(idx >= (mel_lengths[:, None] + 1), torch.full_like(codes, stop), codes)
        ref = codes.clone()
        for b in range(n):                         # the reference's loop
            actual_end = int(mel_lengths[b]) + 1
            ref[b, actual_end:] = stop
        assert out.shape == (n, M) and torch.equal(out, ref)
