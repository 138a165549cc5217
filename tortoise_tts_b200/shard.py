"""Multi-GPU sharding of the AR sampler: one process per GPU, a full weight replica each, no collective
on the decode path.

Every (utterance, candidate) sequence is independent after its prompt (reference
stream_generator.py:618-623 expands with repeat_interleave and samples row by row, :1160-1171), so the
global sequence index space  i = u * n_cand + c  is cut into per-rank blocks:

  * n_utt >= world : utterance-major contiguous blocks, so all candidates of an utterance share their
    prompt KV pages on one GPU;
  * n_utt <  world : each utterance's candidates are split over world // n_utt ranks; every one of those
    ranks recomputes the (cheap) prompt.

The Philox counter is keyed by the GLOBAL index (TtsArSampling.seq_index_base), so the sampled codes do
not depend on the world size.  The only exchange is the final all-gather of the finished codes (and,
optionally, of per-step latents) for the downstream CLVP / diffusion stage (reference
inference.py:392-413 consumes them on one device).
"""
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class Shard:
    utt_lo: int          # utterances [utt_lo, utt_hi)
    utt_hi: int
    cand_lo: int         # candidates [cand_lo, cand_hi) of each of those utterances
    cand_hi: int
    n_cand_total: int

    @property
    def n_utt(self):
        return self.utt_hi - self.utt_lo

    @property
    def n_cand(self):
        return self.cand_hi - self.cand_lo

    @property
    def n_seqs(self):
        return self.n_utt * self.n_cand

    @property
    def seq_index_base(self):
        """Global index of local sequence 0 (only meaningful when the local block is contiguous in the
        global order: one utterance slice, or whole utterances)."""
        return self.utt_lo * self.n_cand_total + self.cand_lo

    def global_indices(self):
        return [u * self.n_cand_total + c for u in range(self.utt_lo, self.utt_hi)
                for c in range(self.cand_lo, self.cand_hi)]


def plan_shards(n_utt, n_cand, world):
    """Returns one Shard per rank (possibly empty: n_seqs == 0)."""
    if n_utt < 1 or n_cand < 1 or world < 1:
        raise ValueError("n_utt, n_cand and world must be >= 1")
    out = []
    if n_utt >= world:
        base, extra = divmod(n_utt, world)
        lo = 0
        for r in range(world):
            n = base + (1 if r < extra else 0)
            out.append(Shard(lo, lo + n, 0, n_cand, n_cand))
            lo += n
        return out
    per = world // n_utt                       # ranks per utterance
    for r in range(world):
        u, j = divmod(r, per)
        if u >= n_utt:
            out.append(Shard(n_utt, n_utt, 0, 0, n_cand))      # idle rank
            continue
        base, extra = divmod(n_cand, per)
        c_lo = j * base + min(j, extra)
        c_hi = c_lo + base + (1 if j < extra else 0)
        out.append(Shard(u, u + 1, c_lo, c_hi, n_cand))
    return out


def gather_codes(codes, pad_token, shards, group=None, latents=None):
    """All-gathers per-rank code matrices [n_local, steps_local] (int64) into the global
    [n_utt * n_cand, max_steps] matrix in global sequence order, padding with `pad_token` (finished rows
    are already padded with it, reference quirk Q7).  `latents` [n_local, steps_local, d] (optional) is
    gathered the same way (zero padded).  Works on CUDA tensors with NCCL and CPU tensors with gloo."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    assert len(shards) == world
    dev = codes.device
    shape = torch.tensor([codes.shape[0], codes.shape[1]], dtype=torch.int64, device=dev)
    shapes = [torch.zeros_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape, group=group)
    shapes = torch.stack(shapes).cpu()
    max_rows, max_steps = int(shapes[:, 0].max()), int(shapes[:, 1].max())
    buf = torch.full((max_rows, max_steps), pad_token, dtype=torch.int64, device=dev)
    buf[: codes.shape[0], : codes.shape[1]] = codes
    allbuf = torch.empty((world * max_rows, max_steps), dtype=torch.int64, device=dev)   # concatenated form
    dist.all_gather_into_tensor(allbuf, buf, group=group)
    allbuf = allbuf.view(world, max_rows, max_steps)
    n_total = sum(s.n_seqs for s in shards)
    out = torch.full((n_total, max_steps), pad_token, dtype=torch.int64, device=dev)
    out_lat = None
    if latents is not None:
        d = latents.shape[-1]
        lbuf = torch.zeros((max_rows, max_steps, d), dtype=latents.dtype, device=dev)
        lbuf[: latents.shape[0], : latents.shape[1]] = latents
        alll = torch.empty((world * max_rows, max_steps, d), dtype=latents.dtype, device=dev)
        dist.all_gather_into_tensor(alll, lbuf, group=group)
        alll = alll.view(world, max_rows, max_steps, d)
        out_lat = torch.zeros((n_total, max_steps, d), dtype=latents.dtype, device=dev)
    for r, s in enumerate(shards):
        if s.n_seqs == 0:
            continue
        idx = torch.tensor(s.global_indices(), dtype=torch.long, device=dev)
        out[idx] = allbuf[r, : s.n_seqs]
        if out_lat is not None:
            out_lat[idx] = alll[r, : s.n_seqs]
    assert int(shapes[rank, 0]) == codes.shape[0]
    return (out, out_lat) if latents is not None else out


# ---- length-balanced utterance shards (BASELINE config C4: ragged text lengths) ---------------------------------------
def plan_balanced_utterance_shards(text_lens, world):
    """Utterance-major shards with equal utterance counts (+-1) and balanced prompt lengths: utterances sorted by length,
    longest first, each handed to the least-loaded rank that still has room (LPT).  All candidates of an utterance stay
    on one rank (they share its prompt KV pages).  Returns one ascending list of utterance indices per rank.  The shards
    are not contiguous in the global order: pass `global_sequence_indices(...)` as `seq_indices` so that the sampled
    codes stay invariant to the world size."""
    n = len(text_lens)
    if world < 1:
        raise ValueError("world must be >= 1")
    cap = [n // world + (1 if r < n % world else 0) for r in range(world)]
    load = [0] * world
    out = [[] for _ in range(world)]
    for u in sorted(range(n), key=lambda i: (-int(text_lens[i]), i)):
        r = min((r for r in range(world) if len(out[r]) < cap[r]), key=lambda r: (load[r], r))
        out[r].append(u)
        load[r] += int(text_lens[u])
    return [sorted(x) for x in out]


def global_sequence_indices(utts, n_cand):
    """Global sequence index i = u * n_cand + c of every local sequence of a shard holding utterances `utts`."""
    return [u * n_cand + c for u in utts for c in range(n_cand)]


def gather_rows(local, index_lists, pad_value, group=None):
    """All-gathers per-rank row blocks [n_r, ...] (same trailing shape up to dim 1, which is padded to the longest with
    `pad_value`) and returns the global tensor with row index_lists[r][k] = rank r's row k."""
    world = dist.get_world_size(group)
    dev = local.device
    shape = torch.tensor([local.shape[0], local.shape[1] if local.dim() > 1 else 1], dtype=torch.int64, device=dev)
    shapes = [torch.zeros_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape, group=group)
    shapes = torch.stack(shapes).cpu()
    max_rows, max_cols = int(shapes[:, 0].max()), int(shapes[:, 1].max())
    tail = tuple(local.shape[2:])
    buf = torch.full((max_rows, max_cols) + tail, pad_value, dtype=local.dtype, device=dev)
    if local.shape[0]:
        buf[: local.shape[0], : local.shape[1]] = local
    allbuf = torch.empty((world * max_rows, max_cols) + tail, dtype=local.dtype, device=dev)
    dist.all_gather_into_tensor(allbuf, buf, group=group)
    allbuf = allbuf.view((world, max_rows, max_cols) + tail)
    total = sum(len(x) for x in index_lists)
    out = torch.full((total, max_cols) + tail, pad_value, dtype=local.dtype, device=dev)
    for r, idx in enumerate(index_lists):
        if idx:
            out[torch.tensor(idx, dtype=torch.long, device=dev)] = allbuf[r, : len(idx)]
    return out


def allgather_inplace(buf, group=None):
    """`buf` [world, ...]: every rank has already written ITS block buf[rank] (e.g. the latent re-forward writes its
    output straight into it, `UnifiedVoiceB200.forward(out=buf[rank])`); one NCCL all-gather fills the other blocks in
    place -- no pack / pad / copy kernel precedes the collective."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if buf.shape[0] != world or not buf.is_contiguous():
        raise ValueError("buf must be a contiguous [world, ...] tensor")
    dist.all_gather_into_tensor(buf.view((-1,) + tuple(buf.shape[2:])) if buf.dim() > 2 else buf.view(-1), buf[rank], group=group)
    return buf
