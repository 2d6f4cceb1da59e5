"""Utterance sharding across GPUs: utterances are independent, so the batch is
partitioned by utterance with NO collective on the data path; NCCL/gloo is used only to gather results.

  cost model   : algorithmic FLOPs of one utterance (SURVEY.md 8d formulas)
  lpt_assign   : longest-processing-time-first greedy assignment to ranks
  make_batches : length-bucketed batches under a row budget (rows = varlen-packed frames)
  assign_batches : batches formed over the WHOLE job, then LPT over batch costs (each rank keeps full batches)
  gather_by_index : optional result gather (all_gather of per-utterance tensors, padded)
"""
import math
from typing import Dict, List, Sequence

import torch


def flow_flops(T: int, v2: bool = False, n_steps: int = 10) -> float:
    """Estimator FLOPs per CFG-pair call x steps (measured with FlopCounterMode on the reference)."""
    if v2:
        per_call = 264.4e6 * T + 229376.0 * T * T
    else:
        per_call = 172.4e6 * T + 4096.0 * (8 * T * T + 56 * math.ceil(T / 2) ** 2)
    return n_steps * per_call


def hift_flops(T_gen: int, v2: bool = False) -> float:
    return (612.3e6 if v2 else 602.1e6) * T_gen


def utterance_cost(total_frames: int, gen_frames: int, v2: bool = False) -> float:
    return flow_flops(total_frames, v2) + hift_flops(gen_frames, v2)


def batch_seconds(total_frames: Sequence[int], gen_frames: Sequence[int], v2: bool = False) -> float:
    """Predicted B200 time of one varlen batch through flow.inference + hift.inference (bf16): least-squares fit to the
    84 batches of config 5 on one GPU (profiles/r01_config5_batches_x1.json, rms error 4.5 %):
    1.39 ms per algorithmic TFLOP + 3.25 us per frame + 21 ms per batch.  The weight assign_batches balances."""
    flops = sum(utterance_cost(t, g, v2) for t, g in zip(total_frames, gen_frames))
    return 1.39e-15 * flops + 3.25e-6 * sum(total_frames) + 0.021


def lpt_assign(costs: Sequence[float], n_ranks: int) -> List[List[int]]:
    """Greedy LPT: heaviest utterance first onto the least-loaded rank.  Deterministic (ties by index)."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    loads = [0.0] * n_ranks
    out: List[List[int]] = [[] for _ in range(n_ranks)]
    for i in order:
        r = min(range(n_ranks), key=lambda k: (loads[k], k))
        out[r].append(i)
        loads[r] += costs[i]
    return out


def make_batches(indices: Sequence[int], lens: Sequence[int], row_budget: int = 48 * 1024, align: int = 256,
                 max_batch: int = 64) -> List[List[int]]:
    """Group a rank's utterances into batches of ONE stride bucket each (stride = round_up(len + 8, align)), at most
    row_budget // stride (and max_batch) utterances per batch.  Uniform strides mean every full batch of a bucket
    has the same layout pattern, so the estimator program and its CUDA graph are built once per bucket and re-bound
    to each batch's lengths (SeqLayout.set_lens); only the last, partial batch of a bucket needs a second one."""
    buckets: Dict[int, List[int]] = {}
    for i in sorted(indices, key=lambda i: (lens[i], i)):
        stride = (lens[i] + 8 + align - 1) // align * align
        buckets.setdefault(stride, []).append(i)
    batches: List[List[int]] = []
    for stride in sorted(buckets):
        nb = max(1, min(max_batch, row_budget // stride))
        idx = buckets[stride]
        batches += [idx[k:k + nb] for k in range(0, len(idx), nb)]
    return batches


def assign_batches(costs: Sequence[float], lens: Sequence[int], n_ranks: int, row_budget: int = 48 * 1024,
                   align: int = 256, max_batch: int = 64, max_imbalance: float = 1.05) -> List[List[List[int]]]:
    """Per-rank batch lists.  Sharding single utterances (lpt_assign) and batching per rank leaves every rank a thin
    slice of every stride bucket, i.e. more and smaller batches as the job spreads (config 5 on 8 GPUs: 136 batches
    against 84 on one GPU).  Here the batches are formed over the whole job first and LPT runs
    over the batch costs, so a rank's batches are as full as on one GPU.  Falls back to utterance-level sharding when
    there are too few batches to balance (predicted max load / mean load > max_imbalance).  Deterministic: every rank
    computes the same assignment from the same lengths, no communication.  Measured on 8 GPUs the two schemes tie
    (6946 vs 7097 audio-s/s, profiles/r01_config5_sweep2048_x8*.jsonl): the 8 % below 8 x one GPU is each GPU's batches
    running slower with 8 ranks on the box (host side of the ragged path), not the sharding."""
    idx = list(range(len(costs)))
    if n_ranks == 1:
        return [make_batches(idx, lens, row_budget, align, max_batch)]
    batches = make_batches(idx, lens, row_budget, align, max_batch)
    bcost = [1.39e-15 * sum(costs[u] for u in b) + 3.25e-6 * sum(lens[u] for u in b) + 0.021 for b in batches]   # batch_seconds
    asg = lpt_assign(bcost, n_ranks)
    loads = [sum(bcost[i] for i in a) for a in asg]
    if sum(loads) > 0 and max(loads) * n_ranks / sum(loads) <= max_imbalance:
        return [[batches[i] for i in sorted(a)] for a in asg]
    return [make_batches(mine, lens, row_budget, align, max_batch) for mine in lpt_assign(costs, n_ranks)]


def gather_by_index(local: Dict[int, torch.Tensor], n_total: int, group=None) -> List[torch.Tensor]:
    """Collect per-utterance 1-D results from all ranks (result gather only; not on the compute path).
    Works on any backend: lengths first, then one padded all_gather."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [local[i] for i in range(n_total)]
    world = dist.get_world_size(group)
    dev = next(iter(local.values())).device if local else torch.device("cpu")
    lens = torch.zeros(n_total, dtype=torch.int64, device=dev)
    for i, t in local.items():
        lens[i] = t.numel()
    dist.all_reduce(lens, op=dist.ReduceOp.SUM, group=group)
    owner = torch.zeros(n_total, dtype=torch.int64, device=dev)
    rank = dist.get_rank(group)
    for i in local:
        owner[i] = rank
    dist.all_reduce(owner, op=dist.ReduceOp.SUM, group=group)
    per_rank = [int(lens[(owner == r)].sum().item()) for r in range(world)]
    pad = max(per_rank) if per_rank else 0
    mine = torch.zeros(pad, dtype=torch.float32, device=dev)
    off = 0
    for i in sorted(local):
        n = local[i].numel()
        mine[off:off + n] = local[i].reshape(-1).float()
        off += n
    bufs = [torch.zeros(pad, dtype=torch.float32, device=dev) for _ in range(world)]
    dist.all_gather(bufs, mine, group=group)
    out: List[torch.Tensor] = [None] * n_total
    offs = [0] * world
    for i in range(n_total):          # each rank packed its utterances in increasing index order
        r, n = int(owner[i].item()), int(lens[i].item())
        out[i] = bufs[r][offs[r]:offs[r] + n]
        offs[r] += n
    return out
