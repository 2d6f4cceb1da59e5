"""N>1 host logic on CPU: LPT sharding, bucketing, and the result gather over a world_size-2 gloo group."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cosyvoice_b200 import shard


def test_cost_model_matches_survey_checkpoints():
    assert abs(shard.flow_flops(937, n_steps=1) / 240.85e9 - 1) < 0.01        # v1 T=937
    assert abs(shard.flow_flops(1218, n_steps=1) / 343.61e9 - 1) < 0.01
    assert abs(shard.flow_flops(650, v2=True, n_steps=1) / 268.77e9 - 1) < 0.01
    assert abs(shard.utterance_cost(1218, 937) / 4.00e12 - 1) < 0.01


def test_lpt_is_balanced_and_complete():
    g = torch.Generator().manual_seed(0)
    dur = 2 + 28 * torch.rand(2048, generator=g)
    gen = [int(round(d.item() * 25 * 3.75)) for d in dur]
    costs = [shard.utterance_cost(281 + t, t) for t in gen]
    for n in (1, 2, 4, 8):
        parts = shard.lpt_assign(costs, n)
        assert sorted(i for p in parts for i in p) == list(range(2048))
        loads = [sum(costs[i] for i in p) for p in parts]
        assert max(loads) / (sum(loads) / n) < 1.01


def test_batches_respect_row_budget():
    lens = [300 + 37 * i for i in range(60)]
    batches = shard.make_batches(range(60), lens, row_budget=16 * 1024)
    assert sorted(i for b in batches for i in b) == list(range(60))
    for b in batches:
        assert sum((lens[i] + 8 + 255) // 256 * 256 for i in b) <= 16 * 1024 or len(b) == 1


def _worker(rank, world, port, n_total):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    costs = [float(10 + (i * 7) % 13) for i in range(n_total)]
    mine = shard.lpt_assign(costs, world)[rank]
    local = {i: torch.full((5 + i % 4,), float(i)) for i in mine}       # stand-in for per-utterance waveforms
    out = shard.gather_by_index(local, n_total)
    ok = all(o.numel() == 5 + i % 4 and bool((o == i).all()) for i, o in enumerate(out))
    flag = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    assert flag.item() == 1.0


def test_gather_over_gloo_world2():
    mp.spawn(_worker, args=(2, 29533, 23), nprocs=2, join=True)


def test_batches_are_single_stride_bucket_and_patterns_repeat():
    import random
    rnd = random.Random(3)
    lens = [281 + rnd.randint(188, 2812) for _ in range(512)]
    batches = shard.make_batches(range(512), lens, row_budget=48 * 1024, align=256)
    patterns = set()
    for b in batches:
        strides = {(lens[i] + 8 + 255) // 256 * 256 for i in b}
        assert len(strides) == 1
        patterns.add((strides.pop(), len(b)))
    assert len(patterns) <= 2 * 12            # <= one full + one partial pattern per 256-row bucket


def test_batch_level_assignment_keeps_full_batches_and_falls_back():
    """assign_batches: whole-job batches spread by LPT over batch costs -- every utterance exactly once, the same number
    of batches as on one GPU (utterance-level sharding needs ~60 % more on 8 ranks), load within 5 %; with too few
    batches to balance it returns the utterance-level sharding instead."""
    g = torch.Generator().manual_seed(2048)
    dur = 2 + 28 * torch.rand(2048, generator=g)
    gen = [int(round(d.item() * 25) / 25 * 24000 / 256) for d in dur]
    tot = [281 + t for t in gen]
    costs = [shard.utterance_cost(t, f) for t, f in zip(tot, gen)]
    one = shard.assign_batches(costs, tot, 1)[0]
    for n in (2, 4, 8):
        per_rank = shard.assign_batches(costs, tot, n)
        assert sorted(u for r in per_rank for b in r for u in b) == list(range(2048))
        assert sum(len(r) for r in per_rank) == len(one)
        loads = [sum(costs[u] for b in r for u in b) for r in per_rank]
        assert max(loads) / (sum(loads) / n) < 1.05
        old = sum(len(shard.make_batches(m, tot)) for m in shard.lpt_assign(costs, n))
        assert old > len(one)
        for r in per_rank:
            for b in r:
                assert len({(tot[u] + 8 + 255) // 256 * 256 for u in b}) == 1
    # 3 batches cannot be balanced over 2 ranks: utterance-level fallback (each rank batches its own share)
    few_tot, few_cost = [1000] * 96 + [4000] * 8, [1.0] * 96 + [16.0] * 8
    per_rank = shard.assign_batches(few_cost, few_tot, 2, row_budget=48 * 1024)
    assert sorted(u for r in per_rank for b in r for u in b) == list(range(104))
    loads = [sum(few_cost[u] for b in r for u in b) for r in per_rank]
    assert max(loads) / (sum(loads) / 2) < 1.05
