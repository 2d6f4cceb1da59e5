"""`bench.py --workload sweep`: BASELINE config 5 as a bench workload (builder-run; the driver uses the default workload).

`--utterances` utterances of 2-30 s (+ 3 s prompt) go from speech tokens to waveforms ON THE HOST; the total work is fixed, so
"scaling" is "strong".  Whole-job batches are cut per stride bucket and assigned to ranks by LPT over their modelled cost
(shard.assign_batches); there is no collective on the data path.  A step = one pass over the rank's share through
AcousticBackend.synthesize_tokens_batches_to_host (pinned results, device->host copies on a side stream under the next
batch).  Time = max over ranks of the CUDA-event time of K passes / K; `e2e` uses the host wall clock of the same passes.
"""
import json
import os
import time

import torch


def run_sweep(args, bench):
    import torch.distributed as dist
    from cosyvoice_b200 import _lib, shard, synthetic
    from cosyvoice_b200.pipeline import AcousticBackend

    SR, PROMPT_FRAMES, PROMPT_TOKENS = bench.SR, bench.PROMPT_FRAMES, bench.PROMPT_TOKENS
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: cosyvoice_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, K, W = args.utterances, args.steps, max(args.warmup, 1)
    g = torch.Generator().manual_seed(2048)
    dur = 2.0 + 28.0 * torch.rand(N, generator=g)
    n_tok = [int(round(25.0 * float(d))) for d in dur]                       # 25 speech tokens per second
    gen_frames = [int(n / 25 * SR / 256) for n in n_tok]
    tot = [PROMPT_FRAMES + f for f in gen_frames]
    costs = [shard.utterance_cost(t, f) for t, f in zip(tot, gen_frames)]
    batches = shard.assign_batches(costs, tot, world, row_budget=48 * 1024, align=256, max_batch=64)[rank]
    be = AcousticBackend(synthetic.flow_estimator_sd(False, seed=0), synthetic.hift_sd(False, seed=1), v2=False,
                         dtype=torch.bfloat16, device=dev, frontend_sd=synthetic.flow_frontend_sd(False, seed=2))
    # host inputs of the rank's share, built once (pageable host tensors, as a caller would hold them)
    inputs = {}
    for b in batches:
        for u in b:
            t, pt, pf, e = synthetic.frontend_inputs(False, n_tok[u], PROMPT_TOKENS, seed=1234 + u, sample_rate=SR)
            inputs[u] = (t, pt, pf, e, torch.randn(1, 80, tot[u], generator=torch.Generator().manual_seed(4321 + u)))

    def one_pass():
        out = be.synthesize_tokens_batches_to_host(tuple([inputs[u][k] for u in b] for k in range(5)) for b in batches)
        return sum(int(v.shape[-1]) for views in out for v in views)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        one_pass()
    sync()
    sampler = bench.ClockSampler(local)
    sampler.start()
    n0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    host_enq, host_wait, n_samples = 0.0, 0.0, 0
    for _ in range(K):
        n_samples = one_pass()
        host_enq += be.last_sweep["host_enqueue_s"]
        host_wait += be.last_sweep["host_wait_s"]
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    sync()
    clocks = sampler.stop()
    launches = _lib.launch_count() - n0
    stats = torch.tensor([e0.elapsed_time(e1) / K, wall * 1e3 / K, host_enq / K, host_wait / K, float(n_samples),
                          float(len(batches))], dtype=torch.float64, device=dev)
    mx, sm = stats.clone(), stats.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    if rank == 0:
        ms = mx[0].item()
        audio_s = sm[4].item() / SR
        h2d = sum(a.numel() * a.element_size() for v in inputs.values() for a in v)
        print(json.dumps({
            "metric": "audio_seconds_per_second", "value": audio_s / (ms * 1e-3), "unit": "audio-s/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"{N} utterances, generated 2-30 s + 3 s prompt, v1 flow.inference + HiFT from speech tokens to "
                                   "host waveforms, whole-job batches per stride bucket assigned to ranks by LPT (shard.assign_batches)",
                       "utterances": N, "batches_total": int(sm[5].item()), "parallelism": f"batch-sharded x{world}",
                       "l2": "inputs_larger_than_L2", "audio_seconds": audio_s,
                       "host_enqueue_s_per_pass_max_rank": mx[2].item(), "host_wait_s_per_pass_max_rank": mx[3].item(),
                       "wall_ms_per_pass_max_rank": mx[1].item()},
            "e2e": {"value": audio_s / (mx[1].item() * 1e-3), "unit": "audio-s/s", "h2d_bytes_per_step_rank0": h2d,
                    "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": int(sm[4].item()) * 4, "ms_per_step": mx[1].item()},
            "gpu_launches": int(launches), "clocks": clocks,
        }))
    if world > 1:
        dist.destroy_process_group()
