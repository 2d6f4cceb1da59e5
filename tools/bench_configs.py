"""Measurement of BASELINE.json configs 3, 4 and 5 -- bench.py itself measures config 2.

    python tools/bench_configs.py stream  [--hop 25] [--chunks 10] [--reps 5]
    python tools/bench_configs.py bigvgan [--batch 64] [--frames 94,938] [--topology 24k100]
    python tools/bench_configs.py sweep   [--utterances 2048] [--gpus N]      (torchrun for N>1)

Each sub-command prints JSON lines (one per measured case).  Synthetic inputs, random-init weights of the named
architecture (cosyvoice_b200.synthetic); CUDA-event timing after warm-up; no /root/reference at run time.
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from cosyvoice_b200 import _lib, shard, synthetic  # noqa: E402


def _peaks():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return float(d.get("bf16_tflops_sustained", 1392.1)), float(d.get("hbm_gbs", 6546.6))
    except Exception:
        return 1392.1, 6546.6


def _ev():
    return torch.cuda.Event(enable_timing=True)


def _timed(fn, reps):
    """Median / all device times (ms) of fn() over `reps` runs, CUDA events on the current stream."""
    out = []
    for _ in range(reps):
        a, b = _ev(), _ev()
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b))
    return out


def _pct(xs, q):
    xs = sorted(xs)
    k = min(len(xs) - 1, max(0, int(round(q * (len(xs) - 1)))))
    return xs[k]


# ----------------------------------------------------------------------------- config 3: v2 streaming chunks
def run_stream(args):
    """CosyVoice2 streaming (cosyvoice/cli/model.py:296-349): chunk k re-runs the causal flow on
    prompt + (k+1)*hop + 3 tokens (mel T = 2*tokens) and vocodes the 2*(hop+3) new frames (+8 cached frames
    and their cached source after the first chunk).  Latency = host wall clock from the call to the chunk
    waveform being on the host, inputs (encoder output mu, prompt cond, spk) already on the device as they are
    in the reference."""
    from cosyvoice_b200.flow import B200CausalConditionalCFM, B200Estimator
    from cosyvoice_b200.hift import B200HiFTGenerator
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    est = B200Estimator(synthetic.flow_estimator_sd(True, seed=0), causal=True, dtype=dt, device=dev)
    cfm = B200CausalConditionalCFM(est)
    cfm.use_cuda_graph = not args.eager
    hift = B200HiFTGenerator(synthetic.hift_sd(True, seed=1), v2=True, dtype=dt, device=dev)
    hop, prompt_tok, look = args.hop, 75, 3
    mel_cache, hop_samples = 8, hift.hop
    g = torch.Generator().manual_seed(7)
    T_full = 2 * (prompt_tok + args.chunks * hop + look)
    mu_full = torch.randn(1, 80, T_full, generator=g).to(dev)
    spk = torch.randn(1, 80, generator=g).to(dev)
    prompt_mel = (torch.randn(1, 80, 2 * prompt_tok, generator=g) * 2 - 4).to(dev)

    def session(record):
        cache_mel, cache_src = None, torch.zeros(1, 1, 0, device=dev)
        for k in range(args.chunks):
            T = 2 * (prompt_tok + (k + 1) * hop + look)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0, e1 = _ev(), _ev()
            e0.record()
            mu = mu_full[:, :, :T].clone()
            cond = torch.zeros(1, 80, T, device=dev)
            cond[:, :, :2 * prompt_tok] = prompt_mel
            mask = torch.ones(1, 1, T, device=dev)
            mel, _ = cfm(mu, mask, 10, spks=spk, cond=cond)
            new = mel[:, :, 2 * prompt_tok + 2 * k * hop:]
            e1.record()
            feat = new if cache_mel is None else torch.cat([cache_mel, new], dim=2)
            wav, src = hift.inference(feat, cache_src)
            wav_host = wav.cpu()
            t2 = time.perf_counter()
            cache_mel = feat[:, :, -mel_cache:]
            cache_src = src[:, :, -mel_cache * hop_samples:]
            record(k, T, (t2 - t0) * 1e3, e0.elapsed_time(e1), wav_host.shape[-1])

    cold = {}
    session(lambda k, T, ms, fms, n: cold.__setitem__(k, ms))            # first sight of every shape: builds + captures
    session(lambda *a: None)
    warm = {k: [] for k in range(args.chunks)}
    flow_ms = {k: [] for k in range(args.chunks)}
    meta = {}

    def rec(k, T, ms, fms, n):
        warm[k].append(ms)
        flow_ms[k].append(fms)
        meta[k] = (T, n)
    for _ in range(args.reps):
        session(rec)
    allv = [v for k in warm for v in warm[k]]
    line = {
        "config": 3, "workload": f"CosyVoice2 causal flow + HiFT streaming chunks, hop {hop} tokens, prompt {prompt_tok} tokens, "
                                 f"{args.chunks} chunks x {args.reps} sessions", "dtype": args.dtype,
        "metric": "chunk_latency_ms", "p50": _pct(allv, 0.5), "p99": _pct(allv, 0.99), "max": max(allv),
        "per_chunk": [{"k": k, "T": meta[k][0], "samples": meta[k][1], "p50_ms": _pct(warm[k], 0.5),
                       "flow_device_p50_ms": _pct(flow_ms[k], 0.5), "cold_ms": cold[k]} for k in range(args.chunks)],
        "audio_ms_per_chunk": 1000.0 * 2 * hop * hop_samples / 24000.0,
        "graphs_cached": len(cfm._graphs), "kernel_launches_per_flow_call": cfm.last_graph_launches,
    }
    print(json.dumps(line))


def run_stream_tokens(args):
    """Config 3 through the reference-facing call: CosyVoice2Model.token2wav (cli/model.py:334-366) per chunk, i.e. the
    whole flow.inference (token embedding, chunk-masked UpsampleConformerEncoder, causal CFM) + HiFT with its caches, from
    speech tokens on the host to the chunk waveform on the host."""
    from cosyvoice_b200.frontend import B200CausalMaskedDiffWithXvec
    from cosyvoice_b200.hift import B200HiFTGenerator
    from cosyvoice_b200.token2wav import Token2Wav2
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    flow = B200CausalMaskedDiffWithXvec(synthetic.flow_frontend_sd(True, seed=2), dtype=dt, device=dev,
                                        estimator_state_dict=synthetic.flow_estimator_sd(True, seed=0))
    flow.decoder.use_cuda_graph = not args.eager
    hift = B200HiFTGenerator(synthetic.hift_sd(True, seed=1), v2=True, dtype=dt, device=dev)
    m = Token2Wav2(flow, hift, device=dev)
    flow.encoder.static_chunk_size = 2 * flow.input_frame_rate          # cli/model.py:314
    hop, prompt_tok, look = args.hop, 75, 3
    tok, ptok, pfeat, emb = synthetic.frontend_inputs(True, args.chunks * hop + look, prompt_tok, seed=7)

    def session(record):
        m.open_session("s")
        for k in range(args.chunks):
            n = (k + 1) * hop + look
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            wav = m.token2wav(tok[:, :n], ptok, pfeat, emb, "s", k * hop, finalize=False)
            wav_host = wav.cpu()
            record(k, 2 * (prompt_tok + n), (time.perf_counter() - t0) * 1e3, wav_host.shape[-1])
        m.close_session("s")

    cold = {}
    session(lambda k, T, ms, n: cold.__setitem__(k, ms))
    session(lambda *a: None)
    warm = {k: [] for k in range(args.chunks)}
    meta = {}

    def rec(k, T, ms, n):
        warm[k].append(ms)
        meta[k] = (T, n)
    for _ in range(args.reps):
        session(rec)
    allv = [v for k in warm for v in warm[k]]
    # front-end alone (device time) at the last chunk's length
    n = args.chunks * hop + look
    cat = torch.cat([ptok, tok[:, :n]], 1).reshape(-1)
    prog = flow.encode_batch([cat], [prompt_tok], [(2 * prompt_tok, 2 * n)], [emb])
    front = _timed(prog.run, 10)
    print(json.dumps({
        "config": 3, "workload": f"CosyVoice2Model.token2wav per chunk from speech tokens (front-end + causal CFM + HiFT + caches), "
                                 f"hop {hop} tokens, prompt {prompt_tok} tokens, {args.chunks} chunks x {args.reps} sessions",
        "dtype": args.dtype, "metric": "chunk_latency_ms", "p50": _pct(allv, 0.5), "p99": _pct(allv, 0.99), "max": max(allv),
        "per_chunk": [{"k": k, "T": meta[k][0], "samples": meta[k][1], "p50_ms": _pct(warm[k], 0.5), "cold_ms": cold[k]}
                      for k in range(args.chunks)],
        "audio_ms_per_chunk": 1000.0 * 2 * hop * hift.hop / 24000.0,
        "frontend_device_ms_last_chunk": _pct(front, 0.5), "frontend_launches": len(prog.ops) + 2,
    }))


# ----------------------------------------------------------------------------- config 4: BigVGAN
def run_bigvgan(args):
    """BigVGAN trunk (BigVGAN/bigvgan.py:406-436) on `batch` synthetic mels N(0,1)*2-4 of the given frame counts,
    plus the anti-aliased SnakeBeta kernel K1 (alias_free_activation/cuda/anti_alias_activation_cuda.cu) stand-alone
    on [batch, C, T] fp32 for every (C, T) the trunk produces, against the HBM roofline (read + write once)."""
    from cosyvoice_b200 import ops
    from cosyvoice_b200.bigvgan import B200BigVGAN, kaiser_sinc_filter12
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    tf_peak, hbm_peak = _peaks()
    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    gen = B200BigVGAN(synthetic.bigvgan_sd(args.topology, seed=3), args.topology, dtype=dt, device=dev)
    a = gen.arch
    B = args.batch
    g = torch.Generator().manual_seed(5)
    for frames in [int(f) for f in args.frames.split(",")]:
        feat = (torch.randn(B, a["in_ch"], frames, generator=g) * 2 - 4).pin_memory()
        spk = torch.randn(B, a["spk"], 1, generator=g).pin_memory()
        for _ in range(2):
            wav = gen.inference(feat, spk)
        n0 = _lib.launch_count()
        ms = _timed(lambda: gen.inference(feat, spk).cpu(), args.reps)        # public call: H2D + trunk + D2H
        launches = (_lib.launch_count() - n0) // args.reps
        prog = gen.program([frames] * B)
        dms = _timed(prog.run, args.reps)                                     # device-resident trunk only
        med, dmed = _pct(ms, 0.5), _pct(dms, 0.5)
        audio_s = B * frames * prog.P / 24000.0
        print(json.dumps({
            "config": 4, "workload": f"BigVGAN {args.topology} trunk, batch {B} x {frames} frames", "dtype": args.dtype,
            "metric": "audio_seconds_per_second", "value": audio_s / (dmed * 1e-3), "e2e": audio_s / (med * 1e-3),
            "ms_resident": dmed, "ms_e2e": med, "gpu_launches": launches,
            "algorithmic_tflop": prog.flops / 1e12, "achieved_tflops": prog.flops / (dmed * 1e-3) / 1e12,
            "frac_of_bf16_peak": prog.flops / (dmed * 1e-3) / 1e12 / tf_peak,
            "wav_abs_max": float(wav.abs().max().item())}))
        del prog, wav
        gen._programs.clear()
        torch.cuda.empty_cache()
    # K1 stand-alone, NCT fp32 in / fp32 out (the reference extension's fwd_cuda contract)
    filt = kaiser_sinc_filter12(dev)
    chs = [a["ch0"] // 2 ** (i + 1) for i in range(len(a["rates"]))]
    for frames in [int(f) for f in args.frames.split(",")]:
        P = 1
        for u, c in zip(a["rates"], chs):
            P *= u
            T = frames * P
            x = torch.randn(B, c, T, device=dev)
            al = torch.randn(c, device=dev) * 0.3
            be = torch.randn(c, device=dev) * 0.3
            fn = lambda: ops.anti_alias_act_nct(x, filt, filt, al, be)
            fn()
            ms = _pct(_timed(fn, max(3, args.reps)), 0.5)
            gb = 2 * x.numel() * 4 / 1e9
            print(json.dumps({"config": 4, "kernel": "anti_alias_act_nct (K1)", "shape": [B, c, T], "ms": ms,
                              "algorithmic_GB": gb, "achieved_GBps": gb / (ms * 1e-3), "hbm_peak_GBps": hbm_peak,
                              "frac": gb / (ms * 1e-3) / hbm_peak}))
            del x


# ----------------------------------------------------------------------------- config 5: ragged sweep
def run_sweep(args):
    """2048 synthetic utterances, generated duration ~ U[2, 30] s after a 3 s prompt, v1 flow + HiFT, sharded by
    utterance over the ranks (shard.lpt_assign on the algorithmic-FLOP cost; no collective on the data path).
    Each rank runs its own length-bucketed batches through AcousticBackend.synthesize_batch (host inputs in,
    host waveforms out).  Time = max over ranks of the second pass (the first pass builds the per-bucket
    programs / CUDA graphs, as a warmed-up server would have)."""
    import torch.distributed as dist
    from cosyvoice_b200.pipeline import AcousticBackend
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    tf_peak, _ = _peaks()
    SR, HOP, PROMPT = 24000, 256, 281
    g = torch.Generator().manual_seed(2048)
    dur = 2.0 + 28.0 * torch.rand(args.utterances, generator=g)
    n_tok = [int(round(25.0 * float(d))) for d in dur]                    # 25 speech tokens per second
    gen_frames = [int(n / 25 * SR / 256) for n in n_tok] if not args.from_mu else [int(round(float(d) * SR / HOP)) for d in dur]
    tot = [PROMPT + f for f in gen_frames]
    costs = [shard.utterance_cost(t, f) for t, f in zip(tot, gen_frames)]
    if args.shard == "utterance":
        mine = shard.lpt_assign(costs, world)[rank]
        batches = shard.make_batches(mine, tot, row_budget=args.row_budget, align=256, max_batch=args.max_batch)
    else:
        batches = shard.assign_batches(costs, tot, world, row_budget=args.row_budget, align=256, max_batch=args.max_batch)[rank]
        mine = [u for b in batches for u in b]
    be = AcousticBackend(synthetic.flow_estimator_sd(False, seed=0), synthetic.hift_sd(False, seed=1), v2=False,
                         dtype=torch.bfloat16, device=dev,
                         frontend_sd=None if args.from_mu else synthetic.flow_frontend_sd(False, seed=2))
    be.cfm.use_cuda_graph = not args.eager

    def one_pass():
        n_samples = 0
        for b in batches:
            if args.from_mu:
                mus, spks, conds, zs = [], [], [], []
                for u in b:
                    m, s, c, z = synthetic.flow_inputs(1, gen_frames[u], PROMPT, first_index=u)
                    mus += m; spks += s; conds += c; zs += z
            else:
                toks, ptoks, pfeats, embs, zs = [], [], [], [], []
                for u in b:
                    t, pt, pf, e = synthetic.frontend_inputs(False, n_tok[u], 75, seed=1234 + u, sample_rate=SR)
                    toks.append(t), ptoks.append(pt), pfeats.append(pf), embs.append(e)
                    zs.append(torch.randn(1, 80, tot[u], generator=torch.Generator().manual_seed(4321 + u)))
            t_in = time.perf_counter()
            if args.from_mu:
                wavs = be.synthesize_batch(mus, spks, conds, zs, [PROMPT] * len(b))
            else:
                wavs = be.synthesize_tokens_batch(toks, ptoks, pfeats, embs, zs, sample_rate=SR)
            host = [w.cpu() for w in wavs]
            dt = time.perf_counter() - t_in
            one_pass.gpu_s += dt
            one_pass.per_batch.append(dict(n=len(b), max_frames=max(tot[u] for u in b), frames=sum(tot[u] for u in b),
                                           cost=sum(costs[u] for u in b), model=shard.batch_seconds([tot[u] for u in b], [gen_frames[u] for u in b]),
                                           seconds=dt))
            n_samples += sum(int(h.shape[-1]) for h in host)
        return n_samples
    one_pass.gpu_s = 0.0
    one_pass.per_batch = []

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    one_pass()                       # builds programs / graphs per stride bucket
    barrier()
    one_pass.gpu_s = 0.0
    one_pass.per_batch = []
    n0 = _lib.launch_count()
    t0 = time.perf_counter()
    n_samples = one_pass()
    barrier()
    wall = time.perf_counter() - t0
    stats = torch.tensor([one_pass.gpu_s, float(n_samples), sum(costs[u] for u in mine), float(len(batches))],
                         dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    else:
        mx, sm = stats, stats
    t = float(mx[0].item())          # synthesis time (input generation on the host excluded), max over ranks
    if rank == 0:
        audio_s = float(sm[1].item()) / SR
        print(json.dumps({
            "config": 5, "workload": f"{args.utterances} utterances, gen 2-30 s + 3 s prompt, v1 flow + HiFT "
                                     f"{'from the encoder output' if args.from_mu else 'from speech tokens (flow.inference + hift.inference)'}, "
                                     f"utterance-sharded x{world} ({args.shard}-level LPT)", "dtype": "bf16", "n_gpus": world,
            "metric": "audio_seconds_per_second", "value": audio_s / t, "seconds": t, "wall_with_input_synthesis": wall,
            "audio_seconds": audio_s, "batches_total": int(sm[3].item()), "cuda_graph": not args.eager,
            "algorithmic_tflop": float(sm[2].item()) / 1e12,
            "frac_of_bf16_peak_per_gpu": float(sm[2].item()) / 1e12 / t / world / tf_peak,
            "rank_imbalance_cost": float(mx[2].item()) * world / float(sm[2].item()),
            "gpu_launches_rank0": _lib.launch_count() - n0}))
    if args.dump_batches:
        with open(f"{args.dump_batches}.rank{rank}.json", "w") as f:
            json.dump(dict(rank=rank, seconds=one_pass.gpu_s, batches=one_pass.per_batch), f)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="cmd", required=True)
    s = sub.add_parser("stream")
    s.add_argument("--hop", type=int, default=25)
    s.add_argument("--chunks", type=int, default=10)
    s.add_argument("--reps", type=int, default=5)
    s.add_argument("--dtype", default="bf16")
    s.add_argument("--eager", action="store_true")
    s.add_argument("--from-tokens", action="store_true", help="measure through Token2Wav2.token2wav (front-end included)")
    b = sub.add_parser("bigvgan")
    b.add_argument("--batch", type=int, default=64)
    b.add_argument("--frames", default="94,938")
    b.add_argument("--topology", default="24k100")
    b.add_argument("--reps", type=int, default=3)
    b.add_argument("--dtype", default="bf16")
    w = sub.add_parser("sweep")
    w.add_argument("--utterances", type=int, default=2048)
    w.add_argument("--gpus", type=int, default=1)
    w.add_argument("--row-budget", type=int, default=48 * 1024)
    w.add_argument("--max-batch", type=int, default=64)
    w.add_argument("--eager", action="store_true")
    w.add_argument("--shard", default="utterance", choices=["batch", "utterance"],
                   help="batch: whole-job batches, LPT over batch costs (shard.assign_batches); utterance: LPT over utterances, batches per rank")
    w.add_argument("--dump-batches", default="", help="write per-batch (size, cost, seconds) of the timed pass to <path>.rank<r>.json")
    w.add_argument("--from-mu", action="store_true", help="start at the encoder output (the pre-front-end measurement)")
    args = ap.parse_args()
    _lib.load()
    if args.cmd == "stream" and args.from_tokens:
        run_stream_tokens(args)
        return
    {"stream": run_stream, "bigvgan": run_bigvgan, "sweep": run_sweep}[args.cmd](args)


if __name__ == "__main__":
    main()
