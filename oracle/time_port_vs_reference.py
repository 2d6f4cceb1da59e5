"""Container-only check (needs /root/reference): is the oracle PORT as fast as the reference's stock CPU path?

Test infrastructure.  `bench.py --impl reference` and `cpu_baseline` time the port (the reference itself cannot travel to
the GPU box), so the port must cost what the reference costs.  This script times, on identical inputs and weights at the
bench size (T = 1218, one utterance):
  * the REAL `ConditionalCFM.solve_euler` (cosyvoice/flow/flow_matching.py:72-124 over the real ConditionalDecoder, whose
    attention goes through F.scaled_dot_product_attention, components/transformer.py:266-271) against
    `oracle.flow_ref.solve_euler`;
  * the REAL `HiFTGenerator.inference` (cosyvoice/hifigan/generator.py:399-411) against `oracle.hift_ref.hift_inference`;
1 warm-up + median of 3 each, fixed thread count, flush-denormal on, and asserts that the port is
within 10 % of the reference on the sum.  Result lines go to BASELINE.md section 2.

    python -m oracle.time_port_vs_reference [--threads 8] [--steps 10]
"""
import argparse
import json
import statistics
import sys
import time
import warnings

import torch


def _median_time(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return statistics.median(ts), ts


def measure(threads: int = 8, n_steps: int = 10, T: int = 1218, gen_frames: int = 937):
    from cosyvoice_b200 import synthetic as W
    from oracle import flow_ref, hift_ref, ref_import as R
    if not R.available():
        raise RuntimeError("needs the reference tree (/root/reference): build container only")
    torch.set_num_threads(threads)
    torch.set_flush_denormal(True)
    g = torch.Generator().manual_seed(7)
    mu, z = torch.randn(1, 80, T, generator=g), torch.randn(1, 80, T, generator=g)
    spks = torch.randn(1, 80, generator=g)
    cond = torch.zeros(1, 80, T)
    cond[:, :, :281] = torch.randn(1, 80, 281, generator=g) * 2 - 4
    mask = torch.ones(1, 1, T)
    fsd = W.flow_estimator_sd(False, seed=0)
    est = R.build_estimator(False).eval()
    est.load_state_dict(fsd, strict=True)
    cfm = R.build_cfm(False, est).eval()
    t_span = flow_ref.cosine_t_span(n_steps)
    outs = {}

    def ref_flow():
        with torch.inference_mode():
            outs["ref"] = cfm.solve_euler(z.clone(), t_span=t_span, mu=mu, mask=mask, spks=spks, cond=cond)

    def port_flow():
        with torch.inference_mode():
            outs["port"] = flow_ref.solve_euler(fsd, z, t_span, mu, mask, spks, cond)

    t_ref_flow, raw_rf = _median_time(ref_flow)
    t_port_flow, raw_pf = _median_time(port_flow)
    flow_err = (outs["ref"] - outs["port"]).abs().max().item()

    hsd = W.hift_sd(False, seed=1)
    hift = R.build_hift(False).eval()
    hift.load_state_dict(hsd, strict=True)
    mel = torch.randn(1, 80, gen_frames, generator=g) * 2 - 4
    pv, nz = hift_ref.draw_source_rng(1, gen_frames * 256)

    def ref_hift():
        with torch.inference_mode():
            torch.manual_seed(3)
            outs["ref_w"] = hift.inference(mel)[0]

    def port_hift():
        with torch.inference_mode():
            outs["port_w"] = hift_ref.hift_inference(hsd, mel, pv, nz)[0]

    t_ref_hift, raw_rh = _median_time(ref_hift)
    t_port_hift, raw_ph = _median_time(port_hift)
    return dict(threads=threads, T=T, n_steps=n_steps,
                ref_flow_s=t_ref_flow, port_flow_s=t_port_flow, flow_max_abs_diff=flow_err,
                ref_hift_s=t_ref_hift, port_hift_s=t_port_hift,
                ref_total_s=t_ref_flow + t_ref_hift, port_total_s=t_port_flow + t_port_hift,
                port_over_ref=(t_port_flow + t_port_hift) / (t_ref_flow + t_ref_hift),
                raw=dict(ref_flow=raw_rf, port_flow=raw_pf, ref_hift=raw_rh, port_hift=raw_ph))


if __name__ == "__main__":
    warnings.filterwarnings("ignore")
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    r = measure(a.threads, a.steps)
    print(json.dumps(r))
    ok = 0.9 <= r["port_over_ref"] <= 1.1
    print("port / reference =", round(r["port_over_ref"], 3), "OK" if ok else "OUTSIDE 10 %")
    sys.exit(0 if ok else 1)
