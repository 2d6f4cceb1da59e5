"""Where a streaming chunk's time goes (BASELINE config 3, builder probe): host wall time and device time of flow.inference
(front-end + causal CFM) and hift.inference inside CosyVoice2Model.token2wav's call, per chunk.
    python tools/stream_breakdown.py [--hop 25] [--chunks 10]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from cosyvoice_b200 import synthetic
from cosyvoice_b200.frontend import B200CausalMaskedDiffWithXvec
from cosyvoice_b200.hift import B200HiFTGenerator
from cosyvoice_b200.token2wav import Token2Wav2

ap = argparse.ArgumentParser()
ap.add_argument("--hop", type=int, default=25)
ap.add_argument("--chunks", type=int, default=10)
a = ap.parse_args()
dev = torch.device("cuda:0")
flow = B200CausalMaskedDiffWithXvec(synthetic.flow_frontend_sd(True, seed=2), dtype=torch.bfloat16, device=dev,
                                    estimator_state_dict=synthetic.flow_estimator_sd(True, seed=0))
hift = B200HiFTGenerator(synthetic.hift_sd(True, seed=1), v2=True, dtype=torch.bfloat16, device=dev)
m = Token2Wav2(flow, hift, device=dev)
flow.encoder.static_chunk_size = 2 * flow.input_frame_rate
tok, ptok, pfeat, emb = synthetic.frontend_inputs(True, a.chunks * a.hop + 3, 75, seed=7)
rec = {}


def wrap(obj, name, key):
    fn = getattr(obj, name)

    def w(*args, **kw):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        out = fn(*args, **kw)
        e1.record()
        rec.setdefault(key, []).append((time.perf_counter() - t0, e0, e1))
        return out
    setattr(obj, name, w)


wrap(flow, "inference", "flow")
wrap(hift, "inference", "hift")
for rep in range(4):
    rec.clear()
    m.open_session("s")
    walls = []
    for k in range(a.chunks):
        n = (k + 1) * a.hop + 3
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        wav = m.token2wav(tok[:, :n], ptok, pfeat, emb, "s", k * a.hop, finalize=False).cpu()
        walls.append((time.perf_counter() - t0) * 1e3)
    m.close_session("s")
torch.cuda.synchronize()
for k in range(a.chunks):
    fh, f0, f1 = rec["flow"][k]
    hh, h0, h1 = rec["hift"][k]
    print(f"chunk {k}: wall {walls[k]:6.2f} ms | flow host {fh * 1e3:5.2f} device {f0.elapsed_time(f1):6.2f} | "
          f"hift host {hh * 1e3:5.2f} device {h0.elapsed_time(h1):5.2f} | rest {walls[k] - f0.elapsed_time(f1) - h0.elapsed_time(h1):5.2f}")
