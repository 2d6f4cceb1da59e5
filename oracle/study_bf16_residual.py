"""Test infrastructure, not product.  Feasibility study (CPU, oracle only): how much error would a bf16 residual stream inside the transformer blocks add?
Emulation: round x to bf16 after each of the two residual adds of every block; everything else fp32."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
torch.set_flush_denormal(True)
from oracle import flow_ref, init_weights as W, encoder_ref
d = torch.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "frontend_v1.pt"), weights_only=True)
esd = W.flow_estimator_sd(False, seed=0)
c = d["cases"][0]
mu, spks, cond, z = c["mu"], c["spks"], c["cond"], c["z"]
T = mu.shape[2]
def run():
    return flow_ref.solve_euler(esd, z, flow_ref.cosine_t_span(10), mu, torch.ones(1, 1, T), spks, cond)
ref = run()
orig = flow_ref._transformer_block
def tb_bf16res(sd, p, x, attn_mask):
    C = x.shape[-1]
    rb = lambda t: t.to(torch.bfloat16).float()
    n = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"])
    x = rb(flow_ref._attention(sd, p + "attn1.", n, attn_mask) + x)
    n = F.layer_norm(x, (C,), sd[p + "norm3.weight"], sd[p + "norm3.bias"])
    h = F.gelu(F.linear(n, sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"]))
    h = F.linear(h, sd[p + "ff.net.2.weight"], sd[p + "ff.net.2.bias"])
    return rb(h + x)
flow_ref._transformer_block = tb_bf16res
got = run()
flow_ref._transformer_block = orig
rel = ((got - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()).item()
print(f"T={T}: bf16 residual stream alone: mel rel-rms {rel:.3e}, max-abs {(got - ref).abs().max().item():.3e} (bar: 3e-2 / 0.25; whole bf16 engine today: 6e-3 / 0.04)")
