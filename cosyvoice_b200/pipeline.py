"""Flow decoder + HiFT vocoder as one batched call: the part of CosyVoiceModel.token2wav
(cosyvoice/cli/model.py:130-175) that lies inside the hot path -- `flow.decoder(...)` on the encoder
output followed by `hift.inference(...)` on the generated frames -- for a batch of independent utterances.
"""
from typing import List, Optional, Sequence

import torch

from . import ops
from .flow import B200CausalConditionalCFM, B200ConditionalCFM, B200Estimator, N_FEATS
from .hift import B200HiFTGenerator


class AcousticBackend:
    def __init__(self, flow_sd, hift_sd, v2: bool = False, dtype: torch.dtype = torch.bfloat16, device="cuda",
                 hift_dtype: Optional[torch.dtype] = None):
        self.device = torch.device(device)
        self.v2 = v2
        est = B200Estimator(flow_sd, causal=v2, dtype=dtype, device=device)
        self.cfm = (B200CausalConditionalCFM if v2 else B200ConditionalCFM)(est)
        self.hift = B200HiFTGenerator(hift_sd, v2=v2, dtype=hift_dtype or dtype, device=device)
        self.sample_rate = 24000

    @torch.inference_mode()
    def synthesize_batch(self, mus: Sequence[torch.Tensor], spks: Sequence[torch.Tensor], conds: Sequence[torch.Tensor],
                         zs: Sequence[torch.Tensor], prompt_lens: Sequence[int], n_timesteps: int = 10) -> List[torch.Tensor]:
        """Per utterance u: mu/cond/z [1,80,T_u] (host or device), spks [1,80]; the first prompt_lens[u] frames are
        the prompt (flow/flow.py:144-158: they are generated and then dropped).  Returns waveforms [1, L_u] on device."""
        mels = self.cfm.solve_batch(mus, spks, conds, zs, n_timesteps)
        gen = [m[:, :, p:] for m, p in zip(mels, prompt_lens)]
        wavs, _ = self.hift.inference_batch(gen, uniform_stride=True)
        return wavs

    # ---- device-resident variant used by bench.py's `value` leg ---------------------------------
    def prepare_resident(self, mus, spks, conds, zs, prompt_lens):
        dev = self.device
        lens = [int(m.shape[-1]) for m in mus]
        B, T_max = len(lens), max(lens)
        lay1 = self.cfm.estimator.layout_for(lens)
        prog, st = self.cfm._state(lay1)

        def stack(seq):
            buf = torch.zeros(B, N_FEATS, T_max, dtype=torch.float32, device=dev)
            for b, a in enumerate(seq):
                buf[b, :, :lens[b]] = a.reshape(N_FEATS, lens[b]).to(dev, torch.float32)
            return buf
        z_rows = torch.zeros_like(st["x"])
        ops.nct_to_rows(stack(zs), z_rows, lay1)
        ops.nct_to_rows(stack(mus), st["mu"], lay1)
        ops.nct_to_rows(stack(conds), st["cond"], lay1)
        st["spks"].copy_(torch.cat([s.reshape(1, N_FEATS) for s in spks], 0).to(dev, torch.float32))
        gen_lens = [l - p for l, p in zip(lens, prompt_lens)]
        hprog = self.hift.program(gen_lens)
        mel_nct = torch.zeros(B, N_FEATS, T_max, dtype=torch.float32, device=dev)
        return dict(lay1=lay1, st=st, z_rows=z_rows, lens=lens, prompt_lens=list(prompt_lens), gen_lens=gen_lens,
                    hprog=hprog, mel_nct=mel_nct, same_prompt=len(set(prompt_lens)) == 1)

    @torch.inference_mode()
    def step_resident(self, r: dict, n_timesteps: int = 10):
        """One pass over the batch with every input already in HBM (noise, mu, cond, spks)."""
        st, lay1, hprog = r["st"], r["lay1"], r["hprog"]
        st["x"].copy_(r["z_rows"])
        self.cfm.solve_rows(lay1, n_timesteps)
        ops.rows_to_nct(st["x"], r["mel_nct"], lay1)
        assert r["same_prompt"], "resident bench path assumes one prompt length"
        p = r["prompt_lens"][0]
        gen = r["mel_nct"][:, :, p:]
        ops.nct_to_rows(gen, hprog.mel32, hprog.lay0)
        if hprog.mel is not hprog.mel32:
            ops.nct_to_rows(gen, hprog.mel, hprog.lay0)
        self.hift.draw_source_rng(hprog)
        for op in hprog.ops_f0:
            op()
        for op in hprog.ops_src:
            op()
        hprog.run()
        return hprog.wav
