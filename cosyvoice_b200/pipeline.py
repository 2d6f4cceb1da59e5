"""Flow + HiFT vocoder as one batched call: the part of CosyVoiceModel.token2wav (cosyvoice/cli/model.py:130-175) that
lies inside the hot path -- `flow.inference(...)` (front-end + CFM; or `flow.decoder(...)` alone on a given encoder output)
followed by `hift.inference(...)` on the generated frames -- for a batch of independent utterances.
"""
from typing import List, Optional, Sequence

import torch

from . import ops
from .flow import B200CausalConditionalCFM, B200ConditionalCFM, B200Estimator, N_FEATS
from .hift import B200HiFTGenerator


class AcousticBackend:
    def __init__(self, flow_sd, hift_sd, v2: bool = False, dtype: torch.dtype = torch.bfloat16, device="cuda",
                 hift_dtype: Optional[torch.dtype] = None, frontend_sd=None):
        """flow_sd: estimator parameters (`decoder.estimator.*` stripped); frontend_sd: the rest of the flow module
        (embedding, encoder, length regulator, speaker projection) -- enables the token entry points."""
        self.device = torch.device(device)
        self.v2 = v2
        if frontend_sd is not None:
            from .frontend import B200CausalMaskedDiffWithXvec, B200MaskedDiffWithXvec
            cls = B200CausalMaskedDiffWithXvec if v2 else B200MaskedDiffWithXvec
            self.flow = cls(frontend_sd, dtype=dtype, device=device, estimator_state_dict=flow_sd)
            self.cfm = self.flow.decoder
        else:
            self.flow = None
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

    @torch.inference_mode()
    def synthesize_tokens_batch(self, tokens: Sequence[torch.Tensor], prompt_tokens: Sequence[torch.Tensor],
                                prompt_feats: Sequence[torch.Tensor], embeddings: Sequence[torch.Tensor],
                                zs: Optional[Sequence[torch.Tensor]] = None, n_timesteps: int = 10,
                                sample_rate: Optional[int] = None) -> List[torch.Tensor]:
        """The whole back-end from speech tokens (host or device): flow.inference (front-end + CFM) per utterance, then
        hift.inference on the generated frames (cli/model.py:137-161).  Returns waveforms [1, L_u] on device."""
        if self.flow is None:
            raise RuntimeError("AcousticBackend was built without frontend_sd")
        gen = self.flow.inference_batch(tokens, prompt_tokens, prompt_feats, embeddings, zs, n_timesteps,
                                        sample_rate or self.sample_rate)
        wavs, _ = self.hift.inference_batch(gen, uniform_stride=True)
        return wavs

    @torch.inference_mode()
    def synthesize_tokens_batches_to_host(self, batches, n_timesteps: int = 10, sample_rate: Optional[int] = None):
        """A sweep over many batches (the batched replacement of the reference's one-utterance loop,
        cosyvoice/bin/inference.py:57,79) with the results on the HOST: `batches` yields
        (tokens, prompt_tokens, prompt_feats, embeddings, zs) lists; returns, per batch, the waveforms [1, L_u] as views of
        one pinned host buffer.  The device->host copies of batch i run on a side stream into pinned memory while the host
        packs and enqueues batch i+1, so neither the copy nor the packing sits between two batches on the GPU; the only
        host waits are the bounded run-ahead (three batches) and the end.  Host inputs go up through pinned staging
        (layout.to_device_async), so no copy synchronises the stream.  `self.last_sweep` records the host time spent
        packing + enqueueing (waits included) and the final wait."""
        import time
        if not hasattr(self, "_d2h_stream"):
            self._d2h_stream = torch.cuda.Stream(device=self.device)
        side, main = self._d2h_stream, torch.cuda.current_stream(self.device)
        out, events = [], []
        t_host = 0.0
        t0 = time.perf_counter()
        for batch in batches:
            if len(events) >= 3:
                events[-3].synchronize()          # bounded run-ahead: at most three batches of results in flight
            wavs = self.synthesize_tokens_batch(*batch, n_timesteps=n_timesteps, sample_rate=sample_rate)
            ready = torch.cuda.Event()
            ready.record(main)
            sizes = [int(w.shape[-1]) for w in wavs]
            arena = torch.empty(sum(sizes), dtype=torch.float32, pin_memory=True)
            views, o = [], 0
            with torch.cuda.stream(side):
                side.wait_event(ready)
                for w, n in zip(wavs, sizes):
                    v = arena[o:o + n].view(1, n)
                    v.copy_(w, non_blocking=True)
                    w.record_stream(side)
                    views.append(v)
                    o += n
                done = torch.cuda.Event()
                done.record(side)
            out.append(views)
            events.append(done)
        t_host = time.perf_counter() - t0
        for e in events:
            e.synchronize()
        self.last_sweep = dict(host_enqueue_s=t_host, host_wait_s=time.perf_counter() - t0 - t_host, batches=len(out))
        return out

    def prepare_resident_tokens(self, tokens, prompt_tokens, prompt_feats, embeddings, zs, sample_rate: Optional[int] = None):
        """Resident state for step_resident with the front-end inside the step: tokens, prompt mels, speaker vectors and
        noise are copied to HBM here, once."""
        sr = sample_rate or self.sample_rate
        f, dev = self.flow, self.device
        B = len(tokens)
        n_prompt = [int(p.numel()) for p in prompt_tokens]
        mel_lens = [f._mel_lens(n_prompt[b], int(tokens[b].numel()), int(prompt_feats[b].shape[-2]), sr) for b in range(B)]
        cat = [torch.cat([prompt_tokens[b].reshape(-1).to(torch.int32), tokens[b].reshape(-1).to(torch.int32)]) for b in range(B)]
        fprog = f.encode_batch(cat, n_prompt, mel_lens, embeddings, sr)
        lay1 = fprog.mel
        lens = [a + b for a, b in mel_lens]
        prog, st = self.cfm._state(lay1)
        st["cond"].zero_()
        z_rows = torch.zeros_like(st["x"])
        for b in range(B):
            r0, l1 = lay1.row0[b], mel_lens[b][0]
            st["cond"][r0:r0 + l1].copy_(prompt_feats[b].reshape(-1, N_FEATS)[:l1].to(dev, torch.float32))
            z_rows[r0:r0 + lens[b]].copy_(zs[b].reshape(N_FEATS, lens[b]).to(dev, torch.float32).t())
        prompt_lens = [m[0] for m in mel_lens]
        gen_lens = [m[1] for m in mel_lens]
        hprog = self.hift.program(gen_lens)
        mel_nct = torch.zeros(B, N_FEATS, max(lens), dtype=torch.float32, device=dev)
        return dict(lay1=lay1, st=st, z_rows=z_rows, lens=lens, prompt_lens=prompt_lens, gen_lens=gen_lens, hprog=hprog,
                    mel_nct=mel_nct, same_prompt=len(set(prompt_lens)) == 1, fprog=fprog)

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
        with self.cfm.estimator.serial, self.hift.serial:
            return self._step_resident_locked(r, n_timesteps)

    def _step_resident_locked(self, r: dict, n_timesteps: int):
        st, lay1, hprog = r["st"], r["lay1"], r["hprog"]
        if r.get("fprog") is not None:          # front-end inside the step: tokens -> mu, spks (flow/flow.py:126-141)
            r["fprog"].run()
            st["mu"].copy_(r["fprog"].mu)
            st["spks"].copy_(r["fprog"].spks)
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
