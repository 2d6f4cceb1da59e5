"""CPU restatement of the streaming glue around the hot path -- TEST INFRASTRUCTURE, never on the product path.

Follows cosyvoice/cli/model.py:130-175 (CosyVoiceModel.token2wav), :334-366 (CosyVoice2Model.token2wav) and
cosyvoice/utils/common.py:149-157 (fade_in_out).  Pinned by tests/golden/token2wav.pt, which
oracle/make_golden.py generates by running the REAL reference classes over the deterministic stub flow / stub
vocoder defined at the bottom of this file (the stubs only use dyadic constants, so CPU and GPU agree bit for bit).
"""
from types import SimpleNamespace

import numpy as np
import torch


def fade_in_out_ref(fade_in: torch.Tensor, fade_out: torch.Tensor, window: np.ndarray) -> torch.Tensor:
    """utils/common.py:149-157.  fp32 tensor * float64 window promotes to double; the slice store rounds to fp32."""
    n = int(window.shape[0] / 2)
    out = fade_in.detach().cpu().clone()
    head = out[..., :n].double().numpy() * window[:n] + fade_out.detach().cpu()[..., -n:].double().numpy() * window[n:]
    out[..., :n] = torch.from_numpy(head).float()
    return out


def _fma(a: torch.Tensor, b: torch.Tensor, c) -> torch.Tensor:
    """fp32 fused multiply-add emulated in double (the product of two fp32 numbers is exact in double)."""
    c = c.double() if torch.is_tensor(c) else c
    return (a.double() * b.double() + c).float()


def interp_linear_ref(x: torch.Tensor, t_out: int) -> torch.Tensor:
    """F.interpolate(x, size=t_out, mode='linear'), align_corners=False, written out (model.py:171).  The FMA
    placement is the one ATen's CPU kernel compiles to on this image (src = fma(scale, j + 0.5, -0.5),
    out = fma(l0, x0, l1 * x1)): with it the restatement is bit-exact against tests/golden/token2wav.pt."""
    t_in = x.shape[2]
    j = torch.arange(t_out, dtype=torch.float32)
    scale = torch.tensor(t_in, dtype=torch.float32) / torch.tensor(t_out, dtype=torch.float32)
    src = _fma(j + 0.5, scale.expand_as(j), -0.5).clamp(min=0)
    i0 = src.floor().long().clamp(max=t_in - 1)
    i1 = (i0 + 1).clamp(max=t_in - 1)
    l1 = src - i0.float()
    l0 = 1 - l1
    return _fma(l0.expand_as(x[:, :, i0]), x[:, :, i0], l1 * x[:, :, i1])


class Token2WavRef:
    """model.py:29-69 (constants), :130-175 (token2wav), :184-188 / :233-238 (session dictionaries)."""

    def __init__(self, flow, hift, sr: int = 22050):
        self.flow, self.hift, self.sample_rate = flow, hift, sr
        self.mel_overlap_len = int(20 / flow.input_frame_rate * sr / 256)
        self.mel_window = np.hamming(2 * self.mel_overlap_len)
        self.mel_cache_len = 20
        self.source_cache_len = int(self.mel_cache_len * 256)
        self.speech_window = np.hamming(2 * self.source_cache_len)
        self.mel_overlap_dict, self.flow_cache_dict, self.hift_cache_dict = {}, {}, {}

    def open_session(self, uuid):
        self.hift_cache_dict[uuid] = None
        self.mel_overlap_dict[uuid] = torch.zeros(1, 80, 0)
        self.flow_cache_dict[uuid] = torch.zeros(1, 80, 0, 2)

    def _tail(self, mel, uuid, finalize, speed, overlap):
        cache = self.hift_cache_dict[uuid]
        src = torch.zeros(1, 1, 0)
        if cache is not None:
            mel, src = torch.concat([cache["mel"], mel], dim=2), cache["source"]
        if not finalize:
            if overlap:
                self.mel_overlap_dict[uuid] = mel[:, :, -self.mel_overlap_len:]
                mel = mel[:, :, :-self.mel_overlap_len]
            speech, source = self.hift.inference(speech_feat=mel, cache_source=src)
            if cache is not None:
                speech = fade_in_out_ref(speech, cache["speech"], self.speech_window)
            n = self.source_cache_len
            self.hift_cache_dict[uuid] = {"mel": mel[:, :, -self.mel_cache_len:], "source": source[:, :, -n:],
                                          "speech": speech[:, -n:]}
            return speech[:, :-n]
        if speed != 1.0:
            assert cache is None
            mel = interp_linear_ref(mel, int(mel.shape[2] / speed))
        speech, _ = self.hift.inference(speech_feat=mel, cache_source=src)
        if cache is not None:
            speech = fade_in_out_ref(speech, cache["speech"], self.speech_window)
        return speech

    def token2wav(self, token, prompt_token, prompt_feat, embedding, uuid, finalize=False, speed=1.0):
        i32 = lambda t: torch.tensor([t.shape[1]], dtype=torch.int32)
        mel, fc = self.flow.inference(token=token, token_len=i32(token), prompt_token=prompt_token,
                                      prompt_token_len=i32(prompt_token), prompt_feat=prompt_feat,
                                      prompt_feat_len=i32(prompt_feat), embedding=embedding,
                                      flow_cache=self.flow_cache_dict[uuid], sample_rate=self.sample_rate)
        self.flow_cache_dict[uuid] = fc
        if self.mel_overlap_dict[uuid].shape[2] != 0:
            mel = fade_in_out_ref(mel, self.mel_overlap_dict[uuid], self.mel_window)
        return self._tail(mel, uuid, finalize, speed, True)


class Token2Wav2Ref(Token2WavRef):
    """model.py:295-330 (constants), :334-366."""

    def __init__(self, flow, hift):
        self.flow, self.hift = flow, hift
        self.mel_cache_len = 8
        self.source_cache_len = int(self.mel_cache_len * 480)
        self.speech_window = np.hamming(2 * self.source_cache_len)
        self.hift_cache_dict = {}

    def open_session(self, uuid):
        self.hift_cache_dict[uuid] = None

    def token2wav(self, token, prompt_token, prompt_feat, embedding, uuid, token_offset, finalize=False, speed=1.0):
        i32 = lambda t: torch.tensor([t.shape[1]], dtype=torch.int32)
        mel, _ = self.flow.inference(token=token, token_len=i32(token), prompt_token=prompt_token,
                                     prompt_token_len=i32(prompt_token), prompt_feat=prompt_feat,
                                     prompt_feat_len=i32(prompt_feat), embedding=embedding, finalize=finalize)
        mel = mel[:, :, token_offset * self.flow.token_mel_ratio:]
        return self._tail(mel, uuid, finalize, speed, False)


# ------------------------------------------------------------------------------------------------ deterministic stubs
def _stub_mel(token: torch.Tensor, ratio: int, salt: int) -> torch.Tensor:
    """[1, n] int tokens -> [1, 80, n*ratio] fp32 with values k/32 - 1 (exact in fp32 on any device)."""
    dev = token.device
    t = torch.arange(token.shape[1] * ratio, device=dev)
    tok = token[0].long().repeat_interleave(ratio)
    c = torch.arange(80, device=dev)[:, None]
    v = (tok[None, :] * 7 + c * 3 + t[None, :] * 5 + salt) % 64
    return (v.float() / 32 - 1).unsqueeze(0)


class StubFlow:
    """Stands in for MaskedDiffWithXvec.inference (flow/flow.py:109-119): mel from tokens, a growing flow cache."""

    def __init__(self, input_frame_rate: int = 50, ratio: int = 2):
        self.input_frame_rate, self.token_mel_ratio, self.fp16 = input_frame_rate, ratio, False
        self.decoder = SimpleNamespace(estimator=SimpleNamespace(static_chunk_size=0))
        self.encoder = SimpleNamespace(static_chunk_size=0)
        self.calls = []

    def inference(self, token, token_len, prompt_token, prompt_token_len, prompt_feat, prompt_feat_len, embedding,
                  flow_cache=None, sample_rate=None, finalize=None, **kw):
        assert int(token_len[0]) == token.shape[1] and int(prompt_token_len[0]) == prompt_token.shape[1]
        self.calls.append((token.shape[1], None if flow_cache is None else tuple(flow_cache.shape), sample_rate, finalize))
        mel = _stub_mel(token, self.token_mel_ratio, int(prompt_token.shape[1]))
        if flow_cache is None:
            return mel, None
        return mel, torch.zeros(1, 80, flow_cache.shape[2] + token.shape[1], 2, device=token.device)


class StubHift:
    """Stands in for HiFTGenerator.inference (hifigan/generator.py:399-411): hop-fold upsampling, the cached source
    overwrites the head of the new one (:408-409)."""

    def __init__(self, hop: int):
        self.hop = hop

    def inference(self, speech_feat, cache_source=torch.zeros(1, 1, 0)):
        hop, dev = self.hop, speech_feat.device
        L = speech_feat.shape[2] * hop
        ramp = (torch.arange(L, device=dev) % 16).float() / 16
        s = speech_feat[:, 0:1, :].repeat_interleave(hop, dim=2) * 0.5 + ramp
        if cache_source.shape[2] != 0:
            s[:, :, :cache_source.shape[2]] = cache_source.to(dev)
        speech = s[:, 0, :] * 0.25 + speech_feat[:, 1, :].repeat_interleave(hop, dim=1) * 0.125
        return speech, s


SCENARIO_V1 = dict(chunks=[40, 45, 30], prompt_tokens=12, speed_tokens=50, speed=1.3, hop=256, sr=22050)
SCENARIO_V2 = dict(steps=[(30, 0, False), (55, 30, False), (70, 55, True)], prompt_tokens=9, speed_tokens=40,
                   speed=0.8, hop=480)


def scenario_tokens(n: int, seed: int, device="cpu") -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 4096, (1, n), generator=g, dtype=torch.int32).to(device)


def run_scenario_v1(model, device="cpu"):
    """Three streamed chunks of one session + one non-streamed call with a speed change.  `model` is anything with the
    reference's token2wav and session dictionaries (the real CosyVoiceModel, Token2WavRef or the CUDA Token2Wav)."""
    sc = SCENARIO_V1
    pt = scenario_tokens(sc["prompt_tokens"], 100, device)
    pf = torch.zeros(1, 2 * sc["prompt_tokens"], 80, device=device)
    emb = torch.zeros(1, 192, device=device)
    outs = []
    for uuid in ("s0", "s1"):
        model.hift_cache_dict[uuid] = None
        model.mel_overlap_dict[uuid] = torch.zeros(1, 80, 0, device=device)
        model.flow_cache_dict[uuid] = torch.zeros(1, 80, 0, 2, device=device)
    for i, n in enumerate(sc["chunks"]):
        outs.append(model.token2wav(token=scenario_tokens(n, i, device), prompt_token=pt, prompt_feat=pf, embedding=emb,
                                    uuid="s0", finalize=i == len(sc["chunks"]) - 1))
    outs.append(model.token2wav(token=scenario_tokens(sc["speed_tokens"], 7, device), prompt_token=pt, prompt_feat=pf,
                                embedding=emb, uuid="s1", finalize=True, speed=sc["speed"]))
    state = dict(mel_overlap=model.mel_overlap_dict["s0"], flow_cache_shape=tuple(model.flow_cache_dict["s0"].shape),
                 **{"hift_" + k: v for k, v in model.hift_cache_dict["s0"].items()})
    return outs, state


def run_scenario_v2(model, device="cpu"):
    sc = SCENARIO_V2
    pt = scenario_tokens(sc["prompt_tokens"], 100, device)
    pf = torch.zeros(1, 2 * sc["prompt_tokens"], 80, device=device)
    emb = torch.zeros(1, 192, device=device)
    all_tok = scenario_tokens(max(n for n, _, _ in sc["steps"]), 11, device)
    model.hift_cache_dict["s0"], model.hift_cache_dict["s1"] = None, None
    outs = []
    for n, off, fin in sc["steps"]:
        outs.append(model.token2wav(token=all_tok[:, :n], prompt_token=pt, prompt_feat=pf, embedding=emb, uuid="s0",
                                    token_offset=off, finalize=fin))
    outs.append(model.token2wav(token=scenario_tokens(sc["speed_tokens"], 8, device), prompt_token=pt, prompt_feat=pf,
                                embedding=emb, uuid="s1", token_offset=0, finalize=True, speed=sc["speed"]))
    state = {"hift_" + k: v for k, v in model.hift_cache_dict["s0"].items()}
    return outs, state
