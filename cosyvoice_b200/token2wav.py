"""Streaming glue around the hot path: CosyVoiceModel.token2wav (cosyvoice/cli/model.py:130-175) and
CosyVoice2Model.token2wav (cli/model.py:334-366), SURVEY.md section 8 row (f-2).

Same argument names, session dictionaries and return values as the reference.  What changes is where the work runs:
the reference's `fade_in_out` (utils/common.py:149-157) copies both tensors to the host, multiplies by a numpy window
and copies back -- twice per chunk; here the caches (`mel_overlap_dict`, `hift_cache_dict`) stay in HBM as contiguous
tensors and the cross-fades / speed interpolation are kernels of libcvb200 (cvb200_fade_in_out, cvb200_interp_linear).

`flow` is whatever object owns `inference(...)` in the caller (the reference's MaskedDiffWithXvec with its decoder
swapped for B200ConditionalCFM); `hift` is a B200HiFTGenerator (or anything with the reference's
`inference(speech_feat=, cache_source=)`).  There is no CPU path: tensors are moved to `device` and the kernels refuse
host pointers.
"""
import threading
from typing import Dict, Optional

import numpy as np
import torch

from . import ops


class Token2Wav:
    """CosyVoiceModel's token2wav + the per-session state it touches (cli/model.py:29-69, 130-175, 186-188, 236-238)."""

    def __init__(self, flow, hift, sr: int = 22050, device="cuda"):
        self.device = torch.device(device)
        self.flow, self.hift = flow, hift
        self.sample_rate = sr
        self.token_overlap_len = 20                                                     # model.py:48
        # mel fade in out (model.py:52-53)
        self.mel_overlap_len = int(self.token_overlap_len / self.flow.input_frame_rate * sr / 256)
        self.mel_window = self._window(self.mel_overlap_len)
        # hift cache (model.py:55-56) and speech fade (model.py:58)
        self.mel_cache_len = 20
        self.source_cache_len = int(self.mel_cache_len * 256)
        self.speech_window = self._window(self.source_cache_len)
        self.lock = threading.Lock()
        self.mel_overlap_dict: Dict[str, torch.Tensor] = {}
        self.flow_cache_dict: Dict[str, torch.Tensor] = {}
        self.hift_cache_dict: Dict[str, Optional[dict]] = {}

    def _window(self, half: int) -> torch.Tensor:
        return torch.from_numpy(np.hamming(2 * half)).to(self.device)                   # float64, like the reference

    # ---- session bookkeeping: what `tts` / `vc` do around their token2wav calls -------------------
    def open_session(self, uuid: str):
        with self.lock:                                                                 # model.py:184-188
            self.hift_cache_dict[uuid] = None
            self.mel_overlap_dict[uuid] = torch.zeros(1, 80, 0, device=self.device)
            self.flow_cache_dict[uuid] = torch.zeros(1, 80, 0, 2, device=self.device)

    def close_session(self, uuid: str):
        with self.lock:                                                                 # model.py:233-238
            self.mel_overlap_dict.pop(uuid)
            self.hift_cache_dict.pop(uuid)
            self.flow_cache_dict.pop(uuid)

    # ---- shared tail: hift cache, vocoder call, speech cross-fade ---------------------------------
    def _dev(self, t: torch.Tensor, dtype=None) -> torch.Tensor:
        return t.to(self.device) if dtype is None else t.to(self.device, dtype)

    def _fade(self, x: torch.Tensor, prev: torch.Tensor, window: torch.Tensor) -> torch.Tensor:
        # the reference returns a new tensor (it works on a host copy); keep that aliasing behaviour
        return ops.fade_in_out(self._dev(x, torch.float32).contiguous().clone(), prev, window)

    def _vocode(self, tts_mel: torch.Tensor, uuid: str, finalize: bool, speed: float, mel_overlap: bool) -> torch.Tensor:
        cache = self.hift_cache_dict[uuid]
        if cache is not None:                                                           # model.py:152-154 / 345-347
            hift_cache_source = cache["source"]
            tts_mel = torch.concat([cache["mel"], tts_mel], dim=2)
        else:
            hift_cache_source = torch.zeros(1, 1, 0, device=self.device)
        if finalize is False:
            if mel_overlap:                                                             # model.py:159-160 (v1 only)
                self.mel_overlap_dict[uuid] = tts_mel[:, :, -self.mel_overlap_len:].contiguous()
                tts_mel = tts_mel[:, :, :-self.mel_overlap_len]
            tts_speech, tts_source = self.hift.inference(speech_feat=tts_mel, cache_source=hift_cache_source)
            if cache is not None:
                tts_speech = self._fade(tts_speech, cache["speech"], self.speech_window)
            n = self.source_cache_len
            self.hift_cache_dict[uuid] = {"mel": tts_mel[:, :, -self.mel_cache_len:].contiguous(),
                                          "source": tts_source[:, :, -n:].contiguous(),
                                          "speech": tts_speech[:, -n:].contiguous()}
            tts_speech = tts_speech[:, :-n]
        else:
            if speed != 1.0:
                assert cache is None, "speed change only support non-stream inference mode"
                tts_mel = ops.interp_linear(self._dev(tts_mel, torch.float32), int(tts_mel.shape[2] / speed))
            tts_speech, tts_source = self.hift.inference(speech_feat=tts_mel, cache_source=hift_cache_source)
            if cache is not None:
                tts_speech = self._fade(tts_speech, cache["speech"], self.speech_window)
        return tts_speech

    def _len(self, t: torch.Tensor) -> torch.Tensor:
        return torch.tensor([t.shape[1]], dtype=torch.int32).to(self.device)

    @torch.inference_mode()
    def token2wav(self, token, prompt_token=torch.zeros(1, 0, dtype=torch.int32), prompt_feat=torch.zeros(1, 0, 80),
                  embedding=torch.zeros(0, 512), uuid="0", finalize=False, speed=1.0):
        tts_mel, flow_cache = self.flow.inference(token=self._dev(token), token_len=self._len(token),
                                                  prompt_token=self._dev(prompt_token),
                                                  prompt_token_len=self._len(prompt_token),
                                                  prompt_feat=self._dev(prompt_feat),
                                                  prompt_feat_len=self._len(prompt_feat),
                                                  embedding=self._dev(embedding),
                                                  flow_cache=self.flow_cache_dict[uuid],
                                                  sample_rate=self.sample_rate)
        self.flow_cache_dict[uuid] = flow_cache
        tts_mel = self._dev(tts_mel, torch.float32)
        if self.mel_overlap_dict[uuid].shape[2] != 0:                                   # model.py:149-150
            tts_mel = self._fade(tts_mel, self.mel_overlap_dict[uuid], self.mel_window)
        return self._vocode(tts_mel, uuid, finalize, speed, mel_overlap=True)


class Token2Wav2(Token2Wav):
    """CosyVoice2Model's variant (cli/model.py:295-366): no mel overlap, a token_offset into the re-generated mel,
    8-frame hift cache at hop 480."""

    def __init__(self, flow, hift, device="cuda"):
        self.device = torch.device(device)
        self.flow, self.hift = flow, hift
        self.token_hop_len = 2 * self.flow.input_frame_rate                             # model.py:313
        self.mel_cache_len = 8                                                          # model.py:318-321
        self.source_cache_len = int(self.mel_cache_len * 480)
        self.speech_window = self._window(self.source_cache_len)
        self.lock = threading.Lock()
        self.hift_cache_dict: Dict[str, Optional[dict]] = {}

    def open_session(self, uuid: str):
        with self.lock:                                                                 # model.py:375-377
            self.hift_cache_dict[uuid] = None

    def close_session(self, uuid: str):
        with self.lock:
            self.hift_cache_dict.pop(uuid)

    @torch.inference_mode()
    def token2wav(self, token, prompt_token, prompt_feat, embedding, uuid, token_offset, finalize=False, speed=1.0):
        tts_mel, _ = self.flow.inference(token=self._dev(token), token_len=self._len(token),
                                         prompt_token=self._dev(prompt_token), prompt_token_len=self._len(prompt_token),
                                         prompt_feat=self._dev(prompt_feat), prompt_feat_len=self._len(prompt_feat),
                                         embedding=self._dev(embedding), finalize=finalize)
        tts_mel = self._dev(tts_mel, torch.float32)[:, :, token_offset * self.flow.token_mel_ratio:]
        return self._vocode(tts_mel, uuid, finalize, speed, mel_overlap=False)
