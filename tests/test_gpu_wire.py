"""GPU checks of the wire formats: the raw-pointer / TensorRT-style estimator binding, checkpoint files
read from disk, int16 PCM packing.  Tolerance: bit-exact (the same kernels run behind a different host surface)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from cosyvoice_b200 import checkpoint, ops
from cosyvoice_b200.engine import B200EstimatorEngine
from cosyvoice_b200.flow import B200Estimator
from oracle import init_weights

DEV = "cuda"


def _trt_branch(estimator_obj, x, mask, mu, t, spks, cond):
    """The caller's side, restated from ConditionalCFM.forward_estimator's last branch (flow_matching.py:141-165)."""
    [estimator, stream], trt_engine = estimator_obj.acquire_estimator()
    with stream:
        estimator.set_input_shape('x', (2, 80, x.size(2)))
        estimator.set_input_shape('mask', (2, 1, x.size(2)))
        estimator.set_input_shape('mu', (2, 80, x.size(2)))
        estimator.set_input_shape('t', (2,))
        estimator.set_input_shape('spks', (2, 80))
        estimator.set_input_shape('cond', (2, 80, x.size(2)))
        output = torch.zeros_like(x)
        data_ptrs = [x.contiguous().data_ptr(), mask.contiguous().data_ptr(), mu.contiguous().data_ptr(),
                     t.contiguous().data_ptr(), spks.contiguous().data_ptr(), cond.contiguous().data_ptr(),
                     output.contiguous().data_ptr()]
        for i, j in enumerate(data_ptrs):
            estimator.set_tensor_address(trt_engine.get_tensor_name(i), j)
        assert estimator.execute_async_v3(torch.cuda.current_stream().cuda_stream) is True
        torch.cuda.current_stream().synchronize()
    estimator_obj.release_estimator(estimator, stream)
    return output


@pytest.mark.parametrize("causal", [False, True])
def test_raw_pointer_engine_equals_module_call(causal):
    est = B200Estimator(init_weights.flow_estimator_sd(causal, seed=0), causal=causal, dtype=torch.bfloat16, device=DEV)
    eng = B200EstimatorEngine(est)
    assert not isinstance(eng, torch.nn.Module)
    g = torch.Generator().manual_seed(3)
    T = 150
    x, mu, cond = (torch.randn(2, 80, T, generator=g).to(DEV) for _ in range(3))
    mask = torch.ones(2, 1, T, device=DEV)
    t = torch.tensor([0.3, 0.3], device=DEV)
    spks = torch.randn(2, 80, generator=g).to(DEV)
    want = est(x, mask, mu, t, spks, cond).clone()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                    # the reference calls it from its own stream context
        got = _trt_branch(eng, x, mask, mu, t, spks, cond)
    side.synchronize()
    assert torch.equal(got, want)
    got2 = _trt_branch(eng, x, mask, mu, t, spks, cond)      # the released context is re-used
    assert torch.equal(got2, want) and len(eng._free) == 1


def test_backend_from_checkpoint_files(tmp_path):
    est = init_weights.flow_estimator_sd(False, seed=0)
    h = init_weights.hift_sd(False, seed=1)
    flow_pt = {"decoder.estimator." + k: v for k, v in est.items()}
    flow_pt["input_embedding.weight"] = torch.zeros(16, 512)
    hift_pt = {"generator." + k.replace("weight_g", "parametrizations.weight.original0")
               .replace("weight_v", "parametrizations.weight.original1"): v for k, v in h.items()}
    torch.save(flow_pt, tmp_path / "flow.pt")
    torch.save(hift_pt, tmp_path / "hift.pt")
    be = checkpoint.load_acoustic_backend(str(tmp_path / "flow.pt"), str(tmp_path / "hift.pt"), device=DEV)
    assert be.v2 is False and be.hift.hop == 256
    from cosyvoice_b200.pipeline import AcousticBackend
    ref = AcousticBackend(est, h, v2=False, device=DEV)
    g = torch.Generator().manual_seed(5)
    T, P = 90, 20
    mu, cond, z = (torch.randn(1, 80, T, generator=g) for _ in range(3))
    spk = torch.randn(1, 80, generator=g)
    mel_a = be.cfm.solve_batch([mu], [spk], [cond], [z], 2)[0]
    mel_b = ref.cfm.solve_batch([mu], [spk], [cond], [z], 2)[0]
    assert torch.equal(mel_a, mel_b)
    src = torch.randn(1, 1, (T - P) * 256, generator=g) * 0.01
    wa, _ = be.hift.inference_batch([mel_a[:, :, P:]], source=src)
    wb, _ = ref.hift.inference_batch([mel_b[:, :, P:]], source=src)
    assert torch.equal(wa[0], wb[0])


def test_pcm16_matches_numpy_cast():
    g = torch.Generator().manual_seed(6)
    wav = (torch.rand(3, 5001, generator=g) * 2 - 1) * 0.99
    wav[0, :4] = torch.tensor([0.0, -0.0, 0.5 / 32768, -0.5 / 32768])
    got = ops.pcm16(wav.to(DEV)).cpu().numpy()
    want = (wav * 32768.0).numpy().astype("int16")              # BigVGAN/utils.py:97-101
    assert got.dtype == np.int16 and np.array_equal(got, want)
    edge = torch.tensor([1.0, -1.0, 1.5, -1.5, 32767.4 / 32768])
    assert ops.pcm16(edge.to(DEV)).cpu().tolist() == [32767, -32768, 32767, -32768, 32767]     # saturate, never wrap
    assert ops.pcm16(torch.zeros(0, device=DEV)).shape == (0,)
