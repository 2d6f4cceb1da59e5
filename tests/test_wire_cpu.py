"""Host-side logic of the wire formats: checkpoint key handling and the TensorRT-style binding
surface of the estimator.  No GPU, no compute."""
import pytest
import torch

from cosyvoice_b200 import checkpoint, engine
from cosyvoice_b200.packing import fold_weight_norm
from oracle import init_weights


def _flow_pt(causal):
    est = init_weights.flow_estimator_sd(causal, seed=0)
    sd = {"decoder.estimator." + k: v for k, v in est.items()}
    sd["input_embedding.weight"] = torch.zeros(4096, 512)
    sd["encoder.embed.out.0.weight"] = torch.zeros(512, 512)
    sd["spk_embed_affine_layer.weight"] = torch.zeros(80, 192)
    return sd, est


@pytest.mark.parametrize("causal", [False, True])
def test_flow_checkpoint_is_cut_down_to_the_estimator(tmp_path, causal):
    sd, est = _flow_pt(causal)
    got, c = checkpoint.flow_estimator_state(sd)
    assert c == causal and set(got) == set(est) and all(torch.equal(got[k], est[k]) for k in est)
    path = tmp_path / "flow.pt"
    torch.save({**sd, "epoch": 3, "step": 100}, path)       # training checkpoints carry scalars next to the tensors
    got2, c2 = checkpoint.flow_estimator_state(str(path))
    assert c2 == causal and set(got2) == set(est)
    # an already-cut dict and a ConditionalCFM dict are accepted as they are
    assert set(checkpoint.flow_estimator_state(est)[0]) == set(est)
    assert set(checkpoint.flow_estimator_state({"estimator." + k: v for k, v in est.items()})[0]) == set(est)
    with pytest.raises(KeyError):
        checkpoint.flow_estimator_state({"encoder.x": torch.zeros(1)})
    bad = dict(est)
    bad["down_blocks.0.0.block1.block.0.weight"] = torch.zeros(256, 160, 3)
    with pytest.raises(ValueError):
        checkpoint.flow_estimator_state(bad)


@pytest.mark.parametrize("v2", [False, True])
def test_hift_checkpoint_prefix_and_weight_norm_styles(v2):
    h = init_weights.hift_sd(v2, seed=1)
    wrapped = {"generator." + k: v for k, v in h.items()}
    wrapped["discriminator.mpd.0.weight"] = torch.zeros(3)
    got, is_v2 = checkpoint.hift_state(wrapped)
    assert is_v2 == v2 and set(got) == set(h)
    # parametrized weight-norm keys (released checkpoints) fold to the same weights as the old-style pair
    new = {}
    for k, v in h.items():
        k = k.replace("weight_g", "parametrizations.weight.original0").replace("weight_v", "parametrizations.weight.original1")
        new[k] = v
    got_new, _ = checkpoint.hift_state(new)
    for p in ("conv_pre.", "ups.0.", "resblocks.0.convs1.0.", "f0_predictor.condnet.0."):
        assert torch.equal(fold_weight_norm(got_new, p), fold_weight_norm(h, p)), p
    with pytest.raises(KeyError):
        checkpoint.hift_state({"generator.foo": torch.zeros(1)})


def test_backend_loader_rejects_mismatched_checkpoints():
    est = init_weights.flow_estimator_sd(False, seed=0)
    h2 = init_weights.hift_sd(True, seed=1)
    with pytest.raises(ValueError):
        checkpoint.load_acoustic_backend(est, h2)


def test_engine_binding_surface_follows_the_onnx_export_names():
    eng = engine._Engine()
    assert [eng.get_tensor_name(i) for i in range(eng.num_io_tensors)] == ["x", "mask", "mu", "t", "spks", "cond",
                                                                           "estimator_out"]     # export_onnx.py:68-69
    ctx = engine.ExecutionContext(estimator=None)
    with pytest.raises(KeyError):
        ctx.set_input_shape("estimator_out", (2, 80, 10))
    with pytest.raises(ValueError):
        ctx.set_input_shape("t", (2, 1))
    with pytest.raises(RuntimeError):           # nothing bound yet
        ctx.execute_async_v3(0)
    T = 10
    for n, s in (("x", (2, 80, T)), ("mask", (2, 1, T)), ("mu", (2, 80, T)), ("t", (2,)), ("spks", (2, 80)),
                 ("cond", (2, 80, T + 1))):
        assert ctx.set_input_shape(n, s)
    for n in engine.TENSOR_NAMES:
        assert ctx.set_tensor_address(n, 4096)
    with pytest.raises(ValueError):             # cond disagrees with x
        ctx.execute_async_v3(0)
    ctx.set_input_shape("cond", (2, 80, T))
    ctx.set_tensor_address("mu", 0)
    with pytest.raises(ValueError):             # null pointer
        ctx.execute_async_v3(0)
