"""Import the REAL reference modules from /root/reference (build container only).

Test infrastructure.  Used by `oracle/make_golden.py` to generate the committed
fixtures and by container-only tests that cross-check the restatement.  Nothing
that runs on the GPU box may call this: /root/reference does not exist there.

Missing third-party packages are replaced by the stubs in `oracle/shims`
(SURVEY.md appendix B).
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("CVB_REFERENCE_ROOT", "/root/reference")
_SHIMS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "cosyvoice"))


def setup():
    """Put shims + reference on sys.path; neutralise the import-time nvcc JIT of
    cosyvoice/BigVGAN/alias_free_activation/cuda/activation1d.py:10 (it would
    write into the read-only reference tree)."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    for p in (REFERENCE_ROOT, _SHIMS):
        if p not in sys.path:
            sys.path.insert(0, p)
    name = "cosyvoice.BigVGAN.alias_free_activation.cuda.load"
    if name not in sys.modules:
        stub = types.ModuleType(name)
        stub.load = lambda *a, **k: None
        sys.modules[name] = stub


def build_estimator(causal: bool):
    """ConditionalDecoder with the yaml numbers
    (examples/tts_vc/cosyvoice/conf/cosyvoice.yaml:91-112, cosyvoice2/conf/cosyvoice.yaml)."""
    setup()
    from cosyvoice.flow.decoder import ConditionalDecoder
    kw = dict(in_channels=320, out_channels=80, dropout=0.0, attention_head_dim=64, n_blocks=4,
              num_mid_blocks=12, num_heads=8, act_fn="gelu")
    if causal:
        return ConditionalDecoder(causal=True, channels=[256], **kw)
    return ConditionalDecoder(causal=False, channels=[256, 256], **kw)


def build_cfm(causal: bool, estimator=None):
    setup()
    from omegaconf import DictConfig
    from cosyvoice.flow.flow_matching import ConditionalCFM, CausalConditionalCFM
    cfm_params = DictConfig(sigma_min=1e-6, solver="euler", t_scheduler="cosine",
                            training_cfg_rate=0.2, inference_cfg_rate=0.7, reg_loss_type="l1")
    est = estimator if estimator is not None else build_estimator(causal)
    cls = CausalConditionalCFM if causal else ConditionalCFM
    return cls(in_channels=240, n_spks=1, spk_emb_dim=80, cfm_params=cfm_params, estimator=est)


def build_hift(v2: bool):
    setup()
    from cosyvoice.hifigan.generator import HiFTGenerator
    from cosyvoice.hifigan.f0_predictor import ConvRNNF0Predictor
    kw = dict(in_channels=80, base_channels=512, nb_harmonics=8, sampling_rate=24000, nsf_alpha=0.1,
              nsf_sigma=0.003, nsf_voiced_threshold=10, istft_params={"n_fft": 16, "hop_len": 4},
              resblock_kernel_sizes=[3, 7, 11], resblock_dilation_sizes=[[1, 3, 5]] * 3,
              lrelu_slope=0.1, audio_limit=0.99, f0_predictor=ConvRNNF0Predictor(1, 80, 512))
    if v2:
        return HiFTGenerator(upsample_rates=[8, 5, 3], upsample_kernel_sizes=[16, 11, 7],
                             source_resblock_kernel_sizes=[7, 7, 11],
                             source_resblock_dilation_sizes=[[1, 3, 5]] * 3, **kw)
    return HiFTGenerator(upsample_rates=[8, 8], upsample_kernel_sizes=[16, 16],
                         source_resblock_kernel_sizes=[7, 11],
                         source_resblock_dilation_sizes=[[1, 3, 5]] * 2, **kw)


def build_bigvgan(topology: str = "24k100", resblock: str = "1"):
    """BigVGAN trunk; encoders left None (out of scope)."""
    setup()
    from cosyvoice.BigVGAN.bigvgan import BigVGAN
    if topology == "24k100":
        return BigVGAN(output_size=100, upsample_rates=[4, 4, 2, 2, 2, 2],
                       upsample_kernel_sizes=[8, 8, 4, 4, 4, 4], upsample_initial_channel=1536,
                       speaker_embedding_dim=512, resblock=resblock)
    if topology == "fork":
        return BigVGAN(output_size=1024, upsample_rates=[8, 5, 3, 2],
                       upsample_kernel_sizes=[16, 11, 7, 2], upsample_initial_channel=1536,
                       speaker_embedding_dim=512, resblock=resblock)
    raise ValueError(topology)


def build_flow(v2: bool, estimator=None):
    """MaskedDiffWithXvec (v1) / CausalMaskedDiffWithXvec (v2) with the yaml numbers
    (examples/tts_vc/cosyvoice/conf/cosyvoice.yaml:65-112, cosyvoice2/conf/cosyvoice.yaml:31-80)."""
    setup()
    from cosyvoice.flow.flow import MaskedDiffWithXvec, CausalMaskedDiffWithXvec
    from cosyvoice.flow.length_regulator import InterpolateRegulator
    from cosyvoice.transformer.encoder import ConformerEncoder
    from cosyvoice.transformer.upsample_encoder import UpsampleConformerEncoder
    enc_kw = dict(output_size=512, attention_heads=8, linear_units=2048, num_blocks=6, dropout_rate=0.1,
                  positional_dropout_rate=0.1, attention_dropout_rate=0.1, normalize_before=True, input_layer="linear",
                  pos_enc_layer_type="rel_pos_espnet", selfattention_layer_type="rel_selfattn", input_size=512,
                  use_cnn_module=False, macaron_style=False)
    dec = build_cfm(v2, estimator)
    if v2:
        return CausalMaskedDiffWithXvec(input_size=512, output_size=80, spk_embed_dim=192, output_type="mel", vocab_size=6561,
                                        input_frame_rate=25, only_mask_loss=True, token_mel_ratio=2, pre_lookahead_len=3,
                                        encoder=UpsampleConformerEncoder(**enc_kw), decoder=dec)
    return MaskedDiffWithXvec(input_size=512, output_size=80, spk_embed_dim=512, output_type="mel", vocab_size=4096,
                              input_frame_rate=25, only_mask_loss=True, encoder=ConformerEncoder(**enc_kw),
                              length_regulator=InterpolateRegulator(channels=80, sampling_ratios=[1, 1, 1, 1]), decoder=dec)
