"""cosyvoice_b200 -- B200-native (sm_100a) acoustic back-end for CosyVoice: flow-matching mel decoder
+ HiFT / BigVGAN vocoders behind the reference's flow / hift call signatures.  See DESIGN.md."""
__version__ = "0.1.0"
