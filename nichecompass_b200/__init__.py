"""nichecompass_b200 -- B200-native VGPGAE training / latent-inference path of NicheCompass."""
__version__ = "0.1.0"
