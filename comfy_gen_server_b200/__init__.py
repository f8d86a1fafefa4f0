"""B200-native per-step UNet denoise path (attention / LayerNorm / GEGLU / CFG / sampler tail)."""
__version__ = "0.1.0"
