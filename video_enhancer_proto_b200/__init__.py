"""B200-native drop-in for the reference's per-frame 2x upscaler hot path (CNN + bicubic).

Public surface: ``B200VideoEnhancer`` (mirror of ``GPUVideoEnhancer`` / the frame-handler
protocol), ``shard`` (collective-free frame-range sharding) and ``synth`` (synthetic frames).
"""
from .upscaler import B200VideoEnhancer  # noqa: F401
from . import shard, synth, weights  # noqa: F401

GPUVideoEnhancer = B200VideoEnhancer       # reference class name (app.py:242)
from .realesrgan import B200RealESRGAN  # noqa: F401,E402

RealESRGANHandler = B200RealESRGAN         # reference class name (models/enhancement/vsr/realesrgan_handler.py:101)
__all__ = ["B200VideoEnhancer", "GPUVideoEnhancer", "B200RealESRGAN", "RealESRGANHandler", "shard", "synth", "weights"]
