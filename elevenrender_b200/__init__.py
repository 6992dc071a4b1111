"""elevenrender_b200 — B200-native (sm_100a) path-tracing hot path of ElevenRender behind a C ABI.

capi    ctypes binding of include/er_cuda.h (elevenrender_b200/liber_cuda.so); no CPU fallback
host    Scene / RenderParameters / RenderingManager mirror of the reference's host interface
scenes  procedural scenes of the BASELINE.json configs
"""
from . import capi, host, scenes, workloads  # noqa: F401
from .host import (HDRI, RenderingManager, RenderParameters, Scene, Texture, camera_words,  # noqa: F401
                   material_words)
