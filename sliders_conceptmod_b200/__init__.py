"""sliders_conceptmod_b200 — B200-native concept-slider hot path.

Host-side mirror of the reference's Python API (LoRANetwork / PromptEmbedsPair /
train_util) over hand-written sm_100a CUDA kernels behind a C ABI
(include/slider_b200.h, libslider_b200.so).
"""
__version__ = "0.1.0"
