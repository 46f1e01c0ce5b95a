"""ash_renderer_b200 — B200-native (sm_100a CUDA) backend for Ash-Renderer's per-frame compute path.

Only what the path needs: `csrc/` (CUDA kernels + the C ABI, built into libashcuda.so),
`backend` (ctypes binding, no CPU fallback), `host` (mirror of the reference's CPU-side state and
plugin API), `types` (POD layouts), `synth` (synthetic depth / colour / scene inputs),
`views` (view sharding across GPUs).
"""
__version__ = "0.1.0"
