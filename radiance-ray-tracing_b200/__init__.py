"""radiance_b200: B200-native backend for the Radiance ray tracer's TraceRays hot path.

rd      -- host mirror of the reference's `namespace RD` API over the C ABI (libradiance_b200.so)
multigpu -- tile partition / framebuffer gather plumbing of the multi-GPU path (torch.distributed)
scenes  -- procedural scenes / descriptor bundles of the benchmark configurations (synthetic data)
"""
from . import multigpu, rd, scenes  # noqa: F401
