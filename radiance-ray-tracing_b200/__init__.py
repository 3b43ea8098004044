"""radiance_b200: B200-native backend for the Radiance ray tracer's TraceRays hot path.

rd      -- host mirror of the reference's `namespace RD` API over the C ABI (libradiance_b200.so)
multigpu -- tile partition / framebuffer gather plumbing of the multi-GPU path (torch.distributed)
scene_loader -- glTF 2.0 / GLB ingestion, the mirror of RD::Scene::Load (tools/sceneBuilder.cpp)
inspector -- headless progressive viewer loop + the sample's render callback (visualization/inspector.h)
scenes  -- procedural scenes / descriptor bundles of the benchmark configurations (synthetic data)
"""
from . import multigpu, rd, scenes  # noqa: F401
