"""Single-GPU operators: the fused stencil engine and the CPU oracles."""
from .oracle import blur_cpu, numpy_oracle, reference_oracle, sharded_oracle, threaded_halo_model, threaded_persist_model, threaded_resident_model
from .stencil import BlurPipeline, Engine, blur, blur_file, clear_engine_cache, load_torch_ops, stencil_step

__all__ = ["blur_cpu", "Engine", "BlurPipeline", "blur", "blur_file", "stencil_step", "load_torch_ops", "clear_engine_cache", "reference_oracle",
           "numpy_oracle", "sharded_oracle", "threaded_halo_model", "threaded_persist_model", "threaded_resident_model"]
