"""B200-native pipeline-parallel training framework (import name: ``ppb200``).

Layers:
  ppb200.models    -- ModelArgs / Transformer (reference-compatible API + state_dict), explicit fwd/bwd stage programs
  ppb200.ops       -- ctypes bindings of libppb200.so (sm_100a kernels) + fp32 torch references for CPU
  ppb200.parallel  -- schedule IR/generators/simulator, partitioner, mesh, stage runtime, stream runner, launcher
  ppb200.utils     -- config, timing/bubble analysis, chrome trace, checkpoint, analysis/plots, clocks

The reference-compatible names (``ModelArgs``, ``Transformer``, ``manual_model_split``,
``run_train_iterations``, ``worker_process``) are re-exported lazily from ``ppb200.api``.
"""
__version__ = "0.1.0"

_API = ("ModelArgs", "Transformer", "manual_model_split", "run_train_iterations", "worker_process",
        "run_one_experiment", "run_all_experiments", "compute_speedup_and_efficiency", "RunConfig")


def __getattr__(name):
    if name in _API:
        import importlib
        api = importlib.import_module(__name__ + ".api")
        return getattr(api, name)
    raise AttributeError(name)
