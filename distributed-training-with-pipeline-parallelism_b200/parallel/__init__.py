"""Parallel runtime: schedule IR / generators / simulator, partitioner, device mesh, stage runtime,
CUDA-stream schedule runner, launcher."""
