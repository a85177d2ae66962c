"""Config, analysis/plots, measured-timeline tracing, checkpointing, GPU clock sampling."""
