"""flex-gpu-scheduler_b200 — B200-native batched Filter+Score engine behind the libfgs C ABI.

The directory name follows the reference repo (``flex-gpu-scheduler``) and is not a valid Python
identifier; import it as ``flex_gpu_scheduler_b200`` (the alias package at the repo root).
"""
from . import _lib, distributed
from .api import Engine, FgsError, Packer, Results, Synth, plugin_cfg
from .configs import CONFIGS, config_plugins, config_synth

__all__ = ["Engine", "FgsError", "Packer", "Results", "Synth", "plugin_cfg", "CONFIGS", "config_plugins", "config_synth", "_lib"]
