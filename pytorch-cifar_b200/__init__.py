"""pytorch-cifar_b200 — B200-native (sm_100a) training hot path for the pytorch-cifar models.

The directory name follows the project convention and is not a valid Python identifier; import it as
`import b200cifar` (root-level alias module) or `importlib.import_module("pytorch-cifar_b200")`.
"""
from . import _lib  # noqa: F401
from . import ops  # noqa: F401
from . import data, engine, trainer, accelerate  # noqa: F401

__all__ = ["_lib", "ops", "data", "engine", "trainer", "accelerate"]
