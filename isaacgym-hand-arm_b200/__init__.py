"""handarm_b200: B200-native task-side hot path of the Isaac Gym hand / hand-arm environments.

Sub-modules: ``config`` (Hydra-YAML loader), ``sim`` (synthetic PhysX state producer),
``vec_task`` (VecTask surface), ``tasks`` (task classes), ``_lib`` (C-ABI binding).
"""
__version__ = "0.1.0"
