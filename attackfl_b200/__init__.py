"""attackfl_b200 — a Blackwell-native federated-learning poisoning testbed.

Capabilities of filrg/attackFL (same ``server.py`` / ``client.py`` CLI, ``config.yaml`` schema and
``*.pth`` files) on a from-scratch design for 8xB200: clients are slots on one process per GPU,
updates live in a flat fp32 arena that never leaves device memory, and the hot paths are
hand-written sm_100a kernels (``csrc/``) reached through a C ABI (``include/attackfl_b200.h``).
"""
__version__ = "0.1.0"

from .config import Config, load_config, from_dict  # noqa: F401
from .layout import FlatLayout  # noqa: F401
