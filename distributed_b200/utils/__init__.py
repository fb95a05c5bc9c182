"""Utilities: version/installation checks, host discovery, timers, clock sampling."""
from ..parallel.cluster import host_ip  # noqa: F401
from .env import tf_version, install_check  # noqa: F401
