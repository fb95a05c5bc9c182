"""Parallelism: cluster spec (TF_CONFIG), strategy, communicators, symmetric memory."""
from .cluster import ClusterSpec, ClusterSpecError, host_ip, local_cluster  # noqa: F401
