"""Launchers: barrier gang-spawn (README.md:170-224) and the manual one-snippet-per-worker flow."""
from .barrier import BarrierContext, BarrierError, BarrierResult, barrier_apply  # noqa: F401
from .sparklike import sdf_len, spark_connect  # noqa: F401
