"""Datasets (README.md:286-287: ``tf.keras.datasets.mnist.load_data()``)."""
from . import mnist  # noqa: F401
