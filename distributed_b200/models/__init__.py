"""Model zoo: the reference's MNIST CNN (README.md:292-298) and the scaled configurations
named in BASELINE.json."""
from .cnn import mnist_cnn, scaled_cnn  # noqa: F401
from .resnet import bottleneck, resnet, resnet50  # noqa: F401
