"""ResNet-50 (BASELINE.json config 4: "ResNet-50 synthetic 224x224 through the same strategy/fit API").

Built from the composite ``Residual`` layer so that it fits the Sequential container.  v1.5
arrangement (stride on the 3x3 convolution of the first block of stages 2-4); convolutions carry
no bias (BatchNorm follows), BatchNorm + ReLU are fused kernels, the classifier is Dense(classes).
The reference README contains no ResNet; this model exists because the driver's baseline names it.
"""
from __future__ import annotations

from ..keras import layers as L
from ..keras.losses import SparseCategoricalCrossentropy
from ..keras.models import Sequential
from ..keras.optimizers import SGD


def _bn_relu(relu=True):
    bn = L.BatchNormalization(momentum=0.9, epsilon=1e-5)
    if relu:
        bn.activation = "relu"
    return bn


def bottleneck(filters: int, stride: int, project: bool) -> L.Residual:
    main = [
        L.Conv2D(filters, 1, use_bias=False, kernel_initializer="he_normal"), _bn_relu(),
        L.Conv2D(filters, 3, strides=stride, padding="same", use_bias=False, kernel_initializer="he_normal"), _bn_relu(),
        L.Conv2D(4 * filters, 1, use_bias=False, kernel_initializer="he_normal"), _bn_relu(relu=False),
    ]
    shortcut = None
    if project:
        shortcut = [L.Conv2D(4 * filters, 1, strides=stride, use_bias=False, kernel_initializer="he_normal"),
                    _bn_relu(relu=False)]
    return L.Residual(main, shortcut)


def resnet(blocks=(3, 4, 6, 3), input_shape=(224, 224, 3), classes=1000, width=64, compile=True,
           learning_rate=0.1, name="resnet50") -> Sequential:
    layers = [
        L.Conv2D(width, 7, strides=2, padding="same", use_bias=False, kernel_initializer="he_normal",
                 input_shape=input_shape),
        _bn_relu(),
        L.MaxPooling2D(3, strides=2, padding="same"),
    ]
    for stage, n in enumerate(blocks):
        for b in range(n):
            stride = 2 if (b == 0 and stage > 0) else 1
            layers.append(bottleneck(width * 2 ** stage, stride, project=(b == 0)))
    layers += [L.GlobalAveragePooling2D(), L.Dense(classes)]
    model = Sequential(layers, name=name)
    if compile:
        model.compile(loss=SparseCategoricalCrossentropy(from_logits=True), optimizer=SGD(learning_rate=learning_rate),
                      metrics=["accuracy"])
    return model


def resnet50(**kw) -> Sequential:
    return resnet((3, 4, 6, 3), **kw)
