"""Shared test helper: replays a framework layer list with PyTorch ops (the autograd oracle)."""
import torch
import torch.nn.functional as F


class _RoundBf16(torch.autograd.Function):
    """Storage rounding of the mixed_bfloat16 policy: values AND their gradients pass through bf16."""

    @staticmethod
    def forward(ctx, t):
        return t.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).float()


def torch_forward(w, x, model, training=True, running=None, emulate_bf16=False):
    """Replays the framework's layer list with torch ops; `w` maps weight name -> tensor.
    ``emulate_bf16``: round every 4-D activation (and, through autograd, its gradient) and every
    convolution kernel to bf16, which is what the mixed_bfloat16 policy stores; arithmetic stays fp32."""
    q = _RoundBf16.apply if emulate_bf16 else (lambda t: t)
    from distributed_b200.keras import layers as L

    def conv(layer, t):
        k = w[layer.name + "/kernel:0"]
        if emulate_bf16:        # bf16 shadow of the fp32 master weight, straight-through gradient
            k = k.detach().to(torch.bfloat16).float() + (k - k.detach())
        k = k.permute(3, 2, 0, 1)
        b = w.get(layer.name + "/bias:0")
        ph, pw = layer.pad_amounts()
        # TF 'same' with stride 2 pads asymmetrically (extra pixel at the end)
        H, W = t.shape[2], t.shape[3]
        oh, ow = layer.output_shape[0], layer.output_shape[1]
        th = max((oh - 1) * layer.strides[0] + layer.kernel_size[0] - H, 0)
        tw = max((ow - 1) * layer.strides[1] + layer.kernel_size[1] - W, 0)
        t = F.pad(t, (pw, tw - pw, ph, th - ph))
        t = F.conv2d(q(t), k, b, stride=layer.strides)
        return q(torch.relu(t) if layer.activation == "relu" else t)

    def run(layers, t):
        for layer in layers:
            if isinstance(layer, L.Conv2D):
                t = conv(layer, t)
            elif isinstance(layer, L.BatchNormalization):
                if training:
                    rm = rv = None
                    if running is not None:      # expected moving statistics after this step
                        rm = w[layer.name + "/moving_mean:0"].detach().clone()
                        rv = w[layer.name + "/moving_variance:0"].detach().clone()
                        running[layer.name + "/moving_mean:0"] = rm
                        running[layer.name + "/moving_variance:0"] = rv
                    t = F.batch_norm(t, rm, rv, w[layer.name + "/gamma:0"], w[layer.name + "/beta:0"],
                                     training=True, momentum=1.0 - layer.momentum, eps=layer.epsilon)
                else:
                    t = F.batch_norm(t, w[layer.name + "/moving_mean:0"], w[layer.name + "/moving_variance:0"],
                                     w[layer.name + "/gamma:0"], w[layer.name + "/beta:0"], training=False,
                                     eps=layer.epsilon)
                if layer.activation == "relu":
                    t = torch.relu(t)
                if t.dim() == 4:
                    t = q(t)
            elif isinstance(layer, L.MaxPooling2D):
                p = layer.pad_amount()
                k, s = layer.pool_size[0], layer.strides[0]
                H = t.shape[2]
                total = max((layer.output_shape[0] - 1) * s + k - H, 0)
                t = F.pad(t, (p, total - p, p, total - p), value=float("-inf"))
                t = q(F.max_pool2d(t, k, s))
            elif isinstance(layer, L.Residual):
                m = run(layer.main, t)
                sc = run(layer.shortcut, t) if layer.shortcut else t
                t = m + sc
                if layer.activation == "relu":
                    t = torch.relu(t)
                t = q(t)
            elif isinstance(layer, L.Activation):
                if layer.activation == "relu":
                    t = torch.relu(t)
            elif isinstance(layer, L.GlobalAveragePooling2D):
                t = t.mean((2, 3))
            elif isinstance(layer, L.Flatten):
                t = t.permute(0, 2, 3, 1).reshape(t.shape[0], -1)
            elif isinstance(layer, L.Dense):
                t = t @ w[layer.name + "/kernel:0"] + w[layer.name + "/bias:0"]
                if layer.activation == "relu":
                    t = torch.relu(t)
            else:
                raise AssertionError(type(layer))
        return t
    # contiguous NCHW: PyTorch 2.11's CPU convolution backward returns run-to-run different (wrong) weight
    # gradients — and occasionally crashes — for a strided 1x1 convolution fed by a permuted NHWC view
    return run(model.layers, x.contiguous())
