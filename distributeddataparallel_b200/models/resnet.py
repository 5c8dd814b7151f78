"""ResNet-18/34/50 and the reference's ``SimpleCNN``.

The reference model (/root/reference/dpp.py:11-18) is torchvision's ResNet-18 with the
head replaced by ``Linear(512, 10)``, held in an attribute called ``model``.  The
checkpoint contract therefore is: ``DDP(SimpleCNN()).state_dict()`` has
122 keys, all ``module.model.<torchvision resnet18 key>``.  The module tree below
reproduces torchvision's attribute names (conv1, bn1, layer1..4, ``N.conv1``,
``N.downsample.0/1``, fc) and initialisation (Kaiming-normal fan_out for convs, BN
weight 1 / bias 0; TV/models/resnet.py:208-217) so keys, shapes, dtypes and
seeded initial values all line up with the stock model.

What differs is the *execution*: layers subclass the torch containers only for parameter
bookkeeping; ``forward`` routes to the fused sm_100a kernels in ``ops`` --
conv (implicit GEMM) -> BN(+residual)(+ReLU) as two launches per conv instead of the
3-5 ATen/cuDNN launches of the stock model -- and keeps activations NHWC
(``channels_last``) end to end.
"""
from __future__ import annotations

from typing import List, Optional, Type, Union

import torch
import torch.nn as nn

from .. import ops


class Conv2d(nn.Conv2d):
    """Bias-free square conv whose forward is the implicit-GEMM kernel (K7-K9)."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.bias is not None or self.groups != 1 or self.dilation != (1, 1):
            return super().forward(x)
        return ops.conv2d(x, self.weight, self.stride[0], self.padding[0])


class BatchNorm2d(nn.BatchNorm2d):
    """BatchNorm with optional fused residual-add and ReLU (K10-K13)."""

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None,
                relu: bool = False) -> torch.Tensor:
        training = self.training or not self.track_running_stats
        ws = None
        if ops.use_native(x):
            from ..ops import native

            ws = native.bn_workspace(self, self.num_features, x.device)
        return ops.batch_norm_act(
            x, self.weight, self.bias,
            self.running_mean if self.track_running_stats else None,
            self.running_var if self.track_running_stats else None,
            self.num_batches_tracked if (self.training and self.track_running_stats) else None,
            residual=residual, relu=relu, training=training,
            momentum=0.1 if self.momentum is None else self.momentum, eps=self.eps, ws=ws)


def conv_bn(conv: "Conv2d", bn: "BatchNorm2d", x: torch.Tensor, residual: Optional[torch.Tensor] = None,
            relu: bool = False) -> torch.Tensor:
    """conv -> BN (+residual) (+ReLU).  On the native path this is two launches: the conv's
    epilogue already accumulates the per-channel statistics BN needs."""
    if ops.use_native(x) and conv.bias is None and conv.groups == 1 and conv.dilation == (1, 1):
        from ..ops import native

        return native.conv_bn_act(x, conv.weight, conv.stride[0], conv.padding[0], bn, residual, relu)
    return bn(conv(x), residual=residual, relu=relu)


class Linear(nn.Linear):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.linear(x, self.weight, self.bias)


def conv3x3(cin: int, cout: int, stride: int = 1) -> Conv2d:
    return Conv2d(cin, cout, 3, stride=stride, padding=1, bias=False)


def conv1x1(cin: int, cout: int, stride: int = 1) -> Conv2d:
    return Conv2d(cin, cout, 1, stride=stride, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None) -> None:
        super().__init__()
        self.conv1 = conv3x3(cin, planes, stride)
        self.bn1 = BatchNorm2d(planes)
        self.relu = nn.ReLU(inplace=True)      # kept for tree/name parity; fused into bn calls
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = conv_bn(self.conv1, self.bn1, x, relu=True)
        identity = x
        if self.downsample is not None:
            identity = conv_bn(self.downsample[0], self.downsample[1], x)
        return conv_bn(self.conv2, self.bn2, out, residual=identity, relu=True)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None) -> None:
        super().__init__()
        self.conv1 = conv1x1(cin, planes)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = BatchNorm2d(planes)
        self.conv3 = conv1x1(planes, planes * self.expansion)
        self.bn3 = BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = conv_bn(self.conv1, self.bn1, x, relu=True)
        out = conv_bn(self.conv2, self.bn2, out, relu=True)
        identity = x
        if self.downsample is not None:
            identity = conv_bn(self.downsample[0], self.downsample[1], x)
        return conv_bn(self.conv3, self.bn3, out, residual=identity, relu=True)


class ResNet(nn.Module):
    def __init__(self, block: Type[Union[BasicBlock, Bottleneck]], layers: List[int],
                 num_classes: int = 1000) -> None:
        super().__init__()
        self.inplanes = 64
        self.conv1 = Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _make_layer(self, block, planes: int, blocks: int, stride: int = 1) -> nn.Sequential:
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(
                conv1x1(self.inplanes, planes * block.expansion, stride),
                BatchNorm2d(planes * block.expansion),
            )
        seq = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        seq += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*seq)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = conv_bn(self.conv1, self.bn1, x, relu=True)
        x = ops.max_pool2d(x, 3, 2, 1)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = ops.global_avg_pool(x)              # avgpool + flatten in one op
        return self.fc(x)


def _finish(model: nn.Module, channels_last: bool) -> nn.Module:
    # Filters stored KRSC (channels_last): the GEMM-K axis (r,s,cin) is contiguous, which is
    # what the implicit-GEMM kernels and the flat gradient buckets use.  Logical shapes and
    # state_dict keys are unchanged.
    return model.to(memory_format=torch.channels_last) if channels_last else model


def resnet18(num_classes: int = 1000, channels_last: bool = True) -> ResNet:
    return _finish(ResNet(BasicBlock, [2, 2, 2, 2], num_classes), channels_last)


def resnet34(num_classes: int = 1000, channels_last: bool = True) -> ResNet:
    return _finish(ResNet(BasicBlock, [3, 4, 6, 3], num_classes), channels_last)


def resnet50(num_classes: int = 1000, channels_last: bool = True) -> ResNet:
    return _finish(ResNet(Bottleneck, [3, 4, 6, 3], num_classes), channels_last)


def load_pretrained_resnet18(model: ResNet) -> None:
    """dpp.py:14 ``pretrained=True``: needs network access; best effort."""
    from torchvision.models import ResNet18_Weights

    sd = ResNet18_Weights.IMAGENET1K_V1.get_state_dict(progress=False)
    model.load_state_dict(sd)


class SimpleCNN(nn.Module):
    """The reference model (dpp.py:11-18): ResNet-18 with a 10-way head under ``.model``."""

    def __init__(self, num_classes: int = 10, pretrained: bool = False, channels_last: bool = True) -> None:
        super().__init__()
        self.model = ResNet(BasicBlock, [2, 2, 2, 2], 1000)
        if pretrained:
            load_pretrained_resnet18(self.model)
        self.model.fc = Linear(512, num_classes)     # dpp.py:15
        _finish(self, channels_last)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.model(x)
