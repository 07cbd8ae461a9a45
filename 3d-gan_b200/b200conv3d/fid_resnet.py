"""3-D ResNet-50 feature extractor of the reference's 3D-FID (FID_ResNet.py:65-102,105-189,
used by eval_utils.fid_3d, eval_utils.py:41-52) built from the B200 drop-in layers, with the
epilogues the conv path offers folded in: BN + ReLU in one pass and BN + residual add + ReLU
in one pass (FID_ResNet.py:97-100).  Parameter names equal the reference's, so the MedicalNet
checkpoint (`resnet_50.pth`, eval_utils.py:54-63) or a reference state_dict loads unchanged.
MaxPool3d(3,2,1) after the stem is one NDHWC pass (b2c_maxpool3d).
"""
import torch
import torch.nn as tnn

from . import nn as BN


class Bottleneck(tnn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, dilation=1, downsample=None):
        super().__init__()
        self.conv1 = BN.Conv3d(inplanes, planes, kernel_size=1, bias=False)
        self.bn1 = BN.BatchNorm3d(planes)
        self.conv2 = BN.Conv3d(planes, planes, kernel_size=3, stride=stride, dilation=dilation, padding=dilation,
                               bias=False)
        self.bn2 = BN.BatchNorm3d(planes)
        self.conv3 = BN.Conv3d(planes, planes * 4, kernel_size=1, bias=False)
        self.bn3 = BN.BatchNorm3d(planes * 4)
        self.downsample = downsample
        for bn in (self.bn1, self.bn2, self.bn3):
            bn.fused_act, bn.fused_slope = "relu", 0.0
        for conv in (self.conv1, self.conv2, self.conv3):      # each feeds a BatchNorm3d: statistics in the GEMM epilogue
            conv.fused_stats = "channel"

    def forward(self, x):
        out = self.bn1(self.conv1(x))                      # BN + ReLU, one pass
        out = self.bn2(self.conv2(out))
        residual = x if self.downsample is None else self.downsample(x)
        return self.bn3(self.conv3(out), residual=residual)   # BN + add + ReLU, one pass


class ResNet(tnn.Module):
    def __init__(self, layers=(3, 4, 6, 3)):
        super().__init__()
        self.inplanes = 64
        self.conv1 = BN.Conv3d(1, 64, kernel_size=7, stride=(2, 2, 2), padding=(3, 3, 3), bias=False)
        self.bn1 = BN.BatchNorm3d(64)
        self.bn1.fused_act = "relu"
        self.conv1.fused_stats = "channel"
        self.maxpool = BN.MaxPool3d(kernel_size=(3, 3, 3), stride=2, padding=1)
        self.layer1 = self._make_layer(64, layers[0])
        self.layer2 = self._make_layer(128, layers[1], stride=2)
        self.layer3 = self._make_layer(256, layers[2], stride=1, dilation=2)
        self.layer4 = self._make_layer(512, layers[3], stride=1, dilation=4)
        for m in self.modules():                           # FID_ResNet.py:139-144
            if isinstance(m, tnn.Conv3d):
                m.weight = tnn.init.kaiming_normal_(m.weight, mode="fan_out")
            elif isinstance(m, tnn.BatchNorm3d):
                m.weight.data.fill_(1)
                m.bias.data.zero_()

    def _make_layer(self, planes, blocks, stride=1, dilation=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * 4:
            downsample = tnn.Sequential(
                BN.Conv3d(self.inplanes, planes * 4, kernel_size=1, stride=stride, bias=False),
                BN.BatchNorm3d(planes * 4))
            downsample[0].fused_stats = "channel"
        layers = [Bottleneck(self.inplanes, planes, stride=stride, dilation=dilation, downsample=downsample)]
        self.inplanes = planes * 4
        for _ in range(1, blocks):
            layers.append(Bottleneck(self.inplanes, planes, dilation=dilation))
        return tnn.Sequential(*layers)

    def forward(self, x):
        x = self.maxpool(self.bn1(self.conv1(x)))
        return self.layer4(self.layer3(self.layer2(self.layer1(x))))


def resnet50():
    return ResNet((3, 4, 6, 3))


@torch.no_grad()
def features(model, x):
    """eval_utils.py:41-44: spatial mean of the final activations -> [B, 2048] (BN in train mode, F10)."""
    return model(x).float().mean(dim=(2, 3, 4))
