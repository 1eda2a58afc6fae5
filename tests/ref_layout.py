"""The reference's MODULE LAYOUT for ERFNet and Fast-SCNN, restated for the GPU box (which has no /root/reference).

These classes keep the reference's class names, attribute names and — above all — its `forward`s with their ATen
glue (`torch.cat`, `nn.MaxPool2d`, in-place `x += residual`, `nn.BatchNorm2d`, bare `nn.Conv2d`, `F.interpolate`):
reference models/enet.py:38-48, models/erfnet.py:15-82, models/fastscnn.py:16-121.  They are built on the rtseg leaf
classes exactly as the reference's files are after `rtseg_b200.inject.install` (which replaces `models.modules`).
Test infrastructure: the subject of tests/test_gpu_convert.py is what `rtseg_b200.convert` does to such a model.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from rtseg_b200.modules import (Activation, ConvBNAct, DeConvBNAct, DSConvBNAct, DWConvBNAct, PWConvBNAct,
                                PyramidPoolingModule, conv1x1)


class InitialBlock(nn.Module):
    def __init__(self, cin, cout, act_type):
        super().__init__()
        self.conv = ConvBNAct(cin, cout - cin, 3, 2, act_type=act_type)
        self.pool = nn.MaxPool2d(3, 2, 1)

    def forward(self, x):
        return torch.cat([self.conv(x), self.pool(x)], dim=1)


class NonBt1DBlock(nn.Module):
    def __init__(self, c, dilation=1, act_type="relu"):
        super().__init__()
        self.conv = nn.Sequential(
            ConvBNAct(c, c, (3, 1), inplace=True), ConvBNAct(c, c, (1, 3), inplace=True),
            ConvBNAct(c, c, (3, 1), dilation=dilation, inplace=True),
            nn.Conv2d(c, c, (1, 3), dilation=dilation, padding=(0, dilation), bias=False))
        self.bn_act = nn.Sequential(nn.BatchNorm2d(c), Activation(act_type, inplace=True))

    def forward(self, x):
        y = self.conv(x)
        y += x
        return self.bn_act(y)


def _stack(c, dilations, act_type):
    return nn.Sequential(*[NonBt1DBlock(c, d, act_type) for d in dilations])


class ERFNet(nn.Module):
    def __init__(self, num_class=19, n_channel=3, act_type="relu"):
        super().__init__()
        self.layer1 = InitialBlock(n_channel, 16, act_type)
        self.layer2 = InitialBlock(16, 64, act_type)
        self.layer3_7 = _stack(64, [1] * 5, act_type)
        self.layer8 = InitialBlock(64, 128, act_type)
        self.layer9_16 = _stack(128, [2, 4, 8, 16] * 2, act_type)
        self.layer17 = DeConvBNAct(128, 64, act_type=act_type)
        self.layer18_19 = _stack(64, [1, 1], act_type)
        self.layer20 = DeConvBNAct(64, 16, act_type=act_type)
        self.layer21_22 = _stack(16, [1, 1], act_type)
        self.layer23 = DeConvBNAct(16, num_class, act_type=act_type)

    def forward(self, x):
        for name in ("layer1", "layer2", "layer3_7", "layer8", "layer9_16", "layer17", "layer18_19", "layer20",
                     "layer21_22", "layer23"):
            x = getattr(self, name)(x)
        return x


class InvertedResidual(nn.Module):
    def __init__(self, cin, cout, stride, expand_ratio=6, act_type="relu"):
        super().__init__()
        hid = int(round(cin * expand_ratio))
        self.use_res_connect = stride == 1 and cin == cout
        self.conv = nn.Sequential(PWConvBNAct(cin, hid, act_type=act_type), DWConvBNAct(hid, hid, 3, stride, act_type=act_type),
                                  ConvBNAct(hid, cout, 1, act_type="none"))

    def forward(self, x):
        return x + self.conv(x) if self.use_res_connect else self.conv(x)


class LearningToDownsample(nn.Sequential):
    def __init__(self, cin, cout, hid=(32, 48), act_type="relu"):
        super().__init__(ConvBNAct(cin, hid[0], 3, 2, act_type=act_type), DSConvBNAct(hid[0], hid[1], 3, 2, act_type=act_type),
                         DSConvBNAct(hid[1], cout, 3, 2, act_type=act_type))


class GlobalFeatureExtractor(nn.Module):
    def __init__(self, cin, cout, act_type="relu"):
        super().__init__()
        blocks = []
        for t, c, n, s in ((6, 64, 3, 2), (6, 96, 2, 2), (6, 128, 3, 1)):
            for i in range(n):
                blocks.append(InvertedResidual(cin, c, s if i == 0 else 1, t, act_type))
                cin = c
        self.bottlenecks = nn.Sequential(*blocks)
        self.ppm = PyramidPoolingModule(cin, cout, act_type=act_type, bias=True)

    def forward(self, x):
        return self.ppm(self.bottlenecks(x))


class FeatureFusionModule(nn.Module):
    def __init__(self, c_hi, c_lo, cout, act_type="relu"):
        super().__init__()
        self.higher_res_conv = conv1x1(c_hi, cout)
        self.lower_res_conv = nn.Sequential(DWConvBNAct(c_lo, c_lo, 3, 1, act_type=act_type), conv1x1(c_lo, cout))
        self.non_linear = nn.Sequential(nn.BatchNorm2d(cout), Activation(act_type))

    def forward(self, hi, lo):
        lo = F.interpolate(lo, hi.size()[2:], mode="bilinear", align_corners=True)
        return self.non_linear(self.higher_res_conv(hi) + self.lower_res_conv(lo))


class Classifier(nn.Sequential):
    def __init__(self, c, num_class, act_type="relu"):
        super().__init__(DSConvBNAct(c, c, 3, 1, act_type=act_type), DSConvBNAct(c, c, 3, 1, act_type=act_type),
                         PWConvBNAct(c, num_class, act_type=act_type))


class FastSCNN(nn.Module):
    def __init__(self, num_class=19, n_channel=3, act_type="relu"):
        super().__init__()
        self.learning_to_downsample = LearningToDownsample(n_channel, 64, act_type=act_type)
        self.global_feature_extractor = GlobalFeatureExtractor(64, 128, act_type)
        self.feature_fusion = FeatureFusionModule(64, 128, 128, act_type)
        self.classifier = Classifier(128, num_class, act_type)

    def forward(self, x):
        size = x.size()[2:]
        hi = self.learning_to_downsample(x)
        lo = self.global_feature_extractor(hi)
        x = self.classifier(self.feature_fusion(hi, lo))
        return F.interpolate(x, size, mode="bilinear", align_corners=True)


LAYOUTS = {"erfnet": ERFNet, "fastscnn": FastSCNN}
