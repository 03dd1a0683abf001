"""ORACLE (test infrastructure only -- never imported by the product path).

CPU fp32 restatement of the timm pieces the reference's STCConnector is built from.  timm is a third-party
dependency that is NOT vendored under /root/reference and is not installed in this image:
    reference call sites : VideoLLaMA2/videollama2/model/projector.py:22-23 (imports), :153-161 (s1), :176-184 (s2)
    pinned version       : timm==1.0.17 (requirements_rorqual.txt:28; upstream VideoLLaMA2 pins 1.0.3)
    restated algorithm   : timm/models/regnet.py (RegStage, Bottleneck, create_shortcut/downsample_conv),
                           timm/layers/{conv_bn_act.py (ConvNormAct), squeeze_excite.py (SEModule),
                           norm.py (LayerNorm2d), norm_act.py (LayerNormAct2d)}.
Parameter names mirror timm's (conv1.conv / conv1.bn / se.fc1 ...) so reference checkpoints would load.
PARITY UNPINNED for these constants (timm absent here): LayerNormAct2d eps (1e-5 by timm's ConvNormAct mapping of
LayerNorm2d -> LayerNormAct2d), se_ratio=0.25 with rd_channels = round(in_chs*se_ratio), bottle_ratio=1,
group_size=1 (depthwise 3x3), conv bias=False.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

STC_LN_EPS = 1e-5


class LayerNorm2d(nn.LayerNorm):
    """LayerNorm over the channel axis of an NCHW tensor (permute -> F.layer_norm -> permute)."""

    def __init__(self, num_channels, eps=STC_LN_EPS, affine=True):
        super().__init__(num_channels, eps=eps, elementwise_affine=affine)

    def forward(self, x):
        x = x.permute(0, 2, 3, 1)
        x = F.layer_norm(x, self.normalized_shape, self.weight, self.bias, self.eps)
        return x.permute(0, 3, 1, 2)


class ConvNormAct(nn.Module):
    def __init__(self, in_chs, out_chs, kernel_size=1, stride=1, dilation=1, groups=1, apply_act=True,
                 act_layer=nn.SiLU, norm_layer=LayerNorm2d):
        super().__init__()
        pad = ((kernel_size - 1) // 2) * dilation
        self.conv = nn.Conv2d(in_chs, out_chs, kernel_size, stride=stride, padding=pad, dilation=dilation,
                              groups=groups, bias=False)
        self.bn = norm_layer(out_chs)
        self.act = act_layer() if apply_act else nn.Identity()

    def forward(self, x):
        return self.act(self.bn(self.conv(x)))


class SEModule(nn.Module):
    def __init__(self, channels, rd_channels, act_layer=nn.SiLU):
        super().__init__()
        self.fc1 = nn.Conv2d(channels, rd_channels, kernel_size=1, bias=True)
        self.act = act_layer()
        self.fc2 = nn.Conv2d(rd_channels, channels, kernel_size=1, bias=True)

    def forward(self, x):
        x_se = x.mean((2, 3), keepdim=True)
        x_se = self.fc2(self.act(self.fc1(x_se)))
        return x * x_se.sigmoid()


class Bottleneck(nn.Module):
    def __init__(self, in_chs, out_chs, stride=1, dilation=(1, 1), bottle_ratio=1, group_size=1, se_ratio=0.25,
                 act_layer=nn.SiLU, norm_layer=LayerNorm2d):
        super().__init__()
        bottleneck_chs = int(round(out_chs * bottle_ratio))
        groups = bottleneck_chs // group_size
        cargs = dict(act_layer=act_layer, norm_layer=norm_layer)
        self.conv1 = ConvNormAct(in_chs, bottleneck_chs, kernel_size=1, **cargs)
        self.conv2 = ConvNormAct(bottleneck_chs, bottleneck_chs, kernel_size=3, stride=stride, dilation=dilation[0],
                                 groups=groups, **cargs)
        se_channels = int(round(in_chs * se_ratio))
        self.se = SEModule(bottleneck_chs, rd_channels=se_channels, act_layer=act_layer)
        self.conv3 = ConvNormAct(bottleneck_chs, out_chs, kernel_size=1, apply_act=False, **cargs)
        self.act3 = act_layer()
        if in_chs != out_chs or stride != 1 or dilation[0] != dilation[1]:
            self.downsample = ConvNormAct(in_chs, out_chs, kernel_size=1, stride=stride, apply_act=False, **cargs)
        else:
            self.downsample = nn.Identity()

    def forward(self, x):
        shortcut = x
        x = self.conv1(x)
        x = self.conv2(x)
        x = self.se(x)
        x = self.conv3(x)
        x = x + self.downsample(shortcut)
        return self.act3(x)


class RegStage(nn.Module):
    def __init__(self, depth, in_chs, out_chs, stride, dilation, act_layer=nn.SiLU, norm_layer=LayerNorm2d, **kw):
        super().__init__()
        first_dilation = 1 if dilation in (1, 2) else 2
        for i in range(depth):
            block_stride = stride if i == 0 else 1
            block_in = in_chs if i == 0 else out_chs
            self.add_module("b%d" % (i + 1), Bottleneck(block_in, out_chs, stride=block_stride,
                                                        dilation=(first_dilation, dilation), act_layer=act_layer,
                                                        norm_layer=norm_layer, **kw))
            first_dilation = dilation

    def forward(self, x):
        for block in self.children():
            x = block(x)
        return x
