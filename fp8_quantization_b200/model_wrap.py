"""Host-side graph surgery that puts the hot-path operators into a model: the callers on either
side of the path.  Runs once per model, no arithmetic.

Mirrors, with the same names and keyword contract,
  approx/replace_operations_with_approx_ops.py:119-157   QuantizedActivationWrapper
                                              :175-177   Flattener
                                              :180-191   module maps
                                              :263-286   fold_bn
                                              :289-342   quantize_sequential
                                              :345-384   quantize_model
  quantization/base_quantized_model.py:20-140            QuantizedModel
  models/resnet_quantized_approx.py:11-130               QuantizedBlock, QuantizedResNet
  models/mobilenet_v2_quantized_approx.py:12-88          QuantizedInvertedResidual, QuantizedMobileNetV2
  models/mobilenet_v2.py:17-132                          the FP32 MobileNetV2 graph the wrapper expects
Conv2d followed by BatchNorm becomes QCustomBNConv2dTorch (approx path), Conv2d without BN becomes
QCustomConv2dTorch (exact matmul), Linear becomes QCustomLinearTorch.
"""
import copy
import math
import warnings

import torch
from torch import nn
from torch.nn.modules.pooling import _AdaptiveAvgPoolNd, _AvgPoolNd

import torch.nn.functional as F

from .approx_calculation import QCustomBNConv2dTorch, QCustomConv2dTorch, QCustomLinearTorch
from .quantization import (ACTIVATIONS, QuantizationHijacker, QuantizationManager, QuantizedActivation,
                           QuantizedModule)


class QuantLayerNorm(QuantizationHijacker, nn.LayerNorm):
    """approx/replace_operations_with_approx_ops.py:160-168: LayerNorm whose weight is fake-quantized."""

    def run_forward(self, x, weight, bias, offsets=None):
        return F.layer_norm(x.contiguous(), self.normalized_shape, weight.contiguous(), bias.contiguous(), self.eps)


non_bn_module_map = {nn.Conv2d: QCustomConv2dTorch, nn.Linear: QCustomLinearTorch, nn.LayerNorm: QuantLayerNorm}
bn_module_map = {nn.Conv2d: QCustomBNConv2dTorch}
non_param_modules = (_AdaptiveAvgPoolNd, _AvgPoolNd)


class FP32Acts(nn.Module):
    def forward(self, x):
        return x

    def reset_ranges(self):
        pass


class Flattener(nn.Module):
    def forward(self, x):
        return x.view(x.shape[0], -1)


class QuantizedActivationWrapper(QuantizedActivation):
    """Runs `layer`, then quantizes its output; with tied quantizers the producer's activation
    quantizer is reused WITHOUT a range update (average pooling cannot widen the range)."""

    def __init__(self, layer, tie_activation_quantizers=False, input_quantizer=None, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.tie_activation_quantizers = tie_activation_quantizers
        if input_quantizer:
            assert isinstance(input_quantizer, QuantizationManager)
            self.activation_quantizer = input_quantizer
        self.layer = layer

    def forward(self, x):
        x = self.layer(x)
        if not self.tie_activation_quantizers:
            return self.quantize_activations(x)
        return self.activation_quantizer.quantizer(x) if self._quant_a else x


def _has_bn(seq, i):
    return i + 1 < len(seq) and isinstance(seq[i + 1], (nn.BatchNorm2d, nn.BatchNorm1d))


def _following_act(seq, i):
    """conv+act or conv+bn+act -> (act, index); anything else -> (None, None)."""
    j = i + 2 if _has_bn(seq, i) else i + 1
    if j < len(seq) and isinstance(seq[j], ACTIVATIONS):
        return seq[j], j
    return None, None


def _ctor_args(mod, act):
    if isinstance(mod, nn.Conv2d):
        kw = dict(in_channels=mod.in_channels, out_channels=mod.out_channels, kernel_size=mod.kernel_size,
                  stride=mod.stride, padding=mod.padding, dilation=mod.dilation, groups=mod.groups,
                  bias=mod.bias is not None)
    elif isinstance(mod, nn.Linear):
        kw = dict(in_features=mod.in_features, out_features=mod.out_features, bias=mod.bias is not None)
    elif isinstance(mod, nn.LayerNorm):
        kw = dict(normalized_shape=mod.normalized_shape, eps=mod.eps, elementwise_affine=mod.elementwise_affine)
    else:
        raise ValueError(type(mod))
    kw["activation"] = act
    return kw


def fold_bn(module, i, **quant_params):
    bn = _has_bn(module, i)
    act, _ = _following_act(module, i)
    src = module[i]
    if bn and type(src) not in bn_module_map:
        raise NotImplementedError(f"{type(src).__name__} followed by BatchNorm is outside the hot path")
    cls = (bn_module_map if bn else non_bn_module_map)[type(src)]
    new = cls(**_ctor_args(src, act), **quant_params)
    new.weight.data = src.weight.data.clone()
    if bn:
        norm = module[i + 1]
        new.gamma.data = norm.weight.data.clone()
        new.beta.data = norm.bias.data.clone()
        new.running_mean.data = norm.running_mean.data.clone()
        new.running_var.data = norm.running_var.data.clone()
        if src.bias is not None:
            new.running_mean.data -= src.bias.data
        new.epsilon = norm.eps
    elif src.bias is not None:
        new.bias.data = src.bias.data.clone()
    return new, i + 1 + int(bn) + int(act is not None)


def quantize_sequential(model, specials=None, tie_activation_quantizers=False, **quant_params):
    specials = specials or {}
    out = []
    i = 0
    while i < len(model):
        m = model[i]
        if isinstance(m, QuantizedModule):
            out.append(m)
        elif type(m) in non_bn_module_map:
            new, i = fold_bn(model, i, **quant_params)
            out.append(new)
            continue
        elif type(m) in specials:
            out.append(specials[type(m)](m, **quant_params))
        elif isinstance(m, non_param_modules):
            producer = None
            if out and isinstance(out[-1], QuantizedModule):
                producer = out[-1]
            elif out and isinstance(out[-1], nn.Sequential) and isinstance(out[-1][-1], QuantizedModule):
                producer = out[-1][-1]
            if producer is not None and tie_activation_quantizers:
                out.append(QuantizedActivationWrapper(m, tie_activation_quantizers=True,
                                                      input_quantizer=producer.activation_quantizer, **quant_params))
            else:
                out.append(QuantizedActivationWrapper(m, **quant_params))
                if tie_activation_quantizers:
                    warnings.warn("Input quantizer not found, so we do not tie quantizers")
        else:
            out.append(quantize_model(m, specials=specials, **quant_params))
        i += 1
    return nn.Sequential(*out)


def quantize_model(model, specials=None, tie_activation_quantizers=False, **quant_params):
    specials = specials or {}
    if isinstance(model, nn.Sequential):
        return quantize_sequential(model, specials, tie_activation_quantizers, **quant_params)
    if type(model) in specials:
        return specials[type(model)](model, **quant_params)
    if isinstance(model, non_param_modules):
        return QuantizedActivationWrapper(model, **quant_params)
    if type(model) in non_bn_module_map:
        new = non_bn_module_map[type(model)](**_ctor_args(model, None), **quant_params)
        new.weight.data = model.weight.data
        if getattr(model, "bias", None) is not None:
            new.bias.data = model.bias.data
        return new
    clone = copy.deepcopy(model)
    children = enumerate(clone) if isinstance(clone, nn.ModuleList) else list(clone._modules.items())
    for key, child in children:
        new = quantize_model(child, specials=specials, **quant_params)
        if new is not None:
            if isinstance(clone, nn.ModuleList):
                clone[key] = new
            else:
                setattr(clone, key, new)
    return clone


class QuantizedModel(nn.Module):
    """Whole-model switches (base_quantized_model.py:65-140)."""

    def __init__(self, input_size=(1, 3, 224, 224)):
        super().__init__()
        self.input_size = input_size

    def _each(self, fn):
        for m in self.modules():
            if isinstance(m, QuantizedModule):
                fn(m)

    def quantized(self):
        self._each(lambda m: m.quantized())

    def full_precision(self):
        self._each(lambda m: m.full_precision())

    def quantized_weights(self):
        self._each(lambda m: m.quantized_weights())

    def quantized_acts(self):
        self._each(lambda m: m.quantized_acts())

    def estimate_ranges(self):
        self._each(lambda m: m.estimate_ranges())

    def fix_ranges(self):
        self._each(lambda m: m.fix_ranges())

    def set_quant_state(self, weight_quant, act_quant):
        self._each(lambda m: (m.quantized_weights() if weight_quant else m.full_precision_weights(),
                              m.quantized_acts() if act_quant else m.full_precision_acts()))


# ---- ResNet (models/resnet_quantized_approx.py) --------------------------------------------------
class QuantizedBlock(QuantizedActivation):
    def __init__(self, block, **quant_params):
        super().__init__(**quant_params)
        if hasattr(block, "conv3"):      # Bottleneck
            body = [block.conv1, block.bn1, block.relu, block.conv2, block.bn2, block.relu, block.conv3, block.bn3]
        else:                            # BasicBlock
            body = [block.conv1, block.bn1, block.relu, block.conv2, block.bn2]
        self.features = quantize_model(nn.Sequential(*body), **quant_params)
        self.downsample = quantize_model(block.downsample, **quant_params) if block.downsample else None
        self.relu = block.relu

    def forward(self, x):
        residual = x if self.downsample is None else self.downsample(x)
        out = self.features(x)
        out += residual
        out = self.relu(out)
        return self.quantize_activations(out)


class QuantizedResNet(QuantizedModel):
    def __init__(self, resnet, input_size=(1, 3, 224, 224), quant_setup=None, **quant_params):
        super().__init__(input_size)
        from torchvision.models.resnet import BasicBlock, Bottleneck
        specials = {BasicBlock: QuantizedBlock, Bottleneck: QuantizedBlock}
        stem = [resnet.conv1, resnet.bn1, resnet.relu] + ([resnet.maxpool] if hasattr(resnet, "maxpool") else [])
        features = nn.Sequential(*stem, resnet.layer1, resnet.layer2, resnet.layer3, resnet.layer4)
        self.features = quantize_model(features, specials=specials, **quant_params)
        self.avgpool = QuantizedActivationWrapper(resnet.avgpool, tie_activation_quantizers=True,
                                                  input_quantizer=self.features[-1][-1].activation_quantizer,
                                                  **quant_params)
        self.flattener = Flattener()
        self.fc = quantize_model(resnet.fc, **quant_params)
        if quant_setup == "FP_logits":
            self.fc.activation_quantizer = FP32Acts()
        elif quant_setup not in (None, "all"):
            raise ValueError(f"Quantization setup '{quant_setup}' not supported for Resnet")

    def forward(self, x):
        return self.fc(self.flattener(self.avgpool(self.features(x))))


def resnet18_quantized_approx(fp_model=None, **qparams):
    """models/resnet_quantized_approx.py:133-147 with random-init weights (no pretrained blobs here)."""
    if fp_model is None:
        from torchvision.models import resnet18
        fp_model = resnet18(weights=None)
    return QuantizedResNet(fp_model, **qparams)


# ---- MobileNetV2 (models/mobilenet_v2.py, models/mobilenet_v2_quantized_approx.py) -----------------
def _cba(cin, cout, k, stride, groups=1, act=True):
    layers = [nn.Conv2d(cin, cout, k, stride, k // 2, groups=groups, bias=False), nn.BatchNorm2d(cout)]
    return layers + ([nn.ReLU6(inplace=True)] if act else [])


class InvertedResidual(nn.Module):
    def __init__(self, inp, oup, stride, expand_ratio):
        super().__init__()
        hidden = round(inp * expand_ratio)
        self.use_res_connect = stride == 1 and inp == oup
        layers = [] if expand_ratio == 1 else _cba(inp, hidden, 1, 1)
        layers += _cba(hidden, hidden, 3, stride, groups=hidden) + _cba(hidden, oup, 1, 1, act=False)
        self.conv = nn.Sequential(*layers)

    def forward(self, x):
        return x + self.conv(x) if self.use_res_connect else self.conv(x)


class MobileNetV2(nn.Module):
    SETTING = [(1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]

    def __init__(self, n_class=1000, input_size=224, width_mult=1.0, dropout=0.0):
        super().__init__()
        assert input_size % 32 == 0
        cin = int(32 * width_mult)
        self.last_channel = int(1280 * width_mult) if width_mult > 1.0 else 1280
        features = [nn.Sequential(*_cba(3, cin, 3, 2))]
        for t, c, n, s in self.SETTING:
            cout = int(c * width_mult)
            for i in range(n):
                features.append(InvertedResidual(cin, cout, s if i == 0 else 1, t))
                cin = cout
        features.append(nn.Sequential(*_cba(cin, self.last_channel, 1, 1)))
        features.append(nn.AvgPool2d(input_size // 32))
        self.features = nn.Sequential(*features)
        self.classifier = nn.Sequential(nn.Dropout(dropout), nn.Linear(self.last_channel, n_class))
        for m in self.modules():           # models/mobilenet_v2.py:117-132
            if isinstance(m, nn.Conv2d):
                m.weight.data.normal_(0, math.sqrt(2.0 / (m.kernel_size[0] * m.kernel_size[1] * m.out_channels)))
            elif isinstance(m, nn.BatchNorm2d):
                m.weight.data.fill_(1)
                m.bias.data.zero_()
            elif isinstance(m, nn.Linear):
                m.weight.data.normal_(0, 0.01)
                m.bias.data.zero_()

    def forward(self, x):
        x = self.features(x)
        return self.classifier(torch.flatten(x, 1))


class QuantizedInvertedResidual(QuantizedActivation):
    def __init__(self, inv_res_orig, **quant_params):
        super().__init__(**quant_params)
        self.use_res_connect = inv_res_orig.use_res_connect
        self.conv = quantize_sequential(inv_res_orig.conv, **quant_params)

    def forward(self, x):
        if self.use_res_connect:
            return self.quantize_activations(x + self.conv(x))
        return self.conv(x)


class QuantizedMobileNetV2(QuantizedModel):
    def __init__(self, model_fp, input_size=(1, 3, 224, 224), quant_setup=None, **quant_params):
        super().__init__(input_size)
        specials = {type(model_fp.features[1]): QuantizedInvertedResidual}
        self.features = quantize_sequential(model_fp.features, tie_activation_quantizers=True, specials=specials,
                                            **quant_params)
        self.flattener = Flattener()
        self.classifier = quantize_model(model_fp.classifier, **quant_params)
        if quant_setup == "FP_logits":
            self.classifier[1].activation_quantizer = FP32Acts()
        elif quant_setup not in (None, "all"):
            raise ValueError(f"Quantization setup '{quant_setup}' not supported for MobilenetV2")

    def forward(self, x):
        return self.classifier(self.flattener(self.features(x)))


def mobilenetv2_quantized_approx(fp_model=None, **qparams):
    return QuantizedMobileNetV2(fp_model if fp_model is not None else MobileNetV2(), **qparams)
