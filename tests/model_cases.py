"""Deterministic FP32 model construction shared by tests/golden/make_models.py (reference side, build
container) and the GPU parity tests (our side): same seeds -> same random-init weights, BN running
statistics perturbed so that BN is not the identity."""
import torch
from torch import nn

SEED = 10   # scripts/image_net.sh:4


def perturb_bn(model, gen):
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=gen) * 0.1)
            m.running_var.copy_(torch.rand(m.running_var.shape, generator=gen) + 0.5)
            m.weight.data.copy_(torch.rand(m.weight.shape, generator=gen) * 0.5 + 0.75)
            m.bias.data.copy_(torch.randn(m.bias.shape, generator=gen) * 0.1)


def fp_resnet18():
    from torchvision.models import resnet18
    torch.manual_seed(SEED)
    model = resnet18(weights=None)
    with torch.no_grad():
        perturb_bn(model, torch.Generator().manual_seed(SEED + 1))
    return model.eval()


def fp_mobilenetv2(make, input_size):
    """`make` is the MobileNetV2 class (the reference's models/mobilenet_v2.py or our mirror of it:
    identical construction order, hence identical seeded init)."""
    torch.manual_seed(SEED)
    model = make(input_size=input_size)
    with torch.no_grad():
        perturb_bn(model, torch.Generator().manual_seed(SEED + 2))
    return model.eval()


def images(batch, size, seed):
    return torch.randn(batch, 3, size, size, generator=torch.Generator().manual_seed(seed))


def approx_params(E, M, **kw):
    d = dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=True, with_approx=True, with_s2nn2s_opt=True,
             sim_hw_add_OFUF=False, with_OF_opt=False, with_UF_opt=False, golden_clip_OF=False,
             quant_btw_mult_accu=True, debug_mode=False, self_check_mode=False)
    d.update(kw)
    return d


RUN_APPROX = dict(approx_flag=True, quantize_after_mult_and_add=False, res_quantizer_flag=True,
                  original_quantize_res=False)

# name, builder, image size, batch, (E, M)
MODEL_CASES = [
    ("resnet18_64", "resnet18", 64, 2, (4, 3)),
    ("mobilenetv2_64", "mobilenetv2", 64, 2, (3, 4)),
    # BASELINE.json configs[0]: ResNet-18, approx_v9 mant_width=3 expo_width=4, batch 8 synthetic 224x224
    ("resnet18_224_b8", "resnet18", 224, 8, (4, 3)),
    # BASELINE.json configs[2] at its real resolution: MobileNetV2 224x224 (112x112 / 56x56 depthwise tiles, the
    # split-K choices of the 7x7 tail), E4M3 -- the format tools/model_bench.py times
    ("mobilenetv2_224_b4", "mobilenetv2", 224, 4, (4, 3)),
]

# hot-path layers whose (input, output) the golden file records in both passes
TAPPED_LAYERS = {
    "resnet18_64": ["features.0", "features.2.0.features.0", "features.3.0.downsample.0", "features.4.1.features.1",
                    "features.5.1.features.1", "fc"],
    "resnet18_224_b8": ["features.5.1.features.1", "fc"],
    "mobilenetv2_224_b4": [],      # probes of all 53 hot-path layers instead (full taps would be tens of MB)
    "mobilenetv2_64": ["features.0.0", "features.1.conv.0", "features.2.conv.0", "features.2.conv.1",
                       "features.4.conv.2", "features.17.conv.1", "features.18.0", "classifier.1"],
}

# cases for which the golden file also holds, for EVERY hot-path layer, a fixed sample of the layer output
# in the end-to-end pass (keys f"{case}_{layer}_p{cal|eval}")
PROBED_CASES = ["resnet18_224_b8", "mobilenetv2_224_b4"]
PROBE_N = 4096


def probe_indices(numel, layer_index):
    g = torch.Generator().manual_seed(1000 + layer_index)
    return torch.randint(0, numel, (min(PROBE_N, numel),), generator=g)
