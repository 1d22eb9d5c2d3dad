import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, torch, torch.nn.functional as F
import model_cases as mc
import test_gpu_models as tm
name, kind, size, batch, (E, M) = [c for c in mc.MODEL_CASES if c[0]==sys.argv[1]][0]
model = tm.build(kind, size, E, M)
info = []
def hook(mod, inp, out, n=None):
    x = inp[0]
    xq = mod.activation_quantizer(x) if mod.quantize_input and mod._quant_a else x
    wq, b = mod.get_params()
    bR = mod.get_res_fp_bias()
    bRi = int(bR.item()) if bR is not None else 2 ** (E - 1)
    u = 2.0 ** (1 - bRi - M)
    if wq.dim() == 4:
        ab = F.conv2d(xq.abs().double(), wq.abs().double(), None, mod.stride, mod.padding, mod.dilation, mod.groups)
        pl = F.conv2d(xq.double(), wq.double(), None, mod.stride, mod.padding, mod.dilation, mod.groups)
    else:
        ab = xq.abs().double().reshape(-1, xq.shape[-1]) @ wq.abs().double().t()
        pl = xq.double().reshape(-1, xq.shape[-1]) @ wq.double().t()
    info.append((n, bRi, float(ab.max() / u), float(pl.abs().max() / u), float(xq.abs().max()), int(mod.get_acts_fp_bias().item())))
for n, m in model.named_modules():
    if hasattr(m, "run_forward"):
        m.register_forward_hook(lambda mod, i, o, n=n: hook(mod, i, o, n))
with torch.inference_mode():
    model.quantized(); model.estimate_ranges()
    y_cal = model(mc.images(batch, size, 1).cuda())
    print("---- cal pass")
    for r in info: print("%-40s bR=%3d sum|addend|/u=%.3e |sum|/u=%.3e (2^24=1.68e7) xmax=%.3g bA=%d" % r)
    info.clear()
    model.fix_ranges()
    y_eval = model(mc.images(batch, size, 2).cuda())
    print("---- eval pass")
    for r in info: print("%-40s bR=%3d sum|addend|/u=%.3e |sum|/u=%.3e xmax=%.3g bA=%d" % r)
