import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import torch
import model_cases as mc
from fp8_quantization_b200 import model_wrap as mw, ops, eval_harness as eh
import tools.model_bench as mb

def run(force):
    orig = ops.make_params
    if force:
        ops.make_params = lambda *a, **k: orig(*a, **{**k, "force_generic": True})
    torch.manual_seed(0)
    model = mb.build("resnet18", 4, 3, torch.device("cuda", 0))
    acts = {}
    hooks = []
    for n, m in model.named_modules():
        if hasattr(m, "run_forward"):
            hooks.append(m.register_forward_hook(lambda mod, i, o, n=n: acts.setdefault(n, []).append(o.detach().clone())))
    x1 = mc.images(8, 224, 1).cuda(); x2 = mc.images(8, 224, 2).cuda()
    with torch.inference_mode():
        model.quantized(); model.estimate_ranges(); ycal = model(x1); model.fix_ranges(); yev = model(x2)
    ops.make_params = orig
    return ycal, yev, acts

a = run(False); b = run(True)
print("cal diff", float((a[0]-b[0]).abs().max()), "eval diff", float((a[1]-b[1]).abs().max()))
for n in a[2]:
    for i,(u,v) in enumerate(zip(a[2][n], b[2][n])):
        d = float((u-v).abs().max())
        if d != 0:
            print("first diff at", n, "call", i, d, tuple(u.shape)); raise SystemExit
print("no diffs")
