import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, torch
import model_cases as mc
import test_gpu_models as tm
name, kind, size, batch, (E, M) = [c for c in mc.MODEL_CASES if c[0]=="resnet18_224_b8"][0]
z = np.load(tm.GOLD)
model = tm.build(kind, size, E, M)
phase = ["cal"]
res = []
mods = dict(model.named_modules())
for li, (lname, mod) in enumerate((n, m) for n, m in mods.items() if hasattr(m, "run_forward")):
    def probe(mod, inp, outp, lname=lname, li=li):
        flat = outp.detach().contiguous().reshape(-1)
        idx = mc.probe_indices(flat.numel(), li).to(flat.device)
        got = flat[idx].cpu().numpy()
        key = f"{name}_{lname}_p{phase[0]}"
        if key in z.files:
            want = z[key]
            nbad = int((got != want).sum())
            res.append((phase[0], lname, nbad, len(want), float(np.abs(got - want).max()), float(np.abs(want).max())))
        else:
            res.append((phase[0], lname, -1, 0, 0, 0))
    mod.register_forward_hook(probe)
with torch.inference_mode():
    model.quantized(); model.estimate_ranges()
    y_cal = model(mc.images(batch, size, 1).cuda())
    model.fix_ranges(); phase[0] = "eval"
    y_eval = model(mc.images(batch, size, 2).cuda())
for r in res: print("%-5s %-28s mismatches %5d / %5d  max|d| %.3e  (scale %.3e)" % r)
