import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, torch
import model_cases as mc
from fp8_quantization_b200 import ops
import test_gpu_models as tm
calls = []
orig = ops.fq_matmul
def spy(x, y, **k):
    calls.append((tuple(x.shape), tuple(y.shape)))
    if os.environ.get("USE_TORCH"): return x @ y
    return orig(x, y, **k)
ops.fq_matmul = spy
name, kind, size, batch, (E, M) = [c for c in mc.MODEL_CASES if c[0]==sys.argv[1]][0]
print(name)
z = np.load(tm.GOLD)
model = tm.build(kind, size, E, M)
with torch.inference_mode():
    model.quantized(); model.estimate_ranges()
    y_cal = model(mc.images(batch, size, 1).cuda())
    model.fix_ranges()
    y_eval = model(mc.images(batch, size, 2).cuda())
print("fq_matmul calls:", calls[:5], len(calls))
for tag, got in (("cal", y_cal), ("eval", y_eval)):
    want = z[f"{name}_y{tag}"]
    print(tag, float(np.abs(got.cpu().numpy().astype(np.float64) - want).max()))
