"""Generates tests/golden/fakequant_special.npz from the UNMODIFIED reference's quantize_to_fp8_ste_MM
(quantization/quantizers/fp8_quantizer.py:97-173) on non-finite and extreme inputs: NaN (torch.min / torch.max
propagate it), +-Inf (clamped to +-maxval), signed zeros, fp32 sub-normals, values far above maxval.

    python tests/golden/make_fakequant_special.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402


def main():
    fp8q = ref_shim.load().fp8q
    base = np.array([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e-42, -1e-42, 1e-30, 3e38, -3e38, 0.3, -0.3, 1.0, 1.4999999,
                     1.5, 1.5000001], np.float32)
    x = np.stack([base, base * np.float32(0.5), base[::-1].copy(), -base])       # [4, 16]
    out, cfgs = {"x": x}, []
    i = 0
    for mbits in (2, 3, 4, 5):
        for sbits in (1, 0):
            for mv in (np.array([1.5], np.float32), np.array([0.7, 1.5, 3.0, 100.0], np.float32)):
                y, b = fp8q.quantize_to_fp8_ste_MM(torch.from_numpy(x), 8, torch.from_numpy(mv),
                                                   torch.tensor([float(mbits)]), sbits)
                out[f"mv{i}"], out[f"y{i}"], out[f"b{i}"] = mv, y.numpy(), b.numpy()
                cfgs.append((mbits, sbits))
                i += 1
    out["cfgs"] = np.array(cfgs, np.int32)
    np.savez_compressed(os.path.join(HERE, "fakequant_special.npz"), **out)
    print("fakequant_special:", i, "cases; NaNs per case:", int(np.isnan(out["y0"]).sum()))


if __name__ == "__main__":
    main()
