"""Generates tests/golden/bn.npz: eval-mode F.batch_norm outputs computed by torch on CPU tensors -- the op
the reference's BNFusedHijacker.forward calls (quantization/quantized_folded_bn.py:66-75) when the model runs
on CPU (BASELINE.json configs[0]).  Pins the op sequence of the fused BN epilogue kernel (csrc/elementwise.cu).

Build container only:   python tests/golden/make_bn.py
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    g = torch.Generator().manual_seed(77)
    out = {}
    cfgs = []
    for i, (B, C, H, W, eps) in enumerate([(2, 20, 5, 7, 1e-5), (3, 64, 14, 14, 1e-5), (1, 33, 1, 1, 1e-3), (2, 96, 9, 3, 1e-5)]):
        x = torch.randn(B, C, H, W, generator=g) * torch.exp(torch.randn(1, C, 1, 1, generator=g))
        mean = torch.randn(C, generator=g) * 0.3
        var = torch.rand(C, generator=g) * 2 + 0.05
        gamma = torch.rand(C, generator=g) + 0.5
        beta = torch.randn(C, generator=g) * 0.2
        y = F.batch_norm(x, mean, var, gamma, beta, False, 0.1, eps)
        y_cl = F.batch_norm(x.contiguous(memory_format=torch.channels_last), mean, var, gamma, beta, False, 0.1, eps)
        assert torch.equal(y, y_cl)            # ATen's CPU result does not depend on the memory format
        for k, v in dict(x=x, mean=mean, var=var, gamma=gamma, beta=beta, y=y).items():
            out[f"{k}{i}"] = v.numpy()
        cfgs.append((B, C, H, W))
        out[f"eps{i}"] = np.float32(eps)
    out["cfgs"] = np.array(cfgs, np.int32)
    np.savez_compressed(os.path.join(HERE, "bn.npz"), **out)
    print("wrote bn.npz", len(cfgs), "cases")


if __name__ == "__main__":
    main()
