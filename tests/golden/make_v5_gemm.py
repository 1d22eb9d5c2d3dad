"""Generates tests/golden/v5_gemm.npz by driving the UNMODIFIED reference's v5 GEMM,
custom_matmul_vectorize of approx/approx_matmul_whole_v5.py:10-152 -- encode (clip_OF), integer add of the codes
+ comp table, the hardware-adder overflow / underflow simulation of approx_mult_new (:155-182), decode, fp32 sum --
on K > 1 GEMMs: 3 formats x 2 biases x {v5 comp table, zero table} x all 8 (sim_hw_add_OFUF, with_OF_opt, with_UF_opt)
combinations, operands on the format's grid (zeros, sub-normals, both signs, top binade) and off it.

approx_mult_new's with_OF_opt line names an undefined module-level tensor `A` (:173) and raises NameError as
shipped; as in make_v5_ofuf.py the generator defines `A` in the module's namespace so that the unmodified line
executes with its evident meaning.

    python tests/golden/make_v5_gemm.py
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402


def grid_values(rng, E, M, bias, shape, zero_frac=0.15):
    """Random values of the (E, M, bias) format: codes uniform over the whole value space, random signs."""
    expo = rng.integers(0, 2 ** E, size=shape)
    mant = rng.integers(0, 2 ** M, size=shape)
    v = np.where(expo == 0, 2.0 ** (1 - bias) * mant / 2 ** M, 2.0 ** (expo - bias) * (1 + mant / 2 ** M))
    v = v * rng.choice([-1.0, 1.0], size=shape)
    v[rng.random(shape) < zero_frac] = 0.0
    return v.astype(np.float32)


def main():
    ref_shim.load()
    v5 = importlib.import_module("approx.approx_matmul_whole_v5")
    v5.A = torch.zeros(1)
    rng = np.random.default_rng(505)
    out, cfgs = {}, []
    i = g = 0
    for (E, M) in [(4, 3), (3, 4), (2, 5)]:
        for bias in (2 ** (E - 1), 2 ** (E - 1) + 2):
            for ongrid in (1, 0):
                A = grid_values(rng, E, M, bias, (24, 40))
                B = grid_values(rng, E, M, bias, (40, 12))
                if not ongrid:
                    A = (A * (1 + rng.standard_normal(A.shape) * 0.01)).astype(np.float32)
                    B = (B * (1 + rng.standard_normal(B.shape) * 0.01)).astype(np.float32)
                out[f"A{g}"], out[f"B{g}"] = A, B
                for wc in (1, 0):
                    table = (v5.get_comp_table_NN(E, M, True, 3, "cpu").cpu().to(torch.int32) if wc
                             else torch.zeros((2 ** M, 2 ** M), dtype=torch.int32))
                    for sim in (0, 1):
                        for of in (0, 1):
                            for uf in (0, 1):
                                r = v5.custom_matmul_vectorize(torch.from_numpy(A), torch.from_numpy(B), E, M, bias, table,
                                                               sim_hw_add_OFUF=bool(sim), with_OF_opt=bool(of),
                                                               with_UF_opt=bool(uf))
                                out[f"r{i}"] = r.numpy().astype(np.float32)
                                cfgs.append((E, M, bias, ongrid, wc, sim, of, uf, g))
                                i += 1
                g += 1
    out["cfgs"] = np.array(cfgs, np.int32)
    np.savez_compressed(os.path.join(HERE, "v5_gemm.npz"), **out)
    print("v5_gemm:", i, "cases over", g, "operand sets")


if __name__ == "__main__":
    main()
