"""Generates tests/golden/v5_ofuf.npz: outputs of the UNMODIFIED reference function
approx_mult_new (approx/approx_matmul_whole_v5.py:155-182) -- the only live definition of the hardware
overflow / underflow clipping (sim_hw_add_OFUF, with_OF_opt, with_UF_opt).

The with_OF_opt branch of that function names a module-level tensor `A` that does not exist
(:173, `device=A.device`): called as is, it raises NameError.  The generator records that fact and
then defines `A` in the module's namespace (a CPU tensor) so that the unmodified line executes with
its evident meaning (clamp overflowed codes to max_norm_int).

Build container only:   python tests/golden/make_v5_ofuf.py
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402


def main():
    ref_shim.load()
    v5 = importlib.import_module("approx.approx_matmul_whole_v5")
    rng = np.random.default_rng(55)
    out = {}
    cfgs = []
    # the reference's own undefined-name behaviour
    try:
        v5.approx_mult_new(torch.zeros(2, dtype=torch.int32), torch.zeros(2, dtype=torch.int32), 0,
                           torch.zeros(2, dtype=torch.int32), torch.zeros(2, dtype=torch.int32),
                           torch.zeros((8, 8), dtype=torch.int8), True, True, False, 128, 127, 8)
        out["of_opt_raises_nameerror"] = np.array(0)
    except NameError:
        out["of_opt_raises_nameerror"] = np.array(1)
    v5.A = torch.zeros(1)
    i = g = 0
    for (E, M) in [(4, 3), (3, 4), (2, 5)]:
        for (wc, dn) in [(False, 3), (True, 3), (True, 5)]:
            # v5's own table choice (:516-542; it raises for withComp=False, where the table is all zero)
            table = (v5.get_comp_table_NN(E, M, wc, dn, "cpu").cpu().to(torch.int8) if wc
                     else torch.zeros((2 ** M, 2 ** M), dtype=torch.int8))
            pd = v5.param_prepare(E, M)
            for bias in (2 ** (E - 1), 2 ** (E - 1) + 3):
                n = 2048
                x_int = rng.integers(0, 2 ** (E + M), size=n).astype(np.int32)
                y_int = rng.integers(0, 2 ** (E + M), size=n).astype(np.int32)
                B_neg = -(bias << M)
                out[f"x{g}"], out[f"y{g}"], out[f"t{g}"] = x_int, y_int, table.numpy()
                for sim in (0, 1):
                    for of in (0, 1):
                        for uf in (0, 1):
                            xi, yi = torch.from_numpy(x_int), torch.from_numpy(y_int)
                            r = v5.approx_mult_new(xi, yi, B_neg, xi % (2 ** M), yi % (2 ** M), table, bool(sim),
                                                   bool(of), bool(uf), pd["OF_UF_mod"], pd["max_norm_int"],
                                                   pd["mant_scale"])
                            out[f"r{i}"] = r.numpy().astype(np.int32)
                            cfgs.append((E, M, int(wc), dn, B_neg, sim, of, uf, g))
                            i += 1
                g += 1
    out["cfgs"] = np.array(cfgs, np.int32)
    path = os.path.join(HERE, "v5_ofuf.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, i, "cases; with_OF_opt raises NameError in the reference:", int(out["of_opt_raises_nameerror"]))


if __name__ == "__main__":
    main()
