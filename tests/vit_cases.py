"""ViT-B/16-shaped operator cases (BASELINE.json configs[3]): 16 images x 197 tokens = 3152 rows of the flattened
[B*T, K] GEMM (the reference asserts on 3-D activations, SURVEY.md 0.9: parity is defined on x.reshape(-1, K)),
K -> N in {768 -> 768, 768 -> 3072, 3072 -> 768}, E3M4.  The reference computes output columns independently
(approx/approx_calculation.py:946-971), so a COLS-column sample of each layer pins the whole layer.
Shared by tests/golden/make_vit_ops.py (reference side, build container) and the GPU tests (same seeds)."""
import torch

ROWS, COLS = 3152, 16
ROW_STEP, LN_ROW_STEP = 4, 48     # the golden file keeps every 4th (48th) output row
SHAPES = [(768, 768), (768, 3072), (3072, 768)]
E, M = 3, 4
RUN_APPROX = dict(approx_flag=True, quantize_after_mult_and_add=False, res_quantizer_flag=True,
                  original_quantize_res=False)
RUN_PLAIN = dict(approx_flag=False, quantize_after_mult_and_add=False, res_quantizer_flag=True,
                 original_quantize_res=True)        # shipped image_net.sh: res_quantizer(x @ w.T + b)
RUN_PLAIN_NOQ = dict(approx_flag=False, quantize_after_mult_and_add=False, res_quantizer_flag=False,
                     original_quantize_res=True)    # x @ w.T + b, result not re-quantized


def approx_params(**kw):
    d = dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=True, with_approx=True, with_s2nn2s_opt=True,
             sim_hw_add_OFUF=False, with_OF_opt=False, with_UF_opt=False, golden_clip_OF=False,
             quant_btw_mult_accu=True, debug_mode=False, self_check_mode=False)
    d.update(kw)
    return d


def operands(K, N, rows=ROWS):
    """x: token activations (LayerNorm-like scale, or post-GELU for the 3072-wide input: ~half non-positive values
    squashed towards zero); w: [N, K] at the ViT init scale; b: [N]."""
    g = torch.Generator().manual_seed(7000 + K + N)
    x = torch.randn(rows, K, generator=g)
    if K == 3072:
        x = torch.nn.functional.gelu(x)
    w = torch.randn(N, K, generator=g) * 0.02
    b = torch.randn(N, generator=g) * 0.02
    return x, w, b


def layernorm_operands(rows=ROWS, hidden=768):
    g = torch.Generator().manual_seed(7100)
    x = torch.randn(rows, hidden, generator=g) * 1.7 + 0.1
    w = torch.rand(hidden, generator=g) * 0.5 + 0.75
    b = torch.randn(hidden, generator=g) * 0.1
    return x, w, b


# BASELINE.json configs[4] (BERT-base, non-approx / tcgen05 path): seq 128 x batch 32 = 4096 rows
BERT_ROWS = 4096
BERT_SHAPES = [(768, 768), (768, 3072), (3072, 768)]
