/*
 * fp8approx_b200.h -- C ABI of libfp8approx_b200.so: the B200 (sm_100a) implementation of the
 * approx-FP8 simulated matmul hot path of revollllt/FP8_quantization.
 *
 * This is the drop-in boundary.  The reference is pure Python/PyTorch, so there is no existing FFI:
 * each entry point below replaces a *Python call edge* of the reference (file:line cited per
 * function; paths relative to the reference root) and is bound from Python with ctypes
 * (fp8_quantization_b200/_lib.py; the reference-side stub is shown in INTEGRATION.md).
 *
 * Conventions
 *   - plain C types only; every `const float*` / `float*` / `int32_t*` is a DEVICE pointer unless the
 *     parameter name starts with `host_`;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); every call is
 *     asynchronous on that stream and performs no host synchronisation and no allocation, except
 *     the first call for a new (format, flags) combination, which builds and caches the pair
 *     table (one small cudaMalloc + one kernel + one 4-byte read-back);
 *   - the caller owns all buffers, including the scratch `workspace` whose size is returned by the
 *     matching *_workspace_bytes function (only if the caller opts in with b200_set_autotune(1) does the table
 *     GEMM's first call per problem shape also time candidate tile configurations);
 *   - return value: B200_OK or a negative B200_ERR_* code; b200_last_error() gives the text.
 *     The Python host maps codes to the exceptions the reference raises at the same condition.
 */
#ifndef FP8APPROX_B200_H
#define FP8APPROX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_INVALID_FORMAT (-1) /* ValueError("Invalid combination of expo_width and mant_width"), approx_matmul_whole_v9.py:590 */
#define B200_ERR_UNBOUND_TABLE (-2)  /* UnboundLocalError: withComp and dnsmp_factor < 3, approx_matmul_whole_v9.py:566-572,575-586 */
#define B200_ERR_SHAPE (-3)          /* AssertionError: A.shape[1] == B.shape[0], approx_matmul_whole_v9.py:20 */
#define B200_ERR_ARGUMENT (-4)
#define B200_ERR_WORKSPACE (-5)
#define B200_ERR_CUDA (-6)           /* RuntimeError */
#define B200_ERR_NO_DEVICE (-7)      /* no sm_100 device: there is NO CPU fallback */

/* custom_approx_params, utils/click_options.py:589-606 (13 keys) + run-mode bits.
 * With ofuf_mode == 0 (the default) sim_hw_add_OFUF / with_OF_opt / with_UF_opt are accepted and IGNORED exactly
 * like approx_v9 does (approx_matmul_whole_v9.py:15 vs :20-169).  ofuf_mode == 1 opts into the last LIVE definition
 * of those flags, the v5 integer-add GEMM (see b200_approx_gemm below). */
typedef struct b200_approx_params {
    int32_t expo_width;
    int32_t mant_width;
    int32_t dnsmp_factor;
    uint8_t withComp;
    uint8_t with_approx;
    uint8_t with_s2nn2s_opt;
    uint8_t sim_hw_add_OFUF;
    uint8_t with_OF_opt;
    uint8_t with_UF_opt;
    uint8_t golden_clip_OF;
    uint8_t quant_btw_mult_accu;
    uint8_t self_check_mode;     /* prints nothing here; kept for struct parity */
    uint8_t single_column_quirk; /* 1: int32-tensor-bias numerics of the N==1 branch, approx_calculation.py:801-809 */
    uint8_t allow_unpinned_format; /* 1: accept (E,M) outside {(4,3),(3,4),(2,5)} with an all-zero table (no reference answer) */
    uint8_t force_generic;       /* 1: skip the table kernel, use the per-pair arithmetic kernel (tests) */
    uint8_t ofuf_mode;           /* 0: approx_v9 (OF/UF flags ignored).  1: the v5 integer-add GEMM with the simulated
                                    hardware-adder overflow / underflow, approx_matmul_whole_v5.py:10-182 */
    uint8_t force_table16;       /* 1: never take the byte-cell variant of the table kernel (tests, before / after runs).
                                    The variant computes the same bits; it is picked on the device for mant_width >= 4 */
    uint8_t prepared_b_ready;    /* see prepared_b */
    const int8_t *custom_table;  /* HOST pointer or NULL.  Non-NULL: use this 2^M x 2^M int8 table (the
                                    error_table_NN argument of custom_matmul_vectorize,
                                    approx_matmul_whole_v9.py:12) instead of selecting by withComp/dnsmp_factor */
    /* Optional (opt-in; NULL = off): a caller-owned DEVICE buffer of b200_approx_gemm_prepared_b_bytes(K, N)
     * (b200_approx_conv2d_prepared_b_bytes(geom) for the conv form) in which the table path keeps the decoded right
     * operand -- the weights of a layer whose ranges are fixed (hijacker.py:117-123 re-quantizes and the reference's
     * loop re-decodes them in every forward).  prepared_b_ready = 0: this call decodes B into it; 1: this call reuses
     * what an earlier call with the SAME B, exponent biases, format and flags left there and skips the B passes.  The
     * caller answers for staleness.  Same bits either way. */
    void *prepared_b;
    uint64_t prepared_b_bytes;
} b200_approx_params;

/* ---- library / device ---------------------------------------------------------------------- */
const char *b200_version(void);
const char *b200_last_error(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches claim) */
uint64_t b200_kernel_launches(void);
/* name of the kernel path the last b200_approx_gemm / conv call took: "lut", "generic", ... */
const char *b200_last_path(void);

/* Measured (tile, split-K) choice of the table GEMM.  OFF by default: the cost model alone decides and no entry
 * point ever synchronises with the host.  With autotune on (b200_set_autotune(1), or B200_AUTOTUNE=1 in the
 * environment) the FIRST call for a new problem shape times the cost model's best few configurations on the
 * caller's operands and caches the winner -- a handful of extra runs and host synchronisations, once per shape,
 * never while the stream is being captured into a CUDA graph.  All configurations compute identical bits.
 * Returns the previous setting. */
int b200_set_autotune(int on);
/* number of candidate configurations timed so far (0 = the cost model alone has decided everything) */
uint64_t b200_autotune_runs(void);

/* ---- A5: get_error_table_NN, approx/approx_matmul_whole_v9.py:555-592 ------------------------ */
/* host_out: 2^M * 2^M int8, row-major, HOST memory. */
int b200_error_table(int expo_width, int mant_width, int withComp, int dnsmp_factor,
                     int allow_unpinned_format, int8_t *host_out);

/* get_comp_table_NN of the v5 integer-add multiplier, approx/approx_matmul_whole_v5.py:516-542 (used when
 * ofuf_mode == 1): 2^M * 2^M int8 to HOST memory; without withComp the all-zero table the reference assigns before
 * it raises. */
int b200_comp_table_v5(int expo_width, int mant_width, int withComp, int dnsmp_factor, int8_t *host_out);

/* ---- A8: quantize_to_fp8_ste_MM, quantization/quantizers/fp8_quantizer.py:97-173 ------------- */
/* maxval: device [n_channels]; n_channels == 1 -> per tensor, else channel = (i / inner) % n_channels
 * (maxval viewed as [C,1,...,1], :114-115).  bias_out: device [n_channels] fp32 (integer valued),
 * the second return value of the reference function (-> FPQuantizer.custom_bias, :244). */
int b200_fp8_fakequant(const float *x, float *y, int64_t n, const float *maxval, int64_t n_channels,
                       int64_t inner, int n_bits, float mantissa_bits, int sign_bits, float *bias_out,
                       void *stream);

/* ---- A2/A3/A4/A11: field codec (elementwise; used by tests and tools) ------------------------ */
/* mode 0: float_to_fpany_absint_torch (approx_matmul_whole_v9.py:233-291)
 * mode 1: float_to_fpany_absint_torch_allnorm (approx/fp_any_int_v5.py:55-110) */
int b200_fpany_encode(const float *v, int64_t n, int expo_width, int mant_width, int bias, int mode,
                      int tensor_bias_quirk, int clip_OF, int32_t *expo_out, int32_t *mant_out,
                      void *stream);
/* fpany_absint_to_float_torch (:295-329) / _allnorm (fp_any_int_v5.py:116-162); sign may be NULL */
int b200_fpany_decode(const int32_t *expo, const int32_t *mant, const float *sign, int64_t n,
                      int expo_width, int mant_width, int bias, int mode, float *out, void *stream);
/* quant_to_fp_any_vectorize_torch (:333-362) */
int b200_fpany_quant(const float *v, int64_t n, int expo_width, int mant_width, int bias,
                     int tensor_bias_quirk, int clip_OF, float *out, void *stream);

/* ---- A7: hardware adder overflow / underflow clipping (only live in v5) ------------------------ */
/* Inside the GEMM / conv: b200_approx_params.ofuf_mode = 1 (above).  Elementwise, on integer codes: */
/* approx_mult_new, approx/approx_matmul_whole_v5.py:155-182, elementwise on the integer code domain:
 *   t = x_int + y_int + B_neg + table[x_mant, y_mant];
 *   sim_hw_add_OFUF: of = t > 2^(E+M)-1, uf = t < 0; t = t mod 2^(E+M) (floor-mod);
 *   with_OF_opt: t = of ? 2^(E+M)-1 : t;   with_UF_opt: t = uf ? t mod 2^M : t.
 * approx_v9 accepts these three flags and never reads them (approx_matmul_whole_v9.py:15 vs :20-169), so
 * b200_approx_gemm ignores them too unless ofuf_mode == 1; this entry point is the same clip on raw codes.  x_mant / y_mant may
 * be NULL (-> x_int mod 2^M); table_dev: DEVICE int8 [2^M, 2^M] or NULL (zeros).  (The with_OF_opt line of
 * the reference names an undefined tensor, :173; its evident meaning is implemented -- see
 * tests/golden/make_v5_ofuf.py.) */
int b200_approx_mult_v5(const int32_t *x_int, const int32_t *y_int, const int32_t *x_mant,
                        const int32_t *y_mant, const int8_t *table_dev, int64_t n, int B_neg, int expo_width,
                        int mant_width, int sim_hw_add_OFUF, int with_OF_opt, int with_UF_opt, int32_t *out,
                        void *stream);

/* ---- A1/A6: approx_multiply -> custom_matmul_vectorize ---------------------------------------- */
/* Replaces the whole per-column Python loop of approx_multiply
 * (approx/approx_calculation.py:946-971 Linear, :774-799 BN-conv; one column =
 * approx/approx_matmul_whole_v9.py:10-169) by ONE call.
 *   A  : [Mrows, K] fp32, element (m,k) at A[m*lda + k]           (activations / im2col rows)
 *   B  : [K, N] fp32, element (k,n) at B[k*sb_k + n*sb_n]          (weight.t(): sb_k=1, sb_n=ldw)
 *   C  : [Mrows, N] fp32, element (m,n) at C[m*ldc + n]
 *   bB : device int32 [N], per-column weight exponent bias (y_bias[i], :952); bA,bR scalars.
 * C[m,n] = sum_{k ascending, fp32} addend(A[m,k], B[k,n]).
 *
 * p->ofuf_mode == 1 (opt-in): the addend is the v5 GEMM's, custom_matmul_vectorize of
 * approx/approx_matmul_whole_v5.py:10-152 with approx_mult_new (:155-182):
 *   (eA,mA) = enc(a), (eB,mB) = enc(b)  [clip_OF = True];   t = (eA*2^M+mA) + (eB*2^M+mB) + B_neg + comp[mA,mB];
 *   sim_hw_add_OFUF: of = t > 2^(E+M)-1, uf = t < 0, t = t mod 2^(E+M); with_OF_opt: of -> 2^(E+M)-1;
 *   with_UF_opt: uf -> t mod 2^M;   addend = dec(t) * sign(a) * sign(b);   fp32 sum, k ascending.
 * v5 has ONE custom_bias; with the layers' per-operand biases B_neg = -((bA + bB[n] - bR) << M) (approx_v9's exponent
 * formula, v9 :66-67) and the result decodes with bR -- equal biases reproduce v5 bit for bit (tests/golden/
 * v5_gemm.npz).  comp = p->custom_table, else v5's own table choice (:516-542; all zero without withComp).  A zero
 * operand is an ordinary code in v5 (its product is NOT zero); with_approx / with_s2nn2s_opt / quant_btw_mult_accu /
 * golden_clip_OF / single_column_quirk do not exist in v5 and are ignored in this mode.  Operands on the format's grid
 * run on the table kernel (a second pair-table flavour), anything else on the arithmetic kernel. */
size_t b200_approx_gemm_workspace_bytes(int64_t Mrows, int64_t K, int64_t N);
int b200_approx_gemm(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n,
                     float *C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N, int bA,
                     const int32_t *bB, int bR, const b200_approx_params *p, void *workspace,
                     size_t workspace_bytes, void *stream);
/* Same, with the biases left ON THE DEVICE as the fp32 tensors the fake-quantizer produced
 * (FPQuantizer.custom_bias, fp8_quantizer.py:244): removes the three .item() host syncs per column
 * of approx_calculation.py:952.  bA_dev / bR_dev: device [1] fp32 or NULL (-> 2^(E-1), :938-939);
 * bB_dev: device fp32, column n at bB_dev[n*bB_stride]; values are truncated like .to(torch.int32). */
int b200_approx_gemm_fbias(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n,
                           float *C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N,
                           const float *bA_dev, const float *bB_dev, int64_t bB_stride,
                           const float *bR_dev, const b200_approx_params *p, void *workspace,
                           size_t workspace_bytes, void *stream);

/* ---- A6 (second mode): quantize_after_mult_and_add -------------------------------------------- */
/* C[m,n] = Q(sum_k Q(A[m,k]*B[k,n])), Q = quantize_to_fp8_ste_MM(., n_bits, maxval, mantissa_bits,
 * sign_bits)[0] with the RESULT quantizer's per-tensor maxval (device [1]); replaces the column loop
 * approx/approx_calculation.py:959-967 (Linear) / :787-795 (BN-conv).  Same operand layout as
 * b200_approx_gemm.  fp32 accumulation, k ascending. */
int b200_qamaa_gemm(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, float *C,
                    int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *maxval_dev, int n_bits,
                    float mantissa_bits, int sign_bits, void *stream);

/* Same function on the TABLE kernel: the operand exponent biases (the fake-quantizers' custom_bias
 * tensors, as in b200_approx_gemm_fbias; bA_dev may be NULL -> 2^(E-1)) fix the index range of a pair
 * table that is rebuilt on the device from `maxval` at every call (maxval moves during calibration).
 * Operands off that grid, an unsigned grid or n_bits != 8 -> the arithmetic kernel computes the result
 * (decided on the device / from the arguments; never on the host CPU). */
size_t b200_qamaa_gemm_workspace_bytes(int64_t Mrows, int64_t K, int64_t N);
int b200_qamaa_gemm_fbias(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, float *C,
                          int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *maxval_dev, int n_bits,
                          float mantissa_bits, int sign_bits, const float *bA_dev, const float *bB_dev,
                          int64_t bB_stride, void *workspace, size_t workspace_bytes, void *stream);

/* Allocation-free variant of b200_qamaa_gemm_fbias.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  The scratch (flags, decoded operand planes, pair table, split-K partial sums) is a
 * file-scope `static __device__` array of b200_static_scratch_bytes() = 64 MiB inside the library -- every GPU that
 * loads the library holds it until the process exits -- so there is no workspace argument and the call performs no
 * cudaMalloc / cudaFree / allocating library routine.  Same arguments, same results (bit for bit) as
 * b200_qamaa_gemm_fbias given a workspace of b200_qamaa_gemm_workspace_bytes(Mrows, K, N); bB_dev is required.
 * Limit: b200_qamaa_gemm_workspace_bytes(Mrows, K, N) <= b200_static_scratch_bytes() (roughly
 * round_up(K,32) * (round_up(Mrows,256) + round_up(N,128)) <= 5.9 M words, e.g. 1024 x 1024 x 1024 or
 * 3152 x 768 x 768); larger problems return B200_ERR_WORKSPACE with a message naming both sizes.
 * All calls on a device share the one array, which keeps the previous call's contents: what must start at zero is
 * cleared on the stream at the start of each call, everything else is rewritten before it is read.  Calls that
 * OVERLAP on the GPU (different streams, or host threads) race on it: serialise them. */
size_t b200_static_scratch_bytes(void);
/* bytes of b200_approx_params.prepared_b for a [K, N] right operand (the conv form: b200_approx_conv2d_prepared_b_bytes) */
size_t b200_approx_gemm_prepared_b_bytes(int64_t K, int64_t N);
int b200_qamaa_gemm_fbias_static(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, float *C,
                                 int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *maxval_dev, int n_bits,
                                 float mantissa_bits, int sign_bits, const float *bA_dev, const float *bB_dev,
                                 int64_t bB_stride, void *stream);

/* ---- K3: the non-approx matmul on the tensor cores (tcgen05) ---------------------------------- */
/* C = A @ B (+ bias) for FAKE-QUANTIZED operands; replaces `x @ y[:, i:i+1]` per column
 * (approx/approx_calculation.py:968-969 Linear, :796-797 BN-conv, :996 / :811 single column) and
 * `self.multiply` = torch.matmul (:704-705, conv without BN).  Same operand layout as b200_approx_gemm.
 * Operands are converted to bf16 on the device (exact for every FP8-grid value); tcgen05.mma kind::f16
 * accumulates in fp32 in tensor memory.  If ANY operand is not exactly representable in bf16 the call
 * computes the product with fp32 CUDA-core arithmetic instead (decided on the device, no host sync).
 * bias_dev: device [N] fp32 added to every row, or NULL.
 * Tolerance vs the reference's fp32 matmul: |C - ref| <= 1e-6 * sum_k |a_k * b_k| (summation order). */
size_t b200_fq_gemm_tc_workspace_bytes(int64_t Mrows, int64_t K, int64_t N);
int b200_fq_gemm_tc(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, float *C,
                    int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *bias_dev,
                    void *workspace, size_t workspace_bytes, void *stream);

/* Prepared right operand (opt-in): once the ranges are fixed the fake-quantized weights are the same tensor in every
 * call, so their bf16 plane can be built once.  b200_fq_gemm_tc_prepare_b writes it (and whether it is exact in bf16)
 * to `bplane` (device, 256-byte aligned, b200_fq_gemm_tc_bplane_bytes(K, N) bytes); b200_fq_gemm_tc_pre is
 * b200_fq_gemm_tc without the convert pass over B (B itself is still read by the fp32 fallback kernel). */
size_t b200_fq_gemm_tc_bplane_bytes(int64_t K, int64_t N);
int b200_fq_gemm_tc_prepare_b(const float *B, int64_t sb_k, int64_t sb_n, int64_t K, int64_t N, void *bplane,
                              size_t bplane_bytes, void *stream);
int b200_fq_gemm_tc_pre(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, const void *bplane,
                        float *C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *bias_dev,
                        void *workspace, size_t workspace_bytes, void *stream);

/* Fused input fake-quantizer: the hijacker's `x = self.activation_quantizer(x)`
 * (quantization/hijacker.py:82-88 -> fp8_quantizer.py:97-173) evaluated inside the operand decode pass of the
 * GEMM / conv instead of as a separate pass over the activations.  A / x then hold the RAW activations;
 * maxval_dev is the quantizer's per-tensor range, bias_dev the exponent bias the stand-alone quantizer derived
 * from it (FPQuantizer.custom_bias; also what bA_dev must point at).  Results are bit-identical to
 * quantizing first.  Single-output-channel (depthwise) groups quantize while the input is decoded into the word plane
 * of their table kernel. */
typedef struct b200_act_quant {
    const float *maxval_dev; /* device [1] */
    const float *bias_dev;   /* device [1] */
    float mantissa_bits;
    int32_t n_bits;
    int32_t sign_bits;
} b200_act_quant;
int b200_approx_gemm_fbias_fq(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, float *C,
                              int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *bA_dev,
                              const float *bB_dev, int64_t bB_stride, const float *bR_dev,
                              const b200_approx_params *p, const b200_act_quant *aq, void *workspace,
                              size_t workspace_bytes, void *stream);

/* K3 with the neighbouring fake-quantizers fused (opt-in): bplane (may be NULL) as in b200_fq_gemm_tc_pre; aq_in: A
 * holds RAW activations and `x = self.activation_quantizer(x)` (quantization/hijacker.py:82-88) runs inside the bf16
 * convert pass; aq_out: `res = self.res_quantizer(res)` (:93-94) runs in the epilogue after the bias.  Either may be
 * NULL.  Bit-identical to running the quantizers as separate passes. */
int b200_fq_gemm_tc_ex(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, const void *bplane,
                       const b200_act_quant *aq_in, const b200_act_quant *aq_out, float *C, int64_t ldc,
                       int64_t Mrows, int64_t K, int64_t N, const float *bias_dev, void *workspace,
                       size_t workspace_bytes, void *stream);

/* ---- N2: fused epilogue (OPT-IN: the entry points above keep the reference's pass structure) --- */
/* What the reference does to a layer's product sums before the next layer reads them, evaluated in the epilogue of
 * the GEMM / conv kernels instead of as separate passes over the output, in the reference's operation order
 * (bit-identical to the unfused sequence):
 *   v = acc;  v += bias[c]           `output += bias`, approx/approx_calculation.py:1020-1021 (Linear), :911-912 (conv)
 *   v = fma(v, alpha[c], beta'[c])   F.batch_norm(eval), quantization/quantized_folded_bn.py:66-72, op order of ATen's CPU
 *                                    kernel: invstd = 1/sqrt(var+eps), alpha = gamma*invstd, beta' = fma(-mean, alpha, beta)
 *                                    (== b200_bn_act below)
 *   v = relu / relu6 (v)             quantized_folded_bn.py:74-75, quantization/hijacker.py:108-109
 *   store                            GEMM form: C[m*ldc + n];  conv form: contiguous NCHW [batch, C_out, Ho, Wo] (the
 *                                    view + permute of approx_calculation.py:909-910 made real)
 * All pointers are DEVICE pointers to [N] (GEMM) / [C_out] (conv) fp32 arrays, or NULL: bias NULL = no bias;
 * bn_mean NULL = no batch norm (bn_var then ignored); bn_gamma / bn_beta NULL = 1 / 0. */
typedef struct b200_epilogue {
    const float *bias;
    const float *bn_mean, *bn_var, *bn_gamma, *bn_beta;
    float bn_eps;
    int32_t act; /* 0 none, 1 ReLU, 2 ReLU6 */
} b200_epilogue;
/* b200_approx_gemm_fbias_fq with the epilogue (aq and ep may each be NULL) */
int b200_approx_gemm_fbias_ep(const float *A, int64_t lda, const float *B, int64_t sb_k, int64_t sb_n, float *C,
                              int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float *bA_dev,
                              const float *bB_dev, int64_t bB_stride, const float *bR_dev,
                              const b200_approx_params *p, const b200_act_quant *aq, const b200_epilogue *ep,
                              void *workspace, size_t workspace_bytes, void *stream);

/* ---- A9: conv form, QCustomBNConv2dTorch.run_forward, approx/approx_calculation.py:822-917 -------- */
typedef struct b200_conv_geom {
    int64_t batch;
    int32_t C_in, H, W;          /* input  [batch, C_in, H, W], contiguous NCHW fp32 */
    int32_t C_out, kh, kw;       /* weight [C_out, C_in/groups, kh, kw], contiguous   */
    int32_t stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, groups;
} b200_conv_geom;
/* rows = batch*Ho*Wo, Kg = (C_in/groups)*kh*kw, Ng = C_out/groups; any out pointer may be NULL */
int b200_conv_out_shape(const b200_conv_geom *g, int64_t *rows, int32_t *Ho, int32_t *Wo, int32_t *Kg,
                        int32_t *Ng);
/* im2col (:724-747): out [rows, C_in*kh*kw] row-major, K order (c, ky, kx), zero padded borders */
/* bytes of b200_approx_params.prepared_b for every group of this conv (0 where the path has no decoded right operand:
 * single-column groups) */
size_t b200_approx_conv2d_prepared_b_bytes(const b200_conv_geom *g);
int b200_im2col(const float *x, float *out, const b200_conv_geom *g, void *stream);
/* exact (non-approx, non-qamaa) branch of the group loop for groups > 1, every group in one launch:
 * out[m, n] = sum_k col[m, grp(n)*Kg + k] * w[n, k]  (`input_col_group @ weight_group.T`,
 * approx/approx_calculation.py:796-797, :810-811), implicit im2col, fp32 FMA with k ascending.
 * out is [rows, C_out] row-major.  (groups == 1 goes through b200_im2col + b200_fq_gemm_tc.) */
int b200_exact_grouped_conv2d(const float *x, const float *w, float *out, const b200_conv_geom *g,
                              void *stream);
/* im2col + group loop + approx_multiply (:846-906) in one call; out is [rows, C_out] row-major (the
 * tensor the reference then views as [B,Ho,Wo,C_out] and permutes, :909-910).  Groups with more than
 * one output channel: table GEMM over an implicit im2col view (no im2col matrix in HBM).  Groups
 * with ONE output channel (depthwise): the reference's int32-tensor-bias numerics (:800-809), one
 * launch for all groups; p->single_column_quirk is set by the library from the geometry.
 * bB_dev: per-output-channel weight bias, channel c at bB_dev[c*bB_stride] (stride 0 = per tensor). */
size_t b200_approx_conv2d_workspace_bytes(const b200_conv_geom *g);
int b200_approx_conv2d_fbias(const float *x, const float *w, float *out, const b200_conv_geom *g,
                             const float *bA_dev, const float *bB_dev, int64_t bB_stride,
                             const float *bR_dev, const b200_approx_params *p, void *workspace,
                             size_t workspace_bytes, void *stream);
/* same, with the input fake-quantizer fused (b200_act_quant above) */
int b200_approx_conv2d_fbias_fq(const float *x, const float *w, float *out, const b200_conv_geom *g,
                                const float *bA_dev, const float *bB_dev, int64_t bB_stride,
                                const float *bR_dev, const b200_approx_params *p, const b200_act_quant *aq,
                                void *workspace, size_t workspace_bytes, void *stream);
/* b200_approx_conv2d_fbias_fq with the fused epilogue (b200_epilogue above; aq and ep may each be NULL).  With
 * ep != NULL `out` is the FINAL tensor: contiguous NCHW [batch, C_out, Ho, Wo] holding act(bn(conv + bias)); depthwise
 * groups are supported (aq must then be NULL).  With ep == NULL: [rows, C_out] row-major as above. */
int b200_approx_conv2d_fbias_ep(const float *x, const float *w, float *out, const b200_conv_geom *g,
                                const float *bA_dev, const float *bB_dev, int64_t bB_stride,
                                const float *bR_dev, const b200_approx_params *p, const b200_act_quant *aq,
                                const b200_epilogue *ep, void *workspace, size_t workspace_bytes, void *stream);
/* quantize_after_mult_and_add conv (:787-795); groups with one output channel are refused
 * (B200_ERR_ARGUMENT): the reference computes those exactly (:810-811). */
int b200_qamaa_conv2d(const float *x, const float *w, float *out, const b200_conv_geom *g,
                      const float *maxval_dev, int n_bits, float mantissa_bits, int sign_bits, void *stream);
/* table-kernel variant (see b200_qamaa_gemm_fbias); bB_dev: per-output-channel weight bias */
size_t b200_qamaa_conv2d_workspace_bytes(const b200_conv_geom *g);
int b200_qamaa_conv2d_fbias(const float *x, const float *w, float *out, const b200_conv_geom *g,
                            const float *maxval_dev, int n_bits, float mantissa_bits, int sign_bits,
                            const float *bA_dev, const float *bB_dev, int64_t bB_stride, void *workspace,
                            size_t workspace_bytes, void *stream);

/* ---- N2: eval-mode BatchNorm + activation epilogue, quantization/quantized_folded_bn.py:66-78 ------- */
/* y = act(fma(x, alpha, beta')) with invstd = 1/sqrt(var+eps), alpha = gamma*invstd,
 * beta' = fma(-mean, alpha, beta): the op sequence of ATen's CPU batch_norm in inference mode (bit-identical
 * to the reference run on CPU tensors).  x: 4-D tensor given by element strides (batch sb, channel sc,
 * pixel sp with pixel = h*W + w, i.e. stride(h) == W*stride(w)); y: contiguous NCHW.  gamma / beta may be
 * NULL (1 / 0).  act: 0 none, 1 ReLU, 2 ReLU6. */
int b200_bn_act(const float *x, int64_t sb, int64_t sc, int64_t sp, float *y, int64_t B, int C, int64_t HW,
                const float *mean, const float *var, const float *gamma, const float *beta, float eps, int act,
                void *stream);

/* ---- measurement tool: CUDA-core integer issue peak (roofline denominator) ------ */
/* Synchronous.  tiops_out: IADD3/LOP3 lane-ops per second in Tiop/s over all SMs; ms_out may be NULL. */
int b200_int_peak_probe(int iters, double *tiops_out, double *ms_out);
/* Shared-memory lookup peak, the bound that binds the table GEMM (one LDS.U16 per MAC): conflict-free 16-bit
 * shared-memory loads per second over all SMs, in T lookups/s.  Synchronous. */
int b200_lds_peak_probe(int iters, double *tlookups_out, double *ms_out);
/* Between begin and end, every table-GEMM launch is bracketed by CUDA events on its own stream;
 * end synchronises those events and returns their summed duration and the launch count. */
int b200_profile_begin(void);
int b200_profile_end(double *total_ms, int *launches);

#ifdef __cplusplus
}
#endif
#endif /* FP8APPROX_B200_H */
