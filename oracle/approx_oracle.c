/*
 * approx_oracle.c -- CPU restatement (plain C) of the reference's approx-FP8 hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the shipped product (fp8_quantization_b200/) imports,
 * links or executes this file.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker / timed CPU baseline.
 *
 * Parity status: PINNED against outputs of the unmodified reference run in the build container
 * (tests/golden/make_golden.py imports /root/reference and writes the npz fixtures under tests/golden; see
 * tests/test_oracle_golden.py).  The reference itself has no asserted tests.
 *
 * Every function cites the reference file:line (paths relative to the reference root) it follows.
 * All arithmetic is done in IEEE fp32 exactly like the reference's torch elementwise ops; the
 * compile flags (-O2 -ffp-contract=off, no -ffast-math) forbid fused multiply-adds.
 *
 * Summation order: the reference uses torch.sum(dim=1) over an [M,K,1] fp32 tensor whose order is
 * backend specific.  The oracle sums k = 0..K-1 sequentially in fp32 (sum_mode 0, what the CUDA
 * kernels do) or in double (sum_mode 1, used by the tests to bound the order effect).  Whenever
 * every partial sum is exactly representable, all orders agree bit for bit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int expo_width;      /* E */
    int mant_width;      /* M */
    int with_approx;     /* approx_matmul_whole_v9.py:13 */
    int with_s2nn2s_opt; /* :14 */
    int golden_clip_OF;  /* :15 (only live OF flag in v9) */
    int quant_btw_mult_accu; /* :16 */
    int tensor_bias_quirk;   /* 1 = biases are int32 tensors (approx_calculation.py:801-809):
                                integer pow in param_prepare (approx_matmul_whole_v9.py:201-202) */
    int sum_mode;        /* 0 fp32 sequential, 1 double */
} orc_cfg;

/* torch `2.0 ** int_tensor` in fp32: exact powers of two, inf above, 0 far below. */
static float pow2f(int n)
{
    if (n > 127) return INFINITY;
    if (n < -149) return 0.0f;
    return ldexpf(1.0f, n);
}

/* torch integer pow `2 ** int_tensor` (int32): negative exponent -> 0, wraps above 2^31. */
/* torch.min(torch.max(x, minval), maxval), fp8_quantizer.py:120: torch's min / max propagate NaN */
static float clamp_nan(float x, float mn, float mx)
{
    return (x != x) ? x : fminf(fmaxf(x, mn), mx);
}

static int32_t ipow2(int n)
{
    if (n < 0) return 0;
    if (n >= 32) return 0;
    return (int32_t)(1u << n);
}

typedef struct {
    int mant_width, fp_bias, max_expo, max_mant, mant_scale;
    float max_norm, min_norm; /* as the fp32 comparison operands the reference ends up using */
} orc_param;

/* param_prepare, approx/approx_matmul_whole_v9.py:189-229 */
static orc_param orc_param_prepare(int E, int M, int bias, int tensor_bias_quirk)
{
    orc_param p;
    p.mant_width = M;
    p.fp_bias = bias;
    p.max_expo = (1 << E) - 1;
    p.max_mant = (1 << M) - 1;
    p.mant_scale = 1 << M;
    float top = 2.0f - ldexpf(1.0f, -M);
    if (!tensor_bias_quirk) {
        /* python ints: 2**(negative) is a python float, exact */
        p.max_norm = (float)(ldexp(1.0, p.max_expo - bias) * (2.0 - ldexp(1.0, -M)));
        p.min_norm = (float)ldexp(1.0, 1 - bias);
    } else {
        /* int32 tensors: integer pow, 2**negative == 0 */
        p.max_norm = (float)ipow2(p.max_expo - bias) * top;
        p.min_norm = (float)ipow2(1 - bias);
    }
    return p;
}

/* float_to_fpany_absint_torch, approx/approx_matmul_whole_v9.py:233-291 */
static void orc_encode(const orc_param *p, float v, int clip_OF, int32_t *expo_out, int32_t *mant_out)
{
    int ex;
    float m = frexpf(v, &ex);            /* :261 */
    if (v == 0.0f) { m = 0.0f; ex = 0; }
    int sub = fabsf(v) < p->min_norm;    /* :264 */
    int extra = p->fp_bias - 1 + p->mant_width; /* :269 */
    float x;
    if (sub) x = fabsf(m) * pow2f(ex + extra);              /* :274 ldexp == x * 2**n */
    else     x = (fabsf(m) * 2.0f - 1.0f) * pow2f(p->mant_width); /* :275 */
    float r = rintf(x);                  /* torch.round: half to even */
    if (r > (float)p->max_mant) r = (float)p->max_mant;     /* clamp(max=max_mant) :272-276 */
    int32_t mant = (int32_t)r;
    int32_t expo = sub ? 0 : ex + (p->fp_bias - 1);         /* :280 */
    if (clip_OF) {                        /* :283-286 */
        if (v < -p->max_norm || v > p->max_norm) { expo = p->max_expo; mant = p->max_mant; }
    }
    *expo_out = expo;
    *mant_out = mant;
}

/* fpany_absint_to_float_torch, approx/approx_matmul_whole_v9.py:295-329 */
static float orc_decode(const orc_param *p, float sign, int32_t expo, int32_t mant)
{
    float frac = (float)mant / (float)p->mant_scale;         /* :323 */
    float v;
    if (expo == 0) v = pow2f(1 - p->fp_bias) * frac;         /* :321-323 */
    else           v = pow2f(expo - p->fp_bias) * (1.0f + frac);
    return v * sign;                                         /* :325-327 */
}

/* quant_to_fp_any_vectorize_torch, approx/approx_matmul_whole_v9.py:333-362 */
static float orc_quant(const orc_param *p, float v, int clip_OF)
{
    int32_t e, m;
    orc_encode(p, v, clip_OF, &e, &m);
    float sign = (v < 0.0f) ? -1.0f : 1.0f;                  /* :357 */
    return orc_decode(p, sign, e, m);
}

/* mult_result_mant, approx/approx_matmul_whole_v9.py:178-184.  Negative indices wrap (python). */
static float orc_mult_mant(int M, int32_t mA, int32_t mB, const int8_t *T, int with_approx)
{
    float s = ldexpf(1.0f, -M);
    float fa = 1.0f + (float)mA * s;
    float fb = 1.0f + (float)mB * s;
    float prod = fa * fb;
    if (!with_approx) return prod;
    int n = 1 << M;
    int ia = mA < 0 ? mA + n : mA;
    int ib = mB < 0 ? mB + n : mB;
    float t = (ia >= 0 && ia < n && ib >= 0 && ib < n) ? (float)T[ia * n + ib] : NAN; /* IndexError */
    return prod - s * t;
}

/* one operand pair of custom_matmul_vectorize, approx/approx_matmul_whole_v9.py:29-110 */
static float orc_pair(const orc_cfg *c, const orc_param *pA, const orc_param *pB, const orc_param *pR,
                      const int8_t *T, float a, float b)
{
    const int M = c->mant_width;
    float g = a * b;                                         /* :30-31 */
    int zero = (g == 0.0f);                                  /* :32 */
    if (c->quant_btw_mult_accu) g = orc_quant(pR, g, c->golden_clip_OF); /* :35-36 */
    int subA = fabsf(a) < pA->min_norm;                      /* :47 */
    int subB = fabsf(b) < pB->min_norm;                      /* :48 */
    float a2 = a, b2 = b;
    if (c->with_s2nn2s_opt) {                                /* :52-54 */
        if (subA) a2 = a * (float)pA->mant_scale;
        if (subB) b2 = b * (float)pB->mant_scale;
    }
    int32_t eA, mA, eB, mB;
    orc_encode(pA, a2, 0, &eA, &mA);                         /* :58 */
    orc_encode(pB, b2, 0, &eB, &mB);                         /* :59 */
    int32_t e = eA + eB - (pA->fp_bias + pB->fp_bias - pR->fp_bias); /* :66-67 */
    float sgn = (g < 0.0f) ? -1.0f : 1.0f;                   /* :68 */
    float r = pow2f(e - pR->fp_bias) * orc_mult_mant(M, mA, mB, T, c->with_approx) * sgn; /* :73-77 */
    if (c->with_s2nn2s_opt) {
        if (subA) r = r / (float)pA->mant_scale;             /* :79 */
        if (subB) r = r / (float)pB->mant_scale;             /* :80 */
        if (zero) r = 0.0f;                                  /* :81 */
    } else {
        int norm = (eA > 0) && (eB > 0) && (fabsf(g) >= pR->min_norm); /* :86-87 */
        if (!norm) r = g;                                    /* :89-98 */
    }
    if (c->quant_btw_mult_accu) r = orc_quant(pR, r, c->golden_clip_OF); /* :107-108 */
    return r;
}

/* ------------------------------------------------------------------------------------------ */
/* Exported entry points (ctypes).  All matrices are fp32; element (i,j) of X is X[i*sx0+j*sx1]. */
/* ------------------------------------------------------------------------------------------ */

/* elementwise codec, for the K5 tests: approx_matmul_whole_v9.py:233-291 / :295-329 / :333-362 */
void orc_encode_array(const float *v, int64_t n, int E, int M, int bias, int quirk, int clip_OF,
                      int32_t *expo, int32_t *mant)
{
    orc_param p = orc_param_prepare(E, M, bias, quirk);
    for (int64_t i = 0; i < n; ++i) orc_encode(&p, v[i], clip_OF, &expo[i], &mant[i]);
}

void orc_decode_array(const int32_t *expo, const int32_t *mant, const float *sign, int64_t n,
                      int E, int M, int bias, float *out)
{
    orc_param p = orc_param_prepare(E, M, bias, 0);
    for (int64_t i = 0; i < n; ++i) out[i] = orc_decode(&p, sign ? sign[i] : 1.0f, expo[i], mant[i]);
}

void orc_quant_array(const float *v, int64_t n, int E, int M, int bias, int quirk, int clip_OF, float *out)
{
    orc_param p = orc_param_prepare(E, M, bias, quirk);
    for (int64_t i = 0; i < n; ++i) out[i] = orc_quant(&p, v[i], clip_OF);
}

/* The per-pair addend alone (K=1 view), used by the exhaustive pair tests. */
void orc_pair_array(const orc_cfg *c, const int8_t *T, const float *a, const float *b, int64_t n,
                    int bA, int bB, int bR, float *out)
{
    orc_param pA = orc_param_prepare(c->expo_width, c->mant_width, bA, c->tensor_bias_quirk);
    orc_param pB = orc_param_prepare(c->expo_width, c->mant_width, bB, c->tensor_bias_quirk);
    orc_param pR = orc_param_prepare(c->expo_width, c->mant_width, bR, c->tensor_bias_quirk);
    for (int64_t i = 0; i < n; ++i) out[i] = orc_pair(c, &pA, &pB, &pR, T, a[i], b[i]);
}

/*
 * approx_multiply's column loop over custom_matmul_vectorize:
 *   approx/approx_calculation.py:946-971 (Linear), :774-799 (BN-conv), one column = v9 :10-169.
 * C[m,n] = sum_k pair(A[m,k], B[k,n]) with per-column weight bias bB[n] (bB_stride 0 = shared).
 */
void orc_approx_matmul(const orc_cfg *c, const int8_t *T,
                       const float *A, int64_t sa0, int64_t sa1,
                       const float *B, int64_t sb0, int64_t sb1,
                       float *C, int64_t sc0, int64_t sc1,
                       int64_t Mrows, int64_t K, int64_t N,
                       int bA, const int32_t *bB, int bB_stride, int bR,
                       int64_t n_begin, int64_t n_end)
{
    if (n_end <= 0 || n_end > N) n_end = N;
    orc_param pA = orc_param_prepare(c->expo_width, c->mant_width, bA, c->tensor_bias_quirk);
    orc_param pR = orc_param_prepare(c->expo_width, c->mant_width, bR, c->tensor_bias_quirk);
#pragma omp parallel for collapse(2) schedule(static)
    for (int64_t n = n_begin; n < n_end; ++n) {
        for (int64_t m = 0; m < Mrows; ++m) {
            orc_param pB = orc_param_prepare(c->expo_width, c->mant_width, bB[n * bB_stride],
                                             c->tensor_bias_quirk);
            float accf = 0.0f;
            double accd = 0.0;
            for (int64_t k = 0; k < K; ++k) {
                float r = orc_pair(c, &pA, &pB, &pR, T, A[m * sa0 + k * sa1], B[k * sb0 + n * sb1]);
                accf += r;
                accd += (double)r;
            }
            C[m * sc0 + n * sc1] = c->sum_mode ? (float)accd : accf;
        }
    }
}

/* Sum of |addend| per output, for the tests' exactness / norm-wise tolerance bound. */
void orc_approx_matmul_abs(const orc_cfg *c, const int8_t *T,
                           const float *A, int64_t sa0, int64_t sa1,
                           const float *B, int64_t sb0, int64_t sb1,
                           double *Cabs, int64_t Mrows, int64_t K, int64_t N,
                           int bA, const int32_t *bB, int bB_stride, int bR)
{
    orc_param pA = orc_param_prepare(c->expo_width, c->mant_width, bA, c->tensor_bias_quirk);
    orc_param pR = orc_param_prepare(c->expo_width, c->mant_width, bR, c->tensor_bias_quirk);
#pragma omp parallel for collapse(2) schedule(static)
    for (int64_t n = 0; n < N; ++n)
        for (int64_t m = 0; m < Mrows; ++m) {
            orc_param pB = orc_param_prepare(c->expo_width, c->mant_width, bB[n * bB_stride],
                                             c->tensor_bias_quirk);
            double acc = 0.0;
            for (int64_t k = 0; k < K; ++k)
                acc += fabs((double)orc_pair(c, &pA, &pB, &pR, T, A[m * sa0 + k * sa1],
                                             B[k * sb0 + n * sb1]));
            Cabs[m * N + n] = acc;
        }
}

/* ------------------------------------------------------------------------------------------ */
/* FP8 fake-quantizer: quantize_to_fp8_ste_MM, quantization/quantizers/fp8_quantizer.py:97-173 */
/* ------------------------------------------------------------------------------------------ */

/* bias = round(2**E - log2(maxval) + log2(2 - 2**-M) - 1), fp8_quantizer.py:111-117, fp32 ops */
float orc_fp8_bias(float maxval, int n_bits, float mantissa_bits, int sign_bits)
{
    float M = rintf(mantissa_bits);
    if (M < 1.0f) M = 1.0f;
    if (M > (float)(n_bits - sign_bits)) M = (float)(n_bits - sign_bits);
    float E = (float)(n_bits - sign_bits) - M;
    float t0 = exp2f(E);
    float t1 = t0 - log2f(maxval);
    float t2 = t1 + log2f(2.0f - exp2f(-M));
    float t3 = t2 - 1.0f;
    return rintf(t3);
}

/* elementwise body, fp8_quantizer.py:119-152.  n_channels==1: per-tensor maxval; otherwise
 * element i belongs to channel (i / inner) % n_channels (maxval viewed as [C,1,..,1], :114-115). */
void orc_fp8_fakequant(const float *x, float *y, int64_t n, const float *maxval, int64_t n_channels,
                       int64_t inner, int n_bits, float mantissa_bits, int sign_bits, float *bias_out)
{
    float M = rintf(mantissa_bits);
    if (M < 1.0f) M = 1.0f;
    if (M > (float)(n_bits - sign_bits)) M = (float)(n_bits - sign_bits);
    for (int64_t c = 0; c < n_channels; ++c)
        bias_out[c] = orc_fp8_bias(maxval[c], n_bits, mantissa_bits, sign_bits);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        int64_t c = (n_channels == 1) ? 0 : (i / inner) % n_channels;
        float mx = maxval[c], bias = bias_out[c];
        float mn = (sign_bits == 1) ? -mx : 0.0f;                 /* :119 */
        float xc = clamp_nan(x[i], mn, mx);                        /* :120 */
        float ls = floorf(log2f(fabsf(xc)) + bias);                /* :144 */
        if (!(ls >= 1.0f)) ls = 1.0f;                              /* clamp(min=1); -inf -> 1 */
        float scale = exp2f(ls - M - bias);                        /* :151 */
        y[i] = rintf(xc / scale) * scale;                          /* :154 */
    }
}

/* quantize_after_mult_and_add column loop: approx/approx_calculation.py:959-967 (Linear),
 * :787-795 (BN-conv): q(sum_k q(x[m,k]*y[k,n])) with q = quantize_to_fp8_ste_MM(res maxval). */
void orc_qamaa_matmul(const float *A, int64_t sa0, int64_t sa1,
                      const float *B, int64_t sb0, int64_t sb1,
                      float *C, int64_t sc0, int64_t sc1,
                      int64_t Mrows, int64_t K, int64_t N,
                      float maxval, int n_bits, float mantissa_bits, int sign_bits, int sum_mode)
{
    float M = rintf(mantissa_bits);
    if (M < 1.0f) M = 1.0f;
    if (M > (float)(n_bits - sign_bits)) M = (float)(n_bits - sign_bits);
    float bias = orc_fp8_bias(maxval, n_bits, mantissa_bits, sign_bits);
    float mn = (sign_bits == 1) ? -maxval : 0.0f;
#pragma omp parallel for collapse(2) schedule(static)
    for (int64_t n = 0; n < N; ++n)
        for (int64_t m = 0; m < Mrows; ++m) {
            float accf = 0.0f;
            double accd = 0.0;
            for (int64_t k = 0; k < K; ++k) {
                float g = A[m * sa0 + k * sa1] * B[k * sb0 + n * sb1];
                float xc = clamp_nan(g, mn, maxval);
                float ls = floorf(log2f(fabsf(xc)) + bias);
                if (!(ls >= 1.0f)) ls = 1.0f;
                float scale = exp2f(ls - M - bias);
                float q = rintf(xc / scale) * scale;
                accf += q;
                accd += (double)q;
            }
            float s = sum_mode ? (float)accd : accf;
            float xc = clamp_nan(s, mn, maxval);
            float ls = floorf(log2f(fabsf(xc)) + bias);
            if (!(ls >= 1.0f)) ls = 1.0f;
            float scale = exp2f(ls - M - bias);
            C[m * sc0 + n * sc1] = rintf(xc / scale) * scale;
        }
}

/* ------------------------------------------------------------------------------------------ */
/* all-normal codec, approx/fp_any_int_v5.py:6-53 (params), :55-110 (encode), :116-162 (decode) */
/* ------------------------------------------------------------------------------------------ */
void orc_allnorm_encode_array(const float *v, int64_t n, int E, int M, int bias, int clip_OF,
                              int32_t *expo, int32_t *mant)
{
    int max_expo = (1 << E) - 1, max_mant = (1 << M) - 1;
    float max_value = (float)(ldexp(1.0, max_expo - bias) * (2.0 - ldexp(1.0, -M)));
    float min_value = (float)(ldexp(1.0, -bias) * (1.0 + ldexp(1.0, -M)));
    for (int64_t i = 0; i < n; ++i) {
        int ex;
        float m = frexpf(v[i], &ex);
        if (v[i] == 0.0f) { m = 0.0f; ex = 0; }
        int zero = (v[i] > -min_value) && (v[i] < min_value);       /* :92 */
        float x = zero ? 0.0f : (fabsf(m) * 2.0f - 1.0f) * pow2f(M); /* :95-99 */
        float r = rintf(x);
        if (r > (float)max_mant) r = (float)max_mant;
        int32_t mm = (int32_t)r;
        int32_t ee = zero ? 0 : ex + (bias - 1);                     /* :102 */
        if (clip_OF && (v[i] < -max_value || v[i] > max_value)) { ee = max_expo; mm = max_mant; }
        expo[i] = ee;
        mant[i] = mm;
    }
}

void orc_allnorm_decode_array(const int32_t *expo, const int32_t *mant, const float *sign, int64_t n,
                              int M, int bias, float *out)
{
    for (int64_t i = 0; i < n; ++i) {
        int zero = (expo[i] == 0) && (mant[i] == 0);                 /* :153 */
        float v = zero ? 0.0f
                       : pow2f(expo[i] - bias) * (1.0f + (float)mant[i] / (float)(1 << M)); /* :156 */
        out[i] = sign ? v * sign[i] : v;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* OF/UF clipping on the integer code domain -- only live in v5:                               */
/* approx_mult_new, approx/approx_matmul_whole_v5.py:155-182.          */
/* The with_OF_opt line of v5 names an undefined variable (:173); the intended clamp is used.  */
/* ------------------------------------------------------------------------------------------ */
void orc_v5_ofuf_array(const int32_t *x_int, const int32_t *y_int, const int32_t *comp, int64_t n,
                       int B_neg, int E, int M, int sim_hw_add_OFUF, int with_OF_opt, int with_UF_opt,
                       int32_t *out)
{
    const int32_t mod = 1 << (E + M), max_norm_int = mod - 1, mant_scale = 1 << M;
    for (int64_t i = 0; i < n; ++i) {
        int32_t t = x_int[i] + y_int[i] + B_neg + comp[i];           /* :160 */
        if (sim_hw_add_OFUF) {
            int of = t > max_norm_int, uf = t < 0;                   /* :166-167 */
            t = ((t % mod) + mod) % mod;                             /* floor-mod :170 */
            if (with_OF_opt && of) t = max_norm_int;                 /* :172-173 */
            if (with_UF_opt && uf) t = t % mant_scale;               /* :175-176 */
        }
        out[i] = t;
    }
}

/*
 * The v5 GEMM, custom_matmul_vectorize of approx/approx_matmul_whole_v5.py:10-152 -- the last version whose
 * sim_hw_add_OFUF / with_OF_opt / with_UF_opt are live (approx_v9 ignores them):
 *   (eA,mA) = enc(a), (eB,mB) = enc(b) with clip_OF = True (:22,:26-50; python-number param_prepare :204-243);
 *   t = (eA*2^M + mA) + (eB*2^M + mB) + B_neg + comp[mA,mB]  (:52-85,:159), then the adder simulation (:162-176);
 *   r = dec(t) * sign(a) * sign(b), sign(v) = v < 0 ? -1 : 1  (:106-116, :285-313: expo = t // 2^M floor, mant = t mod 2^M);
 *   C = sum_k r in fp32 (:131).
 * v5 has ONE custom_bias for all three formats; with per-operand biases (the layers have them) the code offset
 * is B_neg = -((bA + bB[n] - bR) << M) -- approx_v9's exponent formula (v9 :66-67) -- and the result decodes with bR;
 * bA == bB == bR reproduces v5 exactly.  A zero operand is NOT special in v5: its code (0,0) takes part in the add.
 */
static float orc_v5_pair(int E, int M, const orc_param *pA, const orc_param *pB, const orc_param *pR,
                         const int32_t *comp, int sim, int of_opt, int uf_opt, float a, float b)
{
    int32_t eA, mA, eB, mB;
    orc_encode(pA, a, 1, &eA, &mA);
    orc_encode(pB, b, 1, &eB, &mB);
    const int32_t mant_scale = 1 << M, mod = 1 << (E + M), max_norm_int = mod - 1;
    const int32_t B_neg = -((pA->fp_bias + pB->fp_bias - pR->fp_bias) * mant_scale);
    int32_t t = (eA * mant_scale + mA) + (eB * mant_scale + mB) + B_neg + (comp ? comp[mA * mant_scale + mB] : 0);
    if (sim) {
        int of = t > max_norm_int, uf = t < 0;
        t = ((t % mod) + mod) % mod;
        if (of_opt && of) t = max_norm_int;
        if (uf_opt && uf) t = t % mant_scale;
    }
    int32_t expo = (t >= 0) ? t / mant_scale : -((-t + mant_scale - 1) / mant_scale);   /* floor division */
    int32_t mant = t - expo * mant_scale;
    const float sign = ((a < 0.0f) ? -1.0f : 1.0f) * ((b < 0.0f) ? -1.0f : 1.0f);
    return orc_decode(pR, sign, expo, mant);
}

void orc_v5_matmul(int E, int M, const int32_t *comp, const float *A, int64_t sa0, int64_t sa1, const float *B,
                   int64_t sb0, int64_t sb1, float *C, int64_t sc0, int64_t sc1, int64_t Mrows, int64_t K, int64_t N,
                   int bA, const int32_t *bB, int bB_stride, int bR, int sim, int of_opt, int uf_opt, int sum_mode)
{
    orc_param pA = orc_param_prepare(E, M, bA, 0);
    orc_param pR = orc_param_prepare(E, M, bR, 0);
#pragma omp parallel for collapse(2) schedule(static)
    for (int64_t n = 0; n < N; ++n) {
        for (int64_t m = 0; m < Mrows; ++m) {
            orc_param pB = orc_param_prepare(E, M, bB[n * bB_stride], 0);
            float accf = 0.0f;
            double accd = 0.0;
            for (int64_t k = 0; k < K; ++k) {
                float r = orc_v5_pair(E, M, &pA, &pB, &pR, comp, sim, of_opt, uf_opt, A[m * sa0 + k * sa1],
                                      B[k * sb0 + n * sb1]);
                accf += r;
                accd += (double)r;
            }
            C[m * sc0 + n * sc1] = sum_mode ? (float)accd : accf;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* im2col, approx/approx_calculation.py:724-747: [B,C,H,W] -> [B*Ho*Wo, C*kh*kw], K order (c,ky,kx) */
/* ------------------------------------------------------------------------------------------ */
void orc_im2col(const float *x, int B, int C, int H, int W, int kh, int kw, int sh, int sw,
                int ph, int pw, int dh, int dw, float *col)
{
    int Ho = (H + 2 * ph - dh * (kh - 1) - 1) / sh + 1;
    int Wo = (W + 2 * pw - dw * (kw - 1) - 1) / sw + 1;
    int64_t Kc = (int64_t)C * kh * kw;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b)
        for (int oy = 0; oy < Ho; ++oy)
            for (int ox = 0; ox < Wo; ++ox) {
                float *row = col + (((int64_t)b * Ho + oy) * Wo + ox) * Kc;
                for (int c = 0; c < C; ++c)
                    for (int ky = 0; ky < kh; ++ky)
                        for (int kx = 0; kx < kw; ++kx) {
                            int iy = oy * sh + ky * dh - ph, ix = ox * sw + kx * dw - pw;
                            float v = 0.0f;
                            if (iy >= 0 && iy < H && ix >= 0 && ix < W)
                                v = x[(((int64_t)b * C + c) * H + iy) * W + ix];
                            row[((int64_t)c * kh + ky) * kw + kx] = v;
                        }
            }
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers: the timed CPU baseline sets its thread count explicitly */
void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
