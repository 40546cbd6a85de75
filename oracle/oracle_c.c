/*
 * oracle_c.c — CPU restatement (plain C, integer/IEEE arithmetic) of the FP8 linear-layer
 * hot path's quantisation rules.  TEST INFRASTRUCTURE ONLY: imported by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference leg.  The product
 * (llm_fp8_b200/) never loads this file.
 *
 * PARITY UNPINNED at the Transformer-Engine boundary: the reference (/root/reference) holds no
 * arithmetic for this path — it calls the third-party `transformer_engine` package
 * (install.sh:4, "2.5.0" per paper/conference_101719.tex:236), which is absent here.  The rules
 * below follow BASELINE.json's north star (saturating RNE casts, amax/scale rules, E8M0 round-up)
 * and SURVEY.md Appendix A/C, anchored on the reference's call sites:
 *   te_llama.py:39-40   DelayedScaling(HYBRID|E4M3, amax_history_len=16, "max")
 *   te_llama_mxfp8.py:28-29   MXFP8BlockScaling(E4M3)
 *   te_llama.py:58-63   LayerNormMLP(RMSNorm, swiglu) — gate rows first (te_llama.py:225-233)
 * and pinned on the only fixtures the reference holds: the three valid FP8 encodings worked in
 * thesis/chapters/c2/c2_quantization.tex:100,102,103 (tests/test_oracle.py).
 *
 * Every floating-point step is written with an explicit rounding (no contraction: build with
 * -ffp-contract=off) so that the CUDA kernels can reproduce it bit for bit with
 * __fmul_rn/__fadd_rn/__fmaf_rn/__fdiv_rn/__fsqrt_rn.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define FMT_E4M3 0
#define FMT_E5M2 1

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline float bf16_to_f32(uint16_t h) { return u2f(((uint32_t)h) << 16); }

/* RNE fp32 -> bf16 (NaN kept quiet) */
static inline uint16_t f32_to_bf16(float f) {
    uint32_t u = f2u(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

/* Saturating round-to-nearest-even fp32 -> fp8 (e4m3fn: max 448, no inf; e5m2: max 57344).
 * Equals torch: v.clamp(-fmax, fmax).to(float8)   [SURVEY.md Appendix C sat_cast] */
uint8_t orc_f32_to_fp8(float f, int fmt) {
    const int mbits = fmt == FMT_E4M3 ? 3 : 2;
    const int bias = fmt == FMT_E4M3 ? 7 : 15;
    const float fmax = fmt == FMT_E4M3 ? 448.0f : 57344.0f;
    const uint8_t maxbyte = fmt == FMT_E4M3 ? 0x7E : 0x7B;
    uint32_t bits = f2u(f);
    uint8_t sign = (uint8_t)((bits >> 31) << 7);
    uint32_t a = bits & 0x7fffffffu;
    if (a > 0x7f800000u) return sign | 0x7F;          /* NaN stays NaN */
    float af = u2f(a);
    if (af >= fmax) return sign | maxbyte;            /* satfinite (covers inf) */
    int e = (int)(a >> 23) - 127;
    int emin = 1 - bias;
    if (e < emin) {                                   /* target subnormal: quantum 2^(emin-mbits) */
        float q = af * u2f((uint32_t)(127 - (emin - mbits)) << 23);   /* exact power-of-two scale */
        int m = (int)nearbyintf(q);                   /* RNE (default rounding mode) */
        return sign | (uint8_t)m;                     /* m==2^mbits encodes the min normal */
    }
    int shift = 23 - mbits;
    uint32_t lsb = (a >> shift) & 1u;
    uint32_t a2 = a + ((1u << (shift - 1)) - 1u) + lsb;
    int E = (int)(a2 >> 23) - 127 + bias;
    uint32_t M = (a2 >> shift) & ((1u << mbits) - 1u);
    return sign | (uint8_t)((E << mbits) | M);
}

float orc_fp8_to_f32(uint8_t b, int fmt) {
    const int mbits = fmt == FMT_E4M3 ? 3 : 2;
    const int ebits = fmt == FMT_E4M3 ? 4 : 5;
    const int bias = fmt == FMT_E4M3 ? 7 : 15;
    int s = b >> 7;
    int E = (b >> mbits) & ((1 << ebits) - 1);
    int M = b & ((1 << mbits) - 1);
    float v;
    if (fmt == FMT_E4M3) {
        if (E == 15 && M == 7) return u2f(0x7fc00000u | ((uint32_t)s << 31));
    } else {
        if (E == 31) return M ? u2f(0x7fc00000u | ((uint32_t)s << 31)) : (s ? -INFINITY : INFINITY);
    }
    if (E == 0) v = ldexpf((float)M, 1 - bias - mbits);
    else v = ldexpf((float)((1 << mbits) | M), E - bias - mbits);
    return s ? -v : v;
}

/* E8M0 of v = amax/448, rounded UP to a power of two, satfinite  [SURVEY.md Appendix C] */
uint8_t orc_e8m0_roundup(float v) {
    uint32_t bits = f2u(v);
    uint32_t a = bits & 0x7fffffffu;
    if (a > 0x7f800000u) return 0xFF;
    if (a == 0) return 0;
    uint32_t exp = a >> 23, man = a & 0x7fffffu;
    if (man > 0 && exp != 0xFE && !(exp == 0 && man <= 0x400000u)) exp += 1;
    return (uint8_t)exp;
}

/* 2^(127-e) as fp32: what a block is multiplied by before the cast */
float orc_e8m0_inv_scale(uint8_t e) {
    if (e <= 253) return u2f((uint32_t)(254 - e) << 23);
    return u2f(e == 254 ? 0x00400000u : 0x00200000u);
}

/* NaN-ignoring |x| max, like fmaxf */
static inline float amax_step(float m, float x) { float a = fabsf(x); return (a > m) ? a : m; }

/* ---- K2: per-tensor cast + amax ------------------------------------------------------------ */
void orc_quantize_bf16(const uint16_t* x, int64_t n, float scale, int fmt, uint8_t* q, float* amax_out) {
    float amax = 0.f;
    #pragma omp parallel for reduction(max:amax) schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float v = bf16_to_f32(x[i]);
        amax = amax_step(amax, v);
        q[i] = orc_f32_to_fp8(v * scale, fmt);
    }
    *amax_out = amax;
}

void orc_cast_f32(const float* x, int64_t n, int fmt, uint8_t* q) {
    for (int64_t i = 0; i < n; i++) q[i] = orc_f32_to_fp8(x[i], fmt);
}
void orc_decode(const uint8_t* q, int64_t n, int fmt, float* out) {
    for (int64_t i = 0; i < n; i++) out[i] = orc_fp8_to_f32(q[i], fmt);
}

/* column sums of a bf16 [R,C] matrix in fp32 (sequential over rows), rounded to bf16: bgrad */
void orc_colsum_bf16(const uint16_t* x, int64_t R, int64_t C, float* out) {
    #pragma omp parallel for schedule(static)
    for (int64_t c = 0; c < C; c++) {
        double acc = 0.0;   /* reference value; kernel is checked to tolerance, not bits */
        for (int64_t r = 0; r < R; r++) acc += (double)bf16_to_f32(x[r * C + c]);
        out[c] = (float)acc;
    }
}

/* ---- K4: delayed-scaling amax-history roll + scale update  [SURVEY.md Appendix A] ---------- */
void orc_amax_scale_update(float* hist /*[L,T]*/, float* scale, float* scale_inv, const float* fmax /*[T]*/,
                           int L, int T, int margin) {
    for (int t = 0; t < T; t++) {
        float am = 0.f;
        int nonfinite = 0;
        for (int l = 0; l < L; l++) {
            float h = hist[l * T + t];
            if (!isfinite(h)) nonfinite = 1;
            am = (h > am) ? h : am;
        }
        float sf = (fmax[t] / am) / ldexpf(1.0f, margin);
        if (!(am > 0.f)) sf = scale[t];
        if (nonfinite) sf = scale[t];
        if (isinf(sf)) sf = 3.402823466e38f;
        scale[t] = sf;
        scale_inv[t] = 1.0f / sf;
    }
    /* roll(-1, dim 0) then zero row 0: new[l] = old[l+1], new[L-1] = old[0], then new[0] = 0 */
    for (int t = 0; t < T; t++) {
        float first = hist[t];
        for (int l = 0; l + 1 < L; l++) hist[l * T + t] = hist[(l + 1) * T + t];
        hist[(L - 1) * T + t] = first;
        hist[t] = 0.f;
    }
}

/* ---- K1: RMSNorm (+amax +cast) with a pinned reduction order -------------------------------
 * Row of H bf16 values, H % 8 == 0.  ORDER (the kernel is written to this):
 *   128 virtual threads; 8-element chunk c belongs to thread c % 128; a thread walks its chunks
 *   in increasing c and within a chunk j = 0..7:  acc = fma(x, x, acc).
 *   Warp butterfly over lanes (xor 16, 8, 4, 2, 1), then ((w0 + w1) + w2) + w3.
 *   mean = total / H ; rs = 1 / sqrtf(mean + eps) ; y = (x * rs) * gamma   (each op rounded). */
#define RMS_T 128
float orc_rms_rsigma(const uint16_t* x, int64_t H, float eps) {
    float acc[RMS_T];
    for (int t = 0; t < RMS_T; t++) acc[t] = 0.f;
    int64_t nchunk = H / 8;
    for (int64_t c = 0; c < nchunk; c++) {
        int t = (int)(c % RMS_T);
        for (int j = 0; j < 8; j++) {
            float v = bf16_to_f32(x[c * 8 + j]);
            acc[t] = fmaf(v, v, acc[t]);
        }
    }
    float w[4];
    for (int wi = 0; wi < 4; wi++) {
        float v[32], n[32];
        for (int l = 0; l < 32; l++) v[l] = acc[wi * 32 + l];
        for (int off = 16; off >= 1; off >>= 1) {
            for (int l = 0; l < 32; l++) n[l] = v[l] + v[l ^ off];
            memcpy(v, n, sizeof v);
        }
        w[wi] = v[0];
    }
    float total = ((w[0] + w[1]) + w[2]) + w[3];
    float mean = total / (float)H;
    return 1.0f / sqrtf(mean + eps);
}

/* y (fp32) for one row */
static void rms_row_y(const uint16_t* x, const uint16_t* gamma, int64_t H, float rs, float* y) {
    for (int64_t i = 0; i < H; i++) {
        float t = bf16_to_f32(x[i]) * rs;
        y[i] = t * bf16_to_f32(gamma[i]);
    }
}

void orc_rmsnorm_quant(const uint16_t* x, const uint16_t* gamma, int64_t M, int64_t H, float eps, float scale,
                       int fmt, uint8_t* q, float* rsigma, float* amax_out, float* y_out /*nullable [M,H]*/) {
    float amax = 0.f;
    #pragma omp parallel for reduction(max:amax) schedule(static)
    for (int64_t m = 0; m < M; m++) {
        float* y = (float*)malloc(sizeof(float) * H);
        float rs = orc_rms_rsigma(x + m * H, H, eps);
        rsigma[m] = rs;
        rms_row_y(x + m * H, gamma, H, rs, y);
        for (int64_t i = 0; i < H; i++) {
            amax = amax_step(amax, y[i]);
            q[m * H + i] = orc_f32_to_fp8(y[i] * scale, fmt);
        }
        if (y_out) memcpy(y_out + m * H, y, sizeof(float) * H);
        free(y);
    }
    *amax_out = amax;
}

/* ---- K3: SwiGLU with a pinned exp ----------------------------------------------------------- */
/* sigmoid(g) = 1 / (1 + E),  E = exp(-g) evaluated as 2^(x * log2e), x = clamp(-g, -80, 80):
 *   m = fma(x, log2e, 1.5 * 2^23)        -> the low mantissa bits of m hold n = rint(x * log2e) (ties to even)
 *   f = fma(x, log2e, 1.5 * 2^23 - m)    -> x * log2e - n, |f| <= 0.5, one rounding
 *   p = 2^f by a degree-5 Horner chain of fmas (max rel. error 1.7e-7)
 *   E = bits(p) + (bits(m) << 23)        -> 2^n * p by integer exponent arithmetic
 * Every step is one IEEE fp32 operation; csrc/common.cuh: sigmoid2_pinned performs the same steps (two lanes at a time). */
#define SIG_CLAMP 80.0f
#define SIG_LOG2E 1.44269502162933349609375f
#define SIG_MAGIC 12582912.0f
float orc_exp_neg(float g) {
    float x = fminf(fmaxf(-g, -SIG_CLAMP), SIG_CLAMP);
    float m = fmaf(x, SIG_LOG2E, SIG_MAGIC);
    float nneg = SIG_MAGIC - m;
    float f = fmaf(x, SIG_LOG2E, nneg);
    float p = 0.001326472731307149f;
    p = fmaf(p, f, 0.009671512991189957f);
    p = fmaf(p, f, 0.05550733581185341f);
    p = fmaf(p, f, 0.24022242426872253f);
    p = fmaf(p, f, 0.6931470036506653f);
    p = fmaf(p, f, 1.0f);
    return u2f(f2u(p) + (f2u(m) << 23));
}
float orc_sigmoid(float g) { return 1.0f / (1.0f + orc_exp_neg(g)); }

/* fc1_out [M, 2I] bf16, gate = [:, :I], up = [:, I:]  (te_llama.py:225-233: gate rows first) */
void orc_swiglu_f32(const uint16_t* fc1, int64_t M, int64_t I, float* a) {
    #pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; m++)
        for (int64_t i = 0; i < I; i++) {
            float g = bf16_to_f32(fc1[m * 2 * I + i]);
            float u = bf16_to_f32(fc1[m * 2 * I + I + i]);
            float s = g * orc_sigmoid(g);
            a[m * I + i] = s * u;
        }
}

void orc_swiglu_quant(const uint16_t* fc1, int64_t M, int64_t I, float scale, int fmt, uint8_t* q, float* amax_out) {
    float amax = 0.f;
    #pragma omp parallel for reduction(max:amax) schedule(static)
    for (int64_t m = 0; m < M; m++)
        for (int64_t i = 0; i < I; i++) {
            float g = bf16_to_f32(fc1[m * 2 * I + i]);
            float u = bf16_to_f32(fc1[m * 2 * I + I + i]);
            float s = g * orc_sigmoid(g);
            float a = s * u;
            amax = amax_step(amax, a);
            q[m * I + i] = orc_f32_to_fp8(a * scale, fmt);
        }
    *amax_out = amax;
}

/* dSwiGLU: da [M,I] bf16, fc1 [M,2I] bf16 -> d_fc1 [M,2I] fp32 (dg | du)
 *   sig = sigmoid(g) ; dsilu = sig * (1 + g * (1 - sig)) ; dg = (da * u) * dsilu ; du = da * (g * sig) */
void orc_dswiglu_f32(const uint16_t* da, const uint16_t* fc1, int64_t M, int64_t I, float* d) {
    #pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; m++)
        for (int64_t i = 0; i < I; i++) {
            float g = bf16_to_f32(fc1[m * 2 * I + i]);
            float u = bf16_to_f32(fc1[m * 2 * I + I + i]);
            float dy = bf16_to_f32(da[m * I + i]);
            float sig = orc_sigmoid(g);
            float one_m = 1.0f - sig;
            float t = fmaf(g, one_m, 1.0f);
            float dsilu = sig * t;
            float dg = (dy * u) * dsilu;
            float du = dy * (g * sig);
            d[m * 2 * I + i] = dg;
            d[m * 2 * I + I + i] = du;
        }
}

/* ---- K2m: MXFP8 1x32 block quantiser + 128x4 scale swizzle ----------------------------------
 * swizzled offset of scale (r, kb) for a [R, KB] scale matrix (R padded to 128, KB padded to 4):
 *   tile = (r/128) * (KBp/4) + kb/4 ; off = tile*512 + (r%32)*16 + ((r%128)/32)*4 + kb%4      */
int64_t orc_sf_offset(int64_t r, int64_t kb, int64_t KBp) {
    int64_t tile = (r / 128) * (KBp / 4) + kb / 4;
    return tile * 512 + (r % 32) * 16 + ((r % 128) / 32) * 4 + (kb % 4);
}

/* x fp32 [R,C].  rowwise: blocks of 32 along C -> q_row [R,C], scales (r, c/32).
 *                colwise: blocks of 32 along R -> q_col [R,C] (same layout), scales (c, r/32).
 * sf buffers must be pre-zeroed by the caller and sized for the padded swizzled layout. */
void orc_mx_quantize_f32(const float* x, int64_t R, int64_t C, int rowwise, int colwise,
                         uint8_t* q_row, uint8_t* sf_row, uint8_t* q_col, uint8_t* sf_col) {
    if (rowwise) {
        int64_t KB = C / 32, KBp = (KB + 3) / 4 * 4;
        #pragma omp parallel for schedule(static)
        for (int64_t r = 0; r < R; r++)
            for (int64_t kb = 0; kb < KB; kb++) {
                float am = 0.f;
                for (int j = 0; j < 32; j++) am = amax_step(am, x[r * C + kb * 32 + j]);
                uint8_t e = orc_e8m0_roundup(am / 448.0f);
                float inv = orc_e8m0_inv_scale(e);
                sf_row[orc_sf_offset(r, kb, KBp)] = e;
                for (int j = 0; j < 32; j++)
                    q_row[r * C + kb * 32 + j] = orc_f32_to_fp8(x[r * C + kb * 32 + j] * inv, FMT_E4M3);
            }
    }
    if (colwise) {
        int64_t KB = R / 32, KBp = (KB + 3) / 4 * 4;
        #pragma omp parallel for schedule(static)
        for (int64_t c = 0; c < C; c++)
            for (int64_t kb = 0; kb < KB; kb++) {
                float am = 0.f;
                for (int j = 0; j < 32; j++) am = amax_step(am, x[(kb * 32 + j) * C + c]);
                uint8_t e = orc_e8m0_roundup(am / 448.0f);
                float inv = orc_e8m0_inv_scale(e);
                sf_col[orc_sf_offset(c, kb, KBp)] = e;
                for (int j = 0; j < 32; j++)
                    q_col[(kb * 32 + j) * C + c] = orc_f32_to_fp8(x[(kb * 32 + j) * C + c] * inv, FMT_E4M3);
            }
    }
}

void orc_bf16_to_f32(const uint16_t* x, int64_t n, float* out) {
    for (int64_t i = 0; i < n; i++) out[i] = bf16_to_f32(x[i]);
}
void orc_f32_to_bf16(const float* x, int64_t n, uint16_t* out) {
    for (int64_t i = 0; i < n; i++) out[i] = f32_to_bf16(x[i]);
}

/* RMSNorm y in fp32 for the MX path (K1m = orc_rmsnorm_y then orc_mx_quantize_f32) */
void orc_rmsnorm_y(const uint16_t* x, const uint16_t* gamma, int64_t M, int64_t H, float eps, float* y, float* rsigma) {
    #pragma omp parallel for schedule(static)
    for (int64_t m = 0; m < M; m++) {
        float rs = orc_rms_rsigma(x + m * H, H, eps);
        rsigma[m] = rs;
        rms_row_y(x + m * H, gamma, H, rs, y + m * H);
    }
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
