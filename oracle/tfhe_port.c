/*
 * tfhe_port.c -- TEST INFRASTRUCTURE ONLY. See tfhe_port.h for scope, citations and pinning.
 * "T/" = /root/reference/libs/tfhe/src/.
 */
#define _GNU_SOURCE
#include "tfhe_port.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define PN 1024 /* every reference FFT processor asserts N == 1024 (T/libtfhe/fft_processors/spqlios/lagrangehalfc_impl.cpp:8) */
#define PM 512

struct port_key {
    port_params p;
    int32_t kpl, base, n1;
    int32_t *bk;    /* [n][kpl][k+1][N] */
    double *bkfft;  /* [n][kpl][k+1][N] (re[M] | im[M]) -- init_LweBootstrappingKeyFFT, T/libtfhe/lwe-bootstrapping-functions-fft.cpp:23-52 */
    int32_t *ks;    /* [N*k][t][base][n+1] */
};

/* ---------------------------------------------------------------------------------------------
 * scalars
 * --------------------------------------------------------------------------------------------- */
int32_t port_modswitch_from_torus32(int32_t phase, int32_t Msize) {
    /* T/libtfhe/numeric-functions.cpp:59-65 */
    uint64_t interv = ((UINT64_C(1) << 63) / (uint64_t)Msize) * 2;
    uint64_t half_interval = interv / 2;
    uint64_t phase64 = ((uint64_t)(int64_t)phase << 32) + half_interval;
    return (int32_t)(phase64 / interv);
}
int32_t port_modswitch_to_torus32(int32_t mu, int32_t Msize) {
    /* T/libtfhe/numeric-functions.cpp:71-76 */
    uint64_t interv = ((UINT64_C(1) << 63) / (uint64_t)Msize) * 2;
    uint64_t phase64 = (uint64_t)(int64_t)mu * interv;
    return (int32_t)(phase64 >> 32);
}
uint32_t port_decomp_offset(const port_params *p) {
    /* T/libtfhe/tgsw.cpp:7-29: offset = Bg/2 * sum_p 2^(32 - (p+1)*Bgbit) */
    uint32_t half = 1u << (p->Bgbit - 1), off = 0;
    for (int i = 1; i <= p->l; ++i) off += half << (32 - i * p->Bgbit);
    return off;
}

/* ---------------------------------------------------------------------------------------------
 * polynomials
 * --------------------------------------------------------------------------------------------- */
void port_mul_by_xai(int32_t *out, int32_t a, const int32_t *in, int32_t N) {
    /* T/libtfhe/toruspolynomial-functions.cpp:140-160 */
    if (a < N) {
        for (int i = 0; i < a; ++i) out[i] = (int32_t)(0u - (uint32_t)in[i - a + N]);
        for (int i = a; i < N; ++i) out[i] = in[i - a];
    } else {
        const int aa = a - N;
        for (int i = 0; i < aa; ++i) out[i] = in[i - aa + N];
        for (int i = aa; i < N; ++i) out[i] = (int32_t)(0u - (uint32_t)in[i - aa]);
    }
}
void port_mul_by_xai_minus_one(int32_t *out, int32_t a, const int32_t *in, int32_t N) {
    /* T/libtfhe/toruspolynomial-functions.cpp:117-136 */
    if (a < N) {
        for (int i = 0; i < a; ++i) out[i] = (int32_t)(0u - (uint32_t)in[i - a + N] - (uint32_t)in[i]);
        for (int i = a; i < N; ++i) out[i] = (int32_t)((uint32_t)in[i - a] - (uint32_t)in[i]);
    } else {
        const int aa = a - N;
        for (int i = 0; i < aa; ++i) out[i] = (int32_t)((uint32_t)in[i - aa + N] - (uint32_t)in[i]);
        for (int i = aa; i < N; ++i) out[i] = (int32_t)(0u - (uint32_t)in[i - aa] - (uint32_t)in[i]);
    }
}
void port_decompose(int32_t *digits, const int32_t *poly, const port_params *p) {
    /* T/libtfhe/tgsw-functions.cpp:298-348 */
    const uint32_t offset = port_decomp_offset(p), mask = (1u << p->Bgbit) - 1;
    const int32_t half = 1 << (p->Bgbit - 1);
    for (int j = 0; j < p->N; ++j) {
        const uint32_t v = (uint32_t)poly[j] + offset;
        for (int q = 0; q < p->l; ++q)
            digits[q * p->N + j] = (int32_t)((v >> (32 - (q + 1) * p->Bgbit)) & mask) - half;
    }
}

/* ---------------------------------------------------------------------------------------------
 * negacyclic transform, N = 1024.  A real polynomial a(X) mod X^N+1 is folded to
 * z_j = a_j + i a_{j+N/2} and evaluated at the N/2 roots with zeta^(N/2) = i, i.e. the "Lagrange
 * half-complex" domain (T/libtfhe/fft_processors/nayuki/lagrangehalfc_impl.h:36-41): twist by
 * e^{i pi j/N}, then an N/2-point DFT (the folding used by spqlios, ifft_model
 * T/libtfhe/fft_processors/spqlios/spqlios-fft-impl.cpp:469-641).  Ordering and scaling of the domain
 * are private to each reference processor (T/include/polynomials.h:36-47); only int32 results are
 * observable.  Conversions follow fft_processor_spqlios.cpp:38-115: inputs enter as raw
 * double(int32), the way back pre-scales by 2/N and converts with Torus32(int64_t(x)).
 * --------------------------------------------------------------------------------------------- */
static double g_twr[PM], g_twi[PM];   /* e^{i pi j / N} */
static double g_wr[PM / 2], g_wi[PM / 2]; /* e^{2 pi i j / M} */
static pthread_once_t g_once = PTHREAD_ONCE_INIT;
static void init_tables(void) {
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int j = 0; j < PM; ++j) {
        g_twr[j] = (double)cosl(pi * j / PN);
        g_twi[j] = (double)sinl(pi * j / PN);
    }
    for (int j = 0; j < PM / 2; ++j) {
        g_wr[j] = (double)cosl(2 * pi * j / PM);
        g_wi[j] = (double)sinl(2 * pi * j / PM);
    }
}
/* in-place radix-2 DFT of M points, exponent sign s (+1 / -1), natural order in and out */
static void dft512(double *re, double *im, int s) {
    for (int i = 1, j = 0; i < PM; ++i) {
        int bit = PM >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) {
            double t = re[i]; re[i] = re[j]; re[j] = t;
            t = im[i]; im[i] = im[j]; im[j] = t;
        }
    }
    for (int len = 2; len <= PM; len <<= 1) {
        const int step = PM / len;
        for (int i = 0; i < PM; i += len)
            for (int j = 0; j < len / 2; ++j) {
                const double wr = g_wr[j * step], wi = s * g_wi[j * step];
                const int u = i + j, v = i + j + len / 2;
                const double xr = re[v] * wr - im[v] * wi, xi = re[v] * wi + im[v] * wr;
                re[v] = re[u] - xr; im[v] = im[u] - xi;
                re[u] += xr; im[u] += xi;
            }
    }
}
void port_ifft_int(double *out, const int32_t *in) {
    /* IntPolynomial_ifft / TorusPolynomial_ifft -> execute_reverse_int (fft_processor_spqlios.cpp:38-86) */
    pthread_once(&g_once, init_tables);
    double *re = out, *im = out + PM;
    for (int j = 0; j < PM; ++j) {
        const double a = (double)in[j], b = (double)in[j + PM];
        re[j] = a * g_twr[j] - b * g_twi[j];
        im[j] = a * g_twi[j] + b * g_twr[j];
    }
    dft512(re, im, +1);
}
void port_fft_torus(int32_t *out, const double *in) {
    /* TorusPolynomial_fft -> execute_direct_torus32 (fft_processor_spqlios.cpp:88-115): scale 2/N,
     * transform, Torus32(int64_t(x)) i.e. truncation toward zero then wrap to 32 bits */
    pthread_once(&g_once, init_tables);
    double re[PM], im[PM];
    const double s = 2.0 / PN;
    for (int j = 0; j < PM; ++j) { re[j] = in[j] * s; im[j] = in[j + PM] * s; }
    dft512(re, im, -1);
    for (int j = 0; j < PM; ++j) {
        const double a = re[j] * g_twr[j] + im[j] * g_twi[j];
        const double b = im[j] * g_twr[j] - re[j] * g_twi[j];
        out[j] = (int32_t)(uint32_t)(uint64_t)(int64_t)a;
        out[j + PM] = (int32_t)(uint32_t)(uint64_t)(int64_t)b;
    }
}

/* ---------------------------------------------------------------------------------------------
 * key object
 * --------------------------------------------------------------------------------------------- */
port_key *port_key_new(const port_params *p, const int32_t *bk, const int32_t *ks) {
    if (!p || p->N != PN || p->k != 1 || !bk || !ks) return NULL;
    port_key *key = (port_key *)calloc(1, sizeof(port_key));
    key->p = *p;
    key->kpl = (p->k + 1) * p->l;
    key->base = 1 << p->ks_basebit;
    key->n1 = p->n + 1;
    const size_t bk_ints = (size_t)p->n * key->kpl * (p->k + 1) * PN;
    const size_t ks_ints = (size_t)PN * p->k * p->ks_t * key->base * key->n1;
    key->bk = (int32_t *)malloc(bk_ints * 4);
    key->ks = (int32_t *)malloc(ks_ints * 4);
    key->bkfft = (double *)malloc(bk_ints * 8);
    memcpy(key->bk, bk, bk_ints * 4);
    memcpy(key->ks, ks, ks_ints * 4);
    /* tGswToFFTConvert (T/libtfhe/tgsw-fft-operations.cpp:49-54) for every polynomial of the key */
    for (size_t poly = 0; poly < bk_ints / PN; ++poly) port_ifft_int(key->bkfft + poly * PN, key->bk + poly * PN);
    return key;
}
void port_key_free(port_key *key) {
    if (!key) return;
    free(key->bk); free(key->ks); free(key->bkfft); free(key);
}

/* ---------------------------------------------------------------------------------------------
 * external product and CMux
 * --------------------------------------------------------------------------------------------- */
void port_extern_product_fft(const port_key *key, int32_t *acc, int32_t bk_index) {
    /* tGswFFTExternMulToTLwe, T/libtfhe/tgsw-fft-operations.cpp:89-114:
     * decompose every accumulator polynomial (:100-101), transform the kpl digit polynomials
     * (:102-103), tmpa = sum_p decaFFT[p] * bk.row[p] (tLweFFTAddMulRTo, :105-108 ->
     * T/libtfhe/tlwe-fft-operations.cpp:92-100), back to Torus32 (:109) */
    const port_params *p = &key->p;
    const int kpl = key->kpl, k1 = p->k + 1;
    int32_t digits[4 * PN]; /* l <= 4 */
    double dec[PN], tmpa[2][PN];
    memset(tmpa, 0, sizeof(tmpa));
    const double *bki = key->bkfft + (size_t)bk_index * kpl * k1 * PN;
    for (int q = 0; q < k1; ++q) {
        port_decompose(digits, acc + q * PN, p);
        for (int d = 0; d < p->l; ++d) {
            port_ifft_int(dec, digits + d * PN);
            const int row = q * p->l + d; /* row order q*l + p, tgsw-fft-operations.cpp:100-101 */
            for (int c = 0; c < k1; ++c) {
                /* LagrangeHalfCPolynomialAddMul (T/libtfhe/fft_processors/nayuki/lagrangehalfc_impl.cpp:91-103) */
                const double *b = bki + ((size_t)row * k1 + c) * PN;
                for (int f = 0; f < PM; ++f) {
                    tmpa[c][f] += dec[f] * b[f] - dec[f + PM] * b[f + PM];
                    tmpa[c][f + PM] += dec[f] * b[f + PM] + dec[f + PM] * b[f];
                }
            }
        }
    }
    for (int c = 0; c < k1; ++c) port_fft_torus(acc + c * PN, tmpa[c]);
}

void port_extern_product_exact(const port_key *key, int32_t *acc, int32_t bk_index) {
    /* tGswExternMulToTLwe, T/libtfhe/tgsw-functions.cpp:156-170: same sum with exact integer
     * negacyclic products (the reference uses Karatsuba, T/libtfhe/multiplication.cpp:126; schoolbook
     * gives the identical wrap-around result) */
    const port_params *p = &key->p;
    const int kpl = key->kpl, k1 = p->k + 1;
    int32_t digits[4 * PN];
    uint32_t res[2][PN];
    memset(res, 0, sizeof(res));
    const int32_t *bki = key->bk + (size_t)bk_index * kpl * k1 * PN;
    for (int q = 0; q < k1; ++q) {
        port_decompose(digits, acc + q * PN, p);
        for (int d = 0; d < p->l; ++d) {
            const int row = q * p->l + d;
            const int32_t *dg = digits + d * PN;
            for (int c = 0; c < k1; ++c) {
                const int32_t *b = bki + ((size_t)row * k1 + c) * PN;
                for (int i = 0; i < PN; ++i) {
                    const uint32_t di = (uint32_t)dg[i];
                    if (!di) continue;
                    for (int j = 0; j < PN - i; ++j) res[c][i + j] += di * (uint32_t)b[j];
                    for (int j = PN - i; j < PN; ++j) res[c][i + j - PN] -= di * (uint32_t)b[j];
                }
            }
        }
    }
    memcpy(acc, res, sizeof(res));
}

void port_cmux_step(const port_key *key, int32_t *acc, int32_t bk_index, int32_t barai, int exact) {
    /* tfhe_MuxRotate_FFT, T/libtfhe/lwe-bootstrapping-functions-fft.cpp:67-76:
     * result = (X^barai - 1)*accum (:71), result = bki (x) result (:73), result += accum (:75).
     * barai == 0 is skipped by the caller (:101) */
    if (barai == 0) return;
    int32_t tmp[2 * PN];
    for (int q = 0; q < 2; ++q) port_mul_by_xai_minus_one(tmp + q * PN, barai, acc + q * PN, PN);
    if (exact) port_extern_product_exact(key, tmp, bk_index);
    else port_extern_product_fft(key, tmp, bk_index);
    for (int j = 0; j < 2 * PN; ++j) acc[j] = (int32_t)((uint32_t)acc[j] + (uint32_t)tmp[j]);
}

/* ---------------------------------------------------------------------------------------------
 * bootstrap without key switch, key switch, gates
 * --------------------------------------------------------------------------------------------- */
void port_bootstrap_woks(const port_key *key, int32_t *u, const int32_t *x, int32_t mu) {
    const port_params *p = &key->p;
    const int n = p->n, N2 = 2 * PN;
    /* tfhe_bootstrap_woKS_FFT, T/libtfhe/lwe-bootstrapping-functions-fft.cpp:169-200 */
    const int32_t barb = port_modswitch_from_torus32(x[n], N2);                  /* :186 */
    int32_t testvect[PN], acc[2 * PN];
    for (int j = 0; j < PN; ++j) testvect[j] = mu;                              /* :192 */
    /* tfhe_blindRotateAndExtract_FFT :127-156 */
    memset(acc, 0, PN * sizeof(int32_t));                                       /* tLweNoiselessTrivial, T/libtfhe/tlwe-functions.cpp:130-136 */
    if (barb != 0) port_mul_by_xai(acc + PN, N2 - barb, testvect, PN);          /* :146 */
    else memcpy(acc + PN, testvect, sizeof(testvect));
    for (int i = 0; i < n; ++i) {                                               /* tfhe_blindRotate_FFT :88-112 */
        const int32_t barai = port_modswitch_from_torus32(x[i], N2);            /* :187-189 */
        if (barai == 0) continue;                                               /* :101 */
        port_cmux_step(key, acc, i, barai, 0);
    }
    /* tLweExtractLweSampleIndex(index 0), T/libtfhe/lwe.cpp:41-60 */
    u[0] = acc[0];
    for (int j = 1; j < PN; ++j) u[j] = (int32_t)(0u - (uint32_t)acc[PN - j]);
    u[PN] = acc[PN];
}

int32_t port_keyswitch(const port_key *key, int32_t *out, const int32_t *u) {
    /* lweKeySwitch, T/libtfhe/lwe-keyswitch-functions.cpp:228-238 and
     * lweKeySwitchTranslate_fromArray :114-129 */
    const port_params *p = &key->p;
    const int n = p->n, n1 = key->n1, t = p->ks_t, basebit = p->ks_basebit, base = key->base;
    const uint32_t prec_offset = 1u << (32 - (1 + basebit * t));                /* :119 */
    const uint32_t mask = (uint32_t)base - 1;
    int32_t nonzero = 0;
    for (int c = 0; c < n; ++c) out[c] = 0;                                     /* lweNoiselessTrivial :234 */
    out[n] = u[PN];
    for (int i = 0; i < PN; ++i) {
        const uint32_t aibar = (uint32_t)u[i] + prec_offset;                    /* :123 */
        for (int j = 0; j < t; ++j) {
            const uint32_t aij = (aibar >> (32 - (j + 1) * basebit)) & mask;    /* :125 */
            if (aij != 0) {                                                     /* :126 lweSubTo */
                const int32_t *row = key->ks + (((size_t)i * t + j) * base + aij) * n1;
                for (int c = 0; c < n1; ++c) out[c] = (int32_t)((uint32_t)out[c] - (uint32_t)row[c]);
                ++nonzero;
            }
        }
    }
    return nonzero;
}

void port_gate_prologue(int op, int half, int32_t *x, const int32_t *a, const int32_t *b, const int32_t *c, int32_t n) {
    /* T/libtfhe/boot-gates.cpp: NAND :42-45, OR :68-71, AND :94-97, XOR :120-123, XNOR :146-149,
     * NOR :205-208, ANDNY :231-234, ANDYN :257-260, ORNY :283-286, ORYN :309-312, MUX :343-355 */
    const int32_t e8 = port_modswitch_to_torus32(1, 8), e4 = port_modswitch_to_torus32(1, 4);
    int32_t s0 = 0, s1 = 0, cst = 0;
    const int32_t *second = b;
    switch (op) {
        case PORT_NAND:  cst = e8;  s0 = -1; s1 = -1; break;
        case PORT_OR:    cst = e8;  s0 = 1;  s1 = 1;  break;
        case PORT_AND:   cst = -e8; s0 = 1;  s1 = 1;  break;
        case PORT_XOR:   cst = e4;  s0 = 2;  s1 = 2;  break;
        case PORT_XNOR:  cst = -e4; s0 = -2; s1 = -2; break;
        case PORT_NOR:   cst = -e8; s0 = -1; s1 = -1; break;
        case PORT_ANDNY: cst = -e8; s0 = -1; s1 = 1;  break;
        case PORT_ANDYN: cst = -e8; s0 = 1;  s1 = -1; break;
        case PORT_ORNY:  cst = e8;  s0 = -1; s1 = 1;  break;
        case PORT_ORYN:  cst = e8;  s0 = 1;  s1 = -1; break;
        case PORT_MUX:
            cst = -e8; s1 = 1;
            if (half == 0) s0 = 1; else { s0 = -1; second = c; }
            break;
        default: break;
    }
    for (int i = 0; i <= n; ++i)
        x[i] = (int32_t)((uint32_t)s0 * (uint32_t)a[i] + (uint32_t)s1 * (uint32_t)second[i]);
    x[n] = (int32_t)((uint32_t)x[n] + (uint32_t)cst);
}

void port_gate(const port_key *key, int op, int32_t *out, const int32_t *a, const int32_t *b, const int32_t *c) {
    const int n = key->p.n;
    const int32_t mu = port_modswitch_to_torus32(1, 8);                          /* boot-gates.cpp:36 */
    int32_t x[1024], u[PN + 1], u2[PN + 1];
    switch (op) {
        case PORT_NOT:    for (int i = 0; i <= n; ++i) out[i] = (int32_t)(0u - (uint32_t)a[i]); return; /* :164-171 */
        case PORT_COPY:   memmove(out, a, (n + 1) * sizeof(int32_t)); return;                           /* :175-181 */
        case PORT_CONST0:
        case PORT_CONST1: memset(out, 0, n * sizeof(int32_t)); out[n] = op == PORT_CONST1 ? mu : -mu; return; /* :185-189 */
        default: break;
    }
    port_gate_prologue(op, 0, x, a, b, c, n);
    port_bootstrap_woks(key, u, x, mu);
    if (op == PORT_MUX) {
        /* boot-gates.cpp:351-363: second rotation, sum + 1/8 in the extracted dimension, one key switch */
        port_gate_prologue(op, 1, x, a, b, c, n);
        port_bootstrap_woks(key, u2, x, mu);
        for (int j = 0; j <= PN; ++j) u[j] = (int32_t)((uint32_t)u[j] + (uint32_t)u2[j]);
        u[PN] = (int32_t)((uint32_t)u[PN] + (uint32_t)mu);
    }
    port_keyswitch(key, out, u);
}

typedef struct {
    const port_key *key; int op; int32_t *out; const int32_t *a, *b, *c; int64_t begin, end;
} batch_job;
static void *batch_worker(void *arg) {
    batch_job *j = (batch_job *)arg;
    const int n1 = j->key->n1;
    for (int64_t g = j->begin; g < j->end; ++g)
        port_gate(j->key, j->op, j->out + g * n1, j->a ? j->a + g * n1 : NULL, j->b ? j->b + g * n1 : NULL,
                  j->c ? j->c + g * n1 : NULL);
    return NULL;
}
void port_gate_batch(const port_key *key, int op, int32_t *out, const int32_t *a, const int32_t *b,
                     const int32_t *c, int64_t count, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    batch_job jobs[256];
    for (int t = 0; t < nthreads; ++t) {
        jobs[t] = (batch_job){key, op, out, a, b, c, count * t / nthreads, count * (t + 1) / nthreads};
        pthread_create(&th[t], NULL, batch_worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], NULL);
}
