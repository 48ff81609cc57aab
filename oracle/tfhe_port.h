/*
 * tfhe_port.h -- TEST INFRASTRUCTURE ONLY (never linked into or called from the product path).
 *
 * Plain-C CPU restatement of the libtfhe gate-bootstrapping path that eoc-tfhe wraps
 * (reference: /root/reference/libs/tfhe/src, tfhe/tfhe v1.1).  Every function cites the reference
 * file:line it follows ("T/" = libs/tfhe/src/).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library, and only as the checker.
 *
 * Parity pinning: the reference holds no golden vectors for this path; this
 * port is pinned against the UNMODIFIED reference compiled by oracle/Makefile into oracle/_ref/
 * (tests/test_oracle_vs_reference.py) and against fixtures that the reference generated
 * (tests/golden/, made by tests/golden/make_golden.py).
 *
 * Integer parts are exact (wrap-around int32).  The FFT-domain external product uses this port's
 * own double-precision negacyclic transform and the reference's spqlios conversion rules
 * (unscaled int input, 2/N pre-scale, int64 truncation); like any two reference FFT processors it
 * may differ from another implementation by an LSB per coefficient (SURVEY.md finding 4).  An exact
 * integer external product (schoolbook) is provided as the anchor.
 */
#ifndef TFHE_PORT_H
#define TFHE_PORT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t n, N, k, l, Bgbit, ks_t, ks_basebit;
} port_params;

/* key material in wire-format order (T/libtfhe/tfhe_io.cpp:757-930):
 *   bk [n][(k+1)*l][k+1][N] int32, ks [N*k][t][base][n+1] int32 */
typedef struct port_key port_key;
port_key *port_key_new(const port_params *p, const int32_t *bk, const int32_t *ks);
void port_key_free(port_key *key);

enum { PORT_NAND = 0, PORT_OR, PORT_AND, PORT_XOR, PORT_XNOR, PORT_NOR, PORT_ANDNY, PORT_ANDYN,
       PORT_ORNY, PORT_ORYN, PORT_MUX, PORT_NOT, PORT_COPY, PORT_CONST0, PORT_CONST1 };

/* scalars */
int32_t  port_modswitch_from_torus32(int32_t phase, int32_t Msize);   /* T/libtfhe/numeric-functions.cpp:59-65 */
int32_t  port_modswitch_to_torus32(int32_t mu, int32_t Msize);        /* T/libtfhe/numeric-functions.cpp:71-76 */
uint32_t port_decomp_offset(const port_params *p);                   /* T/libtfhe/tgsw.cpp:7-29 */

/* polynomials mod X^N+1, N int32 each */
void port_mul_by_xai(int32_t *out, int32_t a, const int32_t *in, int32_t N);            /* T/libtfhe/toruspolynomial-functions.cpp:140-160 */
void port_mul_by_xai_minus_one(int32_t *out, int32_t a, const int32_t *in, int32_t N);  /* T/libtfhe/toruspolynomial-functions.cpp:117-136 */
void port_decompose(int32_t *digits /*[l][N]*/, const int32_t *poly, const port_params *p); /* T/libtfhe/tgsw-functions.cpp:298-348 */

/* negacyclic transform pair, N = 1024: int32[N] -> double[N] (re[N/2] | im[N/2]) and back
 * (T/libtfhe/fft_processors/spqlios/fft_processor_spqlios.cpp:38-115) */
void port_ifft_int(double *out, const int32_t *in);
void port_fft_torus(int32_t *out, const double *in);

/* acc[k+1][N] <- BK_i (x) acc : FFT domain (T/libtfhe/tgsw-fft-operations.cpp:89-114) and exact
 * integer arithmetic (T/libtfhe/tgsw-functions.cpp:156-170) */
void port_extern_product_fft(const port_key *key, int32_t *acc, int32_t bk_index);
void port_extern_product_exact(const port_key *key, int32_t *acc, int32_t bk_index);
/* acc <- acc + BK_i (x) ((X^barai - 1) acc)   (T/libtfhe/lwe-bootstrapping-functions-fft.cpp:67-76) */
void port_cmux_step(const port_key *key, int32_t *acc, int32_t bk_index, int32_t barai, int exact);

/* u[N+1] = tfhe_bootstrap_woKS_FFT(x[n+1], mu)   (T/libtfhe/lwe-bootstrapping-functions-fft.cpp:88-200, lwe.cpp:41-60) */
void port_bootstrap_woks(const port_key *key, int32_t *u, const int32_t *x, int32_t mu);
/* out[n+1] = lweKeySwitch(u[N+1]); returns the number of non-zero digits
 * (T/libtfhe/lwe-keyswitch-functions.cpp:114-129,228-238) */
int32_t port_keyswitch(const port_key *key, int32_t *out, const int32_t *u);
/* gate prologue x = (0,cst) + s0*a + s1*b (T/libtfhe/boot-gates.cpp:35-319); half selects the MUX half */
void port_gate_prologue(int op, int half, int32_t *x, const int32_t *a, const int32_t *b, const int32_t *c, int32_t n);
/* out[n+1] = boots<op>(a, b, c)   (T/libtfhe/boot-gates.cpp:35-370) */
void port_gate(const port_key *key, int op, int32_t *out, const int32_t *a, const int32_t *b, const int32_t *c);
/* count gates on contiguous [count][n+1] batches, nthreads POSIX threads (one gate stream per core) */
void port_gate_batch(const port_key *key, int op, int32_t *out, const int32_t *a, const int32_t *b,
                     const int32_t *c, int64_t count, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
