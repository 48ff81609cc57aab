/*
 * b200tfhe.h -- C ABI of libb200tfhe.so, a B200 (sm_100a) engine for the TFHE gate-bootstrapping
 * path of libtfhe (tfhe/tfhe v1.1, vendored by EntityOfCode/eoc-tfhe under libs/tfhe).
 *
 * Three groups of entry points:
 *   1. libtfhe drop-in symbols   -- the exact names/signatures of the reference gate API
 *                                   (libs/tfhe/src/include/tfhe_gate_bootstrapping_functions.h:14-87),
 *                                   so a program linked against libb200tfhe.so *before* libtfhe gets
 *                                   GPU gates while keygen / encrypt / IO stay in libtfhe.
 *   2. bootsGate_batch           -- the batched form of (1) over arrays of LweSample pointers.
 *   3. b200_*                    -- key objects, contiguous host/device batches, the pieces of the
 *                                   path (blind rotation, key switch, one CMux step) for parity
 *                                   tests, and levelised circuits.
 *
 * Only plain pointers and sizes cross this boundary (no C++ / torch types).
 * All `int`-returning functions return 0 on success, non-zero on failure; the message is
 * available from b200_last_error().  The void libtfhe-named gates abort() with a message on
 * failure, like the reference's die_dramatically (libs/tfhe/src/libtfhe/tfhe_gate_bootstrapping.cpp:11-15).
 *
 * There is no CPU fallback: every compute entry fails if no CUDA device is usable.
 */
#ifndef B200TFHE_H
#define B200TFHE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * libtfhe-compatible data layouts.  These mirror (field order, types, padding) the reference
 * structs so that pointers produced by libtfhe can be passed here unchanged.  Define
 * B200TFHE_NO_LIBTFHE_TYPES before including this header when <tfhe.h> is also included.
 * ------------------------------------------------------------------------------------------ */
#ifndef B200TFHE_NO_LIBTFHE_TYPES
typedef int32_t Torus32;                       /* libs/tfhe/src/include/tfhe_core.h:28 */

struct LweParams {                             /* include/lweparams.h:8-19 */
    int32_t n;
    double  alpha_min;
    double  alpha_max;
};
struct LweSample {                             /* include/lwesamples.h:9-19 */
    Torus32 *a;                                /* n mask coefficients, own heap block */
    Torus32  b;
    double   current_variance;
};
struct TLweParams {                            /* include/tlwe.h:10-28 */
    int32_t N;
    int32_t k;
    double  alpha_min;
    double  alpha_max;
    struct LweParams extracted_lweparams;
};
struct TGswParams {                            /* include/tgsw.h:10-32 */
    int32_t  l;
    int32_t  Bgbit;
    int32_t  Bg;
    int32_t  halfBg;
    uint32_t maskMod;
    const struct TLweParams *tlwe_params;
    int32_t  kpl;
    Torus32 *h;
    uint32_t offset;
};
struct TorusPolynomial {                       /* include/polynomials.h:22-32 */
    int32_t  N;
    Torus32 *coefsT;
};
struct TLweSample {                            /* include/tlwe.h:47-63 */
    struct TorusPolynomial *a;                 /* k+1 polynomials */
    struct TorusPolynomial *b;                 /* alias of a[k] */
    double  current_variance;
    int32_t k;
};
struct TGswSample {                            /* include/tgsw.h:55-76 */
    struct TLweSample  *all_sample;            /* (k+1)*l rows */
    struct TLweSample **bloc_sample;
    int32_t k;
    int32_t l;
};
struct LweKeySwitchKey {                       /* include/lwekeyswitch.h:11-28 */
    int32_t n;                                 /* input dimension (N*k) */
    int32_t t;
    int32_t basebit;
    int32_t base;
    const struct LweParams *out_params;
    struct LweSample   *ks0_raw;
    struct LweSample  **ks1_raw;
    struct LweSample ***ks;                    /* ks[i][j][h] */
};
struct LweBootstrappingKey {                   /* include/lwebootstrappingkey.h:7-30 */
    const struct LweParams  *in_out_params;
    const struct TGswParams *bk_params;
    const struct TLweParams *accum_params;
    const struct LweParams  *extract_params;
    struct TGswSample       *bk;               /* n TGSW samples, integer domain */
    struct LweKeySwitchKey  *ks;
};
struct LweBootstrappingKeyFFT {                /* include/lwebootstrappingkey.h:36-59 */
    const struct LweParams  *in_out_params;
    const struct TGswParams *bk_params;
    const struct TLweParams *accum_params;
    const struct LweParams  *extract_params;
    const void              *bkFFT;            /* FFT-processor private; never read here */
    const struct LweKeySwitchKey *ks;
};
struct TFheGateBootstrappingParameterSet {     /* include/tfhe_gate_bootstrapping_structures.h:9-24 */
    int32_t ks_t;
    int32_t ks_basebit;
    const struct LweParams  *in_out_params;
    const struct TGswParams *tgsw_params;
};
struct TFheGateBootstrappingCloudKeySet {      /* include/tfhe_gate_bootstrapping_structures.h:26-42 */
    const struct TFheGateBootstrappingParameterSet *params;
    const struct LweBootstrappingKey    *bk;
    const struct LweBootstrappingKeyFFT *bkFFT;
};
#ifndef __cplusplus
typedef struct LweParams LweParams;
typedef struct LweSample LweSample;
typedef struct TFheGateBootstrappingParameterSet TFheGateBootstrappingParameterSet;
typedef struct TFheGateBootstrappingCloudKeySet TFheGateBootstrappingCloudKeySet;
#endif
#endif /* B200TFHE_NO_LIBTFHE_TYPES */

/* ------------------------------------------------------------------------------------------
 * 1. libtfhe drop-in gate symbols (libs/tfhe/src/libtfhe/boot-gates.cpp:35-370).
 *    `bk` is uploaded to the GPU on first use and cached by pointer (b200_cloudkey_upload does
 *    it eagerly).  `result` may alias an input.  result->current_variance is written as
 *    (#non-zero keyswitch digits) * variance-of-a-KS-row, as lweKeySwitch does
 *    (libs/tfhe/src/libtfhe/lwe-functions.cpp:222-234).
 * ------------------------------------------------------------------------------------------ */
void bootsNAND (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:35  */
void bootsOR   (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:61  */
void bootsAND  (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:87  */
void bootsXOR  (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:113 */
void bootsXNOR (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:139 */
void bootsNOR  (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:198 */
void bootsANDNY(struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:224 */
void bootsANDYN(struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:250 */
void bootsORNY (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:276 */
void bootsORYN (struct LweSample *result, const struct LweSample *ca, const struct LweSample *cb, const struct TFheGateBootstrappingCloudKeySet *bk);  /* boot-gates.cpp:302 */
void bootsMUX  (struct LweSample *result, const struct LweSample *a, const struct LweSample *b, const struct LweSample *c, const struct TFheGateBootstrappingCloudKeySet *bk); /* boot-gates.cpp:329 */
void bootsNOT  (struct LweSample *result, const struct LweSample *ca, const struct TFheGateBootstrappingCloudKeySet *bk);                              /* boot-gates.cpp:164 */
void bootsCOPY (struct LweSample *result, const struct LweSample *ca, const struct TFheGateBootstrappingCloudKeySet *bk);                              /* boot-gates.cpp:175 */
void bootsCONSTANT(struct LweSample *result, int32_t value, const struct TFheGateBootstrappingCloudKeySet *bk);                                        /* boot-gates.cpp:185 */

/* Lower-level public functions of the same path (include/tfhe.h:42-45, include/lwe-functions.h:79).
 * They take the reference's LweBootstrappingKeyFFT / LweKeySwitchKey handles; the device key is
 * found through the cloud key they belong to, which must have been passed to
 * b200_cloudkey_upload() before. */
void tfhe_bootstrap_FFT     (struct LweSample *result, const struct LweBootstrappingKeyFFT *bk, Torus32 mu, const struct LweSample *x);  /* lwe-bootstrapping-functions-fft.cpp:213 */
void tfhe_bootstrap_woKS_FFT(struct LweSample *result, const struct LweBootstrappingKeyFFT *bk, Torus32 mu, const struct LweSample *x);  /* lwe-bootstrapping-functions-fft.cpp:169 */
void lweKeySwitch           (struct LweSample *result, const struct LweKeySwitchKey *ks, const struct LweSample *sample);               /* lwe-keyswitch-functions.cpp:228 */

/* libtfhe's key destructors, interposed so that the device key cached for a host key dies with it: they drop the
 * cache entry (b200_cloudkey_forget) and then call libtfhe's own definition, found with dlsym(RTLD_NEXT) -- so
 * they do their job only in the drop-in link order (this library before libtfhe).  Independently of them a cache
 * hit is re-validated against a fingerprint of the key material, so a new key set that malloc placed at a deleted
 * one's address is uploaded afresh instead of being evaluated with the old device key. */
void delete_gate_bootstrapping_cloud_keyset(struct TFheGateBootstrappingCloudKeySet *keyset);  /* tfhe_gate_bootstrapping.cpp:124 */
void delete_LweBootstrappingKeyFFT(struct LweBootstrappingKeyFFT *obj);                        /* lwe-bootstrapping-functions-fft.cpp:290 */

/* ------------------------------------------------------------------------------------------
 * 2. Batched gate over libtfhe samples (new).
 *    Snapshot semantics: every input of the batch is read before any output is written, so
 *    out[] may alias inputs.  b/c may be NULL when the op does not use them.
 * ------------------------------------------------------------------------------------------ */
enum b200_gate_op {
    B200_NAND = 0, B200_OR = 1, B200_AND = 2, B200_XOR = 3, B200_XNOR = 4, B200_NOR = 5,
    B200_ANDNY = 6, B200_ANDYN = 7, B200_ORNY = 8, B200_ORYN = 9,
    B200_MUX = 10,                 /* out = a ? b : c  (two blind rotations, one key switch) */
    B200_NOT = 11, B200_COPY = 12, /* linear, no bootstrap */
    B200_CONST0 = 13, B200_CONST1 = 14,
    B200_GATE_OP_COUNT = 15
};

int bootsGate_batch(int op, struct LweSample *const *out, const struct LweSample *const *a,
                    const struct LweSample *const *b, const struct LweSample *const *c,
                    int32_t count, const struct TFheGateBootstrappingCloudKeySet *bk);

/* ------------------------------------------------------------------------------------------
 * 3. b200_* : keys, contiguous batches, path pieces, circuits.
 *    A ciphertext batch is int32 [count][n+1] row-major: a[0..n) then b (the order of the
 *    reference's wire format, libs/tfhe/src/libtfhe/tfhe_io.cpp:101-107).
 *    An extracted-dimension batch is int32 [count][N*k+1].
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_key b200_key;

typedef struct b200_params {
    int32_t n;          /* LWE dimension (500 | 630) */
    int32_t N;          /* ring degree, must be 1024 (as in every reference FFT processor) */
    int32_t k;          /* must be 1 */
    int32_t l;          /* gadget length (2 | 3), 1..4 supported */
    int32_t Bgbit;      /* log2 gadget base (10 | 7) */
    int32_t ks_t;       /* key-switch digits (8) */
    int32_t ks_basebit; /* key-switch log2 base (2) */
    double  ks_variance;/* variance carried by every KS row (current_variance bookkeeping) */
} b200_params;

const char *b200_last_error(void);
int  b200_device_count(void);
int  b200_set_device(int device);            /* device used by subsequently created keys */

/* Key from a live libtfhe cloud key (reads bk->bk[i].all_sample[p].a[q].coefsT and
 * bk->ks->ks[i][j][h], cf. lwe-bootstrapping-functions-fft.cpp:23-52) -- converts the
 * bootstrapping key to the FFT domain ON THE GPU and uploads the key-switch table. The handle is
 * cached by `bk` pointer so that the drop-in gates find it. */
int  b200_cloudkey_upload(const struct TFheGateBootstrappingCloudKeySet *bk, b200_key **out);
void b200_cloudkey_forget(const struct TFheGateBootstrappingCloudKeySet *bk); /* frees the cached key */

/* Key from raw integer arrays (what the wire format carries, tfhe_io.cpp:757-930):
 *   bk_coefs : int32 [n][(k+1)*l][k+1][N]     bk[i].all_sample[row].a[col].coefsT
 *   ks_rows  : int32 [N*k][t][base][n+1]       ks[i][j][h] as (a[0..n), b); h=0 rows are ignored */
int  b200_key_create(const b200_params *p, const int32_t *bk_coefs, const int32_t *ks_rows, b200_key **out);
/* Key with device storage allocated but not filled (multi-GPU: non-root ranks, then
 * b200_key_device_blobs + a broadcast into them). */
int  b200_key_create_empty(const b200_params *p, b200_key **out);
void b200_key_free(b200_key *key);
int  b200_key_params(const b200_key *key, b200_params *out);
/* Device pointers/sizes of the two resident blobs: FFT-domain bootstrapping key and KS table. */
int  b200_key_device_blobs(const b200_key *key, void **d_bkfft, size_t *bkfft_bytes,
                           void **d_ks, size_t *ks_bytes);

/* Gates on contiguous HOST batches (H2D + kernels + D2H inside the call). */
int  b200_gate_batch_host(b200_key *key, int op, int32_t *out, const int32_t *a, const int32_t *b,
                          const int32_t *c, int64_t count);
/* Gates on DEVICE-resident batches; `stream` is a cudaStream_t (NULL = default stream).
 * Asynchronous: returns after enqueueing. out must not overlap inputs.
 * Streams: a key owns one set of device workspaces (extracted samples, descriptors).  Calls on one key may come
 * from any stream and any host thread: the enqueue is serialised by a mutex, and a call on a different stream
 * than the previous call on that key first waits -- on the device, via an event -- for the previous call's last
 * kernel.  Calls on DIFFERENT keys never wait for each other.  The *_host entries use the key's own stream and
 * follow the same rule, so mixing *_dev and *_host calls on one key is safe.
 * Device: every entry makes the key's device current while it runs and restores the caller's device on return.
 * Timing (b200_last_timing_ms / b200_timing_drain): covers the first blind-rotation + key-switch launch pair of
 * a call; a call that is split into chunks (> 2^18 rotations) or pipelined slices reports its first chunk. */
int  b200_gate_batch_dev(b200_key *key, int op, int32_t *d_out, const int32_t *d_a,
                         const int32_t *d_b, const int32_t *d_c, int64_t count, void *stream);
/* The same gates, same arguments and same results as b200_gate_batch_dev, for callers that cannot allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The call allocates nothing -- no cudaMalloc / cudaFree, no stream,
 * event or lane creation: the extracted samples that travel from the blind rotation to the key switch live in a
 * static device array inside the library, B200_NOALLOC_MAX_ROTATIONS x 1025 words = 16.8 MB, which every GPU that
 * loads the library holds until exit.
 * Limit: count (MUX: 2 x count) must not exceed B200_NOALLOC_MAX_ROTATIONS blind rotations; larger batches are
 * rejected with rc 1 and a b200_last_error() that names the limit.  Keys with N * k != 1024 are rejected.
 * Sharing: every call on a device uses the same array, whatever the key, and it keeps the previous call's contents
 * (nothing in it needs to start at zero: each call writes the rows it later reads).  Calls ordered on one stream are
 * safe; calls that overlap on the GPU (different streams without ordering, several host threads) RACE on the array
 * and return wrong ciphertexts -- b200_gate_batch_dev has per-key workspaces and no such restriction.
 * No timing is recorded for these calls (b200_last_timing_ms / b200_timing_drain do not see them). */
#define B200_NOALLOC_MAX_ROTATIONS 4096
int  b200_gate_batch_dev_noalloc(b200_key *key, int op, int32_t *d_out, const int32_t *d_a,
                                 const int32_t *d_b, const int32_t *d_c, int64_t count, void *stream);

/* Several devices behind one handle (north_star: gates partitioned over the GPUs of a box, cloud key replicated).
 * b200_multi_create replicates `key` (which stays owned by the caller and must outlive the handle) onto the listed
 * devices -- devices[0] must be the key's own device; devices == NULL takes the first ndev devices of the box
 * (ndev <= 0: all), the key's own first -- by device-to-device copies of the two resident blobs.
 * b200_multi_gate_batch_host cuts a host batch into contiguous slices of whole blind-rotation waves, one per
 * device, each evaluated by its own host thread through b200_gate_batch_host's path; results are identical to the
 * single-device call word for word.  The drop-in bootsGate_batch does the same when the environment variable
 * B200TFHE_DEVICES is "all" or a device count. */
typedef struct b200_multi b200_multi;
int  b200_multi_create(b200_key *key, const int32_t *devices, int32_t ndev, b200_multi **out);
void b200_multi_free(b200_multi *m);
int  b200_multi_device_count(const b200_multi *m);
int  b200_multi_gate_batch_host(b200_multi *m, int op, int32_t *out, const int32_t *a, const int32_t *b,
                                const int32_t *c, int64_t count);

/* Pieces of the path, for parity tests and for callers that hold extracted samples. */
/* u[count][N+1] = tfhe_bootstrap_woKS_FFT(x[count][n+1], mu)       (fft.cpp:169-200) */
int  b200_bootstrap_woks_host(b200_key *key, int32_t *u, const int32_t *x, int32_t mu, int64_t count);
/* out[count][n+1] = lweKeySwitch(u[count][N+1]); nonzero_digits[count] optional (may be NULL) */
int  b200_keyswitch_host(b200_key *key, int32_t *out, const int32_t *u, int32_t *nonzero_digits, int64_t count);
int  b200_keyswitch_dev (b200_key *key, int32_t *d_out, const int32_t *d_u, int64_t count, void *stream);
/* The negacyclic transform alone (replaces the reference's FFT processors, fft_processor_spqlios.cpp:38-115), with
 * the checks the reference keeps for them (src/test/lagrangehalfc_test.cpp:16-36, 130-218).  Polynomials are
 * int32 [count][1024] mod X^1024 + 1; the transform is the one the selected blind-rotation kernel uses.
 *   b200_fft_roundtrip_host: out = round(inverse(forward(a)) / 512)         -- the identity on any int32 input
 *   b200_poly_mul_host:      out = a * b mod 2^32, a small (|a_j| <= 512: gadget digits), b any Torus32 */
int  b200_fft_roundtrip_host(int32_t *out, const int32_t *a, int64_t count);
int  b200_poly_mul_host(int32_t *out, const int32_t *a, const int32_t *b, int64_t count);
/* One CMux step: acc[count][k+1][N] <- acc + BK_i (x) ((X^barai - 1) * acc)   (fft.cpp:67-76).
 * With barai == 0 the accumulator is returned unchanged, as tfhe_blindRotate_FFT skips (fft.cpp:101). */
int  b200_cmux_step_host(b200_key *key, int32_t *acc, int32_t bk_index, int32_t barai, int64_t count);
/* tGswFFTExternMulToTLwe alone: acc[count][k+1][N] <- BK_i (x) acc   (tgsw-fft-operations.cpp:89-114) */
int  b200_extern_product_host(b200_key *key, int32_t *acc, int32_t bk_index, int64_t count);

/* Levelised circuits: a netlist over a pool of wires, all independent gates of
 * one depth level run as one launch pair.
 *   gates : int32 [num_gates][5] = {op, out_wire, in_a, in_b, in_c} (unused inputs = -1)
 *   wires : int32 [num_wires][n+1] host pool, inputs pre-filled, outputs written back.
 * Gates must be listed in a valid topological order; each wire is written at most once. */
int  b200_circuit_run_host(b200_key *key, const int32_t *gates, int64_t num_gates,
                           int32_t *wires, int64_t num_wires, int32_t *levels_out);
int  b200_circuit_run_dev (b200_key *key, const int32_t *gates, int64_t num_gates,
                           int32_t *d_wires, int64_t num_wires, int32_t *levels_out, void *stream);
/* The same circuit with packed host I/O: only the rows of the listed input wires go up (in_rows
 * [n_in][n+1] -> wires in_wires[i]) and only the listed output wires come back (out_rows [n_out][n+1]);
 * the pool lives on the device only.  Wires that are neither written by a gate nor listed as inputs read
 * as all-zero samples. */
int  b200_circuit_run_host_io(b200_key *key, const int32_t *gates, int64_t num_gates, int64_t num_wires,
                              const int32_t *in_wires, int64_t n_in, const int32_t *in_rows,
                              const int32_t *out_wires, int64_t n_out, int32_t *out_rows, int32_t *levels_out);

/* Host-only netlist helpers (no device needed): level of every gate (number of bootstrapped stages
 * before its output is ready; linear gates share the level of their input) and validation. */
int  b200_circuit_levelize(const int32_t *gates, int64_t num_gates, int64_t num_wires,
                           int32_t *gate_level, int32_t *num_levels);
/* Expand bootstrapped gates into launch descriptors:
 *   rot_desc rows {w0, w1, s0 | s1<<16, cst}   x = (0,cst) + s0*W[w0] + s1*W[w1]   (MUX: two rows)
 *   ks_desc  rows {u0, u1, add_b, out_wire}    out = KS((0,add_b) + U[u0] (+ U[u1])) */
int  b200_expand_gates(const int32_t *gates, int64_t num_gates, int32_t *rot_desc, int64_t *n_rot,
                       int32_t *ks_desc, int64_t *n_ks);
/* One level given as descriptors (host arrays) over a device wire pool: one blind-rotation launch
 * and one key-switch launch on `stream`. Used by multi-GPU drivers that shard a level by rank. */
int  b200_level_run_dev(b200_key *key, int32_t *d_wires, const int32_t *rot_desc, int64_t n_rot,
                        const int32_t *ks_desc, int64_t n_ks, void *stream);

/* ------------------------------------------------------------------------------------------
 * eoc evaluation-side entry points (ao-tfhe/eoc-tfhe-run.cpp).  Key generation,
 * encryption and decryption need the secret key and stay in the reference.
 * ------------------------------------------------------------------------------------------ */
/* eoc-tfhe-run.cpp:427-469 / :472-513: base64(LweSample record) x 2 -> base64(record of the sum /
 * difference): lweCopy + lweAddTo / lweSubTo, int32 wrap-around, variances add.  Host integer work,
 * bit-exact.  The third argument is ignored, as in the reference.  Returns a new[] string (release it
 * with b200_eoc_free) or NULL with the message on stderr and in b200_last_error().  Unlike the
 * reference no key has to be loaded first: the dimension n comes from the record length. */
const char *addCiphertexts     (const char *base64Ciphertext1, const char *base64Ciphertext2, const char *base64PublicKey);
const char *subtractCiphertexts(const char *base64Ciphertext1, const char *base64Ciphertext2, const char *base64PublicKey);
void b200_eoc_free(const char *s);
/* Evaluation key of the bit-sliced entry points below: base64 of what
 * export_tfheGateBootstrappingCloudKeySet_toStream writes (the library owns the device key), or an
 * existing handle (not owned).  Mirrors eoc's process-global key (eoc-tfhe-run.cpp:38-39). */
int  b200_eoc_set_cloud_key(const char *base64_cloud_key);
int  b200_eoc_set_key(b200_key *key);
/* Integers encrypted bit by bit (LSB first, one gate-bootstrapping ciphertext per bit, as
 * test-addition-boot.cpp:190-197) added / subtracted with the bootstrapped ripple-carry adder
 * (full_adder, test-addition-boot.cpp:70-94; x - y = x + ~y + 1), every depth level of all `count`
 * operand pairs as one launch pair.  x, y: [count][nb_bits][n+1]; sum: [count][nb_bits+1][n+1], bit
 * nb_bits = carry out (subtract: 1 = no borrow). */
int  b200_add_bits_host(b200_key *key, int32_t *sum, const int32_t *x, const int32_t *y, int32_t nb_bits,
                        int64_t count, int subtract);
/* The same for one pair of base64 strings holding nb_bits concatenated sample records; returns
 * nb_bits + 1 records (variance fields 0).  NULL + stderr message when no key is set (as
 * eoc-tfhe-run.cpp:464-468) or the operands do not match the key's dimension. */
const char *addCiphertextBits     (const char *base64X, const char *base64Y, int32_t nb_bits);
const char *subtractCiphertextBits(const char *base64X, const char *base64Y, int32_t nb_bits);

/* ------------------------------------------------------------------------------------------
 * libtfhe wire format (host only, no device needed except b200_key_from_cloudkey_*): what
 * export_tfheGateBootstrappingCloudKeySet_toFile and export_gate_bootstrapping_ciphertext_toFile
 * write (libs/tfhe/src/libtfhe/tfhe_io.cpp:90-107,757-785,883-905,1031-1035,1099-1103).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_wire_cloudkey_info {
    double lwe_alpha_min, lwe_alpha_max;     /* LWEPARAMS */
    double tlwe_alpha_min, tlwe_alpha_max;   /* TLWEPARAMS */
    double bk_variance;                      /* variance stored with the bootstrapping-key block */
    size_t ks_offset, ks_bytes;              /* int32 [N*k][t][base][n+1] inside the buffer */
    size_t bk_offset, bk_bytes;              /* int32 [n][(k+1)l][k+1][N] inside the buffer */
    size_t total_bytes;                      /* bytes consumed by the whole cloud key */
} b200_wire_cloudkey_info;

/* Parse the text sections of a serialized cloud key and locate its two binary blocks. */
int  b200_wire_parse_cloudkey(const void *buf, size_t len, b200_params *p, b200_wire_cloudkey_info *info);
/* Device key straight from a serialized cloud key (new_tfheGateBootstrappingCloudKeySet_fromFile, tfhe_io.cpp:1127). */
int  b200_key_from_cloudkey_bytes(const void *buf, size_t len, b200_key **out);
int  b200_key_from_cloudkey_file(const char *path, b200_key **out);
/* Serialize a cloud key from raw arrays (the layouts of b200_key_create); `alphas` may be NULL. */
int64_t b200_wire_cloudkey_size(const b200_params *p);
int  b200_wire_write_cloudkey(void *buf, size_t cap, size_t *written, const b200_params *p,
                              const b200_wire_cloudkey_info *alphas, const int32_t *bk_coefs, const int32_t *ks_rows);
/* LWE sample records (uid 42, a[n], b, double variance) <-> contiguous batch int32 [count][n+1]. */
int64_t b200_wire_sample_bytes(int32_t n);
int  b200_wire_read_samples(const void *buf, size_t len, int32_t n, int64_t count, int32_t *cts, double *variances);
int  b200_wire_write_samples(void *buf, size_t cap, int32_t n, int64_t count, const int32_t *cts, const double *variances);

/* Timing helper: milliseconds of GPU time (cudaEvent, on the key's stream) spent in the kernels
 * of the most recent *_host / *_dev call on this key: [0]=blind rotation, [1]=key switch,
 * [2]=total incl. copies. Synchronises the stream. */
int  b200_last_timing_ms(b200_key *key, float out_ms[3]);
/* Sum of the per-kernel GPU times ([0] blind rotation, [1] key switch, ms) of every gate-batch call
 * since the previous drain (at most the last 64 calls), and how many calls that was. Synchronises. */
int  b200_timing_drain(b200_key *key, double sum_ms[2], int32_t *calls);
/* FP64 FMA throughput of `device` measured by a DFMA-chain kernel (TFLOP/s): the roofline
 * denominator bench.py reports the blind rotation against. */
int  b200_fp64_peak_tflops(int device, double *tflops);
/* 32-bit integer add throughput of `device` measured by a chain kernel (10^12 adds/s): the roofline denominator of
 * the key switch, whose work is 6,144 row subtractions of n+1 words per sample (lwe-keyswitch-functions.cpp:114-129). */
int  b200_int32_add_peak(int device, double *tera_adds_per_s);
/* Blind-rotation kernel used by subsequent calls: 0 = 64 threads per ciphertext, accumulators in
 * registers (default); 1 = one warp per ciphertext, 16-point transforms, accumulators in tensor memory
 * (tcgen05.ld/st); 2..6 = the 64-thread layout with accumulators in tensor memory (2: 6 groups per SM,
 * 3: 4 groups of two ciphertexts sharing every key read, 4..6: other group shapes); 7 = key rows staged
 * TMA -> shared memory -> tcgen05.cp -> tensor memory, read by the spectral MAC with tcgen05.ld; 8 = two
 * warps per ciphertext, warp h owning accumulator polynomial h, 16-point transforms (l = 2 only, wide
 * batches; falls back to 0 otherwise).  Every variant is bit-identical to 0; the
 * choice is a performance knob that bench.py and the parity tests exercise.  Batches of at most one
 * ciphertext per SM use the latency kernel (one ciphertext per CTA) while the variant is 0.  Also settable with B200_BR_KERNEL=v2|w16|<number> before first use. */
int  b200_set_blind_rotate_kernel(int variant);
/* 1 when the library was compiled with -DB200_EXPERIMENTS, i.e. variants 1..8 above exist; the shipped build has
 * variant 0 only and b200_set_blind_rotate_kernel(v != 0) fails. */
int  b200_experiments_built(void);
/* Key-switch kernel: 0 = chosen by batch size (default), 1 = gather kernel (8-sample tiles, any batch),
 * 2 = table-streaming kernel (64 samples per CTA, table rows staged by TMA).  Results are identical. */
int  b200_set_keyswitch_kernel(int variant);
/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
int64_t b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200TFHE_H */
