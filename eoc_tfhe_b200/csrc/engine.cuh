// engine.cuh -- internal C++ interface between the C ABI (abi.cpp) and the CUDA kernels (kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace b200 {

// Rotation-item descriptor for netlists (one blind rotation): x = cst + s0*W[w0] + s1*W[w1]
struct RotDesc {
    int32_t w0, w1;   // wire indices into the pool (w1 < 0: unused)
    int32_t s0s1;     // (s0 & 0xffff) | (s1 << 16), signed 16-bit each
    int32_t cst;      // added to the b term
};
// Key-switch item descriptor: out_wire <- KS( (0,add_b) + U[u0] (+ U[u1]) )
struct KsDesc {
    int32_t u0, u1;   // rows of the extracted buffer (u1 < 0: unused)
    int32_t add_b;
    int32_t out;      // destination row (wire index or batch row)
};

struct DeviceKey {
    int n, N, k, l, Bgbit, ks_t, ks_basebit;
    int ks_stride;            // n+1 rounded up to a multiple of 4 (int4 loads)
    double ks_variance;
    double2 *bkfft;           // [n][kpl][k+1][8][64]  evaluation domain, pass-3 layout of the 64-thread transform
    double2 *bkfft_tm;        // same shape, evaluation order of the tensor-memory transform (negacyclic_fft_tm.cuh)
    double2 *bkfft16;         // [n][kpl][k+1][16][32] evaluation domain, [r][lane] layout of the warp transform (experiments)
    int32_t *ks;              // [N*k][t][base-1][ks_stride]
    size_t bkfft_bytes, ks_bytes;
    int device;
};

struct BlindRotateArgs {
    // implicit batch form (desc == nullptr): item g uses rows g of in0/in1/in2 and `op`
    const int32_t *in0, *in1, *in2;
    int op;                   // b200_gate_op, or kOpRaw: x = in0[g] as is
    // descriptor form
    const RotDesc *desc;
    const int32_t *pool;      // wires [*, n+1]
    int32_t mu;               // test-vector value (2^29 for gates)
    int32_t *u;               // out: [items][N+1]
    int64_t items;
    // balanced grid of the multi-ciphertext CTA (set by the launcher): CTA b owns bal_q + (b < bal_r)
    // consecutive items; bal_q == 0: CTA b owns items [b*G, b*G + G)
    int bal_q = 0, bal_r = 0;
    int stagger = 0;          // clk: groups that share warp schedulers start (g >> 1) * stagger apart (A/B knob)
};
constexpr int kOpRaw = 100;

struct KeySwitchArgs {
    const int32_t *u;         // [*, N+1]
    const KsDesc *desc;       // nullptr: item g reads row g (pair==0) or rows 2g,2g+1 (+2^29) (pair==1)
    int pair;
    int32_t *out;             // rows of n+1
    int32_t *nonzero;         // optional per-item count of non-zero digits
    int64_t items;
};

// host-callable launchers (kernels.cu)
cudaError_t upload_twiddles();
cudaError_t launch_bk_convert(const DeviceKey &key, const int32_t *d_bk_coefs, cudaStream_t s);
cudaError_t launch_ks_pack(const DeviceKey &key, const int32_t *d_ks_rows, cudaStream_t s);
cudaError_t launch_blind_rotate(const DeviceKey &key, const BlindRotateArgs &a, cudaStream_t s);
cudaError_t launch_keyswitch(const DeviceKey &key, const KeySwitchArgs &a, cudaStream_t s);
// blind rotation + key switch of one gate batch with the extracted samples in a static device array (kernels.cu);
// at most kNoAllocMaxRotations rotations (MUX: two per gate); the key's device must be current
constexpr int64_t kNoAllocMaxRotations = 4096;
cudaError_t launch_gates_noalloc(const DeviceKey &key, int op, int32_t *d_out, const int32_t *d_a, const int32_t *d_b,
                                 const int32_t *d_c, int64_t count, cudaStream_t s);
// mode 0: acc <- acc + BK_i (x) ((X^a - 1) acc)   mode 1: acc <- BK_i (x) acc
cudaError_t launch_cmux_test(const DeviceKey &key, int32_t *d_acc, int bk_index, int barai, int mode,
                             int64_t count, cudaStream_t s);
// the transform alone: mode 0 out = inverse(forward(a)); mode 1 out = a * b mod (X^1024 + 1, 2^32)
cudaError_t launch_fft_test(const int32_t *d_a, const int32_t *d_b, int32_t *d_out, int mode, int64_t count, cudaStream_t s);
// linear gates on device batches: out = s*in (+cst on b)
cudaError_t launch_linear(const DeviceKey &key, int32_t *d_out, const int32_t *d_in, int sign, int32_t cst,
                          int64_t count, cudaStream_t s);
cudaError_t launch_linear_desc(const DeviceKey &key, int32_t *pool, const int32_t *d_desc /*[items][4]: op,out,in,unused*/,
                               int64_t items, cudaStream_t s);
int64_t launch_count();
void set_blind_rotate_variant(int v);  // 0: 64 threads per ciphertext, 1: warp per ciphertext (TMEM)
bool experiments_built();              // compiled with -DB200_EXPERIMENTS (round-1 blind-rotation variants 1..8)
void set_keyswitch_variant(int v);     // 0: chosen by batch size, 1: gather kernel, 2: table-streaming kernel
cudaError_t measure_fp64_peak(double *tflops);
cudaError_t measure_int_add_peak(double *tera_adds_per_s);

}  // namespace b200
