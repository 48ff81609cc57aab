// host_state.h -- host-side state shared by abi.cpp and circuit.cpp (not part of the public ABI).
#pragma once
#include "engine.cuh"

#include <mutex>
#include <string>

int b200_fail(const std::string &msg);                 // records the message, returns 1
int b200_fail_cuda(const char *what, cudaError_t e);   // same, with the CUDA error string
#define B200_CK(call)                                              \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return b200_fail_cuda(#call, e__); \
    } while (0)

namespace b200 {
// gate prologue coefficients (kernels.cu, shared with the device code)
void b200_gate_coeffs(int op, int half, int *s0, int *s1, int32_t *cst, int *second);
}  // namespace b200

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 8;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) return b200_fail_cuda("cudaMalloc(workspace)", e);
        cap = want;
        return 0;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};
struct PinnedBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMallocHost(&p, bytes);
        if (e != cudaSuccess) return b200_fail_cuda("cudaMallocHost", e);
        cap = bytes;
        return 0;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

struct b200_key {
    b200::DeviceKey dk{};
    bool filled = false;
    cudaStream_t stream = nullptr;      // used by the *_host entries
    // host batches are pipelined in slices: uploads, kernels and downloads run on three streams
    static constexpr int kSlices = 8;
    cudaStream_t stream_h2d = nullptr, stream_d2h = nullptr;
    cudaEvent_t ev_up[kSlices] = {}, ev_done[kSlices] = {};
    std::mutex mu;                      // one call at a time per key (shared workspace)
    DevBuf u, in0, in1, in2, out, nz, desc_rot, desc_ks, desc_lin, desc_io, acc;
    PinnedBuf stage_in, stage_out;
    // ring of CUDA-event triplets (before blind rotation, between, after key switch), one per gate call
    static constexpr int kRing = 64;
    cudaEvent_t ring[kRing][3] = {};
    int64_t ring_head = 0, ring_drained = 0;
    cudaEvent_t ev_start = nullptr;     // start of the most recent *_host call (incl. copies)
    float last_ms[3] = {0, 0, 0};
    bool timing_pending = false;
};
