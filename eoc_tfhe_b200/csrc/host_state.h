// host_state.h -- host-side state shared by abi.cpp and circuit.cpp (not part of the public ABI).
#pragma once
#include "engine.cuh"

#include <exception>
#include <functional>
#include <mutex>
#include <thread>
#include <new>
#include <string>

int b200_fail(const std::string &msg);                 // records the message, returns 1
int b200_fail_cuda(const char *what, cudaError_t e);   // same, with the CUDA error string
#define B200_CK(call)                                              \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return b200_fail_cuda(#call, e__); \
    } while (0)

// No C++ exception may cross the C ABI: entries whose bodies allocate (std::vector, std::string) run inside this.
template <class F>
static inline int b200_guarded(F &&f) noexcept {
    try {
        return f();
    } catch (const std::bad_alloc &) {
        return b200_fail("out of host memory");
    } catch (const std::exception &e) {
        return b200_fail(std::string("internal error: ") + e.what());
    } catch (...) {
        return b200_fail("internal error");
    }
}

namespace b200 {
// gate prologue coefficients (kernels.cu, shared with the device code)
void b200_gate_coeffs(int op, int half, int *s0, int *s1, int32_t *cst, int *second);
}  // namespace b200

// Every entry makes the key's device current for its own duration and puts the caller's device back on return
// (a multi-GPU host such as torch keeps allocating on the device it selected).
struct DeviceGuard {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); }
        if (prev != device) err = cudaSetDevice(device);
        else prev = -1;   // nothing to restore
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;
};
#define B200_ON_DEVICE(dev)                                                     \
    DeviceGuard dev_guard__(dev);                                               \
    if (dev_guard__.err != cudaSuccess) return b200_fail_cuda("cudaSetDevice", dev_guard__.err)

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
    std::mutex mu;                      // one ENQUEUE at a time per key (shared workspace)
    DevBuf u, in0, in1, in2, out, nz, desc_rot, desc_ks, desc_lin, desc_io, acc;
    // The workspaces above are shared by all calls on this key, and the mutex orders only the enqueue, not the
    // execution: a call on another stream than the previous one first waits (on the device) for the event the
    // previous call recorded after its last kernel.  scratch_begin / scratch_end bracket every use.
    cudaEvent_t ev_scratch = nullptr;
    cudaStream_t scratch_stream = nullptr;
    bool scratch_used = false;
    int scratch_begin(cudaStream_t s) {
        if (scratch_used && scratch_stream != s) B200_CK(cudaStreamWaitEvent(s, ev_scratch, 0));
        return 0;
    }
    int scratch_end(cudaStream_t s) {
        B200_CK(cudaEventRecord(ev_scratch, s));
        scratch_stream = s;
        scratch_used = true;
        return 0;
    }
    uint64_t fingerprint = 0;           // of the libtfhe cloud key this device key was uploaded from (cache validation)
    // Lanes: up to kMaxLanes - 1 shallow replicas of this key on the SAME device -- the same resident blobs, their
    // own streams, workspaces, staging buffers and mutex -- created when a call finds the key busy, so that calls
    // from several host threads (one gate stream per thread, as the reference's users run libtfhe) overlap on the
    // GPU instead of queueing on one mutex.  A lane is never visible to the caller.
    static constexpr int kMaxLanes = 4;
    b200_key *lanes[kMaxLanes] = {};    // [0] unused
    std::mutex lanes_mu;
    b200_key *master = nullptr;         // set in a lane
    bool owns_blobs = true;
    PinnedBuf stage_in, stage_out;
    // ring of CUDA-event triplets (before blind rotation, between, after key switch), one per gate call
    static constexpr int kRing = 64;
    cudaEvent_t ring[kRing][3] = {};
    int64_t ring_head = 0, ring_drained = 0;
    cudaEvent_t ev_start = nullptr;     // start of the most recent *_host call (incl. copies)
    float last_ms[3] = {0, 0, 0};
    bool timing_pending = false;
};

b200_key *b200_lane_create(b200_key *master);   // abi.cpp; nullptr on failure (the caller then waits for a busy lane)

// Picks a free lane of the key (the key itself first), creating one if all existing ones are busy and fewer than
// kMaxLanes exist; otherwise blocks on this thread's home lane.  Holds the lane's mutex for its lifetime.
struct LaneGuard {
    b200_key *k;
    explicit LaneGuard(b200_key *m) : k(nullptr) {
        if (m->mu.try_lock()) { k = m; return; }
        for (int i = 1; i < b200_key::kMaxLanes; ++i) {
            b200_key *l;
            {
                std::lock_guard<std::mutex> g(m->lanes_mu);
                if (!m->lanes[i]) m->lanes[i] = b200_lane_create(m);
                l = m->lanes[i];
            }
            if (l && l->mu.try_lock()) { k = l; return; }
        }
        const size_t h = std::hash<std::thread::id>()(std::this_thread::get_id()) % b200_key::kMaxLanes;
        k = (h == 0 || !m->lanes[h]) ? m : m->lanes[h];
        k->mu.lock();
    }
    ~LaneGuard() { k->mu.unlock(); }
    LaneGuard(const LaneGuard &) = delete;
    LaneGuard &operator=(const LaneGuard &) = delete;
};
