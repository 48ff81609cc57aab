// abi.cpp -- host side of libb200tfhe.so: the C ABI declared in include/b200tfhe.h.
//
// Mirrors the reference's operator interface for the gate-bootstrapping path
// (libs/tfhe/src/include/tfhe_gate_bootstrapping_functions.h:14-87, include/tfhe.h:42-45,
// include/lwe-functions.h:79): same names, argument meaning, aliasing and error behaviour.
// No CPU implementation of the path exists in this library: every entry needs a CUDA device.
#include "../../include/b200tfhe.h"
#include "engine.cuh"
#include "host_state.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <map>
#include <new>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace b200;

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string t_last_error;
int b200_fail(const std::string &msg) {
    t_last_error = msg;
    return 1;
}
int b200_fail_cuda(const char *what, cudaError_t e) {
    return b200_fail(std::string(what) + ": " + cudaGetErrorString(e));
}
static int fail(const std::string &msg) { return b200_fail(msg); }
static int fail_cuda(const char *what, cudaError_t e) { return b200_fail_cuda(what, e); }
#define CK(call) B200_CK(call)

extern "C" const char *b200_last_error(void) { return t_last_error.c_str(); }

static void die(const char *fn) {
    // the reference's void gates abort on fatal errors (tfhe_gate_bootstrapping.cpp:11-15)
    std::fprintf(stderr, "libb200tfhe: %s failed: %s\n", fn, t_last_error.c_str());
    std::abort();
}

// ------------------------------------------------------------------------------------------------
// key object
// ------------------------------------------------------------------------------------------------
static int g_device = 0;
static std::mutex g_cache_mu;
static std::map<const void *, b200_key *> g_by_cloud, g_by_bkfft, g_by_ks;

extern "C" int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
extern "C" int b200_set_device(int device) {
    CK(cudaSetDevice(device));
    g_device = device;
    return 0;
}
extern "C" int64_t b200_launch_count(void) { return launch_count(); }
extern "C" int b200_set_blind_rotate_kernel(int variant) {
    if (variant < 0 || variant > 12) return b200_fail("blind-rotation kernel variant must be 0..12");
    if (variant != 0 && variant < 9 && !experiments_built())
        return b200_fail("blind-rotation variants 1..8 are round-1 experiments: rebuild with -DB200_EXPERIMENTS");
    set_blind_rotate_variant(variant);
    return 0;
}
extern "C" int b200_experiments_built(void) { return experiments_built() ? 1 : 0; }

extern "C" int b200_set_keyswitch_kernel(int variant) {
    if (variant < 0 || variant > 2) return b200_fail("key-switch kernel variant must be 0 (auto), 1 (gather) or 2 (streaming)");
    set_keyswitch_variant(variant);
    return 0;
}

static int check_params(const b200_params *p) {
    if (!p) return fail("params is NULL");
    if (p->N != 1024) return fail("only N=1024 is supported (as in every reference FFT processor)");
    if (p->k != 1) return fail("only k=1 is supported");
    if (p->n < 1 || p->n > 639) return fail("n out of range (1..639)");
    const bool ok = (p->l == 2 && p->Bgbit == 10) || (p->l == 3 && p->Bgbit == 7) || (p->l == 2 && p->Bgbit == 8) ||
                    (p->l == 1 && p->Bgbit == 16) || (p->l == 4 && p->Bgbit == 6);
    if (!ok) return fail("unsupported (l, Bgbit); built: (2,10) (3,7) (2,8) (1,16) (4,6)");
    if (p->ks_t != 8 || p->ks_basebit != 2) return fail("only ks_t=8, ks_basebit=2 is supported (both reference sets)");
    return 0;
}

// streams, events: everything of a key object but the resident blobs (also what a lane owns)
static cudaError_t key_init_runtime(b200_key *k) {
    cudaError_t e = cudaStreamCreateWithFlags(&k->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&k->stream_h2d, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&k->stream_d2h, cudaStreamNonBlocking);
    for (int i = 0; i < b200_key::kSlices && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&k->ev_up[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&k->ev_done[i], cudaEventDisableTiming);
    }
    for (int i = 0; i < b200_key::kRing * 3 && e == cudaSuccess; ++i) e = cudaEventCreate(&k->ring[i / 3][i % 3]);
    if (e == cudaSuccess) e = cudaEventCreate(&k->ev_start);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&k->ev_scratch, cudaEventDisableTiming);
    return e;
}
b200_key *b200_lane_create(b200_key *master) {
    DeviceGuard dg(master->dk.device);
    if (dg.err != cudaSuccess) return nullptr;
    b200_key *l = new (std::nothrow) b200_key();
    if (!l) return nullptr;
    l->dk = master->dk;          // the same resident blobs
    l->owns_blobs = false;
    l->master = master;
    l->filled = true;
    if (key_init_runtime(l) != cudaSuccess) {
        cudaGetLastError();
        b200_key_free(l);
        return nullptr;
    }
    return l;
}

static int key_alloc(const b200_params *p, b200_key **out, int device = -1) {
    if (!out) return fail("out is NULL");
    *out = nullptr;
    if (check_params(p)) return 1;
    if (b200_device_count() <= 0) return fail("no CUDA device: libb200tfhe has no CPU fallback");
    if (device < 0) device = g_device;
    if (device >= b200_device_count()) return fail("device index out of range");
    B200_ON_DEVICE(device);
    b200_key *k = new (std::nothrow) b200_key();
    if (!k) return fail("out of host memory");
    DeviceKey &d = k->dk;
    d.n = p->n; d.N = p->N; d.k = p->k; d.l = p->l; d.Bgbit = p->Bgbit;
    d.ks_t = p->ks_t; d.ks_basebit = p->ks_basebit; d.ks_variance = p->ks_variance;
    d.ks_stride = (p->n + 1 + 3) & ~3;
    d.device = device;
    const int kpl = (d.k + 1) * d.l, base = 1 << d.ks_basebit;
    // the evaluation-domain key is one blob (one broadcast replicates it); an experiments build appends the
    // layout of the warp-per-ciphertext transform
    const size_t bk_elems = (size_t)d.n * kpl * (d.k + 1) * (d.N / 2);
    d.bkfft_bytes = (experiments_built() ? 3 : 2) * bk_elems * sizeof(double2);
    d.ks_bytes = (size_t)d.N * d.k * d.ks_t * (base - 1) * d.ks_stride * sizeof(int32_t);
    cudaError_t e = cudaMalloc((void **)&d.bkfft, d.bkfft_bytes);
    if (e == cudaSuccess) { d.bkfft_tm = d.bkfft + bk_elems; d.bkfft16 = d.bkfft + 2 * bk_elems; }
    if (e == cudaSuccess) e = cudaMalloc((void **)&d.ks, d.ks_bytes);
    if (e == cudaSuccess) e = key_init_runtime(k);
    if (e == cudaSuccess) e = upload_twiddles();
    if (e != cudaSuccess) {
        fail_cuda("key allocation", e);
        b200_key_free(k);
        return 1;
    }
    *out = k;
    return 0;
}

extern "C" int b200_key_create_empty(const b200_params *p, b200_key **out) {
    if (key_alloc(p, out)) return 1;
    (*out)->filled = true;  // caller fills the blobs (broadcast) before use
    return 0;
}

extern "C" int b200_key_create(const b200_params *p, const int32_t *bk_coefs, const int32_t *ks_rows, b200_key **out) {
    if (!bk_coefs || !ks_rows) return fail("bk_coefs / ks_rows is NULL");
    if (key_alloc(p, out)) return 1;
    b200_key *k = *out;
    const DeviceKey &d = k->dk;
    B200_ON_DEVICE(d.device);
    const int kpl = (d.k + 1) * d.l, base = 1 << d.ks_basebit;
    const size_t bk_ints = (size_t)d.n * kpl * (d.k + 1) * d.N;
    const size_t ks_ints = (size_t)d.N * d.k * d.ks_t * base * (d.n + 1);
    DevBuf tmp;
    int rc = 0;
    do {
        if ((rc = tmp.ensure(std::max(bk_ints, ks_ints) * sizeof(int32_t)))) break;
        cudaError_t e = cudaMemcpyAsync(tmp.p, bk_coefs, bk_ints * sizeof(int32_t), cudaMemcpyHostToDevice, k->stream);
        if (e == cudaSuccess) e = launch_bk_convert(d, (const int32_t *)tmp.p, k->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(k->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(tmp.p, ks_rows, ks_ints * sizeof(int32_t), cudaMemcpyHostToDevice, k->stream);
        if (e == cudaSuccess) e = launch_ks_pack(d, (const int32_t *)tmp.p, k->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(k->stream);
        if (e != cudaSuccess) rc = fail_cuda("key upload/convert", e);
    } while (0);
    tmp.release();
    if (rc) {
        b200_key_free(k);
        *out = nullptr;
        return 1;
    }
    k->filled = true;
    return 0;
}

struct b200_multi;
extern "C" void b200_multi_free(b200_multi *m);
static std::map<b200_key *, b200_multi *> &multi_cache();
extern "C" void b200_key_free(b200_key *k) {
    if (!k) return;
    {
        b200_multi *m = nullptr;
        {
            std::lock_guard<std::mutex> g(g_cache_mu);
            auto &mc = multi_cache();
            auto it = mc.find(k);
            if (it != mc.end()) { m = it->second; mc.erase(it); }
        }
        if (m) b200_multi_free(m);
    }
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        for (auto *m : {&g_by_cloud, &g_by_bkfft, &g_by_ks})
            for (auto it = m->begin(); it != m->end();) it = (it->second == k) ? m->erase(it) : std::next(it);
    }
    for (int i = 1; i < b200_key::kMaxLanes; ++i)
        if (k->lanes[i]) { b200_key_free(k->lanes[i]); k->lanes[i] = nullptr; }
    DeviceGuard dev_guard(k->dk.device);
    cudaDeviceSynchronize();            // calls enqueued on caller streams may still be using the workspaces
    if (k->owns_blobs && k->dk.bkfft) cudaFree(k->dk.bkfft);
    if (k->owns_blobs && k->dk.ks) cudaFree(k->dk.ks);
    for (DevBuf *b : {&k->u, &k->in0, &k->in1, &k->in2, &k->out, &k->nz, &k->desc_rot, &k->desc_ks, &k->desc_lin, &k->desc_io, &k->acc}) b->release();
    k->stage_in.release();
    k->stage_out.release();
    for (auto &slot : k->ring)
        for (auto &e : slot)
            if (e) cudaEventDestroy(e);
    if (k->ev_start) cudaEventDestroy(k->ev_start);
    if (k->ev_scratch) cudaEventDestroy(k->ev_scratch);
    if (k->stream) cudaStreamDestroy(k->stream);
    if (k->stream_h2d) cudaStreamDestroy(k->stream_h2d);
    if (k->stream_d2h) cudaStreamDestroy(k->stream_d2h);
    for (int i = 0; i < b200_key::kSlices; ++i) {
        if (k->ev_up[i]) cudaEventDestroy(k->ev_up[i]);
        if (k->ev_done[i]) cudaEventDestroy(k->ev_done[i]);
    }
    delete k;
}

extern "C" int b200_key_params(const b200_key *k, b200_params *out) {
    if (!k || !out) return fail("NULL argument");
    out->n = k->dk.n; out->N = k->dk.N; out->k = k->dk.k; out->l = k->dk.l; out->Bgbit = k->dk.Bgbit;
    out->ks_t = k->dk.ks_t; out->ks_basebit = k->dk.ks_basebit; out->ks_variance = k->dk.ks_variance;
    return 0;
}

extern "C" int b200_key_device_blobs(const b200_key *k, void **d_bkfft, size_t *bkfft_bytes, void **d_ks, size_t *ks_bytes) {
    if (!k) return fail("key is NULL");
    if (d_bkfft) *d_bkfft = k->dk.bkfft;
    if (bkfft_bytes) *bkfft_bytes = k->dk.bkfft_bytes;
    if (d_ks) *d_ks = k->dk.ks;
    if (ks_bytes) *ks_bytes = k->dk.ks_bytes;
    return 0;
}

// ---- from a live libtfhe cloud key ---------------------------------------------------------------
// The drop-in cache is keyed by host pointers, and malloc may hand the address of a deleted key set to the next
// one.  Two defences: (1) delete_gate_bootstrapping_cloud_keyset / delete_LweBootstrappingKeyFFT are interposed
// below and drop the cache entry; (2) a hit is trusted only if a fingerprint of the key material (parameters and
// a sample of bootstrapping-key and key-switch coefficients) still matches -- otherwise the device key is freed
// and the new key uploaded.
static uint64_t fnv1a(uint64_t h, const void *p, size_t bytes) {
    const unsigned char *c = (const unsigned char *)p;
    for (size_t i = 0; i < bytes; ++i) { h ^= c[i]; h *= 1099511628211ull; }
    return h;
}
static uint64_t cloudkey_fingerprint(const TFheGateBootstrappingCloudKeySet *ck) {
    const TFheGateBootstrappingParameterSet *ps = ck->params;
    const TGswParams *gp = ps->tgsw_params;
    const TLweParams *tp = gp->tlwe_params;
    const LweKeySwitchKey *ks = (ck->bkFFT && ck->bkFFT->ks) ? ck->bkFFT->ks : ck->bk->ks;
    const int32_t shape[7] = {ps->in_out_params->n, tp->N, tp->k, gp->l, gp->Bgbit, ps->ks_t, ps->ks_basebit};
    uint64_t h = fnv1a(14695981039346656037ull, shape, sizeof shape);
    const int n = shape[0], N = shape[1], kpl = (shape[2] + 1) * shape[3];
    if (n > 0 && N >= 16 && kpl > 0 && ck->bk->bk) {
        const int words = 16;
        for (int i : {0, n / 2, n - 1})
            for (int r : {0, kpl - 1}) {
                const TLweSample &row = ck->bk->bk[i].all_sample[r];
                h = fnv1a(h, row.a[0].coefsT, words * sizeof(int32_t));
                h = fnv1a(h, row.a[shape[2]].coefsT + N - words, words * sizeof(int32_t));
            }
    }
    if (ks && ks->ks && ks->n > 0 && ks->t > 0 && n > 0) {
        const int base = 1 << ks->basebit, w = std::min(8, n);
        for (int i : {0, ks->n - 1}) {
            const LweSample &a = ks->ks[i][0][1], &b = ks->ks[i][ks->t - 1][base - 1];
            h = fnv1a(h, a.a, w * sizeof(int32_t));
            h = fnv1a(h, &a.b, sizeof a.b);
            h = fnv1a(h, b.a + (n - w), w * sizeof(int32_t));
            h = fnv1a(h, &b.b, sizeof b.b);
        }
    }
    return h ? h : 1;
}

extern "C" int b200_cloudkey_upload(const TFheGateBootstrappingCloudKeySet *ck, b200_key **out) {
    return b200_guarded([&]() -> int {
        if (!ck || !ck->params || !ck->bk) return fail("cloud key (or its params / bk) is NULL");
        const uint64_t fp = cloudkey_fingerprint(ck);
        {
            b200_key *stale = nullptr;
            {
                std::lock_guard<std::mutex> g(g_cache_mu);
                auto it = g_by_cloud.find(ck);
                if (it != g_by_cloud.end()) {
                    if (it->second->fingerprint == fp) {
                        if (out) *out = it->second;
                        return 0;
                    }
                    stale = it->second;   // same address, different key: the old key set was deleted behind our back
                }
            }
            if (stale) b200_key_free(stale);
        }
        const TFheGateBootstrappingParameterSet *ps = ck->params;
        const TGswParams *gp = ps->tgsw_params;
        const TLweParams *tp = gp->tlwe_params;
        const LweKeySwitchKey *ks = (ck->bkFFT && ck->bkFFT->ks) ? ck->bkFFT->ks : ck->bk->ks;
        if (!ks) return fail("cloud key has no key-switch key");
        b200_params p;
        p.n = ps->in_out_params->n; p.N = tp->N; p.k = tp->k; p.l = gp->l; p.Bgbit = gp->Bgbit;
        p.ks_t = ps->ks_t; p.ks_basebit = ps->ks_basebit;
        if (check_params(&p)) return 1;
        if (ks->n != p.N * p.k || ks->t != p.ks_t || ks->basebit != p.ks_basebit) return fail("key-switch key shape mismatch");
        const int kpl = (p.k + 1) * p.l, base = 1 << p.ks_basebit, n1 = p.n + 1;
        p.ks_variance = ks->ks[0][0][1].current_variance;
        std::vector<int32_t> bk((size_t)p.n * kpl * (p.k + 1) * p.N), rows((size_t)ks->n * p.ks_t * base * n1);
        for (int i = 0; i < p.n; ++i)
            for (int r = 0; r < kpl; ++r)
                for (int c = 0; c <= p.k; ++c)
                    std::memcpy(&bk[(((size_t)i * kpl + r) * (p.k + 1) + c) * p.N], ck->bk->bk[i].all_sample[r].a[c].coefsT,
                                p.N * sizeof(int32_t));
        for (int i = 0; i < ks->n; ++i)
            for (int j = 0; j < p.ks_t; ++j)
                for (int h = 0; h < base; ++h) {
                    int32_t *dst = &rows[(((size_t)i * p.ks_t + j) * base + h) * n1];
                    const LweSample &s = ks->ks[i][j][h];
                    std::memcpy(dst, s.a, p.n * sizeof(int32_t));
                    dst[p.n] = s.b;
                }
        b200_key *k = nullptr;
        if (b200_key_create(&p, bk.data(), rows.data(), &k)) return 1;
        {
            std::lock_guard<std::mutex> g(g_cache_mu);
            k->fingerprint = fp;
            g_by_cloud[ck] = k;
            if (ck->bkFFT) g_by_bkfft[ck->bkFFT] = k;
            g_by_ks[ks] = k;
            if (ck->bk->ks) g_by_ks[ck->bk->ks] = k;
        }
        if (out) *out = k;
        return 0;
    });
}

extern "C" void b200_cloudkey_forget(const TFheGateBootstrappingCloudKeySet *ck) {
    b200_key *k = nullptr;
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        auto it = g_by_cloud.find(ck);
        if (it == g_by_cloud.end()) return;
        k = it->second;
    }
    b200_key_free(k);
}

// libtfhe's destructors, interposed (this library is linked / preloaded BEFORE libtfhe):
// forget the device key, then run libtfhe's own function (the next definition in link order).
extern "C" void delete_gate_bootstrapping_cloud_keyset(TFheGateBootstrappingCloudKeySet *keyset) {
    using fn_t = void (*)(TFheGateBootstrappingCloudKeySet *);
    static fn_t next = (fn_t)dlsym(RTLD_NEXT, "delete_gate_bootstrapping_cloud_keyset");
    if (keyset) b200_cloudkey_forget(keyset);
    if (next) next(keyset);
    else std::fprintf(stderr, "libb200tfhe: delete_gate_bootstrapping_cloud_keyset: libtfhe's definition not found after this library; host key set leaked\n");
}
extern "C" void delete_LweBootstrappingKeyFFT(LweBootstrappingKeyFFT *obj) {
    using fn_t = void (*)(LweBootstrappingKeyFFT *);
    static fn_t next = (fn_t)dlsym(RTLD_NEXT, "delete_LweBootstrappingKeyFFT");
    if (obj) {
        b200_key *k = nullptr;
        {
            std::lock_guard<std::mutex> g(g_cache_mu);
            auto it = g_by_bkfft.find(obj);
            if (it != g_by_bkfft.end()) k = it->second;
        }
        if (k) b200_key_free(k);
    }
    if (next) next(obj);
    else std::fprintf(stderr, "libb200tfhe: delete_LweBootstrappingKeyFFT: libtfhe's definition not found after this library; host key leaked\n");
}

// ------------------------------------------------------------------------------------------------
// gate batches
// ------------------------------------------------------------------------------------------------
static bool op_is_linear(int op) { return op == B200_NOT || op == B200_COPY || op == B200_CONST0 || op == B200_CONST1; }
static bool op_valid(int op) { return op >= 0 && op < B200_GATE_OP_COUNT; }
static constexpr int64_t kChunkItems = 1 << 18;  // rotations per launch: bounds the extracted-sample workspace to ~1 GiB

// enqueue: out[count] = op(a, b, c) on device pointers. Caller holds key->mu.
static int enqueue_gates(b200_key *k, int op, int32_t *d_out, const int32_t *d_a, const int32_t *d_b, const int32_t *d_c,
                         int64_t count, int32_t *d_nonzero, cudaStream_t s, bool timed) {
    const DeviceKey &d = k->dk;
    const int n1 = d.n + 1;
    if (op_is_linear(op)) {
        const int sign = op == B200_NOT ? -1 : 1;
        const int32_t cst = op == B200_CONST1 ? (1 << 29) : (op == B200_CONST0 ? -(1 << 29) : 0);
        const int32_t *in = (op == B200_NOT || op == B200_COPY) ? d_a : nullptr;
        if ((op == B200_NOT || op == B200_COPY) && !d_a) return fail("input a is NULL");
        CK(launch_linear(d, d_out, in, sign, cst, count, s));
        return 0;
    }
    if (!d_a || !d_b || (op == B200_MUX && !d_c)) return fail("gate input pointer is NULL");
    const int per = op == B200_MUX ? 2 : 1;
    const int64_t chunk_gates = kChunkItems / per;
    if (k->u.ensure((size_t)std::min<int64_t>(count, chunk_gates) * per * (d.N + 1) * sizeof(int32_t))) return 1;
    if (k->scratch_begin(s)) return 1;
    for (int64_t g0 = 0; g0 < count; g0 += chunk_gates) {
        const int64_t g = std::min(chunk_gates, count - g0);
        BlindRotateArgs br{};
        br.in0 = d_a + g0 * n1;
        br.in1 = d_b + g0 * n1;
        br.in2 = d_c ? d_c + g0 * n1 : nullptr;
        br.op = op;
        br.mu = 1 << 29;
        br.u = (int32_t *)k->u.p;
        br.items = g * per;
        const bool t = timed && g0 == 0;
        cudaEvent_t *ev = k->ring[k->ring_head % b200_key::kRing];
        if (t) CK(cudaEventRecord(ev[0], s));
        CK(launch_blind_rotate(d, br, s));
        if (t) CK(cudaEventRecord(ev[1], s));
        KeySwitchArgs ks{};
        ks.u = (const int32_t *)k->u.p;
        ks.pair = per == 2;
        ks.out = d_out + g0 * n1;
        ks.nonzero = d_nonzero ? d_nonzero + g0 : nullptr;
        ks.items = g;
        CK(launch_keyswitch(d, ks, s));
        if (t) {
            CK(cudaEventRecord(ev[2], s));
            k->ring_head++;
        }
    }
    return k->scratch_end(s);
}

extern "C" int b200_gate_batch_dev(b200_key *k, int op, int32_t *d_out, const int32_t *d_a, const int32_t *d_b,
                                   const int32_t *d_c, int64_t count, void *stream) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (!op_valid(op)) return fail("invalid gate op");
    if (count < 0) return fail("negative count");
    if (count == 0) return 0;
    if (!d_out) return fail("d_out is NULL");
    LaneGuard lane(k);
    k = lane.k;
    B200_ON_DEVICE(k->dk.device);
    k->timing_pending = true;
    return enqueue_gates(k, op, d_out, d_a, d_b, d_c, count, nullptr, (cudaStream_t)stream, true);
}

// Same gates as b200_gate_batch_dev with no allocation anywhere on the call: the extracted samples go through the
// static device array of kernels.cu (launch_gates_noalloc), the key's own mutex serialises the enqueue (no lane is
// created: that would allocate), and no timing events are recorded.
extern "C" int b200_gate_batch_dev_noalloc(b200_key *k, int op, int32_t *d_out, const int32_t *d_a, const int32_t *d_b,
                                           const int32_t *d_c, int64_t count, void *stream) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (!op_valid(op)) return fail("invalid gate op");
    if (count < 0) return fail("negative count");
    if (count == 0) return 0;
    if (!d_out) return fail("d_out is NULL");
    const DeviceKey &d = k->dk;
    static_assert(kNoAllocMaxRotations == B200_NOALLOC_MAX_ROTATIONS, "header and kernels.cu disagree on the scratch size");
    const int per = op == B200_MUX ? 2 : 1;
    if (!op_is_linear(op) && count * per > kNoAllocMaxRotations) {
        char msg[256];
        snprintf(msg, sizeof msg, "b200_gate_batch_dev_noalloc: %lld gates need %lld blind rotations, the static scratch holds %lld "
                 "(B200_NOALLOC_MAX_ROTATIONS); use b200_gate_batch_dev or split the batch",
                 (long long)count, (long long)(count * per), (long long)kNoAllocMaxRotations);
        return fail(msg);
    }
    std::lock_guard<std::mutex> g(k->mu);
    B200_ON_DEVICE(d.device);
    cudaStream_t s = (cudaStream_t)stream;
    if (op_is_linear(op)) {          // no scratch involved
        const int sign = op == B200_NOT ? -1 : 1;
        const int32_t cst = op == B200_CONST1 ? (1 << 29) : (op == B200_CONST0 ? -(1 << 29) : 0);
        const int32_t *in = (op == B200_NOT || op == B200_COPY) ? d_a : nullptr;
        if ((op == B200_NOT || op == B200_COPY) && !d_a) return fail("input a is NULL");
        CK(launch_linear(d, d_out, in, sign, cst, count, s));
        return 0;
    }
    if (!d_a || !d_b || (op == B200_MUX && !d_c)) return fail("gate input pointer is NULL");
    if (d.N * d.k != 1024) return fail("b200_gate_batch_dev_noalloc: the static scratch is laid out for N * k = 1024");
    CK(launch_gates_noalloc(d, op, d_out, d_a, d_b, d_c, count, s));
    return 0;
}

static int gate_batch_host_locked(b200_key *k, int op, int32_t *out, const int32_t *a, const int32_t *b, const int32_t *c,
                                  int64_t count, int32_t *nonzero) {
    const DeviceKey &d = k->dk;
    const size_t bytes = (size_t)count * (d.n + 1) * sizeof(int32_t);
    cudaStream_t s = k->stream;
    B200_ON_DEVICE(d.device);
    const bool need_a = !(op == B200_CONST0 || op == B200_CONST1);
    const bool need_b = !op_is_linear(op);
    const bool need_c = op == B200_MUX;
    if ((need_a && !a) || (need_b && !b) || (need_c && !c) || !out) return fail("NULL batch pointer");
    CK(cudaEventRecord(k->ev_start, s));
    if (k->out.ensure(bytes)) return 1;
    // Large bootstrapped batches: slice the batch so that the upload of slice i+1 and the download of slice
    // i-1 overlap the kernels of slice i (three streams).  A slice is a whole number of blind-rotation waves
    // (148 SMs x 6 ciphertexts) and of 64-sample key-switch tiles.  Snapshot semantics hold: every input
    // byte is read from the caller's buffers into the device staging copies, never from `out`.
    // whole waves of the 6-ciphertext blind-rotation CTAs (888) and, for big batches, of the 64-sample key-switch
    // CTAs as well (148 x lcm(6, 64) = 28,416); smaller batches settle for whole 64-sample tiles (7,104)
    constexpr int64_t kSliceSmall = 148 * 6 * 8, kSliceBig = 148 * 192;
    const int64_t kSliceQuantum = count >= 2 * kSliceBig ? kSliceBig : kSliceSmall;
    // (an output that overlaps an input at a different offset would break that, so such calls are not sliced)
    auto shifted_overlap = [&](const int32_t *in) {
        if (!in || in == out) return false;
        const char *i0 = (const char *)in, *o0 = (const char *)out;
        return i0 < o0 + bytes && o0 < i0 + bytes;
    };
    if (!op_is_linear(op) && count >= 2 * kSliceSmall && !shifted_overlap(a) && !shifted_overlap(b) &&
        !shifted_overlap(need_c ? c : nullptr)) {
        const int n1 = d.n + 1;
        int32_t *d_nz_all = nullptr;      // per-gate count of non-zero key-switch digits (-> current_variance)
        if (nonzero) {
            if (k->nz.ensure((size_t)count * sizeof(int32_t))) return 1;
            d_nz_all = (int32_t *)k->nz.p;
            CK(cudaMemsetAsync(d_nz_all, 0, (size_t)count * sizeof(int32_t), s));
        }
        int64_t slice = std::max<int64_t>(1, (count / 4 + kSliceQuantum / 2) / kSliceQuantum) * kSliceQuantum;   // about a quarter, at least one quantum
        while ((count + slice - 1) / slice > b200_key::kSlices) slice += kSliceQuantum;
        if (k->in0.ensure(bytes) || k->in1.ensure(bytes) || (need_c && k->in2.ensure(bytes))) return 1;
        CK(cudaEventRecord(k->ev_up[0], s));                 // order after earlier work on the key's stream
        CK(cudaStreamWaitEvent(k->stream_h2d, k->ev_up[0], 0));
        CK(cudaStreamWaitEvent(k->stream_d2h, k->ev_up[0], 0));
        // one slice: upload on the copy stream, kernels on the key's stream, download on the third stream
        auto run_slice = [&](int i, int64_t g0) -> int {
            const int64_t g = std::min(slice, count - g0);
            const size_t off = (size_t)g0 * n1, sb = (size_t)g * n1 * sizeof(int32_t);
            CK(cudaMemcpyAsync((int32_t *)k->in0.p + off, a + off, sb, cudaMemcpyHostToDevice, k->stream_h2d));
            CK(cudaMemcpyAsync((int32_t *)k->in1.p + off, b + off, sb, cudaMemcpyHostToDevice, k->stream_h2d));
            if (need_c) CK(cudaMemcpyAsync((int32_t *)k->in2.p + off, c + off, sb, cudaMemcpyHostToDevice, k->stream_h2d));
            CK(cudaEventRecord(k->ev_up[i], k->stream_h2d));
            CK(cudaStreamWaitEvent(s, k->ev_up[i], 0));
            if (enqueue_gates(k, op, (int32_t *)k->out.p + off, (const int32_t *)k->in0.p + off, (const int32_t *)k->in1.p + off,
                              need_c ? (const int32_t *)k->in2.p + off : nullptr, g, d_nz_all ? d_nz_all + g0 : nullptr, s, i == 0))
                return 1;
            CK(cudaEventRecord(k->ev_done[i], s));
            CK(cudaStreamWaitEvent(k->stream_d2h, k->ev_done[i], 0));
            CK(cudaMemcpyAsync(out + off, (int32_t *)k->out.p + off, sb, cudaMemcpyDeviceToHost, k->stream_d2h));
            if (d_nz_all) CK(cudaMemcpyAsync(nonzero + g0, d_nz_all + g0, (size_t)g * sizeof(int32_t), cudaMemcpyDeviceToHost, k->stream_d2h));
            return 0;
        };
        int i = 0, rc = 0;
        for (int64_t g0 = 0; g0 < count && !rc; g0 += slice, ++i) rc = run_slice(i, g0);
        // Also on failure: copies that read the caller's a / b / c and write the caller's out may still be in
        // flight on the copy streams; the caller may free those buffers as soon as this call returns.
        const cudaError_t e1 = cudaStreamSynchronize(k->stream_h2d), e2 = cudaStreamSynchronize(s),
                          e3 = cudaStreamSynchronize(k->stream_d2h);
        if (rc) return 1;
        if (e1 != cudaSuccess) return fail_cuda("host batch pipeline (upload stream)", e1);
        if (e2 != cudaSuccess) return fail_cuda("host batch pipeline (kernel stream)", e2);
        if (e3 != cudaSuccess) return fail_cuda("host batch pipeline (download stream)", e3);
        k->timing_pending = true;
        return 0;
    }
    if (need_a) { if (k->in0.ensure(bytes)) return 1; CK(cudaMemcpyAsync(k->in0.p, a, bytes, cudaMemcpyHostToDevice, s)); }
    if (need_b) { if (k->in1.ensure(bytes)) return 1; CK(cudaMemcpyAsync(k->in1.p, b, bytes, cudaMemcpyHostToDevice, s)); }
    if (need_c) { if (k->in2.ensure(bytes)) return 1; CK(cudaMemcpyAsync(k->in2.p, c, bytes, cudaMemcpyHostToDevice, s)); }
    int32_t *d_nz = nullptr;
    if (nonzero && !op_is_linear(op)) {
        if (k->nz.ensure((size_t)count * sizeof(int32_t))) return 1;
        d_nz = (int32_t *)k->nz.p;
        CK(cudaMemsetAsync(d_nz, 0, (size_t)count * sizeof(int32_t), s));
    }
    if (enqueue_gates(k, op, (int32_t *)k->out.p, (const int32_t *)k->in0.p, (const int32_t *)k->in1.p,
                      (const int32_t *)k->in2.p, count, d_nz, s, true))
        return 1;
    CK(cudaMemcpyAsync(out, k->out.p, bytes, cudaMemcpyDeviceToHost, s));
    if (d_nz) CK(cudaMemcpyAsync(nonzero, d_nz, (size_t)count * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    else if (nonzero) std::memset(nonzero, 0, (size_t)count * sizeof(int32_t));
    CK(cudaStreamSynchronize(s));
    k->timing_pending = !op_is_linear(op);
    return 0;
}

extern "C" int b200_gate_batch_host(b200_key *k, int op, int32_t *out, const int32_t *a, const int32_t *b,
                                    const int32_t *c, int64_t count) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (!op_valid(op)) return fail("invalid gate op");
    if (count < 0) return fail("negative count");
    if (count == 0) return 0;
    LaneGuard lane(k);
    return gate_batch_host_locked(lane.k, op, out, a, b, c, count, nullptr);
}

extern "C" int b200_last_timing_ms(b200_key *k, float out_ms[3]) {
    if (!k || !out_ms) return fail("NULL argument");
    std::lock_guard<std::mutex> g(k->mu);
    B200_ON_DEVICE(k->dk.device);
    if (k->timing_pending && k->ring_head > 0) {
        cudaEvent_t *ev = k->ring[(k->ring_head - 1) % b200_key::kRing];
        CK(cudaEventSynchronize(ev[2]));
        CK(cudaEventElapsedTime(&k->last_ms[0], ev[0], ev[1]));
        CK(cudaEventElapsedTime(&k->last_ms[1], ev[1], ev[2]));
        float tot = 0;
        if (cudaEventElapsedTime(&tot, k->ev_start, ev[2]) == cudaSuccess && tot > 0) k->last_ms[2] = tot;
        else { cudaGetLastError(); k->last_ms[2] = k->last_ms[0] + k->last_ms[1]; }
        k->timing_pending = false;
    }
    out_ms[0] = k->last_ms[0]; out_ms[1] = k->last_ms[1]; out_ms[2] = k->last_ms[2];
    return 0;
}

static int timing_drain_one(b200_key *k, double sum_ms[2], int32_t *calls) {
    std::lock_guard<std::mutex> g(k->mu);
    const int64_t begin = std::max(k->ring_drained, k->ring_head - b200_key::kRing);
    for (int64_t c = begin; c < k->ring_head; ++c) {
        cudaEvent_t *ev = k->ring[c % b200_key::kRing];
        float a = 0, b = 0;
        CK(cudaEventSynchronize(ev[2]));
        CK(cudaEventElapsedTime(&a, ev[0], ev[1]));
        CK(cudaEventElapsedTime(&b, ev[1], ev[2]));
        sum_ms[0] += a;
        sum_ms[1] += b;
        ++*calls;
    }
    k->ring_drained = k->ring_head;
    return 0;
}
extern "C" int b200_timing_drain(b200_key *k, double sum_ms[2], int32_t *calls) {
    if (!k || !sum_ms || !calls) return fail("NULL argument");
    B200_ON_DEVICE(k->dk.device);
    sum_ms[0] = sum_ms[1] = 0;
    *calls = 0;
    if (timing_drain_one(k, sum_ms, calls)) return 1;
    for (int i = 1; i < b200_key::kMaxLanes; ++i) {     // calls that ran on a lane while the key was busy
        b200_key *l;
        {
            std::lock_guard<std::mutex> g(k->lanes_mu);
            l = k->lanes[i];
        }
        if (l && timing_drain_one(l, sum_ms, calls)) return 1;
    }
    return 0;
}

extern "C" int b200_fp64_peak_tflops(int device, double *tflops) {
    if (!tflops) return fail("NULL argument");
    if (b200_device_count() <= 0) return fail("no CUDA device");
    B200_ON_DEVICE(device);
    CK(measure_fp64_peak(tflops));
    return 0;
}

// ---- several devices behind one handle -------------------------------------------------------------
// The path shards by independent gates (no data-path collective): a b200_multi holds one replica of the cloud key
// per device; a host batch is cut into contiguous slices, one per device, each evaluated by its own host thread
// through the single-device entry (its own streams, workspaces and mutex).  The reference has no counterpart
// (zero collectives); this is the C-ABI form of north_star's multi-GPU split.
struct b200_multi {
    std::vector<b200_key *> keys;     // keys[0] is the caller's key (not owned), the others are replicas (owned)
};

extern "C" int b200_multi_create(b200_key *key, const int32_t *devices, int32_t ndev, b200_multi **out) {
    return b200_guarded([&]() -> int {
        if (!key || !key->filled) return fail("key is NULL or not filled");
        if (!out) return fail("out is NULL");
        *out = nullptr;
        const int have = b200_device_count();
        std::vector<int> devs;
        if (!devices) {                       // all devices, the key's own first
            devs.push_back(key->dk.device);
            for (int d = 0; d < have && (ndev <= 0 || (int)devs.size() < ndev); ++d)
                if (d != key->dk.device) devs.push_back(d);
        } else {
            if (ndev < 1) return fail("ndev must be >= 1");
            for (int i = 0; i < ndev; ++i) {
                if (devices[i] < 0 || devices[i] >= have) return fail("device index out of range");
                for (int d : devs) if (d == devices[i]) return fail("device listed twice");
                devs.push_back(devices[i]);
            }
            if (devs[0] != key->dk.device) return fail("devices[0] must be the device the key lives on");
        }
        b200_multi *m = new b200_multi();
        m->keys.push_back(key);
        b200_params p;
        b200_key_params(key, &p);
        for (size_t i = 1; i < devs.size(); ++i) {
            b200_key *r = nullptr;
            int rc = key_alloc(&p, &r, devs[i]);
            if (!rc) {
                // replicate the two resident blobs device to device (NVLink when peers, through the host otherwise)
                cudaError_t e = cudaMemcpyPeer(r->dk.bkfft, devs[i], key->dk.bkfft, key->dk.device, key->dk.bkfft_bytes);
                if (e == cudaSuccess) e = cudaMemcpyPeer(r->dk.ks, devs[i], key->dk.ks, key->dk.device, key->dk.ks_bytes);
                if (e != cudaSuccess) rc = fail_cuda("key replication", e);
                else r->filled = true;
            }
            if (rc) {
                if (r) b200_key_free(r);
                for (size_t j = 1; j < m->keys.size(); ++j) b200_key_free(m->keys[j]);
                delete m;
                return 1;
            }
            m->keys.push_back(r);
        }
        *out = m;
        return 0;
    });
}
extern "C" void b200_multi_free(b200_multi *m) {
    if (!m) return;
    for (size_t j = 1; j < m->keys.size(); ++j) b200_key_free(m->keys[j]);
    delete m;
}
extern "C" int b200_multi_device_count(const b200_multi *m) { return m ? (int)m->keys.size() : 0; }

static int gate_batch_host_locked(b200_key *k, int op, int32_t *out, const int32_t *a, const int32_t *b, const int32_t *c,
                                  int64_t count, int32_t *nonzero);
static bool op_is_linear(int op);
// slices of whole blind-rotation waves (148 SMs x 6), the remainder on the last device
static int multi_gate_batch(b200_multi *m, int op, int32_t *out, const int32_t *a, const int32_t *b, const int32_t *c,
                            int64_t count, int32_t *nonzero) {
    const int nd = (int)m->keys.size();
    const int n1 = m->keys[0]->dk.n + 1;
    const int64_t quantum = 148 * 6;
    int64_t per = (count + nd - 1) / nd;
    per = (per + quantum - 1) / quantum * quantum;
    if (nd == 1 || count <= quantum) {
        std::lock_guard<std::mutex> g(m->keys[0]->mu);
        return gate_batch_host_locked(m->keys[0], op, out, a, b, c, count, nonzero);
    }
    std::vector<int> rcs(nd, 0);
    std::vector<std::string> errs(nd);
    std::vector<std::thread> th;
    for (int d = 0; d < nd; ++d) {
        const int64_t g0 = std::min<int64_t>(count, d * per), g = std::min<int64_t>(per, count - g0);
        if (g <= 0) break;
        th.emplace_back([&, d, g0, g]() {
            const size_t off = (size_t)g0 * n1;
            b200_key *k = m->keys[d];
            std::lock_guard<std::mutex> lk(k->mu);
            rcs[d] = gate_batch_host_locked(k, op, out + off, a ? a + off : nullptr, b ? b + off : nullptr, c ? c + off : nullptr,
                                            g, nonzero ? nonzero + g0 : nullptr);
            if (rcs[d]) errs[d] = b200_last_error();    // the error message is thread local: carry it to the caller
        });
    }
    for (auto &t : th) t.join();
    for (int d = 0; d < nd; ++d)
        if (rcs[d]) return fail("device " + std::to_string(m->keys[d]->dk.device) + ": " + errs[d]);
    return 0;
}
extern "C" int b200_multi_gate_batch_host(b200_multi *m, int op, int32_t *out, const int32_t *a, const int32_t *b,
                                          const int32_t *c, int64_t count) {
    return b200_guarded([&]() -> int {
        if (!m || m->keys.empty()) return fail("multi-device handle is NULL");
        if (op < 0 || op >= B200_GATE_OP_COUNT) return fail("invalid gate op");
        if (count < 0) return fail("negative count");
        if (count == 0) return 0;
        if (!out) return fail("out is NULL");
        const char *o0 = (const char *)out;
        const size_t bytes = (size_t)count * (m->keys[0]->dk.n + 1) * sizeof(int32_t);
        for (const int32_t *in : {a, b, c}) {   // slices run concurrently: an output slice must not be another slice's input
            const char *i0 = (const char *)in;
            if (in && in != out && i0 < o0 + bytes && o0 < i0 + bytes) return fail("out overlaps an input at a different offset");
        }
        return multi_gate_batch(m, op, out, a, b, c, count, nullptr);
    });
}

extern "C" int b200_int32_add_peak(int device, double *tera_adds_per_s) {
    if (!tera_adds_per_s) return fail("NULL argument");
    if (b200_device_count() <= 0) return fail("no CUDA device");
    B200_ON_DEVICE(device);
    CK(measure_int_add_peak(tera_adds_per_s));
    return 0;
}

// ---- batched gate over libtfhe samples -----------------------------------------------------------
// B200TFHE_DEVICES = "all" | N: the drop-in batched gate (bootsGate_batch) spreads batches of at least two waves
// over that many devices; the replicas are made on first use and die with the key.
static std::map<b200_key *, b200_multi *> g_multi;
static std::map<b200_key *, b200_multi *> &multi_cache() { return g_multi; }
static b200_multi *multi_for_key(b200_key *k) {
    static const int want = [] {
        const char *e = getenv("B200TFHE_DEVICES");
        if (!e || !*e) return 1;
        if (!std::strcmp(e, "all")) return b200_device_count();
        return std::max(1, std::min(atoi(e), b200_device_count()));
    }();
    if (want <= 1) return nullptr;
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        auto it = g_multi.find(k);
        if (it != g_multi.end()) return it->second;
    }
    // (the caller holds k->mu, so nobody else creates replicas of this key now; the cache lock is not held while
    // the replicas are made because a failing replica is freed through b200_key_free, which takes it)
    b200_multi *m = nullptr;
    if (b200_multi_create(k, nullptr, want, &m)) m = nullptr;   // fall back to one device, the error stays readable
    std::lock_guard<std::mutex> g(g_cache_mu);
    g_multi[k] = m;
    return m;
}
// multi_gate_batch for a caller that already holds keys[0]->mu
static int multi_gate_batch_locked0(b200_multi *m, int op, int32_t *out, const int32_t *a, const int32_t *b, const int32_t *c,
                                    int64_t count, int32_t *nonzero) {
    const int nd = (int)m->keys.size();
    const int n1 = m->keys[0]->dk.n + 1;
    const int64_t quantum = 148 * 6;
    int64_t per = (count + nd - 1) / nd;
    per = (per + quantum - 1) / quantum * quantum;
    if (count < 2 * quantum || op_is_linear(op)) return gate_batch_host_locked(m->keys[0], op, out, a, b, c, count, nonzero);
    std::vector<int> rcs(nd, 0);
    std::vector<std::string> errs(nd);
    std::vector<std::thread> th;
    for (int d = 1; d < nd; ++d) {
        const int64_t g0 = std::min<int64_t>(count, d * per), g = std::min<int64_t>(per, count - g0);
        if (g <= 0) break;
        th.emplace_back([&, d, g0, g]() {
            const size_t off = (size_t)g0 * n1;
            b200_key *k = m->keys[d];
            std::lock_guard<std::mutex> lk(k->mu);
            rcs[d] = gate_batch_host_locked(k, op, out + off, a ? a + off : nullptr, b ? b + off : nullptr, c ? c + off : nullptr,
                                            g, nonzero ? nonzero + g0 : nullptr);
            if (rcs[d]) errs[d] = b200_last_error();
        });
    }
    rcs[0] = gate_batch_host_locked(m->keys[0], op, out, a, b, c, std::min(per, count), nonzero);
    if (rcs[0]) errs[0] = b200_last_error();
    for (auto &t : th) t.join();
    for (int d = 0; d < nd; ++d)
        if (rcs[d]) return fail("device " + std::to_string(m->keys[d]->dk.device) + ": " + errs[d]);
    return 0;
}
static int gate_batch_samples(b200_key *k, int op, LweSample *const *out, const LweSample *const *a,
                              const LweSample *const *b, const LweSample *const *c, int32_t count) {
    const int n = k->dk.n, n1 = n + 1;
    const bool need_a = !(op == B200_CONST0 || op == B200_CONST1);
    const bool need_b = !op_is_linear(op), need_c = op == B200_MUX;
    if (!out || (need_a && !a) || (need_b && !b) || (need_c && !c)) return fail("NULL sample array");
    b200_key *const master = k;
    LaneGuard lane(k);
    k = lane.k;
    const size_t row = (size_t)n1 * sizeof(int32_t), bytes = row * count;
    if (k->stage_in.ensure(3 * bytes) || k->stage_out.ensure(bytes + (size_t)count * sizeof(int32_t))) return 1;
    int32_t *ha = (int32_t *)k->stage_in.p, *hb = ha + (size_t)count * n1, *hc = hb + (size_t)count * n1;
    int32_t *ho = (int32_t *)k->stage_out.p, *hnz = ho + (size_t)count * n1;
    auto gather = [&](int32_t *dst, const LweSample *const *src) {
        for (int32_t i = 0; i < count; ++i) {
            std::memcpy(dst + (size_t)i * n1, src[i]->a, n * sizeof(int32_t));
            dst[(size_t)i * n1 + n] = src[i]->b;
        }
    };
    // snapshot semantics: all inputs are staged before any output is written
    if (need_a) gather(ha, a);
    if (need_b) gather(hb, b);
    if (need_c) gather(hc, c);
    if (b200_multi *m = (k == master ? multi_for_key(master) : nullptr)) {   // replicas hang off the key itself, not its lanes
        // B200TFHE_DEVICES: the staged batch is cut over the replicas (this key's mutex is held: the replicas'
        // mutexes are taken by their own threads, device 0's slice by re-entering with the lock already ours)
        if (multi_gate_batch_locked0(m, op, ho, ha, hb, hc, count, hnz)) return 1;
    } else if (gate_batch_host_locked(k, op, ho, ha, hb, hc, count, hnz)) return 1;
    for (int32_t i = 0; i < count; ++i) {
        std::memcpy(out[i]->a, ho + (size_t)i * n1, n * sizeof(int32_t));
        out[i]->b = ho[(size_t)i * n1 + n];
        if (op_is_linear(op)) {
            // lweNegate / lweCopy keep the variance, lweNoiselessTrivial sets 0 (lwe-functions.cpp:120-135)
            out[i]->current_variance = need_a ? a[i]->current_variance : 0.0;
        } else {
            out[i]->current_variance = hnz[i] * k->dk.ks_variance;
        }
    }
    return 0;
}

static b200_key *key_for_cloud(const TFheGateBootstrappingCloudKeySet *ck) {
    b200_key *k = nullptr;
    if (b200_cloudkey_upload(ck, &k)) return nullptr;
    return k;
}

extern "C" int bootsGate_batch(int op, LweSample *const *out, const LweSample *const *a, const LweSample *const *b,
                               const LweSample *const *c, int32_t count, const TFheGateBootstrappingCloudKeySet *bk) {
    return b200_guarded([&]() -> int {
        if (!op_valid(op)) return fail("invalid gate op");
        if (count < 0) return fail("negative count");
        if (count == 0) return 0;
        b200_key *k = key_for_cloud(bk);
        if (!k) return 1;
        return gate_batch_samples(k, op, out, a, b, c, count);
    });
}

// ---- libtfhe drop-in gates (batch of one: the latency path) ----------------------------------------
static void gate1(const char *name, int op, LweSample *r, const LweSample *a, const LweSample *b, const LweSample *c,
                  const TFheGateBootstrappingCloudKeySet *bk) {
    LweSample *outs[1] = {r};
    const LweSample *as[1] = {a}, *bs[1] = {b}, *cs[1] = {c};
    if (bootsGate_batch(op, outs, a ? as : nullptr, b ? bs : nullptr, c ? cs : nullptr, 1, bk)) die(name);
}
#define BINARY_GATE(NAME, OP)                                                                                     \
    extern "C" void NAME(LweSample *r, const LweSample *a, const LweSample *b, const TFheGateBootstrappingCloudKeySet *bk) { \
        gate1(#NAME, OP, r, a, b, nullptr, bk);                                                                   \
    }
BINARY_GATE(bootsNAND, B200_NAND)
BINARY_GATE(bootsOR, B200_OR)
BINARY_GATE(bootsAND, B200_AND)
BINARY_GATE(bootsXOR, B200_XOR)
BINARY_GATE(bootsXNOR, B200_XNOR)
BINARY_GATE(bootsNOR, B200_NOR)
BINARY_GATE(bootsANDNY, B200_ANDNY)
BINARY_GATE(bootsANDYN, B200_ANDYN)
BINARY_GATE(bootsORNY, B200_ORNY)
BINARY_GATE(bootsORYN, B200_ORYN)
extern "C" void bootsMUX(LweSample *r, const LweSample *a, const LweSample *b, const LweSample *c,
                         const TFheGateBootstrappingCloudKeySet *bk) {
    gate1("bootsMUX", B200_MUX, r, a, b, c, bk);
}
// The linear gates never touch the device in the reference either (boot-gates.cpp:164-189); they are
// n+1 integer moves on caller memory, done in place here so that result may alias ca.
extern "C" void bootsNOT(LweSample *r, const LweSample *a, const TFheGateBootstrappingCloudKeySet *bk) {
    const int n = bk->params->in_out_params->n;
    for (int i = 0; i < n; ++i) r->a[i] = -a->a[i];
    r->b = -a->b;
    r->current_variance = a->current_variance;
}
extern "C" void bootsCOPY(LweSample *r, const LweSample *a, const TFheGateBootstrappingCloudKeySet *bk) {
    const int n = bk->params->in_out_params->n;
    for (int i = 0; i < n; ++i) r->a[i] = a->a[i];
    r->b = a->b;
    r->current_variance = a->current_variance;
}
extern "C" void bootsCONSTANT(LweSample *r, int32_t value, const TFheGateBootstrappingCloudKeySet *bk) {
    const int n = bk->params->in_out_params->n;
    for (int i = 0; i < n; ++i) r->a[i] = 0;
    r->b = value ? (1 << 29) : -(1 << 29);
    r->current_variance = 0.0;
}

// ------------------------------------------------------------------------------------------------
// pieces of the path
// ------------------------------------------------------------------------------------------------
static int woks_locked(b200_key *k, int32_t *u, const int32_t *x, int32_t mu, int64_t count) {
    const DeviceKey &d = k->dk;
    cudaStream_t s = k->stream;
    B200_ON_DEVICE(d.device);
    const size_t in_bytes = (size_t)count * (d.n + 1) * 4, u_bytes = (size_t)count * (d.N + 1) * 4;
    if (k->in0.ensure(in_bytes) || k->u.ensure(u_bytes)) return 1;
    if (k->scratch_begin(s)) return 1;
    CK(cudaMemcpyAsync(k->in0.p, x, in_bytes, cudaMemcpyHostToDevice, s));
    BlindRotateArgs br{};
    br.in0 = (const int32_t *)k->in0.p;
    br.op = kOpRaw;
    br.mu = mu;
    br.u = (int32_t *)k->u.p;
    br.items = count;
    CK(launch_blind_rotate(d, br, s));
    CK(cudaMemcpyAsync(u, k->u.p, u_bytes, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}
extern "C" int b200_bootstrap_woks_host(b200_key *k, int32_t *u, const int32_t *x, int32_t mu, int64_t count) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (count < 0 || count > kChunkItems) return fail("count out of range");
    if (count == 0) return 0;
    if (!u || !x) return fail("NULL batch pointer");
    LaneGuard lane(k);
    return woks_locked(lane.k, u, x, mu, count);
}

extern "C" int b200_keyswitch_dev(b200_key *k, int32_t *d_out, const int32_t *d_u, int64_t count, void *stream) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (count < 0) return fail("negative count");
    if (count == 0) return 0;
    if (!d_out || !d_u) return fail("NULL device pointer");
    LaneGuard lane(k);
    k = lane.k;
    B200_ON_DEVICE(k->dk.device);
    KeySwitchArgs ks{};
    ks.u = d_u;
    ks.out = d_out;
    ks.items = count;
    CK(launch_keyswitch(k->dk, ks, (cudaStream_t)stream));
    return 0;
}

static int keyswitch_locked(b200_key *k, int32_t *out, const int32_t *u, int32_t *nonzero, int64_t count) {
    const DeviceKey &d = k->dk;
    cudaStream_t s = k->stream;
    B200_ON_DEVICE(d.device);
    const size_t out_bytes = (size_t)count * (d.n + 1) * 4, u_bytes = (size_t)count * (d.N + 1) * 4;
    if (k->u.ensure(u_bytes) || k->out.ensure(out_bytes) || k->nz.ensure((size_t)count * 4)) return 1;
    if (k->scratch_begin(s)) return 1;
    CK(cudaMemcpyAsync(k->u.p, u, u_bytes, cudaMemcpyHostToDevice, s));
    CK(cudaMemsetAsync(k->nz.p, 0, (size_t)count * 4, s));
    KeySwitchArgs ks{};
    ks.u = (const int32_t *)k->u.p;
    ks.out = (int32_t *)k->out.p;
    ks.nonzero = (int32_t *)k->nz.p;
    ks.items = count;
    CK(launch_keyswitch(d, ks, s));
    CK(cudaMemcpyAsync(out, k->out.p, out_bytes, cudaMemcpyDeviceToHost, s));
    if (nonzero) CK(cudaMemcpyAsync(nonzero, k->nz.p, (size_t)count * 4, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}
extern "C" int b200_keyswitch_host(b200_key *k, int32_t *out, const int32_t *u, int32_t *nonzero, int64_t count) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (count < 0 || count > kChunkItems) return fail("count out of range");
    if (count == 0) return 0;
    if (!out || !u) return fail("NULL batch pointer");
    LaneGuard lane(k);
    return keyswitch_locked(lane.k, out, u, nonzero, count);
}

static int cmux_host(b200_key *k, int32_t *acc, int32_t bk_index, int32_t barai, int mode, int64_t count) {
    if (!k || !k->filled) return fail("key is NULL or not filled");
    if (count < 0) return fail("negative count");
    if (count == 0) return 0;
    if (!acc) return fail("acc is NULL");
    if (bk_index < 0 || bk_index >= k->dk.n) return fail("bk_index out of range");
    if (barai < 0 || barai >= 2 * k->dk.N) return fail("barai out of range");
    LaneGuard lane(k);
    k = lane.k;
    const DeviceKey &d = k->dk;
    cudaStream_t s = k->stream;
    B200_ON_DEVICE(d.device);
    const size_t bytes = (size_t)count * (d.k + 1) * d.N * 4;
    if (k->acc.ensure(bytes)) return 1;
    CK(cudaMemcpyAsync(k->acc.p, acc, bytes, cudaMemcpyHostToDevice, s));
    CK(launch_cmux_test(d, (int32_t *)k->acc.p, bk_index, barai, mode, count, s));
    CK(cudaMemcpyAsync(acc, k->acc.p, bytes, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}
extern "C" int b200_cmux_step_host(b200_key *k, int32_t *acc, int32_t bk_index, int32_t barai, int64_t count) {
    return cmux_host(k, acc, bk_index, barai, 0, count);
}
extern "C" int b200_extern_product_host(b200_key *k, int32_t *acc, int32_t bk_index, int64_t count) {
    return cmux_host(k, acc, bk_index, 0, 1, count);
}

// ---- the transform alone (parity tests of the FFT processor replacement) -------------------------------
static int fft_test_host(int32_t *out, const int32_t *a, const int32_t *b, int mode, int64_t count) {
    if (count < 0 || count > (1 << 20)) return fail("count out of range");
    if (count == 0) return 0;
    if (!out || !a || (mode == 1 && !b)) return fail("NULL polynomial pointer");
    if (b200_device_count() <= 0) return fail("no CUDA device: libb200tfhe has no CPU fallback");
    B200_ON_DEVICE(g_device);
    CK(upload_twiddles());
    const size_t bytes = (size_t)count * 1024 * sizeof(int32_t);
    DevBuf da, db, dout;
    int rc = 0;
    do {
        if ((rc = da.ensure(bytes)) || (rc = dout.ensure(bytes)) || (mode == 1 && (rc = db.ensure(bytes)))) break;
        cudaError_t e = cudaMemcpy(da.p, a, bytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && mode == 1) e = cudaMemcpy(db.p, b, bytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = launch_fft_test((const int32_t *)da.p, (const int32_t *)db.p, (int32_t *)dout.p, mode, count, nullptr);
        if (e == cudaSuccess) e = cudaMemcpy(out, dout.p, bytes, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail_cuda("transform test", e);
    } while (0);
    da.release(); db.release(); dout.release();
    return rc;
}
extern "C" int b200_fft_roundtrip_host(int32_t *out, const int32_t *a, int64_t count) { return fft_test_host(out, a, nullptr, 0, count); }
extern "C" int b200_poly_mul_host(int32_t *out, const int32_t *a, const int32_t *b, int64_t count) { return fft_test_host(out, a, b, 1, count); }

// ---- lower-level libtfhe names ----------------------------------------------------------------------
static b200_key *lookup(std::map<const void *, b200_key *> &m, const void *p) {
    std::lock_guard<std::mutex> g(g_cache_mu);
    auto it = m.find(p);
    return it == m.end() ? nullptr : it->second;
}
extern "C" void tfhe_bootstrap_woKS_FFT(LweSample *result, const LweBootstrappingKeyFFT *bk, Torus32 mu, const LweSample *x) {
    b200_key *k = lookup(g_by_bkfft, bk);
    if (!k) { fail("bootstrapping key unknown: call b200_cloudkey_upload on its cloud key first"); die("tfhe_bootstrap_woKS_FFT"); }
    const int n = k->dk.n, N = k->dk.N;
    std::vector<int32_t> xin(n + 1), u(N + 1);
    std::memcpy(xin.data(), x->a, n * 4);
    xin[n] = x->b;
    {
        LaneGuard lane(k);
        if (woks_locked(lane.k, u.data(), xin.data(), mu, 1)) die("tfhe_bootstrap_woKS_FFT");
    }
    std::memcpy(result->a, u.data(), N * 4);
    result->b = u[N];
    result->current_variance = 0.0;
}
extern "C" void lweKeySwitch(LweSample *result, const LweKeySwitchKey *ks, const LweSample *sample) {
    b200_key *k = lookup(g_by_ks, ks);
    if (!k) { fail("key-switch key unknown: call b200_cloudkey_upload on its cloud key first"); die("lweKeySwitch"); }
    const int n = k->dk.n, N = k->dk.N;
    std::vector<int32_t> u(N + 1), out(n + 1);
    int32_t nz = 0;
    std::memcpy(u.data(), sample->a, N * 4);
    u[N] = sample->b;
    {
        LaneGuard lane(k);
        if (keyswitch_locked(lane.k, out.data(), u.data(), &nz, 1)) die("lweKeySwitch");
    }
    std::memcpy(result->a, out.data(), n * 4);
    result->b = out[n];
    result->current_variance = nz * k->dk.ks_variance;
}
extern "C" void tfhe_bootstrap_FFT(LweSample *result, const LweBootstrappingKeyFFT *bk, Torus32 mu, const LweSample *x) {
    b200_key *k = lookup(g_by_bkfft, bk);
    if (!k) { fail("bootstrapping key unknown: call b200_cloudkey_upload on its cloud key first"); die("tfhe_bootstrap_FFT"); }
    const int n = k->dk.n, N = k->dk.N;
    std::vector<int32_t> xin(n + 1), u(N + 1), out(n + 1);
    int32_t nz = 0;
    std::memcpy(xin.data(), x->a, n * 4);
    xin[n] = x->b;
    {
        LaneGuard lane(k);
        if (woks_locked(lane.k, u.data(), xin.data(), mu, 1)) die("tfhe_bootstrap_FFT");
        if (keyswitch_locked(lane.k, out.data(), u.data(), &nz, 1)) die("tfhe_bootstrap_FFT");
    }
    std::memcpy(result->a, out.data(), n * 4);
    result->b = out[n];
    result->current_variance = nz * k->dk.ks_variance;
}
