// circuit.cpp -- depth-batched evaluation of gate netlists (north_star item 5).
//
// The reference evaluates circuits one gate at a time (libs/tfhe/src/test/test-addition-boot.cpp:30-115:
// full_adder, full_adder_MUX, comparison_MUX).  Here a netlist is levelised on the host and every
// bootstrapped gate of one depth level becomes one blind-rotation launch plus one key-switch launch;
// linear gates (NOT/COPY/CONSTANT, boot-gates.cpp:164-189) ride between levels.
#include "../../include/b200tfhe.h"
#include "engine.cuh"
#include "host_state.h"

#include <algorithm>
#include <cstring>
#include <vector>

using namespace b200;

static bool lin_op(int op) { return op == B200_NOT || op == B200_COPY || op == B200_CONST0 || op == B200_CONST1; }

extern "C" int b200_circuit_levelize(const int32_t *gates, int64_t num_gates, int64_t num_wires, int32_t *gate_level,
                                     int32_t *num_levels) {
    return b200_guarded([&]() -> int {
        if (!gates || num_gates < 0 || num_wires <= 0) return b200_fail("bad netlist arguments");
        std::vector<int32_t> ready(num_wires, 0);
        std::vector<uint8_t> written(num_wires, 0);
        int32_t maxl = 0;
        for (int64_t g = 0; g < num_gates; ++g) {
            const int32_t *G = gates + g * 5;
            const int op = G[0];
            if (op < 0 || op >= B200_GATE_OP_COUNT) return b200_fail("netlist: invalid op");
            const int nin = lin_op(op) ? ((op == B200_NOT || op == B200_COPY) ? 1 : 0) : (op == B200_MUX ? 3 : 2);
            if (G[1] < 0 || G[1] >= num_wires) return b200_fail("netlist: output wire out of range");
            if (written[G[1]]) return b200_fail("netlist: wire written twice");
            int32_t lvl = 0;
            for (int i = 0; i < nin; ++i) {
                const int32_t w = G[2 + i];
                if (w < 0 || w >= num_wires) return b200_fail("netlist: input wire out of range");
                if (w == G[1]) return b200_fail("netlist: gate reads its own output");
                lvl = std::max(lvl, ready[w]);
            }
            if (!lin_op(op)) lvl += 1;
            ready[G[1]] = lvl;
            written[G[1]] = 1;
            if (gate_level) gate_level[g] = lvl;
            maxl = std::max(maxl, lvl);
        }
        // a gate may only read wires that are inputs or were written by an EARLIER gate
        std::fill(written.begin(), written.end(), 0);
        std::vector<uint8_t> is_out(num_wires, 0);
        for (int64_t g = 0; g < num_gates; ++g) is_out[gates[g * 5 + 1]] = 1;
        for (int64_t g = 0; g < num_gates; ++g) {
            const int32_t *G = gates + g * 5;
            const int op = G[0];
            const int nin = lin_op(op) ? ((op == B200_NOT || op == B200_COPY) ? 1 : 0) : (op == B200_MUX ? 3 : 2);
            for (int i = 0; i < nin; ++i)
                if (is_out[G[2 + i]] && !written[G[2 + i]]) return b200_fail("netlist: not in topological order");
            written[G[1]] = 1;
        }
        if (num_levels) *num_levels = maxl;
        return 0;
    });
}

extern "C" int b200_expand_gates(const int32_t *gates, int64_t num_gates, int32_t *rot_desc, int64_t *n_rot,
                                 int32_t *ks_desc, int64_t *n_ks) {
    // rot_desc rows {w0, w1, s0 | s1<<16, cst}; ks_desc rows {u0, u1, add_b, out_wire}
    if (!gates || !rot_desc || !ks_desc || !n_rot || !n_ks) return b200_fail("NULL argument");
    int64_t r = 0, k = 0;
    for (int64_t g = 0; g < num_gates; ++g) {
        const int32_t *G = gates + g * 5;
        const int op = G[0];
        if (op < 0 || op > B200_MUX) return b200_fail("expand: only bootstrapped gates");
        const int halves = op == B200_MUX ? 2 : 1;
        for (int h = 0; h < halves; ++h) {
            int s0, s1, second;
            int32_t cst;
            b200_gate_coeffs(op, h, &s0, &s1, &cst, &second);
            int32_t *R = rot_desc + (r + h) * 4;
            R[0] = G[2];
            R[1] = second == 2 ? G[4] : G[3];
            R[2] = (int32_t)((uint32_t)(uint16_t)(int16_t)s0 | ((uint32_t)(uint16_t)(int16_t)s1 << 16));
            R[3] = cst;
        }
        int32_t *K = ks_desc + k * 4;
        K[0] = (int32_t)r;
        K[1] = halves == 2 ? (int32_t)(r + 1) : -1;
        K[2] = halves == 2 ? (1 << 29) : 0;
        K[3] = G[1];
        r += halves;
        k += 1;
    }
    *n_rot = r;
    *n_ks = k;
    return 0;
}

// one level, descriptors on the host, wires on the device
static int level_run_locked(b200_key *key, int32_t *d_wires, const int32_t *rot_desc, int64_t n_rot,
                            const int32_t *ks_desc, int64_t n_ks, cudaStream_t s) {
    const DeviceKey &d = key->dk;
    if (n_rot <= 0 || n_ks <= 0) return 0;
    if (n_rot > (1 << 18)) return b200_fail("level too large: split it (max 2^18 rotations per call)");
    if (key->desc_rot.ensure(n_rot * 16) || key->desc_ks.ensure(n_ks * 16) ||
        key->u.ensure((size_t)n_rot * (d.N + 1) * 4))
        return 1;
    if (key->scratch_begin(s)) return 1;
    B200_CK(cudaMemcpyAsync(key->desc_rot.p, rot_desc, n_rot * 16, cudaMemcpyHostToDevice, s));
    B200_CK(cudaMemcpyAsync(key->desc_ks.p, ks_desc, n_ks * 16, cudaMemcpyHostToDevice, s));
    BlindRotateArgs br{};
    br.desc = (const RotDesc *)key->desc_rot.p;
    br.pool = d_wires;
    br.mu = 1 << 29;
    br.u = (int32_t *)key->u.p;
    br.items = n_rot;
    B200_CK(launch_blind_rotate(d, br, s));
    KeySwitchArgs ks{};
    ks.u = (const int32_t *)key->u.p;
    ks.desc = (const KsDesc *)key->desc_ks.p;
    ks.out = d_wires;
    ks.items = n_ks;
    B200_CK(launch_keyswitch(d, ks, s));
    // descriptors were copied from pageable host memory: the copy is complete (staged) on return
    return key->scratch_end(s);
}

extern "C" int b200_level_run_dev(b200_key *key, int32_t *d_wires, const int32_t *rot_desc, int64_t n_rot,
                                  const int32_t *ks_desc, int64_t n_ks, void *stream) {
    if (!key || !key->filled) return b200_fail("key is NULL or not filled");
    if (!d_wires || !rot_desc || !ks_desc) return b200_fail("NULL argument");
    LaneGuard lane(key);
    key = lane.k;
    B200_ON_DEVICE(key->dk.device);
    return level_run_locked(key, d_wires, rot_desc, n_rot, ks_desc, n_ks, (cudaStream_t)stream);
}

static int circuit_run_locked(b200_key *key, const int32_t *gates, int64_t num_gates, int32_t *d_wires,
                              int64_t num_wires, int32_t *levels_out, cudaStream_t s) {
    std::vector<int32_t> glevel(num_gates);
    int32_t nlev = 0;
    if (b200_circuit_levelize(gates, num_gates, num_wires, glevel.data(), &nlev)) return 1;
    // bucket gates by level; linear gates additionally by their chain depth inside the level
    std::vector<std::vector<int64_t>> boot(nlev + 1);
    std::vector<std::vector<std::vector<int64_t>>> lin(nlev + 1);
    std::vector<int32_t> lin_depth(num_wires, 0), wire_level(num_wires, 0);
    for (int64_t g = 0; g < num_gates; ++g) {
        const int32_t *G = gates + g * 5;
        const int l = glevel[g];
        if (lin_op(G[0])) {
            int dep = 1;
            if ((G[0] == B200_NOT || G[0] == B200_COPY) && wire_level[G[2]] == l) dep = lin_depth[G[2]] + 1;
            lin_depth[G[1]] = dep;
            wire_level[G[1]] = l;
            if ((int)lin[l].size() < dep) lin[l].resize(dep);
            lin[l][dep - 1].push_back(g);
        } else {
            boot[l].push_back(g);
            wire_level[G[1]] = l;
            lin_depth[G[1]] = 0;
        }
    }
    std::vector<int32_t> sub, rot, ks, ldesc;
    for (int l = 0; l <= nlev; ++l) {
        // bootstrapped gates of this level, in chunks that bound the extracted-sample workspace
        const std::vector<int64_t> &B = boot[l];
        size_t pos = 0;
        while (pos < B.size()) {
            size_t end = pos;
            int64_t rots = 0;
            while (end < B.size()) {
                const int per = gates[B[end] * 5] == B200_MUX ? 2 : 1;
                if (rots + per > (1 << 18)) break;
                rots += per;
                ++end;
            }
            sub.resize((end - pos) * 5);
            for (size_t i = pos; i < end; ++i) std::memcpy(&sub[(i - pos) * 5], gates + B[i] * 5, 5 * sizeof(int32_t));
            rot.resize(rots * 4);
            ks.resize((end - pos) * 4);
            int64_t nr = 0, nk = 0;
            if (b200_expand_gates(sub.data(), (int64_t)(end - pos), rot.data(), &nr, ks.data(), &nk)) return 1;
            if (level_run_locked(key, d_wires, rot.data(), nr, ks.data(), nk, s)) return 1;
            // the host descriptor vectors are reused for the next chunk
            B200_CK(cudaStreamSynchronize(s));
            pos = end;
        }
        for (const auto &L : lin[l]) {
            if (L.empty()) continue;
            ldesc.resize(L.size() * 4);
            for (size_t i = 0; i < L.size(); ++i) {
                const int32_t *G = gates + L[i] * 5;
                ldesc[i * 4 + 0] = G[0];
                ldesc[i * 4 + 1] = G[1];
                ldesc[i * 4 + 2] = (G[0] == B200_NOT || G[0] == B200_COPY) ? G[2] : -1;
                ldesc[i * 4 + 3] = 0;
            }
            if (key->desc_lin.ensure(ldesc.size() * 4)) return 1;
            B200_CK(cudaMemcpyAsync(key->desc_lin.p, ldesc.data(), ldesc.size() * 4, cudaMemcpyHostToDevice, s));
            B200_CK(launch_linear_desc(key->dk, d_wires, (const int32_t *)key->desc_lin.p, (int64_t)L.size(), s));
            B200_CK(cudaStreamSynchronize(s));
        }
    }
    if (levels_out) *levels_out = nlev;
    return 0;
}

extern "C" int b200_circuit_run_dev(b200_key *key, const int32_t *gates, int64_t num_gates, int32_t *d_wires,
                                    int64_t num_wires, int32_t *levels_out, void *stream) {
    return b200_guarded([&]() -> int {
        if (!key || !key->filled) return b200_fail("key is NULL or not filled");
        if (!gates || !d_wires) return b200_fail("NULL argument");
        LaneGuard lane(key);
    key = lane.k;
        B200_ON_DEVICE(key->dk.device);
        return circuit_run_locked(key, gates, num_gates, d_wires, num_wires, levels_out, (cudaStream_t)stream);
    });
}

extern "C" int b200_circuit_run_host(b200_key *key, const int32_t *gates, int64_t num_gates, int32_t *wires,
                                     int64_t num_wires, int32_t *levels_out) {
    return b200_guarded([&]() -> int {
        if (!key || !key->filled) return b200_fail("key is NULL or not filled");
        if (!gates || !wires) return b200_fail("NULL argument");
        LaneGuard lane(key);
    key = lane.k;
        B200_ON_DEVICE(key->dk.device);
        const size_t bytes = (size_t)num_wires * (key->dk.n + 1) * 4;
        if (key->out.ensure(bytes)) return 1;
        cudaStream_t s = key->stream;
        B200_CK(cudaMemcpyAsync(key->out.p, wires, bytes, cudaMemcpyHostToDevice, s));
        if (circuit_run_locked(key, gates, num_gates, (int32_t *)key->out.p, num_wires, levels_out, s)) return 1;
        B200_CK(cudaMemcpyAsync(wires, key->out.p, bytes, cudaMemcpyDeviceToHost, s));
        B200_CK(cudaStreamSynchronize(s));
        return 0;
    });
}

// Packed I/O: the device pool is [num_wires | n_in staged input rows | n_out staged output rows]; inputs
// are scattered into their wires and outputs gathered from theirs by row copies on the device, so the
// host moves n_in + n_out rows instead of 2 x num_wires.
extern "C" int b200_circuit_run_host_io(b200_key *key, const int32_t *gates, int64_t num_gates, int64_t num_wires,
                                        const int32_t *in_wires, int64_t n_in, const int32_t *in_rows,
                                        const int32_t *out_wires, int64_t n_out, int32_t *out_rows,
                                        int32_t *levels_out) {
    return b200_guarded([&]() -> int {
        if (!key || !key->filled) return b200_fail("key is NULL or not filled");
        if (!gates || num_wires <= 0 || n_in < 0 || n_out < 0 || (n_in && (!in_wires || !in_rows)) ||
            (n_out && (!out_wires || !out_rows)))
            return b200_fail("bad argument");
        if (num_wires + n_in + n_out > 0x7fffffffLL) return b200_fail("wire pool too large");
        std::vector<int32_t> desc((size_t)(n_in + n_out) * 4);
        for (int64_t i = 0; i < n_in; ++i) {
            if (in_wires[i] < 0 || in_wires[i] >= num_wires) return b200_fail("input wire out of range");
            int32_t *d = &desc[(size_t)i * 4];
            d[0] = B200_COPY; d[1] = in_wires[i]; d[2] = (int32_t)(num_wires + i); d[3] = 0;
        }
        for (int64_t i = 0; i < n_out; ++i) {
            if (out_wires[i] < 0 || out_wires[i] >= num_wires) return b200_fail("output wire out of range");
            int32_t *d = &desc[(size_t)(n_in + i) * 4];
            d[0] = B200_COPY; d[1] = (int32_t)(num_wires + n_in + i); d[2] = out_wires[i]; d[3] = 0;
        }
        LaneGuard lane(key);
    key = lane.k;
        B200_ON_DEVICE(key->dk.device);
        const size_t row = (size_t)(key->dk.n + 1) * 4;
        if (key->out.ensure((size_t)(num_wires + n_in + n_out) * row)) return 1;
        if (key->desc_io.ensure(desc.size() * 4 + 16)) return 1;
        cudaStream_t s = key->stream;
        int32_t *pool = (int32_t *)key->out.p;
        int32_t *d_desc = (int32_t *)key->desc_io.p;
        B200_CK(cudaMemsetAsync(pool, 0, (size_t)num_wires * row, s));
        if (!desc.empty()) B200_CK(cudaMemcpyAsync(d_desc, desc.data(), desc.size() * 4, cudaMemcpyHostToDevice, s));
        if (n_in) {
            B200_CK(cudaMemcpyAsync((char *)pool + (size_t)num_wires * row, in_rows, (size_t)n_in * row, cudaMemcpyHostToDevice, s));
            B200_CK(b200::launch_linear_desc(key->dk, pool, d_desc, n_in, s));
        }
        if (circuit_run_locked(key, gates, num_gates, pool, num_wires, levels_out, s)) return 1;
        if (n_out) {
            B200_CK(b200::launch_linear_desc(key->dk, pool, d_desc + (size_t)n_in * 4, n_out, s));
            B200_CK(cudaMemcpyAsync(out_rows, (char *)pool + (size_t)(num_wires + n_in) * row, (size_t)n_out * row, cudaMemcpyDeviceToHost, s));
        }
        B200_CK(cudaStreamSynchronize(s));
        return 0;
    });
}
