// eoc_abi.cpp -- evaluation side of eoc's C entry points over the engine.
//
// eoc-tfhe (ao-tfhe/eoc-tfhe-run.cpp) wraps libtfhe in nine extern "C" functions that exchange base64
// strings.  Seven of them need the secret key (key generation, encryption, decryption) and stay in the
// reference.  The two that evaluate on ciphertexts are restated here:
//
//   addCiphertexts / subtractCiphertexts (eoc-tfhe-run.cpp:427-469, :472-513): base64 -> LweSample wire
//   record (tfhe_io.cpp:90-107) -> lweCopy + lweAddTo / lweSubTo (lwe-functions.cpp:110-116, 139-145,
//   222-235: wrap-around int32 on a[] and b, variances ADD in both cases) -> wire record -> base64.
//   One 2.5 KB sample: host integer work, bit-exact, no device involved.
//
// and one capability is added next to them: integers encrypted bit by bit (one gate-bootstrapping
// ciphertext per bit, as test-addition-boot.cpp:190-197 does) are added or subtracted with the
// bootstrapped ripple-carry adder (full_adder, test-addition-boot.cpp:70-94) evaluated level by level
// on the GPU (b200_circuit_run_host): b200_add_bits_host for contiguous int32 arrays, any number of
// operand pairs per call, and addCiphertextBits / subtractCiphertextBits for base64 strings.
#include "../../include/b200tfhe.h"
#include "host_state.h"

#include <cstring>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

namespace {

const char kAlphabet[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";

// Standard alphabet, '=' padded -- the strings eoc-tfhe-run.cpp:48-68 produces.
std::string b64_encode(const unsigned char *p, size_t len) {
    std::string out;
    out.reserve((len + 2) / 3 * 4);
    size_t i = 0;
    for (; i + 3 <= len; i += 3) {
        const uint32_t v = (uint32_t)p[i] << 16 | (uint32_t)p[i + 1] << 8 | p[i + 2];
        out += kAlphabet[v >> 18];
        out += kAlphabet[(v >> 12) & 63];
        out += kAlphabet[(v >> 6) & 63];
        out += kAlphabet[v & 63];
    }
    if (len - i == 1) {
        const uint32_t v = (uint32_t)p[i] << 16;
        out += kAlphabet[v >> 18];
        out += kAlphabet[(v >> 12) & 63];
        out += "==";
    } else if (len - i == 2) {
        const uint32_t v = (uint32_t)p[i] << 16 | (uint32_t)p[i + 1] << 8;
        out += kAlphabet[v >> 18];
        out += kAlphabet[(v >> 12) & 63];
        out += kAlphabet[(v >> 6) & 63];
        out += '=';
    }
    return out;
}

// Decoding stops at the first character outside the alphabet (padding included), like
// eoc-tfhe-run.cpp:70-92.
std::vector<unsigned char> b64_decode(const char *s) {
    static const std::vector<int> T = [] {
        std::vector<int> t(256, -1);
        for (int i = 0; i < 64; ++i) t[(unsigned char)kAlphabet[i]] = i;
        return t;
    }();
    std::vector<unsigned char> out;
    uint32_t val = 0;
    int bits = 0;
    for (; *s; ++s) {
        const int d = T[(unsigned char)*s];
        if (d < 0) break;
        val = (val << 6) | (uint32_t)d;
        bits += 6;
        if (bits >= 8) {
            bits -= 8;
            out.push_back((unsigned char)((val >> bits) & 0xff));
        }
    }
    return out;
}

const char *to_c_string(const std::string &s) {
    char *p = new char[s.size() + 1];
    std::memcpy(p, s.c_str(), s.size() + 1);
    return p;
}

const char *fail_null(const std::string &msg) {
    b200_fail(msg);
    std::cerr << msg << std::endl;   // eoc-tfhe-run.cpp:466,510: message on cerr, nullptr to the caller
    return nullptr;
}

// dimension of the records in a decoded stream that holds `count` of them (0 if it does not fit)
int32_t record_dim(size_t len, int64_t count) {
    if (count <= 0 || len % (size_t)count) return 0;
    const size_t rec = len / (size_t)count;
    if (rec < 12 + 4 || (rec - 12) % 4) return 0;
    return (int32_t)((rec - 12) / 4 - 1);
}

b200_key *g_eval_key = nullptr;   // eoc keeps its keys in globals too (eoc-tfhe-run.cpp:38-39)
bool g_eval_key_owned = false;
std::mutex g_eval_mu;

const char *linear(const char *b64_1, const char *b64_2, int sign, const char *what) {
    if (!b64_1 || !b64_2) return fail_null(std::string(what) + ": NULL ciphertext string");
    const std::vector<unsigned char> r1 = b64_decode(b64_1), r2 = b64_decode(b64_2);
    const int32_t n = record_dim(r1.size(), 1);
    if (n <= 0 || r1.size() != r2.size()) return fail_null(std::string(what) + ": not two LWE sample records of one dimension");
    std::vector<int32_t> c1(n + 1), c2(n + 1);
    double v1 = 0, v2 = 0;
    if (b200_wire_read_samples(r1.data(), r1.size(), n, 1, c1.data(), &v1) ||
        b200_wire_read_samples(r2.data(), r2.size(), n, 1, c2.data(), &v2))
        return fail_null(std::string(what) + ": " + b200_last_error());
    for (int32_t i = 0; i <= n; ++i)
        c1[i] = (int32_t)((uint32_t)c1[i] + (uint32_t)sign * (uint32_t)c2[i]);
    v1 += v2;
    std::vector<unsigned char> out(r1.size());
    if (b200_wire_write_samples(out.data(), out.size(), n, 1, c1.data(), &v1)) return fail_null(b200_last_error());
    return to_c_string(b64_encode(out.data(), out.size()));
}

// x + y (subtract = 0) or x - y = x + ~y + 1 (subtract = 1) as a single-assignment netlist over
// wires [x bits | y bits | gates...]; out_wires receives the nb_bits + 1 result wires.
void ripple_netlist(int nb_bits, int subtract, std::vector<int32_t> &gates, int32_t &num_wires, std::vector<int32_t> &out_wires) {
    int32_t next = 2 * nb_bits;
    auto gate = [&](int op, int32_t a, int32_t b) {
        const int32_t row[5] = {op, next, a, b, -1};
        gates.insert(gates.end(), row, row + 5);
        return next++;
    };
    int32_t carry = gate(subtract ? B200_CONST1 : B200_CONST0, -1, -1);
    for (int i = 0; i < nb_bits; ++i) {
        const int32_t x = i, y = subtract ? gate(B200_NOT, nb_bits + i, -1) : nb_bits + i;
        const int32_t t0 = gate(B200_XOR, x, y);
        out_wires.push_back(gate(B200_XOR, t0, carry));
        const int32_t t1 = gate(B200_AND, x, y);
        const int32_t t2 = gate(B200_AND, carry, t0);
        carry = gate(B200_XOR, t1, t2);
    }
    out_wires.push_back(gate(B200_COPY, carry, -1));
    num_wires = next;
}

}  // namespace

extern "C" int b200_add_bits_host(b200_key *key, int32_t *sum, const int32_t *x, const int32_t *y, int32_t nb_bits,
                                  int64_t count, int subtract) {
    return b200_guarded([&]() -> int {
        if (!key) return b200_fail("key is NULL");
        if (!sum || !x || !y || nb_bits <= 0 || nb_bits > 4096 || count < 0) return b200_fail("bad argument");
        if (count == 0) return 0;
        b200_params P;
        if (b200_key_params(key, &P)) return 1;
        const size_t n1 = (size_t)P.n + 1;
        std::vector<int32_t> one, outw;
        int32_t nw = 0;
        ripple_netlist(nb_bits, subtract != 0, one, nw, outw);
        const int64_t per = (int64_t)one.size() / 5;
        if ((int64_t)nw * count > 0x7fffffffLL) return b200_fail("too many operand pairs for one call");
        // `count` independent instances over one pool, instance-major, so that a level holds the gates of
        // that depth of every operand pair
        std::vector<int32_t> gates((size_t)per * 5 * count);
        for (int64_t c = 0; c < count; ++c)
            for (int64_t g = 0; g < per; ++g) {
                int32_t *row = &gates[((size_t)c * per + g) * 5];
                row[0] = one[g * 5];
                for (int f = 1; f < 5; ++f) row[f] = one[g * 5 + f] < 0 ? -1 : (int32_t)(one[g * 5 + f] + c * nw);
            }
        // packed I/O: operand rows go up as they are (x bits, then y bits, per pair), sum rows come back
        std::vector<int32_t> in_wires((size_t)count * 2 * nb_bits), out_wires((size_t)count * (nb_bits + 1));
        std::vector<int32_t> in_rows((size_t)count * 2 * nb_bits * n1);
        for (int64_t c = 0; c < count; ++c) {
            for (int i = 0; i < 2 * nb_bits; ++i) in_wires[(size_t)c * 2 * nb_bits + i] = (int32_t)(c * nw + i);
            for (int i = 0; i <= nb_bits; ++i) out_wires[(size_t)c * (nb_bits + 1) + i] = (int32_t)(c * nw + outw[i]);
            std::memcpy(&in_rows[(size_t)c * 2 * nb_bits * n1], x + (size_t)c * nb_bits * n1, (size_t)nb_bits * n1 * 4);
            std::memcpy(&in_rows[((size_t)c * 2 + 1) * nb_bits * n1], y + (size_t)c * nb_bits * n1, (size_t)nb_bits * n1 * 4);
        }
        if (b200_circuit_run_host_io(key, gates.data(), per * count, (int64_t)nw * count, in_wires.data(), (int64_t)in_wires.size(),
                                     in_rows.data(), out_wires.data(), (int64_t)out_wires.size(), sum, nullptr))
            return 1;
        return 0;
    });
}

extern "C" int b200_eoc_set_key(b200_key *key) {
    std::lock_guard<std::mutex> g(g_eval_mu);
    if (g_eval_key && g_eval_key_owned) b200_key_free(g_eval_key);
    g_eval_key = key;
    g_eval_key_owned = false;
    return 0;
}

extern "C" int b200_eoc_set_cloud_key(const char *base64_cloud_key) {
    return b200_guarded([&]() -> int {
        if (!base64_cloud_key) return b200_fail("NULL cloud key string");
        const std::vector<unsigned char> raw = b64_decode(base64_cloud_key);
        b200_key *k = nullptr;
        if (b200_key_from_cloudkey_bytes(raw.data(), raw.size(), &k)) return 1;
        std::lock_guard<std::mutex> g(g_eval_mu);
        if (g_eval_key && g_eval_key_owned) b200_key_free(g_eval_key);
        g_eval_key = k;
        g_eval_key_owned = true;
        return 0;
    });
}

extern "C" void b200_eoc_free(const char *s) { delete[] s; }

extern "C" const char *addCiphertexts(const char *base64Ciphertext1, const char *base64Ciphertext2, const char *) {
    return linear(base64Ciphertext1, base64Ciphertext2, +1, "addCiphertexts");
}

extern "C" const char *subtractCiphertexts(const char *base64Ciphertext1, const char *base64Ciphertext2, const char *) {
    return linear(base64Ciphertext1, base64Ciphertext2, -1, "subtractCiphertexts");
}

static const char *bits_op(const char *b64x, const char *b64y, int nb_bits, int subtract, const char *what) {
    std::lock_guard<std::mutex> g(g_eval_mu);
    if (!g_eval_key) return fail_null("Public key not initialized. Generate the public key first.");  // eoc-tfhe-run.cpp:466
    if (!b64x || !b64y || nb_bits <= 0) return fail_null(std::string(what) + ": bad argument");
    b200_params P;
    if (b200_key_params(g_eval_key, &P)) return nullptr;
    const std::vector<unsigned char> rx = b64_decode(b64x), ry = b64_decode(b64y);
    if (record_dim(rx.size(), nb_bits) != P.n || ry.size() != rx.size())
        return fail_null(std::string(what) + ": operands are not nb_bits sample records of the key's dimension");
    const size_t n1 = (size_t)P.n + 1;
    std::vector<int32_t> x((size_t)nb_bits * n1), y(x.size()), s((size_t)(nb_bits + 1) * n1);
    if (b200_wire_read_samples(rx.data(), rx.size(), P.n, nb_bits, x.data(), nullptr) ||
        b200_wire_read_samples(ry.data(), ry.size(), P.n, nb_bits, y.data(), nullptr))
        return fail_null(std::string(what) + ": " + b200_last_error());
    if (b200_add_bits_host(g_eval_key, s.data(), x.data(), y.data(), nb_bits, 1, subtract)) {
        std::cerr << what << ": " << b200_last_error() << std::endl;
        return nullptr;
    }
    std::vector<unsigned char> out((size_t)b200_wire_sample_bytes(P.n) * (nb_bits + 1));
    if (b200_wire_write_samples(out.data(), out.size(), P.n, nb_bits + 1, s.data(), nullptr)) return fail_null(b200_last_error());
    return to_c_string(b64_encode(out.data(), out.size()));
}

extern "C" const char *addCiphertextBits(const char *x, const char *y, int32_t nb_bits) { return bits_op(x, y, nb_bits, 0, "addCiphertextBits"); }
extern "C" const char *subtractCiphertextBits(const char *x, const char *y, int32_t nb_bits) { return bits_op(x, y, nb_bits, 1, "subtractCiphertextBits"); }
