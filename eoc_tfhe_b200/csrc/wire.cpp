// wire.cpp -- reader / writer of libtfhe's serialisation for the objects that cross the hot path:
// the gate-bootstrapping cloud key and LWE samples.  Lets the engine ingest keys and ciphertexts
// produced by libtfhe (export_tfheGateBootstrappingCloudKeySet_toFile, export_gate_bootstrapping_
// ciphertext_toFile) without linking it.  Format restated from the reference:
//
//   text sections   "-----BEGIN <TITLE>-----\n" then "name: value\n" lines, then "-----END <TITLE>-----\n"
//                   (libs/tfhe/src/libtfhe/tfhe_generic_streams.cpp:112-173; integers printed "%10ld",
//                   doubles "%.8lf", :46-56; properties in std::map order, i.e. sorted by name)
//   cloud key       GATEBOOTSPARAMS{ks_basebit,ks_t}  LWEPARAMS{alpha_max,alpha_min,n}
//                   TLWEPARAMS{N,alpha_max,alpha_min,k}  TGSWPARAMS{Bgbit,l}  LWEKSPARAMS{basebit,n,t}
//                   then two binary blocks (tfhe_io.cpp:1031-1035,492-498,937-944,1099-1103):
//     KS block      int32 uid 200, double max variance, then for i<N, j<t, h<base: a[n] int32, b int32
//                   (tfhe_io.cpp:757-785)
//     BK block      int32 uid 201, double max variance, then for i<n, row<kpl, c<=k: N int32
//                   (tfhe_io.cpp:883-905)
//   LWE sample      int32 uid 42, a[n] int32, b int32, double current_variance  (tfhe_io.cpp:90-107)
// All binary fields are native endian, as the reference writes them with fwrite.
#include "../../include/b200tfhe.h"
#include "host_state.h"

#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace {

constexpr int32_t kUidLweSample = 42, kUidKs = 200, kUidBk = 201;

struct Cursor {
    const unsigned char *p;
    size_t len, pos;
    bool line(std::string &out) {  // getLine of CIstream (tfhe_generic_streams.cpp:66-73): '\r' dropped, '\n' ends
        out.clear();
        if (pos >= len) return false;
        while (pos < len) {
            const char c = (char)p[pos++];
            if (c == '\r') continue;
            if (c == '\n') return true;
            out.push_back(c);
        }
        return true;
    }
    bool bytes(void *dst, size_t n) {
        if (len - pos < n) return false;
        std::memcpy(dst, p + pos, n);
        pos += n;
        return true;
    }
};

// one text section -> title + properties (new_TextModeProperties_fromIstream, tfhe_generic_streams.cpp:112-150)
bool read_section(Cursor &c, std::string &title, std::map<std::string, std::string> &props) {
    std::string line, end;
    bool started = false;
    props.clear();
    while (c.line(line)) {
        const size_t n = line.size();
        if (n >= 16 && line.compare(0, 11, "-----BEGIN ") == 0 && line.compare(n - 5, 5, "-----") == 0) {
            title = line.substr(11, n - 16);
            end = "-----END " + title + "-----";
            started = true;
            continue;
        }
        if (!started) continue;  // the reference ignores anything before the body
        if (line == end) return true;
        const size_t pos = line.find(": ");
        if (pos == std::string::npos) continue;
        props[line.substr(0, pos)] = line.substr(pos + 2);
    }
    return false;
}

bool get_int(const std::map<std::string, std::string> &m, const char *name, int64_t &v) {
    auto it = m.find(name);
    if (it == m.end()) return false;
    v = std::strtol(it->second.c_str(), nullptr, 10);
    return true;
}
bool get_double(const std::map<std::string, std::string> &m, const char *name, double &v) {
    auto it = m.find(name);
    if (it == m.end()) return false;
    v = (double)std::strtold(it->second.c_str(), nullptr);
    return true;
}

void put_section(std::string &out, const char *title, const std::map<std::string, std::string> &props) {
    out += "-----BEGIN ";
    out += title;
    out += "-----\n";
    for (const auto &kv : props) out += kv.first + ": " + kv.second + "\n";
    out += "-----END ";
    out += title;
    out += "-----\n";
}
std::string fmt_int(int64_t v) {
    char buf[64];
    std::snprintf(buf, sizeof buf, "%10" PRId64, v);
    return buf;
}
std::string fmt_double(double v) {
    char buf[64];
    std::snprintf(buf, sizeof buf, "%.8lf", v);
    return buf;
}

}  // namespace

extern "C" int b200_wire_parse_cloudkey(const void *buf, size_t len, b200_params *p, b200_wire_cloudkey_info *info) {
    return b200_guarded([&]() -> int {
        if (!buf || !p || !info) return b200_fail("NULL argument");
        Cursor c{(const unsigned char *)buf, len, 0};
        std::string title;
        std::map<std::string, std::string> m;
        int64_t iv = 0;
        double dv = 0;
        std::memset(info, 0, sizeof *info);
        // GATEBOOTSPARAMS (tfhe_io.cpp:1014-1029; ks_basebit is read back through getProperty_double there)
        if (!read_section(c, title, m) || title != "GATEBOOTSPARAMS") return b200_fail("cloud key: GATEBOOTSPARAMS section missing");
        if (!get_int(m, "ks_t", iv)) return b200_fail("cloud key: ks_t missing");
        p->ks_t = (int32_t)iv;
        if (!get_double(m, "ks_basebit", dv)) return b200_fail("cloud key: ks_basebit missing");
        p->ks_basebit = (int32_t)dv;
        if (!read_section(c, title, m) || title != "LWEPARAMS") return b200_fail("cloud key: LWEPARAMS section missing");
        if (!get_int(m, "n", iv)) return b200_fail("cloud key: n missing");
        p->n = (int32_t)iv;
        get_double(m, "alpha_min", info->lwe_alpha_min);
        get_double(m, "alpha_max", info->lwe_alpha_max);
        if (!read_section(c, title, m) || title != "TLWEPARAMS") return b200_fail("cloud key: TLWEPARAMS section missing");
        if (!get_int(m, "N", iv)) return b200_fail("cloud key: N missing");
        p->N = (int32_t)iv;
        if (!get_int(m, "k", iv)) return b200_fail("cloud key: k missing");
        p->k = (int32_t)iv;
        get_double(m, "alpha_min", info->tlwe_alpha_min);
        get_double(m, "alpha_max", info->tlwe_alpha_max);
        if (!read_section(c, title, m) || title != "TGSWPARAMS") return b200_fail("cloud key: TGSWPARAMS section missing");
        if (!get_int(m, "l", iv)) return b200_fail("cloud key: l missing");
        p->l = (int32_t)iv;
        if (!get_int(m, "Bgbit", iv)) return b200_fail("cloud key: Bgbit missing");
        p->Bgbit = (int32_t)iv;
        if (!read_section(c, title, m) || title != "LWEKSPARAMS") return b200_fail("cloud key: LWEKSPARAMS section missing");
        int64_t ksn = 0, kst = 0, ksb = 0;
        if (!get_int(m, "n", ksn) || !get_int(m, "t", kst) || !get_int(m, "basebit", ksb)) return b200_fail("cloud key: LWEKSPARAMS incomplete");
        if (p->n <= 0 || p->N <= 0 || p->k <= 0 || p->l <= 0 || p->ks_t <= 0 || p->ks_basebit <= 0 || p->ks_basebit > 8)
            return b200_fail("cloud key: non-positive parameter");
        if (ksn != (int64_t)p->N * p->k || kst != p->ks_t || ksb != p->ks_basebit)
            return b200_fail("cloud key: key-switch section does not match the parameter set");
        // binary blocks
        int32_t uid = 0;
        double var = 0;
        if (!c.bytes(&uid, 4) || uid != kUidKs) return b200_fail("cloud key: key-switch block (uid 200) missing");
        if (!c.bytes(&var, 8)) return b200_fail("cloud key: truncated");
        p->ks_variance = var;
        const size_t ks_ints = (size_t)p->N * p->k * p->ks_t * ((size_t)1 << p->ks_basebit) * (p->n + 1);
        info->ks_offset = c.pos;
        info->ks_bytes = ks_ints * 4;
        if (len - c.pos < info->ks_bytes) return b200_fail("cloud key: key-switch block truncated");
        c.pos += info->ks_bytes;
        if (!c.bytes(&uid, 4) || uid != kUidBk) return b200_fail("cloud key: bootstrapping-key block (uid 201) missing");
        if (!c.bytes(&var, 8)) return b200_fail("cloud key: truncated");
        info->bk_variance = var;
        const size_t bk_ints = (size_t)p->n * (p->k + 1) * p->l * (p->k + 1) * p->N;
        info->bk_offset = c.pos;
        info->bk_bytes = bk_ints * 4;
        if (len - c.pos < info->bk_bytes) return b200_fail("cloud key: bootstrapping-key block truncated");
        info->total_bytes = c.pos + info->bk_bytes;
        return 0;
    });
}

extern "C" int b200_key_from_cloudkey_bytes(const void *buf, size_t len, b200_key **out) {
    return b200_guarded([&]() -> int {
        b200_params p;
        b200_wire_cloudkey_info info;
        if (b200_wire_parse_cloudkey(buf, len, &p, &info)) return 1;
        // the payloads are int32 arrays at arbitrary byte offsets after the text sections: realign
        std::vector<int32_t> ks(info.ks_bytes / 4), bk(info.bk_bytes / 4);
        std::memcpy(ks.data(), (const unsigned char *)buf + info.ks_offset, info.ks_bytes);
        std::memcpy(bk.data(), (const unsigned char *)buf + info.bk_offset, info.bk_bytes);
        return b200_key_create(&p, bk.data(), ks.data(), out);
    });
}

extern "C" int b200_key_from_cloudkey_file(const char *path, b200_key **out) {
    return b200_guarded([&]() -> int {
        if (!path) return b200_fail("path is NULL");
        FILE *f = std::fopen(path, "rb");
        if (!f) return b200_fail(std::string("cannot open ") + path);
        std::vector<unsigned char> data;
        unsigned char chunk[1 << 16];
        size_t got;
        while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) data.insert(data.end(), chunk, chunk + got);
        std::fclose(f);
        return b200_key_from_cloudkey_bytes(data.data(), data.size(), out);
    });
}

extern "C" int64_t b200_wire_cloudkey_size(const b200_params *p) {
    if (!p) return -1;
    std::vector<unsigned char> tmp;
    // header is small; compute by writing it
    std::string h;
    put_section(h, "GATEBOOTSPARAMS", {{"ks_t", fmt_int(p->ks_t)}, {"ks_basebit", fmt_int(p->ks_basebit)}});
    put_section(h, "LWEPARAMS", {{"n", fmt_int(p->n)}, {"alpha_min", fmt_double(0)}, {"alpha_max", fmt_double(0)}});
    put_section(h, "TLWEPARAMS", {{"N", fmt_int(p->N)}, {"k", fmt_int(p->k)}, {"alpha_min", fmt_double(0)}, {"alpha_max", fmt_double(0)}});
    put_section(h, "TGSWPARAMS", {{"l", fmt_int(p->l)}, {"Bgbit", fmt_int(p->Bgbit)}});
    put_section(h, "LWEKSPARAMS", {{"n", fmt_int((int64_t)p->N * p->k)}, {"t", fmt_int(p->ks_t)}, {"basebit", fmt_int(p->ks_basebit)}});
    const size_t ks_ints = (size_t)p->N * p->k * p->ks_t * ((size_t)1 << p->ks_basebit) * (p->n + 1);
    const size_t bk_ints = (size_t)p->n * (p->k + 1) * p->l * (p->k + 1) * p->N;
    return (int64_t)(h.size() + 12 + ks_ints * 4 + 12 + bk_ints * 4);
}

extern "C" int b200_wire_write_cloudkey(void *buf, size_t cap, size_t *written, const b200_params *p,
                                        const b200_wire_cloudkey_info *alphas, const int32_t *bk_coefs, const int32_t *ks_rows) {
    return b200_guarded([&]() -> int {
        if (!buf || !p || !bk_coefs || !ks_rows) return b200_fail("NULL argument");
        const double a0 = alphas ? alphas->lwe_alpha_min : 0, a1 = alphas ? alphas->lwe_alpha_max : 0;
        const double t0 = alphas ? alphas->tlwe_alpha_min : 0, t1 = alphas ? alphas->tlwe_alpha_max : 0;
        const double bkvar = alphas ? alphas->bk_variance : 0;
        std::string h;
        put_section(h, "GATEBOOTSPARAMS", {{"ks_t", fmt_int(p->ks_t)}, {"ks_basebit", fmt_int(p->ks_basebit)}});
        put_section(h, "LWEPARAMS", {{"n", fmt_int(p->n)}, {"alpha_min", fmt_double(a0)}, {"alpha_max", fmt_double(a1)}});
        put_section(h, "TLWEPARAMS", {{"N", fmt_int(p->N)}, {"k", fmt_int(p->k)}, {"alpha_min", fmt_double(t0)}, {"alpha_max", fmt_double(t1)}});
        put_section(h, "TGSWPARAMS", {{"l", fmt_int(p->l)}, {"Bgbit", fmt_int(p->Bgbit)}});
        put_section(h, "LWEKSPARAMS", {{"n", fmt_int((int64_t)p->N * p->k)}, {"t", fmt_int(p->ks_t)}, {"basebit", fmt_int(p->ks_basebit)}});
        const size_t ks_bytes = (size_t)p->N * p->k * p->ks_t * ((size_t)1 << p->ks_basebit) * (p->n + 1) * 4;
        const size_t bk_bytes = (size_t)p->n * (p->k + 1) * p->l * (p->k + 1) * p->N * 4;
        const size_t total = h.size() + 12 + ks_bytes + 12 + bk_bytes;
        if (written) *written = total;
        if (cap < total) return b200_fail("buffer too small for the cloud key");
        unsigned char *o = (unsigned char *)buf;
        std::memcpy(o, h.data(), h.size());
        o += h.size();
        std::memcpy(o, &kUidKs, 4);
        std::memcpy(o + 4, &p->ks_variance, 8);
        std::memcpy(o + 12, ks_rows, ks_bytes);
        o += 12 + ks_bytes;
        std::memcpy(o, &kUidBk, 4);
        std::memcpy(o + 4, &bkvar, 8);
        std::memcpy(o + 12, bk_coefs, bk_bytes);
        return 0;
    });
}

// ---- LWE samples ------------------------------------------------------------------------------------
extern "C" int64_t b200_wire_sample_bytes(int32_t n) { return 4 + (int64_t)(n + 1) * 4 + 8; }

extern "C" int b200_wire_read_samples(const void *buf, size_t len, int32_t n, int64_t count, int32_t *cts, double *variances) {
    if (!buf || !cts || n <= 0 || count < 0) return b200_fail("bad argument");
    const size_t rec = (size_t)b200_wire_sample_bytes(n);
    if (len < rec * (size_t)count) return b200_fail("sample stream truncated");
    const unsigned char *p = (const unsigned char *)buf;
    for (int64_t i = 0; i < count; ++i, p += rec) {
        int32_t uid;
        std::memcpy(&uid, p, 4);
        if (uid != kUidLweSample) return b200_fail("not an LWE sample record (uid != 42)");  // the reference aborts (tfhe_io.cpp:94)
        std::memcpy(cts + (size_t)i * (n + 1), p + 4, (size_t)(n + 1) * 4);
        if (variances) std::memcpy(variances + i, p + 4 + (size_t)(n + 1) * 4, 8);
    }
    return 0;
}

extern "C" int b200_wire_write_samples(void *buf, size_t cap, int32_t n, int64_t count, const int32_t *cts, const double *variances) {
    if (!buf || !cts || n <= 0 || count < 0) return b200_fail("bad argument");
    const size_t rec = (size_t)b200_wire_sample_bytes(n);
    if (cap < rec * (size_t)count) return b200_fail("buffer too small for the samples");
    unsigned char *p = (unsigned char *)buf;
    const double zero = 0;
    for (int64_t i = 0; i < count; ++i, p += rec) {
        std::memcpy(p, &kUidLweSample, 4);
        std::memcpy(p + 4, cts + (size_t)i * (n + 1), (size_t)(n + 1) * 4);
        std::memcpy(p + 4 + (size_t)(n + 1) * 4, variances ? &variances[i] : &zero, 8);
    }
    return 0;
}
