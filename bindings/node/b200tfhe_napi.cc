// b200tfhe_napi.cc -- Node-API addon over the C ABI of libb200tfhe.so (include/b200tfhe.h).
//
// The reference's JavaScript harness (tests/tfhe.test.js:56-186) reaches native code only through the
// AO loader: WASM -> Lua -> ao-tfhe/eoc-tfhe-bindings.c:128-148.  This addon is the GPU-side equivalent
// of that glue for the gate-bootstrapping path: the JS test loads it with `require()` and passes typed
// arrays; nothing but pointers and sizes reaches the C ABI.  It is a thin marshalling layer: no TFHE
// arithmetic happens here.
//
//   const b200 = require('./build/b200tfhe.node')
//   b200.info()                                   -> { devices, launches }
//   const key = b200.uploadCloudKey(buffer)       // bytes of export_tfheGateBootstrappingCloudKeySet_toFile
//   const p   = b200.keyParams(key)               -> { n, N, k, l, Bgbit, ks_t, ks_basebit }
//   const out = b200.gateBatch(key, op, a, b, c)  // Int32Array [count*(n+1)] each; c only for MUX (op 10)
//   const lv  = b200.runCircuit(key, gates, wires)// Int32Array [G*5], Int32Array [W*(n+1)] updated in place
//   const cts = b200.readSamples(buffer, n)       // libtfhe LWE sample records -> Int32Array
//   b200.freeKey(key)
//
// Build (where Node is installed; not in the round-1 image, where this file is compile-checked against
// node_api_min.h by __graft_entry__.build()):
//   g++ -shared -fPIC -DB200_HAVE_NODE_API_H -I<node prefix>/include/node -I../../include
//       b200tfhe_napi.cc -L../../eoc_tfhe_b200 -lb200tfhe -o build/b200tfhe.node
#include "node_api_min.h"

#define B200TFHE_NO_LIBTFHE_TYPES
struct LweSample;
struct TFheGateBootstrappingCloudKeySet;
struct LweBootstrappingKeyFFT;
struct LweKeySwitchKey;
typedef int32_t Torus32;
#include "../../include/b200tfhe.h"

#include <string.h>

#include <string>

namespace {

napi_value fail(napi_env env, const char *what) {
    const char *detail = b200_last_error();
    char msg[512];
    size_t n = strlen(what);
    if (n > 200) n = 200;
    memcpy(msg, what, n);
    msg[n] = 0;
    if (detail && detail[0]) {
        strncat(msg, ": ", sizeof msg - strlen(msg) - 1);
        strncat(msg, detail, sizeof msg - strlen(msg) - 1);
    }
    napi_throw_error(env, "B200TFHE", msg);
    return nullptr;
}

bool int32_array(napi_env env, napi_value v, int32_t **data, size_t *len) {
    napi_typedarray_type t;
    void *p = nullptr;
    if (napi_get_typedarray_info(env, v, &t, len, &p, nullptr, nullptr) != napi_ok || t != napi_int32_array) return false;
    *data = static_cast<int32_t *>(p);
    return true;
}

napi_value new_int32_array(napi_env env, size_t len, int32_t **data) {
    napi_value ab, ta;
    void *p = nullptr;
    if (napi_create_arraybuffer(env, len * 4, &p, &ab) != napi_ok) return nullptr;
    if (napi_create_typedarray(env, napi_int32_array, len, ab, 0, &ta) != napi_ok) return nullptr;
    *data = static_cast<int32_t *>(p);
    return ta;
}

void set_int(napi_env env, napi_value obj, const char *name, int32_t v) {
    napi_value x;
    napi_create_int32(env, v, &x);
    napi_set_named_property(env, obj, name, x);
}

b200_key *key_arg(napi_env env, napi_value v) {
    void *p = nullptr;
    if (napi_get_value_external(env, v, &p) != napi_ok) return nullptr;
    return static_cast<b200_key *>(p);
}

napi_value Info(napi_env env, napi_callback_info) {
    napi_value obj;
    napi_create_object(env, &obj);
    set_int(env, obj, "devices", b200_device_count());
    napi_value l;
    napi_create_double(env, (double)b200_launch_count(), &l);
    napi_set_named_property(env, obj, "launches", l);
    return obj;
}

napi_value UploadCloudKey(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    void *data = nullptr;
    size_t len = 0;
    if (argc < 1 || napi_get_buffer_info(env, argv[0], &data, &len) != napi_ok) return fail(env, "uploadCloudKey(buffer)");
    b200_key *key = nullptr;
    if (b200_key_from_cloudkey_bytes(data, len, &key)) return fail(env, "uploadCloudKey");
    napi_value ext;
    napi_create_external(env, key, nullptr, nullptr, &ext);   // freed explicitly with freeKey (device memory)
    return ext;
}

napi_value FreeKey(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1], undef;
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    if (argc == 1) b200_key_free(key_arg(env, argv[0]));
    napi_get_undefined(env, &undef);
    return undef;
}

napi_value KeyParams(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    b200_key *key = argc == 1 ? key_arg(env, argv[0]) : nullptr;
    b200_params p;
    if (!key || b200_key_params(key, &p)) return fail(env, "keyParams(key)");
    napi_value obj;
    napi_create_object(env, &obj);
    set_int(env, obj, "n", p.n); set_int(env, obj, "N", p.N); set_int(env, obj, "k", p.k); set_int(env, obj, "l", p.l);
    set_int(env, obj, "Bgbit", p.Bgbit); set_int(env, obj, "ks_t", p.ks_t); set_int(env, obj, "ks_basebit", p.ks_basebit);
    return obj;
}

// gateBatch(key, op, a, b?, c?) -> Int32Array
napi_value GateBatch(napi_env env, napi_callback_info info) {
    size_t argc = 5;
    napi_value argv[5];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    if (argc < 3) return fail(env, "gateBatch(key, op, a, b, c)");
    b200_key *key = key_arg(env, argv[0]);
    int32_t op = 0;
    b200_params p;
    if (!key || napi_get_value_int32(env, argv[1], &op) != napi_ok || b200_key_params(key, &p)) return fail(env, "gateBatch: bad key or op");
    int32_t *in[3] = {nullptr, nullptr, nullptr};
    size_t len[3] = {0, 0, 0};
    for (size_t i = 0; i + 2 < argc && i < 3; ++i)
        if (!int32_array(env, argv[i + 2], &in[i], &len[i])) in[i] = nullptr;
    const size_t row = (size_t)p.n + 1;
    if (!in[0] || len[0] % row) return fail(env, "gateBatch: a must be an Int32Array of count*(n+1) values");
    for (int i = 1; i < 3; ++i)
        if (in[i] && len[i] != len[0]) return fail(env, "gateBatch: batches differ in length");
    int32_t *out = nullptr;
    napi_value result = new_int32_array(env, len[0], &out);
    if (!result) return fail(env, "gateBatch: allocation failed");
    if (b200_gate_batch_host(key, op, out, in[0], in[1], in[2], (int64_t)(len[0] / row))) return fail(env, "gateBatch");
    return result;
}

// runCircuit(key, gates, wires) -> number of bootstrapped levels; wires updated in place
napi_value RunCircuit(napi_env env, napi_callback_info info) {
    size_t argc = 3;
    napi_value argv[3];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    if (argc < 3) return fail(env, "runCircuit(key, gates, wires)");
    b200_key *key = key_arg(env, argv[0]);
    b200_params p;
    int32_t *gates = nullptr, *wires = nullptr;
    size_t ng = 0, nw = 0;
    if (!key || b200_key_params(key, &p) || !int32_array(env, argv[1], &gates, &ng) || !int32_array(env, argv[2], &wires, &nw))
        return fail(env, "runCircuit: bad arguments");
    const size_t row = (size_t)p.n + 1;
    if (ng % 5 || nw % row) return fail(env, "runCircuit: gates must be [G*5], wires [W*(n+1)]");
    int32_t levels = 0;
    if (b200_circuit_run_host(key, gates, (int64_t)(ng / 5), wires, (int64_t)(nw / row), &levels)) return fail(env, "runCircuit");
    napi_value v;
    napi_create_int32(env, levels, &v);
    return v;
}

// readSamples(buffer, n) -> Int32Array [count*(n+1)]
napi_value ReadSamples(napi_env env, napi_callback_info info) {
    size_t argc = 2;
    napi_value argv[2];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    void *data = nullptr;
    size_t len = 0;
    int32_t n = 0;
    if (argc < 2 || napi_get_buffer_info(env, argv[0], &data, &len) != napi_ok || napi_get_value_int32(env, argv[1], &n) != napi_ok || n <= 0)
        return fail(env, "readSamples(buffer, n)");
    const size_t rec = (size_t)b200_wire_sample_bytes(n);
    const size_t count = len / rec;
    int32_t *out = nullptr;
    napi_value result = new_int32_array(env, count * ((size_t)n + 1), &out);
    if (!result) return fail(env, "readSamples: allocation failed");
    if (b200_wire_read_samples(data, len, n, (int64_t)count, out, nullptr)) return fail(env, "readSamples");
    return result;
}

// ---- eoc's string entry points (ao-tfhe/eoc-tfhe-run.cpp:427-513 and the bit-sliced counterparts) -------
bool get_string(napi_env env, napi_value v, std::string *out) {
    size_t len = 0;
    if (napi_get_value_string_utf8(env, v, nullptr, 0, &len) != napi_ok) return false;
    out->resize(len + 1);
    if (napi_get_value_string_utf8(env, v, &(*out)[0], len + 1, &len) != napi_ok) return false;
    out->resize(len);
    return true;
}

// the reference returns nullptr on failure and the Lua/JS side sees nil (eoc-tfhe-bindings.c:128-148): null here
napi_value take_string(napi_env env, const char *r) {
    napi_value v;
    if (!r) {
        napi_get_null(env, &v);
        return v;
    }
    napi_create_string_utf8(env, r, NAPI_AUTO_LENGTH, &v);
    b200_eoc_free(r);
    return v;
}

template <const char *(*FN)(const char *, const char *, const char *)>
napi_value Linear(napi_env env, napi_callback_info info) {
    size_t argc = 3;
    napi_value argv[3];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    std::string a, b;
    if (argc < 2 || !get_string(env, argv[0], &a) || !get_string(env, argv[1], &b)) return fail(env, "(base64Ciphertext1, base64Ciphertext2[, base64PublicKey])");
    return take_string(env, FN(a.c_str(), b.c_str(), ""));
}

template <const char *(*FN)(const char *, const char *, int32_t)>
napi_value Bits(napi_env env, napi_callback_info info) {
    size_t argc = 3;
    napi_value argv[3];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    std::string x, y;
    int32_t nb = 0;
    if (argc < 3 || !get_string(env, argv[0], &x) || !get_string(env, argv[1], &y) || napi_get_value_int32(env, argv[2], &nb) != napi_ok)
        return fail(env, "(base64XBits, base64YBits, nbBits)");
    return take_string(env, FN(x.c_str(), y.c_str(), nb));
}

// setCloudKeyBase64(string): evaluation key of the bit-sliced calls
napi_value SetCloudKeyBase64(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    std::string k;
    if (argc < 1 || !get_string(env, argv[0], &k)) return fail(env, "setCloudKeyBase64(base64CloudKey)");
    if (b200_eoc_set_cloud_key(k.c_str())) return fail(env, "setCloudKeyBase64");
    napi_value u;
    napi_get_undefined(env, &u);
    return u;
}

void def(napi_env env, napi_value exports, const char *name, napi_callback cb) {
    napi_value fn;
    napi_create_function(env, name, NAPI_AUTO_LENGTH, cb, nullptr, &fn);
    napi_set_named_property(env, exports, name, fn);
}

}  // namespace

// Node looks this symbol up when it loads the addon (NAPI_MODULE_INIT in node_api.h expands to it).
extern "C" __attribute__((visibility("default"))) napi_value napi_register_module_v1(napi_env env, napi_value exports) {
    def(env, exports, "info", Info);
    def(env, exports, "uploadCloudKey", UploadCloudKey);
    def(env, exports, "freeKey", FreeKey);
    def(env, exports, "keyParams", KeyParams);
    def(env, exports, "gateBatch", GateBatch);
    def(env, exports, "runCircuit", RunCircuit);
    def(env, exports, "readSamples", ReadSamples);
    def(env, exports, "addCiphertexts", Linear<addCiphertexts>);
    def(env, exports, "subtractCiphertexts", Linear<subtractCiphertexts>);
    def(env, exports, "setCloudKeyBase64", SetCloudKeyBase64);
    def(env, exports, "addCiphertextBits", Bits<addCiphertextBits>);
    def(env, exports, "subtractCiphertextBits", Bits<subtractCiphertextBits>);
    return exports;
}
