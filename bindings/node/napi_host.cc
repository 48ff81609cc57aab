// napi_host.cc -- TEST INFRASTRUCTURE: a minimal Node-API host, just enough of node's C ABI to load the addon
// (b200tfhe.node) and call its exported functions without a JavaScript engine.
//
// `node` is not installed in the build image nor on the GPU box, so tests/test_node_addon.py drives the addon
// through this host instead: it implements the napi_* functions the addon imports (the signatures of Node's stable
// C ABI as declared in node_api_min.h) over a small tagged value model, loads the .node file with dlopen -- the
// addon's undefined napi_* symbols bind to the definitions here -- calls napi_register_module_v1 like node does,
// and offers a C interface (host_*) to build arguments, call an export by name and inspect the result.
// What it proves: the addon links against libb200tfhe.so, registers its exports, marshals typed arrays / buffers /
// strings / externals correctly and reports failures as pending JavaScript exceptions.  What it cannot prove: V8's
// garbage collection and real Buffer semantics.
//
// build: g++ -std=c++17 -shared -fPIC -o build/libnapi_host.so napi_host.cc -ldl
#include "node_api_min.h"

#include <dlfcn.h>
#include <string.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

struct napi_value__ {
    enum Kind { UNDEFINED, NUL, INT32, DOUBLE, STRING, EXTERNAL, ARRAYBUFFER, TYPEDARRAY, BUFFER, OBJECT, FUNCTION } kind = UNDEFINED;
    int32_t i = 0;
    double d = 0;
    std::string s;
    void *ext = nullptr;
    std::shared_ptr<std::vector<unsigned char>> bytes;       // ARRAYBUFFER / BUFFER storage; TYPEDARRAY shares it
    napi_typedarray_type ta_type = napi_int32_array;
    size_t ta_len = 0, ta_off = 0;
    std::map<std::string, napi_value> props;
    napi_callback fn = nullptr;
    void *fn_data = nullptr;
};
struct napi_env__ {
    std::vector<std::unique_ptr<napi_value__>> heap;          // everything lives until host_reset()
    bool pending = false;
    std::string err_code, err_msg;
    void *dl = nullptr;
    napi_value exports = nullptr;
};
struct napi_callback_info__ {
    std::vector<napi_value> args;
    void *data = nullptr;
};

static napi_value mk(napi_env env, napi_value__::Kind k) {
    env->heap.emplace_back(new napi_value__());
    env->heap.back()->kind = k;
    return env->heap.back().get();
}

extern "C" {
// ---- the part of node's ABI the addon imports -------------------------------------------------------------------
napi_status napi_create_function(napi_env env, const char *, size_t, napi_callback cb, void *data, napi_value *result) {
    napi_value v = mk(env, napi_value__::FUNCTION);
    v->fn = cb;
    v->fn_data = data;
    *result = v;
    return napi_ok;
}
napi_status napi_set_named_property(napi_env, napi_value object, const char *name, napi_value value) {
    if (!object || object->kind != napi_value__::OBJECT) return napi_invalid_arg;
    object->props[name] = value;
    return napi_ok;
}
napi_status napi_get_cb_info(napi_env, napi_callback_info info, size_t *argc, napi_value *argv, napi_value *this_arg, void **data) {
    if (argc) {
        const size_t cap = *argc;
        for (size_t i = 0; i < cap && argv; ++i) argv[i] = i < info->args.size() ? info->args[i] : nullptr;
        *argc = info->args.size();      // node reports the actual count, which may exceed the capacity
    }
    if (this_arg) *this_arg = nullptr;
    if (data) *data = info->data;
    return napi_ok;
}
napi_status napi_get_typedarray_info(napi_env, napi_value v, napi_typedarray_type *type, size_t *length, void **data, napi_value *ab, size_t *off) {
    if (!v || v->kind != napi_value__::TYPEDARRAY) return napi_invalid_arg;
    if (type) *type = v->ta_type;
    if (length) *length = v->ta_len;
    if (data) *data = v->bytes->data() + v->ta_off;
    if (ab) *ab = nullptr;
    if (off) *off = v->ta_off;
    return napi_ok;
}
napi_status napi_get_buffer_info(napi_env, napi_value v, void **data, size_t *length) {
    if (!v || v->kind != napi_value__::BUFFER) return napi_invalid_arg;
    if (data) *data = v->bytes->data();
    if (length) *length = v->bytes->size();
    return napi_ok;
}
napi_status napi_create_arraybuffer(napi_env env, size_t byte_length, void **data, napi_value *result) {
    napi_value v = mk(env, napi_value__::ARRAYBUFFER);
    v->bytes = std::make_shared<std::vector<unsigned char>>(byte_length);
    if (data) *data = v->bytes->data();
    *result = v;
    return napi_ok;
}
napi_status napi_create_typedarray(napi_env env, napi_typedarray_type type, size_t length, napi_value ab, size_t off, napi_value *result) {
    if (!ab || ab->kind != napi_value__::ARRAYBUFFER) return napi_invalid_arg;
    const size_t elem = (type == napi_int32_array || type == napi_uint32_array || type == napi_float32_array) ? 4
                        : (type == napi_float64_array) ? 8 : (type == napi_int16_array || type == napi_uint16_array) ? 2 : 1;
    if (off + length * elem > ab->bytes->size()) return napi_invalid_arg;
    napi_value v = mk(env, napi_value__::TYPEDARRAY);
    v->bytes = ab->bytes;
    v->ta_type = type;
    v->ta_len = length;
    v->ta_off = off;
    *result = v;
    return napi_ok;
}
napi_status napi_create_int32(napi_env env, int32_t value, napi_value *result) { *result = mk(env, napi_value__::INT32); (*result)->i = value; return napi_ok; }
napi_status napi_create_double(napi_env env, double value, napi_value *result) { *result = mk(env, napi_value__::DOUBLE); (*result)->d = value; return napi_ok; }
napi_status napi_get_value_int32(napi_env, napi_value v, int32_t *result) {
    if (!v || (v->kind != napi_value__::INT32 && v->kind != napi_value__::DOUBLE)) return napi_invalid_arg;
    *result = v->kind == napi_value__::INT32 ? v->i : (int32_t)v->d;
    return napi_ok;
}
napi_status napi_create_external(napi_env env, void *data, napi_finalize, void *, napi_value *result) { *result = mk(env, napi_value__::EXTERNAL); (*result)->ext = data; return napi_ok; }
napi_status napi_get_value_external(napi_env, napi_value v, void **result) {
    if (!v || v->kind != napi_value__::EXTERNAL) return napi_invalid_arg;
    *result = v->ext;
    return napi_ok;
}
napi_status napi_create_object(napi_env env, napi_value *result) { *result = mk(env, napi_value__::OBJECT); return napi_ok; }
napi_status napi_create_string_utf8(napi_env env, const char *str, size_t length, napi_value *result) {
    *result = mk(env, napi_value__::STRING);
    (*result)->s = length == NAPI_AUTO_LENGTH ? std::string(str) : std::string(str, length);
    return napi_ok;
}
napi_status napi_get_value_string_utf8(napi_env, napi_value v, char *buf, size_t bufsize, size_t *result) {
    if (!v || v->kind != napi_value__::STRING) return napi_invalid_arg;
    if (!buf) { if (result) *result = v->s.size(); return napi_ok; }       // length query
    const size_t n = bufsize == 0 ? 0 : (v->s.size() < bufsize - 1 ? v->s.size() : bufsize - 1);
    memcpy(buf, v->s.data(), n);
    if (bufsize) buf[n] = 0;
    if (result) *result = n;
    return napi_ok;
}
napi_status napi_get_null(napi_env env, napi_value *result) { *result = mk(env, napi_value__::NUL); return napi_ok; }
napi_status napi_get_undefined(napi_env env, napi_value *result) { *result = mk(env, napi_value__::UNDEFINED); return napi_ok; }
napi_status napi_throw_error(napi_env env, const char *code, const char *msg) {
    env->pending = true;
    env->err_code = code ? code : "";
    env->err_msg = msg ? msg : "";
    return napi_ok;
}

// ---- the driver side (ctypes) ------------------------------------------------------------------------------------
napi_env host_load(const char *addon_path, char *err, size_t errlen) {
    napi_env env = new napi_env__();
    env->dl = dlopen(addon_path, RTLD_NOW | RTLD_LOCAL);
    auto init = env->dl ? (napi_value(*)(napi_env, napi_value))dlsym(env->dl, "napi_register_module_v1") : nullptr;
    if (!init) {
        if (err && errlen) { strncpy(err, env->dl ? "napi_register_module_v1 not exported" : dlerror(), errlen - 1); err[errlen - 1] = 0; }
        delete env;
        return nullptr;
    }
    env->exports = init(env, mk(env, napi_value__::OBJECT));
    return env;
}
int host_export_count(napi_env env) { return env->exports ? (int)env->exports->props.size() : 0; }
int host_export_name(napi_env env, int i, char *buf, size_t len) {
    int k = 0;
    for (auto &kv : env->exports->props)
        if (k++ == i) { strncpy(buf, kv.first.c_str(), len - 1); buf[len - 1] = 0; return kv.second->kind == napi_value__::FUNCTION; }
    return -1;
}
napi_value host_int32(napi_env env, int32_t v) { napi_value r; napi_create_int32(env, v, &r); return r; }
napi_value host_string(napi_env env, const char *s) { napi_value r; napi_create_string_utf8(env, s, NAPI_AUTO_LENGTH, &r); return r; }
napi_value host_buffer(napi_env env, const void *p, size_t n) {
    napi_value v = mk(env, napi_value__::BUFFER);
    v->bytes = std::make_shared<std::vector<unsigned char>>((const unsigned char *)p, (const unsigned char *)p + n);
    return v;
}
napi_value host_int32array(napi_env env, const int32_t *p, size_t n) {
    napi_value ab, ta;
    void *data;
    napi_create_arraybuffer(env, n * 4, &data, &ab);
    memcpy(data, p, n * 4);
    napi_create_typedarray(env, napi_int32_array, n, ab, 0, &ta);
    return ta;
}
// call exports[name](argv...); returns the result (NULL when the call threw: see host_error)
napi_value host_call(napi_env env, const char *name, int argc, napi_value *argv) {
    env->pending = false;
    env->err_code.clear();
    env->err_msg.clear();
    auto it = env->exports->props.find(name);
    if (it == env->exports->props.end() || it->second->kind != napi_value__::FUNCTION) {
        napi_throw_error(env, "HOST", "no such export");
        return nullptr;
    }
    napi_callback_info__ info;
    info.args.assign(argv, argv + argc);
    info.data = it->second->fn_data;
    napi_value r = it->second->fn(env, &info);
    return env->pending ? nullptr : r;
}
const char *host_error(napi_env env) { return env->pending ? env->err_msg.c_str() : nullptr; }
int host_kind(napi_value v) { return v ? (int)v->kind : -1; }
int32_t host_get_int32(napi_value v) { return v->kind == napi_value__::INT32 ? v->i : (int32_t)v->d; }
double host_get_double(napi_value v) { return v->kind == napi_value__::DOUBLE ? v->d : (double)v->i; }
const char *host_get_string(napi_value v) { return v->s.c_str(); }
napi_value host_get_property(napi_value v, const char *name) {
    auto it = v->props.find(name);
    return it == v->props.end() ? nullptr : it->second;
}
size_t host_typedarray_length(napi_value v) { return v->ta_len; }
void *host_typedarray_data(napi_value v) { return v->bytes->data() + v->ta_off; }
void host_reset(napi_env env) {      // drop every value but the exports (and what hangs off them)
    std::vector<std::unique_ptr<napi_value__>> keep;
    for (auto &p : env->heap) {
        bool is_export = p.get() == env->exports;
        for (auto &kv : env->exports->props) is_export = is_export || kv.second == p.get();
        if (is_export) keep.push_back(std::move(p));
    }
    env->heap.swap(keep);
}
void host_unload(napi_env env) {
    if (env->dl) dlclose(env->dl);
    delete env;
}
}  // extern "C"
