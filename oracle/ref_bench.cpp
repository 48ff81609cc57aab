// ref_bench.cpp -- TEST/BENCH INFRASTRUCTURE ONLY: times the UNMODIFIED reference (libtfhe,
// spqlios-fma) on the host cores: one gate stream per thread against one shared cloud key
// (libtfhe's FFT state is thread_local, libs/tfhe/src/libtfhe/fft_processors/spqlios/fft_processor_spqlios.cpp:123).
//
//   ref_bench <lambda 80|128> <gate NAND|MUX|...> <threads> <gates-per-thread> [seed]
//
// Inputs are fresh encryptions of random bits (encrypted before the clock starts); every output is
// decrypted after the clock stops and compared with the boolean model, like the reference's own
// integration tests (libs/tfhe/src/test/test-gate-bootstrapping.cpp:56-93).
// Prints one JSON line: {"gates": G, "threads": T, "seconds": S, "gates_per_s": R, "ms_per_gate_per_thread": M,
//                        "errors": E, "lambda": L, "gate": "..."}
#include <tfhe.h>
#include <tfhe_io.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

int main(int argc, char **argv) {
    if (argc < 5) {
        std::fprintf(stderr, "usage: %s <lambda> <gate> <threads> <gates-per-thread> [seed]\n", argv[0]);
        return 2;
    }
    const int lambda = std::atoi(argv[1]);
    const std::string gate = argv[2];
    const int threads = std::max(1, std::atoi(argv[3]));
    const int per = std::max(1, std::atoi(argv[4]));
    uint32_t seed = argc > 5 ? (uint32_t)std::atoi(argv[5]) : 42u;
    tfhe_random_generator_setSeed(&seed, 1);

    TFheGateBootstrappingParameterSet *params = new_default_gate_bootstrapping_parameters(lambda);
    TFheGateBootstrappingSecretKeySet *sk = new_random_gate_bootstrapping_secret_keyset(params);
    const TFheGateBootstrappingCloudKeySet *ck = &sk->cloud;

    const int total = threads * per;
    LweSample *a = new_gate_bootstrapping_ciphertext_array(total, params);
    LweSample *b = new_gate_bootstrapping_ciphertext_array(total, params);
    LweSample *c = new_gate_bootstrapping_ciphertext_array(total, params);
    LweSample *o = new_gate_bootstrapping_ciphertext_array(total, params);
    std::mt19937 rng(1234);
    std::vector<int> ba(total), bb(total), bc(total);
    for (int i = 0; i < total; ++i) {
        ba[i] = rng() & 1; bb[i] = rng() & 1; bc[i] = rng() & 1;
        bootsSymEncrypt(a + i, ba[i], sk);
        bootsSymEncrypt(b + i, bb[i], sk);
        bootsSymEncrypt(c + i, bc[i], sk);
    }
    auto run = [&](int t) {
        for (int i = t * per; i < (t + 1) * per; ++i) {
            if (gate == "MUX") bootsMUX(o + i, a + i, b + i, c + i, ck);
            else if (gate == "AND") bootsAND(o + i, a + i, b + i, ck);
            else if (gate == "OR") bootsOR(o + i, a + i, b + i, ck);
            else if (gate == "XOR") bootsXOR(o + i, a + i, b + i, ck);
            else bootsNAND(o + i, a + i, b + i, ck);
        }
    };
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back(run, t);
    for (auto &th : pool) th.join();
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    int errors = 0;
    for (int i = 0; i < total; ++i) {
        int want;
        if (gate == "MUX") want = ba[i] ? bb[i] : bc[i];
        else if (gate == "AND") want = ba[i] & bb[i];
        else if (gate == "OR") want = ba[i] | bb[i];
        else if (gate == "XOR") want = ba[i] ^ bb[i];
        else want = 1 - (ba[i] & bb[i]);
        if (bootsSymDecrypt(o + i, sk) != want) ++errors;
    }
    std::printf("{\"gates\": %d, \"threads\": %d, \"seconds\": %.6f, \"gates_per_s\": %.3f, "
                "\"ms_per_gate_per_thread\": %.4f, \"errors\": %d, \"lambda\": %d, \"gate\": \"%s\"}\n",
                total, threads, secs, total / secs, secs * 1e3 / per, errors, lambda, gate.c_str());
    return errors ? 1 : 0;
}
