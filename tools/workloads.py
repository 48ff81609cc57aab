#!/usr/bin/env python
"""BASELINE.json's secondary workloads at full size on one GPU, checked and timed (one JSON line each):
  configs[2]  4,096 pairs of 16-bit encrypted integers through the ripple-carry adder netlist of
              test-addition-boot.cpp (depth-batched levels), sums decrypted and compared with x + y
  configs[3]  2^18 bootsMUX at the n=630 set, decrypted bits compared with a ? b : c
Keys, encryption and decryption come from the unmodified reference (oracle/_ref, test infrastructure);
the timed region is the GPU circuit / gate call only (device-resident inputs).
usage: python tools/workloads.py [adder|mux|all]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import reflib
from eoc_tfhe_b200 import CloudKey, GateOp, circuits

def phase_bits(cts, key_t):
    n = key_t.shape[0]
    acc = (cts[:, :n].to(torch.int64) * key_t).sum(dim=1)
    ph = (cts[:, n].to(torch.int64) - acc) & 0xFFFFFFFF
    return (torch.where(ph >= 2 ** 31, ph - 2 ** 32, ph) > 0).to(torch.int32)

def adder():
    ref = reflib.RefLib(); rk = ref.keygen(80, seed=99); ck = CloudKey.from_libtfhe(rk.cloud_ptr)
    nb, count = 16, 4096
    nl = circuits.full_adder(nb)
    rng = np.random.default_rng(99)      # operands as in SURVEY 8d config 3
    x = rng.integers(0, 1 << nb, size=count); y = rng.integers(0, 1 << nb, size=count)
    cx = rk.encrypt_bits(circuits.int_to_bits(x, nb).ravel()).reshape(count, nb, -1)
    cy = rk.encrypt_bits(circuits.int_to_bits(y, nb).ravel()).reshape(count, nb, -1)
    circuits.run_batch(ck, nl, {"x": cx[:64], "y": cy[:64]}, 64)   # warm-up
    torch.cuda.synchronize(); t0 = time.perf_counter()
    outs, levels = circuits.run_batch(ck, nl, {"x": cx, "y": cy}, count)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    s = rk.decrypt_bits(outs["sum"].reshape(count * (nb + 1), -1)).reshape(count, nb + 1)
    ok = bool(np.array_equal(circuits.bits_to_int(s), x + y))
    boots = sum(1 for g in nl.gates if GateOp(g[0]) not in (GateOp.CONST0, GateOp.CONST1, GateOp.COPY, GateOp.NOT))
    print(json.dumps({"workload": "addCiphertexts: 4096 pairs x 16-bit ripple-carry adder (test-addition-boot.cpp full_adder netlist)",
                      "bootstrapped_gates": boots * count, "levels": levels, "seconds": dt, "gates_per_s": boots * count / dt,
                      "sums_correct": ok, "includes": "host->device upload of the 2 x 4096 x 16 input ciphertexts and download of the sums"}), flush=True)
    ck.close()
    return ok

def mux():
    ref = reflib.RefLib(); rk = ref.keygen(128, seed=7); ck = CloudKey.from_libtfhe(rk.cloud_ptr)
    rng = np.random.default_rng(7)
    pool_bits = rng.integers(0, 2, size=1024).astype(np.int32)
    pool = torch.from_numpy(rk.encrypt_bits(pool_bits)).cuda()
    count = 1 << 18
    idx = [torch.from_numpy(rng.integers(0, 1024, size=count)).cuda() for _ in range(3)]
    a, b, c = (pool[i].contiguous() for i in idx)
    out = torch.empty_like(a)
    ck.gate(GateOp.MUX, a[:4096], b[:4096], c[:4096], out=out[:4096]); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ck.gate(GateOp.MUX, a, b, c, out=out); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    key_t = torch.from_numpy(rk.lwe_key().astype(np.int64)).cuda()
    got = torch.cat([phase_bits(out[i:i + 65536], key_t) for i in range(0, count, 65536)]).cpu().numpy()
    ba, bb, bc = (pool_bits[i.cpu().numpy()] for i in idx)
    ok = bool(np.array_equal(got, np.where(ba > 0, bb, bc)))
    print(json.dumps({"workload": "2^18 bootsMUX, n=630 N=1024 l=3 Bgbit=7 (2^19 blind rotations + 2^18 key switches)",
                      "gates": count, "ms": ms, "mux_per_s": count / ms * 1e3, "bootstraps_per_s": 2 * count / ms * 1e3,
                      "fp64_tflops_algorithmic": 2 * count * 162.6e6 / (ms * 1e-3) / 1e12, "bits_correct": ok}), flush=True)
    ck.close()
    return ok

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    good = True
    if which in ("adder", "all"): good &= adder()
    if which in ("mux", "all"): good &= mux()
    sys.exit(0 if good else 1)
