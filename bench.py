"""bench.py -- bootstrapped NAND gates/s on N B200s (BASELINE.json metric), one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): one step = 65,536 independent
bootsNAND gates per GPU at the 80-bit default set (n=500, N=1024, k=1, l=2, Bgbit=10, t=8, basebit=2)
on synthetic uniformly random ciphertexts and a random cloud key (the path's run time is data
independent apart from the 1/2048 barai==0 skip, lwe-bootstrapping-functions-fft.cpp:101).  Gates
shard by rank with no data-path collective; the cloud key is created on rank 0 and replicated with
one NCCL broadcast (outside the timed region); every step ends with one all-gather of the outputs.

  value     gates/s, inputs resident in HBM, timed with CUDA events on the launching stream
  e2e       gates/s through the C ABI host entry (b200_gate_batch_host): pinned host buffers,
            H2D of both input batches and D2H of the output batch inside the timed region
  roofline  the dominant kernel (k_blind_rotate) against the FP64 FMA peak measured live by a
            DFMA-chain kernel of this library; the key-switch kernel against measured HBM bandwidth
  cpu_baseline  the UNMODIFIED reference (oracle/_ref/ref_bench, libtfhe spqlios-fma) on the host
            cores, one gate stream per core, bounded sample
`--impl reference` times only that CPU reference and prints the same line shape.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GATES_PER_GPU = 65536
METRIC = "bootstrapped_nand_gates_per_s"
UNIT = "gates/s"
# per CMux step (kpl + k+1) transforms x (5 M log2 M + 6 M) flop, M = 512, plus
# kpl (k+1) M complex FMAs x 8 flop; 80-bit set: 189,440 flop x n=500 steps
FLOP_PER_BOOTSTRAP_P80 = 189440 * 500
WORKLOAD = "65536 independent bootsNAND per GPU, 80-bit default params (n=500,N=1024,k=1,l=2,Bgbit=10)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gates", type=int, default=GATES_PER_GPU, help="gates per GPU per step (default: the BASELINE config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel", default=None, choices=["auto", "v2", "tm", "tmacc", "tms", "w16"], help="blind-rotation kernel (default: the library's)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# CPU reference arm (the unmodified libtfhe compiled by oracle/Makefile)
# ------------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` in this same workload, from the
    committed ncu capture (profiles/traffic.json, written by tools/ncu_traffic.py); None when not captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(kernel, {}).get("dram_bytes")
    except Exception:
        return None


def run_ref_bench(threads, per_thread, lam=80, gate="NAND"):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
    if not os.path.exists(exe):
        return None
    r = subprocess.run([exe, str(lam), gate, str(threads), str(per_thread)], stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=900)
    for line in r.stdout.splitlines():
        if line.startswith("{"):
            return json.loads(line)
    return None


def run_port_bench(threads, per_thread):
    """Fallback when oracle/_ref is absent: the C restatement (oracle/tfhe_port.c) on a random key."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import portlib
    rng = np.random.default_rng(0)
    n = 500
    bk = rng.integers(-2 ** 31, 2 ** 31, size=(n, 4, 2, 1024), dtype=np.int64).astype(np.int32)
    ks = rng.integers(-2 ** 31, 2 ** 31, size=(1024, 8, 4, n + 1), dtype=np.int64).astype(np.int32)
    key = portlib.PortKey(n, 2, 10, bk, ks)
    count = threads * per_thread
    a = rng.integers(-2 ** 31, 2 ** 31, size=(count, n + 1), dtype=np.int64).astype(np.int32)
    b = rng.integers(-2 ** 31, 2 ** 31, size=(count, n + 1), dtype=np.int64).astype(np.int32)
    t0 = time.perf_counter()
    key.gate("NAND", a, b, threads=threads)
    dt = time.perf_counter() - t0
    key.close()
    return {"gates": count, "threads": threads, "seconds": dt, "gates_per_s": count / dt, "errors": 0}


def cpu_reference(per_thread):
    cores = host_cores()
    res = run_ref_bench(cores, per_thread)
    kind = "reference"
    if res is None:
        res = run_port_bench(cores, per_thread)
        kind = "port"
    return {"value": res["gates_per_s"], "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%d bootsNAND (%d per thread x %d threads, one gate stream per core, 80-bit set, "
                      "%s), %.2f s wall" % (res["gates"], per_thread, cores,
                                            "libtfhe spqlios-fma from oracle/_ref" if kind == "reference" else "oracle/tfhe_port.c",
                                            res["seconds"]),
            "_seconds": res["seconds"], "_gates": res["gates"]}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    per_thread = 64
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_reference(2)
    secs, gates, last = 0.0, 0, None
    for _ in range(args.steps):
        last = cpu_reference(per_thread)
        secs += last["_seconds"]
        gates += last["_gates"]
    value = gates / secs
    cb = {k: v for k, v in last.items() if not k.startswith("_")}
    cb["value"] = value
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample_per_step": "%d gates per host thread" % per_thread},
            "cpu_baseline": cb,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons of one GPU sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.thread.join(timeout=2)
        mhz, mx, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                mhz.append(float(r[0]))
                mx = float(r[1])
                power.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        mhz.sort()
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(mhz), "power_w_max": max(power) if power else None}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def main_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from eoc_tfhe_b200 import CloudKey, GateOp, PARAMS_80, launch_count, fp64_peak_tflops, set_blind_rotate_kernel, shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the GPU path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # stdout carries exactly one JSON line: keep NCCL's version banner (NCCL_DEBUG=VERSION/INFO) off it
        os.environ["NCCL_DEBUG"] = os.environ.get("B200_NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=dev)
    if args.kernel:
        set_blind_rotate_kernel(args.kernel)
    P = PARAMS_80
    count = args.gates
    n1 = P.n + 1

    # ---- cloud key: created (uploaded + converted to the FFT domain on the GPU) on rank 0, broadcast over NCCL
    t0 = time.perf_counter()
    if rank == 0:
        rng = np.random.default_rng(0)
        bk = rng.integers(-2 ** 31, 2 ** 31, size=P.bk_shape, dtype=np.int64).astype(np.int32)
        ks = rng.integers(-2 ** 31, 2 ** 31, size=P.ks_shape, dtype=np.int64).astype(np.int32)
        ck = CloudKey.from_arrays(P, bk, ks, device=local)
        del bk, ks
    else:
        ck = CloudKey.empty(P, device=local)
    torch.cuda.synchronize()
    key_upload_s = time.perf_counter() - t0
    key_bcast_ms = 0.0
    if world > 1:
        dist.barrier()
        key_bcast_ms = shard.broadcast_key(ck, src=0)     # one NCCL broadcast per resident blob

    # ---- inputs: synthetic uniformly random ciphertexts, resident in HBM; 2 x 131 MB per GPU > 126 MB L2
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    a = torch.randint(-2 ** 31, 2 ** 31 - 1, (count, n1), dtype=torch.int32, device=dev, generator=g)
    b = torch.randint(-2 ** 31, 2 ** 31 - 1, (count, n1), dtype=torch.int32, device=dev, generator=g)
    out = torch.empty_like(a)
    gathered = torch.empty((world * count, n1), dtype=torch.int32, device=dev) if world > 1 else None

    def step():
        ck.gate(GateOp.NAND, a, b, out=out)            # two launches on torch's current stream
        if world > 1:
            dist.all_gather_into_tensor(gathered, out)  # results gathered once per level

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    ck.timing_drain()          # forget the warm-up launches
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    br_ms, ks_ms = [], []
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = launch_count() - l0
    total_ms = e0.elapsed_time(e1)
    tm = ck.timing_drain()     # CUDA events around the two kernels of every timed step, same stream
    br_ms.append(tm["blind_rotate"] / max(1, tm["calls"]))
    ks_ms.append(tm["keyswitch"] / max(1, tm["calls"]))
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * count / (ms_per_step * 1e-3)

    # ---- end to end through the C ABI host entry: pinned host buffers, copies inside the timed region
    ha = torch.empty((count, n1), dtype=torch.int32).pin_memory()
    hb = torch.empty((count, n1), dtype=torch.int32).pin_memory()
    ho = torch.empty((count, n1), dtype=torch.int32).pin_memory()
    ha.copy_(a)
    hb.copy_(b)
    na, nb, no = ha.numpy(), hb.numpy(), ho.numpy()
    e2e_steps = max(1, min(args.steps, 3))
    ck.gate(GateOp.NAND, na, nb, out=no)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ck.gate(GateOp.NAND, na, nb, out=no)   # synchronous: returns after the D2H copy completed
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * count * e2e_steps / float(t.item())
    same = bool(torch.equal(ho.to(dev), out))   # host-path and device-path results are identical

    # ---- single-gate latency (second half of BASELINE's metric): one bootsNAND through the host entry
    lat_ms = None
    if rank == 0:
        one_a, one_b, one_o = na[:1].copy(), nb[:1].copy(), no[:1].copy()
        for _ in range(3):
            ck.gate(GateOp.NAND, one_a, one_b, out=one_o)
        samples = []
        for _ in range(15):
            t0 = time.perf_counter()
            ck.gate(GateOp.NAND, one_a, one_b, out=one_o)
            samples.append((time.perf_counter() - t0) * 1e3)
        lat_ms = sorted(samples)[len(samples) // 2]

    if rank == 0:
        hbm_peak, hbm_src = measured_peaks()
        fp64_peak = fp64_peak_tflops(local)
        br = sum(br_ms) / len(br_ms)
        ksm = sum(ks_ms) / len(ks_ms)
        achieved = count * FLOP_PER_BOOTSTRAP_P80 / (br * 1e-3) / 1e12
        ks_bytes = count * (6144 * n1 * 4 + (1024 + 1) * 4 + n1 * 4)  # 6144 table rows gathered per gate
        # bootstrapping-key streaming (north_star: "achieved HBM/L2 GB/s on bootstrapping-key streaming"): every CTA of
        # the blind rotation pulls the whole FFT-domain key (n x 2l row pairs of 16 KiB) from L2 through its TMA ring,
        # once for the 6 ciphertexts it holds; the grid is whole waves of one CTA per SM (kernels.cu, br_launch_st)
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        ctas = -(-(-(-count // 6)) // sms) * sms
        key_bytes = P.n * 2 * P.l * 16384
        tr = ncu_traffic("k_blind_rotate")
        key_stream = {"kernel": "k_blind_rotate", "path": "L2 -> shared memory, cp.async.bulk ring", "ctas": ctas,
                      "key_bytes": key_bytes, "bytes_per_launch": ctas * key_bytes,
                      "achieved": ctas * key_bytes / (br * 1e-3) / 1e9, "unit": "GB/s",
                      "unamortised": count * key_bytes / (br * 1e-3) / 1e9,
                      "hbm_peak": hbm_peak, "frac_of_hbm_peak": ctas * key_bytes / (br * 1e-3) / 1e9 / hbm_peak,
                      "dram_traffic_per_launch": tr,
                      "note": "one key read serves 6 ciphertexts; the key (32.8 MB) stays L2 resident, so DRAM sees it about once per launch (traffic = inputs + extracted samples + key, ncu); 'unamortised' is what one read per ciphertext would need"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "gates_per_gpu_per_step": count, "parallelism": "gates sharded by rank, key replicated (NCCL broadcast), outputs all-gathered per step",
                       "l2": "input batches 2 x %.0f MB per GPU exceed the 126 MB L2; the 32.8 MB FFT-domain key is meant to stay L2 resident" % (count * n1 * 4 / 1e6)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * count * n1 * 4, "d2h_bytes_per_step": count * n1 * 4,
                    "steps": e2e_steps, "api": "b200_gate_batch_host (pinned host buffers)", "matches_device_path": same},
            "gpu_launches": int(launches),
            "latency": {"single_gate_ms": lat_ms, "api": "b200_gate_batch_host, count = 1 (H2D, two kernels, D2H, synchronous)"},
            "roofline": {"bound": "fp64", "kernel": "k_blind_rotate", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp64_peak if fp64_peak else None, "traffic": ncu_traffic("k_blind_rotate"),
                         "ms_per_launch": br, "share_of_step": br / ms_per_step,
                         "peak_source": "DFMA-chain kernel of this library timed live (MEASURED_PEAKS.json has no FP64 figure)",
                         "algorithmic_flop_per_bootstrap": FLOP_PER_BOOTSTRAP_P80},
            "roofline_keyswitch": {"bound": "hbm", "kernel": "k_keyswitch_tma", "achieved": ks_bytes / (ksm * 1e-3) / 1e9, "peak": hbm_peak,
                                   "unit": "GB/s", "frac": ks_bytes / (ksm * 1e-3) / 1e9 / hbm_peak, "traffic": ncu_traffic("k_keyswitch_tma"),
                                   "ms_per_launch": ksm, "share_of_step": ksm / ms_per_step, "peak_source": hbm_src,
                                   "note": "algorithmic bytes = naive per-gate gather of 6144 rows (SURVEY 8d); the kernel streams the table once per 64 samples through shared memory, so frac > 1 is expected: it is instruction-issue bound, see DESIGN.md 4.4"},
            "key_stream": key_stream,
            "clocks": clocks,
            "key": {"upload_convert_s": key_upload_s, "nccl_broadcast_ms": key_bcast_ms,
                    "bytes": sum(n for _, n in ck.device_blobs())},
        }
        if not args.no_cpu_baseline and world == 1:
            cb = cpu_reference(96)   # ~1.2 s wall on every host core: 15-20 core-seconds of libtfhe
            line["cpu_baseline"] = {k: v for k, v in cb.items() if not k.startswith("_")}
            one = run_ref_bench(1, 16)
            if one:
                line["latency"]["cpu_single_gate_ms"] = one["ms_per_gate_per_thread"]
        print(json.dumps(line), flush=True)
    ck.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    _args = parse_args()
    sys.exit(main_reference(_args) if _args.impl == "reference" else main_b200(_args))
