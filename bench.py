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
    ap.add_argument("--no-p128", action="store_true", help="skip the 2^18-MUX / 128-bit secondary workload (N = 1)")
    ap.add_argument("--no-config5", action="store_true", help="skip the 2^20-gate sharded batch (N > 1)")
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


def run_ref_bench(threads, per_thread, lam=80, gate="NAND", variant="spqlios-fma"):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench" if variant == "spqlios-fma" else "ref_bench_nayuki")
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


# gates per host thread of one CPU sample: at least 3 s of wall time per point on hosts where libtfhe
# spqlios-fma takes 9-16 ms per NAND and thread; the nayuki portable processor is about 8x slower, a 128-bit MUX 4x
CPU_SAMPLE = {("spqlios-fma", 80, "NAND"): 320, ("nayuki-portable", 80, "NAND"): 40, ("spqlios-fma", 128, "MUX"): 80}
CPU_NAMES = {"spqlios-fma": "libtfhe spqlios-fma (the fastest FFT processor in the reference tree) from oracle/_ref",
             "nayuki-portable": "libtfhe nayuki-portable (the FFT processor eoc ships, config.yml:21-25) from oracle/_ref"}


def cpu_reference(per_thread=None, variant="spqlios-fma", lam=80, gate="NAND"):
    """The UNMODIFIED reference on every host core, one gate stream per core; repeats with more gates until the
    sample lasted at least 3 s."""
    cores = host_cores()
    per_thread = per_thread or CPU_SAMPLE[(variant, lam, gate)]
    kind = "reference"
    for _ in range(3):
        res = run_ref_bench(cores, per_thread, lam, gate, variant)
        if res is None or res["seconds"] >= 3.0:
            break
        per_thread = int(per_thread * 3.3 / max(res["seconds"], 0.05)) + 1
    if res is None:
        if variant != "spqlios-fma" or lam != 80 or gate != "NAND":
            return None
        res = run_port_bench(cores, 8)
        kind = "port"
    return {"value": res["gates_per_s"], "unit": UNIT if gate == "NAND" else "MUX gates/s", "cores": cores, "kind": kind,
            "sample": "%d boots%s (%d per thread x %d threads, one gate stream per core, %d-bit set, %s), %.2f s wall, %d wrong bits"
                      % (res["gates"], gate, res["gates"] // cores, cores, lam,
                         CPU_NAMES[variant] if kind == "reference" else "oracle/tfhe_port.c", res["seconds"], res.get("errors", 0)),
            "_seconds": res["seconds"], "_gates": res["gates"]}


def bench_config(world, count):
    """The `config` object: identical in the GPU arm and in the reference arm."""
    return {"workload": WORKLOAD, "gates_per_gpu_per_step": count,
            "parallelism": "gates sharded by rank, key replicated (NCCL broadcast), outputs all-gathered per step",
            "l2": "input batches 2 x %.0f MB per GPU exceed the 126 MB L2; the 32.8 MB FFT-domain key is meant to stay L2 resident"
                  % (count * 501 * 4 / 1e6)}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    for _ in range(max(0, min(args.warmup, 1))):
        run_ref_bench(host_cores(), 2)
    secs, gates, last = 0.0, 0, None
    for _ in range(args.steps):
        last = cpu_reference()
        secs += last["_seconds"]
        gates += last["_gates"]
    value = gates / secs
    cb = {k: v for k, v in last.items() if not k.startswith("_")}
    cb["value"] = value
    cb["sample"] += " (last of %d steps; a step of this arm is one such sample of the workload, the CPU rate does not depend on the batch size)" % args.steps
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(args.gpus, args.gates),
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


class NcclLog:
    """NCCL's INFO log is left on (the driver reads the communicator size from it) without letting it reach this
    process's stdout, which carries exactly one JSON line: before torch is imported, file descriptor 1 is pointed at
    a per-process file (NCCL logs to the C stdout unless NCCL_DEBUG_FILE says otherwise), the JSON line is written to
    the saved descriptor, and rank 0 relays the lines that show the communicator (ranks, transports, NVLS) to stderr
    after the process group is gone."""

    def __init__(self):
        if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
            os.environ["NCCL_DEBUG"] = "INFO"          # a preset VERSION / WARN level hides the communicator lines
        os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT,ENV")
        self.debug_file = os.environ.get("NCCL_DEBUG_FILE")       # respected when the launcher set one
        sys.stdout.flush()
        self.real_stdout = os.dup(1)
        self.path = "/tmp/b200_bench_stdout_%d.log" % os.getpid()
        fd = os.open(self.path, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o600)
        os.dup2(fd, 1)
        os.close(fd)

    def emit(self, text):
        os.write(self.real_stdout, (text + "\n").encode())

    def relay(self, world):
        import ctypes
        import glob
        import re
        sys.stdout.flush()
        ctypes.CDLL(None).fflush(None)
        files = [self.path]
        if self.debug_file:
            files += glob.glob(self.debug_file.replace("%h", "*").replace("%p", str(os.getpid())))
        out = {"nranks": None, "lines": 0, "nvls": False}
        for f in files:
            try:
                with open(f, errors="replace") as fh:
                    for line in fh:
                        if "NCCL" not in line:
                            continue
                        m = re.search(r"nranks (\d+)", line)
                        if m:
                            out["nranks"] = int(m.group(1))
                        if "NVLS" in line and "comm" in line:
                            out["nvls"] = True
                        if re.search(r"nranks|Init COMPLETE|NVLS|Channel 00|via P2P|NCCL version", line):
                            out["lines"] += 1
                            if out["lines"] <= 40:
                                sys.stderr.write(line if line.endswith("\n") else line + "\n")
            except OSError:
                pass
        sys.stderr.flush()
        out["comm_nranks_ok"] = out["nranks"] == world
        return out


def main_b200(args):
    nccl_log = NcclLog() if int(os.environ.get("WORLD_SIZE", "1")) > 1 else None      # before torch loads NCCL
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the GPU path has no CPU fallback; use --impl reference for the CPU arm)")
    from eoc_tfhe_b200 import (CloudKey, GateOp, PARAMS_80, PARAMS_128, launch_count, fp64_peak_tflops, int32_add_peak,
                               set_blind_rotate_kernel, shard)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
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

    def make_inputs(seed, rows):
        g = torch.Generator(device=dev)
        g.manual_seed(seed)
        x = torch.randint(-2 ** 31, 2 ** 31 - 1, (rows, n1), dtype=torch.int32, device=dev, generator=g)
        y = torch.randint(-2 ** 31, 2 ** 31 - 1, (rows, n1), dtype=torch.int32, device=dev, generator=g)
        return x, y

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        """W untimed + K timed steps, CUDA events on the launching stream, barrier + synchronize on both sides,
        max over ranks; returns ms per step."""
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    # ---- inputs: synthetic uniformly random ciphertexts, resident in HBM; 2 x 131 MB per GPU > 126 MB L2
    a, b = make_inputs(1234 + rank, count)
    out = torch.empty_like(a)
    gathered = torch.empty((world * count, n1), dtype=torch.int32, device=dev) if world > 1 else None

    def step():
        ck.gate(GateOp.NAND, a, b, out=out)            # two launches on torch's current stream
        if world > 1:
            dist.all_gather_into_tensor(gathered, out)  # results gathered once per level

    for _ in range(args.warmup):
        step()
    barrier()
    ck.timing_drain()          # forget the warm-up launches
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = launch_count()
    ms_per_step = timed(step, args.steps, 0)
    launches = launch_count() - l0
    tm = ck.timing_drain()     # CUDA events around the two kernels of every timed step, same stream
    br = tm["blind_rotate"] / max(1, tm["calls"])
    ksm = tm["keyswitch"] / max(1, tm["calls"])
    clocks = sampler.stop() if rank == 0 else None
    value = world * count / (ms_per_step * 1e-3)

    # ---- N > 1: the sharded result is what each rank computed (outside the timed region)
    sharded_equal = None
    if world > 1:
        ok = bool(torch.equal(gathered[rank * count:(rank + 1) * count], out))
        # and rank r's slice is what THIS rank gets when it evaluates rank r's inputs itself (r = the next rank)
        r = (rank + 1) % world
        ra, rb = make_inputs(1234 + r, count)
        ro = ck.gate(GateOp.NAND, ra, rb)
        ok = ok and bool(torch.equal(gathered[r * count:(r + 1) * count], ro))
        del ra, rb, ro
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        sharded_equal = bool(t.item())
    # ---- end to end through the C ABI host entry: pinned host buffers, copies inside the timed region, same K steps
    ha = torch.empty((count, n1), dtype=torch.int32).pin_memory()
    hb = torch.empty((count, n1), dtype=torch.int32).pin_memory()
    ho = torch.empty((count, n1), dtype=torch.int32).pin_memory()
    ha.copy_(a)
    hb.copy_(b)
    na, nb, no = ha.numpy(), hb.numpy(), ho.numpy()
    for _ in range(min(args.warmup, 2)):
        ck.gate(GateOp.NAND, na, nb, out=no)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ck.gate(GateOp.NAND, na, nb, out=no)   # synchronous: returns after the D2H copy completed
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * count * args.steps / float(t.item())
    same = bool(torch.equal(ho.to(dev), out))   # host-path and device-path results are identical
    del ha, hb, ho

    # ---- BASELINE configs[4]: ONE batch of 2^20 NAND gates sharded contiguously over the ranks, the all-gather of
    # the outputs inside the time, the key already resident (its upload / broadcast is reported separately)
    config5 = None
    if world > 1 and not args.no_config5:
        total5 = 1 << 20
        c5 = total5 // world
        a5, b5 = make_inputs(777 + rank, c5)
        o5 = torch.empty_like(a5)
        g5 = torch.empty((total5, n1), dtype=torch.int32, device=dev)

        def step5():
            ck.gate(GateOp.NAND, a5, b5, out=o5)
            dist.all_gather_into_tensor(g5, o5)
        ms5 = timed(step5, 2, 1)
        ok5 = bool(torch.equal(g5[rank * c5:(rank + 1) * c5], o5))
        config5 = {"workload": "2^20 bootsNAND in one batch, sharded contiguously by rank, outputs all-gathered (BASELINE configs[4])",
                   "gates": total5, "gates_per_rank": c5, "ms": ms5, "value": total5 / (ms5 * 1e-3), "unit": UNIT,
                   "steps": 2, "warmup": 1, "scaling": "strong", "own_slice_intact_on_rank0": ok5,
                   "key_excluded": {"upload_convert_s": key_upload_s, "nccl_broadcast_ms": key_bcast_ms}}
        del a5, b5, o5, g5

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
        # the same call from 4 host threads at once (one gate stream per thread, like libtfhe's users): lanes of the key
        from concurrent.futures import ThreadPoolExecutor

        def stream(_):
            xa, xb, xo = one_a.copy(), one_b.copy(), one_o.copy()
            for _ in range(40):
                ck.gate(GateOp.NAND, xa, xb, out=xo)
        with ThreadPoolExecutor(max_workers=4) as ex:
            list(ex.map(stream, range(4)))
            t0 = time.perf_counter()
            list(ex.map(stream, range(4)))
            lat_threads4 = 160 / (time.perf_counter() - t0)

    # ---- BASELINE configs[3] on one GPU: 2^18 bootsMUX at the 128-bit set (n=630, l=3, Bgbit=7), incl. key switch
    p128 = None
    if rank == 0 and world == 1 and not args.no_p128:
        P2 = PARAMS_128
        rng = np.random.default_rng(1)
        ck2 = CloudKey.from_arrays(P2, rng.integers(-2 ** 31, 2 ** 31, size=P2.bk_shape, dtype=np.int64).astype(np.int32),
                                   rng.integers(-2 ** 31, 2 ** 31, size=P2.ks_shape, dtype=np.int64).astype(np.int32), device=local)
        cnt2 = 1 << 18
        g = torch.Generator(device=dev)
        g.manual_seed(99)
        xa, xb, xc = (torch.randint(-2 ** 31, 2 ** 31 - 1, (cnt2, P2.n + 1), dtype=torch.int32, device=dev, generator=g) for _ in range(3))
        xo = torch.empty_like(xa)
        ms2 = timed(lambda: ck2.gate(GateOp.MUX, xa, xb, xc, out=xo), 1, 1)
        p128 = {"workload": "2^18 bootsMUX on 1 GPU, 128-bit params (n=630,N=1024,l=3,Bgbit=7), two blind rotations + one key switch per gate (BASELINE configs[3])",
                "gates": cnt2, "ms": ms2, "value": cnt2 / (ms2 * 1e-3), "unit": "MUX gates/s", "bootstraps_per_s": 2 * cnt2 / (ms2 * 1e-3),
                "steps": 1, "warmup": 1}
        ck2.close()
        del xa, xb, xc, xo

    if rank == 0:
        hbm_peak, hbm_src = measured_peaks()
        fp64_peak = fp64_peak_tflops(local)
        iadd_peak = int32_add_peak(local)
        achieved = count * FLOP_PER_BOOTSTRAP_P80 / (br * 1e-3) / 1e12
        # key switch: 1024 coefficients x 8 digits, 3 of 4 non-zero, each one row of n+1 words subtracted
        ks_adds = count * 1024 * 8 * 0.75 * n1
        # bootstrapping-key streaming (north_star: "achieved HBM/L2 GB/s on bootstrapping-key streaming"): every CTA of
        # the blind rotation pulls the whole FFT-domain key (n x 2l row pairs of 16 KiB) from L2 through its TMA ring,
        # once for the 6 ciphertexts it holds; the grid is whole waves of one CTA per SM (kernels.cu, br_launch_st)
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        ctas = -(-(-(-count // 6)) // sms) * sms
        key_bytes = P.n * 2 * P.l * 16384
        key_stream = {"kernel": "k_blind_rotate", "path": "L2 -> shared memory, cp.async.bulk ring", "ctas": ctas,
                      "key_bytes": key_bytes, "bytes_per_launch": ctas * key_bytes,
                      "achieved": ctas * key_bytes / (br * 1e-3) / 1e9, "unit": "GB/s",
                      "unamortised": count * key_bytes / (br * 1e-3) / 1e9,
                      "hbm_peak": hbm_peak, "frac_of_hbm_peak": ctas * key_bytes / (br * 1e-3) / 1e9 / hbm_peak,
                      "dram_traffic_per_launch": ncu_traffic("k_blind_rotate"),
                      "note": "one key read serves 6 ciphertexts; the key (32.8 MB) stays L2 resident, so DRAM sees it about once per launch (traffic = inputs + extracted samples + key, ncu); 'unamortised' is what one read per ciphertext would need"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": bench_config(world, count),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * count * n1 * 4, "d2h_bytes_per_step": count * n1 * 4,
                    "steps": args.steps, "api": "b200_gate_batch_host (pinned host buffers)", "matches_device_path": same},
            "gpu_launches": int(launches),
            "latency": {"single_gate_ms": lat_ms, "api": "b200_gate_batch_host, count = 1 (H2D, two kernels, D2H, synchronous)",
                        "single_gate_streams_4_host_threads_gates_per_s": lat_threads4},
            "roofline": {"bound": "fp64", "kernel": "k_blind_rotate<2,10,6,4,FftTmAcc>", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp64_peak if fp64_peak else None, "traffic": ncu_traffic("k_blind_rotate"),
                         "ms_per_launch": br, "share_of_step": br / ms_per_step,
                         "peak_source": "DFMA-chain kernel of this library timed live (MEASURED_PEAKS.json has no FP64 figure)",
                         "algorithmic_flop_per_bootstrap": FLOP_PER_BOOTSTRAP_P80},
            "roofline_keyswitch": {"bound": "int32 add issue", "kernel": "k_keyswitch_tma<64,8,8,2,4,126>", "achieved": ks_adds / (ksm * 1e-3) / 1e12,
                                   "peak": iadd_peak, "unit": "Tadd/s", "frac": ks_adds / (ksm * 1e-3) / 1e12 / iadd_peak if iadd_peak else None,
                                   "traffic": ncu_traffic("k_keyswitch_tma"), "ms_per_launch": ksm, "share_of_step": ksm / ms_per_step,
                                   "peak_source": "int32 add-chain kernel of this library timed live",
                                   "algorithmic_adds_per_gate": 1024 * 8 * 0.75 * n1,
                                   "note": "the table (49.5 MB, L2 resident) is streamed once per 64 samples, so HBM is not the bound (traffic = ncu DRAM bytes); the work is 32-bit subtractions, reported against the measured add issue rate"},
            "key_stream": key_stream,
            "clocks": clocks,
            "key": {"upload_convert_s": key_upload_s, "nccl_broadcast_ms": key_bcast_ms,
                    "bytes": sum(n for _, n in ck.device_blobs())},
        }
        if world > 1:
            line["sharded_equal"] = sharded_equal
        if config5:
            line["config5"] = config5
        if p128:
            line["p128_mux"] = p128
        if not args.no_cpu_baseline and world == 1:
            cb = cpu_reference()
            line["cpu_baseline"] = {k: v for k, v in cb.items() if not k.startswith("_")}
            cn = cpu_reference(variant="nayuki-portable")
            if cn:
                line["cpu_baseline_nayuki"] = {k: v for k, v in cn.items() if not k.startswith("_")}
            if p128:
                cm = cpu_reference(variant="spqlios-fma", lam=128, gate="MUX")
                if cm:
                    p128["cpu_baseline"] = {k: v for k, v in cm.items() if not k.startswith("_")}
            one = run_ref_bench(1, 16)
            if one:
                line["latency"]["cpu_single_gate_ms"] = one["ms_per_gate_per_thread"]
    ck.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        if nccl_log:
            line["nccl"] = nccl_log.relay(world)
            nccl_log.emit(json.dumps(line))
        else:
            print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    _args = parse_args()
    sys.exit(main_reference(_args) if _args.impl == "reference" else main_b200(_args))
