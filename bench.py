#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 CKKS backend (contract: see the task statement / DESIGN.md).

Metric (BASELINE.json): key switches per second at logN=16.  One step = one batch of independent hybrid key
switches (the unit of work inside rotation, relinearisation, every BSGS step and every bootstrap DFT layer) at
logN=16, L=24 Q-limbs, k=3 special primes (beta=8 digits) -- the upper end of BASELINE.json's primitive sweep.
Ciphertexts are sharded across ranks (no data-path collective): weak scaling.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # before torch creates the CUDA context (see orion_b200/__init__.py)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # NCCL's version banner must not precede the JSON line on stdout

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOGN = 16
LOGQ = [60] + [45] * 23          # 24 Q-limbs
LOGP = [61, 61, 61]              # 3 special primes -> beta = 8
BATCH = 32                       # key switches per step per GPU (independent ciphertexts)
N_EVK = 4                        # distinct evaluation keys cycled through (4 x 216 MB >> 126 MB L2)
METRIC = "keyswitches_per_s_logN16"
UNIT = "keyswitch/s"


def workload_name():
    return "hybrid key switch, logN=16, L=24 Q-limbs (60+23x45 bit), k=3 P-limbs (61 bit), beta=8, batch %d/GPU" % BATCH


def algorithmic_bytes_keyswitch(N, L, K):
    beta = (L + K - 1) // K
    return 8 * N * (4 * L + 2 * beta * (L + K))          # ct in + evk + ct out


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        for r in rows:
            try:
                sm.append(float(r[0])); out["sm_max_mhz"] = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        if sm:
            out["sm_mhz"] = statistics.median(sm)
        out["reasons"] = sorted(reasons)
        return out


def cpu_keyswitch_rate(threads, budget_s=12.0):
    """Oracle (CPU restatement, OpenMP) timed on the same workload: a bounded sample of whole key switches."""
    import oracle
    Q, P = oracle.gen_moduli(LOGQ, LOGP, LOGN + 1)
    oc = oracle.Ctx(LOGN, Q + P, len(Q), len(P))
    oc.L.orc_set_threads(threads)
    rng = np.random.default_rng(1234)
    level = len(Q) - 1
    nmod = len(Q) + len(P)
    beta = oc.num_digits(level)
    evk = rng.integers(0, 1 << 44, size=(beta, 2, nmod, oc.N), dtype=np.uint64)
    c1 = oc.uniform_poly(rng, list(range(level + 1)))
    oc.keyswitch(level, c1, evk)                      # warm-up (page faults, table caches)
    n, t0 = 0, time.perf_counter()
    while True:
        oc.keyswitch(level, c1, evk)
        n += 1
        dt = time.perf_counter() - t0
        if dt > budget_s or n >= 64:
            break
    return n / dt, n, dt


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    steps_total = max(1, args.steps)
    rate, n, dt = cpu_keyswitch_rate(threads, budget_s=min(60.0, 4.0 * steps_total))
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * dt / n, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_name(), "note": "CPU oracle (C restatement of the Lattigo algorithm, OpenMP); "
                   "Lattigo itself cannot be built here (no Go toolchain, module not vendored)"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%d whole key switches at logN=16, L=24, k=3 in %.1f s" % (n, dt)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit_json(line)


_ORIG_AFFINITY = None
_JSON_FD = None


def emit_json(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def bind_to_gpu_numa_node(local):
    """Pin this rank to the CPUs next to its GPU (NVML's ideal affinity) BEFORE any pinned host buffer is allocated, so the
    host-buffer (e2e) path copies over the local PCIe root and local DRAM.  ORION_B200_AFFINITY=0 disables it."""
    if os.environ.get("ORION_B200_AFFINITY", "1") == "0":
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else local
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        global _ORIG_AFFINITY
        _ORIG_AFFINITY = os.sched_getaffinity(0)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        return "%d cpus (%d-%d)" % (len(cpus), cpus[0], cpus[-1])
    except Exception as e:                                        # not fatal: the benchmark runs unpinned
        return "unavailable: %s" % (str(e)[:60],)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-resnet", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON result: everything else a library prints there (NCCL's version banner under
    # torchrun) is sent to stderr by pointing fd 1 at fd 2 for the duration of the run
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    os.environ.setdefault("ORION_B200_DEVICE", str(local))
    affinity = bind_to_gpu_numa_node(local)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from orion_b200 import raw
    lib = raw.load()
    Q, P = raw.gen_moduli(LOGQ, LOGP, LOGN + 1)
    ctx = raw.RawContext(LOGN, Q, P)
    N, L, K = ctx.N, len(Q), len(P)
    level, nmod, beta = L - 1, L + K, ctx.num_digits(L - 1)
    LN = L * N
    rng = np.random.default_rng(1234 + rank)

    # synthetic operands: limbs uniform below 2^44 (< every modulus), seeded -- values do not affect timing
    evks = []
    for e in range(N_EVK):
        h = rng.integers(0, 1 << 44, size=beta * 2 * nmod * N, dtype=np.uint64)
        evks.append(ctx.evk_to_device(h))
        del h
    host_in = [np.ctypeslib.as_array(ctypes_cast(lib.B200_HostAllocPinned(2 * LN), 2 * LN)) for _ in range(BATCH)]
    host_out = [np.ctypeslib.as_array(ctypes_cast(lib.B200_HostAllocPinned(2 * LN), 2 * LN)) for _ in range(BATCH)]
    for b in range(BATCH):
        host_in[b][:] = rng.integers(0, 1 << 44, size=2 * LN, dtype=np.uint64)
    dev_in = lib.B200_DevAlloc(BATCH * 2 * LN)          # [BATCH][2][L][N] = (c0, c1) per ciphertext
    dev_out = lib.B200_DevAlloc(BATCH * 2 * LN)
    for b in range(BATCH):
        lib.B200_Upload(dev_in + b * 2 * LN * 8, raw._p(host_in[b]), 2 * LN)

    GROUP = 8                                           # ciphertexts per launch sequence (digit buffers: 100 MB each)

    def step_resident(s):
        for b0 in range(0, BATCH, GROUP):
            ek = (ctypes.c_void_p * GROUP)(*[evks[(s * BATCH + b0 + b) % N_EVK] for b in range(GROUP)])
            lib.B200_KeySwitchBatchDev(ctx.h, level, GROUP, dev_in + b0 * 2 * LN * 8, 1, ek, dev_out + b0 * 2 * LN * 8)

    vp_arr = ctypes.c_void_p * BATCH
    in_ptrs = vp_arr(*[h.ctypes.data for h in host_in])
    out_ptrs = vp_arr(*[h.ctypes.data for h in host_out])

    def step_e2e(s):
        # the reference-facing host-buffer call: (c0, c1) in pinned host memory -> key-switched ciphertext in host memory;
        # H2D, compute and D2H of consecutive ciphertexts overlap inside the library (three streams)
        evk_ptrs = vp_arr(*[evks[(s * BATCH + b) % N_EVK] for b in range(BATCH)])
        lib.B200_KeySwitchBatch(ctx.h, level, BATCH, in_ptrs, 1, evk_ptrs, out_ptrs)

    def barrier():
        lib.B200_Sync(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        lib.B200_Sync(); torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        l0 = lib.B200_LaunchCount()
        lib.B200_TimerStart()
        for s in range(steps):
            fn(s)
        ms = lib.B200_TimerStopMs()
        launches = lib.B200_LaunchCount() - l0
        barrier()
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    for s in range(args.warmup):
        step_resident(s)
    sampler = ClockSampler(local) if rank == 0 else None
    lib.B200_ProbeEnable(1)                             # CUDA events around the dominant kernel, on its own stream
    ms, launches = timed(step_resident, args.steps)
    lib.B200_ProbeEnable(0)
    probe_n = ctypes.c_long(0)
    probe_ms = lib.B200_ProbeReadMs(ctypes.byref(probe_n))
    clocks = sampler.stop() if sampler else None
    for s in range(max(1, args.warmup // 2)):
        step_e2e(s)
    e2e_steps = max(2, args.steps // 4)
    ms_e2e, _ = timed(step_e2e, e2e_steps)

    # roofline of the dominant kernel: ntt2_ip_kernel<ArithF64> -- second NTT pass of the converted digits fused with the
    # evaluation-key inner product, over the nf limbs of at most 46 bits.  Algorithmic bytes per launch (GROUP
    # ciphertexts): per limb beta digit tiles in (own-digit tiles come from c1), 2*beta key tiles in, 2 accumulator tiles out.
    nf = sum(1 for q in list(Q) + list(P) if int(q).bit_length() <= 46)
    ip_bytes = 8 * N * GROUP * nf * (3 * beta + 2)
    ip_ms = probe_ms / max(1, probe_n.value)
    ip_gbs = ip_bytes / (ip_ms * 1e-3) / 1e9 if probe_n.value else 0.0
    # secondary: the stand-alone batched NTT over the converted digit limbs (beta*W - L limbs), timed alone with CUDA
    # events on the library stream, inputs 99 MB per launch pair cycled over > L2
    nl = beta * nmod - L
    mods = [(i % nmod) for i in range(nl)]
    big = [lib.B200_DevAlloc(nl * N) for _ in range(3)]            # 3 x 100 MB working sets
    midx = raw._ints(mods)
    reps = 12
    lib.B200_Sync()
    for i in range(3):
        lib.B200_NTTDev(ctx.h, big[i], big[i], midx, nl, 0)
    lib.B200_Sync()
    lib.B200_TimerStart()
    for i in range(reps):
        lib.B200_NTTDev(ctx.h, big[i % 3], big[i % 3], midx, nl, 0)
    ntt_ms = lib.B200_TimerStopMs() / reps
    ntt_bytes = 2 * 16 * N * nl                                       # two passes, each reads + writes every limb
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = ntt_traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ks_fused_traffic.json")))["dram_bytes_per_launch"]
        ntt_traffic = json.load(open(os.path.join(ROOT, "profiles", "ntt_traffic.json")))["dram_bytes_per_launch"]
    except Exception:
        pass
    ntt_gbs = ntt_bytes / (ntt_ms * 1e-3) / 1e9

    total_ks = BATCH * args.steps * world
    value = total_ks / (ms * 1e-3)
    e2e_value = BATCH * e2e_steps * world / (ms_e2e * 1e-3)
    ks_bytes = algorithmic_bytes_keyswitch(N, L, K)

    # ---- second half of BASELINE's metric: ResNet-20 encrypted inference, seconds per image ----
    # The reference's exact backend-call sequence for ResNet-20 (configs/resnet.yml as shipped: logN=13, 33+2 limbs,
    # boot LogP 8x61; 147 linear transforms, 138 polynomial evaluations, 28 bootstraps), recorded from the unmodified
    # reference and replayed through the drop-in boundary; every rank runs its own image (image-sharded, no collective).
    resnet = None
    lola = None
    if not args.no_resnet:
        for d in [dev_in, dev_out] + big + evks:
            lib.B200_DevFree(d)
        ctx.close()
        resnet = run_resnet(lib, rank, world, dist if world > 1 else None, torch)
        lola = run_lola(lib, rank, world, dist if world > 1 else None, torch)

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            if _ORIG_AFFINITY:
                os.sched_setaffinity(0, _ORIG_AFFINITY)      # the CPU baseline gets every host core back (OpenMP starts its threads now)
            threads = os.cpu_count() or 1
            rate, n, dt = cpu_keyswitch_rate(threads)
            cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "%d whole key switches at logN=16, L=24, k=3 in %.1f s (oracle, OpenMP)" % (n, dt)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": workload_name(), "l2": "inputs larger than L2: %d evaluation keys x 216 MB cycled" % N_EVK,
                       "keyswitch_algorithmic_bytes": ks_bytes,
                       "keyswitch_hbm_floor_frac": (ks_bytes * total_ks / (ms * 1e-3) / 1e9 / world) / peak},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": BATCH * 2 * LN * 8,
                    "d2h_bytes_per_step": BATCH * 2 * LN * 8, "cpu_affinity": affinity},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "ntt2_ip_kernel<ArithF64> (second NTT pass of the digits fused with the evaluation-key inner product; %d limbs x %d ciphertexts per launch)" % (nf, GROUP),
                         "achieved": ip_gbs, "peak": peak, "unit": "GB/s", "frac": ip_gbs / peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                         "algorithmic_bytes_per_launch": ip_bytes, "launches": int(probe_n.value), "avg_launch_us": ip_ms * 1e3,
                         "share_of_step": probe_ms / ms if ms else None,
                         "traffic": traffic, "traffic_note": "DRAM bytes per launch from the committed ncu --set full capture (profiles/ks_fused_traffic.json)"},
            "ntt_roofline": {"bound": "hbm", "kernel": "ntt_pass (stand-alone forward NTT of %d limbs, 2 launches)" % nl,
                             "achieved": ntt_gbs, "peak": peak, "unit": "GB/s", "frac": ntt_gbs / peak,
                             "algorithmic_bytes_per_launch": ntt_bytes // 2, "launches": 2, "traffic": ntt_traffic},
            "clocks": clocks,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        if resnet:
            line["resnet20"] = resnet
        if lola:
            line["lola2"] = lola
        emit_json(line)
    if world > 1:
        dist.destroy_process_group()


def run_trace_timed(name, lib, rank, world, dist, torch, reps=3):
    """Replay the reference's recorded call trace `name`, one image per rank; returns timing fields or None."""
    from orion_b200.bindings import B200Library
    from orion_b200.replay import Replayer, Trace
    gold = os.path.join(ROOT, "tests", "golden")
    jp, npz = os.path.join(gold, "trace_%s.json" % name), os.path.join(gold, "trace_%s.npz" % name)
    if not os.path.exists(jp):
        return None
    B = B200Library()
    B.B200_SetSeed(42 + rank)
    t = Trace(jp, npz)
    r = Replayer(B, t, seed=1234 + rank)
    t0 = time.perf_counter(); r.setup(); lib.B200_Sync(); setup_s = time.perf_counter() - t0
    for _ in range(2):
        r.infer(); lib.B200_Sync()                               # warm-up (plans, key caches, pool growth, side streams)
    l0 = lib.B200_LaunchCount()
    times = []
    for _ in range(reps):
        lib.B200_Sync()
        if dist is not None:
            dist.barrier()
        lib.B200_TimerStart()
        r.infer()
        times.append(lib.B200_TimerStopMs() / 1e3)
    launches = (lib.B200_LaunchCount() - l0) // reps
    sec = min(times)
    if dist is not None:
        tt = torch.tensor([sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt.item())
    B.DeleteScheme()
    return {"s_per_image": sec, "images_per_s": world / sec, "n_gpus": world, "setup_s": setup_s,
            "launches_per_image": int(launches), "counts": {k: r.counts.get(k, 0) // (reps + 2) for k in
                                                           ("EvaluateLinearTransform", "EvaluatePolynomial", "Bootstrap", "MulRelinCiphertextNew")}}


def run_resnet(lib, rank, world, dist, torch, reps=3):
    res = run_trace_timed("resnet20", lib, rank, world, dist, torch, reps)
    if res is None:
        return None
    res.update({"bootstraps": 28, "linear_transforms": 147, "poly_evals": 138,
            "config": "configs/resnet.yml as shipped (logN=13, LogQ 60+32x30, LogP 60x2, boot LogP 61x8), one image per GPU",
            "data": "synthetic payloads of the recorded shapes (call trace of the unmodified reference)",
            "e2e": "includes encode+encrypt (H2D) of the image and decrypt+decode (D2H) of the logits"})
    return res


def run_lola(lib, rank, world, dist, torch):
    """BASELINE configs[2]: LoLA with configs/lola2.yml (logN=14, 7+2 limbs), independent images sharded over the ranks."""
    res = run_trace_timed("lola", lib, rank, world, dist, torch, reps=5)
    if res is None:
        return None
    res.update({"config": "configs/lola2.yml (logN=14, LogQ 60+6x30, LogP 60x2), one image per GPU per pass",
                "data": "synthetic MNIST-shaped image, full recorded payloads (call trace of the unmodified reference)"})
    return res


def ctypes_cast(ptr, words):
    import ctypes
    return ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint64 * words)).contents


if __name__ == "__main__":
    main()
