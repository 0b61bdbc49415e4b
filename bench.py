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
BATCH = 256                      # independent ciphertexts resident per GPU (BASELINE configs[4]: "1-256 ciphertexts")
PASSES = 3                       # a step walks the resident batch PASSES times -> 768 key switches per step per GPU,
KS_PER_STEP = BATCH * PASSES     #   so that the default 20 steps time ~3 s of GPU work
GROUP = 8                        # ciphertexts per launch sequence (digit buffers: 100 MB each)
E2E_HOST_CTS = 64                # distinct pinned host ciphertexts the host-buffer arm cycles through (1.6 GB each way)
N_EVK = 4                        # distinct evaluation keys cycled through (4 x 216 MB >> 126 MB L2)
LOGQ_INT = [60] + [52] * 23      # second line: the same shape with 52-bit scale primes (every limb on the integer path)
METRIC = "keyswitches_per_s_logN16"
UNIT = "keyswitch/s"


def workload_name():
    return ("hybrid key switch, logN=16, L=24 Q-limbs (60+23x45 bit), k=3 P-limbs (61 bit), beta=8, %d resident ciphertexts/GPU, "
            "%d key switches per step" % (BATCH, KS_PER_STEP))


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
                                       "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f, stderr=subprocess.DEVNULL)
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


def make_keyswitch_workload(lib, raw, logq, logp, rng, n_cts, n_evk, keep_host=False):
    """Context + evaluation keys + n_cts resident ciphertexts.  Host RNG is kept short (8 distinct random ciphertexts tiled
    over the batch, limbs uniform below 2^44 < every modulus): operand values do not change what the kernels do."""
    Q, P = raw.gen_moduli(logq, logp, LOGN + 1)
    ctx = raw.RawContext(LOGN, Q, P)
    N, L, K = ctx.N, len(Q), len(P)
    nmod, beta = L + K, ctx.num_digits(L - 1)
    LN = L * N
    evks, evk_host0 = [], None
    for e in range(n_evk):
        h = rng.integers(0, 1 << 44, size=beta * 2 * nmod * N, dtype=np.uint64)
        evks.append(ctx.evk_to_device(h))
        if e == 0 and keep_host:
            evk_host0 = h.reshape(beta, 2, nmod, N)
    distinct = [rng.integers(0, 1 << 44, size=2 * LN, dtype=np.uint64) for _ in range(8)]
    dev_in = lib.B200_DevAlloc(n_cts * 2 * LN)          # [n][2][L][N] = (c0, c1) per ciphertext
    dev_out = lib.B200_DevAlloc(n_cts * 2 * LN)
    for b in range(n_cts):
        lib.B200_Upload(dev_in + b * 2 * LN * 8, raw._p(distinct[b % 8]), 2 * LN)
    return dict(ctx=ctx, Q=Q, P=P, N=N, L=L, K=K, nmod=nmod, beta=beta, LN=LN, evks=evks, evk_host0=evk_host0,
                distinct=distinct, dev_in=dev_in, dev_out=dev_out, n=n_cts)


def resident_step(lib, w, s, passes=PASSES):
    """One step: `passes` walks over the resident batch in groups of GROUP; ciphertext b of pass p uses key (b // GROUP + p + s) % N_EVK
    for its whole group, so consecutive groups stream different keys (working set 4 x 216 MB of keys + 12.9 GB of ciphertexts)."""
    LN, level, ne = w["LN"], w["L"] - 1, len(w["evks"])
    for p_ in range(passes):
        for g, b0 in enumerate(range(0, w["n"], GROUP)):
            m = min(GROUP, w["n"] - b0)
            ek = (ctypes.c_void_p * m)(*([w["evks"][(g + p_ + s) % ne]] * m))
            lib.B200_KeySwitchBatchDev(w["ctx"].h, level, m, w["dev_in"] + b0 * 2 * LN * 8, 1, ek, w["dev_out"] + b0 * 2 * LN * 8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-resnet", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the integer-chain line, the stage split and the stand-alone NTT")
    ap.add_argument("--ks-only", action="store_true", help="key-switch timing only (profiling runs)")
    args = ap.parse_args()
    if args.ks_only:
        args.no_cpu_baseline = args.no_resnet = args.no_extras = True
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
    meta_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        meta_group = dist.new_group(backend="gloo")          # ciphertext metadata of the sharded inference travels host-to-host

    from orion_b200 import raw
    lib = raw.load()
    rng = np.random.default_rng(1234 + rank)

    # ---- everything the host has to prepare happens BEFORE the warm-up, so the GPU does not idle between warm-up and timing ----
    w = make_keyswitch_workload(lib, raw, LOGQ, LOGP, rng, BATCH, N_EVK, keep_host=(rank == 0))
    ctx, N, L, K, LN, level, nmod, beta = w["ctx"], w["N"], w["L"], w["K"], w["LN"], w["L"] - 1, w["nmod"], w["beta"]
    host_in = [np.ctypeslib.as_array(ctypes_cast(lib.B200_HostAllocPinned(2 * LN), 2 * LN)) for _ in range(E2E_HOST_CTS)]
    host_out = [np.ctypeslib.as_array(ctypes_cast(lib.B200_HostAllocPinned(2 * LN), 2 * LN)) for _ in range(E2E_HOST_CTS)]
    for b in range(E2E_HOST_CTS):
        host_in[b][:] = w["distinct"][b % 8]
    vp_arr = ctypes.c_void_p * KS_PER_STEP
    in_ptrs = vp_arr(*[host_in[i % E2E_HOST_CTS].ctypes.data for i in range(KS_PER_STEP)])
    out_ptrs = vp_arr(*[host_out[i % E2E_HOST_CTS].ctypes.data for i in range(KS_PER_STEP)])

    def step_resident(s):
        resident_step(lib, w, s)

    def step_e2e(s):
        # the reference-facing host-buffer call: (c0, c1) in pinned host memory -> key-switched ciphertext in host memory;
        # H2D, compute and D2H of consecutive ciphertexts overlap inside the library (three streams).  KS_PER_STEP key
        # switches per step, cycling over E2E_HOST_CTS pinned host ciphertexts (every one is uploaded and downloaded each time).
        evk_ptrs = vp_arr(*[w["evks"][(s + i // GROUP) % N_EVK] for i in range(KS_PER_STEP)])
        lib.B200_KeySwitchBatch(ctx.h, level, KS_PER_STEP, in_ptrs, 1, evk_ptrs, out_ptrs)

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

    # warm-up: W steps, and in any case >= 0.6 s of the same kernels right before the timed region (SM clocks ramped)
    t0 = time.perf_counter()
    nwarm = 0
    while nwarm < args.warmup or time.perf_counter() - t0 < 0.6:
        step_resident(nwarm); nwarm += 1
        if nwarm >= args.warmup:
            lib.B200_Sync()
    sampler = ClockSampler(local) if rank == 0 else None
    lib.B200_ProbeEnable(1)                             # CUDA events around the dominant kernel, on its own stream
    ms, launches = timed(step_resident, args.steps)
    lib.B200_ProbeEnable(0)
    probe_n = ctypes.c_long(0)
    probe_ms = lib.B200_ProbeReadMs(ctypes.byref(probe_n))
    clocks = sampler.stop() if sampler else None

    # ---- self-check: the last step's output of ciphertext 0 against the CPU oracle (the checker, not the product) ----
    parity = None
    if rank == 0 and not args.no_cpu_baseline:
        parity = check_against_oracle(lib, raw, w, args.steps - 1)

    step_e2e(0)                                          # warm-up of the host-buffer path (staging slots, copy streams)
    e2e_steps = max(2, args.steps // 4)
    ms_e2e, _ = timed(step_e2e, e2e_steps)

    # ---- per-stage split: one more step with an event pair around every stage of the key switch (not inside the timed region:
    #      an event between two launches keeps the second from starting early under PDL) ----
    stage_split = None
    if not args.no_extras or args.ks_only:
        lib.B200_ProbeEnable(2)
        lib.B200_Sync(); lib.B200_TimerStart()
        resident_step(lib, w, 0, passes=1)
        tot_ms = lib.B200_TimerStopMs()
        lib.B200_ProbeEnable(0)
        stage_split = {"unit": "us per key switch (one pass over the resident batch, groups of %d)" % GROUP,
                       "whole_key_switch": 1e3 * tot_ms / BATCH, "stages": {}}
        for sl in range(lib.B200_ProbeSlots()):
            n_ = ctypes.c_long(0)
            t_ = lib.B200_ProbeReadSlotMs(sl, ctypes.byref(n_))
            if n_.value:
                stage_split["stages"][lib.B200_ProbeSlotName(sl).decode()] = round(1e3 * t_ / BATCH, 3)

    # roofline of the dominant kernel: ntt2_ip_kernel<ArithF64> -- second NTT pass of the converted digits fused with the
    # evaluation-key inner product, over the nf limbs of at most 46 bits.  Algorithmic bytes per launch (GROUP
    # ciphertexts): per limb beta digit tiles in (own-digit tiles come from c1), 2*beta key tiles in, 2 accumulator tiles out.
    # The GROUP ciphertexts of a launch share ONE evaluation key (a batch of images through the same rotation), and the launch reads
    # it from DRAM once (CTA order: ciphertext fastest, the other GROUP-1 reads are L2 hits), so the key counts once per launch.
    # Round 1 counted it once per ciphertext (3*beta+2 tiles per limb and ciphertext); that figure is kept as frac_key_per_ciphertext.
    nf = sum(1 for q in list(w["Q"]) + list(w["P"]) if int(q).bit_length() <= 46)
    ip_bytes = 8 * N * nf * (GROUP * (beta + 2) + 2 * beta)
    ip_bytes_r1 = 8 * N * GROUP * nf * (3 * beta + 2)
    ip_ms = probe_ms / max(1, probe_n.value)
    ip_gbs = ip_bytes / (ip_ms * 1e-3) / 1e9 if probe_n.value else 0.0
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

    # secondary: the stand-alone batched forward NTT over beta*W - L limbs (the digit limbs of one group), timed alone with CUDA
    # events on the library stream; six 100 MB working sets cycled (> 4x L2), algorithmic bytes 16 N per limb (SURVEY 8(d):
    # one read + one write of the limb for the WHOLE transform -- the two-launch implementation moves twice that)
    ntt_line = None
    if not args.no_extras:
        nl = (beta * nmod - L)
        mods = [(i % nmod) for i in range(nl)]
        big = [lib.B200_DevAlloc(nl * N) for _ in range(6)]
        midx = raw._ints(mods)
        reps = 24
        lib.B200_Sync()
        for i in range(6):
            lib.B200_NTTDev(ctx.h, big[i], big[i], midx, nl, 0)
        lib.B200_Sync()
        lib.B200_TimerStart()
        for i in range(reps):
            lib.B200_NTTDev(ctx.h, big[i % 6], big[i % 6], midx, nl, 0)
        ntt_ms = lib.B200_TimerStopMs() / reps
        ntt_bytes = 16 * N * nl
        ntt_gbs = ntt_bytes / (ntt_ms * 1e-3) / 1e9
        for b in big:
            lib.B200_DevFree(b)
        ntt_line = {"bound": "hbm", "kernel": "ntt_pass x2 (stand-alone forward NTT of %d limbs; six 100 MB working sets cycled)" % nl,
                    "achieved": ntt_gbs, "peak": peak, "unit": "GB/s", "frac": ntt_gbs / peak,
                    "algorithmic_bytes_per_transform": ntt_bytes, "us_per_transform": ntt_ms * 1e3, "launches": 2,
                    "traffic": ntt_traffic, "traffic_note": "DRAM bytes of BOTH launches from the ncu --set full capture"}

    total_ks = KS_PER_STEP * args.steps * world
    value = total_ks / (ms * 1e-3)
    e2e_value = KS_PER_STEP * e2e_steps * world / (ms_e2e * 1e-3)
    ks_bytes = algorithmic_bytes_keyswitch(N, L, K)

    for d in [w["dev_in"], w["dev_out"]] + w["evks"]:
        lib.B200_DevFree(d)
    for hbuf in host_in + host_out:
        lib.B200_HostFreePinned(hbuf.ctypes.data)
    ctx.close()

    # ---- second key-switch line: the same shape on a chain of 52-bit scale primes (all 27 limbs on the integer path) ----
    int_chain = None
    if not args.no_extras:
        wi = make_keyswitch_workload(lib, raw, LOGQ_INT, LOGP, rng, 64, 2)
        for s_ in range(2):
            resident_step(lib, wi, s_, passes=1)
        lib.B200_Sync(); lib.B200_TimerStart()
        reps_i = 6
        for s_ in range(reps_i):
            resident_step(lib, wi, s_, passes=1)
        ms_i = lib.B200_TimerStopMs()
        int_chain = {"workload": "hybrid key switch, logN=16, L=24 Q-limbs (60+23x52 bit), k=3 (61 bit): every limb on the integer path",
                     "value": 64 * reps_i / (ms_i * 1e-3), "unit": UNIT + " per GPU", "key_switches_timed": 64 * reps_i}
        for d in [wi["dev_in"], wi["dev_out"]] + wi["evks"]:
            lib.B200_DevFree(d)
        wi["ctx"].close()

    # ---- second half of BASELINE's metric: ResNet-20 encrypted inference, seconds per image ----
    # The reference's exact backend-call sequence for ResNet-20 (examples/run_resnet.py flow: default-init weights, fit on the
    # input), recorded from the unmodified reference WITH its payloads and replayed through the drop-in boundary; every rank
    # runs its own image (image-sharded, no collective).  configs/resnet.yml as shipped (logN=13) and the logN=16 variant.
    resnet = resnet16 = lola = None
    if not args.no_resnet:
        resnet = run_resnet(lib, rank, world, dist if world > 1 else None, torch, peak)
        if world > 1 and resnet is not None:
            resnet["strong_scaling"] = run_resnet_sharded(lib, rank, world, dist, torch, meta_group)
        if world == 1 and resnet is not None and not args.no_extras:
            resnet["lockstep"] = run_resnet_lockstep(lib, rank)
        lola = run_lola(lib, rank, world, dist if world > 1 else None, torch)
        if os.environ.get("ORION_B200_BENCH_RESNET16", "1") != "0":                   # last: it needs the whole GPU's memory
            resnet16 = run_resnet16(lib, rank, world, dist if world > 1 else None, torch, peak)

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
            "config": {"workload": workload_name(),
                       "l2": "inputs larger than L2: %d evaluation keys x 216 MB cycled per group, 12.9 GB of resident ciphertexts walked" % N_EVK,
                       "timed_region_s": ms * 1e-3, "warmup_steps_run": nwarm,
                       "keyswitch_algorithmic_bytes": ks_bytes,
                       "keyswitch_hbm_floor_frac": (ks_bytes * total_ks / (ms * 1e-3) / 1e9 / world) / peak},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": KS_PER_STEP * 2 * LN * 8,
                    "d2h_bytes_per_step": KS_PER_STEP * 2 * LN * 8, "steps": e2e_steps, "cpu_affinity": affinity,
                    "note": "B200_KeySwitchBatch on pinned HOST buffers: every key switch uploads its ciphertext (25 MB) and downloads the result (25 MB); "
                            "PCIe-bound by construction.  The API's own use (one ciphertext in, hundreds of key switches, logits out) is the resnet20 / lola2 "
                            "objects below: their timed regions include encode+encrypt (H2D) and decrypt+decode (D2H) of every image."},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "ntt2_ip_kernel<ArithF64> (second NTT pass of the digits fused with the evaluation-key inner product; %d limbs x %d ciphertexts per launch, one shared key)" % (nf, GROUP),
                         "achieved": ip_gbs, "peak": peak, "unit": "GB/s", "frac": ip_gbs / peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                         "frac_key_per_ciphertext": (ip_bytes_r1 / (ip_ms * 1e-3) / 1e9 / peak) if probe_n.value else None,
                         "fp64_pipe_floor_us": 1e6 * (GROUP * nf * (N // 1024) * ((beta - 1) * 4096 * 8 + beta * 2048 * 7)) / (148 * 64 * 1.965e9),
                         "bound_note": "key counted once per launch (shared by the group).  The kernel is not HBM-bound any more: its arithmetic -- "
                                       "(beta-1) x 8 butterfly stages (8 DFMA-class ops per butterfly) + 2 beta multiply-accumulates (7 ops) per coefficient, "
                                       "all on the FP64 pipe at 64 lanes/clk/SM -- sets fp64_pipe_floor_us at the maximum SM clock",
                         "algorithmic_bytes_per_launch": ip_bytes, "launches": int(probe_n.value), "avg_launch_us": ip_ms * 1e3,
                         "share_of_step": probe_ms / ms if ms else None,
                         "traffic": traffic, "traffic_note": "DRAM bytes per launch from the committed ncu --set full capture (profiles/ks_fused_traffic.json)"},
            "clocks": clocks,
        }
        if parity is not None:
            line["parity_checked"] = bool(parity["equal"])
            line["parity"] = parity
        if stage_split:
            line["keyswitch_stage_split"] = stage_split
        if ntt_line:
            line["ntt_roofline"] = ntt_line
        if int_chain:
            line["keyswitch_integer_chain"] = int_chain
        if cpu:
            line["cpu_baseline"] = cpu
        if resnet:
            line["resnet20"] = resnet
        if resnet16:
            line["resnet20_logn16"] = resnet16
        if lola:
            line["lola2"] = lola
        emit_json(line)
    if world > 1:
        dist.destroy_process_group()


def check_against_oracle(lib, raw, w, last_step):
    """bench.py checks its own output: ciphertext 0 of the last timed step (key evks[(0 + PASSES-1 + last_step) % N_EVK] in its last
    pass) is downloaded and compared word for word with oracle.Ctx.keyswitch on the same input; when that key is not key 0 (the one
    whose host copy was kept), ciphertext 0 is switched once more with key 0 through the same entry point.  oracle/ is used as
    the checker only."""
    import oracle
    L, N, LN = w["L"], w["N"], w["LN"]
    level = L - 1
    oc = oracle.Ctx(LOGN, list(w["Q"]) + list(w["P"]), len(w["Q"]), len(w["P"]))
    oc.L.orc_set_threads(os.cpu_count() or 1)
    ct = w["distinct"][0].reshape(2, L, N)
    used = (0 + PASSES - 1 + last_step) % len(w["evks"])
    how = "ciphertext 0 of the last timed step"
    if used != 0:
        m = GROUP
        ek = (ctypes.c_void_p * m)(*([w["evks"][0]] * m))
        lib.B200_KeySwitchBatchDev(w["ctx"].h, level, m, w["dev_in"], 1, ek, w["dev_out"])
        how = "group 0 re-run with key 0 through the timed entry point after the timed region"
    out = np.empty((2, L, N), dtype=np.uint64)
    lib.B200_Download(raw._p(out), w["dev_out"], 2 * LN)
    ref = oc.keyswitch(level, ct[1], w["evk_host0"])
    ref[0] = oc.elementwise("add", ref[0], ct[0])
    return {"equal": bool(np.array_equal(out, ref)), "what": how + ": B200_KeySwitchBatchDev (n=%d, +c0) vs oracle.Ctx.keyswitch, %d words" % (GROUP, out.size)}


def run_trace_timed(name, lib, rank, world, dist, torch, reps=3, peak=None, warm=4):
    """Replay the reference's recorded call trace `name`, one image per rank; returns timing fields or None.
    The timed region of an inference includes encode + encrypt of the image (H2D) and decrypt + decode of the logits (D2H)."""
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
    out = None
    for _ in range(warm):
        out = r.infer(); lib.B200_Sync()                         # warm-up (plans, key caches, pool growth, side streams, graph captures)
    l0 = lib.B200_LaunchCount()
    a0 = lib.B200_AlgorithmicBytes()
    times = []
    for _ in range(reps):
        lib.B200_Sync()
        if dist is not None:
            dist.barrier()
        lib.B200_TimerStart()
        out = r.infer()
        times.append(lib.B200_TimerStopMs() / 1e3)
    launches = (lib.B200_LaunchCount() - l0) // reps
    alg = (lib.B200_AlgorithmicBytes() - a0) / reps
    sec = min(times)
    if dist is not None:
        tt = torch.tensor([sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt.item())
    B.DeleteScheme()
    res = {"s_per_image": sec, "images_per_s": world / sec, "n_gpus": world, "setup_s": setup_s,
           "launches_per_image": int(launches), "counts": {k: r.counts.get(k, 0) // (reps + warm) for k in
                                                          ("EvaluateLinearTransform", "EvaluatePolynomial", "Bootstrap", "MulRelinCiphertextNew")}}
    clear = np.array(t.meta.get("clear_output", []), dtype=np.float64)
    if out is not None and clear.size and t.meta.get("payloads") == "full":
        fhe = np.asarray(out, dtype=np.float64)[: clear.size]
        res["numerics"] = {"mae_vs_cleartext_network": float(np.mean(np.abs(fhe - clear))), "max_abs_err": float(np.max(np.abs(fhe - clear))),
                           "argmax_equal": bool(int(np.argmax(fhe)) == int(np.argmax(clear)))}
    if peak:
        gbs = alg / sec / 1e9
        res["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                           "algorithmic_bytes_per_image": alg,
                           "note": "sum of SURVEY 8(d) algorithmic bytes of every evaluator op of one inference (counted by the library as the ops "
                                   "are issued: key switches, hoisted BSGS transforms, rescales, element-wise ops) / time / measured HBM peak"}
    return res


def run_resnet(lib, rank, world, dist, torch, peak=None, reps=3):
    res = run_trace_timed("resnet20", lib, rank, world, dist, torch, reps, peak)
    if res is None:
        return None
    res.update({"config": "configs/resnet.yml as shipped (logN=13, LogQ 60+32x30, LogP 60x2, boot LogP 61x8), one image per GPU",
                "data": "examples/run_resnet.py flow on a synthetic CIFAR-shaped image: default-init weights, fit on the input; real payloads "
                        "(diagonals, biases, sign polynomials) recorded from the unmodified reference",
                "e2e": "includes encode+encrypt (H2D) of the image and decrypt+decode (D2H) of the logits"})
    return res


def run_resnet_lockstep(lib, rank, k=4, reps=2):
    """K images per GPU in lockstep over ONE scheme (orion_b200.replay.BatchReplayer): call i of image 0, call i of image 1, ... --
    the images' ciphertexts are independent, so the library's dataflow lanes overlap their small-ring launches.  images/s per GPU."""
    from orion_b200.bindings import B200Library
    from orion_b200.replay import BatchReplayer, Trace
    gold = os.path.join(ROOT, "tests", "golden")
    t = Trace(os.path.join(gold, "trace_resnet20.json"), os.path.join(gold, "trace_resnet20.npz"))
    B = B200Library()
    B.B200_SetSeed(42 + rank)
    r = BatchReplayer(B, t, k)
    r.setup(); lib.B200_Sync()
    outs = r.infer(); lib.B200_Sync()
    best = float("inf")
    for _ in range(reps):
        lib.B200_Sync(); lib.B200_TimerStart()
        outs = r.infer()
        best = min(best, lib.B200_TimerStopMs() / 1e3)
    clear = np.array(t.meta["clear_output"], dtype=np.float64)
    err = max(float(np.max(np.abs(np.asarray(o, dtype=np.float64)[: clear.size] - clear))) for o in outs)
    B.DeleteScheme()
    return {"images_in_flight": k, "s_per_batch": best, "images_per_s_per_gpu": k / best, "max_abs_err": err}


def run_resnet_sharded(lib, rank, world, dist, torch, meta_group, name="resnet20", reps=3):
    """ONE image across all ranks (strong scaling): block rows of every linear layer round-robin over the ranks, activations and
    bootstraps where their ciphertext lives, ciphertexts broadcast over NCCL (device limbs) on first remote use."""
    from orion_b200.bindings import B200Library
    from orion_b200.parallel import DeviceExchange, ShardedReplayer
    from orion_b200.replay import Trace
    gold = os.path.join(ROOT, "tests", "golden")
    t = Trace(os.path.join(gold, "trace_%s.json" % name), os.path.join(gold, "trace_%s.npz" % name))
    B = B200Library()
    B.B200_SetSeed(42)                                          # the same keys on every rank (ranks exchange ciphertexts)
    x = DeviceExchange(B, 1 << t.meta["config"]["ckks_params"]["LogN"], rank, world, meta_group=meta_group)
    r = ShardedReplayer(B, t, x, seed=1234)
    r.setup(); lib.B200_Sync()
    out = None
    for _ in range(2):
        out = r.infer(); lib.B200_Sync()
    times = []
    b0, n0, ops0 = x.bytes_moved, x.transfers, r.ops_run
    for _ in range(reps):
        lib.B200_Sync(); torch.cuda.synchronize(); dist.barrier(); lib.B200_Sync()
        lib.B200_TimerStart()
        out = r.infer()
        ms = lib.B200_TimerStopMs()
        torch.cuda.synchronize()
        tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)               # the image is done when the slowest rank is
        times.append(float(tt.item()) / 1e3)
    ops = torch.tensor([float(r.ops_run - ops0) / reps], device="cuda", dtype=torch.float64)
    allops = [torch.zeros_like(ops) for _ in range(world)]
    dist.all_gather(allops, ops)
    res = {"s_per_image": min(times), "n_gpus": world, "scaling": "strong",
           "ciphertext_broadcasts_per_image": (x.transfers - n0) // reps, "broadcast_bytes_per_image": (x.bytes_moved - b0) // reps,
           "evaluator_calls_per_rank": [int(o.item()) for o in allops],
           "note": "limited by the layers whose tensors are a single ciphertext (64-channel stage: 1 ciphertext of 4096 slots at logN=13) -- "
                   "only one rank works there -- not by NVLink: a broadcast moves 2-4 MB"}
    if rank == 0 and out is not None:
        clear = np.array(t.meta["clear_output"], dtype=np.float64)
        fhe = np.asarray(out, dtype=np.float64)[: clear.size]
        res["max_abs_err_vs_cleartext"] = float(np.max(np.abs(fhe - clear)))
    B.DeleteScheme()
    return res


def run_resnet16(lib, rank, world, dist, torch, peak=None):
    """BASELINE configs[3]: the same network with LogN forced to 16 (32768 slots; sparse-slot bootstraps at 16384/8192/4096).  Its
    keys, encoded diagonals and bootstrap factors take 160 GB, so every rank runs it in a child process of its own on its GPU
    (tools/resnet16.py) after handing this process's pooled memory back to the driver; a failure there is reported, never fatal."""
    import subprocess
    lib.B200_TrimPool()
    env = dict(os.environ, ORION_B200_DEVICE=os.environ.get("ORION_B200_DEVICE", "0"))
    res = None
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "resnet16.py"), "2", str(42 + rank)], capture_output=True, text=True,
                           env=env, timeout=900)
        if r.returncode == 0:
            res = json.loads(r.stdout.strip().splitlines()[-1])
        else:
            res = {"status": "failed: rc %d: %s" % (r.returncode, r.stderr.strip()[-200:])}
    except Exception as e:
        res = {"status": "failed: %s" % (str(e)[:200],)}
    sec = res.get("s_per_image", float("inf"))
    if dist is not None:
        tt = torch.tensor([sec], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt.item())
    if "s_per_image" not in res or sec == float("inf"):
        return {"status": res.get("status", "failed on another rank")}
    res["s_per_image"] = sec
    res.update({"images_per_s": world / sec, "n_gpus": world})
    if peak:
        gbs = res["algorithmic_bytes_per_image"] / sec / 1e9
        res["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                           "algorithmic_bytes_per_image": res.pop("algorithmic_bytes_per_image"),
                           "note": "sum of SURVEY 8(d) algorithmic bytes of every evaluator op of one inference / time / measured HBM peak"}
    res.update({"config": "configs/resnet.yml with LogN=16 (LogQ 60+32x30, LogP 60x2, boot LogP 61x8; BASELINE configs[3]), one image per GPU",
                "data": "same flow and payload recording as resnet20 (default-init weights, fit on the input; recorded from the unmodified reference)",
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
