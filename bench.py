#!/usr/bin/env python
"""bench.py -- throughput of the belief-propagation decoder hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--trials T]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic input: decode T independent trials (priors ->
hard decisions) under the iteration cap.  Default workload = BASELINE.json configs[1]: Kyber-512 q-ary BP
(eta1 = 3, 7-value alphabet), 10^4 batched trials on one B200, sum_weight 6, check_blocks 1, 10 iterations.

One JSON line on stdout (rank 0):
  value      whole-job trials/s over all N GPUs with the priors already resident in HBM (device-timed, CUDA events,
             max over ranks); trials are independent, so each rank decodes its own T trials ("weak" scaling) and the
             only exchange is an all-reduce of the success counters.
  e2e        the same metric through the reference-facing API with HOST buffers: every step copies that step's
             priors from pinned host memory to the device and the hard decisions back.
  roofline   the dominant kernel (check-node pass) against the measured HBM copy peak (MEASURED_PEAKS.json).
  cpu_baseline  the CPU oracle (restatement of the reference's brute-force min-sum, one decode per thread as
             simulate/decode.py:247-262 does) on a bounded sample of the same workload.
--impl reference times that CPU path alone (the reference's Rust crate cannot be built here: no cargo/rustc).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (kind, kwargs, default trials, iterations)
    "kyber512": ("qary", dict(variant="kyber512", sum_weight=6, check_blocks=1, accuracy=0.98, nbits=4, nbits_sum=5), 10_000, 10),
    "kyber768": ("qary", dict(variant="kyber768", sum_weight=6, check_blocks=2, accuracy=0.95, nbits=3, nbits_sum=4), 10_000, 10),
    "kyber768cb1": ("qary", dict(variant="kyber768", sum_weight=6, check_blocks=1, accuracy=0.95, nbits=3, nbits_sum=4), 10_000, 10),
    "kyber1024": ("qary", dict(variant="kyber1024", sum_weight=8, check_blocks=1, accuracy=0.98, nbits=5, nbits_sum=6), 12_500, 10),
    # binary batches: large enough that the ~8 % of trials that never converge (and run all 100 iterations on a few
    # tiles after compaction) do not dominate: 4096 -> 250k, 16384 -> 342k, 32768 -> 357k trials/s on hqc128 (measured)
    # the reference's own criterion case (simulate_rs/benches/decoder.rs "medium decoder"): generic Decoder<450,150,3,7,15,7>
    "decoder450": ("qary", dict(generic=True), 10_000, 10),
    "hqc192": ("bp", dict(variant="hqc192", W=50, R=5000, rho=0.9942), 8192, 100),
    "hqc128": ("bp", dict(variant="hqc128", W=50, R=3000, rho=0.9942), 16384, 100),
    "hqc256": ("bp", dict(variant="hqc256", W=50, R=8000, rho=0.9942), 8192, 100),
}


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (the B200_PROFILING.md clocks line), read in-process
    through NVML (nvidia_ml_py) from a thread every 20 ms.  A polling `nvidia-smi -lms` subprocess was measured to
    stall kernel launches for ~1 ms per 6 ms step; plain NVML getters do not.  Falls back to one nvidia-smi query
    after the timed region when NVML is not importable."""

    def __init__(self, index=0):
        self.index, self.samples, self.thread, self.stop_flag, self.nv, self.handle = index, [], None, False, None, None

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv, self.handle = nv, nv.nvmlDeviceGetHandleByIndex(self.index)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.nv = None

    def _pump(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM)
                rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.samples.append((time.perf_counter(), sm, mx, rs))
            except Exception:
                pass
            time.sleep(0.02)

    def stop(self, t0, t1):
        self.stop_flag = True
        if self.nv is None:
            return self._smi_once()
        self.thread.join(timeout=1.0)
        nv = self.nv
        rows = [x for x in self.samples if t0 <= x[0] <= t1] or self.samples[-3:]
        bits = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        reasons = sorted({name for (_, _, _, rs) in rows for name, bit in bits.items() if rs & bit})
        sm = [x[1] for x in rows]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max((x[2] for x in rows), default=None),
                "reasons": reasons, "samples": len(rows), "source": "nvml, 20 ms period, inside the timed region"}

    def _smi_once(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=20).stdout.strip().splitlines()[0]
            f = [x.strip() for x in out.split(",")]
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            return {"sm_mhz": float(f[0]), "sm_max_mhz": float(f[1]),
                    "reasons": [n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")], "samples": 1,
                    "source": "nvidia-smi, one query right after the timed region"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# =====================================================================================================================
# reference arm / cpu_baseline: the CPU oracle on the host cores
# =====================================================================================================================
def host_cores():
    """Host threads the CPU arm uses: the cores this process may run on.  NOT omp_get_max_threads(): torchrun exports
    OMP_NUM_THREADS=1 to its workers, which throttled the round-1 reference arm to one thread at N > 1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_path(kind, wl, iters, trials, threads, warm_only=False, algo=None):
    """Times the oracle on `trials` trials with `threads` host threads.  Returns (trials/s, seconds, successes).
    algo: check-node algorithm of the q-ary oracle (default: the reference's enumeration)."""
    from oracle import oracle as O

    if kind == "qary":
        b = wl.generate(trials, seed=12345)
        orc = O.QaryOracle(wl.H, 1 if warm_only else iters, B=wl.B, BSUM=wl.BSUM, DC=wl.sum_weight + 1,
                           algo=O.ALGO_ENUMERATE if algo is None else algo)
        t0 = time.perf_counter()
        hard, _, _, st = orc.decode_batch(b.priors, b.priors_sum, from_probs=True, threads=threads, want_post=False)
        dt = time.perf_counter() - t0
        return trials / dt, dt, int(wl.success(b, hard).sum())
    b = wl.generate(trials, seed=12345)
    orc = O.BinaryBpOracle(csr=(wl.row_ptr, wl.col), shape=(wl.m, wl.n), max_iter=iters, channel_probs=wl.channel_probs)
    t0 = time.perf_counter()
    out = orc.decode_batch(b.syndromes, threads=threads)
    dt = time.perf_counter() - t0
    return trials / dt, dt, int(wl.success(b, out["decoding"]).sum())


def cpu_sample_size(kind, cores, per_step=False):
    # q-ary: the reference's brute-force check node costs 1-3.5 s per decode and core -> 4 decodes per thread is a
    # 4-14 s sample.  binary: ~5 ms per decode and core -> 512 decodes per thread (~3 s + generating the syndromes).
    if kind == "qary-generic":  # near-deterministic priors: the finite-support enumeration is ~1.5 ms per decode
        return (500 if per_step else 4000) * cores
    if per_step:  # --impl reference: one step of K; K steps must stay within a few minutes
        return cores if kind == "qary" else 128 * cores
    return 4 * cores if kind == "qary" else 512 * cores


def run_reference(args, kind, wl, iters, cfg):
    from oracle import oracle as O

    rank, world, _ = dist_env()
    if rank != 0:
        return
    cores = host_cores()
    if not O.has_openmp() and cores > 1:  # an oracle built without OpenMP would silently run on one thread
        raise SystemExit("the CPU oracle was built without OpenMP: refusing to time the reference arm on 1 of "
                         f"{cores} cores (rebuild oracle/ with gcc -fopenmp)")
    n = cpu_sample_size("qary-generic" if getattr(wl, "BSUM", 1) == 0 else kind, cores, per_step=True)
    for _ in range(args.warmup):  # warm-up: thread pool, page-in (one iteration of one decode per thread)
        cpu_path(kind, wl, iters, min(n, cores), cores, warm_only=True)
    total_t, total_n = 0.0, 0
    for _ in range(args.steps):
        _, dt, _ = cpu_path(kind, wl, iters, n, cores)
        total_t += dt
        total_n += n
    value = total_n / total_t
    sample = f"{n} trials per step ({'1 decode' if kind == 'qary' else '128 decodes'} per thread), {args.steps} steps"
    cfg = dict(cfg, workload=cfg["workload"].replace(f"{args.trials} trials per GPU per step", f"{n} trials per step on the host"),
               trials_per_gpu=None, trials_per_step=n)
    line = {
        "impl": "reference", "metric": "decoded trials/sec", "value": value, "unit": "trials/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32" if kind == "qary" else "f64", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": "trials/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "edge_updates_per_sec": value * cfg["edges"] * (iters if kind == "qary" else 1),
        "note": "CPU restatement of the reference algorithm (oracle/, brute-force check node, one decode per thread); "
                "the reference's own Rust crate cannot be built in this image (no cargo/rustc)",
    }
    print(json.dumps(line), flush=True)


# =====================================================================================================================
# our arm
# =====================================================================================================================
def run_ours(args, kind, wl, iters, cfg):
    import torch
    import torch.distributed as dist

    from sca_ldpc_b200 import QaryDecoder
    from sca_ldpc_b200.ldpc import bp_decoder

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    T = args.trials
    seed = 1000 + rank  # every rank decodes its own independent trials
    # nvidia-smi sampler: started long before the timed region, so that its NVML start-up (hundreds of ms of driver
    # calls that stall kernel launches) is over by then; during the timed region it only polls every 100 ms
    clocks = ClockSampler(local)
    if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS"):
        clocks.start()

    # ---- inputs: generated on the host, resident in HBM for `value`, in pinned host memory for `e2e`
    batch = wl.generate(T, seed=seed)
    if kind == "qary":
        dec = QaryDecoder(wl.H, iters, device=local, **wl.decoder_kwargs)
        special = dec.special
        h_in = [torch.from_numpy(batch.priors).pin_memory()]
        if special:
            h_in.append(torch.from_numpy(batch.priors_sum).pin_memory())
        d_in = [t.to(dev) for t in h_in] + [None]
        h_np = [t.numpy() for t in h_in] + [None]
        secret = torch.from_numpy(batch.secret).to(dev)
        nssum = torch.from_numpy((-batch.ssum).astype(np.int8)).to(dev) if special else None
        bv = wl.bvars
        t_T = torch.tensor(T, device=dev)

        def count(hard):
            ok = (hard[:, :bv] == secret).all(1)
            if special:
                ok = ok & (hard[:, bv:] == nssum).all(1)
            return torch.stack((ok.sum(), t_T))

        def step_dev():
            hard, _, _ = dec.min_sum_batch(d_in[0], d_in[1], want_posterior=False)
            return count(hard)

        h_out = torch.empty((T, wl.N), dtype=torch.int8).pin_memory().numpy()  # pinned result buffer (D2H target)

        def step_host():
            hard, _, _ = dec.min_sum_batch(h_np[0], h_np[1], want_posterior=False, out=h_out)
            return hard

        # the same trials as oracle observations: row indices + prior tables instead of vectors
        has_observed = hasattr(wl, "generate_observed") and not args.brief
        ob = wl.generate_observed(T, seed=seed) if has_observed else None
        if has_observed:
            h_obs = [torch.from_numpy(ob.obs.view(np.int16)).pin_memory(), torch.from_numpy(ob.obs_sum.view(np.int16)).pin_memory()]
            d_obs = [t.to(dev) for t in h_obs]
            d_tab = [torch.from_numpy(wl.table).to(dev), torch.from_numpy(wl.table_sum).to(dev)]

            def step_dev_obs():
                hard, _, _ = dec.min_sum_observed(d_tab[0], d_obs[0], d_tab[1], d_obs[1])
                return count(hard)

            def step_host_obs():
                hard, _, _ = dec.min_sum_observed(wl.table, h_obs[0].numpy().view(np.uint16), wl.table_sum,
                                                  h_obs[1].numpy().view(np.uint16), out=h_out)
                return hard

        h2d = sum(t.numel() * t.element_size() for t in h_in)
        d2h = T * wl.N
        edges = dec.E
        iters_per_trial = lambda: float(max(1, iters))  # noqa: E731  (no early termination on the q-ary path)
        bytes_iter = dec.algorithmic_bytes_per_trial_iteration()
        # dominant kernel = check-node pass (middle iteration): reads + writes every B-edge message slot once and
        # reads the sum-variable priors; the sum-variables' own messages never travel
        cn_bytes = 8 * (dec.msg_slots - (dec.R * dec.QS if dec.special else 0)) + 4 * (dec.R * dec.QS if dec.special else 0)
        vn_bytes = 8 * (dec.msg_slots - (dec.R * dec.QS if dec.special else 0)) + 4 * dec.bvars * dec.Q
        cn_name = (f"k_cn_fast<{wl.sum_weight},{wl.B}> (check-node pass)" if special else
                   f"k_cn_uniform<{dec.Q}> (check-node pass, generic decoder, reachable-window tables in shared memory)")
    else:
        dec = bp_decoder(csr=(wl.row_ptr, wl.col), shape=(wl.m, wl.n), max_iter=iters, channel_probs=wl.channel_probs,
                         device=local)
        h_in = [torch.from_numpy(batch.syndromes).pin_memory()]
        d_in = [t.to(dev) for t in h_in]
        truth = torch.zeros((T, wl.N), dtype=torch.uint8)
        truth.scatter_(1, torch.from_numpy(batch.y), 1)
        truth = truth.to(dev)
        last = {}

        t_T = torch.tensor(T, device=dev)

        def step_dev():
            out = dec.decode_batch(d_in[0])
            last["iters"] = out["iters"]
            ok = (out["decoding"][:, : wl.N] == truth).all(1)
            return torch.stack((ok.sum(), t_T))

        h_out = torch.empty((T, wl.n), dtype=torch.uint8).pin_memory().numpy()  # pinned result buffer (D2H target)

        def step_host():
            return dec.decode_batch(h_in[0].numpy(), out=h_out)["decoding"]

        h2d = T * wl.m
        d2h = T * wl.n
        edges = dec.E
        iters_per_trial = lambda: float(last["iters"].float().mean().item())  # noqa: E731
        bytes_iter = dec.algorithmic_bytes_per_trial_iteration()
        cn_bytes = 8 * dec.E + (wl.m + 7) // 8
        vn_bytes = 8 * dec.E + 4 * wl.n

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident inputs, CUDA events, max over ranks
    counters = None
    for _ in range(max(3, args.warmup)):  # warm-up runs everything the timed loop runs (CUDA loads kernels lazily: the
        c = step_dev()                    # first launch of any kernel, torch's included, stalls the device ~10 ms)
        counters = c if counters is None else counters + c
    sync_all()
    l0 = dec.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    tw0 = time.perf_counter()
    e0.record()
    counters = None
    dbg = [] if os.environ.get("BENCH_DEBUG_STEPS") else None
    for _ in range(args.steps):
        c = step_dev()
        counters = c if counters is None else counters + c
        if dbg is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            dbg.append(ev)
    if world > 1:
        dist.all_reduce(counters)  # the path's only exchange: success / trial counters, once per job
    e1.record()
    sync_all()
    tw1 = time.perf_counter()
    launches = dec.launch_count - l0
    if dbg:
        print("per-step ms:", [round(a.elapsed_time(b), 3) for a, b in zip([e0] + dbg[:-1], dbg)], file=sys.stderr)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clk = clocks.stop(tw0, tw1) if rank == 0 else None
    successes, trials_total = (int(x) for x in counters.tolist())
    ipt = iters_per_trial()

    # ---- e2e: host buffers through the public API (H2D of the priors + D2H of the decisions inside the timed region)
    for _ in range(2):
        step_host()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = world * T * args.steps / float(dt.item())

    # ---- the same trials fed as oracle observations (q-ary only): device-resident and host-buffer throughput
    observed = None
    if kind == "qary" and has_observed:
        for _ in range(3):
            c_obs = step_dev_obs()
        sync_all()
        e0.record()
        for _ in range(args.steps):
            c_obs = step_dev_obs()
        e1.record()
        sync_all()
        ms_obs = torch.tensor([e0.elapsed_time(e1)], device=dev)
        for _ in range(2):
            step_host_obs()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_host_obs()
        torch.cuda.synchronize()
        dt_obs = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms_obs, op=dist.ReduceOp.MAX)
            dist.all_reduce(dt_obs, op=dist.ReduceOp.MAX)
        observed = {
            "api": "scaldpc_qary_decode_observed: uint16 row indices into the oracle's posterior tables instead of "
                   "prior vectors (SURVEY.md 8f.1); results bit-identical to the prior-vector API",
            "value": world * T * args.steps / (float(ms_obs.item()) * 1e-3), "unit": "trials/s",
            "e2e": {"value": world * T * args.steps / float(dt_obs.item()), "unit": "trials/s",
                    "h2d_bytes_per_step": sum(t.numel() * t.element_size() for t in h_obs) + wl.table.nbytes + wl.table_sum.nbytes,
                    "d2h_bytes_per_step": T * wl.N},
            "success_rate": int(c_obs[0]) / max(1, int(c_obs[1])),
        }

    # ---- binary path: host buffers in, hqc.decode's statistics out (scaldpc_bp_decode_scored) instead of T x n decisions
    scored = None
    if kind == "bp":
        y32 = np.ascontiguousarray(batch.y, dtype=np.int32)
        synd_h = h_in[0].numpy()
        for _ in range(2):
            sc = dec.decode_scored(synd_h, y32)
        sync_all()
        t0 = time.perf_counter()
        ok = 0
        for _ in range(args.steps):
            sc = dec.decode_scored(synd_h, y32)
            ok += int(sc["stats"][:, 5].sum())
        dt_sc = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        cnt_sc = torch.tensor([ok], device=dev, dtype=torch.int64)
        if world > 1:
            dist.all_reduce(dt_sc, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt_sc)
        scored = {
            "api": "scaldpc_bp_decode_scored: syndromes + true supports in, the statistics of hqc.py:709-758 out "
                   "(computed on the device, SURVEY.md 8f.2) instead of the T x n decision bytes",
            "e2e": {"value": world * T * args.steps / float(dt_sc.item()), "unit": "trials/s",
                    "h2d_bytes_per_step": T * wl.m + y32.nbytes + wl.n * 8, "d2h_bytes_per_step": T * (24 + 8)},
            "success_rate": int(cnt_sc.item()) / (world * T * args.steps)}

    # ---- the whole Monte-Carlo loop on the device (scaldpc_qary_simulate): trial synthesis + decode + scoring
    if observed is not None and special:
        first = rank * T * (args.steps + 3)
        for i in range(3):
            sim = dec.simulate(T, 20260, first_trial=first + i * T, **wl.simulate_kwargs)
        sync_all()
        t0 = time.perf_counter()
        ok = 0
        for i in range(args.steps):
            sim = dec.simulate(T, 20260, first_trial=first + (3 + i) * T, **wl.simulate_kwargs)
            ok += sim["successes"]
        dt_sim = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        cnt_sim = torch.tensor([ok], device=dev, dtype=torch.int64)
        if world > 1:
            dist.all_reduce(dt_sim, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt_sim)
        observed["simulate"] = {
            "api": "scaldpc_qary_simulate: secrets, check sums, oracle bit flips (Philox4x32-10 keyed by global trial "
                   "index), decode and success counting all on the device; 40 bytes of counters per call cross PCIe",
            "value": world * T * args.steps / float(dt_sim.item()), "unit": "trials/s (host wall clock)",
            "success_rate": int(cnt_sim.item()) / (world * T * args.steps)}

    # ---- latency of ONE decode through the reference-shaped call (min_sum / decode on host arrays, T = 1)
    single = None
    if rank == 0 and not args.brief:
        if kind == "qary":
            one = (lambda: dec.min_sum(batch.priors[0], batch.priors_sum[0])) if special else (lambda: dec.min_sum(batch.priors[0]))
        else:
            one = lambda: dec.decode(batch.syndromes[0])  # noqa: E731
        for _ in range(3):
            one()
        t0 = time.perf_counter()
        n_one = 50
        for _ in range(n_one):
            one()
        ms_one = 1e3 * (time.perf_counter() - t0) / n_one
        single = {"call": "min_sum(channel_output, channel_output_sum)" if kind == "qary" else "decode(syndrome)",
                  "ms_per_call": ms_one, "trials_per_s": 1e3 / ms_one}
        if kind == "qary":  # the reference's FER loop: one decoder object, a ThreadPoolExecutor (decode.py:247-262)
            from concurrent.futures import ThreadPoolExecutor

            nthr, ncall = 16, 1024
            with ThreadPoolExecutor(max_workers=nthr) as ex:
                call = (lambda i: dec.min_sum(batch.priors[i], batch.priors_sum[i])) if special else (lambda i: dec.min_sum(batch.priors[i]))
                list(ex.map(call, range(64)))
                r0, b0 = dec.combining_stats()
                t0 = time.perf_counter()
                list(ex.map(call, range(ncall)))
                dt = time.perf_counter() - t0
                r1, b1 = dec.combining_stats()
                single["threadpool"] = {"threads": nthr, "calls": ncall, "trials_per_s": ncall / dt,
                                        "calls_per_batch": (r1 - r0) / max(1, b1 - b0),
                                        "note": "concurrent min_sum() calls on one decoder are combined into batches inside "
                                                "the library (scaldpc_qary_decode_one)"}

    # ---- roofline of the dominant kernel (rank 0, N=1 semantics: one GPU's kernel against one GPU's peak)
    line = None
    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        st = torch.cuda.current_stream(dev).cuda_stream
        step_dev()
        torch.cuda.synchronize()
        cn_ms = dec.time_kernel(0, T, reps=20, stream=st)
        vn_ms = dec.time_kernel(1, T, reps=20, stream=st)
        achieved = T * cn_bytes / (cn_ms * 1e-3) / 1e9
        vn_achieved = T * vn_bytes / (vn_ms * 1e-3) / 1e9
        k_cn = {"kernel": cn_name if kind == "qary" else "k_bp_cn (check-to-bit pass)", "achieved": achieved,
                "frac": achieved / peak, "ms_per_launch": cn_ms, "algorithmic_bytes_per_launch": T * cn_bytes}
        k_vn = {"kernel": ("k_vn_reg / k_vn (variable-node pass)" if kind == "qary" else "k_bp_vn (bit-to-check pass)"),
                "achieved": vn_achieved, "frac": vn_achieved / peak, "ms_per_launch": vn_ms,
                "algorithmic_bytes_per_launch": T * vn_bytes}
        # the dominant kernel = the one the step spends more time in: the check pass on the q-ary path, the BIT pass on
        # the binary path
        dom, sec = (k_cn, k_vn) if cn_ms >= vn_ms else (k_vn, k_cn)
        roofline = {
            # the ROOFLINE is HBM bandwidth (irregular gather + small (min,+) convolutions); what ncu says
            # actually limits the q-ary check pass is instruction issue (profiles/r2zn_ncu_cn_fast_k512.txt), see `limiter`
            "bound": "hbm", "kernel": dom["kernel"],
            "achieved": dom["achieved"], "peak": peak, "unit": "GB/s", "frac": dom["frac"], "traffic": None,
            "peak_source": peak_src, "ms_per_launch": dom["ms_per_launch"],
            "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_launch"],
            "limiter": ("issue (two warps per scheduler at 255 registers; ncu profiles/r2zn_ncu_cn_fast_k512.txt: issue active 55 %, "
                         "dispatch stall + math-pipe throttle, ~1 870 warp-instructions per item of which ~1 530 are the FADD2 / "
                         "FMNMX3 the (min,+) candidates need)"
                        if kind == "qary" and special else
                        ("shared-memory latency, one warp per scheduler" if kind == "qary" else
                         "latency of the scattered column gather (ncu: long_scoreboard)")),
            "second_kernel": sec,
            "note": ("q-ary: the sum-variables' messages are not moved in middle iterations, 38 % fewer bytes than SURVEY.md "
                     "8(d) assumes for this pass; see `iteration` for the 8(d) figure"
                     if kind == "qary" else "binary: the bit pass is the slower of the two"),
            "iteration": {"bytes_per_trial_iteration_survey_8d": bytes_iter,
                          "achieved": T * bytes_iter / ((cn_ms + vn_ms) * 1e-3) / 1e9,
                          "frac": T * bytes_iter / ((cn_ms + vn_ms) * 1e-3) / 1e9 / peak},
        }
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file):
            try:
                tr = json.load(open(traffic_file)).get(cfg["workload"].split(":")[0])
                if tr and tr.get("match", "k_cn") in roofline["kernel"]:
                    roofline["traffic"] = tr["dram_bytes_per_launch"] * (T / tr["trials"])
                    roofline["traffic_source"] = tr.get("source")
            except Exception:
                pass

        # ---- parity of the timed path on the FULL batch (not timed): the default decode (fast check node + verification)
        # against the exact-order kernel (= the reference's arithmetic bit for bit, tests/test_gpu_qary.py) on every trial
        parity = None
        if kind == "qary":
            hd, pd, psd = dec.min_sum_batch(d_in[0], d_in[1])
            run = dec.last_run
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            hx, px, psx = dec.min_sum_batch(d_in[0], d_in[1], exact=True)
            torch.cuda.synchronize()
            exact_ms = 1e3 * (time.perf_counter() - t0)
            hu, _, _ = dec.min_sum_batch(d_in[0], d_in[1], want_posterior=False, verify=False)

            def maxdiff(a, b_):
                if a is None:
                    return None
                fin = torch.isfinite(a) & torch.isfinite(b_)
                same = bool((torch.isnan(a) == torch.isnan(b_)).all() and (torch.isinf(a) == torch.isinf(b_)).all())
                return {"max_abs": float((a - b_)[fin].abs().max().item()) if bool(fin.any()) else 0.0, "same_inf_nan_masks": same}

            e0.record()
            for _ in range(args.steps):
                dec.min_sum_batch(d_in[0], d_in[1], want_posterior=False, verify=False)
            e1.record()
            torch.cuda.synchronize()
            unverified_ms = e0.elapsed_time(e1) / args.steps
            parity = {
                "against": "exact-order check node on the GPU (SCALDPC_CN_EXACT: the reference's sequential f32 sums, "
                           "decoder_special.rs:539-553, bit-identical to the oracle's enumeration)",
                "trials": T, "hard_mismatch_trials": int((hd != hx).any(1).sum().item()),
                "max_abs_post_diff": maxdiff(pd, px)["max_abs"], "max_abs_post_sum_diff": (maxdiff(psd, psx) or {}).get("max_abs"),
                "same_inf_nan_masks": maxdiff(pd, px)["same_inf_nan_masks"],
                "redecoded": run["redecoded"], "risk_margin": 1e-2,
                "unverified_hard_mismatch_trials": int((hu != hx).any(1).sum().item()),
                "exact_mode_ms_per_step": exact_ms,
                "unverified_ms_per_step": unverified_ms, "verified_ms_per_step": ms / args.steps,
            }

        # ---- opt-in early termination (SCALDPC_QARY_EARLY_STOP; an extension -- the reference always runs every iteration,
        # so `value` above is NOT this mode): throughput, iterations per trial, and how many trials' decisions differ from
        # the fixed-iteration decode
        early = None
        if kind == "qary" and special:
            for _ in range(3):
                he, _, _ = dec.min_sum_batch(d_in[0], d_in[1], want_posterior=False, early_stop=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(args.steps):
                he, _, _ = dec.min_sum_batch(d_in[0], d_in[1], want_posterior=False, early_stop=True)
            e1.record()
            torch.cuda.synchronize()
            es_ms = e0.elapsed_time(e1) / args.steps
            run = dec.last_run
            its = dec.last_iterations(T)
            early = {
                "flag": "SCALDPC_QARY_EARLY_STOP: per-trial freeze on the row test + compacted restart of the stragglers; "
                        "never-converging trials re-decoded exact-order with the fixed iteration count",
                "value": T / (es_ms * 1e-3), "unit": "trials/s (one GPU, device-resident priors)", "ms_per_step": es_ms,
                "iterations_per_trial": float(its.mean()), "max_iterations": int(its.max()),
                "trial_iterations_launched_per_trial": run["trial_iterations"] / T, "redecoded": run["redecoded"],
                "differing_trials": int((he != hd).any(1).sum().item()),
                "differing_trials_that_ran_to_the_cap": int(((he != hd).any(1).cpu().numpy() & (its >= iters)).sum()),
                "success_rate": float(count(he)[0].item()) / T,
            }

        # ---- cpu_baseline (N=1 only): bounded sample of the same workload on the host cores
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            from oracle import oracle as O

            cores = host_cores()
            n = cpu_sample_size("qary-generic" if getattr(wl, "BSUM", 1) == 0 else kind, cores)
            v, secs, _ = cpu_path(kind, wl, iters, n, cores)
            cpu = {"value": v, "unit": "trials/s", "cores": cores, "kind": "port",
                   "sample": f"{n} trials of this workload, one decode at a time per thread "
                             f"({'brute-force check node as decoder_special.rs:531-554' if kind == 'qary' else 'fp64 product-sum'}), "
                             f"{secs:.1f} s"}
            if kind == "qary":
                # SURVEY.md 8d / BASELINE.md 3.3: an OPTIMISED CPU figure (forward/backward check node, the same algorithm
                # class as the GPU kernel) so that the speed-up is not mostly algorithmic
                n2 = 16 * cores
                v2, secs2, _ = cpu_path(kind, wl, iters, n2, cores, algo=O.ALGO_FWD_BWD)
                cpu["optimised"] = {"value": v2, "unit": "trials/s", "cores": cores,
                                    "algorithm": "forward/backward (min,+) check node (oracle algo 2), one decode per thread",
                                    "sample": f"{n2} trials, {secs2:.1f} s"}
        value = world * T * args.steps / (ms * 1e-3)
        line = {
            "metric": "decoded trials/sec", "value": value, "unit": "trials/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "edge_updates_per_sec": value * edges * ipt, "iterations_per_trial": ipt,
            "success_rate": successes / max(1, trials_total),
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": "trials/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            **({"observed": observed} if observed else {}),
            **({"scored": scored} if scored else {}),
            "single_decode": single,
            **({"compactions_last_step": dec.last_run()[1], "iterations_launched_last_step": dec.last_run()[0]}
               if kind == "bp" else {}),
            "roofline": roofline,
            **({"parity": parity} if parity else {}),
            **({"early_stop": early} if early else {}),
            "cpu_baseline": cpu,
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None and world == 1 and args.workload == "kyber512" and not args.no_sub:
        # north_star quotes its >= 100x / >= 60 % target on Kyber-768: the same measurement on that workload, run after
        # every timed region of this process is over (a second process of this script, its JSON line embedded)
        line["kyber768"] = sub_bench("kyber768", args)
    if line is not None:
        print(json.dumps(line), flush=True)


def sub_bench(workload, args):
    keep = ("value", "unit", "ms_per_step", "steps", "config", "success_rate", "e2e", "gpu_launches", "roofline", "parity", "early_stop",
            "cpu_baseline", "clocks", "edge_updates_per_sec")
    cmd = [sys.executable, os.path.abspath(__file__), "--workload", workload, "--steps", str(args.steps), "--warmup",
           str(args.warmup), "--no-sub", "--brief"]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        rows = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
        if out.returncode != 0 or not rows:
            return {"error": (out.stderr or out.stdout)[-400:]}
        full = json.loads(rows[-1])
        return {k: full[k] for k in keep if k in full}
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kyber512", choices=sorted(WORKLOADS))
    ap.add_argument("--trials", type=int, default=0, help="trials per GPU per step (default: the workload's)")
    ap.add_argument("--iters", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="do not append the kyber768 sub-object to the default line")
    ap.add_argument("--brief", action="store_true", help="skip the observed / simulate / single-decode legs")
    args = ap.parse_args()

    from sca_ldpc_b200.workloads import HqcWorkload, KyberWorkload

    kind, kw, T, iters = WORKLOADS[args.workload]
    args.trials = args.trials or T
    iters = args.iters or iters
    if kw.get("generic"):
        from sca_ldpc_b200.workloads import DecoderBenchWorkload

        z = np.load(os.path.join(ROOT, "tests", "golden", "rust_unit_tests.npz"))  # benches/parity_check_150_450.txt
        Hm = np.zeros(tuple(int(x) for x in z["Hm_shape"]), np.int8)
        Hm[z["Hm_rows"], z["Hm_cols"]] = z["Hm_vals"]
        wl = DecoderBenchWorkload(Hm, B=7)
    else:
        wl = KyberWorkload(seed=0, **kw) if kind == "qary" else HqcWorkload(seed=0, **kw)
    if kw.get("generic"):
        cfg = {"workload": f"{args.workload}: generic q-ary min-sum decoder Decoder<450,150,3,7,15,7> on the reference's criterion "
                           f"input (benches/decoder.rs:38-62), {args.trials} trials per GPU per step, {iters} iterations",
               "trials_per_gpu": args.trials, "iterations": iters, "edges": int((wl.H != 0).sum()),
               "check_node": "forward/backward (min,+), k_cn_uniform<Q>: Q templated, row weights at run time", "inputs": "larger than L2", "oracle_accuracy": None}
    elif kind == "qary":
        cfg = {"workload": f"{args.workload}: q-ary min-sum BP, H {wl.R}x{wl.N} (sum_weight {wl.sum_weight}, "
                           f"check_blocks {wl.check_blocks}), alphabets {wl.Q}/{wl.QS}, {args.trials} trials per GPU per step, "
                           f"{iters} iterations, no early termination (reference semantics)",
               "trials_per_gpu": args.trials, "iterations": iters, "edges": int((wl.H != 0).sum()),
               "check_node": "forward/backward (min,+)", "inputs": "larger than L2 (priors + messages > 1 GB per step)",
               "oracle_accuracy": wl.accuracy}
    else:
        cfg = {"workload": f"{args.workload}: binary product-sum BP, n = {wl.N}+{wl.R}, row weight {wl.W}+1, "
                           f"{args.trials} trials per GPU per step, max_iter {iters}, syndrome early stop",
               "trials_per_gpu": args.trials, "iterations": iters, "edges": int(wl.row_ptr[-1]),
               "inputs": "larger than L2 (messages > 1 GB per step)", "rho": wl.rho}
    if args.impl == "reference":
        run_reference(args, kind, wl, iters, cfg)
    else:
        run_ours(args, kind, wl, iters, cfg)


if __name__ == "__main__":
    main()
