#!/usr/bin/env python
"""bench.py -- path-months/s of the Monte Carlo portfolio hot path on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3compat|c4|c4boot] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A "step" is one pass of the hot path over one batch of synthetic paths on every GPU: the fused kernel (RNG ->
compounding in registers -> per-block mean/variance partials -> warp/CTA/grid reduction -> histogram + P(target))
followed, for N > 1, by the ONE all-reduce of the O(KB) statistics vector (NCCL over NVLink).  Per-GPU work is fixed
as N grows (weak scaling); rank r of step s simulates path ids [(s*N + r) * B, +B) -- disjoint RNG streams, never a
cached result.

Workloads (BASELINE.json configs):
  c2        360 months x 100 M paths per GPU, Gaussian(0.5 %, 10/12 %), per-256-path-block mean/variance partials,
            100-bin histogram, P(final < V0)                     <- default; the config the metric is quoted on
  c3        360 x 100 M bootstrap (Philox) from the synthetic 1200-month S&P-500-shaped table (shared-memory resident)
  c3compat  same through the reference's own RNG stream (bit-exact drop-in of src/simulations.cu)
  c4        1000 x 5 M Sobol (Joe-Kuo) Gaussian, histogram + P(target);  c4boot: Sobol bootstrap
  c5        360 x 10^10 Gaussian + histogram: ONE run's paths sharded over the N GPUs (strong scaling; 1.8 s per step on 1 GPU)

Printed: ONE JSON line (rank 0).  `value` is device-timed (CUDA events on the launching stream, summed over the K
steps, max over ranks) with everything resident in HBM; `e2e` is the same metric through the public host-buffer
entry point (smmc_run via simulate(): table H2D, results D2H inside the timed region); `roofline` is the simulation
kernel against the instruction-issue roofline (this path is ALU/RNG-bound: SURVEY 8(d)); `cpu_baseline` is the
reference's own CPU implementation (oracle/_ref, compiled from the reference sources) on this box's host cores.

`--impl reference` times that CPU implementation alone (all host threads it uses, bounded sample per step).
oracle/ is only ever the checker / the CPU baseline here, never the thing measured as the product.
"""
import argparse
import json
import os
import re
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "path-months/sec (device-timed)"
UNIT = "path-months/s"

# Algorithmic thread-instructions per path-month of the IMPLEMENTED recipes (minimal SASS-level count, DESIGN.md
# section 4; restates SURVEY 8(d) for the stream contract actually shipped).  SURVEY 8(d) budgeted 19 (bootstrap:
# 4 draws per Philox call + IMAD.HI index) and 22 (gaussian: 2 normals per 2 words); the shipped recipes draw 6
# fields / 8 normals per call and need 12, so the fraction below stays an issue-slot utilisation (<= 1).  The
# SURVEY-budget fraction is reported next to it as `frac_survey_budget`.
I_ALG = {"bootstrap_compat": 12, "bootstrap": 12, "gaussian": 12, "sobol_bootstrap": 6, "sobol_gaussian": 9}
# Philox4x32-7 (smmc_config.philox_rounds = 7): three rounds = 6 IMAD.WIDE + 6 LOP3 less per call of 6 / 8 months
I_ALG_PHILOX7 = {"bootstrap": 12 - 12 / 6.0, "gaussian": 12 - 12 / 8.0}
# Pipe floors (cycles per warp-month per SM sub-partition) from the measured reciprocal throughputs in
# profiles/r02_pipe_microbench.txt and profiles/r01_pipe_mix_microbench.txt: the busiest single pipe
# of the shipped instruction mix.  (The minimal-count recipes behind I_ALG are the all-MUFU ones; the shipped Gaussian
# kernels trade MUFU for a few more FMA/ALU instructions, so I_ALG is NOT raised for them.)
#   gaussian   pairs 0 and 2 of a call take sin/cos from the shared-memory table: (2 x 4 + 2 x 2) MUFU x 8.0 / 8 months = 12.0 on
#              the XU pipe; Philox keeps fmaheavy at 9.0 and the ALU pipe at 8.5 -- no single pipe binds (ncu: XU 65 %, issue 72 %)
#   sobol_gaussian  six of a lane's eight path slots from the table, two by MUFU.SIN/COS: XU (2 + 0.25 x 2) x 8.0 / 2 = 10.0; the
#              kernel is issue-bound (ncu: 11.5 executed instructions per path-month, issue 84.7 %, XU 73.7 %)
#   bootstrap  Philox: 8 per-lane rounds x 8.39 cycles (2 IMAD.WIDE + 2 LOP3, measured as a unit) per 6 months,
#              + 16 further ALU-pipe instructions per call (fields, addresses) at their measured marginal 2.0 cycles
PIPE_FLOOR = {"gaussian": ("xu", 12.0), "sobol_gaussian": ("xu", 10.0),
              "bootstrap": ("fmaheavy+alu (Philox)", (8 * 8.39 + 16 * 2.0) / 6.0)}
I_ALG_SURVEY = {"bootstrap_compat": 12, "bootstrap": 19, "gaussian": 22, "sobol_bootstrap": None, "sobol_gaussian": None}

WORKLOADS = {
    "c2": dict(mode="gaussian", paths=100_000_000, periods=360, block_partials=True, hist_range=(0.0, 50_000.0),
               desc="360 months x 100M paths Gaussian(0.5%,0.8333%) per-block mean/variance partials + histogram + P(target)"),
    "c3": dict(mode="bootstrap", paths=100_000_000, periods=360, block_partials=True, hist_range=(0.0, 200_000.0),
               desc="360 months x 100M paths Philox bootstrap from synthetic 1200-month S&P500-shaped table (smem-resident)"),
    "c3compat": dict(mode="bootstrap_compat", paths=100_000_000, periods=360, block_partials=True,
                     hist_range=(0.0, 200_000.0),
                     desc="360 months x 100M paths bootstrap, reference RNG stream (drop-in of src/simulations.cu)"),
    "c5": dict(mode="gaussian", paths=10_000_000_000, periods=360, block_partials=False, hist_range=(0.0, 50_000.0),
               strong=True,
               desc="360 months x 10^10 paths Gaussian + histogram + P(target), the 10^10 paths of ONE run sharded over the GPUs (strong scaling)"),
    "c4": dict(mode="sobol_gaussian", paths=5_000_000, periods=1000, block_partials=False, hist_range=(0.0, 500_000.0),
               desc="1000 months x 5M paths Sobol (Joe-Kuo) Gaussian (Box-Muller on dimension pairs), histogram + P(target)"),
    "c4boot": dict(mode="sobol_bootstrap", paths=5_000_000, periods=1000, block_partials=False,
                   hist_range=(0.0, 1.0e8),
                   desc="1000 months x 5M paths Sobol (Joe-Kuo) bootstrap, histogram + P(target)"),
}
HIST_BINS = 100
V0 = 1000.0
GAUSS = dict(gauss_mean=0.5, gauss_std=10.0 / 12.0)


def ncu_traffic(mode):
    """dram bytes per launch of the simulation kernel from the committed ncu capture (profiles/ncu_traffic.json)"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["bytes_per_launch"].get(mode)
    except Exception:
        return None


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


# ---- clocks ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi sampled every 100 ms while the timed region runs (B200_PROFILING.md 'clocks' recipe)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_indices):
        self.rows = []
        self.proc = None
        self.mark0 = self.mark1 = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", ",".join(str(i) for i in gpu_indices)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def begin(self):
        self.mark0 = time.perf_counter()

    def end(self):
        self.mark1 = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        for t, line in self.rows:
            if self.mark0 is None or not (self.mark0 <= t <= (self.mark1 or t) + 0.1):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_max": max(power) if power else None}


# ---- CPU baseline: the reference's own CPU implementation (oracle/_ref) or, failing that, the oracle port -----------
class CpuReference:
    def __init__(self, mode, periods):
        from oracle import binding as O
        import stock_market_monte_carlo_b200 as S
        self.O, self.mode, self.periods = O, mode, periods
        self.table = S.synthetic_sp500_table(1200)
        self.kind = "reference" if O.ref_available("libref_cpu.so") else "port"
        self.cores = None
        self.desc = None

    def run(self, n_paths):
        """simulate n_paths paths x periods months on the host cores; returns seconds"""
        O = self.O
        n_paths = max(int(n_paths), 10_000)  # the reference divides by n_blocks/10 (src/simulations.cpp:255)
        t0 = time.perf_counter()
        if self.kind == "reference":
            if "gaussian" in self.mode:
                _, self.cores = O.ref_gaussian_finals(n_paths, self.periods, V0, GAUSS["gauss_mean"], GAUSS["gauss_std"])
                self.desc = "reference sample_returns_gaussian+many_updates (src/simulations.cpp:41-55,:24-39) under OpenMP"
            else:
                _, self.cores = O.ref_mc_simulations(n_paths, self.periods, V0, self.table)
                self.desc = "reference CPU v2 mc_simulations (src/simulations.cpp:204-266), OpenMP"
        else:
            kw = dict(returns=self.table) if "bootstrap" in self.mode else {}
            O.simulate(self.mode, n_paths, self.periods, seed=1, **kw)
            self.cores = int(O.load().orc_num_threads())
            self.desc = "oracle/smmc_oracle.c port (OpenMP)"
        return time.perf_counter() - t0, n_paths

    def sized_sample(self, target_seconds):
        dt, n = self.run(20_000)
        rate = n / max(dt, 1e-6)
        return int(min(max(rate * target_seconds, 20_000), 50_000_000))


def run_reference_arm(args, wl):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cpu = CpuReference(wl["mode"], wl["periods"])
    sample = cpu.sized_sample(1.0)  # ~1 s of host work per step
    for _ in range(args.warmup):
        cpu.run(sample)
    total = 0.0
    for _ in range(args.steps):
        dt, n = cpu.run(sample)
        total += dt
    value = args.steps * sample * wl["periods"] / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "mode": wl["mode"], "periods": wl["periods"],
                   "paths_per_step": sample, "note": "bounded sample of the workload per step, host cores only"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.cores, "kind": cpu.kind,
                         "sample": f"{sample} paths x {wl['periods']} months per step; {cpu.desc}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---- auxiliary workload: reduce_mean_gpu, the one HBM-bound kernel of the path ------------------
def run_reduce_mean(args):
    """Mean of 2^28 device-resident floats (1 GiB, larger than the 126 MB L2) through smmc_reduce_mean_device.
    Algorithmic traffic: 4 bytes read per element, nothing written.  Peak: MEASURED_PEAKS.json hbm_gbs (a copy: read+write
    bytes), else the profiling guide's fallback."""
    import torch
    import stock_market_monte_carlo_b200 as S
    if S.load().smmc_device_count() < 1:
        raise SystemExit("bench.py needs a B200: the product has no CPU fallback")
    torch.cuda.set_device(0)
    n = args.paths or (1 << 28)
    steps, warm = (args.steps or 2000), max(args.warmup, 3)  # ~0.2 ms per step: long enough for the 100 ms clock samples
    bufs = [torch.full((n,), 1.0 + 0.25 * i, dtype=torch.float32, device="cuda") for i in range(2)]  # alternate: 2 GiB > L2
    stream = torch.cuda.current_stream().cuda_stream
    for i in range(warm):
        S.reduce_mean_device(bufs[i & 1].data_ptr(), n, stream=stream)
    sampler = ClockSampler([0])
    sampler.begin()
    ms, ok = [], True
    for i in range(steps):
        m, k = S.reduce_mean_device(bufs[i & 1].data_ptr(), n, stream=stream)
        ok = ok and (m == 1.0 + 0.25 * (i & 1))
        ms.append(k)
    sampler.end()
    clocks = sampler.stop()
    # end to end through the reference-named call with a HOST vector (H2D copy of 4 B/element inside the timed region)
    h = torch.ones(1 << 26, dtype=torch.float32).numpy()
    S.reduce_mean_gpu(h, h.size)
    t0 = time.perf_counter()
    for _ in range(3):
        S.reduce_mean_gpu(h, h.size)
    e2e_s = (time.perf_counter() - t0) / 3
    peaks = measured_peaks()
    peak = float(peaks.get("hbm_gbs") or 6650.0)
    k_ms = sum(ms) / len(ms)
    gbs = 4.0 * n / (k_ms * 1e-3) / 1e9
    line = {
        "metric": "elements/sec (reduce_mean_gpu, device-timed)", "value": n / (k_ms * 1e-3), "unit": "elements/s", "n_gpus": 1,
        "steps": steps, "warmup": warm, "ms_per_step": k_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 in, f64 accumulation", "data": "synthetic",
        "config": {"workload": f"reduce_mean_gpu over {n} device-resident floats ({4 * n / 2**30:.2f} GiB)",
                   "l2": "two 1 GiB buffers alternate between steps: every step reads from HBM"},
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                     "traffic": ncu_traffic("reduce_mean") if n == (1 << 28) else None, "algorithmic_bytes": 4 * n,
                     "kernel": "smmc::reduce_sum_tma_kernel" if n >= 148 * 2 * 4 * 8192 else "smmc::reduce_sum_kernel", "kernel_ms": k_ms, "bytes_per_element": 4,
                     "peak_source": ("MEASURED_PEAKS.json hbm_gbs (copy: read + write bytes)" if peaks.get("hbm_gbs")
                                     else "fallback 6650 GB/s (B200_PROFILING.md)")},
        "e2e": {"value": h.size / e2e_s, "unit": "elements/s", "h2d_bytes_per_step": int(h.nbytes), "d2h_bytes_per_step": 8,
                "api": "reduce_mean_gpu(host vector) -> smmc_reduce_mean: pageable H2D + kernel"},
        "gpu_launches": steps, "clocks": clocks, "result": {"exact": bool(ok)},
    }
    print(json.dumps(line), flush=True)
    return 0


def run_vector_add(args):
    """out = a + b over 2^27 device-resident floats per vector (3 x 512 MiB per step, larger than L2) through
    smmc_vector_add_device.  Algorithmic traffic: 12 bytes per element (two reads, one write)."""
    import torch
    import stock_market_monte_carlo_b200 as S
    if S.load().smmc_device_count() < 1:
        raise SystemExit("bench.py needs a B200: the product has no CPU fallback")
    torch.cuda.set_device(0)
    n = args.paths or (1 << 27)
    steps, warm = (args.steps or 2000), max(args.warmup, 3)  # ~0.2 ms per step: long enough for the 100 ms clock samples
    a = torch.full((n,), 1.25, dtype=torch.float32, device="cuda")
    b = torch.full((n,), 2.5, dtype=torch.float32, device="cuda")
    out = torch.empty_like(a)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(warm):
        S.vector_add_device(out.data_ptr(), a.data_ptr(), b.data_ptr(), n, stream=stream)
    sampler = ClockSampler([0])
    sampler.begin()
    ms = [S.vector_add_device(out.data_ptr(), a.data_ptr(), b.data_ptr(), n, stream=stream) for _ in range(steps)]
    sampler.end()
    clocks = sampler.stop()
    ok = bool(torch.all(out == 3.75).item())
    ha, hb = torch.ones(1 << 24).numpy(), torch.ones(1 << 24).numpy()
    ho = ha.copy()
    S.vector_add_gpu(ho, ha, hb, ha.size)
    t0 = time.perf_counter()
    for _ in range(3):
        S.vector_add_gpu(ho, ha, hb, ha.size)
    e2e_s = (time.perf_counter() - t0) / 3
    peaks = measured_peaks()
    peak = float(peaks.get("hbm_gbs") or 6650.0)
    k_ms = sum(ms) / len(ms)
    gbs = 12.0 * n / (k_ms * 1e-3) / 1e9
    line = {
        "metric": "elements/sec (vector_add_gpu, device-timed)", "value": n / (k_ms * 1e-3), "unit": "elements/s", "n_gpus": 1,
        "steps": steps, "warmup": warm, "ms_per_step": k_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"vector_add_gpu over {n} device-resident floats per vector ({12 * n / 2**30:.2f} GiB of traffic)",
                   "l2": "1.5 GiB touched per step: larger than L2"},
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                     "traffic": ncu_traffic("vector_add") if n == (1 << 27) else None, "algorithmic_bytes": 12 * n, "kernel": "smmc::vector_add_kernel", "kernel_ms": k_ms, "bytes_per_element": 12,
                     "peak_source": ("MEASURED_PEAKS.json hbm_gbs (copy: read + write bytes)" if peaks.get("hbm_gbs")
                                     else "fallback 6650 GB/s (B200_PROFILING.md)")},
        "e2e": {"value": ha.size / e2e_s, "unit": "elements/s", "h2d_bytes_per_step": int(2 * ha.nbytes),
                "d2h_bytes_per_step": int(ha.nbytes), "api": "vector_add_gpu(host arrays) -> smmc_vector_add: pageable H2D x2 + kernel + D2H"},
        "gpu_launches": steps, "clocks": clocks, "result": {"exact": ok},
    }
    print(json.dumps(line), flush=True)
    return 0


# ---- extras: every other claim of the build, device-timed in the same driver-run process (outside the headline's region) ---
def issue_peak(torch, dev):
    peaks = measured_peaks()
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    return n_sm * 128 * float(peaks.get("sm_max_mhz") or 1965.0) * 1e6, n_sm


def short_leg(torch, S, name, dev, local, warm=3, steps=5, philox_rounds=0, with_e2e=False):
    """3 warm-up + 5 timed steps of workload `name` on one GPU: value from CUDA events around every step on the launching
    stream, kernel_ms from the library's own event pairs, issue-roofline fraction; fresh RNG offsets every step."""
    wl = WORKLOADS[name]
    mode, B, P = wl["mode"], wl["paths"], wl["periods"]
    table = S.synthetic_sp500_table(1200) if "bootstrap" in mode else None
    table_dev = torch.from_numpy(table).to(dev) if table is not None else None
    seed = 0 if mode == "bootstrap_compat" else 1234
    nb = (B + 255) // 256
    stats = torch.zeros(S.stats_len(HIST_BINS), dtype=torch.float64, device=dev)
    bm = torch.empty(nb, dtype=torch.float32, device=dev) if wl["block_partials"] else None
    bv = torch.empty(nb, dtype=torch.float32, device=dev) if wl["block_partials"] else None
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    span = (1 << 31) - 256 - B if mode == "bootstrap_compat" else ((1 << 32) - 512 - B if mode.startswith("sobol") else 1 << 62)

    def off(i):
        return ((7_000 + i) * B % max(span, 1)) // 256 * 256

    def step(i):
        return S.simulate_device(mode, B, P, stats.data_ptr(), initial_capital=V0, seed=seed, stream_offset=off(i),
                                 returns_ptr=table_dev.data_ptr() if table_dev is not None else None,
                                 n_returns=table.size if table is not None else 0, hist_bins=HIST_BINS, hist_range=wl["hist_range"],
                                 block_means_ptr=bm.data_ptr() if bm is not None else None,
                                 block_vars_ptr=bv.data_ptr() if bv is not None else None, device=local, stream=stream.cuda_stream,
                                 philox_rounds=philox_rounds, **GAUSS)

    for i in range(warm):
        step(i)
        flush.zero_()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler([local])
    sampler.begin()
    ev, launches = [], 0
    for i in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        launches += step(warm + i)
        b.record(stream)
        flush.zero_()
        ev.append((a, b))
    torch.cuda.synchronize(dev)
    sampler.end()
    clocks = sampler.stop()
    ms = [a.elapsed_time(b) for a, b in ev]
    k_ms = float(np.mean(S.kernel_times(local, steps)))
    peak, _ = issue_peak(torch, dev)
    i_alg = I_ALG_PHILOX7[mode] if philox_rounds == 7 else I_ALG[mode]
    fin = S.finalize_stats(stats.cpu().numpy(), HIST_BINS, V0)
    leg = {"workload": wl["desc"] + (" [philox_rounds=7]" if philox_rounds == 7 else ""), "mode": mode, "paths": B, "periods": P,
           "value": B * P * steps / (sum(ms) * 1e-3), "unit": UNIT, "ms_per_step": sum(ms) / steps, "kernel_ms": k_ms,
           "steps": steps, "warmup": warm, "gpu_launches": launches,
           "roofline": {"bound": "issue", "i_alg": i_alg, "frac": B * P * i_alg / (k_ms * 1e-3) / peak},
           "sm_mhz": clocks.get("sm_mhz"), "clock_reasons": clocks.get("reasons"),
           "result": {"mean": fin.mean, "std": fin.std, "p_below_target": fin.p_below_target}}
    if mode in PIPE_FLOOR and philox_rounds != 7:
        leg["roofline"]["pipe_floor_frac"] = B * P * PIPE_FLOOR[mode][1] / peak * 1e3 / k_ms
    if with_e2e:  # through the host-buffer entry point: the table really is copied H2D every step
        kw = dict(initial_capital=V0, returns=table, seed=seed, hist_bins=HIST_BINS, hist_range=wl["hist_range"],
                  want_block_partials=wl["block_partials"], device=local, philox_rounds=philox_rounds, **GAUSS)
        for i in range(2):
            S.simulate(mode, B, P, stream_offset=off(100 + i), **kw)
        t0 = time.perf_counter()
        for i in range(steps):
            S.simulate(mode, B, P, stream_offset=off(200 + i), **kw)
        dt = time.perf_counter() - t0
        leg["e2e"] = {"value": B * P * steps / dt, "unit": UNIT, "h2d_bytes_per_step": int(table.nbytes) if table is not None else 0,
                      "d2h_bytes_per_step": int(S.stats_len(HIST_BINS) * 8 + (2 * nb * 4 if wl["block_partials"] else 0))}
    del flush
    return leg


def hbm_leg(torch, S, which, dev, local, steps=200):
    """reduce_mean_gpu / vector_add_gpu over device-resident vectors larger than L2: GB/s against MEASURED_PEAKS.json hbm_gbs"""
    peaks = measured_peaks()
    peak = float(peaks.get("hbm_gbs") or 6650.0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    if which == "reduce_mean":
        n, bpe = 1 << 28, 4
        bufs = [torch.full((n,), 1.0 + 0.25 * i, dtype=torch.float32, device=dev) for i in range(2)]  # alternate: 2 GiB > L2
        for i in range(3):
            S.reduce_mean_device(bufs[i & 1].data_ptr(), n, device=local, stream=stream)
        ms, ok = [], True
        for i in range(steps):
            m, k = S.reduce_mean_device(bufs[i & 1].data_ptr(), n, device=local, stream=stream)
            ok = ok and (m == 1.0 + 0.25 * (i & 1))
            ms.append(k)
        kernel = "smmc::reduce_sum_tma_kernel"
    else:
        n, bpe = 1 << 27, 12
        a = torch.full((n,), 1.25, dtype=torch.float32, device=dev)
        b = torch.full((n,), 2.5, dtype=torch.float32, device=dev)
        out = torch.empty_like(a)
        for _ in range(3):
            S.vector_add_device(out.data_ptr(), a.data_ptr(), b.data_ptr(), n, device=local, stream=stream)
        ms = [S.vector_add_device(out.data_ptr(), a.data_ptr(), b.data_ptr(), n, device=local, stream=stream) for _ in range(steps)]
        ok = bool(torch.all(out == 3.75).item())
        kernel = "smmc::vector_add_kernel"
    k_ms = float(np.mean(ms))
    gbs = bpe * n / (k_ms * 1e-3) / 1e9
    return {"workload": f"{which}_gpu over {n} device-resident floats", "value": n / (k_ms * 1e-3), "unit": "elements/s", "kernel_ms": k_ms,
            "steps": steps, "exact": bool(ok),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "bytes_per_element": bpe,
                         "kernel": kernel, "traffic": ncu_traffic(which),
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks.get("hbm_gbs") else "fallback 6650 GB/s (B200_PROFILING.md)"}}


def ref_gpu_same_box(S):
    """The reference's own kernels (src/simulations.cu unmodified, compiled for sm_100a into oracle/_ref/libref_gpu.so) through its
    own entry points at 10^8 x 360 on the 1127-entry table, next to this build's BOOTSTRAP_COMPAT on the same inputs (bit-identical
    results, tests/test_gpu_parity.py).  A baseline leg like cpu_baseline: the only place bench.py runs the reference's GPU code."""
    from oracle import binding as O
    if not O.ref_available("libref_gpu.so"):
        return {"unavailable": "oracle/_ref/libref_gpu.so did not travel with the snapshot"}
    N, P = 100_000_000, 360
    table = S.synthetic_sp500_table(1127, seed=7)

    def best(fn, reps=3):
        fn()
        ts, extra = [], None
        for _ in range(reps):
            t0 = time.perf_counter()
            extra = fn()
            ts.append(time.perf_counter() - t0)
        return min(ts), extra

    out = {"paths": N, "periods": P, "n_returns": 1127, "unit": UNIT,
           "how": "wall clock around the blocking entry point, best of 3 after one warm-up call; reference_own_kernel_clock = the "
                  "reference library's 'GPU kernel itself took' line (host steady_clock, 1 ms resolution, src/simulations.cu:463)"}
    t, _ = best(lambda: O.ref_mc_simulations_gpu_reduceBlock(N, P, V0, table))
    k = O.ref_gpu_last_kernel_seconds()
    out["reference_reduceBlock"] = {"wall_ms": 1e3 * t, "value": N * P / t, "reference_own_kernel_clock_ms": 1e3 * k if k > 0 else None,
                                    "value_by_own_kernel_clock": N * P / k if k > 0 else None}
    t, _ = best(lambda: O.ref_mc_simulations_gpu(N, P, V0, table), reps=2)
    out["reference_all_finals"] = {"wall_ms": 1e3 * t, "value": N * P / t}
    # the reference's kernels by a real clock: CUDA events around the unmodified kernels (oracle/ref_shim/ref_gpu_event_driver.cu)
    if O.ref_available("libref_gpu_events.so"):
        ev_rb = O.ref_kernel_event_ms("reduceBlock", N, P, V0, table, reps=5)
        ev_fin = O.ref_kernel_event_ms("finals", N, P, V0, table, reps=3)
        out["reference_kernel_event_ms"] = ev_rb
        out["reference_finals_kernel_event_ms"] = ev_fin
        out["reference_reduceBlock"]["value_by_event"] = N * P / (ev_rb * 1e-3)
        out["event_how"] = ("mc_simulations_gpu_kernel_reduceBlock<256> / mc_simulations_gpu_kernel<256> (src/simulations.cu:185-247, "
                            ":112-152), compiled unmodified for sm_100a, between cudaEvents; mean of 5 / 3 launches after a warm-up")
    t, r = best(lambda: S.simulate("bootstrap_compat", N, P, returns=table, want_block_partials=True))
    out["this_build_compat_block_partials"] = {"wall_ms": 1e3 * t, "value": N * P / t, "kernel_ms": r.device_ms, "value_kernel": N * P / (r.device_ms * 1e-3)}
    t, r = best(lambda: S.simulate("bootstrap_compat", N, P, returns=table, want_finals=True))
    out["this_build_compat_all_finals"] = {"wall_ms": 1e3 * t, "value": N * P / t, "device_span_ms": r.device_ms,
                                           "how": "smmc_run, host buffers: chunked launches, D2H and host copy-out pipelined"}
    return out


def cpu_v2_leg():
    """BASELINE config 1 itself: the reference's CPU v2 (`mc_simulations`, src/simulations.cpp:204-266, the call of
    examples/benchmark_mc_cpu_v2.cpp:32) at 360 months x 1 M paths on this box's host cores, from oracle/_ref/libref_cpu.so
    (the unmodified reference source).  One warm-up run, then three timed runs."""
    cpu = CpuReference("bootstrap", 360)
    n = 1_000_000
    cpu.run(n)
    ts = [cpu.run(n)[0] for _ in range(3)]
    return {"workload": "C1: 360 months x 1 M bootstrap paths, reference CPU v2", "kind": cpu.kind, "cores": cpu.cores,
            "seconds_best": min(ts), "seconds_mean": sum(ts) / len(ts), "value": n * 360 / min(ts), "unit": UNIT, "how": cpu.desc}


def r4(x):
    """4 significant digits (the summary has to fit the 1500-character tail the driver keeps)"""
    return None if x is None else float(f"{float(x):.4g}")


def make_summary(extras):
    """The numbers BASELINE.md section 5 quotes, compact, for the END of the JSON line."""
    out = {}
    legs = (extras or {}).get("workloads") or {}
    for k in ("c3", "c3compat", "c4", "c4boot", "c2_philox7", "c3_philox7"):
        if k in legs:
            out[k] = [r4(legs[k]["value"] / 1e9), r4(legs[k]["kernel_ms"]), r4(legs[k]["roofline"]["frac"])]
    for k in ("reduce_mean", "vector_add"):
        if k in legs:
            out[k] = [r4(legs[k]["value"] / 1e9), r4(legs[k]["kernel_ms"]), r4(legs[k]["roofline"]["frac"])]
    if "c3" in legs and "e2e" in legs["c3"]:
        out["c3_e2e"] = [r4(legs["c3"]["e2e"]["value"] / 1e9), legs["c3"]["e2e"]["h2d_bytes_per_step"]]
    c5 = (extras or {}).get("c5")
    if c5:
        out["c5"] = [r4(c5["ms_per_step"]), c5["hist_sha256"][:12], r4(c5["closed_form"]["mean_rel_err"]),
                     r4(c5["closed_form"]["std_rel_err"])]
    v2 = (extras or {}).get("cpu_v2")
    if v2 and "value" in v2:
        out["cpu_v2"] = [r4(v2["value"] / 1e9), v2["cores"]]
    rg = (extras or {}).get("ref_gpu_same_box") or {}
    if "reference_reduceBlock" in rg:
        out["ref_gpu"] = [r4(rg["reference_reduceBlock"]["wall_ms"]), r4(rg["reference_all_finals"]["wall_ms"]),
                          r4(rg["this_build_compat_block_partials"]["wall_ms"]), r4(rg["this_build_compat_all_finals"]["wall_ms"])]
        if rg.get("reference_kernel_event_ms"):
            out["ref_kernel_event_ms"] = r4(rg["reference_kernel_event_ms"])
    cp = (extras or {}).get("cold_process") or {}
    if "call_s_min" in (cp.get("this_build") or {}) and "call_s_min" in (cp.get("reference") or {}):
        out["cold_call_s"] = [cp["this_build"]["call_s_min"], cp["reference"]["call_s_min"]]
    se = (extras or {}).get("sharded_equals_single")
    if se:
        out["all_modes_identical"] = se["all_modes_identical"]
        out["modes"] = sorted(k for k, v in se["modes"].items() if v)
    if "quartiles_device_ms" in (extras or {}):
        out["quartiles_ms"] = [r4(extras["quartiles_device_ms"]), r4(extras.get("quartiles_riding_along_extra_ms"))]
    out["keys"] = ("legs [G path-months/s (G elements/s: reduce_mean, vector_add), kernel_ms, roofline frac]; c3_e2e [G/s, h2d bytes]; "
                   "c5 [ms_per_step, sha256[:12], mean_rel_err, std_rel_err]; cpu_v2 [G/s, threads]; "
                   "ref_gpu wall ms [ref reduceBlock, ref finals, ours partials, ours finals]; cold_call_s [ours, reference]: benchmark_mc_gpu 1 360 1e8, fresh process")
    return out


def cold_process(S, reps=3):
    """The way every reference program uses the library: ONE call in a fresh process.  The reference's own benchmark_mc_gpu
    (examples/benchmark_mc_gpu.cpp, unmodified) at `1 360 100000000`, three fresh processes each of (i) the example relinked
    against libsmmc.so and (ii) the example built with the reference's own src/simulations.cu for sm_100a; the figure is the
    wall clock the example itself prints around mc_simulations_gpu (:68-71), i.e. context creation, module load, allocation,
    kernel, 400 MB download.  Also the whole process by this script's clock."""
    import pathlib
    import shutil
    import subprocess
    import tempfile
    root = os.path.dirname(os.path.abspath(__file__))
    exes = {"this_build": os.path.join(root, "oracle", "_ref", "benchmark_mc_gpu"),
            "reference": os.path.join(root, "oracle", "_ref", "benchmark_mc_gpu_ref")}
    out = {"args": "1 360 100000000", "n_returns": 1127, "how": "seconds printed by the example around its one mc_simulations_gpu call; fresh process each"}
    table = S.synthetic_sp500_table(1127, seed=7)
    with tempfile.TemporaryDirectory() as d:
        d = pathlib.Path(d)
        (d / "data").mkdir()
        with open(d / "data" / "SP500_monthly_returns.csv", "w") as f:
            f.write("Date,returns\n")
            for x in table:
                f.write(f"1900-01-01,{float(x)!r}\n")
        env = dict(os.environ)
        if os.path.isdir("/usr/lib/locale/C.utf8"):  # the example asks for en_US.UTF-8, which this image does not ship
            shutil.copytree("/usr/lib/locale/C.utf8", d / "locale" / "en_US.UTF-8")
            env["LOCPATH"] = str(d / "locale")
        for name, exe in exes.items():
            if not os.path.exists(exe):
                out[name] = {"unavailable": f"{os.path.relpath(exe, root)} did not travel with the snapshot"}
                continue
            call_s, proc_s = [], []
            for _ in range(reps):
                t0 = time.perf_counter()
                p = subprocess.run([exe, "1", "360", "100000000"], cwd=d, env=env, capture_output=True, text=True)
                proc_s.append(time.perf_counter() - t0)
                m = re.search(r"simulation done in ([\d.]+) s", p.stdout)
                if p.returncode != 0 or not m:
                    out[name] = {"unavailable": f"rc={p.returncode}: {(p.stderr or p.stdout)[-200:]}"}
                    break
                call_s.append(float(m.group(1)))
            else:
                out[name] = {"call_s": call_s, "call_s_min": min(call_s), "process_s": [round(x, 3) for x in proc_s]}
    return out


def hist_sha256(hist):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(np.asarray(hist, dtype=np.uint64)).tobytes()).hexdigest()


# ---- the GPU arm --------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (default 100; 5 for the 1.8 s steps of c5)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + ["reduce_mean", "vector_add"])
    ap.add_argument("--paths", type=int, default=0, help="override paths per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extras legs (other workloads, C5, same-box reference kernels)")
    ap.add_argument("--no-c5", action="store_true")
    args = ap.parse_args()
    if args.workload == "reduce_mean":
        return run_reduce_mean(args)
    if args.workload == "vector_add":
        return run_vector_add(args)
    wl = dict(WORKLOADS[args.workload])
    if args.paths:
        wl["paths"] = args.paths
    if args.steps <= 0:
        args.steps = 5 if wl.get("strong") else 100
    if args.impl == "reference":
        return run_reference_arm(args, wl)
    if args.warmup < 3:
        args.warmup = 3  # timing rule: >= 3 warm-up steps

    import torch
    import torch.distributed as dist
    import stock_market_monte_carlo_b200 as S

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if S.load().smmc_device_count() <= local:
        raise SystemExit("bench.py needs a B200 per rank: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm, nccl_log = None, None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL's communicator lines are evidence (rank count, transport).  Unless the caller already asked for INFO or TRACE,
        # ask NCCL for its INIT log into a scratch file and replay the communicator lines on stderr at the end
        # (NCCL_DEBUG_FILE=/dev/stderr would reopen -- and truncate -- a redirected stderr; NCCL's default is stdout, which
        # must keep the JSON line).  libsmmc.so additionally prints one "rank r nranks N cudaDev d" line per rank to stderr.
        asked = os.environ.get("NCCL_DEBUG", "").upper()
        if asked not in ("INFO", "TRACE"):
            os.environ["NCCL_DEBUG"] = "INFO"
            os.environ["NCCL_DEBUG_SUBSYS"] = "INIT"
        if "NCCL_DEBUG_FILE" not in os.environ:
            nccl_log = f"/tmp/smmc_nccl_{os.getpid()}.log"
            os.environ["NCCL_DEBUG_FILE"] = nccl_log
        print(f"bench: rank {rank} NCCL_DEBUG={os.environ['NCCL_DEBUG']} (caller had {asked or 'none'}) "
              f"NCCL_DEBUG_FILE={os.environ['NCCL_DEBUG_FILE']}", file=sys.stderr)
        dist.init_process_group("nccl", device_id=dev)  # launcher plumbing: barrier + max-over-ranks of the timings
        from stock_market_monte_carlo_b200.distributed import init_comm
        S.set_variant("verbose", 1)  # libsmmc's own "rank r nranks N cudaDev d" line (silent by default)
        comm = init_comm(local)  # the path's own communicator: ncclCommInitRank inside libsmmc.so (smmc_run_rank)

    mode, B, P = wl["mode"], wl["paths"], wl["periods"]
    strong = bool(wl.get("strong"))
    job_paths = B if strong else world * B  # paths of one step over all ranks
    # one step = ONE global run of job_paths paths, sharded over the ranks by RNG offset inside smmc_run_rank (whole 256-path
    # blocks per rank); B = this rank's share of it
    _, B = S.rank_range(job_paths, world, rank)
    nb = (B + 255) // 256
    table = S.synthetic_sp500_table(1200) if "bootstrap" in mode else None
    seed = 0 if mode == "bootstrap_compat" else 1234
    sobol = mode.startswith("sobol")
    slen = S.stats_len(HIST_BINS)

    # device-resident inputs / outputs of the timed region
    table_dev = torch.from_numpy(table).to(dev) if table is not None else None
    stats = torch.zeros(slen, dtype=torch.float64, device=dev)
    bmeans = torch.empty(nb, dtype=torch.float32, device=dev) if wl["block_partials"] else None
    bvars = torch.empty(nb, dtype=torch.float32, device=dev) if wl["block_partials"] else None
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream(dev)

    def offset_of(step):
        """stream offset of the GLOBAL run of this step (fresh, disjoint streams every step)"""
        if sobol or mode == "bootstrap_compat":  # 32-bit point / path ids: wrap the offsets
            span = (1 << 31) - 256 - job_paths if mode == "bootstrap_compat" else (1 << 32) - 512 - job_paths
            return ((step * job_paths) % max(span, 1)) // 256 * 256
        return step * ((job_paths + 255) // 256 * 256)

    dev_kw = dict(initial_capital=V0, seed=seed, returns_ptr=table_dev.data_ptr() if table_dev is not None else None,
                  n_returns=table.size if table is not None else 0, hist_bins=HIST_BINS, hist_range=wl["hist_range"],
                  block_means_ptr=bmeans.data_ptr() if bmeans is not None else None,
                  block_vars_ptr=bvars.data_ptr() if bvars is not None else None, stream=stream.cuda_stream, **GAUSS)

    def device_step(step):
        if comm is not None:  # kernel -> ncclAllReduce of the statistics vector, both enqueued by smmc_run_rank on this stream
            return comm.simulate_device(mode, job_paths, P, stats.data_ptr(), stream_offset=offset_of(step), **dev_kw)
        return S.simulate_device(mode, B, P, stats.data_ptr(), stream_offset=offset_of(step), device=local, **dev_kw)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    sampler = ClockSampler(list(range(world))) if rank == 0 else None
    for s in range(args.warmup):
        device_step(s)
        flush.zero_()
    barrier()

    # ---- timed region: exactly K steps; CUDA events on the launching stream around every step; the L2 flush
    # between steps sits outside the event pairs
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches = 0
    if sampler:
        sampler.begin()
    t_wall0 = time.perf_counter()
    for s in range(args.steps):
        ev[s][0].record(stream)
        launches += device_step(args.warmup + s)
        ev[s][1].record(stream)
        flush.zero_()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    if sampler:
        sampler.end()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    kernel_ms = S.kernel_times(local, args.steps)  # the simulation kernel alone, same launches
    final = S.finalize_stats(stats.cpu().numpy(), HIST_BINS, V0)  # result of the last step (all ranks merged)
    clocks = sampler.stop() if sampler else None

    # ---- e2e: the public host-buffer entry point, H2D of the table and D2H of every result inside the timed region
    e2e = None
    if not args.no_e2e:
        k_e2e = min(args.steps, 50)

        host_kw = dict(initial_capital=V0, returns=table, seed=seed, hist_bins=HIST_BINS, hist_range=wl["hist_range"],
                       want_block_partials=wl["block_partials"], **GAUSS)

        def host_step(step):
            if comm is not None:  # smmc_run_rank with host buffers: table H2D, kernel, all-reduce, merged statistics + partials D2H
                return comm.simulate(mode, job_paths, P, stream_offset=offset_of(step), **host_kw)
            return S.simulate(mode, B, P, stream_offset=offset_of(step), device=local, **host_kw)

        for s in range(3):
            host_step(10_000 + s)
        barrier()
        ar0 = comm.allreduce_count() if comm is not None else 0
        t0 = time.perf_counter()
        for s in range(k_e2e):
            host_step(20_000 + s)
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        ar_per_step = (comm.allreduce_count() - ar0) / k_e2e if comm is not None else 0
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": job_paths * P * k_e2e / float(dt.item()), "unit": UNIT,
               "h2d_bytes_per_step": int(table.nbytes) if table is not None else 0,
               "d2h_bytes_per_step": int(slen * 8 + (2 * nb * 4 if wl["block_partials"] else 0)),
               "allreduces_per_step": ar_per_step,
               "steps": k_e2e, "api": ("stock_market_monte_carlo_b200.simulate -> smmc_run (host buffers, blocking)" if comm is None else
                                       "Comm.simulate -> smmc_run_rank (host buffers, blocking; ncclAllReduce inside the C ABI)"),
               "inputs": ("return table copied host->device every step" if table is not None else
                          "none to copy: the Gaussian modes take scalars only (mean, std, seed, offset)")}

    # ---- extras (outside the timed region).  SURVEY 8(f) rank 1: the quartiles of the same workload by replay radix select
    # (two replays + one pass over the candidate keys) instead of a finals download + host sort
    extras = {}
    do_extras = not args.no_e2e and not args.no_extras
    if world == 1 and do_extras and not sobol:
        q, q_ms = S.quantiles(mode, B, P, initial_capital=V0, returns=table, seed=seed, device=local, return_ms=True, **GAUSS)
        _, q_f, q_extra_ms = S.simulate_with_quantiles(mode, B, P, initial_capital=V0, returns=table, seed=seed, device=local,
                                                       hist_bins=HIST_BINS, hist_range=wl["hist_range"], target=V0,
                                                       want_block_partials=wl["block_partials"], **GAUSS)
        extras.update({"quartiles_min_q1_med_q3_max": [float(x) for x in q], "quartiles_device_ms": q_ms,
                       "quartiles_how": "smmc_quantiles: 2 replays of the workload (11 + 11 key bits; the second also collects the "
                                        "candidate keys) + one HBM pass over the candidates (10 bits); exact",
                       "quartiles_riding_along_extra_ms": q_extra_ms, "quartiles_riding_along_equal": bool(np.array_equal(q, q_f)),
                       "quartiles_riding_along_how": "smmc_run_select: the first select pass shares the statistics launch; "
                                                     "1 extra replay + the candidate pass"})

    # ---- C5 (BASELINE config 5): ONE run of 10^10 paths x 360 months sharded over the ranks (strong scaling), two steps; the
    # histogram hash of the step at stream offset 0 must be the same string for every number of GPUs
    if do_extras and args.workload == "c2" and not args.no_c5:
        c5 = WORKLOADS["c5"]
        Nc5, Pc5 = c5["paths"], c5["periods"]
        st5 = torch.zeros(slen, dtype=torch.float64, device=dev)
        kw5 = dict(initial_capital=V0, seed=1234, hist_bins=HIST_BINS, hist_range=c5["hist_range"], stream=stream.cuda_stream, **GAUSS)

        def c5_step(i):
            if comm is not None:
                comm.simulate_device("gaussian", Nc5, Pc5, st5.data_ptr(), stream_offset=i * Nc5, **kw5)
            else:
                S.simulate_device("gaussian", Nc5, Pc5, st5.data_ptr(), stream_offset=i * Nc5, device=local, **kw5)

        barrier()
        ev5, res5 = [], None
        for i in (1, 0):  # the second (reported) step is the run at offset 0
            a5, b5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a5.record(stream)
            c5_step(i)
            b5.record(stream)
            ev5.append((a5, b5))
        barrier()
        ms5 = torch.tensor([sum(a5.elapsed_time(b5) for a5, b5 in ev5)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms5, op=dist.ReduceOp.MAX)
        res5 = S.finalize_stats(st5.cpu().numpy(), HIST_BINS, V0)
        mu, sg = GAUSS["gauss_mean"] * 0.01, GAUSS["gauss_std"] * 0.01
        cf_mean = V0 * (1.0 + mu) ** Pc5  # E[V] = V0 (1+mu)^P, E[V^2] = V0^2 ((1+mu)^2 + sigma^2)^P (RNG-independent)
        cf_std = (V0 ** 2 * ((1.0 + mu) ** 2 + sg ** 2) ** Pc5 - cf_mean ** 2) ** 0.5
        extras["c5"] = {"workload": c5["desc"], "scaling": "strong", "n_gpus": world, "paths": Nc5, "periods": Pc5, "steps": 2,
                        "ms_per_step": float(ms5.item()) / 2, "value": 2.0 * Nc5 * Pc5 / (float(ms5.item()) * 1e-3), "unit": UNIT,
                        "count": res5.count, "mean": res5.mean, "std": res5.std, "below_count": res5.below_count,
                        "hist_sha256": hist_sha256(res5.hist),
                        "closed_form": {"mean": float(cf_mean), "std": float(cf_std), "mean_rel_err": float(res5.mean / cf_mean - 1.0),
                                        "std_rel_err": float(res5.std / cf_std - 1.0)},
                        "collective": "1 ncclAllReduce per step inside smmc_run_rank" if world > 1 else "none"}

    # ---- sharded == single, every mode: a 1 M-path run through the rank path (finals + block partials + histogram +
    # quartiles) against the same run on ONE GPU.  Every rank repeats the single-GPU run on its own device and compares its
    # own path range bit for bit; the verdicts are AND-ed over the ranks.  1 000 003 paths do not divide into equal shards;
    # the Sobol / COMPAT runs start at a non-zero 256-aligned stream offset.
    if do_extras and world > 1:
        Ns, Ps = 1_000_003, 120
        tbl_s = S.synthetic_sp500_table(1200)
        first_s, count_s = S.rank_range(Ns, world, rank)
        per_mode, worst_mean, worst_m2 = {}, 0.0, 0.0
        for m in ("bootstrap_compat", "bootstrap", "gaussian", "sobol_bootstrap", "sobol_gaussian"):
            kws = dict(initial_capital=V0, seed=0 if m == "bootstrap_compat" else 77, hist_bins=HIST_BINS, hist_range=(0.0, 20000.0),
                       stream_offset=0 if m in ("bootstrap", "gaussian") else 7 * 256, **GAUSS)
            if "bootstrap" in m:
                kws["returns"] = tbl_s
            sharded = comm.simulate(m, Ns, Ps, want_finals=True, want_block_partials=True, **kws)
            kwq = {k: v for k, v in kws.items() if k not in ("hist_bins", "hist_range")}
            q_sh = comm.quantiles(m, Ns, Ps, **kwq)
            single = S.simulate(m, Ns, Ps, device=local, want_finals=True, want_block_partials=True, **kws)
            q_1 = S.quantiles(m, Ns, Ps, device=local, **kwq)
            b0 = first_s // 256
            ok = (np.array_equal(sharded.finals, single.finals[first_s:first_s + count_s])
                  and np.array_equal(sharded.block_means, single.block_means[b0:b0 + sharded.block_means.size])
                  and np.array_equal(sharded.block_vars, single.block_vars[b0:b0 + sharded.block_vars.size])
                  and np.array_equal(sharded.hist, single.hist) and sharded.count == single.count
                  and sharded.below_count == single.below_count and np.array_equal(q_sh, q_1))
            worst_mean = max(worst_mean, abs(sharded.mean - single.mean) / abs(single.mean))
            worst_m2 = max(worst_m2, abs(sharded.m2 - single.m2) / abs(single.m2))
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            per_mode[m] = bool(flag.item())
        extras["sharded_equals_single"] = {
            "paths": Ns, "periods": Ps, "n_ranks": world, "modes": per_mode, "all_modes_identical": all(per_mode.values()),
            "compared": "finals and block mean/variance partials of each rank's range, histogram, count, below, quartiles: bitwise",
            "mean_rel_diff_max": worst_mean, "m2_rel_diff_max": worst_m2,
            "how": "Comm.simulate / Comm.quantiles (smmc_run_rank / smmc_quantiles_rank, NCCL inside the C ABI) vs smmc_run / "
                   "smmc_quantiles of the whole run on each rank's own GPU; AND over ranks"}
        barrier()

    def replay_nccl_log():
        if world > 1 and nccl_log and os.path.exists(nccl_log):
            p2p = 0
            for ln in open(nccl_log, errors="replace"):
                if "via P2P" in ln or "via SHM" in ln or "via NET" in ln:
                    p2p += 1
                    if p2p > 2:  # one line per channel and peer: the first two name the transport
                        continue
                elif not any(k in ln for k in ("nranks", "nRanks", "Init COMPLETE", "NVLS", "NCCL version", "Connected all")):
                    continue
                print(ln.rstrip(), file=sys.stderr)
            sys.stderr.flush()

    if rank != 0:
        if world > 1:
            replay_nccl_log()
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- derived numbers
    peaks = measured_peaks()
    props = torch.cuda.get_device_properties(dev)
    n_sm = props.multi_processor_count
    f_max = float(peaks.get("sm_max_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0) * 1e6
    peak = n_sm * 128 * f_max  # thread-instructions/s: 4 schedulers x 32 lanes per SM per clock
    k_ms = float(np.mean(kernel_ms)) if len(kernel_ms) else total_ms / args.steps
    achieved = B * P * I_ALG[mode] / (k_ms * 1e-3)
    roofline = {
        "bound": "issue", "achieved": achieved / 1e12, "peak": peak / 1e12, "unit": "Tinstr/s", "frac": achieved / peak,
        "traffic": ncu_traffic(mode), "traffic_unit": "dram bytes per launch (ncu); algorithmic HBM bytes: table + partials only",
        "kernel": "smmc::mc_kernel<%s>" % mode, "kernel_ms": k_ms,
        "i_alg": I_ALG[mode], "survey_i_alg": I_ALG_SURVEY[mode],
        "frac_survey_budget": (achieved / I_ALG[mode] * I_ALG_SURVEY[mode] / peak) if I_ALG_SURVEY[mode] else None,
        "peak_source": f"{n_sm} SMs x 128 thread-instr/clk x sm_max_mhz {f_max / 1e6:.0f} "
                       f"({'MEASURED_PEAKS.json' if peaks.get('sm_max_mhz') else 'nvidia-smi'}); HBM/tensor peaks do not bound this path",
    }
    if mode in PIPE_FLOOR:
        pipe, cyc = PIPE_FLOOR[mode]
        floor_ms = B * P * cyc / (n_sm * 128 * f_max) * 1e3
        roofline["pipe_floor"] = {"pipe": pipe, "cycles_per_warp_month": cyc, "floor_ms": floor_ms, "frac": floor_ms / k_ms}
    if clocks and clocks.get("sm_mhz"):
        roofline["frac_at_sampled_clock"] = achieved / (n_sm * 128 * clocks["sm_mhz"] * 1e6)

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        cpu = CpuReference(mode, P)
        sample = cpu.sized_sample(12.0)
        dt, n = cpu.run(sample)
        cpu_baseline = {"value": n * P / dt, "unit": UNIT, "cores": cpu.cores, "kind": cpu.kind,
                        "sample": f"{n} paths x {P} months in {dt:.1f} s; {cpu.desc}"}

    # ---- extras: the other workloads, the HBM-bound kernels and the same-box reference kernels, each a short device-timed leg
    if world == 1 and do_extras and args.workload == "c2":
        del flush, bmeans, bvars
        torch.cuda.empty_cache()
        legs = {}
        for name in ("c3", "c3compat", "c4", "c4boot"):
            legs[name] = short_leg(torch, S, name, dev, local, with_e2e=(name == "c3"))
        legs["c2_philox7"] = short_leg(torch, S, "c2", dev, local, philox_rounds=7)
        legs["c3_philox7"] = short_leg(torch, S, "c3", dev, local, philox_rounds=7)
        legs["reduce_mean"] = hbm_leg(torch, S, "reduce_mean", dev, local)
        legs["vector_add"] = hbm_leg(torch, S, "vector_add", dev, local)
        extras["workloads"] = legs
        try:
            extras["cold_process"] = cold_process(S)
        except Exception as e:
            extras["cold_process"] = {"unavailable": f"{type(e).__name__}: {e}"}
        try:
            extras["cpu_v2"] = cpu_v2_leg()
        except Exception as e:
            extras["cpu_v2"] = {"unavailable": f"{type(e).__name__}: {e}"}
        try:
            extras["ref_gpu_same_box"] = ref_gpu_same_box(S)
            rb = extras["ref_gpu_same_box"].get("reference_reduceBlock")
            if rb:
                for k in ("c3", "c3_philox7", "c3compat"):
                    legs[k]["vs_ref_gpu_same_box"] = {"e2e_wall": legs[k]["value"] / rb["value"],
                                                      "kernel_event": (extras["ref_gpu_same_box"]["reference_kernel_event_ms"] / legs[k]["kernel_ms"]
                                                                       if extras["ref_gpu_same_box"].get("reference_kernel_event_ms") else None),
                                                      "reference_own_kernel_clock": (legs[k]["value"] / rb["value_by_own_kernel_clock"]
                                                                                     if rb.get("value_by_own_kernel_clock") else None)}
        except Exception as e:  # the baseline leg must never take the headline down with it
            extras["ref_gpu_same_box"] = {"unavailable": f"{type(e).__name__}: {e}"}

    value = job_paths * P * args.steps / (total_ms * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "mode": mode, "paths_per_gpu_per_step": B, "periods": P, "seed": seed,
                   "hist_bins": HIST_BINS, "block_partials": wl["block_partials"], "parallelism": f"paths sharded by RNG offset x{world}",
                   "collective": ("1 ncclAllReduce(ncclFloat64, ncclSum) of %d doubles per step, enqueued by smmc_run_rank (C ABI) "
                                  "on the kernel's stream; the host-buffer e2e leg (two launches per step) also issues one" % slen)
                   if world > 1 else "none",
                   "l2": "flushed between timed steps (256 MiB memset outside the event pairs); the path has no HBM input larger than 4.8 KB"},
        "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        "wall_ms_per_step_incl_flush": 1e3 * t_wall / args.steps,
        "result": {"count": final.count, "mean": final.mean, "std": final.std, "p_below_target": final.p_below_target},
        "extras": extras or None,
    }
    line["summary"] = make_summary(extras)  # LAST key: survives in the tail of the line that the driver's record keeps
    replay_nccl_log()
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
