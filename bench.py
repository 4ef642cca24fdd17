#!/usr/bin/env python
"""Benchmark of the EEG -> spectrogram hot path (BASELINE.json metric: records -> spectrograms / s).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [--gpus N] ...                # the reference's CPU path (oracle port)

A step is one pass of the transform over one batch of synthetic EEG.  The headline workload is
BASELINE.json configs[1]: 1 024 records of 20 ch x 10 000 samples -> (1024, 128, 256, 4), the
4-chain x 128-mel image ("mode N").  Extra objects on the same JSON line:

  time_major     the same records in the reference's parquet layout (L, 20)
  mode_r         the reference-pinned ``compute_stft`` image on (1024, 20, 10000) and (64, 19, 2000)
  eeg_transform  row (f)1, ``_EEGTransformer.transform`` (+ the fused transform -> STFT image)
  next_rows      rows (f)2-(f)4: DiffEEG condition front-end, classifier-side EEG / spectrogram preprocessing
  cfg4           BASELINE configs[3]: one 1-hour record, 1 776 overlapping windows (mode N) / one call (mode R)
  cfg3           BASELINE configs[2]: the 106 800-segment set sharded over the ranks, STRONG scaling,
                 chunks of 4 096 records per rank, every image retained in HBM
  gather         N > 1: the optional final NCCL all-gather of the images (bytes, ms, bus GB/s) and a
                 check that gathered shards equal rank 0's own single-GPU result
  sustained      the headline launch (and mode R) looped for >= 2 s with clocks / power sampled
  e2e            host buffers in and out through the public API
  cpu_baseline   the oracle port on all host cores (bounded sample)

Under torchrun every rank transforms its own 1 024 records (weak scaling, no collective on the
data path); timing is CUDA events per rank, max over ranks.
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

# rank 0 prints ONE JSON line on stdout.  Libraries write there too (NCCL's version banner is a
# C-level printf on fd 1 and ignores NCCL_DEBUG_FILE), so keep a private handle on the real stdout
# for the JSON line and point fd 1 at stderr for everything else.
_JSON_OUT = None


def claim_stdout():
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(obj):
    claim_stdout()
    _JSON_OUT.write(json.dumps(obj) + "\n")
    _JSON_OUT.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "eeg_records_to_spectrograms_per_sec"
UNIT = "records/s"
RECORDS = 1024
N_DATASET = 106_800                                     # HMS label rows: BASELINE configs[2]
CHUNK_DATASET = 4096
L_HOUR, WINDOW, STRIDE = 720_000, 10_000, 400           # BASELINE configs[3]
WORKLOAD_N = ("BASELINE configs[1]: 1024 synthetic 50-s EEG records (20 ch x 10000 @200 Hz) -> "
              "(1024,128,256,4) 4-chain x 128-mel log spectrogram stack")
BYTES_N = 4 * 20 * 10000 + 4 * 128 * 256 * 4            # 1 324 288 B / record
BYTES_R50 = 4 * 20 * 10000 + 4 * 20 * 33 * 2000         # 6 080 000 B / record
BYTES_R10 = 4 * 19 * 2000 + 4 * 19 * 33 * 2000          # 5 168 000 B / record
CPU_RECORDS_PER_WORKER = 24                             # every CPU worker gets >= 24 records per timed call


def config_n(world):
    """The `config` object of BOTH arms (the reference arm runs a bounded sample of it)."""
    return {"workload": WORKLOAD_N, "mode": "N", "records_per_gpu": RECORDS,
            "l2": "inputs are 819 MB per step per GPU (> 126 MB L2); no flush needed",
            "parallelism": f"records sharded over {world} GPU(s), no collective"}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def committed_traffic(key):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            return json.load(fh).get(key)
    except Exception:
        return None


def roofline(bytes_per_launch, seconds_per_launch, peak, kernel, traffic_key=None, **extra):
    achieved = bytes_per_launch / seconds_per_launch / 1e9
    out = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "traffic": committed_traffic(traffic_key) if traffic_key else None, "kernel": kernel}
    out.update(extra)
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_ids):
        self.gpu_ids = set(gpu_ids)
        self.rows, self.proc, self.thread = [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line))

    def window(self, t0, t1):
        """Clock statistics of the samples taken in [t0, t1]."""
        sm, smax, power, reasons = [], [], [], set()
        for ts, line in list(self.rows):
            p = [s.strip() for s in line.split(",")]
            if len(p) < 8 or not p[0].isdigit() or int(p[0]) not in self.gpu_ids:
                continue
            if not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            try:
                sm.append(float(p[1]))
                smax.append(float(p[2]))
                power.append(float(p[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "reasons": sorted(reasons), "samples": len(sm)}

    def stop(self):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()


def reference_arm(args, rank, world):
    """The reference's CPU implementation of the path, all host cores, rank 0 only.

    ``kind: "port"``: the reference is pure Python with unimportable modules and cannot travel to the
    GPU box, so the timed code is the oracle port -- mode R makes the same SciPy / NumPy calls the same
    number of times as DiffEEG_utils.py:27-76, mode N is the librosa restatement.  Every worker holds
    >= 24 records per step and inputs are synthesised before the clock starts (oracle/cpu_bench.py)."""
    if rank != 0:
        return
    from oracle import cpu_bench

    arm = cpu_bench.CpuArm()
    per_step = arm.cores * CPU_RECORDS_PER_WORKER
    for _ in range(min(args.warmup, 1)):            # one untimed pass warms caches / FFT plans in every worker
        arm.run("n", arm.cores * 2, seed=99)
    steps = max(1, min(args.steps, 8))              # bounded: ~1.2 s per step and core at 20 records/s
    wall, done = 0.0, 0
    for k in range(steps):
        r = arm.run("n", per_step, seed=2)          # the cpu_baseline leg's sample of the same workload
        wall += r["wall_s"]
        done += r["records"]
    r10 = arm.run("r10", arm.cores * 64, seed=5)
    r50 = arm.run("r50", arm.cores * 16, seed=6)
    arm.close()
    value = done / wall
    sample = (f"{per_step} records/step ({CPU_RECORDS_PER_WORKER} per worker) x {steps} steps of the "
              "(20,10000)->(128,256,4) mel image through oracle/melspec_oracle.py (librosa restatement), one process "
              "per core, inputs resident before the clock starts")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * wall / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": config_n(world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mode_r": {
            "what": "compute_stft through oracle/stft_oracle.py::compute_stft_port (same SciPy/NumPy calls, the same "
                    "number of times, as DiffEEG_utils.py:27-76) -- the reference's real call structure",
            "diffeeg": {"value": r10["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                        "sample": f"{r10['records']} records (19,2000)"},
            "hms50s": {"value": r50["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                       "sample": f"{r50['records']} records (20,10000)"}},
        "gpu_launches": 0,
    }
    emit(line)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--quick", action="store_true", help="headline leg only (profiling runs)")
    ap.add_argument("--legs", default="n,tm,r,xf,next,cfg4,cfg3,gather,sustained,e2e,cpu",
                    help="comma list of legs to run (n always runs); profiling runs use a subset")
    args = ap.parse_args()
    legs = set(args.legs.split(",")) if not args.quick else {"n"}
    if args.no_cpu_baseline:
        legs.discard("cpu")
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, rank, world)

    import torch

    import b200spec

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU path)"
    torch.cuda.set_device(local)
    bound = []
    if world > 1:   # one process per GPU: sit on the socket the GPU's PCIe link hangs off (host-buffer leg)
        from b200spec.sharding import bind_host_to_gpu
        bound = bind_host_to_gpu(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(v, op):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    def reduce_max(v):
        return reduce(v, dist.ReduceOp.MAX) if dist is not None else v

    def reduce_sum(v):
        return reduce(v, dist.ReduceOp.SUM) if dist is not None else v

    def synth(shape, seed, per_rank=True):
        g = torch.Generator(device="cuda").manual_seed(seed + (7919 * rank if per_rank else 0))
        x = torch.randn(shape, generator=g, device="cuda")
        x = x.cumsum(-1).mul_(0.3).add_(torch.randn(shape, generator=g, device="cuda"))
        return x.clamp_(-1024, 1024)

    class Cfg:
        diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

    def timed(fn, steps, warmup, flush=None):
        """Exactly `steps` timed calls between barriers; CUDA events on the launching stream."""
        for _ in range(warmup):
            fn()
        barrier()
        n0 = b200spec.launch_count()
        if flush is None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            barrier()
            ms = e0.elapsed_time(e1)
        else:                       # working set fits L2: flush between steps, time each step alone
            ms = 0.0
            for _ in range(steps):
                flush.add_(1.0)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1)
            barrier()
        return reduce_max(ms), b200spec.launch_count() - n0

    peak, peak_src = measured_peak()
    gpu_ids = list(range(world)) if world > 1 else [torch.cuda.current_device()]
    sampler = ClockSampler(gpu_ids) if rank == 0 else None
    if sampler:
        sampler.start()

    # ---------------------------------------------------------------- headline: mode N, device resident
    x_n = synth((RECORDS, 20, 10000), seed=2)
    params = b200spec.mel_params()
    keep = {}

    def step_n():
        keep["n"] = b200spec.eeg_to_spectrogram(x_n, params)

    t_start = time.perf_counter()
    ms_n, launches_n = timed(step_n, args.steps, args.warmup)
    t_end = time.perf_counter()
    if sampler:
        time.sleep(0.2)            # let the sample that covers the (short) timed region arrive
    clocks = sampler.window(t_start, t_end) if sampler else None
    launches_total = reduce_sum(launches_n)
    value_n = world * RECORDS * args.steps / (ms_n * 1e-3)
    per_launch_s = ms_n * 1e-3 / args.steps
    KERNEL_N = ("melspec_ts_kernel (engine 3, frame operand in tensor memory: one persistent clustered launch per step; "
                "algorithmic bytes 1 324 288 B/record x 1024 records per launch)")
    roof_n = roofline(RECORDS * BYTES_N, per_launch_s, peak, KERNEL_N, "melspec_ts_dram_bytes_per_launch",
                      peak_source=peak_src, timing="burst: %d back-to-back launches" % args.steps,
                      note="co-bound by issue slots / MUFU / tensor pipe / hand-off latencies, not by HBM alone (DESIGN.md 3.2)")

    result = {
        "metric": METRIC, "value": value_n, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_n / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_n(world), "host_cores_bound_rank0": len(bound),
        "roofline": roof_n, "gpu_launches": int(launches_total), "clocks": clocks,
    }

    # ---------------------------------------------------------------- the optional final gather (N > 1)
    if "gather" in legs and dist is not None:
        from b200spec.sharding import gather_spectrograms, shard_bounds

        local_out = keep["n"]
        for _ in range(2):
            g_all = gather_spectrograms(local_out, world * RECORDS)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g_steps = 5
        e0.record()
        for _ in range(g_steps):
            g_all = gather_spectrograms(local_out, world * RECORDS)
        e1.record()
        barrier()
        g_ms = reduce_max(e0.elapsed_time(e1)) / g_steps
        total_b = world * local_out.numel() * 4
        del g_all
        # shards reassembled == the one-GPU result: a common record set, rank r transforms its slice
        n_chk = 24 * world + 5                                     # uneven shards on purpose
        x_chk = synth((n_chk, 20, 10000), seed=11, per_rank=False)
        lo, hi = shard_bounds(n_chk, world, rank)
        got = gather_spectrograms(b200spec.eeg_to_spectrogram(x_chk[lo:hi].contiguous(), params), n_chk)
        same = bool(torch.equal(got, b200spec.eeg_to_spectrogram(x_chk, params)))
        same = reduce(1.0 if same else 0.0, dist.ReduceOp.MIN) == 1.0
        result["gather"] = {
            "what": "b200spec.gather_spectrograms: all_gather_into_tensor (NCCL over NVLink / NVSwitch) of every "
                    "rank's (1024,128,256,4) images into record order; optional -- outputs normally stay sharded",
            "bytes_total": total_b, "ms": g_ms,
            "algbw_GBps": total_b / (g_ms * 1e-3) / 1e9,
            "busbw_GBps": total_b * (world - 1) / world / (g_ms * 1e-3) / 1e9,
            "nvlink_GBps_per_direction_nominal": 900.0,
            "ms_per_transform_step": ms_n / args.steps,
            "shards_equal_single_gpu": same,
            "check": f"{n_chk} common records: shards of {world} ranks gathered == rank-local transform of all, bit for bit"}
        del x_chk, got
    del keep["n"]

    if "sustained" in legs:
        # ------------------------------------------------------------ a sustained number beside the burst
        def sustained(fn, seconds, records_per_call):
            barrier()
            t0 = time.perf_counter()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            calls = 0
            while True:
                for _ in range(50):
                    fn()
                calls += 50
                torch.cuda.synchronize()
                if time.perf_counter() - t0 >= seconds:
                    break
            e1.record()
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            ms = reduce_max(e0.elapsed_time(e1))
            return calls, ms, (sampler.window(t0, t1) if sampler else None)

        calls, ms_s, clk = sustained(step_n, 2.5, RECORDS)
        keep.pop("n", None)
        result["sustained"] = {
            "what": "the headline launch looped for >= 2.5 s (one synchronize per 50 launches), clocks / power sampled "
                    "every 50 ms",
            "mode_n": {"value": world * RECORDS * calls / (ms_s * 1e-3), "unit": UNIT, "launches": calls,
                       "ms_per_step": ms_s / calls, "seconds": ms_s * 1e-3, "clocks": clk,
                       "roofline": roofline(RECORDS * BYTES_N, ms_s * 1e-3 / calls, peak, KERNEL_N,
                                            "melspec_ts_dram_bytes_per_launch", timing="sustained")}}

    if "tm" in legs:
        # ------------------------------------------------------------ mode N on the parquet layout (L, 20)
        x_tm = x_n.transpose(1, 2).contiguous()
        p_tm = b200spec.mel_params(time_major=True)

        def step_tm():
            keep["tm"] = b200spec.eeg_to_spectrogram(x_tm, p_tm)

        ms_tm, _ = timed(step_tm, args.steps, args.warmup)
        result["time_major"] = {
            "what": "the same records as (1024,10000,20) rows (the reference's parquet layout): row gather into the "
                    "workspace + the default engine, bit-identical output",
            "value": world * RECORDS * args.steps / (ms_tm * 1e-3), "unit": UNIT, "ms_per_step": ms_tm / args.steps}
        del keep["tm"], x_tm

    if "r" in legs:
        # ------------------------------------------------------------ mode R, device resident
        def step_r50():
            keep["r"] = b200spec.compute_stft(x_n, Cfg)

        steps_r = max(3, args.steps // 2)
        ms_r, _ = timed(step_r50, steps_r, 3)
        if "sustained" in legs:
            calls, ms_s, clk = sustained(step_r50, 2.0, RECORDS)
            result["sustained"]["mode_r"] = {
                "value": world * RECORDS * calls / (ms_s * 1e-3), "unit": UNIT, "launches": calls,
                "ms_per_step": ms_s / calls, "seconds": ms_s * 1e-3, "clocks": clk,
                "roofline": roofline(RECORDS * BYTES_R50, ms_s * 1e-3 / calls, peak, "stft_image_kernel<6,0>",
                                     "stft_hms50s_dram_bytes_per_launch", timing="sustained")}
        keep.pop("r", None)
        torch.cuda.empty_cache()
        x_r10 = synth((64, 19, 2000), seed=5)
        flush = torch.zeros(256 << 20 >> 2, device="cuda")

        def step_r10():
            keep["r"] = b200spec.compute_stft(x_r10, Cfg)

        ms_r10, _ = timed(step_r10, args.steps, 3, flush=flush)
        del flush
        keep.pop("r", None)
        result["mode_r"] = {
            "what": "compute_stft drop-in (reference-pinned parity), one fused kernel per call",
            "hms50s": {"workload": "(1024,20,10000)->(1024,20,33,2000) per GPU", "value": world * RECORDS * steps_r / (ms_r * 1e-3),
                       "unit": UNIT, "ms_per_step": ms_r / steps_r,
                       "roofline": roofline(RECORDS * BYTES_R50, ms_r * 1e-3 / steps_r, peak,
                                            "stft_image_kernel<6,0>, 6 080 000 B/record", "stft_hms50s_dram_bytes_per_launch")},
            "diffeeg": {"workload": "(64,19,2000)->(64,19,33,2000) per GPU, L2 flushed between steps",
                        "value": world * 64 * args.steps / (ms_r10 * 1e-3), "unit": UNIT, "ms_per_step": ms_r10 / args.steps,
                        "roofline": roofline(64 * BYTES_R10, ms_r10 * 1e-3 / args.steps, peak,
                                             "stft_image_kernel<6,0>, 5 168 000 B/record", "stft_diffeeg_dram_bytes_per_launch")},
        }

    if "xf" in legs:
        result["eeg_transform"] = leg_eeg_transform(torch, b200spec, timed, keep, peak, world, args)
    if "next" in legs:
        nxt = leg_next_rows(torch, b200spec, timed, keep, peak, world, args)
        if nxt:
            result["next_rows"] = nxt

    if "cfg4" in legs:
        # ------------------------------------------------------------ BASELINE configs[3]: one 1-hour record
        x_h = synth((20, L_HOUR), seed=4)
        n_win = b200spec.sliding_window_count(L_HOUR, WINDOW, STRIDE)

        def step_h():
            keep["h"] = b200spec.eeg_to_spectrogram_sliding(x_h, WINDOW, STRIDE, params)

        ms_h, _ = timed(step_h, 10, 3)
        keep.pop("h", None)
        bytes_h = 4 * 20 * L_HOUR + n_win * 4 * 128 * 256 * 4
        x_h3 = x_h.unsqueeze(0)

        def step_hr():
            keep["h"] = b200spec.compute_stft(x_h3, Cfg)

        flush = torch.zeros(256 << 20 >> 2, device="cuda")
        ms_hr, _ = timed(step_hr, 10, 3, flush=flush)
        del flush
        keep.pop("h", None)
        bytes_hr = 4 * 20 * L_HOUR + 4 * 20 * 33 * 2000
        result["cfg4"] = {
            "what": "BASELINE configs[3]: one synthetic 1-hour record (20 ch x 720000) per GPU",
            "mode_n": {"workload": f"{n_win} overlapping 50-s windows at stride {STRIDE} read in place (57.6 MB in, not 1.4 GB) "
                                   f"-> ({n_win},128,256,4)", "windows": n_win,
                       "value": world * n_win * 10 / (ms_h * 1e-3), "unit": "windows/s", "ms_per_step": ms_h / 10,
                       "roofline": roofline(bytes_h, ms_h * 1e-3 / 10, peak,
                                            "melspec_ts_kernel over overlapping records; algorithmic bytes = the record once + "
                                            "524 288 B per window")},
            "mode_r": {"workload": "(1,20,720000) -> (1,20,33,2000): 22 501 frames per row, decimating lerp; L2 flushed",
                       "value": world * 10 / (ms_hr * 1e-3), "unit": UNIT, "ms_per_step": ms_hr / 10,
                       "roofline": roofline(bytes_hr, ms_hr * 1e-3 / 10, peak,
                                            "stft_long_pixels_kernel + stft_long_normalise_kernel (decimating plan: column chunks, 16 x 20 CTAs), "
                                            "three launches of ~15 us: launch-latency bound")}}
        del x_h, x_h3

    if "cfg3" in legs:
        # ------------------------------------------------------------ BASELINE configs[2]: the 106 800-segment set
        from b200spec.sharding import shard_bounds

        x_n = None
        keep.clear()
        torch.cuda.empty_cache()
        lo, hi = shard_bounds(N_DATASET, world, rank)
        n_mine = hi - lo
        images = torch.empty((n_mine, 128, 256, 4), dtype=torch.float32, device="cuda")   # retained: 56 GB / world
        g = torch.Generator(device="cuda").manual_seed(3 + 7919 * rank)
        buf = torch.empty((CHUNK_DATASET, 20, 10000), device="cuda")
        ms_c, chunks = 0.0, 0
        barrier()
        for c0 in range(0, n_mine, CHUNK_DATASET):
            nb = min(CHUNK_DATASET, n_mine - c0)
            xb = buf[:nb]
            torch.randn(xb.shape, generator=g, device="cuda", out=xb)          # generated per chunk, outside the clock
            xb.cumsum_(-1).mul_(0.3).add_(torch.randn(xb.shape, generator=g, device="cuda")).clamp_(-1024, 1024)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            b200spec.eeg_to_spectrogram(xb, params, out=images[c0:c0 + nb])
            e1.record()
            torch.cuda.synchronize()
            ms_c += e0.elapsed_time(e1)
            chunks += 1
        ms_c = reduce_max(ms_c)
        finite = bool(torch.isfinite(images[::997]).all())
        result["cfg3"] = {
            "what": "BASELINE configs[2]: the full HMS-scale synthetic set, rank r transforms records "
                    "[r N/G, (r+1) N/G) in chunks of 4096 and keeps every image in HBM",
            "records_total": N_DATASET, "records_this_rank": n_mine, "chunks_this_rank": chunks,
            "images_resident_GB_per_gpu": n_mine * 524288 / 1e9, "scaling": "strong",
            "value": N_DATASET / (ms_c * 1e-3), "unit": UNIT, "ms_total": ms_c,
            "roofline": roofline(n_mine * BYTES_N, ms_c * 1e-3, peak,
                                 "melspec_ts_kernel, one launch per 4096-record chunk (sum of CUDA-event times of the "
                                 "launches, max over ranks; input synthesis between launches is not timed)"),
            "outputs_finite": finite}
        del images, buf

    if "e2e" in legs:
        # ------------------------------------------------------------ end to end: host buffers in and out
        x_n = None
        keep.clear()
        torch.cuda.empty_cache()
        e2e_steps = max(2, min(args.steps, 5))
        h_n = synth((RECORDS, 20, 10000), seed=3).cpu().pin_memory()
        # warm-up the way the timed loop runs: the previous result stays alive while the next one
        # is produced, so BOTH pinned result buffers exist before the clock starts (pinning 537 MB
        # takes ~0.2 s once; torch's host allocator then recycles the two blocks)
        for _ in range(3):
            out = b200spec.eeg_to_spectrogram(h_n, params)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            out = b200spec.eeg_to_spectrogram(h_n, params)
        barrier()
        e2e_s = reduce_max(time.perf_counter() - t0)
        import ctypes
        from b200spec import _native as nat
        h2d_b, d2h_b = ctypes.c_int64(), ctypes.c_int64()   # the bytes the call copies: 16 of the 20 electrode rows in
        nat.check(nat.lib.eegspec_melspec_host_bytes(RECORDS, 20, 10000, ctypes.byref(params), ctypes.byref(h2d_b),
                                                     ctypes.byref(d2h_b)), "melspec_host_bytes")
        assert d2h_b.value == out.numel() * 4
        result["e2e"] = {"value": world * RECORDS * e2e_steps / e2e_s, "unit": UNIT,
                         "h2d_bytes_per_step": h2d_b.value, "d2h_bytes_per_step": d2h_b.value,
                         "steps": e2e_steps,
                         "link_GBps_both_directions": world * (h2d_b.value + d2h_b.value) * e2e_steps / e2e_s / 1e9,
                         "what": "b200spec.eeg_to_spectrogram(pinned host tensor) -> pinned host tensor; "
                                 "H2D (the electrode rows the montage reads), kernels and D2H pipelined in 64-record chunks "
                                 "inside the C ABI call"}
        pcie = committed_traffic(f"pcie_{world}rank_GBps_both_directions_per_gpu")
        if pcie:
            result["e2e"]["fraction_of_measured_link_ceiling"] = result["e2e"]["link_GBps_both_directions"] / world / pcie
        h_r = h_n[:256]
        for _ in range(3):
            out_r = b200spec.compute_stft(h_r, Cfg)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            out_r = b200spec.compute_stft(h_r, Cfg)
        barrier()
        e2e_r = reduce_max(time.perf_counter() - t0)
        result.setdefault("mode_r", {})["e2e"] = {"value": world * 256 * e2e_steps / e2e_r, "unit": UNIT,
                                   "h2d_bytes_per_step": h_r.numel() * 4, "d2h_bytes_per_step": out_r.numel() * 4,
                                   "workload": "(256,20,10000) host -> (256,20,33,2000) host; output is 33x the input, PCIe-bound"}
        del h_n, h_r, out, out_r

    if sampler:
        sampler.stop()
    if dist is not None:
        dist.barrier()
    if rank == 0:
        if world == 1 and "cpu" in legs:
            from oracle import cpu_bench

            arm = cpu_bench.CpuArm()
            n_rec = arm.cores * CPU_RECORDS_PER_WORKER          # ~20 core-seconds of the oracle port (20 records/s per core)
            arm.run("n", arm.cores * 2, seed=99)
            c_n = arm.run("n", n_rec, seed=2)
            c_r = arm.run("r10", arm.cores * 64, seed=5)
            c_r50 = arm.run("r50", arm.cores * 16, seed=6)
            c_xf = arm.run("xf", arm.cores * 32, seed=7)
            c_next = {k: arm.run(k, n * arm.cores, seed=21) for k, n in (("f2", 2), ("f3", 16), ("f4", 8))} \
                if "next_rows" in result else {}
            arm.close()
            for k, name in (("f2", "f2_diffeeg_condition"), ("f3", "f3_process_spectrogram"), ("f4", "f4_process_eeg")):
                if k in c_next and name in result["next_rows"]:
                    result["next_rows"][name]["cpu_baseline"] = {
                        "value": c_next[k]["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                        "sample": f"{c_next[k]['records']} records through oracle/"
                                  + {"f2": "diffeeg_oracle.py (torch CPU conv_transpose2d + interpolate, 1 thread per worker)",
                                     "f3": "classifier_oracle.py::process_spectrogram_oracle",
                                     "f4": "classifier_oracle.py::process_eeg_oracle"}[k],
                        "per_core": c_next[k]["per_core_records_per_s"]}
            if "eeg_transform" in result:
                result["eeg_transform"]["cpu_baseline"] = {
                    "value": c_xf["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                    "sample": f"{c_xf['records']} records (10000,20) through oracle/eegtransform_oracle.py "
                              "(the reference's own NumPy / SciPy calls)", "per_core": c_xf["per_core_records_per_s"]}
            result["cpu_baseline"] = {
                "value": c_n["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                "sample": f"{c_n['records']} of the 1024 records ({c_n['per_worker']} per worker), mel image via "
                          f"oracle/melspec_oracle.py (librosa restatement; librosa is not installed), one process per core, "
                          f"inputs resident before the clock starts, {c_n['wall_s']:.1f} s",
                "per_core": c_n["per_core_records_per_s"]}
            result.setdefault("mode_r", {})["cpu_baseline"] = {
                "diffeeg": {"value": c_r["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                            "sample": f"{c_r['records']} records (19,2000) through compute_stft_port", "per_core": c_r["per_core_records_per_s"]},
                "hms50s": {"value": c_r50["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                           "sample": f"{c_r50['records']} records (20,10000) through compute_stft_port", "per_core": c_r50["per_core_records_per_s"]}}
        emit(result)
    if dist is not None:
        dist.destroy_process_group()


def leg_eeg_transform(torch, b200spec, timed, keep, peak, world, args):
    """Row (f)1: the raw-EEG transformer on the reference's real input, (L, 19) windows of _get_eeg_window."""
    out = {"what": "_EEGTransformer.transform drop-in (row f1; reference-pinned parity): select 19 / clip / NaN->0 / "
                   "/32 / Butterworth-4 lfilter / decimate 5, one kernel per call"}
    xf = b200spec.EEGTransformer(n_feats=19, downsample=5)
    gen = torch.Generator(device="cuda").manual_seed(7)
    for n_ch, key in ((20, "cols20"), (19, "cols19")):
        x_xf = torch.randn((RECORDS, 10000, n_ch), device="cuda", generator=gen) * 40.0

        def step_xf():
            keep["xf"] = xf.transform(x_xf)

        ms_xf, _ = timed(step_xf, args.steps, 3)
        keep.pop("xf", None)
        bytes_xf = 4 * 10000 * n_ch + 4 * 2000 * 19
        out[key] = {
            "workload": f"(1024,10000,{n_ch})->(1024,2000,19) per GPU, {RECORDS * 40000 * n_ch / 1e6:.0f} MB in per step (> L2)"
                        + (" -- the window cfg_utils.py:366-373 hands the transformer" if n_ch == 19 else
                           " -- all 20 parquet columns, EKG ignored"),
            "value": world * RECORDS * args.steps / (ms_xf * 1e-3), "unit": UNIT, "ms_per_step": ms_xf / args.steps,
            "roofline": roofline(RECORDS * bytes_xf, ms_xf * 1e-3 / args.steps, peak,
                                 f"eeg_transform_lane_kernel, {bytes_xf} B/record",
                                 "eeg_transform_dram_bytes_per_launch" if n_ch == 20 else None)}
        del x_xf
    # headline of the row = the reference's own layout
    out.update({k: out["cols19"][k] for k in ("workload", "value", "unit", "ms_per_step", "roofline")})
    if hasattr(b200spec, "eeg_transform_stft"):
        x19 = torch.randn((RECORDS, 10000, 19), device="cuda", generator=gen) * 40.0

        class Cfg:
            diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

        def step_fused():
            keep["xf"] = b200spec.eeg_transform_stft(x19, Cfg, xf)

        xf_cm = b200spec.EEGTransformer(n_feats=19, downsample=5, channel_major=True)

        def step_two():
            keep["xf"] = b200spec.compute_stft(xf_cm.transform(x19), Cfg)

        ms_f, _ = timed(step_fused, max(3, args.steps // 2), 3)
        ms_2, _ = timed(step_two, max(3, args.steps // 2), 3)
        keep.pop("xf", None)
        st = max(3, args.steps // 2)
        bytes_f = 4 * 10000 * 19 + 4 * 19 * 33 * 2000
        out["fused_stft"] = {
            "what": "raw (B,10000,19) -> (B,19,33,2000) in one entry point, eegspec_eeg_transform_stft_f32 (dataset.py:73-104 -> "
                    "DiffEEG_utils.py:27-76): two launches, the 2000-sample rows (3 % of the bytes) in a workspace between them; "
                    "L2-sized passes were measured slower (DESIGN.md)",
            "value": world * RECORDS * st / (ms_f * 1e-3), "unit": UNIT, "ms_per_step": ms_f / st,
            "two_kernel_ms_per_step": ms_2 / st,
            "roofline": roofline(RECORDS * bytes_f, ms_f * 1e-3 / st, peak, f"eeg_transform_stft, {bytes_f} B/record")}
        del x19
    torch.cuda.empty_cache()
    return out


class DiffEegCfg:
    """The diffEEG_trainer keys compute_stft reads (configs/config.yaml)."""
    diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}


def leg_next_rows(torch, b200spec, timed, keep, peak, world, args):
    """Rows (f)2-(f)4: the consumer front end of DiffEEG and the classifier-side preprocessing."""
    out = {}
    gen = torch.Generator(device="cuda").manual_seed(13)
    steps = max(3, args.steps // 2)

    # ---- (f)2: spectrogram condition of DiffEEG on the DiffEEG micro-batch and on 1 024 records
    w = torch.randn((19, 16, 3, 3), device="cuda", generator=gen) * 0.1
    b = torch.randn((16,), device="cuda", generator=gen) * 0.1
    f2 = {"what": "DiffEEG.forward condition branch (diffEEG.py:47-53, 108-110): ConvTranspose2d(19->16, (3,3), stride (1,8)) "
                  "-> relu -> flatten -> F.interpolate(2000) as ONE gather kernel; the (B,16,33,15991) intermediate is never formed"}
    for batch, key in ((64, "diffeeg_batch"), (1024, "records1024")):
        img = torch.rand((batch, 19, 33, 2000), device="cuda", generator=gen)
        flush = torch.zeros(256 << 20 >> 2, device="cuda") if batch == 64 else None

        def step_c():
            keep["c"] = b200spec.spectrogram_condition(img, w, b, 2000)

        ms, _ = timed(step_c, steps, 3, flush=flush)
        keep.pop("c", None)
        # bytes the operator has to touch: the 32-byte sectors of the gathered image values + the output
        live = 1500 * 19 * 3 * 32                      # 3/8 of 4000 positions x C x 3 rows, one sector each
        bytes_alg = batch * (live + 4 * 16 * 2000)
        f2[key] = {"workload": f"({batch},19,33,2000) image -> ({batch},16,2000); "
                               + ("L2 flushed between steps" if batch == 64 else "5.1 GB of images resident (> L2)"),
                   "value": world * batch * steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
                   "roofline": roofline(bytes_alg, ms * 1e-3 / steps, peak,
                                        "spec_condition_kernel; algorithmic bytes = 85 500 gathered image values as 32-byte "
                                        "sectors (2.7 MB) + 128 KB out per record",
                                        image_GBps_if_read_whole=batch * 4 * 19 * 33 * 2000 / (ms * 1e-3 / steps) / 1e9)}
        del img, flush

        # the same condition straight from the (B,19,2000) rows: compute_stft's kernel hands over its normalised knots
        # (33 x 64 {value, delta} per row) and the gather evaluates the 0.76 % of the image it needs from them
        x = torch.randn((batch, 19, 2000), device="cuda", generator=gen)
        flush = torch.zeros(256 << 20 >> 2, device="cuda") if batch == 64 else None

        def step_k():
            keep["c"] = b200spec.stft_condition(x, DiffEegCfg, w, b, 2000)

        def step_2():
            keep["c"] = b200spec.spectrogram_condition(b200spec.compute_stft(x, DiffEegCfg), w, b, 2000)

        ms_k, _ = timed(step_k, steps, 3, flush=flush)
        ms_2, _ = timed(step_2, steps, 3, flush=flush)
        keep.pop("c", None)
        bytes_k = 4 * 19 * 2000 + 4 * 16 * 2000
        f2[key + "_from_rows"] = {
            "what": "b200spec.stft_condition = eegspec_stft_knots_f32 + eegspec_spec_condition_knots_f32: compute_stft and the "
                    "condition without the (B,19,33,2000) image in between; bit-identical to the two-call path",
            "workload": f"({batch},19,2000) rows -> ({batch},16,2000)",
            "value": world * batch * steps / (ms_k * 1e-3), "unit": UNIT, "ms_per_step": ms_k / steps,
            "two_call_ms_per_step": ms_2 / steps,
            "roofline": roofline(batch * bytes_k, ms_k * 1e-3 / steps, peak,
                                 f"stft_image_kernel (knots only) + spec_condition_kernel, {bytes_k} B/record algorithmic "
                                 "(rows in, condition out); the 321 KB of knots per record in between stay mostly in L2")}
        del x, flush
    out["f2_diffeeg_condition"] = f2

    # ---- (f)4: classifier-side EEG preprocessing
    x = torch.randn((RECORDS, 10000, 20), device="cuda", generator=gen).cumsum_(1).mul_(0.5)
    x.add_(torch.randn((RECORDS, 10000, 20), device="cuda", generator=gen) * 5.0)

    def step_e():
        keep["e"] = b200spec.process_eeg(x, check=False)

    ms, _ = timed(step_e, steps, 2)
    keep.pop("e", None)
    bytes_e = 4 * 10000 * 20 + 4 * 37 * 3000
    out["f4_process_eeg"] = {
        "what": "CombinedDataset.process_eeg (XAI_Multimodality.py:1148-1164): band-pass 5 -> NaN fill -> 18 differences -> "
                "band-pass 6 + 4-tap mean + decimate 4 -> z-score -> (1,37,3000); SciPy-exact float64 recurrences "
                "(20 + 38 per record, one thread each): bound by the latency of the float64 recurrences, not by HBM",
        "workload": "(1024,10000,20) parquet rows -> (1024,1,37,3000)",
        "value": world * RECORDS * steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
        "fp64_Gops_per_s": RECORDS * 10000 * (20 * 31 + 38 * 37) / (ms * 1e-3 / steps) / 1e9,
        "roofline": roofline(RECORDS * bytes_e, ms * 1e-3 / steps, peak,
                             f"classifier_eeg_first / second / finish kernels, {bytes_e} B/record algorithmic")}
    del x

    # ---- (f)3: classifier-side spectrogram post-processing
    win = torch.rand((RECORDS, 400, 300), device="cuda", generator=gen) * 40.0
    win[:, 300:] = 0.0

    def step_s():
        keep["s"] = b200spec.process_spectrogram(win, check=False)

    ms, _ = timed(step_s, steps, 2)
    keep.pop("s", None)
    bytes_s = 4 * 400 * 300 * 4
    out["f3_process_spectrogram"] = {
        "what": "CombinedDataset.process_spectrogram (XAI_Multimodality.py:1183-1202): NaN fill -> baseline -> 60 Hz notch "
                "filtfilt -> Gaussian -> min-max -> 3 planes; four small kernels with fp32 intermediates",
        "workload": "(1024,400,300) windows -> (1024,3,400,300)",
        "value": world * RECORDS * steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
        "roofline": roofline(RECORDS * bytes_s, ms * 1e-3 / steps, peak,
                             f"cls_spec_rowfill / column / row / planes kernels, {bytes_s} B/record algorithmic")}
    del win
    torch.cuda.empty_cache()
    return out


if __name__ == "__main__":
    main()
