#!/usr/bin/env python
"""Benchmark of the EEG -> spectrogram hot path (BASELINE.json metric: records -> spectrograms / s).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [--gpus N] ...                # the reference's CPU path (oracle port)

A step is one pass of the transform over one batch of synthetic EEG.  The headline workload is
BASELINE.json configs[1]: 1 024 records of 20 ch x 10 000 samples -> (1024, 128, 256, 4), the
4-chain x 128-mel image ("mode N").  The same run also measures the reference-pinned
``compute_stft`` image ("mode R", object ``mode_r``) on (1024, 20, 10000) and on the DiffEEG
micro-batch (64, 19, 2000).  Under torchrun every rank transforms its own 1 024 records (weak
scaling, no collective on the data path); timing is CUDA events per rank, max over ranks.
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
WORKLOAD_N = ("BASELINE configs[1]: 1024 synthetic 50-s EEG records (20 ch x 10000 @200 Hz) -> "
              "(1024,128,256,4) 4-chain x 128-mel log spectrogram stack")
BYTES_N = 4 * 20 * 10000 + 4 * 128 * 256 * 4            # 1 324 288 B / record
BYTES_R50 = 4 * 20 * 10000 + 4 * 20 * 33 * 2000         # 6 080 000 B / record
BYTES_R10 = 4 * 19 * 2000 + 4 * 19 * 33 * 2000          # 5 168 000 B / record


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


def melspec_step_traffic():
    """One step = 4 chunks of 256 records, each one pair-kernel + one combine-kernel launch."""
    a = committed_traffic("melspec_pair_dram_bytes_per_launch")
    b = committed_traffic("melspec_combine_dram_bytes_per_launch")
    return None if a is None or b is None else 4 * (a + b)


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
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for ts, line in self.rows:
            p = [s.strip() for s in line.split(",")]
            if len(p) < 8 or not p[0].isdigit() or int(p[0]) not in self.gpu_ids:
                continue
            if not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            try:
                sm.append(float(p[1]))
                smax.append(float(p[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def reference_arm(args, rank, world):
    """The reference's CPU implementation of the path (oracle port), all host cores, rank 0 only."""
    if rank != 0:
        return
    from oracle import cpu_bench

    arm = cpu_bench.CpuArm()
    per_step = max(arm.cores * 2, 16)
    for _ in range(args.warmup):
        arm.run("n", max(arm.cores, 8), seed=99)
    t0 = time.perf_counter()
    done = 0
    for k in range(args.steps):
        done += arm.run("n", per_step, seed=1000 + 97 * k)["records"]
    wall = time.perf_counter() - t0
    r = arm.run("r10", max(arm.cores * 4, 32), seed=5)
    arm.close()
    value = done / wall
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD_N, "mode": "N", "records_per_step": per_step,
                   "note": "each step is a bounded sample of the 1024-record workload"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": "port",
                         "sample": f"{per_step} records/step x {args.steps} steps of the (20,10000)->(128,256,4) "
                                   "mel image through oracle/melspec_oracle.py (librosa restatement), one process per core"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mode_r": {"value": r["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                   "workload": "compute_stft (B,19,2000)->(B,19,33,2000), oracle/stft_oracle.py::compute_stft_port "
                               "(same SciPy/NumPy calls as DiffEEG_utils.py:27-76)",
                   "sample": f"{r['records']} records"},
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
    ap.add_argument("--quick", action="store_true", help="skip the mode-R and e2e legs (profiling runs)")
    ap.add_argument("--legs", default="n,r,e2e,cpu",
                    help="comma list of legs to run: n (always), r, e2e, cpu; profiling runs use a subset")
    args = ap.parse_args()
    legs = set(args.legs.split(",")) - ({"r", "e2e", "cpu"} if args.quick else set())
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

    def reduce_max(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def synth(shape, seed):
        g = torch.Generator(device="cuda").manual_seed(seed + 7919 * rank)
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

    # ---------------------------------------------------------------- headline: mode N, device resident
    x_n = synth((RECORDS, 20, 10000), seed=2)
    params = b200spec.mel_params()
    keep = {}

    def step_n():
        keep["n"] = b200spec.eeg_to_spectrogram(x_n, params)

    if sampler:
        sampler.start()
    t_start = time.perf_counter()
    ms_n, launches_n = timed(step_n, args.steps, args.warmup)
    t_end = time.perf_counter()
    clocks = sampler.stop(t_start, t_end) if sampler else None
    launches_total = reduce_sum(launches_n)
    value_n = world * RECORDS * args.steps / (ms_n * 1e-3)
    per_launch_s = ms_n * 1e-3 / args.steps
    achieved_n = RECORDS * BYTES_N / per_launch_s / 1e9
    roofline = {
        "bound": "hbm", "achieved": achieved_n, "peak": peak, "unit": "GB/s", "frac": achieved_n / peak,
        "traffic": committed_traffic("melspec_ts_dram_bytes_per_launch"),
        "peak_source": peak_src,
        "kernel": "melspec_ts_kernel (engine 3, frame operand in tensor memory: one persistent clustered launch per step; "
                  "algorithmic bytes 1 324 288 B/record x 1024 records per launch)",
        "note": "co-bound by issue slots / MUFU / tensor pipe, not by HBM alone (DESIGN.md 3.2)",
    }
    del keep["n"]

    result = {
        "metric": METRIC, "value": value_n, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_n / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD_N, "mode": "N", "records_per_gpu": RECORDS,
                   "l2": "inputs are 819 MB per step per GPU (> 126 MB L2); no flush needed",
                   "parallelism": f"records sharded over {world} GPU(s), no collective",
                   "host_cores_bound_rank0": len(bound)},
        "roofline": roofline, "gpu_launches": int(launches_total), "clocks": clocks,
    }

    if "r" in legs:
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

        # ------------------------------------------------------------ mode R, device resident
        def step_r50():
            keep["r"] = b200spec.compute_stft(x_n, Cfg)

        ms_r, _ = timed(step_r50, max(3, args.steps // 2), 3)
        steps_r = max(3, args.steps // 2)
        keep.pop("r", None)
        torch.cuda.empty_cache()
        x_r10 = synth((64, 19, 2000), seed=5)
        flush = torch.zeros(256 << 20 >> 2, device="cuda")

        def step_r10():
            keep["r"] = b200spec.compute_stft(x_r10, Cfg)

        ms_r10, _ = timed(step_r10, args.steps, 3, flush=flush)
        del flush
        keep.pop("r", None)
        a50 = RECORDS * BYTES_R50 / (ms_r * 1e-3 / steps_r) / 1e9
        a10 = 64 * BYTES_R10 / (ms_r10 * 1e-3 / args.steps) / 1e9
        result["mode_r"] = {
            "what": "compute_stft drop-in (reference-pinned parity), one fused kernel per call",
            "hms50s": {"workload": "(1024,20,10000)->(1024,20,33,2000) per GPU", "value": world * RECORDS * steps_r / (ms_r * 1e-3),
                       "unit": UNIT, "ms_per_step": ms_r / steps_r,
                       "roofline": {"bound": "hbm", "achieved": a50, "peak": peak, "unit": "GB/s", "frac": a50 / peak,
                                    "traffic": committed_traffic("stft_hms50s_dram_bytes_per_launch"),
                                    "kernel": "stft_image_kernel<6,0>, 6 080 000 B/record"}},
            "diffeeg": {"workload": "(64,19,2000)->(64,19,33,2000) per GPU, L2 flushed between steps",
                        "value": world * 64 * args.steps / (ms_r10 * 1e-3), "unit": UNIT, "ms_per_step": ms_r10 / args.steps,
                        "roofline": {"bound": "hbm", "achieved": a10, "peak": peak, "unit": "GB/s", "frac": a10 / peak,
                                     "traffic": committed_traffic("stft_diffeeg_dram_bytes_per_launch"),
                                     "kernel": "stft_image_kernel<6,0>, 5 168 000 B/record"}},
        }

    if "r" in legs:
        # ------------------------------------------------------------ row (f)1: raw-EEG transformer
        x_xf = torch.randn((RECORDS, 10000, 20), device="cuda", generator=torch.Generator(device="cuda").manual_seed(7)) * 40.0
        xf = b200spec.EEGTransformer(n_feats=19, downsample=5)

        def step_xf():
            keep["xf"] = xf.transform(x_xf)

        ms_xf, _ = timed(step_xf, args.steps, 3)
        keep.pop("xf", None)
        del x_xf
        torch.cuda.empty_cache()
        bytes_xf = 4 * 10000 * 20 + 4 * 2000 * 19
        a_xf = RECORDS * bytes_xf / (ms_xf * 1e-3 / args.steps) / 1e9
        result["eeg_transform"] = {
            "what": "_EEGTransformer.transform drop-in (row f1; reference-pinned parity): select 19 / clip / NaN->0 / "
                    "/32 / Butterworth-4 lfilter / decimate 5, one kernel per call",
            "workload": "(1024,10000,20)->(1024,2000,19) per GPU, 819 MB in per step (> L2)",
            "value": world * RECORDS * args.steps / (ms_xf * 1e-3), "unit": UNIT, "ms_per_step": ms_xf / args.steps,
            "roofline": {"bound": "hbm", "achieved": a_xf, "peak": peak, "unit": "GB/s", "frac": a_xf / peak, "traffic": committed_traffic("eeg_transform_dram_bytes_per_launch"),
                         "kernel": f"eeg_transform_lane_kernel<20,0>, {bytes_xf} B/record"}}

    if "e2e" in legs:
        # ------------------------------------------------------------ end to end: host buffers in and out
        x_n = None
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
                         "what": "b200spec.eeg_to_spectrogram(pinned host tensor) -> pinned host tensor; "
                                 "H2D (the electrode rows the montage reads), kernels and D2H pipelined in 64-record chunks "
                                 "inside the C ABI call"}
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

    if dist is not None:
        dist.barrier()
    if rank == 0:
        if world == 1 and "cpu" in legs:
            from oracle import cpu_bench

            arm = cpu_bench.CpuArm()
            n_rec = max(arm.cores * 24, 192)    # ~20 core-seconds of the oracle port (20 records/s per core)
            c_n = arm.run("n", n_rec, seed=2)
            c_r = arm.run("r10", max(arm.cores * 64, 512), seed=5)
            c_r50 = arm.run("r50", max(arm.cores * 16, 128), seed=6)
            c_xf = arm.run("xf", max(arm.cores * 32, 256), seed=7)
            arm.close()
            if "eeg_transform" in result:
                result["eeg_transform"]["cpu_baseline"] = {
                    "value": c_xf["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                    "sample": f"{c_xf['records']} records (10000,20) through oracle/eegtransform_oracle.py "
                              "(the reference's own NumPy / SciPy calls)", "per_core": c_xf["per_core_records_per_s"]}
            result["cpu_baseline"] = {
                "value": c_n["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                "sample": f"{c_n['records']} of the 1024 records, mel image via oracle/melspec_oracle.py "
                          f"(librosa restatement; librosa is not installed), one process per core, {c_n['wall_s']:.1f} s",
                "per_core": c_n["per_core_records_per_s"]}
            result.setdefault("mode_r", {})["cpu_baseline"] = {
                "diffeeg": {"value": c_r["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                            "sample": f"{c_r['records']} records (19,2000) through compute_stft_port", "per_core": c_r["per_core_records_per_s"]},
                "hms50s": {"value": c_r50["records_per_s"], "unit": UNIT, "cores": arm.cores, "kind": "port",
                           "sample": f"{c_r50['records']} records (20,10000) through compute_stft_port", "per_core": c_r50["per_core_records_per_s"]}}
        emit(result)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
