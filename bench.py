#!/usr/bin/env python
"""Benchmark of the hot path: lmfe (40 filters, 20 ms / 10 ms, NFFT 512) + cmvnw (301,
variance) over a VoxCeleb1-dev-shaped synthetic corpus (BASELINE.json config 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one pass of the fused path over the
whole per-GPU corpus, inputs resident in HBM (``value``); ``e2e`` is the same metric
through the host-pointer C-ABI call with pinned host buffers (H2D + D2H inside the timed
region) on a bounded slice of the same corpus.  Weak scaling: every rank featurises its
own corpus of the same shape (different seed); there is no data-path collective.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FS = 16000
L, S, M, NFFT, WIN = 320, 160, 40, 512, 301
CFG3_UTTS = 148642
METRIC = "audio-hours/sec featurized (lmfe+cmvnw)"
UNIT = "audio-hours/s"


def frames_of(lengths):
    return np.maximum((lengths - L) // S, 0)


def algorithmic_bytes(lengths):
    """read 4N (fp32 samples) + write 4*M*F (fp32 features)."""
    return int(4 * lengths.sum() + 4 * M * frames_of(lengths).sum())


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self.stop_flag = threading.Event()

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            names = {
                pynvml.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                pynvml.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                pynvml.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                pynvml.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            }
            while not self.stop_flag.is_set():
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.02)
        except Exception as e:  # NVML missing: report that rather than fail the bench
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        self.stop_flag.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's speechpy (the reference is pure Python and
# cannot travel to the GPU box), one process per host core, OMP threads pinned to 1.
# ---------------------------------------------------------------------------------------
def _cpu_worker(args):
    seed, n = args
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle import speechpy_oracle as o
    from speaker_verification_b200 import synth
    x = synth.pcm16(seed, int(n)).astype(np.float32) * np.float32(1.0 / 32768.0)  # == librosa.load of a PCM16 wav
    out = o.cmvnw(o.lmfe(x, FS), WIN, True)
    return out.shape[0]


def cpu_pool(cores):
    import multiprocessing as mp
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[k] = "1"
    return mp.get_context("fork").Pool(cores)


def cpu_rate(pool, lengths, seed0):
    """audio-hours/s of the oracle port over `lengths` on the pool (wall clock)."""
    jobs = [(seed0 + i, int(n)) for i, n in enumerate(lengths)]
    t0 = time.perf_counter()
    pool.map(_cpu_worker, jobs, chunksize=1)
    dt = time.perf_counter() - t0
    return float(lengths.sum()) / FS / 3600.0 / dt, dt


# ---------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (oracle port,
    kind "port") on all host cores; each step is a bounded sample of the cfg3 workload."""
    if rank != 0:
        return
    from speaker_verification_b200 import synth
    cores = os.cpu_count() or 1
    pool = cpu_pool(cores)
    per_step = max(32, 2 * cores)
    lengths_all = synth.cfg3_lengths(per_step * (args.steps + args.warmup), seed=3)
    k = 0
    for _ in range(args.warmup):
        cpu_rate(pool, lengths_all[k:k + per_step], 10_000 + k)
        k += per_step
    t_total, audio = 0.0, 0.0
    for _ in range(args.steps):
        ls = lengths_all[k:k + per_step]
        _, dt = cpu_rate(pool, ls, 10_000 + k)
        t_total += dt
        audio += float(ls.sum()) / FS / 3600.0
        k += per_step
    pool.close()
    value = audio / t_total
    sample = f"{per_step} cfg3-shaped utterances per step ({args.steps} timed steps)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic 16 kHz audio, 16-bit PCM round(N(0,0.1)*32768) scaled by 1/32768 to float32 (what "
                "librosa.load returns), Philox-seeded",
        "config": workload_config(per_step, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(utts_per_gpu, gpus):
    return {"workload": "cfg3: ragged 4-20 s 16 kHz utterances (VoxCeleb1-dev shape, "
                        "clip(lognormal(ln 7.2 s, 0.5), 4, 20)), lmfe(40, 20ms/10ms, NFFT 512) + cmvnw(301, var)",
            "utterances_per_gpu": int(utts_per_gpu), "cfg3_total_utterances": CFG3_UTTS,
            "sharding": f"utterances, {gpus} rank(s), no collective",
            "l2": "inputs larger than L2 (no flush needed)"}


def make_device_corpus(torch, lengths, seed, device):
    align = 4
    padded = (lengths + align - 1) // align * align
    starts = np.zeros(len(lengths), dtype=np.int64)
    np.cumsum(padded[:-1], out=starts[1:])
    total = int(padded.sum())
    arena = torch.empty(total, dtype=torch.float32, device=device)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    chunk = 1 << 28
    for s in range(0, total, chunk):
        # 16-bit PCM values / 32768: the float32 audio librosa.load (utils.py:170-173) decodes from a wav
        arena[s:s + chunk].normal_(0.0, 0.1 * 32768.0, generator=gen).round_().clamp_(-32768.0, 32767.0).mul_(1.0 / 32768.0)
    return arena, starts


def run_ours(args, rank, world, local_rank):
    from speaker_verification_b200 import synth
    # ---- CPU baseline first (rank 0, N=1 only), before CUDA is initialised in this process
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        pool = cpu_pool(cores)
        ls = synth.cfg3_lengths(max(64, 4 * cores), seed=3)
        cpu_rate(pool, ls[:cores], 20_000)  # warm the workers
        rate, dt = cpu_rate(pool, ls, 30_000)
        pool.close()
        cpu_baseline = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{len(ls)} cfg3-shaped utterances ({ls.sum() / FS / 3600:.3f} audio-h), "
                                  f"oracle port of speechpy lmfe+cmvnw, one process per core, {dt:.1f} s wall"}

    # pin this rank to the CPUs next to its GPU so that the pinned host buffers of the e2e leg
    # are first-touched on the GPU's NUMA node (matters when several ranks share the host)
    numa = "unset"
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        numa = f"{len(os.sched_getaffinity(0))} cpus near gpu {local_rank}"
    except Exception as e:  # not fatal: the bench still runs, just without NUMA placement
        numa = f"unavailable:{type(e).__name__}"
    import torch
    import torch.distributed as dist
    from speaker_verification_b200 import _lib, batch
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    U = args.utts
    lengths = synth.cfg3_lengths(U, seed=3 + 1000 * rank)
    arena, starts = make_device_corpus(torch, lengths, 77 + rank, device)
    fb = batch.FeatureBatch(starts, lengths, device=local_rank)
    ctx = fb.ctx
    out = torch.empty((fb.total_frames, M), dtype=torch.float32, device=device)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=device)
    audio_h = float(lengths.sum()) / FS / 3600.0
    alg_bytes = algorithmic_bytes(lengths)

    def step():
        fb.lmfe_cmvnw_dev(arena.data_ptr(), arena.numel(), out.data_ptr())

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    ctx.timing_enable(True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    n0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    clocks = sampler.result()
    launches = ctx.launch_count() - n0
    ms = e0.elapsed_time(e1)
    k_ms = [ctx.timing_read(k) for k in range(3)]
    ctx.timing_enable(False)
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * audio_h * args.steps / (ms_max / 1e3)

    # ---- e2e: host-pointer C-ABI call, pinned host buffers, H2D and D2H inside the timed region
    Ue = max(1, U // args.e2e_fraction)
    e_len = lengths[:Ue]
    e_end = int(starts[Ue - 1] + lengths[Ue - 1])
    fbe = batch.FeatureBatch(starts[:Ue], e_len, device=local_rank)
    h_out = torch.empty((fbe.total_frames, M), dtype=torch.float32, pin_memory=True)
    e_steps = max(3, min(args.steps, 10))
    e_audio_h = float(e_len.sum()) / FS / 3600.0

    def e2e_leg(host_dtype):
        h_in = torch.empty(e_end, dtype=host_dtype, pin_memory=True)
        pcm = host_dtype == torch.int16
        h_in.copy_(arena[:e_end] * 32768.0 if pcm else arena[:e_end])  # exact: the corpus is PCM / 32768
        scale = batch.PCM16_SCALE if pcm else 1.0
        a_np, o_np = h_in.numpy(), h_out.numpy()
        for _ in range(2):
            fbe.lmfe_cmvnw(a_np, o_np, sample_scale=scale)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            fbe.lmfe_cmvnw(a_np, o_np, sample_scale=scale)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        te = torch.tensor([dt], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        return world * e_audio_h * e_steps / float(te.item())

    e2e_f32 = e2e_leg(torch.float32)
    e2e_value = e2e_leg(torch.int16)  # last: h_out holds the int16 leg's result for the check below
    # sanity: device path and host path agree bit for bit on the shared prefix
    same = bool(torch.equal(out[:fbe.total_frames].cpu(), h_out))

    # ---- config 5 beside it (rank 0): the cosine score matrix on the tensor cores + EER
    scoring_line = None
    if rank == 0:
        from speaker_verification_b200 import scoring
        gen = torch.Generator(device=device)
        gen.manual_seed(5)
        m5 = n5 = 4874
        d5 = 512
        E5 = torch.randn((m5, d5), generator=gen, device=device)
        T5 = torch.randn((n5, d5), generator=gen, device=device)
        S5 = torch.empty((m5, n5), dtype=torch.float32, device=device)
        for _ in range(3):
            scoring.cosine_scores_dev(E5, T5, S5)
        ctx.timing_enable(True)
        for _ in range(10):
            scoring.cosine_scores_dev(E5, T5, S5)
        g_ms, g_n = ctx.timing_read(3)
        ctx.timing_enable(False)
        ti5 = torch.randint(0, m5, (37720,), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
        tj5 = torch.randint(0, n5, (37720,), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
        lab5 = (torch.rand(37720, generator=gen, device=device) < 0.5).to(torch.int32)
        tr5 = scoring.trial_scores_dev(S5, ti5, tj5) + 0.05 * lab5
        scoring.eer_auc_dev(tr5, lab5)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eer5, auc5 = scoring.eer_auc_dev(tr5, lab5)
        eer_ms = (time.perf_counter() - t0) * 1e3
        per = g_ms / max(g_n, 1) / 1e3
        flops = 2.0 * m5 * n5 * d5
        tf_peak = None
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                tf_peak = float(json.load(fh)["bf16_tflops"]) / 2.0  # TF32 runs at half the bf16 rate
        except Exception:
            pass
        scoring_line = {"workload": "cfg5: cosine scores 4874 x 4874 x 512 (3xTF32 on tcgen05) + EER of 37720 trials",
                        "gemm_us": per * 1e6, "useful_tflops": flops / per / 1e12,
                        "issued_tf32_tflops": 3 * flops / per / 1e12,
                        "tensor_frac_of_measured_bf16_over_2": (3 * flops / per / 1e12 / tf_peak) if tf_peak else None,
                        "eer_call_ms": eer_ms, "eer": eer5, "auc": auc5}
        del E5, T5, S5

    if rank == 0:
        peak, peak_src = peaks()
        feat_ms, feat_n = k_ms[0]
        achieved = alg_bytes * args.steps / (feat_ms / 1e3) / 1e9 if feat_ms > 0 else None
        step_gbs = alg_bytes * args.steps / (ms / 1e3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (f64 window accumulators in cmvnw)",
            "data": "synthetic 16 kHz audio, 16-bit PCM round(N(0,0.1)*32768) scaled by 1/32768 to float32 (what "
                    "librosa.load returns), generated on device (torch Philox), random lengths seeded",
            "config": workload_config(U, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(e_end * 2),
                    "d2h_bytes_per_step": int(fbe.total_frames * M * 4),
                    "sample": f"first {Ue} utterances of the corpus ({e_audio_h:.2f} audio-h) per step, "
                              f"{e_steps} steps, pinned host buffers, int16 PCM arena in (the wav payload; scaled by "
                              f"1/32768 on the device like librosa.load), float32 features out, spb_lmfe_cmvnw_batch_dtype "
                              f"(H2D / kernels / D2H pipelined over chunks)",
                    "host_affinity": numa,
                    "float32_arena_value": e2e_f32, "float32_arena_h2d_bytes_per_step": int(e_end * 4),
                    "matches_device_path": same},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {
                "bound": "hbm", "kernel": "features_tc_kernel (framing + hop-DFT on tcgen05 + |X|^2 + mel + log; "
                                          "timed with its flag memset and the guarded no-op FFT launch)",
                "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None,
                # ncu --set full (profiles/r01g_ncu_tc_summary.csv, 4874 x 8 s launch): dram read 2.501 GB +
                # write 0.606 GB for 3.118 GB algorithmic (samples read once, rows written once)
                # -> 0.997 B of DRAM traffic per algorithmic byte; scaled to this run's launches
                "traffic": 0.997 * alg_bytes / max(feat_n // max(args.steps, 1), 1),
                "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum of features_tc_kernel in "
                                  "profiles/r01g_ncu_tc_summary.csv (0.997 x algorithmic bytes), per launch",
                "achieved_per_launch_bytes": alg_bytes / max(feat_n // max(args.steps, 1), 1),
                "peak_source": peak_src,
                "algorithmic_bytes_per_step": alg_bytes,
                "kernel_ms_per_step": {"features": feat_ms / args.steps,
                                       "cmvnw_stage1": k_ms[1][0] / args.steps,
                                       "cmvnw_stage2": k_ms[2][0] / args.steps},
                "launches_per_step": {"features": feat_n // args.steps,
                                      "cmvnw_stage1": k_ms[1][1] // args.steps,
                                      "cmvnw_stage2": k_ms[2][1] // args.steps},
                "whole_step_achieved": step_gbs, "whole_step_frac": step_gbs / peak,
                "nominal_8000_frac_whole_step": step_gbs / 8000.0,
            },
            "audio_hours_per_step_per_gpu": audio_h,
        }
        if scoring_line is not None:
            line["scoring"] = scoring_line
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--utts", type=int, default=int(os.environ.get("SPB_BENCH_UTTS", CFG3_UTTS)),
                    help="utterances per GPU (default: all of config 3)")
    ap.add_argument("--e2e-fraction", type=int, default=16,
                    help="the e2e leg uses the first 1/this of the corpus")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
