#!/usr/bin/env python
"""Benchmark of the hot path: lmfe (40 filters, 20 ms / 10 ms, NFFT 512) + cmvnw (301,
variance) over a VoxCeleb1-dev-shaped synthetic corpus (BASELINE.json config 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one pass of the fused path over the
corpus, inputs resident in HBM (``value``); ``e2e`` is the same metric through the
host-pointer C-ABI call with pinned host buffers (H2D + D2H inside the timed region) on a
bounded slice of the same corpus.  Strong scaling: the ONE 148 642-utterance corpus is dealt
to the N ranks (``batch.shard_indices``), there is no data-path collective; the weak-scaling
figure (every rank a full-size corpus of its own) is reported beside it at N > 1.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

# The CPU legs run one single-threaded process per host core: the BLAS / OpenMP pools must be sized
# BEFORE numpy is imported (a pool created at import time is inherited by every forked worker and
# oversubscribes the cores: round 1's CPU arm was 3-5x too slow at N=1 for exactly this reason).
for _k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS"):
    os.environ[_k] = "1"
# stdout carries ONE JSON line: whatever NCCL has to say (its version banner, NCCL_DEBUG=INFO of the caller) goes to
# stderr.  NCCL honours NCCL_DEBUG_FILE only above the VERSION level, so VERSION (this image's default) becomes WARN.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"

import numpy as np  # noqa: E402

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
# CPU arm: the reference's own speechpy (oracle/_ref, staged by oracle/make_ref.py: kind
# "reference") or, when that is absent, the oracle port of it (kind "port"); one single-threaded
# process per host core over a FIXED subset of cfg3-shaped utterances whose audio is generated
# before the clock starts (a bounded subset, not the full 3 CPU-hours).
# ---------------------------------------------------------------------------------------
CPU_SAMPLE_UTTS = 2048
_CPU_PCM = None      # list of int16 arrays, set in the parent before the timed pool is forked
_CPU_IMPL = None     # per-process cache of (lmfe, cmvnw)


def _cpu_impl():
    global _CPU_IMPL
    if _CPU_IMPL is None:
        from oracle import ref_loader
        ref = ref_loader.load()
        if ref is not None:
            _CPU_IMPL = (lambda x: ref.feature.lmfe(x, FS, frame_length=0.020, frame_stride=0.01, num_filters=M,
                                                     fft_length=NFFT), ref.processing.cmvnw)
        else:
            from oracle import speechpy_oracle as o
            _CPU_IMPL = (lambda x: o.lmfe(x, FS), o.cmvnw)
    return _CPU_IMPL


def cpu_kind():
    from oracle import ref_loader
    return "reference" if ref_loader.available() else "port"


def _cpu_gen(args):
    seed, n = args
    from speaker_verification_b200 import synth
    return synth.pcm16(seed, int(n))


def _cpu_worker(i):
    lmfe, cmvnw = _cpu_impl()
    x = _CPU_PCM[i].astype(np.float32) * np.float32(1.0 / 32768.0)  # == librosa.load of a PCM16 wav (utils.py:170-173)
    out = cmvnw(lmfe(x), WIN, True)
    return out.shape[0]


class CpuArm:
    """Pool of `cores` single-threaded workers over a fixed sample of cfg3 utterances."""

    def __init__(self, utts=CPU_SAMPLE_UTTS, cores=None):
        import multiprocessing as mp
        from speaker_verification_b200 import synth
        global _CPU_PCM
        self.cores = cores or (os.cpu_count() or 1)
        self.lengths = synth.cfg3_lengths(utts, seed=3)
        ctx = mp.get_context("fork")
        with ctx.Pool(self.cores) as gen:  # audio first, outside every timed region
            _CPU_PCM = gen.map(_cpu_gen, [(40_000 + i, int(n)) for i, n in enumerate(self.lengths)], chunksize=8)
        self.pool = ctx.Pool(self.cores)   # forked AFTER the audio exists: workers read it copy-on-write
        self.audio_h = float(self.lengths.sum()) / FS / 3600.0
        self.kind = cpu_kind()
        self.pool.map(_cpu_worker, range(min(2 * self.cores, utts)), chunksize=1)  # imports, page-in

    def step(self):
        """one pass over the sample; returns wall seconds"""
        t0 = time.perf_counter()
        self.pool.map(_cpu_worker, range(len(self.lengths)), chunksize=8)
        return time.perf_counter() - t0

    def describe(self, dt=None):
        what = ("the reference's own speechpy feature.lmfe + processing.cmvnw (oracle/_ref)" if self.kind == "reference"
                else "oracle port of speechpy lmfe + cmvnw")
        tail = f", {dt:.1f} s wall" if dt is not None else ""
        return (f"fixed subset of {len(self.lengths)} cfg3-shaped utterances ({self.audio_h:.3f} audio-h) per step, "
                f"{what}, one single-threaded process per core, audio generated before the clock starts{tail}")

    def close(self):
        self.pool.close()
        self.pool.join()


# ---------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on all host cores; each
    step is one pass over a fixed bounded sample of the cfg3 workload."""
    if rank != 0:
        return
    arm = CpuArm(utts=args.cpu_utts)
    for _ in range(args.warmup):
        arm.step()
    t_total = 0.0
    for _ in range(args.steps):
        t_total += arm.step()
    arm.close()
    value = arm.audio_h * args.steps / t_total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
        "higher_is_better": True, "scaling": scaling_of(args), "vs_baseline": None, "dtype": "f64",
        "data": "synthetic 16 kHz audio, 16-bit PCM round(N(0,0.1)*32768) scaled by 1/32768 to float32 (what "
                "librosa.load returns), Philox-seeded",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
                         "per_core": value / arm.cores, "sample": arm.describe()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def scaling_of(args):
    return "strong"


def workload_config(args, gpus):
    return {"workload": "cfg3: ragged 4-20 s 16 kHz utterances (VoxCeleb1-dev shape, "
                        "clip(lognormal(ln 7.2 s, 0.5), 4, 20)), lmfe(40, 20ms/10ms, NFFT 512) + cmvnw(301, var)",
            "total_utterances": int(args.utts), "cfg3_total_utterances": CFG3_UTTS,
            "sharding": f"ONE corpus dealt to {gpus} rank(s) by batch.shard_indices (length-bucketed, balanced by "
                        f"samples), no collective in the data path",
            "l2": "inputs larger than L2 (no flush needed)"}


def make_device_corpus(torch, lengths, seed, device):
    # utterance starts on 32-byte boundaries (8 samples): the 64-byte runs the converters fetch then cover whole
    # sectors (measured at 20 000 utterances: features 4.24 ms with 16-byte starts, 3.92 ms with 32-byte starts)
    align = int(os.environ.get("SPB_BENCH_ALIGN", "8"))
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


KERNEL_NAMES = {0: "features", 1: "cmvnw_stage1", 2: "cmvnw_stage2", 4: "cmvnw_streaming", 5: "cmvnw_resident"}


def newest_ncu_traffic(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per ALGORITHMIC byte of the newest committed
    `ncu --set full` summary that carries the kernel (profiles/*_ncu_traffic.json, written by
    tools/ncu_summary.py from the .ncu-rep of the same command).  None when there is none."""
    pdir = os.path.join(ROOT, "profiles")
    best = None
    try:
        for name in sorted(os.listdir(pdir)):
            if name.endswith("_ncu_traffic.json"):
                with open(os.path.join(pdir, name)) as fh:
                    for rec in json.load(fh):
                        if kernel_substr in rec.get("kernel", ""):
                            best = dict(rec, source=f"profiles/{name}")
    except Exception:
        return None
    return best


def timed_steps(torch, dist, ctx, stream, step, steps, warmup, world, device, local_rank, sample_clocks):
    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    ctx.timing_enable(True)
    sampler = ClockSampler(local_rank) if sample_clocks else None
    if sampler:
        sampler.start()
    n0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        step()
    e1.record(stream)
    barrier()
    clocks = sampler.result() if sampler else None
    launches = ctx.launch_count() - n0
    ms = e0.elapsed_time(e1)
    k_ms = {name: ctx.timing_read(k) for k, name in KERNEL_NAMES.items()}
    ctx.timing_enable(False)
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), ms, k_ms, launches, clocks


def run_ours(args, rank, world, local_rank):
    from speaker_verification_b200 import synth
    # ---- CPU baseline first (rank 0, N=1 only), before CUDA is initialised in this process
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        arm = CpuArm(utts=args.cpu_utts)
        dt = arm.step()
        arm.close()
        rate = arm.audio_h / dt
        cpu_baseline = {"value": rate, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "per_core": rate / arm.cores,
                        "sample": arm.describe(dt)}

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
    from speaker_verification_b200 import batch
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- the workload: ONE corpus description, dealt to the ranks (strong scaling, no collective)
    U = args.utts
    lengths_all = synth.cfg3_lengths(U, seed=3)
    shards = batch.shard_indices(lengths_all, world)
    lengths = lengths_all[shards[rank]]
    per_rank_samples = [int(lengths_all[ix].sum()) for ix in shards]
    total_audio_h = float(lengths_all.sum()) / FS / 3600.0
    arena, starts = make_device_corpus(torch, lengths, 77 + rank, device)
    fb = batch.FeatureBatch(starts, lengths, device=local_rank)
    ctx = fb.ctx
    out = torch.empty((fb.total_frames, M), dtype=torch.float32, device=device)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=device)
    audio_h = float(lengths.sum()) / FS / 3600.0
    alg_bytes = algorithmic_bytes(lengths)

    def step():
        fb.lmfe_cmvnw_dev(arena.data_ptr(), arena.numel(), out.data_ptr())

    ms_max, ms, k_ms, launches, clocks = timed_steps(torch, dist, ctx, stream, step, args.steps, args.warmup, world,
                                                     device, local_rank, True)
    value = total_audio_h * args.steps / (ms_max / 1e3)

    # ---- e2e: host-pointer C-ABI call, pinned host buffers, H2D and D2H inside the timed region
    Ue = max(1, min(U // args.e2e_fraction, len(lengths)))
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
    # the device path and the host path must agree bit for bit on the shared prefix
    same = bool(torch.equal(out[:fbe.total_frames].cpu(), h_out))
    if not same:
        raise SystemExit("bench: the host-pointer (e2e) result differs from the device-resident result")
    e_total_frames = int(fbe.total_frames)
    fbe.close()
    del h_out

    # ---- the integer-scaled arena (signal.astype(float) of int16 PCM, |x| >= 2): the tensor-core kernel cannot
    # split such samples, the guarded FFT kernel redoes the launch group -- reported beside the headline
    fallback = None
    if rank == 0 and not args.no_extras:
        Uf = max(1, min(len(lengths), U // 16))
        f_end = int(starts[Uf - 1] + lengths[Uf - 1])
        fbf = batch.FeatureBatch(starts[:Uf], lengths[:Uf], device=local_rank)
        big = (arena[:f_end] * 32768.0).contiguous()
        f_out = torch.empty((fbf.total_frames, M), dtype=torch.float32, device=device)
        for _ in range(2):
            fbf.lmfe_cmvnw_dev(big.data_ptr(), big.numel(), f_out.data_ptr())
        fbf.ctx.sync()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(3):
            fbf.lmfe_cmvnw_dev(big.data_ptr(), big.numel(), f_out.data_ptr())
        f1.record(stream)
        fbf.ctx.sync()
        fallback = {"value": float(lengths[:Uf].sum()) / FS / 3600.0 * 3 / (f0.elapsed_time(f1) / 1e3), "unit": UNIT,
                    "what": f"first {Uf} utterances scaled by 32768 (int16-valued floats): tensor-core launch + "
                            f"flagged redo by features_fast512, device-resident"}
        fbf.close()
        del big, f_out

    # ---- weak scaling beside it (N > 1): every rank featurises a full-size corpus of its own
    weak = None
    if world > 1 and not args.no_weak:
        fb.close()
        del arena, out, fb
        torch.cuda.empty_cache()
        w_len = synth.cfg3_lengths(U, seed=3 + 1000 * rank)
        w_arena, w_starts = make_device_corpus(torch, w_len, 177 + rank, device)
        fbw = batch.FeatureBatch(w_starts, w_len, device=local_rank)
        w_out = torch.empty((fbw.total_frames, M), dtype=torch.float32, device=device)

        def w_step():
            fbw.lmfe_cmvnw_dev(w_arena.data_ptr(), w_arena.numel(), w_out.data_ptr())

        w_ms_max, _, _, _, _ = timed_steps(torch, dist, fbw.ctx, stream, w_step, args.steps, args.warmup, world, device,
                                           local_rank, False)
        w_audio = torch.tensor([float(w_len.sum()) / FS / 3600.0], dtype=torch.float64, device=device)
        dist.all_reduce(w_audio)
        weak = {"value": float(w_audio.item()) * args.steps / (w_ms_max / 1e3), "unit": UNIT,
                "ms_per_step": w_ms_max / args.steps, "utterances_per_gpu": int(U),
                "what": "every rank featurises its own full-size cfg3 corpus (different seeds), max over ranks"}
        fbw.close()
        del w_arena, w_out
        torch.cuda.empty_cache()

    # ---- config 5 beside it: the cosine score matrix on the tensor cores + EER; rows sharded over the ranks
    scoring_line = bench_scoring(torch, dist, ctx, rank, world, device)

    if rank == 0:
        peak, peak_src = peaks()
        feat_ms, feat_n = k_ms["features"]
        per_launch = max(feat_n // max(args.steps, 1), 1)
        achieved = alg_bytes * args.steps / (feat_ms / 1e3) / 1e9 if feat_ms > 0 else None
        step_gbs = alg_bytes * args.steps / (ms / 1e3) / 1e9
        tr = newest_ncu_traffic("features_tc2_kernel")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": scaling_of(args), "vs_baseline": None, "dtype": "f32 (f64 window accumulators in cmvnw)",
            "data": "synthetic 16 kHz audio, 16-bit PCM round(N(0,0.1)*32768) scaled by 1/32768 to float32 (what "
                    "librosa.load returns), generated on device (torch Philox), random lengths seeded",
            "config": workload_config(args, world),
            "shard_balance": {"samples_per_rank_min": min(per_rank_samples), "samples_per_rank_max": max(per_rank_samples),
                              "utterances_rank0": int(len(lengths))},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(e_end * 2),
                    "d2h_bytes_per_step": int(e_total_frames * M * 4),
                    "sample": f"first {Ue} utterances of each rank's shard ({e_audio_h:.2f} audio-h) per step, "
                              f"{e_steps} steps, pinned host buffers, int16 PCM arena in (the wav payload; scaled by "
                              f"1/32768 on the device like librosa.load), float32 features out, spb_lmfe_cmvnw_batch_dtype "
                              f"(H2D / kernels / D2H pipelined over chunks)",
                    "host_affinity": numa,
                    "float32_arena_value": e2e_f32, "float32_arena_h2d_bytes_per_step": int(e_end * 4),
                    "matches_device_path": same},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {
                "bound": "hbm", "kernel": "features_tc2_kernel (framing + folded hop-DFT on tcgen05 + |X|^2 + mel + log; "
                                          "timed with its flag memset, the float64 redo of flagged filters and the "
                                          "guarded no-op FFT launch)",
                "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None,
                "traffic": (tr["dram_bytes_per_alg_byte"] * alg_bytes / per_launch) if tr else None,
                "traffic_source": (f"{tr['source']}: dram__bytes_read.sum + dram__bytes_write.sum = "
                                   f"{tr['dram_bytes_per_alg_byte']:.3f} x algorithmic bytes in the ncu --set full capture "
                                   f"({tr.get('what', '')}), scaled to this run's launch size") if tr else None,
                "achieved_per_launch_bytes": alg_bytes / per_launch,
                "peak_source": peak_src,
                "algorithmic_bytes_per_step": alg_bytes,
                "kernel_ms_per_step": {k: v[0] / args.steps for k, v in k_ms.items() if v[1]},
                "launches_per_step": {k: v[1] // args.steps for k, v in k_ms.items() if v[1]},
                "whole_step_achieved": step_gbs, "whole_step_frac": step_gbs / peak,
                "nominal_8000_frac_whole_step": step_gbs / 8000.0,
            },
            "audio_hours_per_step": total_audio_h, "audio_hours_per_step_rank0": audio_h,
        }
        if weak is not None:
            line["weak_scaling"] = weak
        if fallback is not None:
            line["int16_scale_arena"] = fallback
        if scoring_line is not None:
            line["scoring"] = scoring_line
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_scoring(torch, dist, ctx, rank, world, device):
    """cfg5: 4 874 x 4 874 x 512 cosine matrix + EER of 37 720 trials.  Rank 0 times the single-GPU
    path; at N > 1 every rank also runs the row-sharded path (one NCCL broadcast of T, one small
    reduce of the trial scores), timed with the broadcast inside and outside the call."""
    from speaker_verification_b200 import scoring
    gen = torch.Generator(device=device)
    gen.manual_seed(5)
    m5 = n5 = 4874
    d5 = 512
    nt = 37720
    E5 = torch.randn((m5, d5), generator=gen, device=device)   # same seed on every rank: identical tensors
    T5 = torch.randn((n5, d5), generator=gen, device=device)
    ti5 = torch.randint(0, m5, (nt,), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
    tj5 = torch.randint(0, n5, (nt,), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
    lab5 = (torch.rand(nt, generator=gen, device=device) < 0.5).to(torch.int32)
    # target trials get a test embedding that is a noisy copy of the enrolment one: a non-degenerate EER
    tgt = torch.nonzero(lab5).flatten()
    T5[tj5[tgt].long()] = E5[ti5[tgt].long()] + 6.0 * torch.randn((tgt.numel(), d5), generator=gen, device=device)
    line = None
    single = None
    if rank == 0:
        S5 = torch.empty((m5, n5), dtype=torch.float32, device=device)
        for _ in range(3):
            scoring.cosine_scores_dev(E5, T5, S5)
        ctx.timing_enable(True)
        for _ in range(10):
            scoring.cosine_scores_dev(E5, T5, S5)
        g_ms, g_n = ctx.timing_read(3)
        ctx.timing_enable(False)
        single = scoring.trial_scores_dev(S5, ti5, tj5)
        scoring.eer_auc_dev(single, lab5)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eer5, auc5 = scoring.eer_auc_dev(single, lab5)
        eer_ms = (time.perf_counter() - t0) * 1e3
        t0 = time.perf_counter()
        for _ in range(10):
            scoring.cosine_scores_dev(E5, T5, S5)
            tr = scoring.trial_scores_dev(S5, ti5, tj5)
            scoring.eer_auc_dev(tr, lab5)
        whole_ms = (time.perf_counter() - t0) * 1e2
        per = g_ms / max(g_n, 1) / 1e3
        flops = 2.0 * m5 * n5 * d5
        tf_peak = None
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                tf_peak = float(json.load(fh)["bf16_tflops"]) / 2.0  # TF32 runs at half the bf16 rate
        except Exception:
            pass
        line = {"workload": "cfg5: cosine scores 4874 x 4874 x 512 (3xTF32 on tcgen05) + EER of 37720 trials",
                "gemm_us": per * 1e6, "useful_tflops": flops / per / 1e12,
                "issued_tf32_tflops": 3 * flops / per / 1e12,
                "tensor_frac_of_measured_bf16_over_2": (3 * flops / per / 1e12 / tf_peak) if tf_peak else None,
                "eer_call_ms": eer_ms, "single_gpu_scores_trials_eer_ms_per_call": whole_ms,
                "eer": round(eer5, 4), "auc": auc5}
        del S5
    if world > 1:
        b = scoring.row_partition(m5, world)
        r0, r1 = b[rank], b[rank + 1]
        E_blk = E5[r0:r1].contiguous()
        T_dev = T5 if rank == 0 else torch.zeros_like(T5)

        plan = scoring.ShardedTrials(r0, r1 - r0, ti5, tj5, lab5)  # trial ownership, computed once

        def timed(resident):
            res = None
            for _ in range(3):
                res = plan.score(E_blk, T_dev, t_resident=resident)
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            K = 10
            t0 = time.perf_counter()
            for _ in range(K):
                res = plan.score(E_blk, T_dev, t_resident=resident)
            torch.cuda.synchronize()
            t = torch.tensor([(time.perf_counter() - t0) / K * 1e3], dtype=torch.float64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item()), res

        ms_in, res = timed(False)       # T broadcast from rank 0 inside every call
        ms_out, res2 = timed(True)      # T already resident on every rank (broadcast once, outside)
        if rank == 0:
            eer_s, auc_s, full = res
            line["sharded"] = {
                "ranks": world, "rows_per_rank": [b[i + 1] - b[i] for i in range(world)],
                "ms_per_call_broadcast_inside": ms_in, "ms_per_call_T_resident": ms_out,
                "timing": "host wall clock around the whole call (NCCL broadcast of T 9.98 MB, normalise + GEMM of the "
                          "row block, trial gather, NCCL reduce of 37 720 scores to rank 0, EER sort on rank 0), max over ranks",
                "scores_equal_single_gpu": bool(torch.equal(full, single)),
                "scores_equal_T_resident": bool(torch.equal(res2[2], single)),
                "eer": round(eer_s, 4), "eer_equal_single_gpu_4dp": round(eer_s, 4) == line["eer"],
                "limiter": "per-call latency of the two NCCL collectives and of the host-synchronous steps; the GEMM of a "
                           "610-row block is ~20 us of work"}
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--utts", type=int, default=int(os.environ.get("SPB_BENCH_UTTS", CFG3_UTTS)),
                    help="utterances of the corpus (default: all of config 3); dealt to the ranks")
    ap.add_argument("--e2e-fraction", type=int, default=16,
                    help="the e2e leg uses the first 1/this of the corpus per rank")
    ap.add_argument("--cpu-utts", type=int, default=int(os.environ.get("SPB_BENCH_CPU_UTTS", CPU_SAMPLE_UTTS)),
                    help="fixed subset of cfg3 utterances the CPU legs time per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-weak", action="store_true", help="skip the weak-scaling leg at N > 1")
    ap.add_argument("--no-extras", action="store_true", help="skip the int16-scale arena leg")
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
