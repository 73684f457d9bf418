#!/usr/bin/env python
"""Benchmark of the MFCC hot path (BASELINE.json: MFCC frames/s per B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step is one pass of the fused integer-mode kernel over the configs[1] workload of
BASELINE.json: 10 000 synthetic 60 s int16 clips (960 000 samples each, 5645 frames per
clip, 32 filters, 32 coefficients) per GPU.  `value` is frames/s with the clips
resident in HBM; `e2e` is the same metric through MFCC.run_host() (the C-ABI call with
host buffers): pinned host samples in, cepstra back in host memory, copies inside
the timed region.  One process per GPU; no collective on the data path (clips are
independent), NCCL only for the barriers and the max-over-ranks of the times.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_SAMPLES = 960000          # 60 s at 16 kHz
N_CLIPS = 10000             # BASELINE.json configs[1]
NFILT = NCEP = 32
BYTES_PER_FRAME = 170 * 2 + NCEP * 2    # 340 B read + 64 B written
SEED0 = 20260000
WORKLOAD = ("C2: %d synthetic 60 s int16 clips per GPU (960000 samples, 5645 frames each), 512/170 frames, "
            "32 mel, 32 cepstra, integer (bit-exact) mode")


def synth_clips(n_clips, n_samples, device, first_clip=0, out=None):
    """The synthetic mixture of SURVEY.md 8(d) C2, by global clip index i:
    i%4 == 0 uniform full-scale noise (drives the 16-bit mel wrap), 1 Gaussian
    sigma 2000, 2 three sinusoids of amplitude 8000 + Gaussian sigma 50, 3 Gaussian
    sigma 2000 gated in 0.5 s blocks (silence -> log2(0)).  Seeded per clip."""
    import torch
    if out is None:
        out = torch.empty((n_clips, n_samples), dtype=torch.int16, device=device)
    g = torch.Generator(device=device)
    t = torch.arange(n_samples, device=device, dtype=torch.float32)
    for j in range(n_clips):
        i = first_clip + j
        g.manual_seed(SEED0 + i)
        c = i % 4
        if c == 0:
            out[j] = torch.randint(-32768, 32768, (n_samples,), generator=g, device=device,
                                   dtype=torch.int32).to(torch.int16)
            continue
        if c == 2:
            f = torch.rand(3, generator=g, device=device) * 7900.0 + 50.0
            ph = torch.rand(3, generator=g, device=device) * 6.2831853
            v = 8000.0 * torch.sin(t[None, :] * (6.2831853 / 16000.0) * f[:, None] + ph[:, None]).sum(0)
            v = v + 50.0 * torch.randn(n_samples, generator=g, device=device)
        else:
            v = 2000.0 * torch.randn(n_samples, generator=g, device=device)
            if c == 3:
                v = v * ((torch.arange(n_samples, device=device) // 8000) % 2).to(torch.float32)
        out[j] = v.clamp_(-32768, 32767).to(torch.int16)
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def mark(self):
        """Rows from here on belong to the timed region (nvidia-smi is started earlier:
        its start-up stalls the driver for ~100 ms and must not fall inside it)."""
        self.first = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        rows = self.rows[getattr(self, "first", 0):] or self.rows
        for r in rows:
            p = [s.strip() for s in r.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1]))
                mx = float(p[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_oracle_rate(n_clips, threads, repeats=1, fast=True):
    """Frames/s of the CPU restatement of the RTL on `n_clips` clips of the same workload:
    fast=True is oracle/mfcc_ref_fast.c built here with gcc -O3 -march=native -fopenmp (the
    timed baseline, bit-identical to the checker: tests/test_oracle.py), fast=False the
    checker oracle/mfcc_ref.c (-O2).  This is the only place bench.py runs oracle/."""
    import numpy as np
    import torch
    from oracle import ref_c as RC
    cfg = RC.RefConfig(nfilters=NFILT, nceptrums=NCEP)
    x = synth_clips(n_clips, N_SAMPLES, torch.device("cpu")).numpy()
    frames = n_clips * cfg.num_frames(N_SAMPLES)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        RC.run(cfg, x, threads=threads, fast=fast)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return frames / best, frames, best


def run_reference(args, rank, world):
    """--impl reference: the reference's integer pipeline on the host cores.  The
    reference hot path is nMigen RTL and cannot run here (no nmigen, no simulator), so
    this arm times the CPU restatement of it ("port": oracle/mfcc_ref_fast.c, gcc -O3
    -march=native -fopenmp, built on this machine, bit-identical to the checker) with all
    host threads, on a bounded sample of the same workload per step."""
    if rank != 0:
        return
    from oracle import ref_c as RC
    import numpy as np
    import torch
    threads = RC.max_threads()
    n_clips = min(128, 4 * threads)
    cfg = RC.RefConfig(nfilters=NFILT, nceptrums=NCEP)
    x = synth_clips(n_clips, N_SAMPLES, torch.device("cpu")).numpy()
    frames = n_clips * cfg.num_frames(N_SAMPLES)
    RC.build_fast()
    for _ in range(args.warmup):
        RC.run(cfg, x[: max(1, n_clips // 8)], threads=threads, fast=True)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        RC.run(cfg, x, threads=threads, fast=True)
    dt = (time.perf_counter() - t0) / args.steps
    v = frames / dt
    sample = "%d clips x 60 s (%d frames) per step" % (n_clips, frames)
    print(json.dumps({
        "impl": "reference", "metric": "mfcc_frames_per_sec", "value": v, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16",
        "data": "synthetic", "audio_sec_per_sec": v * 170 / 16000,
        # the same workload as the GPU arm's line; each timed step is a bounded sample of it
        "config": {"workload": WORKLOAD % args.clips, "clips_per_gpu": args.clips,
                   "frames_per_step_per_gpu": args.clips * cfg.num_frames(N_SAMPLES), "sample": sample},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample,
                         "build_flags": "gcc " + RC.FAST_FLAGS, "source": "oracle/mfcc_ref_fast.c"},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def stream_latency(op, dev, n_streams=4096, steps=10000, warmup=200, groups=1, zero_copy=False):
    """BASELINE.json configs[4]: `n_streams` live streams, one
    170-sample hop per stream per step, the step captured once in a CUDA graph (H2D of the
    hops from pinned memory, the streaming kernel, D2H of the cepstrum columns).  Returns
    p50/p99 of the device time per step (CUDA events around the replay) and of the host
    wall time per step (perf_counter around replay + stream synchronize).

    groups > 1: the streams are cut into that many groups, each a branch of the graph on its
    own CUDA stream (copy-in, kernel, copy-out of its rows).  Measured SLOWER on B200 (2
    groups 66-74 us, 4 groups 79 us against 63 us): per-copy latency outweighs the overlap.
    zero_copy=True: the kernel reads the pinned hops and writes the pinned columns itself
    (same p50, tighter p99); zero_copy="out": hops by DMA, columns written by the kernel
    straight into pinned host memory (57 us p50: one copy operation less)."""
    import torch
    hop_h = torch.randint(-3000, 3000, (n_streams, 170), dtype=torch.int16).pin_memory()
    out_h = torch.empty((n_streams, op.nceptrums), dtype=op.out_dtype).pin_memory()
    hop_d = torch.empty((n_streams, 170), dtype=torch.int16, device=dev)
    out_d = torch.empty((n_streams, op.nceptrums), dtype=op.out_dtype, device=dev)
    op.reset(n_streams)
    st = torch.cuda.Stream(device=dev)
    side = [torch.cuda.Stream(device=dev) for _ in range(groups)] if groups > 1 else []
    cuts = [n_streams * i // groups for i in range(groups + 1)]

    def one_step():
        if zero_copy == "out":  # DMA the hops in, the kernel writes the columns to pinned host memory itself
            hop_d.copy_(hop_h, non_blocking=True)
            op.step(hop_d, out=out_h)
            return
        if zero_copy:           # the kernel reads the pinned hops and writes the pinned columns itself
            op.step(hop_h, out=out_h)
            return
        if groups == 1:
            hop_d.copy_(hop_h, non_blocking=True)
            op.step(hop_d, out=out_d)
            out_h.copy_(out_d, non_blocking=True)
            return
        for i, s in enumerate(side):
            a, b = cuts[i], cuts[i + 1]
            s.wait_stream(st)
            with torch.cuda.stream(s):
                hop_d[a:b].copy_(hop_h[a:b], non_blocking=True)
                op.step(hop_d[a:b], out=out_d[a:b], first_stream=a)
                out_h[a:b].copy_(out_d[a:b], non_blocking=True)
        for s in side:
            st.wait_stream(s)

    with torch.cuda.stream(st):
        for _ in range(3):
            one_step()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            one_step()
        for _ in range(warmup):
            g.replay()
        st.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        wall = []
        ev[0].record(st)
        for i in range(steps):
            t0 = time.perf_counter()
            g.replay()
            ev[i + 1].record(st)
            st.synchronize()
            wall.append(time.perf_counter() - t0)
    dev_us = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(steps))
    wall_us = sorted(w * 1e6 for w in wall)
    pick = lambda a, q: a[min(len(a) - 1, int(q * len(a)))]
    return {"workload": "C5: %d live streams, one 170-sample hop per stream per step, CUDA graph "
                        "(H2D hops + step kernel + D2H columns)" % n_streams,
            "groups": groups, "zero_copy": zero_copy, "steps": steps, "device_us_p50": pick(dev_us, 0.5), "device_us_p99": pick(dev_us, 0.99),
            "host_us_p50": pick(wall_us, 0.5), "host_us_p99": pick(wall_us, 0.99),
            "realtime_budget_us": 170 / 16000 * 1e6,
            "frames_per_sec_at_p50": n_streams / (pick(wall_us, 0.5) * 1e-6)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--clips", type=int, default=N_CLIPS, help="clips per GPU (default: configs[1])")
    ap.add_argument("--e2e-clips", type=int, default=2048)
    ap.add_argument("--e2e-chunk-mb", type=float, default=None, help="run_host chunk size (MiB of samples)")
    ap.add_argument("--e2e-slots", type=int, default=None, help="run_host chunks in flight (2..8)")
    ap.add_argument("--e2e-wc", action="store_true", help="write-combined pinned sample buffer")
    ap.add_argument("--e2e-sweep", action="store_true",
                    help="also time run_host over a grid of chunk sizes / slots (extra key e2e.sweep)")
    ap.add_argument("--no-c4", action="store_true", help="skip the C4 (10^6 audio-seconds, strong scaling) leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-numa", action="store_true",
                    help="do not pin the process to the GPU's local CPU cores (NUMA node)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the fp32-mode and streaming-latency figures (rank 0, N=1 only)")
    ap.add_argument("--stream-steps", type=int, default=10000)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    import mfcc_b200
    from mfcc_b200 import _lib
    if args.warmup < 3:
        args.warmup = 3
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from mfcc_b200 import shard
    all_cores = os.sched_getaffinity(0)
    numa_cores = shard.bind_to_gpu_numa(dev) if not args.no_numa else []
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    op = mfcc_b200.MFCC(nfilters=NFILT, nceptrums=NCEP, device=dev)
    K = op.num_frames(N_SAMPLES)
    x = torch.empty((args.clips, N_SAMPLES), dtype=torch.int16, device=dev)
    synth_clips(args.clips, N_SAMPLES, dev, first_clip=rank * args.clips, out=x)
    out = torch.empty((args.clips, K, NCEP), dtype=torch.int16, device=dev)
    frames_rank = args.clips * K

    sampler = ClockSampler(local).start()
    for _ in range(args.warmup):
        op(x, out=out)
    barrier()
    time.sleep(0.3)
    sampler.mark()
    l0 = _lib.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for s in range(args.steps):
        op(x, out=out)          # one kernel launch on torch's current stream
        ev[s + 1].record()
    barrier()
    launches = _lib.launch_count() - l0
    clocks = sampler.stop()
    t_total = ev[0].elapsed_time(ev[-1]) * 1e-3
    kern = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e-3 for i in range(args.steps))
    t_kernel_avg = sum(kern) / len(kern)
    tt = torch.tensor([t_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_max = float(tt.item())
    value = frames_rank * world * args.steps / t_max

    # ---- end to end: pinned host buffers through the C-ABI host entry point
    from mfcc_b200.mfcc import pinned_empty
    ne = min(args.e2e_clips, args.clips)
    xh = pinned_empty((ne, N_SAMPLES), torch.int16, write_combined=args.e2e_wc)
    xh.copy_(x[:ne])
    oh = pinned_empty((ne, K, NCEP), torch.int16)
    e_steps = max(1, min(args.steps, 5))
    op.host_config(args.e2e_chunk_mb, args.e2e_slots)

    def timed_host(fn, steps):
        """Wall time of `steps` synchronous calls, max over ranks (fn returns with the device idle)."""
        fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    t_e2e = timed_host(lambda: op.run_host(xh, out=oh), e_steps)
    e2e_value = ne * K * world / t_e2e
    e2e_ok = bool(torch.equal(oh.to(dev), out[:ne]))

    # ---- copy-only ceiling of the same end-to-end step: the same pinned buffers and byte
    # counts, plain cudaMemcpyAsync host->device and device->host on two streams, no kernel
    cs_in, cs_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    xd, od = x[:ne], out[:ne]                       # same contents as xh / oh: nothing changes

    chunk_clips = max(1, int((args.e2e_chunk_mb or 16) * (1 << 20)) // (N_SAMPLES * 2))

    def copy_both(h2d=True, d2h=True, chunked=False):
        step = chunk_clips if chunked else ne
        for a in range(0, ne, step):
            b = min(ne, a + step)
            if h2d:
                with torch.cuda.stream(cs_in):
                    xd[a:b].copy_(xh[a:b], non_blocking=True)
            if d2h:
                with torch.cuda.stream(cs_out):
                    oh[a:b].copy_(od[a:b], non_blocking=True)
        cs_in.synchronize()
        cs_out.synchronize()

    h2d_b, d2h_b = ne * N_SAMPLES * 2, ne * K * NCEP * 2
    t_mono = timed_host(copy_both, e_steps)
    t_chunk = timed_host(lambda: copy_both(chunked=True), e_steps)
    t_both = min(t_mono, t_chunk)
    t_h2d = timed_host(lambda: copy_both(d2h=False), e_steps)
    t_d2h = timed_host(lambda: copy_both(h2d=False), e_steps)
    copy_ceiling = {"value": ne * K * world / t_both, "unit": "frames/s",
                    "what": "same pinned buffers and bytes as e2e, cudaMemcpyAsync H2D + D2H on two streams, "
                            "no kernel; all ranks at once, max over ranks; the better of one copy per direction "
                            "and the e2e path's chunks",
                    "ms_per_step": t_both * 1e3, "ms_one_copy_per_direction": t_mono * 1e3,
                    "ms_chunked": t_chunk * 1e3, "chunk_clips": chunk_clips,
                    "h2d_gbs_all_gpus": h2d_b * world / t_both / 1e9, "d2h_gbs_all_gpus": d2h_b * world / t_both / 1e9,
                    "h2d_alone_gbs_all_gpus": h2d_b * world / t_h2d / 1e9,
                    "d2h_alone_gbs_all_gpus": d2h_b * world / t_d2h / 1e9,
                    "h2d_alone_gbs_per_gpu": h2d_b / t_h2d / 1e9,
                    "e2e_over_ceiling": e2e_value / (ne * K * world / t_both)}
    e2e_sweep = None
    if args.e2e_sweep:
        e2e_sweep = []
        for mb, sl in ((8, 4), (16, 4), (32, 2), (32, 3), (32, 4), (32, 8), (64, 4), (128, 4), (256, 3)):
            op.host_config(mb, sl)
            ts = timed_host(lambda: op.run_host(xh, out=oh), max(2, e_steps // 2))
            e2e_sweep.append({"chunk_mb": mb, "slots": sl, "value": ne * K * world / ts})
        op.host_config(args.e2e_chunk_mb or 16, args.e2e_slots or 4)
    del xd, od, xh, oh

    # ---- C4 (BASELINE.json configs[3]): 10^6 audio-seconds = 16 667 clips of 60 s, rank r of G
    # owns clips r::G (strong scaling: the total is fixed, the per-GPU share shrinks with G)
    c4 = None
    if not args.no_c4:
        C4_CLIPS = 16667
        mine = list(range(rank, C4_CLIPS, world))
        n4 = len(mine)
        if n4 <= args.clips:
            x4, o4 = x[:n4], out[:n4]               # the device never looks at the seed: reuse the resident clips
        else:
            x4 = torch.empty((n4, N_SAMPLES), dtype=torch.int16, device=dev)
            x4[:args.clips] = x
            synth_clips(n4 - args.clips, N_SAMPLES, dev, first_clip=(rank + world) * args.clips, out=x4[args.clips:])
            o4 = torch.empty((n4, K, NCEP), dtype=torch.int16, device=dev)
        c_steps = max(3, min(args.steps, 10))
        for _ in range(3):
            op(x4, out=o4)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(c_steps):
            op(x4, out=o4)
        c1.record()
        barrier()
        t4 = torch.tensor([c0.elapsed_time(c1) * 1e-3 / c_steps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t4, op=dist.ReduceOp.MAX)
        t4 = float(t4.item())
        c4 = {"workload": "C4: 10^6 audio-seconds = %d clips x 60 s (%d frames) over %d GPU(s), rank r owns clips r::G"
                          % (C4_CLIPS, C4_CLIPS * K, world),
              "scaling": "strong", "clips_this_rank": n4, "steps": c_steps, "ms_per_pass": t4 * 1e3,
              "value": C4_CLIPS * K / t4, "unit": "frames/s", "audio_sec_per_sec": C4_CLIPS * 60.0 / t4}
        try:
            ref1 = json.load(open(os.path.join(ROOT, "profiles", "r02_c4_n1.json")))["ms_per_pass"]
            c4["efficiency_vs_recorded_n1"] = ref1 / (t4 * 1e3) / world
            c4["recorded_n1_ms_per_pass"] = ref1
        except Exception:
            pass
        if n4 > args.clips:
            del x4, o4
            torch.cuda.empty_cache()

    extras = {}
    if world == 1 and not args.no_extras:
        # Extra legs of the N=1 line.  Each is guarded: a failure is recorded in the line
        # and can never cost the headline numbers measured above.
        fsteps = max(3, min(args.steps, 5))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def timed(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(fsteps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) * 1e-3 / fsteps

        def leg_fp32():
            # fp32 mode on the same clips (config C3's throughput leg; parity is in tests/)
            opf = mfcc_b200.MFCC(nfilters=NFILT, nceptrums=NCEP, mode="fp32", mel_floor=1e-3, device=dev)
            outf = torch.empty((args.clips, K, NCEP), dtype=torch.float32, device=dev)
            tf = timed(lambda: opf(x, out=outf))
            return {"value": frames_rank / tf, "unit": "frames/s", "ms_per_step": tf * 1e3,
                    "steps": fsteps, "bytes_per_frame": 170 * 2 + NCEP * 4,
                    "achieved_gbs": frames_rank * (170 * 2 + NCEP * 4) / tf / 1e9,
                    "kernel": "mfcc_f32_fused<32>", "finite": bool(torch.isfinite(outf[:64]).all())}

        def leg_librosa():
            # the librosa feature set of software/genlibrosa.py (BASELINE.json configs[2]'s
            # comparison arm): 2048-point frames, 128 mel filters, dB floor, 32 coefficients
            opl = mfcc_b200.LibrosaMFCC(sr=16000, hop_length=170, n_mfcc=32, device=dev)
            nl = min(args.clips, 4096)
            xl = x[:nl]
            Kl = opl.num_frames(xl.shape[1])
            outl = torch.empty((nl, Kl, 32), dtype=torch.float32, device=dev)
            tl = timed(lambda: opl(xl, out=outl))
            return {"value": nl * Kl / tl, "unit": "frames/s", "ms_per_step": tl * 1e3,
                    "clips": nl, "frames_per_clip": Kl, "steps": fsteps,
                    "kernels": ["lr_logmel", "lr_dct_c"], "finite": bool(torch.isfinite(outl[:64]).all())}

        def leg_stream_fp32():
            opf = mfcc_b200.MFCC(nfilters=NFILT, nceptrums=NCEP, mode="fp32", mel_floor=1e-3, device=dev)
            return stream_latency(opf, dev, 4096, max(1000, args.stream_steps // 5))

        for name, leg in (("fp32", leg_fp32), ("librosa_set", leg_librosa),
                          ("streaming", lambda: stream_latency(op, dev, 4096, args.stream_steps)),
                          # same step, but the kernel writes the columns straight into pinned host memory
                          ("streaming_hostwrite", lambda: stream_latency(op, dev, 4096, max(1000, args.stream_steps // 2),
                                                                        zero_copy="out")),
                          # the kernel reads the pinned hops and writes the pinned columns over PCIe itself: one launch
                          ("streaming_zerocopy", lambda: stream_latency(op, dev, 4096, max(1000, args.stream_steps // 2),
                                                                       zero_copy=True)),
                          ("streaming_fp32", leg_stream_fp32)):
            try:
                extras[name] = leg()
            except Exception as e:      # noqa: BLE001 - reported, never fatal
                extras[name] = {"error": "%s: %s" % (type(e).__name__, e)}
            torch.cuda.empty_cache()

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = frames_rank * BYTES_PER_FRAME / t_kernel_avg / 1e9
        line = {
            "metric": "mfcc_frames_per_sec", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_max / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16",
            "data": "synthetic", "audio_sec_per_sec": value * 170 / 16000,
            "config": {"workload": WORKLOAD % args.clips,
                       "clips_per_gpu": args.clips, "frames_per_step_per_gpu": frames_rank,
                       "l2": "inputs (%.1f GB per step) larger than L2, no flush needed"
                             % (args.clips * N_SAMPLES * 2 / 1e9),
                       "e2e_clips": ne, "e2e_steps": e_steps, "e2e_matches_device_path": e2e_ok,
                       "host_cores_bound_to_gpu_numa": len(numa_cores)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                         "kernel": "mfcc_int_fused<32>", "kernel_ms": t_kernel_avg * 1e3,
                         "bytes_per_frame": BYTES_PER_FRAME,
                         "note": "integer-issue bound, not HBM bound: see DESIGN.md"},
            "e2e": {"value": e2e_value, "unit": "frames/s",
                    "h2d_bytes_per_step": ne * N_SAMPLES * 2, "d2h_bytes_per_step": ne * K * NCEP * 2,
                    "ms_per_step": t_e2e * 1e3, "clips_per_gpu": ne, "steps": e_steps,
                    "h2d_gbs_all_gpus": ne * N_SAMPLES * 2 * world / t_e2e / 1e9,
                    "pipeline": "mfcc_b200_run_host: %s MiB chunks, %s in flight, 3 streams (H2D / kernel / D2H), "
                                "1-D copies" % (args.e2e_chunk_mb or 16, args.e2e_slots or 4),
                    "write_combined_source": bool(args.e2e_wc),
                    "copy_ceiling": copy_ceiling},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        # instruction-issue view of the same kernel: warp-instructions per frame and DRAM bytes
        # per frame are constants of the build, measured once with ncu (profiles/collect.py);
        # the peak is the mixed IMAD/ALU issue rate profiles/micro/pipe_bench.cu measured
        try:
            import glob
            iss_path = sorted(glob.glob(os.path.join(ROOT, "profiles", "r[0-9][0-9]_issue.json")))[-1]   # latest round
            iss = json.load(open(iss_path))
            k = iss["mfcc_int_fused"]
            lane_rate = k["warp_inst_per_frame"] * 32 * frames_rank / t_kernel_avg
            pk = iss["pipe_bench"]["mixed_lane_inst_per_s"]
            line["roofline"]["traffic"] = k["dram_bytes_per_frame"] * frames_rank
            line["roofline"]["traffic_note"] = ("%.1f B/frame from one ncu --set full capture of a 300-clip "
                                                "launch, scaled to this launch's frames" % k["dram_bytes_per_frame"])
            line["roofline"]["issue"] = {"achieved": lane_rate, "peak": pk, "unit": "lane-instr/s",
                                         "frac": lane_rate / pk, "warp_inst_per_frame": k["warp_inst_per_frame"],
                                         "peak_source": "profiles/micro/pipe_bench.cu IMAD+SHF 1:1 on this GPU model",
                                         "counts_from": os.path.basename(iss_path)}
        except Exception:
            pass
        if e2e_sweep:
            line["e2e"]["sweep"] = e2e_sweep
        if c4:
            line["c4"] = c4
        if extras:
            if "achieved_gbs" in extras.get("fp32", {}):
                extras["fp32"]["hbm_frac"] = extras["fp32"]["achieved_gbs"] / peak
            line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            try:
                os.sched_setaffinity(0, all_cores)      # the CPU baseline gets every host core back
                from oracle import ref_c as RC
                thr = RC.max_threads()
                ncl = min(256, 16 * thr)        # ~20 core-seconds of CPU work
                v, fr, dt = cpu_oracle_rate(ncl, thr, repeats=2)
                v1, fr1, dt1 = cpu_oracle_rate(16, 1)       # also a single-thread figure
                vc, frc, dtc = cpu_oracle_rate(min(ncl, 4 * thr), thr, fast=False)
                line["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": thr, "kind": "port",
                                        "sample": "%d clips x 60 s (%d frames), %.1f s" % (ncl, fr, dt),
                                        "build_flags": "gcc " + RC.FAST_FLAGS, "source": "oracle/mfcc_ref_fast.c",
                                        "single_thread": {"value": v1, "sample": "16 clips x 60 s, %.1f s" % dt1},
                                        "checker_build": {"value": vc, "build_flags": "gcc -O2 -fopenmp",
                                                          "source": "oracle/mfcc_ref.c", "cores": thr}}
                # the oracle as the checker of the timed launch itself: eight clips spread over the
                # full-size batch (all four classes), bit for bit against the GPU result
                idx = sorted({i for i in (0, 1, 2, 3, args.clips // 4 - 1, args.clips // 2 + 2, 3 * args.clips // 4 + 1,
                                          args.clips - 1) if 0 <= i < args.clips})
                it = torch.tensor(idx, device=dev)
                want = RC.run(RC.RefConfig(nfilters=NFILT, nceptrums=NCEP), x[it].cpu().numpy(), threads=thr)
                got = out[it].cpu().numpy()
                line["cpu_baseline"]["spot_check"] = {"clips": idx, "frames": int(want.shape[0] * want.shape[1]),
                                                      "bit_exact": bool((got == want).all())}
            except Exception as e:      # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": 0, "kind": "port",
                                        "sample": "failed: %s: %s" % (type(e).__name__, e)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
