#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: 1080p ENCODE frames/s with a .bit byte-identical to the reference encoder's.

Workload (BASELINE.json configs[1]): `config_LDB_medium_complexity.txt`, 1920x1080 8-bit synthetic (seeded hard clip: two
texture layers drifting against each other, a patch moving on its own, sensor noise), qp 32.  The encoder is the
reference's Thorenc with every block decision of encode_frame on the B200 (integration/_build/Thorenc_gpu ->
thorcuda_enc_frame + thorcuda_loopfilter_frame; entropy coding = the reference's own write_block on the host).

An LDB stream is one unbroken P chain and one frame's superblock wavefront keeps at most 15 of the 148 SMs busy
(DESIGN.md 4.10), so the unit of parallelism is the INDEPENDENT STREAM: `--streams` (36) encoder processes per GPU
share it through CUDA MPS.  A step = one frame of every stream.

  value : frames/s from the CUDA-event time of the frames' device work (frame upload .. decisions .. loop filter .. download),
          max over streams and ranks -- what the device sustains.
  e2e   : frames/s from the encoders' own clocks over the same frames: file read from tmpfs, pinned-host upload, device work,
          download of the decided trees + coefficients + reconstruction, entropy coding, bitstream write.
  cpu_baseline / --impl reference : the UNMODIFIED reference encoder (oracle/_ref/Thorenc, SSE4 build) on the same clip and
          configuration: one process (cpu_baseline, cores 1) / one process per host thread (--impl reference).
  encoder_e2e / decoder_e2e : one stream, stock binary vs hooked binary, outputs compared byte for byte.
  segments : ONE RA stream cut into closed-GOP segments, sharded over the ranks, joined and checked with the stock decoder.

One rank per GPU under torchrun; ranks exchange nothing but the start barrier and the MAX of their times.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "1080p encode frames/s (bit-identical .bit)"
REF_ENC = os.path.join(ROOT, "oracle", "_ref", "Thorenc")
REF_DEC = os.path.join(ROOT, "oracle", "_ref", "Thordec")
GPU_ENC = os.path.join(ROOT, "integration", "_build", "Thorenc_gpu")
GPU_DEC = os.path.join(ROOT, "integration", "_build", "Thordec_gpu")


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def shm_dir():
    return "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None


def enc_cmd(exe, cfg, yuv, w, h, frames, qp, prefix, rec=True):
    from thorref import THOR_CONFIGS          # pinned to the reference's config_*.txt by tests/test_hooks.py
    cmd = [exe, "-if", yuv, "-of", prefix + ".bit", "-stat", prefix + "_stat.txt", "-width", str(w), "-height", str(h), "-n", str(frames),
           "-f", "60", "-qp", str(qp), "-snrcalc", "0"] + THOR_CONFIGS[cfg]
    if rec:
        cmd += ["-rf", prefix + "_rec.yuv"]
    return cmd


class ClockSampler:
    """nvidia-smi clocks / throttle reasons of one GPU while the timed region runs"""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.stop_flag, self.thread = index, [], threading.Event(), None

    def _run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:                                   # noqa: BLE001
                pass
            self.stop_flag.wait(0.25)

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        self.stop_flag.set()
        if self.thread:
            self.thread.join(timeout=6)
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        mx = max((float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------- MPS ----
class Mps:
    """CUDA MPS control daemon for the duration of the run: without it the contexts of the encoder processes time-slice the
    GPU instead of sharing its SMs (measured, profiles/r2/: 8 processes 1.7 frames/s without, 12.4 with)."""

    def __init__(self, enable):
        self.on, self.env = False, {}
        self.enable = enable and shutil.which("nvidia-cuda-mps-control") is not None

    def start(self):
        if not self.enable:
            return self
        d = getattr(self, "env_dir", None) or tempfile.mkdtemp(prefix="thor_mps_")
        self.env = {"CUDA_MPS_PIPE_DIRECTORY": os.path.join(d, "pipe"), "CUDA_MPS_LOG_DIRECTORY": os.path.join(d, "log")}
        for v in self.env.values():
            os.makedirs(v, exist_ok=True)
        r = subprocess.run(["nvidia-cuda-mps-control", "-d"], env=dict(os.environ, **self.env), capture_output=True, text=True)
        self.on = r.returncode == 0
        if not self.on:
            self.env = {}
        time.sleep(0.5)
        return self

    def stop(self):
        if self.on:
            subprocess.run(["nvidia-cuda-mps-control"], input="quit\n", env=dict(os.environ, **self.env), capture_output=True, text=True, timeout=60)
            self.on = False


# ---------------------------------------------------------------------- encoder streams ----
def make_clip(path, w, h, frames, seed=42):
    from thorref import write_hard_clip
    if not os.path.exists(path):
        write_hard_clip(path, w, h, frames, seed, noise=5.0)
    return path


def run_streams(exe, cfg, yuv, w, h, frames, qp, nstreams, workdir, env, tag):
    """nstreams concurrent encoder processes on the same clip; returns per-stream dicts (wall seconds, frame log, outputs)"""
    procs = []
    t0 = time.perf_counter()
    for i in range(nstreams):
        pre = os.path.join(workdir, f"{tag}{i}")
        e = dict(os.environ, **env, THORCUDA_HOOK_TIMING="1", THORCUDA_HOOK_FRAMELOG=pre + ".frames", THORCUDA_BLOCKING_SYNC="1")
        procs.append((pre, subprocess.Popen(enc_cmd(exe, cfg, yuv, w, h, frames, qp, pre, rec=(i == 0)), stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, env=e)))
    out = []
    for pre, p in procs:
        _, err = p.communicate(timeout=3000)
        if p.returncode != 0:
            raise RuntimeError(f"{exe} failed ({p.returncode}): {err[-1500:]}")
        log = []
        if os.path.exists(pre + ".frames"):
            log = [[float(x) for x in ln.split()] for ln in open(pre + ".frames") if ln.strip()]
        out.append({"prefix": pre, "wall": time.perf_counter() - t0, "frames": log, "stderr": err})
    return out


def reference_leg(cfg, yuv, w, h, frames, qp, nproc, workdir):
    """stock Thorenc, nproc concurrent single-threaded processes; rate over frames 2..n (a 2-frame run is subtracted)"""
    def wall(n, k):
        t0 = time.perf_counter()
        ps = [subprocess.Popen(enc_cmd(REF_ENC, cfg, yuv, w, h, n, qp, os.path.join(workdir, f"ref{n}_{i}"), rec=(i == 0)), stdout=subprocess.DEVNULL,
                               stderr=subprocess.DEVNULL) for i in range(k)]
        for p in ps:
            if p.wait(timeout=3000) != 0:
                raise RuntimeError("stock Thorenc failed")
        return time.perf_counter() - t0
    t_short = wall(2, nproc)
    t_long = wall(frames, nproc)
    return {"fps": nproc * (frames - 2) / max(t_long - t_short, 1e-6), "seconds": {"2": round(t_short, 2), str(frames): round(t_long, 2)}}


def encoder_e2e_leg(args, workdir, env):
    """one stream: stock Thorenc vs Thorenc_gpu on the same frames, outputs compared byte for byte"""
    n = args.e2e_frames
    yuv = make_clip(os.path.join(workdir, f"e2e_{args.width}x{args.height}_{n}.yuv"), args.width, args.height, n)
    ref = reference_leg(args.config, yuv, args.width, args.height, n, args.qp, 1, workdir)
    g = run_streams(GPU_ENC, args.config, yuv, args.width, args.height, n, args.qp, 1, workdir, env, "e2e_gpu")[0]
    log = g["frames"]
    gpu_fps = (n - 2) / max((log[n - 1][1] - log[1][1]) * 1e-3, 1e-9) if len(log) >= n else None
    rd = lambda p: open(p, "rb").read()                       # noqa: E731
    res = {"config": f"{args.config} {args.width}x{args.height} qp {args.qp}, hard clip, {n} frames, one stream",
           "stock_fps": round(ref["fps"], 3), "stock_seconds": ref["seconds"], "gpu_fps": round(gpu_fps, 3) if gpu_fps else None,
           "gpu_wall_seconds": round(g["wall"], 2),
           "identical_bit": rd(os.path.join(workdir, f"ref{n}_0.bit")) == rd(g["prefix"] + ".bit"),
           "identical_rec": rd(os.path.join(workdir, f"ref{n}_0_rec.yuv")) == rd(g["prefix"] + "_rec.yuv"),
           "bitstream_bytes": os.path.getsize(g["prefix"] + ".bit"),
           "gpu_hook": [ln for ln in g["stderr"].splitlines() if "thorcuda frame encoder" in ln and "Mcycles" not in ln],
           "note": "frames 3..n: stock = difference of an n-frame and a 2-frame run (process wall clock), gpu = the hooked encoder's own clock "
                   "(CUDA start-up outside); single host thread in both"}
    return res, ref


def cut_stream(bit, chunks, out):
    """the first `chunks` coded frames of a stream (frames are 4-byte-length-prefixed chunks, enc/putbits.c:57-95)"""
    import struct
    data, i = open(bit, "rb").read(), 0
    for _ in range(chunks):
        i += 4 + struct.unpack(">I", data[i:i + 4])[0]
    open(out, "wb").write(data[:i])


def decoder_e2e_leg(args, workdir):
    """stock Thordec vs Thordec_gpu (whole reconstruction on the device) on RA streams the stock encoder makes here; rate between
    a short and a long stream (both end on a sub-GOP boundary), which cancels process and CUDA start-up"""
    if not all(os.path.exists(x) for x in (REF_ENC, REF_DEC, GPU_DEC)):
        return {"unavailable": "oracle/_ref/Thorenc, Thordec or integration/_build/Thordec_gpu not built"}
    from thorref import write_hard_clip
    out = {}
    for (w, h, short, n, skip) in ((1920, 1080, 41, 89, 36), (3840, 2160, 17, 33, 12)):
        yuv = os.path.join(workdir, f"dec_{w}x{h}.yuv")
        write_hard_clip(yuv, w, h, n, 43, noise=2.0)
        pre = os.path.join(workdir, f"dec_{w}x{h}")
        subprocess.run(enc_cmd(REF_ENC, "RA_low_complexity", yuv, w, h, n, 32, pre, rec=True), check=True, capture_output=True, timeout=2400)
        os.remove(yuv)
        cut_stream(pre + ".bit", short, pre + "_short.bit")
        res = {"stream": f"RA_low_complexity {w}x{h} qp 32, {short} and {n} frames, {os.path.getsize(pre + '.bit')} bytes"}
        outs = {}
        for name, exe in (("stock", REF_DEC), ("gpu", GPU_DEC)):
            best, steady = {}, []
            for key, bit in (("short", pre + "_short.bit"), ("long", pre + ".bit")):
                for _ in range(3):
                    t0 = time.perf_counter()
                    r = subprocess.run([exe, bit, pre + f"_{name}.yuv"], check=True, capture_output=True, text=True, timeout=900,
                                       env=dict(os.environ, THORCUDA_HOOK_TIMING="1", THORCUDA_HOOK_STEADY_FROM=str(skip)))
                    dt = time.perf_counter() - t0
                    best[key] = min(best.get(key, dt), dt)
                    if key == "long":
                        steady += [float(ln.split("=")[1].split("frames/s")[0]) for ln in r.stderr.splitlines() if "steady state" in ln]
            outs[name] = open(pre + f"_{name}.yuv", "rb").read()
            os.remove(pre + f"_{name}.yuv")
            res[name + "_fps"] = round((n - short) / max(best["long"] - best["short"], 1e-6), 2)
            res[name + "_seconds"] = {k: round(v, 3) for k, v in best.items()}
            if steady:
                res[name + "_steady_fps"] = round(sorted(steady)[len(steady) // 2], 1)
        res["identical_yuv"] = outs["stock"] == outs["gpu"] == open(pre + "_rec.yuv", "rb").read()
        os.remove(pre + "_rec.yuv")
        out[f"{w}x{h}"] = res
    out["note"] = ("*_fps: frames per second between the short and the long stream (process wall clock, best of 3 each, files in tmpfs); the hooked "
                   "decoder's CUDA start-up (1-2 s) varies by tenths of a second, which at these stream lengths swamps the difference, so "
                   "gpu_steady_fps -- its own wall clock from its 37th (2160p: 13th) frame until the output file is complete, median of 3 -- is the figure to read")
    return out


# ------------------------------------------------------------ closed-GOP segments of ONE stream ----
def segments_leg(args, rank, world, env, shared, barrier, allmax):
    """BASELINE.json configs[2]'s partitioning: ONE RA stream cut at its intra frames (config: -intra_period 64), its segments
    sharded contiguously over the ranks (weak: --segments-per-gpu each), one Thorenc_gpu process per segment (-seg_start /
    -seg_end, integration/enc_segment_hook.c), the processes of a rank sharing its GPU.  Ranks exchange nothing; rank 0 joins
    the chunks (thor_b200/segments.py) and checks the join with the STOCK decoder: the stream up to the end of rank 0's shard
    (which needs the next rank's intra frame) must decode to exactly the reconstructions rank 0's encoders wrote."""
    from thorref import THOR_CONFIGS
    from thor_b200 import segments as sg
    cfg, w, h = args.segments_config, args.width, args.height
    opts = THOR_CONFIGS[cfg]
    period, reorder = int(opts[opts.index("-intra_period") + 1]), int(opts[opts.index("-num_reorder_pics") + 1])
    spg = args.segments_per_gpu
    nseg = world * spg
    frames = nseg * period + 1
    fsize = w * h * 3 // 2
    ranges = sg.segment_ranges(frames, period, nseg)
    mine = sg.shard(list(enumerate(ranges)), rank, world)
    lo, hi = mine[0][1][0], mine[-1][1][1]
    hi = frames - 1 if hi is None else hi                    # the range's closing intra frame is encoded too
    base = os.path.join(shared, f"seg_base_{w}x{h}.yuv")
    if rank == 0:
        make_clip(base, w, h, period, seed=44)
    barrier()
    # the rank's input: a sparse file of the whole stream's length holding only the frames its encoders read
    yuv = os.path.join(shared, f"seg_in_rank{rank}.yuv")
    with open(base, "rb") as src, open(yuv, "wb") as dst:
        dst.truncate(frames * fsize)
        for n in range(lo, hi + 1):
            src.seek((n % period) * fsize)
            dst.seek(n * fsize)
            dst.write(src.read(fsize))
    sargs = ["-if", yuv, "-width", str(w), "-height", str(h), "-n", str(frames), "-f", "60", "-qp", str(args.qp), "-snrcalc", "0"] + opts
    penv = dict(os.environ, **env, THORCUDA_ENC_CTAS=str(args.segments_ctas), THORCUDA_BLOCKING_SYNC="1")
    prefix = os.path.join(shared, f"seg_rank{rank}")
    barrier()
    t0 = time.perf_counter()
    files = sg.encode_segments(GPU_ENC, sargs, [r for _, r in mine], prefix, env=penv, rec=(rank == 0), timeout=3000)
    dt = time.perf_counter() - t0
    os.remove(yuv)
    barrier()
    dt = allmax(dt)
    if rank != 0:
        barrier()
        return None
    res = {"config": f"Thorenc config_{cfg}.txt {w}x{h} qp {args.qp}: ONE stream of {frames} frames = {nseg} closed-GOP segments of {period}, "
                     f"{spg} per GPU, one encoder process per segment ({args.segments_ctas} CTAs per frame each)",
           "frames": frames, "segments": nseg, "n_gpus": world, "scaling": "weak", "seconds": round(dt, 2), "fps": round(frames / dt, 3),
           "redundant_intra_frames": nseg - 1,
           "timing": "wall clock from the start barrier to the last encoder process's exit, max over ranks: process and CUDA start-up, "
                     "file reads, device work, entropy coding, bitstream and reconstruction writes"}
    streams = []
    for r in range(world):
        for k in range(len(sg.shard(ranges, r, world))):
            streams.append(open(os.path.join(shared, f"seg_rank{r}_{k}.bit"), "rb").read())
    joined = sg.join_segment_streams(ranges, streams, frames, reorder, period)
    import hashlib
    res["joined_bytes"], res["joined_sha256"] = len(joined), hashlib.sha256(joined).hexdigest()
    # decode check with the stock decoder: every coded frame up to the last one of rank 0's shard
    end = mine[-1][1][1]
    order = sg.coding_order(frames, reorder, period)
    last = max(f.num_encoded_before for f in order if end is None or f.frame_num < end)
    cut, pos = 0, 0
    for _ in range(last + 1):
        pos += 4 + int.from_bytes(joined[pos:pos + 4], "big")
    open(prefix + "_joined.bit", "wb").write(joined[:pos])
    if os.path.exists(REF_DEC):
        subprocess.run([REF_DEC, prefix + "_joined.bit", prefix + "_dec.yuv"], check=True, capture_output=True, timeout=3000)
        nshow = frames if end is None else end
        ok, off = True, 0
        with open(prefix + "_dec.yuv", "rb") as dec:
            for _, rec in files:
                with open(rec, "rb") as rf:
                    while True:
                        a = rf.read(fsize)
                        if not a:
                            break
                        ok = ok and a == dec.read(fsize)
                        off += 1
        res["stock_decoder_reproduces_the_encoders_reconstruction"] = bool(ok and off == nshow)
        res["decode_check"] = f"stock Thordec on the joined stream's first {last + 1} coded frames vs the {off} frames rank 0's segment encoders reconstructed"
    for p in [prefix + "_joined.bit", prefix + "_dec.yuv"] + [x for fl in files for x in fl]:
        if os.path.exists(p):
            os.remove(p)
    barrier()
    return res


# ------------------------------------------------------------------------------ arms ----
def workload_config(args, streams):
    return {"workload": f"Thorenc config_{args.config}.txt {args.width}x{args.height} 8-bit synthetic (hard clip, noise 5), qp {args.qp}: "
                        f"whole encode, decisions on the device, entropy coding on the host",
            "streams_per_gpu": streams, "frames_per_step": streams, "ctas_per_frame": args.ctas_per_frame, "qp": args.qp,
            "l2": "every frame's working set (original, 2 references, reconstruction, decisions: ~23 MB) is new data per step and stream; "
                  "the streams' frames together exceed the 126 MB L2",
            "parallelism": "independent streams (one encoder process each, CUDA MPS), one rank per GPU, no collective"}


def run_reference_arm(args, rank, world):
    """--impl reference: the unmodified reference encoder on all host threads, same clip and configuration"""
    if rank != 0:
        return
    threads = host_threads()
    frames = 2 + min(args.steps, 16)
    with tempfile.TemporaryDirectory(dir=shm_dir()) as d:
        yuv = make_clip(os.path.join(d, "clip.yuv"), args.width, args.height, frames)
        r = reference_leg(args.config, yuv, args.width, args.height, frames, args.qp, threads, d)
    fps = r["fps"]
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * threads / fps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args, threads),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": "reference",
                             "sample": f"{threads} concurrent stock Thorenc processes (single-threaded each), frames 3..{frames} of the clip "
                                       f"(an n-frame run minus a 2-frame run), {r['seconds']} s"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_gpu_arm(args, rank, world, local_rank):
    for need in (GPU_ENC, REF_ENC):
        if not os.path.exists(need):
            raise SystemExit(f"{need} is missing: run __graft_entry__.build() where /root/reference is mounted")
    dist = None
    if world > 1:
        os.environ["NCCL_DEBUG"] = os.environ.get("THOR_BENCH_NCCL_DEBUG", "WARN")     # rank 0's stdout is ONE JSON line: no NCCL version banner
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    S, W, K = args.streams, max(args.warmup, 3), args.steps
    frames = 1 + W + K                                  # the intra frame, W warm-up P frames, K timed P frames
    mps_dir = f"/tmp/thor_mps_{os.environ.get('MASTER_PORT', os.getpid())}"
    mps = Mps(not args.no_mps)
    mps.env_dir = mps_dir

    def barrier():
        if dist is not None:
            dist.barrier()

    def allmax(v):
        if dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v):
        if dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    with tempfile.TemporaryDirectory(dir=shm_dir()) as d:
        yuv = make_clip(os.path.join(d, "clip.yuv"), args.width, args.height, frames)
        if rank == 0:
            mps.start()
        barrier()
        env = {"CUDA_VISIBLE_DEVICES": str(local_rank)} if world > 1 or "CUDA_VISIBLE_DEVICES" not in os.environ else {}
        mps_on = allmax(1.0 if mps.on else 0.0) > 0          # rank 0 knows; everybody uses the same pipe directory
        senv = dict(env, THORCUDA_ENC_CTAS=str(args.ctas_per_frame))
        if mps_on:
            senv.update({"CUDA_MPS_PIPE_DIRECTORY": os.path.join(mps_dir, "pipe"), "CUDA_MPS_LOG_DIRECTORY": os.path.join(mps_dir, "log")})
            env.update({"CUDA_MPS_PIPE_DIRECTORY": os.path.join(mps_dir, "pipe"), "CUDA_MPS_LOG_DIRECTORY": os.path.join(mps_dir, "log")})
        sampler = ClockSampler(local_rank)
        try:
            barrier()
            sampler.start()
            streams = run_streams(GPU_ENC, args.config, yuv, args.width, args.height, frames, args.qp, S, d, senv, "s")
            clocks = sampler.stop()
            barrier()
            dev_s, host_s, launches = 0.0, 0.0, 0
            for st in streams:
                log = st["frames"]
                if len(log) < frames:
                    raise SystemExit("a stream wrote no frame log: is integration/_build/Thorenc_gpu current?")
                dev_s = max(dev_s, sum(r[2] for r in log[W + 1:W + 1 + K]) * 1e-3)
                host_s = max(host_s, (log[W + K][1] - log[W][1]) * 1e-3)
                launches += int(log[W + K][3] - log[W][3])
            dev_s, host_s, launches = allmax(dev_s), allmax(host_s), int(allsum(launches))
            total = world * S * K
            fbytes = args.width * args.height * 3 // 2
            nsb = ((args.width + 63) // 64) * ((args.height + 63) // 64)
            d2h = nsb * (85 * 56 + 12288) + fbytes + nsb
            # dominant kernel: k_encode_frame.  Algorithmic bytes of one launch: original + 2 references read, reconstruction
            # written, neighbour units (16 B per 4x4), decided trees and coefficients written
            alg = 4 * fbytes + (args.width // 4) * (args.height // 4) * 16 + nsb * (85 * 56 + 12288)
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
            peak = peaks.get("hbm_gbs", 6650.0)
            launch_ms = dev_s * 1e3 / K                      # a frame's device work is >99% this kernel (profiles/r2_launches_*.csv)
            traffic = None
            tp = os.path.join(ROOT, "profiles", "r2_ncu_encode_frame.json")
            if os.path.exists(tp):
                traffic = json.load(open(tp)).get("dram_bytes_per_launch")
            line = {"metric": METRIC, "value": total / dev_s, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": 1e3 * dev_s / K,
                    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                    "config": workload_config(args, S), "clocks": clocks, "gpu_launches": launches, "mps": mps_on,
                    "e2e": {"value": total / host_s, "unit": "frames/s", "h2d_bytes_per_step": S * fbytes, "d2h_bytes_per_step": S * d2h,
                            "ms_per_step": 1e3 * host_s / K,
                            "timing": "the encoders' own clocks from the end of frame W to the end of frame W+K, max over streams and ranks: file read "
                                      "(tmpfs), upload from page-locked frames, device work, download of trees / coefficients / reconstruction, "
                                      "entropy coding with the reference's write_block, bitstream write"},
                    "roofline": {"bound": "hbm", "kernel": "k_encode_frame", "achieved": alg / (launch_ms * 1e-3) / 1e9, "peak": peak,
                                 "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                                 "unit": "GB/s", "frac": alg / (launch_ms * 1e-3) / 1e9 / peak, "traffic": traffic, "launch_ms": launch_ms,
                                 "algorithmic_bytes_per_launch": alg,
                                 "note": "not a bandwidth kernel: a frame is a 2-superblock-lag wavefront of serial rate-distortion decisions, at most "
                                         "min(rows, cols/2) = 15 CTAs of one frame run at a time (DESIGN.md 4.10); launch_ms is the CUDA-event time of a "
                                         "frame with the streams sharing the device"}}
            if rank == 0 and not args.no_cpu:
                e2e, ref1 = encoder_e2e_leg(args, d, {k: v for k, v in env.items() if k.startswith("CUDA_MPS")})
                line["encoder_e2e"] = e2e
                line["cpu_baseline"] = {"value": ref1["fps"], "unit": "frames/s", "cores": 1, "kind": "reference",
                                        "sample": f"stock Thorenc (SSE4 build, single-threaded by design), frames 3..{args.e2e_frames} of the same clip and "
                                                  f"configuration, {ref1['seconds']} s"}
            barrier()
            if not args.no_segments:
                shared = f"/dev/shm/thor_seg_{os.environ.get('MASTER_PORT', os.getpid())}" if shm_dir() else f"/tmp/thor_seg_{os.environ.get('MASTER_PORT', os.getpid())}"
                need = (args.segments_per_gpu * 64 + 1) * fbytes * (2 if rank == 0 else 1) * world
                if shutil.disk_usage(os.path.dirname(shared)).free < 1.3 * need:
                    shared = shared.replace("/dev/shm/", "/tmp/")
                os.makedirs(shared, exist_ok=True)
                try:
                    seg = segments_leg(args, rank, world, {k: v for k, v in env.items()}, shared, barrier, allmax)
                finally:
                    barrier()
                    if rank == 0:
                        shutil.rmtree(shared, ignore_errors=True)
                if rank == 0:
                    line["segments"] = seg
        finally:
            if rank == 0:
                mps.stop()
        if rank == 0 and not args.no_decoder:
            line["decoder_e2e"] = decoder_e2e_leg(args, d)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="thorcuda", choices=["thorcuda", "reference"])
    ap.add_argument("--streams", type=int, default=36, help="independent encoder processes per GPU = frames per step")
    ap.add_argument("--ctas-per-frame", type=int, default=4, help="CTAs k_encode_frame uses per frame in the multi-stream run (THORCUDA_ENC_CTAS): "
                    "streams x this should not exceed the 148 SMs; the one-stream legs use the library default, min(rows, cols/2)")
    ap.add_argument("--allow-oversubscription", action="store_true", help="run although streams x CTAs per frame exceeds the 148 resident CTAs of a B200 "
                    "(one k_encode_frame CTA per SM): such runs died with an illegal memory access about every second time (DESIGN.md 4.10, open issue)")
    ap.add_argument("--config", default="LDB_medium_complexity")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--qp", type=int, default=32)
    ap.add_argument("--e2e-frames", type=int, default=10, help="frames of the one-stream stock-vs-hooked comparison")
    ap.add_argument("--no-mps", action="store_true", help="do not start the CUDA MPS daemon (the streams then time-slice the GPU)")
    ap.add_argument("--no-cpu", action="store_true", help="skip cpu_baseline / encoder_e2e")
    ap.add_argument("--no-decoder", action="store_true", help="skip decoder_e2e (about a minute)")
    ap.add_argument("--no-segments", action="store_true", help="skip the closed-GOP segment leg (one RA stream sharded over the GPUs)")
    ap.add_argument("--segments-config", default="RA_medium_complexity")
    ap.add_argument("--segments-per-gpu", type=int, default=8, help="segments (= encoder processes) per GPU; the stream has gpus x this x 64 + 1 frames")
    ap.add_argument("--segments-ctas", type=int, default=15, help="CTAs per frame of each segment encoder")
    args = ap.parse_args()
    if args.impl != "reference" and not args.allow_oversubscription and max(args.streams * args.ctas_per_frame, args.segments_per_gpu * args.segments_ctas) > 148:
        raise SystemExit("streams x ctas-per-frame (or segments-per-gpu x segments-ctas) exceeds the 148 SMs: see --allow-oversubscription")
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    else:
        run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
