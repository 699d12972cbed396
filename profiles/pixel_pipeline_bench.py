#!/usr/bin/env python
"""bench.py -- throughput of the Thor pixel pipeline on B200 (libthorcuda) beside the reference's CPU kernels.

Workload (BASELINE.json configs[1]): the pixel-pipeline job of one 1920x1080 8-bit P-frame coded with the
LDB medium-complexity settings (thor_b200/workload.py): SAD tables against 2 references (range +-16), bilinear
sub-pel costs, inter prediction, residual/DCT/quant/dequant/IDCT/reconstruct, deblock + CLPF + reference
creation.  A step = that job for one frame of each of `--streams` independent streams (the unit of
parallelism SURVEY 8e allows for LDB: independent streams, no exchange), all frames of a step in the same
launches.  metric = frames/s.

  value : inputs (pictures, side information, task lists) resident in HBM, outputs left in HBM.
  e2e   : the same job through the C ABI from pinned HOST buffers: original frame, side information and task
          lists go host->device every step; SAD tables, sub-pel results, coefficients/cbp/ssd, CLPF bits and
          the reconstructed frame come back every step.
  cpu_baseline / --impl reference : the same job with the UNMODIFIED reference's own functions
          (oracle/_ref/libthorpipe_ref.so; the oracle port if that is absent) on all host cores.

One process per GPU (torchrun for N>1), streams sharded over ranks, no collective on the data path.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))   # round 2: moved from the repo root (bench.py now measures the encoder)
sys.path.insert(0, ROOT)

ST_SAD, ST_SUBPEL, ST_PRED, ST_TB, ST_FILTER = 1, 2, 4, 8, 16


# --------------------------------------------------------------------------------- CPU arm ----
class PipeJob(C.Structure):   # oracle/pipe_driver.c
    _fields_ = ([(n, C.c_int32) for n in ("w", "h", "qp", "range", "bipred_filter", "stages", "band_y0", "band_y1")] +
                [("org", C.c_void_p * 3), ("ref", (C.c_void_p * 3) * 2), ("ref_stride_y", C.c_int32), ("ref_stride_c", C.c_int32),
                 ("rec", C.c_void_p * 3), ("newref", C.c_void_p * 3), ("newref_stride_y", C.c_int32), ("newref_stride_c", C.c_int32),
                 ("pred", C.c_void_p), ("npred", C.c_int32), ("tb", C.c_void_p), ("ntb", C.c_int32),
                 ("cost", C.c_void_p), ("ncost", C.c_int32), ("blockinfo", C.c_void_p),
                 ("t8", C.c_void_p * 2), ("t16", C.c_void_p * 2), ("t32", C.c_void_p * 2), ("t64", C.c_void_p * 2),
                 ("half", C.c_void_p), ("quarter", C.c_void_p), ("coeffq", C.c_void_p), ("cbp", C.c_void_p), ("ssd", C.c_void_p),
                 ("clpf_bits", C.c_void_p)])


def load_cpu_pipe():
    """(library, kind): the reference build when oracle/_ref travelled here, else the oracle port."""
    p = os.path.join(ROOT, "oracle", "_ref", "libthorpipe_ref.so")
    if os.path.exists(p):
        return C.CDLL(p), "reference"
    p = os.path.join(ROOT, "oracle", "libthorpipe_port.so")
    if not os.path.exists(p):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)
    return C.CDLL(p), "port"


class CpuFrame:
    """Host-side buffers + banded execution of one frame job on P threads."""

    def __init__(self, job, lib, threads):
        from thor_b200.workload import HostPicture, PAD_C, PAD_Y, aligned_zeros
        from thor_b200.lib import COST_RESULT
        self.job, self.lib, self.threads = job, lib, threads
        w, h, R = job["width"], job["height"], job["range"]
        n2 = (2 * R) ** 2
        self.rec = HostPicture(w, h)
        self.newref = HostPicture(w, h, PAD_Y, PAD_C)
        self.tab = [{bs: aligned_zeros(((w // bs) * (h // bs) * n2,), np.uint16 if bs <= 16 else np.uint32) for bs in (8, 16, 32, 64)}
                    for _ in range(2)]
        self.half = np.zeros(job["cost"].size, COST_RESULT)
        self.quarter = np.zeros(job["cost"].size, COST_RESULT)
        self.coeffq = np.zeros(job["pool_entries"], np.int16)
        self.cbp = np.zeros(job["tb"].size, np.uint8)
        self.ssd = np.zeros(job["tb"].size, np.uint32)
        self.bits = np.full(max(1, (w // 64) * (h // 64)), 0xFF, np.uint8)
        self.pool = ThreadPoolExecutor(threads)

    def _job(self, stages, y0, y1, pred, tb, cost, tb_first, cost_first):
        j, J = self.job, PipeJob()
        J.w, J.h, J.qp, J.range, J.bipred_filter, J.stages = j["width"], j["height"], j["qp"], j["range"], j["bipred_filter"], stages
        J.band_y0, J.band_y1 = y0, y1
        for i in range(3):
            J.org[i] = j["org"].origin_ptr(i)
            J.rec[i] = self.rec.origin_ptr(i)
            J.newref[i] = self.newref.origin_ptr(i)
            for r in range(2):
                J.ref[r][i] = j["refs"][r].origin_ptr(i)
        J.ref_stride_y, J.ref_stride_c = j["refs"][0].stride_y, j["refs"][0].stride_c
        J.newref_stride_y, J.newref_stride_c = self.newref.stride_y, self.newref.stride_c
        J.pred, J.npred = pred.ctypes.data, pred.size
        J.tb, J.ntb = tb.ctypes.data, tb.size
        J.cost, J.ncost = cost.ctypes.data, cost.size
        J.blockinfo = j["blockinfo"].ctypes.data
        for r in range(2):
            J.t8[r], J.t16[r] = self.tab[r][8].ctypes.data, self.tab[r][16].ctypes.data
            J.t32[r], J.t64[r] = self.tab[r][32].ctypes.data, self.tab[r][64].ctypes.data
        J.half = self.half.ctypes.data + 8 * cost_first
        J.quarter = self.quarter.ctypes.data + 8 * cost_first
        J.coeffq = self.coeffq.ctypes.data
        J.cbp = self.cbp.ctypes.data + tb_first
        J.ssd = self.ssd.ctypes.data + 4 * tb_first
        J.clpf_bits = self.bits.ctypes.data
        return J

    def run(self, stages=31):
        """Stages 1-4 banded over 64-row bands on P threads (tasks are sorted by band); stage 5 is the
        reference's whole-frame functions on one thread.  Returns seconds per stage group."""
        j = self.job
        h = j["height"]
        nb = (h + 63) // 64
        bands = [(b * 64, min(h, (b + 1) * 64)) for b in range(nb)]
        pb = j["pred"]["y"] // 64
        tbb = (j["tb"]["y"].astype(np.int64) * np.where(j["tb"]["plane"] > 0, 2, 1)) // 64
        cb = j["cost"]["y"] // 64

        def band_jobs(st):
            out = []
            for b, (y0, y1) in enumerate(bands):
                ps, pe = np.searchsorted(pb, [b, b + 1])
                ts, te = np.searchsorted(tbb, [b, b + 1])
                cs, ce = np.searchsorted(cb, [b, b + 1])
                out.append(self._job(st, y0, y1, j["pred"][ps:pe], j["tb"][ts:te], j["cost"][cs:ce], int(ts), int(cs)))
            return out

        t = {}
        # prediction of all bands must finish before the in-place transform stage reads it; SAD/sub-pel are independent
        for name, st in (("sad_subpel", stages & (ST_SAD | ST_SUBPEL)), ("predict", stages & ST_PRED), ("transform", stages & ST_TB)):
            t0 = time.perf_counter()
            if st:
                jobs = band_jobs(st)
                list(self.pool.map(lambda J: self.lib.pipe_run(C.byref(J)), jobs))
            t[name] = time.perf_counter() - t0
        t0 = time.perf_counter()
        if stages & ST_FILTER:
            J = self._job(ST_FILTER, 0, h, j["pred"][:0], j["tb"][:0], j["cost"][:0], 0, 0)
            self.lib.pipe_run(C.byref(J))
        t["loopfilter"] = time.perf_counter() - t0
        return t


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU kernels on the same frame job, all host threads."""
    if rank != 0:
        return
    from thor_b200.workload import make_job
    lib, kind = load_cpu_pipe()
    lib.pipe_init()
    threads = host_threads()
    job = make_job(args.width, args.height, seed=42, search_range=args.range)
    fr = CpuFrame(job, lib, threads)
    for _ in range(args.warmup):
        fr.run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fr.run()
    dt = time.perf_counter() - t0
    fps = args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": workload_config(args, 1),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                             "sample": f"{args.steps} frame(s) of the {args.width}x{args.height} frame job, one frame per step, "
                                       f"stages 1-4 banded over {threads} threads, stage 5 single-threaded"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------- decoder leg (optional) ----
LDB_LOW = ["-HQperiod", "12", "-mqpP", "1.2", "-dqpI", "-2", "-lambda_coeffI", "0.8", "-lambda_coeffP", "1.2", "-max_num_ref", "2",
           "-use_block_contexts", "1", "-intra_rdo", "0", "-enable_tb_split", "0", "-enable_pb_split", "0", "-early_skip_thr", "1.0",
           "-enable_bipred", "0", "-encoder_speed", "2"]          # config_LDB_low_complexity.txt


_RA = ["-HQperiod", "1", "-num_reorder_pics", "7", "-interp_ref", "1", "-intra_period", "64", "-dqpI", "-2", "-dqpB0", "3", "-dqpB1", "1",
       "-dqpB2", "0", "-mqpP", "1.2", "-mqpB", "1.2", "-mqpB0", "1.1", "-mqpB1", "1.2", "-mqpB2", "1.3", "-lambda_coeffI", "0.8",
       "-lambda_coeffP", "1.2", "-lambda_coeffB", "1.2", "-lambda_coeffB0", "1.2", "-lambda_coeffB1", "1.2", "-lambda_coeffB2", "1.2"]
RA_LOW = _RA + ["-intra_rdo", "0", "-enable_tb_split", "0", "-enable_pb_split", "0", "-early_skip_thr", "1.0", "-max_num_ref", "2",
                "-use_block_contexts", "1", "-enable_bipred", "1", "-encoder_speed", "2"]       # config_RA_low_complexity.txt
DEC_CONFIGS = {"LDB_low_complexity": LDB_LOW, "RA_low_complexity": RA_LOW}


def write_clip(path, width, height, frames, seed=42, noise=2.0, obj=False):
    """Seeded 8-bit I420 clip: box-blurred noise texture translating by (1, 2) samples per frame, slow chroma waves,
    Gaussian sensor noise of `noise` grey levels; obj: a 320x256 patch that moves on its own by (5, 13) per frame
    (SURVEY 8d's recipe)."""
    rng = np.random.default_rng(seed)
    big = rng.integers(0, 256, (height + 96, width + 96)).astype(np.float32)
    for ax in (0, 1):                                   # 5-tap box blur through cumulative sums
        c = np.cumsum(np.pad(big, [(3, 2) if a == ax else (0, 0) for a in (0, 1)], mode="edge"), axis=ax)
        big = (np.take(c, range(5, c.shape[ax]), axis=ax) - np.take(c, range(0, c.shape[ax] - 5), axis=ax)) / 5
    big = np.clip((big - 128) * 3 + 128, 0, 255)
    yy, xx = np.mgrid[0:height // 2, 0:width // 2]
    with open(path, "wb") as f:
        for n in range(frames):
            dy, dx = n % 64, (2 * n) % 64                         # stays inside the texture margin
            y = big[32 + dy:32 + dy + height, 32 + dx:32 + dx + width]
            if obj and height >= 512 and width >= 640:
                # a second layer drifting the other way under half of the picture (no single vector predicts the blend),
                # and a patch that moves on its own
                y = y.copy()
                ey, ex = (3 * n) % 64, 63 - n % 64
                y[:, width // 2:] = 0.5 * y[:, width // 2:] + 0.5 * big[32 + ey:32 + ey + height, 32 + ex:32 + ex + width][::-1, width // 2:]
                oy, ox = (5 * n) % (height - 256), (13 * n) % (width - 320)
                y[oy:oy + 256, ox:ox + 320] = big[:256, :320][::-1, ::-1]
            f.write(np.clip(y + rng.normal(0, noise, y.shape), 0, 255).astype(np.uint8).tobytes())
            f.write((128 + 60 * np.sin((xx + 2 * n) / 23.0) * np.cos(yy / 29.0)).astype(np.uint8).tobytes())
            f.write((128 + 60 * np.cos(xx / 27.0) * np.sin((yy + n) / 25.0)).astype(np.uint8).tobytes())


def decoder_leg(args):
    """End-to-end DECODE rate of the reference's Thordec against the same decoder with libthorcuda's fused frame tail
    hooked in (integration/): a 1080p LDB stream encoded by the reference's Thorenc here, decoded
    by both, outputs compared byte for byte.  Wall clock of the whole process, file I/O included, best of 5."""
    import tempfile
    enc, dec = os.path.join(ROOT, "oracle", "_ref", "Thorenc"), os.path.join(ROOT, "oracle", "_ref", "Thordec")
    tail, full = os.path.join(ROOT, "integration", "_build", "Thordec_tail"), os.path.join(ROOT, "integration", "_build", "Thordec_gpu")
    if not all(os.path.exists(x) for x in (enc, dec, tail, full)):
        return {"unavailable": "oracle/_ref/Thorenc, Thordec or integration/_build/Thordec_tail, Thordec_gpu not built"}
    frames = args.decoder_frames
    # the decoder cycles through its 33 reference slots before it reuses one, and the hook page-locks each slot the
    # first time it sees it: both stream lengths lie beyond that warm-up
    short = 40
    frames = max(frames, short + 16)
    # streams and decoded pictures live in memory (tmpfs) when the box has it: at 2160p the decoders otherwise measure the
    # scratch disk (12 MB of output per frame)
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=shm) as d:
        yuv = os.path.join(d, "clip.yuv")
        write_clip(yuv, args.width, args.height, frames, noise=5.0 if args.decoder_hard else 2.0, obj=args.decoder_hard)
        bits = {}
        for n in (short, frames):                      # two stream lengths: their difference cancels process start-up
            bits[n] = os.path.join(d, f"clip{n}.bit")
            subprocess.run([enc, "-if", yuv, "-of", bits[n], "-rf", os.path.join(d, "rec.yuv"), "-stat", os.path.join(d, "stat.txt"),
                            "-width", str(args.width), "-height", str(args.height), "-n", str(n), "-f", "60", "-qp", "32"]
                           + DEC_CONFIGS[args.decoder_config],
                           check=True, capture_output=True, timeout=900)
        res, outs, steady = {}, {}, {}
        for name, exe in (("reference", dec), ("gpu_tail", tail), ("gpu_recon", full)):
            best = {}
            for n in (short, frames):
                for _ in range(5):                     # CUDA start-up varies by tenths of a second from run to run
                    out = os.path.join(d, name + ".yuv")
                    t0 = time.perf_counter()
                    r = subprocess.run([exe, bits[n], out], check=True, capture_output=True, text=True, timeout=600,
                                       env=dict(os.environ, THORCUDA_HOOK_TIMING="1"))
                    dt = time.perf_counter() - t0
                    best[n] = min(best.get(n, dt), dt)
                    for ln in r.stderr.splitlines():          # the hooked decoder's own clock: frame 37 .. output complete
                        if "steady state" in ln and n == frames:
                            steady.setdefault(name, []).append(float(ln.split("=")[1].split("frames/s")[0]))
            outs[name] = open(out, "rb").read()
            res[name + "_fps"] = round((frames - short) / (best[frames] - best[short]), 2)
            res[name + "_startup_s"] = round(best[short] - short * (best[frames] - best[short]) / (frames - short), 3)
            res[name + "_seconds"] = {str(n): round(best[n], 3) for n in (short, frames)}
        for name, v in steady.items():
            res[name + "_steady_fps"] = round(sorted(v)[len(v) // 2], 1)       # median of the 5 runs of the long stream
        res.update({"frames": frames, "stream": f"{args.decoder_config} {args.width}x{args.height} qp 32{', hard clip' if args.decoder_hard else ''}, {os.path.getsize(bits[frames])} bytes",
                    "identical_yuv": outs["reference"] == outs["gpu_tail"] == outs["gpu_recon"], "scratch": "tmpfs (/dev/shm)" if shm else "default temporary directory",
                    "note": "*_fps: marginal decode rate between a short and a long stream (process wall clock, file I/O included; the CUDA "
                            "start-up of the hooked decoders varies by +-0.4 s between runs, which this difference only cancels on average); "
                            "gpu_recon_steady_fps: the hooked decoder's own wall clock from its 37th frame until the output file is complete; "
                            "gpu_tail: deblock + CLPF + reference creation on the GPU; gpu_recon: prediction, inverse transform and "
                            "reconstruction as well (references device-resident), the host parses the stream and writes the output"})
        return res


LDB_MEDIUM = LDB_LOW[:-12] + ["-intra_rdo", "1", "-enable_tb_split", "0", "-enable_pb_split", "0", "-early_skip_thr", "0.8",
                             "-enable_bipred", "1", "-encoder_speed", "1"]      # config_LDB_medium_complexity.txt


def encoder_leg(args):
    """End-to-end ENCODE rate (BASELINE metric 1, configs[1]): the reference's Thorenc against the same encoder with
    libthorcuda's frame tail hooked in (integration/Thorenc_tail: deblocking, CLPF decision, reference creation on the
    device), LDB medium complexity, .bit and reconstruction compared byte for byte.  Marginal rate between a 2-frame and an
    n-frame run (cancels start-up and the intra frame).  The motion search, mode decision and entropy coding stay on
    the host in both (SURVEY Tier A), so this is NOT where the device kernels show: it is reported so that the pixel-pipeline
    metric above is not mistaken for an encoder speed-up."""
    import tempfile
    enc, tail = os.path.join(ROOT, "oracle", "_ref", "Thorenc"), os.path.join(ROOT, "integration", "_build", "Thorenc_tail")
    if not all(os.path.exists(x) for x in (enc, tail)):
        return {"unavailable": "oracle/_ref/Thorenc or integration/_build/Thorenc_tail not built"}
    frames, short = max(args.encoder_frames, 4), 2
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=shm) as d:
        yuv = os.path.join(d, "clip.yuv")
        write_clip(yuv, args.width, args.height, frames, noise=args.encoder_noise, obj=True)
        res, outs, steady = {}, {}, []
        for name, exe in (("reference", enc), ("gpu_tail", tail)):
            secs = {}
            for n in (short, frames):
                pre = os.path.join(d, f"{name}{n}")
                for _ in range(2):                     # best of 2
                    t0 = time.perf_counter()
                    r = subprocess.run([exe, "-if", yuv, "-of", pre + ".bit", "-rf", pre + "_rec.yuv", "-stat", pre + "_stat.txt",
                                        "-width", str(args.width), "-height", str(args.height), "-n", str(n), "-f", "60", "-qp", "32"] + LDB_MEDIUM,
                                       check=True, capture_output=True, text=True, timeout=1800, env=dict(os.environ, THORCUDA_HOOK_TIMING="1"))
                    dt = time.perf_counter() - t0
                    secs[n] = min(secs.get(n, dt), dt)
                    for ln in r.stderr.splitlines():      # the hooked encoder's own clock: frames 3 .. n
                        if "steady state" in ln and n == frames:
                            steady.append(float(ln.split("=")[1].split("frames/s")[0]))
            outs[name] = (open(pre + ".bit", "rb").read(), open(pre + "_rec.yuv", "rb").read())
            res[name + "_fps"] = round((frames - short) / (secs[frames] - secs[short]), 3)
            res[name + "_seconds"] = {str(n): round(v, 2) for n, v in secs.items()}
        if steady:
            res["gpu_tail_steady_fps"] = round(max(steady), 3)
        res.update({"frames": frames, "config": f"LDB_medium_complexity {args.width}x{args.height} qp 32, clip noise {args.encoder_noise}, moving object", "bitstream_bytes": len(outs["reference"][0]),
                    "identical_bit": outs["reference"][0] == outs["gpu_tail"][0], "identical_rec": outs["reference"][1] == outs["gpu_tail"][1],
                    "note": "*_fps: marginal rate between a 2-frame and an n-frame run (process wall clock); for the hooked encoder that difference "
                            "carries the run-to-run variation of the CUDA start-up (1.1-1.6 s), so gpu_tail_steady_fps -- its own clock over the "
                            "same frames 3..n -- is the figure to compare with reference_fps; single host thread in both, host RDO / "
                            "motion-search driver / entropy coding dominate (SURVEY 3.3)"})
        return res


# --------------------------------------------------------------------------------- GPU arm ----
METRIC = "1080p pixel-pipeline frames/s (LDB medium-complexity frame job, bit-identical to the reference kernels)"


def workload_config(args, streams):
    return {"workload": f"Thorenc config_LDB_medium_complexity.txt {args.width}x{args.height} 8-bit synthetic: per-frame pixel-pipeline "
                        f"job (SAD tables +-{args.range} vs 2 refs, sub-pel costs, inter prediction, DCT/quant/recon, deblock+CLPF+reference)",
            "streams_per_gpu": streams, "frames_per_step": streams, "search_range": args.range, "qp": 32,
            "l2": "working set per step exceeds the 126 MB L2 (SAD tables alone are ~100 MB per frame and reference)",
            "parallelism": "independent streams, one process per GPU, no collective"}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed `ncu --set full`
# capture of this very command line (profiles/); only quoted when the workload is the one that was captured.
NCU_TRAFFIC = {("sad_table", 1920, 1080, 16, 16): {"bytes": 140482560 + 2917670000, "source": "profiles/r1_ncu_sad_table_v4.txt",
                                                  "alu_pipe_pct": 79.9}}


def ncu_traffic(kernel, job, B):
    t = NCU_TRAFFIC.get((kernel, job["width"], job["height"], job["range"], B))
    return t["bytes"] if t else None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thr.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class Stream:
    """Device state of one independent stream: original, two references, work picture, next reference."""

    def __init__(self, tb, job, index):
        from thor_b200.workload import PAD_C, PAD_Y
        w, h = job["width"], job["height"]
        self.job, self.index = job, index
        self.org = tb.Picture(w, h)
        self.refs = [tb.Picture(w, h, PAD_Y, PAD_C) for _ in range(2)]
        self.work = tb.Picture(w, h)
        self.newref = tb.Picture(w, h, PAD_Y, PAD_C)
        self.org.upload(*[job["org"].plane(i) for i in range(3)])
        for d, hp in zip(self.refs, job["refs"]):
            d.upload_padded(*hp.buf)
        tb.ops.blockinfo_upload(self.work, job["blockinfo"])
        self.pics = [self.org, self.refs[0], self.refs[1], self.work]
        self.slots = [2 * index, 2 * index + 1]


def merged_lists(tb, streams):
    """One resident list per kind covering every stream of the step (picture indices rebased per stream)."""
    pics, pred, tbs, cost = [], [], [], []
    pool = 0
    for k, s in enumerate(streams):
        base = 4 * k
        pics += s.pics
        p = s.job["pred"].copy()
        p["dst"] += base
        p["ref0"] += base
        p["ref1"] = np.where(p["ref1"] >= 0, p["ref1"] + base, -1)
        t = s.job["tb"].copy()
        for f in ("org", "pred", "rec"):
            t[f] += base
        t["coeff_offset"] += pool
        pool += s.job["pool_entries"]
        c = s.job["cost"].copy()
        c["org"] += base
        c["ref"] += base
        pred.append(p); tbs.append(t); cost.append(c)
    return (pics, np.concatenate(pred), np.concatenate(tbs), np.concatenate(cost))


def bind_to_gpu_numa_node(index):
    """Run this process (and so first-touch its pinned staging buffers) on the CPUs local to GPU `index`: the e2e leg
    moves 0.77 GB per step over PCIe and a cross-socket hop shows.  Returns a note for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} cpus local to gpu {index}"
    except Exception as e:          # no NVML / restricted container: keep the inherited affinity
        return f"unchanged ({type(e).__name__})"
    return "unchanged"


def run_gpu_arm(args, rank, world, local_rank):
    import thor_b200 as tb
    from thor_b200.workload import algorithmic_bytes, make_job, shard_streams
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    affinity = bind_to_gpu_numa_node(local_rank)
    tb.init(local_rank)
    lib = tb.cdll()

    # weak scaling: every rank owns `--streams` streams; stream ids are sharded over ranks for the seeds
    total_streams = args.streams * world
    my_ids = shard_streams(total_streams, rank, world)
    njobs = min(args.distinct_jobs, len(my_ids))
    jobs = [make_job(args.width, args.height, seed=42 + 1000 * rank + k, search_range=args.range) for k in range(njobs)]
    if len(my_ids) * 4 > 128:
        raise SystemExit("--streams must be <= 32 (128 pictures per task list)")
    streams = [Stream(tb, jobs[k % njobs], k) for k in range(len(my_ids))]
    B = len(streams)
    pics, pred, tbs, cost = merged_lists(tb, streams)
    l_pred = tb.ops.TaskList(tb.LIST_PRED, pics, pred)
    l_tb = tb.ops.TaskList(tb.LIST_TB, pics, tbs)
    l_cost = tb.ops.TaskList(tb.LIST_COST, pics, cost)
    slots = [s for st in streams for s in st.slots]
    orgs = [st.org for st in streams for _ in range(2)]
    refs = [r for st in streams for r in st.refs]
    # resident run: the CLPF bits stay in HBM (sb_bits_out = NULL); the e2e leg below fetches them
    lf_jobs = (tb.ops.LfJob * B)(*[tb.ops.LfJob(st.work.handle, st.org.handle, st.newref.handle, st.job["qp"], 15, None, None)
                                   for st in streams])

    def stage_sad():
        tb.ops.sad_table_build(slots, orgs, refs, args.range)

    def stage_subpel():
        l_cost.cost(tb.COST_FASTHALF)
        l_cost.cost(tb.COST_FASTQUARTER)

    def stage_pred():
        l_pred.predict(1)

    def stage_tb():
        l_tb.encode()

    def stage_filter():
        tb.check(lib.thorcuda_loopfilter_frames(B, C.byref(lf_jobs)))

    stages = [("sad_table", stage_sad), ("subpel", stage_subpel), ("predict", stage_pred), ("transform", stage_tb),
              ("loopfilter", stage_filter)]

    def step():
        for _, f in stages:
            f()

    def barrier():
        tb.sync()
        if dist is not None:
            dist.barrier()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = tb.launch_count()
    tid = lib.thorcuda_timer_begin()
    for _ in range(args.steps):
        step()
    lib.thorcuda_timer_end(tid)
    ms = lib.thorcuda_timer_elapsed_ms(tid)
    launches = tb.launch_count() - launches0
    if dist is not None:
        import torch
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = B * world * args.steps / (ms / 1e3)

    # per-stage device time (CUDA events on the library's stream), separate pass so the main loop is undisturbed;
    # each stage is timed IN SITU, between its neighbours in full steps: a stage repeated back to back would see its own
    # output as input (the transform stage would find an already reconstructed picture: all-zero residuals)
    stage_ms = {name: 0.0 for name, _ in stages}
    reps = max(1, min(args.steps, 50))
    for _ in range(reps):
        ids = []
        for name, f in stages:
            tid = lib.thorcuda_timer_begin()
            f()
            lib.thorcuda_timer_end(tid)
            ids.append((name, tid))
        tb.sync()
        for name, tid in ids:
            stage_ms[name] += lib.thorcuda_timer_elapsed_ms(tid) / reps
    clocks = sampler.stop()          # sampled over the timed steps and the per-stage timed passes (all under load)

    # outside the metric (the reference arm has no such stage): the write_coeff bit counts of every transform block of the
    # step, on the levels the transform stage has just left in the pool
    l_tb.coeff_bits()
    tb.sync()
    tid = lib.thorcuda_timer_begin()
    for _ in range(20):
        l_tb.coeff_bits()
    lib.thorcuda_timer_end(tid)
    bits_ms = lib.thorcuda_timer_elapsed_ms(tid) / 20
    nbits = np.zeros(l_tb.ntasks, np.uint32)
    l_tb.coeff_bits(nbits)
    tb.sync()

    # ---- e2e: host buffers in, host buffers out, every step ----
    e2e = measure_e2e(args, tb, lib, streams, barrier, dist, world)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        ab = algorithmic_bytes(jobs[0])
        dom = max(stage_ms, key=stage_ms.get)
        launches_per_stage = {"sad_table": 1, "subpel": 2, "predict": 1, "transform": None, "loopfilter": 2}
        per_stage = {k: {"ms": round(v, 4), "algorithmic_GBps": round(B * ab[k] / (v * 1e-3) / 1e9, 1),
                         "frac_of_hbm_peak": round(B * ab[k] / (v * 1e-3) / 1e9 / peak, 4)} for k, v in stage_ms.items()}
        n_launch = launches_per_stage[dom] or 1
        achieved = B * ab[dom] / (stage_ms[dom] * 1e-3) / 1e9
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic", "config": workload_config(args, B), "clocks": clocks, "gpu_launches": int(launches), "cpu_affinity": affinity,
                "e2e": e2e,
                "roofline": {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                             "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": ncu_traffic(dom, jobs[0], B),
                             "launch_ms": round(stage_ms[dom] / n_launch, 4), "algorithmic_bytes_per_launch": int(B * ab[dom] / n_launch)},
                "stages": per_stage,
                "extra_stages": {"coeff_bits": {"ms": round(bits_ms, 4), "transform_blocks": int(l_tb.ntasks), "coded": int((nbits != 0).sum()),
                                                "mean_bits_per_coded_block": round(float(nbits.sum()) / max(1, int((nbits != 0).sum())), 1),
                                                "note": "not part of `value`: exact write_coeff bit counts of the step's transform blocks"}}}
        # The SAD table is bound by the integer ALU pipe, not by HBM: W*H*(2R)^2 abs-diffs per (frame, ref)
        # pair against the rate of a register-only VABSDIFF4 loop measured on this device just now.
        alu_peak = lib.thorcuda_calibrate_absdiff(0)
        if alu_peak > 0:
            pairs = len(slots)
            absdiffs = float(args.width) * args.height * (2 * args.range) ** 2 * pairs
            alu = absdiffs / (stage_ms["sad_table"] * 1e-3)
            line["roofline"]["alu"] = {"kernel": "sad_table", "achieved": round(alu / 1e12, 2), "peak": round(alu_peak / 1e12, 2),
                                       "unit": "T abs-diff/s", "frac": round(alu / alu_peak, 4),
                                       "peak_source": "thorcuda_calibrate_absdiff (register-only VABSDIFF4.ACC loop, measured in this run)"}
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_and_parity(args, tb, streams[0])
        if world == 1 and args.encoder:
            line["encoder_e2e"] = encoder_leg(args)
        if world == 1 and args.decoder:
            line["decoder_e2e"] = decoder_leg(args)
        print(json.dumps(line), flush=True)
    barrier()
    if dist is not None:
        dist.destroy_process_group()
    tb.shutdown()


def measure_e2e(args, tb, lib, streams, barrier, dist, world):
    """The frame job through the C ABI with HOST buffers: per step and stream, H2D of the original frame, the
    side information and the three task lists; D2H of the SAD tables (4 levels x 2 refs), sub-pel results,
    coefficients / cbp / ssd, CLPF bits and the reconstructed frame."""
    from thor_b200.lib import COST_RESULT
    # Two independent groups of streams are kept in flight: group B is submitted while group A's results (0.77 GB of SAD
    # tables per 4 frames) are still crossing PCIe, and a group's next frame is only submitted once its previous results
    # have arrived in host memory (thorcuda_fence_wait) -- what an encoder serving several streams does.
    E = max(1, min(args.e2e_streams, len(streams)))
    depth = 2 if len(streams) >= 2 * E and not args.e2e_serial else 1
    groups = [streams[i * E:(i + 1) * E] for i in range(depth)]
    streams = groups[0]
    B = E
    R = args.range
    n2 = (2 * R) ** 2
    w, h = args.width, args.height

    def pinned(nbytes, dtype=np.uint8):
        p = lib.thorcuda_host_alloc(nbytes)
        return np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(p)).view(dtype), p

    hosts = [[] for _ in groups]
    for gi, s in [(gi, s) for gi, g in enumerate(groups) for s in g]:
        host = hosts[gi]
        j = s.job
        d = {"allocs": []}

        def take(name, arr_bytes, dtype=np.uint8):
            a, p = pinned(arr_bytes, dtype)
            d[name] = a
            d["allocs"].append(p)
            return a
        for i, nm in enumerate(("oy", "ou", "ov")):
            src = np.ascontiguousarray(j["org"].plane(i))
            take(nm, src.size)[:] = src.reshape(-1)
        for i, nm in enumerate(("ry", "ru", "rv")):
            take(nm, (w * h) >> (0 if i == 0 else 2))
        for r in range(2):
            for bs in (8, 16, 32, 64):
                take(f"t{r}_{bs}", (w // bs) * (h // bs) * n2 * (2 if bs <= 16 else 4))
        take("half", j["cost"].size * 8)
        take("quarter", j["cost"].size * 8)
        take("coeffq", j["pool_entries"] * 2 + 64)
        take("cbp", j["tb"].size)
        take("ssd", j["tb"].size * 4)
        take("bits", max(1, (w // 64) * (h // 64)))
        host.append(d)

    h2d = d2h = 0

    phase = {}

    def mark(name, t_prev):
        t = time.perf_counter()
        phase[name] = phase.get(name, 0.0) + (t - t_prev)
        return t

    pending = [None] * depth

    def retire(g):
        if pending[g] is None:
            return
        tp = time.perf_counter()
        fence, lists = pending[g]
        tb.check(lib.thorcuda_fence_wait(fence))            # this group's results are in host memory
        for l in lists:
            l.close()
        pending[g] = None
        mark("wait", tp)

    def step(g=0):
        nonlocal h2d, d2h
        h2d = d2h = 0
        lists = []
        streams, host = groups[g], hosts[g]
        retire(g)
        tp = time.perf_counter()
        for s, d in zip(streams, host):
            j = s.job
            tb.check(lib.thorcuda_pic_upload(s.org.handle, d["oy"].ctypes.data, d["ou"].ctypes.data, d["ov"].ctypes.data, w, w // 2))
            h2d += w * h * 3 // 2 + (w // 4) * (h // 4) * 2
        tp = mark("upload_submit", tp)
        tb.ops.sad_table_build([x for s in streams for x in s.slots], [s.org for s in streams for _ in range(2)],
                               [r for s in streams for r in s.refs], R)
        # the table fetches go first: they run on the library's copy stream and drain while the kernels below execute
        for s, d in zip(streams, host):
            for r in range(2):
                for bs in (8, 16, 32, 64):
                    a = d[f"t{r}_{bs}"]
                    tb.check(lib.thorcuda_sad_table_fetch(s.slots[r], bs, a.ctypes.data, a.nbytes))
                    d2h += a.nbytes
        tp = mark("tables_submit", tp)
        for s in streams:                       # side information is only needed by the frame tail: pack it while the tables drain
            tb.ops.blockinfo_upload(s.work, s.job["blockinfo"])
        for s, d in zip(streams, host):
            j = s.job
            lc = tb.ops.TaskList(tb.LIST_COST, s.pics, j["cost"])
            lp = tb.ops.TaskList(tb.LIST_PRED, s.pics, j["pred"])
            lt = tb.ops.TaskList(tb.LIST_TB, s.pics, j["tb"])
            lists += [lc, lp, lt]
            h2d += j["cost"].nbytes + j["pred"].nbytes + j["tb"].nbytes
            lc.cost(tb.COST_FASTHALF, d["half"])
            lc.cost(tb.COST_FASTQUARTER, d["quarter"])
            lp.predict(1)
            lt.encode(d["coeffq"], d["cbp"], d["ssd"])
            d2h += 16 * j["cost"].size + 2 * lt.pool_entries + 5 * j["tb"].size
        tp = mark("lists_submit", tp)
        jobs = (tb.ops.LfJob * B)()
        for k, (s, d) in enumerate(zip(streams, host)):
            jobs[k] = tb.ops.LfJob(s.work.handle, s.org.handle, s.newref.handle, s.job["qp"], 15, None, d["bits"].ctypes.data)
        tb.check(lib.thorcuda_loopfilter_frames(B, C.byref(jobs)))
        for s, d in zip(streams, host):
            tb.check(lib.thorcuda_pic_download(s.work.handle, d["ry"].ctypes.data, d["ru"].ctypes.data, d["rv"].ctypes.data, w, w // 2))
            d2h += w * h * 3 // 2 + d["bits"].size
        tp = mark("tail_submit", tp)
        fence = lib.thorcuda_fence_record()
        if fence < 0:
            tb.check(-1)
        pending[g] = (fence, lists)

    def drain():
        for g in range(depth):
            retire(g)

    for _ in range(2):
        for g in range(depth):
            step(g)
    drain()
    barrier()
    phase.clear()
    t0 = time.perf_counter()
    n = max(2, args.e2e_steps) * depth
    for it in range(n):
        step(it % depth)
    drain()
    dt = time.perf_counter() - t0
    if dist is not None:
        import torch
        t = torch.tensor([dt], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    host = hosts[0]
    for d in host:
        d["keep"] = True
    measure_e2e.host = host          # kept for the parity check
    return {"value": B * world * n / dt, "unit": "frames/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "frames_per_step": B, "steps": n, "ms_per_step": 1e3 * dt / n, "groups_in_flight": depth,
            "host_ms_per_step": {k: round(1e3 * v / n, 3) for k, v in phase.items()},
            # the step moves this many bytes over PCIe per second: compare with the link (Gen5 x16: ~57 GB/s measured D2H)
            "pcie_GBps": {"d2h": round(d2h * n / dt / 1e9, 1), "h2d": round(h2d * n / dt / 1e9, 2)},
            "timing": "host wall clock from the first submit until the last group's results are in host memory, max over ranks (includes "
                      "PCIe copies and list building); a group's step is submitted only after its previous results have arrived"}


def cpu_baseline_and_parity(args, tb, stream):
    """Time the CPU arm on a bounded sample (rank 0, N=1) and compare its outputs with the GPU's byte for byte."""
    lib, kind = load_cpu_pipe()
    lib.pipe_init()
    threads = host_threads()
    fr = CpuFrame(stream.job, lib, threads)
    frames = max(1, args.cpu_frames)
    t0 = time.perf_counter()
    per = None
    for _ in range(frames):
        per = fr.run()
    dt = time.perf_counter() - t0
    d = measure_e2e.host[0]          # stream 0's host outputs of the last e2e step
    j = stream.job
    checks = {}
    for r in range(2):
        for bs in (8, 16, 32, 64):
            checks[f"sad_table_ref{r}_{bs}"] = bool(np.array_equal(d[f"t{r}_{bs}"].view(fr.tab[r][bs].dtype), fr.tab[r][bs]))
    checks["fasthalf"] = bool(np.array_equal(d["half"].view(fr.half.dtype), fr.half))
    checks["fastquarter"] = bool(np.array_equal(d["quarter"].view(fr.quarter.dtype), fr.quarter))
    checks["coeffq"] = bool(np.array_equal(d["coeffq"].view(np.int16)[:fr.coeffq.size], fr.coeffq))
    checks["cbp"] = bool(np.array_equal(d["cbp"], fr.cbp))
    checks["ssd"] = bool(np.array_equal(d["ssd"].view(np.uint32), fr.ssd))
    checks["clpf_bits"] = bool(np.array_equal(d["bits"], fr.bits))
    w, h = j["width"], j["height"]
    checks["rec_y"] = bool(np.array_equal(d["ry"].reshape(h, w), fr.rec.plane(0)))
    checks["rec_u"] = bool(np.array_equal(d["ru"].reshape(h // 2, w // 2), fr.rec.plane(1)))
    checks["rec_v"] = bool(np.array_equal(d["rv"].reshape(h // 2, w // 2), fr.rec.plane(2)))
    ny, nu, nv = stream.newref.download_padded()
    wy, wc = w + 192, w // 2 + 96
    checks["new_reference"] = bool(np.array_equal(ny[:, :wy], fr.newref.buf[0][:, :wy]) and np.array_equal(nu[:, :wc], fr.newref.buf[1][:, :wc])
                                   and np.array_equal(nv[:, :wc], fr.newref.buf[2][:, :wc]))
    ok = all(checks.values())
    out = {"value": frames / dt, "unit": "frames/s", "cores": threads, "kind": kind,
           "sample": f"{frames} frame(s) of the same {w}x{h} frame job; stages 1-4 banded over {threads} threads, stage 5 single-threaded",
           "seconds_per_stage": {k: round(v, 4) for k, v in per.items()},
           "parity_vs_gpu": "bit-identical" if ok else "MISMATCH: " + ",".join(k for k, v in checks.items() if not v)}
    if not ok:
        sys.stderr.write("PARITY FAILURE: " + json.dumps(checks) + "\n")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="thorcuda", choices=["thorcuda", "reference"])
    ap.add_argument("--streams", type=int, default=16, help="independent 1080p streams per GPU = frames per step")
    ap.add_argument("--distinct-jobs", type=int, default=2, help="distinct synthetic frame jobs cycled over the streams")
    ap.add_argument("--e2e-streams", type=int, default=4)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--e2e-serial", action="store_true", help="e2e leg: one group of streams, wait for every step before the next is submitted")
    ap.add_argument("--decoder", action="store_true", help="also time Thordec vs Thordec with the GPU frame tail / the whole reconstruction on the GPU (adds ~1 min)")
    ap.add_argument("--encoder", action="store_true", help="also time Thorenc vs Thorenc with the GPU frame tail (LDB medium, adds ~1 min)")
    ap.add_argument("--encoder-frames", type=int, default=12)
    ap.add_argument("--encoder-noise", type=float, default=5.0, help="sensor noise (grey levels) of the encoder leg's clip: with the decoder "
                    "leg's clean clip (2.0) the P frames are skipped almost entirely")
    ap.add_argument("--decoder-frames", type=int, default=168)
    ap.add_argument("--decoder-hard", action="store_true", help="decoder leg on the encoder leg's hard clip (layers drifting against each "
                    "other, moving patch, noise 5): several times the bits per frame of the default clip")
    ap.add_argument("--decoder-config", default="LDB_low_complexity", choices=sorted(DEC_CONFIGS),
                    help="stream of the decoder leg; RA_low_complexity has B frames with interpolated references (one interpolate_frames per B frame)")
    ap.add_argument("--cpu-frames", type=int, default=2)
    ap.add_argument("--range", type=int, default=16)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    else:
        run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
