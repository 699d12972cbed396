"""Test-only helpers: ctypes views of the two CPU checkers and synthetic input generators.

* ``oracle/_ref/libthorref.so`` -- the UNMODIFIED reference (awakecoding/thor) compiled by
  ``oracle/Makefile`` from /root/reference (kind "reference"); may be absent.
* ``oracle/libthor_oracle.so``  -- our C restatement (kind "port"); always buildable.

Nothing in the product package imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

# reference common/types.h:127-135 (sizeof == 44, probed)
BLOCKINFO = np.dtype([
    ("mode", "<i4"), ("cbp_y", "<i4"), ("cbp_u", "<i4"), ("cbp_v", "<i4"),
    ("size", "u1"), ("tb_split", "u1"), ("_pad", "u1", (2,)), ("pb_part", "<i4"),
    ("mv0", "<i2", (2,)), ("mv1", "<i2", (2,)),
    ("ref_idx0", "<u4"), ("ref_idx1", "<u4"), ("bipred_flag", "<u4")], align=False)
assert BLOCKINFO.itemsize == 44

MODE_SKIP, MODE_INTRA, MODE_INTER, MODE_BIPRED, MODE_MERGE = range(5)


class YuvFrame(C.Structure):  # reference common/types.h:41-59
    _fields_ = [("y", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p)] + [
        (n, C.c_int) for n in ("width", "height", "stride_y", "stride_c", "offset_y", "offset_c",
                               "pad_hor_y", "pad_hor_c", "pad_ver_y", "pad_ver_c", "area_y", "area_c",
                               "frame_num")]


def build_oracle():
    subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, capture_output=True)


_cache = {}


def load_oracle():
    if "orc" not in _cache:
        p = os.path.join(ORACLE_DIR, "libthor_oracle.so")
        if not os.path.exists(p):
            build_oracle()
        _cache["orc"] = C.CDLL(p)
    return _cache["orc"]


def load_ref():
    """The compiled reference, or None when oracle/_ref was not built (no /root/reference)."""
    if "ref" not in _cache:
        p = os.path.join(ORACLE_DIR, "_ref", "libthorref.so")
        _cache["ref"] = C.CDLL(p) if os.path.exists(p) else None
    return _cache["ref"]


def set_ref_simd(lib, on):
    C.c_int.in_dll(lib, "use_simd").value = int(on)


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Picture:
    """A padded I420 picture with the reference's geometry (common/common_frame.c:324-351)."""

    def __init__(self, width, height, pad_y=0, pad_c=0, rng=None):
        self.width, self.height, self.pad_y, self.pad_c = width, height, pad_y, pad_c
        self.stride_y = (width + 2 * pad_y + 15) & ~15
        self.stride_c = (width // 2 + 2 * pad_c + 15) & ~15
        self.rows_y, self.rows_c = height + 2 * pad_y, height // 2 + 2 * pad_c
        self.buf_y = np.zeros((self.rows_y, self.stride_y), np.uint8)
        self.buf_u = np.zeros((self.rows_c, self.stride_c), np.uint8)
        self.buf_v = np.zeros((self.rows_c, self.stride_c), np.uint8)
        if rng is not None:
            for b in (self.buf_y, self.buf_u, self.buf_v):
                b[...] = rng.integers(0, 256, b.shape, dtype=np.uint8)

    def plane(self, i):
        b, p = ((self.buf_y, self.pad_y), (self.buf_u, self.pad_c), (self.buf_v, self.pad_c))[i]
        h, w = (self.height, self.width) if i == 0 else (self.height // 2, self.width // 2)
        return b[p:p + h, p:p + w]

    def origin_ptr(self, i):
        b, p = ((self.buf_y, self.pad_y), (self.buf_u, self.pad_c), (self.buf_v, self.pad_c))[i]
        return b.ctypes.data + p * b.shape[1] + p

    def set_visible(self, y, u, v):
        self.plane(0)[...] = y
        self.plane(1)[...] = u
        self.plane(2)[...] = v

    def as_ref_struct(self):
        f = YuvFrame()
        f.y, f.u, f.v = self.origin_ptr(0), self.origin_ptr(1), self.origin_ptr(2)
        f.width, f.height, f.stride_y, f.stride_c = self.width, self.height, self.stride_y, self.stride_c
        f.pad_hor_y = f.pad_ver_y = self.pad_y
        f.pad_hor_c = f.pad_ver_c = self.pad_c
        return f

    def pad_edges(self):
        """edge replication into the border (reference pad_yuv_frame, common/common_frame.c:405-462)"""
        for i, (b, p) in enumerate(((self.buf_y, self.pad_y), (self.buf_u, self.pad_c), (self.buf_v, self.pad_c))):
            if p:
                b[...] = np.pad(self.plane(i), p, mode="edge")[:b.shape[0], :b.shape[1]] if b.shape[1] == self.plane(i).shape[1] + 2 * p \
                    else np.pad(np.pad(self.plane(i), p, mode="edge"), ((0, 0), (0, b.shape[1] - self.plane(i).shape[1] - 2 * p)))

    def copy(self):
        q = Picture(self.width, self.height, self.pad_y, self.pad_c)
        q.buf_y[...], q.buf_u[...], q.buf_v[...] = self.buf_y, self.buf_u, self.buf_v
        return q


def smooth_frame(width, height, seed, shift=(0, 0)):
    """Seeded synthetic I420 frame: blurred noise texture (SURVEY 8d recipe, simplified)."""
    rng = np.random.default_rng(seed)
    big = rng.integers(0, 256, (height + 64, width + 64)).astype(np.float32)
    k = np.ones(5, np.float32) / 5
    for ax in (0, 1):
        big = np.apply_along_axis(lambda m: np.convolve(m, k, mode="same"), ax, big)
    big = np.clip((big - 128) * 3 + 128, 0, 255)
    oy, ox = 32 + shift[0], 32 + shift[1]
    y = big[oy:oy + height, ox:ox + width]
    y = np.clip(y + rng.normal(0, 2, y.shape), 0, 255).astype(np.uint8)
    yy, xx = np.mgrid[0:height // 2, 0:width // 2]
    u = (128 + 60 * np.sin((xx + shift[1]) / 23.0) * np.cos(yy / 29.0)).astype(np.uint8)
    v = (128 + 60 * np.cos(xx / 27.0) * np.sin((yy + shift[0]) / 25.0)).astype(np.uint8)
    return y, u, v


def interp_pair(width, height, seed, motion=(3, 5), kind="scene"):
    """Two padded reference pictures (pad 96/48 like the codec's reference ring) for the temporal-interpolation tests:
    a textured background moving by `motion` (dy, dx) between them, a static band on the left (skip path), an object
    moving the other way, noise.  kind="noise": two unrelated pictures (the search wanders, vectors leave the picture)."""
    rng = np.random.default_rng(seed)
    pics = []
    for k in range(2):
        if kind == "noise":
            y = rng.integers(0, 256, (height, width), dtype=np.uint8)
            u = rng.integers(0, 256, (height // 2, width // 2), dtype=np.uint8)
            v = rng.integers(0, 256, (height // 2, width // 2), dtype=np.uint8)
        else:
            y, u, v = smooth_frame(width, height, seed)
            y, u, v = (np.roll(a, (k * motion[0] // d, k * motion[1] // d), (0, 1)).copy() for a, d in ((y, 1), (u, 2), (v, 2)))
            if k == 1:
                y0 = pics[0].plane(0)
                band = max(16, width // 4)
                y[:, :band] = y0[:, :band]                                   # static part
                oh, ow = height // 3, width // 3
                oy, ox = height // 3, width // 2
                y[oy:oy + oh, ox:ox + ow] = y0[oy + 2:oy + 2 + oh, ox - 6:ox - 6 + ow]   # an object moving against the background
            y = np.clip(y.astype(np.int16) + rng.integers(-2, 3, y.shape), 0, 255).astype(np.uint8)
        p = Picture(width, height, 96, 48)
        p.set_visible(y, u, v)
        p.pad_edges()
        pics.append(p)
    return pics


def random_blockinfo(width, height, rng, p_split=0.55, intra_frac=0.2):
    """Quad-tree consistent per-4x4 side information with a realistic mode mix."""
    bw, bh = width // 4, height // 4
    bi = np.zeros((bh, bw), BLOCKINFO)

    def leaf(y, x, size):
        r = rng.random()
        mode = (MODE_SKIP if r < 0.3 else MODE_INTER if r < 0.6 else MODE_BIPRED if r < 0.7 else
                MODE_MERGE if r < 1.0 - intra_frac else MODE_INTRA)
        e = np.zeros((), BLOCKINFO)
        e["mode"] = mode
        e["size"] = size
        if mode != MODE_SKIP:
            e["cbp_y"], e["cbp_u"], e["cbp_v"] = rng.random() < 0.5, rng.random() < 0.4, rng.random() < 0.4
            e["tb_split"] = rng.random() < 0.15
        if mode == MODE_INTER:
            e["pb_part"] = rng.integers(0, 4) if rng.random() < 0.3 else 0
        if mode != MODE_INTRA:
            e["mv0"] = np.clip(rng.normal(0, 4, 2).round(), -200, 200).astype(np.int16)
            if mode == MODE_BIPRED:
                e["mv1"] = np.clip(rng.normal(0, 4, 2).round(), -200, 200).astype(np.int16)
                e["bipred_flag"] = 1
        bi[y // 4:min(y + size, height) // 4, x // 4:min(x + size, width) // 4] = e

    def rec(y, x, size):
        if y >= height or x >= width:
            return
        fits = y + size <= height and x + size <= width
        if size > 8 and (not fits or rng.random() < p_split):
            h = size // 2
            for dy, dx in ((0, 0), (0, h), (h, 0), (h, h)):
                rec(y + dy, x + dx, h)
        else:
            leaf(y, x, size)

    for y in range(0, height, 64):
        for x in range(0, width, 64):
            rec(y, x, 64)
    return bi.reshape(-1)


def blocky_frame(width, height, seed):
    """Smooth gradients + per-8x8 DC steps + mild noise: a decoded-looking picture whose block
    edges straddle the deblocking activity threshold (so filtered and unfiltered edges both occur)."""
    rng = np.random.default_rng(seed)
    out = []
    for (w, h, b) in ((width, height, 8), (width // 2, height // 2, 4)):
        yy, xx = np.mgrid[0:h, 0:w]
        base = 128 + 50 * np.sin(xx / 37.0 + seed) + 40 * np.cos(yy / 23.0)
        steps = rng.integers(-6, 7, (h // b + 1, w // b + 1))
        base = base + np.kron(steps, np.ones((b, b)))[:h, :w]
        tex = rng.random((h // 32 + 1, w // 32 + 1)) < 0.3          # some busy regions (d >= beta)
        noise = rng.normal(0, 1.2, (h, w)) + np.kron(tex, np.ones((32, 32)))[:h, :w] * rng.normal(0, 12, (h, w))
        out.append(np.clip(base + noise, 0, 255).astype(np.uint8))
    return out[0], out[1], out[1][::-1].copy()


# ------------------------------------------------------------------------------------------------
# End-to-end helpers: the reference binaries built by oracle/Makefile, its shipped operating points as
# command-line flags (values of /root/reference/config_*.txt, which does not exist on the GPU box), and a
# seeded synthetic clip.
REF_DIR = os.path.join(ORACLE_DIR, "_ref")

# The 15 shipped operating points (config_<family>_<level>.txt of the reference) as command-line options: the files do not
# exist on the GPU box.  tests/test_hooks.py::test_thor_configs_match_reference compares them with the files where the
# reference tree is mounted.  "HDB16" / "RA16" are 16-picture B hierarchies (GOP structures), 8-bit like the rest.
_B8 = ["-dqpI", "-2", "-dqpB0", "3", "-dqpB1", "1", "-dqpB2", "0", "-mqpP", "1.2", "-mqpB", "1.2", "-mqpB0", "1.1", "-mqpB1", "1.2",
       "-mqpB2", "1.3", "-lambda_coeffI", "0.8", "-lambda_coeffP", "1.2", "-lambda_coeffB", "1.2", "-lambda_coeffB0", "1.2",
       "-lambda_coeffB1", "1.2", "-lambda_coeffB2", "1.2"]
_B16 = ["-dqpI", "-2", "-dqpB0", "2", "-dqpB1", "1", "-dqpB2", "0", "-dqpB3", "0", "-mqpP", "1.2", "-mqpB", "1.2", "-mqpB0", "1.075",
        "-mqpB1", "1.15", "-mqpB2", "1.225", "-mqpB3", "1.3", "-lambda_coeffI", "0.8", "-lambda_coeffP", "1.2", "-lambda_coeffB", "1.2",
        "-lambda_coeffB0", "1.2", "-lambda_coeffB1", "1.2", "-lambda_coeffB2", "1.2", "-lambda_coeffB3", "1.2"]
_FAMILY = {
    "LDB": ["-HQperiod", "12", "-mqpP", "1.2", "-dqpI", "-2", "-lambda_coeffI", "0.8", "-lambda_coeffP", "1.2"],
    "HDB": ["-HQperiod", "1", "-num_reorder_pics", "7", "-interp_ref", "1"] + _B8,
    "RA": ["-HQperiod", "1", "-num_reorder_pics", "7", "-interp_ref", "1", "-intra_period", "64"] + _B8,
    "HDB16": ["-HQperiod", "1", "-num_reorder_pics", "15", "-interp_ref", "1"] + _B16,
    "RA16": ["-HQperiod", "1", "-num_reorder_pics", "15", "-interp_ref", "1", "-intra_period", "64"] + _B16,
}


def _level(level, family):
    if level == "low_complexity":
        return ["-intra_rdo", "0", "-enable_tb_split", "0", "-enable_pb_split", "0", "-early_skip_thr", "1.0", "-max_num_ref", "2",
                "-use_block_contexts", "1", "-enable_bipred", "0" if family == "LDB" else "1", "-encoder_speed", "2"]
    if level == "medium_complexity":
        return ["-intra_rdo", "1", "-enable_tb_split", "0", "-enable_pb_split", "0", "-early_skip_thr", "0.8", "-max_num_ref", "2",
                "-use_block_contexts", "1", "-enable_bipred", "1", "-encoder_speed", "1"]
    return ["-intra_rdo", "1", "-enable_tb_split", "1", "-enable_pb_split", "1", "-early_skip_thr", "0.3", "-max_num_ref", "4",
            "-use_block_contexts", "1", "-enable_bipred", "1", "-encoder_speed", "0", "-max_delta_qp", "1"]


THOR_CONFIGS = {f"{fam}_{lvl}": opts + _level(lvl, fam) for fam, opts in _FAMILY.items()
                for lvl in ("low_complexity", "medium_complexity", "high_efficiency")}
assert len(THOR_CONFIGS) == 15


def write_clip(path, width, height, frames, seed):
    """Seeded I420 clip: a textured background translating by (1,2) px per frame plus noise."""
    with open(path, "wb") as f:
        for n in range(frames):
            y, u, v = smooth_frame(width, height, seed, shift=(n % 16, (2 * n) % 32))   # stays inside the texture margin
            for p in (y, u, v):
                f.write(np.ascontiguousarray(p).tobytes())


def write_hard_clip(path, width, height, frames, seed, noise=3.0):
    """Seeded I420 clip that makes an encoder work: a contrast-stretched texture translating by (1, 2) per frame, a second
    layer drifting the other way under the right half (no single vector predicts the blend), a patch of a quarter of the
    picture moving on its own by (5, 13), sensor noise, slow chroma waves.  Any size (multiples of 8)."""
    rng = np.random.default_rng(seed)
    big = rng.integers(0, 256, (height + 96, width + 96)).astype(np.float32)
    for ax in (0, 1):
        c = np.cumsum(np.pad(big, [(3, 2) if a == ax else (0, 0) for a in (0, 1)], mode="edge"), axis=ax)
        big = (np.take(c, range(5, c.shape[ax]), axis=ax) - np.take(c, range(0, c.shape[ax] - 5), axis=ax)) / 5
    big = np.clip((big - 128) * 3 + 128, 0, 255)
    ph, pw = max(16, height // 4 // 8 * 8), max(16, width // 4 // 8 * 8)
    yy, xx = np.mgrid[0:height // 2, 0:width // 2]
    with open(path, "wb") as f:
        for n in range(frames):
            dy, dx = n % 64, (2 * n) % 64
            y = big[32 + dy:32 + dy + height, 32 + dx:32 + dx + width].copy()
            ey, ex = (3 * n) % 64, 63 - n % 64
            y[:, width // 2:] = 0.5 * y[:, width // 2:] + 0.5 * big[32 + ey:32 + ey + height, 32 + ex:32 + ex + width][::-1, width // 2:]
            oy, ox = (5 * n) % (height - ph), (13 * n) % (width - pw)
            y[oy:oy + ph, ox:ox + pw] = big[:ph, :pw][::-1, ::-1]
            f.write(np.clip(y + rng.normal(0, noise, y.shape), 0, 255).astype(np.uint8).tobytes())
            f.write((128 + 60 * np.sin((xx + 2 * n) / 23.0) * np.cos(yy / 29.0)).astype(np.uint8).tobytes())
            f.write((128 + 60 * np.cos(xx / 27.0) * np.sin((yy + n) / 25.0)).astype(np.uint8).tobytes())


def run_thorenc(exe, cfg, yuv, width, height, frames, qp, out_prefix, timeout=600):
    """Run a Thorenc binary of oracle/_ref; returns (bitstream bytes, reconstruction bytes, seconds)."""
    import time
    bit, rec = out_prefix + ".bit", out_prefix + "_rec.yuv"
    cmd = [os.path.join(REF_DIR, exe), "-if", yuv, "-of", bit, "-rf", rec, "-stat", out_prefix + "_stat.txt", "-width", str(width),
           "-height", str(height), "-n", str(frames), "-f", "30", "-qp", str(qp)] + THOR_CONFIGS[cfg]
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        raise RuntimeError(f"{exe} failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
    return open(bit, "rb").read(), open(rec, "rb").read(), dt


def run_thordec(exe, bit, out_yuv, timeout=600):
    r = subprocess.run([os.path.join(REF_DIR, exe), bit, out_yuv], capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"{exe} failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
    return open(out_yuv, "rb").read()


# ---------------------------------------------------------------- bit-cost model helpers ----
class RefStream(C.Structure):      # enc/putbits.h:36-43
    _fields_ = [("bytesize", C.c_uint32), ("bytepos", C.c_uint32), ("bitstream", C.c_void_p), ("bitbuf", C.c_uint32), ("bitrest", C.c_uint32)]


def ref_stream(nbytes=1 << 16):
    buf = np.zeros(nbytes, np.uint8)
    s = RefStream(nbytes, 0, buf.ctypes.data, 0, 32)
    s._keep = buf
    return s


def ref_bitpos(s):                 # get_bit_pos, enc/putbits.c:131-134
    return 8 * s.bytepos + 32 - s.bitrest


def coeff_block(rng, size, kind):
    """quantised levels as the encoder leaves them: n x n array, only the low min(n,16)^2 corner populated"""
    q = min(size, 16)
    c = np.zeros((size, size), np.int16)
    if kind == "dc1":
        c[0, 0] = rng.choice([-1, 1])
    elif kind == "dc":
        c[0, 0] = rng.integers(-40, 41) or 3
    elif kind == "sparse":
        k = rng.integers(1, 6)
        for _ in range(k):
            c[rng.integers(0, q), rng.integers(0, q)] = rng.choice([-3, -2, -1, 1, 1, 1, 2, 5])
    elif kind == "lowfreq":
        m = min(q, 4)
        c[:m, :m] = rng.integers(-6, 7, (m, m))
    elif kind == "dense":
        c[:q, :q] = rng.integers(-2, 3, (q, q))
    elif kind == "large":
        c[:q, :q] = np.where(rng.random((q, q)) < 0.3, rng.integers(-3000, 3001, (q, q)), 0)
    elif kind == "last":
        c[q - 1, q - 1] = rng.choice([-1, 1, 2, -7])
        c[0, 0] = rng.integers(-2, 3)
    if not c.any():
        c[0, 1] = 2
    return c


COEFF_KINDS = ["dc1", "dc", "sparse", "lowfreq", "dense", "large", "last"]


# ---- write_block records (oracle/thor_oracle.h ORC_BS_*) ----
BS = {n: i for i, n in enumerate(
    "frame_type num_ref enable_bipred interp_ref num_intra_modes max_delta_qp size ctx_index ctx_cbp max_num_pb_part "
    "max_num_tb_part num_skip_vec num_merge_vec delta_qp mode intra_mode pb_part tb_split skip_idx ref_idx0 ref_idx1 "
    "cbp_y cbp_u cbp_v".split())}
BS["mvp"], BS["mv0"], BS["mv1"], BS_FIELDS = 24, 26, 34, 42
MODE_SKIP, MODE_INTRA, MODE_INTER, MODE_BIPRED, MODE_MERGE = range(5)


def load_ref_bits():
    lib = C.CDLL(os.path.join(REF_DIR, "libthorbits_ref.so"))
    lib.ref_write_block_bits.restype = C.c_int
    lib.ref_write_block_bits.argtypes = [C.c_void_p] * 4
    return lib


def random_block_syntax(rng, frame=None):
    """One block as write_block (enc/write_bits.c:364-650) may meet it: (record int32[42], coeff_y, coeff_u, coeff_v) with
    the coefficient arrays in block_param_t's layout (quadrants contiguous under tb_split).  Only combinations the
    reference encoder can produce: no intra mode outside the frame's set, tb_split only where max_num_tb_part allows it,
    a level wherever a cbp bit says so."""
    d = np.zeros(BS_FIELDS, np.int32)
    if frame is None:
        ft = int(rng.integers(0, 3))
        frame = dict(frame_type=ft, num_ref=0 if ft == 0 else int(rng.integers(1, 5)), enable_bipred=int(rng.integers(0, 2)),
                     interp_ref=0 if ft == 0 else int(rng.integers(0, 2)), num_intra_modes=int(rng.choice([4, 8, 10])),
                     max_delta_qp=int(rng.choice([0, 0, 4])))
    for k, v in frame.items():
        d[BS[k]] = v
    size = int(rng.choice([8, 16, 32, 64]))
    d[BS["size"]] = size
    d[BS["ctx_index"]] = rng.integers(0, 6)
    d[BS["ctx_cbp"]] = rng.integers(-1, 3)
    d[BS["max_num_pb_part"]] = rng.choice([1, 4])
    d[BS["max_num_tb_part"]] = rng.choice([1, 2])
    d[BS["num_skip_vec"]] = rng.integers(1, 5)
    d[BS["num_merge_vec"]] = rng.integers(1, 5)
    d[BS["delta_qp"]] = rng.integers(-4, 5)
    nref = frame["num_ref"]
    modes = [MODE_INTRA] if frame["frame_type"] == 0 else [MODE_SKIP, MODE_INTRA, MODE_INTER, MODE_MERGE] + ([MODE_BIPRED] if nref > 1 and frame["enable_bipred"] else [])
    mode = int(rng.choice(modes))
    d[BS["mode"]] = mode
    nim = frame["num_intra_modes"]
    d[BS["intra_mode"]] = rng.integers(0, 4) if nim <= 4 else rng.choice([0, 2, 3, 4, 6, 7, 8, 9]) if nim <= 8 else rng.integers(0, 10)
    d[BS["pb_part"]] = rng.integers(0, 4) if d[BS["max_num_pb_part"]] > 1 and mode in (MODE_INTER, MODE_BIPRED) else 0
    d[BS["skip_idx"]] = rng.integers(0, max(1, d[BS["num_skip_vec" if mode == MODE_SKIP else "num_merge_vec"]]))
    d[BS["ref_idx0"]] = rng.integers(0, max(1, nref))
    d[BS["ref_idx1"]] = rng.integers(0, max(1, nref))
    d[BS["mvp"]:BS["mvp"] + 18] = np.where(rng.random(18) < 0.3, 0, rng.integers(-300, 301, 18))
    if rng.random() < 0.2:
        d[BS["mv0"]:BS["mv0"] + 2] = d[BS["mvp"]:BS["mvp"] + 2]
    split = mode in (MODE_INTRA, MODE_INTER) and d[BS["max_num_tb_part"]] > 1 and rng.random() < 0.4
    d[BS["tb_split"]] = int(split)
    ty = np.zeros(size * size, np.int16)
    tu, tv = np.zeros(size * size // 4, np.int16), np.zeros(size * size // 4, np.int16)
    kind = lambda: COEFF_KINDS[int(rng.integers(0, len(COEFF_KINDS)))]
    if not split:
        cbp = rng.integers(0, 2, 3)
        for flag, arr, n in zip(cbp, (ty, tu, tv), (size, size // 2, size // 2)):
            if flag:
                arr[:] = coeff_block(rng, n, kind()).ravel()
        d[BS["cbp_y"]:BS["cbp_y"] + 3] = cbp
    elif size > 8:
        h, q = size // 2, size // 4
        cbp = rng.integers(0, 16, 3)
        for mask, arr, n in zip(cbp, (ty, tu, tv), (h, q, q)):
            for i in range(4):
                if (mask >> (3 - i)) & 1:
                    arr[i * n * n:(i + 1) * n * n] = coeff_block(rng, n, kind()).ravel()
        d[BS["cbp_y"]:BS["cbp_y"] + 3] = cbp
    else:
        cy = int(rng.integers(0, 16))
        for i in range(4):
            if (cy >> (3 - i)) & 1:
                ty[i * 16:(i + 1) * 16] = coeff_block(rng, 4, kind()).ravel()
        cuv = rng.integers(0, 2, 2)
        for flag, arr in zip(cuv, (tu, tv)):
            if flag:
                arr[:] = coeff_block(rng, 4, kind()).ravel()
        d[BS["cbp_y"]] = cy
        d[BS["cbp_u"]:BS["cbp_u"] + 2] = cuv
    return d, ty, tu, tv


def block_tbs(d):
    """The transform blocks of a write_block record in the order write_block codes them (enc/write_bits.c:572-641):
    list of (plane 0/1/2, quadrant, transform size, coded) -- empty for a skip block."""
    size, split = int(d[BS["size"]]), int(d[BS["tb_split"]])
    cbp = [int(d[BS["cbp_y"]]), int(d[BS["cbp_u"]]), int(d[BS["cbp_v"]])]
    if d[BS["mode"]] == MODE_SKIP:
        return []
    if not split:
        return [(p, 0, size if p == 0 else size // 2, cbp[p] & 1) for p in range(3)]
    if size > 8:
        return [(p, q, size // 2 if p == 0 else size // 4, (cbp[p] >> (3 - q)) & 1) for q in range(4) for p in range(3)]
    return [(0, q, 4, (cbp[0] >> (3 - q)) & 1) for q in range(4)] + [(1, 0, 4, cbp[1] & 1), (2, 0, 4, cbp[2] & 1)]


def pack_blocks(blocks, block_dtype, tb_dtype):
    """[(record, coeff_y, coeff_u, coeff_v)] -> (thorcuda_block_syntax array, thorcuda_tb_task array with coeff_offset
    assigned, level pool in libthorcuda's layout: the low q x q corner of every transform block, 8-entry aligned)"""
    recs = np.zeros(len(blocks), block_dtype)
    tasks, levels, off = [], [], 0
    for i, (d, cy, cu, cv) in enumerate(blocks):
        r = recs[i]
        for name in block_dtype.names:
            if name in ("mvp", "mv0", "mv1"):
                r[name] = d[BS[name]:BS[name] + r[name].size].reshape(r[name].shape)
            elif name != "tb_first":
                r[name] = d[BS[name]]
        tbs = block_tbs(d)
        r["tb_first"] = len(tasks) if tbs else 0xffffffff
        intra = int(d[BS["mode"]] == MODE_INTRA) << 1
        for plane, quad, n, coded in tbs:
            src = (cy, cu, cv)[plane]
            q = min(n, 16)
            c = src[quad * n * n:(quad + 1) * n * n].reshape(n, n)[:q, :q]
            t = np.zeros((), tb_dtype)
            t["size"], t["coeff_type"], t["coeff_offset"] = n, intra | (plane > 0), off
            tasks.append(t)
            levels.append((off, c.reshape(-1).copy()))
            off += (q * q + 7) & ~7
    pool = np.zeros(max(off, 8), np.int16)
    for o, c in levels:
        pool[o:o + c.size] = c
    return recs, np.array(tasks, tb_dtype) if tasks else np.zeros(0, tb_dtype), pool


BETA_TABLE = [0] * 16 + [6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50,
                         52, 54, 56, 58, 60, 62, 64]     # common/common_frame.c:36-39


def deblock_threshold_case(width, height, qp):
    """Known-answer picture for the luma deblocking decision `d < beta` (common/common_frame.c:128,214): flat 8x8 tiles
    in a 100/112 checkerboard, every tile a coded inter block, and in every tile one pixel that makes the activity measure
    of the edge to its right (pixel (2,7) = p0 of row 2) resp. below it (pixel (7,2) = p0 of column 2) EXACTLY
    beta-2 .. beta+2, cycling over the tiles.  Neither pixel is read by the other direction's measure nor written by the
    other direction's filter, so the decisions do not depend on the order the edges are visited in.
    Returns (y, u, v, blockinfo, dv, dh) with dv/dh the designed d of every tile's right / bottom edge."""
    beta = BETA_TABLE[qp]
    th, tw = height // 8, width // 8
    ty, tx = np.mgrid[0:th, 0:tw]
    dv = np.maximum(beta + (ty * 3 + tx) % 5 - 2, 0)
    dh = np.maximum(beta + (ty + tx * 2 + 1) % 5 - 2, 0)
    y = np.kron(100 + 12 * ((ty + tx) & 1), np.ones((8, 8), int))
    y[2::8, 7::8] += dv
    y[7::8, 2::8] += dh
    y = y.astype(np.uint8)
    u = np.full((height // 2, width // 2), 128, np.uint8)
    bi = np.zeros((height // 4) * (width // 4), BLOCKINFO)
    bi["mode"], bi["size"], bi["cbp_y"] = MODE_INTER, 8, 1
    return y, u, u.copy(), bi, dv, dh


# streams of tests/test_gpu_integration.py::test_thordec_with_gpu_frame_tail (clip seed 77); tests/golden/make_golden.py
# commits the reference's hashes of each
DECODER_STREAMS = [
    ("LDB_medium_complexity", 416, 240, 5, 30),     # P frames, bi-pred blocks, CLPF bits per superblock; 240 = 3*64 + 48
    ("LDB_low_complexity", 200, 136, 4, 36),        # partial superblocks right and bottom
    ("RA_high_efficiency", 128, 128, 9, 32),        # B frames, interpolated references, tb/pb split, delta qp
    ("RA_high_efficiency", 192, 136, 17, 38),       # two sub-GOPs of the B hierarchy, partial superblock row, coarse qp
    ("RA_low_complexity", 416, 240, 17, 32),        # 4-level interpolation pyramid (device-side in Thordec_gpu), 240 = 3*64 + 48
    ("HDB16_low_complexity", 256, 192, 33, 34),     # two 16-frame sub-GOPs: interpolation ratios 2..16, four B levels
    ("RA16_medium_complexity", 192, 128, 17, 30),
    ("LDB_medium_complexity", 136, 72, 14, 22),     # fine qp (dense coefficients), HQ-period frame, 33-slot ring not yet wrapped
    # the rest of the 15 shipped configurations ("all 15 configs"), small pictures
    ("LDB_high_efficiency", 128, 72, 6, 30),        # four references, tb/pb split, delta qp, full search
    ("HDB_low_complexity", 128, 128, 17, 34),
    ("HDB_medium_complexity", 128, 72, 9, 30),
    ("HDB_high_efficiency", 128, 128, 9, 33),
    ("RA_medium_complexity", 200, 136, 9, 31),
    ("HDB16_medium_complexity", 128, 72, 17, 32),
    ("HDB16_high_efficiency", 128, 72, 17, 35),
    ("RA16_low_complexity", 136, 128, 33, 36),
    ("RA16_high_efficiency", 128, 128, 17, 29),
]
