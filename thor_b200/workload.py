"""Synthetic frame jobs for the pixel pipeline (host logic, pure numpy; no pixels are computed here
that the pipeline is supposed to compute).

A *frame job* is everything the reference's encoder hands to its pixel kernels while coding one
1080p P-frame with the LDB medium-complexity settings (2 references, bi-prediction on, qp 32,
`encoder_speed 1`), with the encoder's decisions (partition, modes, vectors) replaced by a seeded
synthetic draw so that the same job can be given to libthorcuda and to the CPU reference:

  stage 1  integer-pel SAD tables of the frame against both references   (sad_calc)
  stage 2  bilinear half-/quarter-pel costs of every 16x16 block          (sad_calc_fasthalf/fastquarter)
  stage 3  inter prediction of every prediction block, Y/U/V, uni and bi  (get_inter_prediction_*)
  stage 4  residual -> DCT -> quant -> dequant -> IDCT -> reconstruct     (encode_and_reconstruct_block_inter)
  stage 5  deblock Y/UV -> CLPF with encoder decision -> new reference    (encode_frame.c:150-177)

Picture index convention of all task lists: 0 original, 1 ref0, 2 ref1, 3 work picture.
"""
import numpy as np

from .lib import COST_TASK, PRED_TASK, TB_TASK

# reference common/types.h:127-135 (sizeof == 44)
BLOCKINFO = np.dtype([
    ("mode", "<i4"), ("cbp_y", "<i4"), ("cbp_u", "<i4"), ("cbp_v", "<i4"),
    ("size", "u1"), ("tb_split", "u1"), ("_pad", "u1", (2,)), ("pb_part", "<i4"),
    ("mv0", "<i2", (2,)), ("mv1", "<i2", (2,)),
    ("ref_idx0", "<u4"), ("ref_idx1", "<u4"), ("bipred_flag", "<u4")], align=False)
MODE_SKIP, MODE_INTRA, MODE_INTER, MODE_BIPRED, MODE_MERGE = range(5)
PAD_Y, PAD_C = 96, 48          # reference enc/mainenc.c:165, common/global.h:64


def aligned_zeros(shape, dtype=np.uint8, align=64):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    raw = np.zeros(n + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n].view(dtype).reshape(shape)


def _box_blur(a, k):
    c = np.cumsum(np.pad(a, ((k, k), (0, 0)), mode="edge"), axis=0)
    a = (c[2 * k:] - c[:-2 * k]) / (2 * k)
    c = np.cumsum(np.pad(a, ((0, 0), (k, k)), mode="edge"), axis=1)
    return (c[:, 2 * k:] - c[:, :-2 * k]) / (2 * k)


def texture(width, height, seed):
    """Large blurred-noise luma texture the frames of one stream are cut from (SURVEY 8d recipe)."""
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (height + 128, width + 128)).astype(np.float32)
    t = _box_blur(_box_blur(t, 2), 1)
    return np.clip((t - 128) * 3 + 128, 0, 255)


def frame_from_texture(tex, width, height, shift, rng, noise=2.0):
    oy, ox = 64 + shift[0], 64 + shift[1]
    y = tex[oy:oy + height, ox:ox + width]
    y = np.clip(y + rng.normal(0, noise, y.shape), 0, 255).astype(np.uint8)
    yy, xx = np.mgrid[0:height // 2, 0:width // 2]
    u = (128 + 60 * np.sin((xx + shift[1]) / 23.0) * np.cos(yy / 29.0) + rng.normal(0, 1, yy.shape)).clip(0, 255).astype(np.uint8)
    v = (128 + 60 * np.cos(xx / 27.0) * np.sin((yy + shift[0]) / 25.0) + rng.normal(0, 1, yy.shape)).clip(0, 255).astype(np.uint8)
    return y, u, v


class HostPicture:
    """Padded I420 picture in host memory with the reference geometry (common/common_frame.c:324-351)."""

    def __init__(self, width, height, pad_y=0, pad_c=0):
        self.width, self.height, self.pad_y, self.pad_c = width, height, pad_y, pad_c
        self.stride_y = (width + 2 * pad_y + 15) & ~15
        self.stride_c = (width // 2 + 2 * pad_c + 15) & ~15
        self.rows_y, self.rows_c = height + 2 * pad_y, height // 2 + 2 * pad_c
        self.buf = [aligned_zeros((self.rows_y, self.stride_y)), aligned_zeros((self.rows_c, self.stride_c)),
                    aligned_zeros((self.rows_c, self.stride_c))]

    def plane(self, i):
        p = self.pad_y if i == 0 else self.pad_c
        h, w = (self.height, self.width) if i == 0 else (self.height // 2, self.width // 2)
        return self.buf[i][p:p + h, p:p + w]

    def origin_ptr(self, i):
        p = self.pad_y if i == 0 else self.pad_c
        return self.buf[i].ctypes.data + p * self.buf[i].shape[1] + p

    def set_and_pad(self, planes):
        """Fill the visible area and replicate it into the borders (what pad_yuv_frame produces)."""
        for i, pl in enumerate(planes):
            p = self.pad_y if i == 0 else self.pad_c
            h, w = pl.shape
            self.buf[i][:h + 2 * p, :w + 2 * p] = np.pad(pl, p, mode="edge") if p else pl


def _quadtree(width, height, rng, p_split):
    out = []

    def rec(y, x, size):
        if y >= height or x >= width:
            return
        fits = y + size <= height and x + size <= width
        if size > 8 and (not fits or rng.random() < p_split[size]):
            h = size // 2
            for dy, dx in ((0, 0), (0, h), (h, 0), (h, h)):
                rec(y + dy, x + dx, h)
        elif fits:
            out.append((x, y, size))

    for y in range(0, height, 64):
        for x in range(0, width, 64):
            rec(y, x, 64)
    return out


def make_job(width=1920, height=1080, seed=42, qp=32, search_range=16):
    """One synthetic frame job (see module docstring).  Everything is deterministic in `seed`."""
    rng = np.random.default_rng(seed)
    tex = texture(width, height, seed)
    org = frame_from_texture(tex, width, height, (0, 0), rng)
    refs = []
    for k, sh in enumerate(((-2, -4), (-4, -8))):                   # the two previous frames of a slow pan
        hp = HostPicture(width, height, PAD_Y, PAD_C)
        y, u, v = frame_from_texture(tex, width, height, sh, rng, noise=2.5)
        hp.set_and_pad((y, u, v))
        refs.append(hp)
    horg = HostPicture(width, height)
    horg.set_and_pad(org)

    cbs = _quadtree(width, height, rng, {64: 0.75, 32: 0.6, 16: 0.45})
    bi = np.zeros((height // 4, width // 4), BLOCKINFO)
    pred, tbs, cost = [], [], []

    def clip_mv(v, pos, size, lim):                                    # keep the block within 80 px of the frame, like clip_mv()
        return int(np.clip(v, -4 * (pos + 80), 4 * (lim + 80 - pos - size)))

    for (x, y, s) in cbs:
        r = rng.random()
        mode = MODE_SKIP if r < 0.30 else MODE_INTER if r < 0.68 else MODE_BIPRED if r < 0.83 else MODE_MERGE if r < 0.96 else MODE_INTRA
        e = np.zeros((), BLOCKINFO)
        e["mode"], e["size"] = mode, s
        coded = mode != MODE_SKIP
        tb_split = coded and s <= 32 and rng.random() < 0.15
        part = int(rng.integers(1, 4)) if (mode == MODE_INTER and s <= 32 and rng.random() < 0.25) else 0
        if coded:
            e["cbp_y"], e["cbp_u"], e["cbp_v"] = rng.random() < 0.6, rng.random() < 0.4, rng.random() < 0.4
            e["tb_split"], e["pb_part"] = tb_split, part
        # vectors: true pan (8,16)/(16,32) quarter-pel + jitter
        mvs = []
        for k in range(2):
            base = (16 + 16 * k, 8 + 8 * k)
            mvs.append((clip_mv(base[0] + rng.normal(0, 6), x, s, width), clip_mv(base[1] + rng.normal(0, 6), y, s, height)))
        bipred = mode == MODE_BIPRED
        ref0 = int(rng.integers(0, 2))
        e["mv0"] = mvs[ref0]
        if bipred:
            e["mv1"] = mvs[1 - ref0]
            e["bipred_flag"] = 1
        e["ref_idx0"], e["ref_idx1"] = ref0, 1 - ref0
        bi[y // 4:(y + s) // 4, x // 4:(x + s) // 4] = e
        # prediction blocks (intra-marked blocks are predicted as inter here: the pixel job has no intra stage)
        if part == 1:
            pbs = [(x, y, s, s // 2), (x, y + s // 2, s, s // 2)]
        elif part == 2:
            pbs = [(x, y, s // 2, s), (x + s // 2, y, s // 2, s)]
        elif part == 3:
            pbs = [(x + dx, y + dy, s // 2, s // 2) for dy in (0, s // 2) for dx in (0, s // 2)]
        else:
            pbs = [(x, y, s, s)]
        for (px, py, pw, ph) in pbs:
            jit = (int(rng.integers(-2, 3)), int(rng.integers(-2, 3))) if part else (0, 0)
            m0 = (clip_mv(mvs[ref0][0] + jit[0], px, pw, width), clip_mv(mvs[ref0][1] + jit[1], py, ph, height))
            m1 = mvs[1 - ref0]
            pred.append((px, py, pw, ph, 1 + ref0, (2 - ref0) if bipred else -1, m0[0], m0[1], m1[0], m1[1], 0, 0, 3, 0))
        # transform blocks (enc/encode_block.c:1469-1532): luma CB or its four quarters, chroma half size, min 4
        if coded:
            if tb_split:
                fast = int(s == 64)
                tbs += [(x + dx, y + dy, s // 2, 0, qp, 0, fast, 0, 3, 3, 0) for dy in (0, s // 2) for dx in (0, s // 2)]
            else:
                tbs.append((x, y, s, 0, qp, 0, int(s == 64), 0, 3, 3, 0))
            cs = s // 2
            qpc = _CHROMA_QP[qp]
            if tb_split and cs > 4:
                for pl in (1, 2):
                    tbs += [(x // 2 + dx, y // 2 + dy, cs // 2, pl, qpc, 1, 0, 0, 3, 3, 0) for dy in (0, cs // 2) for dx in (0, cs // 2)]
            else:
                tbs += [(x // 2, y // 2, cs, pl, qpc, 1, 0, 0, 3, 3, 0) for pl in (1, 2)]
    for by in range(height // 16):
        for bx in range(width // 16):
            x, y = bx * 16, by * 16
            ix = int(np.clip(4 + rng.integers(-3, 4), -(x + 80), width + 80 - x - 16))
            iy = int(np.clip(2 + rng.integers(-3, 4), -(y + 80), height + 80 - y - 16))
            cost.append((x, y, 16, 16, 0, 1 + int(rng.integers(0, 2)), int(rng.choice([-2, 0, 2])), int(rng.choice([-2, 0, 2])), ix, iy, 0))
    pred = np.array(pred, PRED_TASK)
    tbs = np.array(tbs, TB_TASK)
    # stable order by row band so that a CPU thread can take a contiguous slice (bench.py cpu arm)
    pred = pred[np.argsort(pred["y"] // 64, kind="stable")]
    tbs = tbs[np.argsort((tbs["y"] * np.where(tbs["plane"] > 0, 2, 1)) // 64, kind="stable")]
    off = 0
    for t in tbs:
        q = min(int(t["size"]), 16)
        t["coeff_offset"] = off
        off += (q * q + 7) & ~7
    return dict(width=width, height=height, qp=qp, range=search_range, org=horg, refs=refs, pred=pred, tb=tbs,
                cost=np.array(cost, COST_TASK), blockinfo=bi.reshape(-1), pool_entries=off, bipred_filter=1)


# reference common/common_block.c:78-83
_CHROMA_QP = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25,
              26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37, 38, 39, 40, 41, 42, 43, 44, 45]


def algorithmic_bytes(job):
    """Per-frame algorithmic HBM traffic of each stage."""
    w, h, R = job["width"], job["height"], job["range"]
    F = w * h * 3 // 2
    n2 = (2 * R) ** 2
    padded = lambda hp: hp.rows_y * hp.stride_y + 2 * hp.rows_c * hp.stride_c   # noqa: E731
    table = sum((w // bs) * (h // bs) * n2 * (2 if bs <= 16 else 4) for bs in (8, 16, 32, 64))
    ref_luma = (w + 2 * R) * (h + 2 * R)
    coded = int(sum(min(int(s), 16) ** 2 for s in job["tb"]["size"])) * 2
    tb_px = int((job["tb"]["size"].astype(np.int64) ** 2).sum())
    bi_frac = float((job["pred"]["ref1"] >= 0).mean())
    return {
        "sad_table": 2 * (w * h + ref_luma + table),
        "subpel": 2 * job["cost"].size * (16 * 16 + 22 * 22) + 2 * 8 * job["cost"].size,
        "predict": int(F * (2 + bi_frac)),
        "transform": 3 * tb_px + coded,
        "loopfilter": 3 * F + w * h // 8 + padded(job["refs"][0]),
    }


def shard_streams(n_streams, rank, world):
    """Streams are independent: rank r takes streams r, r+world, ... -- no collective."""
    return list(range(rank, n_streams, world))
