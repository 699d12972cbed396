"""Closed-GOP segment sharding with REAL encoders, on the CPU.

One RA stream is encoded serially by the stock Thorenc and as several segments by separate processes of the encoder whose
main() takes -seg_start / -seg_end (integration/enc_segment_hook.c; Thorenc_seg = the stock pixel path, Thorenc_emu = the
device decision source compiled for the host).  thor_b200/segments.py joins the chunks; the joined .bit and the
concatenated reconstructions must equal the serial run's byte for byte.  A world_size-2 gloo run does the same with the
segments sharded over two ranks, as bench.py does over GPUs."""
import os
import socket
import subprocess
import sys

import pytest

from thorref import REF_DIR, ROOT, THOR_CONFIGS, write_hard_clip

sys.path.insert(0, ROOT)
from thor_b200 import segments as sg  # noqa: E402

SEG = os.path.join(ROOT, "integration", "_build", "Thorenc_seg")
EMU = os.path.join(ROOT, "integration", "_build", "Thorenc_emu")
W, H, FRAMES, PERIOD = 128, 72, 67, 16          # intra frames at 0, 16, 32, 48, 64; two frames of PPP tail after the last


@pytest.fixture(scope="module")
def clip(tmp_path_factory):
    for need in (os.path.join(REF_DIR, "Thorenc"), SEG, EMU):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")
    path = str(tmp_path_factory.mktemp("seg") / "clip.yuv")
    write_hard_clip(path, W, H, FRAMES, seed=5)
    return path


def stream_args(cfg, yuv, frames=FRAMES, period=PERIOD):
    return ["-if", yuv, "-width", str(W), "-height", str(H), "-n", str(frames), "-f", "30", "-qp", "32"] + THOR_CONFIGS[cfg] + ["-intra_period", str(period)]


def serial(cfg, yuv, out, frames=FRAMES, period=PERIOD):
    subprocess.run([os.path.join(REF_DIR, "Thorenc")] + stream_args(cfg, yuv, frames, period) + ["-of", out + ".bit", "-rf", out + ".yuv"],
                   check=True, capture_output=True, timeout=600)
    return open(out + ".bit", "rb").read(), open(out + ".yuv", "rb").read()


def joined(exe, cfg, yuv, parts, out, frames=FRAMES, period=PERIOD):
    ranges = sg.segment_ranges(frames, period, parts)
    files = sg.encode_segments(exe, stream_args(cfg, yuv, frames, period), ranges, out, timeout=600)
    bit = sg.join_segment_streams(ranges, [open(f[0], "rb").read() for f in files], frames, 7, period)
    return ranges, bit, b"".join(open(f[1], "rb").read() for f in files)


@pytest.mark.parametrize("cfg,parts", [("RA_medium_complexity", 3), ("RA_medium_complexity", 5), ("RA_low_complexity", 4)])
def test_segments_of_the_stock_encoder_join_into_the_serial_stream(tmp_path, clip, cfg, parts):
    bit0, rec0 = serial(cfg, clip, str(tmp_path / "serial"))
    ranges, bit1, rec1 = joined(SEG, cfg, clip, parts, str(tmp_path / "seg"))
    assert len(ranges) == parts
    assert bit1 == bit0, f"{cfg}: joined bitstream of {ranges} differs from the serial one"
    assert rec1 == rec0, f"{cfg}: joined reconstruction differs"


def test_segments_of_the_device_decision_logic_join_into_the_serial_stream(tmp_path, clip):
    """the same with every block decision taken by thor_b200/csrc/enc (host emulation): what Thorenc_gpu runs per segment"""
    frames = 35
    bit0, rec0 = serial("RA_medium_complexity", clip, str(tmp_path / "serial"), frames)
    ranges, bit1, rec1 = joined(EMU, "RA_medium_complexity", clip, 3, str(tmp_path / "emu"), frames)
    assert ranges == [(0, 16), (16, 32), (32, None)]
    assert bit1 == bit0 and rec1 == rec0


def test_without_segment_arguments_the_hooked_main_is_the_stock_encoder(tmp_path, clip):
    bit0, rec0 = serial("RA_low_complexity", clip, str(tmp_path / "serial"), 20)
    out = str(tmp_path / "plain")
    subprocess.run([SEG] + stream_args("RA_low_complexity", clip, 20) + ["-of", out + ".bit", "-rf", out + ".yuv"], check=True, capture_output=True)
    assert open(out + ".bit", "rb").read() == bit0 and open(out + ".yuv", "rb").read() == rec0


@pytest.mark.parametrize("extra,why", [
    (["-seg_start", "8"], "not an intra frame"),
    (["-seg_start", "16", "-seg_end", "16"], "not a later intra frame"),
    (["-seg_start", "16", "-skip", "1"], "-skip rebases"),
    (["-seg_start", "x"], "not a frame number"),
])
def test_bad_segment_arguments_are_refused(tmp_path, clip, extra, why):
    r = subprocess.run([SEG] + stream_args("RA_low_complexity", clip) + ["-of", str(tmp_path / "o.bit")] + extra, capture_output=True, text=True)
    assert r.returncode == 2 and why in r.stderr


def test_segments_need_an_intra_period(tmp_path, clip):
    args = ["-if", clip, "-width", str(W), "-height", str(H), "-n", "9", "-qp", "32"] + THOR_CONFIGS["LDB_low_complexity"]
    r = subprocess.run([SEG] + args + ["-of", str(tmp_path / "o.bit"), "-seg_start", "0", "-seg_end", "16"], capture_output=True, text=True)
    assert r.returncode == 2 and "closed-GOP" in r.stderr
    with pytest.raises(ValueError):
        sg.segment_ranges(60, 0, 4)


def test_segment_ranges():
    r = sg.segment_ranges(513, 64, 8)                     # nine intra frames (0 .. 512) -> eight ranges, one holds two periods
    assert r[0][0] == 0 and r[-1][1] is None and len(r) == 8
    assert all(a % 64 == 0 and (b is None or b % 64 == 0) for a, b in r)
    assert all(r[k][1] == r[k + 1][0] for k in range(7))
    assert sg.segment_ranges(20, 64, 4) == [(0, None)]    # a stream shorter than one period is one segment
    assert sg.segment_ranges(129, 64, 2) == [(0, 64), (64, None)]


def test_join_refuses_a_missing_or_doubled_frame():
    chunk = lambda n: (n).to_bytes(4, "big") + bytes(n)   # noqa: E731
    order = sg.coding_order(17, 7, 16)
    ranges = [(0, 16), (16, None)]
    first = b"".join(chunk(3) for f in order if f.frame_num < 16)
    second = chunk(5)
    assert len(sg.split_chunks(sg.join_segment_streams(ranges, [first, second], 17, 7, 16))) == 17
    with pytest.raises(ValueError):
        sg.join_segment_streams(ranges, [first, second + chunk(5)], 17, 7, 16)
    with pytest.raises(ValueError):
        sg.join_segment_streams(ranges, [first[:-7], second], 17, 7, 16)
    with pytest.raises(ValueError):
        sg.split_chunks(chunk(9)[:-1])


WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import torch.distributed as dist
from thor_b200 import segments as sg
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
ranges = sg.segment_ranges({frames}, {period}, 4)
mine = sg.shard(list(enumerate(ranges)), rank, world)                  # contiguous: neighbours share an intra frame
files = sg.encode_segments({exe!r}, {args!r}, [r for _, r in mine], os.path.join({out!r}, f"rank{{rank}}"))
local = [(k, open(f[0], "rb").read(), open(f[1], "rb").read()) for (k, _), f in zip(mine, files)]
gathered = [None] * world
dist.all_gather_object(gathered, local)                                # the only exchange: finished chunks, for the join
if rank == 0:
    flat = sorted(x for part in gathered for x in part)
    assert [k for k, _, _ in flat] == list(range(len(ranges)))
    bit = sg.join_segment_streams(ranges, [b for _, b, _ in flat], {frames}, 7, {period})
    open(os.path.join({out!r}, "joined.bit"), "wb").write(bit)
    open(os.path.join({out!r}, "joined.yuv"), "wb").write(b"".join(r for _, _, r in flat))
dist.barrier()
dist.destroy_process_group()
"""


def test_two_ranks_shard_the_segments_of_one_stream(tmp_path, clip):
    bit0, rec0 = serial("RA_low_complexity", clip, str(tmp_path / "serial"))
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, frames=FRAMES, period=PERIOD, exe=SEG, args=stream_args("RA_low_complexity", clip), out=str(tmp_path)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out[-3000:]
    assert (tmp_path / "joined.bit").read_bytes() == bit0
    assert (tmp_path / "joined.yuv").read_bytes() == rec0
