"""bench.py's closed-GOP segment leg (the N-GPU workload of BASELINE configs[2]) on the CPU: the same function, with the
segment encoder swapped for integration/_build/Thorenc_seg (stock pixel path) and three "ranks" run one after the other --
what it shards, what it joins and what it asks the stock decoder to confirm."""
import importlib.util
import os
import types

import pytest

from thorref import REF_DIR, ROOT

SEG = os.path.join(ROOT, "integration", "_build", "Thorenc_seg")


def test_segment_leg_shards_joins_and_checks(tmp_path):
    for need in (SEG, os.path.join(REF_DIR, "Thordec")):
        if not os.path.exists(need):
            pytest.skip(f"{need} not built (needs /root/reference at build time)")
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    bench.GPU_ENC = SEG                                   # TEST ONLY: the leg itself never runs a CPU encoder
    args = types.SimpleNamespace(segments_config="RA_low_complexity", width=128, height=72, segments_per_gpu=2, segments_ctas=15, qp=32)
    shared = str(tmp_path)
    bench.make_clip(os.path.join(shared, "seg_base_128x72.yuv"), 128, 72, 64, seed=44)
    res = None
    for rank in (2, 1, 0):                                # rank 0 last: it joins what the others left behind
        res = bench.segments_leg(args, rank, 3, {}, shared, lambda: None, lambda v: v)
    assert res["segments"] == 6 and res["frames"] == 6 * 64 + 1 and res["n_gpus"] == 3
    assert res["stock_decoder_reproduces_the_encoders_reconstruction"] is True
    assert "first 129 coded frames vs the 128 frames" in res["decode_check"]     # rank 0's shard + the next rank's intra frame
    assert res["redundant_intra_frames"] == 5
