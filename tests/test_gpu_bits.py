"""GPU parity (bit-exact, through the C ABI) of the coefficient bit counts (first slice:
thorcuda_coeff_bits_list = write_coeff of enc/write_bits.c:110-253 as a length) against the CPU oracle (oracle/orc_bits.c,
itself pinned to the reference's write_coeff writing into a real stream): synthetic level blocks of every size, type
and density, and the levels the device's own encoder leg leaves in the pool."""
import numpy as np
import pytest

from test_gpu_blocks import make_tb_tasks, tb_pictures
from thorref import COEFF_KINDS, coeff_block, load_oracle, ptr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import thor_b200
    thor_b200.init(0)
    yield thor_b200
    thor_b200.shutdown()


@pytest.fixture(scope="module")
def orc():
    return load_oracle()


def oracle_bits(orc, tasks, pool):
    out = np.zeros(tasks.size, np.uint32)
    for i, t in enumerate(tasks):
        q = min(int(t["size"]), 16)
        c = np.ascontiguousarray(pool[int(t["coeff_offset"]):int(t["coeff_offset"]) + q * q])
        out[i] = orc.orc_coeff_bits(ptr(c), q, int(t["size"]), int(t["coeff_type"]))
    return out


@pytest.mark.parametrize("seed", [1, 2])
def test_coeff_bits_synthetic(tb, orc, seed):
    """every size x block type x level pattern, incl. all-zero blocks (0 bits), a single +-1 (the 2-bit chroma short cut),
    a level at the last scan position (no end-of-block code) and levels up to +-3000 (escape codes)"""
    rng = np.random.default_rng(seed)
    blocks = []
    for size in (4, 8, 16, 32, 64):
        for kind in COEFF_KINDS + ["zero"]:
            for rep in range(40 if size <= 16 else 12):
                c = np.zeros((size, size), np.int16) if kind == "zero" else coeff_block(rng, size, kind)
                blocks.append((size, int(rng.integers(0, 4)), c))
    order = rng.permutation(len(blocks))
    tasks = np.zeros(len(blocks), tb.TB_TASK)
    for i, k in enumerate(order):
        tasks[i]["size"], tasks[i]["coeff_type"] = blocks[k][0], blocks[k][1]
    entries = tb.ops.pool_layout(tasks)
    pool = np.zeros(entries, np.int16)
    for i, k in enumerate(order):
        q = min(blocks[k][0], 16)
        off = int(tasks[i]["coeff_offset"])
        pool[off:off + q * q] = blocks[k][2][:q, :q].reshape(-1)
    pic = tb.Picture(64, 64)
    got = tb.ops.coeff_bits([pic], tasks, pool)
    want = oracle_bits(orc, tasks, pool)
    bad = np.flatnonzero(got != want)
    assert bad.size == 0, f"{bad.size} of {want.size} differ; first: task {bad[0]} {tasks[bad[0]]} got {got[bad[0]]} want {want[bad[0]]}"
    assert (want == 0).sum() >= 5 * 12 and want.max() > 2000
    pic.close()


@pytest.mark.parametrize("wh,qps,intra", [((256, 192), (22, 32, 37), False), ((416, 240), (8, 30), True), ((1920, 1080), (32,), False)])
def test_coeff_bits_after_encode(tb, orc, wh, qps, intra):
    """the encoder leg and the bit count on ONE resident list: levels never leave the device between the two launches"""
    w, h = wh
    rng = np.random.default_rng(w + qps[0])
    opl, ppl, org, pred = tb_pictures(tb, w, h, 5 + qps[0], noise=14)
    rec = tb.Picture(w, h)
    tasks = make_tb_tasks(tb, w, h, rng, qps, sizes_luma=(16, 64) if w >= 1920 else (8, 64), intra_frame=intra)
    tb.ops.pool_layout(tasks)
    lst = tb.ops.TaskList(tb.LIST_TB, [org, pred, rec], tasks)
    coeffq = np.zeros(lst.pool_entries, np.int16)
    cbp = np.zeros(lst.ntasks, np.uint8)
    bits = np.zeros(lst.ntasks, np.uint32)
    lst.encode(coeffq, cbp, None)
    lst.coeff_bits(bits)
    tb.sync()
    lst.close()
    want = oracle_bits(orc, tasks, coeffq)
    assert np.array_equal(bits, want)
    assert np.array_equal(bits != 0, cbp != 0), "a block has bits exactly when it has coefficients"
    assert 0 < (cbp != 0).sum() < cbp.size
    for p in (org, pred, rec):
        p.close()


@pytest.mark.parametrize("seed,n", [(21, 3000), (22, 20000)])
def test_block_bits(tb, orc, seed, n):
    """thorcuda_block_bits = write_block (enc/write_bits.c:364-650) as a length, against the oracle (pinned to the
    reference's write_block): random blocks of every frame type, mode, partition, context, transform split and cbp
    pattern; the coefficient bits come from the resident transform-block list of the same blocks."""
    from thorref import BS, pack_blocks, random_block_syntax
    rng = np.random.default_rng(seed)
    blocks = [random_block_syntax(rng) for _ in range(n)]
    recs, tasks, pool = pack_blocks(blocks, tb.BLOCK_SYNTAX, tb.TB_TASK)
    want = np.array([orc.orc_block_bits(ptr(d), ptr(cy), ptr(cu), ptr(cv)) for d, cy, cu, cv in blocks], np.uint32)
    want_syntax = np.array([orc.orc_block_syntax_bits(ptr(d)) for d, _, _, _ in blocks], np.uint32)
    pic = tb.Picture(64, 64)
    lst = tb.ops.TaskList(tb.LIST_TB, [pic], tasks)
    cq = np.zeros(lst.pool_entries, np.int16)
    cq[:min(cq.size, pool.size)] = pool[:cq.size]
    lst.coeff_bits(None, cq)
    got = tb.ops.block_bits(recs, lst)
    bad = np.flatnonzero(got != want)
    assert bad.size == 0, f"{bad.size} of {n} differ; first: {recs[bad[0]]} got {got[bad[0]]} want {want[bad[0]]}"
    got_syntax = tb.ops.block_bits(recs, None)             # no list: the syntax alone
    assert np.array_equal(got_syntax, want_syntax)
    nolist = recs.copy()
    nolist["tb_first"] = tb.NO_TB                          # per block opt-out gives the same
    assert np.array_equal(tb.ops.block_bits(nolist, lst), want_syntax)
    modes = {int(d[BS["mode"]]) for d, _, _, _ in blocks}
    assert modes == {0, 1, 2, 3, 4} and (want > want_syntax).any()
    lst.close(); pic.close()


def test_block_bits_rejects(tb):
    """malformed records are refused on the host: no launch, an error"""
    pic = tb.Picture(64, 64)
    tasks = np.zeros(3, tb.TB_TASK)
    tasks["size"] = (16, 8, 8)
    tb.ops.pool_layout(tasks)
    lst = tb.ops.TaskList(tb.LIST_TB, [pic], tasks)
    good = np.zeros(1, tb.BLOCK_SYNTAX)
    good["frame_type"], good["num_ref"], good["size"], good["mode"], good["num_intra_modes"] = 1, 1, 16, 1, 10
    lst.coeff_bits(None, np.zeros(lst.pool_entries, np.int16))
    assert tb.ops.block_bits(good, lst)[0] > 0
    for field, value in (("size", 24), ("mode", 5), ("frame_type", 3), ("pb_part", 4), ("intra_mode", 10), ("cbp_y", 5), ("tb_first", 1)):
        bad = good.copy()
        bad[field] = value
        with pytest.raises(tb.ThorCudaError):
            tb.ops.block_bits(bad, lst)
    lst.close(); pic.close()
