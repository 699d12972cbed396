"""thorcuda_block_bits_static: the allocation-free variant of thorcuda_block_bits (scratch in a static device array of the
library; include/thorcuda.h).  Called through the C ABI and compared bit for bit (integer results: tolerance 0, as in
tests/test_gpu_bits.py) with the existing entry point, on two calls in a row that share the scratch.  Collected last on
purpose: a failure here hides no other test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import thor_b200
    thor_b200.init(0)
    yield thor_b200
    thor_b200.shutdown()


def _records(tb, seed, n):
    from thorref import pack_blocks, random_block_syntax
    rng = np.random.default_rng(seed)
    recs, _, _ = pack_blocks([random_block_syntax(rng) for _ in range(n)], tb.BLOCK_SYNTAX, tb.TB_TASK)
    recs["tb_first"] = tb.NO_TB
    return recs


def test_static_variant_matches_on_two_calls_in_a_row(tb):
    first, second = _records(tb, 11, 3000), _records(tb, 12, 700)     # the second call is shorter: stale entries stay behind it
    want1, want2 = tb.ops.block_bits(first, None), tb.ops.block_bits(second, None)
    before = tb.launch_count()
    got1 = tb.ops.block_bits_static(first)
    got2 = tb.ops.block_bits_static(second)
    assert tb.launch_count() == before + 2, "the static variant must launch its kernel (no host path)"
    assert np.array_equal(got1, want1)
    assert np.array_equal(got2, want2)
    assert (want1 > 0).all() and len(set(want1.tolist())) > 10
    # and once more with the first input: the shared scratch holds nothing that a call depends on
    assert np.array_equal(tb.ops.block_bits_static(first), want1)


def test_static_variant_limits(tb):
    lib = tb.cdll()
    import ctypes as C
    n = 4096
    recs = _records(tb, 13, 64)
    big = np.resize(recs, n + 1)
    out = np.zeros(n + 1, np.uint32)
    assert np.array_equal(tb.ops.block_bits_static(big[:n]), tb.ops.block_bits(big[:n], None))      # exactly the maximum
    rc = lib.thorcuda_block_bits_static(big.ctypes.data_as(C.c_void_p), n + 1, out.ctypes.data_as(C.c_void_p))
    assert rc != 0
    lib.thorcuda_last_error.restype = C.c_char_p
    assert b"exceed the static scratch" in lib.thorcuda_last_error()
    bad = recs.copy()
    bad["size"][0] = 24
    with pytest.raises(tb.ThorCudaError):
        tb.ops.block_bits_static(bad)
    assert tb.ops.block_bits_static(recs[:0]).size == 0
