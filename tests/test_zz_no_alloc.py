"""The entry point for callers that may not allocate device memory (pte_partition_u64_static).

Such a caller -- a server whose memory pool owns the whole GPU, say -- brings its own input and
output buffers; the library's scratch is a fixed array in its device image.  This file checks that
the entry point gives exactly what the ordinary sort entry point (pte_debug_sort_u64, which uses
engine-allocated scratch) gives on the same bit range, on two calls in a row, so that whatever the
first call leaves behind in the shared array cannot leak into the second.  It is named to be
collected last: a run that stops at the first failure has then already run everything else.
"""
import numpy as np
import pytest

from k8s_watcher_b200 import PhaseTransitionEngine, PteError
from k8s_watcher_b200.engine import STATIC_MAX_RECORDS, PinnedBuffer, partition_u64_static

pytestmark = pytest.mark.gpu


def _case(eng, n, bit_lo, n_bits, seed):
    rng = np.random.default_rng(seed)
    src, dst = PinnedBuffer(n, np.uint64), PinnedBuffer(n, np.uint64)
    try:
        src.array[:] = rng.integers(0, 1 << 63, n, dtype=np.uint64)
        dst.array[:] = 0
        partition_u64_static(src.array, dst.array, bit_lo, n_bits)
        want = eng.debug_sort(src.array, bit_lo, bit_lo + n_bits)      # the existing entry point
        assert np.array_equal(dst.array, want)                         # integer work: bit-exact
        # and against numpy directly: a stable sort of the records on the digit
        digit = (src.array >> np.uint64(bit_lo)) & np.uint64((1 << n_bits) - 1)
        assert np.array_equal(dst.array, src.array[np.argsort(digit, kind="stable")])
    finally:
        src.free()
        dst.free()


def test_static_partition_matches_engine_sort_twice_in_a_row():
    with PhaseTransitionEngine(1 << 20, 1000) as eng:
        # second call: other size (fewer tiles), other digit, right after the first
        _case(eng, 1_000_003, 24, 8, seed=1)
        _case(eng, 300_001, 3, 5, seed=2)
        # and once more at the first shape: the scratch now holds the second call's leftovers
        _case(eng, 1_000_003, 24, 8, seed=3)


def test_static_partition_edge_sizes():
    with PhaseTransitionEngine(1 << 16, 100) as eng:
        for n in (1, 31, 4095, 4096, 4097):
            _case(eng, n, 0, 8, seed=n)
    empty = np.zeros(0, np.uint64)
    partition_u64_static(empty, empty.copy(), 0, 8)                    # nothing to do, no error


def test_static_partition_refuses_oversize_and_bad_digits():
    a = np.zeros(8, np.uint64)
    big = np.zeros(STATIC_MAX_RECORDS + 1, np.uint64)                  # never touched: refused up front
    with pytest.raises(PteError) as ei:
        partition_u64_static(big, big.copy(), 0, 8)
    assert ei.value.code == -4 and "PTE_STATIC_MAX_RECORDS" in str(ei.value)
    for lo, bits in ((0, 9), (0, 0), (60, 8), (-1, 4)):
        with pytest.raises(PteError) as ei:
            partition_u64_static(a, a.copy(), lo, bits)
        assert ei.value.code == -1


def test_static_partition_on_a_stream_of_the_callers():
    """The work is issued on the stream the caller names: here one it created itself (not the default stream)."""
    import ctypes as C
    rt = None
    for name in ("libcudart.so.12", "libcudart.so", "/usr/local/cuda/lib64/libcudart.so"):
        try:
            rt = C.CDLL(name)
            break
        except OSError:
            continue
    if rt is None:
        pytest.skip("no shared CUDA runtime to create a stream with")
    stream = C.c_void_p()
    assert rt.cudaStreamCreate(C.byref(stream)) == 0
    try:
        rng = np.random.default_rng(11)
        n = 123_457
        src, dst = PinnedBuffer(n, np.uint64), PinnedBuffer(n, np.uint64)
        try:
            for lo, bits in ((7, 8), (40, 3)):                         # two calls in a row on that stream
                src.array[:] = rng.integers(0, 1 << 63, n, dtype=np.uint64)
                partition_u64_static(src.array, dst.array, lo, bits, stream=stream.value)
                digit = (src.array >> np.uint64(lo)) & np.uint64((1 << bits) - 1)
                assert np.array_equal(dst.array, src.array[np.argsort(digit, kind="stable")])
        finally:
            src.free()
            dst.free()
    finally:
        rt.cudaStreamDestroy(stream)
