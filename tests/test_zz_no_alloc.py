"""ps_running_median_noalloc: the allocation-free variant of the running-median stage.

Some users of this library cannot allocate device memory at the point where they whiten a
spectrum (a cudaMalloc stalls every stream of the device, or the host application owns the
whole memory budget), and have no scratch buffer to hand over.  The variant keeps its scratch
in an array built into the library.  This file checks, through the C ABI, that it returns
exactly what ps_running_median returns -- on two calls in a row, because the built-in array is
shared and keeps the previous call's contents -- and that it refuses sizes beyond its limit.
The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test."""
import numpy as np
import pytest

from peasoup_b200 import check, ptr

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

MAX_BINS = 4194305  # PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS


def spectrum(nbins, seed):
    rng = np.random.default_rng(seed)
    red = 1 + 80.0 / (1 + np.arange(nbins) / 40.0)
    return (np.abs(rng.standard_normal(nbins)) * red).astype(np.float32)


def both_ways(lib, p, bin_width):
    nb = len(p)
    d = torch.from_numpy(p).cuda()
    want = torch.zeros(nb, dtype=torch.float32, device="cuda")
    work = torch.zeros(nb // 5 + nb // 25 + nb // 125 + 8, dtype=torch.float32, device="cuda")
    check(lib.ps_running_median(ptr(d), nb, bin_width, 0.05, 0.5, ptr(want), ptr(work), None))
    got = torch.full((nb,), -1.0, dtype=torch.float32, device="cuda")
    check(lib.ps_running_median_noalloc(ptr(d), nb, bin_width, 0.05, 0.5, ptr(got), None))
    torch.cuda.synchronize()
    return got, want


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(lib):
    # first call: the largest supported size; second call: a smaller spectrum, so the shared
    # scratch still holds stale values from the first call beyond what the second one rewrites
    for nbins, tsamp, seed in ((MAX_BINS, 64e-6, 1), (65537, 0.00032, 2)):
        size = 2 * (nbins - 1)
        bin_width = float(np.float32(1.0 / float(np.float32(np.float32(size) * np.float32(tsamp)))))
        got, want = both_ways(lib, spectrum(nbins, seed), bin_width)
        assert torch.equal(got, want)  # medians are selections + IEEE lerp: bit-exact, as in test_gpu_stages
        assert float(want.min()) > 0


def test_noalloc_rejects_sizes_beyond_its_scratch(lib):
    d = torch.zeros(16, dtype=torch.float32, device="cuda")
    assert lib.ps_running_median_noalloc(ptr(d), MAX_BINS + 1, 1e-3, 0.05, 0.5, ptr(d), None) == -2
    msg = lib.ps_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(MAX_BINS) in msg
