"""Host-side logic of the product library (no GPU needed) against the oracle: AccelerationPlan,
peak clustering, peak ranges and the three distillers, incl. empty / tie / edge inputs."""
import numpy as np
import pytest

import peasoup_b200 as pb
from peasoup_b200.host import CAND_DTYPE


@pytest.mark.parametrize("args", [
    (-5, 5, 1.1, 64.0, 1 << 17, 0.00032, 1475.12, -1.09, 0.0),
    (-50, 50, 1.1, 64.0, 1 << 23, 64e-6, 1382.0, -0.390625, 0.0),
    (-50, 50, 1.1, 64.0, 1 << 23, 64e-6, 1382.0, -0.390625, 500.0),
    (0, 50, 1.25, 40.0, 1 << 22, 64e-6, 1382.0, 0.390625, 12.5),
    (-500, 500, 1.1, 64.0, 1 << 24, 64e-6, 1382.0, -0.390625, 100.0),
    (0, 0, 1.1, 64.0, 1 << 17, 0.00032, 1475.12, -1.09, 3.0),
    (7, 7, 1.1, 64.0, 1 << 17, 0.00032, 1475.12, -1.09, 3.0),
])
def test_accel_plan(lib, oracle, args):
    a = pb.accel_plan(*args)
    b = oracle.accel_plan(*args)
    assert np.array_equal(a, b)
    if args[0] == args[1]:
        assert list(a) == [0.0]
    elif args[0] != 0 and args[1] != 0:
        assert a[0] == 0.0 and a[1] == np.float32(args[0]) and a[-1] == np.float32(args[1])


def test_accel_plan_config2_count(lib):
    a = pb.accel_plan(-50, 50, 1.1, 64.0, 1 << 23, 64e-6, 1382.0, -0.390625, 0.0)
    assert 60 < len(a) < 80  # ~71 trials at +/-50 m/s^2, default tolerance


@pytest.mark.parametrize("seed", range(6))
def test_cluster_peaks(lib, oracle, seed):
    rng = np.random.default_rng(seed)
    n = [0, 1, 2, 50, 500, 5000][seed]
    idx = np.sort(rng.choice(200000, n, replace=False)).astype(np.int32) if n else np.zeros(0, np.int32)
    if n > 10:
        idx[n // 2:n // 2 + 8] = idx[n // 2] + np.arange(8) * 29  # runs just inside the 30-bin gap
        idx = np.unique(idx).astype(np.int32)
    snr = (9 + rng.random(len(idx)) * 20).astype(np.float32)
    if len(idx) > 4:
        snr[3] = snr[2]  # exact tie: the later one must not replace the earlier
    a = pb.cluster_peaks(idx, snr)
    b = oracle.cluster_peaks(idx, snr)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("size,tsamp", [(1 << 17, 0.00032), (1 << 23, 64e-6), (1 << 24, 64e-6), (1 << 20, 1e-4)])
def test_peak_range(lib, oracle, size, tsamp):
    cfg = oracle.SearchConfig(size, tsamp)
    for nh in range(6):
        assert pb.peak_range(cfg.nbins, cfg.bin_width, nh, 0.1, 1100.0) == oracle.peak_range(cfg.nbins, cfg.bin_width, nh, 0.1, 1100.0)
    s, l, f = pb.peak_range(cfg.nbins, cfg.bin_width, 0, 0.1, 1100.0)
    # factor-2 quirk of peakfinder.hpp:83: start_idx ~ 2*min_freq/bin_width
    assert abs(s - 2 * 0.1 / cfg.bin_width) <= 2


def random_cands(rng, n, ties=False):
    c = np.zeros(n, CAND_DTYPE)
    base = rng.choice([1.3333, 4.0, 7.77, 50.0, 133.3], n)
    mult = rng.choice([0.25, 0.5, 1, 2, 3, 1.5, 5, 16], n)
    c["freq"] = (base * mult * (1 + rng.normal(0, 3e-5, n))).astype(np.float32)
    c["snr"] = (9 + rng.random(n) * 30).astype(np.float32)
    if ties and n > 6:
        c["snr"][1::3] = c["snr"][0]
    c["nh"] = rng.integers(0, 5, n)
    c["acc"] = rng.choice([-5.0, 0.0, 5.0, 31.0], n)
    c["dm"] = 10.0
    c["dm_idx"] = rng.integers(0, 4, n)
    return c


@pytest.mark.parametrize("kind,p2,keep,frac", [("harmonic", 16.0, False, True), ("harmonic", 16.0, True, False),
                                               ("harmonic", 4.0, True, True), ("acceleration", 536.87, True, True),
                                               ("acceleration", 41.9, False, True), ("dm", 0.0, True, True)])
@pytest.mark.parametrize("n", [0, 1, 2, 17, 200])
def test_distillers(lib, oracle, kind, p2, keep, frac, n):
    rng = np.random.default_rng(n + len(kind))
    for ties in (False, True):
        c = random_cands(rng, n, ties)
        a = pb.distill(kind, c, 1e-4, p2, keep, frac)
        b = oracle.distill(kind, c, 1e-4, p2, keep, frac)
        assert len(a) == len(b)
        for name in CAND_DTYPE.names:
            assert np.array_equal(a[name], b[name]), name
        if not keep:
            assert a["is_top"].all()
        assert len(a) >= (1 if n else 0)


@pytest.mark.parametrize("n", [0, 1, 3, 40, 300])
def test_distill_merged_vs_oracle_finalise(lib, oracle, n):
    """ps_distill_merged = the two distillations after the workers joined (pipeline_multi.cu:353-365) on
    a merged per-DM list whose entries carry acceleration associates, against the oracle's finalise()
    (itself compared with the golden tutorial run through test_tutorial_final_list_vs_oracle_postprocess).
    Distinct S/N values: the reference's std::sort is unstable, ties have no defined order."""
    from peasoup_b200.host import distill_merged
    rng = np.random.default_rng(100 + n)
    c = random_cands(rng, n)
    c["snr"] = (9 + np.arange(n)[rng.permutation(n)] * 0.37).astype(np.float32)
    c["is_top"] = 1
    if n > 3:                                    # every third record becomes an associate of its predecessor
        c["is_top"][2::3] = 0
        c["dm_idx"] = np.sort(c["dm_idx"])
    a = distill_merged(c, 1e-4, 16)
    b = oracle.finalise(c, 1e-4, 16, 10 ** 9)
    # nothing is dropped: non-unique candidates become associates -- more than once when several harmonic
    # ratios match, because the reference's HarmonicDistiller does not break after a match
    assert len(a) == len(b) >= n
    for name in CAND_DTYPE.names:
        assert np.array_equal(a[name], b[name]), name
    if n:
        tops = a[a["is_top"] == 1]
        assert np.all(np.diff(tops["snr"]) <= 0)  # survivors come out in descending S/N
