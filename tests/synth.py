"""Seeded synthetic DM trials for parity tests: Gaussian baseline (mean 128,
sigma 16), a slow red-noise walk, two RFI birdies (50/100 Hz) and one pulsar (P0 = 7.5 ms,
5 % duty) with a constant line-of-sight acceleration, brightest in the middle DM trial."""
import numpy as np

SEED = 20140213
C_LIGHT = 299792458.0


def make_trials(ndm, nsamps, size, tsamp=64e-6, a_inj=31.4, p0=7.5e-3, amp=6.0, seed=SEED, birdies=True):
    rng = np.random.default_rng(seed)
    j = np.arange(nsamps, dtype=np.float64)
    af = np.float32(a_inj) * np.float32(tsamp) / (2 * C_LIGHT)
    # resample_kernelII reads in[i + i*af*(i-size)]; a pulse train periodic in i therefore sits at
    # phase(j) ~ (j - af*j*(j-size)) * tsamp / p0 in the input series
    phase = ((j - af * j * (j - size)) * tsamp / p0) % 1.0
    pulse = (phase < 0.05).astype(np.float64)
    t = j * tsamp
    out = np.zeros((ndm, nsamps), np.uint8)
    for d in range(ndm):
        x = 128.0 + 16.0 * rng.standard_normal(nsamps)
        walk = np.cumsum(rng.standard_normal(nsamps // 1024 + 2))
        x += 8.0 * np.interp(j, np.arange(len(walk)) * 1024.0, walk) / max(1.0, np.abs(walk).max())
        if birdies:
            x += 3.0 * np.sin(2 * np.pi * 50.0 * t) + 2.0 * np.sin(2 * np.pi * 100.0 * t + 1.0)
        w = max(0.0, 1.0 - abs(d - ndm // 2) / 8.0)
        x += amp * w * pulse
        out[d] = np.clip(np.rint(x), 0, 255).astype(np.uint8)
    return out


def cand_keys(recs, nbins, bin_width_f, with_top=False):
    """(dm_idx, acc, nh, bin) -> snr for flattened candidate records."""
    out = {}
    for r in recs:
        nh = int(r["nh"])
        factor = np.float32(1.0 / nbins * np.float32(bin_width_f * nbins) / 2.0 ** nh)
        b = int(round(float(r["freq"]) / float(factor)))
        k = (int(r["dm_idx"]), float(r["acc"]), nh, b)
        if with_top:
            k = k + (int(r["is_top"]),)
        out[k] = float(r["snr"])
    return out


def write_sigproc(path, data, tsamp, fch1, foff, nbits=8, source="synthetic"):
    """Minimal SIGPROC filterbank writer (header.hpp:171-336 keys the reader needs).  data:
    [nsamps][nchans] uint8 values already packed for nbits = 8."""
    import struct

    def s(x):
        b = x.encode()
        return struct.pack("<i", len(b)) + b

    nsamps, nchans = data.shape
    hdr = s("HEADER_START") + s("source_name") + s(source)
    hdr += s("machine_id") + struct.pack("<i", 0) + s("telescope_id") + struct.pack("<i", 0)
    hdr += s("data_type") + struct.pack("<i", 1)
    hdr += s("fch1") + struct.pack("<d", fch1) + s("foff") + struct.pack("<d", foff)
    hdr += s("nchans") + struct.pack("<i", nchans) + s("nbits") + struct.pack("<i", nbits)
    hdr += s("tstart") + struct.pack("<d", 56000.0) + s("tsamp") + struct.pack("<d", tsamp)
    hdr += s("nifs") + struct.pack("<i", 1) + s("HEADER_END")
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(np.ascontiguousarray(data, np.uint8).tobytes())
