"""oracle.py -- CPU restatement of peasoup's per-DM acceleration search (numpy + oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs as the checker, never by the product (peasoup_b200/).

Per-element arithmetic lives in oracle.cpp (exact float semantics); this module adds
  * the FFTs (numpy pocketfft in double precision, rounded to float32 -- the reference uses
    cuFFT, so FFT-derived values are compared by tolerance),
  * the per-DM loop of Worker::start (src/pipeline_multi.cu:145-244),
  * a restatement of the un-vendored `dedisp` library's dedispersion (the step before the hot
    path) so that the reference's golden example output can pin the whole chain,
  * parsers for SIGPROC filterbanks and the golden candidates.peasoup / overview.xml.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import struct
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")
_lib = None

CAND_DTYPE = np.dtype([("dm", "f4"), ("dm_idx", "i4"), ("acc", "f4"), ("nh", "i4"), ("snr", "f4"),
                       ("freq", "f4"), ("is_top", "i4")])
POD_DTYPE = np.dtype([("dm", "f4"), ("dm_idx", "i4"), ("acc", "f4"), ("nh", "i4"), ("snr", "f4"), ("freq", "f4")])


def build():
    src = os.path.join(_HERE, "oracle.cpp")
    if os.path.exists(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(src):
        return
    subprocess.run(["g++", "-O2", "-std=c++11", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
                    "-o", _LIB, src, "-lm"], check=True)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_distill.restype = C.c_long
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


f32 = np.float32


# ------------------------------------------------------------------ per-stage wrappers
def resample_index(n, a, tsamp, which=2):
    out = np.zeros(n, np.uint64)
    lib().orc_resample_index(which, C.c_size_t(n), C.c_float(a), C.c_float(tsamp), _p(out))
    return out


def resample(x, a, tsamp, which=2):
    x = np.ascontiguousarray(x, f32)
    out = np.empty_like(x)
    lib().orc_resample(which, _p(x), _p(out), C.c_size_t(len(x)), C.c_float(a), C.c_float(tsamp))
    return out


def form_power(X, way):
    X = np.ascontiguousarray(X, np.complex64)
    out = np.empty(len(X), f32)
    lib().orc_form_power(_p(X), _p(out), C.c_size_t(len(X)), way)
    return out


def normalise(p, mean, sigma):
    p = np.array(p, f32)
    lib().orc_normalise(_p(p), C.c_float(mean), C.c_float(sigma), C.c_size_t(len(p)))
    return p


def harmonic_sum(p, nfolds):
    p = np.ascontiguousarray(p, f32)
    out = np.zeros((nfolds, len(p)), f32)
    lib().orc_harmonic_sum(_p(p), _p(out), C.c_size_t(len(p)), nfolds)
    return out


def find_peaks(d, n, start, thresh):
    d = np.ascontiguousarray(d, f32)
    cap = max(1, n - max(start, 0))
    idx = np.zeros(cap, np.int32)
    snr = np.zeros(cap, f32)
    c = lib().orc_find_peaks(n, start, _p(d), C.c_float(thresh), _p(idx), _p(snr), cap)
    return idx[:c], snr[:c]


def cluster_peaks(idxs, snrs, min_gap=30):
    i = np.ascontiguousarray(idxs, np.int32)
    s = np.ascontiguousarray(snrs, f32)
    oi, os_ = np.zeros_like(i), np.zeros_like(s)
    n = lib().orc_cluster_peaks(_p(i), _p(s), len(i), min_gap, _p(oi), _p(os_))
    return oi[:n], os_[:n]


def peak_range(nbins, bin_width, nh, min_freq, max_freq):
    a, b, f = C.c_int(), C.c_int(), C.c_float()
    lib().orc_peak_range(nbins, C.c_double(bin_width), nh, C.c_float(min_freq), C.c_float(max_freq), C.byref(a),
                         C.byref(b), C.byref(f))
    return a.value, b.value, f.value


def median_scrunch5(x):
    x = np.ascontiguousarray(x, f32)
    out = np.zeros(max(1, len(x) // 5 if len(x) >= 5 else 1), f32)
    lib().orc_median_scrunch5(_p(x), C.c_size_t(len(x)), _p(out))
    return out


def linear_stretch(x, out_count):
    x = np.ascontiguousarray(x, f32)
    out = np.zeros(out_count, f32)
    lib().orc_linear_stretch(_p(x), C.c_size_t(len(x)), _p(out), C.c_size_t(out_count))
    return out


def running_median(pspec, bin_width, b5=0.05, b25=0.5):
    pspec = np.ascontiguousarray(pspec, f32)
    out = np.zeros(len(pspec), f32)
    lib().orc_running_median(_p(pspec), len(pspec), C.c_double(bin_width), C.c_float(b5), C.c_float(b25), _p(out))
    return out


def divide_c_by_f(X, f):
    X = np.array(X, np.complex64)
    f = np.ascontiguousarray(f, f32)
    lib().orc_divide_c_by_f(_p(X), _p(f), len(X))
    return X


def zap(X, birdies, widths, bin_width):
    X = np.array(X, np.complex64)
    b = np.ascontiguousarray(birdies, f32)
    w = np.ascontiguousarray(widths, f32)
    lib().orc_zap(_p(X), _p(b), _p(w), C.c_float(bin_width), len(b), len(X))
    return X


def stats(d):
    d = np.ascontiguousarray(d, f32)
    m, r, s = C.c_float(), C.c_float(), C.c_float()
    lib().orc_stats(_p(d), len(d), C.byref(m), C.byref(r), C.byref(s))
    return f32(m.value), f32(r.value), f32(s.value)


def accel_plan(acc_lo, acc_hi, tol, pulse_width, nsamps, tsamp, cfreq, bw, dm):
    out = np.zeros(1 << 16, f32)
    n = lib().orc_accel_plan(C.c_float(acc_lo), C.c_float(acc_hi), C.c_float(tol), C.c_float(pulse_width), nsamps,
                             C.c_float(tsamp), C.c_float(cfreq), C.c_float(bw), C.c_float(dm), _p(out), len(out))
    return out[:n].copy()


def distill(kind, cands, tol, max_harm_or_tobs, keep_related, fractional=True):
    kinds = {"harmonic": 0, "acceleration": 1, "dm": 2}
    c = np.ascontiguousarray(cands, CAND_DTYPE)
    cap = max(16, 4 * len(c) + 16)
    while True:
        out = np.zeros(cap, CAND_DTYPE)
        n = lib().orc_distill(kinds[kind], C.c_float(tol), C.c_float(max_harm_or_tobs), int(keep_related),
                              int(fractional), _p(c), C.c_long(len(c)), _p(out), C.c_long(cap))
        if n <= cap:
            return out[:n].copy()
        cap = n


# ------------------------------------------------------------------ FFTs (cuFFT stand-ins)
def r2c(x):
    """cufftExecR2C (ffter.hpp:57-61): unnormalised, n/2+1 bins."""
    return np.fft.rfft(np.asarray(x, np.float64)).astype(np.complex64)


def c2r(X, n):
    """cufftExecC2R (ffter.hpp:74-78): unnormalised (x n)."""
    return (np.fft.irfft(np.asarray(X, np.complex128), n) * n).astype(f32)


# ------------------------------------------------------------------ Worker::start
class SearchConfig:
    def __init__(self, size, tsamp, nharmonics=4, min_snr=9.0, min_freq=0.1, max_freq=1100.0, freq_tol=1e-4,
                 max_harm=16, zap=None):
        self.size = int(size)
        self.tsamp = f32(tsamp)
        self.nharmonics = nharmonics
        self.min_snr = f32(min_snr)
        self.min_freq = f32(min_freq)
        self.max_freq = f32(max_freq)
        self.freq_tol = f32(freq_tol)
        self.max_harm = max_harm
        self.zap = zap
        self.nbins = self.size // 2 + 1
        self.tobs = f32(f32(self.size) * self.tsamp)          # pipeline_multi.cu:118
        self.bin_width_f = f32(1.0 / float(self.tobs))         # :119 (double division narrowed)
        self.bin_width = float(self.bin_width_f)               # stored as double


def whiten(trial_u8, cfg: SearchConfig, return_intermediates=False):
    """pipeline_multi.cu:157-204 -> (whitened series [x size], mean, std)."""
    size = cfg.size
    n = min(len(trial_u8), size)
    tim = np.zeros(size, f32)
    tim[:n] = trial_u8[:n]
    if size > len(trial_u8):  # padding path :160-163
        pm = f32(f32(np.sum(tim[:n], dtype=np.float64)) / f32(n))
        tim[n:] = pm
    X = r2c(tim)
    p = form_power(X, 0)
    med = running_median(p, cfg.bin_width)
    X = divide_c_by_f(X, med)
    if cfg.zap is not None and len(cfg.zap[0]):
        X = zap(X, cfg.zap[0], cfg.zap[1], cfg.bin_width_f)
    p = form_power(X, 1)
    mean, rms, std = stats(p)
    tim_w = c2r(X, size)
    if return_intermediates:
        return tim_w, mean, std, dict(X=X, pspec=p, median=med)
    return tim_w, mean, std


def accel_spectrum(tim_w, acc, mean, std, cfg: SearchConfig):
    """pipeline_multi.cu:212-224 -> normalised interbinned spectrum."""
    r = resample(tim_w, acc, cfg.tsamp, which=2)
    X = r2c(r)
    p = form_power(X, 1)
    return normalise(p, f32(mean * f32(cfg.size)), f32(std * f32(cfg.size)))


def spectrum_raw_peaks(p, cfg: SearchConfig):
    """pipeline_multi.cu:228-234 up to device_find_peaks: per level (idx, snr) ascending."""
    sums = harmonic_sum(p, cfg.nharmonics)
    out = []
    for nh in range(cfg.nharmonics + 1):
        spec = p if nh == 0 else sums[nh - 1]
        start, limit, factor = peak_range(cfg.nbins, cfg.bin_width, nh, cfg.min_freq, cfg.max_freq)
        idx, snr = find_peaks(spec, limit, start, cfg.min_snr)
        out.append((idx, snr, factor))
    return out


def trial_candidates(raw, dm, dm_idx, acc):
    recs = []
    for nh, (idx, snr, factor) in enumerate(raw):
        pi, ps_ = cluster_peaks(idx, snr)
        for i, s in zip(pi, ps_):
            recs.append((dm, dm_idx, acc, nh, s, f32(f32(i) * f32(factor)), 1))  # int -> float, float multiply
    return np.array(recs, CAND_DTYPE) if recs else np.zeros(0, CAND_DTYPE)


def search_dm_trial(trial_u8, dm, dm_idx, accs, cfg: SearchConfig):
    """One iteration of the DM loop (pipeline_multi.cu:145-244): flattened dm_trial_cands."""
    tim_w, mean, std = whiten(trial_u8, cfg)
    accel_cands = []
    for acc in accs:
        p = accel_spectrum(tim_w, acc, mean, std, cfg)
        c = trial_candidates(spectrum_raw_peaks(p, cfg), dm, dm_idx, acc)
        d = distill("harmonic", c, cfg.freq_tol, cfg.max_harm, False, True)
        accel_cands.append(d)
    allc = np.concatenate(accel_cands) if accel_cands else np.zeros(0, CAND_DTYPE)
    return distill("acceleration", allc, cfg.freq_tol, cfg.tobs, True)


def search(trials_u8, dm_list, accs, cfg: SearchConfig, acc_plan=None):
    """Worker::start over all trials in ascending DM order (the reference's -t 1 order)."""
    out = []
    for ii, dm in enumerate(dm_list):
        a = accs if accs is not None else acc_plan(dm)
        out.append(search_dm_trial(trials_u8[ii], f32(dm), ii, a, cfg))
    return np.concatenate(out) if out else np.zeros(0, CAND_DTYPE)


# ------------------------------------------------------------------ SIGPROC + dedisp restatement
_INT_KEYS = {"telescope_id", "machine_id", "data_type", "barycentric", "pulsarcentric", "nbits", "nsamples",
             "nchans", "nifs", "nbeams", "ibeam", "signed"}
_DBL_KEYS = {"az_start", "za_start", "src_raj", "src_dej", "tstart", "tsamp", "fch1", "foff", "refdm", "period"}
_STR_KEYS = {"rawdatafile", "source_name"}


def read_sigproc(path):
    """SIGPROC filterbank reader (header.hpp:171-403, filterbank.hpp:218-238)."""
    with open(path, "rb") as f:
        buf = f.read()
    pos = 0

    def rstr():
        nonlocal pos
        (n,) = struct.unpack_from("<i", buf, pos)
        pos += 4
        s = buf[pos:pos + n].decode("latin1")
        pos += n
        return s

    assert rstr() == "HEADER_START"
    hdr = {}
    while True:
        key = rstr()
        if key == "HEADER_END":
            break
        if key in _INT_KEYS:
            if key == "signed":
                hdr[key] = buf[pos]
                pos += 1
            else:
                (hdr[key],) = struct.unpack_from("<i", buf, pos)
                pos += 4
        elif key in _DBL_KEYS:
            (hdr[key],) = struct.unpack_from("<d", buf, pos)
            pos += 8
        elif key in _STR_KEYS:
            hdr[key] = rstr()
        else:
            raise ValueError("unknown sigproc key " + key)
    data = np.frombuffer(buf, np.uint8, offset=pos)
    nbits, nchans = hdr["nbits"], hdr["nchans"]
    nsamps = len(data) * 8 // nbits // nchans
    hdr["nsamples"] = nsamps
    return hdr, data


def unpack(data, nbits, nchans, nsamps):
    """LSB-first sub-byte unpacking -> [nsamps][nchans] uint8."""
    if nbits == 8:
        out = data
    else:
        per = 8 // nbits
        shifts = (np.arange(per, dtype=np.uint8) * nbits)[None, :]
        out = ((data[:, None] >> shifts) & ((1 << nbits) - 1)).reshape(-1)
    return out[:nsamps * nchans].reshape(nsamps, nchans)


def dedisp_dm_list(dm_start, dm_end, dt, ti, f0, df, nchans, tol):
    """dedisp_generate_dm_list (un-vendored library; recurrence as published in dedisp):
    pinned by the 59 DM values in example_output/overview.xml:63-123."""
    dt = dt * 1e6
    f = (f0 + ((nchans / 2) - 0.5) * df) * 1e-3
    tol2 = tol * tol
    a = 8.3 * df / (f * f * f)
    a2 = a * a
    b2 = a2 * (nchans * nchans / 16.0)
    c = (dt * dt + ti * ti) * (tol2 - 1.0)
    dms = [float(dm_start)]
    while dms[-1] < dm_end:
        prev = dms[-1]
        prev2 = prev * prev
        k = c + tol2 * a2 * prev2
        dms.append((b2 * prev + np.sqrt(-a2 * b2 * prev2 + (a2 + b2) * k)) / (a2 + b2))
    return np.array(dms, f32)


def dedisperse(fil_data, hdr, dm_list, killmask=None):
    """Restatement of dedisp_execute as used by dedisperser.hpp:98-113 (8-bit output).

    dedisp is not in the reference tree ("third-party, un-vendored, no pinned version"); the
    delay table, delay rounding and 8-bit output scaling below follow the published dedisp
    algorithm and are pinned only end-to-end through the golden candidate file."""
    nbits, nchans = hdr["nbits"], hdr["nchans"]
    nsamps = hdr["nsamples"]
    x = unpack(fil_data, nbits, nchans, nsamps).astype(np.int32)
    if killmask is not None:
        x = x * np.asarray(killmask, np.int32)[None, :]
    dt = f32(hdr["tsamp"])
    f0, df = f32(hdr["fch1"]), f32(hdr["foff"])
    c = np.arange(nchans, dtype=f32)
    a = (f32(1.0) / (f0 + c * df)).astype(f32)
    b = f32(1.0) / f0
    delay_table = ((4.15e3 / np.float64(dt)) * (a * a - b * b).astype(f32).astype(np.float64)).astype(f32)
    dm_list = np.asarray(dm_list, f32)
    max_delay = int(np.float32(dm_list[-1]) * delay_table[-1] + f32(0.5))
    nout = nsamps - max_delay
    in_range = f32((1 << nbits) - 1)
    out_range = f32(255.0)
    factor = f32(f32(f32(3.0) * f32(1024.0)) / f32(255.0)) / f32(16.0)
    out = np.zeros((len(dm_list), nout), np.uint8)
    for i, dm in enumerate(dm_list):
        delays = np.rint((f32(dm) * delay_table).astype(f32)).astype(np.int64)
        s = np.zeros(nout, np.int64)
        for ch in range(nchans):
            s += x[delays[ch]:delays[ch] + nout, ch]
        scaled = (s.astype(f32) * out_range / f32(in_range * f32(nchans))).astype(f32) * factor
        out[i] = np.clip(scaled, 0.0, out_range).astype(np.uint8)
    return out, max_delay


# ------------------------------------------------------------------ golden parsers
def parse_candidates_peasoup(path):
    """candidates.peasoup (output_stats.hpp:238-268): per candidate optional "FOLD"+nbins+nints+
    floats, then int ndets and ndets 24-byte PODs (candidates.hpp:10-17)."""
    with open(path, "rb") as f:
        buf = f.read()
    pos, cands = 0, []
    while pos < len(buf):
        fold = None
        if buf[pos:pos + 4] == b"FOLD":
            nb, ni = struct.unpack_from("<ii", buf, pos + 4)
            fold = np.frombuffer(buf, f32, nb * ni, pos + 12).reshape(ni, nb)
            pos += 12 + 4 * nb * ni
        (nd,) = struct.unpack_from("<i", buf, pos)
        pos += 4
        dets = np.frombuffer(buf, POD_DTYPE, nd, pos).copy()
        pos += nd * POD_DTYPE.itemsize
        cands.append(dict(fold=fold, dets=dets))
    return cands


def parse_overview_xml(path):
    txt = open(path, encoding="latin1").read()
    dms = [float(v) for v in re.findall(r"<trial id='\d+'>([^<]+)</trial>",
                                        txt[txt.index("<dedispersion_trials"):txt.index("</dedispersion_trials>")])]
    accs = [float(v) for v in re.findall(r"<trial id='\d+'>([^<]+)</trial>",
                                         txt[txt.index("<acceleration_trials"):txt.index("</acceleration_trials>")])]
    params = dict(re.findall(r"<(\w+)>([^<]*)</\1>", txt[txt.index("<search_parameters>"):txt.index("</search_parameters>")]))
    return dict(dm_list=np.array(dms, f32), acc_list=np.array(accs, f32), params=params)


def parse_overview_candidates(path):
    """<candidates> of overview.xml (output_stats.hpp:170-222) as a list of dicts of floats."""
    txt = open(path, encoding="latin1").read()
    out = []
    for block in re.findall(r"<candidate id='\d+'>(.*?)</candidate>", txt, re.S):
        out.append({k: float(v) for k, v in re.findall(r"<(\w+)>([^<]*)</\1>", block)})
    return out


# ------------------------------------------------------------------ after the workers join (row f2)
def _tree(recs):
    """flattened records (is_top marks an entry, associates follow) -> list of [rec, assoc-list]"""
    out = []
    for r in recs:
        node = [r.copy(), []]
        if r["is_top"] or not out:
            out.append(node)
        else:
            out[-1][1].append(node)
    return out


def _flatten_tree(nodes, top=True):
    recs = []
    for rec, assoc in nodes:
        r = rec.copy()
        r["is_top"] = 1 if top else 0
        recs.append(r)
        recs.extend(_flatten_tree(assoc, False))
    return recs


def _distill_tree(nodes, cond):
    """BaseDistiller::distill (distiller.hpp:26-59) on candidates that carry associates."""
    import copy
    nodes = sorted(nodes, key=lambda n: -float(n[0]["snr"]))  # stable; std::sort may order ties differently
    unique = [True] * len(nodes)
    for i in range(len(nodes)):
        if not unique[i]:
            continue
        for j in range(i + 1, len(nodes)):
            for _ in range(cond(nodes[i][0], nodes[j][0])):
                nodes[i][1].append(copy.deepcopy(nodes[j]))
                unique[j] = False
    return [n for n, u in zip(nodes, unique) if u]


def finalise(recs, freq_tol=1e-4, max_harm=16, limit=1000):
    """pipeline_multi.cu:353-365,390-391: DMDistiller(tol, true), HarmonicDistiller(tol, max_harm,
    true, false), truncation to --limit (scores are not part of the record)."""
    upper = float(f32(1) + f32(freq_tol))
    lower = float(f32(1) - f32(freq_tol))

    def dm_cond(a, b):
        ratio = float(b["freq"]) / float(a["freq"])
        return 1 if lower < ratio < upper else 0

    def harm_cond(a, b):
        n = 0
        for jj in range(1, int(max_harm) + 1):
            ratio = float(b["freq"]) / (jj * float(a["freq"]))
            if lower < ratio < upper:
                n += 1
        return n

    nodes = _distill_tree(_tree(recs), dm_cond)
    nodes = _distill_tree(nodes, harm_cond)
    nodes = nodes[:limit]
    flat = _flatten_tree(nodes)
    return np.array(flat, CAND_DTYPE) if flat else np.zeros(0, CAND_DTYPE)
