"""CPU restatement of the reference's candidate folding and fold optimisation.

TEST INFRASTRUCTURE ONLY, like oracle.py: nothing under peasoup_b200/ may import this.

Follows include/transforms/folder.hpp (TimeSeriesFolder :34-62, FoldOptimiser :65-352, MultiFolder
:354-460) and the kernels behind it (src/kernels.cu:597-865).  cuFFT C2C is stood in for by numpy's
FFT in double precision rounded to complex64 after every transform, so values agree with the
reference to float rounding, not bit for bit.

Parity pinned: tests/test_oracle_golden.py folds the ten golden candidates of
example_output/candidates.peasoup from the golden DM trials and compares the FOLD blocks,
folded_snr and opt_period with the golden files.
"""
import numpy as np

from . import oracle as o

f32 = np.float32
c64 = np.complex64

NBINS = 64   # folder.hpp:383-384
NINTS = 16
FOLD_CHUNK = 2048   # csrc/fold.cu: samples per partial sum


def prev_power_of_two(n):
    """Utils::prev_power_of_two (utils.hpp): largest power of two <= n."""
    return 1 << (int(n).bit_length() - 1)


def fold_whiten(trial_u8, nsamps, tsamp):
    """folder.hpp:406-413: u8 -> float, R2C, |X|, running median, deredden, C2R.  No zapping, no
    interbinning, no normalisation; the series comes back scaled by nsamps (unnormalised C2R)."""
    tim = np.asarray(trial_u8[:nsamps], f32)
    tobs = f32(f32(nsamps) * f32(tsamp))          # :375  float tobs = nsamps*tsamp
    bin_width = 1.0 / float(tobs)                  # :378  1.0/tobs kept as double
    X = o.r2c(tim)
    p = o.form_power(X, 0)
    med = o.running_median(p, bin_width)
    X = o.divide_c_by_f(X, med)
    return o.c2r(X, nsamps)


def fold_series(x, period, tsamp, nbins=NBINS, nints=NINTS):
    """fold_time_series_kernel (kernels.cu:597-633), one block per sub-integration, nbins threads:
    thread t of sub-integration s adds samples s*per + t + nbins*m for m < ceil(per/nbins); the
    per-bin hit count starts at ONE (so each bin is sum/(hits+1)).  The reference's shared-memory
    atomics have no fixed order; the order here is the one csrc/fold.cu uses: chunks of FOLD_CHUNK
    samples summed in sample order, then the chunk sums added in chunk order (float32 throughout).
    Reads are clamped to the series."""
    x = np.asarray(x, f32)
    n = len(x)
    per = n // nints
    tsamp_by_period = float(tsamp) / float(period)            # :642 double / double
    out = np.zeros((nints, nbins), f32)
    m = -(-per // nbins)
    for s in range(nints):
        jj = (s * per + np.arange(nbins)[None, :] + nbins * np.arange(m)[:, None]).ravel()
        jj = np.sort(jj[jj < n])
        frac = np.modf(jj.astype(np.float64) * tsamp_by_period)[0]
        idx = np.floor(frac * nbins).astype(np.int64)
        sums = np.zeros(nbins, f32)
        for c0 in range(0, len(jj), FOLD_CHUNK):
            cj, ci = jj[c0:c0 + FOLD_CHUNK], idx[c0:c0 + FOLD_CHUNK]
            part = np.zeros(nbins, f32)
            for b in np.unique(ci):
                part[b] = np.cumsum(x[cj[ci == b]], dtype=f32)[-1]   # cumsum adds strictly left to right
            sums = (sums + part).astype(f32)
        count = np.bincount(idx, minlength=nbins).astype(np.int64) + 1
        out[s] = sums / count.astype(f32)
    return out


def _fft(a, inverse=False):
    a = np.asarray(a, np.complex128)
    r = np.fft.ifft(a, axis=-1) * a.shape[-1] if inverse else np.fft.fft(a, axis=-1)
    return r.astype(c64)


class FoldOptimiser:
    """folder.hpp:65-352: template bank (boxcars of width 1..nbins-1), drift bank (nbins linear phase
    drifts across the sub-integrations), both applied in the Fourier domain of the phase axis."""

    def __init__(self, nbins=NBINS, nints=NINTS):
        self.nbins, self.nints = nbins, nints
        self.ntemplates = nbins - 1                               # :113 generate_templates(step=1)
        self.nshifts = nbins                                      # :123
        t = np.arange(self.ntemplates)[:, None]
        b = np.arange(nbins)[None, :]
        self.templates = _fft((b <= t).astype(c64))               # kernels.cu:680-691 + forward FFT
        # shift array, kernels.cu:659-678
        shifts = (np.arange(self.nshifts) - self.nshifts // 2).astype(f32)   # :129
        two_pi = f32(2 * 3.14159265359)                           # kernels.cu:786
        subint = np.arange(nints, dtype=f32)
        frac = (subint / f32(nints)).astype(f32)                  # subint/nints, float
        shift = (frac[None, :] * shifts[:, None]).astype(f32)     # [nshifts][nints]
        ramp = (np.arange(nbins, dtype=f32) * two_pi / f32(nbins)).astype(f32)
        ramp = np.where(np.arange(nbins) > nbins // 2, (ramp - two_pi).astype(f32), ramp)
        arg = (f32(-1) * ramp[None, None, :] * shift[:, :, None]).astype(f32)
        self.shiftar = (np.cos(arg.astype(np.float64)) + 1j * np.sin(arg.astype(np.float64))).astype(c64)

    @staticmethod
    def calculate_sn(prof, bin_, width, nbins):
        """folder.hpp:143-191, with its integer/float conversions."""
        edge = int(width * 0.3 + 0.5)
        width_by_2 = int(width / 2.0 + 0.5)
        jj = (bin_ - nbins // 2 + np.arange(nbins)) % nbins       # unsigned wrap then % nbins
        rprof = np.asarray(prof, f32)[jj]
        centre = nbins // 2 - 1
        upper, lower = centre + (width_by_2 + edge), centre - (width_by_2 + edge)
        ii = np.arange(nbins)
        on = (ii <= upper) & (ii >= lower)
        on_pulse, off_pulse = rprof[on], rprof[~on]
        with np.errstate(all="ignore"):
            on_mean = f32(np.sum(on_pulse, dtype=np.float64) / len(on_pulse)) if len(on_pulse) else f32(np.nan)
            off_mean = f32(np.sum(off_pulse, dtype=np.float64) / len(off_pulse)) if len(off_pulse) else f32(np.nan)
            acc = f32(0)
            for v in off_pulse:                                    # float acc += double pow
                acc = f32(float(acc) + float(f32(v - off_mean)) ** 2)
            off_std = np.sqrt(f32(acc / f32(len(off_pulse)))) if len(off_pulse) else f32(np.nan)
            sn1 = f32(float(f32(on_mean - off_mean)) * np.sqrt(float(width)) / float(off_std))
            r = ((rprof - off_mean).astype(f32) / off_std).astype(f32)
            sn2 = f32(np.sum(r, dtype=np.float64) / np.sqrt(float(width)))
        if sn1 > 99999:
            sn1 = f32(0)
        if sn2 > 99999:
            sn2 = f32(0)
        return sn1, sn2

    def optimise(self, fold, period, tobs):
        """folder.hpp:241-351.  fold [nints][nbins] float; period double; tobs float."""
        nbins, nints, nshifts, nt = self.nbins, self.nints, self.nshifts, self.ntemplates
        inp = _fft(np.asarray(fold, f32).astype(c64))                              # :249-254
        post = (inp[None, :, :] * self.shiftar).astype(c64)                        # :258 [nshifts][nints][nbins]
        prof = np.zeros((nshifts, nbins), c64)                                     # :263 collapse, subint order
        for i in range(nints):
            prof = (prof + post[:, i, :]).astype(c64)
        norm = np.sqrt((np.arange(nt) + 1.0).astype(f32)).astype(f32)              # kernels.cu:731-742
        final = (prof[None, :, :] * self.templates[:, None, :]).astype(c64) / norm[:, None, None]
        final = final.astype(c64)
        final[:, :, 0] = 0
        mag = np.abs(_fft(final, inverse=True)).astype(f32)                        # :276-280
        argmax = int(np.argmax(mag.ravel()))                                       # first maximum
        opt_template = argmax // (nbins * nshifts)
        opt_bin = argmax % nbins - opt_template // 2
        opt_shift = argmax // nbins % nbins
        opt_fold = _fft(post[opt_shift], inverse=True).real.astype(f32)            # :292-295
        opt_prof = _fft(prof[opt_shift], inverse=True).real.astype(f32)            # :304-311
        sn1, sn2 = self.calculate_sn(opt_prof, opt_bin, opt_template, nbins)       # width = opt_template
        p = float(period)
        opt_period = p * ((((32.0 - opt_shift) * p) / float(f32(f32(nbins) * f32(tobs)))) + 1)   # :346
        return dict(opt_sn=f32(max(sn1, sn2)), opt_period=f32(opt_period),         # get_opt_period() is float
                    opt_width=opt_template + 1, opt_bin=opt_bin, opt_shift=opt_shift,
                    opt_fold=opt_fold, opt_prof=opt_prof, sn=(sn1, sn2), peak=float(mag.ravel()[argmax]))


def fold_candidates(tops, trials_u8, tsamp, npdmp, min_period=0.001, max_period=10.0):
    """MultiFolder (folder.hpp:354-460).  `tops`: list of dicts with freq, dm_idx, acc, snr (the
    final candidate list, descending S/N).  Returns (per-candidate results or None, new order)."""
    nsamps = prev_power_of_two(trials_u8.shape[1])
    tsamp = f32(tsamp)
    tobs = f32(f32(nsamps) * tsamp)                       # folder.hpp:50-51: float product
    opt = FoldOptimiser()
    results = [None] * len(tops)
    by_dm = {}
    for i in range(min(npdmp, len(tops))):
        p = f32(1.0 / float(tops[i]["freq"]))
        if f32(min_period) < p < f32(max_period):
            by_dm.setdefault(int(tops[i]["dm_idx"]), []).append(i)
    for dm_idx in sorted(by_dm):
        tim = fold_whiten(trials_u8[dm_idx], nsamps, tsamp)
        for i in by_dm[dm_idx]:
            period = f32(1.0 / float(tops[i]["freq"]))    # float period (folder.hpp:387,419)
            r = o.resample(tim, f32(tops[i]["acc"]), tsamp, which=1)
            fold = fold_series(r, float(period), float(tsamp))
            res = opt.optimise(fold, float(period), tobs)
            res["fold"] = fold
            results[i] = res
    key = [max(float(t["snr"]), float(results[i]["opt_sn"]) if results[i] else 0.0) for i, t in enumerate(tops)]
    order = sorted(range(len(tops)), key=lambda i: -key[i])   # less_than_key, folder.hpp:25-31
    return results, order
