"""CPU restatement of the reference's multibeam coincidencer.

TEST INFRASTRUCTURE ONLY: nothing under peasoup_b200/ may import this.

Follows src/coincidencer.cpp:122-214 (per-beam whitening and normalisation), coincidence_kernel
(src/kernels.cu:1073-1085) and Coincidencer::write_birdie_list / write_samp_mask
(include/transforms/coincidencer.hpp:41-77).  The transforms are numpy's (any length, double,
rounded to float32) standing in for cuFFT.

Pinned by the reference run here: match() against the unmodified coincidence_kernel of
oracle/_ref on the same arrays (tests/test_gpu_coincidence.py::
test_coincidencer_kernel_equals_reference_kernel, bit-exact including threshold ties and NaNs);
every per-beam stage of whiten_beam() is one of the staged functions that tests/test_gpu_stages.py
pins against the same reference build, and by the golden candidate file (oracle.py).  The reference
holds no golden output for the program as a whole, so the file writers are checked against the
reference's format only (coincidencer.hpp:41-77).
"""
import numpy as np

from . import oracle as o

f32 = np.float32


def whiten_beam(trial_u8, tsamp):
    """coincidencer.cpp:160-177 -> (normalised time series [size], normalised interbinned spectrum)."""
    size = len(trial_u8)
    tim = np.asarray(trial_u8, f32)
    tobs = f32(f32(size) * f32(tsamp))                 # :140
    bin_width = f32(1.0 / float(tobs))                 # :141 (float)
    X = o.r2c(tim)
    p = o.form_power(X, 0)
    med = o.running_median(p, float(bin_width))        # default boundaries: -l/-a are never applied (:170)
    X = o.divide_c_by_f(X, med)
    p = o.form_power(X, 1)
    mean, _, std = o.stats(p)
    p = o.normalise(p, mean, std)
    t = o.c2r(X, size)
    mean, _, std = o.stats(t)
    t = o.normalise(t, mean, std)
    return t, p, bin_width


def match(arrays, thresh, beam_thresh):
    """coincidence_kernel: 1 where fewer than beam_thresh arrays exceed thresh."""
    count = np.zeros(len(arrays[0]), np.int64)
    for a in arrays:
        count += np.asarray(a, f32) > f32(thresh)
    return (count < beam_thresh).astype(f32)


def birdies_from_mask(mask, bin_width):
    """coincidencer.hpp:54-77: one (freq, width) per run of zeros."""
    out = []
    ii, size = 0, len(mask)
    bw = float(f32(bin_width))
    while ii < size:
        if mask[ii] == 0:
            count = 0
            while ii < size and mask[ii] == 0:
                count += 1
                ii += 1
            out.append((f32(((ii - 1) - count / 2.0) * bw), f32(count * f32(bin_width))))
        else:
            ii += 1
    return out


def birdie_text(birdies):
    return "".join("%.9f\t%.6f\n" % (float(f), float(w)) for f, w in birdies)


def samp_mask_text(mask):
    return "#0 1\n" + "".join("%d\n" % int(v) for v in mask)
