/* oracle.cpp -- CPU restatement of the reference's per-DM acceleration-search arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, imported by or executed
 * from the product library (libpeasoup_b200.so); only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may call it, as the checker.
 *
 * Each function restates one piece of /root/reference and cites the lines it follows.  The
 * operand types, evaluation order and the fused multiply-adds are those nvcc emits for the
 * reference (verified in the SASS of the reference's kernels.cu built for sm_100a, see
 * DESIGN.md "Oracle"); the file is compiled with -ffp-contract=off so that only the explicit
 * fma()/fmaf() calls fuse.  FFTs are not here: oracle.py uses numpy's double-precision
 * pocketfft and rounds to float32 (the reference uses cuFFT; parity for FFT-derived values is
 * by tolerance, never bit-exact).
 *
 * Pinning: oracle.py's end-to-end run of example_data/tutorial.fil reproduces the reference's
 * golden example_output/candidates.peasoup (tests/test_oracle_golden.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>

extern "C" {

/* ---- resampling: kernels.cu:308-317, 354, 369-370 ---- */
static double accel_fact(float a, float tsamp) { return ((a * tsamp) / (2 * 299792458.0)); }

void orc_resample_index(int which, size_t n, float a, float tsamp, uint64_t* out) {
  double af = accel_fact(a, tsamp);
  double size = (double)n, h = (double)n / 2.0;
  for (size_t i = 0; i < n; i++) {
    double di = (double)i, r;
    if (which == 2)
      r = fma(di * af, di - size, di); /* DMUL, DADD, DFMA in the reference's SASS */
    else
      r = fma(af, fma(di - h, di - h, -(h * h)), di);
    out[i] = (uint64_t)nearbyint(r); /* __double2ull_rn: ties to even */
  }
}

void orc_resample(int which, const float* in, float* out, size_t n, float a, float tsamp) {
  std::vector<uint64_t> idx(n);
  orc_resample_index(which, n, a, tsamp, idx.data());
  for (size_t i = 0; i < n; i++) out[i] = in[idx[i]];
}

/* ---- spectra: kernels.cu:215-252 ---- */
void orc_form_power(const float* c, float* out, size_t nbins, int way) {
  for (size_t i = 0; i < nbins; i++) {
    float re = c[2 * i], im = c[2 * i + 1];
    float ampsq = fmaf(re, re, im * im);
    if (way == 0) {
      out[i] = ampsq * (1.0f / sqrtf(ampsq)); /* z*rsqrtf(z): NaN at z == 0 like the reference */
    } else {
      float re_l = i ? c[2 * i - 2] : 0.f, im_l = i ? c[2 * i - 1] : 0.f;
      float dr = re - re_l, di = im - im_l;
      float diff = fmaf(dr, dr, di * di) * 0.5f;
      out[i] = sqrtf(fmaxf(ampsq, diff));
    }
  }
}

/* ---- kernels.cu:469-480 ---- */
void orc_normalise(float* d, float mean, float sigma, size_t n) {
  for (size_t i = 0; i < n; i++) d[i] = (d[i] - mean) / sigma;
}

/* ---- harmonic sums: kernels.cu:33-99.  out = nfolds arrays of n floats ---- */
void orc_harmonic_sum(const float* in, float* out, size_t n, int nfolds) {
  const double s2 = 1.0 / sqrt(2.0), s8 = 1.0 / sqrt(8.0), s32 = 1.0 / sqrt(32.0);
  for (size_t i = 0; i < n; i++) {
    int idx = (int)i;
    float val = in[idx];
    if (nfolds > 0) {
      val += in[(int)(idx * 0.5 + 0.5)];
      out[0 * n + i] = (float)(val * s2);
    }
    if (nfolds > 1) {
      val += in[(int)(idx * 0.75 + 0.5)];
      val += in[(int)(idx * 0.25 + 0.5)];
      out[1 * n + i] = (float)(val * 0.5);
    }
    if (nfolds > 2) {
      for (int m = 1; m < 8; m += 2) val += in[(int)(idx * (m / 8.0) + 0.5)];
      out[2 * n + i] = (float)(val * s8);
    }
    if (nfolds > 3) {
      for (int m = 1; m < 16; m += 2) val += in[(int)(idx * (m / 16.0) + 0.5)];
      out[3 * n + i] = (float)(val * 0.25);
    }
    if (nfolds > 4) {
      for (int m = 1; m < 32; m += 2) val += in[(int)(idx * (m / 32.0) + 0.5)];
      out[4 * n + i] = (float)(val * s32);
    }
  }
}

/* ---- peaks: kernels.cu:384-416 (copy_if, strict >, ascending index) ---- */
int orc_find_peaks(int n, int start, const float* d, float thresh, int* idx, float* snr, int max_out) {
  int c = 0;
  for (int i = start; i < n; i++)
    if (d[i] > thresh) {
      if (c < max_out) {
        idx[c] = i;
        snr[c] = d[i];
      }
      c++;
    }
  return c;
}

/* ---- peakfinder.hpp:27-56 ---- */
int orc_cluster_peaks(const int* idxs, const float* snrs, int count, int min_gap, int* oidx, float* osnr) {
  int np = 0, ii = 0;
  while (ii < count) {
    float cpeak = snrs[ii];
    int cidx = idxs[ii], last = idxs[ii];
    ii++;
    while (ii < count && (idxs[ii] - last) < min_gap) {
      if (snrs[ii] > cpeak) {
        cpeak = snrs[ii];
        cidx = idxs[ii];
        last = idxs[ii];
      }
      ii++;
    }
    oidx[np] = cidx;
    osnr[np] = cpeak;
    np++;
  }
  return np;
}

/* ---- peakfinder.hpp:77-92 ---- */
void orc_peak_range(int size, double bin_width, int nh, float min_freq, float max_freq, int* start_idx, int* limit,
                    float* factor) {
  float nyquist = bin_width * size;
  int orig_size = 2.0 * (size - 1.0);
  int max_bin = (int)((max_freq / bin_width) * pow(2.0, nh));
  *start_idx = (int)(orig_size * (min_freq / nyquist) * pow(2.0, nh));
  *limit = size < max_bin ? size : max_bin;
  *factor = 1.0 / size * nyquist / pow(2.0, (float)nh);
}

/* ---- rednoise: kernels.cu:875-1011, dereddener.hpp:41-62 ---- */
static float med5(const float* p) {
  float v[5] = {p[0], p[1], p[2], p[3], p[4]};
  std::sort(v, v + 5);
  return v[2];
}
void orc_median_scrunch5(const float* in, size_t count, float* out) {
  if (count == 1) {
    out[0] = in[0];
  } else if (count == 2) {
    out[0] = 0.5f * (in[0] + in[1]);
  } else if (count == 3) {
    float v[3] = {in[0], in[1], in[2]};
    std::sort(v, v + 3);
    out[0] = v[1];
  } else if (count == 4) {
    float v[4] = {in[0], in[1], in[2], in[3]};
    std::sort(v, v + 4);
    out[0] = 0.5f * (v[1] + v[2]);
  } else {
    size_t oc = count / 5; /* tail dropped */
    for (size_t i = 0; i < oc; i++) out[i] = med5(in + 5 * i);
  }
}

void orc_linear_stretch(const float* in, size_t in_count, float* out, size_t out_count) {
  float step = (float)(in_count - 1) / (out_count - 1);
  for (size_t i = 0; i < out_count; i++) {
    float x = (unsigned)i * step;
    unsigned j = (unsigned)x;
    float frac = x - j;
    /* separate multiply and add in the reference's SASS; the reference may read in[in_count]
       at the last element (out of bounds) -- clamped here */
    float b = (size_t)j + 1 < in_count ? in[j + 1] : in[j];
    out[i] = in[j] + ((frac > 1e-5f) ? frac * (b - in[j]) : 0.f);
  }
}

void orc_running_median(const float* pspec, unsigned nbins, double bin_width, float b5, float b25, float* median) {
  int pos5 = (int)(b5 / bin_width), pos25 = (int)(b25 / bin_width);
  size_t n5 = nbins / 5, n25 = n5 / 5, n125 = n25 / 5;
  std::vector<float> m5(n5 + 1), m25(n25 + 1), m125(n125 + 1), tmp(nbins);
  orc_median_scrunch5(pspec, nbins, m5.data());
  orc_median_scrunch5(m5.data(), n5, m25.data());
  orc_median_scrunch5(m25.data(), n25, m125.data());
  orc_linear_stretch(m5.data(), n5, tmp.data(), nbins);
  memcpy(median, tmp.data(), sizeof(float) * pos5);
  orc_linear_stretch(m25.data(), n25, tmp.data(), nbins);
  memcpy(median + pos5, tmp.data() + pos5, sizeof(float) * pos25);
  orc_linear_stretch(m125.data(), n125, tmp.data(), nbins);
  memcpy(median + pos25, tmp.data() + pos25, sizeof(float) * (nbins - pos25));
}

/* kernels.cu:1013-1023 with cuCdivf (cuComplex.h) written out for a real divisor */
void orc_divide_c_by_f(float* c, const float* f, unsigned n) {
  for (unsigned i = 0; i < n; i++) {
    if (i < 5) {
      c[2 * i] = 0.f;
      c[2 * i + 1] = 0.f;
      continue;
    }
    float yr = f[i], yi = 0.f;
    float s = fabsf(yr) + fabsf(yi);
    float oos = 1.0f / s;
    float ars = c[2 * i] * oos, ais = c[2 * i + 1] * oos;
    float brs = yr * oos, bis = yi * oos;
    s = fmaf(brs, brs, bis * bis);
    oos = 1.0f / s;
    c[2 * i] = fmaf(ars, brs, ais * bis) * oos;
    c[2 * i + 1] = fmaf(ais, brs, -(ars * bis)) * oos;
  }
}

/* kernels.cu:1036-1058 */
void orc_zap(float* c, const float* birdies, const float* widths, float bin_width, unsigned nb, unsigned nbins) {
  for (unsigned b = 0; b < nb; b++) {
    int lo = (int)floorf((birdies[b] - widths[b]) / bin_width);
    int hi = (int)ceilf((birdies[b] + widths[b]) / bin_width);
    if (lo < 0) lo = 0;
    if ((unsigned)lo >= nbins) continue;
    if ((unsigned)hi >= nbins) hi = nbins - 1;
    for (int i = lo; i < hi; i++) {
      c[2 * i] = 1.f;
      c[2 * i + 1] = 0.f;
    }
  }
}

/* kernels.cu:426-454 + stats.hpp:20-33.  The reference accumulates in float in Thrust's
   (unpinned) tree order; the sums are taken in double here and narrowed where the reference
   narrows. */
void orc_stats(const float* d, unsigned n, float* mean, float* rms, float* std) {
  double s = 0, q = 0;
  for (unsigned i = 0; i < n; i++) {
    s += d[i];
    q += (double)d[i] * d[i];
  }
  *rms = sqrtf((float)q / (float)(int)n);
  *mean = (float)s / (float)(int)n;
  *std = sqrtf(*rms * *rms - *mean * *mean);
}

/* kernels.cu:1144-1152 */
void orc_convert_u8(const unsigned char* in, float* out, size_t n) {
  for (size_t i = 0; i < n; i++) out[i] = in[i];
}

/* ---- utils.hpp:140-193 ---- */
int orc_accel_plan(float acc_lo, float acc_hi, float tol, float pulse_width, unsigned nsamps, float tsamp, float cfreq,
                   float bw_in, float dm, float* out, int max_out) {
  float bw = fabs(bw_in);
  float tobs = nsamps * tsamp;
  int n = 0;
#define PUSH(v)                \
  do {                         \
    if (n < max_out) out[n] = (v); \
    n++;                       \
  } while (0)
  if (acc_hi == acc_lo) {
    PUSH(0.0f);
    return n;
  }
  float tdm = pow(8.3 * bw / pow(cfreq, 3.0) * dm, 2.0);
  float tpulse = pulse_width * pulse_width; /* member keeps the CLI value: utils.hpp:158-165 */
  float ttsamp = tsamp * tsamp;
  float w_us = sqrt(tdm + tpulse + ttsamp);
  float alt_a = 2.0 * w_us * 1.0e-6 * 24.0 * 299792458.0 / tobs / tobs * sqrt((tol * tol) - 1.0);
  if (acc_hi != 0 && acc_lo != 0) PUSH(0.0f);
  float acc = acc_lo;
  while (acc < acc_hi) {
    PUSH(acc);
    acc += alt_a;
  }
  PUSH(acc_hi);
#undef PUSH
  return n;
}

/* ---- distillers: distiller.hpp:16-197.  Records: dm, dm_idx, acc, nh, snr, freq, is_top ---- */
struct OrcCand {
  float dm;
  int32_t dm_idx;
  float acc;
  int32_t nh;
  float snr;
  float freq;
  int32_t is_top;
};

long orc_distill(int kind, float tol, float max_harm_or_tobs, int keep_related, int fractional, const OrcCand* in, long n,
                 OrcCand* out, long max_out) {
  std::vector<OrcCand> c(in, in + n);
  std::sort(c.begin(), c.end(), [](const OrcCand& x, const OrcCand& y) { return x.snr > y.snr; });
  std::vector<char> unique(n, 1);
  std::vector<std::vector<long> > assoc(n);
  double upper_tol = 1 + tol, lower_tol = 1 - tol;
  for (long idx = 0; idx < n; idx++) {
    if (!unique[idx]) continue;
    double fundi_freq = c[idx].freq;
    if (kind == 0) {
      float max_harm = max_harm_or_tobs;
      for (long ii = idx + 1; ii < n; ii++) {
        double freq = c[ii].freq;
        float max_den = fractional ? (float)pow(2.0, c[ii].nh) : 1.f;
        for (int jj = 1; jj <= max_harm; jj++)
          for (int kk = 1; kk <= max_den; kk++) {
            double ratio = kk * freq / (jj * fundi_freq);
            if (ratio > lower_tol && ratio < upper_tol) {
              if (keep_related) assoc[idx].push_back(ii);
              unique[ii] = 0;
            }
          }
      }
    } else if (kind == 1) {
      float tobs = max_harm_or_tobs;
      double tobs_over_c = tobs / 299792458.0;
      double fundi_acc = c[idx].acc;
      double edge = fundi_freq * tol;
      for (long ii = idx + 1; ii < n; ii++) {
        double delta_acc = fundi_acc - c[ii].acc;
        double acc_freq = (float)(fundi_freq + delta_acc * fundi_freq * tobs_over_c);
        double f = c[ii].freq;
        bool hit = acc_freq > fundi_freq ? (f > fundi_freq - edge && f < acc_freq + edge)
                                         : (f < fundi_freq + edge && f > acc_freq - edge);
        if (hit) {
          if (keep_related) assoc[idx].push_back(ii);
          unique[ii] = 0;
        }
      }
    } else {
      for (long ii = idx + 1; ii < n; ii++) {
        double ratio = c[ii].freq / fundi_freq;
        if (ratio > lower_tol && ratio < upper_tol) {
          if (keep_related) assoc[idx].push_back(ii);
          unique[ii] = 0;
        }
      }
    }
  }
  long m = 0;
  for (long i = 0; i < n; i++) {
    if (!unique[i]) continue;
    if (m < max_out) {
      out[m] = c[i];
      out[m].is_top = 1;
    }
    m++;
    for (size_t k = 0; k < assoc[i].size(); k++, m++)
      if (m < max_out) {
        out[m] = c[assoc[i][k]];
        out[m].is_top = 0;
      }
  }
  return m;
}

}  /* extern "C" */
