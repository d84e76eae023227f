/* ref_harness.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * A thin extern "C" shim around the UNMODIFIED reference sources, compiled
 * from where they lie under /root/reference (see oracle/Makefile) into
 * oracle/_ref/libpeasoup_ref.so.  It lets the GPU parity tests and
 * `bench.py --impl reference` run the reference's own kernels (src/kernels.cu),
 * its own transform classes (include/transforms/\*.hpp) and cuFFT on the same
 * buffers as the product library.
 *
 * Nothing here re-implements reference arithmetic: every function forwards to
 * a reference symbol.  `ref_worker_search` replays the statement order of
 * Worker::start (src/pipeline_multi.cu:100-252) with the reference's own
 * objects, because that class lives in the same file as main() and cannot be
 * linked on its own.
 */
#include <data_types/timeseries.hpp>
#include <data_types/fourierseries.hpp>
#include <data_types/candidates.hpp>
#include <transforms/resampler.hpp>
#include <transforms/ffter.hpp>
#include <transforms/spectrumformer.hpp>
#include <transforms/birdiezapper.hpp>
#include <transforms/peakfinder.hpp>
#include <transforms/distiller.hpp>
#include <transforms/harmonicfolder.hpp>
#include <data_types/folded.hpp>
#include <utils/exceptions.hpp>
#include <utils/utils.hpp>
#include <utils/stats.hpp>
#include <kernels/kernels.h>
#include <kernels/defaults.h>
#include <cstring>
/* Dereddener keeps its spliced median private (dereddener.hpp:9-18); the stage test
 * needs to read it, so the class is opened up for this translation unit only. */
#define private public
#include <transforms/dereddener.hpp>
#undef private
#include <transforms/folder.hpp>
#include <string>
#include <vector>

static std::string g_err;
#define REF_TRY try {
#define REF_CATCH                                                     \
  }                                                                   \
  catch (std::exception & e) {                                        \
    g_err = e.what();                                                 \
    return -1;                                                        \
  }                                                                   \
  return 0;

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

/* kernels.h:126 */
int ref_resampleII(float* d_in, float* d_out, size_t n, float a, float tsamp) {
  REF_TRY device_resampleII(d_in, d_out, n, a, tsamp, MAX_THREADS, MAX_BLOCKS);
  REF_CATCH
}
/* kernels.h:118 */
int ref_resample(float* d_in, float* d_out, size_t n, float a, float tsamp) {
  REF_TRY device_resample(d_in, d_out, n, a, tsamp, MAX_THREADS, MAX_BLOCKS);
  REF_CATCH
}
/* kernels.h:111 */
int ref_form_power_series(void* d_c, float* d_out, size_t nbins, int way) {
  REF_TRY device_form_power_series((cufftComplex*)d_c, d_out, nbins, way, MAX_BLOCKS, MAX_THREADS);
  REF_CATCH
}
/* kernels.h:144 */
int ref_normalise(float* d, float mean, float sigma, unsigned nbins) {
  REF_TRY device_normalise(d, mean, sigma, nbins, MAX_BLOCKS, MAX_THREADS);
  REF_CATCH
}
/* kernels.h:104; d_out is nfolds contiguous arrays of nbins floats */
int ref_harmonic_sum(float* d_in, float* d_out, size_t nbins, unsigned nfolds) {
  REF_TRY
  std::vector<float*> h(nfolds);
  for (unsigned i = 0; i < nfolds; i++) h[i] = d_out + (size_t)i * nbins;
  float** d_tab;
  Utils::device_malloc<float*>(&d_tab, nfolds);
  Utils::h2dcpy<float*>(d_tab, &h[0], nfolds);
  device_harmonic_sum(d_in, d_tab, nbins, nfolds, MAX_BLOCKS, MAX_THREADS);
  Utils::device_free(d_tab);
  REF_CATCH
}
/* kernels.h:134; returns count via *count, host outputs sized >= n */
int ref_find_peaks(int n, int start, float* d_dat, float thresh, int* h_idx, float* h_snr,
                   int* count) {
  REF_TRY
  thrust::device_vector<int> d_i(n);
  thrust::device_vector<float> d_s(n);
  cached_allocator alloc;
  *count = device_find_peaks(n, start, d_dat, thresh, h_idx, h_snr, d_i, d_s, alloc);
  REF_CATCH
}
/* kernels.h:252,256 */
int ref_median_scrunch5(const float* d_in, size_t count, float* d_out) {
  REF_TRY median_scrunch5(d_in, count, d_out);
  ErrorChecker::check_cuda_error("ref_median_scrunch5");
  REF_CATCH
}
int ref_linear_stretch(const float* d_in, size_t in_count, float* d_out, size_t out_count) {
  REF_TRY linear_stretch(d_in, in_count, d_out, out_count);
  ErrorChecker::check_cuda_error("ref_linear_stretch");
  REF_CATCH
}
/* kernels.h:262,268 */
int ref_divide_c_by_f(void* d_c, float* d_f, unsigned nbins) {
  REF_TRY device_divide_c_by_f((cuComplex*)d_c, d_f, nbins, MAX_BLOCKS, MAX_THREADS);
  REF_CATCH
}
int ref_zap_birdies(void* d_c, float* d_birdies, float* d_widths, float bin_width,
                    unsigned nbirdies, unsigned nbins) {
  REF_TRY device_zap_birdies((cuComplex*)d_c, d_birdies, d_widths, bin_width, nbirdies, nbins,
                             MAX_BLOCKS, MAX_THREADS);
  REF_CATCH
}
/* kernels.h:279-292 */
int ref_stats(float* d, unsigned n, float* mean, float* rms, float* std) {
  REF_TRY stats::stats<float>(d, n, mean, rms, std);
  REF_CATCH
}
/* kernels.h:318 */
int ref_conversion_u8(unsigned char* d_in, float* d_out, unsigned n) {
  REF_TRY device_conversion<unsigned char, float>(d_in, d_out, n, MAX_BLOCKS, MAX_THREADS);
  REF_CATCH
}
/* ffter.hpp:46-78 */
int ref_r2c(unsigned n, float* d_in, void* d_out) {
  REF_TRY
  CuFFTerR2C f(n);
  f.execute(d_in, (cufftComplex*)d_out);
  ErrorChecker::check_cuda_error("ref_r2c");
  REF_CATCH
}
int ref_c2r(unsigned n, void* d_in, float* d_out) {
  REF_TRY
  CuFFTerC2R f(n);
  f.execute((cufftComplex*)d_in, d_out);
  ErrorChecker::check_cuda_error("ref_c2r");
  REF_CATCH
}
/* dereddener.hpp:41-62; copies the spliced median out to d_median_out (nbins floats) */
int ref_running_median(float* d_pspec_in, unsigned nbins, double bin_width, float* d_median_out,
                       void* d_fseries_inout) {
  REF_TRY
  DevicePowerSpectrum<float> p(nbins, bin_width);
  Utils::d2dcpy(p.get_data(), d_pspec_in, nbins);
  Dereddener red(nbins);
  red.calculate_median(p);
  if (d_fseries_inout) {
    DeviceFourierSeries<cufftComplex> f(nbins, bin_width);
    Utils::d2dcpy(f.get_data(), (cufftComplex*)d_fseries_inout, nbins);
    red.deredden(f);
    Utils::d2dcpy((cufftComplex*)d_fseries_inout, f.get_data(), nbins);
  }
  if (d_median_out) Utils::d2dcpy(d_median_out, red.median, nbins);
  REF_CATCH
}

struct RefSearchParams {
  unsigned size;        /* transform length (pipeline_multi.cu:326-331) */
  unsigned nsamps;      /* samples per dedispersed trial */
  float tsamp;
  int nharmonics;       /* number of folds */
  float min_snr, min_freq, max_freq, freq_tol;
  int max_harm;
  const char* zapfile;  /* may be NULL or "" */
};

/* Replays Worker::start (pipeline_multi.cu:100-252) on host trials with an
 * explicit acceleration list per DM trial (acc_list, n_acc shared by all DMs).
 * Output: flattened dm_trial_cands; for each top-level candidate its assoc
 * list follows it (is_top = 0).  Returns number of records or -1. */
long ref_worker_search(const RefSearchParams* p, unsigned char* h_trials, const float* dm_list,
                       int ndm, const float* acc_list, int n_acc, long max_out, float* o_dm,
                       int* o_dm_idx, float* o_acc, int* o_nh, float* o_snr, float* o_freq,
                       int* o_is_top, double* seconds_out) {
  try {
    std::vector<float> dms(dm_list, dm_list + ndm);
    DispersionTrials<unsigned char> trials(h_trials, p->nsamps, p->tsamp, dms);
    unsigned size = p->size;
    bool padding = size > trials.get_nsamps();
    CuFFTerR2C r2cfft(size);
    CuFFTerC2R c2rfft(size);
    float tobs = size * trials.get_tsamp();
    float bin_width = 1.0 / tobs;
    DeviceFourierSeries<cufftComplex> d_fseries(size / 2 + 1, bin_width);
    DedispersedTimeSeries<unsigned char> tim;
    ReusableDeviceTimeSeries<float, unsigned char> d_tim(size);
    DeviceTimeSeries<float> d_tim_r(size);
    TimeDomainResampler resampler;
    DevicePowerSpectrum<float> pspec(d_fseries);
    Zapper* bzap = NULL;
    std::string zapfile = p->zapfile ? p->zapfile : "";
    if (zapfile != "") bzap = new Zapper(zapfile);
    Dereddener rednoise(size / 2 + 1);
    SpectrumFormer former;
    PeakFinder cand_finder(p->min_snr, p->min_freq, p->max_freq, size);
    HarmonicSums<float> sums(pspec, p->nharmonics);
    HarmonicFolder harm_folder(sums);
    std::vector<float> accs(acc_list, acc_list + n_acc);
    HarmonicDistiller harm_finder(p->freq_tol, p->max_harm, false);
    AccelerationDistiller acc_still(tobs, p->freq_tol, true);
    CandidateCollection dm_trial_cands;
    float mean, std, rms, padding_mean;

    cudaDeviceSynchronize();
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int ii = 0; ii < ndm; ii++) {
      trials.get_idx(ii, tim);
      d_tim.copy_from_host(tim);
      if (padding) {
        padding_mean = stats::mean<float>(d_tim.get_data(), trials.get_nsamps());
        d_tim.fill(trials.get_nsamps(), d_tim.get_nsamps(), padding_mean);
      }
      r2cfft.execute(d_tim.get_data(), d_fseries.get_data());
      former.form(d_fseries, pspec);
      rednoise.calculate_median(pspec);
      rednoise.deredden(d_fseries);
      if (bzap) bzap->zap(d_fseries);
      former.form_interpolated(d_fseries, pspec);
      stats::stats<float>(pspec.get_data(), size / 2 + 1, &mean, &rms, &std);
      c2rfft.execute(d_fseries.get_data(), d_tim.get_data());
      CandidateCollection accel_trial_cands;
      for (size_t jj = 0; jj < accs.size(); jj++) {
        resampler.resampleII(d_tim, d_tim_r, size, accs[jj]);
        r2cfft.execute(d_tim_r.get_data(), d_fseries.get_data());
        former.form_interpolated(d_fseries, pspec);
        stats::normalise(pspec.get_data(), mean * size, std * size, size / 2 + 1);
        harm_folder.fold(pspec);
        SpectrumCandidates trial_cands(tim.get_dm(), ii, accs[jj]);
        cand_finder.find_candidates(pspec, trial_cands);
        cand_finder.find_candidates(sums, trial_cands);
        accel_trial_cands.append(harm_finder.distill(trial_cands.cands));
      }
      dm_trial_cands.append(acc_still.distill(accel_trial_cands.cands));
    }
    cudaDeviceSynchronize();
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (seconds_out) *seconds_out = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
    if (bzap) delete bzap;

    long n = 0;
    for (size_t i = 0; i < dm_trial_cands.cands.size(); i++) {
      std::vector<CandidatePOD> flat;
      dm_trial_cands.cands[i].collect_candidates(flat);
      for (size_t k = 0; k < flat.size(); k++) {
        if (n < max_out) {
          o_dm[n] = flat[k].dm; o_dm_idx[n] = flat[k].dm_idx; o_acc[n] = flat[k].acc;
          o_nh[n] = flat[k].nh; o_snr[n] = flat[k].snr; o_freq[n] = flat[k].freq;
          o_is_top[n] = (k == 0);
        }
        n++;
      }
    }
    return n;
  } catch (std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

/* Per-DM whitening chain only (pipeline_multi.cu:157-204): returns the whitened
 * time series (size floats, x size scaling of unnormalised C2R) and mean/std. */
int ref_whiten(const RefSearchParams* p, unsigned char* h_trial, float* d_tim_out, float* mean_out,
               float* std_out, void* d_fseries_out, float* d_pspec_out) {
  REF_TRY
  unsigned size = p->size;
  std::vector<float> dms(1, 0.f);
  DispersionTrials<unsigned char> trials(h_trial, p->nsamps, p->tsamp, dms);
  bool padding = size > trials.get_nsamps();
  CuFFTerR2C r2cfft(size);
  CuFFTerC2R c2rfft(size);
  float tobs = size * trials.get_tsamp();
  float bin_width = 1.0 / tobs;
  DeviceFourierSeries<cufftComplex> d_fseries(size / 2 + 1, bin_width);
  DedispersedTimeSeries<unsigned char> tim;
  ReusableDeviceTimeSeries<float, unsigned char> d_tim(size);
  DevicePowerSpectrum<float> pspec(d_fseries);
  Zapper* bzap = NULL;
  std::string zapfile = p->zapfile ? p->zapfile : "";
  if (zapfile != "") bzap = new Zapper(zapfile);
  Dereddener rednoise(size / 2 + 1);
  SpectrumFormer former;
  float mean, std, rms;
  trials.get_idx(0, tim);
  d_tim.copy_from_host(tim);
  if (padding) {
    float pm = stats::mean<float>(d_tim.get_data(), trials.get_nsamps());
    d_tim.fill(trials.get_nsamps(), d_tim.get_nsamps(), pm);
  }
  r2cfft.execute(d_tim.get_data(), d_fseries.get_data());
  former.form(d_fseries, pspec);
  rednoise.calculate_median(pspec);
  rednoise.deredden(d_fseries);
  if (bzap) bzap->zap(d_fseries);
  former.form_interpolated(d_fseries, pspec);
  stats::stats<float>(pspec.get_data(), size / 2 + 1, &mean, &rms, &std);
  if (d_fseries_out) Utils::d2dcpy((cufftComplex*)d_fseries_out, d_fseries.get_data(), size / 2 + 1);
  if (d_pspec_out) Utils::d2dcpy(d_pspec_out, pspec.get_data(), size / 2 + 1);
  c2rfft.execute(d_fseries.get_data(), d_tim.get_data());
  ErrorChecker::check_cuda_error("ref_whiten");
  Utils::d2dcpy(d_tim_out, d_tim.get_data(), size);
  *mean_out = mean;
  *std_out = std;
  if (bzap) delete bzap;
  REF_CATCH
}

/* ---- rows f3 / f4 of SURVEY 8: folding, fold optimisation, coincidence matching -------------- */

/* device_fold_timeseries (src/kernels.cu:634-651), launched the way TimeSeriesFolder::fold does
 * (folder.hpp:48-61: max_blocks = nints, max_threads = nbins) */
int ref_fold_timeseries(float* d_in, float* d_out, size_t nsamps, size_t nsubints, double period, double tsamp,
                        int nbins) {
  REF_TRY
  device_fold_timeseries(d_in, d_out, nsamps, nsubints, period, tsamp, nbins, nsubints, nbins);
  REF_CATCH
}

/* FoldOptimiser::optimise (folder.hpp:241-351) with the reference's own object on d_fold[nints][nbins] */
int ref_fold_optimise(const float* d_fold, unsigned nbins, unsigned nints, double period, float tobs, float* opt_sn,
                      double* opt_period, int* opt_width, int* opt_bin, float* h_opt_fold, float* h_opt_prof) {
  REF_TRY
  FoldedSubints<float> fold(nbins, nints);
  Utils::d2dcpy(fold.get_data(), const_cast<float*>(d_fold), (size_t)nbins * nints);
  fold.set_period(period);
  fold.set_tobs(tobs);
  FoldOptimiser opt(nbins, nints);
  opt.optimise(fold);
  *opt_sn = fold.get_opt_sn();
  *opt_period = fold.get_opt_period();  /* NB: the getter narrows to float (folded.hpp:43) */
  *opt_width = fold.get_opt_width();
  *opt_bin = fold.get_opt_bin();
  if (h_opt_fold) memcpy(h_opt_fold, &fold.opt_fold[0], sizeof(float) * nbins * nints);
  if (h_opt_prof) memcpy(h_opt_prof, &fold.opt_prof[0], sizeof(float) * nbins);
  REF_CATCH
}

/* device_coincidencer (src/kernels.cu:1073-1100); d_arrays = device table of narrays device pointers.
 * The reference kernel has no bounds check: callers pad the arrays to a multiple of MAX_THREADS. */
int ref_coincidencer(float** d_arrays, float* d_out, int narrays, size_t size, float thresh, int beam_thresh) {
  REF_TRY
  device_coincidencer(d_arrays, d_out, narrays, size, thresh, beam_thresh, MAX_BLOCKS, MAX_THREADS);
  REF_CATCH
}

}  // extern "C"
