/* peasoup_b200.h -- C ABI of libpeasoup_b200.so
 *
 * Drop-in boundary for the per-DM acceleration-search hot path of peasoup.  Every entry
 * point replaces one free function of the reference's kernel API
 * (/root/reference/include/kernels/kernels.h) or one cuFFT call site
 * (include/transforms/ffter.hpp); the citation beside each declaration names it.
 *
 * Conventions
 *  - plain pointers and sizes only; all `d_*` pointers are device pointers on the current
 *    CUDA device, `h_*` are host pointers;
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream).  Unlike the
 *    reference (which ends every wrapper in cudaDeviceSynchronize, utils/exceptions.hpp:64-74)
 *    calls are asynchronous unless they return a value to the host;
 *  - return value: 0 on success, a negative PS_E* code on failure; ps_last_error() returns a
 *    thread-local message with the same prefixes the reference's ErrorChecker uses
 *    ("CUDA failed with error: ...").  The C++ wrappers in include/peasoup_b200/ rethrow it
 *    as std::runtime_error like the reference does;
 *  - there is no CPU fallback: without a CUDA device every compute call fails with PS_ECUDA.
 */
#ifndef PEASOUP_B200_H
#define PEASOUP_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PS_OK 0
#define PS_ECUDA (-1)   /* CUDA runtime error (message in ps_last_error) */
#define PS_EINVAL (-2)  /* bad argument / unsupported size */
#define PS_EOVERFLOW (-3) /* more peaks than the workspace holds */

typedef struct { float re, im; } ps_complex; /* layout of cufftComplex / float2 */

const char* ps_last_error(void);
int ps_version(void);

/* ---------------- staged kernel API (one function per reference function) ------------- */

/* device_resampleII, kernels.h:126 / kernels.cu:314-317,335-362.
 * d_out[i] = d_in[rn(i + i*af*(i-n))], af = double(a*tsamp)/(2c); bit-exact index map. */
int ps_resampleII(const float* d_in, float* d_out, size_t n, float a, float tsamp, void* stream);
/* device_resample, kernels.h:118 / kernels.cu:308-311,320-332,364-379 (v1 index map). */
int ps_resample(const float* d_in, float* d_out, size_t n, float a, float tsamp, void* stream);
/* Index maps only (uint32 read index for every output sample); test/diagnostic aid for the
 * bit-exact parity gate.  which = 1 -> resample (v1), 2 -> resampleII. */
int ps_resample_index_map(uint32_t* d_idx, size_t n, float a, float tsamp, int which, void* stream);

/* device_form_power_series, kernels.h:111 / kernels.cu:215-252,287-304.
 * way = 0: |X| (z*rsqrtf(z)); way = 1: interbinned amplitude. */
int ps_form_power_series(const ps_complex* d_in, float* d_out, size_t nbins, int way, void* stream);

/* device_normalise, kernels.h:144 / kernels.cu:469-494: d[i] = (d[i]-mean)/sigma. */
int ps_normalise(float* d_powers, float mean, float sigma, unsigned nbins, void* stream);

/* device_harmonic_sum, kernels.h:104 / kernels.cu:33-99,199-208.
 * h_out_ptrs: HOST array of nfolds DEVICE pointers (the reference re-uploads a device table
 * per call, harmonicfolder.hpp:29-33; here the pointers travel as kernel arguments). nfolds<=5. */
int ps_harmonic_sum(const float* d_in, float* const* h_out_ptrs, size_t nbins, unsigned nfolds,
                    void* stream);

/* device_find_peaks, kernels.h:134 / kernels.cu:384-416.  Stream-compacts (idx,value) with
 * value > thresh for idx in [start,n) in ascending idx order into the host arrays (capacity
 * max_out each).  The Thrust vectors + cached_allocator of the reference become an opaque
 * workspace.  Returns the count (>=0) or a negative error; synchronises the stream. */
typedef struct ps_peak_workspace ps_peak_workspace;
int ps_peak_workspace_create(size_t max_peaks, ps_peak_workspace** out);
void ps_peak_workspace_destroy(ps_peak_workspace* ws);
int ps_find_peaks(int n, int start, const float* d_dat, float thresh, int* h_idx, float* h_snr,
                  size_t max_out, ps_peak_workspace* ws, void* stream);
/* Number of above-threshold bins the last ps_find_peaks call on this workspace counted -- after
 * PS_EOVERFLOW, the capacity a new workspace needs (the reference has no such path: its
 * PeakFinder writes past its fixed 100000-entry host arrays, peakfinder.hpp:59-70,84). */
size_t ps_peak_workspace_needed(const ps_peak_workspace* ws);

/* median_scrunch5 / linear_stretch, kernels.h:252,256 / kernels.cu:875-1011. */
int ps_median_scrunch5(const float* d_in, size_t count, float* d_out, void* stream);
int ps_linear_stretch(const float* d_in, size_t in_count, float* d_out, size_t out_count,
                      void* stream);
/* Dereddener::calculate_median, dereddener.hpp:41-62, as ONE call: three scrunch levels and
 * the spliced 5/25/125 stretch evaluated on the fly.  d_work: >= nbins/5+nbins/25+nbins/125+3
 * floats. */
int ps_running_median(const float* d_pspec, unsigned nbins, double bin_width, float boundary_5_freq,
                      float boundary_25_freq, float* d_median, float* d_work, void* stream);
/* ps_running_median without any allocation and without a work buffer (no counterpart in the
 * reference; added for this library's users).
 *
 * For callers that must not allocate device memory and have no spare buffer to lend: a
 * real-time acquisition loop where a cudaMalloc would stall every stream of the device, a
 * plugin running under a host that owns the whole device-memory budget, or code that runs
 * while other work on the device must not be disturbed.  The three median levels are kept in
 * an array built into the library (about 4 MiB per GPU, held from library load to exit), so
 * the call issues only kernel launches: no cudaMalloc, cudaFree or allocating library routine.
 *
 *  - nbins above PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS (a 2^23-point search) is rejected with
 *    PS_EINVAL and a message naming the limit;
 *  - the first call on each device looks up the array's device address (one
 *    cudaGetSymbolAddress, with that device current); make that first call before any stream
 *    capture begins;
 *  - all calls on a device share the one array: calls whose kernels may overlap on the GPU
 *    (different streams without ordering between them) race -- serialise them.  Nothing has to
 *    be cleared between calls: every level is rewritten before it is read;
 *  - the array used is that of the CURRENT device: d_pspec and d_median must live on it (the call
 *    does not check where they live);
 *  - results are identical to ps_running_median on the same input. */
#define PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS 4194305u
int ps_running_median_noalloc(const float* d_pspec, unsigned nbins, double bin_width,
                              float boundary_5_freq, float boundary_25_freq, float* d_median,
                              void* stream);
/* device_divide_c_by_f, kernels.h:262 / kernels.cu:1013-1034. */
int ps_divide_c_by_f(ps_complex* d_c, const float* d_f, unsigned nbins, void* stream);
/* device_zap_birdies, kernels.h:268 / kernels.cu:1036-1069 (width is a half-width). */
int ps_zap_birdies(ps_complex* d_c, const float* d_birdies, const float* d_widths, float bin_width,
                   unsigned nbirdies, unsigned nbins, void* stream);

/* GPU_rms / GPU_mean / GPU_fill, kernels.h:279-292 / kernels.cu:426-463.  Synchronous. */
int ps_rms(const float* d, int nsamps, int min_bin, float* out);
int ps_mean(const float* d, int nsamps, int min_bin, float* out);
int ps_fill(float* d_start, float* d_end, float value, void* stream);
/* device_conversion<unsigned char,float>, kernels.h:318 / kernels.cu:1144-1170. */
int ps_convert_u8_f32(const unsigned char* d_in, float* d_out, unsigned n, void* stream);

/* ---------------- FFT (replaces cuFFT, ffter.hpp:46-78) ------------------------------- */
typedef struct ps_fft_plan ps_fft_plan;
/* cufftPlan1d(&plan, n, CUFFT_R2C/C2R, 1): powers of two 2^17..2^25 run on the two-pass engine, any
 * other n in [2, 2^24] as a chirp-z convolution on top of it (bluestein.cu). */
int ps_fft_plan_create(unsigned n, ps_fft_plan** out);
void ps_fft_plan_destroy(ps_fft_plan* plan);
/* cufftExecR2C: n floats -> n/2+1 complex, unnormalised. */
int ps_fft_r2c(ps_fft_plan* plan, const float* d_in, ps_complex* d_out, void* stream);
/* cufftExecC2R: n/2+1 complex -> n floats, unnormalised (x n). */
int ps_fft_c2r(ps_fft_plan* plan, const ps_complex* d_in, float* d_out, void* stream);

/* ---------------- fused search path (Worker::start, pipeline_multi.cu:100-252) -------- */
typedef struct {
  unsigned size;        /* transform length, --fft_size / prev_power_of_two (pipeline_multi.cu:326-331) */
  unsigned nsamps;      /* samples per dedispersed trial (may be < or > size: padding path :112-114) */
  float tsamp;          /* seconds, float as in timeseries.hpp:54 */
  int nharmonics;       /* number of folds 0..5 (--nharmonics) */
  float min_snr;        /* --min_snr */
  float min_freq;       /* --min_freq */
  float max_freq;       /* --max_freq */
  float boundary_5_freq;  /* 0.05 (the reference never forwards the CLI value, dereddener.hpp:41-43) */
  float boundary_25_freq; /* 0.5 */
  unsigned max_peaks;   /* capacity of the per-DM raw peak list (0 -> default 1<<20) */
} ps_search_params;

/* One record per above-threshold bin, before clustering (peakfinder.hpp:27-56 runs on host). */
typedef struct {
  int32_t acc_idx;  /* index into the acceleration list of this DM trial */
  int32_t nh;       /* 0 = fundamental spectrum, k = k-th fold */
  int32_t bin;      /* spectrum bin */
  float snr;        /* normalised amplitude */
} ps_raw_peak;

typedef struct ps_search ps_search;
int ps_search_create(const ps_search_params* p, ps_search** out);
void ps_search_destroy(ps_search* s);
/* zaplist (birdiezapper.hpp:34-57): nbirdies host floats each; copies to the device. */
int ps_search_set_zaplist(ps_search* s, const float* h_freqs, const float* h_widths, unsigned n);
/* Per-DM whitening chain, pipeline_multi.cu:157-204: u8 trial (DEVICE pointer, nsamps bytes)
 * -> whitened time series kept inside the handle, plus mean/std of the whitened interbinned
 * spectrum (device-resident; also copied to *h_mean,*h_std if non-NULL, which synchronises). */
int ps_search_whiten(ps_search* s, const unsigned char* d_trial, float* h_mean, float* h_std,
                     void* stream);
/* Acceleration loop, pipeline_multi.cu:209-239 up to device_find_peaks: for each acceleration
 * resample -> R2C -> interbin -> normalise -> harmonic sums -> threshold.  Appends raw peaks to
 * the handle's device list.  Asynchronous. */
int ps_search_accel_batch(ps_search* s, const float* h_acc, int n_acc, int acc_idx0, void* stream);
/* Reset / fetch the device peak list (fetch synchronises; records are sorted by
 * (acc_idx, nh, bin), i.e. the order the reference's per-spectrum copy_if calls produce). */
int ps_search_reset_peaks(ps_search* s, void* stream);
long ps_search_fetch_peaks(ps_search* s, ps_raw_peak* h_out, size_t max_out, void* stream);
/* After ps_search_fetch_peaks returned PS_EOVERFLOW: the number of records the last batch produced,
 * and a way to enlarge the device list to hold them (synchronises the device; the list's content is
 * lost, so the caller resets the peaks and repeats ps_search_accel_batch -- the whitened series is
 * still there).  The pipeline driver does exactly that, so a DM trial full of interference costs
 * time instead of failing; the reference has no overflow handling: beyond 100000 above-threshold
 * bins per spectrum it writes past its host arrays (peakfinder.hpp:59-70,84). */
unsigned ps_search_peaks_needed(ps_search* s);
int ps_search_reserve_peaks(ps_search* s, unsigned max_peaks);
/* The harmonic-sum + threshold + compaction stage on its own (HarmonicFolder::fold +
 * PeakFinder::find_candidates, harmonicfolder.hpp:26-38 + peakfinder.hpp:72-94, without
 * materialising the sums): nbatch normalised spectra of size/2+1 bins, `stride` floats apart,
 * spectrum b reported as acceleration index acc_idx0 + b.  Peaks join the handle's peak list
 * (ps_search_reset_peaks / ps_search_fetch_peaks).  Spectra that are 16-byte aligned, a multiple of
 * 4 floats apart and followed by >= 4 readable floats each are summed from TMA-staged operands
 * (harmonic_rows.cuh); any other layout takes the gather kernel.  Same output either way.
 * Also BASELINE.json configs[2]'s microbench. */
int ps_search_harmonic_peaks(ps_search* s, const float* d_spectra, size_t stride, int nbatch,
                             int acc_idx0, void* stream);
/* Diagnostics for stage-by-stage parity tests: device pointers owned by the handle. */
const float* ps_search_whitened_series(ps_search* s);   /* size floats */
const float* ps_search_last_spectrum(ps_search* s);     /* size/2+1 floats (last accel trial) */
/* Index map of the resampling gather that ps_search_accel_batch runs (TimeDomainResampler::resampleII,
 * resampler.hpp:26-31 -> resample_kernelII / getAcceleratedIndexII, kernels.cu:314-317,335-346): the
 * search path's own loader (fft.cuh LoadResampled: float32 shift estimate, FP64 expression near a
 * rounding boundary) is executed over a series whose sample i holds the integer i, so d_idx[i]
 * (DEVICE, size uint32) is the index that loader reads for output sample i at acceleration `acc`.
 * *h_exact_pairs (optional) receives how many sample pairs evaluated the FP64 expression instead of
 * the float32 estimate.  Diagnostics: allocates a scratch series and synchronises. */
int ps_search_resample_index_map(ps_search* s, float acc, uint32_t* d_idx, unsigned* h_exact_pairs,
                                 void* stream);
/* kernel launch counter (all ps_* kernels launched by this process; for bench.py) */
unsigned long long ps_launch_count(void);
/* host<->device traffic issued by the pipeline driver: add to / read the process-wide counters */
void ps_count_transfer(unsigned long long h2d_bytes, unsigned long long d2h_bytes);
void ps_transfer_bytes(unsigned long long* h2d_bytes, unsigned long long* d2h_bytes);

/* ---------------- pipeline driver + host-side logic ----------------------------------- */
typedef struct {
  ps_search_params search;
  float freq_tol;        /* --freq_tol */
  int max_harm;          /* --max_harm_match */
  /* AccelerationPlan (utils.hpp:140-193); used when no explicit acceleration list is given */
  float acc_start, acc_end, acc_tol, acc_pulse_width;
  float cfreq, foff;     /* filterbank centre frequency / channel width (MHz) */
} ps_pipeline_params;

typedef struct {
  float dm;
  int32_t dm_idx;
  float acc;
  int32_t nh;
  float snr;
  float freq;
  int32_t is_top; /* 1: entry of dm_trial_cands; 0: member of the preceding entry's assoc list */
} ps_candidate;

/* Worker::start + thread fan-out (pipeline_multi.cu:100-252, 342-359) over `ndm` host-resident
 * 8-bit trials [ndm][nsamps], on devices device0 .. device0+nworkers-1 (one host thread each,
 * dynamic DM dispatch).  Candidates of all workers are concatenated in ascending dm_idx order
 * (the reference's -t 1 order) and flattened depth-first like Candidate::collect_candidates.
 * dm_idx_offset is added to every dm_idx (for DM-sharded multi-process runs).
 * Returns the number of records (may exceed max_out: then only max_out were written) or <0.
 * trials_done (optional) receives the number of DM x acceleration trials searched. */
long ps_pipeline_search(const ps_pipeline_params* p, const unsigned char* h_trials, const float* h_dm_list,
                        int ndm, int dm_idx_offset, const float* h_acc_explicit, int n_acc_explicit,
                        const float* h_zap_freqs, const float* h_zap_widths, int nzap, int device0,
                        int nworkers, ps_candidate* out, long max_out, unsigned long long* trials_done);

/* The same with a long-lived handle: ps_pipeline_open builds every worker's device state once
 * (search handles, FFT plans, streams, staging); ps_pipeline_run may then be called repeatedly
 * (e.g. per batch of DM trials as dedispersion delivers them); results are those of
 * ps_pipeline_search.  Not re-entrant per handle. */
typedef struct ps_pipeline ps_pipeline;
int ps_pipeline_open(const ps_pipeline_params* p, const float* h_zap_freqs, const float* h_zap_widths, int nzap,
                     int device0, int nworkers, ps_pipeline** out);
long ps_pipeline_run(ps_pipeline* h, const unsigned char* h_trials, const float* h_dm_list, int ndm,
                     int dm_idx_offset, const float* h_acc_explicit, int n_acc_explicit, ps_candidate* out,
                     long max_out, unsigned long long* trials_done);
void ps_pipeline_close(ps_pipeline* h);

/* ---------------- the stages either side of the hot path ---------- */
/* dedisp_generate_dm_list as called at dedisperser.hpp:54-62 (external dedisp library, restated).
 * dt seconds, ti = pulse width (us), f0/df MHz.  Returns the list length (out may be NULL). */
int ps_dedisp_generate_dm_list(float dm_start, float dm_end, float dt, float ti, float f0, float df,
                               int nchans, float tol, float* out, int max_out);
/* dedisp_get_max_delay (dedisperser.hpp:100). */
long ps_dedisp_max_delay(const float* h_dm_list, int ndm, int nchans, float dt, float f0, float df);
/* dedisp_execute with 8-bit output (dedisperser.hpp:98-113): h_in = packed filterbank
 * [nsamps][nchans] (nbits 1/2/4/8), h_out = [ndm][nout] with nout = nsamps - max_delay.  Host
 * pointers; synchronous. */
int ps_dedisperse(const unsigned char* h_in, size_t nsamps, int nchans, int nbits, float dt, float f0,
                  float df, const float* h_dm_list, int ndm, const int* h_killmask,
                  unsigned char* h_out, size_t nout);
/* The same on a given device (dedisp_create_plan_multi's device index, dedisperser.hpp:21-30;
 * device < 0: the current one) and with a cap on the device memory it may hold (0: 60 % of what is
 * free).  The observation is processed in time gulps of as many samples as fit -- packed input,
 * unpacked copy and ndm output rows per gulp -- like dedisp's own gulping, so long files and long DM
 * lists take more gulps instead of failing to allocate.  Channel frequencies must descend from f0
 * (df < 0) and the DM list must be non-negative and ascending: delays are unsigned sample counts. */
int ps_dedisperse_on(int device, size_t max_device_bytes, const unsigned char* h_in, size_t nsamps,
                     int nchans, int nbits, float dt, float f0, float df, const float* h_dm_list, int ndm,
                     const int* h_killmask, unsigned char* h_out, size_t nout);

typedef struct {
  ps_pipeline_params pipe;  /* search.size = --fft_size (0: prev_power_of_two); nsamps, tsamp, cfreq, foff come from the file */
  float dm_start, dm_end, dm_tol, dm_pulse_width; /* --dm_start --dm_end --dm_tol --dm_pulse_width */
  int limit;                /* --limit (< 0: no limit) */
  int nworkers, device0;    /* -t */
  const char* killfile;     /* -k, may be NULL */
  const char* zapfile;      /* -z, may be NULL */
  const char* outdir;       /* -o: candidates.peasoup + overview.xml are written when non-NULL */
  int npdmp;                /* --npdmp: fold and optimise this many top candidates (0: none) */
} ps_filterbank_params;
/* main() of src/pipeline_multi.cu:262-419: read the SIGPROC filterbank, dedisperse, search every
 * DM trial, DM + harmonic distillation, scoring, folding of the top npdmp candidates
 * (MultiFolder, folder.hpp:354-460; re-sorts by max(snr, folded_snr)), --limit.
 * out: final candidates flattened depth-first (is_top = 1 starts a candidate).  seconds4
 * (optional): reading, dedispersion, searching, total.  folded_snr / opt_period and the FOLD
 * blocks go to the output files; ps_fold_results_last() returns them to the caller. */
long ps_search_filterbank(const char* filename, const ps_filterbank_params* p, ps_candidate* out,
                          long max_out, float* dm_list_out, int max_dm, int* ndm_out, double* seconds4);

/* ---- folding ------------------------------------------------------------ */
/* device_fold_timeseries (kernels.h; src/kernels.cu:597-653): d_out[nsubints][nbins], each bin the
 * sum of its samples / (hits + 1).  Deterministic (sample order), see csrc/fold.cu. */
int ps_fold_timeseries(const float* d_in, float* d_out, size_t nsamps, size_t nsubints, double period,
                       double tsamp, int nbins, void* stream);
typedef struct ps_fold_optimiser ps_fold_optimiser;
typedef struct {
  float opt_sn, sn1, sn2;   /* max(sn1, sn2) and the two estimators of folder.hpp:143-191 */
  double opt_period;        /* folder.hpp:346 */
  int opt_width, opt_bin, opt_shift;
  unsigned argmax;          /* flat index into [template][drift][bin] */
  float peak;               /* value at argmax */
} ps_fold_result;
/* FoldOptimiser::FoldOptimiser (folder.hpp:195-217): template and drift banks for nbins x nints. */
int ps_fold_optimiser_create(unsigned nbins, unsigned nints, ps_fold_optimiser** out);
void ps_fold_optimiser_destroy(ps_fold_optimiser* o);
/* FoldOptimiser::optimise (folder.hpp:241-351): d_fold [nints][nbins] on the device; h_opt_fold
 * [nints][nbins] and h_opt_prof [nbins] (host, optional) receive the optimised sub-integrations
 * and profile.  Synchronises the stream. */
int ps_fold_optimise(ps_fold_optimiser* o, const float* d_fold, double period, float tobs,
                     float* h_opt_fold, float* h_opt_prof, ps_fold_result* res, void* stream);
/* Fold results of the last ps_search_filterbank call of this process: for final candidate i
 * (i < returned count) folded_snr[i], opt_period[i]; folds [count][16*64] when non-NULL (zeros for
 * candidates that were not folded).  Returns the number of final candidates. */
long ps_fold_results_last(float* folded_snr, double* opt_period, float* folds, long max_count);

/* ---- multibeam coincidencer ---------------------------------------------- */
/* device_coincidencer (kernels.h; src/kernels.cu:1073-1100): d_arrays = DEVICE table of narrays
 * device pointers; d_out[i] = 1 when fewer than beam_thresh arrays exceed thresh at i, else 0. */
int ps_coincidencer(float* const* d_arrays, float* d_out, int narrays, size_t size, float thresh,
                    int beam_thresh, void* stream);
typedef struct {
  float threshold;               /* --thresh (4.0) */
  int beam_threshold;            /* --beam_thresh (4) */
  const char* samp_outfilename;  /* --o: sample mask file, NULL or "" = do not write */
  const char* spec_outfilename;  /* --o2: birdie list file, NULL or "" = do not write */
} ps_coincidence_params;
/* main() of src/coincidencer.cpp:122-214: every filterbank dedispersed at DM 0, whitened,
 * normalised; sample mask [nsamps] and spectral mask [nsamps/2+1] by ps_coincidencer; files as the
 * reference writes them.  Masks are also copied to h_samp_mask / h_spec_mask when non-NULL (up to
 * max_samp / max_spec values).  Returns nsamps, or a negative error. */
long ps_coincidence_files(const char* const* filenames, int nfiles, const ps_coincidence_params* p,
                          float* h_samp_mask, size_t max_samp, float* h_spec_mask, size_t max_spec);

/* Diagnostics of the last ps_pipeline_search call of this process, summed over workers:
 * seconds spent queueing device work, waiting for the device, and in host-side clustering +
 * distillation, then the number of raw above-threshold bins fetched. */
void ps_pipeline_profile(double* out4);

/* AccelerationPlan::generate_accel_list (utils.hpp:168-192).  Returns the list length. */
int ps_accel_plan_generate(float acc_lo, float acc_hi, float tol, float pulse_width, unsigned nsamps,
                           float tsamp, float cfreq, float bw, float dm, float* out, int max_out);
/* PeakFinder::identify_unique_peaks (peakfinder.hpp:27-56). Returns the number of clusters. */
int ps_cluster_peaks(const int* idxs, const float* snrs, int count, int min_gap, int* out_idx,
                     float* out_snr);
/* PeakFinder::find_candidates bin range / frequency factor (peakfinder.hpp:77-92). */
int ps_peak_range(int nbins, double bin_width, int nh, float min_freq, float max_freq, int* start_idx,
                  int* limit, float* factor);
/* Distillers (distiller.hpp): kind 0 = HarmonicDistiller(tol,max_harm,keep_related,fractional),
 * 1 = AccelerationDistiller(tobs,tol,keep_related), 2 = DMDistiller(tol,keep_related).
 * in: n candidates (is_top ignored); out: flattened survivors (+ assoc if keep_related). */
long ps_distill(int kind, float tol, float max_harm_or_tobs, int keep_related, int fractional_harms,
                const ps_candidate* in, long n, ps_candidate* out, long max_out);
/* The two distillations after the workers have joined (src/pipeline_multi.cu:353-365:
 * DMDistiller(freq_tol, true) then HarmonicDistiller(freq_tol, max_harm, true, false)) on the MERGED
 * per-DM candidate list.  in: flattened records as ps_pipeline_run / ps_pipeline_search return them
 * (is_top = 1 starts a candidate, its acceleration associates follow), in ascending dm_idx order --
 * after a multi-GPU run that is the gathered list in canonical order.  out: the surviving
 * candidates flattened depth-first.  Returns the number of records (may exceed max_out: call again). */
long ps_distill_merged(const ps_candidate* in, long n, float freq_tol, int max_harm, ps_candidate* out,
                       long max_out);

#ifdef __cplusplus
}
#endif
#endif /* PEASOUP_B200_H */
