// pipeline.hpp -- host-side mirror of the reference's multi-GPU search driver
// (src/pipeline_multi.cu:33-259: DMDispenser, Worker, launch_worker_thread).
//
// Worker::start keeps the reference's statement order per DM trial
//   copy trial -> [pad] -> accel list -> whiten (R2C, |X|, running median, deredden, zap,
//   interbin, stats, C2R) -> for each acceleration: resample, R2C, interbin, normalise,
//   harmonic sums, peaks, harmonic distill -> acceleration distill
// but drives the fused device path (ps_search_*): all device work of a DM trial is queued on
// one stream without host synchronisation, two DM trials are in flight per worker (double
// buffering), and the host-side clustering/distillation of trial i overlaps the device work of
// trial i+1.  Candidate output is identical in content and order to the staged path.
#pragma once
#include <chrono>
#include <cstdlib>
#include <mutex>
#include <thread>
#include "transforms.hpp"

namespace peasoup_b200 {

// subset of CmdLineOptions (utils/cmdline.hpp:12-47) read by the Worker
struct SearchArgs {
  unsigned int size = 0;       // --fft_size (0: prev_power_of_two(nsamps))
  int nharmonics = 4;          // -n: number of folds
  float min_snr = 9.0f;        // -m
  float min_freq = 0.1f;       // --min_freq
  float max_freq = 1100.0f;    // --max_freq
  float freq_tol = 1e-4f;      // --freq_tol
  int max_harm = 16;           // --max_harm_match
  std::string zapfilename;     // -z
  std::vector<float> zap_freqs, zap_widths;  // alternative to zapfilename
  unsigned int max_peaks = 0;  // raw above-threshold bins kept per DM trial (0 -> 1<<20)
  bool verbose = false;
};

class DMDispenser {  // pipeline_multi.cu:33-81 (progress bar omitted)
  std::mutex mutex;
  int dm_idx;
  int count;

 public:
  template <class Trials>
  explicit DMDispenser(Trials& trials) : dm_idx(0), count((int)trials.get_count()) {}
  explicit DMDispenser(int count) : dm_idx(0), count(count) {}
  int get_dm_trial_idx() {
    std::lock_guard<std::mutex> lk(mutex);
    if (dm_idx >= count) return -1;
    return dm_idx++;
  }
};

// Device-side state of one worker: up to kMaxSlots DM trials in flight (search handle, stream, staging).
// Creating it costs ~40 cudaMalloc calls and table uploads, so a long-lived pipeline keeps it
// across calls (ps_pipeline_open); a one-shot Worker builds its own.
class WorkerSlots {
 public:
  struct Slot {
    ps_search* search = 0;
    cudaStream_t stream = 0;
    unsigned char* d_trial = 0;
    std::vector<ps_raw_peak> peaks;
    std::vector<float> acc_list;
    int ii = -1;
    float dm = 0.f;
  };
  enum { kMaxSlots = 4 };
  Slot slots[kMaxSlots];
  // recorded behind every slot's device work: the next slot's device work waits for it.  The kernels
  // are persistent over the whole GPU with a static work partition; kernels of two DM trials side by
  // side (the other trial's small whitening kernels taking SMs for a few microseconds at a time)
  // delay single CTAs and with them the whole kernel: measured 5.59 ms per DM trial with two free-running
  // slots against 5.16 ms = whitening + chain with one.  The slots still overlap the host's share.
  cudaEvent_t chain_done = 0;
  bool chain_recorded = false;
  bool serial_chains = true;
  int nslots;
  int device;
  ps_search_params prm;
  size_t ncopy;

  WorkerSlots(int device, const ps_search_params& prm, const std::vector<float>& zf, const std::vector<float>& zw)
      : device(device), prm(prm), ncopy(std::min(prm.nsamps, prm.size)) {
    // DM trials in flight per worker: the next trial's copy is queued and its kernels wait behind the
    // previous trial's (chain_done) while the host distills earlier ones.  With the kernels serialised,
    // ms per DM trial at 2/3/4 slots, one box: 5.43 / 5.35 / 5.37 over 64 DM trials, 5.47 / 5.18 over
    // 16 (the third slot hides the host's share of the first and last trials); PS_SLOTS overrides.
    nslots = 3;
    if (const char* e = getenv("PS_SLOTS")) nslots = std::max(1, std::min((int)kMaxSlots, atoi(e)));
    if (const char* e = getenv("PS_SERIAL")) serial_chains = atoi(e) != 0;
    ErrorChecker::check_cuda_error(cudaSetDevice(device), "cudaSetDevice");
    ErrorChecker::check_cuda_error(cudaEventCreateWithFlags(&chain_done, cudaEventDisableTiming), "cudaEventCreate");
    const unsigned cap = prm.max_peaks ? prm.max_peaks : (1u << 20);
    try {
      for (int i = 0; i < nslots; i++) {
        ErrorChecker::check(ps_search_create(&this->prm, &slots[i].search));
        if (!zf.empty()) ErrorChecker::check(ps_search_set_zaplist(slots[i].search, &zf[0], &zw[0], (unsigned)zf.size()));
        ErrorChecker::check_cuda_error(cudaStreamCreateWithFlags(&slots[i].stream, cudaStreamNonBlocking), "cudaStreamCreate");
        Utils::device_malloc(&slots[i].d_trial, ncopy);
        slots[i].peaks.resize(cap);
      }
    } catch (...) {
      release();
      throw;
    }
  }
  ~WorkerSlots() { release(); }
  WorkerSlots(const WorkerSlots&) = delete;
  bool matches(const ps_search_params& o) const {
    return o.size == prm.size && o.nsamps == prm.nsamps && o.tsamp == prm.tsamp && o.nharmonics == prm.nharmonics &&
           o.min_snr == prm.min_snr && o.min_freq == prm.min_freq && o.max_freq == prm.max_freq && o.max_peaks == prm.max_peaks;
  }

 private:
  void release() {
    cudaSetDevice(device);
    for (int i = 0; i < kMaxSlots; i++) {
      if (slots[i].stream) cudaStreamSynchronize(slots[i].stream);
      ps_search_destroy(slots[i].search);
      cudaFree(slots[i].d_trial);
      if (slots[i].stream) cudaStreamDestroy(slots[i].stream);
      slots[i] = Slot();
    }
    if (chain_done) cudaEventDestroy(chain_done);
    chain_done = 0;
  }
};

class Worker {  // pipeline_multi.cu:83-254
  DispersionTrials<unsigned char>& trials;
  DMDispenser& manager;
  SearchArgs& args;
  AccelerationPlan& acc_plan;
  unsigned int size;
  int device;
  const std::vector<float>* fixed_accs;  // when set, replaces acc_plan (benchmarks / parity harness)
  int dm_idx_offset;

  typedef WorkerSlots::Slot Slot;
  WorkerSlots* shared_slots;

 public:
  CandidateCollection dm_trial_cands;
  unsigned long long trials_done = 0;  // DM x acceleration trials searched
  // wall-clock split of start(): queueing device work, waiting for the device, host-side
  // clustering + distillation (diagnostics for bench.py)
  double t_submit = 0, t_wait = 0, t_host = 0;
  unsigned long long raw_peaks = 0;

  Worker(DispersionTrials<unsigned char>& trials, DMDispenser& manager, AccelerationPlan& acc_plan, SearchArgs& args,
         unsigned int size, int device, const std::vector<float>* fixed_accs = 0, int dm_idx_offset = 0,
         WorkerSlots* shared_slots = 0)
      : trials(trials), manager(manager), args(args), acc_plan(acc_plan), size(size), device(device), fixed_accs(fixed_accs),
        dm_idx_offset(dm_idx_offset), shared_slots(shared_slots) {}

  static ps_search_params search_params(SearchArgs& args, unsigned int size, unsigned int nsamps, float tsamp) {
    ps_search_params prm;
    prm.size = size;
    prm.nsamps = nsamps;
    prm.tsamp = tsamp;
    prm.nharmonics = args.nharmonics;
    prm.min_snr = args.min_snr;
    prm.min_freq = args.min_freq;
    prm.max_freq = args.max_freq;
    prm.boundary_5_freq = 0.05f;
    prm.boundary_25_freq = 0.5f;
    prm.max_peaks = args.max_peaks;
    return prm;
  }

  // transform lengths the fused device path takes (fft.cu: two-pass engine)
  static bool fused_size(unsigned int n) { return n >= (1u << 17) && n <= (1u << 25) && !(n & (n - 1)); }

  // Any other length (--fft_size smaller than 2^17 or not a power of two): the loop of
  // pipeline_multi.cu:145-244 statement by statement on the staged entry points -- one kernel per
  // reference call, transforms through the chirp-z path of ps_fft_*.  Slower, same candidates.
  void start_staged() {
    ErrorChecker::check_cuda_error(cudaSetDevice(device), "cudaSetDevice");
    const unsigned nsamps = trials.get_nsamps();
    ReusableDeviceTimeSeries<float, unsigned char> d_tim(size);
    DeviceTimeSeries<float> d_tim_r(size);
    TimeDomainResampler resampler;
    float tobs = size * trials.get_tsamp();
    float bin_width = 1.0 / tobs;
    DeviceFourierSeries<cufftComplex_t> d_fseries(size / 2 + 1, bin_width);
    DevicePowerSpectrum<float> pspec(d_fseries);
    std::vector<float> zf = args.zap_freqs, zw = args.zap_widths;
    if (args.zapfilename != "") Zapper::parse(args.zapfilename, zf, zw);
    Zapper bzap(zf, zw);
    Dereddener rednoise(size / 2 + 1);
    SpectrumFormer former;
    PeakFinder cand_finder(args.min_snr, args.min_freq, args.max_freq, size);
    HarmonicSums<float> sums(pspec, args.nharmonics);
    HarmonicFolder harm_folder(sums);
    HarmonicDistiller harm_finder(args.freq_tol, args.max_harm, false);
    AccelerationDistiller acc_still(tobs, args.freq_tol, true);
    FFTerR2C r2cfft(size);
    FFTerC2R c2rfft(size);
    std::vector<float> acc_list;
    float mean, std, rms;
    DedispersedTimeSeries<unsigned char> tim;
    while (true) {
      int ii = manager.get_dm_trial_idx();
      if (ii == -1) break;
      trials.get_idx(ii, tim);
      d_tim.copy_from_host(tim);
      if (size > nsamps) {  // pad with the mean of the valid samples, :160-163
        float pad = stats::mean<float>(d_tim.get_data(), nsamps);
        d_tim.fill(nsamps, size, pad);
      }
      r2cfft.execute(d_tim.get_data(), d_fseries.get_data());
      former.form(d_fseries, pspec);
      rednoise.calculate_median(pspec);
      rednoise.deredden(d_fseries);
      if (!zf.empty()) bzap.zap(d_fseries);
      former.form_interpolated(d_fseries, pspec);
      stats::stats<float>(pspec.get_data(), size / 2 + 1, &mean, &rms, &std);
      c2rfft.execute(d_fseries.get_data(), d_tim.get_data());
      if (fixed_accs)
        acc_list = *fixed_accs;
      else
        acc_plan.generate_accel_list(tim.get_dm(), acc_list);
      CandidateCollection accel_trial_cands;
      for (size_t jj = 0; jj < acc_list.size(); jj++) {
        resampler.resampleII(d_tim, d_tim_r, size, acc_list[jj]);
        r2cfft.execute(d_tim_r.get_data(), d_fseries.get_data());
        former.form_interpolated(d_fseries, pspec);
        stats::normalise(pspec.get_data(), mean * size, std * size, size / 2 + 1);
        harm_folder.fold(pspec);
        SpectrumCandidates trial_cands(tim.get_dm(), ii + dm_idx_offset, acc_list[jj]);
        cand_finder.find_candidates(pspec, trial_cands);
        cand_finder.find_candidates(sums, trial_cands);
        accel_trial_cands.append(harm_finder.distill(trial_cands.cands));
      }
      dm_trial_cands.append(acc_still.distill(accel_trial_cands.cands));
      trials_done += acc_list.size();
    }
    ErrorChecker::check_cuda_error(cudaDeviceSynchronize(), "Worker::start_staged");
  }

  void start() {
    if (!fused_size(size) || (getenv("PS_STAGED") && atoi(getenv("PS_STAGED")) == 1)) {  // PS_STAGED=1: tests
      start_staged();
      return;
    }
    ErrorChecker::check_cuda_error(cudaSetDevice(device), "cudaSetDevice");
    ps_search_params prm = search_params(args, size, trials.get_nsamps(), trials.get_tsamp());

    std::vector<float> zf = args.zap_freqs, zw = args.zap_widths;
    if (args.zapfilename != "") Zapper::parse(args.zapfilename, zf, zw);

    float tobs = size * trials.get_tsamp();  // pipeline_multi.cu:118
    float bin_width_f = 1.0 / tobs;          // :119
    double bin_width = bin_width_f;
    const unsigned nbins = size / 2 + 1;
    std::vector<PeakRange> ranges;
    for (int nh = 0; nh <= args.nharmonics; nh++) ranges.push_back(peak_range((int)nbins, bin_width, nh, args.min_freq, args.max_freq));
    HarmonicDistiller harm_finder(args.freq_tol, args.max_harm, false);  // :138
    AccelerationDistiller acc_still(tobs, args.freq_tol, true);          // :139
    const size_t ncopy = std::min(trials.get_nsamps(), size);

    WorkerSlots* own = 0;
    struct Owner {
      WorkerSlots*& p;
      ~Owner() { delete p; }
    } owner{own};
    WorkerSlots* ws = shared_slots;
    if (!ws || !ws->matches(prm) || ws->device != device) ws = own = new WorkerSlots(device, prm, zf, zw);
    Slot* slots = ws->slots;

    std::vector<int> idxs, peakidxs;
    std::vector<float> snrs, peaksnrs, peakfreqs;

    typedef std::chrono::steady_clock clk;
    auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };

    auto submit = [&](Slot& s) -> bool {
      clk::time_point t0 = clk::now();
      s.ii = manager.get_dm_trial_idx();
      if (s.ii == -1) return false;
      DedispersedTimeSeries<unsigned char> tim;
      trials.get_idx(s.ii, tim);
      s.dm = tim.get_dm();
      // cudaMemcpyDefault: the trials may live in (pinned or pageable) host memory, as in the
      // reference, or already in device memory (bench.py's HBM-resident measurement)
      ErrorChecker::check_cuda_error(
          cudaMemcpyAsync(s.d_trial, tim.get_data(), ncopy, cudaMemcpyDefault, s.stream), "Error from h2dcpy");
      ps_count_transfer(ncopy, 0);
      if (fixed_accs)
        s.acc_list = *fixed_accs;
      else
        acc_plan.generate_accel_list(s.dm, s.acc_list);
      // the KERNELS of this trial start when the previous trial's are done (see chain_done): the slots
      // overlap the host's share (peak fetch, clustering, distillation, submission) and the copy engine's
      // (the trial's H2D above) with the other slot's kernels, not kernels with kernels
      if (ws->serial_chains && ws->chain_recorded)
        ErrorChecker::check_cuda_error(cudaStreamWaitEvent(s.stream, ws->chain_done, 0), "cudaStreamWaitEvent");
      ErrorChecker::check(ps_search_reset_peaks(s.search, s.stream));
      ErrorChecker::check(ps_search_whiten(s.search, s.d_trial, 0, 0, s.stream));
      ErrorChecker::check(ps_search_accel_batch(s.search, &s.acc_list[0], (int)s.acc_list.size(), 0, s.stream));
      if (ws->serial_chains) {
        ErrorChecker::check_cuda_error(cudaEventRecord(ws->chain_done, s.stream), "cudaEventRecord");
        ws->chain_recorded = true;
      }
      t_submit += secs(t0, clk::now());
      return true;
    };

    auto finish = [&](Slot& s) {
      clk::time_point t0 = clk::now();
      long n = ps_search_fetch_peaks(s.search, &s.peaks[0], s.peaks.size(), s.stream);
      for (int attempt = 0; n == PS_EOVERFLOW && attempt < 4; attempt++) {
        // more above-threshold bins than the list holds (interference): enlarge it and search this DM
        // trial's accelerations again -- its whitened series is still on the device
        const size_t want = (size_t)ps_search_peaks_needed(s.search) * 3 / 2 + 1024;
        if (want > 0x7fffffffu) break;
        ErrorChecker::check(ps_search_reserve_peaks(s.search, (unsigned)want));
        s.peaks.resize(want);
        ErrorChecker::check(ps_search_reset_peaks(s.search, s.stream));
        ErrorChecker::check(ps_search_accel_batch(s.search, &s.acc_list[0], (int)s.acc_list.size(), 0, s.stream));
        n = ps_search_fetch_peaks(s.search, &s.peaks[0], s.peaks.size(), s.stream);
      }
      ErrorChecker::check((int)std::min<long>(n, 0));
      clk::time_point t1 = clk::now();
      t_wait += secs(t0, t1);
      raw_peaks += (unsigned long long)n;
      CandidateCollection accel_trial_cands;
      long pos = 0;
      for (int jj = 0; jj < (int)s.acc_list.size(); jj++) {
        SpectrumCandidates trial_cands(s.dm, s.ii + dm_idx_offset, s.acc_list[jj]);
        for (int nh = 0; nh <= args.nharmonics; nh++) {
          idxs.clear();
          snrs.clear();
          for (; pos < n && s.peaks[pos].acc_idx == jj && s.peaks[pos].nh == nh; pos++) {
            idxs.push_back(s.peaks[pos].bin);
            snrs.push_back(s.peaks[pos].snr);
          }
          if (idxs.empty()) continue;
          peakidxs.resize(idxs.size());
          peaksnrs.resize(idxs.size());
          peakfreqs.resize(idxs.size());
          int npeaks = cluster_peaks(&idxs[0], &snrs[0], (int)idxs.size(), 30, &peakidxs[0], &peaksnrs[0]);
          for (int k = 0; k < npeaks; k++) peakfreqs[k] = peakidxs[k] * ranges[nh].factor;
          trial_cands.append(&peaksnrs[0], &peakfreqs[0], nh, npeaks);
        }
        accel_trial_cands.append(harm_finder.distill(trial_cands.cands));
      }
      dm_trial_cands.append(acc_still.distill(accel_trial_cands.cands));
      trials_done += s.acc_list.size();
      t_host += secs(t1, clk::now());
    };

    // ring of DM trials in flight: keep every slot busy, retire in submission order
    const int K = ws->nslots;
    int head = 0, tail = 0, inflight = 0;
    bool more = true;
    while (more && inflight < K) {
      more = submit(slots[head]);
      if (more) {
        head = (head + 1) % K;
        inflight++;
      }
    }
    while (inflight) {
      finish(slots[tail]);
      tail = (tail + 1) % K;
      inflight--;
      if (more) {
        more = submit(slots[head]);
        if (more) {
          head = (head + 1) % K;
          inflight++;
        }
      }
    }
  }
};

inline void* launch_worker_thread(void* ptr) {  // pipeline_multi.cu:256-259
  reinterpret_cast<Worker*>(ptr)->start();
  return NULL;
}

}  // namespace peasoup_b200
