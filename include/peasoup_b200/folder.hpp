// folder.hpp -- host-side mirror of the reference's candidate folder (include/transforms/folder.hpp,
// include/data_types/folded.hpp): TimeSeriesFolder, FoldedSubints, FoldOptimiser
// and MultiFolder with the reference's names and call order.  Device work goes through the C ABI
// (ps_fold_timeseries, ps_fold_optimise and the staged whitening entry points).
#pragma once
#include <algorithm>
#include <map>
#include "transforms.hpp"

namespace peasoup_b200 {

struct less_than_key {  // folder.hpp:25-31: final ordering by the better of the two S/N values
  inline bool operator()(const Candidate& x, const Candidate& y) {
    return (std::max(x.snr, x.folded_snr) > std::max(y.snr, y.folded_snr));
  }
};

template <class T>
class FoldedSubints {  // folded.hpp:10-74
  T* data_ptr;
  unsigned int nbins, nints;
  double period;
  float accel;
  double opt_period;
  int opt_width, opt_bin;
  float opt_sn, tobs;

 public:
  std::vector<float> opt_fold, opt_prof;
  FoldedSubints(unsigned int nbins, unsigned int nints)
      : data_ptr(0), nbins(nbins), nints(nints), period(0), accel(0), opt_period(0), opt_width(0), opt_bin(0), opt_sn(0), tobs(0) {
    Utils::device_malloc<T>(&data_ptr, nbins * nints);
  }
  ~FoldedSubints() { cudaFree(data_ptr); }
  FoldedSubints(const FoldedSubints&) = delete;
  T* get_data() { return data_ptr; }
  unsigned int get_nbins() { return nbins; }
  unsigned int get_nints() { return nints; }
  double get_period() { return period; }
  float get_accel() { return accel; }
  void set_period(double p) { period = p; }
  void set_accel(float a) { accel = a; }
  void set_opt_period(double p) { opt_period = p; }
  void set_opt_width(int w) { opt_width = w; }
  void set_opt_bin(int b) { opt_bin = b; }
  float get_opt_period() { return opt_period; }  // narrowed to float, folded.hpp:43
  int get_opt_width() { return opt_width; }
  int get_opt_bin() { return opt_bin; }
  void set_opt_sn(float sn) { opt_sn = sn; }
  float get_opt_sn() { return opt_sn; }
  void set_tobs(float t) { tobs = t; }
  float get_tobs() { return tobs; }
};

class TimeSeriesFolder {  // folder.hpp:34-62
  unsigned int size;

 public:
  explicit TimeSeriesFolder(unsigned int size, unsigned int = MAX_BLOCKS, unsigned int = MAX_THREADS) : size(size) {}
  void fold(DeviceTimeSeries<float>& input, FoldedSubints<float>& output, double period) {
    output.set_period(period);
    double tobs = input.get_nsamps() * input.get_tsamp();  // float product widened, folder.hpp:50
    output.set_tobs((float)tobs);
    ErrorChecker::check(ps_fold_timeseries(input.get_data(), output.get_data(), input.get_nsamps(), output.get_nints(), period,
                                           input.get_tsamp(), output.get_nbins(), 0));
  }
};

class FoldOptimiser {  // folder.hpp:65-352
  unsigned int nbins, nints;
  ps_fold_optimiser* handle;

 public:
  ps_fold_result last;
  FoldOptimiser(unsigned int nbins, unsigned int nints, unsigned int = MAX_BLOCKS, unsigned int = MAX_THREADS)
      : nbins(nbins), nints(nints), handle(0) {
    ErrorChecker::check(ps_fold_optimiser_create(nbins, nints, &handle));
  }
  ~FoldOptimiser() { ps_fold_optimiser_destroy(handle); }
  FoldOptimiser(const FoldOptimiser&) = delete;
  void optimise(FoldedSubints<float>& fold) {
    if (nbins != fold.get_nbins() || nints != fold.get_nints())
      ErrorChecker::throw_error("FoldedSubints instance has wrong dimensions");
    fold.opt_fold.resize(nbins * nints);
    fold.opt_prof.resize(nbins);
    ErrorChecker::check(
        ps_fold_optimise(handle, fold.get_data(), fold.get_period(), fold.get_tobs(), &fold.opt_fold[0], &fold.opt_prof[0], &last, 0));
    fold.set_opt_sn(last.opt_sn);
    fold.set_opt_period(last.opt_period);
    fold.set_opt_width(last.opt_width);
    fold.set_opt_bin(last.opt_bin);
  }
};

// PROVENANCE: interface mirror, not implementation.  The constructor, fold_all_mapped() and fold_n()
// below keep the statement order and the local names of the reference's MultiFolder
// (include/transforms/folder.hpp:352-434), and FoldedSubints above those of data_types/folded.hpp:10-74,
// because north_star requires the reference's API and call order; only the progress bar is gone.
// They contain no arithmetic: folding and optimisation are csrc/fold.cu behind the C ABI.
class MultiFolder {  // folder.hpp:354-460
  std::vector<Candidate>& cands;
  DispersionTrials<unsigned char>& dm_trials;
  unsigned int nsamps;
  float tsamp;
  std::map<unsigned int, std::vector<unsigned int> > dm_to_cand_map;
  FoldedSubints<float>* subints;
  FoldOptimiser* optimiser;
  float min_period, max_period;

  void fold_all_mapped() {
    if (dm_to_cand_map.empty()) return;
    ReusableDeviceTimeSeries<float, unsigned char> device_tim(nsamps);
    DeviceTimeSeries<float> d_tim_r(nsamps);
    Dereddener rednoise(nsamps / 2 + 1);
    TimeDomainResampler resampler;
    SpectrumFormer former;
    float tobs = nsamps * tsamp;
    FFTerR2C r2cfft(nsamps);
    FFTerC2R c2rfft(nsamps);
    DeviceFourierSeries<cufftComplex_t> d_fseries(nsamps / 2 + 1, 1.0 / tobs);
    DevicePowerSpectrum<float> pspec(d_fseries);
    const int nbins = subints->get_nbins(), nints = subints->get_nints();
    TimeSeriesFolder folder(nsamps);
    for (std::map<unsigned int, std::vector<unsigned int> >::iterator iter = dm_to_cand_map.begin(); iter != dm_to_cand_map.end();
         ++iter) {
      DedispersedTimeSeries<unsigned char> h_tim = dm_trials[iter->first];
      device_tim.copy_from_host(h_tim);
      d_tim_r.set_tsamp(h_tim.get_tsamp());
      // whitening for the fold: deredden only (no zapping, no normalisation), folder.hpp:406-413
      r2cfft.execute(device_tim.get_data(), d_fseries.get_data());
      former.form(d_fseries, pspec);
      rednoise.calculate_median(pspec);
      rednoise.deredden(d_fseries);
      c2rfft.execute(d_fseries.get_data(), device_tim.get_data());
      for (size_t ii = 0; ii < iter->second.size(); ii++) {
        int cand_idx = iter->second[ii];
        float period = 1.0 / cands[cand_idx].freq;
        resampler.resample(device_tim, d_tim_r, nsamps, cands[cand_idx].acc);
        folder.fold(d_tim_r, *subints, period);
        optimiser->optimise(*subints);
        cands[cand_idx].folded_snr = subints->get_opt_sn();
        cands[cand_idx].set_fold(&subints->opt_fold[0], nbins, nints);
        cands[cand_idx].opt_period = subints->get_opt_period();
      }
    }
  }

 public:
  MultiFolder(std::vector<Candidate>& cands, DispersionTrials<unsigned char>& dm_trials)
      : cands(cands), dm_trials(dm_trials), subints(0), optimiser(0) {
    nsamps = Utils::prev_power_of_two(dm_trials.get_nsamps());
    tsamp = dm_trials.get_tsamp();
    subints = new FoldedSubints<float>(64, 16);
    optimiser = new FoldOptimiser(64, 16);
    min_period = 0.001;
    max_period = 10.00;
  }
  ~MultiFolder() {
    delete subints;
    delete optimiser;
  }
  MultiFolder(const MultiFolder&) = delete;
  void fold_n(unsigned int n_to_fold) {
    int count = std::min(n_to_fold, (unsigned int)cands.size());
    for (int ii = 0; ii < count; ii++) {
      float p = 1.0 / cands[ii].freq;
      if (p > min_period && p < max_period) dm_to_cand_map[cands[ii].dm_idx].push_back(ii);
    }
    fold_all_mapped();
    std::sort(cands.begin(), cands.end(), less_than_key());
  }
};

}  // namespace peasoup_b200
