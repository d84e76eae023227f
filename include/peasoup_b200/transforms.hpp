// transforms.hpp -- host-side mirror of the reference's search transforms
// (include/transforms/{resampler,ffter,spectrumformer,dereddener,birdiezapper,harmonicfolder,
// peakfinder,distiller}.hpp, include/utils/{stats,utils}.hpp).  Same class names, method
// names, argument meaning and error behaviour (std::runtime_error); each method forwards to
// the C ABI of libpeasoup_b200.so instead of kernels.h / cuFFT.  MAX_BLOCKS / MAX_THREADS
// constructor arguments are accepted and ignored (launch shapes are chosen per kernel).
#pragma once
#include <cmath>
#include <fstream>
#include <iterator>
#include "data_types.hpp"

#define MAX_BLOCKS 65535  // kernels/defaults.h:9-10
#define MAX_THREADS 512

namespace peasoup_b200 {

typedef ps_complex cufftComplex_t;  // same layout as cufftComplex

// ---- resampler.hpp:7-34 ----
class TimeDomainResampler {
 public:
  TimeDomainResampler(unsigned int = MAX_THREADS, unsigned int = MAX_BLOCKS) {}
  // tsamp is taken from the INPUT series (resampler.hpp:22,29)
  void resample(DeviceTimeSeries<float>& input, DeviceTimeSeries<float>& output, unsigned int size, float acc) {
    ErrorChecker::check(ps_resample(input.get_data(), output.get_data(), size, acc, input.get_tsamp(), 0));
    sync("Error from device_resample");
  }
  void resampleII(DeviceTimeSeries<float>& input, DeviceTimeSeries<float>& output, unsigned int size, float acc) {
    ErrorChecker::check(ps_resampleII(input.get_data(), output.get_data(), size, acc, input.get_tsamp(), 0));
    sync("Error from device_resampleII");
  }

 private:
  static void sync(const char* where) { ErrorChecker::check_cuda_error(cudaDeviceSynchronize(), where); }
};

// ---- ffter.hpp:7-78 (cuFFT replaced by the hand-written plan) ----
class FFTer {
 protected:
  ps_fft_plan* fft_plan;
  unsigned int size;
  explicit FFTer(unsigned int size) : fft_plan(0), size(size) { ErrorChecker::check(ps_fft_plan_create(size, &fft_plan)); }
  ~FFTer() { ps_fft_plan_destroy(fft_plan); }
  FFTer(const FFTer&) = delete;

 public:
  double get_resolution(float tsamp) { return (double)1.0 / (size * tsamp); }
  unsigned int get_output_size() { return size / 2 + 1; }
};

class FFTerR2C : public FFTer {
 public:
  explicit FFTerR2C(unsigned int size, unsigned int batch = 1) : FFTer(size) {
    if (batch != 1) ErrorChecker::throw_error("FFTerR2C: only batch=1 is used by the search path");
  }
  void execute(float* tim, cufftComplex_t* fseries, cudaStream_t st = 0) {
    ErrorChecker::check(ps_fft_r2c(fft_plan, tim, fseries, st));
  }
};

class FFTerC2R : public FFTer {
 public:
  explicit FFTerC2R(unsigned int size, unsigned int batch = 1) : FFTer(size) {
    if (batch != 1) ErrorChecker::throw_error("FFTerC2R: only batch=1 is used by the search path");
  }
  void execute(cufftComplex_t* input, float* output, cudaStream_t st = 0) {
    ErrorChecker::check(ps_fft_c2r(fft_plan, input, output, st));
  }
};
typedef FFTerR2C CuFFTerR2C;  // reference spelling
typedef FFTerC2R CuFFTerC2R;

// ---- spectrumformer.hpp:6-24 ----
class SpectrumFormer {
 public:
  void form_interpolated(DeviceFourierSeries<cufftComplex_t>& input, DevicePowerSpectrum<float>& output) {
    ErrorChecker::check(ps_form_power_series(input.get_data(), output.get_data(), input.get_nbins(), 1, 0));
  }
  void form(DeviceFourierSeries<cufftComplex_t>& input, DevicePowerSpectrum<float>& output) {
    ErrorChecker::check(ps_form_power_series(input.get_data(), output.get_data(), input.get_nbins(), 0, 0));
  }
};

// ---- dereddener.hpp:9-68 ----
class Dereddener {
  unsigned int size;
  float* median;
  float* work;

 public:
  explicit Dereddener(unsigned int size) : size(size), median(0), work(0) {
    Utils::device_malloc(&median, size);
    Utils::device_malloc(&work, size / 5 + size / 25 + size / 125 + 8);
  }
  ~Dereddener() {
    cudaFree(median);
    cudaFree(work);
  }
  Dereddener(const Dereddener&) = delete;
  void calculate_median(DevicePowerSpectrum<float>& powers, float boundary_5_freq = 0.05, float boundary_25_freq = 0.5) {
    if (powers.get_nbins() != size) ErrorChecker::throw_error("Bad data length given to running_median()");
    ErrorChecker::check(ps_running_median(powers.get_data(), size, powers.get_bin_width(), boundary_5_freq, boundary_25_freq,
                                          median, work, 0));
  }
  void deredden(DeviceFourierSeries<cufftComplex_t>& spectrum) {
    ErrorChecker::check(ps_divide_c_by_f(spectrum.get_data(), median, spectrum.get_nbins(), 0));
  }
  float* get_median() { return median; }
};

// ---- birdiezapper.hpp:11-73 ----
class Zapper {
  std::vector<float> birdies, widths;
  float* d_birdies;
  float* d_widths;

 public:
  explicit Zapper(std::string zaplist) : d_birdies(0), d_widths(0) { append_from_file(zaplist); }
  Zapper(const std::vector<float>& f, const std::vector<float>& w) : birdies(f), widths(w), d_birdies(0), d_widths(0) { upload(); }
  ~Zapper() {
    cudaFree(d_birdies);
    cudaFree(d_widths);
  }
  Zapper(const Zapper&) = delete;
  static void parse(const std::string& path, std::vector<float>& f, std::vector<float>& w) {
    std::ifstream infile(path.c_str());
    if (!infile.good()) ErrorChecker::throw_error("File " + path + " could not be opened: ");
    std::string line;
    while (std::getline(infile, line)) {  // whitespace-separated "freq width"; blank lines skipped
      std::stringstream ss(line);
      std::vector<std::string> tok((std::istream_iterator<std::string>(ss)), std::istream_iterator<std::string>());
      if (tok.size() > 1) {
        f.push_back((float)::atof(tok[0].c_str()));
        w.push_back((float)::atof(tok[1].c_str()));
      }
    }
  }
  void append_from_file(std::string zaplist) {
    parse(zaplist, birdies, widths);
    upload();
  }
  const std::vector<float>& get_birdies() const { return birdies; }
  const std::vector<float>& get_widths() const { return widths; }
  void zap(DeviceFourierSeries<cufftComplex_t>& fseries) {
    float bin_width = fseries.get_bin_width();  // narrowed to float: birdiezapper.hpp:62
    zap(fseries.get_data(), bin_width, fseries.get_nbins());
  }
  void zap(cufftComplex_t* fseries, float bin_width, unsigned int nbins) {
    ErrorChecker::check(ps_zap_birdies(fseries, d_birdies, d_widths, bin_width, (unsigned)birdies.size(), nbins, 0));
  }

 private:
  void upload() {
    cudaFree(d_birdies);
    cudaFree(d_widths);
    d_birdies = d_widths = 0;
    if (birdies.empty()) return;
    Utils::device_malloc(&d_birdies, birdies.size());
    Utils::device_malloc(&d_widths, widths.size());
    Utils::h2dcpy(d_birdies, &birdies[0], birdies.size());
    Utils::h2dcpy(d_widths, &widths[0], widths.size());
  }
};

// ---- harmonicfolder.hpp:9-45 ----
class HarmonicFolder {
  HarmonicSums<float>& sums;

 public:
  HarmonicFolder(HarmonicSums<float>& sums, unsigned int = MAX_BLOCKS, unsigned int = MAX_THREADS) : sums(sums) {}
  void fold(DevicePowerSpectrum<float>& fold0) {
    float* ptrs[5] = {0, 0, 0, 0, 0};
    if (sums.size() > 5) ErrorChecker::throw_error("HarmonicFolder: at most 5 folds");
    for (size_t i = 0; i < sums.size(); i++) ptrs[i] = sums[(int)i]->get_data();
    ErrorChecker::check(ps_harmonic_sum(fold0.get_data(), ptrs, fold0.get_nbins(), (unsigned)sums.size(), 0));
  }
};

// peakfinder.hpp:27-56: greedy clustering of ascending (idx, snr); a cluster is extended while the
// next index is closer than min_gap to the index of the cluster's current maximum.
inline int cluster_peaks(const int* idxs, const float* snrs, int count, int min_gap, int* peakidxs, float* peaksnrs) {
  int npeaks = 0, ii = 0;
  while (ii < count) {
    float best = snrs[ii];
    int best_idx = idxs[ii], anchor = idxs[ii];
    for (ii++; ii < count && idxs[ii] - anchor < min_gap; ii++)
      if (snrs[ii] > best) {
        best = snrs[ii];
        best_idx = anchor = idxs[ii];
      }
    peakidxs[npeaks] = best_idx;
    peaksnrs[npeaks] = best;
    npeaks++;
  }
  return npeaks;
}

// peakfinder.hpp:77-92 bin range and bin->frequency factor, evaluated with the reference's types
struct PeakRange {
  int start_idx, limit;
  float factor;
};
inline PeakRange peak_range(int size /*nbins*/, double bin_width, int nh, float min_freq, float max_freq) {
  float nyquist = bin_width * size;
  int orig_size = 2.0 * (size - 1.0);
  int max_bin = (int)((max_freq / bin_width) * pow(2.0, nh));
  int start_idx = (int)(orig_size * (min_freq / nyquist) * pow(2.0, nh));
  PeakRange r;
  r.start_idx = start_idx;
  r.limit = std::min(size, max_bin);
  r.factor = 1.0 / size * nyquist / pow(2.0, (float)nh);
  return r;
}

// ---- peakfinder.hpp:11-95 ----
class PeakFinder {
  float threshold, min_freq, max_freq;
  int min_gap;
  unsigned int max_cands;
  std::vector<int> idxs, peakidxs;
  std::vector<float> snrs, peaksnrs, peakfreqs;
  ps_peak_workspace* ws;

  void reserve(unsigned int n) {
    if (ws) ps_peak_workspace_destroy(ws);
    ws = 0;
    max_cands = n;
    idxs.resize(max_cands);
    snrs.resize(max_cands);
    peakidxs.resize(max_cands);
    peaksnrs.resize(max_cands);
    peakfreqs.resize(max_cands);
    ErrorChecker::check(ps_peak_workspace_create(max_cands, &ws));
  }

 public:
  PeakFinder(float threshold, float min_freq, float max_freq, unsigned int /*size*/, int min_gap = 30, unsigned int max_cands_ = 100000)
      : threshold(threshold), min_freq(min_freq), max_freq(max_freq), min_gap(min_gap), max_cands(max_cands_), ws(0) {
    reserve(max_cands_);
  }
  ~PeakFinder() { ps_peak_workspace_destroy(ws); }
  PeakFinder(const PeakFinder&) = delete;

  void find_candidates(HarmonicSums<float>& sums, SpectrumCandidates& cands) {
    for (size_t i = 0; i < sums.size(); i++) find_candidates(*sums[(int)i], cands);
  }
  void find_candidates(DevicePowerSpectrum<float>& pspec, SpectrumCandidates& cands) {
    int nh = pspec.get_nh();
    PeakRange r = peak_range((int)pspec.get_nbins(), pspec.get_bin_width(), nh, min_freq, max_freq);
    // the reference overruns its 100000-entry host arrays silently; here the workspace grows to 1.5x
    // the count found and the spectrum is searched again (as the fused path does with its peak list)
    int count = ps_find_peaks(r.limit, r.start_idx, pspec.get_data(), threshold, &idxs[0], &snrs[0], max_cands, ws, 0);
    if (count == PS_EOVERFLOW) {
      reserve((unsigned int)(ps_peak_workspace_needed(ws) + ps_peak_workspace_needed(ws) / 2 + 16));
      count = ps_find_peaks(r.limit, r.start_idx, pspec.get_data(), threshold, &idxs[0], &snrs[0], max_cands, ws, 0);
    }
    ErrorChecker::check(count);
    int npeaks = cluster_peaks(&idxs[0], &snrs[0], count, min_gap, &peakidxs[0], &peaksnrs[0]);
    for (int i = 0; i < npeaks; i++) peakfreqs[i] = peakidxs[i] * r.factor;
    cands.append(&peaksnrs[0], &peakfreqs[0], nh, npeaks);
  }
};

// ---- distiller.hpp ----
#define PS_SPEED_OF_LIGHT 299792458.0

class BaseDistiller {
 protected:
  std::vector<bool> unique;
  int size;
  bool keep_related;
  virtual void condition(std::vector<Candidate>& cands, int idx) = 0;
  explicit BaseDistiller(bool keep_related) : size(0), keep_related(keep_related) {}
  virtual ~BaseDistiller() {}
  void knock_out(std::vector<Candidate>& cands, int idx, int ii) {
    if (keep_related) cands[idx].append(cands[ii]);
    unique[ii] = false;
  }

 public:
  // distiller.hpp:26-59: sort by descending S/N (std::sort, as the reference), then every still-unique
  // candidate in that order knocks out the later ones its condition matches.
  std::vector<Candidate> distill(std::vector<Candidate>& cands) {
    size = (int)cands.size();
    unique.assign(size, true);
    std::sort(cands.begin(), cands.end(), [](const Candidate& x, const Candidate& y) { return x.snr > y.snr; });
    int count = 0;
    for (int idx = 0; idx < size; idx++)
      if (unique[idx]) {
        count++;
        condition(cands, idx);
      }
    std::vector<Candidate> out;
    out.reserve(count);
    for (int i = 0; i < size; i++)
      if (unique[i]) out.push_back(cands[i]);
    return out;
  }
};

class HarmonicDistiller : public BaseDistiller {  // distiller.hpp:63-108
  float tolerance;
  float max_harm;
  bool fractional_harms;
  void condition(std::vector<Candidate>& cands, int idx) override {
    double upper_tol = 1 + tolerance;
    double lower_tol = 1 - tolerance;
    double fundi_freq = cands[idx].freq;
    for (int ii = idx + 1; ii < size; ii++) {
      double freq = cands[ii].freq;
      float max_denominator = fractional_harms ? (float)pow(2.0, cands[ii].nh) : 1.f;
      for (int jj = 1; jj <= max_harm; jj++)
        for (int kk = 1; kk <= max_denominator; kk++) {
          double ratio = kk * freq / (jj * fundi_freq);
          if (ratio > lower_tol && ratio < upper_tol) knock_out(cands, idx, ii);  // no break: may append twice
        }
    }
  }

 public:
  HarmonicDistiller(float tol, float max_harm, bool keep_related, bool fractional_harms = true)
      : BaseDistiller(keep_related), tolerance(tol), max_harm(max_harm), fractional_harms(fractional_harms) {}
};

class AccelerationDistiller : public BaseDistiller {  // distiller.hpp:115-164
  float tobs;
  double tobs_over_c;
  float tolerance;
  float correct_for_acceleration(double freq, double delta_acc) { return freq + delta_acc * freq * tobs_over_c; }
  void condition(std::vector<Candidate>& cands, int idx) override {
    double fundi_freq = cands[idx].freq;
    double fundi_acc = cands[idx].acc;
    double edge = fundi_freq * tolerance;
    for (int ii = idx + 1; ii < size; ii++) {
      double delta_acc = fundi_acc - cands[ii].acc;
      double acc_freq = correct_for_acceleration(fundi_freq, delta_acc);
      double f = cands[ii].freq;
      bool hit = acc_freq > fundi_freq ? (f > fundi_freq - edge && f < acc_freq + edge)
                                       : (f < fundi_freq + edge && f > acc_freq - edge);
      if (hit) knock_out(cands, idx, ii);
    }
  }

 public:
  AccelerationDistiller(float tobs, float tolerance, bool keep_related)
      : BaseDistiller(keep_related), tobs(tobs), tolerance(tolerance) {
    tobs_over_c = tobs / PS_SPEED_OF_LIGHT;
  }
};

class DMDistiller : public BaseDistiller {  // distiller.hpp:168-197
  float tolerance;
  void condition(std::vector<Candidate>& cands, int idx) override {
    double fundi_freq = cands[idx].freq;
    double upper_tol = 1 + tolerance;
    double lower_tol = 1 - tolerance;
    for (int ii = idx + 1; ii < size; ii++) {
      double ratio = cands[ii].freq / fundi_freq;
      if (ratio > lower_tol && ratio < upper_tol) knock_out(cands, idx, ii);
    }
  }

 public:
  DMDistiller(float tolerance, bool keep_related) : BaseDistiller(keep_related), tolerance(tolerance) {}
};

// ---- stats.hpp ----
namespace stats {
template <class T>
float rms(T* ptr, unsigned int nsamps, unsigned int first_samp = 0) {
  float r;
  ErrorChecker::check(ps_rms(ptr, nsamps, first_samp, &r));
  return r;
}
template <class T>
float mean(T* ptr, unsigned int nsamps, unsigned int first_samp = 0) {
  float m;
  ErrorChecker::check(ps_mean(ptr, nsamps, first_samp, &m));
  return m;
}
inline float std(float mean, float rms) { return sqrt(rms * rms - mean * mean); }
template <class T>
void stats(T* ptr, unsigned int nsamps, float* mean_, float* rms_, float* std_, unsigned int first_samp = 0) {
  *rms_ = rms<T>(ptr, nsamps, first_samp);
  *mean_ = mean<T>(ptr, nsamps, first_samp);
  *std_ = std(*mean_, *rms_);
}
inline void normalise(float* ptr, float mean, float std, unsigned int size, unsigned int = MAX_BLOCKS, unsigned int = MAX_THREADS) {
  ErrorChecker::check(ps_normalise(ptr, mean, std, size, 0));
}
}  // namespace stats

// ---- utils.hpp:140-193 ----
class AccelerationPlan {
  float acc_lo, acc_hi, tol, pulse_width;
  unsigned int nsamps;
  float tsamp, cfreq, bw, tobs;

 public:
  // pulse_width keeps its command-line value: the reference divides the constructor PARAMETER by
  // 1e3 after the member was initialised (utils.hpp:158-165), which has no effect.
  AccelerationPlan(float acc_lo, float acc_hi, float tol, float pulse_width, unsigned int nsamps, float tsamp, float cfreq, float bw)
      : acc_lo(acc_lo), acc_hi(acc_hi), tol(tol), pulse_width(pulse_width), nsamps(nsamps), tsamp(tsamp), cfreq(cfreq), bw(fabs(bw)) {
    tobs = nsamps * tsamp;
  }
  void generate_accel_list(float dm, std::vector<float>& acc_list) {
    acc_list.clear();
    if (acc_hi == acc_lo) {
      acc_list.push_back(0.0);
      return;
    }
    float tdm = pow(8.3 * bw / pow(cfreq, 3.0) * dm, 2.0);
    float tpulse = pulse_width * pulse_width;
    float ttsamp = tsamp * tsamp;
    float w_us = sqrt(tdm + tpulse + ttsamp);
    float alt_a = 2.0 * w_us * 1.0e-6 * 24.0 * 299792458.0 / tobs / tobs * sqrt((tol * tol) - 1.0);
    if (acc_hi != 0 && acc_lo != 0) acc_list.push_back(0.0);
    for (float acc = acc_lo; acc < acc_hi; acc += alt_a) acc_list.push_back(acc);
    acc_list.push_back(acc_hi);
  }
};

}  // namespace peasoup_b200
