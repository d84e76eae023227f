// data_types.hpp -- host-side mirror of the reference's search-side containers
// (include/data_types/{timeseries,fourierseries,candidates}.hpp), same class and method
// names so that pipeline code written against the reference compiles against this header.
// Device buffers are owned RAII-style exactly like the reference (ctor cudaMalloc, dtor
// cudaFree: fourierseries.hpp:123-132, timeseries.hpp:239-243,294-297).
// PROVENANCE: interface mirror.  Class, method and member names are the reference's by requirement
// ("keep the API"); the containers hold no arithmetic.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdio>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include "../peasoup_b200.h"

namespace peasoup_b200 {

// ---- error handling: utils/exceptions.hpp:13-153 (runtime_error with the same prefixes) ----
class ErrorChecker {
 public:
  static void throw_error(const std::string& msg) { throw std::runtime_error(msg.c_str()); }
  // C-ABI status -> exception
  static void check(int rc) {
    if (rc < 0) throw std::runtime_error(ps_last_error());
  }
  static void check_cuda_error(cudaError_t e, const std::string& msg = "Unspecified location") {
    if (e != cudaSuccess) {
      std::stringstream s;
      s << "CUDA failed with error: " << cudaGetErrorString(e) << std::endl << "Additional: " << msg << std::endl;
      throw std::runtime_error(s.str());
    }
  }
};

// ---- utils/utils.hpp:10-96 ----
class Utils {
 public:
  static unsigned int prev_power_of_two(unsigned int val) {  // strict: utils.hpp:12-18
    unsigned int n = 1;
    while (n * 2 < val) n *= 2;
    return n;
  }
  template <class T>
  static void device_malloc(T** ptr, size_t units) {
    ErrorChecker::check_cuda_error(cudaMalloc((void**)ptr, sizeof(T) * units), "Error from device_malloc");
  }
  template <class T>
  static void host_malloc(T** ptr, size_t units) {
    ErrorChecker::check_cuda_error(cudaMallocHost((void**)ptr, sizeof(T) * units), "Error from host_malloc");
  }
  template <class T>
  static void device_free(T* ptr) {
    ErrorChecker::check_cuda_error(cudaFree(ptr), "Error from device_free");
  }
  template <class T>
  static void host_free(T* ptr) {
    ErrorChecker::check_cuda_error(cudaFreeHost((void*)ptr), "Error from host_free.");
  }
  template <class T>
  static void h2dcpy(T* d, const T* h, size_t units) {
    ErrorChecker::check_cuda_error(cudaMemcpy(d, h, sizeof(T) * units, cudaMemcpyHostToDevice), "Error from h2dcpy");
  }
  template <class T>
  static void d2hcpy(T* h, const T* d, size_t units) {
    ErrorChecker::check_cuda_error(cudaMemcpy(h, d, sizeof(T) * units, cudaMemcpyDeviceToHost), "Error from d2hcpy");
  }
  template <class T>
  static void d2dcpy(T* dst, const T* src, size_t units) {
    ErrorChecker::check_cuda_error(cudaMemcpy(dst, src, sizeof(T) * units, cudaMemcpyDeviceToDevice), "Error from d2dcpy");
  }
  static int gpu_count() {
    int c = 0;
    cudaGetDeviceCount(&c);
    return c;
  }
};

// ---- timeseries.hpp ----
template <class T>
class TimeSeries {
 protected:
  T* data_ptr;
  unsigned int nsamps;
  float tsamp;  // float everywhere on the search side (timeseries.hpp:54)

 public:
  TimeSeries(T* data_ptr, unsigned int nsamps, float tsamp) : data_ptr(data_ptr), nsamps(nsamps), tsamp(tsamp) {}
  TimeSeries() : data_ptr(0), nsamps(0), tsamp(0.f) {}
  explicit TimeSeries(unsigned int nsamps) : data_ptr(0), nsamps(nsamps), tsamp(0.f) {}  // tsamp = 0: timeseries.hpp:77-78
  T operator[](int n) { return data_ptr[n]; }
  T* get_data() { return data_ptr; }
  void set_data(T* p) { data_ptr = p; }
  unsigned int get_nsamps() { return nsamps; }
  void set_nsamps(unsigned int n) { nsamps = n; }
  float get_tsamp() { return tsamp; }
  void set_tsamp(float t) { tsamp = t; }
};

template <class T>
class DedispersedTimeSeries : public TimeSeries<T> {
  float dm;

 public:
  DedispersedTimeSeries() : TimeSeries<T>(), dm(0.f) {}
  DedispersedTimeSeries(T* p, unsigned int nsamps, float tsamp, float dm) : TimeSeries<T>(p, nsamps, tsamp), dm(dm) {}
  float get_dm() { return dm; }
  void set_dm(float d) { dm = d; }
};

template <class OnDeviceType>
class DeviceTimeSeries : public TimeSeries<OnDeviceType> {
 public:
  explicit DeviceTimeSeries(unsigned int nsamps) : TimeSeries<OnDeviceType>(nsamps) {
    Utils::device_malloc<OnDeviceType>(&this->data_ptr, nsamps);
  }
  void fill(size_t start, size_t end, OnDeviceType value) {
    if (end > this->nsamps) ErrorChecker::throw_error("DeviceTimeSeries::fill bad end value requested");
    ErrorChecker::check(ps_fill(this->data_ptr + start, this->data_ptr + end, value, 0));
  }
  ~DeviceTimeSeries() { cudaFree(this->data_ptr); }
  DeviceTimeSeries(const DeviceTimeSeries&) = delete;
  DeviceTimeSeries& operator=(const DeviceTimeSeries&) = delete;
};

template <class OnDeviceType, class OnHostType>
class ReusableDeviceTimeSeries : public DeviceTimeSeries<OnDeviceType> {
  OnHostType* copy_buffer;

 public:
  explicit ReusableDeviceTimeSeries(unsigned int nsamps) : DeviceTimeSeries<OnDeviceType>(nsamps) {
    Utils::device_malloc<OnHostType>(&copy_buffer, nsamps);
  }
  // timeseries.hpp:335-344
  void copy_from_host(TimeSeries<OnHostType>& host_tim) {
    size_t size = std::min(host_tim.get_nsamps(), this->nsamps);
    this->tsamp = host_tim.get_tsamp();
    Utils::h2dcpy(copy_buffer, host_tim.get_data(), size);
    ErrorChecker::check(ps_convert_u8_f32(copy_buffer, this->data_ptr, (unsigned)size, 0));
  }
  ~ReusableDeviceTimeSeries() { cudaFree(copy_buffer); }
};

template <class T>
class DispersionTrials {
  T* data_ptr;
  unsigned int nsamps;
  float tsamp;
  std::vector<float> dm_list;

 public:
  DispersionTrials(T* data_ptr, unsigned int nsamps, float tsamp, std::vector<float> dm_list_in)
      : data_ptr(data_ptr), nsamps(nsamps), tsamp(tsamp), dm_list(std::move(dm_list_in)) {}
  unsigned int get_count() { return (unsigned int)dm_list.size(); }
  unsigned int get_nsamps() { return nsamps; }
  float get_tsamp() { return tsamp; }
  T* get_data() { return data_ptr; }
  std::vector<float>& get_dm_list() { return dm_list; }
  DedispersedTimeSeries<T> operator[](int idx) {
    return DedispersedTimeSeries<T>(data_ptr + (size_t)idx * (size_t)nsamps, nsamps, tsamp, dm_list[idx]);
  }
  void get_idx(unsigned int idx, DedispersedTimeSeries<T>& tim) {  // timeseries.hpp:473-479
    tim.set_data(data_ptr + (size_t)idx * (size_t)nsamps);
    tim.set_dm(dm_list[idx]);
    tim.set_nsamps(nsamps);
    tim.set_tsamp(tsamp);
  }
};

// ---- fourierseries.hpp ----
template <class T>
class FrequencySeries {
 protected:
  T* data_ptr;
  unsigned int nbins;
  double bin_width;
  FrequencySeries(unsigned int nbins, double bin_width) : data_ptr(0), nbins(nbins), bin_width(bin_width) {}

 public:
  T* get_data() { return data_ptr; }
  void set_data(T* p) { data_ptr = p; }
  double get_bin_width() { return bin_width; }
  void set_bin_width(double b) { bin_width = b; }
  unsigned int get_nbins() { return nbins; }
  void set_nbins(unsigned int n) { nbins = n; }
};

template <class T>
class DeviceFrequencySeries : public FrequencySeries<T> {
 protected:
  DeviceFrequencySeries(unsigned int nbins, double bin_width) : FrequencySeries<T>(nbins, bin_width) {
    Utils::device_malloc<T>(&this->data_ptr, nbins);
  }
  ~DeviceFrequencySeries() { cudaFree(this->data_ptr); }
  DeviceFrequencySeries(const DeviceFrequencySeries&) = delete;
};

template <class T>  // T = ps_complex (layout of cufftComplex)
class DeviceFourierSeries : public DeviceFrequencySeries<T> {
 public:
  DeviceFourierSeries(unsigned int nbins, double bin_width) : DeviceFrequencySeries<T>(nbins, bin_width) {}
};

template <class T>
class DevicePowerSpectrum : public DeviceFrequencySeries<T> {
  unsigned int nh;

 public:
  DevicePowerSpectrum(unsigned int nbins, double bin_width, unsigned int nh = 0)
      : DeviceFrequencySeries<T>(nbins, bin_width), nh(nh) {}
  template <class U>
  DevicePowerSpectrum(FrequencySeries<U>& other, unsigned int nh = 0)
      : DeviceFrequencySeries<T>(other.get_nbins(), other.get_bin_width()), nh(nh) {}
  unsigned int get_nh() { return nh; }
  void set_nh(unsigned int n) { nh = n; }
};

template <class T>
class HarmonicSums {  // fourierseries.hpp:162-188: fold k (nh = k+1) sums 2^(k+1) harmonics
  std::vector<DevicePowerSpectrum<T>*> folds;

 public:
  HarmonicSums(DevicePowerSpectrum<T>& fold0, unsigned int nfolds) {
    folds.reserve(nfolds);
    for (unsigned int i = 0; i < nfolds; i++) folds.push_back(new DevicePowerSpectrum<T>(fold0, i + 1));
  }
  size_t size() { return folds.size(); }
  DevicePowerSpectrum<T>* operator[](int i) { return folds[i]; }
  ~HarmonicSums() {
    for (size_t i = 0; i < folds.size(); i++) delete folds[i];
  }
  HarmonicSums(const HarmonicSums&) = delete;
};

// ---- candidates.hpp ----
struct CandidatePOD {  // 24-byte record of candidates.peasoup (candidates.hpp:10-17)
  float dm;
  int dm_idx;
  float acc;
  int nh;
  float snr;
  float freq;
};

struct Candidate {
  float dm;
  int dm_idx;
  float acc;
  int nh;
  float snr;
  float freq;
  float folded_snr;
  double opt_period;
  bool is_adjacent;
  bool is_physical;
  float ddm_count_ratio;
  float ddm_snr_ratio;
  std::vector<Candidate> assoc;
  std::vector<float> fold;
  int nbins;
  int nints;

  Candidate(float dm = 0.f, int dm_idx = 0, float acc = 0.f, int nh = 0, float snr = 0.f, float freq = 0.f)
      : dm(dm), dm_idx(dm_idx), acc(acc), nh(nh), snr(snr), freq(freq), folded_snr(0.f), opt_period(0.0),
        is_adjacent(false), is_physical(false), ddm_count_ratio(0.f), ddm_snr_ratio(0.f), nbins(0), nints(0) {}

  void append(Candidate& other) { assoc.push_back(other); }
  void set_fold(float* ar, int nbins_, int nints_) {  // candidates.hpp:69-76
    nbins = nbins_;
    nints = nints_;
    fold.assign(ar, ar + (size_t)nbins_ * nints_);
  }
  int count_assoc() {
    int count = 0;
    for (size_t i = 0; i < assoc.size(); i++) count += 1 + assoc[i].count_assoc();
    return count;
  }
  void collect_candidates(std::vector<CandidatePOD>& out) {  // depth-first, self first (candidates.hpp:83-89)
    CandidatePOD p = {dm, dm_idx, acc, nh, snr, freq};
    out.push_back(p);
    for (size_t i = 0; i < assoc.size(); i++) assoc[i].collect_candidates(out);
  }
};

class CandidateCollection {
 public:
  std::vector<Candidate> cands;
  void append(CandidateCollection& other) { cands.insert(cands.end(), other.cands.begin(), other.cands.end()); }
  void append(std::vector<Candidate> other) { cands.insert(cands.end(), other.begin(), other.end()); }
  void reset() { cands.clear(); }
};

class SpectrumCandidates : public CandidateCollection {
 public:
  float dm;
  int dm_idx;
  float acc;
  SpectrumCandidates(float dm, int dm_idx, float acc) : dm(dm), dm_idx(dm_idx), acc(acc) {}
  void append(float* snrs, float* freqs, int nh, int size) {
    cands.reserve(size + cands.size());
    for (int i = 0; i < size; i++) cands.push_back(Candidate(dm, dm_idx, acc, nh, snrs[i], freqs[i]));
  }
  using CandidateCollection::append;
};

}  // namespace peasoup_b200
