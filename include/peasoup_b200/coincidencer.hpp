// coincidencer.hpp -- host-side mirror of include/transforms/coincidencer.hpp:9-85 and of the
// program around it (src/coincidencer.cpp:122-214): whiten every beam's DM-0
// time series, mark the samples and Fourier bins that exceed the threshold in at least
// `beam_threshold` beams, write a sample mask and a birdie list (a zaplist for the search).
#pragma once
#include "filterbank.hpp"

namespace peasoup_b200 {

struct Birdie {
  float freq, width;
  Birdie(float freq, float width) : freq(freq), width(width) {}
};

class Coincidencer {
  float** device_ptr_array;
  int narrays;

 public:
  explicit Coincidencer(int narrays, unsigned int = MAX_BLOCKS, unsigned int = MAX_THREADS) : device_ptr_array(0), narrays(narrays) {
    Utils::device_malloc<float*>(&device_ptr_array, narrays);
  }
  ~Coincidencer() { cudaFree(device_ptr_array); }
  Coincidencer(const Coincidencer&) = delete;

  void match(float** host_ptr_array, float* mask, size_t size, float threshold, int beam_threshold) {
    Utils::h2dcpy(device_ptr_array, host_ptr_array, narrays);
    ErrorChecker::check(ps_coincidencer(device_ptr_array, mask, narrays, size, threshold, beam_threshold, 0));
  }

  void write_samp_mask(float* mask, size_t size, std::string filename) {  // "#0 1" then one 0/1 per line
    std::vector<float> h_mask(size);
    Utils::d2hcpy(&h_mask[0], mask, size);
    FILE* fo = fopen(filename.c_str(), "w");
    if (!fo) ErrorChecker::throw_error("File " + filename + " could not be opened: ");
    fprintf(fo, "#0 1\n");
    for (size_t ii = 0; ii < size; ii++) fprintf(fo, "%d\n", (int)h_mask[ii]);
    fclose(fo);
  }

  // every run of masked bins becomes one birdie: centre (last - count/2) * bin_width, width count * bin_width
  // (coincidencer.hpp:54-77; the run scan is bounded here, the reference's can read past the end)
  static std::vector<Birdie> birdies_from_mask(const std::vector<float>& h_mask, float bin_width) {
    std::vector<Birdie> birdies;
    size_t ii = 0, size = h_mask.size();
    while (ii < size) {
      if (h_mask[ii] == 0) {
        int count = 0;
        while (ii < size && h_mask[ii] == 0) {
          count++;
          ii++;
        }
        birdies.push_back(Birdie((((int)ii - 1) - (count / 2.0)) * bin_width, count * bin_width));
      } else {
        ii++;
      }
    }
    return birdies;
  }

  void write_birdie_list(float* mask, size_t size, float bin_width, std::string filename) {
    std::vector<float> h_mask(size);
    Utils::d2hcpy(&h_mask[0], mask, size);
    std::vector<Birdie> birdies = birdies_from_mask(h_mask, bin_width);
    FILE* fo = fopen(filename.c_str(), "w");
    if (!fo) ErrorChecker::throw_error("File " + filename + " could not be opened: ");
    for (size_t ii = 0; ii < birdies.size(); ii++) fprintf(fo, "%.9f\t%.6f\n", birdies[ii].freq, birdies[ii].width);
    fclose(fo);
  }
};

struct CoincidenceOptions {  // CmdLineOptions of coincidencer.cpp:30-44 (the fields the program uses)
  std::string samp_outfilename = "rfi.eb_mask", spec_outfilename = "birdies.txt";
  float boundary_5_freq = 0.05f, boundary_25_freq = 0.5f, threshold = 4.0f;
  int beam_threshold = 4;
};

// main() of src/coincidencer.cpp:122-214.  Masks are also handed back when the pointers are set.
// PROVENANCE: interface mirror, not implementation.  Statement order and local names
// (tim_host_ptr_array, ccder, pspecs, ...) follow the reference's main() line for line -- it is the
// calling sequence of the transform classes that tests/test_boundary_compile.py compiles verbatim;
// the work is csrc/coincidence.cu, csrc/bluestein.cu and the staged entry points behind the C ABI.
inline void coincidence_match_files(const std::vector<std::string>& files, const CoincidenceOptions& args, std::vector<float>* samp_mask_out,
                                    std::vector<float>* spec_mask_out) {
  int nfiles = (int)files.size();
  if (nfiles < 1) ErrorChecker::throw_error("coincidencer: no input files");
  std::vector<std::vector<unsigned char> > storage(nfiles);
  std::vector<DedispersedTimeSeries<unsigned char> > tims;
  for (int ii = 0; ii < nfiles; ii++) {
    SigprocFilterbank filterbank(files[ii]);
    Dedisperser dedisperser(filterbank, 1);
    dedisperser.generate_dm_list(0.0, 0.0, 0.4, 1.1);
    DispersionTrials<unsigned char> trial = dedisperser.dedisperse();
    // the Dedisperser owns the trial bytes: keep a copy that outlives it
    storage[ii].assign(trial.get_data(), trial.get_data() + trial.get_nsamps());
    tims.push_back(DedispersedTimeSeries<unsigned char>(storage[ii].data(), trial.get_nsamps(), trial.get_tsamp(), 0.f));
  }
  size_t size = tims[0].get_nsamps();
  for (int ii = 0; ii < nfiles; ii++)
    if (tims[ii].get_nsamps() != size) ErrorChecker::throw_error("Not all filterbanks the same length");
  float tsamp = tims[0].get_tsamp();
  FFTerR2C r2cfft(size);
  FFTerC2R c2rfft(size);
  float tobs = size * tsamp;
  float bin_width = 1.0 / tobs;
  std::vector<ReusableDeviceTimeSeries<float, unsigned char>*> d_tims;
  std::vector<DevicePowerSpectrum<float>*> pspecs;
  DeviceFourierSeries<cufftComplex_t> d_fseries(size / 2 + 1, bin_width);
  for (int ii = 0; ii < nfiles; ii++) {
    d_tims.push_back(new ReusableDeviceTimeSeries<float, unsigned char>(size));
    pspecs.push_back(new DevicePowerSpectrum<float>(d_fseries));
  }
  Dereddener rednoise(size / 2 + 1);
  SpectrumFormer former;
  float mean, std, rms;
  for (int ii = 0; ii < nfiles; ii++) {  // baseline every beam: coincidencer.cpp:160-177
    d_tims[ii]->copy_from_host(tims[ii]);
    r2cfft.execute(d_tims[ii]->get_data(), d_fseries.get_data());
    former.form(d_fseries, *pspecs[ii]);
    rednoise.calculate_median(*pspecs[ii]);  // -l / -a are parsed but never applied (coincidencer.cpp:170)
    rednoise.deredden(d_fseries);
    former.form_interpolated(d_fseries, *pspecs[ii]);
    stats::stats<float>(pspecs[ii]->get_data(), size / 2 + 1, &mean, &rms, &std);
    stats::normalise(pspecs[ii]->get_data(), mean, std, size / 2 + 1);
    c2rfft.execute(d_fseries.get_data(), d_tims[ii]->get_data());
    stats::stats<float>(d_tims[ii]->get_data(), size, &mean, &rms, &std);
    stats::normalise(d_tims[ii]->get_data(), mean, std, size);
  }
  std::vector<float*> tim_host_ptr_array, spec_host_ptr_array;
  for (int ii = 0; ii < nfiles; ii++) {
    tim_host_ptr_array.push_back(d_tims[ii]->get_data());
    spec_host_ptr_array.push_back(pspecs[ii]->get_data());
  }
  DevicePowerSpectrum<float> spec_mask(d_fseries);
  DeviceTimeSeries<float> samp_mask(size);
  Coincidencer ccder(nfiles);
  ccder.match(&tim_host_ptr_array[0], samp_mask.get_data(), size, args.threshold, args.beam_threshold);
  ccder.match(&spec_host_ptr_array[0], spec_mask.get_data(), size / 2 + 1, args.threshold, args.beam_threshold);
  if (!args.samp_outfilename.empty()) ccder.write_samp_mask(samp_mask.get_data(), size, args.samp_outfilename);
  if (!args.spec_outfilename.empty())
    ccder.write_birdie_list(spec_mask.get_data(), size / 2 + 1, spec_mask.get_bin_width(), args.spec_outfilename);
  if (samp_mask_out) {
    samp_mask_out->resize(size);
    Utils::d2hcpy(&(*samp_mask_out)[0], samp_mask.get_data(), size);
  }
  if (spec_mask_out) {
    spec_mask_out->resize(size / 2 + 1);
    Utils::d2hcpy(&(*spec_mask_out)[0], spec_mask.get_data(), size / 2 + 1);
  }
  for (int ii = 0; ii < nfiles; ii++) {
    delete d_tims[ii];
    delete pspecs[ii];
  }
}

}  // namespace peasoup_b200
