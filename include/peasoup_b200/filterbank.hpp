// filterbank.hpp -- host-side mirror of the reference's filterbank reader and dedispersion wrapper
// (include/data_types/filterbank.hpp:44-250, include/data_types/header.hpp:171-403,
// include/transforms/dedisperser.hpp:12-114): the stage immediately before the search hot path
// (SURVEY 8f row f1).  Same class and method names; dedisp_* calls become ps_dedisp_* / ps_dedisperse.
#pragma once
#include <cstdint>
#include <cstring>
#include <iostream>
#include "transforms.hpp"

namespace peasoup_b200 {

// SIGPROC header (header.hpp:171-222); only the fields the pipeline reads are kept by name
struct SigprocHeader {
  std::string source_name, rawdatafile;
  double az_start = 0, za_start = 0, src_raj = 0, src_dej = 0, tstart = 0, tsamp = 0, period = 0, fch1 = 0, foff = 0, refdm = 0;
  int nchans = 0, telescope_id = 0, machine_id = 0, data_type = 0, ibeam = 0, nbeams = 0, nbits = 0, barycentric = 0,
      pulsarcentric = 0, nbins = 0, nsamples = 0, nifs = 0, npuls = 0;
  unsigned char signed_data = 0;
  size_t size = 0;  // header length in bytes
};

inline void read_header(std::ifstream& stream, SigprocHeader& h) {  // header.hpp:338-403
  auto rstr = [&](std::string& s) {
    int32_t n = 0;
    stream.read((char*)&n, 4);
    if (!stream.good() || n <= 0 || n > 80) ErrorChecker::throw_error("read_header: not a SIGPROC header");
    s.resize(n);
    stream.read(&s[0], n);
  };
  std::string s;
  rstr(s);
  if (s != "HEADER_START") ErrorChecker::throw_error("read_header: HEADER_START not found");
  while (true) {
    rstr(s);
    if (s == "HEADER_END") break;
    auto rd_i = [&](int& v) { stream.read((char*)&v, 4); };
    auto rd_d = [&](double& v) { stream.read((char*)&v, 8); };
    if (s == "source_name") rstr(h.source_name);
    else if (s == "rawdatafile") rstr(h.rawdatafile);
    else if (s == "az_start") rd_d(h.az_start);
    else if (s == "za_start") rd_d(h.za_start);
    else if (s == "src_raj") rd_d(h.src_raj);
    else if (s == "src_dej") rd_d(h.src_dej);
    else if (s == "tstart") rd_d(h.tstart);
    else if (s == "tsamp") rd_d(h.tsamp);
    else if (s == "period") rd_d(h.period);
    else if (s == "fch1") rd_d(h.fch1);
    else if (s == "foff") rd_d(h.foff);
    else if (s == "refdm") rd_d(h.refdm);
    else if (s == "nchans") rd_i(h.nchans);
    else if (s == "telescope_id") rd_i(h.telescope_id);
    else if (s == "machine_id") rd_i(h.machine_id);
    else if (s == "data_type") rd_i(h.data_type);
    else if (s == "ibeam") rd_i(h.ibeam);
    else if (s == "nbeams") rd_i(h.nbeams);
    else if (s == "nbits") rd_i(h.nbits);
    else if (s == "barycentric") rd_i(h.barycentric);
    else if (s == "pulsarcentric") rd_i(h.pulsarcentric);
    else if (s == "nbins") rd_i(h.nbins);
    else if (s == "nsamples") rd_i(h.nsamples);
    else if (s == "nifs") rd_i(h.nifs);
    else if (s == "npuls") rd_i(h.npuls);
    else if (s == "signed") stream.read((char*)&h.signed_data, 1);
    else std::cerr << "Warning: read_header: unknown parameter " << s << std::endl;
    if (!stream.good()) ErrorChecker::throw_error("read_header: truncated header");
  }
  h.size = (size_t)stream.tellg();
  if (h.nsamples == 0) {  // from the file size, header.hpp:394-401
    stream.seekg(0, std::ios::end);
    size_t total = (size_t)stream.tellg();
    h.nsamples = (int)((total - h.size) / h.nchans * 8 / h.nbits);
    stream.seekg(h.size, std::ios::beg);
  }
}

class Filterbank {  // filterbank.hpp:44-197
 protected:
  unsigned char* data;
  unsigned int nsamps, nchans;
  unsigned char nbits;
  float fch1, foff, tsamp;
  Filterbank() : data(0), nsamps(0), nchans(0), nbits(0), fch1(0.f), foff(0.f), tsamp(0.f) {}

 public:
  Filterbank(unsigned char* data, unsigned int nsamps, unsigned int nchans, unsigned char nbits, float fch1, float foff, float tsamp)
      : data(data), nsamps(nsamps), nchans(nchans), nbits(nbits), fch1(fch1), foff(foff), tsamp(tsamp) {}
  virtual ~Filterbank() {}
  float get_tsamp() { return tsamp; }
  float get_foff() { return foff; }
  float get_fch1() { return fch1; }
  unsigned int get_nchans() { return nchans; }
  unsigned int get_nsamps() { return nsamps; }
  unsigned char get_nbits() { return nbits; }
  unsigned char* get_data() { return data; }
  float get_cfreq() { return foff < 0 ? fch1 + foff * nchans / 2 : fch1 - foff * nchans / 2; }  // filterbank.hpp:190-196
};

class SigprocFilterbank : public Filterbank {  // filterbank.hpp:207-250: whole file into host RAM
  std::vector<unsigned char> storage;

 public:
  SigprocHeader header;
  explicit SigprocFilterbank(std::string filename) {
    std::ifstream infile(filename.c_str(), std::ifstream::in | std::ifstream::binary);
    if (!infile.good()) ErrorChecker::throw_error("File " + filename + " could not be opened: ");
    read_header(infile, header);
    size_t input_size = (size_t)header.nsamples * header.nbits * header.nchans / 8;
    storage.resize(input_size);
    infile.seekg(header.size, std::ios::beg);
    infile.read(reinterpret_cast<char*>(storage.data()), input_size);
    data = storage.data();
    nsamps = header.nsamples;
    nchans = header.nchans;
    tsamp = header.tsamp;  // double narrowed to float, filterbank.hpp:234
    nbits = header.nbits;
    fch1 = header.fch1;
    foff = header.foff;
  }
};

class Dedisperser {  // dedisperser.hpp:12-114
  Filterbank& filterbank;
  std::vector<float> dm_list;
  std::vector<int> killmask;
  std::vector<unsigned char> trials_storage;
  int device = -1;  // -1: the current device

 public:
  Dedisperser(Filterbank& filterbank, unsigned int /*num_gpus*/ = 1) : filterbank(filterbank) {
    killmask.assign(filterbank.get_nchans(), 1);
  }
  // the reference's plan takes device indices 0..num_gpus-1 (dedisp_create_plan_multi); here the caller
  // names the device the trials are dedispersed on
  void set_device(int d) { device = d; }
  void set_dm_list(std::vector<float> v) { dm_list = v; }
  void set_dm_list(float* p, unsigned int n) { dm_list.assign(p, p + n); }
  std::vector<float> get_dm_list() { return dm_list; }
  void generate_dm_list(float dm_start, float dm_end, float width, float tolerance) {
    int n = ps_dedisp_generate_dm_list(dm_start, dm_end, filterbank.get_tsamp(), width, filterbank.get_fch1(), filterbank.get_foff(),
                                       filterbank.get_nchans(), tolerance, 0, 0);
    ErrorChecker::check(n);
    dm_list.resize(n);
    ErrorChecker::check(ps_dedisp_generate_dm_list(dm_start, dm_end, filterbank.get_tsamp(), width, filterbank.get_fch1(),
                                                   filterbank.get_foff(), filterbank.get_nchans(), tolerance, &dm_list[0], n));
  }
  void set_killmask(std::vector<int> k) { killmask.swap(k); }
  void set_killmask(std::string filename) {  // one 0/1 per line, dedisperser.hpp:71-95
    std::ifstream infile(filename.c_str());
    if (!infile.good()) ErrorChecker::throw_error("File " + filename + " could not be opened: ");
    std::vector<int> k;
    std::string str;
    while (k.size() < filterbank.get_nchans() && std::getline(infile, str)) k.push_back(std::atoi(str.c_str()));
    if (k.size() != filterbank.get_nchans()) {
      std::cerr << "WARNING: killmask is not the same size as nchans" << std::endl;
      killmask.assign(filterbank.get_nchans(), 1);
    } else {
      killmask.swap(k);
    }
  }
  // 8-bit trials [ndm][nsamps - max_delay] in host memory owned by this object (dedisperser.hpp:98-113)
  DispersionTrials<unsigned char> dedisperse() {
    if (dm_list.empty()) ErrorChecker::throw_error("Dedisperser: empty DM list");
    long max_delay = ps_dedisp_max_delay(&dm_list[0], (int)dm_list.size(), filterbank.get_nchans(), filterbank.get_tsamp(),
                                         filterbank.get_fch1(), filterbank.get_foff());
    if (max_delay >= (long)filterbank.get_nsamps()) ErrorChecker::throw_error("Dedisperser: max delay exceeds the data length");
    unsigned int out_nsamps = filterbank.get_nsamps() - (unsigned int)max_delay;
    trials_storage.resize((size_t)out_nsamps * dm_list.size());
    ErrorChecker::check(ps_dedisperse_on(device, 0, filterbank.get_data(), filterbank.get_nsamps(), filterbank.get_nchans(), filterbank.get_nbits(),
                                      filterbank.get_tsamp(), filterbank.get_fch1(), filterbank.get_foff(), &dm_list[0],
                                      (int)dm_list.size(), &killmask[0], trials_storage.data(), out_nsamps));
    return DispersionTrials<unsigned char>(trials_storage.data(), out_nsamps, filterbank.get_tsamp(), dm_list);
  }
};

}  // namespace peasoup_b200
