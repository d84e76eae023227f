// postprocess.hpp -- host-side mirror of what the reference's main() does AFTER the workers join
// (src/pipeline_multi.cu:353-416): DM and harmonic distillation of the merged
// candidate list, candidate scoring (include/transforms/scorer.hpp), truncation to --limit and the
// output files (include/utils/output_stats.hpp: candidates.peasoup binary + overview.xml).
// Folding (--npdmp, row f3) sits between scoring and truncation: folder.hpp.
#pragma once
#include <sys/stat.h>
#include <cstdio>
#include <ctime>
#include <iomanip>
#include <iostream>
#include <map>
#include "filterbank.hpp"
#include "folder.hpp"
#include "pipeline.hpp"

namespace peasoup_b200 {

class CandidateScorer {  // scorer.hpp:8-87
  float tsamp, cfreq, foff;
  float tdm_chan_partial, tdm_band_partial;

 public:
  CandidateScorer(float tsamp, float cfreq, float foff, float bw) : tsamp(tsamp), cfreq(cfreq), foff(foff) {
    float ftop = cfreq + bw / 2.0;
    float fbottom = cfreq - bw / 2.0;
    tdm_chan_partial = 8300.0 * foff / std::pow(cfreq, 3.0);
    tdm_band_partial = 4150.0 * (1.0 / std::pow(fbottom, 2) - 1.0 / std::pow(ftop, 2));
  }
  void score(Candidate& cand) {
    cand.is_physical = 1.0 / cand.freq > cand.dm * tdm_chan_partial;
    // adjacent: some associate lies in a neighbouring DM trial, or all lie in the candidate's own
    bool adjacent = false, unique = true;
    for (size_t i = 0; i < cand.assoc.size(); i++) {
      if (cand.assoc[i].dm_idx != cand.dm_idx) unique = false;
      if (cand.assoc[i].dm_idx == cand.dm_idx + 1 || cand.assoc[i].dm_idx == cand.dm_idx - 1) {
        adjacent = true;
        break;
      }
    }
    cand.is_adjacent = adjacent || unique;
    int inside_count = 1, total_count = 1;
    float inside_snr = cand.snr, total_snr = cand.snr;
    float ddm = 1.0 / (cand.freq * tdm_band_partial);
    for (size_t i = 0; i < cand.assoc.size(); i++) {
      total_count++;
      total_snr += cand.assoc[i].snr;
      if (fabs(cand.dm - cand.assoc[i].dm) <= ddm) {
        inside_count++;
        inside_snr += cand.assoc[i].snr;
      }
    }
    cand.ddm_count_ratio = (float)inside_count / total_count;
    cand.ddm_snr_ratio = (float)inside_snr / total_snr;
  }
  void score_all(std::vector<Candidate>& cands) {
    for (size_t i = 0; i < cands.size(); i++) score(cands[i]);
  }
};

// pipeline_multi.cu:353-391: distill, score, fold the top npdmp (re-sorting the list), truncate
inline void finalise_candidates(std::vector<Candidate>& cands, float freq_tol, float max_harm, CandidateScorer& scorer, int limit,
                                DispersionTrials<unsigned char>* trials = 0, int npdmp = 0) {
  DMDistiller dm_still(freq_tol, true);
  HarmonicDistiller harm_still(freq_tol, max_harm, true, false);
  cands = dm_still.distill(cands);
  cands = harm_still.distill(cands);
  scorer.score_all(cands);
  if (trials && npdmp > 0) {
    MultiFolder folder(cands, *trials);
    folder.fold_n(npdmp);
  }
  if (limit >= 0 && (size_t)limit < cands.size()) cands.resize(limit);
}

// candidates.peasoup (output_stats.hpp:238-268): per candidate an optional "FOLD" + nbins + nints +
// nbins*nints floats block, then int ndets + ndets 24-byte PODs (the candidate itself first, then
// its associates depth-first)
class CandidateFileWriter {
 public:
  std::map<int, long> byte_mapping;
  std::string output_dir;
  explicit CandidateFileWriter(std::string output_directory) : output_dir(output_directory) {
    struct stat st;
    if (stat(output_dir.c_str(), &st) == -1) mkdir(output_dir.c_str(), 0777);
  }
  void write_binary(std::vector<Candidate>& candidates, std::string filename) {
    std::string path = output_dir + "/" + filename;
    FILE* fo = fopen(path.c_str(), "w");
    if (!fo) ErrorChecker::throw_error("File " + path + " could not be opened: ");
    for (size_t i = 0; i < candidates.size(); i++) {
      byte_mapping[(int)i] = ftell(fo);
      if (!candidates[i].fold.empty()) {
        fputs("FOLD", fo);
        fwrite(&candidates[i].nbins, sizeof(int), 1, fo);
        fwrite(&candidates[i].nints, sizeof(int), 1, fo);
        fwrite(&candidates[i].fold[0], sizeof(float), (size_t)candidates[i].nbins * candidates[i].nints, fo);
      }
      std::vector<CandidatePOD> dets;
      candidates[i].collect_candidates(dets);
      int ndets = (int)dets.size();
      fwrite(&ndets, sizeof(int), 1, fo);
      fwrite(&dets[0], sizeof(CandidatePOD), ndets, fo);
    }
    fclose(fo);
  }
};

// overview.xml with the element names of output_stats.hpp:17-222 that downstream parsers read
// (search_parameters, dedispersion_trials, acceleration_trials, candidates, execution_times)
class OutputFileWriter {
  std::stringstream x;
  template <class T>
  void el(const char* name, const T& v, int indent = 4) {
    x << std::string(indent, ' ') << "<" << name << ">" << std::setprecision(15) << v << "</" << name << ">\n";
  }

 public:
  OutputFileWriter() { x << "<?xml version='1.0' encoding='ISO-8859-1'?>\n<peasoup_search>\n"; }
  void add_header(SigprocHeader& h) {
    x << "  <header_parameters>\n";
    el("source_name", h.source_name);
    el("tstart", h.tstart);
    el("tsamp", h.tsamp);
    el("fch1", h.fch1);
    el("foff", h.foff);
    el("nchans", h.nchans);
    el("nbits", h.nbits);
    el("nsamples", h.nsamples);
    x << "  </header_parameters>\n";
  }
  void add_search_parameters(const std::map<std::string, std::string>& p) {
    x << "  <search_parameters>\n";
    for (std::map<std::string, std::string>::const_iterator it = p.begin(); it != p.end(); ++it) el(it->first.c_str(), it->second);
    x << "  </search_parameters>\n";
  }
  void add_dm_list(std::vector<float>& dms) {
    x << "  <dedispersion_trials count='" << dms.size() << "'>\n";
    for (size_t i = 0; i < dms.size(); i++) x << "    <trial id='" << i << "'>" << std::setprecision(15) << dms[i] << "</trial>\n";
    x << "  </dedispersion_trials>\n";
  }
  void add_acc_list(std::vector<float>& accs) {
    x << "  <acceleration_trials DM='0' count='" << accs.size() << "'>\n";
    for (size_t i = 0; i < accs.size(); i++) x << "    <trial id='" << i << "'>" << std::setprecision(15) << accs[i] << "</trial>\n";
    x << "  </acceleration_trials>\n";
  }
  void add_candidates(std::vector<Candidate>& cands, std::map<int, long>& byte_map) {
    x << "  <candidates>\n";
    for (size_t i = 0; i < cands.size(); i++) {
      Candidate& c = cands[i];
      x << "    <candidate id='" << i << "'>\n";
      el("period", 1.0 / c.freq, 6);
      el("opt_period", c.opt_period, 6);
      el("dm", c.dm, 6);
      el("acc", c.acc, 6);
      el("nh", c.nh, 6);
      el("snr", c.snr, 6);
      el("folded_snr", c.folded_snr, 6);
      el("is_adjacent", c.is_adjacent, 6);
      el("is_physical", c.is_physical, 6);
      el("ddm_count_ratio", c.ddm_count_ratio, 6);
      el("ddm_snr_ratio", c.ddm_snr_ratio, 6);
      el("nassoc", c.count_assoc(), 6);
      el("byte_offset", byte_map[(int)i], 6);
      x << "    </candidate>\n";
    }
    x << "  </candidates>\n";
  }
  void add_timing_info(std::map<std::string, double>& t) {
    x << "  <execution_times>\n";
    for (std::map<std::string, double>::iterator it = t.begin(); it != t.end(); ++it) el(it->first.c_str(), it->second);
    x << "  </execution_times>\n";
  }
  void to_file(std::string path) {
    std::ofstream f(path.c_str());
    if (!f.good()) ErrorChecker::throw_error("File " + path + " could not be opened: ");
    f << x.str() << "</peasoup_search>\n";
  }
};

}  // namespace peasoup_b200
