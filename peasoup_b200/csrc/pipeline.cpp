// pipeline.cpp -- C-ABI entry points over the C++ host mirror (include/peasoup_b200/*.hpp):
// the Worker fan-out of src/pipeline_multi.cu:342-359 and test hooks for the host-side logic.
#include <exception>
#include <string>
#include <thread>
#include <chrono>
#include "../../include/peasoup_b200/postprocess.hpp"
#include "../../include/peasoup_b200/coincidencer.hpp"

namespace ps {
void set_error(const std::string& msg);
}
using namespace peasoup_b200;

static double g_prof[4] = {0, 0, 0, 0};  // submit, wait, host seconds; raw peaks (last call, summed over workers)

static long flatten(std::vector<Candidate>& cands, ps_candidate* out, long max_out) {
  long n = 0;
  for (size_t i = 0; i < cands.size(); i++) {
    std::vector<CandidatePOD> flat;
    cands[i].collect_candidates(flat);
    for (size_t k = 0; k < flat.size(); k++, n++) {
      if (n >= max_out) continue;
      out[n].dm = flat[k].dm;
      out[n].dm_idx = flat[k].dm_idx;
      out[n].acc = flat[k].acc;
      out[n].nh = flat[k].nh;
      out[n].snr = flat[k].snr;
      out[n].freq = flat[k].freq;
      out[n].is_top = k == 0;
    }
  }
  return n;
}

extern "C" {

struct ps_pipeline {
  ps_pipeline_params prm;
  SearchArgs args;
  unsigned size;
  int device0, nworkers;
  std::vector<WorkerSlots*> slots;
  ~ps_pipeline() {
    for (size_t i = 0; i < slots.size(); i++) delete slots[i];
  }
};

int ps_pipeline_open(const ps_pipeline_params* p, const float* h_zap_freqs, const float* h_zap_widths, int nzap, int device0,
                     int nworkers, ps_pipeline** out) {
  ps_pipeline* h = 0;
  try {
    if (!p || !out) ErrorChecker::throw_error("ps_pipeline_open: null argument");
    if (nworkers < 1) nworkers = 1;
    h = new ps_pipeline();
    h->prm = *p;
    h->device0 = device0;
    h->nworkers = nworkers;
    SearchArgs& args = h->args;
    args.size = p->search.size;
    args.nharmonics = p->search.nharmonics;
    args.min_snr = p->search.min_snr;
    args.min_freq = p->search.min_freq;
    args.max_freq = p->search.max_freq;
    args.freq_tol = p->freq_tol;
    args.max_harm = p->max_harm;
    args.max_peaks = p->search.max_peaks;
    if (nzap > 0) {
      args.zap_freqs.assign(h_zap_freqs, h_zap_freqs + nzap);
      args.zap_widths.assign(h_zap_widths, h_zap_widths + nzap);
    }
    h->size = args.size ? args.size : Utils::prev_power_of_two(p->search.nsamps);
    ps_search_params sp = Worker::search_params(args, h->size, p->search.nsamps, p->search.tsamp);
    // (other transform lengths run Worker::start_staged, which owns its device objects)
    for (int i = 0; i < nworkers; i++)
      h->slots.push_back(Worker::fused_size(h->size) ? new WorkerSlots(device0 + i, sp, args.zap_freqs, args.zap_widths) : nullptr);
    *out = h;
    return PS_OK;
  } catch (std::exception& e) {
    delete h;
    ps::set_error(e.what());
    return PS_ECUDA;
  }
}

void ps_pipeline_close(ps_pipeline* h) { delete h; }

long ps_pipeline_run(ps_pipeline* h, const unsigned char* h_trials, const float* h_dm_list, int ndm, int dm_idx_offset,
                     const float* h_acc_explicit, int n_acc_explicit, ps_candidate* out, long max_out,
                     unsigned long long* trials_done) {
  try {
    if (!h || (!h_trials && ndm > 0)) ErrorChecker::throw_error("ps_pipeline_run: null argument");
    const ps_pipeline_params* p = &h->prm;
    const int nworkers = h->nworkers;
    std::vector<float> dms(h_dm_list, h_dm_list + ndm);
    DispersionTrials<unsigned char> trials(const_cast<unsigned char*>(h_trials), p->search.nsamps, p->search.tsamp, dms);
    AccelerationPlan acc_plan(p->acc_start, p->acc_end, p->acc_tol, p->acc_pulse_width, h->size, p->search.tsamp, p->cfreq, p->foff);
    std::vector<float> fixed;
    if (n_acc_explicit > 0) fixed.assign(h_acc_explicit, h_acc_explicit + n_acc_explicit);
    DMDispenser dispenser(trials);
    std::vector<Worker*> workers;
    std::vector<std::thread> threads;
    std::vector<std::exception_ptr> errors(nworkers);
    for (int i = 0; i < nworkers; i++)
      workers.push_back(new Worker(trials, dispenser, acc_plan, h->args, h->size, h->device0 + i, n_acc_explicit > 0 ? &fixed : 0,
                                   dm_idx_offset, h->slots[i]));
    if (nworkers == 1) {
      try {
        workers[0]->start();
      } catch (...) {
        errors[0] = std::current_exception();
      }
    } else {
      for (int i = 0; i < nworkers; i++)
        threads.emplace_back([&, i]() {
          try {
            workers[i]->start();
          } catch (...) {
            errors[i] = std::current_exception();
          }
        });
      for (auto& t : threads) t.join();
    }
    CandidateCollection dm_cands;
    unsigned long long done = 0;
    g_prof[0] = g_prof[1] = g_prof[2] = g_prof[3] = 0;
    for (int i = 0; i < nworkers; i++) {
      dm_cands.append(workers[i]->dm_trial_cands.cands);
      done += workers[i]->trials_done;
      g_prof[0] += workers[i]->t_submit;
      g_prof[1] += workers[i]->t_wait;
      g_prof[2] += workers[i]->t_host;
      g_prof[3] += (double)workers[i]->raw_peaks;
    }
    for (int i = 0; i < nworkers; i++) delete workers[i];
    for (int i = 0; i < nworkers; i++)
      if (errors[i]) std::rethrow_exception(errors[i]);
    // canonical order = ascending dm_idx, emission order within a DM trial (the reference's -t 1 run)
    std::stable_sort(dm_cands.cands.begin(), dm_cands.cands.end(),
                     [](const Candidate& a, const Candidate& b) { return a.dm_idx < b.dm_idx; });
    if (trials_done) *trials_done = done;
    return flatten(dm_cands.cands, out, max_out);
  } catch (std::exception& e) {
    ps::set_error(e.what());
    return PS_ECUDA;
  }
}

long ps_pipeline_search(const ps_pipeline_params* p, const unsigned char* h_trials, const float* h_dm_list, int ndm,
                        int dm_idx_offset, const float* h_acc_explicit, int n_acc_explicit, const float* h_zap_freqs,
                        const float* h_zap_widths, int nzap, int device0, int nworkers, ps_candidate* out, long max_out,
                        unsigned long long* trials_done) {
  if (ndm <= 0) {
    if (trials_done) *trials_done = 0;
    return 0;
  }
  ps_pipeline* h = 0;
  int rc = ps_pipeline_open(p, h_zap_freqs, h_zap_widths, nzap, device0, nworkers, &h);
  if (rc) return rc;
  long n = ps_pipeline_run(h, h_trials, h_dm_list, ndm, dm_idx_offset, h_acc_explicit, n_acc_explicit, out, max_out, trials_done);
  ps_pipeline_close(h);
  return n;
}

// fold results of the last ps_search_filterbank call (ps_fold_results_last)
static std::mutex g_fold_mu;
static std::vector<float> g_fold_snr, g_fold_data;
static std::vector<double> g_fold_period;

long ps_fold_results_last(float* folded_snr, double* opt_period, float* folds, long max_count) {
  std::lock_guard<std::mutex> lk(g_fold_mu);
  long n = (long)g_fold_snr.size();
  for (long i = 0; i < n && i < max_count; i++) {
    if (folded_snr) folded_snr[i] = g_fold_snr[i];
    if (opt_period) opt_period[i] = g_fold_period[i];
    if (folds) memcpy(folds + i * 1024, &g_fold_data[i * 1024], sizeof(float) * 1024);
  }
  return n;
}

// main() of src/pipeline_multi.cu:262-419
long ps_search_filterbank(const char* filename, const ps_filterbank_params* fp, ps_candidate* out, long max_out,
                          float* dm_list_out, int max_dm, int* ndm_out, double* seconds4) {
  typedef std::chrono::steady_clock clk;
  auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
  try {
    if (!filename || !fp) ErrorChecker::throw_error("ps_search_filterbank: null argument");
    clk::time_point t0 = clk::now();
    SigprocFilterbank filobj(filename);
    clk::time_point t1 = clk::now();
    const int nworkers = fp->nworkers < 1 ? 1 : fp->nworkers;
    Dedisperser dedisperser(filobj, nworkers);
    dedisperser.set_device(fp->device0);
    if (fp->killfile && fp->killfile[0]) dedisperser.set_killmask(std::string(fp->killfile));
    dedisperser.generate_dm_list(fp->dm_start, fp->dm_end, fp->dm_pulse_width, fp->dm_tol);
    std::vector<float> dm_list = dedisperser.get_dm_list();
    DispersionTrials<unsigned char> trials = dedisperser.dedisperse();
    clk::time_point t2 = clk::now();
    if (ndm_out) *ndm_out = (int)dm_list.size();
    for (size_t i = 0; dm_list_out && i < dm_list.size() && (int)i < max_dm; i++) dm_list_out[i] = dm_list[i];

    ps_pipeline_params pp = fp->pipe;
    unsigned size = pp.search.size ? pp.search.size : Utils::prev_power_of_two(filobj.get_nsamps());  // :326-331
    pp.search.size = size;
    pp.search.nsamps = trials.get_nsamps();
    pp.search.tsamp = trials.get_tsamp();
    pp.cfreq = filobj.get_cfreq();
    pp.foff = filobj.get_foff();
    std::vector<float> zf, zw;
    if (fp->zapfile && fp->zapfile[0]) Zapper::parse(std::string(fp->zapfile), zf, zw);
    ps_pipeline* h = 0;
    if (ps_pipeline_open(&pp, zf.empty() ? 0 : &zf[0], zw.empty() ? 0 : &zw[0], (int)zf.size(), fp->device0, nworkers, &h))
      ErrorChecker::throw_error(ps_last_error());
    std::vector<ps_candidate> flat(1 << 20);
    long n = ps_pipeline_run(h, trials.get_data(), &dm_list[0], (int)dm_list.size(), 0, 0, 0, flat.data(), (long)flat.size(), 0);
    ps_pipeline_close(h);
    if (n < 0) ErrorChecker::throw_error(ps_last_error());
    if (n > (long)flat.size()) ErrorChecker::throw_error("ps_search_filterbank: too many per-DM candidates");
    clk::time_point t3 = clk::now();
    // rebuild dm_cands (entries + their acceleration associates) from the flattened records
    std::vector<Candidate> dm_cands;
    for (long i = 0; i < n; i++) {
      Candidate c(flat[i].dm, flat[i].dm_idx, flat[i].acc, flat[i].nh, flat[i].snr, flat[i].freq);
      if (flat[i].is_top || dm_cands.empty())
        dm_cands.push_back(c);
      else
        dm_cands.back().append(c);
    }
    CandidateScorer scorer(filobj.get_tsamp(), filobj.get_cfreq(), filobj.get_foff(), fabs(filobj.get_foff()) * filobj.get_nchans());
    finalise_candidates(dm_cands, fp->pipe.freq_tol, (float)fp->pipe.max_harm, scorer, fp->limit, &trials, fp->npdmp);
    clk::time_point t4 = clk::now();
    {
      std::lock_guard<std::mutex> lk(g_fold_mu);
      g_fold_snr.assign(dm_cands.size(), 0.f);
      g_fold_period.assign(dm_cands.size(), 0.0);
      g_fold_data.assign(dm_cands.size() * 1024, 0.f);
      for (size_t i = 0; i < dm_cands.size(); i++) {
        g_fold_snr[i] = dm_cands[i].folded_snr;
        g_fold_period[i] = dm_cands[i].opt_period;
        if (dm_cands[i].fold.size() == 1024) memcpy(&g_fold_data[i * 1024], &dm_cands[i].fold[0], sizeof(float) * 1024);
      }
    }
    if (fp->outdir && fp->outdir[0]) {
      CandidateFileWriter cand_files(fp->outdir);
      cand_files.write_binary(dm_cands, "candidates.peasoup");
      OutputFileWriter stats;
      stats.add_header(filobj.header);
      std::map<std::string, std::string> prm;
      auto put = [&](const char* k, double v) {
        std::stringstream ss;
        ss << std::setprecision(15) << v;
        prm[k] = ss.str();
      };
      prm["infilename"] = filename;
      put("size", size); put("dm_start", fp->dm_start); put("dm_end", fp->dm_end); put("dm_tol", fp->dm_tol);
      put("dm_pulse_width", fp->dm_pulse_width); put("acc_start", pp.acc_start); put("acc_end", pp.acc_end);
      put("acc_tol", pp.acc_tol); put("acc_pulse_width", pp.acc_pulse_width); put("nharmonics", pp.search.nharmonics);
      put("min_snr", pp.search.min_snr); put("min_freq", pp.search.min_freq); put("max_freq", pp.search.max_freq);
      put("max_harm", pp.max_harm); put("freq_tol", pp.freq_tol); put("limit", fp->limit); put("npdmp", fp->npdmp);
      stats.add_search_parameters(prm);
      stats.add_dm_list(dm_list);
      AccelerationPlan acc_plan(pp.acc_start, pp.acc_end, pp.acc_tol, pp.acc_pulse_width, size, pp.search.tsamp, pp.cfreq, pp.foff);
      std::vector<float> acc_list;
      acc_plan.generate_accel_list(0.0, acc_list);
      stats.add_acc_list(acc_list);
      stats.add_candidates(dm_cands, cand_files.byte_mapping);
      std::map<std::string, double> tm;
      tm["reading"] = secs(t0, t1); tm["dedispersion"] = secs(t1, t2); tm["searching"] = secs(t2, t3); tm["total"] = secs(t0, clk::now());
      stats.add_timing_info(tm);
      stats.to_file(std::string(fp->outdir) + "/overview.xml");
    }
    if (seconds4) {
      seconds4[0] = secs(t0, t1); seconds4[1] = secs(t1, t2); seconds4[2] = secs(t2, t3); seconds4[3] = secs(t0, t4);
    }
    return flatten(dm_cands, out, max_out);
  } catch (std::exception& e) {
    ps::set_error(e.what());
    return PS_ECUDA;
  }
}

// pipeline_multi.cu:353-365 on a merged, flattened per-DM list
long ps_distill_merged(const ps_candidate* in, long n, float freq_tol, int max_harm, ps_candidate* out, long max_out) {
  try {
    if (n < 0 || (n > 0 && !in)) ErrorChecker::throw_error("ps_distill_merged: null argument");
    std::vector<Candidate> dm_cands;
    for (long i = 0; i < n; i++) {
      Candidate c(in[i].dm, in[i].dm_idx, in[i].acc, in[i].nh, in[i].snr, in[i].freq);
      if (in[i].is_top || dm_cands.empty())
        dm_cands.push_back(c);
      else
        dm_cands.back().append(c);
    }
    DMDistiller dm_still(freq_tol, true);
    HarmonicDistiller harm_still(freq_tol, (float)max_harm, true, false);
    dm_cands = dm_still.distill(dm_cands);
    dm_cands = harm_still.distill(dm_cands);
    return flatten(dm_cands, out, max_out);
  } catch (std::exception& e) {
    ps::set_error(e.what());
    return PS_ECUDA;
  }
}

// main() of src/coincidencer.cpp
long ps_coincidence_files(const char* const* filenames, int nfiles, const ps_coincidence_params* p, float* h_samp_mask,
                          size_t max_samp, float* h_spec_mask, size_t max_spec) {
  try {
    if (!filenames || nfiles < 1 || !p) ErrorChecker::throw_error("ps_coincidence_files: null argument");
    std::vector<std::string> files(filenames, filenames + nfiles);
    CoincidenceOptions args;
    args.threshold = p->threshold;
    args.beam_threshold = p->beam_threshold;
    args.samp_outfilename = p->samp_outfilename ? p->samp_outfilename : "";
    args.spec_outfilename = p->spec_outfilename ? p->spec_outfilename : "";
    std::vector<float> samp, spec;
    coincidence_match_files(files, args, &samp, &spec);
    if (h_samp_mask) memcpy(h_samp_mask, samp.data(), sizeof(float) * std::min(max_samp, samp.size()));
    if (h_spec_mask) memcpy(h_spec_mask, spec.data(), sizeof(float) * std::min(max_spec, spec.size()));
    return (long)samp.size();
  } catch (std::exception& e) {
    ps::set_error(e.what());
    return PS_ECUDA;
  }
}

void ps_pipeline_profile(double* out4) {
  for (int i = 0; i < 4; i++) out4[i] = g_prof[i];
}

int ps_accel_plan_generate(float acc_lo, float acc_hi, float tol, float pulse_width, unsigned nsamps, float tsamp,
                           float cfreq, float bw, float dm, float* out, int max_out) {
  AccelerationPlan plan(acc_lo, acc_hi, tol, pulse_width, nsamps, tsamp, cfreq, bw);
  std::vector<float> l;
  plan.generate_accel_list(dm, l);
  for (int i = 0; i < (int)l.size() && i < max_out; i++) out[i] = l[i];
  return (int)l.size();
}

int ps_cluster_peaks(const int* idxs, const float* snrs, int count, int min_gap, int* out_idx, float* out_snr) {
  return cluster_peaks(idxs, snrs, count, min_gap, out_idx, out_snr);
}

int ps_peak_range(int nbins, double bin_width, int nh, float min_freq, float max_freq, int* start_idx, int* limit,
                  float* factor) {
  PeakRange r = peak_range(nbins, bin_width, nh, min_freq, max_freq);
  *start_idx = r.start_idx;
  *limit = r.limit;
  *factor = r.factor;
  return PS_OK;
}

long ps_distill(int kind, float tol, float max_harm_or_tobs, int keep_related, int fractional_harms, const ps_candidate* in,
                long n, ps_candidate* out, long max_out) {
  try {
    std::vector<Candidate> c;
    c.reserve(n);
    for (long i = 0; i < n; i++) c.push_back(Candidate(in[i].dm, in[i].dm_idx, in[i].acc, in[i].nh, in[i].snr, in[i].freq));
    std::vector<Candidate> r;
    if (kind == 0) {
      HarmonicDistiller d(tol, max_harm_or_tobs, keep_related != 0, fractional_harms != 0);
      r = d.distill(c);
    } else if (kind == 1) {
      AccelerationDistiller d(max_harm_or_tobs, tol, keep_related != 0);
      r = d.distill(c);
    } else if (kind == 2) {
      DMDistiller d(tol, keep_related != 0);
      r = d.distill(c);
    } else {
      ps::set_error("ps_distill: unknown kind");
      return PS_EINVAL;
    }
    return flatten(r, out, max_out);
  } catch (std::exception& e) {
    ps::set_error(e.what());
    return PS_ECUDA;
  }
}

}  // extern "C"
