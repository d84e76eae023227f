// search.cu -- fused per-DM acceleration search (device side of Worker::start,
// /root/reference/src/pipeline_multi.cu:100-252).
//
// ps_search_whiten      : :157-204  u8 trial -> whitened time series + (mean, std), all on one
//                         stream with no host synchronisation (the reference syncs ~25 times).
// ps_search_accel_batch : :209-234  per acceleration: resampling gather fused into the FFT
//                         pass-1 load, R2C, then ONE pass doing real-FFT split + interbinning +
//                         normalisation, then ONE pass doing the 1/2/4/8/16(/32)-harmonic sums,
//                         thresholding and warp-aggregated compaction of candidate bins.  The
//                         harmonic-sum arrays of the reference are never materialised.
// Host-side clustering / distillation (peakfinder.hpp:27-56, distiller.hpp) consumes the raw
// peak list: include/peasoup_b200/pipeline.hpp.
#include <algorithm>
#include <cmath>
#include <vector>
#include "fft.cuh"

namespace ps {

int plan_create(unsigned n, ps_fft_plan_t** out);
void plan_destroy(ps_fft_plan_t* p);

struct LevelRanges {
  int start[6];
  int limit[6];
};

// Z (packed complex spectrum) -> normalised interbinned amplitude spectrum P[0..M]
// (r2c split, kernels.cu:231-252 interbin, kernels.cu:469-480 normalise with mean*size,std*size
// as in pipeline_multi.cu:224).
__global__ void __launch_bounds__(256) spectrum_kernel(const float2* __restrict__ Z, float* __restrict__ P, unsigned M,
                                                       const float* __restrict__ stats, float fsize) {
  const float mean = __fmul_rn(stats[0], fsize);
  const float sigma = __fmul_rn(stats[2], fsize);
  const unsigned lane = threadIdx.x & 31;
  const unsigned nb = M + 1;
  // whole warps iterate together so the shuffles are convergent
  for (unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) - lane; base < nb; base += gridDim.x * blockDim.x) {
    unsigned k = base + lane;
    float2 x = make_float2(0.f, 0.f);
    if (k < nb) {
      float2 zk = Z[k == M ? 0 : k];
      float2 zm = Z[(k == 0 || k == M) ? 0 : M - k];
      float wr, wi;
      sincospif((float)k / (float)M, &wi, &wr);
      float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);
      float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x);
      x = make_float2(er + (wr * orr + wi * oi), ei + (wr * oi - wi * orr));
    }
    float lx = __shfl_up_sync(0xffffffffu, x.x, 1);
    float ly = __shfl_up_sync(0xffffffffu, x.y, 1);
    if (lane == 0) {
      lx = 0.f;
      ly = 0.f;
      if (k > 0 && k < nb) {
        unsigned kl = k - 1;
        float2 zk = Z[kl];
        float2 zm = Z[kl == 0 ? 0 : M - kl];
        float wr, wi;
        sincospif((float)kl / (float)M, &wi, &wr);
        float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);
        float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x);
        lx = er + (wr * orr + wi * oi);
        ly = ei + (wr * oi - wi * orr);
      }
    }
    if (k < nb) P[k] = normalise_value(interbin_amplitude(x.x, x.y, lx, ly), mean, sigma);
  }
}

// harmonic sums + threshold + compaction (kernels.cu:33-99 + :384-416 + peakfinder.hpp:77-86)
__global__ void __launch_bounds__(256) harmonic_peaks_kernel(const float* __restrict__ P, int nbins, int nfolds,
                                                             LevelRanges r, float thresh, int acc_idx,
                                                             ps_raw_peak* __restrict__ peaks, unsigned* __restrict__ count,
                                                             unsigned cap) {
  const unsigned lane = threadIdx.x & 31;
  for (int base = (blockIdx.x * blockDim.x + threadIdx.x) - (int)lane; base < nbins; base += gridDim.x * blockDim.x) {
    int idx = base + (int)lane;
    float v[6];
    bool in = idx < nbins;
    if (in) {
      v[0] = __ldg(P + idx);
      harmonic_folds(P, idx, nfolds, v + 1);
    }
#pragma unroll
    for (int nh = 0; nh < 6; nh++) {
      if (nh > nfolds) break;
      bool hit = in && idx >= r.start[nh] && idx < r.limit[nh] && v[nh] > thresh;
      unsigned m = __ballot_sync(0xffffffffu, hit);
      if (m) {
        unsigned pos = 0;
        if (lane == 0) pos = atomicAdd(count, (unsigned)__popc(m));
        pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1));
        if (hit && pos < cap) {
          ps_raw_peak pk;
          pk.acc_idx = acc_idx;
          pk.nh = nh;
          pk.bin = idx;
          pk.snr = v[nh];
          peaks[pos] = pk;
        }
      }
    }
  }
}

// mean of the first n u8 samples (GPU_mean on the converted floats, pipeline_multi.cu:161)
__global__ void __launch_bounds__(256) u8_sum_kernel(const unsigned char* __restrict__ b, size_t n, unsigned long long* __restrict__ acc) {
  unsigned long long s = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) s += b[i];
  for (int o = 16; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(acc, s);
}
__global__ void u8_mean_finalize_kernel(unsigned long long* acc, float n, float* out) {
  out[0] = __fdiv_rn((float)acc[0], n);
  acc[0] = 0;
}

struct Search {
  ps_search_params prm;
  ps_fft_plan_t* plan = nullptr;
  unsigned nbins = 0;
  float* d_tim = nullptr;      // size floats: whitened series
  float2* d_X = nullptr;       // nbins complex
  float* d_pspec = nullptr;    // nbins
  float* d_median = nullptr;   // nbins
  float* d_work = nullptr;     // scrunch levels
  float* d_stats = nullptr;    // mean, rms, std, pad mean
  double* d_partials = nullptr;
  unsigned long long* d_u8sum = nullptr;
  float* d_birdies = nullptr;
  float* d_widths = nullptr;
  unsigned nbirdies = 0;
  ps_raw_peak* d_peaks = nullptr;
  unsigned* d_count = nullptr;
  unsigned cap = 0;
  LevelRanges ranges;
  float bin_width_f = 0.f;   // float bin_width = 1.0/tobs   (pipeline_multi.cu:118-119)
  double bin_width = 0.0;    // stored as double in FrequencySeries (fourierseries.hpp:41)
};

static void free_search(Search* s) {
  if (!s) return;
  plan_destroy(s->plan);
  cudaFree(s->d_tim); cudaFree(s->d_X); cudaFree(s->d_pspec); cudaFree(s->d_median); cudaFree(s->d_work);
  cudaFree(s->d_stats); cudaFree(s->d_partials); cudaFree(s->d_u8sum); cudaFree(s->d_birdies); cudaFree(s->d_widths);
  cudaFree(s->d_peaks); cudaFree(s->d_count);
  delete s;
}

// peakfinder.hpp:77-86 evaluated with the reference's operand types
static void level_ranges(const ps_search_params& p, unsigned nbins, double bin_width, LevelRanges* r) {
  for (int nh = 0; nh < 6; nh++) {
    int size = (int)nbins;
    float nyquist = bin_width * size;
    int orig_size = 2.0 * (size - 1.0);
    int max_bin = (int)((p.max_freq / bin_width) * pow(2.0, nh));
    int start_idx = (int)(orig_size * (p.min_freq / nyquist) * pow(2.0, nh));
    r->start[nh] = start_idx < 0 ? 0 : start_idx;
    r->limit[nh] = std::min(size, max_bin);
  }
}

}  // namespace ps

using namespace ps;

struct ps_search {
  Search impl;
};
static Search* SS(ps_search* s) { return reinterpret_cast<Search*>(s); }

extern "C" {

int ps_search_create(const ps_search_params* p, ps_search** out) {
  if (!p || !out) return invalid("ps_search_create: null argument");
  if (p->nharmonics < 0 || p->nharmonics > 5) return invalid("ps_search_create: nharmonics (folds) must be 0..5");
  if (!p->nsamps) return invalid("ps_search_create: nsamps == 0");
  Search* s = new Search();
  s->prm = *p;
  if (s->prm.boundary_5_freq <= 0) s->prm.boundary_5_freq = 0.05f;
  if (s->prm.boundary_25_freq <= 0) s->prm.boundary_25_freq = 0.5f;
  int rc = plan_create(p->size, &s->plan);
  if (rc) {
    delete s;
    return rc;
  }
  unsigned size = p->size;
  s->nbins = size / 2 + 1;
  float tobs = size * p->tsamp;            // pipeline_multi.cu:118
  float bin_width = 1.0 / tobs;            // :119
  s->bin_width_f = bin_width;
  s->bin_width = bin_width;
  s->cap = p->max_peaks ? p->max_peaks : (1u << 20);
  level_ranges(s->prm, s->nbins, s->bin_width, &s->ranges);
  size_t nb = s->nbins;
  cudaError_t e = cudaMalloc(&s->d_tim, sizeof(float) * size);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_X, sizeof(float2) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_pspec, sizeof(float) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_median, sizeof(float) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_work, sizeof(float) * (nb / 5 + nb / 25 + nb / 125 + 8));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_stats, sizeof(float) * 4);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_partials, sizeof(double) * 2 * kMaxStatParts);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_u8sum, sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(s->d_u8sum, 0, sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(s->d_stats, 0, sizeof(float) * 4);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_peaks, sizeof(ps_raw_peak) * (size_t)s->cap);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_count, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMemset(s->d_count, 0, sizeof(unsigned));
  if (e != cudaSuccess) {
    free_search(s);
    return cuda_fail(e, "Error from device_malloc");
  }
  *out = reinterpret_cast<ps_search*>(s);
  return PS_OK;
}

void ps_search_destroy(ps_search* s) { free_search(SS(s)); }

int ps_search_set_zaplist(ps_search* hs, const float* h_freqs, const float* h_widths, unsigned n) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_set_zaplist: null handle");
  cudaFree(s->d_birdies);
  cudaFree(s->d_widths);
  s->d_birdies = s->d_widths = nullptr;
  s->nbirdies = 0;
  if (!n) return PS_OK;
  PS_CUDA(cudaMalloc(&s->d_birdies, sizeof(float) * n), "Error from device_malloc");
  PS_CUDA(cudaMalloc(&s->d_widths, sizeof(float) * n), "Error from device_malloc");
  PS_CUDA(cudaMemcpy(s->d_birdies, h_freqs, sizeof(float) * n, cudaMemcpyHostToDevice), "Error from h2dcpy");
  PS_CUDA(cudaMemcpy(s->d_widths, h_widths, sizeof(float) * n, cudaMemcpyHostToDevice), "Error from h2dcpy");
  s->nbirdies = n;
  return PS_OK;
}

int ps_search_whiten(ps_search* hs, const unsigned char* d_trial, float* h_mean, float* h_std, void* stream) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_whiten: null handle");
  cudaStream_t st = S(stream);
  const ps_fft_plan_t* p = s->plan;
  unsigned size = s->prm.size;
  unsigned valid = std::min(s->prm.nsamps, size);
  int rc;
  if (valid < size) {  // padding path, pipeline_multi.cu:160-163
    u8_sum_kernel<<<ew_grid(valid, 256, 16), 256, 0, st>>>(d_trial, valid, s->d_u8sum);
    PS_LAUNCHED("u8_sum_kernel");
    u8_mean_finalize_kernel<<<1, 1, 0, st>>>(s->d_u8sum, (float)(int)valid, s->d_stats + 3);
    PS_LAUNCHED("u8_mean_finalize_kernel");
  }
  LoadU8Padded ld{d_trial, s->d_stats + 3, valid, p->n2, 0};
  if ((rc = launch_pass1<false>(p, ld, st))) return rc;
  StoreNatural sn{reinterpret_cast<float2*>(p->Z), p->n1, 0};
  if ((rc = launch_pass2<false>(p, sn, st))) return rc;
  if ((rc = launch_r2c_post(p, p->Z, s->d_X, st))) return rc;
  if ((rc = ps_form_power_series((const ps_complex*)s->d_X, s->d_pspec, s->nbins, 0, st))) return rc;
  if ((rc = launch_running_median(s->d_pspec, s->nbins, s->bin_width, s->prm.boundary_5_freq, s->prm.boundary_25_freq,
                                  s->d_median, s->d_work, st)))
    return rc;
  if ((rc = ps_divide_c_by_f((ps_complex*)s->d_X, s->d_median, s->nbins, st))) return rc;
  if (s->nbirdies)
    if ((rc = ps_zap_birdies((ps_complex*)s->d_X, s->d_birdies, s->d_widths, s->bin_width_f, s->nbirdies, s->nbins, st)))
      return rc;
  if ((rc = ps_form_power_series((const ps_complex*)s->d_X, s->d_pspec, s->nbins, 1, st))) return rc;
  if ((rc = launch_stats(s->d_pspec, s->nbins, s->d_partials, s->d_stats, st))) return rc;
  LoadC2RPre lc{s->d_X, p->M, p->n2, 0};
  if ((rc = launch_pass1<true>(p, lc, st))) return rc;
  StoreNatural so{reinterpret_cast<float2*>(s->d_tim), p->n1, 0};
  if ((rc = launch_pass2<true>(p, so, st))) return rc;
  if (h_mean || h_std) {
    float h[3];
    PS_CUDA(cudaMemcpyAsync(h, s->d_stats, sizeof(h), cudaMemcpyDeviceToHost, st), "Error from d2hcpy");
    PS_CUDA(cudaStreamSynchronize(st), "ps_search_whiten");
    if (h_mean) *h_mean = h[0];
    if (h_std) *h_std = h[2];
  }
  return PS_OK;
}

int ps_search_accel_batch(ps_search* hs, const float* h_acc, int n_acc, int acc_idx0, void* stream) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_accel_batch: null handle");
  cudaStream_t st = S(stream);
  const ps_fft_plan_t* p = s->plan;
  unsigned size = s->prm.size;
  int rc;
  for (int j = 0; j < n_acc; j++) {
    LoadResampled ld{s->d_tim, accel_fact(h_acc[j], s->prm.tsamp), (double)size, p->n2, 0};
    if ((rc = launch_pass1<false>(p, ld, st))) return rc;
    StoreNatural sn{reinterpret_cast<float2*>(p->Z), p->n1, 0};
    if ((rc = launch_pass2<false>(p, sn, st))) return rc;
    spectrum_kernel<<<ew_grid(s->nbins, 256, 2), 256, 0, st>>>(reinterpret_cast<const float2*>(p->Z), s->d_pspec, p->M,
                                                               s->d_stats, (float)size);
    PS_LAUNCHED("spectrum_kernel");
    harmonic_peaks_kernel<<<ew_grid(s->nbins, 256, 2), 256, 0, st>>>(s->d_pspec, (int)s->nbins, s->prm.nharmonics, s->ranges,
                                                                     s->prm.min_snr, acc_idx0 + j, s->d_peaks, s->d_count,
                                                                     s->cap);
    PS_LAUNCHED("harmonic_peaks_kernel");
  }
  return PS_OK;
}

int ps_search_reset_peaks(ps_search* hs, void* stream) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_reset_peaks: null handle");
  PS_CUDA(cudaMemsetAsync(s->d_count, 0, sizeof(unsigned), S(stream)), "ps_search_reset_peaks");
  return PS_OK;
}

long ps_search_fetch_peaks(ps_search* hs, ps_raw_peak* h_out, size_t max_out, void* stream) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_fetch_peaks: null handle");
  cudaStream_t st = S(stream);
  unsigned count = 0;
  PS_CUDA(cudaMemcpyAsync(&count, s->d_count, sizeof(unsigned), cudaMemcpyDeviceToHost, st), "Error from d2hcpy");
  PS_CUDA(cudaStreamSynchronize(st), "ps_search_fetch_peaks");
  if (count > s->cap || count > max_out) {
    set_error("ps_search_fetch_peaks: more above-threshold bins than the peak list holds (raise max_peaks)");
    return PS_EOVERFLOW;
  }
  if (!count) {
    ps_count_transfer(0, sizeof(unsigned));
    return 0;
  }
  PS_CUDA(cudaMemcpyAsync(h_out, s->d_peaks, sizeof(ps_raw_peak) * count, cudaMemcpyDeviceToHost, st), "Error from d2hcpy");
  ps_count_transfer(0, sizeof(unsigned) + sizeof(ps_raw_peak) * (unsigned long long)count);
  PS_CUDA(cudaStreamSynchronize(st), "ps_search_fetch_peaks");
  std::sort(h_out, h_out + count, [](const ps_raw_peak& a, const ps_raw_peak& b) {
    if (a.acc_idx != b.acc_idx) return a.acc_idx < b.acc_idx;
    if (a.nh != b.nh) return a.nh < b.nh;
    return a.bin < b.bin;
  });
  return (long)count;
}

const float* ps_search_whitened_series(ps_search* s) { return s ? SS(s)->d_tim : nullptr; }
const float* ps_search_last_spectrum(ps_search* s) { return s ? SS(s)->d_pspec : nullptr; }

}  // extern "C"
