// stages.cu -- element-wise / gather / reduction stages of the search hot path, one
// extern "C" entry per reference function in include/kernels/kernels.h (citations in
// include/peasoup_b200.h).  All kernels are HBM-bound streaming kernels: grid-stride loops
// sized to whole waves of the 148 SMs, float4/float2 vector accesses where alignment
// allows, no shared-memory staging (no reuse), warp-aggregated atomics for compaction.
#include <cuComplex.h>
#include <algorithm>
#include <map>
#include <mutex>
#include <vector>
#include "common.cuh"
#include "stages.cuh"

namespace ps {

thread_local std::string t_err;
std::atomic<unsigned long long> g_launches{0};
void set_error(const std::string& msg) { t_err = msg; }

// ---------------------------------------------------------------- resampling
// kernels.cu:354: af = double(float(a*tsamp)) / (2*c)
double accel_fact(float a, float tsamp) { return ((a * tsamp) / (2 * 299792458.0)); }

template <int WHICH>
__global__ void __launch_bounds__(256) resample_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                       uint32_t* __restrict__ map, double af, double size,
                                                       double size_by_2, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    unsigned long long r = WHICH == 2 ? resample_index_II(af, size, i) : resample_index_I(af, size_by_2, i);
    if (map) map[i] = (uint32_t)r;
    if (out) out[i] = in[r];
  }
}

// ---------------------------------------------------------------- spectra
__global__ void __launch_bounds__(256) power_kernel(const float2* __restrict__ in, float* __restrict__ out, size_t nbins) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nbins; i += (size_t)gridDim.x * blockDim.x) {
    float2 x = in[i];
    out[i] = amplitude_rsqrt(x.x, x.y);
  }
}

__global__ void __launch_bounds__(256) interbin_kernel(const float2* __restrict__ in, float* __restrict__ out, size_t nbins) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nbins; i += (size_t)gridDim.x * blockDim.x) {
    float2 x = in[i];
    float2 l = i > 0 ? in[i - 1] : make_float2(0.f, 0.f);
    out[i] = interbin_amplitude(x.x, x.y, l.x, l.y);
  }
}

__global__ void __launch_bounds__(256) normalise_kernel(float* __restrict__ d, float mean, float sigma, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    d[i] = normalise_value(d[i], mean, sigma);
  }
}

// ---------------------------------------------------------------- harmonic sums
struct FoldPtrs {
  float* p[5];
};

__global__ void __launch_bounds__(256) harmonic_sum_kernel(const float* __restrict__ in, FoldPtrs out, int nbins, int nfolds) {
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nbins; idx += gridDim.x * blockDim.x) {
    float f[5];
    harmonic_folds(in, idx, nfolds, f);
#pragma unroll
    for (int k = 0; k < 5; k++)
      if (k < nfolds) out.p[k][idx] = f[k];
  }
}

// ---------------------------------------------------------------- peaks
__global__ void __launch_bounds__(256) find_peaks_kernel(const float* __restrict__ d, int n, int start, float thresh,
                                                         int* __restrict__ o_idx, float* __restrict__ o_snr,
                                                         unsigned* __restrict__ counter, unsigned cap) {
  int base = start + (blockIdx.x * blockDim.x);
  for (int i0 = base; i0 < n; i0 += gridDim.x * blockDim.x) {
    int i = i0 + threadIdx.x;
    float v = i < n ? d[i] : 0.f;
    bool hit = i < n && v > thresh;
    unsigned m = __ballot_sync(0xffffffffu, hit);
    if (m) {
      int lane = threadIdx.x & 31;
      unsigned pos = 0;
      if (lane == 0) pos = atomicAdd(counter, (unsigned)__popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1));
      if (hit && pos < cap) {
        o_idx[pos] = i;
        o_snr[pos] = v;
      }
    }
  }
}

// ---------------------------------------------------------------- rednoise
__global__ void __launch_bounds__(256) scrunch5_kernel(const float* __restrict__ in, size_t out_count, float* __restrict__ out) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < out_count; i += (size_t)gridDim.x * blockDim.x) {
    const float* p = in + 5 * i;
    out[i] = median_of_5(p[0], p[1], p[2], p[3], p[4]);
  }
}

__global__ void scrunch5_small_kernel(const float* __restrict__ in, int count, float* __restrict__ out) {
  // kernels.cu:954-970 special cases
  if (count == 1) out[0] = in[0];
  if (count == 2) out[0] = 0.5f * (in[0] + in[1]);
  if (count == 3) out[0] = median_of_3(in[0], in[1], in[2]);
  if (count == 4) out[0] = median_of_4(in[0], in[1], in[2], in[3]);
}

__global__ void __launch_bounds__(256) stretch_kernel(const float* __restrict__ in, size_t in_count, float step,
                                                      float* __restrict__ out, size_t out_count) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < out_count; i += (size_t)gridDim.x * blockDim.x)
    out[i] = stretch_value(in, in_count, step, (unsigned)i);
}

__global__ void __launch_bounds__(256) spliced_median_kernel(const float* __restrict__ m5, size_t n5, float s5,
                                                             const float* __restrict__ m25, size_t n25, float s25,
                                                             const float* __restrict__ m125, size_t n125, float s125,
                                                             int pos5, int pos25, float* __restrict__ out, size_t nbins) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nbins; i += (size_t)gridDim.x * blockDim.x)
    out[i] = spliced_median_value(m5, n5, s5, m25, n25, s25, m125, n125, s125, pos5, pos25, (unsigned)i);
}

__global__ void __launch_bounds__(256) divide_c_by_f_kernel(float2* __restrict__ c, const float* __restrict__ f, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    c[i] = deredden_value(c[i], f[i], (unsigned)i);
}

__global__ void zap_kernel(float2* __restrict__ c, const float* __restrict__ birdies, const float* __restrict__ widths,
                           float bin_width, unsigned nbirdies, unsigned nbins) {
  // one warp per birdie; lanes stride over the zapped bins (the reference loops in one thread)
  unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= nbirdies) return;
  int lo, hi;
  if (!zap_range(birdies[w], widths[w], bin_width, nbins, &lo, &hi)) return;
  for (int i = lo + (int)lane; i < hi; i += 32) c[i] = make_float2(1.f, 0.f);
}

// ---------------------------------------------------------------- stats
// two-stage deterministic reduction with double accumulators: partials[2*b] = sum, [2*b+1] = sum sq
__global__ void __launch_bounds__(256) sums_kernel(const float* __restrict__ d, size_t n, double* __restrict__ partials) {
  double s = 0.0, q = 0.0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double v = d[i];
    s += v;
    q += v * v;
  }
  __shared__ double sh[2][8];
  for (int o = 16; o; o >>= 1) {
    s += __shfl_down_sync(0xffffffffu, s, o);
    q += __shfl_down_sync(0xffffffffu, q, o);
  }
  int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    sh[0][w] = s;
    sh[1][w] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < 8; i++) {
      a += sh[0][i];
      b += sh[1][i];
    }
    partials[2 * blockIdx.x] = a;
    partials[2 * blockIdx.x + 1] = b;
  }
}

// stats.hpp:20-33 + kernels.cu:426-454 in float, no contraction: out = {mean, rms, std}
__global__ void __launch_bounds__(256) finalize_stats_kernel(const double* __restrict__ partials, int nparts, float count,
                                                             float* __restrict__ out) {
  // one CTA: strided partial sums, then a fixed-order tree (deterministic; double accumulators)
  __shared__ double sh[2][256];
  double s = 0, q = 0;
  for (int i = threadIdx.x; i < nparts; i += 256) {
    s += partials[2 * i];
    q += partials[2 * i + 1];
  }
  sh[0][threadIdx.x] = s;
  sh[1][threadIdx.x] = q;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if ((int)threadIdx.x < o) {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
      sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x) return;
  s = sh[0][0];
  q = sh[1][0];
  float rms = sqrtf(__fdiv_rn((float)q, count));
  float mean = __fdiv_rn((float)s, count);
  float sd = sqrtf(__fsub_rn(__fmul_rn(rms, rms), __fmul_rn(mean, mean)));
  out[0] = mean;
  out[1] = rms;
  out[2] = sd;
}

__global__ void __launch_bounds__(256) fill_kernel(float* __restrict__ d, size_t n, float v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) d[i] = v;
}

__global__ void __launch_bounds__(256) convert_u8_kernel(const unsigned char* __restrict__ in, float* __restrict__ out, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = (float)in[i];
}

// ---------------------------------------------------------------- internal launch helpers
int launch_sums(const float* d, size_t n, double* d_partials, int* nparts, cudaStream_t st) {
  unsigned g = ew_grid(n, 256, 8);
  if (g > kMaxStatParts) g = kMaxStatParts;
  sums_kernel<<<g, 256, 0, st>>>(d, n, d_partials);
  PS_LAUNCHED("sums_kernel");
  *nparts = (int)g;
  return PS_OK;
}

int launch_finalize_stats(const double* d_partials, int nparts, size_t n, float* d_out3, cudaStream_t st) {
  finalize_stats_kernel<<<1, 256, 0, st>>>(d_partials, nparts, (float)(int)n, d_out3);
  PS_LAUNCHED("finalize_stats_kernel");
  return PS_OK;
}

int launch_stats(const float* d, size_t n, double* d_partials, float* d_out3, cudaStream_t st) {
  int np = 0;
  int rc = launch_sums(d, n, d_partials, &np, st);
  if (rc) return rc;
  return launch_finalize_stats(d_partials, np, n, d_out3, st);
}

int launch_running_median(const float* d_pspec, unsigned nbins, double bin_width, float b5, float b25, float* d_median,
                          float* d_work, cudaStream_t st) {
  // dereddener.hpp:48-52
  int pos5 = (int)(b5 / bin_width);
  int pos25 = (int)(b25 / bin_width);
  size_t n5 = nbins / 5, n25 = n5 / 5, n125 = n25 / 5;
  if (n125 < 2 || (size_t)pos25 > nbins || (size_t)pos5 + (size_t)pos25 > nbins)
    return invalid("Bad data length given to running_median()");
  float* m5 = d_work;
  float* m25 = m5 + n5 + 1;
  float* m125 = m25 + n25 + 1;
  scrunch5_kernel<<<ew_grid(n5, 256), 256, 0, st>>>(d_pspec, n5, m5);
  PS_LAUNCHED("median_scrunch5");
  scrunch5_kernel<<<ew_grid(n25, 256), 256, 0, st>>>(m5, n25, m25);
  PS_LAUNCHED("median_scrunch5");
  scrunch5_kernel<<<ew_grid(n125, 256), 256, 0, st>>>(m25, n125, m125);
  PS_LAUNCHED("median_scrunch5");
  // kernels.cu:989: step = float(in_count-1)/(out_count-1)
  float s5 = float(n5 - 1) / (size_t)(nbins - 1);
  float s25 = float(n25 - 1) / (size_t)(nbins - 1);
  float s125 = float(n125 - 1) / (size_t)(nbins - 1);
  spliced_median_kernel<<<ew_grid(nbins, 256, 4), 256, 0, st>>>(m5, n5, s5, m25, n25, s25, m125, n125, s125, pos5, pos25,
                                                               d_median, nbins);
  PS_LAUNCHED("spliced_median_kernel");
  return PS_OK;
}

struct Scratch {
  double* partials = nullptr;
  float* out3 = nullptr;
};
static std::mutex g_scratch_mu;
static std::map<int, Scratch> g_scratch;

static int get_scratch(Scratch* out) {
  int dev = 0;
  PS_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  auto it = g_scratch.find(dev);
  if (it == g_scratch.end()) {
    Scratch s;
    PS_CUDA(cudaMalloc(&s.partials, sizeof(double) * 2 * kMaxStatParts), "Error from device_malloc");
    PS_CUDA(cudaMalloc(&s.out3, sizeof(float) * 4), "Error from device_malloc");
    it = g_scratch.insert({dev, s}).first;
  }
  *out = it->second;
  return PS_OK;
}

}  // namespace ps

using namespace ps;

extern "C" {

const char* ps_last_error(void) { return ps::t_err.c_str(); }
int ps_version(void) { return 100; }
unsigned long long ps_launch_count(void) { return ps::g_launches.load(); }
static std::atomic<unsigned long long> g_h2d{0}, g_d2h{0};
void ps_count_transfer(unsigned long long h2d, unsigned long long d2h) {
  g_h2d.fetch_add(h2d, std::memory_order_relaxed);
  g_d2h.fetch_add(d2h, std::memory_order_relaxed);
}
void ps_transfer_bytes(unsigned long long* h2d, unsigned long long* d2h) {
  if (h2d) *h2d = g_h2d.load();
  if (d2h) *d2h = g_d2h.load();
}

int ps_resampleII(const float* d_in, float* d_out, size_t n, float a, float tsamp, void* stream) {
  if (!n) return PS_OK;
  double af = accel_fact(a, tsamp);
  resample_kernel<2><<<ew_grid(n, 256, 4), 256, 0, S(stream)>>>(d_in, d_out, nullptr, af, (double)n, (double)n / 2.0, n);
  PS_LAUNCHED("Error from device_resampleII");
  return PS_OK;
}

int ps_resample(const float* d_in, float* d_out, size_t n, float a, float tsamp, void* stream) {
  if (!n) return PS_OK;
  double af = accel_fact(a, tsamp);
  resample_kernel<1><<<ew_grid(n, 256, 4), 256, 0, S(stream)>>>(d_in, d_out, nullptr, af, (double)n, (double)n / 2.0, n);
  PS_LAUNCHED("Error from device_resample");
  return PS_OK;
}

int ps_resample_index_map(uint32_t* d_idx, size_t n, float a, float tsamp, int which, void* stream) {
  if (!n) return PS_OK;
  if (which != 1 && which != 2) return invalid("ps_resample_index_map: which must be 1 or 2");
  double af = accel_fact(a, tsamp);
  if (which == 2)
    resample_kernel<2><<<ew_grid(n, 256, 4), 256, 0, S(stream)>>>(nullptr, nullptr, d_idx, af, (double)n, (double)n / 2.0, n);
  else
    resample_kernel<1><<<ew_grid(n, 256, 4), 256, 0, S(stream)>>>(nullptr, nullptr, d_idx, af, (double)n, (double)n / 2.0, n);
  PS_LAUNCHED("ps_resample_index_map");
  return PS_OK;
}

int ps_form_power_series(const ps_complex* d_in, float* d_out, size_t nbins, int way, void* stream) {
  if (!nbins) return PS_OK;
  if (way == 1)
    interbin_kernel<<<ew_grid(nbins, 256, 4), 256, 0, S(stream)>>>((const float2*)d_in, d_out, nbins);
  else
    power_kernel<<<ew_grid(nbins, 256, 4), 256, 0, S(stream)>>>((const float2*)d_in, d_out, nbins);
  PS_LAUNCHED("Error from device_form_power_series");
  return PS_OK;
}

int ps_normalise(float* d_powers, float mean, float sigma, unsigned nbins, void* stream) {
  if (!nbins) return PS_OK;
  normalise_kernel<<<ew_grid(nbins, 256, 4), 256, 0, S(stream)>>>(d_powers, mean, sigma, nbins);
  PS_LAUNCHED("Error from device_normalise");
  return PS_OK;
}

int ps_harmonic_sum(const float* d_in, float* const* h_out_ptrs, size_t nbins, unsigned nfolds, void* stream) {
  if (nfolds > 5) return invalid("ps_harmonic_sum: at most 5 folds are defined (kernels.cu:42-96)");
  if (!nbins || !nfolds) return PS_OK;
  if (nbins > 0x7fffffffu / 32) return invalid("ps_harmonic_sum: nbins too large for 32-bit harmonic indices");
  FoldPtrs fp;
  for (unsigned i = 0; i < 5; i++) fp.p[i] = i < nfolds ? h_out_ptrs[i] : nullptr;
  harmonic_sum_kernel<<<ew_grid(nbins, 256, 2), 256, 0, S(stream)>>>(d_in, fp, (int)nbins, (int)nfolds);
  PS_LAUNCHED("Error from device_harmonic_sum");
  return PS_OK;
}

struct ps_peak_workspace {
  size_t cap;
  int* d_idx;
  float* d_snr;
  unsigned* d_count;
  std::vector<std::pair<int, float>> tmp;
  std::vector<int> hi;
  std::vector<float> hs;
  size_t needed = 0;  // above-threshold count of the last ps_find_peaks call
};

int ps_peak_workspace_create(size_t max_peaks, ps_peak_workspace** out) {
  if (!max_peaks) max_peaks = 100000;  // peakfinder.hpp:61 max_cands
  ps_peak_workspace* ws = new ps_peak_workspace();
  ws->cap = max_peaks;
  ws->d_idx = nullptr;
  ws->d_snr = nullptr;
  ws->d_count = nullptr;
  cudaError_t e = cudaMalloc(&ws->d_idx, sizeof(int) * max_peaks);
  if (e == cudaSuccess) e = cudaMalloc(&ws->d_snr, sizeof(float) * max_peaks);
  if (e == cudaSuccess) e = cudaMalloc(&ws->d_count, sizeof(unsigned));
  if (e != cudaSuccess) {
    ps_peak_workspace_destroy(ws);
    return cuda_fail(e, "Error from device_malloc");
  }
  ws->hi.resize(max_peaks);
  ws->hs.resize(max_peaks);
  *out = ws;
  return PS_OK;
}

void ps_peak_workspace_destroy(ps_peak_workspace* ws) {
  if (!ws) return;
  cudaFree(ws->d_idx);
  cudaFree(ws->d_snr);
  cudaFree(ws->d_count);
  delete ws;
}

int ps_find_peaks(int n, int start, const float* d_dat, float thresh, int* h_idx, float* h_snr, size_t max_out,
                  ps_peak_workspace* ws, void* stream) {
  if (!ws) return invalid("ps_find_peaks: null workspace");
  if (start < 0) start = 0;
  if (n <= start) return 0;
  cudaStream_t st = S(stream);
  PS_CUDA(cudaMemsetAsync(ws->d_count, 0, sizeof(unsigned), st), "ps_find_peaks memset");
  find_peaks_kernel<<<ew_grid((size_t)(n - start), 256, 2), 256, 0, st>>>(d_dat, n, start, thresh, ws->d_idx, ws->d_snr,
                                                                        ws->d_count, (unsigned)ws->cap);
  PS_LAUNCHED("Error from device_find_peaks;");
  unsigned count = 0;
  PS_CUDA(cudaMemcpyAsync(&count, ws->d_count, sizeof(unsigned), cudaMemcpyDeviceToHost, st), "ps_find_peaks d2h");
  PS_CUDA(cudaStreamSynchronize(st), "Error from device_find_peaks;");
  ws->needed = count;
  if (count > ws->cap || count > max_out) {
    set_error("ps_find_peaks: more peaks above threshold than the workspace/output can hold");
    return PS_EOVERFLOW;
  }
  if (!count) return 0;
  PS_CUDA(cudaMemcpyAsync(ws->hi.data(), ws->d_idx, sizeof(int) * count, cudaMemcpyDeviceToHost, st), "ps_find_peaks d2h");
  PS_CUDA(cudaMemcpyAsync(ws->hs.data(), ws->d_snr, sizeof(float) * count, cudaMemcpyDeviceToHost, st), "ps_find_peaks d2h");
  PS_CUDA(cudaStreamSynchronize(st), "Error from device_find_peaks;");
  ws->tmp.resize(count);
  for (unsigned i = 0; i < count; i++) ws->tmp[i] = {ws->hi[i], ws->hs[i]};
  std::sort(ws->tmp.begin(), ws->tmp.end(), [](const std::pair<int, float>& a, const std::pair<int, float>& b) { return a.first < b.first; });
  for (unsigned i = 0; i < count; i++) {
    h_idx[i] = ws->tmp[i].first;
    h_snr[i] = ws->tmp[i].second;
  }
  return (int)count;
}

size_t ps_peak_workspace_needed(const ps_peak_workspace* ws) { return ws ? ws->needed : 0; }

int ps_median_scrunch5(const float* d_in, size_t count, float* d_out, void* stream) {
  if (!count) return PS_OK;
  if (count < 5) {
    scrunch5_small_kernel<<<1, 1, 0, S(stream)>>>(d_in, (int)count, d_out);
  } else {
    size_t oc = count / 5;
    scrunch5_kernel<<<ew_grid(oc, 256), 256, 0, S(stream)>>>(d_in, oc, d_out);
  }
  PS_LAUNCHED("median_scrunch5");
  return PS_OK;
}

int ps_linear_stretch(const float* d_in, size_t in_count, float* d_out, size_t out_count, void* stream) {
  if (!out_count) return PS_OK;
  if (!in_count) return invalid("ps_linear_stretch: empty input");
  float step = float(in_count - 1) / (size_t)(out_count - 1);  // kernels.cu:989
  stretch_kernel<<<ew_grid(out_count, 256, 4), 256, 0, S(stream)>>>(d_in, in_count, step, d_out, out_count);
  PS_LAUNCHED("linear_stretch");
  return PS_OK;
}

int ps_running_median(const float* d_pspec, unsigned nbins, double bin_width, float b5, float b25, float* d_median,
                      float* d_work, void* stream) {
  return launch_running_median(d_pspec, nbins, bin_width, b5, b25, d_median, d_work, S(stream));
}

// ---- allocation-free variant of ps_running_median -------------------------------------
// Scratch for the three median levels lives in this file-scope device array instead of a
// caller-supplied buffer.  PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS bins (a 2^23-point search) need
// nbins/5 + nbins/25 + nbins/125 + 3 floats: about 4 MiB, resident on every GPU that loads the
// library until the process exits.
namespace ps {
constexpr size_t kNoAllocScratchWords = PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS / 5 + PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS / 25 +
                                        PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS / 125 + 8;
static __device__ __align__(256) float g_noalloc_median_scratch[kNoAllocScratchWords];

// In host code `g_noalloc_median_scratch` is only a handle for the runtime, not an address a
// kernel may dereference; the real device address is looked up once per device (with that
// device current) and cached.
static std::mutex g_noalloc_mu;
static float* g_noalloc_ptr[64] = {};

static int noalloc_scratch(float** out) {
  int dev = 0;
  PS_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
  if (dev < 0 || dev >= 64) return invalid("ps_running_median_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    float* p = nullptr;
    PS_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_median_scratch), "cudaGetSymbolAddress(median scratch)");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return PS_OK;
}
}  // namespace ps

int ps_running_median_noalloc(const float* d_pspec, unsigned nbins, double bin_width, float b5, float b25, float* d_median,
                              void* stream) {
  if (nbins > PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS) {
    set_error("ps_running_median_noalloc: nbins = " + std::to_string(nbins) + " exceeds the built-in scratch (max " +
              std::to_string((unsigned)PS_RUNNING_MEDIAN_NOALLOC_MAX_BINS) +
              " bins); use ps_running_median with a caller-provided work buffer");
    return PS_EINVAL;
  }
  float* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  // every level is fully rewritten before it is read, so nothing has to be cleared between calls
  return launch_running_median(d_pspec, nbins, bin_width, b5, b25, d_median, scratch, S(stream));
}

int ps_divide_c_by_f(ps_complex* d_c, const float* d_f, unsigned nbins, void* stream) {
  if (!nbins) return PS_OK;
  divide_c_by_f_kernel<<<ew_grid(nbins, 256, 4), 256, 0, S(stream)>>>((float2*)d_c, d_f, nbins);
  PS_LAUNCHED("Error from device_divide_c_by_f");
  return PS_OK;
}

int ps_zap_birdies(ps_complex* d_c, const float* d_birdies, const float* d_widths, float bin_width, unsigned nbirdies,
                   unsigned nbins, void* stream) {
  if (!nbirdies) return PS_OK;
  unsigned threads = 128, blocks = (nbirdies * 32 + threads - 1) / threads;
  zap_kernel<<<blocks, threads, 0, S(stream)>>>((float2*)d_c, d_birdies, d_widths, bin_width, nbirdies, nbins);
  PS_LAUNCHED("Error from device_zap_birdies");
  return PS_OK;
}

static int stats3(const float* d, int nsamps, int min_bin, float* out3) {
  if (nsamps <= min_bin) return invalid("ps_rms/ps_mean: empty range");
  Scratch sc;
  int rc = get_scratch(&sc);
  if (rc) return rc;
  rc = launch_stats(d + min_bin, (size_t)(nsamps - min_bin), sc.partials, sc.out3, 0);
  if (rc) return rc;
  PS_CUDA(cudaMemcpy(out3, sc.out3, sizeof(float) * 3, cudaMemcpyDeviceToHost), "Error from d2hcpy");
  return PS_OK;
}

int ps_rms(const float* d, int nsamps, int min_bin, float* out) {
  float o[3];
  int rc = stats3(d, nsamps, min_bin, o);
  if (rc) return rc;
  *out = o[1];
  return PS_OK;
}

int ps_mean(const float* d, int nsamps, int min_bin, float* out) {
  float o[3];
  int rc = stats3(d, nsamps, min_bin, o);
  if (rc) return rc;
  *out = o[0];
  return PS_OK;
}

int ps_fill(float* d_start, float* d_end, float value, void* stream) {
  if (d_end <= d_start) return PS_OK;
  size_t n = (size_t)(d_end - d_start);
  fill_kernel<<<ew_grid(n, 256, 4), 256, 0, S(stream)>>>(d_start, n, value);
  PS_LAUNCHED("Error in GPU_fill");
  return PS_OK;
}

int ps_convert_u8_f32(const unsigned char* d_in, float* d_out, unsigned n, void* stream) {
  if (!n) return PS_OK;
  convert_u8_kernel<<<ew_grid(n, 256, 4), 256, 0, S(stream)>>>(d_in, d_out, n);
  PS_LAUNCHED("Error from device_conversion");
  return PS_OK;
}

}  // extern "C"
