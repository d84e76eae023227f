// fold.cu -- candidate folding and fold optimisation: the stage that follows the post-join
// distillation when --npdmp > 0.
//
// Reference: include/transforms/folder.hpp (TimeSeriesFolder :34-62, FoldOptimiser :65-352) over
// the kernels of src/kernels.cu:597-865 and batched cuFFT C2C plans of nbins points.  Here the
// nbins-point transforms run inside the kernels that need them (radix-2 in shared memory, one
// transform per CTA), which removes the 4032 x 64 complex intermediate array, its |.| copy and
// the thrust::max_element pass: the template search writes one (value, index) pair per CTA.
//
// Difference kept on purpose: the reference's fold kernel zeroes, accumulates (shared-memory
// atomics) and reads back its phase bins without a barrier between the three steps, so a sample
// can be lost when the two warps of a block drift apart -- the golden FOLD blocks show it in one
// or two of the 16 sub-integrations.  The fold here adds every sample in a fixed order (chunks of
// 2048 samples in sample order, then the chunk sums in order), so the result is deterministic.
#include <cuComplex.h>
#include <cmath>
#include <vector>
#include "common.cuh"

namespace ps {

constexpr int kFoldChunk = 2048;  // samples per partial sum

// Sub-integration s covers samples [s*per, s*per + ceil(per/nbins)*nbins) (what the reference's
// nbins threads striding by nbins cover, kernels.cu:620-627), clamped to the series.  The order
// of the additions is fixed: every chunk of kFoldChunk samples is summed per phase bin in sample
// order (fold_partial_kernel, one CTA per chunk), then the chunk sums of a sub-integration are
// added in chunk order (fold_finish_kernel).  Phase bin = floor(frac(j * tsamp/period) * nbins)
// in double; out = sum / (hits + 1).  oracle/fold.py adds in the same order: bit-identical.
__global__ void __launch_bounds__(256) fold_partial_kernel(const float* __restrict__ in, float* __restrict__ psum,
                                                           int* __restrict__ pcount, size_t nsamps, size_t per, int nbins,
                                                           int nchunks, double tsamp_by_period) {
  __shared__ float val[kFoldChunk];
  __shared__ unsigned short bin[kFoldChunk];
  const int s = blockIdx.y, chunk = blockIdx.x;
  const size_t first = per * s;
  size_t last = first + (per + nbins - 1) / nbins * nbins;
  if (last > nsamps) last = nsamps;
  const size_t c0 = first + (size_t)chunk * kFoldChunk;
  const int len = c0 >= last ? 0 : (int)((last - c0) < (size_t)kFoldChunk ? (last - c0) : (size_t)kFoldChunk);
  for (int i = threadIdx.x; i < len; i += blockDim.x) {
    const size_t j = c0 + i;
    double ip;
    const double fp = modf((double)j * tsamp_by_period, &ip);
    bin[i] = (unsigned short)__double2int_rd(fp * nbins);
    val[i] = in[j];
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) {
    float sum = 0.f;
    int count = 0;
    for (int i = 0; i < len; i++)
      if (bin[i] == b) {
        sum += val[i];
        count++;
      }
    psum[((size_t)s * nchunks + chunk) * nbins + b] = sum;
    pcount[((size_t)s * nchunks + chunk) * nbins + b] = count;
  }
}

__global__ void fold_finish_kernel(const float* __restrict__ psum, const int* __restrict__ pcount, float* __restrict__ out, int nbins,
                                   int nchunks) {
  const int s = blockIdx.x;
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) {
    float sum = 0.f;
    int count = 1;  // the reference's hit counter starts at one (kernels.cu:616)
    for (int c = 0; c < nchunks; c++) {
      sum += psum[((size_t)s * nchunks + c) * nbins + b];
      count += pcount[((size_t)s * nchunks + c) * nbins + b];
    }
    out[(size_t)s * nbins + b] = sum / count;
  }
}

__device__ __forceinline__ int brev(int v, int lg) { return (int)(__brev((unsigned)v) >> (32 - lg)); }

// In-place radix-2 transform of n = blockDim.x points held bit-reversed in shared memory.
// tw[k] = exp(-2 pi i k / n), k < n/2; INV conjugates it (unnormalised inverse, like cuFFT).
template <bool INV>
__device__ __forceinline__ void cta_fft(cuComplex* s, const cuComplex* __restrict__ tw, int n, int lg) {
  const int t = threadIdx.x;
  for (int st = 1; st <= lg; st++) {
    __syncthreads();
    if (t < n / 2) {
      const int half = 1 << (st - 1);
      const int j = t & (half - 1);
      const int base = ((t >> (st - 1)) << st) + j;
      cuComplex w = tw[j * (n >> st)];
      if (INV) w.y = -w.y;
      cuComplex a = s[base], b = cuCmulf(s[base + half], w);
      s[base] = cuCaddf(a, b);
      s[base + half] = cuCsubf(a, b);
    }
  }
  __syncthreads();
}

// boxcar templates (kernels.cu:680-691) transformed to the Fourier domain: one CTA per width
__global__ void template_kernel(cuComplex* __restrict__ templates, const cuComplex* __restrict__ tw, int nbins, int lg) {
  extern __shared__ cuComplex sm[];
  const int b = threadIdx.x, t = blockIdx.x;
  sm[brev(b, lg)] = make_cuComplex(b <= t ? 1.f : 0.f, 0.f);
  cta_fft<false>(sm, tw, nbins, lg);
  templates[(size_t)t * nbins + b] = sm[b];
}

// phase ramps exp(-i * ramp(bin) * subint/nints * shift) (kernels.cu:659-678), float arithmetic
__global__ void shift_array_kernel(cuComplex* __restrict__ shift_ar, unsigned size, unsigned nbins, unsigned nints,
                                   float two_pi) {
  unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= size) return;
  float subint = (float)(idx / nbins % nints);
  unsigned shift_idx = idx / (nbins * nints);
  unsigned bin = idx % nbins;
  float mag = (float)((int)shift_idx - (int)(nbins / 2));  // folder.hpp:129 with nshifts = nbins
  float shift = __fmul_rn(__fdiv_rn(subint, (float)nints), mag);
  float ramp = __fdiv_rn(__fmul_rn((float)bin, two_pi), (float)nbins);
  if (bin > nbins / 2) ramp = __fsub_rn(ramp, two_pi);
  float sn, cs;
  sincosf(-__fmul_rn(ramp, shift), &sn, &cs);
  shift_ar[idx] = make_cuComplex(cs, sn);
}

// forward transform of every sub-integration (folder.hpp:249-254): one CTA per row
__global__ void subint_spectra_kernel(const float* __restrict__ fold, cuComplex* __restrict__ spectra,
                                      const cuComplex* __restrict__ tw, int nbins, int lg) {
  extern __shared__ cuComplex sm[];
  const int b = threadIdx.x;
  sm[brev(b, lg)] = make_cuComplex(fold[(size_t)blockIdx.x * nbins + b], 0.f);
  cta_fft<false>(sm, tw, nbins, lg);
  spectra[(size_t)blockIdx.x * nbins + b] = sm[b];
}

// device_multiply_by_shift + device_collapse_subints (folder.hpp:258-264): one CTA per drift
__global__ void shift_collapse_kernel(const cuComplex* __restrict__ spectra, const cuComplex* __restrict__ shift_ar,
                                      cuComplex* __restrict__ post, cuComplex* __restrict__ profiles, int nbins, int nints) {
  const int b = threadIdx.x;
  const size_t base = (size_t)blockIdx.x * nints * nbins;
  cuComplex acc = make_cuComplex(0.f, 0.f);
  for (int i = 0; i < nints; i++) {
    cuComplex v = cuCmulf(spectra[i * nbins + b], shift_ar[base + i * nbins + b]);
    post[base + i * nbins + b] = v;
    acc = cuCaddf(acc, v);
  }
  profiles[(size_t)blockIdx.x * nbins + b] = acc;
}

struct BlockBest {
  float value;
  unsigned index;
};

// first maximum (lowest index among equal values), like thrust::max_element
__device__ __forceinline__ void best_of(float& v, unsigned& i, float ov, unsigned oi) {
  if (ov > v || (ov == v && oi < i)) {
    v = ov;
    i = oi;
  }
}

__device__ __forceinline__ void cta_best(float& v, unsigned& i, float* sv, unsigned* si) {
  for (int o = 16; o > 0; o >>= 1) best_of(v, i, __shfl_down_sync(0xffffffffu, v, o), __shfl_down_sync(0xffffffffu, i, o));
  const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if ((threadIdx.x & 31) == 0) {
    sv[w] = v;
    si[w] = i;
  }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int k = 1; k < nw; k++) best_of(v, i, sv[k], si[k]);
}

// device_multiply_by_templates + inverse transform + |.| + argmax (folder.hpp:270-283):
// CTA (template t, drift s) convolves collapsed profile s with boxcar t+1 and keeps its best bin.
// NaN magnitudes are left out of the comparison.
__global__ void template_search_kernel(const cuComplex* __restrict__ profiles, const cuComplex* __restrict__ templates,
                                       const cuComplex* __restrict__ tw, BlockBest* __restrict__ best, int nbins, int lg,
                                       int nshifts) {
  extern __shared__ cuComplex sm[];
  __shared__ float sv[8];
  __shared__ unsigned si[8];
  const int b = threadIdx.x;
  const int t = blockIdx.x / nshifts, s = blockIdx.x % nshifts;
  cuComplex v = make_cuComplex(0.f, 0.f);
  if (b != 0) {
    cuComplex norm = make_cuComplex(sqrtf((float)t + 1.0f), 0.f);
    v = cuCdivf(cuCmulf(profiles[(size_t)s * nbins + b], templates[(size_t)t * nbins + b]), norm);
  }
  sm[brev(b, lg)] = v;
  cta_fft<true>(sm, tw, nbins, lg);
  float mag = cuCabsf(sm[b]);
  unsigned idx = (unsigned)blockIdx.x * nbins + b;
  if (!(mag == mag)) mag = -INFINITY;
  cta_best(mag, idx, sv, si);
  if (b == 0) best[blockIdx.x] = {mag, idx};
}

// out[0] = argmax over the CTAs' bests; then the optimised sub-integrations and profile
// (folder.hpp:285-311) are transformed back by optimised_kernel, which reads the argmax on device.
__global__ void __launch_bounds__(1024) argmax_kernel(const BlockBest* __restrict__ best, int n, BlockBest* __restrict__ out) {
  __shared__ float sv[32];
  __shared__ unsigned si[32];
  float v = -INFINITY;
  unsigned i = 0xffffffffu;
  for (int k = threadIdx.x; k < n; k += blockDim.x) best_of(v, i, best[k].value, best[k].index);
  cta_best(v, i, sv, si);
  if (threadIdx.x == 0) *out = {v, i == 0xffffffffu ? 0u : i};
}

// CTA r < nints: sub-integration r of the best drift back to the phase domain (real part);
// CTA nints: the collapsed profile of the best drift
__global__ void optimised_kernel(const cuComplex* __restrict__ post, const cuComplex* __restrict__ profiles,
                                 const cuComplex* __restrict__ tw, const BlockBest* __restrict__ arg,
                                 float* __restrict__ opt_fold, float* __restrict__ opt_prof, int nbins, int lg, int nints) {
  extern __shared__ cuComplex sm[];
  const int b = threadIdx.x, r = blockIdx.x;
  const unsigned opt_shift = arg->index / nbins % nbins;
  const cuComplex* src = r < nints ? post + ((size_t)opt_shift * nints + r) * nbins : profiles + (size_t)opt_shift * nbins;
  sm[brev(b, lg)] = src[b];
  cta_fft<true>(sm, tw, nbins, lg);
  if (r < nints)
    opt_fold[(size_t)r * nbins + b] = sm[b].x;
  else
    opt_prof[b] = sm[b].x;
}

}  // namespace ps

using namespace ps;

struct ps_fold_optimiser {
  int nbins = 0, nints = 0, lg = 0, nshifts = 0, ntemplates = 0, device = 0;
  cuComplex *tw = nullptr, *templates = nullptr, *shiftar = nullptr, *spectra = nullptr, *post = nullptr, *profiles = nullptr;
  BlockBest* best = nullptr;
  char* d_result = nullptr;  // BlockBest argmax | opt_fold [nints*nbins] | opt_prof [nbins]
  char* h_result = nullptr;  // pinned mirror
  size_t result_bytes = 0;
};

// S/N of the optimised profile, both estimators of folder.hpp:143-191 with their C++ promotions:
// sn1 = (on-pulse mean - off-pulse mean) * sqrt(width) / off-pulse rms, sn2 = sum of the
// normalised profile / sqrt(width); the pulse is rotated to bin nbins/2-1 first.
static void profile_sn(const float* prof, int bin, unsigned width, unsigned nbins, float* sn1, float* sn2) {
  const int reach = (int)(width / 2.0 + 0.5) + (int)(width * 0.3 + 0.5);
  const int centre = (int)(nbins / 2) - 1;
  std::vector<float> rot(nbins), off;
  for (unsigned i = 0; i < nbins; i++) rot[i] = prof[(bin - nbins / 2 + i) % nbins];  // unsigned wrap, then modulo
  double on_sum = 0.0, off_sum = 0.0;
  size_t n_on = 0;
  for (int i = 0; i < (int)nbins; i++) {
    if (i <= centre + reach && i >= centre - reach) {
      on_sum += rot[i];
      n_on++;
    } else {
      off_sum += rot[i];
      off.push_back(rot[i]);
    }
  }
  float on_mean = on_sum / n_on;
  float off_mean = off_sum / off.size();
  float acc = 0;
  for (size_t i = 0; i < off.size(); i++) acc += std::pow(off[i] - off_mean, 2.0);
  float off_std = std::sqrt(acc / off.size());
  *sn1 = (on_mean - off_mean) * std::sqrt((double)width) / off_std;
  double total = 0.0;
  for (unsigned i = 0; i < nbins; i++) total += (rot[i] - off_mean) / off_std;
  *sn2 = total / std::sqrt((double)width);
  if (*sn1 > 99999) *sn1 = 0.0;
  if (*sn2 > 99999) *sn2 = 0.0;
}

extern "C" {

int ps_fold_timeseries(const float* d_in, float* d_out, size_t nsamps, size_t nsubints, double period, double tsamp, int nbins,
                       void* stream) {
  if (nbins < 1 || nbins > 256) return invalid("ps_fold_timeseries: nbins must be 1..256");
  if (!nsubints || nsamps < nsubints) return invalid("ps_fold_timeseries: nsamps must be >= nsubints > 0");
  if (!(period > 0)) return invalid("ps_fold_timeseries: period must be positive");
  const size_t per = nsamps / nsubints;
  const size_t span = (per + nbins - 1) / nbins * nbins;
  const int nchunks = (int)((span + kFoldChunk - 1) / kFoldChunk);
  if (nsubints > 65535) return invalid("ps_fold_timeseries: too many sub-integrations");
  cudaStream_t st = S(stream);
  float* psum = nullptr;  // [nsubints][nchunks][nbins] partial sums, then the same shape of hit counts
  const size_t cells = nsubints * (size_t)nchunks * nbins;
  PS_CUDA(cudaMallocAsync(&psum, cells * (sizeof(float) + sizeof(int)), st), "ps_fold_timeseries: scratch");
  int* pcount = reinterpret_cast<int*>(psum + cells);
  fold_partial_kernel<<<dim3(nchunks, (unsigned)nsubints), 256, 0, st>>>(d_in, psum, pcount, nsamps, per, nbins, nchunks, tsamp / period);
  g_launches.fetch_add(1);
  fold_finish_kernel<<<(unsigned)nsubints, 256, 0, st>>>(psum, pcount, d_out, nbins, nchunks);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  cudaFreeAsync(psum, st);
  if (e != cudaSuccess) return cuda_fail(e, "ps_fold_timeseries");
  return PS_OK;
}

void ps_fold_optimiser_destroy(ps_fold_optimiser* o) {
  if (!o) return;
  cudaFree(o->tw); cudaFree(o->templates); cudaFree(o->shiftar); cudaFree(o->spectra); cudaFree(o->post);
  cudaFree(o->profiles); cudaFree(o->best); cudaFree(o->d_result);
  if (o->h_result) cudaFreeHost(o->h_result);
  delete o;
}

int ps_fold_optimiser_create(unsigned nbins, unsigned nints, ps_fold_optimiser** out) {
  if (!out) return invalid("ps_fold_optimiser_create: null output");
  *out = nullptr;
  int lg = 0;
  while ((1u << lg) < nbins) lg++;
  if (nbins < 32 || nbins > 256 || (1u << lg) != nbins) return invalid("ps_fold_optimiser_create: nbins must be 32, 64, 128 or 256");
  if (nints < 1 || nints > 1024) return invalid("ps_fold_optimiser_create: nints must be 1..1024");
  ps_fold_optimiser* o = new ps_fold_optimiser;
  o->nbins = nbins; o->nints = nints; o->lg = lg; o->nshifts = nbins; o->ntemplates = nbins - 1;
  cudaGetDevice(&o->device);
  const size_t cell = sizeof(cuComplex);
  o->result_bytes = sizeof(BlockBest) + sizeof(float) * ((size_t)nints * nbins + nbins);
  std::vector<cuComplex> tw(nbins / 2);
  for (unsigned k = 0; k < nbins / 2; k++) {
    double a = -2.0 * M_PI * k / nbins;
    tw[k] = make_cuComplex((float)cos(a), (float)sin(a));
  }
  cudaError_t e = cudaMalloc(&o->tw, cell * nbins / 2);
  if (e == cudaSuccess) e = cudaMalloc(&o->templates, cell * o->ntemplates * nbins);
  if (e == cudaSuccess) e = cudaMalloc(&o->shiftar, cell * o->nshifts * nints * nbins);
  if (e == cudaSuccess) e = cudaMalloc(&o->spectra, cell * nints * nbins);
  if (e == cudaSuccess) e = cudaMalloc(&o->post, cell * o->nshifts * nints * nbins);
  if (e == cudaSuccess) e = cudaMalloc(&o->profiles, cell * o->nshifts * nbins);
  if (e == cudaSuccess) e = cudaMalloc(&o->best, sizeof(BlockBest) * o->ntemplates * o->nshifts);
  if (e == cudaSuccess) e = cudaMalloc(&o->d_result, o->result_bytes);
  if (e == cudaSuccess) e = cudaMallocHost(&o->h_result, o->result_bytes);
  if (e == cudaSuccess) e = cudaMemcpy(o->tw, tw.data(), cell * nbins / 2, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    template_kernel<<<o->ntemplates, nbins, cell * nbins>>>(o->templates, o->tw, nbins, lg);
    g_launches.fetch_add(1);
    unsigned size = o->nshifts * nints * nbins;
    const float two_pi = 2 * 3.14159265359;  // kernels.cu:786
    shift_array_kernel<<<(size + 255) / 256, 256>>>(o->shiftar, size, nbins, nints, two_pi);
    g_launches.fetch_add(1);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
  }
  if (e != cudaSuccess) {
    ps_fold_optimiser_destroy(o);
    return cuda_fail(e, "ps_fold_optimiser_create");
  }
  *out = o;
  return PS_OK;
}

int ps_fold_optimise(ps_fold_optimiser* o, const float* d_fold, double period, float tobs, float* h_opt_fold, float* h_opt_prof,
                     ps_fold_result* res, void* stream) {
  if (!o || !d_fold || !res) return invalid("ps_fold_optimise: null argument");
  const int nbins = o->nbins, nints = o->nints, lg = o->lg;
  const size_t sm = sizeof(cuComplex) * nbins;
  cudaStream_t st = S(stream);
  BlockBest* d_arg = (BlockBest*)o->d_result;
  float* d_opt_fold = (float*)(o->d_result + sizeof(BlockBest));
  float* d_opt_prof = d_opt_fold + (size_t)nints * nbins;
  subint_spectra_kernel<<<nints, nbins, sm, st>>>(d_fold, o->spectra, o->tw, nbins, lg);
  PS_LAUNCHED("ps_fold_optimise: subint spectra");
  shift_collapse_kernel<<<o->nshifts, nbins, 0, st>>>(o->spectra, o->shiftar, o->post, o->profiles, nbins, nints);
  PS_LAUNCHED("ps_fold_optimise: shift + collapse");
  const int nblocks = o->ntemplates * o->nshifts;
  template_search_kernel<<<nblocks, nbins, sm, st>>>(o->profiles, o->templates, o->tw, o->best, nbins, lg, o->nshifts);
  PS_LAUNCHED("ps_fold_optimise: template search");
  argmax_kernel<<<1, 1024, 0, st>>>(o->best, nblocks, d_arg);
  PS_LAUNCHED("ps_fold_optimise: argmax");
  optimised_kernel<<<nints + 1, nbins, sm, st>>>(o->post, o->profiles, o->tw, d_arg, d_opt_fold, d_opt_prof, nbins, lg, nints);
  PS_LAUNCHED("ps_fold_optimise: optimised profile");
  PS_CUDA(cudaMemcpyAsync(o->h_result, o->d_result, o->result_bytes, cudaMemcpyDeviceToHost, st), "ps_fold_optimise: D2H");
  PS_CUDA(cudaStreamSynchronize(st), "ps_fold_optimise: sync");
  ps_count_transfer(0, o->result_bytes);
  const BlockBest* arg = (const BlockBest*)o->h_result;
  const float* opt_fold = (const float*)(o->h_result + sizeof(BlockBest));
  const float* opt_prof = opt_fold + (size_t)nints * nbins;
  const unsigned argmax = arg->index;
  const unsigned opt_template = argmax / (nbins * o->nshifts);  // folder.hpp:285-287
  const int opt_bin = (int)(argmax % nbins) - (int)(opt_template / 2);
  const unsigned opt_shift = argmax / nbins % nbins;
  res->argmax = argmax;
  res->peak = arg->value;
  profile_sn(opt_prof, opt_bin, opt_template, nbins, &res->sn1, &res->sn2);  // width = template index, :320
  res->opt_sn = std::max(res->sn1, res->sn2);
  res->opt_period = period * ((((32.0 - opt_shift) * period) / ((unsigned)nbins * tobs)) + 1);  // :346
  res->opt_width = (int)opt_template + 1;
  res->opt_bin = opt_bin;
  res->opt_shift = (int)opt_shift;
  if (h_opt_fold) memcpy(h_opt_fold, opt_fold, sizeof(float) * nints * nbins);
  if (h_opt_prof) memcpy(h_opt_prof, opt_prof, sizeof(float) * nbins);
  return PS_OK;
}

}  // extern "C"
