// bluestein.cu -- R2C / C2R of arbitrary length n on top of the power-of-two transforms of fft.cu.
//
// The reference plans cuFFT for whatever length it is given (ffter.hpp:46-78); the search path
// always asks for a power of two, but the coincidencer transforms the whole observation
// (coincidencer.cpp:137-139: size = nsamps).  Chirp-z (Bluestein): with w_k = exp(-i pi k^2 / n),
//     X_j = w_j * sum_k (x_k w_k) conj(w_{j-k}),
// a circular convolution of length M >= 2n-1 (M a power of two) evaluated with the transforms we
// already have.  A complex length-M transform is two real ones: FFT(ar + i ai) = R2C(ar) + i R2C(ai)
// and, going back, IFFT(C) = C2R(E) + i C2R(O) with E/O the Hermitian parts of C.
#include <vector>
#include "fft.cuh"

namespace ps {

struct Bluestein {
  unsigned n = 0, M = 0;
  float2* w = nullptr;      // chirp w_k, k < n
  float2* B = nullptr;      // spectrum of the wrapped conj chirp, M/2+1 bins (b is even: B_{M-k} = B_k)
  float *ar = nullptr, *ai = nullptr;    // M reals each: convolution input, then output
  float2 *SA = nullptr, *SB = nullptr;   // M/2+1 bins each
  double* part = nullptr;   // kMeanParts partial sums of the R2C input
  float* mean = nullptr;    // its mean: taken out before the convolution, put back into bin 0
};
constexpr int kMeanParts = 256;

// The rounding error of the convolution scales with the norm of its input, DC included; 8-bit
// samples sit on a mean of ~8 sigma.  A constant only feeds bin 0 of a length-n transform, so the
// mean is subtracted here and n * mean added to X_0 afterwards.  Fixed grid, fixed order.
__global__ void __launch_bounds__(256) partial_sums_kernel(const float* __restrict__ x, unsigned n, double* __restrict__ part) {
  __shared__ double sh[256];
  double acc = 0.0;
  for (unsigned k = blockIdx.x * 256 + threadIdx.x; k < n; k += gridDim.x * 256) acc += (double)x[k];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}

__global__ void __launch_bounds__(kMeanParts) mean_kernel(const double* __restrict__ part, unsigned n, float* __restrict__ mean) {
  __shared__ double sh[kMeanParts];
  sh[threadIdx.x] = part[threadIdx.x];
  __syncthreads();
  for (int s = kMeanParts / 2; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) *mean = (float)(sh[0] / (double)n);
}

// w_k = exp(-i pi (k^2 mod 2n) / n): the phase is reduced in integers, evaluated in double
__global__ void chirp_kernel(float2* __restrict__ w, unsigned n) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    unsigned long long r = ((unsigned long long)k * k) % (2ull * n);
    double s, c;
    sincospi((double)r / (double)n, &s, &c);
    w[k] = make_float2((float)c, (float)-s);
  }
}

// b_k = conj(w_k) for |k| < n (wrapped modulo M), 0 elsewhere; real and imaginary parts apart
__global__ void wrapped_chirp_kernel(const float2* __restrict__ w, float* __restrict__ br, float* __restrict__ bi, unsigned n,
                                     unsigned M) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < M; k += gridDim.x * blockDim.x) {
    unsigned d = k < n ? k : (M - k < n ? M - k : 0xffffffffu);
    float2 v = d == 0xffffffffu ? make_float2(0.f, 0.f) : w[d];
    br[k] = v.x;
    bi[k] = -v.y;
  }
}

__global__ void merge_spectra_kernel(float2* __restrict__ B, const float2* __restrict__ Br, const float2* __restrict__ Bi, unsigned nb) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < nb; k += gridDim.x * blockDim.x) {
    float2 r = Br[k], i = Bi[k];
    B[k] = make_float2(r.x - i.y, r.y + i.x);  // Br + i Bi
  }
}

// a_k = (x_k - mean) w_k (real input), zero padded to M
__global__ void pre_real_kernel(const float* __restrict__ x, const float2* __restrict__ w, float* __restrict__ ar, float* __restrict__ ai,
                                unsigned n, unsigned M, const float* __restrict__ mean) {
  const float mu = *mean;
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < M; k += gridDim.x * blockDim.x) {
    float r = 0.f, i = 0.f;
    if (k < n) {
      float v = x[k] - mu;
      float2 c = w[k];
      r = v * c.x;
      i = v * c.y;
    }
    ar[k] = r;
    ai[k] = i;
  }
}

// a_k = conj(Xfull_k) w_k with Xfull the Hermitian extension of the n/2+1 input bins
__global__ void pre_hermitian_kernel(const float2* __restrict__ X, const float2* __restrict__ w, float* __restrict__ ar,
                                     float* __restrict__ ai, unsigned n, unsigned M) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < M; k += gridDim.x * blockDim.x) {
    float r = 0.f, i = 0.f;
    if (k < n) {
      float2 v = k <= n / 2 ? X[k] : X[n - k];
      float vy = k <= n / 2 ? -v.y : v.y;  // conj of X_k, or conj(conj X_{n-k})
      if (k == 0 || 2 * k == n) vy = 0.f;  // the imaginary parts of DC and Nyquist are ignored, as cuFFT does
      float2 c = w[k];
      r = v.x * c.x - vy * c.y;
      i = v.x * c.y + vy * c.x;
    }
    ar[k] = r;
    ai[k] = i;
  }
}

// C = A .* B on the full circle from the half spectra; written back as its Hermitian parts E, O
__global__ void convolve_kernel(float2* __restrict__ SA, float2* __restrict__ SB, const float2* __restrict__ B, unsigned nb) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k < nb; k += gridDim.x * blockDim.x) {
    float2 p = SA[k], q = SB[k], b = B[k];
    // A_k = p + i q ;  A_{M-k} = conj(p) + i conj(q)
    float2 a1 = make_float2(p.x - q.y, p.y + q.x);
    float2 a2 = make_float2(p.x + q.y, -p.y + q.x);
    float2 c1 = make_float2(a1.x * b.x - a1.y * b.y, a1.x * b.y + a1.y * b.x);
    float2 c2 = make_float2(a2.x * b.x - a2.y * b.y, a2.x * b.y + a2.y * b.x);
    // E = (C_k + conj C_{M-k}) / 2 ;  O = (C_k - conj C_{M-k}) / (2i)
    SA[k] = make_float2(0.5f * (c1.x + c2.x), 0.5f * (c1.y - c2.y));
    SB[k] = make_float2(0.5f * (c1.y + c2.y), -0.5f * (c1.x - c2.x));
  }
}

// X_j = w_j (cr_j + i ci_j) / M, and the mean taken out by pre_real_kernel back into X_0
__global__ void post_complex_kernel(const float* __restrict__ cr, const float* __restrict__ ci, const float2* __restrict__ w,
                                    float2* __restrict__ X, unsigned count, float scale, const float* __restrict__ mean, unsigned n) {
  for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
    float2 c = w[j];
    float r = cr[j] * scale, i = ci[j] * scale;
    float2 v = make_float2(r * c.x - i * c.y, r * c.y + i * c.x);
    if (j == 0) v.x = (float)((double)v.x + (double)n * (double)*mean);
    X[j] = v;
  }
}

__global__ void post_real_kernel(const float* __restrict__ cr, const float* __restrict__ ci, const float2* __restrict__ w,
                                 float* __restrict__ x, unsigned count, float scale) {
  for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
    float2 c = w[j];
    x[j] = (cr[j] * c.x - ci[j] * c.y) * scale;
  }
}

void bluestein_destroy(Bluestein* b) {
  if (!b) return;
  cudaFree(b->w); cudaFree(b->B); cudaFree(b->ar); cudaFree(b->ai); cudaFree(b->SA); cudaFree(b->SB); cudaFree(b->part); cudaFree(b->mean);
  delete b;
}

unsigned bluestein_length(unsigned n) {
  unsigned M = 1u << 17;
  while (M < 2ull * n - 1 && M < (1u << 25)) M <<= 1;
  return M >= 2ull * n - 1 ? M : 0;
}

static int convolve(const Bluestein* b, const ps_fft_plan_t* core, cudaStream_t st) {
  const unsigned nb = b->M / 2 + 1;
  int rc = fft_r2c_core(core, b->ar, b->SA, st);
  if (!rc) rc = fft_r2c_core(core, b->ai, b->SB, st);
  if (rc) return rc;
  convolve_kernel<<<ew_grid(nb, 256), 256, 0, st>>>(b->SA, b->SB, b->B, nb);
  PS_LAUNCHED("bluestein: convolve");
  rc = fft_c2r_core(core, b->SA, b->ar, st);
  if (!rc) rc = fft_c2r_core(core, b->SB, b->ai, st);
  return rc;
}

int bluestein_create(unsigned n, const ps_fft_plan_t* core, Bluestein** out) {
  Bluestein* b = new Bluestein;
  b->n = n;
  b->M = core->n;
  const unsigned nb = b->M / 2 + 1;
  cudaError_t e = cudaMalloc(&b->w, sizeof(float2) * n);
  if (e == cudaSuccess) e = cudaMalloc(&b->B, sizeof(float2) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&b->ar, sizeof(float) * b->M);
  if (e == cudaSuccess) e = cudaMalloc(&b->ai, sizeof(float) * b->M);
  if (e == cudaSuccess) e = cudaMalloc(&b->SA, sizeof(float2) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&b->SB, sizeof(float2) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&b->part, sizeof(double) * kMeanParts);
  if (e == cudaSuccess) e = cudaMalloc(&b->mean, sizeof(float));
  if (e != cudaSuccess) {
    bluestein_destroy(b);
    return cuda_fail(e, "Error from device_malloc");
  }
  chirp_kernel<<<ew_grid(n, 256), 256>>>(b->w, n);
  g_launches.fetch_add(1);
  wrapped_chirp_kernel<<<ew_grid(b->M, 256), 256>>>(b->w, b->ar, b->ai, n, b->M);
  g_launches.fetch_add(1);
  int rc = fft_r2c_core(core, b->ar, b->SA, 0);
  if (!rc) rc = fft_r2c_core(core, b->ai, b->SB, 0);
  if (!rc) {
    merge_spectra_kernel<<<ew_grid(nb, 256), 256>>>(b->B, b->SA, b->SB, nb);
    g_launches.fetch_add(1);
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) rc = cuda_fail(e, "bluestein_create");
  }
  if (rc) {
    bluestein_destroy(b);
    return rc;
  }
  *out = b;
  return PS_OK;
}

int bluestein_r2c(const Bluestein* b, const ps_fft_plan_t* core, const float* d_in, float2* d_out, cudaStream_t st) {
  partial_sums_kernel<<<kMeanParts, 256, 0, st>>>(d_in, b->n, b->part);
  PS_LAUNCHED("bluestein: partial sums");
  mean_kernel<<<1, kMeanParts, 0, st>>>(b->part, b->n, b->mean);
  PS_LAUNCHED("bluestein: mean");
  pre_real_kernel<<<ew_grid(b->M, 256), 256, 0, st>>>(d_in, b->w, b->ar, b->ai, b->n, b->M, b->mean);
  PS_LAUNCHED("bluestein: pre");
  int rc = convolve(b, core, st);
  if (rc) return rc;
  // the two C2Rs are unnormalised (x M)
  post_complex_kernel<<<ew_grid(b->n / 2 + 1, 256), 256, 0, st>>>(b->ar, b->ai, b->w, d_out, b->n / 2 + 1, 1.0f / (float)b->M, b->mean,
                                                                  b->n);
  PS_LAUNCHED("bluestein: post");
  return PS_OK;
}

int bluestein_c2r(const Bluestein* b, const ps_fft_plan_t* core, const float2* d_in, float* d_out, cudaStream_t st) {
  pre_hermitian_kernel<<<ew_grid(b->M, 256), 256, 0, st>>>(d_in, b->w, b->ar, b->ai, b->n, b->M);
  PS_LAUNCHED("bluestein: pre");
  int rc = convolve(b, core, st);
  if (rc) return rc;
  post_real_kernel<<<ew_grid(b->n, 256), 256, 0, st>>>(b->ar, b->ai, b->w, d_out, b->n, 1.0f / (float)b->M);
  PS_LAUNCHED("bluestein: post");
  return PS_OK;
}

}  // namespace ps
