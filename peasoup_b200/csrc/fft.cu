// fft.cu -- plan management and the R2C / C2R entry points that replace cuFFT
// (reference: include/transforms/ffter.hpp:46-78, cufftPlan1d + cufftExecR2C/C2R).
//
// R2C: the n reals are packed as M = n/2 complex points, transformed by the two-pass complex
// FFT of fft.cuh, then split into the n/2+1 Hermitian bins by r2c_post_kernel.
// C2R: the Hermitian pre-process is fused into the pass-1 loader (LoadC2RPre); the inverse
// complex transform writes float pairs, so no separate post kernel is needed.
#include <cmath>
#include <vector>
#include "fft.cuh"

namespace ps {

// X[k] = E[k] + W_N^k O[k],  E = (Z[k] + conj Z[M-k])/2,  O = -i (Z[k] - conj Z[M-k])/2
__device__ __forceinline__ float2 r2c_split(float2 zk, float2 zm, float wr, float wi) {
  float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);
  float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x);
  return make_float2(er + (wr * orr - wi * oi), ei + (wr * oi + wi * orr));
}

__global__ void __launch_bounds__(256) r2c_post_kernel(const float2* __restrict__ Z, float2* __restrict__ X, unsigned M) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k <= M; k += gridDim.x * blockDim.x) {
    float2 zk = Z[k == M ? 0 : k];
    float2 zm = Z[k == 0 || k == M ? 0 : M - k];
    float wr, wi;
    sincospif((float)k / (float)M, &wi, &wr);  // W_N^k = (cos, -sin)(pi k/M)
    X[k] = r2c_split(zk, zm, wr, -wi);
  }
}

int launch_r2c_post(const ps_fft_plan_t* p, const cf* Z, float2* d_out, cudaStream_t st) {
  r2c_post_kernel<<<ew_grid(p->M + 1, 256, 2), 256, 0, st>>>(reinterpret_cast<const float2*>(Z), d_out, p->M);
  PS_LAUNCHED("r2c_post_kernel");
  return PS_OK;
}

static int pick_split(unsigned M, int* n1, int* n2) {
  // balanced split with both factors in [256, 4096]; n1 >= n2: the LONGER transform goes to pass 1,
  // so that pass 2 works on F = 32/R3 >= 4 rows at a time whenever possible -- its output runs
  // (F x 8 B) and the inter-pass tiles (tw = F) then cover whole 32-byte sectors (at 2^24 samples:
  // 4096 x 2048 instead of 2048 x 4096: measured 181 -> 122 us per acceleration trial)
  int lg = 0;
  while ((1u << lg) < M) lg++;
  if ((1u << lg) != M) return -1;
  int l2 = lg / 2, l1 = lg - l2;
  if (l2 < 8 || l1 > 12) return -1;
  *n1 = 1 << l1;
  *n2 = 1 << l2;
  return 0;
}

static std::vector<cf> roots(unsigned count, unsigned step, unsigned period) {
  // exp(-2 pi i (q*step) / period), generated in double precision
  std::vector<cf> v(count);
  for (unsigned q = 0; q < count; q++) {
    double ang = -2.0 * M_PI * (double)(((unsigned long long)q * step) % period) / (double)period;
    v[q].x = (float)cos(ang);
    v[q].y = (float)sin(ang);
  }
  return v;
}

int plan_create(unsigned n, ps_fft_plan_t** out, int max_batch) {
  if (max_batch < 1 || max_batch > kMaxBatch) return invalid("plan_create: bad batch");
  if (n < (1u << 17) || n > (1u << 25) || (n & (n - 1)))
    return invalid("ps_fft_plan_create: n must be a power of two with 2^17 <= n <= 2^25");
  ps_fft_plan_t* p = new ps_fft_plan_t();
  p->n = n;
  p->M = n / 2;
  p->max_batch = max_batch;
  if (pick_split(p->M, &p->n1, &p->n2)) {
    delete p;
    return invalid("ps_fft_plan_create: unsupported length");
  }
  p->r1 = p->n1 / 256;
  p->r2 = p->n2 / 256;
  cudaError_t e = cudaGetDevice(&p->device);
  if (e != cudaSuccess) {
    delete p;
    return cuda_fail(e, "cudaGetDevice");
  }
  unsigned na = p->M / 4096 ? p->M / 4096 : 1;
  std::vector<cf> w = roots(4096, 1, 4096), ta = roots(na, 4096, p->M), tb = roots(4096, 1, p->M);
  std::vector<cf> tc = roots(na + 1, 4096, 2 * p->M), td = roots(4096, 1, 2 * p->M);
  p->W4096 = p->TA = p->TB = p->Yt = p->Z = p->TC = p->TD = nullptr;
  e = cudaMalloc(&p->W4096, sizeof(cf) * 4096);
  if (e == cudaSuccess) e = cudaMalloc(&p->TA, sizeof(cf) * na);
  if (e == cudaSuccess) e = cudaMalloc(&p->TB, sizeof(cf) * 4096);
  if (e == cudaSuccess) e = cudaMalloc(&p->Yt, sizeof(cf) * (size_t)p->M * max_batch);
  if (e == cudaSuccess) e = cudaMalloc(&p->Z, sizeof(cf) * (size_t)p->M * max_batch);
  if (e == cudaSuccess) e = cudaMalloc(&p->TC, sizeof(cf) * (na + 1));
  if (e == cudaSuccess) e = cudaMalloc(&p->TD, sizeof(cf) * 4096);
  if (e == cudaSuccess) e = cudaMemcpy(p->TC, tc.data(), sizeof(cf) * (na + 1), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(p->TD, td.data(), sizeof(cf) * 4096, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(p->W4096, w.data(), sizeof(cf) * 4096, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(p->TA, ta.data(), sizeof(cf) * na, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(p->TB, tb.data(), sizeof(cf) * 4096, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(p->W4096); cudaFree(p->TA); cudaFree(p->TB); cudaFree(p->Yt); cudaFree(p->Z); cudaFree(p->TC); cudaFree(p->TD);
    delete p;
    return cuda_fail(e, "Error from device_malloc");
  }
  *out = p;
  return PS_OK;
}

void plan_destroy(ps_fft_plan_t* p) {
  if (!p) return;
  cudaFree(p->W4096); cudaFree(p->TA); cudaFree(p->TB); cudaFree(p->Yt); cudaFree(p->Z); cudaFree(p->TC); cudaFree(p->TD);
  delete p;
}

}  // namespace ps

using namespace ps;

// power-of-two R2C / C2R on a core plan (also the engine under bluestein.cu)
namespace ps {
int fft_r2c_core(const ps_fft_plan_t* p, const float* d_in, float2* d_out, cudaStream_t st) {
  LoadPackedReal ld{reinterpret_cast<const float2*>(d_in), 0, 0, 0, nullptr};
  int rc = launch_pass1<false>(p, ld, st);
  if (rc) return rc;
  StoreNatural sn{reinterpret_cast<float2*>(p->Z), 0, 0, 0, nullptr};
  rc = launch_pass2<false>(p, sn, st);
  if (rc) return rc;
  return launch_r2c_post(p, p->Z, d_out, st);
}

int fft_c2r_core(const ps_fft_plan_t* p, const float2* d_in, float* d_out, cudaStream_t st) {
  LoadC2RPre ld{d_in, p->M, 0, 0};
  int rc = launch_pass1<true>(p, ld, st);
  if (rc) return rc;
  StoreNatural sn{reinterpret_cast<float2*>(d_out), 0, 0, 0, nullptr};
  return launch_pass2<true>(p, sn, st);
}
}  // namespace ps

struct ps_fft_plan {
  ps_fft_plan_t* core = nullptr;  // transform of length n, or of the chirp-z convolution length when blu is set
  Bluestein* blu = nullptr;       // lengths that are not a power of two in [2^17, 2^25]
};

extern "C" {

int ps_fft_plan_create(unsigned n, ps_fft_plan** out) {
  if (!out) return invalid("ps_fft_plan_create: null output");
  *out = nullptr;
  const bool pow2 = n >= (1u << 17) && n <= (1u << 25) && !(n & (n - 1));
  unsigned core_n = n;
  if (!pow2) {
    if (n < 2 || n > (1u << 24)) return invalid("ps_fft_plan_create: n must be in [2, 2^24], or a power of two up to 2^25");
    core_n = bluestein_length(n);
  }
  ps_fft_plan* plan = new ps_fft_plan;
  int rc = plan_create(core_n, &plan->core, 1);
  if (!rc && !pow2) rc = bluestein_create(n, plan->core, &plan->blu);
  if (rc) {
    plan_destroy(plan->core);
    delete plan;
    return rc;
  }
  *out = plan;
  return PS_OK;
}

void ps_fft_plan_destroy(ps_fft_plan* plan) {
  if (!plan) return;
  bluestein_destroy(plan->blu);
  plan_destroy(plan->core);
  delete plan;
}

int ps_fft_r2c(ps_fft_plan* plan, const float* d_in, ps_complex* d_out, void* stream) {
  if (!plan) return invalid("ps_fft_r2c: null plan");
  if (plan->blu) return bluestein_r2c(plan->blu, plan->core, d_in, reinterpret_cast<float2*>(d_out), S(stream));
  return fft_r2c_core(plan->core, d_in, reinterpret_cast<float2*>(d_out), S(stream));
}

int ps_fft_c2r(ps_fft_plan* plan, const ps_complex* d_in, float* d_out, void* stream) {
  if (!plan) return invalid("ps_fft_c2r: null plan");
  if (plan->blu) return bluestein_c2r(plan->blu, plan->core, reinterpret_cast<const float2*>(d_in), d_out, S(stream));
  return fft_c2r_core(plan->core, reinterpret_cast<const float2*>(d_in), d_out, S(stream));
}

}  // extern "C"
