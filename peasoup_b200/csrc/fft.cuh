// fft.cuh -- two-pass (four-step) FFT kernels built on fft_core.cuh, shared between the
// plan API (fft.cu) and the fused search path (search.cu).
//
// Complex transform of M = n1*n2 points, in[j1*n2 + j2] -> out[k1 + n1*k2]:
//   pass 1 (one CTA = F1 adjacent columns j2): length-n1 FFT over j1, multiply by W_M^(j2*k1),
//          store transposed  Yt[j2*n1 + k1]                    (contiguous 64 B runs per column)
//   pass 2 (one CTA = F2 adjacent rows k1):    length-n2 FFT over j2, Storer epilogue at
//          natural index k = k1 + n1*k2                        (32 B aligned runs)
// Loaders/Storers are functors so that the u8->float conversion, the acceleration resampling
// gather and the C2R pre-processing fuse into the pass-1 load stage and the real-FFT
// post-processing into later passes.
#pragma once
#include <cuda_runtime.h>
#include "fft_core.cuh"
#include "common.cuh"
#include "stages.cuh"

namespace ps {
using psfft::cf;

struct FftTables {
  const cf* W4096;  // exp(-2 pi i q/4096)
  const cf* TA;     // exp(-2 pi i 4096 q / M), q < max(1, M/4096)
  const cf* TB;     // exp(-2 pi i q / M), q < 4096
};

struct ps_fft_plan_t {
  int device;
  unsigned n;   // real length
  unsigned M;   // n/2 complex points
  int n1, n2, r1, r2;
  cf* W4096;
  cf* TA;
  cf* TB;
  cf* Yt;  // M complex: transposed intermediate between the passes
  cf* Z;   // M complex: packed-complex spectrum before the real post-process
  FftTables tables() const { return FftTables{W4096, TA, TB}; }
};

// W_M^p for p < M from the two-level table (one complex multiply)
__device__ __forceinline__ cf twiddle_M(const FftTables& t, unsigned p) {
  return psfft::cmul(psfft::ldg_cf(&t.TA[p >> 12]), psfft::ldg_cf(&t.TB[p & 4095u]));
}

// ------------------------------------------------------------------ pass-1 loaders
// complex element e = idx*n2 + c0 + f of the packed series
struct LoadPackedReal {  // z[e] = (x[2e], x[2e+1])
  const float2* x;
  int n2, c0;
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    float2 v = __ldg(&x[(size_t)idx * n2 + c0 + f]);
    return psfft::mk(v.x, v.y);
  }
};

struct LoadResampled {  // z[e] = (x[map(2e)], x[map(2e+1)])   (kernels.cu:335-346 fused)
  const float* x;
  double af, size;
  int n2, c0;
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    unsigned long long i = 2ull * ((unsigned long long)idx * n2 + c0 + f);
    float a = __ldg(&x[resample_index_II(af, size, i)]);
    float b = __ldg(&x[resample_index_II(af, size, i + 1)]);
    return psfft::mk(a, b);
  }
};

struct LoadU8Padded {  // timeseries.hpp:335-344 + pipeline_multi.cu:160-163 fused
  const unsigned char* b;
  const float* pad;  // device scalar: mean of the first nsamps samples
  unsigned nsamps;   // valid samples (<= size)
  int n2, c0;
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    size_t i = 2 * ((size_t)idx * n2 + c0 + f);
    float p = 0.f;
    if (i + 1 >= nsamps) p = __ldg(pad);
    float a = i < nsamps ? (float)__ldg(&b[i]) : p;
    float c = i + 1 < nsamps ? (float)__ldg(&b[i + 1]) : p;
    return psfft::mk(a, c);
  }
};

// C2R pre-process: Z'[k] = (X[k] + conj X[M-k]) + i (X[k] - conj X[M-k]) W_N^-k
struct LoadC2RPre {
  const float2* X;  // M+1 bins
  unsigned M;
  int n2, c0;
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    unsigned k = (unsigned)idx * n2 + c0 + f;
    float2 xa = __ldg(&X[k]);
    float2 xb = __ldg(&X[M - k]);
    float wr, wi;
    sincospif((float)k / (float)M, &wi, &wr);  // W_N^-k = (cos, +sin)(pi k / M)
    float sr = xa.x + xb.x, si = xa.y - xb.y;
    float dr = xa.x - xb.x, di = xa.y + xb.y;
    float tr = dr * wr - di * wi, ti = dr * wi + di * wr;
    return psfft::mk(sr - ti, si + tr);
  }
};

template <bool INV>
struct StorePass1 {
  cf* Yt;
  FftTables t;
  int n1, c0;
  __device__ __forceinline__ void operator()(int f, int k, cf v) const {
    unsigned j2 = (unsigned)(c0 + f);
    cf w = psfft::dirw<INV>(twiddle_M(t, j2 * (unsigned)k));
    v = psfft::cmul(v, w);
    reinterpret_cast<float2*>(Yt)[(size_t)j2 * n1 + k] = make_float2(v.x, v.y);
  }
};

struct LoadPass2 {
  const cf* Yt;
  int n1, r0;
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    float2 v = reinterpret_cast<const float2*>(Yt)[(size_t)idx * n1 + r0 + f];
    return psfft::mk(v.x, v.y);
  }
};

struct StoreNatural {  // out[k1 + n1*k2]
  float2* out;
  int n1, r0;
  __device__ __forceinline__ void operator()(int f, int k, cf v) const {
    out[(size_t)k * n1 + r0 + f] = make_float2(v.x, v.y);
  }
};

template <int R3, bool INV, class Loader>
__global__ void __launch_bounds__(512, 1) fft_pass1_kernel(Loader proto, cf* Yt, FftTables t, int n1, int n2) {
  using C = psfft::CtaFft<R3, INV>;
  extern __shared__ float smem[];
  const int ntiles = n2 / C::F;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    Loader ld = proto;
    ld.c0 = tile * C::F;
    ld.n2 = n2;
    StorePass1<INV> st{Yt, t, n1, tile * C::F};
    C::phaseA(threadIdx.x, smem, t.W4096, ld);
    __syncthreads();
    C::phaseB(threadIdx.x, smem, t.W4096, st);
    __syncthreads();
    C::phaseC(threadIdx.x, smem, st);
  }
}

template <int R3, bool INV, class Storer>
__global__ void __launch_bounds__(512, 1) fft_pass2_kernel(const cf* Yt, Storer proto, FftTables t, int n1, int n2) {
  using C = psfft::CtaFft<R3, INV>;
  extern __shared__ float smem[];
  const int ntiles = n1 / C::F;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    LoadPass2 ld{Yt, n1, tile * C::F};
    Storer st = proto;
    st.r0 = tile * C::F;
    st.n1 = n1;
    C::phaseA(threadIdx.x, smem, t.W4096, ld);
    __syncthreads();
    C::phaseB(threadIdx.x, smem, t.W4096, st);
    __syncthreads();
    C::phaseC(threadIdx.x, smem, st);
  }
}

// ------------------------------------------------------------------ host-side launchers
// opt in to > 48 KB dynamic shared memory once per (kernel instantiation, device); `done` must be
// a static of the CALLING template (one per kernel), not of a helper keyed on the pointer type
template <class K>
inline int configure_smem(K kernel, size_t bytes, int device, bool* done) {
  if (!done[device & 63]) {
    PS_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes),
            "cudaFuncSetAttribute");
    done[device & 63] = true;
  }
  return PS_OK;
}

inline int fft_grid(int ntiles) { return ntiles < 148 ? ntiles : 148; }

template <int R3, bool INV, class Loader>
inline int launch_pass1_r(const ps_fft_plan_t* p, const Loader& ld, cudaStream_t st) {
  using C = psfft::CtaFft<R3, INV>;
  auto k = fft_pass1_kernel<R3, INV, Loader>;
  static bool done[64] = {};
  int rc = configure_smem(k, C::SMEM_BYTES, p->device, done);
  if (rc) return rc;
  k<<<fft_grid(p->n2 / C::F), C::THREADS, C::SMEM_BYTES, st>>>(ld, p->Yt, p->tables(), p->n1, p->n2);
  PS_LAUNCHED("fft_pass1_kernel");
  return PS_OK;
}

template <int R3, bool INV, class Storer>
inline int launch_pass2_r(const ps_fft_plan_t* p, const Storer& stor, cudaStream_t st) {
  using C = psfft::CtaFft<R3, INV>;
  auto k = fft_pass2_kernel<R3, INV, Storer>;
  static bool done[64] = {};
  int rc = configure_smem(k, C::SMEM_BYTES, p->device, done);
  if (rc) return rc;
  k<<<fft_grid(p->n1 / C::F), C::THREADS, C::SMEM_BYTES, st>>>(p->Yt, stor, p->tables(), p->n1, p->n2);
  PS_LAUNCHED("fft_pass2_kernel");
  return PS_OK;
}

#define PS_DISPATCH_R(r, CALL)                  \
  switch (r) {                                  \
    case 1: return CALL(1);                     \
    case 2: return CALL(2);                     \
    case 4: return CALL(4);                     \
    case 8: return CALL(8);                     \
    case 16: return CALL(16);                   \
    default: return invalid("unsupported FFT radix plan"); \
  }

template <bool INV, class Loader>
inline int launch_pass1(const ps_fft_plan_t* p, const Loader& ld, cudaStream_t st) {
#define PS_CALL1(R) launch_pass1_r<R, INV, Loader>(p, ld, st)
  PS_DISPATCH_R(p->r1, PS_CALL1)
#undef PS_CALL1
}

template <bool INV, class Storer>
inline int launch_pass2(const ps_fft_plan_t* p, const Storer& stor, cudaStream_t st) {
#define PS_CALL2(R) launch_pass2_r<R, INV, Storer>(p, stor, st)
  PS_DISPATCH_R(p->r2, PS_CALL2)
#undef PS_CALL2
}

// Z (M packed-complex bins in p->Z) -> X[0..M] (fft.cu)
int launch_r2c_post(const ps_fft_plan_t* p, const cf* Z, float2* d_out, cudaStream_t st);

}  // namespace ps
