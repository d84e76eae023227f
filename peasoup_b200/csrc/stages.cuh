// stages.cuh -- per-element arithmetic of the search path as inline device functions, shared
// between the staged entry points (stages.cu) and the fused search kernels (search.cu) so that
// both produce identical bits.  Each function cites the reference lines whose arithmetic
// (operand types, evaluation order, FMA contraction as nvcc emits it for the reference) it
// restates; the code itself is written from that description, not copied.
#pragma once
#include <cuComplex.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ps {

constexpr unsigned kMaxStatParts = 1184;  // 148 SMs x 8 blocks

double accel_fact(float a, float tsamp);
int launch_sums(const float* d, size_t n, double* d_partials, int* nparts, cudaStream_t st);
int launch_stats(const float* d, size_t n, double* d_partials, float* d_out3, cudaStream_t st);
// second half of launch_stats: {mean, rms, std} of n values from nparts (sum, sum of squares) pairs
int launch_finalize_stats(const double* d_partials, int nparts, size_t n, float* d_out3, cudaStream_t st);
int launch_running_median(const float* d_pspec, unsigned nbins, double bin_width, float b5, float b25, float* d_median,
                          float* d_work, cudaStream_t st);

// kernels.cu:314-317: rn( i + (i*af)*(i-size) ), the outer multiply-add fused (DFMA in the
// reference's SASS), ties to even.
__device__ __forceinline__ unsigned long long resample_index_II(double af, double size, unsigned long long i) {
  double di = (double)i;
  return __double2ull_rn(__fma_rn(__dmul_rn(di, af), __dadd_rn(di, -size), di));
}

// kernels.cu:308-311: rn( i + af*((i-h)*(i-h) - h*h) ); nvcc fuses (i-h)*(i-h) - h*h and the
// outer af*(..) + i.
__device__ __forceinline__ unsigned long long resample_index_I(double af, double h, unsigned long long i) {
  double di = (double)i;
  double d = __dadd_rn(di, -h);
  double hh = __dmul_rn(h, h);
  double inner = __fma_rn(d, d, -hh);
  return __double2ull_rn(__fma_rn(af, inner, di));
}

// kernels.cu:223-224
__device__ __forceinline__ float amplitude_rsqrt(float re, float im) {
  float z = __fmaf_rn(re, re, __fmul_rn(im, im));
  return __fmul_rn(z, rsqrtf(z));
}

// kernels.cu:246-249
__device__ __forceinline__ float interbin_amplitude(float re, float im, float re_l, float im_l) {
  float ampsq = __fmaf_rn(re, re, __fmul_rn(im, im));
  float dr = __fsub_rn(re, re_l), di = __fsub_rn(im, im_l);
  float diff = __fmul_rn(__fmaf_rn(dr, dr, __fmul_rn(di, di)), 0.5f);
  return sqrtf(fmaxf(ampsq, diff));
}

// kernels.cu:476-479
__device__ __forceinline__ float normalise_value(float v, float mean, float sigma) {
  return __fdiv_rn(__fsub_rn(v, mean), sigma);
}

// kernels.cu:40-95.  (int)(idx*m/2^n + 0.5) in double == (idx*m + 2^(n-1)) >> n for idx >= 0.
// Accumulation order and the double-precision scale factors follow the reference.
__device__ __forceinline__ void harmonic_folds(const float* __restrict__ in, int idx, int nfolds, float* f) {
  float val = __ldg(in + idx);
  if (nfolds > 0) {
    val += __ldg(in + ((idx + 1) >> 1));
    f[0] = (float)((double)val * rsqrt(2.0));
  }
  if (nfolds > 1) {
    val += __ldg(in + ((idx * 3 + 2) >> 2));
    val += __ldg(in + ((idx + 2) >> 2));
    f[1] = (float)((double)val * 0.5);
  }
  if (nfolds > 2) {
#pragma unroll
    for (int m = 1; m < 8; m += 2) val += __ldg(in + ((idx * m + 4) >> 3));
    f[2] = (float)((double)val * rsqrt(8.0));
  }
  if (nfolds > 3) {
#pragma unroll
    for (int m = 1; m < 16; m += 2) val += __ldg(in + ((idx * m + 8) >> 4));
    f[3] = (float)((double)val * 0.25);
  }
  if (nfolds > 4) {
#pragma unroll
    for (int m = 1; m < 32; m += 2) val += __ldg(in + ((idx * m + 16) >> 5));
    f[4] = (float)((double)val * rsqrt(32.0));
  }
}

// Medians as min/max selection networks; for non-NaN input any correct selection returns the
// same element as the comparison trees at kernels.cu:875-929.
__device__ __forceinline__ float median_of_3(float a, float b, float c) {
  return fmaxf(fminf(a, b), fminf(fmaxf(a, b), c));
}
__device__ __forceinline__ float median_of_4(float a, float b, float c, float d) {
  float lo = fmaxf(fminf(a, b), fminf(c, d));
  float hi = fminf(fmaxf(a, b), fmaxf(c, d));
  return 0.5f * (lo + hi);
}
__device__ __forceinline__ float median_of_5(float a, float b, float c, float d, float e) {
  float f = fmaxf(fminf(a, b), fminf(c, d));  // drops the smallest of the first four
  float g = fminf(fmaxf(a, b), fmaxf(c, d));  // drops the largest of the first four
  return median_of_3(e, f, g);
}

// kernels.cu:991-995 (separate multiply and add in the reference's SASS)
__device__ __forceinline__ float stretch_value(const float* __restrict__ in, size_t in_count, float step, unsigned i) {
  float x = __fmul_rn((float)i, step);
  unsigned j = (unsigned)x;
  float frac = __fsub_rn(x, (float)j);
  float a = in[j];
  if (frac > 1e-5f && (size_t)j + 1 < in_count) {
    float b = in[j + 1];
    return __fadd_rn(a, __fmul_rn(frac, __fsub_rn(b, a)));
  }
  return a;
}

// dereddener.hpp:55-61 splice: [0,pos5) 5-bin level, [pos5,pos25) 25-bin, [pos25,nbins) 125-bin
__device__ __forceinline__ float spliced_median_value(const float* m5, size_t n5, float s5, const float* m25, size_t n25,
                                                      float s25, const float* m125, size_t n125, float s125, int pos5,
                                                      int pos25, unsigned i) {
  if ((int)i >= pos25) return stretch_value(m125, n125, s125, i);
  if ((int)i >= pos5) return stretch_value(m25, n25, s25, i);
  return stretch_value(m5, n5, s5, i);
}

// kernels.cu:1019-1022
__device__ __forceinline__ float2 deredden_value(float2 c, float f, unsigned idx) {
  if (idx < 5) return make_float2(0.f, 0.f);
  return cuCdivf(c, make_cuFloatComplex(f, 0.f));
}

// kernels.cu:1045-1056; returns false when nothing is to be written
__device__ __forceinline__ bool zap_range(float freq, float width, float bin_width, unsigned nbins, int* lo, int* hi) {
  int low_bin = __float2int_rd(__fdiv_rn(__fsub_rn(freq, width), bin_width));
  int high_bin = __float2int_ru(__fdiv_rn(__fadd_rn(freq, width), bin_width));
  if (low_bin < 0) low_bin = 0;
  if ((unsigned)low_bin >= nbins) return false;
  if ((unsigned)high_bin >= nbins) high_bin = (int)nbins - 1;  // int vs unsigned compare, as the reference
  *lo = low_bin;
  *hi = high_bin;
  return high_bin > low_bin;
}

}  // namespace ps
