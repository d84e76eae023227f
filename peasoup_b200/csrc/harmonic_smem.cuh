// harmonic_smem.cuh -- harmonic sums + peaks with the gather operands staged in shared memory by
// TMA bulk copies (included by search.cu after hp_emit_slow).
//
// For a tile of W consecutive output bins [k0, k0+W) the operands of fraction m/2^s are the
// CONTIGUOUS input range [p(k0), p(k0+W-1)], p(k) = (k m + 2^(s-1)) >> s: fifteen ranges
// (1/2; 1/4, 3/4; odd/8; odd/16) of 7.5 W words in total, plus the tile itself.  One thread issues
// sixteen cp.async.bulk copies (16-byte aligned starts, completion on an mbarrier); the sums are
// then gathered from shared memory with the same lane layout as harmonic_peaks_kernel (lane l
// holds bins base + l + 32 i, so a warp's 32 addresses span < 32 consecutive words: conflict-free,
// equal addresses broadcast).  The L2 sees sixteen streaming reads instead of 16 scattered 4-byte
// loads per bin, and the latency of the gathers drops from an L2 round trip to a shared-memory
// access.  Six 128-thread CTAs (1024-bin tiles, 35 KB of staging each) per SM take turns loading
// and summing (chain us/trial for tiles of 512 / 1024 / 1536 / 2048 bins with 12 / 6 / 4 / 3 CTAs per
// SM: 58.0 / 55.4 / 56.2 / 57.0); a second staging buffer per CTA (kHsBufs = 2) was measured slower.
//
// Both harmonic kernels pull 8.5 words per bin out of the L2 and run close to its throughput cap
// (143 MB per 2^22-bin spectrum in 14.1 us = 10 TB/s against ~12 TB/s); this one is ~5 % faster
// than harmonic_peaks_kernel (14.1 vs 14.8 us per spectrum inside the chain).  Staging a
// half-precision copy of the spectrum instead (half the L2 bytes, filter bounds lowered by the
// rounding bound) was measured too: 12.9 us, but the spectrum kernel pays 2.2 us to write the copy.
//
// The filter / exact re-evaluation split is the one of harmonic_peaks_kernel (flagged bins go
// through hp_emit_slow on the spectrum in global memory), so the output is bit-identical.
// Requires the spectra to be padded by >= 4 readable floats (the pipeline's own buffers are).
#pragma once

namespace ps {

constexpr int kHsW = 1024;   // output bins per tile; W/8 threads (8 bins per lane)
constexpr int kHsCtas = 6;   // CTAs per SM (35 KB of staging and 10k registers each): enough tiles in flight to hide the copies
constexpr int kHsBufs = 1;   // staging buffers per CTA (2 = the next tile's copies fly while this one is summed)

// fractions in the accumulation order of hp_chunk: index, numerator, shift
#define PS_HS_FRACTIONS(X)                                                                              \
  X(0, 1, 1) X(1, 3, 2) X(2, 1, 2) X(3, 1, 3) X(4, 3, 3) X(5, 5, 3) X(6, 7, 3) X(7, 1, 4) X(8, 3, 4) \
  X(9, 5, 4) X(10, 7, 4) X(11, 9, 4) X(12, 11, 4) X(13, 13, 4) X(14, 15, 4)
__host__ __device__ constexpr int hs_nfrac(int nf) { return nf <= 0 ? 0 : nf == 1 ? 1 : nf == 2 ? 3 : nf == 3 ? 7 : 15; }
__host__ __device__ constexpr int hs_m(int f) {
  constexpr int t[15] = {1, 3, 1, 1, 3, 5, 7, 1, 3, 5, 7, 9, 11, 13, 15};
  return t[f];
}
__host__ __device__ constexpr int hs_s(int f) {
  constexpr int t[15] = {1, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 4, 4, 4, 4};
  return t[f];
}
// words reserved for fraction f: the range is at most (W m >> s) + 2 long, + 3 for the aligned start, rounded to 4
__host__ __device__ constexpr int hs_words(int f) { return ((((kHsW * hs_m(f)) >> hs_s(f)) + 2 + 3) + 3) & ~3; }
constexpr int kHsThreads = kHsW / 8;
__host__ __device__ constexpr int hs_offset(int f) { return f == 0 ? kHsW : hs_offset(f - 1) + hs_words(f - 1); }  // region 0 = the tile itself
__host__ __device__ constexpr int hs_total(int nf) { return hs_nfrac(nf) == 0 ? kHsW : hs_offset(hs_nfrac(nf) - 1) + hs_words(hs_nfrac(nf) - 1); }

__device__ const int c_hs_m[15] = {1, 3, 1, 1, 3, 5, 7, 1, 3, 5, 7, 9, 11, 13, 15};
__device__ const int c_hs_s[15] = {1, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 4, 4, 4, 4};
__device__ const int c_hs_off[15] = {hs_offset(0), hs_offset(1), hs_offset(2),  hs_offset(3),  hs_offset(4),
                                     hs_offset(5), hs_offset(6), hs_offset(7),  hs_offset(8),  hs_offset(9),
                                     hs_offset(10), hs_offset(11), hs_offset(12), hs_offset(13), hs_offset(14)};

// read index of bin k for the fraction m / 2^s (kernels.cu:45-93: (int)(k m / 2^s + 0.5))
__device__ __forceinline__ int hs_p(int k, int m, int s) { return (int)(((long long)k * m + (1 << (s - 1))) >> s); }

template <int NF>
__global__ void __launch_bounds__(kHsThreads, kHsCtas) harmonic_peaks_smem_kernel(const float* __restrict__ Pall, int nbins, size_t pstride,
                                                                            LevelRanges r, LevelScales ls, float thresh, int acc_idx0,
                                                                            PeakSink sink, int ntiles, int nbatch) {
  static_assert(NF >= 1 && NF <= 4, "shared-memory staging is sized for up to 4 folds");
  constexpr int NFR = hs_nfrac(NF);
  constexpr int BUF = hs_total(NF);  // words per staging buffer
  extern __shared__ __align__(128) float sm_all[];
  __shared__ __align__(8) unsigned long long bar[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwork = ntiles * nbatch;

  // threads 0..NFR: thread 0 fetches the tile of work item w, thread f+1 the range of fraction f,
  // into staging buffer `buf`; each arrives on the buffer's mbarrier with its own byte count
  auto fetch = [&](int w, int buf) {
    const int b = w / ntiles, tile = w - b * ntiles;
    const float* __restrict__ P = Pall + (size_t)b * pstride;
    float* sm = sm_all + buf * BUF;
    const int k0 = tile * kHsW;
    const int klast = min(k0 + kHsW, nbins) - 1;
    int first = k0, last = klast, off = 0;
    if (threadIdx.x > 0) {
      const int f = threadIdx.x - 1;
      first = hs_p(k0, c_hs_m[f], c_hs_s[f]) & ~3;
      last = hs_p(klast, c_hs_m[f], c_hs_s[f]);
      off = c_hs_off[f];
    }
    const unsigned bytes = (unsigned)(((last - first + 1) + 3) & ~3) * 4u;
    fence_proxy_async();  // earlier generic-proxy reads of this buffer are ordered before the async writes
    mbar_expect_tx(&bar[buf], bytes);
    bulk_g2s(sm + off, P + first, bytes, &bar[buf]);
  };

  if (threadIdx.x == 0) {
    mbar_init(&bar[0], NFR + 1);
    mbar_init(&bar[1], NFR + 1);
    fence_proxy_async();
  }
  __syncthreads();
  if (kHsBufs == 2 && (int)threadIdx.x <= NFR && (int)blockIdx.x < nwork) fetch(blockIdx.x, 0);
  unsigned parity[2] = {0, 0};
  int it = 0;
  for (int work = blockIdx.x; work < nwork; work += gridDim.x, it++) {
    const int buf = kHsBufs == 2 ? (it & 1) : 0;
    // (the buffer written here was released by the barrier at the end of the previous iteration)
    if (kHsBufs == 2) {
      if ((int)threadIdx.x <= NFR && work + (int)gridDim.x < nwork) fetch(work + gridDim.x, buf ^ 1);
    } else if ((int)threadIdx.x <= NFR) {
      fetch(work, 0);
    }
    const int b = work / ntiles, tile = work - b * ntiles;
    const float* __restrict__ P = Pall + (size_t)b * pstride;
    const float* sm = sm_all + buf * BUF;
    const int k0 = tile * kHsW;
#define PS_HS_LO(I, M_, S_) const int lo##I = hs_p(k0, M_, S_) & ~3;
    PS_HS_FRACTIONS(PS_HS_LO)
#undef PS_HS_LO
    mbar_wait(&bar[buf], parity[buf]);
    parity[buf] ^= 1u;

    const int base = k0 + warp * 256 + lane;  // lane's bins: base + 32 i, i < 8
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = sm[warp * 256 + lane + 32 * i];
    auto emit = [&](int nh) {
      float mx = acc[0];
#pragma unroll
      for (int i = 1; i < 8; i++) mx = fmaxf(mx, acc[i]);
      if (!__any_sync(0xffffffffu, mx > ls.lo[nh])) return;
      unsigned mask = 0;
#pragma unroll
      for (int i = 0; i < 8; i++) mask |= (acc[i] > ls.lo[nh]) ? (1u << i) : 0u;
      hp_emit_slow(P, nbins, base, lane, mask, nh, NF, r, thresh, acc_idx0 + b, sink);
    };
    emit(0);
#define PS_HS_ADD(I, M_, S_)                                                       \
  if ((I) < NFR) {                                                                 \
    constexpr int off_ = hs_offset(I);                                             \
    const float* q_ = sm + off_ + (hs_p(base, M_, S_) - lo##I);                    \
    _Pragma("unroll") for (int i = 0; i < 8; i++) acc[i] += q_[i * ((32 * (M_)) >> (S_))]; \
    if ((I) == 0) emit(1);                                                         \
    if ((I) == 2) emit(2);                                                         \
    if ((I) == 6) emit(3);                                                         \
    if ((I) == 14) emit(4);                                                        \
  }
    PS_HS_FRACTIONS(PS_HS_ADD)
#undef PS_HS_ADD
    __syncthreads();  // all warps are done with this buffer: the next iteration's fetch may overwrite it
  }
}

}  // namespace ps
