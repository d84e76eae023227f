// harmonic_rows.cuh -- harmonic sums + peaks with the gather operands staged in shared memory by TMA
// bulk copies, a lane owning SIXTEEN CONSECUTIVE bins (included by search.cu after hp_emit_rows).
//
// For a tile of W consecutive output bins [k0, k0+W) the operands of the fraction m/2^s are the
// CONTIGUOUS input range [p(k0), p(k0+W-1)], p(k) = (k m + 2^(s-1)) >> s: fifteen ranges (1/2; 1/4,
// 3/4; odd/8; odd/16) of 7.5 W words in total.  Fifteen threads issue one cp.async.bulk copy each
// (completion on an mbarrier): the L2 sees fifteen streaming reads per tile instead of 15 scattered
// 4-byte loads per bin, and the gathers read shared memory.  Eight 64-thread CTAs (1024-bin tiles,
// 26.3 KB of staging each) per SM take turns copying and summing (chain us / trial for tiles of
// 512 / 1024 / 2048 bins with 14 / 7 / 3 CTAs per SM: 51.7 / 45.9 / 47.1; six CTAs of 1024: 46.1).
// The five smallest fractions -- 1/2, 1/4, 1/8, 1/16 and 3/16, 1.1 of the 7.5 words per bin -- are not
// staged: a lane reads their 2..9 consecutive words from global memory together with its own bins
// (23 registers), which is what lets the eighth CTA fit (chain 45.0 -> 44.6 us / trial; the same with
// seven CTAs: 45.0).
//
// With bins k0 + 16 j + i (i < 16) in lane j of a tile that starts at a multiple of 1024, the read
// index of the fraction m / 2^s is  p = P0 + j M16 + c_i,  P0 = k0 m >> s (a multiple of 4),
// M16 = 16 m / 2^s,  c_i = (i M16 + 8) >> 4 -- the c_i are compile-time constants.  A lane therefore
// needs the max(c) + 1 <= 16 CONSECUTIVE staged words from j M16 on, and every one of its sixteen
// sums takes its operand from a fixed register: 6 loaded words per bin instead of 16, no per-bin
// address, and vector loads where the lane stride allows (8-byte loads for M16 = 2, 6, 10, 14,
// 16-byte loads for 4, 8, 12; an odd M16 is conflict-free with 4-byte loads).  The accumulation order per
// bin is that of harmonic_peaks_kernel, so the sums -- and the peaks -- are bit-identical.
//
// The tile itself is not staged: a lane's own sixteen bins come straight from global memory with
// four 16-byte loads issued BEFORE the wait for the staged ranges (7.5 instead of 8.5 staged words per bin;
// from shared memory those loads would be 4-way bank-conflicted, the lane stride being 64 bytes).
// Measured alternatives: the tile staged by a sixteenth bulk copy and read with those conflicted loads
// (six CTAs per SM): 13.7 us per spectrum against 11.9, although the DRAM reads drop from 34 MB to the
// ideal 17.6 MB per spectrum (the loads and the copies of neighbouring tiles fetch the same lines
// separately) -- the copy engine, not DRAM, is what the eighth-and-a-half word would queue on; an L2
// bulk prefetch of the CTA's next tile: no change; the own bins requested one work item ahead in
// registers: no change; two groups of ranges with a barrier each, the first re-requested during the
// sums over the second: slower (10.5 against 9.9 us per spectrum of zeros).
#pragma once

namespace ps {

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
__device__ const int c_hs_m[15] = {1, 3, 1, 1, 3, 5, 7, 1, 3, 5, 7, 9, 11, 13, 15};
__device__ const int c_hs_s[15] = {1, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 4, 4, 4, 4};

// read index of bin k for the fraction m / 2^s (kernels.cu:45-93: (int)(k m / 2^s + 0.5))
__device__ __forceinline__ int hs_p(int k, int m, int s) { return (int)(((long long)k * m + (1 << (s - 1))) >> s); }

#ifdef PS_PROFILING
__device__ int g_hs_probe;  // 1: no staging copies after a CTA's first tile, 2: copies but no sums
#endif

constexpr int kHrW = 1024;            // bins per tile
constexpr int kHrThreads = kHrW / 16;  // 64
constexpr int kHrCtas = 8;            // CTAs per SM (26.3 KB of staging each)

__host__ __device__ constexpr int hr_m16(int f) { return hs_m(f) << (4 - hs_s(f)); }
__host__ __device__ constexpr int hr_cmax(int m16) { return (15 * m16 + 8) >> 4; }
// words reserved for fraction f: W M16 / 16 + 1, rounded up to a multiple of 4 (16-byte copies and 16-byte aligned regions);
// fractions in DMASK are not staged (their few words per lane come straight from global memory)
template <int W, int DMASK = 0>
__host__ __device__ constexpr int hr_words(int f) { return ((DMASK >> f) & 1) ? 0 : (((W * hr_m16(f)) >> 4) + 1 + 3) & ~3; }
template <int W, int DMASK = 0>
__host__ __device__ constexpr int hr_offset(int f) { return f == 0 ? 0 : hr_offset<W, DMASK>(f - 1) + hr_words<W, DMASK>(f - 1); }
template <int W, int DMASK = 0>
__host__ __device__ constexpr int hr_total(int nf) {
  return hs_nfrac(nf) == 0 ? 4 : hr_offset<W, DMASK>(hs_nfrac(nf) - 1) + hr_words<W, DMASK>(hs_nfrac(nf) - 1) + 4;
}
// fractions 1/2, 1/4, 1/8, 1/16, 3/16 (indices 0, 2, 3, 7, 8 of the accumulation order): M16 = 8, 4, 2, 1, 3
constexpr int kHrDirect = (1 << 0) | (1 << 2) | (1 << 3) | (1 << 7) | (1 << 8);

// the last word any lane reads of fraction f's range -- lane W/16 - 1, operand max c, rounded up to the
// vector load that fetches it -- lies inside the words reserved (and copied) for that range
template <int W>
__host__ __device__ constexpr bool hr_reads_fit() {
  for (int f = 0; f < 15; f++) {
    const int m16 = hr_m16(f), nw = hr_cmax(m16) + 1;
    const int vec = m16 % 2 == 1 ? 1 : m16 % 4 == 2 ? 2 : 4;
    const int last = (W / 16 - 1) * m16 + ((nw + vec - 1) / vec) * vec - 1;
    // (a remainder word behind full vectors is a single 4-byte load of word nw - 1)
    const int last_exact = (W / 16 - 1) * m16 + nw - 1;
    if ((nw % vec == 0 ? last : last_exact) >= hr_words<W>(f)) return false;
  }
  return true;
}
static_assert(hr_reads_fit<kHrW>(), "a lane would read past its staged range");
static_assert((sizeof(float) * hr_total<kHrW, kHrDirect>(4) + 1152 + 1024) * kHrCtas <= 228 * 1024, "kHrCtas tiles of staging do not fit a SM");

// the words of fraction M16 a lane needs: w[c] = q[c], q = the lane's first word (range base + j M16), in
// shared memory (staged ranges) or global memory (direct fractions)
template <int M16>
__device__ __forceinline__ void hr_load(float* w, const float* __restrict__ q, int lane) {
  constexpr int NW = hr_cmax(M16) + 1;
  if constexpr (M16 % 2 == 1) {
#pragma unroll
    for (int c = 0; c < NW; c++) w[c] = q[c];
  } else if constexpr (M16 % 4 == 2) {
#pragma unroll
    for (int c = 0; c + 1 < NW; c += 2) {
      const float2 v = *reinterpret_cast<const float2*>(q + c);
      w[c] = v.x;
      w[c + 1] = v.y;
    }
    if constexpr (NW % 2 == 1) w[NW - 1] = q[NW - 1];  // (a 2-way conflict: the lane stride is 2 mod 4 words)
  } else {
#pragma unroll
    for (int c = 0; c + 3 < NW; c += 4) {
      const float4 v = *reinterpret_cast<const float4*>(q + c);
      w[c] = v.x;
      w[c + 1] = v.y;
      w[c + 2] = v.z;
      w[c + 3] = v.w;
    }
    if constexpr (NW % 4 != 0) {
      // the one word left over is the next lane's first word
      static_assert(NW % 4 == 1 && NW - 1 == M16, "remainder is word M16");
      float last = __shfl_down_sync(0xffffffffu, w[0], 1);
      if (lane == 31) last = q[NW - 1];
      w[NW - 1] = last;
    }
  }
}
// acc[i] += w[c_i]
template <int M16>
__device__ __forceinline__ void hr_accum(float* acc, const float* w) {
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] += w[(i * M16 + 8) >> 4];
}
template <int M16>
__device__ __forceinline__ void hr_add(float* acc, const float* __restrict__ q, int lane) {
  float w[hr_cmax(M16) + 1];
  hr_load<M16>(w, q, lane);
  hr_accum<M16>(acc, w);
}

template <int NF, int W = kHrW, int CTAS = kHrCtas, int DMASK = kHrDirect>
__global__ void __launch_bounds__(W / 16, CTAS) harmonic_peaks_rows_kernel(const float* __restrict__ Pall, int nbins, size_t pstride,
                                                                            LevelRanges r, LevelScales ls, float thresh, int acc_idx0,
                                                                            PeakSink sink, int ntiles, int nbatch,
                                                                            unsigned long long* __restrict__ dispenser,
                                                                            unsigned long long dispenser_base) {
  static_assert(NF >= 1 && NF <= 4, "shared-memory staging is sized for up to 4 folds");
  constexpr int NFR = hs_nfrac(NF);
  // staged fractions (one fold: the only fraction, 1/2, is read directly and nothing is staged)
  constexpr int NSTAGED = NFR - ((DMASK & 1) + ((DMASK >> 2) & 1 & (NFR > 2)) + ((DMASK >> 3) & 1 & (NFR > 3)) +
                                 ((DMASK >> 7) & 1 & (NFR > 7)) + ((DMASK >> 8) & 1 & (NFR > 8)));
  extern __shared__ __align__(128) float sm[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ int s_next[2];
  const int lane = threadIdx.x & 31;
  const int j = threadIdx.x;  // lane of the tile: bins k0 + 16 j + i
  const int nwork = ntiles * nbatch;
#ifdef PS_PROFILING
  const int probe = g_hs_probe;  // 1: no staging copies after the first tile, 2: copies but no sums
#else
  constexpr int probe = 0;
#endif

  if (threadIdx.x == 0) {
    mbar_init(&bar, NSTAGED > 0 ? NSTAGED : 1);
    fence_proxy_async();
  }
  __syncthreads();
  // Tiles are handed out dynamically: tiles with peaks cost microseconds more (hp_emit_rows) and a static
  // partition left 8 % of the SM time idle behind the slowest CTA.  A CTA starts with tile blockIdx.x and
  // draws the next one from a counter that only grows (this launch's numbers start at dispenser_base;
  // nwork >= gridDim.x, so a launch draws nwork numbers in all: nwork - gridDim.x tiles and one miss per CTA).
  // Thread 0 draws while the current tile is summed; the result crosses the end-of-tile barrier in shared memory.
  unsigned parity = 0;
  int it = 0;
  for (int work = blockIdx.x; work < nwork; it++) {
    unsigned long long drawn = 0;
    if (threadIdx.x == 0) drawn = atomicAdd(dispenser, 1ULL);
    const bool staged_item = probe != 1 || it == 0;
    const int b = work / ntiles, tile = work - b * ntiles;
    const float* __restrict__ P = Pall + (size_t)b * pstride;
    const int k0 = tile * W;
    if (NSTAGED > 0 && (int)threadIdx.x < NFR && !((DMASK >> threadIdx.x) & 1) && staged_item) {  // thread f stages the range of fraction f
      // (offset, length and first word from compile-time constants: no table read in front of the copy)
      const int f = threadIdx.x;
      int off = 0, m16 = 0, words = 0;
#pragma unroll
      for (int g = 0; g < NFR; g++) {
        if (g < f) off += hr_words<W, DMASK>(g);
        if (g == f) {
          m16 = hr_m16(g);
          words = hr_words<W, DMASK>(g);
        }
      }
      const int first = (k0 >> 4) * m16;  // = hs_p(k0, m, s), a multiple of 4: k0 is a multiple of 512
      unsigned bytes = (unsigned)words * 4u;
      if (k0 + W > nbins) {  // ragged last tile
        const int last = hs_p(nbins - 1, c_hs_m[f], c_hs_s[f]);
        bytes = (unsigned)(((last - first + 1) + 3) & ~3) * 4u;
      }
      fence_proxy_async();  // earlier generic-proxy reads of this buffer are ordered before the async writes
      mbar_expect_tx(&bar, bytes);
      bulk_g2s(sm + off, P + first, bytes, &bar);
    }
    // the lane's own bins, straight from global memory (in flight while the ranges are staged)
    float acc[16];
    const int base = k0 + 16 * j;
    if (k0 + W <= nbins) {
#pragma unroll
      for (int v = 0; v < 4; v++) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(P + base) + v);
        acc[4 * v] = x.x;
        acc[4 * v + 1] = x.y;
        acc[4 * v + 2] = x.z;
        acc[4 * v + 3] = x.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; i++) acc[i] = base + i < nbins ? __ldg(P + base + i) : 0.f;
    }
    // the direct fractions' words, requested with the own bins (in flight while the ranges are staged)
    float wd0[9], wd2[5], wd3[3], wd7[2], wd8[4];
    {
      const float* __restrict__ Pg = P + (k0 >> 4);  // P0 = (k0 / 16) M16
      if ((DMASK >> 0) & 1 && 0 < NFR) hr_load<8>(wd0, Pg + (size_t)(k0 >> 4) * 7 + j * 8, lane);
      if ((DMASK >> 2) & 1 && 2 < NFR) hr_load<4>(wd2, Pg + (size_t)(k0 >> 4) * 3 + j * 4, lane);
      if ((DMASK >> 3) & 1 && 3 < NFR) hr_load<2>(wd3, Pg + (size_t)(k0 >> 4) * 1 + j * 2, lane);
      if ((DMASK >> 7) & 1 && 7 < NFR) hr_load<1>(wd7, Pg + j, lane);
      if ((DMASK >> 8) & 1 && 8 < NFR) hr_load<3>(wd8, Pg + (size_t)(k0 >> 4) * 2 + j * 3, lane);
    }
    auto emit = [&](int nh) {
      float mx = acc[0];
#pragma unroll
      for (int i = 1; i < 16; i++) mx = fmaxf(mx, acc[i]);
      if (!__any_sync(0xffffffffu, mx > ls.lo[nh])) return;
      unsigned mask = 0;
#pragma unroll
      for (int i = 0; i < 16; i++) mask |= (acc[i] > ls.lo[nh]) ? (1u << i) : 0u;
      // bins outside the level's search range (the first bins of a spectrum are large) never reach the slow path
      const int lo_i = r.start[nh] - base, hi_i = r.limit[nh] - base;
      if (lo_i > 0) mask &= lo_i >= 16 ? 0u : ~((1u << lo_i) - 1u);
      if (hi_i < 16) mask &= hi_i <= 0 ? 0u : (1u << hi_i) - 1u;
      if (!__any_sync(0xffffffffu, mask != 0u)) return;
      hp_emit_rows(P, nbins, base - 16 * lane, lane, mask, nh, NF, r, thresh, acc_idx0 + b, sink);
    };
    emit(0);
    if (NSTAGED > 0 && staged_item) {
      mbar_wait(&bar, parity);
      parity ^= 1u;
    }
#define PS_HR_ADD(I, M_, S_)                                                 \
  if ((I) < NFR) {                                                           \
    if (probe != 2) {                                                        \
      constexpr int m16_ = (M_) << (4 - (S_));                               \
      if ((DMASK >> (I)) & 1)                                                \
        hr_accum<m16_>(acc, (I) == 0 ? wd0 : (I) == 2 ? wd2 : (I) == 3 ? wd3 : (I) == 7 ? wd7 : wd8); \
      else                                                                   \
        hr_add<m16_>(acc, sm + hr_offset<W, DMASK>(I) + j * m16_, lane);     \
    }                                                                        \
    if ((I) == 0) emit(1);                                                   \
    if ((I) == 2) emit(2);                                                   \
    if ((I) == 6) emit(3);                                                   \
    if ((I) == 14) emit(4);                                                  \
  }
    PS_HS_FRACTIONS(PS_HR_ADD)
#undef PS_HR_ADD
    if (threadIdx.x == 0) s_next[it & 1] = (int)(drawn - dispenser_base) + (int)gridDim.x;
    __syncthreads();  // all warps are done with the buffer: the next iteration's copies may overwrite it
    work = s_next[it & 1];
  }
}

}  // namespace ps
