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
#include <cstdint>
#include <cstdlib>
#include <vector>
#include "fft.cuh"

namespace ps {

int plan_create(unsigned n, ps_fft_plan_t** out, int max_batch);
// acceleration trials per launch: amortises kernel ramp-up/drain and the 148-SM wave tail (measured chain
// us/trial at 2^23 -- r1: 1: 94, 2: 74, 4: 65, 8: 61, 16: 59, 32: 57.7; r2: 26: 49.7, 34: 49.5, 51: 49.1; 34: 46.1, 51: 45.8,
// 101: 45.0).  Up to 128,
// fewer for long transforms: the inter-pass buffer, the packed spectra and the spectra of a launch
// (20 bytes per sample and trial) are kept under kBatchBytes per handle.
constexpr int kAccelBatch = 128;
constexpr size_t kBatchBytes = (size_t)24 << 30;
static int accel_batch_for(unsigned size) {
  const size_t per_trial = (size_t)20 * size;
  return (int)std::max<size_t>(8, std::min<size_t>(kAccelBatch, kBatchBytes / per_trial));
}
void plan_destroy(ps_fft_plan_t* p);

struct LevelRanges {
  int start[6];
  int limit[6];
};

// W_N^k = exp(-i pi k / M) for k in [0, M] from the plan's two-level table
__device__ __forceinline__ float2 twiddle_N(const cf* __restrict__ TC, const cf* __restrict__ TD, unsigned k) {
  cf w = psfft::cmul(psfft::ldg_cf(&TC[k >> 12]), psfft::ldg_cf(&TD[k & 4095u]));
  return make_float2(w.x, w.y);
}

// X[k] = E + W_N^k O from the packed spectrum (E, O: even/odd-sample transforms)
__device__ __forceinline__ float2 split_bin(float2 zk, float2 zm, float2 w) {
  float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);
  float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x);
  return make_float2(er + (w.x * orr - w.y * oi), ei + (w.x * oi + w.y * orr));
}

// Z (packed complex spectrum) -> normalised interbinned amplitude spectrum P[0..M]
// (r2c split, kernels.cu:231-252 interbin, kernels.cu:469-480 normalise with mean*size,std*size
// as in pipeline_multi.cu:224).  One thread = 4 consecutive bins (+ the left neighbour of the
// first); blockIdx.y = batch item.  The split twiddle advances by the constant W_N^1 inside a
// thread; the square root is MUFU.SQRT (<= 1 ulp) and the division a multiplication by 1/sigma
// (<= 1.5 ulp) -- within 2 ulp of the IEEE results of the staged ps_form_power_series/ps_normalise.
__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// normalised interbinned amplitude of one bin given its left neighbour (kernels.cu:231-252, :469-480)
__device__ __forceinline__ float interbin_norm(float2 x, float2 left, float mean, float inv_sigma) {
  const float ampsq = x.x * x.x + x.y * x.y;
  const float dr = x.x - left.x, di = x.y - left.y;
  const float diff = 0.5f * (dr * dr + di * di);
  return (sqrt_approx(fmaxf(ampsq, diff)) - mean) * inv_sigma;
}

// The eleven pairs of one thread (see spectrum_kernel): zk[t] = Z[k0-1+t], zm[t] = Z[M-(k0-1+t)];
// wk0 = W_N^k0, w1 = W_N^1.  Writes bins [k0, k0+8), [M-k0-8, M-k0) and, for k0 = 0, bin M.
__device__ __forceinline__ void spectrum_pairs(const float2* zk, const float2* zm, float2 wk0, float2 w1, unsigned k0, unsigned M,
                                               float mean, float inv_sigma, float* __restrict__ P) {
  float2 w = make_float2(wk0.x * w1.x + wk0.y * w1.y, wk0.y * w1.x - wk0.x * w1.y);  // W_N^(k0-1)
  float2 xa_prev = make_float2(0.f, 0.f), xb_prev = make_float2(0.f, 0.f);
  float outA[8], outB[8];
#pragma unroll
  for (int t = 0; t < 11; t++) {
    const float2 zkk = zk[t], zmm = zm[t];
    const float er = 0.5f * (zkk.x + zmm.x), ei = 0.5f * (zkk.y - zmm.y);
    const float orr = 0.5f * (zkk.y + zmm.y), oi = -0.5f * (zkk.x - zmm.x);
    const float tr = w.x * orr - w.y * oi, ti = w.x * oi + w.y * orr;
    float2 xa = make_float2(er + tr, ei + ti);      // X[k0-1+t]
    const float2 xb = make_float2(er - tr, ti - ei);  // X[M-(k0-1+t)] = conj(E - T)
    if (t == 0 && k0 == 0) xa = make_float2(0.f, 0.f);  // left neighbour of bin 0 is zero
    if (t >= 1 && t <= 8) outA[t - 1] = interbin_norm(xa, xa_prev, mean, inv_sigma);
    // value X[M-(k0-1+(t-1))] with left neighbour X[M-(k0-1+t)]
    if (t >= 3) outB[10 - t] = interbin_norm(xb_prev, xb, mean, inv_sigma);
    if (t == 2 && k0 == 0) P[M] = interbin_norm(xb_prev, xb, mean, inv_sigma);  // Nyquist bin from the pair k = 0
    xa_prev = xa;
    xb_prev = xb;
    w = make_float2(w.x * w1.x - w.y * w1.y, w.x * w1.y + w.y * w1.x);
  }
  *reinterpret_cast<float4*>(&P[k0]) = make_float4(outA[0], outA[1], outA[2], outA[3]);
  *reinterpret_cast<float4*>(&P[k0 + 4]) = make_float4(outA[4], outA[5], outA[6], outA[7]);
  *reinterpret_cast<float4*>(&P[M - k0 - 8]) = make_float4(outB[0], outB[1], outB[2], outB[3]);
  *reinterpret_cast<float4*>(&P[M - k0 - 4]) = make_float4(outB[4], outB[5], outB[6], outB[7]);
}

// One thread = bins [8g, 8g+8) AND their mirror images [M-8g-8, M-8g): the split of bin k and of
// bin M-k share the same two packed values, E_{M-k} = conj E_k, O_{M-k} = conj O_k and
// W_N^{M-k} = -conj W_N^k, so X_k = E + T and X_{M-k} = conj(E - T) with one T = W_N^k O.  Eleven
// pairs k = 8g-1 .. 8g+9 give the 16 outputs and both left neighbours: 11 B of loads per bin
// where the one-sided version read 20.  Thread 0 also writes the Nyquist bin M.
__global__ void __launch_bounds__(256) spectrum_kernel(const float2* __restrict__ Zall, float* __restrict__ Pall, unsigned M,
                                                       size_t pstride, const float* __restrict__ stats, float fsize,
                                                       const cf* __restrict__ TC, const cf* __restrict__ TD, float2 w1) {
  const float2* __restrict__ Z = Zall + (size_t)blockIdx.y * M;
  float* __restrict__ P = Pall + (size_t)blockIdx.y * pstride;
  const float mean = __fmul_rn(stats[0], fsize);
  const float inv_sigma = __fdiv_rn(1.0f, __fmul_rn(stats[2], fsize));
  const unsigned ngroups = M / 16;
  for (unsigned g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
    const unsigned k0 = 8 * g;
    // zk[t] = Z[k0-1+t], zm[t] = Z[M-(k0-1+t)], t = 0..10
    float2 zk[11], zm[11];
    zk[0] = k0 ? Z[k0 - 1] : make_float2(0.f, 0.f);
    zm[0] = k0 ? Z[M - k0 + 1] : make_float2(0.f, 0.f);
#pragma unroll
    for (int q = 0; q < 5; q++) {
      const float4 a = *reinterpret_cast<const float4*>(&Z[k0 + 2 * q]);
      zk[1 + 2 * q] = make_float2(a.x, a.y);
      zk[2 + 2 * q] = make_float2(a.z, a.w);
    }
#pragma unroll
    for (int q = 1; q < 5; q++) {  // Z[M-k0-2q], Z[M-k0-2q+1]
      const float4 a = *reinterpret_cast<const float4*>(&Z[M - k0 - 2 * q]);
      zm[1 + 2 * q] = make_float2(a.x, a.y);
      zm[2 * q] = make_float2(a.z, a.w);
    }
    zm[1] = Z[k0 ? M - k0 : 0];  // Z[M] == Z[0]
    zm[10] = Z[M - k0 - 9];
    spectrum_pairs(zk, zm, twiddle_N(TC, TD, k0), w1, k0, M, mean, inv_sigma, P);
  }
}


struct PeakSink {
  ps_raw_peak* peaks;
  unsigned* count;
  unsigned cap;
};

// harmonic sums + threshold + compaction (kernels.cu:33-99 + :384-416 + peakfinder.hpp:77-86).
// A warp owns 512 consecutive bins; lane l holds bins base+32i (i < 16, base = chunk*512+l), so
// for the fraction m/2^sh the read index is p0 + i*(32m >> sh): one index computation per
// (thread, fraction), 16 loads at immediate offsets, and the 32 lanes of each load touch <= 2m+1
// consecutive words (coalesced).  Sums are compared in the accumulator domain against a float
// lower bound of thresh/scale: only bins that may exceed the threshold go through the reference's
// double-precision scaling, so emitted values and decisions are bit-identical to it.
struct LevelScales {
  float lo[6];  // lower bound on the unscaled sum that can reach thresh at each level
};

template <bool FULL>
__device__ __forceinline__ float hp_load(const float* __restrict__ q, int off, int idx, int nbins) {
  if (FULL) return __ldg(q + off);
  return idx < nbins ? __ldg(q + off) : 0.f;
}

#define PS_ADD_FRACTION(M_, SH_)                                                              \
  {                                                                                           \
    const float* __restrict__ q_ = P + ((base * (M_) + (1 << ((SH_)-1))) >> (SH_));           \
    _Pragma("unroll") for (int i = 0; i < 16; i++)                                            \
        acc[i] += hp_load<FULL>(q_, i * ((32 * (M_)) >> (SH_)), base + 32 * i, nbins);        \
  }

// Rare path: some of this warp's bins may exceed the threshold at level nh.  The flagged bins are
// re-evaluated with the reference's per-bin arithmetic (harmonic_folds: same accumulation order,
// double-precision scale factors), tested with the strict '>' and the level's bin range, and
// appended to the peak list with one atomic per warp.
__device__ __noinline__ void hp_emit_slow(const float* __restrict__ P, int nbins, int base, int lane, unsigned mask, int nh,
                                          int nfolds, const LevelRanges& r, float thresh, int acc_idx, const PeakSink& sink) {
#pragma unroll 1
  for (int i = 0; i < 16; i++) {
    if (!__any_sync(0xffffffffu, (mask >> i) & 1u)) continue;
    const int idx = base + 32 * i;
    bool hit = false;
    float v = 0.f;
    if (((mask >> i) & 1u) && idx < nbins) {
      float f[5];
      harmonic_folds(P, idx, nfolds, f);
      v = nh == 0 ? __ldg(P + idx) : f[nh - 1];
      hit = idx >= r.start[nh] && idx < r.limit[nh] && v > thresh;
    }
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (m) {
      unsigned pos = 0;
      if (lane == 0) pos = atomicAdd(sink.count, (unsigned)__popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1));
      if (hit && pos < sink.cap) {
        ps_raw_peak pk;
        pk.acc_idx = acc_idx;
        pk.nh = nh;
        pk.bin = idx;
        pk.snr = v;
        sink.peaks[pos] = pk;
      }
    }
  }
}

// The same for lanes that own sixteen CONSECUTIVE bins (harmonic_rows.cuh): bit i of a lane's mask = bin
// warp_base + 16 lane + i.  A cluster of neighbouring bins sits in one lane there, so the flagged bins of the
// warp are first spread over its lanes (sixteen ballots, the n-th flagged bin goes to lane n % 32): one
// round of the exact re-evaluation per 32 flagged bins, however they are distributed.
__device__ __noinline__ void hp_emit_rows(const float* __restrict__ P, int nbins, int warp_base, int lane, unsigned mask, int nh,
                                          int nfolds, const LevelRanges& r, float thresh, int acc_idx, const PeakSink& sink) {
  unsigned bal[16];
  int total = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) {
    bal[i] = __ballot_sync(0xffffffffu, (mask >> i) & 1u);
    total += __popc(bal[i]);
  }
#pragma unroll 1
  for (int n0 = 0; n0 < total; n0 += 32) {
    const int n = n0 + lane;
    int idx = -1;
    if (n < total) {
      int c = 0;
#pragma unroll
      for (int i = 0; i < 16; i++) {
        const int pc = __popc(bal[i]);
        if (idx < 0 && n < c + pc) idx = warp_base + 16 * (int)__fns(bal[i], 0, n - c + 1) + i;
        c += pc;
      }
    }
    bool hit = false;
    float v = 0.f;
    if (idx >= 0 && idx < nbins) {
      float f[5];
      harmonic_folds(P, idx, nfolds, f);
      v = nh == 0 ? __ldg(P + idx) : f[nh - 1];
      hit = idx >= r.start[nh] && idx < r.limit[nh] && v > thresh;
    }
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (m) {
      unsigned pos = 0;
      if (lane == 0) pos = atomicAdd(sink.count, (unsigned)__popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1));
      if (hit && pos < sink.cap) {
        ps_raw_peak pk;
        pk.acc_idx = acc_idx;
        pk.nh = nh;
        pk.bin = idx;
        pk.snr = v;
        sink.peaks[pos] = pk;
      }
    }
  }
}

}  // namespace ps
#include "harmonic_rows.cuh"
namespace ps {

template <int NF, bool FULL>
__device__ __forceinline__ void hp_chunk(const float* __restrict__ P, int nbins, int base, int lane, const LevelRanges& r,
                                         const LevelScales& ls, float thresh, int acc_idx, const PeakSink& sink) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = hp_load<FULL>(P + base, 32 * i, base + 32 * i, nbins);

  auto emit = [&](int nh) {
    float mx = acc[0];
#pragma unroll
    for (int i = 1; i < 16; i++) mx = fmaxf(mx, acc[i]);
    if (!__any_sync(0xffffffffu, mx > ls.lo[nh])) return;
    unsigned mask = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) mask |= (acc[i] > ls.lo[nh]) ? (1u << i) : 0u;
    hp_emit_slow(P, nbins, base, lane, mask, nh, NF, r, thresh, acc_idx, sink);
  };

  emit(0);
  if (NF > 0) {
    PS_ADD_FRACTION(1, 1)
    emit(1);
  }
  if (NF > 1) {
    PS_ADD_FRACTION(3, 2)
    PS_ADD_FRACTION(1, 2)
    emit(2);
  }
  if (NF > 2) {
    PS_ADD_FRACTION(1, 3)
    PS_ADD_FRACTION(3, 3)
    PS_ADD_FRACTION(5, 3)
    PS_ADD_FRACTION(7, 3)
    emit(3);
  }
  if (NF > 3) {
    PS_ADD_FRACTION(1, 4)
    PS_ADD_FRACTION(3, 4)
    PS_ADD_FRACTION(5, 4)
    PS_ADD_FRACTION(7, 4)
    PS_ADD_FRACTION(9, 4)
    PS_ADD_FRACTION(11, 4)
    PS_ADD_FRACTION(13, 4)
    PS_ADD_FRACTION(15, 4)
    emit(4);
  }
  if (NF > 4) {
    PS_ADD_FRACTION(1, 5)
    PS_ADD_FRACTION(3, 5)
    PS_ADD_FRACTION(5, 5)
    PS_ADD_FRACTION(7, 5)
    PS_ADD_FRACTION(9, 5)
    PS_ADD_FRACTION(11, 5)
    PS_ADD_FRACTION(13, 5)
    PS_ADD_FRACTION(15, 5)
    PS_ADD_FRACTION(17, 5)
    PS_ADD_FRACTION(19, 5)
    PS_ADD_FRACTION(21, 5)
    PS_ADD_FRACTION(23, 5)
    PS_ADD_FRACTION(25, 5)
    PS_ADD_FRACTION(27, 5)
    PS_ADD_FRACTION(29, 5)
    PS_ADD_FRACTION(31, 5)
    emit(5);
  }
}

// one warp = one chunk of 512 bins; blockIdx.y = batch item
// 128 registers per thread (4 blocks of 128 threads per SM): batching ~64 independent loads per
// thread hides L2 latency better here than more resident warps do (measured: 4 > 8 > 12 > 16 blocks)
template <int NF>
__global__ void __launch_bounds__(128, 4) harmonic_peaks_kernel(const float* __restrict__ Pall, int nbins, size_t pstride,
                                                             LevelRanges r, LevelScales ls, float thresh, int acc_idx0,
                                                             PeakSink sink) {
  const float* __restrict__ P = Pall + (size_t)blockIdx.y * pstride;
  const int acc_idx = acc_idx0 + (int)blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (chunk * 512 >= nbins) return;
  const int base = chunk * 512 + lane;
  if (chunk * 512 + 512 <= nbins)
    hp_chunk<NF, true>(P, nbins, base, lane, r, ls, thresh, acc_idx, sink);
  else
    hp_chunk<NF, false>(P, nbins, base, lane, r, ls, thresh, acc_idx, sink);
}

__global__ void __launch_bounds__(256) untranspose_pairs_kernel(const float2* __restrict__ D, float2* __restrict__ out, unsigned M,
                                                                int n1, int lg_n2) {
  for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < M; e += gridDim.x * blockDim.x)
    out[e] = D[(size_t)(e & ((1u << lg_n2) - 1u)) * n1 + (e >> lg_n2)];
}

// Two-phase form of the transposed series (LoadResampled): pair e = (x[2e], x[2e+1]) sits at q = c * n1 + row of D
// (c = e % n2, row = e / n2); plane 2c of Dph is column c itself, plane 2c + 1 the pairs one sample on,
// (x[2e+1], x[2e+2]).  Pair e+1 is one column on (n1 words), or in row+1 of column 0 after the last column;
// the very last pair has no successor.
__global__ void __launch_bounds__(256) phase_planes_kernel(const float2* __restrict__ D, float2* __restrict__ Dph, unsigned M, int n1) {
  for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < M; q += gridDim.x * blockDim.x) {
    const unsigned row = q % (unsigned)n1;
    const bool last_col = q + (unsigned)n1 >= M;
    const unsigned nxt = last_col ? row + 1u : q + (unsigned)n1;
    float2 b = make_float2(0.f, 0.f);
    if (!(last_col && row + 1u >= (unsigned)n1)) b = D[nxt];
    const float2 a = D[q];
    const size_t o = (size_t)q + (q - row);  // 2 c n1 + row
    Dph[o] = a;
    Dph[o + (unsigned)n1] = make_float2(a.y, b.x);
  }
}

// diagnostics (ps_search_resample_index_map): a series whose sample i is the bit pattern of i, in the
// transposed pair layout, and the search path's loader run over all of its pairs
__global__ void __launch_bounds__(256) ramp_pairs_kernel(float2* __restrict__ D, unsigned M, int n1, int lg_n2) {
  for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < M; e += gridDim.x * blockDim.x)
    D[(size_t)(e & ((1u << lg_n2) - 1u)) * n1 + (e >> lg_n2)] = make_float2(__uint_as_float(2u * e), __uint_as_float(2u * e + 1u));
}
__global__ void __launch_bounds__(256) loader_dump_kernel(LoadResampled ld, float2* __restrict__ out, unsigned M, int n2,
                                                          unsigned* __restrict__ exact) {
  ld.bind(ld);  // (a by-value parameter: its table sits in this thread's local memory, fine for a diagnostic)
  ld.begin(0, 0, n2);
  for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < M; e += gridDim.x * blockDim.x) {
    const cf v = ld.at(2u * e);
    out[e] = make_float2(v.x, v.y);
    // the test at() applies: count the pairs that went through the FP64 expression
    const float fi = (float)(2u * e);
    const float g = (fi * (fi - ld.fsize)) * ld.af32;
    const float r = (g + 12582912.f) - 12582912.f;
    if (!(fabsf(g - r) < ld.thr)) atomicAdd(exact, 1u);
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

// ---- per-DM whitening, fused passes (the staged entry points keep one kernel per reference call)
// Z (packed transform) -> X (Hermitian bins, r2c_post_kernel's arithmetic) and |X| (kernels.cu:215-227)
// in one pass: the median filter needs |X|, the de-reddening needs X.
__global__ void __launch_bounds__(256) whiten_split_power_kernel(const float2* __restrict__ Z, float2* __restrict__ X,
                                                                 float* __restrict__ pspec, unsigned M) {
  for (unsigned k = blockIdx.x * blockDim.x + threadIdx.x; k <= M; k += gridDim.x * blockDim.x) {
    const float2 zk = Z[k == M ? 0 : k];
    const float2 zm = Z[k == 0 || k == M ? 0 : M - k];
    float wr, wi;
    sincospif((float)k / (float)M, &wi, &wr);  // W_N^k = (cos, -sin)(pi k/M)
    const float2 x = split_bin(zk, zm, make_float2(wr, -wi));
    X[k] = x;
    pspec[k] = amplitude_rsqrt(x.x, x.y);
  }
}

// zapping (kernels.cu:1036-1069) expressed on the MEDIAN: a negative median marks a zapped bin, and
// whiten_deredden_stats_kernel writes (1, 0) there.  One warp per birdie.
__global__ void zap_mark_kernel(float* __restrict__ median, const float* __restrict__ birdies, const float* __restrict__ widths,
                                float bin_width, unsigned nbirdies, unsigned nbins) {
  unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= nbirdies) return;
  int lo, hi;
  if (!zap_range(birdies[w], widths[w], bin_width, nbins, &lo, &hi)) return;
  for (int i = lo + (int)lane; i < hi; i += 32) median[i] = -1.0f;
}

__device__ __forceinline__ float2 whitened_bin(const float2* __restrict__ X, const float* __restrict__ median, unsigned i) {
  const float m = median[i];
  if (m < 0.f) return make_float2(1.f, 0.f);       // zapped
  return deredden_value(X[i], m, i);               // bins 0-4 zeroed, the rest divided by the median
}

// de-redden + zap + interbinned amplitude + its sum and sum of squares (dereddener.hpp:64-66,
// birdiezapper.hpp, spectrumformer.hpp:8-14, stats.hpp:25-33) in one pass: the whitened spectrum is
// written once (to Xw, read by the C2R), the interbinned spectrum is only needed for its mean and
// standard deviation and is never stored.  Four bins per thread + the left neighbour.
__global__ void __launch_bounds__(256) whiten_deredden_stats_kernel(const float2* __restrict__ X, const float* __restrict__ median,
                                                                    float2* __restrict__ Xw, unsigned nbins,
                                                                    double* __restrict__ partials) {
  double s = 0.0, q = 0.0;
  const unsigned ngroups = (nbins + 3) / 4;
  for (unsigned g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
    const unsigned i0 = 4 * g;
    float2 left = i0 ? whitened_bin(X, median, i0 - 1) : make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const unsigned i = i0 + j;
      if (i < nbins) {
        const float2 x = whitened_bin(X, median, i);
        Xw[i] = x;
        const double v = interbin_amplitude(x.x, x.y, left.x, left.y);
        s += v;
        q += v * v;
        left = x;
      }
    }
  }
  __shared__ double sh[2][8];
  for (int o = 16; o; o >>= 1) {
    s += __shfl_down_sync(0xffffffffu, s, o);
    q += __shfl_down_sync(0xffffffffu, q, o);
  }
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
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

struct Search {
  ps_search_params prm;
  ps_fft_plan_t* plan = nullptr;
  unsigned nbins = 0;
  float* d_tim = nullptr;      // whitened series as size/2 complex pairs, TRANSPOSED (see LoadResampled)
  float* d_tim_nat = nullptr;  // natural-order copy, made on request (ps_search_whitened_series)
  float2* d_tim_ph = nullptr;  // two-phase form read by the resampling loader (phase_planes_kernel): 2 * size/2 pairs
  int lg_n2 = 0;    // log2(pair columns) of the transposed whitened series
  int col_len = 0;  // pairs per column
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
  // tile dispenser of harmonic_peaks_rows_kernel: a counter that only grows; a launch starts at hwork_base
  unsigned long long* d_hwork = nullptr;
  unsigned long long hwork_base = 0;
  unsigned cap = 0;
  unsigned needed = 0;  // record count of the last fetch (ps_search_peaks_needed)
  LevelRanges ranges;
  LevelScales scales;
  int batch = 0;               // acceleration trials per launch (accel_batch_for(size))
  float* d_pbatch = nullptr;   // `batch` spectra (stride pstride floats)
  size_t pstride = 0;
  const float* last_spectrum = nullptr;
  float bin_width_f = 0.f;   // float bin_width = 1.0/tobs   (pipeline_multi.cu:118-119)
  double bin_width = 0.0;    // stored as double in FrequencySeries (fourierseries.hpp:41)
};

static void free_search(Search* s) {
  if (!s) return;
  plan_destroy(s->plan);
  cudaFree(s->d_tim); cudaFree(s->d_tim_nat); cudaFree(s->d_tim_ph); cudaFree(s->d_X); cudaFree(s->d_pspec); cudaFree(s->d_median); cudaFree(s->d_work);
  cudaFree(s->d_stats); cudaFree(s->d_partials); cudaFree(s->d_u8sum); cudaFree(s->d_birdies); cudaFree(s->d_widths);
  cudaFree(s->d_peaks); cudaFree(s->d_count); cudaFree(s->d_hwork); cudaFree(s->d_pbatch);
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

// harmonic sums + peaks with TMA-staged operands (harmonic_rows.cuh); spectra must be 16-byte aligned and padded
template <int NF, int W = kHrW, int CTAS = kHrCtas, int DMASK = kHrDirect>
static int launch_harmonic_rows_nf(Search* s, const float* d_spectra, size_t stride, int nb, int acc_idx0, cudaStream_t st) {
  auto k = harmonic_peaks_rows_kernel<NF, W, CTAS, DMASK>;
  static SmemOptIn done[64];
  const size_t smem = sizeof(float) * hr_total<W, DMASK>(NF);
  int rc = configure_smem(k, smem, s->plan->device, done);
  if (rc) return rc;
  PeakSink sink{s->d_peaks, s->d_count, s->cap};
#ifdef PS_PROFILING
  {
    const int probe = getenv("PS_HS_PROBE") ? atoi(getenv("PS_HS_PROBE")) : 0;
    cudaMemcpyToSymbolAsync(g_hs_probe, &probe, sizeof(int), 0, cudaMemcpyHostToDevice, st);
  }
#endif
  const int ntiles = ((int)s->nbins + W - 1) / W;
  const int work = ntiles * nb, cap = 148 * CTAS;
  k<<<work < cap ? work : cap, W / 16, smem, st>>>(d_spectra, (int)s->nbins, stride, s->ranges, s->scales, s->prm.min_snr, acc_idx0,
                                                   sink, ntiles, nb, s->d_hwork, s->hwork_base);
  PS_LAUNCHED("harmonic_peaks_rows_kernel");
  s->hwork_base += (unsigned long long)work;  // a launch that was accepted draws exactly `work` numbers (harmonic_rows.cuh)
  return PS_OK;
}

static int launch_harmonic_staged(Search* s, const float* sp, size_t stride, int nb, int acc_idx0, cudaStream_t st) {
  const int nf = s->prm.nharmonics;
  return nf == 1 ? launch_harmonic_rows_nf<1>(s, sp, stride, nb, acc_idx0, st)
       : nf == 2 ? launch_harmonic_rows_nf<2>(s, sp, stride, nb, acc_idx0, st)
       : nf == 3 ? launch_harmonic_rows_nf<3>(s, sp, stride, nb, acc_idx0, st)
                 : launch_harmonic_rows_nf<4>(s, sp, stride, nb, acc_idx0, st);
}

// loader of one launch: the acceleration factors of its nb trials and the bounds the float32 estimate needs
static void setup_resampler(LoadResampled& ld, const Search* s, const float2* pairs, const float2* planes, const float* h_acc, int nb) {
  const double size = (double)s->prm.size;
  ld.D = pairs;
  ld.Dph = planes;
  ld.n1 = s->col_len;
  ld.lg_n2 = s->lg_n2;
  double amax = 0;
  for (int b = 0; b < nb; b++) {
    ld.afs.v[b] = accel_fact(h_acc[b], s->prm.tsamp);
    amax = std::max(amax, fabs(ld.afs.v[b]));
  }
  ld.size = size;
  ld.max_shift = (float)(amax * 0.25 * size * size * 1.001);  // |i(i-N)| <= N^2/4
  ld.max_step = (float)(amax * size * 1.001);                 // |d/di| <= N
}

extern "C" {

int ps_search_create(const ps_search_params* p, ps_search** out) {
  if (!p || !out) return invalid("ps_search_create: null argument");
  if (p->nharmonics < 0 || p->nharmonics > 5) return invalid("ps_search_create: nharmonics (folds) must be 0..5");
  if (!p->nsamps) return invalid("ps_search_create: nsamps == 0");
  Search* s = new Search();
  s->prm = *p;
  if (s->prm.boundary_5_freq <= 0) s->prm.boundary_5_freq = 0.05f;
  if (s->prm.boundary_25_freq <= 0) s->prm.boundary_25_freq = 0.5f;
  s->batch = accel_batch_for(p->size);
  if (const char* e = getenv("PS_BATCH")) s->batch = std::max(1, std::min(s->batch, atoi(e)));  // (also bounds the handle's memory: 20 bytes per sample and trial)
  int rc = plan_create(p->size, &s->plan, s->batch);
  if (rc) {
    delete s;
    return rc;
  }
  unsigned size = p->size;
  s->nbins = size / 2 + 1;
  while ((1 << s->lg_n2) < s->plan->n2) s->lg_n2++;
  s->col_len = s->plan->n1;
  float tobs = size * p->tsamp;            // pipeline_multi.cu:118
  float bin_width = 1.0 / tobs;            // :119
  s->bin_width_f = bin_width;
  s->bin_width = bin_width;
  s->cap = p->max_peaks ? p->max_peaks : (1u << 20);
  level_ranges(s->prm, s->nbins, s->bin_width, &s->ranges);
  {
    // float32 lower bounds: a sum below thresh/scale*(1-4e-7) cannot exceed thresh after the
    // reference's (float)((double)sum*scale); bounds are rounded toward -inf
    const double sc[6] = {1.0, 1.0 / sqrt(2.0), 0.5, 1.0 / sqrt(8.0), 0.25, 1.0 / sqrt(32.0)};
    for (int i = 0; i < 6; i++) {
      double lo = (double)p->min_snr / sc[i];
      lo -= fabs(lo) * 1e-6;
      s->scales.lo[i] = nextafterf((float)lo, -INFINITY);
    }
    s->pstride = ((size_t)s->nbins + 63) & ~(size_t)63;
  }
  size_t nb = s->nbins;
  cudaError_t e = cudaMalloc(&s->d_tim, sizeof(float) * size);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_tim_ph, sizeof(float2) * (size_t)size);
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
  if (e == cudaSuccess) e = cudaMalloc(&s->d_pbatch, sizeof(float) * s->pstride * s->batch);
  if (e == cudaSuccess) e = cudaMemset(s->d_count, 0, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_hwork, sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(s->d_hwork, 0, sizeof(unsigned long long));
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
  StoreNatural sn{reinterpret_cast<float2*>(p->Z), 0, 0, 0, nullptr};
  if ((rc = launch_pass2<false>(p, sn, st))) return rc;
  // X and |X| in one pass, running median, then de-redden + zap + interbin + stats in one pass; the
  // whitened spectrum goes to the plan's Z buffer (free here: X still holds the unwhitened bins)
  whiten_split_power_kernel<<<ew_grid(p->M + 1, 256, 2), 256, 0, st>>>(reinterpret_cast<const float2*>(p->Z), s->d_X, s->d_pspec, p->M);
  PS_LAUNCHED("whiten_split_power_kernel");
  if ((rc = launch_running_median(s->d_pspec, s->nbins, s->bin_width, s->prm.boundary_5_freq, s->prm.boundary_25_freq,
                                  s->d_median, s->d_work, st)))
    return rc;
  if (s->nbirdies) {
    const unsigned threads = 128, blocks = (s->nbirdies * 32 + threads - 1) / threads;
    zap_mark_kernel<<<blocks, threads, 0, st>>>(s->d_median, s->d_birdies, s->d_widths, s->bin_width_f, s->nbirdies, s->nbins);
    PS_LAUNCHED("zap_mark_kernel");
  }
  float2* Xw = reinterpret_cast<float2*>(p->Z);
  {
    unsigned g = ew_grid((s->nbins + 3) / 4, 256, 2);
    if (g > kMaxStatParts) g = kMaxStatParts;
    whiten_deredden_stats_kernel<<<g, 256, 0, st>>>(s->d_X, s->d_median, Xw, s->nbins, s->d_partials);
    PS_LAUNCHED("whiten_deredden_stats_kernel");
    if ((rc = launch_finalize_stats(s->d_partials, (int)g, s->nbins, s->d_stats, st))) return rc;
  }
  LoadC2RPre lc{Xw, p->M, p->n2, 0};
  if ((rc = launch_pass1<true>(p, lc, st))) return rc;
  StoreTransposedPairs so{reinterpret_cast<float2*>(s->d_tim), 0, 0, s->lg_n2, s->col_len};
  if ((rc = launch_pass2<true>(p, so, st))) return rc;
  phase_planes_kernel<<<ew_grid(p->M, 256, 4), 256, 0, st>>>(reinterpret_cast<const float2*>(s->d_tim), s->d_tim_ph, p->M, s->col_len);
  PS_LAUNCHED("phase_planes_kernel");
  if (h_mean || h_std) {
    float h[3];
    PS_CUDA(cudaMemcpyAsync(h, s->d_stats, sizeof(h), cudaMemcpyDeviceToHost, st), "Error from d2hcpy");
    PS_CUDA(cudaStreamSynchronize(st), "ps_search_whiten");
    if (h_mean) *h_mean = h[0];
    if (h_std) *h_std = h[2];
  }
  return PS_OK;
}

// Z -> normalised interbinned spectra of nb trials
static int launch_spectrum(Search* s, int nb, cudaStream_t st) {
  const ps_fft_plan_t* p = s->plan;
  dim3 g1(ew_grid(p->M / 16, 256, 1), nb);
  spectrum_kernel<<<g1, 256, 0, st>>>(reinterpret_cast<const float2*>(p->Z), s->d_pbatch, p->M, s->pstride, s->d_stats,
                                      (float)s->prm.size, p->TC, p->TD,
                                      make_float2((float)cos(M_PI / (double)p->M), (float)-sin(M_PI / (double)p->M)));
  PS_LAUNCHED("spectrum_kernel");
  return PS_OK;
}

// harmonic sums + peaks of nb spectra (blockIdx.y = spectrum)
static int launch_harmonic_peaks(Search* s, const float* d_spectra, size_t stride, int nb, int acc_idx0, cudaStream_t st) {
  PeakSink sink{s->d_peaks, s->d_count, s->cap};
  dim3 g2((((unsigned)s->nbins + 511) / 512 + 3) / 4, nb);
#define PS_HP(NF)                                                                                                          \
  harmonic_peaks_kernel<NF><<<g2, 128, 0, st>>>(d_spectra, (int)s->nbins, stride, s->ranges, s->scales, s->prm.min_snr, \
                                                 acc_idx0, sink)
  switch (s->prm.nharmonics) {
    case 0: PS_HP(0); break;
    case 1: PS_HP(1); break;
    case 2: PS_HP(2); break;
    case 3: PS_HP(3); break;
    case 4: PS_HP(4); break;
    default: PS_HP(5); break;
  }
#undef PS_HP
  PS_LAUNCHED("harmonic_peaks_kernel");
  return PS_OK;
}

int ps_search_harmonic_peaks(ps_search* hs, const float* d_spectra, size_t stride, int nbatch, int acc_idx0, void* stream) {
  Search* s = SS(hs);
  if (!s || !d_spectra) return invalid("ps_search_harmonic_peaks: null argument");
  if (nbatch < 0 || stride < s->nbins) return invalid("ps_search_harmonic_peaks: stride must be >= size/2+1");
  // spectra that are 16-byte aligned and padded by >= 4 floats can take the TMA-staged kernel
  static const int env_hsmem = getenv("PS_HSMEM") ? atoi(getenv("PS_HSMEM")) : 1;
  const int nf = s->prm.nharmonics;
  const bool staged = env_hsmem && nf >= 1 && nf <= 4 && stride % 4 == 0 && stride >= (size_t)s->nbins + 4 &&
                      (reinterpret_cast<uintptr_t>(d_spectra) & 15u) == 0;
  for (int j = 0; j < nbatch; j += 32768) {  // grid.y limit of the LDG kernel
    int nb = std::min(32768, nbatch - j);
    const float* sp = d_spectra + (size_t)j * stride;
    int rc = !staged ? launch_harmonic_peaks(s, sp, stride, nb, acc_idx0 + j, S(stream))
                     : launch_harmonic_staged(s, sp, stride, nb, acc_idx0 + j, S(stream));
    if (rc) return rc;
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
  static const int env_batch_cap = getenv("PS_BATCH") ? std::max(1, atoi(getenv("PS_BATCH"))) : kAccelBatch;
  const int env_batch = std::min(env_batch_cap, s->batch);
#ifdef PS_PROFILING  // profiling builds only (make EXTRA=-DPS_PROFILING): the shipped library has no such switch
  static const int env_skip = getenv("PS_SKIP") ? atoi(getenv("PS_SKIP")) : 0;
#else
  constexpr int env_skip = 0;
#endif
  static const int env_rev = getenv("PS_REVERSE") ? atoi(getenv("PS_REVERSE")) : 1;
  static const double band_bytes = (getenv("PS_BAND_MB") ? atof(getenv("PS_BAND_MB")) : 24.0) * 1048576.0;
  // balanced launches: 201 trials at up to 71 per launch run as 67+67+67, not 71+71+59
  int launches_left = (n_acc + env_batch - 1) / env_batch;
  int last_nb = 1;
  for (int j = 0, nb = 0; j < n_acc; j += nb, launches_left--) {
    nb = (n_acc - j + launches_left - 1) / launches_left;
    last_nb = nb;
    if (!(env_skip & 1)) {
      // Pass 1 of the launch's nb trials in groups of neighbouring accelerations: the trials of a group are
      // transformed tile by tile together, and at any time they read a band of the series as wide as the
      // difference of their shifts (up to (af_hi - af_lo) N^2 / 8 pairs = plane columns of n1 pairs).  The
      // band has to stay in the L2 from one tile to the next: at 2^24 and +-500 m/s^2 a 67-trial launch
      // spans 1250 columns = 80 MB and 41 % of the loads went to DRAM; groups are cut at PS_BAND_MB (24 MB:
      // chain us / trial for 201 trials at 6 / 10 / 14 / 20 / 40 / 60 MB and uncut: 106.4 / 104.2 / 103.5 / 102.0 / 103.0 / 106.8 / 109.6).
      const float2* pairs = reinterpret_cast<const float2*>(s->d_tim);
      const double col_bytes = 16.0 * (double)s->col_len, quarter = 0.125 * (double)size * (double)size;
      for (int g0 = 0; g0 < nb;) {
        double lo = accel_fact(h_acc[j + g0], s->prm.tsamp), hi = lo;
        int g1 = g0 + 1;
        for (; g1 < nb; g1++) {
          const double a = accel_fact(h_acc[j + g1], s->prm.tsamp);
          const double nlo = std::min(lo, a), nhi = std::max(hi, a);
          if ((nhi - nlo) * quarter * col_bytes > band_bytes) break;
          lo = nlo;
          hi = nhi;
        }
        LoadResampled ld;
        setup_resampler(ld, s, pairs, s->d_tim_ph, h_acc + j + g0, g1 - g0);
        rc = launch_pass1<false>(p, ld, st, g1 - g0, g0);
        if (rc) return rc;
        g0 = g1;
      }
    }
    {
      StoreNatural sn{reinterpret_cast<float2*>(p->Z), (size_t)p->M, 0, 0, nullptr};
      if (!(env_skip & 2) && (rc = launch_pass2<false>(p, sn, st, nb, env_rev))) return rc;
      if (!(env_skip & 4) && (rc = launch_spectrum(s, nb, st))) return rc;
    }
    if (env_skip & 8) continue;
    static const int env_hsmem = getenv("PS_HSMEM") ? atoi(getenv("PS_HSMEM")) : 1;
    const int nf = s->prm.nharmonics;
    if (env_hsmem && nf >= 1 && nf <= 4) {  // the pipeline's own spectra are padded: TMA-staged variants
      if ((rc = launch_harmonic_staged(s, s->d_pbatch, s->pstride, nb, acc_idx0 + j, st))) return rc;
      continue;
    }
    if ((rc = launch_harmonic_peaks(s, s->d_pbatch, s->pstride, nb, acc_idx0 + j, st))) return rc;
  }
  s->last_spectrum = s->d_pbatch + (size_t)(last_nb - 1) * s->pstride;
  return PS_OK;
}

int ps_search_reset_peaks(ps_search* hs, void* stream) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_reset_peaks: null handle");
  PS_CUDA(cudaMemsetAsync(s->d_count, 0, sizeof(unsigned), S(stream)), "ps_search_reset_peaks");
  return PS_OK;
}

unsigned ps_search_peaks_needed(ps_search* hs) {
  Search* s = SS(hs);
  return s ? s->needed : 0;
}

int ps_search_reserve_peaks(ps_search* hs, unsigned max_peaks) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_reserve_peaks: null handle");
  if (max_peaks <= s->cap) return PS_OK;
  PS_CUDA(cudaDeviceSynchronize(), "ps_search_reserve_peaks");
  ps_raw_peak* p = nullptr;
  PS_CUDA(cudaMalloc(&p, sizeof(ps_raw_peak) * (size_t)max_peaks), "Error from device_malloc");
  cudaFree(s->d_peaks);
  s->d_peaks = p;
  s->cap = max_peaks;
  return PS_OK;
}

long ps_search_fetch_peaks(ps_search* hs, ps_raw_peak* h_out, size_t max_out, void* stream) {
  Search* s = SS(hs);
  if (!s) return invalid("ps_search_fetch_peaks: null handle");
  cudaStream_t st = S(stream);
  unsigned count = 0;
  PS_CUDA(cudaMemcpyAsync(&count, s->d_count, sizeof(unsigned), cudaMemcpyDeviceToHost, st), "Error from d2hcpy");
  PS_CUDA(cudaStreamSynchronize(st), "ps_search_fetch_peaks");
  s->needed = count;
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

// natural-order copy of the (internally transposed) whitened series; synchronous, diagnostics only
const float* ps_search_whitened_series(ps_search* hs) {
  Search* s = SS(hs);
  if (!s) return nullptr;
  const ps_fft_plan_t* p = s->plan;
  if (!s->d_tim_nat && cudaMalloc(&s->d_tim_nat, sizeof(float) * s->prm.size) != cudaSuccess) return nullptr;
  // the whitening may still be running on a caller's non-blocking stream: the copy below is launched on the
  // default stream, so wait for the device first
  if (cudaDeviceSynchronize() != cudaSuccess) return nullptr;
  untranspose_pairs_kernel<<<ew_grid(p->M, 256, 2), 256>>>(reinterpret_cast<const float2*>(s->d_tim),
                                                           reinterpret_cast<float2*>(s->d_tim_nat), p->M, s->col_len, s->lg_n2);
  g_launches.fetch_add(1);
  if (cudaDeviceSynchronize() != cudaSuccess) return nullptr;
  return s->d_tim_nat;
}
const float* ps_search_last_spectrum(ps_search* s) { return s ? SS(s)->last_spectrum : nullptr; }

// Index map of the PRODUCT resampling gather (diagnostics): the loader of ps_search_accel_batch is
// run over a series whose sample i holds the 32-bit pattern of i, so d_idx[i] = map(i).
int ps_search_resample_index_map(ps_search* hs, float acc, uint32_t* d_idx, unsigned* h_exact_pairs, void* stream) {
  Search* s = SS(hs);
  if (!s || !d_idx) return invalid("ps_search_resample_index_map: null argument");
  cudaStream_t st = S(stream);
  const ps_fft_plan_t* p = s->plan;
  const unsigned size = s->prm.size;
  float2* ramp = nullptr;
  float2* ramp_ph = nullptr;
  unsigned* d_cnt = nullptr;
  PS_CUDA(cudaMalloc(&ramp, sizeof(float) * (size_t)size), "Error from device_malloc");
  if (cudaMalloc(&ramp_ph, sizeof(float2) * (size_t)size) != cudaSuccess || cudaMalloc(&d_cnt, sizeof(unsigned)) != cudaSuccess) {
    cudaFree(ramp);
    cudaFree(ramp_ph);
    return cuda_fail(cudaGetLastError(), "Error from device_malloc");
  }
  cudaMemsetAsync(d_cnt, 0, sizeof(unsigned), st);
  ramp_pairs_kernel<<<ew_grid(p->M, 256, 2), 256, 0, st>>>(ramp, p->M, s->col_len, s->lg_n2);
  phase_planes_kernel<<<ew_grid(p->M, 256, 4), 256, 0, st>>>(ramp, ramp_ph, p->M, s->col_len);  // as ps_search_whiten does
  g_launches.fetch_add(1);
  LoadResampled ld;  // set up exactly as ps_search_accel_batch does for a one-trial batch
  setup_resampler(ld, s, ramp, ramp_ph, &acc, 1);
  loader_dump_kernel<<<ew_grid(p->M, 256, 2), 256, 0, st>>>(ld, reinterpret_cast<float2*>(d_idx), p->M, p->n2, d_cnt);
  g_launches.fetch_add(1);
  unsigned cnt = 0;
  cudaError_t e = cudaMemcpyAsync(&cnt, d_cnt, sizeof(unsigned), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(ramp);
  cudaFree(ramp_ph);
  cudaFree(d_cnt);
  if (e != cudaSuccess) return cuda_fail(e, "ps_search_resample_index_map");
  if (h_exact_pairs) *h_exact_pairs = cnt;
  return PS_OK;
}

}  // extern "C"
