// rfft.cuh -- fused real-input transform of the acceleration search: resampling gather -> R2C ->
// interbinned, normalised amplitude spectrum in TWO kernels, with no complex spectrum in memory.
//
// The N = L1*L2 real samples are viewed as an L1 x L2 matrix (sample i = j1*L2 + j2).
//   pass 1 (rfft_pass1_kernel): two adjacent real columns (j2, j2+1) form one complex column
//     -- exactly the sample pairs of the transposed whitened series -- so a length-L1 complex FFT
//     over j1 transforms both; the two real-column spectra are separated inside the CTA from the
//     pair (k1, L1-k1) of the SAME column (shared memory), multiplied by W_N^(k1*j2) and stored for
//     rows k1 = 0..L1/2 only (Hermitian symmetry) in the tiled inter-pass layout.
//   pass 2 (rfft_pass2_kernel): a length-L2 complex FFT over j2 of row k1 yields X[k1 + L1*k2]
//     directly: bins of residue k1 for k2 < L2/2 and, conjugated, bins of residue L1-k1 for
//     k2 >= L2/2 (N - k = (L1-k1) + L1*(L2-1-k2)).  No Hermitian pairing across tiles is needed; the
//     interbin term needs the left neighbour bin only, i.e. the adjacent ROW of the same tile, or
//     -- for the first (direct part) / last (mirror part) row of a tile -- the neighbouring tile,
//     which the persistent CTA carries in shared memory from one tile to the next.
// Compared with the packed transform + spectrum_kernel this removes the 4N-byte packed spectrum
// (written and read back) and one kernel per launch.
#pragma once
#include "fft.cuh"

namespace ps {

struct RfftGeom {
  int L1, L2;      // real rows / real columns, N = L1*L2
  int C, lgC;      // complex (pair) columns = L2/2
  int r1, r2;      // CtaFft radices: L1 = 256*r1, L2 = 256*r2
  int tw;          // rows per pass-2 tile = 32/r2
  int ntiles;      // pass-2 tiles per transform = L1/2/tw + 1 (the last holds row L1/2 only)
  size_t ystride;  // complex words of the tiled inter-pass buffer per transform = ntiles*L2*tw
};

// supported: 2^20 <= N <= 2^23 (L2 in {1024, 2048}, L1 in {1024 .. 4096}); others use the packed path
inline bool rfft_geometry(unsigned n, RfftGeom* g) {
  int lg = 0;
  while ((1u << lg) < n) lg++;
  if ((1u << lg) != n || lg < 20 || lg > 23) return false;
  int l2 = lg / 2;
  if (l2 > 11) l2 = 11;
  g->L2 = 1 << l2;
  g->L1 = 1 << (lg - l2);
  g->C = g->L2 / 2;
  g->lgC = l2 - 1;
  g->r1 = g->L1 / 256;
  g->r2 = g->L2 / 256;
  g->tw = 32 / g->r2;
  g->ntiles = g->L1 / 2 / g->tw + 1;
  g->ystride = (size_t)g->ntiles * g->L2 * g->tw;
  return true;
}

// W_N^p from the plan's W_N tables (TC: W_N^(4096 q), TD: W_N^q)
__device__ __forceinline__ cf twiddle_Nr(const cf* __restrict__ TC, const cf* __restrict__ TD, unsigned p) {
  return psfft::cmul(psfft::ldg_cf(&TC[p >> 12]), psfft::ldg_cf(&TD[p & 4095u]));
}

// phase-C storer of pass 1: the column spectra W[k] go to the staging tile S[k*F + f] (E1 region)
struct StageColumns {
  cf* S;
  int nf;
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
#pragma unroll
    for (int j = 0; j < R; j++) S[(k0 + stride * j) * nf + f] = v[j];
  }
};

// One (k1, f) item of the separation: W[k1], W[L1-k1] of complex column c0+f -> the spectra of the
// real columns j2 = 2(c0+f) and j2+1 at row k1, times W_N^(k1*j2), into Yr[k1/tw][j2][k1%tw].
__device__ __forceinline__ void separate_item(const cf* __restrict__ S, int nf, int L1, int k1, int f, int c0, cf* __restrict__ Yr,
                                              int L2, int tw, const cf* __restrict__ TC, const cf* __restrict__ TD) {
  const cf w = S[k1 * nf + f];
  const cf m = S[((L1 - k1) & (L1 - 1)) * nf + f];
  // A = (W + conj M)/2 ; B = -i (W - conj M)/2
  const cf a = psfft::mk(0.5f * (w.x + m.x), 0.5f * (w.y - m.y));
  const cf b = psfft::mk(0.5f * (w.y + m.y), -0.5f * (w.x - m.x));
  const unsigned j2 = 2u * (unsigned)(c0 + f);
  const cf ta = twiddle_Nr(TC, TD, (unsigned)k1 * j2);
  const cf tb = psfft::cmul(ta, twiddle_Nr(TC, TD, (unsigned)k1));
  const cf ya = psfft::cmul(a, ta), yb = psfft::cmul(b, tb);
  float2* o = reinterpret_cast<float2*>(Yr) + ((size_t)(k1 / tw) * L2 + j2) * tw + (k1 % tw);
  o[0] = make_float2(ya.x, ya.y);
  o[tw] = make_float2(yb.x, yb.y);
}

template <int R3>
__global__ void __launch_bounds__(512, 1) rfft_pass1_kernel(LoadResampled ld, cf* Yr, RfftGeom g, const cf* W4096, const cf* TC,
                                                            const cf* TD, int nbatch) {
  using C = psfft::CtaFft<R3, false>;
  extern __shared__ float smem[];
  const int ntiles = g.C / C::F;
  const int nwork = ntiles * nbatch;
  int work = blockIdx.x;
  if (work >= nwork) return;
  cf* S = reinterpret_cast<cf*>(smem);  // E1 region, free after phase B
  cf v[16];
  ld.begin(work / ntiles, (work % ntiles) * C::F, g.C);
  load_operands<C>(ld, v);
  for (; work < nwork; work += gridDim.x) {
    const int b = work / ntiles, tile = work - b * ntiles;
    const int c0 = tile * C::F;
    StageColumns st{S, C::F};
    C::phaseA_compute(threadIdx.x, W4096, v);
    __syncthreads();  // the previous item's separation has finished reading S
    C::phaseA_write(threadIdx.x, smem, v);
    __syncthreads();
    C::phaseB(threadIdx.x, smem, W4096, st);
    __syncthreads();
    const int next = work + gridDim.x;
    if (next < nwork) {
      ld.begin(next / ntiles, (next % ntiles) * C::F, g.C);
      load_operands<C>(ld, v);
    }
    C::phaseC(threadIdx.x, smem, st);
    __syncthreads();
    cf* Yb = Yr + (size_t)b * g.ystride;
    // (L1/2)*F items in rounds of 512, then row L1/2
    const int items = (g.L1 / 2) * C::F;
    for (int id = threadIdx.x; id < items; id += C::THREADS)
      separate_item(S, C::F, g.L1, id / C::F, id % C::F, c0, Yb, g.L2, g.tw, TC, TD);
    if (threadIdx.x < C::F) separate_item(S, C::F, g.L1, g.L1 / 2, threadIdx.x, c0, Yb, g.L2, g.tw, TC, TD);
  }
}

// phase-C storer of pass 2: X of row f at column k2 -> XS[k2*F + f] (E1 region)
struct StageRows {
  cf* XS;
  int nf;
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
#pragma unroll
    for (int j = 0; j < R; j++) XS[(k0 + stride * j) * nf + f] = v[j];
  }
};

__device__ __forceinline__ float sqrt_mufu(float x) {
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// normalised interbinned amplitude of bin X with left neighbour L (kernels.cu:231-252, :469-480);
// MUFU square root and reciprocal multiply: <= 2 ulp from the staged IEEE entry points
__device__ __forceinline__ float bin_value(cf x, cf l, float mean, float inv_sigma) {
  const float amp = x.x * x.x + x.y * x.y;
  const float dr = x.x - l.x, di = x.y - l.y;
  return (sqrt_mufu(fmaxf(amp, 0.5f * (dr * dr + di * di))) - mean) * inv_sigma;
}

// Spectrum epilogue of one pass-2 tile (rows k1 = kt*tw + f).  XS holds X[k1 + L1*k2] of the tile;
// carryD / carryM hold the previous tile's last row (direct half k2 < H / mirror half, conjugated).
// halo = true only refreshes the carries (run prologue).  Every thread owns fixed columns k2, so the
// carries need no barrier.
template <int F>
__device__ __forceinline__ void tile_epilogue(const cf* __restrict__ XS, cf* __restrict__ carryD, cf* __restrict__ carryM,
                                              float* __restrict__ P, const RfftGeom& g, int kt, bool halo, float mean,
                                              float inv_sigma) {
  const int L1 = g.L1, L2 = g.L2, H = L2 / 2;
  const bool last = kt == g.ntiles - 1;
  const int nrows = last ? 1 : F;
  const int k1_0 = kt * F;
  // ---- direct half: bins k1 + L1*k2, k2 < H (plus the Nyquist bin: row 0, k2 = H)
  for (int k2 = threadIdx.x; k2 < H + (kt == 0 ? 1 : 0); k2 += 512) {
    cf x[F];
#pragma unroll
    for (int f = 0; f < F; f++) x[f] = XS[k2 * F + f];
    cf left;
    if (kt == 0) {
      // bin L1*k2 - 1 is the conjugate of row 1 at column L2 - k2; bin 0 has no left neighbour
      left = k2 ? psfft::cconj(XS[(L2 - k2) * F + 1]) : psfft::mk(0.f, 0.f);
    } else {
      left = carryD[k2];
    }
    if (!halo) {
      const int nr = k2 == H ? 1 : nrows;
      float* o = P + (size_t)k1_0 + (size_t)L1 * k2;
#pragma unroll
      for (int f = 0; f < F; f++) {
        if (f < nr) o[f] = bin_value(x[f], left, mean, inv_sigma);
        left = x[f];
      }
    }
    if (!last && k2 < H) carryD[k2] = x[F - 1];
  }
  // ---- mirror half: column k2 >= H of row k1 >= 1 is the conjugate of bin (L1-k1) + L1*(L2-1-k2)
  for (int k2 = H + threadIdx.x; k2 < L2; k2 += 512) {
    cf x[F];
#pragma unroll
    for (int f = 0; f < F; f++) x[f] = psfft::cconj(XS[k2 * F + f]);
    float* o = P + (size_t)L1 * (L2 - 1 - k2);
    if (!halo) {
      if (kt > 0) {
        // the previous tile's last row: its left neighbour is this tile's row 0
        o[L1 - k1_0 + 1] = bin_value(carryM[k2 - H], x[0], mean, inv_sigma);
      }
      if (last) {
        // row L1/2: left neighbour = direct bin of row L1/2-1 at column L2-1-k2
        o[L1 - k1_0] = bin_value(x[0], carryD[L2 - 1 - k2], mean, inv_sigma);
      } else {
#pragma unroll
        for (int f = 0; f < F - 1; f++)
          if (k1_0 + f > 0) o[L1 - k1_0 - f] = bin_value(x[f], x[f + 1], mean, inv_sigma);
      }
    }
    if (!last) carryM[k2 - H] = x[F - 1];
  }
}

// Persistent CTAs over contiguous runs of (transform, tile) items; a run that starts inside a
// transform first recomputes the preceding tile in halo mode to obtain the carries.
template <int R3>
__global__ void __launch_bounds__(512, 1) rfft_pass2_kernel(const cf* __restrict__ Yr, float* __restrict__ Pall, size_t pstride,
                                                            RfftGeom g, const cf* W4096, const float* __restrict__ stats,
                                                            float fsize, int nbatch) {
  using C = psfft::CtaFft<R3, false>;
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long bar;
  constexpr int F = C::F;
  constexpr unsigned kTileBytes = sizeof(cf) * 16 * C::THREADS;
  cf* in_stage = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);
  cf* carryD = in_stage + 16 * C::THREADS;
  cf* carryM = carryD + g.L2 / 2;
  cf* XS = reinterpret_cast<cf*>(smem);  // E1 region
  const long T = (long)nbatch * g.ntiles;
  const long s = (long)blockIdx.x * T / gridDim.x, e = (long)(blockIdx.x + 1) * T / gridDim.x;
  if (s >= e) return;
  const long first = (s % g.ntiles) ? s - 1 : s;  // halo prologue when the run starts mid-transform
  const size_t tile_elems = (size_t)g.L2 * F;
  auto tile_src = [&](long it) { return Yr + (size_t)(it / g.ntiles) * g.ystride + (size_t)(it % g.ntiles) * tile_elems; };
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_proxy_async();
    mbar_expect_tx(&bar, kTileBytes);
    bulk_g2s(in_stage, tile_src(first), kTileBytes, &bar);
  }
  __syncthreads();
  const float mean = __fmul_rn(stats[0], fsize);
  const float inv_sigma = __fdiv_rn(1.0f, __fmul_rn(stats[2], fsize));
  unsigned parity = 0;
  StageRows st{XS, F};
  for (long it = first; it < e; it++) {
    const int b = (int)(it / g.ntiles), kt = (int)(it % g.ntiles);
    mbar_wait(&bar, parity);
    parity ^= 1u;
    cf v[16];
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = in_stage[a * C::THREADS + threadIdx.x];
    C::phaseA_compute(threadIdx.x, W4096, v);
    __syncthreads();  // operands taken; the previous tile's epilogue has finished reading XS
    if (threadIdx.x == 0 && it + 1 < e) {
      fence_proxy_async();
      mbar_expect_tx(&bar, kTileBytes);
      bulk_g2s(in_stage, tile_src(it + 1), kTileBytes, &bar);
    }
    C::phaseA_write(threadIdx.x, smem, v);
    __syncthreads();
    C::phaseB(threadIdx.x, smem, W4096, st);
    __syncthreads();
    C::phaseC(threadIdx.x, smem, st);
    __syncthreads();
    tile_epilogue<F>(XS, carryD, carryM, Pall + (size_t)b * pstride, g, kt, it < s, mean, inv_sigma);
  }
}

// ------------------------------------------------------------------ launchers
template <int R3>
inline int launch_rfft_pass1_r(const ps_fft_plan_t* p, const RfftGeom& g, const LoadResampled& ld, cf* Yr, cudaStream_t st,
                               int nbatch) {
  using C = psfft::CtaFft<R3, false>;
  auto k = rfft_pass1_kernel<R3>;
  static SmemOptIn done[64];
  int rc = configure_smem(k, C::SMEM_BYTES, p->device, done);
  if (rc) return rc;
  k<<<fft_grid(g.C / C::F * nbatch), C::THREADS, C::SMEM_BYTES, st>>>(ld, Yr, g, p->W4096, p->TC, p->TD, nbatch);
  PS_LAUNCHED("rfft_pass1_kernel");
  return PS_OK;
}

template <int R3>
inline int launch_rfft_pass2_r(const ps_fft_plan_t* p, const RfftGeom& g, const cf* Yr, float* P, size_t pstride,
                               const float* stats, float fsize, cudaStream_t st, int nbatch) {
  using C = psfft::CtaFft<R3, false>;
  auto k = rfft_pass2_kernel<R3>;
  static SmemOptIn done[64];
  const size_t smem = C::SMEM_BYTES + sizeof(cf) * 16 * C::THREADS + sizeof(cf) * g.L2;
  int rc = configure_smem(k, smem, p->device, done);
  if (rc) return rc;
  k<<<fft_grid(g.ntiles * nbatch), C::THREADS, smem, st>>>(Yr, P, pstride, g, p->W4096, stats, fsize, nbatch);
  PS_LAUNCHED("rfft_pass2_kernel");
  return PS_OK;
}

inline int launch_rfft_pass1(const ps_fft_plan_t* p, const RfftGeom& g, const LoadResampled& ld, cf* Yr, cudaStream_t st,
                             int nbatch) {
  switch (g.r1) {
    case 4: return launch_rfft_pass1_r<4>(p, g, ld, Yr, st, nbatch);
    case 8: return launch_rfft_pass1_r<8>(p, g, ld, Yr, st, nbatch);
    case 16: return launch_rfft_pass1_r<16>(p, g, ld, Yr, st, nbatch);
    default: return invalid("rfft pass 1: unsupported column length");
  }
}

inline int launch_rfft_pass2(const ps_fft_plan_t* p, const RfftGeom& g, const cf* Yr, float* P, size_t pstride,
                             const float* stats, float fsize, cudaStream_t st, int nbatch) {
  switch (g.r2) {
    case 4: return launch_rfft_pass2_r<4>(p, g, Yr, P, pstride, stats, fsize, st, nbatch);
    case 8: return launch_rfft_pass2_r<8>(p, g, Yr, P, pstride, stats, fsize, st, nbatch);
    default: return invalid("rfft pass 2: unsupported row length");
  }
}

}  // namespace ps
