// pass2spec.cuh -- pass 2 of the packed transform and the spectrum stage in ONE persistent kernel,
// so that the packed spectrum Z never travels to DRAM (included by search.cu after spectrum_pairs).
//
// The split X_k = f(Z_k, Z_{M-k}) needs rows k1 and n1-k1 of the pass-2 output and the interbin
// needs row k1-1, so a pass-2 tile cannot finish its own bins.  Instead the tiles are produced in
// zigzag order (0, last, 1, last-1, ...) into a RING of tile slots stored row-major (Z'[k1][k2],
// each slot = F rows), and every work position also runs one quarter of a spectrum item whose
// rows were produced `lag` positions earlier:
//
//   position Q:  pass-2 tile  zigzag(Q mod ntiles) of trial Q / ntiles      -> ring slot Q mod R
//                spectrum quarter-item Q - lag - 5 (item j = 16 rows [8j,8j+8) + mirror, all k2 of
//                one quarter), which reads the tiles of positions 4j-2 .. 4j+5 of its trial
//
// A CTA publishes "all my positions < p are complete" in progress[cta]; before the spectrum part
// of position Q every CTA checks that all positions <= Q - lag are complete (148 loads spread over
// 148 threads, issued before the pass-2 part so their latency is hidden).  lag = 2 * gridDim, so
// in steady state nobody waits, and with R >= 2 lag + gridDim + 16 the same check (made one
// iteration earlier) also guarantees that the slot a tile overwrites has no reader left.  The ring
// (R x 64 KiB) is rewritten in place while it is still in L2.
//
// STATUS: experiment, opt-in (PS_P2S=1).  Parity-green (the 2^23 comparison against the reference
// build passes through it) and ncu confirms the intent -- 1 MB of DRAM writes and 36 MB/trial of
// DRAM reads per launch, the ring never leaves L2 -- but it is SLOWER than the two separate
// kernels: 31.5 vs 21.0 us/trial for this portion of the chain (pass-2 part alone 17.5 vs 12.5,
// spectrum part alone 19.9 vs 8.5).  With one 512-thread CTA per SM nothing overlaps the serial
// latencies of a position: the progress probe, the 22 L2 loads per thread of the spectrum part,
// its scattered 32-byte stores and the fence before publishing.  The standalone spectrum kernel
// hides the same latencies with 32 warps per SM and runs at 85 % of its own DRAM roofline.
// dbg: 1 = no spectrum part, 2 = no wait, 4 = no pass-2 part (timing experiments only).
#pragma once

namespace ps {

struct SpecArgs {
  float* P;        // normalised spectra, trial b at P + b * pstride
  size_t pstride;
  const float* stats;
  float fsize;
  const cf* TC;
  const cf* TD;
  float2 w1;
};

struct StoreRing {  // rows of one tile, row-major: base[f * n2 + k2]
  float2* base;
  int n2;
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
    float2* o = base + (size_t)f * n2 + k0;
#pragma unroll
    for (int j = 0; j < R; j++) o[stride * j] = make_float2(v[j].x, v[j].y);
  }
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float2 ld_cg(const float2* p) { return __ldcg(p); }

// position of tile T in the zigzag order, and back
__device__ __forceinline__ int zz_pos(int T, int ntiles) { return T < ntiles / 2 ? 2 * T : 2 * (ntiles - 1 - T) + 1; }
__device__ __forceinline__ int zz_tile(int q, int ntiles) { return (q & 1) ? ntiles - 1 - (q >> 1) : (q >> 1); }

template <int R3>
__global__ void __launch_bounds__(512, 1) fft_pass2_spec_kernel(const cf* Yt, float2* ring, int ring_slots, int* progress, FftTables t,
                                                                int n1, int n2, int nbatch, int lag, SpecArgs sa, int dbg) {
  using C = psfft::CtaFft<R3, false>;
  constexpr int F = C::F;
  static_assert(F == 4, "the spectrum part indexes 4-row tiles");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ int slots[8];
  const int ntiles = n1 / F;
  const int npos2 = ntiles * nbatch;        // positions carrying a pass-2 tile
  const int total = npos2 + lag + 5;        // the last spectrum quarter-item sits at npos2 - 1 + lag + 5
  const size_t tile_elems = (size_t)n2 * F;
  constexpr unsigned kTileBytes = sizeof(cf) * 16 * C::THREADS;
  cf* stage = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);
  const int G = gridDim.x;
  const unsigned M = (unsigned)n1 * (unsigned)n2;
  const float mean = __fmul_rn(sa.stats[0], sa.fsize);
  const float inv_sigma = __fdiv_rn(1.0f, __fmul_rn(sa.stats[2], sa.fsize));
  int lg_n1 = 0;
  while ((1 << lg_n1) < n1) lg_n1++;

  auto yt_tile = [&](int Q) {
    const int b = Q / ntiles, q = Q - b * ntiles;
    return Yt + ((size_t)b * ntiles + zz_tile(q, ntiles)) * tile_elems;
  };
  int Q = blockIdx.x;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_proxy_async();
    if (Q < npos2) {
      mbar_expect_tx(&bar, kTileBytes);
      bulk_g2s(stage, yt_tile(Q), kTileBytes, &bar);
    }
  }
  __syncthreads();
  unsigned parity = 0;
  for (; Q < total; Q += G) {
    // all positions <= need must be complete before this position's spectrum part
    const int need = Q - lag;
    int want = 0, seen = 0x7fffffff;
    if ((int)threadIdx.x < G && need >= (int)threadIdx.x) {
      const int c = threadIdx.x;
      want = need - ((need - c) % G) + 1;  // one past the last position <= need that CTA c owns
      seen = ld_acquire(&progress[c]);
    }
    if (Q < npos2 && !(dbg & 4)) {  // ---- pass-2 tile
      StoreRing st{ring + (size_t)(Q % ring_slots) * tile_elems, n2};
      mbar_wait(&bar, parity);
      parity ^= 1u;
      cf v[16];
#pragma unroll
      for (int a = 0; a < 16; a++) v[a] = stage[a * C::THREADS + threadIdx.x];
      C::phaseA_regs(threadIdx.x, smem, t.W4096, v);
      __syncthreads();
      const int nextQ = Q + G;
      if (threadIdx.x == 0 && nextQ < npos2) {
        fence_proxy_async();
        mbar_expect_tx(&bar, kTileBytes);
        bulk_g2s(stage, yt_tile(nextQ), kTileBytes, &bar);
      }
      C::phaseB(threadIdx.x, smem, t.W4096, st);
      __syncthreads();
      C::phaseC(threadIdx.x, smem, st);
    }
    // ---- wait for the producers of this position's spectrum rows (normally already true)
    while (!(dbg & 2) && !__syncthreads_and(seen >= want)) {
      __nanosleep(200);
      if ((int)threadIdx.x < G && need >= (int)threadIdx.x) seen = ld_acquire(&progress[threadIdx.x]);
    }
    const int Qs = Q - lag - 5;  // spectrum quarter-item index
    if (Qs >= 0 && Qs < npos2 && !(dbg & 1)) {
      const int b = Qs / ntiles, qs = Qs - b * ntiles;
      const int j = qs >> 2, quarter = qs & 3;
      float* __restrict__ P = sa.P + (size_t)b * sa.pstride;
      // ring slots of the four ascending tiles 2j-1 .. 2j+2 (slots[0..3]) and of the four
      // descending tiles ntiles-2j .. ntiles-2j-3 (slots[4..7]), indices modulo ntiles
      if (threadIdx.x < 8) {
        int T = threadIdx.x < 4 ? 2 * j - 1 + (int)threadIdx.x : ntiles - 2 * j - ((int)threadIdx.x - 4);
        T = (T + ntiles) % ntiles;
        slots[threadIdx.x] = (b * ntiles + zz_pos(T, ntiles)) % ring_slots;
      }
      __syncthreads();
      const int kq = n2 / 4;
      for (int k2 = quarter * kq + (int)threadIdx.x; k2 < (quarter + 1) * kq; k2 += C::THREADS) {
        const unsigned k0 = 8u * (unsigned)j + ((unsigned)k2 << lg_n1);
        float2 zk[11], zm[11];
        if (j > 0) {
          // row of pair t: k1 = 8j-1+t at column k2; mirror row n1-8j+1-t at column n2-1-k2
          const int k2m = n2 - 1 - k2;
#pragma unroll
          for (int tt = 0; tt < 11; tt++) {
            const int ia = (tt + 3) >> 2, ra = (tt + 3) & 3;  // ascending tile index 0..3, row in tile
            zk[tt] = ld_cg(ring + ((size_t)slots[ia] * F + ra) * n2 + k2);
            const int im = (tt + 2) >> 2, rm = (5 - tt) & 3;  // descending tile index 0..3, row in tile
            zm[tt] = ld_cg(ring + ((size_t)slots[4 + im] * F + rm) * n2 + k2m);
          }
        } else {
          // first item: rows wrap (bin k0-1 is row n1-1 of column k2-1, row 0 mirrors onto itself)
          auto zat = [&](unsigned k) {
            const int k1 = (int)(k & (unsigned)(n1 - 1)), kk2 = (int)(k >> lg_n1);
            const int slot = (b * ntiles + zz_pos(k1 / F, ntiles)) % ring_slots;
            return ld_cg(ring + ((size_t)slot * F + (k1 % F)) * n2 + kk2);
          };
#pragma unroll
          for (int tt = 0; tt < 11; tt++) {
            if (tt == 0 && k0 == 0) {
              zk[0] = zm[0] = make_float2(0.f, 0.f);
              continue;
            }
            zk[tt] = zat(k0 - 1 + tt);
            const unsigned km = M - (k0 - 1 + tt);
            zm[tt] = zat(km == M ? 0u : km);
          }
        }
        spectrum_pairs(zk, zm, twiddle_N(sa.TC, sa.TD, k0), sa.w1, k0, M, mean, inv_sigma, P);
      }
    }
    // ---- publish: everything this CTA wrote for positions <= Q is visible
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      st_release(&progress[blockIdx.x], Q + 1);
    }
  }
}

}  // namespace ps
