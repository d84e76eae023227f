// resample_tile.cuh -- pass 1 of the search transform with the acceleration resampling gather
// (kernels.cu:314-317, 335-346: out[i] = in[rn(i + i*af*(i-N))]) staged through shared memory.
//
// A pass-1 work item is F adjacent columns of the packed series: row r of the tile holds the 2F
// consecutive samples i0 .. i0+2F-1, i0 = 2*(r*n2 + c0).  The shift i -> map(i) - i is smooth
// (|d shift / d i| <= |af| N, below 1e-3 for every survey configuration), so one float32 estimate
// per ROW -- taken at the row's middle sample -- fixes the rounded shift of all 2F samples unless it
// lies within its error bound of a half-integer; those rows (and only those) evaluate the
// reference's FP64 expression sample by sample.  The row is then copied global -> shared with
// cp.async (LDGSTS): F 8-byte copies when the shift is even, 2F 4-byte copies when it is odd.
// Thread = row, so a warp's copies read 32 consecutive rows of one column of the TRANSPOSED whitened
// series (one 256-byte run) and write 32 consecutive shared-memory words.  The copies of the next work
// item are in flight while the current one is transformed; the FFT warps read their 16 operands
// from the tile with 16 LDS.64.  Per work item and thread this replaces 16 estimates + 16..32
// dependent global loads by n1/512 estimates and 16 shared loads.
#pragma once
#include "fft.cuh"

namespace ps {

struct ResampleArgs {
  const float2* D;   // whitened series as pairs, transposed: pair e at D[(e % n2) * n1 + e / n2]
  const float2* Dp;  // the same series one sample on: pair e = (x[2e+1], x[2e+2]), same layout; or null
  AccelFacts afs;    // accel_fact per batch item
  double size;       // N
  float max_shift;   // host bound on |i*af*(i-N)|, all batch items
  float max_step;    // host bound on |shift(i+1) - shift(i)|
  int n1, lg_n2;
  unsigned* exact_rows;  // diagnostics: counts rows that took the FP64 path (null on the product path)
};

template <int R3>
struct ResampleTile {
  using C = psfft::CtaFft<R3, false>;
  static constexpr int F = C::F;
  static constexpr int N = C::N;                      // rows of the tile = n1
  static constexpr int PAD = F >= 16 ? 1 : 16 / F;    // column stride = PAD mod 16 words: the FFT warps'
  static constexpr int COL = N + PAD;                 // reads (F columns x 32/F rows) are conflict-free
  static constexpr int WORDS = F * COL;               // complex words
  static constexpr size_t BYTES = sizeof(cf) * WORDS;
  static constexpr size_t SMEM_BYTES = C::SMEM_BYTES + BYTES;
  static constexpr bool kFits = SMEM_BYTES <= 227 * 1024;

  // exact path: every sample of row r on its own (the reference's FP64 expression)
  static __device__ __noinline__ void row_exact(const float* Df, int lg_n2, double size, unsigned* exact_rows, double af, unsigned i0,
                                                float* dst) {
    const unsigned col_mask = (1u << lg_n2) - 1u;
    if (exact_rows) atomicAdd(exact_rows, 1u);
#pragma unroll 1
    for (int j = 0; j < 2 * F; j++) {
      const unsigned idx = (unsigned)resample_index_II(af, size, i0 + (unsigned)j);
      const unsigned e = idx >> 1;
      const unsigned p = (e & col_mask) * (unsigned)N + (e >> lg_n2);
      cp_async4(dst + 2 * ((j >> 1) * COL) + (j & 1), Df + 2u * p + (idx & 1u));
    }
  }

  // uniform shift, but the row's source pairs run past the last column of the transposed layout
  static __device__ __noinline__ void row_wrapped(const float* Df, int lg_n2, unsigned s0, float* dst) {
    const unsigned col_mask = (1u << lg_n2) - 1u;
#pragma unroll 1
    for (int j = 0; j < 2 * F; j++) {
      const unsigned idx = s0 + (unsigned)j;
      const unsigned e = idx >> 1;
      const unsigned p = (e & col_mask) * (unsigned)N + (e >> lg_n2);
      cp_async4(dst + 2 * ((j >> 1) * COL) + (j & 1), Df + 2u * p + (idx & 1u));
    }
  }

  // Queue the copies of tile (columns c0 .. c0+F-1) of the series resampled with af; one commit group.
  static __device__ __forceinline__ void issue(const ResampleArgs& a, double af, int c0, int n2, cf* tile) {
    issue_rows<512>(a, af, c0, n2, tile, (int)threadIdx.x);
  }
  // NT threads (tid = 0 .. NT-1) share the rows of the tile
  template <int NT>
  static __device__ __forceinline__ void issue_rows(const ResampleArgs& a, double af, int c0, int n2, cf* tile, int tid) {
    const float af32 = (float)af, fsize = (float)a.size;
    // float32 error of the estimate: the middle sample index and (index - N) are exact (even, below
    // 2^25), three roundings give <= 1.8e-7 |shift| (3e-7 with margin); the other samples of the row
    // lie within F steps of the middle one; the reference's own FP64 evaluation is good to ~1e-12 of
    // the shift (1e-5 covers it)
    const float thr = 0.5f - (1e-5f + 3e-7f * a.max_shift + (float)F * a.max_step);
    const bool dual = a.Dp != nullptr;
    const unsigned col_mask = (1u << a.lg_n2) - 1u;
    constexpr int ROWS = (N + NT - 1) / NT;   // rows per thread
    constexpr int GB = ROWS < 4 ? ROWS : 4;   // rows per group: all source addresses of a group are
                                              // computed before its copies are queued (a queued copy
                                              // holds its address registers until it is served)
#pragma unroll 1
    for (int k0 = 0; k0 < ROWS; k0 += GB) {
      const float2* src[GB];
      unsigned mode[GB];  // 0: F 8-byte copies from src; 1: odd shift, 2F 4-byte copies; 2: wrapped (s0); 3: exact
      unsigned s0v[GB];
#pragma unroll
      for (int g = 0; g < GB; g++) {
        const int r = tid + NT * (k0 + g);
        const unsigned i0 = 2u * ((unsigned)r * (unsigned)n2 + (unsigned)c0);
        const float fm = (float)(i0 + (unsigned)F);
        const float gs = (fm * (fm - fsize)) * af32;
        const float tm = gs + 12582912.f;  // 1.5*2^23: rounds the estimate to an integer (ties to even)
        const float rr = tm - 12582912.f;
        const unsigned s0 = i0 + (unsigned)(__float_as_int(tm) - 0x4B400000);  // first source sample
        const unsigned e = s0 >> 1, col = e & col_mask;
        const unsigned off = col * (unsigned)N + (e >> a.lg_n2);
        const bool odd = (s0 & 1u) != 0;
        src[g] = ((dual && odd) ? a.Dp : a.D) + off;
        s0v[g] = s0;
        mode[g] = !(fabsf(gs - rr) < thr) ? 3u : (col + (unsigned)F > col_mask) ? 2u : (odd && !dual) ? 1u : 0u;
        if (N < NT * ROWS && r >= N) mode[g] = 4u;  // no such row
      }
#pragma unroll
      for (int g = 0; g < GB; g++) {
        const int r = tid + NT * (k0 + g);
        cf* dst = tile + r;
        if (mode[g] == 0u) {
#pragma unroll
          for (int c = 0; c < F; c++) cp_async8(dst + c * COL, src[g] + c * N);
        } else if (mode[g] == 1u) {
#pragma unroll
          for (int c = 0; c < F; c++) {
            cp_async4(&dst[c * COL].x, &src[g][c * N].y);
            cp_async4(&dst[c * COL].y, &src[g][(c + 1) * N].x);
          }
        } else if (mode[g] == 2u) {
          row_wrapped(reinterpret_cast<const float*>(a.D), a.lg_n2, s0v[g], reinterpret_cast<float*>(dst));
        } else if (mode[g] == 3u) {
          const unsigned i0 = 2u * ((unsigned)r * (unsigned)n2 + (unsigned)c0);
          row_exact(reinterpret_cast<const float*>(a.D), a.lg_n2, a.size, a.exact_rows, af, i0, reinterpret_cast<float*>(dst));
        }
      }
    }
    cp_async_commit();
  }

  // the 16 operands of FFT thread t = u*F + f: rows u + a*G of column f
  static __device__ __forceinline__ void take(const cf* tile, cf* v) {
    const int f = threadIdx.x % F, u = threadIdx.x / F;
    const cf* p = tile + f * COL + u;
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = p[a * C::G];
  }
};

// pass 1 over (batch item, tile) work items with the resampling gather staged in shared memory
template <int R3>
__global__ void __launch_bounds__(512, 1) fft_pass1_resample_kernel(ResampleArgs ra, cf* Yt, FftTables t, int n2, int nbatch, int tw2) {
  using T = ResampleTile<R3>;
  using C = typename T::C;
  extern __shared__ __align__(16) float smem[];
  cf* tile = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);
  const int ntiles = n2 / C::F;
  const int nwork = ntiles * nbatch;
  const size_t M = (size_t)C::N * n2;
  int work = blockIdx.x;
  if (work >= nwork) return;
  T::issue(ra, ra.afs.v[work % nbatch], (work / nbatch) * C::F, n2, tile);
  cp_async_wait_all();
  __syncthreads();
  for (;;) {
    const int tl = work / nbatch, b = work - tl * nbatch;
    StorePass1<false> st{Yt + (size_t)b * M, t, C::N, tl * C::F, n2, tw2};
    cf v[16];
    T::take(tile, v);
    C::phaseA_regs(threadIdx.x, smem, t.W4096, v);
    __syncthreads();  // E1 complete; every thread has taken its operands out of the tile
    const int next = work + gridDim.x;
    if (next < nwork) T::issue(ra, ra.afs.v[next % nbatch], (next / nbatch) * C::F, n2, tile);
    C::phaseB(threadIdx.x, smem, t.W4096, st);
    cp_async_wait_all();
    __syncthreads();  // E2 complete; the next tile has landed
    C::phaseC(threadIdx.x, smem, st);
    if (next >= nwork) break;
    work = next;
  }
}

// ---- warp-specialised version: 16 FFT warps + 4 loader warps.  The loader warps run the gather
// of work item w+1 (estimates, cp.async, completion wait) while the FFT warps transform item w, so
// neither the copies' queueing in the load/store unit nor their L2 round trips sit on the FFT warps.
// Hand-over through two named barriers (FULL: loaders arrive, FFT warps wait; EMPTY: the reverse);
// the FFT warps synchronise among themselves on a third.  Registers are re-balanced with setmaxnreg
// (loader warps give theirs up).
__device__ __forceinline__ void named_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void named_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

constexpr int kWsFftThreads = 512, kWsLoadThreads = 128, kWsThreads = kWsFftThreads + kWsLoadThreads;
constexpr int kBarFull = 1, kBarEmpty = 2, kBarFft = 3;

#ifdef PS_PROF
// per-warp cycle counters of the phases (profiling builds only: make EXTRA=-DPS_PROF)
__device__ unsigned g_prof_buf[148 * 20 * 8];
#define PS_TICK(k)                                   \
  {                                                  \
    const unsigned now_ = (unsigned)clock();         \
    pacc[k] += now_ - plast;                         \
    plast = now_;                                    \
  }
#define PS_PROF_BEGIN unsigned pacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, plast = (unsigned)clock();
#define PS_PROF_END                                                                                      \
  if ((threadIdx.x & 31) == 0 && blockIdx.x < 148)                                                       \
    for (int k_ = 0; k_ < 8; k_++) g_prof_buf[(blockIdx.x * 20 + threadIdx.x / 32) * 8 + k_] = pacc[k_];
#else
#define PS_TICK(k)
#define PS_PROF_BEGIN
#define PS_PROF_END
#endif

template <int R3>
__global__ void __launch_bounds__(kWsThreads, 1) fft_pass1_resample_ws_kernel(ResampleArgs ra, cf* Yt, FftTables t, int n2, int nbatch,
                                                                               int tw2) {
  using T = ResampleTile<R3>;
  using C = typename T::C;
  extern __shared__ __align__(16) float smem[];
  cf* tile = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);
  const int ntiles = n2 / C::F;
  const int nwork = ntiles * nbatch;
  const size_t M = (size_t)C::N * n2;
  if ((int)blockIdx.x >= nwork) return;
  if (threadIdx.x >= kWsFftThreads) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
    const int lt = (int)threadIdx.x - kWsFftThreads;
    bool first = true;
    PS_PROF_BEGIN
    for (int work = blockIdx.x; work < nwork; work += gridDim.x) {
      if (!first) named_sync(kBarEmpty, kWsThreads);  // the FFT warps have taken the previous tile
      first = false;
      PS_TICK(0)
      T::template issue_rows<kWsLoadThreads>(ra, ra.afs.v[work % nbatch], (work / nbatch) * C::F, n2, tile, lt);
      PS_TICK(1)
      cp_async_wait_all();
      PS_TICK(2)
      named_arrive(kBarFull, kWsThreads);
    }
    PS_PROF_END
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
    PS_PROF_BEGIN
    for (int work = blockIdx.x; work < nwork; work += gridDim.x) {
      const int tl = work / nbatch, b = work - tl * nbatch;
      StorePass1<false> st{Yt + (size_t)b * M, t, C::N, tl * C::F, n2, tw2};
      cf v[16];
      named_sync(kBarFull, kWsThreads);
      PS_TICK(0)
      T::take(tile, v);
      if (work + (int)gridDim.x < nwork) named_arrive(kBarEmpty, kWsThreads);
      PS_TICK(1)
      C::phaseA_regs(threadIdx.x, smem, t.W4096, v);
      PS_TICK(2)
      named_sync(kBarFft, kWsFftThreads);
      PS_TICK(3)
      C::phaseB(threadIdx.x, smem, t.W4096, st);
      PS_TICK(4)
      named_sync(kBarFft, kWsFftThreads);
      PS_TICK(5)
      C::phaseC(threadIdx.x, smem, st);
      PS_TICK(6)
    }
    PS_PROF_END
  }
}

template <int R3>
inline int launch_pass1_resample_ws_r(const ps_fft_plan_t* p, const ResampleArgs& ra, cudaStream_t st, int nbatch) {
  using T = ResampleTile<R3>;
  if constexpr (!T::kFits) {
    return invalid("staged resampling pass needs R3 != 2");
  } else {
    auto k = fft_pass1_resample_ws_kernel<R3>;
    static bool done[64] = {};
    int rc = configure_smem(k, T::SMEM_BYTES, p->device, done);
    if (rc) return rc;
    k<<<fft_grid(p->n2 / T::F * nbatch), kWsThreads, T::SMEM_BYTES, st>>>(ra, p->Yt, p->tables(), p->n2, nbatch, 32 / p->r2);
    PS_LAUNCHED("fft_pass1_resample_ws_kernel");
    return PS_OK;
  }
}
inline int launch_pass1_resample_ws(const ps_fft_plan_t* p, const ResampleArgs& ra, cudaStream_t st, int nbatch) {
  switch (p->r1) {
    case 1: return launch_pass1_resample_ws_r<1>(p, ra, st, nbatch);
    case 4: return launch_pass1_resample_ws_r<4>(p, ra, st, nbatch);
    case 8: return launch_pass1_resample_ws_r<8>(p, ra, st, nbatch);
    case 16: return launch_pass1_resample_ws_r<16>(p, ra, st, nbatch);
    default: return invalid("staged resampling pass: unsupported radix plan");
  }
}

// diagnostic: the same gather, tile written back in natural pair order (out pair e = r*n2 + c0 + c)
template <int R3>
__global__ void __launch_bounds__(512, 1) resample_tile_dump_kernel(ResampleArgs ra, float2* out, int n2) {
  using T = ResampleTile<R3>;
  using C = typename T::C;
  extern __shared__ __align__(16) float smem[];
  cf* tile = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);
  const int ntiles = n2 / C::F;
  for (int work = blockIdx.x; work < ntiles; work += gridDim.x) {
    T::issue(ra, ra.afs.v[0], work * C::F, n2, tile);
    cp_async_wait_all();
    __syncthreads();
    for (int q = threadIdx.x; q < C::N * C::F; q += 512) {
      const int r = q / C::F, c = q % C::F;
      const cf v = tile[c * T::COL + r];
      out[(size_t)r * n2 + work * C::F + c] = make_float2(v.x, v.y);
    }
    __syncthreads();
  }
}

template <int R3>
inline int launch_pass1_resample_r(const ps_fft_plan_t* p, const ResampleArgs& ra, cudaStream_t st, int nbatch) {
  using T = ResampleTile<R3>;
  if constexpr (!T::kFits) {
    return invalid("staged resampling pass needs R3 != 2");
  } else {
    auto k = fft_pass1_resample_kernel<R3>;
    static bool done[64] = {};
    int rc = configure_smem(k, T::SMEM_BYTES, p->device, done);
    if (rc) return rc;
    k<<<fft_grid(p->n2 / T::F * nbatch), 512, T::SMEM_BYTES, st>>>(ra, p->Yt, p->tables(), p->n2, nbatch, 32 / p->r2);
    PS_LAUNCHED("fft_pass1_resample_kernel");
    return PS_OK;
  }
}

template <int R3>
inline int launch_tile_dump_r(const ps_fft_plan_t* p, const ResampleArgs& ra, float2* out, cudaStream_t st) {
  using T = ResampleTile<R3>;
  if constexpr (!T::kFits) {
    return invalid("staged resampling pass needs R3 != 2");
  } else {
    auto k = resample_tile_dump_kernel<R3>;
    static bool done[64] = {};
    int rc = configure_smem(k, T::SMEM_BYTES, p->device, done);
    if (rc) return rc;
    k<<<fft_grid(p->n2 / T::F), 512, T::SMEM_BYTES, st>>>(ra, out, p->n2);
    PS_LAUNCHED("resample_tile_dump_kernel");
    return PS_OK;
  }
}

// whether the plan's pass 1 can take the staged gather
inline bool resample_tile_ok(const ps_fft_plan_t* p) { return p->r1 != 2; }

inline int launch_pass1_resample(const ps_fft_plan_t* p, const ResampleArgs& ra, cudaStream_t st, int nbatch) {
  switch (p->r1) {
    case 1: return launch_pass1_resample_r<1>(p, ra, st, nbatch);
    case 4: return launch_pass1_resample_r<4>(p, ra, st, nbatch);
    case 8: return launch_pass1_resample_r<8>(p, ra, st, nbatch);
    case 16: return launch_pass1_resample_r<16>(p, ra, st, nbatch);
    default: return invalid("staged resampling pass: unsupported radix plan");
  }
}
inline int launch_tile_dump(const ps_fft_plan_t* p, const ResampleArgs& ra, float2* out, cudaStream_t st) {
  switch (p->r1) {
    case 1: return launch_tile_dump_r<1>(p, ra, out, st);
    case 4: return launch_tile_dump_r<4>(p, ra, out, st);
    case 8: return launch_tile_dump_r<8>(p, ra, out, st);
    case 16: return launch_tile_dump_r<16>(p, ra, out, st);
    default: return invalid("staged resampling pass: unsupported radix plan");
  }
}

}  // namespace ps
