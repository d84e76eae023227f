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
#include <cuda.h>
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
  int max_batch;  // Yt and Z hold max_batch transforms
  cf* Yt;  // max_batch x M complex: transposed intermediate between the passes
  cf* Z;   // max_batch x M complex: packed-complex spectrum before the real post-process
  cf* TC;  // exp(-i pi 4096 q / M), q <= M/4096   (real-FFT split twiddle W_N^k, N = 2M)
  cf* TD;  // exp(-i pi q / M), q < 4096
  FftTables tables() const { return FftTables{W4096, TA, TB}; }
};

// fft.cu: power-of-two transforms on a core plan; bluestein.cu: any length on top of them
int fft_r2c_core(const ps_fft_plan_t* p, const float* d_in, float2* d_out, cudaStream_t st);
int fft_c2r_core(const ps_fft_plan_t* p, const float2* d_in, float* d_out, cudaStream_t st);
struct Bluestein;
unsigned bluestein_length(unsigned n);
int bluestein_create(unsigned n, const ps_fft_plan_t* core, Bluestein** out);
void bluestein_destroy(Bluestein* b);
int bluestein_r2c(const Bluestein* b, const ps_fft_plan_t* core, const float* d_in, float2* d_out, cudaStream_t st);
int bluestein_c2r(const Bluestein* b, const ps_fft_plan_t* core, const float2* d_in, float* d_out, cudaStream_t st);

// W_M^p for p < M from the two-level table (one complex multiply)
__device__ __forceinline__ cf twiddle_M(const FftTables& t, unsigned p) {
  return psfft::cmul(psfft::ldg_cf(&t.TA[p >> 12]), psfft::ldg_cf(&t.TB[p & 4095u]));
}

// cp.async (LDGSTS): global -> shared without staging through registers
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc));
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ------------------------------------------------------------------ pass-1 loaders
// Loaders with kAsync = true also provide prefetch(f, idx, dst): an asynchronous copy of the
// element into the thread's own staging slot, so the next tile's operands are in flight while
// the current tile is being transformed.
// Every loader/storer has begin(b, first, n): b = batch item (independent transform sharing the
// plan), first = first column (pass 1) / row (pass 2) of the CTA tile, n = n2 (pass 1) / n1 (pass 2).
// complex element e = idx*n2 + c0 + f of the packed series
constexpr int kMaxBatch = 128;

struct LoadPackedReal {  // z[e] = (x[2e], x[2e+1])
  static constexpr bool kAsync = true;
  __device__ __forceinline__ void prefetch(int f, int idx, cf* dst) const {
    cp_async8(dst, &xb[(size_t)idx * n2 + c0 + f]);
  }
  const float2* x;
  size_t item_stride;  // in float2
  int n2, c0;
  const float2* xb;
  __device__ __forceinline__ void begin(int b, int first, int n) {
    xb = x + (size_t)b * item_stride;
    c0 = first;
    n2 = n;
  }
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    float2 v = __ldg(&xb[(size_t)idx * n2 + c0 + f]);
    return psfft::mk(v.x, v.y);
  }
};

struct AccelFacts {
  double v[kMaxBatch];
};

// z[e] = (x[map(2e)], x[map(2e+1)]): the resampling gather of kernels.cu:335-346 fused into the
// FFT load.  The read index is bit-identical to resample_kernelII: a float32 evaluation of the
// shift i*af*(i-N) decides it whenever it lies safely away from a rounding boundary (the shift is
// then rn(shift32)); otherwise the reference's FP64 expression is evaluated.
struct LoadResampled {
  // The whitened series is kept TRANSPOSED as complex pairs: pair e = (x[2e], x[2e+1]) of the natural
  // series lives at D[(e % n2) * n1 + e / n2], so the column a pass-1 tile transforms is a contiguous
  // run in memory.  A warp's load then touches 4 runs of 64 bytes instead of 8 scattered sectors.
  // The loader reads the TWO-PHASE form of that layout: plane 2c holds column c as it is (the pair for an
  // even shift), plane 2c + 1 the same column one sample on, (x[2e+1], x[2e+2]) (the pair for an odd
  // shift): Dph[((2c + parity) * n1) + row].  The low lg(n2) + 1 bits of the first sample's index i0 are
  // exactly 2c + parity and the rest is the row, so either pair is ONE aligned 8-byte load behind a
  // mask, a shift and a multiply-add -- no second address, no divergence on the parity -- and the 8
  // rows x 4 columns a warp reads (nearly always of one parity) are 4 runs of 64 bytes.
  // (Measured against it: the same two planes in natural order, Dph[parity * M + (i0 >> 1)] -- at 2^24 and
  // +-500 m/s^2, where the shift changes by 3.6 samples from one row of a column to the next and every
  // operand of a transposed column lies in a plane column of its own: 111.1 against 107.9 us / trial, and
  // 100.1 against 92.6 at small accelerations; 45.5 against 45.1 at 2^23.  What costs at large
  // accelerations is the width of the band the trials of a launch read, see ps_search_accel_batch.)
  const float2* D;
  const float2* Dph;
  AccelFacts afs;
  // the kernel's own (grid-constant) parameter copy of afs.v: indexed per work item straight from the
  // parameter bank -- indexing the by-value member would make the compiler keep the whole table in
  // local memory (328-byte stack frame, one local load per work item on the critical path)
  const double* afp = nullptr;
  double size;
  float max_shift;  // host bound on |i*af*(i-N)| over the series, all batch items
  float max_step;   // host bound on |shift(i+1) - shift(i)|
  int n1, lg_n2;
  double af;
  float af32, fsize, thr;
  int n2, c0;
  unsigned col_mask, plane_mask, plane_shift;
  __device__ __forceinline__ void bind(const LoadResampled& param) { afp = param.afs.v; }
  __device__ __forceinline__ void begin(int b, int first, int n) {
    af = afp[b];
    af32 = (float)af;
    fsize = (float)size;
    // float32 evaluation error of the shift: fi and fi - N are exact (even integers up to 2^25), so three
    // roundings (the product, af32 = float(af), the final product) give <= 3 * 2^-24 = 1.79e-7 |shift|
    // (2e-7 used); the second sample of the pair adds at most max_step; the reference's own FP64
    // evaluation of i + i*af*(i-N) is good to (|shift| + i) * 2^-52 < 1e-8 samples, and this float
    // expression to ~3e-8: 1e-5 covers both.  A shift estimate this far from a half-integer fixes the
    // rounding of both samples of a pair; the index maps are pinned bit for bit by
    // test_product_resample_index_map_bit_exact.
    thr = 0.5f - (1e-5f + 2e-7f * max_shift + max_step);
    c0 = first;
    n2 = n;
    col_mask = (1u << lg_n2) - 1u;
    plane_mask = (2u << lg_n2) - 1u;
    plane_shift = (unsigned)lg_n2 + 1u;
  }
  // pair e of the natural series -> its place in the transposed layout (at most 2^24 pairs: 32-bit math)
  __device__ __forceinline__ unsigned pair_index(unsigned e) const { return (e & col_mask) * (unsigned)n1 + (e >> lg_n2); }
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    return at(2u * ((unsigned)idx * (unsigned)n2 + (unsigned)(c0 + f)));
  }
  // samples map(i), map(i + 1) for an even sample index i
  __device__ __forceinline__ cf at(unsigned i) const {
    const float fi = (float)i;                          // exact: i is even and below 2^25
    const float g = (fi * (fi - fsize)) * af32;         // shift estimate
    const float tm = g + 12582912.f;                    // 1.5*2^23: rounds g to an integer (ties to even)
    const float r = tm - 12582912.f;
    if (fabsf(g - r) < thr) {  // samples i0, i0 + 1
      const unsigned i0 = i + (unsigned)(__float_as_int(tm) - 0x4B400000);
      const float2 v = __ldg(Dph + ((i0 & plane_mask) * (unsigned)n1 + (i0 >> plane_shift)));
      return psfft::mk(v.x, v.y);
    }
    // the reference's FP64 expression (kernels.cu:314-317), each sample on its own
    return exact(i);
  }
  __device__ __forceinline__ cf exact(unsigned i) const {
    const float* Df = reinterpret_cast<const float*>(D);
    const unsigned i0 = (unsigned)resample_index_II(af, size, i);
    const unsigned i1 = (unsigned)resample_index_II(af, size, i + 1);
    return psfft::mk(__ldg(&Df[2u * pair_index(i0 >> 1) + (i0 & 1u)]), __ldg(&Df[2u * pair_index(i1 >> 1) + (i1 & 1u)]));
  }

  // The 16 operands of thread (f, u): rows u + a*G of the tile, a < 16 -- operator() with the sample
  // index stepped by a constant instead of rebuilt from (f, idx) per element.
  static constexpr bool kWholeColumn = true;
  template <int G>
  __device__ __forceinline__ void load_column(int f, int u, cf* v) const {
    const unsigned ibase = 2u * ((unsigned)u * (unsigned)n2 + (unsigned)(c0 + f));
    const unsigned di = 2u * (unsigned)G * (unsigned)n2;
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = at(ibase + (unsigned)a * di);
  }
};

// C2R output storer writing the transposed pair layout LoadResampled reads: time pair e = k1 + n1*k2
struct StoreTransposedPairs {
  float2* D;
  int n1, r0;    // packed plan rows / first row of the tile (set by begin)
  int lg_cols;   // log2 of the number of pair columns of the transposed layout
  int col_len;   // pairs per column
  __device__ __forceinline__ void begin(int, int first, int n) {
    r0 = first;
    n1 = n;
  }
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
#pragma unroll
    for (int j = 0; j < R; j++) {
      const unsigned e = (unsigned)(r0 + f) + (unsigned)n1 * (unsigned)(k0 + stride * j);
      D[(size_t)(e & ((1u << lg_cols) - 1u)) * col_len + (e >> lg_cols)] = make_float2(v[j].x, v[j].y);
    }
  }
};

struct LoadU8Padded {  // timeseries.hpp:335-344 + pipeline_multi.cu:160-163 fused
  static constexpr bool kAsync = false;
  __device__ __forceinline__ void prefetch(int, int, cf*) const {}
  const unsigned char* b;
  const float* pad;  // device scalar: mean of the first nsamps samples
  unsigned nsamps;   // valid samples (<= size)
  int n2, c0;
  __device__ __forceinline__ void begin(int, int first, int n) {
    c0 = first;
    n2 = n;
  }
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
  static constexpr bool kAsync = false;
  __device__ __forceinline__ void prefetch(int, int, cf*) const {}
  const float2* X;  // M+1 bins
  unsigned M;
  int n2, c0;
  __device__ __forceinline__ void begin(int, int first, int n) {
    c0 = first;
    n2 = n;
  }
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
  int n2, tw;  // Yt is tiled for pass 2: Yt[k1 / tw][j2][k1 % tw], tw = rows per pass-2 tile
  // W_M^(f u) of this thread (direction applied): depends on the thread only, so a kernel looks it up once
  cf wfu = psfft::mk(1.f, 0.f);
  bool wfu_set = false;
  // outputs k = k0 + stride*j of column j2: multiply by W_M^(j2*k) = base * step^j (two
  // interleaved chains, <= R/2 roundings deep) and store the column run
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
    const unsigned j2 = (unsigned)(c0 + f);
    store_with<R>(f, k0, stride, v, psfft::dirw<INV>(twiddle_M(t, j2 * (unsigned)k0)),
                  psfft::dirw<INV>(twiddle_M(t, j2 * (unsigned)stride)));
  }

  // Phase C hands over the outputs q_m = u + G m, m = 0, 1, ... of thread row u in sequence.  Looking
  // W_M^(j2 q_m) up per call costs a scattered table read each (32 distinct entries per warp: ~30 L1
  // wavefronts); instead W_M^(j2 u) = W_M^(c0 u) W_M^(f u) (8 distinct entries per warp and a
  // thread constant), advanced by W_M^(j2 G) (4 distinct entries) from call to call, and the step
  // W_M^(256 j2) is that factor's square when 256 = 2 G.  One or two extra roundings on the twiddle.
  static constexpr bool kSequenced = true;
  struct Seq {
    cf cur, g, step;
  };
  template <int G>
  __device__ __forceinline__ Seq seq_begin(int f, int u, int stride) const {
    const unsigned j2 = (unsigned)(c0 + f);
    Seq q;
    const cf wf = wfu_set ? wfu : psfft::dirw<INV>(twiddle_M(t, (unsigned)f * (unsigned)u));
    q.cur = psfft::cmul(psfft::dirw<INV>(twiddle_M(t, (unsigned)c0 * (unsigned)u)), wf);
    q.g = psfft::dirw<INV>(twiddle_M(t, j2 * (unsigned)G));
    if (stride == G)
      q.step = q.g;
    else if (stride == 2 * G)
      q.step = psfft::cmul(q.g, q.g);
    else
      q.step = psfft::dirw<INV>(twiddle_M(t, j2 * (unsigned)stride));
    return q;
  }
  template <int R>
  __device__ __forceinline__ void store_seq(int f, int k0, int stride, const cf* v, Seq& q) const {
    store_with<R>(f, k0, stride, v, q.cur, q.step);
    q.cur = psfft::cmul(q.cur, q.g);
  }

  template <int R>
  __device__ __forceinline__ void store_with(int f, int k0, int stride, const cf* v, cf base, cf step1) const {
    const unsigned j2 = (unsigned)(c0 + f);
    cf pw[R > 1 ? R : 2];
    psfft::chain_powers<R>(pw, base, step1);
    // tw (rows per pass-2 tile, a power of two <= 32) divides the stride of every call but the 16-point one:
    // the run position k % tw is then the same for all j and the R stores are one base address plus
    // multiples of a constant -- 32-bit element offsets inside this item's slice (M <= 2^24 elements)
    // instead of a 64-bit index product per store.
    const unsigned twm = ((unsigned)tw - 1u) & 31u;
    float2* out = reinterpret_cast<float2*>(Yt);
    if (((unsigned)stride & twm) == 0u) {
      const unsigned o0 = j2 * (unsigned)tw + ((unsigned)k0 & ~twm) * (unsigned)n2 + ((unsigned)k0 & twm);
      const unsigned step = (unsigned)stride * (unsigned)n2;
#pragma unroll
      for (int j = 0; j < R; j++) {
        const cf o = psfft::cmul(v[j], pw[j]);
        // streaming store: the inter-pass data is read once, much later -- it should not displace the
        // whitened series (re-read by every acceleration trial) from the L2
        __stcs(&out[o0 + step * (unsigned)j], make_float2(o.x, o.y));
      }
      return;
    }
    out += (size_t)j2 * tw;
#pragma unroll
    for (int j = 0; j < R; j++) {
      const cf o = psfft::cmul(v[j], pw[j]);
      const unsigned k = (unsigned)(k0 + stride * j);
      __stcs(&out[(size_t)(k & ~twm) * n2 + (k & twm)], make_float2(o.x, o.y));
    }
  }
};

struct LoadPass2 {
  const cf* Yt;
  int n1, r0;
  int n2, tw;
  // tile r0/tw of Yt[k1 / tw][j2][k1 % tw]: the whole tile is one contiguous run
  __device__ __forceinline__ cf operator()(int f, int idx) const {
    float2 v = reinterpret_cast<const float2*>(Yt)[((size_t)(r0 / tw) * n2 + idx) * tw + f];
    return psfft::mk(v.x, v.y);
  }
};

struct StoreNatural {  // out[k1 + n1*k2]
  float2* out;
  size_t item_stride;  // in float2
  int n1, r0;
  float2* ob;
  __device__ __forceinline__ void begin(int b, int first, int n) {
    ob = out + (size_t)b * item_stride;
    r0 = first;
    n1 = n;
  }
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
    float2* o = ob + (size_t)k0 * n1 + r0 + f;
    const size_t step = (size_t)stride * n1;
#pragma unroll
    for (int j = 0; j < R; j++) o[step * j] = make_float2(v[j].x, v[j].y);
  }
};

// One work item = F transforms of one batch item.  The 16 operands of every thread for the NEXT
// work item are loaded into registers between phase B and phase C of the current one (phase C
// is the light radix-R3 stage), so their latency overlaps phase C and the stores.
template <class L, class = void>
struct loads_whole_column { static constexpr bool value = false; };
template <class L>
struct loads_whole_column<L, decltype((void)L::kWholeColumn)> { static constexpr bool value = true; };

// loaders that index tables of their kernel parameter bind to the grid-constant copy (LoadResampled)
template <class L, class = void>
struct binds_param { static constexpr bool value = false; };
template <class L>
struct binds_param<L, decltype((void)&L::afp)> { static constexpr bool value = true; };
template <class Loader>
__device__ __forceinline__ void bind_loader(Loader& ld, const Loader& param) {
  if constexpr (binds_param<Loader>::value) ld.bind(param);
}

template <class C, class Loader>
__device__ __forceinline__ void load_operands(const Loader& ld, cf* v) {
  const int t = threadIdx.x, f = t % C::F, u = t / C::F;
  if constexpr (loads_whole_column<Loader>::value) {
    ld.template load_column<C::G>(f, u, v);
  } else {
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = ld(f, a * C::G + u);
  }
}

// grid-stride over (batch item, tile); items are independent transforms with their own Yt slice
template <int R3, bool INV, class Loader>
__global__ void __launch_bounds__(512, 1) fft_pass1_kernel(const __grid_constant__ Loader ldp, cf* Yt, FftTables t, int n1, int n2,
                                                           int nbatch, int tw2) {
  using C = psfft::CtaFft<R3, INV>;
  Loader ld = ldp;
  bind_loader(ld, ldp);
  extern __shared__ float smem[];
  const int ntiles = n2 / C::F;
  const int nwork = ntiles * nbatch;
  const size_t M = (size_t)n1 * n2;
  int work = blockIdx.x;
  if (work >= nwork) return;
  cf v[16];
  // work items are ordered tile-major (all batch items of a tile, then the next tile): the CTAs
  // running at any moment read the same few columns of the shared input, whatever the batch size,
  // so the input needs no L2 residency across batch items
  ld.begin(work % nbatch, (work / nbatch) * C::F, n2);
  load_operands<C>(ld, v);
  const cf wA = C::twiddleA(threadIdx.x, t.W4096), wB = R3 > 1 ? C::twiddleB(threadIdx.x, t.W4096) : psfft::mk(1.f, 0.f);
  const cf wfu = psfft::dirw<INV>(twiddle_M(t, (threadIdx.x % C::F) * (threadIdx.x / C::F)));  // StorePass1::wfu
  // (tile, b) of the work item whose operands are in v, stepped by the grid size without a division
  const int dtile = (int)gridDim.x / nbatch, db = (int)gridDim.x - dtile * nbatch;
  int tile = work / nbatch, b = work - tile * nbatch;
  for (; work < nwork; work += gridDim.x) {
    cf* Yb = Yt + (size_t)b * M;
    asm volatile("" : "+l"(Yb));  // one pointer per item: keeps the 64-bit product b * M out of every store's address
    StorePass1<INV> st{Yb, t, n1, tile * C::F, n2, tw2, wfu, true};
    C::phaseA_regs_w(threadIdx.x, smem, wA, v);
    __syncthreads();
    C::phaseB_w(threadIdx.x, smem, wB, st);
    // (issued before the barrier: the loads are in flight while the CTA waits for its slowest warp)
    auto sq = st.template seq_begin<C::G>(threadIdx.x % C::F, threadIdx.x / C::F, 256);
    tile += dtile;
    b += db;
    if (b >= nbatch) {
      b -= nbatch;
      tile++;
    }
    if (work + (int)gridDim.x < nwork) {
      ld.begin(b, tile * C::F, n2);
      load_operands<C>(ld, v);
    }
    __syncthreads();
    C::phaseC_seq(threadIdx.x, smem, st, sq);
  }
}

// ---- TMA bulk copy (cp.async.bulk) + mbarrier helpers: one thread moves a whole contiguous tile
// global -> shared asynchronously; no LSU instructions, no registers, completion on an mbarrier.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Pass 2: a work item's operands are one contiguous 64 KiB tile of the tiled inter-pass buffer, so
// thread 0 fetches the NEXT tile with a single bulk copy into a staging buffer while the current
// tile is transformed (issued once every thread has taken its operands out of the staging buffer).
template <int R3, bool INV, class Storer>
__global__ void __launch_bounds__(512, 1) fft_pass2_kernel(const cf* Yt, Storer st, FftTables t, int n1, int n2, int nbatch, int reverse) {
  using C = psfft::CtaFft<R3, INV>;
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long bar;
  const int ntiles = n1 / C::F;
  const int nwork = ntiles * nbatch;
  const size_t tile_elems = (size_t)n2 * C::F;  // complex words per tile, contiguous in Yt
  constexpr unsigned kTileBytes = sizeof(cf) * 16 * C::THREADS;
  cf* stage = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);
  // reverse = 1 walks the work items backwards: pass 1 wrote the LAST batch item most recently, so
  // starting there finds its inter-pass data still in L2 (the batch does not fit as a whole)
  int slot = blockIdx.x;
  if (slot >= nwork) return;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_proxy_async();
    mbar_expect_tx(&bar, kTileBytes);
    bulk_g2s(stage, Yt + (size_t)(reverse ? nwork - 1 - slot : slot) * tile_elems, kTileBytes, &bar);
  }
  __syncthreads();
  unsigned parity = 0;
  // the thread's two twiddle bases do not depend on the work item (80 registers: room to keep them)
  const cf wA = C::twiddleA(threadIdx.x, t.W4096), wB = R3 > 1 ? C::twiddleB(threadIdx.x, t.W4096) : psfft::mk(1.f, 0.f);
  for (; slot < nwork; slot += gridDim.x) {
    const int work = reverse ? nwork - 1 - slot : slot;
    const int b = work / ntiles, tile = work - b * ntiles;
    st.begin(b, tile * C::F, n1);
    mbar_wait(&bar, parity);
    parity ^= 1u;
    cf v[16];
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = stage[a * C::THREADS + threadIdx.x];
    C::phaseA_regs_w(threadIdx.x, smem, wA, v);
    __syncthreads();  // E1 complete; every thread has taken its operands out of the staging buffer
    const int next_slot = slot + gridDim.x;
    if (threadIdx.x == 0 && next_slot < nwork) {
      const int next = reverse ? nwork - 1 - next_slot : next_slot;
      fence_proxy_async();
      mbar_expect_tx(&bar, kTileBytes);
      bulk_g2s(stage, Yt + (size_t)next * tile_elems, kTileBytes, &bar);
    }
    C::phaseB_w(threadIdx.x, smem, wB, st);
    __syncthreads();
    C::phaseC(threadIdx.x, smem, st);
  }
}

// ------------------------------------------------------------------ host-side launchers
// opt in to > 48 KB dynamic shared memory once per (kernel instantiation, device); `done` must be
// a static of the CALLING template (one per kernel), not of a helper keyed on the pointer type.
// Worker threads of different devices call this concurrently: the flag is atomic, and a thread that
// finds it clear sets the attribute itself (setting it twice is harmless, skipping it is not).
struct SmemOptIn {
  std::atomic<bool> set{false};
};
template <class K>
inline int configure_smem(K kernel, size_t bytes, int device, SmemOptIn* done) {
  SmemOptIn& d = done[device & 63];
  if (!d.set.load(std::memory_order_acquire)) {
    PS_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes),
            "cudaFuncSetAttribute");
    d.set.store(true, std::memory_order_release);
  }
  return PS_OK;
}

inline int fft_grid(int ntiles) { return ntiles < 148 ? ntiles : 148; }

template <int R3, bool INV, class Loader>
inline int launch_pass1_r(const ps_fft_plan_t* p, const Loader& ld, cudaStream_t st, int nbatch, int batch0) {
  using C = psfft::CtaFft<R3, INV>;
  auto k = fft_pass1_kernel<R3, INV, Loader>;
  static SmemOptIn done[64];
  const size_t smem = C::SMEM_BYTES;
  int rc = configure_smem(k, smem, p->device, done);
  if (rc) return rc;
  k<<<fft_grid(p->n2 / C::F * nbatch), C::THREADS, smem, st>>>(ld, p->Yt + (size_t)batch0 * p->M, p->tables(), p->n1, p->n2, nbatch, 32 / p->r2);
  PS_LAUNCHED("fft_pass1_kernel");
  return PS_OK;
}

template <int R3, bool INV, class Storer>
inline int launch_pass2_r(const ps_fft_plan_t* p, const Storer& stor, cudaStream_t st, int nbatch, int reverse) {
  using C = psfft::CtaFft<R3, INV>;
  auto k = fft_pass2_kernel<R3, INV, Storer>;
  static SmemOptIn done[64];
  const size_t smem = C::SMEM_BYTES + sizeof(cf) * 16 * C::THREADS;  // exchange buffers + TMA staging tile
  int rc = configure_smem(k, smem, p->device, done);
  if (rc) return rc;
  k<<<fft_grid(p->n1 / C::F * nbatch), C::THREADS, smem, st>>>(p->Yt, stor, p->tables(), p->n1, p->n2, nbatch, reverse);
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
// nbatch transforms into the inter-pass slices batch0 .. batch0 + nbatch - 1
inline int launch_pass1(const ps_fft_plan_t* p, const Loader& ld, cudaStream_t st, int nbatch = 1, int batch0 = 0) {
#define PS_CALL1(R) launch_pass1_r<R, INV, Loader>(p, ld, st, nbatch, batch0)
  PS_DISPATCH_R(p->r1, PS_CALL1)
#undef PS_CALL1
}

template <bool INV, class Storer>
inline int launch_pass2(const ps_fft_plan_t* p, const Storer& stor, cudaStream_t st, int nbatch = 1, int reverse = 0) {
#define PS_CALL2(R) launch_pass2_r<R, INV, Storer>(p, stor, st, nbatch, reverse)
  PS_DISPATCH_R(p->r2, PS_CALL2)
#undef PS_CALL2
}

// Z (M packed-complex bins in p->Z) -> X[0..M] (fft.cu)
int launch_r2c_post(const ps_fft_plan_t* p, const cf* Z, float2* d_out, cudaStream_t st);

}  // namespace ps
