// pass2x.cuh -- pass 2 of the search transform with the real-FFT split, interbinning and
// normalisation as its epilogue (SpectrumFormer::form_interpolated, kernels.cu:231-252, and
// stats::normalise, kernels.cu:469-494, on top of the row transforms): the packed spectrum Z never
// leaves the SM, so the chain has one kernel and one 8N-byte DRAM round trip less.
// Included inside namespace ps by search.cu (after split/interbin helpers are defined).
//
// Packed transform Z[k], k = k1 + n1*k2: pass 2 produces rows k1 (all k2).  The split needs Z[k] and
// Z[M-k]: row k1 pairs with row n1-k1 (columns reversed), and interbinning needs X[k-1] = row k1-1.
// A CTA therefore WALKS over steps t = ta .. tb-1 of one transform; step t transforms two natural
// 4-row tiles of the inter-pass buffer,
//     lower  L(t) = rows 4t .. 4t+3            upper  U(t) = rows n1-4t-4 .. n1-4t-1,
// pairs rows (4t+f, n1-4t-f), f = 1..3, inside the step, row 4t with row n1-4t (the lowest row of
// U(t-1), carried as Z), and carries three rows to the next step through a per-CTA scratch in
// global memory (it stays in the L2): C0 = Z[n1-4t-4] (becomes X[n1-4(t+1)] there), C1 = X[4t+3]
// (left neighbour of row 4t+4), C2 = X[n1-4t-3] (its own left neighbour X[n1-4t-4] only exists one
// step later, so that row is emitted then).  Step t writes the spectrum rows 4t..4t+3 and
// n1-4t-2 .. n1-4t+1: 16-byte and 8-byte runs that complete their 32-byte sectors within two or
// three consecutive steps of the same CTA, i.e. while the line is still in the L2.
// Row 0 pairs with itself (k2 <-> n2-k2; k = 0 also yields the Nyquist bin M), row n1/2 with itself
// (k2 <-> n2-1-k2); the last step of a transform emits those two extra rows.  A walk that does not
// start at t = 0 first transforms the two tiles of step ta-1 to obtain its three carries (one extra
// step per walk; walks are ~45 steps long).
//
// Shared memory: E1 | E2 (the exchanges of fft_core.cuh) | B.  U(t) is transformed first, its rows go
// to B; L(t) second, its rows go to E1 (free after phase B).  Both are [4][2048] complex words with
// the column index XOR-ed by 4*row, which keeps the phase-C stores (4 rows x 4 columns per
// half-warp) and the row-wise passes conflict-free.
#pragma once

struct P2xArgs {
  const cf* Yt;        // nbatch x M, tiled [k1/4][j2][k1%4]
  float* P;            // nbatch x pstride
  size_t pstride;
  cf* scratch;         // gridDim.x x 3 x n2 complex: the carries C0, C1, C2 of every CTA
  const float* stats;  // device: mean, rms, std of the DM trial's whitened spectrum
  float fsize;         // transform length N (normalisation with mean*N, std*N, pipeline_multi.cu:224)
  const cf* TC;        // W_N^(4096 q)
  const cf* TD;        // W_N^q, q < 4096
  int n1, nbatch;
};

constexpr int kP2xN2 = 2048;  // row length the kernel is written for (R3 = 8, F = 4)

__device__ __forceinline__ cf& p2x_at(cf* rows, int row, int col) { return rows[row * kP2xN2 + (col ^ (row << 2))]; }

// phase-C storer: output k2 = k0 + stride*j of FFT f -> rows[f][k2]
struct P2xRowStore {
  cf* rows;
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
#pragma unroll
    for (int j = 0; j < R; j++) p2x_at(rows, f, k0 + stride * j) = v[j];
  }
};

// X[k] and X[M-k] from the packed pair (zk = Z[k], zm = Z[M-k]) and w = W_N^k: X_k = E + T,
// X_{M-k} = conj(E - T), T = W_N^k O (same arithmetic as spectrum_pairs)
__device__ __forceinline__ void p2x_split(cf zk, cf zm, float2 w, cf* xa, cf* xb) {
  const float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);
  const float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x);
  const float tr = w.x * orr - w.y * oi, ti = w.x * oi + w.y * orr;
  *xa = psfft::mk(er + tr, ei + ti);
  *xb = psfft::mk(er - tr, ti - ei);
}

__device__ __forceinline__ float p2x_bin(cf x, cf left, float mean, float inv_sigma) {
  return interbin_norm(make_float2(x.x, x.y), make_float2(left.x, left.y), mean, inv_sigma);
}

#ifdef PS_PROF
// per-warp cycle counters of the phases (profiling builds only: make EXTRA=-DPS_PROF; tools/prof_phases.py)
__device__ unsigned g_prof_buf[148 * 16 * 16];
#define PS_TICK(k)                                   \
  {                                                  \
    const unsigned now_ = (unsigned)clock();         \
    pacc[k] += now_ - plast;                         \
    plast = now_;                                    \
  }
#define PS_PROF_BEGIN unsigned pacc[16] = {0}, plast = (unsigned)clock();
#define PS_PROF_END                                                                                      \
  if ((threadIdx.x & 31) == 0 && blockIdx.x < 148)                                                       \
    for (int k_ = 0; k_ < 16; k_++) g_prof_buf[(blockIdx.x * 16 + threadIdx.x / 32) * 16 + k_] = pacc[k_];
#else
#define PS_TICK(k)
#define PS_PROF_BEGIN
#define PS_PROF_END
#endif

template <int R3>
__global__ void __launch_bounds__(512, 1) fft_pass2x_kernel(P2xArgs a, FftTables t) {
  using C = psfft::CtaFft<R3, false>;
  static_assert(C::N == kP2xN2 && C::F == 4, "written for 2048-point rows, four per tile");
  extern __shared__ __align__(128) float smem[];
  __shared__ cf x_nyquist;
  __shared__ __align__(8) unsigned long long bar[2];  // [0]: U tile landed in E2, [1]: L tile landed in B
  cf* E1 = reinterpret_cast<cf*>(smem);                         // L rows (Z, then X)
  cf* E2s = reinterpret_cast<cf*>(smem) + C::E1_WORDS;          // exchange 2; between uses: staging of the next U tile
  cf* B = reinterpret_cast<cf*>(smem + C::SMEM_WORDS);          // staging of the L tile, then U rows (Z, then X)
  constexpr int n2 = kP2xN2;
  constexpr unsigned kTileBytes = sizeof(cf) * 4 * kP2xN2;
  const int n1 = a.n1, T = n1 / 8, tid = threadIdx.x;
  const unsigned M = (unsigned)n1 * (unsigned)n2;
  cf* C0 = a.scratch + (size_t)blockIdx.x * 3 * n2;
  cf* C1 = C0 + n2;
  cf* C2 = C1 + n2;
  const float mean = __fmul_rn(a.stats[0], a.fsize);
  const float inv_sigma = __fdiv_rn(1.0f, __fmul_rn(a.stats[2], a.fsize));
  // this CTA's share of the nbatch*T steps is one contiguous run.  Only its first step can lie inside a
  // transform (ta > 0): then the pair of tiles of step ta-1 is transformed first, for the carries.
  const long total = (long)a.nbatch * T;
  const long g0 = total * blockIdx.x / gridDim.x, g_end = total * (blockIdx.x + 1) / gridDim.x;
  if (g0 >= g_end) return;
  const int has_prime = (g0 % T) != 0;
  const int npairs = (int)(g_end - g0) + has_prime;
  // pair n -> (transform b, step st); the priming pair is step ta-1 of the first transform
  auto pair_of = [&](int n, int* b, int* st) {
    const long g = g0 + n - has_prime;
    *b = (int)(g / T);
    *st = (int)(g - (long)*b * T);
  };
  auto tile_ptr = [&](int b, int tile) { return a.Yt + (size_t)b * M + (size_t)tile * (kP2xN2 * 4); };
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_proxy_async();
    int b, st;
    pair_of(0, &b, &st);
    mbar_expect_tx(&bar[0], kTileBytes);
    bulk_g2s(E2s, tile_ptr(b, n1 / 4 - 1 - st), kTileBytes, &bar[0]);
  }
  __syncthreads();
  unsigned parU = 0, parL = 0;
  PS_PROF_BEGIN
  for (int n = 0; n < npairs; n++) {
    int b, st;
    pair_of(n, &b, &st);
    const bool prime = has_prime && n == 0;
    float* __restrict__ P = a.P + (size_t)b * a.pstride;
    // ---- U tile: operands from the staging in E2; the L tile starts flying into B
    if (tid == 0) {
      fence_proxy_async();
      mbar_expect_tx(&bar[1], kTileBytes);
      bulk_g2s(B, tile_ptr(b, st), kTileBytes, &bar[1]);
    }
    cf v[16];
    PS_TICK(0)
    mbar_wait(&bar[0], parU);
    parU ^= 1u;
    PS_TICK(1)
#pragma unroll
    for (int q = 0; q < 16; q++) v[q] = E2s[q * C::THREADS + tid];
    C::phaseA_regs(tid, smem, t.W4096, v);
    PS_TICK(2)
    __syncthreads();  // E1 complete; every thread has taken its operands out of E2
    {
      P2xRowStore rs{B};
      C::phaseB(tid, smem, t.W4096, rs);
      PS_TICK(3)
      mbar_wait(&bar[1], parL);
      parL ^= 1u;
      PS_TICK(4)
#pragma unroll
      for (int q = 0; q < 16; q++) v[q] = B[q * C::THREADS + tid];   // the L tile's operands, before B is overwritten
      __syncthreads();
      C::phaseC(tid, smem, rs);     // U rows -> B
    }
    PS_TICK(5)
    // ---- L tile
    C::phaseA_regs(tid, smem, t.W4096, v);
    PS_TICK(6)
    __syncthreads();
    {
      P2xRowStore rs{E1};
      C::phaseB(tid, smem, t.W4096, rs);
      PS_TICK(7)
      __syncthreads();
      C::phaseC(tid, smem, rs);     // L rows -> E1
    }
    PS_TICK(8)
    __syncthreads();
    if (tid == 0 && n + 1 < npairs) {  // E2 is idle until the next pair: its U tile lands there meanwhile
      int nb_, ns_;
      pair_of(n + 1, &nb_, &ns_);
      fence_proxy_async();
      mbar_expect_tx(&bar[0], kTileBytes);
      bulk_g2s(E2s, tile_ptr(nb_, n1 / 4 - 1 - ns_), kTileBytes, &bar[0]);
    }
    if (prime) {
      // carries of step st = ta-1: split only the pair of rows (4ta-1, n1-4ta+1)
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int k2 = tid + 512 * i;
        const unsigned k = (unsigned)(4 * st + 3) + (unsigned)n1 * (unsigned)k2;
        cf xa, xb;
        p2x_split(p2x_at(E1, 3, k2), p2x_at(B, 1, n2 - 1 - k2), twiddle_N(a.TC, a.TD, k), &xa, &xb);
        C1[k2] = xa;
        C2[n2 - 1 - k2] = xb;
        C0[k2] = p2x_at(B, 0, k2);
      }
      __syncthreads();
      continue;
    }
    const bool more = n + 1 < npairs && st + 1 < T;  // the next step belongs to the same transform: hand over the carries
    // ---- split: Z -> X in place
#pragma unroll 4
    for (int i = 0; i < 16; i++) {
      const int p = tid + 512 * i, f = p >> 11, k2 = p & (n2 - 1);
      const unsigned k = (unsigned)(4 * st + f) + (unsigned)n1 * (unsigned)k2;
      if (f >= 1) {
        cf xa, xb;
        p2x_split(p2x_at(E1, f, k2), p2x_at(B, 4 - f, n2 - 1 - k2), twiddle_N(a.TC, a.TD, k), &xa, &xb);
        p2x_at(E1, f, k2) = xa;
        p2x_at(B, 4 - f, n2 - 1 - k2) = xb;
      } else if (st >= 1) {  // row 4st with the carried row n1-4st
        cf xa, xb;
        p2x_split(p2x_at(E1, 0, k2), C0[n2 - 1 - k2], twiddle_N(a.TC, a.TD, k), &xa, &xb);
        p2x_at(E1, 0, k2) = xa;
        C0[n2 - 1 - k2] = xb;
      } else if (k2 <= n2 / 2) {  // row 0 pairs with itself: k2 <-> n2-k2; k2 = 0 also gives bin M
        const int kp = (n2 - k2) & (n2 - 1);
        cf xa, xb;
        p2x_split(p2x_at(E1, 0, k2), p2x_at(E1, 0, kp), twiddle_N(a.TC, a.TD, k), &xa, &xb);
        p2x_at(E1, 0, k2) = xa;
        if (k2 == 0)
          x_nyquist = xb;
        else if (kp != k2)
          p2x_at(E1, 0, kp) = xb;
      }
    }
    if (st == T - 1) {  // row n1/2 (slot 0 of U) pairs with itself: k2 <-> n2-1-k2
#pragma unroll
      for (int i = 0; i < 2; i++) {
        const int k2 = tid + 512 * i;
        const unsigned k = (unsigned)(n1 / 2) + (unsigned)n1 * (unsigned)k2;
        cf xa, xb;
        p2x_split(p2x_at(B, 0, k2), p2x_at(B, 0, n2 - 1 - k2), twiddle_N(a.TC, a.TD, k), &xa, &xb);
        p2x_at(B, 0, k2) = xa;
        p2x_at(B, 0, n2 - 1 - k2) = xb;
      }
    }
    PS_TICK(9)
    __syncthreads();
    // ---- interbin + normalise, spectrum rows out, carries for the next step
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int k2 = tid + 512 * i;
      const cf x0 = p2x_at(E1, 0, k2), x1 = p2x_at(E1, 1, k2), x2 = p2x_at(E1, 2, k2), x3 = p2x_at(E1, 3, k2);
      const cf b0 = p2x_at(B, 0, k2), b1 = p2x_at(B, 1, k2), b2 = p2x_at(B, 2, k2), b3 = p2x_at(B, 3, k2);
      cf left;
      if (st >= 1)
        left = C1[k2];                                                   // X[4st-1]
      else
        left = k2 >= 1 ? p2x_at(B, 3, k2 - 1) : psfft::mk(0.f, 0.f);    // bin n1*k2 - 1 = row n1-1 of column k2-1
      float* lo = P + (size_t)(4 * st) + (size_t)n1 * k2;
      *reinterpret_cast<float4*>(lo) = make_float4(p2x_bin(x0, left, mean, inv_sigma), p2x_bin(x1, x0, mean, inv_sigma),
                                                   p2x_bin(x2, x1, mean, inv_sigma), p2x_bin(x3, x2, mean, inv_sigma));
      float* up = P + (size_t)(n1 - 4 * st - 2) + (size_t)n1 * k2;      // rows n1-4st-2 .. n1-4st+1
      *reinterpret_cast<float2*>(up) = make_float2(p2x_bin(b2, b1, mean, inv_sigma), p2x_bin(b3, b2, mean, inv_sigma));
      if (st >= 1) {
        const cf c0 = C0[k2], c2 = C2[k2];                               // X[n1-4st], X[n1-4st+1]
        *reinterpret_cast<float2*>(up + 2) = make_float2(p2x_bin(c0, b3, mean, inv_sigma), p2x_bin(c2, c0, mean, inv_sigma));
      }
      if (st == T - 1) {  // rows n1/2 and n1/2+1 have both neighbours now
        float* mid = P + (size_t)(n1 / 2) + (size_t)n1 * k2;
        *reinterpret_cast<float2*>(mid) = make_float2(p2x_bin(b0, x3, mean, inv_sigma), p2x_bin(b1, b0, mean, inv_sigma));
      }
      if (more) {
        C0[k2] = b0;   // Z[n1-4st-4]
        C1[k2] = x3;   // X[4st+3]
        C2[k2] = b1;   // X[n1-4st-3]
      }
    }
    if (st == 0 && tid == 0) P[M] = p2x_bin(x_nyquist, p2x_at(B, 3, n2 - 1), mean, inv_sigma);  // left neighbour: bin M-1
    PS_TICK(10)
    __syncthreads();
  }
  PS_PROF_END
}

template <int R3>
inline int launch_pass2x_r(const ps_fft_plan_t* p, const P2xArgs& a, int grid, cudaStream_t st) {
  using C = psfft::CtaFft<R3, false>;
  auto k = fft_pass2x_kernel<R3>;
  static SmemOptIn done[64];
  const size_t smem = C::SMEM_BYTES + sizeof(cf) * 4 * kP2xN2;
  int rc = configure_smem(k, smem, p->device, done);
  if (rc) return rc;
  k<<<grid, C::THREADS, smem, st>>>(a, p->tables());
  PS_LAUNCHED("fft_pass2x_kernel");
  return PS_OK;
}

// sizes the fused pass 2 takes: 2048-point rows in four-row tiles (2^22 < N <= 2^24 samples)
inline bool pass2x_ok(const ps_fft_plan_t* p) { return p->n2 == kP2xN2 && p->r2 == 8 && p->n1 >= 16; }
constexpr int kP2xGrid = 148;
