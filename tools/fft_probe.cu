// fft_probe.cu -- ablation timing of the pass-2 FFT kernel structure (not part of the product).
// Variants: full | noload (operands synthesised) | nostore (results discarded unless NaN) |
// noA / noB / noC (one phase skipped) to see what the tile time is made of.
#include <cstdio>
#include <vector>
#include "../peasoup_b200/csrc/fft.cuh"
using namespace ps;
using psfft::cf;

namespace ps { void set_error(const std::string&) {} std::atomic<unsigned long long> g_launches{0}; }

struct NullStore {
  float2* out; int n1, r0;
  __device__ __forceinline__ void begin(int, int first, int n) { r0 = first; n1 = n; }
  template <int R>
  __device__ __forceinline__ void store(int f, int k0, int stride, const cf* v) const {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < R; j++) acc += v[j].x * v[j].y;
    if (acc == 123.456f) out[k0] = make_float2(acc, acc);  // never true: keeps the math alive
  }
};

template <int MODE, class Storer>
__global__ void __launch_bounds__(512, 1) probe_kernel(const cf* Yt, Storer st, FftTables t, int n1, int n2, int nbatch) {
  using C = psfft::CtaFft<8, false>;
  extern __shared__ float smem[];
  const int ntiles = n1 / C::F, nwork = ntiles * nbatch;
  const size_t M = (size_t)n1 * n2;
  int work = blockIdx.x;
  if (work >= nwork) return;
  cf v[16];
  auto load = [&](int w) {
    if (MODE == 1) {
#pragma unroll
      for (int a = 0; a < 16; a++) v[a] = psfft::mk((float)(threadIdx.x + a + w), (float)(a * 3 - w));
    } else {
      LoadPass2 ld{Yt + (size_t)(w / ntiles) * M, n1, (w % ntiles) * C::F, n2, C::F};
      load_operands<C>(ld, v);
    }
  };
  load(work);
  for (; work < nwork; work += gridDim.x) {
    const int b = work / ntiles, tile = work - b * ntiles;
    st.begin(b, tile * C::F, n1);
    if (MODE != 3) C::phaseA_regs(threadIdx.x, smem, t.W4096, v);
    __syncthreads();
    if (MODE != 4) C::phaseB(threadIdx.x, smem, t.W4096, st);
    __syncthreads();
    const int next = work + gridDim.x;
    if (next < nwork) load(next);
    if (MODE != 5) C::phaseC(threadIdx.x, smem, st);
  }
}

template <int MODE, class Storer>
float run(const char* name, const cf* Yt, Storer st, FftTables t, int n1, int n2, int nbatch) {
  using C = psfft::CtaFft<8, false>;
  auto k = probe_kernel<MODE, Storer>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; i++) k<<<148, 512, C::SMEM_BYTES>>>(Yt, st, t, n1, n2, nbatch);
  cudaEventRecord(e0);
  const int reps = 20;
  for (int i = 0; i < reps; i++) k<<<148, 512, C::SMEM_BYTES>>>(Yt, st, t, n1, n2, nbatch);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("%-10s %7.1f us per launch (%d trials)  err=%s\n", name, 1e3f * ms / reps, nbatch, cudaGetErrorString(cudaGetLastError()));
  return ms;
}

template <int MODE, class Storer>
__global__ void __launch_bounds__(512, 1) probe1_kernel(LoadResampled ld, Storer st0, cf* Yt, FftTables t, int n1, int n2, int nbatch, int tw2) {
  ld.bind(ld);
  using C = psfft::CtaFft<8, false>;
  extern __shared__ float smem[];
  const int ntiles = n2 / C::F, nwork = ntiles * nbatch;
  const size_t M = (size_t)n1 * n2;
  int work = blockIdx.x;
  if (work >= nwork) return;
  cf v[16];
  auto load = [&](int w) {
    if (MODE == 1) {
#pragma unroll
      for (int a = 0; a < 16; a++) v[a] = psfft::mk((float)(threadIdx.x + a + w), (float)(a * 3 - w));
    } else {
      ld.begin(w / ntiles, (w % ntiles) * C::F, n2);
      load_operands<C>(ld, v);
    }
  };
  load(work);
  for (; work < nwork; work += gridDim.x) {
    const int b = work / ntiles, tile = work - b * ntiles;
    StorePass1<false> st{Yt + (size_t)b * M, t, n1, tile * C::F, n2, tw2};
    Storer sx = st0;
    sx.begin(b, tile * C::F, n1);
    C::phaseA_regs(threadIdx.x, smem, t.W4096, v);
    __syncthreads();
    if (MODE == 2) C::phaseB(threadIdx.x, smem, t.W4096, sx); else C::phaseB(threadIdx.x, smem, t.W4096, st);
    __syncthreads();
    const int next = work + gridDim.x;
    if (next < nwork) load(next);
    if (MODE == 2) C::phaseC(threadIdx.x, smem, sx); else C::phaseC(threadIdx.x, smem, st);
  }
}

template <int MODE>
void run1(const char* name, LoadResampled ld, cf* Yt, FftTables t, int n1, int n2, int nbatch) {
  using C = psfft::CtaFft<8, false>;
  auto k = probe1_kernel<MODE, NullStore>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES);
  NullStore ns{reinterpret_cast<float2*>(Yt), 0, 0};
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; i++) k<<<148, 512, C::SMEM_BYTES>>>(ld, ns, Yt, t, n1, n2, nbatch, 4);
  cudaEventRecord(e0);
  const int reps = 20;
  for (int i = 0; i < reps; i++) k<<<148, 512, C::SMEM_BYTES>>>(ld, ns, Yt, t, n1, n2, nbatch, 4);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("pass1 %-10s %7.1f us per launch (%d trials)  err=%s\n", name, 1e3f * ms / reps, nbatch, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int n1 = 2048, n2 = 2048, nbatch = 8;
  const size_t M = (size_t)n1 * n2;
  cf *Yt, *Z, *W, *TA, *TB;
  cudaMalloc(&Yt, sizeof(cf) * M * nbatch); cudaMalloc(&Z, sizeof(cf) * M * nbatch);
  cudaMemset(Yt, 0, sizeof(cf) * M * nbatch);
  std::vector<cf> w(4096);
  for (int q = 0; q < 4096; q++) { w[q].x = (float)cos(-2 * M_PI * q / 4096); w[q].y = (float)sin(-2 * M_PI * q / 4096); }
  cudaMalloc(&W, sizeof(cf) * 4096); cudaMalloc(&TA, sizeof(cf) * 4096); cudaMalloc(&TB, sizeof(cf) * 4096);
  cudaMemcpy(W, w.data(), sizeof(cf) * 4096, cudaMemcpyHostToDevice);
  cudaMemcpy(TA, w.data(), sizeof(cf) * 4096, cudaMemcpyHostToDevice);
  cudaMemcpy(TB, w.data(), sizeof(cf) * 4096, cudaMemcpyHostToDevice);
  FftTables t{W, TA, TB};
  StoreNatural sn{reinterpret_cast<float2*>(Z), M, 0, 0, nullptr};
  NullStore ns{reinterpret_cast<float2*>(Z), 0, 0};
  run<0>("full", Yt, sn, t, n1, n2, nbatch);
  run<1>("noload", Yt, sn, t, n1, n2, nbatch);
  run<0>("nostore", Yt, ns, t, n1, n2, nbatch);
  run<1>("compute", Yt, ns, t, n1, n2, nbatch);
  run<3>("noA", Yt, sn, t, n1, n2, nbatch);
  run<4>("noB", Yt, sn, t, n1, n2, nbatch);
  run<5>("noC", Yt, sn, t, n1, n2, nbatch);
  {
    float* D; cudaMalloc(&D, sizeof(cf) * M); cudaMemset(D, 0, sizeof(cf) * M);
    float2* Dph; cudaMalloc(&Dph, sizeof(float2) * 2 * M); cudaMemset(Dph, 0, sizeof(float2) * 2 * M);  // two-phase form the loader reads
    LoadResampled ld;
    ld.D = reinterpret_cast<const float2*>(D);
    ld.Dph = Dph;
    double af = ((30.0f * 64e-6f) / (2 * 299792458.0));
    for (int b = 0; b < 8; b++) ld.afs.v[b] = af * (b + 1) / 2;
    ld.size = 2.0 * M; ld.n1 = n1; ld.lg_n2 = 11;
    ld.max_shift = (float)(af * 4 * 0.25 * ld.size * ld.size); ld.max_step = (float)(af * 4 * ld.size);
    run1<0>("full", ld, Yt, t, n1, n2, nbatch);
    run1<1>("noload", ld, Yt, t, n1, n2, nbatch);
    run1<2>("nostore", ld, Yt, t, n1, n2, nbatch);
    LoadResampled hot = ld;
    hot.n1 = 0;  // every pair address collapses to its row index: the loads stay in a 16 KB region (index math kept, memory cheap)
    run1<0>("hotload", hot, Yt, t, n1, n2, nbatch);
    run1<2>("hot+nostore", hot, Yt, t, n1, n2, nbatch);
    LoadResampled zero = ld;
    for (int b = 0; b < 8; b++) zero.afs.v[b] = 0.0;  // no shift: every pair is one aligned 8-byte load
    zero.max_shift = 0.f; zero.max_step = 0.f;
    run1<0>("acc0", zero, Yt, t, n1, n2, nbatch);
  }
  return 0;
}
