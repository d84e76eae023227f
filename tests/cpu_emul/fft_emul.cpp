// fft_emul.cpp -- host execution of the CTA FFT phases in peasoup_b200/csrc/fft_core.cuh.
//
// fft_core.cuh compiles as plain C++; this file drives its phaseA/phaseB/phaseC with a loop
// over the 512 "threads" of a CTA (a barrier between phases becomes the end of a loop) and
// restates the pass-1/pass-2 loaders and storers of fft.cuh on the host, so the digit maps,
// shared-memory layouts, twiddle tables and the four-step algebra are checked against
// numpy.fft in the CPU test-suite (tests/test_fft_emul.py) before any GPU time is spent.
#include <cmath>
#include <cstring>
#include <vector>
#include "../../peasoup_b200/csrc/fft_core.cuh"

using psfft::cf;

namespace {

std::vector<cf> roots(unsigned count, unsigned step, unsigned period) {
  std::vector<cf> v(count);
  for (unsigned q = 0; q < count; q++) {
    double ang = -2.0 * M_PI * (double)(((unsigned long long)q * step) % period) / (double)period;
    v[q].x = (float)cos(ang);
    v[q].y = (float)sin(ang);
  }
  return v;
}

struct Tables {
  std::vector<cf> W4096, TA, TB;
  unsigned M;
  explicit Tables(unsigned M_) : M(M_), W4096(roots(4096, 1, 4096)), TA(roots(M_ / 4096 ? M_ / 4096 : 1, 4096, M_)), TB(roots(4096, 1, M_)) {}
  cf twiddle(unsigned p) const { return psfft::cmul(TA[p >> 12], TB[p & 4095u]); }
};

struct LoadCols {
  const cf* in;
  int n2, c0;
  cf operator()(int f, int idx) const { return in[(size_t)idx * n2 + c0 + f]; }
};
template <bool INV>
struct StoreP1 {
  cf* Yt;
  const Tables* t;
  int n1, c0;
  int n2, tw;
  // same recurrence as StorePass1 in fft.cuh: W^(j2*(k0+stride*j)) = base * step^j
  template <int R>
  void store(int f, int k0, int stride, const cf* v) const {
    unsigned j2 = (unsigned)(c0 + f);
    cf base = psfft::dirw<INV>(t->twiddle(j2 * (unsigned)k0));
    cf pw[R > 1 ? R : 2];
    psfft::chain_powers<R>(pw, base, psfft::dirw<INV>(t->twiddle((j2 * (unsigned)stride) % t->M)));
    for (int j = 0; j < R; j++) {
      int k = k0 + stride * j;
      Yt[((size_t)(k / tw) * n2 + j2) * tw + k % tw] = psfft::cmul(v[j], pw[j]);
    }
  }
};
struct LoadP2 {
  const cf* Yt;
  int n1, r0;
  int n2, tw;
  cf operator()(int f, int idx) const { return Yt[((size_t)(r0 / tw) * n2 + idx) * tw + f]; }
};
struct StoreNat {
  cf* out;
  int n1, r0;
  template <int R>
  void store(int f, int k0, int stride, const cf* v) const {
    for (int j = 0; j < R; j++) out[(size_t)(k0 + stride * j) * n1 + r0 + f] = v[j];
  }
};

template <int R3, bool INV, class L, class S>
void run_tile(std::vector<float>& smem, const Tables& t, const L& ld, const S& st) {
  using C = psfft::CtaFft<R3, INV>;
  smem.assign(C::SMEM_WORDS, 0.f);
  for (int th = 0; th < C::THREADS; th++) C::phaseA(th, smem.data(), t.W4096.data(), ld);
  for (int th = 0; th < C::THREADS; th++) C::phaseB(th, smem.data(), t.W4096.data(), st);
  for (int th = 0; th < C::THREADS; th++) C::phaseC(th, smem.data(), st);
}

template <int R3, bool INV>
void pass1(const cf* in, cf* Yt, const Tables& t, int n1, int n2, int tw2) {
  using C = psfft::CtaFft<R3, INV>;
  std::vector<float> smem;
  for (int tile = 0; tile < n2 / C::F; tile++) {
    LoadCols ld{in, n2, tile * C::F};
    StoreP1<INV> st{Yt, &t, n1, tile * C::F, n2, tw2};
    run_tile<R3, INV>(smem, t, ld, st);
  }
}
template <int R3, bool INV>
void pass2(const cf* Yt, cf* out, const Tables& t, int n1, int n2) {
  using C = psfft::CtaFft<R3, INV>;
  std::vector<float> smem;
  for (int tile = 0; tile < n1 / C::F; tile++) {
    LoadP2 ld{Yt, n1, tile * C::F, n2, C::F};
    StoreNat st{out, n1, tile * C::F};
    run_tile<R3, INV>(smem, t, ld, st);
  }
}

#define DISPATCH(r, FN, ...)                 \
  switch (r) {                               \
    case 1: FN<1, INV>(__VA_ARGS__); break;  \
    case 2: FN<2, INV>(__VA_ARGS__); break;  \
    case 4: FN<4, INV>(__VA_ARGS__); break;  \
    case 8: FN<8, INV>(__VA_ARGS__); break;  \
    case 16: FN<16, INV>(__VA_ARGS__); break; \
    default: return -1;                      \
  }

template <bool INV>
int cfft(int n1, int n2, const cf* in, cf* out) {
  unsigned M = (unsigned)n1 * n2;
  Tables t(M);
  std::vector<cf> Yt(M);
  DISPATCH(n1 / 256, pass1, in, Yt.data(), t, n1, n2, 32 / (n2 / 256))
  DISPATCH(n2 / 256, pass2, Yt.data(), out, t, n1, n2)
  return 0;
}

int split(unsigned M, int* n1, int* n2) {
  int lg = 0;
  while ((1u << lg) < M) lg++;
  if ((1u << lg) != M) return -1;
  int l1 = lg / 2, l2 = lg - l1;
  if (l1 < 8 || l2 > 12) return -1;
  *n1 = 1 << l1;
  *n2 = 1 << l2;
  return 0;
}

}  // namespace

extern "C" {

int emul_cfft(int n1, int n2, int inverse, const cf* in, cf* out) {
  return inverse ? cfft<true>(n1, n2, in, out) : cfft<false>(n1, n2, in, out);
}

// same arithmetic as r2c_post_kernel (fft.cu); sinpi/cospi of the exact ratio k/M
int emul_r2c(unsigned n, const float* x, cf* X) {
  unsigned M = n / 2;
  int n1, n2;
  if (split(M, &n1, &n2)) return -1;
  std::vector<cf> Z(M);
  if (cfft<false>(n1, n2, reinterpret_cast<const cf*>(x), Z.data())) return -1;
  for (unsigned k = 0; k <= M; k++) {
    cf zk = Z[k == M ? 0 : k], zm = Z[(k == 0 || k == M) ? 0 : M - k];
    double a = M_PI * (double)k / (double)M;
    float wr = (float)cos(a), wi = (float)-sin(a);
    float er = 0.5f * (zk.x + zm.x), ei = 0.5f * (zk.y - zm.y);
    float orr = 0.5f * (zk.y + zm.y), oi = -0.5f * (zk.x - zm.x);
    X[k] = psfft::mk(er + (wr * orr - wi * oi), ei + (wr * oi + wi * orr));
  }
  return 0;
}

// same arithmetic as LoadC2RPre (fft.cuh)
int emul_c2r(unsigned n, const cf* X, float* x) {
  unsigned M = n / 2;
  int n1, n2;
  if (split(M, &n1, &n2)) return -1;
  std::vector<cf> Zp(M);
  for (unsigned k = 0; k < M; k++) {
    cf xa = X[k], xb = X[M - k];
    double a = M_PI * (double)k / (double)M;
    float wr = (float)cos(a), wi = (float)sin(a);
    float sr = xa.x + xb.x, si = xa.y - xb.y;
    float dr = xa.x - xb.x, di = xa.y + xb.y;
    float tr = dr * wr - di * wi, ti = dr * wi + di * wr;
    Zp[k] = psfft::mk(sr - ti, si + tr);
  }
  return cfft<true>(n1, n2, Zp.data(), reinterpret_cast<cf*>(x));
}
}
