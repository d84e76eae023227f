// fft_core.cuh -- CTA-level complex FFT building blocks (sm_100a; also compiles as plain
// C++ so that tests/cpu_emul can execute the exact same index math on the host).
//
// Replaces the cuFFT plans of the reference (include/transforms/ffter.hpp:46-78) with a
// hand-written two-pass ("four-step") transform.  One CTA of 512 threads performs F
// independent length-n FFTs at a time, n = 16*16*R3 (R3 in {1,2,4,8,16} => n = 256..4096),
// F = 32/R3, so that F*R3 = 32 and every shared-memory access of a warp is 32 consecutive
// words (bank-conflict free, see DESIGN.md "FFT shared-memory layout").
//
// Each thread owns 16 complex values.  Decimation in frequency with digits
//   j = a*(n/16) + b*R3 + c         (a,b in [0,16), c in [0,R3))
//   k = ka + 16*kb + 256*kc
// phase A: radix-16 over a, twiddle W_n^(ka*(b*R3+c))        -> exchange E1
// phase B: radix-16 over b, twiddle W_(16*R3)^(kb*c)          -> exchange E2
// phase C: radix-R3 over c                                    -> Storer(f, k, value)
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <type_traits>

#if defined(__CUDACC__)
#define PS_HD __host__ __device__ __forceinline__
#else
#define PS_HD inline
#endif

namespace psfft {

struct alignas(8) cf {
  float x, y;
};

PS_HD cf mk(float x, float y) {
  cf r;
  r.x = x;
  r.y = y;
  return r;
}
// read-only (non-coherent) load of a table entry
PS_HD cf ldg_cf(const cf* p) {
#if defined(__CUDA_ARCH__)
  float2 v = __ldg(reinterpret_cast<const float2*>(p));
  return mk(v.x, v.y);
#else
  return *p;
#endif
}
PS_HD cf cadd(cf a, cf b) { return mk(a.x + b.x, a.y + b.y); }
PS_HD cf csub(cf a, cf b) { return mk(a.x - b.x, a.y - b.y); }
PS_HD cf cmul(cf a, cf b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
PS_HD cf cconj(cf a) { return mk(a.x, -a.y); }
// multiply by -i (forward transform) or +i (inverse transform)
template <bool INV>
PS_HD cf rot(cf a) {
  return INV ? mk(-a.y, a.x) : mk(a.y, -a.x);
}
// twiddle tables hold forward roots exp(-2*pi*i*q/n); inverse transforms conjugate them
template <bool INV>
PS_HD cf dirw(cf w) {
  return INV ? mk(w.x, -w.y) : w;
}

template <bool INV>
PS_HD void dft2(cf& a, cf& b) {
  cf t = a;
  a = cadd(t, b);
  b = csub(t, b);
}

template <bool INV>
PS_HD void dft4(cf& a, cf& b, cf& c, cf& d) {
  cf t0 = cadd(a, c), t1 = csub(a, c), t2 = cadd(b, d), t3 = rot<INV>(csub(b, d));
  a = cadd(t0, t2);
  c = csub(t0, t2);
  b = cadd(t1, t3);
  d = csub(t1, t3);
}

#define PS_SQH 0.70710678118654752440f
#define PS_C16 0.92387953251128675613f
#define PS_S16 0.38268343236508977173f

// v[k] <- sum_a v[a] W8^(a*k), natural order in and out
template <bool INV>
PS_HD void dft8(cf* v) {
  // a = a0 + 2*a1: radix-4 over a1 for a0 = 0,1
  dft4<INV>(v[0], v[2], v[4], v[6]);  // T[0][ka] at v[2*ka]
  dft4<INV>(v[1], v[3], v[5], v[7]);  // T[1][ka] at v[2*ka+1]
  // twiddle T[1][ka] *= W8^ka
  const float s = INV ? PS_SQH : -PS_SQH;
  v[3] = cmul(v[3], mk(PS_SQH, s));
  v[5] = rot<INV>(v[5]);
  v[7] = cmul(v[7], mk(-PS_SQH, s));
  cf o[8];
#pragma unroll
  for (int ka = 0; ka < 4; ka++) {
    o[ka] = cadd(v[2 * ka], v[2 * ka + 1]);
    o[ka + 4] = csub(v[2 * ka], v[2 * ka + 1]);
  }
#pragma unroll
  for (int i = 0; i < 8; i++) v[i] = o[i];
}

template <bool INV>
PS_HD void dft16(cf* v) {
  // a = a0 + 4*a1, k = ka + 4*kb : radix-4 over a1 (per a0), twiddle W16^(a0*ka), radix-4 over a0
#pragma unroll
  for (int a0 = 0; a0 < 4; a0++) dft4<INV>(v[a0], v[a0 + 4], v[a0 + 8], v[a0 + 12]);
  // now T[a0][ka] at v[a0 + 4*ka]
  const float sg = INV ? 1.f : -1.f;
  v[1 + 4 * 1] = cmul(v[1 + 4 * 1], mk(PS_C16, sg * PS_S16));   // W16^1
  v[1 + 4 * 2] = cmul(v[1 + 4 * 2], mk(PS_SQH, sg * PS_SQH));   // W16^2
  v[1 + 4 * 3] = cmul(v[1 + 4 * 3], mk(PS_S16, sg * PS_C16));   // W16^3
  v[2 + 4 * 1] = cmul(v[2 + 4 * 1], mk(PS_SQH, sg * PS_SQH));   // W16^2
  v[2 + 4 * 2] = rot<INV>(v[2 + 4 * 2]);                        // W16^4
  v[2 + 4 * 3] = cmul(v[2 + 4 * 3], mk(-PS_SQH, sg * PS_SQH));  // W16^6
  v[3 + 4 * 1] = cmul(v[3 + 4 * 1], mk(PS_S16, sg * PS_C16));   // W16^3
  v[3 + 4 * 2] = cmul(v[3 + 4 * 2], mk(-PS_SQH, sg * PS_SQH));  // W16^6
  v[3 + 4 * 3] = cmul(v[3 + 4 * 3], mk(-PS_C16, -sg * PS_S16)); // W16^9
  cf o[16];
#pragma unroll
  for (int ka = 0; ka < 4; ka++) {
    cf a = v[0 + 4 * ka], b = v[1 + 4 * ka], c = v[2 + 4 * ka], d = v[3 + 4 * ka];
    dft4<INV>(a, b, c, d);  // over a0 -> kb
    o[ka + 4 * 0] = a;
    o[ka + 4 * 1] = b;
    o[ka + 4 * 2] = c;
    o[ka + 4 * 3] = d;
  }
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = o[i];
}

template <int R, bool INV>
PS_HD void dftR(cf* v) {
  if constexpr (R == 2) dft2<INV>(v[0], v[1]);
  if constexpr (R == 4) dft4<INV>(v[0], v[1], v[2], v[3]);
  if constexpr (R == 8) dft8<INV>(v);
  if constexpr (R == 16) dft16<INV>(v);
}

// v[k] *= w^k for k = 1..15, powers built as a balanced product tree (error ~log2(k) ulp)
PS_HD void twiddle_powers16(cf* v, cf w1) {
  cf w[16];
  w[1] = w1;
#pragma unroll
  for (int k = 2; k < 16; k++) w[k] = cmul(w[k / 2], w[k - k / 2]);
#pragma unroll
  for (int k = 1; k < 16; k++) v[k] = cmul(v[k], w[k]);
}

// w[k] = w1^k for k = 1..R-1 as a balanced product tree; w[0] is not touched
template <int R>
PS_HD void power_tree(cf* w, cf w1) {
  w[1] = w1;
#pragma unroll
  for (int k = 2; k < R; k++) w[k] = cmul(w[k / 2], w[k - k / 2]);
}

// w[j] = base * s^j, j < R, as two interleaved chains (even/odd j) of step s^2: R/2 roundings deep
template <int R>
PS_HD void chain_powers(cf* w, cf base, cf s1) {
  w[0] = base;
  if (R > 1) w[1] = cmul(base, s1);
  if (R > 2) {
    const cf s2 = cmul(s1, s1);
#pragma unroll
    for (int j = 2; j < R; j++) w[j] = cmul(w[j - 2], s2);
  }
}

// storers with kSequenced = true take phase C's calls through seq_begin / store_seq
template <class S, class = void>
struct is_sequenced { static constexpr bool value = false; };
template <class S>
struct is_sequenced<S, typename std::enable_if<S::kSequenced>::type> { static constexpr bool value = true; };

// One CTA-wide batch of F FFTs of length n = 256*R3.  512 threads; thread t = u*F + f.
// Shared memory holds complex words (8 bytes): a warp always touches 32 consecutive complex
// words (256 contiguous bytes, two conflict-free half-warp phases).
// Storer interface: st.template store<R>(f, k0, stride, v) receives outputs k0 + stride*j, j < R.
template <int R3, bool INV>
struct CtaFft {
  static constexpr int F = 32 / R3;
  static constexpr int G = 16 * R3;      // threads per FFT
  static constexpr int N = 256 * R3;     // FFT length
  static constexpr int THREADS = F * G;  // 512
  static constexpr int E1_WORDS = 16 * G * F;                          // complex words
  static constexpr int E2_WORDS = (R3 > 1) ? 256 * (R3 + 1) * F : 0;   // complex words
  static constexpr int SMEM_WORDS = 2 * (E1_WORDS + E2_WORDS);         // float words
  static constexpr size_t SMEM_BYTES = sizeof(float) * SMEM_WORDS;

  // W4096[q] = exp(-2*pi*i*q/4096), q in [0,4096)
  template <class Loader>
  PS_HD static void phaseA(int t, float* s, const cf* W4096, const Loader& ld) {
    const int f = t % F, u = t / F;
    cf v[16];
#pragma unroll
    for (int a = 0; a < 16; a++) v[a] = ld(f, a * G + u);
    phaseA_regs(t, s, W4096, v);
  }

  // v[a] = element a*G + u of FFT f, already in registers
  PS_HD static void phaseA_regs(int t, float* s, const cf* W4096, cf* v) {
    phaseA_compute(t, W4096, v);
    phaseA_write(t, s, v);
  }
  PS_HD static void phaseA_compute(int t, const cf* W4096, cf* v) {
    dft16<INV>(v);
    twiddle_powers16(v, dirw<INV>(twiddleA(t, W4096)));
  }
  // The two twiddle bases of a thread depend on its position only, not on the work item: a kernel
  // with registers to spare loads them once (twiddleA / twiddleB) and passes them in (phaseA_regs_w,
  // phaseB_w).
  PS_HD static cf twiddleA(int t, const cf* W4096) { return ldg_cf(&W4096[(t / F) * (4096 / N)]); }         // W_n^u
  PS_HD static cf twiddleB(int t, const cf* W4096) { return ldg_cf(&W4096[((t / F) % R3) * (256 / R3)]); }  // W_(16*R3)^c
  PS_HD static void phaseA_regs_w(int t, float* s, cf wA, cf* v) {
    dft16<INV>(v);
    twiddle_powers16(v, dirw<INV>(wA));
    phaseA_write(t, s, v);
  }
  PS_HD static void phaseA_write(int t, float* s, const cf* v) {
    const int f = t % F, u = t / F;
    cf* e1 = reinterpret_cast<cf*>(s);
#pragma unroll
    for (int ka = 0; ka < 16; ka++) e1[(ka * G + u) * F + f] = v[ka];
  }

  template <class Storer>
  PS_HD static void phaseB(int t, float* s, const cf* W4096, const Storer& st) {
    phaseB_w(t, s, R3 == 1 ? mk(1.f, 0.f) : twiddleB(t, W4096), st);
  }
  template <class Storer>
  PS_HD static void phaseB_w(int t, float* s, cf wB, const Storer& st) {
    const int f = t % F, u = t / F;
    const int ka = u / R3, c = u % R3;
    const cf* e1 = reinterpret_cast<const cf*>(s);
    cf v[16];
#pragma unroll
    for (int b = 0; b < 16; b++) v[b] = e1[(ka * G + b * R3 + c) * F + f];
    dft16<INV>(v);
    if (R3 == 1) {
      st.template store<16>(f, ka, 16, v);
      return;
    }
    twiddle_powers16(v, dirw<INV>(wB));
    cf* e2 = reinterpret_cast<cf*>(s) + E1_WORDS;
#pragma unroll
    for (int kb = 0; kb < 16; kb++) e2[((ka + 16 * kb) * (R3 + 1) + c) * F + f] = v[kb];
  }

  // phase C of a sequenced storer whose state was begun earlier (its table reads then fly during the
  // barrier in front of phase C): sq = st.seq_begin<G>(t % F, t / F, 256)
  template <class Storer, class Seq>
  PS_HD static void phaseC_seq(int t, const float* s, const Storer& st, Seq& sq) {
    if (R3 == 1) return;
    const int f = t % F, u = t / F;
    const cf* e2 = reinterpret_cast<const cf*>(s) + E1_WORDS;
    constexpr int RR = (R3 > 1) ? R3 : 2;
#pragma unroll
    for (int m = 0; m < 16 / RR; m++) {
      const int q = u + G * m;
      cf w[RR];
#pragma unroll
      for (int c = 0; c < RR; c++) w[c] = e2[(q * (R3 + 1) + c) * F + f];
      dftR<RR, INV>(w);
      st.template store_seq<RR>(f, q, 256, w, sq);
    }
  }

  template <class Storer>
  PS_HD static void phaseC(int t, const float* s, const Storer& st) {
    if (R3 == 1) return;
    const int f = t % F, u = t / F;
    const cf* e2 = reinterpret_cast<const cf*>(s) + E1_WORDS;
    constexpr int RR = (R3 > 1) ? R3 : 2;
    if constexpr (is_sequenced<Storer>::value) {
      // the storer derives the per-call state of call m + 1 from call m's (fft.cuh: StorePass1::store_seq)
      auto sq = st.template seq_begin<G>(f, u, 256);
      phaseC_seq(t, s, st, sq);
    } else {
#pragma unroll
      for (int m = 0; m < 16 / RR; m++) {
        const int q = u + G * m;
        cf w[RR];
#pragma unroll
        for (int c = 0; c < RR; c++) w[c] = e2[(q * (R3 + 1) + c) * F + f];
        dftR<RR, INV>(w);
        st.template store<RR>(f, q, 256, w);
      }
    }
  }
};

}  // namespace psfft
