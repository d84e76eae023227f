// dedisp.cu -- incoherent dedispersion of a SIGPROC filterbank to 8-bit DM trials: the step
// immediately BEFORE the search hot path.
//
// The reference delegates this to the external `dedisp` library through
// include/transforms/dedisperser.hpp:21-113 (dedisp_create_plan_multi, dedisp_generate_dm_list,
// dedisp_set_killmask, dedisp_get_max_delay, dedisp_execute with out_nbits = 8).  dedisp is not in
// the reference tree and has no pinned version; the arithmetic here restates its published
// algorithm (delay table 4.15e3/dt*(f^-2 - f0^-2), round-to-nearest delays, channel sum, the
// "2-bit HTRU" 8-bit output scaling) and is pinned end to end by the reference's golden
// candidate file (tests/test_gpu_filterbank.py; the CPU oracle carries the same restatement).
//
// Two kernels: unpack_transpose_kernel turns the packed [time][channel] samples into one byte per
// sample, channel-major, so that the dedispersion kernel's loads are contiguous in time;
// dedisperse_kernel sums the delayed channels for (DM trial, eight time samples) with packed
// 16-bit adds and scales to 8 bits.
#include <algorithm>
#include <cmath>
#include <vector>
#include "common.cuh"

namespace ps {

// [nsamps][nchans] packed LSB-first  ->  U[chan][pitch] bytes (killed channels zeroed)
__global__ void __launch_bounds__(256) unpack_transpose_kernel(const unsigned char* __restrict__ in, unsigned char* __restrict__ U,
                                                               const int* __restrict__ killmask, int nbits, int nchans,
                                                               size_t nsamps, size_t pitch) {
  __shared__ unsigned char tile[32][33];
  // tile of 32 channels x 32 samples
  const int c0 = blockIdx.y * 32, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 8 rows of 32
  const size_t t0 = (size_t)blockIdx.x * 32;
  const int per = 8 / nbits, mask = (1 << nbits) - 1;
  for (int r = ty; r < 32; r += 8) {
    size_t t = t0 + r;
    int c = c0 + tx;
    unsigned char v = 0;
    if (t < nsamps && c < nchans) {
      size_t s = t * nchans + c;  // sample index
      unsigned char byte = in[s / per];
      v = (byte >> ((s % per) * nbits)) & mask;
      if (killmask && !killmask[c]) v = 0;
    }
    tile[r][tx] = v;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    int c = c0 + r;
    size_t t = t0 + tx;
    if (t < nsamps && c < nchans) U[(size_t)c * pitch + t] = tile[tx][r];
  }
}

// out[dm][t] = scale( sum_c U[c][t + rn(dm * delay_table[c])] ), eight consecutive samples per thread.
// For every channel the eight bytes at t + delay are cut out of three aligned 32-bit words with two
// funnel shifts, split into even / odd bytes and added as packed 16-bit lanes (255 * 256 channels
// fit in 16 bits); every 256 channels the lanes are widened into 32-bit sums.  U rows are `pitch`
// bytes apart (a multiple of 16, >= nsamps + 16) so every row starts word-aligned and the last
// group may read past the row's end; out rows are `opitch` bytes apart (a multiple of 8).
__global__ void __launch_bounds__(256) dedisperse_kernel(const unsigned char* __restrict__ U, const float* __restrict__ delay_table,
                                                         const float* __restrict__ dm_list, unsigned char* __restrict__ out,
                                                         int nchans, size_t pitch, size_t nout, size_t opitch, float in_range,
                                                         float factor) {
  extern __shared__ unsigned delays[];
  // blockIdx.x = DM trial (the fast grid dimension): CTAs that run at the same time work on the same
  // stretch of samples for neighbouring DM trials, so the unpacked filterbank is read from DRAM about
  // once instead of once per DM trial (measured 549 GB -> see profiles for 256 trials of a 2.1 GB file)
  const int dm_idx = blockIdx.x;
  const float dm = dm_list[dm_idx];
  for (int c = threadIdx.x; c < nchans; c += blockDim.x) delays[c] = __float2uint_rn(__fmul_rn(dm, delay_table[c]));
  __syncthreads();
  const unsigned* __restrict__ W = reinterpret_cast<const unsigned*>(U);
  const size_t ngroups = (nout + 7) / 8;
  for (size_t g = (size_t)blockIdx.y * blockDim.x + threadIdx.x; g < ngroups; g += (size_t)gridDim.y * blockDim.x) {
    const size_t t = g * 8;
    unsigned sum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c0 = 0; c0 < nchans; c0 += 256) {
      unsigned e0 = 0, o0 = 0, e1 = 0, o1 = 0;  // packed 16-bit lanes: samples (0,2) (1,3) (4,6) (5,7)
      const int c1 = min(c0 + 256, nchans);
      for (int c = c0; c < c1; c++) {
        const size_t byte = (size_t)c * pitch + t + delays[c];
        const unsigned* w = W + (byte >> 2);
        const unsigned sh = (unsigned)(byte & 3) * 8;
        const unsigned w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
        const unsigned a = __funnelshift_r(w0, w1, sh), b = __funnelshift_r(w1, w2, sh);
        e0 += a & 0x00ff00ffu;
        o0 += (a >> 8) & 0x00ff00ffu;
        e1 += b & 0x00ff00ffu;
        o1 += (b >> 8) & 0x00ff00ffu;
      }
      sum[0] += e0 & 0xffffu; sum[2] += e0 >> 16; sum[1] += o0 & 0xffffu; sum[3] += o0 >> 16;
      sum[4] += e1 & 0xffffu; sum[6] += e1 >> 16; sum[5] += o1 & 0xffffu; sum[7] += o1 >> 16;
    }
    // (float)sum * out_range / (in_range * nchans) * factor, clipped to [0, 255], truncated
    unsigned lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      float scaled = __fmul_rn(__fdiv_rn(__fmul_rn((float)sum[i], 255.0f), __fmul_rn(in_range, (float)nchans)), factor);
      scaled = fminf(fmaxf(scaled, 0.0f), 255.0f);
      const unsigned v = (unsigned)(unsigned char)scaled;
      if (i < 4) lo |= v << (8 * i); else hi |= v << (8 * (i - 4));
    }
    *reinterpret_cast<uint2*>(out + (size_t)dm_idx * opitch + t) = make_uint2(lo, hi);
  }
}

}  // namespace ps

using namespace ps;

extern "C" {

int ps_dedisp_generate_dm_list(float dm_start, float dm_end, float dt, float ti, float f0, float df, int nchans, float tol,
                               float* out, int max_out) {
  // dedisp_generate_dm_list (called at dedisperser.hpp:57): published recurrence, double precision
  double dtu = (double)dt * 1e6;
  double f = ((double)f0 + ((nchans / 2) - 0.5) * (double)df) * 1e-3;
  double tol2 = (double)tol * tol;
  double a = 8.3 * df / (f * f * f);
  double a2 = a * a;
  double b2 = a2 * (double)(nchans * nchans / 16.0);
  double c = (dtu * dtu + (double)ti * ti) * (tol2 - 1.0);
  std::vector<double> dms;
  dms.push_back(dm_start);
  while (dms.back() < dm_end) {
    double prev = dms.back(), prev2 = prev * prev;
    double k = c + tol2 * a2 * prev2;
    double dm = (b2 * prev + sqrt(-a2 * b2 * prev2 + (a2 + b2) * k)) / (a2 + b2);
    dms.push_back(dm);
    if (dms.size() > (1u << 22)) return invalid("ps_dedisp_generate_dm_list: DM list does not terminate");
  }
  for (size_t i = 0; i < dms.size() && (int)i < max_out; i++) out[i] = (float)dms[i];
  return (int)dms.size();
}

// delay table and max delay as dedisp computes them (float arithmetic)
static void delay_table_host(std::vector<float>& tab, int nchans, float dt, float f0, float df) {
  tab.resize(nchans);
  for (int c = 0; c < nchans; c++) {
    float a = 1.f / (f0 + c * df);
    float b = 1.f / f0;
    tab[c] = (float)(4.15e3 / dt * (a * a - b * b));
  }
}

long ps_dedisp_max_delay(const float* h_dm_list, int ndm, int nchans, float dt, float f0, float df) {
  if (ndm <= 0 || nchans <= 0) return 0;
  std::vector<float> tab;
  delay_table_host(tab, nchans, dt, f0, df);
  return (long)(h_dm_list[ndm - 1] * tab[nchans - 1] + 0.5f);
}

// One time gulp at a time: gulp g covers output samples [t0, t0 + G) and needs input samples
// [t0, t0 + G + max_delay).  Device memory held: the packed gulp, its one-byte-per-sample unpacked
// copy (4-8x the packed size for 1/2-bit data) and the ndm x G output block -- sized from the free
// memory of the device (or max_device_bytes), so a long observation or a long DM list costs more
// gulps, not a failed cudaMalloc.  dedisp gulps the same way (its gulp_size); every output sample is
// computed exactly as in the single-gulp case.
int ps_dedisperse_on(int device, size_t max_device_bytes, const unsigned char* h_in, size_t nsamps, int nchans, int nbits, float dt,
                     float f0, float df, const float* h_dm_list, int ndm, const int* h_killmask, unsigned char* h_out, size_t nout) {
  if (nbits != 1 && nbits != 2 && nbits != 4 && nbits != 8) return invalid("ps_dedisperse: nbits must be 1, 2, 4 or 8");
  if (ndm <= 0 || nchans <= 0 || !nsamps) return invalid("ps_dedisperse: empty input");
  std::vector<float> tab;
  delay_table_host(tab, nchans, dt, f0, df);
  // delays are unsigned sample counts (dedisp's delay table is relative to the FIRST, highest-frequency
  // channel): a band given in ascending order, or a negative DM, would need samples before the start
  for (int i = 0; i < ndm; i++)
    for (int c = 0; c < nchans; c++)
      if (h_dm_list[i] * tab[c] < -0.5f)
        return invalid("ps_dedisperse: negative delay (channel frequencies must descend from fch1, DM trials must be >= 0)");
  long max_delay = ps_dedisp_max_delay(h_dm_list, ndm, nchans, dt, f0, df);
  for (int i = 0; i < ndm; i++)  // the list need not be sorted: the largest delay of any trial and channel
    for (int c = 0; c < nchans; c++) max_delay = std::max(max_delay, (long)(h_dm_list[i] * tab[c] + 0.5f));
  if (max_delay != ps_dedisp_max_delay(h_dm_list, ndm, nchans, dt, f0, df))
    return invalid("ps_dedisperse: the DM list must ascend (dedisp takes max_delay from its last entry)");
  if ((size_t)max_delay >= nsamps || nout != nsamps - (size_t)max_delay)
    return invalid("ps_dedisperse: nout must equal nsamps - max_delay");
  int prev_device = -1;
  if (device >= 0) {
    PS_CUDA(cudaGetDevice(&prev_device), "ps_dedisperse");
    PS_CUDA(cudaSetDevice(device), "ps_dedisperse");
  }
  struct Restore {
    int d;
    ~Restore() { if (d >= 0) cudaSetDevice(d); }
  } restore{prev_device};

  // gulp length: multiple of 64 output samples (packed bytes and 8-sample groups stay aligned)
  size_t budget = max_device_bytes;
  if (!budget) {
    size_t free_b = 0, total_b = 0;
    PS_CUDA(cudaMemGetInfo(&free_b, &total_b), "ps_dedisperse");
    budget = free_b / 10 * 6;
  }
  const double per_sample = (double)nchans * nbits / 8.0 + (double)nchans + (double)ndm;   // packed + unpacked + output
  const double fixed = (double)max_delay * ((double)nchans * nbits / 8.0 + (double)nchans) + 64.0 * nchans + (1 << 20);
  size_t gulp = nout;
  if (fixed + per_sample * (double)gulp > (double)budget) {
    if ((double)budget <= fixed + per_sample * 64.0) return invalid("ps_dedisperse: not enough device memory for one 64-sample gulp");
    gulp = (size_t)(((double)budget - fixed) / per_sample) / 64 * 64;
  }
  const size_t gin = gulp + (size_t)max_delay;                  // input samples per gulp
  const size_t pitch = (gin + 16 + 15) & ~(size_t)15;           // U rows: word-aligned, 16 readable bytes past the end
  const size_t opitch = (gulp + 7) & ~(size_t)7;                // trial rows on the device: whole 8-sample groups
  const size_t in_cap = (gin * nchans * nbits + 7) / 8;
  unsigned char *d_in = 0, *d_U = 0, *d_out = 0;
  float *d_tab = 0, *d_dm = 0;
  int* d_kill = 0;
  cudaError_t e = cudaMalloc(&d_in, in_cap);
  if (e == cudaSuccess) e = cudaMalloc(&d_U, (size_t)nchans * pitch + 16);
  if (e == cudaSuccess) e = cudaMalloc(&d_out, (size_t)ndm * opitch);
  if (e == cudaSuccess) e = cudaMalloc(&d_tab, sizeof(float) * nchans);
  if (e == cudaSuccess) e = cudaMalloc(&d_dm, sizeof(float) * ndm);
  if (e == cudaSuccess && h_killmask) e = cudaMalloc(&d_kill, sizeof(int) * nchans);
  if (e == cudaSuccess) e = cudaMemcpy(d_tab, tab.data(), sizeof(float) * nchans, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_dm, h_dm_list, sizeof(float) * ndm, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && h_killmask) e = cudaMemcpy(d_kill, h_killmask, sizeof(int) * nchans, cudaMemcpyHostToDevice);
  const float in_range = (float)((1 << nbits) - 1);
  const float factor = ((3.f * 1024.f) / 255.f) / 16.f;
  for (size_t t0 = 0; e == cudaSuccess && t0 < nout; t0 += gulp) {
    const size_t g_out = std::min(gulp, nout - t0);            // output samples of this gulp
    const size_t g_in = g_out + (size_t)max_delay;              // <= nsamps - t0
    const size_t first_bit = t0 * (size_t)nchans * nbits;       // t0 is a multiple of 64: byte-aligned
    const size_t bytes = (g_in * nchans * nbits + 7) / 8;
    e = cudaMemcpy(d_in, h_in + first_bit / 8, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_U, 0, (size_t)nchans * pitch + 16);
    if (e != cudaSuccess) break;
    dim3 g1((unsigned)((g_in + 31) / 32), (unsigned)((nchans + 31) / 32));
    unpack_transpose_kernel<<<g1, 256>>>(d_in, d_U, d_kill, nbits, nchans, g_in, pitch);
    g_launches.fetch_add(1);
    dim3 g2((unsigned)ndm, (unsigned)std::min<size_t>((g_out / 8 + 255) / 256 + 1, 65535));
    dedisperse_kernel<<<g2, 256, sizeof(unsigned) * nchans>>>(d_U, d_tab, d_dm, d_out, nchans, pitch, g_out, opitch, in_range, factor);
    g_launches.fetch_add(1);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy2D(h_out + t0, nout, d_out, opitch, g_out, ndm, cudaMemcpyDeviceToHost);
  }
  int rc = PS_OK;
  if (e != cudaSuccess) rc = cuda_fail(e, "ps_dedisperse");
  cudaFree(d_in); cudaFree(d_U); cudaFree(d_out); cudaFree(d_tab); cudaFree(d_dm); cudaFree(d_kill);
  return rc;
}

int ps_dedisperse(const unsigned char* h_in, size_t nsamps, int nchans, int nbits, float dt, float f0, float df,
                  const float* h_dm_list, int ndm, const int* h_killmask, unsigned char* h_out, size_t nout) {
  return ps_dedisperse_on(-1, 0, h_in, nsamps, nchans, nbits, dt, f0, df, h_dm_list, ndm, h_killmask, h_out, nout);
}

}  // extern "C"
