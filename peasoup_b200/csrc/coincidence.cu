// coincidence.cu -- multibeam coincidence matching.
// Reference: device_coincidencer / coincidence_kernel (src/kernels.cu:1073-1100): for every sample
// (or Fourier bin) count the beams above `thresh`; the mask is 1 where fewer than `beam_thresh`
// beams fire, 0 where the signal is seen in many beams at once (terrestrial interference).
// The reference kernel has no bounds check on the tail block; this one does.
#include "common.cuh"

namespace ps {

__global__ void __launch_bounds__(256) coincidence_kernel(float* const* __restrict__ arrays, float* __restrict__ out, int narrays,
                                                          size_t size, float thresh, int beam_thresh) {
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < size; idx += (size_t)gridDim.x * blockDim.x) {
    int count = 0;
    for (int b = 0; b < narrays; b++)
      if (arrays[b][idx] > thresh) count++;
    out[idx] = count < beam_thresh ? 1.0f : 0.0f;
  }
}

}  // namespace ps

extern "C" int ps_coincidencer(float* const* d_arrays, float* d_out, int narrays, size_t size, float thresh, int beam_thresh,
                               void* stream) {
  if (narrays < 1 || !d_arrays || !d_out) return ps::invalid("ps_coincidencer: need at least one array");
  if (!size) return PS_OK;
  ps::coincidence_kernel<<<ps::ew_grid(size, 256), 256, 0, ps::S(stream)>>>(d_arrays, d_out, narrays, size, thresh, beam_thresh);
  PS_LAUNCHED("Error from device_coincidencer");
  return PS_OK;
}
