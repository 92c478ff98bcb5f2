// output.cuh — the output stage of camera::render on the device.
//
// The reference writes an ASCII PPM: "P3\nW H\n255\n" (src/camera.h:149-151) and then, pixel by pixel,
// write_color (src/color.h:22-36): gamma 2.2, int(255.999 * c) UNCLAMPED, `out << r << ' ' << g << ' ' << b << '\n'`.
// At 3840x2160 that is 25 M integer-to-text conversions through one std::ostream, which dominates the end-to-end time once
// the render itself takes tens of milliseconds.  Here the text is produced by two kernels: measure (three integers and
// the line length per pixel, per-CTA totals), then after a scan of the CTA totals, write (each CTA re-scans its 1024 line
// lengths in shared memory and prints its lines at their final byte offsets).  The host receives the finished file image.
#pragma once
#include <cstdint>

namespace b200 {

#define B200_PPM_CTA 256
#define B200_PPM_PER_THREAD 4

__device__ inline int32_t ppm_component(double sum, double spp) { // pixel_color /= spp; linear_to_gamma; int(255.999 * c)
  const double c = sum / spp;
  const double g = c > 0 ? pow(c, 1 / 2.2) : 0.0;
  const double v = 255.999 * g;
  return (v > -2147483649.0 && v < 2147483648.0) ? (int32_t)v : (-2147483647 - 1); // what x86 cvttsd2si gives the reference
}
__device__ inline int ppm_digits(int32_t v) { // characters operator<<(int) prints
  uint32_t a = v < 0 ? 0u - (uint32_t)v : (uint32_t)v;
  int n = v < 0 ? 2 : 1;
  while (a >= 10u) a /= 10u, n++;
  return n;
}
__device__ inline char *ppm_print(char *p, int32_t v) {
  uint32_t a = v < 0 ? 0u - (uint32_t)v : (uint32_t)v;
  char tmp[10];
  int n = 0;
  do tmp[n++] = (char)('0' + a % 10u), a /= 10u;
  while (a);
  if (v < 0) *p++ = '-';
  while (n) *p++ = tmp[--n];
  return p;
}

// CTA-wide exclusive scan of one value per thread (256 threads); returns the exclusive prefix, *total = CTA sum
__device__ inline uint32_t ppm_cta_scan(uint32_t v, uint32_t *total) {
  __shared__ uint32_t s_warp[8];
  __shared__ uint32_t s_total;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t x = v;
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, off);
    if (lane >= off) x += y;
  }
  if (lane == 31) s_warp[w] = x;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t acc = 0;
    for (int k = 0; k < 8; k++) {
      const uint32_t t = s_warp[k];
      s_warp[k] = acc;
      acc += t;
    }
    s_total = acc;
  }
  __syncthreads();
  const uint32_t r = s_warp[w] + x - v;
  *total = s_total;
  __syncthreads();
  return r;
}

// pass 1: integers + line lengths; one CTA covers 1024 consecutive pixels
__global__ void __launch_bounds__(B200_PPM_CTA) k_ppm_measure(const double *sum, size_t npix, double spp, int32_t *rgb, unsigned long long *cta_bytes) {
  const size_t base = ((size_t)blockIdx.x * B200_PPM_CTA + threadIdx.x) * B200_PPM_PER_THREAD;
  uint32_t len = 0;
  for (int k = 0; k < B200_PPM_PER_THREAD; k++) {
    const size_t i = base + k;
    if (i >= npix) break;
    const int32_t r = ppm_component(sum[3 * i], spp), g = ppm_component(sum[3 * i + 1], spp), b = ppm_component(sum[3 * i + 2], spp);
    rgb[3 * i] = r, rgb[3 * i + 1] = g, rgb[3 * i + 2] = b;
    len += (uint32_t)(ppm_digits(r) + ppm_digits(g) + ppm_digits(b) + 3);
  }
  uint32_t total;
  ppm_cta_scan(len, &total);
  if (threadIdx.x == 0) cta_bytes[blockIdx.x] = total;
}

// exclusive scan of the CTA totals (a few thousand values): one CTA, sequential over chunks of 256
__global__ void __launch_bounds__(B200_PPM_CTA) k_ppm_scan(unsigned long long *cta_bytes, uint32_t n_cta, unsigned long long header, unsigned long long *file_bytes) {
  __shared__ unsigned long long s_carry;
  if (threadIdx.x == 0) s_carry = header;
  __syncthreads();
  for (uint32_t c0 = 0; c0 < n_cta; c0 += B200_PPM_CTA) {
    const uint32_t i = c0 + threadIdx.x;
    const unsigned long long v = i < n_cta ? cta_bytes[i] : 0ull;
    uint32_t total;
    const uint32_t ex = ppm_cta_scan((uint32_t)v, &total); // a CTA writes at most 1024 * 36 bytes: 32 bits are plenty per chunk
    const unsigned long long carry = s_carry;
    if (i < n_cta) cta_bytes[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *file_bytes = s_carry;
}

// pass 2: print
__global__ void __launch_bounds__(B200_PPM_CTA) k_ppm_write(const int32_t *rgb, size_t npix, const unsigned long long *cta_offset, char *out, size_t capacity) {
  const size_t base = ((size_t)blockIdx.x * B200_PPM_CTA + threadIdx.x) * B200_PPM_PER_THREAD;
  int32_t v[B200_PPM_PER_THREAD][3];
  uint32_t len = 0;
  int cnt = 0;
  for (int k = 0; k < B200_PPM_PER_THREAD; k++) {
    const size_t i = base + k;
    if (i >= npix) break;
    for (int c = 0; c < 3; c++) v[k][c] = rgb[3 * i + c];
    len += (uint32_t)(ppm_digits(v[k][0]) + ppm_digits(v[k][1]) + ppm_digits(v[k][2]) + 3);
    cnt++;
  }
  uint32_t total;
  const uint32_t ex = ppm_cta_scan(len, &total);
  const unsigned long long at = cta_offset[blockIdx.x] + ex;
  if (at + len > capacity) return; // the host reports the required size
  char *p = out + at;
  for (int k = 0; k < cnt; k++) {
    p = ppm_print(p, v[k][0]);
    *p++ = ' ';
    p = ppm_print(p, v[k][1]);
    *p++ = ' ';
    p = ppm_print(p, v[k][2]);
    *p++ = '\n';
  }
}

// ------------------------------------------------------------------------------------------------ binary writers
// Beside the reference's P3 text: the same pixels as a binary PPM and as an OpenEXR file.
//   P6 : "P6\nW H\n255\n" + one byte per component = the integer write_color prints (src/color.h:22-36), clamped to
//        [0, 255] because a byte cannot hold the reference's unclamped values (emitters print > 255, NaN prints INT_MIN).
//   EXR: the LINEAR image pixel_color / spp (src/camera.h:169) as 32-bit floats, no gamma, no clamp: single-part scan-line
//        file, NO_COMPRESSION, channels B, G, R (alphabetical, as the format requires), increasing y.  The device writes
//        every scan-line block {y, byte count, B row, G row, R row} at its final offset; the host writes header + offsets.
__global__ void __launch_bounds__(256) k_pack_p6(const double *sum, size_t n_comp, double spp, unsigned char *out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_comp; i += (size_t)gridDim.x * blockDim.x) {
    const int32_t v = ppm_component(sum[i], spp);
    out[i] = (unsigned char)(v < 0 ? 0 : (v > 255 ? 255 : v));
  }
}
__global__ void __launch_bounds__(256) k_pack_exr(const double *sum, int W, int H, double spp, unsigned char *blocks) {
  const size_t block_bytes = 8 + (size_t)W * 12, npix = (size_t)W * H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
    const int y = (int)(i / W), x = (int)(i % W);
    unsigned char *b = blocks + (size_t)y * block_bytes;
    if (x == 0) {
      reinterpret_cast<int32_t *>(b)[0] = y;
      reinterpret_cast<int32_t *>(b)[1] = W * 12;
    }
    float *row = reinterpret_cast<float *>(b + 8); // blocks start at a multiple of 4 (the host pads the header)
    row[x] = (float)(sum[3 * i + 2] / spp);         // B
    row[(size_t)W + x] = (float)(sum[3 * i + 1] / spp);     // G
    row[2 * (size_t)W + x] = (float)(sum[3 * i] / spp);     // R
  }
}

} // namespace b200
