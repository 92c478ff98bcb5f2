// noalloc.cuh — b200rt_format_ppm_noalloc: the P3 formatter of output.cuh for callers that must not reach cudaMalloc.
// Included at the end of api.cu (same translation unit: the __device__ array below and the cudaGetSymbolAddress call that
// resolves it have to live in one .cu file).  No counterpart in the reference.
//
// The two-pass formatter needs scratch: three integers per pixel, one byte count per CTA, the file length.  b200rt_format_ppm
// takes that from cudaMalloc per call; here it is a file-scope array of fixed size, so the entry point allocates nothing:
//
//   [0, 256)                       file length (8 bytes used)
//   [256, 256 + 8 * MAX_CTA)       per-CTA byte counts, then (after k_ppm_scan) per-CTA file offsets
//   [256 + 8 * MAX_CTA, end)       int32 r, g, b per pixel
//
// Every GPU that loads the library holds the array (12 MiB + 8.25 KiB) until the process exits.  All calls on a device share
// it and it keeps the last call's contents; a call writes every word it later reads (k_ppm_measure: all rgb of its npix pixels
// and all n_cta counts; k_ppm_scan: the total), so nothing has to be cleared, but calls that overlap on one GPU race.
#pragma once

namespace b200 {

constexpr size_t NOALLOC_MAX_CTA = (B200RT_NOALLOC_MAX_PIXELS + B200_PPM_CTA * B200_PPM_PER_THREAD - 1) / (B200_PPM_CTA * B200_PPM_PER_THREAD);
constexpr size_t NOALLOC_OFF_CTA = 256;
constexpr size_t NOALLOC_OFF_RGB = NOALLOC_OFF_CTA + 8 * NOALLOC_MAX_CTA;
constexpr size_t NOALLOC_BYTES = NOALLOC_OFF_RGB + (size_t)B200RT_NOALLOC_MAX_PIXELS * 3 * sizeof(int32_t);
static_assert(NOALLOC_OFF_RGB % 256 == 0, "scratch sections stay 256-byte aligned");

// In host code this name is a host-side stand-in, not a device address: never pass it to a kernel; use noalloc_scratch().
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

struct PpmHeader {
  char text[64];
};
__global__ void k_ppm_header(PpmHeader h, int len, char *out) {
  if ((int)threadIdx.x < len) out[threadIdx.x] = h.text[threadIdx.x];
}

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mutex;
static unsigned char *g_noalloc_ptr[NOALLOC_MAX_DEVICES]; // the array's address on each device, resolved once per device

// device address of g_noalloc_scratch on `device` (which must be current); nullptr + *err on failure
static unsigned char *noalloc_scratch(int device, cudaError_t *err) {
  *err = cudaSuccess;
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
    *err = cudaErrorInvalidDevice;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device]) {
    void *p = nullptr;
    *err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (*err != cudaSuccess) return nullptr;
    g_noalloc_ptr[device] = static_cast<unsigned char *>(p);
  }
  return g_noalloc_ptr[device];
}

} // namespace b200

extern "C" int b200rt_format_ppm_noalloc(b200rt_ctx *c, const double *d_rgb_sum, int total_spp, char *d_out, size_t capacity, size_t *out_len) {
  using namespace b200;
  if (!c || !d_rgb_sum || !d_out || !out_len) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!c->have_camera) return fail(B200RT_ERR_STATE, "format_ppm_noalloc before b200rt_set_camera");
  if (total_spp <= 0) return fail(B200RT_ERR_INVALID, "total_spp must be positive");
  const size_t npix = (size_t)c->cam.image_width * c->cam.image_height;
  if (npix > (size_t)B200RT_NOALLOC_MAX_PIXELS)
    return fail(B200RT_ERR_UNSUPPORTED, "format_ppm_noalloc: %d x %d = %zu pixels exceed the static scratch (B200RT_NOALLOC_MAX_PIXELS = %u); use b200rt_format_ppm",
                c->cam.image_width, c->cam.image_height, npix, (unsigned)B200RT_NOALLOC_MAX_PIXELS);
  PpmHeader header;
  const int hl = snprintf(header.text, sizeof header.text, "P3\n%d %d\n255\n", c->cam.image_width, c->cam.image_height); // src/camera.h:149-151
  if (capacity < (size_t)hl) return fail(B200RT_ERR_INVALID, "format_ppm_noalloc: capacity %zu is smaller than the header", capacity);
  CU(cudaSetDevice(c->device));
  cudaError_t err;
  unsigned char *scratch = noalloc_scratch(c->device, &err);
  if (!scratch) return fail(B200RT_ERR_CUDA, "format_ppm_noalloc: cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", c->device, cudaGetErrorString(err));
  unsigned long long *d_total = reinterpret_cast<unsigned long long *>(scratch);
  unsigned long long *d_cta = reinterpret_cast<unsigned long long *>(scratch + NOALLOC_OFF_CTA);
  int32_t *d_rgb = reinterpret_cast<int32_t *>(scratch + NOALLOC_OFF_RGB);
  const uint32_t n_cta = (uint32_t)((npix + B200_PPM_CTA * B200_PPM_PER_THREAD - 1) / (B200_PPM_CTA * B200_PPM_PER_THREAD));
  k_ppm_header<<<1, 64, 0, c->stream>>>(header, hl, d_out);
  k_ppm_measure<<<n_cta, B200_PPM_CTA, 0, c->stream>>>(d_rgb_sum, npix, (double)total_spp, d_rgb, d_cta);
  k_ppm_scan<<<1, B200_PPM_CTA, 0, c->stream>>>(d_cta, n_cta, (unsigned long long)hl, d_total);
  k_ppm_write<<<n_cta, B200_PPM_CTA, 0, c->stream>>>(d_rgb, npix, d_cta, d_out, capacity);
  CU(cudaGetLastError());
  unsigned long long total = 0;
  CU(cudaMemcpyAsync(&total, d_total, 8, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  *out_len = (size_t)total;
  if (total > capacity) return fail(B200RT_ERR_INVALID, "format_ppm_noalloc: the file needs %llu bytes, capacity is %zu (b200rt_ppm_bound() always suffices)", total, capacity);
  return B200RT_OK;
}
