// Stand-in for third_party/tinyexr (submodule pinned at 5fcb4dcb6e3abf96214b67e5c54db1ceec6a455c,
// absent from /root/reference: .SUBMODULES.json:9-15).  The reference calls exactly three names,
// src/image.h:7,54,56,59.  assets/bathroom.exr is missing too (.MISSING_LARGE_BLOBS), so LoadEXR
// synthesises a deterministic equirectangular RGBA float environment instead of decoding a file;
// image::convert_to_bytes (src/image.h:99-117) then quantises it to 8 bits exactly as it would a
// decoded EXR.  EXR decoding itself is therefore "parity unpinned"  — both sides of
// every comparison consume the same decoded bytes.  Integer/rational arithmetic only, so the bytes
// do not depend on libm.
#pragma once
#include <cstdlib>
#define TINYEXR_SUCCESS 0
static const int kStandinEnvWidth = 1024, kStandinEnvHeight = 512;
inline int LoadEXR(float **out_rgba, int *width, int *height, const char *filename, const char **err) {
  (void)filename;
  if (err) *err = nullptr;
  const int W = kStandinEnvWidth, H = kStandinEnvHeight;
  float *px = static_cast<float *>(malloc(sizeof(float) * 4 * W * H)); // freed by STBI_FREE == free, src/image.h:31
  if (!px) return -1;
  for (int y = 0; y < H; y++) {
    for (int x = 0; x < W; x++) {
      float *q = px + 4 * (y * W + x);
      float v = float(y) / float(H - 1);       // 0 = top of the image (straight up), 1 = bottom
      float u = float(x) / float(W);
      bool sky = y < H / 2;
      float r, g, b;
      if (sky) { // gradient sky, brighter toward the horizon
        r = 0.25f + 0.5f * v * 2.0f * 0.6f;
        g = 0.40f + 0.5f * v * 2.0f * 0.5f;
        b = 0.85f - 0.2f * v;
      } else { // tiled floor
        int cx = (x / 32) & 1, cy = (y / 32) & 1;
        float k = (cx ^ cy) ? 0.55f : 0.20f;
        r = k;
        g = k * 0.9f;
        b = k * 0.7f;
      }
      // a bright window (values > 1 clamp to 255 like any HDR texel) and two coloured panels
      if (x >= 200 && x < 264 && y >= 96 && y < 160) r = g = b = 4.0f;
      if (x >= 600 && x < 700 && y >= 180 && y < 230) { r = 0.9f; g = 0.15f; b = 0.1f; }
      if (x >= 820 && x < 900 && y >= 120 && y < 220) { r = 0.1f; g = 0.7f; b = 0.2f; }
      r += 0.1f * u;
      q[0] = r;
      q[1] = g;
      q[2] = b;
      q[3] = 1.0f;
    }
  }
  *out_rgba = px;
  *width = W;
  *height = H;
  return TINYEXR_SUCCESS;
}
inline void FreeEXRErrorMessage(const char *) {}
