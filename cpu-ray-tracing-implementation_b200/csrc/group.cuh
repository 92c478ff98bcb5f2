// group.cuh — the frame split across the GPUs of one box, inside the C ABI: b200rt_group_*.
//
// The reference's only parallel construct is the row loop of camera::render (src/camera.h:158); samples are independent
// (:163-170) and the scene is small and read-only.  So: one context per device with the scene replicated, device g of G
// renders the global sample indices [lo_g, hi_g) of every pixel (Philox draws are keyed by the global sample index: the set
// of samples does not depend on G), each on its own host thread and stream, and the fp64 accumulators (sums, not means) are
// combined onto the first device by ONE exchange step:
//   B200RT_GROUP_REDUCE_NCCL  ncclReduce(sum, root 0) over NVLink / NVSwitch, communicators from ncclCommInitAll
//                             (single process, one communicator per device); libnccl is loaded with dlopen when the first
//                             group is created, so a single-GPU user of libb200rt.so never needs it;
//   B200RT_GROUP_REDUCE_PEER  one kernel on the root device that reads the other devices' accumulators through peer
//                             (NVLink) loads and adds them in device order: a deterministic sum, no library.
// There is nothing to overlap the exchange with: its input is complete only when a device's last wave has retired.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <thread>

namespace {

struct NcclApi {
  void *handle = nullptr;
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclReduce) Reduce = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
  std::string error;
  bool load() {
    if (handle) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names)
      if ((handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!handle) {
      error = std::string("dlopen libnccl.so.2: ") + dlerror();
      return false;
    }
#define B200_NCCL_SYM(field, name)                                                                                     \
  field = (decltype(field))dlsym(handle, #name);                                                                        \
  if (!field) {                                                                                                        \
    error = "libnccl does not export " #name;                                                                          \
    return false;                                                                                                      \
  }
    B200_NCCL_SYM(CommInitAll, ncclCommInitAll)
    B200_NCCL_SYM(CommDestroy, ncclCommDestroy)
    B200_NCCL_SYM(Reduce, ncclReduce)
    B200_NCCL_SYM(GetErrorString, ncclGetErrorString)
    B200_NCCL_SYM(GetVersion, ncclGetVersion)
#undef B200_NCCL_SYM
    return true;
  }
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

#define B200_MAX_GROUP 16
struct PeerSrc {
  const double *p[B200_MAX_GROUP];
};
// dst += src[1] + src[2] + ... in device order; src[k] are peer-mapped pointers (loads travel over NVLink)
__global__ void __launch_bounds__(256) k_reduce_peers(double *dst, PeerSrc src, int n_src, size_t count) {
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; i < count; i += (size_t)gridDim.x * blockDim.x * 2) {
    if (i + 1 < count) {
      double2 a = *reinterpret_cast<const double2 *>(dst + i);
      for (int k = 0; k < n_src; k++) {
        const double2 b = *reinterpret_cast<const double2 *>(src.p[k] + i);
        a.x += b.x, a.y += b.y;
      }
      *reinterpret_cast<double2 *>(dst + i) = a;
    } else {
      double a = dst[i];
      for (int k = 0; k < n_src; k++) a += src.p[k][i];
      dst[i] = a;
    }
  }
}

} // namespace

struct b200rt_group {
  std::vector<b200rt_ctx *> ctx;
  std::vector<int> devices;
  std::vector<ncclComm_t> comms;
  std::vector<double *> accum; // per device, W*H*3
  std::vector<cudaStream_t> streams;
  std::vector<cudaEvent_t> done;
  size_t accum_elems = 0;
  int reduce_mode = B200RT_GROUP_REDUCE_NCCL;
  bool peer_ok = false;
  bool have_camera = false;
  b200rt_camera cam{};
};

namespace {

// runs fn(i) for every member on its own host thread; first failure wins (its message becomes the caller's last error)
template <class F> int group_parallel(b200rt_group *g, F fn) {
  const int n = (int)g->ctx.size();
  std::vector<int> rc(n, B200RT_OK);
  std::vector<std::string> msg(n);
  auto body = [&](int i) {
    rc[i] = fn(i);
    if (rc[i] != B200RT_OK) msg[i] = t_err;
  };
  if (n == 1) {
    body(0);
  } else {
    std::vector<std::thread> th;
    for (int i = 0; i < n; i++) th.emplace_back(body, i);
    for (auto &t : th) t.join();
  }
  for (int i = 0; i < n; i++)
    if (rc[i] != B200RT_OK) return fail(rc[i], "device %d: %s", g->devices[i], msg[i].c_str());
  return B200RT_OK;
}

} // namespace

extern "C" {

int b200rt_group_create(const int *devices, int n, b200rt_group **out) {
  if (!out) return fail(B200RT_ERR_INVALID, "b200rt_group_create: out is NULL");
  *out = nullptr;
  if (!devices || n < 1 || n > B200_MAX_GROUP) return fail(B200RT_ERR_INVALID, "b200rt_group_create: need 1..%d devices", B200_MAX_GROUP);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < i; j++)
      if (devices[i] == devices[j]) return fail(B200RT_ERR_INVALID, "b200rt_group_create: device %d listed twice", devices[i]);
  std::unique_ptr<b200rt_group> g(new b200rt_group);
  auto cleanup = [&](int rc) {
    const std::string keep = t_err;
    b200rt_group_destroy(g.release());
    t_err = keep;
    return rc;
  };
  for (int i = 0; i < n; i++) {
    b200rt_ctx *c = nullptr;
    int rc = b200rt_create(devices[i], &c);
    if (rc != B200RT_OK) return cleanup(rc);
    g->ctx.push_back(c);
    g->devices.push_back(devices[i]);
    cudaStream_t s = nullptr;
    cudaEvent_t e = nullptr;
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess)
      return cleanup(fail(B200RT_ERR_CUDA, "b200rt_group_create: stream / event on device %d", devices[i]));
    g->streams.push_back(s);
    g->done.push_back(e);
    g->accum.push_back(nullptr);
  }
  if (n > 1) {
    // peer access root <- others (for the peer reduce); absence is not an error, the NCCL path does not need it
    g->peer_ok = true;
    cudaSetDevice(devices[0]);
    for (int i = 1; i < n; i++) {
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devices[0], devices[i]);
      if (!can) {
        g->peer_ok = false;
        continue;
      }
      const cudaError_t e = cudaDeviceEnablePeerAccess(devices[i], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) g->peer_ok = false;
      cudaGetLastError();
    }
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (!g_nccl.load()) return cleanup(fail(B200RT_ERR_UNSUPPORTED, "b200rt_group_create: %s", g_nccl.error.c_str()));
    g->comms.resize(n);
    const ncclResult_t r = g_nccl.CommInitAll(g->comms.data(), n, devices);
    if (r != ncclSuccess) {
      g->comms.clear();
      return cleanup(fail(B200RT_ERR_CUDA, "ncclCommInitAll: %s", g_nccl.GetErrorString(r)));
    }
  }
  *out = g.release();
  return B200RT_OK;
}

int b200rt_group_destroy(b200rt_group *g) {
  if (!g) return B200RT_OK;
  for (size_t i = 0; i < g->ctx.size(); i++) {
    cudaSetDevice(g->devices[i]);
    if (i < g->streams.size() && g->streams[i]) cudaStreamSynchronize(g->streams[i]);
  }
  for (ncclComm_t cm : g->comms) g_nccl.CommDestroy(cm);
  for (size_t i = 0; i < g->ctx.size(); i++) {
    cudaSetDevice(g->devices[i]);
    if (i < g->accum.size() && g->accum[i]) cudaFree(g->accum[i]);
    if (i < g->streams.size() && g->streams[i]) cudaStreamDestroy(g->streams[i]);
    if (i < g->done.size() && g->done[i]) cudaEventDestroy(g->done[i]);
    b200rt_destroy(g->ctx[i]);
  }
  delete g;
  return B200RT_OK;
}

int b200rt_group_size(const b200rt_group *g) { return g ? (int)g->ctx.size() : 0; }

int b200rt_group_context(b200rt_group *g, int i, b200rt_ctx **out) {
  if (!g || !out || i < 0 || i >= (int)g->ctx.size()) return fail(B200RT_ERR_INVALID, "b200rt_group_context: bad argument");
  *out = g->ctx[i];
  return B200RT_OK;
}

int b200rt_group_set_option(b200rt_group *g, int option, double value) {
  if (!g) return fail(B200RT_ERR_INVALID, "group is NULL");
  if (option == B200RT_GROUP_OPT_REDUCE) {
    if (value != B200RT_GROUP_REDUCE_NCCL && value != B200RT_GROUP_REDUCE_PEER) return fail(B200RT_ERR_INVALID, "group reduce must be 0 (NCCL) or 1 (peer kernel)");
    if (value == B200RT_GROUP_REDUCE_PEER && g->ctx.size() > 1 && !g->peer_ok) return fail(B200RT_ERR_UNSUPPORTED, "peer access between the devices of the group is not available");
    g->reduce_mode = (int)value;
    return B200RT_OK;
  }
  for (b200rt_ctx *c : g->ctx) {
    const int rc = b200rt_set_option(c, option, value);
    if (rc != B200RT_OK) return rc;
  }
  return B200RT_OK;
}

int b200rt_group_upload_scene(b200rt_group *g, const b200rt_scene_desc *scene) {
  if (!g) return fail(B200RT_ERR_INVALID, "group is NULL");
  return group_parallel(g, [&](int i) { return b200rt_upload_scene(g->ctx[i], scene); });
}

int b200rt_group_set_camera(b200rt_group *g, const b200rt_camera *cam) {
  if (!g || !cam) return fail(B200RT_ERR_INVALID, "NULL argument");
  for (b200rt_ctx *c : g->ctx) {
    const int rc = b200rt_set_camera(c, cam);
    if (rc != B200RT_OK) return rc;
  }
  g->cam = *cam;
  g->have_camera = true;
  return B200RT_OK;
}

int b200rt_group_render(b200rt_group *g, uint64_t seed, int spp_begin, int spp_end, double *rgb_sum, b200rt_stats *stats) {
  if (!g || !rgb_sum) return fail(B200RT_ERR_INVALID, "NULL argument");
  if (!g->have_camera) return fail(B200RT_ERR_STATE, "render before b200rt_group_set_camera");
  if (spp_end < spp_begin || spp_begin < 0) return fail(B200RT_ERR_INVALID, "bad sample range [%d, %d)", spp_begin, spp_end);
  const int n = (int)g->ctx.size();
  const size_t elems = (size_t)g->cam.image_width * g->cam.image_height * 3;
  if (g->accum_elems < elems) {
    for (int i = 0; i < n; i++) {
      CU(cudaSetDevice(g->devices[i]));
      if (g->accum[i]) cudaFree(g->accum[i]);
      g->accum[i] = nullptr;
      CU(cudaMalloc((void **)&g->accum[i], elems * sizeof(double)));
    }
    g->accum_elems = elems;
  }
  std::vector<b200rt_stats> st(n);
  const int span = spp_end - spp_begin, base = span / n, extra = span % n;
  // every device: zero its accumulator, render its contiguous share of the sample range, hand the sums to the exchange step
  int rc = group_parallel(g, [&](int i) -> int {
    CU(cudaSetDevice(g->devices[i]));
    cudaStream_t s = g->streams[i];
    CU(cudaMemsetAsync(g->accum[i], 0, elems * sizeof(double), s));
    const int lo = spp_begin + i * base + std::min(i, extra), hi = lo + base + (i < extra ? 1 : 0);
    std::memset(&st[i], 0, sizeof st[i]);
    if (hi > lo) {
      const int r = b200rt_render_device(g->ctx[i], seed, lo, hi, g->accum[i], s, stats ? &st[i] : nullptr);
      if (r != B200RT_OK) return r;
    }
    if (n > 1 && g->reduce_mode == B200RT_GROUP_REDUCE_NCCL) {
      const ncclResult_t r = g_nccl.Reduce(g->accum[i], g->accum[i], elems, ncclDouble, ncclSum, 0, g->comms[i], s);
      if (r != ncclSuccess) return fail(B200RT_ERR_CUDA, "ncclReduce: %s", g_nccl.GetErrorString(r));
    }
    CU(cudaEventRecord(g->done[i], s));
    return B200RT_OK;
  });
  if (rc != B200RT_OK) return rc;
  CU(cudaSetDevice(g->devices[0]));
  cudaStream_t s0 = g->streams[0];
  if (n > 1 && g->reduce_mode == B200RT_GROUP_REDUCE_PEER) {
    PeerSrc src{};
    for (int i = 1; i < n; i++) {
      CU(cudaStreamWaitEvent(s0, g->done[i], 0));
      src.p[i - 1] = g->accum[i];
    }
    const int grid = (int)std::min<size_t>((elems / 2 + 255) / 256 + 1, (size_t)g->ctx[0]->sm_count * 8);
    k_reduce_peers<<<grid, 256, 0, s0>>>(g->accum[0], src, n - 1, elems);
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(rgb_sum, g->accum[0], elems * sizeof(double), cudaMemcpyDeviceToHost, s0));
  CU(cudaStreamSynchronize(s0));
  for (int i = 1; i < n; i++) { // the other devices' streams are idle by now (NCCL) or were waited for (peer); make it explicit
    CU(cudaSetDevice(g->devices[i]));
    CU(cudaStreamSynchronize(g->streams[i]));
  }
  if (stats) {
    std::memset(stats, 0, sizeof *stats);
    for (int i = 0; i < n; i++) {
      stats->samples += st[i].samples, stats->rays += st[i].rays, stats->nan_samples += st[i].nan_samples, stats->kernel_launches += st[i].kernel_launches;
      stats->gpu_ms = std::max(stats->gpu_ms, st[i].gpu_ms);
      stats->extend_ms = std::max(stats->extend_ms, st[i].extend_ms);
      stats->waves = std::max(stats->waves, st[i].waves);
      stats->h2d_bytes += g->ctx[i]->scene_h2d_bytes + sizeof(b200rt_camera);
    }
    stats->d2h_bytes = (double)(elems * sizeof(double));
  }
  return B200RT_OK;
}

const char *b200rt_group_reduce_backend(void) {
  static thread_local char buf[96];
  int v = 0;
  if (g_nccl.handle && g_nccl.GetVersion(&v) == ncclSuccess)
    snprintf(buf, sizeof buf, "NCCL %d.%d.%d (dlopen)", v / 10000, (v / 100) % 100, v % 100);
  else
    snprintf(buf, sizeof buf, "NCCL not loaded");
  return buf;
}

} // extern "C"
