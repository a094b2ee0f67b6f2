// C ABI of libmxrcnn_b200.so: handle / workspace management and error reporting.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace mxr {

static thread_local char g_err[768] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int ensure_device_ws(Handle* h, size_t bytes) {
  if (bytes <= h->d_bytes) return MXR_OK;
  // growth: nothing may still be using the old block
  MXR_CUDA(cudaDeviceSynchronize());
  if (h->d_ws) MXR_CUDA(cudaFree(h->d_ws));
  h->d_ws = nullptr;
  h->d_bytes = 0;
  const size_t want = align_up(bytes + bytes / 8, 1 << 20);
  cudaError_t e = cudaMalloc(&h->d_ws, want);
  if (e != cudaSuccess) {
    set_error("cudaMalloc(%zu) for the workspace failed: %s", want, cudaGetErrorString(e));
    return MXR_ERR_NOMEM;
  }
  h->d_bytes = want;
  return MXR_OK;
}

int ensure_pinned_ws(Handle* h, size_t bytes) {
  if (bytes <= h->h_bytes) return MXR_OK;
  MXR_CUDA(cudaDeviceSynchronize());
  if (h->h_pin) MXR_CUDA(cudaFreeHost(h->h_pin));
  h->h_pin = nullptr;
  h->h_bytes = 0;
  const size_t want = align_up(bytes + bytes / 8, 1 << 16);
  cudaError_t e = cudaHostAlloc(&h->h_pin, want, cudaHostAllocDefault);
  if (e != cudaSuccess) {
    set_error("cudaHostAlloc(%zu) failed: %s", want, cudaGetErrorString(e));
    return MXR_ERR_NOMEM;
  }
  h->h_bytes = want;
  return MXR_OK;
}

int ws_order(Handle* h, cudaStream_t s) {
  const bool s_private = (s == h->stream) || (h->stream2 && s == h->stream2);
  if (!h->ws_used) {
    h->ws_used = 1;
  } else if (h->ws_last_private && s_private) {
    // the handle's own streams: the two slots of the pipelined host path use disjoint halves, everything else runs on
    // h->stream in order
  } else if (!h->ws_last_private && h->ws_last == s) {
    // same caller stream as last time: stream order
  } else if (h->ws_last_private) {
    if (cudaEventRecord(h->ev, h->stream) == cudaSuccess) cudaStreamWaitEvent(s, h->ev, 0);
    if (h->stream2 && cudaEventRecord(h->ev2, h->stream2) == cudaSuccess) cudaStreamWaitEvent(s, h->ev2, 0);
    (void)cudaGetLastError();
  } else {
    // the previous caller stream may have been destroyed since (its work then completes on its own): errors are dropped
    if (cudaEventRecord(h->ev, h->ws_last) == cudaSuccess) cudaStreamWaitEvent(s, h->ev, 0);
    (void)cudaGetLastError();
  }
  h->ws_last = s;
  h->ws_last_private = s_private ? 1 : 0;
  return MXR_OK;
}

// Kernel attributes and the occupancy figures derived from them, per device, shared by all handles.
static const int kMaxDevices = 64;
struct DeviceCache {
  int smem_attr[12];
  int coop_grid[4];
  int cluster_cap[32];
};
static DeviceCache g_dev_cache[kMaxDevices];
static std::mutex g_dev_mutex;

int raise_dyn_smem(Handle* h, int slot, const void* func, size_t bytes, int* raised) {
  std::lock_guard<std::mutex> lock(g_dev_mutex);
  if (raised) *raised = 0;
  if ((int)bytes <= h->smem_attr[slot]) return MXR_OK;
  MXR_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  h->smem_attr[slot] = (int)bytes;
  if (raised) *raised = 1;
  return MXR_OK;
}

}  // namespace mxr

using namespace mxr;

extern "C" {

const char* mxr_last_error_string(void) { return g_err; }

int mxr_version(void) { return 100; }

int mxr_create(int device, mxr_handle_t* out) {
  MXR_REQUIRE(out != nullptr, MXR_ERR_INVALID, "mxr_create: out is NULL");
  int count = 0;
  MXR_CUDA(cudaGetDeviceCount(&count));
  MXR_REQUIRE(device >= 0 && device < count && device < kMaxDevices, MXR_ERR_INVALID, "mxr_create: device %d of %d", device,
              count);
  MXR_CUDA(cudaSetDevice(device));
  Handle* h = new Handle();
  memset(h, 0, sizeof(*h));
  h->device = device;
  h->smem_attr = g_dev_cache[device].smem_attr;
  h->coop_grid = g_dev_cache[device].coop_grid;
  h->cluster_cap = g_dev_cache[device].cluster_cap;
  MXR_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  MXR_CUDA(cudaEventCreateWithFlags(&h->ev, cudaEventDisableTiming));
  MXR_CUDA(cudaEventCreateWithFlags(&h->ev2, cudaEventDisableTiming));
  const char* nc = getenv("MXRCNN_B200_NO_CLUSTER");
  h->no_cluster = (nc && nc[0] == '1') ? 1 : 0;
  const char* c8 = getenv("MXRCNN_B200_CLUSTER8");
  h->cluster_override = (c8 && c8[0] == '1') ? 8 : 0;
  const char* np = getenv("MXRCNN_B200_NO_PDL");
  h->no_pdl = (np && np[0] == '1') ? 1 : 0;
  const char* v1 = getenv("MXRCNN_B200_V1");
  h->use_v1 = (v1 && v1[0] == '1') ? 1 : 0;
  const char* cs = getenv("MXRCNN_B200_CLUSTER");
  if (cs) {
    const int v = atoi(cs);
    if (v >= 2 && v <= 16 && v % 2 == 0) h->cluster_override = v;
  }
  *out = reinterpret_cast<mxr_handle_t>(h);
  return MXR_OK;
}

int mxr_destroy(mxr_handle_t hh) {
  if (!hh) return MXR_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  if (h->d_ws) cudaFree(h->d_ws);
  if (h->h_pin) cudaFreeHost(h->h_pin);
  if (h->at_state) free(h->at_state);
  if (h->pt_state) free(h->pt_state);
  if (h->prop_state) free(h->prop_state);
  if (h->stream2) cudaStreamDestroy(h->stream2);
  if (h->ev) cudaEventDestroy(h->ev);
  if (h->ev2) cudaEventDestroy(h->ev2);
  if (h->prof) {
    for (int i = 0; i < h->prof_cap; ++i) {
      cudaEventDestroy(h->prof[i].a);
      cudaEventDestroy(h->prof[i].b);
    }
    delete[] h->prof;
  }
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return MXR_OK;
}

size_t mxr_workspace_bytes(mxr_handle_t hh) { return hh ? reinterpret_cast<Handle*>(hh)->d_bytes : 0; }

int mxr_profile_enable(mxr_handle_t hh, int on) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_profile_enable: NULL handle");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  if (on && !h->prof) {
    h->prof_cap = 1 << 15;
    h->prof = new ProfRec[h->prof_cap];
    for (int i = 0; i < h->prof_cap; ++i) {
      MXR_CUDA(cudaEventCreate(&h->prof[i].a));
      MXR_CUDA(cudaEventCreate(&h->prof[i].b));
    }
  }
  h->prof_used = 0;
  h->prof_on = on ? 1 : 0;
  return MXR_OK;
}

int mxr_profile_read(mxr_handle_t hh, int max_stages, double* total_ms, int32_t* counts) {
  MXR_REQUIRE(hh != nullptr && total_ms && counts && max_stages >= 1, MXR_ERR_INVALID, "mxr_profile_read: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  MXR_CUDA(cudaDeviceSynchronize());
  for (int i = 0; i < max_stages; ++i) {
    total_ms[i] = 0;
    counts[i] = 0;
  }
  for (int i = 0; i < h->prof_used; ++i) {
    float ms = 0.f;
    MXR_CUDA(cudaEventElapsedTime(&ms, h->prof[i].a, h->prof[i].b));
    const int st = h->prof[i].stage;
    if (st >= 0 && st < max_stages) {
      total_ms[st] += ms;
      counts[st] += 1;
    }
  }
  h->prof_used = 0;
  return MXR_OK;
}

const char* mxr_profile_stage_name(int stage) {
  static const char* names[ST_COUNT] = {"memset",          "decode_filter_kernel", "topk_sort_kernel", "nms_stripe_kernel",
                                        "nms_scan_kernel", "nms_prepare_kernel",   "boxes_kernels",    "overlaps_kernels",
                                        "anchor_target_kernels", "proposal_target_kernels", "nms_coop_kernel",
                                        "proposal_cluster_kernel", "nms_cluster_kernel", "box_head_kernel",
                                        "proposal_fused_kernel", "nms_fused_kernel", "anchor_phase2_kernel"};
  return (stage >= 0 && stage < ST_COUNT) ? names[stage] : "";
}

int mxr_last_launch_count(mxr_handle_t hh) { return hh ? reinterpret_cast<Handle*>(hh)->last_launches : 0; }

int mxr_proposal_phase_times(mxr_handle_t hh, int max_images, double* us_out, int32_t* images_out) {
  MXR_REQUIRE(hh != nullptr && us_out != nullptr && images_out != nullptr && max_images >= 1, MXR_ERR_INVALID,
              "mxr_proposal_phase_times: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  *images_out = 0;
  if (!h->dbg_ptr || h->dbg_images <= 0) return MXR_OK;
  MXR_CUDA(cudaDeviceSynchronize());
  const int n = h->dbg_images < max_images ? h->dbg_images : max_images;
  unsigned long long* tmp = new unsigned long long[(size_t)n * 16];
  cudaError_t e = cudaMemcpy(tmp, h->dbg_ptr, (size_t)n * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    delete[] tmp;
    set_error("mxr_proposal_phase_times: %s", cudaGetErrorString(e));
    return MXR_ERR_CUDA;
  }
  for (int b = 0; b < n; ++b) {
    for (int i = 0; i < 8; ++i) us_out[b * 9 + i] = (double)(tmp[b * 16 + i + 1] - tmp[b * 16 + i]) * 1e-3;
    us_out[b * 9 + 8] = (double)tmp[b * 16 + 9];
  }
  delete[] tmp;
  *images_out = n;
  return MXR_OK;
}

int mxr_proposal_stamps(mxr_handle_t hh, int max_images, uint64_t* ns_out, int32_t* images_out) {
  MXR_REQUIRE(hh != nullptr && ns_out != nullptr && images_out != nullptr && max_images >= 1, MXR_ERR_INVALID,
              "mxr_proposal_stamps: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  *images_out = 0;
  if (!h->dbg_ptr || h->dbg_images <= 0) return MXR_OK;
  MXR_CUDA(cudaDeviceSynchronize());
  const int n = h->dbg_images < max_images ? h->dbg_images : max_images;
  // device layout: [images][16] phase stamps, then [images][32] cycle stamps; out: 48 values per image
  uint64_t* tmp = new uint64_t[(size_t)h->dbg_images * 48];
  cudaError_t e = cudaMemcpy(tmp, h->dbg_ptr, (size_t)h->dbg_images * 48 * sizeof(uint64_t), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    delete[] tmp;
    set_error("mxr_proposal_stamps: %s", cudaGetErrorString(e));
    return MXR_ERR_CUDA;
  }
  for (int b = 0; b < n; ++b) {
    memcpy(ns_out + (size_t)b * 48, tmp + (size_t)b * 16, 16 * sizeof(uint64_t));
    memcpy(ns_out + (size_t)b * 48 + 16, tmp + (size_t)h->dbg_images * 16 + (size_t)b * 32, 32 * sizeof(uint64_t));
  }
  delete[] tmp;
  *images_out = n;
  return MXR_OK;
}

int mxr_proposal_count_pairs(mxr_handle_t hh, int on) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_proposal_count_pairs: NULL handle");
  reinterpret_cast<Handle*>(hh)->count_pairs = on ? 1 : 0;
  return MXR_OK;
}

int mxr_proposal_pairs_read(mxr_handle_t hh, int max_images, uint32_t* pairs_out, int32_t* images_out) {
  MXR_REQUIRE(hh != nullptr && pairs_out != nullptr && images_out != nullptr && max_images >= 1, MXR_ERR_INVALID,
              "mxr_proposal_pairs_read: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  *images_out = 0;
  if (!h->pair_ptr || h->pair_images <= 0) return MXR_OK;
  MXR_CUDA(cudaDeviceSynchronize());
  const int n = h->pair_images < max_images ? h->pair_images : max_images;
  MXR_CUDA(cudaMemcpy(pairs_out, h->pair_ptr, (size_t)n * 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  *images_out = n;
  return MXR_OK;
}

int mxr_synchronize(mxr_handle_t hh, void* stream) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_synchronize: NULL handle");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);  // the stream belongs to the handle's device
  MXR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return MXR_OK;
}

// helper/processing/generate_anchor.py:8-72 (host-side constant table; float64 like numpy).
int mxr_generate_anchors(double base_size, const double* ratios, int n_ratios, const double* scales, int n_scales,
                         double* out) {
  MXR_REQUIRE(ratios && scales && out && n_ratios > 0 && n_scales > 0 && n_ratios * n_scales <= 64, MXR_ERR_INVALID,
              "mxr_generate_anchors: need 1..64 anchors (got %d x %d)", n_ratios, n_scales);
  // base window [0, 0, base-1, base-1]   (:14)
  const double bx1 = 0.0, by1 = 0.0, bx2 = base_size - 1.0, by2 = base_size - 1.0;
  const double w = bx2 - bx1 + 1.0, h = by2 - by1 + 1.0;
  const double cx = bx1 + 0.5 * (w - 1.0), cy = by1 + 0.5 * (h - 1.0);
  const double size = w * h;
  for (int r = 0; r < n_ratios; ++r) {
    // ratio enumeration (:49-60); numpy's round is round-half-to-even = nearbyint in the default mode
    const double ws = nearbyint(sqrt(size / ratios[r]));
    const double hs = nearbyint(ws * ratios[r]);
    const double rx1 = cx - 0.5 * (ws - 1.0), ry1 = cy - 0.5 * (hs - 1.0);
    const double rx2 = cx + 0.5 * (ws - 1.0), ry2 = cy + 0.5 * (hs - 1.0);
    const double rw = rx2 - rx1 + 1.0, rh = ry2 - ry1 + 1.0;
    const double rcx = rx1 + 0.5 * (rw - 1.0), rcy = ry1 + 0.5 * (rh - 1.0);
    for (int s = 0; s < n_scales; ++s) {
      // scale enumeration (:63-72)
      const double sw = rw * scales[s], sh = rh * scales[s];
      double* o = out + (size_t)(r * n_scales + s) * 4;
      o[0] = rcx - 0.5 * (sw - 1.0);
      o[1] = rcy - 0.5 * (sh - 1.0);
      o[2] = rcx + 0.5 * (sw - 1.0);
      o[3] = rcy + 0.5 * (sh - 1.0);
    }
  }
  return MXR_OK;
}

}  // extern "C"
