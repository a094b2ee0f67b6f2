// helper/processing/nms.py:4-38 -- nms(dets, thresh), single and batched (per-class) forms.
#include <string.h>

#include <atomic>

#include "nms_cluster.cuh"
#include "nms_fused.cuh"
#include "nms_host.cuh"
#include "sort.cuh"

using namespace mxr;

namespace {

const int kSmemSortMax = 16384;

template <typename T>
struct PrepArrays {
  typename Ar<T>::Box* box;
  T* area;
  int* orig;
  int cap;
  u64* g_sort_keys;
  u32* g_sort_pl;
  int g_sort_cap;
};

template <typename T>
__device__ __forceinline__ u64 det_key(T s);
template <>
__device__ __forceinline__ u64 det_key<float>(float s) { return (u64)score_key32(s) << 32; }
template <>
__device__ __forceinline__ u64 det_key<double>(double s) { return score_key64(s); }

// One CTA per problem: order = scores.argsort()[::-1] (nms.py:19; ties: higher index first),
// then boxes / areas (nms.py:18) are gathered into processing order.
template <typename T>
__global__ void __launch_bounds__(1024) nms_prepare_kernel(const T* __restrict__ dets, const int* __restrict__ offsets,
                                                           PrepArrays<T> pa, int n_pow2_max, int use_smem_sort,
                                                           Ctrl* __restrict__ ctrl) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int row0 = offsets[b];
  const int n = offsets[b + 1] - row0;
  if (tid == 0) ctrl[b].n = n;
  if (n <= 0) return;
  const T* d = dets + (size_t)row0 * 5;
  const int n_pow2 = max(2, (int)(1u << (32 - __clz(n - 1))));
  u64* skeys;
  u32* spl;
  if (use_smem_sort) {
    skeys = reinterpret_cast<u64*>(smem_raw);
    spl = reinterpret_cast<u32*>(smem_raw + (size_t)n_pow2_max * 8);
  } else {
    skeys = pa.g_sort_keys + (size_t)b * pa.g_sort_cap;
    spl = pa.g_sort_pl + (size_t)b * pa.g_sort_cap;
  }
  bool nan = false;
  for (int i = tid; i < n_pow2; i += blockDim.x) {
    if (i < n) {
      const T s = d[(size_t)i * 5 + 4];
      nan = nan || (s != s);
      skeys[i] = det_key<T>(s);
      spl[i] = (u32)i + 1u;  // +1 keeps every real pair above the (0,0) padding
    } else {
      skeys[i] = 0;
      spl[i] = 0;
    }
  }
  __syncthreads();
  bitonic_sort_desc(skeys, spl, n_pow2);
  for (int i = tid; i < n; i += blockDim.x) {
    const int src = (int)spl[i] - 1;
    const T x1 = d[(size_t)src * 5], y1 = d[(size_t)src * 5 + 1], x2 = d[(size_t)src * 5 + 2], y2 = d[(size_t)src * 5 + 3];
    nan = nan || (x1 != x1) || (y1 != y1) || (x2 != x2) || (y2 != y2);
    const size_t o = (size_t)b * pa.cap + i;
    typename Ar<T>::Box bx;
    bx.x = x1;
    bx.y = y1;
    bx.z = x2;
    bx.w = y2;
    pa.box[o] = bx;
    pa.area[o] = box_area<T>(x1, y1, x2, y2);
    pa.orig[o] = src;
  }
  if (__syncthreads_or(nan) && tid == 0) ctrl[b].nan_flag = 1;
}

struct NmsLayout {
  int P, n_max, cap, keep_cap, n_pow2_max, use_smem_sort;
  StripePlan plan;
  Ctrl* ctrl;
  u64* R;
  size_t zero_bytes;
  int* offsets;
  void* box;
  void* area;
  int* orig;
  int* kept_slot;
  float4* kbox;  // fused kernel: kept boxes as the float32 screen sees them
  float4* kaux;
  u64* M;
  u64* g_sort_keys;
  u32* g_sort_pl;
  // host-io staging
  void* d_dets;
  int* d_keep;
  int* d_nkeep;
  size_t total;
};

void make_layout(int P, int n_max, int n_total, int max_keep, size_t esize, bool host_io, char* base, NmsLayout* L,
                 bool fused = true) {
  L->P = P;
  L->n_max = n_max;
  L->cap = round_up_int(n_max < 1 ? 1 : n_max, 64);
  L->keep_cap = (max_keep > 0 && max_keep < n_max) ? max_keep : (n_max < 1 ? 1 : n_max);
  L->n_pow2_max = pow2_ceil(n_max < 2 ? 2 : n_max);
  L->use_smem_sort = L->n_pow2_max <= kSmemSortMax;
  L->plan = plan_stripes(n_max, max_keep > 0 ? L->keep_cap : 0, 0);
  Carver c(base);
  L->ctrl = c.take<Ctrl>(P);
  L->R = c.take<u64>((size_t)P * (L->cap / 64));
  L->zero_bytes = c.off;
  L->offsets = c.take<int>(P + 1);
  L->box = c.take<char>((size_t)P * L->cap * 4 * esize);
  L->area = c.take<char>((size_t)P * L->cap * esize);
  L->orig = c.take<int>((size_t)P * L->cap);
  L->kept_slot = c.take<int>((size_t)P * L->keep_cap);
  L->kbox = fused ? c.take<float4>((size_t)P * L->keep_cap) : nullptr;  // mxr_nms_noalloc runs the striped kernels
  L->kaux = fused ? c.take<float4>((size_t)P * L->keep_cap) : nullptr;
  L->M = c.take<u64>((size_t)P * L->plan.max_width * (L->plan.max_width / 64));
  if (!L->use_smem_sort) {
    L->g_sort_keys = c.take<u64>((size_t)P * L->n_pow2_max);
    L->g_sort_pl = c.take<u32>((size_t)P * L->n_pow2_max);
  } else {
    L->g_sort_keys = nullptr;
    L->g_sort_pl = nullptr;
  }
  if (host_io) {
    L->d_dets = c.take<char>((size_t)(n_total < 1 ? 1 : n_total) * 5 * esize);
    L->d_keep = c.take<int>(n_total < 1 ? 1 : n_total);
    L->d_nkeep = c.take<int>((size_t)P * 2);
  }
  L->total = align_up(c.off, 256);
}

template <typename T>
int enqueue_t(Handle* h, cudaStream_t s, const NmsLayout& L, const T* dets, double thresh, int max_keep, int32_t* keep_out,
              int32_t* nkeep_out) {
  PrepArrays<T> pa;
  pa.box = reinterpret_cast<typename Ar<T>::Box*>(L.box);
  pa.area = reinterpret_cast<T*>(L.area);
  pa.orig = L.orig;
  pa.cap = L.cap;
  pa.g_sort_keys = L.g_sort_keys;
  pa.g_sort_pl = L.g_sort_pl;
  pa.g_sort_cap = L.n_pow2_max;
  size_t smem = L.use_smem_sort ? (size_t)L.n_pow2_max * 12 : 16;
  const int slot = sizeof(T) == 8 ? 2 : 1;
  MXR_TRY(raise_dyn_smem(h, slot, (const void*)nms_prepare_kernel<T>, smem, nullptr));
  prof_begin(h, s, ST_NMS_PREP);
  nms_prepare_kernel<T><<<L.P, 1024, smem, s>>>(dets, L.offsets, pa, L.n_pow2_max, L.use_smem_sort, L.ctrl);
  prof_end(h, s);
  NmsArrays<T> na;
  na.box = pa.box;
  na.area = pa.area;
  na.score = nullptr;
  na.orig = L.orig;
  na.cap = L.cap;
  na.M = L.M;
  na.m_pitch = L.plan.max_width / 64;
  na.m_stride = (size_t)L.plan.max_width * na.m_pitch;
  na.R = L.R;
  na.r_words = L.cap / 64;
  na.kept_slot = L.kept_slot;
  na.keep_cap = L.keep_cap;
  NmsOutputs o;
  o.rois = nullptr;
  o.scores = nullptr;
  o.keep_pos = nullptr;
  o.keep_orig = keep_out;
  o.nkeep = nkeep_out;
  o.rows = 0;
  o.row_offsets = L.offsets;
  // numpy compares `ovr <= thresh` in the dtype of dets: a python float is cast to f32
  // for f32 dets (NEP 50 weak scalar; value-based casting before numpy 2 agrees)
  return run_nms_stripes<T>(h, s, L.ctrl, na, o, L.P, L.n_max, max_keep > 0 ? L.keep_cap : 0, L.plan, (T)thresh);
}

// One launch of nms_cluster_kernel: a cluster of 2, 4 or 8 CTAs per problem.
template <typename T>
int enqueue_cluster_t(Handle* h, cudaStream_t s, const NmsLayout& L, const T* dets, double thresh, int max_keep,
                      int32_t* keep_out, int32_t* nkeep_out) {
  NmsClusterArgs<T> g;
  g.dets = dets;
  g.offsets = L.offsets;
  g.a.box = reinterpret_cast<typename Ar<T>::Box*>(L.box);
  g.a.area = reinterpret_cast<T*>(L.area);
  g.a.score = nullptr;
  g.a.orig = L.orig;
  g.a.cap = L.cap;
  g.a.M = nullptr;
  g.a.m_pitch = 0;
  g.a.m_stride = 0;
  g.a.R = nullptr;
  g.a.r_words = 0;
  g.a.kept_slot = L.kept_slot;
  g.a.keep_cap = L.keep_cap;
  g.o.rois = nullptr;
  g.o.scores = nullptr;
  g.o.keep_pos = nullptr;
  g.o.keep_orig = keep_out;
  g.o.nkeep = nkeep_out;
  g.o.rows = 0;
  g.o.row_offsets = L.offsets;
  g.ctrl = L.ctrl;
  g.thr = (T)thresh;  // compared in the dtype of dets, like numpy (see enqueue_t)
  g.max_keep = max_keep > 0 ? L.keep_cap : 0;
  g.sort_cap = L.n_pow2_max < 1024 ? 1024 : L.n_pow2_max;
  const size_t smem = nms_cluster_smem_bytes<T>(g.sort_cap);
  const int slot = sizeof(T) == 8 ? 7 : 6;
  MXR_TRY(raise_dyn_smem(h, slot, (const void*)nms_cluster_kernel<T>, smem, nullptr));
  const int csize = L.n_max > 1024 ? 8 : (L.n_max > 256 ? 4 : 2);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(L.P * csize);
  cfg.blockDim = dim3(1024);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  prof_begin(h, s, ST_NMS_CLUSTER);
  MXR_CUDA(cudaLaunchKernelEx(&cfg, nms_cluster_kernel<T>, g));
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

// One launch of fz::nms_fused_kernel (nms_fused.cuh): a cluster per problem, the stripe engine of the fused proposal kernel.
template <typename T>
int enqueue_fused_t(Handle* h, cudaStream_t s, const NmsLayout& L, const T* dets, double thresh, int max_keep,
                    int32_t* keep_out, int32_t* nkeep_out) {
  fz::NmsFusedArgs<T> g;
  g.dets = dets;
  g.offsets = L.offsets;
  g.box = reinterpret_cast<typename Ar<T>::Box*>(L.box);
  g.area = reinterpret_cast<T*>(L.area);
  g.orig = L.orig;
  g.cap = L.cap;
  g.kbox = L.kbox;
  g.kaux = L.kaux;
  g.keep_cap = L.keep_cap;
  g.o.rois = nullptr;
  g.o.scores = nullptr;
  g.o.keep_pos = nullptr;
  g.o.keep_orig = keep_out;
  g.o.nkeep = nkeep_out;
  g.o.rows = 0;
  g.o.row_offsets = L.offsets;
  g.ctrl = L.ctrl;
  g.thr = (T)thresh;  // compared in the dtype of dets, like numpy (see enqueue_t)
  g.max_keep = max_keep > 0 ? L.keep_cap : 0;
  g.sort_cap = L.n_pow2_max < 1024 ? 1024 : L.n_pow2_max;
  const size_t smem = fz::nms_fused_smem_bytes<T>(g.sort_cap);
  const int slot = sizeof(T) == 8 ? 11 : 10;
  MXR_TRY(raise_dyn_smem(h, slot, (const void*)fz::nms_fused_kernel<T>, smem, nullptr));
  // Blocks of 32 x 32 pairs per stripe: up to 136 inside the stripe plus (kept / 32) x 16 against the kept list; the
  // cluster shares them, the per-stripe resolve does not shrink with it.  Many problems: keep the whole grid resident.
  // Measured (tools/nms_probe.py): one list of 2000 / 6000: 16 CTAs 61 / 239 us, 8 CTAs 68 / 286 us; eight lists of 6000:
  // 8 CTAs 250 us, 16 CTAs 385 us (not all clusters resident); twenty lists of 300: 4 CTAs 28.6 us, 8 CTAs 39 us.
  // eighty lists of 100: one CTA per list 24.7 us, two 32.5 us (ten blocks of pairs: nothing to share, no barrier to pay).
  int csize = L.n_max > 1024 ? (L.P <= 4 ? 16 : 8) : (L.n_max > 256 ? 4 : (L.n_max > 128 ? 2 : 1));
  while (csize > 2 && (long long)L.P * csize > 148) csize >>= 1;  // (never below two for the longer lists)
  if (h->cluster_override >= 2 && h->cluster_override <= 16) csize = h->cluster_override;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(csize, L.P);
  cfg.blockDim = dim3(1024);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (csize > 8)
    MXR_CUDA(cudaFuncSetAttribute((const void*)fz::nms_fused_kernel<T>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  prof_begin(h, s, ST_NMS_FUSED);
  cudaError_t le = cudaLaunchKernelEx(&cfg, fz::nms_fused_kernel<T>, g);
  if (le != cudaSuccess && csize > 8) {
    // a device (partition) that cannot host a 16-CTA cluster: the portable size does the same work
    (void)cudaGetLastError();
    cfg.gridDim = dim3(8, L.P);
    attr[0].val.clusterDim.x = 8;
    le = cudaLaunchKernelEx(&cfg, fz::nms_fused_kernel<T>, g);
  }
  MXR_CUDA(le);
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

// offsets: HOST array of P+1 ints.
int run(Handle* h, cudaStream_t s, const void* dets, bool dets_on_host, const int32_t* offsets, int P, int dtype,
        double thresh, int max_keep, int32_t* keep_out, int32_t* nkeep_out) {
  MXR_REQUIRE(dtype == MXR_DTYPE_F32 || dtype == MXR_DTYPE_F64, MXR_ERR_INVALID, "nms: dtype must be f32 or f64");
  MXR_REQUIRE(P >= 1 && offsets != nullptr && nkeep_out != nullptr, MXR_ERR_INVALID, "nms: bad problem list");
  int n_max = 0;
  for (int i = 0; i < P; ++i) {
    MXR_REQUIRE(offsets[i + 1] >= offsets[i], MXR_ERR_INVALID, "nms: offsets must be non-decreasing");
    if (offsets[i + 1] - offsets[i] > n_max) n_max = offsets[i + 1] - offsets[i];
  }
  const int n_total = offsets[P] - offsets[0];
  MXR_REQUIRE(offsets[0] == 0, MXR_ERR_INVALID, "nms: offsets[0] must be 0");
  MXR_REQUIRE(n_total == 0 || (dets != nullptr && keep_out != nullptr), MXR_ERR_INVALID, "nms: NULL dets/keep");
  MXR_REQUIRE(n_max <= (1 << 24), MXR_ERR_LIMIT, "nms: %d boxes in one problem exceeds the maximum of 16777216", n_max);
  const size_t esize = dtype == MXR_DTYPE_F64 ? 8 : 4;
  NmsLayout L;
  make_layout(P, n_max, n_total, max_keep, esize, dets_on_host, nullptr, &L);
  MXR_TRY(ws_order(h, s));
  MXR_TRY(ensure_device_ws(h, L.total));
  make_layout(P, n_max, n_total, max_keep, esize, dets_on_host, h->d_ws, &L);
  const size_t off_bytes = (size_t)(P + 1) * sizeof(int);
  const size_t pin_need = off_bytes + (dets_on_host ? ((size_t)n_total + 2 * (size_t)P) * sizeof(int) : 0);
  MXR_TRY(ensure_pinned_ws(h, pin_need));
  MXR_CUDA(cudaStreamSynchronize(s));  // the pinned block may still feed an earlier copy
  memcpy(h->h_pin, offsets, off_bytes);
  h->last_launches = 0;
  MXR_CUDA(cudaMemsetAsync(L.ctrl, 0, L.zero_bytes, s));
  MXR_CUDA(cudaMemcpyAsync(L.offsets, h->h_pin, off_bytes, cudaMemcpyHostToDevice, s));
  const void* d_dets = dets;
  int32_t* d_keep = keep_out;
  int32_t* d_nkeep = nkeep_out;
  if (dets_on_host) {
    if (n_total > 0) MXR_CUDA(cudaMemcpyAsync(L.d_dets, dets, (size_t)n_total * 5 * esize, cudaMemcpyHostToDevice, s));
    d_dets = L.d_dets;
    d_keep = L.d_keep;
    d_nkeep = L.d_nkeep;
  }
  // one cluster launch when a problem's sort fits the leader's shared memory, else the striped kernels
  const bool use_cluster = !h->no_cluster && n_max <= (dtype == MXR_DTYPE_F64 ? 8192 : 16384);
  const bool use_fused = use_cluster && !h->use_v1 && P <= 65535;
  if (use_fused) {
    if (dtype == MXR_DTYPE_F64)
      MXR_TRY(enqueue_fused_t<double>(h, s, L, reinterpret_cast<const double*>(d_dets), thresh, max_keep, d_keep, d_nkeep));
    else
      MXR_TRY(enqueue_fused_t<float>(h, s, L, reinterpret_cast<const float*>(d_dets), thresh, max_keep, d_keep, d_nkeep));
  } else if (dtype == MXR_DTYPE_F64) {
    if (use_cluster)
      MXR_TRY(enqueue_cluster_t<double>(h, s, L, reinterpret_cast<const double*>(d_dets), thresh, max_keep, d_keep, d_nkeep));
    else
      MXR_TRY(enqueue_t<double>(h, s, L, reinterpret_cast<const double*>(d_dets), thresh, max_keep, d_keep, d_nkeep));
  } else {
    if (use_cluster)
      MXR_TRY(enqueue_cluster_t<float>(h, s, L, reinterpret_cast<const float*>(d_dets), thresh, max_keep, d_keep, d_nkeep));
    else
      MXR_TRY(enqueue_t<float>(h, s, L, reinterpret_cast<const float*>(d_dets), thresh, max_keep, d_keep, d_nkeep));
  }
  if (dets_on_host) {
    char* hp = h->h_pin + off_bytes;
    if (n_total > 0) MXR_CUDA(cudaMemcpyAsync(hp, d_keep, (size_t)n_total * sizeof(int), cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpyAsync(hp + (size_t)n_total * sizeof(int), d_nkeep, (size_t)P * 2 * sizeof(int),
                             cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaStreamSynchronize(s));
    if (n_total > 0) memcpy(keep_out, hp, (size_t)n_total * sizeof(int));
    memcpy(nkeep_out, hp + (size_t)n_total * sizeof(int), (size_t)P * 2 * sizeof(int));
    for (int i = 0; i < P; ++i) {
      if (nkeep_out[i * 2 + 1] & 1) {
        set_error("nms: NaN in dets (problem %d); the reference's result for NaN input is not defined", i);
        return MXR_ERR_NAN;
      }
    }
  }
  return MXR_OK;
}

}  // namespace

extern "C" {

int mxr_nms(mxr_handle_t hh, void* stream, const void* dets, int n, int dtype, double thresh, int max_keep,
            int32_t* keep_out, int32_t* nkeep_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_nms: NULL handle");
  MXR_REQUIRE(n >= 0, MXR_ERR_INVALID, "mxr_nms: n < 0");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const int32_t offs[2] = {0, n};
  return run(h, (cudaStream_t)stream, dets, false, offs, 1, dtype, thresh, max_keep, keep_out,
             nkeep_out);
}

int mxr_nms_host(mxr_handle_t hh, const void* dets, int n, int dtype, double thresh, int max_keep, int32_t* keep_out,
                 int32_t* nkeep_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_nms_host: NULL handle");
  MXR_REQUIRE(n >= 0, MXR_ERR_INVALID, "mxr_nms_host: n < 0");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const int32_t offs[2] = {0, n};
  return run(h, h->stream, dets, true, offs, 1, dtype, thresh, max_keep, keep_out, nkeep_out);
}

int mxr_nms_batched(mxr_handle_t hh, void* stream, const void* dets, const int32_t* offsets, int n_problems, int dtype,
                    double thresh, int max_keep, int32_t* keep_out, int32_t* nkeep_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_nms_batched: NULL handle");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  return run(h, (cudaStream_t)stream, dets, false, offsets, n_problems, dtype, thresh, max_keep,
             keep_out, nkeep_out);
}

int mxr_nms_batched_host(mxr_handle_t hh, const void* dets, const int32_t* offsets, int n_problems, int dtype, double thresh,
                         int max_keep, int32_t* keep_out, int32_t* nkeep_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_nms_batched_host: NULL handle");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  return run(h, h->stream, dets, true, offsets, n_problems, dtype, thresh, max_keep, keep_out, nkeep_out);
}

}  // extern "C"

// =========================================================================================
// mxr_nms_noalloc -- nms(dets, thresh) for callers that must not allocate device memory.
//
// Who needs this: a plug-in that runs inside a host framework which has already handed the
// whole GPU to its own memory pool (a cudaMalloc from the plug-in then fails, or steals from
// and fragments that pool), and a serving loop that cannot afford cudaMalloc/cudaFree on its
// request path because both synchronise the whole device and stall every other stream.
// mxr_nms() keeps its scratch in the handle's workspace, which it obtains with cudaMalloc the
// first time (and whenever a larger problem arrives).  This variant takes no handle and never
// calls cudaMalloc, cudaFree, cudaHostAlloc or any other allocating routine: its scratch is
// the fixed array below, which is part of the library image -- every GPU that loads the
// library holds these 640 KiB until the process exits.
//
// Limits and rules (also in include/mxrcnn_b200.h):
//   * at most MXR_NMS_NOALLOC_MAX_BOXES (2048) boxes per call; larger n -> MXR_ERR_LIMIT.
//   * the first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress, once per device, with that device current).  Make that first
//     call during set-up, before any stream capture begins.
//   * all calls on one device share the one array, and it keeps whatever the previous call
//     left in it.  Each call therefore clears the words that must start at zero (control
//     block, suppression bit words) in its first kernel.  Two calls whose GPU work overlaps
//     (different streams, or different host threads) race on the array: serialise them.
// =========================================================================================
namespace {

const int kNoAllocMaxBoxes = 2048;
const size_t kNoAllocBytes = 640 * 1024;
const int kNoAllocMaxDevices = 64;

// Scratch of mxr_nms_noalloc.  In HOST code this name is only a handle for
// cudaGetSymbolAddress -- it is not a device address and must never be passed to a kernel.
static __device__ __align__(256) unsigned char g_nms_noalloc_scratch[kNoAllocBytes];

// Device address of g_nms_noalloc_scratch on each device, filled on first use (any host thread may be the first).
std::atomic<unsigned char*> g_nms_noalloc_addr[kNoAllocMaxDevices];

// n == 0: nothing to do but report it (no kernel of the pipeline is given a null pointer).
__global__ void nms_noalloc_empty_kernel(int* nkeep) {
  nkeep[0] = 0;
  nkeep[1] = 0;
}

// First kernel of every call: the array still holds the previous call's data.
__global__ void nms_noalloc_reset_kernel(unsigned char* zero_from, size_t zero_bytes, int* offsets, int n) {
  const size_t words = zero_bytes / sizeof(u32);
  u32* z = reinterpret_cast<u32*>(zero_from);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x) z[i] = 0u;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    offsets[0] = 0;
    offsets[1] = n;
  }
}

int nms_noalloc_scratch(unsigned char** out, int* device_out) {
  int dev = -1;
  MXR_CUDA(cudaGetDevice(&dev));
  MXR_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, MXR_ERR_LIMIT, "mxr_nms_noalloc: device index %d not supported", dev);
  unsigned char* known = g_nms_noalloc_addr[dev].load(std::memory_order_acquire);
  if (known == nullptr) {
    // once per device, with that device current (it is: we just asked which one is)
    unsigned char* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_nms_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("mxr_nms_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return MXR_ERR_CUDA;
    }
    g_nms_noalloc_addr[dev].store(ptr, std::memory_order_release);  // every thread computes the same address
    known = ptr;
  }
  *out = known;
  *device_out = dev;
  return MXR_OK;
}

}  // namespace

extern "C" {

int mxr_nms_noalloc(void* stream, const void* dets, int n, int dtype, double thresh, int max_keep, int32_t* keep_out,
                    int32_t* nkeep_out) {
  MXR_REQUIRE(dtype == MXR_DTYPE_F32 || dtype == MXR_DTYPE_F64, MXR_ERR_INVALID, "mxr_nms_noalloc: dtype must be f32 or f64");
  MXR_REQUIRE(n >= 0 && nkeep_out != nullptr, MXR_ERR_INVALID, "mxr_nms_noalloc: bad n or NULL nkeep_out");
  MXR_REQUIRE(n == 0 || (dets != nullptr && keep_out != nullptr), MXR_ERR_INVALID, "mxr_nms_noalloc: NULL dets/keep_out");
  MXR_REQUIRE(n <= kNoAllocMaxBoxes, MXR_ERR_LIMIT,
              "mxr_nms_noalloc: %d boxes exceed the fixed scratch (maximum %d per call); use mxr_nms", n,
              kNoAllocMaxBoxes);
  unsigned char* scratch = nullptr;
  int dev = 0;
  MXR_TRY(nms_noalloc_scratch(&scratch, &dev));
  if (n == 0) {
    nms_noalloc_empty_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(nkeep_out);
    MXR_CUDA(cudaGetLastError());
    return MXR_OK;
  }
  const size_t esize = dtype == MXR_DTYPE_F64 ? 8 : 4;
  NmsLayout L;
  make_layout(1, n, n, max_keep, esize, false, reinterpret_cast<char*>(scratch), &L, false);
  MXR_REQUIRE(L.total <= kNoAllocBytes, MXR_ERR_LIMIT, "mxr_nms_noalloc: layout of %zu bytes exceeds the %zu-byte scratch",
              L.total, kNoAllocBytes);
  cudaStream_t s = (cudaStream_t)stream;
  Handle tmp;  // plain host struct for the shared launch helpers: no workspace, profiling off
  memset(&tmp, 0, sizeof(tmp));
  tmp.device = dev;
  int no_opt_in[12] = {0, 1 << 30, 1 << 30, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // <= 24 KiB of dynamic smem here: no opt-in needed
  tmp.smem_attr = no_opt_in;
  nms_noalloc_reset_kernel<<<4, 256, 0, s>>>(reinterpret_cast<unsigned char*>(L.ctrl), L.zero_bytes, L.offsets, n);
  MXR_CUDA(cudaGetLastError());
  if (dtype == MXR_DTYPE_F64) {
    MXR_TRY(enqueue_t<double>(&tmp, s, L, reinterpret_cast<const double*>(dets), thresh, max_keep, keep_out, nkeep_out));
  } else {
    MXR_TRY(enqueue_t<float>(&tmp, s, L, reinterpret_cast<const float*>(dets), thresh, max_keep, keep_out, nkeep_out));
  }
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

}  // extern "C"
