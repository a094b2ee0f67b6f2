// Test-time box head: rcnn/detector.py:73-74 (bbox_pred + clip_boxes on the (R, 4C) deltas) and the class loop
// of rcnn/tester.py:50-65 (score threshold, per-class greedy NMS, cap at max_per_image) as ONE launch:
// one 1024-thread CTA per (image, foreground class); the CTA that finishes an image last applies the cap.
#include "box_kernels.cuh"
#include "common.cuh"
#include "sort.cuh"

namespace mxr {
namespace {

struct BoxHeadArgs {
  const float* rois;    // (B, R, 4)
  const float* deltas;  // (B, R, 4C)
  const float* scores;  // (B, R, C)
  const double* im;     // (B, 3): network-input height, width, scale
  int R, C;
  float score_thr;
  double nms_thr;
  int max_per_image;
  int sort_cap;         // power of two >= R
  double* dets_out;     // (B, C, R, 5)
  int* rows_out;        // (B, C, R)
  int* count_out;       // (B, C, 2): after the cap, after NMS
  unsigned int* done;   // (B) zero before the launch
};

// k-th largest (1-based) of the f32 score keys of every kept detection of one image; whole CTA.
// prefix[c] = first flat index of class c (prefix[C] = total).  Four 8-bit radix passes over global memory.
__device__ u32 kth_largest_key(const BoxHeadArgs& g, int b, const int* prefix, int total, int k, u32* hist, u32* st) {
  const int tid = threadIdx.x;
  u32 pref = 0;
  int k_rem = k;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    for (int f = tid; f < total; f += blockDim.x) {
      int lo = 0, hi = g.C;  // class of flat index f: prefix[lo] <= f < prefix[lo+1]
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (prefix[mid] <= f) lo = mid; else hi = mid;
      }
      const double sc = __ldcg(g.dets_out + (((size_t)b * g.C + lo) * g.R + (f - prefix[lo])) * 5 + 4);
      const u32 key = score_key32((float)sc);
      if (pass == 0 || ((key ^ pref) >> (shift + 8)) == 0u) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {  // descending scan of the 256 bins: 8 bins per lane, then a shuffle scan over the lanes
      u32 mine = 0;
      for (int q = 0; q < 8; ++q) mine += hist[255 - 8 * tid - q];
      u32 incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
        if (tid >= o) incl += v;
      }
      const u32 excl = incl - mine;
      if (excl < (u32)k_rem && (u32)k_rem <= incl) {
        u32 cum = excl;
        for (int q = 0; q < 8; ++q) {
          const int d = 255 - 8 * tid - q;
          const u32 c = hist[d];
          if (cum + c >= (u32)k_rem) {
            st[0] = (u32)d;
            st[1] = (u32)k_rem - cum;
            break;
          }
          cum += c;
        }
      }
    }
    __syncthreads();
    pref |= st[0] << shift;
    k_rem = (int)st[1];
    __syncthreads();
  }
  return pref;
}

__global__ void __launch_bounds__(1024, 1) box_head_kernel(BoxHeadArgs g) {
  extern __shared__ __align__(32) unsigned char dyn_smem[];
  __shared__ u64 s_removed[64];  // one bit per sorted position (<= 4096)
  __shared__ u64 s_M[64];        // lower-triangle rows of the current 64x64 diagonal tile
  __shared__ u64 s_keptbits;
  __shared__ int s_n, s_last;
  __shared__ u32 s_hist[256];
  __shared__ u32 s_st[4];
  typedef Ar<double> A;
  const int tid = threadIdx.x, lane = tid & 31;
  const int j = blockIdx.x + 1, b = blockIdx.y;  // class 0 is the background: skipped (tester.py:50)
  const int R = g.R, C = g.C;
  u64* sortbuf = reinterpret_cast<u64*>(dyn_smem);                         // [sort_cap] score key << 32 | roi row
  double4* sbox = reinterpret_cast<double4*>(sortbuf + g.sort_cap);       // [sort_cap] boxes in processing order
  double* sarea = reinterpret_cast<double*>(sbox + g.sort_cap);           // [sort_cap]
  int* skept = reinterpret_cast<int*>(sarea + g.sort_cap);                // [sort_cap] sorted positions kept so far
  const float* scores = g.scores + (size_t)b * R * C;
  if (tid == 0) s_n = 0;
  if (tid < 64) s_removed[tid] = 0ull;
  __syncthreads();

  // ---- tester.py:51  indexes = where(scores[:, j] > thresh): float32 compare; order is restored by the sort ----
  for (int r0 = 0; r0 < R; r0 += blockDim.x) {
    const int r = r0 + tid;
    float s = 0.f;
    bool sel = false;
    if (r < R) {
      s = __ldg(scores + (size_t)r * C + j);
      sel = s > g.score_thr;
    }
    const u32 m = __ballot_sync(0xffffffffu, sel);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(&s_n, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (sel) sortbuf[base + __popc(m & ((1u << lane) - 1u))] = ((u64)score_key32(s) << 32) | (u32)r;
  }
  __syncthreads();
  const int n = s_n;
  int nk = 0;
  if (n > 0) {
    // ---- nms.py:19  descending score; equal scores: higher row first (argsort(kind='stable')[::-1]) ----
    const int n_pow2 = max(2, (int)(1u << (32 - __clz(max(n - 1, 1)))));
    for (int i = n + tid; i < n_pow2; i += blockDim.x) sortbuf[i] = 0ull;
    __syncthreads();
    bitonic_sort_desc_u64(sortbuf, n_pow2);
    // ---- detector.py:73-74 + tester.py:53 for the selected rows only: bbox_pred, clip_boxes, / scale ----
    const double lim_x = A::sub(__ldg(g.im + b * 3 + 1), 1.0), lim_y = A::sub(__ldg(g.im + b * 3), 1.0);
    const double scale = __ldg(g.im + b * 3 + 2);
    for (int i = tid; i < n; i += blockDim.x) {
      const int r = (int)(u32)sortbuf[i];
      const float4 roi = __ldg(reinterpret_cast<const float4*>(g.rois + ((size_t)b * R + r) * 4));
      const float4 d = __ldg(reinterpret_cast<const float4*>(g.deltas + ((size_t)b * R + r) * 4 * C + 4 * j));
      const double x1 = roi.x, y1 = roi.y, x2 = roi.z, y2 = roi.w;
      const double w = A::add(A::sub(x2, x1), 1.0), h = A::add(A::sub(y2, y1), 1.0);
      const double cx = A::add(x1, A::mul(0.5, A::sub(w, 1.0))), cy = A::add(y1, A::mul(0.5, A::sub(h, 1.0)));
      const double pcx = A::add(A::mul((double)d.x, w), cx), pcy = A::add(A::mul((double)d.y, h), cy);
      const double pw = A::mul((double)expf(d.z), w), ph = A::mul((double)expf(d.w), h);
      const double hx = A::mul(0.5, A::sub(pw, 1.0)), hy = A::mul(0.5, A::sub(ph, 1.0));
      double bx[4] = {A::sub(pcx, hx), A::sub(pcy, hy), A::add(pcx, hx), A::add(pcy, hy)};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const double lim = (q & 1) ? lim_y : lim_x;
        bx[q] = A::div(clip_0_lim(bx[q], lim), scale);
      }
      sbox[i] = make_double4(bx[0], bx[1], bx[2], bx[3]);
      sarea[i] = box_area<double>(bx[0], bx[1], bx[2], bx[3]);
    }
    __syncthreads();

    // ---- nms.py:22-36  greedy NMS, 64 sorted positions at a time ----
    const double thr = g.nms_thr;
    for (int c0 = 0; c0 < n; c0 += 64) {
      const int m = min(64, n - c0), blk = c0 >> 6;
      {  // lower triangle of the diagonal tile: thread = (row, 4 columns); bit jj of row i: "jj suppresses i"
        const int i = tid >> 4, sub = tid & 15;
        u64 bits = 0ull;
        if (i < m) {
          const double4 bi = sbox[c0 + i];
          const double ai = sarea[c0 + i];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int jj = sub * 4 + q;
            if (jj < i) {
              const double4 bj = sbox[c0 + jj];
              if (iou_exceeds<double>(bj.x, bj.y, bj.z, bj.w, sarea[c0 + jj], bi.x, bi.y, bi.z, bi.w, ai, thr)) bits |= 1ull << jj;
            }
          }
        }
        u32 lo = (u32)bits, hi = (u32)(bits >> 32);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
          lo |= __shfl_xor_sync(0xffffffffu, lo, o);
          hi |= __shfl_xor_sync(0xffffffffu, hi, o);
        }
        if (sub == 0) s_M[i] = ((u64)hi << 32) | lo;
      }
      __syncthreads();
      if (tid < 32) {
        // greedy keep = the fixed point of  keep_i = alive_i and no kept jj < i suppresses i
        const u64 valid = m == 64 ? ~0ull : ((1ull << m) - 1ull);
        const u64 alive = ~s_removed[blk] & valid;
        const u64 m0 = s_M[lane], m1 = s_M[lane + 32];
        const bool a0 = (alive >> lane) & 1ull, a1 = (alive >> (lane + 32)) & 1ull;
        u64 keep = alive;
        for (int it = 0; it < 64; ++it) {
          const u32 k0 = __ballot_sync(0xffffffffu, a0 && (m0 & keep) == 0ull);
          const u32 k1 = __ballot_sync(0xffffffffu, a1 && (m1 & keep) == 0ull);
          const u64 nxt = ((u64)k1 << 32) | k0;
          if (nxt == keep) break;
          keep = nxt;
        }
        if (lane == 0) s_keptbits = keep;
      }
      __syncthreads();
      const u64 kept = s_keptbits;
      if (tid < 64 && ((kept >> tid) & 1ull)) skept[nk + __popcll(kept & ((1ull << tid) - 1ull))] = c0 + tid;
      nk += __popcll(kept);
      // boxes kept in this tile suppress the later positions (one position per thread, whole warps)
      for (int t0 = c0 + 64; t0 < n; t0 += blockDim.x) {
        const int t = t0 + tid;
        bool sup = false;
        if (t < n && !((s_removed[t >> 6] >> (t & 63)) & 1ull)) {
          const double4 bt = sbox[t];
          const double at = sarea[t];
          u64 rest = kept;
          while (rest) {
            const int jj = __ffsll((long long)rest) - 1;
            rest &= rest - 1ull;
            const double4 bj = sbox[c0 + jj];
            if (iou_exceeds<double>(bj.x, bj.y, bj.z, bj.w, sarea[c0 + jj], bt.x, bt.y, bt.z, bt.w, at, thr)) {
              sup = true;
              break;
            }
          }
        }
        const u32 bal = __ballot_sync(0xffffffffu, sup);
        if (lane == 0 && bal) atomicOr(reinterpret_cast<u32*>(s_removed) + (t >> 5), bal);
      }
      __syncthreads();
    }
    // ---- tester.py:54-56  all_boxes[j][i] = cls_dets[keep, :] ----
    for (int t = tid; t < nk; t += blockDim.x) {
      const int i = skept[t];
      const int r = (int)(u32)sortbuf[i];
      const double4 bx = sbox[i];
      double* o = g.dets_out + (((size_t)b * C + j) * R + t) * 5;
      o[0] = bx.x;
      o[1] = bx.y;
      o[2] = bx.z;
      o[3] = bx.w;
      o[4] = (double)__ldg(scores + (size_t)r * C + j);
      g.rows_out[((size_t)b * C + j) * R + t] = r;
    }
  }
  if (tid == 0) {
    g.count_out[((size_t)b * C + j) * 2] = nk;
    g.count_out[((size_t)b * C + j) * 2 + 1] = nk;
    if (j == 1) g.count_out[(size_t)b * C * 2] = g.count_out[(size_t)b * C * 2 + 1] = 0;
  }
  // ---- tester.py:58-65  cap at max_per_image over all classes: done by the image's last CTA ----
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(g.done + b, 1u) == (unsigned)(C - 2)) ? 1 : 0;
  __syncthreads();
  if (!s_last || g.max_per_image <= 0) return;
  __threadfence();
  int* prefix = reinterpret_cast<int*>(dyn_smem);  // [C + 1]; the sort buffer is free now
  if (tid == 0) {
    int run = 0;
    for (int c = 0; c < C; ++c) {
      prefix[c] = run;
      run += c == 0 ? 0 : __ldcg(g.count_out + ((size_t)b * C + c) * 2 + 1);
    }
    prefix[C] = run;
  }
  __syncthreads();
  const int total = prefix[C];
  if (total <= g.max_per_image) return;
  // image_thresh = sort(image_scores)[-max_per_image]; keep score >= image_thresh (a prefix of each sorted list)
  const u32 T = kth_largest_key(g, b, prefix, total, g.max_per_image, s_hist, s_st);
  for (int c = 1 + tid; c < C; c += blockDim.x) {
    const int cnt = prefix[c + 1] - prefix[c];
    const double* col = g.dets_out + ((size_t)b * C + c) * R * 5 + 4;
    int lo = 0, hi = cnt;  // first position whose key is below T
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (score_key32((float)__ldcg(col + (size_t)mid * 5)) >= T) lo = mid + 1; else hi = mid;
    }
    g.count_out[((size_t)b * C + c) * 2] = lo;
  }
}

}  // namespace
}  // namespace mxr

using namespace mxr;

namespace {

int pow2_at_least(int v, int floor_) {
  int p = floor_;
  while (p < v) p <<= 1;
  return p;
}

size_t box_head_smem(int sort_cap, int C) {
  const size_t lists = (size_t)sort_cap * (8 + 32 + 8 + 4);
  const size_t pre = (size_t)(C + 1) * 4;
  return lists > pre ? lists : pre;
}

int check_params(const mxr_box_head_params* p, const char* who) {
  MXR_REQUIRE(p != nullptr, MXR_ERR_INVALID, "%s: NULL params", who);
  MXR_REQUIRE(p->n_images >= 1 && p->n_rois >= 0 && p->n_classes >= 1, MXR_ERR_INVALID,
              "%s: bad shape: %d images, %d rois, %d classes", who, p->n_images, p->n_rois, p->n_classes);
  MXR_REQUIRE(p->n_rois <= MXR_BOX_HEAD_MAX_ROIS, MXR_ERR_INVALID, "%s: at most %d rois per image (got %d)", who,
              MXR_BOX_HEAD_MAX_ROIS, p->n_rois);
  MXR_REQUIRE(p->n_classes <= 65535 && p->n_images <= 65535, MXR_ERR_INVALID, "%s: at most 65535 classes / images", who);
  MXR_REQUIRE(p->nms_thresh == p->nms_thresh && p->score_thresh == p->score_thresh, MXR_ERR_INVALID, "%s: NaN threshold", who);
  return MXR_OK;
}

}  // namespace

extern "C" {

int mxr_box_head(mxr_handle_t hh, void* stream, const mxr_box_head_params* p, const float* rois, const float* bbox_deltas,
                 const float* scores, const double* im, double* dets_out, int32_t* rows_out, int32_t* count_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_box_head: NULL handle");
  MXR_TRY(check_params(p, "mxr_box_head"));
  MXR_REQUIRE(count_out != nullptr, MXR_ERR_INVALID, "mxr_box_head: NULL count_out");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  const int B = p->n_images, R = p->n_rois, C = p->n_classes;
  h->last_launches = 0;
  if (R == 0 || C < 2) {  // nothing to detect: every list is empty
    MXR_CUDA(cudaMemsetAsync(count_out, 0, (size_t)B * C * 2 * sizeof(int32_t), s));
    return MXR_OK;
  }
  MXR_REQUIRE(rois && bbox_deltas && scores && im && dets_out && rows_out, MXR_ERR_INVALID, "mxr_box_head: NULL array");
  MXR_TRY(ensure_device_ws(h, (size_t)B * sizeof(unsigned int)));
  unsigned int* done = reinterpret_cast<unsigned int*>(h->d_ws);
  MXR_CUDA(cudaMemsetAsync(done, 0, (size_t)B * sizeof(unsigned int), s));
  BoxHeadArgs g;
  g.rois = rois;
  g.deltas = bbox_deltas;
  g.scores = scores;
  g.im = im;
  g.R = R;
  g.C = C;
  g.score_thr = (float)p->score_thresh;  // tester.py:51: float32 array > python float compares in float32
  g.nms_thr = p->nms_thresh;
  g.max_per_image = p->max_per_image;
  g.sort_cap = pow2_at_least(R, 64);
  g.dets_out = dets_out;
  g.rows_out = rows_out;
  g.count_out = count_out;
  g.done = done;
  const size_t smem = box_head_smem(g.sort_cap, C);
  MXR_TRY(raise_dyn_smem(h, 0, (const void*)box_head_kernel, smem, nullptr));
  prof_begin(h, s, ST_BOX_HEAD);
  box_head_kernel<<<dim3(C - 1, B), 1024, smem, s>>>(g);
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_box_head_host(mxr_handle_t hh, const mxr_box_head_params* p, const float* rois, const float* bbox_deltas,
                      const float* scores, const double* im, double* dets_out, int32_t* rows_out, int32_t* count_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_box_head_host: NULL handle");
  MXR_TRY(check_params(p, "mxr_box_head_host"));
  MXR_REQUIRE(count_out != nullptr, MXR_ERR_INVALID, "mxr_box_head_host: NULL count_out");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t B = p->n_images, R = p->n_rois, C = p->n_classes;
  if (R == 0 || C < 2) {
    memset(count_out, 0, B * C * 2 * sizeof(int32_t));
    return MXR_OK;
  }
  MXR_REQUIRE(rois && bbox_deltas && scores && im && dets_out && rows_out, MXR_ERR_INVALID, "mxr_box_head_host: NULL array");
  cudaStream_t s = h->stream;
  MXR_TRY(ws_order(h, s));
  // workspace: [done counters | inputs | outputs]
  Carver probe(nullptr);
  probe.take<unsigned int>(B);
  probe.take<float>(B * R * 4);
  probe.take<float>(B * R * 4 * C);
  probe.take<float>(B * R * C);
  probe.take<double>(B * 3);
  probe.take<double>(B * C * R * 5);
  probe.take<int32_t>(B * C * R);
  probe.take<int32_t>(B * C * 2);
  MXR_TRY(ensure_device_ws(h, probe.off));
  Carver cv(h->d_ws);
  cv.take<unsigned int>(B);
  float* d_rois = cv.take<float>(B * R * 4);
  float* d_deltas = cv.take<float>(B * R * 4 * C);
  float* d_scores = cv.take<float>(B * R * C);
  double* d_im = cv.take<double>(B * 3);
  double* d_dets = cv.take<double>(B * C * R * 5);
  int32_t* d_rows = cv.take<int32_t>(B * C * R);
  int32_t* d_count = cv.take<int32_t>(B * C * 2);
  MXR_CUDA(cudaMemcpyAsync(d_rois, rois, B * R * 4 * sizeof(float), cudaMemcpyHostToDevice, s));
  MXR_CUDA(cudaMemcpyAsync(d_deltas, bbox_deltas, B * R * 4 * C * sizeof(float), cudaMemcpyHostToDevice, s));
  MXR_CUDA(cudaMemcpyAsync(d_scores, scores, B * R * C * sizeof(float), cudaMemcpyHostToDevice, s));
  MXR_CUDA(cudaMemcpyAsync(d_im, im, B * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
  MXR_TRY(mxr_box_head(hh, s, p, d_rois, d_deltas, d_scores, d_im, d_dets, d_rows, d_count));
  MXR_CUDA(cudaMemcpyAsync(count_out, d_count, B * C * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaStreamSynchronize(s));
  // Only the filled head of every class list travels back: one strided copy per array, `most` rows per list
  // (typically ~10 of R rows: max_per_image detections spread over the classes).
  size_t most = 0;
  for (size_t bc = 0; bc < B * C; ++bc) most = (size_t)count_out[bc * 2] > most ? (size_t)count_out[bc * 2] : most;
  if (most > 0) {
    MXR_CUDA(cudaMemcpy2DAsync(dets_out, R * 5 * sizeof(double), d_dets, R * 5 * sizeof(double), most * 5 * sizeof(double), B * C,
                               cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpy2DAsync(rows_out, R * sizeof(int32_t), d_rows, R * sizeof(int32_t), most * sizeof(int32_t), B * C,
                               cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaStreamSynchronize(s));
  }
  return MXR_OK;
}

}  // extern "C"
