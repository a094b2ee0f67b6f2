// Training-target generators: assign_anchor (rcnn/minibatch.py:204-395) and the device side of
// _sample_rois / sample_rois (rcnn/rpn/proposal_target.py:135-196, rcnn/minibatch.py:151-201).
#include <stdlib.h>
#include <string.h>

#include "anchor_fused.cuh"
#include "target_kernels.cuh"

using namespace mxr;

namespace {

struct AnchorTargetState {
  AnchorBatchGeom g;
  AnchorBatchWs w;
  int B, sum_g, csize;
  char* d_in;        // phase 1 inputs, one block: gt offsets (B+1) | im heights and widths (2B) | gt boxes (sum G, 5)
  char* d_p2;        // phase 2 inputs, one block: disable counts (2B) | rank offsets (B) | weights (2B) | ranks
  size_t p2_ints;    // capacity of d_p2 in 4-byte words
  char* d_out;       // staging of the four output maps for the *_host variant
  bool valid;
};

// State carried from mxr_proposal_target_phase1_host to _phase2_host: the rois, gt boxes and assignments stay on the
// device, only the index lists and the sampled indices cross the bus.
struct ProposalTargetState {
  int R, G;
  float* d_rois;
  float* d_gt;
  int* d_assign;
  double* d_maxov;
  int* d_lists;    // [3][R] + counts[3]
  char* d_out;     // phase 2 outputs, contiguous
  int* d_keep;
  size_t out_cap;  // bytes of d_out
  bool valid;
};

size_t pt_layout(int R, int G, int K, int C, char* base, ProposalTargetState* st) {
  Carver c(base);
  st->d_rois = c.take<float>((size_t)R * 5);
  st->d_gt = c.take<float>((size_t)(G > 0 ? G : 1) * 5);
  st->d_assign = c.take<int>(R);
  st->d_maxov = c.take<double>(R);
  st->d_lists = c.take<int>((size_t)3 * R + 4);
  st->d_keep = c.take<int>(K > 0 ? K : 1);
  st->out_cap = (size_t)K * (5 + 1 + 3 * 4 * (size_t)C) * sizeof(float);
  st->d_out = c.take<char>(st->out_cap + 16);
  return align_up(c.off, 256);
}

size_t at_layout(const AnchorBatchGeom& g, int B, int sum_g, char* base, AnchorTargetState* st) {
  Carver c(base);
  const size_t N = (size_t)g.N, G = (size_t)g.g_cap, nb = (size_t)B;
  st->w.counts = c.take<int>(nb * 4);
  st->w.colmax = c.take<double>(nb * G);
  st->w.inside_rank = c.take<int>(nb * N);
  st->w.anchors_in = c.take<double>(nb * N * 4);
  st->w.gt64 = c.take<double>(nb * G * 4);
  st->w.argmax = c.take<int>(nb * N);
  st->w.maxov = c.take<double>(nb * N);
  st->w.labels = c.take<float>(nb * N);
  st->w.fg_list = c.take<int>(nb * N);
  st->w.bg_list = c.take<int>(nb * N);
  st->d_in = c.take<char>(((size_t)B + 1 + 2 * nb + (size_t)(sum_g > 0 ? sum_g : 1) * 5) * 4);
  st->p2_ints = 5 * nb + 2 * nb * N;
  st->d_p2 = c.take<char>(st->p2_ints * 4);
  st->d_out = c.take<char>(nb * N * 13 * sizeof(float));  // label N + 3 maps of 4N per image
  return align_up(c.off, 256);
}

// Cluster size of the fused anchor-target kernels: the smallest that leaves <= 3 anchors per thread, 16 CTAs at most
// (4 anchors per thread there); 0: the map is too large, the single-image kernels run instead.
int at_cluster_size(int N) {
  for (int c = 1; c <= 16; c *= 2)
    if ((N + c * kAtThreads - 1) / (c * kAtThreads) <= (c == 16 ? kAtItems : kAtItems - 1)) return c;
  return 0;
}

int at_launch_cfg(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, const void* func, int csize, int B, cudaStream_t s,
                  int threads = kAtThreads) {
  memset(cfg, 0, sizeof(*cfg));
  cfg->gridDim = dim3(csize, B);
  cfg->blockDim = dim3(threads);
  cfg->dynamicSmemBytes = 0;
  cfg->stream = s;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg->attrs = attr;
  cfg->numAttrs = 1;
  if (csize > 8) MXR_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  return MXR_OK;
}

int at_phase1(Handle* h, const mxr_anchor_target_params* p, const double* anchors, int B, const float* im_hw,
              const float* gt_boxes, const int32_t* gt_off, int32_t* counts_out) {
  MXR_REQUIRE(p->feat_height >= 1 && p->feat_width >= 1 && p->num_anchors >= 1 && p->num_anchors <= 64, MXR_ERR_INVALID,
              "mxr_anchor_target_phase1: bad shape");
  MXR_REQUIRE(B >= 1 && B <= 65535 && gt_off[0] == 0, MXR_ERR_INVALID, "mxr_anchor_target_phase1: bad batch");
  int g_max = 0;
  for (int b = 0; b < B; ++b) {
    MXR_REQUIRE(gt_off[b + 1] >= gt_off[b], MXR_ERR_INVALID, "mxr_anchor_target_phase1: gt offsets must not decrease");
    g_max = gt_off[b + 1] - gt_off[b] > g_max ? gt_off[b + 1] - gt_off[b] : g_max;
  }
  const int sum_g = gt_off[B];
  MXR_REQUIRE(sum_g == 0 || gt_boxes != nullptr, MXR_ERR_INVALID, "mxr_anchor_target_phase1: NULL gt_boxes");
  const long long N = (long long)p->num_anchors * p->feat_height * p->feat_width;
  MXR_REQUIRE(N <= (1 << 24), MXR_ERR_LIMIT, "mxr_anchor_target_phase1: %lld anchors exceed the maximum of 16777216", N);
  MXR_REQUIRE(N * B <= (1ll << 27), MXR_ERR_LIMIT, "mxr_anchor_target_phase1: %lld anchors in the batch exceed the maximum", N * B);
  if (!h->at_state) {
    h->at_state = calloc(1, sizeof(AnchorTargetState));
    MXR_REQUIRE(h->at_state != nullptr, MXR_ERR_NOMEM, "mxr_anchor_target_phase1: out of host memory");
  }
  AnchorTargetState* st = reinterpret_cast<AnchorTargetState*>(h->at_state);
  st->valid = false;
  AnchorBatchGeom& g = st->g;
  g.fh = p->feat_height;
  g.fw = p->feat_width;
  g.A = p->num_anchors;
  g.N = (int)N;
  g.feat_stride = p->feat_stride;
  g.allowed_border = p->allowed_border;
  g.g_cap = g_max > 0 ? g_max : 1;
  g.neg_ov = p->negative_overlap;
  g.pos_ov = p->positive_overlap;
  g.clobber = p->clobber_positives;
  st->B = B;
  st->sum_g = sum_g;
  st->csize = h->use_v1 ? 0 : at_cluster_size(g.N);
  g.items = st->csize ? (g.N + st->csize * kAtThreads - 1) / (st->csize * kAtThreads) : 0;
  cudaStream_t s = h->stream;
  MXR_TRY(ws_order(h, s));
  MXR_TRY(ensure_device_ws(h, at_layout(g, B, sum_g, nullptr, st)));
  at_layout(g, B, sum_g, h->d_ws, st);
  const size_t in_bytes = ((size_t)B + 1 + 2 * (size_t)B + (size_t)sum_g * 5) * 4;
  MXR_TRY(ensure_pinned_ws(h, in_bytes > (size_t)B * 16 ? in_bytes : (size_t)B * 16));
  MXR_CUDA(cudaStreamSynchronize(s));  // the pinned block may still feed an earlier copy
  memcpy(h->h_pin, gt_off, ((size_t)B + 1) * 4);
  memcpy(h->h_pin + ((size_t)B + 1) * 4, im_hw, (size_t)B * 8);
  if (sum_g > 0) memcpy(h->h_pin + ((size_t)B + 1 + 2 * (size_t)B) * 4, gt_boxes, (size_t)sum_g * 20);
  MXR_CUDA(cudaMemcpyAsync(st->d_in, h->h_pin, in_bytes, cudaMemcpyHostToDevice, s));
  st->w.gt_off = reinterpret_cast<const int*>(st->d_in);
  st->w.im_hw = reinterpret_cast<const float*>(st->d_in + ((size_t)B + 1) * 4);
  st->w.gt = reinterpret_cast<const float*>(st->d_in + ((size_t)B + 1 + 2 * (size_t)B) * 4);
  AnchorTable at;
  memset(&at, 0, sizeof(at));
  memcpy(at.a, anchors, sizeof(double) * 4 * g.A);
  h->last_launches = 0;
  if (st->csize) {
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    MXR_TRY(at_launch_cfg(&cfg, attr, (const void*)anchor_phase1_kernel, st->csize, B, s));
    prof_begin(h, s, ST_ANCHOR_TARGET);
    const cudaError_t le = cudaLaunchKernelEx(&cfg, anchor_phase1_kernel, g, at, st->w);
    if (le != cudaSuccess && st->csize > 8) {
      // a device (partition) that cannot host a 16-CTA cluster: the single-image kernels below take over
      (void)cudaGetLastError();
      st->csize = 0;
      g.items = 0;
    } else {
      MXR_CUDA(le);
      prof_end(h, s);
    }
  }
  if (!st->csize) {
    // maps beyond one cluster's registers: the single-image kernels, image by image
    MXR_CUDA(cudaMemsetAsync(st->w.counts, 0, (char*)st->w.inside_rank - (char*)st->w.counts, s));  // counts, colmax
    for (int b = 0; b < B; ++b) {
      AnchorTargetGeom g1;
      g1.fh = g.fh;
      g1.fw = g.fw;
      g1.A = g.A;
      g1.N = g.N;
      g1.feat_stride = g.feat_stride;
      g1.allowed_border = g.allowed_border;
      g1.im_h = im_hw[b * 2];
      g1.im_w = im_hw[b * 2 + 1];
      g1.G = gt_off[b + 1] - gt_off[b];
      AnchorTargetWs w1;
      w1.inside_rank = st->w.inside_rank + (size_t)b * g.N;
      w1.anchors_in = st->w.anchors_in + (size_t)b * g.N * 4;
      w1.gt64 = st->w.gt64 + (size_t)b * g.g_cap * 4;
      w1.argmax = st->w.argmax + (size_t)b * g.N;
      w1.maxov = st->w.maxov + (size_t)b * g.N;
      w1.colmax = st->w.colmax + (size_t)b * g.g_cap;
      w1.labels = st->w.labels + (size_t)b * g.N;
      w1.fg_list = st->w.fg_list + (size_t)b * g.N;
      w1.bg_list = st->w.bg_list + (size_t)b * g.N;
      w1.counts = st->w.counts + (size_t)b * 4;
      const float* d_gt = st->w.gt + (size_t)gt_off[b] * 5;
      prof_begin(h, s, ST_ANCHOR_TARGET);
      anchor_inside_kernel<<<1, 1024, 0, s>>>(g1, at, d_gt, w1);
      prof_end(h, s);
      const int grid = (g.N + 127) / 128;
      if (g1.G > 0) {
        prof_begin(h, s, ST_ANCHOR_TARGET);
        anchor_overlaps_kernel<<<grid, 128, 0, s>>>(g1, w1);
        prof_end(h, s);
      }
      prof_begin(h, s, ST_ANCHOR_TARGET);
      anchor_label_kernel<<<grid, 128, 0, s>>>(g1, w1, g.neg_ov, g.pos_ov, g.clobber);
      prof_end(h, s);
      prof_begin(h, s, ST_ANCHOR_TARGET);
      anchor_lists_kernel<<<1, 1024, 0, s>>>(w1);
      prof_end(h, s);
    }
  }
  MXR_CUDA(cudaGetLastError());
  MXR_CUDA(cudaMemcpyAsync(h->h_pin, st->w.counts, (size_t)B * 4 * sizeof(int), cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaStreamSynchronize(s));
  for (int b = 0; b < B; ++b) memcpy(counts_out + b * 3, h->h_pin + (size_t)b * 16, 3 * sizeof(int));
  st->valid = true;
  return MXR_OK;
}

// fg[b] / bg[b]: image b's ranks to disable (host); n_dis: [B][2]; weights: [B][2] positive, negative.
int at_phase2(Handle* h, cudaStream_t s, AnchorTargetState* st, const int32_t* const* fg, const int32_t* const* bg,
              const int32_t* n_dis, const double* weights, const float* iw4, float* label, float* target, float* inside,
              float* outside) {
  const AnchorBatchGeom& g = st->g;
  const int B = st->B;
  MXR_TRY(ws_order(h, s));  // `s` may be the caller's stream: the next user of the workspace must wait for this work
  size_t total = 0;
  for (int b = 0; b < B; ++b) {
    MXR_REQUIRE(n_dis[b * 2] >= 0 && n_dis[b * 2 + 1] >= 0 && n_dis[b * 2] + n_dis[b * 2 + 1] <= 2 * g.N, MXR_ERR_INVALID,
                "mxr_anchor_target_phase2: bad disable counts %d/%d", n_dis[b * 2], n_dis[b * 2 + 1]);
    total += (size_t)n_dis[b * 2] + n_dis[b * 2 + 1];
  }
  const size_t words = 5 * (size_t)B + total;
  MXR_TRY(ensure_pinned_ws(h, words * 4));
  MXR_CUDA(cudaStreamSynchronize(h->stream));  // the pinned block may still feed an earlier copy
  int* pi = reinterpret_cast<int*>(h->h_pin);
  float* pf = reinterpret_cast<float*>(h->h_pin) + 3 * (size_t)B;
  int* pr = pi + 5 * (size_t)B;
  size_t off = 0;
  for (int b = 0; b < B; ++b) {
    pi[b * 2] = n_dis[b * 2];
    pi[b * 2 + 1] = n_dis[b * 2 + 1];
    pi[2 * B + b] = (int)off;
    pf[b * 2] = (float)weights[b * 2];
    pf[b * 2 + 1] = (float)weights[b * 2 + 1];
    if (n_dis[b * 2]) memcpy(pr + off, fg[b], (size_t)n_dis[b * 2] * 4);
    off += n_dis[b * 2];
    if (n_dis[b * 2 + 1]) memcpy(pr + off, bg[b], (size_t)n_dis[b * 2 + 1] * 4);
    off += n_dis[b * 2 + 1];
  }
  MXR_CUDA(cudaMemcpyAsync(st->d_p2, h->h_pin, words * 4, cudaMemcpyHostToDevice, s));
  AnchorPhase2 a;
  a.dis_counts = reinterpret_cast<const int*>(st->d_p2);
  a.dis_off = a.dis_counts + 2 * (size_t)B;
  a.weights = reinterpret_cast<const float*>(st->d_p2) + 3 * (size_t)B;
  a.dis_ranks = a.dis_counts + 5 * (size_t)B;
  a.inside_w = make_float4(iw4[0], iw4[1], iw4[2], iw4[3]);
  a.label = label;
  a.target = target;
  a.inside = inside;
  a.outside = outside;
  const int csize = st->csize ? st->csize : 8;  // the output pass strides over any number of anchors
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  MXR_TRY(at_launch_cfg(&cfg, attr, (const void*)anchor_phase2_kernel, csize, B, s));
  prof_begin(h, s, ST_ANCHOR_PHASE2);
  MXR_CUDA(cudaLaunchKernelEx(&cfg, anchor_phase2_kernel, g, st->w, a));
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

}  // namespace

extern "C" {

int mxr_anchor_target_phase1(mxr_handle_t hh, const mxr_anchor_target_params* p, const double* anchors,
                             const float* gt_boxes, int32_t* counts_out) {
  MXR_REQUIRE(hh != nullptr && p != nullptr && anchors != nullptr && counts_out != nullptr, MXR_ERR_INVALID,
              "mxr_anchor_target_phase1: NULL argument");
  MXR_REQUIRE(p->num_gt >= 0, MXR_ERR_INVALID, "mxr_anchor_target_phase1: bad shape");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const float im_hw[2] = {p->im_height, p->im_width};
  const int32_t off[2] = {0, p->num_gt};
  return at_phase1(h, p, anchors, 1, im_hw, gt_boxes, off, counts_out);
}

int mxr_anchor_target_batch_phase1(mxr_handle_t hh, const mxr_anchor_target_params* p, const double* anchors, int n_images,
                                   const float* im_heights_widths, const float* gt_boxes, const int32_t* gt_offsets,
                                   int32_t* counts_out) {
  MXR_REQUIRE(hh != nullptr && p != nullptr && anchors != nullptr && counts_out != nullptr && im_heights_widths != nullptr &&
                  gt_offsets != nullptr && n_images >= 1,
              MXR_ERR_INVALID, "mxr_anchor_target_batch_phase1: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  return at_phase1(h, p, anchors, n_images, im_heights_widths, gt_boxes, gt_offsets, counts_out);
}

int mxr_anchor_target_phase2(mxr_handle_t hh, void* stream, const int32_t* fg_disable_ranks, int n_fg_disable,
                             const int32_t* bg_disable_ranks, int n_bg_disable, double positive_weight,
                             double negative_weight, const float* inside_weights4, float* label, float* bbox_target,
                             float* bbox_inside_weight, float* bbox_outside_weight) {
  MXR_REQUIRE(hh != nullptr && inside_weights4 && label && bbox_target && bbox_inside_weight && bbox_outside_weight,
              MXR_ERR_INVALID, "mxr_anchor_target_phase2: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  AnchorTargetState* st = reinterpret_cast<AnchorTargetState*>(h->at_state);
  MXR_REQUIRE(st && st->valid && st->B == 1, MXR_ERR_INVALID, "mxr_anchor_target_phase2: call mxr_anchor_target_phase1 first");
  MXR_DEVICE(h);
  const int32_t n_dis[2] = {n_fg_disable, n_bg_disable};
  const double wts[2] = {positive_weight, negative_weight};
  // phase 1 ran on the handle's stream and was synchronised, so its results are visible to any stream
  MXR_TRY(at_phase2(h, (cudaStream_t)stream, st, &fg_disable_ranks, &bg_disable_ranks, n_dis, wts, inside_weights4, label,
                    bbox_target, bbox_inside_weight, bbox_outside_weight));
  st->valid = false;
  return MXR_OK;
}

int mxr_anchor_target_batch_phase2(mxr_handle_t hh, void* stream, const int32_t* disable_ranks, const int32_t* disable_counts,
                                   const double* weights, const float* inside_weights4, float* label, float* bbox_target,
                                   float* bbox_inside_weight, float* bbox_outside_weight) {
  MXR_REQUIRE(hh != nullptr && disable_counts && weights && inside_weights4 && label && bbox_target && bbox_inside_weight &&
                  bbox_outside_weight,
              MXR_ERR_INVALID, "mxr_anchor_target_batch_phase2: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  AnchorTargetState* st = reinterpret_cast<AnchorTargetState*>(h->at_state);
  MXR_REQUIRE(st && st->valid, MXR_ERR_INVALID, "mxr_anchor_target_batch_phase2: call mxr_anchor_target_batch_phase1 first");
  MXR_DEVICE(h);
  const int B = st->B;
  const int32_t** ptrs = reinterpret_cast<const int32_t**>(malloc(sizeof(int32_t*) * 2 * (size_t)B));
  MXR_REQUIRE(ptrs != nullptr, MXR_ERR_NOMEM, "mxr_anchor_target_batch_phase2: out of host memory");
  size_t off = 0;
  bool ok = true;
  for (int b = 0; b < B; ++b) {
    ok = ok && disable_counts[b * 2] >= 0 && disable_counts[b * 2 + 1] >= 0;
    ptrs[b] = disable_ranks + off;
    off += ok ? (size_t)disable_counts[b * 2] : 0;
    ptrs[B + b] = disable_ranks + off;
    off += ok ? (size_t)disable_counts[b * 2 + 1] : 0;
  }
  int rc = MXR_ERR_INVALID;
  if (ok && (off == 0 || disable_ranks != nullptr))
    rc = at_phase2(h, (cudaStream_t)stream, st, ptrs, ptrs + B, disable_counts, weights, inside_weights4, label, bbox_target,
                   bbox_inside_weight, bbox_outside_weight);
  else
    set_error("mxr_anchor_target_batch_phase2: bad disable counts or NULL ranks");
  free(ptrs);
  if (rc == MXR_OK) st->valid = false;
  return rc;
}

int mxr_anchor_target_phase2_host(mxr_handle_t hh, const int32_t* fg_disable_ranks, int n_fg_disable,
                                  const int32_t* bg_disable_ranks, int n_bg_disable, double positive_weight,
                                  double negative_weight, const float* inside_weights4, float* label, float* bbox_target,
                                  float* bbox_inside_weight, float* bbox_outside_weight) {
  MXR_REQUIRE(hh != nullptr && inside_weights4 && label && bbox_target && bbox_inside_weight && bbox_outside_weight,
              MXR_ERR_INVALID, "mxr_anchor_target_phase2_host: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  AnchorTargetState* st = reinterpret_cast<AnchorTargetState*>(h->at_state);
  MXR_REQUIRE(st && st->valid && st->B == 1, MXR_ERR_INVALID, "mxr_anchor_target_phase2_host: call mxr_anchor_target_phase1 first");
  MXR_DEVICE(h);
  const size_t N = st->g.N;
  float* d = reinterpret_cast<float*>(st->d_out);
  cudaStream_t s = h->stream;
  MXR_TRY(ws_order(h, s));
  const int32_t n_dis[2] = {n_fg_disable, n_bg_disable};
  const double wts[2] = {positive_weight, negative_weight};
  MXR_TRY(at_phase2(h, s, st, &fg_disable_ranks, &bg_disable_ranks, n_dis, wts, inside_weights4, d, d + N, d + 5 * N, d + 9 * N));
  MXR_CUDA(cudaMemcpyAsync(label, d, N * 4, cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaMemcpyAsync(bbox_target, d + N, 4 * N * 4, cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaMemcpyAsync(bbox_inside_weight, d + 5 * N, 4 * N * 4, cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaMemcpyAsync(bbox_outside_weight, d + 9 * N, 4 * N * 4, cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaStreamSynchronize(s));
  st->valid = false;
  return MXR_OK;
}

// ---- fg / bg candidate lists -------------------------------------------------------------
int mxr_fg_bg_lists(mxr_handle_t hh, void* stream, const void* overlaps, int dtype, int n, double fg_thresh, double bg_hi,
                    double bg_lo, int32_t* fg_list, int32_t* bg_list, int32_t* counts) {
  MXR_REQUIRE(hh != nullptr && counts != nullptr, MXR_ERR_INVALID, "mxr_fg_bg_lists: NULL argument");
  MXR_REQUIRE(n >= 0 && (n == 0 || (overlaps && fg_list && bg_list)), MXR_ERR_INVALID, "mxr_fg_bg_lists: bad arrays");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  h->last_launches = 0;
  prof_begin(h, s, ST_PROPOSAL_TARGET);
  if (dtype == MXR_DTYPE_F64)
    fg_bg_lists_kernel<double><<<1, 1024, 0, s>>>((const double*)overlaps, n, fg_thresh, bg_hi, bg_lo, fg_list, bg_list, counts);
  else if (dtype == MXR_DTYPE_F32)
    fg_bg_lists_kernel<float><<<1, 1024, 0, s>>>((const float*)overlaps, n, (float)fg_thresh, (float)bg_hi, (float)bg_lo, fg_list,
                                                 bg_list, counts);
  else {
    set_error("mxr_fg_bg_lists: dtype must be f32 or f64");
    return MXR_ERR_INVALID;
  }
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_fg_bg_lists_host(mxr_handle_t hh, const void* overlaps, int dtype, int n, double fg_thresh, double bg_hi, double bg_lo,
                         int32_t* fg_list, int32_t* bg_list, int32_t* counts) {
  MXR_REQUIRE(hh != nullptr && counts != nullptr && n >= 0, MXR_ERR_INVALID, "mxr_fg_bg_lists_host: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t es = dtype == MXR_DTYPE_F64 ? 8 : 4;
  Carver c(nullptr);
  c.take<char>((size_t)n * es + 8);
  c.take<int>(n + 1);
  c.take<int>(n + 1);
  c.take<int>(2);
  MXR_TRY(ensure_device_ws(h, align_up(c.off, 256)));
  Carver d(h->d_ws);
  char* d_ov = d.take<char>((size_t)n * es + 8);
  int* d_fg = d.take<int>(n + 1);
  int* d_bg = d.take<int>(n + 1);
  int* d_cnt = d.take<int>(2);
  cudaStream_t s = h->stream;
  MXR_TRY(ws_order(h, s));
  if (n) MXR_CUDA(cudaMemcpyAsync(d_ov, overlaps, (size_t)n * es, cudaMemcpyHostToDevice, s));
  MXR_TRY(mxr_fg_bg_lists(hh, s, d_ov, dtype, n, fg_thresh, bg_hi, bg_lo, d_fg, d_bg, d_cnt));
  if (n) {
    MXR_CUDA(cudaMemcpyAsync(fg_list, d_fg, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpyAsync(bg_list, d_bg, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
  }
  MXR_CUDA(cudaMemcpyAsync(counts, d_cnt, 8, cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaStreamSynchronize(s));
  return MXR_OK;
}

// ---- proposal-target gather --------------------------------------------------------------
static int pt_args(const mxr_proposal_target_params* p, PtGather* a) {
  MXR_REQUIRE(p != nullptr, MXR_ERR_INVALID, "proposal_target: NULL params");
  MXR_REQUIRE(p->num_rois >= 1 && p->num_gt >= 1 && p->num_keep >= 0 && p->num_classes >= 1 && p->fg_this >= 0,
              MXR_ERR_INVALID, "proposal_target: bad sizes R=%d G=%d K=%d C=%d", p->num_rois, p->num_gt, p->num_keep,
              p->num_classes);
  a->K = p->num_keep;
  a->fg_this = p->fg_this;
  a->num_classes = p->num_classes;
  a->normalize = p->normalize;
  for (int i = 0; i < 4; ++i) {
    a->mean[i] = p->means[i];
    a->stdv[i] = p->stds[i];
    a->inside_w[i] = p->inside_weights[i];
  }
  return MXR_OK;
}

int mxr_proposal_target_gather(mxr_handle_t hh, void* stream, const mxr_proposal_target_params* p, const float* rois,
                               const float* gt_boxes, const int32_t* gt_assignment, const int32_t* keep_inds,
                               float* out_rois, float* labels, float* bbox_targets, float* inside, float* outside) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_proposal_target_gather: NULL handle");
  PtGather a;
  MXR_TRY(pt_args(p, &a));
  if (a.K == 0) return MXR_OK;
  MXR_REQUIRE(rois && gt_boxes && gt_assignment && keep_inds && out_rois && labels && bbox_targets && inside && outside,
              MXR_ERR_INVALID, "mxr_proposal_target_gather: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  a.rois = rois;
  a.gt = gt_boxes;
  a.assignment = gt_assignment;
  a.keep = keep_inds;
  a.out_rois = out_rois;
  a.labels = labels;
  a.targets = bbox_targets;
  a.inside = inside;
  a.outside = outside;
  const long long total = (long long)a.K * a.num_classes;
  h->last_launches = 0;
  prof_begin(h, s, ST_PROPOSAL_TARGET);
  proposal_target_gather_kernel<<<(int)((total + 255) / 256), 256, 0, s>>>(a);
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_proposal_target_gather_host(mxr_handle_t hh, const mxr_proposal_target_params* p, const float* rois,
                                    const float* gt_boxes, const int32_t* gt_assignment, const int32_t* keep_inds,
                                    float* out_rois, float* labels, float* bbox_targets, float* inside, float* outside) {
  MXR_REQUIRE(hh != nullptr && p != nullptr, MXR_ERR_INVALID, "mxr_proposal_target_gather_host: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t R = p->num_rois, G = p->num_gt, K = p->num_keep, C4 = 4 * (size_t)p->num_classes;
  if (K == 0) return MXR_OK;
  for (int pass = 0; pass < 2; ++pass) {
    Carver c(pass ? h->d_ws : nullptr);
    float* d_rois = c.take<float>(R * 5);
    float* d_gt = c.take<float>(G * 5);
    int* d_as = c.take<int>(R);
    int* d_keep = c.take<int>(K);
    float* d_or = c.take<float>(K * 5);
    float* d_lab = c.take<float>(K);
    float* d_t = c.take<float>(K * C4);
    float* d_i = c.take<float>(K * C4);
    float* d_o = c.take<float>(K * C4);
    if (!pass) {
      MXR_TRY(ensure_device_ws(h, align_up(c.off, 256)));
      continue;
    }
    cudaStream_t s = h->stream;
    MXR_TRY(ws_order(h, s));
    MXR_CUDA(cudaMemcpyAsync(d_rois, rois, R * 5 * 4, cudaMemcpyHostToDevice, s));
    MXR_CUDA(cudaMemcpyAsync(d_gt, gt_boxes, G * 5 * 4, cudaMemcpyHostToDevice, s));
    MXR_CUDA(cudaMemcpyAsync(d_as, gt_assignment, R * 4, cudaMemcpyHostToDevice, s));
    MXR_CUDA(cudaMemcpyAsync(d_keep, keep_inds, K * 4, cudaMemcpyHostToDevice, s));
    MXR_TRY(mxr_proposal_target_gather(hh, s, p, d_rois, d_gt, d_as, d_keep, d_or, d_lab, d_t, d_i, d_o));
    MXR_CUDA(cudaMemcpyAsync(out_rois, d_or, K * 5 * 4, cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpyAsync(labels, d_lab, K * 4, cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpyAsync(bbox_targets, d_t, K * C4 * 4, cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpyAsync(inside, d_i, K * C4 * 4, cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaMemcpyAsync(outside, d_o, K * C4 * 4, cudaMemcpyDeviceToHost, s));
    MXR_CUDA(cudaStreamSynchronize(s));
  }
  return MXR_OK;
}

// ---- _sample_rois in two calls (proposal_target.py:135-196) ------------------------------------
int mxr_proposal_target_phase1_host(mxr_handle_t hh, const float* rois, int num_rois, const float* gt_boxes, int num_gt,
                                    int max_keep, int num_classes, double fg_thresh, double bg_hi, double bg_lo,
                                    double fallback_hi, double fallback_lo, int32_t* fg_list, int32_t* bg_list,
                                    int32_t* fallback_list, int32_t* counts3, int32_t* gt_assignment, double* max_overlaps) {
  MXR_REQUIRE(hh != nullptr && rois && gt_boxes && fg_list && bg_list && fallback_list && counts3, MXR_ERR_INVALID,
              "mxr_proposal_target_phase1_host: NULL argument");
  MXR_REQUIRE(num_rois >= 1 && num_gt >= 1 && max_keep >= 1 && num_classes >= 1, MXR_ERR_INVALID,
              "mxr_proposal_target_phase1_host: bad sizes R=%d G=%d K=%d C=%d", num_rois, num_gt, max_keep, num_classes);
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  if (!h->pt_state) {
    h->pt_state = calloc(1, sizeof(ProposalTargetState));
    MXR_REQUIRE(h->pt_state != nullptr, MXR_ERR_NOMEM, "mxr_proposal_target_phase1_host: out of host memory");
  }
  ProposalTargetState* st = reinterpret_cast<ProposalTargetState*>(h->pt_state);
  st->valid = false;
  const int R = num_rois, G = num_gt;
  cudaStream_t s = h->stream;
  MXR_TRY(ws_order(h, s));
  MXR_TRY(ensure_device_ws(h, pt_layout(R, G, max_keep, num_classes, nullptr, st)));
  pt_layout(R, G, max_keep, num_classes, h->d_ws, st);
  const size_t list_bytes = ((size_t)3 * R + 4) * sizeof(int);
  MXR_TRY(ensure_pinned_ws(h, list_bytes + (size_t)R * 12 + st->out_cap + 64));
  st->R = R;
  st->G = G;
  MXR_CUDA(cudaMemcpyAsync(st->d_rois, rois, (size_t)R * 5 * sizeof(float), cudaMemcpyHostToDevice, s));
  MXR_CUDA(cudaMemcpyAsync(st->d_gt, gt_boxes, (size_t)G * 5 * sizeof(float), cudaMemcpyHostToDevice, s));
  h->last_launches = 0;
  prof_begin(h, s, ST_PROPOSAL_TARGET);
  proposal_target_phase1_kernel<<<1, 1024, 0, s>>>(st->d_rois, R, st->d_gt, G, fg_thresh, bg_hi, bg_lo, fallback_hi, fallback_lo,
                                                  st->d_assign, st->d_maxov, st->d_lists, st->d_lists + 3 * R);
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  // one copy back: the three lists and their counts (+ the assignments / overlaps when the caller wants them)
  char* pin = h->h_pin;
  MXR_CUDA(cudaMemcpyAsync(pin, st->d_lists, list_bytes, cudaMemcpyDeviceToHost, s));
  if (gt_assignment) MXR_CUDA(cudaMemcpyAsync(pin + list_bytes, st->d_assign, (size_t)R * 4, cudaMemcpyDeviceToHost, s));
  if (max_overlaps) MXR_CUDA(cudaMemcpyAsync(pin + list_bytes + (size_t)R * 4, st->d_maxov, (size_t)R * 8, cudaMemcpyDeviceToHost, s));
  MXR_CUDA(cudaStreamSynchronize(s));
  const int* hl = reinterpret_cast<const int*>(pin);
  memcpy(counts3, hl + 3 * R, 3 * sizeof(int));
  memcpy(fg_list, hl, (size_t)counts3[0] * sizeof(int));
  memcpy(bg_list, hl + R, (size_t)counts3[1] * sizeof(int));
  memcpy(fallback_list, hl + 2 * R, (size_t)counts3[2] * sizeof(int));
  if (gt_assignment) memcpy(gt_assignment, pin + list_bytes, (size_t)R * 4);
  if (max_overlaps) memcpy(max_overlaps, pin + list_bytes + (size_t)R * 4, (size_t)R * 8);
  st->valid = true;
  return MXR_OK;
}

int mxr_proposal_target_phase2_host(mxr_handle_t hh, const mxr_proposal_target_params* p, const int32_t* keep_inds,
                                    float* out_rois, float* labels, float* bbox_targets, float* inside, float* outside) {
  MXR_REQUIRE(hh != nullptr && p != nullptr, MXR_ERR_INVALID, "mxr_proposal_target_phase2_host: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  ProposalTargetState* st = reinterpret_cast<ProposalTargetState*>(h->pt_state);
  MXR_REQUIRE(st && st->valid, MXR_ERR_INVALID, "mxr_proposal_target_phase2_host: call mxr_proposal_target_phase1_host first");
  MXR_REQUIRE(p->num_rois == st->R && p->num_gt == st->G, MXR_ERR_INVALID,
              "mxr_proposal_target_phase2_host: sizes differ from phase 1 (R %d vs %d, G %d vs %d)", p->num_rois, st->R,
              p->num_gt, st->G);
  MXR_DEVICE(h);
  const size_t K = p->num_keep, C4 = 4 * (size_t)p->num_classes;
  st->valid = false;
  if (K == 0) return MXR_OK;
  MXR_REQUIRE(keep_inds && out_rois && labels && bbox_targets && inside && outside, MXR_ERR_INVALID,
              "mxr_proposal_target_phase2_host: NULL array");
  const size_t need = K * (5 + 1 + 3 * C4) * sizeof(float);
  MXR_REQUIRE(need <= st->out_cap, MXR_ERR_INVALID, "mxr_proposal_target_phase2_host: num_keep %zu exceeds phase 1's max_keep", K);
  cudaStream_t s = h->stream;
  MXR_TRY(ws_order(h, s));
  MXR_CUDA(cudaMemcpyAsync(st->d_keep, keep_inds, K * sizeof(int), cudaMemcpyHostToDevice, s));
  float* d = reinterpret_cast<float*>(st->d_out);
  float* d_or = d;
  float* d_lab = d_or + K * 5;
  float* d_t = d_lab + K;
  float* d_i = d_t + K * C4;
  float* d_o = d_i + K * C4;
  MXR_TRY(mxr_proposal_target_gather(hh, s, p, st->d_rois, st->d_gt, st->d_assign, st->d_keep, d_or, d_lab, d_t, d_i, d_o));
  MXR_CUDA(cudaMemcpyAsync(h->h_pin, d, need, cudaMemcpyDeviceToHost, s));  // ONE copy, then split on the host
  MXR_CUDA(cudaStreamSynchronize(s));
  const float* hp = reinterpret_cast<const float*>(h->h_pin);
  memcpy(out_rois, hp, K * 5 * sizeof(float));
  memcpy(labels, hp + K * 5, K * sizeof(float));
  memcpy(bbox_targets, hp + K * 6, K * C4 * sizeof(float));
  memcpy(inside, hp + K * 6 + K * C4, K * C4 * sizeof(float));
  memcpy(outside, hp + K * 6 + 2 * K * C4, K * C4 * sizeof(float));
  return MXR_OK;
}

}  // extern "C"
