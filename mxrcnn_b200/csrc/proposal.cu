// rcnn/rpn/proposal.py:39-149 -- host side of the proposal operator.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "nms_coop.cuh"
#include "nms_host.cuh"
#include "proposal_cluster.cuh"
#include "proposal_fused.cuh"
#include "proposal_kernels.cuh"

using namespace mxr;

namespace {

const int kSmemSortMax = 16384;  // pairs that fit the 227 KB of shared memory (12 B each)

struct Layout {
  // derived sizes
  int B, A, H, W, hc, wc, n_crop, n_full, cropped;
  int n_max, cap, keep_cap, out_rows, n_pow2_max, use_smem_sort, head;
  StripePlan plan;
  // device carve
  Ctrl* ctrl;
  u64* R;
  size_t zero_bytes;  // ctrl + R are contiguous and zeroed per call
  u32* keys;
  double4* box;
  double* area;
  float* score;
  int* orig;
  int* kept_slot;
  u64* M;
  u64* g_sort_keys;
  u64* dbg;  // [B][16] phase time stamps of the fused kernel, then [B][32] cycle stamps
  // second-generation fused kernel (proposal_fused.cuh)
  u64* comp;      // [B][n_crop] bucketed composites
  float4* kbox;   // [B][keep_cap]
  float4* kaux;   // [B][keep_cap]
  u64* fsort;     // [B][fsort_cap] scratch for crowded buckets
  int fsort_cap;
  u32* pair_count;  // [B]
  // staging for the *_host entry point
  float* d_cls;
  float* d_box;
  float* d_info;
  char* d_out;
  size_t out_bytes, off_scores, off_keep, off_anchor, off_nkeep;
  size_t total;
};

int make_layout(const mxr_proposal_params* p, int out_capacity, bool host_io, char* base, Layout* L) {
  MXR_REQUIRE(p != nullptr, MXR_ERR_INVALID, "proposal: params is NULL");
  MXR_REQUIRE(p->batch >= 1 && p->num_anchors >= 1 && p->num_anchors <= 64 && p->height >= 1 && p->width >= 1,
              MXR_ERR_INVALID, "proposal: bad shape B=%d A=%d H=%d W=%d (A must be 1..64)", p->batch, p->num_anchors,
              p->height, p->width);
  MXR_REQUIRE(p->feat_stride > 0, MXR_ERR_INVALID, "proposal: feat_stride must be > 0");
  MXR_REQUIRE(p->batch <= 65535, MXR_ERR_LIMIT, "proposal: at most 65535 images per call (got %d)", p->batch);
  L->B = p->batch;
  L->A = p->num_anchors;
  L->H = p->height;
  L->W = p->width;
  L->hc = (p->crop_h > 0 && p->crop_h < p->height) ? p->crop_h : p->height;
  L->wc = (p->crop_w > 0 && p->crop_w < p->width) ? p->crop_w : p->width;
  L->cropped = (L->hc < L->H || L->wc < L->W) ? 1 : 0;
  const long long n_crop = (long long)L->A * L->hc * L->wc, n_full = (long long)L->A * L->H * L->W;
  MXR_REQUIRE(n_full <= (1 << 24), MXR_ERR_LIMIT, "proposal: A*H*W = %lld exceeds the maximum of 16777216", n_full);
  L->n_crop = (int)n_crop;
  L->n_full = (int)n_full;
  L->n_max = (p->pre_nms_top_n > 0 && p->pre_nms_top_n < L->n_crop) ? p->pre_nms_top_n : L->n_crop;
  L->cap = round_up_int(L->n_max, 64);
  L->keep_cap = (p->post_nms_top_n > 0 && p->post_nms_top_n < L->n_max) ? p->post_nms_top_n : L->n_max;
  L->out_rows = p->post_nms_top_n > 0 ? p->post_nms_top_n : out_capacity;
  MXR_REQUIRE(L->out_rows >= 1, MXR_ERR_INVALID, "proposal: post_nms_top_n <= 0 needs out_capacity >= 1");
  L->n_pow2_max = pow2_ceil(L->n_max < 2 ? 2 : L->n_max);
  L->use_smem_sort = L->n_pow2_max <= kSmemSortMax;
  L->plan = plan_stripes(L->n_max, p->post_nms_top_n > 0 ? L->keep_cap : 0, p->first_stripe);
  // "head": the prefix of the candidate order that is sorted up front (whole 1024-box stripes);
  // the rest is sorted by the deferred phase of the NMS kernel only if some image gets that far
  L->head = round_up_int(L->plan.first, kCoopW);
  if (L->head > kMaxStripe) L->head = kMaxStripe;
  if (L->head >= L->n_max) L->head = round_up_int(L->n_max, 64);
  L->plan.max_width = kCoopW;
  Carver c(base);
  const int r_words = L->cap / 64;
  L->ctrl = c.take<Ctrl>(L->B);
  L->R = c.take<u64>((size_t)L->B * r_words);
  L->zero_bytes = c.off;
  L->keys = c.take<u32>((size_t)L->B * L->n_crop);
  L->box = c.take<double4>((size_t)L->B * L->cap);
  L->area = c.take<double>((size_t)L->B * L->cap);
  L->score = c.take<float>((size_t)L->B * L->cap);
  L->orig = c.take<int>((size_t)L->B * L->cap);
  L->kept_slot = c.take<int>((size_t)L->B * L->keep_cap);
  const size_t m_words = (size_t)L->plan.max_width * (L->plan.max_width / 64);
  L->M = c.take<u64>((size_t)L->B * m_words);
  L->g_sort_keys = L->use_smem_sort ? nullptr : c.take<u64>((size_t)L->B * L->n_pow2_max);
  L->dbg = c.take<u64>((size_t)L->B * 48);  // 16 phase stamps + 32 cycle stamps (second-generation kernel)
  L->comp = c.take<u64>((size_t)L->B * L->n_crop);
  L->kbox = c.take<float4>((size_t)L->B * L->keep_cap);
  L->kaux = c.take<float4>((size_t)L->B * L->keep_cap);
  L->fsort_cap = pow2_ceil(L->n_crop < 2 ? 2 : L->n_crop);
  L->fsort = c.take<u64>((size_t)L->B * L->fsort_cap);
  L->pair_count = c.take<u32>((size_t)L->B * 2);
  if (host_io) {
    const size_t hw = (size_t)L->H * L->W;
    // ONE block -- foreground scores | deltas | im_info -- so that a caller-packed host buffer arrives in one copy
    L->d_cls = c.take<float>((size_t)L->B * L->A * hw + (size_t)L->B * 4 * L->A * hw + (size_t)L->B * 3);
    L->d_box = L->d_cls + (size_t)L->B * L->A * hw;
    L->d_info = L->d_box + (size_t)L->B * 4 * L->A * hw;
    const size_t rows = (size_t)L->B * L->out_rows;
    size_t o = 0;
    o += align_up(rows * 5 * sizeof(float), 256);
    L->off_scores = o;
    o += align_up(rows * sizeof(float), 256);
    L->off_keep = o;
    o += align_up(rows * sizeof(int), 256);
    L->off_anchor = o;
    o += align_up(rows * sizeof(int), 256);
    L->off_nkeep = o;
    o += align_up((size_t)L->B * 2 * sizeof(int), 256);
    L->out_bytes = o;
    L->d_out = c.take<char>(o);
  }
  L->total = align_up(c.off, 256);
  return MXR_OK;
}

// Deferred phase of the NMS kernel: full top-k (select + sort + re-decode) of one image.
struct ProposalTier2 {
  TopkCall tc;
  AnchorTable at;
  int head;
  __device__ __forceinline__ bool wanted(int) const { return true; }
  __device__ __forceinline__ void run(int b, Ctrl* ctrl, unsigned char* smem, u32* st, int* cnt) const {
    topk_body<0>(tc, at, 0, ctrl, b, smem, st, cnt);
  }
};



// The float32 anchor tables of the fused kernel's screens (proposal_fused.cuh: decode32_core) and their conversion errors.
void fill_decode_tables(fz::FusedArgs* fa, const mxr_proposal_params* p, const ProposalGeom& pg, const AnchorTable& at, int A) {
  fa->pg = pg;
  fa->at = at;
  const double fs = p->feat_stride;
  fa->fs32 = (float)fs;
  fa->fs_err = (float)fabs(fs - (double)fa->fs32) * 1.0000002f;
  for (int a = 0; a < A; ++a) {
    // bbox_transform.py:47-50 on the base anchor: width, height, centre (the shift w * feat_stride is added in the kernel)
    const double bw = at.a[a][2] - at.a[a][0] + 1.0, bh = at.a[a][3] - at.a[a][1] + 1.0;
    const double cx = at.a[a][0] + 0.5 * (bw - 1.0), cy = at.a[a][1] + 0.5 * (bh - 1.0);
    const double v[4] = {bw, bh, cx, cy};
    float f[4];
    double err = 0.0;
    for (int k = 0; k < 4; ++k) {
      f[k] = (float)v[k];
      const double d = fabs(v[k] - (double)f[k]);
      if (d > err) err = d;
    }
    fa->at32[a] = make_float4(f[0], f[1], f[2], f[3]);
    // the kernel multiplies deltas (|dx| can be large) into bw / bh: a converted width carries a RELATIVE error, which
    // the kernel's bound covers through its 2^-23 |dx * bw| term as long as the conversion is within 2^-24 relative (it is)
    fa->at_err[a] = (float)err * 1.0000002f + (err > 0.0 ? 1e-30f : 0.0f);
  }
  fa->min_size = p->min_size;
}

// Second-generation fused kernel (proposal_fused.cuh): ONE launch, any list length.  Returns MXR_ERR_UNSUPPORTED when an
// image does not fit 16 CTAs x 12 keys x 1024 threads (the caller then takes the multi-kernel path).
static const int MXR_ERR_UNSUPPORTED = -1000;
int enqueue_fused(Handle* h, cudaStream_t s, const mxr_proposal_params* p, const Layout& L, const ProposalGeom& pg,
                  const AnchorTable& at, const float* cls_prob, const float* bbox_pred, const float* im_info, float* out_rois,
                  float* out_scores, int32_t* out_keep, int32_t* out_anchor, int32_t* out_nkeep) {
  static const int kNumSizes = 8;
  static const int kSizes[kNumSizes] = {16, 14, 12, 10, 8, 6, 4, 2};
  static const double kWaveCost[kNumSizes] = {0.883, 0.89, 0.925, 0.99, 1.0, 1.075, 1.35, 1.85};
  if ((L.n_crop + 16 * 1024 - 1) / (16 * 1024) > 12) return MXR_ERR_UNSUPPORTED;
  fz::FusedArgs* fa = new fz::FusedArgs();
  struct Guard {
    fz::FusedArgs* p;
    ~Guard() { delete p; }
  } guard{fa};
  memset(fa, 0, sizeof(*fa));
  fill_decode_tables(fa, p, pg, at, L.A);
  fa->cls_prob = cls_prob;
  fa->bbox_pred = bbox_pred;
  fa->im_info = im_info;
  fa->pre_nms = p->pre_nms_top_n;
  fa->max_keep = p->post_nms_top_n > 0 ? L.keep_cap : 0;
  // stripe widths: the first stripe should settle a small quota on its own (the 300th keep sits near sorted position
  // 390 at the BASELINE configuration); later stripes and quota-less walks use the widest stripe
  int w0 = fz::kWMax, w1 = fz::kWMax;
  if (fa->max_keep > 0 && fa->max_keep < 360) {
    // 1.4 x the quota + 28: at the BASELINE configuration the 300th keep sits at sorted position 380 +- 11 (max 408 over
    // 120 seeds), so 448 candidates settle the image in one stripe; when they do not, the next stripe carries on
    w0 = round_up_int(fa->max_keep + (fa->max_keep * 2) / 5 + 28, 32);
    if (w0 > fz::kWMax) w0 = fz::kWMax;
    w1 = 256;
  }
  if (p->first_stripe > 0) {
    w0 = round_up_int(p->first_stripe, 32);
    if (w0 > fz::kWMax) w0 = fz::kWMax;
  }
  fa->first_stripe = w0;
  fa->next_stripe = w1;
  fa->window0 = fa->max_keep > 0 ? 2 * fa->max_keep + 256 : 0;
  fa->thr = p->nms_thresh;
  fa->comp = L.comp;
  fa->kbox = L.kbox;
  fa->kaux = L.kaux;
  fa->keep_cap = L.keep_cap;
  fa->sortbuf = L.fsort;
  fa->sort_cap = L.fsort_cap;
  fa->o.rois = out_rois;
  fa->o.scores = out_scores;
  fa->o.keep_pos = out_keep;
  fa->o.keep_orig = out_anchor;
  fa->o.nkeep = out_nkeep;
  fa->o.rows = L.out_rows;
  fa->o.row_offsets = nullptr;
  fa->ctrl = L.ctrl;
  h->ctrl_ptr = L.ctrl;
  h->ctrl_images = L.B;
  fa->dbg = h->prof_on ? L.dbg : nullptr;
  fa->clk = h->prof_on ? L.dbg + (size_t)L.B * 16 : nullptr;
  h->dbg_ptr = fa->dbg;
  h->dbg_images = fa->dbg ? L.B : 0;
  fa->pair_count = h->count_pairs ? L.pair_count : nullptr;
  h->pair_ptr = fa->pair_count;
  h->pair_images = fa->pair_count ? L.B : 0;
  if (fa->pair_count) MXR_CUDA(cudaMemsetAsync(L.pair_count, 0, (size_t)L.B * 2 * sizeof(u32), s));
  int csize = 0, items = 0;
  double best = 0;
  int pinned = h->cluster_override;  // ignored when the image does not fit 12 keys per thread at that size
  if (pinned && (L.n_crop + pinned * 1024 - 1) / (pinned * 1024) > 12) pinned = 0;
  void (*kq)(const fz::FusedArgs) = fz::proposal_fused_kernel;
  for (int c = 0; c < kNumSizes; ++c) {
    const int cs = kSizes[c];
    if (pinned && pinned != cs) continue;
    const int it = (L.n_crop + cs * 1024 - 1) / (cs * 1024);
    if (it > 12) continue;
    // the occupancy figure depends on the shared-memory size, which depends on the keys per thread: two buckets
    const int v = it <= 4 ? 0 : 1;
    const size_t smem_v = fz::fused_smem_bytes(v == 0 ? 4 : 12);
    MXR_TRY(raise_dyn_smem(h, 8, (const void*)kq, fz::fused_smem_bytes(12), nullptr));
    int& cap = h->cluster_cap[(2 + v) * kNumSizes + c];
    if (cap == 0) {
      cudaLaunchConfig_t q;
      memset(&q, 0, sizeof(q));
      q.gridDim = dim3(cs);
      q.blockDim = dim3(1024);
      q.dynamicSmemBytes = smem_v;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = cs;
      qa[0].val.clusterDim.y = 1;
      qa[0].val.clusterDim.z = 1;
      q.attrs = qa;
      q.numAttrs = 1;
      int n = 0;
      if (cs > 8 && cudaFuncSetAttribute(kq, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) n = -1;
      if (n == 0 && cudaOccupancyMaxActiveClusters(&n, kq, &q) != cudaSuccess) n = -1;
      (void)cudaGetLastError();
      cap = n > 0 ? n : -1;
    }
    if (cap <= 0) continue;
    const double cost = (double)((L.B + cap - 1) / cap) * kWaveCost[c];
    if (csize == 0 || cost < best) {
      csize = cs;
      items = it;
      best = cost;
    }
  }
  MXR_REQUIRE(csize != 0, MXR_ERR_CUDA, "proposal: no cluster size of the fused kernel fits this device");
  void (*kern)(const fz::FusedArgs) = fz::proposal_fused_kernel;
  fa->items = items;
  fa->chunk = (L.n_crop + csize - 1) / csize;
  fa->hwc = L.hc * L.wc;
  fa->hw32 = L.H * L.W;
  fa->rcp_hwc = 1.0f / (float)fa->hwc;
  fa->rcp_wc = 1.0f / (float)L.wc;
  fa->rcp_A = 1.0f / (float)L.A;
  fa->a_step = 1024 / fa->hwc;
  fa->r_step = 1024 % fa->hwc;
  fa->h_step = fa->r_step / L.wc;
  fa->w_step = fa->r_step % L.wc;
  {
    const char* nr = getenv("MXRCNN_B200_NULL");
    fa->null_run = (nr && nr[0] == '1') ? 1 : 0;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(csize, L.B);  // blockIdx.y = image: no division in the kernel
  cfg.blockDim = dim3(1024);
  cfg.dynamicSmemBytes = fz::fused_smem_bytes(items <= 4 ? 4 : 12);
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (h->no_pdl || fa->pair_count) ? 1 : 2;
  prof_begin(h, s, ST_FUSED);
  MXR_CUDA(cudaLaunchKernelEx(&cfg, kern, *fa));
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int enqueue(Handle* h, cudaStream_t s, const mxr_proposal_params* p, const Layout& L, const float* cls_prob,
            size_t cls_img_stride, size_t cls_fg_offset, const float* bbox_pred, const float* im_info,
            const double* anchors, float* out_rois, float* out_scores, int32_t* out_keep, int32_t* out_anchor,
            int32_t* out_nkeep) {
  AnchorTable at;
  memset(&at, 0, sizeof(at));
  memcpy(at.a, anchors, sizeof(double) * 4 * L.A);
  ProposalGeom pg;
  pg.g.A = L.A;
  pg.g.H = L.H;
  pg.g.W = L.W;
  pg.g.hc = L.hc;
  pg.g.wc = L.wc;
  pg.g.fs = p->feat_stride;
  pg.n_crop = L.n_crop;
  pg.n_full = L.n_full;
  pg.cropped = L.cropped;
  pg.cls_img_stride = cls_img_stride;
  pg.cls_fg_offset = cls_fg_offset;
  h->last_launches = 0;
  // Fast path: the fused cluster kernel (ONE launch for the whole op, one 8-CTA cluster per image).  It needs
  // an image's anchors to fit 8 CTAs x 12 keys x 1024 threads and the pre-NMS list to fit the leader's
  // shared-memory sort buffer; otherwise, or with MXRCNN_B200_NO_CLUSTER=1 (used by the tests to cover both),
  // the multi-kernel path runs.
  if (!h->no_cluster && !h->use_v1) {
    const int st = enqueue_fused(h, s, p, L, pg, at, cls_prob, bbox_pred, im_info, out_rois, out_scores, out_keep, out_anchor,
                                 out_nkeep);
    if (st != MXR_ERR_UNSUPPORTED) return st;  // an image too large for 16 CTAs x 12 keys x 1024 threads: multi-kernel path
  }
  const int cluster_items = (L.n_crop + kClusterSize * 1024 - 1) / (kClusterSize * 1024);
  int sort_cap = pow2_ceil(L.n_max > L.head ? L.n_max : L.head);
  if (sort_cap < 1024) sort_cap = 1024;
  const bool use_cluster = cluster_items <= 12 && sort_cap <= kClusterSortMax && !h->no_cluster;
  // the fused kernel initialises everything it uses; only the multi-kernel path needs zeroed control blocks
  if (!use_cluster) MXR_CUDA(cudaMemsetAsync(L.ctrl, 0, L.zero_bytes, s));
  if (!use_cluster) {
    const int n_threads = L.n_crop > L.n_full ? L.n_crop : L.n_full;
    dim3 g1((n_threads + 255) / 256, L.B);
    prof_begin(h, s, ST_DECODE);
    decode_filter_kernel<<<g1, 256, 0, s>>>(pg, at, cls_prob, bbox_pred, im_info, p->min_size, L.keys, L.ctrl);
    prof_end(h, s);
  }
  TopkCall tc;
  tc.pg = pg;
  tc.ta.keys = L.keys;
  tc.ta.box = L.box;
  tc.ta.area = L.area;
  tc.ta.score = L.score;
  tc.ta.orig = L.orig;
  tc.ta.cap = L.cap;
  tc.ta.g_sort_keys = L.g_sort_keys;
  tc.ta.g_sort_cap = L.n_pow2_max;
  tc.cls_prob = cls_prob;
  tc.bbox_pred = bbox_pred;
  tc.im_info = im_info;
  tc.pre_nms = p->pre_nms_top_n;
  // ---- tier 1: sort only the head of the candidate order (<= 4096 boxes: 32 KiB of smem) ----
  const int head_pow2 = pow2_ceil(L.head < 2 ? 2 : L.head);
  size_t smem1 = (size_t)head_pow2 * 8;
  if (smem1 < 32 * 256 * sizeof(u32)) smem1 = 32 * 256 * sizeof(u32);
  tc.smem_pairs = (int)(smem1 / 8);
  if (!use_cluster) {
    prof_begin(h, s, ST_TOPK);
    if (L.n_crop <= 24 * 1024) {
      topk_kernel<24><<<L.B, 1024, smem1, s>>>(tc, at, L.head, 0, L.ctrl);
    } else {
      topk_kernel<0><<<L.B, 1024, smem1, s>>>(tc, at, L.head, 0, L.ctrl);
    }
    prof_end(h, s);
  }
  // ---- NMS: one cooperative kernel; its deferred phase sorts the rest if an image needs it ----
  CoopArgs<double> ca;
  ca.ctrl = L.ctrl;
  ca.a.box = L.box;
  ca.a.area = L.area;
  ca.a.score = L.score;
  ca.a.orig = L.orig;
  ca.a.cap = L.cap;
  ca.a.M = L.M;
  ca.a.m_pitch = kCoopSW;
  ca.a.m_stride = (size_t)kCoopW * kCoopSW;
  ca.a.R = L.R;
  ca.a.r_words = L.cap / 64;
  ca.a.kept_slot = L.kept_slot;
  ca.a.keep_cap = L.keep_cap;
  ca.o.rois = out_rois;
  ca.o.scores = out_scores;
  ca.o.keep_pos = out_keep;
  ca.o.keep_orig = out_anchor;
  ca.o.nkeep = out_nkeep;
  ca.o.rows = L.out_rows;
  ca.o.row_offsets = nullptr;
  ca.B = L.B;
  ca.n_max = L.n_max;
  ca.max_keep = p->post_nms_top_n > 0 ? L.keep_cap : 0;
  ca.tier2_at = (L.head < L.n_max) ? L.head / kCoopW : -1;
  ca.thr = p->nms_thresh;
  if (use_cluster) {
    ClusterArgs cl;
    cl.pg = pg;
    cl.at = at;
    cl.cls_prob = cls_prob;
    cl.bbox_pred = bbox_pred;
    cl.im_info = im_info;
    cl.min_size = p->min_size;
    cl.pre_nms = p->pre_nms_top_n;
    cl.head = L.head;
    cl.max_keep = ca.max_keep;
    cl.thr = p->nms_thresh;
    cl.sort_cap = sort_cap;
    cl.a = ca.a;
    cl.o = ca.o;
    cl.ctrl = L.ctrl;
    // phase time stamps (%globaltimer reads on the leader's critical path) only while profiling is switched on
    cl.dbg = h->prof_on ? L.dbg : nullptr;
    h->dbg_ptr = cl.dbg;
    h->dbg_images = cl.dbg ? L.B : 0;
    const size_t smem_c = cluster_smem_bytes(sort_cap);
    // Cluster size per call.  A wide cluster shortens every phase of one image (measured per wave at 38x63x9:
    // 47 / 50 / 54 / 58 / 72 / 98 us for 16 / 12 / 8 / 6 / 4 / 2 CTAs) but fewer of them are resident at once
    // (7 / 7 / 14 / 16+ / 36 / 74 on a B200: cudaOccupancyMaxActiveClusters), so a small batch wants wide clusters
    // and a large one narrow clusters: pick the size that minimises waves x relative wave time.
    // MXRCNN_B200_CLUSTER pins it (tests run several sizes).
    static const int kNumSizes = 8;
    static const int kSizes[kNumSizes] = {16, 14, 12, 10, 8, 6, 4, 2};
    static const double kWaveCost[kNumSizes] = {0.883, 0.89, 0.925, 0.99, 1.0, 1.075, 1.35, 1.85};
    int csize = 0, items = 0;
    double best = 0;
    int pinned = h->cluster_override;  // ignored when the image does not fit 12 keys per thread at that size
    if (pinned && (L.n_crop + pinned * 1024 - 1) / (pinned * 1024) > 12) pinned = 0;
    for (int c = 0; c < kNumSizes; ++c) {
      const int cs = kSizes[c];
      if (pinned && pinned != cs) continue;
      const int it = (L.n_crop + cs * 1024 - 1) / (cs * 1024);
      if (it > 12) continue;
      const int v = it <= 4 ? 0 : 1;  // kernel variant: 4 or 12 keys per thread
      void (*kq)(ClusterArgs) = v == 0 ? proposal_cluster_kernel<4> : proposal_cluster_kernel<12>;
      MXR_TRY(raise_dyn_smem(h, 4 + v, (const void*)kq, smem_c, nullptr));
      int& cap = h->cluster_cap[v * kNumSizes + c];
      if (cap == 0) {  // how many clusters of this size the device holds at once (0: not asked yet, -1: none)
        cudaLaunchConfig_t q;
        memset(&q, 0, sizeof(q));
        q.gridDim = dim3(cs);
        q.blockDim = dim3(1024);
        q.dynamicSmemBytes = smem_c;
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = cs;
        qa[0].val.clusterDim.y = 1;
        qa[0].val.clusterDim.z = 1;
        q.attrs = qa;
        q.numAttrs = 1;
        int n = 0;
        if (cs > 8 && cudaFuncSetAttribute(kq, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) n = -1;
        if (n == 0 && cudaOccupancyMaxActiveClusters(&n, kq, &q) != cudaSuccess) n = -1;
        (void)cudaGetLastError();
        cap = n > 0 ? n : -1;
      }
      if (cap <= 0) continue;
      const double cost = (double)((L.B + cap - 1) / cap) * kWaveCost[c];
      if (csize == 0 || cost < best) {
        csize = cs;
        items = it;
        best = cost;
      }
    }
    MXR_REQUIRE(csize != 0, MXR_ERR_CUDA, "proposal: no cluster size of the fused kernel fits this device");
    void (*kern)(ClusterArgs) = items <= 4 ? proposal_cluster_kernel<4> : proposal_cluster_kernel<12>;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(L.B * csize);
    cfg.blockDim = dim3(1024);
    cfg.dynamicSmemBytes = smem_c;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = csize;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    // Programmatic dependent launch: the kernel lets its successor be scheduled as soon as all of its own CTAs run
    // and itself waits (griddepcontrol.wait, first thing) until everything before it in the stream has completed, so
    // back-to-back calls do not pay the launch latency between them.  Stream order of the results is unchanged.
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = h->no_pdl ? 1 : 2;
    prof_begin(h, s, ST_CLUSTER);
    MXR_CUDA(cudaLaunchKernelEx(&cfg, kern, cl));
    prof_end(h, s);
    MXR_CUDA(cudaGetLastError());
    return MXR_OK;  // the fused kernel settles every image itself
  }
  ProposalTier2 t2;
  t2.tc = tc;
  t2.at = at;
  t2.head = L.head;
  size_t tier2_smem = 0;
  if (ca.tier2_at != -1) tier2_smem = L.use_smem_sort ? (size_t)L.n_pow2_max * 8 : 32 * 256 * sizeof(u32);
  const size_t smem2 = coop_smem_bytes(sizeof(double4), sizeof(double), tier2_smem);
  t2.tc.smem_pairs = L.use_smem_sort ? (int)(smem2 / 8) : 0;
  int raised2 = 0;
  MXR_TRY(raise_dyn_smem(h, 3, (const void*)nms_coop_kernel<double, ProposalTier2>, smem2, &raised2));
  if (h->coop_grid[0] == 0 || raised2) {
    MXR_CUDA(cudaFuncSetAttribute(nms_coop_kernel<double, ProposalTier2>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    int per_sm = 0, sms = 0;
    MXR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nms_coop_kernel<double, ProposalTier2>, 1024, smem2));
    MXR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
    MXR_REQUIRE(per_sm >= 1, MXR_ERR_CUDA, "proposal: the cooperative NMS kernel does not fit on an SM");
    h->coop_grid[0] = per_sm * sms;
  }
  void* kargs[] = {(void*)&ca, (void*)&t2};
  prof_begin(h, s, ST_NMS_COOP);
  MXR_CUDA(cudaLaunchCooperativeKernel((const void*)nms_coop_kernel<double, ProposalTier2>, dim3(h->coop_grid[0]), dim3(1024),
                                       kargs, smem2, s));
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

// proposal.py:135-141: rows nkeep ... nkeep + n - 1 of an image repeat kept rows drawn by the host.  The draws travel as a
// kernel argument (up to kPadChunk per launch): no staging buffer, nothing to synchronise with.
static const int kPadChunk = 768;
struct PadRows {
  int src[kPadChunk];
};
__global__ void pad_rows_kernel(float* rois, float* scores, int row0, int nkeep, int first, int n, const __grid_constant__ PadRows pr) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int sidx = pr.src[t];
  if (sidx < 0 || sidx >= nkeep) return;
  for (int k = 0; k < 5; ++k) rois[(size_t)(row0 + nkeep + first + t) * 5 + k] = rois[(size_t)(row0 + sidx) * 5 + k];
  if (scores) scores[row0 + nkeep + first + t] = scores[row0 + sidx];
}

// mxr_topk_sorted: out_order[image][position in the pre-NMS list] = anchor row, from the walk-order outputs.
__global__ void topk_unwalk_kernel(const int32_t* pos, const int32_t* rows, const int32_t* count, int per_image, int32_t* out_order) {
  const int b = blockIdx.y, t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count[b * 2] || t >= per_image) return;
  const int p = pos[(size_t)b * per_image + t];
  if (p >= 0 && p < per_image) out_order[(size_t)b * per_image + p] = rows[(size_t)b * per_image + t];
}

}  // namespace

extern "C" {

int mxr_reserve_proposal(mxr_handle_t hh, const mxr_proposal_params* p) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_reserve_proposal: NULL handle");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  Layout L;
  MXR_TRY(make_layout(p, p->post_nms_top_n > 0 ? p->post_nms_top_n : p->num_anchors * p->height * p->width, true,
                      nullptr, &L));
  MXR_TRY(ensure_device_ws(h, 2 * L.total));
  MXR_TRY(ensure_pinned_ws(h, 2 * align_up(L.out_bytes, 256)));
  return MXR_OK;
}

int mxr_proposal_forward(mxr_handle_t hh, void* stream, const mxr_proposal_params* p, const float* cls_prob,
                         const float* bbox_pred, const float* im_info, const double* anchors, int out_capacity,
                         float* out_rois, float* out_scores, int32_t* out_keep, int32_t* out_anchor_index,
                         int32_t* out_nkeep) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_proposal_forward: NULL handle");
  MXR_REQUIRE(cls_prob && bbox_pred && im_info && anchors && out_rois && out_nkeep, MXR_ERR_INVALID,
              "mxr_proposal_forward: NULL array argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  Layout L;
  MXR_TRY(make_layout(p, out_capacity, false, nullptr, &L));
  MXR_TRY(ensure_device_ws(h, L.total));
  MXR_TRY(make_layout(p, out_capacity, false, h->d_ws, &L));
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  const size_t hw = (size_t)L.H * L.W;
  return enqueue(h, s, p, L, cls_prob, 2 * L.A * hw, L.A * hw, bbox_pred, im_info, anchors, out_rois, out_scores,
                 out_keep, out_anchor_index, out_nkeep);
}

// Pipelined host path: two slots, each with its own device staging, stream and pinned result block,
// so that the H2D copy of one batch overlaps the kernels of the previous one.
struct HostSlots {
  Layout L[2];
  int busy[2];
};

static int submit_host_impl(mxr_handle_t hh, int slot, const mxr_proposal_params* p, const float* cls_prob,
                            const float* bbox_pred, const float* im_info, const float* packed, const double* anchors,
                            int out_capacity) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_proposal_submit_host: NULL handle");
  MXR_REQUIRE(slot == 0 || slot == 1, MXR_ERR_INVALID, "mxr_proposal_submit_host: slot must be 0 or 1");
  MXR_REQUIRE(((cls_prob && bbox_pred && im_info) || packed) && anchors, MXR_ERR_INVALID,
              "mxr_proposal_submit_host: NULL array argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  if (!h->prop_state) {
    h->prop_state = calloc(1, sizeof(HostSlots));
    MXR_REQUIRE(h->prop_state != nullptr, MXR_ERR_NOMEM, "mxr_proposal_submit_host: out of host memory");
  }
  HostSlots* hs = reinterpret_cast<HostSlots*>(h->prop_state);
  MXR_REQUIRE(!hs->busy[slot], MXR_ERR_INVALID, "mxr_proposal_submit_host: slot %d still holds results (call wait)", slot);
  if (!h->stream2) MXR_CUDA(cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
  Layout& L = hs->L[slot];
  MXR_TRY(make_layout(p, out_capacity, true, nullptr, &L));
  const size_t total = L.total, out_bytes = align_up(L.out_bytes, 256);
  if (2 * total > h->d_bytes || 2 * out_bytes > h->h_bytes) {
    // growing the blocks would pull the rug from under the other slot
    MXR_REQUIRE(!hs->busy[1 - slot], MXR_ERR_INVALID,
                "mxr_proposal_submit_host: workspace must grow while slot %d is in flight; call mxr_reserve_proposal first",
                1 - slot);
    MXR_TRY(ensure_device_ws(h, 2 * total));
    MXR_TRY(ensure_pinned_ws(h, 2 * out_bytes));
  }
  MXR_TRY(make_layout(p, out_capacity, true, h->d_ws + (size_t)slot * (h->d_bytes / 2 / 256 * 256), &L));
  cudaStream_t s = slot ? h->stream2 : h->stream;
  MXR_TRY(ws_order(h, s));
  const size_t hw = (size_t)L.H * L.W;
  const size_t fg_bytes = (size_t)L.A * hw * sizeof(float);
  if (packed) {  // foreground scores | deltas | im_info, as the device block is laid out: one copy
    MXR_CUDA(cudaMemcpyAsync(L.d_cls, packed, (size_t)L.B * 5 * fg_bytes + (size_t)L.B * 3 * sizeof(float),
                             cudaMemcpyHostToDevice, s));
  } else {
    // only the foreground half of cls_prob crosses the bus (proposal.py:57)
    MXR_CUDA(cudaMemcpy2DAsync(L.d_cls, fg_bytes, cls_prob + (size_t)L.A * hw, 2 * fg_bytes, fg_bytes, L.B,
                               cudaMemcpyHostToDevice, s));
    MXR_CUDA(cudaMemcpyAsync(L.d_box, bbox_pred, (size_t)L.B * 4 * fg_bytes, cudaMemcpyHostToDevice, s));
    MXR_CUDA(cudaMemcpyAsync(L.d_info, im_info, (size_t)L.B * 3 * sizeof(float), cudaMemcpyHostToDevice, s));
  }
  float* d_rois = reinterpret_cast<float*>(L.d_out);
  float* d_scores = reinterpret_cast<float*>(L.d_out + L.off_scores);
  int* d_keep = reinterpret_cast<int*>(L.d_out + L.off_keep);
  int* d_anchor = reinterpret_cast<int*>(L.d_out + L.off_anchor);
  int* d_nkeep = reinterpret_cast<int*>(L.d_out + L.off_nkeep);
  MXR_TRY(enqueue(h, s, p, L, L.d_cls, (size_t)L.A * hw, 0, L.d_box, L.d_info, anchors, d_rois, d_scores, d_keep,
                  d_anchor, d_nkeep));
  MXR_CUDA(cudaMemcpyAsync(h->h_pin + (size_t)slot * (h->h_bytes / 2 / 256 * 256), L.d_out, L.out_bytes,
                           cudaMemcpyDeviceToHost, s));
  hs->busy[slot] = 1;
  return MXR_OK;
}

int mxr_proposal_submit_host(mxr_handle_t hh, int slot, const mxr_proposal_params* p, const float* cls_prob,
                             const float* bbox_pred, const float* im_info, const double* anchors, int out_capacity) {
  MXR_REQUIRE(cls_prob && bbox_pred && im_info, MXR_ERR_INVALID, "mxr_proposal_submit_host: NULL array argument");
  return submit_host_impl(hh, slot, p, cls_prob, bbox_pred, im_info, nullptr, anchors, out_capacity);
}

int mxr_proposal_submit_host_packed(mxr_handle_t hh, int slot, const mxr_proposal_params* p, const float* packed,
                                    const double* anchors, int out_capacity) {
  MXR_REQUIRE(packed != nullptr, MXR_ERR_INVALID, "mxr_proposal_submit_host_packed: NULL array argument");
  return submit_host_impl(hh, slot, p, nullptr, nullptr, nullptr, packed, anchors, out_capacity);
}

int mxr_proposal_wait_host(mxr_handle_t hh, int slot, float* out_rois, float* out_scores, int32_t* out_keep,
                           int32_t* out_anchor_index, int32_t* out_nkeep) {
  MXR_REQUIRE(hh != nullptr && out_rois != nullptr && out_nkeep != nullptr, MXR_ERR_INVALID,
              "mxr_proposal_wait_host: NULL argument");
  MXR_REQUIRE(slot == 0 || slot == 1, MXR_ERR_INVALID, "mxr_proposal_wait_host: slot must be 0 or 1");
  Handle* h = reinterpret_cast<Handle*>(hh);
  HostSlots* hs = reinterpret_cast<HostSlots*>(h->prop_state);
  MXR_REQUIRE(hs && hs->busy[slot], MXR_ERR_INVALID, "mxr_proposal_wait_host: nothing was submitted to slot %d", slot);
  MXR_DEVICE(h);
  const Layout& L = hs->L[slot];
  MXR_CUDA(cudaStreamSynchronize(slot ? h->stream2 : h->stream));
  hs->busy[slot] = 0;
  const char* src = h->h_pin + (size_t)slot * (h->h_bytes / 2 / 256 * 256);
  const size_t rows = (size_t)L.B * L.out_rows;
  memcpy(out_rois, src, rows * 5 * sizeof(float));
  if (out_scores) memcpy(out_scores, src + L.off_scores, rows * sizeof(float));
  if (out_keep) memcpy(out_keep, src + L.off_keep, rows * sizeof(int));
  if (out_anchor_index) memcpy(out_anchor_index, src + L.off_anchor, rows * sizeof(int));
  memcpy(out_nkeep, src + L.off_nkeep, (size_t)L.B * 2 * sizeof(int));
  for (int b = 0; b < L.B; ++b) {
    if (out_nkeep[b * 2 + 1] & 1) {
      set_error("there is nan in input scores or bbox_deltas (image %d)", b);
      return MXR_ERR_NAN;
    }
  }
  return MXR_OK;
}

int mxr_proposal_forward_host(mxr_handle_t hh, const mxr_proposal_params* p, const float* cls_prob,
                              const float* bbox_pred, const float* im_info, const double* anchors, int out_capacity,
                              float* out_rois, float* out_scores, int32_t* out_keep, int32_t* out_anchor_index,
                              int32_t* out_nkeep) {
  MXR_REQUIRE(out_rois && out_nkeep, MXR_ERR_INVALID, "mxr_proposal_forward_host: NULL array argument");
  MXR_TRY(mxr_proposal_submit_host(hh, 0, p, cls_prob, bbox_pred, im_info, anchors, out_capacity));
  return mxr_proposal_wait_host(hh, 0, out_rois, out_scores, out_keep, out_anchor_index, out_nkeep);
}

// Stage entry (SURVEY 8b mxr_topk_sorted; rcnn/rpn/proposal.py:113-127): the complete pre-NMS list in processing order.
// It is the op itself with a threshold nothing exceeds (an IoU is at most 1) and no post-NMS cut: every candidate is
// "kept", so the anchor-index output is the ordered list the NMS walks.  The float32 screen is off at that threshold,
// every pair takes the float64 test: milliseconds, a stage for tests, not for the hot path.
int mxr_topk_sorted(mxr_handle_t hh, void* stream, const mxr_proposal_params* p, const float* cls_prob, const float* bbox_pred,
                    const float* im_info, const double* anchors, int out_capacity, int32_t* out_order, int32_t* out_count) {
  MXR_REQUIRE(hh != nullptr && p && cls_prob && bbox_pred && im_info && anchors && out_order && out_count, MXR_ERR_INVALID,
              "mxr_topk_sorted: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  mxr_proposal_params q = *p;
  q.nms_thresh = 1.0;
  q.post_nms_top_n = 0;
  Layout L;
  MXR_TRY(make_layout(&q, out_capacity, true, nullptr, &L));
  MXR_REQUIRE(out_capacity >= L.n_max, MXR_ERR_INVALID, "mxr_topk_sorted: out_capacity %d < list length %d", out_capacity, L.n_max);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  MXR_TRY(ensure_device_ws(h, L.total));
  MXR_TRY(make_layout(&q, out_capacity, true, h->d_ws, &L));
  const size_t hw = (size_t)L.H * L.W;
  // everything goes to the layout's own output staging; the rows come out in the order the NMS walks them (nms.py:19
  // re-sorts the list, which turns every run of equal scores around), next to their positions in the list itself:
  // one scatter puts them back
  int32_t* pos = reinterpret_cast<int32_t*>(L.d_out + L.off_keep);
  int32_t* rows = reinterpret_cast<int32_t*>(L.d_out + L.off_anchor);
  MXR_TRY(enqueue(h, s, &q, L, cls_prob, 2 * L.A * hw, L.A * hw, bbox_pred, im_info, anchors, reinterpret_cast<float*>(L.d_out),
                  reinterpret_cast<float*>(L.d_out + L.off_scores), pos, rows, out_count));
  topk_unwalk_kernel<<<dim3((L.out_rows + 255) / 256, L.B), 256, 0, s>>>(pos, rows, out_count, L.out_rows, out_order);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_decode_clip(mxr_handle_t hh, void* stream, const mxr_proposal_params* p, const float* bbox_pred, const float* im_info,
                    const double* anchors, double* out_boxes, int32_t* out_valid) {
  MXR_REQUIRE(hh != nullptr && bbox_pred && im_info && anchors && out_boxes && out_valid, MXR_ERR_INVALID,
              "mxr_decode_clip: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  Layout L;
  MXR_TRY(make_layout(p, 1, false, nullptr, &L));
  AnchorTable at;
  memset(&at, 0, sizeof(at));
  memcpy(at.a, anchors, sizeof(double) * 4 * L.A);
  ProposalGeom pg;
  memset(&pg, 0, sizeof(pg));
  pg.g.A = L.A;
  pg.g.H = L.H;
  pg.g.W = L.W;
  pg.g.hc = L.hc;
  pg.g.wc = L.wc;
  pg.g.fs = p->feat_stride;
  pg.n_crop = L.n_crop;
  pg.n_full = L.n_full;
  pg.cropped = L.cropped;
  fz::FusedArgs* fa = new fz::FusedArgs();
  memset(fa, 0, sizeof(*fa));
  fill_decode_tables(fa, p, pg, at, L.A);
  fa->bbox_pred = bbox_pred;
  fa->im_info = im_info;
  dim3 grid((L.n_crop + 255) / 256, L.B);
  fz::decode_stage_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(*fa, out_boxes, out_valid);
  delete fa;
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_proposal_counts(mxr_handle_t hh, int max_images, int32_t* counts_out, int32_t* images_out) {
  MXR_REQUIRE(hh != nullptr && counts_out != nullptr && images_out != nullptr && max_images >= 1, MXR_ERR_INVALID,
              "mxr_proposal_counts: bad argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  *images_out = 0;
  if (!h->ctrl_ptr || h->ctrl_images <= 0) return MXR_OK;
  MXR_CUDA(cudaDeviceSynchronize());
  const int n = h->ctrl_images < max_images ? h->ctrl_images : max_images;
  Ctrl* tmp = new Ctrl[n];
  cudaError_t e = cudaMemcpy(tmp, h->ctrl_ptr, (size_t)n * sizeof(Ctrl), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    delete[] tmp;
    set_error("mxr_proposal_counts: %s", cudaGetErrorString(e));
    return MXR_ERR_CUDA;
  }
  for (int b = 0; b < n; ++b) {
    counts_out[b * 4] = tmp[b].n_valid;
    counts_out[b * 4 + 1] = tmp[b].n;
    counts_out[b * 4 + 2] = tmp[b].nkept;
    counts_out[b * 4 + 3] = tmp[b].nan_flag;
  }
  delete[] tmp;
  *images_out = n;
  return MXR_OK;
}

int mxr_proposal_pad(mxr_handle_t hh, void* stream, int image, int rows_per_image, int nkeep, const int32_t* pad_rows,
                     int n_pad, float* out_rois, float* out_scores) {
  MXR_REQUIRE(hh != nullptr && out_rois != nullptr, MXR_ERR_INVALID, "mxr_proposal_pad: NULL argument");
  if (n_pad <= 0) return MXR_OK;
  MXR_REQUIRE(nkeep > 0, MXR_ERR_EMPTY, "mxr_proposal_pad: nothing kept, cannot pad (proposal.py:136-138)");
  MXR_REQUIRE(pad_rows != nullptr && nkeep + n_pad <= rows_per_image && image >= 0, MXR_ERR_INVALID,
              "mxr_proposal_pad: bad pad request nkeep=%d n_pad=%d rows=%d", nkeep, n_pad, rows_per_image);
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  PadRows pr;
  for (int first = 0; first < n_pad; first += kPadChunk) {
    const int n = n_pad - first < kPadChunk ? n_pad - first : kPadChunk;
    memcpy(pr.src, pad_rows + first, (size_t)n * sizeof(int));
    pad_rows_kernel<<<(n + 127) / 128, 128, 0, s>>>(out_rois, out_scores, image * rows_per_image, nkeep, first, n, pr);
  }
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

}  // extern "C"
