// assign_anchor (rcnn/minibatch.py:204-395) for a BATCH of images in two launches, one thread-block cluster per image:
//   anchor_phase1_kernel   :260-316  inside test and ranks, bbox_overlaps with its row / column reductions, labels,
//                                    ordered fg / bg lists and their counts -- everything the host needs before it
//                                    draws from numpy's RNG (:320-334)
//   anchor_phase2_kernel   :321-385  applies the draws, writes labels / targets / weights in the output layouts
// The single-image kernels of target_kernels.cuh remain for feature maps too large for one cluster's registers.
#pragma once
#include <cooperative_groups.h>

#include "target_kernels.cuh"

namespace mxr {

static const int kAtThreads = 512;  // 128 registers per thread: up to kAtItems anchors live in registers
static const int kAtItems = 4;      // anchors per thread
static const int kAtChunk = 64;     // gt boxes per round: one 64-bit mask of intersecting gts per anchor

struct AnchorBatchGeom {
  int fh, fw, A, N;
  int feat_stride, allowed_border;
  int items;             // anchors per thread: ceil(N / (cluster size * kAtThreads)) <= kAtItems
  int g_cap;             // per-image capacity of the gt-indexed workspace arrays
  double neg_ov, pos_ov;
  int clobber;
};

struct AnchorBatchWs {
  // per-image slices: [B][N] (gt-indexed ones [B][g_cap]); same meaning as AnchorTargetWs
  int* inside_rank;
  double* anchors_in;
  double* gt64;
  int* argmax;
  double* maxov;
  double* colmax;
  float* labels;
  int* fg_list;
  int* bg_list;
  int* counts;           // [B][4]
  const float* gt;       // concatenated (sum G, 5) float32
  const int* gt_off;     // [B + 1]
  const float* im_hw;    // [B][2] im_info[0][0], im_info[0][1]
};

__device__ __forceinline__ AnchorTargetWs anchor_ws_of(const AnchorBatchWs& w, int b, int N, int g_cap) {
  AnchorTargetWs r;
  r.inside_rank = w.inside_rank + (size_t)b * N;
  r.anchors_in = w.anchors_in + (size_t)b * N * 4;
  r.gt64 = w.gt64 + (size_t)b * g_cap * 4;
  r.argmax = w.argmax + (size_t)b * N;
  r.maxov = w.maxov + (size_t)b * N;
  r.colmax = w.colmax + (size_t)b * g_cap;
  r.labels = w.labels + (size_t)b * N;
  r.fg_list = w.fg_list + (size_t)b * N;
  r.bg_list = w.bg_list + (size_t)b * N;
  r.counts = w.counts + (size_t)b * 4;
  return r;
}

__device__ __forceinline__ void anchor_box(const AnchorBatchGeom& g, const AnchorTable& at, int p, double b[4]) {
  typedef Ar<double> A;
  const int a = p % g.A, cell = p / g.A;
  const int wx = cell % g.fw, hy = cell / g.fw;
  const double sx = (double)((long long)wx * g.feat_stride), sy = (double)((long long)hy * g.feat_stride);
  b[0] = A::add(at.a[a][0], sx);
  b[1] = A::add(at.a[a][1], sy);
  b[2] = A::add(at.a[a][2], sx);
  b[3] = A::add(at.a[a][3], sy);
}

// grid (csize, B), cluster (csize,1,1), block kAtThreads.  CTA `rank` owns the anchors [rank * items * 512, ...) in
// (h,w,a) order; thread t of it the anchors q0 + k * 512 + t, k < items.
__global__ void __launch_bounds__(kAtThreads, 1) anchor_phase1_kernel(const __grid_constant__ AnchorBatchGeom g,
                                                                      const __grid_constant__ AnchorTable at,
                                                                      const AnchorBatchWs wb) {
  namespace cgx = cooperative_groups;
  __shared__ int s_warp[33];
  __shared__ int s_tot[4];
  __shared__ double sq[kAtChunk][6];
  __shared__ unsigned long long s_colmax[kAtChunk];
  __shared__ float4 sq32[kAtChunk];
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank(), csize = (int)cluster.num_blocks();
  const int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
  const AnchorTargetWs w = anchor_ws_of(wb, b, g.N, g.g_cap);
  const int g0 = wb.gt_off[b], G = wb.gt_off[b + 1] - g0;
  const float* gt = wb.gt + (size_t)g0 * 5;
  const float im_h = wb.im_hw[b * 2], im_w = wb.im_hw[b * 2 + 1];
  const int q0 = rank * g.items * kAtThreads;
  // ---- :260-278 inside test, ordered ranks ----
  // im_info[k] + allowed_border is a float32 sum (NEP 50), compared against the float64 anchors
  const double lim_w = (double)__fadd_rn(im_w, (float)g.allowed_border);
  const double lim_h = (double)__fadd_rn(im_h, (float)g.allowed_border);
  const double lo = -(double)g.allowed_border;
  int rk[kAtItems];
  {
    int base = 0;
#pragma unroll
    for (int k = 0; k < kAtItems; ++k) {
      rk[k] = -1;
      if (k < g.items) {
        const int p = q0 + k * kAtThreads + tid;
        bool inside = false;
        if (p < g.N) {
          double bx[4];
          anchor_box(g, at, p, bx);
          inside = (bx[0] >= lo) && (bx[1] >= lo) && (bx[2] < lim_w) && (bx[3] < lim_h);
        }
        int total;
        const int r = block_exclusive_rank(inside, s_warp, &total);
        if (inside) rk[k] = base + r;
        base += total;
      }
    }
    if (tid == 0) s_tot[0] = base;
  }
  if (rank == 0) {
    for (int t = tid; t < G; t += kAtThreads) w.colmax[t] = 0.0;  // overlaps are >= 0
    for (int t = tid; t < G * 4; t += kAtThreads) w.gt64[t] = (double)gt[(t >> 2) * 5 + (t & 3)];
  }
  cluster.sync();
  int n_inside = 0;
  {
    // one remote read per lane (a cluster has at most 16 CTAs), then warp sums: one round trip instead of csize
    const int v = lane < csize ? *cluster.map_shared_rank(&s_tot[0], lane) : 0;
    const int before = __reduce_add_sync(0xffffffffu, lane < rank ? v : 0);
    n_inside = __reduce_add_sync(0xffffffffu, v);
#pragma unroll
    for (int k = 0; k < kAtItems; ++k) {
      if (k < g.items) {
        const int p = q0 + k * kAtThreads + tid;
        if (rk[k] >= 0) rk[k] += before;
        if (p < g.N) w.inside_rank[p] = rk[k];
        if (rk[k] >= 0) {
          double bx[4];
          anchor_box(g, at, p, bx);
          double* o = w.anchors_in + (size_t)rk[k] * 4;
          o[0] = bx[0];
          o[1] = bx[1];
          o[2] = bx[2];
          o[3] = bx[3];
        }
      }
    }
  }
  // From here on the work is spread over the INSIDE anchors in rank order (their boxes were just written in that
  // order): every thread of the cluster gets the same share, whatever part of the map its CTA scanned above.
  cluster.sync();  // anchors_in of every CTA is visible
  const int items_r = (n_inside + csize * kAtThreads - 1) / (csize * kAtThreads);  // <= g.items
  const int r0 = rank * items_r * kAtThreads;
#pragma unroll
  for (int k = 0; k < kAtItems; ++k) {
    const int r = r0 + k * kAtThreads + tid;
    rk[k] = (k < items_r && r < n_inside) ? r : -1;
  }
  // ---- :295-300 overlaps, argmax / max per anchor, max per gt; :307 anchors that tie a gt's maximum ----
  double best[kAtItems];
  int arg[kAtItems];
  unsigned tie = 0;  // bit k: anchor k equals the column maximum of some gt (quirk Q4: also a maximum of 0)
#pragma unroll
  for (int k = 0; k < kAtItems; ++k) {
    best[k] = 0.0;
    arg[k] = 0;
  }
  for (int j0 = 0; j0 < G; j0 += kAtChunk) {
    const int jn = min(kAtChunk, G - j0);
    __syncthreads();
    if (tid < jn) {
      const float4 q = make_float4(gt[(size_t)(j0 + tid) * 5], gt[(size_t)(j0 + tid) * 5 + 1], gt[(size_t)(j0 + tid) * 5 + 2],
                                   gt[(size_t)(j0 + tid) * 5 + 3]);
      const double x1 = (double)q.x, y1 = (double)q.y, x2 = (double)q.z, y2 = (double)q.w;
      sq[tid][0] = x1;
      sq[tid][1] = y1;
      sq[tid][2] = x2;
      sq[tid][3] = y2;
      sq[tid][4] = Ar<double>::mul(Ar<double>::add(Ar<double>::sub(x2, x1), 1.0), Ar<double>::add(Ar<double>::sub(y2, y1), 1.0));
      s_colmax[tid] = 0ull;
      sq32[tid] = q;  // the gt boxes ARE float32 (:295 casts them up)
    }
    __syncthreads();
    // Two passes per anchor.  Pass 1 finds the gts that may intersect it, in float32: iw > 0 means
    // min(x2) - max(x1) > -1, and a float32 difference is within 2^-23 (|a| + |b|) of the exact one (the anchor's own
    // conversion included), so "d > -1 - slack" never misses a pair.  Pass 2 walks only those (a handful per anchor)
    // through the reference's float64 arithmetic (overlap_one: exactly 0 for the pairs pass 1 let through in vain).
    // A pair can equal the gt's final maximum only if it was at least the CTA's running maximum when it was
    // computed: those few are remembered for the tie test below.
    unsigned long long cand[kAtItems];
#pragma unroll
    for (int k = 0; k < kAtItems; ++k) {
      cand[k] = 0ull;
      if (k < items_r && __any_sync(0xffffffffu, rk[k] >= 0)) {
        const bool live = rk[k] >= 0;
        double bx[4] = {0, 0, 0, 0};
        if (live) {
          const double2 lo2 = __ldcg(reinterpret_cast<const double2*>(w.anchors_in + (size_t)rk[k] * 4));
          const double2 hi2 = __ldcg(reinterpret_cast<const double2*>(w.anchors_in + (size_t)rk[k] * 4) + 1);
          bx[0] = lo2.x;
          bx[1] = lo2.y;
          bx[2] = hi2.x;
          bx[3] = hi2.y;
        }
        const float fx1 = (float)bx[0], fy1 = (float)bx[1], fx2 = (float)bx[2], fy2 = (float)bx[3];
        const float slack = 2.4e-7f;  // 2^-22
        unsigned long long m = 0ull;
#pragma unroll 8
        for (int j = 0; j < jn; ++j) {
          const float4 q = sq32[j];
          const float ax = fminf(fx2, q.z), bxm = fmaxf(fx1, q.x), ay = fminf(fy2, q.w), bym = fmaxf(fy1, q.y);
          const bool hx = (ax - bxm) > -1.0f - slack * (fabsf(ax) + fabsf(bxm) + 1.0f);
          const bool hy = (ay - bym) > -1.0f - slack * (fabsf(ay) + fabsf(bym) + 1.0f);
          m |= (unsigned long long)(hx && hy ? 1 : 0) << j;
        }
        if (!live) m = 0ull;
        while (m) {  // ascending j: argmax keeps the first maximum (a row of zeros keeps best 0, argmax 0)
          const int j = __ffsll((long long)m) - 1;
          m &= m - 1ull;
          // bbox_overlaps for this pair (box_kernels.cuh overlap_one, same operations in the same order), with the
          // division skipped when its result can matter neither as this anchor's new maximum nor for the gt's:
          // inter < x * union * (1 - 1e-15) puts the rounded quotient strictly below x.
          typedef Ar<double> D;
          const double iw = D::add(D::sub(fmin(bx[2], sq[j][2]), fmax(bx[0], sq[j][0])), 1.0);
          const double ih = D::add(D::sub(fmin(bx[3], sq[j][3]), fmax(bx[1], sq[j][1])), 1.0);
          if (!(iw > 0.0) || !(ih > 0.0)) continue;  // overlap 0: neither a new maximum (argmax keeps the first) nor a column maximum
          const double barea = D::mul(D::add(D::sub(bx[2], bx[0]), 1.0), D::add(D::sub(bx[3], bx[1]), 1.0));
          const double inter = D::mul(iw, ih);
          const double uni = D::sub(D::add(barea, sq[j][4]), inter);
          const unsigned long long seen = *reinterpret_cast<volatile unsigned long long*>(&s_colmax[j]);
          const double lim = fmin(best[k], __longlong_as_double((long long)seen));
          if (inter < lim * uni * 0.999999999999999) continue;
          const double ov = D::div(inter, uni);
          if (ov > best[k]) {
            best[k] = ov;
            arg[k] = j0 + j;
          }
          if (ov > 0.0) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(ov);  // ov > 0: bit order = value order
            if (bits >= seen) {
              cand[k] |= 1ull << j;
              if (bits > seen) atomicMax(&s_colmax[j], bits);
            }
          }
        }
      }
    }
    __syncthreads();
    if (tid < jn && s_colmax[tid] != 0ull) atomicMax(reinterpret_cast<unsigned long long*>(w.colmax) + j0 + tid, s_colmax[tid]);
    cluster.sync();  // the column maxima of this round are complete
    int zero_col = 0;
    if (tid < jn) {
      const double cm = __ldcg(w.colmax + j0 + tid);
      sq[tid][5] = cm;
      zero_col = cm == 0.0 ? 1 : 0;
    }
    zero_col = __syncthreads_or(zero_col);  // a gt no inside anchor touches: every anchor "equals" its maximum of 0
#pragma unroll
    for (int k = 0; k < kAtItems; ++k) {
      if (k < items_r && rk[k] >= 0) {
        if (zero_col) tie |= 1u << k;
        unsigned long long m = cand[k];
        if (m) {
          double bx[4];
          const double2 lo2 = __ldcg(reinterpret_cast<const double2*>(w.anchors_in + (size_t)rk[k] * 4));
          const double2 hi2 = __ldcg(reinterpret_cast<const double2*>(w.anchors_in + (size_t)rk[k] * 4) + 1);
          bx[0] = lo2.x;
          bx[1] = lo2.y;
          bx[2] = hi2.x;
          bx[3] = hi2.y;
          while (m) {
            const int j = __ffsll((long long)m) - 1;
            m &= m - 1ull;
            if (overlap_one(bx[0], bx[1], bx[2], bx[3], sq[j][0], sq[j][1], sq[j][2], sq[j][3], sq[j][4]) == sq[j][5]) tie |= 1u << k;
          }
        }
      }
    }
  }
  // ---- :289-316 labels; ordered fg / bg lists (np.where(labels == 1)[0], :320, :329) ----
  int lf[kAtItems], lb[kAtItems];
  {
    int nf = 0, nb = 0;
#pragma unroll
    for (int k = 0; k < kAtItems; ++k) {
      lf[k] = lb[k] = -1;
      if (k < items_r) {
        float label = -1.f;
        if (rk[k] >= 0) {
          if (G > 0) {
            const double mo = best[k];
            if (!g.clobber && mo < g.neg_ov) label = 0.f;  // :302-304
            if ((tie >> k) & 1u) label = 1.f;              // :307
            if (mo >= g.pos_ov) label = 1.f;               // :310
            if (g.clobber && mo < g.neg_ov) label = 0.f;   // :312-314
          } else {
            label = 0.f;                                   // :316
          }
          w.labels[rk[k]] = label;
          w.argmax[rk[k]] = arg[k];
          w.maxov[rk[k]] = best[k];
        }
        int total;
        int r = block_exclusive_rank(label == 1.f, s_warp, &total);
        if (label == 1.f) lf[k] = nf + r;
        nf += total;
        r = block_exclusive_rank(label == 0.f, s_warp, &total);
        if (label == 0.f) lb[k] = nb + r;
        nb += total;
      }
    }
    if (tid == 0) {
      s_tot[1] = nf;
      s_tot[2] = nb;
    }
  }
  cluster.sync();
  {
    const int vf = lane < csize ? *cluster.map_shared_rank(&s_tot[1], lane) : 0;
    const int vb = lane < csize ? *cluster.map_shared_rank(&s_tot[2], lane) : 0;
    const int fbase = __reduce_add_sync(0xffffffffu, lane < rank ? vf : 0);
    const int bbase = __reduce_add_sync(0xffffffffu, lane < rank ? vb : 0);
    const int n_fg = __reduce_add_sync(0xffffffffu, vf), n_bg = __reduce_add_sync(0xffffffffu, vb);
#pragma unroll
    for (int k = 0; k < kAtItems; ++k) {
      if (lf[k] >= 0) w.fg_list[fbase + lf[k]] = rk[k];
      if (lb[k] >= 0) w.bg_list[bbase + lb[k]] = rk[k];
    }
    if (rank == 0 && tid == 0) {
      w.counts[0] = n_inside;
      w.counts[1] = n_fg;
      w.counts[2] = n_bg;
    }
  }
  cluster.sync();  // nobody leaves while its totals may still be read by a neighbour
}

struct AnchorPhase2 {
  const int* dis_counts;   // [B][2] fg / bg ranks to disable
  const int* dis_off;      // [B] start of image b's ranks (fg first, then bg) in dis_ranks
  const int* dis_ranks;
  const float* weights;    // [B][2] positive / negative weight (:343-354, computed by the caller)
  float4 inside_w;
  float* label;            // (B, N)
  float* target;           // (B, 4A, H, W)
  float* inside;
  float* outside;
};

// grid (csize, B), cluster (csize,1,1), block kAtThreads.
__global__ void __launch_bounds__(kAtThreads, 1) anchor_phase2_kernel(const __grid_constant__ AnchorBatchGeom g, const AnchorBatchWs wb,
                                                                      const AnchorPhase2 a) {
  namespace cgx = cooperative_groups;
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank(), csize = (int)cluster.num_blocks();
  const int b = blockIdx.y;
  const AnchorTargetWs w = anchor_ws_of(wb, b, g.N, g.g_cap);
  const int g0 = wb.gt_off[b], G = wb.gt_off[b + 1] - g0;
  const float* gt = wb.gt + (size_t)g0 * 5;
  // :321-325, :330-334  labels[list[rank]] = -1 for the ranks the host drew
  {
    const int n_fg = a.dis_counts[b * 2], n_bg = a.dis_counts[b * 2 + 1];
    const int* ranks = a.dis_ranks + a.dis_off[b];
    for (int t = rank * kAtThreads + threadIdx.x; t < n_fg + n_bg; t += csize * kAtThreads) {
      if (t < n_fg) w.labels[w.fg_list[ranks[t]]] = -1.f;
      else w.labels[w.bg_list[ranks[t]]] = -1.f;
    }
  }
  cluster.sync();
  // :336-385 targets for every inside anchor, weights, _unmap to all N anchors and the NCHW transposes, written
  // directly in the output layouts: one thread per output anchor (a,h,w)
  const float pos_w = a.weights[b * 2], neg_w = a.weights[b * 2 + 1];
  const int hw = g.fh * g.fw;
  float* label = a.label + (size_t)b * g.N;
  float* target = a.target + (size_t)b * 4 * g.N;
  float* inside = a.inside + (size_t)b * 4 * g.N;
  float* outside = a.outside + (size_t)b * 4 * g.N;
  for (int t = rank * kAtThreads + threadIdx.x; t < g.N; t += csize * kAtThreads) {
    const int an = t / hw, cell = t - an * hw;
    const int p = cell * g.A + an;
    const int r = w.inside_rank[p];
    float lab = -1.f;
    float tg[4] = {0.f, 0.f, 0.f, 0.f}, iw[4] = {0.f, 0.f, 0.f, 0.f}, ow[4] = {0.f, 0.f, 0.f, 0.f};
    if (r >= 0) {
      lab = __ldcg(w.labels + r);
      if (G > 0) {
        const double* e = w.anchors_in + (size_t)r * 4;
        const float* q = gt + (size_t)w.argmax[r] * 5;  // gt_boxes[argmax_overlaps, :4] stays float32 (:338)
        double o[4];
        bbox_transform_one<double, float, double>(e[0], e[1], e[2], e[3], q[0], q[1], q[2], q[3], o);
        for (int k = 0; k < 4; ++k) tg[k] = (float)o[k];
      }
      if (lab == 1.f) {
        iw[0] = a.inside_w.x;
        iw[1] = a.inside_w.y;
        iw[2] = a.inside_w.z;
        iw[3] = a.inside_w.w;
        for (int k = 0; k < 4; ++k) ow[k] = pos_w;
      } else if (lab == 0.f) {
        for (int k = 0; k < 4; ++k) ow[k] = neg_w;
      }
    }
    label[t] = lab;
    for (int k = 0; k < 4; ++k) {
      const size_t o = (size_t)(an * 4 + k) * hw + cell;
      target[o] = tg[k];
      inside[o] = iw[k];
      outside[o] = ow[k];
    }
  }
}

}  // namespace mxr
