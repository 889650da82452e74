// dibr_softmask.cu -- DIB-R soft mask forward / backward (north_star (c), (d)).
//
// Semantics (the reference's code, not its prose): for every pixel NOT covered by the
// rasterizer, walk the faces in ASCENDING INDEX order, keep the first `knum` whose
// enlarged box contains the pixel centre (half-open test), and for each take
//   d^2  = min over 3 edges (perpendicular foot inside the edge) and 3 vertices,
//   prob = exp(-sigmainv * d^2 / mult / mult);   soft_mask = 1 - prod(1 - prob)
// (dibr_soft_mask_cuda.cu:80-182).  Covered pixels get 1.
//
// Mapping: one CTA per screen block, one WARP per 16x16 tile, one pixel per lane in each
// of the tile's eight 8x4 regions.  The bin's face list (ascending, from dibr_binning.cu)
// is staged through shared memory in chunks; each warp compacts the chunk to the faces
// touching its tile with __ballot_sync/__popc (order preserved), then per region tests
// 32 faces per instruction against the region rectangle and only the survivors are
// evaluated per pixel.  A region stops as soon as all its uncovered pixels hold `knum`
// faces; the CTA stops when every region has.
//
// Backward re-walks the same lists (nothing but soft_mask itself is saved by the fast
// path: the reference keeps 13*knum bytes per pixel), sums each face's contributions
// over the warp with shuffles, accumulates them per staged face in shared memory and
// flushes one red.global.add per touched (face, coordinate) per chunk.
//
// Arithmetic follows the reference's float/double mix literally (its bare literals
// EPS = 1e-7 and 1.0 are doubles), with the FMA shapes nvcc emits for the reference
// (decoded from its sm_100a SASS) written as explicit intrinsics.
#include "dibr_common.cuh"

namespace dibr {

#define DIBR_SOFT_EPS 1e-7


constexpr int kSoftChunk = 256;

// Correctly rounded double quotient a / b from the correctly rounded reciprocal
// inv = RN(1 / b) (Markstein): q0 = RN(a * inv), r = a - b * q0 (exact, FMA),
// q = RN(q0 + r * inv).  Three FP64 ops instead of a ~25-instruction division; the
// reciprocal is taken once per face edge when the face is staged.
__device__ __forceinline__ double div_by(double a, double b, double inv) {
  const double q0 = __dmul_rn(a, inv);
  const double r = __fma_rn(-b, q0, a);
  return __fma_rn(r, inv, q0);
}

// prob and nearest-feature id (0-2 edge, 3-5 vertex) of one (pixel, face) pair
// (dibr_soft_mask_cuda.cu:103-163).  inv[i] = RN(1 / ((double)down_i + EPS)).
__device__ __forceinline__ void soft_eval(float x0, float y0, const float (&X)[6],
                                          const double (&inv)[3], float four_mm, float sigmainv,
                                          float mult, float& prob, int& edgeid) {
  float pdis[6];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const float x1 = X[2 * i], y1 = X[2 * i + 1];
    const float x2 = X[2 * ((i + 1) % 3)], y2 = X[2 * ((i + 1) % 3) + 1];
    const float A = __fsub_rn(y2, y1), B = __fsub_rn(x1, x2);
    const float C = __fmaf_rn(x2, y1, -__fmul_rn(x1, y2));
    const float AA = __fmul_rn(A, A), BB = __fmul_rn(B, B), AB = __fmul_rn(A, B);
    const float down = __fadd_rn(AA, BB);
    const float up = __fadd_rn(__fmaf_rn(B, y0, __fmul_rn(A, x0)), C);
    const float x3n = __fmaf_rn(-A, C, __fmaf_rn(BB, x0, -__fmul_rn(AB, y0)));
    const float y3n = __fmaf_rn(-B, C, __fmaf_rn(AA, y0, -__fmul_rn(AB, x0)));
    const double dd = __dadd_rn((double)down, DIBR_SOFT_EPS);
    const float x3 = __double2float_rn(div_by((double)x3n, dd, inv[i]));
    const float y3 = __double2float_rn(div_by((double)y3n, dd, inv[i]));
    const float direct = __fmaf_rn(__fsub_rn(x3, x1), __fsub_rn(x3, x2),
                                   __fmul_rn(__fsub_rn(y3, y1), __fsub_rn(y3, y2)));
    if (direct > 0.f)
      pdis[i] = four_mm;
    else
      pdis[i] = __double2float_rn(div_by((double)__fmul_rn(up, up), dd, inv[i]));
    const float dx = __fsub_rn(x0, x1), dy = __fsub_rn(y0, y1);
    pdis[i + 3] = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
  }
  edgeid = 0;
  float dis = pdis[0];
#pragma unroll
  for (int i = 1; i < 6; i++)
    if (dis > pdis[i]) { dis = pdis[i]; edgeid = i; }
  const float z = __fdiv_rn(__fdiv_rn(__fmul_rn(sigmainv, dis), mult), mult);
  prob = expf(-z);
}

// RN(1 / ((double)(A*A + B*B) + EPS)) of edge i of a staged face.
__device__ __forceinline__ double edge_inv(float x1, float y1, float x2, float y2) {
  const float A = __fsub_rn(y2, y1), B = __fsub_rn(x1, x2);
  const float down = __fadd_rn(__fmul_rn(A, A), __fmul_rn(B, B));
  return __drcp_rn(__dadd_rn((double)down, DIBR_SOFT_EPS));
}

__device__ __forceinline__ float pick3(int v, float a, float b, float c) {
  return v == 0 ? a : (v == 1 ? b : c);
}
__device__ __forceinline__ int pick3i(int v, int a, int b, int c) {
  return v == 0 ? a : (v == 1 ? b : c);
}
__device__ __forceinline__ double pick3d(int v, double a, double b, double c) {
  return v == 0 ? a : (v == 1 ? b : c);
}

// dibr_soft_mask_cuda.cu:283-347: gradient of one (pixel, face) pair w.r.t. the face's
// six UNscaled coordinates, added into g[6] (register array: static indices only).
// Core: the gradient lands on vertex v1 (always) and v2 (edge case, else -1).
// Evaluated in double from the saved state (pixel centre, scaled coordinates, prob,
// soft_mask): what the reference's kernel computes when instantiated for double on that
// state.  Its float instantiation loses up to 1e-2 of a face's gradient to float
// differences; the pair count is small enough for FP64 not to show in the time.
__device__ __forceinline__ void soft_grad_core_sel(double x0, double y0, const double (&X)[6],
                                                 int edgeid, double prob, double allprob,
                                                 double dLdp, double sigmainv, double mult,
                                                 int& v1, double& g1x, double& g1y, int& v2,
                                                 double& g2x, double& g2y) {
  const double im = 1.0 / mult;
  const double dLdz = -1.0 * sigmainv * dLdp * (1.0 - allprob) / (1.0 - prob + DIBR_SOFT_EPS) * prob;
  v2 = -1;
  g2x = 0.0;
  g2y = 0.0;
  if (edgeid >= 3) {
    v1 = edgeid - 3;
    const double x1 = pick3d(v1, X[0], X[2], X[4]), y1 = pick3d(v1, X[1], X[3], X[5]);
    g1x = dLdz * 2.0 * (x1 - x0) * im;
    g1y = dLdz * 2.0 * (y1 - y0) * im;
  } else {
    v1 = edgeid;
    v2 = (edgeid + 1) % 3;
    const double x1 = pick3d(v1, X[0], X[2], X[4]), y1 = pick3d(v1, X[1], X[3], X[5]);
    const double x2 = pick3d(v2, X[0], X[2], X[4]), y2 = pick3d(v2, X[1], X[3], X[5]);
    const double A = y2 - y1, B = x1 - x2, C = x2 * y1 - x1 * y2;
    const double up = A * x0 + B * y0 + C;
    const double down = A * A + B * B;
    const double id = 1.0 / (down + DIBR_SOFT_EPS);   // the four "/ (down + EPS)" (:318-330)
    const double dissquare = up * up * id;
    const double dzdA = 2.0 * (x0 * up - dissquare * A) * id;
    const double dzdB = 2.0 * (y0 * up - dissquare * B) * id;
    const double dzdC = 2.0 * up * id;
    g1x = dLdz * (dzdB - y2 * dzdC) * im;
    g1y = dLdz * (x2 * dzdC - dzdA) * im;
    g2x = dLdz * (y1 * dzdC - dzdB) * im;
    g2y = dLdz * (dzdA - x1 * dzdC) * im;
  }
}

__device__ __forceinline__ void soft_grad_core(float x0f, float y0f, const float (&X)[6], int edgeid,
                                               float prob, float allprob, float dLdp,
                                               float sigmainv, float mult, double inv_e, int& v1,
                                               float& g1x, float& g1y, int& v2, float& g2x,
                                               float& g2y) {
  (void)inv_e;
  // X[2 * v] with a runtime v would push the array to local memory: select instead
  const double Xd[6] = {X[0], X[1], X[2], X[3], X[4], X[5]};
  double a1, b1, a2, b2;
  soft_grad_core_sel(x0f, y0f, Xd, edgeid, prob, allprob, dLdp, sigmainv, mult, v1, a1, b1, v2, a2, b2);
  g1x = (float)a1; g1y = (float)b1; g2x = (float)a2; g2y = (float)b2;
}

__device__ __forceinline__ void soft_grad(float x0, float y0, const float (&X)[6], int edgeid,
                                          float prob, float allprob, float dLdp, float sigmainv,
                                          float mult, double inv_e, float (&g)[6]) {
  int v1, v2;
  float g1x, g1y, g2x, g2y;
  soft_grad_core(x0, y0, X, edgeid, prob, allprob, dLdp, sigmainv, mult, inv_e, v1, g1x, g1y, v2,
                 g2x, g2y);
#pragma unroll
  for (int v = 0; v < 3; v++) {
    if (v == v1) { g[2 * v] += g1x; g[2 * v + 1] += g1y; }
    if (v == v2) { g[2 * v] += g2x; g[2 * v + 1] += g2y; }
  }
}

// ---------------------------------------------------------------------------
// Block kernel.  8 warps per CTA; the block's 8x4 regions are handed to warps through a
// shared-memory queue, so warps stay busy whatever part of the block is uncovered.
// Per (chunk of staged faces, region) a warp works in two phases:
//   1. gather: walk the chunk in face order; 32 faces per instruction are tested against
//      the region rectangle (__ballot_sync), survivors are tested per pixel with the
//      reference's half-open box test and recorded in the pixel's hit segment until the
//      pixel holds knum faces (order = ascending face index, the reference's rule);
//   2. evaluate: the (pixel, face) pairs gathered by the whole region are dealt out
//      evenly over the 32 lanes, whatever pixel they belong to, so the expensive
//      distance/exp (or gradient) arithmetic runs with every lane busy; each pixel then
//      folds its probabilities in order.
// ---------------------------------------------------------------------------
constexpr int kSoftWarps = 8;
constexpr int kSoftThreads = kSoftWarps * 32;
constexpr int kSeg = 16;  // hit-segment length per pixel between two evaluate phases

template <int BLK_LOG2, int MODE>
struct SoftSmem {
  static constexpr int BLK = 1 << BLK_LOG2;
  static constexpr int NPX = BLK * BLK;
  static constexpr int NR = NPX / 32;
  double inv[3][kSoftChunk];
  float x[6][kSoftChunk];
  float bb[4][kSoftChunk];
  int face[kSoftChunk];
  float ap[NPX];
  unsigned short cnt[NPX];       // faces held; 0xffff = covered / outside the image
  float pr[kSoftWarps][kSeg][32];
  unsigned short hit[kSoftWarps][kSeg][32];
  unsigned char ty[kSoftWarps][kSeg][32];
  int off[kSoftWarps][33];
  unsigned char ropen[NR];
  int next;
  float acc[MODE == SOFT_BWD ? kSoftChunk : 1][6];
};

template <int BLK_LOG2, int MODE>
__device__ __forceinline__ void soft_block(const SoftArgs& a, const Plan& p,
                                           SoftSmem<BLK_LOG2, MODE>& sm, const int bxi,
                                           const int byi, const int b) {
  using Smem = SoftSmem<BLK_LOG2, MODE>;
  constexpr int BLK = Smem::BLK, NPX = Smem::NPX, NR = Smem::NR, RX = BLK / 8;
  constexpr int CH = kSoftChunk;
  const int tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;
  const int bx0 = bxi << BLK_LOG2, by0 = byi << BLK_LOG2;
  const Grid g = plan_grid(p, a.job.mult);
  const float mult = a.job.mult;
  const float four_mm = __fmul_rn(__fmul_rn(mult, 4.f), mult);
  const int K = min(a.K, 0xfffe);
  const int lx = lane & 7, ly = lane >> 3;

  // ---- pixel state
  for (int i = tid; i < NPX; i += kSoftThreads) {
    const int ix = bx0 + (i & (BLK - 1)), iy = by0 + (i >> BLK_LOG2);
    bool uncovered = false;
    if (ix < p.W && iy < p.H) uncovered = a.sel[((long long)b * p.H + iy) * p.W + ix] < 0;
    sm.cnt[i] = uncovered ? 0 : 0xffff;
    sm.ap[i] = 1.0f;
  }
  __syncthreads();
  for (int r = warp; r < NR; r += kSoftWarps) {
    const int pl = (((r / RX) * 4 + ly) << BLK_LOG2) + (r % RX) * 8 + lx;
    const bool open = (int)sm.cnt[pl] < K;
    const unsigned m = __ballot_sync(DIBR_FULL_MASK, open);
    if (lane == 0) sm.ropen[r] = m != 0;
  }

  const int shift = p.bin_log2 - BLK_LOG2;
  const int bin = (byi >> shift) * p.nbx + (bxi >> shift);
  const long long row = (long long)b * p.NB + bin;
  const long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);

  for (long long base = 0; base < n; base += CH) {
    __syncthreads();
    bool mine = false;
    for (int r = tid; r < NR; r += kSoftThreads) mine |= sm.ropen[r] != 0;
    if (tid == 0) sm.next = 0;
    if (!__syncthreads_or(mine)) break;
    const int nc = (int)min((long long)CH, n - base);
    // ---- stage the chunk: scaled coordinates, enlarged box, edge reciprocals
    for (int t = tid; t < nc; t += kSoftThreads) {
      const int f = a.job.list[start + base + t];
      Face2D fc;
      load_face2d(a.job, b, f, fc);
      sm.x[0][t] = fc.ax; sm.x[1][t] = fc.ay; sm.x[2][t] = fc.bx;
      sm.x[3][t] = fc.by; sm.x[4][t] = fc.cx; sm.x[5][t] = fc.cy;
      sm.bb[0][t] = fc.xmin; sm.bb[1][t] = fc.ymin; sm.bb[2][t] = fc.xmax; sm.bb[3][t] = fc.ymax;
      sm.inv[0][t] = edge_inv(fc.ax, fc.ay, fc.bx, fc.by);
      sm.inv[1][t] = edge_inv(fc.bx, fc.by, fc.cx, fc.cy);
      sm.inv[2][t] = edge_inv(fc.cx, fc.cy, fc.ax, fc.ay);
      sm.face[t] = f;
      if (MODE == SOFT_BWD) {
#pragma unroll
        for (int i = 0; i < 6; i++) sm.acc[t][i] = 0.f;
      }
    }
    __syncthreads();

    while (true) {
      int r = 0;
      if (lane == 0) r = atomicAdd(&sm.next, 1);
      r = __shfl_sync(DIBR_FULL_MASK, r, 0);
      if (r >= NR) break;
      if (!sm.ropen[r]) continue;
      const int rx0 = bx0 + (r % RX) * 8, ry0 = by0 + (r / RX) * 4;
      const int ix = rx0 + lx, iy = ry0 + ly;
      const int pl = ((iy - by0) << BLK_LOG2) + (ix - bx0);
      const float r_xlo = pix_x(g, rx0), r_xhi = pix_x(g, min(rx0 + 7, p.W - 1));
      const float r_yhi = pix_y(g, ry0), r_ylo = pix_y(g, min(ry0 + 3, p.H - 1));
      const float x0 = pix_x(g, ix), y0 = pix_y(g, iy);
      const long long pix = ((long long)b * p.H + min(iy, p.H - 1)) * p.W + min(ix, p.W - 1);
      int cnt = sm.cnt[pl];
      float ap = sm.ap[pl];
      bool open = cnt < K;
      int nbuf = 0;

      // evaluate everything gathered so far for this region (all lanes call it together)
      auto flush = [&]() {
        int incl = nbuf;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
          if (lane >= d) incl += t;
        }
        const int total = __shfl_sync(DIBR_FULL_MASK, incl, 31);
        sm.off[warp][lane] = incl - nbuf;
        if (lane == 31) sm.off[warp][32] = total;
        __syncwarp();
        for (int i = lane; i < total; i += 32) {
          int l = 0;  // owner pixel: the largest l with off[l] <= i
#pragma unroll
          for (int st = 16; st > 0; st >>= 1)
            if (sm.off[warp][l + st] <= i) l += st;
          const int j = i - sm.off[warp][l];
          const int s = sm.hit[warp][j][l];
          const int pix_x_i = rx0 + (l & 7), pix_y_i = ry0 + (l >> 3);
          const float px = pix_x(g, pix_x_i), py = pix_y(g, pix_y_i);
          const float X[6] = {sm.x[0][s], sm.x[1][s], sm.x[2][s], sm.x[3][s], sm.x[4][s], sm.x[5][s]};
          const double inv[3] = {sm.inv[0][s], sm.inv[1][s], sm.inv[2][s]};
          float prob;
          int edgeid;
          soft_eval(px, py, X, inv, four_mm, a.sigmainv, mult, prob, edgeid);
          if (MODE == SOFT_BWD) {
            const long long q = ((long long)b * p.H + pix_y_i) * p.W + pix_x_i;
            float gv[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            soft_grad(px, py, X, edgeid, prob, a.mask_in[q], a.grad_mask[q], a.sigmainv, mult,
                      edgeid < 3 ? (edgeid == 0 ? inv[0] : (edgeid == 1 ? inv[1] : inv[2])) : 0.0,
                      gv);
#pragma unroll
            for (int k = 0; k < 6; k++)
              if (gv[k] != 0.f) atomicAdd(&sm.acc[s][k], gv[k]);
          } else {
            sm.pr[warp][j][l] = prob;
            sm.ty[warp][j][l] = (unsigned char)(edgeid + 1);
          }
        }
        __syncwarp();
        if (MODE != SOFT_BWD) {
          for (int j = 0; j < nbuf; j++) {
            const float prob = sm.pr[warp][j][lane];
            ap = __double2float_rn(__dmul_rn((double)ap, __dsub_rn(1.0, (double)prob)));  // :176-177
            if (MODE == SOFT_FWD_DENSE) {
              const long long o = pix * a.K + (cnt - nbuf + j);
              a.prob[o] = prob;
              a.cidx[o] = sm.face[sm.hit[warp][j][lane]];
              a.type[o] = sm.ty[warp][j][lane];
            }
          }
        }
        nbuf = 0;
        __syncwarp();
      };

      for (int j0 = 0; j0 < nc; j0 += 32) {
        const int t = j0 + lane;
        bool cand = false;
        if (t < nc)
          cand = sm.bb[0][t] <= r_xhi && sm.bb[2][t] > r_xlo && sm.bb[1][t] <= r_yhi &&
                 sm.bb[3][t] > r_ylo;
        unsigned m = __ballot_sync(DIBR_FULL_MASK, cand);
        while (m) {
          const int s = j0 + __ffs(m) - 1;
          m &= m - 1;
          const bool in = open && !(x0 < sm.bb[0][s] || x0 >= sm.bb[2][s] || y0 < sm.bb[1][s] ||
                                    y0 >= sm.bb[3][s]);  // :95-97
          if (in) {
            sm.hit[warp][nbuf][lane] = (unsigned short)s;
            nbuf++;
            if (++cnt >= K) open = false;  // :170-171
          }
          if (__any_sync(DIBR_FULL_MASK, nbuf == kSeg)) flush();
          if (!__any_sync(DIBR_FULL_MASK, open)) { m = 0; j0 = nc; }
        }
      }
      if (__any_sync(DIBR_FULL_MASK, nbuf > 0)) flush();
      sm.cnt[pl] = (unsigned short)cnt;
      sm.ap[pl] = ap;
      const unsigned still = __ballot_sync(DIBR_FULL_MASK, open);
      if (lane == 0) sm.ropen[r] = still != 0;
    }
    if (MODE == SOFT_BWD) {
      __syncthreads();
      for (int t = tid; t < nc * 6; t += kSoftThreads) {
        const float v = sm.acc[t / 6][t % 6];
        if (v == 0.f) continue;
        if (a.job.faces) {   // mesh mode: vertex gradients (B, V, 2)
          int vid[3];
          mesh_face(a.job, sm.face[t / 6], vid);
          const int vv = pick3i((t % 6) >> 1, vid[0], vid[1], vid[2]);
          atomicAdd(a.grad_img + ((long long)b * a.job.V + vv) * 2 + (t & 1), v);
        } else {
          atomicAdd(a.grad_img + ((long long)b * a.job.view_stride + sm.face[t / 6]) * 6 + t % 6, v);
        }
      }
    }
  }

  if (MODE != SOFT_BWD) {
    __syncthreads();
    for (int i = tid; i < NPX; i += kSoftThreads) {
      const int ix = bx0 + (i & (BLK - 1)), iy = by0 + (i >> BLK_LOG2);
      if (ix < p.W && iy < p.H)
        a.mask[((long long)b * p.H + iy) * p.W + ix] =
            sm.cnt[i] == 0xffff ? 1.0f                                            // :68-70
                                : __double2float_rn(__dsub_rn(1.0, (double)sm.ap[i]));  // :181-182
    }
  }
}

// Screen blocks are taken from a 1-D grid with a stride loop: when the kernel is only the
// overflow fallback of the pair path (the usual case: nothing to do) it is launched with a
// few CTAs per SM that read the pair cursors and leave.
template <int BLK_LOG2, int MODE>
__global__ void __launch_bounds__(kSoftThreads)
softmask_kernel(const SoftArgs a, const Plan p) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  SoftSmem<BLK_LOG2, MODE>& sm = *reinterpret_cast<SoftSmem<BLK_LOG2, MODE>*>(s_raw);
  pdl_prologue();
  if ((MODE == SOFT_FWD || MODE == SOFT_BWD) && a.pairs.cap > 0 && !pairs_overflowed(a.pairs))
    return;
  const int per_view = p.nkx * p.nky;
  const long long nblk = (long long)per_view * p.B;
  for (long long blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
    const int b = (int)(blk / per_view), rr = (int)(blk - (long long)b * per_view);
    soft_block<BLK_LOG2, MODE>(a, p, sm, rr % p.nkx, rr / p.nkx, b);
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// Pair kernels.  SOFT_GATHER left, for every active region, the first-knum face ids of
// each pixel in the pair buffer, and cut every region's pairs into chunks of kPairChunk.
// Evaluation and gradient run with ONE WARP PER CHUNK, i.e. spread over the whole GPU
// however unevenly the silhouette is distributed over the screen (the pairs live in a
// few thousand regions of very different weight).
// ---------------------------------------------------------------------------
#ifndef DIBR_PAIR_WARPS
#define DIBR_PAIR_WARPS 8
#endif
constexpr int kPairWarps = DIBR_PAIR_WARPS;

__device__ __forceinline__ void load_scaled_coords(const BinJob& j, int b, int f, float (&X)[6]) {
  float2 v0, v1, v2;
  face_corners(j, b, f, v0, v1, v2);   // pair faces come out of the list: ids are valid
  if (!j.prescaled) {  // fl(v * mult), as torch does at dibr.py:32
    v0.x = __fmul_rn(v0.x, j.mult); v0.y = __fmul_rn(v0.y, j.mult);
    v1.x = __fmul_rn(v1.x, j.mult); v1.y = __fmul_rn(v1.y, j.mult);
    v2.x = __fmul_rn(v2.x, j.mult); v2.y = __fmul_rn(v2.y, j.mult);
  }
  X[0] = v0.x; X[1] = v0.y; X[2] = v1.x; X[3] = v1.y; X[4] = v2.x; X[5] = v2.y;
}

// Decodes the active region `act` of the pair buffer for this lane.  Returns false (for
// the whole warp) when there is nothing to do.
__device__ __forceinline__ bool pair_region(const PairBuf& pb, const Plan& p, long long act,
                                            int lane, int& b, int& ix, int& iy, int& cnt) {
  if (pairs_overflowed(pb)) return false;
  if (act >= (long long)pb.cursor[1]) return false;
  const int nrx = (p.W + 7) >> 3, nry = (p.H + 3) >> 2;
  const int rid = pb.act_region[act];
  b = rid / (nrx * nry);
  const int rr = rid - b * nrx * nry;
  ix = (rr % nrx) * 8 + (lane & 7);
  iy = (rr / nrx) * 4 + (lane >> 3);
  cnt = pb.act_cnt[act * 32 + lane];
  return true;
}

// The (pixel, hit) pairs of a region are dealt out evenly over the 32 lanes whatever
// pixel they belong to (partially covered regions would otherwise leave most lanes idle):
// pair i belongs to the pixel lane l with off[l] <= i < off[l + 1], off = exclusive scan of
// the per-pixel hit counts, and is that pixel's hit h = i - off[l].
struct PairDeal {
  int off;     // exclusive prefix of cnt over lanes (this lane's first pair)
  int total;   // pairs of the region
  int segreg;  // lane q holds the region's segment q (q < 32)
};

__device__ __forceinline__ PairDeal pair_deal(const PairBuf& pb, long long act, int cnt, int lane) {
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
    if (lane >= d) incl += t;
  }
  PairDeal pd;
  pd.off = incl - cnt;
  pd.total = __shfl_sync(DIBR_FULL_MASK, incl, 31);
  pd.segreg = lane < pb.nseg ? pb.act_seg[act * pb.nseg + lane] : -1;
  return pd;
}

// (owner lane, hit index, slot) of pair i; all lanes call it, i may differ per lane
__device__ __forceinline__ long long pair_locate(const PairBuf& pb, long long act,
                                                 const PairDeal& pd, int i, int& l, int& h) {
  l = 0;
#pragma unroll
  for (int st = 16; st > 0; st >>= 1) {
    const int o = __shfl_sync(DIBR_FULL_MASK, pd.off, l + st);
    if (o <= i) l += st;
  }
  h = i - __shfl_sync(DIBR_FULL_MASK, pd.off, l);
  const int q = h / kPairSegHits;
  int sg = __shfl_sync(DIBR_FULL_MASK, pd.segreg, q & 31);
  if (q >= 32) sg = pb.act_seg[act * pb.nseg + q];
  return (long long)sg * kPairSegSlots + (h % kPairSegHits) * 32 + l;
}

// Work item of both kernels: one chunk (kPairChunk pairs, dealt order) of one active region,
// taken by one warp of a persistent grid.  Returns false when the chunk list is exhausted.
struct PairWork {
  long long act;
  int b, ix, iy, cnt, i_begin, i_end;
  PairDeal pd;
};

__device__ __forceinline__ bool pair_work(const PairBuf& pb, const Plan& p, unsigned w, int lane,
                                          PairWork& k) {
  const int code = pb.chunk[w];
  k.act = code >> 8;
  if (!pair_region(pb, p, k.act, lane, k.b, k.ix, k.iy, k.cnt)) return false;
  k.pd = pair_deal(pb, k.act, k.cnt, lane);
  k.i_begin = (code & 255) * kPairChunk;
  k.i_end = min(k.pd.total, k.i_begin + kPairChunk);
  return true;
}

#ifndef DIBR_EVAL_MINB
#define DIBR_EVAL_MINB 4
#endif
__global__ void __launch_bounds__(kPairWarps * 32, DIBR_EVAL_MINB)
softmask_pairs_eval_kernel(const SoftArgs a, const Plan p) {
  __shared__ float sX[kPairWarps][6][kPairChunk];
  pdl_prologue();
  const PairBuf& pb = a.pairs;
  if (a.report && (blockIdx.x | threadIdx.x) == 0) {   // pair-buffer needs, see BinJobs::report
    volatile unsigned long long* r = a.report;
    r[2] = pb.cursor[0];
    r[3] = pb.cursor[1];
    __threadfence_system();
  }
  if (pairs_overflowed(pb)) return;
  const int lane = lane_id();
  const unsigned nchunk = min(*pb.nchunk, 3u * (unsigned)pb.cap);
  const Grid g = plan_grid(p, a.job.mult);
  const float mult = a.job.mult;
  const float four_mm = __fmul_rn(__fmul_rn(mult, 4.f), mult);
  for (unsigned w = blockIdx.x * kPairWarps + (threadIdx.x >> 5); w < nchunk;
       w += gridDim.x * kPairWarps) {
    PairWork k;
    if (!pair_work(pb, p, w, lane, k)) continue;
    const int rx0 = k.ix - (lane & 7), ry0 = k.iy - (lane >> 3);
    // ---- evaluate, all lanes busy.  The kernel is bound by the latency of its dependent
    // loads (face id -> coordinates), so all of the chunk's loads are issued back to back
    // and parked in shared memory (each lane reads back only what it wrote) before any
    // arithmetic; lanes past the end redo the chunk's last pair and skip the stores.
    constexpr int U = kPairChunk / 32;
    unsigned slot[U];
    int lpack = 0;
    {
      int f[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int i = k.i_begin + 32 * u + lane;
        int l, h;
        slot[u] = (unsigned)pair_locate(pb, k.act, k.pd, min(i, k.i_end - 1), l, h);
        lpack |= l << (8 * u);
        f[u] = __ldcg(pb.face + slot[u]);
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        float X[6];
        load_scaled_coords(a.job, k.b, f[u], X);
#pragma unroll
        for (int c = 0; c < 6; c++) sX[threadIdx.x >> 5][c][32 * u + lane] = X[c];
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (k.i_begin + 32 * u >= k.i_end) break;
      const int l = (lpack >> (8 * u)) & 31;
      float X[6];
#pragma unroll
      for (int c = 0; c < 6; c++) X[c] = sX[threadIdx.x >> 5][c][32 * u + lane];
      const float x0 = pix_x(g, rx0 + (l & 7)), y0 = pix_y(g, ry0 + (l >> 3));
      const double inv[3] = {edge_inv(X[0], X[1], X[2], X[3]), edge_inv(X[2], X[3], X[4], X[5]),
                             edge_inv(X[4], X[5], X[0], X[1])};
      float prob;
      int edgeid;
      soft_eval(x0, y0, X, inv, four_mm, a.sigmainv, mult, prob, edgeid);
      if (k.i_begin + 32 * u + lane < k.i_end) {
        pb.prob[slot[u]] = prob;
        pb.type[slot[u]] = (unsigned char)(edgeid + 1);
      }
    }
  }
}

// Fold of each pixel's probabilities in face order (:174-182), one warp per active region
// (lane = pixel) once every chunk is evaluated: a chain of dependent FP64 multiplies per
// pixel that only many resident warps can hide, hence its own small-footprint kernel.
// Rows are coalesced, eight independent loads at a time (a missing hit multiplies by
// exactly 1).
__global__ void __launch_bounds__(256) softmask_pairs_fold_kernel(const SoftArgs a, const Plan p) {
  pdl_prologue();
  const PairBuf& pb = a.pairs;
  if (pairs_overflowed(pb)) return;
  const int lane = lane_id();
  const long long nact = (long long)pb.cursor[1];
  for (long long act = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); act < nact;
       act += (long long)gridDim.x * 8) {
    int b, ix, iy, cnt;
    pair_region(pb, p, act, lane, b, ix, iy, cnt);
    const int segreg = lane < pb.nseg ? pb.act_seg[act * pb.nseg + lane] : -1;
    const int mx = __reduce_max_sync(DIBR_FULL_MASK, cnt);
    float ap = 1.0f;
    for (int q = 0; q * kPairSegHits < mx; q++) {
      int sg = __shfl_sync(DIBR_FULL_MASK, segreg, q & 31);
      if (q >= 32) sg = pb.act_seg[act * pb.nseg + q];
      const float* row = pb.prob + (long long)sg * kPairSegSlots + lane;
      float pr[kPairSegHits];
#pragma unroll
      for (int j = 0; j < kPairSegHits; j++)
        pr[j] = q * kPairSegHits + j < cnt ? __ldcg(row + j * 32) : 0.f;
#pragma unroll
      for (int j = 0; j < kPairSegHits; j++)
        ap = __double2float_rn(__dmul_rn((double)ap, __dsub_rn(1.0, (double)pr[j])));  // :176-177
    }
    if (cnt > 0)
      a.mask[((long long)b * p.H + iy) * p.W + ix] =
          __double2float_rn(__dsub_rn(1.0, (double)ap));  // :181-182
  }
}

#ifndef DIBR_GRAD_MINB
#define DIBR_GRAD_MINB 4
#endif
__global__ void __launch_bounds__(kPairWarps * 32, DIBR_GRAD_MINB)
softmask_pairs_grad_kernel(const SoftArgs a, const Plan p) {
  __shared__ float sX[kPairWarps][6][kPairChunk];
  pdl_prologue();
  const PairBuf& pb = a.pairs;
  if (pairs_overflowed(pb)) return;
  const int lane = lane_id();
  const unsigned nchunk = min(*pb.nchunk, 3u * (unsigned)pb.cap);
  const Grid g = plan_grid(p, a.job.mult);
  const float mult = a.job.mult;
  for (unsigned w = blockIdx.x * kPairWarps + (threadIdx.x >> 5); w < nchunk;
       w += gridDim.x * kPairWarps) {
    PairWork k;
    if (!pair_work(pb, p, w, lane, k)) continue;
    const int rx0 = k.ix - (lane & 7), ry0 = k.iy - (lane >> 3);
    float my_dLdp = 0.f, my_allprob = 0.f;
    if (k.cnt > 0) {
      const long long pix = ((long long)k.b * p.H + k.iy) * p.W + k.ix;
      my_dLdp = a.grad_mask[pix];
      my_allprob = a.mask_in[pix];
    }
    constexpr int U = kPairChunk / 32;
    int f[U], lpack = 0, epack = 0;
    float pr[U];
#pragma unroll
    for (int u = 0; u < U; u++) {      // all loads of the chunk first (see the eval kernel)
      const int i = k.i_begin + 32 * u + lane;
      int l, h;
      const long long slot = pair_locate(pb, k.act, k.pd, min(i, k.i_end - 1), l, h);
      lpack |= l << (8 * u);
      f[u] = __ldcg(pb.face + slot);
      epack |= (int)__ldcg(pb.type + slot) << (8 * u);
      pr[u] = __ldcg(pb.prob + slot);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      float X[6];
      load_scaled_coords(a.job, k.b, f[u], X);
#pragma unroll
      for (int c = 0; c < 6; c++) sX[threadIdx.x >> 5][c][32 * u + lane] = X[c];
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (k.i_begin + 32 * u >= k.i_end) break;
      const int l = (lpack >> (8 * u)) & 31;
      const int eid = ((epack >> (8 * u)) & 255) - 1;
      const float dLdp = __shfl_sync(DIBR_FULL_MASK, my_dLdp, l);
      const float allprob = __shfl_sync(DIBR_FULL_MASK, my_allprob, l);
      if (k.i_begin + 32 * u + lane >= k.i_end) continue;
      float X[6];
#pragma unroll
      for (int c = 0; c < 6; c++) X[c] = sX[threadIdx.x >> 5][c][32 * u + lane];
      const float x0 = pix_x(g, rx0 + (l & 7)), y0 = pix_y(g, ry0 + (l >> 3));
      double inv_e = 0.0;
      if (eid < 3) {
        const int w2 = (eid + 1) % 3;
        inv_e = edge_inv(pick3(eid, X[0], X[2], X[4]), pick3(eid, X[1], X[3], X[5]),
                         pick3(w2, X[0], X[2], X[4]), pick3(w2, X[1], X[3], X[5]));
      }
      int v1, v2;
      float g1x, g1y, g2x, g2y;
      soft_grad_core(x0, y0, X, eid, pr[u], allprob, dLdp, a.sigmainv, mult, inv_e, v1, g1x, g1y,
                     v2, g2x, g2y);
      // one 8-byte vector reduction per touched vertex ((x, y) of a vertex are adjacent
      // and 8-byte aligned in grad_face_vertices_image) instead of up to four scalar ones
      if (a.job.faces) {   // mesh mode: straight into the vertex gradients (B, V, 2)
        int vid[3];
        mesh_face(a.job, f[u], vid);
        float2* gv = reinterpret_cast<float2*>(a.grad_img) + (long long)k.b * a.job.V;
        if (g1x != 0.f || g1y != 0.f) atomicAdd(gv + pick3i(v1, vid[0], vid[1], vid[2]), make_float2(g1x, g1y));
        if (v2 >= 0 && (g2x != 0.f || g2y != 0.f))
          atomicAdd(gv + pick3i(v2, vid[0], vid[1], vid[2]), make_float2(g2x, g2y));
      } else {
        float2* gp = reinterpret_cast<float2*>(
            a.grad_img + ((long long)k.b * a.job.view_stride + f[u]) * 6);
        if (g1x != 0.f || g1y != 0.f) atomicAdd(gp + v1, make_float2(g1x, g1y));
        if (v2 >= 0 && (g2x != 0.f || g2y != 0.f)) atomicAdd(gp + v2, make_float2(g2x, g2y));
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Dense-state backward: the raw _C op (dibr_soft_mask_cuda.cu:230-353) on the per-pixel
// lists (B, H, W, K) the dense forward stored; float and double (the reference dispatches
// both).  Only the reference-shaped shim uses it; the autograd path runs over the pair
// buffer.  Mapping: one warp per 8x4-pixel region; the region's (pixel, slot) pairs are
// dealt out evenly over the lanes whatever pixel they belong to (an exclusive scan of the
// per-pixel slot counts, owner found by a search over the lanes), each pair takes the
// gradient arithmetic in double, and lanes that hit the same (face, vertex) are grouped with
// __match_any_sync and summed with shuffles, so that ONE atomic per (face, vertex) and step
// leaves the warp.
// ---------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void add_grouped(T* base, long long key, bool has, T gx, T gy,
                                            int lane) {
  // lanes with the same key (the same (face, vertex) slot) sum their (gx, gy); lanes
  // without a contribution get unique keys
  const long long k = has ? key : -1ll - lane;
  const unsigned peers = __match_any_sync(DIBR_FULL_MASK, k);
  T sx = gx, sy = gy;
  unsigned rest = peers & ~(1u << lane);
  // at most a handful of lanes share a key: walk the peer set (all lanes iterate together)
  while (__any_sync(DIBR_FULL_MASK, rest != 0)) {
    const int src = rest ? __ffs(rest) - 1 : lane;
    const T ox = __shfl_sync(DIBR_FULL_MASK, gx, src), oy = __shfl_sync(DIBR_FULL_MASK, gy, src);
    if (rest) { sx += ox; sy += oy; rest &= rest - 1; }
  }
  if (has && (peers & lanemask_lt()) == 0 && (sx != (T)0 || sy != (T)0)) {
    atomicAdd(base + key * 2, sx);
    atomicAdd(base + key * 2 + 1, sy);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) softmask_dense_backward_kernel(const SoftDenseBwdArgsT<T> a) {
  pdl_prologue();
  const int lane = lane_id();
  const int nrx = (a.W + 7) >> 3, nry = (a.H + 3) >> 2;
  const long long nreg = (long long)a.B * nrx * nry;
  const Grid g = make_grid(a.mult, a.W, a.H);
  for (long long reg = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); reg < nreg;
       reg += (long long)gridDim.x * 8) {
    const int b = (int)(reg / (nrx * nry)), rr = (int)(reg - (long long)b * nrx * nry);
    const int ix = (rr % nrx) * 8 + (lane & 7), iy = (rr / nrx) * 4 + (lane >> 3);
    const bool inimg = ix < a.W && iy < a.H;
    const long long pix = ((long long)b * a.H + min(iy, a.H - 1)) * a.W + min(ix, a.W - 1);
    int cnt = 0;
    double dLdp = 0.0, allprob = 0.0;
    if (inimg && a.sel[pix] < 0) {
      while (cnt < a.K && a.cidx[pix * a.K + cnt] >= 0) cnt++;   // slots are filled in order (:271-276)
      dLdp = (double)a.grad_mask[pix];
      allprob = (double)a.mask[pix];
    }
    int total;
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
      if (lane >= d) incl += t;
    }
    const int off = incl - cnt;
    total = __shfl_sync(DIBR_FULL_MASK, incl, 31);
    const double x0 = (double)pix_x(g, ix), y0 = (double)pix_y(g, iy);   // float products (:262-263)
    for (int i0 = 0; i0 < total; i0 += 32) {
      const int i = i0 + lane;
      const bool act = i < total;
      int l = 0;   // owner pixel lane: the largest l with off[l] <= i
#pragma unroll
      for (int st = 16; st > 0; st >>= 1) {
        const int o = __shfl_sync(DIBR_FULL_MASK, off, l + st);
        if (o <= i) l += st;
      }
      const int h = i - __shfl_sync(DIBR_FULL_MASK, off, l);
      const long long opix = __shfl_sync(DIBR_FULL_MASK, pix, l);
      const double ox0 = __shfl_sync(DIBR_FULL_MASK, x0, l), oy0 = __shfl_sync(DIBR_FULL_MASK, y0, l);
      const double odL = __shfl_sync(DIBR_FULL_MASK, dLdp, l), oap = __shfl_sync(DIBR_FULL_MASK, allprob, l);
      int v1 = 0, v2 = -1;
      double g1x = 0., g1y = 0., g2x = 0., g2y = 0.;
      long long fb = 0;
      if (act) {
        const long long o = opix * a.K + h;
        const int f = (int)a.cidx[o];
        fb = (long long)b * a.F + f;
        const T* ip = a.img + fb * 6;
        const double X[6] = {(double)ip[0], (double)ip[1], (double)ip[2], (double)ip[3],
                             (double)ip[4], (double)ip[5]};
        soft_grad_core_sel(ox0, oy0, X, (int)a.type[o] - 1, (double)a.prob[o], oap, odL,
                           (double)a.sigmainv, (double)a.mult, v1, g1x, g1y, v2, g2x, g2y);
      }
      add_grouped<T>(a.grad_img, fb * 3 + v1, act, (T)g1x, (T)g1y, lane);
      add_grouped<T>(a.grad_img, fb * 3 + max(v2, 0), act && v2 >= 0, (T)g2x, (T)g2y, lane);
    }
  }
}

// ---------------------------------------------------------------------------- host
template <int MODE>
static cudaError_t launch_soft_mode(const SoftArgs& a, const Plan& p, cudaStream_t st) {
  const long long nblk = (long long)p.nkx * p.nky * p.B;
  const bool fallback = MODE != SOFT_FWD_DENSE && a.pairs.cap > 0;
  dim3 grid((unsigned)(fallback ? min(nblk, (long long)148 * 3) : nblk));
  cudaError_t e = cudaSuccess;
  if (p.blk_log2 == 5) {
    const int smem = (int)sizeof(SoftSmem<5, MODE>);
    e = smem_optin((const void*)softmask_kernel<5, MODE>, smem);
    if (e != cudaSuccess) return e;
    prof_begin(MODE == SOFT_BWD ? K_SOFT_BWD : K_SOFT_FWD, st);
    launch_k((softmask_kernel<5, MODE>), grid, kSoftThreads, smem, st, a, p);
  } else {
    const int smem = (int)sizeof(SoftSmem<6, MODE>);
    e = smem_optin((const void*)softmask_kernel<6, MODE>, smem);
    if (e != cudaSuccess) return e;
    prof_begin(MODE == SOFT_BWD ? K_SOFT_BWD : K_SOFT_FWD, st);
    launch_k((softmask_kernel<6, MODE>), grid, kSoftThreads, smem, st, a, p);
  }
  prof_end(st);
  return cudaGetLastError();
}

// persistent grid: every warp strides over the chunk list (its length is only known on the
// device); enough CTAs to fill the GPU, never more warps than there can be chunks
static int pair_grid(const PairBuf& pb) {
  const long long most = (3LL * pb.cap + kPairWarps - 1) / kPairWarps;
  const long long fill = 148LL * (2048 / (kPairWarps * 32));
  return (int)(most < fill ? most : fill);
}

cudaError_t launch_softmask_pairs_eval(const SoftArgs& a, const Plan& p, cudaStream_t st) {
  if (a.pairs.cap <= 0) return cudaSuccess;
  prof_begin(K_SOFT_EVAL, st);
  launch_k(softmask_pairs_eval_kernel, pair_grid(a.pairs), kPairWarps * 32, 0, st, a, p);
  prof_end(st);
  prof_begin(K_SOFT_FOLD, st);
  const long long regions = (a.pairs.cap + 7) / 8;
  launch_k(softmask_pairs_fold_kernel, (int)(regions < 148 * 8 ? regions : 148 * 8), 256, 0, st, a, p);
  prof_end(st);
  return cudaGetLastError();
}

cudaError_t launch_softmask_pairs_grad(const SoftArgs& a, const Plan& p, cudaStream_t st) {
  if (a.pairs.cap <= 0) return cudaSuccess;
  prof_begin(K_SOFT_GRAD, st);
  launch_k(softmask_pairs_grad_kernel, pair_grid(a.pairs), kPairWarps * 32, 0, st, a, p);
  prof_end(st);
  return cudaGetLastError();
}

cudaError_t launch_softmask(const SoftArgs& a, int mode, const Plan& p, cudaStream_t st) {
  switch (mode) {
    case SOFT_GATHER: return launch_softmask_gather(a, p, st);
    case SOFT_FWD: return launch_soft_mode<SOFT_FWD>(a, p, st);
    case SOFT_FWD_DENSE: return launch_soft_mode<SOFT_FWD_DENSE>(a, p, st);
    default: return launch_soft_mode<SOFT_BWD>(a, p, st);
  }
}

template <typename T>
static cudaError_t launch_dense_backward_t(const SoftDenseBwdArgsT<T>& a, cudaStream_t st) {
  const long long nreg = (long long)a.B * ((a.W + 7) >> 3) * ((a.H + 3) >> 2);
  if (nreg == 0) return cudaSuccess;
  const int blocks = (int)min((nreg + 7) / 8, (long long)148 * 16);
  prof_begin(K_SOFT_BWD_DENSE, st);
  launch_k(softmask_dense_backward_kernel<T>, blocks, 256, 0, st, a);
  prof_end(st);
  return cudaGetLastError();
}
cudaError_t launch_softmask_dense_backward(const SoftDenseBwdArgs& a, cudaStream_t st) {
  return launch_dense_backward_t<float>(a, st);
}
cudaError_t launch_softmask_backward_f64(const SoftBwdF64Args& a, cudaStream_t st) {
  return launch_dense_backward_t<double>(a, st);
}

}  // namespace dibr
