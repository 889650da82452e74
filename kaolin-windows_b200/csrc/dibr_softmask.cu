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

// prob and nearest-feature id (0-2 edge, 3-5 vertex) of one (pixel, face) pair.
__device__ __forceinline__ void soft_eval(float x0, float y0, const float (&X)[6], float four_mm,
                                          float sigmainv, float mult, float& prob, int& edgeid) {
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
    const float x3 = __double2float_rn(__ddiv_rn((double)x3n, dd));
    const float y3 = __double2float_rn(__ddiv_rn((double)y3n, dd));
    const float direct = __fmaf_rn(__fsub_rn(x3, x1), __fsub_rn(x3, x2),
                                   __fmul_rn(__fsub_rn(y3, y1), __fsub_rn(y3, y2)));
    if (direct > 0.f)
      pdis[i] = four_mm;
    else
      pdis[i] = __double2float_rn(__ddiv_rn((double)__fmul_rn(up, up), dd));
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

__device__ __forceinline__ float pick3(int v, float a, float b, float c) {
  return v == 0 ? a : (v == 1 ? b : c);
}

// dibr_soft_mask_cuda.cu:283-347: gradient of one (pixel, face) pair w.r.t. the face's
// six UNscaled coordinates, added into g[6] (register array: static indices only).
__device__ __forceinline__ void soft_grad(float x0, float y0, const float (&X)[6], int edgeid,
                                          float prob, float allprob, float dLdp, float sigmainv,
                                          float mult, float (&g)[6]) {
  const float dLdz = (float)(-1.0 * sigmainv * dLdp * (1.0 - allprob) /
                             (1.0 - prob + DIBR_SOFT_EPS) * prob);
  int v1, v2 = -1;
  float g1x, g1y, g2x = 0.f, g2y = 0.f;
  if (edgeid >= 3) {
    v1 = edgeid - 3;
    const float x1 = pick3(v1, X[0], X[2], X[4]), y1 = pick3(v1, X[1], X[3], X[5]);
    const float dLdx1 = dLdz * 2 * (x1 - x0);
    const float dLdy1 = dLdz * 2 * (y1 - y0);
    g1x = dLdx1 / mult;
    g1y = dLdy1 / mult;
  } else {
    v1 = edgeid;
    v2 = (edgeid + 1) % 3;
    const float x1 = pick3(v1, X[0], X[2], X[4]), y1 = pick3(v1, X[1], X[3], X[5]);
    const float x2 = pick3(v2, X[0], X[2], X[4]), y2 = pick3(v2, X[1], X[3], X[5]);
    const float A = y2 - y1, B = x1 - x2, C = x2 * y1 - x1 * y2;
    const float up = A * x0 + B * y0 + C;
    const float down = A * A + B * B;
    const float dissquare = up * up / (down + DIBR_SOFT_EPS);
    const float dzdA = 2 * (x0 * up - dissquare * A) / (down + DIBR_SOFT_EPS);
    const float dzdB = 2 * (y0 * up - dissquare * B) / (down + DIBR_SOFT_EPS);
    const float dzdC = 2 * up / (down + DIBR_SOFT_EPS);
    const float dLdx1 = dLdz * (dzdB - y2 * dzdC);
    const float dLdy1 = dLdz * (x2 * dzdC - dzdA);
    const float dLdx2 = dLdz * (y1 * dzdC - dzdB);
    const float dLdy2 = dLdz * (dzdA - x1 * dzdC);
    g1x = dLdx1 / mult;
    g1y = dLdy1 / mult;
    g2x = dLdx2 / mult;
    g2y = dLdy2 / mult;
  }
#pragma unroll
  for (int v = 0; v < 3; v++) {
    if (v == v1) { g[2 * v] += g1x; g[2 * v + 1] += g1y; }
    if (v == v2) { g[2 * v] += g2x; g[2 * v + 1] += g2y; }
  }
}

template <int BLK_LOG2, int MODE>
__global__ void __launch_bounds__((1 << (2 * (BLK_LOG2 - 4))) * 32)
softmask_kernel(const SoftArgs a, const Plan p) {
  constexpr int BLK = 1 << BLK_LOG2;
  constexpr int TPB = BLK / 16;          // tiles per block row
  constexpr int NW = TPB * TPB;          // warps = tiles
  constexpr int NT = NW * 32;
  constexpr int CH = kSoftChunk;
  __shared__ float s_x[6][CH];           // scaled coordinates ax ay bx by cx cy
  __shared__ float s_bb[4][CH];          // enlarged box
  __shared__ int s_face[CH];
  __shared__ unsigned short s_tl[NW][CH];  // per-warp compacted tile list (chunk slots)
  __shared__ float s_acc[MODE == SOFT_BWD ? CH : 1][6];

  const int tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;
  const int b = blockIdx.z;
  const int bx0 = blockIdx.x << BLK_LOG2, by0 = blockIdx.y << BLK_LOG2;
  const Grid g = make_grid(a.job.mult, p.W, p.H);
  const float mult = a.job.mult;
  const float four_mm = __fmul_rn(__fmul_rn(mult, 4.f), mult);

  // this warp's tile and this lane's pixel in each of the 8 regions
  const int tx0 = bx0 + (warp % TPB) * 16, ty0 = by0 + (warp / TPB) * 16;
  const int lx = lane & 7, ly = lane >> 3;
  unsigned open_bits = 0;                // bit r: pixel of region r still collects faces
  int cnt[8];
  float ap[8];
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const int ix = tx0 + (r & 1) * 8 + lx, iy = ty0 + (r >> 1) * 4 + ly;
    cnt[r] = 0;
    ap[r] = 1.0f;
    if (ix < p.W && iy < p.H) {
      const long long pix = ((long long)b * p.H + iy) * p.W + ix;
      if (a.sel[pix] < 0) open_bits |= 1u << r;
    }
  }
  const unsigned uncov_bits = open_bits;
  if (a.K <= 0) open_bits = 0;

  // tile rectangle in pixel-centre coordinates (x grows with ix, y shrinks with iy)
  const float t_xlo = pix_x(g, tx0), t_xhi = pix_x(g, min(tx0 + 15, p.W - 1));
  const float t_yhi = pix_y(g, ty0), t_ylo = pix_y(g, min(ty0 + 15, p.H - 1));

  const int shift = p.bin_log2 - BLK_LOG2;
  const int bin = (blockIdx.y >> shift) * p.nbx + (blockIdx.x >> shift);
  const long long row = (long long)b * p.NB + bin;
  const long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);

  for (long long base = 0; base < n; base += CH) {
    if (!__syncthreads_or(open_bits != 0)) break;
    const int nc = (int)min((long long)CH, n - base);
    // ---- stage the chunk (scaled coordinates + enlarged box), zero the accumulators
    for (int t = tid; t < nc; t += NT) {
      const int f = a.job.list[start + base + t];
      Face2D fc;
      load_face2d(a.job, b, f, fc);
      s_x[0][t] = fc.ax; s_x[1][t] = fc.ay; s_x[2][t] = fc.bx;
      s_x[3][t] = fc.by; s_x[4][t] = fc.cx; s_x[5][t] = fc.cy;
      s_bb[0][t] = fc.xmin; s_bb[1][t] = fc.ymin; s_bb[2][t] = fc.xmax; s_bb[3][t] = fc.ymax;
      s_face[t] = f;
      if (MODE == SOFT_BWD) {
#pragma unroll
        for (int i = 0; i < 6; i++) s_acc[t][i] = 0.f;
      }
    }
    __syncthreads();

    if (__any_sync(DIBR_FULL_MASK, open_bits != 0)) {
      // ---- compact the chunk to this warp's tile, keeping the order
      int ntl = 0;
      for (int t0 = 0; t0 < nc; t0 += 32) {
        const int t = t0 + lane;
        bool hit = false;
        if (t < nc)
          hit = s_bb[0][t] <= t_xhi && s_bb[2][t] > t_xlo && s_bb[1][t] <= t_yhi &&
                s_bb[3][t] > t_ylo;
        const unsigned m = __ballot_sync(DIBR_FULL_MASK, hit);
        if (hit) s_tl[warp][ntl + __popc(m & lanemask_lt())] = (unsigned short)t;
        ntl += __popc(m);
      }
      __syncwarp();

      // ---- regions
      for (int r = 0; r < 8; r++) {
        if (!__any_sync(DIBR_FULL_MASK, (open_bits >> r) & 1)) continue;
        const int ix = tx0 + (r & 1) * 8 + lx, iy = ty0 + (r >> 1) * 4 + ly;
        const int rx0 = tx0 + (r & 1) * 8, ry0 = ty0 + (r >> 1) * 4;
        const float r_xlo = pix_x(g, rx0), r_xhi = pix_x(g, min(rx0 + 7, p.W - 1));
        const float r_yhi = pix_y(g, ry0), r_ylo = pix_y(g, min(ry0 + 3, p.H - 1));
        const float x0 = pix_x(g, ix), y0 = pix_y(g, iy);
        const long long pix = ((long long)b * p.H + min(iy, p.H - 1)) * p.W + min(ix, p.W - 1);
        bool open = (open_bits >> r) & 1;
        int my_cnt = 0;
        float my_ap = 1.f;
#pragma unroll
        for (int rr = 0; rr < 8; rr++)
          if (rr == r) { my_cnt = cnt[rr]; my_ap = ap[rr]; }
        float dLdp = 0.f, apf = 0.f;
        if (MODE == SOFT_BWD && open) { dLdp = a.grad_mask[pix]; apf = a.mask_in[pix]; }

        for (int j0 = 0; j0 < ntl; j0 += 32) {
          int slot = 0;
          bool cand = false;
          if (j0 + lane < ntl) {
            slot = s_tl[warp][j0 + lane];
            cand = s_bb[0][slot] <= r_xhi && s_bb[2][slot] > r_xlo && s_bb[1][slot] <= r_yhi &&
                   s_bb[3][slot] > r_ylo;
          }
          unsigned m = __ballot_sync(DIBR_FULL_MASK, cand);
          while (m) {
            const int k = __ffs(m) - 1;
            m &= m - 1;
            const int s = __shfl_sync(DIBR_FULL_MASK, slot, k);
            const bool in = open && !(x0 < s_bb[0][s] || x0 >= s_bb[2][s] || y0 < s_bb[1][s] ||
                                      y0 >= s_bb[3][s]);  // :95-97
            float gacc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (in) {
              const float X[6] = {s_x[0][s], s_x[1][s], s_x[2][s], s_x[3][s], s_x[4][s], s_x[5][s]};
              float prob;
              int edgeid;
              soft_eval(x0, y0, X, four_mm, a.sigmainv, mult, prob, edgeid);
              if (MODE == SOFT_FWD_DENSE) {
                const long long o = pix * a.K + my_cnt;
                a.prob[o] = prob;
                a.cidx[o] = s_face[s];
                a.type[o] = (uint8_t)(edgeid + 1);
              }
              if (MODE == SOFT_BWD) {
                soft_grad(x0, y0, X, edgeid, prob, apf, dLdp, a.sigmainv, mult, gacc);
              } else {
                my_ap = __double2float_rn(__dmul_rn((double)my_ap, __dsub_rn(1.0, (double)prob)));
              }
              if (++my_cnt >= a.K) open = false;  // :170-171
            }
            if (MODE == SOFT_BWD) {
              if (__any_sync(DIBR_FULL_MASK, in)) {
#pragma unroll
                for (int i = 0; i < 6; i++) {
                  float v = gacc[i];
#pragma unroll
                  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(DIBR_FULL_MASK, v, d);
                  if (lane == i && v != 0.f) atomicAdd(&s_acc[s][i], v);
                }
              }
            }
            if (!__any_sync(DIBR_FULL_MASK, open)) { m = 0; j0 = ntl; }
          }
        }
#pragma unroll
        for (int rr = 0; rr < 8; rr++)
          if (rr == r) { cnt[rr] = my_cnt; ap[rr] = my_ap; }
        if (!open) open_bits &= ~(1u << r);
      }
    }
    __syncthreads();
    if (MODE == SOFT_BWD) {
      for (int t = tid; t < nc * 6; t += NT) {
        const float v = s_acc[t / 6][t % 6];
        if (v != 0.f)
          atomicAdd(a.grad_img + ((long long)b * a.job.view_stride + s_face[t / 6]) * 6 + t % 6, v);
      }
      __syncthreads();
    }
  }

  if (MODE != SOFT_BWD) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      const int ix = tx0 + (r & 1) * 8 + lx, iy = ty0 + (r >> 1) * 4 + ly;
      if (ix < p.W && iy < p.H) {
        const long long pix = ((long long)b * p.H + iy) * p.W + ix;
        a.mask[pix] = ((uncov_bits >> r) & 1)
                          ? __double2float_rn(__dsub_rn(1.0, (double)ap[r]))  // :181-182
                          : 1.0f;                                             // :68-70
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Dense-state backward: the raw _C op (dibr_soft_mask_cuda.cu:230-353), one thread per
// pixel reading the per-pixel lists the dense forward stored.  Only the reference-shaped
// shim uses it; the autograd path uses the recompute form above.
// ---------------------------------------------------------------------------

__global__ void __launch_bounds__(256) softmask_dense_backward_kernel(const SoftDenseBwdArgs a) {
  const long long npix = (long long)a.B * a.H * a.W;
  const Grid g = make_grid(a.mult, a.W, a.H);
  for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < npix;
       pix += (long long)gridDim.x * blockDim.x) {
    if (a.sel[pix] >= 0) continue;
    const int ix = (int)(pix % a.W), iy = (int)((pix / a.W) % a.H), b = (int)(pix / ((long long)a.W * a.H));
    const float x0 = pix_x(g, ix), y0 = pix_y(g, iy);
    const float dLdp = a.grad_mask[pix], allprob = a.mask[pix];
    for (int kid = 0; kid < a.K; kid++) {
      const long long o = pix * a.K + kid;
      const int f = (int)a.cidx[o];
      if (f < 0) break;
      const float* ip = a.img + ((long long)b * a.F + f) * 6;
      const float X[6] = {ip[0], ip[1], ip[2], ip[3], ip[4], ip[5]};
      float gv[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      soft_grad(x0, y0, X, (int)a.type[o] - 1, a.prob[o], allprob, dLdp, a.sigmainv, a.mult, gv);
      float* gp = a.grad_img + ((long long)b * a.F + f) * 6;
#pragma unroll
      for (int i = 0; i < 6; i++)
        if (gv[i] != 0.f) atomicAdd(gp + i, gv[i]);
    }
  }
}

// ---------------------------------------------------------------------------- host
template <int MODE>
static cudaError_t launch_soft_mode(const SoftArgs& a, const Plan& p, cudaStream_t st) {
  dim3 grid(p.nkx, p.nky, p.B);
  prof_begin(MODE == SOFT_BWD ? K_SOFT_BWD : K_SOFT_FWD, st);
  if (p.blk_log2 == 5)
    softmask_kernel<5, MODE><<<grid, 128, 0, st>>>(a, p);
  else
    softmask_kernel<6, MODE><<<grid, 512, 0, st>>>(a, p);
  prof_end(st);
  return cudaGetLastError();
}

cudaError_t launch_softmask(const SoftArgs& a, int mode, const Plan& p, cudaStream_t st) {
  switch (mode) {
    case SOFT_FWD: return launch_soft_mode<SOFT_FWD>(a, p, st);
    case SOFT_FWD_DENSE: return launch_soft_mode<SOFT_FWD_DENSE>(a, p, st);
    default: return launch_soft_mode<SOFT_BWD>(a, p, st);
  }
}

cudaError_t launch_softmask_dense_backward(const SoftDenseBwdArgs& a, cudaStream_t st) {
  const long long npix = (long long)a.B * a.H * a.W;
  if (npix == 0) return cudaSuccess;
  const int blocks = (int)min((npix + 255) / 256, (long long)148 * 16);
  prof_begin(K_SOFT_BWD_DENSE, st);
  softmask_dense_backward_kernel<<<blocks, 256, 0, st>>>(a);
  prof_end(st);
  return cudaGetLastError();
}

}  // namespace dibr
