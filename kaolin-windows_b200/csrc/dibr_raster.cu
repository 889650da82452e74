// dibr_raster.cu -- block rasterizer forward (north_star (b)) and its backward ((d)).
//
// Forward: one CTA per screen block (32x32 or 64x64 pixels) of one view, blocks taken
// longest face list first.  The block's depth buffer lives in shared memory as 64-bit keys
// (order-preserving depth | ~face).  The bin's face list is consumed in chunks of 256:
//   A   one face per thread: the EXACT pixel rectangle of its box inside the block (the
//       reference's half-open float test on the range ends) and, per rectangle row, a
//       CONSERVATIVE span [xa, xb] outside of which the face cannot cover a pixel centre
//       (edge functions in a block-local frame, widened by one full pixel in x and y);
//   B   the (face, row, x) candidates inside the spans are dealt out evenly over the
//       threads, a contiguous run each, and every candidate takes the reference's exact
//       edge / barycentric / depth arithmetic; hits are published with a shared-memory
//       64-bit max.
// Candidates are thus ~the covered pixels plus a rim, instead of whole bounding boxes, and
// every lane is busy whatever the face sizes are.  A final pixel-parallel pass decodes the
// winner of every pixel, recomputes its weights with the same instruction sequence
// (bit-identical) and writes selected_face_idx / interpolated_features (/ output_weights)
// with coalesced stores; background pixels are written too, so the caller needs no memset.
#include "dibr_common.cuh"

namespace dibr {


constexpr int kRasterThreads = 256;
constexpr int kRasterWarps = kRasterThreads / 32;
constexpr int kRowBatch = 2048;  // row spans expanded per pass (+ one face's rows of slack)

template <int BLK_LOG2>
struct RasterSmem {
  static constexpr int BLK = 1 << BLK_LOG2;
  static constexpr int ROWS = kRowBatch + BLK;
  unsigned long long zbuf[BLK * BLK];
  float x[6][kRasterThreads];   // staged faces of the current chunk (scaled coordinates)
  float z[3][kRasterThreads];
  int face[kRasterThreads];
  int rpre[kRasterThreads + 1]; // exclusive prefix of the faces' row counts
  int cpre[kRasterThreads];     // candidates before the face's first row (within the batch)
  int wsum[kRasterWarps];
  unsigned span[ROWS];          // slot | ry << 8 | xa << 16 | len << 24 (len <= 64 -> 7 bits)
  int spre[ROWS + 1];           // exclusive prefix of the span lengths within the batch
  int unc4[4];   // uncovered pixels per quarter block (the gather's CTA footprint)
};

// block-wide exclusive scan of one int per thread; returns the exclusive prefix, total in
// `total`.  Two barriers.
__device__ __forceinline__ int block_scan(int v, int* wsum, int lane, int warp, int& total) {
  int incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
    if (lane >= d) incl += t;
  }
  __syncthreads();  // wsum free to overwrite
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int woff = 0;
  total = 0;
#pragma unroll
  for (int i = 0; i < kRasterWarps; i++) {
    const int t = wsum[i];
    if (i < warp) woff += t;
    total += t;
  }
  return woff + incl - v;
}

template <int BLK_LOG2>
#ifndef DIBR_RASTER_MINB
#define DIBR_RASTER_MINB 4
#endif
__global__ void __launch_bounds__(kRasterThreads, DIBR_RASTER_MINB)
raster_forward_kernel(const RasterArgs a, const Plan p) {
  constexpr int BLK = 1 << BLK_LOG2;
  extern __shared__ __align__(16) unsigned char s_raw[];
  RasterSmem<BLK_LOG2>& sm = *reinterpret_cast<RasterSmem<BLK_LOG2>*>(s_raw);
  unsigned long long* zbuf = sm.zbuf;

  pdl_prologue();
  DIBR_TRACE_BEGIN();
  // blocks are taken longest face list first (lpt_block)
  int b, bxi, byi;
  if (!lpt_block(a.job, p, BLK_LOG2, blockIdx.x, b, bxi, byi)) return;
  const int bx0 = bxi << BLK_LOG2, by0 = byi << BLK_LOG2;
  const int bx1 = min(bx0 + BLK, p.W) - 1, by1 = min(by0 + BLK, p.H) - 1;
  const Grid g = plan_grid(p, a.job.mult);
  const int tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;
  // local frame of the conservative span culling: origin = centre of the block's first
  // pixel, pixel (rx, ry) at (rx * pitch_x, -ry * pitch_y)
  const float X0 = pix_x(g, bx0), Y0 = pix_y(g, by0);
  const float pitch_x = 2.f * g.xs, pitch_y = 2.f * g.ys;

  for (int i = tid; i < BLK * BLK; i += kRasterThreads) zbuf[i] = 0ull;
  if (tid < 4) sm.unc4[tid] = 0;

  // parent bin of this block
  const int shift = p.bin_log2 - BLK_LOG2;
  const int bin = (byi >> shift) * p.nbx + (bxi >> shift);
  const long long row = (long long)b * p.NB + bin;
  long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);  // overflowed call

#ifdef DIBR_TRACE_PHASES
  long long ph_a = 0, ph_a2 = 0, ph_b = 0, ph_t = clock64();
#define PH_MARK(acc) do { const long long t_ = clock64(); acc += t_ - ph_t; ph_t = t_; } while (0)
#else
#define PH_MARK(acc)
#endif
  for (long long base = 0; base < n; base += kRasterThreads) {
    // ---- phase A: one face per thread -> exact pixel rectangle inside this block and the
    // coefficients of a CONSERVATIVE per-row span test (see DESIGN.md "Span culling")
    int x_lo = 0, x_hi = -1, y_lo = 0, y_hi = -1;
    float ia[3], sb[3], scm[3];
    bool cull = false;
    {
      const long long li = base + tid;
      if (li < n) {
        int f;
        float z_a, z_b, z_c;
        Face2D fc;
        bool have;
        if (a.rec) {
          // the list as compact records written by the binning pass: three 16-byte loads
          // from consecutive addresses, no dependent gather through the face id
          const float4* r = a.rec + (start + li) * 3;
          const float4 q0 = __ldg(r), q1 = __ldg(r + 1), q2 = __ldg(r + 2);
          if (li + kRasterThreads < n) prefetch_l1(r + 3 * kRasterThreads);
          fc.ax = q0.x; fc.ay = q0.y; fc.bx = q0.z; fc.by = q0.w; fc.cx = q1.x; fc.cy = q1.y;
          z_a = q1.z; z_b = q1.w; z_c = q2.x;
          f = __float_as_int(q2.y);
          // the box of the scaled coordinates (rasterization.py:342-344)
          fc.xmin = fminf(fminf(fc.ax, fc.bx), fc.cx); fc.ymin = fminf(fminf(fc.ay, fc.by), fc.cy);
          fc.xmax = fmaxf(fmaxf(fc.ax, fc.bx), fc.cx); fc.ymax = fmaxf(fmaxf(fc.ay, fc.by), fc.cy);
          have = true;
        } else {
          f = a.job.list[start + li];
          // the next chunk's entry now, its face record into L1 while this chunk is worked on
          if (!a.job.faces && li + kRasterThreads < n) {
            const long long gn = (long long)b * a.job.view_stride +
                                 a.job.list[start + li + kRasterThreads];
            prefetch_l1(a.job.img + gn * 6);
            prefetch_l1(a.z + gn * 3);
          }
          // depths requested together with the coordinates (one round trip, not two)
          face_depths(a.job, a.z, b, f, z_a, z_b, z_c);
          have = load_face2d<true>(a.job, b, f, fc);
        }
        if (have &&
            pixel_range(g, fc.xmin, fc.ymin, fc.xmax, fc.ymax, x_lo, x_hi, y_lo, y_hi)) {
          x_lo = max(x_lo, bx0); x_hi = min(x_hi, bx1);
          y_lo = max(y_lo, by0); y_hi = min(y_hi, by1);
          // trim the conservative range to the reference's half-open test
          // (rasterization_cuda.cu:115); pix_x / pix_y are monotone in ix / iy
          while (x_lo <= x_hi && pix_x(g, x_lo) < fc.xmin) x_lo++;
          while (x_lo <= x_hi && pix_x(g, x_hi) >= fc.xmax) x_hi--;
          while (y_lo <= y_hi && pix_y(g, y_lo) >= fc.ymax) y_lo++;
          while (y_lo <= y_hi && pix_y(g, y_hi) < fc.ymin) y_hi--;
          if (x_lo <= x_hi && y_lo <= y_hi) {
            sm.x[0][tid] = fc.ax; sm.x[1][tid] = fc.ay; sm.x[2][tid] = fc.bx;
            sm.x[3][tid] = fc.by; sm.x[4][tid] = fc.cx; sm.x[5][tid] = fc.cy;
            sm.z[0][tid] = z_a;
            sm.z[1][tid] = z_b;
            sm.z[2][tid] = z_c;
            sm.face[tid] = a.job.first ? f - (int)a.job.first[b] : f;
            // edge functions E_k(x, y) = A_k x + B_k y + C_k in the local frame; inside
            // <=> s E_k >= 0 for all k with s = sign(E_0 + E_1 + E_2) (twice the area)
            const float ax = fc.ax - X0, ay = fc.ay - Y0, bx = fc.bx - X0, by = fc.by - Y0;
            const float cx = fc.cx - X0, cy = fc.cy - Y0;
            const float A[3] = {by - cy, cy - ay, ay - by};
            const float Bc[3] = {cx - bx, ax - cx, bx - ax};
            const float C[3] = {bx * cy - by * cx, cx * ay - cy * ax, ax * by - ay * bx};
            const float area2 = (bx - ax) * (cy - ay) - (by - ay) * (cx - ax);
            const float sA = fabsf(A[0]) + fabsf(A[1]) + fabsf(A[2]);
            const float sB = fabsf(Bc[0]) + fabsf(Bc[1]) + fabsf(Bc[2]);
            const float R = fmaxf(fmaxf(fmaxf(fabsf(ax), fabsf(bx)), fabsf(cx)),
                                  fmaxf(fmaxf(fabsf(ay), fabsf(by)), fabsf(cy))) +
                            (float)BLK * fmaxf(pitch_x, pitch_y);
            // cull only where the arithmetic below is far inside its error bounds: a clear
            // orientation, coordinates of moderate size
            cull = fabsf(area2) > 0x1p-10f * sA * sB + 0x1p-6f &&
                   R < 4096.f * fminf(pitch_x, pitch_y);
            const float s = area2 < 0.f ? -1.f : 1.f;
#pragma unroll
            for (int k = 0; k < 3; k++) {
              // keep a pixel unless s E_k < -m_k: m_k = an eighth of a pixel step of E_k in x
              // and in y, plus 64 ulp of R^2, which bounds (with a factor > 8 to spare) both
              // the rounding of the reference's own w_k and of this evaluation
              const float m = 0.125f * (fabsf(A[k]) * pitch_x + fabsf(Bc[k]) * pitch_y) +
                              0x1p-18f * R * R + 0x1p-10f;
              ia[k] = 1.0f / (s * A[k] * pitch_x);
              sb[k] = s * Bc[k];
              scm[k] = s * C[k] + m;
            }
          } else {
            x_hi = x_lo - 1;  // empty
          }
        } else {
          x_lo = 0; x_hi = -1; y_lo = 0; y_hi = -1;
        }
      }
    }
    const int h = (x_lo <= x_hi && y_lo <= y_hi) ? y_hi - y_lo + 1 : 0;
    int rtotal;
    const int rbase = block_scan(h, sm.wsum, lane, warp, rtotal);
    sm.rpre[tid] = rbase;
    if (tid == 0) sm.rpre[kRasterThreads] = rtotal;
    __syncthreads();
    PH_MARK(ph_a);
    if (rtotal == 0) continue;  // uniform

    // ---- row batches: the faces whose first row falls into [bi, bi + 1) * kRowBatch
    for (int rb = 0; rb < rtotal; rb += kRowBatch) {
      // first / one-past-last face of the batch (binary search in the row prefix)
      int f_lo = 0, f_hi = 0;
#pragma unroll
      for (int st = kRasterThreads / 2; st > 0; st >>= 1) {
        if (sm.rpre[f_lo + st] < rb) f_lo += st;
        if (sm.rpre[f_hi + st] < rb + kRowBatch) f_hi += st;
      }
      if (sm.rpre[f_lo] < rb) f_lo++;
      if (sm.rpre[f_hi] < rb + kRowBatch) f_hi++;
      const int row0 = sm.rpre[f_lo];
      const int nrows = sm.rpre[f_hi] - row0;  // <= kRowBatch + BLK
      // ---- phase A2: the owning thread turns each of its rows into a span [xa, xa + len)
      const bool mine = tid >= f_lo && tid < f_hi && h > 0;
      int cand = 0;
      if (mine) {
        const int rxl = x_lo - bx0, rxh = x_hi - bx0;
        for (int i = 0; i < h; i++) {
          const int ry = y_lo - by0 + i;
          int xa = rxl, xb = rxh;
          if (cull) {
            const float yl = -(float)ry * pitch_y;
            float L = (float)rxl, U = (float)rxh;
#pragma unroll
            for (int k = 0; k < 3; k++) {
              const float t = -(sb[k] * yl + scm[k]) * ia[k];  // s E_k = -m_k crossing [pixels]
              if (ia[k] > 0.f) L = fmaxf(L, t); else U = fminf(U, t);
            }
            L = fminf(L, 1e6f);
            U = fmaxf(U, -1e6f);
            // xa stays within the block (it is packed into 8 bits below): a left crossing
            // beyond the rectangle means an empty span
            xa = min(max(rxl, (int)ceilf(L - 0.01f)), rxh + 1);
            xb = min(rxh, (int)floorf(U + 0.01f));
          }
          const int len = max(xb - xa + 1, 0);
          sm.span[rbase - row0 + i] = (unsigned)tid | ((unsigned)ry << 8) | ((unsigned)xa << 16) |
                                      ((unsigned)len << 24);
          cand += len;
        }
      }
      int ctotal;
      int crun = block_scan(cand, sm.wsum, lane, warp, ctotal);
      if (mine) {
        for (int i = 0; i < h; i++) {
          sm.spre[rbase - row0 + i] = crun;
          crun += (int)(sm.span[rbase - row0 + i] >> 24);
        }
      }
      if (tid == 0) sm.spre[nrows] = ctotal;
      __syncthreads();
      PH_MARK(ph_a2);

      // ---- phase B: the batch's candidates dealt out evenly, a contiguous run per thread;
      // every candidate takes the reference's exact test (the spans hold few misses)
      if (ctotal > 0) {
        const int per = (ctotal + kRasterThreads - 1) / kRasterThreads;
        int k = tid * per;
        const int k_end = min(k + per, ctotal);
        if (k < k_end) {
          int r = 0;
          for (int st = 2048; st > 0; st >>= 1)  // nrows < 4096
            if (r + st <= nrows && sm.spre[r + st] <= k) r += st;
          unsigned sp = sm.span[r];
          int slot = sp & 0xff, ry = (sp >> 8) & 0xff;
          int rx = ((sp >> 16) & 0xff) + (k - sm.spre[r]);
          int left = sm.spre[r + 1] - k;  // candidates left in this row, this one included
          float ax = sm.x[0][slot], ay = sm.x[1][slot], bx = sm.x[2][slot];
          float by = sm.x[3][slot], cx = sm.x[4][slot], cy = sm.x[5][slot];
          float y0 = pix_y(g, by0 + ry);
          while (true) {
            const float x0 = pix_x(g, bx0 + rx);
            float w0, w1, w2;
            if (raster_weights(x0, y0, ax, ay, bx, by, cx, cy, a.eps_d, w0, w1, w2)) {
              const unsigned long long key = depth_key(
                  raster_depth(w0, w1, w2, sm.z[0][slot], sm.z[1][slot], sm.z[2][slot]),
                  sm.face[slot]);
              if (key) smem_max64(&zbuf[(ry << BLK_LOG2) + rx], key);
            }
            if (++k >= k_end) break;
            rx++;
            if (--left == 0) {
              do { r++; } while (sm.spre[r + 1] == sm.spre[r]);
              sp = sm.span[r];
              const int ns = sp & 0xff;
              ry = (sp >> 8) & 0xff;
              rx = (sp >> 16) & 0xff;
              left = sp >> 24;
              y0 = pix_y(g, by0 + ry);
              if (ns != slot) {
                slot = ns;
                ax = sm.x[0][slot]; ay = sm.x[1][slot]; bx = sm.x[2][slot];
                by = sm.x[3][slot]; cx = sm.x[4][slot]; cy = sm.x[5][slot];
              }
            }
          }
        }
      }
      __syncthreads();  // spans / staged faces free for the next batch / chunk
      PH_MARK(ph_b);
    }
  }
  __syncthreads();

  // ---- resolve: one pixel per thread-iteration, rows of the block are contiguous in x.
  // Everything a pixel needs hangs off its winner's id: the records of all of a thread's
  // pixels are requested into L1 first, so that the pass pays one memory round trip and
  // not one per pixel and tensor.
  for (int i = tid; i < BLK * BLK; i += kRasterThreads) {
    const unsigned long long key = zbuf[i];
    if (key && !a.job.faces) {
      const int face = key_face(key);
      const long long gi = (long long)b * a.job.view_stride +
                           (a.job.first ? face + (int)a.job.first[b] : face);
      prefetch_l1(a.job.img + gi * 6);
      const float* fp = a.feat + gi * 3 * a.D;
      prefetch_l1(fp);
      if (a.D > 2) prefetch_l1(fp + 3 * a.D - 1);
    }
  }
  // uncovered pixels by quarter row; the quarter COLUMN of a thread's pixels is fixed
  // (i = tid + k * kRasterThreads and BLK divides kRasterThreads)
  static_assert(kRasterThreads % BLK == 0, "quarter column must be fixed per thread");
  int unc_top = 0, unc_bot = 0;
  for (int i = tid; i < BLK * BLK; i += kRasterThreads) {
    const int ix = bx0 + (i & (BLK - 1)), iy = by0 + (i >> BLK_LOG2);
    if (ix >= p.W || iy >= p.H) continue;
    const long long pix = ((long long)b * p.H + iy) * p.W + ix;
    const unsigned long long key = zbuf[i];
    if (!key) {
      a.out_idx[pix] = -1;
      for (int d = 0; d < a.D; d++) a.out_feat[pix * a.D + d] = 0.f;
      if (a.out_w) { a.out_w[pix * 3] = 0.f; a.out_w[pix * 3 + 1] = 0.f; a.out_w[pix * 3 + 2] = 0.f; }
      if (a.mask) a.mask[pix] = 0.f;   // 1 - 1 (dibr_soft_mask_cuda.cu:181-182 with no face held)
      if ((i >> BLK_LOG2) >> (BLK_LOG2 - 1)) unc_bot++; else unc_top++;
      continue;
    }
    const int face = key_face(key);
    const int f = a.job.first ? face + (int)a.job.first[b] : face;
    const float *fp0, *fp1, *fp2;   // features of the three corners
    if (a.job.faces) {
      int v[3];
      mesh_face(a.job, f, v);   // listed faces have valid ids
      const float* vb = a.feat + (a.feat_shared ? 0ll : (long long)b * a.job.V) * a.D;
      fp0 = vb + (long long)v[0] * a.D; fp1 = vb + (long long)v[1] * a.D;
      fp2 = vb + (long long)v[2] * a.D;
    } else {
      fp0 = a.feat + ((long long)b * a.job.view_stride + f) * 3 * a.D;
      fp1 = fp0 + a.D;
      fp2 = fp1 + a.D;
    }
    // features before any store (the stores may alias them as far as the compiler knows)
    float r[3][4];
    const int dn = a.D < 4 ? a.D : 4;
#pragma unroll
    for (int d = 0; d < 4; d++) {
      if (d < dn) {
        r[0][d] = __ldg(fp0 + d); r[1][d] = __ldg(fp1 + d); r[2][d] = __ldg(fp2 + d);
      }
    }
    Face2D fc;
    load_face2d<true>(a.job, b, f, fc);
    float w0, w1, w2;
    raster_weights(pix_x(g, ix), pix_y(g, iy), fc.ax, fc.ay, fc.bx, fc.by, fc.cx, fc.cy, a.eps_d,
                   w0, w1, w2);
    a.out_idx[pix] = face;
    if (a.mask) a.mask[pix] = 1.f;     // covered (dibr_soft_mask_cuda.cu:68-70)
    if (a.out_w) { a.out_w[pix * 3] = w0; a.out_w[pix * 3 + 1] = w1; a.out_w[pix * 3 + 2] = w2; }
#pragma unroll
    for (int d = 0; d < 4; d++)
      if (d < dn)
        a.out_feat[pix * a.D + d] =
            __fmaf_rn(w2, r[2][d], __fmaf_rn(w0, r[0][d], __fmul_rn(w1, r[1][d])));
    for (int d = 4; d < a.D; d++) {
      const float r0 = __ldg(fp0 + d), r1 = __ldg(fp1 + d), r2 = __ldg(fp2 + d);
      a.out_feat[pix * a.D + d] = __fmaf_rn(w2, r2, __fmaf_rn(w0, r0, __fmul_rn(w1, r1)));
    }
  }
  if (a.block_uncovered) {
    const int qx = (tid & (BLK - 1)) >> (BLK_LOG2 - 1);
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int u = __reduce_add_sync(DIBR_FULL_MASK, qx == (q & 1) ? ((q >> 1) ? unc_bot : unc_top) : 0);
      if (lane == 0 && u) atomicAdd(&sm.unc4[q], u);
    }
    __syncthreads();
    if (tid == 0) {
      const int total = sm.unc4[0] + sm.unc4[1] + sm.unc4[2] + sm.unc4[3];
      a.block_uncovered[(long long)b * p.nkx * p.nky + byi * p.nkx + bxi] = total;
      // hand the quarters that hold uncovered pixels to the soft mask's gather -- unless no
      // face box reaches this bin: those pixels keep the 0 written above
      if (a.work.item && total > 0 && a.soft_count[row] > 0) {
        int nq = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) nq += sm.unc4[q] > 0;
        unsigned slot = atomicAdd(a.work.count, (unsigned)nq);
#pragma unroll
        for (int q = 0; q < 4; q++)
          if (sm.unc4[q] > 0) {
            if (slot < a.work.cap)
              a.work.item[slot] =
                  make_int2(b, (bxi * 2 + (q & 1)) | ((byi * 2 + (q >> 1)) << 16));
            slot++;
          }
      }
    }
  }
#ifdef DIBR_TRACE_PHASES
  {
    __syncthreads();
    const long long ph_r = clock64() - ph_t;
    const unsigned long long u = ((unsigned long long)min(ph_a >> 6, 16383ll) << 42) |
                                 ((unsigned long long)min(ph_a2 >> 6, 16383ll) << 28) |
                                 ((unsigned long long)min(ph_b >> 6, 16383ll) << 14) |
                                 (unsigned long long)min(ph_r >> 6, 16383ll);
    DIBR_TRACE_END(u);
  }
#elif !defined(DIBR_TRACE_GATHER)   // (the gather's trace shares the buffer: one kernel at a time)
  DIBR_TRACE_END(n);
#endif
}

cudaError_t launch_raster_forward(const RasterArgs& a, const Plan& p, cudaStream_t st) {
  const int shift = p.bin_log2 - p.blk_log2;
  dim3 grid((unsigned)(((long long)p.B * p.NB) << (2 * shift)));
  prof_begin(K_RASTER_FWD, st);
  if (p.blk_log2 == 5) {
    launch_k(raster_forward_kernel<5>, grid, kRasterThreads, sizeof(RasterSmem<5>), st, a, p);
  } else {
    const size_t smem = sizeof(RasterSmem<6>);  // above the 48 KB default: opt in
    cudaError_t e = smem_optin((const void*)raster_forward_kernel<6>, (int)smem);
    if (e != cudaSuccess) return e;
    launch_k(raster_forward_kernel<6>, grid, kRasterThreads, smem, st, a, p);
  }
  prof_end(st);
  return cudaGetLastError();
}

// ===========================================================================
// Backward (rasterization_cuda.cu:238-402).
//
// The reference issues 9*D global atomics per covered pixel and ~200 flops.  Its
// per-pixel terms are, however, affine in the pixel position: with (s, t) = pixel - vertex
// a (unscaled coordinates) and the face constants m, p, n, q, k3 (:305-322),
//     w1 = (s q - n t) / k3,  w2 = (m t - s p) / k3,  w0 = 1 - w1 - w2,
// and every d(w1|w2)/d(vertex) (:324-374) is (const + const s + const t) / k3^2.  So all a
// face needs from its pixels are the 3*D MOMENTS  M0_d = sum g_d, Ms_d = sum g_d s,
// Mt_d = sum g_d t  of the incoming gradient g over the pixels that selected it -- and the
// map from moments to both gradients is LINEAR, so it can be applied to partial sums.
// One pass, pixel-parallel: a warp owns an 8x4 region; lanes with the same face are grouped
// with __match_any_sync, their moments summed with shuffles, and the group's first lane
// turns the partial moments into gradient contributions and issues ONE
// red.global.add per output element: 3*D + 6 per (warp, face) instead of 9*D per pixel, no
// per-face accumulator, no second pass.  In mesh mode the contributions go straight to the
// vertex gradients through `faces`.
// Precision.  The gradient is a function of the pixel CENTRE; the reference reconstructs the
// centre from the fp32 weights its forward saved (:302-303), whose rounding -- and that of
// fl32(v * multiplier) -- is a 1e-4 .. 1e-2 relative perturbation of (s, t) for sub-pixel
// faces (it is the reference's fp32 noise against its own fp64 run).  Here (s, t) come from
// the pixel centre itself, formed as the kernel forms it (:85-86) and differenced in double,
// which is what the reference computes when run in double.  The eps of the forward's
// normalisation (:140-142) scales all three weights by rho = |N| / (|N| + eps), N = the
// face's area term in multiplied coordinates; the reconstructed centre is rho * X: both
// are carried through (they matter for small multipliers).
// Faces for which fp32 moment sums would not stay within a few ulp -- slivers / zero area
// (|k3| < 10% of |m q| + |n p|: k3 itself cancels) and faces under two pixels
// (|k3| < 4 pixel areas; they own one or two pixels and gain nothing from grouping) -- take
// the reference's per-pixel formula, evaluated in DOUBLE from the unrounded inputs; so does
// every face when the caller hands the forward's weights (the raw _C op).
// The same kernel instantiated for double serves the reference's double dispatch
// (rasterization_cuda.cu:30-39): moments and sums in double, pixel centre from the saved
// double weights (that entry point has no image geometry).
// ===========================================================================

// face constants of the backward (unscaled coordinates); regular = moment form applies
struct BwdFace {
  float ax, ay, bx, by, cx, cy;
  float m, p, n, q, k3raw;
  int v[3];        // mesh mode: vertex ids
  bool regular;
};
__device__ __forceinline__ BwdFace bwd_face(const RasterBwdArgs& a, int b, int face, float min_k3) {
  BwdFace f;
  float2 A, Bv, C;
  bool ok = true;
  if (a.faces) {   // mesh mode: a.img is vertices_image (B, V, 2)
    const int* fp = a.faces + (long long)face * 3;
    f.v[0] = __ldg(fp); f.v[1] = __ldg(fp + 1); f.v[2] = __ldg(fp + 2);
    ok = (unsigned)f.v[0] < (unsigned)a.V && (unsigned)f.v[1] < (unsigned)a.V &&
         (unsigned)f.v[2] < (unsigned)a.V;
    if (!ok) f.v[0] = f.v[1] = f.v[2] = 0;
    const float2* vp = reinterpret_cast<const float2*>(a.img) + (long long)b * a.V;
    A = __ldg(vp + f.v[0]); Bv = __ldg(vp + f.v[1]); C = __ldg(vp + f.v[2]);
  } else {
    f.v[0] = f.v[1] = f.v[2] = 0;
    const float2* ip = reinterpret_cast<const float2*>(a.img + ((long long)b * a.F + face) * 6);
    A = __ldg(ip); Bv = __ldg(ip + 1); C = __ldg(ip + 2);
  }
  f.ax = A.x; f.ay = A.y; f.bx = Bv.x; f.by = Bv.y; f.cx = C.x; f.cy = C.y;
  f.m = Bv.x - A.x; f.p = Bv.y - A.y; f.n = C.x - A.x; f.q = C.y - A.y;
  f.k3raw = f.m * f.q - f.n * f.p;
  // (false for NaN / Inf coordinates as well)
  f.regular = ok && fabsf(f.k3raw) >= 0.1f * (fabsf(f.m * f.q) + fabsf(f.n * f.p)) &&
              fabsf(f.k3raw) >= min_k3 && fabsf(f.k3raw) < 3e38f;
  return f;
}
// features / gradient slots of corner k of a face, either layout.  The view's base (uniform
// over the CTA: b = blockIdx.z) is kept apart from the per-corner offset, a 32 x 32 -> 64 bit
// product: the nine pointers of a face were 15 % of the kernel's instructions as 64-bit
// multiplies of (view offset + vertex id).
__device__ __forceinline__ const float* bwd_feat(const RasterBwdArgs& a, int b, int face,
                                                 const BwdFace& f, int k) {
  if (a.faces)
    return a.feat + (a.feat_shared ? 0ll : (long long)b * a.V * a.D) + (long long)f.v[k] * a.D;
  return a.feat + (long long)b * a.F * 3 * a.D + (long long)(face * 3 + k) * a.D;
}
__device__ __forceinline__ float* bwd_gfeat(const RasterBwdArgs& a, int b, int face,
                                            const BwdFace& f, int k) {
  if (a.faces)
    return a.grad_feat + (a.feat_shared ? 0ll : (long long)b * a.V * a.D) +
           (long long)f.v[k] * a.D;
  return a.grad_feat + (long long)b * a.F * 3 * a.D + (long long)(face * 3 + k) * a.D;
}
__device__ __forceinline__ float* bwd_gimg(const RasterBwdArgs& a, int b, int face,
                                           const BwdFace& f, int k) {
  if (a.faces) return a.grad_img + (long long)b * a.V * 2 + f.v[k] * 2ll;
  return a.grad_img + (long long)b * a.F * 6 + (face * 3 + k) * 2ll;
}

// The reference's per-pixel backward (:273-399) for one pixel, in double, from the weights w:
// grad_face_features += g w_k, grad_face_vertices_image += (sum over d) dldI d w_{1,2} / d v.
// T = storage type of the tensors (float, or double for the reference's double dispatch).
template <typename T>
__device__ __forceinline__ void per_pixel_backward(const T* g, const double (&w)[3], double ax,
                                                   double ay, double bx, double by, double cx,
                                                   double cy, double eps_d, int D, const T* fp0,
                                                   const T* fp1, const T* fp2, T* gf0, T* gf1,
                                                   T* gf2, T* gi0, T* gi1, T* gi2) {
  const double x0 = w[0] * ax + w[1] * bx + w[2] * cx;   // :302-303
  const double y0 = w[0] * ay + w[1] * by + w[2] * cy;
  const double m = bx - ax, pp = by - ay, n = cx - ax, q = cy - ay, ss = x0 - ax, tt = y0 - ay;
  const double k1 = ss * q - n * tt;
  const double k2 = m * tt - ss * pp;
  double k3 = m * q - n * pp;
  k3 += copysign(eps_d, k3);
  const double dw1dm = -q * k1, dw1dn = -tt * k3 + pp * k1;
  const double dw1dp = n * k1, dw1dq = ss * k3 - m * k1;
  const double dw1ds = q * k3, dw1dt = -n * k3;
  const double dw2dm = tt * k3 - q * k2, dw2dn = pp * k2;
  const double dw2dp = -ss * k3 + n * k2, dw2dq = -m * k2;
  const double dw2ds = -pp * k3, dw2dt = m * k3;
  double e1 = 0.0, e2 = 0.0;                 // sum over d of g_d (c1 - c0), g_d (c2 - c0)
  for (int d = 0; d < D; d++) {
    const double gd = (double)g[d];
    atomicAdd(gf0 + d, (T)(gd * w[0]));
    atomicAdd(gf1 + d, (T)(gd * w[1]));
    atomicAdd(gf2 + d, (T)(gd * w[2]));
    const double c0 = (double)fp0[d], c1 = (double)fp1[d], c2 = (double)fp2[d];
    e1 += gd * (c1 - c0);
    e2 += gd * (c2 - c0);
  }
  const double r = 1.0 / (k3 * k3);
  atomicAdd(gi0, (T)((e1 * -(dw1dm + dw1dn + dw1ds) + e2 * -(dw2dm + dw2dn + dw2ds)) * r));
  atomicAdd(gi0 + 1, (T)((e1 * -(dw1dp + dw1dq + dw1dt) + e2 * -(dw2dp + dw2dq + dw2dt)) * r));
  atomicAdd(gi1, (T)((e1 * dw1dm + e2 * dw2dm) * r));
  atomicAdd(gi1 + 1, (T)((e1 * dw1dp + e2 * dw2dp) * r));
  atomicAdd(gi2, (T)((e1 * dw1dn + e2 * dw2dn) * r));
  atomicAdd(gi2 + 1, (T)((e1 * dw1dq + e2 * dw2dq) * r));
}

template <typename T, int N>
__device__ __forceinline__ void reduce_peers(unsigned peers, T (&v)[N], int lane) {
  int rel_pos = __popc(peers & lanemask_lt());
  unsigned rem = peers & ~(lanemask_lt() | (1u << lane));  // peers with a higher lane id
  while (__any_sync(DIBR_FULL_MASK, rem != 0)) {
    const int next = __ffs(rem);
    const int src = next ? next - 1 : lane;
#pragma unroll
    for (int i = 0; i < N; i++) {
      const T t = __shfl_sync(DIBR_FULL_MASK, v[i], src);
      if (next) v[i] += t;
    }
    const bool done = rel_pos & 1;
    rem &= __ballot_sync(DIBR_FULL_MASK, !done);
    rel_pos >>= 1;
  }
}

// moments of a (warp, face) group -> gradient contributions (:324-374 with (1, s, t) replaced
// by their sums; all in double) and one atomic per output element.  M[0..D) = sum g_d,
// M[DC..) = sum g_d s, M[2 DC..) = sum g_d t with (s, t) measured from vertex a WITHOUT the
// eps scaling (sl, tl); rho folds it in.
template <int DC>
__device__ __forceinline__ void emit_group(const RasterBwdArgs& a, int b, int face,
                                           const BwdFace& fc, const float (&M)[3 * DC], int d0) {
  // Regular faces only (|k3| >= 10% of |m q| + |n p|, >= 4 pixel areas): every quotient below
  // is well conditioned, so float carries it (the cancellation inside the K and d terms
  // amplifies float rounding by at most ~1e2: 1e-5 relative).  Face constants in float from
  // the float inputs; only k3's eps and rho, which differ from 1 by ~1e-8, go through double.
  // (edge vectors and k3 differenced in double first: a float difference of two coordinates
  // near 1 would already cost 1e-5 of a two-pixel face at 2048^2)
  const double md = (double)fc.bx - (double)fc.ax, pd = (double)fc.by - (double)fc.ay;
  const double nd = (double)fc.cx - (double)fc.ax, qd = (double)fc.cy - (double)fc.ay;
  const double k3d = md * qd - nd * pd;
  const float m = (float)md, p = (float)pd, n = (float)nd, q = (float)qd, k3raw = (float)k3d;
  const float k3 = (float)(k3d + copysign(a.eps_d, k3d));                          // :321-322
  const float inv_k3raw = 1.0f / k3raw;
  // eps of the forward's normalisation (:140-142), in multiplied coordinates:
  // rho = N / (N + eps), N = |k3raw| mult^2; rho - 1 = -eps / (N + eps) kept separately
  // (rho differs from 1 by eps / N ~ 1e-8 .. 1e-4: a float quotient carries it)
  const float Nf = fabsf(k3raw) * a.mult * a.mult;
  const float rho1 = -a.eps_f / (Nf + a.eps_f), rho = 1.0f + rho1;
  const float *fp0 = bwd_feat(a, b, face, fc, 0), *fp1 = bwd_feat(a, b, face, fc, 1);
  const float* fp2 = bwd_feat(a, b, face, fc, 2);
  float *gf0 = bwd_gfeat(a, b, face, fc, 0), *gf1 = bwd_gfeat(a, b, face, fc, 1);
  float* gf2 = bwd_gfeat(a, b, face, fc, 2);
  float P0 = 0.f, Ps = 0.f, Pt = 0.f, Q0 = 0.f, Qs = 0.f, Qt = 0.f;
#pragma unroll
  for (int dd = 0; dd < DC; dd++) {
    const int d = d0 + dd;
    if (d >= a.D) break;
    const float M0 = M[dd], Msl = M[DC + dd], Mtl = M[2 * DC + dd];
    // moments of the reconstructed centre rho * X - a = rho (X - a) + (rho - 1) a
    const float Ms = rho * Msl + rho1 * fc.ax * M0;
    const float Mt = rho * Mtl + rho1 * fc.ay * M0;
    const float c0 = __ldg(fp0 + d), c1 = __ldg(fp1 + d), c2 = __ldg(fp2 + d);
    const float e1 = c1 - c0, e2 = c2 - c0;
    P0 += e1 * M0; Ps += e1 * Ms; Pt += e1 * Mt;
    Q0 += e2 * M0; Qs += e2 * Ms; Qt += e2 * Mt;
    // grad_face_features: sum g w_k with the forward's weights w1 = rho (sl q - n tl) / k3raw,
    // w2 = rho (m tl - sl p) / k3raw, w0 = rho - w1 - w2
    const float g1 = rho * (Msl * q - n * Mtl) * inv_k3raw;
    const float g2 = rho * (m * Mtl - Msl * p) * inv_k3raw;
    atomicAdd(gf0 + d, rho * M0 - g1 - g2);
    atomicAdd(gf1 + d, g1);
    atomicAdd(gf2 + d, g2);
  }
  const float K1 = Ps * q - n * Pt, K2 = m * Qt - Qs * p;
  const float dm = -q * K1 + (Qt * k3 - q * K2);
  const float dn = (-Pt * k3 + p * K1) + p * K2;
  const float dp = n * K1 + (-Qs * k3 + n * K2);
  const float dq = (Ps * k3 - m * K1) - m * K2;
  const float ds = q * k3 * P0 - p * k3 * Q0;
  const float dt = -n * k3 * P0 + m * k3 * Q0;
  const float r = 1.0f / (k3 * k3);
  atomicAdd(reinterpret_cast<float2*>(bwd_gimg(a, b, face, fc, 0)),
            make_float2(-(dm + dn + ds) * r, -(dp + dq + dt) * r));
  atomicAdd(reinterpret_cast<float2*>(bwd_gimg(a, b, face, fc, 1)), make_float2(dm * r, dp * r));
  atomicAdd(reinterpret_cast<float2*>(bwd_gimg(a, b, face, fc, 2)), make_float2(dn * r, dq * r));
}

// DIBR_BWD_CTA=1 (experiment, off): the (warp, face) groups of a CTA are merged per face in a
// shared-memory table and converted one face per thread with dense lanes.  Measured on cfg 2:
// 98.6 us against 71.0 us for the default (every group leader converts its own partial
// moments): three barriers and a conversion phase with one or two busy warps per CTA cost
// more than the saved instructions.
#ifndef DIBR_BWD_CTA
#define DIBR_BWD_CTA 0
#endif
#if DIBR_BWD_CTA
constexpr int kBwdSlots = 512;   // >= 2 x the pixels of a CTA: at most 256 distinct faces
#endif

#ifndef DIBR_BWD_MINB
#define DIBR_BWD_MINB 4
#endif
template <int DC>  // features handled per pass (D <= 4: one pass)
__global__ void __launch_bounds__(256, DIBR_BWD_MINB) raster_backward_kernel(const RasterBwdArgs a) {
#if DIBR_BWD_CTA
  // per-CTA face table: open addressing on the face id; `s_list` holds the claimed slots in
  // claim order so that the conversion runs one face per THREAD with dense lanes
  __shared__ int s_key[kBwdSlots];
  __shared__ float s_mom[kBwdSlots * 3 * DC];
  __shared__ int s_list[256];
  __shared__ int s_count;
#endif
  pdl_prologue();
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const int ix = (blockIdx.x << 4) + ((warp & 1) << 3) + (lane & 7);
  const int iy = (blockIdx.y << 4) + ((warp >> 1) << 2) + (lane >> 3);
  const int b = blockIdx.z;
  const bool inimg = ix < a.W && iy < a.H;
  const long long pix = ((long long)b * a.H + iy) * a.W + ix;
  int face = -1;
  if (inimg) face = (int)a.idx[pix];
#if DIBR_BWD_CTA
  if (!__syncthreads_or(face >= 0)) return;   // nothing covered in this 16 x 16 block
#else
  if (!__any_sync(DIBR_FULL_MASK, face >= 0)) return;
#endif
  const unsigned peers = __match_any_sync(DIBR_FULL_MASK, face);
  const bool leader = (peers & lanemask_lt()) == 0;

  float s = 0.f, t = 0.f;
  bool regular = false;
  BwdFace fc;
  if (face >= 0) {
    fc = bwd_face(a, b, face, 16.f / ((float)a.W * (float)a.H));
    regular = fc.regular && !a.weights;   // given weights (raw op): per pixel, as the reference
    // pixel centre as the forward forms it (:85-86), back in unscaled coordinates; minus
    // vertex a in double: for faces much smaller than the image the float rounding of the
    // centre itself (~6e-8) would be a visible fraction of (s, t)
    s = (float)((double)pix_x(a.grid, ix) * a.inv_mult - (double)fc.ax);
    t = (float)((double)pix_y(a.grid, iy) * a.inv_mult - (double)fc.ay);
  }
  for (int d0 = 0; d0 < a.D; d0 += DC) {
    float mom[3 * DC];
#pragma unroll
    for (int i = 0; i < 3 * DC; i++) mom[i] = 0.f;
    if (regular) {
#pragma unroll
      for (int dd = 0; dd < DC; dd++) {
        if (d0 + dd < a.D) {
          const float gd = a.grad[pix * a.D + d0 + dd];
          mom[dd] = gd;
          mom[DC + dd] = gd * s;
          mom[2 * DC + dd] = gd * t;
        }
      }
    }
    reduce_peers<float, 3 * DC>(peers, mom, lane);
#if DIBR_BWD_CTA
    // ---- merge the CTA's groups per face: leaders add their partial moments into the
    // face's slot (shared-memory atomics), then one thread per distinct face converts
    for (int i = threadIdx.x; i < kBwdSlots; i += 256) s_key[i] = -1;
    for (int i = threadIdx.x; i < kBwdSlots * 3 * DC; i += 256) s_mom[i] = 0.f;
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    if (leader && regular) {
      unsigned h = ((unsigned)face * 2654435761u) >> 23;   // 9 bits
      while (true) {
        const int prev = atomicCAS(&s_key[h], -1, face);
        if (prev == -1) { s_list[atomicAdd(&s_count, 1)] = (int)h; break; }
        if (prev == face) break;
        h = (h + 1) & (kBwdSlots - 1);
      }
#pragma unroll
      for (int i = 0; i < 3 * DC; i++)
        if (mom[i] != 0.f) atomicAdd(&s_mom[h * 3 * DC + i], mom[i]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < s_count; i += 256) {
      const int slot = s_list[i], f2 = s_key[slot];
      const BwdFace fc2 = bwd_face(a, b, f2, 16.f / ((float)a.W * (float)a.H));
      float M[3 * DC];
#pragma unroll
      for (int j = 0; j < 3 * DC; j++) M[j] = s_mom[slot * 3 * DC + j];
      emit_group<DC>(a, b, f2, fc2, M, d0);
    }
    __syncthreads();   // table free for the next feature chunk
#else
    if (leader && regular) emit_group<DC>(a, b, face, fc, mom, d0);
#endif
  }

  // ---- thin / tiny faces: the reference's per-pixel formula (:273-399) in double
  if (face >= 0 && !regular) {
    const double mult = (double)a.mult;
    const double ax = fc.ax, ay = fc.ay, bx = fc.bx, by = fc.by, cx = fc.cx, cy = fc.cy;
    // forward weights (:131-145) in multiplied coordinates, eps included.  Where the area
    // term is eps-dominated or pure cancellation (collinear corners) the weights ARE the fp32
    // rounding noise the reference's forward saved: those are recomputed bit-identically in
    // float; everywhere else in double.
    double w[3];
    if (a.weights) {
      w[0] = a.weights[pix * 3]; w[1] = a.weights[pix * 3 + 1]; w[2] = a.weights[pix * 3 + 2];
    } else {
      const double x0s = (double)pix_x(a.grid, ix), y0s = (double)pix_y(a.grid, iy);
      const double aex = ax * mult - x0s, aey = ay * mult - y0s;
      const double bex = bx * mult - x0s, bey = by * mult - y0s;
      const double cex = cx * mult - x0s, cey = cy * mult - y0s;
      w[0] = bex * cey - bey * cex; w[1] = cex * aey - cey * aex; w[2] = aex * bey - aey * bex;
      double norm = w[0] + w[1] + w[2];
      if (fabs(norm) > 1e-4 * (fabs(w[0]) + fabs(w[1]) + fabs(w[2])) && fabs(norm) > 4.0 * a.eps_d) {
        norm += copysign(a.eps_d, norm);
        w[0] /= norm; w[1] /= norm; w[2] /= norm;
      } else {
        float f0, f1, f2;
        raster_weights(pix_x(a.grid, ix), pix_y(a.grid, iy), __fmul_rn(fc.ax, a.mult),
                       __fmul_rn(fc.ay, a.mult), __fmul_rn(fc.bx, a.mult), __fmul_rn(fc.by, a.mult),
                       __fmul_rn(fc.cx, a.mult), __fmul_rn(fc.cy, a.mult), a.eps_d, f0, f1, f2);
        w[0] = f0; w[1] = f1; w[2] = f2;
      }
    }
    per_pixel_backward<float>(a.grad + pix * a.D, w, ax, ay, bx, by, cx, cy, a.eps_d, a.D,
                              bwd_feat(a, b, face, fc, 0), bwd_feat(a, b, face, fc, 1),
                              bwd_feat(a, b, face, fc, 2), bwd_gfeat(a, b, face, fc, 0),
                              bwd_gfeat(a, b, face, fc, 1), bwd_gfeat(a, b, face, fc, 2),
                              bwd_gimg(a, b, face, fc, 0), bwd_gimg(a, b, face, fc, 1),
                              bwd_gimg(a, b, face, fc, 2));
  }
}

// ---- double dispatch of the raw op (rasterization_cuda.cu:30-39 with scalar_t = double):
// unscaled double coordinates and the forward's saved double weights, no image geometry.
// Same design as above with everything in double: per pixel the centre (s, t) is rebuilt
// from the weights (:302-303, exact to 1e-16 in double); lanes of a warp that hold the same
// face sum 6 numbers -- E1 = sum_d g_d (c1 - c0) and E2 = sum_d g_d (c2 - c0), each times
// (1, s, t) -- and 3 * D feature-gradient terms g_d w_k with shuffles; the group's first
// lane converts the six sums into the coordinate gradient (the formula is affine in (s, t))
// and issues one atomicAdd(double) per output element.
constexpr int kF64Chunk = 4;   // feature channels reduced per round

__global__ void __launch_bounds__(256) raster_backward_f64_kernel(const RasterBwdF64Args a) {
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const int ix = (blockIdx.x << 4) + ((warp & 1) << 3) + (lane & 7);
  const int iy = (blockIdx.y << 4) + ((warp >> 1) << 2) + (lane >> 3);
  const int b = blockIdx.z;
  const bool inimg = ix < a.W && iy < a.H;
  const long long pix = ((long long)b * a.H + iy) * a.W + ix;
  int face = -1;
  if (inimg) face = (int)a.idx[pix];
  if (!__any_sync(DIBR_FULL_MASK, face >= 0)) return;
  const unsigned peers = __match_any_sync(DIBR_FULL_MASK, face);
  const bool leader = (peers & lanemask_lt()) == 0 && face >= 0;
  const long long gi = (long long)b * a.F + max(face, 0);
  const double* v = a.img + gi * 6;
  const double* fp = a.feat + gi * 3 * a.D;
  const double ax = v[0], ay = v[1], bx = v[2], by = v[3], cx = v[4], cy = v[5];
  double w[3] = {0., 0., 0.}, ss = 0., tt = 0.;
  if (face >= 0) {
    w[0] = a.weights[pix * 3]; w[1] = a.weights[pix * 3 + 1]; w[2] = a.weights[pix * 3 + 2];
    ss = w[0] * ax + w[1] * bx + w[2] * cx - ax;   // :302-303, :311-312
    tt = w[0] * ay + w[1] * by + w[2] * cy - ay;
  }
  // ---- feature gradient: sum over the group's pixels of g_d w_k, kF64Chunk channels a round
  double E1 = 0., E2 = 0.;
  for (int d0 = 0; d0 < a.D; d0 += kF64Chunk) {
    double acc[3 * kF64Chunk];
#pragma unroll
    for (int i = 0; i < 3 * kF64Chunk; i++) acc[i] = 0.;
    if (face >= 0) {
#pragma unroll
      for (int dd = 0; dd < kF64Chunk; dd++) {
        const int d = d0 + dd;
        if (d < a.D) {
          const double gd = a.grad[pix * a.D + d];
          acc[dd] = gd * w[0];
          acc[kF64Chunk + dd] = gd * w[1];
          acc[2 * kF64Chunk + dd] = gd * w[2];
          const double c0 = fp[d];
          E1 += gd * (fp[a.D + d] - c0);
          E2 += gd * (fp[2 * a.D + d] - c0);
        }
      }
    }
    reduce_peers<double, 3 * kF64Chunk>(peers, acc, lane);
    if (leader) {
      double* gf = a.grad_feat + gi * 3 * a.D;
#pragma unroll
      for (int dd = 0; dd < kF64Chunk; dd++) {
        const int d = d0 + dd;
        if (d < a.D) {
          atomicAdd(gf + d, acc[dd]);
          atomicAdd(gf + a.D + d, acc[kF64Chunk + dd]);
          atomicAdd(gf + 2 * a.D + d, acc[2 * kF64Chunk + dd]);
        }
      }
    }
  }
  // ---- coordinate gradient from the six sums
  double S[6] = {E1, E1 * ss, E1 * tt, E2, E2 * ss, E2 * tt};
  reduce_peers<double, 6>(peers, S, lane);
  if (leader) {
    const double m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay;
    double k3 = m * q - n * p;
    k3 += copysign(a.eps_d, k3);                                            // :319-322
    const double P0 = S[0], Ps = S[1], Pt = S[2], Q0 = S[3], Qs = S[4], Qt = S[5];
    const double K1 = Ps * q - n * Pt, K2 = m * Qt - Qs * p;               // sums of e k1, e k2
    const double dm = -q * K1 + (Qt * k3 - q * K2);
    const double dn = (-Pt * k3 + p * K1) + p * K2;
    const double dp = n * K1 + (-Qs * k3 + n * K2);
    const double dq = (Ps * k3 - m * K1) - m * K2;
    const double ds = q * k3 * P0 - p * k3 * Q0;
    const double dt = -n * k3 * P0 + m * k3 * Q0;
    const double r = 1.0 / (k3 * k3);
    double* gv = a.grad_img + gi * 6;
    atomicAdd(gv + 0, -(dm + dn + ds) * r);
    atomicAdd(gv + 1, -(dp + dq + dt) * r);
    atomicAdd(gv + 2, dm * r);
    atomicAdd(gv + 3, dp * r);
    atomicAdd(gv + 4, dn * r);
    atomicAdd(gv + 5, dq * r);
  }
}

cudaError_t launch_raster_backward_f64(const RasterBwdF64Args& a, cudaStream_t st) {
  if ((long long)a.B * a.H * a.W == 0) return cudaSuccess;
  dim3 grid((a.W + 15) / 16, (a.H + 15) / 16, a.B);
  raster_backward_f64_kernel<<<grid, 256, 0, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_raster_backward(const RasterBwdArgs& a, cudaStream_t st) {
  dim3 grid((a.W + 15) / 16, (a.H + 15) / 16, a.B);
  prof_begin(K_RASTER_BWD, st);
  switch (a.D >= 4 ? 4 : a.D) {
    case 1: launch_k(raster_backward_kernel<1>, grid, 256, 0, st, a); break;
    case 2: launch_k(raster_backward_kernel<2>, grid, 256, 0, st, a); break;
    case 3: launch_k(raster_backward_kernel<3>, grid, 256, 0, st, a); break;
    default: launch_k(raster_backward_kernel<4>, grid, 256, 0, st, a); break;
  }
  prof_end(st);
  return cudaGetLastError();
}

}  // namespace dibr
