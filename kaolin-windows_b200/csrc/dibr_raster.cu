// dibr_raster.cu -- block rasterizer forward (north_star (b)) and its backward ((d)).
//
// Forward: one CTA per screen block (32x32 or 64x64 pixels) of one view.  The block's
// depth buffer lives in shared memory as 64-bit keys (order-preserving depth | ~face).
// The bin's face list is consumed FACE-PARALLEL: each thread takes one face, scales it,
// finds the handful of pixel centres inside its box and runs the reference's exact
// edge/barycentric test only on those, publishing hits with a shared-memory 64-bit max.
// With meshes whose faces cover a few pixels this does ~10x less arithmetic than
// testing every pixel of the block against every face.  Faces whose box covers many
// pixels are queued in shared memory and rasterized by a whole warp each, lanes striding
// over the box.  A final pixel-parallel pass decodes the winner of every pixel,
// recomputes its weights with the same instruction sequence (bit-identical) and writes
// selected_face_idx / interpolated_features (/ output_weights) with coalesced stores;
// background pixels are written too, so the caller needs no memset.
#include "dibr_common.cuh"

namespace dibr {


constexpr int kRasterThreads = 256;
constexpr int kRasterWarps = kRasterThreads / 32;
constexpr int kQueue = 64;  // survivor queue entries per warp (drained 32 at a time)

template <int BLK_LOG2>
struct RasterSmem {
  static constexpr int BLK = 1 << BLK_LOG2;
  unsigned long long zbuf[BLK * BLK];
  float x[6][kRasterThreads];   // staged faces of the current chunk (scaled coordinates)
  float z[3][kRasterThreads];
  int rect[kRasterThreads];     // x_lo | y_lo << 8 | (w - 1) << 16 | (h - 1) << 24, block-relative
  int face[kRasterThreads];
  int pre[kRasterThreads + 1];  // exclusive prefix of the candidate counts w * h
  int wsum[kRasterWarps];
  unsigned queue[kRasterWarps][kQueue];
  int uncovered;
};

// Sign-only rejection of a candidate before the normalisation (exactness argument in
// DESIGN.md "Numerics"): the reference rejects when RN(w_k / norm) < 0.  norm has the sign
// bit of n32 = (w0 + w1) + w2 and |n32| <= |norm| <= 2^100 when |n32| <= 2^99 and
// |eps| <= 2^90, so for |w_k| >= 2^-20 with the opposite sign bit the quotient is a
// negative NORMAL number whatever its rounding: rejected.  Everything else (zeros, tiny
// values, NaN/Inf, huge norms) goes through the full division path.
__device__ __forceinline__ bool surely_outside(float w0, float w1, float w2, float n32) {
  const int sn = __float_as_int(n32);
  const bool o0 = ((__float_as_int(w0) ^ sn) < 0) && fabsf(w0) >= 0x1p-20f;
  const bool o1 = ((__float_as_int(w1) ^ sn) < 0) && fabsf(w1) >= 0x1p-20f;
  const bool o2 = ((__float_as_int(w2) ^ sn) < 0) && fabsf(w2) >= 0x1p-20f;
  return (o0 || o1 || o2) && fabsf(n32) <= 0x1p99f;
}

template <int BLK_LOG2>
__global__ void __launch_bounds__(kRasterThreads)
raster_forward_kernel(const RasterArgs a, const Plan p) {
  constexpr int BLK = 1 << BLK_LOG2;
  extern __shared__ __align__(16) unsigned char s_raw[];
  RasterSmem<BLK_LOG2>& sm = *reinterpret_cast<RasterSmem<BLK_LOG2>*>(s_raw);
  unsigned long long* zbuf = sm.zbuf;

  DIBR_TRACE_BEGIN();
  // blocks are taken longest face list first (lpt_block)
  int b, bxi, byi;
  if (!lpt_block(a.job, p, BLK_LOG2, blockIdx.x, b, bxi, byi)) return;
  const int bx0 = bxi << BLK_LOG2, by0 = byi << BLK_LOG2;
  const int bx1 = min(bx0 + BLK, p.W) - 1, by1 = min(by0 + BLK, p.H) - 1;
  const Grid g = make_grid(a.job.mult, p.W, p.H);
  const int tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;
  const bool fast_ok = fabs(a.eps_d) <= 0x1p90;

  for (int i = tid; i < BLK * BLK; i += kRasterThreads) zbuf[i] = 0ull;
  if (tid == 0) sm.uncovered = 0;

  // parent bin of this block
  const int shift = p.bin_log2 - BLK_LOG2;
  const int bin = (byi >> shift) * p.nbx + (bxi >> shift);
  const long long row = (long long)b * p.NB + bin;
  long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);  // overflowed call

  unsigned* queue = sm.queue[warp];

  // Exact test + depth + publish of 32 (or `cnt` < 32) queued candidates, one per lane.
  auto drain = [&](int first, int cnt) {
    if (lane < cnt) {
      const unsigned e = queue[(first + lane) & (kQueue - 1)];
      const int s = e & 0xff, rx = (e >> 8) & 0xff, ry = (e >> 16) & 0xff;
      const float x0 = pix_x(g, bx0 + rx), y0 = pix_y(g, by0 + ry);
      float w0, w1, w2;
      if (raster_weights(x0, y0, sm.x[0][s], sm.x[1][s], sm.x[2][s], sm.x[3][s], sm.x[4][s],
                         sm.x[5][s], a.eps_d, w0, w1, w2)) {
        const unsigned long long key =
            depth_key(raster_depth(w0, w1, w2, sm.z[0][s], sm.z[1][s], sm.z[2][s]), sm.face[s]);
        if (key) smem_max64(&zbuf[(ry << BLK_LOG2) + rx], key);
      }
    }
  };

  for (long long base = 0; base < n; base += kRasterThreads) {
    __syncthreads();  // previous chunk fully consumed (and zbuf initialised)
    // ---- phase A: one face per thread -> exact pixel rectangle inside this block
    int cand = 0;
    {
      const long long li = base + tid;
      int rect = 0;
      if (li < n) {
        const int f = a.job.list[start + li];
        Face2D fc;
        if (load_face2d(a.job, b, f, fc)) {
          int x_lo, x_hi, y_lo, y_hi;
          if (pixel_range(g, fc.xmin, fc.ymin, fc.xmax, fc.ymax, x_lo, x_hi, y_lo, y_hi)) {
            x_lo = max(x_lo, bx0); x_hi = min(x_hi, bx1);
            y_lo = max(y_lo, by0); y_hi = min(y_hi, by1);
            // trim the conservative range to the reference's half-open test
            // (rasterization_cuda.cu:115); pix_x / pix_y are monotone in ix / iy
            while (x_lo <= x_hi && pix_x(g, x_lo) < fc.xmin) x_lo++;
            while (x_lo <= x_hi && pix_x(g, x_hi) >= fc.xmax) x_hi--;
            while (y_lo <= y_hi && pix_y(g, y_lo) >= fc.ymax) y_lo++;
            while (y_lo <= y_hi && pix_y(g, y_hi) < fc.ymin) y_hi--;
            const int w = x_hi - x_lo + 1, h = y_hi - y_lo + 1;
            if (w > 0 && h > 0) {
              cand = w * h;
              rect = (x_lo - bx0) | ((y_lo - by0) << 8) | ((w - 1) << 16) | ((h - 1) << 24);
              const long long gi = (long long)b * a.job.view_stride + f;
              sm.x[0][tid] = fc.ax; sm.x[1][tid] = fc.ay; sm.x[2][tid] = fc.bx;
              sm.x[3][tid] = fc.by; sm.x[4][tid] = fc.cx; sm.x[5][tid] = fc.cy;
              sm.z[0][tid] = __ldg(a.z + gi * 3);
              sm.z[1][tid] = __ldg(a.z + gi * 3 + 1);
              sm.z[2][tid] = __ldg(a.z + gi * 3 + 2);
              sm.face[tid] = a.job.first ? f - (int)a.job.first[b] : f;
            }
          }
        }
      }
      sm.rect[tid] = rect;
    }
    // block-wide exclusive scan of the candidate counts
    int incl = cand;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) sm.wsum[warp] = incl;
    __syncthreads();
    int woff = 0, total = 0;
#pragma unroll
    for (int i = 0; i < kRasterWarps; i++) {
      const int v = sm.wsum[i];
      if (i < warp) woff += v;
      total += v;
    }
    sm.pre[tid] = woff + incl - cand;
    if (tid == 0) sm.pre[kRasterThreads] = total;
    __syncthreads();
    if (total == 0) continue;

    // ---- phase B: the chunk's (face, pixel) candidates dealt out evenly, a contiguous
    // run per thread; cheap sign test for all, survivors compacted into the warp's queue
    const int per = (total + kRasterThreads - 1) / kRasterThreads;
    int k = tid * per;
    const int k_end = min(k + per, total);
    int s = 0, rx0 = 0, rw = 1, rx = 0, ry = 0, left = 0;
    float ax = 0.f, ay = 0.f, bx = 0.f, by = 0.f, cx = 0.f, cy = 0.f, y0 = 0.f;
    if (k < k_end) {
#pragma unroll
      for (int st = kRasterThreads / 2; st > 0; st >>= 1)
        if (sm.pre[s + st] <= k) s += st;
      const int r = sm.rect[s];
      rx0 = r & 0xff; rw = ((r >> 16) & 0xff) + 1;
      const int off = k - sm.pre[s];
      ry = (r >> 8 & 0xff) + off / rw;
      rx = rx0 + off % rw;
      left = sm.pre[s + 1] - k;  // candidates left in this face, this one included
      ax = sm.x[0][s]; ay = sm.x[1][s]; bx = sm.x[2][s];
      by = sm.x[3][s]; cx = sm.x[4][s]; cy = sm.x[5][s];
      y0 = pix_y(g, by0 + ry);
    }
    int qhead = 0, qn = 0;  // warp-uniform
    for (int it = 0; it < per; it++) {
      bool keep = false;
      if (k < k_end) {
        const float x0 = pix_x(g, bx0 + rx);
        const float a_ex = __fsub_rn(ax, x0), a_ey = __fsub_rn(ay, y0);
        const float b_ex = __fsub_rn(bx, x0), b_ey = __fsub_rn(by, y0);
        const float c_ex = __fsub_rn(cx, x0), c_ey = __fsub_rn(cy, y0);
        const float w0 = __fmaf_rn(b_ex, c_ey, -__fmul_rn(b_ey, c_ex));
        const float w1 = __fmaf_rn(c_ex, a_ey, -__fmul_rn(c_ey, a_ex));
        const float w2 = __fmaf_rn(a_ex, b_ey, -__fmul_rn(a_ey, b_ex));
        const float n32 = __fadd_rn(__fadd_rn(w0, w1), w2);
        keep = !(fast_ok && surely_outside(w0, w1, w2, n32));
      }
      const unsigned m = __ballot_sync(DIBR_FULL_MASK, keep);
      if (keep)
        queue[(qhead + qn + __popc(m & lanemask_lt())) & (kQueue - 1)] =
            (unsigned)s | ((unsigned)rx << 8) | ((unsigned)ry << 16);
      qn += __popc(m);
      __syncwarp();
      if (qn >= 32) {
        drain(qhead, 32);
        qhead = (qhead + 32) & (kQueue - 1);
        qn -= 32;
        __syncwarp();
      }
      // advance to the next candidate of this thread's run
      if (++k < k_end) {
        if (--left == 0) {
          do { s++; } while (sm.pre[s + 1] == sm.pre[s]);
          const int r = sm.rect[s];
          rx0 = r & 0xff; rw = ((r >> 16) & 0xff) + 1;
          rx = rx0; ry = (r >> 8) & 0xff;
          left = sm.pre[s + 1] - sm.pre[s];
          ax = sm.x[0][s]; ay = sm.x[1][s]; bx = sm.x[2][s];
          by = sm.x[3][s]; cx = sm.x[4][s]; cy = sm.x[5][s];
          y0 = pix_y(g, by0 + ry);
        } else if (++rx == rx0 + rw) {
          rx = rx0;
          ry++;
          y0 = pix_y(g, by0 + ry);
        }
      }
    }
    if (qn > 0) drain(qhead, qn);
  }
  __syncthreads();

  // ---- resolve: one pixel per thread-iteration, rows of the block are contiguous in x
  int uncovered = 0;
  for (int i = tid; i < BLK * BLK; i += kRasterThreads) {
    const int ix = bx0 + (i & (BLK - 1)), iy = by0 + (i >> BLK_LOG2);
    if (ix >= p.W || iy >= p.H) continue;
    const long long pix = ((long long)b * p.H + iy) * p.W + ix;
    const unsigned long long key = zbuf[i];
    if (!key) {
      a.out_idx[pix] = -1;
      for (int d = 0; d < a.D; d++) a.out_feat[pix * a.D + d] = 0.f;
      if (a.out_w) { a.out_w[pix * 3] = 0.f; a.out_w[pix * 3 + 1] = 0.f; a.out_w[pix * 3 + 2] = 0.f; }
      uncovered++;
      continue;
    }
    const int face = key_face(key);
    const int f = a.job.first ? face + (int)a.job.first[b] : face;
    Face2D fc;
    load_face2d(a.job, b, f, fc);
    float w0, w1, w2;
    raster_weights(pix_x(g, ix), pix_y(g, iy), fc.ax, fc.ay, fc.bx, fc.by, fc.cx, fc.cy, a.eps_d,
                   w0, w1, w2);
    a.out_idx[pix] = face;
    if (a.out_w) { a.out_w[pix * 3] = w0; a.out_w[pix * 3 + 1] = w1; a.out_w[pix * 3 + 2] = w2; }
    const float* fp = a.feat + ((long long)b * a.job.view_stride + f) * 3 * a.D;
    for (int d = 0; d < a.D; d++) {
      const float r0 = __ldg(fp + d), r1 = __ldg(fp + a.D + d), r2 = __ldg(fp + 2 * a.D + d);
      a.out_feat[pix * a.D + d] = __fmaf_rn(w2, r2, __fmaf_rn(w0, r0, __fmul_rn(w1, r1)));
    }
  }
  if (a.block_uncovered) {
    uncovered = __reduce_add_sync(DIBR_FULL_MASK, uncovered);
    if (lane == 0 && uncovered) atomicAdd(&sm.uncovered, uncovered);
    __syncthreads();
    if (tid == 0)
      a.block_uncovered[(long long)b * p.nkx * p.nky + byi * p.nkx + bxi] =
          sm.uncovered;
  }
  DIBR_TRACE_END(n);
}

cudaError_t launch_raster_forward(const RasterArgs& a, const Plan& p, cudaStream_t st) {
  const int shift = p.bin_log2 - p.blk_log2;
  dim3 grid((unsigned)(((long long)p.B * p.NB) << (2 * shift)));
  prof_begin(K_RASTER_FWD, st);
  if (p.blk_log2 == 5) {
    raster_forward_kernel<5><<<grid, kRasterThreads, sizeof(RasterSmem<5>), st>>>(a, p);
  } else {
    const size_t smem = sizeof(RasterSmem<6>);  // above the 48 KB default: opt in
    static bool once = false;
    if (!once) {
      cudaError_t e = cudaFuncSetAttribute(raster_forward_kernel<6>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      once = true;
    }
    raster_forward_kernel<6><<<grid, kRasterThreads, smem, st>>>(a, p);
  }
  prof_end(st);
  return cudaGetLastError();
}

// ===========================================================================
// Backward (rasterization_cuda.cu:238-402), pixel-parallel.
//
// The reference issues 9*D global atomics per covered pixel.  Here a warp owns an
// 8x4 pixel region; lanes that selected the same face are grouped with
// __match_any_sync and their contributions are summed with shuffles first, so one
// red.global.add per (face, component) leaves the warp, and the 6 image-coordinate
// terms are summed over the feature dimension in registers before that.
// ===========================================================================

template <int N>
__device__ __forceinline__ void reduce_peers(unsigned peers, float (&v)[N], int lane) {
  int rel_pos = __popc(peers & lanemask_lt());
  unsigned rem = peers & ~(lanemask_lt() | (1u << lane));  // peers with a higher lane id
  while (__any_sync(DIBR_FULL_MASK, rem != 0)) {
    const int next = __ffs(rem);
    const int src = next ? next - 1 : lane;
#pragma unroll
    for (int i = 0; i < N; i++) {
      const float t = __shfl_sync(DIBR_FULL_MASK, v[i], src);
      if (next) v[i] += t;
    }
    const bool done = rel_pos & 1;
    rem &= __ballot_sync(DIBR_FULL_MASK, !done);
    rel_pos >>= 1;
  }
}

template <int DC>  // features handled per pass (D <= 4: one pass)
__global__ void __launch_bounds__(256) raster_backward_kernel(const RasterBwdArgs a) {
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
  const bool leader = (peers & lanemask_lt()) == 0;

  float gimg[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float w[3] = {0.f, 0.f, 0.f};
  float dw1[6], dw2[6];  // dw*: d(w1|w2 * k3^2)/d(ax,ay,bx,by,cx,cy)
  float inv_k3sq = 0.f;
#pragma unroll
  for (int i = 0; i < 6; i++) dw1[i] = dw2[i] = 0.f;
  long long fbase = 0;
  if (face >= 0) {
    fbase = (long long)b * a.F + face;
    const float2* ip = reinterpret_cast<const float2*>(a.img + fbase * 6);
    const float2 A = __ldg(ip), Bv = __ldg(ip + 1), C = __ldg(ip + 2);
    const float ax = A.x, ay = A.y, bx = Bv.x, by = Bv.y, cx = C.x, cy = C.y;
    if (a.weights) {
      w[0] = a.weights[pix * 3]; w[1] = a.weights[pix * 3 + 1]; w[2] = a.weights[pix * 3 + 2];
    } else {
      const Grid g = make_grid(a.mult, a.W, a.H);
      raster_weights(pix_x(g, ix), pix_y(g, iy), __fmul_rn(ax, a.mult), __fmul_rn(ay, a.mult),
                     __fmul_rn(bx, a.mult), __fmul_rn(by, a.mult), __fmul_rn(cx, a.mult),
                     __fmul_rn(cy, a.mult), a.eps_d, w[0], w[1], w[2]);
    }
    // rasterization_cuda.cu:302-374
    const float x0 = w[0] * ax + w[1] * bx + w[2] * cx;
    const float y0 = w[0] * ay + w[1] * by + w[2] * cy;
    const float m = bx - ax, pp = by - ay, n = cx - ax, q = cy - ay, s = x0 - ax, t = y0 - ay;
    const float k1 = s * q - n * t;
    const float k2 = m * t - s * pp;
    float k3 = m * q - n * pp;
    k3 = (float)((double)k3 + copysign(a.eps_d, (double)k3));
    const float dw1dm = 0.f * k3 - q * k1, dw1dn = (-t) * k3 - (-pp) * k1;
    const float dw1dp = 0.f * k3 - (-n) * k1, dw1dq = s * k3 - m * k1;
    const float dw1ds = q * k3 - 0.f * k1, dw1dt = (-n) * k3 - 0.f * k1;
    const float dw2dm = t * k3 - q * k2, dw2dn = 0.f * k3 - (-pp) * k2;
    const float dw2dp = (-s) * k3 - (-n) * k2, dw2dq = 0.f * k3 - m * k2;
    const float dw2ds = (-pp) * k3 - 0.f * k2, dw2dt = m * k3 - 0.f * k2;
    dw1[0] = -(dw1dm + dw1dn + dw1ds); dw1[1] = -(dw1dp + dw1dq + dw1dt);
    dw1[2] = dw1dm; dw1[3] = dw1dp; dw1[4] = dw1dn; dw1[5] = dw1dq;
    dw2[0] = -(dw2dm + dw2dn + dw2ds); dw2[1] = -(dw2dp + dw2dq + dw2dt);
    dw2[2] = dw2dm; dw2[3] = dw2dp; dw2[4] = dw2dn; dw2[5] = dw2dq;
    inv_k3sq = 1.0f / (k3 * k3);
  }

  for (int d0 = 0; d0 < a.D; d0 += DC) {
    float gf[3 * DC];
#pragma unroll
    for (int i = 0; i < 3 * DC; i++) gf[i] = 0.f;
    if (face >= 0) {
      const float* fp = a.feat + fbase * 3 * a.D;
#pragma unroll
      for (int dd = 0; dd < DC; dd++) {
        const int d = d0 + dd;
        if (d < a.D) {
          const float gd = a.grad[pix * a.D + d];
          gf[dd] = gd * w[0];            // :273-285
          gf[DC + dd] = gd * w[1];
          gf[2 * DC + dd] = gd * w[2];
          const float c0 = __ldg(fp + d), c1 = __ldg(fp + a.D + d), c2 = __ldg(fp + 2 * a.D + d);
          const float dldI = gd * inv_k3sq;  // :389 (gd / k3^2; one reciprocal per pixel)
          const float e1 = c1 - c0, e2 = c2 - c0;
#pragma unroll
          for (int i = 0; i < 6; i++) gimg[i] += dldI * (e1 * dw1[i] + e2 * dw2[i]);
        }
      }
    }
    reduce_peers<3 * DC>(peers, gf, lane);
    if (leader && face >= 0) {
      float* gp = a.grad_feat + fbase * 3 * a.D;
#pragma unroll
      for (int dd = 0; dd < DC; dd++) {
        const int d = d0 + dd;
        if (d < a.D) {
          atomicAdd(gp + d, gf[dd]);
          atomicAdd(gp + a.D + d, gf[DC + dd]);
          atomicAdd(gp + 2 * a.D + d, gf[2 * DC + dd]);
        }
      }
    }
  }
  reduce_peers<6>(peers, gimg, lane);
  if (leader && face >= 0) {
    // (x, y) of a vertex are adjacent and 8-byte aligned: three vector reductions
    float2* gp = reinterpret_cast<float2*>(a.grad_img + fbase * 6);
#pragma unroll
    for (int i = 0; i < 3; i++) atomicAdd(gp + i, make_float2(gimg[2 * i], gimg[2 * i + 1]));
  }
}

cudaError_t launch_raster_backward(const RasterBwdArgs& a, cudaStream_t st) {
  dim3 grid((a.W + 15) / 16, (a.H + 15) / 16, a.B);
  prof_begin(K_RASTER_BWD, st);
  switch (a.D >= 4 ? 4 : a.D) {
    case 1: raster_backward_kernel<1><<<grid, 256, 0, st>>>(a); break;
    case 2: raster_backward_kernel<2><<<grid, 256, 0, st>>>(a); break;
    case 3: raster_backward_kernel<3><<<grid, 256, 0, st>>>(a); break;
    default: raster_backward_kernel<4><<<grid, 256, 0, st>>>(a); break;
  }
  prof_end(st);
  return cudaGetLastError();
}

}  // namespace dibr
