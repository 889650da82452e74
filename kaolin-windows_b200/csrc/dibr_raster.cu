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
constexpr int kSmallBox = 24;  // boxes with more candidate pixels go to the warp queue

struct BigFace {
  float ax, ay, bx, by, cx, cy, xmin, ymin, xmax, ymax, az, bz, cz;
  int face, x_lo, y_lo, w, n;
};

__device__ __forceinline__ void raster_pixel(const Grid& g, int ix, int iy, float ax, float ay,
                                             float bx, float by, float cx, float cy, float xmin,
                                             float ymin, float xmax, float ymax, float az,
                                             float bz, float cz, double eps_d, int face,
                                             unsigned long long* zbuf, int bx0, int by0,
                                             int blk_log2) {
  const float x0 = pix_x(g, ix), y0 = pix_y(g, iy);
  if (x0 < xmin || x0 >= xmax || y0 < ymin || y0 >= ymax) return;  // :115
  float w0, w1, w2;
  if (!raster_weights(x0, y0, ax, ay, bx, by, cx, cy, eps_d, w0, w1, w2)) return;
  const unsigned long long key = depth_key(raster_depth(w0, w1, w2, az, bz, cz), face);
  if (key) smem_max64(&zbuf[((iy - by0) << blk_log2) + (ix - bx0)], key);
}

template <int BLK_LOG2>
__global__ void __launch_bounds__(kRasterThreads)
raster_forward_kernel(const RasterArgs a, const Plan p) {
  constexpr int BLK = 1 << BLK_LOG2;
  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned long long* zbuf = reinterpret_cast<unsigned long long*>(s_raw);        // BLK*BLK
  BigFace* s_big = reinterpret_cast<BigFace*>(s_raw + sizeof(unsigned long long) * BLK * BLK);
  __shared__ int s_nbig;
  __shared__ int s_uncovered;

  const int b = blockIdx.z;
  const int bx0 = blockIdx.x << BLK_LOG2, by0 = blockIdx.y << BLK_LOG2;
  const int bx1 = min(bx0 + BLK, p.W) - 1, by1 = min(by0 + BLK, p.H) - 1;
  const Grid g = make_grid(a.job.mult, p.W, p.H);
  const int tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;

  for (int i = tid; i < BLK * BLK; i += kRasterThreads) zbuf[i] = 0ull;
  if (tid == 0) { s_nbig = 0; s_uncovered = 0; }

  // parent bin of this block
  const int shift = p.bin_log2 - BLK_LOG2;
  const int bin = (blockIdx.y >> shift) * p.nbx + (blockIdx.x >> shift);
  const long long row = (long long)b * p.NB + bin;
  long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);  // overflowed call
  __syncthreads();

  for (long long base = 0; base < n; base += kRasterThreads) {
    const long long li = base + tid;
    if (li < n) {
      const int f = a.job.list[start + li];
      Face2D fc;
      // list entries were accepted by load_face2d during binning; redo it for the data
      if (load_face2d(a.job, b, f, fc)) {
        int x_lo, x_hi, y_lo, y_hi;
        if (pixel_range(g, fc.xmin, fc.ymin, fc.xmax, fc.ymax, x_lo, x_hi, y_lo, y_hi)) {
          x_lo = max(x_lo, bx0); x_hi = min(x_hi, bx1);
          y_lo = max(y_lo, by0); y_hi = min(y_hi, by1);
          const int w = x_hi - x_lo + 1, h = y_hi - y_lo + 1;
          if (w > 0 && h > 0) {
            const long long gi = (long long)b * a.job.view_stride + f;
            const float az = __ldg(a.z + gi * 3), bz = __ldg(a.z + gi * 3 + 1),
                        cz = __ldg(a.z + gi * 3 + 2);
            const int face = a.job.first ? f - (int)a.job.first[b] : f;
            if (w * h <= kSmallBox) {
              for (int iy = y_lo; iy <= y_hi; iy++)
                for (int ix = x_lo; ix <= x_hi; ix++)
                  raster_pixel(g, ix, iy, fc.ax, fc.ay, fc.bx, fc.by, fc.cx, fc.cy, fc.xmin,
                               fc.ymin, fc.xmax, fc.ymax, az, bz, cz, a.eps_d, face, zbuf, bx0,
                               by0, BLK_LOG2);
            } else {
              const int s = atomicAdd(&s_nbig, 1);
              BigFace& q = s_big[s];
              q.ax = fc.ax; q.ay = fc.ay; q.bx = fc.bx; q.by = fc.by; q.cx = fc.cx; q.cy = fc.cy;
              q.xmin = fc.xmin; q.ymin = fc.ymin; q.xmax = fc.xmax; q.ymax = fc.ymax;
              q.az = az; q.bz = bz; q.cz = cz;
              q.face = face; q.x_lo = x_lo; q.y_lo = y_lo; q.w = w; q.n = w * h;
            }
          }
        }
      }
    }
    __syncthreads();
    const int nbig = s_nbig;
    for (int qi = warp; qi < nbig; qi += kRasterThreads / 32) {
      const BigFace& q = s_big[qi];
      for (int k = lane; k < q.n; k += 32) {
        const int iy = q.y_lo + k / q.w, ix = q.x_lo + k % q.w;
        raster_pixel(g, ix, iy, q.ax, q.ay, q.bx, q.by, q.cx, q.cy, q.xmin, q.ymin, q.xmax,
                     q.ymax, q.az, q.bz, q.cz, a.eps_d, q.face, zbuf, bx0, by0, BLK_LOG2);
      }
    }
    __syncthreads();
    if (tid == 0) s_nbig = 0;
    // (the next iteration's atomicAdd on s_nbig happens after its own loads; the barrier
    //  at the end of that iteration orders it against this reset through the next sync)
    __syncthreads();
  }

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
    if (lane == 0 && uncovered) atomicAdd(&s_uncovered, uncovered);
    __syncthreads();
    if (tid == 0)
      a.block_uncovered[(long long)b * p.nkx * p.nky + blockIdx.y * p.nkx + blockIdx.x] =
          s_uncovered;
  }
}

cudaError_t launch_raster_forward(const RasterArgs& a, const Plan& p, cudaStream_t st) {
  dim3 grid(p.nkx, p.nky, p.B);
  const size_t big = sizeof(BigFace) * kRasterThreads;
  prof_begin(K_RASTER_FWD, st);
  if (p.blk_log2 == 5) {
    raster_forward_kernel<5><<<grid, kRasterThreads, 8 * 32 * 32 + big, st>>>(a, p);
  } else {
    const size_t smem = 8 * 64 * 64 + big;  // 50 KB: above the 48 KB default, opt in
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
  float dw1[6], dw2[6], inv_k3sq_num = 0.f;  // dw*: d(w1|w2 * k3^2)/d(ax,ay,bx,by,cx,cy)
  float k3sq = 1.f;
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
    k3sq = k3 * k3;
  }
  (void)inv_k3sq_num;

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
          const float dldI = gd / k3sq;  // :389
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
    float* gp = a.grad_img + fbase * 6;
#pragma unroll
    for (int i = 0; i < 6; i++) atomicAdd(gp + i, gimg[i]);
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
