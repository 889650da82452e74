// dibr_f64.cu -- the two FORWARD kernels of the reference's double dispatch
// (AT_DISPATCH_FLOATING_TYPES at rasterization_cuda.cu:30-39, dibr_soft_mask_cuda.cu:211;
// the reference's tests run both dtypes).  The two backward kernels for double are the
// double instantiations of the float path's designs: raster_backward_f64_kernel
// (dibr_raster.cu: warp-grouped sums, group-level conversion) and
// softmask_dense_backward_kernel<double> (dibr_softmask.cu: pairs dealt over the lanes,
// (face, vertex)-grouped atomics).  Test-sized workloads: one pixel per thread, but over
// the same binned face lists as the float path instead of the reference's loop over every
// face -- the lists come from the float binning run on float
// copies of the coordinates (a conservative superset: its pixel ranges carry 1e-3 pixel of
// slack, the float rounding of a coordinate is < 1e-4 pixel), every listed face then takes
// the exact double test below.
//
// Arithmetic = the reference's expressions with scalar_t = double, written with the FMA
// contraction nvcc emits for them (same shapes as the float path, see dibr_common.cuh).
#include "dibr_common.cuh"

namespace dibr {
namespace {

constexpr int kTile = 16;       // pixels per CTA side; divides every bin size (>= 32)
constexpr int kStage = 64;      // faces staged per round

__device__ __forceinline__ int tile_bin(const Plan& p, int x0, int y0) {
  return (y0 >> p.bin_log2) * p.nbx + (x0 >> p.bin_log2);
}

// ------------------------------------------------------------------ rasterize forward
// rasterization_cuda.cu:43-192 (packed layout: faces of mesh b are [first[b], first[b+1])).
__global__ void __launch_bounds__(kTile * kTile)
raster_forward_f64_kernel(const RasterF64Args a, const Plan p) {
  __shared__ double sx[6][kStage], sz[3][kStage], sb[4][kStage];
  __shared__ int sf[kStage];
  const int b = blockIdx.z, tid = threadIdx.x;
  const int tx0 = blockIdx.x * kTile, ty0 = blockIdx.y * kTile;
  const int ix = tx0 + (tid & (kTile - 1)), iy = ty0 + (tid / kTile);
  const bool inside = ix < p.W && iy < p.H;
  const Grid g = plan_grid(p, a.job.mult);
  // :85-86 float / int division and float * int product, widened afterwards
  const double x0 = (double)pix_x(g, min(ix, p.W - 1)), y0 = (double)pix_y(g, min(iy, p.H - 1));
  const long long row = (long long)b * p.NB + tile_bin(p, tx0, ty0);
  const long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);
  const int first = (int)a.job.first[b];

  double max_z0 = -INFINITY, mw0 = 0., mw1 = 0., mw2 = 0.;
  int max_face = -1;
  for (long long base = 0; base < n; base += kStage) {
    const int cnt = (int)min((long long)kStage, n - base);
    __syncthreads();
    if (tid < cnt) {
      const int f = a.job.list[start + base + tid];
      sf[tid] = f;
#pragma unroll
      for (int k = 0; k < 6; k++) sx[k][tid] = a.img[(long long)f * 6 + k];
#pragma unroll
      for (int k = 0; k < 3; k++) sz[k][tid] = a.z[(long long)f * 3 + k];
#pragma unroll
      for (int k = 0; k < 4; k++) sb[k][tid] = a.bbox[(long long)f * 4 + k];
    }
    __syncthreads();
    if (!inside) continue;
    for (int j = 0; j < cnt; j++) {
      // :115 half-open box on the pixel centre
      if (x0 < sb[0][j] || x0 >= sb[2][j] || y0 < sb[1][j] || y0 >= sb[3][j]) continue;
      const double a_ex = __dsub_rn(sx[0][j], x0), a_ey = __dsub_rn(sx[1][j], y0);
      const double b_ex = __dsub_rn(sx[2][j], x0), b_ey = __dsub_rn(sx[3][j], y0);
      const double c_ex = __dsub_rn(sx[4][j], x0), c_ey = __dsub_rn(sx[5][j], y0);
      double w0 = __fma_rn(b_ex, c_ey, -__dmul_rn(b_ey, c_ex));   // :137-139
      double w1 = __fma_rn(c_ex, a_ey, -__dmul_rn(c_ey, a_ex));
      double w2 = __fma_rn(a_ex, b_ey, -__dmul_rn(a_ey, b_ex));
      double norm = __dadd_rn(__dadd_rn(w0, w1), w2);
      norm = __dadd_rn(norm, copysign(a.eps_d, norm));             // :141-142
      w0 = __ddiv_rn(w0, norm);
      w1 = __ddiv_rn(w1, norm);
      w2 = __ddiv_rn(w2, norm);
      if (w0 < 0. || w1 < 0. || w2 < 0.) continue;                 // :148
      const double z0 = __fma_rn(w2, sz[2][j], __fma_rn(w0, sz[0][j], __dmul_rn(w1, sz[1][j])));
      // :162 ascending scan with a strict '>' == the largest depth, lowest index on ties
      // (the list is not ordered here)
      const int f = sf[j];
      if (z0 > max_z0 || (z0 == max_z0 && max_face >= 0 && f < max_face)) {
        max_z0 = z0; max_face = f; mw0 = w0; mw1 = w1; mw2 = w2;
      }
    }
  }
  if (!inside) return;
  const long long pix = ((long long)b * p.H + iy) * p.W + ix;
  a.out_idx[pix] = max_face >= 0 ? max_face - first : -1;
  if (a.out_w) {
    a.out_w[pix * 3] = mw0; a.out_w[pix * 3 + 1] = mw1; a.out_w[pix * 3 + 2] = mw2;
  }
  const double* fp = a.feat + (long long)max(max_face, 0) * 3 * a.D;
  for (int d = 0; d < a.D; d++) {
    double v = 0.;
    if (max_face >= 0)   // :183-188
      v = __fma_rn(mw2, fp[2 * a.D + d], __fma_rn(mw0, fp[d], __dmul_rn(mw1, fp[a.D + d])));
    a.out_feat[pix * a.D + d] = v;
  }
}

// ------------------------------------------------------------------ soft mask forward
// dibr_soft_mask_cuda.cu:27-184; scaled coordinates, enlarged boxes, dense per-pixel state
// (pre-filled by the caller as dibr_soft_mask.cpp:84-97 does).
#define DIBR_F64_EPS 1e-7

__global__ void __launch_bounds__(kTile * kTile)
softmask_forward_f64_kernel(const SoftF64Args a, const Plan p) {
  __shared__ double sx[6][kStage], sb[4][kStage];
  __shared__ int sf[kStage];
  const int b = blockIdx.z, tid = threadIdx.x;
  const int tx0 = blockIdx.x * kTile, ty0 = blockIdx.y * kTile;
  const int ix = tx0 + (tid & (kTile - 1)), iy = ty0 + (tid / kTile);
  const bool inside = ix < p.W && iy < p.H;
  const Grid g = plan_grid(p, a.job.mult);
  const double x0 = (double)pix_x(g, min(ix, p.W - 1)), y0 = (double)pix_y(g, min(iy, p.H - 1));
  const long long pix = ((long long)b * p.H + min(iy, p.H - 1)) * p.W + min(ix, p.W - 1);
  bool open = false;
  if (inside) {
    open = a.sel[pix] < 0;
    if (!open) a.mask[pix] = 1.0;     // :68-70
  }
  const float mult = a.job.mult;
  const double four_mm = (double)__fmul_rn(__fmul_rn(4.f, mult), mult);   // float product
  const long long row = (long long)b * p.NB + tile_bin(p, tx0, ty0);
  const long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);
  const long long vbase = (long long)b * a.job.view_stride;
  int kid = 0;
  if (a.K <= 0) open = false;
  for (long long base = 0; base < n; base += kStage) {
    if (!__syncthreads_or(open)) break;
    const int cnt = (int)min((long long)kStage, n - base);
    if (tid < cnt) {
      const int f = a.job.list[start + base + tid];
      sf[tid] = f;
#pragma unroll
      for (int k = 0; k < 6; k++) sx[k][tid] = a.img[(vbase + f) * 6 + k];
#pragma unroll
      for (int k = 0; k < 4; k++) sb[k][tid] = a.bbox[(vbase + f) * 4 + k];
    }
    __syncthreads();
    if (!open) continue;
    for (int j = 0; j < cnt; j++) {     // ascending face order (the soft list is ordered)
      if (x0 < sb[0][j] || x0 >= sb[2][j] || y0 < sb[1][j] || y0 >= sb[3][j]) continue;  // :95-97
      double pdis[6];
#pragma unroll
      for (int i = 0; i < 3; i++) {     // :103-141
        const double x1 = sx[2 * i][j], y1 = sx[2 * i + 1][j];
        const double x2 = sx[2 * ((i + 1) % 3)][j], y2 = sx[2 * ((i + 1) % 3) + 1][j];
        const double A = __dsub_rn(y2, y1), B = __dsub_rn(x1, x2);
        const double C = __fma_rn(x2, y1, -__dmul_rn(x1, y2));
        const double AA = __dmul_rn(A, A), BB = __dmul_rn(B, B), AB = __dmul_rn(A, B);
        const double up = __dadd_rn(__fma_rn(B, y0, __dmul_rn(A, x0)), C);
        const double dd = __dadd_rn(__dadd_rn(AA, BB), DIBR_F64_EPS);
        const double x3 = __ddiv_rn(__fma_rn(-A, C, __fma_rn(BB, x0, -__dmul_rn(AB, y0))), dd);
        const double y3 = __ddiv_rn(__fma_rn(-B, C, __fma_rn(AA, y0, -__dmul_rn(AB, x0))), dd);
        const double direct = __fma_rn(__dsub_rn(x3, x1), __dsub_rn(x3, x2),
                                       __dmul_rn(__dsub_rn(y3, y1), __dsub_rn(y3, y2)));
        pdis[i] = direct > 0. ? four_mm : __ddiv_rn(__dmul_rn(up, up), dd);
        const double dx = __dsub_rn(x0, x1), dy = __dsub_rn(y0, y1);     // :144-149
        pdis[i + 3] = __fma_rn(dx, dx, __dmul_rn(dy, dy));
      }
      int edgeid = 0;
      double dis = pdis[0];
#pragma unroll
      for (int i = 1; i < 6; i++)
        if (dis > pdis[i]) { dis = pdis[i]; edgeid = i; }              // :154-159
      // :161-163  sigmainv * dissquare / multiplier / multiplier (float scalars widen)
      const double z = __ddiv_rn(__ddiv_rn(__dmul_rn((double)a.sigmainv, dis), (double)mult),
                                 (double)mult);
      const long long o = pix * a.K + kid;
      a.prob[o] = exp(-z);
      a.cidx[o] = sf[j];
      a.type[o] = (uint8_t)(edgeid + 1);
      if (++kid >= a.K) { open = false; break; }                        // :165-171
    }
  }
  if (inside && a.sel[pix] < 0) {       // :174-182
    double allprob = 1.0;
    for (int i = 0; i < kid; i++) allprob = __dmul_rn(allprob, __dsub_rn(1.0, a.prob[pix * a.K + i]));
    a.mask[pix] = __dsub_rn(1.0, allprob);
  }
}

}  // namespace

cudaError_t launch_raster_forward_f64(const RasterF64Args& a, const Plan& p, cudaStream_t st) {
  dim3 grid((p.W + kTile - 1) / kTile, (p.H + kTile - 1) / kTile, p.B);
  raster_forward_f64_kernel<<<grid, kTile * kTile, 0, st>>>(a, p);
  return cudaGetLastError();
}

cudaError_t launch_softmask_forward_f64(const SoftF64Args& a, const Plan& p, cudaStream_t st) {
  dim3 grid((p.W + kTile - 1) / kTile, (p.H + kTile - 1) / kTile, p.B);
  softmask_forward_f64_kernel<<<grid, kTile * kTile, 0, st>>>(a, p);
  return cudaGetLastError();
}

}  // namespace dibr
