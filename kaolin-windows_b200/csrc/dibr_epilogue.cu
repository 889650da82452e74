// dibr_epilogue.cu -- the consumer step of the DIB-R loop:
//   texture_mapping   kaolin/render/mesh/utils.py:23-76  (clamp to [0,1], to [-1,1], flip y,
//                     torch.nn.functional.grid_sample(align_corners=False, padding 'border'),
//                     NCHW -> channels-last) as ONE kernel each way instead of six torch
//                     kernels and a permuted copy;
//   mask_iou          kaolin/metrics/render.py:18-40: both per-view sums in one pass.
// The sampling arithmetic follows ATen's grid sampler (unnormalize ((c + 1) * size - 1) / 2,
// clip to [0, size - 1], nearbyint for 'nearest', the four corner weights
// (x_se - x)(y_se - y) ... accumulated nw, ne, sw, se for 'bilinear').
#include "../../include/dibr_b200.h"
#include "dibr_common.cuh"

namespace dibr {

int fail(int code, const char* fmt, ...);

namespace {

struct TexArgs {
  const float* coords;   // (B, N, 2)  in [0, 1], y up
  const float* tex;      // (B, C, Ht, Wt)
  float* out;            // (B, N, C)
  const float* grad_out; // (B, N, C)
  float* grad_tex;       // (B, C, Ht, Wt) accumulated
  float* grad_coords;    // (B, N, 2) written (bilinear) or null
  long long N;
  int B, C, Ht, Wt, bilinear;
};

// source index of one coordinate and d(source)/d(coordinate) through the reference's chain
// clamp(0,1) -> *2-1 -> (y: negate) -> unnormalize -> clip to the border
__device__ __forceinline__ float source_index(float c, int size, bool flip, float& dsrc) {
  const float cl = fminf(fmaxf(c, 0.f), 1.f);
  float d = (c >= 0.f && c <= 1.f) ? 2.f : 0.f;        // torch.clamp passes grad on [min, max]
  float g = cl * 2.f - 1.f;
  if (flip) { g = -g; d = -d; }
  float x = ((g + 1.f) * (float)size - 1.f) / 2.f;
  d *= (float)size / 2.f;
  // clip_coordinates_set_grad: zero outside (and on) the border
  if (x <= 0.f) { x = 0.f; d = 0.f; }
  else if (x >= (float)(size - 1)) { x = (float)(size - 1); d = 0.f; }
  dsrc = d;
  return x;
}

__global__ void __launch_bounds__(256) texture_forward_kernel(const TexArgs a) {
  const long long total = (long long)a.B * a.N;
  const long long plane = (long long)a.Ht * a.Wt;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / a.N);
    const float2 uv = __ldg(reinterpret_cast<const float2*>(a.coords) + i);
    float dx, dy;
    const float x = source_index(uv.x, a.Wt, false, dx);
    const float y = source_index(uv.y, a.Ht, true, dy);
    const float* t = a.tex + (long long)b * a.C * plane;
    float* o = a.out + i * a.C;
    if (!a.bilinear) {
      const int xi = (int)nearbyintf(x), yi = (int)nearbyintf(y);
      const bool ok = xi >= 0 && xi < a.Wt && yi >= 0 && yi < a.Ht;
      for (int c = 0; c < a.C; c++) o[c] = ok ? __ldg(t + c * plane + (long long)yi * a.Wt + xi) : 0.f;
      continue;
    }
    const float xf = floorf(x), yf = floorf(y);
    const int x0 = (int)xf, y0 = (int)yf, x1 = x0 + 1, y1 = y0 + 1;
    const float nw = ((float)x1 - x) * ((float)y1 - y), ne = (x - xf) * ((float)y1 - y);
    const float sw = ((float)x1 - x) * (y - yf), se = (x - xf) * (y - yf);
    const bool vx0 = x0 >= 0 && x0 < a.Wt, vx1 = x1 >= 0 && x1 < a.Wt;
    const bool vy0 = y0 >= 0 && y0 < a.Ht, vy1 = y1 >= 0 && y1 < a.Ht;
    for (int c = 0; c < a.C; c++) {
      const float* pc = t + c * plane;
      float v = 0.f;
      if (vx0 && vy0) v += __ldg(pc + (long long)y0 * a.Wt + x0) * nw;
      if (vx1 && vy0) v += __ldg(pc + (long long)y0 * a.Wt + x1) * ne;
      if (vx0 && vy1) v += __ldg(pc + (long long)y1 * a.Wt + x0) * sw;
      if (vx1 && vy1) v += __ldg(pc + (long long)y1 * a.Wt + x1) * se;
      o[c] = v;
    }
  }
}

__global__ void __launch_bounds__(256) texture_backward_kernel(const TexArgs a) {
  const long long total = (long long)a.B * a.N;
  const long long plane = (long long)a.Ht * a.Wt;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / a.N);
    const float2 uv = __ldg(reinterpret_cast<const float2*>(a.coords) + i);
    float dx, dy;
    const float x = source_index(uv.x, a.Wt, false, dx);
    const float y = source_index(uv.y, a.Ht, true, dy);
    const float* t = a.tex + (long long)b * a.C * plane;
    float* gt = a.grad_tex ? a.grad_tex + (long long)b * a.C * plane : nullptr;
    const float* go = a.grad_out + i * a.C;
    if (!a.bilinear) {
      const int xi = (int)nearbyintf(x), yi = (int)nearbyintf(y);
      if (gt && xi >= 0 && xi < a.Wt && yi >= 0 && yi < a.Ht)
        for (int c = 0; c < a.C; c++) {
          const float g = __ldg(go + c);
          if (g != 0.f) atomicAdd(gt + c * plane + (long long)yi * a.Wt + xi, g);
        }
      if (a.grad_coords) reinterpret_cast<float2*>(a.grad_coords)[i] = make_float2(0.f, 0.f);
      continue;
    }
    const float xf = floorf(x), yf = floorf(y);
    const int x0 = (int)xf, y0 = (int)yf, x1 = x0 + 1, y1 = y0 + 1;
    const float wx1 = x - xf, wx0 = (float)x1 - x, wy1 = y - yf, wy0 = (float)y1 - y;
    const bool vx0 = x0 >= 0 && x0 < a.Wt, vx1 = x1 >= 0 && x1 < a.Wt;
    const bool vy0 = y0 >= 0 && y0 < a.Ht, vy1 = y1 >= 0 && y1 < a.Ht;
    float gx = 0.f, gy = 0.f;
    for (int c = 0; c < a.C; c++) {
      const float g = __ldg(go + c);
      const long long pc = c * plane;
      const long long o00 = pc + (long long)y0 * a.Wt + x0, o01 = o00 + 1;
      const long long o10 = pc + (long long)y1 * a.Wt + x0, o11 = o10 + 1;
      if (gt && g != 0.f) {
        if (vx0 && vy0) atomicAdd(gt + o00, g * wx0 * wy0);
        if (vx1 && vy0) atomicAdd(gt + o01, g * wx1 * wy0);
        if (vx0 && vy1) atomicAdd(gt + o10, g * wx0 * wy1);
        if (vx1 && vy1) atomicAdd(gt + o11, g * wx1 * wy1);
      }
      if (a.grad_coords) {   // d out / d x = -nw' ... as ATen's grid_sampler_2d_backward
        const float v00 = vx0 && vy0 ? __ldg(t + o00) : 0.f, v01 = vx1 && vy0 ? __ldg(t + o01) : 0.f;
        const float v10 = vx0 && vy1 ? __ldg(t + o10) : 0.f, v11 = vx1 && vy1 ? __ldg(t + o11) : 0.f;
        gx += g * ((v01 - v00) * wy0 + (v11 - v10) * wy1);
        gy += g * ((v10 - v00) * wx0 + (v11 - v01) * wx1);
      }
    }
    if (a.grad_coords) reinterpret_cast<float2*>(a.grad_coords)[i] = make_float2(gx * dx, gy * dy);
  }
}

}  // namespace
}  // namespace dibr

namespace dibr {
namespace {

// sums[b] = {sum lhs*rhs, sum (lhs + rhs - lhs*rhs)} over the pixels of view b
// (render.py:33-36); one CTA-reduction and two atomics per CTA.
__global__ void __launch_bounds__(256)
mask_iou_sums_kernel(const float* __restrict__ lhs, const float* __restrict__ rhs,
                     float* __restrict__ sums, long long npix) {
  const int b = blockIdx.y;
  const float* l = lhs + (long long)b * npix;
  const float* r = rhs + (long long)b * npix;
  float up = 0.f, down = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix;
       i += (long long)gridDim.x * blockDim.x) {
    const float a = __ldg(l + i), c = __ldg(r + i), m = a * c;
    up += m;
    down += (a + c) - m;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    up += __shfl_xor_sync(DIBR_FULL_MASK, up, d);
    down += __shfl_xor_sync(DIBR_FULL_MASK, down, d);
  }
  __shared__ float s_up[8], s_down[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_up[warp] = up; s_down[warp] = down; }
  __syncthreads();
  if (warp == 0) {
    up = lane < 8 ? s_up[lane] : 0.f;
    down = lane < 8 ? s_down[lane] : 0.f;
#pragma unroll
    for (int d = 4; d > 0; d >>= 1) {
      up += __shfl_xor_sync(DIBR_FULL_MASK, up, d);
      down += __shfl_xor_sync(DIBR_FULL_MASK, down, d);
    }
    if (lane == 0) { atomicAdd(sums + 2 * b, up); atomicAdd(sums + 2 * b + 1, down); }
  }
}

// loss = 1 - mean_b up_b / (down_b + 1e-10):
//   d loss / d lhs = -(g / B) * (rhs (down + eps) - up (1 - rhs)) / (down + eps)^2, same for rhs
__global__ void __launch_bounds__(256)
mask_iou_backward_kernel(const float* __restrict__ lhs, const float* __restrict__ rhs,
                         const float* __restrict__ sums, const float* __restrict__ grad_loss,
                         float* __restrict__ grad_lhs, float* __restrict__ grad_rhs,
                         long long npix, int B) {
  const int b = blockIdx.y;
  const float up = sums[2 * b], den = sums[2 * b + 1] + 1e-10f;
  const float k = -__ldg(grad_loss) / (float)B / (den * den);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix;
       i += (long long)gridDim.x * blockDim.x) {
    const long long o = (long long)b * npix + i;
    const float a = __ldg(lhs + o), c = __ldg(rhs + o);
    if (grad_lhs) grad_lhs[o] = k * (c * den - up * (1.f - c));
    if (grad_rhs) grad_rhs[o] = k * (a * den - up * (1.f - a));
  }
}

// the B-sized tail of mask_iou (render.py:37-40) on the device: 1 - mean_b up_b / (down_b + 1e-10)
__global__ void mask_iou_loss_kernel(const float* __restrict__ sums, float* __restrict__ loss, int B) {
  float acc = 0.f;
  for (int b = threadIdx.x; b < B; b += 32) acc += sums[2 * b] / (sums[2 * b + 1] + 1e-10f);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(DIBR_FULL_MASK, acc, d);
  if (threadIdx.x == 0) loss[0] = 1.0f - acc / (float)B;
}

inline int blocks_for(long long n, int cap) {
  long long g = (n + 255) / 256;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace dibr

using namespace dibr;

#define EPI_CUDA(expr)                                                               \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess)                                                           \
      return fail(DIBR_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));      \
  } while (0)

static int tex_check(int B, long long N, int C, int Ht, int Wt, int mode) {
  if (B < 0 || N < 0 || C < 0 || Ht <= 0 || Wt <= 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (mode != 0 && mode != 1) return fail(DIBR_E_INVALID, "mode must be 0 (nearest) or 1 (bilinear)");
  return DIBR_OK;
}

extern "C" {

DIBR_API int dibr_texture_mapping_fwd(int batch, int64_t num_points, int channels, int tex_height,
                                      int tex_width, int mode, const float* texture_coordinates,
                                      const float* texture_maps, float* out, void* stream) {
  int rc = tex_check(batch, num_points, channels, tex_height, tex_width, mode);
  if (rc) return rc;
  if (batch == 0 || num_points == 0 || channels == 0) return DIBR_OK;
  if (!texture_coordinates || !texture_maps || !out) return fail(DIBR_E_INVALID, "null pointer");
  TexArgs a = {};
  a.coords = texture_coordinates; a.tex = texture_maps; a.out = out;
  a.N = num_points; a.B = batch; a.C = channels; a.Ht = tex_height; a.Wt = tex_width;
  a.bilinear = mode;
  texture_forward_kernel<<<blocks_for((long long)batch * num_points, 148 * 32), 256, 0,
                           (cudaStream_t)stream>>>(a);
  EPI_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_texture_mapping_bwd(int batch, int64_t num_points, int channels, int tex_height,
                                      int tex_width, int mode, const float* texture_coordinates,
                                      const float* texture_maps, const float* grad_out,
                                      float* grad_texture_maps, float* grad_texture_coordinates,
                                      void* stream) {
  int rc = tex_check(batch, num_points, channels, tex_height, tex_width, mode);
  if (rc) return rc;
  if (batch == 0 || num_points == 0) return DIBR_OK;
  if (!texture_coordinates || !texture_maps || (!grad_out && channels > 0))
    return fail(DIBR_E_INVALID, "null pointer");
  TexArgs a = {};
  a.coords = texture_coordinates; a.tex = texture_maps; a.grad_out = grad_out;
  a.grad_tex = grad_texture_maps; a.grad_coords = grad_texture_coordinates;
  a.N = num_points; a.B = batch; a.C = channels; a.Ht = tex_height; a.Wt = tex_width;
  a.bilinear = mode;
  texture_backward_kernel<<<blocks_for((long long)batch * num_points, 148 * 32), 256, 0,
                            (cudaStream_t)stream>>>(a);
  EPI_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_mask_iou_sums(int batch, int64_t num_pixels, const float* lhs_mask,
                                const float* rhs_mask, float* sums, void* stream) {
  if (batch < 0 || num_pixels < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  if (batch == 0) return DIBR_OK;
  if (!lhs_mask || !rhs_mask || !sums) return fail(DIBR_E_INVALID, "null pointer");
  EPI_CUDA(cudaMemsetAsync(sums, 0, (size_t)batch * 2 * sizeof(float), (cudaStream_t)stream));
  if (num_pixels == 0) return DIBR_OK;
  const int per_view = blocks_for(num_pixels, (148 * 8 + batch - 1) / batch);
  mask_iou_sums_kernel<<<dim3(per_view, batch), 256, 0, (cudaStream_t)stream>>>(
      lhs_mask, rhs_mask, sums, num_pixels);
  EPI_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_mask_iou_loss(int batch, const float* sums, float* loss, void* stream) {
  if (batch <= 0) return fail(DIBR_E_INVALID, "mask_iou of an empty batch");
  if (!sums || !loss) return fail(DIBR_E_INVALID, "null pointer");
  mask_iou_loss_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(sums, loss, batch);
  EPI_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_mask_iou_bwd(int batch, int64_t num_pixels, const float* lhs_mask,
                               const float* rhs_mask, const float* sums, const float* grad_loss,
                               float* grad_lhs, float* grad_rhs, void* stream) {
  if (batch < 0 || num_pixels < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  if (batch == 0 || num_pixels == 0) return DIBR_OK;
  if (!lhs_mask || !rhs_mask || !sums || !grad_loss) return fail(DIBR_E_INVALID, "null pointer");
  const int per_view = blocks_for(num_pixels, (148 * 16 + batch - 1) / batch);
  mask_iou_backward_kernel<<<dim3(per_view, batch), 256, 0, (cudaStream_t)stream>>>(
      lhs_mask, rhs_mask, sums, grad_loss, grad_lhs, grad_rhs, num_pixels, batch);
  EPI_CUDA(cudaGetLastError());
  return DIBR_OK;
}

}  // extern "C"
