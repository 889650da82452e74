// ===========================================================================
// TEST INFRASTRUCTURE ONLY.  This file is the CPU oracle ("port") of the
// reference's DIB-R hot path.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load it.  The product
// (kaolin-windows_b200/) never imports, links or calls anything in oracle/.
//
// It restates, loop for loop, the four CUDA kernels of the reference
// (kaolin 0.14.0, /root/reference):
//   packed_rasterize_forward_cuda_kernel  kaolin/csrc/render/mesh/rasterization_cuda.cu:43-192
//   rasterize_backward_cuda_kernel        kaolin/csrc/render/mesh/rasterization_cuda.cu:238-402
//   dibr_soft_mask_forward_cuda_kernel    kaolin/csrc/render/mesh/dibr_soft_mask_cuda.cu:27-184
//   dibr_soft_mask_backward_cuda_kernel   kaolin/csrc/render/mesh/dibr_soft_mask_cuda.cu:230-353
// with the same brute-force O(pixels x faces) algorithm: blockIdx/threadIdx
// loops become for-loops over pixels, atomicAdd becomes an add into a
// per-thread private buffer that is summed at the end.  The torch-side
// preparation (rasterization.py:299-363, dibr.py:31-39) lives in
// oracle/dibr_oracle.py.
//
// Arithmetic: scalar_t is float or double exactly as the reference dispatches
// (rasterization_cuda.cu:30-39).  The reference is built by nvcc with the
// default -fmad=true and no fast-math (setup.py:207-211), so nvcc contracts
// a*b+c into FMAs.  Where that decides the *integer* output (the face index:
// edge functions, normalisation, depth) the contraction nvcc 12.9 emits for
// sm_100a (read from cuobjdump -sass of oracle/_ref) is written
// out explicitly with fma(); this file must be compiled with
// -ffp-contract=off so the host compiler adds none of its own.
//
// Parity status: PINNED -- against the reference's golden tensors
// (tests/samples/dibr/**, re-packed under tests/golden/) on the CPU, and
// against the reference's own kernels (oracle/_ref) on the B200 box.
// ===========================================================================
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

#define EPS 1e-7  // dibr_soft_mask_cuda.cu:22 (a double literal, on purpose)

namespace {

inline float fma_(float a, float b, float c) { return std::fmaf(a, b, c); }
inline double fma_(double a, double b, double c) { return std::fma(a, b, c); }

static int g_threads = 0;  // 0 = OpenMP default (all cores)
// bench.py's bounded CPU sample: only pixel rows [g_row_lo, g_row_hi) are processed (the
// per-pixel work is independent, so throughput in pixels/s is unaffected by the band).
static int g_row_lo = 0, g_row_hi = 0x7fffffff;
inline bool row_skipped(int heiidx) { return heiidx < g_row_lo || heiidx >= g_row_hi; }

inline int nthreads() {
#ifdef _OPENMP
  return g_threads > 0 ? g_threads : omp_get_max_threads();
#else
  return 1;
#endif
}

// ---------------------------------------------------------------------------
// rasterization_cuda.cu:43-192
// One (batch, pixel) pair per loop iteration; faces visited in ascending packed
// order, so on equal depth the lowest face index wins (strict '>' at :162).
// ---------------------------------------------------------------------------
template <typename scalar_t>
void packed_rasterize_forward(
    const scalar_t* face_vertices_z,      // (Fv, 3)
    const scalar_t* face_vertices_image,  // (Fv, 3, 2)  already x multiplier
    const scalar_t* face_bboxes,          // (Fv, 4)     xmin ymin xmax ymax
    const scalar_t* face_features,        // (Fv, 3, D)
    const int64_t* first_idx_face_per_mesh,  // (B + 1)
    int64_t* selected_face_idx,           // (B, H, W)   pre-filled with -1
    scalar_t* output_weights,             // (B, H, W, 3) pre-filled with 0
    scalar_t* interpolated_features,      // (B, H, W, D) pre-filled with 0
    int batch_size, int height, int width, int num_features,
    float multiplier, float eps, int64_t* pair_count) {
  int64_t pairs = 0;
#pragma omp parallel for collapse(2) schedule(dynamic, 1) num_threads(nthreads()) reduction(+ : pairs)
  for (int bidx = 0; bidx < batch_size; bidx++) {
    for (int heiidx = 0; heiidx < height; heiidx++) {
      if (row_skipped(heiidx)) continue;
      const int first_id_faces = (int)first_idx_face_per_mesh[bidx];
      const int last_id_faces = (int)first_idx_face_per_mesh[bidx + 1];
      for (int wididx = 0; wididx < width; wididx++) {
        const int pixel_idx = heiidx * width + wididx;
        scalar_t max_z0 = -INFINITY;
        int max_face_idx = -1;
        scalar_t max_w0 = 0., max_w1 = 0., max_w2 = 0.;
        const int64_t totalidx1 = (int64_t)bidx * height * width + pixel_idx;
        const int64_t totalidx3 = totalidx1 * 3;
        const int64_t totalidxd = totalidx1 * num_features;

        // :85-86  float/int division, then int -> scalar_t product
        scalar_t x0 = multiplier / width * (2 * wididx + 1 - width);
        scalar_t y0 = multiplier / height * (height - 2 * heiidx - 1);

        for (int face_idx = first_id_faces; face_idx < last_id_faces; face_idx++) {
          const scalar_t xmin = face_bboxes[face_idx * 4 + 0];
          const scalar_t ymin = face_bboxes[face_idx * 4 + 1];
          const scalar_t xmax = face_bboxes[face_idx * 4 + 2];
          const scalar_t ymax = face_bboxes[face_idx * 4 + 3];
          // :115 half-open box on the pixel centre
          if (x0 < xmin || x0 >= xmax || y0 < ymin || y0 >= ymax) continue;
          pairs++;
          const int64_t shift3 = (int64_t)face_idx * 3;
          const int64_t shift6 = (int64_t)face_idx * 6;
          const scalar_t ax = face_vertices_image[shift6 + 0];
          const scalar_t ay = face_vertices_image[shift6 + 1];
          const scalar_t bx = face_vertices_image[shift6 + 2];
          const scalar_t by = face_vertices_image[shift6 + 3];
          const scalar_t cx = face_vertices_image[shift6 + 4];
          const scalar_t cy = face_vertices_image[shift6 + 5];

          const scalar_t a_edge_x = ax - x0;
          const scalar_t a_edge_y = ay - y0;
          const scalar_t b_edge_x = bx - x0;
          const scalar_t b_edge_y = by - y0;
          const scalar_t c_edge_x = cx - x0;
          const scalar_t c_edge_y = cy - y0;
          // :137-139 as contracted by nvcc: p*q - r*s -> fma(p, q, -(r*s))
          scalar_t w0 = fma_(b_edge_x, c_edge_y, -(b_edge_y * c_edge_x));
          scalar_t w1 = fma_(c_edge_x, a_edge_y, -(c_edge_y * a_edge_x));
          scalar_t w2 = fma_(a_edge_x, b_edge_y, -(a_edge_y * b_edge_x));
          scalar_t norm = w0 + w1 + w2;
          // :141-142 the += runs in double, then rounds back to scalar_t
          norm = (scalar_t)((double)norm +
                            std::copysign(static_cast<double>(eps), static_cast<double>(norm)));
          w0 /= norm;
          w1 /= norm;
          w2 /= norm;
          // :148 (-0.0 and NaN pass)
          if (w0 < 0. || w1 < 0. || w2 < 0.) continue;

          const scalar_t az = face_vertices_z[shift3 + 0];
          const scalar_t bz = face_vertices_z[shift3 + 1];
          const scalar_t cz = face_vertices_z[shift3 + 2];
          // :159 w0*az + w1*bz + w2*cz as contracted by nvcc 12.9 (SASS: FMUL bz*w1;
          // FFMA az*w0 + .; FFMA cz*w2 + .): in "p*q + r*s" the LEFT product is fused.
          const scalar_t z0 = fma_(w2, cz, fma_(w0, az, w1 * bz));
          if (z0 <= max_z0) continue;  // :162
          max_z0 = z0;
          max_face_idx = face_idx;
          max_w0 = w0;
          max_w1 = w1;
          max_w2 = w2;
        }
        if (max_face_idx > -1) {
          selected_face_idx[totalidx1] = max_face_idx - first_id_faces;
          const int64_t shift3d = (int64_t)max_face_idx * 3 * num_features;
          output_weights[totalidx3 + 0] = max_w0;
          output_weights[totalidx3 + 1] = max_w1;
          output_weights[totalidx3 + 2] = max_w2;
          for (int d = 0; d < num_features; d++) {
            const scalar_t r0 = face_features[shift3d + d];
            const scalar_t r1 = face_features[shift3d + num_features + d];
            const scalar_t r2 = face_features[shift3d + num_features + num_features + d];
            // :187 same contraction shape as z0 (SASS: FMUL r1*w1; FFMA r0*w0; FFMA r2*w2)
            interpolated_features[totalidxd + d] =
                fma_(max_w2, r2, fma_(max_w0, r0, max_w1 * r1));
          }
        }
      }
    }
  }
  if (pair_count) *pair_count = pairs;
}

// ---------------------------------------------------------------------------
// rasterization_cuda.cu:238-402.  atomicAdd -> private accumulators per thread
// (acc_t = double by default so the oracle's gradients are the low-noise
// reference the parity tests compare against; the per-pixel terms themselves
// are computed in scalar_t exactly as the kernel does).
// ---------------------------------------------------------------------------
template <typename scalar_t>
void rasterize_backward(
    const scalar_t* grad_interpolated_features,  // (B, H, W, D)
    const int64_t* selected_face_idx,             // (B, H, W)
    const scalar_t* output_weights,               // (B, H, W, 3)
    const scalar_t* face_vertices_image,          // (B, F, 3, 2)  UNscaled
    const scalar_t* face_features,                // (B, F, 3, D)
    scalar_t* grad_face_vertices_image,           // (B, F, 3, 2)
    scalar_t* grad_face_features,                 // (B, F, 3, D)
    int batch_size, int height, int width, int num_faces, int feat_dim, float eps) {
  const int64_t num_pixels = (int64_t)height * width;
  const int64_t n_img = (int64_t)num_faces * 6;
  const int64_t n_feat = (int64_t)num_faces * 3 * feat_dim;
  for (int batch_idx = 0; batch_idx < batch_size; batch_idx++) {
    const int nt = nthreads();
    std::vector<std::vector<double>> acc_img(nt), acc_feat(nt);
#pragma omp parallel num_threads(nt)
    {
#ifdef _OPENMP
      const int tid = omp_get_thread_num();
#else
      const int tid = 0;
#endif
      std::vector<double>& gi = acc_img[tid];
      std::vector<double>& gf = acc_feat[tid];
      gi.assign(n_img, 0.0);
      gf.assign(n_feat, 0.0);
#pragma omp for schedule(dynamic, 1024)
      for (int64_t pixel_idx = 0; pixel_idx < num_pixels; pixel_idx++) {
        if (row_skipped((int)(pixel_idx / width))) continue;
        const int64_t true_pixel_idx = batch_idx * num_pixels + pixel_idx;
        const int64_t start_weight_idx = true_pixel_idx * 3;
        const int64_t start_feat_idx = true_pixel_idx * feat_dim;
        const int face_idx = (int)selected_face_idx[true_pixel_idx];
        if (face_idx < 0) continue;
        const int64_t l_img = (int64_t)face_idx * 6;
        const int64_t l_feat = (int64_t)face_idx * 3 * feat_dim;
        const int64_t start_image_idx = ((int64_t)batch_idx * num_faces + face_idx) * 6;
        const int64_t start_features_idx =
            ((int64_t)batch_idx * num_faces + face_idx) * 3 * feat_dim;

        for (int ii = 0; ii < 3; ii++) {  // :273-285
          scalar_t w = output_weights[start_weight_idx + ii];
          for (int feat_idx = 0; feat_idx < feat_dim; feat_idx++) {
            scalar_t val = grad_interpolated_features[start_feat_idx + feat_idx] * w;
            gf[l_feat + ii * feat_dim + feat_idx] += (double)val;
          }
        }

        const scalar_t ax = face_vertices_image[start_image_idx + 0];
        const scalar_t ay = face_vertices_image[start_image_idx + 1];
        const scalar_t bx = face_vertices_image[start_image_idx + 2];
        const scalar_t by = face_vertices_image[start_image_idx + 3];
        const scalar_t cx = face_vertices_image[start_image_idx + 4];
        const scalar_t cy = face_vertices_image[start_image_idx + 5];

        const scalar_t aw = output_weights[start_weight_idx + 0];
        const scalar_t bw = output_weights[start_weight_idx + 1];
        const scalar_t cw = output_weights[start_weight_idx + 2];

        const scalar_t x0 = aw * ax + bw * bx + cw * cx;  // :302-303
        const scalar_t y0 = aw * ay + bw * by + cw * cy;

        const scalar_t m = bx - ax;
        const scalar_t p = by - ay;
        const scalar_t n = cx - ax;
        const scalar_t q = cy - ay;
        const scalar_t s = x0 - ax;
        const scalar_t t = y0 - ay;

        const scalar_t k1 = s * q - n * t;
        const scalar_t k2 = m * t - s * p;
        scalar_t k3 = m * q - n * p;
        k3 = (scalar_t)((double)k3 +
                        std::copysign(static_cast<double>(eps), static_cast<double>(k3)));

        const scalar_t dk1dm = 0, dk1dn = -t, dk1dp = 0, dk1dq = s, dk1ds = q, dk1dt = -n;
        const scalar_t dk2dm = t, dk2dn = 0, dk2dp = -s, dk2dq = 0, dk2ds = -p, dk2dt = m;
        const scalar_t dk3dm = q, dk3dn = -p, dk3dp = -n, dk3dq = m, dk3ds = 0, dk3dt = 0;

        const scalar_t dw1dm = dk1dm * k3 - dk3dm * k1;
        const scalar_t dw1dn = dk1dn * k3 - dk3dn * k1;
        const scalar_t dw1dp = dk1dp * k3 - dk3dp * k1;
        const scalar_t dw1dq = dk1dq * k3 - dk3dq * k1;
        const scalar_t dw1ds = dk1ds * k3 - dk3ds * k1;
        const scalar_t dw1dt = dk1dt * k3 - dk3dt * k1;

        const scalar_t dw2dm = dk2dm * k3 - dk3dm * k2;
        const scalar_t dw2dn = dk2dn * k3 - dk3dn * k2;
        const scalar_t dw2dp = dk2dp * k3 - dk3dp * k2;
        const scalar_t dw2dq = dk2dq * k3 - dk3dq * k2;
        const scalar_t dw2ds = dk2ds * k3 - dk3ds * k2;
        const scalar_t dw2dt = dk2dt * k3 - dk3dt * k2;

        const scalar_t dw1dax = -(dw1dm + dw1dn + dw1ds);
        const scalar_t dw1day = -(dw1dp + dw1dq + dw1dt);
        const scalar_t dw1dbx = dw1dm, dw1dby = dw1dp, dw1dcx = dw1dn, dw1dcy = dw1dq;
        const scalar_t dw2dax = -(dw2dm + dw2dn + dw2ds);
        const scalar_t dw2day = -(dw2dp + dw2dq + dw2dt);
        const scalar_t dw2dbx = dw2dm, dw2dby = dw2dp, dw2dcx = dw2dn, dw2dcy = dw2dq;

        for (int feat_idx = 0; feat_idx < feat_dim; feat_idx++) {  // :376-399
          const scalar_t c0 = face_features[start_features_idx + feat_idx];
          const scalar_t c1 = face_features[start_features_idx + feat_dim + feat_idx];
          const scalar_t c2 = face_features[start_features_idx + feat_dim + feat_dim + feat_idx];

          const scalar_t dIdax = (c1 - c0) * dw1dax + (c2 - c0) * dw2dax;
          const scalar_t dIday = (c1 - c0) * dw1day + (c2 - c0) * dw2day;
          const scalar_t dIdbx = (c1 - c0) * dw1dbx + (c2 - c0) * dw2dbx;
          const scalar_t dIdby = (c1 - c0) * dw1dby + (c2 - c0) * dw2dby;
          const scalar_t dIdcx = (c1 - c0) * dw1dcx + (c2 - c0) * dw2dcx;
          const scalar_t dIdcy = (c1 - c0) * dw1dcy + (c2 - c0) * dw2dcy;

          const scalar_t dldI = grad_interpolated_features[start_feat_idx + feat_idx] / (k3 * k3);

          gi[l_img + 0] += (double)(scalar_t)(dldI * dIdax);
          gi[l_img + 1] += (double)(scalar_t)(dldI * dIday);
          gi[l_img + 2] += (double)(scalar_t)(dldI * dIdbx);
          gi[l_img + 3] += (double)(scalar_t)(dldI * dIdby);
          gi[l_img + 4] += (double)(scalar_t)(dldI * dIdcx);
          gi[l_img + 5] += (double)(scalar_t)(dldI * dIdcy);
        }
      }
    }
    scalar_t* out_i = grad_face_vertices_image + (int64_t)batch_idx * n_img;
    scalar_t* out_f = grad_face_features + (int64_t)batch_idx * n_feat;
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int64_t i = 0; i < n_img; i++) {
      double v = 0;
      for (int t = 0; t < nt; t++) v += acc_img[t][i];
      out_i[i] = (scalar_t)v;
    }
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int64_t i = 0; i < n_feat; i++) {
      double v = 0;
      for (int t = 0; t < nt; t++) v += acc_feat[t][i];
      out_f[i] = (scalar_t)v;
    }
  }
}

// ---------------------------------------------------------------------------
// dibr_soft_mask_cuda.cu:27-184.  Types follow the reference literally: the
// bare literals EPS / 1.0 are double, so with scalar_t = float the divisions
// by (down + EPS) and the product over (1.0 - prob) run in double and round
// back to float on assignment.
// ---------------------------------------------------------------------------
template <typename scalar_t>
void dibr_soft_mask_forward(
    const scalar_t* face_vertices_image,  // (B, F, 3, 2) already x multiplier
    const scalar_t* face_bboxes,          // (B, F, 4) enlarged boxes
    const int64_t* selected_face_idx,     // (B, H, W)
    scalar_t* close_face_prob,            // (B, H, W, K) pre-filled 0
    int64_t* close_face_idx,              // (B, H, W, K) pre-filled -1
    uint8_t* close_face_dist_type,        // (B, H, W, K) pre-filled 0
    scalar_t* soft_mask,                  // (B, H, W)    pre-filled 0
    int batch_size, int height, int width, int num_faces, int knum,
    float sigmainv, float multiplier) {
#pragma omp parallel for collapse(2) schedule(dynamic, 1) num_threads(nthreads())
  for (int bidx = 0; bidx < batch_size; bidx++) {
    for (int heiidx = 0; heiidx < height; heiidx++) {
      if (row_skipped(heiidx)) continue;
      for (int wididx = 0; wididx < width; wididx++) {
        const int64_t totalidx1 = ((int64_t)bidx * height + heiidx) * width + wididx;
        const int64_t totalidxk = totalidx1 * knum;
        int fidxcover = (int)selected_face_idx[totalidx1];
        if (fidxcover >= 0) {  // :68-70
          soft_mask[totalidx1] = 1.0;
          continue;
        }
        scalar_t x0 = multiplier / width * (2 * wididx + 1 - width);
        scalar_t y0 = multiplier / height * (height - 2 * heiidx - 1);
        int kid = 0;
        for (int fidxint = 0; fidxint < num_faces; fidxint++) {
          const int64_t shift1 = (int64_t)bidx * num_faces + fidxint;
          const int64_t shift4 = shift1 * 4;
          const int64_t shift6 = shift1 * 6;
          scalar_t xmin = face_bboxes[shift4 + 0];
          scalar_t ymin = face_bboxes[shift4 + 1];
          scalar_t xmax = face_bboxes[shift4 + 2];
          scalar_t ymax = face_bboxes[shift4 + 3];
          if (x0 < xmin || x0 >= xmax || y0 < ymin || y0 >= ymax) continue;  // :95-97

          scalar_t pdis[6];
          for (int i = 0; i < 3; i++) {  // :103-141
            int64_t pshift = shift6 + i * 2;
            scalar_t x1 = face_vertices_image[pshift + 0];
            scalar_t y1 = face_vertices_image[pshift + 1];
            int64_t pshift2 = shift6 + ((i + 1) % 3) * 2;
            scalar_t x2 = face_vertices_image[pshift2 + 0];
            scalar_t y2 = face_vertices_image[pshift2 + 1];

            // :113-134.  Each expression is written with the FMA contraction nvcc 12.9
            // emits for it on sm_100a (decoded from the SASS of oracle/_ref; products
            // that have several uses -- A*A, B*B, A*B -- stay separately rounded).
            scalar_t A = y2 - y1;
            scalar_t B = x1 - x2;
            scalar_t C = fma_(x2, y1, -(x1 * y2));            // x2*y1 - x1*y2
            const scalar_t AA = A * A, BB = B * B, AB = A * B;

            scalar_t up = fma_(B, y0, A * x0) + C;            // A*x0 + B*y0 + C
            scalar_t down = AA + BB;                          // A*A + B*B

            // B*B*x0 - A*B*y0 - A*C   and   A*A*y0 - A*B*x0 - B*C
            scalar_t x3 = fma_(-A, C, fma_(BB, x0, -(AB * y0)));
            scalar_t y3 = fma_(-B, C, fma_(AA, y0, -(AB * x0)));
            x3 = x3 / (down + EPS);
            y3 = y3 / (down + EPS);

            scalar_t direct = fma_(x3 - x1, x3 - x2, (y3 - y1) * (y3 - y2));

            if (direct > 0) {
              pdis[i] = 4 * multiplier * multiplier;
            } else {
              pdis[i] = up * up / (down + EPS);
            }
          }
          for (int i = 0; i < 3; i++) {  // :144-149
            int64_t pshift = shift6 + i * 2;
            scalar_t x1 = face_vertices_image[pshift + 0];
            scalar_t y1 = face_vertices_image[pshift + 1];
            pdis[i + 3] = fma_(x0 - x1, x0 - x1, (y0 - y1) * (y0 - y1));  // as contracted
          }
          int edgeid = 0;
          scalar_t dissquare = pdis[0];
          for (int i = 1; i < 6; i++) {  // :154-159 strict '>' keeps the lowest id on ties
            if (dissquare > pdis[i]) {
              dissquare = pdis[i];
              edgeid = i;
            }
          }
          scalar_t z = sigmainv * dissquare / multiplier / multiplier;
          scalar_t prob = std::exp(-z);

          close_face_prob[totalidxk + kid] = prob;
          close_face_idx[totalidxk + kid] = fidxint;
          close_face_dist_type[totalidxk + kid] = (uint8_t)(edgeid + 1);
          kid++;
          if (kid >= knum) break;
        }
        scalar_t allprob = 1.0;
        for (int i = 0; i < kid; i++) {  // :174-178
          scalar_t prob = close_face_prob[totalidxk + i];
          allprob *= (1.0 - prob);
        }
        allprob = 1.0 - allprob;
        soft_mask[totalidx1] = allprob;
      }
    }
  }
}

// ---------------------------------------------------------------------------
// dibr_soft_mask_cuda.cu:230-353
// ---------------------------------------------------------------------------
template <typename scalar_t>
void dibr_soft_mask_backward(
    const scalar_t* grad_soft_mask,       // (B, H, W)
    const scalar_t* soft_mask,            // (B, H, W)
    const int64_t* selected_face_idx,     // (B, H, W)
    const scalar_t* close_face_prob,      // (B, H, W, K)
    const int64_t* close_face_idx,        // (B, H, W, K)
    const uint8_t* close_face_dist_type,  // (B, H, W, K)
    const scalar_t* face_vertices_image,  // (B, F, 3, 2) already x multiplier
    scalar_t* grad_face_vertices_image,   // (B, F, 3, 2)
    int batch_size, int height, int width, int num_faces, int knum,
    float sigmainv, float multiplier) {
  const int64_t n_img = (int64_t)num_faces * 6;
  for (int bidx = 0; bidx < batch_size; bidx++) {
    const int nt = nthreads();
    std::vector<std::vector<double>> acc(nt);
#pragma omp parallel num_threads(nt)
    {
#ifdef _OPENMP
      const int tid = omp_get_thread_num();
#else
      const int tid = 0;
#endif
      std::vector<double>& g = acc[tid];
      g.assign(n_img, 0.0);
#pragma omp for schedule(dynamic, 1)
      for (int heiidx = 0; heiidx < height; heiidx++) {
        if (row_skipped(heiidx)) continue;
        for (int wididx = 0; wididx < width; wididx++) {
          const int64_t totalidx1 = ((int64_t)bidx * height + heiidx) * width + wididx;
          const int64_t totalidxk = totalidx1 * knum;
          scalar_t x0 = multiplier / width * (2 * wididx + 1 - width);
          scalar_t y0 = multiplier / height * (height - 2 * heiidx - 1);
          if ((int)selected_face_idx[totalidx1] >= 0) continue;

          scalar_t dLdp = grad_soft_mask[totalidx1];
          scalar_t allprob = soft_mask[totalidx1];
          for (int kid = 0; kid < knum; kid++) {
            int fidxint = (int)close_face_idx[totalidxk + kid];
            if (fidxint < 0) break;
            const int64_t l6 = (int64_t)fidxint * 6;
            const int64_t shift6 = ((int64_t)bidx * num_faces + fidxint) * 6;
            scalar_t prob = close_face_prob[totalidxk + kid];
            scalar_t dLdz = -1.0 * sigmainv * dLdp * (1.0 - allprob) / (1.0 - prob + EPS) * prob;
            int edgecase = close_face_dist_type[totalidxk + kid];
            int edgeid = edgecase - 1;
            if (edgeid >= 3) {  // :287-302 nearest feature is a vertex
              int pl = (edgeid - 3) * 2;
              scalar_t x1 = face_vertices_image[shift6 + pl + 0];
              scalar_t y1 = face_vertices_image[shift6 + pl + 1];
              scalar_t dLdx1 = dLdz * 2 * (x1 - x0);
              scalar_t dLdy1 = dLdz * 2 * (y1 - y0);
              g[l6 + pl + 0] += (double)(scalar_t)(dLdx1 / multiplier);
              g[l6 + pl + 1] += (double)(scalar_t)(dLdy1 / multiplier);
            } else {  // :304-347 nearest feature is an edge
              int pl = edgeid * 2;
              scalar_t x1 = face_vertices_image[shift6 + pl + 0];
              scalar_t y1 = face_vertices_image[shift6 + pl + 1];
              int pl2 = ((edgeid + 1) % 3) * 2;
              scalar_t x2 = face_vertices_image[shift6 + pl2 + 0];
              scalar_t y2 = face_vertices_image[shift6 + pl2 + 1];

              scalar_t A = y2 - y1;
              scalar_t B = x1 - x2;
              scalar_t C = x2 * y1 - x1 * y2;
              scalar_t up = A * x0 + B * y0 + C;
              scalar_t down = A * A + B * B;
              scalar_t dissquare = up * up / (down + EPS);

              scalar_t dzdA = 2 * (x0 * up - dissquare * A) / (down + EPS);
              scalar_t dzdB = 2 * (y0 * up - dissquare * B) / (down + EPS);
              scalar_t dzdC = 2 * up / (down + EPS);

              scalar_t dLdx1 = dLdz * (dzdB - y2 * dzdC);
              scalar_t dLdy1 = dLdz * (x2 * dzdC - dzdA);
              scalar_t dLdx2 = dLdz * (y1 * dzdC - dzdB);
              scalar_t dLdy2 = dLdz * (dzdA - x1 * dzdC);

              g[l6 + pl + 0] += (double)(scalar_t)(dLdx1 / multiplier);
              g[l6 + pl + 1] += (double)(scalar_t)(dLdy1 / multiplier);
              g[l6 + pl2 + 0] += (double)(scalar_t)(dLdx2 / multiplier);
              g[l6 + pl2 + 1] += (double)(scalar_t)(dLdy2 / multiplier);
            }
          }
        }
      }
    }
    scalar_t* out = grad_face_vertices_image + (int64_t)bidx * n_img;
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int64_t i = 0; i < n_img; i++) {
      double v = 0;
      for (int t = 0; t < nt; t++) v += acc[t][i];
      out[i] = (scalar_t)v;
    }
  }
}

}  // namespace

// ---------------------------------------------------------------------------
// C ABI (host pointers), bound by oracle/dibr_oracle.py with ctypes.
// ---------------------------------------------------------------------------
#define ORACLE_EXPORT(SUFFIX, T)                                                              \
  extern "C" void oracle_packed_rasterize_forward_##SUFFIX(                                   \
      const T* z, const T* img, const T* bbox, const T* feat, const int64_t* first,           \
      int64_t* idx, T* weights, T* interp, int B, int H, int W, int D, float multiplier,      \
      float eps, int64_t* pair_count) {                                                       \
    packed_rasterize_forward<T>(z, img, bbox, feat, first, idx, weights, interp, B, H, W, D,  \
                                multiplier, eps, pair_count);                                 \
  }                                                                                           \
  extern "C" void oracle_rasterize_backward_##SUFFIX(                                         \
      const T* grad, const int64_t* idx, const T* weights, const T* img, const T* feat,       \
      T* grad_img, T* grad_feat, int B, int H, int W, int F, int D, float eps) {              \
    rasterize_backward<T>(grad, idx, weights, img, feat, grad_img, grad_feat, B, H, W, F, D,  \
                          eps);                                                               \
  }                                                                                           \
  extern "C" void oracle_dibr_soft_mask_forward_##SUFFIX(                                     \
      const T* img, const T* bbox, const int64_t* idx, T* prob, int64_t* cidx, uint8_t* type, \
      T* mask, int B, int H, int W, int F, int K, float sigmainv, float multiplier) {         \
    dibr_soft_mask_forward<T>(img, bbox, idx, prob, cidx, type, mask, B, H, W, F, K,          \
                              sigmainv, multiplier);                                          \
  }                                                                                           \
  extern "C" void oracle_dibr_soft_mask_backward_##SUFFIX(                                    \
      const T* grad, const T* mask, const int64_t* idx, const T* prob, const int64_t* cidx,   \
      const uint8_t* type, const T* img, T* grad_img, int B, int H, int W, int F, int K,      \
      float sigmainv, float multiplier) {                                                     \
    dibr_soft_mask_backward<T>(grad, mask, idx, prob, cidx, type, img, grad_img, B, H, W, F,  \
                               K, sigmainv, multiplier);                                      \
  }

ORACLE_EXPORT(f32, float)
ORACLE_EXPORT(f64, double)

extern "C" void oracle_set_threads(int n) { g_threads = n; }
extern "C" void oracle_set_row_band(int lo, int hi) { g_row_lo = lo; g_row_hi = hi; }
extern "C" int oracle_get_threads() { return nthreads(); }
