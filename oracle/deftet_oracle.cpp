// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's deftet_sparse_render
// kernels.  Never imported, linked or executed by the product.
//
// Follows /root/reference (kaolin 0.14.0), loop for loop:
//   kaolin/csrc/render/mesh/deftet_cuda.cu:31-193   forward kernel  -> deftet_forward_*
//   kaolin/csrc/render/mesh/deftet_cuda.cu:239-401  backward kernel -> deftet_backward_*
//   kaolin/csrc/render/mesh/deftet.cpp:84-92        output fills (-1, -inf, 0, 0)
// The warp of the reference (32 faces per step, ballot-compacted insertion) visits faces in
// ascending index and keeps the first knum hits; a serial loop over faces does the same.
// Brute force over all faces per point, like the reference.
//
// Floating point: built with -ffp-contract=off; the FMA shapes below are the ones nvcc 12.9
// emits for the reference source on sm_100a (cuobjdump -sass of oracle/_ref, the deftet
// kernels): a*b - c*d -> fma(a, b, -(c*d)); w0*az + w1*bz + w2*cz -> fma(w2, cz, fma(w0, az,
// w1*bz)); copysignf(eps) added in scalar_t; IEEE divisions.
//
// Parity: pinned by the known answers of the reference's tests
// (tests/python/kaolin/render/mesh/test_deftet.py:104-336) and by outputs of the
// reference's pure-torch _naive_deftet_sparse_render (deftet.py:25-267) generated in the
// build container (tests/golden/deftet_naive.npz) -- see tests/test_deftet_golden.py.
#include <cmath>
#include <cstdint>
#include <limits>

namespace {

inline float fma_t(float a, float b, float c) { return std::fmaf(a, b, c); }
inline double fma_t(double a, double b, double c) { return std::fma(a, b, c); }

template <typename T>
void forward(const T* z, const T* img, const T* bbox, const T* pix, const T* range,
             int64_t* face_ids, T* depths, T* w0_arr, T* w1_arr, int B, int F, int P, int K,
             float eps) {
  const int64_t total = (int64_t)B * P * K;
  for (int64_t i = 0; i < total; i++) {     // deftet.cpp:84-92
    face_ids[i] = -1;
    depths[i] = -std::numeric_limits<T>::infinity();
    w0_arr[i] = 0;
    w1_arr[i] = 0;
  }
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t gp = 0; gp < (int64_t)B * P; gp++) {
    const int b = (int)(gp / P);
    const T x0 = pix[gp * 2], y0 = pix[gp * 2 + 1];
    const T min_depth = range[gp * 2], max_depth = range[gp * 2 + 1];
    int num_depths = 0;
    for (int f = 0; f < F && num_depths < K; f++) {
      const int64_t s1 = (int64_t)b * F + f;
      const T xmin = bbox[s1 * 4], ymin = bbox[s1 * 4 + 1];
      const T xmax = bbox[s1 * 4 + 2], ymax = bbox[s1 * 4 + 3];
      if (!(x0 >= xmin && x0 < xmax && y0 >= ymin && y0 < ymax)) continue;   // :126
      const T* v = img + s1 * 6;
      const T aex = v[0] - x0, aey = v[1] - y0;
      const T bex = v[2] - x0, bey = v[3] - y0;
      const T cex = v[4] - x0, cey = v[5] - y0;
      const T _w0 = fma_t(bex, cey, -(bey * cex));                           // :141
      const T _w1 = fma_t(cex, aey, -(cey * aex));                           // :142
      const T _w2 = fma_t(aex, bey, -(aey * bex));                           // :143
      const T norm = (_w0 + _w1) + _w2;
      const T norm_eps = (T)copysignf(eps, std::signbit(norm) ? -1.f : 1.f); // :145-146
      const T den = norm + norm_eps;
      const T w0 = _w0 / den, w1 = _w1 / den, w2 = _w2 / den;
      if (!(w0 >= 0. && w1 >= 0. && w2 >= 0.)) continue;                     // :152
      const T* zz = z + s1 * 3;
      const T depth = fma_t(w2, zz[2], fma_t(w0, zz[0], w1 * zz[1]));        // :161
      if (!(depth < max_depth && depth >= min_depth)) continue;              // :163
      const int64_t o = gp * K + num_depths;                                 // :183-187
      face_ids[o] = f;
      w0_arr[o] = w0;
      w1_arr[o] = w1;
      depths[o] = depth;
      num_depths++;
    }
  }
}

// serial accumulation in slot order (the reference's atomics have no defined order)
template <typename T>
void backward(const T* grad, const int64_t* face_ids, const T* weights, const T* img,
              const T* feat, T* grad_img, T* grad_feat, int B, int F, int P, int K, int D,
              float eps) {
  for (int64_t i = 0; i < (int64_t)B * F * 6; i++) grad_img[i] = 0;
  for (int64_t i = 0; i < (int64_t)B * F * 3 * D; i++) grad_feat[i] = 0;
  const int64_t total = (int64_t)B * P * K;
  for (int64_t idx = 0; idx < total; idx++) {
    const int64_t face_idx = face_ids[idx];
    if (face_idx < 0) continue;
    const int64_t b = idx / ((int64_t)P * K);
    const int64_t tf = b * F + face_idx;
    const T* w = weights + idx * 3;
    const T* g = grad + idx * D;
    T* gf = grad_feat + tf * 3 * D;
    for (int ii = 0; ii < 3; ii++)
      for (int d = 0; d < D; d++) gf[ii * D + d] += g[d] * w[ii];            // :270-283
    const T* v = img + tf * 6;
    const T ax = v[0], ay = v[1], bx = v[2], by = v[3], cx = v[4], cy = v[5];
    const T aw = w[0], bw = w[1], cw = w[2];
    const T x0 = aw * ax + bw * bx + cw * cx;
    const T y0 = aw * ay + bw * by + cw * cy;
    const T m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay, s = x0 - ax, t = y0 - ay;
    const T k1 = s * q - n * t;
    const T k2 = m * t - s * p;
    T k3 = m * q - n * p;
    k3 = (T)((double)k3 + std::copysign((double)eps, (double)k3));           // :321
    const T dw1dm = 0 * k3 - q * k1, dw1dn = -t * k3 - (-p) * k1, dw1dp = 0 * k3 - (-n) * k1;
    const T dw1dq = s * k3 - m * k1, dw1ds = q * k3, dw1dt = -n * k3;
    const T dw2dm = t * k3 - q * k2, dw2dn = 0 * k3 - (-p) * k2, dw2dp = -s * k3 - (-n) * k2;
    const T dw2dq = 0 * k3 - m * k2, dw2ds = -p * k3, dw2dt = m * k3;
    const T dw1dax = -(dw1dm + dw1dn + dw1ds), dw1day = -(dw1dp + dw1dq + dw1dt);
    const T dw2dax = -(dw2dm + dw2dn + dw2ds), dw2day = -(dw2dp + dw2dq + dw2dt);
    const T* fp = feat + tf * 3 * D;
    T* gi = grad_img + tf * 6;
    for (int d = 0; d < D; d++) {                                            // :376-399
      const T c0 = fp[d], c1 = fp[D + d], c2 = fp[2 * D + d];
      const T dldI = g[d] / (k3 * k3);
      gi[0] += dldI * ((c1 - c0) * dw1dax + (c2 - c0) * dw2dax);
      gi[1] += dldI * ((c1 - c0) * dw1day + (c2 - c0) * dw2day);
      gi[2] += dldI * ((c1 - c0) * dw1dm + (c2 - c0) * dw2dm);
      gi[3] += dldI * ((c1 - c0) * dw1dp + (c2 - c0) * dw2dp);
      gi[4] += dldI * ((c1 - c0) * dw1dn + (c2 - c0) * dw2dn);
      gi[5] += dldI * ((c1 - c0) * dw1dq + (c2 - c0) * dw2dq);
    }
  }
}

}  // namespace

extern "C" {
void deftet_forward_f32(const float* z, const float* img, const float* bbox, const float* pix,
                        const float* range, int64_t* ids, float* depths, float* w0, float* w1,
                        int B, int F, int P, int K, float eps) {
  forward<float>(z, img, bbox, pix, range, ids, depths, w0, w1, B, F, P, K, eps);
}
void deftet_forward_f64(const double* z, const double* img, const double* bbox,
                        const double* pix, const double* range, int64_t* ids, double* depths,
                        double* w0, double* w1, int B, int F, int P, int K, float eps) {
  forward<double>(z, img, bbox, pix, range, ids, depths, w0, w1, B, F, P, K, eps);
}
void deftet_backward_f32(const float* grad, const int64_t* ids, const float* weights,
                         const float* img, const float* feat, float* grad_img, float* grad_feat,
                         int B, int F, int P, int K, int D, float eps) {
  backward<float>(grad, ids, weights, img, feat, grad_img, grad_feat, B, F, P, K, D, eps);
}
void deftet_backward_f64(const double* grad, const int64_t* ids, const double* weights,
                         const double* img, const double* feat, double* grad_img,
                         double* grad_feat, int B, int F, int P, int K, int D, float eps) {
  backward<double>(grad, ids, weights, img, feat, grad_img, grad_feat, B, F, P, K, D, eps);
}
}
