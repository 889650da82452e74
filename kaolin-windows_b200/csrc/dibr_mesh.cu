// dibr_mesh.cu -- the steps either side of the DIB-R path that touch (views x faces)-sized
// data: vertex -> face-corner gathers, unit face normals, and their
// backward scatter-adds (optionally summed over the views of a shared mesh, which is the
// message of the multi-GPU all-reduce).  Everything V-sized (camera transform, projection)
// stays with the caller.
//
//   index_vertices_by_faces          kaolin/ops/mesh/mesh.py:25-47
//   face_normals(unit=True)          kaolin/ops/mesh/trianglemesh.py:313-337
//   prepare_vertices (gather part)   kaolin/render/mesh/utils.py:171-174
#include "../../include/dibr_b200.h"
#include "dibr_common.cuh"

namespace dibr {

int fail(int code, const char* fmt, ...);

namespace {

constexpr int kThreads = 256;

// out[b, f, c, k] = feat[b (or 0 when shared), faces[f, c], k].  One thread per output float
// (coalesced stores; the index and feature reads hit L1/L2); 32-bit index arithmetic inside
// a view (the launcher refuses views of 2^31 or more floats).
__global__ void __launch_bounds__(kThreads)
gather_corners_kernel(const float* __restrict__ feat, const int64_t* __restrict__ faces,
                      float* __restrict__ out, int V, int F, int C, long long feat_view_stride,
                      unsigned per_view) {
  const int b = blockIdx.y;
  const float* src = feat + b * feat_view_stride;
  float* dst = out + (long long)b * per_view;
  for (unsigned e = blockIdx.x * kThreads + threadIdx.x; e < per_view; e += gridDim.x * kThreads) {
    const unsigned fc = e / (unsigned)C, k = e - fc * (unsigned)C;
    const long long v = __ldg(faces + fc);
    dst[e] = (v >= 0 && v < V) ? __ldg(src + v * C + k) : 0.f;
  }
}

// grad_vertex[b or 0, faces[f, c], col + k] += sum over views (when shared) of
// grad_face[b, f, c, k]; one thread per (corner, k), one atomic per thread and view group.
__global__ void __launch_bounds__(kThreads)
scatter_corners_kernel(const float* __restrict__ grad_face, const int64_t* __restrict__ faces,
                       float* __restrict__ grad_vertex, int B, int V, int F, int C, int shared,
                       int out_stride, int out_col, unsigned per_view) {
  const int b0 = shared ? 0 : blockIdx.y, b1 = shared ? B : blockIdx.y + 1;
  for (unsigned e = blockIdx.x * kThreads + threadIdx.x; e < per_view; e += gridDim.x * kThreads) {
    const unsigned fc = e / (unsigned)C, k = e - fc * (unsigned)C;
    const long long v = __ldg(faces + fc);
    if (v < 0 || v >= V) continue;
    float s = 0.f;
    for (int b = b0; b < b1; b++) s += __ldg(grad_face + (long long)b * per_view + e);
    if (s != 0.f)
      atomicAdd(grad_vertex + ((long long)(shared ? 0 : b0) * V + v) * out_stride + out_col + k, s);
  }
}

struct Vec3 { float x, y, z; };
__device__ __forceinline__ Vec3 sub(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ Vec3 cross(Vec3 a, Vec3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ Vec3 load3(const float* p) { return {__ldg(p), __ldg(p + 1), __ldg(p + 2)}; }

constexpr int kSetupFaces = 128;   // faces per CTA; the tile is written out coalesced

// One thread per (view, face): gather the three corners from the per-view vertex arrays,
// unit normal as the reference computes it (cross of the two edges from corner 0, divided by
// length + 1e-10).  Results go through shared memory so that each of the three outputs
// is stored as one contiguous, fully coalesced range per CTA.
__global__ void __launch_bounds__(kSetupFaces)
face_setup_forward_kernel(const float* __restrict__ vcam, const float* __restrict__ vimg,
                          const int64_t* __restrict__ faces, float* __restrict__ fvc,
                          float* __restrict__ fvi, float* __restrict__ fn, int V, int F) {
  __shared__ float s_cam[kSetupFaces * 9], s_img[kSetupFaces * 6], s_n[kSetupFaces * 3];
  const int b = blockIdx.y, f0 = blockIdx.x * kSetupFaces, t = threadIdx.x;
  const int nf = min(kSetupFaces, F - f0);
  if (t < nf) {
    Vec3 p[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
      long long v = __ldg(faces + (long long)(f0 + t) * 3 + c);
      const bool ok = v >= 0 && v < V;
      v = ok ? v : 0;
      p[c] = ok ? load3(vcam + ((long long)b * V + v) * 3) : Vec3{0.f, 0.f, 0.f};
      s_cam[t * 9 + c * 3 + 0] = p[c].x;
      s_cam[t * 9 + c * 3 + 1] = p[c].y;
      s_cam[t * 9 + c * 3 + 2] = p[c].z;
      if (vimg) {
        s_img[t * 6 + c * 2 + 0] = ok ? __ldg(vimg + ((long long)b * V + v) * 2) : 0.f;
        s_img[t * 6 + c * 2 + 1] = ok ? __ldg(vimg + ((long long)b * V + v) * 2 + 1) : 0.f;
      }
    }
    // one definition shared with the mesh-mode renderer's validity test (dibr_common.cuh)
    const float3 n = unit_face_normal(&p[0].x, &p[1].x, &p[2].x);
    s_n[t * 3 + 0] = n.x;
    s_n[t * 3 + 1] = n.y;
    s_n[t * 3 + 2] = n.z;
  }
  __syncthreads();
  const long long base = (long long)b * F + f0;
  if (fvc) for (int i = t; i < nf * 9; i += kSetupFaces) fvc[base * 9 + i] = s_cam[i];
  if (fvi) for (int i = t; i < nf * 6; i += kSetupFaces) fvi[base * 6 + i] = s_img[i];
  if (fn) for (int i = t; i < nf * 3; i += kSetupFaces) fn[base * 3 + i] = s_n[i];
}

// Backward of the above: cotangents of the three outputs (any may be null) scattered into
// the per-view vertex gradients.  n = c / (|c| + eps), c = e0 x e1:
//   dc = g / (L + eps) - c (g . c) / (L (L + eps)^2)   (second term 0 at L = 0, as torch's norm)
//   de0 = e1 x dc, de1 = dc x e0; corner 1 += de0, corner 2 += de1, corner 0 -= de0 + de1.
__global__ void __launch_bounds__(kThreads)
face_setup_backward_kernel(const float* __restrict__ vcam, const int64_t* __restrict__ faces,
                           const float* __restrict__ g_fvc, const float* __restrict__ g_fvi,
                           const float* __restrict__ g_fn, float* __restrict__ g_vcam,
                           float* __restrict__ g_vimg, int V, int F) {
  const int b = blockIdx.y;
  const int f = blockIdx.x * kThreads + threadIdx.x;
  if (f >= F) return;
  long long v[3];
  bool ok = true;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    v[c] = __ldg(faces + (long long)f * 3 + c);
    ok = ok && v[c] >= 0 && v[c] < V;
  }
  if (!ok) return;
  const long long bf = (long long)b * F + f;
  Vec3 g[3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  if (g_fvc) {
#pragma unroll
    for (int c = 0; c < 3; c++) g[c] = load3(g_fvc + bf * 9 + c * 3);
  }
  if (g_fn) {
    const Vec3 gn = load3(g_fn + bf * 3);
    if (gn.x != 0.f || gn.y != 0.f || gn.z != 0.f) {
      Vec3 p[3];
#pragma unroll
      for (int c = 0; c < 3; c++) p[c] = load3(vcam + ((long long)b * V + v[c]) * 3);
      const Vec3 e0 = sub(p[1], p[0]), e1 = sub(p[2], p[0]);
      const Vec3 c = cross(e0, e1);
      const float L = sqrtf(dot(c, c)), Le = L + 1e-10f;
      const float k = L > 0.f ? dot(gn, c) / (L * Le * Le) : 0.f;
      const Vec3 dc = {gn.x / Le - c.x * k, gn.y / Le - c.y * k, gn.z / Le - c.z * k};
      const Vec3 de0 = cross(e1, dc), de1 = cross(dc, e0);
      g[1].x += de0.x; g[1].y += de0.y; g[1].z += de0.z;
      g[2].x += de1.x; g[2].y += de1.y; g[2].z += de1.z;
      g[0].x -= de0.x + de1.x; g[0].y -= de0.y + de1.y; g[0].z -= de0.z + de1.z;
    }
  }
  if (g_vcam && (g_fvc || g_fn)) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float* dst = g_vcam + ((long long)b * V + v[c]) * 3;
      if (g[c].x != 0.f) atomicAdd(dst, g[c].x);
      if (g[c].y != 0.f) atomicAdd(dst + 1, g[c].y);
      if (g[c].z != 0.f) atomicAdd(dst + 2, g[c].z);
    }
  }
  if (g_vimg && g_fvi) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float gx = __ldg(g_fvi + bf * 6 + c * 2), gy = __ldg(g_fvi + bf * 6 + c * 2 + 1);
      if (gx != 0.f || gy != 0.f)      // 8-byte aligned: (b V + v) * 2 floats
        atomicAdd(reinterpret_cast<float2*>(g_vimg + ((long long)b * V + v[c]) * 2),
                  make_float2(gx, gy));
    }
  }
}

inline int grid_for(long long n, int threads) {
  long long g = (n + threads - 1) / threads;
  return (int)(g < 1 ? 1 : (g > 148 * 64 ? 148 * 64 : g));
}


// ---- camera transform + perspective projection (prepare_vertices before the gathers:
// kaolin/render/mesh/utils.py:158-170, camera/legacy.py:22-37,120-139), V-sized.
//   mode 0: vcam = (v - trans_b) @ rot_b^T        (rotate_translate_points)
//   mode 1: vcam = [v, 1] @ transform_b (4 x 3)   (camera_transform)
//   vimg = (vcam.xy * proj.xy) / (vcam.z * proj.z)
// One thread per (view, vertex); `shared`: one (V, 3) vertex array for all views.
__global__ void __launch_bounds__(kThreads)
vertex_transform_forward_kernel(const float* __restrict__ vertices, int shared,
                                const float* __restrict__ mat, const float* __restrict__ trans,
                                const float* __restrict__ proj, int mode,
                                float* __restrict__ vcam, float* __restrict__ vimg, int V) {
  const int b = blockIdx.y;
  const int v = blockIdx.x * kThreads + threadIdx.x;
  if (v >= V) return;
  const float* p = vertices + ((shared ? 0ll : (long long)b * V) + v) * 3;
  float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
  float cx, cy, cz;
  if (mode == 0) {
    const float* R = mat + (long long)b * 9;
    const float* t = trans + (long long)b * 3;
    x -= __ldg(t); y -= __ldg(t + 1); z -= __ldg(t + 2);
    cx = x * __ldg(R) + y * __ldg(R + 1) + z * __ldg(R + 2);
    cy = x * __ldg(R + 3) + y * __ldg(R + 4) + z * __ldg(R + 5);
    cz = x * __ldg(R + 6) + y * __ldg(R + 7) + z * __ldg(R + 8);
  } else {
    const float* T = mat + (long long)b * 12;
    cx = x * __ldg(T) + y * __ldg(T + 3) + z * __ldg(T + 6) + __ldg(T + 9);
    cy = x * __ldg(T + 1) + y * __ldg(T + 4) + z * __ldg(T + 7) + __ldg(T + 10);
    cz = x * __ldg(T + 2) + y * __ldg(T + 5) + z * __ldg(T + 8) + __ldg(T + 11);
  }
  const long long o = (long long)b * V + v;
  vcam[o * 3] = cx; vcam[o * 3 + 1] = cy; vcam[o * 3 + 2] = cz;
  const float den = cz * __ldg(proj + 2);
  vimg[o * 2] = cx * __ldg(proj) / den;
  vimg[o * 2 + 1] = cy * __ldg(proj + 1) / den;
}

// Backward: cotangents of vcam (may be null) and vimg -> grad_vertices.  64 vertices x 4
// view lanes per CTA; `shared`: the view lanes stride over all views and their partial sums
// meet in shared memory (no atomics); else blockIdx.y selects the view and lane 0 works.
__global__ void __launch_bounds__(kThreads)
vertex_transform_backward_kernel(const float* __restrict__ vcam, const float* __restrict__ mat,
                                 const float* __restrict__ proj, int mode,
                                 const float* __restrict__ g_vcam,
                                 const float* __restrict__ g_vimg, float* __restrict__ g_vertices,
                                 int B, int V, int shared) {
  __shared__ float s_part[4][64][3];
  const int vl = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int v = blockIdx.x * 64 + vl;
  const int b0 = shared ? ty : blockIdx.y, b1 = shared ? B : blockIdx.y + 1;
  const int bstep = shared ? 4 : 1;
  const float p0 = __ldg(proj), p1 = __ldg(proj + 1), p2 = __ldg(proj + 2);
  float gx = 0.f, gy = 0.f, gz = 0.f;
  if (v < V && (shared || ty == 0)) {
    for (int b = b0; b < b1; b += bstep) {
      const long long o = (long long)b * V + v;
      const float X = __ldg(vcam + o * 3), Y = __ldg(vcam + o * 3 + 1), Z = __ldg(vcam + o * 3 + 2);
      float cx = 0.f, cy = 0.f, cz = 0.f;
      if (g_vimg) {
        const float gu = __ldg(g_vimg + o * 2), gv = __ldg(g_vimg + o * 2 + 1);
        const float inv = 1.0f / (Z * p2);
        cx = gu * p0 * inv;
        cy = gv * p1 * inv;
        cz = -(gu * X * p0 + gv * Y * p1) * inv / Z;
      }
      if (g_vcam) { cx += __ldg(g_vcam + o * 3); cy += __ldg(g_vcam + o * 3 + 1); cz += __ldg(g_vcam + o * 3 + 2); }
      if (mode == 0) {
        const float* R = mat + (long long)b * 9;
        gx += cx * __ldg(R) + cy * __ldg(R + 3) + cz * __ldg(R + 6);
        gy += cx * __ldg(R + 1) + cy * __ldg(R + 4) + cz * __ldg(R + 7);
        gz += cx * __ldg(R + 2) + cy * __ldg(R + 5) + cz * __ldg(R + 8);
      } else {
        const float* T = mat + (long long)b * 12;
        gx += cx * __ldg(T) + cy * __ldg(T + 1) + cz * __ldg(T + 2);
        gy += cx * __ldg(T + 3) + cy * __ldg(T + 4) + cz * __ldg(T + 5);
        gz += cx * __ldg(T + 6) + cy * __ldg(T + 7) + cz * __ldg(T + 8);
      }
    }
  }
  s_part[ty][vl][0] = gx; s_part[ty][vl][1] = gy; s_part[ty][vl][2] = gz;
  __syncthreads();
  if (ty == 0 && v < V) {
    if (shared) {
#pragma unroll
      for (int k = 1; k < 4; k++) { gx += s_part[k][vl][0]; gy += s_part[k][vl][1]; gz += s_part[k][vl][2]; }
    }
    float* g = g_vertices + ((shared ? 0ll : (long long)blockIdx.y * V) + v) * 3;
    g[0] = gx; g[1] = gy; g[2] = gz;
  }
}

}  // namespace
}  // namespace dibr

using namespace dibr;

#define MESH_CUDA(expr)                                                              \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess)                                                           \
      return fail(DIBR_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));      \
  } while (0)

extern "C" {

DIBR_API int dibr_index_vertices_by_faces(int batch, int num_vertices, int num_faces, int dim,
                                          int shared, const float* vertex_features,
                                          const int64_t* faces, float* face_features,
                                          void* stream) {
  if (batch < 0 || num_vertices < 0 || num_faces < 0 || dim < 0)
    return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  const long long per_view = (long long)num_faces * 3 * dim;
  if (batch == 0 || per_view == 0) return DIBR_OK;
  if (per_view >= (1ll << 31)) return fail(DIBR_E_INVALID, "more than 2^31 values per view");
  if (!vertex_features || !faces || !face_features) return fail(DIBR_E_INVALID, "null pointer");
  gather_corners_kernel<<<dim3(grid_for(per_view, kThreads), batch), kThreads, 0,
                          (cudaStream_t)stream>>>(
      vertex_features, faces, face_features, num_vertices, num_faces, dim,
      shared ? 0 : (long long)num_vertices * dim, (unsigned)per_view);
  MESH_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_scatter_face_grads(int batch, int num_vertices, int num_faces, int dim,
                                     int shared, const float* grad_face_features,
                                     const int64_t* faces, float* grad_vertex_features,
                                     int out_stride, int out_col, void* stream) {
  if (batch < 0 || num_vertices < 0 || num_faces < 0 || dim < 0 || out_col < 0 ||
      out_stride < out_col + dim)
    return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  const long long per_view = (long long)num_faces * 3 * dim;
  if (batch == 0 || per_view == 0) return DIBR_OK;
  if (per_view >= (1ll << 31)) return fail(DIBR_E_INVALID, "more than 2^31 values per view");
  if (!grad_face_features || !faces || !grad_vertex_features)
    return fail(DIBR_E_INVALID, "null pointer");
  scatter_corners_kernel<<<dim3(grid_for(per_view, kThreads), shared ? 1 : batch), kThreads, 0,
                           (cudaStream_t)stream>>>(
      grad_face_features, faces, grad_vertex_features, batch, num_vertices, num_faces, dim,
      shared, out_stride, out_col, (unsigned)per_view);
  MESH_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_face_setup_fwd(int batch, int num_vertices, int num_faces,
                                 const float* vertices_camera, const float* vertices_image,
                                 const int64_t* faces, float* face_vertices_camera,
                                 float* face_vertices_image, float* face_normals, void* stream) {
  if (batch < 0 || num_vertices < 0 || num_faces < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  if (batch == 0 || num_faces == 0) return DIBR_OK;
  if (!vertices_camera || !faces) return fail(DIBR_E_INVALID, "null pointer");
  if (face_vertices_image && !vertices_image)
    return fail(DIBR_E_INVALID, "face_vertices_image wanted without vertices_image");
  face_setup_forward_kernel<<<dim3((num_faces + kSetupFaces - 1) / kSetupFaces, batch),
                              kSetupFaces, 0, (cudaStream_t)stream>>>(
      vertices_camera, face_vertices_image ? vertices_image : nullptr, faces,
      face_vertices_camera, face_vertices_image, face_normals, num_vertices, num_faces);
  MESH_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_face_setup_bwd(int batch, int num_vertices, int num_faces,
                                 const float* vertices_camera, const int64_t* faces,
                                 const float* grad_face_vertices_camera,
                                 const float* grad_face_vertices_image,
                                 const float* grad_face_normals, float* grad_vertices_camera,
                                 float* grad_vertices_image, void* stream) {
  if (batch < 0 || num_vertices < 0 || num_faces < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  if (batch == 0 || num_faces == 0) return DIBR_OK;
  if (!vertices_camera || !faces) return fail(DIBR_E_INVALID, "null pointer");
  if ((grad_face_vertices_camera || grad_face_normals) && !grad_vertices_camera)
    return fail(DIBR_E_INVALID, "null grad_vertices_camera");
  if (grad_face_vertices_image && !grad_vertices_image)
    return fail(DIBR_E_INVALID, "null grad_vertices_image");
  face_setup_backward_kernel<<<dim3((num_faces + kThreads - 1) / kThreads, batch), kThreads, 0,
                               (cudaStream_t)stream>>>(
      vertices_camera, faces, grad_face_vertices_camera, grad_face_vertices_image,
      grad_face_normals, grad_vertices_camera, grad_vertices_image, num_vertices, num_faces);
  MESH_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_vertex_transform_fwd(int batch, int num_vertices, int vertices_shared,
                                       const float* vertices, const float* camera_rot,
                                       const float* camera_trans, const float* camera_transform,
                                       const float* camera_proj, float* vertices_camera,
                                       float* vertices_image, void* stream) {
  if (batch < 0 || num_vertices < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  if (batch == 0 || num_vertices == 0) return DIBR_OK;
  const bool rt = camera_rot && camera_trans && !camera_transform;
  const bool tf = camera_transform && !camera_rot && !camera_trans;
  if (!rt && !tf)
    return fail(DIBR_E_INVALID, "camera_transform or camera_trans and camera_rot must be defined");
  if (!vertices || !camera_proj || !vertices_camera || !vertices_image)
    return fail(DIBR_E_INVALID, "null pointer");
  vertex_transform_forward_kernel<<<dim3((num_vertices + kThreads - 1) / kThreads, batch),
                                    kThreads, 0, (cudaStream_t)stream>>>(
      vertices, vertices_shared, rt ? camera_rot : camera_transform, camera_trans, camera_proj,
      rt ? 0 : 1, vertices_camera, vertices_image, num_vertices);
  MESH_CUDA(cudaGetLastError());
  return DIBR_OK;
}

DIBR_API int dibr_vertex_transform_bwd(int batch, int num_vertices, int vertices_shared,
                                       const float* vertices_camera, const float* camera_rot,
                                       const float* camera_transform, const float* camera_proj,
                                       const float* grad_vertices_camera,
                                       const float* grad_vertices_image, float* grad_vertices,
                                       void* stream) {
  if (batch < 0 || num_vertices < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (batch > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", batch);
  if (num_vertices == 0) return DIBR_OK;
  if ((camera_rot != nullptr) == (camera_transform != nullptr))
    return fail(DIBR_E_INVALID, "exactly one of camera_rot / camera_transform");
  if (!vertices_camera || !camera_proj || !grad_vertices)
    return fail(DIBR_E_INVALID, "null pointer");
  if (batch == 0) {
    if (vertices_shared)
      MESH_CUDA(cudaMemsetAsync(grad_vertices, 0, (size_t)num_vertices * 3 * sizeof(float),
                                (cudaStream_t)stream));
    return DIBR_OK;
  }
  vertex_transform_backward_kernel<<<dim3((num_vertices + 63) / 64,
                                          vertices_shared ? 1 : batch),
                                     kThreads, 0, (cudaStream_t)stream>>>(
      vertices_camera, camera_rot ? camera_rot : camera_transform, camera_proj,
      camera_rot ? 0 : 1, grad_vertices_camera, grad_vertices_image, grad_vertices, batch,
      num_vertices, vertices_shared);
  MESH_CUDA(cudaGetLastError());
  return DIBR_OK;
}

}  // extern "C"
