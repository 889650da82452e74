/*
 * dibr_b200.h -- C ABI of libdibr_b200.so: the B200-native (sm_100a) DIB-R
 * differentiable-rasterization hot path.
 *
 * Drop-in boundary.  These entry points replace the four functions the
 * reference registers with pybind11 at kaolin/csrc/bindings.cpp:76-81
 *   kaolin._C.render.mesh.packed_rasterize_forward_cuda   (csrc/render/mesh/rasterization.h:23-32)
 *   kaolin._C.render.mesh.rasterize_backward_cuda         (csrc/render/mesh/rasterization.h:34-41)
 *   kaolin._C.render.mesh.dibr_soft_mask_forward_cuda     (csrc/render/mesh/dibr_soft_mask.h:23-29)
 *   kaolin._C.render.mesh.dibr_soft_mask_backward_cuda    (csrc/render/mesh/dibr_soft_mask.h:31-41)
 * and the torch-side preparation around them in kaolin/render/mesh/rasterization.py:299-363
 * and kaolin/render/mesh/dibr.py:31-39, which the kernels here fold in.
 *
 * Conventions
 *  - plain C: raw DEVICE pointers, sizes and scalars; no ATen / torch types.
 *  - every tensor is C-contiguous fp32 unless stated; int64 where the reference uses
 *    int64 (selected_face_idx, close_face_idx); uint8 for bool masks / dist types.
 *  - the library never allocates or frees device memory and keeps no state between
 *    calls: outputs and the scratch workspace are allocated by the caller.
 *  - all work is enqueued on `stream` (a cudaStream_t passed as void*); no host
 *    synchronisation happens inside unless documented (needed_host == NULL).
 *  - return value: DIBR_OK (0) or a negative DIBR_E_* code; dibr_last_error() returns
 *    a thread-local message for the last failure.
 *  - there is no CPU fallback: without a CUDA device every entry point fails.
 *
 * Workspace protocol.  The face-binning pass builds per-bin face lists whose total
 * length is data dependent.  The caller passes `ws` of `ws_bytes` bytes laid out as
 *    [ fixed part | raster list: raster_capacity entries | soft list: soft_capacity int32 ]
 * (a raster entry is a 48-byte face record, or an int32 face id with DIBR_B200_RASTER_REC=0)
 * with ws_bytes >= dibr_workspace_bytes_for(..., raster_capacity, soft_capacity); a call
 * that does not build a list ignores that list's capacity.  The number of list entries
 * the call needed is written as int64[2] = {raster, soft} to `needed_host`:
 *    needed_host != NULL : pinned host memory, written asynchronously on `stream`; the
 *                          caller checks it after its own event/stream sync and, if a
 *                          value exceeds the capacity it passed, grows `ws` and calls
 *                          again (outputs of the overflowing call are invalid, but no
 *                          out-of-bounds access happens).
 *    needed_host == NULL : the call synchronises `stream` once internally and returns
 *                          DIBR_E_WORKSPACE if a capacity was insufficient
 *                          (dibr_last_error() then names the needed entry counts).
 * The backward of the soft mask re-reads the soft list, so `ws` must be kept unchanged
 * between dibr_softmask_fwd / dibr_render_fwd and dibr_softmask_bwd, which takes the
 * same capacities.
 *
 * Pair buffer (optional, pair_capacity > 0 and 0 < knum <= 1024).  The soft-mask forward
 * then also leaves in `ws`, after the two lists, the compact per-pixel face lists it found
 * ("pairs": face id, probability, nearest-feature type per (pixel, face)), allocated in
 * segments of 8 hits x 32 pixels; pair_capacity is the number of such segments (and the
 * maximum number of 8x4-pixel regions holding any hit).  The backward then runs over the
 * saved pairs instead of re-walking the lists.  If the buffer is too small nothing breaks:
 * both directions fall back, decided on the device, to the self-contained kernels.  The
 * segments / regions the call wanted are reported as needed_host[2], needed_host[3]
 * (needed_host is int64[4]; entries 2-3 are written when the gather kernel has finished, so
 * a caller reads them later, e.g. before its next call, to size the buffer).
 */
#ifndef DIBR_B200_H_
#define DIBR_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define DIBR_API __attribute__((visibility("default")))
#else
#define DIBR_API
#endif

#define DIBR_OK 0
#define DIBR_E_INVALID (-1)   /* bad argument (shape, null pointer, unsupported size) */
#define DIBR_E_WORKSPACE (-2) /* workspace too small (see protocol above)             */
#define DIBR_E_CUDA (-3)      /* CUDA runtime error; message in dibr_last_error()      */

DIBR_API const char* dibr_last_error(void);
DIBR_API int dibr_version(void);
/* needed_host reports (protocol above) are normally a 16-byte device-to-host copy between two
 * kernels.  If every needed_host the caller passes is pinned memory the device can address
 * (cudaHostAlloc / cudaMallocHost under unified addressing -- torch's pin_memory() is), turn
 * this on and the kernels store the report themselves: no copy node breaks the kernel chain.
 * Process-wide; returns the previous setting. */
DIBR_API int dibr_host_report_direct(int on);

/* Measurement aid (used by bench.py only): between dibr_profile_begin() and
 * dibr_profile_end() every kernel launch of this library is bracketed by CUDA events on
 * its launching stream.  dibr_profile_end synchronises those events and returns the number
 * of records written: kernel_ids[i] in {0 bin_count, 1 bin_scan, 2 bin_fill, 3 raster_fwd,
 * 4 softmask_fwd (fallback), 5 raster_bwd, 6 softmask_bwd (fallback), 7 softmask_bwd_dense,
 * 8 softmask_gather, 9 softmask_eval, 10 softmask_grad, 11 softmask_fold}, ms[i] its duration. */
DIBR_API int dibr_profile_begin(void);
DIBR_API int dibr_profile_end(int* kernel_ids, float* ms, int capacity);

/* Bytes of the fixed (data-independent) part of the workspace for a call over `batch`
 * views of `num_faces` faces (for the packed mode: the total number of packed faces) at
 * height x width.  Returns < 0 on invalid sizes. */
DIBR_API int64_t dibr_workspace_fixed_bytes(int batch, int num_faces, int height, int width);
/* fixed bytes + room for raster_capacity + soft_capacity list entries + a pair buffer of
 * pair_capacity segments for `knum` (0 / 0 for calls that use none). */
DIBR_API int64_t dibr_workspace_bytes_for(int batch, int num_faces, int height, int width,
                                 int64_t raster_capacity, int64_t soft_capacity,
                                 int64_t pair_capacity, int knum);

/*
 * rasterize forward.  Replaces RasterizeCuda.forward (rasterization.py:289-369) +
 * packed_rasterize_forward_cuda (rasterization.cpp:49-104, rasterization_cuda.cu:43-192).
 *
 *   face_vertices_z      (B, F, 3)
 *   face_vertices_image  (B, F, 3, 2)   NDC, NOT multiplied (the kernel multiplies)
 *   face_features        (B, F, 3, D)
 *   valid_faces          (B, F) uint8/bool or NULL
 *   face_normals_z       (B, F) or NULL; when given, valid = (normal_z >= 0)
 *                        (dibr.py:190-199); at most one of the two may be non-NULL
 * Packed / pre-scaled mode (the raw _C op signature): `prescaled` != 0 means
 *   face_vertices_image is already multiplied, `face_bboxes` (.,4) is given and
 *   `first_idx_face_per_mesh` (B+1) int64 delimits each mesh's faces inside packed
 *   arrays of `num_faces` faces in total; selected_face_idx is then relative to the mesh.
 * Outputs (every element is written):
 *   selected_face_idx      (B, H, W) int64, -1 = background
 *   interpolated_features  (B, H, W, D)
 *   output_weights         (B, H, W, 3) or NULL (the fast path recomputes them in backward)
 */
DIBR_API int dibr_rasterize_fwd(int batch, int num_faces, int height, int width, int feat_dim,
                       const float* face_vertices_z, const float* face_vertices_image,
                       const float* face_features, const uint8_t* valid_faces,
                       const float* face_normals_z, int prescaled, const float* face_bboxes,
                       const int64_t* first_idx_face_per_mesh, float multiplier, float eps,
                       int64_t* selected_face_idx, float* interpolated_features,
                       float* output_weights, void* ws, int64_t ws_bytes,
                       int64_t raster_capacity, int64_t* needed_host, void* stream);

/*
 * rasterize forward WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot allocate device memory (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same inputs,
 * outputs and arithmetic as dibr_rasterize_fwd in its plain mode (unscaled
 * face_vertices_image, boxes computed; valid_faces / face_normals_z optional), but the
 * scratch workspace lives in a file-scope device array of DIBR_STATIC_SCRATCH_BYTES inside
 * the library (every GPU that loads the library holds it until exit); the call makes no
 * cudaMalloc / cudaFree and no allocating library call.  The caller still passes inputs and
 * outputs.
 *  - Sizes whose tables plus 2 list entries per face-view (+ 4096) do not fit are rejected
 *    with DIBR_E_INVALID (the message names the bytes needed and the maximum): use
 *    dibr_rasterize_fwd with a caller-allocated workspace for those.
 *  - The list capacity is everything that fits: (DIBR_STATIC_SCRATCH_BYTES - 512 -
 *    dibr_workspace_fixed_bytes(batch, num_faces, height, width)) / 48 entries (4-byte entries
 *    with DIBR_B200_RASTER_REC=0).  needed_host as in the workspace protocol: NULL = one
 *    internal stream synchronisation and DIBR_E_WORKSPACE if the data needed more entries;
 *    else int64[2] pinned memory written asynchronously, needed_host[0] > capacity means the
 *    outputs are invalid.
 *  - ALL calls share the one array, which keeps the previous call's contents: each call
 *    clears what must start at zero itself, but calls that overlap on the same GPU (two
 *    streams, two threads) race -- serialise them.
 */
#define DIBR_STATIC_SCRATCH_BYTES (16 * 1024 * 1024)
DIBR_API int dibr_rasterize_fwd_static(int batch, int num_faces, int height, int width,
                              int feat_dim, const float* face_vertices_z,
                              const float* face_vertices_image, const float* face_features,
                              const uint8_t* valid_faces, const float* face_normals_z,
                              float multiplier, float eps, int64_t* selected_face_idx,
                              float* interpolated_features, float* output_weights,
                              int64_t* needed_host, void* stream);

/*
 * rasterize backward.  Replaces rasterize_backward_cuda (rasterization.cpp:106-168,
 * rasterization_cuda.cu:238-402).  face_vertices_image is UNscaled (B, F, 3, 2).
 * output_weights may be NULL: the weights are then recomputed bit-identically from
 * (face_vertices_image * multiplier, pixel centre, eps) as the forward did.
 * Both gradients are ACCUMULATED into (the caller zero-fills them, as at::zeros_like at
 * rasterization.cpp:156-157 does).  With output_weights the per-pixel formula of the
 * reference is evaluated (in double) on exactly those weights; without, the pixel-centre
 * form (DESIGN.md "Backward").
 * face_vertices_z receives no gradient (rasterization.py:387-388).
 */
DIBR_API int dibr_rasterize_bwd(int batch, int num_faces, int height, int width, int feat_dim,
                       const float* grad_interpolated_features, const int64_t* selected_face_idx,
                       const float* output_weights, const float* face_vertices_image,
                       const float* face_features, float multiplier, float eps,
                       float* grad_face_vertices_image, float* grad_face_features, void* stream);

/*
 * soft mask forward.  Replaces DibrSoftMaskCuda.forward (dibr.py:29-55) +
 * dibr_soft_mask_forward_cuda (dibr_soft_mask.cpp:48-108, dibr_soft_mask_cuda.cu:27-184).
 *   face_vertices_image (B, F, 3, 2): unscaled when prescaled == 0 (the kernel multiplies
 *   and enlarges the boxes by box_margin = (float)(boxlen * multiplier), the product
 *   taken in double by the caller exactly as dibr.py:37-38 does); when prescaled != 0 it is
 *   already multiplied and face_large_bboxes (B, F, 4) is used as given (raw _C op).
 * Outputs: soft_mask (B,H,W) always; the dense per-pixel state of the reference
 *   close_face_prob (B,H,W,K) f32 / close_face_idx (B,H,W,K) int64 / close_face_dist_type
 *   (B,H,W,K) uint8 only when all three are non-NULL (they must be pre-filled 0 / -1 / 0
 *   as dibr_soft_mask.cpp:84-97 does; only used slots are written).
 */
DIBR_API int dibr_softmask_fwd(int batch, int num_faces, int height, int width,
                      const float* face_vertices_image, int prescaled,
                      const float* face_large_bboxes, const int64_t* selected_face_idx,
                      float sigmainv, float box_margin, int knum, float multiplier,
                      float* soft_mask, float* close_face_prob, int64_t* close_face_idx,
                      uint8_t* close_face_dist_type, void* ws, int64_t ws_bytes,
                      int64_t raster_capacity, int64_t soft_capacity, int64_t pair_capacity,
                      int64_t* needed_host, void* stream);

/*
 * soft mask backward (fast path): runs over the pairs left in `ws` by dibr_softmask_fwd /
 * dibr_render_fwd or, without a (large enough) pair buffer, re-walks the face lists left
 * there -- instead of reading the reference's 13*K bytes/pixel of saved state.
 * Same arguments as the forward (they must be the SAME values) + grad_soft_mask, the
 * forward's soft_mask, and grad_face_vertices_image (B,F,3,2), accumulated into, w.r.t.
 * the UNscaled coordinates (every term is divided by multiplier, dibr_soft_mask_cuda.cu:298-346).
 */
DIBR_API int dibr_softmask_bwd(int batch, int num_faces, int height, int width,
                      const float* face_vertices_image, int prescaled,
                      const float* face_large_bboxes, const int64_t* selected_face_idx,
                      float sigmainv, float box_margin, int knum, float multiplier,
                      const float* grad_soft_mask, const float* soft_mask,
                      float* grad_face_vertices_image, const void* ws, int64_t ws_bytes,
                      int64_t raster_capacity, int64_t soft_capacity, int64_t pair_capacity,
                      void* stream);

/*
 * soft mask backward, dense-state form: the raw _C op dibr_soft_mask_backward_cuda
 * (dibr_soft_mask.cpp:110-183, dibr_soft_mask_cuda.cu:230-353).  face_vertices_image is
 * the SCALED tensor the reference saves (dibr.py:51-54).
 */
DIBR_API int dibr_softmask_bwd_dense(int batch, int num_faces, int height, int width, int knum,
                            const float* grad_soft_mask, const float* soft_mask,
                            const int64_t* selected_face_idx, const float* close_face_prob,
                            const int64_t* close_face_idx, const uint8_t* close_face_dist_type,
                            const float* face_vertices_image_scaled, float sigmainv,
                            float multiplier, float* grad_face_vertices_image, void* stream);

/*
 * dibr_rasterization forward, fused (dibr.py:119-209): one binning pass builds both
 * lists, then the rasterizer and the soft mask run back to back.  valid = normal_z >= 0
 * for the rasterizer, ALL faces for the soft mask (dibr.py:201-203).
 */
DIBR_API int dibr_render_fwd(int batch, int num_faces, int height, int width, int feat_dim,
                    const float* face_vertices_z, const float* face_vertices_image,
                    const float* face_features, const float* face_normals_z, float sigmainv,
                    float box_margin, int knum, float multiplier, float eps,
                    int64_t* selected_face_idx, float* interpolated_features, float* soft_mask,
                    void* ws, int64_t ws_bytes, int64_t raster_capacity, int64_t soft_capacity,
                    int64_t pair_capacity, int64_t* needed_host, void* stream);

/* Backward of dibr_render_fwd in one call: zero-fills both gradients, then runs the
 * rasterizer backward (cotangent grad_interpolated_features, may be NULL) and the soft-mask
 * backward (cotangent grad_soft_mask, may be NULL), which both add into the gradients.
 * Same workspace and capacities as the forward call.  Replaces the two backward halves of DibrRasterization's autograd graph
 * (kaolin/render/mesh/rasterization.py:371-388, dibr.py:56-73). */
DIBR_API int dibr_render_bwd(int batch, int num_faces, int height, int width, int feat_dim,
                    const float* grad_interpolated_features, const float* grad_soft_mask,
                    const int64_t* selected_face_idx, const float* face_vertices_image,
                    const float* face_features, const float* soft_mask, float sigmainv,
                    float box_margin, int knum, float multiplier, float eps,
                    float* grad_face_vertices_image, float* grad_face_features, const void* ws,
                    int64_t ws_bytes, int64_t raster_capacity, int64_t soft_capacity,
                    int64_t pair_capacity, void* stream);

/* ---------------------------------------------------------------------------------------
 * Next to the path: the (views x faces)-sized steps of
 * kaolin.render.mesh.prepare_vertices (kaolin/render/mesh/utils.py:128-175) and of
 * kaolin.ops.mesh.index_vertices_by_faces (kaolin/ops/mesh/mesh.py:25-47), and their
 * backward.  V-sized work (camera transform, projection) stays with the caller.
 * faces: int64 (F, 3).  Out-of-range vertex indices read as 0 / are skipped (the
 * reference's torch.gather raises a device assert there).
 * ------------------------------------------------------------------------------------- */

/* face_features (B,F,3,C)[b,f,c,:] = vertex_features[b, faces[f,c], :]; vertex_features is
 * (B,V,C), or (1,V,C) broadcast over the views when shared != 0. */
DIBR_API int dibr_index_vertices_by_faces(int batch, int num_vertices, int num_faces, int dim,
                                 int shared, const float* vertex_features, const int64_t* faces,
                                 float* face_features, void* stream);

/* Backward of the gather: grad_vertex_features[b, faces[f,c], out_col + k] +=
 * grad_face_features[b,f,c,k].  The destination has rows of out_stride floats (so several
 * gradients can share one flat (V, out_stride) buffer) and is ACCUMULATED into.  With
 * shared != 0 the B views are summed first and the destination is (V, out_stride): the
 * vertex-level message of the multi-GPU all-reduce. */
DIBR_API int dibr_scatter_face_grads(int batch, int num_vertices, int num_faces, int dim,
                            int shared, const float* grad_face_features, const int64_t* faces,
                            float* grad_vertex_features, int out_stride, int out_col,
                            void* stream);

/* prepare_vertices after the camera: vertices_camera (B,V,3), vertices_image (B,V,2) ->
 * face_vertices_camera (B,F,3,3), face_vertices_image (B,F,3,2), unit face_normals (B,F,3)
 * (cross of the edges from corner 0 divided by length + 1e-10, trianglemesh.py:328-335).
 * Any output may be NULL. */
DIBR_API int dibr_face_setup_fwd(int batch, int num_vertices, int num_faces,
                        const float* vertices_camera, const float* vertices_image,
                        const int64_t* faces, float* face_vertices_camera,
                        float* face_vertices_image, float* face_normals, void* stream);

/* Its backward: cotangents of the three outputs (any may be NULL) are ACCUMULATED into
 * grad_vertices_camera (B,V,3) and grad_vertices_image (B,V,2). */
DIBR_API int dibr_face_setup_bwd(int batch, int num_vertices, int num_faces,
                        const float* vertices_camera, const int64_t* faces,
                        const float* grad_face_vertices_camera,
                        const float* grad_face_vertices_image, const float* grad_face_normals,
                        float* grad_vertices_camera, float* grad_vertices_image, void* stream);

/* ---------------------------------------------------------------------------------------
 * Mesh-level entry points: dibr_rasterization straight from the mesh.
 * The reference builds face_vertices_camera (B,F,3,3), face_vertices_image (B,F,3,2),
 * face_normals (B,F,3) and face_features (B,F,3,D) in torch (prepare_vertices,
 * kaolin/render/mesh/utils.py:128-175; index_vertices_by_faces, kaolin/ops/mesh/mesh.py:25-47;
 * face_normals, kaolin/ops/mesh/trianglemesh.py:313-337) and hands them to
 * dibr_rasterization (dibr.py:119-209).  Here every kernel of the path fetches the corners
 * through `faces`, so those tensors never exist, and the backward adds straight into
 * vertex-level gradients.
 *   vertices_camera  (B, V, 3)   camera-space vertices: z = [..., 2], and the unit face normal
 *                                whose z >= 0 selects the rasterized faces (dibr.py:190-199)
 *   vertices_image   (B, V, 2)   NDC, not multiplied
 *   faces            (F, 3)      int32 vertex ids (faces with an id outside [0, V) are skipped)
 *   vertex_features  (B, V, D), or (1, V, D) shared by all views when features_shared != 0
 *   valid_faces      (B, F) uint8 or NULL; when given it replaces the normal test
 * Results are bit-identical to dibr_render_fwd on the face tensors dibr_face_setup_fwd /
 * dibr_index_vertices_by_faces produce from the same arrays.  Workspace protocol and pair
 * buffer as for dibr_render_fwd (same dibr_workspace_bytes_for).
 * The backward zero-fills and then accumulates grad_vertices_image (B, V, 2) and
 * grad_vertex_features (B or 1, V, D) -- with features_shared the views are summed, which
 * makes it the vertex-level message of the multi-GPU all-reduce; vertices_camera receives
 * no gradient (as face_vertices_z / face_normals_z in the reference, rasterization.py:387-388).
 * ------------------------------------------------------------------------------------- */
DIBR_API int dibr_render_mesh_fwd(int batch, int num_vertices, int num_faces, int height,
                         int width, int feat_dim, const float* vertices_camera,
                         const float* vertices_image, const int32_t* faces,
                         const float* vertex_features, int features_shared,
                         const uint8_t* valid_faces, float sigmainv, float box_margin, int knum,
                         float multiplier, float eps, int64_t* selected_face_idx,
                         float* interpolated_features, float* soft_mask, void* ws,
                         int64_t ws_bytes, int64_t raster_capacity, int64_t soft_capacity,
                         int64_t pair_capacity, int64_t* needed_host, void* stream);
DIBR_API int dibr_render_mesh_bwd(int batch, int num_vertices, int num_faces, int height,
                         int width, int feat_dim, const float* grad_interpolated_features,
                         const float* grad_soft_mask, const int64_t* selected_face_idx,
                         const float* vertices_image, const int32_t* faces,
                         const float* vertex_features, int features_shared,
                         const float* soft_mask, float sigmainv, float box_margin, int knum,
                         float multiplier, float eps, float* grad_vertices_image,
                         float* grad_vertex_features, const void* ws, int64_t ws_bytes,
                         int64_t raster_capacity, int64_t soft_capacity, int64_t pair_capacity,
                         void* stream);

/* The V-sized head of prepare_vertices (kaolin/render/mesh/utils.py:158-170):
 * rotate_translate_points (camera/legacy.py:22-37: (v - trans) @ rot^T) or the 4x3
 * camera_transform ([v, 1] @ T), then perspective_camera (legacy.py:120-139:
 * (xy * proj.xy) / (z * proj.z)).  vertices (B, V, 3), or (V, 3) used by every view when
 * vertices_shared != 0; camera_rot (B,3,3) + camera_trans (B,3), or camera_transform (B,4,3)
 * (the others NULL); camera_proj 3 floats.  Outputs vertices_camera (B,V,3), vertices_image
 * (B,V,2) -- the inputs of dibr_render_mesh_fwd.  The backward takes the cotangents of both
 * outputs (either may be NULL) and WRITES grad_vertices of the shape of `vertices` (summed
 * over the views when shared).  Camera parameters get no gradient here (they do through the
 * torch form, render.mesh.prepare_vertices). */
DIBR_API int dibr_vertex_transform_fwd(int batch, int num_vertices, int vertices_shared,
                              const float* vertices, const float* camera_rot,
                              const float* camera_trans, const float* camera_transform,
                              const float* camera_proj, float* vertices_camera,
                              float* vertices_image, void* stream);
DIBR_API int dibr_vertex_transform_bwd(int batch, int num_vertices, int vertices_shared,
                              const float* vertices_camera, const float* camera_rot,
                              const float* camera_transform, const float* camera_proj,
                              const float* grad_vertices_camera,
                              const float* grad_vertices_image, float* grad_vertices,
                              void* stream);

/* ---------------------------------------------------------------------------------------
 * Double precision: the reference dispatches float and double (rasterization_cuda.cu:30-39,
 * dibr_soft_mask_cuda.cu:211,382) and tests both.  These are its four raw ops
 * (bindings.cpp:76-81) for double, in its own layouts: packed faces + boxes + first index
 * for the rasterizer (rasterization.cpp:49-104), scaled coordinates + enlarged boxes +
 * dense per-pixel state for the soft mask (dibr_soft_mask.cpp:48-183).  `*32` arguments are
 * float copies of the same arrays, used only to bin the faces; the double arrays decide.
 * The soft-mask outputs prob / cidx / type must be pre-filled with 0 / -1 / 0.
 * ------------------------------------------------------------------------------------- */
DIBR_API int dibr_packed_rasterize_fwd_f64(
    int batch, int num_packed_faces, int height, int width, int feat_dim,
    const double* face_vertices_z, const double* face_vertices_image_scaled,
    const double* face_bboxes, const double* face_features,
    const float* face_vertices_image_scaled32, const float* face_bboxes32,
    const int64_t* first_idx_face_per_mesh, float multiplier, float eps,
    int64_t* selected_face_idx, double* interpolated_features, double* output_weights, void* ws,
    int64_t ws_bytes, int64_t raster_capacity, int64_t* needed_host, void* stream);

DIBR_API int dibr_rasterize_bwd_f64(int batch, int num_faces, int height, int width, int feat_dim,
                           const double* grad_interpolated_features,
                           const int64_t* selected_face_idx, const double* output_weights,
                           const double* face_vertices_image, const double* face_features,
                           float eps, double* grad_face_vertices_image,
                           double* grad_face_features, void* stream);

DIBR_API int dibr_softmask_fwd_f64(int batch, int num_faces, int height, int width,
                          const double* face_vertices_image_scaled,
                          const double* face_large_bboxes,
                          const float* face_vertices_image_scaled32,
                          const float* face_large_bboxes32, const int64_t* selected_face_idx,
                          float sigmainv, int knum, float multiplier, double* soft_mask,
                          double* close_face_prob, int64_t* close_face_idx,
                          uint8_t* close_face_dist_type, void* ws, int64_t ws_bytes,
                          int64_t soft_capacity, int64_t* needed_host, void* stream);

DIBR_API int dibr_softmask_bwd_dense_f64(int batch, int num_faces, int height, int width, int knum,
                                const double* grad_soft_mask, const double* soft_mask,
                                const int64_t* selected_face_idx, const double* close_face_prob,
                                const int64_t* close_face_idx,
                                const uint8_t* close_face_dist_type,
                                const double* face_vertices_image_scaled, float sigmainv,
                                float multiplier, double* grad_face_vertices_image,
                                void* stream);

/* ---------------------------------------------------------------------------------------
 * After the path: the consumer step of the DIB-R loop.
 * texture_mapping (kaolin/render/mesh/utils.py:23-76): coordinates (B,N,2) in [0,1], y up,
 * clamped; textures (B,C,Ht,Wt); out (B,N,C) channels-last; mode 0 = nearest, 1 = bilinear
 * (grid_sample, align_corners=False, padding 'border').  The backward ACCUMULATES into
 * grad_texture_maps (may be NULL) and WRITES grad_texture_coordinates (may be NULL).
 * mask_iou (kaolin/metrics/render.py:18-40): sums (B,2) = per view {sum lhs*rhs,
 * sum lhs+rhs-lhs*rhs}; loss = 1 - mean(sums[:,0] / (sums[:,1] + 1e-10)) is left to the
 * caller (B numbers); the backward takes d loss (one float on the device).
 * ------------------------------------------------------------------------------------- */
DIBR_API int dibr_texture_mapping_fwd(int batch, int64_t num_points, int channels, int tex_height,
                             int tex_width, int mode, const float* texture_coordinates,
                             const float* texture_maps, float* out, void* stream);
DIBR_API int dibr_texture_mapping_bwd(int batch, int64_t num_points, int channels, int tex_height,
                             int tex_width, int mode, const float* texture_coordinates,
                             const float* texture_maps, const float* grad_out,
                             float* grad_texture_maps, float* grad_texture_coordinates,
                             void* stream);
DIBR_API int dibr_mask_iou_sums(int batch, int64_t num_pixels, const float* lhs_mask,
                       const float* rhs_mask, float* sums, void* stream);
/* loss[0] = 1 - mean_b sums[b][0] / (sums[b][1] + 1e-10): the B-sized tail of mask_iou
 * (kaolin/metrics/render.py:37-40) on the device, so that the metric is two launches. */
DIBR_API int dibr_mask_iou_loss(int batch, const float* sums, float* loss, void* stream);
DIBR_API int dibr_mask_iou_bwd(int batch, int64_t num_pixels, const float* lhs_mask,
                      const float* rhs_mask, const float* sums, const float* grad_loss,
                      float* grad_lhs, float* grad_rhs, void* stream);

/* ---------------------------------------------------------------------------------------
 * Sibling op: deftet_sparse_render, kaolin/render/mesh/deftet.py:269-417,
 * replacing kaolin/csrc/render/mesh/deftet.h:25-44 (deftet_sparse_render_forward_cuda /
 * deftet_sparse_render_backward_cuda, bound at bindings.cpp:82-83).  dtype = DIBR_F32 or
 * DIBR_F64 selects the element type behind every `void*` tensor.
 *   pixel_coords, render_ranges (B, P, 2); face_vertices_z (B, F, 3);
 *   face_vertices_image (B, F, 3, 2); face_features (B, F, 3, D).
 * dibr_deftet_fwd does the whole forward of DeftetSparseRenderer (deftet.py:273-323): the
 * first knum faces hit per point in ascending face order, ordered by depth (descending, ties
 * by face index), face_idx (B, P, K) int64 with -1 = void, weights (B, P, K, 3) with
 * w2 = valid - (w0 + w1), interpolated_features (B, P, K, D).  Every output element is written.
 * dibr_deftet_fwd_raw is the reference's raw op (deftet.cpp:47-107): unsorted slots in
 * ascending face order, fills -1 / -inf / 0 / 0; face_bboxes (B, F, 4) as passed by the
 * caller (may be NULL = computed from the vertices).
 * dibr_deftet_bwd zero-fills and accumulates grad_face_vertices_image (B, F, 3, 2) and
 * grad_face_features (B, F, 3, D) (deftet.cpp:109-165).
 * The workspace (dibr_deftet_workspace_bytes, 256-byte aligned) holds the per-view cell grid
 * and face lists; its size is a bound, there is no capacity protocol and no host sync.
 * knum is limited by shared memory: knum * (4 + 3 * sizeof(T)) + 4 * pow2ceil(knum) + 64 * (4 + 6 * sizeof(T))
 * <= 200 KiB.
 * ------------------------------------------------------------------------------------- */
#define DIBR_F32 0
#define DIBR_F64 1
DIBR_API int64_t dibr_deftet_workspace_bytes(int batch, int num_faces, int num_points);
DIBR_API int dibr_deftet_fwd(int dtype, int batch, int num_faces, int num_points, int knum,
                    int feat_dim, const void* face_vertices_z, const void* face_vertices_image,
                    const void* face_features, const void* pixel_coords,
                    const void* render_ranges, float eps, int64_t* face_idx,
                    void* interpolated_features, void* weights, void* ws, int64_t ws_bytes,
                    void* stream);
DIBR_API int dibr_deftet_fwd_raw(int dtype, int batch, int num_faces, int num_points, int knum,
                        const void* face_vertices_z, const void* face_vertices_image,
                        const void* face_bboxes, const void* pixel_coords,
                        const void* render_ranges, float eps, int64_t* face_idx,
                        void* pixel_depths, void* w0, void* w1, void* ws, int64_t ws_bytes,
                        void* stream);
DIBR_API int dibr_deftet_bwd(int dtype, int batch, int num_faces, int num_points, int knum,
                    int feat_dim, const void* grad_interpolated_features,
                    const int64_t* face_idx, const void* weights,
                    const void* face_vertices_image, const void* face_features, float eps,
                    void* grad_face_vertices_image, void* grad_face_features, void* stream);

/* ---------------------------------------------------------------------------------------
 * The path's only exchange: SUM all-reduce of the shared mesh's vertex-level
 * gradient buffer across the GPUs of one NVSwitch box, as our own one-shot kernel over peer
 * memory (csrc/dibr_peer.cu) instead of a collective-library call.  `buffers[r]` /
 * `signal_pads[r]` (host arrays of `world` addresses) are rank r's buffer of a symmetric
 * allocation and its pad of 32-bit flags as mapped into THIS process (the Python binding
 * takes them from torch.distributed._symmetric_memory); flag rows 8 and 9 of the pads are
 * used.  Per exchange: dibr_peer_signal(mode 1) waits until every peer has finished reading
 * this rank's buffer of the previous exchange, the caller refills buffers[rank], and
 * dibr_peer_allreduce writes out[i] = sum over r of buffers[r][i] (same order on every rank:
 * bit-identical results) and tells the peers it is done reading.  dibr_peer_signal(mode 0)
 * primes the "done" flags once after set-up.  `sync` = int[2] of zeros in local device
 * memory.  Every rank must issue the same sequence of calls (as with any collective).
 * ------------------------------------------------------------------------------------- */
DIBR_API int dibr_peer_signal(int rank, int world, const int64_t* signal_pads, int mode,
                     void* stream);
DIBR_API int dibr_peer_allreduce(int rank, int world, const int64_t* buffers,
                        const int64_t* signal_pads, int64_t num_floats, float* out, int* sync,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DIBR_B200_H_ */
