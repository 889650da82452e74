// dibr_common.cuh -- shared device code of libdibr_b200 (sm_100a).
//
// Exact arithmetic.  selected_face_idx has to be bit-exact against the reference, so the
// per-(pixel, face) test is written with explicit round-to-nearest intrinsics in the
// very shapes nvcc 12.9 contracts the reference source into for sm_100a (decoded from
// the SASS of the reference build, see DESIGN.md "Numerics"):
//   rasterization_cuda.cu:85-86   x0 = fl(fl(mult / W) * float(2*ix + 1 - W))
//   rasterization_cuda.cu:137-139 w = fma(p, q, -fl(r * s))
//   rasterization_cuda.cu:140-142 norm = (w0 + w1) + w2; += copysign(eps) in double
//   rasterization_cuda.cu:143-145 three IEEE divisions
//   rasterization_cuda.cu:159     z0 = fma(w2, cz, fma(w0, az, fl(w1 * bz)))
//   rasterization_cuda.cu:187     feat = fma(w2, r2, fma(w0, r0, fl(w1 * r1)))
// Using intrinsics (never a*b+c) keeps nvcc from re-contracting them differently here.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define DIBR_FULL_MASK 0xffffffffu

namespace dibr {

// ------------------------------------------------------------------ plan / workspace
// Screen hierarchy: bin (bin_px = 32 << k, <= 1024 bins per view, the binning
// granularity) >= block (min(bin_px, 64): one CTA of the rasterizer / soft mask)
// >= tile 16x16 (one warp of the soft mask) >= region 8x4 (one pixel per lane).
struct Grid {
  float xs, ys;        // fl(mult / W), fl(mult / H)       (rasterization_cuda.cu:85-86)
  float inv_xs, inv_ys;
  int W, H;
};

__host__ __device__ inline Grid make_grid(float mult, int W, int H) {
  Grid g;
  g.W = W;
  g.H = H;
#ifdef __CUDA_ARCH__
  g.xs = __fdiv_rn(mult, (float)W);
  g.ys = __fdiv_rn(mult, (float)H);
#else
  g.xs = mult / (float)W;
  g.ys = mult / (float)H;
#endif
  g.inv_xs = 1.0f / g.xs;
  g.inv_ys = 1.0f / g.ys;
  return g;
}

struct Plan {
  int B, F, H, W;
  Grid grid;               // make_grid(grid_mult, W, H), formed once on the host (IEEE
  float grid_mult;         // divisions, bit-identical to the device's); NaN = not set
  int bin_log2, nbx, nby, NB;  // bins
  int blk_log2, nkx, nky;      // blocks (CTA footprint)
  int S, nchunks;              // faces per binning chunk (one warp), chunks per view
  int mstride;                 // row stride of the count matrix: nchunks rounded up to 4 (int4 rows)
  int lists;
  // byte offsets inside the workspace (per list kind l = 0 raster, 1 soft)
  int64_t off_work;       // int2 [B * nkx * nky * 4]  gather work list (rasterizer -> gather)
  int64_t off_mat[2];     // int [B][NB][mstride]  chunk counts -> exclusive offsets
  int64_t off_range[2];   // int2 [B][nchunks]     touched bin range of each chunk
  int64_t off_start[2];   // int [B][NB]           list segment start of each bin
  int64_t off_count[2];   // int [B][NB]           list segment length
  int64_t off_total;      // header: u64 [2] entries needed per list kind, [2] pair cursors;
                          // then int [2][16] bucket sizes of the heaviest-first order
  int64_t off_order[2];   // int [16][B * NB]    (view, bin) rows bucketed by list length
  int64_t off_flags;      // int [B][nkx*nky]      uncovered-pixel count per block
  int64_t off_rect;       // ushort4 [B][F]        exact pixel rectangle of each face's soft box
  int64_t fixed_bytes;    // everything above
  unsigned long long* trace;  // -DDIBR_TRACE builds only, else null
};

// One binning job = one list kind.  Passed by value to the kernels.
struct BinJob {
  // Mesh mode (faces != nullptr): the face corners are fetched through `faces` from the
  // per-view vertex arrays -- no (B, F, 3, .) face tensor exists; img / valid / normal_z /
  // first / bbox are null and a face is valid iff the unit normal's z computed from vcam is
  // >= 0 (mesh_cull) or always (soft list).
  const int* faces;        // (F, 3) int32 vertex ids
  const float* vimg;       // (B, V, 2) vertices_image
  const float* vcam;       // (B, V, 3) vertices_camera (z = [..., 2])
  int V;
  int mesh_cull;           // raster list: keep faces whose normal z >= 0
  const float* img;        // (B or 1, F, 3, 2)
  const float* bbox;       // explicit boxes (., F, 4) or nullptr (computed)
  const uint8_t* valid;    // (B, F) or nullptr
  const float* normal_z;   // (B, F) or nullptr (valid = nz >= 0)
  const int64_t* first;    // (B + 1) packed ranges or nullptr
  long long view_stride;   // faces between consecutive views in img/bbox (F, or 0 if packed)
  float mult;              // multiplier (applied unless prescaled)
  float margin;            // box enlargement, 0 for the raster list
  int prescaled;
  int ordered;             // list must come out in ascending face order (soft mask: yes)
  int* mat;
  int2* range;
  int* start;
  int* count;
  int* list;
  // soft list only: the EXACT pixel rectangle {x_lo, x_hi, y_lo, y_hi} of every face's
  // (enlarged) box -- the pixel centres passing the reference's half-open box test
  // (dibr_soft_mask_cuda.cu:89-97), x_lo > x_hi = none.  Written once per (view, face) by the
  // counting pass; the fill pass and the gather read it instead of redoing the geometry.
  ushort4* rect;
  // raster list as compact records instead of face ids (fused path): 3 x float4 per entry,
  // {ax, ay, bx, by} {cx, cy, za, zb} {zc, face id bits, 0, 0} with scaled coordinates;
  // written by the fill pass (which needs `z`), streamed by the rasterizer
  float4* rec;
  const float* z;          // (., F, 3), only read when rec != nullptr
  int* order;              // [16][B * NB] rows of each length bucket (filled by bin_scan)
  int* bucket;             // [16] rows per bucket
  unsigned long long* total;
  long long capacity;      // entries available in `list`
};


// ------------------------------------------------------------- kernel argument packs
struct BinJobs {
  BinJob j[2];
  // list totals reported straight into device-visible pinned host memory by the fill
  // kernel (dibr_host_report_direct), instead of a copy node between the kernels
  const unsigned long long* totals;
  unsigned long long* report;
};

// Work list of the gather kernel, filled by the rasterizer of the same call: one entry per
// quarter block (the gather's CTA footprint) with uncovered pixels whose bin has a non-empty
// soft list.  x = view, y = block x | block y << 16 (gather-block units).  `count` and the
// queue cursor `next` live in the workspace header (zeroed with it at the start of a call).
struct GatherWork {
  int2* item;
  unsigned* count;
  unsigned* next;
  unsigned cap;
};

struct RasterArgs {
  BinJob job;              // raster list + how to read faces
  const float4* rec;       // = job.rec: the list as records, or nullptr (list of face ids)
  const float* z;          // (., F, 3)
  const float* feat;       // (., F, 3, D); mesh mode: vertex features (B or 1, V, D)
  int feat_shared;         // mesh mode: one (V, D) feature array for all views
  int D;
  double eps_d;
  int64_t* out_idx;        // (B, H, W)
  float* out_feat;         // (B, H, W, D)
  float* out_w;            // (B, H, W, 3) or nullptr
  int* block_uncovered;    // (B, nkx*nky) or nullptr
  // fused forward only (else null): the rasterizer hands the soft mask its work
  float* mask;             // (B, H, W): base value of soft_mask, 1 covered / 0 uncovered
  const int* soft_count;   // soft list length per (view, bin)
  GatherWork work;         // quarter blocks that hold uncovered pixels and have a soft list
};

struct RasterBwdArgs {
  const float* grad;       // (B, H, W, D)
  const int64_t* idx;      // (B, H, W)
  const float* weights;    // (B, H, W, 3) or nullptr -> recompute
  const float* img;        // (B, F, 3, 2) unscaled
  const float* feat;       // (B, F, 3, D)
  float* grad_img;         // (B, F, 3, 2)
  float* grad_feat;        // (B, F, 3, D)
  // mesh mode (faces != nullptr): img = vertices_image (B, V, 2), feat = vertex features
  // (B or 1, V, D), grad_img / grad_feat the vertex gradients of the same shapes
  const int* faces;        // (F, 3) int32 or nullptr
  int V;
  int feat_shared;
  int B, F, H, W, D;
  float mult;
  float eps_f;
  double eps_d;
  double inv_mult;         // 1 / multiplier
  Grid grid;               // make_grid(mult, W, H): pixel centres as the forward forms them
};

// Compact per-pixel face lists of the soft mask ("pairs"), kept between forward and
// backward.  A region (8x4 pixels, one per lane) that holds at least one hit becomes
// "active" and owns segments of 8 hits x 32 lanes: slot (hit h of lane l) lives at
// seg[h / 8] * 256 + (h % 8) * 32 + l.  Everything is sized by `cap` segments.
constexpr int kPairSegHits = 8;
constexpr int kPairSegSlots = kPairSegHits * 32;
constexpr int kPairMaxK = 1024;  // above this knum the pair path is not used
struct PairBuf {
  int cap;                    // segments (and active regions) available; 0 = pair path off
  int nseg;                   // segments per region = ceil(knum / 8)
  unsigned long long* cursor; // [0] segments requested, [1] active regions requested
  int* act_region;            // [cap]        global region id
  unsigned short* act_cnt;    // [cap][32]    hits held by each pixel of the region
  int* act_seg;               // [cap][nseg]  segment ids
  int* face;                  // [cap * 256]
  float* prob;                // [cap * 256]
  unsigned char* type;        // [cap * 256]  1-3 edge, 4-6 vertex (reference numbering)
  // work items of the evaluation / gradient kernels: a region's pairs in chunks of
  // kPairChunk, so that the silhouette's few thousand heavy regions spread over all SMs
  unsigned int* nchunk;       // chunks written (header)
  int* chunk;                 // [3 * cap]    act * 256 + chunk index within the region
};
constexpr int kPairChunk = 128;
// list capacities stay this far below 2^31 so that the fill kernel's int slot counters
// (segment start + up to S increments per chunk) cannot wrap
constexpr long long kMaxListEntries = 0x7fffffffll - 65536;
constexpr int kRecBytes = 48;   // one raster-list record (BinJob::rec)

__device__ __forceinline__ bool pairs_overflowed(const PairBuf& pb) {
  return pb.cursor[0] > (unsigned long long)pb.cap || pb.cursor[1] > (unsigned long long)pb.cap;
}

struct SoftArgs {
  BinJob job;              // soft list + how to read faces (mult, margin, prescaled, boxes)
  const int64_t* sel;      // (B, H, W)
  float sigmainv;
  int K;
  float* mask;             // (B, H, W) forward output
  float* prob;             // dense (B, H, W, K) or nullptr
  int64_t* cidx;
  uint8_t* type;
  const float* grad_mask;  // backward inputs
  const float* mask_in;
  float* grad_img;         // (B, F, 3, 2) accumulated, unscaled coordinates
  PairBuf pairs;
  unsigned long long* report;  // device-visible host slot for the pair-buffer needs, or null
  const int* block_uncovered;  // (B, nkx*nky) uncovered pixels per raster block (written by
                               // the rasterizer of the same call), or null
  GatherWork work;             // item != null: persistent gather over the rasterizer's work
                               // list; soft_mask's base value is already written
};

// dense-state backward of the raw _C op, float and double
template <typename T>
struct SoftDenseBwdArgsT {
  const T* grad_mask;
  const T* mask;
  const int64_t* sel;
  const T* prob;
  const int64_t* cidx;
  const uint8_t* type;
  const T* img;  // (B, F, 3, 2) scaled
  T* grad_img;
  int B, F, H, W, K;
  float sigmainv, mult;
};
using SoftDenseBwdArgs = SoftDenseBwdArgsT<float>;
using SoftBwdF64Args = SoftDenseBwdArgsT<double>;

// ---- double-precision instantiation (dibr_f64.cu): the reference's raw-op layouts --
// packed faces + boxes + first index for the rasterizer, scaled coordinates + enlarged boxes
// + dense per-pixel state for the soft mask.  `job` is a float binning job over float copies.
struct RasterF64Args {
  BinJob job;
  const double *z, *img, *bbox, *feat;   // (Fv,3) (Fv,3,2) scaled (Fv,4) (Fv,3,D)
  double eps_d;
  int D;
  int64_t* out_idx;
  double *out_feat, *out_w;
};
struct RasterBwdF64Args {
  const double *grad, *weights, *img, *feat;
  const int64_t* idx;
  double *grad_img, *grad_feat;
  int B, F, H, W, D;
  double eps_d;
};
struct SoftF64Args {
  BinJob job;
  const double *img, *bbox;              // (B,F,3,2) scaled, (B,F,4) enlarged
  const int64_t* sel;
  float sigmainv;
  int K;
  double *mask, *prob;
  int64_t* cidx;
  uint8_t* type;
};

// SOFT_FWD / SOFT_BWD: self-contained kernels (gather + evaluate per block); when a pair
// buffer is attached they only run if it overflowed.  SOFT_GATHER: gather only, fills the
// pair buffer for launch_softmask_pairs_{eval,grad}.
enum { SOFT_FWD = 0, SOFT_FWD_DENSE = 1, SOFT_BWD = 2, SOFT_GATHER = 3 };

// Optional per-kernel timing (dibr_profile_begin/end in the C ABI): CUDA events recorded on
// the launching stream right before and after a kernel.  No-ops unless enabled.
enum { K_BIN_COUNT = 0, K_BIN_SCAN, K_BIN_FILL, K_RASTER_FWD, K_SOFT_FWD, K_RASTER_BWD,
       K_SOFT_BWD, K_SOFT_BWD_DENSE, K_SOFT_GATHER, K_SOFT_EVAL, K_SOFT_GRAD, K_SOFT_FOLD, K_RASTER_BWD_FINISH, K_NUM };
void prof_begin(int kernel_id, cudaStream_t st);
void prof_end(cudaStream_t st);

cudaError_t launch_binning_count(const BinJobs&, int njobs, const Plan&, cudaStream_t);
cudaError_t launch_binning_fill(const BinJobs&, int njobs, const Plan&, cudaStream_t);
cudaError_t launch_raster_forward(const RasterArgs&, const Plan&, cudaStream_t);
cudaError_t launch_raster_backward(const RasterBwdArgs&, cudaStream_t);
cudaError_t launch_softmask(const SoftArgs&, int mode, const Plan&, cudaStream_t);
cudaError_t launch_softmask_dense_backward(const SoftDenseBwdArgs&, cudaStream_t);
cudaError_t launch_softmask_gather(const SoftArgs&, const Plan&, cudaStream_t);
cudaError_t launch_softmask_pairs_eval(const SoftArgs&, const Plan&, cudaStream_t);
cudaError_t launch_softmask_pairs_grad(const SoftArgs&, const Plan&, cudaStream_t);
cudaError_t launch_raster_forward_f64(const RasterF64Args&, const Plan&, cudaStream_t);
cudaError_t launch_raster_backward_f64(const RasterBwdF64Args&, cudaStream_t);
cudaError_t launch_softmask_forward_f64(const SoftF64Args&, const Plan&, cudaStream_t);
cudaError_t launch_softmask_backward_f64(const SoftBwdF64Args&, cudaStream_t);

// ------------------------------------------------- heaviest-first block scheduling
// bin_scan buckets the (view, bin) rows by floor(log2(list length)) + 1.  Consumers whose
// CTA cost grows with the list length take their block through this order, longest lists
// first, so the long CTAs start early and the short ones fill the tail (LPT scheduling).
constexpr int kOrderBuckets = 16;
// Warp-cooperative (every lane of a converged warp calls it with the same i): lane k holds
// bucket k's size, a suffix sum over the 16 buckets finds the bucket that position i of the
// heaviest-first order falls into.  ~20 instructions per warp; the per-thread version (16
// loads and 64-bit compares in every thread) was a third of the gather kernel's instructions.
// nrows = B * NB < 2^27.
__device__ __forceinline__ int lpt_row(const BinJob& j, int nrows, int i) {
  const int lane = threadIdx.x & 31;
  const int c = lane < kOrderBuckets ? j.bucket[lane] : 0;
  int incl = c;                       // -> sum of the sizes of buckets >= lane
#pragma unroll
  for (int d = 1; d < kOrderBuckets; d <<= 1) {
    const int t = __shfl_down_sync(DIBR_FULL_MASK, incl, d);
    if (lane + d < 32) incl += t;
  }
  const int before = incl - c;        // rows in heavier buckets
  const unsigned hit = __ballot_sync(DIBR_FULL_MASK, c > 0 && i >= before && i < incl);
  if (!hit) return -1;
  const int k = __ffs(hit) - 1;
  const int acc = __shfl_sync(DIBR_FULL_MASK, before, k);
  return j.order[(long long)k * nrows + (i - acc)];
}

// CTA `id` of a 1-D grid of B * NB * 4^(bin_log2 - sub_log2) CTAs -> (view, block x, block y)
// in units of 2^sub_log2 pixels.  Returns false for blocks outside the image.  Called by all
// threads of the CTA at kernel start (see lpt_row).
__device__ __forceinline__ bool lpt_block(const BinJob& j, const Plan& p, int sub_log2,
                                          unsigned id, int& b, int& bxi, int& byi) {
  const int shift = p.bin_log2 - sub_log2;
  const int row = lpt_row(j, p.B * p.NB, (int)(id >> (2 * shift)));
  if (row < 0) return false;
  const int sub = (int)(id & ((1u << (2 * shift)) - 1u));
  b = row / p.NB;
  const int bin = row - b * p.NB;
  const int by = bin / p.nbx;
  bxi = ((bin - by * p.nbx) << shift) + (sub & ((1 << shift) - 1));
  byi = (by << shift) + (sub >> shift);
  return (bxi << sub_log2) < p.W && (byi << sub_log2) < p.H;
}

// ------------------------------------------------------------------------ pixel grid
// the pixel grid of a kernel: the plan's if it was formed for this multiplier
__device__ __forceinline__ Grid plan_grid(const Plan& p, float mult) {
  if (p.grid_mult == mult) return p.grid;
  return make_grid(mult, p.W, p.H);
}

__device__ __forceinline__ float pix_x(const Grid& g, int ix) {
  return __fmul_rn(g.xs, (float)(2 * ix + 1 - g.W));
}
__device__ __forceinline__ float pix_y(const Grid& g, int iy) {
  return __fmul_rn(g.ys, (float)(g.H - 2 * iy - 1));
}

// Conservative pixel range [x_lo, x_hi] x [y_lo, y_hi] (clamped to the image) of the
// pixel centres that can satisfy  xmin <= x0 < xmax && ymin <= y0 < ymax  (the
// reference's half-open test, rasterization_cuda.cu:115).  In real arithmetic
// x0(ix) >= xmin  <=>  ix >= fx_lo = (xmin / xs + W - 1) / 2   and   x0(ix) < xmax  <=>
// ix < fx_hi; the float evaluation of fx is off by < W * 4e-7 pixels for on-screen boxes
// (three roundings of magnitudes <= 2W), so a slack of W * 1e-6 + 1e-3 pixels makes the
// range a superset.  The exact float test is redone per pixel by the consumers.
// Returns false when the range is empty.
__device__ __forceinline__ bool pixel_range(const Grid& g, float xmin, float ymin, float xmax,
                                            float ymax, int& x_lo, int& x_hi, int& y_lo,
                                            int& y_hi) {
  if (!(xmin < xmax) || !(ymin < ymax)) return false;  // also rejects NaN boxes
  const float Wf = (float)(g.W - 1), Hf = (float)(g.H - 1);
  const float dx = (float)g.W * 1e-6f + 1e-3f, dy = (float)g.H * 1e-6f + 1e-3f;
  float fx_lo = (xmin * g.inv_xs + Wf) * 0.5f;
  float fx_hi = (xmax * g.inv_xs + Wf) * 0.5f;
  float fy_lo = (Hf - ymax * g.inv_ys) * 0.5f;  // y grows downwards in pixel rows
  float fy_hi = (Hf - ymin * g.inv_ys) * 0.5f;
  fx_lo = fminf(fmaxf(fx_lo, -4.f), Wf + 4.f);
  fx_hi = fminf(fmaxf(fx_hi, -4.f), Wf + 4.f);
  fy_lo = fminf(fmaxf(fy_lo, -4.f), Hf + 4.f);
  fy_hi = fminf(fmaxf(fy_hi, -4.f), Hf + 4.f);
  // x: xmin <= x0 < xmax  ->  ix in [ceil(fx_lo), ceil(fx_hi) - 1]
  x_lo = max((int)ceilf(fx_lo - dx), 0);
  x_hi = min((int)ceilf(fx_hi + dx) - 1, g.W - 1);
  // y: y0 = ys * (H - 2 iy - 1) decreases with iy:  y0 < ymax <=> iy > fy_lo,
  //    y0 >= ymin <=> iy <= fy_hi  ->  iy in [floor(fy_lo) + 1, floor(fy_hi)]
  y_lo = max((int)floorf(fy_lo - dy) + 1, 0);
  y_hi = min((int)floorf(fy_hi + dy), g.H - 1);
  return x_lo <= x_hi && y_lo <= y_hi;
}

// The exact rectangle: the conservative range trimmed with the reference's own float
// comparisons (pix_x / pix_y are monotone in ix / iy).  Returns false when empty.
__device__ __forceinline__ bool exact_pixel_rect(const Grid& g, float xmin, float ymin, float xmax,
                                                 float ymax, int& x_lo, int& x_hi, int& y_lo,
                                                 int& y_hi) {
  if (!pixel_range(g, xmin, ymin, xmax, ymax, x_lo, x_hi, y_lo, y_hi)) return false;
  while (x_lo <= x_hi && pix_x(g, x_lo) < xmin) x_lo++;
  while (x_lo <= x_hi && pix_x(g, x_hi) >= xmax) x_hi--;
  while (y_lo <= y_hi && pix_y(g, y_lo) >= ymax) y_lo++;
  while (y_lo <= y_hi && pix_y(g, y_hi) < ymin) y_hi--;
  return x_lo <= x_hi && y_lo <= y_hi;
}

// ----------------------------------------------------------------------- face records
struct Face2D {
  float ax, ay, bx, by, cx, cy;     // scaled image coordinates
  float xmin, ymin, xmax, ymax;     // (possibly enlarged) box
};

// Loads face `f` of view `b` for a binning job / consumer: scales (fl(v * mult) as torch
// does at rasterization.py:337 / dibr.py:32) and builds the box (rasterization.py:342-344,
// dibr.py:34-39) unless explicit boxes were given.  Returns false for faces that are not
// part of the list (invalid, outside the packed range, or non-finite coordinates).
// LISTED: `f` comes out of the job's own face list, which binning only fills with faces
// that passed these checks -- consumers skip the (dependent) validity loads.
// z component of the unit face normal as prepare_vertices forms it (face_normals(unit=True),
// trianglemesh.py:328-335: cross of the edges from corner 0, divided by length + 1e-10);
// ONE definition for the face-setup kernel (dibr_mesh.cu) and the mesh-mode validity test,
// with explicit roundings so that both give the same bits
__device__ __forceinline__ float3 unit_face_normal(const float* p0, const float* p1,
                                                   const float* p2) {
  const float e0x = __fsub_rn(p1[0], p0[0]), e0y = __fsub_rn(p1[1], p0[1]), e0z = __fsub_rn(p1[2], p0[2]);
  const float e1x = __fsub_rn(p2[0], p0[0]), e1y = __fsub_rn(p2[1], p0[1]), e1z = __fsub_rn(p2[2], p0[2]);
  const float nx = __fmaf_rn(e0y, e1z, -__fmul_rn(e0z, e1y));
  const float ny = __fmaf_rn(e0z, e1x, -__fmul_rn(e0x, e1z));
  const float nz = __fmaf_rn(e0x, e1y, -__fmul_rn(e0y, e1x));
  const float len = __fadd_rn(__fsqrt_rn(__fmaf_rn(nz, nz, __fmaf_rn(ny, ny, __fmul_rn(nx, nx)))),
                              1e-10f);
  return make_float3(__fdiv_rn(nx, len), __fdiv_rn(ny, len), __fdiv_rn(nz, len));
}

// the three vertex ids of face f (mesh mode); false if any is out of range
__device__ __forceinline__ bool mesh_face(const BinJob& j, int f, int (&v)[3]) {
  const int* fp = j.faces + (long long)f * 3;
  v[0] = __ldg(fp); v[1] = __ldg(fp + 1); v[2] = __ldg(fp + 2);
  return (unsigned)v[0] < (unsigned)j.V && (unsigned)v[1] < (unsigned)j.V &&
         (unsigned)v[2] < (unsigned)j.V;
}

// unscaled image coordinates of the three corners of face f of view b, either layout
__device__ __forceinline__ bool face_corners(const BinJob& j, int b, int f, float2& a, float2& bb,
                                             float2& c) {
  if (j.faces) {
    int v[3];
    if (!mesh_face(j, f, v)) return false;
    const float2* vp = reinterpret_cast<const float2*>(j.vimg) + (long long)b * j.V;
    a = __ldg(vp + v[0]); bb = __ldg(vp + v[1]); c = __ldg(vp + v[2]);
    return true;
  }
  const float2* p = reinterpret_cast<const float2*>(j.img + ((long long)b * j.view_stride + f) * 6);
  a = __ldg(p); bb = __ldg(p + 1); c = __ldg(p + 2);
  return true;
}

// depths of the three corners (rasterizer), either layout
__device__ __forceinline__ void face_depths(const BinJob& j, const float* z, int b, int f,
                                            float& za, float& zb, float& zc) {
  if (j.faces) {
    int v[3];
    za = zb = zc = 0.f;
    if (!mesh_face(j, f, v)) return;
    const float* vp = j.vcam + (long long)b * j.V * 3;
    za = __ldg(vp + v[0] * 3 + 2); zb = __ldg(vp + v[1] * 3 + 2); zc = __ldg(vp + v[2] * 3 + 2);
    return;
  }
  const float* zp = z + ((long long)b * j.view_stride + f) * 3;
  za = __ldg(zp); zb = __ldg(zp + 1); zc = __ldg(zp + 2);
}

template <bool LISTED = false>
__device__ __forceinline__ bool load_face2d(const BinJob& j, int b, int f, Face2D& o) {
  const long long gi = (long long)b * j.view_stride + f;
  if (!LISTED) {
    if (j.first) {
      if (f < (int)j.first[b] || f >= (int)j.first[b + 1]) return false;
    }
    if (j.valid && !j.valid[gi]) return false;
    if (j.normal_z && !(j.normal_z[gi] >= 0.f)) return false;
    if (j.faces && j.mesh_cull) {
      int v[3];
      if (!mesh_face(j, f, v)) return false;
      const float* vp = j.vcam + (long long)b * j.V * 3;
      if (!(unit_face_normal(vp + v[0] * 3, vp + v[1] * 3, vp + v[2] * 3).z >= 0.f)) return false;
    }
  }
  float2 a, bb, c;
  if (!face_corners(j, b, f, a, bb, c)) return false;
  if (!j.prescaled) {
    a.x = __fmul_rn(a.x, j.mult); a.y = __fmul_rn(a.y, j.mult);
    bb.x = __fmul_rn(bb.x, j.mult); bb.y = __fmul_rn(bb.y, j.mult);
    c.x = __fmul_rn(c.x, j.mult); c.y = __fmul_rn(c.y, j.mult);
  }
  o.ax = a.x; o.ay = a.y; o.bx = bb.x; o.by = bb.y; o.cx = c.x; o.cy = c.y;
  if (j.bbox) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(j.bbox + gi * 4));
    o.xmin = q.x; o.ymin = q.y; o.xmax = q.z; o.ymax = q.w;
  } else {
    o.xmin = __fsub_rn(fminf(fminf(a.x, bb.x), c.x), j.margin);
    o.ymin = __fsub_rn(fminf(fminf(a.y, bb.y), c.y), j.margin);
    o.xmax = __fadd_rn(fmaxf(fmaxf(a.x, bb.x), c.x), j.margin);
    o.ymax = __fadd_rn(fmaxf(fmaxf(a.y, bb.y), c.y), j.margin);
  }
  const float s = a.x + a.y + bb.x + bb.y + c.x + c.y;
  return isfinite(s);
}

// -------------------------------------------------- exact rasterizer test (see header)
// Returns true when pixel (x0, y0) is inside the face by the reference's rule
// (none of the three normalised weights < 0, rasterization_cuda.cu:148).
__device__ __forceinline__ bool raster_weights(float x0, float y0, float ax, float ay, float bx,
                                               float by, float cx, float cy, double eps_d,
                                               float& w0, float& w1, float& w2) {
  const float a_ex = __fsub_rn(ax, x0), a_ey = __fsub_rn(ay, y0);
  const float b_ex = __fsub_rn(bx, x0), b_ey = __fsub_rn(by, y0);
  const float c_ex = __fsub_rn(cx, x0), c_ey = __fsub_rn(cy, y0);
  w0 = __fmaf_rn(b_ex, c_ey, -__fmul_rn(b_ey, c_ex));
  w1 = __fmaf_rn(c_ex, a_ey, -__fmul_rn(c_ey, a_ex));
  w2 = __fmaf_rn(a_ex, b_ey, -__fmul_rn(a_ey, b_ex));
  float norm = __fadd_rn(__fadd_rn(w0, w1), w2);
  norm = __double2float_rn(__dadd_rn((double)norm, copysign(eps_d, (double)norm)));
  w0 = __fdiv_rn(w0, norm);
  w1 = __fdiv_rn(w1, norm);
  w2 = __fdiv_rn(w2, norm);
  return !(w0 < 0.f || w1 < 0.f || w2 < 0.f);
}

__device__ __forceinline__ float raster_depth(float w0, float w1, float w2, float az, float bz,
                                              float cz) {
  return __fmaf_rn(w2, cz, __fmaf_rn(w0, az, __fmul_rn(w1, bz)));
}

// Depth key: larger wins.  High word = order-preserving map of z0 (after -0 -> +0), low
// word = ~face index, so that on equal depth the LOWEST face index wins exactly as the
// reference's ascending scan with a strict '>' does (rasterization_cuda.cu:162).
// z0 = -inf can never beat the initial max_z0 = -inf there, and NaN depths are not
// supported: both yield key 0 (= empty).
__device__ __forceinline__ unsigned long long depth_key(float z0, int face) {
  if (!(z0 > -INFINITY)) return 0ull;
  z0 = __fadd_rn(z0, 0.0f);  // -0.0 -> +0.0
  const unsigned u = __float_as_uint(z0);
  const unsigned ord = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)ord << 32) | (unsigned)(~(unsigned)face);
}
__device__ __forceinline__ int key_face(unsigned long long key) {
  return (int)(~(unsigned)(key & 0xffffffffull));
}

__device__ __forceinline__ void smem_max64(unsigned long long* addr, unsigned long long key) {
  unsigned long long old = *addr;
  while (key > old) {
    const unsigned long long assumed = old;
    old = atomicCAS(addr, assumed, key);
    if (old == assumed) break;
  }
}

// ----------------------------------------------------------------------- CTA tracing
// Development aid, compiled in only with -DDIBR_TRACE: every CTA of an instrumented kernel
// records {start ns, end ns, SM id, user value} at plan.trace[4 * linear block id]
// (a kernel argument, set through dibr_trace_set; `p` must be the kernel's Plan).
#ifdef DIBR_TRACE
__device__ __forceinline__ unsigned long long trace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define DIBR_TRACE_BEGIN() const unsigned long long trace_t0_ = trace_now()
#define DIBR_TRACE_END_AT(slot, user)                                                      \
  do {                                                                                     \
    __syncthreads();                                                                       \
    if (p.trace && threadIdx.x == 0) {                                                        \
      unsigned smid_;                                                                      \
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid_));                                   \
      unsigned long long* o_ = p.trace + 4ull * (slot);                                    \
      o_[0] = trace_t0_; o_[1] = trace_now(); o_[2] = smid_; o_[3] = (unsigned long long)(user); \
    }                                                                                      \
  } while (0)
#define DIBR_TRACE_END(user) \
  DIBR_TRACE_END_AT((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x, user)
#else
#define DIBR_TRACE_BEGIN()
#define DIBR_TRACE_END(user)
#define DIBR_TRACE_END_AT(slot, user)
#endif

__device__ __forceinline__ void prefetch_l1(const void* p) {
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}

// ------------------------------------------------------- programmatic dependent launch
// Every kernel of the path starts with pdl_prologue(), which waits until the PREVIOUS kernel
// of the stream has completed and flushed.  Without the launch attribute (launch_k, off by
// default) it is a no-op.  Measured on the cfg-2 step inside a CUDA graph (B200): with the
// attribute and an early griddepcontrol.launch_dependents the step got SLOWER (0.314 ->
// 0.335 ms: the next kernel's parked CTAs take SM slots from the running one), with the
// attribute and no early trigger it was unchanged (0.315 ms) -- graph replay already leaves
// no launch gap worth hiding.  Kept as an experiment switch: DIBR_B200_PDL=1, and
// -DDIBR_PDL_TRIGGER=1 at build time for the early trigger.
#ifndef DIBR_PDL_TRIGGER
#define DIBR_PDL_TRIGGER 0
#endif
__device__ __forceinline__ void pdl_prologue() {
#if DIBR_PDL_TRIGGER
  asm volatile("griddepcontrol.launch_dependents;");
#endif
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

#ifdef __CUDACC__
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel: opted
// in once per (kernel, device) -- a process may render on several GPUs (dibr_api.cu)
cudaError_t smem_optin(const void* kernel, int bytes);
bool pdl_enabled();   // dibr_api.cu: DIBR_B200_PDL=0 turns the attribute off

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                            cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ------------------------------------------------------------------------------ misc
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

}  // namespace dibr
