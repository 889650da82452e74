// dibr_deftet.cu -- deftet_sparse_render: the k-layer sibling of the
// rasterizer for ARBITRARY sample points instead of a pixel grid.
//   kaolin/render/mesh/deftet.py:269-417          DeftetSparseRenderer (torch side)
//   kaolin/csrc/render/mesh/deftet_cuda.cu:31-193 forward kernel, :239-401 backward kernel
//
// Semantics kept from the reference: per sample point the FIRST knum faces in ascending
// face index that (a) contain the point in their half-open bounding box, (b) have all three
// normalised edge functions >= 0 and (c) whose interpolated depth lies in
// [range_min, range_max) are kept (deftet_cuda.cu:120-168), then ordered by depth,
// nearest (largest z) first (deftet.py:303-306), w2 := valid - (w0 + w1) (:307) and the
// features interpolated (:308-317).  Arithmetic of the test in the shapes nvcc contracts
// the reference into for sm_100a (read from cuobjdump -sass of the reference build):
//   _w0 = fma(bex, cey, -fl(bey * cex)), _w1 = fma(cex, aey, -fl(cey * aex)),
//   _w2 = fma(aex, bey, -fl(aey * bex)), norm = (_w0 + _w1) + _w2,
//   den = norm + copysign(eps_float, norm), w_k = _w_k / den (IEEE),
//   depth = fma(w2, cz, fma(w0, az, fl(w1 * bz))).
//
// B200 design.  The reference tests every face against every point (one warp per point,
// F / 32 steps).  Here the faces of a view are dealt into a uniform G x G grid laid over the
// bounding box of the view's sample points; faces spanning more than kWideCells cells go to
// one per-view "wide" list instead.  A point only tests its cell's list plus the wide list.
// G is chosen ON THE DEVICE (no host round trip): one pass counts, for every G = 1, 2, 4 ..
// 512 at once, the list entries E_G and wide faces W_G the mesh would produce, and a
// one-thread-per-view kernel takes the G minimising
//     E_G * (c_build + c_test * P / G^2) + c_test * P * W_G
// (E_G / G^2 is exactly the expected list length under a uniformly placed sample).
// Lists are filled with atomics (unordered): the kept set is order independent (all hits,
// or the knum smallest face ids among them) and so is the final order (depth descending,
// ties by face id = the reference's insertion order under a stable sort).
// float and double are both instantiated (the reference dispatches both).
#include <algorithm>

#include "../../include/dibr_b200.h"
#include "dibr_common.cuh"

namespace dibr {

int fail(int code, const char* fmt, ...);

namespace {

constexpr int kLevels = 10;       // G = 1 << level, level 0..9
constexpr int kWideCells = 64;    // a face spanning more cells goes to the wide list
constexpr int kListCap = 16;      // list entries available per face and view (bounds the workspace)
constexpr float kCostBuild = 30.f, kCostTest = 40.f;   // thread-instructions per entry / test

// ------------------------------------------------------------------ exact arithmetic
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fma_rn(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float to_f(float a) { return a; }
__device__ __forceinline__ float to_f(double a) { return __double2float_rn(a); }  // monotone
__device__ __forceinline__ bool sign_of(float a) { return (__float_as_uint(a) >> 31) != 0; }
__device__ __forceinline__ bool sign_of(double a) { return (__double2hiint(a) >> 31) != 0; }
// min / max that drop a NaN operand: a box computed with them is a superset of the
// reference's torch.min / torch.max box, and a face with a NaN coordinate fails the
// edge test on its own (every weight becomes NaN), so results do not change.
__device__ __forceinline__ float min3(float a, float b, float c) { return fminf(fminf(a, b), c); }
__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
__device__ __forceinline__ double min3(double a, double b, double c) { return fmin(fmin(a, b), c); }
__device__ __forceinline__ double max3(double a, double b, double c) { return fmax(fmax(a, b), c); }

template <typename T> struct Vec2;
template <> struct Vec2<float> { typedef float2 type; };
template <> struct Vec2<double> { typedef double2 type; };

// float <-> unsigned with the order of the floats; every non-NaN value encodes to >= 1
__device__ __forceinline__ unsigned enc_f(float f) {
  const unsigned u = __float_as_uint(f);
  return (u >> 31) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float dec_f(unsigned e) {
  return __uint_as_float((e >> 31) ? (e & 0x7fffffffu) : ~e);
}

// per-view statistics (zero-initialised by a memset)
struct ViewStats {
  unsigned nlo_x, nlo_y;          // ~enc(min sample coordinate)  (atomicMax)
  unsigned hi_x, hi_y;            // enc(max sample coordinate)   (atomicMax); 0 = no sample
  unsigned wide;                  // entries of the wide list (fill cursor)
  unsigned pad[3];
  unsigned long long entries[kLevels];  // list entries a G = 1 << l grid would hold
  unsigned wides[kLevels];              // faces that grid would send to the wide list
  unsigned pad2[2];
};

// the chosen grid of a view (written by deftet_choose_kernel, read by everything after)
struct CellGrid {
  float lox, loy, hix, hiy, invx, invy;
  int g;          // cells per axis
  int empty;      // no sample point (or none that can be hit)
};

// sample domain of a view and 1 / extent (0 where the extent is 0 or not finite)
__device__ __forceinline__ CellGrid domain_of(const ViewStats& s) {
  CellGrid g;
  g.empty = (s.hi_x == 0u);
  g.g = 1;
  g.invx = g.invy = 0.f;
  g.lox = g.loy = g.hix = g.hiy = 0.f;
  if (g.empty) return g;
  g.lox = dec_f(~s.nlo_x); g.loy = dec_f(~s.nlo_y);
  g.hix = dec_f(s.hi_x); g.hiy = dec_f(s.hi_y);
  const float dx = g.hix - g.lox, dy = g.hiy - g.loy;
  g.invx = (dx > 0.f && dx < 3.0e38f) ? 1.f / dx : 0.f;
  g.invy = (dy > 0.f && dy < 3.0e38f) ? 1.f / dy : 0.f;
  if (!(g.invx < 1.0e30f)) g.invx = 0.f;
  if (!(g.invy < 1.0e30f)) g.invy = 0.f;
  return g;
}

// monotone non-decreasing in v (sub, mul by >= 0, truncation and clamp all are), so a
// point inside [lo, hi) of a box falls in a cell between cell(lo) and cell(hi).
// inv = (1 / extent) * G with G a power of two: exact, the same value wherever formed.
__device__ __forceinline__ int cell_of(float v, float lo, float inv, int n) {
  const float t = (v - lo) * inv;
  return (t >= 0.f) ? ((t < (float)n) ? (int)t : n - 1) : 0;
}

template <typename T>
struct DeftetArgs {
  const T* z;          // (B, F, 3)
  const T* img;        // (B, F, 3, 2)
  const T* bbox;       // (B, F, 4) or nullptr (computed from img)
  const T* feat;       // (B, F, 3, D)
  const T* pix;        // (B, P, 2)
  const T* range;      // (B, P, 2)
  int64_t* out_idx;    // (B, P, K)
  T* out_feat;         // (B, P, K, D)        sorted mode
  T* out_w;            // (B, P, K, 3)        sorted mode
  T* out_depth;        // (B, P, K)           raw mode
  T* out_w0;           // (B, P, K)           raw mode
  T* out_w1;           // (B, P, K)           raw mode
  ViewStats* stats;    // [B]
  CellGrid* grids;     // [B]
  int* cell_start;     // [B][cells_cap]
  int* cell_count;     // [B][cells_cap]
  int* wide;           // [B][F]
  int* list;           // [B][kListCap * F]
  int B, F, P, K, D, cells_cap, n2;
  float eps;
};

template <typename T>
__device__ __forceinline__ void load_img6(const T* img, long long bf, T (&v)[6]) {
  typedef typename Vec2<T>::type V2;
  const V2* p = reinterpret_cast<const V2*>(img) + bf * 3;
  const V2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y;
}

// float image of a face's box for the binning; false if it cannot contain any sample
template <typename T>
__device__ __forceinline__ bool face_box_f(const DeftetArgs<T>& a, const CellGrid& g, int b,
                                           int f, float& fx0, float& fy0, float& fx1,
                                           float& fy1) {
  const long long bf = (long long)b * a.F + f;
  if (a.bbox) {
    const T* q = a.bbox + bf * 4;
    fx0 = to_f(__ldg(q)); fy0 = to_f(__ldg(q + 1));
    fx1 = to_f(__ldg(q + 2)); fy1 = to_f(__ldg(q + 3));
  } else {
    T v[6];
    load_img6(a.img, bf, v);
    fx0 = to_f(min3(v[0], v[2], v[4])); fx1 = to_f(max3(v[0], v[2], v[4]));
    fy0 = to_f(min3(v[1], v[3], v[5])); fy1 = to_f(max3(v[1], v[3], v[5]));
  }
  // a hit needs xmin <= x < xmax with lox <= x <= hix (to_f is monotone, so the float
  // images keep the non-strict forms of these inequalities); NaN fails
  return fx1 >= g.lox && fx0 <= g.hix && fy1 >= g.loy && fy0 <= g.hiy;
}

// ------------------------------------------------------------------ grid construction
template <typename T>
__global__ void __launch_bounds__(256) deftet_bounds_kernel(const DeftetArgs<T> a) {
  typedef typename Vec2<T>::type V2;
  const int b = blockIdx.y;
  unsigned nlx = 0, nly = 0, hx = 0, hy = 0;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < a.P; p += gridDim.x * blockDim.x) {
    const V2 c = __ldg(reinterpret_cast<const V2*>(a.pix) + (long long)b * a.P + p);
    if (c.x != c.x || c.y != c.y) continue;   // such a sample hits nothing
    const unsigned ex = enc_f(to_f(c.x)), ey = enc_f(to_f(c.y));
    nlx = max(nlx, ~ex); nly = max(nly, ~ey); hx = max(hx, ex); hy = max(hy, ey);
  }
  nlx = __reduce_max_sync(DIBR_FULL_MASK, nlx); nly = __reduce_max_sync(DIBR_FULL_MASK, nly);
  hx = __reduce_max_sync(DIBR_FULL_MASK, hx); hy = __reduce_max_sync(DIBR_FULL_MASK, hy);
  if ((threadIdx.x & 31) == 0 && hx != 0u) {
    ViewStats* s = a.stats + b;
    atomicMax(&s->nlo_x, nlx); atomicMax(&s->nlo_y, nly);
    atomicMax(&s->hi_x, hx); atomicMax(&s->hi_y, hy);
  }
}

// what every grid resolution would cost: entries and wide faces per level
template <typename T>
__global__ void __launch_bounds__(256) deftet_levels_kernel(const DeftetArgs<T> a) {
  const int b = blockIdx.y;
  const CellGrid g = domain_of(a.stats[b]);
  if (g.empty) return;
  unsigned ent[kLevels], wid[kLevels];
#pragma unroll
  for (int l = 0; l < kLevels; l++) ent[l] = wid[l] = 0;
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < a.F; f += gridDim.x * blockDim.x) {
    float fx0, fy0, fx1, fy1;
    if (!face_box_f(a, g, b, f, fx0, fy0, fx1, fy1)) continue;
#pragma unroll
    for (int l = 0; l < kLevels; l++) {
      const int G = 1 << l;
      const float ix = g.invx * (float)G, iy = g.invy * (float)G;
      const int nx = cell_of(fx1, g.lox, ix, G) - cell_of(fx0, g.lox, ix, G) + 1;
      const int ny = cell_of(fy1, g.loy, iy, G) - cell_of(fy0, g.loy, iy, G) + 1;
      const int n = nx * ny;
      if (n > kWideCells) wid[l]++; else ent[l] += n;
    }
  }
  // warp sums -> shared atomics -> one global atomic per level and CTA
  __shared__ unsigned sh[2 * kLevels];
  if (threadIdx.x < 2 * kLevels) sh[threadIdx.x] = 0;
  __syncthreads();
#pragma unroll
  for (int l = 0; l < kLevels; l++) {
    const unsigned e = __reduce_add_sync(DIBR_FULL_MASK, ent[l]);
    const unsigned w = __reduce_add_sync(DIBR_FULL_MASK, wid[l]);
    if ((threadIdx.x & 31) == 0) {
      if (e) atomicAdd(&sh[l], e);
      if (w) atomicAdd(&sh[kLevels + l], w);
    }
  }
  __syncthreads();
  if (threadIdx.x < kLevels) {
    if (sh[threadIdx.x]) atomicAdd(&a.stats[b].entries[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
  } else if (threadIdx.x < 2 * kLevels) {
    if (sh[threadIdx.x]) atomicAdd(&a.stats[b].wides[threadIdx.x - kLevels], sh[threadIdx.x]);
  }
}

// one thread per view: the cheapest level (see the header), grid parameters stored
__global__ void deftet_choose_kernel(const ViewStats* stats, CellGrid* grids, int B, int P,
                                     int F, int cells_cap) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const ViewStats s = stats[b];
  CellGrid g = domain_of(s);
  int best = 0;
  float best_cost = 3.4e38f;
  for (int l = 0; l < kLevels; l++) {
    const long long cells = 1ll << (2 * l);
    if (cells > cells_cap) break;
    if (s.entries[l] > (unsigned long long)kListCap * F) continue;   // must fit the list area
    const float cost = (float)s.entries[l] * (kCostBuild + kCostTest * (float)P / (float)cells) +
                       kCostTest * (float)P * (float)s.wides[l];
    if (cost < best_cost) { best_cost = cost; best = l; }
  }
  g.g = 1 << best;
  g.invx *= (float)g.g;
  g.invy *= (float)g.g;
  grids[b] = g;
}

// FILL = false: count entries per cell, wide faces go straight to the wide list;
// FILL = true : write the face ids (cell_count doubles as the cursor, rebuilt to the count).
// Each CTA owns a contiguous run of a view's faces.  Grids of up to kSmemCells cells are
// binned through a CTA-private histogram in shared memory: shared atomics per (face, cell),
// then ONE global atomic per touched cell and CTA, issued for consecutive cells by
// consecutive lanes (a warp's 32 atomics on one line travel as one transaction) -- scattered
// global atomics on a few thousand counters were the whole cost of the direct form (6 %
// issue-active).  In the fill the CTA reserves its share of every cell's segment the same
// way and deals the slots out with shared cursors.  Finer grids have counters enough to
// take the atomics directly.
constexpr int kSmemCells = 16384;
template <typename T, bool FILL>
__global__ void __launch_bounds__(256) deftet_bin_kernel(const DeftetArgs<T> a) {
  extern __shared__ int s_hist[];
  const int b = blockIdx.y;
  const CellGrid g = a.grids[b];
  if (g.empty) return;
  int* cnt = a.cell_count + (long long)b * a.cells_cap;
  const int* start = a.cell_start + (long long)b * a.cells_cap;
  int* list = a.list + (long long)b * kListCap * a.F;
  const int ncell = g.g * g.g;
  const bool priv = ncell <= kSmemCells;
  const int per = (a.F + gridDim.x - 1) / gridDim.x;
  const int f_lo = blockIdx.x * per, f_hi = min(f_lo + per, a.F);
  if (f_lo >= f_hi) return;
  __shared__ int s_lo, s_hi;       // range of cells this CTA touched (the flush walks only it)
  if (priv) {
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) s_hist[c] = 0;
    if (threadIdx.x == 0) { s_lo = ncell; s_hi = -1; }
    __syncthreads();
  }
  int c_lo = ncell, c_hi = -1;
  // pass 0 (both modes): count into the private histogram (or straight into the counters);
  // pass 1 (fill, private form only): deal the reserved slots out
  for (int pass = 0; pass < ((FILL && priv) ? 2 : 1); pass++) {
    for (int f = f_lo + threadIdx.x; f < f_hi; f += blockDim.x) {
      float fx0, fy0, fx1, fy1;
      if (!face_box_f(a, g, b, f, fx0, fy0, fx1, fy1)) continue;
      const int cx0 = cell_of(fx0, g.lox, g.invx, g.g), cx1 = cell_of(fx1, g.lox, g.invx, g.g);
      const int cy0 = cell_of(fy0, g.loy, g.invy, g.g), cy1 = cell_of(fy1, g.loy, g.invy, g.g);
      if ((cx1 - cx0 + 1) * (cy1 - cy0 + 1) > kWideCells) {
        if (!FILL) a.wide[(long long)b * a.F + atomicAdd(&a.stats[b].wide, 1u)] = f;
        continue;
      }
      c_lo = min(c_lo, cy0 * g.g + cx0);
      c_hi = max(c_hi, cy1 * g.g + cx1);
      for (int cy = cy0; cy <= cy1; cy++)
        for (int cx = cx0; cx <= cx1; cx++) {
          const int c = cy * g.g + cx;
          if (priv) {
            if (pass == 0) atomicAdd(&s_hist[c], 1);
            else list[atomicAdd(&s_hist[c], 1)] = f;
          } else {
            if (FILL) list[start[c] + atomicAdd(cnt + c, 1)] = f;
            else atomicAdd(cnt + c, 1);
          }
        }
    }
    if (!priv || pass == 1) break;
    c_lo = __reduce_min_sync(DIBR_FULL_MASK, c_lo);
    c_hi = __reduce_max_sync(DIBR_FULL_MASK, c_hi);
    if ((threadIdx.x & 31) == 0 && c_hi >= 0) { atomicMin(&s_lo, c_lo); atomicMax(&s_hi, c_hi); }
    __syncthreads();
    // flush: consecutive lanes, consecutive cells, over the touched range only
    for (int c = (s_lo & ~31) + threadIdx.x; c <= s_hi; c += blockDim.x) {
      const int h = s_hist[c];
      if (h == 0) continue;
      if (FILL) s_hist[c] = start[c] + atomicAdd(cnt + c, h);   // first slot of this CTA's share
      else atomicAdd(cnt + c, h);
    }
    if (FILL) __syncthreads();
  }
}

// one CTA per view: exclusive scan of the cell counts, counts reset for the fill's cursors
__global__ void __launch_bounds__(1024) deftet_scan_kernel(const CellGrid* grids, int* cell_start,
                                                           int* cell_count, int cells_cap) {
  __shared__ int warp_sum[32];
  __shared__ int carry;
  const int b = blockIdx.x;
  const int n = grids[b].g * grids[b].g;
  int* cnt = cell_count + (long long)b * cells_cap;
  int* start = cell_start + (long long)b * cells_cap;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // n is a power of four (or 1); cells_cap rows are 16-byte aligned: 4 cells per thread
  for (int base = 0; base < n; base += 4096) {
    const int i = base + 4 * threadIdx.x;
    int4 v4 = make_int4(0, 0, 0, 0);
    if (i + 3 < n) v4 = *reinterpret_cast<const int4*>(cnt + i);
    else if (i < n) { v4.x = cnt[i]; if (i + 1 < n) v4.y = cnt[i + 1]; if (i + 2 < n) v4.z = cnt[i + 2]; }
    const int v = v4.x + v4.y + v4.z + v4.w;
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(DIBR_FULL_MASK, s, o);
      if (lane >= o) s += t;
    }
    if (lane == 31) warp_sum[warp] = s;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(DIBR_FULL_MASK, w, o);
        if (lane >= o) w += t;
      }
      warp_sum[lane] = w;
    }
    __syncthreads();
    const int prefix = carry + (warp ? warp_sum[warp - 1] : 0) + s - v;
    if (i + 3 < n) {
      *reinterpret_cast<int4*>(start + i) =
          make_int4(prefix, prefix + v4.x, prefix + v4.x + v4.y, prefix + v4.x + v4.y + v4.z);
      *reinterpret_cast<int4*>(cnt + i) = make_int4(0, 0, 0, 0);
    } else if (i < n) {
      start[i] = prefix; cnt[i] = 0;
      if (i + 1 < n) { start[i + 1] = prefix + v4.x; cnt[i + 1] = 0; }
      if (i + 2 < n) { start[i + 2] = prefix + v4.x + v4.y; cnt[i + 2] = 0; }
    }
    __syncthreads();
    if (threadIdx.x == 1023) carry = prefix + v;
    __syncthreads();
  }
}

// ------------------------------------------------------------------ per-sample work
// shared-memory staging of one warp: the hits it keeps, and a queue of the candidates that
// passed the box test and wait for the edge test (run 32 at a time, every lane busy)
constexpr int kQueue = 64;
template <typename T>
struct HitBuf {
  T* depth; T* w0; T* w1; T* qv; int* face; int* perm; int* qf;
};

template <typename T>
__host__ __device__ inline size_t hitbuf_bytes(int K, int n2) {
  const size_t b = (size_t)K * (3 * sizeof(T) + sizeof(int)) + (size_t)n2 * sizeof(int) +
                   (size_t)kQueue * (6 * sizeof(T) + sizeof(int));
  return (b + 15) & ~(size_t)15;
}

template <typename T>
__device__ __forceinline__ HitBuf<T> carve(unsigned char* base, int warp, int K, int n2) {
  unsigned char* p = base + (size_t)warp * hitbuf_bytes<T>(K, n2);
  HitBuf<T> h;
  h.depth = reinterpret_cast<T*>(p);
  h.w0 = h.depth + K;
  h.w1 = h.w0 + K;
  h.qv = h.w1 + K;                       // [6][kQueue]
  h.face = reinterpret_cast<int*>(h.qv + 6 * kQueue);
  h.perm = h.face + K;
  h.qf = h.perm + n2;
  return h;
}

// the reference's test of one face against one sample (deftet_cuda.cu:120-163), in two
// parts: the half-open box test (:126) ...
template <typename T>
__device__ __forceinline__ bool deftet_box_test(const DeftetArgs<T>& a, long long bf,
                                                const T (&v)[6], T x0, T y0) {
  if (a.bbox) {
    const T* q = a.bbox + bf * 4;
    return x0 >= __ldg(q) && x0 < __ldg(q + 2) && y0 >= __ldg(q + 1) && y0 < __ldg(q + 3);
  }
  // min3 / max3 drop NaN where torch.min / torch.max keep it: a face with a NaN coordinate
  // that slips through here gets NaN weights in the edge test and is rejected there
  return x0 >= min3(v[0], v[2], v[4]) && x0 < max3(v[0], v[2], v[4]) &&
         y0 >= min3(v[1], v[3], v[5]) && y0 < max3(v[1], v[3], v[5]);
}

// ... and the edge functions, depth and range test (:131-163)
template <typename T>
__device__ __forceinline__ bool deftet_edge_test(const DeftetArgs<T>& a, long long bf,
                                                 const T (&v)[6], T x0, T y0, T dmin, T dmax,
                                                 T eps_t, T& w0, T& w1, T& depth) {
  const T aex = sub_rn(v[0], x0), aey = sub_rn(v[1], y0);
  const T bex = sub_rn(v[2], x0), bey = sub_rn(v[3], y0);
  const T cex = sub_rn(v[4], x0), cey = sub_rn(v[5], y0);
  const T r0 = fma_rn(bex, cey, -mul_rn(bey, cex));
  const T r1 = fma_rn(cex, aey, -mul_rn(cey, aex));
  const T r2 = fma_rn(aex, bey, -mul_rn(aey, bex));
  const T norm = add_rn(add_rn(r0, r1), r2);
  const T den = add_rn(norm, sign_of(norm) ? -eps_t : eps_t);
  w0 = div_rn(r0, den);
  w1 = div_rn(r1, den);
  const T w2 = div_rn(r2, den);
  if (!(w0 >= (T)0 && w1 >= (T)0 && w2 >= (T)0)) return false;
  const T* zp = a.z + bf * 3;
  const T az = __ldg(zp), bz = __ldg(zp + 1), cz = __ldg(zp + 2);
  depth = fma_rn(w2, cz, fma_rn(w0, az, mul_rn(w1, bz)));
  return depth < dmax && depth >= dmin;
}

// order of the output slots: RAW = the reference kernel's insertion order (ascending face),
// else depth descending, ties by ascending face (a stable sort of the insertion order)
template <typename T, bool RAW>
__device__ __forceinline__ bool slot_before(const HitBuf<T>& h, int i, int j) {
  if (j < 0) return i >= 0;
  if (i < 0) return false;
  if (RAW) return h.face[i] < h.face[j];
  const T di = h.depth[i], dj = h.depth[j];
  return di > dj || (di == dj && h.face[i] < h.face[j]);
}

// one warp per sample point; blockIdx.y = view
template <typename T, bool RAW>
__global__ void __launch_bounds__(128) deftet_render_kernel(const DeftetArgs<T> a) {
  typedef typename Vec2<T>::type V2;
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = blockIdx.x * (blockDim.x >> 5) + warp;
  if (p >= a.P) return;
  const int b = blockIdx.y;
  const long long gp = (long long)b * a.P + p;
  const int K = a.K;
  HitBuf<T> h = carve<T>(smem, warp, K, a.n2);
  const V2 c = __ldg(reinterpret_cast<const V2*>(a.pix) + gp);
  const V2 r = __ldg(reinterpret_cast<const V2*>(a.range) + gp);
  const T x0 = c.x, y0 = c.y, dmin = r.x, dmax = r.y;
  const T eps_t = (T)a.eps;
  const CellGrid g = a.grids[b];
  const long long face0 = (long long)b * a.F;

  int n = 0;            // hits held (warp uniform)
  int maxf = -1, maxs = -1;   // largest held face id and its slot, valid once the buffer is full
  int qn = 0;           // candidates queued for the edge test (warp uniform)
  const bool live = !g.empty && !(x0 != x0) && !(y0 != y0);

  // edge test of `cnt` queued candidates starting at queue position q0, hits appended
  auto drain = [&](int q0, int cnt) {
    int f = -1;
    T w0 = 0, w1 = 0, depth = 0;
    bool hit = false;
    if (lane < cnt) {
      const int q = q0 + lane;
      f = h.qf[q];
      T v[6];
#pragma unroll
      for (int k = 0; k < 6; k++) v[k] = h.qv[k * kQueue + q];
      hit = deftet_edge_test(a, face0 + f, v, x0, y0, dmin, dmax, eps_t, w0, w1, depth);
    }
    unsigned m = __ballot_sync(DIBR_FULL_MASK, hit);
    if (m == 0u) return;
    const int c = __popc(m);
    if (n + c <= K) {
      if (hit) {
        const int s = n + __popc(m & lanemask_lt());
        h.face[s] = f; h.depth[s] = depth; h.w0[s] = w0; h.w1[s] = w1;
      }
      n += c;
      return;
    }
    // more hits than slots: keep the K smallest face ids (the reference keeps the first
    // K in ascending face order)
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const int bf_ = __shfl_sync(DIBR_FULL_MASK, f, src);
      const T bd = __shfl_sync(DIBR_FULL_MASK, depth, src);
      const T b0 = __shfl_sync(DIBR_FULL_MASK, w0, src);
      const T b1 = __shfl_sync(DIBR_FULL_MASK, w1, src);
      int s = -1;
      if (n < K) {
        s = n++;
        maxs = -1;
      } else {
        if (maxs < 0) {
          __syncwarp();
          int mf = -1, ms = -1;
          for (int k = lane; k < K; k += 32) {
            const int t = h.face[k];
            if (t > mf) { mf = t; ms = k; }
          }
          const int top = __reduce_max_sync(DIBR_FULL_MASK, mf);
          const unsigned who = __ballot_sync(DIBR_FULL_MASK, mf == top);
          maxs = __shfl_sync(DIBR_FULL_MASK, ms, __ffs(who) - 1);
          maxf = top;
        }
        if (bf_ < maxf) { s = maxs; maxs = -1; }
      }
      if (s >= 0 && lane == 0) { h.face[s] = bf_; h.depth[s] = bd; h.w0[s] = b0; h.w1[s] = b1; }
      __syncwarp();
    }
  };

  for (int pass = 0; pass < 2 && live; pass++) {
    const int* lst;
    int len;
    if (pass == 0) {
      const int cell = cell_of(to_f(y0), g.loy, g.invy, g.g) * g.g +
                       cell_of(to_f(x0), g.lox, g.invx, g.g);
      lst = a.list + face0 * kListCap + a.cell_start[(long long)b * a.cells_cap + cell];
      len = a.cell_count[(long long)b * a.cells_cap + cell];
    } else {
      lst = a.wide + face0;
      len = (int)a.stats[b].wide;
    }
    for (int base = 0; base < len; base += 32) {
      const int i = base + lane;
      int f = -1;
      T v[6];
      bool in_box = false;
      if (i < len) {
        f = __ldg(lst + i);
        load_img6(a.img, face0 + f, v);
        in_box = deftet_box_test(a, face0 + f, v, x0, y0);
      }
      const unsigned m = __ballot_sync(DIBR_FULL_MASK, in_box);
      if (m == 0u) continue;
      if (in_box) {
        const int q = qn + __popc(m & lanemask_lt());
        h.qf[q] = f;
#pragma unroll
        for (int k = 0; k < 6; k++) h.qv[k * kQueue + q] = v[k];
      }
      qn += __popc(m);
      __syncwarp();
      if (qn >= 32) {
        qn -= 32;
        drain(qn, 32);
        __syncwarp();
      }
    }
  }
  if (qn > 0) drain(0, qn);
  __syncwarp();

  // order the held hits.  Up to 128: every lane ranks its own hits against all the others
  // (broadcast reads from shared memory, n compares per hit); beyond: bitonic network over
  // slot numbers in shared memory (padding -1 last)
  if (n > 1 && n <= 128) {
    for (int i = lane; i < n; i += 32) {
      const T di = h.depth[i];
      const int fi = h.face[i];
      int rank = 0;
      for (int j = 0; j < n; j++) {
        const T dj = h.depth[j];
        const int fj = h.face[j];
        rank += RAW ? (fj < fi) : (dj > di || (dj == di && fj < fi));
      }
      h.perm[rank] = i;
    }
    __syncwarp();
  } else if (n > 128) {
    int n2 = 256;
    while (n2 < n) n2 <<= 1;
    for (int k = lane; k < n2; k += 32) h.perm[k] = k < n ? k : -1;
    __syncwarp();
    for (int k = 2; k <= n2; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = lane; t < (n2 >> 1); t += 32) {
          const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
          const int l = i | j;
          const int pi = h.perm[i], pl = h.perm[l];
          const bool up = (i & k) == 0;
          if (up ? slot_before<T, RAW>(h, pl, pi) : slot_before<T, RAW>(h, pi, pl)) {
            h.perm[i] = pl; h.perm[l] = pi;
          }
        }
        __syncwarp();
      }
  }

  // write every slot of the sample (fills included: the library writes all outputs);
  // one slot per lane
  const long long row = gp * K;
  const int D = a.D;
  for (int k = lane; k < K; k += 32) {
    const bool ok = k < n;
    const int s = ok ? (n > 1 ? h.perm[k] : 0) : 0;
    const long long o = row + k;
    const int f = ok ? h.face[s] : -1;
    const T w0 = ok ? h.w0[s] : (T)0, w1 = ok ? h.w1[s] : (T)0;
    a.out_idx[o] = f;
    if (RAW) {
      a.out_depth[o] = ok ? h.depth[s] : -(T)INFINITY; a.out_w0[o] = w0; a.out_w1[o] = w1;
      continue;
    }
    // deftet.py:307  w2 = (face_idx != -1) - (w0 + w1)
    const T w2 = sub_rn(ok ? (T)1 : (T)0, add_rn(w0, w1));
    T* ow = a.out_w + o * 3;
    ow[0] = w0; ow[1] = w1; ow[2] = w2;
    const T* fp = a.feat + (face0 + (ok ? f : 0)) * 3 * D;
    T* of = a.out_feat + o * D;
    // deftet.py:315-317  sum over the three vertices of weight * feature
    for (int d = 0; d < D; d++) {
      T v = (T)0;
      if (ok) v = add_rn(add_rn(mul_rn(w0, __ldg(fp + d)), mul_rn(w1, __ldg(fp + D + d))),
                         mul_rn(w2, __ldg(fp + 2 * D + d)));
      of[d] = v;
    }
  }
}

// ------------------------------------------------------------------ small knum
// knum <= 32 (surface meshes: a few layers per point, short cell lists): a full warp per
// point spends most of its instructions on per-point overhead executed by 32 lanes for a
// handful of hits.  Here a warp takes TWO points, 16 lanes each; every warp-wide vote is
// taken by all 32 lanes and split by half.  Same tests, same kept set, same order.
constexpr int kHalfK = 32;
template <typename T>
struct HalfBuf {
  T depth[kHalfK], w0[kHalfK], w1[kHalfK];
  int face[kHalfK], perm[kHalfK];
  T tdepth[16], tw0[16], tw1[16];     // a step's hits when they do not all fit any more
  int tface[16];
};

template <typename T, bool RAW>
__global__ void __launch_bounds__(128) deftet_render_half_kernel(const DeftetArgs<T> a) {
  typedef typename Vec2<T>::type V2;
  __shared__ HalfBuf<T> sbuf[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane >> 4, sl = lane & 15;
  const int p0 = (blockIdx.x * 4 + warp) * 2;
  if (p0 >= a.P) return;                       // whole warp
  const int p = p0 + sub;
  const bool exists = p < a.P;
  const int b = blockIdx.y;
  const long long gp = (long long)b * a.P + (exists ? p : p0);
  const int K = a.K;
  HalfBuf<T>& h = sbuf[warp * 2 + sub];
  const V2 c = __ldg(reinterpret_cast<const V2*>(a.pix) + gp);
  const V2 r = __ldg(reinterpret_cast<const V2*>(a.range) + gp);
  const T x0 = c.x, y0 = c.y, dmin = r.x, dmax = r.y;
  const T eps_t = (T)a.eps;
  const CellGrid g = a.grids[b];
  const long long face0 = (long long)b * a.F;
  const unsigned below = (1u << sl) - 1u;

  int n = 0;                  // hits held by this half's point (uniform within the half)
  int maxf = -1, maxs = -1;   // lane sl == 0 only: largest held face id and its slot
  const bool live = exists && !g.empty && !(x0 != x0) && !(y0 != y0);
  for (int pass = 0; pass < 2; pass++) {
    const int* lst = a.wide + face0;
    int len = 0;
    if (live) {
      if (pass == 0) {
        const int cell = cell_of(to_f(y0), g.loy, g.invy, g.g) * g.g +
                         cell_of(to_f(x0), g.lox, g.invx, g.g);
        lst = a.list + face0 * kListCap + a.cell_start[(long long)b * a.cells_cap + cell];
        len = a.cell_count[(long long)b * a.cells_cap + cell];
      } else {
        len = (int)a.stats[b].wide;
      }
    }
    const int maxlen = max(len, __shfl_xor_sync(DIBR_FULL_MASK, len, 16));
    for (int base = 0; base < maxlen; base += 16) {
      const int i = base + sl;
      int f = -1;
      T w0 = 0, w1 = 0, depth = 0;
      bool hit = false;
      if (i < len) {
        f = __ldg(lst + i);
        T v[6];
        load_img6(a.img, face0 + f, v);
        hit = deftet_box_test(a, face0 + f, v, x0, y0) &&
              deftet_edge_test(a, face0 + f, v, x0, y0, dmin, dmax, eps_t, w0, w1, depth);
      }
      const unsigned m = (__ballot_sync(DIBR_FULL_MASK, hit) >> (16 * sub)) & 0xffffu;
      const int cnt = __popc(m);
      const bool fits = n + cnt <= K;
      if (hit) {
        const int t = __popc(m & below);
        if (fits) {
          h.face[n + t] = f; h.depth[n + t] = depth; h.w0[n + t] = w0; h.w1[n + t] = w1;
        } else {
          h.tface[t] = f; h.tdepth[t] = depth; h.tw0[t] = w0; h.tw1[t] = w1;
        }
      }
      __syncwarp();
      if (cnt && !fits && sl == 0) {
        // more hits than slots: keep the K smallest face ids (the reference keeps the
        // first K in ascending face order); one lane, K <= 32
        int nn = n;
        for (int t = 0; t < cnt; t++) {
          const int tf = h.tface[t];
          int s = -1;
          if (nn < K) {
            s = nn++;
            maxs = -1;
          } else {
            if (maxs < 0) {
              maxf = -1;
              for (int k = 0; k < K; k++)
                if (h.face[k] > maxf) { maxf = h.face[k]; maxs = k; }
            }
            if (tf < maxf) { s = maxs; maxs = -1; }
          }
          if (s >= 0) {
            h.face[s] = tf; h.depth[s] = h.tdepth[t]; h.w0[s] = h.tw0[t]; h.w1[s] = h.tw1[t];
          }
        }
      }
      n = min(n + cnt, K);
      __syncwarp();
    }
  }

  // order: every lane ranks its hits against all others (n <= 32, broadcast reads)
  if (n > 1) {
    for (int i = sl; i < n; i += 16) {
      const T di = h.depth[i];
      const int fi = h.face[i];
      int rank = 0;
      for (int j = 0; j < n; j++) {
        const T dj = h.depth[j];
        const int fj = h.face[j];
        rank += RAW ? (fj < fi) : (dj > di || (dj == di && fj < fi));
      }
      h.perm[rank] = i;
    }
  }
  __syncwarp();
  if (!exists) return;

  const long long row = gp * K;
  const int D = a.D;
  for (int k = sl; k < K; k += 16) {
    const bool ok = k < n;
    const int s = ok ? (n > 1 ? h.perm[k] : 0) : 0;
    const long long o = row + k;
    const int f = ok ? h.face[s] : -1;
    const T w0 = ok ? h.w0[s] : (T)0, w1 = ok ? h.w1[s] : (T)0;
    a.out_idx[o] = f;
    if (RAW) {
      a.out_depth[o] = ok ? h.depth[s] : -(T)INFINITY; a.out_w0[o] = w0; a.out_w1[o] = w1;
      continue;
    }
    const T w2 = sub_rn(ok ? (T)1 : (T)0, add_rn(w0, w1));     // deftet.py:307
    T* ow = a.out_w + o * 3;
    ow[0] = w0; ow[1] = w1; ow[2] = w2;
    const T* fp = a.feat + (face0 + (ok ? f : 0)) * 3 * D;
    T* of = a.out_feat + o * D;
    for (int d = 0; d < D; d++) {                                // deftet.py:315-317
      T v = (T)0;
      if (ok) v = add_rn(add_rn(mul_rn(w0, __ldg(fp + d)), mul_rn(w1, __ldg(fp + D + d))),
                         mul_rn(w2, __ldg(fp + 2 * D + d)));
      of[d] = v;
    }
  }
}

// ------------------------------------------------------------------ backward
// deftet_cuda.cu:239-401: one thread per (sample, layer) slot; the six image-coordinate
// terms are summed over the features before the atomics (the reference issues 6 per feature).
template <typename T>
struct DeftetBwdArgs {
  const T* grad;         // (B, P, K, D)
  const int64_t* idx;    // (B, P, K)
  const T* weights;      // (B, P, K, 3)
  const T* img;          // (B, F, 3, 2)
  const T* feat;         // (B, F, 3, D)
  T* grad_img;           // (B, F, 3, 2)  accumulated
  T* grad_feat;          // (B, F, 3, D)  accumulated
  long long slots;       // B * P * K
  long long per_view;    // P * K
  int F, D;
  float eps;
};

// the gradient of one valid slot s holding face f
template <typename T>
__device__ __forceinline__ void deftet_backward_slot(const DeftetBwdArgs<T>& a, long long s,
                                                     long long f) {
  {
    const long long bf = (s / a.per_view) * a.F + f;
    const int D = a.D;
    const T* g = a.grad + s * D;
    const T aw = __ldg(a.weights + s * 3), bw = __ldg(a.weights + s * 3 + 1),
            cw = __ldg(a.weights + s * 3 + 2);
    T* gf = a.grad_feat + bf * 3 * D;
    for (int d = 0; d < D; d++) {
      const T gd = __ldg(g + d);
      atomicAdd(gf + d, gd * aw);
      atomicAdd(gf + D + d, gd * bw);
      atomicAdd(gf + 2 * D + d, gd * cw);
    }
    T v[6];
    load_img6(a.img, bf, v);
    const T ax = v[0], ay = v[1], bx = v[2], by = v[3], cx = v[4], cy = v[5];
    const T x0 = aw * ax + bw * bx + cw * cx;
    const T y0 = aw * ay + bw * by + cw * cy;
    const T m = bx - ax, p = by - ay, n = cx - ax, q = cy - ay, sx = x0 - ax, t = y0 - ay;
    const T k1 = sx * q - n * t;
    const T k2 = m * t - sx * p;
    T k3 = m * q - n * p;
    k3 = (T)((double)k3 + copysign((double)a.eps, (double)k3));   // :321
    // d(w1 k3^2), d(w2 k3^2) with respect to m, n, p, q, s, t (:324-349)
    const T dw1dm = -q * k1, dw1dn = -t * k3 + p * k1, dw1dp = n * k1, dw1dq = sx * k3 - m * k1;
    const T dw1ds = q * k3, dw1dt = -n * k3;
    const T dw2dm = t * k3 - q * k2, dw2dn = p * k2, dw2dp = -sx * k3 + n * k2, dw2dq = -m * k2;
    const T dw2ds = -p * k3, dw2dt = m * k3;
    const T dw1dax = -(dw1dm + dw1dn + dw1ds), dw1day = -(dw1dp + dw1dq + dw1dt);
    const T dw2dax = -(dw2dm + dw2dn + dw2ds), dw2day = -(dw2dp + dw2dq + dw2dt);
    const T inv = (T)1 / (k3 * k3);
    const T* fp = a.feat + bf * 3 * D;
    T s_ax = 0, s_ay = 0, s_bx = 0, s_by = 0, s_cx = 0, s_cy = 0;
    for (int d = 0; d < D; d++) {
      const T c0 = __ldg(fp + d), c1 = __ldg(fp + D + d), c2 = __ldg(fp + 2 * D + d);
      const T e1 = c1 - c0, e2 = c2 - c0;
      const T dl = __ldg(g + d) * inv;
      s_ax += dl * (e1 * dw1dax + e2 * dw2dax);
      s_ay += dl * (e1 * dw1day + e2 * dw2day);
      s_bx += dl * (e1 * dw1dm + e2 * dw2dm);
      s_by += dl * (e1 * dw1dp + e2 * dw2dp);
      s_cx += dl * (e1 * dw1dn + e2 * dw2dn);
      s_cy += dl * (e1 * dw1dq + e2 * dw2dq);
    }
    T* gi = a.grad_img + bf * 6;
    atomicAdd(gi + 0, s_ax); atomicAdd(gi + 1, s_ay);
    atomicAdd(gi + 2, s_bx); atomicAdd(gi + 3, s_by);
    atomicAdd(gi + 4, s_cx); atomicAdd(gi + 5, s_cy);
  }
}

// Valid slots are few (a handful of layers out of knum per point) and sit at the front of
// every point's row, so one thread per slot would leave ~90 % of the lanes idle in the heavy
// part.  Each warp walks a contiguous range of slots 32 at a time (coalesced index reads),
// queues the valid ones in shared memory (ballot compaction) and runs the gradient 32 valid
// slots at a time with every lane busy.  No assumption on where the -1 entries are.
constexpr int kBwdWarps = 8;
template <typename T>
__global__ void __launch_bounds__(kBwdWarps * 32) deftet_backward_kernel(const DeftetBwdArgs<T> a) {
  __shared__ long long q_slot[kBwdWarps][64];
  __shared__ int q_face[kBwdWarps][64];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long nwarps = (long long)gridDim.x * kBwdWarps;
  const long long gw = (long long)blockIdx.x * kBwdWarps + warp;
  // ranges in multiples of 32 slots
  const long long tiles = (a.slots + 31) / 32;
  const long long per = (tiles + nwarps - 1) / nwarps;
  const long long t0 = gw * per, t1 = min(t0 + per, tiles);
  int queued = 0;
  for (long long t = t0; t < t1; t++) {
    const long long s = t * 32 + lane;
    const long long f = s < a.slots ? a.idx[s] : -1;
    const unsigned m = __ballot_sync(DIBR_FULL_MASK, f >= 0);
    if (f >= 0) {
      const int q = queued + __popc(m & lanemask_lt());
      q_slot[warp][q] = s;
      q_face[warp][q] = (int)f;
    }
    queued += __popc(m);
    __syncwarp();
    if (queued >= 32) {
      queued -= 32;
      deftet_backward_slot(a, q_slot[warp][queued + lane], (long long)q_face[warp][queued + lane]);
      __syncwarp();
    }
  }
  if (lane < queued) deftet_backward_slot(a, q_slot[warp][lane], (long long)q_face[warp][lane]);
}

inline int64_t up256(int64_t v) { return (v + 255) / 256 * 256; }

struct DeftetLayout {
  int cells_cap;
  int64_t off_stats, off_grids, off_start, off_count, off_wide, off_list, bytes, zero_bytes;
};

DeftetLayout deftet_layout(int B, int F, int P) {
  DeftetLayout L;
  int64_t cap = 4 * (int64_t)P + 16;      // no point in many more cells than samples
  if (cap > (1ll << (2 * (kLevels - 1)))) cap = 1ll << (2 * (kLevels - 1));
  L.cells_cap = (int)((cap + 3) / 4 * 4);      // rows stay 16-byte aligned
  int64_t off = 0;
  L.off_stats = off; off = up256(off + (int64_t)B * sizeof(ViewStats));
  L.off_count = off; off = up256(off + (int64_t)B * cap * 4);
  L.zero_bytes = off;                       // stats and counts are memset, the rest written
  L.off_grids = off; off = up256(off + (int64_t)B * sizeof(CellGrid));
  L.off_start = off; off = up256(off + (int64_t)B * cap * 4);
  L.off_wide = off; off = up256(off + (int64_t)B * F * 4);
  L.off_list = off; off = up256(off + (int64_t)B * F * 4 * kListCap);
  L.bytes = off;
  return L;
}

#define DEF_CUDA(expr)                                                               \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess)                                                           \
      return fail(DIBR_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));      \
  } while (0)

template <typename T>
int deftet_forward(bool raw, int B, int F, int P, int K, int D, const void* z, const void* img,
                   const void* bbox, const void* feat, const void* pix, const void* range,
                   float eps, int64_t* out_idx, void* o1, void* o2, void* o3, void* ws,
                   int64_t ws_bytes, cudaStream_t st) {
  if (B < 0 || F < 0 || P < 0 || K < 0 || D < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (B > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", B);
  if ((int64_t)F * kListCap > 0x7fffffff) return fail(DIBR_E_INVALID, "too many faces");
  if (B == 0 || P == 0 || K == 0) return DIBR_OK;
  if (!pix || !range || !out_idx) return fail(DIBR_E_INVALID, "null pointer");
  if (F > 0 && (!z || !img)) return fail(DIBR_E_INVALID, "null input");
  if (raw ? (!o1 || !o2 || !o3) : (!o2 || (D > 0 && (!o1 || (F > 0 && !feat)))))
    return fail(DIBR_E_INVALID, "null output");
  const DeftetLayout L = deftet_layout(B, F, P);
  if (!ws || ws_bytes < L.bytes)
    return fail(DIBR_E_WORKSPACE, "workspace of %lld bytes given, %lld needed",
                (long long)ws_bytes, (long long)L.bytes);
  if (((uintptr_t)ws & 255) != 0) return fail(DIBR_E_INVALID, "workspace must be 256-byte aligned");
  DeftetArgs<T> a = {};
  a.z = (const T*)z; a.img = (const T*)img; a.bbox = (const T*)bbox; a.feat = (const T*)feat;
  a.pix = (const T*)pix; a.range = (const T*)range;
  a.out_idx = out_idx;
  if (raw) { a.out_depth = (T*)o1; a.out_w0 = (T*)o2; a.out_w1 = (T*)o3; }
  else { a.out_feat = (T*)o1; a.out_w = (T*)o2; }
  char* base = (char*)ws;
  a.stats = (ViewStats*)(base + L.off_stats);
  a.grids = (CellGrid*)(base + L.off_grids);
  a.cell_count = (int*)(base + L.off_count);
  a.cell_start = (int*)(base + L.off_start);
  a.wide = (int*)(base + L.off_wide);
  a.list = (int*)(base + L.off_list);
  a.B = B; a.F = F; a.P = P; a.K = K; a.D = D; a.cells_cap = L.cells_cap; a.eps = eps;
  int n2 = 1;
  while (n2 < K) n2 <<= 1;
  a.n2 = n2;

  DEF_CUDA(cudaMemsetAsync(base, 0, (size_t)L.zero_bytes, st));
  {
    const int pb = (int)std::min<int64_t>((P + 255) / 256, 64);
    if (F > 0) {
      deftet_bounds_kernel<T><<<dim3(pb, B), 256, 0, st>>>(a);
      const int lb = (int)std::max<int64_t>(1, std::min<int64_t>((F + 1023) / 1024, (148 * 4 + B - 1) / B));
      deftet_levels_kernel<T><<<dim3(lb, B), 256, 0, st>>>(a);
    }
    deftet_choose_kernel<<<(B + 127) / 128, 128, 0, st>>>(a.stats, a.grids, B, P, F, a.cells_cap);
    if (F > 0) {
      const size_t hist = (size_t)kSmemCells * sizeof(int);
      // per device and cheap: set on every call
      DEF_CUDA(cudaFuncSetAttribute(deftet_bin_kernel<T, false>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist));
      DEF_CUDA(cudaFuncSetAttribute(deftet_bin_kernel<T, true>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist));
      // three 64-KB CTAs per SM; a CTA's run of faces stays long enough to amortise its flush
      const int bb = (int)std::max<int64_t>(1, std::min<int64_t>((F + 1023) / 1024, (148 * 3 + B - 1) / B));
      deftet_bin_kernel<T, false><<<dim3(bb, B), 256, hist, st>>>(a);
      deftet_scan_kernel<<<B, 1024, 0, st>>>(a.grids, a.cell_start, a.cell_count, a.cells_cap);
      deftet_bin_kernel<T, true><<<dim3(bb, B), 256, hist, st>>>(a);
    }
    DEF_CUDA(cudaGetLastError());
  }
  if (K <= kHalfK) {            // two points per warp (deftet_render_half_kernel)
    auto kern = raw ? deftet_render_half_kernel<T, true> : deftet_render_half_kernel<T, false>;
    kern<<<dim3((P + 7) / 8, B), 128, 0, st>>>(a);
    DEF_CUDA(cudaGetLastError());
    return DIBR_OK;
  }
  // one warp per sample; warps per CTA limited by the staging of K hits per warp
  const size_t per_warp = hitbuf_bytes<T>(K, n2);
  int warps = (int)std::min<size_t>(4, (200 * 1024) / per_warp);
  if (warps < 1) return fail(DIBR_E_INVALID, "knum %d needs %zu bytes of shared memory per sample, "
                             "more than an SM has", K, per_warp);
  const size_t smem = per_warp * warps;
  auto kern = raw ? deftet_render_kernel<T, true> : deftet_render_kernel<T, false>;
  if (smem > 48 * 1024)
    DEF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int blocks = (P + warps - 1) / warps;
  kern<<<dim3(blocks, B), warps * 32, smem, st>>>(a);
  DEF_CUDA(cudaGetLastError());
  return DIBR_OK;
}

template <typename T>
int deftet_backward(int B, int F, int P, int K, int D, const void* grad, const int64_t* idx,
                    const void* weights, const void* img, const void* feat, float eps,
                    void* grad_img, void* grad_feat, cudaStream_t st) {
  if (B < 0 || F < 0 || P < 0 || K < 0 || D < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (B == 0 || F == 0) return DIBR_OK;
  if (!grad_img || (D > 0 && !grad_feat)) return fail(DIBR_E_INVALID, "null output");
  DEF_CUDA(cudaMemsetAsync(grad_img, 0, (size_t)B * F * 6 * sizeof(T), st));
  if (D > 0) DEF_CUDA(cudaMemsetAsync(grad_feat, 0, (size_t)B * F * 3 * D * sizeof(T), st));
  const long long slots = (long long)B * P * K;
  if (slots == 0 || D == 0) return DIBR_OK;
  if (!grad || !idx || !weights || !img || !feat) return fail(DIBR_E_INVALID, "null input");
  DeftetBwdArgs<T> a = {};
  a.grad = (const T*)grad; a.idx = idx; a.weights = (const T*)weights; a.img = (const T*)img;
  a.feat = (const T*)feat; a.grad_img = (T*)grad_img; a.grad_feat = (T*)grad_feat;
  a.slots = slots; a.per_view = (long long)P * K; a.F = F; a.D = D; a.eps = eps;
  const long long blocks = std::min<long long>((slots + 32 * kBwdWarps - 1) / (32 * kBwdWarps), 148 * 8);
  deftet_backward_kernel<T><<<(unsigned)blocks, kBwdWarps * 32, 0, st>>>(a);
  DEF_CUDA(cudaGetLastError());
  return DIBR_OK;
}

}  // namespace
}  // namespace dibr

using namespace dibr;

extern "C" {

DIBR_API int64_t dibr_deftet_workspace_bytes(int batch, int num_faces, int num_points) {
  if (batch < 0 || num_faces < 0 || num_points < 0)
    return fail(DIBR_E_INVALID, "invalid sizes");
  return deftet_layout(batch, num_faces, num_points).bytes;
}

DIBR_API int dibr_deftet_fwd(int dtype, int batch, int num_faces, int num_points, int knum,
                             int feat_dim, const void* face_vertices_z,
                             const void* face_vertices_image, const void* face_features,
                             const void* pixel_coords, const void* render_ranges, float eps,
                             int64_t* face_idx, void* interpolated_features, void* weights,
                             void* ws, int64_t ws_bytes, void* stream) {
  if (dtype == DIBR_F32)
    return deftet_forward<float>(false, batch, num_faces, num_points, knum, feat_dim,
                                 face_vertices_z, face_vertices_image, nullptr, face_features,
                                 pixel_coords, render_ranges, eps, face_idx,
                                 interpolated_features, weights, nullptr, ws, ws_bytes,
                                 (cudaStream_t)stream);
  if (dtype == DIBR_F64)
    return deftet_forward<double>(false, batch, num_faces, num_points, knum, feat_dim,
                                  face_vertices_z, face_vertices_image, nullptr, face_features,
                                  pixel_coords, render_ranges, eps, face_idx,
                                  interpolated_features, weights, nullptr, ws, ws_bytes,
                                  (cudaStream_t)stream);
  return fail(DIBR_E_INVALID, "dtype must be DIBR_F32 or DIBR_F64");
}

DIBR_API int dibr_deftet_fwd_raw(int dtype, int batch, int num_faces, int num_points, int knum,
                                 const void* face_vertices_z, const void* face_vertices_image,
                                 const void* face_bboxes, const void* pixel_coords,
                                 const void* render_ranges, float eps, int64_t* face_idx,
                                 void* pixel_depths, void* w0, void* w1, void* ws,
                                 int64_t ws_bytes, void* stream) {
  if (dtype == DIBR_F32)
    return deftet_forward<float>(true, batch, num_faces, num_points, knum, 0, face_vertices_z,
                                 face_vertices_image, face_bboxes, nullptr, pixel_coords,
                                 render_ranges, eps, face_idx, pixel_depths, w0, w1, ws, ws_bytes,
                                 (cudaStream_t)stream);
  if (dtype == DIBR_F64)
    return deftet_forward<double>(true, batch, num_faces, num_points, knum, 0, face_vertices_z,
                                  face_vertices_image, face_bboxes, nullptr, pixel_coords,
                                  render_ranges, eps, face_idx, pixel_depths, w0, w1, ws,
                                  ws_bytes, (cudaStream_t)stream);
  return fail(DIBR_E_INVALID, "dtype must be DIBR_F32 or DIBR_F64");
}

DIBR_API int dibr_deftet_bwd(int dtype, int batch, int num_faces, int num_points, int knum,
                             int feat_dim, const void* grad_interpolated_features,
                             const int64_t* face_idx, const void* weights,
                             const void* face_vertices_image, const void* face_features,
                             float eps, void* grad_face_vertices_image,
                             void* grad_face_features, void* stream) {
  if (dtype == DIBR_F32)
    return deftet_backward<float>(batch, num_faces, num_points, knum, feat_dim,
                                  grad_interpolated_features, face_idx, weights,
                                  face_vertices_image, face_features, eps,
                                  grad_face_vertices_image, grad_face_features,
                                  (cudaStream_t)stream);
  if (dtype == DIBR_F64)
    return deftet_backward<double>(batch, num_faces, num_points, knum, feat_dim,
                                   grad_interpolated_features, face_idx, weights,
                                   face_vertices_image, face_features, eps,
                                   grad_face_vertices_image, grad_face_features,
                                   (cudaStream_t)stream);
  return fail(DIBR_E_INVALID, "dtype must be DIBR_F32 or DIBR_F64");
}

}  // extern "C"
