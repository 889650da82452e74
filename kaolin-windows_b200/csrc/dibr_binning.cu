// dibr_binning.cu -- per-view face binning (north_star subsystem (a)).
//
// Builds, for every (view, screen bin), the compacted list of faces whose box touches
// the bin, IN ASCENDING FACE ORDER (the soft mask's "first knum faces" rule,
// dibr_soft_mask_cuda.cu:80,165-171, depends on that order; the rasterizer does not).
// It is a stable multi-bin partition done in three launches, without any global atomics
// on the data path and without a sort:
//
//   bin_count  one WARP per chunk of S consecutive faces of one view.  Per step the warp
//              loads 32 faces, derives each face's bin rectangle and bumps the warp's
//              private shared-memory histogram (native integer shared atomics; the order
//              is irrelevant for counting).  Non-zero counts go to mat[view][bin][chunk].
//   bin_scan   one warp per (view, bin): exclusive scan of mat[view][bin][:] over chunks
//              (rows padded to int4, four independent 16-byte loads per lane in flight, then
//              shuffle scans), then one list segment per bin is reserved from a global
//              cursor (one atomic per CTA of 8 bins).
//   bin_fill   soft-mask list (ordered), up to kMaskBins bins: per step of 32 faces every lane
//              ORs its lane bit into a per-warp shared-memory mask word of each bin of its
//              rectangle; a lane's slot in a bin is then the bin's counter +
//              __popc(mask & lanemask_lt), so every bin's list comes out sorted by face index
//              with all lanes working in parallel; the lowest lane of each mask advances the
//              counter and clears it.  More bins (the mask words would double the shared
//              memory per warp and halve the resident warps: measured slower at 1024 bins):
//              the warp walks the distinct bins its 32 faces touch in ascending bin id (warp
//              min-reduce) and ranks the faces touching the current bin with a ballot.
//              rasterizer list (order free, its depth test is order independent): slots
//              are taken with shared-memory atomics on the same per-warp counters.
//
// All three kernels take up to two jobs (raster list, soft-mask list) via blockIdx.z so
// that the fused forward bins both in the same launches.
#include "dibr_common.cuh"

namespace dibr {


constexpr int kBinWarps = 4;  // warps (= chunks) per CTA of bin_count / bin_fill
constexpr int kMaskBins = 256; // ordered fill: lane masks up to this many bins, else the bin walk

// Bin rectangle of a face box; false if it touches no pixel.
__device__ __forceinline__ bool bin_rect(const Grid& g, const Face2D& f, int bin_log2, int& bx0,
                                         int& bx1, int& by0, int& by1) {
  int x_lo, x_hi, y_lo, y_hi;
  if (!pixel_range(g, f.xmin, f.ymin, f.xmax, f.ymax, x_lo, x_hi, y_lo, y_hi)) return false;
  bx0 = x_lo >> bin_log2;
  bx1 = x_hi >> bin_log2;
  by0 = y_lo >> bin_log2;
  by1 = y_hi >> bin_log2;
  return true;
}

template <bool FILL>
__global__ void __launch_bounds__(kBinWarps * 32)
bin_walk_kernel(const BinJobs jobs, const Plan p) {
  pdl_prologue();
  if (FILL && jobs.report && (blockIdx.x | blockIdx.y | blockIdx.z | threadIdx.x) == 0) {
    volatile unsigned long long* r = jobs.report;
    r[0] = jobs.totals[0];
    r[1] = jobs.totals[1];
    __threadfence_system();
  }
  extern __shared__ int s_hist[];  // [kBinWarps][NB] counters (+ [kBinWarps][NB] lane masks: fill)
  const BinJob& job = jobs.j[blockIdx.z];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const int b = blockIdx.y;
  const int chunk = blockIdx.x * kBinWarps + warp;
  if (chunk >= p.nchunks) return;  // whole warp leaves; no block-wide barrier below
  int* hist = s_hist + warp * p.NB;
  unsigned* lmask = reinterpret_cast<unsigned*>(s_hist + (kBinWarps + warp) * p.NB);  // FILL only
  const Grid g = plan_grid(p, job.mult);
  const long long row = (long long)b * p.NB;  // first bin row of this view in mat/start

  int lo_bin = p.NB, hi_bin = -1;
  if (FILL) {
    const int2 r = job.range[(long long)b * p.nchunks + chunk];
    lo_bin = r.x;
    hi_bin = r.y;
    // 64-bit sum clamped below 2^31: a segment beyond the capacity (an under-provisioned
    // first attempt) must not wrap into a valid slot
    for (int i = lo_bin + lane; i <= hi_bin; i += 32)
      hist[i] = (int)min((long long)job.start[row + i] +
                             job.mat[(row + i) * p.mstride + chunk], kMaxListEntries);
    if (job.ordered && p.NB <= kMaskBins)
      for (int i = lo_bin + lane; i <= hi_bin; i += 32) lmask[i] = 0u;
  } else {
    for (int i = lane; i < p.NB; i += 32) hist[i] = 0;
  }
  __syncwarp();

  const int f_begin = chunk * p.S;
  const int f_end = min(f_begin + p.S, p.F);
  for (int f0 = f_begin; f0 < f_end; f0 += 32) {
    const int f = f0 + lane;
    int bx0 = 0, bx1 = -1, by0 = 0, by1 = -1;
    bool act = false;
    Face2D fc;
    if (f < f_end) {
      if (job.rect) {
        // soft list: the exact pixel rectangle, computed once (counting pass) and kept
        ushort4* rp = job.rect + (long long)b * p.F + f;
        int x_lo = 1, x_hi = 0, y_lo = 1, y_hi = 0;
        if (FILL) {
          const ushort4 r = *rp;
          x_lo = r.x; x_hi = r.y; y_lo = r.z; y_hi = r.w;
          act = x_lo <= x_hi;
        } else {
          act = load_face2d(job, b, f, fc) &&
                exact_pixel_rect(g, fc.xmin, fc.ymin, fc.xmax, fc.ymax, x_lo, x_hi, y_lo, y_hi);
          *rp = act ? make_ushort4((unsigned short)x_lo, (unsigned short)x_hi,
                                   (unsigned short)y_lo, (unsigned short)y_hi)
                    : make_ushort4(1, 0, 1, 0);
        }
        if (act) {
          bx0 = x_lo >> p.bin_log2; bx1 = x_hi >> p.bin_log2;
          by0 = y_lo >> p.bin_log2; by1 = y_hi >> p.bin_log2;
        }
      } else if (load_face2d(job, b, f, fc)) {
        act = bin_rect(g, fc, p.bin_log2, bx0, bx1, by0, by1);
      }
    }
    // Where the order inside a bin's list does not matter -- the counting pass, and the
    // fill of the rasterizer's list (its depth test is order independent) -- every lane
    // simply bumps the shared-memory counters of the bins of its rectangle.
    if (!FILL || !job.ordered) {
      float4 q0, q1, q2;
      if (FILL && job.rec && by0 <= by1) {
        float za, zb, zc;
        face_depths(job, job.z, b, f, za, zb, zc);
        q0 = make_float4(fc.ax, fc.ay, fc.bx, fc.by);
        q1 = make_float4(fc.cx, fc.cy, za, zb);
        q2 = make_float4(zc, __int_as_float(f), 0.f, 0.f);
      }
      for (int cy = by0; cy <= by1; cy++)
        for (int cx = bx0; cx <= bx1; cx++) {
          const int bin = cy * p.nbx + cx;
          const int slot = atomicAdd(&hist[bin], 1);
          if (FILL) {
            if (slot >= 0 && slot < job.capacity) {
              if (job.rec) {
                float4* r = job.rec + (long long)slot * 3;
                r[0] = q0; r[1] = q1; r[2] = q2;
              } else {
                job.list[slot] = f;
              }
            }
          } else {
            lo_bin = min(lo_bin, bin);
            hi_bin = max(hi_bin, bin);
          }
        }
      continue;
    }
    // Ordered emission (soft-mask list, fill pass only: the counting pass took the branch
    // above).
    if (FILL && p.NB <= kMaskBins) {
      // Three lane-parallel sweeps over the lane's own bin rectangle:
      //   mark   OR the lane bit into the mask word of every bin of the rectangle
      //   place  slot = bin counter + number of lower lanes (= lower face ids) in the mask
      //   close  the lowest lane of a mask adds its population to the counter and clears it
      // (a lane that reads an already cleared mask in `close` sees 0 and is not its lowest
      // lane; counters only move in `close`, masks only in `mark` and `close`)
      if (act)
        for (int cy = by0; cy <= by1; cy++)
          for (int cx = bx0; cx <= bx1; cx++) atomicOr(&lmask[cy * p.nbx + cx], 1u << lane);
      __syncwarp();
      if (act)
        for (int cy = by0; cy <= by1; cy++)
          for (int cx = bx0; cx <= bx1; cx++) {
            const int bin = cy * p.nbx + cx;
            const long long slot = (long long)hist[bin] + __popc(lmask[bin] & lanemask_lt());
            if (slot < job.capacity) job.list[slot] = f;
          }
      __syncwarp();
      if (act)
        for (int cy = by0; cy <= by1; cy++)
          for (int cx = bx0; cx <= bx1; cx++) {
            const int bin = cy * p.nbx + cx;
            const unsigned mk = lmask[bin];
            if ((mk & ((2u << lane) - 1u)) == (1u << lane)) {   // own bit is the lowest set
              hist[bin] += __popc(mk);
              lmask[bin] = 0u;
            }
          }
      __syncwarp();
    } else if (FILL) {
      // Each lane walks its rectangle row-major (ascending bin id); the warp repeatedly takes
      // the smallest pending bin id, so all faces touching that bin are handled together, in
      // lane (= face) order.
      int cx = bx0, cy = by0;
      while (true) {
        const int cur = act ? cy * p.nbx + cx : 0x7fffffff;
        const int m = __reduce_min_sync(DIBR_FULL_MASK, cur);
        if (m == 0x7fffffff) break;
        const bool hit = (cur == m);
        const unsigned mask = __ballot_sync(DIBR_FULL_MASK, hit);
        const int base = hist[m];
        __syncwarp();
        if (lane == 0) hist[m] = base + __popc(mask);
        if (hit) {
          const long long slot = (long long)base + __popc(mask & lanemask_lt());
          if (slot < job.capacity) job.list[slot] = f;
          if (++cx > bx1) {
            cx = bx0;
            if (++cy > by1) act = false;
          }
        }
        __syncwarp();
      }
    }
  }
  if (!FILL) {
    lo_bin = __reduce_min_sync(DIBR_FULL_MASK, lo_bin);
    hi_bin = __reduce_max_sync(DIBR_FULL_MASK, hi_bin);
    __syncwarp();
    for (int i = lo_bin + lane; i <= hi_bin; i += 32) {
      const int c = hist[i];
      if (c) job.mat[(row + i) * p.mstride + chunk] = c;
    }
    if (lane == 0) job.range[(long long)b * p.nchunks + chunk] = make_int2(lo_bin, hi_bin);
  }
}

// One warp per (view, bin); 8 warps per CTA share one cursor atomic.
__global__ void __launch_bounds__(256) bin_scan_kernel(const BinJobs jobs, const Plan p) {
  pdl_prologue();
  __shared__ int s_tot[8];
  __shared__ long long s_base;
  const BinJob& job = jobs.j[blockIdx.z];
  const int warp = threadIdx.x >> 5, lane = lane_id();
  const long long nrows = (long long)p.B * p.NB;
  const long long rowi = (long long)blockIdx.x * 8 + warp;
  int carry = 0;
  if (rowi < nrows) {
    // rows are mstride = 4 n4 ints, 16-byte aligned, zero beyond nchunks: each lane scans
    // int4 groups; four groups per lane are loaded before the first shuffle scan so that one
    // L2 round trip covers 512 chunks
    int4* m4 = reinterpret_cast<int4*>(job.mat + rowi * p.mstride);
    const int n4 = p.mstride >> 2;
    constexpr int U = 4;
    for (int i0 = 0; i0 < n4; i0 += 32 * U) {
      int4 v[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int idx = i0 + u * 32 + lane;
        v[u] = idx < n4 ? m4[idx] : make_int4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        if (i0 + u * 32 >= n4) break;   // warp-uniform
        const int idx = i0 + u * 32 + lane;
        const int sum = v[u].x + v[u].y + v[u].z + v[u].w;
        int incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
          if (lane >= d) incl += t;
        }
        const int e = carry + incl - sum;
        if (idx < n4)
          m4[idx] = make_int4(e, e + v[u].x, e + v[u].x + v[u].y, e + v[u].x + v[u].y + v[u].z);
        carry += __shfl_sync(DIBR_FULL_MASK, incl, 31);
      }
    }
  }
  if (lane == 0) s_tot[warp] = carry;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int i = 0; i < 8; i++) t += s_tot[i];
    s_base = t ? (long long)atomicAdd(job.total, (unsigned long long)t) : 0ll;
  }
  __syncthreads();
  if (rowi < nrows && lane == 0) {
    long long st = s_base;
    for (int i = 0; i < warp; i++) st += s_tot[i];
    // Segments beyond the capacity are clamped by the consumers (start + count <= capacity
    // is re-checked there); the host sees the overflow through job.total.
    job.start[rowi] = (int)min(st, (long long)0x7fffffff);
    job.count[rowi] = carry;
    // heaviest-first order (lpt_row in dibr_common.cuh)
    const int bk = carry > 0 ? min(kOrderBuckets - 1, 32 - __clz(carry)) : 0;
    job.order[(long long)bk * nrows + atomicAdd(job.bucket + bk, 1)] = (int)rowi;
  }
}

// ---------------------------------------------------------------------------- host
cudaError_t launch_binning_count(const BinJobs& jobs, int njobs, const Plan& p, cudaStream_t st) {
  const size_t smem = (size_t)kBinWarps * p.NB * sizeof(int);
  if (p.nchunks > 0) {
    dim3 grid((p.nchunks + kBinWarps - 1) / kBinWarps, p.B, njobs);
    prof_begin(K_BIN_COUNT, st);
    launch_k(bin_walk_kernel<false>, grid, kBinWarps * 32, smem, st, jobs, p);
    prof_end(st);
  }
  const long long nrows = (long long)p.B * p.NB;
  dim3 grid2((unsigned)((nrows + 7) / 8), 1, njobs);
  prof_begin(K_BIN_SCAN, st);
  launch_k(bin_scan_kernel, grid2, 256, 0, st, jobs, p);
  prof_end(st);
  return cudaGetLastError();
}

cudaError_t launch_binning_fill(const BinJobs& jobs, int njobs, const Plan& p, cudaStream_t st) {
  // counters (+ lane masks of the ordered fill up to kMaskBins bins)
  const size_t smem = (size_t)(p.NB <= kMaskBins ? 2 : 1) * kBinWarps * p.NB * sizeof(int);
  dim3 grid((p.nchunks + kBinWarps - 1) / kBinWarps, p.B, njobs);
  prof_begin(K_BIN_FILL, st);
  launch_k(bin_walk_kernel<true>, grid, kBinWarps * 32, smem, st, jobs, p);
  prof_end(st);
  return cudaGetLastError();
}

}  // namespace dibr
