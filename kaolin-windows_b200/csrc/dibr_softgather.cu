// dibr_softgather.cu -- soft mask, step 1: which faces influence which pixel.
//
// For every pixel NOT covered by the rasterizer the reference keeps the first `knum` faces,
// in ascending face order, whose enlarged box contains the pixel centre
// (dibr_soft_mask_cuda.cu:80-97,165-171).  This kernel finds exactly those (pixel, face)
// pairs and appends them to the pair buffer (PairBuf, dibr_common.cuh); the arithmetic of
// the mask itself runs afterwards, pair-parallel, in softmask_pairs_eval_kernel.
//
// One CTA per screen block.  Per chunk of the bin's (ascending) face list:
//   stage      one face per thread: the EXACT pixel rectangle of its enlarged box inside
//              the block (conservative range trimmed with the reference's half-open float
//              test; pix_x/pix_y are monotone), and from it the set of 8x4-pixel regions
//              the face touches as a bit mask;
//   transpose  32 faces x 32 regions bit matrices are transposed with __ballot_sync, so
//              every region gets, per 32 staged faces, one word of candidate faces -- the
//              faces x regions pre-test costs one ballot per (face, 32 regions) instead of
//              a compare per (face, region);
//   gather     warps pull regions from a shared queue; the region's candidates are compacted
//              and handled 32 at a time, bit-parallel: each candidate lane turns its
//              rectangle into a mask over the region's 32 pixels, 32 ballots transpose these
//              into per-pixel candidate masks (bit order = face order), of which each pixel
//              keeps the first knum - held; hits are staged in shared memory and leave as
//              whole 128-byte rows.
// A region closes when all its uncovered pixels hold knum faces; the CTA stops when every
// region is closed.
#include "dibr_common.cuh"

namespace dibr {

constexpr int kGWarps = 8;
constexpr int kGThreads = kGWarps * 32;
constexpr int kGChunk = 512;            // staged faces per chunk
constexpr int kGWords = kGChunk / 32;   // candidate words per region and chunk
constexpr int kGTileK = 32;             // knum up to which a region's hits are staged in smem

template <int BLK_LOG2>
struct GatherSmem {
  static constexpr int BLK = 1 << BLK_LOG2;
  static constexpr int NPX = BLK * BLK;
  static constexpr int NR = NPX / 32;   // regions (8 wide, 4 high)
  static constexpr int RX = BLK / 8;    // regions per row
  static constexpr int NW = NR >= 32 ? NR / 32 : 1;  // 32-region words per face
  static constexpr int NRW = NR >= 32 ? 32 : NR;     // regions per word
  unsigned short cnt[NPX];              // faces held; 0xffff = covered / outside the image
  int ract[NR];                         // active-region slot in the pair buffer, -1 = none
  unsigned char rseg[NR];               // pair segments owned
  unsigned char ropen[NR];
  int face[kGChunk];
  unsigned rect[kGChunk];               // x_lo | x_hi << 8 | y_lo << 16 | y_hi << 24, block-relative
  unsigned cand[NR][kGWords];
  int tile[kGWarps][kGTileK][32];       // hits of the region a warp works on: [hit][pixel]
  unsigned short clist[kGWarps][kGChunk];  // the region's candidates of the chunk, compacted
  int next;
  unsigned item_next;                   // persistent mode: the CTA's next work item
};

// One block (b, bxi, byi) of 2^BLK_LOG2 pixels, by the whole CTA.  prefilled: the rasterizer
// of the same call has written soft_mask's base value (1 covered / 0 uncovered) and only
// hands over blocks that hold uncovered pixels.
template <int BLK_LOG2>
__device__ __forceinline__ void gather_block(const SoftArgs& a, const Plan& p,
                                             GatherSmem<BLK_LOG2>& sm, int b, int bxi, int byi,
                                             bool prefilled, unsigned trace_slot) {
  using Smem = GatherSmem<BLK_LOG2>;
  constexpr int BLK = Smem::BLK, NPX = Smem::NPX, NR = Smem::NR, RX = Smem::RX, NW = Smem::NW;
  constexpr int NRW = Smem::NRW;
  const PairBuf& pb = a.pairs;
  const int tid = threadIdx.x, lane = lane_id(), warp = tid >> 5;
#ifdef DIBR_TRACE_GATHER
  DIBR_TRACE_BEGIN();
  long long ph_t = clock64(), ph_init = 0, ph_stage = 0, ph_gather = 0;   // thread 0's wall clock
#define GATHER_PHASE(acc) do { const long long now_ = clock64(); acc += now_ - ph_t; ph_t = now_; } while (0)
#else
#define GATHER_PHASE(acc)
#endif
  const int bx0 = bxi << BLK_LOG2, by0 = byi << BLK_LOG2;
  const int bx1 = min(bx0 + BLK, p.W) - 1, by1 = min(by0 + BLK, p.H) - 1;
  const Grid g = plan_grid(p, a.job.mult);
  const int K = min(a.K, 0xfffe);
  const int lx = lane & 7, ly = lane >> 3;
  const int nrx = (p.W + 7) >> 3, nry = (p.H + 3) >> 2;  // regions of a view

  // a raster block without a single uncovered pixel (the rasterizer of the same call counted
  // them): every pixel of this CTA is covered, soft_mask = 1, nothing to gather and no need
  // to read selected_face_idx
  if (!prefilled && a.block_uncovered &&
      a.block_uncovered[(long long)b * p.nkx * p.nky + (by0 >> p.blk_log2) * p.nkx +
                        (bx0 >> p.blk_log2)] == 0) {
    for (int i = tid; i < NPX; i += kGThreads) {
      const int ix = bx0 + (i & (BLK - 1)), iy = by0 + (i >> BLK_LOG2);
      if (ix < p.W && iy < p.H) a.mask[((long long)b * p.H + iy) * p.W + ix] = 1.0f;
    }
#ifdef DIBR_TRACE_GATHER
    DIBR_TRACE_END_AT(trace_slot, -1);
#endif
    return;
  }
  // ---- pixel state; covered pixels get their final value right away
  for (int i = tid; i < NPX; i += kGThreads) {
    const int ix = bx0 + (i & (BLK - 1)), iy = by0 + (i >> BLK_LOG2);
    bool uncovered = false;
    if (ix < p.W && iy < p.H) {
      const long long pix = ((long long)b * p.H + iy) * p.W + ix;
      uncovered = a.sel[pix] < 0;
      // 1 for covered pixels (:68-70); 1 - 1 = 0 for uncovered ones until the pair
      // evaluation overwrites those that hold faces
      if (!prefilled) a.mask[pix] = uncovered ? 0.0f : 1.0f;
    }
    sm.cnt[i] = uncovered ? 0 : 0xffff;
  }
  __syncthreads();
  bool any_open = false;
  for (int r = warp; r < NR; r += kGWarps) {
    const int pl = (((r / RX) * 4 + ly) << BLK_LOG2) + (r % RX) * 8 + lx;
    const unsigned m = __ballot_sync(DIBR_FULL_MASK, (int)sm.cnt[pl] < K);
    if (lane == 0) {
      sm.ropen[r] = m != 0;
      sm.rseg[r] = 0;
      sm.ract[r] = -1;
    }
    any_open |= m != 0;
  }
  if (!__syncthreads_or(any_open)) {
#ifdef DIBR_TRACE_GATHER
    DIBR_TRACE_END_AT(trace_slot, 0);
#endif
    return;
  }

  // the gather CTA covers a quarter (or less) of a raster block: more, shorter CTAs over
  // the few blocks the silhouette crosses; each re-stages its parent bin's list
  const int bin = (by0 >> p.bin_log2) * p.nbx + (bx0 >> p.bin_log2);
  const long long row = (long long)b * p.NB + bin;
  const long long start = a.job.start[row];
  long long n = a.job.count[row];
  if (start + n > a.job.capacity) n = max(0ll, a.job.capacity - start);

  // The list entries of the NEXT chunk are loaded (coalesced, independent of everything) when
  // a chunk starts and stay in flight through its stage and gather phases; within a chunk all
  // rectangle loads (the dependent gather through the face id) are issued before the first
  // is used.  Leaves one exposed global round trip per chunk instead of four.
  constexpr int NIT = kGChunk / kGThreads;
  int fpre[NIT];
#pragma unroll
  for (int k = 0; k < NIT; k++) {
    const long long e = (long long)k * kGThreads + tid;
    fpre[k] = e < n ? a.job.list[start + e] : -1;
  }
  GATHER_PHASE(ph_init);
  for (long long base = 0; base < n; base += kGChunk) {
    __syncthreads();
    GATHER_PHASE(ph_gather);
    bool mine = false;
    for (int r = tid; r < NR; r += kGThreads) mine |= sm.ropen[r] != 0;
    if (tid == 0) sm.next = 0;
    if (!__syncthreads_or(mine)) break;

    // ---- stage + transpose (t / 32 is warp-uniform)
    // the face's exact pixel rectangle (the reference's half-open test, :89-97), kept by the
    // binning pass: 8 bytes instead of the corner loads and the float geometry
    int fcur[NIT];
    ushort4 rcur[NIT];
#pragma unroll
    for (int k = 0; k < NIT; k++) {
      fcur[k] = fpre[k];
      rcur[k] = fcur[k] >= 0 ? __ldg(a.job.rect + (long long)b * p.F + fcur[k])
                             : make_ushort4(1, 0, 1, 0);
    }
#pragma unroll
    for (int k = 0; k < NIT; k++) {
      const long long e = base + kGChunk + (long long)k * kGThreads + tid;
      fpre[k] = e < n ? a.job.list[start + e] : -1;
    }
#pragma unroll
    for (int k = 0; k < NIT; k++) {
      const int t = k * kGThreads + tid;
      unsigned rmask[NW];
#pragma unroll
      for (int w = 0; w < NW; w++) rmask[w] = 0u;
      if (fcur[k] >= 0) {
        const int f = fcur[k];
        const ushort4 r = rcur[k];
        int x_lo = r.x, x_hi = r.y, y_lo = r.z, y_hi = r.w;
        {
          x_lo = max(x_lo, bx0); x_hi = min(x_hi, bx1);
          y_lo = max(y_lo, by0); y_hi = min(y_hi, by1);
          if (x_lo <= x_hi && y_lo <= y_hi) {
            const int xl = x_lo - bx0, xh = x_hi - bx0, yl = y_lo - by0, yh = y_hi - by0;
            sm.rect[t] = (unsigned)xl | ((unsigned)xh << 8) | ((unsigned)yl << 16) |
                         ((unsigned)yh << 24);
            sm.face[t] = f;
            const unsigned cols = (2u << (xh >> 3)) - (1u << (xl >> 3));  // RX bits
            for (int ry = yl >> 2; ry <= (yh >> 2); ry++) {
              const int bit = ry * RX;  // RX divides 32: a row of regions never straddles words
#pragma unroll
              for (int w = 0; w < NW; w++)
                if ((bit >> 5) == w) rmask[w] |= cols << (bit & 31);
            }
          }
        }
      }
      const int grp = t >> 5;
#pragma unroll
      for (int w = 0; w < NW; w++) {
        unsigned keep = 0u;
        if (__any_sync(DIBR_FULL_MASK, rmask[w] != 0u)) {
#pragma unroll
          for (int r = 0; r < NRW; r++) {
            const unsigned v = __ballot_sync(DIBR_FULL_MASK, (rmask[w] >> r) & 1u);
            if (lane == r) keep = v;
          }
        }
        if (lane < NRW) sm.cand[w * 32 + lane][grp] = keep;
      }
    }
    __syncthreads();
    GATHER_PHASE(ph_stage);

    // ---- gather
    while (true) {
      int r = 0;
      if (lane == 0) r = atomicAdd(&sm.next, 1);
      r = __shfl_sync(DIBR_FULL_MASK, r, 0);
      if (r >= NR) break;
      if (!sm.ropen[r]) continue;
      const unsigned myw = lane < kGWords ? sm.cand[r][lane] : 0u;
      const unsigned gmask = __ballot_sync(DIBR_FULL_MASK, myw != 0u);
      if (!gmask) continue;
      const int px = (r % RX) * 8 + lx, py = (r / RX) * 4 + ly;  // block-relative
      const int pl = (py << BLK_LOG2) + px;
      int cnt = sm.cnt[pl];
      bool open = cnt < K;
      // Hits are staged in the warp's shared-memory tile (bank = lane: conflict-free) and
      // written out row by row when the item ends -- one 128-byte line per row instead of a
      // 32-line scatter per candidate face, which made the kernel LSU-bound.
      const bool staged = K <= kGTileK;
      const int cnt0 = cnt;
      int (*tile)[32] = sm.tile[warp];
      // segment ids the region owns: lane q holds segment q (regions with more than 32
      // segments, knum > 256, read the rest from the table in global memory)
      int aidx = sm.ract[r];
      int have = sm.rseg[r];
      int segreg = -1;
      if (aidx >= 0 && aidx < pb.cap && lane < have)
        segreg = pb.act_seg[(long long)aidx * pb.nseg + lane];

      // Makes sure the region owns segments for `upto` hits per pixel (allocated on demand;
      // small knum takes all its segments with the first hit, so a region pays the cursor
      // round trip once).
      auto ensure = [&](int upto) {
        int need = (min(upto, K) + kPairSegHits - 1) / kPairSegHits;
        if (need <= have) return;
        need = pb.nseg <= 4 ? pb.nseg : min(pb.nseg, max(need, 2 * have));
        long long sbase = 0;
        if (lane == 0) {
          if (aidx < 0) {
            const unsigned long long v = atomicAdd(pb.cursor + 1, 1ull);
            aidx = v < (unsigned long long)pb.cap ? (int)v : 0x7fffffff;
            if (aidx < pb.cap)
              pb.act_region[aidx] = (b * nry + (by0 >> 2) + r / RX) * nrx + (bx0 >> 3) + r % RX;
          }
          sbase = (long long)atomicAdd(pb.cursor, (unsigned long long)(need - have));
        }
        aidx = __shfl_sync(DIBR_FULL_MASK, aidx, 0);
        sbase = __shfl_sync(DIBR_FULL_MASK, sbase, 0);
        for (int q = have + lane; q < need; q += 32) {
          const long long sg = sbase + (q - have);
          if (aidx < pb.cap) pb.act_seg[(long long)aidx * pb.nseg + q] = sg < pb.cap ? (int)sg : -1;
        }
        if (lane >= have && lane < need) {
          const long long sg = sbase + (lane - have);
          segreg = sg < pb.cap ? (int)sg : -1;
        }
        if (lane == 0) { sm.ract[r] = aidx; sm.rseg[r] = (unsigned char)need; }
        have = need;
      };

      // ---- compact the region's candidates of this chunk (ascending face order)
      const int pc = __popc(myw);
      int incl = pc;
#pragma unroll
      for (int d = 1; d < kGWords; d <<= 1) {
        const int t = __shfl_up_sync(DIBR_FULL_MASK, incl, d);
        if (lane >= d) incl += t;
      }
      const int T = __shfl_sync(DIBR_FULL_MASK, incl, kGWords - 1);
      unsigned short* cl = sm.clist[warp];
      {
        unsigned w = myw;
        int o = incl - pc;
        while (w) {
          cl[o++] = (unsigned short)((lane << 5) + __ffs(w) - 1);
          w &= w - 1;
        }
      }
      __syncwarp();

      // ---- 32 candidates per step, bit-parallel: every candidate lane builds the mask of
      // the region's pixels inside its rectangle; 32 ballots transpose that into, per pixel
      // lane, the mask of candidates containing it (bit order = face order)
      const int rx0 = (r % RX) * 8, ry0 = (r / RX) * 4;
      for (int cb = 0; cb < T; cb += 32) {
        if (!__any_sync(DIBR_FULL_MASK, open)) break;
        const bool cv = cb + lane < T;
        const int s = cv ? cl[cb + lane] : 0;
        const unsigned rc = sm.rect[s];
        const int fid = sm.face[s];
        unsigned pixmask = 0u;
        if (cv) {
          const int c0 = max((int)(rc & 0xff) - rx0, 0), c1 = min((int)((rc >> 8) & 0xff) - rx0, 7);
          const int r0 = max((int)((rc >> 16) & 0xff) - ry0, 0), r1 = min((int)(rc >> 24) - ry0, 3);
          if (c0 <= c1 && r0 <= r1) {
            const unsigned cols = ((2u << c1) - (1u << c0)) & 0xffu;
            const unsigned rows = ((2u << r1) - (1u << r0)) & 0xfu;
            pixmask = (cols * 0x01010101u) & ((((rows * 0x00204081u) & 0x01010101u)) * 0xffu);
          }
        }
        unsigned inmask = 0u;
#pragma unroll
        for (int q = 0; q < 32; q++) {
          const unsigned v = __ballot_sync(DIBR_FULL_MASK, (pixmask >> q) & 1u);
          if (lane == q) inmask = v;
        }
        if (!open) inmask = 0u;
        const int room = K - cnt;           // > 0 for open pixels
        for (int drop = open ? __popc(inmask) - room : 0; drop > 0; drop--)  // first `room` (:170-171)
          inmask &= ~(0x80000000u >> __clz(inmask));
        const int nh = __popc(inmask);
        const int mxh = __reduce_max_sync(DIBR_FULL_MASK, nh);
        if (mxh == 0) continue;
        ensure(__reduce_max_sync(DIBR_FULL_MASK, open ? cnt + nh : 0));
        for (int k = 0; k < mxh; k++) {
          const bool has = inmask != 0u;
          const int c = has ? __ffs(inmask) - 1 : 0;
          inmask &= inmask - 1u;
          const int f = __shfl_sync(DIBR_FULL_MASK, fid, c);
          const int q = cnt / kPairSegHits;
          int sg = __shfl_sync(DIBR_FULL_MASK, segreg, q & 31);
          if (has) {
            if (staged) {
              tile[cnt][lane] = f;
            } else {
              if (q >= 32 && aidx < pb.cap) sg = pb.act_seg[(long long)aidx * pb.nseg + q];
              if (aidx < pb.cap && sg >= 0)
                pb.face[(long long)sg * kPairSegSlots + (cnt % kPairSegHits) * 32 + lane] = f;
            }
            cnt++;
          }
        }
        open = cnt < K;
      }
      __syncwarp();
      if (staged && aidx >= 0 && aidx < pb.cap) {
        const bool got = cnt != 0xffff && cnt > cnt0;
        const int h0 = __reduce_min_sync(DIBR_FULL_MASK, got ? cnt0 : 0x7fff);
        const int h1 = __reduce_max_sync(DIBR_FULL_MASK, got ? cnt : 0);
        for (int h = h0; h < h1; h++) {
          const int sg = __shfl_sync(DIBR_FULL_MASK, segreg, h / kPairSegHits);
          if (got && h >= cnt0 && h < cnt && sg >= 0)
            pb.face[(long long)sg * kPairSegSlots + (h % kPairSegHits) * 32 + lane] = tile[h][lane];
        }
      }
      sm.cnt[pl] = (unsigned short)cnt;
      const unsigned still = __ballot_sync(DIBR_FULL_MASK, open);
      if (lane == 0) sm.ropen[r] = still != 0;
    }
  }

  // ---- per-pixel hit counts of the regions that became active
  __syncthreads();
  GATHER_PHASE(ph_gather);
  for (int r = warp; r < NR; r += kGWarps) {
    const int aidx = sm.ract[r];
    if (aidx >= 0 && aidx < pb.cap) {
      const int c = sm.cnt[(((r / RX) * 4 + ly) << BLK_LOG2) + (r % RX) * 8 + lx];
      const int held = c == 0xffff ? 0 : c;
      pb.act_cnt[(long long)aidx * 32 + lane] = (unsigned short)held;
      // work items of the evaluation / gradient kernels: the region's pairs in chunks
      const int total = __reduce_add_sync(DIBR_FULL_MASK, held);
      const int nch = (total + kPairChunk - 1) / kPairChunk;
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(pb.nchunk, (unsigned)nch);
      base = __shfl_sync(DIBR_FULL_MASK, base, 0);
      for (int i = lane; i < nch; i += 32)
        if (base + i < 3u * (unsigned)pb.cap) pb.chunk[base + i] = aidx * 256 + i;
    }
  }
#ifdef DIBR_TRACE_GATHER
  {
    __syncthreads();
    const long long ph_tail = clock64() - ph_t;
    DIBR_TRACE_END_AT(trace_slot, ((unsigned long long)min(ph_init >> 6, 16383ll) << 42) |
                   ((unsigned long long)min(ph_stage >> 6, 16383ll) << 28) |
                   ((unsigned long long)min(ph_gather >> 6, 16383ll) << 14) |
                   (unsigned long long)min(ph_tail >> 6, 16383ll));
  }
#endif
}

#undef GATHER_PHASE

constexpr int kGPersistentCtas = 148 * 4;   // 4 CTAs per SM fit (shared memory)

template <int BLK_LOG2>
__global__ void __launch_bounds__(kGThreads)
softmask_gather_kernel(const SoftArgs a, const Plan p) {
  using Smem = GatherSmem<BLK_LOG2>;
  extern __shared__ __align__(16) unsigned char s_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(s_raw);
  pdl_prologue();
  if (a.work.item) {
    // Persistent CTAs over the rasterizer's work list: only blocks that hold uncovered
    // pixels and have a soft list (the silhouette band) reach this kernel, and all of them
    // start at once instead of queueing behind thousands of CTAs that leave immediately.
    // First item = CTA id; further ones from a queue cursor, fetched while the current
    // block is being worked on (the result is only read afterwards).
    const unsigned nwork = min(*a.work.count, a.work.cap);
    unsigned item = blockIdx.x;
    while (item < nwork) {
      unsigned fetched = 0;
      if (threadIdx.x == 0) fetched = atomicAdd(a.work.next, 1u);
      const int2 wk = a.work.item[item];
      gather_block<BLK_LOG2>(a, p, sm, wk.x, wk.y & 0xffff, wk.y >> 16, true, item);
      __syncthreads();   // the block's shared state is free; item_next is not being read
      if (threadIdx.x == 0) sm.item_next = gridDim.x + fetched;
      __syncthreads();
      item = sm.item_next;
    }
    return;
  }
  // CTAs are taken longest face list first, as the rasterizer's (lpt_block): the few
  // silhouette CTAs that walk a long list start early, the many that leave at once fill in
  int b, bxi, byi;
  if (!lpt_block(a.job, p, BLK_LOG2, blockIdx.x, b, bxi, byi)) return;
  gather_block<BLK_LOG2>(a, p, sm, b, bxi, byi, false, blockIdx.x);
}

template <int G_LOG2>
static cudaError_t launch_gather(const SoftArgs& a, const Plan& p, cudaStream_t st) {
  const int smem = (int)sizeof(GatherSmem<G_LOG2>);
  cudaError_t e = smem_optin((const void*)softmask_gather_kernel<G_LOG2>, smem);
  if (e != cudaSuccess) return e;
  const int shift = p.bin_log2 - G_LOG2;
  long long nblk = ((long long)p.B * p.NB) << (2 * shift);
  if (a.work.item && nblk > kGPersistentCtas) nblk = kGPersistentCtas;
  dim3 grid((unsigned)nblk);
  launch_k(softmask_gather_kernel<G_LOG2>, grid, kGThreads, smem, st, a, p);
  return cudaGetLastError();
}

cudaError_t launch_softmask_gather(const SoftArgs& a, const Plan& p, cudaStream_t st) {
  prof_begin(K_SOFT_GATHER, st);
  // a quarter (or less) of a raster block per CTA
  cudaError_t e = p.blk_log2 == 5 ? launch_gather<4>(a, p, st) : launch_gather<5>(a, p, st);
  prof_end(st);
  return e;
}

}  // namespace dibr
