// dibr_api.cu -- C ABI of libdibr_b200.so (see include/dibr_b200.h).
// Plans the screen hierarchy, lays out the caller's workspace and sequences the
// kernels of dibr_binning.cu / dibr_raster.cu / dibr_softmask.cu on the caller's stream.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/dibr_b200.h"
#include "dibr_common.cuh"

namespace dibr {

static thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define DIBR_CUDA(expr)                                                              \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess)                                                           \
      return fail(DIBR_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));      \
  } while (0)

// ------------------------------------------------------------- per-kernel profiling
struct ProfRec { int id; cudaEvent_t e0, e1; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::mutex g_prof_mu;

void prof_begin(int id, cudaStream_t st) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r;
  r.id = id;
  if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return;
  cudaEventRecord(r.e0, st);
  g_prof.push_back(r);
}
void prof_end(cudaStream_t st) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_prof.empty()) cudaEventRecord(g_prof.back().e1, st);
}

#ifdef DIBR_TRACE
static unsigned long long* g_trace_ptr = nullptr;
#endif

// programmatic dependent launch between the kernels of the path (dibr_common.cuh):
// measured to gain nothing under graph replay, so off unless DIBR_B200_PDL=1
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("DIBR_B200_PDL");
    return e && e[0] == '1';
  }();
  return on;
}

cudaError_t smem_optin(const void* kernel, int bytes) {
  struct Done { const void* fn; int dev; int bytes; };
  static std::mutex mu;
  static std::vector<Done> done;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lk(mu);
  for (const Done& d : done)
    if (d.fn == kernel && d.dev == dev && d.bytes >= bytes) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) done.push_back({kernel, dev, bytes});
  return e;
}

// needed_host is device-visible pinned memory (cudaHostAlloc under UVA): the kernels store
// the report themselves, no copy node between them (dibr_host_report_direct)
static bool g_report_direct = false;   // process-wide, set once by the host binding


static inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

static int make_plan(int B, int F, int H, int W, Plan& p, float mult = nanf("")) {
  if (B < 0 || F < 0 || H <= 0 || W <= 0)
    return fail(DIBR_E_INVALID, "invalid sizes B=%d F=%d H=%d W=%d", B, F, H, W);
  if (B > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", B);
  if (H > 32768 || W > 32768) return fail(DIBR_E_INVALID, "image %dx%d too large", H, W);
  memset(&p, 0, sizeof(p));
  p.B = B; p.F = F; p.H = H; p.W = W;
  p.grid_mult = mult;
  if (mult == mult) p.grid = make_grid(mult, W, H);
  int k = 5;
  while ((int64_t)((W + (1 << k) - 1) >> k) * ((H + (1 << k) - 1) >> k) > 1024) k++;
  p.bin_log2 = k;
  p.nbx = (W + (1 << k) - 1) >> k;
  p.nby = (H + (1 << k) - 1) >> k;
  p.NB = p.nbx * p.nby;
  p.blk_log2 = k < 6 ? k : 6;
  p.nkx = (W + (1 << p.blk_log2) - 1) >> p.blk_log2;
  p.nky = (H + (1 << p.blk_log2) - 1) >> p.blk_log2;
  // one warp per chunk of S faces: aim at ~8k warps, keep the count matrix moderate
  int64_t s = ((int64_t)B * F) / 8192;
  if (s < 128) s = 128;
  if (s > 4096) s = 4096;
  p.S = (int)align_up(s, 32);
  p.nchunks = F > 0 ? (F + p.S - 1) / p.S : 0;
  p.mstride = (int)align_up(p.nchunks, 4);
  int64_t off = 0;
  p.off_total = off; off += 256;
  p.off_flags = off; off = align_up(off + (int64_t)B * p.nkx * p.nky * 4, 256);
  for (int l = 0; l < 2; l++) {
    p.off_start[l] = off; off = align_up(off + (int64_t)B * p.NB * 4, 256);
    p.off_count[l] = off; off = align_up(off + (int64_t)B * p.NB * 4, 256);
    p.off_range[l] = off; off = align_up(off + (int64_t)B * p.nchunks * 8, 256);
    p.off_mat[l] = off; off = align_up(off + (int64_t)B * p.NB * p.mstride * 4, 256);
  }
  for (int l = 0; l < 2; l++) {  // not memset: only the bucket sizes in the header are
    p.off_order[l] = off; off = align_up(off + (int64_t)kOrderBuckets * B * p.NB * 4, 256);
  }
  p.off_rect = off; off = align_up(off + (int64_t)B * F * 8, 256);   // (never memset)
  p.off_work = off; off = align_up(off + (int64_t)B * p.nkx * p.nky * 4 * 8, 256);   // (never memset)
  p.fixed_bytes = off;
#ifdef DIBR_TRACE
  p.trace = g_trace_ptr;
#endif
  return DIBR_OK;
}

struct Workspace {
  char* base;
  int64_t cap[2];
  int* list[2];
  float4* rec;     // the raster list area as records (same memory as list[0]) or nullptr
  PairBuf pairs;
};

// The rasterizer's list as 48-byte face records streamed with bulk async copies (default)
// or as face ids gathered through the face tensors (DIBR_B200_RASTER_REC=0: the A/B
// switch of the experiment recorded in DESIGN.md).  Process-wide, read once.
static bool raster_rec_enabled() {
  static const bool on = [] {
    const char* e = getenv("DIBR_B200_RASTER_REC");
    return !(e && e[0] == '0');
  }();
  return on;
}
static int64_t raster_entry_bytes() { return raster_rec_enabled() ? kRecBytes : 4; }
static int64_t lists_bytes(int64_t cap_r, int64_t cap_s) {
  return align_up(cap_r * raster_entry_bytes(), 256) + align_up(cap_s * 4, 256);
}

// Pair area (compact soft-mask face lists kept for the backward) for `cap` segments.
static int pair_nseg(int K) { return (K + kPairSegHits - 1) / kPairSegHits; }
static bool pair_usable(int64_t cap, int K) { return cap > 0 && K > 0 && K <= kPairMaxK; }
static int64_t pair_bytes(int64_t cap, int K) {
  if (!pair_usable(cap, K)) return 0;
  const int64_t n = cap * kPairSegSlots;
  return align_up(cap * 4, 256) + align_up(cap * 64, 256) + align_up(cap * pair_nseg(K) * 4, 256) +
         2 * align_up(n * 4, 256) + align_up(n, 256) + align_up(cap * 12, 256);
}

static int bind_workspace(const Plan& p, const void* ws, int64_t ws_bytes, int64_t cap_r,
                          int64_t cap_s, Workspace& w, int64_t cap_p = 0, int K = 0) {
  if (cap_r < 0) cap_r = 0;
  if (cap_s < 0) cap_s = 0;
  if (cap_r > kMaxListEntries || cap_s > kMaxListEntries)
    return fail(DIBR_E_INVALID, "list capacity above 2^31 - 2^16 entries");
  if (cap_p > (1 << 23)) return fail(DIBR_E_INVALID, "pair capacity above 2^23 segments");
  const int64_t lists_end = p.fixed_bytes + lists_bytes(cap_r, cap_s);
  const int64_t need = lists_end + pair_bytes(cap_p, K);
  if (!ws || ws_bytes < need)
    return fail(DIBR_E_WORKSPACE, "workspace of %lld bytes given, %lld needed",
                (long long)ws_bytes, (long long)need);
  if (((uintptr_t)ws & 255) != 0) return fail(DIBR_E_INVALID, "workspace must be 256-byte aligned");
  w.base = (char*)ws;
  w.cap[0] = cap_r;
  w.cap[1] = cap_s;
  w.list[0] = (int*)(w.base + p.fixed_bytes);
  w.list[1] = (int*)(w.base + p.fixed_bytes + align_up(cap_r * raster_entry_bytes(), 256));
  w.rec = raster_rec_enabled() ? (float4*)w.list[0] : nullptr;
  memset(&w.pairs, 0, sizeof(w.pairs));
  if (pair_usable(cap_p, K)) {
    PairBuf& pb = w.pairs;
    const int64_t n = cap_p * kPairSegSlots;
    char* q = w.base + lists_end;
    pb.cap = (int)cap_p;
    pb.nseg = pair_nseg(K);
    pb.cursor = (unsigned long long*)(w.base + p.off_total) + 2;
    pb.act_region = (int*)q; q += align_up(cap_p * 4, 256);
    pb.act_cnt = (unsigned short*)q; q += align_up(cap_p * 64, 256);
    pb.act_seg = (int*)q; q += align_up(cap_p * pb.nseg * 4, 256);
    pb.face = (int*)q; q += align_up(n * 4, 256);
    pb.prob = (float*)q; q += align_up(n * 4, 256);
    pb.type = (unsigned char*)q; q += align_up(n, 256);
    pb.chunk = (int*)q;
    pb.nchunk = (unsigned int*)(w.base + p.off_total + 192);
  }
  return DIBR_OK;
}

static void attach(BinJob& j, const Plan& p, const Workspace& w, int l) {
  j.mat = (int*)(w.base + p.off_mat[l]);
  j.range = (int2*)(w.base + p.off_range[l]);
  j.start = (int*)(w.base + p.off_start[l]);
  j.count = (int*)(w.base + p.off_count[l]);
  j.list = w.list[l];
  j.order = (int*)(w.base + p.off_order[l]);
  j.bucket = (int*)(w.base + p.off_total + 64) + l * kOrderBuckets;
  j.total = (unsigned long long*)(w.base + p.off_total) + l;
  j.capacity = w.cap[l];
  j.ordered = l == 1;
  j.rect = l == 1 ? (ushort4*)(w.base + p.off_rect) : nullptr;
}

// DIBR_B200_GATHER_WORK=0: the gather takes every block of the screen in list-length order,
// as the stand-alone soft mask does (A/B switch)
static bool gather_handoff_enabled() {
  static const bool on = [] {
    const char* e = getenv("DIBR_B200_GATHER_WORK");
    return !(e && e[0] == '0');
  }();
  return on;
}

// Zero-fill of the two gradient outputs of a backward call: one memset when the second
// starts exactly where the first ends (the Python graph entries allocate them that way: one
// node less in the launch chain), else two.  Exact adjacency only -- no byte that is not part
// of either output is ever written.
static cudaError_t zero_gradients(float* g0, size_t bytes0, float* g1, size_t bytes1,
                                  cudaStream_t st) {
  if (g1 && (char*)g1 == (char*)g0 + bytes0) return cudaMemsetAsync(g0, 0, bytes0 + bytes1, st);
  cudaError_t e = cudaMemsetAsync(g0, 0, bytes0, st);
  if (e == cudaSuccess && g1) e = cudaMemsetAsync(g1, 0, bytes1, st);
  return e;
}

// gather work list of the fused forward (rasterizer -> gather); header words 200 / 204
static GatherWork gather_work(const Plan& p, const Workspace& w) {
  GatherWork g;
  g.item = (int2*)(w.base + p.off_work);
  g.count = (unsigned*)(w.base + p.off_total + 200);
  g.next = (unsigned*)(w.base + p.off_total + 204);
  g.cap = (unsigned)((int64_t)p.B * p.nkx * p.nky * 4);
  return g;
}

// memset + count + scan + (capacity check) + fill for the given jobs (kinds[i] = 0/1).
static int run_binning(const Plan& p, const Workspace& w, BinJobs& jobs, const int* kinds,
                       int njobs, int64_t* needed_host, cudaStream_t st) {
  if (njobs == 2) {
    // header + block flags + both lists' start / count / range / count matrix are one
    // contiguous range: one memset node instead of three in the launch chain
    DIBR_CUDA(cudaMemsetAsync(w.base, 0, p.off_order[0], st));
  } else {
    DIBR_CUDA(cudaMemsetAsync(w.base, 0, p.off_start[0], st));  // totals + block flags
    for (int i = 0; i < njobs; i++) {
      const int l = kinds[i];
      DIBR_CUDA(cudaMemsetAsync(w.base + p.off_start[l], 0,
                                (l == 0 ? p.off_start[1] : p.off_order[0]) - p.off_start[l], st));
    }
  }
  const bool direct = needed_host && g_report_direct && p.B > 0 && p.F > 0;
  jobs.totals = (const unsigned long long*)(w.base + p.off_total);
  jobs.report = direct ? (unsigned long long*)needed_host : nullptr;
  if (p.B > 0) DIBR_CUDA(launch_binning_count(jobs, njobs, p, st));
  if (direct) {
    // the fill kernel stores the totals into the host slot itself
  } else if (needed_host) {
    DIBR_CUDA(cudaMemcpyAsync(needed_host, w.base + p.off_total, 16, cudaMemcpyDeviceToHost, st));
  } else {
    unsigned long long tot[2] = {0, 0};
    DIBR_CUDA(cudaMemcpyAsync(tot, w.base + p.off_total, 16, cudaMemcpyDeviceToHost, st));
    DIBR_CUDA(cudaStreamSynchronize(st));
    if ((int64_t)tot[0] > w.cap[0] || (int64_t)tot[1] > w.cap[1])
      return fail(DIBR_E_WORKSPACE, "face lists need %llu raster / %llu soft entries, "
                  "capacity %lld / %lld", tot[0], tot[1], (long long)w.cap[0], (long long)w.cap[1]);
  }
  if (p.B > 0 && p.F > 0) DIBR_CUDA(launch_binning_fill(jobs, njobs, p, st));
  return DIBR_OK;
}

static BinJob face_source(const float* img, const float* bbox, const uint8_t* valid,
                          const float* normal_z, const int64_t* first, long long view_stride,
                          float mult, float margin, int prescaled) {
  BinJob j;
  memset(&j, 0, sizeof(j));
  j.img = img; j.bbox = bbox; j.valid = valid; j.normal_z = normal_z; j.first = first;
  j.view_stride = view_stride; j.mult = mult; j.margin = margin; j.prescaled = prescaled;
  return j;
}

}  // namespace dibr

using namespace dibr;


#ifdef DIBR_TRACE
extern "C" __attribute__((visibility("default"))) int dibr_trace_set(void* ptr) {
  dibr::g_trace_ptr = (unsigned long long*)ptr;
  return 0;
}
#endif

extern "C" const char* dibr_last_error(void) { return g_err; }
extern "C" int dibr_version(void) { return 100; }

extern "C" int dibr_host_report_direct(int on) {
  const int was = g_report_direct ? 1 : 0;
  g_report_direct = on != 0;
  return was;
}

extern "C" int dibr_profile_begin(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& r : g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  g_prof.clear();
  g_prof_on = true;
  return DIBR_OK;
}

extern "C" int dibr_profile_end(int* kernel_ids, float* ms, int capacity) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = false;
  int n = 0;
  for (auto& r : g_prof) {
    float t = 0.f;
    if (cudaEventSynchronize(r.e1) == cudaSuccess &&
        cudaEventElapsedTime(&t, r.e0, r.e1) == cudaSuccess && n < capacity) {
      if (kernel_ids) kernel_ids[n] = r.id;
      if (ms) ms[n] = t;
      n++;
    }
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  g_prof.clear();
  return n;
}

extern "C" int64_t dibr_workspace_fixed_bytes(int B, int F, int H, int W) {
  Plan p;
  int rc = make_plan(B, F, H, W, p);
  return rc ? rc : p.fixed_bytes;
}

extern "C" int64_t dibr_workspace_bytes_for(int B, int F, int H, int W, int64_t cap_r,
                                            int64_t cap_s, int64_t cap_p, int K) {
  Plan p;
  int rc = make_plan(B, F, H, W, p);
  if (rc) return rc;
  return p.fixed_bytes + lists_bytes(cap_r > 0 ? cap_r : 0, cap_s > 0 ? cap_s : 0) +
         pair_bytes(cap_p, K);
}

static int check_common(float multiplier) {
  if (!(multiplier > 0.f) || !isfinite(multiplier))
    return fail(DIBR_E_INVALID, "multiplier must be a positive finite number");
  return DIBR_OK;
}

extern "C" int dibr_rasterize_fwd(int B, int F, int H, int W, int D, const float* z,
                                  const float* img, const float* feat, const uint8_t* valid,
                                  const float* normal_z, int prescaled, const float* bboxes,
                                  const int64_t* first, float multiplier, float eps,
                                  int64_t* out_idx, float* out_feat, float* out_w, void* ws,
                                  int64_t ws_bytes, int64_t cap_r, int64_t* needed_host,
                                  void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  int rc = make_plan(B, F, H, W, p, multiplier);
  if (rc) return rc;
  if ((rc = check_common(multiplier))) return rc;
  if (D < 0) return fail(DIBR_E_INVALID, "feat_dim < 0");
  if (B == 0) return DIBR_OK;
  if (!out_idx || (!out_feat && D > 0)) return fail(DIBR_E_INVALID, "null output");
  if (F > 0 && (!z || !img || (!feat && D > 0))) return fail(DIBR_E_INVALID, "null input");
  if (valid && normal_z) return fail(DIBR_E_INVALID, "valid_faces and face_normals_z both given");
  if (prescaled && (!bboxes || !first))
    return fail(DIBR_E_INVALID, "prescaled mode needs face_bboxes and first_idx_face_per_mesh");
  if (!prescaled && (bboxes || first))
    return fail(DIBR_E_INVALID, "face_bboxes / first_idx only valid in prescaled mode");
  Workspace w;
  if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, 0, w))) return rc;
  BinJobs jobs;
  jobs.j[0] = face_source(img, bboxes, valid, normal_z, first, prescaled ? 0 : F, multiplier,
                          0.f, prescaled);
  attach(jobs.j[0], p, w, 0);
  if (!prescaled) { jobs.j[0].rec = w.rec; jobs.j[0].z = z; }
  jobs.j[1] = jobs.j[0];
  const int kinds[1] = {0};
  if ((rc = run_binning(p, w, jobs, kinds, 1, needed_host, st))) return rc;
  RasterArgs a;
  memset(&a, 0, sizeof(a));
  a.job = jobs.j[0]; a.rec = jobs.j[0].rec; a.z = z; a.feat = feat; a.D = D; a.eps_d = (double)eps;
  a.out_idx = out_idx; a.out_feat = out_feat; a.out_w = out_w;
  a.block_uncovered = (int*)(w.base + p.off_flags);
  DIBR_CUDA(launch_raster_forward(a, p, st));
  return DIBR_OK;
}

extern "C" int dibr_rasterize_bwd(int B, int F, int H, int W, int D, const float* grad,
                                  const int64_t* idx, const float* weights, const float* img,
                                  const float* feat, float multiplier, float eps, float* grad_img,
                                  float* grad_feat, void* stream) {
  if (B < 0 || F < 0 || H <= 0 || W <= 0 || D < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  int rc = check_common(multiplier);
  if (rc) return rc;
  if (B == 0 || F == 0 || D == 0) return DIBR_OK;
  if (B > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", B);
  if (!grad || !idx || !img || !feat || !grad_img || !grad_feat)
    return fail(DIBR_E_INVALID, "null pointer");
  RasterBwdArgs a;
  memset(&a, 0, sizeof(a));
  a.grad = grad; a.idx = idx; a.weights = weights; a.img = img; a.feat = feat;
  a.grad_img = grad_img; a.grad_feat = grad_feat;
  a.B = B; a.F = F; a.H = H; a.W = W; a.D = D;
  a.mult = multiplier; a.eps_f = eps; a.eps_d = (double)eps;
  a.inv_mult = 1.0 / (double)multiplier; a.grid = make_grid(multiplier, W, H);
  DIBR_CUDA(launch_raster_backward(a, (cudaStream_t)stream));
  return DIBR_OK;
}

// Soft-mask forward: with a pair buffer, gather -> pair evaluation, and the self-contained
// kernel as the (device-side decided) fallback when the buffer overflowed.
static int run_soft_forward(const SoftArgs& a, const Plan& p, int64_t* needed_host,
                            cudaStream_t st) {
  if (a.pairs.cap > 0) {
    DIBR_CUDA(launch_softmask(a, SOFT_GATHER, p, st));
    SoftArgs e = a;
    e.report = (needed_host && g_report_direct) ? (unsigned long long*)needed_host : nullptr;
    if (needed_host && !e.report)
      DIBR_CUDA(cudaMemcpyAsync(needed_host + 2, a.pairs.cursor, 16, cudaMemcpyDeviceToHost, st));
    DIBR_CUDA(launch_softmask_pairs_eval(e, p, st));
  }
  DIBR_CUDA(launch_softmask(a, SOFT_FWD, p, st));
  return DIBR_OK;
}

static int soft_common(int B, int F, int H, int W, const float* img, int prescaled,
                       const float* bboxes, const int64_t* sel, int K, float multiplier,
                       float box_margin, Plan& p, BinJob& job) {
  int rc = make_plan(B, F, H, W, p, multiplier);
  if (rc) return rc;
  if ((rc = check_common(multiplier))) return rc;
  if (K < 0) return fail(DIBR_E_INVALID, "knum < 0");
  if (B > 0 && !sel) return fail(DIBR_E_INVALID, "null selected_face_idx");
  if (B > 0 && F > 0 && !img) return fail(DIBR_E_INVALID, "null face_vertices_image");
  if (prescaled && !bboxes) return fail(DIBR_E_INVALID, "prescaled mode needs face_large_bboxes");
  if (!prescaled && bboxes) return fail(DIBR_E_INVALID, "face_large_bboxes only in prescaled mode");
  job = face_source(img, bboxes, nullptr, nullptr, nullptr, F, multiplier, box_margin, prescaled);
  return DIBR_OK;
}

extern "C" int dibr_softmask_fwd(int B, int F, int H, int W, const float* img, int prescaled,
                                 const float* bboxes, const int64_t* sel, float sigmainv,
                                 float box_margin, int K, float multiplier, float* mask,
                                 float* prob, int64_t* cidx, uint8_t* type, void* ws,
                                 int64_t ws_bytes, int64_t cap_r, int64_t cap_s, int64_t cap_p,
                                 int64_t* needed_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  BinJobs jobs;
  int rc = soft_common(B, F, H, W, img, prescaled, bboxes, sel, K, multiplier, box_margin, p,
                       jobs.j[0]);
  if (rc) return rc;
  if (B == 0) return DIBR_OK;
  if (!mask) return fail(DIBR_E_INVALID, "null soft_mask");
  const bool dense = prob && cidx && type;
  if (!dense && (prob || cidx || type))
    return fail(DIBR_E_INVALID, "dense outputs must be given all three or none");
  Workspace w;
  if (dense) cap_p = 0;
  if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, cap_s, w, cap_p, K))) return rc;
  attach(jobs.j[0], p, w, 1);
  jobs.j[1] = jobs.j[0];
  const int kinds[1] = {1};
  if ((rc = run_binning(p, w, jobs, kinds, 1, needed_host, st))) return rc;
  SoftArgs a;
  memset(&a, 0, sizeof(a));
  a.job = jobs.j[0]; a.sel = sel; a.sigmainv = sigmainv; a.K = K; a.mask = mask;
  a.prob = prob; a.cidx = cidx; a.type = type;
  a.pairs = w.pairs;
  if (dense) {
    DIBR_CUDA(launch_softmask(a, SOFT_FWD_DENSE, p, st));
    return DIBR_OK;
  }
  return run_soft_forward(a, p, needed_host, st);
}

extern "C" int dibr_softmask_bwd(int B, int F, int H, int W, const float* img, int prescaled,
                                 const float* bboxes, const int64_t* sel, float sigmainv,
                                 float box_margin, int K, float multiplier,
                                 const float* grad_mask, const float* mask, float* grad_img,
                                 const void* ws, int64_t ws_bytes, int64_t cap_r, int64_t cap_s,
                                 int64_t cap_p, void* stream) {
  Plan p;
  BinJob job;
  int rc = soft_common(B, F, H, W, img, prescaled, bboxes, sel, K, multiplier, box_margin, p, job);
  if (rc) return rc;
  if (B == 0 || F == 0) return DIBR_OK;
  if (!grad_mask || !mask || !grad_img) return fail(DIBR_E_INVALID, "null pointer");
  Workspace w;
  if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, cap_s, w, cap_p, K))) return rc;
  attach(job, p, w, 1);
  SoftArgs a;
  memset(&a, 0, sizeof(a));
  a.job = job; a.sel = sel; a.sigmainv = sigmainv; a.K = K;
  a.grad_mask = grad_mask; a.mask_in = mask; a.grad_img = grad_img;
  a.pairs = w.pairs;
  DIBR_CUDA(launch_softmask_pairs_grad(a, p, (cudaStream_t)stream));
  DIBR_CUDA(launch_softmask(a, SOFT_BWD, p, (cudaStream_t)stream));
  return DIBR_OK;
}

extern "C" int dibr_render_bwd(int B, int F, int H, int W, int D, const float* grad_feat_img,
                               const float* grad_mask, const int64_t* idx, const float* img,
                               const float* feat, const float* mask, float sigmainv,
                               float box_margin, int K, float multiplier, float eps,
                               float* grad_img, float* grad_feat, const void* ws,
                               int64_t ws_bytes, int64_t cap_r, int64_t cap_s, int64_t cap_p,
                               void* stream) {
  if (B < 0 || F < 0 || H <= 0 || W <= 0 || D < 0 || K < 0)
    return fail(DIBR_E_INVALID, "invalid sizes");
  int rc = check_common(multiplier);
  if (rc) return rc;
  if (B > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", B);
  if (B == 0 || F == 0) return DIBR_OK;
  if (!idx || !img || !grad_img || (D > 0 && (!feat || !grad_feat)))
    return fail(DIBR_E_INVALID, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t nfv = (size_t)B * F * 3;
  DIBR_CUDA(zero_gradients(grad_img, nfv * 2 * sizeof(float), D > 0 ? grad_feat : nullptr,
                           nfv * D * sizeof(float), st));
  const bool do_raster = grad_feat_img && D > 0, do_soft = grad_mask && K > 0;
  if (do_soft && !mask) return fail(DIBR_E_INVALID, "null soft_mask");
  RasterBwdArgs ra;
  memset(&ra, 0, sizeof(ra));
  if (do_raster) {
    ra.grad = grad_feat_img; ra.idx = idx; ra.weights = nullptr; ra.img = img; ra.feat = feat;
    ra.grad_img = grad_img; ra.grad_feat = grad_feat;
    ra.B = B; ra.F = F; ra.H = H; ra.W = W; ra.D = D;
    ra.mult = multiplier; ra.eps_f = eps; ra.eps_d = (double)eps;
    ra.inv_mult = 1.0 / (double)multiplier; ra.grid = make_grid(multiplier, W, H);
    DIBR_CUDA(launch_raster_backward(ra, st));
  }
  if (do_soft) {
    // (running this on a second stream beside the rasterizer's backward was measured: both
    // kernels fill the SMs, the overlap gained nothing)
    rc = dibr_softmask_bwd(B, F, H, W, img, 0, nullptr, idx, sigmainv, box_margin, K, multiplier,
                           grad_mask, mask, grad_img, ws, ws_bytes, cap_r, cap_s, cap_p, st);
    if (rc) return rc;
  }
  return DIBR_OK;
}

extern "C" int dibr_softmask_bwd_dense(int B, int F, int H, int W, int K, const float* grad_mask,
                                       const float* mask, const int64_t* sel, const float* prob,
                                       const int64_t* cidx, const uint8_t* type,
                                       const float* img_scaled, float sigmainv, float multiplier,
                                       float* grad_img, void* stream) {
  if (B < 0 || F < 0 || H <= 0 || W <= 0 || K < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  int rc = check_common(multiplier);
  if (rc) return rc;
  if (B == 0 || F == 0 || K == 0) return DIBR_OK;
  if (!grad_mask || !mask || !sel || !prob || !cidx || !type || !img_scaled || !grad_img)
    return fail(DIBR_E_INVALID, "null pointer");
  SoftDenseBwdArgs a;
  a.grad_mask = grad_mask; a.mask = mask; a.sel = sel; a.prob = prob; a.cidx = cidx;
  a.type = type; a.img = img_scaled; a.grad_img = grad_img;
  a.B = B; a.F = F; a.H = H; a.W = W; a.K = K; a.sigmainv = sigmainv; a.mult = multiplier;
  DIBR_CUDA(launch_softmask_dense_backward(a, (cudaStream_t)stream));
  return DIBR_OK;
}

extern "C" int dibr_render_fwd(int B, int F, int H, int W, int D, const float* z,
                               const float* img, const float* feat, const float* normal_z,
                               float sigmainv, float box_margin, int K, float multiplier,
                               float eps, int64_t* out_idx, float* out_feat, float* mask,
                               void* ws, int64_t ws_bytes, int64_t cap_r, int64_t cap_s,
                               int64_t cap_p, int64_t* needed_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  int rc = make_plan(B, F, H, W, p, multiplier);
  if (rc) return rc;
  if ((rc = check_common(multiplier))) return rc;
  if (D < 0 || K < 0) return fail(DIBR_E_INVALID, "feat_dim / knum < 0");
  if (B == 0) return DIBR_OK;
  if (!out_idx || (!out_feat && D > 0) || !mask) return fail(DIBR_E_INVALID, "null output");
  if (F > 0 && (!z || !img || (!feat && D > 0))) return fail(DIBR_E_INVALID, "null input");
  Workspace w;
  if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, cap_s, w, cap_p, K))) return rc;
  BinJobs jobs;
  jobs.j[0] = face_source(img, nullptr, nullptr, normal_z, nullptr, F, multiplier, 0.f, 0);
  jobs.j[1] = face_source(img, nullptr, nullptr, nullptr, nullptr, F, multiplier, box_margin, 0);
  attach(jobs.j[0], p, w, 0);
  attach(jobs.j[1], p, w, 1);
  jobs.j[0].rec = w.rec; jobs.j[0].z = z;
  const int kinds[2] = {0, 1};
  if ((rc = run_binning(p, w, jobs, kinds, 2, needed_host, st))) return rc;
  RasterArgs a;
  memset(&a, 0, sizeof(a));
  a.job = jobs.j[0]; a.rec = jobs.j[0].rec; a.z = z; a.feat = feat; a.D = D; a.eps_d = (double)eps;
  a.out_idx = out_idx; a.out_feat = out_feat; a.out_w = nullptr;
  a.block_uncovered = (int*)(w.base + p.off_flags);
  const bool handoff = gather_handoff_enabled() && w.pairs.cap > 0 && K > 0;
  if (handoff) {   // the rasterizer writes soft_mask's base value and the gather's work list
    a.mask = mask; a.soft_count = jobs.j[1].count; a.work = gather_work(p, w);
  }
  DIBR_CUDA(launch_raster_forward(a, p, st));
  SoftArgs s;
  memset(&s, 0, sizeof(s));
  s.job = jobs.j[1]; s.sel = out_idx; s.sigmainv = sigmainv; s.K = K; s.mask = mask;
  s.pairs = w.pairs;
  s.block_uncovered = (const int*)(w.base + p.off_flags);
  if (handoff) s.work = a.work;
  return run_soft_forward(s, p, needed_host, st);
}

// ---------------------------------------------------------------- double precision
// The reference's four raw ops for scalar_t = double (dibr_f64.cu).  Binning runs on the
// float copies the caller passes beside the double arrays.
extern "C" int dibr_packed_rasterize_fwd_f64(
    int B, int F, int H, int W, int D, const double* z, const double* img, const double* bbox,
    const double* feat, const float* img32, const float* bbox32, const int64_t* first,
    float multiplier, float eps, int64_t* out_idx, double* out_feat, double* out_w, void* ws,
    int64_t ws_bytes, int64_t cap_r, int64_t* needed_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  int rc = make_plan(B, F, H, W, p, multiplier);
  if (rc) return rc;
  if ((rc = check_common(multiplier))) return rc;
  if (D < 0) return fail(DIBR_E_INVALID, "feat_dim < 0");
  if (B == 0) return DIBR_OK;
  if (!out_idx || (!out_feat && D > 0) || !first) return fail(DIBR_E_INVALID, "null pointer");
  if (F > 0 && (!z || !img || !bbox || !img32 || !bbox32 || (!feat && D > 0)))
    return fail(DIBR_E_INVALID, "null input");
  Workspace w;
  if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, 0, w))) return rc;
  BinJobs jobs;
  jobs.j[0] = face_source(img32, bbox32, nullptr, nullptr, first, 0, multiplier, 0.f, 1);
  attach(jobs.j[0], p, w, 0);
  jobs.j[1] = jobs.j[0];
  const int kinds[1] = {0};
  if ((rc = run_binning(p, w, jobs, kinds, 1, needed_host, st))) return rc;
  RasterF64Args a;
  a.job = jobs.j[0]; a.z = z; a.img = img; a.bbox = bbox; a.feat = feat;
  a.eps_d = (double)eps; a.D = D; a.out_idx = out_idx; a.out_feat = out_feat; a.out_w = out_w;
  DIBR_CUDA(launch_raster_forward_f64(a, p, st));
  return DIBR_OK;
}

extern "C" int dibr_rasterize_bwd_f64(int B, int F, int H, int W, int D, const double* grad,
                                      const int64_t* idx, const double* weights,
                                      const double* img, const double* feat, float eps,
                                      double* grad_img, double* grad_feat, void* stream) {
  if (B < 0 || F < 0 || H <= 0 || W <= 0 || D < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  if (B == 0 || F == 0 || D == 0) return DIBR_OK;
  if (B > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", B);
  if (!grad || !idx || !weights || !img || !feat || !grad_img || !grad_feat)
    return fail(DIBR_E_INVALID, "null pointer");
  RasterBwdF64Args a;
  a.grad = grad; a.idx = idx; a.weights = weights; a.img = img; a.feat = feat;
  a.grad_img = grad_img; a.grad_feat = grad_feat;
  a.B = B; a.F = F; a.H = H; a.W = W; a.D = D; a.eps_d = (double)eps;
  DIBR_CUDA(launch_raster_backward_f64(a, (cudaStream_t)stream));
  return DIBR_OK;
}

extern "C" int dibr_softmask_fwd_f64(int B, int F, int H, int W, const double* img,
                                     const double* bbox, const float* img32, const float* bbox32,
                                     const int64_t* sel, float sigmainv, int K, float multiplier,
                                     double* mask, double* prob, int64_t* cidx, uint8_t* type,
                                     void* ws, int64_t ws_bytes, int64_t cap_s,
                                     int64_t* needed_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  BinJobs jobs;
  int rc = soft_common(B, F, H, W, img32, 1, bbox32, sel, K, multiplier, 0.f, p, jobs.j[0]);
  if (rc) return rc;
  if (B == 0) return DIBR_OK;
  if (!mask || !prob || !cidx || !type) return fail(DIBR_E_INVALID, "null output");
  if (F > 0 && (!img || !bbox)) return fail(DIBR_E_INVALID, "null input");
  Workspace w;
  if ((rc = bind_workspace(p, ws, ws_bytes, 0, cap_s, w))) return rc;
  attach(jobs.j[0], p, w, 1);
  jobs.j[1] = jobs.j[0];
  const int kinds[1] = {1};
  if ((rc = run_binning(p, w, jobs, kinds, 1, needed_host, st))) return rc;
  SoftF64Args a;
  a.job = jobs.j[0]; a.img = img; a.bbox = bbox; a.sel = sel; a.sigmainv = sigmainv; a.K = K;
  a.mask = mask; a.prob = prob; a.cidx = cidx; a.type = type;
  DIBR_CUDA(launch_softmask_forward_f64(a, p, st));
  return DIBR_OK;
}

extern "C" int dibr_softmask_bwd_dense_f64(int B, int F, int H, int W, int K,
                                           const double* grad_mask, const double* mask,
                                           const int64_t* sel, const double* prob,
                                           const int64_t* cidx, const uint8_t* type,
                                           const double* img_scaled, float sigmainv,
                                           float multiplier, double* grad_img, void* stream) {
  if (B < 0 || F < 0 || H <= 0 || W <= 0 || K < 0) return fail(DIBR_E_INVALID, "invalid sizes");
  int rc = check_common(multiplier);
  if (rc) return rc;
  if (B == 0 || F == 0 || K == 0) return DIBR_OK;
  if (!grad_mask || !mask || !sel || !prob || !cidx || !type || !img_scaled || !grad_img)
    return fail(DIBR_E_INVALID, "null pointer");
  SoftBwdF64Args a;
  a.grad_mask = grad_mask; a.mask = mask; a.sel = sel; a.prob = prob; a.cidx = cidx;
  a.type = type; a.img = img_scaled; a.grad_img = grad_img;
  a.B = B; a.F = F; a.H = H; a.W = W; a.K = K; a.sigmainv = sigmainv; a.mult = multiplier;
  DIBR_CUDA(launch_softmask_backward_f64(a, (cudaStream_t)stream));
  return DIBR_OK;
}


// ---------------------------------------------------------------- mesh-level entry points
// the renderer reads the face corners THROUGH `faces` from the per-view
// vertex arrays, so the (B, F, 3, .) face tensors of prepare_vertices
// (kaolin/render/mesh/utils.py:128-175, ops/mesh/mesh.py:25-47, trianglemesh.py:313-337)
// never exist, and the backward adds straight into vertex-level gradients.
static BinJob mesh_source(const int* faces, const float* vimg, const float* vcam, int V, int F,
                          const uint8_t* valid, int cull, float mult, float margin) {
  BinJob j;
  memset(&j, 0, sizeof(j));
  j.faces = faces; j.vimg = vimg; j.vcam = vcam; j.V = V; j.mesh_cull = cull;
  j.valid = valid; j.view_stride = F; j.mult = mult; j.margin = margin;
  return j;
}

extern "C" int dibr_render_mesh_fwd(int B, int V, int F, int H, int W, int D,
                                    const float* vertices_camera, const float* vertices_image,
                                    const int32_t* faces, const float* vertex_features,
                                    int features_shared, const uint8_t* valid_faces,
                                    float sigmainv, float box_margin, int K, float multiplier,
                                    float eps, int64_t* out_idx, float* out_feat, float* mask,
                                    void* ws, int64_t ws_bytes, int64_t cap_r, int64_t cap_s,
                                    int64_t cap_p, int64_t* needed_host, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  int rc = make_plan(B, F, H, W, p, multiplier);
  if (rc) return rc;
  if ((rc = check_common(multiplier))) return rc;
  if (D < 0 || K < 0 || V < 0) return fail(DIBR_E_INVALID, "feat_dim / knum / num_vertices < 0");
  if (B == 0) return DIBR_OK;
  if (!out_idx || (!out_feat && D > 0) || !mask) return fail(DIBR_E_INVALID, "null output");
  if (F > 0 && (!faces || !vertices_image || (!vertices_camera) || (!vertex_features && D > 0)))
    return fail(DIBR_E_INVALID, "null input");
  Workspace w;
  if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, cap_s, w, cap_p, K))) return rc;
  BinJobs jobs;
  // rasterizer: valid_faces if given, else normal_z >= 0 from the camera-space corners
  // (dibr.py:190-199); soft mask: every face (dibr.py:201-203)
  jobs.j[0] = mesh_source(faces, vertices_image, vertices_camera, V, F, valid_faces,
                          valid_faces ? 0 : 1, multiplier, 0.f);
  jobs.j[1] = mesh_source(faces, vertices_image, vertices_camera, V, F, nullptr, 0, multiplier,
                          box_margin);
  attach(jobs.j[0], p, w, 0);
  attach(jobs.j[1], p, w, 1);
  jobs.j[0].rec = w.rec; jobs.j[0].z = nullptr;
  const int kinds[2] = {0, 1};
  if ((rc = run_binning(p, w, jobs, kinds, 2, needed_host, st))) return rc;
  RasterArgs a;
  memset(&a, 0, sizeof(a));
  a.job = jobs.j[0]; a.rec = jobs.j[0].rec; a.z = nullptr; a.feat = vertex_features;
  a.feat_shared = features_shared; a.D = D; a.eps_d = (double)eps;
  a.out_idx = out_idx; a.out_feat = out_feat; a.out_w = nullptr;
  a.block_uncovered = (int*)(w.base + p.off_flags);
  const bool handoff = gather_handoff_enabled() && w.pairs.cap > 0 && K > 0;
  if (handoff) {   // the rasterizer writes soft_mask's base value and the gather's work list
    a.mask = mask; a.soft_count = jobs.j[1].count; a.work = gather_work(p, w);
  }
  DIBR_CUDA(launch_raster_forward(a, p, st));
  SoftArgs s;
  memset(&s, 0, sizeof(s));
  s.job = jobs.j[1]; s.sel = out_idx; s.sigmainv = sigmainv; s.K = K; s.mask = mask;
  s.pairs = w.pairs;
  s.block_uncovered = (const int*)(w.base + p.off_flags);
  if (handoff) s.work = a.work;
  return run_soft_forward(s, p, needed_host, st);
}

extern "C" int dibr_render_mesh_bwd(int B, int V, int F, int H, int W, int D,
                                    const float* grad_feat_img, const float* grad_mask,
                                    const int64_t* idx, const float* vertices_image,
                                    const int32_t* faces, const float* vertex_features,
                                    int features_shared, const float* mask, float sigmainv,
                                    float box_margin, int K, float multiplier, float eps,
                                    float* grad_vertices_image, float* grad_vertex_features,
                                    const void* ws, int64_t ws_bytes,
                                    int64_t cap_r, int64_t cap_s, int64_t cap_p, void* stream) {
  if (B < 0 || V < 0 || F < 0 || H <= 0 || W <= 0 || D < 0 || K < 0)
    return fail(DIBR_E_INVALID, "invalid sizes");
  int rc = check_common(multiplier);
  if (rc) return rc;
  if (B > 65535) return fail(DIBR_E_INVALID, "batch %d > 65535 not supported", B);
  if (B == 0) return DIBR_OK;
  if (!grad_vertices_image || (D > 0 && !grad_vertex_features))
    return fail(DIBR_E_INVALID, "null gradient output");
  cudaStream_t st = (cudaStream_t)stream;
  DIBR_CUDA(zero_gradients(grad_vertices_image, (size_t)B * V * 2 * sizeof(float),
                           D > 0 ? grad_vertex_features : nullptr,
                           (size_t)(features_shared ? 1 : B) * V * D * sizeof(float), st));
  if (F == 0 || V == 0) return DIBR_OK;
  if (!idx || !vertices_image || !faces || (D > 0 && !vertex_features))
    return fail(DIBR_E_INVALID, "null pointer");
  const bool do_raster = grad_feat_img && D > 0, do_soft = grad_mask && K > 0;
  if (do_soft && !mask) return fail(DIBR_E_INVALID, "null soft_mask");
  RasterBwdArgs ra;
  memset(&ra, 0, sizeof(ra));
  if (do_raster) {
    ra.grad = grad_feat_img; ra.idx = idx; ra.weights = nullptr; ra.img = vertices_image;
    ra.feat = vertex_features; ra.grad_img = grad_vertices_image;
    ra.grad_feat = grad_vertex_features;
    ra.faces = faces; ra.V = V; ra.feat_shared = features_shared;
    ra.B = B; ra.F = F; ra.H = H; ra.W = W; ra.D = D;
    ra.mult = multiplier; ra.eps_f = eps; ra.eps_d = (double)eps;
    ra.inv_mult = 1.0 / (double)multiplier; ra.grid = make_grid(multiplier, W, H);
    DIBR_CUDA(launch_raster_backward(ra, st));
  }
  if (do_soft) {
    Plan p;
    if ((rc = make_plan(B, F, H, W, p, multiplier))) return rc;
    Workspace w;
    if ((rc = bind_workspace(p, ws, ws_bytes, cap_r, cap_s, w, cap_p, K))) return rc;
    BinJob job = mesh_source(faces, vertices_image, nullptr, V, F, nullptr, 0, multiplier,
                             box_margin);
    attach(job, p, w, 1);
    SoftArgs a;
    memset(&a, 0, sizeof(a));
    a.job = job; a.sel = idx; a.sigmainv = sigmainv; a.K = K;
    a.grad_mask = grad_mask; a.mask_in = mask; a.grad_img = grad_vertices_image;
    a.pairs = w.pairs;
    DIBR_CUDA(launch_softmask_pairs_grad(a, p, st));
    DIBR_CUDA(launch_softmask(a, SOFT_BWD, p, st));
  }
  return DIBR_OK;
}

// ------------------------------------------------------------- no-allocation variant
// dibr_rasterize_fwd with its scratch workspace (binning tables + the per-bin face lists) in
// a file-scope device array instead of caller memory, for callers that cannot allocate
// (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc);
// it has no counterpart in the reference.  Every GPU that
// loads the library holds DIBR_STATIC_SCRATCH_BYTES of it until exit.  All calls share the
// array (overlapping calls on one GPU race); each call re-zeroes the parts that must start
// at zero (run_binning's memsets), the rest is overwritten before it is read.
static __device__ __align__(256) unsigned char g_static_scratch[DIBR_STATIC_SCRATCH_BYTES];

// Device address of the array on the CURRENT device.  In host code `g_static_scratch` is
// only a stand-in for the symbol, never a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_ptr(void** out) {
  constexpr int kMaxDevices = 64;
  static void* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  DIBR_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(DIBR_E_INVALID, "device index %d not supported", dev);
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    DIBR_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p || ((uintptr_t)p & 255) != 0)
      return fail(DIBR_E_CUDA, "static scratch array has no usable device address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return DIBR_OK;
}

extern "C" int dibr_rasterize_fwd_static(int B, int F, int H, int W, int D, const float* z,
                                         const float* img, const float* feat,
                                         const uint8_t* valid, const float* normal_z,
                                         float multiplier, float eps, int64_t* out_idx,
                                         float* out_feat, float* out_w, int64_t* needed_host,
                                         void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  Plan p;
  int rc = make_plan(B, F, H, W, p, multiplier);
  if (rc) return rc;
  // sizes first: the fixed tables plus room for 2 list entries per face-view (+ 4096) must
  // fit into the array; whatever is left beyond that is list capacity as well
  const int64_t want = 2 * (int64_t)B * F + 4096;
  const int64_t room = ((int64_t)DIBR_STATIC_SCRATCH_BYTES - p.fixed_bytes - 512) / raster_entry_bytes();
  if (room < want)
    return fail(DIBR_E_INVALID,
                "dibr_rasterize_fwd_static: batch %d x %d faces at %dx%d needs %lld bytes of "
                "scratch, the static array holds DIBR_STATIC_SCRATCH_BYTES = %lld; use "
                "dibr_rasterize_fwd with a caller-allocated workspace",
                B, F, H, W, (long long)(p.fixed_bytes + 512 + want * raster_entry_bytes()),
                (long long)DIBR_STATIC_SCRATCH_BYTES);
  if ((rc = check_common(multiplier))) return rc;
  if (D < 0) return fail(DIBR_E_INVALID, "feat_dim < 0");
  if (B == 0) return DIBR_OK;
  if (!out_idx || (!out_feat && D > 0)) return fail(DIBR_E_INVALID, "null output");
  if (F > 0 && (!z || !img || (!feat && D > 0))) return fail(DIBR_E_INVALID, "null input");
  if (valid && normal_z) return fail(DIBR_E_INVALID, "valid_faces and face_normals_z both given");
  void* ws = nullptr;
  if ((rc = static_scratch_ptr(&ws))) return rc;
  const int64_t cap_r = room < kMaxListEntries ? room : kMaxListEntries;
  Workspace w;
  if ((rc = bind_workspace(p, ws, DIBR_STATIC_SCRATCH_BYTES, cap_r, 0, w))) return rc;
  BinJobs jobs;
  jobs.j[0] = face_source(img, nullptr, valid, normal_z, nullptr, F, multiplier, 0.f, 0);
  attach(jobs.j[0], p, w, 0);
  jobs.j[0].rec = w.rec; jobs.j[0].z = z;
  jobs.j[1] = jobs.j[0];
  const int kinds[1] = {0};
  if ((rc = run_binning(p, w, jobs, kinds, 1, needed_host, st))) return rc;
  RasterArgs a;
  memset(&a, 0, sizeof(a));
  a.job = jobs.j[0]; a.rec = jobs.j[0].rec; a.z = z; a.feat = feat; a.D = D; a.eps_d = (double)eps;
  a.out_idx = out_idx; a.out_feat = out_feat; a.out_w = out_w;
  a.block_uncovered = (int*)(w.base + p.off_flags);
  DIBR_CUDA(launch_raster_forward(a, p, st));
  return DIBR_OK;
}
