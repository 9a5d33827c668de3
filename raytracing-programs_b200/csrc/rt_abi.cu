// csrc/rt_abi.cu -- implementation of include/rt_abi.h: scene flattening + upload,
// tile scheduling of the wavefront kernels, multi-GPU band gather, statistics.
// One host thread drives all GPUs of the process (asynchronous launches, one stream
// per device).  NCCL is resolved with dlopen only when more than one GPU is in use.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <cmath>
#include <nccl.h>   // types only; symbols come from dlopen("libnccl.so.2")
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <algorithm>
#include <thread>
#include <utility>
#include <vector>

#define RT_ABI_DEBUG 1   /* the rt_dbg_* diagnostics are defined here */
#include "rt_abi.h"
#include "rt_kernels.cuh"
#include "rt_p3.cuh"

#define RT_MAX_GPUS 8

namespace {

char g_err[512] = "";
int set_err(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}
#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return set_err(e_ == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA, "%s failed at %s:%d: %s", \
                     #call, __FILE__, __LINE__, cudaGetErrorString(e_));                      \
  } while (0)

double now_ms() {
  struct timeval tv;
  gettimeofday(&tv, NULL);
  return tv.tv_sec * 1e3 + tv.tv_usec / 1e3;
}

// RT_TRACE=1: wall-clock stamps of the host-side stages on stderr (where does a one-shot CLI run spend its time)
bool g_trace = getenv("RT_TRACE") != nullptr;
double g_trace_t0 = 0;
void trace(const char *what) {
  if (!g_trace) return;
  const double t = now_ms();
  if (g_trace_t0 == 0) g_trace_t0 = t;
  fprintf(stderr, "[rt_trace %9.3f ms] %s\n", t - g_trace_t0, what);
}

struct DeviceCtx {
  int dev = -1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  int sm_count = 0;
  // scene
  void *scene_blob = nullptr;
  DevScene sc{};
  // per-tile working set
  size_t tile_cap = 0;
  float4 *acc = nullptr, *c0 = nullptr;
  float4 *lvl = nullptr; size_t lvl_cap = 0;   // RT_MODE_SHOOT: per-level {colour, reflectivity} of a tile, [depth][tile of the render]
  float4 *qbuf = nullptr;     // 6 * tile_cap float4: two queues x {r0,r1,r2}
  int *qcount = nullptr;      // [RT_MAX_DEPTH_CAP + 1 levels][RT_CTL_STRIDE] queue control words (HitQueue::ctl)
  int *qcls = nullptr;        // 2 queues x RT_NCLS x tile_cap class lists
  unsigned long long *counters = nullptr;
  // output
  uint8_t *rank_buf = nullptr; size_t rank_cap = 0;
  uint8_t *staging = nullptr; size_t staging_cap = 0;   // device 0 only
  uint8_t *frame = nullptr; size_t frame_cap = 0;       // device 0 only
  int grid_primary = 0, grid_shade = 0, grid_fin = 0;
  int lb = 0;                         // lights per k_shade pass for the resident scene (pick_lb)
  void *shade = nullptr;              // the k_shade instantiation of the resident scene (loaded at upload)
  bool p3_attr_set = false;           // k_p3_emit's dynamic shared memory attribute is per device
  // one event before every launch (+1 at the end): per-kernel-class device times
  std::vector<cudaEvent_t> evpool;
  std::vector<char> evkind;   // 'p' primary, 's' shade, 'f' finalize
  // P3 writer (device 0 only)
  char *p3_text = nullptr; size_t p3_text_cap = 0;
  unsigned long long *p3_state = nullptr; size_t p3_state_cap = 0;   // span offsets + span lengths (bytes)
  char *p3_pinned[3] = {nullptr, nullptr, nullptr}; size_t p3_pinned_cap = 0;   // three slabs: copy k+2 is queued while the sink has k
  cudaEvent_t p3_ev[3] = {nullptr, nullptr, nullptr};
  int frame_w = 0, frame_h = 0;   // size of the frame currently held in `frame`
};

DeviceCtx g_dev[RT_MAX_GPUS];
int g_ndev = 0;
int prepare_kernels(DeviceCtx &d);   // launch plan of the resident scene (defined with the launchers below)
bool g_have_scene = false;
int g_n_lights = 0, g_n_shapes = 0;
int g_max_depth = 7;   // MAX_RECURSION, C/raycast.c:14; rt_set_max_depth
int g_mode = 0;    // rt_set_mode: RT_MODE_ITERATIVE / RT_MODE_SHOOT
int g_accel = 0;   // rt_set_accel / RT_ACCEL: light buffer for the sphere part of shadow queries

// ---- NCCL through dlopen ---------------------------------------------------------
struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  ncclComm_t comms[RT_MAX_GPUS] = {};
  bool ready = false;
} g_nccl;

int nccl_load(int ndev) {
  if (g_nccl.ready) return RT_OK;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char *n : names) { g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (g_nccl.lib) break; }
  if (!g_nccl.lib) return set_err(RT_ERR_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                          \
  *(void **)(&g_nccl.field) = dlsym(g_nccl.lib, name);                            \
  if (!g_nccl.field) return set_err(RT_ERR_NCCL, "libnccl lacks symbol %s", name)
  SYM(CommInitAll, "ncclCommInitAll"); SYM(CommDestroy, "ncclCommDestroy");
  SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd");
  SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  // whatever libnccl.so.2 the process resolves (the one torch already mapped, else the system's): point-to-point
  // send/recv inside a group needs 2.7+, and the structs of <nccl.h> we compile against are stable across 2.x
  {
    ncclResult_t (*get_version)(int *) = nullptr;
    *(void **)(&get_version) = dlsym(g_nccl.lib, "ncclGetVersion");
    int v = 0;
    if (get_version && get_version(&v) == ncclSuccess && (v < 2700 || v >= 30000))
      return set_err(RT_ERR_NCCL, "libnccl version %d is outside the supported range [2.7, 3.0)", v);
  }
  int devs[RT_MAX_GPUS];
  for (int i = 0; i < ndev; i++) devs[i] = g_dev[i].dev;
  ncclResult_t r = g_nccl.CommInitAll(g_nccl.comms, ndev, devs);
  if (r != ncclSuccess) return set_err(RT_ERR_NCCL, "ncclCommInitAll: %s", g_nccl.GetErrorString(r));
  g_nccl.ready = true;
  return RT_OK;
}

void nccl_unload() {
  if (g_nccl.ready) for (int i = 0; i < g_ndev; i++) if (g_nccl.comms[i]) g_nccl.CommDestroy(g_nccl.comms[i]);
  g_nccl = NcclApi();   // keep the library mapped: NCCL does not like dlclose
}

int band_rows() {
  const char *e = getenv("RT_BAND_ROWS");
  int v = e ? atoi(e) : 32;
  return v > 0 ? v : 32;
}
size_t tile_pixels_cap() {
  const char *e = getenv("RT_TILE_PIXELS");
  // default 64 Mi pixels (12.9 GB of wavefront state): the deep bounce levels of a tile hold few hit points and their
  // launches are bound by the latency of one warp's scan, not by throughput, so fewer, larger tiles cost less of them
  // (C4 16384^2 on one B200: 862.7 ms per frame with 16 Mi, 842.3 with 32 Mi, 835.7 with 64 Mi, 835.6 with 128 Mi)
  long long v = e ? atoll(e) : (64ll << 20);
  if (v < 1024) v = 1024;
  if (v > (1ll << 30)) v = 1ll << 30;
  return (size_t)v;
}

template <typename T>
int dev_realloc(T *&p, size_t &cap, size_t need) {
  if (need <= cap && p) return RT_OK;
  if (p) { cudaFree(p); p = nullptr; cap = 0; }
  CK(cudaMalloc((void **)&p, need));
  cap = need;
  return RT_OK;
}

void free_device(DeviceCtx &d) {
  if (d.dev < 0) return;
  cudaSetDevice(d.dev);
  if (d.stream) cudaStreamSynchronize(d.stream);
  cudaFree(d.scene_blob); cudaFree(d.lvl); cudaFree(d.acc); /* c0, qbuf, qcls live in the same block */ cudaFree(d.qcount);
  cudaFree(d.counters); cudaFree(d.rank_buf); cudaFree(d.staging); cudaFree(d.frame);
  cudaFree(d.p3_text); cudaFree(d.p3_state);
  for (int k = 0; k < 3; k++) { if (d.p3_pinned[k]) cudaFreeHost(d.p3_pinned[k]); if (d.p3_ev[k]) cudaEventDestroy(d.p3_ev[k]); }
  if (d.ev0) cudaEventDestroy(d.ev0);
  if (d.ev1) cudaEventDestroy(d.ev1);
  if (d.ev2) cudaEventDestroy(d.ev2);
  if (d.stream) cudaStreamDestroy(d.stream);
  for (cudaEvent_t e : d.evpool) cudaEventDestroy(e);
  d = DeviceCtx();
}

// CUDA loads kernels lazily: touch every kernel once at init so the first render does not pay for it.
void warm_modules() {
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k_primary<false>); cudaFuncGetAttributes(&fa, k_finalize); cudaFuncGetAttributes(&fa, k_assemble);
  // (the k_shade instantiations are loaded by shade_fn() at their first use: 20 of them, one is used per scene)
  cudaFuncGetAttributes(&fa, k_p3_count); cudaFuncGetAttributes(&fa, k_p3_scan); cudaFuncGetAttributes(&fa, k_p3_emit);
  cudaGetLastError();
}

// ---- light buffer ----------------------------------------------
// Exact acceleration of the sphere part of shadow queries.  A shadow ray towards light L from P lies
// on the line through L with direction u = (P - L)/|P - L| (the reference sets no t_max, so the part
// beyond the light counts too: direction -u).  The reference can report a hit on sphere (C, r) only
// if its float discriminant is >= 0, which by the reject-filter contract needs the line
// to pass within r_eff = sqrt(r^2 + 2^-18 (T + r^2)) of C, T = |P - C|^2 (the filter's error budget of
// 35u (T + r^2) plus the roundings of restating it in exact arithmetic: 59u < 2^-18); for |P - L| <= Dcap,
// T <= (Dcap + |C - L|)^2.  So per light, per distance level (Dcap = 0.35 .. 16 x the light's scene
// radius: the nearer the hit point, the less the spheres need inflating) and per cube-map cell the list holds every sphere whose cone of half-angle
// asin(r_eff / |C - L|) around +-(C - L), widened by the cell's own angular radius and a 3e-4 rad
// margin (float rounding of the direction and of the cell lookup is ~1e-7), meets the cell.
// Hit points farther than the last level, and degenerate lights, fall back to testing every sphere.
struct LightBuffer {
  std::vector<unsigned> start;
  std::vector<int> items;
  std::vector<float> caps;
};

void lb_dir(int face, double s, double t, double *o) {
  const double sg = (face & 1) ? -1.0 : 1.0;
  if (face < 2) { o[0] = sg; o[1] = s; o[2] = t; }
  else if (face < 4) { o[0] = s; o[1] = sg; o[2] = t; }
  else { o[0] = s; o[1] = t; o[2] = sg; }
  const double n = sqrt(o[0] * o[0] + o[1] * o[1] + o[2] * o[2]);
  o[0] /= n; o[1] /= n; o[2] /= n;
}

// Bounding-sphere hierarchy over sphere slots [0, n): median split along the widest axis of the centres,
// leaves of <= 4 spheres.  Node sphere: centre of the members' bounding box, radius max(|C - c| + |r|),
// computed in double and rounded up.  Returns false (no hierarchy) on non-finite geometry or depth > 28.
struct SphereBvh {
  std::vector<float4> node;
  std::vector<int2> kids;
  std::vector<int> items;
};

bool bvh_build_rec(const std::vector<float4> &sph, std::vector<int> &ids, int lo, int hi, int depth, SphereBvh &out, int *self) {
  if (depth > 28) return false;
  double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
  for (int k = lo; k < hi; k++) {
    const float4 &q = sph[ids[k]];
    const double c[3] = {q.x, q.y, q.z}, r = fabs((double)q.w);
    if (!std::isfinite(c[0]) || !std::isfinite(c[1]) || !std::isfinite(c[2]) || !std::isfinite(r)) return false;
    for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], c[a] - r); mx[a] = std::max(mx[a], c[a] + r); }
  }
  const double cc[3] = {0.5 * (mn[0] + mx[0]), 0.5 * (mn[1] + mx[1]), 0.5 * (mn[2] + mx[2])};
  const float cf[3] = {(float)cc[0], (float)cc[1], (float)cc[2]};
  double R = 0;
  for (int k = lo; k < hi; k++) {
    const float4 &q = sph[ids[k]];
    const double dx = q.x - (double)cf[0], dy = q.y - (double)cf[1], dz = q.z - (double)cf[2];
    R = std::max(R, sqrt(dx * dx + dy * dy + dz * dz) + fabs((double)q.w));
  }
  const float Rf = (float)(R * (1.0 + 1e-6)) + 1e-30f;
  if (!std::isfinite(Rf)) return false;
  const int me = (int)out.node.size();
  *self = me;
  out.node.push_back(make_float4(cf[0], cf[1], cf[2], Rf));
  out.kids.push_back(make_int2(0, 0));
  if (hi - lo <= 4) {
    out.kids[me] = make_int2(-((int)out.items.size() + 1), hi - lo);
    for (int k = lo; k < hi; k++) out.items.push_back(ids[k]);
    return true;
  }
  // Split: sweep the members sorted by centre along each axis and take the cut that minimises
  // n_left * d_left^2 + n_right * d_right^2, d = diagonal of the side's bounding box (a node sphere is built around
  // that box, and the chance that a ray meets it grows with its cross-section).  Where a lopsided tree comes close to
  // the depth limit it falls back to the median along the widest axis (balanced from there on).
  int axis = 0, mid = (lo + hi) / 2;
  const int n = hi - lo;
  int lg = 0;
  while ((1 << lg) < n) lg++;
  if (depth + lg + 2 < 28 && !getenv("RT_BVH_MEDIAN")) {     // a median-split subtree of n members is lg levels deep
    double best = 1e300;
    std::vector<int> tmp(ids.begin() + lo, ids.begin() + hi), keep;
    std::vector<double> suf((size_t)n + 1);
    for (int a = 0; a < 3; a++) {
      std::sort(tmp.begin(), tmp.end(), [&](int x, int y) {
        const float vx = a == 0 ? sph[x].x : a == 1 ? sph[x].y : sph[x].z, vy = a == 0 ? sph[y].x : a == 1 ? sph[y].y : sph[y].z;
        return vx < vy || (vx == vy && x < y);
      });
      double bmn[3] = {1e300, 1e300, 1e300}, bmx[3] = {-1e300, -1e300, -1e300};
      auto grow = [&](int id) {
        const float4 &q = sph[id];
        const double c[3] = {q.x, q.y, q.z}, r = fabs((double)q.w);
        for (int k = 0; k < 3; k++) { bmn[k] = std::min(bmn[k], c[k] - r); bmx[k] = std::max(bmx[k], c[k] + r); }
      };
      auto diag2 = [&]() { return (bmx[0] - bmn[0]) * (bmx[0] - bmn[0]) + (bmx[1] - bmn[1]) * (bmx[1] - bmn[1]) + (bmx[2] - bmn[2]) * (bmx[2] - bmn[2]); };
      for (int k = n - 1; k >= 1; k--) { grow(tmp[k]); suf[k] = diag2() * (n - k); }
      for (int k = 0; k < 3; k++) { bmn[k] = 1e300; bmx[k] = -1e300; }
      bool took = false;
      for (int k = 1; k < n; k++) {
        grow(tmp[k - 1]);
        const double cost = diag2() * k + suf[k];
        if (cost < best) { best = cost; axis = a; mid = lo + k; took = true; }
      }
      if (took) keep = tmp;
    }
    if (!keep.empty()) std::copy(keep.begin(), keep.end(), ids.begin() + lo);
  } else {
    double ext = -1;
    for (int a = 0; a < 3; a++) {
      double cmn = 1e300, cmx = -1e300;
      for (int k = lo; k < hi; k++) { const float4 &q = sph[ids[k]]; const double v = a == 0 ? q.x : a == 1 ? q.y : q.z; cmn = std::min(cmn, v); cmx = std::max(cmx, v); }
      if (cmx - cmn > ext) { ext = cmx - cmn; axis = a; }
    }
    std::nth_element(ids.begin() + lo, ids.begin() + mid, ids.begin() + hi, [&](int x, int y) {
      const float4 &a = sph[x], &b = sph[y];
      return (axis == 0 ? a.x : axis == 1 ? a.y : a.z) < (axis == 0 ? b.x : axis == 1 ? b.y : b.z);
    });
  }
  int l = 0, r = 0;
  if (!bvh_build_rec(sph, ids, lo, mid, depth + 1, out, &l)) return false;
  if (!bvh_build_rec(sph, ids, mid, hi, depth + 1, out, &r)) return false;
  if (r >= (1 << 28)) return false;
  out.kids[me] = make_int2(l, r | (axis << 28));   // left = lower half along `axis`
  return true;
}

// Children's bounding spheres stored with the parent ({left, right} at 2k, 2k+1), so a traversal step fetches both
// with one index; `per_node` (optional) is copied the same way.
void bvh_child_pairs(const SphereBvh &b, const std::vector<float4> &per_node, std::vector<float4> &pair) {
  pair.assign(2 * b.node.size() + 2, make_float4(0.f, 0.f, 0.f, 0.f));
  for (size_t k = 0; k < b.node.size(); k++) {
    const int2 kid = b.kids[k];
    if (kid.x < 0) continue;
    pair[2 * k] = per_node[kid.x];
    pair[2 * k + 1] = per_node[kid.y & 0x0fffffff];
  }
}

bool build_sphere_bvh(const std::vector<float4> &sph, int n_sph, SphereBvh &out) {
  std::vector<int> ids(n_sph);
  for (int i = 0; i < n_sph; i++) ids[i] = i;
  int root = 0;
  return n_sph > 0 && bvh_build_rec(sph, ids, 0, n_sph, 0, out, &root);
}

// Index over the ellipsoids among a list of quadrics (rt_qindex.cuh).  order: positions of the input list, the
// n_idx certified ellipsoids first in the leaf order of their bounding-sphere hierarchy, then everything else in
// input order.  marg: per node {miss_rel, behind_abs, t_rel, t_abs}, the largest margins its members need.
struct QLbItem { double K, lmin, a0, a1, a2, shift; };
struct QuadricIndex {
  std::vector<int> order;
  int n_idx = 0;
  SphereBvh bvh;
  std::vector<float4> marg;
  float behind_rel = 0.f;
  std::vector<float4> slot_ball;     // bounding sphere of indexed slot k (leaf order) and its level-set constants, for the light buffer
  std::vector<QLbItem> slot_q;
};

void build_quadric_index(const std::vector<const float *> &params, QuadricIndex &out) {
  std::vector<int> idxd, rest;
  std::vector<float4> balls, marg;
  std::vector<QLbItem> qits;
  out = QuadricIndex();
  for (size_t k = 0; k < params.size(); k++) {
    QIndexItem it;
    qindex_classify(params[k], it);
    // an ellipsoid whose certified bounding sphere is out of proportion gains nothing from the index
    if (it.ok && it.r_item <= 4.0 * sqrt(it.K / it.lmin)) {
      idxd.push_back((int)k);
      balls.push_back(make_float4(it.centre[0], it.centre[1], it.centre[2], it.r_item));
      marg.push_back(make_float4(it.miss_rel, it.behind_abs, it.t_rel, it.t_abs));
      {
        const double sx = it.c[0] - it.centre[0], sy = it.c[1] - it.centre[1], sz = it.c[2] - it.centre[2];
        qits.push_back(QLbItem{it.K, it.lmin, it.a0, it.a1, it.a2, sqrt(sx * sx + sy * sy + sz * sz)});
      }
      out.behind_rel = std::max(out.behind_rel, it.behind_rel);
    } else rest.push_back((int)k);
  }
  if (idxd.size() >= 8 && build_sphere_bvh(balls, (int)balls.size(), out.bvh)) {
    out.n_idx = (int)idxd.size();
    for (int k : out.bvh.items) { out.order.push_back(idxd[k]); out.slot_ball.push_back(balls[k]); out.slot_q.push_back(qits[k]); }
    out.order.insert(out.order.end(), rest.begin(), rest.end());
    // children are created after their parent, so one backward sweep folds the members' margins upwards
    auto fold = [](float4 x, float4 y) { return make_float4(std::max(x.x, y.x), std::max(x.y, y.y), std::max(x.z, y.z), std::max(x.w, y.w)); };
    out.marg.assign(out.bvh.node.size(), make_float4(0.f, 0.f, 0.f, 0.f));
    for (int k = (int)out.bvh.node.size() - 1; k >= 0; k--) {
      const int2 kid = out.bvh.kids[k];
      if (kid.x >= 0) out.marg[k] = fold(out.marg[kid.x], out.marg[kid.y & 0x0fffffff]);
      else for (int i = 0; i < kid.y; i++) out.marg[k] = fold(out.marg[k], marg[out.bvh.items[-kid.x - 1 + i]]);
    }
  } else {
    out.bvh = SphereBvh();
    out.behind_rel = 0.f;
    out.order.resize(params.size());
    for (size_t k = 0; k < params.size(); k++) out.order[k] = (int)k;
  }
}

// qit = NULL: the items are spheres (inflation of the reject-filter contract, above).  qit != NULL: the items are
// indexed ellipsoids (rt_qindex.cuh): the reference can only report ellipsoid s if the line enters {q <= delta(|P - c|)},
// which lies within sqrt((K + a0 + a1 s + a2 s^2)/lmin) of the centre, s = |P - c| <= Dcap + |c - L|; only the four
// nearest distance levels are built for them (farther out the inflated ellipsoids fill too many cells: hierarchy).
// scene = NULL: the same cube map for the primary rays -- one pseudo light at the camera (0,0,0), one level with
// Dcap = 0 (a primary ray starts AT the apex, so T = |C|^2 exactly): level 0 of "light" 0 lists what a ray leaving
// the origin in that cell's directions can be reported to hit.
bool build_light_buffer(const rt_scene *scene, const std::vector<float4> &sph, int n_sph, const std::vector<QLbItem> *qit, LightBuffer &lb) {
  const bool camera = scene == nullptr;
  const int R = RT_LB_RES, CELLS = 6 * R * R, nl = camera ? 1 : scene->num_lights;
  std::vector<double> cb(3 * (size_t)CELLS), cbeta(CELLS), ccos(CELLS), csin(CELLS);
  for (int face = 0; face < 6; face++)
    for (int j = 0; j < R; j++)
      for (int i = 0; i < R; i++) {
        const int c = (face * R + j) * R + i;
        const double s0 = -1.0 + 2.0 * i / R, s1 = -1.0 + 2.0 * (i + 1) / R, t0 = -1.0 + 2.0 * j / R, t1 = -1.0 + 2.0 * (j + 1) / R;
        double b[3], q[3], beta = 0;
        lb_dir(face, 0.5 * (s0 + s1), 0.5 * (t0 + t1), b);
        const double cs[4][2] = {{s0, t0}, {s1, t0}, {s0, t1}, {s1, t1}};
        for (auto &k : cs) {
          lb_dir(face, k[0], k[1], q);
          double d = b[0] * q[0] + b[1] * q[1] + b[2] * q[2];
          d = d > 1 ? 1 : d < -1 ? -1 : d;
          beta = std::max(beta, acos(d));
        }
        cb[3 * c] = b[0]; cb[3 * c + 1] = b[1]; cb[3 * c + 2] = b[2]; cbeta[c] = beta;
        ccos[c] = cos(beta); csin[c] = sin(beta);
      }
  // 4 x 4 blocks of cells: centre direction and angular radius, to skip 16 cells with one test
  const int RB = R / 4;
  std::vector<double> bb(3 * (size_t)6 * RB * RB), bbeta((size_t)6 * RB * RB);
  for (int face = 0; face < 6; face++)
    for (int j = 0; j < RB; j++)
      for (int i = 0; i < RB; i++) {
        const double s0 = -1.0 + 8.0 * i / R, s1 = -1.0 + 8.0 * (i + 1) / R, t0 = -1.0 + 8.0 * j / R, t1 = -1.0 + 8.0 * (j + 1) / R;
        double b[3], q[3], beta = 0;
        lb_dir(face, 0.5 * (s0 + s1), 0.5 * (t0 + t1), b);
        const double cs[4][2] = {{s0, t0}, {s1, t0}, {s0, t1}, {s1, t1}};
        for (auto &k : cs) {
          lb_dir(face, k[0], k[1], q);
          double d = b[0] * q[0] + b[1] * q[1] + b[2] * q[2];
          d = d > 1 ? 1 : d < -1 ? -1 : d;
          beta = std::max(beta, acos(d));
        }
        const int blk = (face * RB + j) * RB + i;
        bb[3 * blk] = b[0]; bb[3 * blk + 1] = b[1]; bb[3 * blk + 2] = b[2]; bbeta[blk] = beta + 1e-9;
      }
  lb.caps.assign((size_t)nl * RT_LB_LEVELS + 1, -1.f);
  std::vector<std::vector<unsigned>> cnt((size_t)nl * RT_LB_LEVELS);           // per (light, level): CSR counts
  std::vector<std::vector<std::pair<int, int>>> pairs((size_t)nl * RT_LB_LEVELS);   // (cell, sphere)
  auto work = [&](int l) {
    double L[3] = {0.0, 0.0, 0.0};
    if (!camera) { const rt_light &t = scene->lights[l]; L[0] = t.position[0]; L[1] = t.position[1]; L[2] = t.position[2]; }
    if (!std::isfinite(L[0]) || !std::isfinite(L[1]) || !std::isfinite(L[2])) return;
    double Rl = 0;
    for (int s = 0; s < n_sph; s++) {
      const double dx = sph[s].x - L[0], dy = sph[s].y - L[1], dz = sph[s].z - L[2], r = fabs((double)sph[s].w);
      const double d = sqrt(dx * dx + dy * dy + dz * dz) + r;
      if (std::isfinite(d)) Rl = std::max(Rl, d);
    }
    if (!(Rl > 0) || !(Rl < 1e15)) return;
    // (the last level used to reach 32 x: its inflated spheres filled 25-58 cells' worth of list per cell and were 70 % of
    // all entries; the few hit points beyond 16 x walk the hierarchy instead)
    double mult[RT_LB_LEVELS] = {0.35, 0.7, 1.4, 2.8, 6.0, 16.0};
    if (const char *e = getenv("RT_LB_FAR")) { double x = 0, y = 0; if (sscanf(e, "%lf,%lf", &x, &y) == 2 && x > 2.8 && y > x) { mult[4] = x; mult[5] = y; } }
    float capf[RT_LB_LEVELS];
    for (int lv = 0; lv < RT_LB_LEVELS; lv++) {
      const double Dcap = camera ? 0.0 : Rl * mult[lv];
      capf[lv] = (float)((Dcap * (1.0 - 1e-3)) * (Dcap * (1.0 - 1e-3)));
      if ((qit && lv >= 4) || (camera && lv >= 1)) { capf[lv] = -1.f; continue; }
      auto &pv = pairs[(size_t)l * RT_LB_LEVELS + lv];
      for (int s = 0; s < n_sph; s++) {
        const double dx = sph[s].x - L[0], dy = sph[s].y - L[1], dz = sph[s].z - L[2], r = fabs((double)sph[s].w);
        const double dist = sqrt(dx * dx + dy * dy + dz * dz);
        const double Tcap = (Dcap + dist) * (Dcap + dist);
        double reff = sqrt(r * r + ldexp(1.0, -18) * (Tcap + r * r)) + ldexp(8.0, -24) * sqrt(Tcap);
        if (qit) {
          const QLbItem &e = (*qit)[s];
          const double sm = Dcap + dist + 2.0 * e.shift;
          reff = (sqrt((e.K + e.a0 + e.a1 * sm + e.a2 * sm * sm) / e.lmin) + e.shift) * (1.0 + 1e-6) + 1e-7 * sm;
        }
        const bool all = !(dist > reff * (1.0 + 1e-6)) || !std::isfinite(reff);
        double a[3] = {0, 0, 0}, theta = 4.0;
        if (!all) { a[0] = dx / dist; a[1] = dy / dist; a[2] = dz / dist; theta = asin(std::min(1.0, reff / dist)) + 3e-4; }
        for (int face = 0; face < 6; face++) {
          if (!all) {   // whole face out of reach? (a cube face spans acos(1/sqrt 3) = 0.9554 rad around its axis)
            const double ax = fabs(a[face >> 1]), reach = theta + 0.9555;
            if (reach < 1.5707 && ax < cos(reach)) continue;
          }
          const double ct = all ? 0 : cos(theta), st = all ? 0 : sin(theta);
          for (int bj = 0; bj < RB; bj++)
            for (int bi = 0; bi < RB; bi++) {
              if (!all) {   // whole block out of reach of both +a and -a?
                const int blk = (face * RB + bj) * RB + bi;
                const double ang = theta + bbeta[blk];
                if (ang < 3.14159) {
                  const double dd = bb[3 * blk] * a[0] + bb[3 * blk + 1] * a[1] + bb[3 * blk + 2] * a[2];
                  const double ca = cos(ang);
                  if (ca > 0 && fabs(dd) < ca) continue;
                }
              }
              for (int cj = 4 * bj; cj < 4 * bj + 4; cj++)
                for (int ci = 4 * bi; ci < 4 * bi + 4; ci++) {
                  const int c = (face * R + cj) * R + ci;
                  bool in = all;
                  if (!in) {
                    if (theta + cbeta[c] >= 3.14159) in = true;
                    else {
                      const double ca = ct * ccos[c] - st * csin[c];   // cos(theta + beta_c)
                      const double dd = cb[3 * c] * a[0] + cb[3 * c + 1] * a[1] + cb[3 * c + 2] * a[2];
                      in = ca <= 0 || fabs(dd) >= ca - 1e-12;
                    }
                  }
                  if (in) pv.push_back(std::make_pair(c, s));
                }
            }
        }
      }
    }
    for (int lv = 0; lv < RT_LB_LEVELS; lv++) lb.caps[(size_t)l * RT_LB_LEVELS + lv] = capf[lv];
  };
  {
    std::vector<std::thread> th;
    for (int l = 0; l < nl; l++) th.emplace_back(work, l);
    for (auto &x : th) x.join();
  }
  size_t total = 0;
  for (auto &pv : pairs) total += pv.size();
  if (total > ((size_t)64 << 20)) return false;   // degenerate scene (NaN geometry everywhere): no light buffer
  lb.start.assign((size_t)nl * RT_LB_LEVELS * CELLS + 1, 0u);
  lb.items.resize(total + 1);
  size_t base = 0;
  for (size_t g = 0; g < pairs.size(); g++) {
    // within a cell: largest first (a shadow query stops at its first occluder, and the order is free)
    std::stable_sort(pairs[g].begin(), pairs[g].end(), [&](const std::pair<int, int> &x, const std::pair<int, int> &y) {
      return fabsf(sph[x.second].w) > fabsf(sph[y.second].w);
    });
    std::vector<unsigned> c((size_t)CELLS + 1, 0u);
    for (auto &pr : pairs[g]) c[pr.first + 1]++;
    for (int k = 0; k < CELLS; k++) c[k + 1] += c[k];
    for (int k = 0; k < CELLS; k++) lb.start[g * CELLS + k] = (unsigned)(base + c[k]);
    std::vector<unsigned> fill(c.begin(), c.end() - 1);
    for (auto &pr : pairs[g]) lb.items[base + fill[pr.first]++] = pr.second;
    base += pairs[g].size();
  }
  lb.start[(size_t)nl * RT_LB_LEVELS * CELLS] = (unsigned)base;
  return true;
}

}  // namespace

// ==================================================================================
extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }
int rt_set_max_depth(int depth) {
  if (depth < 1 || depth > RT_MAX_DEPTH_CAP) return set_err(RT_ERR_ARG, "rt_set_max_depth: depth must be in 1..%d", RT_MAX_DEPTH_CAP);
  g_max_depth = depth;
  return RT_OK;
}
int rt_dbg_qindex_classify(const float *q10, rt_qindex_info *out) {
  if (!q10 || !out) return 0;
  QIndexItem it;
  qindex_classify(q10, it);
  if (!it.ok) return 0;
  for (int k = 0; k < 3; k++) { out->centre[k] = it.c[k]; out->ball[k] = it.centre[k]; }
  out->K = it.K; out->lmin = it.lmin; out->lmax = it.lmax;
  out->delta[0] = it.a0; out->delta[1] = it.a1; out->delta[2] = it.a2;
  out->ball[3] = it.r_item;
  out->consts[0] = it.miss_rel; out->consts[1] = it.behind_rel; out->consts[2] = it.behind_abs;
  out->consts[3] = it.t_rel; out->consts[4] = it.t_abs;
  return 1;
}

int rt_dbg_qindex_prune(const float *node4, const float *P3, const float *d3, int n_rays, int nearest, const float *best_t,
                        const float *consts5, int *codes) {
  const QIndexConsts k = {consts5[0], consts5[1], consts5[2], consts5[3], consts5[4]};
  const float4 nd = make_float4(node4[0], node4[1], node4[2], node4[3]);
  for (int r = 0; r < n_rays; r++)
    codes[r] = qnode_prune(nd, P3[3 * r], P3[3 * r + 1], P3[3 * r + 2], d3[3 * r], d3[3 * r + 1], d3[3 * r + 2], nearest != 0,
                           nearest ? best_t[r] : 0.0f, k);
  return 0;
}

static long long g_dbg_nodes = 0;
long long rt_dbg_qindex_last_nodes(void) { return g_dbg_nodes; }

int rt_dbg_qindex_visit(const float *q10, int n_quadrics, const float *P3, const float *d3, int n_rays, int nearest,
                        const float *best_t, unsigned char *visited) {
  if (!q10 || n_quadrics <= 0 || !visited) return -1;
  std::vector<const float *> params;
  for (int k = 0; k < n_quadrics; k++) params.push_back(q10 + 10 * (size_t)k);
  QuadricIndex qi;
  build_quadric_index(params, qi);
  g_dbg_nodes = 0;
  for (int r = 0; r < n_rays; r++) {
    unsigned char *v = visited + (size_t)r * n_quadrics;
    memset(v, 0, n_quadrics);
    for (int k = qi.n_idx; k < n_quadrics; k++) v[qi.order[k]] = 1;       // not indexed: always scanned
    if (!qi.n_idx) continue;
    std::vector<int> stack(1, 0);
    while (!stack.empty()) {
      const int k = stack.back(); stack.pop_back();
      g_dbg_nodes++;
      const float4 mg = qi.marg[k];
      const QIndexConsts qk = {mg.x, qi.behind_rel, mg.y, mg.z, mg.w};
      if (qnode_prune(qi.bvh.node[k], P3[3 * r], P3[3 * r + 1], P3[3 * r + 2], d3[3 * r], d3[3 * r + 1], d3[3 * r + 2],
                      nearest != 0, nearest ? best_t[r] : 0.0f, qk)) continue;
      const int2 kid = qi.bvh.kids[k];
      if (kid.x >= 0) { stack.push_back(kid.y & 0x0fffffff); stack.push_back(kid.x); }
      else for (int i = 0; i < kid.y; i++) v[qi.order[-kid.x - 1 + i]] = 1;   // slot -kid.x-1+i holds input quadric order[slot]
    }
  }
  return qi.n_idx;
}

long long rt_dbg_light_buffer_sizes(const rt_scene *scene, long long *per_level) {
  if (!scene || scene->num_lights <= 0) return -1;
  std::vector<float4> sph;
  for (int i = 0; i < scene->num_shapes; i++) {
    const rt_shape &s = scene->shapes[i];
    if (s.type == RT_SPHERE) sph.push_back(make_float4(s.position[0], s.position[1], s.position[2], s.params[0]));
  }
  LightBuffer lb;
  if (sph.empty() || !build_light_buffer(scene, sph, (int)sph.size(), nullptr, lb)) return -1;
  const size_t CELLS = (size_t)6 * RT_LB_RES * RT_LB_RES;
  long long total = 0;
  for (int g = 0; g < scene->num_lights * RT_LB_LEVELS; g++) {
    const long long n = (long long)lb.start[(size_t)(g + 1) * CELLS] - (long long)lb.start[(size_t)g * CELLS];
    if (per_level) per_level[g] = lb.caps[g] >= 0.f ? n : -1;
    total += n;
  }
  return total;
}

int rt_set_mode(int mode) {
  if (mode != RT_MODE_ITERATIVE && mode != RT_MODE_SHOOT) return set_err(RT_ERR_ARG, "rt_set_mode: mode must be RT_MODE_ITERATIVE or RT_MODE_SHOOT");
  g_mode = mode;
  return RT_OK;
}

int rt_set_accel(int mode) {
  if (mode != 0 && mode != 1) return set_err(RT_ERR_ARG, "rt_set_accel: mode must be 0 or 1");
  g_accel = mode;
  return RT_OK;
}
const char *rt_last_error(void) { return g_err; }
int rt_device_count(void) { return g_ndev; }
int rt_band_height(void) { return band_rows(); }

int64_t rt_rank_num_rows(int height, int rank, int nranks) {
  if (height <= 0 || nranks <= 0 || rank < 0 || rank >= nranks) return 0;
  const int bh = band_rows();
  int64_t rows = 0;
  for (int b = rank; (int64_t)b * bh < height; b += nranks) {
    int64_t lo = (int64_t)b * bh, hi = lo + bh;
    rows += (hi > height ? height : hi) - lo;
  }
  return rows;
}

int64_t rt_rank_rows(int height, int rank, int nranks, int32_t *rows_out) {
  if (height <= 0 || nranks <= 0 || rank < 0 || rank >= nranks) return 0;
  const int bh = band_rows();
  int64_t n = 0;
  for (int b = rank; (int64_t)b * bh < height; b += nranks)
    for (int y = b * bh; y < (b + 1) * bh && y < height; y++) rows_out[n++] = y;
  return n;
}

void rt_shutdown(void) {
  nccl_unload();
  for (int i = 0; i < RT_MAX_GPUS; i++) free_device(g_dev[i]);
  g_ndev = 0;
  g_have_scene = false;
}

static int init_one(DeviceCtx &d, int ordinal) {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, ordinal));
  if (prop.major != 10)
    return set_err(RT_ERR_NO_DEVICE, "device %d (%s, sm_%d%d) is not a Blackwell sm_100 part; kernels are sm_100a only",
                   ordinal, prop.name, prop.major, prop.minor);
  CK(cudaSetDevice(ordinal));
  d.dev = ordinal;           // from here on free_device() releases whatever exists
  d.sm_count = prop.multiProcessorCount;
  CK(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&d.ev0)); CK(cudaEventCreate(&d.ev1)); CK(cudaEventCreate(&d.ev2));
  CK(cudaMalloc((void **)&d.qcount, (RT_MAX_DEPTH_CAP + 1) * RT_CTL_STRIDE * sizeof(int)));
  CK(cudaMalloc((void **)&d.counters, CT_COUNT * sizeof(unsigned long long)));
  warm_modules();
  return RT_OK;
}

// A failed init leaves nothing behind: devices already set up are released, the first error message is kept.
static int init_fail(int code) {
  char msg[sizeof g_err];
  memcpy(msg, g_err, sizeof msg);
  g_ndev = RT_MAX_GPUS;      // rt_shutdown walks every slot
  rt_shutdown();
  memcpy(g_err, msg, sizeof g_err);
  return code;
}

int rt_init(int n_gpus) {
  if (g_ndev) rt_shutdown();
  trace("rt_init: begin");
  int avail = 0;
  cudaError_t e = cudaGetDeviceCount(&avail);
  if (e != cudaSuccess || avail <= 0)
    return set_err(RT_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU path",
                   e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  if (n_gpus < 0 || n_gpus > RT_MAX_GPUS) return set_err(RT_ERR_ARG, "n_gpus=%d out of range", n_gpus);
  int n = n_gpus == 0 ? (avail > RT_MAX_GPUS ? RT_MAX_GPUS : avail) : n_gpus;
  if (n > avail) return set_err(RT_ERR_NO_DEVICE, "%d GPUs requested, %d visible", n, avail);
  for (int i = 0; i < n; i++) {
    const int r = init_one(g_dev[i], i);
    if (r != RT_OK) return init_fail(r);
  }
  g_ndev = n;
  if (n > 1) {
    for (int i = 0; i < n; i++) {       // peer access lets NCCL use NVLink P2P
      cudaSetDevice(i);
      for (int j = 0; j < n; j++) if (i != j) { cudaDeviceEnablePeerAccess(j, 0); cudaGetLastError(); }
    }
    int r = nccl_load(n);
    if (r != RT_OK) return init_fail(r);
  }
  trace("rt_init: done");
  return RT_OK;
}

// ---- scene upload ------------------------------------------------------------------
int rt_scene_upload(const rt_scene *scene) {
  if (!g_ndev) return set_err(RT_ERR_NO_DEVICE, "rt_scene_upload before rt_init");
  if (!scene || scene->num_shapes < 0 || scene->num_lights < 0 ||
      (scene->num_shapes > 0 && !scene->shapes) || (scene->num_lights > 0 && !scene->lights))
    return set_err(RT_ERR_ARG, "bad scene argument");
  trace("rt_scene_upload: begin");
  const int n = scene->num_shapes, nl = scene->num_lights;
  std::vector<float4> sph, sph_hot, sph_prim, pln, quad, quad_hot, quad_prim, mat(2 * (size_t)n + 1), lights(4 * (size_t)nl + 1);
  std::vector<int> sph_idx, pln_idx, quad_idx;
  std::vector<int2> slot((size_t)n + 1);
  std::vector<int> sphere_order, quad_order;
  for (int i = 0; i < n; i++) {
    const rt_shape &s = scene->shapes[i];
    // opacity, C/raycast.c:383: 1.0 - reflectivity - refractivity in double, stored float
    const float opacity = (float)(1.0 - (double)s.reflectivity - (double)s.refractivity);
    mat[2 * i] = make_float4(s.diffuse[0], s.diffuse[1], s.diffuse[2], s.reflectivity);
    mat[2 * i + 1] = make_float4(s.specular[0], s.specular[1], s.specular[2], opacity);
    if (s.type == RT_SPHERE) {
      sphere_order.push_back(i);
    } else if (s.type == RT_PLANE) {
      slot[i] = make_int2(1, (int)pln_idx.size());
      pln.push_back(make_float4(s.position[0], s.position[1], s.position[2], 0.f));
      pln.push_back(make_float4(s.params[0], s.params[1], s.params[2], 0.f));
      pln_idx.push_back(i);
    } else if (s.type == RT_QUADRIC) {
      quad_order.push_back(i);                                 // laid out below, once the index is decided
    } else {
      return set_err(RT_ERR_ARG, "shape %d has unknown type %d", i, s.type);
    }
  }
  // Spheres in descending radius (stable): big occluders first, so shadow rays die early in the
  // scan and k_shade can drop to a smaller scan variant.  Any order is parity-safe: nearest hit
  // is the lexicographic (t, original index) minimum and a shadow query is any-hit.
  std::stable_sort(sphere_order.begin(), sphere_order.end(), [&](int x, int y) {
    const float rx = scene->shapes[x].params[0], ry = scene->shapes[y].params[0];
    return (rx == rx ? rx : -INFINITY) > (ry == ry ? ry : -INFINITY);   // NaN radii last
  });
  if (getenv("RT_NO_SORT")) std::sort(sphere_order.begin(), sphere_order.end());
  // accel mode: leaf order of the bounding-sphere hierarchy instead, so a leaf is a run of consecutive slots
  const char *env_accel = getenv("RT_ACCEL");
  const bool want_accel = env_accel ? atoi(env_accel) != 0 : g_accel != 0;
  SphereBvh bvh;
  bool have_bvh = false;
  if (want_accel && sphere_order.empty() && !quad_order.empty()) {
    // no spheres, but quadrics the index may hold: an empty hierarchy (one leaf without members) keeps the mode on
    bvh.node.push_back(make_float4(0.f, 0.f, 0.f, 0.f));
    bvh.kids.push_back(make_int2(-1, 0));
    have_bvh = true;
  }
  if (want_accel && !sphere_order.empty()) {
    std::vector<float4> tmp;
    for (int i : sphere_order) { const rt_shape &s = scene->shapes[i]; tmp.push_back(make_float4(s.position[0], s.position[1], s.position[2], s.params[0])); }
    have_bvh = build_sphere_bvh(tmp, (int)tmp.size(), bvh);
    if (have_bvh) {
      std::vector<int> reordered;
      for (int k : bvh.items) reordered.push_back(sphere_order[k]);
      sphere_order.swap(reordered);
    }
  }
  for (int i : sphere_order) {
    const rt_shape &s = scene->shapes[i];
    slot[i] = make_int2(0, (int)sph.size());
    const float r = s.params[0];
    sph.push_back(make_float4(s.position[0], s.position[1], s.position[2], r));
    { volatile float rr = r * r; sph_hot.push_back(make_float4(s.position[0], s.position[1], s.position[2], rr)); }
    // primary rays start at exactly (0,0,0): temp = o - centre and c of C/raycast.c:581,585
    volatile float t0 = 0.0f - s.position[0], t1 = 0.0f - s.position[1], t2 = 0.0f - s.position[2];
    volatile float p0 = t0 * t0, p1 = t1 * t1, p2 = t2 * t2;
    volatile float s01 = p0 + p1;
    volatile float tt = s01 + p2;
    // reject-filter constant k = 2^-18 (T' + r^2) - (T' - r^2), same formula as k_shade's
    (void)tt;
    volatile float rr2 = r * r;
    volatile float q0 = t0 * t0;
    volatile float Tp = fmaf(t2, t2, fmaf(t1, t1, q0));
    volatile float su = Tp + rr2, di = rr2 - Tp;
    sph_prim.push_back(make_float4(t0, t1, t2, fmaf(RT_FILTER_EPS, su, di)));
    sph_idx.push_back(i);
  }
  const size_t slot_count_spheres = sph_idx.size();
  for (int l = 0; l < nl; l++) {
    const rt_light &t = scene->lights[l];
    if (t.type != RT_POINT && t.type != RT_SPOTLIGHT) return set_err(RT_ERR_ARG, "light %d has unknown type %d", l, t.type);
    int ty = t.type; float tyf; memcpy(&tyf, &ty, 4);
    lights[4 * l] = make_float4(t.position[0], t.position[1], t.position[2], tyf);
    lights[4 * l + 1] = make_float4(t.color[0], t.color[1], t.color[2], t.cos_theta);
    lights[4 * l + 2] = make_float4(t.radial[0], t.radial[1], t.radial[2], t.a0);
    // cull flag: a light behind the surface contributes (0 + 0) * rad * ang (C/raycast.c:414-416),
    // which is exactly +0 only when rad and ang are certainly finite: radial a0 > 0, a1, a2 >= 0
    // (denominator >= a0), and for spots an angular exponent evaluated by the integer chain.
    bool cull_ok = t.radial[0] > 0.0f && t.radial[1] >= 0.0f && t.radial[2] >= 0.0f &&
                   std::isfinite(t.radial[0]) && std::isfinite(t.radial[1]) && std::isfinite(t.radial[2]);
    if (t.type == RT_SPOTLIGHT)
      cull_ok = cull_ok && t.a0 >= 0.0f && t.a0 <= 64.0f && (float)(int)t.a0 == t.a0;
    if (getenv("RT_NO_CULL")) cull_ok = false;   // debugging aid: trace every shadow ray
    lights[4 * l + 3] = make_float4(t.direction[0], t.direction[1], t.direction[2], cull_ok ? 1.0f : 0.0f);
  }
  // pad the per-sphere arrays to a multiple of 32 with entries the reject filter never passes
  // (r*r = -inf => k = -inf) so the hot loops can run unrolled without a tail
  const size_t padded = (sph_hot.size() + 31) / 32 * 32 + 4;   // whole 32-sphere words (the scans record survivors per word) + one prefetch group
  while (sph_hot.size() < padded) {
    sph_hot.push_back(make_float4(0.f, 0.f, 0.f, -INFINITY));
    sph.push_back(make_float4(0.f, 0.f, 0.f, 0.f));
    sph_prim.push_back(make_float4(0.f, 0.f, 0.f, -INFINITY));
    sph_idx.push_back(-2);
  }
  const int n_sph_real = (int)slot_count_spheres;
  // Constant-bank image of the reject-filter operands (rt_scene.cuh, c_sph / c_prim): {cx, cy, cz, rho} with
  // rho = fl(r^2) (1 + 2^-18) + 2^-19 (1 + 2^-16) |C|^2, evaluated in double and rounded UP to float, so the
  // filter's margin can only grow; every other entry of the bank never passes (rho = -inf => kappa' = +inf).
  // All spheres when they fit, else the leading whole chunks (the kernels scan the rest from global memory).
  // (the scan reads whole 32-sphere words and prefetches one trip ahead: 40 spare entries)
  const int n_csph = n_sph_real <= RT_CSPH_CAP - 40 ? n_sph_real : ((RT_CSPH_CAP - 40) / RT_CHUNK) * RT_CHUNK;
  std::vector<float4> csph(RT_CSPH_CAP, make_float4(0.f, 0.f, 0.f, -INFINITY));
  std::vector<float> cprim(RT_CSPH_CAP, -INFINITY);
  for (int k = 0; k < n_csph; k++) {
    const float4 h = sph_hot[k];
    const double c2 = (double)h.x * h.x + (double)h.y * h.y + (double)h.z * h.z;
    const double rho = (double)h.w * (1.0 + 3.814697265625e-06) + 1.9073486328125e-06 * (1.0 + 1.52587890625e-05) * c2;
    float rf = (float)rho;
    if ((double)rf < rho) rf = nextafterf(rf, INFINITY);
    csph[k] = make_float4(h.x, h.y, h.z, rf);   // NaN radius or centre: NaN rho, the test is false, as before
    cprim[k] = sph_prim[k].w;
  }
  // light buffer (opt-in: rt_set_accel(1) or RT_ACCEL=1)
  LightBuffer lb;
  const bool have_lb = have_bvh && nl > 0 && build_light_buffer(scene, sph, n_sph_real, nullptr, lb);
  // Quadrics: file order; in accel mode the ellipsoids the index can certify (rt_qindex.cuh) come first, in
  // the leaf order of their own bounding-sphere hierarchy.  Any order is parity-safe, as for the spheres.
  QuadricIndex qi;
  if (have_lb && !quad_order.empty() && !getenv("RT_NO_QINDEX")) {
    std::vector<const float *> params;
    for (int i : quad_order) params.push_back(scene->shapes[i].params);
    build_quadric_index(params, qi);
    std::vector<int> reordered;
    for (int k : qi.order) reordered.push_back(quad_order[k]);
    quad_order.swap(reordered);
  }
  const int n_qidx = qi.n_idx;
  // the cube map of the primary rays (apex = the camera) over spheres and ellipsoids -- usable only if its one level
  // was really built: a scene of absurd extent leaves caps[0] at -1 and no lists -- and the light buffer over the
  // indexed ellipsoids; the three builds are independent, one host thread each (plus one per light inside)
  LightBuffer cam, camq, lbq;
  bool have_cam = false, have_camq = false, have_lbq = false;
  {
    const bool camgrid = !getenv("RT_NO_CAMGRID"), qlb = !getenv("RT_NO_QLB");
    std::thread t_cam([&] {
      have_cam = have_lb && n_sph_real > 0 && camgrid && build_light_buffer(nullptr, sph, n_sph_real, nullptr, cam) && cam.caps[0] >= 0.f;
    });
    std::thread t_camq([&] {
      have_camq = have_lb && n_qidx > 0 && camgrid && build_light_buffer(nullptr, qi.slot_ball, n_qidx, &qi.slot_q, camq) && camq.caps[0] >= 0.f;
    });
    std::thread t_lbq([&] { have_lbq = n_qidx > 0 && qlb && build_light_buffer(scene, qi.slot_ball, n_qidx, &qi.slot_q, lbq); });
    t_cam.join(); t_camq.join(); t_lbq.join();
  }
  const SphereBvh &qbvh = qi.bvh;
  const std::vector<float4> &qbvh_marg = qi.marg;
  const float q_behind_rel = qi.behind_rel;
  for (int i : quad_order) {
    const float *q = scene->shapes[i].params;
    slot[i] = make_int2(2, (int)quad_idx.size());
    quad.push_back(make_float4(q[0], q[1], q[2], q[3]));
    quad.push_back(make_float4(q[4], q[5], q[6], q[7]));
    quad.push_back(make_float4(q[8], q[9], 0.f, 0.f));
    const float S2 = fabsf(q[0]) + fabsf(q[1]) + fabsf(q[2]) + fabsf(q[3]) + fabsf(q[4]) + fabsf(q[5]);
    const float S1 = fabsf(q[6]) + fabsf(q[7]) + fabsf(q[8]);
    const float4 h0 = make_float4(q[0], q[1], q[2], q[3]), h1 = make_float4(q[4], q[5], q[6], q[7]),
                 h2 = make_float4(q[8], q[9], S2 * 1.000001f, S1 * 1.000001f);
    quad_hot.push_back(h0); quad_hot.push_back(h1); quad_hot.push_back(h2);
    const QuadFilter f = quad_filter_setup(h0, h1, h2, 0.0f, 0.0f, 0.0f);
    quad_prim.push_back(make_float4(f.As, f.Bs, f.Cs, f.Ds));
    quad_prim.push_back(make_float4(f.Es, f.Fs, f.m0, f.eb));
    quad_prim.push_back(make_float4(f.gx, f.gy, f.gz, 0.f));
    quad_idx.push_back(i);
  }
  for (int k = 0; k < 3; k++) quad_hot.push_back(make_float4(0.f, 0.f, 0.f, 0.f));   // prefetch spare
  // Bounding balls for the pre-test of the quadric scan (rt_kernels.cuh, "Quadrics"; claim (1) of rt_qindex.cuh).  A certified
  // ellipsoid cannot be reported for a line that passes its (float) centre farther than R = r_item (1 + 1e-4) + M reach,
  // reach = |P - c| + r_item >= the distance to the exact centre, M = the largest miss_rel of the scene (float slack of the
  // test + rho) -- the node test of the index, qnode_prune, with the node = the item.  With s = |P - c|: R = a + M s,
  // a = r_item (1 + 1e-4 + M).  The device evaluates s from above (T rsqrt(T) (1 + 2^-20)) and rejects the quadric for every
  // ray of the hit point when  max_j e_j^2 < T (1 - 2^-16) - R^2  (dist(line, c)^2 = T - e^2; 2^-16 T covers the roundings
  // of T and e and |d| != 1, as the 1e-5 Tn of qnode_prune).  Every other quadric gets a = +inf (always passes), pads -inf.
  std::vector<float4> quad_ball, quad_ball_prim;
  float quad_ball_eps = 0.0f;
  {
    std::vector<QIndexItem> items(quad_order.size());
    double M = 0.0;
    const bool use_balls = !getenv("RT_NO_QBALL");
    for (size_t k = 0; k < quad_order.size(); k++) {
      items[k].ok = false;
      if (use_balls) qindex_classify(scene->shapes[quad_order[k]].params, items[k]);
      if (items[k].ok) M = std::max(M, (double)items[k].miss_rel);
    }
    quad_ball_eps = (float)(M * (1.0 + 1e-6));
    if ((double)quad_ball_eps < M) quad_ball_eps = nextafterf(quad_ball_eps, INFINITY);
    for (size_t k = 0; k < quad_order.size(); k++) {
      float4 e = make_float4(0.f, 0.f, 0.f, INFINITY);
      if (items[k].ok) {
        const double a = (double)items[k].r_item * (1.0 + 1e-4 + (double)quad_ball_eps) * (1.0 + 1e-6);
        float af = (float)a;
        if ((double)af < a) af = nextafterf(af, INFINITY);
        if (std::isfinite(af)) e = make_float4(items[k].centre[0], items[k].centre[1], items[k].centre[2], af);
      }
      quad_ball.push_back(e);
      // primary rays leave (0,0,0): T = |c|^2 and R are constants of the quadric; kappa = T (1 - 2^-16) - R^2 rounded
      // down (a smaller kappa lets more through), stored negated; +inf = always passes
      float4 ep = make_float4(0.f, 0.f, 0.f, INFINITY);
      if (std::isfinite(e.w)) {
        const double T = (double)e.x * e.x + (double)e.y * e.y + (double)e.z * e.z;
        const double R = (double)e.w + (double)quad_ball_eps * sqrt(T) * (1.0 + 1e-6);
        const double kap = T * (1.0 - 1.52587890625e-05) * (1.0 - 1e-6) - R * R * (1.0 + 1e-6);
        float nk = (float)(-kap);
        if ((double)nk < -kap) nk = nextafterf(nk, INFINITY);
        if (std::isfinite(nk)) ep = make_float4(e.x, e.y, e.z, nk);
      }
      quad_ball_prim.push_back(ep);
    }
    const size_t padded_q = (quad_ball.size() + 31) / 32 * 32 + 4;
    while (quad_ball.size() < padded_q) quad_ball.push_back(make_float4(0.f, 0.f, 0.f, -INFINITY));
  }
  // one blob, 256-byte aligned sections
  struct Sec { const void *src; size_t bytes; size_t off; };
  std::vector<float4> lb_hot;
  if (have_lb) { lb_hot.resize(lb.items.size()); for (size_t k = 0; k + 1 < lb.items.size(); k++) lb_hot[k] = sph_hot[lb.items[k]]; }
  std::vector<float4> bvh_pair, qbvh_pair, qbvh_mpair;
  if (have_lb) bvh_child_pairs(bvh, bvh.node, bvh_pair);
  if (n_qidx) { bvh_child_pairs(qbvh, qbvh.node, qbvh_pair); bvh_child_pairs(qbvh, qbvh_marg, qbvh_mpair); }
  Sec secs[34] = {
      {sph.data(), sph.size() * 16, 0},       {sph_prim.data(), sph_prim.size() * 16, 0},
      {sph_idx.data(), sph_idx.size() * 4, 0}, {pln.data(), pln.size() * 16, 0},
      {pln_idx.data(), pln_idx.size() * 4, 0}, {quad.data(), quad.size() * 16, 0},
      {quad_idx.data(), quad_idx.size() * 4, 0}, {mat.data(), mat.size() * 16, 0},
      {slot.data(), slot.size() * 8, 0},       {lights.data(), lights.size() * 16, 0},
      {sph_hot.data(), sph_hot.size() * 16, 0}, {quad_hot.data(), quad_hot.size() * 16, 0},
      {quad_prim.data(), quad_prim.size() * 16, 0},
      {lb.start.data(), have_lb ? lb.start.size() * 4 : 0, 0}, {lb.items.data(), have_lb ? lb.items.size() * 4 : 0, 0},
      {lb.caps.data(), have_lb ? lb.caps.size() * 4 : 0, 0},
      {bvh.node.data(), have_lb ? bvh.node.size() * 16 : 0, 0}, {bvh.kids.data(), have_lb ? bvh.kids.size() * 8 : 0, 0},
      {lb_hot.data(), have_lb ? lb_hot.size() * 16 : 0, 0},
      {qbvh.node.data(), n_qidx ? qbvh.node.size() * 16 : 0, 0}, {qbvh.kids.data(), n_qidx ? qbvh.kids.size() * 8 : 0, 0},
      {qbvh_marg.data(), n_qidx ? qbvh_marg.size() * 16 : 0, 0},
      {lbq.start.data(), have_lbq ? lbq.start.size() * 4 : 0, 0}, {lbq.items.data(), have_lbq ? lbq.items.size() * 4 : 0, 0},
      {lbq.caps.data(), have_lbq ? lbq.caps.size() * 4 : 0, 0},
      {bvh_pair.data(), have_lb ? bvh_pair.size() * 16 : 0, 0}, {qbvh_pair.data(), n_qidx ? qbvh_pair.size() * 16 : 0, 0},
      {qbvh_mpair.data(), n_qidx ? qbvh_mpair.size() * 16 : 0, 0},
      {cam.start.data(), have_cam ? cam.start.size() * 4 : 0, 0}, {cam.items.data(), have_cam ? cam.items.size() * 4 : 0, 0},
      {camq.start.data(), have_camq ? camq.start.size() * 4 : 0, 0}, {camq.items.data(), have_camq ? camq.items.size() * 4 : 0, 0},
      {quad_ball.data(), quad_ball.size() * 16, 0}, {quad_ball_prim.data(), quad_ball_prim.size() * 16, 0}};
  size_t total = 0;
  for (auto &s : secs) { s.off = total; total += (s.bytes + 255) & ~(size_t)255; }
  if (total == 0) total = 256;
  std::vector<char> blob(total, 0);
  for (auto &s : secs) if (s.bytes) memcpy(blob.data() + s.off, s.src, s.bytes);

  for (int i = 0; i < g_ndev; i++) {
    DeviceCtx &d = g_dev[i];
    CK(cudaSetDevice(d.dev));
    CK(cudaStreamSynchronize(d.stream));
    if (d.scene_blob) { cudaFree(d.scene_blob); d.scene_blob = nullptr; }
    CK(cudaMalloc(&d.scene_blob, total));
    CK(cudaMemcpyAsync(d.scene_blob, blob.data(), total, cudaMemcpyHostToDevice, d.stream));
    CK(cudaMemcpyToSymbolAsync(c_sph, csph.data(), sizeof(float4) * RT_CSPH_CAP, 0, cudaMemcpyHostToDevice, d.stream));
    CK(cudaMemcpyToSymbolAsync(c_prim, cprim.data(), sizeof(float) * RT_CSPH_CAP, 0, cudaMemcpyHostToDevice, d.stream));
    CK(cudaStreamSynchronize(d.stream));
    char *b = (char *)d.scene_blob;
    DevScene &sc = d.sc;
    sc.n_shapes = n; sc.n_spheres = n_sph_real; sc.n_planes = (int)pln_idx.size();
    sc.n_quadrics = (int)quad_idx.size(); sc.n_lights = nl;
    sc.cam_w = scene->camera_width; sc.cam_h = scene->camera_height;
    sc.sph = (const float4 *)(b + secs[0].off); sc.sph_prim = (const float4 *)(b + secs[1].off);
    sc.sph_idx = (const int *)(b + secs[2].off); sc.pln = (const float4 *)(b + secs[3].off);
    sc.pln_idx = (const int *)(b + secs[4].off); sc.quad = (const float4 *)(b + secs[5].off);
    sc.quad_idx = (const int *)(b + secs[6].off); sc.mat = (const float4 *)(b + secs[7].off);
    sc.slot = (const int2 *)(b + secs[8].off); sc.lights = (const float4 *)(b + secs[9].off);
    sc.sph_hot = (const float4 *)(b + secs[10].off);
    sc.quad_hot = (const float4 *)(b + secs[11].off); sc.quad_prim = (const float4 *)(b + secs[12].off);
    sc.lb_start = have_lb ? (const unsigned *)(b + secs[13].off) : nullptr;
    sc.lb_items = have_lb ? (const int *)(b + secs[14].off) : nullptr;
    sc.lb_caps = have_lb ? (const float *)(b + secs[15].off) : nullptr;
    sc.bvh_node = have_lb ? (const float4 *)(b + secs[16].off) : nullptr;
    sc.bvh_kids = have_lb ? (const int2 *)(b + secs[17].off) : nullptr;
    sc.lb_hot = have_lb ? (const float4 *)(b + secs[18].off) : nullptr;
    sc.qbvh_node = n_qidx ? (const float4 *)(b + secs[19].off) : nullptr;
    sc.qbvh_kids = n_qidx ? (const int2 *)(b + secs[20].off) : nullptr;
    sc.qbvh_marg = n_qidx ? (const float4 *)(b + secs[21].off) : nullptr;
    sc.lbq_start = have_lbq ? (const unsigned *)(b + secs[22].off) : nullptr;
    sc.lbq_items = have_lbq ? (const int *)(b + secs[23].off) : nullptr;
    sc.lbq_caps = have_lbq ? (const float *)(b + secs[24].off) : nullptr;
    sc.bvh_pair = have_lb ? (const float4 *)(b + secs[25].off) : nullptr;
    sc.qbvh_pair = n_qidx ? (const float4 *)(b + secs[26].off) : nullptr;
    sc.qbvh_mpair = n_qidx ? (const float4 *)(b + secs[27].off) : nullptr;
    sc.cam_start = have_cam ? (const unsigned *)(b + secs[28].off) : nullptr;
    sc.cam_items = have_cam ? (const int *)(b + secs[29].off) : nullptr;
    sc.camq_start = have_camq ? (const unsigned *)(b + secs[30].off) : nullptr;
    sc.camq_items = have_camq ? (const int *)(b + secs[31].off) : nullptr;
    sc.n_qidx = n_qidx;
    sc.q_behind_rel = q_behind_rel;
    sc.quad_ball = (const float4 *)(b + secs[32].off);
    sc.quad_ball_prim = (const float4 *)(b + secs[33].off);
    sc.quad_ball_eps = quad_ball_eps;
    sc.n_csph = getenv("RT_NO_CONST_SCAN") ? 0 : n_csph;   // debugging aid: force the global-operand loop
    sc.max_depth = g_max_depth;
    const int pr = prepare_kernels(d);
    if (pr != RT_OK) return pr;
  }
  trace("rt_scene_upload: scene resident, kernels of this scene loaded");
  if (getenv("RT_VERBOSE"))
    fprintf(stderr, "rt_scene_upload: %d spheres, %d planes, %d quadrics (%d indexed), light buffer %s (%zu items), ellipsoid light buffer %s (%zu items), camera grid %zu + %zu items, blob %zu bytes\n",
            n_sph_real, (int)pln_idx.size(), (int)quad_idx.size(), n_qidx, have_lb ? "yes" : "no", have_lb ? lb.items.size() : (size_t)0,
            have_lbq ? "yes" : "no", have_lbq ? lbq.items.size() : (size_t)0, have_cam ? cam.items.size() : (size_t)0,
            have_camq ? camq.items.size() : (size_t)0, total);
  g_have_scene = true;
  g_n_lights = nl;
  g_n_shapes = n;
  return RT_OK;
}

}  // extern "C" (reopened below)

namespace {

typedef void (*shade_fn_t)(DevScene, int, HitQueue, HitQueue, float4 *, float4 *, unsigned long long *, float4 *);

// The k_shade instantiation for this scene: LB lights per pass, with / without the quadric sections,
// with / without the light buffer.
template <int LB>
shade_fn_t shade_fn(const DeviceCtx &d) {
  const bool q = d.sc.n_quadrics > 0, a = d.sc.lb_start != nullptr;
  return q ? (a ? k_shade<LB, true, true> : k_shade<LB, true, false>) : (a ? k_shade<LB, false, true> : k_shade<LB, false, false>);
}

shade_fn_t shade_fn_for(const DeviceCtx &d, int lb) {
  switch (lb) {
    case 1: return shade_fn<1>(d);
    case 2: return shade_fn<2>(d);
    case 4: return shade_fn<4>(d);
    case 8: return shade_fn<8>(d);
    default: return shade_fn<16>(d);
  }
}

int pick_lb(int n_lights) {
  const char *e = getenv("RT_LIGHT_BATCH");   // tuning knob: lights per pass (1,2,4,8,16)
  if (e) { int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) return v; }
  const int c[5] = {1, 2, 4, 8, 16};
  for (int v : c) if (n_lights <= v) return v;
  return 16;
}

// Everything about the launches that depends only on the resident scene: which k_shade instantiation, its
// occupancy grid, the grids of k_primary / k_finalize.  Runs at rt_scene_upload, so the (lazy) module load of the
// scene's k_shade variant and the occupancy queries are paid there and not inside the first render
// (round-1 finding: a one-shot `raytrace` run spent its time in exactly these).
int prepare_kernels(DeviceCtx &d) {
  d.lb = pick_lb(d.sc.n_lights);
  shade_fn_t fn = shade_fn_for(d, d.lb);
  d.shade = (void *)fn;
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 128, 0));
  d.grid_shade = (per_sm > 0 ? per_sm : 1) * d.sm_count;
  if (d.sc.bvh_node) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_primary<true>, 256, 0));
  else CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_primary<false>, 256, 0));
  d.grid_primary = (per_sm > 0 ? per_sm : 1) * d.sm_count;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_finalize, 256, 0));
  d.grid_fin = (per_sm > 0 ? per_sm : 1) * d.sm_count;
  return RT_OK;
}

struct RankPlan { int64_t rows; size_t bytes; int launches; };

int mark(DeviceCtx &d, char kind) {
  size_t i = d.evkind.size();
  if (i >= d.evpool.size()) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    d.evpool.push_back(e);
  }
  d.evkind.push_back(kind);
  CK(cudaEventRecord(d.evpool[i], d.stream));
  return RT_OK;
}

// Enqueue the whole render of `rank`'s rows on device d (asynchronous).
int enqueue_rank(DeviceCtx &d, int W, int H, int rank, int nranks, uint8_t *dst, RankPlan *plan) {
  CK(cudaSetDevice(d.dev));
  const int bh = band_rows();
  const int64_t rows = rt_rank_num_rows(H, rank, nranks);
  const long long px_total = (long long)rows * W;
  plan->rows = rows; plan->bytes = (size_t)px_total * 3; plan->launches = 0;
  size_t tile = tile_pixels_cap();
  if ((long long)tile > px_total) tile = (size_t)(px_total > 0 ? px_total : 1);
  const bool shoot = g_mode == RT_MODE_SHOOT;
  if (shoot) {
    // RT_MODE_SHOOT keeps one {colour, reflectivity} record per level and pixel of a tile: bound that buffer to 2 GiB
    // whatever the depth (rt_set_max_depth(50) with the default 16 M pixel tile would ask for 13 GB) by shrinking the tile
    const size_t cap = (((size_t)2 << 30) / (sizeof(float4) * (size_t)g_max_depth)) & ~(size_t)1023;
    if (tile > cap) tile = cap;
  }
  if (tile > d.tile_cap || !d.acc) {
    // one allocation for the whole per-tile working set (acc, c0, two queues, class lists)
    CK(cudaStreamSynchronize(d.stream));
    if (d.acc) CK(cudaFree(d.acc));
    d.acc = d.c0 = d.qbuf = nullptr; d.qcls = nullptr; d.tile_cap = 0;
    const size_t bytes = tile * (8 * sizeof(float4) + 2 * RT_NCLS * sizeof(int));
    char *base = nullptr;
    CK(cudaMalloc((void **)&base, bytes));
    d.acc = (float4 *)base;
    d.c0 = d.acc + tile;
    d.qbuf = d.c0 + tile;
    d.qcls = (int *)(d.qbuf + 6 * tile);
    d.tile_cap = tile;
    trace("enqueue_rank: tile working set allocated");
  }
  if (!d.shade) return set_err(RT_ERR_NO_SCENE, "render before rt_scene_upload");
  const size_t lvl_stride = tile;   // pixels per level slice of d.lvl (this render's tile, not the allocated capacity)
  if (shoot) {
    const size_t need = (size_t)g_max_depth * lvl_stride * sizeof(float4);
    if (dev_realloc(d.lvl, d.lvl_cap, need) != RT_OK)
      return set_err(RT_ERR_OOM, "RT_MODE_SHOOT: cannot allocate the per-level buffer (%zu MiB = depth %d x %zu pixels x 16 B); "
                     "lower RT_TILE_PIXELS or the depth", need >> 20, g_max_depth, lvl_stride);
  }
  const int lb = d.lb;
  const shade_fn_t shade = (shade_fn_t)d.shade;
  HitQueue q[2];
  for (int k = 0; k < 2; k++) {
    q[k].r0 = d.qbuf + (size_t)(3 * k) * d.tile_cap;
    q[k].r1 = d.qbuf + (size_t)(3 * k + 1) * d.tile_cap;
    q[k].r2 = d.qbuf + (size_t)(3 * k + 2) * d.tile_cap;
    q[k].cls = d.qcls + (size_t)k * RT_NCLS * d.tile_cap; q[k].cap = (int)d.tile_cap; q[k].ctl = nullptr;
  }
  CK(cudaMemsetAsync(d.counters, 0, CT_COUNT * sizeof(unsigned long long), d.stream));
  d.evkind.clear();
  d.sc.max_depth = g_max_depth;   // before the first launch: k_primary files its hit points by it
  int mr;
  CK(cudaEventRecord(d.ev0, d.stream));
  for (long long first = 0; first < px_total; first += (long long)tile) {
    TileGeom tg;
    tg.width = W; tg.height = H; tg.rank = rank; tg.nranks = nranks; tg.band_h = bh;
    tg.first_px = first;
    tg.n_px = (int)((px_total - first) < (long long)tile ? (px_total - first) : (long long)tile);
    CK(cudaMemsetAsync(d.qcount, 0, (RT_MAX_DEPTH_CAP + 1) * RT_CTL_STRIDE * sizeof(int), d.stream));
    if (shoot)                                                     // no hit = black, reflectivity 0: only the pixels of this tile
      for (int level = 0; level < g_max_depth; level++)
        CK(cudaMemsetAsync(d.lvl + (size_t)level * lvl_stride, 0, (size_t)tg.n_px * sizeof(float4), d.stream));
    HitQueue q0 = q[0]; q0.ctl = d.qcount;
    const int ppx = d.sc.bvh_node ? 1 : RT_PRIM_PX_BRUTE;
    int gp = (tg.n_px + 256 * ppx - 1) / (256 * ppx); if (gp > d.grid_primary) gp = d.grid_primary;
    if ((mr = mark(d, 'p')) != RT_OK) return mr;
    if (d.sc.bvh_node) k_primary<true><<<gp, 256, 0, d.stream>>>(d.sc, tg, d.acc, d.c0, q0, lb + 1, d.counters);
    else k_primary<false><<<gp, 256, 0, d.stream>>>(d.sc, tg, d.acc, d.c0, q0, lb + 1, d.counters);
    plan->launches++;
    for (int level = 0; level < g_max_depth; level++) {
      HitQueue qin = q[level & 1], qout = q[(level + 1) & 1];
      qin.ctl = d.qcount + level * RT_CTL_STRIDE; qout.ctl = d.qcount + (level + 1) * RT_CTL_STRIDE;
      if ((mr = mark(d, 's')) != RT_OK) return mr;
      shade<<<d.grid_shade, 128, 0, d.stream>>>(d.sc, level, qin, qout, d.acc, d.c0, d.counters,
                                                shoot ? d.lvl + (size_t)level * lvl_stride : nullptr);
      plan->launches++;
    }
    int gf = (tg.n_px / 4 + 256) / 256; if (gf > d.grid_fin) gf = d.grid_fin;
    if ((mr = mark(d, 'f')) != RT_OK) return mr;
    k_finalize<<<gf, 256, 0, d.stream>>>(tg, d.acc, d.c0, dst, shoot ? d.lvl : nullptr, lvl_stride, g_max_depth);
    plan->launches++;
    CK(cudaGetLastError());
  }
  if ((mr = mark(d, 'e')) != RT_OK) return mr;
  CK(cudaEventRecord(d.ev1, d.stream));
  trace("enqueue_rank: all launches enqueued");
  return RT_OK;
}

int collect_stats(int ndev, int W, int H, const RankPlan *plans, rt_stats *st, double ms_gather, double t0) {
  rt_stats s;
  memset(&s, 0, sizeof s);
  s.pixels = (uint64_t)W * (uint64_t)H;
  s.rays_primary = s.pixels;
  double ms_k = 0;
  for (int i = 0; i < ndev; i++) {
    DeviceCtx &d = g_dev[i];
    CK(cudaSetDevice(d.dev));
    unsigned long long c[CT_COUNT];
    CK(cudaMemcpy(c, d.counters, sizeof c, cudaMemcpyDeviceToHost));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    if (ms > ms_k) ms_k = ms;
    s.rays_reflection += c[CT_RAYS_REFLECTION]; s.rays_shadow += c[CT_RAYS_SHADOW];
    s.hit_points += c[CT_HIT_POINTS];
    s.sphere_tests_o += c[CT_SPH_O]; s.sphere_tests_d += c[CT_SPH_D];
    s.plane_tests_o += c[CT_PLN_O]; s.plane_tests_d += c[CT_PLN_D];
    s.quadric_tests_o += c[CT_QUAD_O]; s.quadric_tests_d += c[CT_QUAD_D];
    s.kernel_launches += plans[i].launches;
    s.scan_slots += c[CT_SLOTS];
    s.ball_tests_o += c[CT_BALL_O]; s.ball_tests_d += c[CT_BALL_D];
    // SURVEY.md section 8a flop convention
    const double all = 10.0 * (c[CT_SPH_O] + c[CT_BALL_O]) + 8.0 * (c[CT_SPH_D] + c[CT_BALL_D]) + 8.0 * c[CT_PLN_O] + 6.0 * c[CT_PLN_D] +
               21.0 * c[CT_QUAD_O] + 37.0 * c[CT_QUAD_D] + 3.0 * c[CT_CANDIDATES] +
               77.0 * c[CT_LIGHT_SAMPLES] + 12.0 * c[CT_RAYS_SHADOW] + 21.0 * c[CT_RAYS_REFLECTION] +
               20.0 * c[CT_HIT_POINTS];
    // k_primary's share: per pixel every sphere (direction part only), plane and quadric once
    const double px = (double)plans[i].rows * W;
    const double prim = px * (17.0 + 8.0 * d.sc.n_spheres + 14.0 * d.sc.n_planes + 8.0 * d.sc.n_quadrics) +   // (quadrics: their bounding-ball tests; the few that pass are left in k_shade's share)
                        20.0 * c[CT_PRIMARY_HITS];
    s.flops += all + px * 17.0 + 20.0 * c[CT_PRIMARY_HITS];
    s.flops_shade += std::max(0.0, all + px * 17.0 + 20.0 * c[CT_PRIMARY_HITS] - prim);   // (with the exact index on, k_primary's share is an over-estimate)
    if (i == 0) {
      for (size_t k = 0; k + 1 < d.evkind.size(); k++) {
        float t = 0;
        CK(cudaEventElapsedTime(&t, d.evpool[k], d.evpool[k + 1]));
        if (d.evkind[k] == 'p') s.ms_primary += t;
        else if (d.evkind[k] == 's') { s.ms_shade += t; s.launches_shade++; }
        else if (d.evkind[k] == 'f') s.ms_finalize += t;
      }
    }
  }
  s.ms_kernels = ms_k;
  s.ms_gather = ms_gather;
  s.ms_total = now_ms() - t0;
  if (st) *st = s;
  return RT_OK;
}

int check_render_args(int W, int H) {
  if (!g_ndev) return set_err(RT_ERR_NO_DEVICE, "render before rt_init (no GPU initialised; there is no CPU path)");
  if (!g_have_scene) return set_err(RT_ERR_NO_SCENE, "render before rt_scene_upload");
  if (W <= 0 || H <= 0) return set_err(RT_ERR_ARG, "bad image size %dx%d", W, H);
  if (W > (1 << 26) || H > (1 << 26)) return set_err(RT_ERR_ARG, "image too large: %dx%d", W, H);
  return RT_OK;
}

}  // namespace

extern "C" {

// Drain every device's stream (error paths: nothing may still be running when the caller sees the failure
// and re-enters the library with buffers that the next call may reallocate).
static void drain_all(void) {
  for (int i = 0; i < g_ndev; i++) {
    if (g_dev[i].dev < 0 || !g_dev[i].stream) continue;
    cudaSetDevice(g_dev[i].dev);
    cudaStreamSynchronize(g_dev[i].stream);
  }
  cudaGetLastError();
}

// Renders every rank's rows on its device (asynchronous); with to_gpu0 the bands are gathered to GPU 0 over
// NCCL and de-interleaved into d0.frame, otherwise they stay in the per-device rank buffers (rank 0's in
// d0.staging) for rt_render's direct band copies.
static int render_all(int W, int H, bool to_gpu0, RankPlan *plans, size_t *off, double *ms_gather) {
  const int n = g_ndev;
  DeviceCtx &d0 = g_dev[0];
  int r;
  *ms_gather = 0;
  if (n == 1) {
    CK(cudaSetDevice(d0.dev));
    size_t need = (size_t)W * H * 3 + 64;
    if ((r = dev_realloc(d0.frame, d0.frame_cap, need)) != RT_OK) return r;
    if ((r = enqueue_rank(d0, W, H, 0, 1, d0.frame, &plans[0])) != RT_OK) return r;
    return RT_OK;
  }
  off[0] = 0;
  for (int i = 0; i < n; i++) {
    size_t b = (size_t)rt_rank_num_rows(H, i, n) * W * 3;
    off[i + 1] = off[i] + ((b + 255) & ~(size_t)255);
  }
  CK(cudaSetDevice(d0.dev));
  if ((r = dev_realloc(d0.staging, d0.staging_cap, (to_gpu0 ? off[n] : off[1]) + 64)) != RT_OK) return r;
  if (to_gpu0 && (r = dev_realloc(d0.frame, d0.frame_cap, (size_t)W * H * 3 + 64)) != RT_OK) return r;
  for (int i = 0; i < n; i++) {
    DeviceCtx &d = g_dev[i];
    uint8_t *dst = d0.staging;   // rank 0 renders straight into its staging slot
    if (i > 0) {
      CK(cudaSetDevice(d.dev));
      if ((r = dev_realloc(d.rank_buf, d.rank_cap, off[i + 1] - off[i] + 64)) != RT_OK) return r;
      dst = d.rank_buf;
    }
    if ((r = enqueue_rank(d, W, H, i, n, dst, &plans[i])) != RT_OK) return r;
  }
  if (!to_gpu0) return RT_OK;
  // band gather: every rank sends its rows to GPU 0 over NVLink (NCCL point-to-point)
  ncclResult_t nr = g_nccl.GroupStart();
  for (int i = 1; i < n && nr == ncclSuccess; i++) {
    nr = g_nccl.Send(g_dev[i].rank_buf, plans[i].bytes, ncclUint8, 0, g_nccl.comms[i], g_dev[i].stream);
    if (nr == ncclSuccess)
      nr = g_nccl.Recv(d0.staging + off[i], plans[i].bytes, ncclUint8, i, g_nccl.comms[0], d0.stream);
  }
  ncclResult_t ne = g_nccl.GroupEnd();
  if (nr != ncclSuccess || ne != ncclSuccess)
    return set_err(RT_ERR_NCCL, "band gather failed: %s", g_nccl.GetErrorString(nr != ncclSuccess ? nr : ne));
  CK(cudaSetDevice(d0.dev));
  AssembleArgs a;
  memset(&a, 0, sizeof a);
  for (int i = 0; i < n; i++) a.rank_off[i] = (long long)off[i];
  a.width = W; a.height = H; a.nranks = n; a.band_h = band_rows();
  k_assemble<<<d0.sm_count * 8, 256, 0, d0.stream>>>(a, d0.staging, d0.frame);
  CK(cudaGetLastError());
  CK(cudaEventRecord(d0.ev2, d0.stream));
  plans[0].launches += 1;
  return RT_OK;
}

static int sync_all(void) {
  for (int i = 0; i < g_ndev; i++) { CK(cudaSetDevice(g_dev[i].dev)); CK(cudaStreamSynchronize(g_dev[i].stream)); }
  CK(cudaSetDevice(g_dev[0].dev));
  return RT_OK;
}

int rt_render_device(int W, int H, void **dev_rgb_out, rt_stats *stats) {
  int r = check_render_args(W, H);
  if (r != RT_OK) return r;
  const double t0 = now_ms();
  const int n = g_ndev;
  RankPlan plans[RT_MAX_GPUS];
  size_t off[RT_MAX_GPUS + 1];
  DeviceCtx &d0 = g_dev[0];
  double ms_gather = 0;
  r = render_all(W, H, true, plans, off, &ms_gather);
  if (r == RT_OK) r = sync_all();
  if (r != RT_OK) { drain_all(); return r; }
  if (n > 1) {
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, d0.ev1, d0.ev2));
    ms_gather = ms;
  }
  if (dev_rgb_out) *dev_rgb_out = d0.frame;
  d0.frame_w = W; d0.frame_h = H;
  return collect_stats(n, W, H, plans, stats, ms_gather, t0);
}

// Rows of `rank` (back to back in dev_rows, band order) -> their places in the row-major host frame.  A band of
// full-width rows is contiguous on both sides, so the whole rank is ONE strided copy (band -> every nranks-th band
// slot of the frame) plus one plain copy for a ragged last band.  Asynchronous on `st`.
static int copy_rank_to_host(int W, int H, int rank, int nranks, const uint8_t *dev_rows, uint8_t *host_frame, cudaStream_t st) {
  const size_t pitch = (size_t)W * 3, bh = (size_t)band_rows(), band_bytes = pitch * bh;
  const size_t n_bands_total = ((size_t)H + bh - 1) / bh;
  if ((size_t)rank >= n_bands_total) return RT_OK;
  size_t mine = (n_bands_total - 1 - (size_t)rank) / (size_t)nranks + 1;          // bands rank, rank + nranks, ...
  const size_t last_band = (size_t)rank + (mine - 1) * (size_t)nranks;
  const size_t last_rows = (last_band == n_bands_total - 1) ? (size_t)H - last_band * bh : bh;
  const size_t full = last_rows == bh ? mine : mine - 1;
  uint8_t *dst0 = host_frame + (size_t)rank * band_bytes;
  if (full > 0) {
    if (nranks == 1) CK(cudaMemcpyAsync(dst0, dev_rows, full * band_bytes, cudaMemcpyDeviceToHost, st));
    else CK(cudaMemcpy2DAsync(dst0, band_bytes * (size_t)nranks, dev_rows, band_bytes, band_bytes, full, cudaMemcpyDeviceToHost, st));
  }
  if (full < mine)
    CK(cudaMemcpyAsync(host_frame + last_band * band_bytes, dev_rows + full * band_bytes, last_rows * pitch, cudaMemcpyDeviceToHost, st));
  return RT_OK;
}

static bool host_ptr_is_pinned(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost;
}

int rt_rank_rows_to_host(int W, int H, int rank, int nranks, const void *dev_rows, uint8_t *host_frame) {
  if (!g_ndev) return set_err(RT_ERR_NO_DEVICE, "rt_rank_rows_to_host before rt_init");
  if (W <= 0 || H <= 0 || nranks < 1 || rank < 0 || rank >= nranks || !dev_rows || !host_frame)
    return set_err(RT_ERR_ARG, "bad rt_rank_rows_to_host argument");
  DeviceCtx &d0 = g_dev[0];
  CK(cudaSetDevice(d0.dev));
  int r = copy_rank_to_host(W, H, rank, nranks, (const uint8_t *)dev_rows, host_frame, d0.stream);
  if (r != RT_OK) return r;
  CK(cudaStreamSynchronize(d0.stream));
  return RT_OK;
}

int rt_render(int W, int H, uint8_t *rgb_out, rt_stats *stats) {
  if (!rgb_out) return set_err(RT_ERR_ARG, "rgb_out is NULL");
  int r = check_render_args(W, H);
  if (r != RT_OK) return r;
  const double t0 = now_ms();
  trace("rt_render: begin");
  const int n = g_ndev;
  DeviceCtx &d0 = g_dev[0];
  // More than one GPU and a pinned (or registered) host pixmap: every GPU copies its own bands straight to their
  // places in the caller's pixmap over its own PCIe link -- no gather, no serial full-frame copy from GPU 0.
  // (The frame is then NOT left on GPU 0: rt_p3_write(NULL, ...) needs rt_render_device.)  RT_HOST_GATHER=1
  // forces the gather path.
  if (n > 1 && host_ptr_is_pinned(rgb_out) && !getenv("RT_HOST_GATHER")) {
    RankPlan plans[RT_MAX_GPUS];
    size_t off[RT_MAX_GPUS + 1];
    double ms_gather = 0;
    r = render_all(W, H, false, plans, off, &ms_gather);
    for (int i = 0; i < n && r == RT_OK; i++) {
      DeviceCtx &d = g_dev[i];
      cudaSetDevice(d.dev);
      r = copy_rank_to_host(W, H, i, n, i == 0 ? d0.staging : d.rank_buf, rgb_out, d.stream);
    }
    if (r == RT_OK) r = sync_all();
    if (r != RT_OK) { drain_all(); return r; }
    d0.frame_w = d0.frame_h = 0;
    r = collect_stats(n, W, H, plans, stats, 0.0, t0);
    if (stats) stats->ms_total = now_ms() - t0;
    return r;
  }
  void *frame = nullptr;
  r = rt_render_device(W, H, &frame, stats);
  if (r != RT_OK) return r;
  trace("rt_render: kernels done");
  CK(cudaSetDevice(d0.dev));
  CK(cudaMemcpyAsync(rgb_out, frame, (size_t)W * H * 3, cudaMemcpyDeviceToHost, d0.stream));
  CK(cudaStreamSynchronize(d0.stream));
  trace("rt_render: frame copied to the host pixmap");
  if (stats) stats->ms_total = now_ms() - t0;
  return RT_OK;
}

int rt_render_rank(int W, int H, int rank, int nranks, void *dev_dst, void **dev_rows_out, size_t *nbytes,
                   rt_stats *stats) {
  int r = check_render_args(W, H);
  if (r != RT_OK) return r;
  if (nranks < 1 || rank < 0 || rank >= nranks) return set_err(RT_ERR_ARG, "bad rank %d of %d", rank, nranks);
  if ((uintptr_t)dev_dst & 3) return set_err(RT_ERR_ARG, "rt_render_rank: dev_dst must be 4-byte aligned (k_finalize stores 32-bit words)");
  const double t0 = now_ms();
  DeviceCtx &d0 = g_dev[0];
  CK(cudaSetDevice(d0.dev));
  uint8_t *dst = (uint8_t *)dev_dst;
  if (!dst) {
    size_t need = (size_t)rt_rank_num_rows(H, rank, nranks) * W * 3 + 64;
    if ((r = dev_realloc(d0.rank_buf, d0.rank_cap, need)) != RT_OK) return r;
    dst = d0.rank_buf;
  }
  RankPlan plan;
  if ((r = enqueue_rank(d0, W, H, rank, nranks, dst, &plan)) != RT_OK) return r;
  CK(cudaStreamSynchronize(d0.stream));
  if (dev_rows_out) *dev_rows_out = dst;
  if (nbytes) *nbytes = plan.bytes;
  r = collect_stats(1, W, H, &plan, stats, 0.0, t0);
  if (stats) { stats->pixels = (uint64_t)plan.rows * W; stats->rays_primary = stats->pixels; }
  return r;
}

int rt_assemble_device(int W, int H, int nranks, const void *dev_staging, const int64_t *rank_offsets,
                       void *dev_frame) {
  if (!g_ndev) return set_err(RT_ERR_NO_DEVICE, "rt_assemble_device before rt_init");
  if (W <= 0 || H <= 0 || nranks < 1 || nranks > RT_MAX_GPUS || !dev_staging || !rank_offsets || !dev_frame)
    return set_err(RT_ERR_ARG, "bad rt_assemble_device argument");
  if (((size_t)W * 3) % 16 == 0) {   // k_assemble then moves 16 bytes at a time
    bool ok = ((uintptr_t)dev_frame & 15) == 0;
    for (int i = 0; i < nranks; i++) ok = ok && (((uintptr_t)dev_staging + (uintptr_t)rank_offsets[i]) & 15) == 0;
    if (!ok) return set_err(RT_ERR_ARG, "rt_assemble_device: dev_frame and dev_staging + rank_offsets[r] must be 16-byte aligned when 3*W is a multiple of 16");
  }
  DeviceCtx &d0 = g_dev[0];
  CK(cudaSetDevice(d0.dev));
  AssembleArgs a;
  memset(&a, 0, sizeof a);
  for (int i = 0; i < nranks; i++) a.rank_off[i] = rank_offsets[i];
  a.width = W; a.height = H; a.nranks = nranks; a.band_h = band_rows();
  k_assemble<<<d0.sm_count * 8, 256, 0, d0.stream>>>(a, (const uint8_t *)dev_staging, (uint8_t *)dev_frame);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(d0.stream));
  return RT_OK;
}

int rt_init_device(int ordinal) {
  if (g_ndev) rt_shutdown();
  int avail = 0;
  cudaError_t e = cudaGetDeviceCount(&avail);
  if (e != cudaSuccess || avail <= 0)
    return set_err(RT_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU path",
                   e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  if (ordinal < 0 || ordinal >= avail) return set_err(RT_ERR_ARG, "device ordinal %d out of range (0..%d)", ordinal, avail - 1);
  const int r = init_one(g_dev[0], ordinal);
  if (r != RT_OK) return init_fail(r);
  g_ndev = 1;
  return RT_OK;
}

}  // extern "C"

// ==================================================================================
// P3 writer (rt_p3.cuh): ppm_WriteOutP3 of C/ppm.c:168-184 with the formatting on the GPU.
// ==================================================================================
namespace {

int p3_header(char *hdr, size_t cap, int W, int H, unsigned max_color) {
  // fprintf(outFile, "P3\n"); fprintf(outFile, "%d %d \n%u\n", width, height, maxColor);  (C/ppm.c:171-172)
  return snprintf(hdr, cap, "P3\n%d %d \n%u\n", W, H, max_color & 0xffu);   // maxColor is a uint8_t there
}

int p3_check(int W, int H) {
  if (!g_ndev) return set_err(RT_ERR_NO_DEVICE, "P3 writer before rt_init (no GPU initialised; there is no CPU path)");
  if (W <= 0 || H <= 0 || W > (1 << 26) || H > (1 << 26)) return set_err(RT_ERR_ARG, "bad image size %dx%d", W, H);
  return RT_OK;
}

// Enqueue header copy + count + scan + emit on d.stream; d.p3_state = [n_spans + 1 offsets (u64)][n_spans lengths (u32)],
// the end offset of the whole text lands in d.p3_state[*n_spans_out].
int p3_enqueue(DeviceCtx &d, const uint8_t *dev_rgb, int W, int H, unsigned max_color, char *dev_text, size_t *n_spans_out) {
  if (!d.p3_attr_set) {   // per device: a process-wide flag breaks after rt_shutdown + rt_init_device(other ordinal)
    CK(cudaFuncSetAttribute(k_p3_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, P3_SMEM_BYTES));
    d.p3_attr_set = true;
  }
  char hdr[64];
  const int hl = p3_header(hdr, sizeof hdr, W, H, max_color);
  const unsigned long long n_values = (unsigned long long)W * (unsigned long long)H * 3ull;
  const unsigned long long n_chunks = (n_values + P3_CHUNK - 1) / P3_CHUNK;
  const unsigned long long n_spans = (n_chunks + P3_SPAN_CHUNKS - 1) / P3_SPAN_CHUNKS;
  if (n_spans >= (1ull << 30)) return set_err(RT_ERR_ARG, "image too large for the P3 writer");
  const size_t need = (size_t)(n_spans + 1) * sizeof(unsigned long long) + (size_t)n_spans * sizeof(unsigned) + 64;
  size_t cap_bytes = d.p3_state_cap;
  int r = dev_realloc(d.p3_state, cap_bytes, need);
  d.p3_state_cap = cap_bytes;
  if (r != RT_OK) return r;
  unsigned long long *span_off = d.p3_state;
  unsigned *span_len = (unsigned *)(d.p3_state + n_spans + 1);
  CK(cudaMemsetAsync(span_len, 0, (size_t)n_spans * sizeof(unsigned), d.stream));
  CK(cudaMemcpyAsync(dev_text, hdr, (size_t)hl, cudaMemcpyHostToDevice, d.stream));   // pageable source: staged before return
  const unsigned long long gmax = 16ull * d.sm_count;
  k_p3_count<<<(unsigned)(n_chunks < gmax ? n_chunks : gmax), P3_THREADS, 0, d.stream>>>(dev_rgb, n_values, n_chunks, span_len);
  k_p3_scan<<<1, 1024, 0, d.stream>>>(span_len, span_off, (unsigned)n_spans, (unsigned long long)hl);
  k_p3_emit<<<(unsigned)n_spans, P3_THREADS, P3_SMEM_BYTES, d.stream>>>(dev_rgb, n_values, n_chunks, span_off, (unsigned)n_spans, dev_text);
  CK(cudaGetLastError());
  if (n_spans_out) *n_spans_out = (size_t)n_spans;
  return RT_OK;
}

}  // namespace

extern "C" {

size_t rt_p3_max_bytes(int W, int H) {
  if (W <= 0 || H <= 0) return 0;
  return (size_t)64 + (size_t)W * (size_t)H * 12u;   // header + 3 values x 4 bytes per pixel
}

int rt_p3_format_device(const void *dev_rgb, int W, int H, unsigned max_color, void *dev_text, size_t text_cap,
                        size_t *text_bytes, rt_p3_stats *stats) {
  int r = p3_check(W, H);
  if (r != RT_OK) return r;
  if (!dev_rgb || !dev_text || ((uintptr_t)dev_rgb & 15) || ((uintptr_t)dev_text & 15))
    return set_err(RT_ERR_ARG, "rt_p3_format_device: NULL or not 16-byte aligned device pointer");
  if (text_cap < rt_p3_max_bytes(W, H)) return set_err(RT_ERR_ARG, "rt_p3_format_device: text_cap below rt_p3_max_bytes");
  const double t0 = now_ms();
  DeviceCtx &d = g_dev[0];
  CK(cudaSetDevice(d.dev));
  CK(cudaEventRecord(d.ev0, d.stream));
  // time the kernel alone: the resets go first
  size_t n_spans = 0;
  if ((r = p3_enqueue(d, (const uint8_t *)dev_rgb, W, H, max_color, (char *)dev_text, &n_spans)) != RT_OK) return r;
  CK(cudaEventRecord(d.ev1, d.stream));
  CK(cudaStreamSynchronize(d.stream));
  unsigned long long total = 0;
  CK(cudaMemcpy(&total, d.p3_state + n_spans, sizeof total, cudaMemcpyDeviceToHost));
  if (text_bytes) *text_bytes = (size_t)total;
  if (stats) {
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    memset(stats, 0, sizeof *stats);
    stats->text_bytes = total; stats->ms_format = ms; stats->ms_total = now_ms() - t0;
  }
  return RT_OK;
}

}  // extern "C"

// ---- P3 formatter for callers that may not allocate (rt_p3_format_device_noalloc) ---------------------------------
// Same three kernels as p3_enqueue, but the span table is not a cudaMalloc'ed buffer: it is the array below, part of
// the library image on every GPU that loads it ((16384 + 1) start offsets of 8 bytes + 16384 lengths of 4 = 196 616
// bytes).  In host code `g_p3_na_table` is only a handle for the runtime, not a device address; the address a kernel
// can use comes from cudaGetSymbolAddress, asked once per device with that device current (p3_na_device).
#define P3_NA_MAX_SPANS ((unsigned)(RT_P3_NOALLOC_MAX_VALUES / P3_SPAN))
static __device__ __align__(256) unsigned long long g_p3_na_table[(P3_NA_MAX_SPANS + 1) + P3_NA_MAX_SPANS / 2];

namespace {

struct P3Header { char b[64]; };

// The file header travels as a kernel argument, so no host-to-device copy (and no staging buffer) is involved.
__global__ void k_p3_header(P3Header h, int n, char *__restrict__ text) {
  if ((int)threadIdx.x < n) text[threadIdx.x] = h.b[threadIdx.x];
}

#define P3_NA_MAX_DEVICES 64
struct P3NaDevice {
  bool ready = false;
  unsigned long long *table = nullptr;   // device address of g_p3_na_table on this device
  int sm_count = 0;
} g_p3_na[P3_NA_MAX_DEVICES];

// First call on a device (that device is current): architecture check, address of the table, shared-memory limit of
// k_p3_emit.  Nothing here allocates on behalf of the caller; loading the kernels is the CUDA runtime's own business.
int p3_na_device(int dev, P3NaDevice **out) {
  if (dev < 0 || dev >= P3_NA_MAX_DEVICES) return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: device ordinal %d not supported", dev);
  P3NaDevice &s = g_p3_na[dev];
  if (!s.ready) {
    int major = 0, sms = 0;
    CK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (major != 10)
      return set_err(RT_ERR_NO_DEVICE, "rt_p3_format_device_noalloc: device %d is sm_%dx; the kernels are sm_100a only", dev, major);
    void *p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_p3_na_table));
    if (!p || ((uintptr_t)p & 255)) return set_err(RT_ERR_CUDA, "rt_p3_format_device_noalloc: bad address %p for the built-in span table", p);
    CK(cudaFuncSetAttribute(k_p3_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, P3_SMEM_BYTES));
    s.table = (unsigned long long *)p;
    s.sm_count = sms;
    s.ready = true;
  }
  *out = &s;
  return RT_OK;
}

// dev_rgb / dev_text must be device (or managed) memory reachable from the current device: a host pointer, or memory
// of another GPU, would otherwise only show up as a fault inside the kernels.
int p3_na_check_pointer(const void *p, int dev, const char *name) {
  cudaPointerAttributes pa;
  const cudaError_t e = cudaPointerGetAttributes(&pa, p);
  if (e != cudaSuccess) { cudaGetLastError(); return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: %s is not a CUDA pointer (%s)", name, cudaGetErrorString(e)); }
  if (pa.type == cudaMemoryTypeManaged) return RT_OK;
  if (pa.type != cudaMemoryTypeDevice) return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: %s is not device memory", name);
  if (pa.device != dev) return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: %s lives on device %d, the current device is %d", name, pa.device, dev);
  return RT_OK;
}

}  // namespace

extern "C" {

int rt_p3_format_device_noalloc(const void *dev_rgb, int W, int H, unsigned max_color, void *dev_text, size_t text_cap,
                                size_t *text_bytes, void *stream_ptr) {
  if (W <= 0 || H <= 0 || W > (1 << 26) || H > (1 << 26)) return set_err(RT_ERR_ARG, "bad image size %dx%d", W, H);
  const unsigned long long n_values = (unsigned long long)W * (unsigned long long)H * 3ull;
  if (n_values > RT_P3_NOALLOC_MAX_VALUES)
    return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: %dx%d has %llu values, the built-in span table ends at "
                   "RT_P3_NOALLOC_MAX_VALUES = %llu; use rt_p3_format_device for this frame", W, H, n_values,
                   (unsigned long long)RT_P3_NOALLOC_MAX_VALUES);
  if (!dev_rgb || !dev_text || ((uintptr_t)dev_rgb & 15) || ((uintptr_t)dev_text & 15))
    return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: NULL or not 16-byte aligned device pointer");
  if (text_cap < rt_p3_max_bytes(W, H)) return set_err(RT_ERR_ARG, "rt_p3_format_device_noalloc: text_cap below rt_p3_max_bytes");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { cudaGetLastError(); return set_err(RT_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e)); }
  P3NaDevice *s = nullptr;
  int r = p3_na_device(dev, &s);
  if (r != RT_OK) return r;
  if ((r = p3_na_check_pointer(dev_rgb, dev, "dev_rgb")) != RT_OK) return r;
  if ((r = p3_na_check_pointer(dev_text, dev, "dev_text")) != RT_OK) return r;
  cudaStream_t stream = (cudaStream_t)stream_ptr;
  const unsigned long long n_chunks = (n_values + P3_CHUNK - 1) / P3_CHUNK;
  const unsigned n_spans = (unsigned)((n_chunks + P3_SPAN_CHUNKS - 1) / P3_SPAN_CHUNKS);   // <= P3_NA_MAX_SPANS by the check above
  unsigned long long *span_off = s->table;                                  // n_spans + 1 entries
  unsigned *span_len = (unsigned *)(s->table + P3_NA_MAX_SPANS + 1);        // fixed place, whatever n_spans is
  // the table still holds the previous call's lengths: k_p3_count adds into them, so they start at zero
  CK(cudaMemsetAsync(span_len, 0, (size_t)n_spans * sizeof(unsigned), stream));
  P3Header hdr;
  memset(&hdr, 0, sizeof hdr);
  const int hl = p3_header(hdr.b, sizeof hdr.b, W, H, max_color);
  k_p3_header<<<1, 64, 0, stream>>>(hdr, hl, (char *)dev_text);
  const unsigned long long gmax = 16ull * (unsigned long long)s->sm_count;
  k_p3_count<<<(unsigned)(n_chunks < gmax ? n_chunks : gmax), P3_THREADS, 0, stream>>>((const uint8_t *)dev_rgb, n_values, n_chunks, span_len);
  k_p3_scan<<<1, 1024, 0, stream>>>(span_len, span_off, n_spans, (unsigned long long)hl);
  k_p3_emit<<<n_spans, P3_THREADS, P3_SMEM_BYTES, stream>>>((const uint8_t *)dev_rgb, n_values, n_chunks, span_off, n_spans, (char *)dev_text);
  CK(cudaGetLastError());
  unsigned long long total = 0;
  CK(cudaMemcpyAsync(&total, span_off + n_spans, sizeof total, cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  if (text_bytes) *text_bytes = (size_t)total;
  return RT_OK;
}

int rt_p3_write(const uint8_t *host_rgb, int W, int H, unsigned max_color, rt_p3_sink sink, void *user, rt_p3_stats *stats) {
  int r = p3_check(W, H);
  if (r != RT_OK) return r;
  if (!sink) return set_err(RT_ERR_ARG, "rt_p3_write: sink is NULL");
  const double t0 = now_ms();
  DeviceCtx &d = g_dev[0];
  CK(cudaSetDevice(d.dev));
  const size_t n_values = (size_t)W * (size_t)H * 3u;
  double ms_h2d = 0;
  if (host_rgb) {
    if ((r = dev_realloc(d.frame, d.frame_cap, n_values + 64)) != RT_OK) return r;
    const double t1 = now_ms();
    CK(cudaMemcpyAsync(d.frame, host_rgb, n_values, cudaMemcpyHostToDevice, d.stream));
    CK(cudaStreamSynchronize(d.stream));
    ms_h2d = now_ms() - t1;
    d.frame_w = W; d.frame_h = H;
  } else if (!d.frame || d.frame_w != W || d.frame_h != H) {
    return set_err(RT_ERR_ARG, "rt_p3_write: no resident %dx%d frame (last render was %dx%d)", W, H, d.frame_w, d.frame_h);
  }
  trace("rt_p3_write: begin");
  if ((r = dev_realloc(d.p3_text, d.p3_text_cap, rt_p3_max_bytes(W, H) + 64)) != RT_OK) return r;
  trace("rt_p3_write: text buffer allocated");
  // Pinned staging slabs sized to the job: pinning costs ~0.4 ms per MiB (the round-1 writer pinned 64 MiB up front,
  // 25 ms of a 30 ms one-shot 1000x1000 run), a big frame wants large slabs to keep the PCIe link busy.  Three of
  // them, each 1/6 of the worst-case text, between 1 and 64 MiB.
  size_t slab = rt_p3_max_bytes(W, H) / 6;
  slab = (slab + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
  if (slab < ((size_t)1 << 20)) slab = (size_t)1 << 20;
  if (slab > ((size_t)64 << 20)) slab = (size_t)64 << 20;
  if (d.p3_pinned_cap < slab) {
    for (int k = 0; k < 3; k++) {
      if (d.p3_pinned[k]) { CK(cudaFreeHost(d.p3_pinned[k])); d.p3_pinned[k] = nullptr; }
      CK(cudaHostAlloc((void **)&d.p3_pinned[k], slab, cudaHostAllocDefault));
      if (!d.p3_ev[k]) CK(cudaEventCreateWithFlags(&d.p3_ev[k], cudaEventDisableTiming));
    }
    d.p3_pinned_cap = slab;
    trace("rt_p3_write: pinned slabs allocated");
  }
  slab = d.p3_pinned_cap;
  CK(cudaEventRecord(d.ev0, d.stream));
  size_t n_spans = 0;
  if ((r = p3_enqueue(d, d.frame, W, H, max_color, d.p3_text, &n_spans)) != RT_OK) return r;
  CK(cudaEventRecord(d.ev1, d.stream));
  CK(cudaStreamSynchronize(d.stream));
  trace("rt_p3_write: text formatted on the GPU");
  unsigned long long total = 0;
  CK(cudaMemcpy(&total, d.p3_state + n_spans, sizeof total, cudaMemcpyDeviceToHost));
  // text -> host through the pinned slabs: copies k+1 and k+2 are in flight while the sink consumes slab k
  const size_t n_slabs = ((size_t)total + slab - 1) / slab;
  int sink_rc = 0;
  auto enqueue_copy = [&](size_t k) -> int {
    const size_t off = k * slab, nb = (size_t)total - off < slab ? (size_t)total - off : slab;
    CK(cudaMemcpyAsync(d.p3_pinned[k % 3], d.p3_text + off, nb, cudaMemcpyDeviceToHost, d.stream));
    CK(cudaEventRecord(d.p3_ev[k % 3], d.stream));
    return RT_OK;
  };
  for (size_t k = 0; k < 2 && k < n_slabs; k++) if ((r = enqueue_copy(k)) != RT_OK) return r;
  for (size_t k = 0; k < n_slabs && !sink_rc; k++) {
    if (k + 2 < n_slabs && (r = enqueue_copy(k + 2)) != RT_OK) return r;   // slab (k+2)%3 was consumed at step k-1
    const size_t off = k * slab, nb = (size_t)total - off < slab ? (size_t)total - off : slab;
    CK(cudaEventSynchronize(d.p3_ev[k % 3]));
    sink_rc = sink(d.p3_pinned[k % 3], nb, user);
  }
  CK(cudaStreamSynchronize(d.stream));
  trace("rt_p3_write: text delivered to the sink");
  if (sink_rc) return set_err(RT_ERR_ARG, "rt_p3_write: the sink reported failure (%d)", sink_rc);
  if (stats) {
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    stats->text_bytes = total; stats->ms_h2d = ms_h2d; stats->ms_format = ms; stats->ms_total = now_ms() - t0;
  }
  return RT_OK;
}

}  // extern "C"
