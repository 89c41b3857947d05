// C ABI of the PowerSGD hot path (see include/psgd_b200.h): host-side plan (tile tables,
// packed offsets), per-step pointer ring and the kernel launches.
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/psgd_b200.h"
#include "psgd_kernels.cuh"
#include "psgd_stream.cuh"
#include "psgd_tc.cuh"

using namespace psgd;

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};

int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}

#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                       __FILE__, __LINE__);                                   \
  } while (0)

long long env_ll(const char* name, long long dflt) {
  const char* s = getenv(name);
  if (!s || !*s) return dflt;
  return atoll(s);
}

int pow2_ceil(long long x) {
  int p = 1;
  while (p < x) p <<= 1;
  return p;
}

constexpr int kSlots = 4;
constexpr int kCtrRing = 8;

struct Timer {
  cudaEvent_t a, b;
  int phase;
};

}  // namespace

struct psgd_plan {
  int device = -1;
  int n_mat = 0, n_vec = 0, rank = 0, R = 1;  // R: compile-time rank the kernels run with
  std::vector<MatDesc> mats;
  std::vector<int4> tiles1, tiles4;
  std::vector<ColTile> tiles3;
  std::vector<long long> v_pre;
  std::vector<VecTile> vtiles;
  std::vector<QTile> qtiles;
  long long p_total = 0, q_total = 0, v_total = 0, ws_total = 0;
  int wc_cap = 0;
  int n_max = 0;
  std::vector<int> orth_small, orth_tall, orth_generic;  // K2 routing by row count
  std::vector<int> orth_tall_s;                          // CTAs cooperating on orth_tall[i] (2 .. orth_tall_cluster)
  // K2 job tables {matrix, first cluster rank of the group, CTAs in the group, rank in the group} per CTA: the tall
  // matrices' groups (largest first, so that a group never straddles a hardware cluster) and, in the merged table, the
  // small matrices behind them
  std::vector<int4> orth_jobs_tall, orth_jobs_merged;
  int4 *d_orth_jobs_tall = nullptr, *d_orth_jobs_merged = nullptr;
  int orth_tall_cluster = kOrthCluster;
  int split_max_sub = 0;        // K1: column-split units with at most this many row groups (PSGD_SPLIT_MAX_SUB, 0..3)
  bool pdl = true;              // programmatic dependent launch between the kernels of a step (PSGD_PDL)
  bool orth_merge = true;       // small matrices ride on the tall launch (PSGD_ORTH_MERGE)
  int orth_merge_max_ctas = 96;
  bool orth16_attr = false, orth16_attr_peer = false;  // 16-CTA clusters allowed on the K2 kernel of this plan's R
  int *d_orth_small = nullptr, *d_orth_tall = nullptr, *d_orth_generic = nullptr;
  size_t smem1 = 0, smem3 = 0, smem4 = 0;
  // streamed (bulk-copy ring) path for K1 / K4
  bool stream_ok = false;              // static: smem configuration valid and something qualifies
  bool stream1_now = false, stream4_now = false, stream1f_now = false;  // per step: every qualifying matrix is 16-byte aligned
  std::vector<char> eligible;
  bool has_chunked = false;  // some matrix streams in two column halves (its Q does not fit the stage in one piece)
  std::vector<StreamTask> tasks1, tasks4;
  std::vector<int4> tiles1_fb, tiles4_fb;  // direct-load tiles of the matrices that do not qualify
  std::vector<int> direct_ids;             // ... except the tiny ones, done as a side job of the streamed launch
  int* d_direct_ids = nullptr;
  std::vector<int> direct_row_pre;
  int* d_direct_row_pre = nullptr;
  StreamTask* d_tasks1 = nullptr;
  StreamTask* d_tasks4 = nullptr;
  // Task / done counter pairs of the streamed kernels.  A launch may start (programmatic dependent launch) while an
  // earlier launch of the same kernel is still draining, and its producers take their first task before they wait
  // for their predecessors, so consecutive launches must not share a pair: each launch takes the next pair of a
  // ring of kCtrRing, K1 from [0, 2 kCtrRing), K4 from [2 kCtrRing, 4 kCtrRing).  The last CTA of a launch re-arms
  // the pair it used.
  int* d_counters = nullptr;
  unsigned seq1 = 0, seq4 = 0;
  int n_ctr3 = 0;
  int* d_ctr3 = nullptr;  // K3: completion counter per (matrix, column tile), zero between launches
  unsigned long long* d_dbg = nullptr;  // PSGD_STREAM_DEBUG: per-CTA cycle breakdown of the last streamed K1
  int4* d_tiles1_fb = nullptr;
  int4* d_tiles4_fb = nullptr;
  int sm_count = 148;
  int stream_grid = 148, ring_slot_bytes = 0, ring_slots = 0, q_bytes = 0;
  int ring_slot_bytes1 = 0, ring_slots1 = 0;  // K1's own ring geometry (same shared-memory budget)
  size_t stream_smem = 0;
  // tensor-pipe (3 x TF32) kernels for R >= 16: their own, larger row tiles over the matrices with m % 4 == 0; the
  // rest stays with the FFMA direct kernels
  bool tc = false;
  int tc_mask = 0;  // PSGD_TC bits: 1 = K1, 2 = K3, 4 = K4
  bool tc_now = false, tcf_now = false;  // per step: every such matrix (and its outputs / fused operands) is 16-byte aligned
  int tc_wc1 = 0, tc_wc4 = 0;            // column chunk widths of the staged Q
  size_t tc_smem1 = 0, tc_smem4 = 0;
  std::vector<int4> tiles1_tc, tiles4_tc, tiles1_ntc, tiles4_ntc;
  int4 *d_tiles1_tc = nullptr, *d_tiles4_tc = nullptr, *d_tiles1_ntc = nullptr, *d_tiles4_ntc = nullptr;
  // device tables
  MatDesc* d_mats = nullptr;
  int4* d_tiles1 = nullptr;
  int4* d_tiles4 = nullptr;
  ColTile* d_tiles3 = nullptr;
  VecTile* d_vtiles = nullptr;
  QTile* d_qtiles = nullptr;
  // pointer ring
  size_t slot_bytes = 0;
  char* h_ring = nullptr;
  char* d_ring = nullptr;
  cudaEvent_t slot_ev[kSlots] = {};
  bool slot_used[kSlots] = {};
  int cur = -1;
  PtrSlot slot{};
  bool have_ge = false, have_vge = false;
  // timing
  unsigned timing = 0;  // bit k set: phase k is bracketed by events
  std::vector<Timer> timers;
  std::vector<Timer> free_timers;
  double ms_sum[PSGD_N_PHASES] = {};
  long long ms_cnt[PSGD_N_PHASES] = {};
};

struct psgd_tb {
  int device = -1;
  int n = 0, elem_size = 4;
  long long total = 0;
  std::vector<long long> pre;
  std::vector<VecTile> tiles;
  VecTile* d_tiles = nullptr;
  size_t slot_bytes = 0;
  char* h_ring = nullptr;
  char* d_ring = nullptr;
  cudaEvent_t slot_ev[kSlots] = {};
  bool slot_used[kSlots] = {};
  int cur = 0;
};

struct psgd_peer {
  int device = -1, world = 1, rank = 0, mode = kPeerPull;
  PeerLoad pf{}, q{};          // templates: P || flat buffer, Q buffer
  unsigned epoch_pf = 0, epoch_q = 0;
  unsigned* d_gsync = nullptr;  // grid-wide arrival counter of the Q prologue (monotonic)
  unsigned gseq = 0;
  // push mode (kPeerSlots): receive buffers with one slot per source rank, as ordinary and as multicast addresses
  float *rx_pf_local = nullptr, *rx_pf_mc = nullptr, *rx_q_local = nullptr, *rx_q_mc = nullptr;
  long long stride_pf = 0, stride_q = 0;
};

namespace {

// can this device hold a 16-CTA cluster of the register-resident Gram-Schmidt kernel (one CTA per SM)?
bool orth_cluster16_ok(int R, int device) {
  const void* fn = nullptr;
  switch (R) {
    case 1: fn = (const void*)k_orth_tall<1, 4>; break;
    case 2: fn = (const void*)k_orth_tall<2, 4>; break;
    case 4: fn = (const void*)k_orth_tall<4, 4>; break;
    case 8: fn = (const void*)k_orth_tall<8, 4>; break;
    default: return false;
  }
  int cur = -1;
  cudaGetDevice(&cur);
  struct Restore {
    int d;
    ~Restore() { if (d >= 0) cudaSetDevice(d); }
  } restore{cur};
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(kOrthClusterMax);
  cfg.blockDim = dim3(orth_threads(R));
  cudaLaunchAttribute at{};
  at.id = cudaLaunchAttributeClusterDimension;
  at.val.clusterDim.x = kOrthClusterMax;
  at.val.clusterDim.y = at.val.clusterDim.z = 1;
  cfg.attrs = &at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, fn, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return n >= 1;
}

DevPlan dev_plan(const psgd_plan* p) {
  DevPlan d;
  d.mats = p->d_mats;
  d.tiles1 = p->d_tiles1;
  d.tiles4 = p->d_tiles4;
  d.tiles3 = p->d_tiles3;
  d.vtiles = p->d_vtiles;
  d.qtiles = p->d_qtiles;
  d.n_mat = p->n_mat;
  d.n_vec = p->n_vec;
  d.wc_cap = p->wc_cap;
  d.pad_ = 0;
  return d;
}

// rows per tile for the row-streaming kernels (K1, K4)
void build_row_tiles(const psgd_plan* p, std::vector<int4>& out, long long target_bytes, int rows_per_warp,
                     int chunked_sets) {
  out.clear();
  const int gran = kWarps * rows_per_warp;
  for (int i = 0; i < p->n_mat; ++i) {
    const MatDesc& md = p->mats[i];
    if (md.n == 0 || md.m == 0) continue;
    long long t = target_bytes / ((long long)md.m * 4);
    t = (t / gran) * gran;
    if (t < gran) t = gran;
    if (md.m > p->wc_cap) t = std::max<long long>(t, chunked_sets * gran);  // a staged Q chunk serves this many row sets
    for (long long r0 = 0; r0 < md.n; r0 += t) {
      const int nr = (int)std::min<long long>(t, md.n - r0);
      out.push_back(make_int4(i, (int)r0, nr, 0));
    }
  }
}

// Tasks of the streamed kernels: runs of rows of one qualifying matrix, a whole number of ring units
// each, with guided-self-scheduling sizes (about remaining / (2 * CTAs) bytes, never below
// `min_units`).  The largest matrix goes last so that the fine-grained tail of the schedule stays inside
// one matrix (one Q stage per CTA).
void build_stream_tasks(const psgd_plan* p, int rows_per_sub, int slot_bytes, std::vector<StreamTask>& tasks) {
  tasks.clear();
  // smallest matrices first, largest last: the long head tasks cover the many small matrices (one Q stage
  // each), the fine-grained tail of the schedule walks through the few largest ones
  std::vector<int> order;
  long long total = 0;
  for (int i = 0; i < p->n_mat; ++i) {
    if (!p->eligible[i]) continue;
    total += (long long)p->mats[i].n * p->mats[i].m * 4;
    order.push_back(i);
  }
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    return (long long)p->mats[a].n * p->mats[a].m < (long long)p->mats[b].n * p->mats[b].m;
  });

  const long long div0 = 2LL * p->stream_grid * env_ll("PSGD_GSS_DIV", 1);
  // small sets (under ~1 MB per CTA, e.g. ResNet-18: 0.3 MB) get longer tasks: with only a handful of ring units per
  // CTA every extra task is an extra Q stage the consumers wait for (ResNet-18 r=2: 104 -> 95 us)
  // (measured, ResNet-18 r=2, 0.3 MB per CTA: floor 2 -> 104 us, 4 -> 95, 8 -> 90, 12 -> 101; ResNet-50, 0.7 MB per CTA:
  // 4 -> 118, 6 -> 126, 8 -> 123)
  const long long per_cta = total / std::max(1, p->stream_grid);
  const long long min_units_dflt = per_cta < (512 << 10) ? 8 : per_cta < (1 << 20) ? 4 : 2;
  const long long min_units0 = std::max<long long>(1, env_ll("PSGD_GSS_MIN_UNITS", min_units_dflt));
  // The end of the schedule lies inside the last (largest) matrix, where a finer cut costs no extra Q stage: a CTA's
  // consecutive tasks are then of the same matrix.  Elsewhere finer tasks mean more matrix switches per CTA.
  // (measured: DistilBERT r=4 step 236.8 -> 232.3 us, r=8 280.4 -> 274.5 us; tasks under ~48 KB cost ResNet-50, whose
  // last matrix has 18 KB rows, 3 us, hence the floor in bytes)
  const long long tail_div = std::max<long long>(1, env_ll("PSGD_GSS_TAIL_DIV", 3));
  const long long tail_min_bytes = std::max<long long>(1, env_ll("PSGD_GSS_TAIL_MIN_BYTES", 48 * 1024));
  long long remaining = total;
  int rot = 0;
  for (int i : order) {
    const bool last = (i == order.back());
    const long long div = last ? div0 * tail_div : div0;
    const MatDesc& md = p->mats[i];
    const long long row_bytes = (long long)md.m * 4;
    const int nchunk = md.pad_ > 1 ? md.pad_ : 1;
    // rows per ring unit of one pass; a column-chunked matrix streams every task once per column half and stages a
    // (large) Q half each time, so its tasks are kept longer
    const int rpu = (int)std::max<long long>(1, slot_bytes / (row_bytes / nchunk));
    const int sub_per_unit = (rpu + rows_per_sub - 1) / rows_per_sub;
    long long min_units = last ? std::min(min_units0, (tail_min_bytes + rpu * row_bytes - 1) / (rpu * row_bytes)) : min_units0;
    if (nchunk > 1) min_units = std::max<long long>(min_units, env_ll("PSGD_CHUNK_MIN_UNITS", 4));
    int row = 0;
    while (row < md.n) {
      const long long want = std::max<long long>(1, remaining / std::max<long long>(1, div));
      long long units = std::max<long long>(min_units, (want + rpu * row_bytes - 1) / (rpu * row_bytes));
      long long rows = std::min<long long>(units * rpu, md.n - row);
      StreamTask t;
      t.mat = i;
      t.row_begin = row;
      t.row_end = row + (int)rows;
      t.rot0 = rot;
      tasks.push_back(t);
      const long long nunits = (rows + rpu - 1) / rpu;
      rot = (int)((rot + nunits * sub_per_unit) % kConsumerWarps);
      row += (int)rows;
      remaining -= rows * row_bytes;
    }
  }
}

template <typename T>
int upload(T** dst, const std::vector<T>& v) {
  *dst = nullptr;
  if (v.empty()) return 0;
  CU(cudaMalloc((void**)dst, v.size() * sizeof(T)));
  CU(cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

struct PhaseScope {
  psgd_plan* p;
  cudaStream_t s;
  Timer t{};
  bool on = false;
  PhaseScope(psgd_plan* plan, int phase, cudaStream_t st) : p(plan), s(st) {
    if (!((p->timing >> phase) & 1u)) return;
    if (!p->free_timers.empty()) {
      t = p->free_timers.back();
      p->free_timers.pop_back();
    } else {
      if (cudaEventCreate(&t.a) != cudaSuccess || cudaEventCreate(&t.b) != cudaSuccess) return;
    }
    t.phase = phase;
    on = cudaEventRecord(t.a, s) == cudaSuccess;
  }
  ~PhaseScope() {
    if (!on) return;
    cudaEventRecord(t.b, s);
    p->timers.push_back(t);
  }
};

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail("launch of %s failed: %s", what, cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

int need_device(const psgd_plan* p) {
  if (!p) return fail("null plan");
  if (p->device < 0) return fail("plan was created host-only (device < 0): no compute entry point is available");
  return 0;
}
int need_ptrs(const psgd_plan* p) {
  if (p->cur < 0) return fail("psgd_set_ptrs has not been called on this plan");
  return 0;
}

// The opt-in dynamic shared-memory limit belongs to the kernel function (per device), not to a plan: several plans
// with different stage sizes share one instantiation.  The limit is therefore only ever raised, process-wide.
std::mutex g_attr_mu;
std::map<std::pair<const void*, int>, int> g_attr_max;

int raise_smem_limit(const void* f, size_t bytes) {
  if (bytes <= 48 * 1024) return 0;
  int device = 0;
  CU(cudaGetDevice(&device));
  std::lock_guard<std::mutex> lk(g_attr_mu);
  int& cur = g_attr_max[std::make_pair(f, device)];
  if ((int)bytes > cur) {
    CU(cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    cur = (int)bytes;
  }
  return 0;
}
template <typename F>
int set_smem(F* f, size_t bytes) {
  return raise_smem_limit((const void*)f, bytes);
}

// Launch with the optional attributes of this library: thread-block cluster size and programmatic stream
// serialisation (the kernel may start while its predecessor in the stream drains; see pdl_wait()).
template <typename... KArgs, typename... Args>
cudaError_t launch_ex(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, int cluster, bool pdl,
                      Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  unsigned na = 0;
  if (cluster > 1) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = (unsigned)cluster;
    at[na].val.clusterDim.y = 1;
    at[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace

extern "C" {

int psgd_abi_version(void) { return PSGD_ABI_VERSION; }
const char* psgd_last_error(void) { return g_err.c_str(); }
int64_t psgd_launch_count(void) { return g_launches.load(); }

int psgd_plan_create(int n_mat, const int64_t* n, const int64_t* m, int rank, int n_vec,
                     const int64_t* vec_len, int device, psgd_plan** out) {
  if (!out) return fail("out is null");
  *out = nullptr;
  if (n_mat < 0 || n_vec < 0) return fail("negative tensor count");
  if (rank < 1) return fail("rank must be >= 1 (got %d)", rank);
  if ((n_mat && (!n || !m)) || (n_vec && !vec_len)) return fail("null shape array");
  psgd_plan* p = new psgd_plan();
  p->device = device;
  p->n_mat = n_mat;
  p->n_vec = n_vec;
  p->rank = rank;

  int rmax = 1;
  for (int i = 0; i < n_mat; ++i) {
    if (n[i] < 0 || m[i] < 0 || n[i] > 0x7fffffffLL || m[i] > 0x7fffffffLL) {
      delete p;
      return fail("matrix %d has unsupported shape %lld x %lld", i, (long long)n[i], (long long)m[i]);
    }
    const int r = (int)std::min<long long>(std::min<long long>(n[i], m[i]), rank);  // reducer.py:78
    rmax = std::max(rmax, r);
  }
  if (rmax > PSGD_MAX_RANK) {
    delete p;
    return fail("effective rank %d exceeds PSGD_MAX_RANK=%d of this build", rmax, PSGD_MAX_RANK);
  }
  p->R = pow2_ceil(rmax);
  const int R = p->R;

  // shared-memory budget of the transposed Q stage -> column chunk width
  const long long q_stage_bytes = env_ll("PSGD_QSTAGE_BYTES", 96 * 1024);
  p->wc_cap = (int)std::max<long long>(128, ((q_stage_bytes / 4 / R) / 128) * 128);

  const long long t1_bytes = env_ll("PSGD_T1_BYTES", 128 * 1024);
  const long long t4_bytes = env_ll("PSGD_T4_BYTES", 96 * 1024);
  // K3 tiles: 256 KB of M each; half that for very small sets (under 0.5 MB per SM), whose few tiles per SM otherwise
  // leave the tail unbalanced (ResNet-18 r=2: 90 -> 85 us; ResNet-50 at 0.7 MB per SM loses 3 us with it)
  long long total_bytes = 0;
  for (int i = 0; i < n_mat; ++i) total_bytes += (long long)n[i] * m[i] * 4;
  const long long t3_bytes = env_ll("PSGD_T3_BYTES", (total_bytes / 148 < (512 << 10)) ? 128 * 1024 : 256 * 1024);
  const long long t3_maxblk = env_ll("PSGD_T3_MAXBLK", 48);
  const int cgw_cap = (int)std::min<long long>(env_ll("PSGD_T3_CGW", 64), kThreads3);

  p->mats.resize(n_mat);
  long long po = 0, qo = 0, wo = 0;
  int n_ctr3 = 0;
  int mmax = 0;
  size_t smem3 = 0;
  for (int i = 0; i < n_mat; ++i) {
    MatDesc& md = p->mats[i];
    md.n = (int)n[i];
    md.m = (int)m[i];
    md.r = (int)std::min<long long>(std::min<long long>(n[i], m[i]), rank);
    md.p_off = po;
    md.q_off = qo;
    md.ws_off = wo;
    md.pad_ = 0;
    po += (long long)md.n * md.r;  // reducer.py:97
    qo += (long long)md.m * md.r;  // reducer.py:98
    for (long long e0 = 0; e0 < (long long)md.m * md.r; e0 += kQTileElems) {
      QTile qt;
      qt.mat = i;
      qt.e0 = (int)e0;
      qt.count = (int)std::min<long long>(kQTileElems, (long long)md.m * md.r - e0);
      qt.pad_ = 0;
      if (md.n > 0) p->qtiles.push_back(qt);
    }
    mmax = std::max(mmax, md.m);
    p->n_max = std::max(p->n_max, md.n);

    // K3 geometry
    md.cpt = (md.m % 4 == 0 && md.m > 0) ? 4 : 1;
    const long long groups = std::max<long long>(1, md.m / md.cpt);
    int cgw = 1;
    while (cgw * 2 <= cgw_cap && groups % (cgw * 2) == 0) cgw *= 2;
    if (cgw < 16) cgw = std::min(cgw_cap, pow2_ceil(groups));  // ragged last tile instead of tiny tiles
    md.cgw = cgw;
    const long long wbytes = (long long)cgw * md.cpt * 4;
    long long t3 = std::max<long long>(32, t3_bytes / wbytes);
    t3 = std::max<long long>(t3, (md.n + t3_maxblk - 1) / t3_maxblk);
    t3 = std::min<long long>(t3, 2048);
    t3 = std::min<long long>(t3, std::max(1, md.n));
    md.t3_rows = (int)t3;
    md.nblk3 = md.n ? (int)((md.n + t3 - 1) / t3) : 0;
    wo += (long long)md.nblk3 * md.m * md.r;
    wo = (wo + 3) & ~3LL;
    const int RL = kThreads3 / cgw;
    size_t s3 = ((size_t)md.t3_rows * R + (size_t)RL * md.cpt * R * (cgw + 1)) * 4;
    if (R >= 16) s3 = std::max(s3, (size_t)((md.t3_rows + 7) & ~7) * (R == 16 ? 24 : 40) * 4);  // tensor-pipe tile: P rows, padded pitch
    smem3 = std::max(smem3, s3);
    const int ncol = (int)((groups + cgw - 1) / cgw);
    if (md.n && md.m) {
      for (int b = 0; b < md.nblk3; ++b)
        for (int ct = 0; ct < ncol; ++ct) {
          ColTile t{};
          t.mat = i;
          t.row0 = b * md.t3_rows;
          t.nrows = std::min(md.t3_rows, md.n - t.row0);
          t.col0 = ct * cgw * md.cpt;
          t.blk = b;
          t.pad0 = n_ctr3 + ct;  // completion counter of this (matrix, column tile)
          p->tiles3.push_back(t);
        }
      n_ctr3 += ncol;
    }
  }
  p->n_ctr3 = n_ctr3;
  p->p_total = po;
  p->q_total = qo;
  p->ws_total = wo;
  p->smem3 = smem3;

  const int wc = std::min(std::max(mmax, 1), p->wc_cap);
  p->smem1 = p->smem4 = (size_t)R * q_pitch(wc, R) * 4;

  const int rows1 = rows_s1(R);
  const int rows4 = rows_s4(R);
  // Matrices whose Q does not fit the stage in one piece: a staged chunk should serve several row sets (up to 4 per
  // warp), as long as that still leaves a tile per SM for their launch.
  auto chunked_sets_for = [&](int rows_per_warp) {
    const long long forced = env_ll("PSGD_CHUNKED_SETS", 0);
    if (forced > 0) return (int)forced;
    int sms = 148;
    if (device >= 0) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    for (int sets = 4; sets > 1; sets /= 2) {
      long long tiles = 0;
      const long long t = (long long)sets * kWarps * rows_per_warp;
      for (int i = 0; i < n_mat; ++i)
        if (p->mats[i].m > p->wc_cap && p->mats[i].n > 0) tiles += (p->mats[i].n + t - 1) / t;
      if (tiles >= sms) return sets;
    }
    return 1;
  };
  build_row_tiles(p, p->tiles1, t1_bytes, rows1, chunked_sets_for(rows1));
  build_row_tiles(p, p->tiles4, t4_bytes, rows4, chunked_sets_for(rows4));
  // off by default: measured slower than the FFMA kernels in all three phases (tensor-pipe evaluation);
  // PSGD_TC = 1 | 2 | 4 switches K1 | K3 | K4 to the tensor pipe
  if (R >= 16 && env_ll("PSGD_TC", 0) != 0) {
    p->tc = true;
    p->tc_mask = (int)env_ll("PSGD_TC", 0);
    p->tc_wc1 = R == 16 ? 768 : 384;
    p->tc_wc4 = R == 16 ? 768 : 384;
    const int wc1 = std::min(std::max((mmax + 15) & ~15, 16), p->tc_wc1);
    p->tc_smem1 = (size_t)2 * R * tc_q_pitch(wc1) * 4;
    p->tc_smem4 = (size_t)2 * p->tc_wc4 * (R + 1) * 4;
    const long long tc_bytes = env_ll("PSGD_TC_TILE_BYTES", 384 * 1024);
    std::vector<int4> all1, all4;
    const int save_cap = p->wc_cap;
    p->wc_cap = 1 << 30;  // no "chunked sets" floor for these tiles
    build_row_tiles(p, all1, tc_bytes, 2, 1);  // granularity: 16 rows per warp pass
    build_row_tiles(p, all4, tc_bytes, 2, 1);
    p->wc_cap = save_cap;
    for (const int4& t : all1) if (p->mats[t.x].m % 4 == 0) p->tiles1_tc.push_back(t);
    for (const int4& t : all4) if (p->mats[t.x].m % 4 == 0) p->tiles4_tc.push_back(t);
    for (const int4& t : p->tiles1) if (p->mats[t.x].m % 4 != 0) p->tiles1_ntc.push_back(t);
    for (const int4& t : p->tiles4) if (p->mats[t.x].m % 4 != 0) p->tiles4_ntc.push_back(t);
  }

  {  // streamed path: which matrices qualify, and the shared-memory split between Q stage and ring
    int smem_optin = 232448, sms = 148;
    if (device >= 0) {
      cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    }
    p->sm_count = sms;
    p->stream_grid = (int)env_ll("PSGD_STREAM_GRID", sms);
    const long long q_cap = env_ll("PSGD_STREAM_QCAP", 100 * 1024);
    long long slot = env_ll("PSGD_SLOT_BYTES", 48 * 1024);
    slot = std::max<long long>(1024, (slot / 128) * 128);
    long long slot1 = env_ll("PSGD_SLOT1_BYTES", slot);
    slot1 = std::max<long long>(1024, (slot1 / 128) * 128);
    p->eligible.assign(n_mat, 0);
    long long qb = 0;
    if (env_ll("PSGD_STREAM", 1) != 0) {
      const bool chunk_ok = env_ll("PSGD_STREAM_CHUNK", 1) != 0;
      for (int i = 0; i < n_mat; ++i) {
        MatDesc& md = p->mats[i];
        md.pad_ = 1;
        if (!(md.n > 0 && md.m > 0 && md.m % 4 == 0)) continue;
        const long long qs = (long long)R * q_pitch(md.m, R) * 4;
        if ((long long)md.m * 4 <= std::min(slot, slot1) && qs <= q_cap) {
          p->eligible[i] = 1;
          qb = std::max(qb, qs);
          continue;
        }
        // Q too large for the stage in one piece (wide matrix, high rank): stream the matrix in two column halves,
        // one Q half staged at a time (psgd_stream.cuh, column-chunked tasks)
        const int w = md.m / 2;
        const long long qs2 = (long long)R * q_pitch(w, R) * 4;
        if (chunk_ok && md.m % 8 == 0 && i < kMatCache && (long long)w * 4 <= std::min(slot, slot1) && qs2 <= q_cap) {
          p->eligible[i] = 1;
          md.pad_ = 2;
          p->has_chunked = true;
          qb = std::max(qb, qs2);
        }
      }
    }
    qb = (qb + 127) & ~127LL;
    const long long ring = (long long)smem_optin - qb - kStreamSmemExtra - 128;
    const int slots = (int)std::min<long long>(kMaxSlots, ring / (slot + kPartBytes));
    if (qb > 0 && slots >= 2) {
      p->stream_ok = true;
      p->ring_slot_bytes = (int)slot;
      p->ring_slots = slots;
      p->q_bytes = (int)qb;
      p->stream_smem = (size_t)qb + (size_t)slots * slot + kStreamSmemExtra + (size_t)slots * kPartBytes;
      p->ring_slot_bytes1 = (int)slot1;
      p->ring_slots1 = (int)std::min<long long>(kMaxSlots, ring / (slot1 + kPartBytes));
      if (p->ring_slots1 < 2) {
        p->ring_slot_bytes1 = (int)slot;
        p->ring_slots1 = slots;
      }
      p->stream_smem = std::max(p->stream_smem, (size_t)qb + (size_t)p->ring_slots1 * (p->ring_slot_bytes1 + kPartBytes) + kStreamSmemExtra);
      build_stream_tasks(p, rows1, p->ring_slot_bytes1, p->tasks1);
      build_stream_tasks(p, rows4, p->ring_slot_bytes, p->tasks4);
      // tiny matrices that do not qualify ride on the streamed launch as a side job; the rest get their own launch
      std::vector<char> side(n_mat, 0);
      const long long side_max = env_ll("PSGD_SIDE_MAX_ELEMS", 64 * 1024);
      for (int i = 0; i < n_mat; ++i) {
        const MatDesc& md = p->mats[i];
        if (!p->eligible[i] && md.n && md.m && (long long)md.n * md.m <= side_max) {
          side[i] = 1;
          if (p->direct_row_pre.empty()) p->direct_row_pre.push_back(0);
          p->direct_ids.push_back(i);
          p->direct_row_pre.push_back(p->direct_row_pre.back() + md.n);
        }
      }
      for (const int4& t : p->tiles1) if (!p->eligible[t.x] && !side[t.x]) p->tiles1_fb.push_back(t);
      for (const int4& t : p->tiles4) if (!p->eligible[t.x] && !side[t.x]) p->tiles4_fb.push_back(t);
    } else {
      p->eligible.assign(n_mat, 0);
    }
  }

  {  // K2 routing: register-resident single CTA, register-resident cluster, or generic cluster
    const int rpt = 4;  // rows per thread of the register-resident Gram-Schmidt kernels (1024 threads, 512 at R = 16)
    // A tall matrix is spread over a cluster of 16 CTAs: half the rows, and so half the warps, of 8 CTAs -- a step is
    // bound by the instructions a CTA issues (DistilBERT r=4 234 -> 232 us, r=8 276 -> 271; K2's embedding cluster
    // 10.4 -> 8.1 us).  8 where the device cannot place such a cluster (R = 16: 512 threads x 4 rows, 16 CTAs as before).
    p->orth_tall_cluster = kOrthCluster;
    if (R == 16 || (env_ll("PSGD_ORTH_TALL_CLUSTER", 16) >= 16 && (device < 0 || orth_cluster16_ok(R, device))))
      p->orth_tall_cluster = kOrthClusterMax;
    p->orth_merge = env_ll("PSGD_ORTH_MERGE", 1) != 0;
    p->pdl = env_ll("PSGD_PDL", 1) != 0;
    // measured: pays at R >= 8 (twice the FMAs per byte: DistilBERT r=8 step 291 -> 284 us), costs 1-3 % at R <= 4
    // where one warp per unit already keeps up with the ring
    p->split_max_sub = (int)std::min<long long>(3, std::max<long long>(0, env_ll("PSGD_SPLIT_MAX_SUB", R >= 8 ? 2 : 0)));
    p->orth_merge_max_ctas = (int)env_ll("PSGD_ORTH_MERGE_MAX", 96);
    for (int i = 0; i < n_mat; ++i) {
      const long long nn = p->mats[i].n;
      if (nn == 0 || p->mats[i].m == 0) continue;
      const long long cap = (long long)orth_threads(R) * rpt;  // rows one CTA holds in registers
      // CTAs per matrix: as many as it takes to hold the rows, and more while a CTA would keep over 1024 rows (8
      // warps): a step is paced by the CTA's slowest warp reaching the barrier before warp 0 can send (rank 4: 30
      // warps per CTA 1.5 us per step, 15 warps 1.15 us; rank 16: 12 warps 1.6 us, 6 warps 1.0 us)
      int grp = 1;
      while (((long long)grp * cap < nn || (long long)grp * 1024 < nn) && grp < p->orth_tall_cluster) grp *= 2;
      if (R <= 16 && nn <= cap) p->orth_small.push_back(i);
      else if (R <= 16 && nn <= (long long)grp * cap) { p->orth_tall.push_back(i); p->orth_tall_s.push_back(grp); }
      else p->orth_generic.push_back(i);
    }
    {  // job tables: groups in descending size (powers of two: every group starts at a multiple of its size)
      std::vector<int> order(p->orth_tall.size());
      for (size_t k = 0; k < order.size(); ++k) order[k] = (int)k;
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return p->orth_tall_s[a] > p->orth_tall_s[b]; });
      const int H = p->orth_tall_cluster;
      for (int k : order) {
        const int g = p->orth_tall_s[k], pos = (int)p->orth_jobs_tall.size();
        for (int c = 0; c < g; ++c) p->orth_jobs_tall.push_back(make_int4(p->orth_tall[k], pos % H, g, c));
      }
      p->orth_jobs_merged = p->orth_jobs_tall;
      for (int i : p->orth_small) {
        const int pos = (int)p->orth_jobs_merged.size();
        p->orth_jobs_merged.push_back(make_int4(i, pos % H, 1, 0));
      }
    }
  }

  p->v_pre.assign(n_vec + 1, 0);
  for (int j = 0; j < n_vec; ++j) {
    if (vec_len[j] < 0) {
      delete p;
      return fail("vector %d has negative length", j);
    }
    p->v_pre[j + 1] = p->v_pre[j] + vec_len[j];  // tensor_buffer.py:11-13
    for (long long st = 0; st < vec_len[j]; st += kVecTileElems) {
      VecTile vt;
      vt.flat_off = p->v_pre[j] + st;
      vt.tensor = j;
      vt.start = (int)st;
      vt.count = (int)std::min<long long>(kVecTileElems, vec_len[j] - st);
      vt.pad_ = 0;
      p->vtiles.push_back(vt);
    }
  }
  p->v_total = p->v_pre[n_vec];

  if (device >= 0) {
    int rc = 0;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
      delete p;
      return fail("cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    }
    rc |= upload(&p->d_mats, p->mats);
    rc |= upload(&p->d_tiles1, p->tiles1);
    rc |= upload(&p->d_tiles4, p->tiles4);
    rc |= upload(&p->d_tiles3, p->tiles3);
    rc |= upload(&p->d_vtiles, p->vtiles);
    rc |= upload(&p->d_qtiles, p->qtiles);
    rc |= upload(&p->d_tasks1, p->tasks1);
    rc |= upload(&p->d_tasks4, p->tasks4);
    if (!rc && cudaMalloc((void**)&p->d_counters, 4 * kCtrRing * sizeof(int)) != cudaSuccess) rc = fail("cudaMalloc(counters) failed");
    if (!rc && cudaMemset(p->d_counters, 0, 4 * kCtrRing * sizeof(int)) != cudaSuccess) rc = fail("cudaMemset(counters) failed");
    if (!rc && p->n_ctr3) {
      if (cudaMalloc((void**)&p->d_ctr3, (size_t)p->n_ctr3 * sizeof(int)) != cudaSuccess ||
          cudaMemset(p->d_ctr3, 0, (size_t)p->n_ctr3 * sizeof(int)) != cudaSuccess) rc = fail("cudaMalloc(ctr3) failed");
    }
    if (!rc && env_ll("PSGD_STREAM_DEBUG", 0)) {
      const size_t nb = (size_t)p->stream_grid * kDbgWords * sizeof(unsigned long long);
      if (cudaMalloc((void**)&p->d_dbg, nb) != cudaSuccess || cudaMemset(p->d_dbg, 0, nb) != cudaSuccess) rc = fail("cudaMalloc(dbg) failed");
    }
    rc |= upload(&p->d_direct_ids, p->direct_ids);
    rc |= upload(&p->d_direct_row_pre, p->direct_row_pre);
    rc |= upload(&p->d_tiles1_fb, p->tiles1_fb);
    rc |= upload(&p->d_tiles4_fb, p->tiles4_fb);
    rc |= upload(&p->d_tiles1_tc, p->tiles1_tc);
    rc |= upload(&p->d_tiles4_tc, p->tiles4_tc);
    rc |= upload(&p->d_tiles1_ntc, p->tiles1_ntc);
    rc |= upload(&p->d_tiles4_ntc, p->tiles4_ntc);
    rc |= upload(&p->d_orth_small, p->orth_small);
    rc |= upload(&p->d_orth_tall, p->orth_tall);
    rc |= upload(&p->d_orth_generic, p->orth_generic);
    rc |= upload(&p->d_orth_jobs_tall, p->orth_jobs_tall);
    rc |= upload(&p->d_orth_jobs_merged, p->orth_jobs_merged);

    // ring slot: 5 pointer arrays + flags per matrix, 2 pointer arrays per vector
    size_t sb = (size_t)n_mat * (5 * sizeof(void*) + sizeof(unsigned)) + (size_t)n_vec * 4 * sizeof(void*);
    sb = (sb + 255) & ~(size_t)255;
    if (sb == 0) sb = 256;
    p->slot_bytes = sb;
    if (!rc && cudaMallocHost((void**)&p->h_ring, sb * kSlots) != cudaSuccess) rc = fail("cudaMallocHost failed");
    if (!rc && cudaMalloc((void**)&p->d_ring, sb * kSlots) != cudaSuccess) rc = fail("cudaMalloc(ring) failed");
    for (int s = 0; s < kSlots && !rc; ++s)
      if (cudaEventCreateWithFlags(&p->slot_ev[s], cudaEventDisableTiming) != cudaSuccess) rc = fail("cudaEventCreate failed");
    if (rc) {
      std::string keep = g_err;
      psgd_plan_destroy(p);
      g_err = keep;
      return 1;
    }
  }
  *out = p;
  return 0;
}

int psgd_plan_destroy(psgd_plan* p) {
  if (!p) return 0;
  if (p->device >= 0) {
    cudaSetDevice(p->device);
    cudaFree(p->d_mats);
    cudaFree(p->d_tiles1);
    cudaFree(p->d_tiles4);
    cudaFree(p->d_tiles3);
    cudaFree(p->d_vtiles);
    cudaFree(p->d_qtiles);
    cudaFree(p->d_tasks1);
    cudaFree(p->d_tasks4);
    cudaFree(p->d_counters);
    cudaFree(p->d_dbg);
    cudaFree(p->d_ctr3);
    cudaFree(p->d_direct_ids);
    cudaFree(p->d_direct_row_pre);
    cudaFree(p->d_tiles1_fb);
    cudaFree(p->d_tiles4_fb);
    cudaFree(p->d_tiles1_tc);
    cudaFree(p->d_tiles4_tc);
    cudaFree(p->d_tiles1_ntc);
    cudaFree(p->d_tiles4_ntc);
    cudaFree(p->d_orth_small);
    cudaFree(p->d_orth_tall);
    cudaFree(p->d_orth_generic);
    cudaFree(p->d_orth_jobs_tall);
    cudaFree(p->d_orth_jobs_merged);
    cudaFree(p->d_ring);
    if (p->h_ring) cudaFreeHost(p->h_ring);
    for (int s = 0; s < kSlots; ++s)
      if (p->slot_ev[s]) cudaEventDestroy(p->slot_ev[s]);
    for (auto& t : p->timers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
    for (auto& t : p->free_timers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
  }
  delete p;
  return 0;
}

int64_t psgd_p_elems(const psgd_plan* p) { return p ? p->p_total : -1; }
int64_t psgd_q_elems(const psgd_plan* p) { return p ? p->q_total : -1; }
int64_t psgd_vec_elems(const psgd_plan* p) { return p ? p->v_total : -1; }
int64_t psgd_workspace_bytes(const psgd_plan* p) { return p ? std::max<long long>(16, p->ws_total * 4) : -1; }

int psgd_matrix_info(const psgd_plan* p, int i, int64_t* o) {
  if (!p || !o || i < 0 || i >= p->n_mat) return fail("psgd_matrix_info: bad argument");
  const MatDesc& md = p->mats[i];
  o[0] = md.n; o[1] = md.m; o[2] = md.r; o[3] = md.p_off; o[4] = md.q_off; o[5] = md.ws_off;
  o[6] = md.nblk3; o[7] = (int64_t)md.cgw * md.cpt;
  return 0;
}

int psgd_tile_count(const psgd_plan* p, int phase) {
  if (!p) return -1;
  switch (phase) {
    case PSGD_PHASE_P_FROM_M: return (int)p->tiles1.size();
    case PSGD_PHASE_FINALIZE: return (int)p->tiles4.size();
    case PSGD_PHASE_Q_FROM_MT: return (int)p->tiles3.size();
    case PSGD_PHASE_ORTH: return (int)(p->orth_jobs_merged.size() + p->orth_generic.size() * kOrthCluster);
    default: return 0;
  }
}

int psgd_tile_get(const psgd_plan* p, int phase, int idx, int32_t* o) {
  if (!p || !o || idx < 0 || idx >= psgd_tile_count(p, phase)) return fail("psgd_tile_get: bad argument");
  if (phase == PSGD_PHASE_P_FROM_M || phase == PSGD_PHASE_FINALIZE) {
    const int4 t = (phase == PSGD_PHASE_P_FROM_M ? p->tiles1 : p->tiles4)[idx];
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = 0; o[4] = p->mats[t.x].m; o[5] = 0;
  } else if (phase == PSGD_PHASE_Q_FROM_MT) {
    const ColTile& t = p->tiles3[idx];
    const MatDesc& md = p->mats[t.mat];
    o[0] = t.mat; o[1] = t.row0; o[2] = t.nrows; o[3] = t.col0;
    o[4] = std::min(md.cgw * md.cpt, md.m - t.col0); o[5] = t.blk;
  } else if (phase == PSGD_PHASE_ORTH) {
    // K2 job of CTA idx of the merged launch: {matrix, first cluster rank of its group, CTAs in the group, rank in
    // the group, CTAs per hardware cluster, 0}; behind them the matrices of the generic kernel (one 8-CTA cluster each)
    const int nj = (int)p->orth_jobs_merged.size();
    if (idx < nj) {
      const int4 j = p->orth_jobs_merged[idx];
      o[0] = j.x; o[1] = j.y; o[2] = j.z; o[3] = j.w; o[4] = p->orth_tall_cluster; o[5] = 0;
    } else {
      const int k = idx - nj;
      o[0] = p->orth_generic[k / kOrthCluster]; o[1] = 0; o[2] = kOrthCluster; o[3] = k % kOrthCluster;
      o[4] = kOrthCluster; o[5] = 1;
    }
  } else {
    return fail("psgd_tile_get: phase %d has no tile table", phase);
  }
  return 0;
}

static int set_ptrs_impl(psgd_plan* p, const void* const* mat_in, void* const* mat_out, void* const* mat_mem,
                         const void* const* vec_in, void* const* vec_out, const void* const* grad,
                         const void* const* err, const void* const* vec_grad, const void* const* vec_err,
                         psgd_stream stream) {
  if (need_device(p)) return 1;
  if (p->n_mat && (!mat_in || !mat_out || !mat_mem)) return fail("psgd_set_ptrs: null matrix pointer array");
  if (p->n_vec && (!vec_in || !vec_out)) return fail("psgd_set_ptrs: null vector pointer array");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const int slot = (p->cur + 1) % kSlots;
  if (p->slot_used[slot]) CU(cudaEventSynchronize(p->slot_ev[slot]));  // its upload has left the host buffer
  char* h = p->h_ring + (size_t)slot * p->slot_bytes;
  char* d = p->d_ring + (size_t)slot * p->slot_bytes;
  const size_t L = p->n_mat, L1 = p->n_vec;
  const void** h_in = (const void**)h;
  void** h_out = (void**)(h + L * 8);
  void** h_mem = (void**)(h + 2 * L * 8);
  const void** h_g = (const void**)(h + 3 * L * 8);
  const void** h_e = (const void**)(h + 4 * L * 8);
  const void** h_vin = (const void**)(h + 5 * L * 8);
  void** h_vout = (void**)(h + 5 * L * 8 + L1 * 8);
  const void** h_vg = (const void**)(h + 5 * L * 8 + 2 * L1 * 8);
  const void** h_ve = (const void**)(h + 5 * L * 8 + 3 * L1 * 8);
  unsigned* h_fl = (unsigned*)(h + 5 * L * 8 + 4 * L1 * 8);
  const size_t used_bytes = 5 * L * 8 + 4 * L1 * 8 + L * 4;
  p->have_ge = grad && err;
  p->have_vge = vec_grad && vec_err;
  bool s1 = p->stream_ok, s4 = p->stream_ok, s1f = p->stream_ok && grad && err;
  bool tc_all = p->tc, tcf_all = p->tc && grad && err;
  for (size_t i = 0; i < L; ++i) {
    const MatDesc& md = p->mats[i];
    const bool nonempty = md.n && md.m;
    if (nonempty && (!mat_in[i] || !mat_out[i] || !mat_mem[i])) return fail("psgd_set_ptrs: matrix %zu has a null pointer", i);
    if (((uintptr_t)mat_in[i] | (uintptr_t)mat_out[i] | (uintptr_t)mat_mem[i]) & 3)
      return fail("psgd_set_ptrs: matrix %zu is not 4-byte aligned", i);
    h_in[i] = mat_in[i];
    h_out[i] = mat_out[i];
    h_mem[i] = mat_mem[i];
    h_g[i] = p->have_ge ? grad[i] : nullptr;
    h_e[i] = p->have_ge ? err[i] : nullptr;
    unsigned fl = 0;
    const bool m4 = md.m % 4 == 0;
    if (m4 && ((uintptr_t)mat_in[i] & 15) == 0) fl |= kFlagInVec;
    if (m4 && (((uintptr_t)mat_out[i] | (uintptr_t)mat_mem[i]) & 15) == 0) fl |= kFlagOutVec;
    if (p->have_ge && m4 && (((uintptr_t)grad[i] | (uintptr_t)err[i]) & 15) == 0) fl |= kFlagGeVec;
    h_fl[i] = fl;
    if (p->tc && m4 && nonempty) {
      if (!(fl & kFlagInVec) || !(fl & kFlagOutVec)) tc_all = false;
      if (!(fl & kFlagInVec) || !(fl & kFlagOutVec) || !(fl & kFlagGeVec)) tcf_all = false;
    }
    if (p->stream_ok && p->eligible[i]) {
      if (!(fl & kFlagInVec)) s1 = false;
      if (!(fl & kFlagInVec) || !(fl & kFlagGeVec)) s1f = false;
      if (!(fl & kFlagInVec) || !(fl & kFlagOutVec)) s4 = false;
    }
  }
  p->tc_now = tc_all;
  p->tcf_now = tcf_all;
  p->stream1_now = s1;
  p->stream1f_now = s1f;
  p->stream4_now = s4;
  for (size_t j = 0; j < L1; ++j) {
    const bool nonempty = p->v_pre[j + 1] > p->v_pre[j];
    if (nonempty && (!vec_in[j] || !vec_out[j])) return fail("psgd_set_ptrs: vector %zu has a null pointer", j);
    h_vin[j] = vec_in[j];
    h_vout[j] = vec_out[j];
    h_vg[j] = p->have_vge ? vec_grad[j] : nullptr;
    h_ve[j] = p->have_vge ? vec_err[j] : nullptr;
    if (p->have_vge && nonempty && (!vec_grad[j] || !vec_err[j])) return fail("psgd_set_ptrs: vector %zu has a null grad/err pointer", j);
  }
  if (p->cur >= 0 && memcmp(h, p->h_ring + (size_t)p->cur * p->slot_bytes, used_bytes) == 0) {
    return 0;  // same addresses as the table already on the device: nothing to upload
  }
  CU(cudaMemcpyAsync(d, h, p->slot_bytes, cudaMemcpyHostToDevice, s));
  CU(cudaEventRecord(p->slot_ev[slot], s));
  p->slot_used[slot] = true;
  p->cur = slot;
  p->slot.in = (const float* const*)d;
  p->slot.out = (float* const*)(d + L * 8);
  p->slot.mem = (float* const*)(d + 2 * L * 8);
  p->slot.g = (const float* const*)(d + 3 * L * 8);
  p->slot.e = (const float* const*)(d + 4 * L * 8);
  p->slot.vin = (const float* const*)(d + 5 * L * 8);
  p->slot.vout = (float* const*)(d + 5 * L * 8 + L1 * 8);
  p->slot.vg = (const float* const*)(d + 5 * L * 8 + 2 * L1 * 8);
  p->slot.ve = (const float* const*)(d + 5 * L * 8 + 3 * L1 * 8);
  p->slot.flags = (const unsigned*)(d + 5 * L * 8 + 4 * L1 * 8);
  return 0;
}

int psgd_set_ptrs(psgd_plan* p, const void* const* mat_in, void* const* mat_out, void* const* mat_mem,
                  const void* const* vec_in, void* const* vec_out, const void* const* grad,
                  const void* const* err, psgd_stream stream) {
  return set_ptrs_impl(p, mat_in, mat_out, mat_mem, vec_in, vec_out, grad, err, nullptr, nullptr, stream);
}

int psgd_set_ptrs_fused(psgd_plan* p, const void* const* mat_send, void* const* mat_out, void* const* mat_mem,
                        const void* const* vec_send, void* const* vec_out, const void* const* mat_grad,
                        const void* const* mat_err, const void* const* vec_grad, const void* const* vec_err,
                        psgd_stream stream) {
  if (p && ((p->n_mat && (!mat_grad || !mat_err)) || (p->n_vec && (!vec_grad || !vec_err))))
    return fail("psgd_set_ptrs_fused: grad / err pointer arrays are required");
  return set_ptrs_impl(p, mat_send, mat_out, mat_mem, vec_send, vec_out, mat_grad, mat_err, vec_grad, vec_err, stream);
}

int psgd_pack_vec_fused(psgd_plan* p, float* flat, psgd_stream stream) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (!p->have_vge) return fail("psgd_pack_vec_fused: call psgd_set_ptrs_fused first");
  if (p->v_total == 0) return 0;
  if (!flat) return fail("psgd_pack_vec_fused: flat is null");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  PhaseScope ts(p, PSGD_PHASE_PACK, s);
  k_gather_sum<<<(unsigned)p->vtiles.size(), kThreads, 0, s>>>(p->d_vtiles, p->slot, flat);
  return check_launch("k_gather_sum");
}

int psgd_pack_vec(psgd_plan* p, float* flat, psgd_stream stream) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (p->v_total == 0) return 0;
  if (!flat) return fail("psgd_pack_vec: flat is null");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  PhaseScope ts(p, PSGD_PHASE_PACK, s);
  k_gather<uint32_t><<<(unsigned)p->vtiles.size(), kThreads, 0, s>>>(p->d_vtiles, (const void* const*)p->slot.vin,
                                                                      (uint32_t*)flat);
  return check_launch("k_gather");
}

#define DISPATCH_R(R_, ...)                    \
  switch (R_) {                                \
    case 1: { constexpr int RR = 1; __VA_ARGS__; } break;   \
    case 2: { constexpr int RR = 2; __VA_ARGS__; } break;   \
    case 4: { constexpr int RR = 4; __VA_ARGS__; } break;   \
    case 8: { constexpr int RR = 8; __VA_ARGS__; } break;   \
    case 16: { constexpr int RR = 16; __VA_ARGS__; } break; \
    case 32: { constexpr int RR = 32; __VA_ARGS__; } break; \
    default: return fail("unsupported compile-time rank %d", R_); \
  }

#define DISPATCH_R8(R_, ...)                   \
  switch (R_) {                                \
    case 1: { constexpr int RR = 1; __VA_ARGS__; } break;   \
    case 2: { constexpr int RR = 2; __VA_ARGS__; } break;   \
    case 4: { constexpr int RR = 4; __VA_ARGS__; } break;   \
    case 8: { constexpr int RR = 8; __VA_ARGS__; } break;   \
    default: return fail("unsupported compile-time rank %d", R_); \
  }
#define DISPATCH_R16(R_, ...)                  \
  switch (R_) {                                \
    case 1: { constexpr int RR = 1; __VA_ARGS__; } break;   \
    case 2: { constexpr int RR = 2; __VA_ARGS__; } break;   \
    case 4: { constexpr int RR = 4; __VA_ARGS__; } break;   \
    case 8: { constexpr int RR = 8; __VA_ARGS__; } break;   \
    case 16: { constexpr int RR = 16; __VA_ARGS__; } break; \
    default: return fail("unsupported compile-time rank %d", R_); \
  }

static int p_from_m_impl(psgd_plan* p, float* p_mem, const float* q_mem, int fuse_add, float q_div, float* flat,
                         psgd_stream stream, psgd_peer* peer = nullptr, bool out_mc = false);
static PeerQ peer_q_args(psgd_plan* p, psgd_peer* peer, float* q_mem);
static int peer_reduce_standalone(psgd_peer* peer, int which, float* out, long long n, float div, long long div_from,
                                  cudaStream_t s);
static int orth_impl(psgd_plan* p, float* p_mem, float n_workers, psgd_stream stream);

int psgd_p_from_m(psgd_plan* p, float* p_mem, const float* q_mem, int fuse_add, float q_div, psgd_stream stream) {
  return p_from_m_impl(p, p_mem, q_mem, fuse_add, q_div, nullptr, stream);
}

int psgd_p_from_m_orth(psgd_plan* p, float* p_mem, const float* q_mem, int fuse_add, float q_div, float* flat,
                       psgd_stream stream) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (fuse_add && flat && !p->have_vge) return fail("psgd_p_from_m_orth: fuse_add needs psgd_set_ptrs_fused");
  if (p_from_m_impl(p, p_mem, q_mem, fuse_add, q_div, flat, stream)) return 1;
  return orth_impl(p, p_mem, 1.0f, stream);
}

int psgd_p_from_m_pack(psgd_plan* p, float* p_mem, const float* q_mem, int fuse_add, float q_div, float* flat,
                       psgd_stream stream) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (p->v_total && !flat) return fail("psgd_p_from_m_pack: flat is null");
  if (fuse_add && !p->have_vge) return fail("psgd_p_from_m_pack: fuse_add needs psgd_set_ptrs_fused");
  return p_from_m_impl(p, p_mem, q_mem, fuse_add, q_div, flat, stream);
}

static int p_from_m_impl(psgd_plan* p, float* p_mem, const float* q_mem, int fuse_add, float q_div, float* flat,
                         psgd_stream stream, psgd_peer* peer, bool out_mc) {
  if (need_device(p) || need_ptrs(p)) return 1;
  const bool tc1 = p->tc && (p->tc_mask & 1) && (fuse_add ? p->tcf_now : p->tc_now);
  const bool fold = flat && p->v_total && !p->tiles1.empty() && !tc1 && (fuse_add ? p->stream1f_now : p->stream1_now);
  if (flat && p->v_total && !fold) {  // no streamed launch to ride on: separate gather
    if (fuse_add ? psgd_pack_vec_fused(p, flat, stream) : psgd_pack_vec(p, flat, stream)) return 1;
  }
  if (p->tiles1.empty()) return 0;
  if (!p_mem || !q_mem) return fail("psgd_p_from_m: null p_mem/q_mem");
  if (fuse_add && !p->have_ge) return fail("psgd_p_from_m: fuse_add needs grad/err pointers in psgd_set_ptrs");
  if (!(q_div > 0.f)) return fail("psgd_p_from_m: q_div must be positive");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const DevPlan dp = dev_plan(p);
  const unsigned grid = (unsigned)p->tiles1.size();
  PhaseScope ts(p, PSGD_PHASE_P_FROM_M, s);
  if (tc1) {  // ranks 16 / 32: tensor pipe (3 x TF32), see psgd_tc.cuh
    if (peer && peer_reduce_standalone(peer, 1, const_cast<float*>(q_mem), p->q_total, (float)peer->world, 0, s)) return 1;
    if (!p->tiles1_tc.empty()) {
      DevPlan tp = dp;
      tp.tiles1 = p->d_tiles1_tc;
      const unsigned g = (unsigned)p->tiles1_tc.size();
      if (p->R == 16) {
        if (fuse_add) {
          if (set_smem(k_p_from_m_tc<16, true>, p->tc_smem1)) return 1;
          k_p_from_m_tc<16, true><<<g, kTcThreads, p->tc_smem1, s>>>(tp, p->slot, q_mem, p_mem, q_div, p->tc_wc1);
        } else {
          if (set_smem(k_p_from_m_tc<16, false>, p->tc_smem1)) return 1;
          k_p_from_m_tc<16, false><<<g, kTcThreads, p->tc_smem1, s>>>(tp, p->slot, q_mem, p_mem, q_div, p->tc_wc1);
        }
      } else {
        if (fuse_add) {
          if (set_smem(k_p_from_m_tc<32, true>, p->tc_smem1)) return 1;
          k_p_from_m_tc<32, true><<<g, kTcThreads, p->tc_smem1, s>>>(tp, p->slot, q_mem, p_mem, q_div, p->tc_wc1);
        } else {
          if (set_smem(k_p_from_m_tc<32, false>, p->tc_smem1)) return 1;
          k_p_from_m_tc<32, false><<<g, kTcThreads, p->tc_smem1, s>>>(tp, p->slot, q_mem, p_mem, q_div, p->tc_wc1);
        }
      }
      if (check_launch("k_p_from_m_tc")) return 1;
    }
    if (!p->tiles1_ntc.empty()) {
      DevPlan fb = dp;
      fb.tiles1 = p->d_tiles1_ntc;
      DISPATCH_R(p->R, {
        if (fuse_add) {
          if (set_smem(k_p_from_m<RR, true>, p->smem1)) return 1;
          k_p_from_m<RR, true><<<(unsigned)p->tiles1_ntc.size(), kThreads, p->smem1, s>>>(fb, p->slot, q_mem, p_mem, q_div);
        } else {
          if (set_smem(k_p_from_m<RR, false>, p->smem1)) return 1;
          k_p_from_m<RR, false><<<(unsigned)p->tiles1_ntc.size(), kThreads, p->smem1, s>>>(fb, p->slot, q_mem, p_mem, q_div);
        }
      });
      if (check_launch("k_p_from_m")) return 1;
    }
    return 0;
  }
  if (fuse_add ? p->stream1f_now : p->stream1_now) {
    StreamPlan sp{p->d_tasks1, p->d_counters + 2 * (p->seq1++ % kCtrRing), (int)p->tasks1.size(), p->ring_slot_bytes1, p->ring_slots1, p->q_bytes, p->d_dbg,
                  p->d_vtiles, flat, fold ? (int)p->vtiles.size() : 0, fold ? (fuse_add ? 2 : 1) : 0, 1.0f,
                  p->d_direct_ids, p->d_direct_row_pre, (int)p->direct_ids.size(), p->split_max_sub, out_mc ? 1 : 0};
    // peer: Q is summed over the ranks first, in the kernel's prologue when its grid is co-resident (one CTA per SM)
    if (peer && p->stream_grid > p->sm_count) {
      if (peer_reduce_standalone(peer, 1, const_cast<float*>(q_mem), p->q_total, (float)peer->world, 0, s)) return 1;
      peer = nullptr;
    }
    const PeerQ pq = peer_q_args(p, peer, const_cast<float*>(q_mem));
    DISPATCH_R(p->R, {
      if (fuse_add) {
        if (set_smem(k_p_from_m_stream<RR, true>, p->stream_smem)) return 1;
        CU(launch_ex(k_p_from_m_stream<RR, true>, dim3(p->stream_grid), dim3(kStreamThreads), p->stream_smem, s, 1, p->pdl, dp, p->slot, sp, q_mem, p_mem, q_div, pq));
      } else {
        if (set_smem(k_p_from_m_stream<RR, false>, p->stream_smem)) return 1;
        CU(launch_ex(k_p_from_m_stream<RR, false>, dim3(p->stream_grid), dim3(kStreamThreads), p->stream_smem, s, 1, p->pdl, dp, p->slot, sp, q_mem, p_mem, q_div, pq));
      }
    });
    if (check_launch("k_p_from_m_stream")) return 1;
    if (!p->tiles1_fb.empty()) {
      DevPlan fb = dp;
      fb.tiles1 = p->d_tiles1_fb;
      DISPATCH_R(p->R, {
        if (fuse_add) {
          if (set_smem(k_p_from_m<RR, true>, p->smem1)) return 1;
          k_p_from_m<RR, true><<<(unsigned)p->tiles1_fb.size(), kThreads, p->smem1, s>>>(fb, p->slot, q_mem, p_mem, q_div);
        } else {
          if (set_smem(k_p_from_m<RR, false>, p->smem1)) return 1;
          k_p_from_m<RR, false><<<(unsigned)p->tiles1_fb.size(), kThreads, p->smem1, s>>>(fb, p->slot, q_mem, p_mem, q_div);
        }
      });
      if (check_launch("k_p_from_m")) return 1;
    }
    return 0;
  }
  if (peer && peer_reduce_standalone(peer, 1, const_cast<float*>(q_mem), p->q_total, (float)peer->world, 0, s)) return 1;
  DISPATCH_R(p->R, {
    if (fuse_add) {
      if (set_smem(k_p_from_m<RR, true>, p->smem1)) return 1;
      k_p_from_m<RR, true><<<grid, kThreads, p->smem1, s>>>(dp, p->slot, q_mem, p_mem, q_div);
    } else {
      if (set_smem(k_p_from_m<RR, false>, p->smem1)) return 1;
      k_p_from_m<RR, false><<<grid, kThreads, p->smem1, s>>>(dp, p->slot, q_mem, p_mem, q_div);
    }
  });
  return check_launch("k_p_from_m");
}

int psgd_orthogonalize(psgd_plan* p, float* p_mem, float n_workers, psgd_stream stream) {
  return orth_impl(p, p_mem, n_workers, stream);
}

static int orth_impl(psgd_plan* p, float* p_mem, float n_workers, psgd_stream stream) {
  if (need_device(p)) return 1;
  if (p->n_mat == 0) return 0;
  if (!p_mem) return fail("psgd_orthogonalize: null p_mem");
  if (!(n_workers >= 1.f)) return fail("psgd_orthogonalize: n_workers must be >= 1");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const DevPlan dp = dev_plan(p);
  if (p->orth_small.empty() && p->orth_tall.empty() && p->orth_generic.empty()) return 0;
  PhaseScope ts(p, PSGD_PHASE_ORTH, s);
  // The small matrices (one CTA each) share the launch of the tall ones (one cluster each) when the whole thing
  // fits a single wave of clusters; a big population of small matrices keeps its own plain launch, where no SM is
  // lost to cluster placement.
  const int tcs = p->orth_tall_cluster;
  const int n_small = (int)p->orth_small.size(), n_tall = (int)p->orth_tall.size();
  const int merged_ctas = ((int)p->orth_jobs_merged.size() + tcs - 1) / tcs * tcs;
  const bool merged = n_tall > 0 && n_small > 0 && p->orth_merge && merged_ctas <= p->orth_merge_max_ctas;
  // 1. matrices that fit one CTA's registers: one CTA each, block barriers only
  if (n_small && !merged) {
    const unsigned grid = (unsigned)n_small;
    if (p->R <= 8) {
      DISPATCH_R8(p->R, { CU(launch_ex(k_orth_small<RR, 4>, dim3(grid), dim3(orth_threads(p->R)), 0, s, 1, p->pdl, dp, (const int*)p->d_orth_small, p_mem, n_workers)); });
    } else {
      CU(launch_ex(k_orth_small<16, 4>, dim3(grid), dim3(orth_threads(p->R)), 0, s, 1, p->pdl, dp, (const int*)p->d_orth_small, p_mem, n_workers));
    }
    if (check_launch("k_orth_small")) return 1;
  }
  // 2. tall matrices: one thread-block cluster each
  for (int pass = 0; pass < 2; ++pass) {
    const std::vector<int>& ids = pass == 0 ? p->orth_tall : p->orth_generic;
    const int* d_ids = pass == 0 ? p->d_orth_tall : p->d_orth_generic;
    if (ids.empty()) continue;
    const int csz = pass == 0 ? tcs : kOrthCluster;
    if (pass == 0) {
      const int n_jobs = (int)(merged ? p->orth_jobs_merged.size() : p->orth_jobs_tall.size());
      const int4* d_jobs = merged ? p->d_orth_jobs_merged : p->d_orth_jobs_tall;
      const dim3 grid((unsigned)((n_jobs + csz - 1) / csz * csz));
      if (p->R <= 8) {
        DISPATCH_R8(p->R, {
          if (csz > 8 && !p->orth16_attr) CU(cudaFuncSetAttribute((const void*)k_orth_tall<RR, 4>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
          p->orth16_attr = true;
          CU(launch_ex(k_orth_tall<RR, 4>, grid, dim3(orth_threads(p->R)), 0, s, csz, p->pdl, dp, d_jobs, n_jobs, p_mem, n_workers));
        });
      } else {
        if (!p->orth16_attr) CU(cudaFuncSetAttribute((const void*)k_orth_tall<16, 4>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        p->orth16_attr = true;
        CU(launch_ex(k_orth_tall<16, 4>, grid, dim3(orth_threads(p->R)), 0, s, csz, p->pdl, dp, d_jobs, n_jobs, p_mem, n_workers));
      }
    } else {
      const dim3 grid((unsigned)ids.size() * csz);
      // the generic kernel has no dependency wait of its own: plain stream order
      DISPATCH_R(p->R, { CU(launch_ex(k_orthogonalize<RR>, grid, dim3(kOrthThreads), 0, s, csz, false, dp, d_ids, p_mem, n_workers)); });
    }
    if (check_launch("k_orth_tall")) return 1;
  }
  return 0;
}

static int q_from_mt_impl(psgd_plan* p, const float* p_mem, float* q_mem, void* workspace, psgd_stream stream, bool q_mc);

int psgd_q_from_mt(psgd_plan* p, const float* p_mem, float* q_mem, void* workspace, psgd_stream stream) {
  return q_from_mt_impl(p, p_mem, q_mem, workspace, stream, false);
}

static bool k3_folds(const psgd_plan* p) { return p->d_ctr3 && env_ll("PSGD_K3_FOLD", 1) != 0; }

// q_mc: q_mem is a multicast address (only with the folded second stage, whose stores then go to every rank)
static int q_from_mt_impl(psgd_plan* p, const float* p_mem, float* q_mem, void* workspace, psgd_stream stream, bool q_mc) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (p->tiles3.empty()) return 0;
  if (!p_mem || !q_mem || !workspace) return fail("psgd_q_from_mt: null p_mem/q_mem/workspace");
  if ((uintptr_t)workspace & 15) return fail("psgd_q_from_mt: workspace must be 16-byte aligned");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const DevPlan dp = dev_plan(p);
  float* ws = (float*)workspace;
  const bool fold = k3_folds(p);
  if (q_mc && !fold) return fail("psgd_q_from_mt: multicast output needs the folded second stage");
  {
    PhaseScope ts(p, PSGD_PHASE_Q_FROM_MT, s);
    const unsigned grid = (unsigned)p->tiles3.size();
    DISPATCH_R(p->R, {
      if (set_smem(k_q_from_mt<RR>, p->smem3)) return 1;
      CU(launch_ex(k_q_from_mt<RR>, dim3(grid), dim3(kThreads3), p->smem3, s, 1, p->pdl, dp, p->slot, p_mem, (float*)ws, q_mem, fold ? p->d_ctr3 : (int*)nullptr, (p->tc && (p->tc_mask & 2) && p->tc_now) ? 1 : 0, q_mc ? 1 : 0));
    });
    if (check_launch("k_q_from_mt")) return 1;
  }
  if (!fold) {
    PhaseScope ts(p, PSGD_PHASE_Q_REDUCE, s);
    k_q_reduce<<<(unsigned)p->qtiles.size(), kThreads, 0, s>>>(dp, ws, q_mem);
    if (check_launch("k_q_reduce")) return 1;
  }
  return 0;
}

static int finalize_impl(psgd_plan* p, const float* p_mem, const float* q_mem, float n_workers, const float* flat,
                         psgd_stream stream, psgd_peer* peer = nullptr);

int psgd_finalize(psgd_plan* p, const float* p_mem, const float* q_mem, float n_workers, psgd_stream stream) {
  return finalize_impl(p, p_mem, q_mem, n_workers, nullptr, stream);
}

int psgd_finalize_unpack(psgd_plan* p, const float* p_mem, const float* q_mem, float n_workers, const float* flat,
                         psgd_stream stream) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (p->v_total && !flat) return fail("psgd_finalize_unpack: flat is null");
  return finalize_impl(p, p_mem, q_mem, n_workers, flat, stream);
}

static int finalize_impl(psgd_plan* p, const float* p_mem, const float* q_mem, float n_workers, const float* flat,
                         psgd_stream stream, psgd_peer* peer) {
  if (need_device(p) || need_ptrs(p)) return 1;
  const bool tc4 = p->tc && (p->tc_mask & 4) && p->tc_now;
  const bool fold = flat && p->v_total && !p->tiles4.empty() && !tc4 && p->stream4_now;
  if (flat && p->v_total && !fold) {  // no streamed launch to ride on: separate scatter
    if (psgd_finish(p, flat, nullptr, n_workers, stream)) return 1;
  }
  if (p->tiles4.empty()) return 0;
  if (!p_mem || !q_mem) return fail("psgd_finalize: null p_mem/q_mem");
  if (!(n_workers >= 1.f)) return fail("psgd_finalize: n_workers must be >= 1");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const DevPlan dp = dev_plan(p);
  const unsigned grid = (unsigned)p->tiles4.size();
  PhaseScope ts(p, PSGD_PHASE_FINALIZE, s);
  if (tc4) {  // ranks 16 / 32: tensor pipe
    if (peer && peer_reduce_standalone(peer, 1, const_cast<float*>(q_mem), p->q_total, (float)peer->world, 0, s)) return 1;
    if (!p->tiles4_tc.empty()) {
      DevPlan tp = dp;
      tp.tiles4 = p->d_tiles4_tc;
      const unsigned g = (unsigned)p->tiles4_tc.size();
      if (p->R == 16) {
        if (set_smem(k_finalize_tc<16>, p->tc_smem4)) return 1;
        k_finalize_tc<16><<<g, kTcThreads, p->tc_smem4, s>>>(tp, p->slot, p_mem, q_mem, n_workers, p->tc_wc4);
      } else {
        if (set_smem(k_finalize_tc<32>, p->tc_smem4)) return 1;
        k_finalize_tc<32><<<g, kTcThreads, p->tc_smem4, s>>>(tp, p->slot, p_mem, q_mem, n_workers, p->tc_wc4);
      }
      if (check_launch("k_finalize_tc")) return 1;
    }
    if (!p->tiles4_ntc.empty()) {
      DevPlan fb = dp;
      fb.tiles4 = p->d_tiles4_ntc;
      DISPATCH_R(p->R, {
        if (set_smem(k_finalize<RR>, p->smem4)) return 1;
        k_finalize<RR><<<(unsigned)p->tiles4_ntc.size(), kThreads, p->smem4, s>>>(fb, p->slot, p_mem, q_mem, n_workers);
      });
      if (check_launch("k_finalize")) return 1;
    }
    return 0;
  }
  if (p->stream4_now) {
    StreamPlan sp{p->d_tasks4, p->d_counters + 2 * kCtrRing + 2 * (p->seq4++ % kCtrRing), (int)p->tasks4.size(), p->ring_slot_bytes, p->ring_slots, p->q_bytes, nullptr,
                  p->d_vtiles, const_cast<float*>(flat), fold ? (int)p->vtiles.size() : 0, fold ? 3 : 0, n_workers,
                  p->d_direct_ids, p->d_direct_row_pre, (int)p->direct_ids.size(), 0, 0};
    if (peer && p->stream_grid > p->sm_count) {
      if (peer_reduce_standalone(peer, 1, const_cast<float*>(q_mem), p->q_total, (float)peer->world, 0, s)) return 1;
      peer = nullptr;
    }
    const PeerQ pq = peer_q_args(p, peer, const_cast<float*>(q_mem));
    DISPATCH_R(p->R, {
      if (set_smem(k_finalize_stream<RR>, p->stream_smem)) return 1;
      CU(launch_ex(k_finalize_stream<RR>, dim3(p->stream_grid), dim3(kStreamThreads), p->stream_smem, s, 1, p->pdl, dp, p->slot, sp, p_mem, q_mem, n_workers, pq));
    });
    if (check_launch("k_finalize_stream")) return 1;
    if (!p->tiles4_fb.empty()) {
      DevPlan fb = dp;
      fb.tiles4 = p->d_tiles4_fb;
      DISPATCH_R(p->R, {
        if (set_smem(k_finalize<RR>, p->smem4)) return 1;
        k_finalize<RR><<<(unsigned)p->tiles4_fb.size(), kThreads, p->smem4, s>>>(fb, p->slot, p_mem, q_mem, n_workers);
      });
      if (check_launch("k_finalize")) return 1;
    }
    return 0;
  }
  if (peer && peer_reduce_standalone(peer, 1, const_cast<float*>(q_mem), p->q_total, (float)peer->world, 0, s)) return 1;
  DISPATCH_R(p->R, {
    if (set_smem(k_finalize<RR>, p->smem4)) return 1;
    k_finalize<RR><<<grid, kThreads, p->smem4, s>>>(dp, p->slot, p_mem, q_mem, n_workers);
  });
  return check_launch("k_finalize");
}

int psgd_finish(psgd_plan* p, const float* flat, float* q_mem, float n_workers, psgd_stream stream) {
  if (need_device(p) || need_ptrs(p)) return 1;
  if (!(n_workers >= 1.f)) return fail("psgd_finish: n_workers must be >= 1");
  const long long qn = q_mem ? p->q_total : 0;
  const size_t nvt = flat ? p->vtiles.size() : 0;  // flat == NULL: the scatter was folded into psgd_finalize_unpack
  if (nvt == 0 && qn == 0) return 0;
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  PhaseScope ts(p, PSGD_PHASE_FINISH, s);
  const unsigned grid = (unsigned)(nvt + (size_t)((qn + kVecTileElems - 1) / kVecTileElems));
  k_finish<<<grid, kThreads, 0, s>>>(dev_plan(p), p->slot, flat, (int)nvt, q_mem, qn, n_workers);
  return check_launch("k_finish");
}

// ---- all-reduces taken into the kernels (SURVEY f-2): symmetric-memory handle and the peer-aware phase calls ----------

static PeerLoad peer_load(psgd_peer* peer, int which) {
  PeerLoad pm = which == 0 ? peer->pf : peer->q;
  pm.epoch = which == 0 ? ++peer->epoch_pf : ++peer->epoch_q;
  if (peer->mode == kPeerSlots) {  // everything was pushed into this rank's receive buffer: W local slots
    pm.in[0] = which == 0 ? peer->rx_pf_local : peer->rx_q_local;
    pm.stride = which == 0 ? peer->stride_pf : peer->stride_q;
  }
  return pm;
}

static PeerQ peer_q_args(psgd_plan* p, psgd_peer* peer, float* q_mem) {
  PeerQ pq{};
  pq.pm.mode = kPeerOff;
  if (!peer) return pq;
  pq.pm = peer_load(peer, 1);
  pq.q_out = q_mem;
  pq.n = p->q_total;
  pq.div = (float)peer->world;
  pq.gsync = peer->d_gsync;
  peer->gseq += (unsigned)p->stream_grid;
  pq.gtarget = peer->gseq;
  return pq;
}

static int peer_reduce_standalone(psgd_peer* peer, int which, float* out, long long n, float div, long long div_from,
                                  cudaStream_t s) {
  if (n <= 0) return 0;
  const PeerLoad pm = peer_load(peer, which);
  const unsigned grid = (unsigned)std::min<long long>(128, std::max<long long>(1, (n + kThreads * 4 - 1) / (kThreads * 4)));
  k_peer_reduce_to_local<<<grid, kThreads, 0, s>>>(pm, out, n, div, div_from);
  return check_launch("k_peer_reduce_to_local");
}

int psgd_peer_create(int world, int rank, int device, const void* const* pf_bufs, void* const* pf_signals,
                     const void* pf_multicast, const void* const* q_bufs, void* const* q_signals,
                     const void* q_multicast, int slot0, int mode, psgd_peer** out) {
  if (!out) return fail("out is null");
  *out = nullptr;
  if (world < 2 || world > kMaxPeers) return fail("psgd_peer_create: world size %d not in 2..%d", world, kMaxPeers);
  if (rank < 0 || rank >= world) return fail("psgd_peer_create: bad rank");
  if (device < 0) return fail("psgd_peer_create: needs a CUDA device");
  if (mode != kPeerMultimem && mode != kPeerPull) return fail("psgd_peer_create: mode must be 1 (multimem) or 2 (pull)");
  if (!pf_bufs || !pf_signals || !q_bufs || !q_signals) return fail("psgd_peer_create: null pointer array");
  if (mode == kPeerMultimem && (!pf_multicast || !q_multicast)) return fail("psgd_peer_create: multimem mode needs the multicast addresses");
  if (slot0 < 0) return fail("psgd_peer_create: bad signal slot");
  psgd_peer* h = new psgd_peer();
  h->device = device; h->world = world; h->rank = rank; h->mode = mode;
  for (int w = 0; w < 2; ++w) {
    PeerLoad& pm = w == 0 ? h->pf : h->q;
    const void* const* bufs = w == 0 ? pf_bufs : q_bufs;
    void* const* sigs = w == 0 ? pf_signals : q_signals;
    for (int r = 0; r < world; ++r) {
      if (!bufs[r] || !sigs[r] || ((uintptr_t)bufs[r] & 15)) {
        delete h;
        return fail("psgd_peer_create: rank %d has a null or misaligned (16 bytes) buffer", r);
      }
      pm.in[r] = (const float*)bufs[r];
      pm.sig[r] = (unsigned*)sigs[r];
    }
    pm.mc = (const float*)(w == 0 ? pf_multicast : q_multicast);
    pm.stride = 0;
    if ((uintptr_t)pm.mc & 15) { delete h; return fail("psgd_peer_create: multicast address must be 16-byte aligned"); }
    pm.rank = rank; pm.world = world; pm.slot0 = slot0; pm.mode = mode; pm.epoch = 0;
    pm.light = env_ll("PSGD_PEER_LIGHT", 0) != 0 ? 1 : 0;
  }
  if (cudaSetDevice(device) != cudaSuccess || cudaMalloc((void**)&h->d_gsync, sizeof(unsigned)) != cudaSuccess ||
      cudaMemset(h->d_gsync, 0, sizeof(unsigned)) != cudaSuccess) {
    delete h;
    return fail("psgd_peer_create: device set-up failed");
  }
  *out = h;
  return 0;
}

int psgd_peer_create_push(int world, int rank, int device, void* rx_pf_local, void* rx_pf_multicast, int64_t pf_stride,
                          void* const* pf_signals, void* rx_q_local, void* rx_q_multicast, int64_t q_stride,
                          void* const* q_signals, int slot0, psgd_peer** out) {
  if (!out) return fail("out is null");
  *out = nullptr;
  if (world < 2 || world > kMaxPeers) return fail("psgd_peer_create_push: world size %d not in 2..%d", world, kMaxPeers);
  if (rank < 0 || rank >= world) return fail("psgd_peer_create_push: bad rank");
  if (device < 0) return fail("psgd_peer_create_push: needs a CUDA device");
  if (!rx_pf_local || !rx_pf_multicast || !rx_q_local || !rx_q_multicast || !pf_signals || !q_signals)
    return fail("psgd_peer_create_push: null pointer (push mode needs the multicast mapping of both receive buffers)");
  if (pf_stride <= 0 || q_stride <= 0 || (pf_stride & 3) || (q_stride & 3)) return fail("psgd_peer_create_push: slot strides must be positive multiples of 4 floats");
  if (((uintptr_t)rx_pf_local | (uintptr_t)rx_pf_multicast | (uintptr_t)rx_q_local | (uintptr_t)rx_q_multicast) & 15)
    return fail("psgd_peer_create_push: buffers must be 16-byte aligned");
  if (slot0 < 0) return fail("psgd_peer_create_push: bad signal slot");
  psgd_peer* h = new psgd_peer();
  h->device = device; h->world = world; h->rank = rank; h->mode = kPeerSlots;
  h->rx_pf_local = (float*)rx_pf_local; h->rx_pf_mc = (float*)rx_pf_multicast; h->stride_pf = pf_stride;
  h->rx_q_local = (float*)rx_q_local; h->rx_q_mc = (float*)rx_q_multicast; h->stride_q = q_stride;
  for (int w = 0; w < 2; ++w) {
    PeerLoad& pm = w == 0 ? h->pf : h->q;
    void* const* sigs = w == 0 ? pf_signals : q_signals;
    for (int r = 0; r < world; ++r) {
      if (!sigs[r]) { delete h; return fail("psgd_peer_create_push: rank %d has a null signal pad", r); }
      pm.in[r] = nullptr;
      pm.sig[r] = (unsigned*)sigs[r];
    }
    pm.mc = nullptr; pm.stride = 0;
    pm.rank = rank; pm.world = world; pm.slot0 = slot0; pm.mode = kPeerSlots; pm.epoch = 0;
    pm.light = env_ll("PSGD_PEER_LIGHT", 0) != 0 ? 1 : 0;
  }
  if (cudaSetDevice(device) != cudaSuccess || cudaMalloc((void**)&h->d_gsync, sizeof(unsigned)) != cudaSuccess ||
      cudaMemset(h->d_gsync, 0, sizeof(unsigned)) != cudaSuccess) {
    delete h;
    return fail("psgd_peer_create_push: device set-up failed");
  }
  *out = h;
  return 0;
}

static int push_to_all(const float* local, float* mc, long long n, cudaStream_t s) {
  if (n <= 0) return 0;
  const unsigned grid = (unsigned)std::min<long long>(148, std::max<long long>(1, (n / 4 + kThreads - 1) / kThreads));
  k_push_to_all<<<grid, kThreads, 0, s>>>(local, mc, n);
  return check_launch("k_push_to_all");
}

int psgd_p_from_m_push(psgd_plan* p, psgd_peer* peer, float* q_mem, int fuse_add, int with_flat, int reduce_q_first,
                       psgd_stream stream) {
  if (!peer || peer->mode != kPeerSlots) return fail("psgd_p_from_m_push: needs a push-mode peer handle");
  if (need_device(p) || need_ptrs(p)) return 1;
  if (peer->device != p->device) return fail("psgd_p_from_m_push: plan and peer handle are on different devices");
  if (p->p_total + p->v_total > peer->stride_pf) return fail("psgd_p_from_m_push: receive slot too small for P || flat");
  if (with_flat && fuse_add && !p->have_vge) return fail("psgd_p_from_m_push: fuse_add needs psgd_set_ptrs_fused");
  cudaStream_t s = (cudaStream_t)stream;
  float* loc = peer->rx_pf_local + (long long)peer->rank * peer->stride_pf;
  float* mcp = peer->rx_pf_mc + (long long)peer->rank * peer->stride_pf;
  const bool want_flat = with_flat && p->v_total > 0;
  const bool tc1 = p->tc && (p->tc_mask & 1) && (fuse_add ? p->tcf_now : p->tc_now);
  // The streamed kernel can store straight through the multicast address when it is the only producer; otherwise
  // the results are written to this rank's own slot and pushed by one small extra launch.
  const bool direct = !tc1 && (fuse_add ? p->stream1f_now : p->stream1_now) && p->tiles1_fb.empty() && !p->tiles1.empty() &&
                      !p->has_chunked;  // (a column-chunked matrix adds to its own P in the second pass: ordinary memory only)
  psgd_peer* pq = reduce_q_first ? peer : nullptr;
  if (direct) return p_from_m_impl(p, mcp, q_mem, fuse_add, 1.0f, want_flat ? mcp + p->p_total : nullptr, stream, pq, true);
  if (p_from_m_impl(p, loc, q_mem, fuse_add, 1.0f, want_flat ? loc + p->p_total : nullptr, stream, pq, false)) return 1;
  return push_to_all(loc, mcp, p->p_total + (want_flat ? p->v_total : 0), s);
}

int psgd_q_from_mt_push(psgd_plan* p, psgd_peer* peer, const float* p_mem, void* workspace, psgd_stream stream) {
  if (!peer || peer->mode != kPeerSlots) return fail("psgd_q_from_mt_push: needs a push-mode peer handle");
  if (need_device(p) || need_ptrs(p)) return 1;
  if (peer->device != p->device) return fail("psgd_q_from_mt_push: plan and peer handle are on different devices");
  if (p->q_total > peer->stride_q) return fail("psgd_q_from_mt_push: receive slot too small for Q");
  float* loc = peer->rx_q_local + (long long)peer->rank * peer->stride_q;
  float* mcq = peer->rx_q_mc + (long long)peer->rank * peer->stride_q;
  if (k3_folds(p)) return q_from_mt_impl(p, p_mem, mcq, workspace, stream, true);
  if (q_from_mt_impl(p, p_mem, loc, workspace, stream, false)) return 1;
  return push_to_all(loc, mcq, p->q_total, (cudaStream_t)stream);
}

int psgd_peer_destroy(psgd_peer* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaFree(h->d_gsync);
  delete h;
  return 0;
}

int psgd_p_from_m_peer(psgd_plan* p, psgd_peer* peer, float* p_mem, float* q_mem, int fuse_add, psgd_stream stream) {
  if (!peer) return fail("psgd_p_from_m_peer: null peer handle");
  if (need_device(p) || need_ptrs(p)) return 1;
  if (p->stream_grid > 0 && peer->device != p->device) return fail("psgd_p_from_m_peer: plan and peer handle are on different devices");
  return p_from_m_impl(p, p_mem, q_mem, fuse_add, 1.0f, nullptr, stream, peer);
}

int psgd_finalize_peer(psgd_plan* p, psgd_peer* peer, const float* p_mem, float* q_mem, const float* flat,
                       psgd_stream stream) {
  if (!peer) return fail("psgd_finalize_peer: null peer handle");
  if (need_device(p) || need_ptrs(p)) return 1;
  if (peer->device != p->device) return fail("psgd_finalize_peer: plan and peer handle are on different devices");
  if (p->v_total && !flat) return fail("psgd_finalize_peer: flat is null");
  if (p->tiles4.empty()) {  // no matrix at all: only the scatter of the (already averaged) flat buffer is left
    return psgd_finish(p, flat, nullptr, 1.0f, stream);
  }
  // q_mem holds the MEAN once the kernel's prologue (or the stand-alone reduction) has run: staged with a world of 1
  return finalize_impl(p, p_mem, q_mem, 1.0f, flat, stream, peer);
}

int psgd_orthogonalize_peer(psgd_plan* p, psgd_peer* peer, float* p_mem, float* flat_out, psgd_stream stream) {
  if (!peer) return fail("psgd_orthogonalize_peer: null peer handle");
  if (need_device(p)) return 1;
  if (peer->device != p->device) return fail("psgd_orthogonalize_peer: plan and peer handle are on different devices");
  const bool want_flat = flat_out != nullptr && p->v_total > 0;
  if (p->n_mat == 0 && !want_flat) return 0;
  if (p->n_mat && !p_mem) return fail("psgd_orthogonalize_peer: null p_mem");
  CU(cudaSetDevice(p->device));
  cudaStream_t s = (cudaStream_t)stream;
  const float world = (float)peer->world;
  const int n_small = (int)p->orth_small.size(), n_tall = (int)p->orth_tall.size();
  // what the register-resident kernels cannot take (very tall matrices, rank 32) goes through the stand-alone
  // reduction into the local buffers followed by the ordinary Gram-Schmidt launches
  if (!p->orth_generic.empty() || p->R > 16 || (n_small == 0 && n_tall == 0)) {
    if (p->p_total && peer_reduce_standalone(peer, 0, p_mem, p->p_total, 1.0f, 0, s)) return 1;
    if (want_flat) {
      // the flat part behind P: a second reduction with its own barrier epoch (scalar loads: any alignment)
      PeerLoad pm = peer_load(peer, 0);
      for (int r = 0; r < peer->world; ++r) pm.in[r] += p->p_total;
      if (pm.mc) pm.mc += p->p_total;
      const unsigned grid = (unsigned)std::min<long long>(64, std::max<long long>(1, (p->v_total + kThreads - 1) / kThreads));
      k_peer_reduce_to_local<<<grid, kThreads, 0, s>>>(pm, flat_out, p->v_total, world, 0);
      if (check_launch("k_peer_reduce_to_local")) return 1;
    }
    return orth_impl(p, p_mem, world, stream);
  }
  PhaseScope ts(p, PSGD_PHASE_ORTH, s);
  const DevPlan dp = dev_plan(p);
  const PeerLoad pm = peer_load(peer, 0);
  PeerFlat pf{};
  pf.out = want_flat ? flat_out : nullptr;
  pf.sym_off = p->p_total;
  pf.n = want_flat ? p->v_total : 0;
  pf.div = world;
  const int n_flat_ctas = want_flat ? (int)((p->v_total + peer_flat_per_cta(p->R) - 1) / peer_flat_per_cta(p->R)) : 0;
  const int tcs = p->orth_tall_cluster;
  const bool merged = n_tall > 0 && ((int)p->orth_jobs_merged.size() + n_flat_ctas + tcs - 1) / tcs * tcs <= std::max(p->orth_merge_max_ctas, 148);
  if (n_tall == 0 || !merged) {  // small matrices (and, without tall ones, the flat buffer) in a plain launch
    const bool flat_here = n_tall == 0;
    PeerFlat pf_small = flat_here ? pf : PeerFlat{};
    const unsigned grid = (unsigned)(n_small + (flat_here ? n_flat_ctas : 0));
    if (grid) {
      if (p->R <= 8) {
        DISPATCH_R8(p->R, { CU(launch_ex(k_orth_small_peer<RR, 4>, dim3(grid), dim3(orth_threads(p->R)), 0, s, 1, p->pdl, dp, (const int*)p->d_orth_small, n_small, p_mem, world, pm, pf_small)); });
      } else {
        CU(launch_ex(k_orth_small_peer<16, 4>, dim3(grid), dim3(orth_threads(p->R)), 0, s, 1, p->pdl, dp, (const int*)p->d_orth_small, n_small, p_mem, world, pm, pf_small));
      }
      if (check_launch("k_orth_small_peer")) return 1;
    }
  }
  if (n_tall) {
    // CTAs behind the job table sum the flat buffer
    const int n_jobs = (int)(merged ? p->orth_jobs_merged.size() : p->orth_jobs_tall.size());
    const int4* d_jobs = merged ? p->d_orth_jobs_merged : p->d_orth_jobs_tall;
    const dim3 grid((unsigned)((n_jobs + n_flat_ctas + tcs - 1) / tcs * tcs));
    if (p->R <= 8) {
      DISPATCH_R8(p->R, {
        if (tcs > 8 && !p->orth16_attr_peer) CU(cudaFuncSetAttribute((const void*)k_orth_tall_peer<RR, 4>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        p->orth16_attr_peer = true;
        CU(launch_ex(k_orth_tall_peer<RR, 4>, grid, dim3(orth_threads(p->R)), 0, s, tcs, p->pdl, dp, d_jobs, n_jobs, p_mem, world, pm, pf));
      });
    } else {
      if (!p->orth16_attr_peer) CU(cudaFuncSetAttribute((const void*)k_orth_tall_peer<16, 4>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      p->orth16_attr_peer = true;
      CU(launch_ex(k_orth_tall_peer<16, 4>, grid, dim3(orth_threads(p->R)), 0, s, tcs, p->pdl, dp, d_jobs, n_jobs, p_mem, world, pm, pf));
    }
    if (check_launch("k_orth_tall_peer")) return 1;
  }
  return 0;
}

int psgd_timing_enable(psgd_plan* p, int on) {
  if (need_device(p)) return 1;
  p->timing = on < 0 ? 0u : (unsigned)on;
  return 0;
}

int psgd_timing_read(psgd_plan* p, double* ms_sum, int64_t* count) {
  if (need_device(p)) return 1;
  CU(cudaSetDevice(p->device));
  for (auto& t : p->timers) {
    CU(cudaEventSynchronize(t.b));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, t.a, t.b));
    p->ms_sum[t.phase] += ms;
    p->ms_cnt[t.phase] += 1;
    p->free_timers.push_back(t);
  }
  p->timers.clear();
  for (int i = 0; i < PSGD_N_PHASES; ++i) {
    if (ms_sum) ms_sum[i] = p->ms_sum[i];
    if (count) count[i] = p->ms_cnt[i];
    p->ms_sum[i] = 0;
    p->ms_cnt[i] = 0;
  }
  return 0;
}

int psgd_debug_read(psgd_plan* p, uint64_t* out, int max_words) {
  if (need_device(p)) return -1;
  if (!p->d_dbg) return 0;
  const int n = std::min(max_words, p->stream_grid * kDbgWords);
  if (cudaMemcpy(out, p->d_dbg, (size_t)n * 8, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return n;
}

// ---- stand-alone TensorBuffer packer -----------------------------------------------------

int psgd_tb_create(int n, const int64_t* nelem, int elem_size, int device, psgd_tb** out) {
  if (!out) return fail("out is null");
  *out = nullptr;
  if (n < 0 || (n && !nelem)) return fail("psgd_tb_create: bad tensor list");
  if (elem_size != 1 && elem_size != 2 && elem_size != 4 && elem_size != 8)
    return fail("psgd_tb_create: element size %d not supported (1, 2, 4, 8)", elem_size);
  if (device < 0) return fail("psgd_tb_create: needs a CUDA device");
  psgd_tb* tb = new psgd_tb();
  tb->device = device;
  tb->n = n;
  tb->elem_size = elem_size;
  tb->pre.assign(n + 1, 0);
  for (int i = 0; i < n; ++i) {
    if (nelem[i] < 0) { delete tb; return fail("psgd_tb_create: negative length"); }
    tb->pre[i + 1] = tb->pre[i] + nelem[i];
    for (long long st = 0; st < nelem[i]; st += kVecTileElems) {
      VecTile vt;
      vt.flat_off = tb->pre[i] + st;
      vt.tensor = i;
      vt.start = (int)st;
      vt.count = (int)std::min<long long>(kVecTileElems, nelem[i] - st);
      vt.pad_ = 0;
      tb->tiles.push_back(vt);
    }
  }
  tb->total = tb->pre[n];
  tb->slot_bytes = std::max<size_t>(256, ((size_t)n * 8 + 255) & ~(size_t)255);
  int rc = 0;
  if (cudaSetDevice(device) != cudaSuccess) rc = fail("cudaSetDevice(%d) failed", device);
  if (!rc) rc = upload(&tb->d_tiles, tb->tiles);
  if (!rc && cudaMallocHost((void**)&tb->h_ring, tb->slot_bytes * kSlots) != cudaSuccess) rc = fail("cudaMallocHost failed");
  if (!rc && cudaMalloc((void**)&tb->d_ring, tb->slot_bytes * kSlots) != cudaSuccess) rc = fail("cudaMalloc failed");
  for (int s = 0; s < kSlots && !rc; ++s)
    if (cudaEventCreateWithFlags(&tb->slot_ev[s], cudaEventDisableTiming) != cudaSuccess) rc = fail("cudaEventCreate failed");
  if (rc) {
    std::string keep = g_err;
    psgd_tb_destroy(tb);
    g_err = keep;
    return 1;
  }
  *out = tb;
  return 0;
}

int psgd_tb_destroy(psgd_tb* tb) {
  if (!tb) return 0;
  cudaSetDevice(tb->device);
  cudaFree(tb->d_tiles);
  cudaFree(tb->d_ring);
  if (tb->h_ring) cudaFreeHost(tb->h_ring);
  for (int s = 0; s < kSlots; ++s)
    if (tb->slot_ev[s]) cudaEventDestroy(tb->slot_ev[s]);
  delete tb;
  return 0;
}

static int tb_upload(psgd_tb* tb, const void* const* ptrs, cudaStream_t s, void*** d_out) {
  const int slot = tb->cur;
  tb->cur = (tb->cur + 1) % kSlots;
  if (tb->slot_used[slot]) CU(cudaEventSynchronize(tb->slot_ev[slot]));
  char* h = tb->h_ring + (size_t)slot * tb->slot_bytes;
  char* d = tb->d_ring + (size_t)slot * tb->slot_bytes;
  for (int i = 0; i < tb->n; ++i) {
    if (!ptrs[i] && tb->pre[i + 1] > tb->pre[i]) return fail("tensor %d has a null pointer", i);
    if ((uintptr_t)ptrs[i] % tb->elem_size) return fail("tensor %d is not aligned to its element size", i);
    ((const void**)h)[i] = ptrs[i];
  }
  CU(cudaMemcpyAsync(d, h, (size_t)tb->n * 8, cudaMemcpyHostToDevice, s));
  CU(cudaEventRecord(tb->slot_ev[slot], s));
  tb->slot_used[slot] = true;
  *d_out = (void**)d;
  return 0;
}

int psgd_tb_pack(psgd_tb* tb, const void* const* src, void* flat, psgd_stream stream) {
  if (!tb) return fail("null tensor buffer");
  if (tb->total == 0) return 0;
  if (!src || !flat) return fail("psgd_tb_pack: null argument");
  CU(cudaSetDevice(tb->device));
  cudaStream_t s = (cudaStream_t)stream;
  void** d = nullptr;
  if (tb_upload(tb, src, s, &d)) return 1;
  const unsigned grid = (unsigned)tb->tiles.size();
  switch (tb->elem_size) {
    case 1: k_gather<uint8_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const void* const*)d, (uint8_t*)flat); break;
    case 2: k_gather<uint16_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const void* const*)d, (uint16_t*)flat); break;
    case 4: k_gather<uint32_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const void* const*)d, (uint32_t*)flat); break;
    default: k_gather<uint64_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const void* const*)d, (uint64_t*)flat); break;
  }
  return check_launch("k_gather");
}

int psgd_tb_unpack(psgd_tb* tb, const void* flat, void* const* dst, psgd_stream stream) {
  if (!tb) return fail("null tensor buffer");
  if (tb->total == 0) return 0;
  if (!dst || !flat) return fail("psgd_tb_unpack: null argument");
  CU(cudaSetDevice(tb->device));
  cudaStream_t s = (cudaStream_t)stream;
  void** d = nullptr;
  if (tb_upload(tb, (const void* const*)dst, s, &d)) return 1;
  const unsigned grid = (unsigned)tb->tiles.size();
  switch (tb->elem_size) {
    case 1: k_scatter<uint8_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const uint8_t*)flat, (void* const*)d); break;
    case 2: k_scatter<uint16_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const uint16_t*)flat, (void* const*)d); break;
    case 4: k_scatter<uint32_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const uint32_t*)flat, (void* const*)d); break;
    default: k_scatter<uint64_t><<<grid, kThreads, 0, s>>>(tb->d_tiles, (const uint64_t*)flat, (void* const*)d); break;
  }
  return check_launch("k_scatter");
}

// ---- fused momentum + SGD step (the caller-side loops after reduce(), ddp_init.py:166-178) ---------------

struct psgd_sgd {
  int device = -1;
  int n = 0;
  std::vector<VecTile> tiles;
  VecTile* d_tiles = nullptr;
  size_t slot_bytes = 0;
  char* h_ring = nullptr;
  char* d_ring = nullptr;
  cudaEvent_t slot_ev[kSlots] = {};
  bool slot_used[kSlots] = {};
  int cur = -1;
};

int psgd_sgd_create(int n, const int64_t* nelem, int device, psgd_sgd** out) {
  if (!out) return fail("out is null");
  *out = nullptr;
  if (n < 0 || (n && !nelem)) return fail("psgd_sgd_create: bad tensor list");
  if (device < 0) return fail("psgd_sgd_create: needs a CUDA device");
  psgd_sgd* h = new psgd_sgd();
  h->device = device;
  h->n = n;
  for (int i = 0; i < n; ++i) {
    if (nelem[i] < 0 || nelem[i] > 0x7fffffffLL) { delete h; return fail("psgd_sgd_create: unsupported length"); }
    for (long long st = 0; st < nelem[i]; st += kEwTileElems) {
      VecTile vt;
      vt.flat_off = 0;
      vt.tensor = i;
      vt.start = (int)st;
      vt.count = (int)std::min<long long>(kEwTileElems, nelem[i] - st);
      vt.pad_ = 0;
      h->tiles.push_back(vt);
    }
  }
  h->slot_bytes = std::max<size_t>(256, ((size_t)n * 3 * 8 + 255) & ~(size_t)255);
  int rc = 0;
  if (cudaSetDevice(device) != cudaSuccess) rc = fail("cudaSetDevice(%d) failed", device);
  if (!rc) rc = upload(&h->d_tiles, h->tiles);
  if (!rc && cudaMallocHost((void**)&h->h_ring, h->slot_bytes * kSlots) != cudaSuccess) rc = fail("cudaMallocHost failed");
  if (!rc && cudaMalloc((void**)&h->d_ring, h->slot_bytes * kSlots) != cudaSuccess) rc = fail("cudaMalloc failed");
  for (int s = 0; s < kSlots && !rc; ++s)
    if (cudaEventCreateWithFlags(&h->slot_ev[s], cudaEventDisableTiming) != cudaSuccess) rc = fail("cudaEventCreate failed");
  if (rc) {
    std::string keep = g_err;
    psgd_sgd_destroy(h);
    g_err = keep;
    return 1;
  }
  *out = h;
  return 0;
}

int psgd_sgd_destroy(psgd_sgd* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaFree(h->d_tiles);
  cudaFree(h->d_ring);
  if (h->h_ring) cudaFreeHost(h->h_ring);
  for (int s = 0; s < kSlots; ++s)
    if (h->slot_ev[s]) cudaEventDestroy(h->slot_ev[s]);
  delete h;
  return 0;
}

int psgd_sgd_step(psgd_sgd* h, void* const* params, void* const* grads, void* const* momenta, float lr,
                  float momentum, int first_step, psgd_stream stream) {
  if (!h) return fail("null handle");
  if (h->tiles.empty()) return 0;
  if (!params || !grads || !momenta) return fail("psgd_sgd_step: null pointer array");
  CU(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  const int slot = (h->cur + 1) % kSlots;
  if (h->slot_used[slot]) CU(cudaEventSynchronize(h->slot_ev[slot]));
  char* hp = h->h_ring + (size_t)slot * h->slot_bytes;
  char* dp = h->d_ring + (size_t)slot * h->slot_bytes;
  const size_t n = h->n;
  for (size_t i = 0; i < n; ++i) {
    if (!params[i] || !grads[i] || !momenta[i]) return fail("psgd_sgd_step: tensor %zu has a null pointer", i);
    if (((uintptr_t)params[i] | (uintptr_t)grads[i] | (uintptr_t)momenta[i]) & 3) return fail("psgd_sgd_step: tensor %zu is not 4-byte aligned", i);
    ((void**)hp)[i] = params[i];
    ((void**)hp)[n + i] = grads[i];
    ((void**)hp)[2 * n + i] = momenta[i];
  }
  if (h->cur < 0 || memcmp(hp, h->h_ring + (size_t)h->cur * h->slot_bytes, 3 * n * 8) != 0) {
    CU(cudaMemcpyAsync(dp, hp, 3 * n * 8, cudaMemcpyHostToDevice, s));
    CU(cudaEventRecord(h->slot_ev[slot], s));
    h->slot_used[slot] = true;
    h->cur = slot;
  } else {
    dp = h->d_ring + (size_t)h->cur * h->slot_bytes;
  }
  float* const* d = (float* const*)dp;
  k_momentum_step<<<(unsigned)h->tiles.size(), kThreads, 0, s>>>(h->d_tiles, d, d + n, d + 2 * n, lr, momentum, first_step);
  return check_launch("k_momentum_step");
}

int psgd_peer_allreduce(const void* const* peer_in, void* const* peer_signal, float* out, int64_t n, int rank,
                        int world, int slot0, uint32_t epoch, int device, psgd_stream stream) {
  if (world < 1 || world > kMaxPeers) return fail("psgd_peer_allreduce: world size %d not in 1..%d", world, kMaxPeers);
  if (rank < 0 || rank >= world) return fail("psgd_peer_allreduce: bad rank");
  if (n < 0 || !peer_in || !peer_signal || (n && !out)) return fail("psgd_peer_allreduce: bad argument");
  if (n == 0) return 0;
  PeerArgs a{};
  for (int r = 0; r < world; ++r) {
    if (!peer_in[r] || !peer_signal[r]) return fail("psgd_peer_allreduce: rank %d has a null buffer", r);
    if (((uintptr_t)peer_in[r] & 15) != 0) return fail("psgd_peer_allreduce: buffers must be 16-byte aligned");
    a.in[r] = (const float*)peer_in[r];
    a.sig[r] = (unsigned*)peer_signal[r];
  }
  if ((uintptr_t)out & 15) return fail("psgd_peer_allreduce: out must be 16-byte aligned");
  CU(cudaSetDevice(device));
  cudaStream_t s = (cudaStream_t)stream;
  const long long n4 = (n + 3) / 4;
  const unsigned grid = (unsigned)std::min<long long>(296, std::max<long long>(1, (n4 + kThreads - 1) / kThreads));
  k_peer_allreduce<<<grid, kThreads, 0, s>>>(a, out, n, rank, world, slot0, epoch);
  return check_launch("k_peer_allreduce");
}

int psgd_peer_allreduce_mc(const void* const* peer_in, void* const* peer_signal, void* multicast, float* out, int64_t n,
                           int rank, int world, int slot0, uint32_t epoch, float out_div, int64_t div_from,
                           uint32_t* blocks_done, int device, psgd_stream stream) {
  if (!blocks_done) return fail("psgd_peer_allreduce_mc: blocks_done is null");
  if (!(out_div > 0.f) || div_from < 0) return fail("psgd_peer_allreduce_mc: out_div must be positive, div_from >= 0");
  if (world < 1 || world > kMaxPeers) return fail("psgd_peer_allreduce_mc: world size %d not in 1..%d", world, kMaxPeers);
  if (rank < 0 || rank >= world) return fail("psgd_peer_allreduce_mc: bad rank");
  if (n < 0 || !peer_in || !peer_signal || !multicast || (n && !out)) return fail("psgd_peer_allreduce_mc: bad argument");
  if (n == 0) return 0;
  PeerArgs a{};
  for (int r = 0; r < world; ++r) {
    if (!peer_in[r] || !peer_signal[r]) return fail("psgd_peer_allreduce_mc: rank %d has a null buffer", r);
    a.in[r] = (const float*)peer_in[r];
    a.sig[r] = (unsigned*)peer_signal[r];
  }
  if (((uintptr_t)peer_in[rank] | (uintptr_t)multicast | (uintptr_t)out) & 15)
    return fail("psgd_peer_allreduce_mc: buffers must be 16-byte aligned");
  CU(cudaSetDevice(device));
  cudaStream_t s = (cudaStream_t)stream;
  const long long n4 = (n + 3) / 4;
  // every CTA takes part in both flag barriers, so the grid must be resident at once: at most one CTA per SM
  const unsigned grid = (unsigned)std::min<long long>(128, std::max<long long>(1, (n4 + kThreads - 1) / kThreads));
  k_peer_allreduce_mc<<<grid, kThreads, 0, s>>>(a, (float*)multicast, out, (long long)n, rank, world, slot0, epoch, out_div,
                                                (long long)div_from, blocks_done);
  return check_launch("k_peer_allreduce_mc");
}

int psgd_orthogonalize_one(float* mat, int64_t n, int r, int device, psgd_stream stream) {
  if (n < 0 || r < 0) return fail("psgd_orthogonalize_one: negative shape");
  if (n == 0 || r == 0) return 0;
  if (!mat) return fail("psgd_orthogonalize_one: null matrix");
  if (r > PSGD_MAX_RANK) return fail("psgd_orthogonalize_one: %d columns exceed PSGD_MAX_RANK=%d", r, PSGD_MAX_RANK);
  if (n > 0x7fffffffLL) return fail("psgd_orthogonalize_one: too many rows");
  CU(cudaSetDevice(device));
  cudaStream_t s = (cudaStream_t)stream;
  const int R = pow2_ceil(r);
  DISPATCH_R(R, { k_orthogonalize_one<RR><<<1, kOrthThreads, 0, s>>>(mat, (int)n, r); });
  return check_launch("k_orthogonalize_one");
}

}  // extern "C"

// =======================================================================================================
// psgd_q_from_mt_noalloc: Q = M^T P for one matrix without a plan and without a single allocation.
//
// Who needs this: code that runs where device memory can no longer be requested.  Typical cases are a
// process whose framework allocator has already claimed the GPU (a raw cudaMalloc from a plug-in then fails
// or eats into a budget somebody else manages), a latency-critical inner loop that cannot afford the
// device-wide synchronisation cudaMalloc / cudaFree imply, and hosts that forbid driver allocations after
// start-up.  Such a caller still owns its inputs and its output; what it cannot supply is the scratch space
// the split-n reduction needs for its per-row-block partial sums.  Here that scratch is a fixed array inside
// the library's device image: PSGD_NOALLOC_SCRATCH_BYTES on every GPU that loads the library, held until
// the process exits.
//
// Rules that follow from sharing one array:
//   * the first call on a device looks the array's device address up (cudaGetSymbolAddress, with that
//     device current) and keeps it; make that first call before any stream capture begins;
//   * requests whose partials do not fit are refused (the message states the bytes needed and the limit);
//   * the array keeps whatever the previous call left in it.  Nothing has to start at zero: the first kernel
//     overwrites every partial the second kernel reads, so stale contents are never observed;
//   * two calls that overlap on the GPU (different streams, or different host threads on one device) would
//     write the same partials: run them back to back on one stream, or order them yourself.
// =======================================================================================================
namespace {

static __device__ __align__(256) float g_noalloc_scratch[PSGD_NOALLOC_SCRATCH_BYTES / sizeof(float)];

constexpr int kNoAllocMaxDevices = 64;
float* g_noalloc_ptr[kNoAllocMaxDevices] = {};  // device address of g_noalloc_scratch, per device
std::atomic<bool> g_noalloc_ready[kNoAllocMaxDevices];

// stage 1 for a single matrix: the tile geometry is derived from blockIdx instead of plan tables
template <int R>
__global__ void __launch_bounds__(kThreads, (R <= 8) ? 2 : 1)
k_q_from_mt_single(const float* __restrict__ M, const float* __restrict__ P, float* __restrict__ scratch, int n, int m,
                   int r, int cgw, int cpt, int t_rows, int ncol, int vec_ok) {
  extern __shared__ __align__(16) float smem[];
  MatDesc md;
  md.p_off = 0; md.q_off = 0; md.ws_off = 0;
  md.n = n; md.m = m; md.r = r;
  md.cgw = cgw; md.cpt = cpt; md.t3_rows = t_rows;
  md.nblk3 = (n + t_rows - 1) / t_rows;
  md.pad_ = 0;
  ColTile t;
  t.mat = 0;
  t.blk = (int)(blockIdx.x / ncol);
  t.row0 = t.blk * t_rows;
  t.nrows = min(t_rows, n - t.row0);
  t.col0 = (int)(blockIdx.x % ncol) * cgw * cpt;
  t.pad0 = t.pad1 = t.pad2 = 0;
  if (cpt == 4) {
    if (vec_ok) q_tile<R, 4, true>(md, t, M, P, scratch, smem);
    else q_tile<R, 4, false>(md, t, M, P, scratch, smem);
  } else {
    q_tile<R, 1, false>(md, t, M, P, scratch, smem);
  }
}

// stage 2 for a single matrix: Q[e] = sum of the row-block partials, in block order
__global__ void __launch_bounds__(kThreads)
k_q_reduce_single(const float* __restrict__ scratch, float* __restrict__ Q, long long mr, int nblk) {
  const long long e = (long long)blockIdx.x * kThreads + threadIdx.x;
  if (e >= mr) return;
  float s = 0.f;
  for (int b = 0; b < nblk; ++b) s += scratch[(size_t)b * mr + e];
  Q[e] = s;
}

}  // namespace

extern "C" int psgd_q_from_mt_noalloc(const float* mat, const float* p_hat, float* q_out, int64_t n, int64_t m, int r,
                                      int device, psgd_stream stream) {
  if (n < 0 || m < 0 || r < 0) return fail("psgd_q_from_mt_noalloc: negative shape");
  if (n > 0x7fffffffLL || m > 0x7fffffffLL) return fail("psgd_q_from_mt_noalloc: shape out of range");
  if (r > PSGD_MAX_RANK) return fail("psgd_q_from_mt_noalloc: r = %d exceeds PSGD_MAX_RANK = %d", r, PSGD_MAX_RANK);
  if (r > n || r > m) return fail("psgd_q_from_mt_noalloc: r = %d exceeds min(n, m)", r);
  if (device < 0 || device >= kNoAllocMaxDevices) return fail("psgd_q_from_mt_noalloc: device %d out of range", device);
  if (m == 0 || r == 0) return 0;
  if (!q_out || (n > 0 && (!mat || !p_hat))) return fail("psgd_q_from_mt_noalloc: null pointer");
  if (((uintptr_t)mat | (uintptr_t)p_hat | (uintptr_t)q_out) & 3) return fail("psgd_q_from_mt_noalloc: pointers must be 4-byte aligned");

  // row blocks of 128..512 rows; every block leaves one m x r partial in the scratch array
  const int t_rows = (int)std::min<int64_t>(512, std::max<int64_t>(128, (n + 31) / 32));
  const int nblk = n ? (int)((n + t_rows - 1) / t_rows) : 0;
  const long long mr = (long long)m * r;
  const long long need = (long long)std::max(nblk, 1) * mr * (long long)sizeof(float);
  if (need > (long long)PSGD_NOALLOC_SCRATCH_BYTES)
    return fail("psgd_q_from_mt_noalloc: %lld x %lld at r = %d needs %lld bytes of scratch (%d row blocks x m x r x 4), "
                "the built-in array holds %lld; use psgd_q_from_mt with a caller-provided workspace instead",
                (long long)n, (long long)m, r, need, std::max(nblk, 1), (long long)PSGD_NOALLOC_SCRATCH_BYTES);

  CU(cudaSetDevice(device));
  if (!g_noalloc_ready[device].load(std::memory_order_acquire)) {
    // set-up, once per device and with that device current: the name of a __device__ array is only a
    // host-side stand-in, the kernels need the address the runtime reports for it
    float* ptr = nullptr;
    CU(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    if (!ptr) return fail("psgd_q_from_mt_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[device] = ptr;
    g_noalloc_ready[device].store(true, std::memory_order_release);
  }
  float* scratch = g_noalloc_ptr[device];
  cudaStream_t s = (cudaStream_t)stream;

  if (n == 0) {  // empty sum: Q = 0
    k_q_reduce_single<<<(unsigned)((mr + kThreads - 1) / kThreads), kThreads, 0, s>>>(scratch, q_out, mr, 0);
    return check_launch("k_q_reduce_single");
  }

  const int R = pow2_ceil(r);
  const int cpt = (m % 4 == 0) ? 4 : 1;
  const long long groups = m / cpt;
  int cgw = 1;
  while (cgw * 2 <= 128 && groups % (cgw * 2) == 0) cgw *= 2;
  if (cgw < 16) cgw = (int)std::min<long long>(128, pow2_ceil(groups));
  const int ncol = (int)((groups + cgw - 1) / cgw);
  const int RL = kThreads / cgw;
  const size_t smem = ((size_t)t_rows * R + (size_t)RL * cpt * R * (cgw + 1)) * sizeof(float);
  const int vec_ok = (cpt == 4 && ((uintptr_t)mat & 15) == 0) ? 1 : 0;
  const unsigned grid = (unsigned)nblk * (unsigned)ncol;
  DISPATCH_R(R, {
    if (set_smem(k_q_from_mt_single<RR>, smem)) return 1;
    k_q_from_mt_single<RR><<<grid, kThreads, smem, s>>>(mat, p_hat, scratch, (int)n, (int)m, r, cgw, cpt, t_rows, ncol, vec_ok);
  });
  if (check_launch("k_q_from_mt_single")) return 1;
  k_q_reduce_single<<<(unsigned)((mr + kThreads - 1) / kThreads), kThreads, 0, s>>>(scratch, q_out, mr, nblk);
  return check_launch("k_q_reduce_single");
}


#ifdef PSGD_ORTH_TS
extern "C" int psgd_debug_orth_ts(unsigned long long* out) {
  return (int)cudaMemcpyFromSymbol(out, psgd::g_orth_ts, sizeof(unsigned long long) * 64 * 12);
}
#endif
