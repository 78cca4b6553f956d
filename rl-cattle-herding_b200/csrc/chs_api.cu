// C ABI of the simulator (include/chs.h): handle management, launches, host-buffer path, state I/O.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cstdlib>
#include <vector>

#include "../../include/chs.h"
#include "chs_kernel.cuh"

using namespace chs;

static thread_local char g_err[512] = "";

// NVTX ranges around the entry points: nvtx3 is header-only and resolves the injection library at
// run time, so without a profiler attached these are a pointer test.
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
static inline void nvtx_push(const char* name) { nvtxRangePushA(name); }
static inline void nvtx_pop() { nvtxRangePop(); }
#else
static inline void nvtx_push(const char*) {}
static inline void nvtx_pop() {}
#endif

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CUDA_TRY(expr)                                                                       \
  do {                                                                                       \
    cudaError_t err__ = (expr);                                                              \
    if (err__ != cudaSuccess)                                                                \
      return fail(CHS_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__), __FILE__, __LINE__); \
  } while (0)

static const double h_curr[k::NUM_LEVELS][k::C_NCOLS] = {CHS_CURRICULUM_ROWS};

struct ChsHandle {
  ChsParams prm;
  KParams kp;
  int E, device;
  bool f64;
  size_t tsz;                 // sizeof(T)
  void *drone, *rpm, *cpos, *cvel, *envf, *epret, *spawn, *path;
  int4 *envi, *envj;
  double* stats;
  double* tab;
  void* own2;
  float* zero_page;      // 32 * stage_stride zero floats (KArgs::zero_page)
  void* pos0;
  int* done_list;
  int* done_count;
  int step_mode;      // 0 auto, 1 fused (one kernel), 2 split (drone_kernel + step_kernel)
  // experiment knobs, read from the environment once at creation (tools/residency_probe.py and A/B runs):
  // CHS_SMEM_PAD=<bytes> extra dynamic shared memory per CTA, CHS_NO_BULK=1 warp copy-out instead of bulk stores,
  // CHS_FULL_KERNEL=1 never use the lean build, CHS_DEFERRED_RESET=1 split mode lists finished envs and rebuilds them in reset_list_kernel (A/B runs)
  unsigned char* d_small;           // host path: count | reward | flags | done idx in one block (chs_api_io.inc)
  unsigned char* h_small;           // pinned mirror
  struct { const void* ptr; size_t bytes; } zeroed_host[4];   // pageable host buffers whose padding rows were zero-filled (chs_api_io.inc)
  unsigned zeroed_next;
  bool knob_early_rows;             // CHS_EARLY_ROWS=1: the host path stores the live rows before phase 3, finished envs included (A/B runs)
  bool knob_host_copy;              // CHS_HOST_COPY=1: the host path returns reward / flags / indices through its staging copy (A/B runs)
  bool knob_no_pdl;                 // CHS_NO_PDL=1: plain launches, no programmatic dependent launch
  bool knob_no_direct;              // CHS_NO_DIRECT=1: the host path stages through HBM even for pinned buffers
  int knob_smem_pad;
  bool knob_full, knob_deferred_reset;
  int *dbg_nb, *dbg_herd;
  long long* dbg_clock;
  int threads, smem;
  int reg_variant;            // 0 auto, 1 latency (128 regs), 2 occupancy (64 regs)
  bool spawn_set;
  unsigned* amask;       // alpha-term in-range masks per spawn row (alpha_mask_kernel)
  void* agrad; void* aw; int aw_stride; double2* acent;   // alpha-term weight table per spawn row (alpha_weight_kernel), valid with amask
  CUtensorMap tm_cpos, tm_cvel;   // fp32: herd tiles of the launch shape in use (configure_launch); valid when kp.tma_ok
  bool knob_no_tma;      // CHS_NO_TMA=1: stage the herd with cp.async chunks (A/B runs)
  bool amask_ok;         // every env's cattle sit on its spawn row: a full reset happened and nobody injected positions
  // device-side buffers of the host path (allocated on first use)
  float *d_actions, *d_obs, *d_reward, *d_termobs;
  uint8_t *d_term, *d_trunc, *d_mask;
  int *d_count, *d_idx;
  bool obs_dst_is_host;             // set by chs_step_host around its launch: obs / terminal obs point into pinned host memory
  int* count_next;                  // host path: the ping-pong partner of the counter in use (zeroed by the launch)
  int* d_cta_done;                  // host path: CTA ticket counter (chs_kernel.cuh: the last CTA publishes the finished-env count)
  int* h_count;                     // host path: pinned word the kernel writes that count to; h_count_dev = its device alias
  int* h_count_dev;
  int* pub_cta_done;                // set by chs_step_host around its launch when the results go straight to pinned host arrays
  int* pub_host_count;
  unsigned host_calls;
  struct { const void* host; void* dev; } alias_cache[8];   // pinned host buffer -> device alias (chs_api_io.inc)
  unsigned alias_next;
};

struct DeviceGuard {
  int prev;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ---- static scratch for chs_step_static_scratch (no counterpart in the reference; added
// for callers that may not call cudaMalloc: code under CUDA graph capture, tight serving loops) -----------------
// One file-scope device array per GPU that loads this library, held until the process exits.  In HOST code the name
// `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per device with
// that device current (resolve_static_scratch, called from chs_create) and that pointer is what the kernels receive.
// Layout used by a call:  [0, 256) finished-env counter | [256, 256 + 4E rounded up to 256) finished-env list |
//                         D*E 4-vectors of the split step's (ang_v, yaw) hand-over between drone_kernel and step_kernel.
static __device__ __align__(256) unsigned char g_static_scratch[CHS_STATIC_SCRATCH_BYTES];

static const int kMaxDevices = 64;
static void* g_static_scratch_addr[kMaxDevices] = {nullptr};

// Caller has made `device` current.  Idempotent; no allocation (the array is part of the loaded module image).
static int resolve_static_scratch(int device) {
  if (device < 0 || device >= kMaxDevices) return fail(CHS_ERR_INVALID, "device %d outside the static-scratch table", device);
  if (g_static_scratch_addr[device]) return CHS_OK;
  void* ptr = nullptr;
  const cudaError_t err = cudaGetSymbolAddress(&ptr, g_static_scratch);
  if (err != cudaSuccess || !ptr)
    return fail(CHS_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch) failed on device %d: %s", device, cudaGetErrorString(err));
  g_static_scratch_addr[device] = ptr;
  return CHS_OK;
}

extern "C" int chs_abi_version(void) { return CHS_ABI_VERSION; }
extern "C" const char* chs_last_error(void) { return g_err; }

static int action_width(int act) {
  switch (act) {   // BaseRLAviary._actionSpace :115-123
    case CHS_ACT_VEL: case CHS_ACT_RPM: return 4;
    case CHS_ACT_PID: return 3;
    case CHS_ACT_ONE_D_RPM: case CHS_ACT_ONE_D_PID: return 1;
  }
  return -1;
}

template <typename T> static SmemLayout layout_for(const KParams& kp, int G) {
  return smem_layout<T>(kp.D, kp.M, G, kp.stage_stride, kp.zero_bytes);
}

static int smem_for(const ChsHandle* h, int G) {
  return h->f64 ? layout_for<double>(h->kp, G).total : layout_for<float>(h->kp, G).total;
}

static const int kSmemCap = 200 * 1024;

// Tensor map over one herd array seen as a row-major [M][2E] float matrix (row = cow, 2 floats per env), box = M rows x 2G
// floats = the [cow][G] shared-memory tile of one CTA.  cuTensorMapEncodeTiled comes from the driver through the runtime
// (no link against libcuda).  Returns false where TMA cannot express the tile (odd E, one env per CTA, more than 256 cows).
typedef CUresult (*TmapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static bool herd_tensor_map(CUtensorMap* out, void* base, int E, int M, int G) {
  static TmapEncodeFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (TmapEncodeFn)sym;
    else
      cudaGetLastError();
  }
  if (!fn || (E & 1) || G < 2 || M < 1 || M > 256) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)2 * (cuuint64_t)E, (cuuint64_t)M};
  const cuuint64_t gstride[1] = {(cuuint64_t)E * 8u};
  const cuuint32_t box[2] = {(cuuint32_t)(2 * G), (cuuint32_t)M};
  const cuuint32_t estr[2] = {1, 1};
  return fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int configure_launch(ChsHandle* h, int gact, int threads) {
  // envs_per_block need not be a power of two: the shared-memory arrays keep a power-of-two stride G >= gact and the CTA
  // leaves the slots beyond gact empty.  An even count keeps the first env of every CTA on a 16-byte boundary of the herd rows:
  // the async 16-byte copies need it, and so do the TMA tile loads (a box that starts 8 bytes off raises "illegal
  // instruction" on the B200 -- measured with 7 envs per CTA).
  if (gact < 1 || gact > 32) return fail(CHS_ERR_INVALID, "envs_per_block must be in 1..32");
  int G = 1;
  while (G < gact) G <<= 1;
  if (gact != G && (gact & 1)) return fail(CHS_ERR_INVALID, "envs_per_block must be a power of two or even");
  if (threads < 32 || threads > 512 || (threads & 31)) return fail(CHS_ERR_INVALID, "threads_per_block must be a multiple of 32 in 32..512");
  int smem = smem_for(h, G);
  smem += h->knob_smem_pad;
  // the opt-in limit is a per-function, process-wide attribute: always raise it to the cap so that handles with
  // different shared-memory needs can coexist (the launch's own size decides occupancy)
  if (smem > kSmemCap) return fail(CHS_ERR_INVALID, "shared memory need %d B exceeds the per-CTA budget", smem);
  DeviceGuard dg(h->device);
#define CHS_SMEM_ATTR(K) CUDA_TRY(cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemCap))
#define CHS_SMEM_ATTR4(TT, MT, MB, FL)                                                                  \
  CHS_SMEM_ATTR((step_kernel<TT, MT, MB, false, false, FL>)); CHS_SMEM_ATTR((step_kernel<TT, MT, MB, true, false, FL>)); \
  CHS_SMEM_ATTR((step_kernel<TT, MT, MB, false, true, FL>));  CHS_SMEM_ATTR((step_kernel<TT, MT, MB, true, true, FL>))
  if (h->f64) {
    CHS_SMEM_ATTR4(double, 512, 1, true);
    CHS_SMEM_ATTR(reset_kernel<double>);
  } else {
    CHS_SMEM_ATTR4(float, 512, 1, true); CHS_SMEM_ATTR4(float, 256, 4, true);
    CHS_SMEM_ATTR4(float, 512, 1, false); CHS_SMEM_ATTR4(float, 256, 4, false);
    CHS_SMEM_ATTR(reset_kernel<float>);
  }
#undef CHS_SMEM_ATTR4
#undef CHS_SMEM_ATTR
  h->kp.G = G;
  h->kp.gact = gact;
  h->kp.lay = h->f64 ? layout_for<double>(h->kp, G) : layout_for<float>(h->kp, G);
  h->kp.cp_ok = h->f64 ? 1 : ((G >= 2 && (h->E & 1) == 0) ? 1 : 0);
  h->kp.tma_ok = (!h->f64 && h->kp.cp_ok && !h->knob_no_tma && herd_tensor_map(&h->tm_cpos, h->cpos, h->E, h->kp.M, G) &&
                  herd_tensor_map(&h->tm_cvel, h->cvel, h->E, h->kp.M, G)) ? 1 : 0;
  h->threads = threads;
  h->smem = smem;
  return CHS_OK;
}

// per-env widths of the reward and flag outputs: the MARL dicts have one entry per drone, the flag dicts one more
// for "__all__" (rllib_envs/BaseAviary.py:426-430)
static inline int reward_width(const ChsHandle* h) { return h->kp.packing == CHS_PACK_MARL ? h->kp.D : 1; }
static inline int flag_width(const ChsHandle* h) { return h->kp.packing == CHS_PACK_MARL ? h->kp.D + 1 : 1; }

extern "C" int chs_set_launch(ChsHandle* h, int32_t envs_per_block, int32_t threads_per_block) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  return configure_launch(h, envs_per_block, threads_per_block);
}

extern "C" int chs_create(const ChsParams* prm, int32_t num_envs, int32_t device, ChsHandle** out) {
  if (!prm || !out) return fail(CHS_ERR_INVALID, "null argument");
  if (prm->struct_size != (int32_t)sizeof(ChsParams))
    return fail(CHS_ERR_INVALID, "ChsParams.struct_size %d != %zu (ABI mismatch)", prm->struct_size, sizeof(ChsParams));
  if (num_envs < 1) return fail(CHS_ERR_INVALID, "num_envs must be >= 1");
  if (prm->num_cattle < 1 || prm->num_cattle > 4096) return fail(CHS_ERR_INVALID, "num_cattle out of range");
  if (prm->max_drones < 1 || prm->max_drones > CHS_MAX_DRONES) return fail(CHS_ERR_INVALID, "max_drones must be 1..12");
  if (prm->ctrl_freq < 1 || prm->pyb_freq < 1 || prm->pyb_freq % prm->ctrl_freq != 0)
    return fail(CHS_ERR_INVALID, "pyb_freq is not divisible by ctrl_freq");   // BaseAviary.py:102-103
  const int A = action_width(prm->act_type);
  if (A < 0) return fail(CHS_ERR_INVALID, "unknown act_type %d", prm->act_type);
  if (prm->start_level < 0 || prm->start_level >= k::NUM_LEVELS) return fail(CHS_ERR_INVALID, "start_level must be 0..7");
  if (prm->packing != CHS_PACK_CTDE && prm->packing != CHS_PACK_MARL) return fail(CHS_ERR_INVALID, "unknown packing %d", prm->packing);
  if (prm->aero_flags < 0 || prm->aero_flags > 7) return fail(CHS_ERR_INVALID, "aero_flags must be a combination of bits 0..2");
  if (prm->marl_wrapper_quirks && prm->packing != CHS_PACK_MARL) return fail(CHS_ERR_INVALID, "marl_wrapper_quirks needs the MARL packing");
  if (prm->fixed_num_drones < 0 || prm->fixed_num_drones > prm->max_drones)
    return fail(CHS_ERR_INVALID, "fixed_num_drones must be 0..max_drones");
  int R = prm->obs_rows == 0 ? (prm->packing == CHS_PACK_MARL ? prm->max_drones : CHS_MAX_DRONES) : prm->obs_rows;
  if (R != CHS_MAX_DRONES && R != prm->max_drones) return fail(CHS_ERR_INVALID, "obs_rows must be 12 or max_drones");
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail(CHS_ERR_INVALID, "device %d not present (%d devices)", device, ndev);

  ChsHandle* h = (ChsHandle*)calloc(1, sizeof(ChsHandle));
  if (!h) return fail(CHS_ERR_NOMEM, "out of host memory");
  h->prm = *prm;
  h->E = num_envs;
  h->device = device;
  h->f64 = prm->precision == 1;
  h->tsz = h->f64 ? 8 : 4;
  KParams& kp = h->kp;
  kp.E = num_envs; kp.D = prm->max_drones; kp.M = prm->num_cattle;
  kp.K = prm->pyb_freq / prm->ctrl_freq; kp.A = A; kp.R = R;
  const int ABS = (int)(0.15 * prm->ctrl_freq);                 // BaseRLAviary.py:69
  kp.L = 10 + 4 * 2 + 16 * 2 + ABS * A;                         // BaseRLAviary.py:262
  kp.act = prm->act_type; kp.packing = prm->packing; kp.auto_reset = prm->auto_reset ? 1 : 0;
  kp.n_sims = 0; kp.spawn_stride = 0;
  const double* row = h_curr[prm->start_level];
  // fixed at construction, never refreshed on level-up (CattleAviary.py:66,72-73)
  kp.fixed_n = prm->fixed_num_drones;
  kp.min_n = (int)row[k::C_MIN_DRONES]; kp.max_n = (int)row[k::C_MAX_DRONES];
  {
    // truncated iff step_counter / CTRL_FREQ > EPISODE_LEN_SEC in float64 (CattleAviary.py:545-550)
    const double ep = row[k::C_EP_LEN];
    long long lim = (long long)floor(ep * prm->ctrl_freq) + 2;
    while (lim > 0 && (double)lim / (double)prm->ctrl_freq > ep) --lim;
    kp.trunc_step_limit = (int)lim;
  }
  kp.ctrl_freq = prm->ctrl_freq;
  kp.env_id_offset = prm->env_id_offset;
  kp.stage_copy = (kp.D * kp.L + 3) & ~3;
  // env stride of the staging area: a multiple of 4 floats (16-byte bulk-store sources) that is 4 mod 8, so that the
  // consecutive envs held by the lanes of a warp fall into 8 different banks (a stride of 0 mod 8 folds them onto 4 or fewer)
  kp.stage_stride = kp.stage_copy;
  while ((kp.stage_stride & 7) != 4) kp.stage_stride += 4;
  kp.vec4_ok = ((kp.R * kp.L) % 4 == 0) ? 1 : 0;
  kp.bulk_ok = (kp.vec4_ok && !getenv("CHS_NO_BULK")) ? 1 : 0;
  // persistent padding needs real padding rows and float4-addressable images (the reset kernel zero-fills them)
  kp.pad_persist = (prm->persistent_padding && R > prm->max_drones && kp.vec4_ok) ? 1 : 0;
  // dead columns of a live row under the same contract: neighbour slots 2..11 of the sb3 observation (cols 14..33)
  kp.dead_lo = kp.dead_hi = 0;
  if (prm->persistent_padding && kp.bulk_ok && kp.packing == CHS_PACK_CTDE && kp.L >= 10 + 2 * CHS_MAX_DRONES && !getenv("CHS_NO_DEAD_SKIP")) {
    kp.dead_lo = 10 + 2 * 2;                 // own state + active_neighbors = 2 slots (BaseAviary.py:117)
    kp.dead_hi = 10 + 2 * CHS_MAX_DRONES;    // padded to GLOBAL_MAX_NUM_DRONES slots (BaseRLAviary.py:313-314)
  }
  {
    const int total = kp.R * kp.L * 4, c1 = kp.stage_copy * 4 < total ? kp.stage_copy * 4 : total;
    kp.zero_bytes = (kp.bulk_ok && !kp.pad_persist) ? total - c1 : 0;
  }
  h->step_mode = prm->reserved0;
  { const char* v = getenv("CHS_SMEM_PAD"); h->knob_smem_pad = v ? atoi(v) : 0; }
  h->knob_full = getenv("CHS_FULL_KERNEL") != nullptr;
  kp.early_ok = getenv("CHS_NO_EARLY") ? 0 : 1;
  h->knob_deferred_reset = getenv("CHS_DEFERRED_RESET") != nullptr;
  h->knob_no_tma = getenv("CHS_NO_TMA") != nullptr;
  h->knob_no_direct = getenv("CHS_NO_DIRECT") != nullptr;
  h->knob_no_pdl = getenv("CHS_NO_PDL") != nullptr;
  h->knob_host_copy = getenv("CHS_HOST_COPY") != nullptr;
  h->knob_early_rows = getenv("CHS_EARLY_ROWS") != nullptr;
  kp.split = 0;
  kp.cattle_integrate = prm->cattle_integrate ? 1 : 0;
  kp.wrapper_quirks = (prm->marl_wrapper_quirks && prm->packing == CHS_PACK_MARL) ? 1 : 0;
  kp.aero_flags = prm->aero_flags;
  kp.act4_ok = (A == 4) ? 1 : 0;
  kp.seed_lo = (uint32_t)(prm->seed & 0xffffffffu); kp.seed_hi = (uint32_t)(prm->seed >> 32);

  DeviceGuard dg(device);
  const size_t E = (size_t)num_envs, D = kp.D, M = kp.M, ts = h->tsz;
  cudaError_t err = cudaSuccess;
#define ALLOC0(ptr, bytes)                                                        \
  if (err == cudaSuccess) { err = cudaMalloc((void**)&(ptr), (bytes));             \
    if (err == cudaSuccess) err = cudaMemset((ptr), 0, (bytes)); }
  ALLOC0(h->drone, F_COUNT * D * E * 4 * ts);
  ALLOC0(h->rpm, D * E * 4 * ts);
  ALLOC0(h->cpos, M * E * 2 * ts);
  ALLOC0(h->zero_page, (size_t)32 * kp.stage_stride * 4);
  ALLOC0(h->cvel, M * E * 2 * ts);
  ALLOC0(h->envf, E * sizeof(double2));
  ALLOC0(h->epret, E * sizeof(double));
  ALLOC0(h->path, D * E * ts);
  ALLOC0(h->envi, E * sizeof(int4));
  ALLOC0(h->envj, E * sizeof(int4));
  ALLOC0(h->stats, (size_t)CHS_STAT_ROWS * CHS_NUM_STATS * sizeof(double));
  ALLOC0(h->tab, 152 * sizeof(double));
  ALLOC0(h->own2, D * E * 4 * ts);
  ALLOC0(h->done_list, E * sizeof(int));
  ALLOC0(h->done_count, sizeof(int));
  if (prm->aero_flags & 4) { ALLOC0(h->pos0, D * E * 4 * ts); }
#undef ALLOC0
  if (err != cudaSuccess) {
    chs_destroy(h);
    return fail(CHS_ERR_CUDA, "device allocation failed: %s", cudaGetErrorString(err));
  }
  // set-up of the static scratch used by chs_step_static_scratch: symbol address for THIS device (current via `dg`)
  { const int src = resolve_static_scratch(device); if (src != CHS_OK) { chs_destroy(h); return src; } }
  // env scalars: level = start level, prev_cent_dist = None (NaN), everything else zero
  {
    std::vector<int4> ej(E);
    for (size_t e = 0; e < E; ++e) ej[e] = make_int4(prm->start_level, 0, 0, 0);
    err = cudaMemcpy(h->envj, ej.data(), E * sizeof(int4), cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
      std::vector<double2> ef(E, make_double2(NAN, 0.0));
      err = cudaMemcpy(h->envf, ef.data(), E * sizeof(double2), cudaMemcpyHostToDevice);
    }
    if (err != cudaSuccess) { chs_destroy(h); return fail(CHS_ERR_CUDA, "state initialisation failed: %s", cudaGetErrorString(err)); }
  }
  {
    std::vector<double> tab(152, 0.0);
    for (int l = 0; l < k::NUM_LEVELS; ++l) for (int c = 0; c < k::C_NCOLS; ++c) tab[l * k::C_NCOLS + c] = h_curr[l][c];
    for (int n = 1; n <= CHS_MAX_DRONES; ++n) tab[CHS_TAB_INV_N + n] = 1.0 / n;
    tab[CHS_TAB_APPROACH] = 5.0 / (k::SPEED_LIMIT / (double)h->kp.ctrl_freq + 1e-6);
    tab[CHS_TAB_INV_M] = h->kp.M > 0 ? 1.0 / (double)h->kp.M : 0.0;
    err = cudaMemcpy(h->tab, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (err != cudaSuccess) { chs_destroy(h); return fail(CHS_ERR_CUDA, "table upload failed: %s", cudaGetErrorString(err)); }
  }
  // launch shape: the largest power-of-two group that still gives every one of the 148 SMs a CTA, shared memory permitting,
  // with 256 threads -- (16, 256) at 3 drones / 20 cattle x 4096 envs, (4, 256) at 8 / 100 x 1024, (32, 256) from 4736 envs
  // up: the best power-of-two shapes measured.  The wave-fitted shapes (28 envs per CTA) are a few per cent
  // better still where the batch is about one wave; the Python layer's autotune_launch() finds them.
  int G = prm->envs_per_block, threads = prm->threads_per_block;
  if (G == 0) {
    G = 32;
    // not below two envs per CTA when that fits: one env per CTA cannot stage the herd with 16-byte async copies and
    // falls to the full build of the kernel (same results up to fp32 contraction, but not bit for bit)
    while (G > 2 && (smem_for(h, G) > 64 * 1024 || (num_envs + G - 1) / G < 148)) G >>= 1;
    if (G == 2 && (num_envs < 2 || smem_for(h, 2) > 64 * 1024)) G = 1;
  }
  if (threads == 0) threads = (G * (kp.D + kp.M) < 128) ? 128 : 256;
  int rc = configure_launch(h, G, threads);
  if (rc != CHS_OK) { chs_destroy(h); return rc; }
  *out = h;
  return CHS_OK;
}

extern "C" int chs_destroy(ChsHandle* h) {
  if (!h) return CHS_OK;
  DeviceGuard dg(h->device);
  void* ptrs[] = {h->drone, h->rpm, h->cpos, h->cvel, h->envf, h->epret, h->path, h->spawn, h->envi, h->envj, h->stats, h->tab, h->own2, h->zero_page, h->pos0, h->done_list, h->done_count, h->amask, h->agrad, h->aw, h->acent,
                  h->d_actions, h->d_obs, h->d_small, h->d_termobs, h->d_mask, h->d_cta_done};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (h->h_small) cudaFreeHost(h->h_small);
  if (h->h_count) cudaFreeHost(h->h_count);
  free(h);
  return CHS_OK;
}

extern "C" int chs_get_dims(const ChsHandle* h, ChsDims* out) {
  if (!h || !out) return fail(CHS_ERR_INVALID, "null argument");
  out->num_envs = h->E; out->max_drones = h->kp.D; out->num_cattle = h->kp.M;
  out->action_width = h->kp.A; out->obs_rows = h->kp.R; out->obs_len = h->kp.L;
  out->substeps = h->kp.K; out->action_buffer_size = (int)(0.15 * h->kp.ctrl_freq);
  out->reward_width = reward_width(h); out->flag_width = flag_width(h);
  out->envs_per_block = h->kp.gact; out->threads_per_block = h->threads; out->smem_bytes = h->smem;
  return CHS_OK;
}

extern "C" int chs_set_spawn_table(ChsHandle* h, const double* xy, int32_t n_sims, int32_t cows_per_sim) {
  if (!h || !xy) return fail(CHS_ERR_INVALID, "null argument");
  if (n_sims < 1 || cows_per_sim < h->kp.M)
    return fail(CHS_ERR_INVALID, "spawn table needs >= 1 simulation and >= num_cattle (%d) cows per simulation, got %d x %d",
                h->kp.M, n_sims, cows_per_sim);   // the reference raises IndexError here
  DeviceGuard dg(h->device);
  CUDA_TRY(cudaDeviceSynchronize());
  if (h->spawn) { cudaFree(h->spawn); h->spawn = nullptr; }
  const size_t n = (size_t)n_sims * cows_per_sim;
  CUDA_TRY(cudaMalloc(&h->spawn, n * 2 * h->tsz));
  if (h->f64) {
    CUDA_TRY(cudaMemcpy(h->spawn, xy, n * 2 * sizeof(double), cudaMemcpyHostToDevice));
  } else {
    std::vector<float> f(n * 2);
    for (size_t i = 0; i < n * 2; ++i) f[i] = (float)xy[i];
    CUDA_TRY(cudaMemcpy(h->spawn, f.data(), n * 2 * sizeof(float), cudaMemcpyHostToDevice));
  }
  h->kp.n_sims = n_sims; h->kp.spawn_stride = cows_per_sim;
  // spawn row cursor keyed by the GLOBAL env id so that results do not depend on the sharding
  std::vector<int4> ei((size_t)h->E);
  CUDA_TRY(cudaMemcpy(ei.data(), h->envi, ei.size() * sizeof(int4), cudaMemcpyDeviceToHost));
  for (int e = 0; e < h->E; ++e) ei[e].w = (int)(((long long)h->kp.env_id_offset + e) % n_sims);
  CUDA_TRY(cudaMemcpy(h->envi, ei.data(), ei.size() * sizeof(int4), cudaMemcpyHostToDevice));
  h->spawn_set = true;
  // who is within the alpha range of whom is a property of the spawn row (cattle do not move on the DYN path)
  if (h->amask) { cudaFree(h->amask); h->amask = nullptr; }
  if (h->agrad) { cudaFree(h->agrad); h->agrad = nullptr; }
  if (h->aw) { cudaFree(h->aw); h->aw = nullptr; }
  if (h->acent) { cudaFree(h->acent); h->acent = nullptr; }
  h->amask_ok = false;                 // until the next full reset puts every herd on its row of the new table
  const size_t words = (size_t)n_sims * h->kp.M * ((h->kp.M + 31) / 32);
  if (h->kp.M > 0 && !getenv("CHS_NO_ALPHA_TABLE") && words * sizeof(unsigned) <= (64u << 20)) {
    CUDA_TRY(cudaMalloc(&h->amask, words * sizeof(unsigned)));
    const unsigned grid = (unsigned)(((size_t)n_sims * h->kp.M + 255) / 256);
    if (h->f64) alpha_mask_kernel<double><<<grid, 256>>>((const double2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->amask);
    else        alpha_mask_kernel<float><<<grid, 256>>>((const float2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->amask);
    CUDA_TRY(cudaDeviceSynchronize());
    // ... and so is everything in the alpha term that depends on positions only: the gradient sum and the consensus weight
    // of every in-range neighbour (chs_flock.cuh: alpha_term_tab).  First pass sizes the weight rows, second pass fills them.
    if (!getenv("CHS_NO_ALPHA_WEIGHTS")) {
      CUDA_TRY(cudaMalloc(&h->acent, (size_t)n_sims * sizeof(double2)));
      if (h->f64) spawn_centroid_kernel<double><<<(n_sims + 127) / 128, 128>>>((const double2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->acent);
      else        spawn_centroid_kernel<float><<<(n_sims + 127) / 128, 128>>>((const float2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->acent);
      int* dmax = nullptr;
      int hmax = 0;
      CUDA_TRY(cudaMalloc(&dmax, sizeof(int)));
      CUDA_TRY(cudaMemset(dmax, 0, sizeof(int)));
      if (h->f64) alpha_weight_kernel<double><<<grid, 256>>>((const double2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->amask, nullptr, nullptr, 0, dmax);
      else        alpha_weight_kernel<float><<<grid, 256>>>((const float2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->amask, nullptr, nullptr, 0, dmax);
      CUDA_TRY(cudaMemcpy(&hmax, dmax, sizeof(int), cudaMemcpyDeviceToHost));
      cudaFree(dmax);
      const int ws = hmax > 0 ? hmax : 1;
      const size_t rows = (size_t)n_sims * h->kp.M;
      if (rows * ws * h->tsz <= (64u << 20)) {
        CUDA_TRY(cudaMalloc(&h->agrad, rows * 2 * h->tsz));
        CUDA_TRY(cudaMalloc(&h->aw, rows * ws * h->tsz));
        CUDA_TRY(cudaMemset(h->aw, 0, rows * ws * h->tsz));
        h->aw_stride = ws;
        if (h->f64) alpha_weight_kernel<double><<<grid, 256>>>((const double2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->amask, (double2*)h->agrad, (double*)h->aw, ws, nullptr);
        else        alpha_weight_kernel<float><<<grid, 256>>>((const float2*)h->spawn, n_sims, cows_per_sim, h->kp.M, h->amask, (float2*)h->agrad, (float*)h->aw, ws, nullptr);
        CUDA_TRY(cudaDeviceSynchronize());
      }
    }
  }
  return CHS_OK;
}

extern "C" int chs_set_register_variant(ChsHandle* h, int32_t variant) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  if (variant < 0 || variant > 2) return fail(CHS_ERR_INVALID, "variant must be 0 (auto), 1 (latency) or 2 (occupancy)");
  if (variant == 2 && h->threads > 256) return fail(CHS_ERR_INVALID, "the occupancy variant needs threads_per_block <= 256");
  h->reg_variant = variant;
  return CHS_OK;
}

extern "C" int chs_set_step_mode(ChsHandle* h, int32_t mode) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  if (mode < 0 || mode > 2) return fail(CHS_ERR_INVALID, "mode must be 0 (auto), 1 (fused) or 2 (split)");
  h->step_mode = mode;
  return CHS_OK;
}

extern "C" int chs_set_phase_clock(ChsHandle* h, int64_t* clock_dev) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  h->dbg_clock = (long long*)clock_dev;
  return CHS_OK;
}

extern "C" int chs_get_num_drones(ChsHandle* h, int32_t* n_drones_dev, void* stream) {
  if (!h || !n_drones_dev) return fail(CHS_ERR_INVALID, "null argument");
  DeviceGuard dg(h->device);
  // envi[e] = (step_counter, step_counter_A, n_drones, spawn_idx): column 2 of an [E][4] int32 matrix
  CUDA_TRY(cudaMemcpy2DAsync(n_drones_dev, sizeof(int32_t), reinterpret_cast<const char*>(h->envi) + 2 * sizeof(int32_t), sizeof(int4),
                             sizeof(int32_t), (size_t)h->E, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return CHS_OK;
}

extern "C" int chs_set_debug_buffers(ChsHandle* h, int32_t* neighbors_dev, int32_t* herded_dev) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  h->dbg_nb = neighbors_dev; h->dbg_herd = herded_dev;
  return CHS_OK;
}

template <typename T> static KArgs<T> base_args(ChsHandle* h) {
  KArgs<T> a;
  memset(&a, 0, sizeof(a));
  a.p = h->kp;
  a.drone = (typename VT<T>::v4*)h->drone; a.rpm = (typename VT<T>::v4*)h->rpm;
  a.cpos = (typename VT<T>::v2*)h->cpos; a.cvel = (typename VT<T>::v2*)h->cvel;
  a.tm_cpos = h->tm_cpos; a.tm_cvel = h->tm_cvel; a.zero_page = h->zero_page;
  a.envi = h->envi; a.envj = h->envj; a.envf = (double2*)h->envf; a.epret = (double*)h->epret; a.path = (T*)h->path;
  a.spawn = (const typename VT<T>::v2*)h->spawn; a.stats = h->stats; a.tab = h->tab;
  a.amask = (h->amask_ok && !h->kp.cattle_integrate) ? h->amask : nullptr;
  if (a.amask && h->acent) a.acent = h->acent;
  if (a.amask && h->aw) { a.agrad = (const typename VT<T>::v2*)h->agrad; a.aw = (const T*)h->aw; a.aw_stride = h->aw_stride; }
  a.own2 = (typename VT<T>::v4*)h->own2; a.pos0 = (const typename VT<T>::v4*)h->pos0;
  a.dbg_neighbors = h->dbg_nb; a.dbg_herded = h->dbg_herd; a.dbg_clock = h->dbg_clock;
  return a;
}

static bool aligned16(const void* p) { return (((uintptr_t)p) & 15) == 0; }

template <typename T>
static int launch_reset(ChsHandle* h, const uint8_t* mask_dev, float* obs_dev, cudaStream_t st) {
  KArgs<T> a = base_args<T>(h);
  a.reset_mask = mask_dev; a.obs = obs_dev;
  const int ngroups = (h->E + h->kp.G - 1) / h->kp.G;
  const int grid = ngroups < 148 * 8 ? ngroups : 148 * 8;
  reset_kernel<T><<<grid, h->threads, h->smem, st>>>(a);
  CUDA_TRY(cudaGetLastError());
  if (!mask_dev && h->amask) h->amask_ok = true;     // every herd now sits on its spawn row (stream-ordered before later steps)
  return CHS_OK;
}

extern "C" int chs_reset(ChsHandle* h, const uint8_t* env_mask_dev, float* obs_dev, void* stream) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  if (!h->spawn_set) return fail(CHS_ERR_STATE, "chs_set_spawn_table must be called before chs_reset");
  if (obs_dev && h->kp.vec4_ok && !aligned16(obs_dev)) return fail(CHS_ERR_INVALID, "obs_dev must be 16-byte aligned");
  DeviceGuard dg(h->device);
  nvtx_push("chs_reset");
  const int rc = h->f64 ? launch_reset<double>(h, env_mask_dev, obs_dev, (cudaStream_t)stream)
                        : launch_reset<float>(h, env_mask_dev, obs_dev, (cudaStream_t)stream);
  nvtx_pop();
  return rc;
}

// Launch of one kernel of the step path with programmatic dependent launch allowed (chs_kernel.cuh: pdl_wait): the launch
// latency and the kernel's prologue overlap the tail of whatever kernel precedes it on the stream; every kernel launched
// through here executes griddepcontrol.wait before its first global-memory access.  CHS_NO_PDL=1 launches plainly (A/B runs).
template <typename K, typename A>
static void launch_pdl(const ChsHandle* h, K kernel, dim3 grid, unsigned threads, size_t smem, cudaStream_t st, const A& a) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = h->knob_no_pdl ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kernel, a);
}

template <typename T> static void launch_step_variant(ChsHandle* h, const KArgs<T>& a, int grid, cudaStream_t st);
#define CHS_LAUNCH(TT, MT, MB, FL)                                                                                 \
  do {                                                                                                               \
    if (marl) { if (split) launch_pdl(h, step_kernel<TT, MT, MB, true, true, FL>, dim3(grid), h->threads, h->smem, st, a);  \
                else launch_pdl(h, step_kernel<TT, MT, MB, true, false, FL>, dim3(grid), h->threads, h->smem, st, a); }     \
    else      { if (split) launch_pdl(h, step_kernel<TT, MT, MB, false, true, FL>, dim3(grid), h->threads, h->smem, st, a); \
                else launch_pdl(h, step_kernel<TT, MT, MB, false, false, FL>, dim3(grid), h->threads, h->smem, st, a); }    \
  } while (0)
template <> void launch_step_variant<double>(ChsHandle* h, const KArgs<double>& a, int grid, cudaStream_t st) {
  const bool marl = h->kp.packing == CHS_PACK_MARL, split = a.p.split != 0;
  CHS_LAUNCH(double, 512, 1, true);
}
template <> void launch_step_variant<float>(ChsHandle* h, const KArgs<float>& a, int grid, cudaStream_t st) {
  // latency variant while the grid is under ~512 threads per SM, occupancy variant beyond
  const bool big = (h->reg_variant == 2 && h->threads <= 256) || (h->reg_variant == 0 && h->threads <= 256 && (long long)grid * h->threads > 148LL * 512);
  const bool marl = h->kp.packing == CHS_PACK_MARL, split = a.p.split != 0;
  // lean build unless a runtime-checked feature is in use
  const bool full = a.p.aero_flags || a.p.cattle_integrate || a.p.wrapper_quirks || !a.p.cp_ok || !a.p.bulk_ok || h->knob_full
#ifndef CHS_PROFILE_LEAN
                    || a.dbg_clock
#endif
      ;
  if (full) { if (big) CHS_LAUNCH(float, 256, 4, true); else CHS_LAUNCH(float, 512, 1, true); }
  else      { if (big) CHS_LAUNCH(float, 256, 4, false); else CHS_LAUNCH(float, 512, 1, false); }
}
#undef CHS_LAUNCH

template <typename T>
static int launch_step(ChsHandle* h, const float* actions, float* obs, float* reward, uint8_t* term, uint8_t* trunc,
                       float* term_obs, int term_mode, int* term_count, int* term_idx, cudaStream_t st) {
  KArgs<T> a = base_args<T>(h);
  a.actions = actions; a.obs = obs; a.reward = reward; a.terminated = term; a.truncated = trunc;
  a.term_obs = term_obs; a.term_mode = term_obs ? term_mode : 0; a.term_count = term_count; a.term_idx = term_idx;
  a.term_count_next = (term_obs && term_mode == 2) ? h->count_next : nullptr;
  a.cta_done = h->pub_cta_done; a.host_count = h->pub_host_count;
  // dead-column skipping pays on the PCIe path only (fewer bytes per env); in HBM the extra, smaller bulk stores cost more
  // than the bytes they save (measured: 243 -> 294 us at 262 144 envs), so device destinations get the rows in one piece
  if (!h->obs_dst_is_host) a.p.dead_lo = a.p.dead_hi = 0;
  a.p.late_rows = (h->obs_dst_is_host && h->kp.auto_reset && obs && !h->knob_early_rows) ? 1 : 0;
  const int grid = (h->E + h->kp.gact - 1) / h->kp.gact;
  // split mode: the drones get their own barrier-free kernel; auto picks it once the fused grid would be more than about one
  // wave of resident CTAs (measured at 3 drones / 20 cattle: 16 384 envs 23.5 us fused / 25.9 split, 24 576 envs 38.3 / 36.5,
  // 49 152 envs 61.4 / 51.7)
  const bool split = h->step_mode == 2 || (h->step_mode == 0 && (long long)h->E * h->kp.D >= 148LL * 448);
  if (split) {
    if (h->pos0) CUDA_TRY(cudaMemcpyAsync(h->pos0, h->drone, (size_t)h->kp.D * h->E * 4 * h->tsz, cudaMemcpyDeviceToDevice, st));
    const long long nthr = (long long)h->kp.D * h->E;
    if (h->kp.auto_reset && h->knob_deferred_reset) a.done_count = h->done_count;
    launch_pdl(h, drone_kernel<T>, dim3((unsigned)((h->E + 255) / 256), (unsigned)h->kp.D), 256, 0, st, a);
    a.p.split = 1;
  }
  // Finished envs are rebuilt inside step_kernel in both modes (phase 5, on the warps that idle while warp 0 finishes phase
  // 3).  Split mode used to list them and rebuild them in a third kernel, which was faster while the rebuild was 2000 cold
  // instructions on the tail of a CTA; since the rebuild was trimmed the third kernel only costs (262 144 envs: 253.8 us
  // deferred, 238.9 us in place; 65 536 envs: 73.4 / 69.4).  chs_step_static_scratch keeps the deferred form.
  const bool defer = split && h->kp.auto_reset && h->knob_deferred_reset;
  if (defer) { a.done_list = h->done_list; a.done_count = h->done_count; }
  launch_step_variant<T>(h, a, grid, st);
  if (defer) {
    KArgs<T> r = base_args<T>(h);
    r.done_list = h->done_list; r.done_count = h->done_count; r.obs = obs;
    launch_pdl(h, reset_list_kernel<T>, dim3(148 * 8), 256, 0, st, r);
  }
  CUDA_TRY(cudaGetLastError());
  return CHS_OK;
}

extern "C" int chs_step(ChsHandle* h, const float* actions_dev, float* obs_dev, float* reward_dev,
                        uint8_t* terminated_dev, uint8_t* truncated_dev, float* terminal_obs_dev, void* stream) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  if (!actions_dev || !reward_dev || !terminated_dev || !truncated_dev)
    return fail(CHS_ERR_INVALID, "actions, reward, terminated and truncated pointers are required");
  if (h->kp.auto_reset && !h->spawn_set) return fail(CHS_ERR_STATE, "auto_reset needs a spawn table");
  if (h->kp.act4_ok && !aligned16(actions_dev)) return fail(CHS_ERR_INVALID, "actions_dev must be 16-byte aligned");
  if (h->kp.vec4_ok && ((obs_dev && !aligned16(obs_dev)) || (terminal_obs_dev && !aligned16(terminal_obs_dev))))
    return fail(CHS_ERR_INVALID, "obs buffers must be 16-byte aligned");
  DeviceGuard dg(h->device);
  nvtx_push("chs_step");
  const int rc = h->f64 ? launch_step<double>(h, actions_dev, obs_dev, reward_dev, terminated_dev, truncated_dev,
                                              terminal_obs_dev, 1, nullptr, nullptr, (cudaStream_t)stream)
                        : launch_step<float>(h, actions_dev, obs_dev, reward_dev, terminated_dev, truncated_dev,
                                             terminal_obs_dev, 1, nullptr, nullptr, (cudaStream_t)stream);
  nvtx_pop();
  return rc;
}

// ---- chs_step_static_scratch ------------------------------------------------------------------------------------------
// The split-mode control step (drone_kernel -> step_kernel<SPLIT> -> reset_list_kernel, the same kernels chs_step launches in
// step mode 2) with its scratch -- the drone-to-task hand-over rows and the finished-env list -- in g_static_scratch instead of
// the handle's cudaMalloc'ed buffers.  Calls no cudaMalloc / cudaFree / allocating routine: kernel launches, one
// cudaMemsetAsync of the 256-byte counter block and (downwash only) one device-to-device copy into a buffer chs_create made.
template <typename T>
static int launch_step_static(ChsHandle* h, unsigned char* scratch, const float* actions, float* obs, float* reward, uint8_t* term,
                              uint8_t* trunc, float* term_obs, cudaStream_t st) {
  const size_t E = (size_t)h->E, D = (size_t)h->kp.D;
  const size_t list_off = 256, list_bytes = (E * sizeof(int) + 255) & ~(size_t)255;
  const size_t own_off = list_off + list_bytes, own_bytes = D * E * 4 * sizeof(T);
  if (own_off + own_bytes > (size_t)CHS_STATIC_SCRATCH_BYTES)
    return fail(CHS_ERR_INVALID, "chs_step_static_scratch: %zu envs x %zu drone slots need %zu bytes of scratch, the static array "
                "holds CHS_STATIC_SCRATCH_BYTES = %d; use chs_step for larger batches", E, D, own_off + own_bytes, CHS_STATIC_SCRATCH_BYTES);
  KArgs<T> a = base_args<T>(h);
  a.actions = actions; a.obs = obs; a.reward = reward; a.terminated = term; a.truncated = trunc;
  a.term_obs = term_obs; a.term_mode = term_obs ? 1 : 0;
  a.own2 = reinterpret_cast<typename VT<T>::v4*>(scratch + own_off);
  int* count = reinterpret_cast<int*>(scratch);
  int* list = reinterpret_cast<int*>(scratch + list_off);
  // the array keeps the previous call's contents: the counter has to start at zero (the hand-over rows and the list are
  // written before they are read, for exactly the entries that are read)
  CUDA_TRY(cudaMemsetAsync(scratch, 0, 256, st));
  if (h->pos0) CUDA_TRY(cudaMemcpyAsync(h->pos0, h->drone, D * E * 4 * h->tsz, cudaMemcpyDeviceToDevice, st));
  const bool defer = h->kp.auto_reset != 0;
  if (defer) a.done_count = count;
  const long long nthr = (long long)D * (long long)E;
  launch_pdl(h, drone_kernel<T>, dim3((unsigned)((h->E + 255) / 256), (unsigned)h->kp.D), 256, 0, st, a);
  a.p.split = 1;
  if (defer) { a.done_list = list; a.done_count = count; }
  const int grid = (h->E + h->kp.gact - 1) / h->kp.gact;
  launch_step_variant<T>(h, a, grid, st);
  if (defer) {
    KArgs<T> r = base_args<T>(h);
    r.done_list = list; r.done_count = count; r.obs = obs;
    launch_pdl(h, reset_list_kernel<T>, dim3(148 * 8), 256, 0, st, r);
  }
  CUDA_TRY(cudaGetLastError());
  return CHS_OK;
}

extern "C" int chs_step_static_scratch(ChsHandle* h, const float* actions_dev, float* obs_dev, float* reward_dev,
                                       uint8_t* terminated_dev, uint8_t* truncated_dev, float* terminal_obs_dev, void* stream) {
  if (!h) return fail(CHS_ERR_INVALID, "null handle");
  if (!actions_dev || !reward_dev || !terminated_dev || !truncated_dev)
    return fail(CHS_ERR_INVALID, "actions, reward, terminated and truncated pointers are required");
  if (h->kp.auto_reset && !h->spawn_set) return fail(CHS_ERR_STATE, "auto_reset needs a spawn table");
  if (h->kp.act4_ok && !aligned16(actions_dev)) return fail(CHS_ERR_INVALID, "actions_dev must be 16-byte aligned");
  if (h->kp.vec4_ok && ((obs_dev && !aligned16(obs_dev)) || (terminal_obs_dev && !aligned16(terminal_obs_dev))))
    return fail(CHS_ERR_INVALID, "obs buffers must be 16-byte aligned");
  if (h->device < 0 || h->device >= kMaxDevices || !g_static_scratch_addr[h->device])
    return fail(CHS_ERR_STATE, "static scratch address of device %d was not resolved at chs_create", h->device);
  DeviceGuard dg(h->device);
  unsigned char* scratch = static_cast<unsigned char*>(g_static_scratch_addr[h->device]);   // the looked-up device address, never the symbol
  return h->f64 ? launch_step_static<double>(h, scratch, actions_dev, obs_dev, reward_dev, terminated_dev, truncated_dev,
                                             terminal_obs_dev, (cudaStream_t)stream)
                : launch_step_static<float>(h, scratch, actions_dev, obs_dev, reward_dev, terminated_dev, truncated_dev,
                                            terminal_obs_dev, (cudaStream_t)stream);
}

#include "chs_api_io.inc"
