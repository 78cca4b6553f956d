// HBM state layout and the kernel argument block.
//
// All persistent state is structure-of-arrays with the ENV INDEX FASTEST, so that a warp whose lanes
// are consecutive environments moves 512 B (float4) / 256 B (float2) fully coalesced lines:
//   drone   v4 [7][D][E]   A: pos.xyz | roll      B: quat.xyzw       C: vel.xyz | pitch
//                          D: rpy_rates.xyz | yaw E: pid_ipos.xyz | ang_v.x
//                          F: pid_last_rpy.xyz | ang_v.y             G: pid_irpy.xyz | ang_v.z
//                          (exactly the 28 per-drone floats the reference keeps: BaseAviary.py:704-721
//                           kinematics + DSLPIDControl.py:65-78 controller state)
//   rpm     v4 [D][E]      last_clipped_action (BaseAviary.py:450), write-only on the DYN path
//   cpos    v2 [M][E]      cattle xy (z is the constant 0.1; cattle never move in DYN)
//   cvel    v2 [M][E]      cattle xy velocity (z = 0)
//   envi  int4 [E]         step_counter, step_counter_A, n_drones, spawn_idx
//   envj  int4 [E]         level, tally, episode, active-agent bit mask of the RLlib wrapper (marl_wrapper_quirks only)
//   envf double2 [E]       prev_cent_dist (NaN = None), spacing_clock -- float64 in BOTH builds: the approach
//                          reward multiplies a difference of two ~10 m centroid distances by 5*ctrl_freq/2.5
//   epret double [E]       running episode return (telemetry)
//   path     T [D][E]      1.7 x xy path length of the running episode per drone (update_evaluation_metrics,
//                          BaseAviary.py:1406-1435); negative = no step since the reset (the first step of an episode is
//                          measured from the origin: last_drones_pos is zeroed by reset, :317)
#pragma once
#include <cuda.h>          // CUtensorMap (the type only: the encoder is fetched through cudaGetDriverEntryPoint)
#include "chs_math.cuh"
#include "../../include/chs.h"

namespace chs {

constexpr int CHS_STAT_ROWS = 1024;   // rows of the statistics accumulator (a power of two); chs_stats sums them

enum DroneField { F_A = 0, F_B, F_C, F_D, F_E, F_F, F_G, F_COUNT };

// shared-memory carve-up of one CTA (byte offsets; smem_layout below), computed by the host for the launch shape in use
struct SmemLayout {
  int off_dpos, off_cpos, off_cvel, off_u, off_redA, off_redB, off_stage, off_ccent, off_ctab, off_zero, off_int, total;
};

struct KParams {
  int E, D, M, K, A, R, L;
  int act, packing, auto_reset;
  int G;                 // envs per CTA (1..32, power of two)
  int n_sims, spawn_stride;
  int fixed_n, min_n, max_n;
  int trunc_step_limit;  // truncated iff step_counter > this  (step_counter/CTRL_FREQ > EPISODE_LEN_SEC)
  int ctrl_freq;
  int env_id_offset;
  int stage_stride;      // floats per env in the obs staging area (>= D*L, multiple of 4, = 4 mod 8 so that the 32 envs of a
                         // CTA spread over 8 banks instead of 2 or 4: lanes of a warp are consecutive envs)
  int stage_copy;        // floats of an env's staged rows that leave for the obs image: D*L rounded up to 4
  int vec4_ok;           // (R*L) % 4 == 0 -> float4 copy-out
  int cattle_integrate;  // non-reference: herd positions advance by vel * CTRL_TIMESTEP every control step
  int aero_flags;        // bit0 drag, bit1 ground effect, bit2 downwash (chs_drone.cuh: dyn_substeps_aero)
  int cp_ok;             // herd rows can be staged with 16-byte cp.async (alignment of row segments)
  int act4_ok;           // A == 4: one float4 load per action row
  int split;             // 1: drones are integrated by drone_kernel, step_kernel does the rest
  int bulk_ok;           // observations leave through bulk (TMA) stores: R*L*4 and the staging stride are multiples of 16 B
  int zero_bytes;        // size of the shared zero block = R*L*4 - min(stage_copy*4, R*L*4)
  int pad_persist;       // rows >= D of an obs image are caller-owned zeros: chs_step does not write them (ChsParams.persistent_padding)
  int dead_lo, dead_hi;  // columns [dead_lo, dead_hi) of every live obs row are caller-owned constant zeros (the unused neighbour
                         // slots, BaseRLAviary.py:313-314): not rewritten by chs_step when persistent_padding is on; 0, 0 = none
  int wrapper_quirks;    // MARL packing: outputs are those of the reference RLlib wrapper's second invocation (ChsParams.marl_wrapper_quirks)
  int late_rows;         // observations go to pinned host memory: live rows are stored after phase 3, finished envs skipped
  int gact;              // envs a CTA of the step kernel handles (<= G, the power-of-two shared-memory stride): wave fitting
  int tma_ok;            // the herd tile of a CTA is fetched by two 2-D TMA loads (KArgs::tm_cpos / tm_cvel) instead of cp.async chunks
  int early_ok;          // warp 0 may start its drone chain before the other warps finish staging (CHS_NO_EARLY=1 clears it)
  uint32_t seed_lo, seed_hi;
  SmemLayout lay;        // = smem_layout<T>(D, M, G, stage_stride, zero_bytes): every thread of every CTA would recompute it otherwise
};

template <typename T> struct KArgs {
  KParams p;
  typename VT<T>::v4* drone;
  typename VT<T>::v4* rpm;
  typename VT<T>::v2* cpos;
  typename VT<T>::v2* cvel;
  typename VT<T>::v4* own2;           // split mode scratch [D][E]: (ang_v, yaw) of the step
  const typename VT<T>::v4* pos0;     // split mode + downwash: start-of-step copy of drone field A [D][E]
  int4* envi;
  int4* envj;
  double2* envf;
  double* epret;
  T* path;                           // [D][E] running 1.7 x path length per drone (BaseAviary.py:1424), < 0 = fresh episode
  const typename VT<T>::v2* spawn;   // [n_sims][spawn_stride]
  const unsigned* amask;             // nullable [n_sims][M][ceil(M/32)]: alpha-term in-range masks per spawn row (chs_flock.cuh)
  // nullable, valid with amask: per (spawn row, cow) the alpha term's position-only part -- .x/.y = sum of the gradient pieces,
  // and aw[(row * M + cow) * aw_stride + k] = consensus weight (the bump value) of its k-th in-range neighbour in mask order
  const typename VT<T>::v2* agrad;
  const T* aw;
  int aw_stride;
  const double2* acent;              // nullable, valid with amask: herd centroid of every spawn row (float64, the step kernel's summation order)
  double* stats;                     // [CHS_STAT_ROWS][CHS_NUM_STATS], row = CTA index mod CHS_STAT_ROWS
  const double* tab;                 // [CHS_TAB_DOUBLES] curriculum rows + 1/n (chs_task.cuh)
  const float* zero_page;            // 32 * stage_stride zero floats: source of the bulk copy that clears a CTA's obs staging area (occupancy build)
  // per-call
  const float* actions;              // [E][D][A]
  float* obs;                        // [E][R][L]
  float* reward;                     // [E] (CTDE) / [E][D] (MARL)
  uint8_t* terminated;
  uint8_t* truncated;
  float* term_obs;                   // nullable: full [E][R][L] (term_mode 1) or compact rows (mode 2)
  int term_mode;
  int* term_count;                   // mode 2: atomic row counter
  int* term_idx;                     // mode 2: env index of each compact row
  int* term_count_next;              // mode 2, nullable: the counter the NEXT call will use; zeroed by this launch (ping-pong)
  int* cta_done;                     // host path, nullable: ticket counter of the CTAs that finished phase 3 (device memory, 0 between launches)
  int* host_count;                   // host path, nullable: pinned host word that receives *term_count from the last CTA
  const uint8_t* reset_mask;         // reset kernel only (nullable = all)
  int* done_list;                    // deferred reset (split mode): env ids finished this step (nullable = reset in the kernel)
  int* done_count;                   // number of entries, zeroed by drone_kernel
  int* dbg_neighbors;                // nullable [E][D][2]
  int* dbg_herded;                   // nullable [E]
  long long* dbg_clock;              // nullable [grid][32]: clock64() of thread 0 at phase and sub-phase boundaries
  // fp32 build: tensor maps over cpos / cvel seen as [M][2E] floats, box = [M][2G]: one TMA load brings the herd positions (or
  // velocities) of the G envs of a CTA into the [cow][G] shared-memory tile, zero-filled past the end of the batch
  alignas(64) CUtensorMap tm_cpos;
  alignas(64) CUtensorMap tm_cvel;
};

// ---- shared-memory carve-up (per CTA of G envs) ----------------------------------------------------

template <typename T> __host__ __device__ inline SmemLayout smem_layout(int D, int M, int G, int stage_stride, int zero_bytes) {
  SmemLayout s;
  int o = 0;
  const int v4 = 4 * (int)sizeof(T), v2 = 2 * (int)sizeof(T);
  s.off_dpos = o; o += D * G * v4;
  s.off_redA = o; o += D * G * v4;
  s.off_redB = o; o += D * G * v4;
  o = (o + 127) & ~127;                 // TMA destinations: 128-byte aligned
  s.off_cpos = o; o += M * G * v2;
  o = (o + 127) & ~127;
  s.off_cvel = o; o += M * G * v2;
  s.off_u = o;    o += M * G * v2;
  o = (o + 15) & ~15;
  s.off_stage = o; o += G * stage_stride * 4;
  s.off_ccent = o; o += G * 16;
  s.off_ctab = o; o += 152 * 8;         // CHS_TAB_DOUBLES (149) rounded up
  s.off_zero = o; o += zero_bytes;      // multiple of 16
  s.off_int = o;  o += (10 * 32 + 8) * 4;  // ten per-env int arrays [32] + misc
  s.total = o;
  return s;
}

enum IntSlot { I_N = 0, I_HERD = 1, I_DONE = 2, I_NEWN = 3, I_ROW = 4, I_FLOCK = 5, I_TERMROW = 6, I_DLIST = 7, I_EPI = 8, I_LEVEL = 9, I_COUNT = 10 };

}  // namespace chs
