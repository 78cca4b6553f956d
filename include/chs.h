/*
 * chs.h -- C ABI of the B200-native batched simulator for the RL-Cattle-Herding env step.
 *
 * The reference (BenCooper305/RL-Cattle-Herding) has no FFI layer: its hot path is the Python
 * method `CattleAviary.step()` / `reset()` (gym_pybullet_drones/sb3_envs/BaseAviary.py:335-465,
 * :280-331) running one environment per process.  This library replaces that path for E
 * independent environments at once; every entry point cites the reference interface it stands in
 * for.  Plain pointers and sizes only (no torch types).  Conventions:
 *   - every function returns 0 on success or a negative ChsStatus; the message is available
 *     from chs_last_error() (thread local);
 *   - `*_dev` pointers are CUDA device pointers owned by the caller, `*_host` are host pointers;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); chs_step and
 *     chs_reset are asynchronous on it and perform no host synchronisation and no allocation;
 *   - a handle is bound to one device and is not thread safe (one handle per process per GPU).
 *
 * Tensor layouts seen by the caller (row-major, C order):
 *   actions     float32 [E][D][A]   D = max_drones slots, A = action width of `act_type`
 *                                   (rows >= the env's current drone count are ignored,
 *                                   BaseRLAviary.py:166)
 *   obs         float32 [E][R][L]   R = obs_rows (12 = reference padding, BaseRLAviary.py:334),
 *                                   L = 10 + 8 + 32 + ABS*A (BaseRLAviary.py:262)
 *   reward      float32 [E]         (CTDE packing, CattleAviary.py:332)
 *   terminated  uint8   [E]         (CattleAviary.py:422)
 *   truncated   uint8   [E]         (CattleAviary.py:497)
 * MARL packing (rllib_envs/BaseAviary.py:425-430, the per-agent dicts as arrays; slots >= the env's
 * drone count are zero):
 *   obs         float32 [E][R][L]   row i = _computeObs(i) (rllib_envs/BaseMARLAviary.py:253-303), R = D
 *   reward      float32 [E][D]      _computeReward(i)      (rllib_envs/MARLCattleAviary.py:110)
 *   terminated  uint8   [E][D+1]    _computeTerminated(i), column D = "__all__"   (:246)
 *   truncated   uint8   [E][D+1]    _computeTruncated(i),  column D = "__all__"   (:326)
 *   auto_reset resets an env when either "__all__" flag is set (what RLlib ends an episode on).
 */
#ifndef CHS_H_
#define CHS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CHS_ABI_VERSION 4   /* 4: ChsStateView.active_agents / path_len, ChsParams.marl_wrapper_quirks, chs_stats slots 11-15 */
#define CHS_MAX_DRONES 12   /* GLOBAL_MAX_NUM_DRONES, BaseAviary.py:112 */
#define CHS_NUM_STATS 16
#define CHS_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)   /* size of the static device array behind chs_step_static_scratch */

typedef enum ChsStatus {
  CHS_OK = 0,
  CHS_ERR_INVALID = -1,   /* bad argument / unsupported configuration */
  CHS_ERR_CUDA = -2,      /* a CUDA runtime call failed */
  CHS_ERR_STATE = -3,     /* call order (e.g. step before a spawn table was set) */
  CHS_ERR_NOMEM = -4
} ChsStatus;

/* utils/enums.py:26-35 (ActionType); values are this ABI's, names the reference's */
typedef enum ChsAct {
  CHS_ACT_VEL = 0,        /* (N,4): xy direction, unused z, speed fraction  BaseRLAviary.py:185-222 */
  CHS_ACT_RPM = 1,        /* (N,4): HOVER_RPM*(1+0.05a)                    BaseRLAviary.py:168-169 */
  CHS_ACT_PID = 2,        /* (N,3): waypoint                                BaseRLAviary.py:170-184 */
  CHS_ACT_ONE_D_RPM = 3,  /* (N,1)                                          BaseRLAviary.py:223-224 */
  CHS_ACT_ONE_D_PID = 4   /* (N,1)                                          BaseRLAviary.py:225-234 */
} ChsAct;

typedef enum ChsPacking {
  CHS_PACK_CTDE = 0,      /* sb3_envs: one reward / flag pair per env */
  CHS_PACK_MARL = 1       /* rllib_envs: per-drone reward / flags (rllib_envs/MARLCattleAviary.py) */
} ChsPacking;

/* Constructor arguments: the kwargs of CattleAviary.__init__ (sb3_envs/CattleAviary.py:14-28) that
 * matter on the DYN path, plus batching/sharding knobs.  Zero-initialise, then set struct_size. */
typedef struct ChsParams {
  int32_t struct_size;        /* sizeof(ChsParams) */
  int32_t num_cattle;         /* num_cattle kwarg (M) */
  int32_t max_drones;         /* drone slots per env, 1..12 (num_drones kwarg sizes the action space) */
  int32_t fixed_num_drones;   /* >0: every episode has this many drones; 0: drawn per episode in
                                 [MIN,MAX] of the start level like BaseAviary.py:307 */
  int32_t pyb_freq;           /* 240 */
  int32_t ctrl_freq;          /* 30 (BASELINE) or 60 (reference default) */
  int32_t act_type;           /* ChsAct */
  int32_t start_level;        /* curriculum level 0..7 (CattleAviary.py:62 hard-codes 7) */
  int32_t obs_rows;           /* R: 12 (reference) or max_drones (compact); 0 -> 12 */
  int32_t packing;            /* ChsPacking */
  int32_t precision;          /* 0 = fp32 (product), 1 = fp64 (parity instrument) */
  int32_t auto_reset;         /* 1: finished envs are reset inside chs_step (VecEnv semantics) */
  int32_t aero_flags;         /* bit0 drag, bit1 ground effect, bit2 downwash (BaseAviary.py:943-1039;
                                 PYB-only in the reference, here forces of the DYN step; 0 = reference DYN) */
  int32_t cattle_integrate;   /* non-reference extension: herd pos += vel*CTRL_TIMESTEP every control step */
  int32_t env_id_offset;      /* global id of local env 0 (multi-GPU sharding by env index) */
  int32_t envs_per_block;     /* launch tuning, 0 = auto */
  int32_t threads_per_block;  /* launch tuning, 0 = auto */
  int32_t reserved0;          /* step mode at creation: 0 auto, 1 fused, 2 split (see chs_set_step_mode) */
  int32_t persistent_padding; /* 1: rows >= max_drones of every (R, L) observation image are constant zeros OWNED BY THE
                                 CALLER: chs_reset writes them for the envs it resets, chs_step never touches them (3/4 of
                                 the bytes of a (12, 66) image at 3 drones).  Buffers handed to chs_step -- obs_dev and
                                 terminal_obs_dev -- must hold zeros there (zero-fill them once, or chs_reset into them).
                                 Granularity is 16 bytes: the store of the live rows may rewrite (with zeros) up to 12 bytes
                                 at the start of row max_drones.  On the pinned-host path (chs_step_host / chs_reset_host)
                                 the same holds for the dead columns of the live rows in the sb3 packing: neighbour slots
                                 2..11 (columns 14..33, BaseRLAviary.py:313-314 pads 2 active neighbours to 12 slots) are
                                 zeros that chs_reset_host writes and chs_step_host skips (fewer bytes across PCIe).
                                 0: every call rewrites the full image (the reference's _computeObs does, BaseRLAviary.py:334) */
  int32_t marl_wrapper_quirks;/* MARL packing only, default 0.  1: reward / terminated / truncated are what the reference's RLlib wrapper
                                 hands to RLlib (rllib_envs/marl_wrapper.py:104-117), not what env.step returned: the wrapper calls the
                                 stateful per-agent functions a SECOND time for every agent still in its `agents` list (approach term
                                 0, two more spacing-clock ticks and possibly another tally bump per agent), agents whose second-pass
                                 `terminated` is set leave the list, and both "__all__" columns mean "the list is empty".  Slots of
                                 agents that left the list read 0.  The list is part of the state (ChsStateView.active_agents) */
  uint64_t seed;              /* Philox key; streams are keyed by (seed, global env id, episode) */
} ChsParams;

typedef struct ChsHandle ChsHandle;

/* Derived sizes, so that callers can allocate tensors (BaseRLAviary._observationSpace :243-267,
 * _actionSpace :106-133). */
typedef struct ChsDims {
  int32_t num_envs, max_drones, num_cattle;
  int32_t action_width;       /* A */
  int32_t obs_rows, obs_len;  /* R, L */
  int32_t substeps;           /* K = pyb_freq / ctrl_freq */
  int32_t action_buffer_size; /* ABS = int(0.15*ctrl_freq), BaseRLAviary.py:69 */
  int32_t reward_width;       /* 1 (CTDE) or max_drones (MARL) */
  int32_t envs_per_block, threads_per_block, smem_bytes;
  int32_t flag_width;         /* 1 (CTDE) or max_drones + 1 (MARL: one flag per drone, then "__all__") */
} ChsDims;

/* Host-side, float64, env-major view of the complete simulator state: the fields
 * `oracle/ref_harness/ref_env.capture_state` reads off the reference env.  NULL members are skipped.
 * Used for teacher-forced parity tests and checkpoint/resume. */
typedef struct ChsStateView {
  double* pos;            /* [n][D][3]  BaseAviary.pos          */
  double* quat;           /* [n][D][4]  BaseAviary.quat (xyzw)  */
  double* rpy;            /* [n][D][3]  BaseAviary.rpy          */
  double* vel;            /* [n][D][3]  BaseAviary.vel          */
  double* ang_v;          /* [n][D][3]  BaseAviary.ang_v        */
  double* rpy_rates;      /* [n][D][3]  BaseAviary.rpy_rates    */
  double* last_rpm;       /* [n][D][4]  BaseAviary.last_clipped_action */
  double* pid_ipos;       /* [n][D][3]  DSLPIDControl.integral_pos_e */
  double* pid_last_rpy;   /* [n][D][3]  DSLPIDControl.last_rpy  */
  double* pid_irpy;       /* [n][D][3]  DSLPIDControl.integral_rpy_e */
  double* cattle_pos;     /* [n][M][3]  */
  double* cattle_vel;     /* [n][M][3]  */
  double* prev_cent_dist; /* [n]  CattleAviary.prev_cent_dists (NaN = None) */
  double* spacing_clock;  /* [n]  CattleAviary.drone_spacing_clock */
  double* ep_return;      /* [n]  running episode return (telemetry) */
  int64_t* step_counter;  /* [n]  BaseAviary.step_counter (physics steps) */
  int32_t* step_counter_A;/* [n]  BaseAviary.step_counter_A (control steps) */
  int32_t* n_drones;      /* [n]  BaseAviary.NUM_DRONES */
  int32_t* level;         /* [n]  CurriculumLearning.curriculum_level */
  int32_t* tally;         /* [n]  CurriculumLearning.curriculum_success_tally */
  int32_t* spawn_idx;     /* [n]  BaseAviary.Cattle_Spawn_Index */
  int32_t* episode;       /* [n]  episodes started so far (Philox counter) */
  int32_t* active_agents; /* [n]  RLlibMultiAgentWrapper.agents as a bit mask, bit i = "agent_i" (marl_wrapper.py:50,113); only
                             maintained with ChsParams.marl_wrapper_quirks, rebuilt on the first step of every episode (:68-72) */
  double* path_len;       /* [n][D]  BaseAviary.episode_drone_distances minus the spawn position it starts from (:683-688): the
                             running 1.7 x xy path length of each drone (:1418-1426).  A NEGATIVE value marks an episode that has
                             not stepped yet: its first step is measured from the origin (last_drones_pos is zeroed by reset, :317) */
} ChsStateView;

int chs_abi_version(void);
const char* chs_last_error(void);

/* CattleAviary.__init__ (sb3_envs/CattleAviary.py:14-105) for `num_envs` environments on CUDA
 * device `device`.  All state lives in HBM owned by the handle. */
int chs_create(const ChsParams* params, int32_t num_envs, int32_t device, ChsHandle** out);
int chs_destroy(ChsHandle* h);
int chs_get_dims(const ChsHandle* h, ChsDims* out);

/* config/cattle_positions.yaml as read at BaseAviary.py:88-94: `xy_host` is
 * float64 [n_sims][cows_per_sim][2]; cows_per_sim must be >= num_cattle. */
int chs_set_spawn_table(ChsHandle* h, const double* xy_host, int32_t n_sims, int32_t cows_per_sim);

/* BaseAviary.reset (:280-331) for the envs whose mask byte is non-zero (NULL = all).  Writes the
 * initial observation of those envs into obs_dev (may be NULL). */
int chs_reset(ChsHandle* h, const uint8_t* env_mask_dev, float* obs_dev, void* stream);

/* BaseAviary.step (:335-465): one control step of every env.  Step mode 1 (fused) is ONE kernel launch; step mode 2
 * (split, picked automatically for batches several waves deep) is drone_kernel + step_kernel, plus one
 * device-to-device copy when the downwash model is on (see chs_set_step_mode).
 * terminal_obs_dev (nullable): [E][R][L]; with auto_reset the pre-reset observation of each
 * finished env is written there (rows of other envs are left untouched) and obs_dev receives the
 * first observation of the new episode, as SB3's VecEnv contract requires. */
int chs_step(ChsHandle* h, const float* actions_dev, float* obs_dev, float* reward_dev,
             uint8_t* terminated_dev, uint8_t* truncated_dev, float* terminal_obs_dev, void* stream);

/* No counterpart in the reference: added for callers that may not call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same control step, same arguments and same
 * results as chs_step in split mode (here in its three-kernel form: drone_kernel, step_kernel, reset_list_kernel), but the scratch those
 * kernels hand data through -- the per-drone (ang_v, yaw) rows between the first two and the list of finished envs between
 * the last two -- lives in ONE file-scope `static __device__` array of CHS_STATIC_SCRATCH_BYTES (8 MiB) per GPU instead of
 * buffers from cudaMalloc.  Every GPU that loads the library holds that array until the process exits.  The call itself
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine): kernel launches, a 256-byte cudaMemsetAsync and,
 * with the downwash model only, one device-to-device copy.  The array's device address is looked up with
 * cudaGetSymbolAddress by chs_create, once per device, with that device current.
 * Limit: 256 + roundup256(4*E) + 16*D*E bytes (32*D*E in the fp64 build) must fit, e.g. E <= 161 000 envs at D = 3 drone
 * slots, E <= 42 700 at D = 12; larger batches are rejected with CHS_ERR_INVALID and a message naming the limit.
 * ALL handles of a process on one GPU share the array, and it keeps the previous call's contents (the call clears the one
 * counter that has to start at zero): calls that overlap on the GPU -- two streams, two handles, or this call racing a graph
 * replay of it -- corrupt each other; issue them in stream order.  Mixing with chs_step on the same handle is fine. */
int chs_step_static_scratch(ChsHandle* h, const float* actions_dev, float* obs_dev, float* reward_dev,
                            uint8_t* terminated_dev, uint8_t* truncated_dev, float* terminal_obs_dev, void* stream);

/* Same step with HOST buffers (what a numpy-facing VecEnv calls); the call returns after the stream has drained.
 * Terminal observations come back compacted: *n_done_host envs finished, their indices are in done_idx_host[0..n)
 * (capacity E) and their terminal observations in terminal_obs_host[i][R][L] (capacity E rows); all three may be NULL.
 * PINNED (page-locked, mapped) buffers are used in place: the kernel reads the actions from actions_host and stores
 * observations, terminal observations, reward, flags and done indices straight into the caller's arrays over PCIe; the
 * finished-env count arrives through a pinned word of the handle.  Nothing is staged and no copy is enqueued: one launch
 * and one stream synchronisation per call.  If any of the buffers is pageable the call takes the staged path (device staging
 * + copies; actions are copied from the caller's array) with the same results.
 * Observation rows >= max_drones are identically zero, so only the first max_drones rows of every image cross PCIe (under
 * ChsParams.persistent_padding also without the dead neighbour columns): the padding of obs_host / terminal_obs_host is
 * written by chs_reset_host (staged path: zero-filled the first time a buffer address is seen) and is not rewritten
 * afterwards -- do not scribble on it. */
int chs_step_host(ChsHandle* h, const float* actions_host, float* obs_host, float* reward_host,
                  uint8_t* terminated_host, uint8_t* truncated_host, int32_t* n_done_host,
                  int32_t* done_idx_host, float* terminal_obs_host, void* stream);
int chs_reset_host(ChsHandle* h, const uint8_t* env_mask_host, float* obs_host, void* stream);

/* State injection / extraction for envs [env0, env0+count): synchronous, host float64. */
int chs_get_state(ChsHandle* h, int32_t env0, int32_t count, ChsStateView* out_host);
int chs_set_state(ChsHandle* h, int32_t env0, int32_t count, const ChsStateView* in_host);

/* BaseAviary.NUM_DRONES of every env (BaseAviary.py:242,307: drawn per episode) as int32 [E] on the device: one
 * stream-ordered device-to-device copy, no kernel, no allocation.  A MARL learner uses it to tell drone slots that exist from
 * padding slots (rllib_envs/BaseAviary.py:425-430 builds its dicts over range(NUM_DRONES) only). */
int chs_get_num_drones(ChsHandle* h, int32_t* n_drones_dev, void* stream);

/* Optional integer diagnostics written by chs_step (bit-exact parity outputs): the indices of the two
 * nearest drones per drone as ranked by _computeObs (BaseRLAviary.py:310-311), int32 [E][D][2],
 * and the herded-cow count of evaluate_herding_effectiveness (utils/evaluation.py:100-138),
 * int32 [E].  Pass NULL to disable. */
int chs_set_debug_buffers(ChsHandle* h, int32_t* neighbors_dev, int32_t* herded_dev);

/* The fp32 step kernel is compiled with two register budgets: 1 = latency (<=128 registers/thread, for grids of
 * about one wave), 2 = occupancy (64 registers, 4 CTAs of 256 threads per SM, for large batches); 0 = chosen
 * by grid size at every launch (default). */
int chs_set_register_variant(ChsHandle* h, int32_t variant);

/* How chs_step maps onto kernels: 1 = fused, one launch per control step (drone chains on the first warps of every
 * CTA, the herd on the others: the shortest path for batches of about one wave); 2 = split, two launches
 * (drone_kernel: one thread per drone, no shared memory, no barrier; step_kernel: herd, task layer, observation);
 * 0 = chosen by batch size (default).  Results are bit-identical between the modes in the fp64 build and equal up to fp32
 * contraction in the fp32 build (two compilations of the same device functions). */
int chs_set_step_mode(ChsHandle* h, int32_t mode);

/* Profiling aid: when non-NULL, thread 0 of every CTA of chs_step writes clock64() at the 7 phase boundaries
 * of the fused kernel into int64 [grid][32] (24 stamps used) (grid = ceil(E / envs_per_block)).  NULL disables. */
int chs_set_phase_clock(ChsHandle* h, int64_t* clock_dev);

/* Episode statistics accumulated on device since the last reset of the accumulator:
 * float64 [CHS_NUM_STATS] = { episodes, sum return, sum length (control steps), terminated,
 * trunc altitude, trunc collision, trunc formation, trunc boundary, trunc time limit,
 * non-finite envs, env steps,
 * [11] sum over finished episodes of the 1.7 x path length of all their drones (update_evaluation_metrics,
 *      BaseAviary.py:1418-1426, first step measured from the origin like the reference),
 * [12] sum over finished episodes of the herding effectiveness in percent at their last step
 *      (evaluation_episode_trigger :1439-1450, utils/evaluation.py:100-138),
 * [13] sum over ALL env-steps of the effectiveness in percent (the log's effectiveness_per_step; mean = [13] / [10]),
 * [14] curriculum successes = increments of curriculum_success_tally, [15] level-ups
 *      (utils/curriculum_learning.py:200-222), both summed over envs }.
 * Copies to stats_host (synchronous); `reset`!=0 zeroes the accumulator afterwards.  This is the vector all-reduced over
 * NCCL once per training iteration. */
int chs_stats(ChsHandle* h, double* stats_host, int32_t reset);

/* Launch-shape override after creation (bench autotuning): envs per CTA (1..32: a power of two, or any even count --
 * the CTA then keeps the next power of two as its shared-memory stride and leaves the other slots empty, which lets the
 * grid be fitted to whole waves of the 148 SMs, e.g. 28 envs x 147 CTAs for 4096 envs) and threads per CTA (multiple of
 * 32, 32..512).  Results do not depend on the shape. */
int chs_set_launch(ChsHandle* h, int32_t envs_per_block, int32_t threads_per_block);

#ifdef __cplusplus
}
#endif
#endif  /* CHS_H_ */
