// The fused env-step kernel and the masked reset kernel.
//
// One CTA owns G consecutive environments (G = 1..32, power of two).  Inside a phase every thread takes
// flat (entity, env) items with the env index fastest, so lanes of a warp touch consecutive environments:
// HBM accesses are coalesced along the env axis and shared-memory accesses are conflict free.
//
//   phase 0  zero the obs staging area, stage the herd (pos, vel) in shared memory, read env scalars,
//            build the list of envs whose herd updates this step (step_counter_A even, BaseAviary.py:454)
//   phase 1  "drone" threads: action -> PID -> K DYN substeps in registers, state written back to HBM
//            "worker" threads (all others), concurrently: alpha-lattice term of the herd update, which
//            does not depend on the drones
//   phase 2  drone threads: neighbour ranking, spacing / cattle-distance reward pieces, obs row pieces
//            all threads: shepherd + predator + gamma terms, herd velocity write-back; winding numbers
//   phase 3  one thread per env: reward, terminated (x2, stateful), truncated, counters, statistics
//   phase 4  coalesced copy of the staged observation rows (and zero padding) to the obs tensor
//   phase 5  (only CTAs with a finished env, auto_reset) reset those envs in place and emit their first obs
#pragma once
#ifndef CHS_EXP
#define CHS_EXP 0
#endif
#include "chs_drone.cuh"
#include "chs_flock.cuh"
#include "chs_task.cuh"

namespace chs {

template <typename T> struct Smem {
  typename VT<T>::v4* dpos;
  typename VT<T>::v4* redA;
  typename VT<T>::v4* redB;
  typename VT<T>::v2* cpos;
  typename VT<T>::v2* cvel;
  typename VT<T>::v2* u;
  float* stage;
  double2* ccent;   // herd centroid per env
  double* ctab;     // curriculum rows + 1/n (chs_task.cuh)
  float* zeros;     // zero block, source of the obs padding for bulk stores
  int* n; int* herd; int* done; int* newn; int* row; int* flist; int* termrow; int* dlist; int* epi; int* level;
  int* misc;   // [0] number of flock-active envs, [1] number of finished envs
};

template <typename T> __device__ __forceinline__ Smem<T> carve(unsigned char* base, const KParams& p) {
  const SmemLayout& l = p.lay;
  Smem<T> s;
  s.dpos = reinterpret_cast<typename VT<T>::v4*>(base + l.off_dpos);
  s.redA = reinterpret_cast<typename VT<T>::v4*>(base + l.off_redA);
  s.redB = reinterpret_cast<typename VT<T>::v4*>(base + l.off_redB);
  s.cpos = reinterpret_cast<typename VT<T>::v2*>(base + l.off_cpos);
  s.cvel = reinterpret_cast<typename VT<T>::v2*>(base + l.off_cvel);
  s.u = reinterpret_cast<typename VT<T>::v2*>(base + l.off_u);
  s.stage = reinterpret_cast<float*>(base + l.off_stage);
  s.ccent = reinterpret_cast<double2*>(base + l.off_ccent);
  s.ctab = reinterpret_cast<double*>(base + l.off_ctab);
  s.zeros = reinterpret_cast<float*>(base + l.off_zero);
  int* ib = reinterpret_cast<int*>(base + l.off_int);
  s.n = ib + 32 * I_N; s.herd = ib + 32 * I_HERD; s.done = ib + 32 * I_DONE; s.newn = ib + 32 * I_NEWN;
  s.row = ib + 32 * I_ROW; s.flist = ib + 32 * I_FLOCK; s.termrow = ib + 32 * I_TERMROW; s.dlist = ib + 32 * I_DLIST; s.epi = ib + 32 * I_EPI; s.level = ib + 32 * I_LEVEL;
  s.misc = ib + 32 * I_COUNT;
  return s;
}

__device__ __forceinline__ int ilog2_pow2(int g) { return 31 - __clz(g); }

// (slot, row) = (it % n, it / n) of a flat item index without an integer division per item: `magic` = 2^32 / n + 1 is
// exact for it < 2^22 and 2 <= n <= 32 (checked exhaustively); n == 1 passes the index through.
__device__ __forceinline__ unsigned div_magic(int n) { return n > 1 ? 0xffffffffu / (unsigned)n + 1u : 0u; }
__device__ __forceinline__ void split_item(int it, int n, unsigned magic, int& slot, int& row) {
  row = (n == 1) ? it : (int)__umulhi((unsigned)it, magic);
  slot = it - row * n;
}

// 16-byte asynchronous global->shared copy (LDGSTS); bytes past `src_bytes` (0..16) are zero-filled.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
// Bulk (TMA) shared->global store: one thread moves a whole staged observation with two instructions, no LSU
// traffic.  Sizes and both addresses are multiples of 16 bytes (KParams::bulk_ok).  The generic-proxy writes that
// filled the staging area are ordered before the async-proxy read by fence.proxy.async after the CTA barrier.
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, unsigned bytes) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(sa), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// One env's observation image = staged rows followed by zero padding (from the CTA's zero block) up to R*L floats.
__device__ __forceinline__ void bulk_copy_rows(const KParams& p, const float* stage_g, const float* zeros, float* dst) {
  const unsigned total = (unsigned)(p.R * p.L) * 4u;
  unsigned c1 = (unsigned)p.stage_copy * 4u;
  if (c1 > total) c1 = total;
  bulk_store(dst, stage_g, c1);
  if (total > c1 && !p.pad_persist) bulk_store(reinterpret_cast<char*>(dst) + c1, zeros, total - c1);
}

// The staged live rows of one env to its image, skipping -- under the persistent-padding contract -- the 16-byte aligned
// interior of each row's dead column range (12 - 2 unused neighbour slots = 80 of 264 bytes per row at L = 66): the row
// tail and the next row's head are contiguous, so D rows leave as D + 1 bulk stores.  26 % fewer bytes per env at 3 drones,
// which is what bounds the host path (the stores cross PCIe) and the large batches (HBM).
__device__ __forceinline__ void bulk_store_live(const KParams& p, const float* stage_g, float* dst) {
  const unsigned total = (unsigned)(p.R * p.L) * 4u;
  unsigned c1 = (unsigned)p.stage_copy * 4u;
  if (c1 > total) c1 = total;
  const char* src = reinterpret_cast<const char*>(stage_g);
  char* out = reinterpret_cast<char*>(dst);
  unsigned cur = 0;
  if (p.dead_hi > p.dead_lo) {
    const unsigned rowb = (unsigned)p.L * 4u;
    for (int d = 0; d < p.D; ++d) {
      const unsigned a = ((unsigned)d * rowb + (unsigned)p.dead_lo * 4u + 15u) & ~15u;
      const unsigned b = ((unsigned)d * rowb + (unsigned)p.dead_hi * 4u) & ~15u;
      if (b > a && b <= c1) {
        if (a > cur) bulk_store(out + cur, src + cur, a - cur);
        cur = b;
      }
    }
  }
  if (c1 > cur) bulk_store(out + cur, src + cur, c1 - cur);
}

// Programmatic dependent launch (PDL): every kernel of the step path is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so its CTAs may become resident -- and run whatever precedes pdl_wait()
// -- while the previous kernel of the stream is still draining.  pdl_wait() returns once that kernel has completed and its
// memory is visible; NOTHING that touches global memory may precede it.  pdl_launch_next() lets the next kernel of the stream
// (if it opted in the same way) start its own prologue.  Both are no-ops for a launch without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_next() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- 2-D TMA tile load (global -> shared) completing on an mbarrier ----
__device__ __forceinline__ void mbar_init(void* mbar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(mbar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* mbar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(mbar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int x, int y, void* mbar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(tmap), "r"(x), "r"(y),
                 "r"((unsigned)__cvta_generic_to_shared(mbar)) : "memory");
}
// 1-D bulk (TMA) load global -> shared completing on an mbarrier; addresses and size are multiples of 16 bytes
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, unsigned bytes, void* mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(bytes),
                 "r"((unsigned)__cvta_generic_to_shared(mbar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* mbar, unsigned parity) {
  const unsigned addr = (unsigned)__cvta_generic_to_shared(mbar);
  unsigned done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.b32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
}

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Copies the staged rows of env g (plus zero padding up to R*L floats) to dst, one warp per env.
__device__ __forceinline__ void copy_rows(const KParams& p, const float* stage_g, float* dst, int lane) {
  const int RL = p.R * p.L;
  if (p.vec4_ok) {
    const float4* src4 = reinterpret_cast<const float4*>(stage_g);
    float4* dst4 = reinterpret_cast<float4*>(dst);
    const int nq = RL >> 2;
    const int sq = min(p.stage_copy >> 2, nq);
    // staged part: two independent LDS -> STG pairs in flight per lane; zero padding: stores only
    int r = lane;
    for (; r + 32 < sq; r += 64) {
      const float4 v0 = src4[r], v1 = src4[r + 32];
      __stcs(dst4 + r, v0);
      __stcs(dst4 + r + 32, v1);
    }
    if (r < sq) { __stcs(dst4 + r, src4[r]); r += 32; }
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!p.pad_persist) {
#pragma unroll 4
      for (; r < nq; r += 32) __stcs(dst4 + r, z);
    }
  } else {
    const int DL = p.D * p.L;
    const int end = p.pad_persist ? (DL < RL ? DL : RL) : RL;
    for (int r = lane; r < end; r += 32) __stcs(dst + r, r < DL ? stage_g[r] : 0.f);
  }
}

// Draws of one reset (BaseAviary.py:307 and :617,631), keyed by (seed, global env id, episode).
__device__ __forceinline__ int draw_num_drones(const KParams& p, int gid, int episode) {
  if (p.fixed_n > 0) return p.fixed_n;
  uint32_t c[4] = {(uint32_t)gid, (uint32_t)episode, 0u, (uint32_t)TAG_NDRONES};
  philox4x32_10(c, p.seed_lo, p.seed_hi);
  const int span = p.max_n - p.min_n + 1;
  int n = p.min_n + (int)(u01_24(c[0]) * span);
  return n > p.D ? p.D : n;
}

// update_evaluation_metrics (BaseAviary.py:1418-1426): the running distance of a drone grows by 1.7 x the xy length of
// this step; `acc` < 0 marks an episode that has not stepped yet, whose first step is measured from the origin
// (last_drones_pos is zeroed by reset, :317).
template <typename T>
__device__ __forceinline__ T path_advance(T acc, T bx, T by, T ax, T ay) {
  const bool fresh = acc < T(0);
  const T dx = ax - (fresh ? T(0) : bx), dy = ay - (fresh ? T(0) : by);
  return (fresh ? T(0) : acc) + Mth<T>::sqrt(dx * dx + dy * dy) * T(1.7);
}

// Phase 5 / reset kernel body: in-place reset of the envs listed in s.dlist[0..nd) and emission of their
// first observation (BaseAviary.reset :280-331, _housekeeping :547-637, initialize_drone_positions
// :251-277).  PID state, prev_cent_dist, the spacing clock and the curriculum survive.
// Must be called by every thread of the CTA; `nd` is CTA-uniform.
template <typename T, bool FULL>
__device__ __forceinline__ void reset_envs(const KArgs<T>& a, Smem<T>& s, int e0, int nd) {
  const KParams& p = a.p;
  const int t = threadIdx.x, nt = blockDim.x, G = p.G, E = p.E;
  for (int i = t; i < (p.zero_bytes >> 4); i += nt) reinterpret_cast<float4*>(s.zeros)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();   // the copy-out of the terminal observations has finished reading the staging area
  const int sq = p.stage_stride >> 2;
  for (int it = t; it < nd * sq; it += nt) {
    const int g = s.dlist[it / sq];
    reinterpret_cast<float4*>(s.stage + (size_t)g * p.stage_stride)[it % sq] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // cows: table position, velocity 0.2*(cos th, sin th), th = pi*(2u-1) (BaseAviary.py:611-637)
  for (int it = t; it < nd * p.M; it += nt) {
    const int g = s.dlist[it % nd], j = it / nd, e = e0 + g;
    const auto xy = a.spawn[(size_t)s.row[g] * p.spawn_stride + j];
    uint32_t c[4] = {(uint32_t)(p.env_id_offset + e), (uint32_t)s.epi[g] /* new episode index */, (uint32_t)j,
                     (uint32_t)TAG_COW};
    philox4x32_10(c, p.seed_lo, p.seed_hi);
    // c[0] stands for the yaw draw (unused on the DYN path), c[1] for the velocity angle
    T sn, cs;
    Mth<T>::sincospi(T(2.0 * u01_24(c[1]) - 1.0), &sn, &cs);
    const auto v = mk2<T>(T(k::MAX_VEL_CATTLE) * cs, T(k::MAX_VEL_CATTLE) * sn);
    a.cpos[(size_t)j * E + e] = xy;
    a.cvel[(size_t)j * E + e] = v;
    s.cpos[j * G + g] = xy;
  }
  // drones: line / two-row spawn at z = 0.45 (BaseAviary.py:251-277); every slot's kinematics zeroed
  for (int it = t; it < nd * p.D; it += nt) {
    const int g = s.dlist[it % nd], d = it / nd, e = e0 + g;
    const int n = s.newn[g];
    DroneState<T> ds;
    load_drone<T>(a, d, e, ds);     // keeps the PID fields
#pragma unroll
    for (int i = 0; i < 3; ++i) { ds.p[i] = 0; ds.v[i] = 0; ds.w[i] = 0; ds.rpy[i] = 0; ds.av[i] = 0; }
    ds.q[0] = ds.q[1] = ds.q[2] = ds.q[3] = 0;
    if (d < n) {
      int col = d;
      T y = 0;
      if (n > 4) { const int n1 = n / 2; if (d >= n1) { col = d - n1; y = T(k::DRONE_SPAWN_SPACING); } }
      ds.p[0] = T(col) * T(k::DRONE_SPAWN_SPACING); ds.p[1] = y; ds.p[2] = T(k::TARGET_ALT);
      ds.q[3] = T(1);
      s.stage[(size_t)g * p.stage_stride + d * p.L] = (float)ds.p[2];
    }
    store_drone<T>(a, d, e, ds);
    a.rpm[(size_t)d * E + e] = mk4<T>(0, 0, 0, 0);
    a.path[(size_t)d * E + e] = T(-1);
    s.dpos[d * G + g] = mk4<T>(ds.p[0], ds.p[1], ds.p[2], T(0));
    if (d == 0) s.n[g] = n;
  }
  __syncthreads();
  for (int it = t; it < nd * p.D; it += nt) {
    const int g = s.dlist[it % nd], d = it / nd;
    const int n = s.n[g];
    if (d < n) {
      drone_neighbors<T>(p, nullptr, g, d, n, 0, false, s.dpos, s.redA, s.redB, s.stage, nullptr);
      drone_cattle<T>(p, g, d, false, s.dpos, s.cpos, s.redB, s.stage);
    }
  }
  __syncthreads();
  if (a.obs && (!FULL || p.bulk_ok)) {
    if (t < nd) {
      const int g = s.dlist[t];
      fence_async_smem();
      bulk_copy_rows(p, s.stage + (size_t)g * p.stage_stride, s.zeros, a.obs + (size_t)(e0 + g) * p.R * p.L);
      bulk_commit();
      bulk_wait_read();
    }
    if (p.pad_persist) {
      // chs_reset is the one call that writes the caller-owned padding rows (ChsParams.persistent_padding)
      const int q0 = p.stage_copy >> 2, nq = (p.R * p.L) >> 2;
      for (int it = t; it < nd * (nq - q0); it += nt) {
        const int g = s.dlist[it / (nq - q0)];
        reinterpret_cast<float4*>(a.obs + (size_t)(e0 + g) * p.R * p.L)[q0 + it % (nq - q0)] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  } else if (FULL && a.obs) {
    const int warp = t >> 5, lane = t & 31, nwarp = nt >> 5;
    for (int i = warp; i < nd; i += nwarp) {
      const int g = s.dlist[i];
      copy_rows(p, s.stage + (size_t)g * p.stage_stride, a.obs + (size_t)(e0 + g) * p.R * p.L, lane);
    }
  }
}

// Masked reset.  Persistent CTAs: each one scans the mask bytes of its groups of G consecutive envs at once (one
// coalesced pass), keeps the groups that hold a selected env in a work list and resets only those -- the usual
// case after a step is that most groups have nothing to do, and launching a CTA per group costs more than the work.
template <typename T>
__global__ void __launch_bounds__(512) reset_kernel(const KArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int work[256];
  __shared__ int nwork;
  const KParams& p = a.p;
  Smem<T> s = carve<T>(smem_raw, p);
  const int t = threadIdx.x, nt = blockDim.x, G = p.G, E = p.E;
  const int ngroups = (E + G - 1) / G;
  const int per_cta = (ngroups + gridDim.x - 1) / gridDim.x;
  for (int base = 0; base < per_cta; base += 256) {
    if (t == 0) nwork = 0;
    __syncthreads();
    for (int k = t; k < 256 && base + k < per_cta; k += nt) {
      const int grp = blockIdx.x + (base + k) * gridDim.x;       // interleaved: neighbouring CTAs read neighbouring bytes
      if (grp < ngroups) {
        bool any = (a.reset_mask == nullptr);
        if (!any) {
          const int eb = grp * G, en = (eb + G < E) ? eb + G : E;
          for (int e = eb; e < en; ++e) any = any || (a.reset_mask[e] != 0);
        }
        if (any) work[atomicAdd(&nwork, 1)] = grp;               // processing order does not affect the results
      }
    }
    __syncthreads();
    const int nw = nwork;
    for (int w = 0; w < nw; ++w) {
      const int e0 = work[w] * G;
      if (t < 32) {
        bool sel = false;
        const int e = e0 + t;
        if (t < G && e < E && (a.reset_mask == nullptr || a.reset_mask[e])) {
          sel = true;
          int4 ei = a.envi[e];
          int4 ej = a.envj[e];
          ej.z += 1;                                         // episode index
          const int n = draw_num_drones(p, p.env_id_offset + e, ej.z);
          int row = ei.w + 1;                                // Cattle_Spawn_Index pre-increment with wrap (:600-603)
          if (row >= p.n_sims) row = 0;
          a.envi[e] = make_int4(0, 0, n, row);
          a.envj[e] = ej;
          a.epret[e] = 0.0;
          s.newn[t] = n; s.row[t] = row; s.epi[t] = ej.z;
        }
        const unsigned m = __ballot_sync(0xffffffffu, sel);
        if (sel) s.dlist[__popc(m & ((1u << t) - 1u))] = t;
        if (t == 0) s.misc[1] = __popc(m);
      }
      __syncthreads();
      const int nd = s.misc[1];
      if (nd > 0) reset_envs<T, true>(a, s, e0, nd);
      __syncthreads();
    }
  }
}

// Deferred reset of the envs the step kernel listed (split mode): one WARP per finished env, no shared memory and
// no CTA barrier, persistent grid.  Same arithmetic as reset_envs (BaseAviary.reset :280-331, _housekeeping :547-637,
// initialize_drone_positions :251-277) -- positions at reset are analytic (line / two-row spawn, table row), so every
// lane can rebuild what it needs; tests/test_cuda_env.py::test_split_mode_equals_fused_mode holds the two
// implementations together bit for bit.
template <typename T>
__device__ __forceinline__ void spawn_xy(int d, int n, T& x, T& y) {
  int col = d;
  y = T(0);
  if (n > 4) { const int n1 = n / 2; if (d >= n1) { col = d - n1; y = T(k::DRONE_SPAWN_SPACING); } }
  x = T(col) * T(k::DRONE_SPAWN_SPACING);
}

// One finished env rebuilt by one warp (all 32 lanes call; n / row / epi are warp-uniform): cows from the spawn row
// with Philox velocity angles, drone slots zeroed except the PID state, and the first observation of the new episode
// written straight from registers -- everything a fresh episode shows is analytic in (n, row).
// `parts` selects what this call does (bit 0 cows, bit 1 drones, bit 2 observation): the three are independent, so the
// fused kernel hands them to different warps.
enum ResetPart { RESET_COWS = 1, RESET_DRONES = 2, RESET_OBS = 4, RESET_ALL = 7 };
// The only state a rebuild reads is the PID memory of the drone slots (fields E, F, G); `PidPre` lets the caller issue
// those loads before it knows n / row / epi, so that their DRAM round trip overlaps the env-scalar one.
template <typename T> struct PidPre { typename VT<T>::v4 e, f, g; };
template <typename T>
__device__ __forceinline__ PidPre<T> reset_preload(const KArgs<T>& a, int e, int lane) {
  PidPre<T> r;
  r.e = r.f = r.g = mk4<T>(0, 0, 0, 0);
  if (lane < a.p.D) {
    const size_t st = (size_t)a.p.D * a.p.E;
    const typename VT<T>::v4* base = a.drone + (size_t)lane * a.p.E + e;
    r.e = base[F_E * st]; r.f = base[F_F * st]; r.g = base[F_G * st];
  }
  return r;
}
template <typename T>
__device__ __forceinline__ void reset_env_warp(const KArgs<T>& a, int e, int n, int row, int epi, int lane, int parts,
                                               const PidPre<T>& pid) {
  const KParams& p = a.p;
  const unsigned full = 0xffffffffu;
  const int E = p.E, D = p.D, M = p.M, L = p.L;
  const int RL = p.R * L;
  const int cbase = (p.packing == CHS_PACK_MARL) ? 10 + 2 * 4 : 10 + 2 * CHS_MAX_DRONES;
  const int nobs = (M < 16) ? M : 16;
  const typename VT<T>::v2* srow = a.spawn + (size_t)row * p.spawn_stride;
#define CHS_COLD_LOOP _Pragma("unroll 1")
  // this code runs once per finished env, cold: every instruction is an instruction-cache miss, so its loops stay rolled
  if (parts & RESET_COWS) CHS_COLD_LOOP for (int j = lane; j < M; j += 32) {                       // cows (BaseAviary.py:611-637)
    const auto xy = srow[j];
    uint32_t c[4] = {(uint32_t)(p.env_id_offset + e), (uint32_t)epi, (uint32_t)j, (uint32_t)TAG_COW};
    philox4x32_10(c, p.seed_lo, p.seed_hi);
    T sn, cs, ang;
    // 2u - 1 with u = k / 2^24: exact in float as well ((k - 2^23) / 2^23 has a 24-bit numerator), so the float build
    // skips the double arithmetic and lands on the same bits
    if (sizeof(T) == 4) ang = T((float)(c[1] >> 8) * 1.1920928955078125e-07f - 1.0f);
    else ang = T(2.0 * u01_24(c[1]) - 1.0);
    Mth<T>::sincospi(ang, &sn, &cs);
    a.cpos[(size_t)j * E + e] = xy;
    a.cvel[(size_t)j * E + e] = mk2<T>(T(k::MAX_VEL_CATTLE) * cs, T(k::MAX_VEL_CATTLE) * sn);
  }
  if ((parts & RESET_DRONES) && lane < D) {   // drones (D <= 12 slots, one lane each): kinematics zeroed, PID kept
    const int d = lane;
    T px = 0, py = 0, pz = 0, qw = 0;
    if (d < n) { spawn_xy<T>(d, n, px, py); pz = T(k::TARGET_ALT); qw = T(1); }
    const size_t st = (size_t)D * E;
    typename VT<T>::v4* base = a.drone + (size_t)d * E + e;
    base[F_A * st] = mk4<T>(px, py, pz, 0);
    base[F_B * st] = mk4<T>(0, 0, 0, qw);
    base[F_C * st] = mk4<T>(0, 0, 0, 0);
    base[F_D * st] = mk4<T>(0, 0, 0, 0);
    base[F_E * st] = mk4<T>(pid.e.x, pid.e.y, pid.e.z, 0);
    base[F_F * st] = mk4<T>(pid.f.x, pid.f.y, pid.f.z, 0);
    base[F_G * st] = mk4<T>(pid.g.x, pid.g.y, pid.g.z, 0);
    a.rpm[(size_t)d * E + e] = mk4<T>(0, 0, 0, 0);
    a.path[(size_t)d * E + e] = T(-1);
  }
  if (a.obs == nullptr || !(parts & RESET_OBS)) return;
  // Lane l holds the spawn position of drone l (line / two-row spawn, BaseAviary.py:251-277); whoever needs another drone's
  // position takes it from that lane with a shuffle.  A rebuild costs ~2000 warp instructions and reset_list_kernel is bound
  // by instruction issue (as many instructions per step as drone_kernel at 7 % finished envs), so nothing is recomputed.
  T sx = T(0), sy = T(0);
  if (lane < n) spawn_xy<T>(lane, n, sx, sy);
  // two nearest other drones of drone `lane` (strict '<' keeps index order on ties, like drone_neighbors)
  int j1 = -1, j2 = -1;
  {
    T d1 = Mth<T>::inf(), d2 = Mth<T>::inf();
    CHS_COLD_LOOP for (int j = 0; j < n; ++j) {
      const T ox = __shfl_sync(full, sx, j), oy = __shfl_sync(full, sy, j);
      const T dx = ox - sx, dy = oy - sy;
      const T dist = (j == lane || lane >= n) ? Mth<T>::inf() : Mth<T>::sqrt(dx * dx + dy * dy);
      const bool lt1 = dist < d1, lt2 = dist < d2;
      d2 = lt1 ? d1 : (lt2 ? dist : d2);
      j2 = lt1 ? j1 : (lt2 ? j : j2);
      d1 = lt1 ? dist : d1;
      j1 = lt1 ? j : j1;
    }
  }
  float* ob = a.obs + (size_t)e * RL;
  // lane l holds component (l & 1) of cow (l >> 1): the 32 cattle columns of every row come from one register
  T cow_c = T(0);
  const bool cow_ok = (lane >> 1) < nobs;
  if (cow_ok) { const auto xy = srow[lane >> 1]; cow_c = (lane & 1) ? xy.y : xy.x; }
  CHS_COLD_LOOP for (int d = 0; d < n && d < p.R; ++d) {
    const T mx = __shfl_sync(full, sx, d), my = __shfl_sync(full, sy, d);
    const int nb1 = __shfl_sync(full, j1, d), nb2 = __shfl_sync(full, j2, d);
    // columns 0..31, one per lane: z | zeros | the two neighbour slots (columns 10..13) | zeros
    const int nbl = (lane < 12) ? nb1 : nb2;
    const T ox = __shfl_sync(full, sx, nbl >= 0 ? nbl : 0), oy = __shfl_sync(full, sy, nbl >= 0 ? nbl : 0);
    float v = 0.f;
    if (lane == 0) v = (float)T(k::TARGET_ALT);
    else if (lane >= 10 && lane < 14 && nbl >= 0) v = (lane & 1) ? (float)(oy - my) : (float)(ox - mx);
    float* orow = ob + d * L;
    if (lane < cbase) orow[lane] = v;
    if (32 + lane < cbase) orow[32 + lane] = 0.f;                        // sb3 packing: columns 32, 33 (cbase <= 34)
    if (cbase + lane < L) orow[cbase + lane] = cow_ok ? (float)(cow_c - ((lane & 1) ? my : mx)) : 0.f;
    CHS_COLD_LOOP for (int c = cbase + 32 + lane; c < L; c += 32) orow[c] = 0.f;
  }
  const int zfrom = (n < p.R ? n : p.R) * L;
  if (p.pad_persist) {   // rows >= D are the caller's constant zeros; only unused drone slots (n < D) are cleared
    const int zend = (D < p.R ? D : p.R) * L;
    CHS_COLD_LOOP for (int idx = zfrom + lane; idx < zend; idx += 32) ob[idx] = 0.f;
  } else if (p.vec4_ok) {       // R*L is a multiple of 4 and the image is 16-byte aligned: the zero rows go out as float4
    const int z4 = (zfrom + 3) >> 2;
    if (zfrom + lane < (z4 << 2)) ob[zfrom + lane] = 0.f;        // up to 3 floats to the next 16-byte boundary
    float4* o4 = reinterpret_cast<float4*>(ob);
    for (int r = z4 + lane; r < (RL >> 2); r += 32) o4[r] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
    for (int idx = zfrom + lane; idx < RL; idx += 32) ob[idx] = 0.f;
  }
}

// Alpha-term in-range masks of every cow of every spawn row (one thread per (row, cow)), same expression as the step kernel.
template <typename T>
__global__ void __launch_bounds__(256) alpha_mask_kernel(const typename VT<T>::v2* spawn, int n_sims, int stride, int M, unsigned* out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_sims * M) return;
  const int row = idx / M, i = idx - row * M, nch = (M + 31) >> 5;
  const typename VT<T>::v2* r = spawn + (size_t)row * stride;
  const auto qi = r[i];
  for (int c = 0; c < nch; ++c)
    out[(size_t)idx * nch + c] = alpha_chunk_mask<T>([&](int j) { return r[j]; }, qi, i, c * 32, M);
}

// Position-only part of the alpha term of every cow of every spawn row (one thread per (row, cow)): gradient sum and the
// consensus weight of each in-range neighbour, walked in the order of the mask table with lattice_term's own arithmetic.
// `*max_cnt` receives the largest neighbour count when `grad` is null (first pass, sizes the weight rows).
template <typename T>
__global__ void __launch_bounds__(256) alpha_weight_kernel(const typename VT<T>::v2* spawn, int n_sims, int stride, int M,
                                                           const unsigned* masks, typename VT<T>::v2* grad, T* wts, int wstride,
                                                           int* max_cnt) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_sims * M) return;
  const int row = idx / M, i = idx - row * M, nch = (M + 31) >> 5;
  const typename VT<T>::v2* r = spawn + (size_t)row * stride;
  const auto qi = r[i];
  T gx = 0, gy = 0, cx = 0, cy = 0;
  int kk = 0;
  for (int c = 0; c < nch; ++c) {
    unsigned mask = masks[(size_t)idx * nch + c];
    while (mask) {
      const int j = c * 32 + __ffs(mask) - 1;
      mask &= mask - 1;
      if (grad) {
        const auto qj = r[j];
        // the gradient sum grows through lattice_term itself (the step kernel's arithmetic); the consensus weight is
        // lattice_term's `b`, taken from the same expressions
        lattice_term<T>(qj.x - qi.x, qj.y - qi.y, T(0), T(0), T(k::SIGMA_ALPHA), T(1.0 / k::SIGMA_ALPHA), gx, gy, cx, cy);
        T w, b;
        lattice_weights<T>(qj.x - qi.x, qj.y - qi.y, T(k::SIGMA_ALPHA), T(1.0 / k::SIGMA_ALPHA), w, b);
        wts[(size_t)idx * wstride + kk] = b;
      }
      ++kk;
    }
  }
  if (grad) grad[idx] = mk2<T>(gx, gy);
  else atomicMax(max_cnt, kk);
}

// Herd centroid of every spawn row (one thread per row) in the summation order of the step kernel: a serial pass for herds
// under 48 cows, 8 strided partial sums folded by an xor tree for larger ones (centroid_wide).
template <typename T>
__global__ void __launch_bounds__(128) spawn_centroid_kernel(const typename VT<T>::v2* spawn, int n_sims, int stride, int M, double2* out) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n_sims) return;
  const typename VT<T>::v2* r = spawn + (size_t)row * stride;
  const double inv_m = 1.0 / (double)M;
  if (M < 48) {
    double cx = 0, cy = 0;
    for (int j = 0; j < M; ++j) { cx += (double)r[j].x; cy += (double)r[j].y; }
    out[row] = make_double2(cx * inv_m, cy * inv_m);
  } else {
    double sx[8], sy[8];
    for (int sub = 0; sub < 8; ++sub) {
      double cx = 0, cy = 0;
      for (int j = sub; j < M; j += 8) { cx += (double)r[j].x; cy += (double)r[j].y; }
      sx[sub] = cx; sy[sub] = cy;
    }
    for (int o = 4; o > 0; o >>= 1)
      for (int k = 0; k < o; ++k) { sx[k] += sx[k + o]; sy[k] += sy[k + o]; }     // lane k of the tree adds its partner k ^ o
    out[row] = make_double2(sx[0] * inv_m, sy[0] * inv_m);
  }
}

// After a state injection: do the herd positions of every env still equal its spawn row bit for bit?  (Then the
// alpha-range table stays valid.)  One thread per (cow, env); *bad counts the mismatches.
template <typename T>
__global__ void __launch_bounds__(256) spawn_row_check_kernel(const KArgs<T> a, int* bad) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const KParams& p = a.p;
  if (idx >= (long long)p.M * p.E) return;
  const int j = (int)(idx / p.E), e = (int)(idx - (long long)j * p.E);
  const int row = a.envi[e].w;
  bool ok = row >= 0 && row < p.n_sims;
  if (ok) {
    const auto c = a.cpos[idx];
    const auto r = a.spawn[(size_t)row * p.spawn_stride + j];
    ok = (c.x == r.x) && (c.y == r.y);
  }
  if (!ok) atomicAdd(bad, 1);
}

template <typename T>
__global__ void __launch_bounds__(256) reset_list_kernel(const KArgs<T> a) {
  const KParams& p = a.p;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  pdl_wait();
  pdl_launch_next();
  const int count = *a.done_count;
  for (int i = gw; i < count; i += nwarps) {
    const int e = a.done_list[i];
    const PidPre<T> pid = reset_preload<T>(a, e, lane);
    int n = 0, row = 0, epi = 0;
    if (lane == 0) {
      const int4 ei = a.envi[e];
      int4 ej = a.envj[e];
      ej.z += 1; epi = ej.z;
      n = draw_num_drones(p, p.env_id_offset + e, epi);
      row = ei.w + 1;
      if (row >= p.n_sims) row = 0;
      a.envi[e] = make_int4(0, 0, n, row);
      a.envj[e] = ej;
      a.epret[e] = 0.0;
    }
    n = __shfl_sync(full, n, 0); row = __shfl_sync(full, row, 0); epi = __shfl_sync(full, epi, 0);
    reset_env_warp<T>(a, e, n, row, epi, lane, RESET_ALL, pid);
  }
}

// Split mode, kernel 1 of 2: one thread per (drone slot, env), env index fastest.  action -> PID -> K DYN substeps in
// registers with no shared memory and no barrier, so the long dependent chain of every drone is hidden by plain
// occupancy; writes the drone state, the RPMs and the (ang_v, yaw) scratch row the second kernel needs for the
// observation.  The fused kernel below does the same work on its "drone" threads (phase 1).
template <typename T>
__global__ void __launch_bounds__(256) drone_kernel(const KArgs<T> a) {
  const KParams& p = a.p;
  const int E = p.E, D = p.D;
  // grid = (ceil(E / 256), D): the drone slot is the block row, so no thread divides a flat index by E
  const int e = blockIdx.x * blockDim.x + threadIdx.x, d = blockIdx.y;
  pdl_wait();
  pdl_launch_next();
  if (e == 0 && d == 0 && a.done_count) *a.done_count = 0;      // the list of finished envs is rebuilt by this step's step_kernel
  if (e >= E || d >= D) return;
  const int n = a.envi[e].z;
  if (d >= n) return;
  const T ctrl_dt = T(1.0) / T(p.ctrl_freq);
  const T pyb_dt = T(1.0) / T(p.ctrl_freq * p.K);
  DroneState<T> ds;
  load_drone<T>(a, d, e, ds);
  const T path0 = a.path[(size_t)d * E + e];
  const T bx = ds.p[0], by = ds.p[1];
  float act[4] = {0.f, 0.f, 0.f, 0.f};
  const float* ap = a.actions + ((size_t)e * D + d) * p.A;
  if (p.A == 4 && p.act4_ok) {
    const float4 v = *reinterpret_cast<const float4*>(ap);
    act[0] = v.x; act[1] = v.y; act[2] = v.z; act[3] = v.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) if (i < p.A) act[i] = ap[i];
  }
  T rpm[4];
  action_to_rpm<T>(p, ds, act, ctrl_dt, rpm);
  if (p.aero_flags) {
    const auto pr = a.rpm[(size_t)d * E + e];
    const T prev[4] = {pr.x, pr.y, pr.z, pr.w};
    DroneState<T> tmp = ds;
    T rp[4] = {rpm[0], rpm[1], rpm[2], rpm[3]};
    dyn_substeps_aero<T>(tmp, rp, prev, p.K, pyb_dt, p.aero_flags, a.pos0 + e, E, n, d);
    ds = tmp;
  } else {
    dyn_substeps<T>(ds, rpm, p.K, pyb_dt);
  }
  store_drone<T>(a, d, e, ds);
  a.rpm[(size_t)d * E + e] = mk4<T>(rpm[0], rpm[1], rpm[2], rpm[3]);
  a.own2[(size_t)d * E + e] = mk4<T>(ds.av[0], ds.av[1], ds.av[2], ds.rpy[2]);
  a.path[(size_t)d * E + e] = path_advance<T>(path0, bx, by, ds.p[0], ds.p[1]);
}

// Two register budgets are compiled: MINB = 1 (up to 128 registers per thread, shortest dependent chains, for
// grids of about one wave where latency decides) and MINB = 4 (64 registers, 4 CTAs of 256 threads per SM, for
// large batches where issue-slot occupancy decides).  The host picks by grid size.
// MARL selects the rllib_envs packing (per-drone reward / flags, sequential reward loop) at compile time.
// SPLIT drops the drone chain (drone_kernel ran it) at compile time: fewer registers, a smaller kernel.
// FULL = false is the lean build for the common configuration (no aero extras, no kinematic herd, async herd staging
// and bulk obs stores usable, no phase stamps): those branches are compiled out, which matters for batches of about
// one wave where the kernel is bound by cold instruction fetch (tools/microbench/icache_cold.cu).
template <typename T, int MAXT, int MINB, bool MARL, bool SPLIT, bool FULL>
__global__ void __launch_bounds__(MAXT, MINB) step_kernel(const __grid_constant__ KArgs<T> a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const KParams& p = a.p;
  Smem<T> s = carve<T>(smem_raw, p);
  const int t = threadIdx.x, nt = blockDim.x, G = p.G, E = p.E, D = p.D, M = p.M;
  const int lg = ilog2_pow2(G);
  // A CTA handles `gact` <= G consecutive envs; G stays the power-of-two stride of the shared-memory arrays (slots >= gact
  // are treated like envs past the end of the batch).  The host picks gact so that the grid fills whole waves of the 148 SMs.
  const int e0 = blockIdx.x * p.gact;
  const int Ev = (e0 + p.gact < E) ? e0 + p.gact : E;      // end of this CTA's envs
  const T ctrl_dt = T(1.0) / T(p.ctrl_freq);
  const T pyb_dt = T(1.0) / T(p.ctrl_freq * p.K);
  const int DG = D * G;
  int ndt = (DG + 31) & ~31;          // "drone" threads: the first warps of the CTA
  if (ndt > nt) ndt = nt;

// CHS_PROFILE_LEAN (a build flag for profiling libraries only): the phase stamps are compiled into the lean build as well, so
// that tools/phase_timing.py times the kernel that actually ships instead of the FULL build (costs 0.4 us at 4096 envs)
#ifdef CHS_PROFILE_LEAN
#define CHS_STAMPS_ON true
#else
#define CHS_STAMPS_ON FULL
#endif
#define CHS_STAMP(i) do { if (CHS_STAMPS_ON && a.dbg_clock && t == 0) a.dbg_clock[(size_t)blockIdx.x * 32 + (i)] = clock64(); } while (0)
#define CHS_GSTAMP(i) do { if (CHS_STAMPS_ON && a.dbg_clock && t == 0) { unsigned long long gt_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_)); a.dbg_clock[(size_t)blockIdx.x * 32 + (i)] = (long long)gt_; } } while (0)
  CHS_STAMP(0);
  CHS_GSTAMP(30);
  // ---------------- phase 0 ----------------
  // Every global read the step needs is issued here, before the first barrier, so that the DRAM round trips
  // of the herd, the env scalars and the first drone item of each drone thread overlap.
  // Early start of the drone chain: with a single drone warp (warp 0) and other warps to do the staging, warp 0 only
  // ARRIVES at the barrier that closes phase 0 and goes straight into its serial chain as soon as its own loads land;
  // it neither stages the herd nor waits for the copies of the others.  It meets the staging warps again on named
  // barrier 1 just before its first shared-memory write (the staging area has to be zeroed by then).
  const bool early = !SPLIT && p.early_ok && ndt == 32 && nt > 32 && (!FULL || (p.cp_ok && !(p.aero_flags & 4)));
  const int sw0 = early ? ndt : 0;           // first thread that takes part in the staging loops
  const int snt = nt - sw0, sw = t - sw0;    // sw < 0: exempt
  // prologue (no global memory): the observation staging area and the zero block are cleared while, under PDL, the
  // previous kernel of the stream is still draining
  const bool tma = sizeof(T) == 4 && p.tma_ok && (!FULL || p.cp_ok);
  // Occupancy build (MINB > 1, batches of several waves, bound by instruction issue): the staging area is cleared and the
  // curriculum table fetched by bulk copies as well -- two more instructions of one thread instead of loops over all of
  // them (16 384 envs 23.3 -> 22.8 us, 65 536 fused 79.9 -> 77.4).  The latency build keeps the loops: its clearing runs in
  // the PDL prologue, under the previous kernel's tail, and a one-wave launch whose CTAs all pull the same page of zeros
  // through L2 at the same moment measured 1 % slower.
  const bool tmaz = tma && MINB > 1;
  const unsigned stage_bytes = (unsigned)(G * p.stage_stride) * 4u;
  if (tma && sw == 0) mbar_init(s.misc + 4, 1);
  if (sw >= 0) {
    if (!tmaz) {
      const int nq = (G * p.stage_stride) >> 2;
      float4* st4 = reinterpret_cast<float4*>(s.stage);
      for (int i = sw; i < nq; i += snt) st4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int i = sw; i < (p.zero_bytes >> 4); i += snt) reinterpret_cast<float4*>(s.zeros)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  pdl_wait();
  int4 pre_i = make_int4(0, 0, 0, 0), pre_j = make_int4(0, 0, 0, 0);
  double2 pre_f = make_double2(0.0, 0.0);
  double pre_ret = 0.0;
  const bool env_lane = (t < G) && (e0 + t < Ev);     // warp 0 only (G <= 32)
  if (env_lane) {
    // env scalars are fetched once, here, and stay in registers until phase 3
    const int e = e0 + t;
    pre_i = a.envi[e]; pre_j = a.envj[e]; pre_f = a.envf[e]; pre_ret = a.epret[e];
  }
  DroneState<T> ds0;
  float act0[4] = {0.f, 0.f, 0.f, 0.f};
  typename VT<T>::v4 sp0 = mk4<T>(0, 0, 0, 0), sp1 = sp0, sp2 = sp0;     // split mode: fields A, C and the own2 row
  T path0 = T(0);                                                        // running path length of the first drone item
  const bool have0 = (t < ndt) && (t < DG) && (e0 + (t & (G - 1)) < Ev);
  if constexpr (SPLIT) {
    if (have0) {
      const int g = t & (G - 1), d = t >> lg, e = e0 + g;
      const size_t st = (size_t)D * E, o = (size_t)d * E + e;
      sp0 = a.drone[F_A * st + o]; sp1 = a.drone[F_C * st + o]; sp2 = a.own2[o]; path0 = a.path[o];
    }
  } else if (have0) {
    const int g = t & (G - 1), d = t >> lg, e = e0 + g;
    load_drone<T>(a, d, e, ds0);
    path0 = a.path[(size_t)d * E + e];
    const float* ap = a.actions + ((size_t)e * D + d) * p.A;
    if (p.A == 4 && p.act4_ok) {
      const float4 v = *reinterpret_cast<const float4*>(ap);
      act0[0] = v.x; act0[1] = v.y; act0[2] = v.z; act0[3] = v.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) if (i < p.A) act0[i] = ap[i];
    }
  }
  {
    using V2 = typename VT<T>::v2;
    if (!FULL || p.cp_ok) {
      // herd rows -> shared memory as 16-byte asynchronous copies (no register staging, all in flight at once);
      // envs past the end of the batch are zero-filled through the src-size operand
      constexpr int EPC = 16 / (int)sizeof(V2);            // envs per 16-byte chunk
      const int cpr = G / EPC;                             // chunks per cow row
      const int lcpr = ilog2_pow2(cpr);
      const int nchunk = 2 * M * cpr;
      if (tma) {
        // fp32: the whole [M][G] tile of positions and the one of velocities arrive through two TMA loads issued by one
        // thread (box = M rows of G envs, out-of-range envs zero-filled by the unit); completion is counted
        // on an mbarrier that the same thread waits on before the barrier that closes phase 0
        if (sw == 0) {
          void* mbar = s.misc + 4;
          mbar_expect_tx(mbar, 2u * (unsigned)(M * G) * (unsigned)sizeof(V2) + (tmaz ? 76u * 16u + stage_bytes : 0u));
          tma_load_2d(s.cpos, &a.tm_cpos, 2 * e0, 0, mbar);
          tma_load_2d(s.cvel, &a.tm_cvel, 2 * e0, 0, mbar);
          if (tmaz) {
            bulk_load(s.ctab, a.tab, 76u * 16u, mbar);
            bulk_load(s.stage, a.zero_page, stage_bytes, mbar);
          }
        }
      } else if (sw >= 0) for (int it = sw; it < nchunk; it += snt) {
        const int arr = (it >= M * cpr) ? 1 : 0;
        const int r = it - arr * M * cpr;
        const int j = r >> lcpr, c = r & (cpr - 1);            // cpr is a power of two
        const int eb = e0 + c * EPC;
        int valid = Ev - eb; valid = valid < 0 ? 0 : (valid > EPC ? EPC : valid);
        const V2* src = (arr ? a.cvel : a.cpos) + (size_t)j * E + (valid > 0 ? eb : e0);
        V2* dst = (arr ? s.cvel : s.cpos) + j * G + c * EPC;
        cp_async16(dst, src, valid * (int)sizeof(V2));
      }
    }
    if (FULL && (p.aero_flags & 4) && !SPLIT) {   // downwash: start-of-step positions of every drone, parked in redA until phase 2
      const size_t st = (size_t)D * E;
      for (int it = t; it < DG; it += nt) {
        const int g = it & (G - 1), d = it >> lg, e = e0 + g;
        if (e < Ev) s.redA[it] = a.drone[F_A * st + (size_t)d * E + e];
      }
    }
    // curriculum table: 76 chunks of 16 bytes
    if (sw >= 0 && !tmaz) for (int it = sw; it < 76; it += snt) cp_async16(s.ctab + 2 * it, a.tab + 2 * it, 16);
    cp_async_commit();
    pdl_launch_next();        // every load of this step is in flight: the next kernel of the stream may start its prologue
    if (FULL && !p.cp_ok) {
      for (int it = t; it < M * G; it += nt) {
        const int g = it & (G - 1), j = it >> lg, e = e0 + g;
        if (e < Ev) { s.cpos[it] = a.cpos[(size_t)j * E + e]; s.cvel[it] = a.cvel[(size_t)j * E + e]; }
        else { s.cpos[it] = mk2<T>(0, 0); s.cvel[it] = mk2<T>(0, 0); }
      }
    }
    if (t < 32) {
      bool fl = false;
      if (t < G) {
        int n = 0;
        if (env_lane) { n = pre_i.z; fl = (((pre_i.y + 1) & 1) == 0); }
        s.n[t] = n; s.herd[t] = 0; s.done[t] = 0; s.level[t] = pre_j.x; s.row[t] = pre_i.w;   // row: spawn row of the running episode
      }
      const unsigned m = __ballot_sync(0xffffffffu, fl);
      if (fl) s.flist[__popc(m & ((1u << t) - 1u))] = t;
      if (t == 0) { s.misc[0] = __popc(m); s.misc[1] = 0; s.misc[2] = (int)div_magic(__popc(m)); }
    }
    cp_async_wait_all();
    if (tma && sw == 0) mbar_wait(s.misc + 4, 0);
  }
  CHS_STAMP(1);
  if (early) {
    if (t < ndt) {
      __syncwarp();
      asm volatile("bar.arrive 0, %0;" ::"r"(nt) : "memory");
    } else {
      asm volatile("bar.sync 0, %0;" ::"r"(nt) : "memory");
      asm volatile("bar.arrive 1, %0;" ::"r"(nt) : "memory");
    }
  } else {
    __syncthreads();
  }
  CHS_STAMP(2);
  // Observation traffic is 3/4 constant zeros (rows >= D of the (R, L) image).  They leave NOW, from the CTA's zero
  // block, issued by the last warp: the write drains while the SMs compute instead of queueing behind the live rows at
  // the end of the kernel, where every CTA of a one-wave launch would hit the memory system at the same moment.  A
  // finished env's image goes to term_obs instead (phase 3 sends both pieces there) and its obs slot receives the
  // first observation of the next episode, whose rows >= D are zeros as well.
  const int ow = t - (nt - 32);             // lane index inside the last warp, < 0 elsewhere
  const bool obs_lane = (!FULL || p.bulk_ok) && a.obs != nullptr && ow >= 0 && ow < G && e0 + ow < Ev;
  if (obs_lane) {
    const unsigned total = (unsigned)(p.R * p.L) * 4u, c1 = (unsigned)p.stage_copy * 4u;
    if (total > c1 && !p.pad_persist) {
      fence_async_smem();
      bulk_store(reinterpret_cast<char*>(a.obs + (size_t)(e0 + ow) * p.R * p.L) + c1, s.zeros, total - c1);
      bulk_commit();
    }
  }

  // ---------------- phase 1 ----------------
  const int nact = s.misc[0];
  const unsigned nmagic = (unsigned)s.misc[2];
  if constexpr (SPLIT) {
    // the drones were integrated by drone_kernel: stage their positions and the own-state part of the obs rows
    if (t < ndt) {
      for (int it = t; it < DG; it += ndt) {
        const int g = it & (G - 1), d = it >> lg, e = e0 + g;
        if (e < Ev && d < s.n[g]) {
          typename VT<T>::v4 A = sp0, C = sp1, O = sp2;
          T pl = path0;
          if (it != t) {
            const size_t st = (size_t)D * E, o = (size_t)d * E + e;
            A = a.drone[F_A * st + o]; C = a.drone[F_C * st + o]; O = a.own2[o]; pl = a.path[o];
          }
          s.dpos[it] = mk4<T>(A.x, A.y, A.z, pl);         // .w: the drone's running path length (phase 3 sums it at episode end)
          float* row = s.stage + (size_t)g * p.stage_stride + d * p.L;
          if ((p.L & 1) == 0) {
            float2* r2 = reinterpret_cast<float2*>(row);
            r2[0] = make_float2((float)A.z, (float)A.w); r2[1] = make_float2((float)C.w, (float)O.w);
            r2[2] = make_float2((float)C.x, (float)C.y); r2[3] = make_float2((float)C.z, (float)O.x);
            r2[4] = make_float2((float)O.y, (float)O.z);
          } else {
            row[0] = (float)A.z; row[1] = (float)A.w; row[2] = (float)C.w; row[3] = (float)O.w;
            row[4] = (float)C.x; row[5] = (float)C.y; row[6] = (float)C.z;
            row[7] = (float)O.x; row[8] = (float)O.y; row[9] = (float)O.z;
          }
        }
      }
    }
  } else if (t < ndt) {
    DroneState<T> held;            // early mode: the one item of this thread, parked until the staging area is ready
    int held_it = -1;
    T held_pl = T(0);
    for (int it = t; it < DG; it += ndt) {
      const int g = it & (G - 1), d = it >> lg, e = e0 + g;
      if (e < Ev && d < s.n[g]) {
        DroneState<T> ds;
        float act[4];
        T pl = path0;
        if (it == t) {
          ds = ds0;
#pragma unroll
          for (int i = 0; i < 4; ++i) act[i] = act0[i];
        } else {
          load_drone<T>(a, d, e, ds);
          pl = a.path[(size_t)d * E + e];
          const float* ap = a.actions + ((size_t)e * D + d) * p.A;
#pragma unroll
          for (int i = 0; i < 4; ++i) act[i] = (i < p.A) ? ap[i] : 0.f;
        }
        const T bx = ds.p[0], by = ds.p[1];
        T rpm[4];
        action_to_rpm<T>(p, ds, act, ctrl_dt, rpm);
        CHS_STAMP(16);
        if (FULL && p.aero_flags) {
          const auto pr = a.rpm[(size_t)d * E + e];
          const T prev[4] = {pr.x, pr.y, pr.z, pr.w};
          DroneState<T> tmp = ds;      // the out-of-line call takes an address: keep `ds` itself in registers
          T rp[4] = {rpm[0], rpm[1], rpm[2], rpm[3]};
          dyn_substeps_aero<T>(tmp, rp, prev, p.K, pyb_dt, p.aero_flags, s.redA + g, G, s.n[g], d);
          ds = tmp;
        } else {
          dyn_substeps<T>(ds, rpm, p.K, pyb_dt);
        }
        CHS_STAMP(17);
        store_drone<T>(a, d, e, ds);
        a.rpm[(size_t)d * E + e] = mk4<T>(rpm[0], rpm[1], rpm[2], rpm[3]);
        pl = path_advance<T>(pl, bx, by, ds.p[0], ds.p[1]);
        a.path[(size_t)d * E + e] = pl;
        if (early) { held = ds; held_it = it; held_pl = pl; continue; }      // DG <= 32: one item per thread
        s.dpos[it] = mk4<T>(ds.p[0], ds.p[1], ds.p[2], pl);    // .w: running path length, summed by phase 3 at episode end
        // own-state part of the obs row (BaseRLAviary.py:286-297): z, rpy, vel, ang_v
        float* row = s.stage + (size_t)g * p.stage_stride + d * p.L;
        row[0] = (float)ds.p[2];
#pragma unroll
        for (int i = 0; i < 3; ++i) { row[1 + i] = (float)ds.rpy[i]; row[4 + i] = (float)ds.v[i]; row[7 + i] = (float)ds.av[i]; }
      }
    }
    if (early) {
      __syncwarp();
      asm volatile("bar.sync 1, %0;" ::"r"(nt) : "memory");     // the staging warps arrived here long ago
      if (held_it >= 0) {
        const int g = held_it & (G - 1), d = held_it >> lg;
        s.dpos[held_it] = mk4<T>(held.p[0], held.p[1], held.p[2], held_pl);
        float* row = s.stage + (size_t)g * p.stage_stride + d * p.L;
        row[0] = (float)held.p[2];
#pragma unroll
        for (int i = 0; i < 3; ++i) { row[1 + i] = (float)held.rpy[i]; row[4 + i] = (float)held.v[i]; row[7 + i] = (float)held.av[i]; }
      }
    }
  }
  {
    // worker threads (everyone when the CTA has no spare warps): herd centroid per env in float64, then the
    // alpha-lattice term of the herds that update this step -- neither depends on the drones
    const int wt0 = (ndt < nt && !SPLIT) ? ndt : 0;
    const int nw = nt - wt0;
    if (t >= wt0) {
      const int w = t - wt0;
      // Herd centroid per env (float64; phase 3 needs it).  Small herds: one thread per env, serial, in front of its alpha
      // items.  Herds of 48 cows and more: AFTER the alpha items, 8 lanes per env taken from the END of the worker range
      // (the threads with the fewest alpha items), cows strided over the lanes, xor-shuffle tree -- the serial 100-cow loop
      // made the first G worker threads the last to reach the barrier (4000 cycles, measured with the lean-build stamps;
      // config 3: 17.4 -> 16.2 us), while for 20 cows the tree costs more instructions than it saves latency.  Either way
      // the summation order depends on M only, not on the launch shape or the step mode.
      const bool wide_herd = M >= 48 && !a.acent;
      if (a.acent) {
        if (w < G) s.ccent[w] = a.acent[s.row[w]];      // a property of the spawn row while the herd sits on it
      } else if (!wide_herd && w < G) {
        double cx = 0, cy = 0;
#pragma unroll 4
        for (int j = 0; j < M; ++j) { const auto c = s.cpos[j * G + w]; cx += (double)c.x; cy += (double)c.y; }
        const double inv_m = 1.0 / (double)M;
        s.ccent[w] = make_double2(cx * inv_m, cy * inv_m);
      }
      for (int it = w; it < nact * M; it += nw) {
        int slot, i;
        split_item(it, nact, nmagic, slot, i);
        const int g = s.flist[slot];
        T ux, uy;
        const unsigned* mw = a.amask ? a.amask + ((size_t)s.row[g] * M + i) * ((M + 31) >> 5) : nullptr;
        if (mw && a.aw) {
          const size_t rc = (size_t)s.row[g] * M + i;
          alpha_term_tab<T>(s.cvel, i, g, G, M, mw, a.agrad[rc], a.aw + rc * a.aw_stride, ux, uy);
        } else if (mw && M > 32) { T u2[2]; alpha_term_wide<T>(s.cpos, s.cvel, i, g, G, M, mw, u2); ux = u2[0]; uy = u2[1]; }
        else alpha_term<T>(s.cpos, s.cvel, i, g, G, M, mw, ux, uy);
        s.u[i * G + g] = mk2<T>(ux, uy);
      }
      if (wide_herd) centroid_wide<T>(s.cpos, s.ccent, G, M, nw - 1 - w, nw);
    }
  }
  CHS_STAMP(3);
  __syncthreads();
  CHS_STAMP(4);

  // ---------------- phase 2 ----------------
  // per-drone pieces: neighbour ranking + spacing rewards on the drone threads, cattle pieces on threads half a
  // CTA away so that the two halves run on different warps
  if (FULL && p.cattle_integrate) {
    // non-reference kinematic herd: velocities first, then pos += vel * CTRL_TIMESTEP for
    // every cow of every env, then the herd centroid again; obs / reward / effectiveness below see the new positions
    for (int it = t; it < nact * M; it += nt) {
      int slot, i;
      split_item(it, nact, nmagic, slot, i);
      const int g = s.flist[slot], e = e0 + g;
      const auto ua = s.u[i * G + g];
      const auto v = herd_finish<T>(s.dpos, s.n[g], g, G, s.cpos[i * G + g], s.cvel[i * G + g], ua.x, ua.y);
      a.cvel[(size_t)i * E + e] = v;
      s.cvel[i * G + g] = v;
    }
    __syncthreads();
    for (int it = t; it < M * G; it += nt) {
      const int g = it & (G - 1), j = it >> lg, e = e0 + g;
      if (e < Ev) {
        const auto c = s.cpos[it];
        const auto v = s.cvel[it];
        const auto nc = mk2<T>(c.x + v.x * ctrl_dt, c.y + v.y * ctrl_dt);
        s.cpos[it] = nc;
        a.cpos[(size_t)j * E + e] = nc;
      }
    }
    __syncthreads();
    if (t < G) {
      double cx = 0, cy = 0;
      for (int j = 0; j < M; ++j) { const auto c = s.cpos[j * G + t]; cx += (double)c.x; cy += (double)c.y; }
      const double inv_m = 1.0 / (double)M;
      s.ccent[t] = make_double2(cx * inv_m, cy * inv_m);
    }
  }
  if (t < ndt) {
    for (int it = t; it < DG; it += ndt) {
      const int g = it & (G - 1), d = it >> lg, e = e0 + g;
      const int n = s.n[g];
      if (e < Ev && d < n)
        drone_neighbors<T>(p, s.ctab, g, d, n, s.level[g], true, s.dpos, s.redA, s.redB, s.stage,
                           a.dbg_neighbors ? a.dbg_neighbors + ((size_t)e * D + d) * 2 : nullptr,
                           (CHS_STAMPS_ON && a.dbg_clock && t == 0) ? a.dbg_clock + (size_t)blockIdx.x * 32 + 24 : nullptr);
    }
  }
  CHS_STAMP(19);
  {
    // cattle pieces: C lanes per (drone, env) item, as many as the CTA has threads for (up to 8), on threads half a CTA away
    // from the drone warps; every thread runs the same number of rounds (the lanes of an item meet in shuffles)
    const int half = (nt >> 1) & ~31;
    const int tb = (t >= half) ? t - half : t + nt - half;
    int lc = 3;
    while (lc > 0 && (DG << lc) > nt) --lc;
    const int C = 1 << lc, items = DG << lc;
    for (int base = 0; base < items; base += nt) {
      const int idx = base + tb, it = idx >> lc, sub = idx & (C - 1);
      const int g = it & (G - 1), d = it >> lg;
      const bool valid = idx < items && e0 + g < Ev && d < s.n[g];
      drone_cattle_coop<T>(p, valid, g, d, sub, C, true, s.dpos, s.cpos, s.redB, s.stage);
    }
  }
  CHS_STAMP(20);
  // The drone warps are the critical path of the CTA (they come out of the neighbour ranking while the others have
  // been idle since the alpha term), so the per-cow work below goes to the other warps when the CTA has any.
  const int pw0 = (ndt < nt) ? ndt : 0;
  const int pnw = nt - pw0;
  const int pw = t - pw0;               // < 0 on the drone warps when they are exempt
  if (!FULL || !p.cattle_integrate) {
    if (pw >= 0) for (int it = pw; it < nact * M; it += pnw) {
      int slot, i;
      split_item(it, nact, nmagic, slot, i);
      const int g = s.flist[slot], e = e0 + g;
      const auto ua = s.u[i * G + g];
      const auto v = herd_finish<T>(s.dpos, s.n[g], g, G, s.cpos[i * G + g], s.cvel[i * G + g], ua.x, ua.y);
      a.cvel[(size_t)i * E + e] = v;       // positions are not advanced on the DYN path
    }
  }
  CHS_STAMP(21);
  if (pw >= 0 && pw < M * G) {
    // winding numbers (evaluate_herding_effectiveness).  The thread stride pnw is a multiple of 32, hence of G: a thread sees
    // the same env in every round, so it takes the bounding box of that env's drone polygon once and only walks the polygon
    // for cows inside the box -- a point outside it has winding number zero.  A non-finite vertex disables the shortcut.
    const int g = pw & (G - 1);
    const int n = s.n[g];
    if (n > 0) {
      T x0 = Mth<T>::inf(), y0 = Mth<T>::inf(), x1 = -Mth<T>::inf(), y1 = -Mth<T>::inf();
      bool finite = true;
#pragma unroll 1
      for (int i = 0; i < n; ++i) {
        const auto P = s.dpos[i * G + g];
        x0 = Mth<T>::min(x0, P.x); x1 = Mth<T>::max(x1, P.x);
        y0 = Mth<T>::min(y0, P.y); y1 = Mth<T>::max(y1, P.y);
        finite = finite && (P.x == P.x) && (P.y == P.y);
      }
      if (!finite) { x0 = y0 = -Mth<T>::inf(); x1 = y1 = Mth<T>::inf(); }
      for (int it = pw; it < M * G; it += pnw) {
        const auto c = s.cpos[it];
        const bool outside = (c.x < x0) || (c.x > x1) || (c.y < y0) || (c.y > y1);
        if (!outside && cow_inside<T>(s.dpos, n, g, G, c.x, c.y)) atomicAdd(&s.herd[g], 1);
      }
    }
  }
  CHS_STAMP(5);
  __syncthreads();
  CHS_STAMP(6);
  // The live rows leave as soon as they are complete, issued by the last warp while warp 0 works through phase 3.
  // In-kernel auto-reset rewrites the obs slot of a finished env with ordinary stores after the next barrier, so
  // there the issuing lane waits until its bulk write has landed (the wait overlaps phase 3 as well).
  if (obs_lane && !p.late_rows) {
    fence_async_smem();
    bulk_store_live(p, s.stage + (size_t)ow * p.stage_stride, a.obs + (size_t)(e0 + ow) * p.R * p.L);
    bulk_commit();
  }

  // ---------------- phase 3 ----------------
  // Hand-off: warp 0 releases the other warps into the rebuild of the finished envs as soon as it knows which envs finished
  // (bar.arrive, below), and goes on with the statistics, the env scalars and the terminal observations while they work:
  // the rebuild is the tail of every CTA that holds a finished env, and the part of phase 3 behind the flags is a third of it.
  const bool handoff = !SPLIT && (!FULL || p.bulk_ok) && nt >= 64;
  if (t < 32) {
    bool fin = false;
    const int e = e0 + t;
    const bool active = (t < G && e < Ev);
    // values that cross the warp-converged statistics block below
    int4 ej = pre_j;
    double2 ef = pre_f;
    const int4 ei = pre_i;
    bool st_term = false, st_trunc = false, st_bad = false, done_env = false;
    int st_cause = -1, sc = 0, scA = 0, nn = 0, row = 0;
    int st_herded = 0, st_succ = 0, st_lvup = 0;      // telemetry: cows inside the drone polygon, curriculum successes / level-ups
    double epret = 0, st_path = 0;
    if (active) {
      const int n = s.n[t];
      int level = ej.x;
      const int level_in = level;
      // sums in the reference's order: drone by drone, nearest neighbour first (CattleAviary.py:231-247).
      // The centroid distance, the approach term and the spacing clock are float64 in both builds: the approach
      // term amplifies centroid-distance differences by 5*ctrl_freq/2.5 and the clock counts in 1/240.
      T ssum = 0, csum = 0, c2d = 0, ms = Mth<T>::inf();
      double dcx = 0, dcy = 0;
      bool alt_bad = false, isolated = false;
      CHS_STAMP(13);
#pragma unroll 1
      for (int d = 0; d < n; ++d) {
        const auto A = s.redA[d * G + t];
        const auto B = s.redB[d * G + t];
        const auto P = s.dpos[d * G + t];
        csum += A.z; ssum += A.x; csum += A.w; ssum += A.y;
        c2d += B.x;
        ms = Mth<T>::min(ms, B.y);
        isolated = isolated || (B.y > T(k::MAX_FORMATION_DISTANCE));
        alt_bad = alt_bad || (B.z != T(0));
        dcx += (double)P.x; dcy += (double)P.y;
        st_path += (double)P.w;
      }
      CHS_STAMP(14);
      const double inv_n = s.ctab[CHS_TAB_INV_N + n];
      const T inv_nT = T(inv_n);
      const double2 cc = s.ccent[t];
      const double ddx = dcx * inv_n - cc.x, ddy = dcy * inv_n - cc.y;
      const double cent = sqrt(ddx * ddx + ddy * ddy + 0.0);
      CHS_STAMP(7);
      const int herded = s.herd[t];
      // effectiveness in percent with the reference's own arithmetic (evaluation.py:136) where a success test compares
      // it with a threshold (levels 4-6; a level-3 env can move up inside this step); the reward uses the fraction
      const double eff = (M > 0 && level_in >= 3 && level_in <= 6) ? (double)herded / (double)M * 100.0 : 0.0;
      const double eff_frac = (double)herded * s.ctab[CHS_TAB_INV_M];
      const double msd = (double)ms;
      double r_global = 0;
      bool term = false, trunc = false;
      int cause = -1, sc_inc = p.K;
      if constexpr (!MARL) {
        double approach = 0;
        if (ef.x == ef.x) {   // prev_cent_dists is not None
          const double change = ef.x - cent;
          approach = clipv<double>(change * s.ctab[CHS_TAB_APPROACH], -1.0, 1.0);
        }
        ef.x = cent;
        csum *= T(0.5) * inv_nT; ssum *= T(0.5) * inv_nT; c2d *= inv_nT;
        r_global = (double)ssum * s.ctab[level * k::C_NCOLS + k::C_W_SIMPLE] + (double)csum * s.ctab[level * k::C_NCOLS + k::C_W_COMPLEX] +
                   0.1 * s.ctab[level * k::C_NCOLS + k::C_W_SURV] + approach * s.ctab[level * k::C_NCOLS + k::C_W_APPROACH] +
                   eff_frac * s.ctab[level * k::C_NCOLS + k::C_W_EFF] + (double)c2d * s.ctab[level * k::C_NCOLS + k::C_W_C2D];
        // first _computeTerminated call, inside _computeReward (:314): only its side effects survive
        const bool term1 = compute_terminated<double>(s.ctab, level, msd, cent, eff, ef.y, 1.0 / 240.0);
        if (term1) {   // CurriculumLearning.evaluate_curriculum_results (curriculum_learning.py:200-222)
          ej.y += 1; st_succ = 1;
          if (ej.y >= (int)s.ctab[level_in * k::C_NCOLS + k::C_REQ_TALLY]) {
            ej.y = 0; st_lvup = 1;
            level = (level + 1 >= k::NUM_LEVELS) ? k::NUM_LEVELS - 1 : level + 1;
          }
        }
        // second call, the one step() returns (BaseAviary.py:458)
        term = compute_terminated<double>(s.ctab, level, msd, cent, eff, ef.y, 1.0 / 240.0);
        // _computeTruncated (:497-552), first hit in the reference's order
        if (alt_bad) cause = 0;
        else if (msd < k::COLLISION_THRESHOLD) cause = 1;
        else if (isolated) cause = 2;
        else if (cent > k::MISSION_BOUNDARY) cause = 3;
        else if (ei.x > p.trunc_step_limit) cause = 4;
        trunc = cause >= 0;
        a.reward[e] = (float)r_global;
        a.terminated[e] = term ? 1 : 0;
        a.truncated[e] = trunc ? 1 : 0;
      } else {
        // rllib_envs/BaseAviary.step (:425-430): {i: _computeReward(i)}, then {i: _computeTerminated(i)}, then
        // {i: _computeTruncated(i)}, each dict built drone by drone.  The reward call is stateful
        // (MARLCattleAviary.py:110-178): prev_cent_dists is refreshed by the first drone, so later drones see a
        // zero approach term; every _computeTerminated call advances the spacing clock by 1/CTRL_FREQ; a success
        // bumps the curriculum tally -- and possibly the level -- before the next drone is rewarded.
        sc_inc = 1;                                                // `step_counter += 1` (rllib_envs/BaseAviary.py:437)
        const double clk = 1.0 / (double)p.ctrl_freq;
        const bool shared_trunc = (cent > k::MISSION_BOUNDARY) || (ei.x > p.trunc_step_limit);
        float* rw = a.reward + (size_t)e * D;
        uint8_t* tf = a.terminated + (size_t)e * (D + 1);
        uint8_t* uf = a.truncated + (size_t)e * (D + 1);
        bool all_term = true, all_trunc = true;
        // one _computeReward(d) call (MARLCattleAviary.py:110-178), state effects included
        auto agent_reward = [&](int d, bool& own_trunc) -> double {
          const auto A = s.redA[d * G + t];
          const auto B = s.redB[d * G + t];
          T s1 = A.x, s2 = A.y, c1 = A.z, c2 = A.w;
          if (level != level_in) {        // the level moved inside this step: phase 2 used the old targets
            const T desired = curr<T>(s.ctab, level, k::C_DESIRED), tolf = curr<T>(s.ctab, level, k::C_TOL);
            s1 = simple_spacing<T>(B.y, desired, tolf); s2 = simple_spacing<T>(B.w, desired, tolf);
            c1 = complex_spacing<T>(B.y, desired); c2 = complex_spacing<T>(B.w, desired);
          }
          double r = (double)((s1 + s2) * T(0.5)) * s.ctab[level * k::C_NCOLS + k::C_W_SIMPLE] +
                     (double)((c1 + c2) * T(0.5)) * s.ctab[level * k::C_NCOLS + k::C_W_COMPLEX] + 0.1 * s.ctab[level * k::C_NCOLS + k::C_W_SURV];
          double approach = 0;
          if (ef.x == ef.x) approach = clipv<double>((ef.x - cent) * s.ctab[CHS_TAB_APPROACH], -1.0, 1.0);
          ef.x = cent;
          r += approach * s.ctab[level * k::C_NCOLS + k::C_W_APPROACH];
          r += eff_frac * s.ctab[level * k::C_NCOLS + k::C_W_EFF];
          r += (double)B.x * s.ctab[level * k::C_NCOLS + k::C_W_C2D];
          own_trunc = (B.z != T(0)) || ((double)B.y < k::COLLISION_THRESHOLD) ||
                      (B.y > T(k::MAX_FORMATION_DISTANCE)) || shared_trunc;
          if (compute_terminated<double>(s.ctab, level, msd, cent, eff, ef.y, clk)) {
            const auto P = s.dpos[d * G + t];
            const double hx = cc.x - (double)P.x, hy = cc.y - (double)P.y;
            r += marl_end_of_episode(s.ctab, level, n, (double)B.y, (double)B.w, cent, eff, sqrt(hx * hx + hy * hy));
            ej.y += 1; st_succ += 1;
            if (ej.y >= (int)s.ctab[level * k::C_NCOLS + k::C_REQ_TALLY]) {
              ej.y = 0; st_lvup += 1;
              level = (level + 1 >= k::NUM_LEVELS) ? k::NUM_LEVELS - 1 : level + 1;
            }
          } else if (own_trunc) {
            r -= 50.0;
          }
          return r;
        };
        for (int d = 0; d < n; ++d) {
          bool own_trunc;
          const double r = agent_reward(d, own_trunc);
          rw[d] = (float)r;
          r_global += r;
          uf[d] = own_trunc ? 1 : 0;
          all_trunc = all_trunc && own_trunc;
          if (d == 0) {
            const auto B = s.redB[t];
            cause = (B.z != T(0)) ? 0 : ((double)B.y < k::COLLISION_THRESHOLD) ? 1
                    : (B.y > T(k::MAX_FORMATION_DISTANCE)) ? 2 : (cent > k::MISSION_BOUNDARY) ? 3 : 4;
          }
        }
        for (int d = 0; d < n; ++d) {
          const bool td = compute_terminated<double>(s.ctab, level, msd, cent, eff, ef.y, clk);
          tf[d] = td ? 1 : 0;
          all_term = all_term && td;
        }
        for (int d = n; d < D; ++d) { rw[d] = 0.f; tf[d] = 0; uf[d] = 0; }
        term = all_term; trunc = all_trunc;          // the "__all__" entries (:428,430)
        if (FULL && p.wrapper_quirks) {
          // Opt-in bug compatibility with the reference's RLlib wrapper (rllib_envs/marl_wrapper.py:104-110): after
          // env.step it DISCARDS the step's outputs and calls the stateful _computeReward / _computeTerminated /
          // _computeTruncated once more for every agent still in its `agents` list, interleaved per agent -- the approach
          // term is then always 0 (prev_cent_dists was refreshed by the step), the spacing clock advances two more ticks
          // per agent and a success bumps the tally again.  Agents whose second-pass `terminated` is set leave the list
          // (:113); both "__all__" flags mean "the list is empty" (:116-117).  The list lives in envj.w as a bit mask and
          // is rebuilt on the first step of an episode (the wrapper's reset, :68-72).
          unsigned act = (ei.y == 0) ? 0xffffffffu : (unsigned)ej.w;
          act &= (n >= 32) ? 0xffffffffu : ((1u << n) - 1u);
          for (int d = 0; d < n; ++d) {
            if (!((act >> d) & 1u)) { rw[d] = 0.f; tf[d] = 0; uf[d] = 0; continue; }
            bool own_trunc;
            rw[d] = (float)agent_reward(d, own_trunc);
            const bool td = compute_terminated<double>(s.ctab, level, msd, cent, eff, ef.y, clk);
            tf[d] = td ? 1 : 0; uf[d] = own_trunc ? 1 : 0;
            if (td) act &= ~(1u << d);
          }
          ej.w = (int)act;
          term = trunc = (act == 0u);
        }
        tf[D] = term ? 1 : 0; uf[D] = trunc ? 1 : 0;
        if (!trunc) cause = -1;
      }
      ej.x = level;
      CHS_STAMP(8);
      if (a.dbg_herded) a.dbg_herded[e] = herded;
      st_herded = herded;
      epret = pre_ret + r_global;
      sc = ei.x + sc_inc; scA = ei.y + 1; nn = n; row = ei.w;
      fin = term || trunc;
      st_term = term; st_trunc = trunc; st_cause = cause;
      st_bad = !(r_global == r_global) || fabs(r_global) == Mth<double>::inf();
      done_env = fin && p.auto_reset;          // the observation of this step goes to term_obs, a fresh one to obs
    }
    // what the rebuild of a finished env needs (in-kernel rebuild only: in split mode reset_list_kernel does all of this)
    int new_n = 0, new_row = 0;
    const bool rebuild = active && done_env && !a.done_list;
    if (rebuild) {
      new_n = draw_num_drones(p, p.env_id_offset + e, ej.z + 1);
      new_row = ei.w + 1;                      // Cattle_Spawn_Index pre-increment with wrap (BaseAviary.py:600-603)
      if (new_row >= p.n_sims) new_row = 0;
      s.newn[t] = new_n; s.row[t] = new_row; s.epi[t] = ej.z + 1; s.done[t] = 1;
    }
    {
      const unsigned m = __ballot_sync(0xffffffffu, rebuild);
      if (rebuild) s.dlist[__popc(m & ((1u << t) - 1u))] = t;
      if (t == 0) s.misc[1] = __popc(m);
    }
    if (handoff) {
      __threadfence_block();
      asm volatile("bar.arrive 0, %0;" ::"r"(nt) : "memory");
    }
    // ---- episode statistics: fire-and-forget reductions (RED, no return value, nothing waits on them) into the row of the
    // [CHS_STAT_ROWS][CHS_NUM_STATS] accumulator picked by the CTA index, so that the thousands of envs that finish in one
    // step of a large batch do not queue on sixteen addresses; chs_stats sums the rows.
    if (active) {
      double* srow = a.stats + (size_t)(blockIdx.x & (CHS_STAT_ROWS - 1)) * CHS_NUM_STATS;
      if (st_bad) atomicAdd(srow + 9, 1.0);
      if (fin) {
        atomicAdd(srow + 0, 1.0);
        atomicAdd(srow + 1, epret);
        atomicAdd(srow + 2, (double)scA);
        if (st_term) atomicAdd(srow + 3, 1.0);
        if (st_trunc && st_cause >= 0) atomicAdd(srow + 4 + st_cause, 1.0);
        // evaluation telemetry of the finished episode (update_evaluation_metrics / evaluation_episode_trigger,
        // BaseAviary.py:1406-1450): 1.7 x path length summed over its drones, effectiveness in percent at its last step
        atomicAdd(srow + 11, st_path);
        if (st_herded) atomicAdd(srow + 12, (double)st_herded * (100.0 * s.ctab[CHS_TAB_INV_M]));
      }
      // curriculum_success_tally increments and level-ups (curriculum_learning.py:200-222), summed over envs
      if (st_succ) atomicAdd(srow + 14, (double)st_succ);
      if (st_lvup) atomicAdd(srow + 15, (double)st_lvup);
    }
    {
      // per-step effectiveness (effectiveness_per_step of the reference's log), summed over the CTA's envs: one reduction
      // and one RED per CTA
      const int hsum = __reduce_add_sync(0xffffffffu, st_herded);
      if (t == 0 && hsum) atomicAdd(a.stats + (size_t)(blockIdx.x & (CHS_STAT_ROWS - 1)) * CHS_NUM_STATS + 13, (double)hsum * (100.0 * s.ctab[CHS_TAB_INV_M]));
    }
    {
      const unsigned full = 0xffffffffu, lt = (1u << t) - 1u;
      const unsigned dm = __ballot_sync(full, done_env);
      if (dm) {
        const int lead = __ffs(dm) - 1, cnt = __popc(dm), off = __popc(dm & lt);
        if (a.term_mode == 2) {
          int base = 0;
          if (t == lead) base = atomicAdd(a.term_count, cnt);
          base = __shfl_sync(full, base, lead);
          if (done_env) { a.term_idx[base + off] = e; s.termrow[t] = base + off; }
        } else if (done_env) {
          s.termrow[t] = 0;
        }
        if (done_env) s.done[t] = 1;
        if (a.done_list) {
          // deferred reset (split mode): finished envs are only listed; reset_list_kernel, launched right after
          // this kernel, advances their episode counters and rebuilds them: CTAs here do not wait on the reset path
          int base = 0;
          if (t == lead) base = atomicAdd(a.done_count, cnt);
          base = __shfl_sync(full, base, lead);
          if (done_env) a.done_list[base + off] = e;
        }
      }
    }
    CHS_STAMP(22);
    if (active) {
      if (a.done_list) {
        fin = false;
      } else if (done_env) {
        ej.z += 1;
        nn = new_n; row = new_row;
        sc = 0; scA = 0; epret = 0;
      } else {
        fin = false;
      }
      a.envi[e] = make_int4(sc, scA, nn, row);
      a.envj[e] = ej;
      a.envf[e] = ef;
      a.epret[e] = epret;
      CHS_STAMP(23);
      if ((!FULL || p.bulk_ok) && done_env && a.term_obs) {
        // terminal observation of a finished env: the whole image (staged rows + zero padding) to its term_obs row
        const size_t RL = (size_t)p.R * p.L;
        fence_async_smem();
        float* trow = a.term_obs + (a.term_mode == 2 ? (size_t)s.termrow[t] : (size_t)e) * RL;
        if (p.pad_persist || p.dead_hi > p.dead_lo) {
          bulk_store_live(p, s.stage + (size_t)t * p.stage_stride, trow);
          if (!p.pad_persist && p.zero_bytes > 0) bulk_store(reinterpret_cast<char*>(trow) + p.stage_copy * 4, s.zeros, (unsigned)p.zero_bytes);
        } else {
          bulk_copy_rows(p, s.stage + (size_t)t * p.stage_stride, s.zeros, trow);
        }
        bulk_commit();
      }
    }
    if (t == 0 && blockIdx.x == 0) { atomicAdd(a.stats + 10, (double)E); if (a.term_count_next) *a.term_count_next = 0; }
  }
  CHS_STAMP(9);
  if (!handoff) __syncthreads();
  else if (t >= 32) asm volatile("bar.sync 0, %0;" ::"r"(nt) : "memory");      // warp 0 arrived at the hand-off
  CHS_STAMP(10);
  // Pinned-host destination (`late_rows`): the live rows leave only now, and not at all for the envs that finished -- their
  // slot receives the first observation of the next episode from the rebuild, and every byte saved is a byte less on PCIe
  // (7 % of the envs finish per step under random actions).
  if (obs_lane && p.late_rows && !s.done[ow]) {
    fence_async_smem();
    bulk_store_live(p, s.stage + (size_t)ow * p.stage_stride, a.obs + (size_t)(e0 + ow) * p.R * p.L);
    bulk_commit();
  }
  // Host path without a result copy (chs_step_host, pinned buffers): the CTA that passes here last hands the number of
  // finished envs to the caller's pinned counter -- every CTA's additions to term_count precede its ticket.
  if (a.host_count && t == 0) {
    __threadfence();
    if (atomicAdd(a.cta_done, 1) == (int)gridDim.x - 1) {
      *a.host_count = *reinterpret_cast<volatile int*>(a.term_count);
      *a.cta_done = 0;
    }
  }

  // ---------------- phase 4 ----------------
  const int nd = s.misc[1];
  if (FULL && !p.bulk_ok) {
    const int warp = t >> 5, lane = t & 31, nwarp = nt >> 5;
    const size_t RL = (size_t)p.R * p.L;
    for (int g = warp; g < G; g += nwarp) {
      const int e = e0 + g;
      if (e >= Ev) break;
      float* dst = a.obs ? a.obs + (size_t)e * RL : nullptr;
      if (s.done[g]) {
        dst = nullptr;
        if (a.term_obs) dst = a.term_obs + (a.term_mode == 2 ? (size_t)s.termrow[g] : (size_t)e) * RL;
      }
      if (dst) copy_rows(p, s.stage + (size_t)g * p.stage_stride, dst, lane);
    }
  }
  CHS_STAMP(11);
  // ---------------- phase 5 ----------------
  if (nd > 0) {
    // (finished env, part) items over the warps, straight to global memory: no shared-memory staging.  The obs slot of a
    // finished env already received this step's live rows through a bulk store of the last warp; that write has to
    // land before the ordinary stores of the new episode's first observation (CTAs without a finished env skip this).
    const int warp = t >> 5, lane = t & 31, nwarp = nt >> 5;
    if (!handoff) {
      if (obs_lane) bulk_wait_all();
      __syncthreads();
      for (int i = warp; i < 3 * nd; i += nwarp) {
        const int g = s.dlist[i / 3], part = 1 << (i % 3);
        PidPre<T> pid;
        if (part == RESET_DRONES) pid = reset_preload<T>(a, e0 + g, lane);
        reset_env_warp<T>(a, e0 + g, s.newn[g], s.row[g], s.epi[g], lane, part, pid);
      }
    } else {
      // Observation parts first: to the warps from the top down, never to warp 0 -- they follow the wait for the early write
      // of the same slot, which is a barrier among warps 1.. only (warp 0 is still busy with the rest of phase 3).  Cow and
      // drone parts: item j to warp j + 1, warp 0 last.  With one finished env the three parts run on three warps at once.
      if (warp > 0) {
        if (obs_lane) bulk_wait_all();
        asm volatile("bar.sync 2, %0;" ::"r"(nt - 32) : "memory");
        for (int k = nwarp - 1 - warp; k < nd; k += nwarp - 1) {
          const int g = s.dlist[k];
          PidPre<T> pid;
          pid.e = pid.f = pid.g = mk4<T>(0, 0, 0, 0);
          reset_env_warp<T>(a, e0 + g, s.newn[g], s.row[g], s.epi[g], lane, RESET_OBS, pid);
        }
      }
      for (int j = (warp + nwarp - 1) % nwarp; j < 2 * nd; j += nwarp) {
        const int g = s.dlist[j >> 1];
        PidPre<T> pid;
        pid.e = pid.f = pid.g = mk4<T>(0, 0, 0, 0);
        if (j & 1) {
          pid = reset_preload<T>(a, e0 + g, lane);
          reset_env_warp<T>(a, e0 + g, s.newn[g], s.row[g], s.epi[g], lane, RESET_DRONES, pid);
        } else {
          reset_env_warp<T>(a, e0 + g, s.newn[g], s.row[g], s.epi[g], lane, RESET_COWS, pid);
        }
      }
    }
  }
  if (!FULL || p.bulk_ok) bulk_wait_read();       // shared memory must outlive the pending bulk reads
  CHS_STAMP(12);
  CHS_GSTAMP(31);
#undef CHS_GSTAMP
#undef CHS_STAMP
}

}  // namespace chs
