// Host-side launch tables: one ChainOps per compiled chain (specialised or generic
// joint bucket).  Each chain lives in its own translation unit (msik_chain_tu.cu
// compiled once per chain with -DMSIK_TU_CHAIN=...) so the build parallelises.
#pragma once
#include <atomic>
#include <cstdio>
#include <cstdlib>

#include "msik_kernels.cuh"
#include "msik_tma.cuh"

namespace msik {

enum CtrlMode { CTRL_FULL = 0, CTRL_MASKED = 1, CTRL_RANGE = 2 };

struct ChainOps {
  const char* name;
  int n;               // joints the kernel is compiled for (bucket size for generic)
  bool generic;
  uint32_t range_mask; // ctrl_mask served by the CTRL_RANGE variant (0 = none)
  const float* pre;    // (n+1) x 12 table of a specialised chain, NULL for generic
  const int* types;
  cudaError_t (*delta)(const StepParams& P, int nc, bool gather, int mode, size_t smem, cudaStream_t s);
  cudaError_t (*target_ik)(const StepParams& P, int mode, size_t smem, cudaStream_t s);
  // persistent TMA-staged path A (identity rows, 16-byte aligned bases); smem from tma_smem_bytes()
  cudaError_t (*delta_tma)(const StepParams& P, int nc, int mode, size_t smem, cudaStream_t s);
  cudaError_t (*fk)(const StepParams& P, int n_real, cudaStream_t s);
  // dual launch: this chain's stock range + the stock range of `dual_partner` (another compiled chain of the same robot) in
  // one pass of the persistent kernel; NULL when there is no partner (msik_delta_step_dual then issues two launches)
  const char* dual_partner;
  cudaError_t (*delta_tma_dual)(const StepParams& P, int nc, size_t smem, cudaStream_t s);
};

extern std::atomic<long long> g_launches;

// contiguous controlled range served without masking (C0, NR); default: none
template <class Chain>
struct ChainRange {
  static constexpr uint32_t mask = 0;
  static constexpr int C0 = 0;
  static constexpr int NR = Chain::N;
};
// Fetch: the arm controller drives the last 7 of the 11 ancestors
// (agents/robots/fetch/fetch.py:77-85,100-113)
template <>
struct ChainRange<MsikChain_fetch_gripper> {
  static constexpr uint32_t mask = 0x7f0u;
  static constexpr int C0 = 4;
  static constexpr int NR = 7;
};

// Realman dual-arm mobile robot: each arm controller drives the last 6 of the 9 ancestors of its end link
// (3 base joints first; agents/robots/realman/mobile_realman.py:137-138,298-350)
template <>
struct ChainRange<MsikChain_realman_r_link6> {
  static constexpr uint32_t mask = 0x1f8u;
  static constexpr int C0 = 3;
  static constexpr int NR = 6;
};
template <>
struct ChainRange<MsikChain_realman_l_link6> {
  static constexpr uint32_t mask = 0x1f8u;
  static constexpr int C0 = 3;
  static constexpr int NR = 6;
};

// the other arm of a dual-arm robot: ChainDual<A>::Partner is solved together with A by msik_delta_step_dual
template <class Chain>
struct ChainDual {
  using Partner = void;
  static constexpr const char* name = nullptr;
};
template <>
struct ChainDual<MsikChain_realman_r_link6> {
  using Partner = MsikChain_realman_l_link6;
  static constexpr const char* name = "realman_l_link6";
};
template <>
struct ChainDual<MsikChain_realman_l_link6> {
  using Partner = MsikChain_realman_r_link6;
  static constexpr const char* name = "realman_r_link6";
};

template <class K>
inline cudaError_t launch_k(K kernel, const StepParams& P, size_t smem, cudaStream_t s) {
  if (smem > 40 * 1024) {  // opt in early: static shared memory counts towards the 48 KB default limit too
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int grid = (P.cfg.n_envs + TPB - 1) / TPB;
  (void)cudaGetLastError();
  kernel<<<grid, TPB, smem, s>>>(P);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError();
}

template <class Chain, int C0, int NR, bool MASKED>
inline cudaError_t delta_mode(const StepParams& P, int nc, bool gather, size_t smem, cudaStream_t s) {
  if (nc == 6) {
    if constexpr (NR >= 6) {
      return gather ? launch_k(delta_step_kernel<Chain, C0, NR, 6, MASKED, true>, P, 0, s)
                    : launch_k(delta_step_kernel<Chain, C0, NR, 6, MASKED, false>, P, smem, s);
    } else {
      return cudaErrorInvalidValue;
    }
  }
  if constexpr (NR >= 3) {
    return gather ? launch_k(delta_step_kernel<Chain, C0, NR, 3, MASKED, true>, P, 0, s)
                  : launch_k(delta_step_kernel<Chain, C0, NR, 3, MASKED, false>, P, smem, s);
  } else {
    return cudaErrorInvalidValue;
  }
}

template <class Chain>
cudaError_t launch_delta(const StepParams& P, int nc, bool gather, int mode, size_t smem, cudaStream_t s) {
  using RG = ChainRange<Chain>;
  if (mode == CTRL_FULL) return delta_mode<Chain, 0, Chain::N, false>(P, nc, gather, smem, s);
  if (mode == CTRL_RANGE) {
    if constexpr (RG::mask != 0) return delta_mode<Chain, RG::C0, RG::NR, false>(P, nc, gather, smem, s);
    else return cudaErrorInvalidValue;
  }
  return delta_mode<Chain, 0, Chain::N, true>(P, nc, gather, smem, s);
}

constexpr int kTmaMaxSmem = 96 * 1024;  // largest dynamic shared memory the TMA kernel is launched with

// persistent launch: one CTA per resident slot (SM count x occupancy), capped by the tile count
template <class K>
inline cudaError_t launch_persistent(K kernel, const StepParams& P, size_t smem, cudaStream_t s) {
  // K is only the function-pointer TYPE (shared by every instantiation of the kernel template), so the cache
  // is keyed by the kernel address as well: both the opt-in attribute and the occupancy belong to one function
  struct Cfg { const void* fn; size_t smem; int dev; int slots; int tpb; };
  const int tpb = 32 * (P.tma_warps >= 1 && P.tma_warps <= WARPS ? P.tma_warps : WARPS);
  static thread_local Cfg cache[32] = {};
  static thread_local int next = 0;
  const void* fn = reinterpret_cast<const void*>(kernel);
  int slots = 0, dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  for (auto& c : cache) if (c.slots && c.fn == fn && c.smem == smem && c.dev == dev && c.tpb == tpb) slots = c.slots;
  if (!slots) {
    // the attribute belongs to the FUNCTION (not to one launch): always raise it to the largest
    // size choose_tma() admits, never to the size of the current configuration
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaMaxSmem);
    if (e != cudaSuccess) return e;
    int sms = 0, occ = 0;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, tpb, smem)) != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    // experiment knob: fewer resident CTAs per SM than the occupancy limit (profiles/README.md, variant sweeps)
    if (const char* v = getenv("MSIK_TMA_CTAS_PER_SM")) { const int want = atoi(v); if (want >= 1 && want < occ) occ = want; }
    slots = sms * occ;
    cache[next] = Cfg{fn, smem, dev, slots, tpb};  // round-robin replacement
    next = (next + 1) % 32;
  }
  const int ntile = (P.cfg.n_envs + tpb - 1) / tpb;  // CTAs needed if every warp took one 32-env tile
  const int grid = ntile < 1 ? 1 : (ntile < slots ? ntile : slots);
  (void)cudaGetLastError();  // a stale error of an earlier, unrelated call must not be blamed on this launch
  // Programmatic dependent launch: the grid may be scheduled while the previous kernel of the stream
  // drains (its tail overlaps our barrier init / first bulk loads); the kernel executes
  // griddepcontrol.wait before its first global access, so memory ordering is unchanged.
  static const bool pdl = [] { const char* v = getenv("MSIK_PDL"); return !(v && *v == '0'); }();
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3(grid); lc.blockDim = dim3(tpb); lc.dynamicSmemBytes = smem; lc.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = at; lc.numAttrs = pdl ? 1 : 0;
  e = cudaLaunchKernelEx(&lc, kernel, P);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) fprintf(stderr, "msik: persistent launch failed: grid=%d block=%d smem=%zu slots=%d n_envs=%d\n", grid, tpb, smem, slots, P.cfg.n_envs);
  return e;
}

// which LEAN instantiation of delta_step_tma_kernel serves this step (0: the general one)
inline int tma_lean(const StepParams& P) {
  static const bool off = [] { const char* v = getenv("MSIK_TMA_LEAN"); return v && *v == '0'; }();
  if (off || (P.cfg.flags & MSIK_F_NO_DRIVE) || P.tq_out || P.sq_out || P.debug != 0) return 0;
  // 1 = exactly the stock arm + mimic-gripper dict: one or two sibling columns fed by ONE normalised action column, arm +
  // siblings covering the whole target row (the kernel then tests neither the column count, the flag nor a reload of the old
  // targets per tile); anything else (no sibling, raw columns, more columns,
  // a velocity controller) takes the general hoisted instantiation
  if (P.vel.n_cols == 0 && P.full_row && P.aux_uniform && P.aux.n_cols >= 1 && P.aux.n_cols <= 2 && (P.aux.flags[0] & MSIK_AUX_NORMALIZE)) return 1;
  return 2;
}

template <class Chain, int C0, int NR, bool MASKED>
inline cudaError_t delta_tma_mode(const StepParams& P, int nc, size_t smem, cudaStream_t s) {
  bool ident = true;  // chain joint i in column i
  for (int i = 0; i < Chain::N; ++i) ident = ident && P.cfg.anc_col[i] == i;
  const int lean = tma_lean(P);
  if (nc == 6) {
    if constexpr (NR >= 6) {
      if constexpr (!MASKED) {
        if (lean == 1)
          return ident ? launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 6, false, true, 1>, P, smem, s)
                       : launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 6, false, false, 1>, P, smem, s);
        if (lean == 2)  // stock Fetch dict: general sibling columns + the velocity controller
          return ident ? launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 6, false, true, 2>, P, smem, s)
                       : launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 6, false, false, 2>, P, smem, s);
      }
      return ident ? launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 6, MASKED, true>, P, smem, s)
                   : launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 6, MASKED, false>, P, smem, s);
    } else {
      return cudaErrorInvalidValue;
    }
  }
  if constexpr (NR >= 3) {
    if constexpr (!MASKED) {
      if (lean == 1)  // pd_ee_delta_pos in the stock stepping configuration
        return ident ? launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 3, false, true, 1>, P, smem, s)
                     : launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 3, false, false, 1>, P, smem, s);
      if (lean == 2)
        return ident ? launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 3, false, true, 2>, P, smem, s)
                     : launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 3, false, false, 2>, P, smem, s);
    }
    return ident ? launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 3, MASKED, true>, P, smem, s)
                 : launch_persistent(delta_step_tma_kernel<Chain, C0, NR, 3, MASKED, false>, P, smem, s);
  } else {
    return cudaErrorInvalidValue;
  }
}

// both arms on their stock (CTRL_RANGE) ranges, stock stepping configuration (tma_lean(P) != 0 checked by the caller)
template <class Chain>
cudaError_t launch_delta_tma_dual(const StepParams& P, int nc, size_t smem, cudaStream_t s) {
  using Partner = typename ChainDual<Chain>::Partner;
  if constexpr (std::is_void<Partner>::value) {
    return cudaErrorInvalidValue;
  } else {
    using RA = ChainRange<Chain>;
    using RB = ChainRange<Partner>;
    static_assert(RA::mask != 0 && RB::mask != 0, "dual partners are driven on their stock ranges");
    if (nc == 6) return launch_persistent(delta_step_tma_kernel<Chain, RA::C0, RA::NR, 6, false, false, 2, Partner, RB::C0, RB::NR>, P, smem, s);
    return launch_persistent(delta_step_tma_kernel<Chain, RA::C0, RA::NR, 3, false, false, 2, Partner, RB::C0, RB::NR>, P, smem, s);
  }
}

template <class Chain>
cudaError_t launch_delta_tma(const StepParams& P, int nc, int mode, size_t smem, cudaStream_t s) {
  using RG = ChainRange<Chain>;
  if (mode == CTRL_FULL) return delta_tma_mode<Chain, 0, Chain::N, false>(P, nc, smem, s);
  if (mode == CTRL_RANGE) {
    if constexpr (RG::mask != 0) return delta_tma_mode<Chain, RG::C0, RG::NR, false>(P, nc, smem, s);
    else return cudaErrorInvalidValue;
  }
  return delta_tma_mode<Chain, 0, Chain::N, true>(P, nc, smem, s);
}

template <class Chain, int C0, int NR, bool MASKED>
inline cudaError_t ik_mode(const StepParams& P, size_t smem, cudaStream_t s) {
  if constexpr (NR >= 6) {
    // SORT = true: the instantiation that regroups a CTA's envs by remaining work after the first iteration (large batches;
    // decided by the host, msik_target_ik).  Unmasked chains only: the other instantiation serves everything else unchanged.
    if constexpr (!MASKED) {
      if (P.ik_sort) return launch_k(target_ik_kernel<Chain, C0, NR, false, true>, P, smem, s);
    }
    return launch_k(target_ik_kernel<Chain, C0, NR, MASKED, false>, P, smem, s);
  } else {
    return cudaErrorInvalidValue;
  }
}

// P.na / P.gather are run-time (they touch only the prologue / epilogue); smem = 0 in gather mode
template <class Chain>
cudaError_t launch_target_ik(const StepParams& P, int mode, size_t smem, cudaStream_t s) {
  using RG = ChainRange<Chain>;
  if (mode == CTRL_FULL) return ik_mode<Chain, 0, Chain::N, false>(P, smem, s);
  if (mode == CTRL_RANGE) {
    if constexpr (RG::mask != 0) return ik_mode<Chain, RG::C0, RG::NR, false>(P, smem, s);
    else return cudaErrorInvalidValue;
  }
  return ik_mode<Chain, 0, Chain::N, true>(P, smem, s);
}

template <class Chain>
cudaError_t launch_fk(const StepParams& P, int n_real, cudaStream_t s) {
  const int grid = (P.cfg.n_envs + TPB - 1) / TPB;
  fk_kernel<Chain><<<grid, TPB, 0, s>>>(P, n_real);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError();
}

}  // namespace msik
