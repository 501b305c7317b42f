// One translation unit per compiled chain.  Build with exactly one of
//   -DMSIK_TU_CHAIN=<name>     (a specialised chain from generated/chains.cuh)
//   -DMSIK_TU_GENERIC=<N>      (table-driven kernel for a joint bucket of N)
#include "msik_launch.cuh"

#define MSIK_CAT_(a, b) a##b
#define MSIK_CAT(a, b) MSIK_CAT_(a, b)
#define MSIK_STR_(a) #a
#define MSIK_STR(a) MSIK_STR_(a)

namespace msik {

#if defined(MSIK_TU_CHAIN)
using ChainT = MSIK_CAT(MsikChain_, MSIK_TU_CHAIN);
extern const ChainOps MSIK_CAT(kOps_, MSIK_TU_CHAIN) = {
    MSIK_STR(MSIK_TU_CHAIN),
    ChainT::N,
    false,
    ChainRange<ChainT>::mask,
    &MSIK_CAT(kMsikPre_, MSIK_TU_CHAIN)[0][0],
    MSIK_CAT(kMsikTypes_, MSIK_TU_CHAIN),
    &launch_delta<ChainT>,
    &launch_target_ik<ChainT>,
    &launch_delta_tma<ChainT>,
    &launch_fk<ChainT>,
    ChainDual<ChainT>::name,
    std::is_void<ChainDual<ChainT>::Partner>::value ? nullptr : &launch_delta_tma_dual<ChainT>,
};
#elif defined(MSIK_TU_GENERIC)
using ChainT = ChainGeneric<MSIK_TU_GENERIC>;
extern const ChainOps MSIK_CAT(kOps_generic, MSIK_TU_GENERIC) = {
    "generic" MSIK_STR(MSIK_TU_GENERIC),
    MSIK_TU_GENERIC,
    true,
    0u,
    nullptr,
    nullptr,
    &launch_delta<ChainT>,
    &launch_target_ik<ChainT>,
    &launch_delta_tma<ChainT>,
    &launch_fk<ChainT>,
    nullptr,
    nullptr,
};
#elif defined(MSIK_TU_GENERIC_REV)
using ChainT = ChainGenericRev<MSIK_TU_GENERIC_REV>;
extern const ChainOps MSIK_CAT(MSIK_CAT(kOps_generic, MSIK_TU_GENERIC_REV), r) = {
    "generic" MSIK_STR(MSIK_TU_GENERIC_REV) "r",
    MSIK_TU_GENERIC_REV,
    true,
    0u,
    nullptr,
    nullptr,
    &launch_delta<ChainT>,
    &launch_target_ik<ChainT>,
    &launch_delta_tma<ChainT>,
    &launch_fk<ChainT>,
    nullptr,
    nullptr,
};
#else
#error "define MSIK_TU_CHAIN, MSIK_TU_GENERIC or MSIK_TU_GENERIC_REV"
#endif

}  // namespace msik
