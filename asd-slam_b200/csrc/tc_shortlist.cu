// placeholder until the tcgen05 kernel lands
#include "internal.cuh"
cudaError_t asdm_tc_shortlist_launch(asdm_ctx*, const PairDesc*, const void*, const TcWork*, int, float, TcPair*,
                                     int32_t*, int, int32_t*) { return cudaErrorNotSupported; }
int asdm_tc_build_tmap(asdm_ctx*, asdm_set*) { return ASDM_ERR_INTERNAL; }
