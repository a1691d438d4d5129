#include "cdx_internal.cuh"
void cdx_pregrasp_free(cdx_ctx* ctx) { (void)ctx; }
#define NOT_YET(name) { cdx_set_error(#name ": not implemented yet"); return CDX_ERR_STATE; }
extern "C" int cdx_hand_load(cdx_ctx*, int, const int*, const int*, const int*, const float*, const float*, const float*, int, const int*, const float*) NOT_YET(cdx_hand_load)
extern "C" int cdx_hand_fk(cdx_ctx*, const double*, const double*, int64_t, double*, void*) NOT_YET(cdx_hand_fk)
extern "C" int cdx_pregrasp_init(cdx_ctx*, const cdx_pregrasp_params*, int64_t, const double*, const double*, const double*, const double*, void*) NOT_YET(cdx_pregrasp_init)
extern "C" int cdx_pregrasp_run(cdx_ctx*, int, void*) NOT_YET(cdx_pregrasp_run)
extern "C" int cdx_pregrasp_eval(cdx_ctx*, double*, double*, double*, double*, double*, double*, void*) NOT_YET(cdx_pregrasp_eval)
extern "C" int cdx_pregrasp_results(cdx_ctx*, double*, double*, double*, double*, double*, double*, void*) NOT_YET(cdx_pregrasp_results)
extern "C" int cdx_pregrasp_state(cdx_ctx*, double*, double*, double*, double*, void*) NOT_YET(cdx_pregrasp_state)
extern "C" int cdx_pregrasp_argmin(cdx_ctx*, double*, void*) NOT_YET(cdx_pregrasp_argmin)
