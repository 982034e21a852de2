// bm.cu -- StereoBM (BASELINE config 2).  Placeholder until the kernels land: every entry
// point fails loudly (no CPU fallback).
#include "svb_common.cuh"
struct svb_bm { svb_bm_params prm; int launches; };
extern "C" int svb_bm_create(const svb_bm_params*, svb_bm_t**) { svb_set_error("StereoBM kernels not built yet"); return SVB_ERR_UNSUPPORTED; }
extern "C" void svb_bm_destroy(svb_bm_t*) {}
extern "C" int svb_bm_compute_dev(svb_bm_t*, const uint8_t*, const uint8_t*, int, int, int, int16_t*, void*) { svb_set_error("StereoBM kernels not built yet"); return SVB_ERR_UNSUPPORTED; }
extern "C" int svb_bm_compute_host(svb_bm_t*, const uint8_t*, const uint8_t*, int, int, int, int16_t*) { svb_set_error("StereoBM kernels not built yet"); return SVB_ERR_UNSUPPORTED; }
extern "C" int svb_bm_last_launches(const svb_bm_t*) { return 0; }
