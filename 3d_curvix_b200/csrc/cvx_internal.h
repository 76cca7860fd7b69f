// cvx_internal.h -- what cvx_comm.cu needs from the context defined in cvx_api.cu (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/curvix_b200.h"

struct CvxResultsView {
    int64_t P; int V; int device;
    int view1, view2, run_begin, run_end;
    const int32_t *h1, *h2rep, *count, *nan, *sup;
    const double *gc, *gw, *tw, *h2c;
    cudaStream_t stream;
};
struct CvxCommState;
// runs the finalize kernel if needed and exposes the device-resident results of the last run
int cvx_internal_results(cvx_ctx* ctx, CvxResultsView* out);
int cvx_internal_device(cvx_ctx* ctx, int* device, cudaStream_t* stream);
CvxCommState** cvx_internal_comm_slot(cvx_ctx* ctx);
int cvx_internal_fail(cvx_ctx* ctx, int code, const char* msg);
// claimed-edge bitmap (allocated on demand)
int cvx_internal_claim_words(cvx_ctx* ctx, uint32_t** words, size_t* n_words);
void cvx_internal_count_launch(cvx_ctx* ctx, int n);
void cvx_comm_release(CvxCommState* s);   // called by cvx_destroy
