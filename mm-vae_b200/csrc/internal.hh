// Definitions of the opaque C-ABI handles (not installed; include/mmvae_b200.h
// is the public face).
#pragma once

#include "../../include/mmvae_b200.h"
#include "kernels.hh"

#include <cuda_runtime.h>
#include <string>
#include <vector>

struct mmv_ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaStream_t copy_stream = nullptr;   // mmv_csr_update copies here, so minibatch k+1 uploads while step k runs
    // Lowest-priority stream for work of a backward pass that only the optimiser consumes (the per-gene sums of dlogits,
    // the hidden layers' weight-gradient GEMMs): it leaves the step's dependency chain and fills what the chain's
    // kernels leave free (model.cu: side_fork / side_join)
    cudaStream_t side_stream = nullptr;
    cudaEvent_t ev_side_fork = nullptr, ev_side_join = nullptr;
    bool side_busy = false;
    // NCCL data-parallel state (comm.cu)
    void *nccl_comm = nullptr;
    cudaStream_t comm_stream = nullptr;          // early gradient buckets are reduced here, under the rest of backward
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int nranks = 1, rank = 0;
    bool in_group = false;
    void *comm_stage = nullptr;                  // bf16 staging of a gradient bucket reduced in bf16 (comm.cu)
    size_t comm_stage_bytes = 0;
    float *comm_ss = nullptr;                    // per-block sums of squares left by the unpack kernels of this step's bf16 buckets
    int comm_ss_used = 0, comm_ss_cap = 0;
    // MMV_COMM_TRACE=<k>: timed marks on both streams for the k-th optimiser step, printed by rank 0 (comm.cu)
    int trace_step = -1, trace_seen = 0;
    bool trace_on = false;
    struct mark_t { std::string name; cudaEvent_t ev; int comm; };
    std::vector<mark_t> marks;
};

struct mmv_csr {
    mmv_ctx *ctx = nullptr;
    int64_t n_cells = 0, n_genes = 0, nnz = 0, c_dim = 0;
    int64_t cap_cells = 0, cap_nnz = 0, max_row_nnz = 0;
    int64_t *indptr = nullptr;
    int32_t *indices = nullptr;
    float *vals = nullptr;
    float *cov = nullptr;      // [n_cells x c_dim]; nullptr / !has_cov = all ones
    bool has_cov = false;
    bool borrowed = false;     // device arrays owned by the caller (mmv_csr_wrap_device)
    size_t h2d_bytes = 0;      // bytes copied by the last update (bench accounting)
    cudaEvent_t ev_ready = nullptr;     // streamed slot: recorded on the copy stream after an update
    cudaEvent_t ev_consumed = nullptr;  // recorded on the ctx stream after the last step that read the slot
    int *d_unsorted = nullptr; // device flag: 1 if some cell's gene ids are not strictly increasing (set on the
                               // ctx stream by every upload / update / wrap; kernels that bisect rows read it)
};

namespace mmv {

// comm.cu: sum-reduce over ranks on the ctx stream (no-ops when nranks == 1)
int comm_allreduce_f32(mmv_ctx *ctx, float *buf, size_t n);
int comm_allreduce_f64(mmv_ctx *ctx, double *buf, size_t n);
// bucketed overlap: reduce buf[0, n) on the comm stream once everything queued so far on the ctx stream
// has finished; comm_join makes the ctx stream wait for it
// as_bf16: the bucket crosses NVLink as bf16 (cast on the comm stream, ncclAllReduce(bf16), cast back) — half the wire
// bytes for gradients that are products of bf16 operands anyway (tensor-core mode; see comm.cu)
// ss_partial / ss_count (optional, bf16 buckets only): the unpack pass leaves sum(g^2) of the REDUCED bucket as *ss_count
// per-block partials at *ss_partial (fixed order, identical on every rank) — the clip norm needs no second pass over it
int comm_allreduce_f32_async(mmv_ctx *ctx, float *buf, size_t n, bool as_bf16 = false, const float **ss_partial = nullptr,
                             int *ss_count = nullptr);
int comm_join(mmv_ctx *ctx);
// reductions queued between begin and end go out as one NCCL launch on the comm stream
int comm_group_begin(mmv_ctx *ctx);
int comm_allreduce_f64_async(mmv_ctx *ctx, double *buf, size_t n);
int comm_group_end(mmv_ctx *ctx);
// gradient buffer: a registered symmetric NCCL window when data-parallel (see comm.cu)
int comm_alloc_grads(mmv_ctx *ctx, size_t bytes, float **ptr, void **win, bool *nccl_owned);
void comm_free_grads(mmv_ctx *ctx, float *ptr, void *win, bool nccl_owned);
// timeline marks (no-ops unless this step is traced)
void trace_begin(mmv_ctx *ctx);
void trace_mark(mmv_ctx *ctx, const char *name, bool on_comm = false);
void trace_end(mmv_ctx *ctx);

} // namespace mmv
