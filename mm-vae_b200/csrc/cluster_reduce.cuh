// Block + thread-block-cluster reductions shared by the K4/K5 kernels: warp shuffles,
// one shared-memory hop, then DSMEM reads of every peer CTA's slot.
#pragma once

#include "common.cuh"

#include <cooperative_groups.h>

namespace mmv {

namespace cg = cooperative_groups;

// Sum N per-thread values over the block, then over the cluster (DSMEM).
template <int N>
__device__ __forceinline__ void
cluster_sum(float (&v)[N], float *slot, float *scratch, cg::cluster_group &cluster, int cs)
{
    block_sum<N>(v, scratch);
    if (cs == 1) return;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) slot[k] = v[k];
    }
    cluster.sync();
#pragma unroll
    for (int k = 0; k < N; ++k) v[k] = 0.f;
    for (int r = 0; r < cs; ++r) {
        const float *remote = cluster.map_shared_rank(slot, r);
#pragma unroll
        for (int k = 0; k < N; ++k) v[k] += remote[k];
    }
}

__device__ __forceinline__ float
cluster_max(float v, float *slot, float *scratch, cg::cluster_group &cluster, int cs)
{
    v = block_max(v, scratch);
    if (cs == 1) return v;
    if (threadIdx.x == 0) slot[0] = v;
    cluster.sync();
    float m = -INFINITY;
    for (int r = 0; r < cs; ++r) m = fmaxf(m, cluster.map_shared_rank(slot, r)[0]);
    return m;
}

} // namespace mmv
