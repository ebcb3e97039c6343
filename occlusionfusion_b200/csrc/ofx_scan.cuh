// ofx_scan.cuh -- exclusive prefix sum of ints on the device, shared by the graph builder (ofx_graph_edges.cu), the
// skin-cache set-up (ofx_skin_cache.cu) and marching cubes (ofx_mc.cu).  Two levels: 2048 elements per block, the block
// sums by one block, one add pass.  Kernels are static: every translation unit that includes this gets its own copy.
#pragma once
#include "ofx_common.h"

namespace ofx {

// ---- exclusive scan of ints (two levels: 2048 elements per block, block sums by one block) ------------------------
constexpr int SCAN_BLOCK = 1024, SCAN_ITEMS = 2;

static __global__ void __launch_bounds__(SCAN_BLOCK) scan_block_kernel(const int *__restrict__ in, int n, int *__restrict__ out,
                                                                int *__restrict__ block_sum)
{
    __shared__ int s_warp[32];
    const int base = blockIdx.x * SCAN_BLOCK * SCAN_ITEMS + threadIdx.x * SCAN_ITEMS;
    int x[SCAN_ITEMS], tot = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { x[k] = base + k < n ? in[base + k] : 0; tot += x[k]; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = tot;
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int t = s_warp[lane];
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += y; }
        s_warp[lane] = t;
    }
    __syncthreads();
    int run = (wid ? s_warp[wid - 1] : 0) + incl - tot;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { if (base + k < n) out[base + k] = run; run += x[k]; }
    if (threadIdx.x == SCAN_BLOCK - 1) block_sum[blockIdx.x] = run;
}

// one block: exclusive scan of the block sums in place (any count), total -> total_out
static __global__ void __launch_bounds__(1024) scan_sums_kernel(int *__restrict__ sums, int nb, int *__restrict__ total_out)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int b0 = 0; b0 < nb; b0 += 1024) {
        const int i = b0 + threadIdx.x;
        const int x = i < nb ? sums[i] : 0;
        int incl = x;
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int t = s_warp[lane];
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += y; }
            s_warp[lane] = t;
        }
        __syncthreads();
        const int excl = s_carry + (wid ? s_warp[wid - 1] : 0) + incl - x;
        if (i < nb) sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = s_carry;
}

static __global__ void scan_add_kernel(int *__restrict__ out, int n, const int *__restrict__ sums)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] += sums[i / (SCAN_BLOCK * SCAN_ITEMS)];
}

// out[i] = sum of in[0..i), out[n] (= *total) written too when total != nullptr; scratch: ceil(n / 2048) + 1 ints
static int exclusive_scan(const int *in, int n, int *out, int *sums, int *total, cudaStream_t s)
{
    if (n <= 0) return 0;
    const int nb = (n + SCAN_BLOCK * SCAN_ITEMS - 1) / (SCAN_BLOCK * SCAN_ITEMS);
    scan_block_kernel<<<nb, SCAN_BLOCK, 0, s>>>(in, n, out, sums);
    scan_sums_kernel<<<1, 1024, 0, s>>>(sums, nb, total ? total : sums + nb);
    if (nb > 1) scan_add_kernel<<<(n + 255) / 256, 256, 0, s>>>(out, n, sums);
    OFX_LAUNCH_CHECK();
    return 0;
}

}  // namespace ofx
