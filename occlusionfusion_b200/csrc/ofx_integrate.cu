// ofx_integrate.cu -- the fused hot kernels of libofx.so (sm_100a):
//     per voxel: anchors (cached ids) -> skinning weights -> linear-blend warp -> projection ->
//     visibility test -> running-average TSDF / weight / colour update, in one pass.
// Replaces, per frame, WarpField.deform_tsdf (warpfield.py:369-380: skin_tsdf + deform) and
// TSDFVolume.integrate (tsdf.py:378-494) -- in the reference five full-volume host arrays and
// two device round trips; here 24 B of HBM traffic per UPDATED voxel plus K (padded to 4 or 8)
// bytes of anchor cache per voxel of a visited brick, and nothing for the rest of the volume.
//
// Work unit: a brick of 4 x 8 x 32 voxels handled by 256 threads, each thread owning four
// consecutive voxels along z = one float4 of each of the three volumes (coalesced 128 B rows).
//
//   fused_warped_persistent  (warped frames, the per-frame hot path)
//       persistent CTAs (a multiple of the SM count) walk the list of active bricks.  A
//       producer warp gathers the next brick's node records (64 B each: position, rotation,
//       translation) into shared memory with the TMA unit (cp.async.bulk, one copy per node)
//       through a two-stage mbarrier pipeline and decides the brick-level frustum cull, while
//       eight consumer warps process the current brick; counters stay in registers until the
//       CTA retires.
//   fused_brick_kernel       (source frame = identity warp; deform_tsdf materialisation)
//       one CTA per brick, same per-thread body.
#include "ofx_brick.cuh"

namespace ofx {

int sm_count();

// ---- generic one-CTA-per-brick kernel (identity warp, deform materialisation) --------------------
template <int K, bool VEC, int MODE>
__global__ void __launch_bounds__(BRICK_THREADS)
fused_brick_kernel(VolumeGeom g, float *__restrict__ tsdf, float *__restrict__ weight, float *__restrict__ colour,
                   GraphArgs ga, FrameArgs fa, unsigned long long *__restrict__ stats, float *__restrict__ out_pts,
                   uint8_t *__restrict__ out_valid)
{
    __shared__ __align__(128) float s_rec[(MODE == MODE_IDENTITY ? 1 : MAX_BRICK_NODES) * REC_STRIDE];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_stat[3];
    __shared__ int s_cull;

    const int b = (int)blockIdx.x;
    const BrickCoord bc = brick_coord(g, b);
    int n_used = 0;
    if constexpr (MODE != MODE_IDENTITY) {
        n_used = ga.cnt[b];
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    }
    if (threadIdx.x < 3) s_stat[threadIdx.x] = 0ull;
    __syncthreads();
    if constexpr (MODE != MODE_IDENTITY) {
        if (n_used > 0) {
            if (threadIdx.x == 0) mbar_arrive_expect_tx(&s_bar, (uint32_t)n_used * NODE_REC_FLOATS * 4u);
            if (threadIdx.x < 32) {
                const int o = ga.off[b];
                for (int i = threadIdx.x; i < n_used; i += 32)
                    bulk_g2s(s_rec + i * REC_STRIDE,
                             ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + i) * NODE_REC_FLOATS,
                             NODE_REC_FLOATS * 4u, &s_bar);
            }
            mbar_wait(&s_bar, 0);
        }
    }
    if constexpr (MODE == MODE_IDENTITY) {
        if (threadIdx.x < 32) {
            const bool out = brick_outside_frustum(g, bc, s_rec, 0, true, fa.cam);
            if (threadIdx.x == 0) s_cull = out ? 1 : 0;
        }
        __syncthreads();
        if (s_cull) return;
    }
    const float rdenom = (MODE != MODE_IDENTITY) ? __frcp_rn(ga.denom) : 0.f;
    unsigned c_upd = 0, c_valid = 0, c_wsum = 0;
    brick_thread_work<K, VEC, MODE>(g, tsdf, weight, colour, ga, fa, s_rec, bc, n_used, rdenom, threadIdx.x, out_pts,
                                    out_valid, c_upd, c_valid, c_wsum);
    if constexpr (MODE != MODE_DEFORM) {
        c_upd = __reduce_add_sync(0xffffffffu, c_upd);
        c_valid = __reduce_add_sync(0xffffffffu, c_valid);
        c_wsum = __reduce_add_sync(0xffffffffu, c_wsum);
        if ((threadIdx.x & 31) == 0) {
            if (c_upd) atomicAdd(&s_stat[0], (unsigned long long)c_upd);
            if (c_valid) atomicAdd(&s_stat[1], (unsigned long long)c_valid);
            if (c_wsum) atomicAdd(&s_stat[2], (unsigned long long)c_wsum);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            if (s_stat[0]) atomicAdd(stats + OFX_STAT_UPDATED, s_stat[0]);
            if (s_stat[1]) atomicAdd(stats + OFX_STAT_SKIN_VALID, s_stat[1]);
            if (s_stat[2]) atomicAdd(stats + OFX_STAT_WEIGHT_SUM, s_stat[2]);
            atomicAdd(stats + OFX_STAT_BRICKS, 1ull);
        }
    }
}

// ---- persistent, warp-specialised kernel for warped frames ----------------------------------------
constexpr int N_STAGES = 2;
constexpr int CONSUMER_WARPS = BRICK_THREADS / 32;          // 8
constexpr int PERSISTENT_THREADS = BRICK_THREADS + 32;      // + 1 producer warp

struct StageInfo {
    int lx0, iy0, iz0;
    int n_used;      // -1 = no more bricks
    int cull;
};

template <int K, bool VEC>
__global__ void __launch_bounds__(PERSISTENT_THREADS, 4)
fused_warped_persistent(VolumeGeom g, float *__restrict__ tsdf, float *__restrict__ weight, float *__restrict__ colour,
                        GraphArgs ga, FrameArgs fa, int n_active, int *__restrict__ work_counter,
                        unsigned long long *__restrict__ stats)
{
    extern __shared__ __align__(128) unsigned char s_raw[];
    float *s_rec = reinterpret_cast<float *>(s_raw);                       // [N_STAGES][MAX_BRICK_NODES * REC_STRIDE]
    constexpr int STAGE_FLOATS = MAX_BRICK_NODES * REC_STRIDE;
    __shared__ __align__(8) uint64_t bar_data[N_STAGES], bar_full[N_STAGES], bar_empty[N_STAGES];
    __shared__ StageInfo s_info[N_STAGES];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < N_STAGES; ++s) {
            mbar_init(&bar_data[s], 1);
            mbar_init(&bar_full[s], 1);
            mbar_init(&bar_empty[s], CONSUMER_WARPS);
        }
    }
    __syncthreads();

    if (warp == CONSUMER_WARPS) {
        // ===== producer warp: TMA gather of the next brick's node records + frustum cull =====
        // Bricks are handed out dynamically (culled bricks cost ~1/50 of live ones and the two kinds
        // are spatially sorted): the first brick is blockIdx.x, later ones come from an atomic
        // counter.  Exactly n_active fetches happen per launch; whoever draws the last one
        // resets the counter for the next launch.
        int idx = blockIdx.x;
        for (int it = 0;; ++it) {
            if (it > 0) {
                int f = 0;
                if (lane == 0) {
                    f = atomicAdd(work_counter, 1);
                    if (f == n_active - 1) atomicExch(work_counter, 0);
                }
                idx = (int)gridDim.x + __shfl_sync(0xffffffffu, f, 0);
            }
            const int st = it % N_STAGES;
            const uint32_t ph = (uint32_t)(it / N_STAGES) & 1u;
            if (it >= N_STAGES) mbar_wait(&bar_empty[st], ph ^ 1u);        // consumers released this stage
            if (idx >= n_active) {
                if (lane == 0) { s_info[st].n_used = -1; mbar_arrive(&bar_full[st]); }
                break;
            }
            const int b = __ldg(ga.active + idx);
            const int n_used = __ldg(ga.cnt + b), o = __ldg(ga.off + b);
            const BrickCoord bc = brick_coord(g, b);
            float *rec = s_rec + st * STAGE_FLOATS;
            if (lane == 0) mbar_arrive_expect_tx(&bar_data[st], (uint32_t)n_used * NODE_REC_FLOATS * 4u);
            for (int i = lane; i < n_used; i += 32)
                bulk_g2s(rec + i * REC_STRIDE, ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + i) * NODE_REC_FLOATS,
                         NODE_REC_FLOATS * 4u, &bar_data[st]);
            mbar_wait(&bar_data[st], ph);
            const bool out = brick_outside_frustum(g, bc, rec, n_used, false, fa.cam);
            if (lane == 0) {
                s_info[st].lx0 = bc.lx0; s_info[st].iy0 = bc.iy0; s_info[st].iz0 = bc.iz0;
                s_info[st].n_used = n_used; s_info[st].cull = out ? 1 : 0;
                mbar_arrive(&bar_full[st]);                                // release: info + records visible
            }
        }
        return;
    }

    // ===== consumer warps =====
    const float rdenom = __frcp_rn(ga.denom);
    unsigned c_upd = 0, c_valid = 0, c_bricks = 0;
    unsigned long long c_wsum = 0;
    for (int it = 0;; ++it) {
        const int st = it % N_STAGES;
        const uint32_t ph = (uint32_t)(it / N_STAGES) & 1u;
        mbar_wait(&bar_full[st], ph);
        const StageInfo info = s_info[st];
        if (info.n_used < 0) break;
        if (!info.cull) {
            mbar_wait(&bar_data[st], ph);          // already complete; orders the TMA writes before our reads
            const BrickCoord bc{info.lx0, info.iy0, info.iz0};
            unsigned w32 = 0;
            brick_thread_work<K, VEC, MODE_WARPED>(g, tsdf, weight, colour, ga, fa, s_rec + st * STAGE_FLOATS, bc,
                                                   info.n_used, rdenom, threadIdx.x, nullptr, nullptr, c_upd, c_valid,
                                                   w32);
            c_wsum += w32;
            if (threadIdx.x == 0) ++c_bricks;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_empty[st]);
    }
    // counters: once per CTA lifetime
    c_upd = __reduce_add_sync(0xffffffffu, c_upd);
    c_valid = __reduce_add_sync(0xffffffffu, c_valid);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c_wsum += __shfl_xor_sync(0xffffffffu, c_wsum, o);
    if (lane == 0) {
        if (c_upd) atomicAdd(stats + OFX_STAT_UPDATED, (unsigned long long)c_upd);
        if (c_valid) atomicAdd(stats + OFX_STAT_SKIN_VALID, (unsigned long long)c_valid);
        if (c_wsum) atomicAdd(stats + OFX_STAT_WEIGHT_SUM, c_wsum);
        if (c_bricks) atomicAdd(stats + OFX_STAT_BRICKS, (unsigned long long)c_bricks);
    }
}

}  // namespace ofx

using namespace ofx;

static bool volume_vec_ok(const ofx_volume *v)
{
    const uintptr_t a = (uintptr_t)v->desc.tsdf | (uintptr_t)v->desc.weight | (uintptr_t)v->desc.colour |
                        (uintptr_t)v->anchor_cache;
    return v->geom.Dz % 4 == 0 && (a & 15) == 0;
}

static GraphArgs graph_args(const ofx_volume *v)
{
    return GraphArgs{v->node_rec, v->active, v->brick_off, v->brick_cnt, v->brick_nodes, v->anchor_cache, v->cutoff, v->denom};
}

template <int K, int MODE>
static int launch_brick(ofx_volume *v, int grid, const GraphArgs &ga, const FrameArgs &fa, unsigned long long *stats,
                        float *out_pts, uint8_t *out_valid, cudaStream_t s)
{
    if (grid <= 0) return 0;
    if (volume_vec_ok(v))
        fused_brick_kernel<K, true, MODE><<<grid, BRICK_THREADS, 0, s>>>(v->geom, v->desc.tsdf, v->desc.weight,
                                                                         v->desc.colour, ga, fa, stats, out_pts, out_valid);
    else
        fused_brick_kernel<K, false, MODE><<<grid, BRICK_THREADS, 0, s>>>(v->geom, v->desc.tsdf, v->desc.weight,
                                                                          v->desc.colour, ga, fa, stats, out_pts, out_valid);
    OFX_LAUNCH_CHECK();
    return 0;
}

template <int K, bool VEC>
static int launch_persistent_t(ofx_volume *v, const FrameArgs &fa, unsigned long long *stats, cudaStream_t s)
{
    static int blocks_per_sm = 0;      // per instantiation
    const size_t smem = (size_t)N_STAGES * MAX_BRICK_NODES * REC_STRIDE * sizeof(float);
    if (blocks_per_sm == 0) {
        OFX_CUDA(cudaFuncSetAttribute(fused_warped_persistent<K, VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        OFX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, fused_warped_persistent<K, VEC>,
                                                               PERSISTENT_THREADS, smem));
        if (blocks_per_sm < 1) blocks_per_sm = 1;
    }
    const int cap = sm_count() * blocks_per_sm;       // persistent grid: a multiple of the SM count
    const int grid = v->n_active < cap ? v->n_active : cap;
    if (grid <= 0) return 0;
    fused_warped_persistent<K, VEC><<<grid, PERSISTENT_THREADS, smem, s>>>(v->geom, v->desc.tsdf, v->desc.weight,
                                                                          v->desc.colour, graph_args(v), fa,
                                                                          v->n_active, v->scratch + 8, stats);
    OFX_LAUNCH_CHECK();
    return 0;
}

template <int K>
static int launch_persistent(ofx_volume *v, const FrameArgs &fa, unsigned long long *stats, cudaStream_t s)
{
    return volume_vec_ok(v) ? launch_persistent_t<K, true>(v, fa, stats, s) : launch_persistent_t<K, false>(v, fa, stats, s);
}

extern "C" int ofx_integrate(ofx_volume *v, const float *depth, const float *colour_packed, int32_t H, int32_t W,
                             const float *intr_host, double obs_weight, int32_t identity_warp, uint64_t *stats,
                             void *stream)
{
    OFX_REQUIRE(v && depth && colour_packed && intr_host && H > 0 && W > 0, "ofx_integrate: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    unsigned long long *st = reinterpret_cast<unsigned long long *>(stats ? stats : v->stats_scratch);
    OFX_CUDA(cudaMemsetAsync(st, 0, OFX_N_STATS * sizeof(uint64_t), s));
    FrameArgs fa{depth, colour_packed, Camera{intr_host[0], intr_host[1], intr_host[2], intr_host[3], H, W},
                 v->desc.trunc_margin, 1.0 / v->desc.trunc_margin, obs_weight, (float)obs_weight};
    if (identity_warp) {
        GraphArgs ga{};
        return launch_brick<4, MODE_IDENTITY>(v, v->n_bricks, ga, fa, st, nullptr, nullptr, s);
    }
    OFX_REQUIRE(v->graph_ready, "ofx_integrate: warped frame before ofx_set_graph");
    OFX_REQUIRE(v->transforms_ready, "ofx_integrate: warped frame before ofx_set_transforms");
    switch (v->K) {
    case 1: return launch_persistent<1>(v, fa, st, s);
    case 2: return launch_persistent<2>(v, fa, st, s);
    case 3: return launch_persistent<3>(v, fa, st, s);
    case 4: return launch_persistent<4>(v, fa, st, s);
    case 5: return launch_persistent<5>(v, fa, st, s);
    case 6: return launch_persistent<6>(v, fa, st, s);
    case 7: return launch_persistent<7>(v, fa, st, s);
    default: return launch_persistent<8>(v, fa, st, s);
    }
}

extern "C" int ofx_volume_deform(ofx_volume *v, float *out_pts, uint8_t *valid, void *stream)
{
    OFX_REQUIRE(v, "ofx_volume_deform: null handle");
    OFX_REQUIRE(v->graph_ready, "ofx_volume_deform: call ofx_set_graph first");
    FrameArgs fa{};
    const GraphArgs ga = graph_args(v);
    cudaStream_t s = (cudaStream_t)stream;
    switch (v->K) {
    case 1: return launch_brick<1, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    case 2: return launch_brick<2, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    case 3: return launch_brick<3, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    case 4: return launch_brick<4, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    case 5: return launch_brick<5, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    case 6: return launch_brick<6, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    case 7: return launch_brick<7, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    default: return launch_brick<8, MODE_DEFORM>(v, v->n_bricks, ga, fa, nullptr, out_pts, valid, s);
    }
}

extern "C" int ofx_volume_skin(ofx_volume *v, int32_t *anchors, float *weights, uint8_t *valid, void *stream)
{
    OFX_REQUIRE(v && anchors && weights && valid, "ofx_volume_skin: null argument");
    OFX_REQUIRE(v->graph_ready, "ofx_volume_skin: call ofx_set_graph first");
    // Full (n,K) arrays as WarpField.skin returns them, including partially-anchored voxels:
    // exact brute force over all nodes; on request only, never on the per-frame path.
    cudaStream_t s = (cudaStream_t)stream;
    float *pts = nullptr, *nodes = nullptr;
    OFX_CUDA(cudaMalloc(&pts, (size_t)v->n_voxels * 3 * sizeof(float)));
    if (cudaMalloc(&nodes, (size_t)v->M * 3 * sizeof(float)) != cudaSuccess) {
        cudaFree(pts);
        return fail("ofx_volume_skin: out of device memory");
    }
    int rc = ofx_volume_world_points(v, pts, stream);
    if (!rc) {
        cudaMemcpy2DAsync(nodes, 3 * sizeof(float), v->node_rec, NODE_REC_FLOATS * sizeof(float), 3 * sizeof(float),
                          v->M, cudaMemcpyDeviceToDevice, s);
        rc = ofx_skin(pts, v->n_voxels, nodes, v->M, v->coverage, v->K, anchors, weights, valid, stream);
    }
    cudaStreamSynchronize(s);
    cudaFree(pts);
    cudaFree(nodes);
    return rc;
}
