// ofx_integrate.cu -- the fused hot kernels of libofx.so (sm_100a):
//     per voxel: cached anchors + weights -> linear-blend warp -> projection -> visibility test ->
//     running-average TSDF / weight / colour update, in one pass.
// Replaces, per frame, WarpField.deform_tsdf (warpfield.py:369-380: skin_tsdf + deform) and
// TSDFVolume.integrate (tsdf.py:378-494) -- in the reference five full-volume host arrays and
// two device round trips; here 24 B of HBM traffic per UPDATED voxel plus 20 B of skin cache per
// voxel of a visited brick (K <= 4), and nothing for the rest of the volume.
//
// Work unit: a brick of 4 x 8 x 32 voxels handled by 256 threads, each thread owning four
// consecutive voxels along z = one float4 of each of the three volumes (coalesced 128 B rows).
//
//   frame_prep_kernel        per frame: interleaved {depth, colour} image, depth tile maxima, node transforms
//   brick_cull_kernel        per frame: active bricks -> live list (exact frustum + depth-range cull)
//   fused_warped_persistent  (ofx_fused.cuh: warped frames, the per-frame hot path)
//       persistent CTAs (a multiple of the SM count) walk the live list: a producer warp fills a ring
//       of shared-memory stages with the TMA unit (a brick's 20 KB skin-cache slot + one 48-byte bulk
//       copy per node), consumer groups of eight warps take bricks from the ring in turn.
//   fused_brick_kernel       (source frame = identity warp; deform_tsdf materialisation; warped frames of
//       ragged volumes / non-float32 observation weights) one CTA per brick, the generic thread body
//       of ofx_brick.cuh.
#include "ofx_fused.cuh"

namespace ofx {

int sm_count();
int graph_scratch_alloc(void **p, size_t bytes, cudaStream_t s);   // ofx_graph.cu: the stream-ordered scratch pool

// ---- generic one-CTA-per-brick kernel (identity warp, deform materialisation) --------------------
// dynamic shared memory (non-identity modes): [skin-cache slot | node transforms]
template <int K, bool VEC, int MODE, bool OBS32>
__global__ void __launch_bounds__(BRICK_THREADS)
fused_brick_kernel(VolumeGeom g, float *__restrict__ tsdf, float *__restrict__ weight, float *__restrict__ colour,
                   GraphArgs ga, FrameArgs fa, unsigned long long *__restrict__ stats, float *__restrict__ out_pts,
                   uint8_t *__restrict__ out_valid)
{
    constexpr int KP = K <= 4 ? 4 : 8;
    constexpr uint32_t SLOT_BYTES = slot_bytes_of(KP);
    extern __shared__ __align__(128) unsigned char s_dyn[];
    unsigned char *s_slot = s_dyn;
    float *s_rec = reinterpret_cast<float *>(s_dyn + SLOT_BYTES);
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_stat[3];
    __shared__ int s_cull;

    const int b = (int)blockIdx.x;
    const BrickCoord bc = brick_coord(g, b);
    int n_used = 0;
    if constexpr (MODE != MODE_IDENTITY) {
        n_used = ga.cnt[b];
        if (MODE == MODE_WARPED && n_used == 0) return;      // no voxel of the brick is skin-valid
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    }
    if (threadIdx.x < 3) s_stat[threadIdx.x] = 0ull;
    __syncthreads();
    if constexpr (MODE != MODE_IDENTITY) {
        if (n_used > 0) {
            if (threadIdx.x == 0) {
                mbar_arrive_expect_tx(&s_bar, (uint32_t)n_used * REC_BYTES + SLOT_BYTES);
                bulk_g2s(s_slot, ga.cache + (int64_t)ga.brick_slot[b] * SLOT_BYTES, SLOT_BYTES, &s_bar);
            }
            __syncthreads();
            if (threadIdx.x < 32) {
                const int o = ga.off[b];
                for (int i = threadIdx.x; i < n_used; i += 32)
                    bulk_g2s(s_rec + i * REC_STRIDE,
                             ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + i) * NODE_REC_FLOATS + NODE_XFORM_OFFSET,
                             REC_BYTES, &s_bar);
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
    unsigned c_upd = 0, c_valid = 0, c_wsum = 0;
    brick_thread_work<K, VEC, MODE, OBS32>(g, tsdf, weight, colour, ga, fa, s_rec, s_slot, nullptr, bc, n_used, threadIdx.x, out_pts,
                                           out_valid, c_upd, c_valid, c_wsum, [] {});
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

// ---- per-frame preparation: interleaved frame image, depth tile maxima, node transforms ----------------
// One launch, three independent jobs (block ranges):
//   * {depth, packed colour} -> ONE interleaved image (the fused kernel gathers 8 bytes per candidate voxel once)
//   * the maximum valid depth of every DEPTH_TILE x DEPTH_TILE pixel tile (0 = no valid depth), which lets the
//     brick cull drop bricks that lie wholly behind the observed surface or off its silhouette
//   * R | T of the M nodes into the frame set's node-record table (ofx_integrate_warped; what ofx_set_transforms does)
constexpr int DEPTH_TILE = 8;
constexpr int PREP_BW = 32, PREP_BH = DEPTH_TILE;            // pixels per CTA: 4 tiles wide, 1 tile tall

__global__ void __launch_bounds__(PREP_BW * PREP_BH)
frame_prep_kernel(const float *__restrict__ depth, const float *__restrict__ colour, int H, int W, float2 *__restrict__ dc,
                  float *__restrict__ tiles, int tw, int blocks_per_row, int n_img_blocks, const float *__restrict__ R,
                  const float *__restrict__ T, int M, float *__restrict__ rec)
{
    grid_dep_launch();           // the cull's CTAs may move in as this grid drains (they wait for it to finish)
    const int b = (int)blockIdx.x;
    if (b < n_img_blocks) {
        __shared__ float s_m[PREP_BH][PREP_BW / DEPTH_TILE];
        const int bx = b % blocks_per_row, by = b / blocks_per_row;
        const int lane = threadIdx.x & 31, row = threadIdx.x >> 5;
        const int x = bx * PREP_BW + lane, y = by * PREP_BH + row;
        float d = 0.0f;
        if (x < W && y < H) {
            const int i = y * W + x;
            d = __ldg(depth + i);
            dc[i] = make_float2(d, __ldg(colour + i));
        }
        float m = d > 0.0f ? d : 0.0f;                       // invalid (<= 0, NaN) never updates (tsdf.py:609)
#pragma unroll
        for (int o = 1; o < DEPTH_TILE; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((lane & (DEPTH_TILE - 1)) == 0) s_m[row][lane / DEPTH_TILE] = m;
        __syncthreads();
        if (threadIdx.x < PREP_BW / DEPTH_TILE) {
            float mm = 0.0f;
#pragma unroll
            for (int r = 0; r < PREP_BH; ++r) mm = fmaxf(mm, s_m[r][threadIdx.x]);
            const int tx = bx * (PREP_BW / DEPTH_TILE) + (int)threadIdx.x;
            if (tx < tw) tiles[by * tw + tx] = mm;
        }
    } else {
        const int t = (b - n_img_blocks) * (int)blockDim.x + (int)threadIdx.x;
        if (t < M * 12) {
            const int j = t / 12, e = t % 12;
            rec[(int64_t)j * NODE_REC_FLOATS + NODE_XFORM_OFFSET + e] = e < 9 ? __ldg(R + 9 * j + e) : __ldg(T + 3 * j + (e - 9));
        }
    }
}

// ---- per-frame brick cull: active bricks -> live list -----------------------------------------------
// A few lanes per active brick walk the brick's node list (records come from L2), bound the pixel box and the
// depth range of the warped brick and drop the bricks that cannot update a voxel.  The bound: every voxel p of a
// brick lies in its axis-aligned box (centre c, half extents h); node k maps the box to an oriented box inside the
// camera-space box R_k c + T_k +- |R_k| h; the blend is a convex combination, up to a scale sigma in (0,1] which
// the projection ignores, of one point from each of those, so the union of their pixel boxes bounds the voxel's
// pixel (ofx_brick.cuh: box_pixbox) and sigma * (the smallest z of the boxes) bounds its depth from below.
//   * frustum: a brick whose pixel box misses the image (1 px margin) never reaches the fused kernel;
//   * depth (tiles != nullptr): a voxel updates only when depth[pixel] > 0 and z <= depth[pixel] + trunc
//     (tsdf.py:609); a brick whose smallest z exceeds trunc + the largest valid depth over the tiles its pixel box
//     touches -- or whose tiles hold no valid depth at all -- cannot update anything.  sigma >= 1 - 1e-6 / (K e^-8)
//     > 0.996 (warpfield.py:114-121: every cached anchor lies inside 4 * coverage).
// Survivors are appended to live[] (one atomic per CTA); the fused kernel then sees only bricks that do arithmetic,
// so its dynamic scheduler balances equal-cost units.
#ifndef OFX_CULL_THREADS
#define OFX_CULL_THREADS 128
#endif
#ifndef OFX_DEPTH_CULL
#define OFX_DEPTH_CULL 1
#endif
constexpr int CULL_THREADS = OFX_CULL_THREADS;
constexpr int CULL_MAX_TILES = 256;      // larger pixel boxes (bricks next to the camera) skip the depth test

// CULL_LANES lanes share a brick: its node list is a chain of dependent loads (list entry -> node record), so the
// kernel's duration is (nodes per lane) x (memory latency) until the grid fills the machine.  Small active lists
// (the band workloads: a few thousand bricks) take a whole warp per brick -- one round of loads --, large ones
// four lanes per brick (profiles/: 16 us -> 6 us at 256^3, unchanged at 512^3 dense).
template <int CULL_LANES>
__global__ void __launch_bounds__(CULL_THREADS)
brick_cull_kernel(GraphArgs ga, const int4 *__restrict__ active_desc, int n_active, Camera cam, const float *__restrict__ tiles,
                  int tw, float trunc, int *__restrict__ live, int *__restrict__ n_live, unsigned long long *__restrict__ acc)
{
    __shared__ int s_mask[CULL_THREADS / 32];
    __shared__ int s_base;
    constexpr int CULL_BRICKS_PER_CTA = CULL_THREADS / CULL_LANES;
    const int slot = threadIdx.x / CULL_LANES, sub = threadIdx.x % CULL_LANES;
    const int idx = blockIdx.x * CULL_BRICKS_PER_CTA + slot;
    grid_dep_launch();           // the fused kernel's CTAs may move in as this grid drains (they wait for it to finish)
    // the frame's counter accumulators start from zero here (nothing else writes them before the fused kernel)
    if (blockIdx.x == 0 && threadIdx.x < OFX_N_STATS) acc[threadIdx.x] = 0ull;
    int n_used = 0, o = 0;
    int4 d0 = make_int4(0, 0, 0, 0), d1 = make_int4(0, 0, -1, 0);
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f), ext = box;
    if (idx < n_active) {
        const int4 *d = active_desc + ACTIVE_DESC_VEC * idx;
        d0 = __ldg(d); d1 = __ldg(d + 1);
        const int4 bx = __ldg(d + 2), ex = __ldg(d + 3);
        box = make_float4(__int_as_float(bx.x), __int_as_float(bx.y), __int_as_float(bx.z), __int_as_float(bx.w));
        ext = make_float4(__int_as_float(ex.x), __int_as_float(ex.y), __int_as_float(ex.z), 0.f);
        n_used = d0.w; o = d1.x;
    }
    const int b = d1.z;
    grid_dep_wait();             // the kernel before this one (node transforms, depth tiles) has finished
    float u_lo = 3.0e38f, u_hi = -3.0e38f, v_lo = 3.0e38f, v_hi = -3.0e38f, z_min = 3.0e38f;
    int unbounded = 0;
    for (int i = sub; i < n_used; i += CULL_LANES) {
        const float4 *r = reinterpret_cast<const float4 *>(ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + i) * NODE_REC_FLOATS);
        const float4 a = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3);   // R00..R22, T0..T2
        const float bx = a.x * box.x + a.y * box.y + a.z * box.z + d.y;
        const float by = a.w * box.x + c.x * box.y + c.y * box.z + d.z;
        const float bz = c.z * box.x + c.w * box.y + d.x * box.z + d.w;
        // the rigid image of the brick's box is an oriented box; its camera-space half extents are |R| h
        const float ex = fabsf(a.x) * ext.x + fabsf(a.y) * ext.y + fabsf(a.z) * ext.z;
        const float ey = fabsf(a.w) * ext.x + fabsf(c.x) * ext.y + fabsf(c.y) * ext.z;
        const float ez = fabsf(c.z) * ext.x + fabsf(c.w) * ext.y + fabsf(d.x) * ext.z;
        const PixBox pb = box_pixbox(bx, by, bz, ex, ey, ez, cam);
        unbounded |= pb.unbounded;
        u_lo = fminf(u_lo, pb.u_lo); u_hi = fmaxf(u_hi, pb.u_hi);
        v_lo = fminf(v_lo, pb.v_lo); v_hi = fmaxf(v_hi, pb.v_hi);
        z_min = fminf(z_min, pb.z_lo);
    }
#pragma unroll
    for (int o2 = CULL_LANES / 2; o2 > 0; o2 >>= 1) {
        unbounded |= __shfl_xor_sync(0xffffffffu, unbounded, o2);
        u_lo = fminf(u_lo, __shfl_xor_sync(0xffffffffu, u_lo, o2));
        u_hi = fmaxf(u_hi, __shfl_xor_sync(0xffffffffu, u_hi, o2));
        v_lo = fminf(v_lo, __shfl_xor_sync(0xffffffffu, v_lo, o2));
        v_hi = fmaxf(v_hi, __shfl_xor_sync(0xffffffffu, v_hi, o2));
        z_min = fminf(z_min, __shfl_xor_sync(0xffffffffu, z_min, o2));
    }
    // rint(u) in [0, W) <=> u in [-0.5, W - 0.5); 1 px margin
    bool out = !unbounded && (u_hi < -1.5f || u_lo > (float)cam.W + 0.5f || v_hi < -1.5f || v_lo > (float)cam.H + 0.5f);
    if (OFX_DEPTH_CULL && tiles != nullptr) {
        // the tiles the pixel box touches (same 1 px margin), the box's lanes share the lookups
        const bool test = b >= 0 && !unbounded && !out;
        const int x0 = max(0, (int)floorf(u_lo - 1.5f)) / DEPTH_TILE, x1 = min(cam.W - 1, (int)ceilf(u_hi + 1.5f)) / DEPTH_TILE;
        const int y0 = max(0, (int)floorf(v_lo - 1.5f)) / DEPTH_TILE, y1 = min(cam.H - 1, (int)ceilf(v_hi + 1.5f)) / DEPTH_TILE;
        const int nx = x1 - x0 + 1, ny = y1 - y0 + 1;
        const bool small = test && nx > 0 && ny > 0 && nx * ny <= CULL_MAX_TILES;
        float tmax = 0.0f;
        if (small)
            for (int i = sub; i < nx * ny; i += CULL_LANES) tmax = fmaxf(tmax, __ldg(tiles + (y0 + i / nx) * tw + x0 + i % nx));
#pragma unroll
        for (int o2 = CULL_LANES / 2; o2 > 0; o2 >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o2));
        if (small && (!(tmax > 0.0f) || z_min * 0.996f - 1e-4f > tmax + trunc)) out = true;
    }
    const bool keep = b >= 0 && !out;
    const unsigned bal = __ballot_sync(0xffffffffu, keep && sub == 0);
    if ((threadIdx.x & 31) == 0) s_mask[threadIdx.x >> 5] = (int)bal;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
        for (int w = 0; w < CULL_THREADS / 32; ++w) tot += __popc((unsigned)s_mask[w]);
        s_base = tot ? atomicAdd(n_live, tot) : 0;
    }
    __syncthreads();
    if (keep && sub == 0) {
        int rank = 0;
        const int w = threadIdx.x >> 5;
        for (int k = 0; k < w; ++k) rank += __popc((unsigned)s_mask[k]);
        rank += __popc((unsigned)s_mask[w] & ((1u << (threadIdx.x & 31)) - 1u));
        // everything the fused kernel's producer needs about the brick, in one 32-byte record
        int4 *d = reinterpret_cast<int4 *>(live) + 2 * (s_base + rank);
        d[0] = d0;
        d[1] = d1;
    }
}

}  // namespace ofx

using namespace ofx;

int ofx_write_transforms(ofx_volume *v, float *table, const float *rotations, const float *translations, cudaStream_t s);

static bool volume_vec_ok(const ofx_volume *v)
{
    const uintptr_t a = (uintptr_t)v->desc.tsdf | (uintptr_t)v->desc.weight | (uintptr_t)v->desc.colour;
    return v->geom.Dz % 4 == 0 && (a & 15) == 0;
}

static GraphArgs graph_args(const ofx_volume *v)
{
    return GraphArgs{v->node_rec, v->active, v->brick_off, v->brick_cnt, v->brick_nodes, v->brick_slot, v->cache,
                     v->cutoff, v->denom};
}

template <int K, int MODE>
static int launch_brick(ofx_volume *v, int grid, const GraphArgs &ga, const FrameArgs &fa, unsigned long long *stats,
                        float *out_pts, uint8_t *out_valid, cudaStream_t s)
{
    if (grid <= 0) return 0;
    const size_t smem = MODE == MODE_IDENTITY ? 16 : (size_t)slot_bytes_of(K <= 4 ? 4 : 8) + MAX_BRICK_NODES * REC_BYTES;
    if (volume_vec_ok(v)) {
        OFX_CUDA(cudaFuncSetAttribute(fused_brick_kernel<K, true, MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        fused_brick_kernel<K, true, MODE, false><<<grid, BRICK_THREADS, smem, s>>>(v->geom, v->desc.tsdf, v->desc.weight,
                                                                            v->desc.colour, ga, fa, stats, out_pts, out_valid);
    } else {
        OFX_CUDA(cudaFuncSetAttribute(fused_brick_kernel<K, false, MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        fused_brick_kernel<K, false, MODE, false><<<grid, BRICK_THREADS, smem, s>>>(v->geom, v->desc.tsdf, v->desc.weight,
                                                                             v->desc.colour, ga, fa, stats, out_pts, out_valid);
    }
    OFX_LAUNCH_CHECK();
    return 0;
}

// Launch configuration with programmatic stream serialisation: the kernel's CTAs may be scheduled while the
// kernel before it in the stream drains; it calls grid_dep_wait() before it touches that kernel's results.
// (The attribute lives in the caller's frame: nothing process-global.)
#ifndef OFX_PDL
#define OFX_PDL 1
#endif
#ifndef OFX_SMALL_REC
#define OFX_SMALL_REC 1
#endif
#ifndef OFX_STAGE_RESERVE
#define OFX_STAGE_RESERVE 0      // CTA slots per SM the fused kernel leaves to the staging launches of the next frame (measured: 1 costs the fused kernel 15-20 %, more than the staging hides; tools/shard_probe.py)
#endif
#ifndef OFX_PDL_SHARDS
#define OFX_PDL_SHARDS 1      // shards of a multi-GPU volume too: 0.0903 vs 0.0915 ms/frame on 8 B200 (round 2; in round 1, with a
#endif                        // separate node-transform kernel in front, it was slower: DESIGN.md section 6)
struct PdlLaunch {
    cudaLaunchAttribute attr;
    cudaLaunchConfig_t cfg;
    PdlLaunch(const ofx_volume *v, int grid, int block, size_t smem, cudaStream_t s)
    {
        attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr.val.programmaticStreamSerializationAllowed = 1;
        cfg = cudaLaunchConfig_t{};
        cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)block);
        cfg.dynamicSmemBytes = smem; cfg.stream = s;
        cfg.attrs = &attr;
        cfg.numAttrs = (OFX_PDL && (v->desc.n_shards == 1 || OFX_PDL_SHARDS)) ? 1 : 0;
    }
    PdlLaunch(const PdlLaunch &) = delete;
};

// the handle's interleaved frame images + depth tiles, one per frame set (grown on demand; the first frame of a size
// allocates): set k at offset k * capacity
static int ensure_frame_buffers(ofx_volume *v, int H, int W)
{
    const int64_t npix = (int64_t)H * W;
    if (npix <= v->frame_capacity) return 0;
    if (v->frame_dc) OFX_CUDA(cudaFree(v->frame_dc));
    if (v->depth_tiles) OFX_CUDA(cudaFree(v->depth_tiles));
    v->frame_dc = nullptr; v->depth_tiles = nullptr; v->frame_capacity = 0; v->tile_capacity = 0;
    OFX_CUDA(cudaMalloc(&v->frame_dc, (size_t)2 * npix * sizeof(float2)));
    const int64_t nt = (int64_t)((W + DEPTH_TILE - 1) / DEPTH_TILE) * ((H + DEPTH_TILE - 1) / DEPTH_TILE);
    OFX_CUDA(cudaMalloc(&v->depth_tiles, (size_t)2 * nt * sizeof(float)));
    v->frame_capacity = npix; v->tile_capacity = nt;
    v->prepared = 0;             // a staged frame lived in the old buffers
    return 0;
}
static inline float2 *frame_dc_of(const ofx_volume *v, int set) { return v->frame_dc + (size_t)set * v->frame_capacity; }
static inline float *tiles_of(const ofx_volume *v, int set) { return v->depth_tiles + (size_t)set * v->tile_capacity; }

// frame_prep_kernel: interleaved image + depth tiles (+ node transforms into `table` when rotations != nullptr)
static int launch_prep(ofx_volume *v, int set, const float *depth, const float *colour, int H, int W, const float *rotations,
                       const float *translations, float *table, cudaStream_t s)
{
    if (int rc = ensure_frame_buffers(v, H, W)) return rc;
    const int bpr = (W + PREP_BW - 1) / PREP_BW, n_img = bpr * ((H + PREP_BH - 1) / PREP_BH);
    const int n_xf = rotations ? (v->M * 12 + PREP_BW * PREP_BH - 1) / (PREP_BW * PREP_BH) : 0;
    frame_prep_kernel<<<n_img + n_xf, PREP_BW * PREP_BH, 0, s>>>(depth, colour, H, W, frame_dc_of(v, set), tiles_of(v, set),
                                                                 (W + DEPTH_TILE - 1) / DEPTH_TILE, bpr, n_img, rotations,
                                                                 translations, v->M, table);
    OFX_LAUNCH_CHECK();
    return 0;
}

// The per-frame brick cull of one of the two frame sets (node transforms, live list, launch counters, counter
// accumulators): on the integrating stream for the current set, or ahead of time for the other one (ofx_prepare_frame;
// no depth image yet, so frustum only: tiles == nullptr).
static int launch_cull(ofx_volume *v, int set, const Camera &cam, const float *tiles, cudaStream_t s)
{
    int *counters = v->scratch + 8 + 4 * set;    // {n_live, work, retired CTAs}: zero between launches (the fused kernel resets them)
    if (v->counters_dirty[set]) OFX_CUDA(cudaMemsetAsync(counters, 0, 3 * sizeof(int), s));
    v->counters_dirty[set] = true;
    GraphArgs ga = graph_args(v);
    ga.node_rec = v->node_rec_set[set];
    const int tw = (cam.W + DEPTH_TILE - 1) / DEPTH_TILE;
    const float trunc = (float)v->desc.trunc_margin;
    int *live = v->live_set[set];
    unsigned long long *acc = v->stats_acc + set * OFX_N_STATS;
    const int4 *desc = (const int4 *)v->active_desc;
    // lanes per brick by the size of the active list (see brick_cull_kernel)
    const int lanes = v->n_active <= 8192 ? 32 : (v->n_active <= 32768 ? 16 : (v->n_active <= 98304 ? 8 : 4));
    PdlLaunch L(v, (int)(((int64_t)v->n_active * lanes + CULL_THREADS - 1) / CULL_THREADS), CULL_THREADS, 0, s);
    switch (lanes) {
    case 32: OFX_CUDA(cudaLaunchKernelEx(&L.cfg, brick_cull_kernel<32>, ga, desc, v->n_active, cam, tiles, tw, trunc, live, counters, acc)); break;
    case 16: OFX_CUDA(cudaLaunchKernelEx(&L.cfg, brick_cull_kernel<16>, ga, desc, v->n_active, cam, tiles, tw, trunc, live, counters, acc)); break;
    case 8: OFX_CUDA(cudaLaunchKernelEx(&L.cfg, brick_cull_kernel<8>, ga, desc, v->n_active, cam, tiles, tw, trunc, live, counters, acc)); break;
    default: OFX_CUDA(cudaLaunchKernelEx(&L.cfg, brick_cull_kernel<4>, ga, desc, v->n_active, cam, tiles, tw, trunc, live, counters, acc)); break;
    }
    return 0;
}

template <int K, bool OBS1, bool SMALL>
static int launch_persistent_t(ofx_volume *v, const FrameArgs &fa, unsigned long long *stats, int staged, cudaStream_t s)
{
    using C = PersistCfg<K, SMALL>;
    // (function attributes are per device: set on every launch, nothing cached per process)
    OFX_CUDA(cudaFuncSetAttribute(fused_warped_persistent<K, OBS1, SMALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
    const int set = v->cur_set;
    const GraphArgs ga = graph_args(v);
    // persistent grid: a multiple of the SM count.  A frame that arrived through ofx_stage_frame belongs to a caller that
    // stages the NEXT frame on another stream while this kernel runs: OFX_STAGE_RESERVE CTA slots per SM stay free for
    // those launches (a persistent grid that owns every slot would hold them back until its tail)
    const int per_sm = staged == 2 && C::MIN_CTAS > OFX_STAGE_RESERVE ? C::MIN_CTAS - OFX_STAGE_RESERVE : C::MIN_CTAS;
    const int cap = v->sm_count * per_sm;
    const int want = (v->n_active + C::GROUPS - 1) / C::GROUPS;
    const int grid = want < cap ? want : cap;         // CTAs beyond the live count retire at once
    // row tables sized for this graph (graph_finish): shared memory the CTAs do not ask for is L1 cache for the gathers
    const int tab_cap = v->tab_cap > 0 && v->tab_cap <= ROW_TAB_CAP ? v->tab_cap : ROW_TAB_CAP;
    PdlLaunch L(v, grid, C::THREADS, C::smem(tab_cap), s);
    OFX_CUDA(cudaLaunchKernelEx(&L.cfg, fused_warped_persistent<K, OBS1, SMALL>, v->geom, v->desc.tsdf, v->desc.weight, v->desc.colour,
                                ga, fa, (const int *)v->live, v->scratch + 8 + 4 * set, v->stats_acc + set * OFX_N_STATS, stats, tab_cap));
    v->counters_dirty[set] = false;
    return 0;
}

template <int K>
static int launch_warped(ofx_volume *v, FrameArgs fa, unsigned long long *stats, const float *rotations,
                         const float *translations, int staged, cudaStream_t s)
{
    const float obsf = fa.fu.obsf;
    const bool fast = volume_vec_ok(v) && fa.fu.obs_is_f32 && obsf >= 1e-18f && obsf <= 1e18f;
    if (!fast) {
        // ragged volumes (Dz % 4 != 0) and observation weights outside the float32 fast path: the generic
        // one-CTA-per-brick kernel over every brick of the shard (same arithmetic, no pipeline)
        if (rotations)
            if (int rc = ofx_write_transforms(v, v->node_rec, rotations, translations, s)) return rc;
        OFX_CUDA(cudaMemsetAsync(stats, 0, OFX_N_STATS * sizeof(uint64_t), s));
        return launch_brick<K, MODE_WARPED>(v, v->n_bricks, graph_args(v), fa, stats, nullptr, nullptr, s);
    }
    if (v->n_active <= 0) {
        if (rotations)
            if (int rc = ofx_write_transforms(v, v->node_rec, rotations, translations, s)) return rc;
        OFX_CUDA(cudaMemsetAsync(stats, 0, OFX_N_STATS * sizeof(uint64_t), s));
        return 0;
    }
    // staged: 0 nothing, 1 node transforms + frustum cull (ofx_prepare_frame), 2 everything but the fused kernel (ofx_stage_frame)
    if (staged < 2)
        if (int rc = launch_prep(v, v->cur_set, fa.depth, fa.colour_im, fa.cam.H, fa.cam.W, rotations, translations, v->node_rec, s)) return rc;
    fa.dc = frame_dc_of(v, v->cur_set);
    if (staged == 0)
        if (int rc = launch_cull(v, v->cur_set, fa.cam, tiles_of(v, v->cur_set), s)) return rc;
#ifdef OFX_CULL_ONLY     // timing builds only: preparation + cull alone
    return 0;
#endif
    // graphs whose bricks use at most 32 nodes each run the variant with the small node-transform area (more L1)
    if (OFX_SMALL_REC && v->max_used > 0 && v->max_used <= 32)
        return obsf == 1.0f ? launch_persistent_t<K, true, true>(v, fa, stats, staged, s) : launch_persistent_t<K, false, true>(v, fa, stats, staged, s);
    return obsf == 1.0f ? launch_persistent_t<K, true, false>(v, fa, stats, staged, s) : launch_persistent_t<K, false, false>(v, fa, stats, staged, s);
}

static int dispatch_warped(ofx_volume *v, const FrameArgs &fa, unsigned long long *st, const float *rotations,
                           const float *translations, int staged, cudaStream_t s)
{
    switch (v->K) {
    case 1: return launch_warped<1>(v, fa, st, rotations, translations, staged, s);
    case 2: return launch_warped<2>(v, fa, st, rotations, translations, staged, s);
    case 3: return launch_warped<3>(v, fa, st, rotations, translations, staged, s);
    case 4: return launch_warped<4>(v, fa, st, rotations, translations, staged, s);
    case 5: return launch_warped<5>(v, fa, st, rotations, translations, staged, s);
    case 6: return launch_warped<6>(v, fa, st, rotations, translations, staged, s);
    case 7: return launch_warped<7>(v, fa, st, rotations, translations, staged, s);
    default: return launch_warped<8>(v, fa, st, rotations, translations, staged, s);
    }
}

// The next frame's node transforms and live-brick list, into the frame set that is NOT in use: may run on another
// stream while the current frame integrates (the caller orders it after the frame before the current one).
extern "C" int ofx_prepare_frame(ofx_volume *v, const float *rotations, const float *translations, int32_t H, int32_t W,
                                 const float *intr_host, void *stream)
{
    OFX_REQUIRE(v && rotations && translations && intr_host && H > 0 && W > 0, "ofx_prepare_frame: bad arguments");
    OFX_REQUIRE(v->graph_ready, "ofx_prepare_frame: call ofx_set_graph first");
    if (int rc = graph_finish(v)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int set = 1 - v->cur_set;
    v->prepared = 0;
    if (int rc = ofx_write_transforms(v, v->node_rec_set[set], rotations, translations, s)) return rc;
    const Camera cam{intr_host[0], intr_host[1], intr_host[2], intr_host[3], H, W};
    if (v->n_active > 0)
        if (int rc = launch_cull(v, set, cam, nullptr, s)) return rc;
    for (int i = 0; i < 4; ++i) v->prepared_intr[i] = intr_host[i];
    v->prepared_hw[0] = H; v->prepared_hw[1] = W;
    v->prepared = 1;
    return 0;
}

// Everything of the next warped frame except the fused kernel -- interleaved image, depth tiles, node transforms, brick cull
// with the depth-range test -- into the frame set that is NOT in use: runs on another stream while the current frame
// integrates (the caller orders it after the frame BEFORE the current one, whose set it overwrites).  The next
// ofx_integrate with the same images and camera launches the fused kernel only.
extern "C" int ofx_stage_frame(ofx_volume *v, const float *rotations, const float *translations, const float *depth,
                               const float *colour_packed, int32_t H, int32_t W, const float *intr_host, void *stream)
{
    OFX_REQUIRE(v && rotations && translations && depth && colour_packed && intr_host && H > 0 && W > 0, "ofx_stage_frame: bad arguments");
    OFX_REQUIRE(v->graph_ready, "ofx_stage_frame: call ofx_set_graph first");
    if (int rc = graph_finish(v)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int set = 1 - v->cur_set;
    v->prepared = 0;
    if (v->n_active <= 0 || !volume_vec_ok(v)) {          // nothing to cull / generic kernel: transforms only
        if (int rc = ofx_write_transforms(v, v->node_rec_set[set], rotations, translations, s)) return rc;
        if (v->n_active > 0) {
            const Camera cam0{intr_host[0], intr_host[1], intr_host[2], intr_host[3], H, W};
            if (int rc = launch_cull(v, set, cam0, nullptr, s)) return rc;
        }
        v->prepared = 1;
    } else {
        if (int rc = launch_prep(v, set, depth, colour_packed, H, W, rotations, translations, v->node_rec_set[set], s)) return rc;
        const Camera cam{intr_host[0], intr_host[1], intr_host[2], intr_host[3], H, W};
        if (int rc = launch_cull(v, set, cam, tiles_of(v, set), s)) return rc;
        v->prepared = 2;
        v->staged_depth = depth; v->staged_colour = colour_packed;
    }
    for (int i = 0; i < 4; ++i) v->prepared_intr[i] = intr_host[i];
    v->prepared_hw[0] = H; v->prepared_hw[1] = W;
    return 0;
}

extern "C" int ofx_integrate(ofx_volume *v, const float *depth, const float *colour_packed, int32_t H, int32_t W,
                             const float *intr_host, double obs_weight, int32_t identity_warp, uint64_t *stats,
                             void *stream)
{
    OFX_REQUIRE(v && depth && colour_packed && intr_host && H > 0 && W > 0, "ofx_integrate: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    unsigned long long *st = reinterpret_cast<unsigned long long *>(stats ? stats : v->stats_scratch);
    const FrameArgs fa = make_frame_args(depth, colour_packed, intr_host, H, W, v->desc.trunc_margin, obs_weight);
    if (identity_warp) {
        OFX_CUDA(cudaMemsetAsync(st, 0, OFX_N_STATS * sizeof(uint64_t), s));
        GraphArgs ga{};
        return launch_brick<4, MODE_IDENTITY>(v, v->n_bricks, ga, fa, st, nullptr, nullptr, s);
    }
    OFX_REQUIRE(v->graph_ready, "ofx_integrate: warped frame before ofx_set_graph");
    if (int rc = graph_finish(v)) return rc;
    int staged = v->prepared;
    if (staged) {
        // ofx_prepare_frame has staged this frame in the other set: switch to it
        const float *pi = v->prepared_intr;
        OFX_REQUIRE(pi[0] == intr_host[0] && pi[1] == intr_host[1] && pi[2] == intr_host[2] && pi[3] == intr_host[3] &&
                        v->prepared_hw[0] == H && v->prepared_hw[1] == W,
                    "ofx_integrate: camera differs from the one given to ofx_prepare_frame");
        v->cur_set ^= 1;
        v->node_rec = v->node_rec_set[v->cur_set];
        v->live = v->live_set[v->cur_set];
        v->prepared = 0;
        v->transforms_ready = true;
        if (staged == 2) {
            OFX_REQUIRE(v->staged_depth == depth && v->staged_colour == colour_packed,
                        "ofx_integrate: images differ from the ones given to ofx_stage_frame");
            // a frame weight outside the float32 fast path takes the generic kernel, which reads the plain images
            const FuseArgs &fu = fa.fu;
            if (!(fu.obs_is_f32 && fu.obsf >= 1e-18f && fu.obsf <= 1e18f)) staged = 1;
        }
    }
    OFX_REQUIRE(v->transforms_ready, "ofx_integrate: warped frame before ofx_set_transforms");
    return dispatch_warped(v, fa, st, nullptr, nullptr, staged, s);
}

// ofx_set_transforms + ofx_integrate of a warped frame as ONE call: the node transforms ride in the frame
// preparation launch (three kernels per frame instead of four).
extern "C" int ofx_integrate_warped(ofx_volume *v, const float *rotations, const float *translations, const float *depth,
                                    const float *colour_packed, int32_t H, int32_t W, const float *intr_host,
                                    double obs_weight, uint64_t *stats, void *stream)
{
    OFX_REQUIRE(v && rotations && translations && depth && colour_packed && intr_host && H > 0 && W > 0,
                "ofx_integrate_warped: bad arguments");
    OFX_REQUIRE(v->graph_ready, "ofx_integrate_warped: warped frame before ofx_set_graph");
    if (int rc = graph_finish(v)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    unsigned long long *st = reinterpret_cast<unsigned long long *>(stats ? stats : v->stats_scratch);
    const FrameArgs fa = make_frame_args(depth, colour_packed, intr_host, H, W, v->desc.trunc_margin, obs_weight);
    v->prepared = 0;             // a staged look-ahead frame is superseded by the transforms given here
    const int rc = dispatch_warped(v, fa, st, rotations, translations, 0, s);
    if (!rc) v->transforms_ready = true;
    return rc;
}

#if OFX_TIMING
extern "C" int ofx_debug_cta(unsigned long long *out_host)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out_host, ofx::g_cta, sizeof(unsigned long long) * 2048 * 8);
    return 0;
}
#endif

extern "C" int ofx_volume_deform(ofx_volume *v, float *out_pts, uint8_t *valid, void *stream)
{
    OFX_REQUIRE(v, "ofx_volume_deform: null handle");
    OFX_REQUIRE(v->graph_ready, "ofx_volume_deform: call ofx_set_graph first");
    if (int rc = graph_finish(v)) return rc;
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
    if (int rc = graph_finish(v)) return rc;
    // Full (n,K) arrays as WarpField.skin returns them, including partially-anchored voxels:
    // exact brute force over all nodes; on request only, never on the per-frame path.
    cudaStream_t s = (cudaStream_t)stream;
    float *pts = nullptr, *nodes = nullptr;                       // stream-ordered pool: no device-wide synchronisation
    if (graph_scratch_alloc((void **)&pts, (size_t)v->n_voxels * 3 * sizeof(float), s)) return 1;
    if (graph_scratch_alloc((void **)&nodes, (size_t)v->M * 3 * sizeof(float), s)) { cudaFreeAsync(pts, s); return 1; }
    int rc = ofx_volume_world_points(v, pts, stream);
    if (!rc) {
        cudaMemcpy2DAsync(nodes, 3 * sizeof(float), v->node_rec, NODE_REC_FLOATS * sizeof(float), 3 * sizeof(float),
                          v->M, cudaMemcpyDeviceToDevice, s);
        rc = ofx_skin(pts, v->n_voxels, nodes, v->M, v->coverage, v->K, anchors, weights, valid, stream);
    }
    cudaFreeAsync(pts, s);
    cudaFreeAsync(nodes, s);
    return rc;
}
