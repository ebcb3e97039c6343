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
//   brick_cull_kernel        per frame: active bricks -> live list (exact frustum cull)
//   fused_warped_persistent  (warped frames, the per-frame hot path)
//       persistent CTAs (a multiple of the SM count) walk the live list.  A producer warp stages
//       the next brick's skin-cache slot (one 20 KB bulk copy) and node transforms (one 48-byte
//       bulk copy per node) in shared memory with the TMA unit; eight consumer warps tabulate the
//       row-invariant part of R p + T, process the brick in three phases (ofx_brick.cuh) and hand
//       the stage back through a named barrier.
//   fused_brick_kernel       (source frame = identity warp; deform_tsdf materialisation)
//       one CTA per brick, same per-thread body.
#include "ofx_brick.cuh"

namespace ofx {

int sm_count();

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

// ---- per-frame brick cull: active bricks -> live list -----------------------------------------------
// A few lanes per active brick walk the brick's node list (records come from L2), bound the pixel box of the
// warped brick and drop the bricks that cannot reach the image.  The bound: every voxel p of a brick lies in its
// axis-aligned box (centre c, half extents h); node k maps the box to an oriented box inside the camera-space
// box R_k c + T_k +- |R_k| h; the blend is a convex combination (up to a scale in (0,1], which the projection
// ignores) of one point from each of those, so the union of their pixel boxes bounds the voxel's pixel
// (ofx_brick.cuh: box_pixbox).  Bricks are long in z (4 x 8 x 32 voxels): the oriented box is several times
// tighter in pixels than the bounding ball used up to v9.  Survivors are appended to live[] (one atomic per
// CTA); the fused kernel then sees only bricks that do arithmetic, so its dynamic scheduler balances equal-cost units.
#ifndef OFX_CULL_BALL
#define OFX_CULL_BALL 0      // 1: bound the warped brick by balls (v4-v9) instead of oriented boxes
#endif
#ifndef OFX_CULL_SUB
#define OFX_CULL_SUB 1       // pieces along z the brick's box is cut into for the bound (2/4/8 drop 0.5-1 % more bricks: no gain)
#endif
constexpr int CULL_SUB = OFX_CULL_SUB;
#ifndef OFX_CULL_LANES
#define OFX_CULL_LANES 4
#endif
#ifndef OFX_CULL_THREADS
#define OFX_CULL_THREADS 128
#endif
constexpr int CULL_LANES = OFX_CULL_LANES;
constexpr int CULL_THREADS = OFX_CULL_THREADS;
constexpr int CULL_BRICKS_PER_CTA = CULL_THREADS / CULL_LANES;

__global__ void __launch_bounds__(CULL_THREADS)
brick_cull_kernel(GraphArgs ga, const int4 *__restrict__ active_desc, int n_active, Camera cam, int *__restrict__ live,
                  int *__restrict__ n_live, unsigned long long *__restrict__ acc)
{
    __shared__ int s_mask[CULL_THREADS / 32];
    __shared__ int s_base;
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
    grid_dep_wait();             // the node-transform kernel before this one has finished
    float u_lo = 3.0e38f, u_hi = -3.0e38f, v_lo = 3.0e38f, v_hi = -3.0e38f;
    int unbounded = 0;
    for (int i = sub; i < n_used; i += CULL_LANES) {
        const float4 *r = reinterpret_cast<const float4 *>(ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + i) * NODE_REC_FLOATS);
        const float4 a = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3);   // R00..R22, T0..T2
        const float bx = a.x * box.x + a.y * box.y + a.z * box.z + d.y;
        const float by = a.w * box.x + c.x * box.y + c.y * box.z + d.z;
        const float bz = c.z * box.x + c.w * box.y + d.x * box.z + d.w;
        // the rigid image of the brick's box is an oriented box; its camera-space half extents are |R| h
#if OFX_CULL_BALL
        const PixBox pb = ball_pixbox(bx, by, bz, box.w, cam);
        unbounded |= pb.unbounded;
        u_lo = fminf(u_lo, pb.u_lo); u_hi = fmaxf(u_hi, pb.u_hi);
        v_lo = fminf(v_lo, pb.v_lo); v_hi = fmaxf(v_hi, pb.v_hi);
#else
        // (optionally cut into CULL_SUB pieces along z, the brick's long side)
        const float hz = ext.z * (1.0f / CULL_SUB);
        const float ex = fabsf(a.x) * ext.x + fabsf(a.y) * ext.y + fabsf(a.z) * hz;
        const float ey = fabsf(a.w) * ext.x + fabsf(c.x) * ext.y + fabsf(c.y) * hz;
        const float ez = fabsf(c.z) * ext.x + fabsf(c.w) * ext.y + fabsf(d.x) * hz;
#pragma unroll
        for (int sb = 0; sb < CULL_SUB; ++sb) {
            const float oz = (float)(2 * sb + 1 - CULL_SUB) * hz;      // centre of piece sb relative to the brick centre
            const PixBox pb = box_pixbox(bx + a.z * oz, by + c.y * oz, bz + d.x * oz, ex, ey, ez, cam);
            unbounded |= pb.unbounded;
            u_lo = fminf(u_lo, pb.u_lo); u_hi = fmaxf(u_hi, pb.u_hi);
            v_lo = fminf(v_lo, pb.v_lo); v_hi = fmaxf(v_hi, pb.v_hi);
        }
#endif
    }
#pragma unroll
    for (int o2 = CULL_LANES / 2; o2 > 0; o2 >>= 1) {
        unbounded |= __shfl_xor_sync(0xffffffffu, unbounded, o2);
        u_lo = fminf(u_lo, __shfl_xor_sync(0xffffffffu, u_lo, o2));
        u_hi = fmaxf(u_hi, __shfl_xor_sync(0xffffffffu, u_hi, o2));
        v_lo = fminf(v_lo, __shfl_xor_sync(0xffffffffu, v_lo, o2));
        v_hi = fmaxf(v_hi, __shfl_xor_sync(0xffffffffu, v_hi, o2));
    }
    // rint(u) in [0, W) <=> u in [-0.5, W - 0.5); 1 px margin
    const bool out = !unbounded && (u_hi < -1.5f || u_lo > (float)cam.W + 0.5f || v_hi < -1.5f || v_lo > (float)cam.H + 0.5f);
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

// ---- persistent, warp-specialised kernel for warped frames ----------------------------------------
// Eight consumer warps process bricks; one producer warp runs ahead of them by up to N_SLOTS bricks.
// The producer draws live bricks (first = blockIdx.x, then an atomic counter), allocates room for the
// brick's node transforms in a ring of RING_RECS records, gathers them with the TMA unit (one 48-byte
// cp.async.bulk per node) and publishes a descriptor.  Three mbarriers per descriptor slot:
//   bar_data   TMA -> consumers: the transforms have landed (transaction bytes)
//   named barrier 1 + slot   consumers -> producer: all eight warps are done with the slot and its records
//                            (bar.arrive / bar.sync: the producer blocks in hardware, it does not poll)
#ifndef OFX_N_SLOTS
#define OFX_N_SLOTS 1
#endif
// 1: the consumers block in hardware on a named barrier ("full") that the producer warp arrives at once the brick's TMA
// bytes have landed, instead of all eight consumer warps polling the mbarrier (N_SLOTS == 1 only)
#ifndef OFX_FULL_BAR
#define OFX_FULL_BAR 0
#endif
#ifndef OFX_MIN_CTAS
#define OFX_MIN_CTAS 4
#endif
#ifndef OFX_TIMING
#define OFX_TIMING 0
#endif
#if OFX_TIMING
// investigation build only: where the persistent kernel's warps wait (ns, summed over warps)
__device__ unsigned long long g_dbg[8];
__device__ unsigned long long g_cta[1024][4];   // per CTA: kernel entry, first data, end, bricks
__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#endif
constexpr int N_SLOTS = OFX_N_SLOTS;

// a CTA's shared-memory counters {updated, skin-valid, weight sum: low 16 bits, weight sum >> 16} into the
// frame's accumulators (one thread)
__device__ __forceinline__ void flush_counters(unsigned *s_cnt, unsigned long long *acc)
{
    if (s_cnt[0]) atomicAdd(acc + OFX_STAT_UPDATED, (unsigned long long)s_cnt[0]);
    if (s_cnt[1]) atomicAdd(acc + OFX_STAT_SKIN_VALID, (unsigned long long)s_cnt[1]);
    const unsigned long long ws = ((unsigned long long)s_cnt[3] << 16) + s_cnt[2];
    if (ws) atomicAdd(acc + OFX_STAT_WEIGHT_SUM, ws);
    s_cnt[0] = s_cnt[1] = s_cnt[2] = s_cnt[3] = 0u;
}
constexpr int RING_RECS = 256;                              // > MAX_BRICK_NODES: the largest brick fits an empty ring
constexpr int CONSUMER_WARPS = BRICK_THREADS / 32;          // 8
constexpr int PERSISTENT_THREADS = BRICK_THREADS + 32;      // + 1 producer warp
constexpr int FULL_BAR_ID = 2 + OFX_N_SLOTS;                // named barriers: 0 __syncthreads, 1..N_SLOTS room, 1+N_SLOTS consumers only
static_assert(!OFX_FULL_BAR || OFX_N_SLOTS == 1, "OFX_FULL_BAR needs a single stage");

struct SlotInfo {
    int lx0, iy0, iz0;
    int n_used;      // -1 = no more bricks
    int rec_off;     // first record of the brick in the ring
};

template <int K, bool VEC, bool OBS32>
__global__ void __launch_bounds__(PERSISTENT_THREADS, K <= 4 ? OFX_MIN_CTAS : 3)
fused_warped_persistent(VolumeGeom g, float *__restrict__ tsdf, float *__restrict__ weight, float *__restrict__ colour,
                        GraphArgs ga, FrameArgs fa, const int *__restrict__ live, int *counters,
                        unsigned long long *acc, unsigned long long *__restrict__ stats)
{
    // dynamic shared memory: [N_SLOTS skin-cache slots | ring of node transforms]
    constexpr int KP = K <= 4 ? 4 : 8;
    constexpr uint32_t SLOT_BYTES = slot_bytes_of(KP);
    extern __shared__ __align__(128) unsigned char s_dyn[];
    float *s_rec = reinterpret_cast<float *>(s_dyn + N_SLOTS * SLOT_BYTES);
    uint8_t *s_tab = s_dyn + N_SLOTS * SLOT_BYTES + RING_RECS * REC_BYTES + (threadIdx.x >> 5) * ROW_TAB_BYTES;   // consumer warps
    __shared__ __align__(8) uint64_t bar_data[N_SLOTS];
    __shared__ SlotInfo s_info[N_SLOTS];
    __shared__ unsigned s_cnt[4];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < N_SLOTS; ++s) mbar_init(&bar_data[s], 1);
        s_cnt[0] = s_cnt[1] = s_cnt[2] = s_cnt[3] = 0u;
    }
    grid_dep_wait();             // the cull before this kernel has finished: live list, n_live, accumulators
    __syncthreads();

    if (warp == CONSUMER_WARPS) {
        // ===== producer warp (all lanes keep the same bookkeeping; lane 0 talks to the barriers) =====
#if OFX_TIMING
        unsigned long long dbg_a = 0, dbg_b = 0, dbg_n = 0;
#endif
        const int n_live = *reinterpret_cast<volatile int *>(counters);          // counters: {n_live, work, retired CTAs}
        int idx = blockIdx.x;
        int rel = 0;                  // iterations [0, rel) have been released by the consumers
        int head = 0, tail = 0;       // ring: next free record / first record of the oldest unreleased brick
        bool wrapped = false;         // in use: [tail, head) when !wrapped, [tail, end) + [0, head) when wrapped
#if OFX_FULL_BAR
        int pending_it = -1;          // brick whose bytes are in flight and whose consumers have not been woken yet
#endif
        for (int it = 0;; ++it) {
            const int slot = it % N_SLOTS;
            // everything that does not depend on the hand-back happens before waiting for it: the brick's
            // descriptor (one 32-byte record written by the cull kernel), the first 32 node ids of its list,
            // the source addresses, and an L2 prefetch of its cache slot
            int n_used = 0, o = 0;
            BrickCoord bc{0, 0, 0};
            const uint8_t *slot_src = nullptr;
            const float *rec_src = nullptr;
            const bool more = idx < n_live;
            if (more) {
                const int4 *d = reinterpret_cast<const int4 *>(live) + 2 * idx;
                const int4 d0 = __ldg(d), d1 = __ldg(d + 1);
                bc = BrickCoord{d0.x, d0.y, d0.z};
                n_used = d0.w; o = d1.x;
                slot_src = ga.cache + (int64_t)d1.y * SLOT_BYTES;
                if (lane < n_used)
                    rec_src = ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + lane) * NODE_REC_FLOATS + NODE_XFORM_OFFSET;
                if (lane < 5 * KP / 4) bulk_prefetch_l2(slot_src + lane * 4096, 4096u);
            }
#if OFX_FULL_BAR
            if (pending_it >= 0) {
                // the previous brick's bytes have landed -> wake its consumers (they block on the named barrier)
                mbar_wait(&bar_data[pending_it % N_SLOTS], (uint32_t)(pending_it / N_SLOTS) & 1u);
                __threadfence_block();
                named_bar_arrive(FULL_BAR_ID, PERSISTENT_THREADS);
                pending_it = -1;
            }
#endif
            // room: a free descriptor slot and n_used contiguous records
            int start = -1;
            while (true) {
                if (it - rel < N_SLOTS) {
                    if (it == rel) { head = 0; tail = 0; wrapped = false; }
                    if (!wrapped) {
                        if (RING_RECS - head >= n_used) { start = head; head += n_used; }
                        else if (tail >= n_used) { start = 0; head = n_used; wrapped = true; }
                    } else if (tail - head >= n_used) { start = head; head += n_used; }
                    if (start >= 0) break;
                }
#if OFX_TIMING
                const unsigned long long t0 = gtime();
#endif
                named_bar_sync(1 + rel % N_SLOTS, PERSISTENT_THREADS);                          // oldest brick done
#if OFX_TIMING
                dbg_a += gtime() - t0; ++dbg_n;
#endif
                ++rel;
                if (rel < it) {
                    const int nt = s_info[rel % N_SLOTS].rec_off;
                    if (nt < tail) wrapped = false;
                    tail = nt;
                }
            }
            if (!more) {
#if OFX_FULL_BAR
                if (lane == 0) s_info[slot].n_used = -1;
                __threadfence_block();
                named_bar_arrive(FULL_BAR_ID, PERSISTENT_THREADS);
#else
                if (lane == 0) { s_info[slot].n_used = -1; mbar_arrive(&bar_data[slot]); }
#endif
                break;
            }
            if (lane == 0) {
                s_info[slot].lx0 = bc.lx0; s_info[slot].iy0 = bc.iy0; s_info[slot].iz0 = bc.iz0;
                s_info[slot].n_used = n_used; s_info[slot].rec_off = start;
                mbar_arrive_expect_tx(&bar_data[slot], (uint32_t)n_used * REC_BYTES + SLOT_BYTES);
                // the brick's skin-cache slot (ids + weights of its 1024 voxels): one bulk copy
                bulk_g2s(s_dyn + slot * SLOT_BYTES, slot_src, SLOT_BYTES, &bar_data[slot]);
            }
            __syncwarp();
            float *rec = s_rec + start * REC_STRIDE;
            if (lane < n_used) bulk_g2s(rec + lane * REC_STRIDE, rec_src, REC_BYTES, &bar_data[slot]);
            for (int i = lane + 32; i < n_used; i += 32)
                bulk_g2s(rec + i * REC_STRIDE,
                         ga.node_rec + (int64_t)__ldg(ga.brick_nodes + o + i) * NODE_REC_FLOATS + NODE_XFORM_OFFSET, REC_BYTES,
                         &bar_data[slot]);
#if OFX_TIMING
            {
                const unsigned long long t2 = gtime();
                mbar_wait(&bar_data[slot], (uint32_t)(it / N_SLOTS) & 1u);
                dbg_b += gtime() - t2;
            }
#endif
#if OFX_FULL_BAR
            pending_it = it;
#endif
            int f = 0;
            if (lane == 0) f = atomicAdd(counters + 1, 1);
            idx = (int)gridDim.x + __shfl_sync(0xffffffffu, f, 0);
        }
#if OFX_TIMING
        if (lane == 0) { atomicAdd(&g_dbg[0], dbg_a); atomicAdd(&g_dbg[1], dbg_n); atomicAdd(&g_dbg[5], dbg_b); }
#endif
        return;
    }

    // ===== consumer warps =====
    // counters: per brick a warp reduction into shared memory (keeps them out of the registers that
    // live across the brick loop), flushed once when the CTA's consumers are done
#if OFX_TIMING
    unsigned long long dbg_a = 0, dbg_b = 0, dbg_n = 0;
    if (threadIdx.x == 0) g_cta[blockIdx.x][0] = gtime();
#endif
    for (int it = 0;; ++it) {
        const int slot = it % N_SLOTS;
        const uint32_t ph = (uint32_t)(it / N_SLOTS) & 1u;
#if OFX_TIMING
        const unsigned long long t0 = gtime();
#endif
#if OFX_FULL_BAR
        (void)ph;
        named_bar_sync(FULL_BAR_ID, PERSISTENT_THREADS);   // the producer has seen this brick's TMA bytes land (or: no more bricks)
#else
        mbar_wait(&bar_data[slot], ph);            // descriptor written and the TMA bytes of this brick have landed
#endif
#if OFX_TIMING
        const unsigned long long t1 = gtime();
        dbg_a += t1 - t0; ++dbg_n;
        if (it == 0 && threadIdx.x == 0) g_cta[blockIdx.x][1] = t1;
#endif
        const SlotInfo info = s_info[slot];
        if (info.n_used < 0) break;
        const BrickCoord bc{info.lx0, info.iy0, info.iz0};
        unsigned c_upd = 0, c_valid = 0, c_wsum = 0;
        brick_thread_work<K, VEC, MODE_WARPED, OBS32>(g, tsdf, weight, colour, ga, fa, s_rec + info.rec_off * REC_STRIDE,
                                                      s_dyn + slot * SLOT_BYTES, s_tab, bc, info.n_used, threadIdx.x,
                                                      nullptr, nullptr, c_upd, c_valid, c_wsum,
                                                      [slot] { named_bar_arrive(1 + slot, PERSISTENT_THREADS); });
#if OFX_TIMING
        dbg_b += gtime() - t1;
#endif
        // one warp reduction for both voxel counts (each at most 128 per warp and brick), one for the weight sum;
        // lane 0 adds them to the CTA's shared-memory counters with result-less 32-bit atomics (nothing waits
        // for them; 64-bit shared atomics would be CAS loops).  A CTA sees far fewer than 2^32 voxels per
        // launch; the weight sum is kept in two halves, 16 bits apart, so that neither can overflow.
        const unsigned both = __reduce_add_sync(0xffffffffu, c_upd | (c_valid << 16));
        c_wsum = __reduce_add_sync(0xffffffffu, c_wsum);
        if (lane == 0) {
            if (both & 0xFFFFu) atomicAdd(&s_cnt[0], both & 0xFFFFu);
            if (both >> 16) atomicAdd(&s_cnt[1], both >> 16);
            if (c_wsum) {
                atomicAdd(&s_cnt[2], c_wsum & 0xFFFFu);
                if (c_wsum >> 16) atomicAdd(&s_cnt[3], c_wsum >> 16);
            }
        }
        if ((it & 4095) == 4095) {
            // (never in practice: > 4096 bricks per CTA and launch) empty the 32-bit counters before they can overflow
            named_bar_sync(1 + N_SLOTS, BRICK_THREADS);
            if (threadIdx.x == 0) flush_counters(s_cnt, acc);
            named_bar_sync(1 + N_SLOTS, BRICK_THREADS);
        }
    }
#if OFX_TIMING
    if (lane == 0) { atomicAdd(&g_dbg[2], dbg_a); atomicAdd(&g_dbg[3], dbg_n); atomicAdd(&g_dbg[4], dbg_b); }
    if (threadIdx.x == 0) { g_cta[blockIdx.x][2] = gtime(); g_cta[blockIdx.x][3] = dbg_n - 1; }
#endif
    named_bar_sync(1 + N_SLOTS, BRICK_THREADS);    // consumers only
    if (threadIdx.x == 0) {
        flush_counters(s_cnt, acc);
        // the last CTA to retire publishes the frame's counters and leaves the launch counters at zero for the
        // next cull of this set (every producer has read n_live and drawn its last work item before its
        // consumers get here)
        const int n_live = *reinterpret_cast<volatile int *>(counters);
        __threadfence();
        if (atomicAdd(counters + 2, 1) == (int)gridDim.x - 1) {
            __threadfence();
            stats[OFX_STAT_UPDATED] = __ldcg(acc + OFX_STAT_UPDATED);
            stats[OFX_STAT_SKIN_VALID] = __ldcg(acc + OFX_STAT_SKIN_VALID);
            stats[OFX_STAT_WEIGHT_SUM] = __ldcg(acc + OFX_STAT_WEIGHT_SUM);
            stats[OFX_STAT_BRICKS] = (unsigned long long)n_live;
            counters[0] = 0; counters[1] = 0; counters[2] = 0;
        }
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
#ifndef OFX_PDL
#define OFX_PDL 1
#endif
static cudaLaunchConfig_t pdl_config(const ofx_volume *v, int grid, int block, size_t smem, cudaStream_t s)
{
    static cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)block);
    cfg.dynamicSmemBytes = smem; cfg.stream = s;
    // Single-GPU volumes only.  On a shard of a multi-GPU volume the early-resident CTAs take the SM room that the
    // per-frame collective's kernel needs at the start of the frame (measured on 8 B200: 0.106 ms/frame with,
    // 0.101 without; the kernels alone 0.098 vs 0.100).
    cfg.attrs = &attr; cfg.numAttrs = (OFX_PDL && v->desc.n_shards == 1) ? 1 : 0;
    return cfg;
}

// The per-frame brick cull of one of the two frame sets (node transforms, live list, launch counters, counter
// accumulators): on the integrating stream for the current set, or ahead of time for the other one (ofx_prepare_frame).
static int launch_cull(ofx_volume *v, int set, const Camera &cam, cudaStream_t s)
{
    int *counters = v->scratch + 8 + 4 * set;    // {n_live, work, retired CTAs}: zero between launches (the fused kernel resets them)
    if (v->counters_dirty[set]) OFX_CUDA(cudaMemsetAsync(counters, 0, 3 * sizeof(int), s));
    v->counters_dirty[set] = true;
    GraphArgs ga = graph_args(v);
    ga.node_rec = v->node_rec_set[set];
    cudaLaunchConfig_t cfg = pdl_config(v, (v->n_active + CULL_BRICKS_PER_CTA - 1) / CULL_BRICKS_PER_CTA, CULL_THREADS, 0, s);
    OFX_CUDA(cudaLaunchKernelEx(&cfg, brick_cull_kernel, ga, (const int4 *)v->active_desc, v->n_active, cam, v->live_set[set],
                                counters, v->stats_acc + set * OFX_N_STATS));
    return 0;
}

template <int K, bool VEC, bool OBS32>
static int launch_persistent_t(ofx_volume *v, const FrameArgs &fa, unsigned long long *stats, bool culled, cudaStream_t s)
{
    static int blocks_per_sm = 0;      // per instantiation
    const size_t smem = (size_t)N_SLOTS * slot_bytes_of(K <= 4 ? 4 : 8) + RING_RECS * REC_BYTES + CONSUMER_WARPS * ROW_TAB_BYTES;
    if (blocks_per_sm == 0) {
        OFX_CUDA(cudaFuncSetAttribute(fused_warped_persistent<K, VEC, OBS32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        OFX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, fused_warped_persistent<K, VEC, OBS32>,
                                                               PERSISTENT_THREADS, smem));
        if (blocks_per_sm < 1) blocks_per_sm = 1;
    }
    if (v->n_active <= 0) {
        OFX_CUDA(cudaMemsetAsync(stats, 0, OFX_N_STATS * sizeof(uint64_t), s));
        return 0;
    }
    const int set = v->cur_set;
    if (!culled)
        if (int rc = launch_cull(v, set, fa.cam, s)) return rc;
#ifdef OFX_CULL_ONLY     // timing builds only: the cull alone
    return 0;
#endif
    const GraphArgs ga = graph_args(v);
    const int cap = sm_count() * blocks_per_sm;       // persistent grid: a multiple of the SM count
    const int grid = v->n_active < cap ? v->n_active : cap;   // CTAs beyond the live count retire at once
    cudaLaunchConfig_t cfg = pdl_config(v, grid, PERSISTENT_THREADS, smem, s);
    OFX_CUDA(cudaLaunchKernelEx(&cfg, fused_warped_persistent<K, VEC, OBS32>, v->geom, v->desc.tsdf, v->desc.weight,
                                v->desc.colour, ga, fa, (const int *)v->live, v->scratch + 8 + 4 * set,
                                v->stats_acc + set * OFX_N_STATS, stats));
    v->counters_dirty[set] = false;
    return 0;
}

template <int K>
static int launch_persistent(ofx_volume *v, const FrameArgs &fa, unsigned long long *stats, bool culled, cudaStream_t s)
{
    // the float32-observation-weight build is the hot one; the general one shares everything else
    if (fa.fu.obs_is_f32)
        return volume_vec_ok(v) ? launch_persistent_t<K, true, true>(v, fa, stats, culled, s)
                                : launch_persistent_t<K, false, true>(v, fa, stats, culled, s);
    return volume_vec_ok(v) ? launch_persistent_t<K, true, false>(v, fa, stats, culled, s)
                            : launch_persistent_t<K, false, false>(v, fa, stats, culled, s);
}

// The next frame's node transforms and live-brick list, into the frame set that is NOT in use: may run on another
// stream while the current frame integrates (the caller orders it after the frame before the current one).
extern "C" int ofx_prepare_frame(ofx_volume *v, const float *rotations, const float *translations, int32_t H, int32_t W,
                                 const float *intr_host, void *stream)
{
    OFX_REQUIRE(v && rotations && translations && intr_host && H > 0 && W > 0, "ofx_prepare_frame: bad arguments");
    OFX_REQUIRE(v->graph_ready, "ofx_prepare_frame: call ofx_set_graph first");
    cudaStream_t s = (cudaStream_t)stream;
    const int set = 1 - v->cur_set;
    v->prepared = false;
    if (int rc = ofx_write_transforms(v, v->node_rec_set[set], rotations, translations, s)) return rc;
    const Camera cam{intr_host[0], intr_host[1], intr_host[2], intr_host[3], H, W};
    if (v->n_active > 0)
        if (int rc = launch_cull(v, set, cam, s)) return rc;
    for (int i = 0; i < 4; ++i) v->prepared_intr[i] = intr_host[i];
    v->prepared_hw[0] = H; v->prepared_hw[1] = W;
    v->prepared = true;
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
    const bool culled = v->prepared;
    if (culled) {
        // ofx_prepare_frame has staged this frame in the other set: switch to it
        const float *pi = v->prepared_intr;
        OFX_REQUIRE(pi[0] == intr_host[0] && pi[1] == intr_host[1] && pi[2] == intr_host[2] && pi[3] == intr_host[3] &&
                        v->prepared_hw[0] == H && v->prepared_hw[1] == W,
                    "ofx_integrate: camera differs from the one given to ofx_prepare_frame");
        v->cur_set ^= 1;
        v->node_rec = v->node_rec_set[v->cur_set];
        v->live = v->live_set[v->cur_set];
        v->prepared = false;
        v->transforms_ready = true;
    }
    OFX_REQUIRE(v->transforms_ready, "ofx_integrate: warped frame before ofx_set_transforms");
    switch (v->K) {
    case 1: return launch_persistent<1>(v, fa, st, culled, s);
    case 2: return launch_persistent<2>(v, fa, st, culled, s);
    case 3: return launch_persistent<3>(v, fa, st, culled, s);
    case 4: return launch_persistent<4>(v, fa, st, culled, s);
    case 5: return launch_persistent<5>(v, fa, st, culled, s);
    case 6: return launch_persistent<6>(v, fa, st, culled, s);
    case 7: return launch_persistent<7>(v, fa, st, culled, s);
    default: return launch_persistent<8>(v, fa, st, culled, s);
    }
}

#if OFX_TIMING
extern "C" int ofx_debug_timing(unsigned long long *out_host, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out_host, ofx::g_dbg, sizeof(unsigned long long) * 8);
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(ofx::g_dbg, z, sizeof(z)); }
    return 0;
}
extern "C" int ofx_debug_cta(unsigned long long *out_host)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out_host, ofx::g_cta, sizeof(unsigned long long) * 1024 * 4);
    return 0;
}
#endif

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
