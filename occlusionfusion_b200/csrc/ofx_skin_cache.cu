// ofx_skin_cache.cu -- ofx_set_graph: builds the per-voxel anchor cache, the device form of
// WarpField.skin_tsdf (warpfield.py:131-141), which the reference recomputes only when the graph
// changes (node insertion, warpfield.py:493,532-533).
//
// The reference caches 33 B/voxel (anchors i32 x4, weights f32 x4, valid) for every voxel.  Here
// the cache exists only for bricks that can hold a skin-valid voxel (one "slot" per such brick,
// brick-major so that a thread's four voxels are one 16-byte word): K bytes of brick-local anchor
// ids per voxel (padded to 4 or 8; an id indexes the brick's list of the nodes its voxels use)
// and the K normalised float32 weights, computed once here with the contract arithmetic
// (ofx_math.cuh: skin_weights_valid) exactly as the reference computes world_weights once.
// Bricks without a single skin-valid voxel never enter the work list, so the 4*coverage cutoff
// (warpfield.py:109) culls whole bricks exactly.
//
// Pipeline (all on `stream`; ends with one small device->host read, i.e. it synchronises):
//   1 brick_bound_kernel   per brick: K-th nearest node distance D_K from the brick centre;
//                          exact candidate radius R = min(D_K + 2h, cutoff + h) (h = half
//                          diagonal); bricks with D_K - h > cutoff are dead
//   2 scan x2              offsets of the candidate lists; cache slots of the bricks that have candidates
//   3 brick_fill_kernel    candidate node ids, ascending
//   4 skin_brick_kernel    per voxel exact top-K over the brick's candidates (registers),
//                          used-node marking, compaction, byte ids + weights -> skin cache
//   5 scan                 work list of active bricks
#include <vector>

#include <algorithm>
#include <cstdlib>
#include "ofx_volume.cuh"

int ofx_upload_nodes(ofx_volume *v, const float *nodes, int M, cudaStream_t s);

namespace ofx {

constexpr int BOUND_NODE_TILE = 1024;
constexpr float RADIUS_SLACK_REL = 1.0002f;   // covers float32 rounding of centre/voxel distances
constexpr float RADIUS_SLACK_ABS = 1e-5f;

struct BrickBox {
    float cx, cy, cz, h;
    float hx, hy, hz;      // half extents of the voxel-centre box
};

__device__ __forceinline__ BrickBox brick_box(const VolumeGeom &g, int b)
{
    const BrickCoord c = brick_coord(g, b);
    const int lx1 = min(c.lx0 + BX, g.nx) - 1, iy1 = min(c.iy0 + BY, g.Dy) - 1, iz1 = min(c.iz0 + BZ, g.Dz) - 1;
    // planes of one brick may be non-contiguous in world x when shards are dealt round-robin
    float x_lo = g.px[c.lx0], x_hi = x_lo;
    for (int lx = c.lx0 + 1; lx <= lx1; ++lx) { x_lo = fminf(x_lo, g.px[lx]); x_hi = fmaxf(x_hi, g.px[lx]); }
    const float y_lo = g.py[c.iy0], y_hi = g.py[iy1], z_lo = g.pz[c.iz0], z_hi = g.pz[iz1];
    BrickBox bb;
    bb.cx = 0.5f * (x_lo + x_hi); bb.cy = 0.5f * (y_lo + y_hi); bb.cz = 0.5f * (z_lo + z_hi);
    const float hx = 0.5f * (x_hi - x_lo), hy = 0.5f * (y_hi - y_lo), hz = 0.5f * (z_hi - z_lo);
    bb.h = sqrtf(hx * hx + hy * hy + hz * hz) * RADIUS_SLACK_REL + RADIUS_SLACK_ABS;
    bb.hx = hx; bb.hy = hy; bb.hz = hz;
    return bb;
}

// Pass 1+2 (count) when cand == nullptr, pass 3 (fill) otherwise.
__global__ void __launch_bounds__(128) brick_bound_kernel(VolumeGeom g, int n_bricks, const float *__restrict__ node_rec,
                                                          int M, int K, float cutoff, float *__restrict__ radius,
                                                          int *__restrict__ cnt, const int *__restrict__ off,
                                                          int *__restrict__ cand, float4 *__restrict__ box_out)
{
    __shared__ float s_n[BOUND_NODE_TILE * 3];
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = b < n_bricks;
    BrickBox bb{0.f, 0.f, 0.f, 0.f};
    if (live) bb = brick_box(g, b);
    const bool fill = cand != nullptr;
    float R = 0.f;
    int w = 0;
    if (fill && live) { R = radius[b]; w = off[b]; }

    if (!fill) {
        float bd[OFX_MAX_K];
        for (int k = 0; k < OFX_MAX_K; ++k) bd[k] = __int_as_float(0x7f800000);
        for (int base = 0; base < M; base += BOUND_NODE_TILE) {
            const int n = min(BOUND_NODE_TILE, M - base);
            __syncthreads();
            for (int t = threadIdx.x; t < n; t += blockDim.x) {
                const float *r = node_rec + (int64_t)(base + t) * NODE_REC_FLOATS;
                s_n[3 * t] = r[0]; s_n[3 * t + 1] = r[1]; s_n[3 * t + 2] = r[2];
            }
            __syncthreads();
            if (live)
                for (int j = 0; j < n; ++j) {
                    const float d2 = dist2(bb.cx, bb.cy, bb.cz, s_n[3 * j], s_n[3 * j + 1], s_n[3 * j + 2]);
                    if (d2 < bd[K - 1]) {
                        int k = K - 1;
                        while (k > 0 && d2 < bd[k - 1]) { bd[k] = bd[k - 1]; --k; }
                        bd[k] = d2;
                    }
                }
        }
        if (live) {
            const float dk = sqrtf(bd[K - 1]) * RADIUS_SLACK_REL + RADIUS_SLACK_ABS;
            const float cut = cutoff * RADIUS_SLACK_REL + RADIUS_SLACK_ABS;
            // dead brick: even the nearest corner is farther than the cutoff from its K-th node
            R = (sqrtf(bd[K - 1]) - bb.h > cut) ? -1.0f : fminf(dk + 2.0f * bb.h, cut + bb.h);
            radius[b] = R;
            // centre + bounding radius and the box half extents for the per-frame cull (brick_cull_kernel); wider
            // slack: the cull also absorbs the float32 rounding of the transformed centre
            box_out[2 * b] = make_float4(bb.cx, bb.cy, bb.cz, bb.h * 1.001f + 1e-5f);
            box_out[2 * b + 1] = make_float4(bb.hx * 1.001f + 1e-5f, bb.hy * 1.001f + 1e-5f, bb.hz * 1.001f + 1e-5f, 0.0f);
        }
    }
    int count = 0;
    for (int base = 0; base < M; base += BOUND_NODE_TILE) {
        const int n = min(BOUND_NODE_TILE, M - base);
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            const float *r = node_rec + (int64_t)(base + t) * NODE_REC_FLOATS;
            s_n[3 * t] = r[0]; s_n[3 * t + 1] = r[1]; s_n[3 * t + 2] = r[2];
        }
        __syncthreads();
        if (live && R >= 0.f)
            for (int j = 0; j < n; ++j) {
                const float d2 = dist2(bb.cx, bb.cy, bb.cz, s_n[3 * j], s_n[3 * j + 1], s_n[3 * j + 2]);
                if (d2 <= R * R) {
                    if (fill) cand[w + count] = base + j;
                    ++count;
                }
            }
    }
    if (live && !fill) cnt[b] = count;
}

// Exclusive scans over the bricks, two levels, every block busy (the single-CTA scan of round 1 walked 131 k bricks of a
// 512^3 volume in 128 serial rounds, three times per ofx_set_graph).  One pass handles both sequences the set-up needs:
// the candidate counts themselves (-> list offsets) and the flags count > 0 (-> ranks: cache slots, or the compact list).
//   brick_scan_local_kernel   per block of 1024 bricks: local exclusive prefixes, block sums, running maximum
//   brick_scan_apply_kernel   per block: base = sum of the block sums before it (re-reduced by every block: at most 1024
//                             entries), prefixes + base -> outputs, compaction, totals by the last block
constexpr int BSCAN = 1024;

__device__ __forceinline__ int2 block_scan2(int a, int b, int2 *s_warp, int2 *total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int ia = a, ib = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int ya = __shfl_up_sync(0xffffffffu, ia, o), yb = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) { ia += ya; ib += yb; }
    }
    __syncthreads();
    if (lane == 31) s_warp[wid] = make_int2(ia, ib);
    __syncthreads();
    if (wid == 0) {
        int2 t = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int ya = __shfl_up_sync(0xffffffffu, t.x, o), yb = __shfl_up_sync(0xffffffffu, t.y, o);
            if (lane >= o) { t.x += ya; t.y += yb; }
        }
        s_warp[lane] = t;
    }
    __syncthreads();
    const int2 before = wid ? s_warp[wid - 1] : make_int2(0, 0);
    *total = s_warp[31];
    return make_int2(before.x + ia - a, before.y + ib - b);
}

__global__ void __launch_bounds__(BSCAN) brick_scan_local_kernel(const int *__restrict__ cnt, int n, int *__restrict__ off_out,
                                                                 int *__restrict__ rank_out, int2 *__restrict__ block_sums,
                                                                 int *__restrict__ max_out)
{
    __shared__ int2 s_warp[32];
    const int i = blockIdx.x * BSCAN + threadIdx.x;
    const int x = i < n ? cnt[i] : 0;
    int2 total;
    const int2 ex = block_scan2(x, x > 0 ? 1 : 0, s_warp, &total);
    if (i < n) {
        if (off_out) off_out[i] = ex.x;
        rank_out[i] = ex.y;
    }
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
    if (max_out) {
        int m = x;
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(max_out, m);
    }
}

// totals[0] = sum of the counts, totals[1] = number of bricks with count > 0
__global__ void __launch_bounds__(BSCAN) brick_scan_apply_kernel(const int *__restrict__ cnt, int n, int *__restrict__ off_out,
                                                                 int *__restrict__ rank_out, int *__restrict__ compact,
                                                                 const int2 *__restrict__ block_sums, int *__restrict__ total_cnt,
                                                                 int *__restrict__ total_flags)
{
    __shared__ int2 s_warp[32];
    int a = 0, b = 0;
    for (int k = threadIdx.x; k < (int)blockIdx.x; k += BSCAN) { const int2 t = block_sums[k]; a += t.x; b += t.y; }
    int2 base;
    block_scan2(a, b, s_warp, &base);
    const int i = blockIdx.x * BSCAN + threadIdx.x;
    if (i < n) {
        if (off_out) off_out[i] += base.x;
        const int r = rank_out[i] + base.y;
        rank_out[i] = r;
        if (compact && cnt[i] > 0) compact[r] = i;
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
        const int2 mine = block_sums[blockIdx.x];
        if (total_cnt) *total_cnt = base.x + mine.x;
        if (total_flags) *total_flags = base.y + mine.y;
    }
}

// Pass 4.  One CTA per brick; dynamic shared memory:
//   s_pos  [3 * CAND_TILE] floats   candidate positions of the current tile
//   s_rank [max_cnt] uint16         used flag, then rank among used
constexpr int CAND_TILE = 512;

template <int K>
__global__ void __launch_bounds__(BRICK_THREADS) skin_brick_kernel(VolumeGeom g, const float *__restrict__ node_rec,
                                                                   float cutoff, float denom,
                                                                   const int *__restrict__ off,
                                                                   int *__restrict__ cnt, const int *__restrict__ cand,
                                                                   int *__restrict__ brick_nodes,
                                                                   const int *__restrict__ brick_slot,
                                                                   uint8_t *__restrict__ cache,
                                                                   int *__restrict__ error_flag)
{
    extern __shared__ __align__(16) unsigned char s_raw[];
    float *s_pos = reinterpret_cast<float *>(s_raw);
    unsigned short *s_rank = reinterpret_cast<unsigned short *>(s_raw + 3 * CAND_TILE * sizeof(float));
    __shared__ int s_warp[BRICK_THREADS / 32];
    __shared__ int s_carry;

    const int b = blockIdx.x;
    const int n_cand = cnt[b], base_off = off[b];
    if (n_cand == 0) return;   // dead brick: never enters the work list, its cache bytes are never read
    const BrickCoord bc = brick_coord(g, b);
    const int tz = threadIdx.x & 7, ty = (threadIdx.x >> 3) & 7, tx = threadIdx.x >> 6;
    const int lx = bc.lx0 + tx, iy = bc.iy0 + ty, iz0 = bc.iz0 + tz * VOX_PER_THREAD;
    const bool row_live = lx < g.nx && iy < g.Dy;
    const int64_t slot = brick_slot[b];

    bool any_valid = false;
    int bi[VOX_PER_THREAD][K];
    bool ok[VOX_PER_THREAD];
    float bd[VOX_PER_THREAD][K];
    if (n_cand > 0) {
#pragma unroll
        for (int e = 0; e < VOX_PER_THREAD; ++e)
#pragma unroll
            for (int k = 0; k < K; ++k) { bd[e][k] = __int_as_float(0x7f800000); bi[e][k] = -1; }
        float px = 0.f, py = 0.f, pz[VOX_PER_THREAD] = {0.f, 0.f, 0.f, 0.f};
        if (row_live) {
            px = g.px[lx]; py = g.py[iy];
#pragma unroll
            for (int e = 0; e < VOX_PER_THREAD; ++e) pz[e] = (iz0 + e < g.Dz) ? g.pz[iz0 + e] : 0.f;
        }
        for (int t = threadIdx.x; t < n_cand; t += blockDim.x) s_rank[t] = 0;
        for (int base = 0; base < n_cand; base += CAND_TILE) {
            const int n = min(CAND_TILE, n_cand - base);
            __syncthreads();
            for (int t = threadIdx.x; t < n; t += blockDim.x) {
                const float *r = node_rec + (int64_t)cand[base_off + base + t] * NODE_REC_FLOATS;
                s_pos[3 * t] = r[0]; s_pos[3 * t + 1] = r[1]; s_pos[3 * t + 2] = r[2];
            }
            __syncthreads();
            if (row_live)
                for (int j = 0; j < n; ++j) {
                    const float gx = s_pos[3 * j], gy = s_pos[3 * j + 1], gz = s_pos[3 * j + 2];
                    const float dx = px - gx, dy = py - gy;
                    const float dxy = (dx * dx) + (dy * dy);          // shared by the 4 voxels of the thread
#pragma unroll
                    for (int e = 0; e < VOX_PER_THREAD; ++e) {
                        const float dz = pz[e] - gz;
                        topk_insert<K>(dxy + (dz * dz), base + j, bd[e], bi[e]);
                    }
                }
        }
#pragma unroll
        for (int e = 0; e < VOX_PER_THREAD; ++e) {
            // all K within the cutoff <=> the K-th is (distances ascending)   warpfield.py:109-120
            ok[e] = row_live && (iz0 + e < g.Dz) && !(sqrtf(bd[e][K - 1]) > cutoff);
            any_valid |= ok[e];
            if (ok[e])
#pragma unroll
                for (int k = 0; k < K; ++k) s_rank[bi[e][k]] = 1;
        }
    } else {
#pragma unroll
        for (int e = 0; e < VOX_PER_THREAD; ++e) ok[e] = false;
    }
    __syncthreads();

    // rank the used candidates (exclusive scan over the flags), emit the brick's node list
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < n_cand; base += BRICK_THREADS) {
        const int i = base + threadIdx.x;
        const int x = i < n_cand ? s_rank[i] : 0;
        int incl = x;
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        int wpre = 0;
        for (int k = 0; k < wid; ++k) wpre += s_warp[k];
        const int excl = s_carry + wpre + incl - x;
        if (i < n_cand) {
            s_rank[i] = (unsigned short)excl;
            if (x) brick_nodes[base_off + excl] = cand[base_off + i];
        }
        __syncthreads();
        if (threadIdx.x == BRICK_THREADS - 1) s_carry = excl + x;
        __syncthreads();
    }
    const int n_used = s_carry;
    if (threadIdx.x == 0) {
        cnt[b] = n_used;
        if (n_used > MAX_BRICK_NODES) atomicExch(error_flag, n_used);
    }

    // byte ids (0xFF = no anchor) and normalised weights -> the brick's cache slot (layout: ofx_volume.cuh);
    // threads outside the volume write "invalid"
    constexpr int KP = K <= 4 ? 4 : 8;
    const float rdenom = __frcp_rn(denom);
    uint8_t *sl = cache + slot * slot_bytes_of(KP);
#pragma unroll
    for (int e = 0; e < VOX_PER_THREAD; ++e) {
        uint32_t words[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
        float w[KP];
#pragma unroll
        for (int k = 0; k < KP; ++k) w[k] = 0.0f;
        if (ok[e]) {
            words[0] = words[1] = 0u;
#pragma unroll
            for (int k = 0; k < KP; ++k) {
                uint32_t id = 0xFFu;
                if (k < K) id = (uint32_t)min((int)s_rank[bi[e][k < K ? k : 0]], 254);
                words[k >> 2] |= id << (8 * (k & 3));
            }
            float wv[K];
            skin_weights_valid<K>(bd[e], denom, rdenom, wv);            // warpfield.py:104-121
#pragma unroll
            for (int k = 0; k < K; ++k) w[k] = wv[k];
        }
#pragma unroll
        for (int gq = 0; gq < KP / 4; ++gq) {
            const int cell = (gq * 4 + e) * BRICK_THREADS + threadIdx.x;
            reinterpret_cast<uint32_t *>(sl)[cell] = words[gq];
            reinterpret_cast<float4 *>(sl + BRICK_VOXELS * KP)[cell] = make_float4(w[4 * gq], w[4 * gq + 1], w[4 * gq + 2], w[4 * gq + 3]);
        }
    }
    (void)any_valid;
}

// Histogram of the used-node counts of the active bricks (bins 0..ROW_TAB_CAP, bin ROW_TAB_CAP + 1 = more, entry
// ROW_TAB_CAP + 2 = the largest count when it is beyond the bins): the host sizes the persistent kernel's row tables
// and node-transform area from it (choose_tab_cap, graph_finish).
__global__ void node_hist_kernel(const int *__restrict__ cnt, int nb, int *__restrict__ hist)
{
    __shared__ int s_h[ROW_TAB_CAP + 2];
    for (int i = threadIdx.x; i < ROW_TAB_CAP + 2; i += blockDim.x) s_h[i] = 0;
    __syncthreads();
    for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += gridDim.x * blockDim.x) {
        const int c = cnt[b];
        if (c > 0) atomicAdd(&s_h[min(c, ROW_TAB_CAP + 1)], 1);
        if (c > ROW_TAB_CAP) atomicMax(hist + ROW_TAB_CAP + 2, c);         // the largest count (rare: beyond the bins)
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ROW_TAB_CAP + 2; i += blockDim.x)
        if (s_h[i]) atomicAdd(hist + i, s_h[i]);
}

// One record per active brick with everything the per-frame cull needs, contiguous in the order of the
// active list: {lx0, iy0, iz0, n_used | list offset, cache slot, brick, 0 | centre + bounding radius | half extents}.
__global__ void active_desc_kernel(VolumeGeom g, const int *__restrict__ active, const int *__restrict__ n_active_dev,
                                   const int *__restrict__ cnt,
                                   const int *__restrict__ off, const int *__restrict__ brick_slot,
                                   const float4 *__restrict__ brick_box, int4 *__restrict__ desc)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= *n_active_dev) return;                  // the grid covers every brick: the host does not know the count yet
    const int b = active[i];
    const BrickCoord bc = brick_coord(g, b);
    const float4 box = brick_box[2 * b], ext = brick_box[2 * b + 1];
    desc[ACTIVE_DESC_VEC * i] = make_int4(bc.lx0, bc.iy0, bc.iz0, cnt[b]);
    desc[ACTIVE_DESC_VEC * i + 1] = make_int4(off[b], brick_slot[b], b, 0);
    desc[ACTIVE_DESC_VEC * i + 2] = make_int4(__float_as_int(box.x), __float_as_int(box.y), __float_as_int(box.z), __float_as_int(box.w));
    desc[ACTIVE_DESC_VEC * i + 3] = make_int4(__float_as_int(ext.x), __float_as_int(ext.y), __float_as_int(ext.z), 0);
}

}  // namespace ofx

using namespace ofx;

template <int K>
static int launch_skin_brick(ofx_volume *v, int max_cnt, cudaStream_t s)
{
    const size_t smem = 3 * CAND_TILE * sizeof(float) + (size_t)max_cnt * sizeof(unsigned short) + 16;
    OFX_REQUIRE(smem <= 200 * 1024, "ofx_set_graph: a brick has %d candidate nodes; unsupported", max_cnt);
    OFX_CUDA(cudaFuncSetAttribute(skin_brick_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    skin_brick_kernel<K><<<v->n_bricks, BRICK_THREADS, smem, s>>>(v->geom, v->node_rec, v->cutoff, v->denom,
                                                                  v->brick_off, v->brick_cnt, v->cand, v->brick_nodes,
                                                                  v->brick_slot, v->cache, v->scratch + 2);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_set_graph(ofx_volume *v, const float *nodes, int32_t M, double node_coverage, int32_t K,
                             void *stream)
{
    OFX_REQUIRE(v && nodes, "ofx_set_graph: null argument");
    OFX_REQUIRE(K >= 1 && K <= OFX_MAX_K && M >= K, "ofx_set_graph: need 1 <= K <= %d and K <= M (K=%d, M=%d)",
                OFX_MAX_K, K, M);
    OFX_REQUIRE(node_coverage > 0, "ofx_set_graph: node_coverage must be positive");
    cudaStream_t s = (cudaStream_t)stream;
    v->graph_ready = false;
    v->transforms_ready = false;
    if (int rc = ofx_upload_nodes(v, nodes, M, s)) return rc;
    v->M = M; v->K = K; v->coverage = node_coverage;
    v->cutoff = (float)(4.0 * node_coverage);                        // warpfield.py:109
    v->denom = (float)(2.0 * (node_coverage * node_coverage));       // warpfield.py:114

    v->cache_kp = K <= 4 ? 4 : 8;

    const int nb = v->n_bricks, nblk = (nb + BSCAN - 1) / BSCAN;
    v->graph_pending = false;
    OFX_CUDA(cudaMemsetAsync(v->scratch, 0, 16 * sizeof(int), s));
    brick_bound_kernel<<<(nb + 127) / 128, 128, 0, s>>>(v->geom, nb, v->node_rec, M, K, v->cutoff, v->brick_radius,
                                                        v->brick_cnt, nullptr, nullptr, v->brick_box);
    OFX_LAUNCH_CHECK();
    // list offsets (scan of the counts) and cache slots (rank among the bricks that have candidates) in one pass
    brick_scan_local_kernel<<<nblk, BSCAN, 0, s>>>(v->brick_cnt, nb, v->brick_off, v->brick_slot, v->scan_sums, v->scratch + 1);
    brick_scan_apply_kernel<<<nblk, BSCAN, 0, s>>>(v->brick_cnt, nb, v->brick_off, v->brick_slot, nullptr, v->scan_sums,
                                                   v->scratch + 0, v->scratch + 6);
    OFX_LAUNCH_CHECK();
    // the one wait of the set-up: the sizes of the cache and of the node lists are needed on the host to provide the memory
    int tm[8] = {0};
    OFX_CUDA(cudaMemcpyAsync(tm, v->scratch, sizeof(tm), cudaMemcpyDeviceToHost, s));
    OFX_CUDA(cudaStreamSynchronize(s));
    const int64_t total = tm[0];
    const int max_cnt = tm[1];
    v->n_slots = tm[6];
    const int64_t need = (int64_t)v->n_slots * slot_bytes_of(v->cache_kp);
    if (need > v->cache_capacity) {
        if (v->cache) OFX_CUDA(cudaFree(v->cache));
        v->cache = nullptr; v->cache_capacity = 0;
        OFX_CUDA(cudaMalloc(&v->cache, (size_t)(need + need / 8)));
        v->cache_capacity = need + need / 8;
    }
    if (total > v->list_capacity) {
        if (v->cand) OFX_CUDA(cudaFree(v->cand));
        if (v->brick_nodes) OFX_CUDA(cudaFree(v->brick_nodes));
        v->cand = v->brick_nodes = nullptr;
        const int64_t cap = total + total / 4 + 1024;
        OFX_CUDA(cudaMalloc(&v->cand, cap * sizeof(int)));
        OFX_CUDA(cudaMalloc(&v->brick_nodes, cap * sizeof(int)));
        v->list_capacity = cap;
    }
    if (total > 0) {
        brick_bound_kernel<<<(nb + 127) / 128, 128, 0, s>>>(v->geom, nb, v->node_rec, M, K, v->cutoff,
                                                            v->brick_radius, v->brick_cnt, v->brick_off, v->cand, v->brick_box);
        OFX_LAUNCH_CHECK();
    }
    int rc = 0;
    switch (K) {
    case 1: rc = launch_skin_brick<1>(v, max_cnt, s); break;
    case 2: rc = launch_skin_brick<2>(v, max_cnt, s); break;
    case 3: rc = launch_skin_brick<3>(v, max_cnt, s); break;
    case 4: rc = launch_skin_brick<4>(v, max_cnt, s); break;
    case 5: rc = launch_skin_brick<5>(v, max_cnt, s); break;
    case 6: rc = launch_skin_brick<6>(v, max_cnt, s); break;
    case 7: rc = launch_skin_brick<7>(v, max_cnt, s); break;
    default: rc = launch_skin_brick<8>(v, max_cnt, s); break;
    }
    if (rc) return rc;
    // active list = the bricks that kept a skin-valid voxel (v->live is free scratch here: the ranks)
    brick_scan_local_kernel<<<nblk, BSCAN, 0, s>>>(v->brick_cnt, nb, nullptr, v->live, v->scan_sums, nullptr);
    brick_scan_apply_kernel<<<nblk, BSCAN, 0, s>>>(v->brick_cnt, nb, nullptr, v->live, v->active, v->scan_sums, nullptr,
                                                   v->scratch + 4);
    active_desc_kernel<<<(nb + 255) / 256, 256, 0, s>>>(v->geom, v->active, v->scratch + 4, v->brick_cnt, v->brick_off,
                                                        v->brick_slot, v->brick_box, v->active_desc);
    OFX_LAUNCH_CHECK();
    OFX_CUDA(cudaMemsetAsync(v->node_hist, 0, 64 * sizeof(int), s));
    node_hist_kernel<<<std::min((nb + 255) / 256, 1024), 256, 0, s>>>(v->brick_cnt, nb, v->node_hist);
    OFX_LAUNCH_CHECK();
    OFX_CUDA(cudaMemcpyAsync(v->host_hist, v->node_hist, 64 * sizeof(int), cudaMemcpyDeviceToHost, s));
    // the results (error flag, number of active bricks) travel to page-locked host memory behind the kernels; nobody waits
    // here -- graph_finish() reads them when the next call needs them
    OFX_CUDA(cudaMemcpyAsync(v->host_res, v->scratch, 8 * sizeof(int), cudaMemcpyDeviceToHost, s));
    OFX_CUDA(cudaEventRecord(v->graph_event, s));
    v->graph_pending = true;
    v->n_active = 0;
    v->counters_dirty[0] = v->counters_dirty[1] = false;       // scratch was cleared above
    v->graph_ready = true;
    return 0;
}

// Row-table capacity of the persistent kernel for a graph.  The kernel's time is a step function of its dynamic shared
// memory (what a CTA does not ask for is L1 cache for the depth gathers: measured on 512^3 dense, 0.469 ms up to 14
// entries per row, 0.477 at 16, 0.498 at 32), while a brick with more nodes than the capacity reads its node transforms
// directly (12 % slower on the band workloads when that is the rule).  So: the smallest capacity, in steps of 2 from
// 12, that covers 97 % of the active bricks.
static int choose_tab_cap(const int *hist, int n_active)
{
    if (n_active <= 0) return ROW_TAB_CAP;
    long long covered = 0;
    int cap = 0;
    for (int c = 0; c <= ROW_TAB_CAP; ++c) {
        covered += hist[c];
        cap = c;
        if (c >= 12 && c % 2 == 0 && covered * 100 >= (long long)n_active * 97) break;
    }
    return std::min(ROW_TAB_CAP, std::max(12, (cap + 1) & ~1));
}

// The tail of ofx_set_graph, run by the first call that needs its results: waits for the set-up kernels (long done in a
// frame loop), reports the brick-node overflow, publishes the number of active bricks.
int ofx::graph_finish(ofx_volume *v)
{
    if (!v->graph_pending) return 0;
    v->graph_pending = false;
    OFX_CUDA(cudaEventSynchronize(v->graph_event));
    const int *res = v->host_res;
    if (res[2] != 0) {
        v->graph_ready = false;
        return fail("ofx_set_graph: a brick of %dx%dx%d voxels uses %d distinct anchor nodes (limit %d); "
                    "nodes are far denser than the voxel grid", BX, BY, BZ, res[2], MAX_BRICK_NODES);
    }
    v->n_active = res[4];
    v->tab_cap = choose_tab_cap(v->host_hist, v->n_active);
    int max_used = v->host_hist[ROW_TAB_CAP + 2];
    for (int c = ROW_TAB_CAP; c > 0 && max_used == 0; --c)
        if (v->host_hist[c]) max_used = c;
    v->max_used = max_used;
    if (const char *e = getenv("OFX_TAB_CAP")) v->tab_cap = std::min(ROW_TAB_CAP, std::max(1, atoi(e)));
    if (getenv("OFX_DEBUG"))
        fprintf(stderr, "[ofx] graph: %d active bricks, largest brick node count %d -> row tables of %d entries\n", v->n_active, max_used, v->tab_cap);
    return 0;
}
