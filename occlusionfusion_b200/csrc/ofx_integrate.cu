// ofx_integrate.cu -- the fused hot kernel of libofx.so (sm_100a):
//     per voxel: anchors (cached ids) -> skinning weights -> linear-blend warp -> projection ->
//     visibility test -> running-average TSDF / weight / colour update, in one pass.
// Replaces, per frame, WarpField.deform_tsdf (warpfield.py:369-380: skin_tsdf + deform) and
// TSDFVolume.integrate (tsdf.py:378-494) -- in the reference five full-volume host arrays and
// two device round trips; here 24 B of HBM traffic per UPDATED voxel plus K (padded to 4 or 8) bytes of
// anchor cache per voxel of an active brick, and nothing at all for the rest of the volume.
//
// Mapping: one CTA per active brick (4 x 8 x 32 voxels), 256 threads, each thread owns four
// consecutive voxels along z = one float4 of each of the three volumes (coalesced 128 B rows).
// The brick's node records (64 B each: position, rotation, translation) are gathered into shared
// memory by the TMA unit (cp.async.bulk, one copy per node, one mbarrier per CTA) while the
// threads fetch their anchor-cache words.  Threads whose four voxels all fail the skin /
// frustum / depth tests touch no volume memory.
#include "ofx_volume.cuh"

namespace ofx {

int sm_count();

enum Mode { MODE_WARPED = 0, MODE_IDENTITY = 1, MODE_DEFORM = 2 };

struct FrameArgs {
    const float *depth;
    const float *colour_im;
    Camera cam;
    double trunc, rtrunc, obs;   // rtrunc = RN(1/trunc)
    float obsf;
};

struct GraphArgs {
    const float *node_rec;
    const int *active;       // work list (MODE_WARPED) or nullptr (all bricks)
    const int *off;
    const int *cnt;
    const int *brick_nodes;
    const uint8_t *cache;
    float cutoff, denom;
};

// Node records sit in shared memory at a stride of 20 floats (80 B): consecutive records start
// 20 banks apart, so the eight float4 loads of a quarter-warp hit eight distinct 4-bank groups.
constexpr int REC_STRIDE = 20;

template <int K>
__device__ __forceinline__ void warp_voxel(const uint32_t (&ids)[2], const float *__restrict__ s_rec, float denom,
                                           float rdenom, float px, float py, float pz, float &qx, float &qy, float &qz)
{
    const float *rec[K];
    float bd[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const uint32_t id = (ids[k >> 2] >> (8 * (k & 3))) & 0xFFu;
        rec[k] = s_rec + id * REC_STRIDE;
        const float4 g4 = *reinterpret_cast<const float4 *>(rec[k]);
        bd[k] = dist2(px, py, pz, g4.x, g4.y, g4.z);
    }
    float w[K];
    skin_weights_valid<K>(bd, denom, rdenom, w);       // cached voxel: all K anchors are inside the cutoff
    float ax = 0.f, ay = 0.f, az = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (w[k] == 0.0f) continue;                                  // warpfield.py:219
        float r12[12];
        const float4 a = *reinterpret_cast<const float4 *>(rec[k] + 4);
        const float4 b = *reinterpret_cast<const float4 *>(rec[k] + 8);
        const float4 c = *reinterpret_cast<const float4 *>(rec[k] + 12);
        r12[0] = a.x; r12[1] = a.y; r12[2] = a.z; r12[3] = a.w; r12[4] = b.x; r12[5] = b.y; r12[6] = b.z; r12[7] = b.w;
        r12[8] = c.x; r12[9] = c.y; r12[10] = c.z; r12[11] = c.w;
        lbs_accumulate(r12, w[k], px, py, pz, ax, ay, az);
    }
    qx = ax; qy = ay; qz = az;
}

// ---- brick-level frustum culling -------------------------------------------------------------------
// Every voxel p of a brick lies within h of the brick centre c, rotations preserve distances,
// and the blend weights are non-negative with sum sigma in (0, 1]: the warped voxel is
// sigma * (a point of the convex hull of the balls B(R_k c + T_k, h) over the brick's nodes).
// Perspective projection is invariant under the positive scale sigma and maps convex sets in
// z > 0 to convex sets, so the pixel bounding box of those balls bounds every voxel's pixel.
// A brick whose box misses the image (1 px margin) cannot update anything and is skipped
// before it reads a byte.  Returns the bounds of ONE ball; the caller reduces over the nodes.
struct PixBox {
    float u_lo, u_hi, v_lo, v_hi, z_lo;
    int unbounded;
};

__device__ __forceinline__ PixBox ball_pixbox(float bx, float by, float bz, float h, const Camera &cam)
{
    PixBox b;
    const float z_lo = bz - h, z_hi = bz + h;
    b.z_lo = z_lo;
    b.unbounded = !(z_lo > 1e-4f);
    const float xl = bx - h, xh = bx + h, yl = by - h, yh = by + h;
    const float zl = b.unbounded ? 1.f : z_lo;
    b.u_lo = cam.fx * (xl / (xl >= 0.f ? z_hi : zl)) + cam.cx;
    b.u_hi = cam.fx * (xh / (xh >= 0.f ? zl : z_hi)) + cam.cx;
    b.v_lo = cam.fy * (yl / (yl >= 0.f ? z_hi : zl)) + cam.cy;
    b.v_hi = cam.fy * (yh / (yh >= 0.f ? zl : z_hi)) + cam.cy;
    return b;
}

__device__ __forceinline__ void brick_centre(const VolumeGeom &g, const BrickCoord &bc, float &cx, float &cy, float &cz,
                                             float &h)
{
    const int lx1 = min(bc.lx0 + BX, g.nx) - 1, iy1 = min(bc.iy0 + BY, g.Dy) - 1, iz1 = min(bc.iz0 + BZ, g.Dz) - 1;
    float x_lo = __ldg(g.px + bc.lx0), x_hi = x_lo;
    for (int lx = bc.lx0 + 1; lx <= lx1; ++lx) { const float x = __ldg(g.px + lx); x_lo = fminf(x_lo, x); x_hi = fmaxf(x_hi, x); }
    const float y_lo = __ldg(g.py + bc.iy0), y_hi = __ldg(g.py + iy1), z_lo = __ldg(g.pz + bc.iz0), z_hi = __ldg(g.pz + iz1);
    cx = 0.5f * (x_lo + x_hi); cy = 0.5f * (y_lo + y_hi); cz = 0.5f * (z_lo + z_hi);
    const float hx = 0.5f * (x_hi - x_lo), hy = 0.5f * (y_hi - y_lo), hz = 0.5f * (z_hi - z_lo);
    h = sqrtf(hx * hx + hy * hy + hz * hz) * 1.001f + 1e-5f;
}

// warp 0 only: true when no voxel of the brick can project into the image
__device__ __forceinline__ bool brick_outside_frustum(const VolumeGeom &g, const BrickCoord &bc, const float *s_rec,
                                                      int n_used, bool identity, const Camera &cam)
{
    float cx, cy, cz, h;
    brick_centre(g, bc, cx, cy, cz, h);
    const int lane = threadIdx.x & 31;
    float u_lo = 3.0e38f, u_hi = -3.0e38f, v_lo = 3.0e38f, v_hi = -3.0e38f;
    int unbounded = 0;
    const int n = identity ? 1 : n_used;
    for (int i = lane; i < n; i += 32) {
        float bx = cx, by = cy, bz = cz;
        if (!identity) {
            const float *r = s_rec + i * REC_STRIDE + 4;
            bx = r[0] * cx + r[1] * cy + r[2] * cz + r[9];
            by = r[3] * cx + r[4] * cy + r[5] * cz + r[10];
            bz = r[6] * cx + r[7] * cy + r[8] * cz + r[11];
        }
        const PixBox b = ball_pixbox(bx, by, bz, h, cam);
        unbounded |= b.unbounded;
        u_lo = fminf(u_lo, b.u_lo); u_hi = fmaxf(u_hi, b.u_hi);
        v_lo = fminf(v_lo, b.v_lo); v_hi = fmaxf(v_hi, b.v_hi);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unbounded |= __shfl_xor_sync(0xffffffffu, unbounded, o);
        u_lo = fminf(u_lo, __shfl_xor_sync(0xffffffffu, u_lo, o));
        u_hi = fmaxf(u_hi, __shfl_xor_sync(0xffffffffu, u_hi, o));
        v_lo = fminf(v_lo, __shfl_xor_sync(0xffffffffu, v_lo, o));
        v_hi = fmaxf(v_hi, __shfl_xor_sync(0xffffffffu, v_hi, o));
    }
    if (unbounded) return false;
    // pixel u is accepted when rint(u) in [0, W): u in [-0.5, W-0.5); keep a 1 px margin
    return u_hi < -1.5f || u_lo > (float)cam.W + 0.5f || v_hi < -1.5f || v_lo > (float)cam.H + 0.5f;
}

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

    const int b = (MODE == MODE_WARPED) ? ga.active[blockIdx.x] : (int)blockIdx.x;
    const BrickCoord bc = brick_coord(g, b);
    const int tz = threadIdx.x & 7, ty = (threadIdx.x >> 3) & 7, tx = threadIdx.x >> 6;
    const int lx = bc.lx0 + tx, iy = bc.iy0 + ty, iz0 = bc.iz0 + tz * VOX_PER_THREAD;
    const bool row_live = lx < g.nx && iy < g.Dy && iz0 < g.Dz;
    const int64_t vox0 = ((int64_t)lx * g.Dy + iy) * g.Dz + iz0;
    const int n_vox = row_live ? min(VOX_PER_THREAD, g.Dz - iz0) : 0;

    int n_used = 0;
    if constexpr (MODE != MODE_IDENTITY) {
        n_used = ga.cnt[b];
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    }
    if (threadIdx.x < 3) s_stat[threadIdx.x] = 0ull;
    __syncthreads();
    if constexpr (MODE != MODE_IDENTITY) {
        if (n_used > 0) {
            // TMA gather of the brick's node records: one 64-byte bulk copy per node
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
    if constexpr (MODE != MODE_DEFORM) {
        // brick-level frustum cull (warp 0 decides, everyone obeys)
        if (threadIdx.x < 32) {
            const bool out = brick_outside_frustum(g, bc, s_rec, n_used, MODE == MODE_IDENTITY, fa.cam);
            if (threadIdx.x == 0) s_cull = out ? 1 : 0;
        }
        __syncthreads();
        if (s_cull) return;
    }

    // anchor-cache words of this thread's voxels
    uint32_t ids[VOX_PER_THREAD][2];
#pragma unroll
    for (int e = 0; e < VOX_PER_THREAD; ++e) ids[e][0] = ids[e][1] = 0xFFFFFFFFu;
    if constexpr (MODE != MODE_IDENTITY) {
        if (n_used > 0 && row_live) {
            if constexpr (K <= 4) {
                const uint32_t *c = reinterpret_cast<const uint32_t *>(ga.cache) + vox0;
                if (VEC) {
                    const uint4 q = __ldg(reinterpret_cast<const uint4 *>(c));
                    ids[0][0] = q.x; ids[1][0] = q.y; ids[2][0] = q.z; ids[3][0] = q.w;
                } else {
                    for (int e = 0; e < n_vox; ++e) ids[e][0] = __ldg(c + e);
                }
            } else {
                const uint2 *c = reinterpret_cast<const uint2 *>(ga.cache) + vox0;
                for (int e = 0; e < n_vox; ++e) {
                    const uint2 q = __ldg(c + e);
                    ids[e][0] = q.x; ids[e][1] = q.y;
                }
            }
        }
    }

    float px = 0.f, py = 0.f;
    if (row_live) { px = __ldg(g.px + lx); py = __ldg(g.py + iy); }
    const float rdenom = (MODE != MODE_IDENTITY) ? __frcp_rn(ga.denom) : 0.f;

    unsigned upd = 0, n_valid = 0;
    double diff[VOX_PER_THREAD];
    float ncol[VOX_PER_THREAD];
#pragma unroll
    for (int e = 0; e < VOX_PER_THREAD; ++e) {
        if (e >= n_vox) continue;
        const float pz = __ldg(g.pz + iz0 + e);
        float qx = px, qy = py, qz = pz;
        bool skin_ok = true;                                          // source frame: all valid, tsdf.py:395-398
        if constexpr (MODE != MODE_IDENTITY) {
            skin_ok = (ids[e][0] & 0xFFu) != 0xFFu;
            if (skin_ok) warp_voxel<K>(ids[e], s_rec, ga.denom, rdenom, px, py, pz, qx, qy, qz);
        }
        if constexpr (MODE == MODE_DEFORM) {
            if (out_pts) { out_pts[3 * (vox0 + e)] = qx; out_pts[3 * (vox0 + e) + 1] = qy; out_pts[3 * (vox0 + e) + 2] = qz; }
            if (out_valid) out_valid[vox0 + e] = skin_ok ? 1 : 0;
        } else {
            if (!skin_ok) continue;                                   // tsdf.py:464
            ++n_valid;
            int u, v;
            if (!project(qx, qy, qz, fa.cam, u, v)) continue;         // tsdf.py:589-593
            const int64_t pix = (int64_t)v * fa.cam.W + u;
            const double d = (double)__ldg(fa.depth + pix);
            const double df = d - (double)qz;                         // tsdf.py:608
            if (!(d > 0.0 && df >= -fa.trunc)) continue;              // tsdf.py:609
            diff[e] = df;
            ncol[e] = __ldg(fa.colour_im + pix);
            upd |= 1u << e;
        }
    }

    if constexpr (MODE != MODE_DEFORM) {
        unsigned long long wsum = 0;
        if (upd) {
            float t[VOX_PER_THREAD], w[VOX_PER_THREAD], c[VOX_PER_THREAD];
            if (VEC) {
                const float4 t4 = *reinterpret_cast<const float4 *>(tsdf + vox0);
                const float4 w4 = *reinterpret_cast<const float4 *>(weight + vox0);
                const float4 c4 = *reinterpret_cast<const float4 *>(colour + vox0);
                t[0] = t4.x; t[1] = t4.y; t[2] = t4.z; t[3] = t4.w;
                w[0] = w4.x; w[1] = w4.y; w[2] = w4.z; w[3] = w4.w;
                c[0] = c4.x; c[1] = c4.y; c[2] = c4.z; c[3] = c4.w;
            } else {
#pragma unroll
                for (int e = 0; e < VOX_PER_THREAD; ++e)
                    if (upd & (1u << e)) { t[e] = tsdf[vox0 + e]; w[e] = weight[vox0 + e]; c[e] = colour[vox0 + e]; }
            }
#pragma unroll
            for (int e = 0; e < VOX_PER_THREAD; ++e)
                if (upd & (1u << e)) {
                    fuse_voxel(t[e], w[e], c[e], diff[e], ncol[e], fa.trunc, fa.rtrunc, fa.obs, fa.obsf);
                    wsum += (unsigned long long)w[e];
                }
            if (VEC) {
                *reinterpret_cast<float4 *>(tsdf + vox0) = make_float4(t[0], t[1], t[2], t[3]);
                *reinterpret_cast<float4 *>(weight + vox0) = make_float4(w[0], w[1], w[2], w[3]);
                *reinterpret_cast<float4 *>(colour + vox0) = make_float4(c[0], c[1], c[2], c[3]);
            } else {
#pragma unroll
                for (int e = 0; e < VOX_PER_THREAD; ++e)
                    if (upd & (1u << e)) { tsdf[vox0 + e] = t[e]; weight[vox0 + e] = w[e]; colour[vox0 + e] = c[e]; }
            }
        }
        // counters: warp shuffle -> shared -> one global atomic per CTA and counter
        unsigned long long s0 = __popc(upd), s1 = n_valid, s2 = wsum;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, o);
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        }
        if ((threadIdx.x & 31) == 0) {
            if (s0) atomicAdd(&s_stat[0], s0);
            if (s1) atomicAdd(&s_stat[1], s1);
            if (s2) atomicAdd(&s_stat[2], s2);
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

}  // namespace ofx

using namespace ofx;

static bool volume_vec_ok(const ofx_volume *v)
{
    const uintptr_t a = (uintptr_t)v->desc.tsdf | (uintptr_t)v->desc.weight | (uintptr_t)v->desc.colour |
                        (uintptr_t)v->anchor_cache;
    return v->geom.Dz % 4 == 0 && (a & 15) == 0;
}

template <int K, int MODE>
static int launch_fused(ofx_volume *v, int grid, const GraphArgs &ga, const FrameArgs &fa, unsigned long long *stats,
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

template <int MODE>
static int dispatch_k(ofx_volume *v, int grid, const GraphArgs &ga, const FrameArgs &fa, unsigned long long *stats,
                      float *out_pts, uint8_t *out_valid, cudaStream_t s)
{
    switch (v->K) {
    case 1: return launch_fused<1, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    case 2: return launch_fused<2, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    case 3: return launch_fused<3, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    case 4: return launch_fused<4, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    case 5: return launch_fused<5, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    case 6: return launch_fused<6, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    case 7: return launch_fused<7, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    default: return launch_fused<8, MODE>(v, grid, ga, fa, stats, out_pts, out_valid, s);
    }
}

static GraphArgs graph_args(const ofx_volume *v)
{
    return GraphArgs{v->node_rec, v->active, v->brick_off, v->brick_cnt, v->brick_nodes, v->anchor_cache, v->cutoff, v->denom};
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
        return launch_fused<4, MODE_IDENTITY>(v, v->n_bricks, ga, fa, st, nullptr, nullptr, s);
    }
    OFX_REQUIRE(v->graph_ready, "ofx_integrate: warped frame before ofx_set_graph");
    OFX_REQUIRE(v->transforms_ready, "ofx_integrate: warped frame before ofx_set_transforms");
    return dispatch_k<MODE_WARPED>(v, v->n_active, graph_args(v), fa, st, nullptr, nullptr, s);
}

extern "C" int ofx_volume_deform(ofx_volume *v, float *out_pts, uint8_t *valid, void *stream)
{
    OFX_REQUIRE(v, "ofx_volume_deform: null handle");
    OFX_REQUIRE(v->graph_ready, "ofx_volume_deform: call ofx_set_graph first");
    FrameArgs fa{};
    return dispatch_k<MODE_DEFORM>(v, v->n_bricks, graph_args(v), fa, nullptr, out_pts, valid, (cudaStream_t)stream);
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
