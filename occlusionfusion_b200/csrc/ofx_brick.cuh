// ofx_brick.cuh -- device code shared by the brick kernels of ofx_integrate.cu: argument structs,
// the per-voxel warp (cached anchors -> weights -> linear blend), brick-level frustum culling and
// the per-thread body (4 consecutive z voxels: warp, project, test, fuse, store).
#pragma once
#include <cmath>
#include <type_traits>

#include "ofx_volume.cuh"

namespace ofx {

enum Mode { MODE_WARPED = 0, MODE_IDENTITY = 1, MODE_DEFORM = 2 };

struct FrameArgs {
    const float *depth;
    const float *colour_im;
    const float2 *dc;            // the same frame interleaved {depth, packed colour} (frame_prep_kernel); persistent kernel only
    Camera cam;
    FastCam fc;                  // the same intrinsics + the error terms of the certified float32 pixel
    double trunc;
    FuseArgs fu;
    float ntrunc_up, ntrunc_dn;  // the two float32 neighbours of -trunc (up >= -trunc >= dn)
};

// host: FrameArgs for one frame
static inline FrameArgs make_frame_args(const float *depth, const float *colour, const float *intr, int H, int W,
                                        double trunc, double obs)
{
    FrameArgs fa{};
    fa.depth = depth; fa.colour_im = colour;
    fa.cam = Camera{intr[0], intr[1], intr[2], intr[3], H, W};
    fa.fc = FastCam{intr[0], intr[1], intr[2], intr[3], 6.2e-8f * fabsf(intr[2]) + 1e-6f,
                    6.2e-8f * fabsf(intr[3]) + 1e-6f, W, H};
    fa.trunc = trunc;
    float tu = (float)trunc;
    if (!((double)tu > trunc)) tu = nextafterf(tu, INFINITY);
    fa.fu = FuseArgs{trunc, 1.0 / trunc, obs, (float)obs, (double)(float)obs == obs ? 1 : 0, tu};
    const float t = (float)(-trunc);
    if ((double)t >= -trunc) { fa.ntrunc_up = t; fa.ntrunc_dn = ((double)t == -trunc) ? t : nextafterf(t, -INFINITY); }
    else { fa.ntrunc_dn = t; fa.ntrunc_up = nextafterf(t, INFINITY); }
    return fa;
}

struct GraphArgs {
    const float *node_rec;
    const int *active;       // work list (MODE_WARPED) or nullptr (all bricks)
    const int *off;
    const int *cnt;
    const int *brick_nodes;
    const int *brick_slot;   // cache slot of a brick
    const uint8_t *cache;    // skin cache, slot_bytes_of(Kp) per slot (layout: ofx_volume.cuh)
    float cutoff, denom;
};

// Node transforms {R (9), T (3)} sit in shared memory at a stride of 12 floats (48 B): consecutive
// records start 12 banks apart, so the eight float4 loads of a quarter-warp on eight consecutive
// records hit eight distinct 4-bank groups.
constexpr int REC_STRIDE = NODE_XFORM_FLOATS;
constexpr uint32_t REC_BYTES = NODE_XFORM_FLOATS * 4u;

// q = sum_k w_k (R_k p + T_k) with the cached anchors and weights of one voxel (warpfield.py:225-226).
// (The reference skips w == 0, warpfield.py:219: cached voxels have w >= exp(-8.0001) / (K + 1e-6) > 0.)
template <int K, int KP>
__device__ __forceinline__ void warp_voxel(const uint32_t (&ids)[2], const float (&w)[KP], uint32_t rec_base, float px,
                                           float py, float pz, float &qx, float &qy, float &qz)
{
    // rec_base: 32-bit shared-memory address of the brick's first record; one PRMT (byte k of the ids
    // word) and one IMAD per anchor, then three 16-byte loads at immediate offsets.  The loads are
    // `asm volatile`: the bytes are written by the TMA unit / by build_row_table, invisibly to the
    // compiler, which must neither reuse a value from the previous brick nor move a load above them.
    float ax = 0.f, ay = 0.f, az = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const uint32_t addr = __byte_perm(ids[k >> 2], 0u, 0x4440u + (k & 3)) * REC_BYTES + rec_base;
        float r12[12];
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r12[0]), "=f"(r12[1]), "=f"(r12[2]), "=f"(r12[3]) : "r"(addr));
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+16];" : "=f"(r12[4]), "=f"(r12[5]), "=f"(r12[6]), "=f"(r12[7]) : "r"(addr));
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+32];" : "=f"(r12[8]), "=f"(r12[9]), "=f"(r12[10]), "=f"(r12[11]) : "r"(addr));
        lbs_accumulate(r12, w[k], px, py, pz, ax, ay, az);
    }
    qx = ax; qy = ay; qz = az;
}

// ---- per-warp row tables ---------------------------------------------------------------------------
// The reference evaluates R p + T as fma(R_2, p_z, fma(R_1, p_y, fma(R_0, p_x, T))) per component
// (ofx_math.cuh: lbs_accumulate).  The inner two FMAs depend only on the node and on the voxel ROW
// (x, y); a warp owns four rows of a brick, eight lanes per row.  So each warp first tabulates, for its
// rows and the brick's nodes, B = fma(R_1, p_y, fma(R_0, p_x, T)) together with the z column of R, and
// the per-voxel work becomes q_k = fma(R_2, p_z, B): the same operations in the same order (bit-identical),
// half the FMAs and half the shared-memory bytes per anchor (24 instead of 48) -- shared-memory wavefronts
// are what bounds the kernel (profiles/: l1tex data pipe at 90 %).
//   tab   float4 [4 rows][ROW_TAB_CAP]   {B_x, B_y, B_z, R_22}
//   col   float2 [ROW_TAB_CAP]           {R_02, R_12}
// ({B_x, B_y} and {R_02, R_12} sit in aligned register pairs after the loads: the x and y components of
// fma(R_2, p_z, B) are ONE packed fma.rn.f32x2, FFMA2 in SASS.)
// Bricks with more than ROW_TAB_CAP nodes take the record path (warp_voxel).
// (ROW_TAB_CAP = 32: ofx_volume.cuh)
constexpr int ROW_TAB_BYTES = 4 * ROW_TAB_CAP * 16 + ROW_TAB_CAP * 8;      // per warp, at the largest capacity
// The persistent kernel sizes its tables per LAUNCH (tab_cap <= ROW_TAB_CAP entries per row, chosen per graph: see
// choose_tab_cap in ofx_skin_cache.cu): every KB of shared memory a CTA does not ask for goes to the L1 cache.
__host__ __device__ constexpr int row_tab_bytes(int cap) { return 4 * cap * 16 + cap * 8; }

__device__ __forceinline__ void build_row_table(uint32_t rec_base, int n_used, float px, float py, uint32_t tab_base, int lane,
                                                int cap = ROW_TAB_CAP)
{
    const int r = lane >> 3;
    for (int n = lane & 7; n < n_used; n += 8) {
        const uint32_t addr = (uint32_t)n * REC_BYTES + rec_base;
        float a[12];
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a[0]), "=f"(a[1]), "=f"(a[2]), "=f"(a[3]) : "r"(addr));
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+16];" : "=f"(a[4]), "=f"(a[5]), "=f"(a[6]), "=f"(a[7]) : "r"(addr));
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+32];" : "=f"(a[8]), "=f"(a[9]), "=f"(a[10]), "=f"(a[11]) : "r"(addr));
        const float bx = __fmaf_rn(a[1], py, __fmaf_rn(a[0], px, a[9]));
        const float by = __fmaf_rn(a[4], py, __fmaf_rn(a[3], px, a[10]));
        const float bz = __fmaf_rn(a[7], py, __fmaf_rn(a[6], px, a[11]));
        const uint32_t ent = tab_base + (uint32_t)(r * cap + n) * 16u;
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(ent), "f"(bx), "f"(by), "f"(bz), "f"(a[8]) : "memory");
        if (r == 0)
            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(tab_base + 4u * (uint32_t)cap * 16u + (uint32_t)n * 8u), "f"(a[2]),
                         "f"(a[5])
                         : "memory");
    }
    __syncwarp();
}

// q = sum_k w_k fma(R_2k, p_z, B_k) through the warp's row table; row_base = table row of this lane
template <int K, int KP>
__device__ __forceinline__ void warp_voxel_tab(const uint32_t (&ids)[2], const float (&w)[KP], uint32_t row_base,
                                               uint32_t col_base, float pz, float &qx, float &qy, float &qz)
{
    // (a scalar operand of a packed instruction is broadcast to both lanes for free: FFMA2 ..., R.F32, ...)
    float2 axy = make_float2(0.f, 0.f);
    float az = 0.f;
    const float2 pz2 = make_float2(pz, pz);
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const uint32_t id = __byte_perm(ids[k >> 2], 0u, 0x4440u + (k & 3));
        float bx, by, bz, r22, r02, r12;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(bx), "=f"(by), "=f"(bz), "=f"(r22) : "r"(id * 16u + row_base));
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(r02), "=f"(r12) : "r"(id * 8u + col_base));
        const float2 txy = __ffma2_rn(make_float2(r02, r12), pz2, make_float2(bx, by));     // fma(R_02, p_z, B_x), fma(R_12, p_z, B_y)
        axy = __ffma2_rn(make_float2(w[k], w[k]), txy, axy);
        az = __fmaf_rn(w[k], __fmaf_rn(r22, pz, bz), az);
    }
    qx = axy.x; qy = axy.y; qz = az;
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

// The same for an axis-aligned box of half extents (ex, ey, ez) around (bx, by, bz) in camera space.
__device__ __forceinline__ PixBox box_pixbox(float bx, float by, float bz, float ex, float ey, float ez, const Camera &cam)
{
    PixBox b;
    const float z_lo = bz - ez, z_hi = bz + ez;
    b.z_lo = z_lo;
    b.unbounded = !(z_lo > 1e-4f);
    const float xl = bx - ex, xh = bx + ex, yl = by - ey, yh = by + ey;
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
            const float *r = s_rec + i * REC_STRIDE;
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


// ---- per-thread body ---------------------------------------------------------------------------------
// Thread t of a brick owns voxels (lx0 + t/64, iy0 + (t/8)%8, iz0 + 4*(t%8) .. +3).  Counters are
// accumulated into c_upd / c_valid / c_wsum (the caller reduces them).
// `release` runs once per thread, warp-convergent, as soon as the staged node transforms and skin-cache
// slot are no longer needed (after phase 1): the persistent kernel hands the stage back to its producer there.
template <int K, bool VEC, int MODE, bool OBS32, typename Release>
__device__ __forceinline__ void brick_thread_work(const VolumeGeom &g, float *__restrict__ tsdf,
                                                  float *__restrict__ weight, float *__restrict__ colour,
                                                  const GraphArgs &ga, const FrameArgs &fa,
                                                  const float *__restrict__ s_rec, const uint8_t *__restrict__ s_slot,
                                                  uint8_t *s_tab, const BrickCoord &bc, int n_used, int t,
                                                  float *__restrict__ out_pts,
                                                  uint8_t *__restrict__ out_valid, unsigned &c_upd, unsigned &c_valid,
                                                  unsigned &c_wsum, Release release)
{
    const int tz = t & 7, ty = (t >> 3) & 7, tx = t >> 6;
    const int lx = bc.lx0 + tx, iy = bc.iy0 + ty, iz0 = bc.iz0 + tz * VOX_PER_THREAD;
    const bool row_live = lx < g.nx && iy < g.Dy && iz0 < g.Dz;
    const int64_t vox0 = ((int64_t)lx * g.Dy + iy) * g.Dz + iz0;
    const int n_vox = row_live ? min(VOX_PER_THREAD, g.Dz - iz0) : 0;

    // the brick's skin-cache slot sits in shared memory (one TMA bulk copy): per voxel one ids word and
    // one weight float4 per group of four anchors, read just in time (conflict-free: lane stride 4 / 16 B)
    constexpr int KP = K <= 4 ? 4 : 8;
    const uint32_t *s_ids = reinterpret_cast<const uint32_t *>(s_slot) + t;
    const float4 *s_w = reinterpret_cast<const float4 *>(s_slot + BRICK_VOXELS * KP) + t;
    // (x, y of the row even for lanes whose z lies beyond the volume: they help build the row table)
    const float px = __ldg(g.px + min(lx, g.nx - 1)), py = __ldg(g.py + min(iy, g.Dy - 1));
    const uint32_t rec_base = smem_u32(s_rec);

    // ---- phase 1: warp + pixel of each voxel (no memory traffic beyond shared memory) ----
    int pixs[VOX_PER_THREAD];
    float zsel[VOX_PER_THREAD];
    unsigned cand = 0;
#pragma unroll
    for (int e = 0; e < VOX_PER_THREAD; ++e) { pixs[e] = 0; zsel[e] = 0.f; }
    // s_tab != nullptr (persistent kernel): row tables when the brick's node list fits (brick-uniform)
    const bool use_tab = MODE != MODE_IDENTITY && s_tab != nullptr && n_used <= ROW_TAB_CAP;
    const uint32_t tab_base = use_tab ? smem_u32(s_tab) : 0u;
    if (use_tab) build_row_table(rec_base, n_used, px, py, tab_base, t & 31);
    const uint32_t row_base = tab_base + (uint32_t)((t & 31) >> 3) * (ROW_TAB_CAP * 16u);
    const uint32_t col_base = tab_base + 4u * ROW_TAB_CAP * 16u;
    auto phase1 = [&](auto tab_tag) {
    constexpr bool TAB = decltype(tab_tag)::value;
    if (row_live)
#pragma unroll
    for (int e = 0; e < VOX_PER_THREAD; ++e) {
        if (!VEC && e >= n_vox) continue;
        const float pz = __ldg(g.pz + iz0 + e);
        float qx = px, qy = py, qz = pz;
        bool skin_ok = true;                                          // source frame: all valid, tsdf.py:395-398
        if constexpr (MODE != MODE_IDENTITY) {
            uint32_t idw[2] = {0u, 0u};
            float wv[KP];
            if (MODE == MODE_WARPED || n_used > 0) {               // live bricks of the per-frame path always have nodes
#pragma unroll
                for (int gq = 0; gq < KP / 4; ++gq) {
                    idw[gq] = s_ids[(gq * 4 + e) * BRICK_THREADS];
                    const float4 w4 = s_w[(gq * 4 + e) * BRICK_THREADS];
                    wv[4 * gq] = w4.x; wv[4 * gq + 1] = w4.y; wv[4 * gq + 2] = w4.z; wv[4 * gq + 3] = w4.w;
                }
            } else {
                idw[0] = 0xFFFFFFFFu;
#pragma unroll
                for (int k = 0; k < KP; ++k) wv[k] = 0.0f;
            }
            skin_ok = (idw[0] & 0xFFu) != 0xFFu;
            // straight-line: a voxel without anchors blends record 0 with zero weights and is masked below
            if (!skin_ok) idw[0] = idw[1] = 0u;
            if constexpr (TAB) warp_voxel_tab<K, KP>(idw, wv, row_base, col_base, pz, qx, qy, qz);
            else warp_voxel<K, KP>(idw, wv, rec_base, px, py, pz, qx, qy, qz);
        }
        if constexpr (MODE == MODE_DEFORM) {
            if (!skin_ok) { qx = px; qy = py; qz = pz; }              // registration_fusion.py:161,180
            if (out_pts) { out_pts[3 * (vox0 + e)] = qx; out_pts[3 * (vox0 + e) + 1] = qy; out_pts[3 * (vox0 + e) + 2] = qz; }
            if (out_valid) out_valid[vox0 + e] = skin_ok ? 1 : 0;
        } else {
            int u, v;
            int st = project_fast(qx, qy, qz, fa.fc, u, v);           // tsdf.py:589-593, certified float32
            if (st == 2 && skin_ok) st = project(qx, qy, qz, fa.cam, u, v) ? 1 : 0;   // undecided (~1e-4): the float64 original
            c_valid += skin_ok ? 1u : 0u;                             // tsdf.py:464
            pixs[e] = v * fa.cam.W + u;
            zsel[e] = qz;
            if (skin_ok && st == 1) cand |= 1u << e;
        }
    }
    };
    if (use_tab) phase1(std::true_type{}); else phase1(std::false_type{});
    __syncwarp();
    release();
    if constexpr (MODE != MODE_DEFORM) {
        // ---- phase 2: the (up to four) depth gathers in flight together, then the visibility test ----
        if (!cand) return;
        float dsel[VOX_PER_THREAD];
#pragma unroll
        for (int e = 0; e < VOX_PER_THREAD; ++e) dsel[e] = (cand & (1u << e)) ? __ldg(fa.depth + pixs[e]) : 0.0f;
        unsigned upd = 0;
#pragma unroll
        for (int e = 0; e < VOX_PER_THREAD; ++e) {
            // depth > 0 and depth - z >= -trunc in float64 (tsdf.py:608-609).  fl32 is monotone: a float32
            // difference above / below the float32 neighbours of -trunc decides; only a tie needs float64
            const float d = dsel[e], df = d - zsel[e];
            bool vis = d > 0.0f && df >= fa.ntrunc_dn;
            if (vis && !(df > fa.ntrunc_up)) vis = visible_exact(d, zsel[e], fa.trunc);
            if (vis) upd |= 1u << e;
        }
        if (!upd) return;

        // ---- phase 3: volume words + new colours in flight together, fuse, store.  Only threads with an
        // update touch the volume (speculative loads before the depth test were measured: no gain) ----
        float tt[VOX_PER_THREAD], ww[VOX_PER_THREAD], cc[VOX_PER_THREAD], ncol[VOX_PER_THREAD];
        if (VEC) {
            const float4 t4 = *reinterpret_cast<const float4 *>(tsdf + vox0);
            const float4 w4 = *reinterpret_cast<const float4 *>(weight + vox0);
            const float4 c4 = *reinterpret_cast<const float4 *>(colour + vox0);
#pragma unroll
            for (int e = 0; e < VOX_PER_THREAD; ++e) ncol[e] = (upd & (1u << e)) ? __ldg(fa.colour_im + pixs[e]) : 0.0f;
            tt[0] = t4.x; tt[1] = t4.y; tt[2] = t4.z; tt[3] = t4.w;
            ww[0] = w4.x; ww[1] = w4.y; ww[2] = w4.z; ww[3] = w4.w;
            cc[0] = c4.x; cc[1] = c4.y; cc[2] = c4.z; cc[3] = c4.w;
        } else {
#pragma unroll
            for (int e = 0; e < VOX_PER_THREAD; ++e)
                if (upd & (1u << e)) {
                    tt[e] = tsdf[vox0 + e]; ww[e] = weight[vox0 + e]; cc[e] = colour[vox0 + e];
                    ncol[e] = __ldg(fa.colour_im + pixs[e]);
                }
        }
        float wsum = 0.0f;
#pragma unroll
        for (int e = 0; e < VOX_PER_THREAD; ++e)
            if (upd & (1u << e)) {
                fuse_voxel<OBS32>(tt[e], ww[e], cc[e], dsel[e], zsel[e], ncol[e], fa.fu);
                wsum += ww[e];
            }
        if (VEC) {
            *reinterpret_cast<float4 *>(tsdf + vox0) = make_float4(tt[0], tt[1], tt[2], tt[3]);
            *reinterpret_cast<float4 *>(weight + vox0) = make_float4(ww[0], ww[1], ww[2], ww[3]);
            *reinterpret_cast<float4 *>(colour + vox0) = make_float4(cc[0], cc[1], cc[2], cc[3]);
        } else {
#pragma unroll
            for (int e = 0; e < VOX_PER_THREAD; ++e)
                if (upd & (1u << e)) { tsdf[vox0 + e] = tt[e]; weight[vox0 + e] = ww[e]; colour[vox0 + e] = cc[e]; }
        }
        c_wsum += (unsigned)wsum;        // weights are integer-valued counters (obs_weight = 1)
        c_upd += __popc(upd);
    }
}

}  // namespace ofx
