// ofx_math.cuh -- the contract arithmetic of the fusion hot path as device functions.
//
// Compiled with --fmad=false: every fused multiply-add below is an explicit __fmaf_rn, every
// other float/double operation is a single IEEE round-to-nearest op (nvcc defaults
// -prec-div=true -prec-sqrt=true -ftz=false), so results are bit-identical to the reference
// op order restated in DESIGN.md ("contract arithmetic").  Reference lines are relative to
// /root/reference/fusion_with_occlusion/.
#pragma once
#ifndef OFX_SLEEP_NS
#define OFX_SLEEP_NS 100
#endif
#ifndef OFX_KEEP_ONE
#define OFX_KEEP_ONE 1
#endif
#include <cstdint>
#include <cuda_runtime.h>

#define OFX_MAX_K 8

namespace ofx {

// float32 exp with a fixed op sequence: Cody-Waite reduction by ln2 (hi/lo), degree-5 Cephes
// polynomial, exact scale by 2^n.  Stands in for numpy's float32 exp in warpfield.py:114
// (neither numpy's nor CUDA's expf is correctly rounded, so the contract fixes its own).
__device__ __forceinline__ float expf_det(float x)
{
    if (!(x > -87.0f)) return 0.0f;
    const float n = rintf(x * 1.44269504088896341f);
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860682030941723e-06f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    const float y = __fmaf_rn(p, r * r, r) + 1.0f;
    return y * __int_as_float(((int)n + 127) << 23);
}

// expf_det for -87 < x <= 0 (the cached-voxel path: x >= -8.0001 because every anchor lies inside
// 4*coverage).  Same value, fewer issue slots: rintf through the 1.5*2^23 magic constant
// (round-half-even, exact for |t| < 2^22) and the power-of-two scale as an exponent-field add
// (y in (0.4, 1], n >= -126 + 2: the product is normal, so adding n << 23 to the bits of y IS y * 2^n).
__device__ __forceinline__ float expf_det_neg(float x)
{
    const float tm = x * 1.44269504088896341f + 12582912.0f;
    const float n = tm - 12582912.0f;
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860682030941723e-06f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    const float y = __fmaf_rn(p, r * r, r) + 1.0f;
    return __int_as_float(__float_as_int(y) + (__float_as_int(tm) << 23));
}

// pykdtree's float32 squared distance: ((dx*dx)+(dy*dy))+(dz*dz)   (warpfield.py:104)
__device__ __forceinline__ float dist2(float px, float py, float pz, float gx, float gy, float gz)
{
    const float dx = px - gx, dy = py - gy, dz = pz - gz;
    return ((dx * dx) + (dy * dy)) + (dz * dz);
}

// Sorted top-K insertion, ascending by (d2, arrival order).  Candidates must arrive in
// ascending node index so that ties keep the lower index first (strict <).
template <int K>
__device__ __forceinline__ void topk_insert(float d2, int j, float (&bd)[K], int (&bi)[K])
{
    if (d2 < bd[K - 1]) {
        bd[K - 1] = d2;
        bi[K - 1] = j;
#pragma unroll
        for (int k = K - 1; k > 0; --k) {
            if (bd[k] < bd[k - 1]) {
                const float td = bd[k]; bd[k] = bd[k - 1]; bd[k - 1] = td;
                const int ti = bi[k]; bi[k] = bi[k - 1]; bi[k - 1] = ti;
            }
        }
    }
}

// np.sum(weights, axis=-1) order (warpfield.py:121): sequential for K < 8, numpy's 8-lane
// pairwise tree for K == 8.
template <int K>
__device__ __forceinline__ float sum_k(const float (&w)[K])
{
    if constexpr (K < 8) {
        float s = w[0];
#pragma unroll
        for (int k = 1; k < K; ++k) s = s + w[k];
        return s;
    } else {
        return ((w[0] + w[1]) + (w[2] + w[3])) + ((w[4] + w[5]) + (w[6] + w[7]));
    }
}

// Skinning weights from the K sorted squared distances (warpfield.py:104-121).
// Returns true when all K anchors lie within cutoff = float32(4*coverage).
template <int K>
__device__ __forceinline__ bool skin_weights(const float (&bd)[K], float cutoff, float denom,
                                             float (&w)[K], bool (&in_range)[K])
{
    bool ok = true;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const float d = sqrtf(bd[k]);
        in_range[k] = !(d > cutoff);
        ok = ok && in_range[k];
        w[k] = in_range[k] ? expf_det(-(d * d) / denom) : 0.0f;
    }
    const float s = sum_k<K>(w) + 1e-6f;
#pragma unroll
    for (int k = 0; k < K; ++k) w[k] = w[k] / s;
    return ok;
}

// ---- exact division through a correctly rounded reciprocal ---------------------------------------
// a / b rounded to nearest, given y = RN(1/b) (__frcp_rn / __drcp_rn, or 1/b computed once):
// Markstein's correction applied twice (q <- q + RN(a - b q) * y; the residual is exact in FMA
// arithmetic).  Bit-identical to IEEE division for normal operands (checked against `/` on
// 2e8 random pairs per type and on the whole colour/weight grid: tests/test_host_logic.py and
// every GPU-vs-oracle test, the oracle uses plain `/`); the result for a = -0 is +0, which no
// caller can observe.  ~5 issue slots instead of ~10 + a slow-path branch, and one reciprocal
// serves every division that shares the divisor.
__device__ __forceinline__ float div_rn(float a, float b, float y)
{
    float q = a * y;
    float r = __fmaf_rn(-b, q, a);
    q = __fmaf_rn(r, y, q);
    r = __fmaf_rn(-b, q, a);
    return __fmaf_rn(r, y, q);
}
__device__ __forceinline__ double div_rn(double a, double b, double y)
{
    double q = a * y;
    double r = __fma_rn(-b, q, a);
    q = __fma_rn(r, y, q);
    r = __fma_rn(-b, q, a);
    return __fma_rn(r, y, q);
}

// skin_weights for a voxel already known to have all K anchors inside the cutoff (anchor cache):
// same values as skin_weights(), divisions through shared reciprocals.  rdenom = RN(1/denom).
template <int K>
__device__ __forceinline__ void skin_weights_valid(const float (&bd)[K], float denom, float rdenom, float (&w)[K])
{
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const float d = sqrtf(bd[k]);
        w[k] = expf_det_neg(div_rn(-(d * d), denom, rdenom));
    }
    const float s = sum_k<K>(w) + 1e-6f;
    const float rs = __frcp_rn(s);
#pragma unroll
    for (int k = 0; k < K; ++k) w[k] = div_rn(w[k], s, rs);
}

// One anchor's contribution to the blend: acc += w * (R p + T)   (warpfield.py:225-226).
// rec = {R00..R22 row-major, T0, T1, T2}.
__device__ __forceinline__ void lbs_accumulate(const float *__restrict__ rec, float w, float px,
                                               float py, float pz, float &ax, float &ay, float &az)
{
    const float qx = __fmaf_rn(rec[2], pz, __fmaf_rn(rec[1], py, __fmaf_rn(rec[0], px, rec[9])));
    const float qy = __fmaf_rn(rec[5], pz, __fmaf_rn(rec[4], py, __fmaf_rn(rec[3], px, rec[10])));
    const float qz = __fmaf_rn(rec[8], pz, __fmaf_rn(rec[7], py, __fmaf_rn(rec[6], px, rec[11])));
    ax = __fmaf_rn(w, qx, ax);
    ay = __fmaf_rn(w, qy, ay);
    az = __fmaf_rn(w, qz, az);
}

struct Camera {
    float fx, fy, cx, cy;   // already float32 (tsdf.py:357-359)
    int H, W;
};

// cam2pix + frustum test (tsdf.py:351-364, 589-593): float64, multiply before divide,
// round half to even.  false when z <= 0 or the pixel falls outside the image.
__device__ __forceinline__ bool project(float qx, float qy, float qz, const Camera &cam, int &u, int &v)
{
    if (!(qz > 0.0f)) return false;
    const double z = (double)qz;
    const double rz = __drcp_rn(z);                  // one reciprocal, two exact divisions
    const double fu = rint(div_rn((double)qx * (double)cam.fx, z, rz) + (double)cam.cx);
    const double fv = rint(div_rn((double)qy * (double)cam.fy, z, rz) + (double)cam.cy);
    if (!(fu >= 0.0 && fu < (double)cam.W && fv >= 0.0 && fv < (double)cam.H)) return false;
    u = (int)fu;
    v = (int)fv;
    return true;
}

// Certified float32 evaluation of the same pixel.  m = fl(fl(qx*fx) * rcp(qz)) carries a relative
// error <= 2^-24 + 2^-23 (MUFU.RCP, 1 ulp) + 2^-24 < 2.5e-7 against qx*fx/qz, u = fl(m + cx) adds
// <= 2^-24 |u| <= 6e-8 (|m| + |c|): the float32 u lies within eps = 3.2e-7 |m| + eps0 of the float64
// value the reference rounds (eps0 = 6.2e-8 |c| + 1e-6 also covers the reference's own 1e-12).
// If u is farther than eps from a half-integer, rint(u) IS the reference pixel; otherwise (about
// 1e-4 of the voxels), and for any NaN/Inf, the caller evaluates project().
struct FastCam {
    float fx, fy, cx, cy;
    float eps0x, eps0y;     // 6.2e-8 |cx| + 1e-6, 6.2e-8 |cy| + 1e-6
    int W, H;
};

__device__ __forceinline__ bool fast_pixel(float q, float f, float c, float eps0, float rz, int lim, int &ni,
                                           bool &inside)
{
    const float m = (q * f) * rz;
    const float u = m + c;
    const float tm = u + 12582912.0f;                         // 1.5 * 2^23: the sum rounds u half-to-even
    const float n = tm - 12582912.0f;                         // rint(u) for |u| < 2^22
    const float frac = u - n;                                 // exact
    const float eps = __fmaf_rn(fabsf(m), 3.2e-7f, eps0);     // > 0.5 long before the magic constant breaks down
    // low mantissa bits of tm = rint(u); anything outside [-0.5, lim - 0.5] (NaN, Inf included) maps
    // to a value that fails the unsigned compare
    ni = __float_as_int(tm) - 0x4B400000;
    inside = (unsigned)ni < (unsigned)lim;
    return fabsf(frac) + eps < 0.5f;                          // false for NaN
}

// 0 = certainly no pixel, 1 = pixel (u, v) certain, 2 = undecided (use project()).
__device__ __forceinline__ int project_fast(float qx, float qy, float qz, const FastCam &fc, int &u, int &v)
{
    // straight-line on purpose: the caller evaluates four voxels back to back and the scheduler
    // interleaves their dependency chains
    float rz;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rz) : "f"(qz));
    bool in_u, in_v;
    const bool ok_u = fast_pixel(qx, fc.fx, fc.cx, fc.eps0x, rz, fc.W, u, in_u);
    const bool ok_v = fast_pixel(qy, fc.fy, fc.cy, fc.eps0y, rz, fc.H, v, in_v);
    // z <= 0 has no pixel; a coordinate that is certainly outside decides the voxel even if the other
    // one is undecided
    const bool dead = !(qz > 0.0f) || (ok_u && !in_u) || (ok_v && !in_v);
    return dead ? 0 : ((ok_u && ok_v) ? 1 : 2);
}

// The same decision as two flags: in_certain = the pixel (u, v) is certain and inside the image; undecided = the
// caller must evaluate project() (a superset of project_fast()'s state 2: a coordinate that is certainly outside
// no longer short-cuts the other one -- a handful of predicate instructions less on the hot path, and project()
// gives the exact answer for every voxel sent to it).
__device__ __forceinline__ void project_fast2(float qx, float qy, float qz, const FastCam &fc, int &u, int &v, bool &in_certain,
                                              bool &undecided)
{
    // fast_pixel() on (u, v) as one pair of packed float32 lanes (FMUL2 / FADD2).  Whether ptxas contracts the second
    // product with the following sum does not matter here: the certificate only needs the float32 value within eps of
    // the float64 one, and a fused multiply-add is closer to it, not farther.
    float rz;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rz) : "f"(qz));
    const float2 m2 = __fmul2_rn(__fmul2_rn(make_float2(qx, qy), make_float2(fc.fx, fc.fy)), make_float2(rz, rz));
    const float2 u2 = __fadd2_rn(m2, make_float2(fc.cx, fc.cy));
    const float2 tm2 = __fadd2_rn(u2, make_float2(12582912.0f, 12582912.0f));     // 1.5 * 2^23: the sum rounds u half-to-even
    const float2 n2 = __fadd2_rn(tm2, make_float2(-12582912.0f, -12582912.0f));   // rint(u) for |u| < 2^22
    const float2 fr2 = __fadd2_rn(u2, make_float2(-n2.x, -n2.y));                 // exact
    const float eps_u = __fmaf_rn(fabsf(m2.x), 3.2e-7f, fc.eps0x), eps_v = __fmaf_rn(fabsf(m2.y), 3.2e-7f, fc.eps0y);
    u = __float_as_int(tm2.x) - 0x4B400000;
    v = __float_as_int(tm2.y) - 0x4B400000;
    const bool in = (unsigned)u < (unsigned)fc.W && (unsigned)v < (unsigned)fc.H;
    const bool ok = fabsf(fr2.x) + eps_u < 0.5f && fabsf(fr2.y) + eps_v < 0.5f;      // false for NaN
    const bool pos = qz > 0.0f;
    in_certain = pos && ok && in;
    undecided = pos && !ok;
}

// depth - z >= -trunc in float64 (tsdf.py:608-609), for the rare float32 ties
__device__ __forceinline__ bool visible_exact(float d, float z, double trunc)
{
    return (double)d - (double)z >= -trunc;
}

// Running-average update of one voxel (tsdf.py:466, 373-375, 479-493).
// diff = depth - z (float64), already known to satisfy depth > 0 && diff >= -trunc.
struct FuseArgs {
    double trunc, rtrunc, obs;   // rtrunc = RN(1/trunc), computed once on the host
    float obsf;
    int obs_is_f32;              // obs == (double)obsf: then f32(f64(w) + obs) == w + obsf (see below)
    float trunc_up;              // smallest float32 > trunc
};

// byte k of an integer-valued colour word as float: bits 0x4B0000bb are 2^23 + bb
__device__ __forceinline__ float byte_to_float(uint32_t word, uint32_t selector)
{
    return __uint_as_float(__byte_perm(word, 0x4B000000u, selector)) - 8388608.0f;
}

// OBS32: obs_weight is a float32 value (fu.obs_is_f32; the reference's default 1.0 is) -- a compile-time
// switch so that the float64 weight sum of the general case does not ride along in the hot loop.
template <bool OBS32>
__device__ __forceinline__ void fuse_voxel(float &tsdf, float &weight, float &colour, float d, float z, float new_colour,
                                           const FuseArgs &fu)
{
    const float w_old = weight;
    // f32(f64(w) + obs): when obs is a float32 value the float64 sum is either exact (exponents within
    // 29 bits) or the small operand is far below half an ulp32 of the large one, so the single float32
    // addition rounds identically
    float w_new;
    if constexpr (OBS32) w_new = w_old + fu.obsf;
    else w_new = (float)((double)w_old + fu.obs);
    // Free space seen again -- the bulk of all updates: dist = 1 and tsdf = 1.  fl32(d - z) above the
    // float32 successor of trunc proves d - z >= trunc (rounding is monotone), i.e. dist = 1; then
    // num = w + obs (exact in float64) and the quotient is exactly 1 iff fl32(w + obs) was exact
    // (Fast2Sum, |w| >= |obs|), otherwise the float64 expression is evaluated.
    const float sum_err = fu.obsf - (w_new - w_old);
    const bool keep_one = OFX_KEEP_ONE && OBS32 && (d - z) > fu.trunc_up && tsdf == 1.0f && w_old >= fabsf(fu.obsf) && sum_err == 0.0f &&
                          w_new != 0.0f;
    if (!keep_one) {
        const double diff = (double)d - (double)z;                      // tsdf.py:608, exact
        const double dist = diff >= fu.trunc ? 1.0 : fmin(1.0, div_rn(diff, fu.trunc, fu.rtrunc));
        const float prod = w_old * tsdf;
        const double num = (double)prod + fu.obs * dist, den = (double)w_new;
        tsdf = (num == den && den != 0.0) ? 1.0f : (float)div_rn(num, den, __drcp_rn(den));
    }
    weight = w_new;

    // colour channels (tsdf.py:479-493).  The packed words are integers below 2^24 with byte-sized
    // channels, so floor(c/65536) etc. are the bytes of the integer; the blend is the reference's
    // float32 expression; rintf through the 1.5*2^23 constant (values < 2^22); repacking is exact.
    const uint32_t oc = (uint32_t)colour, nc = (uint32_t)new_colour;
    const float ob = byte_to_float(oc, 0x7442u), og = byte_to_float(oc, 0x7441u), orr = byte_to_float(oc, 0x7440u);
    const float nb = byte_to_float(nc, 0x7442u), ng = byte_to_float(nc, 0x7441u), nr = byte_to_float(nc, 0x7440u);
    const float rw = __frcp_rn(w_new);
    const float b = fminf(255.0f, (div_rn(w_old * ob + fu.obsf * nb, w_new, rw) + 12582912.0f) - 12582912.0f);
    const float g = fminf(255.0f, (div_rn(w_old * og + fu.obsf * ng, w_new, rw) + 12582912.0f) - 12582912.0f);
    const float r = fminf(255.0f, (div_rn(w_old * orr + fu.obsf * nr, w_new, rw) + 12582912.0f) - 12582912.0f);
    colour = __fmaf_rn(b, 65536.0f, __fmaf_rn(g, 256.0f, r));
}

// ---- programmatic dependent launch (the per-frame sequence node transforms -> cull -> fused kernel) ----
// grid_dep_wait: block until the preceding kernel in the stream has completed and its writes are visible (returns
// at once when the kernel was launched without the attribute); grid_dep_launch: the next kernel's CTAs may be
// scheduled as soon as there is room (they wait in grid_dep_wait).
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- TMA (1-D bulk copy) + mbarrier helpers ------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    // plain try_wait loop.  (A suspend-time hint compiles to NANOSLEEP.SYNCS plus a reload of the
    // parity from local memory per poll -- measured: no faster, and the reloads cost L1 wavefronts.)
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// named barriers (bar.arrive / bar.sync): consumers signal without blocking, the producer blocks in
// hardware without taking issue slots.  `threads` = arriving + waiting threads, a multiple of 32.
__device__ __forceinline__ void named_bar_arrive(int id, int threads)
{
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// the same wait for a warp that expects to wait long (the producer): sleep between polls
__device__ __forceinline__ void mbar_wait_sleep(uint64_t *bar, uint32_t parity)
{
    uint32_t done = 0;
    while (true) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) break;
        __nanosleep(OFX_SLEEP_NS);
    }
}
// global -> shared bulk copy through the TMA unit (SASS: UBLKCP); bytes % 16 == 0, both
// addresses 16-byte aligned; completion is signalled on `bar` as transaction bytes.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// asynchronous DRAM -> L2 prefetch of a contiguous range (bytes % 16 == 0, 16-byte aligned)
__device__ __forceinline__ void bulk_prefetch_l2(const void *src_gmem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}

}  // namespace ofx
