// ofx_volume.cu -- handle lifetime, geometry tables, node-record table, world points.
//   TSDFVolume.__init__ state      tsdf.py:127-141
//   TSDFVolume.vox2world           tsdf.py:338-349
//   WarpField.update_transformations (device side: upload of R and T)   warpfield.py:389-420
#include <cstring>
#include <new>
#include <vector>

#include "ofx_volume.cuh"

namespace ofx {

char *error_buffer()
{
    static thread_local char buf[512] = "";
    return buf;
}

// node_rec[j] = {g, 0, R, T}; g is written by ofx_set_graph, R|T by ofx_set_transforms.
__global__ void write_node_positions_kernel(const float *__restrict__ nodes, int M, float *__restrict__ rec)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= M) return;
    float *r = rec + (int64_t)j * NODE_REC_FLOATS;
    r[0] = nodes[3 * j]; r[1] = nodes[3 * j + 1]; r[2] = nodes[3 * j + 2]; r[3] = 0.0f;
    // identity transform until ofx_set_transforms is called (warpfield.py:52-53: R = I, T = 0)
    r[4] = 1.f; r[5] = 0.f; r[6] = 0.f; r[7] = 0.f; r[8] = 1.f; r[9] = 0.f; r[10] = 0.f; r[11] = 0.f; r[12] = 1.f;
    r[13] = 0.f; r[14] = 0.f; r[15] = 0.f;
}

__global__ void write_node_transforms_kernel(const float *__restrict__ R, const float *__restrict__ T, int M,
                                             float *__restrict__ rec)
{
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");    // the cull after this kernel waits for its completion
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= M * 12) return;
    const int j = t / 12, e = t % 12;
    rec[(int64_t)j * NODE_REC_FLOATS + 4 + e] = e < 9 ? R[9 * j + e] : T[3 * j + (e - 9)];
}

__global__ void fill_volume_kernel(float4 *__restrict__ tsdf, float4 *__restrict__ weight, float4 *__restrict__ colour,
                                   int64_t n4, float *tsdf_s, float *weight_s, float *colour_s, int64_t n)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
        tsdf[i] = make_float4(1.f, 1.f, 1.f, 1.f);
        weight[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        colour[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int64_t i = n4 * 4 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) {
        tsdf_s[i] = 1.f; weight_s[i] = 0.f; colour_s[i] = 0.f;
    }
}

__global__ void world_points_kernel(VolumeGeom g, int64_t n, float *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) {
        const int iz = (int)(i % g.Dz);
        const int iy = (int)((i / g.Dz) % g.Dy);
        const int lx = (int)(i / ((int64_t)g.Dz * g.Dy));
        out[3 * i] = g.px[lx]; out[3 * i + 1] = g.py[iy]; out[3 * i + 2] = g.pz[iz];
    }
}

int sm_count();

}  // namespace ofx

using namespace ofx;

extern "C" int ofx_abi_version(void) { return OFX_ABI_VERSION; }
extern "C" const char *ofx_last_error(void) { return error_buffer(); }

extern "C" int ofx_volume_create(ofx_volume **out, const ofx_volume_desc *d)
{
    OFX_REQUIRE(out && d, "ofx_volume_create: null argument");
    *out = nullptr;
    OFX_REQUIRE(d->dims[0] > 0 && d->dims[1] > 0 && d->dims[2] > 0, "ofx_volume_create: bad dims");
    OFX_REQUIRE(d->n_shards >= 1 && d->shard >= 0 && d->shard < d->n_shards, "ofx_volume_create: bad shard %d/%d",
                d->shard, d->n_shards);
    OFX_REQUIRE(d->x_block >= 1 && d->dims[0] % (d->x_block * d->n_shards) == 0,
                "ofx_volume_create: Dx=%d must be a multiple of x_block*n_shards=%d", d->dims[0],
                d->x_block * d->n_shards);
    OFX_REQUIRE(d->tsdf && d->weight && d->colour, "ofx_volume_create: null volume pointer");
    OFX_REQUIRE(d->voxel_size > 0 && d->trunc_margin > 0, "ofx_volume_create: bad voxel_size/trunc_margin");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
        return fail("ofx_volume_create: no CUDA device (libofx has no CPU fallback)");

    ofx_volume *v = new (std::nothrow) ofx_volume();
    OFX_REQUIRE(v, "ofx_volume_create: out of host memory");
    std::memset(v, 0, sizeof(*v));
    v->desc = *d;
    OFX_CUDA(cudaGetDevice(&v->device));
    v->sm_count = sm_count();
    const int nx = d->dims[0] / d->n_shards, Dy = d->dims[1], Dz = d->dims[2];
    v->n_voxels = (int64_t)nx * Dy * Dz;

    // vox2world tables: float32(float64(origin) + voxel_size * float64(float32(index)))  tsdf.py:345-348
    // (each table starts on a 16-byte boundary: the fused kernel reads four z coordinates as one float4)
    const size_t off_y = ((size_t)nx + 3) & ~(size_t)3, off_z = (off_y + Dy + 3) & ~(size_t)3;
    std::vector<float> tab(off_z + Dz + 4, 0.0f);
    for (int lx = 0; lx < nx; ++lx) {
        const int gx = (lx / d->x_block) * d->x_block * d->n_shards + d->shard * d->x_block + lx % d->x_block;
        tab[lx] = (float)((double)d->origin[0] + d->voxel_size * (double)(float)gx);
    }
    for (int iy = 0; iy < Dy; ++iy) tab[off_y + iy] = (float)((double)d->origin[1] + d->voxel_size * (double)(float)iy);
    for (int iz = 0; iz < Dz; ++iz) tab[off_z + iz] = (float)((double)d->origin[2] + d->voxel_size * (double)(float)iz);
    OFX_CUDA(cudaMalloc(&v->tables, tab.size() * sizeof(float)));
    OFX_CUDA(cudaMemcpy(v->tables, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice));

    VolumeGeom &g = v->geom;
    g.nx = nx; g.Dy = Dy; g.Dz = Dz;
    g.nbx = (nx + BX - 1) / BX; g.nby = (Dy + BY - 1) / BY; g.nbz = (Dz + BZ - 1) / BZ;
    g.px = v->tables; g.py = v->tables + off_y; g.pz = v->tables + off_z;
    const int64_t nb = (int64_t)g.nbx * g.nby * g.nbz;
    OFX_REQUIRE(nb < (1ll << 31), "ofx_volume_create: too many bricks");
    v->n_bricks = (int)nb;
    OFX_CUDA(cudaMalloc(&v->brick_cnt, nb * sizeof(int)));
    OFX_CUDA(cudaMalloc(&v->brick_off, nb * sizeof(int)));
    OFX_CUDA(cudaMalloc(&v->brick_radius, nb * sizeof(float)));
    OFX_CUDA(cudaMalloc(&v->active, nb * sizeof(int)));
    for (int k = 0; k < 2; ++k) OFX_CUDA(cudaMalloc(&v->live_set[k], nb * LIVE_DESC_INTS * sizeof(int)));
    v->live = v->live_set[0];
    OFX_CUDA(cudaMalloc(&v->stats_acc, 2 * OFX_N_STATS * sizeof(unsigned long long)));
    OFX_CUDA(cudaMemset(v->stats_acc, 0, 2 * OFX_N_STATS * sizeof(unsigned long long)));
    OFX_CUDA(cudaMalloc(&v->active_desc, nb * ACTIVE_DESC_VEC * sizeof(int4)));
    OFX_CUDA(cudaMalloc(&v->brick_slot, nb * sizeof(int)));
    OFX_CUDA(cudaMalloc(&v->brick_box, 2 * nb * sizeof(float4)));
    OFX_CUDA(cudaMalloc(&v->scratch, 16 * sizeof(int)));      // [0..7] set_graph results, [8 + 4*set ..] n_live, brick work counter, retired CTAs
    OFX_CUDA(cudaMemset(v->scratch, 0, 16 * sizeof(int)));
    OFX_CUDA(cudaMalloc(&v->stats_scratch, OFX_N_STATS * sizeof(uint64_t)));
    OFX_CUDA(cudaMalloc(&v->scan_sums, ((nb + 1023) / 1024 + 1) * sizeof(int2)));
    OFX_CUDA(cudaHostAlloc(&v->host_res, 8 * sizeof(int), cudaHostAllocDefault));
    OFX_CUDA(cudaMalloc(&v->node_hist, 64 * sizeof(int)));
    OFX_CUDA(cudaHostAlloc(&v->host_hist, 64 * sizeof(int), cudaHostAllocDefault));
    OFX_CUDA(cudaEventCreateWithFlags(&v->graph_event, cudaEventDisableTiming));
    *out = v;
    return 0;
}

extern "C" int ofx_volume_destroy(ofx_volume *v)
{
    if (!v) return 0;
    cudaFree(v->tables); cudaFree(v->node_rec_set[0]); cudaFree(v->cache); cudaFree(v->brick_slot); cudaFree(v->brick_cnt);
    cudaFree(v->brick_off); cudaFree(v->brick_radius); cudaFree(v->cand); cudaFree(v->brick_nodes);
    cudaFree(v->active); cudaFree(v->scratch); cudaFree(v->stats_scratch);
    cudaFree(v->live_set[0]); cudaFree(v->live_set[1]); cudaFree(v->stats_acc); cudaFree(v->active_desc); cudaFree(v->brick_box); cudaFree(v->depth_tiles); cudaFree(v->frame_dc);
    cudaFree(v->scan_sums); cudaFree(v->node_hist);
    if (v->host_hist) cudaFreeHost(v->host_hist);
    if (v->host_res) cudaFreeHost(v->host_res);
    if (v->graph_event) cudaEventDestroy(v->graph_event);
    delete v;
    return 0;
}

extern "C" int64_t ofx_volume_num_voxels(const ofx_volume *v) { return v ? v->n_voxels : 0; }

extern "C" int ofx_volume_reset(ofx_volume *v, void *stream)
{
    OFX_REQUIRE(v, "ofx_volume_reset: null handle");
    const int64_t n = v->n_voxels;
    const bool aligned = (((uintptr_t)v->desc.tsdf | (uintptr_t)v->desc.weight | (uintptr_t)v->desc.colour) & 15) == 0;
    const int64_t n4 = aligned ? n / 4 : 0;
    fill_volume_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(
        (float4 *)v->desc.tsdf, (float4 *)v->desc.weight, (float4 *)v->desc.colour, n4, v->desc.tsdf, v->desc.weight,
        v->desc.colour, n);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_volume_world_points(ofx_volume *v, float *out_pts, void *stream)
{
    OFX_REQUIRE(v && out_pts, "ofx_volume_world_points: null argument");
    world_points_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(v->geom, v->n_voxels, out_pts);
    OFX_LAUNCH_CHECK();
    return 0;
}

// Device part of ofx_set_graph that lives here: (re)allocate and fill node positions.
int ofx_upload_nodes(ofx_volume *v, const float *nodes, int M, cudaStream_t s)
{
    if (M > v->node_rec_capacity) {
        if (v->node_rec_set[0]) OFX_CUDA(cudaFree(v->node_rec_set[0]));
        v->node_rec = v->node_rec_set[0] = v->node_rec_set[1] = nullptr;
        v->node_rec_capacity = 0;
        const int cap = M + M / 4 + 64;   // head-room for node insertion (EDGraph.update appends)
        OFX_CUDA(cudaMalloc(&v->node_rec_set[0], (size_t)2 * cap * NODE_REC_FLOATS * sizeof(float)));
        v->node_rec_set[1] = v->node_rec_set[0] + (size_t)cap * NODE_REC_FLOATS;
        v->node_rec_capacity = cap;
    }
    v->cur_set = 0;
    v->prepared = 0;
    v->node_rec = v->node_rec_set[0];
    v->live = v->live_set[0];
    for (int k = 0; k < 2; ++k) {
        write_node_positions_kernel<<<(M + 255) / 256, 256, 0, s>>>(nodes, M, v->node_rec_set[k]);
        OFX_LAUNCH_CHECK();
    }
    return 0;
}

// R|T of the M nodes into one of the two per-frame tables
int ofx_write_transforms(ofx_volume *v, float *table, const float *rotations, const float *translations, cudaStream_t s)
{
    write_node_transforms_kernel<<<(v->M * 12 + 255) / 256, 256, 0, s>>>(rotations, translations, v->M, table);
    OFX_LAUNCH_CHECK();
    return 0;
}

extern "C" int ofx_set_transforms(ofx_volume *v, const float *rotations, const float *translations, void *stream)
{
    OFX_REQUIRE(v && rotations && translations, "ofx_set_transforms: null argument");
    OFX_REQUIRE(v->graph_ready, "ofx_set_transforms: call ofx_set_graph first");
    if (int rc = graph_finish(v)) return rc;
    if (int rc = ofx_write_transforms(v, v->node_rec, rotations, translations, (cudaStream_t)stream)) return rc;
    v->transforms_ready = true;
    return 0;
}
