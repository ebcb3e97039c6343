// ofx_volume.cuh -- the device-side state behind an ofx_volume handle, shared by the
// translation units that build the anchor cache (ofx_skin_cache.cu) and run the fused
// warp + integrate kernel (ofx_integrate.cu).
#pragma once
#include "../../include/ofx.h"
#include "ofx_common.h"
#include "ofx_math.cuh"

namespace ofx {

// A brick is the unit of work of one CTA: BX x BY x BZ voxels, 256 threads, each thread owns
// VOX_PER_THREAD consecutive voxels along z (the contiguous axis) = one float4 of each volume.
constexpr int BX = 4, BY = 8, BZ = 32;
constexpr int VOX_PER_THREAD = 4;
constexpr int BRICK_THREADS = BX * BY * (BZ / VOX_PER_THREAD);   // 256
constexpr int BRICK_VOXELS = BX * BY * BZ;                       // 1024
constexpr int NODE_REC_FLOATS = 16;   // {gx,gy,gz,0, R00..R22, T0,T1,T2} = 64 B in global memory
constexpr int ACTIVE_DESC_VEC = 4;     // active-brick descriptor in int4 units: the live descriptor + the brick's bounding ball + box half extents
constexpr int LIVE_DESC_INTS = 8;      // live-brick descriptor: {lx0, iy0, iz0, n_used | list offset, cache slot, brick, 0}
constexpr int NODE_XFORM_OFFSET = 4;  // the 48-byte {R, T} tail is what the fused kernels stage (one TMA bulk copy)
constexpr int NODE_XFORM_FLOATS = 12;
__host__ __device__ constexpr int slot_bytes_of(int kp) { return BRICK_VOXELS * 5 * kp; }
#ifndef OFX_ROW_TAB_CAP
#define OFX_ROW_TAB_CAP 32
#endif
constexpr int ROW_TAB_CAP = OFX_ROW_TAB_CAP;   // largest per-warp row table of the fused kernels (ofx_brick.cuh)
constexpr int MAX_BRICK_NODES = 255;  // local anchor ids are bytes, 0xFF = "no anchor"

struct VolumeGeom {
    int nx, Dy, Dz;          // local extent (nx = planes owned by this shard)
    int nbx, nby, nbz;       // bricks per axis
    const float *px;         // world x of local plane lx   (nx entries)   tsdf.py:348
    const float *py;         // world y of iy               (Dy entries)
    const float *pz;         // world z of iz               (Dz entries)
};

struct BrickCoord {
    int lx0, iy0, iz0;
};

__device__ __forceinline__ BrickCoord brick_coord(const VolumeGeom &g, int b)
{
    const int bz = b % g.nbz;
    const int by = (b / g.nbz) % g.nby;
    const int bx = b / (g.nbz * g.nby);
    return BrickCoord{bx * BX, by * BY, bz * BZ};
}

}  // namespace ofx

struct ofx_volume {
    ofx_volume_desc desc;
    int device;
    ofx::VolumeGeom geom;
    int64_t n_voxels;
    int n_bricks;
    float *tables;            // px | py | pz

    // graph
    int M, K;
    double coverage;
    float cutoff, denom;
    float *node_rec;          // M x NODE_REC_FLOATS: the CURRENT per-frame set (= node_rec_set[cur_set])
    float *node_rec_set[2];   // two tables (same node positions): ofx_prepare_frame fills the other one while a frame integrates
    int node_rec_capacity;

    // skin cache (device form of WarpField.skin_tsdf's cache, warpfield.py:131-141), brick-major and
    // compact: one slot of slot_bytes = 1024 * 5 * Kp bytes (Kp = 4 for K <= 4, else 8) per brick that
    // has candidate nodes, laid out for ONE bulk copy into shared memory and conflict-free reads
    // (thread t, voxel e of the thread, anchor group g = k / 4):
    //   ids      u32   at ((g*4 + e)*256 + t) * 4              four brick-local node ids, 0xFF = not skin-valid
    //   weights  f32x4 at 1024*Kp + ((g*4 + e)*256 + t) * 16   normalised weights (warpfield.py:121), 0 beyond K
    int cache_kp;                // Kp
    uint8_t *cache;
    int64_t cache_capacity;      // bytes allocated
    int *brick_slot;             // n_bricks: slot of a brick (meaningful where brick_cnt > 0)
    int n_slots;
    int *brick_cnt;              // n_bricks: candidate count, later used-node count
    int *brick_off;              // n_bricks: offset into brick_nodes / cand
    float *brick_radius;         // n_bricks
    int *cand;                   // conservative candidate lists
    int *brick_nodes;            // used-node lists (ascending node index)
    int64_t list_capacity;
    int *active;                 // bricks with at least one skin-valid voxel
    int n_active;
    float4 *brick_box;           // 2 x n_bricks: {brick centre (world), bounding radius h}, {box half extents, 0}; written by ofx_set_graph
    int4 *active_desc;           // n_active x ACTIVE_DESC_VEC int4: {lx0, iy0, iz0, n_used | list offset, cache slot, brick, 0 | centre + radius | half extents}
    int *live;                   // per frame: descriptors (LIVE_DESC_INTS ints) of the active bricks that survive the cull (= live_set[cur_set])
    int *live_set[2];
    int cur_set;                 // per-frame set (node transforms, live list, launch counters, counter accumulators) in use
    float prepared_intr[4];      // the camera ofx_prepare_frame culled against
    int prepared_hw[2];
    int prepared;                // the other set holds the next frame: 1 = transforms + frustum-culled live list (ofx_prepare_frame), 2 = everything but the fused kernel (ofx_stage_frame)
    const float *staged_depth, *staged_colour;   // the images ofx_stage_frame was given
    unsigned long long *stats_acc;   // 2 x OFX_N_STATS: per-set accumulators of the persistent kernel
    float2 *frame_dc;            // per frame set: the frame interleaved {depth, packed colour} (frame_prep_kernel); set k at k * frame_capacity
    float *depth_tiles;          // per frame set: max valid depth of each DEPTH_TILE x DEPTH_TILE pixel tile (0 = none); set k at k * tile_capacity
    int64_t frame_capacity, tile_capacity;   // pixels / tiles allocated per set
    int sm_count;                // of the handle's device
    int *scratch;                // small device scratch: [0]=total [1]=max [2]=error [8 + 4*set ..]: n_live, work counter, retired CTAs (all zero between launches)
    uint64_t *stats_scratch;     // OFX_N_STATS counters when the caller passes NULL
    bool graph_ready, transforms_ready;
    // ofx_set_graph does not wait for its last kernels: their results land in host_res behind graph_event and are read
    // by graph_finish(), which every call that needs them runs first
    int2 *scan_sums;              // ceil(n_bricks / 1024) block sums of the brick scans
    int *host_res;                // page-locked copy of scratch[0..8)
    cudaEvent_t graph_event;
    bool graph_pending;
    int *node_hist;               // device: histogram of the used-node counts of the active bricks (bins 0..ROW_TAB_CAP, last = more)
    int *host_hist;               // page-locked copy
    int tab_cap;                  // row-table capacity of the persistent kernel for this graph (graph_finish)
    int max_used;                 // largest brick node count of the graph (graph_finish)
    bool counters_dirty[2];       // a launch failed between the cull and the fused kernel: the launch counters need a reset
};

namespace ofx {
int graph_finish(ofx_volume *v);   // ofx_skin_cache.cu
}
