// zsb_internal.h — host-side context layout shared by zsb_host.cu and zsb_zstream.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>
#include "../../include/zsb.h"

namespace zsb { struct Aux; }

struct zsb_dev {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr, back_stream = nullptr;   // decode, host->device, device->host
    cudaStream_t group_stream = nullptr;      // large members (group path) beside the lane path
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    std::vector<cudaStream_t> pipe_streams;   // decode streams of the chunks of a pipelined host batch
    std::vector<cudaEvent_t> pipe_ev;  // per chunk of a host batch: inputs arrived, decoded
    uint16_t *fixed = nullptr;         // prebuilt fixed-code tables (one member slot, group path)
    uint8_t *fixed_lane = nullptr;     // the same in the lane path's slot format
    uint8_t *d_seq = nullptr;          // LZ sequence arena of the lane path
    size_t d_seq_cap = 0;
    size_t arena_auto = 0;             // arena chosen from the device's free memory (option seq_arena_bytes = 0)
    uint8_t *d_plan = nullptr;         // per call: slot order (u32) + per-slot sequence offsets (u64)
    size_t d_plan_cap = 0;
    bool lane_configured = false;
    int lz_occ = 0, lane_small_occ = 0, big_occ = 0;
    uint32_t *counters = nullptr;      // [0] decode work counter, [1] check work counter
    zsb::Aux *aux = nullptr;
    size_t aux_cap = 0;
    // staging arenas for host-pointer batches
    uint8_t *d_in = nullptr, *d_out = nullptr, *d_desc = nullptr;
    size_t d_in_cap = 0, d_out_cap = 0, d_desc_cap = 0;
    uint8_t *d_df = nullptr;           // zsb_deflate_batch: descriptors, token scratch, block scratch, staging
    size_t d_df_cap = 0;
    uint8_t *d_segbuf = nullptr;       // segmented members: plans, segment records, per-slot sequence offsets
    size_t d_segbuf_cap = 0;
    bool seg_configured = false;
    cudaStream_t seg_stream = nullptr;  // the segmented members run beside the lane path's part
    cudaEvent_t seg_fork = nullptr, seg_join = nullptr;
    cudaStream_t seg_stream2 = nullptr; // the global-memory resolver beside the ring resolver
    cudaEvent_t seg_fork2 = nullptr, seg_join2 = nullptr;
    int find_occ = 0;
    uint8_t *d_chkparts = nullptr;     // large wrapped members: per-member first part + the parts' partial check values
    size_t d_chkparts_cap = 0;
    uint8_t *h_desc = nullptr;
    size_t h_desc_cap = 0;
    bool configured[33] = {};
    int occ[33] = {};
    // "profile" option: CUDA events around every decode-kernel launch, on the launching stream
    std::vector<cudaEvent_t> prof_ev;  // 3 per lane-path slice: before k_lane, between, after k_lz
    size_t prof_used = 0;
    std::vector<cudaEvent_t> chk_ev;   // 2 per k_check launch
    size_t chk_used = 0;
    std::vector<cudaEvent_t> seg_ev;   // 2 per segmented part
    size_t seg_used = 0;
};

struct zsb_ctx {
    std::vector<zsb_dev> devs;
    int lanes = 16;                    // group path: lanes of a warp that cooperate on one member
    int path = 0;                      // 0 auto, 1 lane path (k_lane + k_lz), 2 group path (k_inflate<G>)
    size_t seq_arena_bytes = 0;        // most sequence scratch one slice of a batch may use; 0 = from free memory (arena_bytes)
    size_t slab_bytes = 256u << 20;
    int chunks = 0;                    // host batches: pipeline depth (0 = choose from the batch)
    long launches = 0;
    bool force_small_ctas = false;
    int split_big = 1;                 // path 0: 1 = members far larger than the rest get a warp each (k_big,
                                       // split_big()), 0 = never, 2 = every member (tests)
    int segment = 1;                   // members of seg_min compressed bytes or more are decoded in segments (zsb_seg.cuh)
    size_t seg_bytes = 0, seg_min = 131072;       // seg_bytes 0 = chosen per batch (seg_bytes_for)
    int profile = 0;                   // 1: record per-kernel events (zsb_profile_read)
    char err[256] = "";
};
