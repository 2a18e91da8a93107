// gs2d_internal.h -- host-side interfaces between the translation units of libgs2d.so.
#pragma once
#include "gs2d_host.h"

namespace gs2d {

// Byte offsets inside the caller-owned state buffer of one forward (public part mirrors
// gs2d_state_layout; the private part is scratch that is dead once the forward is done).
struct StateOffsets {
    gs2d_state_layout pub;
    size_t off_scan_state;   // uint64[scan_blocks]
    size_t off_ghist;        // uint32[4][256]
    size_t off_lookback;     // uint32[passes][sort_blocks][256]
    size_t off_keys_tmp;     // uint32[cap]  (other half of the ping-pong)
    size_t off_ids_tmp;      // int32[cap]
    int scan_blocks;
    int sort_blocks;
};
int compute_state_offsets(int n, int cap, int tiles_x, int tiles_y, StateOffsets* out);

// memset + scan/map/histogram + onesweep passes + edges/pack, all on `s`, no host sync.
int launch_bin_and_pack(int n, const float* xys, const float* depths, const int32_t* radii,
                        const int32_t* num_tiles_hit, const float* conics, const float* colors, const float* opac,
                        int img_h, int img_w, int cap, const StateOffsets& so, char* state, int32_t* status_host,
                        cudaEvent_t scan_done, cudaStream_t s);

// pack a caller-sorted id list (step-by-step API): tile ids are recovered from tile_bins.
int launch_pack_binned(int num_intersects, const int32_t* ids_sorted, const int32_t* tile_bins, int tiles,
                       const float* xys, const float* conics, const float* colors, const float* opac, int img_h,
                       int img_w, float4* packed, cudaStream_t s);

}  // namespace gs2d

extern "C" {
// fit.cu: layout of the operator path's per-call state (the fit workspace without gradient rows)
size_t gs2d_opstate_bytes(int cap, int img_h, int img_w);
int gs2d_opstate_ranges(int cap, int img_h, int img_w, size_t* off_tile_off, size_t* off_packed, size_t* off_status);
}
