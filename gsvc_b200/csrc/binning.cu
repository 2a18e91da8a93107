// binning.cu -- tile binning: prefix scan of num_tiles_hit, intersect map, stable onesweep
// radix sort sized to the tile-key width, tile bin edges, sorted packed parameter copy.
// Reference: gsplat/gsplat/utils.py:99-167 (cumsum / .item() / torch.sort / gather glue),
// forward.cu:100-163 (map_gaussian_to_intersects, get_tile_bin_edges).
//
// Design (B200): everything is stream-ordered and sync-free.  The scan is a single-pass
// decoupled look-back fused with the intersect map and with the digit histograms of the
// sort; the sort is a least-significant-digit onesweep (one kernel per 8-bit digit, chained
// by decoupled look-back) over only ceil(log2(tiles)) key bits -- 2 passes at 1080p / 4K
// where the reference's torch.sort runs 8 passes over 64-bit keys plus a gather.
#include "gs2d_host.h"
#include "gs2d_math.cuh"

namespace gs2d {

// ------------------------------------------------------------------------------------------
// small device utilities
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

constexpr int kScanThreads = 256;
constexpr unsigned long long kFlagAgg = 1ull << 32, kFlagPrefix = 2ull << 32;

// inclusive scan of one value per thread over a 256-thread block; returns the inclusive
// value, *block_total receives the block sum.  sh: 8 ints of shared memory.
__device__ __forceinline__ int block_incl_scan(int v, int* sh, int* block_total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = warp_incl_scan(v, lane);
    if (lane == 31) sh[warp] = incl;
    __syncthreads();
    int wsum = (lane < kScanThreads / 32) ? sh[lane] : 0;
    const int wincl = warp_incl_scan(wsum, lane);
    const int wexcl = __shfl_sync(0xffffffffu, wincl - wsum, warp);
    *block_total = __shfl_sync(0xffffffffu, wincl, kScanThreads / 32 - 1);
    __syncthreads();
    return incl + wexcl;
}

// decoupled look-back (single pass prefix over blocks); called by all lanes of warp 0.
// state[b] = flag<<32 | value, zero-initialised.  Returns the exclusive prefix of block `bid`.
__device__ __forceinline__ int lookback_exclusive(unsigned long long* state, int bid, int block_total, int lane) {
    if (bid == 0) {
        if (lane == 0) st_relaxed(state, kFlagPrefix | (uint32_t)block_total);
        return 0;
    }
    if (lane == 0) st_relaxed(state + bid, kFlagAgg | (uint32_t)block_total);
    int excl = 0, look = bid - 1;
    while (true) {
        const int idx = look - lane;
        const unsigned long long s = idx >= 0 ? ld_relaxed(state + idx) : kFlagPrefix;
        const unsigned flag = (unsigned)(s >> 32);
        const unsigned inv = __ballot_sync(0xffffffffu, flag == 0);
        const unsigned pre = __ballot_sync(0xffffffffu, flag == 2);
        const unsigned need = pre ? ((2u << (__ffs(pre) - 1)) - 1u) : 0xffffffffu;   // lanes up to first prefix
        if (inv & need) continue;   // a needed predecessor has not published yet
        excl += warp_sum(((need >> lane) & 1u) ? (int)(uint32_t)s : 0);
        if (pre) break;
        look -= 32;
    }
    if (lane == 0) st_relaxed(state + bid, kFlagPrefix | (uint32_t)(excl + block_total));
    return excl;
}

// ------------------------------------------------------------------------------------------
// a4: inclusive cumsum of num_tiles_hit (reference utils.py:116)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kScanThreads)
k_cumsum(int n, const int32_t* __restrict__ nth, int32_t* __restrict__ cum, unsigned long long* state,
         int32_t* total_out, int32_t* total_host) {
    __shared__ int sh[8];
    __shared__ int sh_excl;
    const int g = blockIdx.x * kScanThreads + threadIdx.x;
    const int v = g < n ? nth[g] : 0;
    int total;
    const int incl = block_incl_scan(v, sh, &total);
    if (threadIdx.x < 32) {
        const int e = lookback_exclusive(state, blockIdx.x, total, threadIdx.x);
        if (threadIdx.x == 0) sh_excl = e;
    }
    __syncthreads();
    if (g < n) {
        const int out = sh_excl + incl;
        cum[g] = out;
        if (g == n - 1) {
            if (total_out) *total_out = out;
            if (total_host) *total_host = out;
        }
    }
}

// ------------------------------------------------------------------------------------------
// a5: intersect map with 64-bit keys (reference forward.cu:100-136), step-by-step API
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kScanThreads)
k_map_intersects64(int n, int num_intersects, const float* __restrict__ xys, const float* __restrict__ depths,
                   const int32_t* __restrict__ radii, const int32_t* __restrict__ cum, int tx, int ty,
                   int64_t* __restrict__ isect_ids, int32_t* __restrict__ gaussian_ids) {
    const int g = blockIdx.x * kScanThreads + threadIdx.x;
    if (g >= n) return;
    int cur = g == 0 ? 0 : cum[g - 1];
    const int end = min(cum[g], num_intersects);
    if (radii[g] > 0) {
        const float2 c = reinterpret_cast<const float2*>(xys)[g];
        int x0, y0, x1, y1;
        tile_bbox(c.x, c.y, (float)radii[g], tx, ty, x0, y0, x1, y1);
        const int64_t depth_id = (int64_t)__float_as_int(depths[g]);   // sign-extending, forward.cu:125
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
                if (cur >= end) break;
                const int64_t tile_id = (int64_t)(y * tx + x);
                isect_ids[cur] = (tile_id << 32) | depth_id;
                gaussian_ids[cur] = g;
                ++cur;
            }
    }
    // slots the reference leaves at their torch::zeros value (bindings.cu:293-296)
    for (; cur < end; ++cur) { isect_ids[cur] = 0; gaussian_ids[cur] = 0; }
}

// ------------------------------------------------------------------------------------------
// fused scan + map + sort histograms (hot path, 32-bit tile keys, depths == 0)
// ------------------------------------------------------------------------------------------
constexpr int kMaxPasses32 = 4;

__global__ void __launch_bounds__(kScanThreads)
k_scan_map(int n, const int32_t* __restrict__ nth, const float* __restrict__ xys, const int32_t* __restrict__ radii,
           const float* __restrict__ depths, int tx, int ty, int cap, int passes, unsigned long long* state,
           int32_t* status, int32_t* status_host, uint32_t* __restrict__ keys, int32_t* __restrict__ vals,
           uint32_t* ghist) {
    __shared__ int sh[8];
    __shared__ int sh_excl;
    __shared__ uint32_t sh_hist[kMaxPasses32 * 256];
    for (int i = threadIdx.x; i < passes * 256; i += kScanThreads) sh_hist[i] = 0;
    const int g = blockIdx.x * kScanThreads + threadIdx.x;
    const int cnt = g < n ? max(nth[g], 0) : 0;
    int total;
    const int incl = block_incl_scan(cnt, sh, &total);   // contains __syncthreads (covers sh_hist init)
    if (threadIdx.x < 32) {
        const int e = lookback_exclusive(state, blockIdx.x, total, threadIdx.x);
        if (threadIdx.x == 0) sh_excl = e;
    }
    __syncthreads();
    const int off = sh_excl + incl - cnt;
    if (g == n - 1) {
        const int tot = off + cnt;
        status[0] = tot;
        status[1] = tot > cap ? 1 : 0;
        status[3] = min(tot, cap);
    }
    unsigned dbits = 0;
    if (cnt > 0) {
        int k = 0;
        const int room = min(cnt, max(cap - off, 0));
        if (depths) dbits = (unsigned)__float_as_int(depths[g]);
        if (radii[g] > 0) {
            const float2 c = reinterpret_cast<const float2*>(xys)[g];
            int x0, y0, x1, y1;
            tile_bbox(c.x, c.y, (float)radii[g], tx, ty, x0, y0, x1, y1);
            for (int y = y0; y < y1 && k < room; ++y)
                for (int x = x0; x < x1 && k < room; ++x) {
                    const uint32_t key = (uint32_t)(y * tx + x);
                    keys[off + k] = key;
                    vals[off + k] = g;
                    for (int p = 0; p < passes; ++p) atomicAdd(&sh_hist[p * 256 + ((key >> (8 * p)) & 255u)], 1u);
                    ++k;
                }
        }
        for (; k < room; ++k) {   // reference leaves these slots zero (key 0, id 0)
            keys[off + k] = 0u;
            vals[off + k] = 0;
            for (int p = 0; p < passes; ++p) atomicAdd(&sh_hist[p * 256], 1u);
        }
    }
    if (depths) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dbits |= __shfl_xor_sync(0xffffffffu, dbits, o);
        if ((threadIdx.x & 31) == 0 && dbits) atomicOr((unsigned*)&status[2], dbits);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += kScanThreads) {
        const uint32_t v = sh_hist[i];
        if (v) atomicAdd(&ghist[i], v);
    }
    // the last block to finish mirrors the status words to pinned host memory
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int done = atomicAdd(&status[4], 1);
        if (done == (int)gridDim.x - 1 && status_host) {
            __threadfence();
            volatile int32_t* s = status;
            status_host[0] = s[0]; status_host[1] = s[1]; status_host[2] = s[2]; status_host[3] = s[3];
            __threadfence_system();
        }
    }
}

// ------------------------------------------------------------------------------------------
// a6: stable onesweep LSD radix sort of (key, int32 value) pairs
// ------------------------------------------------------------------------------------------
constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortIPT = 8;                             // items per thread
constexpr int kSortTile = kSortThreads * kSortIPT;      // 2048 items per block
constexpr uint32_t kLbAgg = 1u << 30, kLbPrefix = 2u << 30, kLbMask = (1u << 30) - 1u;

template <typename KeyT>
__device__ __forceinline__ uint32_t digit_of(KeyT key, int shift, uint32_t mask, uint32_t flip) {
    return ((uint32_t)(key >> shift) & mask) ^ flip;
}

// digit histograms of every pass at once (step-by-step API; the hot path builds them in k_scan_map)
template <typename KeyT>
__global__ void __launch_bounds__(kSortThreads)
k_sort_hist(const KeyT* __restrict__ keys, int n, int begin_bit, int end_bit, int passes, uint32_t top_flip,
            uint32_t* ghist) {
    __shared__ uint32_t sh_hist[8 * 256];
    for (int i = threadIdx.x; i < passes * 256; i += kSortThreads) sh_hist[i] = 0;
    __syncthreads();
    const int base = blockIdx.x * kSortTile;
    for (int j = 0; j < kSortIPT; ++j) {
        const int idx = base + j * kSortThreads + threadIdx.x;
        if (idx < n) {
            const KeyT key = keys[idx];
            for (int p = 0; p < passes; ++p) {
                const int shift = begin_bit + 8 * p;
                const int bits = min(8, end_bit - shift);
                const uint32_t flip = (p == passes - 1) ? top_flip : 0u;
                atomicAdd(&sh_hist[p * 256 + digit_of(key, shift, (1u << bits) - 1u, flip)], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += kSortThreads) {
        const uint32_t v = sh_hist[i];
        if (v) atomicAdd(&ghist[i], v);
    }
}

// One digit pass.  n comes from *n_dev when given (hot path: the count lives on the device).
// Items are warp-striped: warp w owns items [w*256, (w+1)*256) of the block's tile, item j of
// lane l is w*256 + j*32 + l, so (warp, j, lane) order == memory order and ranks are stable.
template <typename KeyT>
__global__ void __launch_bounds__(kSortThreads)
k_sort_pass(const KeyT* __restrict__ keys_in, const int32_t* __restrict__ vals_in, KeyT* __restrict__ keys_out,
            int32_t* __restrict__ vals_out, const int32_t* n_dev, int n_static, int shift, int bits, uint32_t flip,
            const uint32_t* __restrict__ ghist, uint32_t* lookback) {
    const int n = n_dev ? *n_dev : n_static;
    const int base = blockIdx.x * kSortTile;
    if (base >= n) return;
    __shared__ uint32_t warp_hist[kSortWarps][256];
    __shared__ uint32_t digit_off[256];
    __shared__ uint32_t sh_scan[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t mask = (1u << bits) - 1u;
    for (int i = tid; i < kSortWarps * 256; i += kSortThreads) (&warp_hist[0][0])[i] = 0;
    __syncthreads();

    KeyT key[kSortIPT];
    uint32_t rank[kSortIPT];
    const int wbase = base + warp * (32 * kSortIPT);
#pragma unroll
    for (int j = 0; j < kSortIPT; ++j) {
        const int idx = wbase + j * 32 + lane;
        key[j] = idx < n ? keys_in[idx] : (KeyT)0;
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < kSortIPT; ++j) {
        const bool valid = (wbase + j * 32 + lane) < n;
        const uint32_t d = digit_of(key[j], shift, mask, flip);
        const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
        uint32_t prev = 0;
        if (valid) prev = warp_hist[warp][d];
        __syncwarp();
        if (valid && lane == __ffs(peers) - 1) warp_hist[warp][d] = prev + __popc(peers);
        __syncwarp();
        rank[j] = prev + __popc(peers & lt_mask);
    }
    __syncthreads();

    // thread d owns digit d: exclusive scan over warps, block count, look-back over blocks
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
        const uint32_t t = warp_hist[w][tid];
        warp_hist[w][tid] = run;
        run += t;
    }
    uint32_t* col = lookback + tid;
    const int bid = blockIdx.x;
    uint32_t prefix = 0;
    if (bid == 0) {
        st_relaxed(col, kLbPrefix | run);
    } else {
        st_relaxed(col + (size_t)bid * 256, kLbAgg | run);
        int look = bid - 1;
        while (look >= 0) {
            const uint32_t v = ld_relaxed(col + (size_t)look * 256);
            const uint32_t flag = v >> 30;
            if (flag == 0) continue;   // predecessor has not published yet
            prefix += v & kLbMask;
            if (flag == 2) break;
            --look;
        }
        st_relaxed(col + (size_t)bid * 256, kLbPrefix | (prefix + run));
    }
    // global digit base = exclusive scan of the whole-array histogram of this pass
    const uint32_t gcount = ghist[tid];
    uint32_t incl = (uint32_t)warp_incl_scan((int)gcount, lane);
    if (lane == 31) sh_scan[warp] = incl;
    __syncthreads();
    uint32_t wsum = lane < kSortWarps ? sh_scan[lane] : 0;
    const uint32_t wincl = (uint32_t)warp_incl_scan((int)wsum, lane);
    const uint32_t wexcl = __shfl_sync(0xffffffffu, wincl - wsum, warp);
    digit_off[tid] = (incl - gcount) + wexcl + prefix;
    __syncthreads();

#pragma unroll
    for (int j = 0; j < kSortIPT; ++j) {
        const int idx = wbase + j * 32 + lane;
        if (idx < n) {
            const uint32_t d = digit_of(key[j], shift, mask, flip);
            const uint32_t pos = digit_off[d] + warp_hist[warp][d] + rank[j];
            keys_out[pos] = key[j];
            vals_out[pos] = vals_in[idx];
        }
    }
}

// ------------------------------------------------------------------------------------------
// a7: tile bin edges on 64-bit keys (reference forward.cu:141-163), step-by-step API
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_tile_bin_edges64(int n, const int64_t* __restrict__ keys_sorted, int32_t* __restrict__ tile_bins) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int32_t cur = (int32_t)(keys_sorted[i] >> 32);
    if (i == 0) tile_bins[2 * cur] = 0;
    if (i == n - 1) tile_bins[2 * cur + 1] = n;
    if (i == 0) return;
    const int32_t prev = (int32_t)(keys_sorted[i - 1] >> 32);
    if (prev != cur) {
        tile_bins[2 * prev + 1] = i;
        tile_bins[2 * cur] = i;
    }
}

// ------------------------------------------------------------------------------------------
// hot path: bin edges + sorted packed parameter copy in one pass over the sorted pairs.
// packed[i] = { x, y, conic.x, conic.y | conic.z, opacity, rect, gaussian id | r, g, b, tile }
// so each tile's batch is one contiguous 48-byte-per-entry run for the rasteriser.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_edges_pack(const int32_t* __restrict__ n_dev, int n_static, const uint32_t* __restrict__ keys_sorted,
             const int32_t* __restrict__ ids_sorted, const float* __restrict__ xys, const float* __restrict__ conics,
             const float* __restrict__ colors, const float* __restrict__ opac, int tx, int W, int H,
             int32_t* __restrict__ tile_bins, float4* __restrict__ packed) {
    const int n = n_dev ? *n_dev : n_static;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const uint32_t cur = keys_sorted[i];
    if (tile_bins) {
        if (i == 0) tile_bins[2 * cur] = 0;
        if (i == n - 1) tile_bins[2 * cur + 1] = n;
        if (i > 0) {
            const uint32_t prev = keys_sorted[i - 1];
            if (prev != cur) {
                tile_bins[2 * prev + 1] = i;
                tile_bins[2 * cur] = i;
            }
        }
    }
    const int g = ids_sorted[i];
    const float2 xy = reinterpret_cast<const float2*>(xys)[g];
    const float cx = conics[3 * g], cy = conics[3 * g + 1], cz = conics[3 * g + 2];
    const float o = opac[g];
    // colors == nullptr: the N-channel rasteriser reads its colours by Gaussian id instead
    const float r = colors ? colors[3 * g] : 0.f, gg = colors ? colors[3 * g + 1] : 0.f,
                b = colors ? colors[3 * g + 2] : 0.f;
    const int ox = (int)(cur % (uint32_t)tx) * GS2D_TILE, oy = (int)(cur / (uint32_t)tx) * GS2D_TILE;
    const uint32_t rect = pair_rect(xy.x, xy.y, o, cx, cy, cz, ox, oy, W, H);
    packed[3 * (size_t)i + 0] = make_float4(xy.x, xy.y, cx, cy);
    packed[3 * (size_t)i + 1] = make_float4(cz, o, __uint_as_float(rect), __int_as_float(g));
    packed[3 * (size_t)i + 2] = make_float4(r, gg, b, __uint_as_float(cur));
}

}  // namespace gs2d

// ==========================================================================================
// host side
// ==========================================================================================
#include <mutex>

#include "gs2d_internal.h"

namespace gs2d {

// step-by-step API: packed copy from (ids_sorted, tile_bins); one warp per tile
__global__ void __launch_bounds__(256)
k_pack_binned(int num_intersects, const int32_t* __restrict__ ids_sorted, const int32_t* __restrict__ tile_bins,
              int tiles, const float* __restrict__ xys, const float* __restrict__ conics,
              const float* __restrict__ colors, const float* __restrict__ opac, int tx, int W, int H,
              float4* __restrict__ packed) {
    const int t = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (t >= tiles) return;
    const int lo = tile_bins[2 * t], hi = min(tile_bins[2 * t + 1], num_intersects);
    const int ox = (t % tx) * GS2D_TILE, oy = (t / tx) * GS2D_TILE;
    for (int i = lo + (threadIdx.x & 31); i < hi; i += 32) {
        const int g = ids_sorted[i];
        const float2 xy = reinterpret_cast<const float2*>(xys)[g];
        const float cx = conics[3 * g], cy = conics[3 * g + 1], cz = conics[3 * g + 2];
        const float o = opac[g];
        const uint32_t rect = pair_rect(xy.x, xy.y, o, cx, cy, cz, ox, oy, W, H);
        packed[3 * (size_t)i + 0] = make_float4(xy.x, xy.y, cx, cy);
        packed[3 * (size_t)i + 1] = make_float4(cz, o, __uint_as_float(rect), __int_as_float(g));
        packed[3 * (size_t)i + 2] =
            colors ? make_float4(colors[3 * g], colors[3 * g + 1], colors[3 * g + 2], __uint_as_float((uint32_t)t))
                   : make_float4(0.f, 0.f, 0.f, __uint_as_float((uint32_t)t));
    }
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int key_bits_for(long long tiles) {
    int b = 1;
    while ((1ll << b) < tiles) ++b;
    return b;
}

int compute_state_offsets(int n, int cap, int tiles_x, int tiles_y, StateOffsets* so) {
    if (n < 0 || cap < 1 || tiles_x < 1 || tiles_y < 1) {
        set_error("state layout: bad sizes (n=%d cap=%d tiles=%dx%d)", n, cap, tiles_x, tiles_y);
        return GS2D_ERR_ARG;
    }
    const long long tiles = (long long)tiles_x * tiles_y;
    if (tiles > (1ll << 31) - 1) { set_error("state layout: too many tiles"); return GS2D_ERR_ARG; }
    gs2d_state_layout& L = so->pub;
    L.key_bits = key_bits_for(tiles);
    L.sort_passes = (L.key_bits + 7) / 8;
    so->scan_blocks = div_up(n > 0 ? n : 1, kScanThreads);
    so->sort_blocks = div_up(cap, kSortTile);
    size_t o = 0;
    L.off_status = o;            o = align_up(o + 8 * sizeof(int32_t), 256);
    L.off_tile_bins = o;         o = align_up(o + (size_t)tiles * 2 * sizeof(int32_t), 256);
    so->off_scan_state = o;      o = align_up(o + (size_t)so->scan_blocks * sizeof(unsigned long long), 256);
    so->off_ghist = o;           o = align_up(o + (size_t)kMaxPasses32 * 256 * sizeof(uint32_t), 256);
    so->off_lookback = o;        o = align_up(o + (size_t)L.sort_passes * so->sort_blocks * 256 * sizeof(uint32_t), 256);
    L.zero_begin = 0;
    L.zero_bytes = o;
    L.off_ids_sorted = o;        o = align_up(o + (size_t)cap * sizeof(int32_t), 256);
    L.off_keys_sorted = o;       o = align_up(o + (size_t)cap * sizeof(uint32_t), 256);
    L.off_packed = o;            o = align_up(o + (size_t)cap * 3 * sizeof(float4), 256);
    L.off_scratch = so->off_scan_state;
    so->off_keys_tmp = o;        o = align_up(o + (size_t)cap * sizeof(uint32_t), 256);
    so->off_ids_tmp = o;         o = align_up(o + (size_t)cap * sizeof(int32_t), 256);
    L.total_bytes = o;
    return GS2D_OK;
}

int launch_bin_and_pack(int n, const float* xys, const float* depths, const int32_t* radii, const int32_t* nth,
                        const float* conics, const float* colors, const float* opac, int img_h, int img_w, int cap,
                        const StateOffsets& so, char* state, int32_t* status_host, cudaEvent_t scan_done,
                        cudaStream_t s) {
    const gs2d_state_layout& L = so.pub;
    const int tx = div_up(img_w, GS2D_TILE), ty = div_up(img_h, GS2D_TILE);
    int32_t* status = (int32_t*)(state + L.off_status);
    int32_t* tile_bins = (int32_t*)(state + L.off_tile_bins);
    unsigned long long* scan_state = (unsigned long long*)(state + so.off_scan_state);
    uint32_t* ghist = (uint32_t*)(state + so.off_ghist);
    uint32_t* lookback = (uint32_t*)(state + so.off_lookback);
    uint32_t* keys_fin = (uint32_t*)(state + L.off_keys_sorted);
    int32_t* ids_fin = (int32_t*)(state + L.off_ids_sorted);
    uint32_t* keys_tmp = (uint32_t*)(state + so.off_keys_tmp);
    int32_t* ids_tmp = (int32_t*)(state + so.off_ids_tmp);
    float4* packed = (float4*)(state + L.off_packed);
    const int passes = L.sort_passes;

    {
        ProfScope ps_(GS2D_STAGE_MEMSET, s);
        GS2D_CHECK_CUDA(cudaMemsetAsync(state + L.zero_begin, 0, L.zero_bytes, s));
    }
    // the map writes into the buffer from which `passes` ping-pong steps end in the final one
    uint32_t* kin = (passes % 2 == 0) ? keys_fin : keys_tmp;
    int32_t* vin = (passes % 2 == 0) ? ids_fin : ids_tmp;
    uint32_t* kout = (passes % 2 == 0) ? keys_tmp : keys_fin;
    int32_t* vout = (passes % 2 == 0) ? ids_tmp : ids_fin;
    if (n > 0) {
        {
            ProfScope ps_(GS2D_STAGE_SCAN_MAP, s);
            k_scan_map<<<so.scan_blocks, kScanThreads, 0, s>>>(n, nth, xys, radii, depths, tx, ty, cap, passes,
                                                             scan_state, status, status_host, kin, vin, ghist);
        }
        GS2D_CHECK_LAUNCH("scan_map");
    } else if (status_host) {
        GS2D_CHECK_CUDA(cudaMemcpyAsync(status_host, status, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    }
    if (scan_done) GS2D_CHECK_CUDA(cudaEventRecord(scan_done, s));
    if (n == 0) return GS2D_OK;
    for (int p = 0; p < passes; ++p) {
        const int shift = 8 * p, bits = min(8, L.key_bits - shift);
        {
            ProfScope ps_(GS2D_STAGE_SORT, s);
            k_sort_pass<uint32_t><<<so.sort_blocks, kSortThreads, 0, s>>>(
            kin, vin, kout, vout, status + 3, 0, shift, bits, 0u, ghist + p * 256,
            lookback + (size_t)p * so.sort_blocks * 256);
        }
        GS2D_CHECK_LAUNCH("sort_pass");
        uint32_t* tk = kin; kin = kout; kout = tk;
        int32_t* tv = vin; vin = vout; vout = tv;
    }
    // kin/vin now alias keys_fin/ids_fin
    {
        ProfScope ps_(GS2D_STAGE_EDGES_PACK, s);
        k_edges_pack<<<div_up(cap, 256), 256, 0, s>>>(status + 3, 0, keys_fin, ids_fin, xys, conics, colors, opac, tx,
                                                  img_w, img_h, tile_bins, packed);
    }
    GS2D_CHECK_LAUNCH("edges_pack");
    return GS2D_OK;
}

int launch_pack_binned(int num_intersects, const int32_t* ids_sorted, const int32_t* tile_bins, int tiles,
                       const float* xys, const float* conics, const float* colors, const float* opac, int img_h,
                       int img_w, float4* packed, cudaStream_t s) {
    const int tx = div_up(img_w, GS2D_TILE);
    (void)img_h;
    {
        ProfScope ps_(GS2D_STAGE_EDGES_PACK, s);
        k_pack_binned<<<div_up(tiles, 8), 256, 0, s>>>(num_intersects, ids_sorted, tile_bins, tiles, xys, conics, colors,
                                                   opac, tx, img_w, img_h, packed);
    }
    GS2D_CHECK_LAUNCH("pack_binned");
    return GS2D_OK;
}

}  // namespace gs2d

using namespace gs2d;

extern "C" {

int gs2d_state_layout_for(int n, int cap, int tiles_x, int tiles_y, gs2d_state_layout* out) {
    GS2D_REQUIRE(out != nullptr, "state_layout_for: null out");
    StateOffsets so;
    const int rc = compute_state_offsets(n, cap, tiles_x, tiles_y, &so);
    if (rc != GS2D_OK) return rc;
    *out = so.pub;
    return GS2D_OK;
}

size_t gs2d_scan_scratch_bytes(int n) {
    return align_up((size_t)div_up(n > 0 ? n : 1, kScanThreads) * sizeof(unsigned long long), 256);
}

int gs2d_cumsum_tiles_hit(int n, const int32_t* nth, int32_t* cum, int32_t* total_out, int32_t* total_host,
                          void* scratch, size_t scratch_bytes, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0, "cumsum: bad size");
    GS2D_REQUIRE(scratch_bytes >= gs2d_scan_scratch_bytes(n), "cumsum: scratch too small");
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {
        if (total_out) GS2D_CHECK_CUDA(cudaMemsetAsync(total_out, 0, sizeof(int32_t), s));
        return GS2D_OK;
    }
    GS2D_CHECK_CUDA(cudaMemsetAsync(scratch, 0, gs2d_scan_scratch_bytes(n), s));
    {
        ProfScope ps_(GS2D_STAGE_SCAN_MAP, s);
        k_cumsum<<<div_up(n, kScanThreads), kScanThreads, 0, s>>>(n, nth, cum, (unsigned long long*)scratch, total_out,
                                                               total_host);
    }
    GS2D_CHECK_LAUNCH("cumsum");
    return GS2D_OK;
}

// ---- the same scan with its scratch in a file-scope device array: allocates nothing -----------------------
// (no counterpart in the reference).  The look-back state of
// k_cumsum -- one 8-byte word per block of 256 values -- lives in g_scan_scratch instead of caller memory:
// GS2D_CUMSUM_STATIC_MAX_N values at most = 16 384 words = 128 KB, held by every GPU that loads the library
// until exit.  In host code `g_scan_scratch` is only a stand-in for the symbol, NOT a device address: the
// address is fetched once per device, with that device current, by cudaGetSymbolAddress.
constexpr int kStaticScanMaxN = GS2D_CUMSUM_STATIC_MAX_N;
constexpr int kStaticScanWords = (kStaticScanMaxN + kScanThreads - 1) / kScanThreads;
static __device__ __align__(256) unsigned long long g_scan_scratch[kStaticScanWords];

constexpr int kMaxDevices = 64;
static unsigned long long* g_scan_scratch_ptr[kMaxDevices] = {};   // device address of the array, per device
static std::mutex g_scan_scratch_mu;

static int static_scan_scratch(unsigned long long** out) {
    int dev = -1;
    GS2D_CHECK_CUDA(cudaGetDevice(&dev));
    GS2D_REQUIRE(dev >= 0 && dev < kMaxDevices, "cumsum_static: device index out of range");
    std::lock_guard<std::mutex> lk(g_scan_scratch_mu);
    if (g_scan_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        GS2D_CHECK_CUDA(cudaGetSymbolAddress(&p, g_scan_scratch));   // this device's copy: `dev` is current
        GS2D_REQUIRE(p != nullptr, "cumsum_static: no address for the scratch symbol");
        g_scan_scratch_ptr[dev] = (unsigned long long*)p;
    }
    *out = g_scan_scratch_ptr[dev];
    return GS2D_OK;
}

int gs2d_cumsum_tiles_hit_static(int n, const int32_t* nth, int32_t* cum, int32_t* total_out, int32_t* total_host,
                                 gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0, "cumsum_static: bad size");
    if (n > kStaticScanMaxN) {
        set_error("cumsum_static: n = %d exceeds the built-in scratch (GS2D_CUMSUM_STATIC_MAX_N = %d values); "
                  "use gs2d_cumsum_tiles_hit with caller-owned scratch", n, kStaticScanMaxN);
        return GS2D_ERR_ARG;
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {
        if (total_out) GS2D_CHECK_CUDA(cudaMemsetAsync(total_out, 0, sizeof(int32_t), s));
        return GS2D_OK;
    }
    GS2D_REQUIRE(nth != nullptr && cum != nullptr, "cumsum_static: null pointer");
    unsigned long long* state = nullptr;
    const int rc = static_scan_scratch(&state);
    if (rc != GS2D_OK) return rc;
    const int blocks = div_up(n, kScanThreads);
    // the array keeps the previous call's flags: the words this call uses must start at zero
    GS2D_CHECK_CUDA(cudaMemsetAsync(state, 0, (size_t)blocks * sizeof(unsigned long long), s));
    {
        ProfScope ps_(GS2D_STAGE_SCAN_MAP, s);
        k_cumsum<<<blocks, kScanThreads, 0, s>>>(n, nth, cum, state, total_out, total_host);
    }
    GS2D_CHECK_LAUNCH("cumsum_static");
    return GS2D_OK;
}

int gs2d_map_intersects(int n, int num_intersects, const float* xys, const float* depths, const int32_t* radii,
                        const int32_t* cum, int tx, int ty, int64_t* isect_ids, int32_t* gaussian_ids,
                        gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && num_intersects >= 0 && tx > 0 && ty > 0, "map_intersects: bad sizes");
    if (n == 0 || num_intersects == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_SCAN_MAP, (cudaStream_t)stream);
        k_map_intersects64<<<div_up(n, kScanThreads), kScanThreads, 0, (cudaStream_t)stream>>>(
        n, num_intersects, xys, depths, radii, cum, tx, ty, isect_ids, gaussian_ids);
    }
    GS2D_CHECK_LAUNCH("map_intersects");
    return GS2D_OK;
}

// scratch = ghist[8][256] | lookback[8][blocks][256] | keys_tmp[n] (u64) | vals_tmp[n]
size_t gs2d_sort_scratch_bytes(int n) {
    const size_t blocks = (size_t)div_up(n > 0 ? n : 1, kSortTile);
    size_t o = align_up(8 * 256 * sizeof(uint32_t), 256);
    o = align_up(o + 8 * blocks * 256 * sizeof(uint32_t), 256);
    o = align_up(o + (size_t)(n > 0 ? n : 1) * sizeof(int64_t), 256);
    o = align_up(o + (size_t)(n > 0 ? n : 1) * sizeof(int32_t), 256);
    return o;
}

int gs2d_sort_pairs(int n, const int64_t* keys_in, const int32_t* vals_in, int64_t* keys_out, int32_t* vals_out,
                    int begin_bit, int end_bit, void* scratch, size_t scratch_bytes, gs2d_stream_t stream) {
    GS2D_REQUIRE(n >= 0 && begin_bit >= 0 && end_bit <= 64 && begin_bit < end_bit, "sort_pairs: bad arguments");
    GS2D_REQUIRE(scratch_bytes >= gs2d_sort_scratch_bytes(n), "sort_pairs: scratch too small");
    if (n == 0) return GS2D_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int passes = (end_bit - begin_bit + 7) / 8;
    const int blocks = div_up(n, kSortTile);
    char* base = (char*)scratch;
    size_t o = 0;
    uint32_t* ghist = (uint32_t*)(base + o);     o = align_up(o + 8 * 256 * sizeof(uint32_t), 256);
    uint32_t* lookback = (uint32_t*)(base + o);  o = align_up(o + (size_t)8 * blocks * 256 * sizeof(uint32_t), 256);
    const size_t zero_bytes = o;
    uint64_t* keys_tmp = (uint64_t*)(base + o);  o = align_up(o + (size_t)n * sizeof(int64_t), 256);
    int32_t* vals_tmp = (int32_t*)(base + o);
    GS2D_CHECK_CUDA(cudaMemsetAsync(base, 0, zero_bytes, s));
    const int top_bits = (end_bit - begin_bit) - 8 * (passes - 1);
    const uint32_t top_flip = end_bit == 64 ? (1u << (top_bits - 1)) : 0u;   // signed int64 order
    {
        ProfScope ps_(GS2D_STAGE_SORT, s);
        k_sort_hist<uint64_t><<<blocks, kSortThreads, 0, s>>>((const uint64_t*)keys_in, n, begin_bit, end_bit, passes,
                                                          top_flip, ghist);
    }
    GS2D_CHECK_LAUNCH("sort_hist");
    // ping-pong so that the last pass lands in keys_out / vals_out
    const uint64_t* kin = (const uint64_t*)keys_in;
    const int32_t* vin = vals_in;
    for (int p = 0; p < passes; ++p) {
        const bool to_out = ((passes - 1 - p) % 2) == 0;
        uint64_t* kout = to_out ? (uint64_t*)keys_out : keys_tmp;
        int32_t* vout = to_out ? vals_out : vals_tmp;
        const int shift = begin_bit + 8 * p, bits = min(8, end_bit - shift);
        {
            ProfScope ps_(GS2D_STAGE_SORT, s);
            k_sort_pass<uint64_t><<<blocks, kSortThreads, 0, s>>>(kin, vin, kout, vout, nullptr, n, shift, bits,
                                                              p == passes - 1 ? top_flip : 0u, ghist + p * 256,
                                                              lookback + (size_t)p * blocks * 256);
        }
        GS2D_CHECK_LAUNCH("sort_pass64");
        kin = kout;
        vin = vout;
    }
    return GS2D_OK;
}

int gs2d_tile_bin_edges(int num_intersects, const int64_t* keys_sorted, int32_t* tile_bins, int rows,
                        gs2d_stream_t stream) {
    GS2D_REQUIRE(num_intersects >= 0 && rows >= 0, "tile_bin_edges: bad sizes");
    cudaStream_t s = (cudaStream_t)stream;
    if (rows > 0) GS2D_CHECK_CUDA(cudaMemsetAsync(tile_bins, 0, (size_t)rows * 2 * sizeof(int32_t), s));
    if (num_intersects == 0) return GS2D_OK;
    {
        ProfScope ps_(GS2D_STAGE_EDGES_PACK, s);
        k_tile_bin_edges64<<<div_up(num_intersects, 256), 256, 0, s>>>(num_intersects, keys_sorted, tile_bins);
    }
    GS2D_CHECK_LAUNCH("tile_bin_edges");
    return GS2D_OK;
}

}  // extern "C"
