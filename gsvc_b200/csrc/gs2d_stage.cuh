// gs2d_stage.cuh -- staging of a tile's batch of packed records into shared memory in ascending
// Gaussian-id order, shared by the fused fit / render tile kernel (fit.cu) and the operator-path
// backward (raster.cu).  The count / scatter binning writes a tile's records in arbitrary order;
// ascending id is the order the reference's stable sort on the tile id produces (SURVEY Q5:
// gsplat/gsplat/utils.py:164-165 of the reference), so ranking the ids restores it exactly.
#pragma once
#include "gs2d_math.cuh"

namespace gs2d {

constexpr int kIdChunk = 1024;

// rank += (a < b), as one predicate set and one predicated add
__device__ __forceinline__ void count_if_less(int& rank, int a, int b) {
    asm("{\n.reg .pred p;\nsetp.lt.s32 p, %1, %2;\n@p add.s32 %0, %0, 1;\n}" : "+r"(rank) : "r"(a), "r"(b));
}

// Stage the tile's batch into shared memory in ascending Gaussian-id order (the scatter wrote
// the range in arbitrary order; ascending id is the order of the reference's stable sort on
// tile id).  cnt <= 256: ids go to shared memory, every thread ranks its records'
// ids against them and stores each record at its rank (the first record per thread is already
// in flight while the ids land).  cnt > 256 (rare; the reference then renders only the 256
// lowest ids): ranks are computed against all cnt ids in chunks and the 256
// lowest-ranked records are kept.
// publish (every Gaussian of a fuller tile is to be rendered, in batches of 256): the rank of each
// record is also stored into the record's spare word (q2.w, the tile id, which the fit kernels never
// read), so that the later batches are staged by stage_tile_window without ranking again.
template <int THREADS>
__device__ __forceinline__ int stage_tile_sorted(const float4* __restrict__ src, int cnt, float4* sq, int* sid,
                                                 int tid, bool publish = false) {
    const int nb = min(cnt, GS2D_MAX_PER_TILE);
    if (cnt <= GS2D_MAX_PER_TILE) {
        float4 r0, r1, r2;
        if (tid < cnt) {
            r0 = __ldg(&src[3 * (size_t)tid + 0]);
            r1 = __ldg(&src[3 * (size_t)tid + 1]);
            r2 = __ldg(&src[3 * (size_t)tid + 2]);
            sid[tid] = __float_as_int(r1.w);
        }
        for (int j = tid + THREADS; j < cnt; j += THREADS)
            sid[j] = __float_as_int(__ldg(&src[3 * (size_t)j + 1]).w);
        __syncthreads();
        for (int j = tid; j < cnt; j += THREADS) {
            const int my = sid[j];
            int rank = 0;
            for (int i = 0; i < cnt; ++i) count_if_less(rank, sid[i], my);
            if (j != tid) {
                r0 = __ldg(&src[3 * (size_t)j + 0]);
                r1 = __ldg(&src[3 * (size_t)j + 1]);
                r2 = __ldg(&src[3 * (size_t)j + 2]);
            }
            GS2D_ASSERT(rank >= 0 && rank < cnt);
            sq[3 * rank + 0] = r0;
            sq[3 * rank + 1] = r1;
            sq[3 * rank + 2] = r2;
        }
        __syncthreads();
    } else {
        for (int r0 = 0; r0 < cnt; r0 += THREADS) {
            const int j = r0 + tid;
            const int my = j < cnt ? __float_as_int(__ldg(&src[3 * (size_t)j + 1]).w) : 0x7fffffff;
            int rank = 0;
            for (int c0 = 0; c0 < cnt; c0 += kIdChunk) {
                const int m = min(kIdChunk, cnt - c0);
                __syncthreads();
                for (int k = tid; k < m; k += THREADS) sid[k] = __float_as_int(__ldg(&src[3 * (size_t)(c0 + k) + 1]).w);
                __syncthreads();
                for (int k = 0; k < m; ++k) count_if_less(rank, sid[k], my);
            }
            if (j < cnt && rank < GS2D_MAX_PER_TILE) {
                sq[3 * rank + 0] = __ldg(&src[3 * (size_t)j + 0]);
                sq[3 * rank + 1] = __ldg(&src[3 * (size_t)j + 1]);
                sq[3 * rank + 2] = __ldg(&src[3 * (size_t)j + 2]);
            }
            if (publish && j < cnt)
                reinterpret_cast<int*>(const_cast<float4*>(src) + 3 * (size_t)j + 2)[3] = rank;
        }
        __syncthreads();
    }
    return nb;
}

// Batch `base / 256` of a tile whose ranks stage_tile_sorted published: the records with rank in
// [base, base + 256), at their place.  The caller has made sure nobody still reads sq.  Ranks are
// read past L1 (another thread of the CTA wrote them, before a barrier); no scratch is used, so this
// also runs while sv holds v_out.
template <int THREADS>
__device__ __forceinline__ int stage_tile_window(const float4* __restrict__ src, int cnt, int base, float4* sq,
                                                 int tid) {
    for (int j = tid; j < cnt; j += THREADS) {
        const int rank = __ldcg(reinterpret_cast<const int*>(src + 3 * (size_t)j + 2) + 3) - base;
        if (rank >= 0 && rank < GS2D_MAX_PER_TILE) {
            sq[3 * rank + 0] = __ldg(&src[3 * (size_t)j + 0]);
            sq[3 * rank + 1] = __ldg(&src[3 * (size_t)j + 1]);
            sq[3 * rank + 2] = __ldg(&src[3 * (size_t)j + 2]);
        }
    }
    __syncthreads();
    return min(cnt - base, GS2D_MAX_PER_TILE);
}

// cnt <= 32: every warp stages the whole batch for itself.  Lane j holds record j in registers (in slot
// mode the three loads were issued together with the count: their address does not depend on it),
// ranks its Gaussian id against the other lanes' by shuffles and stores the record at its rank into the
// warp's own copy of the batch -- no block barrier, and the two warps of a tile never wait for each
// other before the loss.  Same order as stage_tile_sorted (ids are distinct within a tile).
__device__ __forceinline__ void stage_tile_warp(float4 r0, float4 r1, float4 r2, int cnt, float4* sqw, int lane) {
    const int my = lane < cnt ? __float_as_int(r1.w) : 0x7fffffff;
    int rank = 0;
    // lanes past the batch hold INT_MAX, which is below nobody's id: the loop may run on to a multiple of 4
    // fully unrolled in groups of four: the source lane is an immediate of the shuffle, and the count is a
    // predicated add (count_if_less; the compiler's form of `rank += a < b` was add + predicated move) -- 3.5
    // instead of 5.5 instructions per id: 46.3 -> 45.4 us at C2 (profiles/r2_ab_rank_unrolled.txt)
#pragma unroll
    for (int g = 0; g < 8; ++g) {
        if (4 * g >= cnt) break;   // uniform over the warp
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int other = __shfl_sync(0xffffffffu, my, 4 * g + j);
            count_if_less(rank, other, my);
        }
    }
    GS2D_ASSERT(cnt >= 0 && cnt <= 32);
    if (lane < cnt) {
        GS2D_ASSERT(rank >= 0 && rank < cnt);      // ids within a tile are distinct: ranks are a permutation
        sqw[3 * rank + 0] = r0;
        sqw[3 * rank + 1] = r1;
        sqw[3 * rank + 2] = r2;
    }
    __syncwarp();
}

// Sum 8 per-lane values over each group of 8 consecutive lanes with 7 shuffles: lane l of a
// group ends up with the group's total of v[l].
__device__ __forceinline__ float group8_reduce8_t(float (&v)[8], int lane) {
    const bool b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
    float w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float send = b2 ? v[k] : v[4 + k];
        const float keep = b2 ? v[4 + k] : v[k];
        w[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    float u[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float send = b1 ? w[k] : w[2 + k];
        const float keep = b1 ? w[2 + k] : w[k];
        u[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    const float send = b0 ? u[0] : u[1];
    const float keep = b0 ? u[1] : u[0];
    return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

// The same sums left as two 16-byte halves: after 12 shuffles every lane of a group whose bit 2 is clear holds
// the totals of v[0..3] in out[0..3], every lane whose bit 2 is set those of v[4..7] -- for one vectorised
// reduction (red.global.add.v4.f32) per half instead of eight scalar ones.
__device__ __forceinline__ void group8_reduce8_v4(float (&v)[8], int lane, float (&out)[4]) {
    const bool b2 = lane & 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float send = b2 ? v[k] : v[4 + k];
        const float keep = b2 ? v[4 + k] : v[k];
        out[k] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) out[k] += __shfl_xor_sync(0xffffffffu, out[k], 2);
#pragma unroll
    for (int k = 0; k < 4; ++k) out[k] += __shfl_xor_sync(0xffffffffu, out[k], 1);
}

// SYS: the row may live in a peer GPU's memory and other GPUs add to it too (system scope)
template <bool SYS>
__device__ __forceinline__ void red_add_v4(float* p, const float (&t)[4]) {
    if (SYS)
        asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(p), "f"(t[0]), "f"(t[1]), "f"(t[2]), "f"(t[3]) : "memory");
    else
        asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(p), "f"(t[0]), "f"(t[1]), "f"(t[2]), "f"(t[3]) : "memory");
}

}  // namespace gs2d
