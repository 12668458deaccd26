// xyz_tiles.cuh — double-buffered staging of (x,y,z) candidate tiles into shared memory.
//
// Used by query_ball_point and three_nn: every thread of a CTA scans the same candidate
// list, so the list is streamed through shared memory once per CTA and read back as
// warp-wide broadcasts.  Full tiles are moved by the TMA engine (1-D cp.async.bulk,
// SASS UBLKCP) and complete on an mbarrier, so the copy of tile i+1 overlaps the
// distance tests on tile i without occupying any thread.  Tiles the bulk copy cannot
// take (source not 16-byte aligned, or a tail whose byte count is not a multiple of 16)
// fall back to a cooperative ld.global/st.shared copy.
#pragma once
#include "common.cuh"

namespace hf {

template <int TILE>
struct XyzTilePipe {
    static_assert(TILE % 4 == 0, "tile bytes must be a multiple of 16");
    float *stage[2];
    uint64_t *bars;  // [2]
    const float *src;  // scene base, n*3 floats
    int n;
    int ntiles;
    bool tma_ok;
    uint32_t phase[2];
    int pending;  // tile index with a bulk copy in flight issued by thread 0, or -1

    __device__ __forceinline__ void init(float *s0, float *s1, uint64_t *b, const float *scene_xyz, int n_) {
        stage[0] = s0;
        stage[1] = s1;
        bars = b;
        src = scene_xyz;
        n = n_;
        ntiles = (n_ + TILE - 1) / TILE;
        tma_ok = ((reinterpret_cast<uintptr_t>(scene_xyz) & 15) == 0);
        phase[0] = phase[1] = 0;
        pending = -1;
        if (threadIdx.x == 0) {
            mbar_init(&bars[0], 1);
            mbar_init(&bars[1], 1);
            mbar_fence_init();
        }
        __syncthreads();
    }
    __device__ __forceinline__ int count(int i) const {
        int c = n - i * TILE;
        return c < TILE ? c : TILE;
    }
    __device__ __forceinline__ bool uses_tma(int i) const { return tma_ok && ((count(i) * 12) & 15) == 0; }
    // Called by all threads; only thread 0 issues.  The target stage must be free.
    __device__ __forceinline__ void prefetch(int i) {
        if (i < ntiles && uses_tma(i) && threadIdx.x == 0) {
            uint32_t bytes = (uint32_t)count(i) * 12u;
            mbar_arrive_expect_tx(&bars[i & 1], bytes);
            tma_load_1d(stage[i & 1], src + (size_t)i * TILE * 3, bytes, &bars[i & 1]);
            pending = i;
        }
    }
    // Make tile i readable in stage[i&1] for every thread of the CTA.
    __device__ __forceinline__ const float *acquire(int i) {
        int s = i & 1;
        if (uses_tma(i)) {
            mbar_wait(&bars[s], phase[s]);
            phase[s] ^= 1u;
            if (threadIdx.x == 0 && pending == i) pending = -1;
        } else {
            const float *g = src + (size_t)i * TILE * 3;
            int words = count(i) * 3;
            for (int w = threadIdx.x; w < words; w += blockDim.x) stage[s][w] = g[w];
            __syncthreads();
        }
        return stage[s];
    }
    // Before the CTA may exit early: no bulk copy may still be writing its shared memory.
    __device__ __forceinline__ void drain() {
        if (threadIdx.x == 0 && pending >= 0) {
            mbar_wait(&bars[pending & 1], phase[pending & 1]);
            pending = -1;
        }
    }
};

}  // namespace hf
