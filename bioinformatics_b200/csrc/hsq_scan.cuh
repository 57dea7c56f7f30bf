// hsq_scan.cuh -- exclusive prefix sum of int64 (three-phase: tile sums, scan of sums, tile rescan).
// Used for ragged output offsets.  HBM-bound: reads the input twice, writes once.
#pragma once
#include <stdint.h>

namespace hsq {

static constexpr int SCAN_BLOCK = 256;
static constexpr int SCAN_ITEMS = 8;
static constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

__device__ inline int64_t block_exclusive_scan(int64_t v, int64_t* total) {  // SCAN_BLOCK threads
    __shared__ int64_t warp_sums[SCAN_BLOCK / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int64_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
        int64_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        int64_t s = lane < SCAN_BLOCK / 32 ? warp_sums[lane] : 0;
        for (int o = 1; o < SCAN_BLOCK / 32; o <<= 1) {
            int64_t y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += y;
        }
        if (lane < SCAN_BLOCK / 32) warp_sums[lane] = s;
    }
    __syncthreads();
    int64_t base = wid > 0 ? warp_sums[wid - 1] : 0;
    *total = warp_sums[SCAN_BLOCK / 32 - 1];
    __syncthreads();
    return base + x - v;
}

__global__ void scan_tile_sums_kernel(const int64_t* __restrict__ in, int64_t n, int64_t* __restrict__ tile_sums) {
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int64_t s = 0;
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (base + k < n) s += in[base + k];
    int64_t total;
    block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void scan_sums_kernel(int64_t* __restrict__ tile_sums, int64_t n_tiles) {  // one block
    __shared__ int64_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t b = 0; b < n_tiles; b += SCAN_BLOCK) {
        int64_t i = b + threadIdx.x;
        int64_t v = i < n_tiles ? tile_sums[i] : 0;
        int64_t total;
        int64_t ex = block_exclusive_scan(v, &total);
        if (i < n_tiles) tile_sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry += total;
        __syncthreads();
    }
}

// out[i] = sum(in[0..i)), i in [0, n]; out has n+1 entries
__global__ void scan_apply_kernel(const int64_t* __restrict__ in, int64_t n, const int64_t* __restrict__ tile_sums,
                                  int64_t* __restrict__ out) {
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int64_t v[SCAN_ITEMS];
    int64_t s = 0;
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        s += v[k];
    }
    int64_t total;
    int64_t ex = block_exclusive_scan(s, &total) + tile_sums[blockIdx.x];
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k <= n) out[base + k] = ex;
        ex += v[k];
    }
}

}  // namespace hsq
