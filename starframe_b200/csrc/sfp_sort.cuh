// sfp_sort.cuh — device-wide exclusive scan and LSD radix sort (8-bit digits), hand-written.
//
// Both primitives take their element count from DEVICE memory (n_ptr), so the host can queue a
// whole frame without reading anything back: grids are sized for the buffer capacity and blocks
// past the live count retire immediately.  Everything is integer work and order-deterministic:
// no atomics on global memory, stable ranking inside a tile via warp match + warp-private
// shared-memory counters.
//
// Reference: none — the reference's broadphase is a CPU structure (not mounted).
// north_star asks for "a device radix-sort spatial hash".
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sfpk {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 2048

constexpr int RADIX_THREADS = 256;
constexpr int RADIX_ITEMS = 8;
constexpr int RADIX_TILE = RADIX_THREADS * RADIX_ITEMS;  // 2048
constexpr int RADIX_BINS = 256;
constexpr int RADIX_WARPS = RADIX_THREADS / 32;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// exclusive scan of one value per thread across a 256-thread block; returns the exclusive
// prefix and writes the block total to *total (valid in every thread after the call)
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total, uint32_t* smem9) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v, lane);
    if (lane == 31) smem9[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < (int)(blockDim.x >> 5)) ? smem9[lane] : 0u;
        uint32_t winc = warp_incl_scan(w, lane);
        if (lane < (int)(blockDim.x >> 5)) smem9[lane] = winc - w;
        if (lane == 31) smem9[8] = winc;
    }
    __syncthreads();
    uint32_t r = smem9[warp] + inc - v;
    *total = smem9[8];
    __syncthreads();
    return r;
}

// ---- scan, phase 1: per-tile sums ---------------------------------------------------------
static __global__ void __launch_bounds__(SCAN_THREADS)
k_scan_tile_sums(const uint32_t* __restrict__ in, const uint32_t* __restrict__ n_ptr,
                 uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t sm[9];
    const uint32_t n = *n_ptr;
    const uint32_t base = blockIdx.x * SCAN_TILE;
    uint32_t s = 0;
    if (base < n) {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) {
            uint32_t i = base + k * SCAN_THREADS + threadIdx.x;
            if (i < n) s += in[i];
        }
    }
    uint32_t total;
    block_excl_scan(s, &total, sm);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// ---- scan, phase 2: one block scans the tile sums in place --------------------------------
static __global__ void __launch_bounds__(1024)
k_scan_sums(uint32_t* __restrict__ tile_sums, uint32_t n_tiles, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t sm[33];
    __shared__ uint32_t carry_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t v = (i < n_tiles) ? tile_sums[i] : 0u;
        uint32_t inc = warp_incl_scan(v, lane);
        if (lane == 31) sm[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = sm[lane];
            uint32_t winc = warp_incl_scan(w, lane);
            sm[lane] = winc - w;
            if (lane == 31) sm[32] = winc;
        }
        __syncthreads();
        uint32_t carry = carry_s;
        if (i < n_tiles) tile_sums[i] = carry + sm[warp] + inc - v;
        uint32_t chunk_total = sm[32];
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + chunk_total;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}

// ---- scan, phase 3: rescan each tile with its offset; out[n] = total ----------------------
static __global__ void __launch_bounds__(SCAN_THREADS)
k_scan_apply(const uint32_t* __restrict__ in, const uint32_t* __restrict__ n_ptr,
             const uint32_t* __restrict__ tile_sums, uint32_t* __restrict__ out) {
    __shared__ uint32_t sm[9];
    const uint32_t n = *n_ptr;
    const uint32_t base = blockIdx.x * SCAN_TILE;
    if (base > n) return;
    // thread t owns SCAN_ITEMS consecutive elements: blocked arrangement keeps the scan simple
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
    const uint32_t first = base + threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        uint32_t i = first + k;
        v[k] = (i < n) ? in[i] : 0u;
        s += v[k];
    }
    uint32_t total;
    uint32_t ex = block_excl_scan(s, &total, sm) + tile_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        uint32_t i = first + k;
        if (i <= n) out[i] = ex;  // i == n receives the grand total
        ex += v[k];
    }
}

// Exclusive scan of in[0..n) into out[0..n], out[n] = total; *total_out = total.
// `cap` = the largest n this buffer can hold (sizes the grids); tile_sums holds cap/2048+2.
inline cudaError_t scan_exclusive(const uint32_t* in, uint32_t* out, const uint32_t* n_ptr,
                                  uint32_t cap, uint32_t* tile_sums, uint32_t* total_out,
                                  cudaStream_t st) {
    uint32_t tiles = cap / SCAN_TILE + 1;
    k_scan_tile_sums<<<tiles, SCAN_THREADS, 0, st>>>(in, n_ptr, tile_sums);
    k_scan_sums<<<1, 1024, 0, st>>>(tile_sums, tiles, total_out);
    k_scan_apply<<<tiles, SCAN_THREADS, 0, st>>>(in, n_ptr, tile_sums, out);
    return cudaGetLastError();
}

// ---- radix sort pass ----------------------------------------------------------------------
// hist layout: hist[block * 256 + digit]; every scatter block derives its own global bases from
// it (a coalesced sweep over the live blocks + one block scan), so no separate scan launch
static __global__ void __launch_bounds__(RADIX_THREADS)
k_radix_hist(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ n_ptr, int shift,
             uint32_t* __restrict__ hist) {
    __shared__ uint32_t h[RADIX_BINS];
    const uint32_t n = *n_ptr;
    const uint32_t base = blockIdx.x * RADIX_TILE;
    h[threadIdx.x] = 0;
    __syncthreads();
    if (base < n) {
#pragma unroll
        for (int k = 0; k < RADIX_ITEMS; ++k) {
            uint32_t i = base + k * RADIX_THREADS + threadIdx.x;
            if (i < n) atomicAdd(&h[(keys[i] >> shift) & 0xffu], 1u);
        }
    }
    __syncthreads();
    hist[blockIdx.x * RADIX_BINS + threadIdx.x] = h[threadIdx.x];
}

// Above RADIX_INLINE_BLOCKS live blocks the per-block sweep over the histogram would be
// quadratic in the block count, so one block per digit scans that digit's column in place and
// leaves the column total behind the histogram (hist[n_blocks * 256 + digit]).
constexpr uint32_t RADIX_INLINE_BLOCKS = 32;
static __global__ void __launch_bounds__(256)
k_radix_colscan(uint32_t* __restrict__ hist, uint32_t n_blocks, const uint32_t* __restrict__ n_ptr) {
    __shared__ uint32_t sm[9];
    __shared__ uint32_t carry;
    const uint32_t n = *n_ptr;
    const uint32_t live = min(n_blocks, (n + RADIX_TILE - 1) / RADIX_TILE);
    if (live <= RADIX_INLINE_BLOCKS) return;
    const uint32_t d = blockIdx.x;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < live; base += 256) {
        uint32_t b = base + threadIdx.x;
        uint32_t v = (b < live) ? hist[(size_t)b * RADIX_BINS + d] : 0u;
        uint32_t total;
        uint32_t ex = block_excl_scan(v, &total, sm);
        uint32_t c0 = carry;
        if (b < live) hist[(size_t)b * RADIX_BINS + d] = c0 + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry = c0 + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) hist[(size_t)n_blocks * RADIX_BINS + d] = carry;
}

// stable scatter.  vals_in == nullptr means "value = source index" (used to build permutations)
static __global__ void __launch_bounds__(RADIX_THREADS)
k_radix_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                const uint32_t* __restrict__ n_ptr, int shift,
                const uint32_t* __restrict__ hist_scanned) {
    __shared__ uint32_t wcnt[RADIX_WARPS][RADIX_BINS];
    __shared__ uint32_t gbase[RADIX_BINS];
    __shared__ uint32_t scan_sm[9];
    const uint32_t n = *n_ptr;
    const uint32_t base = blockIdx.x * RADIX_TILE;
    if (base >= n) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int w = 0; w < RADIX_WARPS; ++w) wcnt[w][threadIdx.x] = 0;
    __syncthreads();

    uint32_t key[RADIX_ITEMS], val[RADIX_ITEMS], rank[RADIX_ITEMS];
    const uint32_t wbase = base + warp * (32 * RADIX_ITEMS);
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RADIX_ITEMS; ++r) {
        uint32_t i = wbase + r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? keys_in[i] : 0u;
        val[r] = valid ? (vals_in ? vals_in[i] : i) : 0u;
        uint32_t d = valid ? ((key[r] >> shift) & 0xffu) : 0x100u;
        uint32_t peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t b = 0;
        if (lane == leader && valid) {
            b = wcnt[warp][d];
            wcnt[warp][d] = b + __popc(peers);
        }
        b = __shfl_sync(0xffffffffu, b, leader);
        rank[r] = b + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    {
        // exclusive scan over warps for digit = threadIdx.x
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RADIX_WARPS; ++w) {
            uint32_t t = wcnt[w][threadIdx.x];
            wcnt[w][threadIdx.x] = run;
            run += t;
        }
        // global base of (this block, digit): all keys with a smaller digit, plus the keys with
        // this digit in earlier blocks
        const uint32_t live = (n + RADIX_TILE - 1) / RADIX_TILE;
        uint32_t before = 0, total = 0;
        if (live > RADIX_INLINE_BLOCKS) {
            // many blocks: k_radix_colscan already turned each digit's column into running sums
            before = hist_scanned[blockIdx.x * RADIX_BINS + threadIdx.x];
            total = hist_scanned[(size_t)gridDim.x * RADIX_BINS + threadIdx.x];
        } else {
#pragma unroll 8
            for (uint32_t b = 0; b < live; ++b) {
                uint32_t v = hist_scanned[b * RADIX_BINS + threadIdx.x];
                total += v;
                before += (b < blockIdx.x) ? v : 0u;
            }
        }
        uint32_t all;
        uint32_t ex = block_excl_scan(total, &all, scan_sm);
        gbase[threadIdx.x] = ex + before;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RADIX_ITEMS; ++r) {
        uint32_t i = wbase + r * 32 + lane;
        if (i < n) {
            uint32_t d = (key[r] >> shift) & 0xffu;
            uint32_t pos = gbase[d] + wcnt[warp][d] + rank[r];
            keys_out[pos] = key[r];
            vals_out[pos] = val[r];
        }
    }
}

inline uint32_t radix_blocks(uint32_t cap) { return cap / RADIX_TILE + 1; }

// one stable pass over bits [shift, shift+8)
inline cudaError_t radix_pass(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out,
                              uint32_t* vals_out, const uint32_t* n_ptr, uint32_t cap, int shift,
                              uint32_t* hist, cudaStream_t st) {
    uint32_t nb = radix_blocks(cap);
    k_radix_hist<<<nb, RADIX_THREADS, 0, st>>>(keys_in, n_ptr, shift, hist);
    if (nb > RADIX_INLINE_BLOCKS) k_radix_colscan<<<RADIX_BINS, 256, 0, st>>>(hist, nb, n_ptr);
    k_radix_scatter<<<nb, RADIX_THREADS, 0, st>>>(keys_in, vals_in, keys_out, vals_out, n_ptr,
                                                  shift, hist);
    return cudaGetLastError();
}

}  // namespace sfpk
