// sfp_batch.cu — batched independent worlds (BASELINE.json config 5): a thread block runs a world's
// WHOLE frame — broadphase, colouring, all substeps — and one launch runs the frame of all worlds.
//
// Why a different shape from sfp_world.cu: a 512-body world is far too small to fill a B200, and
// its ~350 dependent phases per frame would each cost a grid-wide barrier.  Inside one CTA the
// same phases are separated by __syncthreads() (tens of ns) and the world's state (45 KB of
// bodies in shared memory, a few hundred KB of contacts) stays in that SM / the L2.  Two CTAs
// are resident per SM (SFP_BCTAS); a batch with more worlds than resident CTAs launches exactly
// that many, which take worlds from a queue in device memory (k_batch_frame, BatchQueue).
//
// The stages are the SAME __device__ functions as the single-world path (sfp_device.cuh), called
// with (tid, nth) = (threadIdx.x, blockDim.x) on a per-world view of the buffers, so a world
// stepped in a batch is bit-identical to the same world stepped alone, and to the oracle.  Only
// the broadphase differs in structure: with 512 bodies a sweep along x over boxes sorted in
// shared memory is cheaper than hashing and yields exactly the same pair SET, which is all that
// matters downstream (the colouring depends on the set, not on the order); and the narrowphase
// visits the pairs grouped by class through shared-memory lists (batch_narrowphase).
//
// Reference: none for batching — north_star: "batched independent worlds are partitioned across
// the 8 GPUs of one box, with NCCL used only to gather statistics".
#include <cuda_pipeline.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "sfp.h"
#include "sfp_device.cuh"
#include "sfp_host.h"
#include "sfp_sort.cuh"

using namespace sfpk;

namespace {

#ifndef SFP_BTPB
#define SFP_BTPB 256
#endif
constexpr int BTPB = SFP_BTPB;  // threads per world

struct BatchCtx {
    Ctx base;              // pointers to world 0's slabs; parameters
    uint32_t bpw;          // bodies per world
    uint32_t pair_cap;     // pairs per world
    uint32_t n_worlds;
    Counters* counters;    // [n_worlds]
    uint32_t* scratch;     // [n_worlds][bpw] per-body pair counts / offsets
    uint32_t tab_slots;    // colour-table slots per world (power of two)
    uint32_t world_base;   // first world of this launch
    uint32_t n_launch;     // worlds of this launch: [world_base, world_base + n_launch)
    uint32_t n_whole;      // the first n_whole of them are handed out whole (k_batch_frame)
    struct BatchQueue* queue;  // this launch's queue, or null: CTA i runs world world_base + i
    uint32_t* ring;        // [n_launch - n_whole] ring of worlds between two stages
    uint32_t* progress;    // [n_worlds] stages done of a world that is handed out by the stage
    uint32_t np_list;      // 1: the dynamic shared memory holds the narrowphase list (pair_cap x u16) after the bodies
    uint32_t joint_cap;    // joint slots per world (0: the batch has no joints)
    const uint32_t* n_joints;  // [n_worlds] joints in use per world, or null
};

__device__ __forceinline__ Ctx world_view(const BatchCtx& b, uint32_t w) {
    Ctx c = b.base;
    const size_t bo = (size_t)w * b.bpw, po = (size_t)w * b.pair_cap;
    c.pose += bo;
    c.prev += bo;
    c.vel += bo;
    c.shape += bo;
    c.mat += bo;
    c.minv += bo;
    c.aabb += bo;
    c.meta += bo;
    c.mask += bo;
    c.claim0 += bo;
    c.claim1 += bo;
    c.pairs_raw += po;
    c.pairs += po;
    c.pcol += po;
    c.ca += po;
    c.cb += po;
    c.cp0 += po;
    c.cp1 += po;
    c.cs += po;
    c.lam += po;
    if (c.np_order) c.np_order += po;
    c.tab_old += (size_t)w * b.tab_slots;
    c.tab_new += (size_t)w * b.tab_slots;
    c.cnt = b.counters + w;
    c.n = b.bpw;
    c.nj = 0;
    if (b.joint_cap) {
        const size_t jo = (size_t)w * b.joint_cap;
        c.jbod_in += jo;
        c.janc_in += jo;
        c.jpar_in += jo;
        c.jbod += jo;
        c.janc += jo;
        c.jpar += jo;
        c.jcol += jo;
        c.nj = b.n_joints[w];
    }
    c.cap_pairs = b.pair_cap;
    return c;
}

// Broadphase of one world: every two bodies whose boxes overlap (and whose layers collide, not
// both static), appended to pairs_raw through a shared-memory cursor (integer atomics).  The
// ORDER of pairs_raw is therefore arbitrary, which changes no result: colouring depends on the
// pair set only.
//   s_sort == null (no room in shared memory): all pairs, body a scans b > a; thread t takes
//   bodies t and n-1-t, so every thread runs n-1 tests.
//   otherwise sweep and prune along x: the bodies are sorted by the low edge of their box
//   (bitonic sort of key = 20 bits of the edge | 12 bits of index in shared memory), and a body
//   tests only the bodies behind it in that order whose low edge is not beyond its high edge —
//   ~30 instead of 256 for a 512-body pile.  The few bodies that span the world (floor, walls:
//   wider than 8 cells) are kept out of the sort and tested against everybody.
__device__ void batch_broadphase(const Ctx& c, float dt, uint32_t* s_sort, uint32_t sort_cap) {
    const uint32_t tid = threadIdx.x, nth = blockDim.x, n = c.n;
    for (uint32_t i = tid; i < n; i += nth) {
        f4 bb = sfp::body_aabb(ld4(c.pose, i), ld4(c.vel, i), ld4(c.shape, i), c.aabb_margin,
                               c.vel_expand, dt, reinterpret_cast<const sfp::f4*>(c.polytab));
        int x0, y0, x1, y1;
        if (!sfp::cell_range(bb, c.inv_cell, &x0, &y0, &x1, &y1)) atomicOr(&c.cnt->overflow, 16u);
        st4(c.aabb, i, bb);
        f2 mi = ld2(c.minv, i);
        c.meta[i] = make_int4(x0, y0, (int)sfp::f2u(c.mat[i].w), (mi.x > 0.0f || mi.y > 0.0f) ? 1 : 0);
    }
    constexpr uint32_t MAX_WIDE = 32u;
    __shared__ uint32_t cursor, s_nwide;
    __shared__ uint32_t s_wide[MAX_WIDE];
    if (tid == 0) {
        c.cnt->n_colours = 0;
        c.cnt->colour_rounds = 0;
        c.cnt->n_entries = 0;
        cursor = 0;
        s_nwide = 0;
    }
    __syncthreads();
    auto test = [&](f4 bba, int4 ma, uint32_t b) {
        if (!sfp::aabb_overlap(bba, ld4(c.aabb, b))) return false;  // rejects almost everything
        int4 mb = c.meta[b];
        if ((ma.w | mb.w) == 0) return false;
        return sfp::layers_collide((uint32_t)ma.z, (uint32_t)mb.z);
    };
    auto emit = [&](uint32_t a, uint32_t b) {
        uint32_t pos = atomicAdd(&cursor, 1u);
        if (pos < c.cap_pairs) c.pairs_raw[pos] = a < b ? make_uint2(a, b) : make_uint2(b, a);
    };
    uint32_t N = 1u;
    while (N < n) N <<= 1;
    if (s_sort == nullptr || N > sort_cap || n >= 4095u) {
        const uint32_t half = (n + 1u) / 2u;
        for (uint32_t t = tid; t < half; t += nth) {
            for (int side = 0; side < 2; ++side) {
                const uint32_t a = side ? (n - 1u - t) : t;
                if (side && a == t) break;  // middle body of an odd count
                f4 bba = ld4(c.aabb, a);
                int4 ma = c.meta[a];
                for (uint32_t b = a + 1; b < n; ++b)
                    if (test(bba, ma, b)) emit(a, b);
            }
        }
    } else {
        constexpr uint32_t NONE = 0xffffffffu, IDX = 0xfffu;
        // a float's bits, ordered as unsigned integers
        auto ordered = [](float x) {
            const uint32_t u = sfp::f2u(x);
            return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
        };
        for (uint32_t i = tid; i < N; i += nth) {
            uint32_t key = NONE;
            if (i < n) {
                f4 bb = ld4(c.aabb, i);
                bool wide = !((bb.z - bb.x) * c.inv_cell <= 8.0f);  // (also a box that is not finite)
                if (wide) {
                    const uint32_t slot = atomicAdd(&s_nwide, 1u);
                    if (slot < MAX_WIDE) s_wide[slot] = i;
                    else wide = false;  // no room in the list: it takes part in the sweep
                }
                if (!wide) key = (ordered(bb.x) & ~IDX) | i;
            }
            s_sort[i] = key;
        }
        __syncthreads();
        for (uint32_t k = 2u; k <= N; k <<= 1) {
            for (uint32_t j = k >> 1; j > 0u; j >>= 1) {
                for (uint32_t t = tid; t < (N >> 1); t += nth) {
                    const uint32_t lo = ((t & ~(j - 1u)) << 1) | (t & (j - 1u)), hi = lo + j;
                    const uint32_t x = s_sort[lo], y = s_sort[hi];
                    if ((x > y) == ((lo & k) == 0u)) {
                        s_sort[lo] = y;
                        s_sort[hi] = x;
                    }
                }
                __syncthreads();
            }
        }
        const uint32_t nwide = s_nwide < MAX_WIDE ? s_nwide : MAX_WIDE;
        // the sweep: position p against the positions behind it
        for (uint32_t p = tid; p < N; p += nth) {
            const uint32_t ka = s_sort[p];
            if (ka == NONE) continue;  // (the sentinels sort to the end: nothing behind them either)
            const uint32_t a = ka & IDX;
            const f4 bba = ld4(c.aabb, a);
            const int4 ma = c.meta[a];
            const uint32_t limit = ordered(bba.z);
            for (uint32_t q = p + 1u; q < N; ++q) {
                const uint32_t kb = s_sort[q];
                // kb without the index bits is <= the ordered low edge of b: past the limit, b
                // and everybody behind b start beyond a's high edge
                if (kb == NONE || (kb & ~IDX) > limit) break;
                if (test(bba, ma, kb & IDX)) emit(a, kb & IDX);
            }
        }
        // the wide bodies against everybody (two wide bodies: once, from the higher index)
        for (uint32_t i = tid; i < n; i += nth) {
            bool listed = false;
            for (uint32_t j = 0; j < nwide; ++j) listed = listed || s_wide[j] == i;
            const f4 bbi = ld4(c.aabb, i);
            const int4 mi = c.meta[i];
            for (uint32_t j = 0; j < nwide; ++j) {
                const uint32_t wj = s_wide[j];
                if (wj == i || (listed && i < wj)) continue;
                if (test(bbi, mi, wj)) emit(i, wj);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t P = cursor;
        if (P > c.cap_pairs) {
            atomicOr(&c.cnt->overflow, 2u);
            P = c.cap_pairs;
        }
        c.cnt->n_pairs = P;
    }
    __syncthreads();
}

// bucket pairs by (colour, narrowphase class): counting sort on the 8-bit key written by
// stage_colour_take.  Ranks inside a bin come from a shared-memory cursor (integer atomics): the
// order INSIDE a (colour, type) bin is therefore arbitrary, which changes no result — constraints
// of one colour touch disjoint bodies — only which contact slot a pair happens to use.
constexpr int BUCKET_BINS = SFP_MAX_COLOURS << SFP_TYPE_BITS;
__device__ void batch_bucket(const Ctx& c, uint32_t* smem9) {
    __shared__ uint32_t bins[BUCKET_BINS];
    const uint32_t tid = threadIdx.x, nth = blockDim.x;
    const uint32_t P = c.cnt->n_pairs;
    for (uint32_t k = tid; k < BUCKET_BINS; k += nth) bins[k] = 0;
    __syncthreads();
    for (uint32_t k = tid; k < P; k += nth) atomicAdd(&bins[c.pcol[k] & (BUCKET_BINS - 1)], 1u);
    __syncthreads();
    // exclusive scan of the bins: BUCKET_BINS / nth consecutive bins per thread
    constexpr int PER = (BUCKET_BINS + BTPB - 1) / BTPB;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        v[j] = (tid * PER + j < (uint32_t)BUCKET_BINS) ? bins[tid * PER + j] : 0u;
        sum += v[j];
    }
    uint32_t total;
    uint32_t ex = block_excl_scan(sum, &total, smem9);
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        if (tid * PER + j < (uint32_t)BUCKET_BINS) bins[tid * PER + j] = ex;
        ex += v[j];
    }
    __syncthreads();
    // joints: a world holds at most a few hundred, so each one counts its predecessors — the
    // joints of a lower colour, and those of its own colour that come earlier in upload order
    // (stable: the serial group, whose members may share bodies, keeps upload order)
    const uint32_t nj = c.nj;
    for (uint32_t k = tid; k < nj; k += nth) {
        const uint32_t col = c.jcol[k];
        uint32_t pos = 0;
        for (uint32_t i = 0; i < nj; ++i) {
            const uint32_t ci = c.jcol[i];
            pos += (ci < col || (ci == col && i < k)) ? 1u : 0u;
        }
        c.jbod[pos] = c.jbod_in[k];
        c.janc[pos] = c.janc_in[k];
        c.jpar[pos] = c.jpar_in[k];
    }
    for (uint32_t col = tid; col <= SFP_MAX_COLOURS; col += nth) {
        c.cnt->pair_col_off[col] = (col < SFP_MAX_COLOURS) ? bins[col << SFP_TYPE_BITS] : P;
        uint32_t below = 0;
        for (uint32_t i = 0; i < nj; ++i) below += (c.jcol[i] < col) ? 1u : 0u;
        c.cnt->joint_col_off[col] = below;
    }
    __syncthreads();
    for (uint32_t k = tid; k < P; k += nth) {
        uint32_t pos = atomicAdd(&bins[c.pcol[k] & (BUCKET_BINS - 1)], 1u);
        c.pairs[pos] = c.pairs_raw[k];
    }
    __syncthreads();
    // Narrowphase order: the same positions grouped by class alone.  A (colour, class) bucket
    // of a 512-body world holds a few dozen pairs, so in position order nearly every warp
    // would run two of the three narrowphase routines.  bins[key] is now the END of bucket
    // key; obase[key] = where the bucket starts in class-major order.
    if (c.np_order != nullptr) {
        __shared__ uint32_t obase[BUCKET_BINS];
        constexpr uint32_t NCLS = 1u << SFP_TYPE_BITS;
        if (tid < NCLS) {
            uint32_t run = 0;  // pairs of the lower classes
            for (uint32_t cl = 0; cl < tid; ++cl)
                for (uint32_t col = 0; col < SFP_MAX_COLOURS; ++col) {
                    const uint32_t key = (col << SFP_TYPE_BITS) | cl;
                    run += bins[key] - (key ? bins[key - 1] : 0u);
                }
            if (tid > 0u) c.cnt->np_cls_off[tid - 1u] = run;
            for (uint32_t col = 0; col < SFP_MAX_COLOURS; ++col) {
                const uint32_t key = (col << SFP_TYPE_BITS) | tid;
                obase[key] = run;
                run += bins[key] - (key ? bins[key - 1] : 0u);
            }
        }
        __syncthreads();
        const uint32_t lane = tid & 31u, warp = tid >> 5, nwarp = nth >> 5;
        for (uint32_t key = warp; key < (uint32_t)BUCKET_BINS; key += nwarp) {
            const uint32_t start = key ? bins[key - 1] : 0u, size = bins[key] - start;
            for (uint32_t j = lane; j < size; j += 32u) c.np_order[obase[key] + j] = start + j;
        }
        __syncthreads();
    }
}

// Narrowphase of one batched world, one substep.  Pass 1: every candidate pair (visited in class
// order, c.np_order) tests its bounding circles; a pair out of reach gets its empty record, a
// pair in reach is appended to its class's list in shared memory (s_list is laid out like
// np_order: class cl's list starts at its class offset, and holds at most the class's pairs).
// Pass 2: the SAT over the lists, densely — a thread takes the items tid, tid + nth, ... of the
// concatenated lists, so all warps but those on a class boundary run ONE routine with all lanes
// busy, and the pass takes ceil(pairs in reach / threads) rounds instead of one per 256
// candidates.  Which thread handles which pair changes nothing in the records.
template <bool POLYS>
__device__ __forceinline__ void batch_narrowphase(const Ctx& c, uint16_t* s_list, uint32_t tid,
                                                  uint32_t nth) {
    constexpr uint32_t NCLS = 1u << SFP_TYPE_BITS;
    __shared__ uint32_t s_cur[NCLS];
    const uint32_t P = c.cnt->n_pairs;
    const uint32_t off1 = c.cnt->np_cls_off[0], off2 = c.cnt->np_cls_off[1], off3 = c.cnt->np_cls_off[2];
    const uint32_t lane = tid & 31u;
    if (tid < NCLS) s_cur[tid] = 0u;
    __syncthreads();
    for (uint32_t base = 0; base < P; base += nth) {
        const uint32_t i = base + tid;
        bool near = false;
        uint32_t k = 0u, cls = 0u;
        if (i < P) {
            k = c.np_order[i];
            cls = (i >= off1 ? 1u : 0u) + (i >= off2 ? 1u : 0u) + (i >= off3 ? 1u : 0u);
            uint2 p = c.pairs[k];
            near = !sfp::narrowphase_far(ld4(c.pose, p.x), ld4(c.shape, p.x), ld4(c.pose, p.y),
                                         ld4(c.shape, p.y), c.contact_margin);
            if (!near) c.cb[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
        // a warp's 32 positions are consecutive in class order: one class, two on a boundary
#pragma unroll
        for (uint32_t cl = 0; cl < NCLS; ++cl) {
            const uint32_t m = __ballot_sync(0xffffffffu, near && cls == cl);
            if (m) {
                uint32_t at = 0;
                if (lane == 0) at = atomicAdd(&s_cur[cl], (uint32_t)__popc(m));
                at = __shfl_sync(0xffffffffu, at, 0);
                if (near && cls == cl) {
                    const uint32_t o = cl == 0u ? 0u : (cl == 1u ? off1 : (cl == 2u ? off2 : off3));
                    s_list[o + at + __popc(m & ((1u << lane) - 1u))] = (uint16_t)k;
                }
            }
        }
    }
    __syncthreads();
    const uint32_t n0 = s_cur[0], n1 = n0 + s_cur[1], n2 = n1 + s_cur[2], n3 = n2 + s_cur[3];
    uint32_t npts = 0, nman = 0;
    float maxpen = 0.0f;
    for (uint32_t idx = tid; idx < n3; idx += nth) {
        uint32_t at;
        if (idx < n0) at = idx;
        else if (idx < n1) at = off1 + (idx - n0);
        else if (idx < n2) at = off2 + (idx - n1);
        else at = off3 + (idx - n2);
        const uint32_t k2 = s_list[at];
        uint2 p = c.pairs[k2];
        sfp::ContactRec r = sfp::narrowphase_near<POLYS>(
            ld4(c.pose, p.x), ld4(c.vel, p.x), ld4(c.shape, p.x), ld4(c.mat, p.x), ld4(c.pose, p.y),
            ld4(c.vel, p.y), ld4(c.shape, p.y), ld4(c.mat, p.y), c.contact_margin,
            reinterpret_cast<const sfp::f4*>(c.polytab));
        uint32_t cnt = sfp::f2u(r.b.y);
        st4(c.cb, k2, r.b);
        if (cnt) {  // an empty manifold is fully described by count == 0
            st4(c.ca, k2, r.a);
            st4(c.cp0, k2, r.p0);
            st4(c.cp1, k2, r.p1);
            st4(c.cs, k2, r.s);
        }
        npts += cnt;
        nman += cnt ? 1u : 0u;
        if (cnt > 0) maxpen = fmaxf(maxpen, -r.s.x);
        if (cnt > 1) maxpen = fmaxf(maxpen, -r.s.y);
    }
    // warp-aggregated integer statistics (order-independent)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        npts += __shfl_xor_sync(0xffffffffu, npts, d);
        nman += __shfl_xor_sync(0xffffffffu, nman, d);
        maxpen = fmaxf(maxpen, __shfl_xor_sync(0xffffffffu, maxpen, d));
    }
    if (lane == 0 && (npts | nman)) {
        atomicAdd(&c.cnt->n_contacts, npts);
        atomicAdd(&c.cnt->n_manifolds, nman);
        atomicMax(&c.cnt->maxpen_bits, sfp::f2u(maxpen));
    }
}

// Shared-memory layout of one world (bytes per body): pose 16 | vel 16 | X 16 | meta 16 |
// minv 8 | mask 8 = 80.  X is time-shared: the AABBs during the broadphase, the two claim
// arrays during colouring, the substep record `prev` during the substeps.  Everything the
// solver gathers by body index therefore lives in this SM's shared memory; only the per-pair
// contact records (read and written by the pair's own thread, coalesced) stay in global memory.
// CTAs per SM.  Measured on 8192 worlds of 512 bodies: one 29.5 ms, TWO 19.7 ms, three 21.0 ms
// (256 threads each), four of 128 threads 24.3 ms.  Three need 80 registers (spills), and — the
// larger effect — every resident CTA is in a different phase of the frame, the kernel's hot code
// is ~70 KB against a 32 KB instruction cache level, and ncu shows "no instruction" as the
// largest stall of the working warps (profiles/r2_ncu_batch.md).
#ifndef SFP_BCTAS
#define SFP_BCTAS 2
#endif
// FULL: the variant that can collide general polygons and solve joints (more registers and
// code: two CTAs per SM); the host launches it only for batches that hold a polygon body or
// a joint, so that plain worlds keep the lean kernel
// stages [sa, se) of world w's frame (stage 0: pairs and colours; stage s >= 1: substep s)
template <bool SMEM, bool FULL>
__device__ __forceinline__ void batch_world_stages(const BatchCtx& b, uint32_t w, uint32_t sa, uint32_t se,
                                                float dt, uint32_t substeps, float rest_threshold) {
    extern __shared__ float4 dyn[];
    __shared__ uint32_t smem9[9];
    // colour order as a list, and a per-thread shared-memory slot into which the contact record
    // of this thread's first item of the NEXT colour is copied asynchronously (cp.async, no
    // registers held across the barrier): after the barrier only shared-memory reads remain
    // before the solve, instead of two dependent L2 / HBM round trips.
    __shared__ uint8_t s_order[SFP_MAX_COLOURS];
    __shared__ uint32_t s_nord;
    __shared__ float4 s_rec[4][BTPB];
    __shared__ float4 s_lam[BTPB];
    __shared__ uint2 s_pair[BTPB];
    const uint32_t tid = threadIdx.x, nth = blockDim.x;
    const float h = dt / (float)substeps, inv_h = 1.0f / h, inv_h2 = inv_h * inv_h;
    auto next_bit = [](unsigned long long m, int after) {
        unsigned long long r = (after >= 63) ? 0ull : (m & (~0ull << (after + 1)));
        return r ? (__ffsll((long long)r) - 1) : 64;
    };
    Ctx c = world_view(b, w);
    if (!FULL) c.nj = 0;  // compile-time: the lean variant carries no joint code
    const uint32_t n = c.n;
    float4 *g_pose = c.pose, *g_vel = c.vel;
    long long t_prev = clock64();
    unsigned long long cyc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#define LAP(slot)                          \
do {                                   \
    long long t_now = clock64();       \
    cyc[slot] += (unsigned long long)(t_now - t_prev); \
    t_prev = t_now;                    \
} while (0)
    if (SMEM) {
        float4* s_pose = dyn;
        float4* s_vel = dyn + n;
        float4* s_x = dyn + 2 * (size_t)n;
        int4* s_meta = reinterpret_cast<int4*>(dyn + 3 * (size_t)n);
        float2* s_minv = reinterpret_cast<float2*>(dyn + 4 * (size_t)n);
        unsigned long long* s_mask = reinterpret_cast<unsigned long long*>(s_minv + n);
        for (uint32_t i = tid; i < n; i += nth) {
            s_pose[i] = __ldcg(&g_pose[i]);
            s_vel[i] = __ldcg(&g_vel[i]);
            s_minv[i] = c.minv[i];
        }
        c.pose = s_pose;
        c.vel = s_vel;
        c.minv = s_minv;
        c.aabb = s_x;
        c.claim0 = reinterpret_cast<unsigned long long*>(s_x);
        c.claim1 = reinterpret_cast<unsigned long long*>(s_x) + n;
        c.prev = s_x;
        c.meta = s_meta;
        c.mask = s_mask;
        __syncthreads();
    }
    LAP(0);

    if (sa == 0u) {
        if (SMEM && b.np_list)  // the narrowphase list's room doubles as the sort array
            batch_broadphase(c, dt, reinterpret_cast<uint32_t*>(dyn + 5 * (size_t)n), b.pair_cap / 2u);
        else
            batch_broadphase(c, dt, nullptr, 0u);
        LAP(1);

        // colouring: same rounds as stage_colour_* in the single-world path
        stage_colour_init(c, tid, nth);
        for (uint32_t k = tid; k <= c.tab_mask; k += nth) c.tab_new[k] = 0ull;
        __syncthreads();
        if (c.warm) {
            stage_colour_warm(c, tid, nth);
            __syncthreads();
        }
        uint32_t round = 0;
        for (; round < (uint32_t)MAX_ROUNDS; ++round) {
            stage_colour_bid(c, round, tid, nth);
            __syncthreads();
            stage_colour_take(c, round, tid, nth);
            __syncthreads();
            if (c.cnt->round_flags[round] == 0u) break;
        }
        if (round == (uint32_t)MAX_ROUNDS && tid == 0) atomicOr(&c.cnt->overflow, 8u);
        stage_colour_record(c, tid, nth);
        __syncthreads();
        LAP(2);
        batch_bucket(c, smem9);
        LAP(3);
        if (c.cnt->n_colours > SFP_SERIAL_COLOUR) serial_sort_pairs(c, c.cnt->n_colours);
        __syncthreads();
    }

    // The substeps read the world's counters — above all the colour offsets, twice per colour
    // and pass — from a copy in shared memory: in global memory every first touch of a line
    // is an L2 round trip on the critical path of its colour, and a world that goes through
    // the queue by the stage starts every stage with a cold L1.  What the substeps change
    // (the head of the struct: contact statistics, overflow flags; the stage times) is written
    // back at the end.
    __shared__ Counters s_cnt;
    Counters* const g_cnt = c.cnt;
    {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(g_cnt);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&s_cnt);
        for (uint32_t i = tid; i < (uint32_t)(sizeof(Counters) / 4u); i += nth) dst[i] = __ldcg(src + i);
    }
    __syncthreads();
    c.cnt = &s_cnt;

    const uint32_t ncol = c.cnt->n_colours;
    // solve order: big colours ascending, then small colours ascending (SFP_TAIL_ITEMS)
    const unsigned long long all = (ncol >= 64u) ? ~0ull : ((1ull << ncol) - 1ull);
    const unsigned long long big = big_colour_mask(c, ncol);
    const unsigned long long small = all & ~big & ~(1ull << SFP_SERIAL_COLOUR);
    const bool serial = ncol > SFP_SERIAL_COLOUR;
    if (tid == 0) {
        uint32_t k = 0;
        for (int pass = 0; pass < 2; ++pass) {
            const unsigned long long m = pass ? small : big;
            for (int col = next_bit(m, -1); col < 64; col = next_bit(m, col)) s_order[k++] = (uint8_t)col;
        }
        s_nord = k;
    }
    __syncthreads();
    const uint32_t nord = s_nord;
    auto issue = [&](uint32_t oi, bool with_lam) {
        if (oi >= nord) return false;
        const uint32_t col = s_order[oi];
        // a colour's joints come before its pairs: item `tid` is then not pair `tid`
        if (!with_lam && c.nj != 0u &&
            c.cnt->joint_col_off[col + 1] != c.cnt->joint_col_off[col])
            return false;
        const uint32_t k = c.cnt->pair_col_off[col] + tid;
        if (k >= c.cnt->pair_col_off[col + 1]) return false;
        __pipeline_memcpy_async(&s_rec[0][tid], &c.ca[k], 16);
        __pipeline_memcpy_async(&s_rec[1][tid], &c.cb[k], 16);
        __pipeline_memcpy_async(&s_rec[2][tid], &c.cp0[k], 16);
        __pipeline_memcpy_async(&s_rec[3][tid], &c.cp1[k], 16);
        __pipeline_memcpy_async(&s_pair[tid], &c.pairs[k], 8);
        if (with_lam) __pipeline_memcpy_async(&s_lam[tid], &c.lam[k], 16);
        __pipeline_commit();
        return true;
    };
    auto take = [&]() {
        PairRec r;
        r.p = s_pair[tid];
        float4 a = s_rec[0][tid], b = s_rec[1][tid], q0 = s_rec[2][tid], q1 = s_rec[3][tid];
        r.ca.x = a.x, r.ca.y = a.y, r.ca.z = a.z, r.ca.w = a.w;
        r.cb.x = b.x, r.cb.y = b.y, r.cb.z = b.z, r.cb.w = b.w;
        r.q0.x = q0.x, r.q0.y = q0.y, r.q0.z = q0.z, r.q0.w = q0.w;
        r.q1.x = q1.x, r.q1.y = q1.y, r.q1.z = q1.z, r.q1.w = q1.w;
        return r;
    };
    for (uint32_t s = (sa ? sa : 1u); s < se; ++s) {
        stage_integrate(c, h, tid, nth);
        __syncthreads();
        LAP(4);
        if (SMEM && b.np_list)
            batch_narrowphase<FULL>(c, reinterpret_cast<uint16_t*>(dyn + 5 * (size_t)n), tid, nth);
        else
            stage_narrowphase<FULL, false, false, true>(c, tid, nth);
        __syncthreads();
        LAP(5);
        {
            bool have = issue(0, false);
            for (uint32_t oi = 0; oi < nord; ++oi) {
                const uint32_t col = s_order[oi];
                if (have) {
                    __pipeline_wait_prior(0);
                    solve_pos_pair(c, c.cnt->pair_col_off[col] + tid, take());
                }
                stage_solve_pos(c, col, inv_h2, tid, nth, have);
                have = issue(oi + 1, false);
                __syncthreads();
            }
        }
        if (serial) {
            if (tid == 0) serial_solve_pos(c, ncol, inv_h2);
            __syncthreads();
        }
        LAP(6);
        stage_derive(c, inv_h, tid, nth);
        __syncthreads();
        LAP(7);
        {
            bool have = issue(0, true);
            for (uint32_t oi = 0; oi < nord; ++oi) {
                const uint32_t col = s_order[oi];
                if (have) {
                    __pipeline_wait_prior(0);
                    float4 l = s_lam[tid];
                    f4 lam;
                    lam.x = l.x, lam.y = l.y, lam.z = l.z, lam.w = l.w;
                    solve_vel_pair(c, c.cnt->pair_col_off[col] + tid, take(), lam, h, rest_threshold);
                }
                stage_solve_vel(c, col, h, rest_threshold, tid, nth, have);
                have = issue(oi + 1, true);
                __syncthreads();
            }
        }
        if (serial) {
            if (tid == 0) serial_solve_vel(c, ncol, h, rest_threshold);
            __syncthreads();
        }
        LAP(8);
    }
    if (SMEM && se > 1u) {
        for (uint32_t i = tid; i < n; i += nth) {
            g_pose[i] = c.pose[i];
            g_vel[i] = c.vel[i];
        }
    }
    LAP(0);
    if (tid == 0) {
        if (sa == 0u)
            for (int k = 0; k < 10; ++k) c.cnt->stage_cycles[k] = cyc[k];
        else
            for (int k = 0; k < 10; ++k) c.cnt->stage_cycles[k] += cyc[k];
    }
#undef LAP
    __syncthreads();
    {
        constexpr uint32_t HEAD = (uint32_t)(offsetof(Counters, pair_col_off) / 4u);
        constexpr uint32_t CYC0 = (uint32_t)(offsetof(Counters, stage_cycles) / 4u);
        constexpr uint32_t CYC1 = (uint32_t)(sizeof(Counters) / 4u);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(&s_cnt);
        uint32_t* dst = reinterpret_cast<uint32_t*>(g_cnt);
        if (tid < HEAD) dst[tid] = src[tid];
        if (CYC0 + tid < CYC1) dst[CYC0 + tid] = src[CYC0 + tid];
    }
}

// Scheduling of a launch's worlds over its CTAs.  A world's frame is a chain of substeps + 1
// stages: stage 0 finds and colours the pairs, stage s >= 1 is substep s.
//   queue == null: CTA i runs world i, start to end (as many CTAs as worlds).
//   otherwise the launch has as many CTAs as are resident at once, and they share two lists:
//   (a) the first n_whole worlds are handed out whole, one at a time, to whichever CTA is free
//       (an atomic counter): the world's state stays in that CTA's shared memory for the frame;
//   (b) the other worlds — as many as there are CTAs — are handed out ONE STAGE at a time,
//       first come first served, through a ring of world indices in global memory: after a
//       stage the CTA writes the state back, notes the world's progress and appends the world
//       to the ring; whoever takes it next reloads the state.  These worlds therefore advance
//       together and end together, whatever the speed of the SMs that serve them, so the
//       launch has no ragged last wave: CTAs that leave (a) early simply take more of (b).
// A CTA only ever waits for a ring slot that a RUNNING CTA is about to fill (see the comments
// at the two waits), so the scheme needs the CTAs that hold tasks to be running — which they
// are — and not the whole grid to be resident.
struct BatchQueue {
    uint32_t next_whole;  // (a): next world to hand out
    uint32_t head;        // (b): tasks handed out
    uint32_t tail;        // (b): ring slots filled
    uint32_t exits;       // CTAs that found nothing left; the last one zeroes the queue
};

template <bool SMEM, bool FULL = false>
__global__ void __launch_bounds__(BTPB, FULL ? 2 : SFP_BCTAS)
    k_batch_frame(BatchCtx b, float dt, uint32_t substeps, float rest_threshold) {
    // thread 0 finds the next task; nothing but the loop itself is live across the stages
    __shared__ uint32_t s_seg[4];  // world, first stage, end stage (0: no task left), staged?
    for (uint32_t it = 0;; ++it) {
        if (threadIdx.x == 0) {
            const uint32_t per = substeps + 1u;
            uint32_t w = 0u, sa = 0u, se = 0u, staged = 0u;
            BatchQueue* q = b.queue;
            if (!q) {
                if (it == 0u) w = blockIdx.x, se = per;
            } else {
                const uint32_t n_staged = b.n_launch - b.n_whole;
                uint32_t i = (it == 0u || s_seg[3] == 0u) ? atomicAdd(&q->next_whole, 1u) : b.n_whole;
                if (i < b.n_whole) {
                    w = i, se = per;
                } else {
                    const uint32_t t = atomicAdd(&q->head, 1u);
                    staged = 1u;
                    if (t < n_staged) {
                        w = b.n_whole + t, se = 1u;  // the world's stage 0
                    } else if (t < n_staged * per) {
                        // Task t continues the world in ring slot t - n_staged.  That slot is
                        // filled by the (t - n_staged + 1)-th stage that ends with stages to
                        // go; tasks 0..t-1 are all handed out, and at most n_staged - 1 of them
                        // can be the last stage of their world while task t exists, so that
                        // stage is among them: it has ended or is running.
                        volatile uint32_t* slot = b.ring + (t - n_staged) % n_staged;
                        uint32_t v;
                        while ((v = *slot) == 0u) __nanosleep(100);
                        *slot = 0u;
                        __threadfence();
                        w = v - 1u;
                        sa = *(volatile uint32_t*)(b.progress + b.world_base + w);
                        se = sa + 1u;
                    }
                }
                if (se == 0u && atomicAdd(&q->exits, 1u) == gridDim.x - 1u) {
                    // every CTA has taken its last look at the queue: leave it ready for the
                    // next launch (the ring and the progress notes are already back to zero)
                    q->next_whole = 0u, q->head = 0u, q->tail = 0u, q->exits = 0u;
                }
            }
            s_seg[0] = w + b.world_base;
            s_seg[1] = sa;
            s_seg[2] = se;
            s_seg[3] = staged;
        }
        __syncthreads();
        const uint32_t w = s_seg[0], sa = s_seg[1], se = s_seg[2];
#ifdef SFP_QUEUE_PROBE
        if (threadIdx.x == 0 && (it == 0u || se == 0u)) {
            unsigned long long t;
            unsigned smid;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            reinterpret_cast<unsigned long long*>(b.scratch)[2 * blockIdx.x + (se == 0u ? 1 : 0)] = t;
            reinterpret_cast<uint32_t*>(b.scratch)[4 * 8192 + blockIdx.x] = smid;
        }
#endif
        if (se == 0u) break;
        batch_world_stages<SMEM, FULL>(b, w, sa, se, dt, substeps, rest_threshold);
        // shared memory is reused by the next task; a world that goes back to the ring is
        // published after everything this CTA wrote for it
        __syncthreads();
        if (threadIdx.x == 0 && s_seg[3] != 0u) {
            const uint32_t w_ = s_seg[0], se_ = s_seg[2];
            if (se_ != substeps + 1u) {
                const uint32_t n_staged = b.n_launch - b.n_whole;
                *(volatile uint32_t*)(b.progress + w_) = se_;
                __threadfence();
                // The slot's previous occupant (one lap of the ring earlier) belongs to task
                // `pos`, which has been handed out — pos + 1 stages have ended, so at least as
                // many tasks were taken — and whose CTA clears the slot as soon as it reads it.
                const uint32_t pos = atomicAdd(&b.queue->tail, 1u);
                volatile uint32_t* slot = b.ring + pos % n_staged;
                while (*slot != 0u) __nanosleep(100);
                *slot = (w_ - b.world_base) + 1u;
            } else {
                *(volatile uint32_t*)(b.progress + w_) = 0u;  // ready for the next frame
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_batch_pack(const float4* __restrict__ pose,
                                                    const float4* __restrict__ vel,
                                                    sfp::StateRec* __restrict__ out, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x) {
        float4 p = pose[i], v = vel[i];
        f4 fp, fv;
        fp.x = p.x, fp.y = p.y, fp.z = p.z, fp.w = p.w;
        fv.x = v.x, fv.y = v.y, fv.z = v.z, fv.w = v.w;
        sfp::StateRec r = sfp::state_pack(fp, fv);
        float4* o = reinterpret_cast<float4*>(out + i);
        o[0] = make_float4(r.x, r.y, r.c, r.s);
        o[1] = make_float4(r.vx, r.vy, r.omega, r.angle);
    }
}

// scene-graph states -> pose / vel (statics keep zero velocity; a zero rotor is replaced by
// sincos_det(angle) inside state_unpack)
__global__ void __launch_bounds__(256) k_batch_unpack(const sfp::StateRec* __restrict__ in,
                                                      const float2* __restrict__ minv,
                                                      float4* __restrict__ pose,
                                                      float4* __restrict__ vel, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x) {
        const float4* q = reinterpret_cast<const float4*>(in + i);
        float4 a = q[0], b = q[1];
        sfp::StateRec r;
        r.x = a.x, r.y = a.y, r.c = a.z, r.s = a.w, r.vx = b.x, r.vy = b.y, r.omega = b.z, r.angle = b.w;
        float2 m = minv[i];
        f2 mi;
        mi.x = m.x, mi.y = m.y;
        f4 p, v;
        sfp::state_unpack(r, mi, &p, &v);
        pose[i] = make_float4(p.x, p.y, p.z, p.w);
        vel[i] = make_float4(v.x, v.y, v.z, v.w);
    }
}

// true if p is page-locked host memory the device can DMA from / to directly
bool batch_host_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

}  // namespace

// The dynamic shared-memory limit of k_batch_frame is a per-device property of the FUNCTION,
// shared by every batch in the process: only ever raise it (a later, smaller batch must not
// pull it below what an earlier, larger one launches with).
static cudaError_t raise_batch_smem_limit(int device, size_t bytes) {
    static size_t current[64] = {0};
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    if (bytes <= current[device]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k_batch_frame<true, false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_batch_frame<true, true>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) current[device] = bytes;
    return e;
}

#define BCU(call)                                                        \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) {                                         \
            b->err = std::string(#call) + ": " + cudaGetErrorString(e_); \
            b->sticky = SFP_E_CUDA;                                      \
            return SFP_E_CUDA;                                           \
        }                                                                \
    } while (0)

struct SfpBatch {
    SfpParams p;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false, uploaded = false, stepped = false;
    uint32_t n_worlds = 0, bpw = 0, pair_cap = 0;
    size_t smem_bytes = 0;  // 0: the world's bodies stay in global memory (too big for one SM)
    BatchCtx ctx;
    std::vector<void*> allocs;
    std::vector<sfp::f2> h_minv;
    uint64_t frames = 0, body_substeps = 0;
    unsigned long long* d_tab[2] = {nullptr, nullptr};
    uint32_t colour_age = 0;
    sfp::StateRec* d_state = nullptr;   // device AoS staging for downloads
    SfpBodyState* pin_state = nullptr;  // pinned host staging, pin_states entries
    size_t pin_states = 0;
    uint32_t n_dynamic_total = 0;
    std::vector<uint32_t> n_dynamic;
    cudaStream_t aux[3] = {nullptr, nullptr, nullptr};  // sfp_batch_step_host's pipeline
    cudaEvent_t aux_done[3] = {nullptr, nullptr, nullptr};
    sfp::PolyTable polys;                // general-polygon table shared by all worlds
    float4* d_polytab = nullptr;
    bool has_polygons = false;
    uint32_t joint_cap = 0;              // joint slots per world
    std::vector<uint32_t> n_joints;      // per world
    std::vector<void*> joint_allocs;
    uint32_t resident[2] = {0u, 0u};     // frame-kernel CTAs resident at once (lean, full)
    uint32_t* d_ring = nullptr;          // [n_worlds] rings of the launches (k_batch_frame)
    BatchQueue* d_queue = nullptr;       // [4] one per concurrent launch
    std::string err;
    int32_t sticky = SFP_OK;
};

namespace {
template <typename T>
int balloc(SfpBatch* b, T** out, size_t count) {
    void* p = nullptr;
    size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    BCU(cudaMalloc(&p, bytes));
    BCU(cudaMemset(p, 0, bytes));
    b->allocs.push_back(p);
    *out = reinterpret_cast<T*>(p);
    return SFP_OK;
}
#define BTRY(x)                        \
    do {                               \
        int rc_ = (x);                 \
        if (rc_ != SFP_OK) return rc_; \
    } while (0)
}  // namespace

extern "C" {

int32_t sfp_batch_create(const SfpParams* params, uint32_t n_worlds, uint32_t bodies_per_world,
                         SfpBatch** out) {
    if (!out) return SFP_E_INVALID;
    *out = nullptr;
    SfpParams p;
    if (params) p = *params;
    else sfp::default_params(&p);
    if (!sfp::params_valid(p) || p.solver_order != SFP_ORDER_COLOUR) return SFP_E_INVALID;
    if (n_worlds == 0 || bodies_per_world == 0 || bodies_per_world > 4096) return SFP_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        fprintf(stderr, "libsfp_b200: no CUDA device — this library has no CPU fallback\n");
        return SFP_E_CUDA;
    }
    if (p.device < 0 || p.device >= ndev) return SFP_E_INVALID;
    SfpBatch* b = new (std::nothrow) SfpBatch();
    if (!b) return SFP_E_NOMEM;
    b->p = p;
    b->device = p.device;
    b->n_worlds = n_worlds;
    b->bpw = bodies_per_world;
    b->pair_cap = p.max_pairs ? p.max_pairs : 12u * bodies_per_world + 256u;
    std::memset(&b->ctx, 0, sizeof(b->ctx));
    auto die = [&](int rc) {
        sfp_batch_destroy(b);
        return (int32_t)rc;
    };
    if (cudaSetDevice(b->device) != cudaSuccess) return die(SFP_E_CUDA);
    if (cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking) != cudaSuccess)
        return die(SFP_E_CUDA);
    if (cudaEventCreate(&b->ev0) != cudaSuccess || cudaEventCreate(&b->ev1) != cudaSuccess)
        return die(SFP_E_CUDA);
    const size_t nb = (size_t)n_worlds * bodies_per_world, np = (size_t)n_worlds * b->pair_cap;
    Ctx& c = b->ctx.base;
    int rc = SFP_OK;
#define A(ptr, count) \
    if (rc == SFP_OK) rc = balloc(b, &(ptr), (count))
    A(c.pose, nb);
    A(c.prev, nb);
    A(c.vel, nb);
    float4 *shape = nullptr, *mat = nullptr;
    float2* minv = nullptr;
    A(shape, nb);
    A(mat, nb);
    A(minv, nb);
    c.shape = shape;
    c.mat = mat;
    c.minv = minv;
    A(c.aabb, nb);
    A(c.meta, nb);
    A(c.mask, nb);
    A(c.claim0, nb);
    A(c.claim1, nb);
    A(c.pairs_raw, np);
    A(c.pairs, np);
    A(c.pcol, np);
    A(c.ca, np);
    A(c.cb, np);
    A(c.cp0, np);
    A(c.cp1, np);
    A(c.cs, np);
    A(c.lam, np);
    A(c.np_order, np);
    A(b->ctx.counters, n_worlds);
    A(b->ctx.scratch, nb);
    A(b->ctx.progress, n_worlds);
    A(b->d_ring, n_worlds);
    A(b->d_queue, 4);
    uint32_t slots = 256;
    while (slots < 2u * b->pair_cap) slots <<= 1;
    A(b->d_tab[0], (size_t)n_worlds * slots);
    A(b->d_tab[1], (size_t)n_worlds * slots);
    b->ctx.tab_slots = slots;
    c.tab_mask = slots - 1u;
    A(b->d_state, nb);
    b->pin_states = std::min<size_t>(nb, (size_t)2 << 20);  // <= 64 MiB pinned
    if (rc == SFP_OK && cudaMallocHost(&b->pin_state, sizeof(SfpBodyState) * b->pin_states) != cudaSuccess)
        rc = SFP_E_CUDA;
#undef A
    if (rc != SFP_OK) return die(rc);
    // joints are not part of the batched path: give the unused pointers a valid target
    c.jbod_in = reinterpret_cast<const uint2*>(c.pairs_raw);
    c.jcol = c.pcol;
    c.gx = p.gravity_x;
    c.gy = p.gravity_y;
    c.inv_cell = 1.0f / p.cell_size;
    c.aabb_margin = p.aabb_margin;
    c.vel_expand = p.vel_expand;
    c.contact_margin = p.contact_margin;
    b->ctx.bpw = bodies_per_world;
    b->ctx.pair_cap = b->pair_cap;
    b->ctx.n_worlds = n_worlds;
    {
        size_t need = (size_t)bodies_per_world * 80u;
        // + the narrowphase list (batch_narrowphase): one u16 per candidate pair
        const size_t with_list = need + (((size_t)b->pair_cap * 2u + 15u) & ~(size_t)15u);
        if (b->pair_cap <= 65535u && with_list <= 100u * 1024u) {
            need = with_list;
            b->ctx.np_list = 1u;
        }
        if (need <= 200u * 1024u) {
            if (raise_batch_smem_limit(b->device, need) != cudaSuccess) return die(SFP_E_CUDA);
            b->smem_bytes = need;
        }
    }
    *out = b;
    return SFP_OK;
}

int32_t sfp_batch_destroy(SfpBatch* b) {
    if (!b) return SFP_OK;
    cudaSetDevice(b->device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    for (void* p : b->allocs) cudaFree(p);
    for (void* p : b->joint_allocs) cudaFree(p);
    if (b->d_polytab) cudaFree(b->d_polytab);
    for (int k = 0; k < 3; ++k) {
        if (b->aux_done[k]) cudaEventDestroy(b->aux_done[k]);
        if (b->aux[k]) cudaStreamDestroy(b->aux[k]);
    }
    if (b->pin_state) cudaFreeHost(b->pin_state);
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
    return SFP_OK;
}

int32_t sfp_batch_upload(SfpBatch* b, const SfpBodyDesc* bodies) {
    if (!b || !bodies) return SFP_E_INVALID;
    const size_t nb = (size_t)b->n_worlds * b->bpw;
    for (size_t i = 0; i < nb; ++i)
        if (!sfp::desc_valid(bodies[i], &b->polys)) {
            b->err = "invalid body desc at index " + std::to_string(i);
            return SFP_E_INVALID;
        }
    BCU(cudaSetDevice(b->device));
    std::vector<sfp::f4> pose(nb), vel(nb), shape(nb), mat(nb);
    b->h_minv.resize(nb);
    b->n_dynamic.assign(b->n_worlds, 0);
    b->has_polygons = false;
    for (size_t i = 0; i < nb; ++i) {
        sfp::body_from_desc(bodies[i], &pose[i], &vel[i], &b->h_minv[i], &shape[i], &mat[i], &b->polys);
        b->has_polygons |= (bodies[i].shape & 0xffu) == SFP_SHAPE_POLYGON;
        if (b->h_minv[i].x > 0.0f || b->h_minv[i].y > 0.0f) b->n_dynamic[i / b->bpw] += 1;
        if (!sfp::body_fits_grid(pose[i], shape[i], b->p)) {
            b->err = "body " + std::to_string(i) + " spans more than 16384 broadphase cells (or 4096 along one axis) at this cell_size: split it or raise cell_size";
            return SFP_E_INVALID;
        }
    }
    Ctx& c = b->ctx.base;
    BCU(cudaStreamSynchronize(b->stream));
    BCU(cudaMemcpy(c.pose, pose.data(), sizeof(sfp::f4) * nb, cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(c.prev, pose.data(), sizeof(sfp::f4) * nb, cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(c.vel, vel.data(), sizeof(sfp::f4) * nb, cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(const_cast<float4*>(c.shape), shape.data(), sizeof(sfp::f4) * nb,
                   cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(const_cast<float4*>(c.mat), mat.data(), sizeof(sfp::f4) * nb,
                   cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(const_cast<float2*>(c.minv), b->h_minv.data(), sizeof(sfp::f2) * nb,
                   cudaMemcpyHostToDevice));
    BCU(cudaMemset(b->ctx.counters, 0, sizeof(Counters) * b->n_worlds));
    b->colour_age = 0;
    b->uploaded = true;
    b->stepped = false;
    return SFP_OK;
}

int32_t sfp_batch_set_polygons(SfpBatch* b, const SfpPolygon* polys, uint32_t n) {
    if (!b || (n && !polys)) return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    sfp::PolyTable t;
    long bad = sfp::build_poly_table(polys, n, &t);
    if (bad >= 0) {
        b->err = "invalid polygon at index " + std::to_string(bad) +
                 " (3..8 vertices, counter-clockwise, strictly convex, centroid at the origin)";
        return SFP_E_INVALID;
    }
    BCU(cudaSetDevice(b->device));
    BCU(cudaStreamSynchronize(b->stream));
    if (b->d_polytab) cudaFree(b->d_polytab);
    b->d_polytab = nullptr;
    if (n) {
        BCU(cudaMalloc(&b->d_polytab, sizeof(float4) * 8u * (size_t)n));
        BCU(cudaMemcpy(b->d_polytab, t.recs.data(), sizeof(float4) * 8u * (size_t)n,
                       cudaMemcpyHostToDevice));
    }
    b->ctx.base.polytab = b->d_polytab;
    b->polys = std::move(t);
    return SFP_OK;
}

int32_t sfp_batch_set_joints(SfpBatch* b, const SfpJointDesc* joints, uint32_t joints_per_world,
                             const uint32_t* counts) {
    if (!b || (joints_per_world && !joints) || joints_per_world > 65536u) return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    const uint32_t J = joints_per_world;
    std::vector<uint32_t> nj(b->n_worlds, J);
    for (uint32_t w = 0; w < b->n_worlds && J; ++w) {
        if (counts) nj[w] = counts[w];
        if (nj[w] > J) return SFP_E_INVALID;
        for (uint32_t k = 0; k < nj[w]; ++k)
            if (!sfp::joint_valid(joints[(size_t)w * J + k], b->bpw)) {
                b->err = "invalid joint desc, world " + std::to_string(w) + " joint " + std::to_string(k);
                return SFP_E_INVALID;
            }
    }
    BCU(cudaSetDevice(b->device));
    BCU(cudaStreamSynchronize(b->stream));
    for (void* p : b->joint_allocs) cudaFree(p);
    b->joint_allocs.clear();
    Ctx& c = b->ctx.base;
    b->joint_cap = J;
    b->ctx.joint_cap = J;
    b->ctx.n_joints = nullptr;
    b->n_joints = nj;
    b->colour_age = 0;
    if (J == 0) {
        // as after creation: unused pointers keep a valid target
        c.jbod_in = reinterpret_cast<const uint2*>(c.pairs_raw);
        c.jcol = c.pcol;
        std::fill(b->n_joints.begin(), b->n_joints.end(), 0u);
        return SFP_OK;
    }
    const size_t tot = (size_t)b->n_worlds * J;
    std::vector<uint2> jb(tot, make_uint2(0u, 0u));
    std::vector<float4> ja(tot, make_float4(0, 0, 0, 0)), jp(tot, make_float4(0, 0, 0, 0));
    for (uint32_t w = 0; w < b->n_worlds; ++w)
        for (uint32_t k = 0; k < nj[w]; ++k) {
            const SfpJointDesc& j = joints[(size_t)w * J + k];
            const size_t i = (size_t)w * J + k;
            jb[i] = make_uint2(j.body_a, j.body_b);
            ja[i] = make_float4(j.ax, j.ay, j.bx, j.by);
            jp[i] = make_float4(j.rest, j.compliance, sfp::u2f(j.kind), 0.0f);
        }
    auto jalloc = [&](void** out, size_t bytes) {
        if (cudaMalloc(out, bytes) != cudaSuccess) return false;
        b->joint_allocs.push_back(*out);
        return cudaMemset(*out, 0, bytes) == cudaSuccess;
    };
    void *d_jb_in = nullptr, *d_ja_in = nullptr, *d_jp_in = nullptr, *d_jb = nullptr, *d_ja = nullptr,
         *d_jp = nullptr, *d_jcol = nullptr, *d_nj = nullptr;
    if (!jalloc(&d_jb_in, sizeof(uint2) * tot) || !jalloc(&d_ja_in, sizeof(float4) * tot) ||
        !jalloc(&d_jp_in, sizeof(float4) * tot) || !jalloc(&d_jb, sizeof(uint2) * tot) ||
        !jalloc(&d_ja, sizeof(float4) * tot) || !jalloc(&d_jp, sizeof(float4) * tot) ||
        !jalloc(&d_jcol, sizeof(uint32_t) * tot) || !jalloc(&d_nj, sizeof(uint32_t) * b->n_worlds)) {
        b->err = "cudaMalloc (joints)";
        b->sticky = SFP_E_CUDA;
        return SFP_E_CUDA;
    }
    BCU(cudaMemcpy(d_jb_in, jb.data(), sizeof(uint2) * tot, cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(d_ja_in, ja.data(), sizeof(float4) * tot, cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(d_jp_in, jp.data(), sizeof(float4) * tot, cudaMemcpyHostToDevice));
    BCU(cudaMemcpy(d_nj, nj.data(), sizeof(uint32_t) * b->n_worlds, cudaMemcpyHostToDevice));
    c.jbod_in = static_cast<const uint2*>(d_jb_in);
    c.janc_in = static_cast<const float4*>(d_ja_in);
    c.jpar_in = static_cast<const float4*>(d_jp_in);
    c.jbod = static_cast<uint2*>(d_jb);
    c.janc = static_cast<float4*>(d_ja);
    c.jpar = static_cast<float4*>(d_jp);
    c.jcol = static_cast<uint32_t*>(d_jcol);
    b->ctx.n_joints = static_cast<const uint32_t*>(d_nj);
    return SFP_OK;
}

namespace {
// CTAs of the frame kernel that are resident at once on this batch's device
uint32_t batch_resident_ctas(SfpBatch* b, bool full) {
    uint32_t& cached = b->resident[full ? 1 : 0];
    if (cached) return cached;
    int per_sm = 0, sms = 0;
    cudaError_t e;
    if (b->smem_bytes)
        e = full ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_batch_frame<true, true>, BTPB, b->smem_bytes)
                 : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_batch_frame<true, false>, BTPB, b->smem_bytes);
    else
        e = full ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_batch_frame<false, true>, BTPB, 0)
                 : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_batch_frame<false, false>, BTPB, 0);
    if (e != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, b->device) != cudaSuccess ||
        per_sm <= 0 || sms <= 0) {
        cudaGetLastError();
        return 0u;  // unknown: one CTA per world
    }
    cached = (uint32_t)per_sm * (uint32_t)sms;
    return cached;
}

// worlds [w0, w0 + nw) of the batch on `st`, as launch number `slot` of those that may run side by
// side.  Up to as many worlds as CTAs are resident: one CTA per world.  More: exactly that many
// CTAs and the two lists of k_batch_frame.  `closing`: this launch is the last of a pipeline of
// slices and starts while earlier slices still hold most of the SMs — then ALL its worlds go by
// the stage, so that the CTAs that get an SM late still find work to share and the whole
// pipeline, not just this slice, ends evenly.
void launch_batch_frame(SfpBatch* b, cudaStream_t st, uint32_t w0, uint32_t nw, float dt,
                        uint32_t substeps, float rt, uint32_t slot = 0u, bool closing = false) {
    BatchCtx ctx = b->ctx;
    ctx.world_base = w0;
    ctx.n_launch = nw;
    ctx.n_whole = nw;
    ctx.queue = nullptr;
    ctx.ring = b->d_ring + w0;
    const bool full = b->has_polygons || b->joint_cap != 0u;
    uint32_t grid = nw;
    const uint32_t resident = batch_resident_ctas(b, full);
    if (resident && slot < 4u && !(b->p.flags & SFP_FLAG_BATCH_WHOLE_WORLDS)) {
        if (nw > resident) {
            grid = resident;
            ctx.queue = b->d_queue + slot;
            // as many worlds by the stage as there are CTAs: fewer would leave CTAs without
            // a world at the end (and the ring assumes at least one slot per CTA: two tasks
            // must never wait for the same slot); more cost ~19 % per world for the reloads
            // and for contact records that no longer stay in the L2 from substep to substep
            // (2048 worlds: 296 by the stage 4.34 ms, 568: 4.44 ms, all: 5.00 ms)
            ctx.n_whole = nw - grid;
        }
        if (closing && nw <= 2u * resident) {
            grid = std::min(nw, resident);
            ctx.queue = b->d_queue + slot;
            ctx.n_whole = 0u;
        }
    }
    if (ctx.queue && nw - ctx.n_whole < grid) grid = nw - ctx.n_whole;  // a ring slot per CTA
    if (b->smem_bytes) {
        if (full) k_batch_frame<true, true><<<grid, BTPB, b->smem_bytes, st>>>(ctx, dt, substeps, rt);
        else k_batch_frame<true, false><<<grid, BTPB, b->smem_bytes, st>>>(ctx, dt, substeps, rt);
    } else {
        if (full) k_batch_frame<false, true><<<grid, BTPB, 0, st>>>(ctx, dt, substeps, rt);
        else k_batch_frame<false, false><<<grid, BTPB, 0, st>>>(ctx, dt, substeps, rt);
    }
}
}  // namespace

int32_t sfp_batch_step(SfpBatch* b, float dt, uint32_t substeps) {
    if (!b || !(dt > 0.0f) || substeps == 0 || substeps > 1024 || !b->uploaded)
        return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    BCU(cudaSetDevice(b->device));
    const float h = dt / (float)substeps;
    const float g = sqrtf(b->p.gravity_x * b->p.gravity_x + b->p.gravity_y * b->p.gravity_y);
    b->ctx.base.warm = (b->colour_age % SFP_RECOLOUR_PERIOD) != 0u ? 1u : 0u;
    b->ctx.base.tab_old = b->d_tab[(b->colour_age + 1u) & 1u];
    b->ctx.base.tab_new = b->d_tab[b->colour_age & 1u];
    b->colour_age += 1;
    BCU(cudaEventRecord(b->ev0, b->stream));
    launch_batch_frame(b, b->stream, 0u, b->n_worlds, dt, substeps, 2.0f * g * h);
    BCU(cudaGetLastError());
    BCU(cudaEventRecord(b->ev1, b->stream));
    b->timed = true;
    b->stepped = true;
    b->frames += 1;
    b->body_substeps += (uint64_t)b->n_worlds * b->bpw * substeps;
    return SFP_OK;
}

#ifdef SFP_QUEUE_PROBE
__attribute__((visibility("default"))) int32_t sfp_debug_batch_scratch(SfpBatch* b, void* out, size_t bytes) {
    cudaSetDevice(b->device);
    cudaStreamSynchronize(b->stream);
    return cudaMemcpy(out, b->ctx.scratch, bytes, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -1;
}
#endif

int32_t sfp_batch_sync(SfpBatch* b) {
    if (!b) return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    BCU(cudaSetDevice(b->device));
    BCU(cudaStreamSynchronize(b->stream));
    return SFP_OK;
}

int32_t sfp_batch_download(SfpBatch* b, SfpBodyState* out) {
    if (!b || !out) return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    BCU(cudaSetDevice(b->device));
    const size_t nb = (size_t)b->n_worlds * b->bpw;
    // packed on the device into AoS states (incl. the angle), then one copy per 64 MiB chunk
    // through the pinned staging buffer
    k_batch_pack<<<(unsigned)std::min<size_t>((nb + 255) / 256, 148 * 8), 256, 0, b->stream>>>(
        b->ctx.base.pose, b->ctx.base.vel, b->d_state, nb);
    BCU(cudaGetLastError());
    if (batch_host_pinned(out)) {  // page-locked destination (sfp_alloc_host_states): one DMA
        BCU(cudaMemcpyAsync(out, b->d_state, sizeof(SfpBodyState) * nb, cudaMemcpyDeviceToHost,
                            b->stream));
        BCU(cudaStreamSynchronize(b->stream));
        return SFP_OK;
    }
    const size_t chunk = b->pin_states;
    for (size_t o = 0; o < nb; o += chunk) {
        size_t m = std::min(chunk, nb - o);
        BCU(cudaMemcpyAsync(b->pin_state, b->d_state + o, sizeof(SfpBodyState) * m,
                            cudaMemcpyDeviceToHost, b->stream));
        BCU(cudaStreamSynchronize(b->stream));
        std::memcpy(out + o, b->pin_state, sizeof(SfpBodyState) * m);
    }
    return SFP_OK;
}

int32_t sfp_batch_step_host(SfpBatch* b, const SfpBodyState* in, SfpBodyState* out, float dt,
                            uint32_t substeps) {
    if (!b || !in || !out || !b->uploaded) return SFP_E_INVALID;
    if (!(dt > 0.0f) || substeps == 0 || substeps > 1024) return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    BCU(cudaSetDevice(b->device));
    const size_t nb = (size_t)b->n_worlds * b->bpw;
    if (batch_host_pinned(in) && batch_host_pinned(out) && b->n_worlds >= 64u) {
        // Both arrays page-locked: the worlds go through in four slices on three extra streams,
        // so that a slice's upload and download overlap the other slices' frames (the worlds
        // are independent; each slice is the same kernel over its own world range).
        for (int k = 0; k < 3; ++k) {
            if (!b->aux[k]) BCU(cudaStreamCreateWithFlags(&b->aux[k], cudaStreamNonBlocking));
            if (!b->aux_done[k]) BCU(cudaEventCreateWithFlags(&b->aux_done[k], cudaEventDisableTiming));
        }
        const float h = dt / (float)substeps;
        const float g = sqrtf(b->p.gravity_x * b->p.gravity_x + b->p.gravity_y * b->p.gravity_y);
        b->ctx.base.warm = (b->colour_age % SFP_RECOLOUR_PERIOD) != 0u ? 1u : 0u;
        b->ctx.base.tab_old = b->d_tab[(b->colour_age + 1u) & 1u];
        b->ctx.base.tab_new = b->d_tab[b->colour_age & 1u];
        b->colour_age += 1;
        BCU(cudaEventRecord(b->ev0, b->stream));
        for (int k = 0; k < 3; ++k) BCU(cudaStreamWaitEvent(b->aux[k], b->ev0, 0));
        // Four slices on four streams, the outer two small: what stays exposed is the first
        // slice's upload and the last one's download, and every launch has a fixed cost (the
        // worlds that go through the queue by the stage), so the bulk goes in two big slices.
        const uint32_t slices = 4;
        const uint32_t edge = std::min(b->n_worlds / 4u, std::max(256u, b->n_worlds / 16u));
        const uint32_t mid = (b->n_worlds - 2u * edge + 1u) / 2u;
        const uint32_t first[5] = {0u, edge, edge + mid, b->n_worlds - edge, b->n_worlds};
        for (uint32_t k = 0; k < slices; ++k) {
            const uint32_t w0 = first[k], nw = first[k + 1] - first[k];
            if (nw == 0u) continue;
            const size_t o = (size_t)w0 * b->bpw, m = (size_t)nw * b->bpw;
            cudaStream_t st = (k % 4u == 0u) ? b->stream : b->aux[k % 4u - 1u];
            // 64-thread blocks of <= 32 registers: they fit on an SM beside two CTAs of the frame
            // kernel (which leave 4096 registers free at 120 registers a thread).  With 256-thread blocks a slice's pack
            // kernel waited until the NEXT slice's frame kernel had ended, and so did its
            // download: the downloads were not overlapped at all (8192 worlds: 19.5 ms a frame
            // against 17.2 without the copies).
            const unsigned blocks = (unsigned)std::min<size_t>((m + 63) / 64, 148 * 16);
            BCU(cudaMemcpyAsync(b->d_state + o, in + o, sizeof(SfpBodyState) * m,
                                cudaMemcpyHostToDevice, st));
            k_batch_unpack<<<blocks, 64, 0, st>>>(b->d_state + o, b->ctx.base.minv + o,
                                                   b->ctx.base.pose + o, b->ctx.base.vel + o, m);
            launch_batch_frame(b, st, w0, nw, dt, substeps, 2.0f * g * h, k % 4u, w0 + nw == b->n_worlds);
            k_batch_pack<<<blocks, 64, 0, st>>>(b->ctx.base.pose + o, b->ctx.base.vel + o,
                                                 b->d_state + o, m);
            BCU(cudaGetLastError());
            BCU(cudaMemcpyAsync(out + o, b->d_state + o, sizeof(SfpBodyState) * m,
                                cudaMemcpyDeviceToHost, st));
            if (k % 4u != 0u) BCU(cudaEventRecord(b->aux_done[k % 4u - 1u], st));
        }
        for (int k = 0; k < 3; ++k) BCU(cudaStreamWaitEvent(b->stream, b->aux_done[k], 0));
        BCU(cudaEventRecord(b->ev1, b->stream));
        b->timed = true;
        b->stepped = true;
        b->frames += 1;
        b->body_substeps += (uint64_t)b->n_worlds * b->bpw * substeps;
        BCU(cudaStreamSynchronize(b->stream));
        return SFP_OK;
    }
    if (batch_host_pinned(in)) {
        BCU(cudaMemcpyAsync(b->d_state, in, sizeof(SfpBodyState) * nb, cudaMemcpyHostToDevice,
                            b->stream));
    } else {
        const size_t chunk = b->pin_states;
        for (size_t o = 0; o < nb; o += chunk) {
            size_t m = std::min(chunk, nb - o);
            BCU(cudaStreamSynchronize(b->stream));  // the staging buffer is reused
            std::memcpy(b->pin_state, in + o, sizeof(SfpBodyState) * m);
            BCU(cudaMemcpyAsync(b->d_state + o, b->pin_state, sizeof(SfpBodyState) * m,
                                cudaMemcpyHostToDevice, b->stream));
        }
    }
    k_batch_unpack<<<(unsigned)std::min<size_t>((nb + 255) / 256, 148 * 8), 256, 0, b->stream>>>(
        b->d_state, b->ctx.base.minv, b->ctx.base.pose, b->ctx.base.vel, nb);
    BCU(cudaGetLastError());
    int32_t rc = sfp_batch_step(b, dt, substeps);
    if (rc != SFP_OK) return rc;
    return sfp_batch_download(b, out);
}

int32_t sfp_batch_stats(SfpBatch* b, SfpStats* out) {
    if (!b || !out) return SFP_E_INVALID;
    if (b->sticky != SFP_OK) return b->sticky;
    BCU(cudaSetDevice(b->device));
    std::vector<Counters> k(b->n_worlds);
    BCU(cudaStreamSynchronize(b->stream));
    BCU(cudaMemcpy(k.data(), b->ctx.counters, sizeof(Counters) * b->n_worlds,
                   cudaMemcpyDeviceToHost));
    int32_t rc = SFP_OK;
    for (uint32_t w = 0; w < b->n_worlds; ++w) {
        SfpStats s;
        std::memset(&s, 0, sizeof(s));
        s.frames = b->frames;
        s.body_substeps = b->body_substeps / b->n_worlds;
        s.n_bodies = b->bpw;
        s.n_dynamic = b->n_dynamic.empty() ? 0u : b->n_dynamic[w];
        s.n_joints = b->n_joints.empty() ? 0u : b->n_joints[w];
        s.n_pairs = b->stepped ? k[w].n_pairs : 0u;
        s.n_contacts = k[w].n_contacts;
        s.n_manifolds = k[w].n_manifolds;
        s.n_colours = k[w].n_colours;
        s.colour_rounds = k[w].colour_rounds;
        s.overflow = k[w].overflow;
        s.max_penetration = sfp::u2f(k[w].maxpen_bits);
        s.launches = 1;
        out[w] = s;
        if (k[w].overflow) {
            b->err = "world " + std::to_string(w) + ": capacity exceeded (bits " +
                     std::to_string(k[w].overflow) + ")";
            rc = SFP_E_CAPACITY;
        }
    }
    return rc;
}

/* development aid (not part of include/sfp.h): SM-clock cycles per stage of world `w` */
__attribute__((visibility("default"))) int32_t sfp_batch_debug_stage_cycles(SfpBatch* b, uint32_t w,
                                                                          unsigned long long out[10]) {
    if (!b || !out || w >= b->n_worlds) return SFP_E_INVALID;
    BCU(cudaSetDevice(b->device));
    BCU(cudaStreamSynchronize(b->stream));
    Counters k;
    BCU(cudaMemcpy(&k, b->ctx.counters + w, sizeof(k), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 10; ++i) out[i] = k.stage_cycles[i];
    return SFP_OK;
}

const char* sfp_batch_last_error(SfpBatch* b) { return b ? b->err.c_str() : "null batch"; }

int32_t sfp_batch_last_step_ms(SfpBatch* b, float* ms) {
    if (!b || !ms) return SFP_E_INVALID;
    *ms = 0.0f;
    if (!b->timed) return SFP_OK;
    BCU(cudaSetDevice(b->device));
    BCU(cudaEventSynchronize(b->ev1));
    BCU(cudaEventElapsedTime(ms, b->ev0, b->ev1));
    return SFP_OK;
}


/* ------------------------------------------------------------------------------------------
 * sfp_step_small_noalloc — one small world, stepped with NO allocation of any kind.
 *
 * Who needs this: host code that is not allowed to grow the device heap while it runs.  An
 * engine that steps physics from its render / audio-paced frame thread cannot afford
 * cudaMalloc or cudaFree there — both take the driver's allocator lock and wait for the whole
 * device to go idle, so every other stream of the process (rendering interop copies, other
 * worlds) stalls behind them — and an engine that carves up the GPU's memory once at start-up
 * has no budget left for a library to allocate from later.  sfp_create / sfp_batch_create
 * size their buffers with cudaMalloc; this entry point instead runs the same one-CTA-per-world
 * frame kernel (k_batch_frame) with every intermediate buffer — AABBs, candidate pairs,
 * colours, colour tables, contact records, counters — carved out of one fixed array that is
 * part of the library image.
 *
 * The array: g_small_scratch below, SFP_NOALLOC_SCRATCH_BYTES (about 2.2 MiB) of device memory
 * on every GPU that loads the library, held until the process exits.  It bounds the world size:
 * n <= SFP_NOALLOC_MAX_BODIES (1024) and at most 12 n + 256 candidate pairs (params->max_pairs
 * if given, up to 12 544) — the same pair capacity sfp_batch_create gives such a world, so the
 * two also agree on WHEN a dense world overflows; larger requests are refused with
 * SFP_E_CAPACITY and a message, nothing is truncated silently.
 *
 * In host code the array's name is only a stand-in for the symbol, not a device address, so the
 * address is looked up with cudaGetSymbolAddress — once per device, with that device current,
 * on the FIRST call that names the device (that first call also raises the kernel's dynamic
 * shared-memory limit).  Make that first call before any stream capture begins.
 *
 * All calls on one GPU share the array and it keeps whatever the previous call left in it.
 * What has to start from zero (the counters with their sticky overflow bits; the colour table
 * is re-zeroed by the kernel itself every frame) is cleared at the start of each call; the
 * rest is fully rewritten before it is read.  Two calls that overlap in time on the SAME GPU —
 * from two threads of this process — would trample each other's intermediates on the device
 * and in that GPU's host staging slot: serialise them.  Calls on DIFFERENT GPUs may run at the
 * same time: the host staging area below has one slot per device and the per-device set-up is
 * done under std::call_once.  The call is synchronous (it returns after the result has been
 * copied back); its copies and launches go to the calling thread's cudaStreamPerThread, not to
 * the legacy default stream, so it does not serialise against the host's other streams.
 *
 * It performs: host-side conversion of `bodies` into a fixed host staging area, H2D copies into
 * the array, `frames` launches of the frame kernel (the first one colours from scratch, the
 * following ones carry colours over exactly like sfp_batch_step), D2H copy, conversion to `out`.
 * The result is bit-identical to sfp_batch_create(1 world, n bodies) + sfp_batch_upload +
 * `frames` x sfp_batch_step + sfp_batch_download.
 * ------------------------------------------------------------------------------------------ */
}  // extern "C"  (reopened below)

namespace {

constexpr uint32_t NA_MAX_BODIES = 1024;
constexpr uint32_t NA_PAIR_CAP = 12u * NA_MAX_BODIES + 256u;
constexpr uint32_t NA_TAB_SLOTS = 32768;  // power of two >= 2 * NA_PAIR_CAP
static_assert(NA_TAB_SLOTS >= 2u * NA_PAIR_CAP, "colour table too small");

constexpr size_t na_align(size_t x) { return (x + 255u) & ~(size_t)255u; }
// layout of the scratch array (every region 256-byte aligned)
constexpr size_t NA_OFF_POSE = 0;
constexpr size_t NA_OFF_PREV = NA_OFF_POSE + na_align(sizeof(float4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_VEL = NA_OFF_PREV + na_align(sizeof(float4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_SHAPE = NA_OFF_VEL + na_align(sizeof(float4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_MAT = NA_OFF_SHAPE + na_align(sizeof(float4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_MINV = NA_OFF_MAT + na_align(sizeof(float4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_AABB = NA_OFF_MINV + na_align(sizeof(float2) * NA_MAX_BODIES);
constexpr size_t NA_OFF_META = NA_OFF_AABB + na_align(sizeof(float4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_MASK = NA_OFF_META + na_align(sizeof(int4) * NA_MAX_BODIES);
constexpr size_t NA_OFF_CLAIM0 = NA_OFF_MASK + na_align(8u * NA_MAX_BODIES);
constexpr size_t NA_OFF_CLAIM1 = NA_OFF_CLAIM0 + na_align(8u * NA_MAX_BODIES);
constexpr size_t NA_OFF_CNTOFF = NA_OFF_CLAIM1 + na_align(8u * NA_MAX_BODIES);
constexpr size_t NA_OFF_PRAW = NA_OFF_CNTOFF + na_align(4u * NA_MAX_BODIES);
constexpr size_t NA_OFF_PAIRS = NA_OFF_PRAW + na_align(sizeof(uint2) * NA_PAIR_CAP);
constexpr size_t NA_OFF_PCOL = NA_OFF_PAIRS + na_align(sizeof(uint2) * NA_PAIR_CAP);
constexpr size_t NA_OFF_CREC = NA_OFF_PCOL + na_align(4u * NA_PAIR_CAP);  // 6 float4 arrays
constexpr size_t NA_CREC_STRIDE = na_align(sizeof(float4) * NA_PAIR_CAP);
constexpr size_t NA_OFF_TAB0 = NA_OFF_CREC + 6u * NA_CREC_STRIDE;
constexpr size_t NA_OFF_TAB1 = NA_OFF_TAB0 + na_align(8u * NA_TAB_SLOTS);
constexpr size_t NA_OFF_CNT = NA_OFF_TAB1 + na_align(8u * NA_TAB_SLOTS);
constexpr size_t NA_SCRATCH_BYTES = NA_OFF_CNT + na_align(sizeof(Counters));
static_assert(NA_SCRATCH_BYTES <= (size_t)SFP_NOALLOC_SCRATCH_BYTES, "sfp.h states a smaller maximum");

// the one fixed scratch area (see the comment above)
static __device__ __align__(256) unsigned char g_small_scratch[SFP_NOALLOC_SCRATCH_BYTES];

// per-device set-up state and the host staging area (fixed size: no heap), one slot per device
constexpr int NA_MAX_DEVICES = 64;
struct NaSlot {
    std::once_flag once;
    std::atomic<int> setup_rc{-1};    // cudaError_t of the set-up, -1 = not run yet
    unsigned char* scratch = nullptr;  // device address of g_small_scratch ON THIS DEVICE
    sfp::f4 pose[NA_MAX_BODIES], vel[NA_MAX_BODIES], shape[NA_MAX_BODIES], mat[NA_MAX_BODIES];
    sfp::f2 minv[NA_MAX_BODIES];
};
NaSlot g_na[NA_MAX_DEVICES];

void na_msg(char* err, uint32_t cap, const char* what, const char* detail) {
    if (err && cap) snprintf(err, cap, "%s%s%s", what, detail ? ": " : "", detail ? detail : "");
}

}  // namespace

extern "C" {

int32_t sfp_step_small_noalloc(const SfpParams* params, const SfpBodyDesc* bodies, uint32_t n,
                               float dt, uint32_t substeps, uint32_t frames, SfpBodyState* out,
                               char* err, uint32_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (!bodies || !out || n == 0 || !(dt > 0.0f) || substeps == 0 || substeps > 1024 || frames == 0) {
        na_msg(err, err_cap, "sfp_step_small_noalloc: invalid argument", nullptr);
        return SFP_E_INVALID;
    }
    if (n > NA_MAX_BODIES) {
        if (err && err_cap)
            snprintf(err, err_cap,
                     "sfp_step_small_noalloc: %u bodies requested, the fixed scratch area holds at "
                     "most %u (and %u candidate pairs); use sfp_create / sfp_batch_create for "
                     "larger worlds",
                     n, NA_MAX_BODIES, NA_PAIR_CAP);
        return SFP_E_CAPACITY;
    }
    SfpParams p;
    if (params) p = *params;
    else sfp::default_params(&p);
    if (!sfp::params_valid(p) || p.solver_order != SFP_ORDER_COLOUR || p.device < 0 ||
        p.device >= NA_MAX_DEVICES) {
        na_msg(err, err_cap, "sfp_step_small_noalloc: invalid parameters", nullptr);
        return SFP_E_INVALID;
    }
    for (uint32_t i = 0; i < n; ++i)
        if (!sfp::desc_valid(bodies[i])) {
            if (err && err_cap)
                snprintf(err, err_cap, "sfp_step_small_noalloc: invalid body desc at index %u", i);
            return SFP_E_INVALID;
        }
#define NCU(call)                                                   \
    do {                                                            \
        cudaError_t e_ = (call);                                    \
        if (e_ != cudaSuccess) {                                    \
            na_msg(err, err_cap, #call, cudaGetErrorString(e_));    \
            return SFP_E_CUDA;                                      \
        }                                                           \
    } while (0)
    const uint32_t pair_cap = p.max_pairs ? p.max_pairs : 12u * n + 256u;
    if (pair_cap > NA_PAIR_CAP) {
        if (err && err_cap)
            snprintf(err, err_cap,
                     "sfp_step_small_noalloc: max_pairs %u requested, the fixed scratch area holds at "
                     "most %u candidate pairs",
                     pair_cap, NA_PAIR_CAP);
        return SFP_E_CAPACITY;
    }
    NCU(cudaSetDevice(p.device));
    NaSlot& slot = g_na[p.device];
    // once per device, with that device current: the symbol's address ON THIS DEVICE and the
    // kernel's shared-memory limit (two threads naming the same device for the first time: one
    // runs the set-up, the other waits for it)
    std::call_once(slot.once, [&] {
        unsigned char* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_small_scratch);
        if (e == cudaSuccess) e = raise_batch_smem_limit(p.device, (size_t)NA_MAX_BODIES * 80u);
        slot.scratch = ptr;
        slot.setup_rc.store((int)e, std::memory_order_release);
    });
    if (slot.setup_rc.load(std::memory_order_acquire) != (int)cudaSuccess) {
        na_msg(err, err_cap, "sfp_step_small_noalloc: set-up of the scratch area failed",
               cudaGetErrorString((cudaError_t)slot.setup_rc.load()));
        return SFP_E_CUDA;
    }
    unsigned char* base = slot.scratch;
    const cudaStream_t st = cudaStreamPerThread;  // a constant handle: nothing is created

    BatchCtx ctx;
    std::memset(&ctx, 0, sizeof(ctx));
    Ctx& c = ctx.base;
    c.pose = reinterpret_cast<float4*>(base + NA_OFF_POSE);
    c.prev = reinterpret_cast<float4*>(base + NA_OFF_PREV);
    c.vel = reinterpret_cast<float4*>(base + NA_OFF_VEL);
    c.shape = reinterpret_cast<const float4*>(base + NA_OFF_SHAPE);
    c.mat = reinterpret_cast<const float4*>(base + NA_OFF_MAT);
    c.minv = reinterpret_cast<const float2*>(base + NA_OFF_MINV);
    c.aabb = reinterpret_cast<float4*>(base + NA_OFF_AABB);
    c.meta = reinterpret_cast<int4*>(base + NA_OFF_META);
    c.mask = reinterpret_cast<unsigned long long*>(base + NA_OFF_MASK);
    c.claim0 = reinterpret_cast<unsigned long long*>(base + NA_OFF_CLAIM0);
    c.claim1 = reinterpret_cast<unsigned long long*>(base + NA_OFF_CLAIM1);
    c.pairs_raw = reinterpret_cast<uint2*>(base + NA_OFF_PRAW);
    c.pairs = reinterpret_cast<uint2*>(base + NA_OFF_PAIRS);
    c.pcol = reinterpret_cast<uint32_t*>(base + NA_OFF_PCOL);
    float4* crec = reinterpret_cast<float4*>(base + NA_OFF_CREC);
    const size_t cstride = NA_CREC_STRIDE / sizeof(float4);
    c.ca = crec;
    c.cb = crec + cstride;
    c.cp0 = crec + 2 * cstride;
    c.cp1 = crec + 3 * cstride;
    c.cs = crec + 4 * cstride;
    c.lam = crec + 5 * cstride;
    c.jbod_in = reinterpret_cast<const uint2*>(c.pairs_raw);
    c.jcol = c.pcol;
    c.tab_mask = NA_TAB_SLOTS - 1u;
    c.gx = p.gravity_x;
    c.gy = p.gravity_y;
    c.inv_cell = 1.0f / p.cell_size;
    c.aabb_margin = p.aabb_margin;
    c.vel_expand = p.vel_expand;
    c.contact_margin = p.contact_margin;
    ctx.bpw = n;
    ctx.pair_cap = pair_cap;
    ctx.n_worlds = 1;
    ctx.n_launch = 1;  // one CTA runs the whole world
    ctx.n_whole = 1;
    ctx.queue = nullptr;
    ctx.ring = nullptr;
    ctx.progress = nullptr;
    ctx.counters = reinterpret_cast<Counters*>(base + NA_OFF_CNT);
    ctx.scratch = reinterpret_cast<uint32_t*>(base + NA_OFF_CNTOFF);
    ctx.tab_slots = NA_TAB_SLOTS;
    unsigned long long* tab[2] = {reinterpret_cast<unsigned long long*>(base + NA_OFF_TAB0),
                                  reinterpret_cast<unsigned long long*>(base + NA_OFF_TAB1)};

    for (uint32_t i = 0; i < n; ++i) {
        sfp::body_from_desc(bodies[i], &slot.pose[i], &slot.vel[i], &slot.minv[i], &slot.shape[i],
                            &slot.mat[i]);
        if (!sfp::body_fits_grid(slot.pose[i], slot.shape[i], p)) {
            if (err && err_cap)
                snprintf(err, err_cap,
                         "sfp_step_small_noalloc: body %u spans more than 16384 broadphase cells (or "
                         "4096 along one axis) at this cell_size", i);
            return SFP_E_INVALID;
        }
    }
    NCU(cudaMemcpyAsync(c.pose, slot.pose, sizeof(sfp::f4) * n, cudaMemcpyHostToDevice, st));
    NCU(cudaMemcpyAsync(c.vel, slot.vel, sizeof(sfp::f4) * n, cudaMemcpyHostToDevice, st));
    NCU(cudaMemcpyAsync(const_cast<float4*>(c.shape), slot.shape, sizeof(sfp::f4) * n,
                        cudaMemcpyHostToDevice, st));
    NCU(cudaMemcpyAsync(const_cast<float4*>(c.mat), slot.mat, sizeof(sfp::f4) * n, cudaMemcpyHostToDevice, st));
    NCU(cudaMemcpyAsync(const_cast<float2*>(c.minv), slot.minv, sizeof(sfp::f2) * n,
                        cudaMemcpyHostToDevice, st));
    // the previous call's counters (sticky overflow bits, colour counts) must not leak in
    NCU(cudaMemsetAsync(ctx.counters, 0, sizeof(Counters), st));

    const float h = dt / (float)substeps;
    const float g = sqrtf(p.gravity_x * p.gravity_x + p.gravity_y * p.gravity_y);
    for (uint32_t f = 0; f < frames; ++f) {
        c.warm = (f % SFP_RECOLOUR_PERIOD) != 0u ? 1u : 0u;
        c.tab_old = tab[(f + 1u) & 1u];
        c.tab_new = tab[f & 1u];
        k_batch_frame<true><<<1, BTPB, (size_t)n * 80u, st>>>(ctx, dt, substeps, 2.0f * g * h);
        NCU(cudaGetLastError());
    }
    NCU(cudaMemcpyAsync(slot.pose, c.pose, sizeof(sfp::f4) * n, cudaMemcpyDeviceToHost, st));
    NCU(cudaMemcpyAsync(slot.vel, c.vel, sizeof(sfp::f4) * n, cudaMemcpyDeviceToHost, st));
    Counters k;
    NCU(cudaMemcpyAsync(&k, ctx.counters, sizeof(k), cudaMemcpyDeviceToHost, st));
    NCU(cudaStreamSynchronize(st));
#undef NCU
    for (uint32_t i = 0; i < n; ++i) out[i] = sfp::state_from_records(slot.pose[i], slot.vel[i]);
    if (k.overflow) {
        if (err && err_cap)
            snprintf(err, err_cap,
                     "sfp_step_small_noalloc: capacity exceeded inside the fixed scratch area "
                     "(bits %u: 2 = more than %u candidate pairs, 4 = serial group too large, "
                     "16 = runaway body)",
                     k.overflow, pair_cap);
        return SFP_E_CAPACITY;
    }
    return SFP_OK;
}

}  // extern "C"
