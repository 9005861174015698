// sfp_device.cuh — device-side stages of the physics step for ONE world in global memory.
//
// Each stage is a __device__ function written as a grid-stride loop over (tid, nthreads) so the
// same code runs (a) inside the persistent cooperative kernels, separated by grid.sync(), which
// is the product path (no per-substep host launches, north_star), and (b) as one plain kernel
// per stage (SFP_FLAG_MULTI_KERNEL, debugging / profiling).
//
// Per-item arithmetic comes from include/sfp_math.h (shared with the oracle, FMA contraction
// off).  Data layout (SoA): one float4 array per record, 88 B/body; contacts
// 6 x float4 per candidate pair, stored in colour-sorted pair order so every solver colour
// reads and writes a contiguous range; bodies are the only gathered accesses.
//
// Reference: the reference's step is CPU Rust and is not mounted; stage names
// follow north_star / SURVEY.md §8a rows A2..A8.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "sfp_math.h"

namespace sfpk {

using sfp::f2;
using sfp::f4;
using sfp::v2;

constexpr uint32_t UNCOLOURED = 0xffffffffu;
// overflow bits 1 (entries) and 2 (pairs): the broadphase ran out of room.  Every later kernel
// of the frame — and every kernel of the frames queued behind it — then does nothing, so that
// the world stays exactly as it was before the frame and the host can enlarge the buffers and
// run those frames again (sfp_world.cu, recover_capacity).
constexpr uint32_t OVERFLOW_CAPACITY = 3u;
constexpr int MAX_ROUNDS = 512;

struct Counters {
    uint32_t n_bodies, n_joints;
    uint32_t n_entries, n_pairs;
    uint32_t n_colours, colour_rounds;
    uint32_t overflow;
    uint32_t n_contacts, n_manifolds, maxpen_bits;
    uint32_t n_unc;  // length of Ctx::unc (pairs still uncoloured after the warm start)
    uint32_t n_unc2; // length of Ctx::unc2 (the list the odd rounds read; see stage_colour_take)
    uint32_t n_colours_int;  // interior colours in use (tiles; n_colours counts the boundary ones)
    uint32_t frames_done;    // frames whose substep kernel ran to the end (capacity recovery)
    uint32_t np_cls_off[3];  // batched worlds: where classes 1, 2, 3 start in Ctx::np_order
    int qbox[4];             // spatial queries: cell bounding box of all bodies (x0, y0, x1, y1)
    uint32_t pair_col_off[SFP_MAX_COLOURS + 1];
    uint32_t joint_col_off[SFP_MAX_COLOURS + 1];
    uint32_t round_flags[MAX_ROUNDS];
    // SM-clock cycles spent per stage in the last frame, written by thread 0 (batched path:
    // of each world's CTA): 0 load/store, 1 broadphase, 2 colouring, 3 bucketing, 4 integrate,
    // 5 narrowphase, 6 position solve, 7 derive, 8 velocity solve
    unsigned long long stage_cycles[12];  // [10] / [11]: tile-interior position / velocity solve
};

// step lengths of the current frame.  They live in device memory (written before the frame's
// launches), not in kernel arguments: a captured frame graph then serves ANY dt, where a graph
// with dt baked in missed the cache whenever a host passed its measured frame delta.
struct FrameParams {
    float dt, h, rest_threshold, pad;
};

struct Ctx {
    const FrameParams* fp;
    // body state
    float4 *pose, *prev, *vel;
    const float4 *shape, *mat;
    const float2* minv;
    const float4* polytab;  // general-polygon table, 8 float4 per polygon (sfp_math.h), or null
    // The single-world path keeps its bodies in a spatial order of its own (sfp_world.cu,
    // "re-sort"): ext[i] is the caller's index (upload order) of the body stored at i, inv the
    // inverse.  Everything whose RESULT depends on body numbering — which body of a pair is "a",
    // constraint priorities, the order inside the serial group, ties — uses the caller's index,
    // so the results do not depend on the internal order.  Null (batched worlds): identity.
    const uint32_t *ext, *inv;
    // broadphase
    float4* aabb;
    int4* meta;  // cx0, cy0, layer bits, dynamic flag
    uint32_t *cell_cnt, *cell_off;
    uint32_t *ekey0, *eval0, *ekey1, *eval1;
    uint32_t *pcnt, *poff;
    uint2* pairs_raw;
    uint32_t* np_order;  // batched worlds: pair positions in narrowphase-class order (sfp_batch.cu), or null
    unsigned long long* pprio;  // per raw pair: its colouring priority (from the caller's indices); null: computed on the fly
    // tiles (sfp_math.h, "Tiles"): after the re-sort the dynamic bodies are the indices [0, n_dyn)
    // and tile t is the range [t * tile_b, (t + 1) * tile_b).  tile_b == 0: tiles off, every
    // constraint is a boundary constraint.  Sorted constraint arrays: tile 0's interior
    // constraints by colour, tile 1's, ..., then the boundary constraints by colour.  goff / jgoff:
    // first sorted pair / joint of (group g, colour col) at [g * 64 + col], group n_tiles being the
    // boundary (whose offsets Counters::pair_col_off / joint_col_off repeat).
    uint32_t tile_b, n_dyn, n_tiles;
    uint32_t *goff, *jgoff;
    unsigned long long* mask_i;  // per body: interior colours in use (mask: boundary colours)
    // colouring
    unsigned long long *mask, *claim0, *claim1;
    // dataflow solve (k_substeps_flow): per body, the colours whose constraint has been applied
    // in the current substep's position pass [2 i] and velocity pass [2 i + 1]; null elsewhere
    unsigned long long* done;
    uint32_t *pcol, *jcol;
    // pairs that the warm start left uncoloured, in any order (null: the rounds walk all pairs);
    // after a warm start that is a few per cent of the pairs
    uint32_t* unc;
    uint32_t* unc2;  // second list: a round reads one and appends the pairs it leaves uncoloured to the other
    uint32_t *pcol_sorted, *pperm, *jcol_sorted, *jperm;
    // colour tables (sfp_math.h, SFP_CTAB_*): last frame's (read) and this frame's (built)
    unsigned long long *tab_old, *tab_new;
    uint32_t tab_mask;  // slots - 1 (power of two)
    uint32_t warm;      // 1: carry colours over from tab_old / the previous jcol
    uint2* pairs;  // colour-sorted
    // joints: upload order (…_in) and colour-sorted
    const uint2* jbod_in;
    const float4 *janc_in, *jpar_in;
    uint2* jbod;
    float4 *janc, *jpar;
    // contacts (per colour-sorted pair)
    float4 *ca, *cb, *cp0, *cp1, *cs, *lam;
    // per sorted pair, constant within a frame (written by stage_gather, read by every substep's
    // narrowphase with coalesced loads instead of four gathers): the two shapes, and
    // (mu_s, mu_d, e, reach of the bounding-circle test).  Null: gather from the bodies.
    float4 *psa, *psb, *pmt;
    Counters* cnt;
    // parameters
    float gx, gy, inv_cell, aabb_margin, vel_expand, contact_margin;
    uint32_t n, nj, cap_entries, cap_pairs;
};

__device__ __forceinline__ f4 ld4(const float4* p, uint32_t i) {
    float4 v = p[i];
    f4 r;
    r.x = v.x;
    r.y = v.y;
    r.z = v.z;
    r.w = v.w;
    return r;
}
__device__ __forceinline__ void st4(float4* p, uint32_t i, f4 v) {
    p[i] = make_float4(v.x, v.y, v.z, v.w);
}
__device__ __forceinline__ f2 ld2(const float2* p, uint32_t i) {
    float2 v = p[i];
    f2 r;
    r.x = v.x;
    r.y = v.y;
    return r;
}

__device__ __forceinline__ uint32_t ext_id(const Ctx& c, uint32_t i) { return c.ext ? c.ext[i] : i; }
__device__ __forceinline__ uint32_t tile_of(const Ctx& c, uint32_t i) {
    return (c.tile_b != 0u && i < c.n_dyn) ? i / c.tile_b : SFP_NO_TILE;
}
// sort group of a constraint: its tile if interior, n_tiles for the boundary
__device__ __forceinline__ uint32_t group_of(const Ctx& c, uint32_t a, uint32_t b) {
    const uint32_t t = sfp::constraint_tile(tile_of(c, a), tile_of(c, b));
    return t == SFP_NO_TILE ? c.n_tiles : t;
}

// ---------------------------------------------------------------------------------------
// A2 broadphase
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_aabb_count(const Ctx& c, float dt, uint32_t tid,
                                                 uint32_t nth, bool new_frame = true) {
    for (uint32_t i = tid; i < c.n; i += nth) {
        f4 pose = ld4(c.pose, i), vel = ld4(c.vel, i), shape = ld4(c.shape, i);
        f2 mi = ld2(c.minv, i);
        f4 bb = sfp::body_aabb(pose, vel, shape, c.aabb_margin, c.vel_expand, dt,
                               reinterpret_cast<const sfp::f4*>(c.polytab));
        int x0, y0, x1, y1;
        if (!sfp::cell_range(bb, c.inv_cell, &x0, &y0, &x1, &y1)) atomicOr(&c.cnt->overflow, 16u);
        st4(c.aabb, i, bb);
        int dyn = (mi.x > 0.0f || mi.y > 0.0f) ? 1 : 0;
        c.meta[i] = make_int4(x0, y0, (int)sfp::f2u(c.mat[i].w), dyn);
        c.cell_cnt[i] = (uint32_t)(x1 - x0 + 1) * (uint32_t)(y1 - y0 + 1);
        if (!new_frame && x1 >= x0) {  // query hash: the world's cell bounding box
            atomicMin(&c.cnt->qbox[0], x0);
            atomicMin(&c.cnt->qbox[1], y0);
            atomicMax(&c.cnt->qbox[2], x1);
            atomicMax(&c.cnt->qbox[3], y1);
        }
    }
    if (tid == 0 && new_frame) {
        // per-frame counters
        c.cnt->n_colours = 0;
        c.cnt->n_colours_int = 0;
        c.cnt->colour_rounds = 0;
    }
}

// (cell, body) entries, written to the SECOND buffer pair: the three radix passes that follow
// (24-bit key) ping-pong 1 -> 0 -> 1 -> 0 and leave the sorted list in ekey0 / eval0.
// A body covering few cells writes them itself; a body covering more than 32 is handed to its
// whole warp, 32 cells per step; a body covering more than 512 (the floor and the walls of a
// pile cover a thousand and more: one warp walking them alone was 10 us of this 20 us stage)
// is handed to the whole block.  Every thread of the block must call this.
__device__ __forceinline__ void stage_emit_entries(const Ctx& c, uint32_t tid, uint32_t nth) {
    constexpr uint32_t HUGE_SLOTS = 8;
    __shared__ int s_hx0[HUGE_SLOTS], s_hy0[HUGE_SLOTS];
    __shared__ uint32_t s_hw[HUGE_SLOTS], s_ho[HUGE_SLOTS], s_hcnt[HUGE_SLOTS], s_hi[HUGE_SLOTS];
    __shared__ uint32_t s_nh;
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t bbase = tid - threadIdx.x; bbase < c.n; bbase += nth) {
        if (threadIdx.x == 0) s_nh = 0u;
        __syncthreads();
        const uint32_t i = bbase + threadIdx.x;
        const uint32_t base = i - lane;
        int x0 = 0, y0 = 0, x1 = -1, y1 = -1;
        uint32_t o = 0, cnt = 0;
        if (i < c.n) {
            f4 bb = ld4(c.aabb, i);
            sfp::cell_range(bb, c.inv_cell, &x0, &y0, &x1, &y1);
            o = c.cell_off[i];
            cnt = (uint32_t)(x1 - x0 + 1) * (uint32_t)(y1 - y0 + 1);
        }
        bool taken = false;
        if (cnt > 512u) {
            const uint32_t slot = atomicAdd(&s_nh, 1u);
            if (slot < HUGE_SLOTS) {
                s_hx0[slot] = x0;
                s_hy0[slot] = y0;
                s_hw[slot] = (uint32_t)(x1 - x0 + 1);
                s_ho[slot] = o;
                s_hcnt[slot] = cnt;
                s_hi[slot] = i;
                taken = true;
            }
        }
        const bool big = cnt > 32u && !taken;
        if (cnt <= 32u) {
            for (int cy = y0; cy <= y1; ++cy)
                for (int cx = x0; cx <= x1; ++cx) {
                    if (o < c.cap_entries) {
                        c.ekey1[o] = sfp::cell_key(cx, cy);
                        c.eval1[o] = i;
                    }
                    ++o;
                }
        }
        uint32_t todo = __ballot_sync(0xffffffffu, big);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1u;
            const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
            const int bx1 = __shfl_sync(0xffffffffu, x1, src);
            const uint32_t bo = __shfl_sync(0xffffffffu, o, src);
            const uint32_t bcnt = __shfl_sync(0xffffffffu, cnt, src);
            const uint32_t bi = base + (uint32_t)src;
            const uint32_t w = (uint32_t)(bx1 - bx0 + 1);
            for (uint32_t k = lane; k < bcnt; k += 32u) {
                uint32_t pos = bo + k;
                if (pos < c.cap_entries) {
                    c.ekey1[pos] = sfp::cell_key(bx0 + (int)(k % w), by0 + (int)(k / w));
                    c.eval1[pos] = bi;
                }
            }
        }
        __syncthreads();
        const uint32_t nh = s_nh < HUGE_SLOTS ? s_nh : HUGE_SLOTS;
        for (uint32_t j = 0; j < nh; ++j) {
            const uint32_t w = s_hw[j], bo = s_ho[j], bcnt = s_hcnt[j], bi = s_hi[j];
            const int bx0 = s_hx0[j], by0 = s_hy0[j];
            for (uint32_t k = threadIdx.x; k < bcnt; k += blockDim.x) {
                uint32_t pos = bo + k;
                if (pos < c.cap_entries) {
                    c.ekey1[pos] = sfp::cell_key(bx0 + (int)(k % w), by0 + (int)(k / w));
                    c.eval1[pos] = bi;
                }
            }
        }
        __syncthreads();
    }
    if (tid == 0 && c.cnt->n_entries > c.cap_entries) {
        c.cnt->n_entries = c.cap_entries;
        c.cnt->overflow |= 1u;
    }
}

// pair test shared by the count and the write pass; entries are sorted by cell key
__device__ __forceinline__ bool pair_test(int cx, int cy, uint32_t a, f4 bba, int4 ma, uint32_t b,
                                          int4 mb, f4 bbb) {
    // (a conjunction: the box test goes first because it rejects most candidates of a cell)
    if (!sfp::aabb_overlap(bba, bbb)) return false;
    if (a == b) return false;
    if ((ma.w | mb.w) == 0) return false;
    if (!sfp::same_cell(max(ma.x, mb.x), max(ma.y, mb.y), cx, cy)) return false;
    return sfp::layers_collide((uint32_t)ma.z, (uint32_t)mb.z);
}

// One thread per entry: the rest of the entry's cell is scanned for partners.  The scan is a
// chain of dependent loads (next key -> its body -> that body's box), so it goes four entries
// at a time: the four keys and bodies are loaded together, then the boxes of those that are
// still in the cell — two round trips per four candidates instead of three per candidate.
template <bool WRITE>
__device__ __forceinline__ void stage_pairs(const Ctx& c, const uint32_t* ekey,
                                            const uint32_t* eval, uint32_t tid, uint32_t nth) {
    const uint32_t E = c.cnt->n_entries;
    for (uint32_t e = tid; e < E; e += nth) {
        const uint32_t key = ekey[e], a = eval[e];
        int cx = sfp::key_cx(key), cy = sfp::key_cy(key);
        f4 bba = ld4(c.aabb, a);
        int4 ma = c.meta[a];
        uint32_t k = 0;
        uint32_t o = WRITE ? c.poff[e] : 0u;
        // the write pass knows from the count pass how many partners this entry has: none (close
        // to half of the entries of a pile) needs no scan, and the scan stops at the last one
        const uint32_t want = WRITE ? c.pcnt[e] : 0xffffffffu;
        if (WRITE && want == 0u) continue;
        for (uint32_t f = e + 1; f < E; f += 4u) {
            uint32_t kf[4], bf[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t idx = f + (uint32_t)j;
                const bool in = idx < E;
                kf[j] = in ? ekey[idx] : ~key;
                bf[j] = in ? eval[idx] : 0u;
            }
            bool same[4];
            same[0] = kf[0] == key;
#pragma unroll
            for (int j = 1; j < 4; ++j) same[j] = same[j - 1] && kf[j] == key;
            int4 mb[4];
            f4 bbb[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (same[j]) {
                    mb[j] = c.meta[bf[j]];
                    bbb[j] = ld4(c.aabb, bf[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (same[j] && pair_test(cx, cy, a, bba, ma, bf[j], mb[j], bbb[j])) {
                    if (WRITE) {
                        if (o + k < c.cap_pairs) {
                            // "a" is the body with the lower index in the caller's numbering
                            const uint32_t ea = ext_id(c, a), eb = ext_id(c, bf[j]);
                            const bool fwd = ea < eb;
                            c.pairs_raw[o + k] = fwd ? make_uint2(a, bf[j]) : make_uint2(bf[j], a);
                            if (c.pprio) c.pprio[o + k] = sfp::pair_priority(min(ea, eb), max(ea, eb));
                        }
                    }
                    ++k;
                }
            }
            if (!same[3] || k >= want) break;
        }
        if (!WRITE) c.pcnt[e] = k;
    }
    if (WRITE && tid == 0 && c.cnt->n_pairs > c.cap_pairs) {
        c.cnt->n_pairs = c.cap_pairs;
        c.cnt->overflow |= 2u;
    }
}

// Colours travel as `enc`: 0..62 a boundary colour, 63 the serial group, 64 + c an interior colour
// (what sfp_get_colours reports).
// Sort key of a coloured pair: bits 0-1 narrowphase class, 2-7 colour, 8-19 group (tile, or
// n_tiles for the boundary and the serial group) — the radix passes sort on these —, bit 30 / 31
// set when body a / b is static (rides along into pcol_sorted for the dataflow experiment).
constexpr uint32_t PAIR_KEY_COLOUR_MASK = (uint32_t)SFP_MAX_COLOURS - 1u;
constexpr uint32_t PAIR_KEY_GROUP_SHIFT = 8u, PAIR_KEY_GROUP_MASK = 0xfffu;
constexpr uint32_t PAIR_KEY_STATIC_A = 1u << 30, PAIR_KEY_STATIC_B = 1u << 31;
__device__ __forceinline__ uint32_t pair_sort_key(const Ctx& c, uint32_t enc, uint2 p) {
    const uint32_t col = enc & PAIR_KEY_COLOUR_MASK;
    const uint32_t group = (enc >= SFP_CLASS_INTERIOR) ? group_of(c, p.x, p.y) : c.n_tiles;
    return (col << SFP_TYPE_BITS) | sfp::pair_type_key(ld4(c.shape, p.x), ld4(c.shape, p.y)) |
           (group << PAIR_KEY_GROUP_SHIFT) | (c.meta[p.x].w ? 0u : PAIR_KEY_STATIC_A) |
           (c.meta[p.y].w ? 0u : PAIR_KEY_STATIC_B);
}
__device__ __forceinline__ uint32_t pair_key_colour(uint32_t key) {
    return (key >> SFP_TYPE_BITS) & PAIR_KEY_COLOUR_MASK;
}
__device__ __forceinline__ uint32_t pair_key_group(uint32_t key) {
    return (key >> PAIR_KEY_GROUP_SHIFT) & PAIR_KEY_GROUP_MASK;
}
__device__ __forceinline__ uint32_t pair_key_enc(const Ctx& c, uint32_t key) {
    return pair_key_colour(key) + (pair_key_group(key) != c.n_tiles ? SFP_CLASS_INTERIOR : 0u);
}
// joints: enc in bits 0-7, group above
__device__ __forceinline__ uint32_t joint_sort_key(const Ctx& c, uint32_t enc, uint2 jb) {
    const uint32_t group = (enc >= SFP_CLASS_INTERIOR) ? group_of(c, jb.x, jb.y) : c.n_tiles;
    return enc | (group << PAIR_KEY_GROUP_SHIFT);
}

// ---------------------------------------------------------------------------------------
// A8 colouring: rounds of bid (atomicMin of a unique 64-bit priority on each dynamic body) and
// take (a constraint holding the claim on all its dynamic bodies takes the lowest free colour)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_colour_init(const Ctx& c, uint32_t tid, uint32_t nth) {
    const uint32_t P = c.cnt->n_pairs;
    for (uint32_t i = tid; i < c.n; i += nth) {
        c.mask[i] = 0ull;
        if (c.tile_b) c.mask_i[i] = 0ull;
        c.claim0[i] = SFP_NO_CLAIM;
        c.claim1[i] = SFP_NO_CLAIM;
    }
    for (uint32_t k = tid; k < P; k += nth) c.pcol[k] = UNCOLOURED;
    if (c.unc) {
        // no warm start: every pair is on the list; warm start: stage_colour_warm appends
        if (!c.warm)
            for (uint32_t k = tid; k < P; k += nth) c.unc[k] = k;
        if (tid == 0) c.cnt->n_unc = c.warm ? 0u : P;
    }
    for (uint32_t k = tid; k < c.nj; k += nth) {
        uint2 jb = c.jbod_in[k];
        bool dyn = (c.meta[jb.x].w | c.meta[jb.y].w) != 0;
        if (!dyn) {
            c.jcol[k] = joint_sort_key(c, 0u, jb);
            atomicMax(&c.cnt->n_colours, 1u);
        } else if (!c.warm) {
            c.jcol[k] = UNCOLOURED;
        }
    }
    for (uint32_t k = tid; k < MAX_ROUNDS; k += nth) c.cnt->round_flags[k] = 0u;
}

// warm start: a constraint that existed in the previous frame keeps its colour.  Runs after
// stage_colour_init (masks are zero), before the first round.
// (the colour counts are accumulated per thread and published once per warp: an atomicMax per
// constraint on one of two addresses chosen per lane is 340 k serialised atomics on a 100 k pile)
struct ColourCount {
    uint32_t bnd = 0, inter = 0;
};
__device__ __forceinline__ void colour_count_publish(const Ctx& c, ColourCount n) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        n.bnd = max(n.bnd, __shfl_xor_sync(0xffffffffu, n.bnd, d));
        n.inter = max(n.inter, __shfl_xor_sync(0xffffffffu, n.inter, d));
    }
    if ((threadIdx.x & 31u) == 0u) {
        if (n.bnd) atomicMax(&c.cnt->n_colours, n.bnd);
        if (n.inter) atomicMax(&c.cnt->n_colours_int, n.inter);
    }
}
__device__ __forceinline__ void colour_keep(const Ctx& c, uint32_t a, uint32_t b, uint32_t enc,
                                            ColourCount* n) {
    const uint32_t col = enc & PAIR_KEY_COLOUR_MASK;
    const bool inter = enc >= SFP_CLASS_INTERIOR;
    if (col != SFP_SERIAL_COLOUR) {
        unsigned long long* mk = inter ? c.mask_i : c.mask;
        if (c.meta[a].w) atomicOr(&mk[a], 1ull << col);
        if (c.meta[b].w) atomicOr(&mk[b], 1ull << col);
    }
    if (inter) n->inter = max(n->inter, col + 1u);
    else n->bnd = max(n->bnd, col + 1u);
}
__device__ __forceinline__ void stage_colour_warm(const Ctx& c, uint32_t tid, uint32_t nth) {
    if (!c.warm) return;
    const uint32_t P = c.cnt->n_pairs;
    // (whole warps walk the loop together: the append below is warp-aggregated)
    const uint32_t lane = threadIdx.x & 31u;
    ColourCount ncol;
    for (uint32_t base = tid - lane; base < P; base += nth) {
        const uint32_t k = base + lane;
        bool found = k >= P;  // lanes past the end have nothing to append
        if (k < P) {
            uint2 p = c.pairs_raw[k];
            const unsigned long long key = sfp::ctab_key(p.x, p.y);
            uint32_t h = sfp::ctab_hash(p.x, p.y) & c.tab_mask;
            for (;;) {
                unsigned long long slot = c.tab_old[h];
                if (slot == 0ull) break;
                if ((slot & SFP_CTAB_KEYMASK) == key) {
                    uint32_t col = (uint32_t)(slot >> 44) - 1u;
                    colour_keep(c, p.x, p.y, col, &ncol);
                    c.pcol[k] = pair_sort_key(c, col, p);
                    found = true;
                    break;
                }
                h = (h + 1u) & c.tab_mask;
            }
        }
        if (c.unc) {
            const uint32_t miss = __ballot_sync(0xffffffffu, !found);
            if (miss) {
                uint32_t at = 0;
                if (lane == 0) at = atomicAdd(&c.cnt->n_unc, (uint32_t)__popc(miss));
                at = __shfl_sync(0xffffffffu, at, 0);
                if (!found) c.unc[at + __popc(miss & ((1u << lane) - 1u))] = k;
            }
        }
    }
    for (uint32_t k = tid; k < c.nj; k += nth) {
        uint32_t col = c.jcol[k];
        uint2 jb = c.jbod_in[k];
        if (col != UNCOLOURED && (c.meta[jb.x].w | c.meta[jb.y].w) != 0)
            colour_keep(c, jb.x, jb.y, col & 0xffu, &ncol);
    }
    colour_count_publish(c, ncol);
}

// record this frame's colours for the next one (tab_new was zeroed by the host / the caller)
__device__ __forceinline__ void stage_colour_record(const Ctx& c, uint32_t tid, uint32_t nth) {
    const uint32_t P = c.cnt->n_pairs;
    for (uint32_t k = tid; k < P; k += nth) {
        uint32_t key16 = c.pcol[k];
        if (key16 == UNCOLOURED) continue;
        uint2 p = c.pairs_raw[k];
        const unsigned long long packed = sfp::ctab_pack(p.x, p.y, pair_key_enc(c, key16));
        uint32_t h = sfp::ctab_hash(p.x, p.y) & c.tab_mask;
        while (atomicCAS(&c.tab_new[h], 0ull, packed) != 0ull) h = (h + 1u) & c.tab_mask;
    }
}

// colours with more than SFP_TAIL_ITEMS constraints, as a bit mask (the solve order is: those
// in ascending index, then the others in ascending index)
__device__ __forceinline__ unsigned long long big_colour_mask(const Ctx& c, uint32_t ncol) {
    unsigned long long m = 0ull;
    for (uint32_t col = 0; col < ncol; ++col) {
        uint32_t items = (c.cnt->pair_col_off[col + 1] - c.cnt->pair_col_off[col]) +
                         (c.cnt->joint_col_off[col + 1] - c.cnt->joint_col_off[col]);
        if (items > SFP_TAIL_ITEMS && col != SFP_SERIAL_COLOUR) m |= 1ull << col;
    }
    return m;
}

// ---- the serial group ("colour" 63, sfp_math.h): members may share bodies ----------------------
__device__ __forceinline__ uint32_t serial_pairs(const Ctx& c, uint32_t ncol, uint32_t* first) {
    if (ncol <= SFP_SERIAL_COLOUR) return 0u;
    *first = c.cnt->pair_col_off[SFP_SERIAL_COLOUR];
    return c.cnt->pair_col_off[SFP_SERIAL_COLOUR + 1] - *first;
}
// the serial group's position / velocity pass, executed by ONE thread
__device__ __forceinline__ void serial_solve_pos(const Ctx& c, uint32_t ncol, float inv_h2);
__device__ __forceinline__ void serial_solve_vel(const Ctx& c, uint32_t ncol, float h,
                                                 float rest_threshold);
// put the serial pairs into (a, b) order: odd-even transposition by ONE block (call with the
// block's threadIdx / blockDim; every thread of the block must call it).  Runs once per frame.
__device__ __forceinline__ void serial_sort_pairs(const Ctx& c, uint32_t ncol) {
    uint32_t first = 0;
    const uint32_t m = serial_pairs(c, ncol, &first);
    if (m > SFP_SERIAL_MAX && threadIdx.x == 0) atomicOr(&c.cnt->overflow, 4u);
    if (m < 2u) return;
    for (uint32_t round = 0; round < m; ++round) {
        for (uint32_t i = 2u * threadIdx.x + (round & 1u); i + 1u < m; i += 2u * blockDim.x) {
            uint2 x = c.pairs[first + i], y = c.pairs[first + i + 1u];
            const uint32_t xa = ext_id(c, x.x), xb = ext_id(c, x.y), ya = ext_id(c, y.x), yb = ext_id(c, y.y);
            if (xa > ya || (xa == ya && xb > yb)) {
                c.pairs[first + i] = y;
                c.pairs[first + i + 1u] = x;
                if (c.psa) {  // the pairs' per-frame narrowphase constants move with them
                    const uint32_t u = first + i, v = u + 1u;
                    float4 t = c.psa[u];
                    c.psa[u] = c.psa[v];
                    c.psa[v] = t;
                    t = c.psb[u];
                    c.psb[u] = c.psb[v];
                    c.psb[v] = t;
                    t = c.pmt[u];
                    c.pmt[u] = c.pmt[v];
                    c.pmt[v] = t;
                }
            }
        }
        __syncthreads();
    }
}

__device__ __forceinline__ void colour_bid_one(const Ctx& c, unsigned long long* claim, uint32_t a,
                                               uint32_t b, unsigned long long key) {
    if (c.meta[a].w) atomicMin(&claim[a], key);
    if (c.meta[b].w) atomicMin(&claim[b], key);
}

__device__ __forceinline__ void stage_colour_bid(const Ctx& c, uint32_t round, uint32_t tid,
                                                 uint32_t nth) {
    unsigned long long* claim = (round & 1u) ? c.claim1 : c.claim0;
    // the list of this round (even rounds: unc, odd rounds: unc2); the take stage of this round
    // fills the other one, whose counter is reset here
    const uint32_t* list = c.unc ? ((round & 1u) ? c.unc2 : c.unc) : nullptr;
    const uint32_t P = c.unc ? ((round & 1u) ? c.cnt->n_unc2 : c.cnt->n_unc) : c.cnt->n_pairs;
    if (c.unc && tid == 0) *((round & 1u) ? &c.cnt->n_unc : &c.cnt->n_unc2) = 0u;
    for (uint32_t i = tid; i < P; i += nth) {
        const uint32_t k = list ? list[i] : i;
        if (c.pcol[k] == UNCOLOURED) {
            uint2 p = c.pairs_raw[k];
            colour_bid_one(c, claim, p.x, p.y,
                           c.pprio ? c.pprio[k] : sfp::pair_priority(ext_id(c, p.x), ext_id(c, p.y)));
        }
    }
    for (uint32_t k = tid; k < c.nj; k += nth)
        if (c.jcol[k] == UNCOLOURED) {
            uint2 p = c.jbod_in[k];
            colour_bid_one(c, claim, p.x, p.y, sfp::joint_priority(k));
        }
}

// returns the colour taken, or UNCOLOURED
__device__ __forceinline__ uint32_t colour_take_one(const Ctx& c, const unsigned long long* claim,
                                                    unsigned long long* next_claim, uint32_t a,
                                                    uint32_t b, unsigned long long key,
                                                    uint32_t round, ColourCount* n) {
    bool da = c.meta[a].w != 0, db = c.meta[b].w != 0;
    bool win = (!da || claim[a] == key) && (!db || claim[b] == key);
    if (!win) {
        if (da) next_claim[a] = SFP_NO_CLAIM;
        if (db) next_claim[b] = SFP_NO_CLAIM;
        c.cnt->round_flags[round] = 1u;
        return UNCOLOURED;
    }
    const bool inter = c.tile_b != 0u && group_of(c, a, b) != c.n_tiles;
    unsigned long long* mk = inter ? c.mask_i : c.mask;
    unsigned long long m = (da ? mk[a] : 0ull) | (db ? mk[b] : 0ull);
    int col = sfp::pick_colour(m);
    if (col != SFP_SERIAL_COLOUR) {
        if (da) mk[a] |= 1ull << col;
        if (db) mk[b] |= 1ull << col;
        if (inter) n->inter = max(n->inter, (uint32_t)col + 1u);
        else n->bnd = max(n->bnd, (uint32_t)col + 1u);
        return (uint32_t)col + (inter ? SFP_CLASS_INTERIOR : 0u);
    }
    // every colour of the class is taken on these bodies: the serial group (one for both classes)
    n->bnd = max(n->bnd, (uint32_t)SFP_SERIAL_COLOUR + 1u);
    return (uint32_t)SFP_SERIAL_COLOUR;
}

__device__ __forceinline__ void stage_colour_take(const Ctx& c, uint32_t round, uint32_t tid,
                                                  uint32_t nth) {
    unsigned long long* claim = (round & 1u) ? c.claim1 : c.claim0;
    unsigned long long* next_claim = (round & 1u) ? c.claim0 : c.claim1;
    // With lists: the pairs this round leaves uncoloured go on the OTHER list, which the next
    // round reads — a colouring from scratch takes ~33 rounds for a 100 k-body pile and used to
    // walk all 340 k pairs in every one of them (0.86 ms every SFP_RECOLOUR_PERIOD-th frame; 0.67 now).
    // The order of a list means nothing: a round treats every pair on it independently.
    const uint32_t* list = c.unc ? ((round & 1u) ? c.unc2 : c.unc) : nullptr;
    uint32_t* next_list = (round & 1u) ? c.unc : c.unc2;
    uint32_t* next_n = (round & 1u) ? &c.cnt->n_unc : &c.cnt->n_unc2;
    const uint32_t P = c.unc ? ((round & 1u) ? c.cnt->n_unc2 : c.cnt->n_unc) : c.cnt->n_pairs;
    const uint32_t lane = threadIdx.x & 31u;
    ColourCount ncol;
    // (whole warps walk the loop together: the append is warp-aggregated)
    for (uint32_t base = tid - lane; base < P; base += nth) {
        const uint32_t i = base + lane;
        bool left = false;
        uint32_t k = 0u;
        if (i < P) {
            k = list ? list[i] : i;
            if (c.pcol[k] == UNCOLOURED) {
                uint2 p = c.pairs_raw[k];
                uint32_t col = colour_take_one(
                    c, claim, next_claim, p.x, p.y,
                    c.pprio ? c.pprio[k] : sfp::pair_priority(ext_id(c, p.x), ext_id(c, p.y)), round, &ncol);
                // sort key: colour in the high bits, narrowphase class in the low ones (coherence
                // inside a colour bucket)
                left = col == UNCOLOURED;
                if (!left) col = pair_sort_key(c, col, p);
                c.pcol[k] = col;
            }
        }
        if (list) {
            const uint32_t m = __ballot_sync(0xffffffffu, left);
            if (m) {
                uint32_t at = 0;
                if (lane == 0) at = atomicAdd(next_n, (uint32_t)__popc(m));
                at = __shfl_sync(0xffffffffu, at, 0);
                if (left) next_list[at + __popc(m & ((1u << lane) - 1u))] = k;
            }
        }
    }
    for (uint32_t k = tid; k < c.nj; k += nth)
        if (c.jcol[k] == UNCOLOURED) {
            uint2 p = c.jbod_in[k];
            const uint32_t enc =
                colour_take_one(c, claim, next_claim, p.x, p.y, sfp::joint_priority(k), round, &ncol);
            c.jcol[k] = enc == UNCOLOURED ? UNCOLOURED : joint_sort_key(c, enc, p);
        }
    colour_count_publish(c, ncol);
    if (tid == 0) c.cnt->colour_rounds = round + 1u;
}

// after the stable sort by (group, colour): gather the constraints into that order and find where
// every (group, colour) starts: one binary search per table entry over the sorted keys
__device__ __forceinline__ uint32_t pair_key_dense(uint32_t key) {
    return pair_key_group(key) * (uint32_t)SFP_MAX_COLOURS + pair_key_colour(key);
}
__device__ __forceinline__ uint32_t joint_key_dense(uint32_t key) {
    return (key >> PAIR_KEY_GROUP_SHIFT) * (uint32_t)SFP_MAX_COLOURS + (key & PAIR_KEY_COLOUR_MASK);
}
template <bool JOINTS>
__device__ __forceinline__ void group_offsets(const uint32_t* sorted, uint32_t n, uint32_t n_tiles,
                                              uint32_t* goff, uint32_t* boff, uint32_t tid,
                                              uint32_t nth) {
    const uint32_t kmax = (n_tiles + 1u) * (uint32_t)SFP_MAX_COLOURS, bbase = n_tiles * (uint32_t)SFP_MAX_COLOURS;
    for (uint32_t K = goff ? tid : bbase + tid; K <= kmax; K += nth) {
        uint32_t lo = 0, hi = n;  // first i with dense(sorted[i]) >= K
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t d = JOINTS ? joint_key_dense(sorted[mid]) : pair_key_dense(sorted[mid]);
            if (d < K) lo = mid + 1u;
            else hi = mid;
        }
        if (goff) goff[K] = lo;
        if (K >= bbase) boff[K - bbase] = lo;
    }
}

__device__ __forceinline__ void stage_gather(const Ctx& c, uint32_t tid, uint32_t nth) {
    const uint32_t P = c.cnt->n_pairs;
    for (uint32_t i = tid; i < P; i += nth) {
        const uint2 p = c.pairs_raw[c.pperm[i]];
        c.pairs[i] = p;
        if (c.psa) {  // the narrowphase's per-frame constants of this pair
            const f4 sa = ld4(c.shape, p.x), sb = ld4(c.shape, p.y);
            float mu_s, mu_d, e;
            sfp::pair_material(ld4(c.mat, p.x), ld4(c.mat, p.y), &mu_s, &mu_d, &e);
            st4(c.psa, i, sa);
            st4(c.psb, i, sb);
            c.pmt[i] = make_float4(mu_s, mu_d, e, sfp::narrowphase_reach(sa, sb, c.contact_margin));
        }
    }
    group_offsets<false>(c.pcol_sorted, P, c.n_tiles, c.goff, c.cnt->pair_col_off, tid, nth);
    for (uint32_t i = tid; i < c.nj; i += nth) {
        uint32_t s = c.jperm[i];
        c.jbod[i] = c.jbod_in[s];
        c.janc[i] = c.janc_in[s];
        c.jpar[i] = c.jpar_in[s];
    }
    if (c.nj) {
        group_offsets<true>(c.jcol_sorted, c.nj, c.n_tiles, c.jgoff, c.cnt->joint_col_off, tid, nth);
    } else {
        const uint32_t kmax = (c.n_tiles + 1u) * (uint32_t)SFP_MAX_COLOURS;
        if (c.jgoff)
            for (uint32_t K = tid; K <= kmax; K += nth) c.jgoff[K] = 0u;
        for (uint32_t K = tid; K <= (uint32_t)SFP_MAX_COLOURS; K += nth) c.cnt->joint_col_off[K] = 0u;
    }
}

// ---------------------------------------------------------------------------------------
// A3 integrate
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_integrate(const Ctx& c, float h, uint32_t tid, uint32_t nth) {
    for (uint32_t i = tid; i < c.n; i += nth) {
        f4 pose = ld4(c.pose, i), vel = ld4(c.vel, i);
        f2 mi = ld2(c.minv, i);
        f4 prev;
        sfp::integrate_body(&pose, &prev, &vel, mi, c.gx, c.gy, h);
        st4(c.prev, i, prev);
        st4(c.pose, i, pose);
        st4(c.vel, i, vel);
        if (c.done) reinterpret_cast<ulonglong2*>(c.done)[i] = make_ulonglong2(0ull, 0ull);
    }
    if (tid == 0) {
        c.cnt->n_contacts = 0;
        c.cnt->n_manifolds = 0;
        c.cnt->maxpen_bits = 0;
    }
}

// ---------------------------------------------------------------------------------------
// A4 narrowphase: one thread per candidate pair (colour-sorted order)
// ---------------------------------------------------------------------------------------
// MARK (dataflow solve only): a pair without a contact in this substep is a no-op in both solver
// passes, so it reports itself done on its dynamic bodies here and nobody ever waits for it.
__device__ __forceinline__ void red_or_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("red.relaxed.gpu.global.or.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void flow_mark_noop(const Ctx& c, uint32_t k, uint2 p, bool pos_too) {
    const uint32_t key = c.pcol_sorted[k];
    const uint32_t col = pair_key_colour(key);
    if (col == SFP_SERIAL_COLOUR) return;
    const unsigned long long bit = 1ull << col;
    if (!(key & PAIR_KEY_STATIC_A)) {
        if (pos_too) red_or_u64(&c.done[2u * p.x], bit);
        red_or_u64(&c.done[2u * p.x + 1u], bit);
    }
    if (!(key & PAIR_KEY_STATIC_B)) {
        if (pos_too) red_or_u64(&c.done[2u * p.y], bit);
        red_or_u64(&c.done[2u * p.y + 1u], bit);
    }
}

// ORDERED: pair positions are visited through c.np_order (a permutation of 0..P-1 that groups
// the pairs by narrowphase class), so that the pairs of a warp run the same routine even when
// the (colour, class) buckets are only a few dozen pairs long; which thread handles which pair
// changes nothing in the records.
template <bool POLYS, bool MARK = false, bool STAGED = false, bool ORDERED = false>
__device__ __forceinline__ void stage_narrowphase(const Ctx& c, uint32_t tid, uint32_t nth) {
    // Per block-sized chunk of pairs: (1) every thread tests its pair's bounding circles; a pair
    // that is out of reach gets its empty record at once.  (2) The SAT for the pairs in reach.
    // About a third of a settled pile's candidate pairs are out of reach in a given substep.
    // When a thread has many pairs to go through (big worlds; a batched world's CTA), the pairs
    // in reach are first compacted through a shared-memory list, so that step (2) runs in
    // whole warps less instead of the same warps with a third of their lanes idle (4 M bodies:
    // 11.4 -> 9.3 ms per frame; batched worlds: 2.74 -> 2.60 ms).  With only a few pairs per
    // thread the step is a latency chain per chunk and the two extra barriers cost more than
    // the idle lanes (100 k bodies: 205 -> 227 us), so there each thread keeps its own pair.
    // (The order of the list is whatever the shared-memory atomic gives: each pair is still
    // handled exactly once and writes only its own record.)
    __shared__ uint32_t s_list[512];  // one slot per thread of the block (blocks are <= 512 wide)
    __shared__ uint32_t s_n;
    const uint32_t P = c.cnt->n_pairs;
    const bool compact = P >= 6u * nth;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t NONE = 0xffffffffu;
    uint32_t npts = 0, nman = 0;
    float maxpen = 0.0f;
    const uint32_t* order = ORDERED ? c.np_order : nullptr;
    for (uint32_t base = tid - threadIdx.x; base < P; base += nth) {
        uint32_t k = base + threadIdx.x;
        if (ORDERED && order != nullptr && k < P) k = order[k];
        bool near = false;
        if (compact) {
            if (threadIdx.x == 0) s_n = 0u;
            __syncthreads();
        }
        if (k < P) {
            uint2 p = c.pairs[k];
            if (STAGED)
                near = !sfp::narrowphase_far_reach(ld4(c.pose, p.x), ld4(c.pose, p.y), c.pmt[k].w);
            else
                near = !sfp::narrowphase_far(ld4(c.pose, p.x), ld4(c.shape, p.x), ld4(c.pose, p.y),
                                             ld4(c.shape, p.y), c.contact_margin);
            if (!near) {
                c.cb[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                if (MARK) flow_mark_noop(c, k, p, true);
            }
        }
        uint32_t k2 = near ? k : NONE;
        if (compact) {
            const uint32_t hit = __ballot_sync(0xffffffffu, near);
            if (hit) {
                uint32_t at = 0;
                if (lane == 0) at = atomicAdd(&s_n, (uint32_t)__popc(hit));
                at = __shfl_sync(0xffffffffu, at, 0);
                if (near) s_list[at + __popc(hit & ((1u << lane) - 1u))] = k;
            }
            __syncthreads();
            k2 = (threadIdx.x < s_n) ? s_list[threadIdx.x] : NONE;
        }
        if (k2 != NONE) {
            uint2 p = c.pairs[k2];
            sfp::ContactRec r;
            if (STAGED) {
                const float4 m = c.pmt[k2];
                r = sfp::narrowphase_near_mat<POLYS>(
                    ld4(c.pose, p.x), ld4(c.vel, p.x), ld4(c.psa, k2), ld4(c.pose, p.y), ld4(c.vel, p.y),
                    ld4(c.psb, k2), m.x, m.y, m.z, c.contact_margin,
                    reinterpret_cast<const sfp::f4*>(c.polytab));
            } else {
                r = sfp::narrowphase_near<POLYS>(
                    ld4(c.pose, p.x), ld4(c.vel, p.x), ld4(c.shape, p.x), ld4(c.mat, p.x), ld4(c.pose, p.y),
                    ld4(c.vel, p.y), ld4(c.shape, p.y), ld4(c.mat, p.y), c.contact_margin,
                    reinterpret_cast<const sfp::f4*>(c.polytab));
            }
            uint32_t cnt = sfp::f2u(r.b.y);
            st4(c.cb, k2, r.b);
            if (MARK && cnt == 0u) flow_mark_noop(c, k2, p, true);
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
        if (compact) __syncthreads();
    }
    // warp-aggregated integer statistics (order-independent)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        npts += __shfl_xor_sync(0xffffffffu, npts, d);
        nman += __shfl_xor_sync(0xffffffffu, nman, d);
        maxpen = fmaxf(maxpen, __shfl_xor_sync(0xffffffffu, maxpen, d));
    }
    if ((threadIdx.x & 31) == 0 && (npts | nman)) {
        atomicAdd(&c.cnt->n_contacts, npts);
        atomicAdd(&c.cnt->n_manifolds, nman);
        atomicMax(&c.cnt->maxpen_bits, sfp::f2u(maxpen));
    }
}

// ---------------------------------------------------------------------------------------
// A5 + A6 position projection of one colour
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ sfp::BodyPos load_pos(const Ctx& c, uint32_t i) {
    f4 pose = ld4(c.pose, i), prev = ld4(c.prev, i);
    f2 mi = ld2(c.minv, i);
    sfp::BodyPos b;
    b.x = sfp::mk(pose.x, pose.y);
    b.q = sfp::mk(pose.z, pose.w);
    b.d = sfp::mk(prev.x, prev.y);
    b.qp = sfp::mk(prev.z, prev.w);
    b.im = mi.x;
    b.ii = mi.y;
    return b;
}
__device__ __forceinline__ void store_pos(const Ctx& c, uint32_t i, const sfp::BodyPos& b) {
    if (b.im > 0.0f || b.ii > 0.0f) {
        c.pose[i] = make_float4(b.x.x, b.x.y, b.q.x, b.q.y);
        c.prev[i] = make_float4(b.d.x, b.d.y, b.qp.x, b.qp.y);
    }
}

// what the solver reads of one candidate pair
struct PairRec {
    uint2 p;
    f4 ca, cb, q0, q1;
};
__device__ __forceinline__ PairRec load_pair_rec(const Ctx& c, uint32_t k) {
    PairRec r;
    r.p = c.pairs[k];
    r.cb = ld4(c.cb, k);
    r.ca = ld4(c.ca, k);
    r.q0 = ld4(c.cp0, k);
    r.q1 = ld4(c.cp1, k);
    return r;
}

__device__ __forceinline__ void solve_pos_pair(const Ctx& c, uint32_t k, const PairRec& r) {
    uint32_t cnt = sfp::f2u(r.cb.y);
    f4 lam;
    lam.x = lam.y = lam.z = lam.w = 0.0f;
    if (cnt > 0) {
        sfp::BodyPos A = load_pos(c, r.p.x), B = load_pos(c, r.p.y);
        lam = sfp::solve_contact_manifold(&A, &B, sfp::mk(r.ca.x, r.ca.y), cnt,
                                          sfp::mk(r.q0.x, r.q0.y), sfp::mk(r.q0.z, r.q0.w),
                                          sfp::mk(r.q1.x, r.q1.y), sfp::mk(r.q1.z, r.q1.w), r.ca.z);
        store_pos(c, r.p.x, A);
        store_pos(c, r.p.y, B);
    }
    st4(c.lam, k, lam);
}

__device__ __forceinline__ void solve_pos_joint(const Ctx& c, uint32_t k, float inv_h2) {
    uint2 jb = c.jbod[k];
    float4 an = c.janc[k], pr = c.jpar[k];
    sfp::BodyPos A = load_pos(c, jb.x), B = load_pos(c, jb.y);
    sfp::solve_joint(&A, &B, sfp::mk(an.x, an.y), sfp::mk(an.z, an.w), pr.x, pr.y, sfp::f2u(pr.z),
                     inv_h2);
    store_pos(c, jb.x, A);
    store_pos(c, jb.y, B);
}

// items of one colour: its joints first, then its pairs.  `skip_first`: the caller has
// already handled item `tid` (from a prefetched record).
// `joff` / `poff`: the colour offsets (Counters::joint_col_off / pair_col_off, or a copy of them
// that is cheaper to reach: a load from global memory after a grid-wide barrier is an L2 round
// trip, and it would sit in front of every phase)
__device__ __forceinline__ void stage_solve_pos(const Ctx& c, uint32_t colour, float inv_h2,
                                                uint32_t tid, uint32_t nth, bool skip_first,
                                                const uint32_t* joff, const uint32_t* poff) {
    const uint32_t j0 = joff[colour], j1 = joff[colour + 1];
    const uint32_t p0 = poff[colour], p1 = poff[colour + 1];
    const uint32_t nj = j1 - j0, total = nj + (p1 - p0);
    for (uint32_t t = skip_first ? tid + nth : tid; t < total; t += nth) {
        if (t < nj) {
            solve_pos_joint(c, j0 + t, inv_h2);
        } else {
            uint32_t k = p0 + (t - nj);
            // an empty manifold needs only its count
            f4 cb = ld4(c.cb, k);
            if (sfp::f2u(cb.y) == 0u) {
                c.lam[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                continue;
            }
            solve_pos_pair(c, k, load_pair_rec(c, k));
        }
    }
}

__device__ __forceinline__ void stage_solve_pos(const Ctx& c, uint32_t colour, float inv_h2,
                                                uint32_t tid, uint32_t nth, bool skip_first = false) {
    stage_solve_pos(c, colour, inv_h2, tid, nth, skip_first, c.cnt->joint_col_off,
                    c.cnt->pair_col_off);
}

// ---------------------------------------------------------------------------------------
// A7 (i) velocity derivation, (ii) velocity pass of one colour
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_derive(const Ctx& c, float inv_h, uint32_t tid, uint32_t nth) {
    for (uint32_t i = tid; i < c.n; i += nth) {
        f4 v = sfp::derive_velocity(ld4(c.pose, i), ld4(c.prev, i), ld4(c.vel, i), ld2(c.minv, i),
                                    inv_h);
        st4(c.vel, i, v);
    }
}

__device__ __forceinline__ void solve_vel_pair(const Ctx& c, uint32_t k, const PairRec& r, f4 lam,
                                               float h, float rest_threshold) {
    if (!(lam.x > 0.0f || lam.z > 0.0f)) return;
    uint32_t cnt = sfp::f2u(r.cb.y);
    f4 va = ld4(c.vel, r.p.x), vb = ld4(c.vel, r.p.y);
    f4 pa = ld4(c.pose, r.p.x), pb = ld4(c.pose, r.p.y);
    f2 ma = ld2(c.minv, r.p.x), mb = ld2(c.minv, r.p.y);
    sfp::BodyVel A, B;
    A.v = sfp::mk(va.x, va.y);
    A.w = va.z;
    A.im = ma.x;
    A.ii = ma.y;
    B.v = sfp::mk(vb.x, vb.y);
    B.w = vb.z;
    B.im = mb.x;
    B.ii = mb.y;
    sfp::solve_manifold_velocity(&A, &B, sfp::mk(pa.z, pa.w), sfp::mk(pb.z, pb.w),
                                 sfp::mk(r.ca.x, r.ca.y), cnt, sfp::mk(r.q0.x, r.q0.y),
                                 sfp::mk(r.q0.z, r.q0.w), sfp::mk(r.q1.x, r.q1.y),
                                 sfp::mk(r.q1.z, r.q1.w), lam, r.cb.z, r.cb.w, r.ca.z, r.ca.w, r.cb.x,
                                 h, rest_threshold);
    if (ma.x > 0.0f || ma.y > 0.0f) c.vel[r.p.x] = make_float4(A.v.x, A.v.y, A.w, 0.0f);
    if (mb.x > 0.0f || mb.y > 0.0f) c.vel[r.p.y] = make_float4(B.v.x, B.v.y, B.w, 0.0f);
}

__device__ __forceinline__ void stage_solve_vel(const Ctx& c, uint32_t colour, float h,
                                                float rest_threshold, uint32_t tid, uint32_t nth,
                                                bool skip_first, const uint32_t* poff) {
    const uint32_t p0 = poff[colour], p1 = poff[colour + 1];
    for (uint32_t k = p0 + (skip_first ? tid + nth : tid); k < p1; k += nth) {
        f4 lam = ld4(c.lam, k);
        if (!(lam.x > 0.0f || lam.z > 0.0f)) continue;
        solve_vel_pair(c, k, load_pair_rec(c, k), lam, h, rest_threshold);
    }
}

__device__ __forceinline__ void stage_solve_vel(const Ctx& c, uint32_t colour, float h,
                                                float rest_threshold, uint32_t tid, uint32_t nth,
                                                bool skip_first = false) {
    stage_solve_vel(c, colour, h, rest_threshold, tid, nth, skip_first, c.cnt->pair_col_off);
}

// ---------------------------------------------------------------------------------------
// Tile-resident solve (sfp_math.h, "Tiles"): ONE thread block holds a tile's bodies in shared
// memory and runs all interior colours of a solver pass between __syncthreads — a colour costs a
// shared-memory gather, the solve and a block barrier instead of an L2 gather / scatter and a
// grid-wide barrier.  Interior constraints only touch the tile's own bodies and static ones
// (read from global memory, never written), so tiles need nothing from each other.
// ---------------------------------------------------------------------------------------
struct TileView {
    float4 *pose, *aux;  // aux: `prev` in the position pass, `vel` in the velocity pass
    float2* minv;
    unsigned long long *mask, *done;  // interior colours a body uses / has had applied this pass
    uint32_t base, count;  // the tile is the bodies [base, base + count)
};
constexpr uint32_t TILE_SMEM_PER_BODY = 16u + 16u + 8u + 8u + 8u;

__device__ __forceinline__ sfp::BodyPos tile_load_pos(const Ctx& c, const TileView& tv, uint32_t i) {
    const uint32_t l = i - tv.base;
    float4 pose, prev;
    float2 mi;
    if (l < tv.count) {
        pose = tv.pose[l];
        prev = tv.aux[l];
        mi = tv.minv[l];
    } else {  // a static body
        pose = c.pose[i];
        prev = c.prev[i];
        mi = c.minv[i];
    }
    sfp::BodyPos b;
    b.x = sfp::mk(pose.x, pose.y);
    b.q = sfp::mk(pose.z, pose.w);
    b.d = sfp::mk(prev.x, prev.y);
    b.qp = sfp::mk(prev.z, prev.w);
    b.im = mi.x;
    b.ii = mi.y;
    return b;
}
__device__ __forceinline__ void tile_store_pos(const TileView& tv, uint32_t i, const sfp::BodyPos& b) {
    const uint32_t l = i - tv.base;
    if (l < tv.count && (b.im > 0.0f || b.ii > 0.0f)) {
        tv.pose[l] = make_float4(b.x.x, b.x.y, b.q.x, b.q.y);
        tv.aux[l] = make_float4(b.d.x, b.d.y, b.qp.x, b.qp.y);
    }
}

// Inside a tile the colours are NOT separated by barriers: every body of the tile has a bit mask
// in shared memory of the interior colours already applied to it in this pass (TileView::done),
// and a constraint of colour c waits until the masks of its two bodies hold every interior colour
// those bodies use below c (TileView::mask, the colouring's per-body masks).  The per-body order
// is the colour order, so the result is that of a barrier after every colour; but a hop of a
// dependency chain now costs a shared-memory round trip plus the solve, and the warps of the
// block work on many colours at once.  Items of a colour are cut into runs of 32 (padded, so a
// run never straddles two colours and its lanes never depend on each other); runs are numbered
// in colour order and dealt round-robin to the block's warps, each warp takes its runs in
// ascending order: a run depends only on lower-numbered runs and every warp of the block is
// resident, so the waiting cannot deadlock.  A pair without a contact (or without a normal
// impulse, in the velocity pass) reports itself done without waiting.
#ifndef SFP_TILE_SPIN_LIMIT
#define SFP_TILE_SPIN_LIMIT (1u << 22)
#endif
__device__ __forceinline__ void tile_wait(const Ctx& c, const TileView& tv, uint32_t la,
                                          unsigned long long need_a, uint32_t lb,
                                          unsigned long long need_b) {
    if ((need_a | need_b) == 0ull) return;
    volatile unsigned long long* done = tv.done;
    uint32_t spins = 0;
    for (;;) {
        const unsigned long long da = need_a ? done[la] : 0ull, db = need_b ? done[lb] : 0ull;
        if ((da & need_a) == need_a && (db & need_b) == need_b) break;
        __nanosleep(20);
        if (++spins >= SFP_TILE_SPIN_LIMIT) {  // a bug, not a data condition: give up loudly
            atomicOr(&c.cnt->overflow, 32u);
            break;
        }
    }
    __threadfence_block();
}
__device__ __forceinline__ void tile_mark(const TileView& tv, uint32_t l, unsigned long long bit) {
    if (l < tv.count) atomicOr(&tv.done[l], bit);
}
// colour of run r (cursor `col` only moves forward); s_rb[col] = first run of colour col
__device__ __forceinline__ uint32_t tile_run_colour(const uint32_t* s_rb, uint32_t r, uint32_t col) {
    while (r >= s_rb[col + 1]) ++col;
    return col;
}

__device__ __forceinline__ void tile_solve_pos(const Ctx& c, const TileView& tv, const uint32_t* s_jo,
                                               const uint32_t* s_po, const uint32_t* s_rb,
                                               float inv_h2) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const uint32_t total = s_rb[SFP_SERIAL_COLOUR];
    uint32_t col = 0;
    for (uint32_t r = warp; r < total; r += nwarps) {
        col = tile_run_colour(s_rb, r, col);
        const unsigned long long bit = 1ull << col, below = bit - 1ull;
        const uint32_t t = (r - s_rb[col]) * 32u + lane;
        const uint32_t j0 = s_jo[col], nj = s_jo[col + 1] - j0;
        const uint32_t p0 = s_po[col], np = s_po[col + 1] - p0;
        if (t < nj) {
            const uint32_t k = j0 + t;
            const uint2 jb = c.jbod[k];
            const float4 an = c.janc[k], pr = c.jpar[k];
            const uint32_t la = jb.x - tv.base, lb = jb.y - tv.base;
            tile_wait(c, tv, la, la < tv.count ? tv.mask[la] & below : 0ull, lb,
                      lb < tv.count ? tv.mask[lb] & below : 0ull);
            sfp::BodyPos A = tile_load_pos(c, tv, jb.x), B = tile_load_pos(c, tv, jb.y);
            sfp::solve_joint(&A, &B, sfp::mk(an.x, an.y), sfp::mk(an.z, an.w), pr.x, pr.y,
                             sfp::f2u(pr.z), inv_h2);
            tile_store_pos(tv, jb.x, A);
            tile_store_pos(tv, jb.y, B);
            __threadfence_block();
            tile_mark(tv, la, bit);
            tile_mark(tv, lb, bit);
        } else if (t - nj < np) {
            const uint32_t k = p0 + (t - nj);
#ifdef SFP_TILE_PROBE
            const long long q0_ = clock64();
#endif
            const PairRec r_ = load_pair_rec(c, k);
            const uint32_t cnt = sfp::f2u(r_.cb.y);
            const uint32_t la = r_.p.x - tv.base, lb = r_.p.y - tv.base;
            f4 lam;
            lam.x = lam.y = lam.z = lam.w = 0.0f;
            if (cnt != 0u) {
#ifdef SFP_TILE_PROBE
                const long long q1_ = clock64();
#endif
                tile_wait(c, tv, la, la < tv.count ? tv.mask[la] & below : 0ull, lb,
                          lb < tv.count ? tv.mask[lb] & below : 0ull);
#ifdef SFP_TILE_PROBE
                const long long q2_ = clock64();
#endif
                sfp::BodyPos A = tile_load_pos(c, tv, r_.p.x), B = tile_load_pos(c, tv, r_.p.y);
                lam = sfp::solve_contact_manifold(&A, &B, sfp::mk(r_.ca.x, r_.ca.y), cnt,
                                                  sfp::mk(r_.q0.x, r_.q0.y), sfp::mk(r_.q0.z, r_.q0.w),
                                                  sfp::mk(r_.q1.x, r_.q1.y), sfp::mk(r_.q1.z, r_.q1.w),
                                                  r_.ca.z);
                tile_store_pos(tv, r_.p.x, A);
                tile_store_pos(tv, r_.p.y, B);
                __threadfence_block();
#ifdef SFP_TILE_PROBE
                // development probe: per LANE cycles for record load / wait / solve, sampled
                if (blockIdx.x == 3 && (k & 7u) == 0u) {
                    const long long q3_ = clock64();
                    uint32_t* f = c.cnt->round_flags;
                    atomicAdd(&f[480], (uint32_t)(q1_ - q0_) >> 4);
                    atomicAdd(&f[481], (uint32_t)(q2_ - q1_) >> 4);
                    atomicAdd(&f[482], (uint32_t)(q3_ - q2_) >> 4);
                    atomicMax(&f[484], (uint32_t)(q1_ - q0_));
                    atomicMax(&f[485], (uint32_t)(q2_ - q1_));
                    atomicMax(&f[486], (uint32_t)(q3_ - q2_));
                    atomicAdd(&f[488], 1u);
                    atomicAdd(&f[col], (uint32_t)(q2_ - q1_) >> 4);
                    atomicAdd(&f[64 + col], 1u);
                    atomicAdd(&f[128 + col], (uint32_t)(q3_ - q2_) >> 4);
                }
#endif
            }
            st4(c.lam, k, lam);
            tile_mark(tv, la, bit);
            tile_mark(tv, lb, bit);
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void tile_solve_vel(const Ctx& c, const TileView& tv, const uint32_t* s_jo,
                                               const uint32_t* s_po, const uint32_t* s_rb, float h,
                                               float rest_threshold) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const uint32_t total = s_rb[SFP_SERIAL_COLOUR];
    uint32_t col = 0;
    for (uint32_t r = warp; r < total; r += nwarps) {
        col = tile_run_colour(s_rb, r, col);
        const unsigned long long bit = 1ull << col, below = bit - 1ull;
        const uint32_t t = (r - s_rb[col]) * 32u + lane;
        const uint32_t j0 = s_jo[col], nj = s_jo[col + 1] - j0;
        const uint32_t p0 = s_po[col], np = s_po[col + 1] - p0;
        if (t < nj) {
            // joints have no velocity pass
            const uint2 jb = c.jbod[j0 + t];
            tile_mark(tv, jb.x - tv.base, bit);
            tile_mark(tv, jb.y - tv.base, bit);
        } else if (t - nj < np) {
            const uint32_t k = p0 + (t - nj);
            const f4 lam = ld4(c.lam, k);
            const uint2 p = c.pairs[k];
            const uint32_t la = p.x - tv.base, lb = p.y - tv.base;
            if (lam.x > 0.0f || lam.z > 0.0f) {
                const PairRec r_ = load_pair_rec(c, k);
                tile_wait(c, tv, la, la < tv.count ? tv.mask[la] & below : 0ull, lb,
                          lb < tv.count ? tv.mask[lb] & below : 0ull);
                const bool ina = la < tv.count, inb = lb < tv.count;
                const float4 va = ina ? tv.aux[la] : c.vel[p.x], vb = inb ? tv.aux[lb] : c.vel[p.y];
                const float4 pa = ina ? tv.pose[la] : c.pose[p.x], pb = inb ? tv.pose[lb] : c.pose[p.y];
                const float2 ma = ina ? tv.minv[la] : c.minv[p.x], mb = inb ? tv.minv[lb] : c.minv[p.y];
                sfp::BodyVel A, B;
                A.v = sfp::mk(va.x, va.y);
                A.w = va.z;
                A.im = ma.x;
                A.ii = ma.y;
                B.v = sfp::mk(vb.x, vb.y);
                B.w = vb.z;
                B.im = mb.x;
                B.ii = mb.y;
                sfp::solve_manifold_velocity(
                    &A, &B, sfp::mk(pa.z, pa.w), sfp::mk(pb.z, pb.w), sfp::mk(r_.ca.x, r_.ca.y),
                    sfp::f2u(r_.cb.y), sfp::mk(r_.q0.x, r_.q0.y), sfp::mk(r_.q0.z, r_.q0.w),
                    sfp::mk(r_.q1.x, r_.q1.y), sfp::mk(r_.q1.z, r_.q1.w), lam, r_.cb.z, r_.cb.w, r_.ca.z,
                    r_.ca.w, r_.cb.x, h, rest_threshold);
                if (ina && (ma.x > 0.0f || ma.y > 0.0f)) tv.aux[la] = make_float4(A.v.x, A.v.y, A.w, 0.0f);
                if (inb && (mb.x > 0.0f || mb.y > 0.0f)) tv.aux[lb] = make_float4(B.v.x, B.v.y, B.w, 0.0f);
                __threadfence_block();
            }
            tile_mark(tv, la, bit);
            tile_mark(tv, lb, bit);
        }
    }
    __syncthreads();
}

// one pass over the tiles of this block: load the tile, solve its interior colours, store it.
// smem: tile_b x TILE_SMEM_PER_BODY bytes; s_off: 3 x 65 words.  Every thread of the block must
// call it.
template <bool VEL>
__device__ __forceinline__ void stage_tiles(const Ctx& c, float4* smem, uint32_t* s_off, float h,
                                            float rest_threshold) {
    TileView tv;
    tv.pose = smem;
    tv.aux = smem + c.tile_b;
    tv.minv = reinterpret_cast<float2*>(smem + 2u * (size_t)c.tile_b);
    tv.mask = reinterpret_cast<unsigned long long*>(tv.minv + c.tile_b);
    tv.done = tv.mask + c.tile_b;
    float4* g_aux = VEL ? c.vel : c.prev;
    const float inv_h = 1.0f / h, inv_h2 = inv_h * inv_h;
    uint32_t* s_jo = s_off;
    uint32_t* s_po = s_off + (SFP_MAX_COLOURS + 1);
    uint32_t* s_rb = s_off + 2 * (SFP_MAX_COLOURS + 1);
    for (uint32_t t = blockIdx.x; t < c.n_tiles; t += gridDim.x) {
        tv.base = t * c.tile_b;
        tv.count = min(c.tile_b, c.n_dyn - tv.base);
        for (uint32_t i = threadIdx.x; i <= (uint32_t)SFP_MAX_COLOURS; i += blockDim.x) {
            s_po[i] = c.goff[(size_t)t * SFP_MAX_COLOURS + i];
            s_jo[i] = c.jgoff[(size_t)t * SFP_MAX_COLOURS + i];
        }
        __syncthreads();
        // (a tile without interior constraints has nothing to do)
        if (s_po[0] == s_po[SFP_MAX_COLOURS] && s_jo[0] == s_jo[SFP_MAX_COLOURS]) {
            __syncthreads();
            continue;
        }
        if (threadIdx.x == 0) {
            uint32_t runs = 0;
            for (uint32_t col = 0; col < (uint32_t)SFP_SERIAL_COLOUR; ++col) {
                s_rb[col] = runs;
                runs += ((s_jo[col + 1] - s_jo[col]) + (s_po[col + 1] - s_po[col]) + 31u) / 32u;
            }
            s_rb[SFP_SERIAL_COLOUR] = runs;
        }
        for (uint32_t i = threadIdx.x; i < tv.count; i += blockDim.x) {
            tv.pose[i] = c.pose[tv.base + i];
            tv.aux[i] = g_aux[tv.base + i];
            tv.minv[i] = c.minv[tv.base + i];
            tv.mask[i] = c.mask_i[tv.base + i];
            tv.done[i] = 0ull;
        }
        __syncthreads();
        if (VEL) tile_solve_vel(c, tv, s_jo, s_po, s_rb, h, rest_threshold);
        else tile_solve_pos(c, tv, s_jo, s_po, s_rb, inv_h2);
        for (uint32_t i = threadIdx.x; i < tv.count; i += blockDim.x) {
            if (!VEL) c.pose[tv.base + i] = tv.pose[i];
            g_aux[tv.base + i] = tv.aux[i];
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// Dataflow solve (k_substeps_flow).  Coloured Gauss-Seidel needs, for every body, that its
// constraints are applied one after another in the solve order of their colours — nothing
// more.  A grid-wide barrier after every colour enforces far more than that (264 barriers a
// frame at ~3 us each on a 100 k-body pile: most of the frame).  Here every body carries a
// bit mask of the colours already applied to it in the current pass (Ctx::done).  A constraint
// of colour c waits until the masks of its two bodies hold every colour those bodies use that
// comes before c in the solve order (Ctx::mask, left by the colouring, says which those are),
// reads the bodies, solves, writes them back and sets its bit: release fence + RED on the
// writer's side, relaxed polling + acquire fence on the reader's.  The whole pass is one
// grid-wide phase; the per-body order — and therefore every bit of the result — is that of
// the barrier version and of the oracle.
//
// Work distribution: a colour's items (joints, then pairs) are cut into runs of 32, runs are
// numbered in solve order and dealt round-robin to the grid's warps, each warp takes its runs
// in ascending order.  A run depends only on runs with lower numbers, every warp is resident
// (cooperative launch) and the lanes of a run never depend on each other (one colour), so the
// waiting cannot deadlock.  A wait that does not end (a bug, not a data condition) gives up
// after ~1 s and raises overflow bit 32.
// ---------------------------------------------------------------------------------------
struct FlowOrder {
    uint32_t n;                                   // proper colours in use, in solve order
    uint32_t run_base[SFP_MAX_COLOURS + 1];       // first run of the i-th colour of the order
    unsigned long long before[SFP_MAX_COLOURS];   // colours that precede it
    uint8_t colour[SFP_MAX_COLOURS];
};

// by ONE thread; joff / poff are the colour offsets of joints and pairs
__device__ __forceinline__ void flow_build_order(const Ctx& c, uint32_t ncol, const uint32_t* joff,
                                                 const uint32_t* poff, FlowOrder* fo) {
    const unsigned long long all = (ncol >= 64u) ? ~0ull : ((1ull << ncol) - 1ull);
    const unsigned long long big = big_colour_mask(c, ncol);
    const unsigned long long small = all & ~big & ~(1ull << SFP_SERIAL_COLOUR);
    uint32_t n = 0, runs = 0;
    unsigned long long acc = 0ull;
    for (int pass = 0; pass < 2; ++pass) {
        const unsigned long long m = pass ? small : big;
        for (uint32_t col = 0; col < (uint32_t)SFP_SERIAL_COLOUR; ++col) {
            if (!((m >> col) & 1ull)) continue;
            const uint32_t items = (joff[col + 1] - joff[col]) + (poff[col + 1] - poff[col]);
            fo->colour[n] = (uint8_t)col;
            fo->before[n] = acc;
            fo->run_base[n] = runs;
            runs += (items + 31u) / 32u;
            acc |= 1ull << col;
            ++n;
        }
    }
    fo->run_base[n] = runs;
    fo->n = n;
}

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ f4 ldcg4(const float4* p, uint32_t i) {
    float4 v = __ldcg(p + i);
    f4 r;
    r.x = v.x, r.y = v.y, r.z = v.z, r.w = v.w;
    return r;
}

#ifndef SFP_FLOW_SLEEP
#define SFP_FLOW_SLEEP 40
#endif
#ifndef SFP_FLOW_SPIN_LIMIT
#define SFP_FLOW_SPIN_LIMIT (1u << 21)
#endif
// wait until slot_a holds need_a and slot_b holds need_b
__device__ __forceinline__ void flow_wait(const Ctx& c, const unsigned long long* slot_a,
                                          unsigned long long need_a,
                                          const unsigned long long* slot_b,
                                          unsigned long long need_b) {
    if ((need_a | need_b) == 0ull) return;  // the bodies were last written before a grid barrier
    uint32_t spins = 0;
    for (;;) {
        const unsigned long long da = need_a ? ld_relaxed_u64(slot_a) : 0ull;
        const unsigned long long db = need_b ? ld_relaxed_u64(slot_b) : 0ull;
        if ((da & need_a) == need_a && (db & need_b) == need_b) break;
#if SFP_FLOW_SLEEP > 0
        __nanosleep(SFP_FLOW_SLEEP);
#endif
        ++spins;
        if ((spins & 1023u) == 0u &&
            (spins >= SFP_FLOW_SPIN_LIMIT ||
             (*reinterpret_cast<volatile uint32_t*>(&c.cnt->overflow) & 32u))) {
            atomicOr(&c.cnt->overflow, 32u);
            break;
        }
    }
#ifndef SFP_FLOW_NOACQ
    fence_acq_rel_gpu();
#endif
}

__device__ __forceinline__ sfp::BodyPos load_pos_cg(const Ctx& c, uint32_t i) {
    f4 pose = ldcg4(c.pose, i), prev = ldcg4(c.prev, i);
    f2 mi = ld2(c.minv, i);
    sfp::BodyPos b;
    b.x = sfp::mk(pose.x, pose.y);
    b.q = sfp::mk(pose.z, pose.w);
    b.d = sfp::mk(prev.x, prev.y);
    b.qp = sfp::mk(prev.z, prev.w);
    b.im = mi.x;
    b.ii = mi.y;
    return b;
}

// one pass over all proper colours; gwarp / nwarps: this warp's index and the warp count of the grid
__device__ __forceinline__ void flow_solve_pos(const Ctx& c, const FlowOrder& fo,
                                               const uint32_t* joff, const uint32_t* poff,
                                               float inv_h2, uint32_t gwarp, uint32_t nwarps,
                                               uint32_t lane) {
    uint32_t oi = 0;
    const uint32_t total = fo.run_base[fo.n];
    for (uint32_t r = gwarp; r < total; r += nwarps) {
        while (r >= fo.run_base[oi + 1]) ++oi;
        const uint32_t col = fo.colour[oi];
        const unsigned long long before = fo.before[oi], bit = 1ull << col;
        const uint32_t t = (r - fo.run_base[oi]) * 32u + lane;
        const uint32_t j0 = joff[col], nj = joff[col + 1] - j0;
        const uint32_t p0 = poff[col], np = poff[col + 1] - p0;
#ifdef SFP_FLOW_PROBE
        // development probe (scripts/flow_probe.py): per run, cycles before the wait (record
        // loads), in the wait, in gather + solve + store, in the release; lane 0 reads the clock
        // with the warp converged, so the readings bracket the slowest lane
        long long pt[5];
        __syncwarp();
        pt[0] = pt[1] = pt[2] = pt[3] = pt[4] = clock64();
#define FLOW_PROBE_AT(i) \
    do {                 \
        pt[i] = clock64(); \
    } while (0)
#else
#define FLOW_PROBE_AT(i) \
    do {                 \
    } while (0)
#endif
        if (t < nj) {
            const uint32_t k = j0 + t;
            const uint2 jb = c.jbod[k];
            const float4 an = c.janc[k], pr = c.jpar[k];
            // (a body's mask holds this joint's colour iff the body is dynamic)
            const unsigned long long ma = c.mask[jb.x], mb = c.mask[jb.y];
            flow_wait(c, &c.done[2u * jb.x], ma & before, &c.done[2u * jb.y], mb & before);
            sfp::BodyPos A = load_pos_cg(c, jb.x), B = load_pos_cg(c, jb.y);
            sfp::solve_joint(&A, &B, sfp::mk(an.x, an.y), sfp::mk(an.z, an.w), pr.x, pr.y,
                             sfp::f2u(pr.z), inv_h2);
            store_pos(c, jb.x, A);
            store_pos(c, jb.y, B);
            fence_acq_rel_gpu();
            // joints have no velocity pass: done there as well
            if (ma & bit) {
                red_or_u64(&c.done[2u * jb.x], bit);
                red_or_u64(&c.done[2u * jb.x + 1u], bit);
            }
            if (mb & bit) {
                red_or_u64(&c.done[2u * jb.y], bit);
                red_or_u64(&c.done[2u * jb.y + 1u], bit);
            }
        } else if (t - nj < np) {
            const uint32_t k = p0 + (t - nj);
            const f4 cb = ld4(c.cb, k);
            const uint32_t cnt = sfp::f2u(cb.y);
            if (cnt == 0u) {
                // (reported done by the narrowphase)
                c.lam[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            } else {
                const uint2 p = c.pairs[k];
                const f4 ca = ld4(c.ca, k), q0 = ld4(c.cp0, k), q1 = ld4(c.cp1, k);
                const unsigned long long ma = c.mask[p.x], mb = c.mask[p.y];
                FLOW_PROBE_AT(1);
                flow_wait(c, &c.done[2u * p.x], ma & before, &c.done[2u * p.y], mb & before);
                FLOW_PROBE_AT(2);
                sfp::BodyPos A = load_pos_cg(c, p.x), B = load_pos_cg(c, p.y);
                const f4 lam = sfp::solve_contact_manifold(
                    &A, &B, sfp::mk(ca.x, ca.y), cnt, sfp::mk(q0.x, q0.y), sfp::mk(q0.z, q0.w),
                    sfp::mk(q1.x, q1.y), sfp::mk(q1.z, q1.w), ca.z);
                store_pos(c, p.x, A);
                store_pos(c, p.y, B);
                st4(c.lam, k, lam);
                FLOW_PROBE_AT(3);
                fence_acq_rel_gpu();
                // a manifold that carried no normal impulse is a no-op in the velocity pass
                const bool vel_too = !(lam.x > 0.0f || lam.z > 0.0f);
                if (ma & bit) {
                    red_or_u64(&c.done[2u * p.x], bit);
                    if (vel_too) red_or_u64(&c.done[2u * p.x + 1u], bit);
                }
                if (mb & bit) {
                    red_or_u64(&c.done[2u * p.y], bit);
                    if (vel_too) red_or_u64(&c.done[2u * p.y + 1u], bit);
                }
                FLOW_PROBE_AT(4);
            }
        }
#ifdef SFP_FLOW_PROBE
        {
            // per lane: the four segments of its own path; reduced to the warp's maxima
            uint32_t seg[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                long long d = pt[i + 1] - pt[i];
                uint32_t v = d > 0 ? (uint32_t)d : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
                seg[i] = v;
            }
            __syncwarp();
            const long long tend = clock64();
            if (lane == 0 && (gwarp & 31u) == 0u) {  // a sample of the warps: the counters are shared
                uint32_t* f = c.cnt->round_flags;
                for (int i = 0; i < 4; ++i) {
                    atomicAdd(&f[480 + i], seg[i] >> 4);
                    atomicMax(&f[484 + i], seg[i]);
                }
                atomicAdd(&f[488], 1u);
                atomicAdd(&f[489], (uint32_t)((tend - pt[0]) >> 4));
                atomicAdd(&f[oi], seg[1] >> 4);        // wait per colour of the order
                atomicAdd(&f[64 + oi], 1u);            // runs per colour
                atomicAdd(&f[128 + oi], (uint32_t)((tend - pt[0]) >> 4));
            }
        }
#endif
    }
#undef FLOW_PROBE_AT
}

__device__ __forceinline__ void flow_solve_vel(const Ctx& c, const FlowOrder& fo,
                                               const uint32_t* joff, const uint32_t* poff, float h,
                                               float rest_threshold, uint32_t gwarp,
                                               uint32_t nwarps, uint32_t lane) {
    uint32_t oi = 0;
    const uint32_t total = fo.run_base[fo.n];
    for (uint32_t r = gwarp; r < total; r += nwarps) {
        while (r >= fo.run_base[oi + 1]) ++oi;
        const uint32_t col = fo.colour[oi];
        const unsigned long long before = fo.before[oi], bit = 1ull << col;
        const uint32_t t = (r - fo.run_base[oi]) * 32u + lane;
        const uint32_t nj = joff[col + 1] - joff[col];
        const uint32_t p0 = poff[col], np = poff[col + 1] - p0;
        if (t < nj || t - nj >= np) continue;
        const uint32_t k = p0 + (t - nj);
        const f4 lam = ld4(c.lam, k);
        if (!(lam.x > 0.0f || lam.z > 0.0f)) continue;  // reported done by the position pass
        const PairRec r_ = load_pair_rec(c, k);
        const uint32_t a = r_.p.x, b = r_.p.y;
        const unsigned long long ma = c.mask[a], mb = c.mask[b];
        flow_wait(c, &c.done[2u * a + 1u], ma & before, &c.done[2u * b + 1u], mb & before);
        const uint32_t cnt = sfp::f2u(r_.cb.y);
        f4 va = ldcg4(c.vel, a), vb = ldcg4(c.vel, b);
        f4 pa = ld4(c.pose, a), pb = ld4(c.pose, b);
        f2 mia = ld2(c.minv, a), mib = ld2(c.minv, b);
        sfp::BodyVel A, B;
        A.v = sfp::mk(va.x, va.y);
        A.w = va.z;
        A.im = mia.x;
        A.ii = mia.y;
        B.v = sfp::mk(vb.x, vb.y);
        B.w = vb.z;
        B.im = mib.x;
        B.ii = mib.y;
        sfp::solve_manifold_velocity(&A, &B, sfp::mk(pa.z, pa.w), sfp::mk(pb.z, pb.w),
                                     sfp::mk(r_.ca.x, r_.ca.y), cnt, sfp::mk(r_.q0.x, r_.q0.y),
                                     sfp::mk(r_.q0.z, r_.q0.w), sfp::mk(r_.q1.x, r_.q1.y),
                                     sfp::mk(r_.q1.z, r_.q1.w), lam, r_.cb.z, r_.cb.w, r_.ca.z,
                                     r_.ca.w, r_.cb.x, h, rest_threshold);
        if (mia.x > 0.0f || mia.y > 0.0f) c.vel[a] = make_float4(A.v.x, A.v.y, A.w, 0.0f);
        if (mib.x > 0.0f || mib.y > 0.0f) c.vel[b] = make_float4(B.v.x, B.v.y, B.w, 0.0f);
        fence_acq_rel_gpu();
        if (ma & bit) red_or_u64(&c.done[2u * a + 1u], bit);
        if (mb & bit) red_or_u64(&c.done[2u * b + 1u], bit);
    }
}

__device__ __forceinline__ void serial_solve_pos(const Ctx& c, uint32_t ncol, float inv_h2) {
    if (ncol <= SFP_SERIAL_COLOUR) return;
    stage_solve_pos(c, SFP_SERIAL_COLOUR, inv_h2, 0u, 1u);  // one thread walks the whole group
}
__device__ __forceinline__ void serial_solve_vel(const Ctx& c, uint32_t ncol, float h,
                                                 float rest_threshold) {
    if (ncol <= SFP_SERIAL_COLOUR) return;
    stage_solve_vel(c, SFP_SERIAL_COLOUR, h, rest_threshold, 0u, 1u);
}

}  // namespace sfpk
