// sfp_world.cu — the C ABI of include/sfp.h for ONE world on one B200 (libsfp_b200.so).
//
// Frame = broadphase (radix-sort spatial hash) -> colouring (cooperative kernel, rounds
// separated by grid.sync) -> stable sort of constraints by colour -> ONE persistent cooperative
// kernel that runs all substeps (integrate / narrowphase / per-colour position solve / derive
// velocity / per-colour velocity solve, separated by grid.sync).  The host queues ~35 launches
// per FRAME and none per substep, never reads anything back while stepping, and takes every
// data-dependent count (entries, pairs, colours) from device memory.
//
// Reference: replaces the reference's CPU `Physics` step (not mounted).
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "sfp.h"
#include "sfp_device.cuh"
#include "sfp_host.h"
#include "sfp_sort.cuh"

namespace cg = cooperative_groups;
using namespace sfpk;

namespace {

constexpr int TPB = 256;
#ifdef SFP_NO_STAGED
constexpr bool NP_STAGED = false;
#else
constexpr bool NP_STAGED = true;
#endif

__device__ __forceinline__ uint32_t gtid() { return blockIdx.x * blockDim.x + threadIdx.x; }
__device__ __forceinline__ uint32_t gnth() { return gridDim.x * blockDim.x; }
__device__ __forceinline__ uint32_t vload(const uint32_t* p) {
    return *reinterpret_cast<const volatile uint32_t*>(p);
}

// ---- one kernel per stage (broadphase always; the rest only with SFP_FLAG_MULTI_KERNEL) ----
__global__ void __launch_bounds__(TPB) k_aabb_count(Ctx c) {
    stage_aabb_count(c, c.fp->dt, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_emit_entries(Ctx c) { stage_emit_entries(c, gtid(), gnth()); }
__global__ void __launch_bounds__(TPB) k_pair_count(Ctx c) {
    stage_pairs<false>(c, c.ekey0, c.eval0, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_pair_write(Ctx c) {
    stage_pairs<true>(c, c.ekey0, c.eval0, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_colour_init(Ctx c) { stage_colour_init(c, gtid(), gnth()); }
__global__ void __launch_bounds__(TPB) k_colour_warm(Ctx c) { stage_colour_warm(c, gtid(), gnth()); }
__global__ void __launch_bounds__(TPB) k_colour_record(Ctx c) {
    stage_colour_record(c, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_colour_bid(Ctx c, uint32_t round) {
    stage_colour_bid(c, round, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_colour_take(Ctx c, uint32_t round) {
    stage_colour_take(c, round, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_gather(Ctx c) { stage_gather(c, gtid(), gnth()); }
__global__ void __launch_bounds__(TPB) k_integrate(Ctx c, float h) {
    stage_integrate(c, h, gtid(), gnth());
}
template <bool POLYS>
__global__ void __launch_bounds__(TPB) k_narrowphase(Ctx c) {
    stage_narrowphase<POLYS, false, NP_STAGED>(c, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_solve_pos(Ctx c, uint32_t colour, float inv_h2) {
    stage_solve_pos(c, colour, inv_h2, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_derive(Ctx c, float inv_h) {
    stage_derive(c, inv_h, gtid(), gnth());
}
__global__ void __launch_bounds__(TPB) k_solve_vel(Ctx c, uint32_t colour, float h, float rt) {
    stage_solve_vel(c, colour, h, rt, gtid(), gnth());
}

__global__ void __launch_bounds__(TPB) k_serial_sort(Ctx c) {
    serial_sort_pairs(c, vload(&c.cnt->n_colours));
}
__global__ void k_serial_pos(Ctx c, float inv_h2) {
    if (threadIdx.x == 0 && blockIdx.x == 0) serial_solve_pos(c, vload(&c.cnt->n_colours), inv_h2);
}
__global__ void k_serial_vel(Ctx c, float h, float rt) {
    if (threadIdx.x == 0 && blockIdx.x == 0) serial_solve_vel(c, vload(&c.cnt->n_colours), h, rt);
}

// ---- re-sort: bodies kept in the order of their Morton cell key (statics last) -----------------
// Bodies that touch are then neighbours in memory — the solver's 16-byte gathers land in shared
// sectors whatever order the scene graph uploaded them in — and a tile of the solve order
// (sfp_math.h, "Tiles") is a contiguous index range.  Runs when the colouring starts from scratch
// (every SFP_RECOLOUR_PERIOD-th frame): the colour table, keyed by internal indices, is not read
// in such a frame.  Scratch: the broadphase entry buffers for the sort, the contact record arrays
// (dead between frames) for the permuted copies.
// (keys are laid out in the CALLER'S order, value = where the body is now: the stable sort then
// breaks ties by the caller's index, as the oracle does, whatever the current internal order)
__global__ void __launch_bounds__(TPB) k_tile_keys(Ctx c, uint32_t* __restrict__ keys,
                                                   uint32_t* __restrict__ vals) {
    for (uint32_t e = gtid(); e < c.n; e += gnth()) {
        const uint32_t i = c.inv[e];
        const float4 p = c.pose[i];
        const float2 m = c.minv[i];
        keys[e] = sfp::tile_sort_key(p.x, p.y, c.inv_cell, m.x > 0.0f || m.y > 0.0f);
        vals[e] = i;
    }
}
__global__ void __launch_bounds__(TPB) k_permute_gather(Ctx c, const uint32_t* __restrict__ src) {
    for (uint32_t j = gtid(); j < c.n; j += gnth()) {
        const uint32_t i = src[j];
        c.ca[j] = c.pose[i];
        c.cb[j] = c.prev[i];
        c.cp0[j] = c.vel[i];
        c.cp1[j] = c.shape[i];
        c.cs[j] = c.mat[i];
        const float2 m = c.minv[i];
        c.lam[j] = make_float4(m.x, m.y, __uint_as_float(c.ext[i]), 0.0f);
    }
}
__global__ void __launch_bounds__(TPB) k_permute_store(Ctx c, uint32_t* __restrict__ ext,
                                                       uint32_t* __restrict__ inv) {
    if (vload(&c.cnt->overflow) & OVERFLOW_CAPACITY) return;  // a frame behind a failed one: no-op
    for (uint32_t j = gtid(); j < c.n; j += gnth()) {
        c.pose[j] = c.ca[j];
        c.prev[j] = c.cb[j];
        c.vel[j] = c.cp0[j];
        const_cast<float4*>(c.shape)[j] = c.cp1[j];
        const_cast<float4*>(c.mat)[j] = c.cs[j];
        const float4 m = c.lam[j];
        const_cast<float2*>(c.minv)[j] = make_float2(m.x, m.y);
        const uint32_t e = __float_as_uint(m.z);
        ext[j] = e;
        inv[e] = j;
    }
}
__global__ void __launch_bounds__(TPB) k_remap_joints(Ctx c, const uint2* __restrict__ jext) {
    for (uint32_t k = gtid(); k < c.nj; k += gnth()) {
        const uint2 j = jext[k];
        const_cast<uint2*>(c.jbod_in)[k] = make_uint2(c.inv[j.x], c.inv[j.y]);
    }
}
// state before the frame (after its re-sort), for the capacity recovery; a frame queued behind a
// failed one must not overwrite it
__global__ void __launch_bounds__(TPB) k_snapshot(Ctx c, float4* __restrict__ pose_bak,
                                                  float4* __restrict__ vel_bak) {
    if (vload(&c.cnt->overflow) & OVERFLOW_CAPACITY) return;
    for (uint32_t i = gtid(); i < c.n; i += gnth()) {
        pose_bak[i] = c.pose[i];
        vel_bak[i] = c.vel[i];
    }
}
__global__ void __launch_bounds__(TPB) k_clear_table(Ctx c, unsigned long long* __restrict__ tab,
                                                     uint32_t slots) {
    if (vload(&c.cnt->overflow) & OVERFLOW_CAPACITY) return;
    for (uint32_t i = gtid(); i < slots; i += gnth()) tab[i] = 0ull;
}
// colour table into a larger one (capacity recovery)
__global__ void __launch_bounds__(TPB) k_rehash(const unsigned long long* __restrict__ from,
                                                uint32_t from_slots, unsigned long long* __restrict__ to,
                                                uint32_t to_mask) {
    for (uint32_t i = gtid(); i < from_slots; i += gnth()) {
        const unsigned long long e = from[i];
        if (e == 0ull) continue;
        const uint32_t a = (uint32_t)((e >> 22) & 0x3fffffu), b = (uint32_t)(e & 0x3fffffu);
        uint32_t h = sfp::ctab_hash(a, b) & to_mask;
        while (atomicCAS(&to[h], 0ull, e) != 0ull) h = (h + 1u) & to_mask;
    }
}
__global__ void __launch_bounds__(TPB) k_identity(uint32_t* __restrict__ a, uint32_t* __restrict__ b,
                                                  uint32_t n) {
    for (uint32_t i = gtid(); i < n; i += gnth()) {
        a[i] = i;
        b[i] = i;
    }
}

// ---- spatial queries on the device hash --------------------------------------------------------
__global__ void __launch_bounds__(TPB) k_query_aabbs(Ctx c) {
    stage_aabb_count(c, 0.0f, gtid(), gnth(), false);  // current poses, no sweep
}

// one thread per ray: DDA through the hash cells, exact ray / collider test per entry
__global__ void __launch_bounds__(TPB) k_ray_cast(Ctx c, const sfp::RayRec* __restrict__ rays,
                                                  uint32_t m, sfp::RayHitRec* __restrict__ hits,
                                                  float cell) {
    const uint32_t E = c.cnt->n_entries;
    for (uint32_t k = gtid(); k < m; k += gnth()) {
        const sfp::RayRec r = rays[k];
        const v2 o = sfp::mk(r.ox, r.oy), d = sfp::mk(r.dx, r.dy);
        uint32_t best_body = 0xffffffffu;
        float best_t = r.max_t;
        v2 best_n = sfp::mk(0.0f, 0.0f);
        if (!(d.x != 0.0f || d.y != 0.0f)) {
            hits[k] = sfp::RayHitRec{0xffffffffu, 0.0f, 0.0f, 0.0f};
            continue;
        }
        // The walk is confined to the cell bounding box of the bodies (Counters::qbox): a ray
        // that starts outside enters it at t0 (slab clip), and the walk ends when it leaves —
        // its length is bounded by the world's extent, not by a step budget, however long the
        // ray is.
        const int bx0 = c.cnt->qbox[0], by0 = c.cnt->qbox[1], bx1 = c.cnt->qbox[2], by1 = c.cnt->qbox[3];
        const float inf = 3.0e38f;
        float t0 = 0.0f, t1 = r.max_t;
        {
            // (one cell of slack: the clip only has to be conservative)
            const float lo_x = (float)(bx0 - 1) * cell, hi_x = (float)(bx1 + 2) * cell;
            const float lo_y = (float)(by0 - 1) * cell, hi_y = (float)(by1 + 2) * cell;
            if (d.x != 0.0f) {
                const float a = (lo_x - o.x) / d.x, b = (hi_x - o.x) / d.x;
                t0 = fmaxf(t0, fminf(a, b));
                t1 = fminf(t1, fmaxf(a, b));
            } else if (o.x < lo_x || o.x > hi_x) {
                t1 = -1.0f;
            }
            if (d.y != 0.0f) {
                const float a = (lo_y - o.y) / d.y, b = (hi_y - o.y) / d.y;
                t0 = fmaxf(t0, fminf(a, b));
                t1 = fminf(t1, fmaxf(a, b));
            } else if (o.y < lo_y || o.y > hi_y) {
                t1 = -1.0f;
            }
        }
        if (bx1 < bx0 || t1 < t0) {  // no bodies, or the ray misses the box
            hits[k] = sfp::RayHitRec{0xffffffffu, 0.0f, 0.0f, 0.0f};
            continue;
        }
        // The walk runs in a parameter local to the entry point (p0 = o + t0 d, t' = t - t0): far
        // from the scene t itself is a large float whose steps are coarser than a cell.  The hit
        // tests below still use the caller's ray, so the reported t is the oracle's.
        const float p0x = o.x + t0 * d.x, p0y = o.y + t0 * d.y;
        int cx = min(max(sfp::cell_coord(p0x, c.inv_cell), bx0), bx1);
        int cy = min(max(sfp::cell_coord(p0y, c.inv_cell), by0), by1);
        const int sx = d.x > 0.0f ? 1 : -1, sy = d.y > 0.0f ? 1 : -1;
        float tmx = d.x != 0.0f ? (((float)(cx + (sx > 0 ? 1 : 0))) * cell - p0x) / d.x : inf;
        float tmy = d.y != 0.0f ? (((float)(cy + (sy > 0 ? 1 : 0))) * cell - p0y) / d.y : inf;
        const float tdx = d.x != 0.0f ? fabsf(cell / d.x) : inf;
        const float tdy = d.y != 0.0f ? fabsf(cell / d.y) : inf;
        for (;;) {
            const uint32_t key = sfp::cell_key(cx, cy);
            uint32_t lo = 0, hi = E;  // lower bound of key among the sorted entries
            while (lo < hi) {
                uint32_t mid = (lo + hi) >> 1;
                if (c.ekey0[mid] < key) lo = mid + 1;
                else hi = mid;
            }
            for (uint32_t e = lo; e < E && c.ekey0[e] == key; ++e) {
                const uint32_t b = c.eval0[e];
                if ((((uint32_t)c.meta[b].z & 0xffffu) & r.mask) == 0u) continue;
                float t;
                v2 n;
                if (!sfp::ray_collider(o, d, r.max_t, ld4(c.pose, b), ld4(c.shape, b),
                                       reinterpret_cast<const sfp::f4*>(c.polytab), &t, &n))
                    continue;
                const uint32_t be = ext_id(c, b);
                if (best_body == 0xffffffffu || t < best_t || (t == best_t && be < best_body)) {
                    best_body = be;
                    best_t = t;
                    best_n = n;
                }
            }
            // this cell is left at t0 + t_exit; stop when the best hit lies before that (with a
            // margin for the rounding of the large sum: stopping late only costs a cell)
            const float t_exit = t0 + fminf(tmx, tmy);
            if (best_body != 0xffffffffu && best_t <= t_exit - 2.0e-6f * fabsf(t_exit)) break;
            if (t_exit - 2.0e-6f * fabsf(t_exit) > r.max_t) break;
            if (tmx < tmy) {
                cx += sx;
                tmx += tdx;
            } else {
                cy += sy;
                tmy += tdy;
            }
            if (cx < bx0 || cx > bx1 || cy < by0 || cy > by1) break;  // left the world
        }
        hits[k] = sfp::RayHitRec{best_body, best_body == 0xffffffffu ? 0.0f : best_t, best_n.x,
                                 best_n.y};
    }
}

__global__ void __launch_bounds__(TPB) k_box_flags(Ctx c, float4 box, uint32_t mask,
                                                   uint32_t* __restrict__ flags) {
    f4 bx;
    bx.x = box.x, bx.y = box.y, bx.z = box.z, bx.w = box.w;
    for (uint32_t e = gtid(); e < c.n; e += gnth()) {  // e: the caller's index
        const uint32_t i = c.inv[e];
        bool in = (((sfp::f2u(c.mat[i].w) & 0xffffu) & mask) != 0u) &&
                  sfp::aabb_overlap(sfp::tight_aabb(ld4(c.pose, i), ld4(c.shape, i),
                                                    reinterpret_cast<const sfp::f4*>(c.polytab)), bx);
        flags[e] = in ? 1u : 0u;
    }
}
__global__ void __launch_bounds__(TPB) k_shape_flags(Ctx c, float4 probe_pose, float4 probe_shape,
                                                     uint32_t mask, uint32_t* __restrict__ flags) {
    f4 pp, ps;
    pp.x = probe_pose.x, pp.y = probe_pose.y, pp.z = probe_pose.z, pp.w = probe_pose.w;
    ps.x = probe_shape.x, ps.y = probe_shape.y, ps.z = probe_shape.z, ps.w = probe_shape.w;
    for (uint32_t e = gtid(); e < c.n; e += gnth()) {  // e: the caller's index
        const uint32_t i = c.inv[e];
        bool in = (((sfp::f2u(c.mat[i].w) & 0xffffu) & mask) != 0u) &&
                  sfp::colliders_overlap(ld4(c.pose, i), ld4(c.shape, i), pp, ps,
                                         reinterpret_cast<const sfp::f4*>(c.polytab));
        flags[e] = in ? 1u : 0u;
    }
}
__global__ void __launch_bounds__(TPB) k_box_write(Ctx c, const uint32_t* __restrict__ off,
                                                   uint32_t* __restrict__ out) {
    for (uint32_t i = gtid(); i < c.n; i += gnth())
        if (off[i + 1] != off[i]) out[off[i]] = i;
}

// ---- contact events: flag -> scan -> write, so only the events cross the bus ------------------
__global__ void __launch_bounds__(TPB) k_event_flags(Ctx c, uint32_t* __restrict__ flags) {
    const uint32_t P = c.cnt->n_pairs;
    for (uint32_t k = gtid(); k < P; k += gnth()) {
        float4 cb = c.cb[k], lam = c.lam[k];
        uint32_t cnt = __float_as_uint(cb.y);
        flags[k] = ((cnt > 0u && lam.x > 0.0f) || (cnt > 1u && lam.z > 0.0f)) ? 1u : 0u;
    }
}
__global__ void __launch_bounds__(TPB) k_event_write(Ctx c, const uint32_t* __restrict__ off,
                                                     sfp::EventRec* __restrict__ out, float inv_h) {
    const uint32_t P = c.cnt->n_pairs;
    for (uint32_t k = gtid(); k < P; k += gnth()) {
        if (off[k + 1] == off[k]) continue;
        uint2 p = c.pairs[k];
        sfp::EventRec e;
        if (sfp::pair_event(ext_id(c, p.x), ext_id(c, p.y), ld4(c.ca, k), ld4(c.cb, k), ld4(c.cp0, k),
                            ld4(c.cp1, k), ld4(c.lam, k), ld4(c.pose, p.x), ld4(c.pose, p.y), inv_h,
                            &e)) {
            uint4* o = reinterpret_cast<uint4*>(out + off[k]);
            o[0] = make_uint4(e.a, e.b, __float_as_uint(e.nx), __float_as_uint(e.ny));
            o[1] = make_uint4(__float_as_uint(e.px), __float_as_uint(e.py),
                              __float_as_uint(e.impulse), e.points);
        }
    }
}
// scatter-update of individual bodies from scene-graph states
__global__ void __launch_bounds__(TPB) k_update_bodies(const uint32_t* __restrict__ idx,
                                                       const uint32_t* __restrict__ inv,
                                                       const sfp::StateRec* __restrict__ in,
                                                       const float2* __restrict__ minv,
                                                       float4* __restrict__ pose,
                                                       float4* __restrict__ vel, uint32_t m) {
    for (uint32_t k = gtid(); k < m; k += gnth()) {
        const uint32_t i = inv[idx[k]];
        const float4* q = reinterpret_cast<const float4*>(in + k);
        float4 a = q[0], b = q[1];
        sfp::StateRec r;
        r.x = a.x, r.y = a.y, r.c = a.z, r.s = a.w, r.vx = b.x, r.vy = b.y, r.omega = b.z, r.angle = b.w;
        f4 p, v;
        sfp::state_unpack(r, ld2(minv, i), &p, &v);
        st4(pose, i, p);
        st4(vel, i, v);
    }
}

// ---- scene-graph state <-> solver records, on the device (AoS SfpBodyState, 32 B/body) -------
// (states cross the ABI in the caller's order: e is the caller's index, inv[e] where the body is)
__global__ void __launch_bounds__(TPB) k_pack_states(const float4* __restrict__ pose,
                                                     const float4* __restrict__ vel,
                                                     const uint32_t* __restrict__ inv,
                                                     sfp::StateRec* __restrict__ out, uint32_t n) {
    for (uint32_t e = gtid(); e < n; e += gnth()) {
        const uint32_t i = inv[e];
        sfp::StateRec r = sfp::state_pack(ld4(pose, i), ld4(vel, i));
        float4* o = reinterpret_cast<float4*>(out + e);
        o[0] = make_float4(r.x, r.y, r.c, r.s);
        o[1] = make_float4(r.vx, r.vy, r.omega, r.angle);
    }
}
__global__ void __launch_bounds__(TPB) k_unpack_states(const sfp::StateRec* __restrict__ in,
                                                       const float2* __restrict__ minv,
                                                       const uint32_t* __restrict__ inv,
                                                       float4* __restrict__ pose,
                                                       float4* __restrict__ vel, uint32_t n) {
    for (uint32_t e = gtid(); e < n; e += gnth()) {
        const uint32_t i = inv[e];
        const float4* q = reinterpret_cast<const float4*>(in + e);
        float4 a = q[0], b = q[1];
        sfp::StateRec r;
        r.x = a.x, r.y = a.y, r.c = a.z, r.s = a.w, r.vx = b.x, r.vy = b.y, r.omega = b.z, r.angle = b.w;
        f4 p, v;
        sfp::state_unpack(r, ld2(minv, i), &p, &v);
        st4(pose, i, p);
        st4(vel, i, v);
    }
}

// ---- persistent cooperative kernels (product path) ----------------------------------------
__global__ void __launch_bounds__(TPB) k_colour_coop(Ctx c) {
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = gtid(), nth = gnth();
    // the broadphase ran out of room (this frame or an earlier one): keep last frame's colours
    // and table untouched for the re-run (every thread takes this exit, before any grid barrier)
    if (vload(&c.cnt->overflow) & OVERFLOW_CAPACITY) return;
    // wall time of the parts (ns, thread 0) -> stage_cycles[1..3]: init + warm, rounds, record
    unsigned long long t0 = 0, t1 = 0, t2 = 0;
    auto now = [&]() {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        return t;
    };
    if (tid == 0) t0 = now();
    stage_colour_init(c, tid, nth);
    grid.sync();
    if (c.warm) {
        stage_colour_warm(c, tid, nth);
        grid.sync();
    }
    if (tid == 0) t1 = now();
    uint32_t round = 0;
    for (; round < (uint32_t)MAX_ROUNDS; ++round) {
        stage_colour_bid(c, round, tid, nth);
        grid.sync();
        stage_colour_take(c, round, tid, nth);
        grid.sync();
        if (vload(&c.cnt->round_flags[round]) == 0u) break;
    }
    if (round == (uint32_t)MAX_ROUNDS && tid == 0) atomicOr(&c.cnt->overflow, 8u);
    if (tid == 0) t2 = now();
    stage_colour_record(c, tid, nth);
    if (tid == 0) {
        c.cnt->stage_cycles[1] = t1 - t0;
        c.cnt->stage_cycles[2] = t2 - t1;
        c.cnt->stage_cycles[3] = now() - t2;
    }
}

// Colours are split into BIG ones (more than SFP_TAIL_ITEMS constraints: one grid-wide phase
// each) and SMALL ones.  A small colour is a handful of constraints around the few high-degree
// bodies of the scene; a grid-wide phase for it costs the same ~3-4 us as a phase with 50 000
// constraints.  So all small colours are run by block 0 alone, separated by __syncthreads(),
// inside ONE grid-wide phase.  The solve order (big colours ascending, then small colours
// ascending) is part of the algorithm's definition (sfp_math.h) and the oracle follows it.
__device__ __forceinline__ int next_bit(unsigned long long m, int after) {
    // lowest set bit of m with index > after (after = -1 for the first); 64 if none
    unsigned long long r = (after >= 63) ? 0ull : (m & (~0ull << (after + 1)));
    return r ? (__ffsll((long long)r) - 1) : 64;
}

// one 512-thread block per SM rather than two of 256: the same threads and registers, but a
// small colour of up to 512 constraints is a single pass of block 0 (pile100k: small colours
// 166 -> 136 us per frame)
#ifndef SFP_SUB_BLOCKS
#define SFP_SUB_BLOCKS 1
#endif
#ifndef SFP_SUB_TPB
#define SFP_SUB_TPB 512
#endif
// POLYS: the variant that can collide general polygons (bigger, more registers); the host
// launches it only for worlds that hold at least one polygon body
// TILES: the solve order with tiles (sfp_math.h): every block first runs the interior colours of
// its tiles out of shared memory (stage_tiles), then the boundary colours follow in grid-wide
// phases as before
template <bool POLYS, bool TILES>
__global__ void __launch_bounds__(SFP_SUB_TPB, SFP_SUB_BLOCKS) k_substeps_coop(Ctx c, uint32_t substeps) {
    const float h = c.fp->h, rest_threshold = c.fp->rest_threshold;
    extern __shared__ float4 dyn_smem[];
    __shared__ uint32_t s_toff[3 * (SFP_MAX_COLOURS + 1)];
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = gtid(), nth = gnth();
    if (vload(&c.cnt->overflow) & OVERFLOW_CAPACITY) return;  // see k_colour_coop
    // A colour's constraints go to threads in warp-sized runs dealt round-robin over the BLOCKS
    // (run r -> block r % gridDim, warp r / gridDim), not block after block: a phase lasts as
    // long as its slowest warp, and a colour of a few thousand constraints would otherwise fill
    // every warp of a few SMs (16 warps contending for issue slots and the L1) while the rest of
    // the GPU idles.  Which thread solves which constraint does not change any result.
    const uint32_t stid = ((threadIdx.x >> 5) * gridDim.x + blockIdx.x) * 32u + (threadIdx.x & 31u);
    const uint32_t ncol = vload(&c.cnt->n_colours);
    const unsigned long long all = (ncol >= 64u) ? ~0ull : ((1ull << ncol) - 1ull);
    const unsigned long long big = big_colour_mask(c, ncol);
    const unsigned long long small = all & ~big & ~(1ull << SFP_SERIAL_COLOUR);
    const bool serial = ncol > SFP_SERIAL_COLOUR;
    const float inv_h = 1.0f / h, inv_h2 = inv_h * inv_h;
    // once per frame (substeps == all substeps of the frame, or 1 with per-substep broadphase):
    // the serial group must be in (a, b) order before the first narrowphase
    if (serial && blockIdx.x == 0) serial_sort_pairs(c, ncol);
    // the colour offsets, once per launch, where every phase can reach them without an L2 trip
    __shared__ uint32_t s_poff[SFP_MAX_COLOURS + 1], s_joff[SFP_MAX_COLOURS + 1];
    for (uint32_t i = threadIdx.x; i <= (uint32_t)SFP_MAX_COLOURS; i += blockDim.x) {
        s_poff[i] = c.cnt->pair_col_off[i];
        s_joff[i] = c.cnt->joint_col_off[i];
    }
    __syncthreads();

    // Prefetch: the contact record of this thread's first item of the NEXT big colour does not
    // depend on anything the current phase writes, so it is loaded before the barrier and only
    // the body gather remains on the critical path after it.
    PairRec rec;
    uint32_t rec_k = 0;
    bool have = false;
    auto prefetch_pos = [&](int col, uint32_t t0) {
        have = false;
        if (col >= 64) return;
        if (s_joff[col + 1] != s_joff[col]) return;
        uint32_t k = s_poff[col] + t0;
        if (k < s_poff[col + 1]) {
            rec = load_pair_rec(c, k);
            rec_k = k;
            have = true;
        }
    };
    f4 rec_lam;
    auto prefetch_vel = [&](int col, uint32_t t0) {
        have = false;
        if (col >= 64) return;
        uint32_t k = s_poff[col] + t0;
        if (k < s_poff[col + 1]) {
            rec_lam = ld4(c.lam, k);
            rec = load_pair_rec(c, k);
            rec_k = k;
            have = true;
        }
    };

    // after the last big colour, block 0 fetches its first SMALL colour's record instead, so
    // that the small-colour loop starts from registers as well
    const int first_small = next_bit(small, -1);
    auto prefetch_pos_after = [&](int col) {
        const int nb = next_bit(big, col);
        if (nb < 64) prefetch_pos(nb, stid);
        else if (blockIdx.x == 0) prefetch_pos(first_small, threadIdx.x);
        else have = false;
    };
    auto prefetch_vel_after = [&](int col) {
        const int nb = next_bit(big, col);
        if (nb < 64) prefetch_vel(nb, stid);
        else if (blockIdx.x == 0) prefetch_vel(first_small, threadIdx.x);
        else have = false;
    };

    // per-stage wall time (ns, %globaltimer) seen by thread 0 — the phases are grid-wide, so
    // one observer is representative
    unsigned long long t_prev = 0, acc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    auto lap = [&](int slot) {
        if (tid == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (slot >= 0) acc[slot] += t - t_prev;
            t_prev = t;
        }
    };
#ifdef SFP_PHASE_PROBE
    uint32_t ph = 0;
    long long pr_dur = 0, pr_wait0 = 0;
    for (uint32_t i = tid; i < 512u; i += nth) c.cnt->round_flags[i] = 0u;
    grid.sync();
#endif
#ifdef SFP_TILE_PROBE
    for (uint32_t i = tid; i < 512u; i += nth) c.cnt->round_flags[i] = 0u;
    grid.sync();
#endif
    lap(-1);
    for (uint32_t s = 0; s < substeps; ++s) {
        stage_integrate(c, h, tid, nth);
        grid.sync();
        lap(4);
        stage_narrowphase<POLYS, false, NP_STAGED>(c, tid, nth);
        grid.sync();
        lap(5);
        if (TILES) {
            stage_tiles<false>(c, dyn_smem, s_toff, h, rest_threshold);
            grid.sync();
            lap(10);
        }
        have = false;
#ifdef SFP_PHASE_PROBE
        // development probe (scripts/phase_probe.py).  Per big position phase: the slowest and
        // fastest block's work time (cycles from leaving the previous barrier to arriving at
        // this one) in round_flags[ph] / [256 + ph]; over all busy warps of all phases, the
        // sum and maximum of the gather / compute / store / rest segments in [480..495].
        for (int col = next_bit(big, -1); col < 64; col = next_bit(big, col)) {
            long long ts = clock64(), t1 = ts, t2 = ts, t3 = ts;
            bool busy = false;
            if (have) {
                uint32_t cnt = sfp::f2u(rec.cb.y);
                f4 lam;
                lam.x = lam.y = lam.z = lam.w = 0.0f;
                if (cnt > 0) {
                    busy = true;
                    sfp::BodyPos A = load_pos(c, rec.p.x), B = load_pos(c, rec.p.y);
                    asm volatile("" ::"f"(A.x.x), "f"(B.x.x), "f"(A.d.x), "f"(B.d.x), "f"(A.im), "f"(B.im));
                    t1 = clock64();
                    lam = sfp::solve_contact_manifold(
                        &A, &B, sfp::mk(rec.ca.x, rec.ca.y), cnt, sfp::mk(rec.q0.x, rec.q0.y),
                        sfp::mk(rec.q0.z, rec.q0.w), sfp::mk(rec.q1.x, rec.q1.y),
                        sfp::mk(rec.q1.z, rec.q1.w), rec.ca.z);
                    asm volatile("" ::"f"(A.x.x), "f"(B.x.x), "f"(A.q.x), "f"(B.q.x), "f"(lam.x), "f"(lam.y));
                    t2 = clock64();
                    store_pos(c, rec.p.x, A);
                    store_pos(c, rec.p.y, B);
                }
                st4(c.lam, rec_k, lam);
                t3 = clock64();
            }
            stage_solve_pos(c, (uint32_t)col, inv_h2, stid, nth, have, s_joff, s_poff);
            prefetch_pos_after(col);
            long long t4 = clock64();
            if (__any_sync(0xffffffffu, busy) && (threadIdx.x & 31u) == 0u) {
                // lane 0's clock reads bracket the whole warp's (divergent) execution
                uint32_t seg[4] = {(uint32_t)(t1 - ts), (uint32_t)(t2 - t1), (uint32_t)(t3 - t2),
                                   (uint32_t)(t4 - t3)};
                for (int k = 0; k < 4; ++k) {
                    atomicAdd(&c.cnt->round_flags[480 + k], seg[k] >> 4);
                    atomicMax(&c.cnt->round_flags[484 + k], seg[k]);
                }
                atomicAdd(&c.cnt->round_flags[488], 1u);
            }
            __syncthreads();
            long long ta = clock64();
            if (threadIdx.x == 0 && ph < 224u) {
                uint32_t wk = (uint32_t)(ta - ts);
                atomicMax(&c.cnt->round_flags[ph], wk);
                atomicMax(&c.cnt->round_flags[256 + ph], 0xffffffffu - wk);
            }
            grid.sync();
            long long te = clock64();
            if (tid == 0) {
                pr_dur += te - ts;
                pr_wait0 += te - ta;
            }
            ++ph;
        }
#else
        for (int col = next_bit(big, -1); col < 64; col = next_bit(big, col)) {
            if (have) solve_pos_pair(c, rec_k, rec);
            stage_solve_pos(c, (uint32_t)col, inv_h2, stid, nth, have, s_joff, s_poff);
            prefetch_pos_after(col);
            grid.sync();
        }
#endif
        lap(6);
        if (small || serial) {
            if (blockIdx.x == 0) {
                // same record prefetch as above, one block wide (the first record came with
                // the last big colour, if there was one)
                if (!big) have = false;
                for (int col = next_bit(small, -1); col < 64; col = next_bit(small, col)) {
                    if (have) solve_pos_pair(c, rec_k, rec);
                    stage_solve_pos(c, (uint32_t)col, inv_h2, threadIdx.x, blockDim.x, have, s_joff,
                                    s_poff);
                    prefetch_pos(next_bit(small, col), threadIdx.x);
                    __syncthreads();
                }
                if (serial && threadIdx.x == 0) serial_solve_pos(c, ncol, inv_h2);
            }
            grid.sync();
        }
        lap(9);
        stage_derive(c, inv_h, tid, nth);
        grid.sync();
        lap(7);
        if (TILES) {
            stage_tiles<true>(c, dyn_smem, s_toff, h, rest_threshold);
            grid.sync();
            lap(11);
        }
        have = false;
        for (int col = next_bit(big, -1); col < 64; col = next_bit(big, col)) {
            if (have) solve_vel_pair(c, rec_k, rec, rec_lam, h, rest_threshold);
            stage_solve_vel(c, (uint32_t)col, h, rest_threshold, stid, nth, have, s_poff);
            prefetch_vel_after(col);
            grid.sync();
        }
        lap(8);
        if (small || serial) {
            if (blockIdx.x == 0) {
                if (!big) have = false;
                for (int col = next_bit(small, -1); col < 64; col = next_bit(small, col)) {
                    if (have) solve_vel_pair(c, rec_k, rec, rec_lam, h, rest_threshold);
                    stage_solve_vel(c, (uint32_t)col, h, rest_threshold, threadIdx.x, blockDim.x,
                                    have, s_poff);
                    prefetch_vel(next_bit(small, col), threadIdx.x);
                    __syncthreads();
                }
                if (serial && threadIdx.x == 0) serial_solve_vel(c, ncol, h, rest_threshold);
            }
            grid.sync();
        }
        lap(9);
    }
    if (tid == 0) {
        c.cnt->frames_done += 1u;
        for (int k = 4; k < 12; ++k) c.cnt->stage_cycles[k] = acc[k];
        c.cnt->stage_cycles[0] = (unsigned long long)__popcll(big);
#ifdef SFP_PHASE_PROBE
        unsigned long long smax = 0, smin = 0;
        uint32_t np = ph < 224u ? ph : 224u;
        for (uint32_t i = 0; i < np; ++i) {
            smax += vload(&c.cnt->round_flags[i]);
            smin += 0xffffffffu - vload(&c.cnt->round_flags[256 + i]);
        }
        c.cnt->stage_cycles[1] = smax;
        c.cnt->stage_cycles[2] = smin;
        c.cnt->stage_cycles[3] = np;
        c.cnt->stage_cycles[4] = (unsigned long long)pr_dur;
        c.cnt->stage_cycles[5] = (unsigned long long)pr_wait0;
#endif
    }
}

// Experiment kept behind SFP_FLAG_FLOW_SOLVE (same bits as k_substeps_coop, slower — see
// profiles/r2_flow_experiment.md): each solver pass is ONE grid-wide phase in which constraints
// wait for their own two bodies only (sfp_device.cuh, "Dataflow solve").  Grid barriers per
// substep: 5 (integrate | narrowphase | position pass | derive | velocity pass) instead of
// 3 + 2 x (big colours + 1).
template <bool POLYS>
__global__ void __launch_bounds__(SFP_SUB_TPB, SFP_SUB_BLOCKS)
    k_substeps_flow(Ctx c, uint32_t substeps) {
    const float h = c.fp->h, rest_threshold = c.fp->rest_threshold;
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = gtid(), nth = gnth();
    const uint32_t lane = threadIdx.x & 31u;
    if (vload(&c.cnt->overflow) & OVERFLOW_CAPACITY) return;  // see k_colour_coop
    // consecutive runs go to different SMs
    const uint32_t gwarp = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5);
    const uint32_t ncol = vload(&c.cnt->n_colours);
    const bool serial = ncol > SFP_SERIAL_COLOUR;
    const float inv_h = 1.0f / h, inv_h2 = inv_h * inv_h;
    if (serial && blockIdx.x == 0) serial_sort_pairs(c, ncol);
    __shared__ uint32_t s_poff[SFP_MAX_COLOURS + 1], s_joff[SFP_MAX_COLOURS + 1];
    __shared__ FlowOrder s_fo;
    for (uint32_t i = threadIdx.x; i <= (uint32_t)SFP_MAX_COLOURS; i += blockDim.x) {
        s_poff[i] = c.cnt->pair_col_off[i];
        s_joff[i] = c.cnt->joint_col_off[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) flow_build_order(c, ncol, s_joff, s_poff, &s_fo);
    __syncthreads();
#ifdef SFP_FLOW_PROBE
    for (uint32_t i = tid; i < 512u; i += nth) c.cnt->round_flags[i] = 0u;
    grid.sync();
#endif

    unsigned long long t_prev = 0, acc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    auto lap = [&](int slot) {
        if (tid == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (slot >= 0) acc[slot] += t - t_prev;
            t_prev = t;
        }
    };
    lap(-1);
    for (uint32_t s = 0; s < substeps; ++s) {
        stage_integrate(c, h, tid, nth);
        grid.sync();
        lap(4);
        stage_narrowphase<POLYS, true, NP_STAGED>(c, tid, nth);
        grid.sync();
        lap(5);
        flow_solve_pos(c, s_fo, s_joff, s_poff, inv_h2, gwarp, nwarps, lane);
        if (serial) {
            grid.sync();
            if (tid == 0) serial_solve_pos(c, ncol, inv_h2);
        }
        grid.sync();
        lap(6);
        stage_derive(c, inv_h, tid, nth);
        grid.sync();
        lap(7);
        flow_solve_vel(c, s_fo, s_joff, s_poff, h, rest_threshold, gwarp, nwarps, lane);
        if (serial) {
            grid.sync();
            if (tid == 0) serial_solve_vel(c, ncol, h, rest_threshold);
        }
        grid.sync();
        lap(8);
    }
    if (tid == 0) {
        c.cnt->frames_done += 1u;
        for (int k = 4; k < 10; ++k) c.cnt->stage_cycles[k] = acc[k];
        c.cnt->stage_cycles[0] = (unsigned long long)s_fo.n;
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            w->err = std::string(#call) + ": " + cudaGetErrorString(e_);                  \
            w->sticky = SFP_E_CUDA;                                                       \
            return SFP_E_CUDA;                                                            \
        }                                                                                 \
    } while (0)

struct SfpWorld {
    SfpParams p;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_bp = nullptr, ev_col = nullptr;
    cudaStream_t side = nullptr;                  // frame work that only the colouring waits for
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool timed = false;
    uint32_t launches = 0;
    uint32_t n = 0, nj = 0;
    uint32_t cap_bodies = 0, cap_joints = 0, cap_entries = 0, cap_pairs = 0;
    Ctx ctx;
    std::vector<void*> body_allocs, pair_allocs, joint_allocs;
    // capacity recovery: the state before the last frame, and what the frames since were
    float4 *d_pose_bak = nullptr, *d_vel_bak = nullptr;
    struct FrameLog {
        float dt;
        uint32_t substeps, colour_age;
    } frame_log[1024];
    uint64_t frames_launched = 0;  // sfp_step calls since the last upload
    uint32_t grows = 0;            // times the entry / pair buffers were enlarged
    uint32_t *d_scan_tiles = nullptr, *d_hist = nullptr;
    uint32_t hist_blocks = 0;
    unsigned long long* d_tab[2] = {nullptr, nullptr};
    uint32_t tab_slots = 0;
    uint32_t colour_age = 0;  // colourings since the last upload (warm start, sfp_math.h)
    // the frame's launch sequence captured as CUDA graphs (one per (substeps, table parity, warm)
    // combination; dt travels in FrameParams); falls back to plain launches if capture is refused
    struct FrameGraph {
        cudaGraphExec_t exec = nullptr;
        uint32_t substeps = 0, parity = 0, warm = 0, launches = 0;
    } graphs[6];
    uint32_t graph_next = 0;
    bool graphs_ok = true, capturing = false;
    Counters* d_cnt = nullptr;
    FrameParams* d_fp = nullptr;  // dt / h / rest threshold of the frame being queued
    // internal body order (re-sort): internal -> caller's index and back; joints as uploaded
    uint32_t *d_ext = nullptr, *d_inv = nullptr;
    uint2* d_jext = nullptr;
    // host mirrors
    sfp::PolyTable polys;
    float4* d_polytab = nullptr;
    std::vector<sfp::f2> h_minv;
    std::vector<sfp::f4> h_shape, h_mat;
    uint32_t *d_evt_flag = nullptr, *d_evt_off = nullptr;  // cap_pairs + 1 each
    sfp::EventRec* d_events = nullptr;                    // cap_pairs
    uint32_t* d_idx = nullptr;                            // cap_bodies (sfp_update_bodies)
    sfp::RayRec* d_rays = nullptr;                        // ray_cap (grow-only)
    sfp::RayHitRec* d_hits = nullptr;
    uint32_t ray_cap = 0;
    float last_h = 0.0f;
    SfpBodyState* pin_state = nullptr;   // pinned host staging, cap_bodies states
    sfp::StateRec* d_state = nullptr;    // device staging, cap_bodies states
    uint32_t pin_cap = 0;
    int coop_max_colour = 0, coop_max_substep = 0, coop_max_substep_poly = 0, n_sm = 0;
    int coop_max_flow = 0, coop_max_flow_poly = 0, coop_max_tiles = 0, coop_max_tiles_poly = 0;
    // ping-pong partners of the Ctx's pcol_sorted / pperm / jcol_sorted / jperm
    uint32_t *d_pkeyB = nullptr, *d_ppermB = nullptr, *d_jkeyB = nullptr, *d_jpermB = nullptr;
    int sort_passes = 1;  // radix passes over the (group, colour, class) key
    uint32_t max_tiles = 0;
    bool has_polygons = false;
    SfpStats stats;
    std::string err;
    int32_t sticky = SFP_OK;
    bool stepped = false;
};

namespace {

template <typename T>
int alloc_dev(SfpWorld* w, T** out, size_t count, std::vector<void*>& bag) {
    void* p = nullptr;
    size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    CU(cudaMalloc(&p, bytes));
    CU(cudaMemsetAsync(p, 0, bytes, w->stream));
    bag.push_back(p);
    *out = reinterpret_cast<T*>(p);
    return SFP_OK;
}

void free_bag(std::vector<void*>& bag) {
    for (void* p : bag) cudaFree(p);
    bag.clear();
}

#define TRY(x)                      \
    do {                            \
        int rc_ = (x);              \
        if (rc_ != SFP_OK) return rc_; \
    } while (0)

// entry- and pair-sized buffers (scratch of one frame, plus the two colour tables); separate from
// the body-sized ones so that the capacity recovery can enlarge them without touching the state
int alloc_pair_arrays(SfpWorld* w, uint32_t ce, uint32_t cp) {
    Ctx& c = w->ctx;
    auto& bag = w->pair_allocs;
    TRY(alloc_dev(w, &c.ekey0, ce, bag));
    TRY(alloc_dev(w, &c.eval0, ce, bag));
    TRY(alloc_dev(w, &c.ekey1, ce, bag));
    TRY(alloc_dev(w, &c.eval1, ce, bag));
    TRY(alloc_dev(w, &c.pcnt, ce + 1, bag));
    TRY(alloc_dev(w, &c.poff, ce + 1, bag));
    TRY(alloc_dev(w, &c.pairs_raw, cp, bag));
    TRY(alloc_dev(w, &c.pairs, cp, bag));
    TRY(alloc_dev(w, &c.pprio, cp, bag));
    TRY(alloc_dev(w, &c.pcol, cp, bag));
    TRY(alloc_dev(w, &c.unc, cp, bag));
    TRY(alloc_dev(w, &c.unc2, cp, bag));
    TRY(alloc_dev(w, &c.pcol_sorted, cp, bag));
    TRY(alloc_dev(w, &c.pperm, cp, bag));
    TRY(alloc_dev(w, &w->d_pkeyB, cp, bag));
    TRY(alloc_dev(w, &w->d_ppermB, cp, bag));
    TRY(alloc_dev(w, &c.ca, cp, bag));
    TRY(alloc_dev(w, &c.cb, cp, bag));
    TRY(alloc_dev(w, &c.cp0, cp, bag));
    TRY(alloc_dev(w, &c.cp1, cp, bag));
    TRY(alloc_dev(w, &c.cs, cp, bag));
    TRY(alloc_dev(w, &c.lam, cp, bag));
    TRY(alloc_dev(w, &c.psa, cp, bag));
    TRY(alloc_dev(w, &c.psb, cp, bag));
    TRY(alloc_dev(w, &c.pmt, cp, bag));
    TRY(alloc_dev(w, &w->d_evt_flag, cp + 1, bag));
    TRY(alloc_dev(w, &w->d_evt_off, cp + 1, bag));
    TRY(alloc_dev(w, &w->d_events, cp, bag));
    uint32_t slots = 1024;
    while (slots < 2u * cp) slots <<= 1;
    TRY(alloc_dev(w, &w->d_tab[0], slots, bag));
    TRY(alloc_dev(w, &w->d_tab[1], slots, bag));
    w->tab_slots = slots;
    c.tab_mask = slots - 1u;
    w->cap_entries = ce;
    w->cap_pairs = cp;
    c.cap_entries = ce;
    c.cap_pairs = cp;
    uint32_t biggest = std::max(std::max(ce, cp), std::max(w->cap_bodies, w->cap_joints)) + 1;
    w->hist_blocks = radix_blocks(biggest);
    uint32_t *tiles = nullptr, *hist = nullptr;
    TRY(alloc_dev(w, &tiles, biggest / SCAN_TILE + 4, bag));
    TRY(alloc_dev(w, &hist, RADIX_BINS * ((size_t)w->hist_blocks + 1), bag));
    w->d_scan_tiles = tiles;
    w->d_hist = hist;
    return SFP_OK;
}

int ensure_body_capacity(SfpWorld* w, uint32_t n) {
    if (n <= w->cap_bodies && w->cap_bodies > 0) return SFP_OK;
    CU(cudaStreamSynchronize(w->stream));
    free_bag(w->body_allocs);
    free_bag(w->pair_allocs);
    w->d_scan_tiles = w->d_hist = nullptr;
    uint32_t cb = std::max<uint32_t>(std::max(n, w->p.max_bodies), 1u);
    uint32_t ce = std::max<uint32_t>(w->p.max_entries, 8u * cb + 8192u);
    uint32_t cp = std::max<uint32_t>(w->p.max_pairs, 12u * cb + 4096u);
    Ctx& c = w->ctx;
    auto& bag = w->body_allocs;
    TRY(alloc_dev(w, &c.pose, cb, bag));
    TRY(alloc_dev(w, &c.prev, cb, bag));
    TRY(alloc_dev(w, &c.vel, cb, bag));
    TRY(alloc_dev(w, const_cast<float4**>(&c.shape), cb, bag));
    TRY(alloc_dev(w, const_cast<float4**>(&c.mat), cb, bag));
    TRY(alloc_dev(w, const_cast<float2**>(&c.minv), cb, bag));
    TRY(alloc_dev(w, &c.aabb, cb, bag));
    TRY(alloc_dev(w, &c.meta, cb, bag));
    TRY(alloc_dev(w, &c.cell_cnt, cb + 1, bag));
    TRY(alloc_dev(w, &c.cell_off, cb + 1, bag));
    TRY(alloc_dev(w, &c.mask, cb, bag));
    TRY(alloc_dev(w, &c.claim0, cb, bag));
    TRY(alloc_dev(w, &c.claim1, cb, bag));
    TRY(alloc_dev(w, &c.done, 2 * (size_t)cb, bag));
    TRY(alloc_dev(w, &c.mask_i, cb, bag));
    w->max_tiles = cb / SFP_TILE_MIN + 2u;
    TRY(alloc_dev(w, &c.goff, ((size_t)w->max_tiles + 1u) * SFP_MAX_COLOURS + 1u, bag));
    TRY(alloc_dev(w, &c.jgoff, ((size_t)w->max_tiles + 1u) * SFP_MAX_COLOURS + 1u, bag));
    TRY(alloc_dev(w, &w->d_idx, cb, bag));
    TRY(alloc_dev(w, &w->d_ext, cb, bag));
    TRY(alloc_dev(w, &w->d_inv, cb, bag));
    TRY(alloc_dev(w, &w->d_pose_bak, cb, bag));
    TRY(alloc_dev(w, &w->d_vel_bak, cb, bag));
    c.ext = w->d_ext;
    c.inv = w->d_inv;
    w->cap_bodies = cb;
    TRY(alloc_pair_arrays(w, ce, cp));
    if (w->pin_state) cudaFreeHost(w->pin_state);
    w->pin_state = nullptr;
    CU(cudaMallocHost(&w->pin_state, sizeof(SfpBodyState) * cb));
    TRY(alloc_dev(w, &w->d_state, cb, bag));
    w->pin_cap = cb;
    return SFP_OK;
}

int ensure_joint_capacity(SfpWorld* w, uint32_t nj) {
    if (nj <= w->cap_joints && w->cap_joints > 0) return SFP_OK;
    CU(cudaStreamSynchronize(w->stream));
    free_bag(w->joint_allocs);
    uint32_t cj = std::max<uint32_t>(std::max(nj, w->p.max_joints), 1u);
    Ctx& c = w->ctx;
    auto& bag = w->joint_allocs;
    TRY(alloc_dev(w, const_cast<uint2**>(&c.jbod_in), cj, bag));
    TRY(alloc_dev(w, const_cast<float4**>(&c.janc_in), cj, bag));
    TRY(alloc_dev(w, const_cast<float4**>(&c.jpar_in), cj, bag));
    TRY(alloc_dev(w, &c.jbod, cj, bag));
    TRY(alloc_dev(w, &c.janc, cj, bag));
    TRY(alloc_dev(w, &c.jpar, cj, bag));
    TRY(alloc_dev(w, &c.jcol, cj, bag));
    TRY(alloc_dev(w, &c.jcol_sorted, cj, bag));
    TRY(alloc_dev(w, &c.jperm, cj, bag));
    TRY(alloc_dev(w, &w->d_jext, cj, bag));
    TRY(alloc_dev(w, &w->d_jkeyB, cj, bag));
    TRY(alloc_dev(w, &w->d_jpermB, cj, bag));
    w->cap_joints = cj;
    // the radix histogram (and the scan's tile sums) must cover the joint sort too
    if (radix_blocks(cj + 1) > w->hist_blocks && w->cap_bodies) {
        w->hist_blocks = radix_blocks(cj + 1);
        uint32_t *tiles = nullptr, *hist = nullptr;
        TRY(alloc_dev(w, &tiles, (size_t)cj / SCAN_TILE + 4, w->pair_allocs));
        TRY(alloc_dev(w, &hist, RADIX_BINS * ((size_t)w->hist_blocks + 1), w->pair_allocs));
        w->d_scan_tiles = tiles;
        w->d_hist = hist;
    }
    return SFP_OK;
}

uint32_t blocks_for(const SfpWorld* w, uint64_t items) {
    uint64_t b = (items + TPB - 1) / TPB;
    uint64_t cap = (uint64_t)w->n_sm * 8;
    return (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(b, cap));
}

constexpr int SFP_KEY_PASSES = (2 * SFP_KEY_BITS + 7) / 8;
static_assert(SFP_KEY_PASSES % 2 == 1, "stage_emit_entries writes buffer 1; an odd pass count ends in buffer 0");
int sort_entries(SfpWorld* w) {
    Ctx& c = w->ctx;
    const uint32_t* n_ptr = &w->d_cnt->n_entries;
    // the hist grid must match between hist and scatter: radix_pass derives it from cap
    for (int pass = 0; pass < SFP_KEY_PASSES; ++pass) {
        const bool back = (pass & 1) != 0;  // 1 -> 0 -> 1 -> 0: sorted entries end in ekey0 / eval0
        const uint32_t* ki = back ? c.ekey0 : c.ekey1;
        const uint32_t* vi = back ? c.eval0 : c.eval1;
        uint32_t* ko = back ? c.ekey1 : c.ekey0;
        uint32_t* vo = back ? c.eval1 : c.eval0;
        CU(radix_pass(ki, vi, ko, vo, n_ptr, w->cap_entries, pass * 8, w->d_hist, w->stream));
    }
    return SFP_OK;
}

int launch_coop(SfpWorld* w, const void* fn, uint32_t blocks, void** args, uint32_t tpb = TPB) {
    CU(cudaLaunchCooperativeKernel(fn, dim3(blocks), dim3(tpb), args, 0, w->stream));
    return SFP_OK;
}

// bodies into Morton order of their cells, statics last (see the kernels' comment)
int resort_bodies(SfpWorld* w) {
    Ctx& c = w->ctx;
    cudaStream_t st = w->stream;
    const uint32_t gb = blocks_for(w, w->n);
    k_tile_keys<<<gb, TPB, 0, st>>>(c, c.ekey1, c.eval1);
    // four 8-bit passes, 1 -> 0 -> 1 -> 0 -> 1: the sorted order ends in eval1
    for (int pass = 0; pass < 4; ++pass) {
        const bool back = (pass & 1) != 0;
        const uint32_t* ki = back ? c.ekey0 : c.ekey1;
        const uint32_t* vi = back ? c.eval0 : c.eval1;
        uint32_t* ko = back ? c.ekey1 : c.ekey0;
        uint32_t* vo = back ? c.eval1 : c.eval0;
        CU(radix_pass(ki, vi, ko, vo, &w->d_cnt->n_bodies, w->cap_bodies, pass * 8, w->d_hist, st));
    }
    k_permute_gather<<<gb, TPB, 0, st>>>(c, c.eval1);
    k_permute_store<<<gb, TPB, 0, st>>>(c, w->d_ext, w->d_inv);
    if (w->nj) k_remap_joints<<<blocks_for(w, w->nj), TPB, 0, st>>>(c, w->d_jext);
    CU(cudaGetLastError());
    w->launches += 3 + (w->nj ? 1 : 0) + 4 * (radix_blocks(w->cap_bodies) > RADIX_INLINE_BLOCKS ? 3 : 2);
    return SFP_OK;
}

int broadphase_and_colour(SfpWorld* w, float dt) {
    Ctx& c = w->ctx;
    cudaStream_t st = w->stream;
    const bool multi = (w->p.flags & SFP_FLAG_MULTI_KERNEL) != 0u;
    const uint32_t gb = blocks_for(w, w->n);
    // the frames whose colouring starts from scratch also renew the body order
    if ((w->colour_age % SFP_RECOLOUR_PERIOD) == 0u) TRY(resort_bodies(w));
    const bool mark = !(w->p.flags & SFP_FLAG_BROADPHASE_EVERY_SUBSTEP);
    // warm start: read last frame's table, build this frame's in the other one
    c.warm = (w->colour_age % SFP_RECOLOUR_PERIOD) != 0u ? 1u : 0u;
    c.tab_old = w->d_tab[(w->colour_age + 1u) & 1u];
    c.tab_new = w->d_tab[w->colour_age & 1u];
    w->colour_age += 1;
    // The snapshot for the capacity recovery and the clearing of this frame's colour table feed
    // nothing before the colouring: they run on a side stream (a parallel branch of the captured
    // graph) beside the broadphase kernels, which are each too short to fill the device.
    CU(cudaEventRecord(w->ev_fork, st));
    CU(cudaStreamWaitEvent(w->side, w->ev_fork, 0));
    if (mark) {  // the state this frame starts from
        k_snapshot<<<gb, TPB, 0, w->side>>>(c, w->d_pose_bak, w->d_vel_bak);
        w->launches += 1;
    }
    k_clear_table<<<blocks_for(w, w->tab_slots), TPB, 0, w->side>>>(c, c.tab_new, w->tab_slots);
    w->launches += 1;
    CU(cudaEventRecord(w->ev_join, w->side));
    k_aabb_count<<<gb, TPB, 0, st>>>(c);  // (dt from the frame's parameter block)
    CU(scan_exclusive(c.cell_cnt, c.cell_off, &w->d_cnt->n_bodies, w->cap_bodies, w->d_scan_tiles,
                      &w->d_cnt->n_entries, st));
    k_emit_entries<<<gb, TPB, 0, st>>>(c);
    TRY(sort_entries(w));
    const uint32_t ge = blocks_for(w, w->cap_entries);
    k_pair_count<<<ge, TPB, 0, st>>>(c);
    CU(scan_exclusive(c.pcnt, c.poff, &w->d_cnt->n_entries, w->cap_entries, w->d_scan_tiles,
                      &w->d_cnt->n_pairs, st));
    k_pair_write<<<ge, TPB, 0, st>>>(c);
    CU(cudaGetLastError());
    w->launches += 4 + 3 + 3 + 2 * SFP_KEY_PASSES +
                   (radix_blocks(w->cap_entries) > RADIX_INLINE_BLOCKS ? SFP_KEY_PASSES : 0);
    if (mark) CU(cudaEventRecordWithFlags(w->ev_bp, st, w->capturing ? cudaEventRecordExternal : 0));

    CU(cudaStreamWaitEvent(st, w->ev_join, 0));
    if (!multi) {
        // a grid barrier costs 1.2 us with 2-3 CTAs per SM and ~3 us with 8, and the kernel is a
        // dozen barriers around loops that are short for a 100 k-body world (measured: 3 CTAs
        // per SM best there, 8 best for 4 M bodies) — so the grid grows with the world
        const uint32_t full = blocks_for(w, 4ull * w->n + w->nj);
        const uint32_t want = std::min<uint32_t>(full, std::max<uint32_t>(3u * (uint32_t)w->n_sm, full / 8u));
        uint32_t blocks = std::min<uint32_t>((uint32_t)w->coop_max_colour, want);
        void* args[] = {&c};
        TRY(launch_coop(w, (const void*)k_colour_coop, blocks, args));
        w->launches += 1;
    } else {
        const uint32_t gp = blocks_for(w, (uint64_t)w->cap_pairs + w->nj);
        k_colour_init<<<gp, TPB, 0, st>>>(c);
        if (c.warm) k_colour_warm<<<gp, TPB, 0, st>>>(c);
        w->launches += 1 + c.warm;
        for (uint32_t round = 0; round < (uint32_t)MAX_ROUNDS; ++round) {
            k_colour_bid<<<gp, TPB, 0, st>>>(c, round);
            k_colour_take<<<gp, TPB, 0, st>>>(c, round);
            w->launches += 2;
            uint32_t flag = 0;
            CU(cudaMemcpyAsync(&flag, &w->d_cnt->round_flags[round], sizeof(flag),
                               cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            if (!flag) break;
        }
        k_colour_record<<<gp, TPB, 0, st>>>(c);
        w->launches += 1;
    }
    // stable sort of the pairs by (group, colour, narrowphase class): 8 bits without tiles (one
    // radix pass), 8 more per 256 tiles.  The passes alternate between the Ctx's buffers and their
    // partners such that the LAST one writes the Ctx's.
    {
        const uint32_t *ki = c.pcol, *vi = nullptr;
        for (int pass = 0; pass < w->sort_passes; ++pass) {
            const bool fin = ((w->sort_passes - 1 - pass) & 1) == 0;
            uint32_t* ko = fin ? c.pcol_sorted : w->d_pkeyB;
            uint32_t* vo = fin ? c.pperm : w->d_ppermB;
            CU(radix_pass(ki, vi, ko, vo, &w->d_cnt->n_pairs, w->cap_pairs, pass * 8, w->d_hist, st));
            ki = ko;
            vi = vo;
        }
        if (w->nj) {
            ki = c.jcol, vi = nullptr;
            for (int pass = 0; pass < w->sort_passes; ++pass) {
                const bool fin = ((w->sort_passes - 1 - pass) & 1) == 0;
                uint32_t* ko = fin ? c.jcol_sorted : w->d_jkeyB;
                uint32_t* vo = fin ? c.jperm : w->d_jpermB;
                CU(radix_pass(ki, vi, ko, vo, &w->d_cnt->n_joints, w->cap_joints, pass * 8, w->d_hist, st));
                ki = ko;
                vi = vo;
            }
        }
    }
    k_gather<<<blocks_for(w, (uint64_t)w->cap_pairs + w->nj), TPB, 0, st>>>(c);
    CU(cudaGetLastError());
    w->launches += 1 + w->sort_passes * ((radix_blocks(w->cap_pairs) > RADIX_INLINE_BLOCKS ? 3 : 2) +
                                         (w->nj ? (radix_blocks(w->cap_joints) > RADIX_INLINE_BLOCKS ? 3 : 2) : 0));
    if (mark) CU(cudaEventRecordWithFlags(w->ev_col, st, w->capturing ? cudaEventRecordExternal : 0));
    return SFP_OK;
}

int read_counters(SfpWorld* w, Counters* out);

int run_substeps(SfpWorld* w, float h, uint32_t substeps) {
    Ctx& c = w->ctx;
    cudaStream_t st = w->stream;
    const bool multi = (w->p.flags & SFP_FLAG_MULTI_KERNEL) != 0u;
    const float g = sqrtf(w->p.gravity_x * w->p.gravity_x + w->p.gravity_y * w->p.gravity_y);
    float rest_threshold = 2.0f * g * h;
    if (!multi && (w->p.flags & SFP_FLAG_FLOW_SOLVE)) {
        const uint32_t want = (uint32_t)((4ull * w->n + w->nj + SFP_SUB_TPB - 1) / SFP_SUB_TPB) + 1u;
        void* args[] = {&c, &substeps};
        w->launches += 1;
        if (w->has_polygons)
            return launch_coop(w, (const void*)k_substeps_flow<true>,
                               std::min<uint32_t>(want, (uint32_t)w->coop_max_flow_poly), args, SFP_SUB_TPB);
        return launch_coop(w, (const void*)k_substeps_flow<false>,
                           std::min<uint32_t>(want, (uint32_t)w->coop_max_flow), args, SFP_SUB_TPB);
    }
    if (!multi) {
        uint32_t blocks = std::min<uint32_t>(
            (uint32_t)w->coop_max_substep,
            (uint32_t)((4ull * w->n + w->nj + SFP_SUB_TPB - 1) / SFP_SUB_TPB) + 1u);
        void* args[] = {&c, &substeps};
        w->launches += 1;
        if (c.tile_b) {
            // one block per SM; every block walks its share of the tiles
            const size_t smem = (size_t)c.tile_b * TILE_SMEM_PER_BODY;
            const int cap = w->has_polygons ? w->coop_max_tiles_poly : w->coop_max_tiles;
            const void* fn = w->has_polygons ? (const void*)k_substeps_coop<true, true>
                                             : (const void*)k_substeps_coop<false, true>;
            const unsigned tblocks = (unsigned)std::max<int>(w->n_sm, std::min<int>(cap, (int)c.n_tiles));
            CU(cudaLaunchCooperativeKernel(fn, dim3(std::min<unsigned>(tblocks, (unsigned)cap)), dim3(SFP_SUB_TPB),
                                           args, smem, w->stream));
            return SFP_OK;
        }
        if (w->has_polygons) {
            blocks = std::min<uint32_t>(blocks, (uint32_t)w->coop_max_substep_poly);
            return launch_coop(w, (const void*)k_substeps_coop<true, false>, blocks, args, SFP_SUB_TPB);
        }
        return launch_coop(w, (const void*)k_substeps_coop<false, false>, blocks, args, SFP_SUB_TPB);
    }
    Counters k;
    TRY(read_counters(w, &k));
    const uint32_t ncol = k.n_colours;
    std::vector<uint32_t> order;  // big colours, then small ones (SFP_TAIL_ITEMS)
    for (int pass = 0; pass < 2; ++pass)
        for (uint32_t col = 0; col < ncol; ++col) {
            uint32_t items = (k.pair_col_off[col + 1] - k.pair_col_off[col]) +
                             (k.joint_col_off[col + 1] - k.joint_col_off[col]);
            if (col != SFP_SERIAL_COLOUR && (items > SFP_TAIL_ITEMS) == (pass == 0))
                order.push_back(col);
        }
    const bool serial = ncol > SFP_SERIAL_COLOUR;
    if (serial) k_serial_sort<<<1, TPB, 0, st>>>(c);
    const float inv_h = 1.0f / h, inv_h2 = inv_h * inv_h;
    const uint32_t gb = blocks_for(w, w->n), gp = blocks_for(w, (uint64_t)w->cap_pairs + w->nj);
    for (uint32_t s = 0; s < substeps; ++s) {
        k_integrate<<<gb, TPB, 0, st>>>(c, h);
        if (w->has_polygons) k_narrowphase<true><<<gp, TPB, 0, st>>>(c);
        else k_narrowphase<false><<<gp, TPB, 0, st>>>(c);
        for (uint32_t col : order) k_solve_pos<<<gp, TPB, 0, st>>>(c, col, inv_h2);
        if (serial) k_serial_pos<<<1, 32, 0, st>>>(c, inv_h2);
        k_derive<<<gb, TPB, 0, st>>>(c, inv_h);
        for (uint32_t col : order) k_solve_vel<<<gp, TPB, 0, st>>>(c, col, h, rest_threshold);
        if (serial) k_serial_vel<<<1, 32, 0, st>>>(c, h, rest_threshold);
        w->launches += 3 + 2 * ncol;
    }
    CU(cudaGetLastError());
    return SFP_OK;
}

int frame(SfpWorld* w, float dt, uint32_t substeps) {
    const float h = dt / (float)substeps;
    if (w->n == 0) return SFP_OK;
    if (w->p.flags & SFP_FLAG_BROADPHASE_EVERY_SUBSTEP) {
        for (uint32_t s = 0; s < substeps; ++s) {
            TRY(broadphase_and_colour(w, h));
            TRY(run_substeps(w, h, 1));
        }
        return SFP_OK;
    }
    TRY(broadphase_and_colour(w, dt));
    return run_substeps(w, h, substeps);
}

void drop_graphs(SfpWorld* w) {
    for (auto& g : w->graphs) {
        if (g.exec) cudaGraphExecDestroy(g.exec);
        g = SfpWorld::FrameGraph();
    }
}

// one cudaGraphLaunch per frame instead of ~30 kernel launches
int frame_graphed(SfpWorld* w, float dt, uint32_t substeps) {
    // this frame's step lengths go to device memory ahead of the launches (pageable source: the
    // driver stages it before the call returns), so one captured graph serves every dt
    {
        const float h = dt / (float)substeps;
        const float g = sqrtf(w->p.gravity_x * w->p.gravity_x + w->p.gravity_y * w->p.gravity_y);
        FrameParams fp;
        fp.dt = (w->p.flags & SFP_FLAG_BROADPHASE_EVERY_SUBSTEP) ? h : dt;
        fp.h = h;
        fp.rest_threshold = 2.0f * g * h;
        fp.pad = 0.0f;
        CU(cudaMemcpyAsync(w->d_fp, &fp, sizeof(fp), cudaMemcpyHostToDevice, w->stream));
    }
    const bool direct = !w->graphs_ok || w->n == 0 ||
                        (w->p.flags & (SFP_FLAG_MULTI_KERNEL | SFP_FLAG_BROADPHASE_EVERY_SUBSTEP));
    if (direct) return frame(w, dt, substeps);
    const uint32_t parity = w->colour_age & 1u;
    const uint32_t warm = (w->colour_age % SFP_RECOLOUR_PERIOD) != 0u ? 1u : 0u;
    for (auto& g : w->graphs)
        if (g.exec && g.substeps == substeps && g.parity == parity && g.warm == warm) {
            CU(cudaGraphLaunch(g.exec, w->stream));
            w->colour_age += 1;
            w->launches = g.launches;
            return SFP_OK;
        }
    // capture
    const uint32_t age0 = w->colour_age;
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(w->stream, cudaStreamCaptureModeThreadLocal);
    int rc = SFP_E_CUDA;
    if (e == cudaSuccess) {
        w->capturing = true;
        rc = frame(w, dt, substeps);
        w->capturing = false;
        e = cudaStreamEndCapture(w->stream, &graph);
    }
    cudaGraphExec_t exec = nullptr;
    if (rc == SFP_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (rc != SFP_OK || e != cudaSuccess || !exec) {
        // capture refused (e.g. a driver without cooperative-launch capture): plain launches
        cudaGetLastError();
        w->graphs_ok = false;
        w->sticky = SFP_OK;
        w->err.clear();
        w->colour_age = age0;
        return frame(w, dt, substeps);
    }
    // (substeps, table parity, warm): 2 x 2 combinations per substep count; six slots hold a
    // host that alternates between two substep counts
    auto& slot = w->graphs[w->graph_next++ % 6u];
    if (slot.exec) cudaGraphExecDestroy(slot.exec);
    slot.exec = exec;
    slot.substeps = substeps;
    slot.parity = parity;
    slot.warm = warm;
    slot.launches = w->launches;
    CU(cudaGraphLaunch(exec, w->stream));
    return SFP_OK;
}

int read_counters(SfpWorld* w, Counters* out) {
    CU(cudaMemcpyAsync(out, w->d_cnt, sizeof(Counters), cudaMemcpyDeviceToHost, w->stream));
    CU(cudaStreamSynchronize(w->stream));
    return SFP_OK;
}

int capacity_error(SfpWorld* w, uint32_t bits) {
    w->stats.overflow |= bits;
    char buf[200];
    snprintf(buf, sizeof(buf),
             "device capacity exceeded (bits %u: 1=entries cap %u, 2=pairs cap %u, 4=serial group>1024, "
             "8=colour rounds, 16=runaway or non-finite body, 32=solver wait)",
             bits, w->cap_entries, w->cap_pairs);
    w->err = buf;
    return SFP_E_CAPACITY;
}

// The broadphase ran out of entry / pair room in some frame since the last synchronisation.
// Every kernel after that point was a no-op (OVERFLOW_CAPACITY), so the device still holds the
// state that frame started from (k_snapshot), last frame's colours and the body order.  Enlarge
// the buffers, carry the colour tables over, put the state back and run the lost frames again
// — with the same dt, substeps and colouring age, so the result is what a large enough world
// would have produced (and what the oracle, which has no capacities, produces).
int recover_capacity(SfpWorld* w, Counters* k) {
    for (int attempt = 0; attempt < 12; ++attempt) {
        if (!(k->overflow & OVERFLOW_CAPACITY) || (k->overflow & ~OVERFLOW_CAPACITY)) break;
        const bool fixed_e = (k->overflow & 1u) && w->p.max_entries != 0u;
        const bool fixed_p = (k->overflow & 2u) && w->p.max_pairs != 0u;
        // explicit capacities are a contract, and the debug modes keep the old behaviour
        if (fixed_e || fixed_p || (w->p.flags & (SFP_FLAG_BROADPHASE_EVERY_SUBSTEP | SFP_FLAG_MULTI_KERNEL)))
            break;
        const uint64_t lost = w->frames_launched - (uint64_t)k->frames_done;
        if (lost == 0 || lost > 1024u) break;
        const uint64_t ce64 = (k->overflow & 1u) ? 2ull * w->cap_entries : w->cap_entries;
        const uint64_t cp64 = (k->overflow & 2u) ? 2ull * w->cap_pairs : w->cap_pairs;
        if (ce64 > 0x7fffffffull || cp64 > 0x7fffffffull) break;
        CU(cudaStreamSynchronize(w->stream));
        drop_graphs(w);  // the buffers are kernel arguments
        // keep the old tables until they are rehashed
        std::vector<void*> old;
        old.swap(w->pair_allocs);
        unsigned long long* old_tab[2] = {w->d_tab[0], w->d_tab[1]};
        const uint32_t old_slots = w->tab_slots;
        int rc = alloc_pair_arrays(w, (uint32_t)ce64, (uint32_t)cp64);
        if (rc != SFP_OK) {
            free_bag(old);
            return rc;
        }
        for (int t = 0; t < 2; ++t)
            k_rehash<<<blocks_for(w, old_slots), TPB, 0, w->stream>>>(old_tab[t], old_slots, w->d_tab[t],
                                                                     w->tab_slots - 1u);
        CU(cudaMemcpyAsync(w->ctx.pose, w->d_pose_bak, sizeof(float4) * w->n, cudaMemcpyDeviceToDevice,
                           w->stream));
        CU(cudaMemcpyAsync(w->ctx.vel, w->d_vel_bak, sizeof(float4) * w->n, cudaMemcpyDeviceToDevice,
                           w->stream));
        const uint32_t zero = 0;
        CU(cudaMemcpyAsync(&w->d_cnt->overflow, &zero, sizeof(zero), cudaMemcpyHostToDevice, w->stream));
        CU(cudaStreamSynchronize(w->stream));
        free_bag(old);
        w->grows += 1;
        // run the lost frames again
        const uint64_t first = w->frames_launched - lost;
        w->colour_age = w->frame_log[first % 1024u].colour_age;
        for (uint64_t f = first; f < w->frames_launched; ++f) {
            const SfpWorld::FrameLog& e = w->frame_log[f % 1024u];
            w->launches = 0;
            TRY(frame_graphed(w, e.dt, e.substeps));
        }
        TRY(read_counters(w, k));
    }
    return SFP_OK;
}

// reports errors; recovers from a capacity overflow if there was one
int check_overflow(SfpWorld* w, Counters& k) {
    if (k.overflow & OVERFLOW_CAPACITY) TRY(recover_capacity(w, &k));
    if (k.overflow) return capacity_error(w, k.overflow);
    return SFP_OK;
}
// Waits for the queued frames and makes sure they really happened (capacity recovery) BEFORE
// anything reads their results or changes the world underneath them.
int settle(SfpWorld* w, Counters* k_out = nullptr) {
    Counters k;
    TRY(read_counters(w, &k));
    TRY(check_overflow(w, k));
    if (k_out) *k_out = k;
    return SFP_OK;
}

}  // namespace

extern "C" {

int32_t sfp_default_params(SfpParams* out) {
    if (!out) return SFP_E_INVALID;
    sfp::default_params(out);
    return SFP_OK;
}

int32_t sfp_create(const SfpParams* params, SfpWorld** out) {
    if (!out) return SFP_E_INVALID;
    *out = nullptr;
    SfpParams p;
    if (params) p = *params;
    else sfp::default_params(&p);
    if (!sfp::params_valid(p)) return SFP_E_INVALID;
    if (p.solver_order != SFP_ORDER_COLOUR) return SFP_E_INVALID;  // the device path is coloured GS
    // the tile order lives in the persistent barrier kernel only
    if ((p.flags & SFP_FLAG_TILES) && (p.flags & (SFP_FLAG_MULTI_KERNEL | SFP_FLAG_FLOW_SOLVE)))
        return SFP_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        fprintf(stderr, "libsfp_b200: no CUDA device — this library has no CPU fallback\n");
        return SFP_E_CUDA;
    }
    if (p.device < 0 || p.device >= ndev) return SFP_E_INVALID;
    SfpWorld* w = new (std::nothrow) SfpWorld();
    if (!w) return SFP_E_NOMEM;
    w->p = p;
    w->device = p.device;
    std::memset(&w->ctx, 0, sizeof(w->ctx));
    std::memset(&w->stats, 0, sizeof(w->stats));
    auto fail = [&](const char* what, cudaError_t e) {
        fprintf(stderr, "libsfp_b200: %s: %s\n", what, cudaGetErrorString(e));
        delete w;
        return (int32_t)SFP_E_CUDA;
    };
    cudaError_t e;
    if ((e = cudaSetDevice(w->device)) != cudaSuccess) return fail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, w->device)) != cudaSuccess)
        return fail("cudaGetDeviceProperties", e);
    if (!prop.cooperativeLaunch) return fail("cooperative launch unsupported", cudaErrorNotSupported);
    w->n_sm = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking)) != cudaSuccess)
        return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreate(&w->ev0)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&w->ev1)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&w->ev_bp)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&w->ev_col)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaStreamCreateWithFlags(&w->side, cudaStreamNonBlocking)) != cudaSuccess)
        return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&w->ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&w->ev_join, cudaEventDisableTiming)) != cudaSuccess)
        return fail("cudaEventCreate", e);
    if ((e = cudaMalloc(&w->d_cnt, sizeof(Counters))) != cudaSuccess) return fail("cudaMalloc", e);
    if ((e = cudaMemset(w->d_cnt, 0, sizeof(Counters))) != cudaSuccess) return fail("cudaMemset", e);
    w->ctx.cnt = w->d_cnt;
    if ((e = cudaMalloc(&w->d_fp, sizeof(FrameParams))) != cudaSuccess) return fail("cudaMalloc", e);
    w->ctx.fp = w->d_fp;
    int per_sm = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_colour_coop, TPB, 0)) !=
        cudaSuccess)
        return fail("occupancy(colour)", e);
    w->coop_max_colour = std::max(1, per_sm) * w->n_sm;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_substeps_coop<false, false>, SFP_SUB_TPB, 0)) !=
        cudaSuccess)
        return fail("occupancy(substeps)", e);
    w->coop_max_substep = std::max(1, per_sm) * w->n_sm;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_substeps_coop<true, false>, SFP_SUB_TPB, 0)) !=
        cudaSuccess)
        return fail("occupancy(substeps, polygons)", e);
    w->coop_max_substep_poly = std::max(1, per_sm) * w->n_sm;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_substeps_flow<false>, SFP_SUB_TPB, 0)) !=
        cudaSuccess)
        return fail("occupancy(flow)", e);
    w->coop_max_flow = std::max(1, per_sm) * w->n_sm;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_substeps_flow<true>, SFP_SUB_TPB, 0)) !=
        cudaSuccess)
        return fail("occupancy(flow, polygons)", e);
    w->coop_max_flow_poly = std::max(1, per_sm) * w->n_sm;
    const size_t tile_smem = (size_t)SFP_TILE_MAX * TILE_SMEM_PER_BODY;
    if ((e = cudaFuncSetAttribute(k_substeps_coop<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)tile_smem)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_substeps_coop<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)tile_smem)) != cudaSuccess)
        return fail("cudaFuncSetAttribute(tiles)", e);
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_substeps_coop<false, true>, SFP_SUB_TPB,
                                                           tile_smem)) != cudaSuccess)
        return fail("occupancy(tiles)", e);
    w->coop_max_tiles = std::max(1, per_sm) * w->n_sm;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_substeps_coop<true, true>, SFP_SUB_TPB,
                                                           tile_smem)) != cudaSuccess)
        return fail("occupancy(tiles, polygons)", e);
    w->coop_max_tiles_poly = std::max(1, per_sm) * w->n_sm;
    w->ctx.gx = p.gravity_x;
    w->ctx.gy = p.gravity_y;
    w->ctx.inv_cell = 1.0f / p.cell_size;
    w->ctx.aabb_margin = p.aabb_margin;
    w->ctx.vel_expand = p.vel_expand;
    w->ctx.contact_margin = p.contact_margin;
    *out = w;
    return SFP_OK;
}

int32_t sfp_destroy(SfpWorld* w) {
    if (!w) return SFP_OK;
    cudaSetDevice(w->device);
    if (w->stream) cudaStreamSynchronize(w->stream);
    drop_graphs(w);
    free_bag(w->body_allocs);
    free_bag(w->pair_allocs);
    free_bag(w->joint_allocs);
    if (w->d_cnt) cudaFree(w->d_cnt);
    if (w->d_fp) cudaFree(w->d_fp);
    if (w->pin_state) cudaFreeHost(w->pin_state);
    if (w->d_rays) cudaFree(w->d_rays);
    if (w->d_hits) cudaFree(w->d_hits);
    if (w->d_polytab) cudaFree(w->d_polytab);
    if (w->ev0) cudaEventDestroy(w->ev0);
    if (w->ev1) cudaEventDestroy(w->ev1);
    if (w->ev_bp) cudaEventDestroy(w->ev_bp);
    if (w->ev_col) cudaEventDestroy(w->ev_col);
    if (w->ev_fork) cudaEventDestroy(w->ev_fork);
    if (w->ev_join) cudaEventDestroy(w->ev_join);
    if (w->side) cudaStreamDestroy(w->side);
    if (w->stream) cudaStreamDestroy(w->stream);
    delete w;
    return SFP_OK;
}

int32_t sfp_upload_bodies(SfpWorld* w, const SfpBodyDesc* bodies, uint32_t n) {
    if (!w || (n && !bodies)) return SFP_E_INVALID;
    if (n > SFP_MAX_BODIES || (w->p.max_bodies && n > w->p.max_bodies)) {
        w->err = "too many bodies";
        return SFP_E_CAPACITY;
    }
    for (uint32_t i = 0; i < n; ++i)
        if (!sfp::desc_valid(bodies[i], &w->polys)) {
            w->err = "invalid body desc at index " + std::to_string(i);
            return SFP_E_INVALID;
        }
    if (w->nj && n < w->n) {
        w->err = "shrinking the body set while joints exist: clear the joints first";
        return SFP_E_INVALID;
    }
    CU(cudaSetDevice(w->device));
    drop_graphs(w);
    TRY(ensure_body_capacity(w, n));
    std::vector<sfp::f4> pose(n), vel(n);
    w->h_minv.resize(n);
    w->h_shape.resize(n);
    w->h_mat.resize(n);
    uint32_t ndyn = 0;
    bool polys_used = false;
    for (uint32_t i = 0; i < n; ++i) {
        sfp::body_from_desc(bodies[i], &pose[i], &vel[i], &w->h_minv[i], &w->h_shape[i],
                            &w->h_mat[i], &w->polys);
        ndyn += (w->h_minv[i].x > 0.0f || w->h_minv[i].y > 0.0f) ? 1u : 0u;
        polys_used |= (bodies[i].shape & 0xffu) == SFP_SHAPE_POLYGON;
        if (!sfp::body_fits_grid(pose[i], w->h_shape[i], w->p)) {
            w->err = "body " + std::to_string(i) + " spans more than 16384 broadphase cells (or 4096 along one axis) at this cell_size: split it or raise cell_size";
            return SFP_E_INVALID;
        }
    }
    w->has_polygons = polys_used;
    Ctx& c = w->ctx;
    cudaStream_t st = w->stream;
    CU(cudaStreamSynchronize(st));
    if (n) {
        CU(cudaMemcpyAsync(c.pose, pose.data(), sizeof(sfp::f4) * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(c.prev, pose.data(), sizeof(sfp::f4) * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(c.vel, vel.data(), sizeof(sfp::f4) * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(const_cast<float4*>(c.shape), w->h_shape.data(), sizeof(sfp::f4) * n,
                           cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(const_cast<float4*>(c.mat), w->h_mat.data(), sizeof(sfp::f4) * n,
                           cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(const_cast<float2*>(c.minv), w->h_minv.data(), sizeof(sfp::f2) * n,
                           cudaMemcpyHostToDevice, st));
    }
    w->n = n;
    c.n = n;
    if (n) k_identity<<<blocks_for(w, n), TPB, 0, st>>>(w->d_ext, w->d_inv, n);
    if (w->nj) k_remap_joints<<<blocks_for(w, w->nj), TPB, 0, st>>>(c, w->d_jext);
    CU(cudaGetLastError());
    Counters zero;
    std::memset(&zero, 0, sizeof(zero));
    zero.n_bodies = n;
    zero.n_joints = w->nj;
    CU(cudaMemcpyAsync(w->d_cnt, &zero, sizeof(zero), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));  // the host vectors above die at return
    w->stats.n_bodies = n;
    w->stats.n_dynamic = ndyn;
    w->colour_age = 0;
    w->stepped = false;
    w->frames_launched = 0;  // (the device's frames_done was zeroed with the counters)
    // tiles of the solve order (sfp_math.h): after the first frame's re-sort the dynamic bodies
    // are the indices [0, ndyn)
    c.n_dyn = ndyn;
    c.tile_b = (w->p.flags & SFP_FLAG_TILES) && ndyn ? sfp::tile_size(ndyn) : 0u;
    c.n_tiles = c.tile_b ? (ndyn + c.tile_b - 1u) / c.tile_b : 0u;
    w->sort_passes = c.n_tiles == 0u ? 1 : (c.n_tiles + 1u <= 256u ? 2 : 3);
    return SFP_OK;
}

int32_t sfp_set_polygons(SfpWorld* w, const SfpPolygon* polys, uint32_t n) {
    if (!w || (n && !polys)) return SFP_E_INVALID;
    sfp::PolyTable t;
    long bad = sfp::build_poly_table(polys, n, &t);
    if (bad >= 0) {
        w->err = "invalid polygon at index " + std::to_string(bad) +
                 " (3..8 vertices, counter-clockwise, strictly convex, centroid at the origin)";
        return SFP_E_INVALID;
    }
    CU(cudaSetDevice(w->device));
    CU(cudaStreamSynchronize(w->stream));
    drop_graphs(w);  // the table pointer is a kernel argument
    if (w->d_polytab) cudaFree(w->d_polytab);
    w->d_polytab = nullptr;
    if (n) {
        CU(cudaMalloc(&w->d_polytab, sizeof(float4) * 8u * (size_t)n));
        CU(cudaMemcpy(w->d_polytab, t.recs.data(), sizeof(float4) * 8u * (size_t)n,
                      cudaMemcpyHostToDevice));
    }
    w->ctx.polytab = w->d_polytab;
    w->polys = std::move(t);
    return SFP_OK;
}

int32_t sfp_set_joints(SfpWorld* w, const SfpJointDesc* joints, uint32_t nj) {
    if (!w || (nj && !joints)) return SFP_E_INVALID;
    if (w->p.max_joints && nj > w->p.max_joints) return SFP_E_CAPACITY;
    for (uint32_t i = 0; i < nj; ++i)
        if (!sfp::joint_valid(joints[i], w->n)) {
            w->err = "invalid joint desc at index " + std::to_string(i);
            return SFP_E_INVALID;
        }
    CU(cudaSetDevice(w->device));
    if (w->stepped) TRY(settle(w));
    drop_graphs(w);
    TRY(ensure_joint_capacity(w, nj));
    std::vector<uint2> jb(nj);
    std::vector<float4> ja(nj), jp(nj);
    for (uint32_t i = 0; i < nj; ++i) {
        jb[i] = make_uint2(joints[i].body_a, joints[i].body_b);
        ja[i] = make_float4(joints[i].ax, joints[i].ay, joints[i].bx, joints[i].by);
        jp[i] = make_float4(joints[i].rest, joints[i].compliance, sfp::u2f(joints[i].kind), 0.0f);
    }
    Ctx& c = w->ctx;
    cudaStream_t st = w->stream;
    CU(cudaStreamSynchronize(st));
    if (nj) {
        CU(cudaMemcpyAsync(w->d_jext, jb.data(), sizeof(uint2) * nj, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(const_cast<float4*>(c.janc_in), ja.data(), sizeof(float4) * nj,
                           cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(const_cast<float4*>(c.jpar_in), jp.data(), sizeof(float4) * nj,
                           cudaMemcpyHostToDevice, st));
    }
    w->nj = nj;
    c.nj = nj;
    if (nj) {
        k_remap_joints<<<blocks_for(w, nj), TPB, 0, st>>>(c, w->d_jext);
        CU(cudaGetLastError());
    }
    CU(cudaMemcpyAsync(&w->d_cnt->n_joints, &nj, sizeof(nj), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
    w->stats.n_joints = nj;
    w->colour_age = 0;
    return SFP_OK;
}

int32_t sfp_step(SfpWorld* w, float dt, uint32_t substeps) {
    if (!w || !(dt > 0.0f) || substeps == 0 || substeps > 1024) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    CU(cudaEventRecord(w->ev0, w->stream));
    w->launches = 0;
    w->frame_log[w->frames_launched % 1024u] = {dt, substeps, w->colour_age};
    w->frames_launched += 1;
    int rc = frame_graphed(w, dt, substeps);
    if (rc != SFP_OK) return rc;
    CU(cudaEventRecord(w->ev1, w->stream));
    w->timed = true;
    w->stats.frames += 1;
    w->stats.body_substeps += (uint64_t)w->n * substeps;
    w->stepped = true;
    w->last_h = dt / (float)substeps;
    return SFP_OK;
}

int32_t sfp_sync(SfpWorld* w) {
    if (!w) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    return settle(w);
}

namespace {
// true if p is page-locked host memory the device can DMA from / to directly
bool host_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}
}  // namespace

int32_t sfp_alloc_host_states(uint32_t n, SfpBodyState** out) {
    if (!out || n == 0) return SFP_E_INVALID;
    void* p = nullptr;
    if (cudaMallocHost(&p, sizeof(SfpBodyState) * (size_t)n) != cudaSuccess) return SFP_E_NOMEM;
    *out = static_cast<SfpBodyState*>(p);
    return SFP_OK;
}
int32_t sfp_free_host_states(SfpBodyState* p) {
    if (p && cudaFreeHost(p) != cudaSuccess) return SFP_E_CUDA;
    return SFP_OK;
}

int32_t sfp_download_bodies(SfpWorld* w, SfpBodyState* out, uint32_t n) {
    if (!w || (n && !out) || n > w->n) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    TRY(settle(w));
    const bool direct = n && host_pinned(out);
    if (n) {
        // states are packed on the device (incl. the angle), one D2H copy, no per-body host work
        k_pack_states<<<blocks_for(w, n), TPB, 0, w->stream>>>(w->ctx.pose, w->ctx.vel, w->d_inv, w->d_state, n);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(direct ? out : w->pin_state, w->d_state, sizeof(SfpBodyState) * n,
                           cudaMemcpyDeviceToHost, w->stream));
        CU(cudaStreamSynchronize(w->stream));
    }
    if (n && !direct) std::memcpy(out, w->pin_state, sizeof(SfpBodyState) * n);
    return SFP_OK;
}

int32_t sfp_step_host(SfpWorld* w, const SfpBodyState* in, SfpBodyState* out, uint32_t n, float dt,
                      uint32_t substeps) {
    if (!w || !in || !out || n != w->n || n == 0) return SFP_E_INVALID;
    if (!(dt > 0.0f) || substeps == 0 || substeps > 1024) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    TRY(settle(w));  // (also: the staging buffer is reused)
    // (a state that carries its rotation in `angle` only is given its rotor by the unpack
    // kernel: no pass over the states on the host)
    const SfpBodyState* src = in;
    if (!host_pinned(in)) {
        std::memcpy(w->pin_state, in, sizeof(SfpBodyState) * n);
        src = w->pin_state;
    }
    CU(cudaMemcpyAsync(w->d_state, src, sizeof(SfpBodyState) * n, cudaMemcpyHostToDevice,
                       w->stream));
    k_unpack_states<<<blocks_for(w, n), TPB, 0, w->stream>>>(w->d_state, w->ctx.minv, w->d_inv,
                                                            w->ctx.pose, w->ctx.vel, n);
    CU(cudaGetLastError());
    TRY(sfp_step(w, dt, substeps));
    return sfp_download_bodies(w, out, n);
}

namespace {
struct HostPairs {
    std::vector<uint2> pairs;       // colour-sorted device order, INTERNAL body indices
    std::vector<uint2> ext;         // the same pairs in the caller's numbering (a < b)
    std::vector<uint32_t> colour;   // per pair
    std::vector<uint32_t> order;    // permutation that sorts by the caller's (a,b)
    std::vector<uint32_t> ext_of;   // internal -> caller's index
    Counters k;
};
int fetch_pairs(SfpWorld* w, HostPairs* hp) {
    TRY(settle(w, &hp->k));
    uint32_t P = w->stepped ? hp->k.n_pairs : 0u;
    hp->pairs.resize(P);
    hp->ext_of.resize(w->n);
    if (w->n)
        CU(cudaMemcpyAsync(hp->ext_of.data(), w->d_ext, sizeof(uint32_t) * w->n, cudaMemcpyDeviceToHost,
                           w->stream));
    if (P)
        CU(cudaMemcpyAsync(hp->pairs.data(), w->ctx.pairs, sizeof(uint2) * P, cudaMemcpyDeviceToHost,
                           w->stream));
    CU(cudaStreamSynchronize(w->stream));
    hp->ext.resize(P);
    for (uint32_t i = 0; i < P; ++i)
        hp->ext[i] = make_uint2(hp->ext_of[hp->pairs[i].x], hp->ext_of[hp->pairs[i].y]);
    hp->colour.resize(P);
    if (P) {
        // colour as reported (sfp_math.h): from the sorted keys, interior colours + 64
        std::vector<uint32_t> keys(P);
        CU(cudaMemcpy(keys.data(), w->ctx.pcol_sorted, sizeof(uint32_t) * P, cudaMemcpyDeviceToHost));
        for (uint32_t i = 0; i < P; ++i) {
            const uint32_t col = (keys[i] >> SFP_TYPE_BITS) & 63u, group = (keys[i] >> 8) & 0xfffu;
            hp->colour[i] = col + (group != w->ctx.n_tiles ? SFP_CLASS_INTERIOR : 0u);
        }
    }
    hp->order.resize(P);
    for (uint32_t i = 0; i < P; ++i) hp->order[i] = i;
    std::sort(hp->order.begin(), hp->order.end(), [&](uint32_t x, uint32_t y) {
        const uint2 &a = hp->ext[x], &b = hp->ext[y];
        return a.x != b.x ? a.x < b.x : a.y < b.y;
    });
    return SFP_OK;
}
}  // namespace

int32_t sfp_get_pairs(SfpWorld* w, SfpPair* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    HostPairs hp;
    TRY(fetch_pairs(w, &hp));
    *n_out = (uint32_t)hp.pairs.size();
    for (uint32_t i = 0; i < cap && i < hp.pairs.size(); ++i) {
        out[i].a = hp.ext[hp.order[i]].x;
        out[i].b = hp.ext[hp.order[i]].y;
    }
    return SFP_OK;
}

int32_t sfp_get_colours(SfpWorld* w, uint32_t* pair_colours, uint32_t cap, uint32_t* joint_colours,
                        uint32_t jcap) {
    if (!w) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    HostPairs hp;
    TRY(fetch_pairs(w, &hp));
    if (pair_colours)
        for (uint32_t i = 0; i < cap && i < hp.pairs.size(); ++i)
            pair_colours[i] = hp.colour[hp.order[i]];
    if (joint_colours && w->nj && w->stepped) {
        std::vector<uint32_t> jc(w->nj);
        CU(cudaMemcpy(jc.data(), w->ctx.jcol, sizeof(uint32_t) * w->nj, cudaMemcpyDeviceToHost));
        for (uint32_t i = 0; i < jcap && i < w->nj; ++i) joint_colours[i] = jc[i] & 0xffu;
    }
    return SFP_OK;
}

int32_t sfp_get_contacts(SfpWorld* w, SfpContact* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    HostPairs hp;
    TRY(fetch_pairs(w, &hp));
    const uint32_t P = (uint32_t)hp.pairs.size();
    std::vector<sfp::f4> ca(P), cb(P), cp0(P), cp1(P), cs(P), lam(P), pose(w->n);
    if (P) {
        const Ctx& c = w->ctx;
        const float4* src[6] = {c.ca, c.cb, c.cp0, c.cp1, c.cs, c.lam};
        sfp::f4* dst[6] = {ca.data(), cb.data(), cp0.data(), cp1.data(), cs.data(), lam.data()};
        for (int k = 0; k < 6; ++k)
            CU(cudaMemcpyAsync(dst[k], src[k], sizeof(sfp::f4) * P, cudaMemcpyDeviceToHost,
                               w->stream));
        CU(cudaMemcpyAsync(pose.data(), c.pose, sizeof(sfp::f4) * w->n, cudaMemcpyDeviceToHost,
                           w->stream));
        CU(cudaStreamSynchronize(w->stream));
    }
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < P; ++i) {
        uint32_t k = hp.order[i];
        if (sfp::f2u(cb[k].y) == 0u) continue;
        if (cnt < cap) {
            sfp::ContactRec r;
            r.a = ca[k];
            r.b = cb[k];
            r.p0 = cp0[k];
            r.p1 = cp1[k];
            r.s = cs[k];
            uint32_t a = hp.pairs[k].x, b = hp.pairs[k].y;
            out[cnt] = sfp::contact_from_records(hp.ext[k].x, hp.ext[k].y, r, lam[k], pose[a], pose[b]);
        }
        ++cnt;
    }
    *n_out = cnt;
    return SFP_OK;
}

int32_t sfp_get_contact_events(SfpWorld* w, SfpContactEvent* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    static_assert(sizeof(sfp::EventRec) == sizeof(SfpContactEvent), "EventRec mirrors the ABI");
    *n_out = 0;
    if (!w->stepped || w->n == 0) return SFP_OK;
    CU(cudaSetDevice(w->device));
    TRY(settle(w));
    cudaStream_t st = w->stream;
    const uint32_t gp = blocks_for(w, w->cap_pairs);
    k_event_flags<<<gp, TPB, 0, st>>>(w->ctx, w->d_evt_flag);
    // the scan's total goes to the spare slot behind the flags (device memory, read back below)
    CU(scan_exclusive(w->d_evt_flag, w->d_evt_off, &w->d_cnt->n_pairs, w->cap_pairs,
                      w->d_scan_tiles, w->d_evt_flag + w->cap_pairs, st));
    k_event_write<<<gp, TPB, 0, st>>>(w->ctx, w->d_evt_off, w->d_events, 1.0f / w->last_h);
    CU(cudaGetLastError());
    uint32_t total = 0;
    CU(cudaMemcpyAsync(&total, w->d_evt_flag + w->cap_pairs, sizeof(total), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    *n_out = total;
    const uint32_t m = std::min(total, cap);
    if (m == 0) return SFP_OK;
    std::vector<SfpContactEvent> ev(total);
    CU(cudaMemcpy(ev.data(), w->d_events, sizeof(SfpContactEvent) * total, cudaMemcpyDeviceToHost));
    std::sort(ev.begin(), ev.end(), [](const SfpContactEvent& x, const SfpContactEvent& y) {
        return x.a != y.a ? x.a < y.a : x.b < y.b;
    });
    std::memcpy(out, ev.data(), sizeof(SfpContactEvent) * m);
    return SFP_OK;
}

int32_t sfp_ray_cast(SfpWorld* w, const SfpRay* rays, uint32_t m, SfpRayHit* hits) {
    if (!w || (m && (!rays || !hits))) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    static_assert(sizeof(sfp::RayRec) == sizeof(SfpRay) && sizeof(sfp::RayHitRec) == sizeof(SfpRayHit),
                  "ray records mirror the ABI");
    if (m == 0) return SFP_OK;
    if (w->n == 0) {
        for (uint32_t k = 0; k < m; ++k) hits[k] = SfpRayHit{0xffffffffu, 0.0f, 0.0f, 0.0f};
        return SFP_OK;
    }
    CU(cudaSetDevice(w->device));
    TRY(settle(w));
    cudaStream_t st = w->stream;
    if (m > w->ray_cap) {
        CU(cudaStreamSynchronize(st));
        if (w->d_rays) cudaFree(w->d_rays);
        if (w->d_hits) cudaFree(w->d_hits);
        w->d_rays = nullptr;
        w->d_hits = nullptr;
        w->ray_cap = 0;
        CU(cudaMalloc(&w->d_rays, sizeof(sfp::RayRec) * (size_t)m));
        CU(cudaMalloc(&w->d_hits, sizeof(sfp::RayHitRec) * (size_t)m));
        w->ray_cap = m;
    }
    // spatial hash of the CURRENT poses (the frame's own hash is swept and one frame old);
    // it reuses the broadphase buffers, which are scratch between frames
    Ctx& c = w->ctx;
    const uint32_t gb = blocks_for(w, w->n);
    const int qinit[4] = {0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000};
    CU(cudaMemcpyAsync(w->d_cnt->qbox, qinit, sizeof(qinit), cudaMemcpyHostToDevice, st));
    k_query_aabbs<<<gb, TPB, 0, st>>>(c);
    CU(scan_exclusive(c.cell_cnt, c.cell_off, &w->d_cnt->n_bodies, w->cap_bodies, w->d_scan_tiles,
                      &w->d_cnt->n_entries, st));
    k_emit_entries<<<gb, TPB, 0, st>>>(c);
    TRY(sort_entries(w));
    CU(cudaMemcpyAsync(w->d_rays, rays, sizeof(SfpRay) * (size_t)m, cudaMemcpyHostToDevice, st));
    k_ray_cast<<<blocks_for(w, m), TPB, 0, st>>>(c, w->d_rays, m, w->d_hits, w->p.cell_size);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(hits, w->d_hits, sizeof(SfpRayHit) * (size_t)m, cudaMemcpyDeviceToHost, st));
    // the query's own hash can run out of entry room (bodies have moved since the frame's): the
    // entry buffers are then enlarged by the next frame's recovery; here it is an error
    Counters k;
    TRY(read_counters(w, &k));
    if (k.overflow) {
        const uint32_t zero = 0;
        if (!(k.overflow & ~OVERFLOW_CAPACITY))
            CU(cudaMemcpy(&w->d_cnt->overflow, &zero, sizeof(zero), cudaMemcpyHostToDevice));
        return capacity_error(w, k.overflow);
    }
    return SFP_OK;
}

int32_t sfp_query_aabb(SfpWorld* w, float lo_x, float lo_y, float hi_x, float hi_y, uint32_t mask,
                       uint32_t* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    *n_out = 0;
    if (w->n == 0) return SFP_OK;
    CU(cudaSetDevice(w->device));
    TRY(settle(w));
    cudaStream_t st = w->stream;
    const uint32_t gb = blocks_for(w, w->n);
    k_box_flags<<<gb, TPB, 0, st>>>(w->ctx, make_float4(lo_x, lo_y, hi_x, hi_y), mask, w->d_evt_flag);
    CU(scan_exclusive(w->d_evt_flag, w->d_evt_off, &w->d_cnt->n_bodies, w->cap_bodies,
                      w->d_scan_tiles, w->d_evt_flag + w->cap_pairs, st));
    k_box_write<<<gb, TPB, 0, st>>>(w->ctx, w->d_evt_off, w->d_idx);
    CU(cudaGetLastError());
    uint32_t total = 0;
    CU(cudaMemcpyAsync(&total, w->d_evt_flag + w->cap_pairs, sizeof(total), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    *n_out = total;
    const uint32_t mm = std::min(total, cap);
    if (mm) CU(cudaMemcpy(out, w->d_idx, sizeof(uint32_t) * mm, cudaMemcpyDeviceToHost));
    return SFP_OK;
}

int32_t sfp_query_shape(SfpWorld* w, const SfpBodyDesc* probe, uint32_t mask, uint32_t* out,
                        uint32_t cap, uint32_t* n_out) {
    if (!w || !probe || !n_out || (cap && !out)) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    *n_out = 0;
    sfp::f4 pp, ps;
    if (!sfp::probe_from_desc(*probe, &w->polys, &pp, &ps)) return SFP_E_INVALID;
    if (w->n == 0) return SFP_OK;
    CU(cudaSetDevice(w->device));
    TRY(settle(w));
    cudaStream_t st = w->stream;
    const uint32_t gb = blocks_for(w, w->n);
    k_shape_flags<<<gb, TPB, 0, st>>>(w->ctx, make_float4(pp.x, pp.y, pp.z, pp.w),
                                      make_float4(ps.x, ps.y, ps.z, ps.w), mask, w->d_evt_flag);
    CU(scan_exclusive(w->d_evt_flag, w->d_evt_off, &w->d_cnt->n_bodies, w->cap_bodies,
                      w->d_scan_tiles, w->d_evt_flag + w->cap_pairs, st));
    k_box_write<<<gb, TPB, 0, st>>>(w->ctx, w->d_evt_off, w->d_idx);
    CU(cudaGetLastError());
    uint32_t total = 0;
    CU(cudaMemcpyAsync(&total, w->d_evt_flag + w->cap_pairs, sizeof(total), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    *n_out = total;
    const uint32_t mm = std::min(total, cap);
    if (mm) CU(cudaMemcpy(out, w->d_idx, sizeof(uint32_t) * mm, cudaMemcpyDeviceToHost));
    return SFP_OK;
}

int32_t sfp_update_bodies(SfpWorld* w, const uint32_t* idx, const SfpBodyState* states, uint32_t m) {
    if (!w || (m && (!idx || !states)) || m > w->n) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    for (uint32_t k = 0; k < m; ++k)
        if (idx[k] >= w->n) return SFP_E_INVALID;
    if (m == 0) return SFP_OK;
    CU(cudaSetDevice(w->device));
    TRY(settle(w));  // (also: the staging buffer is reused)
    std::memcpy(w->pin_state, states, sizeof(SfpBodyState) * m);
    CU(cudaMemcpyAsync(w->d_state, w->pin_state, sizeof(SfpBodyState) * m, cudaMemcpyHostToDevice,
                       w->stream));
    CU(cudaMemcpyAsync(w->d_idx, idx, sizeof(uint32_t) * m, cudaMemcpyHostToDevice, w->stream));
    k_update_bodies<<<blocks_for(w, m), TPB, 0, w->stream>>>(w->d_idx, w->d_inv, w->d_state,
                                                            w->ctx.minv, w->ctx.pose, w->ctx.vel, m);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(w->stream));  // idx is the caller's memory
    return SFP_OK;
}

int32_t sfp_get_stats(SfpWorld* w, SfpStats* out) {
    if (!w || !out) return SFP_E_INVALID;
    if (w->sticky != SFP_OK) return w->sticky;
    CU(cudaSetDevice(w->device));
    Counters k;
    TRY(read_counters(w, &k));
    if (k.overflow & OVERFLOW_CAPACITY) TRY(recover_capacity(w, &k));
    w->stats.n_entries = k.n_entries;
    w->stats.n_pairs = w->stepped ? k.n_pairs : 0u;
    w->stats.n_contacts = k.n_contacts;
    w->stats.n_manifolds = k.n_manifolds;
    w->stats.n_colours = k.n_colours + k.n_colours_int;
    w->stats.colour_rounds = k.colour_rounds;
    w->stats.overflow |= k.overflow;
    w->stats.max_penetration = sfp::u2f(k.maxpen_bits);
    w->stats.launches = w->launches;
    std::vector<sfp::f4> pose(w->n), vel(w->n);
    std::vector<sfp::f2> minv(w->n);  // (the device arrays share one internal order)
    if (w->n) {
        CU(cudaMemcpy(pose.data(), w->ctx.pose, sizeof(sfp::f4) * w->n, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(vel.data(), w->ctx.vel, sizeof(sfp::f4) * w->n, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(minv.data(), w->ctx.minv, sizeof(sfp::f2) * w->n, cudaMemcpyDeviceToHost));
    }
    sfp::energies(pose, vel, minv, w->p.gravity_x, w->p.gravity_y, &w->stats.kinetic_energy,
                  &w->stats.potential_energy);
    *out = w->stats;
    return SFP_OK;
}

const char* sfp_last_error(SfpWorld* w) { return w ? w->err.c_str() : "null world"; }

int32_t sfp_last_step_ms(SfpWorld* w, float* ms) {
    if (!w || !ms) return SFP_E_INVALID;
    *ms = 0.0f;
    if (!w->timed) return SFP_OK;
    CU(cudaSetDevice(w->device));
    CU(cudaEventSynchronize(w->ev1));
    CU(cudaEventElapsedTime(ms, w->ev0, w->ev1));
    return SFP_OK;
}

int32_t sfp_last_step_breakdown_ms(SfpWorld* w, float out[4]) {
    if (!w || !out) return SFP_E_INVALID;
    out[0] = out[1] = out[2] = out[3] = 0.0f;
    if (!w->timed || w->n == 0) return SFP_OK;
    CU(cudaSetDevice(w->device));
    CU(cudaEventSynchronize(w->ev1));
    CU(cudaEventElapsedTime(&out[0], w->ev0, w->ev1));
    if (!(w->p.flags & SFP_FLAG_BROADPHASE_EVERY_SUBSTEP)) {
        CU(cudaEventElapsedTime(&out[1], w->ev0, w->ev_bp));
        CU(cudaEventElapsedTime(&out[2], w->ev_bp, w->ev_col));
        CU(cudaEventElapsedTime(&out[3], w->ev_col, w->ev1));
    }
    return SFP_OK;
}

int32_t sfp_last_step_stage_ns(SfpWorld* w, uint64_t out[6]) {
    if (!w || !out) return SFP_E_INVALID;
    for (int i = 0; i < 6; ++i) out[i] = 0;
    if (!w->timed || w->n == 0) return SFP_OK;
    CU(cudaSetDevice(w->device));
    Counters k;
    TRY(read_counters(w, &k));
    // Counters::stage_cycles: 4 integrate, 5 narrowphase, 6 position (+10 tile-interior),
    // 7 derive, 8 velocity (+11), 9 small colours
    out[0] = k.stage_cycles[4];
    out[1] = k.stage_cycles[5];
    out[2] = k.stage_cycles[6] + k.stage_cycles[10];
    out[3] = k.stage_cycles[7];
    out[4] = k.stage_cycles[8] + k.stage_cycles[11];
    out[5] = k.stage_cycles[9];
    return SFP_OK;
}

/* development aid (not part of include/sfp.h): ns per stage of the last substep kernel, as seen
 * by thread 0: [4] integrate [5] narrowphase [6] head position colours [7] derive
 * [8] head velocity colours [9] tail colours (both passes); [0] = index of the first tail colour */
__attribute__((visibility("default"))) int32_t sfp_debug_stage_ns(SfpWorld* w, unsigned long long out[12]) {
    if (!w || !out) return SFP_E_INVALID;
    Counters k;
    TRY(read_counters(w, &k));
    for (int i = 0; i < 12; ++i) out[i] = k.stage_cycles[i];
    return SFP_OK;
}

// development aid: the 512 scratch words the colouring rounds (and the probe build) write
__attribute__((visibility("default"))) int32_t sfp_debug_round_flags(SfpWorld* w, uint32_t out[512]) {
    if (!w || !out) return SFP_E_INVALID;
    Counters k;
    TRY(read_counters(w, &k));
    for (int i = 0; i < 512; ++i) out[i] = k.round_flags[i];
    return SFP_OK;
}

const char* sfp_backend(void) { return "cuda-sm_100a"; }

}  // extern "C"
