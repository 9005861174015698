/* sfp_ref.cpp — CPU ORACLE for the physics substep loop.  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (starframe_b200/csrc, libsfp_b200.so) never links,
 * loads or calls it.
 *
 * PARITY UNPINNED: the reference source tree is not mounted (/root/reference holds only
 * README.md:1-3, "moved to Codeberg"), so there is no reference file:line to
 * follow and no reference golden vector to pin against.  This file restates the step from
 *   - BASELINE.json north_star (stage list: spatial-hash broadphase, narrowphase for circle /
 *     capsule / rounded-polygon colliders, XPBD projection of contacts and distance/rope
 *     joints, velocity update with friction and restitution), and
 *   - the public XPBD rigid-body algorithm (Müller et al. 2020), Algorithm 2 + §3.5/3.6,
 * in plain sequential C++ (Gauss-Seidel; one thread unless SFP_REF_THREADS asks for more, in
 * which case the loops over bodies, pairs and the constraints of ONE colour are split over a
 * small pool — same bits, see tests/test_oracle_props.py — for bench.py's CPU arm).  It is
 * pinned to PHYSICS instead, by the analytic known-answer tests in tests/test_oracle_kat.py.
 *
 * The per-item arithmetic lives in include/sfp_math.h, compiled here with -ffp-contract=off.
 * The data-parallel structure (hashing, sorting, colouring rounds, solve order) is restated
 * here independently of the CUDA code: sequential loops over std::vector.
 */
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "sfp.h"
#include "sfp_host.h"
#include "sfp_math.h"

using namespace sfp;

/* A small persistent fork-join pool for the colour-order step (SFP_REF_THREADS > 1): the
 * caller is worker 0, the others sleep on a condition variable between loops.  Static
 * partition of [0, n) into one contiguous range per worker; no work stealing, no atomics on the
 * data — the loops it runs touch disjoint items by construction. */
class Pool {
   public:
    explicit Pool(int workers) : n_workers_(workers) {
        for (int t = 1; t < workers; ++t) threads_.emplace_back([this, t] { run(t); });
    }
    ~Pool() {
        stop_.store(true, std::memory_order_release);
        epoch_.fetch_add(1, std::memory_order_acq_rel);
        {
            std::unique_lock<std::mutex> lk(m_);
        }
        cv_.notify_all();
        for (auto& th : threads_) th.join();
    }
    int workers() const { return n_workers_; }
    template <typename F>
    void run_loop(size_t n, const F& f) {
        fn_ = [&](int t) {
            size_t lo = n * (size_t)t / (size_t)n_workers_, hi = n * (size_t)(t + 1) / (size_t)n_workers_;
            for (size_t i = lo; i < hi; ++i) f(i, t);
        };
        pending_.store(n_workers_ - 1, std::memory_order_release);
        epoch_.fetch_add(1, std::memory_order_acq_rel);
        if (sleepers_.load(std::memory_order_acquire) > 0) {
            {
                std::unique_lock<std::mutex> lk(m_);
            }
            cv_.notify_all();
        }
        fn_(0);
        while (pending_.load(std::memory_order_acquire) != 0) cpu_relax();
    }

   private:
    static void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    }
    /* the loops of a substep follow each other within microseconds: spin first, sleep only when
     * the owner has gone quiet (between frames, or when the world is idle) */
    void run(int t) {
        unsigned long seen = 0;
        for (;;) {
            int spins = 0;
            while (epoch_.load(std::memory_order_acquire) == seen) {
                if (++spins < 200000) {
                    cpu_relax();
                } else {
                    std::unique_lock<std::mutex> lk(m_);
                    sleepers_.fetch_add(1, std::memory_order_acq_rel);
                    cv_.wait(lk, [&] { return epoch_.load(std::memory_order_acquire) != seen; });
                    sleepers_.fetch_sub(1, std::memory_order_acq_rel);
                }
            }
            seen = epoch_.load(std::memory_order_acquire);
            if (stop_.load(std::memory_order_acquire)) return;
            fn_(t);
            pending_.fetch_sub(1, std::memory_order_acq_rel);
        }
    }
    int n_workers_;
    std::vector<std::thread> threads_;
    std::mutex m_;
    std::condition_variable cv_;
    std::function<void(int)> fn_;
    std::atomic<unsigned long> epoch_{0};
    std::atomic<int> pending_{0}, sleepers_{0};
    std::atomic<bool> stop_{false};
};

struct SfpWorld {
    SfpParams p;
    uint32_t n = 0;
    std::vector<f4> pose, prev, vel, shape, mat;
    std::vector<f2> minv;
    std::vector<SfpJointDesc> joints;
    PolyTable polys; /* general convex polygons (sfp_ref_set_polygons) */
    /* per frame */
    std::vector<SfpPair> pairs;          /* sorted by (a,b) */
    /* colour of a constraint as reported by get_colours: a boundary constraint's colour as it is
     * (0..62), an interior one's + SFP_CLASS_INTERIOR, the serial group's 63 (sfp_math.h, "Tiles") */
    std::vector<uint32_t> pair_colour;   /* aligned with pairs */
    std::vector<uint32_t> joint_colour;  /* aligned with joints */
    std::vector<uint32_t> tile;          /* per body: its tile, SFP_NO_TILE for statics / tiles off */
    std::vector<ContactRec> contacts;    /* aligned with pairs */
    std::vector<f4> lambdas;             /* aligned with pairs: ln0, lt0, ln1, lt1 */
    uint32_t n_colours = 0, colour_rounds = 0, n_entries = 0;
    /* colours of the previous frame, for the warm start */
    std::unordered_map<uint64_t, uint32_t> prev_pair_colour;
    std::vector<uint32_t> prev_joint_colour;
    uint32_t colour_age = 0; /* colourings since the last upload */
    /* constraints of each colour, in index order (rebuilt with every colouring) */
    std::vector<std::vector<uint32_t>> pairs_of, joints_of;
    /* CPU threads for the colour-order step (env SFP_REF_THREADS, default 1).  Constraints of one
     * colour touch disjoint dynamic bodies and every other loop is per body or per pair, so the
     * result does not depend on the thread count (tests/test_oracle_props.py checks). */
    int threads = 1;
    std::unique_ptr<Pool> pool;
    SfpStats stats;
    std::string err;
    float last_ms = 0.0f;
    float last_h = 0.0f; /* substep length of the last frame */
    bool stepped = false;
};

namespace {

struct Entry {
    uint32_t key, body;
};

bool is_dynamic(const SfpWorld& w, uint32_t i) { return w.minv[i].x > 0.0f || w.minv[i].y > 0.0f; }

bool pair_allowed(const SfpWorld& w, uint32_t a, uint32_t b) {
    if (!is_dynamic(w, a) && !is_dynamic(w, b)) return false;
    return layers_collide(f2u(w.mat[a].w), f2u(w.mat[b].w));
}

/* f(i, worker) for i in [0, n); with T == 1 (or a short loop) a plain loop in the caller */
template <typename F>
void parallel_for(SfpWorld& w, int T, size_t n, const F& f) {
    if (T <= 1 || n < 256) {
        for (size_t i = 0; i < n; ++i) f(i, 0);
        return;
    }
    if (!w.pool || w.pool->workers() != T) w.pool.reset(new Pool(T));
    w.pool->run_loop(n, f);
}

void compute_aabbs(const SfpWorld& w, float dt, std::vector<f4>& aabb) {
    aabb.resize(w.n);
    for (uint32_t i = 0; i < w.n; ++i)
        aabb[i] = body_aabb(w.pose[i], w.vel[i], w.shape[i], w.p.aabb_margin, w.p.vel_expand, dt,
                            w.polys.recs.empty() ? nullptr : w.polys.recs.data());
}

/* A2: every body is entered into each cell its expanded AABB touches; inside a cell every two
 * entries are tested; a pair is reported only from the cell that holds the lower corner of
 * the two cell ranges' intersection, which makes the report unique without deduplication. */
int broadphase(SfpWorld& w, float dt) {
    std::vector<f4> aabb;
    compute_aabbs(w, dt, aabb);
    float inv = 1.0f / w.p.cell_size;
    std::vector<Entry> entries;
    std::vector<int> cx0(w.n), cy0(w.n);
    bool runaway = false;
    for (uint32_t i = 0; i < w.n; ++i) {
        int x0, y0, x1, y1;
        if (!cell_range(aabb[i], inv, &x0, &y0, &x1, &y1)) runaway = true;
        cx0[i] = x0;
        cy0[i] = y0;
        for (int cy = y0; cy <= y1; ++cy)
            for (int cx = x0; cx <= x1; ++cx) entries.push_back({cell_key(cx, cy), i});
    }
    w.n_entries = (uint32_t)entries.size();
    if (runaway) {
        w.stats.overflow |= 16u;
        w.err = "a body is not finite or spans more than 16384 cells";
        return SFP_E_CAPACITY;
    }
    if (w.p.max_entries && entries.size() > w.p.max_entries) {
        w.stats.overflow |= 1u;
        w.err = "broadphase entry capacity exceeded";
        return SFP_E_CAPACITY;
    }
    std::stable_sort(entries.begin(), entries.end(),
                     [](const Entry& a, const Entry& b) { return a.key < b.key; });
    w.pairs.clear();
    size_t i = 0;
    while (i < entries.size()) {
        size_t j = i;
        while (j < entries.size() && entries[j].key == entries[i].key) ++j;
        int cx = key_cx(entries[i].key), cy = key_cy(entries[i].key);
        for (size_t u = i; u < j; ++u)
            for (size_t v = u + 1; v < j; ++v) {
                uint32_t a = entries[u].body, b = entries[v].body;
                if (a == b) continue;
                if (!same_cell(std::max(cx0[a], cx0[b]), std::max(cy0[a], cy0[b]), cx, cy)) continue;
                if (!aabb_overlap(aabb[a], aabb[b])) continue;
                if (!pair_allowed(w, a, b)) continue;
                SfpPair p = {std::min(a, b), std::max(a, b)};
                w.pairs.push_back(p);
            }
        i = j;
    }
    std::sort(w.pairs.begin(), w.pairs.end(), [](const SfpPair& x, const SfpPair& y) {
        return x.a != y.a ? x.a < y.a : x.b < y.b;
    });
    if (w.p.max_pairs && w.pairs.size() > w.p.max_pairs) {
        w.stats.overflow |= 2u;
        w.err = "pair capacity exceeded";
        return SFP_E_CAPACITY;
    }
    return SFP_OK;
}

/* tile of every dynamic body (sfp_math.h, "Tiles"): runs of tile_size() bodies in the order of
 * their Morton cell key, ties by body index.  From the poses at the moment of the call. */
void compute_tiles(SfpWorld& w) {
    w.tile.assign(w.n, SFP_NO_TILE);
    if (!(w.p.flags & SFP_FLAG_TILES)) return;
    const float inv = 1.0f / w.p.cell_size;
    std::vector<uint32_t> key(w.n), order(w.n);
    uint32_t ndyn = 0;
    for (uint32_t i = 0; i < w.n; ++i) {
        const bool dyn = is_dynamic(w, i);
        key[i] = tile_sort_key(w.pose[i].x, w.pose[i].y, inv, dyn);
        order[i] = i;
        ndyn += dyn ? 1u : 0u;
    }
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return key[a] < key[b]; });
    const uint32_t B = tile_size(ndyn);
    for (uint32_t r = 0; r < ndyn; ++r) w.tile[order[r]] = r / B;
}

/* A8: rounds of "every body grants its lowest-priority-key uncoloured constraint; a
 * constraint granted by all its dynamic bodies takes the lowest colour free on them" — free in
 * the colour namespace of the constraint's class (interior to a tile / boundary). */
int colour_constraints(SfpWorld& w) {
    const uint32_t P = (uint32_t)w.pairs.size(), J = (uint32_t)w.joints.size();
    const uint32_t UNC = 0xffffffffu;
    /* warm start: a constraint that existed in the previous frame keeps its colour (those
     * colours were proper among themselves and still are); the tiles stay as they are */
    const bool warm = (w.colour_age % SFP_RECOLOUR_PERIOD) != 0u;
    w.colour_age += 1;
    if (!warm || w.tile.size() != w.n) compute_tiles(w);
    w.pair_colour.assign(P, UNC);
    w.joint_colour.assign(J, UNC);
    /* mask[0]: boundary colours in use per body, mask[1]: interior colours */
    std::vector<uint64_t> mask[2] = {std::vector<uint64_t>(w.n, 0ull), std::vector<uint64_t>(w.n, 0ull)};
    std::vector<uint64_t> claim(w.n);
    std::vector<uint64_t> pkey(P), jkey(J);
    for (uint32_t k = 0; k < P; ++k) pkey[k] = pair_priority(w.pairs[k].a, w.pairs[k].b);
    for (uint32_t k = 0; k < J; ++k) jkey[k] = joint_priority(k);
    uint32_t remaining = P + J, rounds = 0, ncol[2] = {0, 0}, serial = 0;
    bool overflow = false;
    auto interior = [&](uint32_t a, uint32_t b) {
        return constraint_tile(w.tile[a], w.tile[b]) != SFP_NO_TILE;
    };
    /* constraints without a dynamic end never conflict: colour 0 */
    for (uint32_t k = 0; k < J; ++k)
        if (!is_dynamic(w, w.joints[k].body_a) && !is_dynamic(w, w.joints[k].body_b)) {
            w.joint_colour[k] = 0;
            --remaining;
            ncol[0] = std::max(ncol[0], 1u);
        }
    if (warm) {
        auto keep = [&](uint32_t a, uint32_t b, uint32_t enc, uint32_t* colour) {
            const uint32_t cls = enc >= SFP_CLASS_INTERIOR ? 1u : 0u, c = enc & 63u;
            if (c != SFP_SERIAL_COLOUR) {
                if (is_dynamic(w, a)) mask[cls][a] |= 1ull << c;
                if (is_dynamic(w, b)) mask[cls][b] |= 1ull << c;
            } else if (++serial > SFP_SERIAL_MAX) {
                overflow = true;
            }
            *colour = enc;
            ncol[cls] = std::max(ncol[cls], c + 1u);
            --remaining;
        };
        for (uint32_t k = 0; k < P; ++k) {
            auto it = w.prev_pair_colour.find(ctab_key(w.pairs[k].a, w.pairs[k].b));
            if (it != w.prev_pair_colour.end())
                keep(w.pairs[k].a, w.pairs[k].b, it->second, &w.pair_colour[k]);
        }
        if (w.prev_joint_colour.size() == J)
            for (uint32_t k = 0; k < J; ++k)
                if (w.joint_colour[k] == UNC && w.prev_joint_colour[k] != UNC)
                    keep(w.joints[k].body_a, w.joints[k].body_b, w.prev_joint_colour[k],
                         &w.joint_colour[k]);
    }
    /* at least one (possibly empty) round, like the device loop */
    bool first = w.n > 0;
    while (remaining > 0 || first) {
        first = false;
        ++rounds;
        std::fill(claim.begin(), claim.end(), SFP_NO_CLAIM);
        auto bid = [&](uint32_t body, uint64_t key) {
            if (is_dynamic(w, body)) claim[body] = std::min(claim[body], key);
        };
        for (uint32_t k = 0; k < P; ++k)
            if (w.pair_colour[k] == UNC) {
                bid(w.pairs[k].a, pkey[k]);
                bid(w.pairs[k].b, pkey[k]);
            }
        for (uint32_t k = 0; k < J; ++k)
            if (w.joint_colour[k] == UNC) {
                bid(w.joints[k].body_a, jkey[k]);
                bid(w.joints[k].body_b, jkey[k]);
            }
        auto try_take = [&](uint32_t a, uint32_t b, uint64_t key, uint32_t* colour) {
            bool da = is_dynamic(w, a), db = is_dynamic(w, b);
            if (da && claim[a] != key) return;
            if (db && claim[b] != key) return;
            const uint32_t cls = interior(a, b) ? 1u : 0u;
            uint64_t m = (da ? mask[cls][a] : 0ull) | (db ? mask[cls][b] : 0ull);
            int c = pick_colour(m);
            uint32_t enc = (uint32_t)c + (cls ? SFP_CLASS_INTERIOR : 0u);
            if (c != SFP_SERIAL_COLOUR) {
                if (da) mask[cls][a] |= 1ull << c;
                if (db) mask[cls][b] |= 1ull << c;
                ncol[cls] = std::max(ncol[cls], (uint32_t)c + 1u);
            } else {
                /* all 63 colours of the class taken: the serial group (one for both classes) */
                enc = SFP_SERIAL_COLOUR;
                ncol[0] = std::max(ncol[0], (uint32_t)SFP_SERIAL_COLOUR + 1u);
                if (++serial > SFP_SERIAL_MAX) overflow = true;
            }
            *colour = enc;
            --remaining;
        };
        for (uint32_t k = 0; k < P; ++k)
            if (w.pair_colour[k] == UNC)
                try_take(w.pairs[k].a, w.pairs[k].b, pkey[k], &w.pair_colour[k]);
        for (uint32_t k = 0; k < J; ++k)
            if (w.joint_colour[k] == UNC)
                try_take(w.joints[k].body_a, w.joints[k].body_b, jkey[k], &w.joint_colour[k]);
    }
    w.n_colours = ncol[0] + ncol[1];
    w.colour_rounds = rounds;
    w.prev_pair_colour.clear();
    w.prev_pair_colour.reserve(P * 2);
    for (uint32_t k = 0; k < P; ++k)
        w.prev_pair_colour[ctab_key(w.pairs[k].a, w.pairs[k].b)] = w.pair_colour[k];
    w.prev_joint_colour = w.joint_colour;
    w.pairs_of.assign(2 * SFP_MAX_COLOURS, {});
    w.joints_of.assign(2 * SFP_MAX_COLOURS, {});
    for (uint32_t k = 0; k < P; ++k)
        if (w.pair_colour[k] < 2 * SFP_MAX_COLOURS) w.pairs_of[w.pair_colour[k]].push_back(k);
    for (uint32_t k = 0; k < (uint32_t)w.joints.size(); ++k)
        if (w.joint_colour[k] < 2 * SFP_MAX_COLOURS) w.joints_of[w.joint_colour[k]].push_back(k);
    if (overflow) {
        w.stats.overflow |= 4u;
        w.err = "more than SFP_SERIAL_MAX constraints left without a colour";
        return SFP_E_CAPACITY;
    }
    return SFP_OK;
}

BodyPos load_pos(const SfpWorld& w, uint32_t i) {
    BodyPos b;
    b.x = mk(w.pose[i].x, w.pose[i].y);
    b.q = mk(w.pose[i].z, w.pose[i].w);
    b.d = mk(w.prev[i].x, w.prev[i].y);
    b.qp = mk(w.prev[i].z, w.prev[i].w);
    b.im = w.minv[i].x;
    b.ii = w.minv[i].y;
    return b;
}
void store_pos(SfpWorld& w, uint32_t i, const BodyPos& b) {
    if (!is_dynamic(w, i)) return;
    w.pose[i].x = b.x.x;
    w.pose[i].y = b.x.y;
    w.pose[i].z = b.q.x;
    w.pose[i].w = b.q.y;
    w.prev[i].x = b.d.x;
    w.prev[i].y = b.d.y;
}

void solve_pair_position(SfpWorld& w, uint32_t k) {
    const ContactRec& c = w.contacts[k];
    uint32_t cnt = f2u(c.b.y);
    f4 lam = {0.0f, 0.0f, 0.0f, 0.0f};
    if (cnt > 0) {
        uint32_t a = w.pairs[k].a, b = w.pairs[k].b;
        BodyPos A = load_pos(w, a), B = load_pos(w, b);
        v2 n = mk(c.a.x, c.a.y);
        lam = solve_contact_manifold(&A, &B, n, cnt, mk(c.p0.x, c.p0.y), mk(c.p0.z, c.p0.w),
                                     mk(c.p1.x, c.p1.y), mk(c.p1.z, c.p1.w), c.a.z);
        store_pos(w, a, A);
        store_pos(w, b, B);
    }
    w.lambdas[k] = lam;
}

void solve_joint_position(SfpWorld& w, uint32_t k, float inv_h2) {
    const SfpJointDesc& j = w.joints[k];
    BodyPos A = load_pos(w, j.body_a), B = load_pos(w, j.body_b);
    solve_joint(&A, &B, mk(j.ax, j.ay), mk(j.bx, j.by), j.rest, j.compliance, j.kind, inv_h2);
    store_pos(w, j.body_a, A);
    store_pos(w, j.body_b, B);
}

void solve_pair_velocity(SfpWorld& w, uint32_t k, float h, float rest_threshold) {
    const ContactRec& c = w.contacts[k];
    uint32_t cnt = f2u(c.b.y);
    f4 lam = w.lambdas[k];
    if (cnt == 0 || !(lam.x > 0.0f || lam.z > 0.0f)) return;
    uint32_t a = w.pairs[k].a, b = w.pairs[k].b;
    BodyVel A, B;
    A.v = mk(w.vel[a].x, w.vel[a].y);
    A.w = w.vel[a].z;
    A.im = w.minv[a].x;
    A.ii = w.minv[a].y;
    B.v = mk(w.vel[b].x, w.vel[b].y);
    B.w = w.vel[b].z;
    B.im = w.minv[b].x;
    B.ii = w.minv[b].y;
    v2 qa = mk(w.pose[a].z, w.pose[a].w), qb = mk(w.pose[b].z, w.pose[b].w);
    v2 n = mk(c.a.x, c.a.y);
    solve_manifold_velocity(&A, &B, qa, qb, n, cnt, mk(c.p0.x, c.p0.y), mk(c.p0.z, c.p0.w),
                            mk(c.p1.x, c.p1.y), mk(c.p1.z, c.p1.w), lam, c.b.z, c.b.w, c.a.z, c.a.w,
                            c.b.x, h, rest_threshold);
    if (is_dynamic(w, a)) {
        w.vel[a].x = A.v.x;
        w.vel[a].y = A.v.y;
        w.vel[a].z = A.w;
    }
    if (is_dynamic(w, b)) {
        w.vel[b].x = B.v.x;
        w.vel[b].y = B.v.y;
        w.vel[b].z = B.w;
    }
}

void substep(SfpWorld& w, float h) {
    const uint32_t P = (uint32_t)w.pairs.size(), J = (uint32_t)w.joints.size();
    const float inv_h = 1.0f / h, inv_h2 = inv_h * inv_h;
    const float g = sqrtf(w.p.gravity_x * w.p.gravity_x + w.p.gravity_y * w.p.gravity_y);
    const float rest_threshold = 2.0f * g * h;
    const int T = w.threads; /* 1: plain sequential loops in the caller */
    /* A3 */
    parallel_for(w, T, w.n, [&](size_t i, int) {
        integrate_body(&w.pose[i], &w.prev[i], &w.vel[i], w.minv[i], w.p.gravity_x, w.p.gravity_y, h);
    });
    /* A4 */
    w.contacts.resize(P);
    w.lambdas.assign(P, f4{0.0f, 0.0f, 0.0f, 0.0f});
    std::vector<uint32_t> t_npts((size_t)T, 0u), t_nman((size_t)T, 0u);
    std::vector<float> t_maxpen((size_t)T, 0.0f);
    parallel_for(w, T, P, [&](size_t k, int t) {
        uint32_t a = w.pairs[k].a, b = w.pairs[k].b;
        w.contacts[k] = narrowphase<true>(w.pose[a], w.vel[a], w.shape[a], w.mat[a], w.pose[b], w.vel[b],
                                    w.shape[b], w.mat[b], w.p.contact_margin, w.polys.recs.data());
        uint32_t cnt = f2u(w.contacts[k].b.y);
        t_npts[t] += cnt;
        t_nman[t] += cnt > 0 ? 1u : 0u;
        if (cnt > 0) t_maxpen[t] = std::max(t_maxpen[t], -w.contacts[k].s.x);
        if (cnt > 1) t_maxpen[t] = std::max(t_maxpen[t], -w.contacts[k].s.y);
    });
    uint32_t npts = 0, nman = 0;
    float maxpen = 0.0f;
    for (int t = 0; t < T; ++t) {
        npts += t_npts[t];
        nman += t_nman[t];
        maxpen = std::max(maxpen, t_maxpen[t]);
    }
    w.stats.n_contacts = npts;
    w.stats.n_manifolds = nman;
    w.stats.max_penetration = maxpen;
    /* solve order (sfp_math.h, "Tiles"): interior colours ascending; boundary colours with more
     * than SFP_TAIL_ITEMS constraints ascending, then the smaller ones; the serial group last */
    std::vector<uint32_t> order;
    if (w.p.solver_order != SFP_ORDER_REFERENCE) {
        for (uint32_t c = 0; c < SFP_SERIAL_COLOUR; ++c)
            if (!w.pairs_of[SFP_CLASS_INTERIOR + c].empty() || !w.joints_of[SFP_CLASS_INTERIOR + c].empty())
                order.push_back(SFP_CLASS_INTERIOR + c);
        for (int pass = 0; pass < 2; ++pass)
            for (uint32_t c = 0; c < SFP_SERIAL_COLOUR; ++c) {
                const size_t items = w.pairs_of[c].size() + w.joints_of[c].size();
                if (items > 0 && (items > SFP_TAIL_ITEMS) == (pass == 0)) order.push_back(c);
            }
        if (!w.pairs_of[SFP_SERIAL_COLOUR].empty() || !w.joints_of[SFP_SERIAL_COLOUR].empty())
            order.push_back(SFP_SERIAL_COLOUR);
    }
    /* A5 + A6 */
    if (w.p.solver_order == SFP_ORDER_REFERENCE) {
        for (uint32_t k = 0; k < J; ++k) solve_joint_position(w, k, inv_h2);
        for (uint32_t k = 0; k < P; ++k) solve_pair_position(w, k);
    } else {
        for (uint32_t c : order) {
            const std::vector<uint32_t>&jl = w.joints_of[c], &pl = w.pairs_of[c];
            /* the serial group's members may share bodies: always in sequence */
            const int Tc = (c == SFP_SERIAL_COLOUR) ? 1 : T;
            parallel_for(w, Tc, jl.size(), [&](size_t i, int) { solve_joint_position(w, jl[i], inv_h2); });
            parallel_for(w, Tc, pl.size(), [&](size_t i, int) { solve_pair_position(w, pl[i]); });
        }
    }
    /* A7 (i) */
    parallel_for(w, T, w.n, [&](size_t i, int) {
        w.vel[i] = derive_velocity(w.pose[i], w.prev[i], w.vel[i], w.minv[i], inv_h);
    });
    /* A7 (ii) */
    if (w.p.solver_order == SFP_ORDER_REFERENCE) {
        for (uint32_t k = 0; k < P; ++k) solve_pair_velocity(w, k, h, rest_threshold);
    } else {
        for (uint32_t c : order) {
            const std::vector<uint32_t>& pl = w.pairs_of[c];
            const int Tc = (c == SFP_SERIAL_COLOUR) ? 1 : T;
            parallel_for(w, Tc, pl.size(), [&](size_t i, int) { solve_pair_velocity(w, pl[i], h, rest_threshold); });
        }
    }
}

int frame(SfpWorld& w, float dt, uint32_t substeps) {
    const float h = dt / (float)substeps;
    const bool every = (w.p.flags & SFP_FLAG_BROADPHASE_EVERY_SUBSTEP) != 0u;
    int rc;
    if (!every) {
        if ((rc = broadphase(w, dt)) != SFP_OK) return rc;
        if ((rc = colour_constraints(w)) != SFP_OK) return rc;
    }
    for (uint32_t s = 0; s < substeps; ++s) {
        if (every) {
            if ((rc = broadphase(w, h)) != SFP_OK) return rc;
            if ((rc = colour_constraints(w)) != SFP_OK) return rc;
        }
        substep(w, h);
    }
    w.stats.frames += 1;
    w.stats.body_substeps += (uint64_t)w.n * substeps;
    w.stepped = true;
    w.last_h = h;
    return SFP_OK;
}

} /* namespace */

extern "C" {

int32_t sfp_ref_default_params(SfpParams* out) {
    if (!out) return SFP_E_INVALID;
    default_params(out);
    return SFP_OK;
}

int32_t sfp_ref_create(const SfpParams* params, SfpWorld** out) {
    if (!out) return SFP_E_INVALID;
    *out = nullptr;
    SfpParams p;
    if (params) p = *params;
    else default_params(&p);
    if (!params_valid(p)) return SFP_E_INVALID;
    SfpWorld* w = new (std::nothrow) SfpWorld();
    if (!w) return SFP_E_NOMEM;
    w->p = p;
    std::memset(&w->stats, 0, sizeof(w->stats));
    if (const char* e = std::getenv("SFP_REF_THREADS")) {
        int t = std::atoi(e);
        w->threads = t < 1 ? 1 : (t > 256 ? 256 : t);
    }
    *out = w;
    return SFP_OK;
}

int32_t sfp_ref_destroy(SfpWorld* w) {
    delete w;
    return SFP_OK;
}

int32_t sfp_ref_upload_bodies(SfpWorld* w, const SfpBodyDesc* bodies, uint32_t n) {
    if (!w || (n && !bodies)) return SFP_E_INVALID;
    if (n > SFP_MAX_BODIES || (w->p.max_bodies && n > w->p.max_bodies)) {
        w->err = "too many bodies";
        return SFP_E_CAPACITY;
    }
    for (uint32_t i = 0; i < n; ++i)
        if (!desc_valid(bodies[i], &w->polys)) {
            w->err = "invalid body desc at index " + std::to_string(i);
            return SFP_E_INVALID;
        }
    for (const SfpJointDesc& j : w->joints)
        if (j.body_a >= n || j.body_b >= n) {
            w->err = "existing joints reference bodies beyond the new count";
            return SFP_E_INVALID;
        }
    w->n = n;
    w->pose.resize(n);
    w->prev.resize(n);
    w->vel.resize(n);
    w->shape.resize(n);
    w->mat.resize(n);
    w->minv.resize(n);
    uint32_t ndyn = 0;
    for (uint32_t i = 0; i < n; ++i) {
        body_from_desc(bodies[i], &w->pose[i], &w->vel[i], &w->minv[i], &w->shape[i], &w->mat[i],
                       &w->polys);
        w->prev[i] = w->pose[i];
        ndyn += is_dynamic(*w, i) ? 1u : 0u;
        if (!body_fits_grid(w->pose[i], w->shape[i], w->p)) {
            w->err = "body " + std::to_string(i) + " spans more than 16384 broadphase cells (or 4096 along one axis) at this cell_size: split it or raise cell_size";
            w->n = 0;
            return SFP_E_INVALID;
        }
    }
    w->stats.n_bodies = n;
    w->stats.n_dynamic = ndyn;
    w->pairs.clear();
    w->contacts.clear();
    w->lambdas.clear();
    w->pair_colour.clear();
    w->prev_pair_colour.clear();
    w->prev_joint_colour.clear();
    w->tile.clear();
    w->colour_age = 0;
    w->stepped = false;
    return SFP_OK;
}

int32_t sfp_ref_set_polygons(SfpWorld* w, const SfpPolygon* polys, uint32_t n) {
    if (!w || (n && !polys)) return SFP_E_INVALID;
    PolyTable t;
    long bad = build_poly_table(polys, n, &t);
    if (bad >= 0) {
        w->err = "invalid polygon at index " + std::to_string(bad) +
                 " (3..8 vertices, counter-clockwise, strictly convex, centroid at the origin)";
        return SFP_E_INVALID;
    }
    w->polys = std::move(t);
    return SFP_OK;
}

int32_t sfp_ref_set_joints(SfpWorld* w, const SfpJointDesc* joints, uint32_t n) {
    if (!w || (n && !joints)) return SFP_E_INVALID;
    if (w->p.max_joints && n > w->p.max_joints) return SFP_E_CAPACITY;
    for (uint32_t i = 0; i < n; ++i)
        if (!joint_valid(joints[i], w->n)) {
            w->err = "invalid joint desc at index " + std::to_string(i);
            return SFP_E_INVALID;
        }
    w->joints.assign(joints, joints + n);
    w->joint_colour.clear();
    w->prev_pair_colour.clear();
    w->prev_joint_colour.clear();
    w->colour_age = 0;
    w->stats.n_joints = n;
    return SFP_OK;
}

int32_t sfp_ref_step(SfpWorld* w, float dt, uint32_t substeps) {
    if (!w || !(dt > 0.0f) || substeps == 0 || substeps > 1024) return SFP_E_INVALID;
    auto t0 = std::chrono::steady_clock::now();
    int rc = frame(*w, dt, substeps);
    auto t1 = std::chrono::steady_clock::now();
    w->last_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    return rc;
}

int32_t sfp_ref_sync(SfpWorld* w) { return w ? SFP_OK : SFP_E_INVALID; }

int32_t sfp_ref_download_bodies(SfpWorld* w, SfpBodyState* out, uint32_t n) {
    if (!w || (n && !out) || n > w->n) return SFP_E_INVALID;
    for (uint32_t i = 0; i < n; ++i) out[i] = state_from_records(w->pose[i], w->vel[i]);
    return SFP_OK;
}

int32_t sfp_ref_get_pairs(SfpWorld* w, SfpPair* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    *n_out = (uint32_t)w->pairs.size();
    uint32_t m = std::min<uint32_t>(cap, *n_out);
    if (m) std::memcpy(out, w->pairs.data(), sizeof(SfpPair) * m);
    return SFP_OK;
}

int32_t sfp_ref_get_contacts(SfpWorld* w, SfpContact* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    uint32_t cnt = 0;
    for (size_t k = 0; k < w->contacts.size() && k < w->pairs.size(); ++k) {
        if (f2u(w->contacts[k].b.y) == 0) continue;
        if (cnt < cap) {
            uint32_t a = w->pairs[k].a, b = w->pairs[k].b;
            out[cnt] = contact_from_records(a, b, w->contacts[k], w->lambdas[k], w->pose[a],
                                            w->pose[b]);
        }
        ++cnt;
    }
    *n_out = cnt;
    return SFP_OK;
}

int32_t sfp_ref_get_colours(SfpWorld* w, uint32_t* pair_colours, uint32_t cap,
                            uint32_t* joint_colours, uint32_t jcap) {
    if (!w) return SFP_E_INVALID;
    if (pair_colours)
        for (uint32_t k = 0; k < cap && k < w->pair_colour.size(); ++k)
            pair_colours[k] = w->pair_colour[k];
    if (joint_colours)
        for (uint32_t k = 0; k < jcap && k < w->joint_colour.size(); ++k)
            joint_colours[k] = w->joint_colour[k];
    return SFP_OK;
}

int32_t sfp_ref_get_stats(SfpWorld* w, SfpStats* out) {
    if (!w || !out) return SFP_E_INVALID;
    w->stats.n_entries = w->n_entries;
    w->stats.n_pairs = (uint32_t)w->pairs.size();
    w->stats.n_colours = w->n_colours;
    w->stats.colour_rounds = w->colour_rounds;
    energies(w->pose, w->vel, w->minv, w->p.gravity_x, w->p.gravity_y, &w->stats.kinetic_energy,
             &w->stats.potential_energy);
    *out = w->stats;
    return SFP_OK;
}

int32_t sfp_ref_get_contact_events(SfpWorld* w, SfpContactEvent* out, uint32_t cap,
                                   uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    static_assert(sizeof(EventRec) == sizeof(SfpContactEvent), "EventRec mirrors SfpContactEvent");
    uint32_t cnt = 0;
    const float inv_h = w->last_h > 0.0f ? 1.0f / w->last_h : 0.0f;
    for (size_t k = 0; k < w->contacts.size() && k < w->pairs.size(); ++k) {
        const ContactRec& c = w->contacts[k];
        uint32_t a = w->pairs[k].a, b = w->pairs[k].b;
        EventRec e;
        if (!pair_event(a, b, c.a, c.b, c.p0, c.p1, w->lambdas[k], w->pose[a], w->pose[b], inv_h, &e))
            continue;
        if (cnt < cap) std::memcpy(&out[cnt], &e, sizeof(e));
        ++cnt;
    }
    *n_out = cnt;
    return SFP_OK;
}

/* brute force over every body: the device's hash walk must find the same closest hit */
int32_t sfp_ref_ray_cast(SfpWorld* w, const SfpRay* rays, uint32_t m, SfpRayHit* hits) {
    if (!w || (m && (!rays || !hits))) return SFP_E_INVALID;
    static_assert(sizeof(RayRec) == sizeof(SfpRay) && sizeof(RayHitRec) == sizeof(SfpRayHit), "");
    for (uint32_t k = 0; k < m; ++k) {
        const SfpRay& r = rays[k];
        SfpRayHit best = {0xffffffffu, r.max_t, 0.0f, 0.0f};
        v2 o = mk(r.ox, r.oy), d = mk(r.dx, r.dy);
        for (uint32_t i = 0; i < w->n; ++i) {
            if (((f2u(w->mat[i].w) & 0xffffu) & r.mask) == 0u) continue;
            float t;
            v2 n;
            if (!ray_collider(o, d, r.max_t, w->pose[i], w->shape[i], w->polys.recs.data(), &t, &n))
                continue;
            if (best.body == 0xffffffffu || t < best.t) best = SfpRayHit{i, t, n.x, n.y};
        }
        if (best.body == 0xffffffffu) best.t = 0.0f;
        hits[k] = best;
    }
    return SFP_OK;
}

int32_t sfp_ref_query_aabb(SfpWorld* w, float lo_x, float lo_y, float hi_x, float hi_y,
                           uint32_t mask, uint32_t* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    f4 box = {lo_x, lo_y, hi_x, hi_y};
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < w->n; ++i) {
        if (((f2u(w->mat[i].w) & 0xffffu) & mask) == 0u) continue;
        if (!aabb_overlap(tight_aabb(w->pose[i], w->shape[i],
                                     w->polys.recs.empty() ? nullptr : w->polys.recs.data()), box))
            continue;
        if (cnt < cap) out[cnt] = i;
        ++cnt;
    }
    *n_out = cnt;
    return SFP_OK;
}

int32_t sfp_ref_query_shape(SfpWorld* w, const SfpBodyDesc* probe, uint32_t mask, uint32_t* out,
                            uint32_t cap, uint32_t* n_out) {
    if (!w || !probe || !n_out || (cap && !out)) return SFP_E_INVALID;
    f4 pp, ps;
    if (!probe_from_desc(*probe, &w->polys, &pp, &ps)) return SFP_E_INVALID;
    const f4* table = w->polys.recs.empty() ? nullptr : w->polys.recs.data();
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < w->n; ++i) {
        if (((f2u(w->mat[i].w) & 0xffffu) & mask) == 0u) continue;
        if (!colliders_overlap(w->pose[i], w->shape[i], pp, ps, table)) continue;
        if (cnt < cap) out[cnt] = i;
        ++cnt;
    }
    *n_out = cnt;
    return SFP_OK;
}

int32_t sfp_ref_update_bodies(SfpWorld* w, const uint32_t* idx, const SfpBodyState* states,
                              uint32_t m) {
    if (!w || (m && (!idx || !states))) return SFP_E_INVALID;
    for (uint32_t k = 0; k < m; ++k)
        if (idx[k] >= w->n) return SFP_E_INVALID;
    for (uint32_t k = 0; k < m; ++k)
        records_from_state(states[k], w->minv[idx[k]], &w->pose[idx[k]], &w->vel[idx[k]]);
    return SFP_OK;
}

const char* sfp_ref_last_error(SfpWorld* w) { return w ? w->err.c_str() : "null world"; }

int32_t sfp_ref_step_host(SfpWorld* w, const SfpBodyState* in, SfpBodyState* out, uint32_t n,
                          float dt, uint32_t substeps) {
    if (!w || !in || !out || n != w->n) return SFP_E_INVALID;
    for (uint32_t i = 0; i < n; ++i) {
        records_from_state(in[i], w->minv[i], &w->pose[i], &w->vel[i]);
        w->prev[i] = w->pose[i];
    }
    int rc = sfp_ref_step(w, dt, substeps);
    if (rc != SFP_OK) return rc;
    return sfp_ref_download_bodies(w, out, n);
}

int32_t sfp_ref_last_step_ms(SfpWorld* w, float* ms) {
    if (!w || !ms) return SFP_E_INVALID;
    *ms = w->last_ms;
    return SFP_OK;
}

const char* sfp_ref_backend(void) { return "cpu-oracle"; }

int32_t sfp_ref_brute_pairs(SfpWorld* w, float dt, SfpPair* out, uint32_t cap, uint32_t* n_out) {
    if (!w || !n_out || (cap && !out)) return SFP_E_INVALID;
    std::vector<f4> aabb;
    compute_aabbs(*w, dt, aabb);
    uint32_t cnt = 0;
    for (uint32_t a = 0; a < w->n; ++a)
        for (uint32_t b = a + 1; b < w->n; ++b) {
            if (!aabb_overlap(aabb[a], aabb[b])) continue;
            if (!pair_allowed(*w, a, b)) continue;
            if (cnt < cap) {
                out[cnt].a = a;
                out[cnt].b = b;
            }
            ++cnt;
        }
    *n_out = cnt;
    return SFP_OK;
}

} /* extern "C" */
