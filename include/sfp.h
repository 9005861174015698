/* sfp.h — C ABI of the starframe physics step (B200-native build).
 *
 * This is the drop-in boundary for ONE path of m0lentum/starframe: the per-frame physics
 * substep loop (spatial-hash broadphase -> narrowphase -> XPBD position projection ->
 * velocity update with friction and restitution; BASELINE.json north_star).
 *
 * Reference citations: the reference source tree is NOT mounted offline
 * (/root/reference holds only README.md:1-3, "moved to Codeberg").  Every entry point below
 * therefore cites the interface it replaces by north_star's own wording ("the reference's
 * Physics/Collider/Body API surface") instead of a file:line; see BASELINE.md §0 and
 * SURVEY.md §8b.  Parity with the reference itself is UNPINNED.
 *
 * Two shared libraries export this API:
 *   libsfp_b200.so  (starframe_b200/csrc)  symbols  sfp_*      — CUDA sm_100a product path
 *   libsfp_ref.so   (oracle/)              symbols  sfp_ref_*  — CPU oracle, TEST INFRASTRUCTURE ONLY
 * The oracle's entry points have identical signatures with the prefix sfp_ref_.
 *
 * Conventions
 *   - every function returns int32_t: 0 = SFP_OK, negative = error (never throws / unwinds)
 *   - the library owns everything behind SfpWorld*; the caller owns every buffer it passes;
 *     no caller pointer is retained after a call returns
 *   - a world is not thread-safe (mirrors `&mut self` on the Rust side); distinct worlds may
 *     be used from distinct threads
 *   - body index == upload order, stable for the life of the world
 *   - all structs are plain C PODs (#[repr(C)]-compatible), little-endian
 */
#ifndef SFP_H_
#define SFP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define SFP_API
#else
#define SFP_API __attribute__((visibility("default")))
#endif

/* ---- error codes ------------------------------------------------------------------ */
enum {
    SFP_OK = 0,
    SFP_E_INVALID = -1,  /* bad argument / bad state */
    SFP_E_CUDA = -2,     /* a CUDA call failed; sfp_last_error() has the string */
    SFP_E_CAPACITY = -3, /* a device buffer (entries / pairs / colours) overflowed */
    /* -4 is unused: the library holds no collective (statistics of batched worlds are per-world
     * records, concatenated on the host — sfp_batch_multi_stats) */
    SFP_E_NOMEM = -5
};

/* ---- collider shapes (north_star: circle / capsule / rounded polygon) --------------- */
enum {
    SFP_SHAPE_CIRCLE = 0,  /* point core + radius                                        */
    SFP_SHAPE_CAPSULE = 1, /* segment core (-hx,0)..(+hx,0) in body frame + radius       */
    SFP_SHAPE_RECT = 2,    /* rectangle core [-hx,hx]x[-hy,hy] + rounding radius         */
    SFP_SHAPE_POLYGON = 3  /* convex polygon core (3..8 vertices) + rounding radius: the
                              vertices come from the table set with sfp_set_polygons;
                              SfpBodyDesc.shape = SFP_SHAPE_POLYGON | (table index << 8);
                              hx, hy are ignored                                          */
};
#define SFP_POLYGON_MAX_VERTS 8

/* one entry of the polygon table: body-frame vertices, counter-clockwise, strictly convex, with
 * the centroid at the origin (within 1e-3 of the polygon's size) */
typedef struct SfpPolygon {
    uint32_t count; /* 3..8 */
    uint32_t reserved;
    float x[SFP_POLYGON_MAX_VERTS];
    float y[SFP_POLYGON_MAX_VERTS];
} SfpPolygon; /* 72 bytes */

enum {
    SFP_BODY_STATIC = 1u /* infinite mass; never moved by the solver */
};

/* solver constraint order (oracle honours both; the CUDA path is COLOUR only) */
enum {
    SFP_ORDER_COLOUR = 0,   /* graph-coloured Gauss-Seidel == sequential GS in colour order */
    SFP_ORDER_REFERENCE = 1 /* sequential GS: joints in upload order, then sorted pairs      */
};

enum {
    SFP_JOINT_DISTANCE = 0, /* bilateral: |pb-pa| == rest                                 */
    SFP_JOINT_ROPE = 1      /* unilateral: |pb-pa| <= rest                                */
};

enum {
    SFP_FLAG_BROADPHASE_EVERY_SUBSTEP = 1u, /* default: once per frame     */
    SFP_FLAG_MULTI_KERNEL = 2u,             /* CUDA: one launch per stage instead of the
                                               persistent cooperative substep kernel (debug) */
    SFP_FLAG_TILES = 8u,                    /* solve order: tile-interior constraints first,
                                               then the boundary ones (sfp_math.h, "Tiles")  */
    SFP_FLAG_FLOW_SOLVE = 4u,               /* CUDA: per-body dependency masks instead of a
                                               grid-wide barrier after every colour (same
                                               results, measured slower; an experiment)      */
    SFP_FLAG_BATCH_WHOLE_WORLDS = 16u       /* batches: always one CTA per world, start to end.
                                               Default: a batch with more worlds than CTAs fit
                                               on the GPU at once runs that many CTAs, which
                                               take worlds from a queue — most of them whole,
                                               the last one per CTA a stage at a time,
                                               so that all CTAs end together (same results)   */
};

/* ---- parameters ------------------------------------------------------------------- */
typedef struct SfpParams {
    float gravity_x, gravity_y; /* default (0, -9.81)                                      */
    float cell_size;            /* spatial-hash cell edge, metres (default 0.5)            */
    float aabb_margin;          /* constant AABB expansion, metres (default 0.02)          */
    float vel_expand;           /* AABB swept by vel_expand*dt*v towards the motion (def. 1) */
    float contact_margin;       /* narrowphase keeps points with separation < this (0.01)  */
    uint32_t solver_order;      /* SFP_ORDER_*                                             */
    uint32_t flags;             /* SFP_FLAG_*                                              */
    uint32_t max_bodies;        /* capacities; 0 = choose from the first upload            */
    uint32_t max_joints;
    uint32_t max_entries;       /* (cell,body) broadphase entries                          */
    uint32_t max_pairs;         /* candidate pairs                                         */
    int32_t device;             /* CUDA device ordinal (ignored by the oracle)             */
    uint32_t reserved[3];
} SfpParams; /* 64 bytes */

/* ---- scene-graph side records (AoS, as a scene-graph walk produces them) ------------ */
typedef struct SfpBodyDesc {
    float x, y, angle;      /* pose                                                        */
    float vx, vy, omega;    /* velocity                                                    */
    float hx, hy, radius;   /* collider dimensions, see SFP_SHAPE_*                        */
    uint32_t shape;         /* SFP_SHAPE_*                                                 */
    float density;          /* mass = density * area(shape); ignored for static bodies     */
    float mu_s, mu_d;       /* static / dynamic friction                                   */
    float restitution;
    uint16_t category;      /* collision layer bits of this body                           */
    uint16_t mask;          /* layers it collides with; pair kept iff both ways match      */
    uint32_t flags;         /* SFP_BODY_*                                                  */
} SfpBodyDesc; /* 64 bytes */

typedef struct SfpJointDesc {
    uint32_t body_a, body_b;
    float ax, ay; /* anchor in body_a's frame */
    float bx, by; /* anchor in body_b's frame */
    float rest;   /* rest (distance) / max (rope) length */
    float compliance; /* XPBD alpha, m/N; 0 = rigid */
    uint32_t kind;    /* SFP_JOINT_* */
    uint32_t reserved[3];
} SfpJointDesc; /* 48 bytes */

/* what the scene graph reads back per body */
typedef struct SfpBodyState {
    float x, y;
    float c, s; /* rotation as (cos, sin) — the solver's native representation */
    float vx, vy, omega;
    float angle; /* atan2(s, c), derived on the host at download */
} SfpBodyState; /* 32 bytes */

typedef struct SfpPair {
    uint32_t a, b; /* a < b */
} SfpPair;

/* one contact manifold per candidate pair that touches (1 or 2 points) */
typedef struct SfpContact {
    uint32_t a, b;
    uint32_t count; /* 1 or 2 */
    float nx, ny;   /* unit normal, from a to b */
    float depth[2]; /* penetration depth at narrowphase time (negative = separated, < margin) */
    float px[2], py[2]; /* world-space contact point (midway between the two surfaces) */
    float lambda_n[2];  /* normal multiplier accumulated in the last substep */
    float lambda_t[2];
} SfpContact; /* 60 bytes */

/* what game code hears about a collision */
typedef struct SfpContactEvent {
    uint32_t a, b;   /* a < b */
    float nx, ny;    /* unit normal from a to b */
    float px, py;    /* world-space contact point (first active point of the manifold) */
    float impulse;   /* normal impulse applied in the last substep, N s: (ln0 + ln1) / h */
    uint32_t points; /* active points: 1 or 2 */
} SfpContactEvent; /* 32 bytes */

/* spatial queries */
typedef struct SfpRay {
    float ox, oy;  /* origin */
    float dx, dy;  /* direction, any non-zero length; t is measured in units of it */
    float max_t;
    uint32_t mask; /* a body can be hit iff (its category & mask) != 0 */
    uint32_t reserved[2];
} SfpRay; /* 32 bytes */

typedef struct SfpRayHit {
    uint32_t body; /* 0xffffffff = no hit */
    float t;       /* hit point = origin + t * direction; 0 if the origin is inside the body */
    float nx, ny;  /* surface normal at the hit (minus the ray direction if inside) */
} SfpRayHit; /* 16 bytes */

typedef struct SfpStats {
    uint64_t frames;        /* frames stepped since creation */
    uint64_t body_substeps; /* sum of n_bodies * substeps */
    uint32_t n_bodies, n_dynamic, n_joints;
    uint32_t n_entries;  /* last frame: (cell,body) entries */
    uint32_t n_pairs;    /* last frame: candidate pairs */
    uint32_t n_contacts; /* last substep: contact points */
    uint32_t n_manifolds;/* last substep: pairs with >=1 point */
    uint32_t n_colours;  /* last frame */
    uint32_t colour_rounds; /* last frame: colouring rounds taken */
    uint32_t overflow;   /* sticky: nonzero if a capacity was exceeded */
    float max_penetration;  /* last substep, over contact points, at narrowphase time */
    float kinetic_energy;   /* after last substep */
    float potential_energy; /* -m g.x, after last substep */
    uint32_t launches;      /* kernels launched by the last sfp_step (0 from the oracle) */
    uint32_t reserved[2];
} SfpStats; /* 80 bytes */

typedef struct SfpWorld SfpWorld;
typedef struct SfpBatch SfpBatch;
typedef struct SfpBatchMulti SfpBatchMulti;

/* ---- single world ----------------------------------------------------------------- */

/* Replaces: construction of the reference's `Physics` object (tuning constants + empty
 * body/collider storage).  Reference file:line unavailable — tree not mounted. */
SFP_API int32_t sfp_default_params(SfpParams* out);
SFP_API int32_t sfp_create(const SfpParams* params, SfpWorld** out);
SFP_API int32_t sfp_destroy(SfpWorld* w);

/* Replaces: the scene-graph -> physics sync that hands every `Body` + `Collider` to the
 * step (AoS in, converted to SoA device buffers inside).  Replaces all bodies.
 * Limit: a collider whose box (plus aabb_margin) covers more than 16384 broadphase cells, or 4096
 * or more cells along one axis — at the default cell_size of 0.5 m a floor of more than 8192 m^2
 * or longer than 2048 m — is refused here with SFP_E_INVALID (it would take no part in the
 * broadphase): split it into several static bodies or raise cell_size.  A DYNAMIC body that
 * grows that large while stepping (a runaway) raises overflow bit 16 -> SFP_E_CAPACITY. */
SFP_API int32_t sfp_upload_bodies(SfpWorld* w, const SfpBodyDesc* bodies, uint32_t n);

/* Replaces: the vertex data of the reference's convex-polygon colliders (reference file:line
 * unavailable).  Replaces the whole table; call it BEFORE uploading bodies that use
 * SFP_SHAPE_POLYGON.  Bodies already uploaded keep the indices they were given. */
SFP_API int32_t sfp_set_polygons(SfpWorld* w, const SfpPolygon* polys, uint32_t n);

/* Replaces: registration of the reference's distance / rope constraints. Replaces all joints. */
SFP_API int32_t sfp_set_joints(SfpWorld* w, const SfpJointDesc* joints, uint32_t n);

/* Replaces: the reference's per-frame `Physics` step (frame dt, substep count): broadphase,
 * then `substeps` x { integrate, narrowphase, position solve, velocity update }.
 * Asynchronous on the world's stream until a download / get / sync call. */
SFP_API int32_t sfp_step(SfpWorld* w, float dt, uint32_t substeps);

/* Blocks until all queued work of this world is done; reports sticky errors. */
SFP_API int32_t sfp_sync(SfpWorld* w);

/* Replaces: physics -> scene-graph pose write-back. `out` has room for n states. */
SFP_API int32_t sfp_download_bodies(SfpWorld* w, SfpBodyState* out, uint32_t n);

/* Candidate pair set of the last frame, sorted by (a,b).  *n_out receives the count even
 * when it exceeds cap (then only cap pairs are written). */
SFP_API int32_t sfp_get_pairs(SfpWorld* w, SfpPair* out, uint32_t cap, uint32_t* n_out);

/* Contact manifolds of the last substep, sorted by (a,b). Same cap semantics. */
SFP_API int32_t sfp_get_contacts(SfpWorld* w, SfpContact* out, uint32_t cap, uint32_t* n_out);

/* Per-pair colour of the last frame's constraint colouring, aligned with sfp_get_pairs
 * order; joint colours follow in upload order if joint_colours != NULL. (Test hook.) */
SFP_API int32_t sfp_get_colours(SfpWorld* w, uint32_t* pair_colours, uint32_t cap,
                                uint32_t* joint_colours, uint32_t jcap);

SFP_API int32_t sfp_get_stats(SfpWorld* w, SfpStats* out);
SFP_API const char* sfp_last_error(SfpWorld* w);

/* Replaces: the contact-event list the reference hands to game code after a step (reference
 * file:line unavailable).  One event per candidate pair that carried a normal impulse in the
 * last substep, compacted ON THE DEVICE (only the events cross the bus), sorted by (a, b).
 * Same cap semantics as sfp_get_pairs. */
SFP_API int32_t sfp_get_contact_events(SfpWorld* w, SfpContactEvent* out, uint32_t cap,
                                       uint32_t* n_out);

/* Replaces: the reference's ray / shape queries against the physics world (reference file:line
 * unavailable).  Both work on the CURRENT body state.
 * sfp_ray_cast: m rays at once, one device thread per ray walking the spatial hash (rebuilt for
 * the current poses); closest hit per ray, ties to the lower body index.
 * sfp_query_aabb: indices (ascending) of the bodies whose tight AABB overlaps the box.
 * sfp_query_shape: indices (ascending) of the bodies whose COLLIDER overlaps the probe collider,
 * by the narrowphase itself (zero margin: a point of negative separation).  The probe is given
 * like a body — shape, x, y, angle, hx, hy, radius; a polygon probe indexes the world's polygon
 * table; every other field is ignored.  A circle of radius 0 is a point query (mouse picking). */
SFP_API int32_t sfp_ray_cast(SfpWorld* w, const SfpRay* rays, uint32_t m, SfpRayHit* hits);
SFP_API int32_t sfp_query_aabb(SfpWorld* w, float lo_x, float lo_y, float hi_x, float hi_y,
                               uint32_t mask, uint32_t* out, uint32_t cap, uint32_t* n_out);
SFP_API int32_t sfp_query_shape(SfpWorld* w, const SfpBodyDesc* probe, uint32_t mask, uint32_t* out,
                                uint32_t cap, uint32_t* n_out);

/* Replaces: game code moving individual bodies between steps (teleports, kinematic control).
 * Overwrites pose and velocity of the m bodies idx[0..m) with states[0..m); everything else
 * (shape, mass, material, joints) is untouched.  Indices must be distinct. */
SFP_API int32_t sfp_update_bodies(SfpWorld* w, const uint32_t* idx, const SfpBodyState* states,
                                  uint32_t m);

/* ---- step with host buffers (the end-to-end call the engine makes each frame) -------- */
/* Uploads poses+velocities for n bodies (shape/material unchanged since the last full
 * upload), steps one frame and downloads the new states.  Host<->device copies included. */
SFP_API int32_t sfp_step_host(SfpWorld* w, const SfpBodyState* in, SfpBodyState* out,
                              uint32_t n, float dt, uint32_t substeps);

/* Page-locked host arrays for sfp_step_host / sfp_download_bodies: when `in` / `out` point into
 * memory obtained here the copies go straight over the bus (no staging memcpy on the host).
 * Any other host memory works too, through the library's own pinned staging buffer. */
SFP_API int32_t sfp_alloc_host_states(uint32_t n, SfpBodyState** out);
SFP_API int32_t sfp_free_host_states(SfpBodyState* p);

/* ---- batched independent worlds (BASELINE.json config 5) --------------------------- */
/* n_worlds worlds of exactly bodies_per_world bodies each, stepped together on ONE device
 * (one process per GPU shards the world range; no collective on the data path). */
SFP_API int32_t sfp_batch_create(const SfpParams* params, uint32_t n_worlds,
                                 uint32_t bodies_per_world, SfpBatch** out);
SFP_API int32_t sfp_batch_destroy(SfpBatch* b);
/* bodies: n_worlds * bodies_per_world descs, world-major */
SFP_API int32_t sfp_batch_upload(SfpBatch* b, const SfpBodyDesc* bodies);
/* The polygon table shared by all worlds of the batch (see sfp_set_polygons); call it BEFORE
 * sfp_batch_upload of bodies that use SFP_SHAPE_POLYGON. */
SFP_API int32_t sfp_batch_set_polygons(SfpBatch* b, const SfpPolygon* polys, uint32_t n);
/* Replaces: registration of distance / rope constraints, per world (reference file:line
 * unavailable).  joints: n_worlds * joints_per_world descs, world-major, body indices local to
 * the world; counts (or NULL = every slot used) says how many of a world's slots are in use.
 * Replaces all joints of the batch; joints_per_world = 0 removes them. */
SFP_API int32_t sfp_batch_set_joints(SfpBatch* b, const SfpJointDesc* joints,
                                     uint32_t joints_per_world, const uint32_t* counts);
SFP_API int32_t sfp_batch_step(SfpBatch* b, float dt, uint32_t substeps);
SFP_API int32_t sfp_batch_sync(SfpBatch* b);
SFP_API int32_t sfp_batch_download(SfpBatch* b, SfpBodyState* out);
/* Replaces: the same round trip as sfp_step_host for every world of the batch (reference
 * file:line unavailable): `in` / `out` hold n_worlds * bodies_per_world states, world-major;
 * upload + unpack on the device + one frame + pack + download.  Page-locked arrays from
 * sfp_alloc_host_states are copied by DMA without staging. */
SFP_API int32_t sfp_batch_step_host(SfpBatch* b, const SfpBodyState* in, SfpBodyState* out,
                                    float dt, uint32_t substeps);
/* per-world stats: out has room for n_worlds entries (n_pairs, n_contacts, n_colours ...) */
SFP_API int32_t sfp_batch_stats(SfpBatch* b, SfpStats* out);
SFP_API const char* sfp_batch_last_error(SfpBatch* b);

/* ---- batched worlds over several GPUs from one process ------------------------------- */
/* Replaces: nothing in the reference (a single-process CPU engine; reference file:line
 * unavailable); BASELINE.json config 5 asks for the batch to be partitioned over the GPUs of one
 * box, and an engine is one process.  n_worlds worlds are cut into n_dev contiguous blocks
 * (the first n_worlds % n_dev devices own one world more); device_ids[d] is the CUDA ordinal of
 * block d (an ordinal may repeat).  The library keeps one SfpBatch and one worker thread per
 * block; every call below fans out to the workers and returns when all have finished, with the
 * first error.  Arrays are world-major over ALL worlds, exactly as for sfp_batch_*; there is no
 * exchange between worlds and therefore no collective.  Not thread-safe (one caller at a time). */
SFP_API int32_t sfp_batch_create_multi(const SfpParams* params, uint32_t n_worlds,
                                       uint32_t bodies_per_world, const int32_t* device_ids,
                                       uint32_t n_dev, SfpBatchMulti** out);
SFP_API int32_t sfp_batch_multi_destroy(SfpBatchMulti* m);
/* which worlds block dev_index holds, and on which device (any out pointer may be NULL) */
SFP_API int32_t sfp_batch_multi_partition(SfpBatchMulti* m, uint32_t dev_index, int32_t* device,
                                          uint32_t* first_world, uint32_t* n_worlds);
SFP_API int32_t sfp_batch_multi_upload(SfpBatchMulti* m, const SfpBodyDesc* bodies);
SFP_API int32_t sfp_batch_multi_set_polygons(SfpBatchMulti* m, const SfpPolygon* polys, uint32_t n);
SFP_API int32_t sfp_batch_multi_set_joints(SfpBatchMulti* m, const SfpJointDesc* joints,
                                           uint32_t joints_per_world, const uint32_t* counts);
SFP_API int32_t sfp_batch_multi_step(SfpBatchMulti* m, float dt, uint32_t substeps);
SFP_API int32_t sfp_batch_multi_sync(SfpBatchMulti* m);
SFP_API int32_t sfp_batch_multi_download(SfpBatchMulti* m, SfpBodyState* out);
SFP_API int32_t sfp_batch_multi_step_host(SfpBatchMulti* m, const SfpBodyState* in, SfpBodyState* out,
                                          float dt, uint32_t substeps);
SFP_API int32_t sfp_batch_multi_stats(SfpBatchMulti* m, SfpStats* out);
/* device time of the last step: the slowest block's */
SFP_API int32_t sfp_batch_multi_last_step_ms(SfpBatchMulti* m, float* ms);
SFP_API const char* sfp_batch_multi_last_error(SfpBatchMulti* m);

/* ---- one small world without any allocation ------------------------------------------ */
/* No counterpart in the reference (its step is CPU code and owns no device memory); added for
 * hosts that must not allocate while they run — see the comment on the definition in
 * starframe_b200/csrc/sfp_batch.cu for who they are and the rules that follow.
 *
 * Steps ONE world of n <= SFP_NOALLOC_MAX_BODIES bodies for `frames` frames of `substeps`
 * substeps and writes the resulting states to out[0..n).  Calls no cudaMalloc / cudaFree /
 * allocating routine, on the device or on the heap: every intermediate lives in a fixed
 * SFP_NOALLOC_SCRATCH_BYTES device array that each GPU loading the library holds until exit.
 * Result: bit-identical to sfp_batch_* on one world, including the pair capacity (12 n + 256, or
 * params->max_pairs up to 12 544).  Worlds beyond the limits are refused with SFP_E_CAPACITY.
 * The first call on a device does a one-off set-up (make it before any stream capture).  Calls
 * that name the SAME GPU must not overlap in time (they share the device array and that GPU's
 * host staging slot); calls on different GPUs may.  The call is synchronous and works on the
 * calling thread's cudaStreamPerThread.  err (optional, err_cap bytes) receives a message on
 * failure. */
#define SFP_NOALLOC_MAX_BODIES 1024u
#define SFP_NOALLOC_SCRATCH_BYTES (2304u * 1024u)
SFP_API int32_t sfp_step_small_noalloc(const SfpParams* params, const SfpBodyDesc* bodies,
                                       uint32_t n, float dt, uint32_t substeps, uint32_t frames,
                                       SfpBodyState* out, char* err, uint32_t err_cap);

/* device-time of the last sfp_step / sfp_batch_step in milliseconds, measured with CUDA
 * events on the world's own stream (0 from the oracle) */
SFP_API int32_t sfp_last_step_ms(SfpWorld* w, float* ms);
SFP_API int32_t sfp_batch_last_step_ms(SfpBatch* b, float* ms);

/* device-time breakdown of the last sfp_step, CUDA events on the world's stream:
 * out[0] = whole step, out[1] = broadphase (AABBs, cell sort, pair emission),
 * out[2] = colouring + bucketing by colour, out[3] = the persistent substep kernel. */
SFP_API int32_t sfp_last_step_breakdown_ms(SfpWorld* w, float out[4]);

/* wall time of the stages INSIDE the persistent substep kernel of the last sfp_step, in
 * nanoseconds summed over the frame's substeps, as seen by one thread (%globaltimer; the stages
 * are grid-wide phases, so one observer is representative): out[0] integrate, out[1] narrowphase,
 * out[2] position solve (all colour phases), out[3] velocity derivation, out[4] velocity solve,
 * out[5] the small colours of both passes (run by one block inside one phase). */
SFP_API int32_t sfp_last_step_stage_ns(SfpWorld* w, uint64_t out[6]);

/* library identification: "cuda-sm_100a" or "cpu-oracle" */
SFP_API const char* sfp_backend(void);

/* ---- oracle twins (libsfp_ref.so) -------------------------------------------------- */
SFP_API int32_t sfp_ref_default_params(SfpParams* out);
SFP_API int32_t sfp_ref_create(const SfpParams* params, SfpWorld** out);
SFP_API int32_t sfp_ref_destroy(SfpWorld* w);
SFP_API int32_t sfp_ref_upload_bodies(SfpWorld* w, const SfpBodyDesc* bodies, uint32_t n);
SFP_API int32_t sfp_ref_set_joints(SfpWorld* w, const SfpJointDesc* joints, uint32_t n);
SFP_API int32_t sfp_ref_set_polygons(SfpWorld* w, const SfpPolygon* polys, uint32_t n);
SFP_API int32_t sfp_ref_step(SfpWorld* w, float dt, uint32_t substeps);
SFP_API int32_t sfp_ref_sync(SfpWorld* w);
SFP_API int32_t sfp_ref_download_bodies(SfpWorld* w, SfpBodyState* out, uint32_t n);
SFP_API int32_t sfp_ref_get_pairs(SfpWorld* w, SfpPair* out, uint32_t cap, uint32_t* n_out);
SFP_API int32_t sfp_ref_get_contacts(SfpWorld* w, SfpContact* out, uint32_t cap,
                                     uint32_t* n_out);
SFP_API int32_t sfp_ref_get_colours(SfpWorld* w, uint32_t* pair_colours, uint32_t cap,
                                    uint32_t* joint_colours, uint32_t jcap);
SFP_API int32_t sfp_ref_get_stats(SfpWorld* w, SfpStats* out);
SFP_API int32_t sfp_ref_get_contact_events(SfpWorld* w, SfpContactEvent* out, uint32_t cap,
                                           uint32_t* n_out);
SFP_API int32_t sfp_ref_ray_cast(SfpWorld* w, const SfpRay* rays, uint32_t m, SfpRayHit* hits);
SFP_API int32_t sfp_ref_query_aabb(SfpWorld* w, float lo_x, float lo_y, float hi_x, float hi_y,
                                   uint32_t mask, uint32_t* out, uint32_t cap, uint32_t* n_out);
SFP_API int32_t sfp_ref_query_shape(SfpWorld* w, const SfpBodyDesc* probe, uint32_t mask,
                                    uint32_t* out, uint32_t cap, uint32_t* n_out);
SFP_API int32_t sfp_ref_update_bodies(SfpWorld* w, const uint32_t* idx, const SfpBodyState* states,
                                      uint32_t m);
SFP_API const char* sfp_ref_last_error(SfpWorld* w);
SFP_API int32_t sfp_ref_step_host(SfpWorld* w, const SfpBodyState* in, SfpBodyState* out,
                                  uint32_t n, float dt, uint32_t substeps);
SFP_API int32_t sfp_ref_last_step_ms(SfpWorld* w, float* ms);
SFP_API const char* sfp_ref_backend(void);
/* oracle-only: brute-force O(n^2) candidate pairs of the CURRENT state (property tests) */
SFP_API int32_t sfp_ref_brute_pairs(SfpWorld* w, float dt, SfpPair* out, uint32_t cap,
                                    uint32_t* n_out);

#ifdef __cplusplus
}
#endif
#endif /* SFP_H_ */
