/* b2col.h — C ABI of the B200 collision-query backend (libb2col.so).
 *
 * Drop-in boundary for ONE path of bailwillharr/engine: the collision system's AABB BVH
 *   world-matrix propagation   reference src/system_transform.cpp:12-38   (TransformSystem::onUpdate)
 *   world-space AABB refit     reference src/system_collisions.cpp:14-41,123-132 (transformBox + refit loop)
 *   BVH build                  reference src/system_collisions.cpp:205-376 (BuildNode; replaced by an LBVH)
 *   nearest-hit ray queries    reference src/system_collisions.cpp:62-100,160-199,379-540
 *                              (RayBoxIntersection, GetRaycast, RaycastTreeNode)
 * The reference has no FFI for this path: the boundary is the C++ class engine::CollisionSystem
 * (reference include/system_collisions.h:28-74).  engine_b200/shim/system_collisions.{h,cpp} is the
 * replacement class a maintainer compiles instead of the reference's; its bodies call the entry
 * points below.  Everything here is POD + plain pointers, extern "C", no
 * exceptions cross the boundary.  There is no CPU fallback: without a CUDA device b2col_create
 * fails with B2COL_ERR_NO_DEVICE.
 *
 * Conventions
 *  - Entities are uint32 IDs, 0 is invalid, first valid ID is 1 (reference include/entity.h:7,
 *    include/scene.h:141).  Per-entity arrays are indexed BY ENTITY ID and therefore hold
 *    n_entities + 1 rows, row 0 ignored — the layout of the reference's ComponentArray<T>
 *    (include/ecs.h:25-50).
 *  - Quaternions are x,y,z,w (glm::quat memory order); matrices are 16 floats column-major
 *    (glm::mat4 memory order).
 *  - A context owns one CUDA stream; calls on one context are serialised on it.  A context is
 *    not thread-safe.  One context per GPU.
 *  - Pointers named h_* are host memory, d_* are device memory on the context's device.  The
 *    caller owns every buffer it passes; nothing is retained after the call returns except
 *    where a *_device call documents that the work is merely enqueued.
 */
#ifndef B2COL_H
#define B2COL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2COL_ABI_VERSION 4 /* 3: b2col_stats grew by key_bits, fused_keys; 4: + b2col_raycast_batch_static, b2col_bind_trs_device */

typedef struct b2col_ctx b2col_ctx;

typedef enum b2col_status {
    B2COL_OK = 0,
    B2COL_ERR_INVALID = 1,     /* bad argument */
    B2COL_ERR_CUDA = 2,        /* a CUDA call failed; see b2col_last_error */
    B2COL_ERR_NO_DEVICE = 3,   /* no usable CUDA device: this library has no CPU path */
    B2COL_ERR_STATE = 4,       /* call sequence error (e.g. update before set_*) */
    B2COL_ERR_UNSUPPORTED = 5, /* e.g. a parent whose ID is not smaller than its child's */
    B2COL_ERR_OVERFLOW = 6,    /* traversal stack overflow in reference-topology mode */
    B2COL_ERR_COMM = 7         /* NCCL could not be loaded, or an NCCL call failed; see b2col_last_error */
} b2col_status;

/* engine::Ray — reference include/system_collisions.h:13-16 (24 bytes) */
typedef struct b2col_ray {
    float origin[3];
    float direction[3]; /* need not be normalised: GetRaycast normalises (:162) */
} b2col_ray;

/* engine::Raycast — reference include/system_collisions.h:18-24 (36 bytes; bool + 3 pad).
 * A miss is all-zero (value-initialised, reference src/system_collisions.cpp:164-165). */
typedef struct b2col_raycast {
    float location[3];
    float normal[3];
    uint32_t hit_entity;
    float distance; /* = tmin of the slab test; NEGATIVE when the origin is inside the box (:99) */
    uint8_t hit;
    uint8_t pad[3]; /* always written as 0 */
} b2col_raycast;

/* Compact result (additive; the reference has no such type): what a Raycast holds that is not a function of the ray itself.
 * 12 instead of 36 bytes per ray leave the device; b2col_expand_hits restores the Raycast on the host, bit for bit. */
typedef struct b2col_hit {
    uint32_t hit_entity;
    float distance;
    uint32_t flags; /* bit 0: hit; bits 8-10: engine::AABBSide of the entry face (Left, Right, Bottom, Top, Front, Back) */
} b2col_hit;

/* engine::AABB — reference include/component_collider.h:7-10 (24 bytes) */
typedef struct b2col_aabb {
    float min[3];
    float max[3];
} b2col_aabb;

/* engine::CollisionSystem::BiTreeNode — reference include/system_collisions.h:40-48 (60 bytes, align 4) */
enum { B2COL_NODE_BOUNDING_VOLUME = 0, B2COL_NODE_ENTITY = 1, B2COL_NODE_EMPTY = 2 };
typedef struct b2col_bitreenode {
    b2col_aabb box1, box2;
    uint32_t index1, index2; /* node index or entity ID, depending on type */
    uint8_t type1, type2;
    uint8_t pad[2];
} b2col_bitreenode;

typedef struct b2col_config {
    int device;              /* CUDA device ordinal */
    float max_distance;      /* reference hard-codes T = 1000.0f (src/system_collisions.cpp:384); 0 => 1000 */
    int morton_bits_per_axis; /* 4..21 (key = 3*bits); 0 => automatic: 32-bit keys up to 2M colliders, 40 up to 32M, else 48 */
    int use_cuda_graph;      /* 1: b2col_update replays a captured CUDA graph when shapes are unchanged */
} b2col_config;

typedef enum b2col_mode {
    B2COL_MODE_LBVH = 0,    /* production: tree built on the GPU by b2col_update */
    B2COL_MODE_REFTOPO = 1, /* walk a reference-format tree uploaded with b2col_upload_ref_bvh
                               (exact reference semantics incl. tie winners; parity mode) */
    B2COL_MODE_LBVH_COUNT = 2 /* diagnostic: B2COL_MODE_LBVH (identical results) that also counts node visits
                               (b2col_get_trace_counters); synchronises, a little slower — never the timed path */
} b2col_mode;

typedef struct b2col_stats {
    uint32_t n_entities, n_colliders, n_levels;
    uint32_t n_internal_nodes;  /* LBVH internal nodes (= max(n_colliders,2) - 1) */
    uint32_t root;              /* index of the LBVH root node */
    uint32_t sort_passes;
    uint64_t kernel_launches;   /* cumulative count of this library's kernel launches on this ctx */
    uint64_t bvh_bytes;         /* device bytes of the traversal structure (nodes + leaf entity table) */
    uint64_t full_builds, refits; /* b2col_update / b2col_refit calls that built or re-bounded a tree */
    uint64_t graph_replays;     /* frames that ran as a replay of a captured CUDA graph (cfg.use_cuda_graph) */
    float scene_min[3], scene_max[3]; /* centroid bounds used for the Morton grid */
    uint32_t key_bits;          /* Morton key width of the last full build (b2col_update adapts it to the scene: one radix pass per 8 bits) */
    uint32_t fused_keys;        /* 1: the last full build took its keys from a Morton plan made before the frame (scene_min / scene_max are that plan's bounds) */
} b2col_stats;

/* ---- lifecycle ------------------------------------------------------------------------------ */
int b2col_abi_version(void);
/* CUDA devices visible to this process (0 and B2COL_ERR_NO_DEVICE when there is none: the library has no CPU path) */
b2col_status b2col_device_count(int* out);
/* replaces: CollisionSystem::CollisionSystem(Scene*) — reference src/system_collisions.cpp:104 */
b2col_status b2col_create(b2col_ctx** out, const b2col_config* cfg /* may be NULL */);
void b2col_destroy(b2col_ctx* ctx);
/* message for the last non-OK status on this ctx (ctx may be NULL for create failures) */
const char* b2col_last_error(const b2col_ctx* ctx);
b2col_status b2col_sync(b2col_ctx* ctx);
/* the context's cudaStream_t, for callers that time with their own events */
void* b2col_stream(b2col_ctx* ctx);

/* ---- scene input ---------------------------------------------------------------------------- */
/* Hierarchy (TransformComponent::parent, reference include/component_transform.h:22).
 * h_parent[0..n_entities]; parent 0 = root.  Every parent ID must be smaller than its child's (true
 * for every entity-creating path of the reference: Scene::CreateEntity hands out ascending IDs,
 * src/scene.cpp:37); otherwise B2COL_ERR_UNSUPPORTED. */
b2col_status b2col_set_hierarchy(b2col_ctx* ctx, uint32_t n_entities, const uint32_t* h_parent);
/* TRS (TransformComponent::position/rotation/scale).  Rows 0..n_entities. */
b2col_status b2col_set_trs(b2col_ctx* ctx, uint32_t n_entities, const float* h_pos3, const float* h_rot_xyzw, const float* h_scl3);
/* Same, from device memory (async on the ctx stream; buffers must stay valid until b2col_sync). */
b2col_status b2col_set_trs_device(b2col_ctx* ctx, uint32_t n_entities, const float* d_pos3, const float* d_rot_xyzw, const float* d_scl3);
/* Zero-copy form for callers whose TransformComponent data already lives on the GPU (a simulation that animates it there):
 * the context does not copy (b2col_set_trs_device: 44 bytes per entity read and written again every frame — 0.05 ms at 1M
 * entities, a third of the update it feeds) but READS THE CALLER'S ARRAYS IN PLACE in every b2col_update / b2col_refit /
 * b2col_propagate until the next b2col_set_trs* / b2col_set_hierarchy.  The arrays (rows 0..n_entities; rotations 16-byte
 * aligned) must stay valid for that long, and the caller orders its writes before the update and the update before its next
 * writes (same stream as b2col_stream, or events).  The sparse row calls below return B2COL_ERR_STATE while arrays are bound. */
b2col_status b2col_bind_trs_device(b2col_ctx* ctx, uint32_t n_entities, const float* d_pos3, const float* d_rot_xyzw, const float* d_scl3);
/* Sparse TRS update for frames in which few entities moved (dirty tracking through
 * TransformComponent::is_static, reference include/component_transform.h:24): row h_entity[i] of the device arrays takes
 * pos / rot / scl [i]; everything else keeps the values of the last b2col_set_trs.  44 bytes per moved entity cross PCIe
 * instead of 44 bytes per entity.  Enqueued; the caller's arrays are free on return.
 * The _stage / _commit pair is the zero-copy form: _stage hands out page-locked arrays of n_rows rows owned by the ctx,
 * the caller gathers its AoS components straight into them, _commit(n <= n_rows) uploads and scatters them. */
b2col_status b2col_update_trs_rows(b2col_ctx* ctx, uint32_t n_rows, const uint32_t* h_entity, const float* h_pos3, const float* h_rot_xyzw, const float* h_scl3);
b2col_status b2col_stage_trs_rows(b2col_ctx* ctx, uint32_t n_rows, uint32_t** h_entity, float** h_pos3, float** h_rot_xyzw, float** h_scl3);
b2col_status b2col_commit_trs_rows(b2col_ctx* ctx, uint32_t n_rows);
/* Alternative to hierarchy+TRS when the caller keeps the reference's CPU TransformSystem:
 * upload TransformComponent::world_matrix directly (rows 0..n_entities, 16 floats each);
 * b2col_update then skips propagation. */
b2col_status b2col_set_world_matrices(b2col_ctx* ctx, uint32_t n_entities, const float* h_world16);
/* Collider set = CollisionSystem::m_entities + ColliderComponent::aabb
 * (reference src/system_collisions.cpp:104,125-127).  Entities must be valid IDs; the reference
 * iterates them in ascending order (std::set) and so should the caller, but any order is accepted. */
b2col_status b2col_set_colliders(b2col_ctx* ctx, uint32_t n_colliders, const uint32_t* h_entity, const b2col_aabb* h_local);

/* ---- per-frame update ----------------------------------------------------------------------- */
/* replaces: TransformSystem::onUpdate (src/system_transform.cpp:12-38) followed by
 * CollisionSystem::onUpdate (src/system_collisions.cpp:112-158): propagate world matrices, refit
 * world AABBs, rebuild the BVH from scratch (the reference rebuilds every frame, :121).
 * Enqueued on the ctx stream; returns without waiting.  n_colliders == 0 yields an empty tree and
 * every query misses (the reference abort()s, :136 -> :208 — deliberate divergence).
 * Successive updates of one collider set adapt the build to the scene (same answers; only tree shape, i.e. the winner among
 * exactly tied hits, and the frame time can differ from one frame to the next): the Morton key width follows the key
 * collisions the previous build counted, and from the second frame on the keys are made inside the refit kernel from a plan
 * based on the bounds earlier frames reported (b2col_stats.key_bits / fused_keys; B2COL_ADAPTIVE_KEYS=0, B2COL_FUSE_KEYS=0
 * pin the static width / the exact per-frame plan; contexts with a communicator always use both). */
b2col_status b2col_update(b2col_ctx* ctx);
/* replaces: TransformSystem::onUpdate alone (src/system_transform.cpp:12-38) — for callers that replace the
 * reference's TransformSystem and CollisionSystem separately (engine_b200/shim/system_transform.cpp): world
 * matrices from the current hierarchy + TRS, enqueued on the ctx stream.  A following b2col_update reuses them
 * (no second propagation) until the TRS or the hierarchy are set again. */
b2col_status b2col_propagate(b2col_ctx* ctx);
/* The refit-only frame the reference leaves as a TODO (src/system_collisions.cpp:116-120: "only rebuild the
 * BVH if an entity has moved / been added"): propagate world matrices, refit world AABBs
 * and re-bound the EXISTING tree bottom-up — same topology and leaf order as the last b2col_update, new boxes —
 * skipping the Morton pass and the radix sort.  Query results stay exact (the tree is still a valid bounding
 * hierarchy of the current boxes and ties keep resolving by leaf order); only traversal cost grows as the
 * colliders drift from where they were sorted, so callers interleave it with b2col_update (the shim's
 * rebuild_interval).  B2COL_ERR_STATE unless b2col_update has built a tree on this ctx for the current
 * collider set.  Enqueued on the ctx stream. */
b2col_status b2col_refit(b2col_ctx* ctx);

/* ---- queries -------------------------------------------------------------------------------- */
/* replaces: CollisionSystem::GetRaycast (src/system_collisions.cpp:160-199), batched.
 * Host buffers; copies in, traces, copies out, returns when h_out is complete. */
b2col_status b2col_raycast_batch(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, b2col_raycast* h_out, int mode);
/* Device buffers; enqueued on the ctx stream, returns without waiting. */
b2col_status b2col_raycast_device(b2col_ctx* ctx, const b2col_ray* d_rays, size_t n, b2col_raycast* d_out, int mode);
/* No counterpart in the reference.  b2col_raycast_batch (production tree, same bytes
 * in h_out) for callers that cannot call cudaMalloc — tight serving loops, harnesses that forbid it: this call allocates NOTHING
 * (no cudaMalloc / cudaFree / allocating library routine).  All device scratch of the trace — the persistent kernel's work counter
 * and the staging of the rays and of the results — lives in one file-scope `static __device__ __align__(256)` array of
 * 256 + 60 * B2COL_STATIC_MAX_RAYS bytes (3.75 MiB) that every GPU which loads the library holds until exit.  n >
 * B2COL_STATIC_MAX_RAYS is rejected with B2COL_ERR_INVALID (split the batch, or use b2col_raycast_batch).  Every call on a GPU
 * shares the array, which keeps the last call's contents (the call clears what must start at zero): calls that overlap on one
 * GPU — two contexts, two host threads — race; calls on one context are serialised as always.  Returns when h_out is complete.
 * The tree must have been built before (b2col_update, which does allocate); before that every ray misses, as in the reference. */
#define B2COL_STATIC_MAX_RAYS 65536
b2col_status b2col_raycast_batch_static(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, b2col_raycast* h_out);

/* Compact-result forms of the two calls above (production tree): 12-byte b2col_hit records.  The host-buffer path is bound by
 * the device->host copy of its results (36 of 60 bytes per ray), so a caller that can consume hits — or expand them on its own
 * cores with b2col_expand_hits — moves 36 instead of 60 bytes per ray over PCIe. */
b2col_status b2col_raycast_hits_batch(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, b2col_hit* h_out);
b2col_status b2col_raycast_hits_device(b2col_ctx* ctx, const b2col_ray* d_rays, size_t n, b2col_hit* d_out);
/* replaces: the epilogue of GetRaycast (src/system_collisions.cpp:162,170-194,489) on the HOST: h_out[i] = the Raycast that
 * b2col_raycast_batch would have returned for h_rays[i] (same bytes: direction normalised again, location = d*t + o, normal from
 * the side).  Pure host function, n_threads worker threads. */
b2col_status b2col_expand_hits(const b2col_ray* h_rays, const b2col_hit* h_hits, size_t n, b2col_raycast* h_out, int n_threads);

/* ---- reference-topology mode / debug overlay ------------------------------------------------- */
/* Upload a tree in the reference's own format (e.g. CollisionSystem::bvh_), root = LAST node
 * (src/system_collisions.cpp:168). */
b2col_status b2col_upload_ref_bvh(b2col_ctx* ctx, const b2col_bitreenode* h_nodes, size_t n_nodes);
/* replaces: the public member CollisionSystem::bvh_ (include/system_collisions.h:58) read by the
 * debug overlay (src/application.cpp:61-62,153-154).  Writes the GPU-built tree in the reference's
 * format: {BoundingVolume,BoundingVolume} / {Entity,Entity} / {Entity,Empty} nodes only, children
 * before parents, root last, so the reference's own GetRaycast can walk it.  capacity in nodes;
 * 2*n_colliders is always enough.  *n_out = nodes written. */
b2col_status b2col_export_bitreenode(b2col_ctx* ctx, b2col_bitreenode* h_out, size_t capacity, size_t* n_out);

/* ---- read-back (parity tests; world_matrix write-back for MeshRenderSystem) ------------------ */
/* rows first_entity .. first_entity+count-1, 16 floats each, column-major */
b2col_status b2col_get_world_matrices(b2col_ctx* ctx, uint32_t first_entity, uint32_t count, float* h_out16);
/* world_matrix write-back of the rows that can have moved (TransformComponent::world_matrix, read by MeshRenderSystem,
 * reference src/system_mesh_render.cpp:62): *h_rows16 = n_rows x 16 floats, row i = world matrix of h_entity[i], in
 * page-locked memory owned by the ctx (valid until the next *_rows call).  Gathered on the device, one D2H copy. */
b2col_status b2col_fetch_world_rows(b2col_ctx* ctx, uint32_t n_rows, const uint32_t* h_entity, const float** h_rows16);
/* world AABBs in the order colliders were given to b2col_set_colliders */
b2col_status b2col_get_world_aabbs(b2col_ctx* ctx, b2col_aabb* h_out);
/* collider indices (into the b2col_set_colliders order) in BVH leaf order == Morton order ==
 * depth-first leaf order of the tree, the order that decides tie winners */
b2col_status b2col_get_leaf_order(b2col_ctx* ctx, uint32_t* h_out);
b2col_status b2col_get_stats(b2col_ctx* ctx, b2col_stats* out);
/* copy `bytes` of device memory owned by this ctx (e.g. the gathered Raycasts of b2col_comm_raycast_gather) to the host, ordered
 * after everything enqueued on the ctx stream; returns when h_dst is complete */
b2col_status b2col_read_device(b2col_ctx* ctx, const void* d_src, size_t bytes, void* h_dst);
/* totals over all B2COL_MODE_LBVH_COUNT calls on this ctx: BVH nodes fetched (one 32-byte sector each — the
 * unit the traversal's binding resource, L1 wavefronts, is counted in) and rays traced */
b2col_status b2col_get_trace_counters(b2col_ctx* ctx, uint64_t* node_visits, uint64_t* rays);

/* ---- multi-GPU: rays shard across the GPUs of one box, the BVH is replicated, hits are gathered back ------------
 * replaces: nothing in the reference, which is single-threaded on one CPU core (src/application.cpp:390); this is the
 * scale-out of CollisionSystem::GetRaycast (include/system_collisions.h:36, src/system_collisions.cpp:160-199) that
 * BASELINE.json's north_star asks for.  Rays are independent and the tree is read-only during
 * queries, so a batch is cut into contiguous slices — b2col_shard_bounds — one per rank, the results land in the same
 * slices and the gathered output is bit-identical to a 1-GPU trace of the whole batch.
 *
 * A "rank" is one context on one GPU.  Two deployments:
 *   - one process per GPU (torchrun): rank 0 calls b2col_comm_get_unique_id, the application distributes the 128 bytes
 *     (any transport), every rank calls b2col_comm_init_rank; afterwards b2col_comm_replicate / _raycast_gather are
 *     collectives: every rank calls them with the same arguments, in the same order.
 *   - one process driving n GPUs (a game engine: engine_b200/shim): b2col_comm_init_all over n contexts, then the
 *     *_all entry points (one NCCL group per call) and b2col_mgpu_raycast_batch for host buffers.
 * NCCL (libnccl.so.2) is loaded on first use; without it these calls fail with B2COL_ERR_COMM and nothing else is
 * affected. */
#define B2COL_COMM_ID_BYTES 128 /* sizeof(ncclUniqueId) */
/* rank's slice [*lo, *hi) of an n-element batch: lo = n*rank/n_ranks, hi = n*(rank+1)/n_ranks (pure function) */
void b2col_shard_bounds(size_t n, int n_ranks, int rank, size_t* lo, size_t* hi);
b2col_status b2col_comm_get_unique_id(void* id128);
b2col_status b2col_comm_init_rank(b2col_ctx* ctx, int n_ranks, int rank, const void* id128);
/* n contexts of this process, each on its own device; context i becomes rank i.  Enables peer access between them. */
b2col_status b2col_comm_init_all(b2col_ctx* const* ctxs, int n);
/* releases the communicator and the gather window (b2col_destroy does it too) */
b2col_status b2col_comm_finalize(b2col_ctx* ctx);

/* Replicate the tree of src_rank (built by b2col_update there) on every rank: ONE ncclBroadcast over NVLink straight
 * out of the source's BVH allocation into the other ranks' — no staging copy, no size exchange, no host round trip;
 * every rank passes the collider count (it sizes the allocation).  Enqueued on the ctx stream.  Receiving ranks can
 * trace the tree but not export / refit it (the sorted keys stay on src_rank). */
b2col_status b2col_comm_replicate(b2col_ctx* ctx, int src_rank, uint32_t n_colliders);
b2col_status b2col_comm_replicate_all(b2col_ctx* const* ctxs, int n, int src_rank);
/* The alternative SURVEY.md 8(e) asks to measure against it: broadcast the 32-byte world records (half the bytes of the
 * tree) and run Morton + sort + hierarchy on every rank; the build is deterministic, so the trees are identical. */
b2col_status b2col_comm_replicate_records(b2col_ctx* ctx, int src_rank, uint32_t n_colliders);

/* Sharded GetRaycast with the gather fused into the traversal: d_rays_shard holds THIS rank's slice of an n_total-ray
 * batch (device memory); the nearest hits of the whole batch end up in batch order in device memory on dst_rank,
 * *d_out_all (set on dst_rank only; owned by the context, valid until its next gather with a larger batch or
 * b2col_comm_finalize).  The trace kernel of every rank writes its Raycasts straight into dst_rank's buffer over
 * NVLink (peer mapping), so nothing is sent after the traversal; the call ends with the completion step of the
 * collective on the ctx stream.  Enqueued; b2col_sync on dst_rank (after the other ranks enqueued theirs) waits for
 * the whole batch. */
b2col_status b2col_comm_raycast_gather(b2col_ctx* ctx, const b2col_ray* d_rays_shard, size_t n_total, int dst_rank, const b2col_raycast** d_out_all);
b2col_status b2col_comm_raycast_gather_all(b2col_ctx* const* ctxs, int n, const b2col_ray* const* d_rays_shards, size_t n_total, int dst_rank,
                                           const b2col_raycast** d_out_all);
/* replaces: CollisionSystem::GetRaycast over a batch in HOST memory, on several GPUs of this process: slice i is
 * uploaded, traced and downloaded by context i, straight from / into the caller's buffers (nothing to gather).
 * Contexts need not have a communicator, only equal trees (b2col_comm_replicate_all, or the same b2col_update). */
b2col_status b2col_mgpu_raycast_batch(b2col_ctx* const* ctxs, int n, const b2col_ray* h_rays, size_t n_rays, b2col_raycast* h_out, int mode);

enum { B2COL_GATHER_NONE = 0, B2COL_GATHER_PEER_WINDOW = 1, B2COL_GATHER_PEER_WINDOW_IPC = 2, B2COL_GATHER_NCCL_SENDRECV = 3 };
typedef struct b2col_comm_info_t {
    int n_ranks, rank;
    int single_process;   /* 1: created by b2col_comm_init_all */
    int gather_mode;      /* B2COL_GATHER_*: how b2col_comm_raycast_gather moves the hits (3 = fallback: IPC mapping unavailable) */
    int nccl_version;
    uint64_t window_bytes, gathers, replications;
} b2col_comm_info_t;
b2col_status b2col_comm_info(b2col_ctx* ctx, b2col_comm_info_t* out);

/* ---- single-allocation BVH blob (tools, tests; the collectives above use it directly) --------------------------- */
/* The traversal structure lives in ONE contiguous device allocation.  b2col_bvh_blob returns its device pointer and
 * size; export copies it into caller-owned device memory; import copies a blob of matching size into this ctx and
 * makes it the current tree (the leaf count follows from the size; the header is checked at the next b2col_sync). */
b2col_status b2col_bvh_blob(b2col_ctx* ctx, void** d_ptr, size_t* bytes);
/* copy the blob into caller-owned device memory (e.g. the tensor a collective will send); enqueued */
b2col_status b2col_bvh_export_blob(b2col_ctx* ctx, void* d_dst, size_t bytes);
b2col_status b2col_bvh_import_blob(b2col_ctx* ctx, const void* d_src, size_t bytes);

/* ---- timing helpers (CUDA events on the ctx stream) ------------------------------------------ */
/* slot in [0,16) */
b2col_status b2col_event_record(b2col_ctx* ctx, int slot);
b2col_status b2col_event_elapsed_ms(b2col_ctx* ctx, int slot_begin, int slot_end, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* B2COL_H */
