// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// C-callable harness around the UNMODIFIED reference collision path. It is compiled
// (see oracle/Makefile) together with the reference's own translation units, taken from
// where they lie under /root/reference, into oracle/_ref/libref_engine.so.  Nothing in
// engine_b200/ links, loads or calls it; only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs do, as the checker and the CPU baseline.
//
// The reference's system_collisions.cpp is pulled in textually so that its file-static
// helpers (transformBox :14, RayBoxIntersection :62) can be called directly for
// known-answer tests and for the O(N) refit at sizes where the reference's O(N^2)
// BuildNode (:205-376) cannot run.  No reference source is copied into this repository.
// Built a second time with -DB2COL_DROPIN (oracle/_ref/libdropin_engine.so): the same harness and the
// same reference Scene / ECS / TransformSystem translation units, but with engine_b200/shim's
// replacement CollisionSystem compiled in place of the reference's — the drop-in test.
#ifdef B2COL_DROPIN
#include "system_collisions.h"  // engine_b200/shim precedes the reference's include directory
#include "component_transform.h"
#include "scene.h"
#else
#include "system_collisions.cpp"  // -I/root/reference/src
#endif

#include "system_transform.h"

#include <chrono>
#include <cstring>
#include <thread>

using namespace engine;

#include "harness_types.h"  // struct Harness: the Scene (+ a headless Application stand-in nothing on this path dereferences)

namespace {
static_assert(sizeof(Ray) == 24, "Ray layout");
static_assert(sizeof(Raycast) == 36, "Raycast layout");
static_assert(sizeof(AABB) == 24, "AABB layout");
static_assert(sizeof(CollisionSystem::BiTreeNode) == 60, "BiTreeNode layout");
}  // namespace

extern "C" {

void* refh_create() { return new Harness(); }
void refh_destroy(void* h) { delete static_cast<Harness*>(h); }

// which: 0 Ray, 1 Raycast, 2 AABB, 3 BiTreeNode, 4 TransformComponent, 5 PrimitiveInfo
size_t refh_sizeof(int which)
{
    switch (which) {
        case 0: return sizeof(Ray);
        case 1: return sizeof(Raycast);
        case 2: return sizeof(AABB);
        case 3: return sizeof(CollisionSystem::BiTreeNode);
        case 4: return sizeof(TransformComponent);
#ifndef B2COL_DROPIN
        case 5: return sizeof(CollisionSystem::PrimitiveInfo);
#endif
    }
    return 0;
}

// Scene::CreateEntity (src/scene.cpp:35-53). parent[i]==0 => root. Quaternions arrive as
// x,y,z,w (glm's memory order); glm::quat's constructor takes (w,x,y,z).
// Returns the ID of the first entity created; IDs are consecutive.
uint32_t refh_create_entities(void* h, size_t n, const uint32_t* parent, const float* pos3, const float* rot_xyzw, const float* scl3)
{
    auto* H = static_cast<Harness*>(h);
    uint32_t first = 0;
    for (size_t i = 0; i < n; ++i) {
        const glm::vec3 p{pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]};
        const glm::quat q{rot_xyzw[4 * i + 3], rot_xyzw[4 * i], rot_xyzw[4 * i + 1], rot_xyzw[4 * i + 2]};
        const glm::vec3 s{scl3[3 * i], scl3[3 * i + 1], scl3[3 * i + 2]};
        const Entity e = H->scene.CreateEntity("e", parent ? parent[i] : 0u, p, q, s);
        if (i == 0) first = e;
    }
    return first;
}

// Scene::AddComponent<ColliderComponent> (include/scene.h:80-102), aabb = min.xyz,max.xyz
void refh_add_colliders(void* h, size_t n, const uint32_t* entity, const float* aabb6)
{
    auto* H = static_cast<Harness*>(h);
    for (size_t i = 0; i < n; ++i) {
        auto* c = H->scene.AddComponent<ColliderComponent>(entity[i]);
        std::memcpy(&c->aabb, aabb6 + 6 * i, sizeof(AABB));
        H->colliders.push_back(entity[i]);
    }
}

void refh_set_trs(void* h, size_t n, const uint32_t* entity, const float* pos3, const float* rot_xyzw, const float* scl3)
{
    auto* H = static_cast<Harness*>(h);
    for (size_t i = 0; i < n; ++i) {
        auto* t = H->scene.GetTransform(entity[i]);
        t->position = glm::vec3{pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]};
        t->rotation = glm::quat{rot_xyzw[4 * i + 3], rot_xyzw[4 * i], rot_xyzw[4 * i + 1], rot_xyzw[4 * i + 2]};
        t->scale = glm::vec3{scl3[3 * i], scl3[3 * i + 1], scl3[3 * i + 2]};
    }
}

// TransformComponent::is_static (include/component_transform.h:24): set by loadGLTF for static scenes; the reference's own
// TransformSystem / CollisionSystem ignore it, the drop-in TransformSystem uploads such rows once.
void refh_set_static(void* h, size_t n, const uint32_t* entity, int is_static)
{
    auto* H = static_cast<Harness*>(h);
    for (size_t i = 0; i < n; ++i) H->scene.GetTransform(entity[i])->is_static = is_static != 0;
}

// knobs / counters of the drop-in shims (no-ops in the reference build). key: 0 collider_check_interval, 1 rebuild_interval,
// 2 skip_static_rows, 3 write_back_world_matrices, 4 static_rescan_interval, 5 UseDevices(0..value-1), 6 multi_gpu_min_rays;
// returns 10: full uploads, 11: sparse uploads, 12: devices in use
long refh_shim_option(void* h, int key, long value)
{
    auto* H = static_cast<Harness*>(h);
    (void)H;
    (void)key;
    (void)value;
#ifdef B2COL_DROPIN
    auto* cs = H->scene.GetSystem<CollisionSystem>();
    if (key == 0) cs->collider_check_interval = static_cast<unsigned>(value);
    if (key == 1) cs->rebuild_interval = static_cast<unsigned>(value);
    if (key == 5) {
        std::vector<int> devices;
        for (long d = 0; d < value; ++d) devices.push_back(static_cast<int>(d));
        cs->UseDevices(devices);
    }
    if (key == 6) cs->multi_gpu_min_rays = static_cast<std::size_t>(value);
    if (key == 12) return cs->devices_in_use();
#ifdef B2COL_DROPIN_TRANSFORM
    auto* ts = H->scene.GetSystem<TransformSystem>();
    if (key == 2) ts->skip_static_rows = value != 0;
    if (key == 3) ts->write_back_world_matrices = value != 0;
    if (key == 4) ts->static_rescan_interval = static_cast<unsigned>(value);
    if (key == 10) return static_cast<long>(ts->full_uploads);
    if (key == 11) return static_cast<long>(ts->sparse_uploads);
#endif
#endif
    return 0;
}

// Full frame: Scene::Update (src/scene.cpp:59-66) = Transform + Collision (refit + BuildNode).
// Returns seconds.
double refh_scene_update(void* h)
{
    auto* H = static_cast<Harness*>(h);
    const auto t0 = std::chrono::steady_clock::now();
    H->scene.Update(0.016f);
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// TransformSystem::onUpdate only (src/system_transform.cpp:12-38). Returns seconds.
double refh_transform_update(void* h)
{
    auto* H = static_cast<Harness*>(h);
    auto* ts = H->scene.GetSystem<TransformSystem>();
    const auto t0 = std::chrono::steady_clock::now();
    ts->onUpdate(0.016f);
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

#ifndef B2COL_DROPIN
// The reference's refit loop (src/system_collisions.cpp:123-132) without the O(N^2) build:
// iterates CollisionSystem::m_entities (ascending entity ID) and calls the reference's own
// transformBox on each collider. Returns the collider count; seconds in *secs.
size_t refh_refit(void* h, uint32_t* out_entity, float* out_aabb6, double* secs)
{
    auto* H = static_cast<Harness*>(h);
    auto* cs = H->scene.GetSystem<CollisionSystem>();
    const auto t0 = std::chrono::steady_clock::now();
    size_t i = 0;
    for (Entity entity : cs->m_entities) {
        const auto t = H->scene.GetComponent<TransformComponent>(entity);
        const auto c = H->scene.GetComponent<ColliderComponent>(entity);
        const AABB box = transformBox(c->aabb, t->world_matrix);
        if (out_entity) out_entity[i] = entity;
        if (out_aabb6) std::memcpy(out_aabb6 + 6 * i, &box, sizeof(AABB));
        ++i;
    }
    if (secs) *secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return i;
}

#endif  // !B2COL_DROPIN

// world_matrix as 16 floats, column-major (glm memory order)
void refh_get_world_matrices(void* h, size_t n, const uint32_t* entity, float* out16)
{
    auto* H = static_cast<Harness*>(h);
    for (size_t i = 0; i < n; ++i) std::memcpy(out16 + 16 * i, &H->scene.GetTransform(entity[i])->world_matrix, 64);
}

size_t refh_bvh_size(void* h)
{
#ifdef B2COL_DROPIN
    static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->SyncDebugBvh();
#endif
    return static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->bvh_.size();
}

void refh_bvh_get(void* h, void* out60)
{
#ifdef B2COL_DROPIN
    static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->SyncDebugBvh();
#endif
    auto& v = static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->bvh_;
    if (!v.empty()) std::memcpy(out60, v.data(), v.size() * sizeof(CollisionSystem::BiTreeNode));
}

// Assign a foreign tree to the public bvh_ member (include/system_collisions.h:58); the root
// must be the last node (src/system_collisions.cpp:168).
void refh_bvh_set(void* h, const void* nodes60, size_t n)
{
    auto& v = static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->bvh_;
    v.resize(n);
    if (n) std::memcpy(v.data(), nodes60, n * sizeof(CollisionSystem::BiTreeNode));
}

// CollisionSystem::GetRaycast (src/system_collisions.cpp:160-199) over a batch. The reference is
// single-threaded; nthreads>1 only shards the loop (GetRaycast reads bvh_ and nothing else).
// Returns seconds.
double refh_raycast(void* h, size_t n, const void* rays24, void* out36, int nthreads)
{
    auto* cs = static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>();
    const Ray* rays = static_cast<const Ray*>(rays24);
    Raycast* out = static_cast<Raycast*>(out36);
    if (nthreads < 1) nthreads = 1;
    const auto t0 = std::chrono::steady_clock::now();
    auto work = [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i) {
            Raycast r = cs->GetRaycast(rays[i]);
            // write field by field so the 3 padding bytes after `hit` are deterministic (zero)
            std::memset(&out[i], 0, sizeof(Raycast));
            out[i].location = r.location;
            out[i].normal = r.normal;
            out[i].hit_entity = r.hit_entity;
            out[i].distance = r.distance;
            out[i].hit = r.hit;
        }
    };
    if (nthreads == 1) {
        work(0, n);
    }
    else {
        std::vector<std::thread> th;
        for (int k = 0; k < nthreads; ++k) th.emplace_back(work, n * k / nthreads, n * (k + 1) / nthreads);
        for (auto& t : th) t.join();
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

#ifndef B2COL_DROPIN
// Direct known-answer access to the file-static helpers.
// RayBoxIntersection (src/system_collisions.cpp:62-100); direction is used as given (NOT normalised).
int refh_ray_box(const float* ray6, const float* aabb6, float T, float* t_out, int* side_out)
{
    Ray r;
    AABB b;
    std::memcpy(&r, ray6, 24);
    std::memcpy(&b, aabb6, 24);
    AABBSide side;
    auto [hit, t] = RayBoxIntersection(r, b, T, side);
    *t_out = t;
    *side_out = static_cast<int>(side);
    return hit ? 1 : 0;
}

// transformBox (src/system_collisions.cpp:14-41); mat16 column-major
void refh_transform_box(const float* aabb6, const float* mat16, float* out6)
{
    AABB b;
    glm::mat4 m;
    std::memcpy(&b, aabb6, 24);
    std::memcpy(&m, mat16, 64);
    const AABB o = transformBox(b, m);
    std::memcpy(out6, &o, 24);
}

// glm::normalize as GetRaycast applies it (src/system_collisions.cpp:162)
void refh_normalize(const float* v3, float* out3)
{
    const glm::vec3 v = glm::normalize(glm::vec3{v3[0], v3[1], v3[2]});
    std::memcpy(out3, &v, 12);
}

#endif  // !B2COL_DROPIN

// batch entry of the drop-in class (additive API); the reference has none
#ifdef B2COL_DROPIN
void refh_raycast_batch(void* h, size_t n, const void* rays24, void* out36)
{
    static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>()->GetRaycasts(static_cast<const Ray*>(rays24), n, static_cast<Raycast*>(out36));
}
// compact results, then the host expansion to Raycasts (both additive)
void refh_raycast_hits_expanded(void* h, size_t n, const void* rays24, void* hits12, void* out36, int threads)
{
    auto* cs = static_cast<Harness*>(h)->scene.GetSystem<CollisionSystem>();
    cs->GetRaycastHits(static_cast<const Ray*>(rays24), n, static_cast<CollisionSystem::RaycastHit*>(hits12));
    CollisionSystem::ExpandHits(static_cast<const Ray*>(rays24), static_cast<const CollisionSystem::RaycastHit*>(hits12), n, static_cast<Raycast*>(out36), threads);
}
#endif

int refh_hw_threads() { return static_cast<int>(std::thread::hardware_concurrency()); }

}  // extern "C"
