/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  ("the oracle port")
 *
 * Plain-C, single-rounding (no FMA: build with -ffp-contract=off, no -march=native)
 * restatement of the arithmetic of the reference collision path, each function citing the
 * reference file:line it follows (paths relative to /root/reference).  It exists so that the
 * CUDA path can be checked on machines where /root/reference is absent (the GPU box) and at
 * sizes where the reference's O(N^2) builder cannot run.  It is pinned against the real
 * reference (oracle/_ref/libref_engine.so, built from the reference's own sources) by
 * tests/test_oracle.py in this container and against tests/golden/ fixtures, which
 * were produced by that real reference (tests/golden/make_golden.py), everywhere else.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.  engine_b200/ never does; the product has no CPU path.
 *
 * Parity status: PINNED to outputs of the reference itself run in this container; the
 * reference ships no tests or golden vectors of its own.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float min[3], max[3]; } orc_aabb;               /* include/component_collider.h:7-10 */
typedef struct { float origin[3], direction[3]; } orc_ray;       /* include/system_collisions.h:13-16 */
typedef struct {                                                  /* include/system_collisions.h:18-24 */
    float location[3], normal[3];
    uint32_t hit_entity;
    float distance;
    uint8_t hit, pad[3];
} orc_raycast;
/* include/system_collisions.h:40-48 — 60 bytes, align 4 */
typedef struct {
    orc_aabb box1, box2;
    uint32_t index1, index2;
    uint8_t type1, type2, pad[2];
} orc_node;
enum { ORC_BV = 0, ORC_ENTITY = 1, ORC_EMPTY = 2 };              /* BiTreeNode::Type */
enum { ORC_LEFT = 0, ORC_RIGHT, ORC_BOTTOM, ORC_TOP, ORC_FRONT, ORC_BACK }; /* AABBSide, :26 */

_Static_assert(sizeof(orc_aabb) == 24, "aabb");
_Static_assert(sizeof(orc_ray) == 24, "ray");
_Static_assert(sizeof(orc_raycast) == 36, "raycast");
_Static_assert(sizeof(orc_node) == 60, "node");

/* ---------------------------------------------------------------------------------------------
 * TransformSystem::onUpdate — src/system_transform.cpp:12-38
 * Matrices are 16 floats, column-major: m[4*c + r] (glm memory order).
 * ------------------------------------------------------------------------------------------- */

/* glm::mat4_cast — vendor/glm/glm/gtc/quaternion.inl:47-78 (via mat3_cast, then mat4(mat3),
 * vendor/glm/glm/detail/type_mat4x4.inl:189-191); then column 3 = position
 * (src/system_transform.cpp:24); then glm::scale — vendor/glm/glm/ext/matrix_transform.inl:78-86:
 * columns 0..2 multiplied by s.x, s.y, s.z, column 3 untouched. q = x,y,z,w. */
void orc_local_matrix(const float* p, const float* q, const float* s, float* m)
{
    const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    const float qxx = qx * qx, qyy = qy * qy, qzz = qz * qz;
    const float qxz = qx * qz, qxy = qx * qy, qyz = qy * qz;
    const float qwx = qw * qx, qwy = qw * qy, qwz = qw * qz;
    float r[16];
    r[0] = 1.0f - 2.0f * (qyy + qzz);
    r[1] = 2.0f * (qxy + qwz);
    r[2] = 2.0f * (qxz - qwy);
    r[3] = 0.0f;
    r[4] = 2.0f * (qxy - qwz);
    r[5] = 1.0f - 2.0f * (qxx + qzz);
    r[6] = 2.0f * (qyz + qwx);
    r[7] = 0.0f;
    r[8] = 2.0f * (qxz + qwy);
    r[9] = 2.0f * (qyz - qwx);
    r[10] = 1.0f - 2.0f * (qxx + qyy);
    r[11] = 0.0f;
    r[12] = p[0];
    r[13] = p[1];
    r[14] = p[2];
    r[15] = 1.0f;
    for (int c = 0; c < 3; ++c)
        for (int k = 0; k < 4; ++k) m[4 * c + k] = r[4 * c + k] * s[c];
    for (int k = 0; k < 4; ++k) m[12 + k] = r[12 + k];
}

/* glm mat4*mat4 — vendor/glm/glm/detail/type_mat4x4.inl:630-658 (the generic, non-SIMD path):
 * column j of the product = ((A0*B[j][0] + A1*B[j][1]) + A2*B[j][2]) + A3*B[j][3]. */
void orc_mat_mul(const float* a, const float* b, float* out)
{
    float t[16];
    for (int j = 0; j < 4; ++j)
        for (int r = 0; r < 4; ++r) {
            float acc = a[r] * b[4 * j];
            acc = acc + a[4 + r] * b[4 * j + 1];
            acc = acc + a[8 + r] * b[4 * j + 2];
            acc = acc + a[12 + r] * b[4 * j + 3];
            t[4 * j + r] = acc;
        }
    memcpy(out, t, sizeof t);
}

/* Whole pass. Arrays are indexed by entity ID; slot 0 is unused (entity 0 is invalid,
 * include/scene.h:141). Entities are visited in ascending ID (std::set, include/ecs.h:56) and a
 * child reads whatever world matrix its parent currently holds (src/system_transform.cpp:28-31). */
void orc_transform_update(size_t n_entities, const uint32_t* parent, const float* pos3, const float* rot_xyzw, const float* scl3,
                          float* world16)
{
    for (size_t e = 1; e <= n_entities; ++e) {
        float l[16];
        orc_local_matrix(pos3 + 3 * e, rot_xyzw + 4 * e, scl3 + 3 * e, l);
        if (parent[e] != 0)
            orc_mat_mul(world16 + 16 * (size_t)parent[e], l, world16 + 16 * e);
        else
            memcpy(world16 + 16 * e, l, sizeof l);
    }
}

/* ---------------------------------------------------------------------------------------------
 * transformBox — src/system_collisions.cpp:14-41
 * mat4*vec4 — vendor/glm/glm/detail/type_mat4x4.inl:540-577 (generic path):
 *   (m[0]*v.x + m[1]*v.y) + (m[2]*v.z + m[3]*v.w), v.w = 1.0f
 * ------------------------------------------------------------------------------------------- */
void orc_transform_box(const orc_aabb* box, const float* m, orc_aabb* out)
{
    orc_aabb nb;
    for (int k = 0; k < 3; ++k) {
        nb.min[k] = INFINITY;
        nb.max[k] = -INFINITY;
    }
    for (int i = 0; i < 8; ++i) { /* corner order of :16-19: x outer, y middle, z inner */
        const float x = (i & 4) ? box->max[0] : box->min[0];
        const float y = (i & 2) ? box->max[1] : box->min[1];
        const float z = (i & 1) ? box->max[2] : box->min[2];
        for (int r = 0; r < 3; ++r) {
            const float mul0 = m[r] * x, mul1 = m[4 + r] * y;
            const float add0 = mul0 + mul1;
            const float mul2 = m[8 + r] * z, mul3 = m[12 + r] * 1.0f;
            const float add1 = mul2 + mul3;
            const float v = add0 + add1;
            nb.min[r] = fminf(nb.min[r], v);
            nb.max[r] = fmaxf(nb.max[r], v);
        }
    }
    *out = nb;
}

/* refit loop — src/system_collisions.cpp:123-132 */
void orc_refit(size_t n, const uint32_t* entity, const orc_aabb* local, const float* world16, orc_aabb* out)
{
    for (size_t i = 0; i < n; ++i) orc_transform_box(&local[i], world16 + 16 * (size_t)entity[i], &out[i]);
}

/* ---------------------------------------------------------------------------------------------
 * RayBoxIntersection — src/system_collisions.cpp:62-100.  Returns hit; *t_out = tmin.
 * ------------------------------------------------------------------------------------------- */
int orc_ray_box(const orc_ray* ray, const orc_aabb* box, float T, float* t_out, int* side_out)
{
    const float nx = 1.0f / ray->direction[0], ny = 1.0f / ray->direction[1], nz = 1.0f / ray->direction[2];
    const float tx1 = (box->min[0] - ray->origin[0]) * nx;
    const float tx2 = (box->max[0] - ray->origin[0]) * nx;
    float tmin = fminf(tx1, tx2);
    float tmax = fmaxf(tx1, tx2);
    const float ty1 = (box->min[1] - ray->origin[1]) * ny;
    const float ty2 = (box->max[1] - ray->origin[1]) * ny;
    tmin = fmaxf(tmin, fminf(ty1, ty2));
    tmax = fminf(tmax, fmaxf(ty1, ty2));
    const float tz1 = (box->min[2] - ray->origin[2]) * nz;
    const float tz2 = (box->max[2] - ray->origin[2]) * nz;
    tmin = fmaxf(tmin, fminf(tz1, tz2));
    tmax = fminf(tmax, fmaxf(tz1, tz2));
    int side = ORC_LEFT;
    if (tmin == tx2) side = ORC_RIGHT;
    else if (tmin == tx1) side = ORC_LEFT;
    else if (tmin == ty2) side = ORC_BACK;
    else if (tmin == ty1) side = ORC_FRONT;
    else if (tmin == tz2) side = ORC_TOP;
    else if (tmin == tz1) side = ORC_BOTTOM;
    *side_out = side;
    *t_out = tmin;
    return (tmax >= fmaxf(0.0f, tmin) && tmin < T) ? 1 : 0;
}

/* glm::normalize — vendor/glm/glm/detail/func_geometric.inl:84-90: v * inversesqrt(dot(v,v)),
 * inversesqrt = 1/sqrt (func_exponential.inl:138), dot = x*x + y*y + z*z left to right (:53). */
void orc_normalize(const float* v, float* out)
{
    const float d = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
    const float inv = 1.0f / sqrtf(d);
    out[0] = v[0] * inv;
    out[1] = v[1] * inv;
    out[2] = v[2] * inv;
}

/* ---------------------------------------------------------------------------------------------
 * CollisionSystem::RaycastTreeNode — src/system_collisions.cpp:379-540 (recursive, no pruning)
 * ------------------------------------------------------------------------------------------- */
typedef struct { float t; uint32_t object; int side; int hit; } orc_tres;

static orc_tres tree_node(const orc_node* nodes, const orc_ray* ray, const orc_node* nd)
{
    const float T = 1000.0f; /* :384 */
    orc_tres res;
    memset(&res, 0, sizeof res);
    int hit1 = 0, hit2 = 0, side1 = 0, side2 = 0;
    float t1 = INFINITY, t2 = INFINITY;
    if (nd->type1 != ORC_EMPTY) hit1 = orc_ray_box(ray, &nd->box1, T, &t1, &side1);
    if (nd->type2 != ORC_EMPTY) hit2 = orc_ray_box(ray, &nd->box2, T, &t2, &side2);
#define LEAF(T_, IDX_, SIDE_) do { res.t = (T_); res.object = (IDX_); res.side = (SIDE_); res.hit = 1; return res; } while (0)
    if (!hit1 && !hit2) return res; /* :409 */
    if (hit1 && hit2) {
        if (nd->type1 == ORC_BV && nd->type2 == ORC_ENTITY) { /* :415-426 */
            if (t2 < t1) LEAF(t2, nd->index2, side2);
            return tree_node(nodes, ray, &nodes[nd->index1]);
        }
        if (nd->type1 == ORC_ENTITY && nd->type2 == ORC_BV) { /* :429-440 */
            if (t1 < t2) LEAF(t1, nd->index1, side1);
            return tree_node(nodes, ray, &nodes[nd->index2]);
        }
        if (nd->type1 == ORC_BV && nd->type2 == ORC_BV) { /* :443-484 */
            const orc_tres a = tree_node(nodes, ray, &nodes[nd->index1]);
            const orc_tres b = tree_node(nodes, ray, &nodes[nd->index2]);
            if (a.hit && b.hit) return (a.t < b.t) ? a : b; /* :447 strict <, else second */
            if (a.hit) return a;
            if (b.hit) return b;
            return res;
        }
        /* Entity + Entity :487-504 */
        if (t1 < t2) LEAF(t1, nd->index1, side1);
        LEAF(t2, nd->index2, side2);
    }
    if (hit1) { /* :506-521 */
        if (nd->type1 == ORC_BV) return tree_node(nodes, ray, &nodes[nd->index1]);
        LEAF(t1, nd->index1, side1);
    }
    /* hit2 :522-537 */
    if (nd->type2 == ORC_BV) return tree_node(nodes, ray, &nodes[nd->index2]);
    LEAF(t2, nd->index2, side2);
#undef LEAF
}

static void side_normal(int side, float* n)
{ /* src/system_collisions.cpp:175-194 */
    n[0] = n[1] = n[2] = 0.0f;
    switch (side) {
        case ORC_LEFT: n[0] = -1.0f; break;
        case ORC_RIGHT: n[0] = 1.0f; break;
        case ORC_BOTTOM: n[2] = -1.0f; break;
        case ORC_TOP: n[2] = 1.0f; break;
        case ORC_FRONT: n[1] = -1.0f; break;
        case ORC_BACK: n[1] = 1.0f; break;
    }
}

static void fill_result(const orc_ray* nray, const orc_tres* r, orc_raycast* out)
{
    memset(out, 0, sizeof *out); /* value-initialised Raycast, :164 */
    if (!r->hit) return;
    out->hit = 1;
    out->distance = r->t;
    for (int k = 0; k < 3; ++k) out->location[k] = (nray->direction[k] * r->t) + nray->origin[k]; /* :489 */
    out->hit_entity = r->object;
    side_normal(r->side, out->normal);
}

/* CollisionSystem::GetRaycast — src/system_collisions.cpp:160-199. Root is the LAST node (:168). */
void orc_get_raycast(const orc_node* nodes, size_t n_nodes, const orc_ray* ray, orc_raycast* out)
{
    orc_ray r = *ray;
    orc_normalize(ray->direction, r.direction);
    orc_tres res;
    memset(&res, 0, sizeof res);
    if (n_nodes) res = tree_node(nodes, &r, &nodes[n_nodes - 1]);
    fill_result(&r, &res, out);
}

/* Brute force: argmin tmin over every box passing the reference slab test; on equal tmin the
 * LAST box in array order wins (== "last leaf in DFS order" when boxes are given in the DFS leaf
 * order of a tree whose nodes only ever combine with strict '<', SURVEY.md section 0 / probe P4B).
 * ties_out[i] (optional) = number of OTHER boxes sharing the winning tmin. */
void orc_brute_raycast(const uint32_t* entity, const orc_aabb* boxes, size_t n_boxes, const orc_ray* ray, orc_raycast* out,
                       uint32_t* ties_out)
{
    orc_ray r = *ray;
    orc_normalize(ray->direction, r.direction);
    orc_tres best;
    memset(&best, 0, sizeof best);
    uint32_t ties = 0;
    for (size_t i = 0; i < n_boxes; ++i) {
        float t;
        int side;
        if (!orc_ray_box(&r, &boxes[i], 1000.0f, &t, &side)) continue;
        if (!best.hit || !(best.t < t)) {
            if (best.hit && best.t == t) ++ties;
            else ties = 0;
            best.hit = 1;
            best.t = t;
            best.side = side;
            best.object = entity[i];
        }
    }
    fill_result(&r, &best, out);
    if (ties_out) *ties_out = ties;
}

/* ---- batch drivers (pthread sharding of the ray loop only; each ray is independent) ---- */
typedef struct {
    const orc_node* nodes; size_t n_nodes;
    const uint32_t* entity; const orc_aabb* boxes; size_t n_boxes;
    const orc_ray* rays; orc_raycast* out; uint32_t* ties; size_t lo, hi;
} job_t;

static void* job_tree(void* p)
{
    job_t* j = (job_t*)p;
    for (size_t i = j->lo; i < j->hi; ++i) orc_get_raycast(j->nodes, j->n_nodes, &j->rays[i], &j->out[i]);
    return NULL;
}
static void* job_brute(void* p)
{
    job_t* j = (job_t*)p;
    for (size_t i = j->lo; i < j->hi; ++i)
        orc_brute_raycast(j->entity, j->boxes, j->n_boxes, &j->rays[i], &j->out[i], j->ties ? &j->ties[i] : NULL);
    return NULL;
}
static void run_jobs(job_t base, size_t n, int nthreads, void* (*fn)(void*))
{
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    job_t jobs[256];
    for (int k = 0; k < nthreads; ++k) {
        jobs[k] = base;
        jobs[k].lo = n * (size_t)k / (size_t)nthreads;
        jobs[k].hi = n * (size_t)(k + 1) / (size_t)nthreads;
        pthread_create(&th[k], NULL, fn, &jobs[k]);
    }
    for (int k = 0; k < nthreads; ++k) pthread_join(th[k], NULL);
}
void orc_raycast_batch(const orc_node* nodes, size_t n_nodes, const orc_ray* rays, size_t n, orc_raycast* out, int nthreads)
{
    job_t b;
    memset(&b, 0, sizeof b);
    b.nodes = nodes; b.n_nodes = n_nodes; b.rays = rays; b.out = out;
    run_jobs(b, n, nthreads, job_tree);
}
void orc_brute_batch(const uint32_t* entity, const orc_aabb* boxes, size_t n_boxes, const orc_ray* rays, size_t n, orc_raycast* out,
                     uint32_t* ties, int nthreads)
{
    job_t b;
    memset(&b, 0, sizeof b);
    b.entity = entity; b.boxes = boxes; b.n_boxes = n_boxes; b.rays = rays; b.out = out; b.ties = ties;
    run_jobs(b, n, nthreads, job_brute);
}

/* ---------------------------------------------------------------------------------------------
 * Tree builder for sizes where the reference's BuildNode (src/system_collisions.cpp:205-376,
 * O(N^2) per level) cannot run.  NOT a restatement of BuildNode's split search: it is a binned
 * SAH builder that minimises the same cost the reference minimises (area1*n1 + area2*n2, :327)
 * over 32 centroid bins per axis, and emits nodes in the reference's shape and order:
 *   >2 prims -> {BoundingVolume, BoundingVolume} with the two child boxes (:343-352)
 *    2 prims -> {Entity, Entity} (:355-362);  1 prim -> {Entity, Empty} (:363-369)
 *   post-order push_back, root last (:372-375).
 * Hit flag / distance / location / normal from the reference's traversal are independent of tree
 * topology; hit_entity differs from the reference-built tree only on exact tmin ties
 * (SURVEY.md section 0).  Used (a) to give the reference's GetRaycast a tree at >=100k colliders
 * for the CPU baseline, (b) in tests as a second, foreign topology.
 * ------------------------------------------------------------------------------------------- */
typedef struct { uint32_t entity; orc_aabb box; float c[3]; } prim_t;

static float box_area(const orc_aabb* b)
{ /* GetBoxArea :52-58 (evaluated in float here; only steers topology) */
    const float dx = fabsf(b->max[0] - b->min[0]), dy = fabsf(b->max[1] - b->min[1]), dz = fabsf(b->max[2] - b->min[2]);
    return dx * dy * 2.0f + dz * dy * 2.0f + dx * dz * 2.0f;
}
static void box_empty(orc_aabb* b)
{
    for (int k = 0; k < 3; ++k) { b->min[k] = INFINITY; b->max[k] = -INFINITY; }
}
static void box_grow(orc_aabb* b, const orc_aabb* o)
{
    for (int k = 0; k < 3; ++k) {
        if (o->min[k] < b->min[k]) b->min[k] = o->min[k];
        if (o->max[k] > b->max[k]) b->max[k] = o->max[k];
    }
}

#define NBINS 32
static uint32_t build_rec(prim_t* p, size_t n, orc_node* nodes, size_t* n_nodes)
{
    orc_node nd;
    memset(&nd, 0, sizeof nd);
    if (n > 2) {
        float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (size_t i = 0; i < n; ++i)
            for (int k = 0; k < 3; ++k) {
                if (p[i].c[k] < cmin[k]) cmin[k] = p[i].c[k];
                if (p[i].c[k] > cmax[k]) cmax[k] = p[i].c[k];
            }
        float best = INFINITY;
        int best_axis = -1, best_bin = 0;
        for (int axis = 0; axis < 3; ++axis) {
            const float ext = cmax[axis] - cmin[axis];
            if (!(ext > 0.0f)) continue;
            orc_aabb bb[NBINS];
            size_t cnt[NBINS];
            for (int b = 0; b < NBINS; ++b) { box_empty(&bb[b]); cnt[b] = 0; }
            const float scale = (float)NBINS / ext;
            for (size_t i = 0; i < n; ++i) {
                int b = (int)((p[i].c[axis] - cmin[axis]) * scale);
                if (b < 0) b = 0;
                if (b >= NBINS) b = NBINS - 1;
                box_grow(&bb[b], &p[i].box);
                cnt[b]++;
            }
            orc_aabb left[NBINS], right[NBINS];
            size_t lc[NBINS], rc[NBINS];
            orc_aabb acc; size_t c = 0;
            box_empty(&acc);
            for (int b = 0; b < NBINS; ++b) { box_grow(&acc, &bb[b]); c += cnt[b]; left[b] = acc; lc[b] = c; }
            box_empty(&acc); c = 0;
            for (int b = NBINS - 1; b >= 0; --b) { box_grow(&acc, &bb[b]); c += cnt[b]; right[b] = acc; rc[b] = c; }
            for (int b = 0; b < NBINS - 1; ++b) {
                if (lc[b] == 0 || rc[b + 1] == 0) continue;
                const float cost = box_area(&left[b]) * (float)lc[b] + box_area(&right[b + 1]) * (float)rc[b + 1];
                if (cost < best) { best = cost; best_axis = axis; best_bin = b; }
            }
        }
        size_t mid;
        if (best_axis >= 0) {
            const float ext = cmax[best_axis] - cmin[best_axis];
            const float scale = (float)NBINS / ext;
            size_t i = 0, j = n;
            while (i < j) {
                int b = (int)((p[i].c[best_axis] - cmin[best_axis]) * scale);
                if (b < 0) b = 0;
                if (b >= NBINS) b = NBINS - 1;
                if (b <= best_bin) ++i;
                else { prim_t tmp = p[i]; p[i] = p[--j]; p[j] = tmp; }
            }
            mid = i;
        }
        else {
            mid = n / 2; /* all centroids coincide: split the array in half */
        }
        box_empty(&nd.box1);
        box_empty(&nd.box2);
        for (size_t i = 0; i < mid; ++i) box_grow(&nd.box1, &p[i].box);
        for (size_t i = mid; i < n; ++i) box_grow(&nd.box2, &p[i].box);
        nd.index1 = build_rec(p, mid, nodes, n_nodes);
        nd.index2 = build_rec(p + mid, n - mid, nodes, n_nodes);
        nd.type1 = ORC_BV;
        nd.type2 = ORC_BV;
    }
    else if (n == 2) {
        nd.box1 = p[0].box; nd.box2 = p[1].box;
        nd.index1 = p[0].entity; nd.index2 = p[1].entity;
        nd.type1 = ORC_ENTITY; nd.type2 = ORC_ENTITY;
    }
    else {
        nd.box1 = p[0].box; nd.index1 = p[0].entity;
        nd.type1 = ORC_ENTITY; nd.type2 = ORC_EMPTY;
    }
    nodes[*n_nodes] = nd;
    return (uint32_t)((*n_nodes)++);
}

/* out_nodes must hold 2n-1 nodes. Returns the node count (0 for n == 0). */
size_t orc_build_binned_sah(const uint32_t* entity, const orc_aabb* boxes, size_t n, orc_node* out_nodes)
{
    if (n == 0) return 0;
    prim_t* p = (prim_t*)malloc(n * sizeof(prim_t));
    for (size_t i = 0; i < n; ++i) {
        p[i].entity = entity[i];
        p[i].box = boxes[i];
        for (int k = 0; k < 3; ++k) p[i].c[k] = (boxes[i].min[k] + boxes[i].max[k]) * 0.5f; /* GetBoxCentroid :43-50 */
    }
    size_t nn = 0;
    build_rec(p, n, out_nodes, &nn);
    free(p);
    return nn;
}

/* Leaves of a reference-shaped tree in DFS order (slot 1 before slot 2, root = last node):
 * the order in which orc_brute_raycast's tie rule reproduces the tree's tie winners.
 * Returns the number of leaves written. */
static void dfs_rec(const orc_node* nodes, uint32_t idx, uint32_t* ent, orc_aabb* box, size_t* k)
{
    const orc_node* nd = &nodes[idx];
    if (nd->type1 == ORC_BV) dfs_rec(nodes, nd->index1, ent, box, k);
    else if (nd->type1 == ORC_ENTITY) { ent[*k] = nd->index1; box[*k] = nd->box1; ++*k; }
    if (nd->type2 == ORC_BV) dfs_rec(nodes, nd->index2, ent, box, k);
    else if (nd->type2 == ORC_ENTITY) { ent[*k] = nd->index2; box[*k] = nd->box2; ++*k; }
}
size_t orc_tree_leaves_dfs(const orc_node* nodes, size_t n_nodes, uint32_t* out_entity, orc_aabb* out_boxes)
{
    size_t k = 0;
    if (n_nodes) dfs_rec(nodes, (uint32_t)(n_nodes - 1), out_entity, out_boxes, &k);
    return k;
}
