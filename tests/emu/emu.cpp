// TEST INFRASTRUCTURE — NOT PRODUCT CODE, never loaded by engine_b200.
//
// Runs the per-thread kernel bodies of engine_b200/csrc/b2col_{math,bvh}.cuh on the host, one
// "thread" after another, so that the LOGIC of the GPU path (hierarchy emission, tie order, export,
// both traversals, transform/refit arithmetic) can be checked against the oracle on machines
// without a GPU (pytest -m "not gpu").  It cannot and does not stand in for the CUDA library:
// the radix sort, reductions and everything launch-related only exist in b2col.cu and are only
// tested on the GPU.  Build: g++ -O2 -ffp-contract=off (tests/test_emu_logic.py does it).
#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../engine_b200/csrc/b2col_bvh.cuh"

using namespace b2;

struct HostOps {
    static void store_slot(Slot* p, const Slot& s) { *p = s; }
    static Slot load_slot(const Slot* p) { return *p; }
    static Slot load_slot_coherent(const Slot* p) { return *p; }
    static uint32_t arrive(uint32_t* p, uint32_t v)
    {
        const uint32_t o = *p;
        *p = v;
        return o;
    }
    static void store_qnode(QNode* p, const QNode& n) { *p = n; }
    static void store_ref(QNode* p, int side, uint32_t ref) { p->w[6 + side] = ref; }
    static void store_qboxes(QNode* p, const uint32_t* left, const uint32_t* right)
    {
        for (int k = 0; k < 3; ++k) {
            p->w[k] = left[k];
            p->w[3 + k] = right[k];
        }
    }
    static Xch exchange(Xch* p, const Xch& v)
    {
        const Xch o = *p;
        *p = v;
        return o;
    }
    static void store_leaf(LeafRec* p, const LeafRec& r) { *p = r; }
    static QNode load_qnode(const QNode* p) { return *p; }
    static LeafRec load_leaf(const LeafRec* p) { return *p; }
};

extern "C" {

void emu_propagate(uint32_t n_entities, const uint32_t* parent, const float* pos3, const float* rot4, const float* scl3, float* world16)
{
    for (uint32_t e = 1; e <= n_entities; ++e) {  // ascending ID == level order because parents precede children
        float L[16];
        local_matrix(pos3 + 3 * (size_t)e, rot4 + 4 * (size_t)e, scl3 + 3 * (size_t)e, L);
        if (parent[e]) mat_mul(world16 + 16 * (size_t)parent[e], L, world16 + 16 * (size_t)e);
        else std::memcpy(world16 + 16 * (size_t)e, L, sizeof L);
    }
}

void emu_refit(uint32_t n, const uint32_t* col_entity, const Aabb24* col_local, const float* world16, Aabb24* out)
{
    for (uint32_t i = 0; i < n; ++i) {
        const Box b = transform_box(col_local[i].lo, col_local[i].hi, world16 + 16 * (size_t)col_entity[i]);
        for (int k = 0; k < 3; ++k) {
            out[i].lo[k] = b.lo[k];
            out[i].hi[k] = b.hi[k];
        }
    }
}

// bounds -> Morton keys -> stable sort -> bottom-up hierarchy.  nodes/qnodes: max(n-1,1) entries, leaves: n.
// `order` (optional, n entries): visit the leaf "threads" in this order instead of 0..n-1, to show
// that the result does not depend on thread scheduling.
void emu_build(uint32_t n, const Aabb24* world_aabb, const uint32_t* col_entity, int bits, int drop, float T, Node* nodes, QNode* qnodes, LeafRec* leaves,
               Grid* grid_out, uint32_t* sorted_prim, uint64_t* sorted_keys, uint32_t* root, const uint32_t* order, int* mismatch, uint32_t* has_wild,
               const float* plan_bounds)  // optional, 6 floats: min.xyz, max.xyz the Morton plan is made from instead of this frame's centroid bounds (b2col_update's plan policy)
{
    if (n == 0) return;
    float smin[3] = {INFINITY, INFINITY, INFINITY}, smax[3] = {-INFINITY, -INFINITY, -INFINITY};
    float bmin[3] = {INFINITY, INFINITY, INFINITY}, bmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            const float lo = world_aabb[i].lo[k], hi = world_aabb[i].hi[k];
            const float c = (lo + hi) * 0.5f;
            if (grid_coord_ok(c)) {
                smin[k] = fminf(smin[k], c);
                smax[k] = fmaxf(smax[k], c);
            }
            if (grid_coord_ok(lo)) { bmin[k] = fminf(bmin[k], lo); bmax[k] = fmaxf(bmax[k], lo); }
            if (grid_coord_ok(hi)) { bmin[k] = fminf(bmin[k], hi); bmax[k] = fmaxf(bmax[k], hi); }
        }
    const MortonPlan plan = plan_bounds ? make_morton_plan(plan_bounds, plan_bounds + 3, 3 * bits - drop) : make_morton_plan(smin, smax, 3 * bits - drop);
    const Grid grid = make_grid(bmin, bmax, T);
    *grid_out = grid;
    std::vector<uint64_t> keys(n);
    for (uint32_t i = 0; i < n; ++i) keys[i] = morton_key(world_aabb[i].lo, world_aabb[i].hi, plan);
    std::vector<uint32_t> idx(n);
    std::iota(idx.begin(), idx.end(), 0u);
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    for (uint32_t i = 0; i < n; ++i) {
        sorted_prim[i] = idx[i];
        sorted_keys[i] = keys[idx[i]];
    }
    // exact two-slot nodes (export path) ...
    std::vector<uint32_t> other(n, UNVISITED);
    std::vector<QNode> q_exact(n > 1 ? n - 1 : 1);
    for (uint32_t t = 0; t < n; ++t)
        build_from_leaf<HostOps>(order ? order[t] : t, n, sorted_keys, sorted_prim, world_aabb, col_entity, nodes, other.data(), grid, q_exact.data(),
                                 (LeafRec*)nullptr, root);
    // ... and the traversal structure by the production builder (grid-space merge, visited in reverse order to
    // vary which child arrives first); the two must agree when no leaf is wild (the exact builder does not flag them)
    Xch none;
    none.w[0] = none.w[1] = none.w[2] = 0u;
    none.w[3] = UNVISITED;
    std::vector<Xch> xch(n, none);
    uint32_t root_q = 0;
    for (uint32_t t = 0; t < n; ++t)
        build_from_leaf_q<HostOps>(order ? order[n - 1 - t] : n - 1 - t, n, sorted_keys, sorted_prim, world_aabb, col_entity, xch.data(), grid, qnodes, leaves, &root_q, has_wild);
    *mismatch = (root_q != *root) ? 1 : 0;
    bool any_wild = false;
    for (uint32_t i = 0; i < n && !any_wild; ++i) any_wild = box_is_wild(world_aabb[i].lo, world_aabb[i].hi);
    if (any_wild != (*has_wild != 0u)) *mismatch |= 4;
    if (!any_wild && std::memcmp(q_exact.data(), qnodes, q_exact.size() * sizeof(QNode)) != 0) *mismatch |= 2;
}

void emu_trace_lbvh(const QNode* qnodes, const LeafRec* leaves, const Grid* grid, uint32_t root, uint32_t n_leaves, const float* rays6, size_t n_rays,
                    uint32_t* out9, float T, uint32_t has_wild)
{
    for (size_t i = 0; i < n_rays; ++i) {
        const RayPre r = ray_prepare(rays6 + 6 * i, rays6 + 6 * i + 3);
        if (n_leaves == 0) {
            write_raycast(out9 + 9 * i, r, false, 0.0f, 0, 0u);
            continue;
        }
        const Hit h = has_wild ? trace_lbvh<HostOps, true>(r, *grid, qnodes, leaves, root, T) : trace_lbvh<HostOps, false>(r, *grid, qnodes, leaves, root, T);
        finish_raycast<HostOps>(out9 + 9 * i, r, h, leaves);
    }
}

int emu_trace_reftopo(const uint32_t* words, uint32_t n_nodes, const float* rays6, size_t n_rays, uint32_t* out9, float T)
{
    int overflow = 0;
    for (size_t i = 0; i < n_rays; ++i) {
        const RayPre r = ray_prepare(rays6 + 6 * i, rays6 + 6 * i + 3);
        const RefHit h = trace_reftopo(r, words, n_nodes, T);
        overflow |= h.overflow ? 1 : 0;
        write_raycast(out9 + 9 * i, r, h.hit, h.t, h.side, h.entity);
    }
    return overflow;
}

// returns number of nodes written to out (capacity 2n)
size_t emu_export(uint32_t n, const Node* nodes, uint32_t root, const LeafRec* leaves, BiTreeNode60* out, size_t capacity)
{
    if (n == 0) return 0;
    if (n == 1) {
        std::memset(out, 0, sizeof *out);
        std::memcpy(out[0].box1, &leaves[0], 24);
        out[0].index1 = leaves[0].entity;
        out[0].type1 = 1;
        out[0].type2 = 2;
        return 1;
    }
    const size_t cand = 2 * (size_t)n - 1;
    std::vector<BiTreeNode60> tmp(cand);
    std::vector<uint8_t> used(cand, 0);
    for (uint32_t i = 0; i < n; ++i) {
        export_wrapper(i, leaves, tmp.data());
        if (i + 1 < n) export_internal(i, n, root, nodes, leaves, tmp.data(), used.data());
    }
    return compact_export(tmp.data(), used.data(), cand, out, capacity);
}

// Morton plan + keys of a set of centroids (boxes lo = hi = centroid), for the tests of the bit allocation.
// plan_out: axis a, b, c, shift_a, nc, d2, drop
void emu_morton(uint32_t n, const float* centroids3, const float* cmin, const float* cmax, int key_bits, int* plan_out, uint64_t* keys_out)
{
    const MortonPlan p = make_morton_plan(cmin, cmax, key_bits);
    plan_out[0] = p.axis[0]; plan_out[1] = p.axis[1]; plan_out[2] = p.axis[2];
    plan_out[3] = p.shift_a; plan_out[4] = p.nc; plan_out[5] = p.d2; plan_out[6] = p.drop;
    for (uint32_t i = 0; i < n; ++i) keys_out[i] = morton_key(centroids3 + 3 * (size_t)i, centroids3 + 3 * (size_t)i, p);
}

}  // extern "C"

// Instrumentation (dev): node visits per tree depth, visits in total and exact leaf tests, for a ray sample.
extern "C" void emu_visit_histogram(const QNode* qnodes, const LeafRec* leaves, const Grid* grid, uint32_t root, const float* rays6, size_t n_rays,
                                    float T, uint64_t* hist, int max_depth, uint64_t* total_visits, uint64_t* leaf_tests)
{
    std::vector<uint32_t> sn(256);
    std::vector<float> stt(256);
    std::vector<int> sd(256);
    for (size_t i = 0; i < n_rays; ++i) {
        const RayPre r = ray_prepare(rays6 + 6 * i, rays6 + 6 * i + 3);
        RayQ q;
        if (!rayq_prepare(r, *grid, q)) continue;
        Hit best;
        best.t = INFINITY;
        best.leaf = -1;
        int sp = 0, depth = 0;
        uint32_t cur = root;
        for (;;) {
            ++*total_visits;
            ++hist[depth < max_depth ? depth : max_depth - 1];
            const QNode nd = qnodes[cur];
            float ta, tb;
            const float lim = best.leaf >= 0 ? best.t : float_pred(T);
            bool ha = qslab(q, nd.w[0], nd.w[1], nd.w[2], lim, ta);
            bool hb = qslab(q, nd.w[3], nd.w[4], nd.w[5], lim, tb);
            if (ha && (nd.w[6] & LEAF_BIT)) { ++*leaf_tests; leaf_candidate(best, r, leaves[nd.w[6] & ~LEAF_BIT], nd.w[6], T); ha = false; }
            if (hb && (nd.w[7] & LEAF_BIT)) { ++*leaf_tests; leaf_candidate(best, r, leaves[nd.w[7] & ~LEAF_BIT], nd.w[7], T); hb = false; }
            ha = ha && ta <= best.t;
            hb = hb && tb <= best.t;
            if (ha && hb) {
                const bool af = !(tb < ta);
                sn[sp] = af ? nd.w[7] : nd.w[6]; stt[sp] = af ? tb : ta; sd[sp] = depth + 1; ++sp;
                cur = af ? nd.w[6] : nd.w[7]; ++depth; continue;
            }
            if (ha || hb) { cur = ha ? nd.w[6] : nd.w[7]; ++depth; continue; }
            bool found = false;
            while (sp > 0) { --sp; if (stt[sp] <= best.t) { cur = sn[sp]; depth = sd[sp]; found = true; break; } }
            if (!found) break;
        }
    }
}
