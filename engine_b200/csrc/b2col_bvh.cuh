// b2col_bvh.cuh — per-thread bodies of the hierarchy, export and traversal kernels, written as
// host/device inline functions over an `Ops` policy (atomics + cache-bypassing loads) so that the
// same logic runs inside the CUDA kernels (b2col.cu) and, one "thread" at a time, inside the
// CPU-side logic tests (tests/emu/emu.cpp — test infrastructure, never part of the product).
#pragma once
#include <stddef.h>

#include "b2col_math.cuh"

namespace b2 {

struct Aabb24 {  // engine::AABB, 24 bytes (reference include/component_collider.h:7-10)
    float lo[3], hi[3];
};

constexpr uint32_t UNVISITED = 0xFFFFFFFFu;

// ---- LBVH hierarchy + bounds in ONE bottom-up pass ---------------------------------------------
// Replaces BuildNode (reference src/system_collisions.cpp:205-376) by a different algorithm: the
// tree of Karras 2012 (radix tree over sorted Morton keys, ties broken by position), emitted
// bottom-up in the manner of Apetrei 2014: the thread of leaf i climbs; at each step its current
// range [l,r] joins the neighbour with which it shares the longer key prefix, i.e. the internal
// node p = r (as left child) or p = l-1 (as right child).  Internal node p is the split between
// sorted positions p and p+1.  The first thread to reach a node parks its box in the node's child
// slot and retires; the second merges and climbs on.  Node boxes are exact min/max of the leaf
// boxes (bit copies of the refit output), so containment is exact.
//
// delta(i) orders the split between sorted keys i and i+1: larger = splits higher in the tree.
struct Delta {
    uint64_t hi;
    uint32_t lo;
};
B2_HD Delta key_delta(const uint64_t* keys, uint32_t i)
{
    Delta d;
    d.hi = keys[i] ^ keys[i + 1];
    d.lo = d.hi ? 0u : (i ^ (i + 1u));  // duplicate keys: fall back to the position bits
    return d;
}
B2_HD bool delta_less(const Delta& a, const Delta& b) { return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo); }

template <class Ops>
B2_HD void build_from_leaf(uint32_t i, uint32_t n, const uint64_t* keys, const uint32_t* sorted_prim, const Aabb24* world_aabb,
                           const uint32_t* col_entity, Node* nodes, uint32_t* other_bound, uint32_t* leaf_entity, uint32_t* root_out)
{
    const uint32_t prim = sorted_prim[i];
    leaf_entity[i] = col_entity[prim];
    Box box;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        box.lo[k] = world_aabb[prim].lo[k];
        box.hi[k] = world_aabb[prim].hi[k];
    }
    if (n == 1) {  // a lone leaf still gets a node: {leaf, nothing} (cf. the reference's {Entity,Empty}, :363-369)
        Slot a, b;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            a.lo[k] = box.lo[k];
            a.hi[k] = box.hi[k];
            b.lo[k] = B2_NAN;
            b.hi[k] = B2_NAN;
        }
        a.ref = LEAF_BIT | 0u;
        a.aux = 0u;
        b.ref = REF_EMPTY;
        b.aux = 0u;
        Ops::store_slot(&nodes[0].s[0], a);
        Ops::store_slot(&nodes[0].s[1], b);
        *root_out = 0u;
        return;
    }
    uint32_t l = i, r = i;
    uint32_t ref = LEAF_BIT | i;
    for (;;) {
        bool is_left;
        if (l == 0u) is_left = true;
        else if (r == n - 1u) is_left = false;
        else is_left = delta_less(key_delta(keys, r), key_delta(keys, l - 1u));
        const uint32_t p = is_left ? r : l - 1u;
        Slot mine;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            mine.lo[k] = box.lo[k];
            mine.hi[k] = box.hi[k];
        }
        mine.ref = ref;
        mine.aux = is_left ? l : r;
        Ops::store_slot(&nodes[p].s[is_left ? 0 : 1], mine);
        Ops::fence();
        const uint32_t other = Ops::exch(&other_bound[p], is_left ? l : r);
        if (other == UNVISITED) return;  // first to arrive: the sibling's thread finishes this node
        Ops::fence();
        const Slot sib = Ops::load_slot_coherent(&nodes[p].s[is_left ? 1 : 0]);
        if (is_left) r = other;
        else l = other;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            box.lo[k] = fminf(box.lo[k], sib.lo[k]);
            box.hi[k] = fmaxf(box.hi[k], sib.hi[k]);
        }
        ref = p;
        if (l == 0u && r == n - 1u) {
            *root_out = p;
            return;
        }
    }
}

// ---- export to the reference's BiTreeNode format -----------------------------------------------
// Layout written (2n-1 candidate nodes, compacted afterwards by the caller):
//   [0, n)        wrapper {Entity, Empty} for leaf position i (used only when the leaf's sibling is
//                 an internal node: the reference drops the entity of a mixed BV+Entity node,
//                 src/system_collisions.cpp:415-440, so mixed nodes are never emitted)
//   [n, 2n-1)     internal node p at n + perm(p), perm swaps the root with the last index so that the
//                 root is bvh_.back() (:168)
// used[k] = 1 for every node that is referenced (or is the root).
struct BiTreeNode60 {
    float box1[6], box2[6];
    uint32_t index1, index2;
    uint8_t type1, type2, pad0, pad1;
};
static_assert(sizeof(BiTreeNode60) == 60, "BiTreeNode layout");

B2_HD uint32_t export_perm(uint32_t p, uint32_t root, uint32_t last) { return p == root ? last : (p == last ? root : p); }

B2_HD void export_internal(uint32_t p, uint32_t n, uint32_t root, const Node* nodes, const uint32_t* leaf_entity, BiTreeNode60* out,
                           uint8_t* used)
{
    const Node nd = nodes[p];
    const uint32_t last = n - 2u;
    BiTreeNode60 o;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        o.box1[k] = nd.s[0].lo[k];
        o.box1[3 + k] = nd.s[0].hi[k];
        o.box2[k] = nd.s[1].lo[k];
        o.box2[3 + k] = nd.s[1].hi[k];
    }
    o.pad0 = 0;
    o.pad1 = 0;
    const bool leaf0 = (nd.s[0].ref & LEAF_BIT) != 0u, leaf1 = (nd.s[1].ref & LEAF_BIT) != 0u;
    if (leaf0 && leaf1) {  // {Entity, Entity}, reference :355-362
        o.type1 = 1;
        o.type2 = 1;
        o.index1 = leaf_entity[nd.s[0].ref & ~LEAF_BIT];
        o.index2 = leaf_entity[nd.s[1].ref & ~LEAF_BIT];
    }
    else {  // {BoundingVolume, BoundingVolume}, reference :349-352; leaf children go through a wrapper
        o.type1 = 0;
        o.type2 = 0;
        if (leaf0) {
            o.index1 = nd.s[0].ref & ~LEAF_BIT;
            used[o.index1] = 1;
        }
        else {
            o.index1 = n + export_perm(nd.s[0].ref, root, last);
        }
        if (leaf1) {
            o.index2 = nd.s[1].ref & ~LEAF_BIT;
            used[o.index2] = 1;
        }
        else {
            o.index2 = n + export_perm(nd.s[1].ref, root, last);
        }
    }
    const uint32_t at = n + export_perm(p, root, last);
    out[at] = o;
    used[at] = 1;
}

B2_HD void export_wrapper(uint32_t i, const Aabb24* world_aabb, const uint32_t* sorted_prim, const uint32_t* leaf_entity, BiTreeNode60* out)
{
    BiTreeNode60 o;
    const Aabb24 b = world_aabb[sorted_prim[i]];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        o.box1[k] = b.lo[k];
        o.box1[3 + k] = b.hi[k];
        o.box2[k] = 0.0f;  // value-initialised, as BiTreeNode node{} at reference :223
        o.box2[3 + k] = 0.0f;
    }
    o.index1 = leaf_entity[i];
    o.index2 = 0u;
    o.type1 = 1;  // Entity
    o.type2 = 2;  // Empty
    o.pad0 = 0;
    o.pad1 = 0;
    out[i] = o;
}

// Host-side tail of the export: drop the wrapper nodes nobody refers to, keeping the order (wrappers,
// then internal nodes, root last) and renumbering child indices.  Returns the number of nodes kept;
// writes nothing when `out` is null or `capacity` is too small.
inline size_t compact_export(const BiTreeNode60* cand, const uint8_t* used, size_t n_cand, BiTreeNode60* out, size_t capacity)
{
    size_t kept = 0;
    for (size_t k = 0; k < n_cand; ++k) kept += used[k] ? 1 : 0;
    if (!out || kept > capacity) return kept;
    uint32_t* remap = new uint32_t[n_cand];
    size_t at = 0;
    for (size_t k = 0; k < n_cand; ++k) {
        remap[k] = (uint32_t)at;
        at += used[k] ? 1 : 0;
    }
    for (size_t k = 0; k < n_cand; ++k) {
        if (!used[k]) continue;
        BiTreeNode60 nd = cand[k];
        if (nd.type1 == 0) nd.index1 = remap[nd.index1];
        if (nd.type2 == 0) nd.index2 = remap[nd.index2];
        out[remap[k]] = nd;
    }
    delete[] remap;
    return kept;
}

// ---- production traversal (LBVH) ----------------------------------------------------------------
// Result contract: the hit leaf with the least tmin among all leaves
// that pass the reference's slab test; on equal tmin the leaf that comes LAST in depth-first order
// wins, which for this tree is the larger Morton position — exactly what the reference's own
// RaycastTreeNode (strict '<' at :447 and :488, else second child) returns when it walks the
// exported tree.  Internal boxes only prune: a child is skipped when its box misses (same slab test,
// monotone in the box, so a hit leaf implies hit ancestors) or when its entry tmin is already
// greater than the best tmin (strictly: an equal tmin may still hide a later-position tie winner).
constexpr int LBVH_STACK = 96;  // depth <= key bits (63) + position bits (<= 31): cannot overflow

struct Hit {
    float t;
    int32_t leaf;  // Morton position of the winning leaf, -1 = none
    int side;
};

// Traversal state of one ray, advanced one node visit at a time so that a persistent warp can
// retire finished rays and fetch new ones between steps (k_trace_lbvh in b2col.cu).  The stack of
// deferred far children (node, entry tmin) is a policy: a plain array on the host, shared memory
// with a local-memory overflow on the device.
struct TraceState {
    uint32_t cur;
    int sp;
    Hit best;
};

struct ArrayStack {
    uint32_t node[LBVH_STACK];
    float t[LBVH_STACK];
    B2_HD void put(int i, uint32_t n, float tt)
    {
        node[i] = n;
        t[i] = tt;
    }
    B2_HD float get_t(int i) const { return t[i]; }
    B2_HD uint32_t get_node(int i) const { return node[i]; }
};

B2_HD void trace_begin(TraceState& st, uint32_t root)
{
    st.cur = root;
    st.sp = 0;
    st.best.t = B2_INF;
    st.best.leaf = -1;
    st.best.side = SIDE_LEFT;
}

B2_HD void leaf_candidate(Hit& best, const SlabT& s, uint32_t ref)
{
    const int32_t leaf = (int32_t)(ref & ~LEAF_BIT);
    if (s.tmin < best.t || (s.tmin == best.t && leaf > best.leaf)) {
        best.t = s.tmin;
        best.leaf = leaf;
        best.side = slab_side(s);
    }
}

// Visit node st.cur (its two slots are passed in), then select the next node. Returns false when the
// traversal of this ray is complete.
template <class Stack>
B2_HD bool trace_visit(TraceState& st, Stack& stk, const RayPre& r, const Slot& a, const Slot& b, float T)
{
    const SlabT sa = slab(r, a.lo, a.hi);
    const SlabT sb = slab(r, b.lo, b.hi);
    bool ha = slab_hit(sa, T);
    bool hb = slab_hit(sb, T);
    if (ha && (a.ref & LEAF_BIT)) {
        leaf_candidate(st.best, sa, a.ref);
        ha = false;
    }
    if (hb && (b.ref & LEAF_BIT)) {
        leaf_candidate(st.best, sb, b.ref);
        hb = false;
    }
    ha = ha && (sa.tmin <= st.best.t);
    hb = hb && (sb.tmin <= st.best.t);
    if (ha && hb) {
        const bool a_first = !(sb.tmin < sa.tmin);
        stk.put(st.sp, a_first ? b.ref : a.ref, a_first ? sb.tmin : sa.tmin);
        ++st.sp;
        st.cur = a_first ? a.ref : b.ref;
        return true;
    }
    if (ha || hb) {
        st.cur = ha ? a.ref : b.ref;
        return true;
    }
    while (st.sp > 0) {
        --st.sp;
        if (stk.get_t(st.sp) <= st.best.t) {
            st.cur = stk.get_node(st.sp);
            return true;
        }
    }
    return false;
}

template <class Ops>
B2_HD Hit trace_lbvh(const RayPre& r, const Node* nodes, uint32_t root, float T)
{
    TraceState st;
    ArrayStack stk;
    trace_begin(st, root);
    for (;;) {
        const Slot a = Ops::load_slot(&nodes[st.cur].s[0]);
        const Slot b = Ops::load_slot(&nodes[st.cur].s[1]);
        if (!trace_visit(st, stk, r, a, b, T)) break;
    }
    return st.best;
}

// ---- reference-topology traversal ---------------------------------------------------------------
// Walks an uploaded reference-format tree with the reference's exact semantics
// (RaycastTreeNode, src/system_collisions.cpp:379-540), iteratively: children are visited slot 1
// before slot 2 and a later candidate replaces the best when best.t < cand.t is false — the
// sequential form of ":447 node1.t < node2.t ? node1 : node2" and ":488 t1 < t2 ? 1 : 2".  The mixed
// BoundingVolume+Entity cases (:415-440) — dead code for trees the reference builds — are honoured
// as written, including that the entity is dropped when the volume is entered.  A volume whose entry
// tmin exceeds the best tmin is skipped (its leaves could never replace the best).
constexpr int REFTOPO_STACK = 192;

struct RefHit {
    float t;
    uint32_t entity;
    int side;
    bool hit;
    bool overflow;
};

B2_HD RefHit trace_reftopo(const RayPre& r, const uint32_t* words, uint32_t n_nodes, float T)
{
    RefHit best;
    best.t = B2_INF;
    best.entity = 0u;
    best.side = SIDE_LEFT;
    best.hit = false;
    best.overflow = false;
    if (n_nodes == 0u) return best;
    uint32_t stack[REFTOPO_STACK];
    int sp = 0;
    uint32_t cur = n_nodes - 1u;  // root = bvh_.back(), :168
    for (;;) {
        const uint32_t* w = words + (size_t)cur * 15u;
        float b1[6], b2[6];
        union {
            uint32_t u;
            float f;
        } cv;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            cv.u = w[k];
            b1[k] = cv.f;
            cv.u = w[6 + k];
            b2[k] = cv.f;
        }
        const uint32_t idx1 = w[12], idx2 = w[13];
        const uint32_t ty1 = w[14] & 0xffu, ty2 = (w[14] >> 8) & 0xffu;
        SlabT s1 = slab(r, b1, b1 + 3), s2 = slab(r, b2, b2 + 3);
        const bool h1 = (ty1 != 2u) && slab_hit(s1, T);
        const bool h2 = (ty2 != 2u) && slab_hit(s2, T);
        // what to do with each slot: 0 nothing, 1 candidate entity, 2 descend
        int act1 = 0, act2 = 0;
        if (h1 && h2) {
            if (ty1 == 0u && ty2 == 1u) {  // :415-426
                if (s2.tmin < s1.tmin) act2 = 1;
                else act1 = 2;
            }
            else if (ty1 == 1u && ty2 == 0u) {  // :429-440
                if (s1.tmin < s2.tmin) act1 = 1;
                else act2 = 2;
            }
            else {
                act1 = (ty1 == 0u) ? 2 : 1;
                act2 = (ty2 == 0u) ? 2 : 1;
            }
        }
        else if (h1) act1 = (ty1 == 0u) ? 2 : 1;
        else if (h2) act2 = (ty2 == 0u) ? 2 : 1;
        if (act1 == 1 && !(best.hit && best.t < s1.tmin)) {
            best.hit = true;
            best.t = s1.tmin;
            best.entity = idx1;
            best.side = slab_side(s1);
        }
        if (act2 == 1 && !(best.hit && best.t < s2.tmin)) {
            best.hit = true;
            best.t = s2.tmin;
            best.entity = idx2;
            best.side = slab_side(s2);
        }
        if (act1 == 2 && best.hit && s1.tmin > best.t) act1 = 0;
        if (act2 == 2 && best.hit && s2.tmin > best.t) act2 = 0;
        if (act1 == 2 && act2 == 2) {
            if (sp >= REFTOPO_STACK) {
                best.overflow = true;
                return best;
            }
            stack[sp++] = idx2;
            cur = idx1;
            continue;
        }
        if (act1 == 2) {
            cur = idx1;
            continue;
        }
        if (act2 == 2) {
            cur = idx2;
            continue;
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return best;
}

}  // namespace b2
