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

// ---- colliders the grid cannot bound ("wild") --------------------------------------------------
// The reference's transformBox / slab test have defined behaviour for NaN and +-inf coordinates (fminf / fmaxf drop NaN,
// src/system_collisions.cpp:21-38,65-83): a world AABB with a NaN plane ignores that plane, one with an infinite plane is
// open on that side.  Such a box has no conservative image on a finite grid, so its child slot — and the slot of every
// ancestor — carries WILD_BIT: the slot is ALWAYS entered, whatever the grid test says, and the exact leaf test decides
// as it does for every other leaf.  Sane scenes have no wild leaf and run the instantiation of the traversal that does
// not look at the bit (hdr->has_wild, decided once per kernel); a scene with some pays ~tree depth extra visits per ray
// and wild leaf.  Coordinates beyond 1e37 count as non-finite so that the grid extent itself stays finite.
constexpr uint32_t WILD_BIT = 0x40000000u;
constexpr uint32_t REF_INDEX_MASK = 0x3FFFFFFFu;  // leaf position / node index of a child ref (n < 2^30)
constexpr float GRID_COORD_MAX = 1.0e37f;
B2_HD bool grid_coord_ok(float x) { return fabsf(x) <= GRID_COORD_MAX; }  // false for NaN, +-inf and |x| > 1e37
B2_HD bool box_is_wild(const float* lo, const float* hi)
{
    return !(grid_coord_ok(lo[0]) && grid_coord_ok(lo[1]) && grid_coord_ok(lo[2]) && grid_coord_ok(hi[0]) && grid_coord_ok(hi[1]) && grid_coord_ok(hi[2]));
}

// ---- traversal structure: 32-byte quantised nodes + 32-byte exact leaf records -------------------
// The production traversal is bound first by L1 wavefronts (one per lane per 32-byte sector: incoherent
// rays never share a line) and then by instruction issue, so a node visit must cost ONE sector and as
// few instructions as possible.  Child boxes are therefore stored on a 15-bit grid spanning the scene's
// box bounds — conservatively: lo rounded down, hi rounded up, plus a 2-cell margin that absorbs every
// float rounding of the grid-space slab test — and only PRUNE.  Exactness lives in the leaf records: a
// leaf slot that passes the grid test is re-tested against the exact world AABB (bit copy of the refit
// output) with the reference's arithmetic.
//
// A cell index q (15 bits) is stored as the 16-bit value 0x8000|q so that ONE byte-permute turns it into
// the float 1 + q*2^-15 in [1,2): bytes (0x00, s.lo, s.hi, 0x3F) = 0x3F800000 | q<<8.  The lo and hi
// planes of one axis share a 32-bit word; the ray picks its near / far plane with a per-ray permute
// selector, so the slab test needs no per-axis min/max.
struct alignas(32) QNode {
    // w[0..2]: slot 0, axis x,y,z: (lo | hi<<16);  w[3..5]: slot 1 likewise;
    // w[6], w[7]: child refs (LEAF_BIT | Morton position, or internal node index)
    uint32_t w[8];
};
struct alignas(32) LeafRec {  // indexed by Morton position
    float lo[3], hi[3];
    uint32_t entity;
    uint32_t prim;  // index into the b2col_set_colliders order
};
static_assert(sizeof(QNode) == 32 && sizeof(LeafRec) == 32, "traversal records are one sector each");

constexpr float GRID_CELLS = 32768.0f;  // 15 bits

struct Grid {
    float gmin[3], cell[3], inv_cell[3];
};

// Grid over [bmin,bmax] (bounds of all finite world-box coordinates).  The cell size is bounded below
// so that (a) a cell is >= 8 ulp of the largest coordinate (quantisation arithmetic stays monotone),
// (b) a cell is >= T*2^-18: the grid-space t of any |t| <= T then carries an error far below one
// cell, and a ray origin more than 2^22 cells away from the grid is more than 16 T away from every
// box, i.e. provably misses (rayq_prepare).
B2_HD Grid make_grid(const float* bmin, const float* bmax, float T)
{
    Grid g;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float lo = bmin[k], hi = bmax[k];
        if (!(lo <= hi)) {  // no finite coordinate on this axis
            lo = 0.0f;
            hi = 0.0f;
        }
        const float ext = hi - lo;
        const float maxabs = fmaxf(fabsf(lo), fabsf(hi));
        float cell = ext / (GRID_CELLS - 9.0f);
        cell = fmaxf(cell, T * 3.814697265625e-06f);       // T * 2^-18
        cell = fmaxf(cell, maxabs * 9.5367431640625e-07f);  // maxabs * 2^-20
        if (!(cell < B2_INF)) cell = 1.0f;
        g.cell[k] = cell;
        g.inv_cell[k] = 1.0f / cell;
        g.gmin[k] = lo - 4.0f * cell;
    }
    return g;
}

B2_HD uint32_t quant_lo(float x, const Grid& g, int k)
{
    float f = floorf((x - g.gmin[k]) * g.inv_cell[k]) - 2.0f;
    f = fminf(fmaxf(f, 0.0f), GRID_CELLS - 1.0f);  // NaN -> 0 (widest)
    return 0x8000u | (uint32_t)f;
}
B2_HD uint32_t quant_hi(float x, const Grid& g, int k)
{
    float f = ceilf((x - g.gmin[k]) * g.inv_cell[k]) + 2.0f;
    f = fmaxf(fminf(f, GRID_CELLS - 1.0f), 0.0f);
    if (!(f == f)) f = GRID_CELLS - 1.0f;
    return 0x8000u | (uint32_t)f;
}

B2_HD QNode quantise_node(const Slot& a, const Slot& b, const Grid& g)
{
    QNode n;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        n.w[k] = quant_lo(a.lo[k], g, k) | (quant_hi(a.hi[k], g, k) << 16);
        n.w[3 + k] = quant_lo(b.lo[k], g, k) | (quant_hi(b.hi[k], g, k) << 16);
    }
    n.w[6] = a.ref;
    n.w[7] = b.ref;
    return n;
}

// One box on the grid: per axis (lo | hi << 16), the layout of a QNode slot.  Quantisation is monotone,
// so min/max commute with it: the grid box of a union of boxes is the per-plane min/max of their grid
// boxes — the hierarchy can therefore be merged entirely in grid space (build_from_leaf_q).
struct QBox {
    uint32_t w[3];
};
B2_HD QBox quantise_box(const float* lo, const float* hi, const Grid& g)
{
    QBox q;
#pragma unroll
    for (int k = 0; k < 3; ++k) q.w[k] = quant_lo(lo[k], g, k) | (quant_hi(hi[k], g, k) << 16);
    return q;
}
B2_HD uint32_t qplane_merge(uint32_t a, uint32_t b)  // min of the low halves, max of the high halves
{
#if defined(__CUDA_ARCH__)
    return (__vminu2(a, b) & 0xFFFFu) | (__vmaxu2(a, b) & 0xFFFF0000u);  // VIMNMX.U16x2
#else
    const uint32_t lo = (a & 0xFFFFu) < (b & 0xFFFFu) ? (a & 0xFFFFu) : (b & 0xFFFFu);
    const uint32_t hi = (a >> 16) > (b >> 16) ? (a >> 16) : (b >> 16);
    return lo | (hi << 16);
#endif
}

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
template <class Key>
B2_HD Delta key_delta(const Key* keys, uint32_t i)
{
    Delta d;
    d.hi = (uint64_t)(keys[i] ^ keys[i + 1]);
    d.lo = d.hi ? 0u : (i ^ (i + 1u));  // duplicate keys: fall back to the position bits
    return d;
}
B2_HD bool delta_less(const Delta& a, const Delta& b) { return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo); }

template <class Ops, class Key, class BoxRec>  // BoxRec: anything with lo[3], hi[3] (Aabb24 on the host, the 32-byte world record on the device)
B2_HD void build_from_leaf(uint32_t i, uint32_t n, const Key* keys, const uint32_t* sorted_prim, const BoxRec* world_aabb,
                           const uint32_t* col_entity, Node* nodes, uint32_t* other_bound, const Grid& grid, QNode* qnodes, LeafRec* leaves,
                           uint32_t* root_out)
{
    const uint32_t prim = sorted_prim[i];
    Box box;
    LeafRec rec;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        box.lo[k] = world_aabb[prim].lo[k];
        box.hi[k] = world_aabb[prim].hi[k];
        rec.lo[k] = box.lo[k];
        rec.hi[k] = box.hi[k];
    }
    rec.entity = col_entity[prim];
    rec.prim = prim;
    if (leaves) Ops::store_leaf(&leaves[i], rec);
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
        if (qnodes) Ops::store_qnode(&qnodes[0], quantise_node(a, a, grid));  // both slots = the leaf: an empty slot has no grid encoding
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
        const uint32_t other = Ops::arrive(&other_bound[p], is_left ? l : r);  // release my slot / acquire the sibling's
        if (other == UNVISITED) return;  // first to arrive: the sibling's thread finishes this node
        const Slot sib = Ops::load_slot_coherent(&nodes[p].s[is_left ? 1 : 0]);
        if (qnodes) Ops::store_qnode(&qnodes[p], is_left ? quantise_node(mine, sib, grid) : quantise_node(sib, mine, grid));
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

// Production form of the same pass: builds ONLY the traversal structure (QNodes + leaf records), merging
// boxes in grid space.  A child's whole message to its sibling fits in 128 bits, so it travels INSIDE the
// arrival atomic (one 128-bit exchange per child per node): the first arriver deposits its message and
// retires, the second receives it as the exchange's return value.  No data is published through ordinary
// memory, hence no fences.  The message holds
//     the child's grid box            6 x 15 bits   (the always-set bit 15 of each plane is reused as payload)
//     the far bound of its range      30 bits
//     whether it holds a wild leaf     1 bit   (WILD_BIT above)
//     the split LEVEL beyond that bound  7 bits
// where level(i) = 64 + msb(key[i] ^ key[i+1]), or msb(i ^ (i+1)) for duplicate keys, 127 = array end, is the
// delta of Karras' tree reduced to what the climb compares: in any state the climb reaches, all splits
// inside [l,r] are lower than both outer splits, and two outer splits with the same top bit would force
// key[l] and key[r] to differ in that bit — so the outer deltas always differ in their top bit and the level
// decides exactly as the full delta does (same tree as build_from_leaf; the tests compare them).  With the
// levels in the message a thread reads keys only once (its own leaf's two neighbours, coalesced); per climbed
// level it issues one exchange, one 4-byte and one 24-byte store, and three packed min/max.
struct alignas(16) Xch {
    uint32_t w[4];  // w[3] == UNVISITED: nobody has arrived yet
};
constexpr uint32_t LEVEL_END = 127u;

B2_HD uint32_t msb32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return 31u - (uint32_t)__clz((int)x);
#else
    return 31u - (uint32_t)__builtin_clz(x);
#endif
}
B2_HD uint32_t msb64(uint64_t x)
{
#if defined(__CUDA_ARCH__)
    return 63u - (uint32_t)__clzll((long long)x);
#else
    return 63u - (uint32_t)__builtin_clzll(x);
#endif
}
// level of the split between sorted positions i and i+1 (ka, kb = their keys)
B2_HD uint32_t split_level(uint64_t ka, uint64_t kb, uint32_t i)
{
    const uint64_t x = ka ^ kb;
    return x ? 64u + msb64(x) : msb32(i ^ (i + 1u));
}

// w[3] = far bound (30 bits) | wild << 30 | level bit 6 << 31.  bound <= n - 1 < 2^30 - 1, so w[3] is never UNVISITED.
B2_HD Xch xch_pack(const QBox& box, uint32_t bound, uint32_t level, bool wild)
{
    Xch m;
    m.w[0] = (box.w[0] & 0x7FFF7FFFu) | ((level & 1u) << 15) | ((level & 2u) << 30);
    m.w[1] = (box.w[1] & 0x7FFF7FFFu) | ((level & 4u) << 13) | ((level & 8u) << 28);
    m.w[2] = (box.w[2] & 0x7FFF7FFFu) | ((level & 16u) << 11) | ((level & 32u) << 26);
    m.w[3] = bound | (wild ? WILD_BIT : 0u) | ((level & 64u) << 25);
    return m;
}
B2_HD void xch_unpack(const Xch& m, QBox& box, uint32_t& bound, uint32_t& level, bool& wild)
{
    level = ((m.w[0] >> 15) & 1u) | ((m.w[0] >> 30) & 2u) | ((m.w[1] >> 13) & 4u) | ((m.w[1] >> 28) & 8u) | ((m.w[2] >> 11) & 16u) |
            ((m.w[2] >> 26) & 32u) | ((m.w[3] >> 25) & 64u);
    box.w[0] = m.w[0] | 0x80008000u;
    box.w[1] = m.w[1] | 0x80008000u;
    box.w[2] = m.w[2] | 0x80008000u;
    bound = m.w[3] & REF_INDEX_MASK;
    wild = (m.w[3] & WILD_BIT) != 0u;
}

template <class Ops, class Key>
B2_HD void build_from_leaf_q(uint32_t i, uint32_t n, const Key* keys, const uint32_t* sorted_prim, const Aabb24* world_aabb,
                             const uint32_t* col_entity, Xch* xch, const Grid& grid, QNode* qnodes, LeafRec* leaves, uint32_t* root_out,
                             uint32_t* has_wild_out)
{
    const uint32_t prim = sorted_prim[i];
    LeafRec rec;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        rec.lo[k] = world_aabb[prim].lo[k];
        rec.hi[k] = world_aabb[prim].hi[k];
    }
    rec.entity = col_entity[prim];
    rec.prim = prim;
    Ops::store_leaf(&leaves[i], rec);
    QBox box = quantise_box(rec.lo, rec.hi, grid);
    bool wild = box_is_wild(rec.lo, rec.hi);
    if (n == 1) {  // a lone leaf still gets a node; both slots = the leaf (an empty slot has no grid encoding)
        QNode nd;
#pragma unroll
        for (int k = 0; k < 3; ++k) nd.w[k] = nd.w[3 + k] = box.w[k];
        nd.w[6] = LEAF_BIT | (wild ? WILD_BIT : 0u) | 0u;
        nd.w[7] = nd.w[6];
        Ops::store_qnode(&qnodes[0], nd);
        *root_out = 0u;
        *has_wild_out = wild ? 1u : 0u;
        return;
    }
    const uint64_t key = (uint64_t)keys[i];
    uint32_t l = i, r = i;
    uint32_t lev_l = i > 0u ? split_level((uint64_t)keys[i - 1u], key, i - 1u) : LEVEL_END;  // split between l-1 and l
    uint32_t lev_r = i + 1u < n ? split_level(key, (uint64_t)keys[i + 1u], i) : LEVEL_END;    // split between r and r+1
    uint32_t ref = LEAF_BIT | i;
    for (;;) {
        const bool is_left = lev_r < lev_l;  // join the neighbour beyond the LOWER split; array ends never win
        const uint32_t p = is_left ? r : l - 1u;
        Ops::store_ref(&qnodes[p], is_left ? 0 : 1, ref | (wild ? WILD_BIT : 0u));
        const Xch mine = xch_pack(box, is_left ? l : r, is_left ? lev_l : lev_r, wild);
        const Xch got = Ops::exchange(&xch[p], mine);
        if (got.w[3] == UNVISITED) return;  // first to arrive: the sibling's thread finishes this node
        QBox sib;
        uint32_t bound, level;
        bool sib_wild;
        xch_unpack(got, sib, bound, level, sib_wild);
        wild = wild || sib_wild;
        Ops::store_qboxes(&qnodes[p], is_left ? box.w : sib.w, is_left ? sib.w : box.w);
        if (is_left) {
            r = bound;
            lev_r = level;
        }
        else {
            l = bound;
            lev_l = level;
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) box.w[k] = qplane_merge(box.w[k], sib.w[k]);
        ref = p;
        if (lev_l == LEVEL_END && lev_r == LEVEL_END) {  // the range is the whole array
            *root_out = p;
            *has_wild_out = wild ? 1u : 0u;
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

B2_HD void export_internal(uint32_t p, uint32_t n, uint32_t root, const Node* nodes, const LeafRec* leaves, BiTreeNode60* out,
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
        o.index1 = leaves[nd.s[0].ref & ~LEAF_BIT].entity;
        o.index2 = leaves[nd.s[1].ref & ~LEAF_BIT].entity;
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

B2_HD void export_wrapper(uint32_t i, const LeafRec* leaves, BiTreeNode60* out)
{
    BiTreeNode60 o;
    const LeafRec b = leaves[i];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        o.box1[k] = b.lo[k];
        o.box1[3 + k] = b.hi[k];
        o.box2[k] = 0.0f;  // value-initialised, as BiTreeNode node{} at reference :223
        o.box2[3 + k] = 0.0f;
    }
    o.index1 = b.entity;
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
// exported tree.  Internal (and quantised leaf) boxes only prune: a child is skipped when its grid box
// misses (superset of the exact box, monotone slab test => a hit leaf implies hit ancestors) or when its
// grid-space entry tmin is greater than the best exact tmin (strictly: an equal tmin may still hide a
// later-position tie winner).
constexpr int LBVH_STACK = 96;  // depth <= key bits (63) + position bits (<= 31): cannot overflow

struct Hit {
    float t;
    int32_t leaf;  // Morton position of the winning leaf, -1 = none
};

// The ray in normalised grid space (a cell plane q sits at x_n = 1 + q*2^-15):
//     t = (x_n - o_n) * n,   n = 2^15*cell / d,   evaluated as fma(x_n, n, c) with c = -(o_n*n).
// |n| is clamped to 2^40: an (almost) axis-parallel ray then keeps a finite, correctly signed slab
// interval instead of inf - inf; 1.5 cells * 2^40 * 2^-15 dwarfs any T, so the clamp can only ever turn
// a provable miss into a (harmless) false positive.  sel_near / sel_far pick the lo or hi plane of each
// axis' word according to the sign of the direction.
constexpr uint32_t SEL_LO16 = 0x7104u, SEL_HI16 = 0x7324u;  // bytes (const.0, w.0|w.2, w.1|w.3, const.3)
struct RayQ {
    float n[3];
    float c[3];
    uint32_t sel_near[3], sel_far[3];
};

// false => the origin is more than 2^22 cells from the grid (or infinite): every box on the grid is then more than
// 16 T away (make_grid), so the reference's slab test (tmin < T) rejects them all: a miss — unless the tree holds wild
// leaves, which the caller must then still visit.  A NaN origin coordinate or reciprocal is NOT a miss: the reference's
// fminf / fmaxf drop the NaN slab values of that axis (src/system_collisions.cpp:67-83) and the other axes decide; here
// c (and n) become NaN, the axis' near / far values become NaN and max3f / min3f drop them likewise.
B2_HD bool rayq_prepare(const RayPre& r, const Grid& g, RayQ& q)
{
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float og = (r.o[k] - g.gmin[k]) * g.inv_cell[k];
        ok = ok && !(fabsf(og) >= 4194304.0f);
        const float on = 1.0f + og * (1.0f / GRID_CELLS);
        float n = r.ninv[k] * (g.cell[k] * GRID_CELLS);
        const float nc = fminf(fmaxf(n, -1.099511627776e12f), 1.099511627776e12f);  // +-2^40
        n = (n == n) ? nc : n;                                                         // NaN stays NaN: the axis drops out
        q.n[k] = n;
        q.c[k] = -(on * n);
        const bool neg = n < 0.0f;
        q.sel_near[k] = neg ? SEL_HI16 : SEL_LO16;
        q.sel_far[k] = neg ? SEL_LO16 : SEL_HI16;
    }
    return ok;
}

B2_HD float cell_plane(uint32_t w, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    // prmt directly (not __byte_perm, which masks the selector first): the selectors only ever hold
    // nibbles < 8, so the generic mode picks plain bytes
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(0x3F000000u), "r"(sel));
    return __uint_as_float(r);
#else
    union { uint32_t u; float f; } cv;
    const uint32_t s16 = (sel == SEL_LO16) ? (w & 0xFFFFu) : (w >> 16);
    cv.u = 0x3F000000u | (s16 << 8);
    return cv.f;
#endif
}

B2_HD float max3f(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
B2_HD float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }

// Largest float below a positive finite T.
B2_HD float float_pred(float T)
{
    union { uint32_t u; float f; } cv;
    cv.f = T;
    cv.u -= 1u;
    return cv.f;
}

// Grid-space slab test of one slot (wx, wy, wz = the slot's three axis words) against the pruning limit
//     lim = best.t  once a hit exists (every hit has tmin < T),  else the largest float below T,
// so that  tmin <= lim  <=>  tmin <= best.t && tmin < T  (the reference's cut-off, :99, and the
// nearest-hit pruning in one comparison; equality must pass: an equal tmin may hide a later tie winner).
B2_HD bool qslab(const RayQ& q, uint32_t wx, uint32_t wy, uint32_t wz, float lim, float& tmin_out)
{
    const float tnx = fmaf(cell_plane(wx, q.sel_near[0]), q.n[0], q.c[0]);
    const float tny = fmaf(cell_plane(wy, q.sel_near[1]), q.n[1], q.c[1]);
    const float tnz = fmaf(cell_plane(wz, q.sel_near[2]), q.n[2], q.c[2]);
    const float tfx = fmaf(cell_plane(wx, q.sel_far[0]), q.n[0], q.c[0]);
    const float tfy = fmaf(cell_plane(wy, q.sel_far[1]), q.n[1], q.c[1]);
    const float tfz = fmaf(cell_plane(wz, q.sel_far[2]), q.n[2], q.c[2]);
    const float tmin = max3f(tnx, tny, tnz);
    const float tmax = min3f(tfx, tfy, tfz);
    tmin_out = tmin;
    return (tmax >= tmin) && (tmax >= 0.0f) && (tmin <= lim);
}

// Exact test of a candidate leaf (the reference's RayBoxIntersection on the exact world AABB).
// Returns true when the leaf became the best hit.
B2_HD bool leaf_candidate(Hit& best, const RayPre& r, const LeafRec& rec, uint32_t ref, float T)
{
    const SlabT s = slab(r, rec.lo, rec.hi);
    if (!slab_hit(s, T)) return false;
    const int32_t leaf = (int32_t)(ref & ~LEAF_BIT);
    if (s.tmin < best.t || (s.tmin == best.t && leaf > best.leaf)) {
        best.t = s.tmin;
        best.leaf = leaf;
        return true;
    }
    return false;
}

// One node visit, stack-free: tests both slots, resolves leaf slots exactly (updating best / lim) and
// tells the caller where to go.  RayFn() yields the exact ray; it is only evaluated when a leaf slot
// passes the grid test — about once per ray.
//   next == REF_EMPTY : nothing to descend into, the caller pops
//   push              : `far` (entry tmin `tfar`) is to be deferred, `next` is the nearer child
struct VisitOut {
    uint32_t next, far;
    float tfar;
    bool push;
};

// WILD: the tree holds wild leaves (see WILD_BIT): a flagged slot is entered whatever its grid box says, as early as possible.
template <class Ops, bool WILD, class RayFn>
B2_HD VisitOut visit_node(Hit& best, float& lim, const RayQ& q, const QNode& nd, const LeafRec* leaves, float T, RayFn exact_ray)
{
    float ta, tb;
    bool ha = qslab(q, nd.w[0], nd.w[1], nd.w[2], lim, ta);
    bool hb = qslab(q, nd.w[3], nd.w[4], nd.w[5], lim, tb);
    uint32_t ra = nd.w[6], rb = nd.w[7];
    if (WILD) {
        if (ra & WILD_BIT) {
            ha = true;
            ta = -B2_INF;
        }
        if (rb & WILD_BIT) {
            hb = true;
            tb = -B2_INF;
        }
        ra &= ~WILD_BIT;
        rb &= ~WILD_BIT;
    }
    const bool la = ha && (ra & LEAF_BIT), lb = hb && (rb & LEAF_BIT);
    if (la || lb) {
        const RayPre r = exact_ray();
        bool better = false;
        if (la) better = leaf_candidate(best, r, Ops::load_leaf(&leaves[ra & ~LEAF_BIT]), ra, T);
        if (lb) better = leaf_candidate(best, r, Ops::load_leaf(&leaves[rb & ~LEAF_BIT]), rb, T) || better;
        // Not below -2^26: an exact tmin of -inf (every axis either dropped as NaN or entered with an infinite reciprocal from
        // inside the slab) has a FINITE image in grid space, where |n| is clamped to 2^40 — at most -2^-14 * 2^40, the
        // origin being at least the 2-cell margin inside the grid box — and a later leaf tying at -inf must still be reached.
        if (better) lim = fmaxf(best.t, -67108864.0f);
        ha = ha && !la && (ta <= lim);
        hb = hb && !lb && (tb <= lim);
    }
    VisitOut o;
    const bool a_first = !(tb < ta);
    o.push = ha && hb;
    o.far = a_first ? rb : ra;
    o.tfar = a_first ? tb : ta;
    o.next = o.push ? (a_first ? ra : rb) : (ha ? ra : (hb ? rb : REF_EMPTY));
    return o;
}

// Raycast epilogue for the winner: the exact slab test is repeated on the winning leaf record to obtain
// the entry side (bit-identical inputs, bit-identical result), src/system_collisions.cpp:85-97,170-194.
template <class Ops>
B2_HD void finish_raycast(uint32_t* out9, const RayPre& r, const Hit& best, const LeafRec* leaves)
{
    if (best.leaf < 0) {
        write_raycast(out9, r, false, 0.0f, 0, 0u);
        return;
    }
    const LeafRec rec = Ops::load_leaf(&leaves[best.leaf]);
    const SlabT s = slab(r, rec.lo, rec.hi);
    write_raycast(out9, r, true, s.tmin, slab_side(s), rec.entity);
}

// The same winner as a 12-byte b2col_hit (include/b2col.h): entity, distance, hit | side << 8 — everything a Raycast holds that
// is not a function of the ray itself (expand_hit below restores the rest on the host, bit for bit).
template <class Ops>
B2_HD void finish_hit(uint32_t* out3, const RayPre& r, const Hit& best, const LeafRec* leaves)
{
    if (best.leaf < 0) {
        out3[0] = out3[1] = out3[2] = 0u;
        return;
    }
    const LeafRec rec = Ops::load_leaf(&leaves[best.leaf]);
    const SlabT s = slab(r, rec.lo, rec.hi);
    union {
        float f;
        uint32_t u;
    } cv;
    cv.f = s.tmin;
    out3[0] = rec.entity;
    out3[1] = cv.u;
    out3[2] = 1u | ((uint32_t)slab_side(s) << 8);
}

// b2col_hit -> Raycast (GetRaycast's epilogue, src/system_collisions.cpp:162,170-194,489): the direction is normalised again and
// location = d * t + o, normal = axis of the side — the arithmetic of write_raycast, so the bytes equal the 36-byte path's.
B2_HD void expand_hit(uint32_t* out9, const float* origin, const float* direction, const uint32_t* hit3)
{
    const RayPre r = ray_prepare(origin, direction);
    union {
        float f;
        uint32_t u;
    } cv;
    cv.u = hit3[1];
    write_raycast(out9, r, (hit3[2] & 1u) != 0u, cv.f, (int)((hit3[2] >> 8) & 7u), hit3[0]);
}

// Whole-ray form with a plain array stack (host emulation; the device kernel, k_trace_lbvh in b2col.cu,
// runs the same visit_node per step with a shared-memory stack and interleaves ray fetching).  A
// deferred child whose entry tmin has meanwhile fallen behind the best hit is dropped when popped.
template <class Ops, bool WILD>
B2_HD Hit trace_lbvh(const RayPre& r, const Grid& grid, const QNode* qnodes, const LeafRec* leaves, uint32_t root, float T)
{
    Hit best;
    best.t = B2_INF;
    best.leaf = -1;
    RayQ q;
    if (!rayq_prepare(r, grid, q) && !WILD) return best;
    float lim = float_pred(T);
    uint32_t s_node[LBVH_STACK];
    float s_t[LBVH_STACK];
    int sp = 0;
    uint32_t cur = root;
    for (;;) {
        if (cur == REF_EMPTY) {
            if (sp == 0) break;
            --sp;
            if (!(s_t[sp] <= lim)) continue;
            cur = s_node[sp];
        }
        const QNode nd = Ops::load_qnode(&qnodes[cur]);
        const VisitOut v = visit_node<Ops, WILD>(best, lim, q, nd, leaves, T, [&]() { return r; });
        if (v.push) {
            s_node[sp] = v.far;
            s_t[sp] = v.tfar;
            ++sp;
        }
        cur = v.next;
    }
    return best;
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
