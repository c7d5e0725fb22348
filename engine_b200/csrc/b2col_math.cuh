// b2col_math.cuh — the parity-critical arithmetic of the collision path, as host/device inline
// functions.  Every operation that the reference rounds once is rounded once here: on the device
// through the _rn intrinsics (never contracted to FMA, whatever the compile flags), on the host
// (tests/emu only, built with -ffp-contract=off) through plain operators.  Evaluation order follows
// SURVEY.md Appendix B, which restates reference src/system_transform.cpp:22-36,
// src/system_collisions.cpp:14-41,62-100,162 and the vendored glm 1.0.1 kernels they call.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline
#endif

#if defined(__CUDA_ARCH__)
#define B2_MUL(a, b) __fmul_rn((a), (b))
#define B2_ADD(a, b) __fadd_rn((a), (b))
#define B2_SUB(a, b) __fsub_rn((a), (b))
#define B2_DIV(a, b) __fdiv_rn((a), (b))
#define B2_SQRT(a) __fsqrt_rn((a))
#define B2_INF __int_as_float(0x7f800000)
#define B2_NAN __int_as_float(0x7fc00000)
#else
#define B2_MUL(a, b) ((a) * (b))
#define B2_ADD(a, b) ((a) + (b))
#define B2_SUB(a, b) ((a) - (b))
#define B2_DIV(a, b) ((a) / (b))
#define B2_SQRT(a) sqrtf((a))
#define B2_INF INFINITY
#define B2_NAN NAN
#endif

namespace b2 {

// ---- traversal node: two child slots of 32 bytes each (one DRAM/L2 sector per slot) ----------
constexpr uint32_t LEAF_BIT = 0x80000000u;
constexpr uint32_t REF_EMPTY = 0xFFFFFFFFu;  // slot holds nothing (its box is all-NaN: never hit)

struct alignas(16) Slot {
    float lo[3];
    float hi[3];
    uint32_t ref;  // LEAF_BIT | leaf position in Morton order, or internal node index, or REF_EMPTY
    uint32_t aux;  // slot 0: first leaf position of the NODE's range; slot 1: last leaf position
};
struct alignas(64) Node {
    Slot s[2];
};
static_assert(sizeof(Slot) == 32 && sizeof(Node) == 64, "node layout");

struct Box {
    float lo[3], hi[3];
};

// AABBSide — reference include/system_collisions.h:26
enum Side : int { SIDE_LEFT = 0, SIDE_RIGHT, SIDE_BOTTOM, SIDE_TOP, SIDE_FRONT, SIDE_BACK };

// ---- TransformSystem::onUpdate arithmetic ------------------------------------------------------
// Local matrix: glm::mat4_cast(q) (vendor/glm/glm/gtc/quaternion.inl:47-78), column 3 = position
// (src/system_transform.cpp:24), glm::scale (vendor/glm/glm/ext/matrix_transform.inl:78-86).
// m is column-major 4x4: m[4*c + r].  The bottom row is 0*s (sign follows s) and 1, as glm computes it.
B2_HD void local_matrix(const float* p, const float* q, const float* s, float* m)
{
    const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    const float qxx = B2_MUL(qx, qx), qyy = B2_MUL(qy, qy), qzz = B2_MUL(qz, qz);
    const float qxz = B2_MUL(qx, qz), qxy = B2_MUL(qx, qy), qyz = B2_MUL(qy, qz);
    const float qwx = B2_MUL(qw, qx), qwy = B2_MUL(qw, qy), qwz = B2_MUL(qw, qz);
    m[0] = B2_MUL(B2_SUB(1.0f, B2_MUL(2.0f, B2_ADD(qyy, qzz))), s[0]);
    m[1] = B2_MUL(B2_MUL(2.0f, B2_ADD(qxy, qwz)), s[0]);
    m[2] = B2_MUL(B2_MUL(2.0f, B2_SUB(qxz, qwy)), s[0]);
    m[3] = B2_MUL(0.0f, s[0]);
    m[4] = B2_MUL(B2_MUL(2.0f, B2_SUB(qxy, qwz)), s[1]);
    m[5] = B2_MUL(B2_SUB(1.0f, B2_MUL(2.0f, B2_ADD(qxx, qzz))), s[1]);
    m[6] = B2_MUL(B2_MUL(2.0f, B2_ADD(qyz, qwx)), s[1]);
    m[7] = B2_MUL(0.0f, s[1]);
    m[8] = B2_MUL(B2_MUL(2.0f, B2_ADD(qxz, qwy)), s[2]);
    m[9] = B2_MUL(B2_MUL(2.0f, B2_SUB(qyz, qwx)), s[2]);
    m[10] = B2_MUL(B2_SUB(1.0f, B2_MUL(2.0f, B2_ADD(qxx, qyy))), s[2]);
    m[11] = B2_MUL(0.0f, s[2]);
    m[12] = p[0];
    m[13] = p[1];
    m[14] = p[2];
    m[15] = 1.0f;
}

// glm mat4 * mat4, generic path (vendor/glm/glm/detail/type_mat4x4.inl:630-658):
// out.col[j] = ((a.col0*b[j][0] + a.col1*b[j][1]) + a.col2*b[j][2]) + a.col3*b[j][3]
B2_HD void mat_mul(const float* a, const float* b, float* out)
{
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float acc = B2_MUL(a[r], b[4 * j]);
            acc = B2_ADD(acc, B2_MUL(a[4 + r], b[4 * j + 1]));
            acc = B2_ADD(acc, B2_MUL(a[8 + r], b[4 * j + 2]));
            acc = B2_ADD(acc, B2_MUL(a[12 + r], b[4 * j + 3]));
            out[4 * j + r] = acc;
        }
    }
}

// transformBox — src/system_collisions.cpp:14-41; mat4*vec4(p,1) generic path
// (vendor/glm/glm/detail/type_mat4x4.inl:540-577): (m0*x + m1*y) + (m2*z + m3*1)
B2_HD Box transform_box(const float* lmin, const float* lmax, const float* m)
{
    Box nb;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        nb.lo[k] = B2_INF;
        nb.hi[k] = -B2_INF;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float x = (i & 4) ? lmax[0] : lmin[0];
        const float y = (i & 2) ? lmax[1] : lmin[1];
        const float z = (i & 1) ? lmax[2] : lmin[2];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const float a0 = B2_ADD(B2_MUL(m[r], x), B2_MUL(m[4 + r], y));
            const float a1 = B2_ADD(B2_MUL(m[8 + r], z), B2_MUL(m[12 + r], 1.0f));
            const float v = B2_ADD(a0, a1);
            nb.lo[r] = fminf(nb.lo[r], v);
            nb.hi[r] = fmaxf(nb.hi[r], v);
        }
    }
    return nb;
}

// ---- query arithmetic --------------------------------------------------------------------------
// glm::normalize as GetRaycast applies it (src/system_collisions.cpp:162;
// vendor/glm/glm/detail/func_geometric.inl:53,84-90; func_exponential.inl:138)
B2_HD void normalize3(const float* v, float* out)
{
    const float d = B2_ADD(B2_ADD(B2_MUL(v[0], v[0]), B2_MUL(v[1], v[1])), B2_MUL(v[2], v[2]));
    const float inv = B2_DIV(1.0f, B2_SQRT(d));
    out[0] = B2_MUL(v[0], inv);
    out[1] = B2_MUL(v[1], inv);
    out[2] = B2_MUL(v[2], inv);
}

struct RayPre {
    float o[3];     // origin
    float d[3];     // normalised direction
    float ninv[3];  // 1/d per axis (+-inf allowed) — RayBoxIntersection :65, hoisted: a pure function of the ray
};

B2_HD RayPre ray_prepare(const float* origin, const float* direction)
{
    RayPre r;
    r.o[0] = origin[0];
    r.o[1] = origin[1];
    r.o[2] = origin[2];
    normalize3(direction, r.d);
    r.ninv[0] = B2_DIV(1.0f, r.d[0]);
    r.ninv[1] = B2_DIV(1.0f, r.d[1]);
    r.ninv[2] = B2_DIV(1.0f, r.d[2]);
    return r;
}

struct SlabT {
    float tx1, tx2, ty1, ty2, tz1, tz2, tmin, tmax;
};

// RayBoxIntersection — src/system_collisions.cpp:67-83
B2_HD SlabT slab(const RayPre& r, const float* lo, const float* hi)
{
    SlabT s;
    s.tx1 = B2_MUL(B2_SUB(lo[0], r.o[0]), r.ninv[0]);
    s.tx2 = B2_MUL(B2_SUB(hi[0], r.o[0]), r.ninv[0]);
    float tmin = fminf(s.tx1, s.tx2);
    float tmax = fmaxf(s.tx1, s.tx2);
    s.ty1 = B2_MUL(B2_SUB(lo[1], r.o[1]), r.ninv[1]);
    s.ty2 = B2_MUL(B2_SUB(hi[1], r.o[1]), r.ninv[1]);
    tmin = fmaxf(tmin, fminf(s.ty1, s.ty2));
    tmax = fminf(tmax, fmaxf(s.ty1, s.ty2));
    s.tz1 = B2_MUL(B2_SUB(lo[2], r.o[2]), r.ninv[2]);
    s.tz2 = B2_MUL(B2_SUB(hi[2], r.o[2]), r.ninv[2]);
    tmin = fmaxf(tmin, fminf(s.tz1, s.tz2));
    tmax = fminf(tmax, fmaxf(s.tz1, s.tz2));
    s.tmin = tmin;
    s.tmax = tmax;
    return s;
}

// hit predicate — src/system_collisions.cpp:99
B2_HD bool slab_hit(const SlabT& s, float T) { return (s.tmax >= fmaxf(0.0f, s.tmin)) && (s.tmin < T); }

// side chain — src/system_collisions.cpp:85-97
B2_HD int slab_side(const SlabT& s)
{
    int side = SIDE_LEFT;
    if (s.tmin == s.tx2) side = SIDE_RIGHT;
    else if (s.tmin == s.tx1) side = SIDE_LEFT;
    else if (s.tmin == s.ty2) side = SIDE_BACK;
    else if (s.tmin == s.ty1) side = SIDE_FRONT;
    else if (s.tmin == s.tz2) side = SIDE_TOP;
    else if (s.tmin == s.tz1) side = SIDE_BOTTOM;
    return side;
}

// Raycast epilogue — src/system_collisions.cpp:164-194 and :489 (location = d*t + o, mul then add).
// out9 = 36 bytes as nine 32-bit words: location.xyz, normal.xyz, hit_entity, distance, hit|pad.
B2_HD void write_raycast(uint32_t* out9, const RayPre& r, bool hit, float t, int side, uint32_t entity)
{
    union {
        float f;
        uint32_t u;
    } cv;
    if (!hit) {
#pragma unroll
        for (int k = 0; k < 9; ++k) out9[k] = 0u;
        return;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        cv.f = B2_ADD(B2_MUL(r.d[k], t), r.o[k]);
        out9[k] = cv.u;
    }
    float n[3] = {0.0f, 0.0f, 0.0f};
    switch (side) {
        case SIDE_LEFT: n[0] = -1.0f; break;
        case SIDE_RIGHT: n[0] = 1.0f; break;
        case SIDE_BOTTOM: n[2] = -1.0f; break;
        case SIDE_TOP: n[2] = 1.0f; break;
        case SIDE_FRONT: n[1] = -1.0f; break;
        case SIDE_BACK: n[1] = 1.0f; break;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        cv.f = n[k];
        out9[3 + k] = cv.u;
    }
    out9[6] = entity;
    cv.f = t;
    out9[7] = cv.u;
    out9[8] = 1u;  // hit = true, 3 padding bytes zero (little endian)
}

// ---- Morton keys ---------------------------------------------------------------------------------
// Spread the low 21 bits of v so that there are two zero bits between consecutive bits.
B2_HD uint64_t spread3(uint32_t v)
{
    uint64_t x = v & 0x1fffffull;
    x = (x | (x << 32)) & 0x1f00000000ffffull;
    x = (x | (x << 16)) & 0x1f0000ff0000ffull;
    x = (x | (x << 8)) & 0x100f00f00f00f00full;
    x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

// Spread the low 32 bits of v so that there is one zero bit between consecutive bits.
B2_HD uint64_t spread2(uint32_t v)
{
    uint64_t x = v;
    x = (x | (x << 16)) & 0x0000ffff0000ffffull;
    x = (x | (x << 8)) & 0x00ff00ff00ff00ffull;
    x = (x | (x << 4)) & 0x0f0f0f0f0f0f0f0full;
    x = (x | (x << 2)) & 0x3333333333333333ull;
    x = (x | (x << 1)) & 0x5555555555555555ull;
    return x;
}

// ---- Morton keys with the bits handed out by extent ---------------------------------------------
// A Morton code that gives every axis the same number of bits cuts a flat scene (a game world: 1000 x 1000 x 50)
// into cells 20x thinner in z than in x and y, and the radix tree then splits along z where that separates
// nothing (measured: 31 % fewer rays/s on BASELINE config 2 than with cubic cells, profiles/r1_e_morton_cells.jsonl).
// The K key bits are therefore distributed so that the cells come out as cubic as powers of two allow: with the
// axes ordered by centroid extent a >= b >= c, axis a gets d1 bits more than b and b gets d2 more than c
// (d = how many times the smaller extent must be doubled to come within sqrt 2 of the larger).  The code is,
// from the most significant bit: d1 bits of a alone, d2 rounds interleaving (a, b), nc rounds interleaving
// (a, b, c); the lowest `drop` bits are discarded so that exactly K bits remain.  For a roughly cubic scene
// d1 = d2 = 0 and this is the plain 3-way interleave.  Only steers tree topology; never affects which box a
// ray hits.  Everything below is exact float / integer arithmetic shared by the kernel and the host-side
// restatement of the tests.
struct MortonPlan {
    float smin[3], sinv[3];  // per ORIGINAL axis: q = (centroid - smin) * sinv, in [0, 2^qbits)
    float qmax[3];           // 2^qbits - 1
    int axis[3];             // a, b, c: original axis indices by decreasing extent
    int shift_a;             // axis a is quantised to at most 21 bits; the code wants na = qbits_a + shift_a
    int nc, d2, drop;
    int fast32;              // plain three-way interleave of at most 11 bits per axis into at most 32 key bits: morton_key's 32-bit path
};

constexpr int MORTON_AXIS_BITS = 21;  // spread3 / float quantisation limit per axis

// number of doublings that bring `small` within sqrt(2) of `big` (0 if already there), capped
B2_HD int morton_ratio_steps(float big, float small, int cap)
{
    if (!(small > 0.0f)) return cap;
    int d = 0;
    float e = small;
    while (d < cap && B2_MUL(e, 1.41421354f) < big) {
        e = B2_MUL(e, 2.0f);
        ++d;
    }
    return d;
}

// cmin / cmax: bounds of the finite centroids (an axis without any: cmin > cmax); key_bits <= 63
B2_HD MortonPlan make_morton_plan(const float* cmin, const float* cmax, int key_bits)
{
    MortonPlan p;
    float ext[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float e = B2_SUB(cmax[k], cmin[k]);
        ext[k] = (e > 0.0f && e < B2_INF) ? e : 0.0f;
        p.smin[k] = cmin[k];
    }
    // stable sort of the three axes by decreasing extent
    int a = 0, b = 1, c = 2;
    if (ext[b] > ext[a]) { const int t = a; a = b; b = t; }
    if (ext[c] > ext[b]) { const int t = b; b = c; c = t; }
    if (ext[b] > ext[a]) { const int t = a; a = b; b = t; }
    int d1 = morton_ratio_steps(ext[a], ext[b], key_bits);
    int d2 = morton_ratio_steps(ext[b], ext[c], key_bits);
    p.axis[0] = a;
    p.axis[1] = b;
    p.axis[2] = c;
    if (d1 > key_bits) d1 = key_bits;
    if (2 * d2 > key_bits - d1) d2 = (key_bits - d1 + 1) / 2;
    if (d2 > MORTON_AXIS_BITS) d2 = MORTON_AXIS_BITS;
    int rest = key_bits - d1 - 2 * d2;
    int nc = rest > 0 ? (rest + 2) / 3 : 0;
    if (3 * nc + 2 * d2 + d1 > 63) nc = rest / 3;  // the undropped code must fit 64 bits
    if (nc + d2 > MORTON_AXIS_BITS) nc = MORTON_AXIS_BITS - d2;  // axis b is limited to 21 bits (only reachable for key_bits > 42 on flat scenes)
    int na = nc + d2 + d1;
    const int total = 3 * nc + 2 * d2 + d1;
    if (total < key_bits) na += key_bits - total;  // b and c are at their limit: the remaining bits go to a (low bits, zero)
    const int qa = na > MORTON_AXIS_BITS ? MORTON_AXIS_BITS : na;
    p.shift_a = na - qa;
    p.nc = nc;
    p.d2 = d2;
    p.drop = total > key_bits ? total - key_bits : 0;
    p.fast32 = (d1 == 0 && d2 == 0 && p.shift_a == 0 && na == nc && nc >= 1 && nc <= 11 && p.drop <= 2 && 3 * nc - p.drop <= 32) ? 1 : 0;
    const int qbits[3] = {qa, nc + d2, nc};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const int k = p.axis[r];
        const float cells = (float)(1u << qbits[r]);
        p.sinv[k] = ext[k] > 0.0f ? B2_DIV(cells, ext[k]) : 0.0f;
        p.qmax[k] = B2_SUB(cells, 1.0f);
    }
    return p;
}

// spread3 for values below 2^11, in 32-bit arithmetic (the same steps as spread3 with the masks' low halves; bit 10 lands on bit 30)
B2_HD uint32_t spread3_11(uint32_t v)
{
    uint32_t x = v & 0x7FFu;
    x = (x | (x << 16)) & 0xFF0000FFu;
    x = (x | (x << 8)) & 0x0F00F00Fu;
    x = (x | (x << 4)) & 0xC30C30C3u;
    x = (x | (x << 2)) & 0x49249249u;
    return x;
}

// Centroid (GetBoxCentroid, src/system_collisions.cpp:43-50: (min+max)*0.5f) -> key
B2_HD uint64_t morton_key(const float* lo, const float* hi, const MortonPlan& p)
{
    uint32_t q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float c = B2_MUL(B2_ADD(lo[k], hi[k]), 0.5f);
        float f = B2_MUL(B2_SUB(c, p.smin[k]), p.sinv[k]);  // in [0, 2^qbits) for finite centroids
        f = fminf(fmaxf(f, 0.0f), p.qmax[k]);               // NaN -> 0 (fmaxf drops NaN)
        q[k] = (uint32_t)f;
    }
    const uint32_t qa32 = p.axis[0] == 0 ? q[0] : (p.axis[0] == 1 ? q[1] : q[2]);
    const uint32_t qb = p.axis[1] == 0 ? q[0] : (p.axis[1] == 1 ? q[1] : q[2]);
    const uint32_t qc = p.axis[2] == 0 ? q[0] : (p.axis[2] == 1 ? q[1] : q[2]);
    if (p.fast32) {
        // cubic scene, key of at most 32 bits (every packed-key build): the code is the plain interleave a b c a b c ... with the lowest
        // `drop` bits cut — c's bit 0, then b's — which 32-bit arithmetic can form directly: a third of the instructions of the
        // general 64-bit path below, bit for bit the same key (tests/test_emu_logic.py checks both against an independent interleave)
        const uint32_t k32 = p.drop == 0   ? (spread3_11(qa32) << 2) | (spread3_11(qb) << 1) | spread3_11(qc)
                             : p.drop == 1 ? (spread3_11(qa32) << 1) | spread3_11(qb) | (spread3_11(qc >> 1) << 2)
                                           : spread3_11(qa32) | (spread3_11(qb >> 1) << 2) | (spread3_11(qc >> 1) << 1);
        return (uint64_t)k32;
    }
    const uint64_t qa = (uint64_t)qa32 << p.shift_a;
    const uint32_t mc = (1u << p.nc) - 1u, m2 = (1u << p.d2) - 1u;
    const uint64_t low = (spread3((uint32_t)qa & mc) << 2) | (spread3(qb & mc) << 1) | spread3(qc);
    const uint64_t mid = (spread2((uint32_t)(qa >> p.nc) & m2) << 1) | spread2(qb >> p.nc);
    const uint64_t top = qa >> (p.nc + p.d2);
    return ((((top << (2 * p.d2)) | mid) << (3 * p.nc)) | low) >> p.drop;
}

// Order-preserving float <-> uint map for atomicMin/atomicMax on floats.
B2_HD uint32_t float_to_ordered(float f)
{
    union {
        float f;
        uint32_t u;
    } cv;
    cv.f = f;
    return (cv.u & 0x80000000u) ? ~cv.u : (cv.u | 0x80000000u);
}
B2_HD float ordered_to_float(uint32_t u)
{
    union {
        float f;
        uint32_t u;
    } cv;
    cv.u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    return cv.f;
}

}  // namespace b2
