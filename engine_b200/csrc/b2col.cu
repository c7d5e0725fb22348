// b2col.cu — kernels and C ABI of the B200 collision-query backend (see include/b2col.h).
// sm_100a only; no CPU path.  Build: engine_b200/build.py (nvcc -gencode arch=compute_100a,code=sm_100a).
#include "../../include/b2col.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <utility>
#include <thread>
#include <vector>

#include "b2col_bvh.cuh"
#include "b2col_math.cuh"
#include "b2col_sort.cuh"

using namespace b2;

static_assert(sizeof(b2col_hit) == 12, "compact result layout");
static_assert(sizeof(b2col_ray) == 24 && sizeof(b2col_raycast) == 36 && sizeof(b2col_aabb) == 24 && sizeof(b2col_bitreenode) == 60,
              "POD layouts must equal the reference's (include/system_collisions.h:13-24,40-48)");

// =================================================================================================
// device-side policy for the shared kernel bodies in b2col_bvh.cuh
// =================================================================================================
// 32 bytes as ONE store (sm_100a STG.E.ENL2.256): a lane that writes a whole 32-byte sector sends one full-sector write down the
// L1 -> L2 path instead of two half-filled ones (B2_STORE256=0: two 16-byte stores, A/B timing)
#ifndef B2_STORE256
#define B2_STORE256 1
#endif
__device__ __forceinline__ void store32(void* p, const float4& a, const float4& b)
{
#if B2_STORE256
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
#else
    reinterpret_cast<float4*>(p)[0] = a;
    reinterpret_cast<float4*>(p)[1] = b;
#endif
}

struct DevOps {
    static __device__ __forceinline__ void store_slot(Slot* p, const Slot& s)
    {
        float4* q = reinterpret_cast<float4*>(p);
        q[0] = make_float4(s.lo[0], s.lo[1], s.lo[2], s.hi[0]);
        q[1] = make_float4(s.hi[1], s.hi[2], __uint_as_float(s.ref), __uint_as_float(s.aux));
    }
    static __device__ __forceinline__ Slot unpack(const float4& a, const float4& b)
    {
        Slot s;
        s.lo[0] = a.x; s.lo[1] = a.y; s.lo[2] = a.z;
        s.hi[0] = a.w; s.hi[1] = b.x; s.hi[2] = b.y;
        s.ref = __float_as_uint(b.z);
        s.aux = __float_as_uint(b.w);
        return s;
    }
    static __device__ __forceinline__ Slot load_slot(const Slot* p)
    {
        const float4* q = reinterpret_cast<const float4*>(p);
        return unpack(__ldg(q), __ldg(q + 1));
    }
    static __device__ __forceinline__ Slot load_slot_coherent(const Slot* p)
    {  // written by another SM during this kernel: read at L2, never from this SM's L1
        const float4* q = reinterpret_cast<const float4*>(p);
        return unpack(__ldcg(q), __ldcg(q + 1));
    }
    // Arrival at a node: one exchange with acquire-release semantics at GPU scope.  The release half
    // publishes this thread's child slot (stored just before) to whoever reads the exchanged word; the
    // acquire half makes the sibling's slot, published the same way, visible to the loads that follow.
    static __device__ __forceinline__ uint32_t arrive(uint32_t* p, uint32_t v)
    {
        uint32_t old;
        asm volatile("atom.acq_rel.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
        return old;
    }
    // 32-byte records move as two 16-byte vectors on the store side and as ONE 256-bit read-only load
    // (sm_100a LDG.E.ENL2.256) on the traversal side: one L1 wavefront per lane per record.
    static __device__ __forceinline__ void store_qnode(QNode* p, const QNode& n)
    {
        uint4* q = reinterpret_cast<uint4*>(p);
        q[0] = make_uint4(n.w[0], n.w[1], n.w[2], n.w[3]);
        q[1] = make_uint4(n.w[4], n.w[5], n.w[6], n.w[7]);
    }
    static __device__ __forceinline__ void store_ref(QNode* p, int side, uint32_t ref) { p->w[6 + side] = ref; }
    static __device__ __forceinline__ void store_qboxes(QNode* p, const uint32_t* left, const uint32_t* right)
    {
        *reinterpret_cast<uint4*>(p) = make_uint4(left[0], left[1], left[2], right[0]);
        *reinterpret_cast<uint2*>(&p->w[4]) = make_uint2(right[1], right[2]);
    }
    // The arrival exchange of build_from_leaf_q: one relaxed 128-bit atomic at GPU scope (ATOMG.E.EXCH.128).
    // Relaxed is enough: everything the sibling needs is the exchanged value itself.
    static __device__ __forceinline__ Xch exchange(Xch* p, const Xch& v)
    {
        const unsigned long long lo = (unsigned long long)v.w[0] | ((unsigned long long)v.w[1] << 32);
        const unsigned long long hi = (unsigned long long)v.w[2] | ((unsigned long long)v.w[3] << 32);
        unsigned long long olo, ohi;
        asm volatile("{\n\t.reg .b128 v, o;\n\tmov.b128 v, {%2, %3};\n\tatom.relaxed.gpu.global.exch.b128 o, [%4], v;\n\tmov.b128 {%0, %1}, o;\n\t}"
                     : "=l"(olo), "=l"(ohi)
                     : "l"(lo), "l"(hi), "l"(p)
                     : "memory");
        Xch o;
        o.w[0] = (uint32_t)olo;
        o.w[1] = (uint32_t)(olo >> 32);
        o.w[2] = (uint32_t)ohi;
        o.w[3] = (uint32_t)(ohi >> 32);
        return o;
    }
    static __device__ __forceinline__ Xch exchange_shared(Xch* p, const Xch& v)
    {
        const unsigned long long lo = (unsigned long long)v.w[0] | ((unsigned long long)v.w[1] << 32);
        const unsigned long long hi = (unsigned long long)v.w[2] | ((unsigned long long)v.w[3] << 32);
        unsigned long long olo, ohi;
        const unsigned sa = (unsigned)__cvta_generic_to_shared(p);
        asm volatile("{\n\t.reg .b128 v, o;\n\tmov.b128 v, {%2, %3};\n\tatom.relaxed.cta.shared.exch.b128 o, [%4], v;\n\tmov.b128 {%0, %1}, o;\n\t}"
                     : "=l"(olo), "=l"(ohi)
                     : "l"(lo), "l"(hi), "r"(sa)
                     : "memory");
        Xch o;
        o.w[0] = (uint32_t)olo;
        o.w[1] = (uint32_t)(olo >> 32);
        o.w[2] = (uint32_t)ohi;
        o.w[3] = (uint32_t)(ohi >> 32);
        return o;
    }
    static __device__ __forceinline__ void store_leaf(LeafRec* p, const LeafRec& r)
    {
        store32(p, make_float4(r.lo[0], r.lo[1], r.lo[2], r.hi[0]), make_float4(r.hi[1], r.hi[2], __uint_as_float(r.entity), __uint_as_float(r.prim)));
    }
    static __device__ __forceinline__ QNode load_qnode(const QNode* p)
    {
        QNode n;
        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(n.w[0]), "=r"(n.w[1]), "=r"(n.w[2]), "=r"(n.w[3]), "=r"(n.w[4]), "=r"(n.w[5]), "=r"(n.w[6]), "=r"(n.w[7])
                     : "l"(p));
        return n;
    }
    static __device__ __forceinline__ LeafRec load_leaf(const LeafRec* p)
    {
        LeafRec r;
        asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(r.lo[0]), "=f"(r.lo[1]), "=f"(r.lo[2]), "=f"(r.hi[0]), "=f"(r.hi[1]), "=f"(r.hi[2]), "=f"(*reinterpret_cast<float*>(&r.entity)),
                       "=f"(*reinterpret_cast<float*>(&r.prim))
                     : "l"(p));
        return r;
    }
};

// Contents of the first 256 bytes of the BVH blob (device memory).
struct BlobHeader {
    uint32_t magic, n_leaves, n_nodes, root;
    uint32_t bounds_min_inv[3];  // ~ordered(min centroid): zero-initialised, raised with atomicMax
    uint32_t bounds_max[3];      //  ordered(max centroid)
    uint32_t box_min_inv[3];     // ~ordered(min finite box coordinate)
    uint32_t box_max[3];         //  ordered(max finite box coordinate)
    Grid grid;                   // quantisation grid of the traversal nodes (written by k_morton)
    uint32_t trace_root;         // where k_trace_lbvh starts: `root`, or the compact copy of the root (k_top_layout)
    uint32_t top_count;          // nodes in the compact copy of the top levels behind the leaf records (0: none)
    uint32_t has_wild;           // 1: some leaf box cannot be bounded on the grid (NaN / inf / > 1e37: WILD_BIT slots in the tree)
    uint32_t key_bits;           // Morton key width of the build (k_morton)
    uint32_t reserved[35];
};
static_assert(sizeof(BlobHeader) == 256, "blob header");
constexpr uint32_t BLOB_MAGIC = 0xB2C0117Eu;

// =================================================================================================
// kernels
// =================================================================================================

// Programmatic dependent launch (frame path, ctx->pdl): consecutive kernels of a frame are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so kernel k+1 is set up and its blocks become resident while the last
// blocks of kernel k drain; each frame kernel therefore waits HERE, before its first access to global memory, until the
// grid before it has completed and its writes are visible.  A no-op when the kernel was launched the ordinary way.
// Measured at 1M colliders (profiles/r1_e_pdl.txt): stream launches 191 -> 178 us per update, the same as one CUDA-graph
// replay (178; graph + PDL: 178).  An early cudaTriggerProgrammaticLaunchCompletion() bought another 1.5 us and was left out.
__device__ __forceinline__ void pdl_wait()
{
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();
#endif
}


// Scene bounds accumulated by the refit (12 words of the per-frame scratch, all zero between frames):
//   [0..2] ~ordered(min centroid)   [3..5] ordered(max centroid)      -> Morton grid
//   [6..8] ~ordered(min finite box coordinate)   [9..11] ordered(max) -> quantisation grid of the traversal nodes
// raised with atomicMax; finish_header copies them into the blob header.
constexpr int BOUNDS_WORDS = 12;
constexpr uint32_t NO_COLLIDER = 0xFFFFFFFFu;

// world matrix of entity e (reference src/system_transform.cpp:22-36): local T*R*S, times the parent's world matrix
// root_level: every entity of hierarchy level 0 has parent 0 (b2col_set_hierarchy), so the parent load — and with it the only
// data-dependent branch of the level — is skipped there.
// the arithmetic half: TRS values and parent ID in hand (the bulk-copy kernel reads them from shared memory)
__device__ __forceinline__ void entity_world_from(const float* p, const float4& q4, const float* s, uint32_t par, const float4* world, float* W)
{
    const float q[4] = {q4.x, q4.y, q4.z, q4.w};
    float L[16];
    local_matrix(p, q, s, L);
    if (par != 0u) {
        float P[16];
        const float4* pw = world + 4 * (size_t)par;  // written by the previous launch
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float4 v = __ldcg(pw + c);
            P[4 * c] = v.x; P[4 * c + 1] = v.y; P[4 * c + 2] = v.z; P[4 * c + 3] = v.w;
        }
        mat_mul(P, L, W);
    }
    else {
#pragma unroll
        for (int k = 0; k < 16; ++k) W[k] = L[k];
    }
}
__device__ __forceinline__ void entity_world(uint32_t e, const uint32_t* __restrict__ parent, const float* __restrict__ pos3, const float4* __restrict__ rot4,
                                             const float* __restrict__ scl3, const float4* world, float* W, bool root_level = false)
{
    const float4 q4 = __ldg(rot4 + e);
    const float p[3] = {__ldg(pos3 + 3 * (size_t)e), __ldg(pos3 + 3 * (size_t)e + 1), __ldg(pos3 + 3 * (size_t)e + 2)};
    const float s[3] = {__ldg(scl3 + 3 * (size_t)e), __ldg(scl3 + 3 * (size_t)e + 1), __ldg(scl3 + 3 * (size_t)e + 2)};
    const float q[4] = {q4.x, q4.y, q4.z, q4.w};
    float L[16];
    local_matrix(p, q, s, L);
    const uint32_t par = root_level ? 0u : __ldg(parent + e);
    if (par != 0u) {
        float P[16];
        const float4* pw = world + 4 * (size_t)par;  // written by the previous launch
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float4 v = __ldcg(pw + c);
            P[4 * c] = v.x; P[4 * c + 1] = v.y; P[4 * c + 2] = v.z; P[4 * c + 3] = v.w;
        }
        mat_mul(P, L, W);
    }
    else {
#pragma unroll
        for (int k = 0; k < 16; ++k) W[k] = L[k];
    }
}

// transformBox of collider c under matrix M (reference src/system_collisions.cpp:14-41,123-132); folds the box into
// the thread's running bounds
// The result is stored as a 32-byte record {lo, hi, entity, collider index} — the layout of a LeafRec, one sector —
// so that the build's gather by primitive index costs ONE sector per leaf (a 24-byte AABB at 24-byte stride
// straddles two sectors half of the time, and the entity ID was a third).
struct LocalBox {
    float2 a, b, c;  // the 24 bytes of a collider's local AABB (min.xyz, max.xyz) as three 8-byte loads
};
__device__ __forceinline__ LocalBox load_local_box(const Aabb24* __restrict__ col_local, uint32_t c)
{
    const float2* lp = reinterpret_cast<const float2*>(col_local + c);
    LocalBox l;
    l.a = __ldg(lp);
    l.b = __ldg(lp + 1);
    l.c = __ldg(lp + 2);
    return l;
}
__device__ __forceinline__ Box refit_one(uint32_t c, uint32_t entity, const float* M, const LocalBox& l, LeafRec* __restrict__ world_rec, float* cmin, float* cmax)
{
    const float lmin[3] = {l.a.x, l.a.y, l.b.x}, lmax[3] = {l.b.y, l.c.x, l.c.y};
    const Box b = transform_box(lmin, lmax, M);
    store32(world_rec + c, make_float4(b.lo[0], b.lo[1], b.lo[2], b.hi[0]), make_float4(b.hi[1], b.hi[2], __uint_as_float(entity), __uint_as_float(c)));
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float ctr = B2_MUL(B2_ADD(b.lo[k], b.hi[k]), 0.5f);
        if (grid_coord_ok(ctr)) {  // NaN, +-inf and |x| > 1e37 stay out of the grids (such boxes become wild leaves)
            cmin[k] = fminf(cmin[k], ctr);  // running bounds: a thread of the grid-stride kernel folds several colliders
            cmax[k] = fmaxf(cmax[k], ctr);
        }
        if (grid_coord_ok(b.lo[k])) {
            cmin[3 + k] = fminf(cmin[3 + k], b.lo[k]);
            cmax[3 + k] = fmaxf(cmax[3 + k], b.lo[k]);
        }
        if (grid_coord_ok(b.hi[k])) {
            cmin[3 + k] = fminf(cmin[3 + k], b.hi[k]);
            cmax[3 + k] = fmaxf(cmax[3 + k], b.hi[k]);
        }
    }
    return b;
}

#ifndef B2_REFIT_BLOCK
#define B2_REFIT_BLOCK 128
#endif
// threads per block of the refit kernels; update at 1M / 10M: 64 -> 179 us, 96 -> 167 / 1612, 128 -> 162 / 1466 (with 8 blocks per SM
// promised to the compiler), 160 -> 166 / 1514, 192 -> 167 / 1526, 256 -> 168 / 1534, 512 -> 182
constexpr int REFIT_BLOCK = B2_REFIT_BLOCK;
// block reduction of the per-thread bounds: warp shuffles, then one atomic per block and word
template <int BLOCK = REFIT_BLOCK>
__device__ __forceinline__ void bounds_reduce(float* cmin, float* cmax, uint32_t* bounds)
{
    // warp level: one REDUX per value on the order-preserving integer image of the float (twelve REDUX.MAX instead of sixty
    // shuffle + min/max pairs: a third of the instructions a one-entity thread of k_propagate_refit_level executes);
    // min = max of the complemented image — the form the scratch words hold anyway
    uint32_t umin_inv[6], umax[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        umin_inv[k] = __reduce_max_sync(0xffffffffu, ~float_to_ordered(cmin[k]));
        umax[k] = __reduce_max_sync(0xffffffffu, float_to_ordered(cmax[k]));
    }
    __shared__ uint32_t smin_inv[BLOCK / 32][6], smax[BLOCK / 32][6];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            smin_inv[warp][k] = umin_inv[k];
            smax[warp][k] = umax[k];
        }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int k = threadIdx.x;
        uint32_t a = smin_inv[0][k], b = smax[0][k];
        for (int w = 1; w < BLOCK / 32; ++w) {
            a = max(a, smin_inv[w][k]);
            b = max(b, smax[w][k]);
        }
        if (~a <= b) {  // ordered(min) <= ordered(max); false when the block saw no finite value (min = +inf, max = -inf)
            const int base = k < 3 ? 0 : 6, kk = k < 3 ? k : k - 3;
            atomicMax(&bounds[base + kk], a);
            atomicMax(&bounds[base + 3 + kk], b);
        }
    }
}

// TransformSystem::onUpdate (reference src/system_transform.cpp:12-38), one hierarchy level per launch:
// entities of depth k only depend on depth k-1, which the previous launch finished.
// world: 16 floats per entity, column-major, indexed by entity ID.
__global__ void __launch_bounds__(256)
k_propagate_level(const uint32_t* __restrict__ level_entities, uint32_t count, const uint32_t* __restrict__ parent,
                  const float* __restrict__ pos3, const float4* __restrict__ rot4, const float* __restrict__ scl3, float4* world)
{
    pdl_wait();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t e = level_entities[i];
    float W[16];
    entity_world(e, parent, pos3, rot4, scl3, world, W);
    float4* ow = world + 4 * (size_t)e;
#pragma unroll
    for (int c = 0; c < 4; ++c) ow[c] = make_float4(W[4 * c], W[4 * c + 1], W[4 * c + 2], W[4 * c + 3]);
}

// The same level, FUSED with the refit of the colliders that sit on this level's entities (b2col_update's normal
// path): the world matrix is still in registers when transformBox needs it, so the refit's 64-byte re-read per
// collider and one launch disappear.
// The kernel is bound by memory LATENCY, not bandwidth (ncu, profiles/r2_ncu_update_counters.txt: 3.0 TB/s at 45 % warps
// active): what a thread has in flight is one link of a chain of dependent loads.  The chain used to be entity list ->
// {TRS, parent, entity's collider index} -> {parent matrix, local box}.  level_col[i] = collider index of level_entities[i]
// (NO_COLLIDER = none) is a second list in LEVEL order, loaded together with the entity ID; and a level whose entities AND
// colliders are consecutive runs (DENSE: e = e0 + i, c = c0 + i — every flat scene, every scene whose colliders were created
// in entity order) needs neither list: all of its loads are independent and issued at once.
#ifndef B2_REFIT_MIN_BLOCKS
#define B2_REFIT_MIN_BLOCKS 8
#endif
template <bool DENSE>
__global__ void __launch_bounds__(REFIT_BLOCK, B2_REFIT_MIN_BLOCKS)
k_propagate_refit_level(const uint32_t* __restrict__ level_entities, const uint32_t* __restrict__ level_col, uint32_t e0, uint32_t c0, uint32_t count, int root_level,
                        const uint32_t* __restrict__ parent, const float* __restrict__ pos3, const float4* __restrict__ rot4, const float* __restrict__ scl3, float4* world,
                        const Aabb24* __restrict__ col_local, LeafRec* __restrict__ world_rec, uint32_t* bounds)
{
    pdl_wait();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float cmin[6] = {B2_INF, B2_INF, B2_INF, B2_INF, B2_INF, B2_INF}, cmax[6] = {-B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF};
    if (i < count) {
        const uint32_t e = DENSE ? e0 + i : level_entities[i];
        const uint32_t c = DENSE ? c0 + i : level_col[i];
        // the local box is requested together with the TRS (behind the matrix stores it would be a second round trip)
        LocalBox lb;
        lb.a = lb.b = lb.c = make_float2(0.0f, 0.0f);
        if (c != NO_COLLIDER) lb = load_local_box(col_local, c);
        float W[16];
        entity_world(e, parent, pos3, rot4, scl3, world, W, root_level != 0);
        float4* ow = world + 4 * (size_t)e;
#pragma unroll
        for (int k = 0; k < 4; k += 2)
            store32(ow + k, make_float4(W[4 * k], W[4 * k + 1], W[4 * k + 2], W[4 * k + 3]), make_float4(W[4 * k + 4], W[4 * k + 5], W[4 * k + 6], W[4 * k + 7]));
        if (c != NO_COLLIDER) refit_one(c, e, W, lb, world_rec, cmin, cmax);
    }
    bounds_reduce(cmin, cmax, bounds);
}

// Refit loop alone (reference src/system_collisions.cpp:123-132): world AABB = transformBox(local, world_matrix), for
// callers that supply world matrices themselves or propagate separately.  Also reduces the scene bounds.
__global__ void __launch_bounds__(REFIT_BLOCK)
k_refit(uint32_t n, const uint32_t* __restrict__ col_entity, const Aabb24* __restrict__ col_local, const float4* __restrict__ world,
        LeafRec* __restrict__ world_rec, uint32_t* bounds)
{
    pdl_wait();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float cmin[6] = {B2_INF, B2_INF, B2_INF, B2_INF, B2_INF, B2_INF}, cmax[6] = {-B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF};
    if (i < n) {
        const uint32_t e = __ldg(col_entity + i);
        float M[16];
        const float4* pw = world + 4 * (size_t)e;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float4 v = pw[c];
            M[4 * c] = v.x; M[4 * c + 1] = v.y; M[4 * c + 2] = v.z; M[4 * c + 3] = v.w;
        }
        refit_one(i, e, M, load_local_box(col_local, i), world_rec, cmin, cmax);
    }
    bounds_reduce(cmin, cmax, bounds);
}

// Sparse TRS update (b2col_commit_trs_rows): rows ids[i] of the entity-indexed SoA take the packed values i.  One thread per row.
__global__ void __launch_bounds__(256)
k_scatter_trs(uint32_t n, const uint32_t* __restrict__ ids, const float* __restrict__ pos3, const float4* __restrict__ rot4, const float* __restrict__ scl3,
              float* d_pos, float4* d_rot, float* d_scl)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t e = ids[i];
    d_rot[e] = rot4[i];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        d_pos[3 * e + k] = pos3[3 * (size_t)i + k];
        d_scl[3 * e + k] = scl3[3 * (size_t)i + k];
    }
}

// world_matrix write-back of the rows that can have moved (b2col_fetch_world_rows): packed[i] = world[ids[i]].  Four threads per row.
__global__ void __launch_bounds__(256)
k_gather_world(uint32_t n, const uint32_t* __restrict__ ids, const float4* __restrict__ world, float4* __restrict__ packed)
{
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 4 * (size_t)n) return;
    packed[t] = world[4 * (size_t)ids[t >> 2] + (t & 3)];
}

// key-width statistic (b2col_update's policy): k_build counts the leaves whose key equals the key KEY_RUN sorted places before
// them — the members of runs the key does not separate — into KEY_CTR_SLOTS counter pairs, one 32-byte sector each (a block adds
// to slot blockIdx % KEY_CTR_SLOTS; scenes the keys resolve add nothing at all); the next frame's k_morton sums and clears them
constexpr int KEY_RUN = 4;
constexpr uint32_t KEY_CTR_SLOTS = 64;

// Header fields every build needs besides the bounds: leaf / node counts and the quantisation grid of the
// traversal nodes, derived from the box bounds k_refit reduced into the header.  One thread.
__device__ void finish_header(BlobHeader* hdr, const uint32_t* bounds, uint32_t n, float T)
{
    hdr->magic = BLOB_MAGIC;
    hdr->n_leaves = n;
    hdr->n_nodes = n > 1u ? n - 1u : 1u;
    float bmin[3], bmax[3];
    for (int k = 0; k < 3; ++k) {
        hdr->bounds_min_inv[k] = bounds[k];
        hdr->bounds_max[k] = bounds[3 + k];
        hdr->box_min_inv[k] = bounds[6 + k];
        hdr->box_max[k] = bounds[9 + k];
        bmin[k] = ordered_to_float(~bounds[6 + k]);
        bmax[k] = ordered_to_float(bounds[9 + k]);
    }
    hdr->grid = make_grid(bmin, bmax, T);
}

// Morton keys + primitive indices + the digit histograms of every radix pass (the keys are in
// registers here anyway).  The bit allocation (make_morton_plan) depends on the centroid bounds the refit has
// just reduced: thread 0 of every block derives it once and shares it.
#ifndef B2_MORTON_BLOCK
#define B2_MORTON_BLOCK 256
#endif
constexpr int MORTON_BLOCK = B2_MORTON_BLOCK;
__global__ void __launch_bounds__(MORTON_BLOCK)
k_morton(uint32_t n, const LeafRec* __restrict__ world_rec, BlobHeader* hdr, const uint32_t* __restrict__ bounds, int key_bits, int passes, int hist_lo, int packed, float T,
         uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* ghist,  // histograms of the radix passes hist_lo .. passes-1 (the MSD pipeline needs the top digit's only)
         uint32_t* key_ctr, uint4* host_stat)
{
    pdl_wait();
    __shared__ uint32_t shist[RS_MAX_PASSES][RS_RADIX];
    __shared__ MortonPlan s_plan;
    for (int p = hist_lo; p < passes; ++p)
        for (int b = threadIdx.x; b < RS_RADIX; b += MORTON_BLOCK) shist[p][b] = 0u;
    if (threadIdx.x == 0) {
        float cmin[3], cmax[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            cmin[k] = ordered_to_float(~bounds[k]);
            cmax[k] = ordered_to_float(bounds[3 + k]);
        }
        s_plan = make_morton_plan(cmin, cmax, key_bits);
    }
    if (blockIdx.x == 0 && threadIdx.x == 32) {
        // (beside thread 0's plan arithmetic) key-width statistic of the PREVIOUS full build — k_build's counters; hdr still describes that
        // build until the end of this kernel — and this frame's centroid bounds: 16-byte stores into page-locked host memory, read by the
        // host's policies before a later build (b2col_update); the counters start this frame at zero
        uint32_t dc = 0u, dn = 0u;
        for (uint32_t sl = 0; sl < KEY_CTR_SLOTS; ++sl) {
            dc += __ldcg(key_ctr + 8u * sl);
            dn += __ldcg(key_ctr + 8u * sl + 1u);
        }
        for (uint32_t sl = 0; sl < KEY_CTR_SLOTS; ++sl) *reinterpret_cast<uint2*>(key_ctr + 8u * sl) = make_uint2(0u, 0u);
        if (host_stat) {
            if (hdr->magic == BLOB_MAGIC) host_stat[0] = make_uint4(dc, dn, hdr->n_leaves, hdr->key_bits);
            host_stat[1] = make_uint4(~bounds[0], ~bounds[1], ~bounds[2], n);
            host_stat[2] = make_uint4(bounds[3], bounds[4], bounds[5], n);
        }
    }
    __syncthreads();
    const MortonPlan plan = s_plan;
    // A modest grid (4 blocks per SM: every block ends with up to 1024 atomics on the same 1024 global counters, and 8 blocks
    // per SM spent 4 us more there) whose threads keep two records in flight.
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 2u * stride) {
        const uint32_t j = i + stride;
        const bool two = j < n;
        const float4* bp = reinterpret_cast<const float4*>(world_rec + i);
        const float4* cp = reinterpret_cast<const float4*>(world_rec + (two ? j : i));
        const float4 b0 = bp[0], b1 = bp[1], c0 = cp[0], c1 = cp[1];
        {
            const float lo[3] = {b0.x, b0.y, b0.z}, hi[3] = {b0.w, b1.x, b1.y};
            const uint64_t key = morton_key(lo, hi, plan);
            if (packed) {  // keys of at most 32 bits travel through the sort as key << 32 | index (k_onesweep MODE 1 / 2)
                keys[i] = (key << 32) | i;
            }
            else {
                keys[i] = key;
                vals[i] = i;
            }
            for (int p = hist_lo; p < passes; ++p) atomicAdd(&shist[p][(uint32_t)(key >> (8 * p)) & 255u], 1u);
        }
        if (two) {
            const float lo[3] = {c0.x, c0.y, c0.z}, hi[3] = {c0.w, c1.x, c1.y};
            const uint64_t key = morton_key(lo, hi, plan);
            if (packed) {  // keys of at most 32 bits travel through the sort as key << 32 | index (k_onesweep MODE 1 / 2)
                keys[j] = (key << 32) | j;
            }
            else {
                keys[j] = key;
                vals[j] = j;
            }
            for (int p = hist_lo; p < passes; ++p) atomicAdd(&shist[p][(uint32_t)(key >> (8 * p)) & 255u], 1u);
        }
    }
    __syncthreads();
    for (int p = hist_lo; p < passes; ++p) {
        for (int b = threadIdx.x; b < RS_RADIX; b += MORTON_BLOCK) {
            const uint32_t c = shist[p][b];
            if (c) atomicAdd(&ghist[p * RS_RADIX + b], c);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        finish_header(hdr, bounds, n, T);
        hdr->key_bits = (uint32_t)key_bits;
    }
}

// ---- the frame's first kernel(s) when the Morton plan is known BEFORE the refit: propagate + refit + keys + histograms ---------
// k_morton exists because the exact plan needs the centroid bounds of ALL boxes of the frame; but a plan only steers tree
// topology (any keys give a valid tree), so a frame may just as well use a plan made from the bounds of an EARLIER frame
// (b2col_update's plan policy: bounds reported by the previous frames, widened by 1/32, kept while the scene stays inside
// them and fills at least half of them; centroids outside clamp to the edge cells).  The key is then computed while the world
// box is still in registers: one kernel and a 32-byte re-read per collider less per frame.
//
// Grid-stride form (a few blocks per SM, every thread folds several entities): the twelve-value bounds reduction — 120 of
// the ~390 instructions a one-entity thread executes — and the histogram flush run once per thread / block, not per entity.
// The LAST block of the last level's launch to finish (ticket through tail.done) finds the frame's bounds complete and does
// what thread 0 of k_morton does: statistics of the previous build, header, quantisation grid.
struct KeyGen {
    MortonPlan plan;
    int passes, hist_lo, packed;
};
struct FrameTail {
    BlobHeader* hdr;
    uint32_t* key_ctr;
    uint4* host_stat;       // page-locked, mapped: [0] key statistic of the previous build, [1] / [2] centroid bounds of this frame
    uint32_t* done;         // zero when the frame starts (per-frame scratch)
    uint32_t n;             // colliders of the frame
    float T;
    int key_bits;           // > 0: a full build with keys of this width (hdr->key_bits, report the previous build's statistic)
    int last_level;         // this launch completes the frame's bounds
    uint32_t plan_min_inv[3], plan_max[3];  // KEYS: what the header records as the Morton bounds (ordered encoding) — the plan's, not the frame's
};

__device__ __noinline__ void frame_tail(const FrameTail& t, const uint32_t* bounds_in, bool keys)  // not inlined: its IEEE divisions (make_grid) are the only FFMA-containing code near the refit arithmetic, and tests/test_capi_cpu.py keeps them apart
{
    uint32_t b[BOUNDS_WORDS];
    for (int k = 0; k < BOUNDS_WORDS; ++k) b[k] = __ldcg(bounds_in + k);
    uint32_t dc = 0u, dn = 0u;
    for (uint32_t sl = 0; sl < KEY_CTR_SLOTS; ++sl) {
        dc += __ldcg(t.key_ctr + 8u * sl);
        dn += __ldcg(t.key_ctr + 8u * sl + 1u);
    }
    for (uint32_t sl = 0; sl < KEY_CTR_SLOTS; ++sl) *reinterpret_cast<uint2*>(t.key_ctr + 8u * sl) = make_uint2(0u, 0u);
    if (t.host_stat) {
        if (t.key_bits > 0 && t.hdr->magic == BLOB_MAGIC) t.host_stat[0] = make_uint4(dc, dn, t.hdr->n_leaves, t.hdr->key_bits);
        t.host_stat[1] = make_uint4(~b[0], ~b[1], ~b[2], t.n);  // ordered(min centroid) / ordered(max centroid) of THIS frame: the plan policy's input
        t.host_stat[2] = make_uint4(b[3], b[4], b[5], t.n);
    }
    if (keys) {
        for (int k = 0; k < 3; ++k) {
            b[k] = t.plan_min_inv[k];
            b[3 + k] = t.plan_max[k];
        }
    }
    else if (t.key_bits == 0 && t.hdr->magic == BLOB_MAGIC) {  // refit-only frame: the keys of the tree are still those of its build
        for (int k = 0; k < 3; ++k) {
            b[k] = t.hdr->bounds_min_inv[k];
            b[3 + k] = t.hdr->bounds_max[k];
        }
    }
    finish_header(t.hdr, b, t.n, t.T);
    if (t.key_bits > 0) t.hdr->key_bits = (uint32_t)t.key_bits;
}

#ifndef B2_RK_BLOCK
#define B2_RK_BLOCK 256
#endif
constexpr int RK_BLOCK = B2_RK_BLOCK;
// resident blocks per SM = grid / SMs (persistent).  Measured, update at 1M flat / 3-level scene / 10M: 4 -> 135.8 / 163.8 / 1088 us,
// 3 -> 135.6 / 157.6, 2 -> 132.8 / 155.4 / 1065 (kept), 1 -> 139.4 / 159.9 (profiles/r2_experiments.txt): every block flushes up to 768 histogram bins and twelve
// bounds with global atomics, and the small levels of a hierarchy want few blocks; below two per SM the stream starves.
#ifndef B2_RK_MIN_BLOCKS
#define B2_RK_MIN_BLOCKS 2
#endif
template <bool DENSE, bool KEYS>
__global__ void __launch_bounds__(RK_BLOCK, B2_RK_MIN_BLOCKS)
k_refit_keys(const uint32_t* __restrict__ level_entities, const uint32_t* __restrict__ level_col, uint32_t e0, uint32_t c0, uint32_t count, int root_level,
             const uint32_t* __restrict__ parent, const float* __restrict__ pos3, const float4* __restrict__ rot4, const float* __restrict__ scl3, float4* world,
             const Aabb24* __restrict__ col_local, LeafRec* __restrict__ world_rec, uint32_t* bounds, KeyGen kg, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
             uint32_t* ghist, FrameTail tail)
{
    pdl_wait();
    __shared__ uint32_t shist[KEYS ? RS_MAX_PASSES : 1][RS_RADIX];
    __shared__ bool s_last;
    if (KEYS) {
        for (int p = kg.hist_lo; p < kg.passes; ++p)
            for (int d = threadIdx.x; d < RS_RADIX; d += RK_BLOCK) shist[p][d] = 0u;
        __syncthreads();
    }
    float cmin[6] = {B2_INF, B2_INF, B2_INF, B2_INF, B2_INF, B2_INF}, cmax[6] = {-B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF};
    const uint32_t stride = gridDim.x * RK_BLOCK;
    for (uint32_t i = blockIdx.x * RK_BLOCK + threadIdx.x; i < count; i += stride) {
        const uint32_t e = DENSE ? e0 + i : level_entities[i];
        const uint32_t c = DENSE ? c0 + i : level_col[i];
        LocalBox lb;
        lb.a = lb.b = lb.c = make_float2(0.0f, 0.0f);
        if (c != NO_COLLIDER) lb = load_local_box(col_local, c);
        float W[16];
        entity_world(e, parent, pos3, rot4, scl3, world, W, root_level != 0);
        float4* ow = world + 4 * (size_t)e;
#pragma unroll
        for (int k = 0; k < 4; k += 2)
            store32(ow + k, make_float4(W[4 * k], W[4 * k + 1], W[4 * k + 2], W[4 * k + 3]), make_float4(W[4 * k + 4], W[4 * k + 5], W[4 * k + 6], W[4 * k + 7]));
        if (c != NO_COLLIDER) {
            const Box b = refit_one(c, e, W, lb, world_rec, cmin, cmax);
            if (KEYS) {
                const uint64_t key = morton_key(b.lo, b.hi, kg.plan);
                if (kg.packed) {
                    keys[c] = (key << 32) | c;
                }
                else {
                    keys[c] = key;
                    vals[c] = c;
                }
                for (int p = kg.hist_lo; p < kg.passes; ++p) atomicAdd(&shist[p][(uint32_t)(key >> (8 * p)) & 255u], 1u);
            }
        }
    }
    bounds_reduce<RK_BLOCK>(cmin, cmax, bounds);
    if (KEYS) {
        __syncthreads();
        for (int p = kg.hist_lo; p < kg.passes; ++p)
            for (int d = threadIdx.x; d < RS_RADIX; d += RK_BLOCK) {
                const uint32_t cnt = shist[p][d];
                if (cnt) atomicAdd(&ghist[p * RS_RADIX + d], cnt);
            }
    }
    if (!tail.last_level) return;
    // The last block to get here sees every block's bounds: the block's atomics (threads < 6, bounds_reduce) are ordered before the
    // barrier, thread 0's fence after it makes them — cumulatively — visible at GPU scope before its ticket.  (A fence in EVERY
    // thread made each warp wait for the acknowledgement of all of its own stores before it could exit: MEMBAR + ERRBAR + CCTL.IVALL
    // were 20 % of this kernel's stall samples, ncu source page.)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(tail.done, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        frame_tail(tail, bounds, KEYS);
    }
}

// ---- the same kernel for DENSE levels, inputs moved by the TMA unit -------------------------------------------------------------
// On a dense level (entities e0 + i, colliders c0 + i) everything a tile of RK_BLOCK entities reads is five contiguous byte
// ranges: rotations (16 B / entity), positions and scales (12 B), parent IDs (4 B) and local boxes (24 B / collider).  One thread
// asks the bulk-copy engine (cp.async.bulk global -> shared, completion counted on an mbarrier) for the NEXT tile's ranges while
// the block computes the current one out of shared memory: two stages, no registers spent on prefetching, no load instructions
// in the compute warps besides the parent-matrix gather of non-root levels.  A grid-stride kernel whose threads load, compute
// and store in lock step leaves the memory system idle while it computes; with the ranges in flight a tile ahead it does not
// (measured at 1M colliders: profiles/r2_experiments.txt).
// cp.async.bulk wants 16-byte aligned addresses and sizes: a range starts at the 16-byte boundary below its first byte (`lead`
// bytes early) and is rounded up (<= 30 bytes of neighbouring rows; every array carries allocation slack, see ensure()).
namespace bulk {
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred P1;\n\tB2_MBAR_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra B2_MBAR_DONE;\n\tbra B2_MBAR_WAIT;\n\tB2_MBAR_DONE:\n\t}" ::"r"(
            smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void g2s(void* sdst, const void* gsrc16, uint32_t bytes16, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sdst)), "l"(gsrc16), "r"(bytes16),
                 "r"(smem_u32(bar))
                 : "memory");
}
// the 16-byte aligned range that covers [p, p + bytes): lead = bytes between its start and p
__device__ __forceinline__ uint32_t lead_of(const void* p) { return (uint32_t)(reinterpret_cast<uintptr_t>(p) & 15u); }
__device__ __forceinline__ uint32_t span_of(const void* p, uint32_t bytes) { return (lead_of(p) + bytes + 15u) & ~15u; }
}  // namespace bulk

// One pipeline per block (tile = its 256 entities, a __syncthreads per tile).  B2_RK_WARP_TILES: one pipeline PER WARP (tile = the
// warp's 32 entities, its own pair of mbarriers, no block barrier inside the loop, so that the warps drift apart) — measured equal
// (136.9 vs 136.1 us per update, profiles/r2_experiments.txt): this kernel does not wait for its inputs either way.
#ifdef B2_RK_WARP_TILES
constexpr int RK_TILE = 32;
constexpr int RK_PIPES = RK_BLOCK / 32;
#else
constexpr int RK_TILE = RK_BLOCK;
constexpr int RK_PIPES = 1;
#endif
struct alignas(16) RkStage {
    unsigned char rot[RK_TILE * 16];
    unsigned char pos[RK_TILE * 12 + 32];
    unsigned char scl[RK_TILE * 12 + 32];
    unsigned char box[RK_TILE * 24 + 32];
    unsigned char par[RK_TILE * 4 + 32];
};
static_assert(sizeof(RkStage) % 16 == 0 && (RK_TILE * 12) % 16 == 0 && (RK_TILE * 24) % 16 == 0 && (RK_TILE * 4) % 16 == 0, "stage ranges are 16-byte multiples");

template <bool KEYS>
__global__ void __launch_bounds__(RK_BLOCK, B2_RK_MIN_BLOCKS)
k_refit_keys_bulk(uint32_t e0, uint32_t c0, uint32_t count, int root_level, const uint32_t* __restrict__ parent, const float* __restrict__ pos3,
                  const float4* __restrict__ rot4, const float* __restrict__ scl3, float4* world, const Aabb24* __restrict__ col_local, LeafRec* __restrict__ world_rec,
                  uint32_t* bounds, KeyGen kg, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* ghist, FrameTail tail)
{
    __shared__ RkStage stage[2][RK_PIPES];
    __shared__ alignas(8) uint64_t full[2][RK_PIPES];
    __shared__ uint32_t shist[KEYS ? RS_MAX_PASSES : 1][RS_RADIX];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    const int pipe = RK_PIPES == 1 ? 0 : tid >> 5;        // which pipeline this thread reads from
    const int slot = RK_PIPES == 1 ? tid : (tid & 31);    // its entity inside the tile
    const bool issuer = slot == 0;                        // the thread that requests the pipeline's copies
    if (tid == 0) {
        for (int st = 0; st < 2; ++st)
            for (int w = 0; w < RK_PIPES; ++w) bulk::mbar_init(&full[st][w], 1u);
        bulk::mbar_init_fence();
    }
    if (KEYS) {
        for (int p = kg.hist_lo; p < kg.passes; ++p)
            for (int d = tid; d < RS_RADIX; d += RK_BLOCK) shist[p][d] = 0u;
    }
    __syncthreads();
    pdl_wait();  // the arrays below are written by work earlier in the stream
    const uint32_t n_tiles = (count + RK_TILE - 1u) / RK_TILE;
    const uint32_t t_stride = gridDim.x * RK_PIPES;
    // ranges of tile t (every thread can name them; the pipeline's issuer requests them)
    auto request = [&](uint32_t t, int st) {
        const uint32_t i0 = t * RK_TILE, valid = min((uint32_t)RK_TILE, count - i0);
        const size_t e = (size_t)e0 + i0, c = (size_t)c0 + i0;
        const void* g_rot = rot4 + e;
        const void* g_pos = pos3 + 3 * e;
        const void* g_scl = scl3 + 3 * e;
        const void* g_box = col_local + c;
        const void* g_par = parent + e;
        const uint32_t b_rot = valid * 16u, b_pos = bulk::span_of(g_pos, valid * 12u), b_scl = bulk::span_of(g_scl, valid * 12u), b_box = bulk::span_of(g_box, valid * 24u),
                       b_par = root_level ? 0u : bulk::span_of(g_par, valid * 4u);
        uint64_t* bar = &full[st][pipe];
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the stage's earlier reads (generic proxy) before the copies' writes (async proxy)
        bulk::mbar_arrive_expect_tx(bar, b_rot + b_pos + b_scl + b_box + b_par);
        RkStage& sg = stage[st][pipe];
        bulk::g2s(sg.rot, g_rot, b_rot, bar);
        bulk::g2s(sg.pos, static_cast<const char*>(g_pos) - bulk::lead_of(g_pos), b_pos, bar);
        bulk::g2s(sg.scl, static_cast<const char*>(g_scl) - bulk::lead_of(g_scl), b_scl, bar);
        bulk::g2s(sg.box, static_cast<const char*>(g_box) - bulk::lead_of(g_box), b_box, bar);
        if (!root_level) bulk::g2s(sg.par, static_cast<const char*>(g_par) - bulk::lead_of(g_par), b_par, bar);
    };
    float cmin[6] = {B2_INF, B2_INF, B2_INF, B2_INF, B2_INF, B2_INF}, cmax[6] = {-B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF, -B2_INF};
    uint32_t t = blockIdx.x * RK_PIPES + (uint32_t)pipe;
    if (issuer && t < n_tiles) request(t, 0);
    for (uint32_t k = 0; t < n_tiles; t += t_stride, ++k) {
        const int st = (int)(k & 1u);
        if (issuer && t + t_stride < n_tiles) request(t + t_stride, st ^ 1);  // that stage was last read in iteration k - 1, which ended with a barrier of the pipeline's threads
        bulk::mbar_wait(&full[st][pipe], (k >> 1) & 1u);
        const uint32_t i0 = t * RK_TILE, i = i0 + (uint32_t)slot;
        if (i < count) {
            const RkStage& sg = stage[st][pipe];
            const size_t ef = (size_t)e0 + i0, cf = (size_t)c0 + i0;
            const uint32_t e = e0 + i, c = c0 + i;
            const float4 q4 = *reinterpret_cast<const float4*>(sg.rot + 16 * slot);
            const float* sp = reinterpret_cast<const float*>(sg.pos + bulk::lead_of(pos3 + 3 * ef) + 12 * slot);
            const float* ss = reinterpret_cast<const float*>(sg.scl + bulk::lead_of(scl3 + 3 * ef) + 12 * slot);
            const float2* sb = reinterpret_cast<const float2*>(sg.box + bulk::lead_of(col_local + cf) + 24 * slot);
            const float p[3] = {sp[0], sp[1], sp[2]}, sc[3] = {ss[0], ss[1], ss[2]};
            LocalBox lb;
            lb.a = sb[0];
            lb.b = sb[1];
            lb.c = sb[2];
            const uint32_t par = root_level ? 0u : *reinterpret_cast<const uint32_t*>(sg.par + bulk::lead_of(parent + ef) + 4 * slot);
            float W[16];
            entity_world_from(p, q4, sc, par, world, W);
            float4* ow = world + 4 * (size_t)e;
#pragma unroll
            for (int j = 0; j < 4; j += 2)
                store32(ow + j, make_float4(W[4 * j], W[4 * j + 1], W[4 * j + 2], W[4 * j + 3]), make_float4(W[4 * j + 4], W[4 * j + 5], W[4 * j + 6], W[4 * j + 7]));
            const Box b = refit_one(c, e, W, lb, world_rec, cmin, cmax);
            if (KEYS) {
                const uint64_t key = morton_key(b.lo, b.hi, kg.plan);
                if (kg.packed) {
                    keys[c] = (key << 32) | c;
                }
                else {
                    keys[c] = key;
                    vals[c] = c;
                }
                for (int pp = kg.hist_lo; pp < kg.passes; ++pp) atomicAdd(&shist[pp][(uint32_t)(key >> (8 * pp)) & 255u], 1u);
            }
        }
        // the stage is free for the request of iteration k + 1 once every thread of the pipeline has read it
        if (RK_PIPES == 1) __syncthreads();
        else __syncwarp();
    }
    if (RK_PIPES != 1) __syncthreads();  // the warps ran at their own pace: meet before the block-level reductions
    bounds_reduce<RK_BLOCK>(cmin, cmax, bounds);
    if (KEYS) {
        __syncthreads();
        for (int p = kg.hist_lo; p < kg.passes; ++p)
            for (int d = tid; d < RS_RADIX; d += RK_BLOCK) {
                const uint32_t cnt = shist[p][d];
                if (cnt) atomicAdd(&ghist[p * RS_RADIX + d], cnt);
            }
    }
    if (!tail.last_level) return;
    __syncthreads();  // (see k_refit_keys: one fence, in the thread that takes the ticket)
    if (tid == 0) {
        __threadfence();
        s_last = atomicAdd(tail.done, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (s_last && tid == 0) {
        __threadfence();
        frame_tail(tail, bounds, KEYS);
    }
}

// b2col_comm_replicate_records: the scene bounds a replica needs for its Morton / quantisation grids arrive inside the
// source rank's blob header; this puts them where k_morton reads them.
__global__ void k_bounds_from_header(const BlobHeader* hdr, uint32_t* bounds)
{
    pdl_wait();
    const int k = threadIdx.x;
    if (k < 3) {
        bounds[k] = hdr->bounds_min_inv[k];
        bounds[3 + k] = hdr->bounds_max[k];
        bounds[6 + k] = hdr->box_min_inv[k];
        bounds[9 + k] = hdr->box_max[k];
    }
}

// Refit-only frames (b2col_refit) skip k_morton: this fills the header fields k_morton would have written,
// from the bounds k_refit has just reduced.
__global__ void k_header(uint32_t n, BlobHeader* hdr, const uint32_t* bounds, float T, uint32_t* key_ctr)
{
    pdl_wait();
    if (threadIdx.x < KEY_CTR_SLOTS) *reinterpret_cast<uint2*>(key_ctr + 8u * threadIdx.x) = make_uint2(0u, 0u);  // the re-bounding pass counts the same keys again
    if (threadIdx.x == 0) finish_header(hdr, bounds, n, T);
}

// LBVH hierarchy + bounds, production form: traversal structure only, merged in grid space — the per-thread
// climb of build_from_leaf_q (b2col_bvh.cuh), with the arrival exchange moved into SHARED memory for every
// node whose whole leaf range lies inside this block's BUILD_BLOCK consecutive leaves (>95 % of all nodes);
// only the few subtrees that straddle a block boundary (~16 per block) meet through the 128-bit global
// exchange, which keeps 2M atomics per frame out of L2.  Measured decomposition at 1M (variant builds, ncu):
// leaf phase (gathers by primitive index + leaf records) 26 us, climb logic 35 us (issue-bound: the loop runs
// at ~7 of 32 lanes, half of the climbers retire per level), node stores 7 us.
//
// Which nodes are block-local must be decided identically by both children, from the node alone: node p
// (the split between sorted positions p and p+1) spans the leaves between the nearest HIGHER splits on either
// side, so it is local iff a higher split exists on both sides within the block, counting the block's two
// boundary splits: prefix-max(level)[p-1] > level[p] < suffix-max(level)[p+1].
// (Measured and dropped, profiles/r1_d_experiments.txt: writing the leaf records from the last radix pass, where the
// gathers are independent — the pass grew by 14 us, this kernel shrank by 4: here the gather hides under the climb.)
#ifndef B2_BUILD_BLOCK
#define B2_BUILD_BLOCK 128
#endif
// the one thread that completes the root: header fields of the finished tree
__device__ __noinline__ void build_finish_root(BlobHeader* hdr, uint32_t root, bool wild)
{
    hdr->root = root;
    hdr->trace_root = root;  // k_top_layout, when it runs, redirects the traversal to its compact copy
    hdr->top_count = 0u;
    hdr->has_wild = wild ? 1u : 0u;
}
#ifndef B2_BUILD_COMPACT
#define B2_BUILD_COMPACT 1  // (measured at 1M colliders: 0 -> 138.6, 1 -> 136.0, 2 -> 137.2, 3 -> 138.5 us per update) levels after which the block's surviving climbers are packed into its first warps (0: never)
#endif
constexpr int BUILD_BLOCK = B2_BUILD_BLOCK;  // update at 1M: 64 -> 170.4 us, 96 -> 168.4, 128 -> 167.8, 256 -> 170.1, 512 -> 178.9, 1024 -> 204.0
__global__ void __launch_bounds__(BUILD_BLOCK)
k_build(uint32_t n, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ sorted_prim, const LeafRec* __restrict__ world_rec, Xch* xch, QNode* qnodes, LeafRec* leaves, BlobHeader* hdr, uint32_t* scratch, uint32_t scratch_words,
        uint32_t* key_ctr)  // KEY_CTR_SLOTS x 8 words, one 32-byte sector per slot: [0] / [1] = leaves in runs of equal keys / of keys equal above their low 8 bits
{
    pdl_wait();
    // The frame's zero-initialised scratch (scene bounds, digit histograms, tile tickets, look-back words) has been
    // consumed by the kernels before this one: wipe it here for the next frame instead of a memset per frame.
    for (uint32_t w = blockIdx.x * BUILD_BLOCK + threadIdx.x; w < scratch_words; w += gridDim.x * BUILD_BLOCK) scratch[w] = 0u;
    __shared__ uint32_t s_lev[BUILD_BLOCK + 1];     // s_lev[j] = level of split b0-1+j (j = 0: left boundary, j = BUILD_BLOCK: right boundary)
    __shared__ uint32_t s_wmax[2][BUILD_BLOCK / 32];
    __shared__ uint8_t s_local[BUILD_BLOCK + 1];
    __shared__ Xch s_xch[BUILD_BLOCK + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t b0 = blockIdx.x * BUILD_BLOCK;
    const uint32_t i = b0 + (uint32_t)tid;
    const bool valid = i < n;
    const Grid grid = hdr->grid;  // written by k_morton

    // ---- leaf record + grid box (gathers by primitive index) and this leaf's left split level
    QBox box;
    box.w[0] = box.w[1] = box.w[2] = 0u;
    uint64_t key = 0ull;
    uint32_t lev_l = LEVEL_END;
    bool wild = false;  // this subtree holds a leaf the grid cannot bound (b2col_bvh.cuh, WILD_BIT)
    bool dup_cur = false, dup_narrow = false;
    if (valid) {
        const LeafRec rec = DevOps::load_leaf(&world_rec[sorted_prim[i]]);  // one 256-bit load: the world record IS the leaf record
        DevOps::store_leaf(&leaves[i], rec);
        box = quantise_box(rec.lo, rec.hi, grid);
        wild = box_is_wild(rec.lo, rec.hi);
        key = keys[i];
        if (i > 0u) lev_l = split_level(keys[i - 1u], key, i - 1u);
        if (i >= (uint32_t)KEY_RUN) {  // sorted: equal to the key KEY_RUN places back = inside a run of more than KEY_RUN leaves the key does not separate
            const uint64_t x = keys[i - (uint32_t)KEY_RUN] ^ key;
            dup_cur = x == 0ull;
            dup_narrow = (x >> 8) == 0ull;  // ... and which 8 bits fewer (one radix pass) would not separate
        }
    }
    if (n == 1u) {  // a lone leaf still gets a node; both slots = the leaf (an empty slot has no grid encoding)
        if (i == 0u) {
            QNode nd;
#pragma unroll
            for (int k = 0; k < 3; ++k) nd.w[k] = nd.w[3 + k] = box.w[k];
            nd.w[6] = LEAF_BIT | (wild ? WILD_BIT : 0u) | 0u;
            nd.w[7] = nd.w[6];
            DevOps::store_qnode(&qnodes[0], nd);
            hdr->root = 0u;
            hdr->trace_root = 0u;
            hdr->top_count = 0u;
            hdr->has_wild = wild ? 1u : 0u;
        }
        return;
    }
    s_lev[tid] = lev_l;
    *reinterpret_cast<uint4*>(&s_xch[tid]) = make_uint4(0u, 0u, 0u, UNVISITED);
    if (tid == BUILD_BLOCK - 1) {
        s_lev[BUILD_BLOCK] = (valid && i + 1u < n) ? split_level(key, keys[i + 1u], i) : LEVEL_END;
        *reinterpret_cast<uint4*>(&s_xch[BUILD_BLOCK]) = make_uint4(0u, 0u, 0u, UNVISITED);
    }
    // ---- prefix / suffix maxima of the split levels (warp shuffles, then across the 8 warps)
    uint32_t pm = lev_l, sx = lev_l;  // inclusive max over s_lev[0..tid] / s_lev[tid..BUILD_BLOCK-1]
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, pm, o), c = __shfl_down_sync(0xffffffffu, sx, o);
        if (lane >= o) pm = max(pm, a);
        if (lane + o < 32) sx = max(sx, c);
    }
    if (lane == 31) s_wmax[0][warp] = pm;
    if (lane == 0) s_wmax[1][warp] = sx;
    const int n_dup_narrow = __syncthreads_count(dup_narrow);  // the two barriers of this prologue also count the block's key collisions
    {
        const uint32_t right_boundary = s_lev[BUILD_BLOCK];
        uint32_t before = 0u, after = right_boundary;
#pragma unroll
        for (int w = 0; w < BUILD_BLOCK / 32; ++w) {
            if (w < warp) before = max(before, s_wmax[0][w]);
            if (w > warp) after = max(after, s_wmax[1][w]);
        }
        // exclusive maxima: strictly left of split tid / strictly right of it
        const uint32_t pm_in = __shfl_up_sync(0xffffffffu, pm, 1), sx_in = __shfl_down_sync(0xffffffffu, sx, 1);
        const uint32_t left_max = max(before, lane > 0 ? pm_in : 0u);
        const uint32_t right_max = max(after, lane < 31 ? sx_in : 0u);
        // split j = tid is node p = b0 - 1 + tid; j = 0 is the left boundary split (belongs to the neighbours too): never local
        s_local[tid] = (tid > 0 && left_max > lev_l && right_max > lev_l) ? 1 : 0;
        if (tid == 0) s_local[BUILD_BLOCK] = 0;
    }
    const int n_dup_cur = __syncthreads_count(dup_cur);
    if (tid == 0) {  // rare in scenes the keys resolve; spread over KEY_CTR_SLOTS sectors of their own where they do not
        uint32_t* slot = key_ctr + 8u * (blockIdx.x & (KEY_CTR_SLOTS - 1u));
        if (n_dup_cur) atomicAdd(slot, (uint32_t)n_dup_cur);
        if (n_dup_narrow) atomicAdd(slot + 1, (uint32_t)n_dup_narrow);
    }

    // ---- climb
    // The loop below is bound by instruction issue and half of the climbers retire at every level: after two levels a warp
    // executes the ~110-instruction step for a quarter of its lanes.  So the block's surviving climbers are packed into its
    // first warps after each of the first BUILD_COMPACT levels (their state — range, outer levels, box, ref: 32 bytes — moves
    // through shared memory; who carries a climber is irrelevant to the arrival protocol): 128 leaves then cost 4 + 2 + 1 + 1 + ...
    // warp-steps instead of 4 + 4 + 4 + 4 + ...  Same exchanges, same stores, same tree.
    uint32_t l = i, r = i;
    uint32_t lev_r = valid ? s_lev[tid + 1] : LEVEL_END;
    uint32_t ref = LEAF_BIT | i;
    bool active = valid;
    auto step = [&]() {  // one level; clears `active` when this climber parks (first to arrive) or has finished the root
        const bool is_left = lev_r < lev_l;  // join the neighbour beyond the LOWER split; array ends never win
        const uint32_t p = is_left ? r : l - 1u;
        const uint32_t j = p - (b0 - 1u);  // split index inside the block (wraps to a huge value left of the block)
        const bool local = j <= (uint32_t)BUILD_BLOCK && s_local[j] != 0;
        QBox sib;
        uint32_t bound, level;
        bool sib_wild;
        DevOps::store_ref(&qnodes[p], is_left ? 0 : 1, ref | (wild ? WILD_BIT : 0u));
        if (local) {
            // 128-bit exchange in shared memory (ATOMS.EXCH.128): the message is the atomic's operand, no fences.  A block-local node's
            // outer splits lie inside the block (or are its boundary splits), so the level beyond the sibling's far bound is read from
            // s_lev instead of travelling in the message: no packing of the 7 level bits into / out of the boxes' flag bits — 24 of the
            // ~120 instructions of a step
            Xch mine;
            mine.w[0] = box.w[0];
            mine.w[1] = box.w[1];
            mine.w[2] = box.w[2];
            mine.w[3] = (is_left ? l : r) | (wild ? WILD_BIT : 0u);  // < 2^31: never UNVISITED
            const Xch got = DevOps::exchange_shared(&s_xch[j], mine);
            if (got.w[3] == UNVISITED) {  // first to arrive: the sibling's climber finishes this node
                active = false;
                return;
            }
            sib.w[0] = got.w[0];
            sib.w[1] = got.w[1];
            sib.w[2] = got.w[2];
            bound = got.w[3] & REF_INDEX_MASK;
            sib_wild = (got.w[3] & WILD_BIT) != 0u;
            level = s_lev[is_left ? bound - (b0 - 1u) : bound - b0];  // split `bound` (right of the new range) / `bound - 1` (left of it)
        }
        else {
            const Xch mine = xch_pack(box, is_left ? l : r, is_left ? lev_l : lev_r, wild);
            const Xch got = DevOps::exchange(&xch[p], mine);
            if (got.w[3] == UNVISITED) {
                active = false;
                return;
            }
            *reinterpret_cast<uint4*>(&xch[p]) = make_uint4(0u, 0u, 0u, UNVISITED);  // self-cleaning: ready for the next frame
            xch_unpack(got, sib, bound, level, sib_wild);
        }
        wild = wild || sib_wild;
        DevOps::store_qboxes(&qnodes[p], is_left ? box.w : sib.w, is_left ? sib.w : box.w);
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
            build_finish_root(hdr, p, wild);  // (a call: kept out of the loop's instruction stream)
            active = false;
        }
    };
#if B2_BUILD_COMPACT > 0
    __shared__ uint4 s_state[BUILD_BLOCK][2];
    __shared__ uint32_t s_cnt[BUILD_BLOCK / 32];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int round = 0; round < B2_BUILD_COMPACT; ++round) {
        if (active) step();
        const unsigned alive = __ballot_sync(0xffffffffu, active);
        if (lane == 0) s_cnt[warp] = (uint32_t)__popc(alive);
        __syncthreads();
        uint32_t before = 0u, total = 0u;
#pragma unroll
        for (int w = 0; w < BUILD_BLOCK / 32; ++w) {
            const uint32_t c = s_cnt[w];
            if (w < warp) before += c;
            total += c;
        }
        if (active) {
            const uint32_t slot = before + (uint32_t)__popc(alive & lt_mask);
            s_state[slot][0] = make_uint4(l, r, ref, lev_l | (lev_r << 8) | (wild ? 0x10000u : 0u));
            s_state[slot][1] = make_uint4(box.w[0], box.w[1], box.w[2], 0u);
        }
        __syncthreads();
        active = (uint32_t)tid < total;
        if (active) {
            const uint4 a = s_state[tid][0], c = s_state[tid][1];
            l = a.x;
            r = a.y;
            ref = a.z;
            lev_l = a.w & 0xFFu;
            lev_r = (a.w >> 8) & 0xFFu;
            wild = (a.w & 0x10000u) != 0u;
            box.w[0] = c.x;
            box.w[1] = c.y;
            box.w[2] = c.z;
        }
        // (the next round writes s_cnt / s_state only after its own first barrier, which every reader above has then passed)
    }
#endif
    if (!active) return;
    for (;;) {
        step();
        if (!active) return;
    }
}

#if B2_BUILD_BLOCK == 128  // the two measured-and-rejected alternatives assume 128 leaves per block (dev sweeps of B2_BUILD_BLOCK build without them)
#define B2_HAVE_BUILD_VARIANTS 1
#include "b2col_build.inl"  // k_build_direct: the same hierarchy, block-local nodes computed from range queries instead of climbed
#endif

// Compact copy of the top levels of the tree, for trees larger than L2.
//
// A node lives at its split position, so the few thousand nodes of the upper levels — visited by every ray — are
// scattered over the whole node array, one L2 line and one 2 MB page each.  Where they fall is an accident of the
// scene: on BASELINE config 5 (10M clustered colliders, 640 MB tree) the same tree traced at 2090 or 2630 Mrays/s
// depending only on which axis led the Morton code (profiles/r1_e_top_layout.jsonl).  This kernel copies the top
// TOP_LEVELS levels in breadth-first order into a contiguous block behind the leaf records (<= 512 KB) and points
// the traversal at it (hdr->trace_root): 2730 Mrays/s whichever axis leads.  Copies refer to copies down to the
// last copied level, whose child refs are the original ones; the original nodes stay where they are (the export
// and the refit use them).  Node indices are relative to `qnodes` and LeafRec has the size of a QNode, so copy j
// is node (n_nodes + n_leaves + j) and k_trace_lbvh needs no change.
//
// One block: level k's nodes are read (through the frontier of original indices in shared memory), their internal children are
// numbered by a block-wide prefix sum — the breadth-first positions of level k+1 — and the copies are written with
// the remapped refs.  ~1 us per level; run only for trees that do not fit L2 (b2col_update decides).
constexpr int TOP_LEVELS = 14;
constexpr uint32_t TOP_CAP = (1u << TOP_LEVELS) - 1u;
constexpr int TOP_BLOCK = 1024;
constexpr int TOP_ITEMS = 4;
constexpr uint32_t TOP_FRONTIER = 1u << (TOP_LEVELS - 1);  // nodes on the deepest copied level
constexpr size_t TOP_SMEM = 2 * TOP_FRONTIER * sizeof(uint32_t);
__global__ void __launch_bounds__(TOP_BLOCK)
k_top_layout(BlobHeader* hdr, const QNode* qnodes, QNode* top, uint32_t top_base, int levels)
{
    pdl_wait();
    // original indices of the current level's nodes and of the next level's, ping-pong (a level has <= 2^13 nodes)
    extern __shared__ uint32_t s_frontier[];
    __shared__ uint32_t s_warp[TOP_BLOCK / 32];
    __shared__ uint32_t s_total;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* cur_f = s_frontier;
    uint32_t* next_f = s_frontier + TOP_FRONTIER;
    if (tid == 0) cur_f[0] = hdr->root;
    __syncthreads();
    uint32_t off = 0u, count = 1u;  // level k occupies copies [off, off + count)
    for (int k = 0; k < levels && count != 0u; ++k) {
        const bool last = k + 1 == levels;
        uint32_t run = 0u;  // internal children numbered so far on this level
        for (uint32_t base = 0u; base < count; base += TOP_BLOCK * TOP_ITEMS) {
            // TOP_ITEMS consecutive nodes per thread (breadth-first order = thread order, then item order); all loads
            // of a round are in flight together — the level is a chain of DRAM latencies otherwise
            const uint32_t first = base + tid * TOP_ITEMS;
            uint32_t src[TOP_ITEMS];
            QNode nd[TOP_ITEMS];
#pragma unroll
            for (int j = 0; j < TOP_ITEMS; ++j) src[j] = first + j < count ? cur_f[first + j] : 0u;
#pragma unroll
            for (int j = 0; j < TOP_ITEMS; ++j)
                if (first + j < count) nd[j] = DevOps::load_qnode(&qnodes[src[j]]);
            uint32_t mine = 0u;
#pragma unroll
            for (int j = 0; j < TOP_ITEMS; ++j)
                if (first + j < count) mine += ((nd[j].w[6] & LEAF_BIT) ? 0u : 1u) + ((nd[j].w[7] & LEAF_BIT) ? 0u : 1u);
            // block-wide exclusive prefix sum of the internal-child counts
            uint32_t incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) s_warp[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                uint32_t w = s_warp[lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t v = __shfl_up_sync(0xffffffffu, w, o);
                    if (lane >= o) w += v;
                }
                s_warp[lane] = w;
                if (lane == 31) s_total = w;
            }
            __syncthreads();
            uint32_t pre = run + (warp ? s_warp[warp - 1] : 0u) + incl - mine;
            const uint32_t next = off + count;  // first copy of level k+1
#pragma unroll
            for (int j = 0; j < TOP_ITEMS; ++j) {
                if (first + j < count) {
                    if (!last) {
                        if (!(nd[j].w[6] & LEAF_BIT)) {
                            next_f[pre] = nd[j].w[6] & REF_INDEX_MASK;
                            nd[j].w[6] = (nd[j].w[6] & WILD_BIT) | (top_base + next + pre);
                            ++pre;
                        }
                        if (!(nd[j].w[7] & LEAF_BIT)) {
                            next_f[pre] = nd[j].w[7] & REF_INDEX_MASK;
                            nd[j].w[7] = (nd[j].w[7] & WILD_BIT) | (top_base + next + pre);
                            ++pre;
                        }
                    }
                    DevOps::store_qnode(&top[off + first + j], nd[j]);
                }
            }
            run += s_total;
            __syncthreads();  // s_warp / s_total are reused; the next level's frontier is complete after the last round
        }
        off += count;
        count = last ? 0u : run;
        uint32_t* t = cur_f;
        cur_f = next_f;
        next_f = t;
    }
    if (tid == 0) {
        hdr->trace_root = top_base;
        hdr->top_count = off;
    }
}

// Same hierarchy with exact (full-precision) two-slot nodes: only run on demand, as the source of the
// BiTreeNode export (b2col_export_bitreenode).  Writes nothing into the traversal structure.
__global__ void __launch_bounds__(256)
k_build_exact(uint32_t n, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ sorted_prim, const LeafRec* __restrict__ world_aabb,
              const uint32_t* __restrict__ col_entity, Node* nodes, uint32_t* other_bound, const BlobHeader* hdr, uint32_t* root_out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Grid grid = hdr->grid;
    build_from_leaf<DevOps>(i, n, keys, sorted_prim, world_aabb, col_entity, nodes, other_bound, grid, (QNode*)nullptr, (LeafRec*)nullptr, root_out);
}

#include "b2col_trace.inl"  // k_trace_lbvh: the production traversal

// Same query, walking an uploaded reference-format tree with the reference's exact semantics.
__global__ void __launch_bounds__(128)
k_trace_reftopo(const b2col_ray* __restrict__ rays, size_t n_rays, b2col_raycast* __restrict__ out, const uint32_t* __restrict__ words,
                uint32_t n_nodes, float T, uint32_t* overflow_flag)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays) return;
    const RayPre r = load_ray(rays, i);
    const RefHit h = trace_reftopo(r, words, n_nodes, T);
    if (h.overflow) atomicExch(overflow_flag, 1u);
    write_raycast(reinterpret_cast<uint32_t*>(out + i), r, h.hit, h.t, h.side, h.entity);
}

// Results to a page-locked host buffer by SM stores instead of the copy engine (host-buffer pipeline, d2h mode "sm").
// With eight processes of one host copying at once and nothing else running, the copy engines of half of the GPUs were
// starved to ~9 GB/s while plain stores from the SMs kept ~15-18 GB/s on every GPU (tools/micro/duplex_bench.cu: slowest
// rank 80 -> 42 ms per step).  Inside the real pipeline the gain shrinks to 4 % at 8 GPUs and turns into a loss on one GPU,
// so this leg is opt-in (B2COL_D2H_MODE=sm, or =auto to let b2col_raycast_batch measure both and keep the faster).
__global__ void __launch_bounds__(256)
k_copy_out(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, size_t n_words)
{
    const size_t n16 = n_words / 4;
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    const size_t stride = (size_t)gridDim.x * blockDim.x, first = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t i = first; i < n16; i += stride) d4[i] = s4[i];
    for (size_t i = 4 * n16 + first; i < n_words; i += stride) dst[i] = src[i];
}

__global__ void __launch_bounds__(256)
k_export(uint32_t n, const BlobHeader* __restrict__ hdr, const Node* __restrict__ nodes, const LeafRec* __restrict__ leaves, BiTreeNode60* out,
         uint8_t* used)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) export_wrapper(i, leaves, out);
    if (i + 1u < n) export_internal(i, n, hdr->root, nodes, leaves, out, used);
}

// Compaction of the export candidates on the device (b2col_export_bitreenode): exclusive prefix sum of the `used` flags
// (tile-local scan, scan of the tile sums by one block, no host round trip), then every used node moves to its final
// place with its child indices renumbered — the order (wrappers, internal nodes, root last) is kept.
constexpr int XS_THREADS = 256, XS_ITEMS = 4, XS_TILE = XS_THREADS * XS_ITEMS;
__global__ void __launch_bounds__(XS_THREADS)
k_export_scan_tiles(const uint8_t* __restrict__ used, uint32_t n, uint32_t* __restrict__ excl, uint32_t* __restrict__ tile_sums)
{
    __shared__ uint32_t s_warp[XS_THREADS / 32];
    const uint32_t base = blockIdx.x * XS_TILE + threadIdx.x * XS_ITEMS;
    uint32_t v[XS_ITEMS], mine = 0u;
#pragma unroll
    for (int k = 0; k < XS_ITEMS; ++k) {
        v[k] = (base + k < n && used[base + k]) ? 1u : 0u;
        mine += v[k];
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0u;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    uint32_t run = before + incl - mine;
#pragma unroll
    for (int k = 0; k < XS_ITEMS; ++k) {
        if (base + k < n) excl[base + k] = run;
        run += v[k];
    }
    if (threadIdx.x == XS_THREADS - 1) tile_sums[blockIdx.x] = run;
}
__global__ void __launch_bounds__(1024)
k_export_scan_sums(uint32_t* tile_sums, uint32_t n_tiles, uint32_t* total)
{
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0u;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n_tiles; base += 1024u) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t mine = i < n_tiles ? tile_sums[i] : 0u;
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t before = s_carry;
        for (int w = 0; w < warp; ++w) before += s_warp[w];
        if (i < n_tiles) tile_sums[i] = before + incl - mine;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = before + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = s_carry;
}
__global__ void __launch_bounds__(256)
k_export_compact(uint32_t n_cand, const BiTreeNode60* __restrict__ cand, const uint8_t* __restrict__ used, const uint32_t* __restrict__ excl,
                 const uint32_t* __restrict__ tile_sums, BiTreeNode60* __restrict__ out)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_cand || !used[k]) return;
    BiTreeNode60 nd = cand[k];
    if (nd.type1 == 0) nd.index1 = excl[nd.index1] + tile_sums[nd.index1 / XS_TILE];
    if (nd.type2 == 0) nd.index2 = excl[nd.index2] + tile_sums[nd.index2 / XS_TILE];
    out[excl[k] + tile_sums[k / XS_TILE]] = nd;
}

// =================================================================================================
// host side
// =================================================================================================
namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    template <class T>
    T* as() const { return static_cast<T*>(p); }
};

thread_local std::string g_create_error;

}  // namespace

struct b2col_comm_state;  // b2col_comm.inl

struct b2col_ctx {
    int device = 0;
    b2col_comm_state* comm = nullptr;  // multi-GPU: communicator, gather window (b2col_comm_*)
    // a host-buffer batch whose work is enqueued but not yet waited for (raycast_batch_begin / _end)
    struct PendingBatch {
        bool active = false, pipeline = false, timed = false, has_dev_out = false;
        int d2h = 0, mode = 0;
        size_t n = 0;
        std::chrono::steady_clock::time_point t_begin;
    } pending;
    cudaStream_t stream = nullptr;
    b2col_config cfg{};
    float T = 1000.0f;
    int cfg_bits = 0;  // 0 = choose the key width from the collider count
    int bits = 11, drop = 1, passes = 4;
    std::string err;
    uint64_t launches = 0;

    // entities
    uint32_t n_entities = 0;
    bool have_hierarchy = false, have_trs = false, world_external = false;
    bool world_fresh = false;  // d_world already holds the matrices of the current hierarchy + TRS (b2col_propagate)
    std::vector<uint32_t> level_offsets;  // host; level k = [off[k], off[k+1]) in d_level_entities
    DevBuf d_parent, d_level_entities, d_pos, d_rot, d_scl, d_world;
    // b2col_bind_trs_device: the frame's kernels read the CALLER's TRS arrays in place (no 44 B/entity hand-over copy per frame)
    bool trs_bound = false;
    bool bound_padded = false;  // the bound arrays' allocations extend >= 16 bytes behind them: the bulk-copy kernel may read them
    const float* bound_pos = nullptr;
    const float4* bound_rot = nullptr;
    const float* bound_scl = nullptr;
    const float* trs_pos() const { return trs_bound ? bound_pos : d_pos.as<float>(); }
    const float4* trs_rot() const { return trs_bound ? bound_rot : d_rot.as<float4>(); }
    const float* trs_scl() const { return trs_bound ? bound_scl : d_scl.as<float>(); }
    // colliders
    uint32_t n_colliders = 0, max_col_entity = 0;
    bool have_colliders = false;
    DevBuf d_col_entity, d_col_local, d_world_aabb;
    // build scratch
    uint64_t counted_visits = 0, counted_rays = 0;  // B2COL_MODE_LBVH_COUNT totals
    bool pdl = true;                             // programmatic dependent launch between the kernels of a frame (B2COL_PDL=0: off)
    int top_levels = -1;                         // -1: automatic (trees larger than L2); dev override B2COL_TOP_LEVELS
    DevBuf d_keys[2], d_vals[2], d_xch, d_zero;  // d_xch: per-node arrival/exchange words; d_zero: histograms, tickets, look-back words
    DevBuf d_nodes_full, d_other_bound;  // export only: full-precision two-slot nodes + their arrival words (k_build_exact)
    DevBuf d_export_cand, d_export_out, d_export_used, d_export_scan;  // export only: candidates, compacted nodes, flags, prefix sums
    bool exact_nodes_valid = false;      // d_nodes_full matches the tree in the blob
    void* xch_ready = nullptr;           // the d_xch allocation whose words are all in the "nobody arrived" state
    size_t zero_bytes = 0;
    bool scratch_clean = false;          // d_zero[0, zero_bytes) is all zero (left so by the last k_build)
    // fused propagate + refit: collider index of every entity (NO_COLLIDER = none); needs one collider per entity at most
    std::vector<uint32_t> h_col_entity, h_level_entities;
    DevBuf d_col_of_entity, d_level_col;  // collider index by entity (host-side product only) / by position in d_level_entities
    struct LevelRun { bool dense = false; uint32_t e0 = 0, c0 = 0; };
    std::vector<LevelRun> level_runs;     // per hierarchy level: entities and colliders are consecutive runs (k_propagate_refit_level<true>)
    bool no_dense_refit = false;          // dev: B2COL_NO_DENSE_REFIT=1 keeps the list-driven kernel for every level (A/B timing)
    bool fuse_map_valid = false, fusable = false;
    int sorted_side = 0;
    // traversal structure: one allocation [BlobHeader | QNode x (N-1) | LeafRec x N]
    DevBuf d_blob;
    size_t blob_bytes = 0;
    uint32_t tree_leaves = 0;  // leaves of the tree currently in the blob
    int tree_key_bits = 0;     // Morton key width that tree was built with (b2col_update; 0: imported / received)
    bool tree_fused_keys = false;  // ... with keys from a plan made before the frame (k_refit_keys)
    bool have_tree = false;
    bool topology_valid = false;  // keys / leaf order in d_keys/d_vals[sorted_side] describe the tree in the blob and the current collider set (b2col_refit)
    uint64_t full_builds = 0, refits = 0;
    // reference-topology tree
    DevBuf d_ref_nodes, d_flag;
    uint32_t n_ref_nodes = 0;
    // staging for host-buffer queries
    DevBuf d_rays, d_hits, d_counter;
    // few-ray calls (GetRaycast: a batch of one): rays and results travel through ONE mapped, page-locked staging
    // block the kernel reads / writes directly, and the work counter comes from a pool of pre-zeroed slots — the call
    // is a single kernel launch + one synchronize, no copy or memset operations
    static constexpr size_t SMALL_MAX = 32, SMALL_SLOTS = 2048;
    void* h_small = nullptr;   // [SMALL_MAX rays | SMALL_MAX results], cudaHostAllocMapped
    void* d_small = nullptr;   // device address of the same block
    DevBuf d_small_counters;   // SMALL_SLOTS x u64, all zero beyond small_next
    size_t small_next = SMALL_SLOTS;
    cudaStream_t stream_in = nullptr, stream_out = nullptr;  // copy streams of the chunked host-buffer pipeline
    // second trace stream of the pipeline (B2COL_E2E_STREAMS=1 turns it off): chunks alternate between ctx->stream and this one, so that the
    // tail of one chunk's persistent kernel (its last rays, most SMs idle) overlaps the head of the next chunk's
    cudaStream_t stream_trace2 = nullptr;
    cudaEvent_t pipe_start_ev = nullptr;
    DevBuf d_pipe_counters;  // one pre-zeroed work counter per chunk (the kernels of two streams must not share one)
    int e2e_streams = 2;  // measured (profiles/r2_experiments.txt): compact results 1428 -> 1559 Mrays/s, 36-byte results unchanged (download-bound)
    std::vector<cudaEvent_t> pipe_events;
    // device->host leg of the pipeline: 0 = copy engine (default), 1 = SM stores (k_copy_out), -1 = measure both and keep the
    // faster (big calls only); B2COL_D2H_MODE=ce|sm|auto.  Measured (profiles/r1_e_d2h_modes.txt): alone on the host the copy
    // engine wins by far (1336 vs 743 Mrays/s: the copy kernel queues behind the persistent trace blocks); with eight
    // processes sharing the host the SM stores win by 4 % (2130 vs 2044) — not worth a default that depends on the neighbours.
    int d2h_policy = 0;
    int n_sm = 148;
    int env_key_bits = 0;         // dev: B2COL_KEY_BITS overrides the automatic key width (0 = automatic)
    // Adaptive key width (b2col_update's policy, prepare_build_inner): the hierarchy pass of every frame reports how many adjacent
    // sorted leaves its keys failed to separate — at the current width and at a width one radix pass (8 bits) narrower — into this
    // page-locked, mapped block {dup_cur, dup_narrow, leaves, key_bits}; the NEXT full build narrows or widens its keys accordingly.
    DevBuf d_key_ctr;             // k_build's counters (KEY_CTR_SLOTS sectors), summed and cleared by the next k_morton
    uint32_t* h_key_stat = nullptr;  // 48 bytes: [0..3] key statistic, [4..7] / [8..11] ordered centroid bounds of the last frame + its collider count
    void* d_key_stat = nullptr;   // device address of the same block
    // Morton plan known before the refit (k_refit_keys): made from centroid bounds the device reported for EARLIER frames
    // (h_key_stat[4..11]), widened by 1/32; kept while the reported bounds stay inside it and fill at least half of it
    struct PlanState {
        bool valid = false;
        float pmin[3] = {0, 0, 0}, pmax[3] = {0, 0, 0};
        int key_bits = 0;
        uint32_t n = 0;
        uint64_t gen = 0;
        MortonPlan plan{};
    } plan_state;
    bool fuse_keys = true;        // B2COL_FUSE_KEYS=0: always the exact plan (k_morton after the refit)
    bool use_fused_keys = false;  // this frame's choice (part of the graph key)
    bool refit_bulk = true;       // B2COL_TMA=0: dense levels through k_refit_keys (per-thread loads) instead of k_refit_keys_bulk (A/B timing)
    bool stride_always = false;   // dev: B2COL_REFIT=stride: the grid-stride kernel for frames without fused keys too (measured slower: 43.2 vs 40.3 us at 1M)
    int rk_waves = 1;             // dev: B2COL_RK_WAVES: grid of k_refit_keys in resident waves (1 = persistent)
    bool refit_stride = true;     // B2COL_REFIT=classic: one thread per entity (k_propagate_refit_level) instead of the grid-stride kernel (A/B timing)
    int adapt_key_bits = 0;       // width chosen by the policy (0: none yet, the static width by collider count)
    bool adaptive_keys = true;    // B2COL_ADAPTIVE_KEYS=0: always the static width
    bool no_packed_sort = false;  // B2COL_NO_PACKED_SORT=1: separate key / value arrays for keys of <= 32 bits too (A/B timing)
    // sort pipeline of a full build: MSD pass + block-local sort (b2col_sort.cuh) where the buckets fit a block, LSD otherwise
    // B2COL_BUILD=climb|direct|warp: k_build (0, default: measured fastest at 1M — 61.5 us vs 71.7 / 73.8, profiles/r2_experiments.txt), k_build_direct (1), k_build_warp (2)
    int build_variant = 0;
    int sort_policy = 1;              // 0 = automatic (MSD where the buckets fit); 1 = always LSD; 2 = MSD whenever the keys are packed  (B2COL_SORT=auto|lsd|msd)
    bool sort_msd = false;            // this frame's choice (part of the graph key)
    uint32_t* h_sort_stat = nullptr;  // page-locked: largest top-digit bucket of the last MSD frame, written by the frame itself
    int morton_blocks_per_sm = 0;  // 0 = by collider count; B2COL_MORTON_BLOCKS_PER_SM overrides (tuning)
    uint64_t d2h_calls = 0;
    double d2h_ns_per_ray[2] = {0.0, 0.0};
    // rays per chunk of the host-buffer pipeline: largest, first (tools/tune_e2e.py on B200, 16M rays: 256K -> 1M: 12.55 ms;
    // 128K -> 2M: 13.1; 256K -> 512K: 14.1 — smaller chunks leave the persistent trace kernel too few rays per launch)
    size_t e2e_chunk = 1u << 20, e2e_first_chunk = 256u << 10, e2e_min_chunk = 64u << 10;
    int trace_grid = 148 * 8, fetch_threshold = 25;
    cudaEvent_t events[16] = {};
    // dev: B2COL_PHASE_EVENTS=1 records an event after every phase of b2col_update / b2col_refit (b2col_debug_phase_ms)
    // CUDA graphs of the frame's kernel sequence (cfg.use_cuda_graph): [0] full update, [1] refit-only frame
    struct GraphSlot {
        cudaGraphExec_t exec = nullptr;
        uint64_t key[4] = {0, 0, 0, 0};
        uint64_t launches = 0;
    } graphs[2];
    bool use_graph = false;
    // a captured graph stays valid while no device buffer moved (buffers_gen), the hierarchy's level structure is the
    // same (shape_gen) and the counts / flags in the key are; the DATA in the buffers may change freely
    uint64_t buffers_gen = 0, shape_gen = 0, graph_replays = 0;
    bool no_small_path = false;  // dev: B2COL_NO_SMALL=1 sends few-ray calls through the copy path (A/B timing)
    bool no_fuse = false;  // dev: B2COL_NO_FUSE=1 keeps propagate and refit as separate kernels (A/B timing)
    bool phase_timing = false;
    // sparse TRS rows / world-matrix rows (the drop-in TransformSystem): page-locked host staging + device staging
    void* h_rows = nullptr;
    size_t h_rows_cap = 0;
    DevBuf d_rows;
    cudaEvent_t rows_ev = nullptr;  // the last copy out of / into h_rows has completed
    uint32_t rows_staged = 0;
    void* h_import_hdr = nullptr;  // page-locked copy of the header of the last imported blob, checked by b2col_sync
    bool import_check_pending = false;
    uint32_t import_check_leaves = 0;
    cudaEvent_t phase_ev[16] = {};
    int n_phase = 0;
};

namespace {

b2col_status fail(b2col_ctx* ctx, b2col_status st, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    else g_create_error = buf;
    return st;
}

#define CU(call)                                                                                             \
    do {                                                                                                     \
        cudaError_t e_ = (call);                                                                             \
        if (e_ != cudaSuccess) return fail(ctx, B2COL_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

#define LAUNCH(ctx_, kern, grid, block, ...)                                              \
    do {                                                                                  \
        kern<<<(grid), (block), 0, (ctx_)->stream>>>(__VA_ARGS__);                        \
        ++(ctx_)->launches;                                                               \
        cudaError_t e_ = cudaGetLastError();                                              \
        if (e_ != cudaSuccess)                                                            \
            return fail((ctx_), B2COL_ERR_CUDA, "launch of %s failed: %s", #kern, cudaGetErrorString(e_)); \
    } while (0)

// Frame-path launch: as LAUNCH, with programmatic stream serialization when ctx->pdl (every kernel launched this way
// starts with pdl_wait()).
#define LAUNCH_FRAME(ctx_, kern, grid, block, smem, ...)                                  \
    do {                                                                                  \
        cudaLaunchConfig_t cfg_ = {};                                                     \
        cfg_.gridDim = dim3(grid);                                                        \
        cfg_.blockDim = dim3(block);                                                      \
        cfg_.dynamicSmemBytes = (smem);                                                   \
        cfg_.stream = (ctx_)->stream;                                                     \
        cudaLaunchAttribute attr_[1];                                                     \
        attr_[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                 \
        attr_[0].val.programmaticStreamSerializationAllowed = 1;                          \
        cfg_.attrs = attr_;                                                               \
        cfg_.numAttrs = (ctx_)->pdl ? 1 : 0;                                              \
        cudaError_t e_ = cudaLaunchKernelEx(&cfg_, kern, __VA_ARGS__);                    \
        ++(ctx_)->launches;                                                               \
        if (e_ != cudaSuccess)                                                            \
            return fail((ctx_), B2COL_ERR_CUDA, "launch of %s failed: %s", #kern, cudaGetErrorString(e_)); \
    } while (0)

b2col_status ensure(b2col_ctx* ctx, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return B2COL_OK;
    if (b.p) {
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    const size_t want = bytes + bytes / 8 + 256;  // slack: scenes grow a little every frame
    CU(cudaMalloc(&b.p, want));
    b.cap = want;
    ++ctx->buffers_gen;  // captured graphs hold the old pointers
    return B2COL_OK;
}

void release(DevBuf& b)
{
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
}

inline uint32_t cdiv(size_t a, size_t b) { return (uint32_t)((a + b - 1) / b); }

// blob = [header | QNode x n_nodes | LeafRec x n_leaves | QNode x top_cap: compact copy of the top levels (k_top_layout)]
inline size_t top_cap_for(uint32_t n_leaves)
{
    const size_t n_nodes = n_leaves > 1 ? n_leaves - 1 : 1;
    return std::min<size_t>(TOP_CAP, n_nodes);
}
size_t blob_size_for(uint32_t n_leaves)
{
    const size_t n_nodes = n_leaves > 1 ? n_leaves - 1 : 1;
    return sizeof(BlobHeader) + n_nodes * sizeof(QNode) + (size_t)n_leaves * sizeof(LeafRec) + top_cap_for(n_leaves) * sizeof(QNode);
}
inline BlobHeader* blob_hdr(const b2col_ctx* c) { return c->d_blob.as<BlobHeader>(); }
inline BlobHeader* blob_hdr_n(const b2col_ctx* c) { return c->d_blob.as<BlobHeader>(); }
inline QNode* blob_qnodes(const b2col_ctx* c) { return reinterpret_cast<QNode*>(static_cast<char*>(c->d_blob.p) + sizeof(BlobHeader)); }
inline LeafRec* blob_leaves(const b2col_ctx* c, uint32_t n_leaves)
{
    const size_t n_nodes = n_leaves > 1 ? n_leaves - 1 : 1;
    return reinterpret_cast<LeafRec*>(static_cast<char*>(c->d_blob.p) + sizeof(BlobHeader) + n_nodes * sizeof(QNode));
}

#define PHASE(ctx_)                                                                                         \
    do {                                                                                                    \
        if ((ctx_)->phase_timing && (ctx_)->n_phase < 16) cudaEventRecord((ctx_)->phase_ev[(ctx_)->n_phase++], (ctx_)->stream); \
    } while (0)

#define GUARD(st_expr)                      \
    do {                                    \
        b2col_status s_ = (st_expr);        \
        if (s_ != B2COL_OK) return s_;      \
    } while (0)

// `on` / `counter`: launch on another stream of the context with a work counter the caller has already zeroed (chunked pipeline)
b2col_status enqueue_trace(b2col_ctx* ctx, const b2col_ray* d_rays, size_t n, b2col_raycast* d_out, int mode, bool remote_out = false, bool compact = false,
                           cudaStream_t on = nullptr, unsigned long long* counter = nullptr)
{
    const size_t out_bytes = compact ? 12 : sizeof(b2col_raycast);  // compact: d_out is an array of b2col_hit
    if (n == 0) return B2COL_OK;
    const bool count = mode == B2COL_MODE_LBVH_COUNT;
    if (mode == B2COL_MODE_LBVH || count) {
        if (!ctx->have_tree) {
            // frame-0 behaviour of the reference: bvh_ is empty before the first onUpdate, every query misses (:167)
            CU(cudaMemsetAsync(d_out, 0, n * out_bytes, ctx->stream));
            return B2COL_OK;
        }
        if (ctx->tree_leaves == 0) {  // empty tree: value-initialised Raycast (:164-167)
            CU(cudaMemsetAsync(d_out, 0, n * out_bytes, ctx->stream));
            return B2COL_OK;
        }
        const uint32_t grid = (uint32_t)std::min<size_t>((size_t)ctx->trace_grid, (n + TRACE_BLOCK - 1) / TRACE_BLOCK);
        if (on && counter && !count) {
            if (compact) k_trace_lbvh<false, false, true><<<grid, TRACE_BLOCK, 0, on>>>(d_rays, (unsigned long long)n, d_out, blob_hdr(ctx), blob_qnodes(ctx),
                                                                                      blob_leaves(ctx, ctx->tree_leaves), ctx->T, counter, ctx->fetch_threshold);
            else k_trace_lbvh<false, false><<<grid, TRACE_BLOCK, 0, on>>>(d_rays, (unsigned long long)n, d_out, blob_hdr(ctx), blob_qnodes(ctx),
                                                                        blob_leaves(ctx, ctx->tree_leaves), ctx->T, counter, ctx->fetch_threshold);
            ++ctx->launches;
            if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) return fail(ctx, B2COL_ERR_CUDA, "launch of k_trace_lbvh failed: %s", cudaGetErrorString(e));
            return B2COL_OK;
        }
        GUARD(ensure(ctx, ctx->d_counter, 16));
        CU(cudaMemsetAsync(ctx->d_counter.p, 0, 16, ctx->stream));
        if (count) {            LAUNCH(ctx, (k_trace_lbvh<true, false>), grid, TRACE_BLOCK, d_rays, (unsigned long long)n, d_out, blob_hdr(ctx), blob_qnodes(ctx),
                   blob_leaves(ctx, ctx->tree_leaves), ctx->T, ctx->d_counter.as<unsigned long long>(), ctx->fetch_threshold);
            unsigned long long v = 0;
            CU(cudaMemcpyAsync(&v, ctx->d_counter.as<unsigned long long>() + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            ctx->counted_visits += v;
            ctx->counted_rays += n;
        }
        else {
            if (compact)
                LAUNCH(ctx, (k_trace_lbvh<false, false, true>), grid, TRACE_BLOCK, d_rays, (unsigned long long)n, d_out, blob_hdr(ctx), blob_qnodes(ctx),
                       blob_leaves(ctx, ctx->tree_leaves), ctx->T, ctx->d_counter.as<unsigned long long>(), ctx->fetch_threshold);
            else if (remote_out)
                LAUNCH(ctx, (k_trace_lbvh<false, true>), grid, TRACE_BLOCK, d_rays, (unsigned long long)n, d_out, blob_hdr(ctx), blob_qnodes(ctx),
                       blob_leaves(ctx, ctx->tree_leaves), ctx->T, ctx->d_counter.as<unsigned long long>(), ctx->fetch_threshold);
            else
                LAUNCH(ctx, (k_trace_lbvh<false, false>), grid, TRACE_BLOCK, d_rays, (unsigned long long)n, d_out, blob_hdr(ctx), blob_qnodes(ctx),
                       blob_leaves(ctx, ctx->tree_leaves), ctx->T, ctx->d_counter.as<unsigned long long>(), ctx->fetch_threshold);
        }
    }
    else if (mode == B2COL_MODE_REFTOPO) {
        GUARD(ensure(ctx, ctx->d_flag, 4));
        CU(cudaMemsetAsync(ctx->d_flag.p, 0, 4, ctx->stream));
        LAUNCH(ctx, k_trace_reftopo, cdiv(n, 128), 128, d_rays, n, d_out, ctx->d_ref_nodes.as<uint32_t>(), ctx->n_ref_nodes, ctx->T,
               ctx->d_flag.as<uint32_t>());
    }
    else {
        return fail(ctx, B2COL_ERR_INVALID, "unknown mode %d", mode);
    }
    return B2COL_OK;
}

}  // namespace

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

int b2col_abi_version(void) { return B2COL_ABI_VERSION; }

b2col_status b2col_device_count(int* out)
{
    if (!out) return B2COL_ERR_INVALID;
    int count = 0;
    const cudaError_t e = cudaGetDeviceCount(&count);
    *out = e == cudaSuccess ? count : 0;
    if (e != cudaSuccess) cudaGetLastError();
    return *out > 0 ? B2COL_OK : B2COL_ERR_NO_DEVICE;
}

b2col_status b2col_create(b2col_ctx** out, const b2col_config* cfg)
{
    b2col_ctx* ctx = nullptr;  // for the CU/fail macros: errors before the ctx exists go to g_create_error
    if (!out) return fail(nullptr, B2COL_ERR_INVALID, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, B2COL_ERR_NO_DEVICE, "no CUDA device (%s); libb2col has no CPU path", e != cudaSuccess ? cudaGetErrorString(e) : "count = 0");
    const int dev = cfg ? cfg->device : 0;
    if (dev < 0 || dev >= count) return fail(nullptr, B2COL_ERR_INVALID, "device %d out of range (%d devices)", dev, count);
    CU(cudaSetDevice(dev));
    b2col_ctx* c = new (std::nothrow) b2col_ctx();
    if (!c) return fail(nullptr, B2COL_ERR_INVALID, "out of host memory");
    c->device = dev;
    if (cfg) c->cfg = *cfg;
    c->T = (cfg && cfg->max_distance > 0.0f) ? cfg->max_distance : 1000.0f;
    c->cfg_bits = (cfg && cfg->morton_bits_per_axis > 0) ? cfg->morton_bits_per_axis : 0;
    if (c->cfg_bits != 0 && (c->cfg_bits < 4 || c->cfg_bits > 21)) {
        delete c;
        return fail(nullptr, B2COL_ERR_INVALID, "morton_bits_per_axis must be 0 (auto) or in [4,21]");
    }
    e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete c;
        return fail(nullptr, B2COL_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
    }
    for (auto& ev : c->events) cudaEventCreate(&ev);
    {   // persistent trace kernel: one resident wave of blocks
        int sms = 148, per_sm = 8;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        c->n_sm = sms;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace_lbvh<false, false>, TRACE_BLOCK, 0) != cudaSuccess || per_sm < 1) per_sm = 8;
        if (const char* e = getenv("B2COL_TRACE_BLOCKS_PER_SM")) per_sm = std::max(1, atoi(e));  // tuning knobs
        if (const char* e = getenv("B2COL_E2E_CHUNK")) c->e2e_chunk = std::max<size_t>(1024, (size_t)atoll(e));
        if (const char* e = getenv("B2COL_E2E_FIRST_CHUNK")) c->e2e_first_chunk = std::max<size_t>(1024, (size_t)atoll(e));
        if (const char* e = getenv("B2COL_E2E_STREAMS")) c->e2e_streams = atoi(e) >= 2 ? 2 : 1;
        if (const char* e = getenv("B2COL_E2E_MIN_CHUNK")) c->e2e_min_chunk = std::max<size_t>(1024, (size_t)atoll(e));
        c->e2e_chunk = std::max(c->e2e_chunk, c->e2e_first_chunk);
        if (const char* e = getenv("B2COL_PHASE_EVENTS")) c->phase_timing = atoi(e) != 0;
        if (const char* e = getenv("B2COL_NO_FUSE")) c->no_fuse = atoi(e) != 0;
        if (const char* e = getenv("B2COL_NO_DENSE_REFIT")) c->no_dense_refit = atoi(e) != 0;
        if (const char* e = getenv("B2COL_NO_SMALL")) c->no_small_path = atoi(e) != 0;
        cudaFuncSetAttribute(k_top_layout, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TOP_SMEM);  // 64 KB frontier; a failure surfaces at the launch
        if (const char* e = getenv("B2COL_PDL")) c->pdl = atoi(e) != 0;
        if (const char* e = getenv("B2COL_NO_PACKED_SORT")) c->no_packed_sort = atoi(e) != 0;
        if (const char* e = getenv("B2COL_KEY_BITS")) c->env_key_bits = std::min(63, std::max(0, atoi(e)));  // dev: total Morton key width (A/B timing)
        if (const char* e = getenv("B2COL_ADAPTIVE_KEYS")) c->adaptive_keys = atoi(e) != 0;
        if (cudaMalloc(&c->d_key_ctr.p, KEY_CTR_SLOTS * 32) != cudaSuccess || cudaMemset(c->d_key_ctr.p, 0, KEY_CTR_SLOTS * 32) != cudaSuccess) {
            cudaGetLastError();
            b2col_destroy(c);
            return fail(nullptr, B2COL_ERR_CUDA, "cudaMalloc of the key statistic counters failed");
        }
        c->d_key_ctr.cap = KEY_CTR_SLOTS * 32;
        if (const char* e = getenv("B2COL_FUSE_KEYS")) c->fuse_keys = atoi(e) != 0;
        if (const char* e = getenv("B2COL_REFIT")) {
            c->refit_stride = strcmp(e, "classic") != 0;
            c->stride_always = strcmp(e, "stride") == 0;
        }
        if (const char* e = getenv("B2COL_TMA")) c->refit_bulk = atoi(e) != 0;
        if (const char* e = getenv("B2COL_RK_WAVES")) c->rk_waves = std::min(64, std::max(1, atoi(e)));
        if (cudaHostAlloc((void**)&c->h_key_stat, 48, cudaHostAllocMapped) == cudaSuccess && cudaHostGetDevicePointer(&c->d_key_stat, c->h_key_stat, 0) == cudaSuccess) {
            memset(c->h_key_stat, 0, 48);
        }
        else {
            cudaGetLastError();
            if (c->h_key_stat) cudaFreeHost(c->h_key_stat);
            c->h_key_stat = nullptr;  // no statistic: the static width only
            c->d_key_stat = nullptr;
        }
        if (const char* e = getenv("B2COL_MORTON_BLOCKS_PER_SM")) c->morton_blocks_per_sm = std::max(0, atoi(e));
        if (const char* e = getenv("B2COL_BUILD")) c->build_variant = !strcmp(e, "direct") ? 1 : (!strcmp(e, "warp") ? 2 : 0);
        if (const char* e = getenv("B2COL_SORT")) c->sort_policy = !strcmp(e, "auto") ? 0 : (!strcmp(e, "msd") ? 2 : 1);
        cudaFuncSetAttribute(k_local_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LocalSortSmem));
        if (cudaHostAlloc((void**)&c->h_sort_stat, sizeof(uint32_t), cudaHostAllocDefault) == cudaSuccess) *c->h_sort_stat = 0u;
        else {
            c->h_sort_stat = nullptr;  // no stat word: the LSD pipeline only
            cudaGetLastError();
        }
        if (const char* e = getenv("B2COL_D2H_MODE")) c->d2h_policy = !strcmp(e, "ce") ? 0 : (!strcmp(e, "sm") ? 1 : -1);
        if (const char* e = getenv("B2COL_TOP_LEVELS")) c->top_levels = std::min(TOP_LEVELS, std::max(0, atoi(e)));
        c->use_graph = cfg && cfg->use_cuda_graph != 0;
        if (const char* e = getenv("B2COL_GRAPH")) c->use_graph = atoi(e) != 0;
        if (c->phase_timing)
            for (auto& ev : c->phase_ev) cudaEventCreate(&ev);
        if (const char* e = getenv("B2COL_TRACE_FETCH_THRESHOLD")) c->fetch_threshold = std::min(32, std::max(1, atoi(e)));
        c->trace_grid = sms * per_sm;
    }
    *out = c;
    return B2COL_OK;
}

void b2col_destroy(b2col_ctx* ctx)
{
    if (!ctx) return;
    b2col_comm_finalize(ctx);
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (DevBuf* b : {&ctx->d_parent, &ctx->d_level_entities, &ctx->d_pos, &ctx->d_rot, &ctx->d_scl, &ctx->d_world, &ctx->d_col_entity,
                      &ctx->d_col_local, &ctx->d_world_aabb, &ctx->d_col_of_entity, &ctx->d_level_col, &ctx->d_key_ctr, &ctx->d_keys[0], &ctx->d_keys[1], &ctx->d_vals[0], &ctx->d_vals[1],
                      &ctx->d_other_bound, &ctx->d_export_cand, &ctx->d_export_out, &ctx->d_export_used, &ctx->d_export_scan, &ctx->d_xch, &ctx->d_zero, &ctx->d_nodes_full, &ctx->d_blob, &ctx->d_ref_nodes, &ctx->d_flag, &ctx->d_rays, &ctx->d_hits, &ctx->d_counter})
        release(*b);
    for (auto& g : ctx->graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    if (ctx->h_import_hdr) cudaFreeHost(ctx->h_import_hdr);
    if (ctx->h_sort_stat) cudaFreeHost(ctx->h_sort_stat);
    if (ctx->h_key_stat) cudaFreeHost(ctx->h_key_stat);
    if (ctx->h_rows) cudaFreeHost(ctx->h_rows);
    if (ctx->rows_ev) cudaEventDestroy(ctx->rows_ev);
    release(ctx->d_rows);
    release(ctx->d_small_counters);
    for (auto& ev : ctx->events)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->pipe_events) cudaEventDestroy(ev);
    if (ctx->stream_trace2) cudaStreamDestroy(ctx->stream_trace2);
    if (ctx->pipe_start_ev) cudaEventDestroy(ctx->pipe_start_ev);
    release(ctx->d_pipe_counters);
    if (ctx->stream_in) cudaStreamDestroy(ctx->stream_in);
    if (ctx->stream_out) cudaStreamDestroy(ctx->stream_out);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* b2col_last_error(const b2col_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

b2col_status b2col_sync(b2col_ctx* ctx)
{
    if (!ctx) return B2COL_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->import_check_pending) {  // header of the last imported blob (b2col_bvh_import_blob), now on the host
        ctx->import_check_pending = false;
        const BlobHeader* h = static_cast<const BlobHeader*>(ctx->h_import_hdr);
        const bool empty = h->magic == 0u && h->n_leaves == 0u && ctx->import_check_leaves == 0u;
        if (!empty && (h->magic != BLOB_MAGIC || h->n_leaves != ctx->import_check_leaves)) {
            ctx->have_tree = false;
            return fail(ctx, B2COL_ERR_INVALID, "imported blob rejected: magic %08x, %u leaves in the header, %u by its size", h->magic, h->n_leaves, ctx->import_check_leaves);
        }
    }
    return B2COL_OK;
}

void* b2col_stream(b2col_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

b2col_status b2col_set_hierarchy(b2col_ctx* ctx, uint32_t n_entities, const uint32_t* h_parent)
{
    if (!ctx || (!h_parent && n_entities)) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    const size_t rows = (size_t)n_entities + 1;
    // depth of every entity; parents precede children, so one ascending pass suffices (this is the
    // order TransformSystem::onUpdate itself relies on, reference src/system_transform.cpp:16,28-31)
    std::vector<uint32_t> depth(rows, 0);
    uint32_t max_depth = 0;
    for (uint32_t e = 1; e <= n_entities; ++e) {
        const uint32_t p = h_parent[e];
        if (p >= e) return fail(ctx, B2COL_ERR_UNSUPPORTED, "entity %u has parent %u: parents must have smaller IDs than their children", e, p);
        depth[e] = p ? depth[p] + 1 : 0;
        max_depth = std::max(max_depth, depth[e]);
    }
    const uint32_t n_levels = n_entities ? max_depth + 1 : 0;
    ctx->level_offsets.assign(n_levels + 1, 0);
    for (uint32_t e = 1; e <= n_entities; ++e) ++ctx->level_offsets[depth[e] + 1];
    for (uint32_t k = 0; k < n_levels; ++k) ctx->level_offsets[k + 1] += ctx->level_offsets[k];
    std::vector<uint32_t> order(n_entities);
    {
        std::vector<uint32_t> cursor(ctx->level_offsets.begin(), ctx->level_offsets.end() - (n_levels ? 1 : 0));
        for (uint32_t e = 1; e <= n_entities; ++e) order[cursor[depth[e]]++] = e;
    }
    GUARD(ensure(ctx, ctx->d_parent, rows * 4));
    GUARD(ensure(ctx, ctx->d_level_entities, (size_t)n_entities * 4 + 4));
    CU(cudaMemcpyAsync(ctx->d_parent.p, h_parent, rows * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (n_entities) CU(cudaMemcpyAsync(ctx->d_level_entities.p, order.data(), (size_t)n_entities * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));  // `order` is a local
    ctx->h_level_entities.swap(order);       // kept for prepare_fuse_map (collider index in level order)
    ctx->n_entities = n_entities;
    ctx->have_hierarchy = true;
    ctx->have_trs = false;
    ctx->trs_bound = false;  // a new hierarchy wants new TRS arrays (b2col_set_trs* / b2col_bind_trs_device)
    ctx->world_fresh = false;
    ctx->fuse_map_valid = false;
    ++ctx->shape_gen;
    return B2COL_OK;
}

static b2col_status set_trs_common(b2col_ctx* ctx, uint32_t n_entities, const float* pos3, const float* rot, const float* scl3, cudaMemcpyKind kind)
{
    if (!ctx || !pos3 || !rot || !scl3) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_hierarchy || n_entities != ctx->n_entities)
        return fail(ctx, B2COL_ERR_STATE, "b2col_set_hierarchy must be called first with the same n_entities");
    CU(cudaSetDevice(ctx->device));
    const size_t rows = (size_t)n_entities + 1;
    GUARD(ensure(ctx, ctx->d_pos, rows * 12));
    GUARD(ensure(ctx, ctx->d_rot, rows * 16));
    GUARD(ensure(ctx, ctx->d_scl, rows * 12));
    CU(cudaMemcpyAsync(ctx->d_pos.p, pos3, rows * 12, kind, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_rot.p, rot, rows * 16, kind, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_scl.p, scl3, rows * 12, kind, ctx->stream));
    if (ctx->trs_bound) {  // back to the context's own copies
        ctx->trs_bound = false;
        ++ctx->buffers_gen;
    }
    ctx->have_trs = true;
    ctx->world_external = false;
    ctx->world_fresh = false;
    return B2COL_OK;
}

b2col_status b2col_set_trs(b2col_ctx* ctx, uint32_t n_entities, const float* h_pos3, const float* h_rot_xyzw, const float* h_scl3)
{
    GUARD(set_trs_common(ctx, n_entities, h_pos3, h_rot_xyzw, h_scl3, cudaMemcpyHostToDevice));
    CU(cudaStreamSynchronize(ctx->stream));  // the caller may reuse its buffers as soon as we return
    return B2COL_OK;
}

b2col_status b2col_set_trs_device(b2col_ctx* ctx, uint32_t n_entities, const float* d_pos3, const float* d_rot_xyzw, const float* d_scl3)
{
    return set_trs_common(ctx, n_entities, d_pos3, d_rot_xyzw, d_scl3, cudaMemcpyDeviceToDevice);
}

b2col_status b2col_bind_trs_device(b2col_ctx* ctx, uint32_t n_entities, const float* d_pos3, const float* d_rot_xyzw, const float* d_scl3)
{
    if (!ctx || !d_pos3 || !d_rot_xyzw || !d_scl3) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_hierarchy || n_entities != ctx->n_entities)
        return fail(ctx, B2COL_ERR_STATE, "b2col_set_hierarchy must be called first with the same n_entities");
    if ((reinterpret_cast<uintptr_t>(d_rot_xyzw) & 15u) || (reinterpret_cast<uintptr_t>(d_pos3) & 3u) || (reinterpret_cast<uintptr_t>(d_scl3) & 3u))
        return fail(ctx, B2COL_ERR_INVALID, "bound TRS arrays: rotations must be 16-byte aligned (read as float4), positions and scales 4-byte aligned");
    CU(cudaSetDevice(ctx->device));
    {   // the arrays must be device (or managed) memory this context's device can read
        cudaPointerAttributes at{};
        for (const void* p : {(const void*)d_pos3, (const void*)d_rot_xyzw, (const void*)d_scl3}) {
            const cudaError_t e = cudaPointerGetAttributes(&at, p);
            if (e != cudaSuccess || (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)) {
                cudaGetLastError();
                return fail(ctx, B2COL_ERR_INVALID, "b2col_bind_trs_device takes device memory (use b2col_set_trs for host arrays)");
            }
        }
    }
    // May the bulk-copy kernel read the arrays?  Its 16-byte granular copies end up to 15 bytes behind a range's last row: allowed when
    // the ALLOCATION each array lives in (cuMemGetAddressRange; a framework's pool block counts) extends that far behind the array
    bool padded = false;
    {
        typedef int (*PFN_range)(unsigned long long*, size_t*, unsigned long long);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qr) == cudaSuccess && fn && qr == cudaDriverEntryPointSuccess) {
            padded = true;
            const size_t rows = (size_t)n_entities + 1;
            const std::pair<const void*, size_t> arrays[3] = {{d_pos3, rows * 12}, {d_rot_xyzw, rows * 16}, {d_scl3, rows * 12}};
            for (const auto& a : arrays) {
                unsigned long long base = 0;
                size_t size = 0;
                const unsigned long long p = reinterpret_cast<unsigned long long>(a.first);
                if (reinterpret_cast<PFN_range>(fn)(&base, &size, p) != 0 || p + a.second + 16 > base + size) padded = false;
            }
        }
        else {
            cudaGetLastError();
        }
    }
    const float4* rot = reinterpret_cast<const float4*>(d_rot_xyzw);
    if (ctx->bound_padded != padded) ++ctx->buffers_gen;  // the captured frame holds the kernel choice
    ctx->bound_padded = padded;
    if (!ctx->trs_bound || ctx->bound_pos != d_pos3 || ctx->bound_rot != rot || ctx->bound_scl != d_scl3) ++ctx->buffers_gen;  // a captured frame holds the old pointers
    ctx->trs_bound = true;
    ctx->bound_pos = d_pos3;
    ctx->bound_rot = rot;
    ctx->bound_scl = d_scl3;
    ctx->have_trs = true;
    ctx->world_external = false;
    ctx->world_fresh = false;  // and never becomes true while bound: the caller changes the arrays without telling us (b2col_propagate)
    return B2COL_OK;
}

// page-locked staging of at least `bytes`, safe to overwrite (the copies that last used it have completed)
static b2col_status rows_staging(b2col_ctx* ctx, size_t bytes)
{
    if (!ctx->rows_ev) CU(cudaEventCreateWithFlags(&ctx->rows_ev, cudaEventDisableTiming));
    else CU(cudaEventSynchronize(ctx->rows_ev));
    if (bytes > ctx->h_rows_cap) {
        if (ctx->h_rows) CU(cudaFreeHost(ctx->h_rows));
        ctx->h_rows = nullptr;
        ctx->h_rows_cap = 0;
        const size_t want = bytes + bytes / 4 + 4096;
        CU(cudaHostAlloc(&ctx->h_rows, want, cudaHostAllocDefault));
        ctx->h_rows_cap = want;
    }
    return ensure(ctx, ctx->d_rows, bytes);
}
// layout of a staged batch of n rows: [ids u32 x n | pad to 16 | rot float4 x n | pos float3 x n | scl float3 x n]
static inline size_t rows_off_rot(uint32_t n) { return ((size_t)n * 4 + 15) & ~(size_t)15; }
static inline size_t rows_bytes(uint32_t n) { return rows_off_rot(n) + (size_t)n * 40; }

b2col_status b2col_stage_trs_rows(b2col_ctx* ctx, uint32_t n_rows, uint32_t** h_entity, float** h_pos3, float** h_rot_xyzw, float** h_scl3)
{
    if (!ctx || !h_entity || !h_pos3 || !h_rot_xyzw || !h_scl3) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_trs) return fail(ctx, B2COL_ERR_STATE, "b2col_set_trs must have uploaded the full arrays once");
    if (ctx->trs_bound) return fail(ctx, B2COL_ERR_STATE, "the TRS arrays are the caller's (b2col_bind_trs_device): update them in place");
    CU(cudaSetDevice(ctx->device));
    GUARD(rows_staging(ctx, rows_bytes(n_rows)));
    char* base = static_cast<char*>(ctx->h_rows);
    *h_entity = reinterpret_cast<uint32_t*>(base);
    *h_rot_xyzw = reinterpret_cast<float*>(base + rows_off_rot(n_rows));
    *h_pos3 = *h_rot_xyzw + 4 * (size_t)n_rows;
    *h_scl3 = *h_pos3 + 3 * (size_t)n_rows;
    ctx->rows_staged = n_rows;
    return B2COL_OK;
}

b2col_status b2col_commit_trs_rows(b2col_ctx* ctx, uint32_t n_rows)
{
    if (!ctx) return B2COL_ERR_INVALID;
    if (n_rows > ctx->rows_staged) return fail(ctx, B2COL_ERR_STATE, "%u rows committed, %u staged", n_rows, ctx->rows_staged);
    if (ctx->trs_bound) return fail(ctx, B2COL_ERR_STATE, "the TRS arrays are the caller's (b2col_bind_trs_device): update them in place");
    const uint32_t staged = ctx->rows_staged;
    ctx->rows_staged = 0;
    if (n_rows == 0) return B2COL_OK;
    CU(cudaSetDevice(ctx->device));
    const uint32_t* ids = static_cast<const uint32_t*>(ctx->h_rows);
    for (uint32_t i = 0; i < n_rows; ++i)
        if (ids[i] == 0u || ids[i] > ctx->n_entities) return fail(ctx, B2COL_ERR_INVALID, "row %u names entity %u of %u", i, ids[i], ctx->n_entities);
    CU(cudaMemcpyAsync(ctx->d_rows.p, ctx->h_rows, rows_bytes(staged), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaEventRecord(ctx->rows_ev, ctx->stream));
    const char* d = static_cast<const char*>(ctx->d_rows.p);
    const float4* rot = reinterpret_cast<const float4*>(d + rows_off_rot(staged));
    const float* pos = reinterpret_cast<const float*>(rot + staged);
    const float* scl = pos + 3 * (size_t)staged;
    LAUNCH(ctx, k_scatter_trs, cdiv(n_rows, 256), 256, n_rows, reinterpret_cast<const uint32_t*>(d), pos, rot, scl, ctx->d_pos.as<float>(), ctx->d_rot.as<float4>(),
           ctx->d_scl.as<float>());
    ctx->world_external = false;
    ctx->world_fresh = false;
    return B2COL_OK;
}

b2col_status b2col_update_trs_rows(b2col_ctx* ctx, uint32_t n_rows, const uint32_t* h_entity, const float* h_pos3, const float* h_rot_xyzw, const float* h_scl3)
{
    if (!ctx || (n_rows && (!h_entity || !h_pos3 || !h_rot_xyzw || !h_scl3))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    uint32_t* e;
    float *p, *r, *s;
    GUARD(b2col_stage_trs_rows(ctx, n_rows, &e, &p, &r, &s));
    memcpy(e, h_entity, (size_t)n_rows * 4);
    memcpy(p, h_pos3, (size_t)n_rows * 12);
    memcpy(r, h_rot_xyzw, (size_t)n_rows * 16);
    memcpy(s, h_scl3, (size_t)n_rows * 12);
    return b2col_commit_trs_rows(ctx, n_rows);
}

b2col_status b2col_fetch_world_rows(b2col_ctx* ctx, uint32_t n_rows, const uint32_t* h_entity, const float** h_rows16)
{
    if (!ctx || !h_rows16 || (n_rows && !h_entity)) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    *h_rows16 = nullptr;
    if (n_rows == 0) return B2COL_OK;
    if (!ctx->d_world.p) return fail(ctx, B2COL_ERR_STATE, "no world matrices on the device: call b2col_propagate / b2col_update first");
    for (uint32_t i = 0; i < n_rows; ++i)
        if (h_entity[i] > ctx->n_entities) return fail(ctx, B2COL_ERR_INVALID, "row %u names entity %u of %u", i, h_entity[i], ctx->n_entities);
    CU(cudaSetDevice(ctx->device));
    const size_t ids_bytes = ((size_t)n_rows * 4 + 63) & ~(size_t)63;
    GUARD(rows_staging(ctx, ids_bytes + (size_t)n_rows * 64));
    memcpy(ctx->h_rows, h_entity, (size_t)n_rows * 4);
    char* d = static_cast<char*>(ctx->d_rows.p);
    CU(cudaMemcpyAsync(d, ctx->h_rows, (size_t)n_rows * 4, cudaMemcpyHostToDevice, ctx->stream));
    LAUNCH(ctx, k_gather_world, cdiv(4 * (size_t)n_rows, 256), 256, n_rows, reinterpret_cast<const uint32_t*>(d), ctx->d_world.as<float4>(), reinterpret_cast<float4*>(d + ids_bytes));
    CU(cudaMemcpyAsync(static_cast<char*>(ctx->h_rows) + ids_bytes, d + ids_bytes, (size_t)n_rows * 64, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaEventRecord(ctx->rows_ev, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    *h_rows16 = reinterpret_cast<const float*>(static_cast<char*>(ctx->h_rows) + ids_bytes);
    return B2COL_OK;
}

b2col_status b2col_set_world_matrices(b2col_ctx* ctx, uint32_t n_entities, const float* h_world16)
{
    if (!ctx || !h_world16) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    const size_t rows = (size_t)n_entities + 1;
    GUARD(ensure(ctx, ctx->d_world, rows * 64));
    CU(cudaMemcpyAsync(ctx->d_world.p, h_world16, rows * 64, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->n_entities = n_entities;
    ctx->world_external = true;
    return B2COL_OK;
}

b2col_status b2col_set_colliders(b2col_ctx* ctx, uint32_t n, const uint32_t* h_entity, const b2col_aabb* h_local)
{
    if (!ctx || (n && (!h_entity || !h_local))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (n >= (1u << 29)) return fail(ctx, B2COL_ERR_INVALID, "too many colliders (child refs hold 30-bit indices)");
    CU(cudaSetDevice(ctx->device));
    uint32_t max_e = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if (h_entity[i] == 0u) return fail(ctx, B2COL_ERR_INVALID, "collider %u refers to entity 0, which is not a valid entity", i);
        max_e = std::max(max_e, h_entity[i]);
    }
    ctx->max_col_entity = max_e;
    ctx->h_col_entity.assign(h_entity, h_entity + n);
    ctx->fuse_map_valid = false;
    GUARD(ensure(ctx, ctx->d_col_entity, (size_t)n * 4 + 4));
    GUARD(ensure(ctx, ctx->d_col_local, (size_t)n * 24 + 8));
    if (n) {
        CU(cudaMemcpyAsync(ctx->d_col_entity.p, h_entity, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->d_col_local.p, h_local, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    ctx->n_colliders = n;
    ctx->have_colliders = true;
    ctx->topology_valid = false;  // a new collider set needs a new tree
    if (ctx->h_sort_stat) *ctx->h_sort_stat = 0u;  // ... and gets a fresh try at the MSD sort pipeline
    ctx->adapt_key_bits = 0;                       // ... and starts from the static key width again
    ctx->plan_state.valid = false;                 // ... and from the exact Morton plan
    if (ctx->h_key_stat) {
        CU(cudaStreamSynchronize(ctx->stream));    // a frame in flight may still report into the block
        memset(ctx->h_key_stat, 0, 48);
    }
    return B2COL_OK;
}

// hierarchy + bounds from the sorted keys in d_keys/d_vals[side]
static b2col_status launch_build(b2col_ctx* ctx, uint32_t n, int side)
{
    if (ctx->build_variant == 0)  // B2COL_BUILD=climb: the per-thread climb with shared-memory arrivals (A/B timing; same tree, same bytes)
        LAUNCH_FRAME(ctx, k_build, cdiv(n, BUILD_BLOCK), BUILD_BLOCK, 0, n, ctx->d_keys[side].as<uint64_t>(), ctx->d_vals[side].as<uint32_t>(),
               ctx->d_world_aabb.as<LeafRec>(), ctx->d_xch.as<Xch>(), blob_qnodes(ctx), blob_leaves(ctx, n), blob_hdr(ctx),
               ctx->d_zero.as<uint32_t>(), (uint32_t)(ctx->zero_bytes / 4), ctx->d_key_ctr.as<uint32_t>());
#ifdef B2_HAVE_BUILD_VARIANTS
    else if (ctx->build_variant == 2)
        LAUNCH_FRAME(ctx, k_build_warp, cdiv(n, BUILD_BLOCK), BUILD_BLOCK, 0, n, ctx->d_keys[side].as<uint64_t>(), ctx->d_vals[side].as<uint32_t>(),
               ctx->d_world_aabb.as<LeafRec>(), ctx->d_xch.as<Xch>(), blob_qnodes(ctx), blob_leaves(ctx, n), blob_hdr(ctx),
               ctx->d_zero.as<uint32_t>(), (uint32_t)(ctx->zero_bytes / 4));
    else
        LAUNCH_FRAME(ctx, k_build_direct, cdiv(n, BUILD_BLOCK), BUILD_BLOCK, 0, n, ctx->d_keys[side].as<uint64_t>(), ctx->d_vals[side].as<uint32_t>(),
               ctx->d_world_aabb.as<LeafRec>(), ctx->d_xch.as<Xch>(), blob_qnodes(ctx), blob_leaves(ctx, n), blob_hdr(ctx),
               ctx->d_zero.as<uint32_t>(), (uint32_t)(ctx->zero_bytes / 4));
#else
    else return fail(ctx, B2COL_ERR_UNSUPPORTED, "B2COL_BUILD variants are not part of this build");
#endif
    ctx->scratch_clean = true;  // k_build leaves every scratch word zero
    // Trees that do not fit L2: compact copy of the top levels (see k_top_layout).  Below ~2M colliders the whole tree is
    // L2-resident and the copy buys nothing (measured at 1M: 1875 Mrays/s either way) but would cost the frame ~15 us.
    const int levels = ctx->top_levels >= 0 ? ctx->top_levels : (n > (1u << 21) ? TOP_LEVELS : 0);
    if (levels > 0 && n > 1u) {
        QNode* top = reinterpret_cast<QNode*>(blob_leaves(ctx, n) + n);
        LAUNCH_FRAME(ctx, k_top_layout, 1, TOP_BLOCK, TOP_SMEM, blob_hdr(ctx), (const QNode*)blob_qnodes(ctx), top, (uint32_t)((n - 1u) + n), std::min(levels, TOP_LEVELS));
    }
    return B2COL_OK;
}

static b2col_status enqueue_propagate(b2col_ctx* ctx)
{
    const size_t rows = (size_t)ctx->n_entities + 1;
    GUARD(ensure(ctx, ctx->d_world, rows * 64));
    const size_t n_levels = ctx->level_offsets.empty() ? 0 : ctx->level_offsets.size() - 1;
    for (size_t k = 0; k < n_levels; ++k) {
        const uint32_t off = ctx->level_offsets[k], cnt = ctx->level_offsets[k + 1] - off;
        if (!cnt) continue;
        LAUNCH_FRAME(ctx, k_propagate_level, cdiv(cnt, 256), 256, 0, ctx->d_level_entities.as<uint32_t>() + off, cnt, ctx->d_parent.as<uint32_t>(),
               ctx->trs_pos(), ctx->trs_rot(), ctx->trs_scl(), ctx->d_world.as<float4>());
    }
    return B2COL_OK;
}

b2col_status b2col_propagate(b2col_ctx* ctx)
{
    if (!ctx) return B2COL_ERR_INVALID;
    if (!(ctx->have_hierarchy && ctx->have_trs)) return fail(ctx, B2COL_ERR_STATE, "need b2col_set_hierarchy + b2col_set_trs");
    CU(cudaSetDevice(ctx->device));
    GUARD(enqueue_propagate(ctx));
    ctx->world_fresh = !ctx->trs_bound;  // bound arrays change without notice: the next update propagates again
    return B2COL_OK;
}

namespace {

// per-frame scratch, all zero when a frame starts (k_build wipes it for the next one):
//   [bounds: 16 words][digit histograms: RS_MAX_PASSES x 256][tile tickets: 16][look-back words of every pass]
constexpr size_t SCRATCH_BOUNDS_WORDS = 16;
static_assert(BOUNDS_WORDS <= (int)SCRATCH_BOUNDS_WORDS, "bounds fit their scratch slot");
inline uint32_t* scratch_bounds(const b2col_ctx* c) { return c->d_zero.as<uint32_t>(); }
inline uint32_t* scratch_ghist(const b2col_ctx* c) { return c->d_zero.as<uint32_t>() + SCRATCH_BOUNDS_WORDS; }
inline uint32_t* scratch_tickets(const b2col_ctx* c) { return scratch_ghist(c) + RS_MAX_PASSES * RS_RADIX; }
inline uint32_t* scratch_lookback(const b2col_ctx* c) { return scratch_tickets(c) + 16; }

// make sure the scratch exists at `bytes` and is zero
b2col_status prepare_scratch(b2col_ctx* ctx, size_t bytes)
{
    void* before = ctx->d_zero.p;
    GUARD(ensure(ctx, ctx->d_zero, bytes));
    if (ctx->d_zero.p != before || bytes > ctx->zero_bytes) ctx->scratch_clean = false;  // new memory, or words k_build never wiped
    ctx->zero_bytes = bytes;
    if (!ctx->scratch_clean) CU(cudaMemsetAsync(ctx->d_zero.p, 0, bytes, ctx->stream));
    ctx->scratch_clean = false;  // about to be used; k_build restores it
    return B2COL_OK;
}

// entity -> collider index map of the fused propagate + refit kernel; fusable only with at most one collider per entity
b2col_status prepare_fuse_map(b2col_ctx* ctx)
{
    if (ctx->fuse_map_valid) return B2COL_OK;
    const uint32_t E = ctx->n_entities, n = ctx->n_colliders;
    ctx->fusable = false;
    ctx->fuse_map_valid = true;
    if (n == 0 || ctx->max_col_entity > E || ctx->h_col_entity.size() != n) return B2COL_OK;
    std::vector<uint32_t> map((size_t)E + 1, NO_COLLIDER);
    for (uint32_t c = 0; c < n; ++c) {
        uint32_t& slot = map[ctx->h_col_entity[c]];
        if (slot != NO_COLLIDER) return B2COL_OK;  // two colliders on one entity: keep the separate refit kernel
        slot = c;
    }
    if (ctx->h_level_entities.size() != E) return B2COL_OK;
    // the same map in LEVEL order (what the kernel reads), and which levels are consecutive runs of entities and colliders
    std::vector<uint32_t> level_col(E);
    for (uint32_t i = 0; i < E; ++i) level_col[i] = map[ctx->h_level_entities[i]];
    const size_t n_levels = ctx->level_offsets.empty() ? 0 : ctx->level_offsets.size() - 1;
    ctx->level_runs.assign(n_levels, b2col_ctx::LevelRun());
    for (size_t k = 0; k < n_levels; ++k) {
        const uint32_t off = ctx->level_offsets[k], cnt = ctx->level_offsets[k + 1] - off;
        if (!cnt) continue;
        b2col_ctx::LevelRun& run = ctx->level_runs[k];
        run.e0 = ctx->h_level_entities[off];
        run.c0 = level_col[off];
        run.dense = run.c0 != NO_COLLIDER;
        for (uint32_t i = 0; run.dense && i < cnt; ++i) run.dense = ctx->h_level_entities[off + i] == run.e0 + i && level_col[off + i] == run.c0 + i;
    }
    GUARD(ensure(ctx, ctx->d_level_col, (size_t)E * 4 + 4));
    if (E) CU(cudaMemcpyAsync(ctx->d_level_col.p, level_col.data(), (size_t)E * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));  // `level_col` is a local
    ++ctx->shape_gen;  // a captured frame holds the per-level kernel choice and run origins
    ctx->fusable = true;
    return B2COL_OK;
}

// a1 + a2: world matrices and world AABBs of the frame (+ scene bounds into the scratch)
// `tail` (out): true when the refit kernel itself finished the frame's header (k_refit_keys' last block): the caller then
// launches neither k_morton (full build with fused keys) nor k_header (refit-only frame)
b2col_status enqueue_transform_refit(b2col_ctx* ctx, bool full, bool* tail_done)
{
    const uint32_t n = ctx->n_colliders;
    const bool propagate = !ctx->world_external && !ctx->world_fresh;
    *tail_done = false;
    if (propagate && ctx->fusable && !ctx->no_fuse && ctx->refit_stride && ((full && ctx->use_fused_keys) || ctx->stride_always)) {
        // grid-stride kernel: with the frame's keys when the plan is known (full && use_fused_keys), and — fused keys or refit-only frame —
        // with the header written by its last block
        const bool keys = full && ctx->use_fused_keys;
        const bool tail = keys || !full;
        const size_t n_levels = ctx->level_offsets.empty() ? 0 : ctx->level_offsets.size() - 1;
        size_t last = n_levels;
        for (size_t k = 0; k < n_levels; ++k)
            if (ctx->level_offsets[k + 1] > ctx->level_offsets[k]) last = k;
        KeyGen kg{};
        FrameTail ft{};
        ft.hdr = blob_hdr(ctx);
        ft.key_ctr = ctx->d_key_ctr.as<uint32_t>();
        ft.host_stat = static_cast<uint4*>(ctx->d_key_stat);
        ft.done = scratch_tickets(ctx) + 13;
        ft.n = n;
        ft.T = ctx->T;
        const bool packed = 3 * ctx->bits - ctx->drop <= 32 && !ctx->no_packed_sort;
        if (keys) {
            kg.plan = ctx->plan_state.plan;
            kg.passes = ctx->passes;
            kg.hist_lo = 0;
            kg.packed = (int)packed;
            ft.key_bits = 3 * ctx->bits - ctx->drop;
            for (int k = 0; k < 3; ++k) {
                ft.plan_min_inv[k] = ~float_to_ordered(ctx->plan_state.pmin[k]);
                ft.plan_max[k] = float_to_ordered(ctx->plan_state.pmax[k]);
            }
        }
        for (size_t k = 0; k < n_levels; ++k) {
            const uint32_t off = ctx->level_offsets[k], cnt = ctx->level_offsets[k + 1] - off;
            if (!cnt) continue;
            const b2col_ctx::LevelRun run = k < ctx->level_runs.size() ? ctx->level_runs[k] : b2col_ctx::LevelRun();
            ft.last_level = (tail && k == last) ? 1 : 0;
            const uint32_t grid = std::min<uint32_t>(cdiv(cnt, RK_BLOCK), (uint32_t)(ctx->n_sm * B2_RK_MIN_BLOCKS * ctx->rk_waves));
#define B2_REFIT_KEYS(D, K)                                                                                                                                   \
    LAUNCH_FRAME(ctx, (k_refit_keys<D, K>), grid, RK_BLOCK, 0, ctx->d_level_entities.as<uint32_t>() + off, ctx->d_level_col.as<uint32_t>() + off, run.e0, run.c0, \
                 cnt, (int)(k == 0), ctx->d_parent.as<uint32_t>(), ctx->trs_pos(), ctx->trs_rot(), ctx->trs_scl(), ctx->d_world.as<float4>(), \
                 ctx->d_col_local.as<Aabb24>(), ctx->d_world_aabb.as<LeafRec>(), scratch_bounds(ctx), kg, ctx->d_keys[0].as<uint64_t>(), ctx->d_vals[0].as<uint32_t>(), \
                 scratch_ghist(ctx), ft)
            const bool dense = run.dense && !ctx->no_dense_refit;
            if (dense && ctx->refit_bulk && (!ctx->trs_bound || ctx->bound_padded)) {  // inputs through the TMA unit, a tile ahead (k_refit_keys_bulk); its 16-byte granular copies read up to 30 bytes around a range: the context's own, padded arrays, or bound arrays whose allocations were checked (b2col_bind_trs_device)
#define B2_REFIT_BULK(K)                                                                                                                                       \
    LAUNCH_FRAME(ctx, (k_refit_keys_bulk<K>), grid, RK_BLOCK, 0, run.e0, run.c0, cnt, (int)(k == 0), ctx->d_parent.as<uint32_t>(), ctx->trs_pos(),               \
                 ctx->trs_rot(), ctx->trs_scl(), ctx->d_world.as<float4>(), ctx->d_col_local.as<Aabb24>(), ctx->d_world_aabb.as<LeafRec>(),      \
                 scratch_bounds(ctx), kg, ctx->d_keys[0].as<uint64_t>(), ctx->d_vals[0].as<uint32_t>(), scratch_ghist(ctx), ft)
                if (keys) B2_REFIT_BULK(true);
                else B2_REFIT_BULK(false);
#undef B2_REFIT_BULK
                continue;
            }
            if (dense && keys) B2_REFIT_KEYS(true, true);
            else if (dense) B2_REFIT_KEYS(true, false);
            else if (keys) B2_REFIT_KEYS(false, true);
            else B2_REFIT_KEYS(false, false);
#undef B2_REFIT_KEYS
        }
        *tail_done = tail && last < n_levels;
        PHASE(ctx);
        return B2COL_OK;
    }
    if (propagate && ctx->fusable && !ctx->no_fuse) {
        const size_t n_levels = ctx->level_offsets.empty() ? 0 : ctx->level_offsets.size() - 1;
        for (size_t k = 0; k < n_levels; ++k) {
            const uint32_t off = ctx->level_offsets[k], cnt = ctx->level_offsets[k + 1] - off;
            if (!cnt) continue;
            const b2col_ctx::LevelRun run = k < ctx->level_runs.size() ? ctx->level_runs[k] : b2col_ctx::LevelRun();
#define B2_REFIT_LEVEL(D)                                                                                                                                              \
    LAUNCH_FRAME(ctx, k_propagate_refit_level<D>, cdiv(cnt, REFIT_BLOCK), REFIT_BLOCK, 0, ctx->d_level_entities.as<uint32_t>() + off, ctx->d_level_col.as<uint32_t>() + off, \
                 run.e0, run.c0, cnt, (int)(k == 0), ctx->d_parent.as<uint32_t>(), ctx->trs_pos(), ctx->trs_rot(), ctx->trs_scl(),                                           \
                 ctx->d_world.as<float4>(), ctx->d_col_local.as<Aabb24>(), ctx->d_world_aabb.as<LeafRec>(), scratch_bounds(ctx))
            if (run.dense && !ctx->no_dense_refit) B2_REFIT_LEVEL(true);
            else B2_REFIT_LEVEL(false);
#undef B2_REFIT_LEVEL
        }
        PHASE(ctx);
        return B2COL_OK;
    }
    if (propagate) GUARD(enqueue_propagate(ctx));
    PHASE(ctx);
    LAUNCH_FRAME(ctx, k_refit, cdiv(n, REFIT_BLOCK), REFIT_BLOCK, 0, n, ctx->d_col_entity.as<uint32_t>(), ctx->d_col_local.as<Aabb24>(), ctx->d_world.as<float4>(),
           ctx->d_world_aabb.as<LeafRec>(), scratch_bounds(ctx));
    PHASE(ctx);
    return B2COL_OK;
}

inline size_t lookback_words_per_pass(uint32_t tiles) { return ((size_t)tiles + (tiles + LB_GROUP_SMALL - 1) / LB_GROUP_SMALL) * RS_RADIX; }  // tile counts + group status

// The frame's kernels, nothing else (no allocation, no memset, no synchronisation): capturable.
//   full:  [propagate+refit per level | propagate per level, refit] , Morton, radix passes, hierarchy
//   !full: [ ... same transform/refit ... ], header, hierarchy on the sorted keys of the last full build
b2col_status enqueue_frame(b2col_ctx* ctx, bool full, uint32_t n_records = 0)
{
    // n_records != 0: the world records (and the scene bounds, in the blob header) were received from another rank
    // (b2col_comm_replicate_records): Morton + sort + hierarchy only
    const uint32_t n = n_records ? n_records : ctx->n_colliders;
    bool tail_done = false;  // the refit kernel's last block has written the header (and, in a full build, keys + histograms)
    if (n_records) LAUNCH_FRAME(ctx, k_bounds_from_header, 1, 32, 0, (const BlobHeader*)blob_hdr_n(ctx), scratch_bounds(ctx));
    else GUARD(enqueue_transform_refit(ctx, full, &tail_done));  // one phase when fused, two otherwise
    if (!full) {
        if (!tail_done) LAUNCH_FRAME(ctx, k_header, 1, 64, 0, n, blob_hdr(ctx), scratch_bounds(ctx), ctx->T, ctx->d_key_ctr.as<uint32_t>());
        GUARD(launch_build(ctx, n, ctx->sorted_side));
        PHASE(ctx);
        return B2COL_OK;
    }
    const uint32_t tiles = cdiv(n, RS_TILE);
    const size_t lb_words_per_pass = lookback_words_per_pass(tiles);
    uint32_t* ghist = scratch_ghist(ctx);
    const bool packed = 3 * ctx->bits - ctx->drop <= 32 && !ctx->no_packed_sort;  // key << 32 | index in one word through the sort
    const bool msd = ctx->sort_msd;  // chosen by b2col_update
    if (!tail_done) {
        // 4 blocks per SM up to 2M colliders (fewer end-of-block atomics on the shared counters: -4 us at 1M), 8 beyond (the
        // stream dominates: 125 -> 119 us at 10M)
        const int per_sm = ctx->morton_blocks_per_sm > 0 ? ctx->morton_blocks_per_sm : (n <= (1u << 21) ? 4 : 8);
        const uint32_t grid = std::min<uint32_t>(cdiv(n, MORTON_BLOCK), (uint32_t)(ctx->n_sm * per_sm * 256 / MORTON_BLOCK));  // per_sm counts 256-thread blocks
        LAUNCH_FRAME(ctx, k_morton, grid, MORTON_BLOCK, 0, n, ctx->d_world_aabb.as<LeafRec>(), blob_hdr(ctx), scratch_bounds(ctx), 3 * ctx->bits - ctx->drop, ctx->passes,
               msd ? ctx->passes - 1 : 0, (int)packed, ctx->T, ctx->d_keys[0].as<uint64_t>(), ctx->d_vals[0].as<uint32_t>(), ghist, ctx->d_key_ctr.as<uint32_t>(),
               static_cast<uint4*>(ctx->d_key_stat));
    }
    PHASE(ctx);  // morton
    if (msd) {
        // MSD pass on the top digit, then every bucket finished inside one block (b2col_sort.cuh): two launches instead of four
        const int top = ctx->passes - 1;
        const uint32_t* gh = ghist + top * RS_RADIX;
        uint32_t* lb = scratch_lookback(ctx);
        uint32_t* tk = scratch_tickets(ctx);
        const uint64_t* kin = ctx->d_keys[0].as<uint64_t>();
        uint64_t* kout = ctx->d_keys[1].as<uint64_t>();
        const uint32_t* vnone = nullptr;
        uint32_t* vnone_out = nullptr;
        const int sh = 32 + top * RS_BITS;
        if (tiles >= 128u) LAUNCH_FRAME(ctx, (k_onesweep<1, LB_GROUP_SMALL>), tiles, RS_THREADS, 0, kin, vnone, kout, vnone_out, n, tiles, sh, gh, lb, tk);
        else LAUNCH_FRAME(ctx, (k_onesweep<1, LB_GROUP_LARGE>), tiles, RS_THREADS, 0, kin, vnone, kout, vnone_out, n, tiles, sh, gh, lb, tk);
        PHASE(ctx);
        LAUNCH_FRAME(ctx, k_local_sort, RS_RADIX, LS_THREADS, sizeof(LocalSortSmem), ctx->d_keys[1].as<uint64_t>(), ctx->d_keys[0].as<uint64_t>(), ctx->d_keys[0].as<uint64_t>(),
                     ctx->d_vals[0].as<uint32_t>(), gh, scratch_tickets(ctx) + 12);
        CU(cudaMemcpyAsync(ctx->h_sort_stat, scratch_tickets(ctx) + 12, 4, cudaMemcpyDeviceToHost, ctx->stream));  // largest bucket: read by the NEXT update
        PHASE(ctx);
        ctx->sorted_side = 0;
        GUARD(launch_build(ctx, n, 0));
        PHASE(ctx);  // last: build
        return B2COL_OK;
    }
    int side = 0;
    for (int p = 0; p < ctx->passes; ++p) {
        const uint64_t* kin = ctx->d_keys[side].as<uint64_t>();
        const uint32_t* vin = ctx->d_vals[side].as<uint32_t>();
        uint64_t* kout = ctx->d_keys[side ^ 1].as<uint64_t>();
        uint32_t* vout = ctx->d_vals[side ^ 1].as<uint32_t>();
        const uint32_t* gh = ghist + p * RS_RADIX;
        uint32_t* lb = scratch_lookback(ctx) + (size_t)p * lb_words_per_pass;
        uint32_t* tk = scratch_tickets(ctx) + p;
        const int sh = p * RS_BITS + (packed ? 32 : 0);
        const int mode = !packed ? 0 : (p + 1 < ctx->passes ? 1 : 2);
#define B2_SORT_PASS(M, G) LAUNCH_FRAME(ctx, (k_onesweep<M, G>), tiles, RS_THREADS, 0, kin, vin, kout, vout, n, tiles, sh, gh, lb, tk)
        if (tiles >= 128u) {
            if (mode == 0) B2_SORT_PASS(0, LB_GROUP_SMALL);
            else if (mode == 1) B2_SORT_PASS(1, LB_GROUP_SMALL);
            else B2_SORT_PASS(2, LB_GROUP_SMALL);
        }
        else {
            if (mode == 0) B2_SORT_PASS(0, LB_GROUP_LARGE);
            else if (mode == 1) B2_SORT_PASS(1, LB_GROUP_LARGE);
            else B2_SORT_PASS(2, LB_GROUP_LARGE);
        }
#undef B2_SORT_PASS
        side ^= 1;
        PHASE(ctx);  // radix pass p
    }
    ctx->sorted_side = side;
    GUARD(launch_build(ctx, n, side));
    PHASE(ctx);  // last: build
    return B2COL_OK;
}

// enqueue_frame, through a CUDA graph when the context asks for one: captured once per (shape, buffers) and replayed
// — one launch per frame instead of 7-10, and no host-side gaps between the short kernels of the sort.
b2col_status run_frame(b2col_ctx* ctx, bool full)
{
    if (!ctx->use_graph || ctx->phase_timing) return enqueue_frame(ctx, full);
    b2col_ctx::GraphSlot& g = ctx->graphs[full ? 0 : 1];
    const bool propagate = !ctx->world_external && !ctx->world_fresh;
    const uint64_t key[4] = {ctx->buffers_gen, ctx->shape_gen,
                             (uint64_t)propagate | ((uint64_t)(ctx->fusable && !ctx->no_fuse) << 1) | ((uint64_t)ctx->passes << 8) | ((uint64_t)ctx->bits << 16) |
                                 ((uint64_t)ctx->drop << 24) | ((uint64_t)ctx->sorted_side << 32) | ((uint64_t)ctx->sort_msd << 33) |
                                 ((uint64_t)(full && ctx->use_fused_keys) << 34) | ((full && ctx->use_fused_keys ? ctx->plan_state.gen & 0xFFFFFFull : 0ull) << 35),
                             ((uint64_t)ctx->n_colliders << 32) | ctx->n_entities};
    if (g.exec && memcmp(g.key, key, sizeof key) == 0) {
        CU(cudaGraphLaunch(g.exec, ctx->stream));
        ctx->launches += g.launches;
        ++ctx->graph_replays;
        if (full) ctx->sorted_side = ctx->sort_msd ? 0 : (ctx->passes & 1);
        ctx->scratch_clean = true;
        return B2COL_OK;
    }
    if (g.exec) {
        cudaGraphExecDestroy(g.exec);
        g.exec = nullptr;
    }
    const uint64_t before = ctx->launches;
    CU(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    const b2col_status st = enqueue_frame(ctx, full);
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
    if (st != B2COL_OK) {
        if (graph) cudaGraphDestroy(graph);
        return st;
    }
    if (e != cudaSuccess || !graph) return fail(ctx, B2COL_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
    const cudaError_t ei = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) {
        g.exec = nullptr;
        return fail(ctx, B2COL_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ei));
    }
    g.launches = ctx->launches - before;
    memcpy(g.key, key, sizeof key);
    CU(cudaGraphLaunch(g.exec, ctx->stream));
    return B2COL_OK;
}

}  // namespace

// Plan policy of the fused key generation (k_refit_keys): this frame's keys come from a Morton plan made BEFORE the frame's boxes
// exist, from the centroid bounds the device reported for earlier frames (h_key_stat[4..11], written by whichever kernel
// finished that frame's header).  The plan is kept while the reported bounds stay inside its own and span more than half of
// them on every axis; otherwise it is re-made around the reported bounds widened by 1/32 of their extent per side.  A plan is
// only about tree quality — centroids outside it clamp to its edge cells — so a scene that outruns it costs one frame of
// slightly coarser keys at the rim.  No report for this collider count yet (the first frames of a collider set), a
// communicator (replicas rebuild from the exact bounds in the header), the MSD pipeline, the list-per-entity kernel or a
// scene the fused kernel cannot take: the exact plan, i.e. k_morton after the refit.
static void choose_morton_plan(b2col_ctx* ctx, uint32_t n)
{
    ctx->use_fused_keys = false;
    const bool propagate = !ctx->world_external && !ctx->world_fresh;
    if (!ctx->fuse_keys || !ctx->refit_stride || !ctx->h_key_stat || ctx->comm || ctx->sort_msd || !(propagate && ctx->fusable && !ctx->no_fuse)) return;
    const volatile uint32_t* st = ctx->h_key_stat;
    uint32_t w[8];
    for (int k = 0; k < 8; ++k) w[k] = st[4 + k];
    if (w[3] != n || w[7] != n) return;
    float amin[3], amax[3];
    for (int k = 0; k < 3; ++k) {
        amin[k] = ordered_to_float(w[k]);
        amax[k] = ordered_to_float(w[4 + k]);
    }
    b2col_ctx::PlanState& ps = ctx->plan_state;
    const int key_bits = 3 * ctx->bits - ctx->drop;
    bool remake = !ps.valid || ps.n != n || ps.key_bits != key_bits;
    for (int k = 0; k < 3 && !remake; ++k) {
        if (!(amin[k] <= amax[k])) continue;  // no finite centroid on this axis
        const float pext = ps.pmax[k] - ps.pmin[k], aext = amax[k] - amin[k];
        if (!(ps.pmin[k] <= ps.pmax[k]) || amin[k] < ps.pmin[k] || amax[k] > ps.pmax[k] || !(pext <= 2.0f * aext || pext == 0.0f)) remake = true;
    }
    if (remake) {
        for (int k = 0; k < 3; ++k) {
            const float margin = (amin[k] <= amax[k]) ? (amax[k] - amin[k]) * 0.03125f : 0.0f;
            ps.pmin[k] = (margin < B2_INF) ? amin[k] - margin : amin[k];
            ps.pmax[k] = (margin < B2_INF) ? amax[k] + margin : amax[k];
        }
        ps.plan = make_morton_plan(ps.pmin, ps.pmax, key_bits);
        ps.key_bits = key_bits;
        ps.n = n;
        ps.valid = true;
        ++ps.gen;
    }
    ctx->use_fused_keys = true;
}

// buffers, key width and scratch of a full build of n colliders (everything but the blob, which the callers size first)
static b2col_status prepare_build_inner(b2col_ctx* ctx, uint32_t n)
{
    GUARD(ensure(ctx, ctx->d_world_aabb, (size_t)n * sizeof(LeafRec)));  // 32-byte records {world AABB, entity, collider index}
    GUARD(ensure(ctx, ctx->d_keys[0], (size_t)n * 8));
    GUARD(ensure(ctx, ctx->d_keys[1], (size_t)n * 8));
    GUARD(ensure(ctx, ctx->d_vals[0], (size_t)n * 4));
    GUARD(ensure(ctx, ctx->d_vals[1], (size_t)n * 4));
    GUARD(ensure(ctx, ctx->d_xch, (size_t)n * sizeof(Xch)));
    if (ctx->d_xch.p != ctx->xch_ready) {  // new allocation: all arrival words "nobody has arrived" (k_build leaves them so)
        CU(cudaMemsetAsync(ctx->d_xch.p, 0xFF, ctx->d_xch.cap, ctx->stream));
        ctx->xch_ready = ctx->d_xch.p;
    }
    // Morton key width: the radix sort costs one pass per 8 key bits, and at ~1M colliders each pass is
    // latency-bound (~20 us), so the key is only as wide as the collider count needs: 32 bits (11,11,10 per
    // axis, 4.3e9 cells) up to 2M colliders, 40 bits up to 32M, 48 beyond; an explicit
    // morton_bits_per_axis overrides (3*bits key bits).
    if (ctx->cfg_bits > 0) {
        ctx->bits = ctx->cfg_bits;
        ctx->drop = 0;
    }
    else {
        const int static_bits = n <= (1u << 21) ? 32 : (n <= (1u << 25) ? 40 : 48);
        int key_bits = ctx->env_key_bits > 0 ? ctx->env_key_bits : static_bits;
        // Adaptive width.  One radix pass costs ~15 us at 1M colliders and the last 8 key bits only order leaves that share a cell
        // of the grid 8 bits coarser: where few do (uniform and city-like scenes: 6 % of the leaves at 1M colliders and 24 bits,
        // traversal rate unchanged) the pass buys nothing; where many do (clustered scenes: 2M colliders traced at 1540 instead of
        // 4050 Mrays/s with 24-bit keys, profiles/r2_experiments.txt) every bit counts.  The hierarchy pass counts both kinds of
        // collision (k_build: leaves inside runs of more than KEY_RUN equal keys; isolated pairs cost nothing — 23 % of the leaves of
        // a 1M-collider hierarchy scene pair up at 24 bits and it traces exactly as fast as at 32); a full build narrows by one pass
        // while fewer than 1/32 of the leaves would sit in such runs and widens by one — up to 48 bits — when more than 1/8 do
        // (2M clustered colliders at 32 bits: 56 % of the leaves; rays aimed at the clusters trace 34 % faster at 40 bits, for 1.5 %
        // more update time).  The first builds of a collider set use the static width (the statistic arrives one frame late).
        const bool adaptive = ctx->adaptive_keys && ctx->env_key_bits == 0 && ctx->h_key_stat && ctx->build_variant == 0 && !ctx->comm;
        if (!adaptive) {
            ctx->adapt_key_bits = 0;
        }
        else {
            if (ctx->adapt_key_bits == 0) ctx->adapt_key_bits = static_bits;
            const volatile uint32_t* st = ctx->h_key_stat;
            const uint32_t dup_cur = st[0], dup_narrow = st[1], leaves = st[2], at_bits = st[3];
            if (leaves == n && (int)at_bits == ctx->adapt_key_bits) {  // the statistic describes this collider count at the width in use
                if ((uint64_t)dup_narrow * 32u < n && ctx->adapt_key_bits - 8 >= 16) ctx->adapt_key_bits -= 8;
                else if ((uint64_t)dup_cur * 8u > n && ctx->adapt_key_bits + 8 <= 48) ctx->adapt_key_bits += 8;
            }
            key_bits = ctx->adapt_key_bits;
        }
        ctx->bits = (key_bits + 2) / 3;
        ctx->drop = 3 * ctx->bits - key_bits;
    }
    ctx->passes = (3 * ctx->bits - ctx->drop + RS_BITS - 1) / RS_BITS;
    // MSD + block-local sort: packed 32-bit keys (exactly four digits), a mean bucket comfortably inside a block, and
    // the largest bucket of the last MSD frame inside one too (a crowded scene falls back to LSD from the next frame
    // on; the oversized bucket itself is still sorted exactly, only slowly)
    {
        const bool packed32 = 3 * ctx->bits - ctx->drop == 32 && !ctx->no_packed_sort;
        const bool fits = ctx->h_sort_stat && (uint64_t)n * 4u <= (uint64_t)LS_TILE * RS_RADIX * 3u && *ctx->h_sort_stat <= (uint32_t)LS_TILE;  // mean <= 3/4 tile
        ctx->sort_msd = ctx->h_sort_stat && packed32 && n >= 4096u && (ctx->sort_policy == 2 || (ctx->sort_policy == 0 && fits));
    }
    const uint32_t tiles = cdiv(n, RS_TILE);
    const size_t lb_words_per_pass = lookback_words_per_pass(tiles);
    GUARD(prepare_scratch(ctx, (SCRATCH_BOUNDS_WORDS + (size_t)RS_MAX_PASSES * RS_RADIX + 16 + (size_t)ctx->passes * lb_words_per_pass) * 4));
    return B2COL_OK;
}
static b2col_status prepare_build(b2col_ctx* ctx, uint32_t n)
{
    const void* before[3] = {ctx->d_world_aabb.p, ctx->d_keys[0].p, ctx->d_vals[0].p};
    const void* before2[2] = {ctx->d_keys[1].p, ctx->d_vals[1].p};
    const b2col_status st = prepare_build_inner(ctx, n);
    if (before[0] != ctx->d_world_aabb.p || before[1] != ctx->d_keys[0].p || before[2] != ctx->d_vals[0].p || before2[0] != ctx->d_keys[1].p ||
        before2[1] != ctx->d_vals[1].p) {  // the sorted keys / world records of the last build are gone: no refit, no export of the old tree
        ctx->topology_valid = false;
        ctx->exact_nodes_valid = false;
    }
    return st;
}

b2col_status b2col_update(b2col_ctx* ctx)
{
    if (!ctx) return B2COL_ERR_INVALID;
    if (!ctx->have_colliders) return fail(ctx, B2COL_ERR_STATE, "b2col_set_colliders has not been called");
    if (!ctx->world_external && !(ctx->have_hierarchy && ctx->have_trs))
        return fail(ctx, B2COL_ERR_STATE, "need b2col_set_hierarchy + b2col_set_trs, or b2col_set_world_matrices");
    CU(cudaSetDevice(ctx->device));
    const uint32_t E = ctx->n_entities, n = ctx->n_colliders;
    ctx->n_phase = 0;
    PHASE(ctx);  // 0: start

    // ---- validate, then allocate the frame's buffers.  Nothing that describes the CURRENT tree (have_tree, tree_leaves,
    // blob_bytes, topology_valid) changes before the new frame is enqueued, except where an allocation has already
    // destroyed the old tree: a failed update leaves the previous tree answering queries, or none at all — never a
    // half-described one.
    if (n != 0 && ctx->max_col_entity > E)
        return fail(ctx, B2COL_ERR_INVALID, "a collider refers to entity %u but only %u entities have transforms", ctx->max_col_entity, E);
    const size_t blob_bytes = blob_size_for(n);
    {
        const void* before = ctx->d_blob.p;
        const b2col_status st = ensure(ctx, ctx->d_blob, blob_bytes);
        if (ctx->d_blob.p != before || st != B2COL_OK) {  // the old blob is gone (ensure frees before it allocates)
            ctx->have_tree = false;
            ctx->topology_valid = false;
            ctx->exact_nodes_valid = false;
            ctx->tree_leaves = 0;
        }
        GUARD(st);
    }
    if (n == 0) {
        if (!ctx->world_external && !ctx->world_fresh) GUARD(enqueue_propagate(ctx));  // world matrices are still the caller's to read back
        ctx->have_tree = false;
        CU(cudaMemsetAsync(ctx->d_blob.p, 0, sizeof(BlobHeader), ctx->stream));
        ctx->blob_bytes = blob_bytes;
        ctx->tree_leaves = 0;
        ctx->have_tree = true;  // empty tree: header says n_leaves = 0, every query misses
        ctx->topology_valid = true;
        ctx->exact_nodes_valid = false;
        return B2COL_OK;
    }
    GUARD(prepare_build(ctx, n));
    if (!ctx->world_external && !ctx->world_fresh) GUARD(prepare_fuse_map(ctx));
    if (!ctx->world_external) GUARD(ensure(ctx, ctx->d_world, ((size_t)E + 1) * 64));
    choose_morton_plan(ctx, n);
    PHASE(ctx);  // 1: set-up
    // from here on the blob is being overwritten: until the frame is enqueued in full there is no tree
    ctx->have_tree = false;
    ctx->topology_valid = false;
    ctx->exact_nodes_valid = false;
    GUARD(run_frame(ctx, true));
    ctx->blob_bytes = blob_bytes;
    ctx->tree_leaves = n;
    ctx->tree_key_bits = 3 * ctx->bits - ctx->drop;
    ctx->tree_fused_keys = ctx->use_fused_keys;
    ctx->have_tree = true;
    ctx->topology_valid = true;
    ++ctx->full_builds;
    return B2COL_OK;
}

b2col_status b2col_refit(b2col_ctx* ctx)
{
    if (!ctx) return B2COL_ERR_INVALID;
    if (!ctx->topology_valid || !ctx->have_tree || ctx->tree_leaves != ctx->n_colliders)
        return fail(ctx, B2COL_ERR_STATE, "b2col_refit needs a tree built by b2col_update on this ctx for the current collider set");
    if (!ctx->world_external && !(ctx->have_hierarchy && ctx->have_trs))
        return fail(ctx, B2COL_ERR_STATE, "need b2col_set_hierarchy + b2col_set_trs, or b2col_set_world_matrices");
    CU(cudaSetDevice(ctx->device));
    const uint32_t n = ctx->n_colliders;
    ctx->n_phase = 0;
    if (n == 0) {
        if (!ctx->world_external && !ctx->world_fresh) GUARD(enqueue_propagate(ctx));
        return B2COL_OK;
    }
    if (ctx->max_col_entity > ctx->n_entities)
        return fail(ctx, B2COL_ERR_INVALID, "a collider refers to entity %u but only %u entities have transforms", ctx->max_col_entity, ctx->n_entities);
    // Same tree shape, new boxes: the hierarchy is a pure function of the SORTED KEYS of the last full build, which
    // are still in d_keys/d_vals[sorted_side], so k_build re-emits the same nodes (same indices, same leaf order —
    // the order that decides tie winners) around the boxes the refit has just written.  Skips Morton + the radix sort.
    GUARD(prepare_scratch(ctx, ctx->zero_bytes));
    if (!ctx->world_external && !ctx->world_fresh) GUARD(prepare_fuse_map(ctx));
    ctx->exact_nodes_valid = false;
    if (!ctx->world_external) GUARD(ensure(ctx, ctx->d_world, ((size_t)ctx->n_entities + 1) * 64));
    PHASE(ctx);
    {
        const b2col_status st = run_frame(ctx, false);
        if (st != B2COL_OK) {  // the blob may be half re-bounded
            ctx->have_tree = false;
            ctx->topology_valid = false;
            return st;
        }
    }
    ++ctx->refits;
    return B2COL_OK;
}

b2col_status b2col_raycast_device(b2col_ctx* ctx, const b2col_ray* d_rays, size_t n, b2col_raycast* d_out, int mode)
{
    if (!ctx || (n && (!d_rays || !d_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    return enqueue_trace(ctx, d_rays, n, d_out, mode);
}

// b2col_raycast_batch in two halves, so that one host thread can keep the pipelines of several GPUs in flight
// (b2col_mgpu_raycast_batch): _begin enqueues every copy and kernel of the batch, _end waits for them.
// `compact`: results are 12-byte b2col_hit records instead of 36-byte Raycasts (h_out_any points at whichever it is).
static b2col_status raycast_batch_begin(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, void* h_out_any, int mode, bool compact = false)
{
    const size_t out_words = compact ? 3 : 9, out_bytes = 4 * out_words;
    b2col_raycast* const h_out = static_cast<b2col_raycast*>(h_out_any);  // byte arithmetic below goes through h_out_b
    char* const h_out_b = static_cast<char*>(h_out_any);
    if (compact && mode != B2COL_MODE_LBVH) return fail(ctx, B2COL_ERR_INVALID, "compact results exist for the production tree only");
    if (!ctx || (n && (!h_rays || !h_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    ctx->pending = b2col_ctx::PendingBatch();
    if (n == 0) return B2COL_OK;
    CU(cudaSetDevice(ctx->device));
    if (!compact && mode == B2COL_MODE_LBVH && n <= b2col_ctx::SMALL_MAX && ctx->have_tree && ctx->tree_leaves > 0 && !ctx->no_small_path) {
        if (!ctx->h_small) {
            CU(cudaHostAlloc(&ctx->h_small, b2col_ctx::SMALL_MAX * (sizeof(b2col_ray) + sizeof(b2col_raycast)), cudaHostAllocMapped));
            CU(cudaHostGetDevicePointer(&ctx->d_small, ctx->h_small, 0));
            GUARD(ensure(ctx, ctx->d_small_counters, b2col_ctx::SMALL_SLOTS * 8));
        }
        if (ctx->small_next >= b2col_ctx::SMALL_SLOTS) {  // every slot used once: zero them all again (once per SMALL_SLOTS calls)
            CU(cudaMemsetAsync(ctx->d_small_counters.p, 0, b2col_ctx::SMALL_SLOTS * 8, ctx->stream));
            ctx->small_next = 0;
        }
        b2col_ray* hr = static_cast<b2col_ray*>(ctx->h_small);
        b2col_raycast* ho = reinterpret_cast<b2col_raycast*>(hr + b2col_ctx::SMALL_MAX);
        const b2col_ray* dr = static_cast<const b2col_ray*>(ctx->d_small);
        b2col_raycast* dout = reinterpret_cast<b2col_raycast*>(const_cast<b2col_ray*>(dr) + b2col_ctx::SMALL_MAX);
        memcpy(hr, h_rays, n * sizeof(b2col_ray));
        LAUNCH(ctx, (k_trace_lbvh<false, false>), 1, TRACE_BLOCK, dr, (unsigned long long)n, dout, blob_hdr(ctx), blob_qnodes(ctx), blob_leaves(ctx, ctx->tree_leaves), ctx->T,
               ctx->d_small_counters.as<unsigned long long>() + ctx->small_next++, ctx->fetch_threshold);
        CU(cudaStreamSynchronize(ctx->stream));
        memcpy(h_out, ho, n * sizeof(b2col_raycast));
        return B2COL_OK;
    }
    GUARD(ensure(ctx, ctx->d_rays, n * sizeof(b2col_ray)));
    GUARD(ensure(ctx, ctx->d_hits, n * out_bytes));
    // first chunk: e2e_first_chunk for large batches, halved down to e2e_min_chunk until the batch holds four of them: a
    // 256K-ray call runs as 64K + 128K + 64K instead of upload -> trace -> download back to back (300 -> 384 Mrays/s, 512K:
    // 521 -> 592; chunks below 64K cost more in launches and events than the overlap returns — tools/tune_e2e_small.py)
    size_t first_chunk = ctx->e2e_first_chunk;
    while (first_chunk > ctx->e2e_min_chunk && n < 4 * first_chunk) first_chunk >>= 1;
    if (mode == B2COL_MODE_LBVH && n >= 4 * first_chunk) {
        // Pipeline in chunks over three streams so that the upload of chunk c+1, the trace of chunk c and the
        // download of chunk c-1 overlap (PCIe is full duplex).  The download is the long pole (36 of the 60 bytes per
        // ray), so the schedule is about starting it early: chunk sizes ramp up from e2e_first_chunk by doubling to
        // e2e_chunk — the first results leave the device after ~0.2 ms instead of after a full-size chunk.  Needs
        // page-locked host buffers to actually overlap; with pageable memory the copies serialise but stay correct.
        if (!ctx->stream_in) {
            CU(cudaStreamCreateWithFlags(&ctx->stream_in, cudaStreamNonBlocking));
            CU(cudaStreamCreateWithFlags(&ctx->stream_out, cudaStreamNonBlocking));
        }
        b2col_ray* d_r = ctx->d_rays.as<b2col_ray>();
        char* d_h = ctx->d_hits.as<char>();
        // device->host leg: copy engine, or SM stores when the buffer is page-locked (device-accessible through UVA) and
        // 16-byte aligned.  Big calls time themselves: the first uses the copy engine, the second the SMs, later ones the
        // faster of the two (every 32nd call tries the other again: what else runs on the host changes).
        uint32_t* h_out_dev = nullptr;
        {
            cudaPointerAttributes pa{};
            if (cudaPointerGetAttributes(&pa, h_out) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer && ((uintptr_t)pa.devicePointer & 15u) == 0)
                h_out_dev = static_cast<uint32_t*>(pa.devicePointer);
            else
                cudaGetLastError();
        }
        const bool timed = n >= ((size_t)1 << 20);
        int d2h = 0;
        if (h_out_dev) {
            if (ctx->d2h_policy >= 0) d2h = ctx->d2h_policy;
            else if (timed) {
                const uint64_t k = ctx->d2h_calls;
                const int best = ctx->d2h_ns_per_ray[1] < ctx->d2h_ns_per_ray[0] ? 1 : 0;
                d2h = k == 0 ? 0 : (k == 1 ? 1 : ((k & 31u) == 0 ? best ^ 1 : best));
            }
            else d2h = (ctx->d2h_calls >= 2 && ctx->d2h_ns_per_ray[1] < ctx->d2h_ns_per_ray[0]) ? 1 : 0;
        }
        const auto t_begin = std::chrono::steady_clock::now();
        size_t off = 0, m = first_chunk, c = 0;
        const bool two = ctx->e2e_streams >= 2 && ctx->have_tree && ctx->tree_leaves > 0;
        if (two) {
            if (!ctx->stream_trace2) {
                CU(cudaStreamCreateWithFlags(&ctx->stream_trace2, cudaStreamNonBlocking));
                CU(cudaEventCreateWithFlags(&ctx->pipe_start_ev, cudaEventDisableTiming));
            }
            const size_t max_chunks = n / first_chunk + 2;
            GUARD(ensure(ctx, ctx->d_pipe_counters, max_chunks * 16));
            CU(cudaMemsetAsync(ctx->d_pipe_counters.p, 0, max_chunks * 16, ctx->stream));
            CU(cudaEventRecord(ctx->pipe_start_ev, ctx->stream));  // behind the tree's build and the counters' memset
            CU(cudaStreamWaitEvent(ctx->stream_trace2, ctx->pipe_start_ev, 0));
        }
        while (off < n) {
            m = std::min(m, n - off);
            if (n - off - m < first_chunk) m = n - off;  // no tiny tail chunk
            while (ctx->pipe_events.size() < 2 * (c + 1)) {
                cudaEvent_t ev;
                CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
                ctx->pipe_events.push_back(ev);
            }
            cudaEvent_t ev_in = ctx->pipe_events[2 * c], ev_k = ctx->pipe_events[2 * c + 1];
            CU(cudaMemcpyAsync(d_r + off, h_rays + off, m * sizeof(b2col_ray), cudaMemcpyHostToDevice, ctx->stream_in));
            CU(cudaEventRecord(ev_in, ctx->stream_in));
            cudaStream_t ts = (two && (c & 1)) ? ctx->stream_trace2 : ctx->stream;
            CU(cudaStreamWaitEvent(ts, ev_in, 0));
            GUARD(enqueue_trace(ctx, d_r + off, m, reinterpret_cast<b2col_raycast*>(d_h + off * out_bytes), mode, false, compact, two ? ts : nullptr,
                                two ? ctx->d_pipe_counters.as<unsigned long long>() + 2 * c : nullptr));
            CU(cudaEventRecord(ev_k, ts));
            CU(cudaStreamWaitEvent(ctx->stream_out, ev_k, 0));
            if (d2h == 1) {
                k_copy_out<<<2 * ctx->n_sm, 256, 0, ctx->stream_out>>>(reinterpret_cast<const uint32_t*>(d_h + off * out_bytes), h_out_dev + out_words * off, out_words * m);
                ++ctx->launches;
                if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) return fail(ctx, B2COL_ERR_CUDA, "launch of k_copy_out failed: %s", cudaGetErrorString(e));
            }
            else {
                CU(cudaMemcpyAsync(h_out_b + off * out_bytes, d_h + off * out_bytes, m * out_bytes, cudaMemcpyDeviceToHost, ctx->stream_out));
            }
            off += m;
            ++c;
            m = std::min(2 * m, ctx->e2e_chunk);
        }
        ctx->pending.active = true;
        ctx->pending.pipeline = true;
        ctx->pending.timed = timed;
        ctx->pending.has_dev_out = h_out_dev != nullptr;
        ctx->pending.d2h = d2h;
        ctx->pending.n = n;
        ctx->pending.t_begin = t_begin;
        return B2COL_OK;
    }
    CU(cudaMemcpyAsync(ctx->d_rays.p, h_rays, n * sizeof(b2col_ray), cudaMemcpyHostToDevice, ctx->stream));
    GUARD(enqueue_trace(ctx, ctx->d_rays.as<b2col_ray>(), n, ctx->d_hits.as<b2col_raycast>(), mode, false, compact));
    CU(cudaMemcpyAsync(h_out_b, ctx->d_hits.p, n * out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->pending.active = true;
    ctx->pending.mode = mode;
    ctx->pending.n = n;
    return B2COL_OK;
}

static b2col_status raycast_batch_end(b2col_ctx* ctx)
{
    if (!ctx->pending.active) return B2COL_OK;
    const b2col_ctx::PendingBatch pb = ctx->pending;
    ctx->pending.active = false;
    CU(cudaSetDevice(ctx->device));
    if (pb.pipeline) {
        CU(cudaStreamSynchronize(ctx->stream_out));
        CU(cudaStreamSynchronize(ctx->stream));
        if (ctx->stream_trace2) CU(cudaStreamSynchronize(ctx->stream_trace2));
        if (pb.timed && pb.has_dev_out && ctx->d2h_policy < 0) {
            const double ns = std::chrono::duration<double, std::nano>(std::chrono::steady_clock::now() - pb.t_begin).count() / (double)pb.n;
            double& est = ctx->d2h_ns_per_ray[pb.d2h];
            est = est == 0.0 ? ns : 0.5 * est + 0.5 * ns;
            ++ctx->d2h_calls;
        }
        return B2COL_OK;
    }
    CU(cudaStreamSynchronize(ctx->stream));
    if (pb.mode == B2COL_MODE_REFTOPO) {
        uint32_t flag = 0;
        CU(cudaMemcpy(&flag, ctx->d_flag.p, 4, cudaMemcpyDeviceToHost));
        if (flag) return fail(ctx, B2COL_ERR_OVERFLOW, "reference-topology traversal stack overflow (tree deeper than %d)", REFTOPO_STACK);
    }
    return B2COL_OK;
}

// ---- b2col_raycast_batch_static: the production trace with ALL of its device scratch in a file-scope array ------------------------
// For callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid it).  b2col_raycast_batch stages the
// rays and the results in device buffers the context allocates on demand (ctx->d_rays / d_hits) and takes the persistent
// kernel's work counter from another one; here all three live in ONE static array that every GPU which loads the library holds
// until exit: [work counter, 256 B | rays, 24 B each | Raycasts, 36 B each] for at most B2COL_STATIC_MAX_RAYS rays.
// In host code `g_static_scratch` is only a host-side stand-in for the symbol, never a device address: the address is asked
// for once per device, with that device current (static_scratch_base), and that pointer is what the copies and the kernel get.
// Every call on a device shares the array (it keeps the last call's bytes: the counter is cleared per call, rays and results are
// overwritten before they are read); calls that overlap on one GPU — two contexts, two host threads — race.
constexpr size_t STATIC_COUNTER_BYTES = 256;
constexpr size_t STATIC_RAYS_OFF = STATIC_COUNTER_BYTES;
constexpr size_t STATIC_HITS_OFF = STATIC_RAYS_OFF + (size_t)B2COL_STATIC_MAX_RAYS * sizeof(b2col_ray);
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_HITS_OFF + (size_t)B2COL_STATIC_MAX_RAYS * sizeof(b2col_raycast);
static_assert(STATIC_RAYS_OFF % 8 == 0 && STATIC_HITS_OFF % 4 == 0 && STATIC_SCRATCH_BYTES <= ((size_t)4 << 20), "static scratch layout (3.75 MiB per GPU)");
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static void* g_static_scratch_addr[STATIC_MAX_DEVICES];  // device address of g_static_scratch on device d (nullptr: not asked yet)
static std::mutex g_static_scratch_mutex;

// set-up, once per device: device address of the array on ctx's device (the current device on entry)
static b2col_status static_scratch_base(b2col_ctx* ctx, unsigned char** out)
{
    if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) return fail(ctx, B2COL_ERR_UNSUPPORTED, "device ordinal %d beyond the %d this entry point keeps addresses for", ctx->device, STATIC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    void*& slot = g_static_scratch_addr[ctx->device];
    if (!slot) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(ctx, B2COL_ERR_CUDA, "cudaGetSymbolAddress(g_static_scratch) on device %d failed: %s", ctx->device, cudaGetErrorString(e));
        }
        slot = p;
    }
    *out = static_cast<unsigned char*>(slot);
    return B2COL_OK;
}

b2col_status b2col_raycast_batch_static(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, b2col_raycast* h_out)
{
    if (!ctx || (n && (!h_rays || !h_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (n > (size_t)B2COL_STATIC_MAX_RAYS)
        return fail(ctx, B2COL_ERR_INVALID, "b2col_raycast_batch_static takes at most B2COL_STATIC_MAX_RAYS = %d rays per call (its device scratch is a fixed array), got %zu: split the batch or use b2col_raycast_batch",
                    (int)B2COL_STATIC_MAX_RAYS, n);
    if (n == 0) return B2COL_OK;
    CU(cudaSetDevice(ctx->device));
    if (!ctx->have_tree || ctx->tree_leaves == 0) {  // before the first update / empty tree: every query misses (reference :164-167)
        CU(cudaStreamSynchronize(ctx->stream));      // same ordering as the allocating path: behind whatever the context has enqueued
        memset(h_out, 0, n * sizeof(b2col_raycast));
        return B2COL_OK;
    }
    unsigned char* base = nullptr;
    GUARD(static_scratch_base(ctx, &base));
    unsigned long long* counter = reinterpret_cast<unsigned long long*>(base);
    b2col_ray* d_rays = reinterpret_cast<b2col_ray*>(base + STATIC_RAYS_OFF);
    b2col_raycast* d_out = reinterpret_cast<b2col_raycast*>(base + STATIC_HITS_OFF);
    CU(cudaMemsetAsync(counter, 0, 16, ctx->stream));  // the array keeps the last call's contents: the work counter must start at zero
    CU(cudaMemcpyAsync(d_rays, h_rays, n * sizeof(b2col_ray), cudaMemcpyHostToDevice, ctx->stream));
    GUARD(enqueue_trace(ctx, d_rays, n, d_out, B2COL_MODE_LBVH, false, false, ctx->stream, counter));  // the launch form that takes its counter from the caller: no ensure()
    CU(cudaMemcpyAsync(h_out, d_out, n * sizeof(b2col_raycast), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return B2COL_OK;
}

b2col_status b2col_raycast_hits_device(b2col_ctx* ctx, const b2col_ray* d_rays, size_t n, b2col_hit* d_out)
{
    if (!ctx || (n && (!d_rays || !d_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    return enqueue_trace(ctx, d_rays, n, reinterpret_cast<b2col_raycast*>(d_out), B2COL_MODE_LBVH, false, true);
}

b2col_status b2col_raycast_hits_batch(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, b2col_hit* h_out)
{
    if (!ctx || (n && (!h_rays || !h_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (n == 0) return B2COL_OK;
    GUARD(raycast_batch_begin(ctx, h_rays, n, h_out, B2COL_MODE_LBVH, true));
    return raycast_batch_end(ctx);
}

// Host side of the compact format: b2col_hit -> Raycast with the arithmetic of the device epilogue (b2col_bvh.cuh expand_hit; this
// translation unit's host code is compiled without FMA contraction), sharded over n_threads host threads.
b2col_status b2col_expand_hits(const b2col_ray* h_rays, const b2col_hit* h_hits, size_t n, b2col_raycast* h_out, int n_threads)
{
    if (n && (!h_rays || !h_hits || !h_out)) return B2COL_ERR_INVALID;
    auto work = [=](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i)
            expand_hit(reinterpret_cast<uint32_t*>(h_out + i), h_rays[i].origin, h_rays[i].direction, reinterpret_cast<const uint32_t*>(h_hits + i));
    };
    const int T = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(1, n_threads), n / 65536 + 1));
    if (T == 1) {
        work(0, n);
        return B2COL_OK;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < T; ++t) pool.emplace_back(work, n * (size_t)t / T, n * (size_t)(t + 1) / T);
    for (auto& th : pool) th.join();
    return B2COL_OK;
}

b2col_status b2col_raycast_batch(b2col_ctx* ctx, const b2col_ray* h_rays, size_t n, b2col_raycast* h_out, int mode)
{
    if (!ctx || (n && (!h_rays || !h_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (n == 0) return B2COL_OK;
    GUARD(raycast_batch_begin(ctx, h_rays, n, h_out, mode));
    return raycast_batch_end(ctx);
}


b2col_status b2col_upload_ref_bvh(b2col_ctx* ctx, const b2col_bitreenode* h_nodes, size_t n_nodes)
{
    if (!ctx || (n_nodes && !h_nodes)) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (n_nodes >= 0xffffffffull) return fail(ctx, B2COL_ERR_INVALID, "too many nodes");
    CU(cudaSetDevice(ctx->device));
    // validate child indices so that a malformed tree cannot make the kernel read out of bounds
    for (size_t i = 0; i < n_nodes; ++i) {
        const b2col_bitreenode& nd = h_nodes[i];
        if (nd.type1 > 2 || nd.type2 > 2) return fail(ctx, B2COL_ERR_INVALID, "node %zu has an invalid type", i);
        if ((nd.type1 == B2COL_NODE_BOUNDING_VOLUME && nd.index1 >= n_nodes) || (nd.type2 == B2COL_NODE_BOUNDING_VOLUME && nd.index2 >= n_nodes))
            return fail(ctx, B2COL_ERR_INVALID, "node %zu refers to a child outside the array", i);
    }
    // ... and that it is a TREE below its root (the last node): a node reachable twice means a cycle or a shared subtree, and a
    // cycle would keep k_trace_reftopo descending forever
    if (n_nodes) {
        std::vector<uint8_t> seen(n_nodes, 0);
        std::vector<uint32_t> todo;
        todo.push_back((uint32_t)(n_nodes - 1));
        seen[n_nodes - 1] = 1;
        while (!todo.empty()) {
            const b2col_bitreenode& nd = h_nodes[todo.back()];
            todo.pop_back();
            const uint32_t child[2] = {nd.index1, nd.index2};
            const uint8_t type[2] = {nd.type1, nd.type2};
            for (int k = 0; k < 2; ++k) {
                if (type[k] != B2COL_NODE_BOUNDING_VOLUME) continue;
                if (seen[child[k]]) return fail(ctx, B2COL_ERR_INVALID, "node %u is reachable twice from the root: not a tree", child[k]);
                seen[child[k]] = 1;
                todo.push_back(child[k]);
            }
        }
    }
    GUARD(ensure(ctx, ctx->d_ref_nodes, n_nodes * 60 + 4));
    if (n_nodes) {
        CU(cudaMemcpyAsync(ctx->d_ref_nodes.p, h_nodes, n_nodes * 60, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    ctx->n_ref_nodes = (uint32_t)n_nodes;
    return B2COL_OK;
}

b2col_status b2col_export_bitreenode(b2col_ctx* ctx, b2col_bitreenode* h_out, size_t capacity, size_t* n_out)
{
    if (!ctx || !n_out) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    *n_out = 0;
    if (!ctx->have_tree || ctx->tree_leaves == 0) return B2COL_OK;  // empty bvh_
    if (!ctx->topology_valid || ctx->tree_leaves != ctx->n_colliders || !ctx->d_world_aabb.p)
        return fail(ctx, B2COL_ERR_STATE, "export needs a tree built by b2col_update on this ctx (an imported tree has no keys / world records here)");
    CU(cudaSetDevice(ctx->device));
    const uint32_t n = ctx->tree_leaves;
    if (n == 1) {
        if (capacity < 1 || !h_out) return fail(ctx, B2COL_ERR_INVALID, "capacity too small");
        LeafRec rec;
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaMemcpy(&rec, blob_leaves(ctx, 1), sizeof rec, cudaMemcpyDeviceToHost));
        memset(h_out, 0, sizeof *h_out);
        memcpy(&h_out[0].box1, &rec, 24);
        h_out[0].index1 = rec.entity;
        h_out[0].type1 = B2COL_NODE_ENTITY;
        h_out[0].type2 = B2COL_NODE_EMPTY;
        *n_out = 1;
        return B2COL_OK;
    }
    if (!ctx->exact_nodes_valid) {  // full-precision nodes are only built when somebody asks for the reference-format tree
        GUARD(ensure(ctx, ctx->d_other_bound, (size_t)n * 4 + 4));
        GUARD(ensure(ctx, ctx->d_nodes_full, (size_t)(n - 1) * sizeof(Node)));
        CU(cudaMemsetAsync(ctx->d_other_bound.p, 0xFF, (size_t)n * 4 + 4, ctx->stream));
        LAUNCH(ctx, k_build_exact, cdiv(n, 256), 256, n, ctx->d_keys[ctx->sorted_side].as<uint64_t>(), ctx->d_vals[ctx->sorted_side].as<uint32_t>(),
               ctx->d_world_aabb.as<LeafRec>(), ctx->d_col_entity.as<uint32_t>(), ctx->d_nodes_full.as<Node>(), ctx->d_other_bound.as<uint32_t>(), blob_hdr(ctx),
               ctx->d_other_bound.as<uint32_t>() + n);
        ctx->exact_nodes_valid = true;
    }
    // candidates -> flags -> prefix sum -> compacted nodes, all on the device and in buffers the context keeps (a debug overlay
    // that exports every frame allocates nothing); the host only learns the node count before the one copy into h_out
    const size_t cand = 2 * (size_t)n - 1;
    const uint32_t tiles = cdiv(cand, XS_TILE);
    GUARD(ensure(ctx, ctx->d_export_cand, cand * 60));
    GUARD(ensure(ctx, ctx->d_export_out, cand * 60));
    GUARD(ensure(ctx, ctx->d_export_used, cand));
    GUARD(ensure(ctx, ctx->d_export_scan, (cand + tiles + 4) * 4));
    uint32_t* excl = ctx->d_export_scan.as<uint32_t>();
    uint32_t* tile_sums = excl + cand;
    uint32_t* total = tile_sums + tiles;
    CU(cudaMemsetAsync(ctx->d_export_used.p, 0, cand, ctx->stream));
    LAUNCH(ctx, k_export, cdiv(n, 256), 256, n, blob_hdr(ctx), ctx->d_nodes_full.as<Node>(), blob_leaves(ctx, n), ctx->d_export_cand.as<BiTreeNode60>(),
           ctx->d_export_used.as<uint8_t>());
    LAUNCH(ctx, k_export_scan_tiles, tiles, XS_THREADS, ctx->d_export_used.as<uint8_t>(), (uint32_t)cand, excl, tile_sums);
    LAUNCH(ctx, k_export_scan_sums, 1, 1024, tile_sums, tiles, total);
    LAUNCH(ctx, k_export_compact, cdiv(cand, 256), 256, (uint32_t)cand, ctx->d_export_cand.as<BiTreeNode60>(), ctx->d_export_used.as<uint8_t>(), excl, tile_sums,
           ctx->d_export_out.as<BiTreeNode60>());
    uint32_t kept = 0;
    CU(cudaMemcpyAsync(&kept, total, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    static_assert(sizeof(BiTreeNode60) == sizeof(b2col_bitreenode), "export node layout");
    if (kept > capacity || !h_out) return fail(ctx, B2COL_ERR_INVALID, "capacity %zu too small, need %u nodes", capacity, kept);
    CU(cudaMemcpyAsync(h_out, ctx->d_export_out.p, (size_t)kept * 60, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    *n_out = kept;
    return B2COL_OK;
}

b2col_status b2col_get_world_matrices(b2col_ctx* ctx, uint32_t first_entity, uint32_t count, float* h_out16)
{
    if (!ctx || (count && !h_out16)) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if ((size_t)first_entity + count > (size_t)ctx->n_entities + 1 || !ctx->d_world.p) return fail(ctx, B2COL_ERR_INVALID, "entity range out of bounds");
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(h_out16, ctx->d_world.as<float>() + 16 * (size_t)first_entity, (size_t)count * 64, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return B2COL_OK;
}

b2col_status b2col_get_world_aabbs(b2col_ctx* ctx, b2col_aabb* h_out)
{
    if (!ctx || (ctx->n_colliders && !h_out)) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_tree || ctx->tree_leaves != ctx->n_colliders) return fail(ctx, B2COL_ERR_STATE, "b2col_update has not run");
    if (!ctx->n_colliders) return B2COL_OK;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpy2DAsync(h_out, 24, ctx->d_world_aabb.p, sizeof(LeafRec), 24, ctx->n_colliders, cudaMemcpyDeviceToHost, ctx->stream));  // the AABB of each 32-byte record
    CU(cudaStreamSynchronize(ctx->stream));
    return B2COL_OK;
}

b2col_status b2col_read_device(b2col_ctx* ctx, const void* d_src, size_t bytes, void* h_dst)
{
    if (!ctx || (bytes && (!d_src || !h_dst))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!bytes) return B2COL_OK;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return B2COL_OK;
}

b2col_status b2col_get_leaf_order(b2col_ctx* ctx, uint32_t* h_out)
{
    if (!ctx || (ctx->n_colliders && !h_out)) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_tree || !ctx->topology_valid || ctx->tree_leaves != ctx->n_colliders)
        return fail(ctx, B2COL_ERR_STATE, "b2col_update has not run on this ctx for the tree it holds");
    if (!ctx->n_colliders) return B2COL_OK;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(h_out, ctx->d_vals[ctx->sorted_side].p, (size_t)ctx->n_colliders * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return B2COL_OK;
}

b2col_status b2col_get_trace_counters(b2col_ctx* ctx, uint64_t* node_visits, uint64_t* rays)
{
    if (!ctx) return B2COL_ERR_INVALID;
    if (node_visits) *node_visits = ctx->counted_visits;
    if (rays) *rays = ctx->counted_rays;
    return B2COL_OK;
}

b2col_status b2col_get_stats(b2col_ctx* ctx, b2col_stats* out)
{
    if (!ctx || !out) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    memset(out, 0, sizeof *out);
    out->n_entities = ctx->n_entities;
    out->n_colliders = ctx->n_colliders;
    out->n_levels = ctx->level_offsets.empty() ? 0 : (uint32_t)ctx->level_offsets.size() - 1;
    out->sort_passes = (uint32_t)ctx->passes;
    out->key_bits = (uint32_t)ctx->tree_key_bits;
    out->fused_keys = ctx->tree_fused_keys ? 1u : 0u;
    out->kernel_launches = ctx->launches;
    out->full_builds = ctx->full_builds;
    out->refits = ctx->refits;
    out->graph_replays = ctx->graph_replays;
    if (ctx->have_tree && ctx->d_blob.p) {
        CU(cudaSetDevice(ctx->device));
        BlobHeader h;
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaMemcpy(&h, ctx->d_blob.p, sizeof h, cudaMemcpyDeviceToHost));
        out->n_internal_nodes = h.n_nodes;
        out->root = h.root;
        out->bvh_bytes = ctx->blob_bytes;
        for (int k = 0; k < 3; ++k) {
            out->scene_min[k] = ordered_to_float(~h.bounds_min_inv[k]);
            out->scene_max[k] = ordered_to_float(h.bounds_max[k]);
        }
    }
    return B2COL_OK;
}

b2col_status b2col_bvh_blob(b2col_ctx* ctx, void** d_ptr, size_t* bytes)
{
    if (!ctx || !d_ptr || !bytes) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_tree) return fail(ctx, B2COL_ERR_STATE, "no tree: call b2col_update first");
    *d_ptr = ctx->d_blob.p;
    *bytes = ctx->blob_bytes;
    return B2COL_OK;
}

b2col_status b2col_bvh_export_blob(b2col_ctx* ctx, void* d_dst, size_t bytes)
{
    if (!ctx || !d_dst) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (!ctx->have_tree) return fail(ctx, B2COL_ERR_STATE, "no tree: call b2col_update first");
    if (bytes != ctx->blob_bytes) return fail(ctx, B2COL_ERR_INVALID, "destination holds %zu bytes, blob is %zu", bytes, ctx->blob_bytes);
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(d_dst, ctx->d_blob.p, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    return B2COL_OK;
}

// leaf count of a blob of `bytes` bytes (blob_size_for is strictly increasing); false if no count matches
static bool leaves_from_blob_size(size_t bytes, uint32_t* n_out)
{
    uint32_t lo = 0, hi = (1u << 29);
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (blob_size_for(mid) < bytes) lo = mid + 1;
        else hi = mid;
    }
    *n_out = lo;
    return blob_size_for(lo) == bytes;
}

b2col_status b2col_bvh_import_blob(b2col_ctx* ctx, const void* d_src, size_t bytes)
{
    if (!ctx || !d_src || bytes < sizeof(BlobHeader)) return fail(ctx, B2COL_ERR_INVALID, "bad blob");
    uint32_t n_leaves = 0;
    if (!leaves_from_blob_size(bytes, &n_leaves)) return fail(ctx, B2COL_ERR_INVALID, "%zu bytes is not the size of a BVH blob", bytes);
    CU(cudaSetDevice(ctx->device));
    // No host round trip on this path: the leaf count follows from the size, and the header (magic, leaf count) is copied
    // to page-locked memory behind the import and checked by the next b2col_sync.
    if (!ctx->h_import_hdr) CU(cudaHostAlloc(&ctx->h_import_hdr, sizeof(BlobHeader), cudaHostAllocDefault));
    {
        const void* before = ctx->d_blob.p;
        const b2col_status st = ensure(ctx, ctx->d_blob, bytes);
        if (ctx->d_blob.p != before || st != B2COL_OK) ctx->have_tree = false;
        GUARD(st);
    }
    ctx->have_tree = false;
    if (ctx->d_blob.p != d_src) CU(cudaMemcpyAsync(ctx->d_blob.p, d_src, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_import_hdr, ctx->d_blob.p, sizeof(BlobHeader), cudaMemcpyDeviceToHost, ctx->stream));
    ctx->import_check_leaves = n_leaves;
    ctx->import_check_pending = true;
    ctx->blob_bytes = bytes;
    ctx->tree_leaves = n_leaves;
    ctx->have_tree = true;
    ctx->topology_valid = false;  // a foreign tree: its sorted keys are not on this ctx
    ctx->exact_nodes_valid = false;
    return B2COL_OK;
}

// dev only (not in include/b2col.h): ms between consecutive phase events of the last b2col_update; returns the count
int b2col_debug_phase_ms(b2col_ctx* ctx, float* out, int cap)
{
    if (!ctx || !ctx->phase_timing || ctx->n_phase < 2) return 0;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    int k = 0;
    for (int i = 1; i < ctx->n_phase && k < cap; ++i, ++k) cudaEventElapsedTime(out + k, ctx->phase_ev[i - 1], ctx->phase_ev[i]);
    return k;
}

#ifdef B2_SORT_TIMING
int b2col_debug_sort_timing(unsigned long long* out, int n_tiles)  // dev builds only
{
    return (int)cudaMemcpyFromSymbol(out, b2::g_sort_timing, sizeof(unsigned long long) * 8 * (size_t)n_tiles);
}
#endif

b2col_status b2col_event_record(b2col_ctx* ctx, int slot)
{
    if (!ctx || slot < 0 || slot >= 16) return fail(ctx, B2COL_ERR_INVALID, "bad event slot");
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventRecord(ctx->events[slot], ctx->stream));
    return B2COL_OK;
}

b2col_status b2col_event_elapsed_ms(b2col_ctx* ctx, int a, int b, float* ms)
{
    if (!ctx || !ms || a < 0 || a >= 16 || b < 0 || b >= 16) return fail(ctx, B2COL_ERR_INVALID, "bad event slot");
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventSynchronize(ctx->events[b]));
    CU(cudaEventElapsedTime(ms, ctx->events[a], ctx->events[b]));
    return B2COL_OK;
}

}  // extern "C"

#include "b2col_comm.inl"
