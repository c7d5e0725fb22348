// The production traversal: k_trace_lbvh and everything only it uses (ray prologue, shared-memory layout, result
// staging).  Included by b2col.cu after DevOps / BlobHeader; a file of its own so that the ncu capture quoted as
// roofline.traffic (bench.py, tools/capture_traffic.py) can be tied to exactly the sources this kernel is compiled
// from: b2col_trace.inl, b2col_bvh.cuh, b2col_math.cuh.
// Ray prologue (GetRaycast :162 + the hoisted reciprocals of :65).  Deliberately NOT inlined: it runs
// three times per ray against ~100 node visits, and keeping its IEEE division / square-root sequences
// (the only FFMA-containing code on this path) out of the traversal loop saves registers there.
__device__ __noinline__ RayPre load_ray(const b2col_ray* rays, size_t i)
{
    const float2* rp = reinterpret_cast<const float2*>(rays + i);  // 24-byte records are 8-byte aligned
    const float2 a = __ldg(rp), b = __ldg(rp + 1), c = __ldg(rp + 2);
    const float o[3] = {a.x, a.y, b.x}, d[3] = {b.y, c.x, c.y};
    return ray_prepare(o, d);
}

constexpr int TRACE_BLOCK = 128;
#ifndef B2_SMEM_STACK
#define B2_SMEM_STACK 12
#endif
constexpr int TRACE_SMEM_STACK = B2_SMEM_STACK;  // deferred children per lane held in shared memory (even)
constexpr int TRACE_SPILL = TRACE_SMEM_STACK / 2;
static_assert(TRACE_SMEM_STACK % 2 == 0 && TRACE_SMEM_STACK >= 4, "spill granularity is half the shared-memory stack");
#ifndef B2_TRACE_MIN_BLOCKS
#define B2_TRACE_MIN_BLOCKS 10
#endif
#ifndef B2_TRACE_STEPS
#define B2_TRACE_STEPS 4  // node visits between two warp votes
#endif

// Shared memory of one trace block.  Everything is laid out [row][thread], so lane l always hits bank
// l whatever row it addresses: every access is one conflict-free wavefront.
#ifndef B2_STAGE_RAYS
#define B2_STAGE_RAYS 11
#endif
constexpr int STAGE_RAYS = B2_STAGE_RAYS;  // finished Raycasts a warp stages per round on their way out (flush_finished)
constexpr int STAGE_WORDS = 11;            // per staged ray: the nine words of the Raycast + its 64-bit index in the batch
struct TraceSmem {
    uint32_t node[TRACE_SMEM_STACK][TRACE_BLOCK];  // deferred far children ...
    float t[TRACE_SMEM_STACK][TRACE_BLOCK];        // ... and their entry tmin
    float ray[9][TRACE_BLOCK];                     // the lane's exact ray: origin, normalised direction, reciprocals
    uint32_t stage[TRACE_BLOCK / 32][STAGE_RAYS * STAGE_WORDS];
};
// 10 blocks per SM must fit the 196 KB shared-memory carve-out (10 x (18 912 + 1 024 reserved) = 199 360 <= 200 704): with the
// 228 KB carve-out only 28 KB of L1 remain for the node gathers and the kernel loses 30 % (measured, profiles/r2_experiments.txt:
// 32 staged rays per warp, 21.5 KB per block: 1867 -> 1300 Mrays/s)
#ifndef B2_ALLOW_BIG_SMEM  // dev variants only
static_assert((sizeof(TraceSmem) + 1024) * B2_TRACE_MIN_BLOCKS <= 196 * 1024, "trace blocks per SM no longer fit the 196 KB shared-memory carve-out");
#endif

// Overflow of the shared-memory stack (a lane more than TRACE_SMEM_STACK deferred children deep — rare):
// the older half moves to a local-memory array and comes back when the shared part runs empty.  Total depth
// is bounded by LBVH_STACK (tree depth <= key bits + position bits), so `l_node/l_t` cannot overflow.
// (sp and lsp are passed and returned by value so that they stay in registers in the caller.)
__device__ __noinline__ int trace_spill(TraceSmem& sm, int tid, uint32_t* l_node, float* l_t, int lsp)
{
    for (int k = 0; k < TRACE_SPILL; ++k) {
        l_node[lsp + k] = sm.node[k][tid];
        l_t[lsp + k] = sm.t[k][tid];
        sm.node[k][tid] = sm.node[k + TRACE_SPILL][tid];
        sm.t[k][tid] = sm.t[k + TRACE_SPILL][tid];
    }
    return lsp + TRACE_SPILL;  // sp -= TRACE_SPILL in the caller
}
__device__ __noinline__ int trace_refill(TraceSmem& sm, int tid, const uint32_t* l_node, const float* l_t, int lsp)
{
    lsp -= TRACE_SPILL;
    for (int k = 0; k < TRACE_SPILL; ++k) {
        sm.node[k][tid] = l_node[lsp + k];
        sm.t[k][tid] = l_t[lsp + k];
    }
    return lsp;  // sp = TRACE_SPILL in the caller
}

// Epilogue of the lanes whose ray has finished (`pending`).  A Raycast is 36 bytes at a 36-byte stride: written by its own
// lane it costs nine 4-byte stores per lane, each instruction touching one sector per ray — 7 % of the kernel's L1
// wavefronts (ncu, profiles/r1_e).  Into LOCAL memory that is nevertheless the fastest form (measured, profiles/r2_experiments.txt:
// the warp-cooperative transposition below costs 2.5 % more in instructions than the wavefronts it saves: 1800 vs 1846 Mrays/s),
// so STAGED is only chosen when `out` is a peer GPU's memory (b2col_comm_raycast_gather), where per-lane stores would become
// nine NVLink write packets of 4 bytes per ray: every finished lane packs its nine words into the warp's staging
// rows, ray after ray (stride 9 words: conflict-free), and the warp writes the staged words out 32 at a time — each
// instruction covers ~3.5 whole Raycasts, contiguous 36-byte runs in global memory.  The staging area holds STAGE_RAYS rays
// per warp (shared memory is what limits the kernel's occupancy); more finished lanes than that take another round.
template <bool STAGED, int OUTW>
__device__ __forceinline__ void flush_finished(TraceSmem& sm, int warp, int lane, uint32_t lt_mask, bool pending, const uint32_t* w9, unsigned long long my,
                                               b2col_raycast* __restrict__ out)
{
    if (!STAGED) {  // local memory: every lane writes its own result, OUTW 4-byte stores (9: Raycast, 3: b2col_hit)
        if (pending) {
#pragma unroll
            for (int k = 0; k < OUTW; ++k) reinterpret_cast<uint32_t*>(out)[(unsigned long long)OUTW * my + k] = w9[k];
        }
        return;
    }
    static_assert(!STAGED || OUTW == 9, "only full Raycasts are staged");
    unsigned pend = __ballot_sync(0xffffffffu, pending);
    if (pend == 0u) return;
    uint32_t* st = sm.stage[warp];
    uint32_t* o32 = reinterpret_cast<uint32_t*>(out);
    int j = pending ? __popc(pend & lt_mask) : -1;  // this lane's place among the finished rays
    int left = __popc(pend);
    for (;;) {  // rounds of at most STAGE_RAYS rays (usually one: 7-12 lanes finish per fetch)
        const int round = min(left, STAGE_RAYS);
        if (j >= 0 && j < round) {  // 11 words per ray: the Raycast and where it goes (stride 11: conflict-free)
            uint32_t* p = st + STAGE_WORDS * j;
#pragma unroll
            for (int k = 0; k < 9; ++k) p[k] = w9[k];
            p[9] = (uint32_t)my;
            p[10] = (uint32_t)(my >> 32);
        }
        __syncwarp();
        const int total = 9 * round;
        for (int e = lane; e < total; e += 32) {
            const int r = (e * 7282) >> 16;  // e / 9 for e < 288
            const uint32_t* p = st + STAGE_WORDS * r;
            const unsigned long long m = (unsigned long long)p[9] | ((unsigned long long)p[10] << 32);
            o32[9ull * m + (unsigned)(e - 9 * r)] = p[e - 9 * r];
        }
        __syncwarp();  // the staging rows are reused by the next round / flush
        left -= round;
        if (left <= 0) break;
        j -= round;
    }
}

// GetRaycast (reference src/system_collisions.cpp:160-199) over a batch, production tree.
//
// Persistent warps with dynamic ray fetch: incoherent rays finish after wildly different numbers of
// node visits, so a lane whose ray is done does not idle until the slowest ray of its warp ends —
// when fewer than `fetch_threshold` lanes are still busy the warp claims new rays for its idle lanes
// from a global counter (one atomicAdd per warp) and carries on.
//
// The kernel is bound by instruction issue (ncu: 79 % issue slots busy, profiles/r1_b), so the step is
// written branch-free where lanes disagree: the pop of a deferred child, the visit and the push are ONE
// straight instruction stream per step (a lane that has nothing to pop / push just predicates the two
// shared-memory accesses off); only the rare events — a leaf slot passing the grid test (about once per
// ray), stack spill / refill — branch.  A popped child that has fallen behind the best hit costs its
// lane one idle step instead of an inner loop.
//
// Per-lane registers: the grid-space ray (6 floats + 6 byte-permute selectors), node, stack pointers, best
// (t, leaf), pruning limit, ray index.  The exact ray lives in shared memory (written once at fetch) for
// the exact leaf tests and the epilogue.
// COUNT (B2COL_MODE_LBVH_COUNT, a diagnostic instantiation): also adds the number of node visits to counter[1] — what
// bench.py turns into the L1 wavefront rate, the resource that actually binds this kernel.
// WILD: the tree holds leaves the grid cannot bound (hdr->has_wild; b2col_bvh.cuh WILD_BIT): flagged slots are always
// entered and no ray is dismissed by its distance from the grid.  Chosen once per kernel, so that sane scenes run the
// instantiation that never looks at the bit.
template <bool COUNT, bool WILD, bool STAGED, bool COMPACT>
__device__ __forceinline__ void trace_body(TraceSmem& sm, const b2col_ray* __restrict__ rays, unsigned long long n_rays, b2col_raycast* __restrict__ out,
                                           const BlobHeader* __restrict__ hdr, const QNode* __restrict__ qnodes, const LeafRec* __restrict__ leaves, float T,
                                           unsigned long long* counter, int fetch_threshold)
{
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t root = hdr->trace_root;
    const Grid grid = hdr->grid;
    const float lim0 = float_pred(T);
    uint32_t l_node[LBVH_STACK];
    float l_t[LBVH_STACK];
    RayQ q;
    Hit best;
    best.t = B2_INF;
    best.leaf = -1;
    float lim = lim0;
    uint32_t cur = REF_EMPTY;
    int sp = 0, lsp = 0;
    unsigned long long my = 0ull;
    bool active = false;
    bool pending = false;    // this lane's ray is finished but its Raycast is not written yet
    bool exhausted = false;  // warp-uniform: the pool has no rays left
    unsigned long long visits = 0ull;  // COUNT only
#ifdef B2_REG_TOS
    uint32_t tos_n = REF_EMPTY;
    float tos_t = 0.0f;
    bool has_tos = false;
#endif
    auto exact_ray = [&]() {
        RayPre r;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            r.o[k] = sm.ray[k][tid];
            r.d[k] = sm.ray[3 + k][tid];
            r.ninv[k] = sm.ray[6 + k][tid];
        }
        return r;
    };
    for (;;) {
        // Epilogues are deferred to this point so that the lanes that finished since the last fetch
        // run them together instead of one at a time inside the traversal loop.
        {
            uint32_t w9[9];
            if (pending) {
                if (COMPACT) finish_hit<DevOps>(w9, exact_ray(), best, leaves);
                else finish_raycast<DevOps>(w9, exact_ray(), best, leaves);
            }
            flush_finished<STAGED, COMPACT ? 3 : 9>(sm, warp, lane, lt_mask, pending, w9, my, out);
            pending = false;
        }
        const unsigned need = __ballot_sync(0xffffffffu, !active);
        if (need != 0u && !exhausted) {
            const int leader = __ffs(need) - 1;
            const unsigned long long cnt = (unsigned long long)__popc(need);
            unsigned long long base = 0ull;
            if (lane == leader) base = atomicAdd(counter, cnt);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!active) {
                my = base + (unsigned long long)__popc(need & lt_mask);
                if (my < n_rays) {
                    const RayPre r = load_ray(rays, my);
                    if (rayq_prepare(r, grid, q) || WILD) {
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            sm.ray[k][tid] = r.o[k];
                            sm.ray[3 + k][tid] = r.d[k];
                            sm.ray[6 + k][tid] = r.ninv[k];
                        }
                        best.t = B2_INF;
                        best.leaf = -1;
                        lim = lim0;
                        cur = root;
                        active = true;
                    }
                    else if (COMPACT) {  // provable miss
                        uint32_t* o3 = reinterpret_cast<uint32_t*>(out) + 3ull * my;
                        o3[0] = o3[1] = o3[2] = 0u;
                    }
                    else {
                        write_raycast(reinterpret_cast<uint32_t*>(out + my), r, false, 0.0f, 0, 0u);  // provable miss
                    }
                }
            }
            exhausted = base + cnt >= n_rays;
        }
        const unsigned busy = __ballot_sync(0xffffffffu, active);
        if (busy == 0u) {
            if (exhausted) break;
            continue;
        }
        const int keep_going = exhausted ? 1 : fetch_threshold;
        for (;;) {
#pragma unroll
            for (int step = 0; step < B2_TRACE_STEPS; ++step) {
                // ---- pop (predicated): a lane without a current node takes its most recent deferred child
#ifdef B2_REG_TOS  // dev variant: the newest deferred child stays in registers; shared memory holds the older ones
                if (active && cur == REF_EMPTY && !has_tos && sp == 0) {  // rare; lsp > 0 here
                    lsp = trace_refill(sm, tid, l_node, l_t, lsp);
                    sp = TRACE_SPILL;
                }
                const bool pop = active && cur == REF_EMPTY;
                if (pop) {
                    uint32_t pn = tos_n;
                    float pt = tos_t;
                    if (!has_tos) {
                        --sp;
                        pn = sm.node[sp][tid];
                        pt = sm.t[sp][tid];
                    }
                    has_tos = false;
                    cur = (pt <= lim) ? pn : REF_EMPTY;
                }
#else
                if (active && cur == REF_EMPTY && sp == 0) {  // rare; lsp > 0 here
                    lsp = trace_refill(sm, tid, l_node, l_t, lsp);
                    sp = TRACE_SPILL;
                }
                const bool pop = active && cur == REF_EMPTY;
                sp -= pop ? 1 : 0;
                const uint32_t pn = sm.node[sp][tid];  // sp < TRACE_SMEM_STACK always: harmless when not popping
                const float pt = sm.t[sp][tid];
                if (pop) cur = (pt <= lim) ? pn : REF_EMPTY;
#endif
                // ---- visit
                if (cur != REF_EMPTY) {
                    if (COUNT) ++visits;
                    const QNode nd = DevOps::load_qnode(&qnodes[cur]);
                    const VisitOut v = visit_node<DevOps, WILD>(best, lim, q, nd, leaves, T, exact_ray);
#ifdef B2_REG_TOS
                    if (v.push) {
                        if (has_tos) {
                            sm.node[sp][tid] = tos_n;
                            sm.t[sp][tid] = tos_t;
                            ++sp;
                        }
                        tos_n = v.far;
                        tos_t = v.tfar;
                        has_tos = true;
                    }
#else
                    if (v.push) {
                        sm.node[sp][tid] = v.far;
                        sm.t[sp][tid] = v.tfar;
                    }
                    sp += v.push ? 1 : 0;
#endif
                    cur = v.next;
                }
                if (sp == TRACE_SMEM_STACK) {  // rare
                    lsp = trace_spill(sm, tid, l_node, l_t, lsp);
                    sp = TRACE_SMEM_STACK - TRACE_SPILL;
                }
#ifdef B2_REG_TOS
                if (active && cur == REF_EMPTY && sp == 0 && lsp == 0 && !has_tos) {
#else
                if (active && cur == REF_EMPTY && sp == 0 && lsp == 0) {
#endif
                    active = false;
                    pending = true;
                }
            }
            if (__popc(__ballot_sync(0xffffffffu, active)) < keep_going) break;
        }
    }
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) visits += __shfl_xor_sync(0xffffffffu, visits, o);
        if (lane == 0 && visits) atomicAdd(counter + 1, visits);
    }
}

// COMPACT: `out` is an array of 12-byte b2col_hit records (entity, distance, hit | side << 8) instead of 36-byte Raycasts.
template <bool COUNT, bool STAGED, bool COMPACT = false>
__global__ void __launch_bounds__(TRACE_BLOCK, B2_TRACE_MIN_BLOCKS)
k_trace_lbvh(const b2col_ray* __restrict__ rays, unsigned long long n_rays, b2col_raycast* __restrict__ out,
             const BlobHeader* __restrict__ hdr, const QNode* __restrict__ qnodes, const LeafRec* __restrict__ leaves, float T,
             unsigned long long* counter, int fetch_threshold)
{
    __shared__ TraceSmem sm;
    if (hdr->has_wild) trace_body<COUNT, true, STAGED, COMPACT>(sm, rays, n_rays, out, hdr, qnodes, leaves, T, counter, fetch_threshold);
    else trace_body<COUNT, false, STAGED, COMPACT>(sm, rays, n_rays, out, hdr, qnodes, leaves, T, counter, fetch_threshold);
}
