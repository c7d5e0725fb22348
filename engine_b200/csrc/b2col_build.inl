// k_build_direct: the LBVH hierarchy of k_build with the block-local part computed DIRECTLY instead of climbed.
// Included by b2col.cu after DevOps / BlobHeader / k_build (shares BUILD_BLOCK and the global exchange words).
//
// The nodes of an LBVH over sorted keys are the Cartesian tree of the split levels: node p (the split between sorted
// positions p and p+1) covers the leaves between the nearest HIGHER splits on either side, its children are the highest
// splits of the two halves (or single leaves).  Inside one block's BUILD_BLOCK consecutive leaves everything a node
// needs is therefore a range query:
//     nearest higher split left / right   -> binary lifting over a sparse table of range-max levels
//     child node of a half                -> arg-max of the same table (levels packed with their position)
//     child box, wild flag                -> range union over a sparse table of the leaf boxes (min/max are idempotent:
//                                            two overlapping power-of-two windows cover any range)
// so every block-local node (>95 % of all nodes) is produced by ONE thread from ~25 shared-memory loads and written
// with one coalesced 32-byte store — no arrival protocol, no divergence: the climb of k_build ran at ~7 of 32 lanes
// and stored every node in three scattered pieces.  Quantised-plane merges are associative and commutative, so the
// boxes are bit-identical to the climbed ones (tests/test_gpu_parity.py::test_build_matches_host_restatement compares
// against the host restatement of the climb).
//
// What is left for the global 128-bit exchange (b2col_bvh.cuh) is exactly what k_build sent there: the subtrees whose
// parent straddles a block boundary — maximal local nodes and boundary leaves, ~30 per block.  They are compacted into
// a shared-memory list first, so one converged warp does the block's global climbing while the other warps retire.
constexpr int BUILD_LOG = 7;
static_assert(BUILD_BLOCK == (1 << BUILD_LOG), "k_build_direct's sparse tables assume BUILD_BLOCK = 2^BUILD_LOG");

struct ClimbTask {  // 32 bytes
    uint32_t l, r, lev_l, lev_r;
    uint32_t w[3];
    uint32_t ref;  // child ref incl. LEAF_BIT / WILD_BIT
};

__device__ __forceinline__ uint4 qentry_merge(const uint4& a, const uint4& b)
{
    return make_uint4(qplane_merge(a.x, b.x), qplane_merge(a.y, b.y), qplane_merge(a.z, b.z), a.w | b.w);
}

__global__ void __launch_bounds__(BUILD_BLOCK)
k_build_direct(uint32_t n, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ sorted_prim, const LeafRec* __restrict__ world_rec, Xch* xch, QNode* qnodes,
               LeafRec* leaves, BlobHeader* hdr, uint32_t* scratch, uint32_t scratch_words)
{
    pdl_wait();
    for (uint32_t w = blockIdx.x * BUILD_BLOCK + threadIdx.x; w < scratch_words; w += gridDim.x * BUILD_BLOCK) scratch[w] = 0u;  // as k_build
    __shared__ uint32_t s_lev[BUILD_BLOCK + 1];  // s_lev[j] = level of split b0-1+j (j = 0: left boundary, j = BUILD_BLOCK: right boundary)
    __shared__ uint32_t s_wmax[2][BUILD_BLOCK / 32];
    __shared__ uint8_t s_local[BUILD_BLOCK + 1];
    __shared__ uint32_t s_stl[BUILD_LOG][BUILD_BLOCK];               // [k][j]: max of (level << 8 | split) over splits [j, j + 2^k)
    __shared__ __align__(16) uint4 s_stb[BUILD_LOG][BUILD_BLOCK];    // [k][i]: union of the grid boxes (x,y,z planes; w = wild) of leaves [i, i + 2^k)
    __shared__ uint32_t s_ntasks;
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
    bool wild = false;
    if (valid) {
        const LeafRec rec = DevOps::load_leaf(&world_rec[sorted_prim[i]]);
        DevOps::store_leaf(&leaves[i], rec);
        box = quantise_box(rec.lo, rec.hi, grid);
        wild = box_is_wild(rec.lo, rec.hi);
        key = keys[i];
        if (i > 0u) lev_l = split_level(keys[i - 1u], key, i - 1u);
    }
    if (n == 1u) {  // a lone leaf still gets a node; both slots = the leaf
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
    if (tid == BUILD_BLOCK - 1) s_lev[BUILD_BLOCK] = (valid && i + 1u < n) ? split_level(key, keys[i + 1u], i) : LEVEL_END;
    if (tid == 0) s_ntasks = 0u;
    uint32_t cur_l = (lev_l << 8) | (uint32_t)tid;
    uint4 cur_b = make_uint4(box.w[0], box.w[1], box.w[2], wild ? 1u : 0u);
    s_stl[0][tid] = cur_l;
    s_stb[0][tid] = cur_b;
    // ---- which splits are block-local nodes: prefix / suffix maxima of the levels (as k_build)
    uint32_t pm = lev_l, sx = lev_l;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, pm, o), c = __shfl_down_sync(0xffffffffu, sx, o);
        if (lane >= o) pm = max(pm, a);
        if (lane + o < 32) sx = max(sx, c);
    }
    if (lane == 31) s_wmax[0][warp] = pm;
    if (lane == 0) s_wmax[1][warp] = sx;
    __syncthreads();
    bool local;
    {
        uint32_t before = 0u, after = s_lev[BUILD_BLOCK];
#pragma unroll
        for (int w = 0; w < BUILD_BLOCK / 32; ++w) {
            if (w < warp) before = max(before, s_wmax[0][w]);
            if (w > warp) after = max(after, s_wmax[1][w]);
        }
        const uint32_t pm_in = __shfl_up_sync(0xffffffffu, pm, 1), sx_in = __shfl_down_sync(0xffffffffu, sx, 1);
        const uint32_t left_max = max(before, lane > 0 ? pm_in : 0u);
        const uint32_t right_max = max(after, lane < 31 ? sx_in : 0u);
        local = tid > 0 && left_max > lev_l && right_max > lev_l;  // split 0 belongs to the neighbours too: never local
        s_local[tid] = local ? 1 : 0;
        if (tid == 0) s_local[BUILD_BLOCK] = 0;
    }
    // ---- sparse tables: level k from level k-1 (entries whose window would leave the block are never read)
#pragma unroll
    for (int k = 1; k < BUILD_LOG; ++k) {
        const int half = 1 << (k - 1);
        if (tid + 2 * half <= BUILD_BLOCK) {
            cur_l = max(cur_l, s_stl[k - 1][tid + half]);
            cur_b = qentry_merge(cur_b, s_stb[k - 1][tid + half]);
            s_stl[k][tid] = cur_l;
            s_stb[k][tid] = cur_b;
        }
        __syncthreads();
    }

    // ---- node role: thread tid finishes node p = b0 - 1 + tid if it is block-local
    ClimbTask task[2];
    bool have[2] = {false, false};
    if (local) {
        const uint32_t thr = lev_l << 8;  // packed entries below this have a lower level
        int L = tid, R = tid;             // grow over the lower splits on either side
#pragma unroll
        for (int k = BUILD_LOG - 1; k >= 0; --k) {
            const int step = 1 << k;
            if (L >= step && s_stl[k][L - step] < thr) L -= step;
            if (R + step <= BUILD_BLOCK - 1 && s_stl[k][R + 1] < thr) R += step;
        }
        L -= 1;  // nearest higher split on the left (exists: the node is local); leaves [L, tid-1] hang on the left
        R += 1;  // ... on the right (BUILD_BLOCK = the right boundary split); leaves [tid, R-1] hang on the right
        const uint32_t p = b0 - 1u + (uint32_t)tid;
        QNode nd;
        uint32_t any_wild = 0u;
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const int a = side ? tid : L, b = side ? R - 1 : tid - 1;  // leaf range of this child, block-relative
            const int k = 31 - __clz(b - a + 1);
            const uint4 u = qentry_merge(s_stb[k][a], s_stb[k][b + 1 - (1 << k)]);
            nd.w[3 * side + 0] = u.x;
            nd.w[3 * side + 1] = u.y;
            nd.w[3 * side + 2] = u.z;
            uint32_t ref;
            if (a == b) ref = LEAF_BIT | (b0 + (uint32_t)a);
            else {  // highest split strictly inside: splits [a+1, b]
                const int ks = 31 - __clz(b - a);
                const uint32_t m = max(s_stl[ks][a + 1], s_stl[ks][b + 1 - (1 << ks)]);
                ref = b0 - 1u + (m & 0xFFu);
            }
            nd.w[6 + side] = ref | (u.w ? WILD_BIT : 0u);
            any_wild |= u.w;
        }
        DevOps::store_qnode(&qnodes[p], nd);
        const uint32_t up_l = s_lev[L], up_r = s_lev[R];
        if (up_l == LEVEL_END && up_r == LEVEL_END) {  // the range is the whole array (a tree of one block)
            hdr->root = p;
            hdr->trace_root = p;
            hdr->top_count = 0u;
            hdr->has_wild = any_wild ? 1u : 0u;
        }
        else if (!s_local[up_r < up_l ? R : L]) {  // the parent straddles a block boundary: this subtree meets its sibling through global memory
            ClimbTask& t = task[1];
            t.l = b0 + (uint32_t)L;
            t.r = b0 + (uint32_t)R - 1u;
            t.lev_l = up_l;
            t.lev_r = up_r;
#pragma unroll
            for (int k = 0; k < 3; ++k) t.w[k] = qplane_merge(nd.w[k], nd.w[3 + k]);
            t.ref = p | (any_wild ? WILD_BIT : 0u);
            have[1] = true;
        }
    }
    // ---- leaf role: a leaf whose parent is not block-local climbs through global memory as well
    if (valid) {
        const uint32_t lev_r = s_lev[tid + 1];
        if (!s_local[lev_r < lev_l ? tid + 1 : tid]) {
            ClimbTask& t = task[0];
            t.l = t.r = i;
            t.lev_l = lev_l;
            t.lev_r = lev_r;
#pragma unroll
            for (int k = 0; k < 3; ++k) t.w[k] = box.w[k];
            t.ref = LEAF_BIT | i | (wild ? WILD_BIT : 0u);
            have[0] = true;
        }
    }
    __syncthreads();  // every read of the box table is done: its first rows become the task list
    ClimbTask* s_tasks = reinterpret_cast<ClimbTask*>(&s_stb[0][0]);  // <= 2 * BUILD_BLOCK tasks of 32 bytes = rows 0..3
    static_assert(sizeof(ClimbTask) * 2 * BUILD_BLOCK <= sizeof(uint4) * BUILD_LOG * BUILD_BLOCK, "task list must fit the box table");
#pragma unroll
    for (int role = 0; role < 2; ++role) {
        if (have[role]) {
            const uint32_t at = atomicAdd(&s_ntasks, 1u);
            uint4* dst = reinterpret_cast<uint4*>(&s_tasks[at]);
            dst[0] = make_uint4(task[role].l, task[role].r, task[role].lev_l, task[role].lev_r);
            dst[1] = make_uint4(task[role].w[0], task[role].w[1], task[role].w[2], task[role].ref);
        }
    }
    __syncthreads();
    const uint32_t n_tasks = s_ntasks;
    for (uint32_t t = (uint32_t)tid; t < n_tasks; t += BUILD_BLOCK) {
        const uint4 t0 = reinterpret_cast<const uint4*>(&s_tasks[t])[0], t1 = reinterpret_cast<const uint4*>(&s_tasks[t])[1];
        uint32_t l = t0.x, r = t0.y, lv_l = t0.z, lv_r = t0.w;
        QBox bx;
        bx.w[0] = t1.x;
        bx.w[1] = t1.y;
        bx.w[2] = t1.z;
        uint32_t ref = t1.w & ~WILD_BIT;
        bool wd = (t1.w & WILD_BIT) != 0u;
        for (;;) {  // the climb of build_from_leaf_q, every step through the global exchange
            const bool is_left = lv_r < lv_l;
            const uint32_t p = is_left ? r : l - 1u;
            DevOps::store_ref(&qnodes[p], is_left ? 0 : 1, ref | (wd ? WILD_BIT : 0u));
            const Xch got = DevOps::exchange(&xch[p], xch_pack(bx, is_left ? l : r, is_left ? lv_l : lv_r, wd));
            if (got.w[3] == UNVISITED) break;  // first to arrive: the sibling's thread finishes this node
            *reinterpret_cast<uint4*>(&xch[p]) = make_uint4(0u, 0u, 0u, UNVISITED);  // self-cleaning: ready for the next frame
            QBox sib;
            uint32_t bound, level;
            bool sib_wild;
            xch_unpack(got, sib, bound, level, sib_wild);
            wd = wd || sib_wild;
            DevOps::store_qboxes(&qnodes[p], is_left ? bx.w : sib.w, is_left ? sib.w : bx.w);
            if (is_left) {
                r = bound;
                lv_r = level;
            }
            else {
                l = bound;
                lv_l = level;
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) bx.w[k] = qplane_merge(bx.w[k], sib.w[k]);
            ref = p;
            if (lv_l == LEVEL_END && lv_r == LEVEL_END) {
                hdr->root = p;
                hdr->trace_root = p;
                hdr->top_count = 0u;
                hdr->has_wild = wd ? 1u : 0u;
                break;
            }
        }
    }
}

// k_build_warp: k_build with the lowest levels of the climb done by WARP ROUNDS instead of arrivals.
//
// The roots of the partial subtrees over a warp's 32 consecutive leaves always partition those leaves into contiguous
// ranges; the root of the range starting at leaf x stays with lane x - warp_base.  Range A = [l..r] joins its right
// neighbour iff the split beyond r is the lower of its two outer splits; the neighbour B = [r+1..] joins to the left iff
// the split before it is its lower one.  When both hold they are siblings (their parent is node r), and A learns that
// with shuffles from lane r+1 - warp_base: no arrival word, no atomic, no divergence — every round all mutual pairs of
// the warp merge at once, A stores the finished 32-byte node in ONE piece and carries the union on, B retires.  A few
// rounds (the height of the forest inside 32 leaves) finish ~29 of every 31 nodes; the roots that are left — subtrees
// whose parent spans a warp boundary — continue with k_build's arrival protocol (shared-memory exchange inside the
// block, global exchange across blocks).  Same nodes, same bytes as k_build.
__global__ void __launch_bounds__(BUILD_BLOCK)
k_build_warp(uint32_t n, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ sorted_prim, const LeafRec* __restrict__ world_rec, Xch* xch, QNode* qnodes,
             LeafRec* leaves, BlobHeader* hdr, uint32_t* scratch, uint32_t scratch_words)
{
    pdl_wait();
    for (uint32_t w = blockIdx.x * BUILD_BLOCK + threadIdx.x; w < scratch_words; w += gridDim.x * BUILD_BLOCK) scratch[w] = 0u;  // as k_build
    __shared__ uint32_t s_lev[BUILD_BLOCK + 1];
    __shared__ uint32_t s_wmax[2][BUILD_BLOCK / 32];
    __shared__ uint8_t s_local[BUILD_BLOCK + 1];
    __shared__ Xch s_xch[BUILD_BLOCK + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t b0 = blockIdx.x * BUILD_BLOCK;
    const uint32_t i = b0 + (uint32_t)tid;
    const bool valid = i < n;
    const Grid grid = hdr->grid;

    QBox box;
    box.w[0] = box.w[1] = box.w[2] = 0u;
    uint64_t key = 0ull;
    uint32_t lev_l = LEVEL_END;
    bool wild = false;
    if (valid) {
        const LeafRec rec = DevOps::load_leaf(&world_rec[sorted_prim[i]]);
        DevOps::store_leaf(&leaves[i], rec);
        box = quantise_box(rec.lo, rec.hi, grid);
        wild = box_is_wild(rec.lo, rec.hi);
        key = keys[i];
        if (i > 0u) lev_l = split_level(keys[i - 1u], key, i - 1u);
    }
    if (n == 1u) {
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
    uint32_t pm = lev_l, sx = lev_l;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, pm, o), c = __shfl_down_sync(0xffffffffu, sx, o);
        if (lane >= o) pm = max(pm, a);
        if (lane + o < 32) sx = max(sx, c);
    }
    if (lane == 31) s_wmax[0][warp] = pm;
    if (lane == 0) s_wmax[1][warp] = sx;
    __syncthreads();
    {
        uint32_t before = 0u, after = s_lev[BUILD_BLOCK];
#pragma unroll
        for (int w = 0; w < BUILD_BLOCK / 32; ++w) {
            if (w < warp) before = max(before, s_wmax[0][w]);
            if (w > warp) after = max(after, s_wmax[1][w]);
        }
        const uint32_t pm_in = __shfl_up_sync(0xffffffffu, pm, 1), sx_in = __shfl_down_sync(0xffffffffu, sx, 1);
        const uint32_t left_max = max(before, lane > 0 ? pm_in : 0u);
        const uint32_t right_max = max(after, lane < 31 ? sx_in : 0u);
        s_local[tid] = (tid > 0 && left_max > lev_l && right_max > lev_l) ? 1 : 0;
        if (tid == 0) s_local[BUILD_BLOCK] = 0;
    }
    __syncthreads();

    // ---- phase 1: warp rounds
    uint32_t l = i, r = i;
    uint32_t lev_r = s_lev[tid + 1];
    uint32_t refw = LEAF_BIT | i | (wild ? WILD_BIT : 0u);  // this subtree's child ref incl. the wild flag
    bool active = valid;
    const uint32_t wbase = i - (uint32_t)lane;
    for (;;) {
        const bool want_right = active && lev_r < lev_l;  // joins the neighbour beyond the LOWER split; array ends never win
        const uint32_t pl = r + 1u - wbase;               // lane holding the root of the range that starts behind mine
        const int src = (int)(pl & 31u);
        const uint32_t o_r = __shfl_sync(0xffffffffu, r, src);
        const uint32_t o_lv = __shfl_sync(0xffffffffu, lev_r | (want_right ? 0x100u : 0u), src);
        const uint32_t o_b0 = __shfl_sync(0xffffffffu, box.w[0], src);
        const uint32_t o_b1 = __shfl_sync(0xffffffffu, box.w[1], src);
        const uint32_t o_b2 = __shfl_sync(0xffffffffu, box.w[2], src);
        const uint32_t o_ref = __shfl_sync(0xffffffffu, refw, src);
        const bool merge = want_right && pl < 32u && !(o_lv & 0x100u);  // the neighbour wants to join to the left: siblings, parent = node r
        const uint32_t taken = __reduce_or_sync(0xffffffffu, merge ? 1u << src : 0u);
        if (taken == 0u) break;
        if (merge) {
            const uint32_t p = r;
            QNode nd;
            nd.w[0] = box.w[0];
            nd.w[1] = box.w[1];
            nd.w[2] = box.w[2];
            nd.w[3] = o_b0;
            nd.w[4] = o_b1;
            nd.w[5] = o_b2;
            nd.w[6] = refw;
            nd.w[7] = o_ref;
            DevOps::store_qnode(&qnodes[p], nd);
            box.w[0] = qplane_merge(box.w[0], o_b0);
            box.w[1] = qplane_merge(box.w[1], o_b1);
            box.w[2] = qplane_merge(box.w[2], o_b2);
            refw = p | ((refw | o_ref) & WILD_BIT);
            r = o_r;
            lev_r = o_lv & 0xFFu;
            if (lev_l == LEVEL_END && lev_r == LEVEL_END) {  // the range is the whole array (a tree inside one warp)
                hdr->root = p;
                hdr->trace_root = p;
                hdr->top_count = 0u;
                hdr->has_wild = (refw & WILD_BIT) ? 1u : 0u;
                active = false;
            }
        }
        if ((taken >> lane) & 1u) active = false;  // merged into the left neighbour's subtree
    }
    if (!active) return;

    // ---- phase 2: the remaining roots climb with k_build's arrival protocol
    wild = (refw & WILD_BIT) != 0u;
    uint32_t ref = refw & ~WILD_BIT;
    for (;;) {
        const bool is_left = lev_r < lev_l;
        const uint32_t p = is_left ? r : l - 1u;
        const uint32_t j = p - (b0 - 1u);
        const bool local = j <= (uint32_t)BUILD_BLOCK && s_local[j] != 0;
        QBox sib;
        uint32_t bound, level;
        bool sib_wild;
        DevOps::store_ref(&qnodes[p], is_left ? 0 : 1, ref | (wild ? WILD_BIT : 0u));
        const Xch mine = xch_pack(box, is_left ? l : r, is_left ? lev_l : lev_r, wild);
        Xch got;
        if (local) {
            got = DevOps::exchange_shared(&s_xch[j], mine);
            if (got.w[3] == UNVISITED) return;
        }
        else {
            got = DevOps::exchange(&xch[p], mine);
            if (got.w[3] == UNVISITED) return;
            *reinterpret_cast<uint4*>(&xch[p]) = make_uint4(0u, 0u, 0u, UNVISITED);
        }
        xch_unpack(got, sib, bound, level, sib_wild);
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
        if (lev_l == LEVEL_END && lev_r == LEVEL_END) {
            hdr->root = p;
            hdr->trace_root = p;
            hdr->top_count = 0u;
            hdr->has_wild = wild ? 1u : 0u;
            return;
        }
    }
}
