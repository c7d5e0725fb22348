// b2col_comm.inl — multi-GPU data plane of libb2col.so (included at the end of b2col.cu).
//
// ray batches shard across the GPUs of one box (contiguous, order-preserving slices), the BVH is
// REPLICATED (rank `src` builds, one ncclBroadcast straight out of / into the blob allocations over NVLink), and the
// hits are gathered back in batch order on one rank.  Two ways to run it, same entry points underneath:
//   * one process per GPU (torchrun, bench.py): b2col_comm_get_unique_id on rank 0, the caller distributes the 128
//     bytes by any means, every rank calls b2col_comm_init_rank;
//   * one process driving several GPUs (the C++ engine: engine_b200/shim): b2col_comm_init_all over n contexts, then the
//     *_all entry points, which wrap the per-context calls in one NCCL group.
//
// The gather is FUSED into the trace kernel: the destination rank owns a window (one cudaMalloc of n_total Raycasts),
// every other rank maps it — cudaIpcOpenMemHandle between processes, peer access inside one — and k_trace_lbvh writes
// each finished Raycast straight into its slot of the window over NVLink (flush_finished: whole 36-byte records, ~3.5
// per store instruction), so the transfer rides under the traversal ray by ray and nothing is sent afterwards.  What
// remains of the collective is its completion: a 4-byte all-reduce behind the kernel on every rank's stream (between
// processes) or an event wait (inside one).  If the window cannot be mapped (IPC unavailable) every rank falls back,
// collectively, to tracing locally and ncclSend / ncclRecv of its slice.
//
// NCCL is loaded with dlopen at the first b2col_comm_* call (libnccl.so.2: inside a PyTorch process that resolves to the
// copy torch already loaded), so that libb2col.so itself carries no NCCL dependency: single-GPU users never need it.

#include <dlfcn.h>

namespace {

// the subset of nccl.h this file uses (ABI-stable since NCCL 2.0); declared here so that the build needs no NCCL headers
typedef struct ncclComm* b2_ncclComm_t;
typedef struct {
    char internal[128];
} b2_ncclUniqueId;
static_assert(sizeof(b2_ncclUniqueId) == B2COL_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
enum { B2_NCCL_UINT8 = 1, B2_NCCL_UINT32 = 3, B2_NCCL_MAX = 2 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(b2_ncclUniqueId*) = nullptr;
    int (*CommInitRank)(b2_ncclComm_t*, int, b2_ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(b2_ncclComm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(b2_ncclComm_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, b2_ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, b2_ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, b2_ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, b2_ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string error;
};

NcclApi& nccl_api()
{
    static NcclApi api;
    return api;
}

bool nccl_load()
{
    NcclApi& a = nccl_api();
    if (a.handle) return true;
    const char* names[] = {getenv("B2COL_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* nm : names) {
        if (!nm || !*nm) continue;
        h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        a.error = std::string("cannot load NCCL (libnccl.so.2): ") + (dlerror() ? dlerror() : "not found");
        return false;
    }
#define B2_SYM(field, name)                                                   \
    do {                                                                      \
        *reinterpret_cast<void**>(&a.field) = dlsym(h, name);                 \
        if (!a.field) {                                                       \
            a.error = std::string("NCCL symbol missing: ") + name;            \
            return false;                                                     \
        }                                                                     \
    } while (0)
    B2_SYM(GetUniqueId, "ncclGetUniqueId");
    B2_SYM(CommInitRank, "ncclCommInitRank");
    B2_SYM(CommInitAll, "ncclCommInitAll");
    B2_SYM(CommDestroy, "ncclCommDestroy");
    B2_SYM(Broadcast, "ncclBroadcast");
    B2_SYM(AllReduce, "ncclAllReduce");
    B2_SYM(Send, "ncclSend");
    B2_SYM(Recv, "ncclRecv");
    B2_SYM(GroupStart, "ncclGroupStart");
    B2_SYM(GroupEnd, "ncclGroupEnd");
    B2_SYM(GetErrorString, "ncclGetErrorString");
    B2_SYM(GetVersion, "ncclGetVersion");
#undef B2_SYM
    a.handle = h;
    return true;
}

#define NC(call)                                                                                                   \
    do {                                                                                                           \
        const int r_ = (call);                                                                                     \
        if (r_ != 0) return fail(ctx, B2COL_ERR_COMM, "%s failed: %s", #call, nccl_api().GetErrorString(r_));     \
    } while (0)

}  // namespace

// Communicator state of one context.  `peers` is only set in the single-process flavour (b2col_comm_init_all): the
// contexts of all ranks, in rank order, so that completion can use CUDA events and the window a plain peer pointer.
struct b2col_comm_state {
    b2_ncclComm_t comm = nullptr;
    int n_ranks = 0, rank = -1;
    std::vector<b2col_ctx*> peers;
    // gather window: n_total Raycasts in batch order, owned by rank win_dst
    void* win = nullptr;         // this rank's address of the window (the allocation itself on win_dst)
    size_t win_bytes = 0;
    int win_dst = -1;
    bool win_ipc = false;        // `win` came from cudaIpcOpenMemHandle
    bool win_fallback = false;   // the window could not be mapped somewhere: every rank sends its slice with NCCL instead
    void* d_word = nullptr;      // 64 bytes of device scratch: completion word, IPC handle exchange
    cudaEvent_t done = nullptr;  // single-process completion
    bool stage_remote = true;    // fused mode: warp-staged 36-byte runs (default from 3 ranks on) or plain per-lane stores (B2COL_REMOTE_STAGE=0|1)
    // How a rank's hits reach a window in ANOTHER GPU's memory (B2COL_GATHER=fused|bulk; default by rank count):
    //   fused  the trace kernel stores them straight into the window (warp-staged 36-byte runs).  Nothing follows the traversal,
    //          but every Raycast is its own small NVLink write: measured fine up to 4 ranks (97.7 % of the no-gather rate) and
    //          at 46 % with 8, where seven peers' small writes saturate the destination's inbound path at ~215 GB/s.
    //   bulk   the slice is traced in chunks into LOCAL memory (alternating between two streams, so that one chunk's tail overlaps
    //          the next one's head) and every finished chunk is shipped into the window by the copy engine — large NVLink
    //          transfers that overlap the following chunks' traversal; only the last chunk's copy is exposed.
    int gather_bulk = -1;        // -1: bulk when n_ranks > 4
    size_t gather_chunk = (size_t)2 << 20;  // rays per chunk of the bulk gather (B2COL_GATHER_CHUNK)
    bool gather_tail = true;                // halving chunk sizes at the end of a slice (B2COL_GATHER_TAIL=0: equal chunks)
    uint64_t gathers = 0, replications = 0;
};

namespace {

b2col_status comm_require(b2col_ctx* ctx)
{
    if (!ctx) return B2COL_ERR_INVALID;
    if (!ctx->comm || !ctx->comm->comm) return fail(ctx, B2COL_ERR_STATE, "no communicator: call b2col_comm_init_rank / b2col_comm_init_all first");
    return B2COL_OK;
}

void comm_release_window(b2col_ctx* ctx)
{
    b2col_comm_state* c = ctx->comm;
    if (!c || !c->win) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (c->win_ipc) cudaIpcCloseMemHandle(c->win);
    else if (c->rank == c->win_dst) cudaFree(c->win);
    c->win = nullptr;
    c->win_bytes = 0;
    c->win_dst = -1;
    c->win_ipc = false;
}

b2col_status comm_attach(b2col_ctx* ctx, b2_ncclComm_t comm, int n_ranks, int rank)
{
    b2col_comm_state* c = new (std::nothrow) b2col_comm_state();
    if (!c) return fail(ctx, B2COL_ERR_INVALID, "out of host memory");
    c->comm = comm;
    c->n_ranks = n_ranks;
    c->rank = rank;
    ctx->comm = c;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMalloc(&c->d_word, 64));
    CU(cudaMemsetAsync(c->d_word, 0, 64, ctx->stream));
    CU(cudaEventCreateWithFlags(&c->done, cudaEventDisableTiming));
    if (const char* e = getenv("B2COL_GATHER")) {
        c->win_fallback = !strcmp(e, "nccl");  // A/B: force the send / recv gather
        if (!strcmp(e, "bulk")) c->gather_bulk = 1;
        if (!strcmp(e, "fused")) c->gather_bulk = 0;
    }
    if (const char* e = getenv("B2COL_GATHER_CHUNK")) c->gather_chunk = std::max<size_t>(65536, (size_t)atoll(e));
    c->gather_tail = n_ranks > 4;  // measured: 8 ranks 12 924 -> 13 374 Mrays/s with the gather inside the timed region; 2 ranks (bulk forced) 3 407 -> 3 364
    if (const char* e = getenv("B2COL_GATHER_TAIL")) c->gather_tail = atoi(e) != 0;
    c->stage_remote = n_ranks > 2;  // one peer's per-lane stores beat the staged form (3665 vs 3605 Mrays/s on 2 GPUs); from three peers on staging wins
    if (const char* e = getenv("B2COL_REMOTE_STAGE")) c->stage_remote = atoi(e) != 0;
    return B2COL_OK;
}

// The window of n_total Raycasts on rank dst, mapped on every rank.  Collective (every rank passes the same arguments);
// a no-op while the existing window is large enough.
b2col_status comm_ensure_window(b2col_ctx* ctx, size_t bytes, int dst)
{
    b2col_comm_state* c = ctx->comm;
    if (c->win && c->win_dst == dst && c->win_bytes >= bytes) return B2COL_OK;
    if (c->win_fallback && c->rank != dst) return B2COL_OK;  // senders need no window
    const size_t want = bytes + bytes / 8 + 4096;
    if (!c->peers.empty()) {  // one process: the first context to get here allocates on dst's device, the others read the pointer
        b2col_ctx* owner = c->peers[(size_t)dst];
        if (!owner || !owner->comm) return fail(ctx, B2COL_ERR_STATE, "the destination rank's context has been destroyed");
        b2col_comm_state* oc = owner->comm;
        if (!(oc->win && oc->win_dst == dst && oc->win_bytes >= bytes)) {
            comm_release_window(owner);
            for (b2col_ctx* p : c->peers)
                if (p && p != owner && p->comm && p->comm->win) {  // stale views of the old window
                    cudaSetDevice(p->device);
                    cudaStreamSynchronize(p->stream);
                    p->comm->win = nullptr;
                    p->comm->win_bytes = 0;
                }
            CU(cudaSetDevice(owner->device));
            CU(cudaMalloc(&oc->win, want));
            oc->win_bytes = want;
            oc->win_dst = dst;
        }
        c->win = oc->win;
        c->win_bytes = oc->win_bytes;
        c->win_dst = dst;
        CU(cudaSetDevice(ctx->device));
        return B2COL_OK;
    }
    // one process per GPU: dst allocates, its IPC handle travels through a 64-byte broadcast, the others open it
    comm_release_window(ctx);
    CU(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t handle;
    memset(&handle, 0, sizeof handle);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    uint32_t failed = 0;
    if (c->rank == dst) {
        CU(cudaMalloc(&c->win, want));
        if (!c->win_fallback && cudaIpcGetMemHandle(&handle, c->win) != cudaSuccess) {
            cudaGetLastError();
            failed = 1;
        }
        CU(cudaMemcpyAsync(c->d_word, &handle, 64, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (c->win_fallback) {
        c->win_bytes = want;
        c->win_dst = dst;
        return B2COL_OK;
    }
    NC(nccl_api().Broadcast(c->d_word, c->d_word, 64, B2_NCCL_UINT8, dst, c->comm, ctx->stream));
    CU(cudaMemcpyAsync(&handle, c->d_word, 64, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (c->rank != dst && !failed) {
        if (cudaIpcOpenMemHandle(&c->win, handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            c->win = nullptr;
            failed = 1;
        }
        else c->win_ipc = true;
    }
    // does the window work everywhere?  (max over ranks of `failed`)
    CU(cudaMemcpyAsync(c->d_word, &failed, 4, cudaMemcpyHostToDevice, ctx->stream));
    NC(nccl_api().AllReduce(c->d_word, c->d_word, 1, B2_NCCL_UINT32, B2_NCCL_MAX, c->comm, ctx->stream));
    CU(cudaMemcpyAsync(&failed, c->d_word, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (failed) {
        if (c->win_ipc) {
            cudaIpcCloseMemHandle(c->win);
            c->win = nullptr;
            c->win_ipc = false;
        }
        c->win_fallback = true;
        if (c->rank != dst) return B2COL_OK;
    }
    c->win_bytes = want;
    c->win_dst = dst;
    return B2COL_OK;
}

// trace this rank's slice into its place in the gathered batch; the completion step follows in the callers
b2col_status comm_enqueue_shard(b2col_ctx* ctx, const b2col_ray* d_rays_shard, size_t n_total, int dst, size_t* lo_out, size_t* n_out)
{
    b2col_comm_state* c = ctx->comm;
    size_t lo = 0, hi = 0;
    b2col_shard_bounds(n_total, c->n_ranks, c->rank, &lo, &hi);
    *lo_out = lo;
    *n_out = hi - lo;
    GUARD(comm_ensure_window(ctx, n_total * sizeof(b2col_raycast), dst));
    CU(cudaSetDevice(ctx->device));
    if (hi == lo) return B2COL_OK;
    if (!d_rays_shard) return fail(ctx, B2COL_ERR_INVALID, "null ray slice");
    if (c->win_fallback && c->rank != dst) {
        GUARD(ensure(ctx, ctx->d_hits, (hi - lo) * sizeof(b2col_raycast)));
        return enqueue_trace(ctx, d_rays_shard, hi - lo, ctx->d_hits.as<b2col_raycast>(), B2COL_MODE_LBVH);
    }
    const bool bulk = c->gather_bulk >= 0 ? c->gather_bulk != 0 : c->n_ranks > 4;
    if (c->rank != dst && bulk && ctx->have_tree && ctx->tree_leaves > 0) {
        const size_t n = hi - lo;
        // at most gather_chunk rays per chunk, at least four chunks per slice (the last chunk's copy is what stays exposed), never tiny ones
        const size_t chunk = std::min(c->gather_chunk, std::max<size_t>((size_t)256 << 10, (n / 4 + 65535) & ~(size_t)65535));
        // chunk schedule: full chunks, then a halving tail (chunk/2, chunk/4, ... down to 256K rays): what stays exposed at the end of
        // the step is the LAST chunk's copy into the window — seven peers' worth of it on the destination's inbound links — so the last
        // chunks are small; the two alternating trace streams hide the extra kernel boundaries (B2COL_GATHER_TAIL=0: equal chunks)
        std::vector<size_t> sizes;
        {
            size_t left = n;
            const size_t floor_chunk = (size_t)256 << 10;
            if (c->gather_tail && chunk >= 4 * floor_chunk && n >= 3 * chunk) {
                size_t tail_total = 0;
                std::vector<size_t> tail;
                for (size_t t = chunk / 2; t >= floor_chunk; t /= 2) {
                    tail.push_back(t);
                    tail_total += t;
                }
                tail.push_back(floor_chunk);
                tail_total += floor_chunk;
                while (left > tail_total + chunk / 2) {
                    const size_t m = std::min(chunk, left - tail_total);
                    sizes.push_back(m);
                    left -= m;
                }
                for (size_t t : tail) {
                    const size_t m = std::min(t, left);
                    if (m) sizes.push_back(m);
                    left -= m;
                }
            }
            while (left) {
                const size_t m = std::min(chunk, left);
                sizes.push_back(m);
                left -= m;
            }
        }
        const size_t n_chunks = sizes.size();
        GUARD(ensure(ctx, ctx->d_hits, n * sizeof(b2col_raycast)));
        if (!ctx->stream_in) {
            CU(cudaStreamCreateWithFlags(&ctx->stream_in, cudaStreamNonBlocking));
            CU(cudaStreamCreateWithFlags(&ctx->stream_out, cudaStreamNonBlocking));
        }
        if (!ctx->stream_trace2) {
            CU(cudaStreamCreateWithFlags(&ctx->stream_trace2, cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&ctx->pipe_start_ev, cudaEventDisableTiming));
        }
        while (ctx->pipe_events.size() < n_chunks + 1) {
            cudaEvent_t ev;
            CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            ctx->pipe_events.push_back(ev);
        }
        GUARD(ensure(ctx, ctx->d_pipe_counters, n_chunks * 16));
        CU(cudaMemsetAsync(ctx->d_pipe_counters.p, 0, n_chunks * 16, ctx->stream));
        CU(cudaEventRecord(ctx->pipe_start_ev, ctx->stream));  // behind the tree (replicate / update), the rays' producer and the memset
        CU(cudaStreamWaitEvent(ctx->stream_trace2, ctx->pipe_start_ev, 0));
        CU(cudaStreamWaitEvent(ctx->stream_out, ctx->pipe_start_ev, 0));  // also: behind the previous gather's completion step
        b2col_raycast* local = ctx->d_hits.as<b2col_raycast>();
        b2col_raycast* remote = static_cast<b2col_raycast*>(c->win) + lo;
        size_t off = 0;
        for (size_t k = 0; k < n_chunks; off += sizes[k], ++k) {
            const size_t m = sizes[k];
            cudaStream_t ts = (k & 1) ? ctx->stream_trace2 : ctx->stream;
            GUARD(enqueue_trace(ctx, d_rays_shard + off, m, local + off, B2COL_MODE_LBVH, false, false, ts, ctx->d_pipe_counters.as<unsigned long long>() + 2 * k));
            CU(cudaEventRecord(ctx->pipe_events[k], ts));
            CU(cudaStreamWaitEvent(ctx->stream_out, ctx->pipe_events[k], 0));
            CU(cudaMemcpyAsync(remote + off, local + off, m * sizeof(b2col_raycast), cudaMemcpyDefault, ctx->stream_out));  // copy engine, over NVLink
        }
        CU(cudaEventRecord(ctx->pipe_events[n_chunks], ctx->stream_out));
        CU(cudaStreamWaitEvent(ctx->stream, ctx->pipe_events[n_chunks], 0));  // the completion step (callers) follows the last copy
        return B2COL_OK;
    }
    // the window is another GPU's memory on every rank but dst: there the kernel writes whole staged Raycasts (see flush_finished)
    return enqueue_trace(ctx, d_rays_shard, hi - lo, static_cast<b2col_raycast*>(c->win) + lo, B2COL_MODE_LBVH, c->rank != dst && c->stage_remote);
}

// fallback gather: slices travel as NCCL point-to-point messages after the trace (caller wraps this in a group when one
// thread drives several ranks)
b2col_status comm_enqueue_sendrecv(b2col_ctx* ctx, size_t n_total, int dst)
{
    b2col_comm_state* c = ctx->comm;
    if (c->rank != dst) {
        size_t lo = 0, hi = 0;
        b2col_shard_bounds(n_total, c->n_ranks, c->rank, &lo, &hi);
        if (hi > lo) NC(nccl_api().Send(ctx->d_hits.p, (hi - lo) * sizeof(b2col_raycast), B2_NCCL_UINT8, dst, c->comm, ctx->stream));
        return B2COL_OK;
    }
    for (int r = 0; r < c->n_ranks; ++r) {
        if (r == dst) continue;
        size_t lo = 0, hi = 0;
        b2col_shard_bounds(n_total, c->n_ranks, r, &lo, &hi);
        if (hi > lo) NC(nccl_api().Recv(static_cast<b2col_raycast*>(c->win) + lo, (hi - lo) * sizeof(b2col_raycast), B2_NCCL_UINT8, r, c->comm, ctx->stream));
    }
    return B2COL_OK;
}

b2col_status comm_enqueue_broadcast_blob(b2col_ctx* ctx, int src, uint32_t n_colliders)
{
    b2col_comm_state* c = ctx->comm;
    const size_t bytes = blob_size_for(n_colliders);
    CU(cudaSetDevice(ctx->device));
    if (c->rank == src) {
        if (!ctx->have_tree || ctx->tree_leaves != n_colliders)
            return fail(ctx, B2COL_ERR_STATE, "source rank holds a tree of %u leaves, the collective was called for %u", ctx->have_tree ? ctx->tree_leaves : 0u, n_colliders);
    }
    else {
        const void* before = ctx->d_blob.p;
        const b2col_status st = ensure(ctx, ctx->d_blob, bytes);
        if (ctx->d_blob.p != before || st != B2COL_OK) ctx->have_tree = false;
        GUARD(st);
        ctx->have_tree = false;  // being overwritten
        ctx->topology_valid = false;
        ctx->exact_nodes_valid = false;
    }
    NC(nccl_api().Broadcast(ctx->d_blob.p, ctx->d_blob.p, bytes, B2_NCCL_UINT8, src, c->comm, ctx->stream));
    if (c->rank != src) {
        ctx->blob_bytes = bytes;
        ctx->tree_leaves = n_colliders;
        ctx->have_tree = true;
    }
    ++c->replications;
    return B2COL_OK;
}

}  // namespace

extern "C" {

void b2col_shard_bounds(size_t n, int n_ranks, int rank, size_t* lo, size_t* hi)
{
    // contiguous, order-preserving slices; the same arithmetic as engine_b200/dist.py shard_bounds (n * r // world)
    const unsigned __int128 a = (unsigned __int128)n * (unsigned)rank / (unsigned)n_ranks;
    const unsigned __int128 b = (unsigned __int128)n * ((unsigned)rank + 1u) / (unsigned)n_ranks;
    if (lo) *lo = (size_t)a;
    if (hi) *hi = (size_t)b;
}

b2col_status b2col_comm_get_unique_id(void* id128)
{
    b2col_ctx* ctx = nullptr;
    if (!id128) return fail(nullptr, B2COL_ERR_INVALID, "null argument");
    if (!nccl_load()) return fail(nullptr, B2COL_ERR_COMM, "%s", nccl_api().error.c_str());
    b2_ncclUniqueId id;
    NC(nccl_api().GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return B2COL_OK;
}

b2col_status b2col_comm_init_rank(b2col_ctx* ctx, int n_ranks, int rank, const void* id128)
{
    if (!ctx || !id128 || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(ctx, B2COL_ERR_INVALID, "bad communicator arguments");
    if (ctx->comm) return fail(ctx, B2COL_ERR_STATE, "this context already has a communicator");
    if (!nccl_load()) return fail(ctx, B2COL_ERR_COMM, "%s", nccl_api().error.c_str());
    CU(cudaSetDevice(ctx->device));
    b2_ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    b2_ncclComm_t comm = nullptr;
    NC(nccl_api().CommInitRank(&comm, n_ranks, id, rank));
    return comm_attach(ctx, comm, n_ranks, rank);
}

b2col_status b2col_comm_init_all(b2col_ctx* const* ctxs, int n)
{
    b2col_ctx* ctx = (ctxs && n > 0) ? ctxs[0] : nullptr;
    if (!ctx || n < 1 || n > 64) return fail(ctx, B2COL_ERR_INVALID, "bad context list");
    for (int i = 0; i < n; ++i) {
        if (!ctxs[i]) return fail(ctx, B2COL_ERR_INVALID, "null context in the list");
        if (ctxs[i]->comm) return fail(ctx, B2COL_ERR_STATE, "context %d already has a communicator", i);
        for (int j = 0; j < i; ++j)
            if (ctxs[j]->device == ctxs[i]->device) return fail(ctx, B2COL_ERR_INVALID, "contexts %d and %d share device %d: one rank per GPU", j, i, ctxs[i]->device);
    }
    if (!nccl_load()) return fail(ctx, B2COL_ERR_COMM, "%s", nccl_api().error.c_str());
    std::vector<int> devs((size_t)n);
    std::vector<b2_ncclComm_t> comms((size_t)n, nullptr);
    for (int i = 0; i < n; ++i) devs[(size_t)i] = ctxs[i]->device;
    NC(nccl_api().CommInitAll(comms.data(), n, devs.data()));
    std::vector<b2col_ctx*> peers(ctxs, ctxs + n);
    for (int i = 0; i < n; ++i) {
        GUARD(comm_attach(ctxs[i], comms[(size_t)i], n, i));
        ctxs[i]->comm->peers = peers;
        cudaSetDevice(ctxs[i]->device);
        for (int j = 0; j < n; ++j) {  // the gather window is written through peer pointers
            if (j == i) continue;
            const cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[j]->device, 0);
            if (e != cudaSuccess) {
                cudaGetLastError();
                if (e != cudaErrorPeerAccessAlreadyEnabled) ctxs[i]->comm->win_fallback = true;
            }
        }
    }
    bool any_fallback = false;
    for (int i = 0; i < n; ++i) any_fallback = any_fallback || ctxs[i]->comm->win_fallback;
    for (int i = 0; i < n; ++i) ctxs[i]->comm->win_fallback = any_fallback;
    return B2COL_OK;
}

b2col_status b2col_comm_finalize(b2col_ctx* ctx)
{
    if (!ctx || !ctx->comm) return B2COL_OK;
    b2col_comm_state* c = ctx->comm;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (!c->peers.empty() && c->win && c->rank == c->win_dst) {  // single process: the other ranks hold plain copies of this pointer
        for (b2col_ctx* p : c->peers)
            if (p && p != ctx && p->comm && p->comm->win == c->win) {
                cudaSetDevice(p->device);
                cudaStreamSynchronize(p->stream);
                p->comm->win = nullptr;
                p->comm->win_bytes = 0;
            }
        cudaSetDevice(ctx->device);
    }
    if (c->peers.empty() || c->rank == c->win_dst) comm_release_window(ctx);
    for (b2col_ctx* p : c->peers)
        if (p && p != ctx && p->comm) {
            auto& v = p->comm->peers;
            for (auto& q : v)
                if (q == ctx) q = nullptr;
        }
    if (c->d_word) cudaFree(c->d_word);
    if (c->done) cudaEventDestroy(c->done);
    if (c->comm) nccl_api().CommDestroy(c->comm);
    delete c;
    ctx->comm = nullptr;
    return B2COL_OK;
}

b2col_status b2col_comm_info(b2col_ctx* ctx, b2col_comm_info_t* out)
{
    if (!ctx || !out) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    memset(out, 0, sizeof *out);
    if (!ctx->comm) return B2COL_OK;
    const b2col_comm_state* c = ctx->comm;
    out->n_ranks = c->n_ranks;
    out->rank = c->rank;
    out->single_process = c->peers.empty() ? 0 : 1;
    out->gather_mode = c->win_fallback ? B2COL_GATHER_NCCL_SENDRECV : (c->peers.empty() ? B2COL_GATHER_PEER_WINDOW_IPC : B2COL_GATHER_PEER_WINDOW);
    out->window_bytes = c->win_bytes;
    out->gathers = c->gathers;
    out->replications = c->replications;
    int v = 0;
    if (nccl_api().GetVersion && nccl_api().GetVersion(&v) == 0) out->nccl_version = v;
    return B2COL_OK;
}

b2col_status b2col_comm_replicate(b2col_ctx* ctx, int src_rank, uint32_t n_colliders)
{
    GUARD(comm_require(ctx));
    if (src_rank < 0 || src_rank >= ctx->comm->n_ranks) return fail(ctx, B2COL_ERR_INVALID, "bad source rank");
    if (!ctx->comm->peers.empty()) return fail(ctx, B2COL_ERR_STATE, "contexts of one process replicate with b2col_comm_replicate_all (one NCCL group)");
    return comm_enqueue_broadcast_blob(ctx, src_rank, n_colliders);
}

b2col_status b2col_comm_replicate_all(b2col_ctx* const* ctxs, int n, int src_rank)
{
    b2col_ctx* ctx = (ctxs && n > 0) ? ctxs[0] : nullptr;
    if (!ctx || src_rank < 0 || src_rank >= n) return fail(ctx, B2COL_ERR_INVALID, "bad arguments");
    for (int i = 0; i < n; ++i) {
        GUARD(comm_require(ctxs[i]));
        if (ctxs[i]->comm->rank != i || ctxs[i]->comm->n_ranks != n) return fail(ctx, B2COL_ERR_INVALID, "contexts must be the list given to b2col_comm_init_all, in the same order");
    }
    b2col_ctx* src = ctxs[src_rank];
    if (!src->have_tree) return fail(ctx, B2COL_ERR_STATE, "the source context holds no tree: call b2col_update first");
    const uint32_t n_colliders = src->tree_leaves;
    NC(nccl_api().GroupStart());
    b2col_status st = B2COL_OK;
    for (int i = 0; i < n && st == B2COL_OK; ++i) st = comm_enqueue_broadcast_blob(ctxs[i], src_rank, n_colliders);
    const int r = nccl_api().GroupEnd();
    if (st != B2COL_OK) {
        if (ctx != ctxs[0]) ctx->err = "replicate failed on another context";
        return st;
    }
    if (r != 0) return fail(ctx, B2COL_ERR_COMM, "ncclGroupEnd failed: %s", nccl_api().GetErrorString(r));
    return B2COL_OK;
}

b2col_status b2col_comm_replicate_records(b2col_ctx* ctx, int src_rank, uint32_t n_colliders)
{
    GUARD(comm_require(ctx));
    b2col_comm_state* c = ctx->comm;
    if (src_rank < 0 || src_rank >= c->n_ranks) return fail(ctx, B2COL_ERR_INVALID, "bad source rank");
    if (!c->peers.empty()) return fail(ctx, B2COL_ERR_STATE, "b2col_comm_replicate_records is a per-process collective; contexts of one process use b2col_comm_replicate_all");
    if (n_colliders < 2) return comm_enqueue_broadcast_blob(ctx, src_rank, n_colliders);
    CU(cudaSetDevice(ctx->device));
    const size_t blob_bytes = blob_size_for(n_colliders);
    if (c->rank == src_rank) {
        if (!ctx->have_tree || !ctx->topology_valid || ctx->tree_leaves != n_colliders)
            return fail(ctx, B2COL_ERR_STATE, "source rank needs a tree built by b2col_update for %u colliders", n_colliders);
        // the receivers rebuild with the static key width (contexts with a communicator never adapt it, prepare_build_inner): a tree
        // built at another width BEFORE the communicator existed would differ from theirs in leaf order, i.e. in tie winners
        const int static_bits = n_colliders <= (1u << 21) ? 32 : (n_colliders <= (1u << 25) ? 40 : 48);
        if (ctx->cfg_bits == 0 && ctx->env_key_bits == 0 && ctx->tree_key_bits != static_bits)
            return fail(ctx, B2COL_ERR_STATE, "the source's tree was built with %d-bit keys before b2col_comm_init; call b2col_update again", ctx->tree_key_bits);
    }
    else {
        const void* before = ctx->d_blob.p;
        const b2col_status st = ensure(ctx, ctx->d_blob, blob_bytes);
        if (ctx->d_blob.p != before || st != B2COL_OK) ctx->have_tree = false;
        GUARD(st);
        GUARD(prepare_build(ctx, n_colliders));
        ctx->have_tree = false;
        ctx->topology_valid = false;
        ctx->exact_nodes_valid = false;
    }
    // the world records (32 B per collider) and the source's blob header, which carries the scene bounds of this frame
    NC(nccl_api().GroupStart());
    const int r1 = nccl_api().Broadcast(ctx->d_world_aabb.p, ctx->d_world_aabb.p, (size_t)n_colliders * sizeof(LeafRec), B2_NCCL_UINT8, src_rank, c->comm, ctx->stream);
    const int r2 = nccl_api().Broadcast(ctx->d_blob.p, ctx->d_blob.p, sizeof(BlobHeader), B2_NCCL_UINT8, src_rank, c->comm, ctx->stream);
    const int r3 = nccl_api().GroupEnd();
    if (r1 || r2 || r3) return fail(ctx, B2COL_ERR_COMM, "ncclBroadcast of the world records failed: %s", nccl_api().GetErrorString(r1 ? r1 : (r2 ? r2 : r3)));
    if (c->rank != src_rank) {
        GUARD(enqueue_frame(ctx, true, n_colliders));  // Morton + sort + hierarchy; deterministic: the same tree as on src_rank
        ctx->blob_bytes = blob_bytes;
        ctx->tree_leaves = n_colliders;
        ctx->have_tree = true;
    }
    ++c->replications;
    return B2COL_OK;
}

b2col_status b2col_comm_raycast_gather(b2col_ctx* ctx, const b2col_ray* d_rays_shard, size_t n_total, int dst_rank, const b2col_raycast** d_out_all)
{
    GUARD(comm_require(ctx));
    b2col_comm_state* c = ctx->comm;
    if (dst_rank < 0 || dst_rank >= c->n_ranks) return fail(ctx, B2COL_ERR_INVALID, "bad destination rank");
    if (!c->peers.empty()) return fail(ctx, B2COL_ERR_STATE, "contexts of one process gather with b2col_comm_raycast_gather_all");
    if (d_out_all) *d_out_all = nullptr;
    if (n_total == 0) return B2COL_OK;
    size_t lo = 0, n_mine = 0;
    GUARD(comm_enqueue_shard(ctx, d_rays_shard, n_total, dst_rank, &lo, &n_mine));
    if (c->win_fallback) {
        NC(nccl_api().GroupStart());
        const b2col_status st = comm_enqueue_sendrecv(ctx, n_total, dst_rank);
        const int r = nccl_api().GroupEnd();
        GUARD(st);
        if (r != 0) return fail(ctx, B2COL_ERR_COMM, "ncclGroupEnd failed: %s", nccl_api().GetErrorString(r));
    }
    else {
        // completion of the fused gather: when this all-reduce has passed on dst's stream, every rank's trace kernel — enqueued
        // before its own contribution — has finished, i.e. its stores into the window have been performed
        NC(nccl_api().AllReduce(c->d_word, c->d_word, 1, B2_NCCL_UINT32, B2_NCCL_MAX, c->comm, ctx->stream));
    }
    ++c->gathers;
    if (d_out_all && c->rank == dst_rank) *d_out_all = static_cast<const b2col_raycast*>(c->win);
    return B2COL_OK;
}

b2col_status b2col_comm_raycast_gather_all(b2col_ctx* const* ctxs, int n, const b2col_ray* const* d_rays_shards, size_t n_total, int dst_rank, const b2col_raycast** d_out_all)
{
    b2col_ctx* ctx = (ctxs && n > 0) ? ctxs[0] : nullptr;
    if (!ctx || !d_rays_shards || dst_rank < 0 || dst_rank >= n) return fail(ctx, B2COL_ERR_INVALID, "bad arguments");
    for (int i = 0; i < n; ++i) {
        GUARD(comm_require(ctxs[i]));
        if (ctxs[i]->comm->rank != i || ctxs[i]->comm->n_ranks != n) return fail(ctx, B2COL_ERR_INVALID, "contexts must be the list given to b2col_comm_init_all, in the same order");
    }
    if (d_out_all) *d_out_all = nullptr;
    if (n_total == 0) return B2COL_OK;
    b2col_ctx* dst = ctxs[dst_rank];
    // the destination first: it allocates the window the others write into
    for (int k = 0; k < n; ++k) {
        const int i = k == 0 ? dst_rank : (k <= dst_rank ? k - 1 : k);
        size_t lo = 0, n_mine = 0;
        const b2col_status st = comm_enqueue_shard(ctxs[i], d_rays_shards[i], n_total, dst_rank, &lo, &n_mine);
        if (st != B2COL_OK) {
            if (ctxs[i] != ctx) ctx->err = ctxs[i]->err;
            return st;
        }
    }
    if (dst->comm->win_fallback) {
        NC(nccl_api().GroupStart());
        b2col_status st = B2COL_OK;
        for (int i = 0; i < n && st == B2COL_OK; ++i) st = comm_enqueue_sendrecv(ctxs[i], n_total, dst_rank);
        const int r = nccl_api().GroupEnd();
        GUARD(st);
        if (r != 0) return fail(ctx, B2COL_ERR_COMM, "ncclGroupEnd failed: %s", nccl_api().GetErrorString(r));
    }
    else {
        // completion inside one process: dst's stream waits for an event behind every other rank's trace kernel
        for (int i = 0; i < n; ++i) {
            if (i == dst_rank) continue;
            CU(cudaSetDevice(ctxs[i]->device));
            CU(cudaEventRecord(ctxs[i]->comm->done, ctxs[i]->stream));
            CU(cudaSetDevice(dst->device));
            CU(cudaStreamWaitEvent(dst->stream, ctxs[i]->comm->done, 0));
        }
    }
    for (int i = 0; i < n; ++i) ++ctxs[i]->comm->gathers;
    if (d_out_all) *d_out_all = static_cast<const b2col_raycast*>(dst->comm->win);
    return B2COL_OK;
}

// Host buffers, several GPUs of one process (the shim's GetRaycasts): the batch is cut into contiguous slices, every
// context runs its own upload / trace / download pipeline on its slice, straight from and into the caller's buffers —
// there is nothing to gather.  All pipelines are enqueued before the first wait.  Needs replicated trees
// (b2col_comm_replicate_all) or contexts that each ran b2col_update on the same scene.
b2col_status b2col_mgpu_raycast_batch(b2col_ctx* const* ctxs, int n, const b2col_ray* h_rays, size_t n_rays, b2col_raycast* h_out, int mode)
{
    b2col_ctx* ctx = (ctxs && n > 0) ? ctxs[0] : nullptr;
    if (!ctx || (n_rays && (!h_rays || !h_out))) return fail(ctx, B2COL_ERR_INVALID, "null argument");
    if (mode != B2COL_MODE_LBVH) return fail(ctx, B2COL_ERR_INVALID, "multi-GPU batches run the production tree only");
    for (int i = 0; i < n; ++i)
        if (!ctxs[i]) return fail(ctx, B2COL_ERR_INVALID, "null context in the list");
    if (n_rays == 0) return B2COL_OK;
    // small batches are not worth a second GPU: one slice must keep a pipeline busy
    int use = n;
    while (use > 1 && n_rays / (size_t)use < (size_t)4 * ctx->e2e_min_chunk) --use;
    if (use == 1) return b2col_raycast_batch(ctx, h_rays, n_rays, h_out, mode);
    b2col_status st = B2COL_OK;
    int begun = 0;
    for (int i = 0; i < use && st == B2COL_OK; ++i) {
        size_t lo = 0, hi = 0;
        b2col_shard_bounds(n_rays, use, i, &lo, &hi);
        st = raycast_batch_begin(ctxs[i], h_rays + lo, hi - lo, h_out + lo, mode);
        if (st == B2COL_OK) ++begun;
        else if (ctxs[i] != ctx) ctx->err = ctxs[i]->err;
    }
    for (int i = 0; i < begun; ++i) {
        const b2col_status s2 = raycast_batch_end(ctxs[i]);
        if (s2 != B2COL_OK && st == B2COL_OK) {
            st = s2;
            if (ctxs[i] != ctx) ctx->err = ctxs[i]->err;
        }
    }
    return st;
}

}  // extern "C"
