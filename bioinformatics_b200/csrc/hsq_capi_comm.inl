// hsq_capi_comm.inl -- SURVEY.md 8(e): the multi-GPU UID dedup behind the C ABI.  One context per GPU; the contexts of a job share an
// NCCL communicator (one process per GPU: hsq_comm_init_rank; one process driving all GPUs, the shape of the reference's single JVM
// fanning probes out to a thread pool, HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:836-846: hsq_comm_init_all).
//   reads never move.  Every rank turns its pairs into 24-byte key records (probe, uid, rank, origin), buckets them by owner
//   (probe mod world) with device kernels, exchanges counts (ncclAllGather) and payload (ncclGroupStart + ncclSend / ncclRecv on device
//   buffers), the owner reduces its (probe, UID) groups (findBestData, FilterByUid.java:281-297) and the answers travel back along
//   the same routes.  Per-probe counters are summed over ranks (each probe is counted on exactly one).
// NCCL is loaded with dlopen at the first hsq_comm_* call: a single-GPU host needs no NCCL, and a process that already holds a copy
// (e.g. the one bundled with PyTorch) shares it instead of loading a second one.
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>

namespace {

struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    std::string error;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) {
            api.error = std::string("NCCL is not available: ") + dlerror();
            return;
        }
        bool ok = true;
        auto sym = [&](const char* n) {
            void* p = dlsym(api.handle, n);
            if (!p) { ok = false; api.error = std::string("NCCL symbol missing: ") + n; }
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        if (!ok) {
            dlclose(api.handle);
            api.handle = nullptr;
        }
    });
    return api.handle ? &api : nullptr;
}

#define HSQ_NCCL(ctx, expr)                                                                                       \
    do {                                                                                                          \
        ncclResult_t _r = (expr);                                                                                 \
        if (_r != ncclSuccess) return hsq::fail(ctx, HSQ_ERR_CUDA, std::string(#expr) + ": " + nccl_api()->GetErrorString(_r)); \
    } while (0)

int need_nccl(hsq_ctx* ctx) {
    if (nccl_api()) return HSQ_OK;
    static NcclApi* dummy = nullptr;
    (void)dummy;
    return fail(ctx, HSQ_ERR_UNSUPPORTED, "NCCL could not be loaded (libnccl.so.2): the multi-GPU entry points are unavailable");
}

void comm_release(hsq_ctx* ctx) {
    if (ctx->comm && nccl_api()) nccl_api()->CommDestroy(static_cast<ncclComm_t>(ctx->comm));
    ctx->comm = nullptr;
    ctx->comm_world = 1;
    ctx->comm_rank = 0;
    if (ctx->h_counts) cudaFreeHost(ctx->h_counts);
    ctx->h_counts = nullptr;
}

int comm_adopt(hsq_ctx* ctx, ncclComm_t comm, int rank, int world) {
    comm_release(ctx);
    ctx->comm = comm;
    ctx->comm_rank = rank;
    ctx->comm_world = world;
    HSQ_CUDA(ctx, cudaMallocHost((void**)&ctx->h_counts, sizeof(int64_t) * (size_t)world * (size_t)(world + 1)));
    return HSQ_OK;
}

}  // namespace

namespace hsq {

// send_cnt[o] = pairs of this rank whose probe is owned by rank o; within[i] = arrival number of pair i inside its owner's bucket
__global__ void owner_hist_kernel(const int32_t* __restrict__ probe, int64_t n, int world, unsigned long long* __restrict__ send_cnt,
                                  uint32_t* __restrict__ within) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (; i < n; i += stride) {
        const int o = probe[i] % world;
        const unsigned peers = __match_any_sync(__activemask(), o);
        const int leader = __ffs(peers) - 1;
        unsigned long long base = 0;
        if (lane == leader) base = atomicAdd(&send_cnt[o], (unsigned long long)__popc(peers));
        base = __shfl_sync(peers, base, leader);
        within[i] = (uint32_t)base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
    }
}

// key columns -> 24-byte records at their place in the send buffer (bucket of the owner + arrival number); place[i] remembers it
__global__ void owner_pack_kernel(const uint64_t* __restrict__ uid, const uint64_t* __restrict__ rank, const int32_t* __restrict__ probe,
                                  int64_t n, int world, const int64_t* __restrict__ send_off, uint32_t* __restrict__ within_then_place,
                                  uint32_t origin_base, hsq_dedup_key* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int32_t p = probe[i];
        const uint32_t at = (uint32_t)send_off[p % world] + within_then_place[i];
        hsq_dedup_key k;
        k.uid = uid[i];
        k.rank = rank[i];
        k.probe = p;
        k.origin = origin_base + (uint32_t)i;
        out[at] = k;
        within_then_place[i] = at;
    }
}

__global__ void answer_pack_kernel(const uint8_t* __restrict__ best, const int32_t* __restrict__ gs, int64_t n, uint32_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = (uint32_t)gs[i] | (best[i] ? 0x80000000u : 0u);
}

__global__ void answer_unpack_kernel(const uint32_t* __restrict__ back, const uint32_t* __restrict__ place, int64_t n, uint8_t* __restrict__ best,
                                     int32_t* __restrict__ gs) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const uint32_t a = back[place[i]];
        best[i] = (uint8_t)(a >> 31);
        gs[i] = (int32_t)(a & 0x7fffffffu);
    }
}

}  // namespace hsq

namespace {

// The collective half of the sharded dedup.  In: this rank's key columns on the device (uid -> sort_a, rank -> sort_b, probe ->
// ctx->probe_id, n pairs).  Out: d_best / d_gs for this rank's pairs (device) and the GLOBAL per-probe counters d_stats[3 * np]
// (device, identical on every rank).  ctx->probe_id is preserved.  Every rank of the communicator must call it.
int dedup_exchange_reduce(hsq_ctx* ctx, int64_t n, int64_t np, int uid_len, uint8_t* d_best, int32_t* d_gs, int32_t* d_stats) {
    NcclApi* nc = nccl_api();
    ncclComm_t comm = static_cast<ncclComm_t>(ctx->comm);
    const int W = ctx->comm_world, me = ctx->comm_rank;
    const size_t nn = (size_t)n;
    int rc;
    if (n >= (1ll << 31)) return fail(ctx, HSQ_ERR_UNSUPPORTED, "more than 2^31 pairs on one rank");
    span_begin(ctx, 1);
    // counts: row of this rank = [pairs for owner 0 .. W-1, n]
    HSQ_CUDA(ctx, ctx->cm_cnt.ensure(sizeof(int64_t) * (size_t)(W + 1) * (size_t)(W + 2) + 64));
    unsigned long long* my_row = ctx->cm_cnt.as<unsigned long long>();
    int64_t* all_rows = ctx->cm_cnt.as<int64_t>() + (W + 1);
    int64_t* d_send_off = all_rows + (size_t)W * (W + 1);
    HSQ_CUDA(ctx, cudaMemsetAsync(my_row, 0, sizeof(int64_t) * (size_t)(W + 1), ctx->stream));
    HSQ_CUDA(ctx, ctx->cm_i32.ensure(nn * 4 + 16));  // within, then place
    uint32_t* place = ctx->cm_i32.as<uint32_t>();
    const int gblocks = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16));
    if (n > 0) {
        owner_hist_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, W, my_row, place);
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaMemcpyAsync(my_row + W, &n, 8, cudaMemcpyHostToDevice, ctx->stream));  // n lives on the caller's stack until the sync below
    HSQ_NCCL(ctx, nc->AllGather(my_row, all_rows, (size_t)(W + 1), ncclInt64, comm, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_counts, all_rows, sizeof(int64_t) * (size_t)W * (size_t)(W + 1), cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t* C = ctx->h_counts;  // C[r * (W + 1) + o] = records rank r sends to owner o; C[r * (W + 1) + W] = pairs of rank r
    std::vector<int64_t> send_off((size_t)W + 1, 0), recv_off((size_t)W + 1, 0);
    int64_t origin_base = 0, total_pairs = 0;
    for (int r = 0; r < W; r++) {
        send_off[(size_t)r + 1] = send_off[(size_t)r] + C[(size_t)me * (W + 1) + r];
        recv_off[(size_t)r + 1] = recv_off[(size_t)r] + C[(size_t)r * (W + 1) + me];
        if (r < me) origin_base += C[(size_t)r * (W + 1) + W];
        total_pairs += C[(size_t)r * (W + 1) + W];
    }
    if (total_pairs >= (1ll << 32)) return fail(ctx, HSQ_ERR_UNSUPPORTED, "more than 2^32 pairs in one sharded batch");
    const int64_t n_recv = recv_off[(size_t)W];
    if (send_off[(size_t)W] != n) return fail(ctx, HSQ_ERR_CUDA, "sharded dedup: bucket counts do not add up");
    HSQ_CUDA(ctx, cudaMemcpyAsync(d_send_off, send_off.data(), sizeof(int64_t) * (size_t)(W + 1), cudaMemcpyHostToDevice, ctx->stream));
    HSQ_CUDA(ctx, ctx->cm_send.ensure(std::max(nn, (size_t)1) * sizeof(hsq_dedup_key)));
    HSQ_CUDA(ctx, ctx->cm_recv.ensure(std::max((size_t)n_recv, (size_t)1) * sizeof(hsq_dedup_key)));
    hsq_dedup_key* sendbuf = ctx->cm_send.as<hsq_dedup_key>();
    hsq_dedup_key* recvbuf = ctx->cm_recv.as<hsq_dedup_key>();
    if (n > 0) {
        owner_pack_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->sort_a.as<uint64_t>(), ctx->sort_b.as<uint64_t>(), ctx->probe_id.as<int32_t>(), n, W,
                                                           d_send_off, place, (uint32_t)origin_base, sendbuf);
        ctx->launches++;
    }
    // this rank's probe ids survive the reduction of the received records (which reuses ctx->probe_id)
    HSQ_CUDA(ctx, ctx->cm_probe.ensure(nn * 4 + 16));
    if (n > 0) HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->cm_probe.p, ctx->probe_id.p, nn * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // send_off (host vector) has been read
    // payload all-to-all
    HSQ_NCCL(ctx, nc->GroupStart());
    for (int r = 0; r < W; r++) {
        const int64_t sc = C[(size_t)me * (W + 1) + r], rcnt = C[(size_t)r * (W + 1) + me];
        if (sc > 0) HSQ_NCCL(ctx, nc->Send(sendbuf + send_off[(size_t)r], (size_t)sc * sizeof(hsq_dedup_key), ncclUint8, r, comm, ctx->stream));
        if (rcnt > 0) HSQ_NCCL(ctx, nc->Recv(recvbuf + recv_off[(size_t)r], (size_t)rcnt * sizeof(hsq_dedup_key), ncclUint8, r, comm, ctx->stream));
    }
    HSQ_NCCL(ctx, nc->GroupEnd());
    // owner side: group + reduce the received records
    const size_t nr = (size_t)n_recv;
    HSQ_CUDA(ctx, ctx->sort_a.ensure(nr * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nr * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_d.ensure(nr * 8 + 16));
    HSQ_CUDA(ctx, ctx->probe_id.ensure(nr * 4 + 16));
    HSQ_CUDA(ctx, ctx->cm_tie.ensure(nr * 4 + 16));
    HSQ_CUDA(ctx, ctx->out_slots.ensure(nr * 9 + nn * 4 + 64));  // owner-side answers: best (u8), group size (i32), packed answers (u32) | answers that came back
    uint8_t* o_best = ctx->out_slots.as<uint8_t>();
    int32_t* o_gs = reinterpret_cast<int32_t*>(o_best + ((nr + 15) & ~(size_t)15));
    uint32_t* o_ans = reinterpret_cast<uint32_t*>(o_gs + nr);
    uint32_t* back = o_ans + nr;
    const int rblocks = (int)std::max<int64_t>(1, std::min<int64_t>((n_recv + 255) / 256, (int64_t)ctx->sm_count * 16));
    if (n_recv > 0) {
        dedup_unpack_keys_kernel<<<rblocks, 256, 0, ctx->stream>>>(recvbuf, n_recv, np, ctx->sort_a.as<uint64_t>(), ctx->sort_b.as<uint64_t>(),
                                                                  ctx->probe_id.as<int32_t>(), ctx->sort_d.as<uint32_t>(), ctx->cm_tie.as<uint32_t>(),
                                                                  d_status(ctx));
        ctx->launches++;
    }
    if ((rc = dedup_reduce_device(ctx, n_recv, np, uid_len, o_best, o_gs, d_stats, ctx->cm_tie.as<uint32_t>()))) return rc;
    if (n_recv > 0) {
        answer_pack_kernel<<<rblocks, 256, 0, ctx->stream>>>(o_best, o_gs, n_recv, o_ans);
        ctx->launches++;
    }
    // answers back along the same routes
    HSQ_NCCL(ctx, nc->GroupStart());
    for (int r = 0; r < W; r++) {
        const int64_t sc = C[(size_t)me * (W + 1) + r], rcnt = C[(size_t)r * (W + 1) + me];
        if (rcnt > 0) HSQ_NCCL(ctx, nc->Send(o_ans + recv_off[(size_t)r], (size_t)rcnt, ncclUint32, r, comm, ctx->stream));
        if (sc > 0) HSQ_NCCL(ctx, nc->Recv(back + send_off[(size_t)r], (size_t)sc, ncclUint32, r, comm, ctx->stream));
    }
    HSQ_NCCL(ctx, nc->GroupEnd());
    if (n > 0) {
        answer_unpack_kernel<<<gblocks, 256, 0, ctx->stream>>>(back, place, n, d_best, d_gs);
        ctx->launches++;
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->probe_id.p, ctx->cm_probe.p, nn * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    // every probe was counted on its owner alone: a sum gives all ranks the whole table
    if (np > 0) HSQ_NCCL(ctx, nc->AllReduce(d_stats, d_stats, (size_t)(3 * np), ncclInt32, ncclSum, comm, ctx->stream));
    span_end(ctx, 1);
    HSQ_CUDA(ctx, cudaGetLastError());
    int w[1];
    if ((rc = read_status(ctx, w, 1))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], "sharded dedup: probe out of range in a key record");
    return HSQ_OK;
}

int comm_ready(hsq_ctx* ctx) {
    if (!ctx->comm) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "this context has no communicator: call hsq_comm_init_rank / hsq_comm_init_all first");
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_comm_unique_id(uint8_t* id_out) {
    if (!id_out) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    int rc = need_nccl(nullptr);
    if (rc) return rc;
    static_assert(HSQ_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "id size");
    ncclUniqueId id;
    HSQ_NCCL(nullptr, nccl_api()->GetUniqueId(&id));
    std::memcpy(id_out, id.internal, NCCL_UNIQUE_ID_BYTES);
    return HSQ_OK;
}

extern "C" int32_t hsq_comm_init_rank(hsq_ctx* ctx, const uint8_t* id_in, int32_t rank, int32_t world) {
    if (!ctx || !id_in || world < 1 || rank < 0 || rank >= world) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "bad argument");
    int rc = need_nccl(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId id;
    std::memcpy(id.internal, id_in, NCCL_UNIQUE_ID_BYTES);
    ncclComm_t comm = nullptr;
    HSQ_NCCL(ctx, nccl_api()->CommInitRank(&comm, world, id, rank));
    return comm_adopt(ctx, comm, rank, world);
}

extern "C" int32_t hsq_comm_init_all(hsq_ctx** ctxs, int32_t n) {
    if (!ctxs || n < 1) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "bad argument");
    for (int k = 0; k < n; k++)
        if (!ctxs[k]) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "null context");
    int rc = need_nccl(ctxs[0]);
    if (rc) return rc;
    std::vector<int> devs((size_t)n);
    for (int k = 0; k < n; k++) devs[(size_t)k] = ctxs[k]->device;
    std::vector<ncclComm_t> comms((size_t)n, nullptr);
    HSQ_NCCL(ctxs[0], nccl_api()->CommInitAll(comms.data(), n, devs.data()));
    for (int k = 0; k < n; k++) {
        cudaSetDevice(ctxs[k]->device);
        if ((rc = comm_adopt(ctxs[k], comms[(size_t)k], k, n))) return rc;
    }
    return HSQ_OK;
}

extern "C" int32_t hsq_comm_destroy(hsq_ctx* ctx) {
    if (!ctx) return HSQ_ERR_INVALID_ARGUMENT;
    cudaSetDevice(ctx->device);
    comm_release(ctx);
    return HSQ_OK;
}

// FilterByUid.reduceReadsByProbeAndUid over the union of all ranks' pairs (collective: every rank of the communicator calls it with its
// own pairs).  is_best / group_size answer this rank's pairs; stats are the counters of the whole batch, identical on every rank.
extern "C" int32_t hsq_dedup_sharded(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t u, int32_t v, uint8_t* is_best, int32_t* group_size,
                                     hsq_probe_stats* stats) {
    if (!ctx || !pairs || (pairs->n > 0 && !is_best)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (pairs->n < 0 || pairs->r1.n != pairs->n || pairs->r2.n != pairs->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (u < 0 || v < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative UID length");
    if (u + v > 32) return fail(ctx, HSQ_ERR_UNSUPPORTED, "UID longer than 32 bases in total is outside the accelerated domain");
    if (!stats || stats->n_probes <= 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe stats (with n_probes) are required");
    int rc = comm_ready(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const size_t nn = (size_t)pairs->n;
    const int64_t np = stats->n_probes;
    auto run = [&]() -> int {
        int rc;
        // a rank whose own batch is bad must still take part in the collectives, or the others hang: it contributes no pairs
        int bad = dedup_upload_keygen(ctx, pairs, u, v, np);
        const std::string bad_msg = ctx->error;
        HSQ_CUDA(ctx, ctx->dd_u8.ensure(nn + 32));
        HSQ_CUDA(ctx, ctx->dd_i32.ensure((nn + 3 * (size_t)np) * 4 + 16));
        uint8_t* d_best = ctx->dd_u8.as<uint8_t>();
        int32_t* d_gs = ctx->dd_i32.as<int32_t>();
        int32_t* d_stats = d_gs + nn;
        if (bad) HSQ_CUDA(ctx, cudaMemsetAsync(d_status(ctx), 0, 4, ctx->stream));
        if ((rc = dedup_exchange_reduce(ctx, bad ? 0 : pairs->n, np, u + v, d_best, d_gs, d_stats))) return rc;
        if (bad) return fail(ctx, bad, bad_msg);
        if ((rc = d2h(ctx, is_best, d_best, nn))) return rc;
        if ((rc = d2h(ctx, group_size, d_gs, nn))) return rc;
        return fetch_stats(ctx, stats, d_stats);
    };
    rc = run();
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}
