// bm_comm.cu -- multi-GPU plumbing of libbucketmap_b200.so (no reference counterpart: the reference is one process).
//
// Reads are independent units: every GPU holds the whole index and maps its own contiguous shard of the reads.
// Two exchanges exist, both outside the per-read work:
//
//   bm_bcast_index / bm_bcast_ref   the q-gram bit matrix (+ kmer_to_index, geometry) and the 2-bit reference go from the
//                                   rank that built / loaded them to all others with ncclBroadcast over NVLink.  NCCL is
//                                   dlopen'ed (libnccl.so.2) at bm_comm_init, so single-GPU users need no NCCL at all.
//
//   bm_gather_results               the location records of every rank end up on one rank, in read order.  This is NOT a
//                                   collective call: the last stage of each rank's locator pass writes its records and
//                                   per-read offsets straight into the root's memory over NVLink (peer mapping: CUDA IPC
//                                   between processes, cudaDeviceEnablePeerAccess inside one process) and then bumps a
//                                   sequence flag there; the root's gather stream waits on the flags with
//                                   cuStreamWaitValue32 (no kernel spins, no host synchronisation, no receive to post), packs
//                                   the slots into one contiguous (loc_off, locs) pair and acknowledges, which frees the slot
//                                   for the batch after next (two slot sets: the gather of batch i runs under the mapper of
//                                   batch i+1 on every rank).
#include <cuda.h>
#include <dlfcn.h>
#include <nccl.h>
#include <unistd.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "bm_internal.cuh"

int bm_internal_index_set_table(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint32_t NB);

namespace {

// ---- NCCL through dlopen ----------------------------------------------------------------------------------------------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi& nccl() {
    static NcclApi api;
    if (api.lib || api.ok) return api;
    // a process that already holds an NCCL (torch brings its own) keeps using that one
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return api;
    api.lib = h;
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
    api.Broadcast = (decltype(api.Broadcast))dlsym(h, "ncclBroadcast");
    api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Broadcast && api.GetErrorString;
    return api;
}

#define BM_NCCL(ctx, call)                                                                            \
    do {                                                                                              \
        ncclResult_t _r = (call);                                                                     \
        if (_r != ncclSuccess) {                                                                      \
            (ctx)->err = std::string(#call) + ": " + nccl().GetErrorString(_r);                       \
            return BM_E_CUDA;                                                                         \
        }                                                                                             \
    } while (0)

// ---- driver entry points for stream memory operations -------------------------------------------------------------------------
typedef int (*StreamWaitValue32_t)(cudaStream_t, unsigned long long /*CUdeviceptr*/, uint32_t, unsigned int);
constexpr unsigned int kWaitGeq = CU_STREAM_WAIT_VALUE_GEQ;   // (int32_t)(*addr - value) >= 0

StreamWaitValue32_t wait_value_fn() {
    static StreamWaitValue32_t fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (StreamWaitValue32_t)p;
        else cudaGetLastError();
    }
    return fn;
}

// ---- gather state ---------------------------------------------------------------------------------------------------------------
constexpr uint32_t kHandleMagic = 0x424D4731u;   // "BMG1"
constexpr int kSlotSets = 2;

struct PeerHandle {                 // BM_PEER_HANDLE_BYTES = 128
    uint32_t magic, pid;
    int32_t device, rank;
    uint64_t ptr, bytes;            // the rank's exported block (root: slots + flags; others: ack word)
    uint32_t max_reads, pad;
    uint64_t max_locs;
    cudaIpcMemHandle_t ipc;         // 64 bytes
    uint8_t fill[128 - 48 - 64];
};
static_assert(sizeof(PeerHandle) == 128, "handle layout");

struct SlotHdr { uint32_t seq, n_reads; };

}  // namespace

struct bm_gather {
    int rank = 0, nranks = 1, root = 0;
    uint32_t max_reads = 0;
    uint64_t max_locs = 0;
    uint32_t seq = 0;                               // batches gathered so far
    cudaStream_t gstream = nullptr;                 // root: waits + pack; every rank: the copies into the root's slots
    cudaEvent_t ev_ready = nullptr, ev_copied = nullptr;
    bool copy_pending = false;
    bool host_async = false;                        // root: host copies of the packed records are not waited for in the call
    // local block: root = [kSlotSets][nranks] slots + flags; every rank = ack word (first 256 bytes of its block)
    DevBuf block;
    uint8_t* root_base = nullptr;                   // root's block as seen from this rank (peer mapping or own pointer)
    bool root_is_ipc = false;
    std::vector<uint8_t*> ack_peer;                 // root: ack word of every rank as seen from the root
    std::vector<bool> ack_is_ipc;
    HostBuf h_words;                                // pinned source words for the small copies (ring)
    uint32_t word_pos = 0;
    // root: packed result
    DevBuf r_off, r_locs, r_tot;
    uint64_t total_reads = 0;
    bool connected = false;
    // layout of the root block
    size_t off_bytes() const { return ((size_t)(max_reads + 1) * 4 + 255) & ~(size_t)255; }
    size_t locs_bytes() const { return ((size_t)max_locs * sizeof(bm_loc) + 255) & ~(size_t)255; }
    size_t slot_bytes() const { return off_bytes() + locs_bytes(); }
    size_t flags_offset() const { return (size_t)kSlotSets * nranks * slot_bytes(); }
    size_t ack_offset() const { return flags_offset() + 1024; }      // flags: kSlotSets * nranks * 8 <= 1024 bytes (nranks <= 64)
    size_t root_block_bytes() const { return ack_offset() + 256; }
    // the word the root bumps when a slot set of this rank is free again (the root keeps its own behind the flags)
    uint8_t* my_ack() const { return (uint8_t*)block.p + (rank == root ? ack_offset() : 0); }
    uint8_t* slot(uint8_t* base, int set, int r) const { return base + ((size_t)set * nranks + r) * slot_bytes(); }
    SlotHdr* flag(uint8_t* base, int set, int r) const {
        return reinterpret_cast<SlotHdr*>(base + flags_offset()) + (size_t)set * nranks + r;
    }
    uint32_t* next_word() {                         // a pinned word the copy engine reads later (ring of 4096)
        uint32_t* w = h_words.as<uint32_t>() + (word_pos & 4095u) * 2;
        word_pos++;
        return w;
    }
};

namespace {

// root: slots of one set -> contiguous (loc_off, locs) in rank order = read order
struct PackArgs {
    const uint8_t* base; size_t slot_bytes, off_bytes, flags_offset;
    int set, nranks;
    uint32_t* out_off; bm_loc* out_locs; unsigned long long* totals;   // totals[0] = reads, [1] = locations
    unsigned long long out_cap;
};

__global__ void __launch_bounds__(256)
k_gather_pack(const PackArgs a) {
    __shared__ unsigned long long s_rbase[65], s_lbase[65];
    if (threadIdx.x == 0) {
        unsigned long long rb = 0, lb = 0;
        for (int r = 0; r < a.nranks; r++) {
            const SlotHdr* h = reinterpret_cast<const SlotHdr*>(a.base + a.flags_offset) + (size_t)a.set * a.nranks + r;
            const uint32_t* off = reinterpret_cast<const uint32_t*>(a.base + ((size_t)a.set * a.nranks + r) * a.slot_bytes);
            s_rbase[r] = rb; s_lbase[r] = lb;
            rb += h->n_reads; lb += off[h->n_reads];
        }
        s_rbase[a.nranks] = rb; s_lbase[a.nranks] = lb;
        if (blockIdx.x == 0) { a.totals[0] = rb; a.totals[1] = lb; a.out_off[rb] = (uint32_t)lb; }
    }
    __syncthreads();
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (int r = 0; r < a.nranks; r++) {
        const uint8_t* sl = a.base + ((size_t)a.set * a.nranks + r) * a.slot_bytes;
        const uint32_t* off = reinterpret_cast<const uint32_t*>(sl);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(sl + a.off_bytes);       // records as 5 words each
        const unsigned long long nr = s_rbase[r + 1] - s_rbase[r], nl = s_lbase[r + 1] - s_lbase[r];
        for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < nr; i += stride)
            a.out_off[s_rbase[r] + i] = (uint32_t)(s_lbase[r] + off[i]);
        if (s_lbase[r] + nl > a.out_cap) continue;                                        // reported through totals
        uint32_t* dst = reinterpret_cast<uint32_t*>(a.out_locs + s_lbase[r]);
        for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < nl * 5ull; i += stride)
            dst[i] = src[i];
    }
}

// The gather stream waits on words other ranks write: a rank that died must not hang this one for ever.  Host waits on
// the gather stream are polled and give up after $BM_GATHER_TIMEOUT_S seconds (default 120).
int wait_stream(bm_ctx* c, cudaStream_t st, const char* what) {
    static const double limit = [] { const char* e = getenv("BM_GATHER_TIMEOUT_S"); return e ? atof(e) : 120.0; }();
    const auto t0 = std::chrono::steady_clock::now();
    for (unsigned spin = 0;; spin++) {
        const cudaError_t e = cudaStreamQuery(st);
        if (e == cudaSuccess) return BM_OK;
        if (e != cudaErrorNotReady) BM_CK(c, e);
        if (spin > 2000) std::this_thread::sleep_for(std::chrono::microseconds(50));
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > limit) {
            c->err = std::string(what) + ": timed out waiting for the other ranks' records";
            return BM_E_CUDA;
        }
    }
}

int open_peer(bm_ctx* c, const PeerHandle& h, uint8_t** out, bool* is_ipc) {
    *is_ipc = false;
    if (h.pid == (uint32_t)getpid()) {                      // several contexts in one process: plain peer access
        if (h.device != c->device) {
            int can = 0;
            BM_CK(c, cudaDeviceCanAccessPeer(&can, c->device, h.device));
            if (!can) BM_FAIL(c, BM_E_CUDA, "bm_gather_connect: no peer access between the GPUs");
            cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) BM_CK(c, e);
            cudaGetLastError();
        }
        *out = reinterpret_cast<uint8_t*>(h.ptr);
        return BM_OK;
    }
    void* p = nullptr;
    BM_CK(c, cudaIpcOpenMemHandle(&p, h.ipc, cudaIpcMemLazyEnablePeerAccess));
    *out = static_cast<uint8_t*>(p);
    *is_ipc = true;
    return BM_OK;
}

}  // namespace

extern "C" {

// ---- NCCL communicator ------------------------------------------------------------------------------------------------------
int bm_comm_unique_id(uint8_t* id) {
    if (!id) return BM_E_ARG;
    if (!nccl().ok) return BM_E_CUDA;
    ncclUniqueId u;
    if (nccl().GetUniqueId(&u) != ncclSuccess) return BM_E_CUDA;
    static_assert(sizeof(u) == BM_COMM_ID_BYTES, "ncclUniqueId size");
    memcpy(id, &u, sizeof u);
    return BM_OK;
}

int bm_comm_init(bm_ctx* c, const uint8_t* id, int nranks, int rank) {
    if (!c) return BM_E_ARG;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks) BM_FAIL(c, BM_E_ARG, "bm_comm_init: bad arguments");
    if (!nccl().ok) BM_FAIL(c, BM_E_CUDA, "bm_comm_init: libnccl.so.2 cannot be loaded");
    cudaSetDevice(c->device);
    if (c->comm && c->comm_owned) nccl().CommDestroy((ncclComm_t)c->comm);
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    ncclComm_t comm = nullptr;
    BM_NCCL(c, nccl().CommInitRank(&comm, nranks, u, rank));
    c->comm = comm; c->comm_owned = true; c->comm_rank = rank; c->comm_nranks = nranks;
    return BM_OK;
}

int bm_comm_attach(bm_ctx* c, void* nccl_comm, int nranks, int rank) {
    if (!c) return BM_E_ARG;
    if (!nccl_comm || nranks < 1 || rank < 0 || rank >= nranks) BM_FAIL(c, BM_E_ARG, "bm_comm_attach: bad arguments");
    if (!nccl().ok) BM_FAIL(c, BM_E_CUDA, "bm_comm_attach: libnccl.so.2 cannot be loaded");
    if (c->comm && c->comm_owned) nccl().CommDestroy((ncclComm_t)c->comm);
    c->comm = nccl_comm; c->comm_owned = false; c->comm_rank = rank; c->comm_nranks = nranks;
    return BM_OK;
}

int bm_comm_destroy(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    if (c->comm && c->comm_owned && nccl().ok) nccl().CommDestroy((ncclComm_t)c->comm);
    c->comm = nullptr; c->comm_owned = false; c->comm_nranks = 0;
    return BM_OK;
}

// geometry + kmer_to_index + bit matrix: root -> every rank
// K1 sharded by buckets: rank r sets the bits of its contiguous share of the buckets in a zeroed matrix,
// then the matrices are summed over NVLink.  A bit belongs to one bucket, so exactly one rank can have set it: the sum of the
// 32-bit words is their OR, no carry can occur.  Every rank ends up with the whole matrix (what bm_index_build on one rank
// followed by bm_bcast_index would give).  Collective: every rank of the communicator calls it with the same table.
int bm_index_build_sharded(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows) {
    if (!c) return BM_E_ARG;
    if (!c->comm) BM_FAIL(c, BM_E_ARG, "bm_index_build_sharded: bm_comm_init first");
    if (!nccl().AllReduce) BM_FAIL(c, BM_E_CUDA, "bm_index_build_sharded: this NCCL has no ncclAllReduce");
    cudaSetDevice(c->device);
    const uint32_t nr = (uint32_t)c->comm_nranks, rk = (uint32_t)c->comm_rank;
    const uint32_t b0 = (uint32_t)((uint64_t)c->n_emit * rk / nr), b1 = (uint32_t)((uint64_t)c->n_emit * (rk + 1) / nr);
    int r = bm_internal_index_build_range(c, k2i, q, n_rows, b0, b1);
    if (r) return r;
    ncclComm_t comm = (ncclComm_t)c->comm;
    const size_t words = ((size_t)n_rows * c->Wp + 64) / 4;
    const size_t chunk = (size_t)1 << 28;                    // 1 GB of words per call
    for (size_t o = 0; o < words; o += chunk) {
        const size_t n = words - o < chunk ? words - o : chunk;
        uint32_t* p = c->d_rows.as<uint32_t>() + o;
        BM_NCCL(c, nccl().AllReduce(p, p, n, ncclUint32, ncclSum, comm, c->stream));
    }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->have_index = true;
    return BM_OK;
}

int bm_bcast_index(bm_ctx* c, int root) {
    if (!c) return BM_E_ARG;
    if (!c->comm) BM_FAIL(c, BM_E_ARG, "bm_bcast_index: bm_comm_init first");
    if (root < 0 || root >= c->comm_nranks) BM_FAIL(c, BM_E_ARG, "bm_bcast_index: bad root");
    cudaSetDevice(c->device);
    const bool is_root = c->comm_rank == root;
    if (is_root && !c->have_index) BM_FAIL(c, BM_E_ARG, "bm_bcast_index: the root has no index resident");
    ncclComm_t comm = (ncclComm_t)c->comm;
    DevBuf hdr;
    BM_CK(c, hdr.ensure(64));
    uint64_t h[4] = {c->q, c->n_rows, c->NB, 0};
    if (is_root) BM_CK(c, cudaMemcpyAsync(hdr.p, h, sizeof h, cudaMemcpyHostToDevice, c->stream));
    BM_NCCL(c, nccl().Broadcast(hdr.p, hdr.p, sizeof h, ncclUint8, root, comm, c->stream));
    BM_CK(c, cudaMemcpyAsync(h, hdr.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    hdr.release();
    const uint32_t q = (uint32_t)h[0], n_rows = (uint32_t)h[1], NB = (uint32_t)h[2];
    if (q == 0 || q > 14 || NB == 0) BM_FAIL(c, BM_E_ARG, "bm_bcast_index: bad geometry from the root");
    const size_t n_codes = (size_t)1 << (2 * q);
    if (!is_root) {
        // the table first (it sizes everything), through a staging buffer; then the matrix into its final place
        DevBuf tab;
        BM_CK(c, tab.ensure(n_codes * 4));
        BM_NCCL(c, nccl().Broadcast(tab.p, tab.p, n_codes * 4, ncclUint8, root, comm, c->stream));
        std::vector<int32_t> k2i(n_codes);
        BM_CK(c, cudaMemcpyAsync(k2i.data(), tab.p, n_codes * 4, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaStreamSynchronize(c->stream));
        tab.release();
        if (c->d_packed.p && c->NB && c->NB != NB) BM_FAIL(c, BM_E_ARG, "bm_bcast_index: NB differs from the uploaded reference");
        int r = bm_internal_index_set_table(c, k2i.data(), q, n_rows, NB);
        if (r) return r;
    } else {
        BM_NCCL(c, nccl().Broadcast(c->d_k2i.p, c->d_k2i.p, n_codes * 4, ncclUint8, root, comm, c->stream));
    }
    const size_t bytes = (size_t)n_rows * c->Wp + 64;        // padded rows + the zero tail word block
    const size_t chunk = (size_t)1 << 30;
    for (size_t o = 0; o < bytes; o += chunk) {
        const size_t n = bytes - o < chunk ? bytes - o : chunk;
        uint8_t* p = (uint8_t*)c->d_rows.p + o;
        BM_NCCL(c, nccl().Broadcast(p, p, n, ncclUint8, root, comm, c->stream));
    }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    if (!is_root) {
        c->have_index = true;
        c->zeros_for_d = -1.f; c->krows_k = 0; c->frows_k = 0; c->fine_tried_k = 0;   // derived tables are rebuilt locally
    }
    return BM_OK;
}

// 2-bit reference + bucket table: root -> every rank
int bm_bcast_ref(bm_ctx* c, int root) {
    if (!c) return BM_E_ARG;
    if (!c->comm) BM_FAIL(c, BM_E_ARG, "bm_bcast_ref: bm_comm_init first");
    if (root < 0 || root >= c->comm_nranks) BM_FAIL(c, BM_E_ARG, "bm_bcast_ref: bad root");
    cudaSetDevice(c->device);
    const bool is_root = c->comm_rank == root;
    if (is_root && !c->d_packed.p) BM_FAIL(c, BM_E_ARG, "bm_bcast_ref: the root has no reference resident");
    ncclComm_t comm = (ncclComm_t)c->comm;
    DevBuf hdr;
    BM_CK(c, hdr.ensure(64));
    uint64_t h[4] = {c->n_bases, c->n_emit, c->NB, c->max_blen};
    if (is_root) BM_CK(c, cudaMemcpyAsync(hdr.p, h, sizeof h, cudaMemcpyHostToDevice, c->stream));
    BM_NCCL(c, nccl().Broadcast(hdr.p, hdr.p, sizeof h, ncclUint8, root, comm, c->stream));
    BM_CK(c, cudaMemcpyAsync(h, hdr.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    hdr.release();
    const uint64_t n_bases = h[0];
    const uint32_t n_emit = (uint32_t)h[1];
    const uint64_t nbytes = (n_bases + 3) / 4;
    if (!is_root) {
        c->d_gidx_pos.release(); c->d_gidx_off.release(); c->gidx_k = 0;
        BM_CK(c, c->d_packed.ensure(nbytes + 64));
        BM_CK(c, c->d_bstart.ensure((size_t)(n_emit + 1) * 8));
        BM_CK(c, c->d_blen.ensure((size_t)(n_emit + 1) * 4));
    }
    BM_NCCL(c, nccl().Broadcast(c->d_packed.p, c->d_packed.p, nbytes + 64, ncclUint8, root, comm, c->stream));
    if (n_emit) {
        BM_NCCL(c, nccl().Broadcast(c->d_bstart.p, c->d_bstart.p, (size_t)n_emit * 8, ncclUint8, root, comm, c->stream));
        BM_NCCL(c, nccl().Broadcast(c->d_blen.p, c->d_blen.p, (size_t)n_emit * 4, ncclUint8, root, comm, c->stream));
    }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    if (!is_root) { c->n_bases = n_bases; c->n_emit = n_emit; c->NB = (uint32_t)h[2]; c->max_blen = (uint32_t)h[3]; }
    return BM_OK;
}

// ---- gather ---------------------------------------------------------------------------------------------------------------------
int bm_gather_open(bm_ctx* c, int rank, int nranks, int root, uint32_t max_reads, uint64_t max_locs, uint8_t* handle) {
    if (!c) return BM_E_ARG;
    if (!handle || nranks < 1 || nranks > 64 || rank < 0 || rank >= nranks || root < 0 || root >= nranks || !max_reads || !max_locs)
        BM_FAIL(c, BM_E_ARG, "bm_gather_open: bad arguments");
    cudaSetDevice(c->device);
    bm_gather_close(c);
    if (!wait_value_fn()) BM_FAIL(c, BM_E_CUDA, "bm_gather_open: cuStreamWaitValue32 is not available on this driver");
    bm_gather* g = new bm_gather();
    c->gather = g;
    g->rank = rank; g->nranks = nranks; g->root = root; g->max_reads = max_reads; g->max_locs = max_locs;
    int lo = 0, hi = 0;
    BM_CK(c, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    BM_CK(c, cudaStreamCreateWithPriority(&g->gstream, cudaStreamNonBlocking, hi));
    BM_CK(c, cudaEventCreateWithFlags(&g->ev_ready, cudaEventDisableTiming));
    BM_CK(c, cudaEventCreateWithFlags(&g->ev_copied, cudaEventDisableTiming));
    BM_CK(c, g->h_words.ensure(4096 * 8));
    const size_t bytes = rank == root ? g->root_block_bytes() : 256;
    BM_CK(c, g->block.ensure_exact(bytes));
    if (rank == root) BM_CK(c, cudaMemset((uint8_t*)g->block.p + g->flags_offset(), 0, bytes - g->flags_offset()));
    else BM_CK(c, cudaMemset(g->block.p, 0, bytes));
    PeerHandle h;
    memset(&h, 0, sizeof h);
    h.magic = kHandleMagic; h.pid = (uint32_t)getpid(); h.device = c->device; h.rank = rank;
    h.ptr = (uint64_t)(uintptr_t)g->block.p; h.bytes = bytes; h.max_reads = max_reads; h.max_locs = max_locs;
    BM_CK(c, cudaIpcGetMemHandle(&h.ipc, g->block.p));
    memcpy(handle, &h, sizeof h);
    return BM_OK;
}

int bm_gather_connect(bm_ctx* c, const uint8_t* handles) {
    if (!c) return BM_E_ARG;
    bm_gather* g = c->gather;
    if (!g || !handles) BM_FAIL(c, BM_E_ARG, "bm_gather_connect: bm_gather_open first");
    cudaSetDevice(c->device);
    const PeerHandle* hs = reinterpret_cast<const PeerHandle*>(handles);
    for (int r = 0; r < g->nranks; r++)
        if (hs[r].magic != kHandleMagic || hs[r].rank != r || hs[r].max_reads != g->max_reads || hs[r].max_locs != g->max_locs)
            BM_FAIL(c, BM_E_ARG, "bm_gather_connect: handles are not the ranks' bm_gather_open handles in rank order");
    if (g->rank == g->root) {
        g->root_base = (uint8_t*)g->block.p;
        g->ack_peer.assign(g->nranks, nullptr);
        g->ack_is_ipc.assign(g->nranks, false);
        for (int r = 0; r < g->nranks; r++) {
            if (r == g->root) { g->ack_peer[r] = g->my_ack(); continue; }
            bool ipc = false;
            int rc = open_peer(c, hs[r], &g->ack_peer[r], &ipc);
            if (rc) return rc;
            g->ack_is_ipc[r] = ipc;
        }
        BM_CK(c, g->r_off.ensure(((size_t)g->nranks * g->max_reads + 2) * 4));
        BM_CK(c, g->r_locs.ensure(((size_t)g->nranks * g->max_locs + 1) * sizeof(bm_loc)));
        BM_CK(c, g->r_tot.ensure(64));
    } else {
        int rc = open_peer(c, hs[g->root], &g->root_base, &g->root_is_ipc);
        if (rc) return rc;
    }
    g->connected = true;
    return BM_OK;
}

// Every rank, after bm_locate_device: its records travel to the root.  Root: also waits for all ranks' batch and packs.
// Host pointers (root only) make the call synchronous: the packed records are copied out and the call returns when they
// are there; with NULL pointers nothing on the host waits (bm_gather_sync / the next call with pointers does).
int bm_gather_results(bm_ctx* c, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* n_reads_total,
                      uint64_t* n_locs_total) {
    if (!c) return BM_E_ARG;
    bm_gather* g = c->gather;
    if (!g || !g->connected) BM_FAIL(c, BM_E_ARG, "bm_gather_results: bm_gather_open / bm_gather_connect first");
    if (!c->have_locs) BM_FAIL(c, BM_E_ARG, "bm_gather_results: bm_locate_device first");
    if (c->n_reads > g->max_reads) BM_FAIL(c, BM_E_ARG, "bm_gather_results: batch larger than max_reads of bm_gather_open");
    cudaSetDevice(c->device);
    StreamWaitValue32_t wait32 = wait_value_fn();
    // records to send: the exact count if the host already knows it, else the candidate count (an upper bound that needs
    // no device->host round trip)
    uint64_t n_send = c->n_locs_known ? (c->locs_filtered ? c->n_locs_best : c->n_locs) : c->n_cand;
    if (n_send > g->max_locs) {
        uint64_t exact = 0;
        int r = bm_locate_count(c, &exact);
        if (r) return r;
        n_send = exact;
        if (n_send > g->max_locs) BM_FAIL(c, BM_E_ARG, "bm_gather_results: more locations than max_locs of bm_gather_open");
    }
    const uint32_t seq = ++g->seq;
    const int set = (int)((seq - 1) % kSlotSets);
    const void* src_off = c->locs_filtered ? c->d_loc_off2.p : c->d_loc_off.p;
    const void* src_locs = c->locs_filtered ? c->d_locs3.p : c->d_locs.p;
    // the copies run on the gather stream, behind this batch's locator and (slot reuse) behind the root's acknowledgement
    // of the batch before last
    BM_CK(c, cudaEventRecord(g->ev_ready, c->stream));
    BM_CK(c, cudaStreamWaitEvent(g->gstream, g->ev_ready, 0));
    if (seq > (uint32_t)kSlotSets) {
        if (wait32(g->gstream, (unsigned long long)(uintptr_t)g->my_ack(), seq - kSlotSets, kWaitGeq) != 0)
            BM_FAIL(c, BM_E_CUDA, "bm_gather_results: cuStreamWaitValue32 failed");
    }
    uint8_t* sl = g->slot(g->root_base, set, g->rank);
    BM_CK(c, cudaMemcpyAsync(sl, src_off, (size_t)(c->n_reads + 1) * 4, cudaMemcpyDefault, g->gstream));
    if (n_send) BM_CK(c, cudaMemcpyAsync(sl + g->off_bytes(), src_locs, (size_t)n_send * sizeof(bm_loc), cudaMemcpyDefault, g->gstream));
    uint32_t* w = g->next_word();
    w[0] = seq; w[1] = c->n_reads;
    BM_CK(c, cudaMemcpyAsync(g->flag(g->root_base, set, g->rank), w, sizeof(SlotHdr), cudaMemcpyDefault, g->gstream));
    BM_CK(c, cudaEventRecord(g->ev_copied, g->gstream));
    g->copy_pending = true;                                  // the next locator pass must not overwrite what is in flight
    if (g->rank != g->root) {
        if (n_reads_total) *n_reads_total = 0;
        if (n_locs_total) *n_locs_total = 0;
        return BM_OK;
    }
    // ---- root: all ranks' batch `seq` landed -> pack -> acknowledge ----
    for (int r = 0; r < g->nranks; r++)
        if (wait32(g->gstream, (unsigned long long)(uintptr_t)g->flag(g->root_base, set, r), seq, kWaitGeq) != 0)
            BM_FAIL(c, BM_E_CUDA, "bm_gather_results: cuStreamWaitValue32 failed");
    PackArgs a;
    a.base = g->root_base; a.slot_bytes = g->slot_bytes(); a.off_bytes = g->off_bytes(); a.flags_offset = g->flags_offset();
    a.set = set; a.nranks = g->nranks;
    a.out_off = g->r_off.as<uint32_t>(); a.out_locs = g->r_locs.as<bm_loc>(); a.totals = g->r_tot.as<unsigned long long>();
    a.out_cap = (unsigned long long)g->nranks * g->max_locs;
    {
        c->kstat[BM_K_GATHER].launches += 1;
        k_gather_pack<<<c->sm_count * 2, 256, 0, g->gstream>>>(a);
    }
    BM_CK(c, cudaGetLastError());
    for (int r = 0; r < g->nranks; r++) {
        uint32_t* aw = g->next_word();
        aw[0] = seq; aw[1] = 0;
        BM_CK(c, cudaMemcpyAsync(g->ack_peer[r], aw, 4, cudaMemcpyDefault, g->gstream));
    }
    if (!loc_off && !locs && !n_reads_total && !n_locs_total) return BM_OK;
    unsigned long long tot[2] = {0, 0};
    BM_CK(c, cudaMemcpyAsync(tot, g->r_tot.p, sizeof tot, cudaMemcpyDeviceToHost, g->gstream));
    { int rw = wait_stream(c, g->gstream, "bm_gather_results"); if (rw) return rw; }
    g->total_reads = tot[0];
    if (n_reads_total) *n_reads_total = tot[0];
    if (n_locs_total) *n_locs_total = tot[1];
    if (loc_off) BM_CK(c, cudaMemcpyAsync(loc_off, g->r_off.p, (size_t)(tot[0] + 1) * 4, cudaMemcpyDeviceToHost, g->gstream));
    if (locs && tot[1]) {
        if (tot[1] > loc_cap) BM_FAIL(c, BM_E_ARG, "bm_gather_results: location buffer too small");
        BM_CK(c, cudaMemcpyAsync(locs, g->r_locs.p, (size_t)tot[1] * sizeof(bm_loc), cudaMemcpyDeviceToHost, g->gstream));
    }
    if (!g->host_async) BM_CK(c, cudaStreamSynchronize(g->gstream));   // else: complete after bm_gather_sync / the next call
    return BM_OK;
}

// Root: with on != 0 a bm_gather_results call with host pointers returns as soon as the totals are known and leaves the
// copy of the packed records to the gather stream (it then runs under the next batch's host->device copy and mapper); the
// host arrays are complete after bm_gather_sync or the next bm_gather_results.  Default 0: the call returns with the data.
int bm_gather_host_async(bm_ctx* c, int on) {
    if (!c) return BM_E_ARG;
    if (!c->gather) BM_FAIL(c, BM_E_ARG, "bm_gather_host_async: bm_gather_open first");
    c->gather->host_async = on != 0;
    return BM_OK;
}

// everything this rank has enqueued for the gather is done (root: the last batch is packed)
int bm_gather_sync(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    bm_gather* g = c->gather;
    if (!g) BM_FAIL(c, BM_E_ARG, "bm_gather_sync: bm_gather_open first");
    cudaSetDevice(c->device);
    return wait_stream(c, g->gstream, "bm_gather_sync");
}

// makes the library stream wait for the gather stream, so that bm_timer_stop measures the gather's tail as well
int bm_gather_join(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    bm_gather* g = c->gather;
    if (!g) BM_FAIL(c, BM_E_ARG, "bm_gather_join: bm_gather_open first");
    cudaSetDevice(c->device);
    BM_CK(c, cudaEventRecord(g->ev_copied, g->gstream));
    BM_CK(c, cudaStreamWaitEvent(c->stream, g->ev_copied, 0));
    g->copy_pending = false;
    return BM_OK;
}

// root: device-resident packed result of the last gathered batch (valid after bm_gather_sync, until the next gather)
int bm_gather_device_buffers(bm_ctx* c, void** loc_off, void** locs, uint64_t* n_reads, uint64_t* n_locs) {
    if (!c) return BM_E_ARG;
    bm_gather* g = c->gather;
    if (!g || g->rank != g->root) BM_FAIL(c, BM_E_ARG, "bm_gather_device_buffers: only on the root after bm_gather_results");
    cudaSetDevice(c->device);
    unsigned long long tot[2] = {0, 0};
    BM_CK(c, cudaMemcpyAsync(tot, g->r_tot.p, sizeof tot, cudaMemcpyDeviceToHost, g->gstream));
    { int rw = wait_stream(c, g->gstream, "bm_gather_device_buffers"); if (rw) return rw; }
    if (loc_off) *loc_off = g->r_off.p;
    if (locs) *locs = g->r_locs.p;
    if (n_reads) *n_reads = tot[0];
    if (n_locs) *n_locs = tot[1];
    return BM_OK;
}

// root: host copy of the last gathered batch (waits for it); NULL pointers: only the totals
int bm_gather_download(bm_ctx* c, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* n_reads, uint64_t* n_locs) {
    void *po = nullptr, *pl = nullptr;
    uint64_t nr = 0, nl = 0;
    int r = bm_gather_device_buffers(c, &po, &pl, &nr, &nl);
    if (r) return r;
    bm_gather* g = c->gather;
    if (n_reads) *n_reads = nr;
    if (n_locs) *n_locs = nl;
    if (loc_off) BM_CK(c, cudaMemcpyAsync(loc_off, po, (size_t)(nr + 1) * 4, cudaMemcpyDeviceToHost, g->gstream));
    if (locs && nl) {
        if (nl > loc_cap) BM_FAIL(c, BM_E_ARG, "bm_gather_download: location buffer too small");
        BM_CK(c, cudaMemcpyAsync(locs, pl, (size_t)nl * sizeof(bm_loc), cudaMemcpyDeviceToHost, g->gstream));
    }
    BM_CK(c, cudaStreamSynchronize(g->gstream));
    return BM_OK;
}

int bm_gather_close(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    bm_gather* g = c->gather;
    if (!g) return BM_OK;
    cudaSetDevice(c->device);
    if (g->gstream) wait_stream(c, g->gstream, "bm_gather_close");   // gives up after the timeout: nothing is freed under a live wait
    if (g->root_is_ipc && g->root_base) cudaIpcCloseMemHandle(g->root_base);
    for (size_t r = 0; r < g->ack_peer.size(); r++)
        if (g->ack_is_ipc[r] && g->ack_peer[r]) cudaIpcCloseMemHandle(g->ack_peer[r]);
    g->block.release(); g->r_off.release(); g->r_locs.release(); g->r_tot.release(); g->h_words.release();
    if (g->ev_ready) cudaEventDestroy(g->ev_ready);
    if (g->ev_copied) cudaEventDestroy(g->ev_copied);
    if (g->gstream) cudaStreamDestroy(g->gstream);
    delete g;
    c->gather = nullptr;
    return BM_OK;
}

}  // extern "C"

// the locator's result buffers must not be overwritten while a gather copy of the previous batch is still reading them
int bm_internal_gather_fence(bm_ctx* c) {
    bm_gather* g = c->gather;
    if (!g || !g->copy_pending) return BM_OK;
    BM_CK(c, cudaStreamWaitEvent(c->stream, g->ev_copied, 0));
    g->copy_pending = false;
    return BM_OK;
}
