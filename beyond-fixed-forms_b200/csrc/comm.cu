// bff_comm: the library's own NCCL communicator for the point-sharded multi-GPU pipelines (include/bff.h).
//
// The device-resident pipelines enqueue their kernels without returning to the host, so the few integer all-reduces
// between them (intersection matrices, popcounts) are enqueued from here, on the same stream, with ncclAllReduce.
// NCCL is NOT a link-time dependency: the symbols are resolved with dlopen at bff_comm_create (first the copy the
// process already holds -- PyTorch's --, then the system's), so a single-GPU user never needs NCCL at all.
// Bootstrap: rank 0 obtains a unique id (bff_comm_unique_id), the host package broadcasts its 128 bytes with
// torch.distributed, every rank calls bff_comm_create.
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"

namespace {
typedef struct { char internal[128]; } nccl_unique_id;        // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES 128)
typedef void *nccl_comm_t;
enum { kNcclInt32 = 2, kNcclUint8 = 1 };                        // ncclDataType_t
enum { kNcclSum = 0, kNcclMax = 2 };                            // ncclRedOp_t

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(nccl_unique_id *) = nullptr;
    int (*CommInitRank)(nccl_comm_t *, int, nccl_unique_id, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
};

NcclApi g_nccl;

int load_nccl() {
    if (g_nccl.handle) return BFF_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *nm : names)
        if ((h = dlopen(nm, RTLD_NOW | RTLD_NOLOAD)) != nullptr) break;      // the copy PyTorch already loaded
    if (!h) {
        const char *env = getenv("BFF_NCCL_PATH");
        if (env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
    }
    for (const char *nm : names)
        if (!h) h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        bff::set_error("bff_comm: libnccl.so.2 not found (%s)", dlerror());
        return BFF_EINVAL;
    }
#define BFF_NCCL_SYM(field, name)                                             \
    *(void **)(&g_nccl.field) = dlsym(h, name);                               \
    if (!g_nccl.field) { bff::set_error("bff_comm: %s missing in libnccl", name); return BFF_EINVAL; }
    BFF_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
    BFF_NCCL_SYM(CommInitRank, "ncclCommInitRank");
    BFF_NCCL_SYM(CommDestroy, "ncclCommDestroy");
    BFF_NCCL_SYM(AllReduce, "ncclAllReduce");
    BFF_NCCL_SYM(AllGather, "ncclAllGather");
    BFF_NCCL_SYM(Send, "ncclSend");
    BFF_NCCL_SYM(Recv, "ncclRecv");
    BFF_NCCL_SYM(GroupStart, "ncclGroupStart");
    BFF_NCCL_SYM(GroupEnd, "ncclGroupEnd");
    BFF_NCCL_SYM(GetErrorString, "ncclGetErrorString");
    BFF_NCCL_SYM(GetVersion, "ncclGetVersion");
#undef BFF_NCCL_SYM
    g_nccl.handle = h;
    return BFF_OK;
}
}  // namespace

struct bff_comm {
    nccl_comm_t comm;
    int device, rank, world;
};

#define BFF_NCCL(call)                                                                             \
    do {                                                                                           \
        const int r_ = (call);                                                                     \
        if (r_ != 0) {                                                                             \
            bff::set_error("%s failed: %s", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "?"); \
            return BFF_ECUDA;                                                                      \
        }                                                                                          \
    } while (0)

extern "C" int bff_comm_unique_id(void *id128) {
    BFF_REQUIRE(id128, "bff_comm_unique_id: null buffer");
    int rc = load_nccl();
    if (rc != BFF_OK) return rc;
    nccl_unique_id id;
    BFF_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return BFF_OK;
}

extern "C" int bff_comm_create(int device, int rank, int world, const void *id128, bff_comm **out) {
    BFF_REQUIRE(out && id128 && world >= 1 && rank >= 0 && rank < world, "bff_comm_create: bad arguments");
    int rc = load_nccl();
    if (rc != BFF_OK) return rc;
    BFF_CUDA(cudaSetDevice(device));
    nccl_unique_id id;
    memcpy(&id, id128, sizeof(id));
    nccl_comm_t comm = nullptr;
    BFF_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
    bff_comm *c = new bff_comm();
    c->comm = comm;
    c->device = device;
    c->rank = rank;
    c->world = world;
    *out = c;
    return BFF_OK;
}

extern "C" void bff_comm_destroy(bff_comm *c) {
    if (!c) return;
    if (g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    delete c;
}

extern "C" int bff_comm_rank(const bff_comm *c) { return c ? c->rank : 0; }
extern "C" int bff_comm_world(const bff_comm *c) { return c ? c->world : 1; }

namespace bff {
int comm_allreduce_i32(bff_comm *c, int32_t *buf, size_t count, int op, cudaStream_t s) {
    if (!c || c->world == 1 || count == 0) return BFF_OK;
    BFF_NCCL(g_nccl.AllReduce(buf, buf, count, kNcclInt32, op == 1 ? kNcclMax : kNcclSum, c->comm, s));
    return BFF_OK;
}
}  // namespace bff

extern "C" int bff_comm_allreduce_i32(bff_comm *c, int32_t *buf, int64_t count, int32_t op, void *stream) {
    BFF_REQUIRE(count >= 0 && (op == 0 || op == 1), "bff_comm_allreduce_i32: bad arguments");
    if (count == 0) return BFF_OK;
    BFF_REQUIRE(buf, "bff_comm_allreduce_i32: null buffer");
    return bff::comm_allreduce_i32(c, buf, (size_t)count, op, (cudaStream_t)stream);
}

extern "C" int bff_comm_allgather_bytes(bff_comm *c, const void *send, void *recv, int64_t bytes_per_rank, void *stream) {
    BFF_REQUIRE(bytes_per_rank >= 0, "bff_comm_allgather_bytes: negative size");
    if (bytes_per_rank == 0) return BFF_OK;
    BFF_REQUIRE(send && recv, "bff_comm_allgather_bytes: null buffer");
    if (!c || c->world == 1) {
        if (send != recv) BFF_CUDA(cudaMemcpyAsync(recv, send, (size_t)bytes_per_rank, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
        return BFF_OK;
    }
    BFF_NCCL(g_nccl.AllGather(send, recv, (size_t)bytes_per_rank, kNcclUint8, c->comm, (cudaStream_t)stream));
    return BFF_OK;
}

extern "C" int bff_comm_exchange_rows(bff_comm *c, const uint32_t *send, int32_t n_local_rows, const int32_t *rows_per_rank,
                                      int64_t slice_words, uint32_t *recv, void *stream_) {
    cudaStream_t s = (cudaStream_t)stream_;
    BFF_REQUIRE(n_local_rows >= 0 && slice_words >= 0 && rows_per_rank, "bff_comm_exchange_rows: bad arguments");
    const int world = c ? c->world : 1, rank = c ? c->rank : 0;
    BFF_REQUIRE(rows_per_rank[rank] == n_local_rows, "bff_comm_exchange_rows: rows_per_rank[%d] = %d, but %d local rows", rank, rows_per_rank[rank], n_local_rows);
    long long total = 0, row0_self = 0;
    for (int p = 0; p < world; ++p) {
        BFF_REQUIRE(rows_per_rank[p] >= 0, "bff_comm_exchange_rows: negative row count");
        if (p == rank) row0_self = total;
        total += rows_per_rank[p];
    }
    if (total == 0 || slice_words == 0) return BFF_OK;
    BFF_REQUIRE(recv && (send || n_local_rows == 0), "bff_comm_exchange_rows: null buffer");
    const size_t block = (size_t)n_local_rows * slice_words;          // words of one outgoing slice
    if (n_local_rows)
        BFF_CUDA(cudaMemcpyAsync(recv + row0_self * slice_words, send + (size_t)rank * block, block * 4, cudaMemcpyDeviceToDevice, s));
    if (world == 1) return BFF_OK;
    BFF_NCCL(g_nccl.GroupStart());
    long long row0 = 0;
    for (int p = 0; p < world; ++p) {
        if (p != rank) {
            if (n_local_rows) BFF_NCCL(g_nccl.Send(send + (size_t)p * block, block * 4, kNcclUint8, p, c->comm, s));
            if (rows_per_rank[p]) BFF_NCCL(g_nccl.Recv(recv + row0 * slice_words, (size_t)rows_per_rank[p] * slice_words * 4, kNcclUint8, p, c->comm, s));
        }
        row0 += rows_per_rank[p];
    }
    BFF_NCCL(g_nccl.GroupEnd());
    return BFF_OK;
}

// ---------------------------------------------------------------------------------------------------
// CUDA IPC buffers (include/bff.h)
extern "C" int bff_ipc_alloc(int64_t bytes, void **dev_ptr, void *handle64) {
    BFF_REQUIRE(bytes > 0 && dev_ptr && handle64, "bff_ipc_alloc: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    void *p = nullptr;
    BFF_CUDA(cudaMalloc(&p, (size_t)bytes));
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        cudaFree(p);
        bff::set_error("bff_ipc_alloc: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
        return BFF_ECUDA;
    }
    memcpy(handle64, &h, 64);
    *dev_ptr = p;
    return BFF_OK;
}

extern "C" int bff_ipc_open(const void *handle64, void **dev_ptr) {
    BFF_REQUIRE(handle64 && dev_ptr, "bff_ipc_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        bff::set_error("bff_ipc_open: cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
        return BFF_ECUDA;
    }
    *dev_ptr = p;
    return BFF_OK;
}

extern "C" int bff_ipc_close(void *dev_ptr) {
    if (dev_ptr) BFF_CUDA(cudaIpcCloseMemHandle(dev_ptr));
    return BFF_OK;
}

extern "C" int bff_ipc_free(void *dev_ptr) {
    if (dev_ptr) BFF_CUDA(cudaFree(dev_ptr));
    return BFF_OK;
}
