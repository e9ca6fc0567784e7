// nccl_merge.cu -- multi-GPU merge of per-GPU result lists (no reference counterpart; the reference is
// single-GPU).  One process per GPU; start pixels are sharded by the caller (row bands via the start bounds),
// psi/phi is replicated.  Every rank contributes its sorted list (optionally only its first top_n entries --
// the global best N are always among the per-rank best N); ncclAllGather over NVLink moves the padded blocks
// (record count in each block's header), then every record computes its own position in the merged list from
// binary searches in the other ranks' sorted runs (results.cu merge_runs_kernel).  Ties keep (rank, position)
// order == (y, x, slot) order, so the merged list equals the single-GPU list.
//
// NCCL is bound at run time with dlopen so the library has no link-time dependency on it.
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"

namespace kb {
cudaStream_t handle_stream(kb_handle h);
int handle_device(kb_handle h);
kb_trajectory *handle_results(kb_handle h, uint64_t *n);
kb_trajectory *handle_results_reserve(kb_handle h, uint64_t capacity);
void handle_set_result_count(kb_handle h, uint64_t n);

typedef struct { char internal[128]; } nccl_unique_id;
typedef void *nccl_comm_t;
enum { NCCL_INT8 = 0, NCCL_UINT64 = 5 };

struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(nccl_unique_id *) = nullptr;
    int (*CommInitRank)(nccl_comm_t *, int, nccl_unique_id, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    nccl_comm_t comm = nullptr;
    int rank = 0, world = 1;
};
static NcclApi g_nccl;

// Grow-only device buffers of the merge (one communicator per process): cudaMalloc/cudaFree synchronise the
// device, which showed up in the 8-GPU step time when they ran every merge.
struct MergeBuf {
    void *p = nullptr;
    size_t cap = 0;
    void *ensure(size_t bytes) {
        if (bytes > cap) {
            if (p) cudaFree(p);
            p = nullptr;
            cap = 0;
            const size_t want = bytes + bytes / 4 + 4096;
            KB_CUDA(cudaMalloc(&p, want));
            cap = want;
        }
        return p;
    }
};
static MergeBuf g_counts, g_send, g_all;

static void nccl_load() {
    if (g_nccl.lib) return;
    const char *env = getenv("KB_NCCL_LIB");
    const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        if (!nm) continue;
        g_nccl.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) throw Error(std::string("Unable to load NCCL: ") + dlerror());
#define KB_SYM(field, name)                                                      \
    *(void **)(&g_nccl.field) = dlsym(g_nccl.lib, name);                         \
    if (!g_nccl.field) throw Error(std::string("NCCL symbol missing: ") + name);
    KB_SYM(GetUniqueId, "ncclGetUniqueId")
    KB_SYM(CommInitRank, "ncclCommInitRank")
    KB_SYM(CommDestroy, "ncclCommDestroy")
    KB_SYM(AllGather, "ncclAllGather")
    KB_SYM(GetErrorString, "ncclGetErrorString")
#undef KB_SYM
}

#define KB_NCCL(expr)                                                                                   \
    do {                                                                                                \
        int _r = (expr);                                                                                \
        if (_r != 0) throw Error(std::string("NCCL error: ") + g_nccl.GetErrorString(_r) + " (" #expr ")"); \
    } while (0)

}  // namespace kb

using namespace kb;

extern "C" {

int kb_nccl_unique_id(uint8_t id[128]) {
    KB_API_BEGIN
    nccl_load();
    nccl_unique_id u;
    KB_NCCL(g_nccl.GetUniqueId(&u));
    memcpy(id, u.internal, 128);
    KB_API_END
}

int kb_nccl_init(const uint8_t id[128], int rank, int world_size) {
    KB_API_BEGIN
    nccl_load();
    if (g_nccl.comm) throw Error("NCCL communicator already initialised.");
    nccl_unique_id u;
    memcpy(u.internal, id, 128);
    KB_NCCL(g_nccl.CommInitRank(&g_nccl.comm, world_size, u, rank));
    g_nccl.rank = rank;
    g_nccl.world = world_size;
    KB_API_END
}

int kb_nccl_finalize(void) {
    KB_API_BEGIN
    if (g_nccl.comm) {
        KB_NCCL(g_nccl.CommDestroy(g_nccl.comm));
        g_nccl.comm = nullptr;
    }
    for (MergeBuf *b : {&g_counts, &g_send, &g_all}) {
        if (b->p) cudaFree(b->p);
        b->p = nullptr;
        b->cap = 0;
    }
    KB_API_END
}

int kb_merge_results_allgather(kb_handle h, uint64_t top_n, uint64_t *n_results) {
    KB_API_BEGIN
    if (!g_nccl.comm) throw Error("kb_nccl_init has not been called.");
    KB_CUDA(cudaSetDevice(handle_device(h)));
    cudaStream_t st = handle_stream(h);
    const int world = g_nccl.world;
    uint64_t n_local = 0;
    kb_trajectory *d_local = handle_results(h, &n_local);
    const uint64_t n_send = (top_n && top_n < n_local) ? top_n : n_local;
    const size_t hdr = (size_t)merge_header_bytes();

    // Records per block.  With a caller-supplied top_n no rank sends more than that, so nothing has to be known
    // about the other ranks before the gather: the counts travel in the block headers and the host never waits.
    // Without it the largest count is exchanged first (8 bytes per rank, one host wait).
    uint64_t cap = top_n;
    uint64_t *d_counts = static_cast<uint64_t *>(g_counts.ensure(sizeof(uint64_t) * (world + 2)));
    if (!top_n) {
        KB_CUDA(cudaMemcpyAsync(d_counts + world, &n_send, sizeof(uint64_t), cudaMemcpyHostToDevice, st));
        KB_NCCL(g_nccl.AllGather(d_counts + world, d_counts, 1, NCCL_UINT64, g_nccl.comm, st));
        std::vector<uint64_t> counts(world);
        KB_CUDA(cudaMemcpyAsync(counts.data(), d_counts, sizeof(uint64_t) * world, cudaMemcpyDeviceToHost, st));
        KB_CUDA(cudaStreamSynchronize(st));
        for (uint64_t c : counts) cap = std::max(cap, c);
    }
    const size_t blk = (hdr + sizeof(kb_trajectory) * cap + 15) / 16 * 16;
    uint8_t *d_send = static_cast<uint8_t *>(g_send.ensure(blk));
    uint8_t *d_all = static_cast<uint8_t *>(g_all.ensure(blk * world));
    merge_pack_block(d_send, d_local, n_send, st);
    KB_NCCL(g_nccl.AllGather(d_send, d_all, blk, NCCL_INT8, g_nccl.comm, st));
    // every record places itself in the merged list (results.cu merge_runs_kernel); the list replaces the handle's
    const uint64_t keep_cap = top_n ? top_n : cap * (uint64_t)world;
    kb_trajectory *d_out = handle_results_reserve(h, keep_cap);   // may reallocate: the send block is packed already
    merge_sorted_runs(d_all, blk, world, cap, d_out, keep_cap, d_counts + world + 1, st);
    uint64_t total = 0;
    KB_CUDA(cudaMemcpyAsync(&total, d_counts + world + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    KB_CUDA(cudaStreamSynchronize(st));
    const uint64_t keep = (top_n && top_n < total) ? top_n : total;
    handle_set_result_count(h, keep);
    if (n_results) *n_results = keep;
    KB_API_END
}

}  // extern "C"
