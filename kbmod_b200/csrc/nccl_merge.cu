// nccl_merge.cu -- multi-GPU merge of per-GPU result lists (no reference counterpart; the reference is
// single-GPU).  One process per GPU; start pixels are sharded by the caller (row bands via the start bounds),
// psi/phi is replicated.  Every rank contributes its sorted list (optionally only its first top_n entries --
// the global best N are always among the per-rank best N); ncclAllGather over NVLink moves the padded blocks,
// then every rank compacts and re-sorts with the same stable radix sort as the single-GPU path, so ties keep
// (rank, position) order == (y, x, slot) order and the merged list equals the single-GPU list.
//
// NCCL is bound at run time with dlopen so the library has no link-time dependency on it.
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"

namespace kb {
cudaStream_t handle_stream(kb_handle h);
int handle_device(kb_handle h);
kb_trajectory *handle_results(kb_handle h, uint64_t *n);
void handle_replace_results(kb_handle h, const kb_trajectory *d_src, uint64_t n, bool sort);

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
static MergeBuf g_counts, g_send, g_all, g_cat;

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
    for (MergeBuf *b : {&g_counts, &g_send, &g_all, &g_cat}) {
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
    uint64_t n_send = (top_n && top_n < n_local) ? top_n : n_local;

    // 1. counts
    uint64_t *d_counts = static_cast<uint64_t *>(g_counts.ensure(sizeof(uint64_t) * (world + 1)));
    KB_CUDA(cudaMemcpyAsync(d_counts + world, &n_send, sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    KB_NCCL(g_nccl.AllGather(d_counts + world, d_counts, 1, NCCL_UINT64, g_nccl.comm, st));
    std::vector<uint64_t> counts(world);
    KB_CUDA(cudaMemcpyAsync(counts.data(), d_counts, sizeof(uint64_t) * world, cudaMemcpyDeviceToHost, st));
    KB_CUDA(cudaStreamSynchronize(st));
    uint64_t max_c = 0, total = 0;
    for (uint64_t c : counts) { max_c = std::max(max_c, c); total += c; }

    if (total == 0) {
        handle_replace_results(h, d_local, 0, false);
        if (n_results) *n_results = 0;
        return 0;
    }
    // 2. padded blocks
    const size_t blk = sizeof(kb_trajectory) * max_c;
    kb_trajectory *d_send = static_cast<kb_trajectory *>(g_send.ensure(blk));
    kb_trajectory *d_all = static_cast<kb_trajectory *>(g_all.ensure(blk * world));
    kb_trajectory *d_cat = static_cast<kb_trajectory *>(g_cat.ensure(sizeof(kb_trajectory) * total));
    KB_CUDA(cudaMemsetAsync(d_send, 0, blk, st));
    if (n_send) KB_CUDA(cudaMemcpyAsync(d_send, d_local, sizeof(kb_trajectory) * n_send, cudaMemcpyDeviceToDevice, st));
    KB_NCCL(g_nccl.AllGather(d_send, d_all, blk, NCCL_INT8, g_nccl.comm, st));
    // 3. concatenate the valid prefixes in rank order, then one stable sort
    uint64_t off = 0;
    for (int r = 0; r < world; ++r) {
        if (counts[r])
            KB_CUDA(cudaMemcpyAsync(d_cat + off, d_all + (size_t)r * max_c, sizeof(kb_trajectory) * counts[r],
                                    cudaMemcpyDeviceToDevice, st));
        off += counts[r];
    }
    handle_replace_results(h, d_cat, total, true);
    uint64_t keep = (top_n && top_n < total) ? top_n : total;
    handle_replace_results(h, handle_results(h, &n_local), keep, false);
    if (n_results) *n_results = keep;
    KB_API_END
}

}  // extern "C"
