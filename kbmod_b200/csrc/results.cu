// results.cu -- device-side replacement of the host post-processing in StackSearch::search_all
// (stack_search.cpp:268-277; trajectory_list.cpp:96-126): drop lh < min_lh, drop obs_count < min_obs, sort by
// lh descending.  The reference ships all K*P records to the host and filters/sorts there; here the slots
// stay in HBM, survivors are compacted with an order-preserving scan, sorted with a stable LSD radix sort
// (hand-written, 8-bit digits, warp match-any ranking) and only then expanded to 28-byte Trajectory records.
// Ties in lh keep slot order (y, x, rank): deterministic, where the reference's parallel sort is not.
#include <float.h>

#include <algorithm>

#include "common.cuh"

namespace kb {

static constexpr int CB = 256, C_ITEMS = 4, C_TILE = CB * C_ITEMS;  // compaction tile
static constexpr int RB = 256, R_WARPS = RB / 32, R_ITEMS = 16, R_TILE = RB * R_ITEMS;  // radix tile

__device__ __forceinline__ bool slot_passes(const Slot &s, float min_lh, int min_obs) {
    return !(s.lh < min_lh) && !(s.obs < min_obs);  // remove_if(lh < min_lh), remove_if(obs < min_obs)
}

// key: ascending unsigned order == descending lh.  -0.0 and +0.0 compare equal in the reference's sort, so both map
// to the key of +0.0 and keep their input order like any other tie.
__device__ __forceinline__ uint32_t lh_key(float lh) {
    const uint32_t u = __float_as_uint(lh == 0.0f ? 0.0f : lh);
    const uint32_t ordered = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ~ordered;
}

// exclusive rank of a 0/1 flag among the CB threads of the block + block total
__device__ __forceinline__ uint32_t block_rank(bool flag, uint32_t *s_warp /*[CB/32]*/, uint32_t &total) {
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t bal = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_warp[w] = __popc(bal);
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < CB / 32; ++i) {
        const uint32_t c = s_warp[i];
        if (i < (int)w) base += c;
        tot += c;
    }
    __syncthreads();
    total = tot;
    return base + __popc(bal & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(CB) count_kernel(const Slot *__restrict__ slots, uint64_t n, float min_lh,
                                                   int min_obs, uint32_t *__restrict__ block_counts) {
    __shared__ uint32_t s_warp[CB / 32];
    const uint64_t base = (uint64_t)blockIdx.x * C_TILE;
    uint32_t cnt = 0;
#pragma unroll
    for (int j = 0; j < C_ITEMS; ++j) {
        const uint64_t i = base + (uint64_t)j * CB + threadIdx.x;
        if (i < n && slot_passes(slots[i], min_lh, min_obs)) ++cnt;
    }
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < CB / 32; ++i) t += s_warp[i];
        block_counts[blockIdx.x] = t;
    }
}

// In-place exclusive scan of `n` uint32 by one CTA; total -> *total_out (may be null).
__global__ void __launch_bounds__(1024) scan_kernel(uint32_t *data, uint64_t n, uint32_t *total_out) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint64_t base = 0; base < n; base += 1024) {
        const uint64_t i = base + threadIdx.x;
        const uint32_t v = (i < n) ? data[i] : 0u;
        uint32_t inc = v;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if ((int)lane >= d) inc += t;
        }
        if (lane == 31) s_warp[w] = inc;
        __syncthreads();
        if (w == 0) {
            uint32_t ws = s_warp[lane], winc = ws;
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, winc, d);
                if ((int)lane >= d) winc += t;
            }
            s_warp[lane] = winc - ws;  // exclusive prefix of warp sums
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t excl = carry + s_warp[w] + inc - v;
        if (i < n) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total_out) *total_out = s_carry;
}

// Large inputs (radix histograms of many-million-item sorts): three steps over tiles of 8192 words -- tile sums, an
// exclusive scan of the tile sums by the single-CTA kernel above, and a per-tile scan seeded with its tile's offset.
static constexpr int ST_ITEMS = 8, ST_TILE = 1024 * ST_ITEMS;
__global__ void __launch_bounds__(1024) scan_tile_sums_kernel(const uint32_t *__restrict__ data, uint64_t n,
                                                              uint32_t *__restrict__ tile_sums) {
    __shared__ uint32_t s_warp[32];
    const uint64_t base = (uint64_t)blockIdx.x * ST_TILE;
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < ST_ITEMS; ++j) {
        const uint64_t i = base + (uint64_t)j * 1024 + threadIdx.x;
        if (i < n) acc += data[i];
    }
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t v = s_warp[threadIdx.x];
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        if (threadIdx.x == 0) tile_sums[blockIdx.x] = v;
    }
}
__global__ void __launch_bounds__(1024) scan_tiles_kernel(uint32_t *data, uint64_t n, const uint32_t *__restrict__ tile_offsets) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = tile_offsets[blockIdx.x];
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t tile0 = (uint64_t)blockIdx.x * ST_TILE;
    for (int j = 0; j < ST_ITEMS; ++j) {
        const uint64_t i = tile0 + (uint64_t)j * 1024 + threadIdx.x;
        const uint32_t v = (i < n) ? data[i] : 0u;
        uint32_t inc = v;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if ((int)lane >= d) inc += t;
        }
        if (lane == 31) s_warp[w] = inc;
        __syncthreads();
        if (w == 0) {
            uint32_t ws = s_warp[lane], winc = ws;
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, winc, d);
                if ((int)lane >= d) winc += t;
            }
            s_warp[lane] = winc - ws;
        }
        __syncthreads();
        const uint32_t excl = s_carry + s_warp[w] + inc - v;
        if (i < n) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
}

// In-place exclusive scan of n words; `tmp` holds one word per 8192-word tile (only used for large n).
static void scan_u32(uint32_t *data, uint64_t n, uint32_t *total_out, uint32_t *tmp, uint64_t tmp_words, cudaStream_t st) {
    const uint64_t tiles = (n + ST_TILE - 1) / ST_TILE;
    if (n <= 4 * (uint64_t)ST_TILE || !tmp || tiles > tmp_words) {
        scan_kernel<<<1, 1024, 0, st>>>(data, n, total_out);
        return;
    }
    scan_tile_sums_kernel<<<(unsigned)tiles, 1024, 0, st>>>(data, n, tmp);
    scan_kernel<<<1, 1024, 0, st>>>(tmp, tiles, total_out);
    scan_tiles_kernel<<<(unsigned)tiles, 1024, 0, st>>>(data, n, tmp);
}

__global__ void __launch_bounds__(CB)
        compact_kernel(const Slot *__restrict__ slots, uint64_t n, float min_lh, int min_obs,
                       const uint32_t *__restrict__ block_offsets, uint32_t *__restrict__ keys,
                       uint32_t *__restrict__ vals) {
    __shared__ uint32_t s_warp[CB / 32];
    const uint64_t base = (uint64_t)blockIdx.x * C_TILE;
    uint32_t out = block_offsets[blockIdx.x];
#pragma unroll
    for (int j = 0; j < C_ITEMS; ++j) {
        const uint64_t i = base + (uint64_t)j * CB + threadIdx.x;
        Slot s{0.f, 0.f, 0, -1};
        bool keep = false;
        if (i < n) {
            s = slots[i];
            keep = slot_passes(s, min_lh, min_obs);
        }
        uint32_t total;
        const uint32_t r = block_rank(keep, s_warp, total);
        if (keep) {
            keys[out + r] = lh_key(s.lh);
            vals[out + r] = (uint32_t)i;
        }
        out += total;
    }
}

// ---- LSD radix sort, one 8-bit digit per pass ----------------------------------------------------------
__global__ void __launch_bounds__(RB) radix_hist_kernel(const uint32_t *__restrict__ keys, uint64_t n, int shift,
                                                        uint32_t *__restrict__ hist, uint32_t nblocks) {
    __shared__ uint32_t s_h[256];
    s_h[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * R_TILE;
#pragma unroll
    for (int j = 0; j < R_ITEMS; ++j) {
        const uint64_t i = base + (uint64_t)j * RB + threadIdx.x;
        if (i < n) atomicAdd(&s_h[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(uint64_t)threadIdx.x * nblocks + blockIdx.x] = s_h[threadIdx.x];
}

__global__ void __launch_bounds__(RB)
        radix_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in, uint64_t n,
                             int shift, const uint32_t *__restrict__ hist_scanned, uint32_t nblocks,
                             uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out) {
    __shared__ uint32_t s_cnt[R_WARPS][256];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < R_WARPS * 256; i += RB) (&s_cnt[0][0])[i] = 0;
    __syncthreads();

    const uint64_t wbase = (uint64_t)blockIdx.x * R_TILE + (uint64_t)w * (R_ITEMS * 32);
    uint32_t k[R_ITEMS], v[R_ITEMS], rk[R_ITEMS];
#pragma unroll
    for (int r = 0; r < R_ITEMS; ++r) {
        const uint64_t i = wbase + (uint64_t)r * 32 + lane;
        const bool valid = i < n;
        k[r] = valid ? keys_in[i] : 0u;
        v[r] = valid ? vals_in[i] : 0u;
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        rk[r] = 0;
        if (valid) {
            const uint32_t d = (k[r] >> shift) & 255u;
            const uint32_t peers = __match_any_sync(vmask, d);
            const int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if ((int)lane == leader) {
                old = s_cnt[w][d];
                s_cnt[w][d] = old + __popc(peers);
            }
            old = __shfl_sync(peers, old, leader);
            rk[r] = old + __popc(peers & ((1u << lane) - 1u));
        }
        __syncwarp();
    }
    __syncthreads();
    {   // thread d: turn per-warp counts of digit d into global destinations
        const uint32_t d = threadIdx.x;
        uint32_t run = hist_scanned[(uint64_t)d * nblocks + blockIdx.x];
#pragma unroll
        for (int ww = 0; ww < R_WARPS; ++ww) {
            const uint32_t c = s_cnt[ww][d];
            s_cnt[ww][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < R_ITEMS; ++r) {
        const uint64_t i = wbase + (uint64_t)r * 32 + lane;
        if (i < n) {
            const uint32_t d = (k[r] >> shift) & 255u;
            const uint32_t dst = s_cnt[w][d] + rk[r];
            keys_out[dst] = k[r];
            vals_out[dst] = v[r];
        }
    }
}

// keys/vals live in scr.keys[0]/vals[0]; after the 4 passes the sorted data is back in buffer 0.
static int radix_sort_pairs(ResultScratch &scr, uint64_t n, cudaStream_t st) {
    if (n <= 1) return 0;
    const uint32_t nblocks = (uint32_t)((n + R_TILE - 1) / R_TILE);
    int launches = 0, cur = 0;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = pass * 8;
        radix_hist_kernel<<<nblocks, RB, 0, st>>>(scr.keys[cur], n, shift, scr.hist, nblocks);
        scan_u32(scr.hist, (uint64_t)256 * nblocks, nullptr, scr.scan_tmp, scr.cap_scan_tmp, st);
        radix_scatter_kernel<<<nblocks, RB, 0, st>>>(scr.keys[cur], scr.vals[cur], n, shift, scr.hist, nblocks,
                                                     scr.keys[cur ^ 1], scr.vals[cur ^ 1]);
        cur ^= 1;
        launches += 3;
    }
    KB_CUDA(cudaGetLastError());
    return launches;
}

__device__ __forceinline__ kb_trajectory expand_slot(const Slot &s, uint64_t slot_index, int K, int sw, int x0,
                                                     int y0, const float *vx, const float *vy) {
    const uint64_t pix = slot_index / (uint64_t)K;
    kb_trajectory t;
    t.x = x0 + (int)(pix % (uint64_t)sw);
    t.y = y0 + (int)(pix / (uint64_t)sw);
    t.lh = s.lh;
    t.flux = s.flux;
    t.obs_count = s.obs;
    t.vx = (s.cand >= 0) ? vx[s.cand] : 0.0f;
    t.vy = (s.cand >= 0) ? vy[s.cand] : 0.0f;
    return t;
}

__global__ void gather_kernel(const Slot *__restrict__ slots, const uint32_t *__restrict__ order, uint64_t n, int K,
                              int sw, int x0, int y0, const float *__restrict__ vx, const float *__restrict__ vy,
                              kb_trajectory *__restrict__ out, uint32_t *invalid_flag) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t si = order ? order[i] : i;
    const kb_trajectory t = expand_slot(slots[si], si, K, sw, x0, y0, vx, vy);
    // Trajectory::is_valid (common.h:100-102): search_all throws if a survivor is not finite.
    if (invalid_flag && !(isfinite(t.lh) && isfinite(t.flux) && isfinite(t.vx) && isfinite(t.vy) && t.obs_count >= 0))
        atomicExch(invalid_flag, 1u);
    out[i] = t;
}

static void ensure_scratch(ResultScratch &s, uint64_t items, uint64_t cblocks) {
    if (!s.d_count) {
        KB_CUDA(cudaMalloc(&s.d_count, sizeof(uint32_t)));
        KB_CUDA(cudaMalloc(&s.d_flag, sizeof(uint32_t)));
    }
    if (cblocks > s.cap_blocks) {
        if (s.block_counts) cudaFree(s.block_counts);
        s.block_counts = nullptr;
        KB_CUDA(cudaMalloc(&s.block_counts, sizeof(uint32_t) * cblocks));
        s.cap_blocks = cblocks;
    }
    if (items > s.cap_items) {
        for (int i = 0; i < 2; ++i) {
            if (s.keys[i]) cudaFree(s.keys[i]);
            if (s.vals[i]) cudaFree(s.vals[i]);
            s.keys[i] = s.vals[i] = nullptr;
        }
        s.cap_items = 0;
        const uint64_t cap = items + items / 8 + 1024;
        for (int i = 0; i < 2; ++i) {
            KB_CUDA(cudaMalloc(&s.keys[i], sizeof(uint32_t) * cap));
            KB_CUDA(cudaMalloc(&s.vals[i], sizeof(uint32_t) * cap));
        }
        s.cap_items = cap;
    }
    const uint64_t need_hist = 256 * ((items + R_TILE - 1) / R_TILE + 1);
    if (need_hist > s.cap_hist) {
        if (s.hist) cudaFree(s.hist);
        s.hist = nullptr;
        KB_CUDA(cudaMalloc(&s.hist, sizeof(uint32_t) * need_hist));
        s.cap_hist = need_hist;
    }
    const uint64_t need_tmp = std::max<uint64_t>(need_hist, cblocks) / ST_TILE + 2;
    if (need_tmp > s.cap_scan_tmp) {
        if (s.scan_tmp) cudaFree(s.scan_tmp);
        s.scan_tmp = nullptr;
        KB_CUDA(cudaMalloc(&s.scan_tmp, sizeof(uint32_t) * need_tmp));
        s.cap_scan_tmp = need_tmp;
    }
}

void result_scratch_reserve(ResultScratch &s, uint64_t items) { ensure_scratch(s, items, 1); }
int radix_sort_u32_pairs(ResultScratch &scr, uint64_t n, cudaStream_t st) { return radix_sort_pairs(scr, n, st); }

void result_scratch_free(ResultScratch &s) {
    for (int i = 0; i < 2; ++i) {
        if (s.keys[i]) cudaFree(s.keys[i]);
        if (s.vals[i]) cudaFree(s.vals[i]);
    }
    if (s.block_counts) cudaFree(s.block_counts);
    if (s.hist) cudaFree(s.hist);
    if (s.scan_tmp) cudaFree(s.scan_tmp);
    if (s.d_count) cudaFree(s.d_count);
    if (s.d_flag) cudaFree(s.d_flag);
    s = ResultScratch();
}

int filter_count(const Slot *d_slots, uint64_t n_slots, float min_lh, int min_obs, ResultScratch &scr,
                 uint64_t *n_out, cudaStream_t st) {
    if (n_slots >= (1ull << 32)) throw Error("Too many result slots for 32-bit indices.");
    const uint64_t cblocks = (n_slots + C_TILE - 1) / C_TILE;
    ensure_scratch(scr, 0, cblocks);
    count_kernel<<<(unsigned)cblocks, CB, 0, st>>>(d_slots, n_slots, min_lh, min_obs, scr.block_counts);
    scan_u32(scr.block_counts, cblocks, scr.d_count, scr.scan_tmp, scr.cap_scan_tmp, st);
    uint32_t h_count = 0;
    KB_CUDA(cudaMemcpyAsync(&h_count, scr.d_count, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    KB_CUDA(cudaStreamSynchronize(st));
    KB_CUDA(cudaGetLastError());
    *n_out = h_count;
    return 2;
}

int filter_finish(const Slot *d_slots, uint64_t n_slots, uint64_t n_keep, int K, int sw, int x0, int y0, float min_lh,
                  int min_obs, const float *d_vx, const float *d_vy, ResultScratch &scr, kb_trajectory *d_out,
                  bool *invalid, cudaStream_t st) {
    *invalid = false;
    if (n_keep == 0) return 0;
    int launches = 0;
    const uint64_t cblocks = (n_slots + C_TILE - 1) / C_TILE;
    ensure_scratch(scr, n_keep, cblocks);
    compact_kernel<<<(unsigned)cblocks, CB, 0, st>>>(d_slots, n_slots, min_lh, min_obs, scr.block_counts,
                                                     scr.keys[0], scr.vals[0]);
    launches += 1;
    launches += radix_sort_pairs(scr, n_keep, st);
    KB_CUDA(cudaMemsetAsync(scr.d_flag, 0, sizeof(uint32_t), st));
    gather_kernel<<<(unsigned)((n_keep + 255) / 256), 256, 0, st>>>(d_slots, scr.vals[0], n_keep, K, sw, x0, y0,
                                                                    d_vx, d_vy, d_out, scr.d_flag);
    launches += 1;
    uint32_t h_flag = 0;
    KB_CUDA(cudaMemcpyAsync(&h_flag, scr.d_flag, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    KB_CUDA(cudaStreamSynchronize(st));
    KB_CUDA(cudaGetLastError());
    *invalid = h_flag != 0;
    return launches;
}

int slots_to_trajectories(const Slot *d_slots, uint64_t n_slots, int K, int sw, int x0, int y0, const float *d_vx,
                          const float *d_vy, kb_trajectory *d_out, cudaStream_t st) {
    if (n_slots == 0) return 0;
    gather_kernel<<<(unsigned)((n_slots + 255) / 256), 256, 0, st>>>(d_slots, nullptr, n_slots, K, sw, x0, y0, d_vx,
                                                                     d_vy, d_out, nullptr);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- sort an existing trajectory list by lh descending (multi-GPU merge) --------------------------------
__global__ void traj_keys_kernel(const kb_trajectory *__restrict__ list, uint64_t n, uint32_t *keys, uint32_t *vals) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = lh_key(list[i].lh);
    vals[i] = (uint32_t)i;
}
__global__ void traj_permute_kernel(const kb_trajectory *__restrict__ in, const uint32_t *__restrict__ order,
                                    uint64_t n, kb_trajectory *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = in[order[i]];
}

int sort_trajectories_by_lh(kb_trajectory *d_list, uint64_t n, ResultScratch &scr, kb_trajectory *d_tmp,
                            cudaStream_t st) {
    if (n <= 1) return 0;
    if (n >= (1ull << 32)) throw Error("Too many results for 32-bit indices.");
    ensure_scratch(scr, n, 1);
    const unsigned gs = (unsigned)((n + 255) / 256);
    traj_keys_kernel<<<gs, 256, 0, st>>>(d_list, n, scr.keys[0], scr.vals[0]);
    int launches = 1 + radix_sort_pairs(scr, n, st);
    traj_permute_kernel<<<gs, 256, 0, st>>>(d_list, scr.vals[0], n, d_tmp);
    KB_CUDA(cudaMemcpyAsync(d_list, d_tmp, sizeof(kb_trajectory) * n, cudaMemcpyDeviceToDevice, st));
    KB_CUDA(cudaGetLastError());
    return launches + 1;
}

// ---- merging the sorted result lists of several GPUs (nccl_merge.cu) --------------------------------------------
// `blocks` holds `world` blocks of `block_bytes` each: a 32-byte header whose first uint64 is the number of records,
// then that many 28-byte records sorted by lh descending (the order kb_search leaves).  Every record finds its place
// in the merged list by itself: its index in its own run, plus the number of records of every earlier run that do
// not rank behind it, plus the number of records of every later run that rank strictly ahead -- so equal lh values
// keep (rank, position) order, which is what a stable sort of the concatenation gives (and, with row bands in rank
// order, the order of a single-GPU search).  One kernel, no passes over the data, nothing to re-sort.
static constexpr int kMergeHeaderBytes = 32;

__device__ __forceinline__ const kb_trajectory *merge_run(const uint8_t *blocks, size_t block_bytes, int r) {
    return reinterpret_cast<const kb_trajectory *>(blocks + (size_t)r * block_bytes + kMergeHeaderBytes);
}
__device__ __forceinline__ uint64_t merge_count(const uint8_t *blocks, size_t block_bytes, int r) {
    return *reinterpret_cast<const uint64_t *>(blocks + (size_t)r * block_bytes);
}

__global__ void merge_runs_kernel(const uint8_t *__restrict__ blocks, size_t block_bytes, int world, uint64_t cap_in,
                                  kb_trajectory *__restrict__ out, uint64_t keep, uint64_t *total_out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y;
    const uint64_t n_r = min(merge_count(blocks, block_bytes, r), cap_in);
    if (i == 0 && r == 0 && total_out) {
        uint64_t tot = 0;
        for (int q = 0; q < world; ++q) tot += min(merge_count(blocks, block_bytes, q), cap_in);
        *total_out = tot;
    }
    if (i >= n_r) return;
    const kb_trajectory me = merge_run(blocks, block_bytes, r)[i];
    const uint32_t key = lh_key(me.lh);
    uint64_t pos = i;
    for (int q = 0; q < world; ++q) {
        if (q == r) continue;
        const kb_trajectory *run = merge_run(blocks, block_bytes, q);
        const uint64_t n_q = min(merge_count(blocks, block_bytes, q), cap_in);
        // earlier run: count keys <= mine; later run: count keys < mine (keys ascend along a run)
        uint64_t lo = 0, hi = n_q;
        while (lo < hi) {
            const uint64_t mid = (lo + hi) >> 1;
            const uint32_t k = lh_key(run[mid].lh);
            const bool before = (q < r) ? (k <= key) : (k < key);
            if (before) lo = mid + 1;
            else hi = mid;
        }
        pos += lo;
    }
    if (pos < keep) out[pos] = me;
}

__global__ void merge_header_kernel(uint8_t *block, uint64_t n) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *reinterpret_cast<uint64_t *>(block) = n;
}

int merge_header_bytes() { return kMergeHeaderBytes; }

int merge_pack_block(void *d_block, const kb_trajectory *d_list, uint64_t n, cudaStream_t st) {
    merge_header_kernel<<<1, 32, 0, st>>>(static_cast<uint8_t *>(d_block), n);
    if (n)
        KB_CUDA(cudaMemcpyAsync(static_cast<uint8_t *>(d_block) + kMergeHeaderBytes, d_list, sizeof(kb_trajectory) * n,
                                cudaMemcpyDeviceToDevice, st));
    KB_CUDA(cudaGetLastError());
    return 1;
}

int merge_sorted_runs(const void *d_blocks, size_t block_bytes, int world, uint64_t cap_in, kb_trajectory *d_out,
                      uint64_t keep, uint64_t *d_total, cudaStream_t st) {
    if (cap_in == 0) cap_in = 1;  // the grid still has to run once for the total
    dim3 grid((unsigned)((cap_in + 255) / 256), (unsigned)world);
    merge_runs_kernel<<<grid, 256, 0, st>>>(static_cast<const uint8_t *>(d_blocks), block_bytes, world, cap_in, d_out, keep,
                                            d_total);
    KB_CUDA(cudaGetLastError());
    return 1;
}

// ---- merging deferred sigma-G survivors into the per-pixel slots ------------------------------------------------
__global__ void surv_keys_kernel(const ClipSurvivor *__restrict__ surv, uint64_t n, uint32_t *keys, uint32_t *vals) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = surv[i].pixel;
    vals[i] = (uint32_t)i;
}

// One thread per run of equal pixel keys (sorted): insert the run's survivors into that pixel's K slots, which are
// kept in descending (lh, then ascending candidate index) order -- the order the tiled kernel uses.
__global__ void surv_merge_kernel(const ClipSurvivor *__restrict__ surv, const uint32_t *__restrict__ keys,
                                  const uint32_t *__restrict__ order, uint64_t n, Slot *slots, int K) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t pix = keys[i];
    if (i > 0 && keys[i - 1] == pix) return;  // not the first of its run
    Slot *my = slots + (size_t)pix * K;
    for (uint64_t j = i; j < n && keys[j] == pix; ++j) {
        const ClipSurvivor sv = surv[order[j]];
        Slot cur{sv.lh, sv.flux, sv.obs, sv.cand};
        for (int r = 0; r < K; ++r) {
            const Slot s = my[r];
            if (cur.lh > s.lh || (cur.lh == s.lh && cur.cand < s.cand)) {
                my[r] = cur;
                cur = s;
            }
        }
    }
}

int merge_clip_survivors(const ClipSurvivor *d_surv, uint64_t n, Slot *d_slots, int K, ResultScratch &scr,
                         cudaStream_t st) {
    if (n == 0) return 0;
    if (n >= (1ull << 32)) throw Error("Too many sigma-G survivors for 32-bit indices.");
    ensure_scratch(scr, n, 1);
    const unsigned gs = (unsigned)((n + 255) / 256);
    surv_keys_kernel<<<gs, 256, 0, st>>>(d_surv, n, scr.keys[0], scr.vals[0]);
    int launches = 1 + radix_sort_pairs(scr, n, st);
    surv_merge_kernel<<<gs, 256, 0, st>>>(d_surv, scr.keys[0], scr.vals[0], n, d_slots, K);
    KB_CUDA(cudaGetLastError());
    return launches + 1;
}

void clip_scratch_free(ClipScratch &c) {
    if (c.queue) cudaFree(c.queue);
    if (c.bar) cudaFree(c.bar);
    if (c.surv) cudaFree(c.surv);
    if (c.counters) cudaFree(c.counters);
    c = ClipScratch();
}

// ---- near-duplicate grid filter (filters/clustering_grid.py:13-176) -------------------------------
// The reference hashes every result into a spatial bin (start cell, end cell after max_time) in a Python dict and
// keeps the best lh per bin (strict '>': the earlier result wins a tie); bins come out in the order they were first
// seen.  Here: stable sort by lh descending, then four stable sorts by the bin coordinates, so every bin is one run
// whose head is its winner; the bins are then ordered by the smallest original index they contain.
__global__ void grid_bins_kernel(const kb_trajectory *__restrict__ list, uint64_t n, double bin_width, double max_time,
                                 int4 *__restrict__ bins, uint32_t *keys, uint32_t *vals) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const kb_trajectory t = list[i];
    // add_trajectory (clustering_grid.py:72-75): python doubles.  add_trajectory_list (:109-113) gets vx/vy from
    // extract_all_trajectory_vx, a std::vector<float> that pybind11 hands over as a list of Python floats, so numpy
    // makes it float64: the product max_time * vx is a double product there too.
    const double ex = __dmul_rn(max_time, (double)t.vx);
    const double ey = __dmul_rn(max_time, (double)t.vy);
    int4 b;
    b.x = (int)__ddiv_rn((double)t.x, bin_width);                       // int(): truncation toward zero
    b.y = (int)__ddiv_rn((double)t.y, bin_width);
    b.z = (int)__ddiv_rn(__dadd_rn((double)t.x, ex), bin_width);
    b.w = (int)__ddiv_rn(__dadd_rn((double)t.y, ey), bin_width);
    bins[i] = b;
    keys[i] = lh_key(t.lh);
    vals[i] = (uint32_t)i;
}
__global__ void grid_comp_kernel(const int4 *__restrict__ bins, const uint32_t *__restrict__ order, int comp, uint64_t n,
                                 uint32_t *__restrict__ keys) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = reinterpret_cast<const uint32_t *>(bins + order[i])[comp];
}
__device__ __forceinline__ bool grid_is_head(const int4 *bins, const uint32_t *order, uint64_t i) {
    if (i == 0) return true;
    const int4 a = bins[order[i]], b = bins[order[i - 1]];
    return a.x != b.x || a.y != b.y || a.z != b.z || a.w != b.w;
}
__global__ void grid_flag_kernel(const int4 *__restrict__ bins, const uint32_t *__restrict__ order, uint64_t n,
                                 uint32_t *__restrict__ flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = grid_is_head(bins, order, i) ? 1u : 0u;
}
__global__ void grid_segment_kernel(const int4 *__restrict__ bins, const uint32_t *__restrict__ order,
                                    const uint32_t *__restrict__ excl, uint64_t n, uint32_t *seg_first,
                                    uint32_t *seg_best) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool head = grid_is_head(bins, order, i);
    const uint32_t seg = excl[i] + (head ? 1u : 0u) - 1u;
    atomicMin(seg_first + seg, order[i]);
    if (head) seg_best[seg] = order[i];
}

int grid_filter_indices(const kb_trajectory *d_list, uint64_t n, double bin_width, double max_time,
                        uint32_t *d_idx_out, uint64_t *n_out, ResultScratch &scr, cudaStream_t st) {
    *n_out = 0;
    if (n == 0) return 0;
    if (n >= (1ull << 31)) throw Error("Too many trajectories for the grid filter.");
    ensure_scratch(scr, n, 1);
    int4 *bins = nullptr;
    uint32_t *flags = nullptr, *total = nullptr;
    KB_CUDA(cudaMalloc(&bins, sizeof(int4) * n));
    KB_CUDA(cudaMalloc(&flags, sizeof(uint32_t) * n));
    KB_CUDA(cudaMalloc(&total, sizeof(uint32_t)));
    int launches = 0;
    try {
        const unsigned gs = (unsigned)((n + 255) / 256);
        grid_bins_kernel<<<gs, 256, 0, st>>>(d_list, n, bin_width, max_time, bins, scr.keys[0], scr.vals[0]);
        launches += 1 + radix_sort_pairs(scr, n, st);
        for (int comp = 3; comp >= 0; --comp) {
            grid_comp_kernel<<<gs, 256, 0, st>>>(bins, scr.vals[0], comp, n, scr.keys[0]);
            launches += 1 + radix_sort_pairs(scr, n, st);
        }
        grid_flag_kernel<<<gs, 256, 0, st>>>(bins, scr.vals[0], n, flags);
        scan_kernel<<<1, 1024, 0, st>>>(flags, n, total);
        uint32_t n_seg = 0;
        KB_CUDA(cudaMemcpyAsync(&n_seg, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        KB_CUDA(cudaStreamSynchronize(st));
        // keys[1] / vals[1] are free between sorts: first-seen index and winner of every bin
        KB_CUDA(cudaMemsetAsync(scr.keys[1], 0xff, sizeof(uint32_t) * n_seg, st));
        grid_segment_kernel<<<gs, 256, 0, st>>>(bins, scr.vals[0], flags, n, scr.keys[1], scr.vals[1]);
        KB_CUDA(cudaMemcpyAsync(scr.keys[0], scr.keys[1], sizeof(uint32_t) * n_seg, cudaMemcpyDeviceToDevice, st));
        KB_CUDA(cudaMemcpyAsync(scr.vals[0], scr.vals[1], sizeof(uint32_t) * n_seg, cudaMemcpyDeviceToDevice, st));
        launches += 3 + radix_sort_pairs(scr, n_seg, st);
        KB_CUDA(cudaMemcpyAsync(d_idx_out, scr.vals[0], sizeof(uint32_t) * n_seg, cudaMemcpyDeviceToDevice, st));
        KB_CUDA(cudaStreamSynchronize(st));
        KB_CUDA(cudaGetLastError());
        *n_out = n_seg;
    } catch (...) {
        cudaFree(bins);
        cudaFree(flags);
        cudaFree(total);
        throw;
    }
    cudaFree(bins);
    cudaFree(flags);
    cudaFree(total);
    return launches;
}

// ---- sorting a trajectory list without allocating anything ----------------------------------------------------
// The scratch the radix sort needs (two key buffers, two index buffers, the digit histograms and a staging copy of
// the list) lives in one file-scope device array instead of cudaMalloc'd buffers.  Every GPU that loads this
// library holds the array (about 12 MB) until the process exits.
static constexpr uint64_t kStaticSortMax = KB_STATIC_SORT_MAX;
static constexpr uint64_t kStaticHistWords = 256 * ((kStaticSortMax + R_TILE - 1) / R_TILE + 1);
static constexpr uint64_t kStaticScratchBytes =
        4 * sizeof(uint32_t) * kStaticSortMax + sizeof(uint32_t) * kStaticHistWords + sizeof(kb_trajectory) * kStaticSortMax;
static __device__ __align__(256) unsigned char g_static_sort_scratch[kStaticScratchBytes];

// In host code `g_static_sort_scratch` is only a stand-in for the symbol, not a device address: the address each
// device gave the array is looked up once per device (with that device current) and cached here.
static unsigned char *g_static_scratch_addr[64] = {nullptr};

static unsigned char *static_scratch_for_current_device() {
    int dev = -1;
    KB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw Error("device ordinal out of range for the static sort scratch");
    if (!g_static_scratch_addr[dev]) {
        void *ptr = nullptr;
        KB_CUDA(cudaGetSymbolAddress(&ptr, g_static_sort_scratch));
        if (!ptr) throw Error("cudaGetSymbolAddress returned a null address for the static sort scratch");
        g_static_scratch_addr[dev] = static_cast<unsigned char *>(ptr);
    }
    return g_static_scratch_addr[dev];
}

int sort_trajectories_by_lh_static(kb_trajectory *d_list, uint64_t n, cudaStream_t st) {
    if (n > kStaticSortMax)
        throw Error("kb_sort_trajectories_static: " + std::to_string(n) + " trajectories exceed the fixed scratch (" +
                    std::to_string(kStaticSortMax) + " at most); use the allocating path for longer lists");
    if (n <= 1) return 0;
    unsigned char *base = static_scratch_for_current_device();
    ResultScratch scr;  // a view into the static array; never freed
    uint32_t *w = reinterpret_cast<uint32_t *>(base);
    scr.keys[0] = w;
    scr.keys[1] = w + kStaticSortMax;
    scr.vals[0] = w + 2 * kStaticSortMax;
    scr.vals[1] = w + 3 * kStaticSortMax;
    scr.hist = w + 4 * kStaticSortMax;
    kb_trajectory *tmp = reinterpret_cast<kb_trajectory *>(w + 4 * kStaticSortMax + kStaticHistWords);
    // Nothing in the scratch has to start at zero: the key/index buffers are written before they are read, every
    // histogram word a pass reads is rewritten by that pass's histogram kernel first, and the staging copy is
    // fully overwritten by the permutation.  (The per-CTA digit counters live in shared memory and are cleared
    // inside the kernel.)
    const unsigned gs = (unsigned)((n + 255) / 256);
    traj_keys_kernel<<<gs, 256, 0, st>>>(d_list, n, scr.keys[0], scr.vals[0]);
    int launches = 1 + radix_sort_pairs(scr, n, st);
    traj_permute_kernel<<<gs, 256, 0, st>>>(d_list, scr.vals[0], n, tmp);
    KB_CUDA(cudaMemcpyAsync(d_list, tmp, sizeof(kb_trajectory) * n, cudaMemcpyDeviceToDevice, st));
    KB_CUDA(cudaGetLastError());
    return launches + 1;
}

}  // namespace kb

extern "C" int kb_sort_trajectories_static(kb_trajectory *d_list, uint64_t n, void *cuda_stream) {
    KB_API_BEGIN
    if (!d_list && n) throw kb::Error("kb_sort_trajectories_static: null list");
    kb::sort_trajectories_by_lh_static(d_list, n, static_cast<cudaStream_t>(cuda_stream));
    KB_API_END
}
