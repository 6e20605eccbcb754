// Batch construction on the device.
//
// b200mp_negative_sample replaces the reference's only native component,
// /root/reference/src/primitives/negative_sampling/negative_sampling.cpp:10-76 (called from
// src/utils/batch_processing.py:145 on Python lists): per positive edge (src, dst), num_neg / 2
// corrupted-destination edges (src, n) followed by num_neg / 2 corrupted-source edges (n, dst), n drawn
// uniformly from [0, |nodeset|) and rejected while n is src, dst or an (undirected) neighbour of src or
// dst in the batch graph.  The reference seeds std::mt19937 from std::random_device (not reproducible),
// so parity is defined on the constraints and the distribution, not the stream; here the draw is a
// counter-based hash of (seed, output slot, attempt) -- reproducible, order independent.
//
// Device algorithm: the 2E directed pairs (u, v), (v, u) are packed into 64-bit keys and radix sorted
// once (CUB); "is v adjacent to u" is a binary search for the key; |nodeset| = number of distinct u
// (every node of the batch graph appears as a first component).  One thread per output slot.
#include "common.cuh"
#include <cub/cub.cuh>

namespace b200mp {

__device__ __forceinline__ uint64_t mix64(uint64_t z) {      // splitmix64 finaliser
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

__global__ void pair_keys_kernel(const int64_t* __restrict__ ei, int64_t E, uint64_t* __restrict__ keys) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= E) return;
    const uint64_t u = (uint64_t)ei[i], v = (uint64_t)ei[E + i];
    keys[2 * i] = (u << 32) | (v & 0xffffffffull);
    keys[2 * i + 1] = (v << 32) | (u & 0xffffffffull);
}

__global__ void count_nodes_kernel(const uint64_t* __restrict__ keys, int64_t M, unsigned long long* __restrict__ count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    if (i == 0 || (keys[i] >> 32) != (keys[i - 1] >> 32)) atomicAdd(count, 1ull);
}

__device__ __forceinline__ bool has_key(const uint64_t* __restrict__ keys, int64_t M, uint64_t key) {
    int64_t lo = 0, hi = M;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo < M && keys[lo] == key;
}

__global__ void negative_sample_kernel(const uint64_t* __restrict__ keys, int64_t M, const int64_t* __restrict__ pos, int64_t B,
                                       int half, int64_t num_nodes_arg, const unsigned long long* __restrict__ count,
                                       uint64_t seed, int64_t* __restrict__ neg) {
    const int64_t per = 2 * (int64_t)half, total = B * per;
    const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= total) return;
    const int64_t i = slot / per;
    const int jj = (int)(slot % per);
    const uint64_t n = num_nodes_arg > 0 ? (uint64_t)num_nodes_arg : (uint64_t)*count;
    const uint64_t src = (uint64_t)pos[i], dst = (uint64_t)pos[B + i];
    int64_t cand = -1;
    if (n > 0) {
        for (uint32_t attempt = 0; attempt < (1u << 20); ++attempt) {
            const uint64_t r = mix64(mix64(seed ^ (uint64_t)slot * 0xd1342543de82ef95ull) + attempt);
            const uint64_t c = __umul64hi(r, n);                  // uniform on [0, n) up to n / 2^64
            if (c == src || c == dst) continue;
            if (has_key(keys, M, (src << 32) | c) || has_key(keys, M, (dst << 32) | c)) continue;
            cand = (int64_t)c;
            break;
        }
    }
    neg[slot] = jj < half ? (int64_t)src : cand;                  // row 0
    neg[total + slot] = jj < half ? cand : (int64_t)dst;          // row 1
}

// ------------------------------------------------------------------ CSC neighbour sampler (row a3)
// One hop of PyG NeighborSampler.sample_from_edges -> pyg-lib neighbor_sample (non-disjoint, no
// replacement) as reached from /root/reference/src/datasets/ibm_transactions_for_aml.py:88-102:
// frontier nodes in order, each one's CSC slice (in-edges) whole if deg <= fanout, else `fanout`
// distinct positions (Floyd's algorithm on a hash stream; pyg-lib's own RNG is not reproducible, so
// only the take-all regime is bit-exact: SURVEY.md App. A.3).
__global__ void hop_count_kernel(const int64_t* __restrict__ colptr, const int64_t* __restrict__ frontier, int64_t F,
                                 int64_t fanout, int64_t* __restrict__ take) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= F) return;
    const int64_t v = frontier[i], deg = colptr[v + 1] - colptr[v];
    take[i] = (fanout >= 0 && deg > fanout) ? fanout : deg;
}

// one warp per frontier node: take-all slices are written cooperatively; a hub runs Floyd on lane 0
__global__ void hop_fill_kernel(const int64_t* __restrict__ colptr, const int64_t* __restrict__ frontier, int64_t F,
                                int64_t fanout, const int64_t* __restrict__ offs, uint64_t seed, int64_t* __restrict__ out_pos) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= F) return;
    const int64_t v = frontier[i], lo = colptr[v], deg = colptr[v + 1] - lo;
    int64_t* out = out_pos + offs[i];
    if (fanout < 0 || deg <= fanout) {
        for (int64_t j = lane; j < deg; j += 32) out[j] = lo + j;
        return;
    }
    if (lane != 0) return;
    int64_t cnt = 0;
    for (int64_t j = deg - fanout; j < deg; ++j) {           // Floyd: fanout distinct values of [0, deg)
        const uint64_t r = mix64(mix64(seed ^ (uint64_t)v * 0x9e3779b97f4a7c15ull) + (uint64_t)j);
        int64_t t = (int64_t)__umul64hi(r, (uint64_t)(j + 1));
        bool dup = false;
        for (int64_t k = 0; k < cnt; ++k) dup |= (out[k] == lo + t);
        out[cnt++] = lo + (dup ? j : t);
    }
}

// first[node] = smallest output position k holding an unvisited `node`
__global__ void first_pos_kernel(const int64_t* __restrict__ cand, int64_t K, const uint8_t* __restrict__ visited,
                                 unsigned long long* __restrict__ first) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const int64_t u = cand[k];
    if (!visited[u]) atomicMin(first + u, (unsigned long long)k);
}
// flags[k] = 1 iff k is that first position; then the node is visited and its scratch entry reset
__global__ void first_flag_kernel(const int64_t* __restrict__ cand, int64_t K, const uint8_t* __restrict__ visited,
                                  const unsigned long long* __restrict__ first, uint8_t* __restrict__ flags) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const int64_t u = cand[k];
    flags[k] = (!visited[u] && first[u] == (unsigned long long)k) ? 1 : 0;
}
__global__ void first_commit_kernel(const int64_t* __restrict__ cand, int64_t K, uint8_t* __restrict__ visited,
                                    unsigned long long* __restrict__ first) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const int64_t u = cand[k];
    visited[u] = 1;
    first[u] = ~0ull;
}

}  // namespace b200mp

extern "C" int b200mp_negative_sample_workspace_bytes(int64_t E, size_t* bytes) {
    if (!bytes || E < 0) return B200MP_EINVAL;
    size_t tmp = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, tmp, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)(2 * E > 0 ? 2 * E : 1));
    *bytes = tmp + 2 * (size_t)(2 * E + 1) * sizeof(uint64_t) + 1024;
    return 0;
}

extern "C" int b200mp_negative_sample(const int64_t* edge_index, int64_t E, int64_t num_nodes, const int64_t* pos, int64_t B,
                                      int num_neg, uint64_t seed, int64_t* neg, void* workspace, size_t workspace_bytes,
                                      void* stream) {
    using namespace b200mp;
    if (num_neg <= 0 || E < 0 || B < 0 || 2 * E >= (int64_t)1 << 31) return B200MP_EINVAL;   // reference: std::invalid_argument
    const int half = num_neg / 2;
    if (B == 0 || half == 0) return 0;
    size_t need = 0;
    b200mp_negative_sample_workspace_bytes(E, &need);
    if (!pos || !neg || !workspace || (E > 0 && !edge_index)) return B200MP_EINVAL;
    if (workspace_bytes < need) return B200MP_EWORKSP;
    cudaStream_t st = (cudaStream_t)stream;
    char* w = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    const int64_t M = 2 * E;
    uint64_t* keys_in = (uint64_t*)w;
    uint64_t* keys = keys_in + (M + 1);
    unsigned long long* count = (unsigned long long*)(keys + M);
    void* tmp = (void*)(((uintptr_t)(count + 1) + 255) & ~(uintptr_t)255);
    size_t tmp_bytes = workspace_bytes - (size_t)((char*)tmp - (char*)workspace);
    B200MP_CUDA(cudaMemsetAsync(count, 0, sizeof(unsigned long long), st));
    if (E > 0) {
        pair_keys_kernel<<<(unsigned)((E + 255) / 256), 256, 0, st>>>(edge_index, E, keys_in);
        B200MP_LAUNCH_CHECK();
        B200MP_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, (const uint64_t*)keys_in, keys, (int)M, 0, 64, st));
        count_nodes_kernel<<<(unsigned)((M + 255) / 256), 256, 0, st>>>(keys, M, count);
        B200MP_LAUNCH_CHECK();
    }
    const int64_t total = B * 2 * (int64_t)half;
    negative_sample_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(keys, M, pos, B, half, num_nodes, count, seed, neg);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_sample_hop_count(const int64_t* colptr, const int64_t* frontier, int64_t F, int64_t fanout, int64_t* take,
                                       void* stream) {
    using namespace b200mp;
    if (F < 0 || (F > 0 && (!colptr || !frontier || !take))) return B200MP_EINVAL;
    if (F == 0) return 0;
    hop_count_kernel<<<(unsigned)((F + 255) / 256), 256, 0, (cudaStream_t)stream>>>(colptr, frontier, F, fanout, take);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_sample_hop_fill(const int64_t* colptr, const int64_t* frontier, int64_t F, int64_t fanout,
                                      const int64_t* offs, uint64_t seed, int64_t* out_pos, void* stream) {
    using namespace b200mp;
    if (F < 0 || (F > 0 && (!colptr || !frontier || !offs || !out_pos))) return B200MP_EINVAL;
    if (F == 0) return 0;
    hop_fill_kernel<<<(unsigned)((F * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(colptr, frontier, F, fanout, offs, seed, out_pos);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_first_occurrence(const int64_t* cand, int64_t K, uint8_t* visited, uint64_t* first, uint8_t* flags,
                                       void* stream) {
    using namespace b200mp;
    if (K < 0 || (K > 0 && (!cand || !visited || !first || !flags))) return B200MP_EINVAL;
    if (K == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((K + 255) / 256);
    first_pos_kernel<<<grid, 256, 0, st>>>(cand, K, visited, (unsigned long long*)first);
    first_flag_kernel<<<grid, 256, 0, st>>>(cand, K, visited, (const unsigned long long*)first, flags);
    first_commit_kernel<<<grid, 256, 0, st>>>(cand, K, visited, (unsigned long long*)first);
    B200MP_LAUNCH_CHECK();
    return 0;
}
