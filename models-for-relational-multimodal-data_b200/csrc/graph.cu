// Graph-structure kernels: CSC/CSR build (stable sort by aggregation node), degree + histogram,
// batch relabel, row permutation.  All integer work, bit-exact against the oracle.
//
// Replaces: torch_geometric `to_csc` / `sampler.edge_permutation` reached from
//   /root/reference/src/datasets/util/graph.py:31-57, ibm_transactions_for_aml.py:91,
//   `degree` + bincount at main.py:284-286 / utils.py:179-182 and the unique/relabel loops of
//   src/utils/batch_processing.py:65-72.
// The stable LSD radix sort itself is CUB's (toolkit library, like cuBLAS for plain GEMMs);
// everything around it is hand-written.
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <atomic>
#include <cstdio>
#include <cstdlib>

namespace b200mp {

const DeviceInfo& device_info() {
    static DeviceInfo info = [] {
        DeviceInfo d{148, 232448};
        int dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) {
            cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
            cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        }
        return d;
    }();
    return info;
}

bool use_tensor_cores() {
    static const bool on = [] { const char* v = getenv("B200MP_SIMT"); return !(v && v[0] == '1'); }();
    return on;
}

int persistent_grid(const void* kernel, size_t smem, int64_t work_items, int threads) {
    // Resident CTAs per SM from the kernel's own resource usage.  cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for every
    // tcgen05 kernel here, whatever its footprint (B200, CUDA 12.9: even at 48 KB of shared memory), while ncu's launch statistics
    // give "Block Limit Registers 2 / Shared Mem 2" for the 97 KB kernels and a grid of 2 x SMs does run them side by side
    // (hi_small: message forward 1.03 -> 0.87 ms, node_linear64 0.140 -> 0.113 ms).
    int per_sm = 1;
    cudaFuncAttributes fa{};
    cudaError_t oe = cudaFuncGetAttributes(&fa, kernel);
    int dev = 0, smem_sm = 233472, smem_res = 1024, regs_sm = 65536;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&smem_res, cudaDevAttrReservedSharedMemoryPerBlock, dev);
    cudaDeviceGetAttribute(&regs_sm, cudaDevAttrMaxRegistersPerMultiprocessor, dev);
    if (oe == cudaSuccess) {
        const int warps = (threads + 31) / 32;
        const int regs_warp = ((fa.numRegs * 32 + 511) / 512) * 512;                 // allocation unit: 512 registers per warp
        const size_t smem_blk = ((smem + fa.sharedSizeBytes + (size_t)smem_res + 127) / 128) * 128;
        const int by_regs = regs_warp > 0 ? regs_sm / (regs_warp * warps) : 32;
        const int by_smem = smem_blk > 0 ? (int)((size_t)smem_sm / smem_blk) : 32;
        const int by_warps = 64 / warps;
        per_sm = by_regs < by_smem ? by_regs : by_smem;
        if (by_warps < per_sm) per_sm = by_warps;
        if (per_sm > 2) per_sm = 2;             // tensor-memory columns (512 per SM) and the kernels' design point: at most two
        if (per_sm < 1) per_sm = 1;
    } else {
        (void)cudaGetLastError();
    }
    static const bool carve = [] { const char* v = getenv("B200MP_CARVEOUT"); return !(v && v[0] == '0'); }();
    if (carve && per_sm > 1)                    // two CTAs only fit with the largest shared-memory carve-out; one-CTA kernels keep their L1
        (void)cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    static const bool dbg = [] { const char* v = getenv("B200MP_DEBUG_OCC"); return v && v[0] == '1'; }();
    if (dbg)
        fprintf(stderr, "[b200mp occ] kernel %p: threads %d dyn smem %zu static smem %zu regs %d -> %d CTA/SM\n", kernel, threads, smem,
                fa.sharedSizeBytes, fa.numRegs, per_sm);
    if (const char* f = getenv("B200MP_FORCE_CTAS_PER_SM")) {      // experiment switch
        const int v = atoi(f);
        if (v >= 1 && v <= 8) per_sm = v;
    }
    int64_t g = (int64_t)device_info().sms * per_sm;
    if (work_items < g) g = work_items;
    return (int)(g < 1 ? 1 : g);
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static int bits_for(int64_t n) {   // number of key bits needed for values in [0, n)
    int b = 1;
    while (b < 63 && ((int64_t)1 << b) < n) ++b;
    return b;
}

__global__ void csc_prepare_kernel(const int64_t* __restrict__ key, int64_t E, uint32_t* __restrict__ k32, int32_t* __restrict__ iota) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < E; i += (int64_t)gridDim.x * blockDim.x) {
        k32[i] = (uint32_t)__ldg(key + i);
        iota[i] = (int32_t)i;
    }
}

__global__ void csc_finish_kernel(const uint32_t* __restrict__ ksorted, const int32_t* __restrict__ psorted,
                                  const int64_t* __restrict__ other_end, int64_t E, int64_t* __restrict__ perm,
                                  int64_t* __restrict__ other, int32_t* __restrict__ other32, int32_t* __restrict__ key32) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < E; i += (int64_t)gridDim.x * blockDim.x) {
        const int32_t p = psorted[i];
        const int64_t o = __ldg(other_end + p);
        perm[i] = p;
        other[i] = o;
        other32[i] = (int32_t)o;
        key32[i] = (int32_t)ksorted[i];
    }
}

// ptr[v] = first position whose sorted key is >= v  (v = 0..N)
__global__ void csc_ptr_kernel(const uint32_t* __restrict__ ksorted, int64_t E, int64_t N, int64_t* __restrict__ ptr) {
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v <= N; v += (int64_t)gridDim.x * blockDim.x) {
        int64_t lo = 0, hi = E;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if ((int64_t)ksorted[mid] < v) lo = mid + 1; else hi = mid;
        }
        ptr[v] = lo;
    }
}

__global__ void degree_kernel(const int64_t* __restrict__ ptr, int64_t N, int64_t* __restrict__ deg, unsigned long long* max_deg) {
    long long local = 0;
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < N; v += (int64_t)gridDim.x * blockDim.x) {
        const int64_t d = ptr[v + 1] - ptr[v];
        deg[v] = d;
        local = d > local ? d : local;
    }
    for (int off = 16; off; off >>= 1) {
        long long o = __shfl_xor_sync(0xffffffffu, local, off);
        local = o > local ? o : local;
    }
    if ((threadIdx.x & 31) == 0 && max_deg) atomicMax(max_deg, (unsigned long long)local);
}

__global__ void bincount_kernel(const int64_t* __restrict__ vals, int64_t n, unsigned long long* __restrict__ hist, int64_t nbins) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = vals[i];
        if (v >= 0 && v < nbins) atomicAdd(hist + v, 1ull);
    }
}

__global__ void unique_flag_kernel(const int64_t* __restrict__ sorted, int64_t M, int32_t* __restrict__ flag) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x)
        flag[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1 : 0;
}

__global__ void unique_scatter_kernel(const int64_t* __restrict__ sorted, const int32_t* __restrict__ flag,
                                      const int32_t* __restrict__ pos /* inclusive scan of flag */, int64_t M,
                                      int64_t* __restrict__ nodes, int64_t* __restrict__ n_unique) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x) {
        if (flag[i]) nodes[pos[i] - 1] = sorted[i];
        if (i == M - 1) *n_unique = pos[i];
    }
}

__global__ void relabel_search_kernel(const int64_t* __restrict__ ids, int64_t M, const int64_t* __restrict__ nodes,
                                      const int64_t* __restrict__ n_unique, int64_t* __restrict__ local) {
    const int64_t n = *n_unique;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = ids[i];
        int64_t lo = 0, hi = n;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (nodes[mid] < v) lo = mid + 1; else hi = mid;
        }
        local[i] = lo;
    }
}

template <bool SCATTER>
__global__ void permute_rows_kernel(const float* __restrict__ in, const int64_t* __restrict__ idx, float* __restrict__ out,
                                    int64_t n, int c4 /* H/4 */) {
    const int64_t total = n * c4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / c4;
        const int cc = (int)(i % c4);
        const int64_t j = __ldg(idx + r);
        const int64_t from = SCATTER ? r : j, to = SCATTER ? j : r;
        st4(out + (to * c4 + cc) * 4, ldg4(in + (from * c4 + cc) * 4));
    }
}

// x = hi + lo with hi = x rounded to tf32 (low 13 mantissa bits zero), lo = x - hi exactly
__global__ void split_tf32_kernel(const float4* __restrict__ x, float4* __restrict__ hi, float4* __restrict__ lo, int64_t n4) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = __ldg(x + i);
        float4 h;
        h.x = __uint_as_float((__float_as_uint(v.x) + 0x1000u) & 0xffffe000u);
        h.y = __uint_as_float((__float_as_uint(v.y) + 0x1000u) & 0xffffe000u);
        h.z = __uint_as_float((__float_as_uint(v.z) + 0x1000u) & 0xffffe000u);
        h.w = __uint_as_float((__float_as_uint(v.w) + 0x1000u) & 0xffffe000u);
        hi[i] = h;
        lo[i] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
    }
}

static inline int grid_for(int64_t n, int block = 256) {
    int64_t g = (n + block - 1) / block;
    int64_t cap = (int64_t)device_info().sms * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200mp

using namespace b200mp;

extern "C" int b200mp_version(void) { return 100; }

extern "C" int b200mp_partial_rows(int* rows) {
    if (!rows) return B200MP_EINVAL;
    *rows = device_info().sms * 2;
    return 0;
}

extern "C" int b200mp_partial_rows_for(int H, int* rows) {
    if (!rows) return B200MP_EINVAL;
    (void)H;
    *rows = device_info().sms * 2;             // tensor-core kernels: two 64-edge pipelines per CTA, one CTA per SM
    return 0;
}

extern "C" int b200mp_tile_edges(int H, int* te) {
    if (!te) return B200MP_EINVAL;
    switch (H) {
        case 16: *te = Tile<16>::TE; return 0;
        case 32: *te = Tile<32>::TE; return 0;
        case 64: *te = use_tensor_cores() ? 128 : Tile<64>::TE; return 0;
        case 128: *te = Tile<128>::TE; return 0;
    }
    return B200MP_EINVAL;
}

extern "C" int b200mp_csc_workspace_bytes(int64_t E, int64_t N, size_t* bytes) {
    if (!bytes || E < 0 || N < 0 || E >= ((int64_t)1 << 31) || N >= ((int64_t)1 << 31)) return B200MP_EINVAL;
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const int32_t*)nullptr,
                                    (int32_t*)nullptr, (int)(E > 0 ? E : 1), 0, bits_for(N > 1 ? N : 2));
    *bytes = 4 * align256(sizeof(uint32_t) * (size_t)(E > 0 ? E : 1)) + align256(tmp) + 256;
    return 0;
}

extern "C" int b200mp_csc_build(const int64_t* key, const int64_t* other_end, int64_t E, int64_t N,
                                int64_t* perm, int64_t* ptr, int64_t* other, int32_t* other32, int32_t* key32,
                                void* workspace, size_t workspace_bytes, void* stream) {
    size_t need = 0;
    int rc = b200mp_csc_workspace_bytes(E, N, &need);
    if (rc) return rc;
    if (!ptr || !workspace || (E > 0 && (!key || !other_end || !perm || !other || !other32 || !key32))) return B200MP_EINVAL;
    if (workspace_bytes < need) return B200MP_EWORKSP;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t seg = align256(sizeof(uint32_t) * (size_t)(E > 0 ? E : 1));
    char* w = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    uint32_t* k_in = (uint32_t*)w;
    uint32_t* k_out = (uint32_t*)(w + seg);
    int32_t* v_in = (int32_t*)(w + 2 * seg);
    int32_t* v_out = (int32_t*)(w + 3 * seg);
    void* tmp = w + 4 * seg;
    size_t tmp_bytes = need - 4 * seg - 256;
    if (E > 0) {
        csc_prepare_kernel<<<grid_for(E), 256, 0, st>>>(key, E, k_in, v_in);
        B200MP_LAUNCH_CHECK();
        B200MP_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, (const uint32_t*)k_in, k_out, (const int32_t*)v_in, v_out,
                                                    (int)E, 0, bits_for(N > 1 ? N : 2), st));
        csc_finish_kernel<<<grid_for(E), 256, 0, st>>>(k_out, v_out, other_end, E, perm, other, other32, key32);
        B200MP_LAUNCH_CHECK();
    }
    csc_ptr_kernel<<<grid_for(N + 1), 256, 0, st>>>(k_out, E, N, ptr);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_degree(const int64_t* ptr, int64_t N, int64_t* deg, int64_t* max_deg, void* stream) {
    if (N < 0 || !ptr || (N > 0 && !deg)) return B200MP_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (max_deg) B200MP_CUDA(cudaMemsetAsync(max_deg, 0, sizeof(int64_t), st));
    if (N == 0) return 0;
    degree_kernel<<<grid_for(N), 256, 0, st>>>(ptr, N, deg, (unsigned long long*)max_deg);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_bincount(const int64_t* vals, int64_t n, int64_t* hist, int64_t nbins, void* stream) {
    if (n < 0 || nbins <= 0 || !hist || (n > 0 && !vals)) return B200MP_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    B200MP_CUDA(cudaMemsetAsync(hist, 0, sizeof(int64_t) * (size_t)nbins, st));
    if (n == 0) return 0;
    bincount_kernel<<<grid_for(n), 256, 0, st>>>(vals, n, (unsigned long long*)hist, nbins);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_relabel_workspace_bytes(int64_t M, size_t* bytes) {
    if (!bytes || M < 0 || M >= ((int64_t)1 << 31)) return B200MP_EINVAL;
    const int m = (int)(M > 0 ? M : 1);
    size_t t1 = 0, t2 = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, t1, (const int64_t*)nullptr, (int64_t*)nullptr, m);
    cub::DeviceScan::InclusiveSum(nullptr, t2, (const int32_t*)nullptr, (int32_t*)nullptr, m);
    *bytes = align256(sizeof(int64_t) * (size_t)m) + 2 * align256(sizeof(int32_t) * (size_t)m) + align256(t1 > t2 ? t1 : t2) + 256;
    return 0;
}

extern "C" int b200mp_relabel(const int64_t* ids, int64_t M, int64_t* nodes, int64_t* n_unique, int64_t* local,
                              void* workspace, size_t workspace_bytes, void* stream) {
    size_t need = 0;
    int rc = b200mp_relabel_workspace_bytes(M, &need);
    if (rc) return rc;
    if (!n_unique || !workspace || (M > 0 && (!ids || !nodes || !local))) return B200MP_EINVAL;
    if (workspace_bytes < need) return B200MP_EWORKSP;
    cudaStream_t st = (cudaStream_t)stream;
    B200MP_CUDA(cudaMemsetAsync(n_unique, 0, sizeof(int64_t), st));
    if (M == 0) return 0;
    const int m = (int)M;
    const size_t s64 = align256(sizeof(int64_t) * (size_t)m), s32 = align256(sizeof(int32_t) * (size_t)m);
    char* w = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    int64_t* sorted = (int64_t*)w;
    int32_t* flag = (int32_t*)(w + s64);
    int32_t* pos = (int32_t*)(w + s64 + s32);
    void* tmp = w + s64 + 2 * s32;
    size_t tmp_bytes = need - (s64 + 2 * s32) - 256;
    B200MP_CUDA(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, ids, sorted, m, 0, 64, st));
    unique_flag_kernel<<<grid_for(M), 256, 0, st>>>(sorted, M, flag);
    B200MP_LAUNCH_CHECK();
    B200MP_CUDA(cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, (const int32_t*)flag, pos, m, st));
    unique_scatter_kernel<<<grid_for(M), 256, 0, st>>>(sorted, flag, pos, M, nodes, n_unique);
    B200MP_LAUNCH_CHECK();
    relabel_search_kernel<<<grid_for(M), 256, 0, st>>>(ids, M, nodes, n_unique, local);
    B200MP_LAUNCH_CHECK();
    return 0;
}

// ---- batch relabel without any allocation (b200mp_relabel_static) -----------------------------------------------------------
// Same launches as b200mp_relabel; the workspace is a file-scope device array instead of a caller buffer, for callers that cannot
// allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the
// B200MP_RELABEL_STATIC_BYTES until exit.  The array is shared by every call and keeps the previous call's contents: nothing in
// it has to start at zero (sorted / flag / pos are fully rewritten for the M ids of a call, the sort's temporary storage is
// initialised by the sort itself), and two calls that overlap on the GPU (different streams or threads) race on it.
namespace b200mp {
constexpr size_t RELABEL_STATIC_BYTES = (size_t)B200MP_RELABEL_STATIC_BYTES;
static __device__ __align__(256) unsigned char relabel_static_scratch[RELABEL_STATIC_BYTES];
constexpr int RELABEL_STATIC_MAX_DEVICES = 64;

// The array's name is a host-side stand-in in host code, not a device address: resolve it once per device (with that device
// current) and hand the kernels the resolved pointer.
static int relabel_static_address(void** out) {
    static std::atomic<void*> cache[RELABEL_STATIC_MAX_DEVICES];
    int dev = -1;
    B200MP_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= RELABEL_STATIC_MAX_DEVICES) return B200MP_EINVAL;
    void* p = cache[dev].load(std::memory_order_acquire);
    if (!p) {
        const cudaError_t e = cudaGetSymbolAddress(&p, relabel_static_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!p) return B200MP_EINVAL;
        cache[dev].store(p, std::memory_order_release);      // idempotent: a concurrent first call resolves the same address
    }
    *out = p;
    return 0;
}
}  // namespace b200mp

extern "C" int b200mp_relabel_static(const int64_t* ids, int64_t M, int64_t* nodes, int64_t* n_unique, int64_t* local, void* stream) {
    if (M < 0) return B200MP_EINVAL;
    if (M > (int64_t)B200MP_RELABEL_STATIC_MAX_IDS) return B200MP_ELIMIT;
    size_t need = 0;
    int rc = b200mp_relabel_workspace_bytes(M, &need);
    if (rc) return rc;
    if (need > RELABEL_STATIC_BYTES) return B200MP_ELIMIT;
    void* scratch = nullptr;
    rc = relabel_static_address(&scratch);
    if (rc) return rc;
    return b200mp_relabel(ids, M, nodes, n_unique, local, scratch, RELABEL_STATIC_BYTES, stream);
}

extern "C" int b200mp_gather_rows(const float* in, const int64_t* idx, float* out, int64_t n, int H, void* stream) {
    if (n < 0 || H <= 0 || (H & 3) || (n > 0 && (!in || !idx || !out))) return B200MP_EINVAL;
    if (n == 0) return 0;
    permute_rows_kernel<false><<<grid_for(n * (H / 4)), 256, 0, (cudaStream_t)stream>>>(in, idx, out, n, H / 4);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_scatter_rows(const float* in, const int64_t* idx, float* out, int64_t n, int H, void* stream) {
    if (n < 0 || H <= 0 || (H & 3) || (n > 0 && (!in || !idx || !out))) return B200MP_EINVAL;
    if (n == 0) return 0;
    permute_rows_kernel<true><<<grid_for(n * (H / 4)), 256, 0, (cudaStream_t)stream>>>(in, idx, out, n, H / 4);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_split_tf32(const float* x, float* hi, float* lo, int64_t n, void* stream) {
    if (n < 0 || (n & 3) || (n > 0 && (!x || !hi || !lo))) return B200MP_EINVAL;
    if (n == 0) return 0;
    split_tf32_kernel<<<grid_for(n / 4), 256, 0, (cudaStream_t)stream>>>((const float4*)x, (float4*)hi, (float4*)lo, n / 4);
    B200MP_LAUNCH_CHECK();
    return 0;
}
