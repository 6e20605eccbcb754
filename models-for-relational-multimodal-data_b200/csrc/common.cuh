// Shared device helpers for the b200mp kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/b200mp.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "b200mp kernels are written for sm_100a (B200) only"
#endif

namespace b200mp {

constexpr int NT = 256;  // threads per CTA of every edge kernel

__host__ __device__ __forceinline__ int64_t min64(int64_t a, int64_t b) { return a < b ? a : b; }

// Tile geometry of the SIMT fp32 tile GEMM [TE x H] * [H x H]:
// thread (tx, ty) owns rows ty*4..ty*4+3 and columns tx*4..tx*4+3.
template <int H>
struct Tile {
    static_assert(H == 16 || H == 32 || H == 64 || H == 128, "H must be 16/32/64/128");
    static constexpr int TX = H / 4;
    static constexpr int TY = NT / TX;
    static constexpr int TE = TY * 4;     // edges per tile: 256 / 128 / 64 / 32
    static constexpr int LDX = H + 4;     // padded row stride of an activation tile (floats)
    static constexpr int XS = TE * LDX;   // floats per activation tile
    static constexpr int WS = H * H;      // floats per weight tile
    static constexpr int MT = H / 16;     // weight-gradient micro tile (MT x MT per thread)
};

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 f4(float a, float b, float c, float d) { return make_float4(a, b, c, d); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return f4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 operator*(float a, float4 b) { return f4(a * b.x, a * b.y, a * b.z, a * b.w); }
__device__ __forceinline__ float4 relu4(float4 a) { return f4(fmaxf(a.x, 0.f), fmaxf(a.y, 0.f), fmaxf(a.z, 0.f), fmaxf(a.w, 0.f)); }

// vector reduction into global memory (sm_90+: one 16-byte RED instead of four)
__device__ __forceinline__ void red_add4(float* p, float4 v) {
    atomicAdd(reinterpret_cast<float4*>(p), v);
}

// Weight tile into shared memory.  TRANSPOSE: Ws[k][c] = W[c][k] (forward, out = x W^T);
// otherwise Ws[k][c] = W[k][c] (data gradient, out = g W).
template <int H, bool TRANSPOSE>
__device__ __forceinline__ void load_weight(float* Ws, const float* __restrict__ W) {
    for (int i = threadIdx.x; i < H * H; i += NT) {
        float v = __ldg(W + i);
        if (TRANSPOSE) Ws[(i % H) * H + (i / H)] = v; else Ws[i] = v;
    }
}

// Coalesced load of TE rows x H floats (row ids row0 .. row0+nrows-1, optionally through a
// gather map) into a padded tile; rows >= nrows are zero-filled.  scale multiplies the values.
template <int H>
__device__ __forceinline__ void load_tile(float* Xs, const float* __restrict__ src, const int32_t* __restrict__ rowmap,
                                          int64_t row0, int nrows) {
    using T = Tile<H>;
    constexpr int C4 = H / 4;
    for (int i = threadIdx.x; i < T::TE * C4; i += NT) {
        int r = i / C4, c4 = i % C4;
        float4 v = f4(0.f, 0.f, 0.f, 0.f);
        if (r < nrows) {
            int64_t row = rowmap ? (int64_t)__ldg(rowmap + row0 + r) : row0 + r;
            v = ldg4(src + row * H + c4 * 4);
        }
        st4(Xs + r * T::LDX + c4 * 4, v);
    }
}

// acc[i][j] += sum_k Xs[ty*4+i][k] * Ws[k][tx*4+j]
template <int H>
__device__ __forceinline__ void tile_gemm(const float* Xs, const float* Ws, float (&acc)[4][4], int tx, int ty) {
    using T = Tile<H>;
    const float* xr = Xs + (ty * 4) * T::LDX;
    const float* wc = Ws + tx * 4;
#pragma unroll 2
    for (int k4 = 0; k4 < H; k4 += 4) {
        float4 xa[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) xa[i] = lds4(xr + i * T::LDX + k4);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            float4 w = lds4(wc + (k4 + kk) * H);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float xv = kk == 0 ? xa[i].x : kk == 1 ? xa[i].y : kk == 2 ? xa[i].z : xa[i].w;
                acc[i][0] = fmaf(xv, w.x, acc[i][0]);
                acc[i][1] = fmaf(xv, w.y, acc[i][1]);
                acc[i][2] = fmaf(xv, w.z, acc[i][2]);
                acc[i][3] = fmaf(xv, w.w, acc[i][3]);
            }
        }
    }
}

// Weight-gradient micro kernel: acc[a][b] += sum_r Gs[r][wy*MT+a] * Xs[r][wx*MT+b]
// (rows beyond the valid count must be zero in Gs).  wx = tid % 16, wy = tid / 16.
template <int H>
__device__ __forceinline__ void tile_wgrad(const float* Gs, const float* Xs, float (&acc)[Tile<H>::MT][Tile<H>::MT]) {
    using T = Tile<H>;
    constexpr int MT = T::MT;
    const int wx = threadIdx.x & 15, wy = threadIdx.x >> 4;
    const float* gp = Gs + wy * MT;
    const float* xp = Xs + wx * MT;
#pragma unroll 4
    for (int r = 0; r < T::TE; ++r) {
        float g[MT], x[MT];
        if constexpr (MT % 4 == 0) {
#pragma unroll
            for (int a = 0; a < MT; a += 4) {
                float4 gv = lds4(gp + r * T::LDX + a), xv = lds4(xp + r * T::LDX + a);
                g[a] = gv.x; g[a + 1] = gv.y; g[a + 2] = gv.z; g[a + 3] = gv.w;
                x[a] = xv.x; x[a + 1] = xv.y; x[a + 2] = xv.z; x[a + 3] = xv.w;
            }
        } else {
#pragma unroll
            for (int a = 0; a < MT; ++a) { g[a] = gp[r * T::LDX + a]; x[a] = xp[r * T::LDX + a]; }
        }
#pragma unroll
        for (int a = 0; a < MT; ++a)
#pragma unroll
            for (int b = 0; b < MT; ++b) acc[a][b] = fmaf(g[a], x[b], acc[a][b]);
    }
}

template <int H>
__device__ __forceinline__ void store_wgrad(float* part /* [H,H] of this CTA */, const float (&acc)[Tile<H>::MT][Tile<H>::MT], float scale) {
    constexpr int MT = Tile<H>::MT;
    const int wx = threadIdx.x & 15, wy = threadIdx.x >> 4;
#pragma unroll
    for (int a = 0; a < MT; ++a)
#pragma unroll
        for (int b = 0; b < MT; ++b) part[(wy * MT + a) * H + wx * MT + b] = scale * acc[a][b];
}

// ------------------------------------------------------------------ host side helpers
struct DeviceInfo { int sms; int max_smem_optin; };
const DeviceInfo& device_info();
int persistent_grid(const void* kernel, size_t smem, int64_t work_items, int threads = NT);
// H = 64 runs on the tcgen05 kernels (*_umma.cu) unless B200MP_SIMT=1 (A/B comparison only)
bool use_tensor_cores();

#define B200MP_LAUNCH_CHECK() do { cudaError_t _e = cudaGetLastError(); if (_e != cudaSuccess) return (int)_e; } while (0)
#define B200MP_CUDA(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return (int)_e; } while (0)

}  // namespace b200mp
