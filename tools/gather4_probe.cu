// tools/gather4_probe.cu: what does cp.async.bulk.tensor.2d...tile::gather4 need from its tensor map, and where do the
// four rows land in shared memory?  Standalone (nvcc -arch=sm_100a tools/gather4_probe.cu -o /tmp/g4 -lcuda is NOT needed:
// the driver entry point is fetched through the runtime).  Usage: g4 BOX_ROWS   (1 or 4).  The kernel polls the mbarrier
// a bounded number of times, so a wrong guess reports "timeout" instead of hanging.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 2; } } while (0)

constexpr int kRowWords = 256;  // 1 KB rows of uint32

__global__ void probe(const __grid_constant__ CUtensorMap map, int r0, int r1, int r2, int r3, uint32_t expect_bytes,
                      uint32_t* out, int* status) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 4 * kRowWords * 4);
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar), dst = (uint32_t)__cvta_generic_to_shared(smem);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 4 * kRowWords; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0xDEADBEEFu;
    __syncthreads();
    asm volatile("fence.proxy.async;" ::: "memory");
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(expect_bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                     ::"r"(dst), "l"(&map), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(bar_a) : "memory");
    }
    int ok = 0;
    for (int spin = 0; spin < 2000000 && !ok; ++spin) {
        uint32_t p;
        asm volatile("{ .reg .pred P; mbarrier.try_wait.parity.shared::cta.b64 P, [%1], 0; selp.u32 %0, 1, 0, P; }" : "=r"(p) : "r"(bar_a) : "memory");
        ok = (int)p;
    }
    if (threadIdx.x == 0) *status = ok;
    __syncthreads();
    for (int i = threadIdx.x; i < 4 * kRowWords; i += blockDim.x) out[i] = reinterpret_cast<uint32_t*>(smem)[i];
}

int main(int argc, char** argv) {
    const int box_rows = argc > 1 ? atoi(argv[1]) : 1;
    const uint32_t expect = argc > 2 ? (uint32_t)atoi(argv[2]) : 4096u;
    const int n_rows = 1000;
    std::vector<uint32_t> h((size_t)n_rows * kRowWords);
    for (int r = 0; r < n_rows; ++r) for (int w = 0; w < kRowWords; ++w) h[(size_t)r * kRowWords + w] = (uint32_t)(r << 16 | w);
    uint32_t* d = nullptr; CK(cudaMalloc(&d, h.size() * 4)); CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    auto encode = reinterpret_cast<CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill)>(fn);
    CUtensorMap map;
    const cuuint64_t gdim[2] = {(cuuint64_t)kRowWords, (cuuint64_t)n_rows};
    const cuuint64_t gstr[1] = {(cuuint64_t)kRowWords * 4};
    const cuuint32_t box[2] = {(cuuint32_t)kRowWords, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("box_rows=%d expect=%u encode=%d\n", box_rows, expect, (int)cr);
    if (cr != CUDA_SUCCESS) return 1;
    uint32_t* out = nullptr; int* st = nullptr;
    CK(cudaMalloc(&out, 4 * kRowWords * 4)); CK(cudaMalloc(&st, 4)); CK(cudaMemset(st, 0xFF, 4));
    const int rows[4] = {7, 900, 3, 512};
    probe<<<1, 128, 4 * kRowWords * 4 + 64>>>(map, rows[0], rows[1], rows[2], rows[3], expect, out, st);
    const cudaError_t e = cudaDeviceSynchronize();
    printf("sync: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 3;
    std::vector<uint32_t> o(4 * kRowWords); int status = -1;
    CK(cudaMemcpy(o.data(), out, o.size() * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&status, st, 4, cudaMemcpyDeviceToHost));
    printf("mbarrier completed: %d\n", status);
    for (int s = 0; s < 4; ++s)
        printf("smem row %d: first word %08x (row %u word %u), last word %08x; want row %d\n", s, o[s * kRowWords], o[s * kRowWords] >> 16,
               o[s * kRowWords] & 0xFFFF, o[s * kRowWords + kRowWords - 1], rows[s]);
    return 0;
}
