// Stand-alone probe used while bringing up the TMA path: loads one 2-D box and writes it back.
// usage: tma_probe <f64|f32> <box0> <box1> <prefetch 0|1> <c0> <c1>
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <typename T>
__global__ void probe(const __grid_constant__ CUtensorMap map, T* out, int box0, int box1, int c0, int c1, int do_prefetch) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (do_prefetch) asm volatile("prefetch.tensormap [%0];" ::"l"(&map) : "memory");
    }
    __syncthreads();
    const uint32_t bytes = box0 * box1 * sizeof(T);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
            ::"r"(smem_u32(smem)), "l"(&map), "r"(smem_u32(&bar)), "r"(c0), "r"(c1) : "memory");
    }
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    }
    const T* s = (const T*)smem;
    for (int e = threadIdx.x; e < box0 * box1; e += blockDim.x) out[e] = s[e];
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <typename T>
int run(CUtensorMapDataType dt, int box0, int box1, int pf, int c0, int c1) {
    const int n0 = 512, n1 = 300;   // n0 innermost
    T* h = (T*)malloc(sizeof(T) * n0 * n1);
    for (int i = 0; i < n0 * n1; ++i) h[i] = (T)i;
    T *d, *o;
    cudaMalloc(&d, sizeof(T) * n0 * n1);
    cudaMalloc(&o, sizeof(T) * box0 * box1);
    cudaMemcpy(d, h, sizeof(T) * n0 * n1, cudaMemcpyHostToDevice);
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    if (!fp) { printf("no encode fn\n"); return 2; }
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)n0, (cuuint64_t)n1};
    cuuint64_t strides[1] = {(cuuint64_t)n0 * sizeof(T)};
    cuuint32_t box[2] = {(cuuint32_t)box0, (cuuint32_t)box1};
    cuuint32_t es[2] = {1, 1};
    CUresult r = ((EncodeFn)fp)(&m, dt, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    if (r) return 3;
    size_t smem = (size_t)box0 * box1 * sizeof(T);
    cudaFuncSetAttribute(probe<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    probe<T><<<1, 128, smem>>>(m, o, box0, box1, c0, c1, pf);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e) return 4;
    T* ho = (T*)malloc(smem);
    cudaMemcpy(ho, o, smem, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int j = 0; j < box1; ++j)
        for (int i = 0; i < box0; ++i) {
            int gi = c0 + i, gj = c1 + j;
            T want = (gi >= 0 && gi < n0 && gj >= 0 && gj < n1) ? (T)(gj * n0 + gi) : (T)0;
            if (ho[j * box0 + i] != want) ++bad;
        }
    printf("mismatches: %d\n", bad);
    return bad ? 5 : 0;
}

int main(int argc, char** argv) {
    if (argc < 7) return 1;
    int box0 = atoi(argv[2]), box1 = atoi(argv[3]), pf = atoi(argv[4]), c0 = atoi(argv[5]), c1 = atoi(argv[6]);
    if (!strcmp(argv[1], "f64")) return run<double>(CU_TENSOR_MAP_DATA_TYPE_FLOAT64, box0, box1, pf, c0, c1);
    return run<float>(CU_TENSOR_MAP_DATA_TYPE_FLOAT32, box0, box1, pf, c0, c1);
}
