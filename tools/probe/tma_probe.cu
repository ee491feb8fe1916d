// Developer probe: one warp per CTA loads a (bw x bh) u8 box of a 3-D tensor at byte-granular x with TMA.
// usage: tma_probe W H pitch frames bw bh x y f
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                           const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
struct alignas(64) Maps { CUtensorMap m[8]; };
__global__ void k(const __grid_constant__ Maps maps, int l, int bw, int bh, int x, int y, int f, uint8_t* out, int mode, const CUtensorMap* gmap, int wmul) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar[4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* tile = smem + warp * 8192;
    uint32_t bsa = (uint32_t)__cvta_generic_to_shared(&bar[warp]);
    if (mode >= 4) { tile = smem + 4 * 8192 + 4096 * warp; bsa = (uint32_t)__cvta_generic_to_shared(smem + 4 * 8192 + 16384 + 8 * warp); }
    if (lane == 0) {
        uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile);
        printf("warp %d dst %x bar %x\n", warp, dst, bsa);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bsa) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bsa), "r"(bw * bh) : "memory");
        const CUtensorMap* mp = mode == 1 ? &maps.m[3] : mode == 2 ? gmap : &maps.m[l];
        if (mode == 3)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(dst), "l"(&maps.m[4]), "r"(bsa), "r"(x + warp * wmul), "r"(y + f * 480) : "memory");
        else
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(dst), "l"(mp), "r"(bsa), "r"(x + warp * wmul), "r"(y), "r"(f) : "memory");
    }
    __syncwarp();
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bsa) : "memory");
    __syncwarp();
    for (int i = lane; i < bw * bh; i += 32) out[warp * 8192 + i] = tile[i];
}
int main(int argc, char** argv) {
    int W = atoi(argv[1]), H = atoi(argv[2]), pitch = atoi(argv[3]), F = atoi(argv[4]), bw = atoi(argv[5]), bh = atoi(argv[6]);
    int mode = argc > 10 ? atoi(argv[10]) : 0;
    int x = atoi(argv[7]), y = atoi(argv[8]), f = atoi(argv[9]);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    enc_fn enc = (enc_fn)p;
    size_t fs = ((size_t)pitch * H + 15) & ~(size_t)15;
    std::vector<uint8_t> h(fs * F);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)(i * 2654435761u >> 24);
    uint8_t *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 4 * 8192);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    Maps maps; memset(&maps, 0, sizeof maps);
    cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)F};
    cuuint64_t str[2] = {(cuuint64_t)pitch, (cuuint64_t)fs};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
    CUresult r = enc(&maps.m[3], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    cuuint64_t dims2[2] = {(cuuint64_t)W, (cuuint64_t)H * F};
    r = enc(&maps.m[4], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims2, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode2d rc=%d\n", (int)r);
    CUtensorMap* gmap; cudaMalloc(&gmap, 128); cudaMemcpy(gmap, &maps.m[3], 128, cudaMemcpyHostToDevice);
    for (int i = 0; i < 16; ++i) printf("%016llx ", ((unsigned long long*)&maps.m[3])[i]);
    printf("\n");
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    k<<<1, mode == 5 ? 32 : 128, 56 * 1024>>>(maps, 3, bw, bh, x, y, f, o, mode, gmap, argc > 11 ? atoi(argv[11]) : 1);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e) return 1;
    std::vector<uint8_t> got(4 * 8192);
    cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int w = 0; w < (mode == 5 ? 1 : 4); ++w)
        for (int r2 = 0; r2 < bh; ++r2)
            for (int c = 0; c < bw; ++c) {
                int gx = x + w * (argc > 11 ? atoi(argv[11]) : 1) + c, gy = y + r2;
                uint8_t want = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? h[f * fs + (size_t)gy * pitch + gx] : 0;
                if (got[w * 8192 + r2 * bw + c] != want) ++bad;
            }
    printf("mismatches=%ld\n", bad);
    return 0;
}
