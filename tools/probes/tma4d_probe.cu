// Probe: 4-D tiled TMA box over (x, y, image, channel) of an NCHW fp32 tensor with SWIZZLE_128B, negative
// start coordinates and clipped extents (what the tap-major im2col SYRK issues).  Prints per-case status.
// nvcc -gencode arch=compute_100a,code=sm_100a -o tma4d_probe tma4d_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const __grid_constant__ CUtensorMap map, int c0, int c1, int c2, int c3, int bytes, float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
            smem_u32(smem)),
        "l"(reinterpret_cast<uint64_t>(&map)), "r"(smem_u32(&bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
  }
  asm volatile(
      "{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(
          smem_u32(&bar))
      : "memory");
  for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) out[i] = reinterpret_cast<float*>(smem)[i];
}

int main(int argc, char** argv) {
  cudaSetDevice(0);
  cudaFree(0);
  const int N = 9, C = 128, H = 28, W = 28;
  std::vector<float> h((size_t)N * C * H * W);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 100003) + 1.f;
  float *d, *out;
  cudaMalloc(&d, h.size() * 4);
  cudaMalloc(&out, 65536);
  cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  struct Case { int bx, by, bn, cb, wt, ht, x, y, n, c; };
  Case all[] = {{32, 1, 1, 128, 28, 28, 0, 0, 0, 0},  {32, 1, 1, 128, 27, 28, 0, 0, 0, 0},   {32, 1, 1, 128, 28, 27, 0, 27, 0, 0},
                {32, 1, 1, 128, 28, 28, -1, 0, 0, 0}, {32, 1, 1, 128, 28, 28, 0, -1, 0, 0},  {32, 1, 1, 128, 27, 27, -1, -1, 0, 0},
                {4, 4, 2, 128, 28, 28, 0, 0, 0, 0},   {4, 4, 2, 128, 27, 27, -1, -1, 8, 0},  {8, 4, 1, 64, 28, 28, 1, 1, 0, 64},
                {16, 2, 1, 128, 28, 28, 16, 26, 8, 0}, {32, 1, 1, 128, 28, 28, 4, 0, 0, 0}};
  const int which = argc > 1 ? atoi(argv[1]) : 0;
  Case cases[1] = {all[which]};
  for (const Case& k : cases) {
    CUtensorMap m;
    cuuint64_t dims[4] = {(cuuint64_t)k.wt, (cuuint64_t)k.ht, (cuuint64_t)N, (cuuint64_t)C};
    cuuint64_t strides[3] = {(cuuint64_t)W * 4, (cuuint64_t)C * H * W * 4, (cuuint64_t)H * W * 4};
    cuuint32_t box[4] = {(cuuint32_t)k.bx, (cuuint32_t)k.by, (cuuint32_t)k.bn, (cuuint32_t)k.cb};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    const int bytes = k.bx * k.by * k.bn * k.cb * 4;
    printf("box %dx%dx%dx%d dims %dx%d start (%d,%d,%d,%d): encode %d", k.bx, k.by, k.bn, k.cb, k.wt, k.ht, k.x, k.y, k.n, k.c, (int)r);
    if (r != CUDA_SUCCESS) { printf("\n"); continue; }
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    probe<<<1, 128, 65536>>>(m, k.x, k.y, k.n, k.c, bytes, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("  run: %s", cudaGetErrorString(e));
    if (e != cudaSuccess) { printf("\n"); return 1; }
    std::vector<float> o(bytes / 4);
    cudaMemcpy(o.data(), out, bytes, cudaMemcpyDeviceToHost);
    // check against the expected swizzled image: element (c, n, y, x) -> row c, k = (n*by + y)*bx + x
    int bad = 0;
    for (int c = 0; c < k.cb; ++c)
      for (int kk = 0; kk < 32; ++kk) {
        const int x = kk % k.bx, y = (kk / k.bx) % k.by, n = kk / (k.bx * k.by);
        const int gx = k.x + x, gy = k.y + y, gn = k.n + n, gc = k.c + c;
        float want = 0.f;
        if (gx >= 0 && gx < k.wt && gy >= 0 && gy < k.ht && gn < N && gc < C) want = h[(((size_t)gn * C + gc) * H + gy) * W + gx];
        const int chunk = (kk >> 2) ^ (c & 7);
        const float got = o[c * 32 + chunk * 4 + (kk & 3)];
        if (got != want) ++bad;
      }
    printf("  mismatches %d of %d\n", bad, k.cb * 32);
  }
  return 0;
}
