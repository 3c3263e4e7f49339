// TMA store throughput per SM as a function of the box shape (round 2: is the bf16 epilogue of the up-projection bound by
// the NUMBER of bulk stores / row requests rather than by bytes?).  148 CTAs x 16 warps, every warp streams boxes from its
// own staging buffer to its own rows of a [rows, 2048] bf16 tensor; the box is rows x cols with cols*2 bytes per row
// (64-byte rows with SWIZZLE_64B = the shape gemm.cu's epilogue uses today, 128-byte rows with SWIZZLE_128B).
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/tsp tools/tma_store_probe.cu -lcuda && /tmp/tsp
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__global__ void __launch_bounds__(512, 1)
store_kernel(const __grid_constant__ CUtensorMap m, int box_rows, int box_cols, int iters, int pending, int rows_per_cta) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int box_bytes = box_rows * box_cols * 2;
  unsigned char* stg = smem + warp * 8192;                    // up to two 4 KB boxes per warp
  for (int i = lane; i < 8192 / 4; i += 32) reinterpret_cast<uint32_t*>(stg)[i] = 0x3f803f80u;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  const int cols_total = 2048;
  const int boxes_per_row_block = cols_total / box_cols;
  // warp w of CTA b owns rows [b*rows_per_cta + (w % 4)*32 .. +32) x a column range; iterate over column boxes and row blocks
  if (lane == 0) {
    for (int it = 0; it < iters; ++it) {
      const int col = (it % boxes_per_row_block) * box_cols;
      const int row = blockIdx.x * rows_per_cta + ((warp + it / boxes_per_row_block * 16) * 32) % rows_per_cta;
      if (pending == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      const uint32_t src = (uint32_t)__cvta_generic_to_shared(stg + (pending ? (it & 1) * box_bytes : 0));
      asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                   :: "l"(&m), "r"(src), "r"(col), "r"(row) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  cuInit(0);
  const int ctas = 148, rows_per_cta = 2048, cols = 2048;
  const size_t rows = (size_t)ctas * rows_per_cta;
  void* d;
  cudaMalloc(&d, rows * cols * 2);                            // 1.24 GB
  cudaFuncSetAttribute(store_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16 * 8192);
  struct Shape { int r, c; CUtensorMapSwizzle s; const char* name; } shapes[] = {
      {32, 32, CU_TENSOR_MAP_SWIZZLE_64B, "32 rows x 64 B (today)"},
      {32, 64, CU_TENSOR_MAP_SWIZZLE_128B, "32 rows x 128 B"},
      {16, 64, CU_TENSOR_MAP_SWIZZLE_128B, "16 rows x 128 B"},
      {32, 16, CU_TENSOR_MAP_SWIZZLE_32B, "32 rows x 32 B"},
  };
  for (auto& sh : shapes) {
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)sh.c, (cuuint32_t)sh.r};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, sh.s, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); continue; }
    for (int pending = 0; pending < 2; ++pending) {
      for (int nwarps : {4, 16}) {
        const int box_bytes = sh.r * sh.c * 2;
        const int iters = (64 << 20) / ctas / nwarps / box_bytes * 4;      // ~256 MB per launch
        cudaEvent_t a, b;
        cudaEventCreate(&a); cudaEventCreate(&b);
        store_kernel<<<ctas, nwarps * 32, 16 * 8192>>>(m, sh.r, sh.c, iters, pending, rows_per_cta);
        cudaEventRecord(a);
        store_kernel<<<ctas, nwarps * 32, 16 * 8192>>>(m, sh.r, sh.c, iters, pending, rows_per_cta);
        cudaEventRecord(b);
        cudaError_t e = cudaDeviceSynchronize();
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        const double bytes = (double)ctas * nwarps * iters * box_bytes;
        const double stores = (double)ctas * nwarps * iters;
        printf("%-24s warps/CTA %2d  boxes in flight per warp %d: %7.1f us  %6.0f GB/s  %5.1f B/clk/SM  %6.1f ns per store per SM (%s)\n",
               sh.name, nwarps, pending + 1, ms * 1e3, bytes / ms / 1e6, bytes / ctas / (ms * 1e-3) / 1.9e9,
               ms * 1e6 / (stores / ctas), cudaGetErrorString(e));
      }
    }
  }
  return 0;
}
