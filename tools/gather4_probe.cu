// Probe of cp.async.bulk.tensor.2d.tile::gather4 on sm_100a (no public documentation is reachable from this box):
//   * is the shared-memory image of 4 gathered rows the same as rows r..r+3 of a tile-mode SWIZZLE_128B box?
//   * what does a negative / past-the-end row index deliver (zero fill?) and does it count towards complete_tx?
// Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/g4 tools/gather4_probe.cu -lcuda && /tmp/g4
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

__global__ void probe(const __grid_constant__ CUtensorMap row_map, const __grid_constant__ CUtensorMap tile_map,
                      int4 idx, int col0, uint16_t* out_g4, uint16_t* out_tile, int* status) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar;
  uint32_t sb = (uint32_t)__cvta_generic_to_shared(smem);
  uint32_t bb = (uint32_t)__cvta_generic_to_shared(&bar);
  for (int i = threadIdx.x; i < 2048 / 2; i += blockDim.x) reinterpret_cast<uint16_t*>(smem)[i] = 0xEEEE;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bb));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;");
  if (threadIdx.x == 0) {
    // 4 gathered rows (512 B) into rows 4..7 of the first 8-row swizzle atom, the 8-row tile box into the second KB
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bb), "r"(512 + 1024));
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, "
        "%6, %7}], [%2];" ::"r"(sb + 512),
        "l"(&row_map), "r"(bb), "r"(col0), "r"(idx.x), "r"(idx.y), "r"(idx.z), "r"(idx.w)
        : "memory");
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            sb + 1024),
        "l"(&tile_map), "r"(bb), "r"(col0), "r"(0)
        : "memory");
  }
  // bounded wait: report instead of hanging when the byte count does not complete
  int ok = 0;
  for (int it = 0; it < 2000000 && !ok; ++it) {
    uint32_t p;
    asm volatile(
        "{.reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0; selp.u32 %0, 1, 0, q;}"
        : "=r"(p)
        : "r"(bb)
        : "memory");
    ok = p;
  }
  if (threadIdx.x == 0) *status = ok;
  __syncthreads();
  for (int i = threadIdx.x; i < 512; i += blockDim.x) {
    out_g4[i] = reinterpret_cast<uint16_t*>(smem)[i];
    out_tile[i] = reinterpret_cast<uint16_t*>(smem + 1024)[i];
  }
}

static CUtensorMap make_map(void* base, int rows, int cols, int box_rows) {
  CUtensorMap m;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
                                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                      CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) printf("cuTensorMapEncodeTiled(box_rows=%d) failed: %d\n", box_rows, (int)r);
  return m;
}

int main() {
  const int rows = 64, cols = 256;
  std::vector<uint16_t> h(rows * cols);
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c) h[r * cols + c] = (uint16_t)((r + 1) << 8 | c);   // row+1 so that 0 means "filled with zeros"
  uint16_t *d, *o1, *o2;
  int* st;
  cudaMalloc(&d, h.size() * 2);
  cudaMalloc(&o1, 1024);
  cudaMalloc(&o2, 1024);
  cudaMalloc(&st, 4);
  cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
  cuInit(0);
  CUtensorMap row_map = make_map(d, rows, cols, 1), tile_map = make_map(d, rows, cols, 8);
  struct Case { int4 idx; const char* what; } cases[] = {
      {{4, 5, 6, 7}, "rows 4..7 (must equal rows 4..7 of the tile box)"},
      {{9, 3, 60, 0}, "arbitrary rows"},
      {{5, -1, 7, 63}, "a negative index"},
      {{5, 64, 7, 100000}, "past-the-end indices"},
  };
  for (auto& c : cases) {
    cudaMemset(o1, 0xAB, 1024);
    probe<<<1, 128, 2048>>>(row_map, tile_map, c.idx, 64, o1, o2, st);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<uint16_t> g(512), t(512);
    int ok = -1;
    cudaMemcpy(g.data(), o1, 1024, cudaMemcpyDeviceToHost);
    cudaMemcpy(t.data(), o2, 1024, cudaMemcpyDeviceToHost);
    cudaMemcpy(&ok, st, 4, cudaMemcpyDeviceToHost);
    printf("== %s: idx {%d,%d,%d,%d}  err=%s  barrier completed=%d\n", c.what, c.idx.x, c.idx.y, c.idx.z, c.idx.w,
           cudaGetErrorString(e), ok);
    // smem rows 0..7 of the gather buffer (rows 0..3 untouched = EEEE), 16-byte chunks: print (row+1, first col) per chunk
    for (int r = 0; r < 8; ++r) {
      printf("  smem row %d:", r);
      for (int ch = 0; ch < 8; ++ch) printf(" %04x", g[r * 64 + ch * 8]);
      printf("   | tile:");
      for (int ch = 0; ch < 8; ++ch) printf(" %04x", t[r * 64 + ch * 8]);
      printf("\n");
    }
    if (e != cudaSuccess) break;
  }
  return 0;
}
