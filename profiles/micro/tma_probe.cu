// Probe: where does a TMA tile::gather4 load (box {Fp, 1}, SWIZZLE_128B / NONE) put each byte in shared memory?
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at line %d: %s\n", #x, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int FP>
__global__ void probe(const __grid_constant__ CUtensorMap tmap, const uint32_t* rows, uint8_t* out, uint32_t* info) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32 * FP);
  const int lane = threadIdx.x;
  for (int i = lane; i < 32 * FP; i += 32) smem[i] = 0xEE;
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    info[0] = smem_u32(smem_raw); info[1] = smem_u32(smem);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(32 * FP) : "memory");
  __syncwarp();
  if (lane < 8) {
    const int r0 = rows[4 * lane], r1 = rows[4 * lane + 1], r2 = rows[4 * lane + 2], r3 = rows[4 * lane + 3];
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                 ::"r"(smem_u32(smem + lane * 4 * FP)), "l"(&tmap), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(smem_u32(bar)) : "memory");
  }
  uint32_t ok = 0;
  int spins = 0;
  while (!ok && spins < 1000000) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(smem_u32(bar)) : "memory");
    spins++;
  }
  if (lane == 0) info[2] = ok;
  for (int i = lane; i < 32 * FP; i += 32) out[i] = smem[i];
}
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
template <int FP>
void run(EncodeFn encode, CUtensorMapSwizzle sw, const char* name) {
  const int N = 4096;
  std::vector<uint8_t> h((size_t)N * FP);
  for (int r = 0; r < N; r++) for (int c = 0; c < FP; c++) h[(size_t)r * FP + c] = (uint8_t)((r * 131 + c * 7) & 0xff);
  uint8_t *d, *d_out; uint32_t *d_rows, *d_info;
  CK(cudaMalloc(&d, h.size())); CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&d_out, 32 * FP)); CK(cudaMalloc(&d_rows, 128)); CK(cudaMalloc(&d_info, 16));
  std::vector<uint32_t> rows(32);
  for (int i = 0; i < 32; i++) rows[i] = (uint32_t)((i * 97 + 5) % N);
  CK(cudaMemcpy(d_rows, rows.data(), 128, cudaMemcpyHostToDevice));
  CUtensorMap tm;
  const cuuint64_t dims[2] = {FP, (cuuint64_t)N}, strides[1] = {FP};
  const cuuint32_t box[2] = {FP, 1}, estr[2] = {1, 1};
  CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("== Fp=%d %s: encode rc=%d\n", FP, name, (int)cr);
  if (cr != CUDA_SUCCESS) return;
  probe<FP><<<1, 32, 32 * FP + 2048>>>(tm, d_rows, d_out, d_info);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); exit(1); }
  std::vector<uint8_t> out(32 * FP); uint32_t info[4];
  CK(cudaMemcpy(out.data(), d_out, out.size(), cudaMemcpyDeviceToHost)); CK(cudaMemcpy(info, d_info, 16, cudaMemcpyDeviceToHost));
  printf("smem_raw 0x%x aligned 0x%x barrier completed %u\n", info[0], info[1], info[2]);
  // for every 16-byte chunk of shared memory: which (row index i in the batch, source chunk c) landed there?
  int plain = 0, swz = 0, unk = 0;
  for (int p = 0; p < 32 * FP / 16; p++) {
    int fi = -1, fc = -1;
    for (int i = 0; i < 32 && fi < 0; i++) for (int c = 0; c < FP / 16; c++) {
      bool eq = true;
      for (int k = 0; k < 16; k++) eq &= out[p * 16 + k] == h[(size_t)rows[i] * FP + c * 16 + k];
      if (eq) { fi = i; fc = c; break; }
    }
    const int logical = fi < 0 ? -1 : (fi * FP + fc * 16) / 16;
    const int a = logical * 16;
    const int expect_swz = logical < 0 ? -2 : ((a ^ (((a >> 7) & 7) << 4)) / 16);
    if (logical == p) plain++;
    if (expect_swz == p) swz++;
    if (fi < 0) unk++;
    if (p < 24) printf("chunk %2d <- batch row %2d src chunk %d (logical chunk %d)\n", p, fi, fc, logical);
  }
  printf("chunks at plain position %d, at 128B-swizzled position %d, unidentified %d of %d\n", plain, swz, unk, 32 * FP / 16);
  cudaFree(d); cudaFree(d_out); cudaFree(d_rows); cudaFree(d_info);
}
int main() {
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &q));
  run<64>(encode, CU_TENSOR_MAP_SWIZZLE_NONE, "no swizzle");
  run<64>(encode, CU_TENSOR_MAP_SWIZZLE_128B, "swizzle 128B");
  run<128>(encode, CU_TENSOR_MAP_SWIZZLE_128B, "swizzle 128B");
  run<128>(encode, CU_TENSOR_MAP_SWIZZLE_NONE, "no swizzle");
  run<80>(encode, CU_TENSOR_MAP_SWIZZLE_NONE, "no swizzle");
  run<32>(encode, CU_TENSOR_MAP_SWIZZLE_128B, "swizzle 128B");
  run<32>(encode, CU_TENSOR_MAP_SWIZZLE_NONE, "no swizzle");
  run<16>(encode, CU_TENSOR_MAP_SWIZZLE_NONE, "no swizzle");
  return 0;
}
