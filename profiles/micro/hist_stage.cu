// Microbenchmark: row staging for the histogram kernel's accumulate loop on B200 (sm_100a).
// One CTA = one tile of one node's entry segment, as in hist_scan_kernel: zero a [8][32][50] u32 shared-memory
// histogram, add F_sub = 8 bins per entry, fold the histogram into a per-node checksum.  Variants differ only in how
// the entries' TreePoint rows (Fp = 64 bytes, row-major u8 matrix of 1M rows) reach shared memory:
//   V0  cooperative LDG.128 -> registers -> transposed STS.32 tile (round-1 kernel)
//   V1  dedicated producer warp(s): TMA tile::gather4 (four rows per instruction, SWIZZLE_128B) into an mbarrier ring,
//       consumers read lane = row (one LDS.U8 per feature)
//   V2  as V1, consumers read lane = row with the feature order rotated per 8-lane group (bank-conflict schedule)
//   V3  no producer warp: every consumer warp issues the gathers of its own next batches (private ring), schedule as V2
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o hist_stage hist_stage.cu ; run: ./hist_stage
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int FP = 64, NB = 32, NC = 50, FSUB = 8, BC = NB * NC, HIST_WORDS = FSUB * BC;
constexpr int NCONS = 256;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at line %d: %s\n", #x, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

struct Node { uint64_t start; uint32_t len; uint32_t pad; int32_t feat[FSUB]; };
struct Tile { uint32_t node; uint32_t first; uint32_t len; uint32_t pad; };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ unsigned int g_timeouts[8];
// bounded wait (microbenchmark only): gives up after ~0.2 s and counts the give-up instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int who = 0) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 400000000ll) { atomicAdd(&g_timeouts[who], 1u); return; }
  }
}
__device__ __forceinline__ void tma_gather4(uint32_t dst, const CUtensorMap* tm, int r0, int r1, int r2, int r3, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
               ::"r"(dst), "l"(tm), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(bar) : "memory");
}
// byte offset of (row r of the batch, feature f) inside a 1024-byte aligned stage written by TMA with SWIZZLE_128B
__device__ __forceinline__ uint32_t swz(uint32_t r, uint32_t f) {
  const uint32_t a = r * FP + f;
  return a ^ (((a >> 7) & (uint32_t)(FP / 16 - 1)) << 4);
}

struct Args {
  const uint8_t* bins; const uint64_t* ent; const Node* nodes; const Tile* tiles; uint32_t* checksum; uint32_t n_tiles;
};

__device__ __forceinline__ void fold(const uint32_t* sh, uint32_t* out) {
  uint32_t acc = 0;
  for (int i = threadIdx.x; i < HIST_WORDS; i += NCONS) acc += sh[i] * (uint32_t)(i + 1);
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

// ---------------------------------------------------------------- V0: round-1 staging
__global__ void __launch_bounds__(NCONS, 3) k_legacy(const Args A) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint32_t* sh = reinterpret_cast<uint32_t*>(smem);
  constexpr int TS = 34, CH = 4;
  uint32_t* tiles_sm = sh + HIST_WORDS;
  __shared__ int sh_off[FSUB];
  const int tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
  const Tile tl = A.tiles[blockIdx.x];
  const Node& nd = A.nodes[tl.node];
  if (tid < FSUB) { const int f = nd.feat[tid]; sh_off[tid] = (((f >> 2) * TS) << 2) + (f & 3); }
  for (int i = tid; i < HIST_WORDS; i += NCONS) sh[i] = 0;
  __syncthreads();
  const uint64_t* ent = A.ent + nd.start + tl.first;
  const uint32_t len = tl.len;
  uint32_t* tile = tiles_sm + wv * (16 * TS);
  const uint8_t* tb = reinterpret_cast<const uint8_t*>(tile) + (lane << 2);
  int offr[FSUB];
#pragma unroll
  for (int i = 0; i < FSUB; i++) offr[i] = sh_off[i];
  const uint32_t n_batches = (len + 31) / 32;
  auto load_entry = [&](uint32_t j) { return j < len ? __ldcs(ent + j) : 0ull; };
  auto load_rows = [&](uint32_t my_row, uint4 (&R)[CH]) {
#pragma unroll
    for (int it = 0; it < CH; it++) {
      const int q = it * 32 + lane, r = q / CH, c = q - r * CH;
      const uint32_t row = __shfl_sync(0xffffffffu, my_row, r & 31);
      R[it] = __ldcg(reinterpret_cast<const uint4*>(A.bins + (size_t)row * FP) + c);
    }
  };
  uint4 R[CH];
  uint32_t b = wv;
  uint64_t e_cur = load_entry(b * 32 + lane), e_nxt = load_entry((b + 8) * 32 + lane);
  if (b < n_batches) load_rows((uint32_t)e_cur, R);
  for (; b < n_batches; b += 8) {
#pragma unroll
    for (int it = 0; it < CH; it++) {
      const int q = it * 32 + lane, r = q / CH, c = q - r * CH;
      uint32_t* d = tile + (4 * c) * TS + r;
      d[0] = R[it].x; d[TS] = R[it].y; d[2 * TS] = R[it].z; d[3 * TS] = R[it].w;
    }
    __syncwarp();
    const uint64_t e_nn = load_entry((b + 16) * 32 + lane);
    if (b + 8 < n_batches) load_rows((uint32_t)e_nxt, R);
    const uint32_t w = (uint32_t)(e_cur >> 32) & 0xffu, lab = (uint32_t)(e_cur >> 40) & 0xffu;
    if (w) {
      uint32_t hl = lab;
#pragma unroll
      for (int i = 0; i < FSUB; i++) { atomicAdd(sh + hl + tb[offr[i]] * NC, w); hl += BC; }
    }
    __syncwarp();
    e_cur = e_nxt; e_nxt = e_nn;
  }
  __syncthreads();
  fold(sh, A.checksum + tl.node);
}

// ---------------------------------------------------------------- V1/V2: producer warps + mbarrier ring
// shared memory: [hist 51200][pad to 1024][S stages of 32*FP][meta S*32 u32][full S u64][empty S u64]
template <int S, int P, bool SCHED>
__global__ void __launch_bounds__(NCONS + 32 * P, (S <= 10 ? 3 : 2)) k_ring(const Args A, const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // TMA swizzle patterns repeat every 1024 B
  uint8_t* stages = smem;
  uint32_t* sh = reinterpret_cast<uint32_t*>(stages + S * 32 * FP);
  uint32_t* meta = sh + HIST_WORDS;
  uint64_t* full = reinterpret_cast<uint64_t*>(meta + S * 32);
  uint64_t* empty = full + S;
  int* sh_feat = reinterpret_cast<int*>(empty + S);
  int* sh_perm = sh_feat + FSUB;
  uint32_t* rowx = reinterpret_cast<uint32_t*>(sh_perm + FSUB);     // [P][4][32] row ids, producer-private
  const int tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
  const Tile tl = A.tiles[blockIdx.x];
  const Node& nd = A.nodes[tl.node];
  if (tid == 0) {
    for (int s = 0; s < S; s++) { mbar_init(smem_u32(full + s), 1); mbar_init(smem_u32(empty + s), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // features ordered by the bank class of their byte ((f >> 2) & 3): neighbours in the order have different classes
    int f[FSUB], k[FSUB];
    for (int i = 0; i < FSUB; i++) { f[i] = nd.feat[i]; k[i] = i; sh_feat[i] = f[i]; }
    for (int i = 1; i < FSUB; i++)
      for (int j = i; j > 0 && ((f[k[j]] >> 2) & 3) < ((f[k[j - 1]] >> 2) & 3); j--) { const int t = k[j]; k[j] = k[j - 1]; k[j - 1] = t; }
    for (int i = 0; i < FSUB; i++) sh_perm[i] = k[i];
  }
  if (tid < NCONS) for (int i = tid; i < HIST_WORDS; i += NCONS) sh[i] = 0;
  __syncthreads();
  const uint64_t* ent = A.ent + nd.start + tl.first;
  const uint32_t len = tl.len;
  const uint32_t n_batches = (len + 31) / 32;
  if (wv >= NCONS / 32) {
    // ---- producer: lane group g = lane / 8 takes one batch, lane j of the group gathers rows 4j .. 4j+3 of it
    const int p = wv - NCONS / 32, g = lane >> 3, j = lane & 7;
    for (uint32_t b0 = 4 * p; b0 < n_batches; b0 += 4 * P) {
      const uint32_t b = b0 + g;
      const bool act = b < n_batches;
      const uint32_t s = b % S, use = b / S;
      // the warp loads the 128 entries of its four batches coalesced (lane l: entry l of batch b0 + i) ...
      uint64_t e[4];
#pragma unroll
      for (int i = 0; i < 4; i++) { const uint32_t jj = (b0 + i) * 32 + lane; e[i] = jj < len ? __ldcs(ent + jj) : 0ull; }
      if (act && j == 0) mbar_wait(smem_u32(empty + s), (use & 1) ^ 1, 1);
      __syncwarp();
      // ... hands the row ids to the gathering lanes through shared memory and the (weight, label) pairs to the consumers
#pragma unroll
      for (int i = 0; i < 4; i++) {
        rowx[p * 128 + i * 32 + lane] = (uint32_t)e[i];
        if (b0 + i < n_batches) meta[((b0 + i) % S) * 32 + lane] = (uint32_t)(e[i] >> 32) & 0xffffu;
      }
      __syncwarp();
      const uint4 r4 = reinterpret_cast<const uint4*>(rowx + p * 128 + g * 32)[j];
      if (act) {
        if (j == 0) mbar_expect_tx(smem_u32(full + s), 32 * FP);
        tma_gather4(smem_u32(stages + s * 32 * FP + j * 4 * FP), &tmap, (int)r4.x, (int)r4.y, (int)r4.z, (int)r4.w, smem_u32(full + s));
      }
    }
  } else {
    // ---- consumers
    const int q = lane >> 3;
    uint32_t off[FSUB], plane[FSUB];
#pragma unroll
    for (int i = 0; i < FSUB; i++) {
      const int k = SCHED ? sh_perm[(i + 2 * q) & 7] : i;
      off[i] = swz(lane, sh_feat[k]);
      plane[i] = k * BC;
    }
    for (uint32_t b = wv; b < n_batches; b += NCONS / 32) {
      const uint32_t s = b % S, use = b / S;
      mbar_wait(smem_u32(full + s), use & 1, 2);
      const uint32_t m = meta[s * 32 + lane];
      const uint32_t w = m & 0xffu, lab = m >> 8;
      const uint8_t* st = stages + s * 32 * FP;
      if (w) {
#pragma unroll
        for (int i = 0; i < FSUB; i++) atomicAdd(sh + plane[i] + st[off[i]] * NC + lab, w);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(empty + s));
    }
  }
  __syncthreads();
  if (tid < NCONS) fold(sh, A.checksum + tl.node);
}

// ---------------------------------------------------------------- V3: every warp gathers for itself (D private stages)
template <int D, int MODE = 0>
__global__ void __launch_bounds__(NCONS, (D <= 1 ? 3 : 2)) k_self(const Args A, const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int NW = NCONS / 32, S = NW * D;
  uint8_t* stages = smem;
  uint32_t* sh = reinterpret_cast<uint32_t*>(stages + S * 32 * FP);
  uint32_t* meta = sh + HIST_WORDS;
  uint64_t* full = reinterpret_cast<uint64_t*>(meta + S * 32);
  int* sh_feat = reinterpret_cast<int*>(full + S);
  int* sh_perm = sh_feat + FSUB;
  const int tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
  const Tile tl = A.tiles[blockIdx.x];
  const Node& nd = A.nodes[tl.node];
  if (tid == 0) {
    for (int s = 0; s < S; s++) mbar_init(smem_u32(full + s), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    int f[FSUB], k[FSUB];
    for (int i = 0; i < FSUB; i++) { f[i] = nd.feat[i]; k[i] = i; sh_feat[i] = f[i]; }
    for (int i = 1; i < FSUB; i++)
      for (int j = i; j > 0 && ((f[k[j]] >> 2) & 3) < ((f[k[j - 1]] >> 2) & 3); j--) { const int t = k[j]; k[j] = k[j - 1]; k[j - 1] = t; }
    for (int i = 0; i < FSUB; i++) sh_perm[i] = k[i];
  }
  for (int i = tid; i < HIST_WORDS; i += NCONS) sh[i] = 0;
  __syncthreads();
  const uint64_t* ent = A.ent + nd.start + tl.first;
  const uint32_t len = tl.len;
  const uint32_t n_batches = (len + 31) / 32;
  const int q = lane >> 3;
  uint32_t off[FSUB], plane[FSUB];
#pragma unroll
  for (int i = 0; i < FSUB; i++) {
    const int k = sh_perm[(i + 2 * q) & 7];
    off[i] = swz(lane, sh_feat[k]);
    plane[i] = k * BC;
  }
  auto load_entry = [&](uint32_t b) { const uint32_t j = b * 32 + lane; return (b < n_batches && j < len) ? __ldcs(ent + j) : 0ull; };
  // issue the gathers of the warp's i-th batch (entry of this lane: e) into private stage i % D
  auto issue = [&](uint32_t i, uint64_t e) {
    const uint32_t d = i % D, s = wv * D + d;
    meta[s * 32 + lane] = (uint32_t)(e >> 32) & 0xffffu;
    const uint32_t row = (uint32_t)e;
    const int src = 4 * (lane & 7);
    const int r0 = __shfl_sync(0xffffffffu, row, src), r1 = __shfl_sync(0xffffffffu, row, src + 1);
    const int r2 = __shfl_sync(0xffffffffu, row, src + 2), r3 = __shfl_sync(0xffffffffu, row, src + 3);
    if (MODE == 2) return;        // consume only: nothing is fetched
    if (lane == 0) mbar_expect_tx(smem_u32(full + s), 32 * FP);
    __syncwarp();
    if (lane < 8) tma_gather4(smem_u32(stages + s * 32 * FP + lane * 4 * FP), &tmap, r0, r1, r2, r3, smem_u32(full + s));
  };
  const uint32_t my_batches = wv < n_batches ? (n_batches - wv + NW - 1) / NW : 0;
#pragma unroll
  for (int i = 0; i < D; i++)
    if ((uint32_t)i < my_batches) issue(i, load_entry(wv + NW * i));
  uint64_t e_pref = load_entry(wv + NW * D);
  for (uint32_t i = 0; i < my_batches; i++) {
    const uint32_t d = i % D, s = wv * D + d;
    if (MODE != 2) mbar_wait(smem_u32(full + s), (i / D) & 1, 3);
    const uint32_t m = meta[s * 32 + lane];
    const uint32_t w = m & 0xffu, lab = m >> 8;
    const uint8_t* st = stages + s * 32 * FP;
    if (MODE == 1) {              // fetch only: one byte read per batch, no histogram work
      if (st[off[0]] == 77 && w == 99) atomicAdd(sh, 1u);
    } else if (w) {
#pragma unroll
      for (int k = 0; k < FSUB; k++) atomicAdd(sh + plane[k] + (MODE == 2 ? (st[off[k]] & 31u) : st[off[k]]) * NC + lab, w);
    }
    __syncwarp();
    const uint64_t e_now = e_pref;
    e_pref = load_entry(wv + NW * (i + D + 1));
    if (i + D < my_batches) issue(i + D, e_now);
  }
  __syncthreads();
  fold(sh, A.checksum + tl.node);
}


// ---------------------------------------------------------------- V4: both engines at once.  Warps [0, NT) stage their batches
// with TMA gathers (one private stage each), warps [NT, 8) with LDG.128 + STS as V0; every warp claims its next batch from
// a shared counter, so the split between the copy engine and the LSU pipe balances itself.
// ISSUE0: lane 0 issues the eight gathers of a batch in straight-line code (row ids handed over through shared memory)
template <int NT, bool ISSUE0>
__global__ void __launch_bounds__(NCONS, 3) k_hybrid(const Args A, const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int NW = NCONS / 32, NL = NW - NT, TS = 34, CH = 4;
  uint8_t* stages = smem;                                                   // [NT][32*FP]
  uint32_t* ltiles = reinterpret_cast<uint32_t*>(stages + NT * 32 * FP);    // [NL][16*TS]
  uint32_t* sh = ltiles + NL * 16 * TS;
  uint32_t* meta = sh + HIST_WORDS;                                         // [NT][32]
  uint64_t* full = reinterpret_cast<uint64_t*>(meta + NT * 32);
  int* sh_feat = reinterpret_cast<int*>(full + NT);
  int* sh_perm = sh_feat + FSUB;
  uint32_t* next_batch = reinterpret_cast<uint32_t*>(sh_perm + FSUB);
  const int tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
  const Tile tl = A.tiles[blockIdx.x];
  const Node& nd = A.nodes[tl.node];
  if (tid == 0) {
    for (int s = 0; s < NT; s++) mbar_init(smem_u32(full + s), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    *next_batch = 0;
    int f[FSUB], k[FSUB];
    for (int i = 0; i < FSUB; i++) { f[i] = nd.feat[i]; k[i] = i; sh_feat[i] = f[i]; }
    for (int i = 1; i < FSUB; i++)
      for (int j = i; j > 0 && ((f[k[j]] >> 2) & 3) < ((f[k[j - 1]] >> 2) & 3); j--) { const int t = k[j]; k[j] = k[j - 1]; k[j - 1] = t; }
    for (int i = 0; i < FSUB; i++) sh_perm[i] = k[i];
  }
  for (int i = tid; i < HIST_WORDS; i += NCONS) sh[i] = 0;
  __syncthreads();
  const uint64_t* ent = A.ent + nd.start + tl.first;
  const uint32_t len = tl.len;
  const uint32_t n_batches = (len + 31) / 32;
  auto claim = [&]() { uint32_t b = 0; if (lane == 0) b = atomicAdd(next_batch, 1u); return __shfl_sync(0xffffffffu, b, 0); };
  auto load_entry = [&](uint32_t b) { const uint32_t j = b * 32 + lane; return (b < n_batches && j < len) ? __ldcs(ent + j) : 0ull; };
  if (wv < NT) {
    const int q = lane >> 3;
    uint32_t off[FSUB], plane[FSUB];
#pragma unroll
    for (int i = 0; i < FSUB; i++) {
      const int k = sh_perm[(i + 2 * q) & 7];
      off[i] = swz(lane, sh_feat[k]);
      plane[i] = k * BC;
    }
    const uint32_t s = wv;
    const uint32_t bar = smem_u32(full + s), st_addr = smem_u32(stages + s * 32 * FP);
    const uint8_t* st = stages + s * 32 * FP;
    uint32_t* my_meta = meta + s * 32;
    uint32_t* my_rows = reinterpret_cast<uint32_t*>(next_batch + 4) + s * 32;   // row ids for the issuing lane (ISSUE0)
    auto issue = [&](uint64_t e) {
      const uint32_t row = (uint32_t)e;
      uint32_t mw = (uint32_t)(e >> 32) & 0xffffu;
      if (ISSUE0) {
        my_meta[lane] = mw;
        my_rows[lane] = row;
        __syncwarp();
        if (lane == 0) {
          mbar_expect_tx(bar, 32 * FP);
#pragma unroll
          for (int g = 0; g < 8; g++) {
            const uint4 m4 = reinterpret_cast<const uint4*>(my_rows)[g];
            tma_gather4(st_addr + g * 4 * FP, &tmap, (int)m4.x, (int)m4.y, (int)m4.z, (int)m4.w, bar);
          }
        }
        __syncwarp();
      } else {
        my_meta[lane] = mw;
        const int src = 4 * (lane & 7);
        const int r0 = __shfl_sync(0xffffffffu, row, src), r1 = __shfl_sync(0xffffffffu, row, src + 1);
        const int r2 = __shfl_sync(0xffffffffu, row, src + 2), r3 = __shfl_sync(0xffffffffu, row, src + 3);
        if (lane == 0) mbar_expect_tx(bar, 32 * FP);
        __syncwarp();
        if (lane < 8) tma_gather4(st_addr + lane * 4 * FP, &tmap, r0, r1, r2, r3, bar);
      }
    };
    uint32_t b = claim();
    if (b < n_batches) issue(load_entry(b));
    uint32_t ph = 0;
    while (b < n_batches) {
      const uint32_t b2 = claim();
      const uint64_t e2 = load_entry(b2);
      mbar_wait(bar, ph, 3);
      ph ^= 1;
      const uint32_t m = my_meta[lane];
      const uint32_t w = m & 0xffu, lab = m >> 8;
      if (w) {
#pragma unroll
        for (int k = 0; k < FSUB; k++) atomicAdd(sh + plane[k] + st[off[k]] * NC + lab, w);
      }
      __syncwarp();
      if (b2 < n_batches) issue(e2);
      b = b2;
    }
  } else {
    uint32_t* tile = ltiles + (wv - NT) * (16 * TS);
    const uint8_t* tb = reinterpret_cast<const uint8_t*>(tile) + (lane << 2);
    int offr[FSUB];
#pragma unroll
    for (int i = 0; i < FSUB; i++) { const int f = sh_feat[i]; offr[i] = (((f >> 2) * TS) << 2) + (f & 3); }
    auto load_rows = [&](uint32_t my_row, uint4 (&R)[CH]) {
#pragma unroll
      for (int it = 0; it < CH; it++) {
        const int qq = it * 32 + lane, r = qq / CH, c = qq - r * CH;
        const uint32_t row = __shfl_sync(0xffffffffu, my_row, r & 31);
        R[it] = __ldcg(reinterpret_cast<const uint4*>(A.bins + (size_t)row * FP) + c);
      }
    };
    uint4 R[CH];
    uint32_t b = claim();
    uint64_t e_cur = load_entry(b);
    if (b < n_batches) load_rows((uint32_t)e_cur, R);
    while (b < n_batches) {
      const uint32_t b2 = claim();
      const uint64_t e_nxt = load_entry(b2);
#pragma unroll
      for (int it = 0; it < CH; it++) {
        const int qq = it * 32 + lane, r = qq / CH, c = qq - r * CH;
        uint32_t* d = tile + (4 * c) * TS + r;
        d[0] = R[it].x; d[TS] = R[it].y; d[2 * TS] = R[it].z; d[3 * TS] = R[it].w;
      }
      __syncwarp();
      if (b2 < n_batches) load_rows((uint32_t)e_nxt, R);
      const uint32_t w = (uint32_t)(e_cur >> 32) & 0xffu, lab = (uint32_t)(e_cur >> 40) & 0xffu;
      if (w) {
        uint32_t hl = lab;
#pragma unroll
        for (int i = 0; i < FSUB; i++) { atomicAdd(sh + hl + tb[offr[i]] * NC, w); hl += BC; }
      }
      __syncwarp();
      e_cur = e_nxt; b = b2;
    }
  }
  __syncthreads();
  fold(sh, A.checksum + tl.node);
}

// ---------------------------------------------------------------- host
static uint64_t mix(uint64_t x) { x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; return x ^ (x >> 31); }

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct Level { std::vector<uint64_t> ent; std::vector<Node> nodes; std::vector<Tile> tiles; };

static void make_level(int L, int N, int T, Level& out) {
  const int per = 1 << L;
  out.nodes.resize((size_t)T * per);
  std::vector<uint32_t> cnt(per);
  out.ent.clear();
  out.ent.reserve((size_t)(0.64 * N) * T);
  for (int t = 0; t < T; t++) {
    std::fill(cnt.begin(), cnt.end(), 0u);
    for (int r = 0; r < N; r++) { const uint64_t h = mix(((uint64_t)t << 32) | (uint32_t)r); if ((h & 0xffff) < 41420) cnt[(h >> 16) & (per - 1)]++; }
    const size_t base = out.ent.size();
    std::vector<uint64_t> pos(per);
    uint64_t run = base;
    for (int n = 0; n < per; n++) {
      Node& nd = out.nodes[(size_t)t * per + n];
      nd.start = run; nd.len = cnt[n]; pos[n] = run; run += cnt[n];
      uint64_t fs = mix(0x1234567ull + (uint64_t)t * per + n);
      int nf = 0;
      while (nf < FSUB) { const int f = (int)(fs % 64); fs = mix(fs); bool dup = false; for (int i = 0; i < nf; i++) dup |= nd.feat[i] == f; if (!dup) nd.feat[nf++] = f; }
    }
    out.ent.resize(run);
    for (int r = 0; r < N; r++) {
      const uint64_t h = mix(((uint64_t)t << 32) | (uint32_t)r);
      if ((h & 0xffff) < 41420) {
        const uint32_t w = 1 + (uint32_t)((h >> 40) % 3), lab = (uint32_t)((h >> 48) % NC);
        out.ent[pos[(h >> 16) & (per - 1)]++] = (uint64_t)(uint32_t)r | ((uint64_t)w << 32) | ((uint64_t)lab << 40);
      }
    }
  }
  const uint64_t total = out.ent.size();
  uint32_t tile_len = (uint32_t)std::max<uint64_t>(2048, (total + 16 * 148 - 1) / (16 * 148));
  tile_len = (tile_len + 31) & ~31u;
  out.tiles.clear();
  for (size_t n = 0; n < out.nodes.size(); n++)
    for (uint32_t f = 0; f < out.nodes[n].len || f == 0; f += tile_len) {
      out.tiles.push_back(Tile{(uint32_t)n, f, std::min(tile_len, out.nodes[n].len - f), 0});
      if (out.nodes[n].len == 0) break;
    }
}

template <class F>
static float time_it(F launch, int reps = 3) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int i = 0; i < reps; i++) {
    CK(cudaEventRecord(a));
    launch();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    best = std::min(best, ms);
  }
  CK(cudaGetLastError());
  return best;
}

int main(int argc, char** argv) {
  setvbuf(stdout, nullptr, _IONBF, 0);
  const int N = 1000000, T = argc > 1 ? atoi(argv[1]) : 100;
  std::vector<uint8_t> bins((size_t)N * FP);
  for (size_t i = 0; i < bins.size(); i += 8) { const uint64_t h = mix(i + 77); for (int j = 0; j < 8; j++) bins[i + j] = (h >> (8 * j)) & 31; }
  uint8_t* d_bins;
  CK(cudaMalloc(&d_bins, bins.size()));
  CK(cudaMemcpy(d_bins, bins.data(), bins.size(), cudaMemcpyHostToDevice));
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres));
  if (!encode || qres != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
  CUtensorMap tmap;
  const cuuint64_t dims[2] = {FP, (cuuint64_t)N}, strides[1] = {FP};
  const cuuint32_t box[2] = {FP, 1}, estr[2] = {1, 1};
  const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d_bins, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             (FP == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : FP == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : FP == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE), CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d\n", (int)cr); return 1; }

  int clk;
  CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
  const int levels[] = {4, 7, 9};
  for (int L : levels) {
    Level lv;
    make_level(L, N, T, lv);
    uint64_t* d_ent; Node* d_nodes; Tile* d_tiles; uint32_t* d_ck;
    CK(cudaMalloc(&d_ent, lv.ent.size() * 8 + 64)); CK(cudaMalloc(&d_nodes, lv.nodes.size() * sizeof(Node)));
    CK(cudaMalloc(&d_tiles, lv.tiles.size() * sizeof(Tile))); CK(cudaMalloc(&d_ck, lv.nodes.size() * 4));
    CK(cudaMemcpy(d_ent, lv.ent.data(), lv.ent.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_nodes, lv.nodes.data(), lv.nodes.size() * sizeof(Node), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_tiles, lv.tiles.data(), lv.tiles.size() * sizeof(Tile), cudaMemcpyHostToDevice));
    Args A{d_bins, d_ent, d_nodes, d_tiles, d_ck, (uint32_t)lv.tiles.size()};
    const uint32_t grid = (uint32_t)lv.tiles.size();
    printf("== level %d: %zu nodes, %zu entries, %u tiles (tile_len %u)\n", L, lv.nodes.size(), lv.ent.size(), grid, lv.tiles[0].len);
    std::vector<uint32_t> ref(lv.nodes.size()), got(lv.nodes.size());
    auto report = [&](const char* name, float ms, bool is_ref) {
      CK(cudaMemcpy(got.data(), d_ck, got.size() * 4, cudaMemcpyDeviceToHost));
      if (is_ref) ref = got;
      unsigned int to[8] = {0};
      CK(cudaMemcpyFromSymbol(to, g_timeouts, sizeof to));
      if (to[1] | to[2] | to[3]) { printf("   !! barrier wait timeouts: producer(empty) %u consumer(full) %u self(full) %u\n", to[1], to[2], to[3]); unsigned int z[8] = {0}; CK(cudaMemcpyToSymbol(g_timeouts, z, sizeof z)); }
      size_t bad = 0;
      for (size_t i = 0; i < got.size(); i++) bad += got[i] != ref[i];
      const double batches_per_sm = (double)lv.ent.size() / 32 / 148;
      printf("%-46s %8.3f ms  %6.1f G entries/s  %6.1f cycles/batch/SM  %s\n", name, ms, lv.ent.size() / ms * 1e-6,
             ms * 1e-3 * clk * 1e3 / batches_per_sm, bad ? "CHECKSUM MISMATCH" : "ok");
    };
    // V0
    {
      const size_t smem = HIST_WORDS * 4 + 8 * 16 * 34 * 4;
      CK(cudaFuncSetAttribute(k_legacy, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      const float ms = time_it([&] { cudaMemsetAsync(d_ck, 0, lv.nodes.size() * 4); k_legacy<<<grid, NCONS, smem>>>(A); });
      report("V0 LDG.128 + STS staging (3 CTAs/SM)", ms, true);
    }
#define RING(S_, P_, SCHED_, label)                                                                                        \
    {                                                                                                                      \
      const size_t smem = 1024 + HIST_WORDS * 4 + (size_t)S_ * 32 * FP + S_ * 128 + 2 * S_ * 8 + 64 + P_ * 512;               \
      CK(cudaFuncSetAttribute(k_ring<S_, P_, SCHED_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));            \
      int occ = 0;                                                                                                         \
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_ring<S_, P_, SCHED_>, NCONS + 32 * P_, smem));              \
      const float ms = time_it([&] { cudaMemsetAsync(d_ck, 0, lv.nodes.size() * 4); k_ring<S_, P_, SCHED_><<<grid, NCONS + 32 * P_, smem>>>(A, tmap); }); \
      char nm[128]; snprintf(nm, sizeof nm, "%s S=%d P=%d (%d CTAs/SM)", label, S_, P_, occ);                               \
      report(nm, ms, false);                                                                                               \
    }
#define SELFM(D_, M_, label)                                                                                               \
    {                                                                                                                      \
      const size_t smem = 1024 + HIST_WORDS * 4 + (size_t)8 * D_ * 32 * FP + 8 * D_ * 128 + 8 * D_ * 8 + 64;                \
      CK(cudaFuncSetAttribute(k_self<D_, M_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                    \
      int occ = 0;                                                                                                         \
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_self<D_, M_>, NCONS, smem));                                \
      const float ms = time_it([&] { cudaMemsetAsync(d_ck, 0, lv.nodes.size() * 4); k_self<D_, M_><<<grid, NCONS, smem>>>(A, tmap); }); \
      char nm[128]; snprintf(nm, sizeof nm, "V3 %s D=%d (%d CTAs/SM)", label, D_, occ);                                     \
      report(nm, ms, false);                                                                                               \
    }
#define SELF(D_)                                                                                                           \
    {                                                                                                                      \
      const size_t smem = 1024 + HIST_WORDS * 4 + (size_t)8 * D_ * 32 * FP + 8 * D_ * 128 + 8 * D_ * 8 + 64;        \
      CK(cudaFuncSetAttribute(k_self<D_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                        \
      int occ = 0;                                                                                                         \
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_self<D_>, NCONS, smem));                                    \
      const float ms = time_it([&] { cudaMemsetAsync(d_ck, 0, lv.nodes.size() * 4); k_self<D_><<<grid, NCONS, smem>>>(A, tmap); }); \
      char nm[128]; snprintf(nm, sizeof nm, "V3 gather4 self-service D=%d (%d CTAs/SM)", D_, occ);                          \
      report(nm, ms, false);                                                                                               \
    }
    SELF(1)

#define HYB(NT_, I0_)                                                                                                      \
    {                                                                                                                      \
      const size_t smem = 1024 + HIST_WORDS * 4 + (size_t)NT_ * 32 * FP + (8 - NT_) * 16 * 34 * 4 + NT_ * 256 + NT_ * 8 + 128; \
      CK(cudaFuncSetAttribute(k_hybrid<NT_, I0_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                \
      int occ = 0;                                                                                                         \
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_hybrid<NT_, I0_>, NCONS, smem));                            \
      const float ms = time_it([&] { cudaMemsetAsync(d_ck, 0, lv.nodes.size() * 4); k_hybrid<NT_, I0_><<<grid, NCONS, smem>>>(A, tmap); }); \
      char nm[128]; snprintf(nm, sizeof nm, "V4 hybrid %d TMA + %d LSU warps%s (%d CTAs/SM)", NT_, 8 - NT_, I0_ ? ", lane-0 issue" : "", occ); \
      report(nm, ms, false);                                                                                               \
    }
    SELF(2)
    SELFM(1, 1, "FETCH ONLY") SELFM(2, 1, "FETCH ONLY") SELFM(3, 1, "FETCH ONLY")
    SELFM(1, 2, "CONSUME ONLY") SELFM(2, 2, "CONSUME ONLY")
    CK(cudaFree(d_ent)); CK(cudaFree(d_nodes)); CK(cudaFree(d_tiles)); CK(cudaFree(d_ck));
  }
  return 0;
}
