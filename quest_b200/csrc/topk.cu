// Per-head top-k page selector: one CTA per head, scores resident in shared memory as order-preserving
// 16-bit keys, two 8-bit MSB radix-select passes, then a stable (ascending position) compaction.
// Replaces decode_select_k -> raft radix_topk_one_block_kernel<half,int,8,512>
// (reference kernels/include/topk/decode_select_k.cuh:25-62; raft select_radix.cuh:82-104 key twiddle,
// :380-399 choose_bucket, :404-440 last_filter).  Unlike raft's atomicAdd race among ties, the tie rule
// here is deterministic: lowest position first.
#include "common.cuh"

namespace quest {

constexpr int kTopkThreads = 1024;
constexpr uint32_t kTopkMaxLen = 96 * 1024;  // 192 KB of u16 keys

__device__ __forceinline__ uint16_t key_of(uint16_t h) { return (h & 0x8000u) ? (uint16_t)~h : (uint16_t)(h | 0x8000u); }
__device__ __forceinline__ uint16_t half_of(uint16_t k) { return (k & 0x8000u) ? (uint16_t)(k & 0x7fffu) : (uint16_t)~k; }

// Warp 0: find the bucket where the descending cumulative count crosses k.  res[0]=bucket, res[1]=k left.
__device__ __forceinline__ void choose_bucket(const uint32_t* hist, uint32_t k, uint32_t* res) {
  const uint32_t lane = threadIdx.x;
  uint32_t c[8], tot = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    c[i] = hist[lane * 8 + i];
    tot += c[i];
  }
  uint32_t suf = tot;  // inclusive suffix sum over lanes
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t t = __shfl_down_sync(0xffffffffu, suf, off);
    if (lane + off < 32) suf += t;
  }
  uint32_t above = suf - tot;
#pragma unroll
  for (int i = 7; i >= 0; --i) {
    if (above < k && above + c[i] >= k) {
      res[0] = lane * 8 + i;
      res[1] = k - above;
    }
    above += c[i];
  }
}

__global__ void __launch_bounds__(kTopkThreads)
topk_kernel(const uint16_t* __restrict__ scores, const int32_t* __restrict__ in_idx, uint32_t in_stride,
            uint16_t* __restrict__ out_val, int32_t* __restrict__ out_idx, uint32_t len, uint32_t k) {
  extern __shared__ uint16_t keys[];  // [len]
  __shared__ uint32_t hist[256];
  __shared__ uint32_t res[2];
  __shared__ uint32_t wsum_gt[32], wsum_eq[32];
  const uint32_t head = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint16_t* s = scores + (size_t)head * len;

  pdl_trigger();
  pdl_wait();
  if (tid < 256) hist[tid] = 0;
  __syncthreads();
  // Pass 1: load + histogram of the high byte.  Warp-aggregated to survive skewed score distributions.
  const uint32_t rounds = (len + kTopkThreads - 1) / kTopkThreads;
  for (uint32_t r = 0; r < rounds; ++r) {
    const uint32_t i = r * kTopkThreads + tid;
    uint32_t bin = 256;
    if (i < len) {
      const uint16_t key = key_of(s[i]);
      keys[i] = key;
      bin = key >> 8;
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, bin);
    if (bin < 256 && lane == (uint32_t)(__ffs(peers) - 1)) atomicAdd(&hist[bin], __popc(peers));
  }
  __syncthreads();
  if (warp == 0) choose_bucket(hist, k, res);
  __syncthreads();
  const uint32_t b1 = res[0], k1 = res[1];
  if (tid < 256) hist[tid] = 0;
  __syncthreads();
  // Pass 2: low byte among keys whose high byte is b1.
  for (uint32_t r = 0; r < rounds; ++r) {
    const uint32_t i = r * kTopkThreads + tid;
    uint32_t bin = 256;
    if (i < len) {
      const uint16_t key = keys[i];
      if ((uint32_t)(key >> 8) == b1) bin = key & 0xff;
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, bin);
    if (bin < 256 && lane == (uint32_t)(__ffs(peers) - 1)) atomicAdd(&hist[bin], __popc(peers));
  }
  __syncthreads();
  if (warp == 0) choose_bucket(hist, k1, res);
  __syncthreads();
  const uint32_t tau = (b1 << 8) | res[0];
  const uint32_t need = res[1];  // number of keys == tau to take, lowest position first

  // Stable compaction: thread t owns positions [t*per, (t+1)*per).
  const uint32_t per = rounds;
  const uint32_t lo = min(tid * per, len), hi = min(lo + per, len);
  uint32_t ngt = 0, neq = 0;
  for (uint32_t i = lo; i < hi; ++i) {
    const uint32_t key = keys[i];
    ngt += key > tau;
    neq += key == tau;
  }
  uint32_t sgt = ngt, seq = neq;  // inclusive warp scans
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, sgt, off);
    const uint32_t b = __shfl_up_sync(0xffffffffu, seq, off);
    if (lane >= (uint32_t)off) {
      sgt += a;
      seq += b;
    }
  }
  if (lane == 31) {
    wsum_gt[warp] = sgt;
    wsum_eq[warp] = seq;
  }
  __syncthreads();
  if (warp == 0) {
    uint32_t a = wsum_gt[lane], b = wsum_eq[lane];
    const uint32_t a0 = a, b0 = b;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, a, off);
      const uint32_t y = __shfl_up_sync(0xffffffffu, b, off);
      if (lane >= (uint32_t)off) {
        a += x;
        b += y;
      }
    }
    wsum_gt[lane] = a - a0;  // exclusive
    wsum_eq[lane] = b - b0;
  }
  __syncthreads();
  uint32_t gt_before = wsum_gt[warp] + sgt - ngt;
  uint32_t eq_before = wsum_eq[warp] + seq - neq;
  const int32_t* idx_row = in_idx ? in_idx + (size_t)head * in_stride : nullptr;
  for (uint32_t i = lo; i < hi; ++i) {
    const uint32_t key = keys[i];
    const bool gt = key > tau, eq = key == tau;
    if (gt || (eq && eq_before < need)) {
      const uint32_t pos = gt_before + min(eq_before, need);
      out_idx[(size_t)head * k + pos] = idx_row ? __ldg(idx_row + i) : (int32_t)i;
      if (out_val) out_val[(size_t)head * k + pos] = half_of((uint16_t)key);
    }
    gt_before += gt;
    eq_before += eq;
  }
}

}  // namespace quest

extern "C" uint32_t quest_topk_max_len(void) { return quest::kTopkMaxLen; }

extern "C" int quest_topk(const void* scores, const int32_t* in_idx, uint32_t in_idx_row_stride, void* out_val,
                          int32_t* out_idx, uint32_t num_heads, uint32_t len, uint32_t k, void* stream) {
  if (!scores || !out_idx || !num_heads) return QUEST_ERR_BAD_ARG;
  if (k == 0) return QUEST_OK;
  if (k > len) return QUEST_ERR_BAD_ARG;
  if (len > quest::kTopkMaxLen) return QUEST_ERR_UNSUPPORTED;
  const size_t smem = ((size_t)len * sizeof(uint16_t) + 15) & ~(size_t)15;
  if (smem > 40 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(quest::topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_heads);
  cfg.blockDim = dim3(quest::kTopkThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, quest::topk_kernel, (const uint16_t*)scores, in_idx, in_idx_row_stride,
                                     (uint16_t*)out_val, out_idx, len, k);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}
