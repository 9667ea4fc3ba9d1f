// All-gather of a small message over NVLink peer memory in ONE kernel: push + signal + wait + unpack.
// The two exchanges of the sequence-sharded decode step (top-k candidates, (o, lse) partials) are
// 16-32 KB per rank -- pure latency for a collective library (two NCCL all-gathers cost ~20 us per layer at 2 GPUs,
// profiles/r01_s15_long_context_multi_gpu.md).  Here every rank's kernel
//   1. stores its message straight into every peer's staging buffer (symmetric memory mapped over NVLink / NVSwitch:
//      16-byte stores, one CTA per peer -- the message is far below the bandwidth-delay product of a link),
//   2. publishes it: CTA barrier, then one cumulative st.release.sys of this step's epoch into the peer's flag word
//      for this rank,
//   3. waits (ld.acquire.sys on its OWN flag words, so the polling never crosses the link) until every rank's epoch
//      arrived, bounded by a wall-clock limit so that a lost peer is an error code, not a hung GPU,
//   4. unpacks the staging area into a private, fixed-address buffer in the layout the consumer wants
//      ([rows][world][row_bytes]: the gathered candidate scores come out as [head][rank][c], ready for quest_topk).
// The epoch lives in device memory and is advanced by the kernel itself, so the call can sit in a CUDA graph; staging
// is double-buffered on the epoch's parity: a rank can only finish exchange e+1 after every peer has entered it, i.e.
// after every peer has left exchange e, so pushes of e+2 never meet readers of e.
#include "common.cuh"

namespace quest {

constexpr int kPeerMax = 8;
constexpr int kPeerThreads = 512;

struct PeerTable {
  uint4* staging[kPeerMax];     // rank p's staging area as mapped into this process: [2][world][msg_vecs]
  uint32_t* flags[kPeerMax];    // rank p's flag words: [world]
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// One CTA per peer: CTA p pushes this rank's message into rank p's staging area, publishes it to p, waits for p's
// message of the same epoch and unpacks it -- the pairwise exchanges are independent, so the remote stores, the fences
// and the waits of the world-1 links overlap instead of queueing in one CTA.  state[0] counts timed-out waits,
// state[1 + p] is the epoch of the exchange with rank p (advanced here, so the launch can sit in a CUDA graph).
__global__ void __launch_bounds__(kPeerThreads)
peer_allgather_kernel(const uint4* __restrict__ src, uint32_t rows, uint32_t row_vecs, const PeerTable tbl,
                      uint32_t* __restrict__ state, uint4* __restrict__ dst, uint32_t world, uint32_t rank,
                      unsigned long long timeout_ns, unsigned int* __restrict__ err_word) {
  __shared__ uint32_t s_epoch, s_lost;
  const uint32_t tid = threadIdx.x, p = blockIdx.x;
  pdl_wait();  // the message comes from the kernel before this one
  if (tid == 0) {
    s_epoch = state[1 + p] + 1;
    state[1 + p] = s_epoch;
    s_lost = 0;
  }
  __syncthreads();
  const uint32_t epoch = s_epoch, par = epoch & 1u;
  const uint32_t msg_vecs = rows * row_vecs;
  // push: my slot in rank p's staging area
  uint4* out = tbl.staging[p] + ((size_t)par * world + rank) * msg_vecs;
  for (uint32_t j = tid; j < msg_vecs; j += kPeerThreads) out[j] = src[j];
  // publish: the CTA barrier orders every thread's stores before thread 0, whose system-scope release store is
  // cumulative over them (the semaphore pattern: barrier, then one st.release) -- no per-thread fence
  __syncthreads();
  if (tid == 0) {
    st_release_sys(tbl.flags[p] + rank, epoch);
    // wait for rank p's message of this epoch (flag words are monotonic; compare modulo 2^32)
    const uint32_t* f = tbl.flags[rank] + p;
    const unsigned long long t0 = global_ns();
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if ((++spins & 1023u) == 0 && global_ns() - t0 > timeout_ns) {
        // rank p never showed up: count it, raise the sticky host-visible error (quest_device_error) and poison p's
        // slot below, so the consumer gets NaNs instead of whatever the staging area held
        atomicAdd(state, 1u);
        device_error_raise(err_word, QUEST_DEVERR_PEER);
        s_lost = 1;
        break;
      }
    }
  }
  __syncthreads();
  // unpack rank p's slot of my staging area: dst[row][p][vec] = staging[par][p][row][vec]
  const uint4* mine = tbl.staging[rank] + ((size_t)par * world + p) * msg_vecs;
  const bool lost = s_lost != 0;
  for (uint32_t j = tid; j < msg_vecs; j += kPeerThreads) {
    const uint32_t row = j / row_vecs, vec = j - row * row_vecs;
    // written by another GPU: read at L2, never a stale L1 line.  All-ones is a NaN as fp16 and as fp32.
    dst[((size_t)row * world + p) * row_vecs + vec] = lost ? make_uint4(~0u, ~0u, ~0u, ~0u) : __ldcg(mine + j);
  }
}

}  // namespace quest

extern "C" int quest_peer_allgather(const void* src, uint32_t rows, uint32_t row_bytes, void* const* peer_staging,
                                    void* const* peer_flags, void* state, void* dst, uint32_t world, uint32_t rank,
                                    uint32_t timeout_ms, void* stream) {
  if (!src || !peer_staging || !peer_flags || !state || !dst || !rows || !row_bytes) return QUEST_ERR_BAD_ARG;
  if (world < 1 || world > quest::kPeerMax || rank >= world) return QUEST_ERR_UNSUPPORTED;
  if (row_bytes % 16 || ((uintptr_t)src & 15) || ((uintptr_t)dst & 15)) return QUEST_ERR_BAD_ARG;
  if (quest::device_error_peek() & QUEST_DEVERR_PEER) return QUEST_ERR_DEVICE;  // an earlier exchange lost a rank
  quest::PeerTable tbl = {};
  for (uint32_t p = 0; p < world; ++p) {
    if (!peer_staging[p] || !peer_flags[p] || ((uintptr_t)peer_staging[p] & 15)) return QUEST_ERR_BAD_ARG;
    tbl.staging[p] = (uint4*)peer_staging[p];
    tbl.flags[p] = (uint32_t*)peer_flags[p];
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(world);
  cfg.blockDim = dim3(quest::kPeerThreads);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const unsigned long long timeout_ns = (unsigned long long)(timeout_ms ? timeout_ms : 2000u) * 1000000ull;
  cudaError_t e = cudaLaunchKernelEx(&cfg, quest::peer_allgather_kernel, (const uint4*)src, rows, row_bytes / 16, tbl,
                                     (uint32_t*)state, (uint4*)dst, world, rank, timeout_ns, quest::device_error_word());
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

extern "C" size_t quest_peer_staging_bytes(uint32_t rows, uint32_t row_bytes, uint32_t world) {
  return (size_t)2 * world * rows * row_bytes;
}
