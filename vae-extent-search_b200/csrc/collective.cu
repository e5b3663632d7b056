// Multi-GPU entry points of the C ABI: the candidate merge of the sharded scoring path, the
// data-parallel gradient all-reduce and a plain all-gather for the sharded latent ascent.  One process per GPU; the
// communicator is NCCL over NVLink / NVSwitch.  NCCL is bound at run time (dlopen of libnccl.so.2, the copy torch has
// already mapped when the caller is a PyTorch process), so the library itself carries no link-time dependency and still
// loads on hosts without NCCL; world == 1 needs no NCCL at all.
//
// The candidate merge is the path's only exchange: every rank owns a contiguous block of programs and contributes its
// k best (score, rank-local index) pairs, already ordered (score desc, index asc) by vaex_topk.
//   pack   : k records of 16 B  {score, valid, global index}                         (1 launch)
//   gather : one ncclAllGather of 16 k bytes per rank  (<= 16 KiB: pure latency)
//   merge  : ONE CTA sorts the world*k records by the unique key (score desc, gathered position asc) -- position order
//            equals global index order because ranks own ascending blocks -- and writes the k best   (1 launch)
// It replaces shard.merge_topk's two all-gathers + ~12 eager torch ops + a second vaex_topk.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <mutex>

#include "common.cuh"

namespace vaex {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mutex;

static int nccl_api(NcclApi** out) {
  std::lock_guard<std::mutex> lock(g_nccl_mutex);
  if (!g_nccl.ok) {
    const char* override_path = getenv("VAEX_NCCL_LIB");
    void* h = override_path ? dlopen(override_path, RTLD_NOW | RTLD_GLOBAL) : nullptr;
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { set_error(std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror()); return VAEX_E_UNSUPPORTED; }
    g_nccl.handle = h;
#define VAEX_NCCL_SYM(field, name)                                                         \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));               \
    if (!g_nccl.field) { set_error(std::string("NCCL symbol missing: ") + name); return VAEX_E_UNSUPPORTED; }
    VAEX_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    VAEX_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    VAEX_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    VAEX_NCCL_SYM(AllGather, "ncclAllGather")
    VAEX_NCCL_SYM(AllReduce, "ncclAllReduce")
    VAEX_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef VAEX_NCCL_SYM
    g_nccl.ok = true;
  }
  *out = &g_nccl;
  return VAEX_OK;
}

#define VAEX_NCCL(api, expr)                                                                \
  do {                                                                                      \
    ncclResult_t r_ = (expr);                                                               \
    if (r_ != ncclSuccess) {                                                                \
      ::vaex::set_error(std::string(#expr) + ": " + (api)->GetErrorString(r_));             \
      return VAEX_E_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

constexpr int MERGE_MAX_K = 1024, MERGE_MAX_RECORDS = 8192, MERGE_THREADS = 1024;

struct __align__(16) MergeRec {
  float score;
  int32_t valid;
  int64_t gidx;
};

}  // namespace vaex

struct vaex_comm {
  int rank = 0, world = 1, device = 0;
  ncclComm_t comm = nullptr;
  vaex::NcclApi* api = nullptr;
  vaex::MergeRec* send = nullptr;    // [MERGE_MAX_K]
  vaex::MergeRec* recv = nullptr;    // [world * MERGE_MAX_K]
};

namespace vaex {

__device__ __forceinline__ unsigned int ordered_u32(float s) {
  unsigned int u = __float_as_uint(s);
  if (s != s) return 0u;                                   // NaN sorts last (as in topk.cu)
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void merge_pack_kernel(const float* __restrict__ vals, const int64_t* __restrict__ idx,
                                  const int32_t* __restrict__ n_local, int k, int64_t base, MergeRec* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= k) return;
  const int n = n_local ? min(*n_local, k) : k;
  MergeRec r;
  r.valid = t < n;
  r.score = r.valid ? vals[t] : 0.f;
  r.gidx = r.valid ? idx[t] + base : -1;
  out[t] = r;
}

// world * k records, rank-major.  key = ordered(score) << 32 | ~position: unique, so the bitonic network yields exactly
// (score desc, position asc); an invalid record gets key 0 and sorts behind every valid one (a valid record's low word
// ~position is never 0 for position < 2^32 - 1, and a valid NaN keeps its high word 0 but a non-zero low word).
__global__ void __launch_bounds__(MERGE_THREADS, 1)
merge_topk_kernel(const MergeRec* __restrict__ recs, int total, int k, float* __restrict__ out_vals,
                  int64_t* __restrict__ out_idx, int32_t* __restrict__ n_out) {
  extern __shared__ unsigned long long keys[];             // [S]
  __shared__ int n_valid;
  int S = 2;
  while (S < total) S <<= 1;
  if (threadIdx.x == 0) n_valid = 0;
  __syncthreads();
  int mine = 0;
  for (int i = threadIdx.x; i < S; i += MERGE_THREADS) {
    unsigned long long key = 0ull;
    if (i < total && recs[i].valid) {
      key = ((unsigned long long)ordered_u32(recs[i].score) << 32) | (unsigned int)(~(unsigned int)i);
      ++mine;
    }
    keys[i] = key;
  }
  if (mine) atomicAdd(&n_valid, mine);
  __syncthreads();
  for (int size = 2; size <= S; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < S; t += MERGE_THREADS) {
        const int j = t ^ stride;
        if (j > t) {
          const bool desc = (t & size) == 0;
          const unsigned long long x = keys[t], y = keys[j];
          if (desc ? (x < y) : (x > y)) { keys[t] = y; keys[j] = x; }
        }
      }
      __syncthreads();
    }
  }
  const int m = min(n_valid, k);
  for (int t = threadIdx.x; t < k; t += MERGE_THREADS) {
    if (t < m) {
      const unsigned int pos = ~(unsigned int)(keys[t] & 0xffffffffull);
      out_vals[t] = recs[pos].score;
      out_idx[t] = recs[pos].gidx;
    } else {
      out_vals[t] = -INFINITY;
      out_idx[t] = -1;
    }
  }
  if (threadIdx.x == 0 && n_out) *n_out = m;
}

}  // namespace vaex

extern "C" {

int vaex_nccl_unique_id(uint8_t* id_out_host) {
  using namespace vaex;
  VAEX_CHECK(id_out_host, "null id buffer");
  NcclApi* api = nullptr;
  TRY(nccl_api(&api));
  ncclUniqueId id;
  VAEX_NCCL(api, api->GetUniqueId(&id));
  static_assert(sizeof(id) == VAEX_NCCL_ID_BYTES, "ncclUniqueId size");
  memcpy(id_out_host, &id, sizeof(id));
  return VAEX_OK;
}

int vaex_nccl_init(int32_t rank, int32_t world, const uint8_t* id_host, vaex_comm_t* out) {
  using namespace vaex;
  VAEX_CHECK(out && world >= 1 && rank >= 0 && rank < world, "bad rank / world");
  VAEX_CHECK(world * 1 <= MERGE_MAX_RECORDS, "world too large");
  vaex_comm* c = new vaex_comm();
  c->rank = rank;
  c->world = world;
  cudaError_t e = cudaGetDevice(&c->device);
  if (e != cudaSuccess) { delete c; set_error(std::string("cudaGetDevice: ") + cudaGetErrorString(e)); return VAEX_E_CUDA; }
  if (world > 1) {
    if (!id_host) { delete c; set_error("world > 1 needs the unique id of rank 0 (vaex_nccl_unique_id)"); return VAEX_E_INVALID; }
    int rc = nccl_api(&c->api);
    if (rc != VAEX_OK) { delete c; return rc; }
    ncclUniqueId id;
    memcpy(&id, id_host, sizeof(id));
    ncclResult_t r = c->api->CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) { set_error(std::string("ncclCommInitRank: ") + c->api->GetErrorString(r)); delete c; return VAEX_E_CUDA; }
  }
  if (cudaMalloc(&c->send, sizeof(MergeRec) * MERGE_MAX_K) != cudaSuccess ||
      cudaMalloc(&c->recv, sizeof(MergeRec) * MERGE_MAX_K * (size_t)world) != cudaSuccess) {
    set_error("cudaMalloc of the merge buffers failed");
    if (c->send) cudaFree(c->send);
    if (c->comm) c->api->CommDestroy(c->comm);
    delete c;
    return VAEX_E_NOMEM;
  }
  cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MERGE_MAX_RECORDS * 8);
  *out = c;
  return VAEX_OK;
}

int vaex_nccl_destroy(vaex_comm_t c) {
  if (!c) return VAEX_OK;
  if (c->send) cudaFree(c->send);
  if (c->recv) cudaFree(c->recv);
  if (c->comm) c->api->CommDestroy(c->comm);
  delete c;
  return VAEX_OK;
}

int vaex_comm_info(vaex_comm_t c, int32_t* rank, int32_t* world) {
  VAEX_CHECK(c, "null communicator");
  if (rank) *rank = c->rank;
  if (world) *world = c->world;
  return VAEX_OK;
}

int vaex_topk_allgather_merge(vaex_comm_t c, const float* local_vals_dev, const int64_t* local_idx_dev,
                              const int32_t* n_local_dev_or_null, int32_t k, int64_t index_base, float* out_vals_dev,
                              int64_t* out_idx_dev, int32_t* n_out_dev_or_null, void* stream) {
  using namespace vaex;
  VAEX_CHECK(c, "null communicator");
  VAEX_CHECK(k >= 0 && k <= MERGE_MAX_K, "k must be in [0, 1024]");
  VAEX_CHECK((int64_t)k * c->world <= MERGE_MAX_RECORDS, "world * k must be <= 8192");
  VAEX_CHECK(k == 0 || (local_vals_dev && local_idx_dev && out_vals_dev && out_idx_dev), "null argument");
  cudaStream_t st = (cudaStream_t)stream;
  if (k == 0) {
    if (n_out_dev_or_null) VAEX_CUDA(cudaMemsetAsync(n_out_dev_or_null, 0, sizeof(int32_t), st));
    return VAEX_OK;
  }
  MergeRec* gathered = c->world > 1 ? c->recv : c->send;
  merge_pack_kernel<<<(unsigned)cdiv(k, 256), 256, 0, st>>>(local_vals_dev, local_idx_dev, n_local_dev_or_null, k, index_base,
                                                          c->send);
  VAEX_LAUNCH_CHECK();
  if (c->world > 1)
    VAEX_NCCL(c->api, c->api->AllGather(c->send, c->recv, sizeof(MergeRec) * (size_t)k, ncclChar, c->comm, st));
  const int total = k * c->world;
  int S = 2;
  while (S < total) S <<= 1;
  merge_topk_kernel<<<1, MERGE_THREADS, (size_t)S * 8, st>>>(gathered, total, k, out_vals_dev, out_idx_dev, n_out_dev_or_null);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

int vaex_allreduce_grads(vaex_comm_t c, float* grads_dev, int64_t n, void* stream) {
  using namespace vaex;
  VAEX_CHECK(c && n >= 0 && (grads_dev || n == 0), "bad argument");
  if (c->world == 1 || n == 0) return VAEX_OK;
  VAEX_NCCL(c->api, c->api->AllReduce(grads_dev, grads_dev, (size_t)n, ncclFloat, ncclSum, c->comm, (cudaStream_t)stream));
  return VAEX_OK;
}

int vaex_allgather(vaex_comm_t c, const void* send_dev, void* recv_dev, int64_t bytes_per_rank, void* stream) {
  using namespace vaex;
  VAEX_CHECK(c && bytes_per_rank >= 0 && (bytes_per_rank == 0 || (send_dev && recv_dev)), "bad argument");
  if (bytes_per_rank == 0) return VAEX_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (c->world == 1) {
    if (send_dev != recv_dev) VAEX_CUDA(cudaMemcpyAsync(recv_dev, send_dev, (size_t)bytes_per_rank, cudaMemcpyDeviceToDevice, st));
    return VAEX_OK;
  }
  VAEX_NCCL(c->api, c->api->AllGather(send_dev, recv_dev, (size_t)bytes_per_rank, ncclChar, c->comm, st));
  return VAEX_OK;
}

}  // extern "C"
