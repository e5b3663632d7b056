// Ansor's packed per-store feature blob -> flat [rows, 164] fp32 + CSR sizes, straight into pinned
// staging and on to the scoring kernels.
//
//   wire format: SerializeFeatures, src/auto_scheduler/feature.cc:1562-1622
//     int32 n; int32 sizes[n + 3];
//     per record i: sizes[i] floats = { float n_stmts; float rows[n_stmts][vec_len] }  (sizes[i] == 0: lowering failed)
//     float throughputs[sizes[n]]; int32 task_ids[sizes[n+1]]; float min_costs[sizes[n+2]]
//   parser restated: unpack_feature, python/tvm/auto_scheduler/feature.py:49-158 (a failed record becomes one
//   all-zero row, :114-116; n_stmts = int(x + 0.5), :119; the vec_len and total-length assertions, :120-127,157)
//
// The reference unpacks with struct.unpack_from per row into Python tuples, re-wraps them as float64 numpy
// arrays, and SegmentDataLoader flattens them again; here one pass copies the rows (dropping the n_stmts
// floats) into pinned memory -- the pageable -> pinned copy an H2D needs anyway -- and the chunks go to the GPU
// while the next one is being parsed.  Host code only; no kernel in this file.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

#include "../../include/vaex_b200.h"
#include "common.cuh"
#include "model.cuh"

namespace vaex {

struct BlobIndex {
  int64_t n = 0, rows = 0;
  std::vector<int64_t> src;     // float offset of record i's first row in the blob (-1: failed record)
  std::vector<int32_t> nrows;   // rows of record i (failed record: 1 zero row)
  int64_t thr_off = 0, task_off = 0, min_off = 0;   // byte offsets of the trailing arrays
  int64_t n_thr = 0, n_task = 0, n_min = 0;
};

static inline int32_t rd_i32(const uint8_t* p) { int32_t v; std::memcpy(&v, p, 4); return v; }
static inline float rd_f32(const uint8_t* p) { float v; std::memcpy(&v, p, 4); return v; }

static int blob_index(const void* blob, int64_t nbytes, int32_t vec_len, BlobIndex& ix) {
  const uint8_t* b = static_cast<const uint8_t*>(blob);
  VAEX_CHECK(blob && nbytes >= 4, "feature blob is empty");
  VAEX_CHECK(vec_len > 0, "bad feature vector length");
  const int64_t n = rd_i32(b);
  VAEX_CHECK(n >= 0 && 4 + 4 * (n + 3) <= nbytes, "feature blob: bad record count");
  ix.n = n;
  ix.src.resize((size_t)n);
  ix.nrows.resize((size_t)n);
  const uint8_t* sizes = b + 4;
  // pass 1, sequential over the contiguous size table: where every record starts
  int64_t off = 4 + 4 * (n + 3);
  int64_t n_ok = n;                       // records [0, n_ok) have a valid extent
  const char* extent_err = nullptr;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t size = rd_i32(sizes + 4 * i);
    if (size < 0) { extent_err = "feature blob: negative record size"; n_ok = i; break; }
    if (size == 0) { ix.src[i] = -1; continue; }   // failed during lowering: one all-zero row (feature.py:114-116)
    if (off + 4 * size > nbytes) { extent_err = "feature blob: record runs past the end"; n_ok = i; break; }
    ix.src[i] = (off + 4) / 4;
    off += 4 * size;
  }
  // pass 2: the n_stmts float in front of every record is one cache miss per record (records are ~10 KB apart):
  // spread over threads for big blobs.  The first bad record in record order decides the error, as in a sequential scan.
  struct Bad { int64_t i = -1; int kind = 0; int64_t got = 0; };
  auto check = [&](int64_t lo, int64_t hi, Bad& bad) {
    for (int64_t i = lo; i < hi; ++i) {
      if (ix.src[i] < 0) { ix.nrows[i] = 1; continue; }
      const int64_t size = rd_i32(sizes + 4 * i);
      const int64_t n_stmts = (int64_t)((double)rd_f32(b + 4 * (ix.src[i] - 1)) + 0.5);   // int(n_stmts[0] + 0.5), feature.py:119
      if (n_stmts <= 0) { bad = {i, 1, 0}; return; }
      const int64_t got = (size - 1) / n_stmts;
      if (got != vec_len || got * n_stmts != size - 1) { bad = {i, 2, got}; return; }
      if (n_stmts >= (1ll << 31)) { bad = {i, 3, 0}; return; }
      ix.nrows[i] = (int32_t)n_stmts;
    }
  };
  const int nt = n_ok >= 32768 ? (int)std::min<unsigned>(8, std::max<unsigned>(1, std::thread::hardware_concurrency())) : 1;
  std::vector<Bad> bads((size_t)nt);
  if (nt == 1) check(0, n_ok, bads[0]);
  else {
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t) th.emplace_back(check, n_ok * t / nt, n_ok * (t + 1) / nt, std::ref(bads[t]));
    for (auto& x : th) x.join();
  }
  for (const Bad& bad : bads) {
    if (bad.i < 0) continue;
    if (bad.kind == 2) {
      set_error("The length of feature vector is wrong. Expected " + std::to_string(vec_len) + " but got " + std::to_string(bad.got) + ".");
      return VAEX_E_INVALID;
    }
    VAEX_CHECK(bad.kind != 1, "feature blob: record without statements");
    VAEX_CHECK(bad.kind != 3, "feature blob: too many statements");
  }
  if (extent_err) { set_error(extent_err); return VAEX_E_INVALID; }
  int64_t rows = 0;
  for (int64_t i = 0; i < n; ++i) rows += ix.nrows[i];
  ix.rows = rows;
  ix.n_thr = rd_i32(sizes + 4 * n);
  ix.n_task = rd_i32(sizes + 4 * (n + 1));
  ix.n_min = rd_i32(sizes + 4 * (n + 2));
  VAEX_CHECK(ix.n_thr >= 0 && ix.n_task >= 0 && ix.n_min >= 0, "feature blob: negative trailer size");
  ix.thr_off = off;
  ix.task_off = off + 4 * ix.n_thr;
  ix.min_off = ix.task_off + 4 * ix.n_task;
  if (ix.min_off + 4 * ix.n_min != nbytes) {   // feature.py:157
    set_error(std::to_string(ix.min_off + 4 * ix.n_min) + " vs " + std::to_string(nbytes));
    return VAEX_E_INVALID;
  }
  return VAEX_OK;
}

// copies records [i0, i1) into dst (row-major [rows, vec_len]); split over threads when it is worth it
static void blob_copy(const uint8_t* b, const BlobIndex& ix, int64_t i0, int64_t i1, int32_t vec_len, float* dst,
                      const std::vector<int64_t>& row_start /* row offset of record i relative to dst, [i - i0] */) {
  const size_t row_bytes = (size_t)vec_len * 4;
  auto work = [&](int64_t a, int64_t z) {
    for (int64_t i = a; i < z; ++i) {
      float* d = dst + (size_t)row_start[i - i0] * vec_len;
      if (ix.src[i] < 0) std::memset(d, 0, row_bytes);
      else std::memcpy(d, b + 4 * ix.src[i], row_bytes * ix.nrows[i]);
    }
  };
  const int64_t total_rows = (i1 > i0) ? row_start[i1 - 1 - i0] + ix.nrows[i1 - 1] : 0;
  const size_t bytes = (size_t)total_rows * row_bytes;
  int nt = (int)std::min<size_t>(8, std::max<size_t>(1, bytes >> 23));   // one thread per 8 MiB, up to 8
  nt = (int)std::min<int64_t>(nt, std::max<unsigned>(1, std::thread::hardware_concurrency()));
  if (nt <= 1 || i1 - i0 < 2 * nt) { work(i0, i1); return; }
  std::vector<std::thread> th;
  // split by rows so the threads copy equal amounts
  int64_t a = i0;
  for (int t = 0; t < nt; ++t) {
    const int64_t want = total_rows * (t + 1) / nt;
    int64_t z = a;
    while (z < i1 && row_start[z - i0] + ix.nrows[z] <= want) ++z;
    if (t == nt - 1) z = i1;
    if (z > a) th.emplace_back(work, a, z);
    a = z;
  }
  for (auto& x : th) x.join();
}

}  // namespace vaex

using namespace vaex;

extern "C" {

int vaex_feature_blob_scan(const void* blob, int64_t nbytes, int32_t vec_len, int64_t* counts) {
  VAEX_CHECK(counts, "null counts");
  BlobIndex ix;
  TRY(blob_index(blob, nbytes, vec_len, ix));
  counts[0] = ix.n; counts[1] = ix.rows; counts[2] = ix.n_thr; counts[3] = ix.n_task; counts[4] = ix.n_min;
  return VAEX_OK;
}

int vaex_feature_blob_unpack(const void* blob, int64_t nbytes, int32_t vec_len, float* feats, int32_t* sizes,
                             uint8_t* lowered, float* throughputs, int32_t* task_ids, float* min_costs) {
  BlobIndex ix;
  TRY(blob_index(blob, nbytes, vec_len, ix));
  const uint8_t* b = static_cast<const uint8_t*>(blob);
  VAEX_CHECK(ix.n == 0 || (feats && sizes), "null output");
  std::vector<int64_t> row_start((size_t)ix.n);
  int64_t r = 0;
  for (int64_t i = 0; i < ix.n; ++i) {
    row_start[i] = r;
    r += ix.nrows[i];
    sizes[i] = ix.nrows[i];
    if (lowered) lowered[i] = ix.src[i] >= 0;
  }
  blob_copy(b, ix, 0, ix.n, vec_len, feats, row_start);
  if (throughputs && ix.n_thr) std::memcpy(throughputs, b + ix.thr_off, 4 * (size_t)ix.n_thr);
  if (task_ids && ix.n_task) std::memcpy(task_ids, b + ix.task_off, 4 * (size_t)ix.n_task);
  if (min_costs && ix.n_min) std::memcpy(min_costs, b + ix.min_off, 4 * (size_t)ix.n_min);
  return VAEX_OK;
}

// Chunk plan of both staging paths (one plan: a program cut by a 256-row tile boundary is summed in two pieces, so
// the last bit of a score depends on where its chunk starts).  Up to 64 MiB of rows: one chunk, like
// vaex_segsum_mlp_predict_host; above: <= 128 MiB per chunk and at least ~8 chunks so that staging, copy and kernels overlap.
static int64_t blob_chunk_rows(const BlobIndex& ix, int F) {
  const int64_t row_bytes = F * (int64_t)sizeof(float), total = ix.rows * row_bytes;
  const int64_t chunk = total < (64ll << 20) ? (128ll << 20) : std::min<int64_t>(128ll << 20, std::max<int64_t>(8ll << 20, total / 8));
  return std::max<int64_t>(1, chunk / row_bytes);
}

// Narrowing feeders for the blob: T worker threads turn the records of chunk 0, 1, 2, ... into fp16 rows in a ring of
// pinned buffers (each thread an equal share of the chunk's rows: records are the unit, found by bisection of the
// row prefix), the calling thread ships every chunk as soon as all T shares are in.  A chunk in which a non-zero
// value would round to an fp16 zero (|x| <= 2^-25; it would flip the "all-zero program" rule) goes the fp32 way.
static bool copy_done(cudaEvent_t e) {
  if (cudaEventQuery(e) == cudaSuccess) return true;
  (void)cudaGetLastError();   // cudaErrorNotReady is not an error of ours
  return false;
}

static int predict_blob_narrowed(vaex_model_t h, const uint8_t* b, const BlobIndex& ix, float* scores_host, int flags, int T) {
  const int F = h->F;
  const int64_t max_rows = blob_chunk_rows(ix, F);
  std::vector<int64_t> rs((size_t)ix.n + 1);           // rows before record i
  rs[0] = 0;
  for (int64_t i = 0; i < ix.n; ++i) rs[i + 1] = rs[i] + ix.nrows[i];
  struct Chunk { int64_t p0, p1; };
  std::vector<Chunk> chunks;
  size_t slot_rows = 0;
  for (int64_t p0 = 0; p0 < ix.n;) {
    int64_t p1 = p0 + 1;
    while (p1 < ix.n && rs[p1 + 1] - rs[p0] <= max_rows) ++p1;
    chunks.push_back({p0, p1});
    slot_rows = std::max<size_t>(slot_rows, (size_t)(rs[p1] - rs[p0]));
    p0 = p1;
  }
  const int NC = (int)chunks.size();
  constexpr int NS = 4;
  const size_t need = std::max<size_t>(slot_rows, (size_t)max_rows) * F * sizeof(uint16_t);
  for (int i = 0; i < NS; ++i) {
    vaex_model::PinnedSlot& pn = h->pack_stage[i];
    if (pn.bytes < need) {
      if (pn.p) cudaFreeHost(pn.p);
      pn.p = nullptr; pn.bytes = 0;
      VAEX_CUDA(cudaHostAlloc(&pn.p, need, cudaHostAllocDefault));
      pn.bytes = need;
    }
    if (!h->pack_events[i]) VAEX_CUDA(cudaEventCreateWithFlags(&h->pack_events[i], cudaEventDisableTiming));
  }
  std::atomic<int> released{0};                        // chunk c may be narrowed once c < released + NS
  std::atomic<bool> abort_flag{false};
  std::vector<std::atomic<int>> done((size_t)NC), tiny((size_t)NC);
  for (int c = 0; c < NC; ++c) { done[c].store(0); tiny[c].store(0); }
  std::vector<std::thread> workers;
  for (int t = 0; t < T; ++t)
    workers.emplace_back([&, t] {
      for (int c = 0; c < NC; ++c) {
        while (c >= released.load(std::memory_order_acquire) + NS) {
          if (abort_flag.load(std::memory_order_relaxed)) return;
          std::this_thread::sleep_for(std::chrono::microseconds(20));
        }
        if (abort_flag.load(std::memory_order_relaxed)) return;
        const Chunk& ck = chunks[c];
        const int64_t r0 = rs[ck.p0], rows = rs[ck.p1] - r0;
        auto cut = [&](int k) {                         // first record of share k
          if (k >= T) return ck.p1;
          return (int64_t)(std::lower_bound(rs.begin() + ck.p0, rs.begin() + ck.p1, r0 + rows * k / T) - rs.begin());
        };
        uint16_t* stage = static_cast<uint16_t*>(h->pack_stage[c % NS].p);
        bool lossy = false;
        for (int64_t i = cut(t), z = cut(t + 1); i < z; ++i) {
          uint16_t* d = stage + (size_t)(rs[i] - r0) * F;
          if (ix.src[i] < 0) std::memset(d, 0, (size_t)F * sizeof(uint16_t));
          else lossy |= host_pack_f16(reinterpret_cast<const float*>(b) + ix.src[i], d, (size_t)ix.nrows[i] * F);
        }
        if (lossy) tiny[c].store(1, std::memory_order_relaxed);
        done[c].fetch_add(1, std::memory_order_release);
      }
    });
  auto stop_workers = [&] {
    abort_flag.store(true);
    for (auto& w : workers) w.join();
    workers.clear();
  };
  struct Joiner { decltype(stop_workers)& f; ~Joiner() { f(); } } joiner{stop_workers};   // every early return joins

  std::vector<int32_t> off_host[2];
  std::vector<int64_t> row_start;
  int rel = 0;
  for (int c = 0; c < NC; ++c) {
    const int s = c & 1;
    cudaStream_t st = h->copy_streams[s];
    VAEX_CUDA(cudaStreamSynchronize(st));   // device slot s (chunk c-2) retired, incl. its D2H and its fp32 staging
    for (;;) {
      while (rel < c && copy_done(h->pack_events[rel % NS])) ++rel;   // ring slots whose copy is over
      released.store(rel, std::memory_order_release);
      if (done[c].load(std::memory_order_acquire) == T) break;
      std::this_thread::sleep_for(std::chrono::microseconds(20));
    }
    const Chunk& ck = chunks[c];
    const int64_t np = ck.p1 - ck.p0, r0 = rs[ck.p0], rows = rs[ck.p1] - r0;
    const bool f16 = tiny[c].load(std::memory_order_relaxed) == 0;
    off_host[s].resize((size_t)np + 1);
    for (int64_t i = 0; i <= np; ++i) off_host[s][i] = (int32_t)(rs[ck.p0 + i] - r0);
    vaex_model::HostSlot& sl = h->slots[s];
    if ((size_t)rows > sl.rows) {
      if (sl.feats) cudaFree(sl.feats);
      sl.rows = std::max<size_t>(rows, max_rows);
      VAEX_CUDA(cudaMalloc(&sl.feats, sl.rows * F * sizeof(float)));
    }
    if ((size_t)np > sl.progs) {
      if (sl.off) cudaFree(sl.off);
      if (sl.scores) cudaFree(sl.scores);
      sl.progs = std::max<size_t>(np, max_rows);
      VAEX_CUDA(cudaMalloc(&sl.off, (sl.progs + 1) * sizeof(int32_t)));
      VAEX_CUDA(cudaMalloc(&sl.scores, sl.progs * sizeof(float)));
    }
    if (f16) {
      VAEX_CUDA(cudaMemcpyAsync(sl.feats, h->pack_stage[c % NS].p, (size_t)rows * F * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
    } else {
      vaex_model::PinnedSlot& pn = h->pinned[s];
      const size_t bytes = (size_t)rows * F * sizeof(float);
      if (bytes > pn.bytes) {
        if (pn.p) cudaFreeHost(pn.p);
        pn.p = nullptr; pn.bytes = 0;
        VAEX_CUDA(cudaHostAlloc(&pn.p, bytes, cudaHostAllocDefault));
        pn.bytes = bytes;
      }
      row_start.resize((size_t)np);
      for (int64_t i = 0; i < np; ++i) row_start[i] = rs[ck.p0 + i] - r0;
      blob_copy(b, ix, ck.p0, ck.p1, F, static_cast<float*>(pn.p), row_start);
      VAEX_CUDA(cudaMemcpyAsync(sl.feats, pn.p, bytes, cudaMemcpyHostToDevice, st));
    }
    VAEX_CUDA(cudaEventRecord(h->pack_events[c % NS], st));   // ring slot c % NS is free once this copy is over
    VAEX_CUDA(cudaMemcpyAsync(sl.off, off_host[s].data(), (size_t)(np + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (c > 0) VAEX_CUDA(cudaStreamWaitEvent(st, h->events[(c - 1) & 1], 0));   // the model's workspaces are shared
    TRY(vaex_segsum_mlp_forward(h, sl.feats, sl.off, np, rows, sl.scores, flags | (f16 ? VAEX_F_FEATS_F16 : 0), st));
    VAEX_CUDA(cudaEventRecord(h->events[c & 1], st));
    VAEX_CUDA(cudaMemcpyAsync(scores_host + ck.p0, sl.scores, (size_t)np * sizeof(float), cudaMemcpyDeviceToHost, st));
    h->pack_stats[f16 ? 0 : 1] += 1;
  }
  VAEX_CUDA(cudaStreamSynchronize(h->copy_streams[0]));
  VAEX_CUDA(cudaStreamSynchronize(h->copy_streams[1]));
  return VAEX_OK;
}

int vaex_segsum_mlp_predict_blob(vaex_model_t h, const void* blob, int64_t nbytes, float* scores_host,
                                 int64_t n_scores, int flags) {
  VAEX_CHECK(h && h->kind == VAEX_KIND_SEGSUM_MLP && h->packed, "bad handle");
  BlobIndex ix;
  const int F = h->F;
  TRY(blob_index(blob, nbytes, F, ix));
  VAEX_CHECK(n_scores == ix.n, "scores buffer does not match the blob's record count");
  if (ix.n == 0) return VAEX_OK;
  VAEX_CHECK(scores_host, "null scores");
  const uint8_t* b = static_cast<const uint8_t*>(blob);
  const int64_t max_rows = blob_chunk_rows(ix, F);
  for (int i = 0; i < 2; ++i) if (!h->copy_streams[i]) VAEX_CUDA(cudaStreamCreateWithFlags(&h->copy_streams[i], cudaStreamNonBlocking));
  for (int i = 0; i < 4; ++i) if (!h->events[i]) VAEX_CUDA(cudaEventCreateWithFlags(&h->events[i], cudaEventDisableTiming));
  {
    // big blobs on a host with spare AVX2/F16C cores: narrow to fp16 while staging (same rule as vaex_segsum_mlp_predict_host)
    const int hw = (int)std::thread::hardware_concurrency();
    int T = std::max(1, std::min(hw - 1, 32));
    if (const char* e = getenv("VAEX_PACK_THREADS")) if (atoi(e) > 0) T = std::min(atoi(e), 64);
    const bool pack = !(flags & (VAEX_F_FORCE_SIMT | VAEX_F_NO_HOST_PACK | VAEX_F_FEATS_F16)) && h->tc.ready && F % 4 == 0 &&
                      hw >= 4 && ix.rows * F * (int64_t)sizeof(float) >= (64ll << 20) && host_pack_supported();
    if (pack) return predict_blob_narrowed(h, b, ix, scores_host, flags, T);
  }
  std::vector<int64_t> row_start;
  int64_t p0 = 0;
  int it = 0;
  while (p0 < ix.n) {
    const int s = it & 1;
    cudaStream_t st = h->copy_streams[s];
    VAEX_CUDA(cudaStreamSynchronize(st));   // slot s (chunk it-2) retired: its staging buffers are free again
    int64_t p1 = p0, rows = 0;
    row_start.clear();
    while (p1 < ix.n && (rows == 0 || rows + ix.nrows[p1] <= max_rows)) {
      row_start.push_back(rows);
      rows += ix.nrows[p1];
      ++p1;
    }
    const int64_t np = p1 - p0;
    vaex_model::HostSlot& sl = h->slots[s];
    if ((size_t)rows > sl.rows) {
      if (sl.feats) cudaFree(sl.feats);
      sl.rows = std::max<size_t>(rows, max_rows);
      VAEX_CUDA(cudaMalloc(&sl.feats, sl.rows * F * sizeof(float)));
    }
    if ((size_t)np > sl.progs) {
      if (sl.off) cudaFree(sl.off);
      if (sl.scores) cudaFree(sl.scores);
      sl.progs = std::max<size_t>(np, max_rows);
      VAEX_CUDA(cudaMalloc(&sl.off, (sl.progs + 1) * sizeof(int32_t)));
      VAEX_CUDA(cudaMalloc(&sl.scores, sl.progs * sizeof(float)));
    }
    vaex_model::PinnedSlot& pn = h->pinned[s];
    const size_t need_bytes = (size_t)rows * F * sizeof(float) + (size_t)(np + 1) * sizeof(int32_t);
    if (need_bytes > pn.bytes) {
      if (pn.p) cudaFreeHost(pn.p);
      pn.p = nullptr;
      pn.bytes = std::max<size_t>(need_bytes, std::min<size_t>((size_t)max_rows * F * sizeof(float) + (1 << 20), need_bytes * 2));
      VAEX_CUDA(cudaHostAlloc(&pn.p, pn.bytes, cudaHostAllocDefault));
    }
    float* stage = static_cast<float*>(pn.p);
    int32_t* off_stage = reinterpret_cast<int32_t*>(stage + (size_t)rows * F);
    blob_copy(b, ix, p0, p1, F, stage, row_start);
    for (int64_t i = 0; i < np; ++i) off_stage[i] = (int32_t)row_start[i];
    off_stage[np] = (int32_t)rows;
    VAEX_CUDA(cudaMemcpyAsync(sl.feats, stage, (size_t)rows * F * sizeof(float), cudaMemcpyHostToDevice, st));
    VAEX_CUDA(cudaMemcpyAsync(sl.off, off_stage, (size_t)(np + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (it > 0) VAEX_CUDA(cudaStreamWaitEvent(st, h->events[(it - 1) & 1], 0));   // the model's workspaces are shared
    TRY(vaex_segsum_mlp_forward(h, sl.feats, sl.off, np, rows, sl.scores, flags, st));
    VAEX_CUDA(cudaEventRecord(h->events[it & 1], st));
    VAEX_CUDA(cudaMemcpyAsync(scores_host + p0, sl.scores, (size_t)np * sizeof(float), cudaMemcpyDeviceToHost, st));
    p0 = p1;
    ++it;
  }
  VAEX_CUDA(cudaStreamSynchronize(h->copy_streams[0]));
  VAEX_CUDA(cudaStreamSynchronize(h->copy_streams[1]));
  return VAEX_OK;
}

}  // extern "C"
