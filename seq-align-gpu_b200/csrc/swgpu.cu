// swgpu.cu -- implementation of the C-ABI in include/swgpu.h: database handle, search orchestration.
// Thin extern "C" layer over the kernels in sw_k1s.cuh (16-bit systolic), sw_kernels.cuh (K3, 32-bit)
// and sw_k3x.cuh (32-bit wavefront).  No CPU fallback: every compute entry point needs a CUDA device.
#include "swgpu_internal.h"
#include "db_pack.h"
#include "sw_kernels.cuh"
#include "sw_k1s.cuh"
#include "sw_k3x.cuh"
#include "sw_k3s.cuh"

// the K1S kernels themselves are instantiated in sw_k1s_inst.cu (one object per gap specialisation and ptxas option set)

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include <map>
#include <mutex>

// ---- guard zones (swgpu_internal.h) ----
namespace swi {
namespace {
constexpr size_t kGuardBytes = 512;
struct GuardRec { size_t bytes; int device; };
std::mutex& guard_mu() { static std::mutex m; return m; }
std::map<void*, GuardRec>& guard_map() { static std::map<void*, GuardRec> m; return m; }
bool guard_on() { static const bool on = [] { const char* e = getenv("SWGPU_GUARD"); return e && atoi(e) > 0; }(); return on; }
}  // namespace

cudaError_t guard_malloc(void** p, size_t bytes) {
  if (!guard_on()) return (cudaMalloc)(p, bytes);
  cudaError_t e = (cudaMalloc)(p, bytes + kGuardBytes);
  if (e != cudaSuccess) return e;
  e = cudaMemset((char*)*p + bytes, 0xA5, kGuardBytes);
  if (e != cudaSuccess) return e;
  int dev = 0; cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(guard_mu());
  guard_map()[*p] = GuardRec{bytes, dev};
  return cudaSuccess;
}

cudaError_t guard_free(void* p) {
  if (guard_on() && p) { std::lock_guard<std::mutex> lk(guard_mu()); guard_map().erase(p); }
  return (cudaFree)(p);
}

int guard_check() {
  if (!guard_on()) return -1;
  std::lock_guard<std::mutex> lk(guard_mu());
  int cur = 0; cudaGetDevice(&cur);
  int bad = 0;
  unsigned char tail[kGuardBytes];
  for (const auto& kv : guard_map()) {
    cudaSetDevice(kv.second.device);
    cudaDeviceSynchronize();
    if (cudaMemcpy(tail, (const char*)kv.first + kv.second.bytes, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) { ++bad; continue; }
    size_t first = kGuardBytes;
    for (size_t i = 0; i < kGuardBytes; ++i) if (tail[i] != 0xA5) { first = i; break; }
    if (first < kGuardBytes) {
      ++bad;
      fprintf(stderr, "SWGPU_GUARD: allocation %p of %zu bytes on device %d was written past its end (first damaged byte at +%zu)\n",
              kv.first, kv.second.bytes, kv.second.device, first);
    }
  }
  cudaSetDevice(cur);
  return bad;
}
}  // namespace swi

extern "C" int swgpu_guard_check(void) { return swi::guard_check(); }

// Proof that the check sees what it is for: one byte written past the end of a guarded allocation must be reported.
// Returns 1 when exactly that allocation was reported, 0 otherwise, -1 when the mode is off.
extern "C" int swgpu_guard_selftest(void) {
  if (swi::guard_check() < 0) return -1;
  const int before = swi::guard_check();
  void* p = nullptr;
  if (swi::guard_malloc(&p, 1000) != cudaSuccess) return 0;
  cudaMemset((char*)p + 1000, 0, 1);                       // the overrun
  const int during = swi::guard_check();
  swi::guard_free(p);
  const int after = swi::guard_check();
  return (during == before + 1 && after == before) ? 1 : 0;
}

namespace {

using swi::fail;

constexpr int kDefaultStreamCols = 2304;
constexpr int kDefaultTaper = 3;             // stream lengths halve at the last 1/16, 1/64 and 1/256 of the database (scripts/ab_taper.py)
constexpr int kWaveMinLen = 4096;          // 32-bit path: targets at least this long take k3x (one warp per 256-row stripe)

template <typename T>
int dalloc(swgpu_db* db, T** p, size_t count) {
  const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
  CU(cudaMalloc((void**)p, bytes));
  db->device_bytes += (int64_t)bytes;
  return SWGPU_OK;
}

}  // namespace

namespace swi {

int slot_prepare(swgpu_db* db, swgpu_slot& S) {
  if (S.ready) return SWGPU_OK;
  int lo = 0, hi = 0, r;
  CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));            // lo = least urgent, hi = most urgent (numerically smaller)
  CU(cudaStreamCreateWithPriority(&S.s_main, cudaStreamNonBlocking, lo));
  CU(cudaStreamCreateWithPriority(&S.s_aux, cudaStreamNonBlocking, hi));
  for (auto& e : S.ev) CU(cudaEventCreate(&e));
  if ((r = dalloc(db, &S.d_scores, (size_t)db->n))) return r;
  if ((r = dalloc(db, &S.d_raw_s, (size_t)db->n_pairs * 32))) return r;
  if ((r = dalloc(db, &S.d_rescore_list, (size_t)db->n))) return r;
  if ((r = dalloc(db, &S.d_row_bound, (size_t)db->n))) return r;
  if ((r = dalloc(db, &S.d_counters, 16 + 256))) return r;           // 16 counters + the per-SM CTA counters of k3_intra's placement
  if ((r = dalloc(db, &S.d_table, 1024))) return r;
  if ((r = dalloc(db, &S.d_bnd_s, (size_t)db->k1s_grid * swk::K1S_WARPS * (size_t)db->bnd_s_stride))) return r;
  if ((r = dalloc(db, &S.d_bnd3, (size_t)db->k3_grid * swk::K3_WARPS * (size_t)db->bnd3_stride))) return r;
  if (db->max_len >= kWaveMinLen) {
    if ((r = dalloc(db, &S.d_x_rows, (size_t)swk::K3X_SLOTS * 2 * (size_t)db->x_row_stride))) return r;
    CU(cudaMemset(S.d_x_rows, 0, sizeof(int4) * (size_t)swk::K3X_SLOTS * 2 * (size_t)db->x_row_stride));
    if ((r = dalloc(db, &S.d_x_done, (size_t)swk::K3X_SLOTS + 256))) return r;      // + the per-SM CTA counters of k3x
  }
  CU(cudaMallocHost((void**)&S.h_ctr, 16 * sizeof(unsigned)));
  S.ready = true;
  return SWGPU_OK;
}

// k3x slot state for a query of lq rows: the per-stripe partial results grow with the query; the completion
// counters are zeroed on `stream` and the tag epoch advances before every k3x launch.
int k3x_prepare(swgpu_db* db, swgpu_slot& S, int lq, cudaStream_t stream) {
  const int ns = std::max(1, (lq + swk::K3X_ROWS - 1) / swk::K3X_ROWS);
  if (!S.d_x_rows) {                                       // top-k coordinates on a database without long targets
    CU(cudaMalloc((void**)&S.d_x_rows, sizeof(int4) * (size_t)swk::K3X_SLOTS * 2 * (size_t)db->x_row_stride));
    CU(cudaMemsetAsync(S.d_x_rows, 0, sizeof(int4) * (size_t)swk::K3X_SLOTS * 2 * (size_t)db->x_row_stride, stream));
    CU(cudaMalloc((void**)&S.d_x_done, sizeof(unsigned) * (swk::K3X_SLOTS + 256)));
  }
  if (ns > S.x_stripes_cap) {
    cudaFree(S.d_x_partial); S.d_x_partial = nullptr;
    S.x_stripes_cap = ns + ns / 2;
    CU(cudaMalloc((void**)&S.d_x_partial, sizeof(int4) * (size_t)swk::K3X_SLOTS * S.x_stripes_cap));
  }
  CU(cudaMemsetAsync(S.d_x_done, 0, sizeof(unsigned) * (swk::K3X_SLOTS + 256), stream));
  ++S.x_epoch;
  return SWGPU_OK;
}

void slot_release(swgpu_slot& S) {
  cudaFree(S.d_scores); cudaFree(S.d_raw_s); cudaFree(S.d_rescore_list); cudaFree(S.d_row_bound); cudaFree(S.d_counters);
  cudaFree(S.d_bnd_s); cudaFree(S.d_bnd3); cudaFree(S.d_bnd3s); cudaFree(S.d_query); cudaFree(S.d_table);
  cudaFree(S.d_sprof); cudaFree(S.d_sprof2);
  cudaFree(S.d_keys); cudaFree(S.d_hit_sidx); cudaFree(S.d_sel); cudaFree(S.d_coords);
  cudaFree(S.d_xs_bnd); cudaFree(S.d_xs_progress); cudaFree(S.d_xs_raw);
  cudaFree(S.d_x_rows); cudaFree(S.d_x_partial); cudaFree(S.d_x_done);
  cudaFreeHost(S.h_ctr); cudaFreeHost(S.h_keys); cudaFreeHost(S.h_coords); cudaFreeHost(S.h_best); cudaFreeHost(S.h_sidx);
  cudaFreeHost(S.h_scores);
  for (auto& e : S.ev) if (e) cudaEventDestroy(e);
  if (S.s_main) cudaStreamDestroy(S.s_main);
  if (S.s_aux) cudaStreamDestroy(S.s_aux);
  S = swgpu_slot();
}

}  // namespace swi

namespace {

// Host -> device copy of a large pageable array through two pinned staging buffers: the host memcpy into one buffer
// overlaps the DMA out of the other (north_star (1): "pinned, HBM-resident batches"; a plain cudaMemcpy from pageable
// memory stages through the driver's own buffer synchronously, chunk after chunk).
static int upload_staged(void* dst, const void* src, size_t bytes) {
  constexpr size_t CHUNK = 16u << 20;
  if (bytes <= CHUNK) { CU(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice)); return SWGPU_OK; }
  void* stage[2] = {nullptr, nullptr}; cudaEvent_t ev[2] = {nullptr, nullptr}; cudaStream_t st = nullptr;
  auto body = [&]() -> int {
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) { CU(cudaMallocHost(&stage[i], CHUNK)); CU(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming)); }
    int b = 0;
    for (size_t o = 0; o < bytes; o += CHUNK, b ^= 1) {
      const size_t n = std::min(CHUNK, bytes - o);
      CU(cudaEventSynchronize(ev[b]));                       // the DMA that last used this buffer is done (no-op the first time)
      memcpy(stage[b], (const char*)src + o, n);
      CU(cudaMemcpyAsync((char*)dst + o, stage[b], n, cudaMemcpyHostToDevice, st));
      CU(cudaEventRecord(ev[b], st));
    }
    CU(cudaStreamSynchronize(st));
    return SWGPU_OK;
  };
  const int rc = body();
  for (int i = 0; i < 2; ++i) { if (stage[i]) cudaFreeHost(stage[i]); if (ev[i]) cudaEventDestroy(ev[i]); }
  if (st) cudaStreamDestroy(st);
  return rc;
}

// Host tables + device buffers of a packed database.  With pk.plain / pk.scols present they are uploaded; without
// (layout only, swgpu_db_create_synthetic) the two buffers are allocated and `fill` generates them on the device.
template <typename Fill>
static int db_from_pack(swgpu_db_t** out, int device, swpack_t& pk, int64_t entry_base, const swgpu_pack_opts_t* opts, Fill fill) {
  swgpu_db* db = new swgpu_db();
  db->device = device;
  db->n = pk.n_targets; db->n_residues = pk.n_residues; db->max_len = pk.max_len; db->entry_base = entry_base;
  db->long_threshold = (opts && opts->long_threshold > 0) ? opts->long_threshold : 0;
  db->h_len_sorted_prefix.assign(pk.plain_off, pk.plain_off + pk.n_targets + 1);
  db->h_inv_order.resize((size_t)pk.n_targets);
  for (int64_t s = 0; s < pk.n_targets; ++s) db->h_inv_order[(size_t)pk.order[s]] = (int32_t)s;
  db->n_pairs = pk.n_pairs; db->n_streams = pk.n_streams; db->max_stream_cols = pk.max_stream_cols;
  db->n_long_streams = pk.n_long_streams; db->n_long_pairs = pk.n_long_pairs;
  db->scols_len = (int64_t)pk.scols_len;
  db->long_cols = pk.n_long_streams < pk.n_streams ? pk.stream_off[pk.n_long_streams] : (int64_t)pk.scols_len;
  // A stream is a serial chain of columns run by ONE warp.  Sharing its scheduler with the bulk warps, that warp
  // takes ~7 cycles per instruction, the whole GPU retires 592 x 0.73 per cycle: a stream longer than
  // scols_len / 3000 columns finishes after everything else (measured on C2: the 35 213-column stream added
  // 8-25 % to 2 launches out of 3, profiles/README.md).  Streams past a third of that get a warp of their own.
  {
    const int64_t crit = std::max<int64_t>(SWPACK_LONG_STREAM_COLS, (int64_t)pk.scols_len / 9000);
    int64_t nv = 0;
    while (nv < pk.n_long_streams && pk.stream_ncols[nv] > crit) ++nv;
    db->n_vlong_streams = nv;
    db->n_vlong_pairs = nv < pk.n_streams ? pk.stream_pair0[nv] : pk.n_pairs;
  }

  auto body = [&]() -> int {
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    db->sm_count = prop.multiProcessorCount;
    for (auto& e : db->ev_batch) CU(cudaEventCreate(&e));
    int r;
    if ((r = dalloc(db, &db->d_plain, (size_t)pk.n_residues))) return r;
    if ((r = dalloc(db, &db->d_plain_off, (size_t)pk.n_targets + 1))) return r;
    if ((r = dalloc(db, &db->d_order, (size_t)pk.n_targets))) return r;
    if ((r = dalloc(db, &db->d_scols, (size_t)pk.scols_len + 64))) return r;
    if ((r = dalloc(db, &db->d_stream_off, (size_t)pk.n_streams + 1))) return r;
    if ((r = dalloc(db, &db->d_stream_ncols, (size_t)pk.n_streams))) return r;
    if ((r = dalloc(db, &db->d_stream_pair0, (size_t)pk.n_streams))) return r;
    CU(cudaMemcpy(db->d_plain_off, pk.plain_off, sizeof(int64_t) * ((size_t)pk.n_targets + 1), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(db->d_order, pk.order, sizeof(int64_t) * (size_t)pk.n_targets, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(db->d_stream_off, pk.stream_off, sizeof(int64_t) * ((size_t)pk.n_streams + 1), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(db->d_stream_ncols, pk.stream_ncols, sizeof(int32_t) * (size_t)pk.n_streams, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(db->d_stream_pair0, pk.stream_pair0, sizeof(int64_t) * (size_t)pk.n_streams, cudaMemcpyHostToDevice));
    if (pk.plain && pk.scols) {
      if ((r = upload_staged(db->d_plain, pk.plain, (size_t)pk.n_residues))) return r;
      if ((r = upload_staged(db->d_scols, pk.scols, sizeof(uint16_t) * (size_t)pk.scols_len))) return r;
    } else if ((r = fill(db))) {
      return r;
    }
    // persistent grids: k1s two CTAs of 8 warps per SM; k3 at most four CTAs of 4 warps per SM
    db->k1s_grid = db->sm_count * 3;                          // upper bound of resident k1s CTAs (carry-row slots)
    db->bnd_s_stride = ((long long)std::max<int32_t>(pk.max_stream_cols, 64) + 63) & ~63ll;
    db->k3_grid = db->sm_count * 4;
    db->bnd3_stride = std::max<int32_t>(std::min<int32_t>(pk.max_len, kWaveMinLen), 32);   // k3_intra takes targets < kWaveMinLen
    db->x_row_stride = ((long long)std::max<int32_t>(pk.max_len, 32) + 31) & ~31ll;
    return swi::slot_prepare(db, db->slots[0]);
  };
  const int rc = body();
  swpack_free(&pk);
  if (rc != SWGPU_OK) { swgpu_db_destroy(db); return rc; }
  *out = db;
  return SWGPU_OK;
}

// shorter streams at the end of the database (db_pack.c::stream_limit); opts->no_taper or SWGPU_PACK_TAPER=0 turn it off
static int pack_taper(const swgpu_pack_opts_t* opts) {
  if (opts && opts->no_taper) return 0;
  if (const char* ev = getenv("SWGPU_PACK_TAPER")) return std::max(0, std::min(8, atoi(ev)));   // number of halvings (profiling knob)
  return kDefaultTaper;
}

// Stream length when the caller does not give one: kDefaultStreamCols for a Swiss-Prot-sized database and up, shorter
// for smaller ones so that the work queue still holds ~10 multi-stripe items (8 streams each) per resident CTA
// (measured at 0.3 x C2, Lq = 2005: 24.1 ms with 2304-column streams, 23.1 ms with 1152; profiles/r1g_stream_cols_small_db.txt).
static int default_stream_cols(int64_t n_residues) {
  const int64_t want_streams = 8 * 296 * 10;
  int64_t c = (n_residues / 2) / want_streams;
  c = (c + 127) / 128 * 128;
  return (int)std::max<int64_t>(1152, std::min<int64_t>(kDefaultStreamCols, c));
}

static int pick_device(int device) {
  int ndev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) return fail(SWGPU_ENODEV, "no CUDA device");
  if (device < 0 || device >= ndev) return fail(SWGPU_EINVAL, "device %d out of range (%d devices)", device, ndev);
  CU(cudaSetDevice(device));
  return SWGPU_OK;
}

}  // namespace

extern "C" {

int swgpu_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

const char* swgpu_last_error(void) { return swi::err_buf(); }
const char* swgpu_version(void) { return "swgpu 0.2 (sm_100a)"; }

void swgpu_db_destroy(swgpu_db_t* db) {
  if (!db) return;
  cudaSetDevice(db->device);
  cudaDeviceSynchronize();
  if (swi::guard_check() > 0) { fprintf(stderr, "SWGPU_GUARD: damaged guard zones at swgpu_db_destroy\n"); abort(); }   // debug mode only
  cudaFree(db->d_plain); cudaFree(db->d_plain_off); cudaFree(db->d_order);
  cudaFree(db->d_scols); cudaFree(db->d_stream_off); cudaFree(db->d_stream_ncols); cudaFree(db->d_stream_pair0);
  for (auto& S : db->slots) swi::slot_release(S);
  for (auto& e : db->ev_batch) if (e) cudaEventDestroy(e);
  delete db;
}

int swgpu_db_create(swgpu_db_t** out, int device, const int8_t* codes, const int64_t* offsets,
                    int64_t n_targets, int64_t entry_base, const swgpu_pack_opts_t* opts) {
  if (!out || n_targets < 0 || (n_targets > 0 && (!codes || !offsets))) return fail(SWGPU_EINVAL, "swgpu_db_create: bad argument");
  if (n_targets > 0x7ffffff0ll) return fail(SWGPU_EINVAL, "swgpu_db_create: more than 2^31 targets per GPU; shard the database");
  *out = nullptr;
  if (int r = pick_device(device)) return r;
  const int stream_cols = (opts && opts->stream_cols > 0) ? opts->stream_cols
                                                          : default_stream_cols(n_targets ? offsets[n_targets] - offsets[0] : 0);
  swpack_t pk;
  static const int64_t zero_off[1] = {0};
  if (swpack_build(&pk, codes, n_targets ? offsets : zero_off, n_targets, stream_cols, pack_taper(opts)) != 0)
    return fail(SWGPU_EINVAL, "database packing failed (residue code outside 0..31, negative length, or out of memory)");
  return db_from_pack(out, device, pk, entry_base, opts, [](swgpu_db*) { return (int)SWGPU_OK; });
}

int swgpu_db_create_synthetic(swgpu_db_t** out, int device, const int32_t* lengths, const int64_t* entry_ids,
                              int64_t n_targets, int64_t entry_base, const swgpu_synth_t* gen, const swgpu_pack_opts_t* opts) {
  if (!out || n_targets < 0 || (n_targets > 0 && !lengths) || !gen) return fail(SWGPU_EINVAL, "swgpu_db_create_synthetic: bad argument");
  if (n_targets > 0x7ffffff0ll) return fail(SWGPU_EINVAL, "swgpu_db_create_synthetic: more than 2^31 targets per GPU; shard the database");
  if (gen->n_letters < 1 || gen->n_letters > 32) return fail(SWGPU_EINVAL, "swgpu_db_create_synthetic: n_letters must be 1..32");
  for (int i = 0; i < gen->n_letters; ++i) {
    if (gen->codes[i] < 1 || gen->codes[i] > 31) return fail(SWGPU_EINVAL, "swgpu_db_create_synthetic: residue code outside 1..31");
    if (gen->cdf[i] > 65536u || (i && gen->cdf[i] < gen->cdf[i - 1])) return fail(SWGPU_EINVAL, "swgpu_db_create_synthetic: cdf must be non-decreasing, <= 65536");
  }
  if (gen->cdf[gen->n_letters - 1] != 65536u) return fail(SWGPU_EINVAL, "swgpu_db_create_synthetic: cdf must end at 65536");
  *out = nullptr;
  if (int r = pick_device(device)) return r;
  int64_t total_len = 0;
  for (int64_t t = 0; t < n_targets; ++t) total_len += lengths[t] > 0 ? lengths[t] : 0;
  const int stream_cols = (opts && opts->stream_cols > 0) ? opts->stream_cols : default_stream_cols(total_len);
  swpack_t pk;
  std::vector<int64_t> pair_col((size_t)(n_targets + 1) / 2 + 1);
  static const int32_t zero_len[1] = {0};
  if (swpack_layout(&pk, n_targets ? lengths : zero_len, n_targets, stream_cols, pack_taper(opts), pair_col.data()) != 0)
    return fail(SWGPU_EINVAL, "database layout failed (negative length or out of memory)");
  auto fill = [&](swgpu_db* db) -> int {
    // the residues exist only on the device: generated straight into `plain` (sorted order), then paired into streams
    int64_t *d_ids = nullptr, *d_pair_col = nullptr;
    auto body = [&]() -> int {
      if (entry_ids) {
        CU(cudaMalloc((void**)&d_ids, sizeof(int64_t) * (size_t)std::max<int64_t>(n_targets, 1)));
        CU(cudaMemcpy(d_ids, entry_ids, sizeof(int64_t) * (size_t)n_targets, cudaMemcpyHostToDevice));
      }
      CU(cudaMalloc((void**)&d_pair_col, sizeof(int64_t) * pair_col.size()));
      CU(cudaMemcpy(d_pair_col, pair_col.data(), sizeof(int64_t) * pair_col.size(), cudaMemcpyHostToDevice));
      CU(cudaMemset(db->d_scols, 0, sizeof(uint16_t) * ((size_t)pk.scols_len + 64)));
      CU(swi::synth_fill(db, *gen, d_ids, entry_base, d_pair_col));
      CU(cudaDeviceSynchronize());
      return SWGPU_OK;
    };
    const int r = body();
    cudaFree(d_ids); cudaFree(d_pair_col);
    return r;
  };
  return db_from_pack(out, device, pk, entry_base, opts, fill);
}

int swgpu_db_fetch_target(swgpu_db_t* db, int64_t local_entry, int8_t* out, int64_t cap, int64_t* len) {
  if (!db || local_entry < 0 || local_entry >= db->n || !len) return fail(SWGPU_EINVAL, "swgpu_db_fetch_target: bad argument");
  const size_t s = (size_t)db->h_inv_order[(size_t)local_entry];
  const int64_t b = db->h_len_sorted_prefix[s], L = db->h_len_sorted_prefix[s + 1] - b;
  *len = L;
  if (!out) return SWGPU_OK;                       // length query
  if (cap < L) return fail(SWGPU_EINVAL, "swgpu_db_fetch_target: buffer too small (%lld needed)", (long long)L);
  CU(cudaSetDevice(db->device));
  CU(cudaMemcpy(out, db->d_plain + b, (size_t)L, cudaMemcpyDeviceToHost));
  return SWGPU_OK;
}

int64_t swgpu_db_num_targets(const swgpu_db_t* db) { return db ? db->n : 0; }
int64_t swgpu_db_num_residues(const swgpu_db_t* db) { return db ? db->n_residues : 0; }
int64_t swgpu_db_device_bytes(const swgpu_db_t* db) { return db ? db->device_bytes : 0; }

}  // extern "C"

// ------------------------------------------------------------------------------------------
// search
// ------------------------------------------------------------------------------------------
namespace {

// Instantiated kernels: every R in 8..32 for g = 8, 16, 32; multi-stripe (g = 32) with R in {16, 20, 24..32}.
bool k1s_has_plan(int g, int R, bool multi) {
  if (R < 8 || R > 32 || (g != 8 && g != 16 && g != 32)) return false;
  return !multi || (g == 32 && (R == 16 || R == 20 || R >= 24));
}

// K1S plan: lanes per target pair (g), query rows per lane (R), stripes (npass).
// Cost model: padded rows x (cell work + per-step overhead / R); few work items -> prefer more lanes per pair.
struct K1SPlan { int g, g_log2, R, npass; };

// Measured deviations of single instantiations from the smooth model (scripts/tune_plan.py on the C2 database,
// profiles/r1f_tune_plan_odd.jsonl): the planner may take one or two rows per lane more than needed to avoid them.
double k1s_penalty(int g, int R) {
  if (R == 21 && g < 32) return 1.09;              // unpaired profile loads, 95 registers (g = 32: no deviation with the round-2 kernel)
  if (R == 11) return 1.06;
  if (R == 8 && g == 8) return 1.25;
  if (R == 25 && g < 32) return 1.05;
  if (R == 24 && g == 32) return 1.05;
  if (R == 19 || R == 18) return 1.01;
  return 1.0;
}

K1SPlan plan_k1s(int64_t lq, int64_t n_streams, int warp_slots) {
  K1SPlan best{32, 5, 32, 1}; double best_cost = -1;
  for (int gl = 3; gl <= 5; ++gl) {
    const int g = 1 << gl;
    const int R0 = std::max((int)((lq + g - 1) / g), 8);
    for (int R = R0; R <= std::min(R0 + 2, 32); ++R) {
      double cost = (double)g * (9.0 * R + 16.0) * k1s_penalty(g, R);
      const double items = (double)n_streams * g / 32.0;
      if (items < 2.0 * warp_slots) cost *= (2.0 * warp_slots) / std::max(items, 1.0);
      if (best_cost < 0 || cost <= best_cost) { best_cost = cost; best = K1SPlan{g, gl, R, 1}; }
    }
  }
  if (best_cost < 0) {
    // multi-stripe: measured time per padded query row on the C2 database (scripts/tune_plan.py, us per row; round-2 kernels:
    // profiles/r2c_tune_plan.jsonl, r2w_ab_nested_per_R.txt).  A database too small to fill the GPU runs every stream as X items on one CTA per SM, a
    // latency regime with its own ranking (profiles/r1f_c4_plans.txt): it keeps the table measured there.
    struct MultiCost { int R; double us_per_row; };
    static const MultiCost kBulk[] = {{16, 39.9}, {20, 35.5}, {24, 34.11}, {25, 34.06}, {26, 33.46}, {27, 33.85},
                                      {28, 33.68}, {29, 33.33}, {30, 33.9}, {31, 33.06}, {32, 32.86}};   // profiles/r2w2_ab_nested_x_slot.txt
    static const MultiCost kFew[] = {{16, 44.2}, {20, 39.7}, {24, 37.63}, {25, 37.96}, {26, 36.94}, {27, 36.85},
                                     {28, 36.3}, {29, 36.55}, {30, 36.8}, {31, 37.8}, {32, 36.72}};
    const bool few = n_streams < 2 * (int64_t)warp_slots;
    const MultiCost (&kMulti)[11] = few ? kFew : kBulk;
    for (const auto& m : kMulti) {
      const int np = (int)((lq + 32 * m.R - 1) / (32 * m.R));
      const double cost = (double)np * 32 * m.R * m.us_per_row;
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = K1SPlan{32, 5, m.R, np}; }
    }
  }
  if (const char* ev = getenv("SWGPU_K1S_PLAN")) {               // "g,R": profiling override of the plan
    int g = 0, R = 0;
    if (sscanf(ev, "%d,%d", &g, &R) == 2 && (g == 8 || g == 16 || g == 32) && R >= 8 && R <= 32) {
      const int gl = g == 8 ? 3 : g == 16 ? 4 : 5;
      const int np = (int)((lq + (int64_t)g * R - 1) / ((int64_t)g * R));
      if (k1s_has_plan(g, R, np > 1)) best = K1SPlan{g, gl, R, np};
    }
  }
  return best;
}

template <int GAPSEL>
cudaError_t launch_k1s_sel(int g, int R, bool multi, const swk::K1SParams& P, cudaStream_t stream, int grid, int device) {
  if (multi && R >= 24) return swk::launch_k1s_o1<GAPSEL>(R, P, stream, grid, device);
  return swk::launch_k1s_o3<GAPSEL>(g, R, multi, P, stream, grid, device);
}

// oe / ge select the gap specialisation of the kernel (immediates instead of registers, sw_k1s.cuh); anything else runs
// the generic kernel.  SWGPU_K1S_GENERIC=1 forces the generic kernel (A/B and test knob).
int launch_k1s(swgpu_db* db, const swk::K1SParams& P, int g, int R, bool multi, int oe, int ge, cudaStream_t stream, int grid) {
  if (!k1s_has_plan(g, R, multi)) return fail(SWGPU_EINVAL, "no k1s instantiation for g=%d R=%d multi=%d", g, R, (int)multi);
  int sel = swk::k1s_gapsel(oe, ge);
  if (sel) { const char* ev = getenv("SWGPU_K1S_GENERIC"); if (ev && atoi(ev) > 0) sel = 0; }
  cudaError_t e;
  switch (sel) {
    case 1:  e = launch_k1s_sel<1>(g, R, multi, P, stream, grid, db->device); break;
    case 2:  e = launch_k1s_sel<2>(g, R, multi, P, stream, grid, db->device); break;
    default: e = launch_k1s_sel<0>(g, R, multi, P, stream, grid, db->device); break;
  }
  if (e != cudaSuccess) return fail(SWGPU_ECUDA, "k1s launch (g=%d R=%d multi=%d gapsel=%d) failed: %s", g, R, (int)multi, sel, cudaGetErrorString(e));
  return SWGPU_OK;
}

// 32-bit scoring of a work list (device-side count) or of all targets (list == nullptr): k3_intra (one warp per
// target) takes the short targets, the wavefront kernel k3x (one warp per stripe, many warps per target) the rest (wave_min below).
int launch_32bit(swgpu_db* db, swgpu_slot& S, cudaStream_t stream, const int* list, const unsigned* list_count,
                 int n_all, int lq, int oe, int ge, unsigned* q_k3, unsigned* q_k3x, int* launches) {
  swk::K3Params P; memset(&P, 0, sizeof(P));
  P.worklist = list; P.work_count = list_count; P.n_work = n_all; P.counter = q_k3;
  P.plain = db->d_plain; P.plain_off = db->d_plain_off; P.query = S.d_query; P.lq = lq; P.table = S.d_table;
  P.oe = oe; P.ge = ge; P.bnd = S.d_bnd3; P.bnd_stride = db->bnd3_stride; P.order = db->d_order;
  // Who takes a target: k3_intra runs a target's stripes one after the other on ONE warp (npass x lt column steps in a row),
  // k3x runs them as a pipeline on npass warps (lt steps + 6 us per stripe).  The longer the query, the shorter the targets
  // that are worth the pipeline: targets whose serial chain would exceed kWaveMinLen steps go to k3x (Lq <= 256: one stripe,
  // lt >= 4096 as before; Lq = 6 000: 24 stripes, lt >= 171).  SWGPU_WAVE_MIN=<len> overrides it (A/B knob).
  const int npass8 = std::max(1, (lq + swk::K3X_ROWS - 1) / swk::K3X_ROWS);
  int wave_min = std::max(128, kWaveMinLen / npass8);
  if (const char* ev = getenv("SWGPU_WAVE_MIN")) wave_min = std::max(32, std::min(kWaveMinLen, atoi(ev)));
  P.scores_file = S.d_scores; P.one = 1u; P.max_len = wave_min - 1;
  P.sm_slots = S.d_counters + 16; P.level_ctas = db->sm_count;
  if (lq <= 128) swk::k3_intra<4><<<db->k3_grid, swk::K3_WARPS * 32, 0, stream>>>(P);
  else           swk::k3_intra<8><<<db->k3_grid, swk::K3_WARPS * 32, 0, stream>>>(P);
  CU(cudaGetLastError());
  ++*launches;
  if (db->max_len >= wave_min) {
    int r = swi::k3x_prepare(db, S, lq, stream);
    if (r) return r;
    swk::K3XParams W; memset(&W, 0, sizeof(W));
    W.sidx_list = list; W.work_count = list_count; W.n_targets = n_all; W.counter = q_k3x;
    W.plain = db->d_plain; W.plain_off = db->d_plain_off; W.query = S.d_query; W.lq = lq; W.nstripes_max = (lq + swk::K3X_ROWS - 1) / swk::K3X_ROWS;
    W.table = S.d_table; W.oe = oe; W.ge = ge; W.rows = S.d_x_rows; W.row_stride = db->x_row_stride;
    W.epoch = S.x_epoch; W.done = S.d_x_done; W.partial = S.d_x_partial; W.sm_slots = S.d_x_done + swk::K3X_SLOTS; W.level_ctas = db->sm_count;
    W.order = db->d_order; W.scores_file = S.d_scores; W.error_flag = S.d_counters + 6; W.min_len = wave_min;
    if ((double)db->n * W.nstripes_max >= 4.0e9) return fail(SWGPU_EINVAL, "query too long for this database size (32-bit path)");
#ifdef K3X_TRACE
    int4* d_trace = nullptr;
    const size_t n_trace = 6 * (size_t)(n_all > 0 ? n_all : 8) * W.nstripes_max;     // device-side list: room for 8 targets
    if (getenv("SWGPU_K3X_TRACE")) { CU(cudaMalloc((void**)&d_trace, sizeof(int4) * n_trace)); CU(cudaMemset(d_trace, 0, sizeof(int4) * n_trace)); }
    W.trace = d_trace;
#endif
    swk::k3x<false><<<db->sm_count * swk::k3x_ctas_per_sm<false>(db->device), swk::K3X_WARPS * 32, 0, stream>>>(W);
    CU(cudaGetLastError());
    ++*launches;
#ifdef K3X_TRACE
    if (d_trace) {                                            // profiling build: (begin, end, waiting; 0.1 us) and SM of every item, appended as text
      CU(cudaStreamSynchronize(stream));
      std::vector<int4> h(n_trace);
      CU(cudaMemcpy(h.data(), d_trace, sizeof(int4) * n_trace, cudaMemcpyDeviceToHost));
      cudaFree(d_trace);
      if (FILE* f = fopen(getenv("SWGPU_K3X_TRACE"), "a")) {
        fprintf(f, "# lq %d stripes %d items %zu\n", lq, W.nstripes_max, n_trace);
        for (size_t i = 0; i + 5 < n_trace; i += 6) { fprintf(f, "%zu", i / 6); for (int q = 0; q < 6; ++q) fprintf(f, " %d %d %d %d", h[i + q].x, h[i + q].y, h[i + q].z, h[i + q].w); fprintf(f, "\n"); }
        fclose(f);
      }
    }
#endif
  }
  return SWGPU_OK;
}

// A positive gap score: the reference's three-state recurrence (sw_k3s.cuh), every target through one 32-bit kernel.
// The two-gap-state form of the other kernels is only score-equivalent for gap_open, gap_extend <= 0.
bool needs_three_state(int gap_open, int gap_extend) { return gap_open > 0 || gap_extend > 0 || gap_open + gap_extend > 0; }

int launch_three_state(swgpu_db* db, swgpu_slot& S, cudaStream_t stream, int lq, int oe, int ge, unsigned* queue, int* launches) {
  const int grid = db->sm_count;                                     // 4 warps per CTA, one carry row of the longest target each
  const int64_t stride = ((int64_t)std::max<int32_t>(db->max_len, 32) + 31) & ~31ll;
  const int64_t need = (int64_t)grid * swk::K3_WARPS * stride;
  if (need > S.bnd3s_cap) {
    cudaFree(S.d_bnd3s); S.d_bnd3s = nullptr; S.bnd3s_cap = need;
    CU(cudaMalloc((void**)&S.d_bnd3s, sizeof(int2) * (size_t)need));
  }
  swk::K3Params P; memset(&P, 0, sizeof(P));
  P.worklist = nullptr; P.work_count = nullptr; P.n_work = (int)db->n; P.counter = queue;
  P.plain = db->d_plain; P.plain_off = db->d_plain_off; P.query = S.d_query; P.lq = lq; P.table = S.d_table;
  P.oe = oe; P.ge = ge; P.bnd = S.d_bnd3s; P.bnd_stride = stride; P.order = db->d_order;
  P.scores_file = S.d_scores; P.one = 1u; P.max_len = 0;
  if (lq <= 128) swk::k3_three<4><<<grid, swk::K3_WARPS * 32, 0, stream>>>(P);
  else           swk::k3_three<8><<<grid, swk::K3_WARPS * 32, 0, stream>>>(P);
  CU(cudaGetLastError());
  ++*launches;
  return SWGPU_OK;
}

int check_args(swgpu_db* db, const int8_t* query, int64_t lq, const int8_t* table, int gap_open, int gap_extend) {
  if (!db) return fail(SWGPU_EINVAL, "null database handle");
  if (lq < 0 || (lq > 0 && !query) || !table) return fail(SWGPU_EINVAL, "bad query/table argument");
  if (lq > 0x7fffffff) return fail(SWGPU_EINVAL, "query too long");
  const int oe = gap_open + gap_extend, ge = gap_extend;
  if (oe < -100000 || ge < -100000 || oe > 100000 || ge > 100000) return fail(SWGPU_EINVAL, "gap score out of range");
  for (int64_t i = 0; i < lq; ++i) if (query[i] < 0 || query[i] > 31) return fail(SWGPU_EINVAL, "query code outside 0..31");
  return SWGPU_OK;
}

// Queue one search on a slot.  Nothing here waits for the GPU: the bulk kernel goes to the slot's low-priority
// stream, the long-stream launch and everything after the 16-bit pass (rescoring, later top-k and read-back) to its
// high-priority stream.  The slot must be idle (collect_search() done for its previous search).
int enqueue_search(swgpu_db* db, swgpu_slot& S, const int8_t* query, int64_t lq, const int8_t* table, int gap_open,
                   int gap_extend) {
  if (S.in_flight) return fail(SWGPU_EINVAL, "internal: slot busy");
  const int oe = gap_open + gap_extend, ge = gap_extend;
  swgpu_stats_t& st = S.st; memset(&st, 0, sizeof(st));
  S.have_scores = false; S.topk_k = 0; S.user_scores = nullptr; S.scores_staged = false;
  cudaEvent_t* e = S.ev;
  cudaStream_t sm = S.s_main, sa = S.s_aux;

  // ---- uploads (query, table) ----
  if (lq > S.query_cap) {
    if (S.d_query) { cudaFree(S.d_query); S.d_query = nullptr; }
    S.query_cap = std::max<int64_t>(lq, 1024);
    CU(cudaMalloc((void**)&S.d_query, (size_t)S.query_cap));
  }
  CU(cudaEventRecord(e[6], sm));
  if (lq > 0) CU(cudaMemcpyAsync(S.d_query, query, (size_t)lq, cudaMemcpyHostToDevice, sm));
  CU(cudaMemcpyAsync(S.d_table, table, 1024, cudaMemcpyHostToDevice, sm));
  CU(cudaEventRecord(e[7], sm));

  // ---- plan ----
  // 16-bit path: biased unsigned arithmetic with B = 128 + |oe| + |ge| (sw_k1s.cuh); huge gap penalties go 32-bit.
  const unsigned B = swk::k1s_bias(oe, ge);
  const bool three = needs_three_state(gap_open, gap_extend);
  const bool use16 = !three && (B <= 16384u) && db->n_streams > 0 && lq > 0;
  const K1SPlan plan = plan_k1s(std::max<int64_t>(lq, 1), db->n_streams, db->k1s_grid * swk::K1S_WARPS);
  // Long streams (one target pair of thousands of columns) are latency-bound on their serial column chain: when the
  // bulk plan uses fewer than 32 lanes per pair they get their own launch with all 32 lanes and the fewest rows per
  // lane, one stream per CTA, on the high-priority CUDA stream.
  // With 32 lanes per pair already, only the streams that would outlast the bulk launch (n_vlong_streams) are split off.
  const int64_t n_split_streams = plan.g < 32 ? db->n_long_streams : db->n_vlong_streams;
  const int64_t n_split_pairs = plan.g < 32 ? db->n_long_pairs : db->n_vlong_pairs;
  const bool split_long = use16 && plan.npass == 1 && n_split_streams > 0;
  K1SPlan plan_long = plan;
  if (split_long) {
    int R = (int)((lq + 31) / 32); if (R < 8) R = 8;
    plan_long = K1SPlan{32, 5, R, 1};
  }
  auto words_of = [](const K1SPlan& p) { return (size_t)p.npass * (size_t)swk::sprof_words(p.R, p.npass > 1); };
  if (use16) {
    st.tile_rows = plan.R; st.lanes_per_pair = plan.g; st.passes = plan.npass;
    if (words_of(plan) > S.sprof_cap) {
      if (S.d_sprof) { cudaFree(S.d_sprof); S.d_sprof = nullptr; }
      S.sprof_cap = words_of(plan) * 3 / 2;
      CU(cudaMalloc((void**)&S.d_sprof, S.sprof_cap * sizeof(unsigned)));
    }
    if (split_long && words_of(plan_long) > S.sprof2_cap) {
      if (S.d_sprof2) { cudaFree(S.d_sprof2); S.d_sprof2 = nullptr; }
      S.sprof2_cap = words_of(plan_long) * 2;
      CU(cudaMalloc((void**)&S.d_sprof2, S.sprof2_cap * sizeof(unsigned)));
    }
  }
  S.three_state = three;
  S.use16 = use16; S.split_long = split_long; S.n_split_pairs = split_long ? n_split_pairs : 0; S.lq = lq; S.oe = oe; S.ge = ge;

  // ---- timed region: first kernel -> last result resident in HBM ----
  CU(cudaEventRecord(e[0], sm));
  CU(cudaMemsetAsync(S.d_counters, 0, (16 + 256) * sizeof(unsigned), sm));
  int r;
  if (lq == 0 || db->n == 0) {
    if (db->n) CU(cudaMemsetAsync(S.d_scores, 0, sizeof(int32_t) * (size_t)db->n, sm));
    CU(cudaEventRecord(e[5], sm));
    CU(cudaStreamWaitEvent(sa, e[5], 0));
    CU(cudaEventRecord(e[1], sa));
  } else if (use16) {
    swk::K1SParams PS; memset(&PS, 0, sizeof(PS));
    PS.scols = db->d_scols; PS.stream_off = db->d_stream_off; PS.stream_ncols = db->d_stream_ncols;
    PS.stream_pair0 = db->d_stream_pair0; PS.bnd = S.d_bnd_s; PS.bnd_stride = db->bnd_s_stride;
    PS.BB = B * 0x10001u; PS.BBM = (B + (unsigned)(-oe)) * 0x10001u;
    PS.OE2 = ((unsigned)oe & 0xffffu) * 0x10001u; PS.GE2 = ((unsigned)ge & 0xffffu) * 0x10001u;
    const unsigned limit = 65535u - 129u;
    const int64_t nls = split_long ? n_split_streams : 0, nlp = split_long ? n_split_pairs : 0;
    CU(cudaEventRecord(e[3], sm));
    CU(cudaStreamWaitEvent(sa, e[3], 0));
    if (split_long) {
      swk::build_sprofile<<<32, 256, 0, sa>>>(S.d_query, (int)lq, S.d_table, plan_long.R, 32, 1, S.d_sprof2);
      swk::K1SParams PL = PS;
      PL.stream_begin = 0; PL.n_streams = (int)nls; PL.n_items = (int)nls; PL.prof = S.d_sprof2; PL.npass = 1;
      PL.g = 32; PL.g_log2 = 5; PL.counter = S.d_counters + 1; PL.raw = S.d_raw_s;
      PL.active_warps = 1;                                     // one long stream per CTA: spread them over the SMs
      if ((r = launch_k1s(db, PL, 32, plan_long.R, false, oe, ge, sa, (int)std::min<int64_t>(nls, db->k1s_grid)))) return r;
      swk::k1s_finalize<<<std::max(1, (int)((nlp + 255) / 256)), 256, 0, sa>>>(
          S.d_raw_s, 0, (long)nlp, 32, (long)db->n, B, limit, db->d_order, db->d_plain_off, db->long_threshold,
          S.d_scores, S.d_rescore_list, S.d_counters + 2, reinterpret_cast<unsigned long long*>(S.d_counters + 12),
          plan_long.R, (int)lq, S.d_row_bound);
      CU(cudaGetLastError()); st.kernel_launches += 3;
    }
    CU(cudaEventRecord(e[4], sa));
    swk::build_sprofile<<<std::min(2048, (int)(words_of(plan) / 256) + 1), 256, 0, sm>>>(
        S.d_query, (int)lq, S.d_table, plan.R, plan.g, plan.npass, S.d_sprof);
    CU(cudaGetLastError()); st.kernel_launches++;
    const int per_warp = 32 / plan.g;
    const bool multi = plan.npass > 1;
    // X items: with several stripes the long streams are cut into (8 streams, one stripe) items that run
    // concurrently on different CTAs (sw_k1s.cuh); their pairs get one set of maxima slots per stripe.
    // A database too small to fill the GPU with (8 streams x all stripes) items goes through X items entirely.
    const int k1s_ctas = db->sm_count * (plan.R <= K1S_R3CTA ? 3 : 2);
    const bool all_x = multi && (db->n_streams - db->n_long_streams + swk::K1S_WARPS - 1) / swk::K1S_WARPS < 2 * (int64_t)k1s_ctas;
    const int64_t nxs = !multi ? 0 : all_x ? db->n_streams : db->n_long_streams;
    const int64_t nxp = !multi ? 0 : all_x ? db->n_pairs : db->n_long_pairs;
    if (nxs > 0) {
      const int64_t need_bnd = 2 * (all_x ? db->scols_len : db->long_cols) + 64;
      if (need_bnd > S.xs_bnd_cap) {
        cudaFree(S.d_xs_bnd); S.d_xs_bnd = nullptr; S.xs_bnd_cap = need_bnd;
        CU(cudaMalloc((void**)&S.d_xs_bnd, sizeof(uint2) * (size_t)need_bnd));
      }
      if (nxs * plan.npass > S.xs_prog_cap) {
        cudaFree(S.d_xs_progress); S.d_xs_progress = nullptr; S.xs_prog_cap = nxs * plan.npass * 3 / 2;
        CU(cudaMalloc((void**)&S.d_xs_progress, sizeof(unsigned) * (size_t)S.xs_prog_cap));
      }
      if (nxp * plan.npass > S.xs_raw_cap) {
        cudaFree(S.d_xs_raw); S.d_xs_raw = nullptr; S.xs_raw_cap = nxp * plan.npass * 3 / 2;
        CU(cudaMalloc((void**)&S.d_xs_raw, sizeof(unsigned) * 32 * (size_t)S.xs_raw_cap));
      }
      CU(cudaMemsetAsync(S.d_xs_progress, 0, sizeof(unsigned) * (size_t)nxs * plan.npass, sm));
    }
    PS.stream_begin = (int)nls; PS.n_streams = (int)(db->n_streams - nls);
    PS.n_items = multi ? (int)((nxs + swk::K1S_WARPS - 1) / swk::K1S_WARPS) * plan.npass +
                             (int)((PS.n_streams - nxs + swk::K1S_WARPS - 1) / swk::K1S_WARPS)
                       : (PS.n_streams + per_warp - 1) / per_warp;
    PS.prof = S.d_sprof; PS.npass = plan.npass; PS.g = plan.g; PS.g_log2 = plan.g_log2;
    PS.counter = S.d_counters + 0; PS.active_warps = swk::K1S_WARPS;
    PS.n_x_streams = (int)nxs; PS.x_bnd = S.d_xs_bnd; PS.x_progress = S.d_xs_progress; PS.x_raw = S.d_xs_raw;
    PS.error_flag = S.d_counters + 6;
    // the bulk's [pair][g] maxima start after the long launch's [pair][32] region
    PS.raw = S.d_raw_s + (size_t)nlp * 32 - ((size_t)nlp << plan.g_log2);
    if (PS.n_streams > 0) {
      // all_x: the stripes of a stream form a pipeline that runs at the pace of its slowest stage; one CTA per SM keeps
      // every stage at the same (higher) step rate (C4: 56 -> 21 ms)
      int ctas = all_x ? 1 : plan.R <= K1S_R3CTA ? 3 : 2;
      if (const char* ev = getenv("SWGPU_K1S_CTAS")) ctas = std::max(1, std::min(3, atoi(ev)));      // A/B knob (profiling)
      if ((r = launch_k1s(db, PS, plan.g, plan.R, multi, oe, ge, sm, db->sm_count * ctas))) return r;
      if (nxp > 0) {
        swk::k1s_finalize<<<std::max(1, (int)((nxp + 255) / 256)), 256, 0, sm>>>(
            S.d_xs_raw, 0, (long)nxp, 32 * plan.npass, (long)db->n, B, limit, db->d_order, db->d_plain_off,
            db->long_threshold, S.d_scores, S.d_rescore_list, S.d_counters + 2, reinterpret_cast<unsigned long long*>(S.d_counters + 12),
            plan.R, (int)lq, S.d_row_bound);
        st.kernel_launches++;
      }
      const int64_t pb = nlp + nxp;                            // first pair scored by the bulk items
      if (db->n_pairs > pb)
        swk::k1s_finalize<<<(int)std::min<int64_t>((db->n_pairs - pb + 255) / 256, 4096), 256, 0, sm>>>(
            PS.raw, (long)pb, (long)db->n_pairs, 1 << plan.g_log2, (long)db->n, B, limit, db->d_order, db->d_plain_off,
            db->long_threshold, S.d_scores, S.d_rescore_list, S.d_counters + 2, reinterpret_cast<unsigned long long*>(S.d_counters + 12),
            multi ? 0 : plan.R, (int)lq, S.d_row_bound);                   // a multi-stripe bulk slot is the maximum over the stripes: no bound
      CU(cudaGetLastError()); st.kernel_launches += 2;
    }
    CU(cudaEventRecord(e[5], sm));
    CU(cudaStreamWaitEvent(sa, e[5], 0));
    CU(cudaEventRecord(e[1], sa));
    // 32-bit rescoring of the saturated (or forced) targets; the count lives on the device
    if ((r = launch_32bit(db, S, sa, S.d_rescore_list, S.d_counters + 2, 0, (int)lq, oe, ge, S.d_counters + 3,
                          S.d_counters + 4, &st.kernel_launches))) return r;
  } else {
    if (db->n) CU(cudaMemsetAsync(S.d_row_bound, 0x7f, sizeof(int) * (size_t)db->n, sm));     // no 16-bit pass: no row bounds
    CU(cudaEventRecord(e[5], sm));
    CU(cudaStreamWaitEvent(sa, e[5], 0));
    CU(cudaEventRecord(e[1], sa));
    if (three) r = launch_three_state(db, S, sa, (int)lq, oe, ge, S.d_counters + 5, &st.kernel_launches);
    else r = launch_32bit(db, S, sa, nullptr, nullptr, (int)db->n, (int)lq, oe, ge, S.d_counters + 5, S.d_counters + 4, &st.kernel_launches);
    if (r) return r;
  }
  CU(cudaEventRecord(e[2], sa));
  S.in_flight = true;
  return SWGPU_OK;
}

// Queue the read-back of a slot's results (counters always; all scores if the caller wants them) and the slot's
// completion event.  `staged`: the caller's buffer may be pageable and the call must not block -> go through the
// slot's pinned staging buffer, copied out in collect_search().
int enqueue_readback(swgpu_db* db, swgpu_slot& S, int32_t* scores, bool staged) {
  cudaStream_t sa = S.s_aux;
  CU(cudaEventRecord(S.ev[10], sa));
  CU(cudaMemcpyAsync(S.h_ctr, S.d_counters, 16 * sizeof(unsigned), cudaMemcpyDeviceToHost, sa));
  if (scores && db->n) {
    int32_t* dst = scores;
    if (staged) {
      cudaPointerAttributes at; memset(&at, 0, sizeof(at));
      const bool pinned = cudaPointerGetAttributes(&at, scores) == cudaSuccess && at.type == cudaMemoryTypeHost;
      cudaGetLastError();
      if (!pinned) {
        if (S.h_scores_cap < db->n) {
          cudaFreeHost(S.h_scores); S.h_scores = nullptr;
          CU(cudaMallocHost((void**)&S.h_scores, sizeof(int32_t) * (size_t)db->n));
          S.h_scores_cap = db->n;
        }
        dst = S.h_scores; S.scores_staged = true;
      }
    }
    S.user_scores = scores;
    CU(cudaMemcpyAsync(dst, S.d_scores, sizeof(int32_t) * (size_t)db->n, cudaMemcpyDeviceToHost, sa));
  }
  CU(cudaEventRecord(S.ev[11], sa));
  return SWGPU_OK;
}

// Wait for a slot's search (and top-k / read-back queued behind it) and fill in its statistics.
int collect_search(swgpu_db* db, swgpu_slot& S, swgpu_stats_t* stats) {
  if (!S.in_flight) return fail(SWGPU_EINVAL, "internal: slot idle");
  cudaEvent_t* e = S.ev;
  CU(cudaEventSynchronize(e[11]));
  CU(cudaStreamSynchronize(S.s_main));
  S.in_flight = false;
  S.have_scores = true;
  swgpu_stats_t& st = S.st;
  const int64_t lq = S.lq;
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, e[0], e[2])); st.ms_total = ms;
  CU(cudaEventElapsedTime(&ms, e[6], e[7])); st.ms_h2d = ms;
  if (S.user_scores) { CU(cudaEventElapsedTime(&ms, e[10], e[11])); st.ms_d2h = ms; }
  if (S.topk_k > 0) { CU(cudaEventElapsedTime(&ms, e[8], e[9])); st.ms_topk = ms; st.topk_launches = S.topk_launches; }
  st.cells = lq * db->n_residues;
  if (S.scores_staged) memcpy(S.user_scores, S.h_scores, sizeof(int32_t) * (size_t)db->n);
  if (lq > 0 && db->n > 0) {
    const unsigned* ctr = S.h_ctr;
    if (ctr[6]) return fail(SWGPU_ECUDA, "32-bit wavefront kernel stalled (internal error)");
    if (S.use16) {
      CU(cudaEventElapsedTime(&ms, e[0], e[1])); st.ms_inter = ms;
      CU(cudaEventElapsedTime(&ms, e[1], e[2])); st.ms_rescore = ms;
      if (S.split_long) { CU(cudaEventElapsedTime(&ms, e[3], e[4])); st.ms_intra = ms; st.n_long = 2 * S.n_split_pairs; }
      st.cells_inter = st.cells;
      st.n_rescored = ctr[2];
      // residues of the rescored targets: summed on the device by k1s_finalize (counters 12..13, one 64-bit word), read back
      // with the other counters -- no extra copy, nothing blocking inside the pipelined swgpu_search_many()
      st.cells_rescore = lq * (int64_t)(((unsigned long long)ctr[13] << 32) | ctr[12]);
    } else {
      st.ms_rescore = st.ms_total; st.n_rescored = db->n; st.cells_rescore = st.cells;
    }
  }
  db->last_stats = st;
  if (stats) *stats = st;
  return SWGPU_OK;
}

int run_single(swgpu_db* db, const int8_t* query, int64_t lq, const int8_t* table, int gap_open, int gap_extend,
               int32_t* scores, swgpu_stats_t* stats) {
  int r = check_args(db, query, lq, table, gap_open, gap_extend);
  if (r) return r;
  CU(cudaSetDevice(db->device));
  swgpu_slot& S = db->slots[0];
  if ((r = enqueue_search(db, S, query, lq, table, gap_open, gap_extend))) { S.in_flight = false; return r; }
  if ((r = enqueue_readback(db, S, scores, false))) { cudaDeviceSynchronize(); S.in_flight = false; return r; }
  return collect_search(db, S, stats);
}

}  // namespace

namespace swi {

// One 16-lane batch in the reference's own layout (GPU twin of alignment_fill_matrices(aligner_t*), src/alignment.c:47).
// The reference calls that function from every OpenMP worker at once (src/alignment_cmdline.c:504-507), thousands of times
// per search; each host thread therefore keeps ONE context per device (stream, device buffers that only grow, pinned
// result staging) and a call is: transpose the batch on the host, two small uploads, one k3_intra launch (one warp per
// lane, 32-bit, exact), one 64-byte read-back.  Round 1 built and destroyed a whole swgpu_db per call.
struct Batch16Ctx {
  bool ready = false;
  cudaStream_t stream = nullptr;
  int8_t *d_query = nullptr, *d_codes = nullptr, *d_table = nullptr;  int64_t query_cap = 0, codes_cap = 0;
  int64_t *d_off = nullptr, *d_order = nullptr;
  int32_t* d_scores = nullptr;  unsigned* d_counter = nullptr;
  int2* d_bnd = nullptr;  int64_t bnd_stride = 0;
  int32_t* h_scores = nullptr;  int64_t* h_off = nullptr;  int8_t* h_codes = nullptr;  int64_t h_codes_cap = 0;
  int8_t table_cache[1024];  bool table_valid = false;
  ~Batch16Ctx() {
    if (!ready) return;
    cudaFree(d_query); cudaFree(d_codes); cudaFree(d_table); cudaFree(d_off); cudaFree(d_order); cudaFree(d_scores);
    cudaFree(d_counter); cudaFree(d_bnd); cudaFreeHost(h_scores); cudaFreeHost(h_off); cudaFreeHost(h_codes);
    if (stream) cudaStreamDestroy(stream);
    cudaGetLastError();                      // at process exit the context may be gone already
  }
};

int fill_batch16(int device, const int8_t* query, int64_t query_len, const int8_t* batch, int64_t padded_len,
                 int lanes, const int8_t* table, int gap_open, int gap_extend, int32_t* scores16) {
  if (lanes < 1 || lanes > 16 || padded_len < 0 || !batch || !scores16 || !table || query_len < 0 || (query_len > 0 && !query))
    return fail(SWGPU_EINVAL, "bad batch argument");
  if (query_len > 0x7fffffff || padded_len > 0x7ffffff0) return fail(SWGPU_EINVAL, "batch too long");
  const int oe = gap_open + gap_extend, ge = gap_extend;
  if (oe < -100000 || ge < -100000 || oe > 100000 || ge > 100000) return fail(SWGPU_EINVAL, "gap score out of range");
  const bool three = needs_three_state(gap_open, gap_extend);      // a positive gap score: the reference's three-state form
  if (device < 0 || device >= 16) return fail(SWGPU_EINVAL, "device %d out of range", device);
  if (int r = pick_device(device)) return r;
  static thread_local Batch16Ctx ctxs[16];
  Batch16Ctx& c = ctxs[device];
  constexpr int kWarps = 16;                                 // 4 CTAs of K3_WARPS: one warp per lane of the batch
  if (!c.ready) {
    CU(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    CU(cudaMalloc((void**)&c.d_table, 1024));
    CU(cudaMalloc((void**)&c.d_off, sizeof(int64_t) * 17));
    CU(cudaMalloc((void**)&c.d_order, sizeof(int64_t) * 16));
    CU(cudaMalloc((void**)&c.d_scores, sizeof(int32_t) * 16));
    CU(cudaMalloc((void**)&c.d_counter, sizeof(unsigned)));
    CU(cudaMallocHost((void**)&c.h_scores, sizeof(int32_t) * 16));
    CU(cudaMallocHost((void**)&c.h_off, sizeof(int64_t) * 17));
    int64_t order[16]; for (int l = 0; l < 16; ++l) order[l] = l;
    CU(cudaMemcpy(c.d_order, order, sizeof(order), cudaMemcpyHostToDevice));
    c.ready = true;
  }
  if (query_len > c.query_cap) {
    cudaFree(c.d_query); c.d_query = nullptr; c.query_cap = std::max<int64_t>(query_len * 2, 1024);
    CU(cudaMalloc((void**)&c.d_query, (size_t)c.query_cap));
  }
  const int64_t total = (int64_t)lanes * padded_len;
  if (total > c.codes_cap || !c.d_codes) {
    cudaFree(c.d_codes); c.d_codes = nullptr; c.codes_cap = std::max<int64_t>(total * 2, 16 * 1024);
    CU(cudaMalloc((void**)&c.d_codes, (size_t)c.codes_cap));
  }
  if (total > c.h_codes_cap || !c.h_codes) {
    cudaFreeHost(c.h_codes); c.h_codes = nullptr; c.h_codes_cap = std::max<int64_t>(total * 2, 16 * 1024);
    CU(cudaMallocHost((void**)&c.h_codes, (size_t)c.h_codes_cap));
  }
  const int64_t stride = std::max<int64_t>((padded_len + 63) & ~63ll, 64);
  if (stride > c.bnd_stride) {
    cudaFree(c.d_bnd); c.d_bnd = nullptr; c.bnd_stride = stride * 2;
    CU(cudaMalloc((void**)&c.d_bnd, sizeof(int2) * (size_t)kWarps * (size_t)c.bnd_stride));
  }
  // The reference scores the '*' padding rows too (alignment.c:88 runs to the padded length),
  // so every lane is handed over at its full padded length: [pos][lane] -> lane after lane.
  for (int l = 0; l < lanes; ++l) {
    c.h_off[l] = (int64_t)l * padded_len;
    int8_t* dst = c.h_codes + c.h_off[l];
    for (int64_t j = 0; j < padded_len; ++j) {
      const int8_t v = batch[j * 16 + l];
      if (v < 0 || v > 31) return fail(SWGPU_EINVAL, "residue code outside 0..31 in the batch");
      dst[j] = v;
    }
  }
  for (int l = lanes; l <= 16; ++l) c.h_off[l] = total;
  for (int64_t i = 0; i < query_len; ++i) if (query[i] < 0 || query[i] > 31) return fail(SWGPU_EINVAL, "query code outside 0..31");
  cudaStream_t st = c.stream;
  if (query_len > 0) CU(cudaMemcpyAsync(c.d_query, query, (size_t)query_len, cudaMemcpyHostToDevice, st));
  if (!c.table_valid || memcmp(c.table_cache, table, 1024) != 0) {
    CU(cudaMemcpyAsync(c.d_table, table, 1024, cudaMemcpyHostToDevice, st));
    memcpy(c.table_cache, table, 1024); c.table_valid = true;
  }
  if (total > 0) CU(cudaMemcpyAsync(c.d_codes, c.h_codes, (size_t)total, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(c.d_off, c.h_off, sizeof(int64_t) * 17, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(c.d_counter, 0, sizeof(unsigned), st));
  CU(cudaMemsetAsync(c.d_scores, 0, sizeof(int32_t) * 16, st));
  if (query_len > 0 && padded_len > 0) {
    swk::K3Params P; memset(&P, 0, sizeof(P));
    P.worklist = nullptr; P.work_count = nullptr; P.n_work = lanes; P.counter = c.d_counter;
    P.plain = c.d_codes; P.plain_off = c.d_off; P.query = c.d_query; P.lq = (int)query_len; P.table = c.d_table;
    P.oe = oe; P.ge = ge; P.bnd = c.d_bnd; P.bnd_stride = c.bnd_stride; P.order = c.d_order;
    P.scores_file = c.d_scores; P.one = 1u; P.max_len = 0;
    if (three && query_len <= 128) swk::k3_three<4><<<kWarps / swk::K3_WARPS, swk::K3_WARPS * 32, 0, st>>>(P);
    else if (three)                swk::k3_three<8><<<kWarps / swk::K3_WARPS, swk::K3_WARPS * 32, 0, st>>>(P);
    else if (query_len <= 128)     swk::k3_intra<4><<<kWarps / swk::K3_WARPS, swk::K3_WARPS * 32, 0, st>>>(P);
    else                           swk::k3_intra<8><<<kWarps / swk::K3_WARPS, swk::K3_WARPS * 32, 0, st>>>(P);
    CU(cudaGetLastError());
  }
  CU(cudaMemcpyAsync(c.h_scores, c.d_scores, sizeof(int32_t) * 16, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (int l = 0; l < lanes; ++l) scores16[l] = c.h_scores[l];
  return SWGPU_OK;
}

}  // namespace swi

extern "C" {

int swgpu_search_resident(swgpu_db_t* db, const int8_t* query, int64_t query_len, const int8_t* table,
                          int gap_open, int gap_extend, swgpu_stats_t* stats) {
  return run_single(db, query, query_len, table, gap_open, gap_extend, nullptr, stats);
}

int swgpu_db_scores_d2h(swgpu_db_t* db, int32_t* scores) {
  if (!db || !scores) return fail(SWGPU_EINVAL, "bad argument");
  swgpu_slot& S = db->slots[0];
  if (!S.have_scores) return fail(SWGPU_EINVAL, "no search result resident");
  CU(cudaSetDevice(db->device));
  if (db->n) CU(cudaMemcpy(scores, S.d_scores, sizeof(int32_t) * (size_t)db->n, cudaMemcpyDeviceToHost));
  return SWGPU_OK;
}

int swgpu_search(swgpu_db_t* db, const int8_t* query, int64_t query_len, const int8_t* table,
                 int gap_open, int gap_extend, int32_t* scores, swgpu_stats_t* stats) {
  if (!scores && db && db->n) return fail(SWGPU_EINVAL, "null scores buffer");
  return run_single(db, query, query_len, table, gap_open, gap_extend, scores, stats);
}

int swgpu_search_many(swgpu_db_t* db, int32_t n_queries, const int8_t* const* queries, const int64_t* query_lens,
                      const int8_t* table, int gap_open, int gap_extend, int32_t* const* scores, int32_t k,
                      swgpu_hit_t* hits, int32_t* n_hits, swgpu_stats_t* per_query, swgpu_batch_stats_t* batch) {
  if (!db || n_queries < 0 || (n_queries > 0 && (!queries || !query_lens))) return fail(SWGPU_EINVAL, "swgpu_search_many: bad argument");
  if (k < 0 || (k > 0 && (!hits || !n_hits))) return fail(SWGPU_EINVAL, "swgpu_search_many: bad top-k argument");
  int r;
  for (int q = 0; q < n_queries; ++q)
    if ((r = check_args(db, queries[q], query_lens[q], table, gap_open, gap_extend))) return r;
  CU(cudaSetDevice(db->device));
  const int kk = (int)std::min<int64_t>(k, db->n);
  const int n_slots = std::min<int>(SWGPU_SLOTS, std::max(1, n_queries));
  for (int s = 0; s < n_slots; ++s) if ((r = swi::slot_prepare(db, db->slots[s]))) return r;
  swgpu_batch_stats_t bs; memset(&bs, 0, sizeof(bs));
  bs.slots = n_slots;
  CU(cudaEventRecord(db->ev_batch[0], db->slots[0].s_main));
  int rc = SWGPU_OK;
  auto collect = [&](int q) -> int {
    swgpu_slot& S = db->slots[q % n_slots];
    swgpu_stats_t st;
    int r2 = collect_search(db, S, &st);
    if (r2) return r2;
    if (kk > 0 && (r2 = swi::collect_topk(db, S, hits + (size_t)q * k, n_hits + q))) return r2;
    else if (k > 0 && kk == 0) n_hits[q] = 0;
    if (per_query) per_query[q] = st;
    bs.cells += st.cells; bs.n_rescored += st.n_rescored; bs.kernel_launches += st.kernel_launches + st.topk_launches;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, db->ev_batch[0], S.ev[11]) == cudaSuccess) bs.ms_total = std::max<double>(bs.ms_total, ms);
    return SWGPU_OK;
  };
  int next_collect = 0;
  for (int q = 0; q < n_queries && rc == SWGPU_OK; ++q) {
    if (q >= n_slots) { rc = collect(next_collect++); if (rc) break; }
    swgpu_slot& S = db->slots[q % n_slots];
    if ((rc = enqueue_search(db, S, queries[q], query_lens[q], table, gap_open, gap_extend))) { S.in_flight = false; break; }
    if (kk > 0 && (rc = swi::enqueue_topk(db, S, kk))) break;
    if ((rc = enqueue_readback(db, S, scores ? scores[q] : nullptr, true))) break;
  }
  if (rc != SWGPU_OK) {                                     // drain whatever is queued, then report the first error
    cudaDeviceSynchronize();
    for (auto& S : db->slots) S.in_flight = false;
    return rc;
  }
  for (; next_collect < n_queries; ++next_collect) if ((rc = collect(next_collect))) {
    cudaDeviceSynchronize();
    for (auto& S : db->slots) S.in_flight = false;
    return rc;
  }
  if (batch) *batch = bs;
  return SWGPU_OK;
}

int swgpu_fill_batch16(int device, const int8_t* query, int64_t query_len, const int8_t* batch, int64_t padded_len,
                       int lanes, const int8_t* table, int gap_open, int gap_extend, int32_t* scores16) {
  return swi::fill_batch16(device, query, query_len, batch, padded_len, lanes, table, gap_open, gap_extend, scores16);
}

}  // extern "C"
