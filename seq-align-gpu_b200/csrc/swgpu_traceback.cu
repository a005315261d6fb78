// swgpu_traceback.cu -- alignment paths (CIGAR) for reported hits.
//
// The reference keeps only a score per target (src/alignment.c:133 of the reference tree); upstream seq-align
// printed alignments, this fork removed them (Report p.10).  Nothing here is pinned by the reference: the rule
// is this file's; the test suite holds an independent scalar restatement and compares operation for operation.
//
// A hit comes with its end cell and start cell (swgpu_topk).  The optimal LOCAL alignment between
// those two cells is an optimal GLOBAL alignment of the two substrings that starts and ends with a match state
// (any other global alignment of the substrings is also a local alignment of the sequences, so it cannot score
// more).  The kernel therefore runs the affine GLOBAL recurrence on the m x n rectangle with the corner forced:
//   D(i,j) = H(i-1,j-1) + s(i,j)              H(-1,-1) = 0, every other H(-1,.) / H(.,-1) = -inf
//   E(i,j) = max(H(i,j-1) + oe, E(i,j-1) + ge)      gap in the query  (consumes a target residue, CIGAR 'D')
//   F(i,j) = max(H(i-1,j) + oe, F(i-1,j) + ge)      gap in the target (consumes a query residue,  CIGAR 'I')
//   H(i,j) = max(D, E, F)
// and keeps one direction byte per cell: bits 0-1 = source of H (0 D, 1 E, 2 F; ties: D, then E, then F),
// bit 2 = E extended (ties: open), bit 3 = F extended (ties: open).  One CTA per hit: thread r owns row r of a
// 256-row stripe, the stripe is swept along anti-diagonals (one __syncthreads per diagonal, neighbours exchange
// (H, F) through shared memory), direction bytes are stored diagonal-major so that the 256 threads write 256
// consecutive bytes; the stripe's bottom row goes through a ping-pong row in global memory.  Thread 0 then walks
// the path back from (m-1, n-1) and emits run-length encoded operations (in reverse; the host turns them round).
#include "swgpu_internal.h"

#include <algorithm>
#include <cstring>
#include <vector>

namespace swk {

constexpr int TB_ROWS = 256;                    // rows per stripe = threads per CTA
constexpr int TB_NEG = -(1 << 29);

struct TbHit {
  long long t_off;      // offset of the target in `plain` (sorted order)
  int q0, t0, m, n;     // rectangle: rows q0..q0+m-1 of the query, columns t0..t0+n-1 of the target
  int score;            // expected H(m-1, n-1)
  long long dir_off;    // offset of this hit's direction bytes
  long long ops_off;    // offset of this hit's operations (uint32), capacity m + n
};

struct TbParams {
  const TbHit* hits; int n_hits;
  unsigned* counter;
  const int8_t* plain; const int8_t* query; const int8_t* table;
  int oe, ge;
  unsigned char* dir;
  int2* rows; long long row_stride;   // per CTA two rows (H, F) of row_stride columns
  unsigned* ops; int* n_ops; int* status;   // status: 0 ok, 1 score mismatch, 2 broken path
};

__device__ __forceinline__ long long tb_dir_index(int i, int j, int n) {
  const int st = i / TB_ROWS, r = i - st * TB_ROWS;
  return (long long)st * (long long)(n + TB_ROWS) * TB_ROWS + (long long)(r + j) * TB_ROWS + r;
}

__global__ void __launch_bounds__(TB_ROWS) tb_fill_trace(const TbParams P) {
  __shared__ int8_t tab[32 * 32];
  __shared__ int2 xch[2][TB_ROWS];               // (H, F) of every row at the previous diagonal
  __shared__ int s_item;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = P.table[i];
  const int r = threadIdx.x;
  int2* row_a = P.rows + (size_t)blockIdx.x * 2 * (size_t)P.row_stride;
  int2* row_b = row_a + P.row_stride;
  const int oe = P.oe, ge = P.ge;

  for (;;) {
    __syncthreads();
    if (r == 0) s_item = (int)atomicAdd(P.counter, 1u);
    __syncthreads();
    const int item = s_item;
    if (item >= P.n_hits) break;
    const TbHit h = P.hits[item];
    const int8_t* qp = P.query + h.q0;
    const int8_t* tp = P.plain + h.t_off + h.t0;
    unsigned char* dir = P.dir + h.dir_off;
    const int m = h.m, n = h.n;
    const int nstripes = (m + TB_ROWS - 1) / TB_ROWS;
    int final_h = TB_NEG;

    for (int st = 0; st < nstripes; ++st) {
      const int i = st * TB_ROWS + r;
      const bool live = i < m;
      const int8_t* trow = tab + (live ? (qp[i] & 31) * 32 : 0);
      const int2* row_in = (st & 1) ? row_b : row_a;      // bottom row of the previous stripe
      int2* row_out = (st & 1) ? row_a : row_b;
      const bool last_row = (r == TB_ROWS - 1) || (i == m - 1);
      unsigned char* dbase = dir + (long long)st * (long long)(n + TB_ROWS) * TB_ROWS + r;
      int Hl = TB_NEG, El = TB_NEG;                       // H(i, j-1), E(i, j-1)
      int up_h_prev = (st == 0 && r == 0) ? 0 : TB_NEG;   // H(i-1, j-1); the forced corner H(-1,-1) = 0
      if (r == 0 && st > 0) up_h_prev = TB_NEG;           // H(i-1, -1) = -inf
      xch[0][r] = make_int2(TB_NEG, TB_NEG); xch[1][r] = make_int2(TB_NEG, TB_NEG);
      __syncthreads();
      const int ndiag = n + min(TB_ROWS, m - st * TB_ROWS) - 1;
      for (int d = 0; d < ndiag; ++d) {
        const int j = d - r;
        int2 mine = make_int2(TB_NEG, TB_NEG);
        if (live && j >= 0 && j < n) {
          int2 up;                                         // (H, F)(i-1, j)
          if (r > 0) up = xch[(d + 1) & 1][r - 1];
          else if (st > 0) up = __ldcg(&row_in[j]);
          else up = make_int2(TB_NEG, TB_NEG);
          const int s = (int)trow[tp[j] & 31];
          const int D = up_h_prev > TB_NEG / 2 ? up_h_prev + s : TB_NEG;
          const int e_open = Hl + oe, e_ext = El + ge;
          const int E = max(e_open, e_ext);
          const int f_open = up.x + oe, f_ext = up.y + ge;
          const int F = max(f_open, f_ext);
          int Hc = D; unsigned char code = 0;
          if (E > Hc) { Hc = E; code = 1; }
          if (F > Hc) { Hc = F; code = 2; }
          if (e_ext > e_open) code |= 4;
          if (f_ext > f_open) code |= 8;
          Hc = max(Hc, TB_NEG);
          dbase[(long long)d * TB_ROWS] = code;
          up_h_prev = up.x;
          Hl = Hc; El = max(E, TB_NEG);
          mine = make_int2(Hc, max(F, TB_NEG));
          if (last_row) {
            if (i == m - 1) { if (j == n - 1) final_h = Hc; }
            else row_out[j] = mine;
          }
        }
        xch[d & 1][r] = mine;
        __syncthreads();
      }
      __threadfence_block();
      __syncthreads();
    }

    // the owner of cell (m-1, n-1) reports the score, then thread 0 walks the path back
    __shared__ int s_final;
    if (final_h != TB_NEG) s_final = final_h;
    __syncthreads();
    if (r == 0) {
      int status = (s_final == h.score) ? 0 : 1;
      unsigned* ops = P.ops + h.ops_off;
      int nops = 0, i = m - 1, j = n - 1, state = 0;      // 0 = H, 1 = E, 2 = F
      unsigned cur_op = 3, cur_len = 0;
      const int cap = m + n;
      while (i >= 0 && j >= 0 && status == 0) {
        const unsigned char c = __ldcg(dir + tb_dir_index(i, j, n));
        unsigned op;
        if (state == 0) {
          const int src = c & 3;
          if (src == 1) { state = 1; continue; }
          if (src == 2) { state = 2; continue; }
          op = 0; --i; --j;                               // 'M'
        } else if (state == 1) {
          op = 2; state = (c & 4) ? 1 : 0; --j;           // 'D': consumes a target residue
        } else {
          op = 1; state = (c & 8) ? 2 : 0; --i;           // 'I': consumes a query residue
        }
        if (op == cur_op) ++cur_len;
        else {
          if (cur_len) { if (nops < cap) ops[nops] = cur_len << 2 | cur_op; ++nops; }
          cur_op = op; cur_len = 1;
        }
      }
      if (cur_len) { if (nops < cap) ops[nops] = cur_len << 2 | cur_op; ++nops; }
      if (status == 0 && (i != -1 || j != -1 || state != 0)) status = 2;   // must end exactly at the forced corner
      P.n_ops[item] = nops;
      P.status[item] = status;
    }
  }
}

}  // namespace swk

using swi::fail;

extern "C" int swgpu_traceback(swgpu_db_t* db, const int8_t* query, int64_t query_len, const int8_t* table,
                               int gap_open, int gap_extend, const swgpu_hit_t* hits, int32_t n_hits,
                               uint32_t* ops, int64_t ops_cap, int64_t* ops_off) {
  if (!db || n_hits < 0 || (n_hits > 0 && (!hits || !ops_off)) || !table || (query_len > 0 && !query))
    return fail(SWGPU_EINVAL, "swgpu_traceback: bad argument");
  const int oe = gap_open + gap_extend, ge = gap_extend;
  if (oe > 0 || ge > 0 || gap_open > 0) return fail(SWGPU_EINVAL, "swgpu_traceback: positive gap scores are not supported");
  if (ops_off) ops_off[0] = 0;
  if (n_hits == 0) return SWGPU_OK;
  CU(cudaSetDevice(db->device));

  // rectangles; hits without coordinates (score 0) get an empty path
  std::vector<swk::TbHit> th;
  std::vector<int> which;
  long long dir_total = 0, ops_total = 0; int max_n = 32;
  for (int k = 0; k < n_hits; ++k) {
    const swgpu_hit_t& h = hits[k];
    if (h.score <= 0 || h.q_start < 0) continue;
    const int64_t e = h.entry - db->entry_base;
    if (e < 0 || e >= db->n) return fail(SWGPU_EINVAL, "swgpu_traceback: hit %d: entry outside this database", k);
    const int64_t sidx = db->h_inv_order[(size_t)e];
    const int64_t lt = db->h_len_sorted_prefix[sidx + 1] - db->h_len_sorted_prefix[sidx];
    if (h.q_end < h.q_start || h.t_end < h.t_start || h.q_end >= query_len || h.t_end >= lt)
      return fail(SWGPU_EINVAL, "swgpu_traceback: hit %d: coordinates outside the sequences", k);
    swk::TbHit t;
    t.t_off = db->h_len_sorted_prefix[sidx]; t.q0 = h.q_start; t.t0 = h.t_start;
    t.m = h.q_end - h.q_start + 1; t.n = h.t_end - h.t_start + 1; t.score = h.score;
    const long long ns = (t.m + swk::TB_ROWS - 1) / swk::TB_ROWS;
    t.dir_off = dir_total; dir_total += ns * (long long)(t.n + swk::TB_ROWS) * swk::TB_ROWS;
    t.ops_off = ops_total; ops_total += (long long)t.m + t.n;
    max_n = std::max(max_n, t.n);
    th.push_back(t); which.push_back(k);
  }
  const int nt = (int)th.size();
  std::vector<int> h_nops(nt, 0), h_status(nt, 0);
  std::vector<unsigned> h_ops((size_t)std::max<long long>(ops_total, 1));
  if (nt > 0) {
    if (dir_total > (16ll << 30)) return fail(SWGPU_ENOMEM, "swgpu_traceback: %lld bytes of direction data needed (limit 16 GiB); pass fewer hits per call", dir_total);
    const int grid = std::min(nt, db->sm_count * 2);
    unsigned char* d_dir = nullptr; swk::TbHit* d_hits = nullptr; unsigned* d_ops = nullptr; int* d_nops = nullptr;
    int* d_status = nullptr; unsigned* d_counter = nullptr; int2* d_rows = nullptr; int8_t* d_query = nullptr; int8_t* d_table = nullptr;
    const long long row_stride = (max_n + 31) & ~31;
    auto body = [&]() -> int {
      CU(cudaMalloc((void**)&d_dir, (size_t)dir_total));
      CU(cudaMalloc((void**)&d_hits, sizeof(swk::TbHit) * nt));
      CU(cudaMalloc((void**)&d_ops, sizeof(unsigned) * (size_t)ops_total));
      CU(cudaMalloc((void**)&d_nops, sizeof(int) * nt));
      CU(cudaMalloc((void**)&d_status, sizeof(int) * nt));
      CU(cudaMalloc((void**)&d_counter, sizeof(unsigned)));
      CU(cudaMalloc((void**)&d_rows, sizeof(int2) * 2 * (size_t)row_stride * grid));
      CU(cudaMalloc((void**)&d_query, (size_t)std::max<int64_t>(query_len, 1)));
      CU(cudaMalloc((void**)&d_table, 1024));
      cudaStream_t s = db->slots[0].s_aux;
      CU(cudaMemcpyAsync(d_hits, th.data(), sizeof(swk::TbHit) * nt, cudaMemcpyHostToDevice, s));
      CU(cudaMemcpyAsync(d_query, query, (size_t)query_len, cudaMemcpyHostToDevice, s));
      CU(cudaMemcpyAsync(d_table, table, 1024, cudaMemcpyHostToDevice, s));
      CU(cudaMemsetAsync(d_counter, 0, sizeof(unsigned), s));
      swk::TbParams P; memset(&P, 0, sizeof(P));
      P.hits = d_hits; P.n_hits = nt; P.counter = d_counter; P.plain = db->d_plain; P.query = d_query; P.table = d_table;
      P.oe = oe; P.ge = ge; P.dir = d_dir; P.rows = d_rows; P.row_stride = row_stride; P.ops = d_ops; P.n_ops = d_nops;
      P.status = d_status;
      swk::tb_fill_trace<<<grid, swk::TB_ROWS, 0, s>>>(P);
      CU(cudaGetLastError());
      CU(cudaMemcpyAsync(h_nops.data(), d_nops, sizeof(int) * nt, cudaMemcpyDeviceToHost, s));
      CU(cudaMemcpyAsync(h_status.data(), d_status, sizeof(int) * nt, cudaMemcpyDeviceToHost, s));
      CU(cudaMemcpyAsync(h_ops.data(), d_ops, sizeof(unsigned) * (size_t)ops_total, cudaMemcpyDeviceToHost, s));
      CU(cudaStreamSynchronize(s));
      return SWGPU_OK;
    };
    const int rc = body();
    cudaFree(d_dir); cudaFree(d_hits); cudaFree(d_ops); cudaFree(d_nops); cudaFree(d_status); cudaFree(d_counter);
    cudaFree(d_rows); cudaFree(d_query); cudaFree(d_table);
    if (rc) return rc;
  }
  // hand over: operations of hit k in path order (start -> end) at ops[ops_off[k] .. ops_off[k+1])
  int64_t w = 0; int t = 0;
  for (int k = 0; k < n_hits; ++k) {
    ops_off[k] = w;
    if (t < nt && which[t] == k) {
      if (h_status[t] == 1) return fail(SWGPU_ECUDA, "swgpu_traceback: hit %d: the rectangle does not score %d (query, matrix or gaps differ from the search?)", k, hits[k].score);
      if (h_status[t] == 2) return fail(SWGPU_ECUDA, "swgpu_traceback: hit %d: broken path (internal error)", k);
      const int no = h_nops[t];
      if (w + no > ops_cap) return fail(SWGPU_EINVAL, "swgpu_traceback: operation buffer too small");
      const unsigned* src = h_ops.data() + th[t].ops_off;
      for (int x = 0; x < no; ++x) ops[w + x] = src[no - 1 - x];
      w += no; ++t;
    }
  }
  ops_off[n_hits] = w;
  return SWGPU_OK;
}
