// Host-side planning for reductions: axis canonicalisation, dim merging and the
// choice of kernel + launch geometry.  Pure C++ (no CUDA types) so the same code
// is unit-tested on a CPU-only box.
//
// Re-states, for one device family, what the reference spreads over
//   details::GetReduceDim            reduce_function.h:100-123   (negative axes, range check)
//   recompute_reduce_all             kernel_utils.h:405-416
//   ReduceConfig::SetReduceDim       reduce_function.h:301-383   (merge adjacent dims)
//   ReduceConfig::SetReduceType      reduce_function.h:410-431   (LastDim / HigherDim / Any)
//   ReduceConfig::SetBlockDim*       reduce_function.h:434-561   (geometry)
// but sizes grids for 148 SMs and never narrows an extent to 32 bits.
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <string>

#include "../../include/b200phi.h"

namespace b200phi {

enum ReduceKind {
  kRowWarp = 0,   // [rows, cols], G < 256 threads per row (several rows per CTA)
  kRowBlock = 1,  // [rows, cols], one CTA (G = 256) per row
  kRowSplit = 2,  // [rows, cols], several CTAs per row + last-block finalise
  kCol = 3,       // [outer, R, inner]
  kGeneric = 4,   // anything else (also the reduce_num == 1 cast/copy)
  kRowSmall = 5,  // [rows, cols] with short rows: G < 32 lanes per row
  kRowTiny = 9,   // [rows, cols], aligned rows of 1..16 vectors: G <= 16 lanes per row, 8 rows per group
  kRowScalar = 8, // [rows, cols], 129..1023 columns of any alignment: G lanes per row, scalar loads
  kRowBatch = 7,  // [rows, cols], aligned rows under ~16 KiB: all loads of R rows in one round
  kTwoStage = 6,  // two or more reduced dims: one of them peeled by a streaming row / column kernel
                  // into scratch, then the (much smaller) rest reduced by whatever its plan says
  kRowStaged = 14,  // [rows, cols], rows of at most 512 bytes that are not whole aligned vectors: spans of 128 rows
                    // bulk-copied into shared memory, a thread per row (rows_staged.cuh)
  kRowPeel = 13,    // [rows, cols] of 1- / 2-byte elements, rows not whole aligned vectors, up to 256 vectors:
                    // G = 8..32 lanes per row, head element + NV vectors + tail element per lane
  kCastFirst = 12,  // out_dtype != x.dtype where the reference's Cast pre-pass is not foldable into the
                    // load (sums into bf16 / fp16, products, narrowing integers): cast into scratch, then reduce
  kTiledKeep = 10,  // any other pattern whose innermost merged dim is kept (tile_keep_kernel)
  kTiledRed = 11    // any other pattern whose innermost merged dim is reduced (tile_red_kernel)
};

constexpr int kPlanTiledMaxDims = 5;

struct ReducePlan {
  int kind;
  int op;         // b200phi_reduce_op with MEAN/AMAX/AMIN kept as given
  int in_dtype;
  int out_dtype;  // resolved (never UNDEFINED)
  int acc;        // 0 float, 1 double, 2 int64, 3 logical(int)
  long long in_numel, out_numel, red_numel;
  // row kinds
  long long rows, cols, seg;
  int G;  // threads per row: 32, 64, 128 or 256
  int aligned;  // every row starts 16-byte aligned and is a whole number of 16-byte vectors
  int nv, rpg;  // kRowBatch: 16-byte loads per row per thread, rows per group per round
  // col kind
  long long outer, R, inner, tiles, rows_per_split;
  int vec;  // elements per 16-byte vector on the col vector path, else 1
  int TX;
  // generic kind
  int nkeep, nred;
  long long keep_extent[B200PHI_MAX_RANK], keep_stride[B200PHI_MAX_RANK];
  long long red_extent[B200PHI_MAX_RANK], red_stride[B200PHI_MAX_RANK];
  // tiled kinds: dims walked with 32-bit divides, innermost last (extents < 2^31)
  int td_nkeep, td_nred;
  unsigned td_keep_extent[kPlanTiledMaxDims], td_red_extent[kPlanTiledMaxDims];
  long long td_keep_stride[kPlanTiledMaxDims], td_red_stride[kPlanTiledMaxDims];
  unsigned nq;        // kTiledKeep: output vectors; (TX, vec, rows_per_split as in the col kind)
  unsigned Lv, W, chunk;  // kTiledRed: units per contiguous run, units per output, units per split; G = 32 or 256
  // launch
  unsigned grid_x, grid_y;
  int splits;
  size_t ws_partial_bytes;  // first region of the workspace
  size_t ws_ticket_bytes;   // second region (256-byte aligned start)
  size_t ws_total_bytes;
  // two-stage kind: the merged dims, which reduced dim stage 1 peels, the op / dtype of each
  // stage, and how the workspace is cut
  int ts_rank, ts_peel;
  long long ts_dims[B200PHI_MAX_RANK];
  int ts_isred[B200PHI_MAX_RANK];
  int ts_op1, ts_op2, ts_tmp_dtype;
  size_t ts_tmp_bytes, ts_ws1_bytes, ts_ws2_bytes;
};

// The two reductions a kTwoStage plan stands for; `tmp` is the intermediate (input without the peeled dim).
void two_stage_descs(const ReducePlan& p, const b200phi_reduce_desc* d, void* tmp,
                     b200phi_reduce_desc* d1, b200phi_reduce_desc* d2);

// Split count of a reduced extent given `bxy` blocks from the outputs, `slots` resident blocks and
// the input-to-partial byte ratio per output (reduce_plan.cc).
long long pick_splits(long long bxy, long long slots, long long max_splits, double ratio);

size_t dtype_size(int dtype);
bool dtype_known(int dtype);
bool dtype_is_float(int dtype);

// Tunables (read once from the environment; see DESIGN.md "knobs").
struct Tunables {
  int sm_count;             // multiprocessors of the current device (148 on B200)
  long long row_bytes_per_thread;  // a row gets ~row_bytes / this many threads (pow2, 32..256)
  int ctas_per_sm;          // resident CTAs per SM the split heuristics aim for
  int force_kind;           // -1, or a ReduceKind to force where legal (testing)
  int allow_narrow_out;     // internal: a 16-bit float result from a wider intermediate (stage 2)
  int no_two_stage;         // env B200PHI_NO_TWO_STAGE: keep [A,K,B] patterns on the generic kernel
  int no_row_batch;         // env B200PHI_NO_ROW_BATCH: keep medium rows on row_reduce_kernel
  int row_batch_waves;      // env B200PHI_ROW_BATCH_WAVES: grid cap = SMs * 8 * this
  int no_row_peel;          // env B200PHI_NO_ROW_PEEL: keep 16-bit unaligned rows on the scalar-load kernel
  int no_row_staged;        // env B200PHI_NO_ROW_STAGED: keep short ragged rows off the shared-memory thread-per-row kernel
  int no_tiled;             // env B200PHI_NO_TILED: keep general keep/reduce patterns on the scalar walk
  long long p2p_timeout_ms; // env B200PHI_P2P_TIMEOUT_MS: how long the peer exchange waits for a flag
};
Tunables default_tunables(int sm_count);

// ---- short rows through shared memory, a thread per row (rows_staged.cuh) --------------------
constexpr int kStThreads = 128;  // rows per span = threads per CTA
// Bytes of one ring stage for rows of `row_bytes`.
inline int staged_stage_bytes(long long row_bytes) {
  return static_cast<int>(((static_cast<long long>(kStThreads) * row_bytes + 15) & ~15LL) + 32);
}
// Can rows of this geometry take the staged kernels?  Short rows only (a span must fit shared
// memory a few times over), enough rows to fill the machine, and a row pitch that does not put the
// threads of a warp on the same shared-memory bank: thread t reads word (t * row_bytes + i) / 4, so
// a pitch that is a multiple of 16 bytes (4 words) maps four or more lanes to one bank.
// Measured limits (tools/stagedbench.py, profiles/r02_stagedbench_raw.txt): 16-bit rows up to 512
// bytes, wider elements up to 832; below 16 / 32 / 64 bytes (2- / 4- / 8-byte elements) the plain
// reductions (sum, max ...) are faster on the sub-warp scalar kernels, while argmax, statistics and
// logsumexp gain at every length.
long long staged_max_row_bytes();  // B200PHI_STAGED_MAX_BYTES overrides both upper limits (reduce_plan.cc)
long long staged_min_row_bytes();  // B200PHI_STAGED_MIN_BYTES overrides the plain reductions' lower limits
int staged_aligned_mode();         // B200PHI_STAGED_ALIGNED: 0 never, 1 (default) see below, 2 wherever possible
enum StagedFamily { kStagedReduce = 0, kStagedArg = 1, kStagedStats = 2, kStagedLse = 3 };
// Rows that are whole 16-byte vectors (and start on 16-byte boundaries) are visited vector by vector
// from staggered starts (rows_staged.cuh).  They have vector kernels of their own, so the staged
// kernels take them only where those measured slower (tools/stagedbench.py ... all,
// profiles/r02_stagedbench_aligned_raw.txt): statistics at every length, argmax / logsumexp at every
// length for 4- and 8-byte elements and up to 416 bytes per row for 16-bit ones, the plain
// reductions when the row is not a power-of-two number of vectors (row_tiny_kernel rounds its lanes
// up) -- from 64 bytes; 16-bit rows up to 256 bytes -- and, for 4- and 8-byte elements, any row of
// 192 bytes or more.
// base16: the tensor starts on a 16-byte boundary (such rows are read with 16-byte shared loads).
inline bool staged_rows_ok(long long rows, long long row_bytes, int esize, int family, bool base16) {
  const bool plain_reduce = family == kStagedReduce;
  const long long env_max = staged_max_row_bytes();
  const long long max_bytes = env_max > 0 ? env_max : (esize <= 2 ? 512 : 832);
  const long long env_min = staged_min_row_bytes();
  const long long min_bytes = plain_reduce ? (env_min > 0 ? env_min : (esize <= 2 ? 16 : (esize == 4 ? 32 : 64))) : 2;
  if (row_bytes < min_bytes || row_bytes > max_bytes || rows < 4096) return false;
  if (row_bytes % 16 == 0) {
    const int mode = staged_aligned_mode();
    if (mode == 0 || !base16) return false;
    if (mode == 1) {
      const long long nvec = row_bytes / 16;
      const bool pow2 = (nvec & (nvec - 1)) == 0;
      if (plain_reduce) {
        const bool ok = esize >= 4 ? ((!pow2 && row_bytes >= 64) || row_bytes >= 192)
                                   : (!pow2 && row_bytes >= 64 && row_bytes <= 256);
        if (!ok) return false;
      } else if (family != kStagedStats && esize < 4 && row_bytes > 416) {
        return false;
      }
    }
  }
  return true;
}

// Fills *plan. Returns B200PHI_OK or a status with *err set.
int plan_reduce(const b200phi_reduce_desc* d, const Tunables& t, ReducePlan* plan,
                std::string* err);

}  // namespace b200phi
