/*
 * spchol.h -- C-ABI of the B200-native sparse Cholesky (libspchol.so).
 *
 * This is the drop-in boundary for the reference's numeric hot path.  The reference
 * (evanwporter/SparseCholesky) has no FFI of its own; its boundary is the public C++ header
 * API.  Each entry point below names the reference interface it replaces (file:line relative
 * to the reference tree).  The C++23 host header include/chol.hpp (same public names as the
 * reference's include/chol.hpp) and the Python binding sparsecholesky_b200/api.py are thin
 * wrappers over exactly these symbols.  INTEGRATION.md shows the reference-side binding.
 *
 * Conventions: plain pointers and sizes only; A is the upper triangle in CSC (row <= col,
 * rows ascending per column) exactly as triplet_to_csc_matrix produces it; L is lower CSC,
 * diagonal first, rows ascending (reference include/chol.hpp:844-855).  All functions return
 * 0 on success, a negative code on error (text via spchol_last_error()).  There is NO CPU
 * fallback: numeric entry points fail with SPCHOL_ENOGPU when no CUDA device is usable.
 */
#ifndef SPCHOL_H
#define SPCHOL_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPCHOL_OK        0
#define SPCHOL_EINVAL   -1   /* malformed argument / pattern */
#define SPCHOL_ENOMEM   -2
#define SPCHOL_ENOGPU   -3   /* no CUDA device: the product path never falls back to the CPU */
#define SPCHOL_ECUDA    -4   /* CUDA runtime error */
#define SPCHOL_ERANGE   -5   /* nnz(L) does not fit the 32-bit CSC the reference API uses */
#define SPCHOL_EIO      -6
#define SPCHOL_ENCCL    -7   /* NCCL missing or failed (multi-GPU entry points only) */

typedef struct spchol_symbolic spchol_symbolic;  /* host symbolic analysis == reference SChol + supernodes */
typedef struct spchol_plan     spchol_plan;      /* device-resident numeric plan for one pattern */
typedef struct spchol_comm     spchol_comm;      /* multi-GPU communicator: one rank per process per GPU */

const char* spchol_last_error(void);
const char* spchol_version(void);
int spchol_device_count(void);   /* 0 when no usable GPU */

/* ---- ingest ------------------------------------------------------------------------------ */
/* replaces triplet_to_csc_matrix<T> (include/chol.hpp:309-369). Ap[n+1], Ai/Ax[nz] caller
 * allocated; *nnz_out = merged entry count. */
int spchol_triplet_to_csc(int64_t nz, const int* ti, const int* tj, const double* tx, int n,
                          int* Ap, int* Ai, double* Ax, int64_t* nnz_out);
/* replaces load_matrix_market_to_csc<T> (include/mtx_reader.hpp:17-62): same parsing rules
 * (skip '%' lines, "rows cols nnz", 1-based triples mirrored to the upper triangle).
 * Two-call protocol: with ti == NULL only *n_out / *nz_out are written. */
int spchol_mtx_read(const char* path, int* n_out, int64_t* nz_out, int* ti, int* tj, double* tx);

/* The same on the GPU: triplets and outputs in DEVICE memory (d_Ap[n+1], d_Ai / d_Ax[nz]); a stable 64-bit
 * radix sort by (column, row) + segmented sum, duplicates added in input order, explicit zeros kept.  The CSC stays on the
 * device (d_Ax feeds spchol_factor_device).  stream: a cudaStream_t or NULL; synchronous on return. */
int spchol_triplet_to_csc_device(int64_t nz, const int* d_ti, const int* d_tj, const double* d_tx, int n,
                                 int* d_Ap, int* d_Ai, double* d_Ax, int64_t* nnz_out, void* stream);
/* The same without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and bit-for-bit the same result as spchol_triplet_to_csc_device, on the current device; the work space
 * of the sort is a static device array inside the library (36.3 MB, held by every GPU that loads the library until the
 * process exits), so the call is limited to SPCHOL_NOALLOC_MAX_NZ triplets and SPCHOL_NOALLOC_MAX_N columns and returns
 * SPCHOL_ERANGE beyond that.  It calls no cudaMalloc / cudaFree.  All calls share the array: calls that overlap on one
 * GPU (two host threads, two streams) race -- run one at a time per GPU. */
#define SPCHOL_NOALLOC_MAX_NZ 1048576
#define SPCHOL_NOALLOC_MAX_N  1048576
int spchol_triplet_to_csc_device_noalloc(int64_t nz, const int* d_ti, const int* d_tj, const double* d_tx, int n,
                                         int* d_Ap, int* d_Ai, double* d_Ax, int64_t* nnz_out, void* stream);
/* host-pointer convenience around it (H2D, device ingest, D2H); *device_ms (optional) = CUDA-event time of the device part */
int spchol_triplet_to_csc_gpu(int64_t nz, const int* ti, const int* tj, const double* tx, int n,
                              int* Ap, int* Ai, double* Ax, int64_t* nnz_out, double* device_ms);

/* The Matrix Market parse on the GPU (the text half of SURVEY.md 8f-3; replaces the `file >> row >> col >> data` loop of
 * load_matrix_market_to_csc, include/mtx_reader.hpp:41-56).  spchol_mtx_parse_device: d_text = the file's bytes from the
 * size line on (leading '%' lines removed) in DEVICE memory; writes n, nnz and nnz 0-based triplets mirrored into the upper
 * triangle to d_ti / d_tj / d_tx [nz_cap] -- the input of spchol_triplet_to_csc_device, so a matrix goes from text to the
 * factorization without its triplets visiting the host.  Decimal -> double is exact (Eisel-Lemire, ties to even = strtod);
 * the few tokens it does not decide (> 19 significant digits, inf / nan) are converted by the host's strtod, their count
 * comes back in *host_converted (optional).  Errors as spchol_mtx_read; SPCHOL_ERANGE if the file holds more than nz_cap
 * entries.  spchol_mtx_read_gpu: spchol_mtx_read's protocol and results with the parse on the device. */
int spchol_mtx_parse_device(const char* d_text, int64_t nbytes, int64_t nz_cap, int* d_ti, int* d_tj, double* d_tx,
                            int* n_out, int64_t* nz_out, int64_t* host_converted, void* stream);
int spchol_mtx_read_gpu(const char* path, int* n_out, int64_t* nz_out, int* ti, int* tj, double* tx, double* device_ms);

/* ---- orderings (host; additive: the reference has no ordering code) -------- */
/* geometric nested dissection of an nx*ny*nz grid, the fixed permutation of the benchmark Laplacians
 * (SURVEY.md 8d): perm[new] = old grid id, id = x + nx*(y + ny*z) */
int spchol_nd_perm_grid(int nx, int ny, int nz, int64_t* perm);
/* approximate-minimum-degree ordering of a symmetric pattern given as its upper triangle (CSC, diagonal
 * optional), composed with an elimination-tree postorder: perm[new] = old */
int spchol_order_amd(int n, const int* Ap, const int* Ai, int* perm);
/* B = triu(P A P^T), rows ascending per column; Bp[n+1], Bi/Bx[nnz(A)].  Ax/Bx may be NULL (pattern only);
 * map (optional, nnz(A)): position in B of every entry of A, to permute further value sets over the same pattern */
int spchol_permute_upper(int n, const int* Ap, const int* Ai, const double* Ax, const int* perm,
                         int* Bp, int* Bi, double* Bx, int64_t* map);

/* ---- symbolic helpers (integer only, host) ------------------------------------------------ */
int spchol_etree(int n, const int* Ap, const int* Ai, int* parent);                 /* chol.hpp:378  */
int spchol_post_order(int n, const int* parent, int* post);                          /* chol.hpp:466  */
int spchol_col_count(int n, const int* Ap, const int* Ai, const int* parent,
                     const int* post, int* colcount);                               /* chol.hpp:567  */
int spchol_ereach(int n, const int* Ap, const int* Ai, const double* Ax, int k,
                  const int* parent, int* s, int* w, double* x, int top);           /* chol.hpp:725/:737, returns new top */
int spchol_compute_levels(int n, const int* parent, int* level_ptr, int* level_nodes); /* src/chol.cpp:7, returns #levels */
int spchol_compute_supernodes(int n, const int* parent, const int64_t* Lp,
                              int* sn_id, int* sn_start);                           /* src/chol.cpp:42, returns #supernodes */
int spchol_atree(int n, const int* parent, const int* sn_id, int ns, const int* sn_start,
                 int* super_parent);                                                /* src/chol.cpp:102 */

/* ---- symbolic phase: replaces schol<T>(A) (include/chol.hpp:874-946) ---------------------- */
int     spchol_symbolic_create(int n, const int* Ap, const int* Ai, spchol_symbolic** out);
/* same, with explicit amalgamation control: relax = 0 keeps the reference's fundamental supernodes, 1 = default,
 * 2 = aggressive chain merging capped at 128 rows per front (for spchol_factor_batched: fewer, wider fronts),
 * 3 = default, except that a chain of >= 1024 columns is not merged into a parent that has other children (multi-GPU
 *     plans: keeps sibling separators as separate fronts so that proportional mapping can split the ranks evenly) */
int     spchol_symbolic_create_ex(int n, const int* Ap, const int* Ai, int relax, spchol_symbolic** out);
void    spchol_symbolic_destroy(spchol_symbolic*);
int     spchol_symbolic_n(const spchol_symbolic*);
int64_t spchol_symbolic_nnzL(const spchol_symbolic*);
int64_t spchol_symbolic_nnzA(const spchol_symbolic*);           /* stored entries of triu(A) the pattern was built from */
double  spchol_symbolic_flops(const spchol_symbolic*);          /* sum_j colcount_j^2 */
int     spchol_symbolic_parent(const spchol_symbolic*, int* parent);          /* SChol::parent */
int     spchol_symbolic_colcount(const spchol_symbolic*, int* colcount);
int     spchol_symbolic_Lp64(const spchol_symbolic*, int64_t* Lp);            /* n+1 */
int     spchol_symbolic_Lp(const spchol_symbolic*, int* Lp);                  /* SChol::p(); SPCHOL_ERANGE if > INT_MAX */
int     spchol_symbolic_Li(const spchol_symbolic*, int* Li);                  /* SChol::i() */
/* supernode partition: which = 0 fundamental (reference rule), 1 = relaxed fronts actually factored */
int     spchol_symbolic_supernodes(const spchol_symbolic*, int which, int* ns_out,
                                   int* sn_start /* n+1 slots */, int* sn_parent /* n slots */);
/* stats[0..7] = ns_fund, ns_relaxed, sum_m, panel_entries, padded_zeros, max_w, max_m, nlevels(relaxed) */
int     spchol_symbolic_stats(const spchol_symbolic*, int64_t* stats);

/* Supernodal export format (for spchol_factor_lbuf / spchol_lbuf_to_host of a single-GPU plan; the way L comes back
 * when nnz(L) > INT_MAX and the reference's 32-bit CSC cannot hold it): front k = relaxed supernode k owns the columns
 * sn_start[k] .. sn_start[k+1]-1 and a dense column-major panel of m[k] rows x w[k] columns with leading dimension
 * ld[k] at Lbuf + panel_off[k]; row i of the panel is global row rows[i] (spchol_symbolic_front_rows: the front's own
 * columns first, then its below-rows ascending); only the lower trapezoid (i >= column) is meaningful.  Entries that
 * relaxed amalgamation added to the pattern are exact zeros. */
int     spchol_symbolic_panel_layout(const spchol_symbolic*, int* w, int* m, int* ld, int64_t* panel_off, int64_t* lbuf_entries);
int     spchol_symbolic_front_rows(const spchol_symbolic*, int front, int* rows /* m[front] */);

/* CSC of L with 64-bit column pointers, gathered on the host from a supernodal panel buffer of the layout above:
 * Lp[n+1] (int64), Li / Lx [nnz(L)] in the reference's order (diagonal first, rows ascending, true pattern only).
 * The reference's containers are 32-bit (include/chol.hpp:52,55), so this is how a factor with nnz(L) > INT_MAX
 * (128^3 Laplacian: 2.83e9) reaches a host caller as CSC.  Lp / Li / Lx may be NULL individually (lbuf may be NULL
 * when Lx is). */
int     spchol_symbolic_gather_csc64(const spchol_symbolic*, const double* lbuf, int64_t* Lp, int* Li, double* Lx);

/* ---- numeric plan + factorization: replaces chol<T>(A) (include/chol.hpp:750-863) --------- */
/* device: CUDA ordinal. The plan owns the device L buffer and contribution workspace. */
int  spchol_plan_create(const spchol_symbolic*, int device, spchol_plan** out);
void spchol_plan_destroy(spchol_plan*);
/* plan_stats[0..9] = ns, nlevels, n_small, n_large, lbuf_entries, ws_entries, n_ops, max_m, max_w, device_bytes */
int  spchol_plan_stats(const spchol_plan*, int64_t* stats);

/* host-only dry run of the planner (no GPU needed): same stats layout, device_bytes = bytes the
 * plan would allocate; extra[0..5] = #small ops, #extadd ops, #potrf ops, #trsm ops, #gemm ops, gemm flops (as int64) */
int  spchol_plan_dryrun(const spchol_symbolic*, int64_t* stats, int64_t* extra);
/* same for rank `rank` of a `world`-rank plan; extra[6] = #transfers, extra[7] = doubles moved;
 * comm_sig (may be NULL, 4 slots): order-sensitive hashes of this rank's transfer sequence, used by the
 * tests to check that the ranks' programs match: [0] #broadcasts, [1] #sends, [2] #recvs, [3] hash */
int  spchol_plan_dryrun_dist(const spchol_symbolic*, int rank, int world, const int* dist_opts,
                             int64_t* stats, int64_t* extra, int64_t* comm_sig);

/* ---- multi-GPU: one matrix factored by `world` ranks, one process per GPU ------------------ */
/* No reference counterpart (the reference is single-threaded).  The host binding
 * moves the 128-byte NCCL unique id from rank 0 to the others (torch.distributed / MPI / a file);
 * libnccl.so.2 is dlopen()ed here, single-GPU callers never need it. */
int  spchol_comm_unique_id(void* id128);
int  spchol_comm_create(int rank, int world, const void* id128, int device, spchol_comm** out);
void spchol_comm_destroy(spchol_comm*);
int  spchol_comm_rank(const spchol_comm*);
int  spchol_comm_world(const spchol_comm*);
/* Collective over the communicator: every rank passes the same symbolic analysis.  The assembly
 * tree is mapped proportionally (subtrees -> ranks, split subtrees -> rank groups); fronts of a
 * rank group are factored block-column cyclic with in-place NCCL broadcasts of factored block
 * columns; contribution blocks move up with NCCL broadcast / send / recv.  dist_opts (may be NULL):
 * [0] block-column width (default 256), [1] narrowest front that is factored by a group (default 1024).
 * spchol_factor_device / _sync / _export_csc_device / _lbuf work as for a single-GPU plan; every
 * rank must call them in the same order.  Exported CSC values are 0 where another rank holds the
 * entry (sum the ranks' vectors); info is already the minimum over all ranks. */
int  spchol_plan_create_dist(const spchol_symbolic*, spchol_comm*, const int* dist_opts, spchol_plan** out);
/* dist_stats[0..7] = rank, world, fronts held, fronts factored by a group, #transfers, doubles moved
 * per factorization by this rank, block-column width, doubles this rank reads back in spchol_factor_lbuf (the fronts it
 * holds alone + the block columns it owns of the fronts its group factors; with pinned host memory the other block
 * columns of such a front are left untouched in Lbuf_host -- their owners deliver them); rank_load[world] = estimated flops per rank */
int  spchol_plan_dist_stats(const spchol_plan*, int64_t* dist_stats, double* rank_load);
/* host-only: the mapping a `world`-rank plan would use (no GPU, no NCCL): own_lo/own_n per relaxed front */
int  spchol_map_fronts(const spchol_symbolic*, int world, int dist_min_w, int* own_lo, int* own_n, double* rank_load);

/* Host-buffer call (what chol(A,S) binds): H2D of Ax, factor, CSC gather, D2H of Lx.
 * info = 0, or (1-based) a column whose pivot was <= 0 -> "A is not positive definite." (chol.hpp:849-850). */
int  spchol_factor(spchol_plan*, const double* Ax_host, double* Lx_host, int* info);
/* Host values in, factor stays on the device (supernodal panels) for spchol_solve / spchol_residual / spchol_lbuf*:
 * H2D of Ax, factor, info back.  The handle-returning form of chol(A, S) for factors too large to read back. */
int  spchol_factor_resident(spchol_plan*, const double* Ax_host, int* info);
/* Batched numeric factorization: nbatch value sets over ONE pattern (BASELINE config 5; the reference
 * would loop chol(A_b), include/chol.hpp:750).  Ax[nbatch][nnzA] -> Lx[nbatch][nnzL] in the CSC order of S,
 * info[nbatch] = 0 or the 1-based failing column of that matrix.  Every launch of the level program covers
 * all matrices of a chunk (chunks sized to the free HBM).  Single-GPU plans whose fronts all fit the
 * shared-memory front kernel (m <= 128, e.g. 1138_bus); across GPUs the batch index is sharded by the caller. */
int  spchol_factor_batched(spchol_plan*, int nbatch, const double* Ax_host, double* Lx_host, int* info_host);
int  spchol_factor_batched_device(spchol_plan*, int nbatch, const double* dAx, double* dLx, int* dinfo);

/* Device-resident call: dAx is a device pointer (nnzA doubles); L stays on the device in
 * supernodal panels.  Asynchronous on the plan's stream; *info is valid after spchol_sync. */
int  spchol_factor_device(spchol_plan*, const double* dAx);
int  spchol_export_csc_device(spchol_plan*, double* dLx);      /* gather panels -> CSC order (device ptr) */
int  spchol_sync(spchol_plan*, int* info);
/* Host-buffer call for patterns with nnz(L) > INT_MAX: L is returned as the supernodal panel
 * buffer (lbuf_entries doubles, layout described by spchol_plan_stats / DESIGN.md) */
int  spchol_factor_lbuf(spchol_plan*, const double* Ax_host, double* Lbuf_host, int* info);
/* pinned host memory for the host-buffer calls (cudaHostAlloc / cudaFreeHost) */
int  spchol_host_alloc(int64_t bytes, void** out);
int  spchol_host_free(void* ptr);
/* measured FP64 tensor-pipe (DMMA) peak of the device in TFLOP/s: the roofline denominator */
int  spchol_fp64_peak(int device, double* tflops);
/* pointer to the device L buffer (supernodal panels), for callers that keep L resident */
int  spchol_lbuf(spchol_plan*, double** dL, int64_t* entries);
/* copy the supernodal L buffer to host (entries doubles) */
int  spchol_lbuf_to_host(spchol_plan*, double* L_host);
/* timing of the last spchol_factor call, milliseconds: [h2d, numeric, export, d2h] */
int  spchol_last_timing(const spchol_plan*, double* ms4);
/* per-op-kind device timing (CUDA events around every launch group; adds a sync, so use it on an
 * extra untimed step): kinds 0 small fronts, 1 zero, 2 extend-add, 3 potrf, 4 trsm, 5 gemm/syrk (DMMA),
 * 6 inter-GPU transfers (NCCL), 7 pack/unpack copies of transferred block columns, 8 memset+assemble
 * (+ the info all-reduce of a multi-GPU plan); arrays have 12 slots.  flops2[0] = algorithmic flops of the kind-5 launches (lower trapezoids only),
 * flops2[1] = flops of their full rectangles. */
int  spchol_set_profile(spchol_plan*, int on);
int  spchol_get_profile(const spchol_plan*, double* kind_ms12, int64_t* kind_launches12, double* flops2);
/* algorithmic flops of the DMMA update tasks by K class (K <= 128, <= 256, larger); their times are profile slots 9..11 */
int  spchol_get_gemm_classes(const spchol_plan*, double* flops3);
/* operand bytes of the update tasks of one factorization if every operand moved exactly once, and the number of tasks */
int  spchol_plan_gemm_bytes(const spchol_plan*, double* bytes_min, int64_t* tasks);
/* stream handle (cudaStream_t) the plan launches on, for CUDA-event timing by the caller */
void* spchol_stream(spchol_plan*);
/* number of kernel launches the last factorization issued */
int64_t spchol_last_launches(const spchol_plan*);

/* ||A - L L^T||_F / ||A||_F of the factor the plan holds (the last spchol_factor* call), for the values Ax_host
 * that were factored.  Estimated on the device with 8 Rademacher probes z: sqrt(mean ||(A - L L^T) z||^2) / ||A||_F
 * (E||Rz||^2 = ||R||_F^2), using two panel-times-vectors passes over the supernodal L buffer -- no CSC, so it also
 * works when nnz(L) > INT_MAX (128^3 Laplacian), where the reference's dense comparison (tests/test_chol.cpp:87)
 * cannot run.  Multi-GPU plans: collective (every rank calls with the same Ax); each rank multiplies the panels of the
 * fronts it holds alone or leads, the probe products are summed with ncclAllReduce, every rank gets the same value. */
int  spchol_residual(spchol_plan*, const double* Ax_host, double* rel_fro);

/* x = A^-1 b with the factor the plan holds: L y = b, then L^T x = y, on the device-resident supernodal panels
 * (level by level; L is streamed once per direction).  b, x: n x nrhs column-major host arrays (may alias).
 * The reference stops at the factor (it has no solve).  Multi-GPU plans: collective (every rank calls
 * with the same b and receives the whole x); a front is solved by the rank that holds it alone or leads its group, the
 * subtrees below the split need no communication, the columns of the shared top of the tree are exchanged once per
 * level (ncclAllReduce of a packed buffer), the pieces of x are summed at the end. */
int  spchol_solve(spchol_plan*, const double* b_host, double* x_host, int nrhs);

#ifdef __cplusplus
}
#endif
#endif /* SPCHOL_H */
