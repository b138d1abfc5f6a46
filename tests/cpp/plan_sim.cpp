// TEST INFRASTRUCTURE ONLY -- never linked into libspchol.so.
//
// Host interpreter of the numeric launch program (csrc/plan.h): builds the plan of every rank of a
// `world`-rank factorization with the product's own planner and executes the ops with plain loops,
// one simulated rank per program counter, transfers resolved when all their participants have
// reached them (a mismatch or a cycle is reported as an error).  It lets the CPU-only test run
// check the planner -- front mapping, relative indices, block-column ownership, transfer order,
// workspace lifetimes -- against the oracle without a GPU.  The op semantics restate this repo's
// own kernels (csrc/kernels.cuh), not reference code.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../sparsecholesky_b200/csrc/plan.h"
#include "../../sparsecholesky_b200/csrc/symbolic.h"

using namespace spchol;

namespace {

inline int even_up(int v) { return (v + 1) & ~1; }

struct Rank {
    HostPlan H;
    std::vector<double> L, W;
    std::vector<double> early;    // what a read-back started at the OP_EXPORT marks would have seen (NaN elsewhere)
    // one program counter per stream (lane): q[s] lists the indices of the ops of stream s in program order
    std::vector<size_t> q[2];
    size_t pc[2] = {0, 0};
    int cl = -1;              // lane whose head is the OP_COMM being resolved
    std::vector<char> ev_done;   // per event id: its OP_EV_RECORD has executed
    std::vector<char> done;   // per task of the current OP_COMM (grouped sends/recvs complete in any order)
    bool finished() const { return pc[0] >= q[0].size() && pc[1] >= q[1].size(); }
    int info = std::numeric_limits<int>::max();
};

void run_small(Rank& R, int s) {
    const HostPlan& H = R.H;
    const int w = H.w[s], m = H.m[s], ld = H.ld[s];
    std::vector<double> F(static_cast<size_t>(m) * m, 0.0);
    double* P = R.L.data() + H.panel_off[s];
    for (int j = 0; j < w; ++j) for (int i = j; i < m; ++i) F[i + static_cast<size_t>(j) * m] = P[i + static_cast<size_t>(j) * ld];
    for (int q = H.child_ptr[s]; q < H.child_ptr[s + 1]; ++q) {
        const int c = H.child_idx[q];
        const int mc = H.m[c] - H.w[c];
        if (!mc) continue;
        const int ldc = even_up(mc);
        const double* C = R.W.data() + H.cb_off[c];
        const int* r = H.rel.data() + H.rel_off[c];
        for (int jj = 0; jj < mc; ++jj) for (int ii = jj; ii < mc; ++ii) F[r[ii] + static_cast<size_t>(r[jj]) * m] += C[ii + static_cast<size_t>(jj) * ldc];
    }
    for (int k = 0; k < w; ++k) {
        const double d = F[k + static_cast<size_t>(k) * m];
        if (d <= 0.0) R.info = std::min(R.info, H.first[s] + k + 1);
        const double piv = std::sqrt(d);
        for (int i = k + 1; i < m; ++i) F[i + static_cast<size_t>(k) * m] /= piv;
        F[k + static_cast<size_t>(k) * m] = piv;
        for (int j = k + 1; j < m; ++j) {
            const double uj = F[j + static_cast<size_t>(k) * m];
            for (int i = j; i < m; ++i) F[i + static_cast<size_t>(j) * m] -= F[i + static_cast<size_t>(k) * m] * uj;
        }
    }
    for (int j = 0; j < w; ++j) for (int i = j; i < m; ++i) P[i + static_cast<size_t>(j) * ld] = F[i + static_cast<size_t>(j) * m];
    const int mb = m - w;
    if (mb > 0) {
        double* C = R.W.data() + H.cb_off[s];
        const int ldc = even_up(mb);
        for (int j = 0; j < mb; ++j) for (int i = j; i < mb; ++i) C[i + static_cast<size_t>(j) * ldc] = F[(w + i) + static_cast<size_t>(w + j) * m];
    }
}

void run_extadd(Rank& R, int c, int mode) {
    const HostPlan& H = R.H;
    const int p = H.parent[c];
    const int mc = H.m[c] - H.w[c], ldc = even_up(mc);
    const int wp = H.w[p], ldp = H.ld[p], ldcb = even_up(H.m[p] - wp);
    const double* C = R.W.data() + H.cb_off[c];
    const int* r = H.rel.data() + H.rel_off[c];
    double* Pp = R.L.data() + H.panel_off[p];
    double* Cp = H.cb_off[p] >= 0 ? R.W.data() + H.cb_off[p] : nullptr;
    const int G = H.own_n[p], me = H.own_idx[p], DNB = H.dist_NB, nbp = (wp + DNB - 1) / DNB;
    for (int jj = 0; jj < mc; ++jj) {
        const int cj = r[jj];
        if ((mode == 1 && cj >= wp) || (mode == 2 && cj < wp)) continue;
        if (G > 1) {
            const int blk = cj < wp ? cj / DNB : nbp + (cj - wp) / DNB;
            if (blk % G != me) continue;
        }
        for (int ii = jj; ii < mc; ++ii) {
            const double v = C[ii + static_cast<size_t>(jj) * ldc];
            if (cj < wp) Pp[r[ii] + static_cast<size_t>(cj) * ldp] += v;
            else Cp[(r[ii] - wp) + static_cast<size_t>(cj - wp) * ldcb] += v;
        }
    }
}

void run_potrf(Rank& R, const PotrfTask& t) {
    double* D = R.L.data() + t.d_off;
    const int nb = t.nb; const size_t ld = t.ld;
    for (int k = 0; k < nb; ++k) {
        const double d = D[k + k * ld];
        if (d <= 0.0) R.info = std::min(R.info, t.col0 + k + 1);
        const double piv = std::sqrt(d);
        for (int i = k + 1; i < nb; ++i) D[i + k * ld] /= piv;
        D[k + k * ld] = piv;
        for (int j = k + 1; j < nb; ++j) for (int i = j; i < nb; ++i) D[i + j * ld] -= D[i + k * ld] * D[j + k * ld];
    }
}

void run_trsm(Rank& R, const TrsmTask& t) {
    const double* D = R.L.data() + t.d_off;
    double* B = R.L.data() + t.b_off;
    const size_t ld = t.ld;
    for (int row = 0; row < t.rows; ++row)
        for (int k = 0; k < t.nb; ++k) {
            const double v = B[row + k * ld] / D[k + k * ld];
            B[row + k * ld] = v;
            for (int j = k + 1; j < t.nb; ++j) B[row + j * ld] -= v * D[j + k * ld];
        }
}

void run_gemm(Rank& R, const GemmTask& t) {
    const double* A = R.L.data() + t.a_off;
    const double* B = R.L.data() + t.b_off;
    double* C = ((t.c_space & 1) ? R.W.data() : R.L.data()) + t.c_off;
    // read all operands first: A/B may alias rows of C's panel only outside the written region
    for (int j = 0; j < t.N; ++j)
        for (int i = 0; i < t.M; ++i) {
            if (i + t.diag < j) continue;
            double acc = 0;
            for (int k = 0; k < t.K; ++k) acc += A[i + static_cast<size_t>(k) * t.lda] * B[j + static_cast<size_t>(k) * t.ldb];
            if (t.c_space & 2) C[i + static_cast<size_t>(j) * t.ldc] = -acc; else C[i + static_cast<size_t>(j) * t.ldc] -= acc;
        }
}

std::string describe(const CommTask& t, int rank) {
    char buf[160];
    std::snprintf(buf, sizeof buf, "rank %d: kind=%d group=(%d,%d) root/peer=%d space=%d count=%lld", rank, t.kind, t.lo, t.n,
                  t.root, t.space, static_cast<long long>(t.count));
    return buf;
}

}  // namespace

// order: 0 = replay in program order (what a single stream does); 1 = let stream 1 run ahead as far
// as its event waits allow; 2 = let stream 0 run ahead.  All three must give the same factor: a
// missing event dependency in the look-ahead program shows up as a wrong result in one of them.
extern "C" int plansim_factor(int n, const int* Ap, const int* Ai, const double* Ax, int relax, int world,
                              int dist_NB, int dist_min_w, int small_max_m, int nb, int NB, int order,
                              double* Lx, int* info_out, int64_t* stats /* 6: #dist fronts, #transfers, doubles moved, max ws, #stream events, #read-back marks */,
                              char* err, int errlen) {
    auto fail = [&](const std::string& msg) { std::snprintf(err, errlen, "%s", msg.c_str()); return -1; };
    try {
        Symbolic S;
        RelaxParams rp; rp.enable = relax != 0;
        if (relax == 3) rp.par_min_w = 8;         // tree-parallel amalgamation at test scale
        analyze(n, Ap, Ai, rp, S);
        std::vector<Rank> R(world);
        for (int r = 0; r < world; ++r) {
            PlanOptions opt;
            opt.rank = r; opt.world = world;
            opt.tree = 0;       // the warp-per-subtree tasks have their own host run (tests/cpp/tree_emu.cpp); here every front goes through the level program
            if (dist_NB > 0) opt.dist_NB = dist_NB;
            if (dist_min_w > 0) opt.dist_min_w = dist_min_w;
            if (small_max_m > 0) opt.small_max_m = small_max_m;
            if (nb > 0) opt.nb = nb;
            if (NB > 0) { opt.NB = NB; opt.NB_wide = 2 * NB; opt.wide_w = 3 * NB; }    // both outer-block widths get exercised
            if (NB > 0) opt.export_mark_w = 2 * NB;                                    // read-back marks at test scale
            build_plan(S, opt, R[r].H);
            const HostPlan& H = R[r].H;
            for (size_t i = 0; i < H.ops.size(); ++i) R[r].q[H.ops[i].stream & 1].push_back(i);
            R[r].ev_done.assign(static_cast<size_t>(H.n_events) + 1, 0);
            // poison instead of zero where the engine does not clear: the workspace is never memset
            R[r].L.assign(static_cast<size_t>(H.lbuf_entries), 0.0);
            R[r].early.assign(static_cast<size_t>(H.lbuf_entries), std::numeric_limits<double>::quiet_NaN());
            R[r].W.assign(static_cast<size_t>(H.ws_entries), std::numeric_limits<double>::quiet_NaN());
            for (int64_t i = 0; i < H.nnzA; ++i) if (H.a_dest[i] >= 0) R[r].L[H.a_dest[i]] = Ax[i];
        }
        if (stats) { stats[0] = stats[1] = stats[2] = stats[3] = stats[4] = stats[5] = 0; }
        for (int r = 0; r < world; ++r) if (stats) {
            stats[0] = std::max<int64_t>(stats[0], R[r].H.n_dist);
            stats[1] += static_cast<int64_t>(R[r].H.comm_tasks.size());
            stats[2] += R[r].H.comm_doubles;
            stats[3] = std::max<int64_t>(stats[3], R[r].H.ws_entries);
            stats[4] = std::max<int64_t>(stats[4], R[r].H.n_events);
        }
        auto head = [&](const Rank& k, int lane) -> const Op* {
            return k.pc[lane] < k.q[lane].size() ? &k.H.ops[k.q[lane][k.pc[lane]]] : nullptr;
        };
        // lanes of rank r in the order the policy wants them tried
        auto lane_order = [&](const Rank& k, int* lanes) {
            if (order == 0) {      // program order: only the lane holding the earliest pending op may move
                const size_t i0 = k.pc[0] < k.q[0].size() ? k.q[0][k.pc[0]] : SIZE_MAX;
                const size_t i1 = k.pc[1] < k.q[1].size() ? k.q[1][k.pc[1]] : SIZE_MAX;
                lanes[0] = i0 <= i1 ? 0 : 1; lanes[1] = -1;
            } else if (order == 1) { lanes[0] = 1; lanes[1] = 0; }
            else { lanes[0] = 0; lanes[1] = 1; }
        };
        auto comm_op = [&](int r) -> const Op* {
            Rank& k = R[r];
            int lanes[2]; lane_order(k, lanes);
            for (int li = 0; li < 2; ++li) {
                const int lane = lanes[li];
                if (lane < 0) continue;
                const Op* op = head(k, lane);
                if (!op || op->kind != OP_COMM) continue;
                if (k.cl != lane) { k.cl = lane; k.done.clear(); }
                if (k.done.size() != static_cast<size_t>(op->count)) k.done.assign(op->count, 0);
                return op;
            }
            return nullptr;
        };
        auto finish_if_done = [&](int r) {
            Rank& k = R[r];
            for (char d : k.done) if (!d) return;
            k.done.clear(); ++k.pc[k.cl]; k.cl = -1;
        };
        auto buf = [&](int r, const CommTask& t) { return (t.space ? R[r].W.data() : R[r].L.data()) + t.off; };
        auto describe_rank = [&](int r) -> std::string {
            const Op* op = comm_op(r);
            if (!op) return "[rank " + std::to_string(r) + (R[r].finished() ? " done]" : " busy]");
            std::string m2 = "[";
            for (int i = 0; i < op->count && i < 4; ++i) if (!R[r].done[i]) m2 += describe(R[r].H.comm_tasks[op->begin + i], r) + "; ";
            return m2 + "]";
        };
        for (;;) {
            bool progress = false, all_done = true;
            for (int r = 0; r < world; ++r) {
                Rank& k = R[r];
                for (;;) {
                    int lanes[2]; lane_order(k, lanes);
                    int lane = -1;
                    for (int li = 0; li < 2 && lane < 0; ++li) {
                        if (lanes[li] < 0) continue;
                        const Op* h = head(k, lanes[li]);
                        if (!h || h->kind == OP_COMM) continue;
                        if (h->kind == OP_EV_WAIT && !k.ev_done[h->begin]) continue;      // blocked on the other stream
                        lane = lanes[li];
                    }
                    if (lane < 0) break;
                    const Op& op = *head(k, lane);
                    const HostPlan& H = k.H;
                    switch (op.kind) {
                    case OP_SMALL: for (int i = 0; i < op.count; ++i) run_small(k, H.small_list[op.begin + i]); break;
                    case OP_ZERO: for (int i = 0; i < op.count; ++i) { const ZeroTask& z = H.zero_tasks[op.begin + i]; std::fill_n(k.W.data() + z.off, z.count, 0.0); } break;
                    case OP_EXTADD: for (int i = 0; i < op.count; ++i) run_extadd(k, H.extadd_list[op.begin + i], op.threads); break;
                    case OP_POTRF: for (int i = 0; i < op.count; ++i) run_potrf(k, H.potrf_tasks[op.begin + i]); break;
                    case OP_TRSM: for (int i = 0; i < op.count; ++i) run_trsm(k, H.trsm_tasks[op.begin + i]); break;
                    case OP_GEMM: for (int i = 0; i < op.count; ++i) run_gemm(k, H.gemm_tasks[op.begin + i]); break;
                    case OP_COPY:
                        for (int i = 0; i < op.count; ++i) {
                            const CopyTask& c = H.copy_tasks[op.begin + i];
                            const double* src = (c.src_space ? k.W.data() : k.L.data()) + c.src_off;
                            double* dst = (c.dst_space ? k.W.data() : k.L.data()) + c.dst_off;
                            for (int j = 0; j < c.cols; ++j) for (int ii = 0; ii < c.rows; ++ii) dst[ii + static_cast<size_t>(j) * c.ldd] = src[ii + static_cast<size_t>(j) * c.lds];
                        }
                        break;
                    case OP_EXPORT:
                        for (int i = 0; i < op.count; ++i) {
                            const ExportTask& t = H.export_tasks[op.begin + i];
                            std::memcpy(k.early.data() + t.off, k.L.data() + t.off, static_cast<size_t>(t.count) * sizeof(double));
                        }
                        break;
                    case OP_EV_RECORD: k.ev_done[op.begin] = 1; break;
                    case OP_EV_WAIT: break;
                    default: return fail("unknown op kind");
                    }
                    ++k.pc[lane]; progress = true;
                }
                if (!k.finished()) all_done = false;
            }
            if (all_done) break;
            for (int r = 0; r < world; ++r) {
                const Op* op = comm_op(r);
                if (!op) continue;
                const CommTask* tasks = R[r].H.comm_tasks.data() + op->begin;
                if (tasks[0].kind == COMM_BCAST) {
                    const CommTask* t = &tasks[0];
                    if (op->count != 1) return fail("broadcasts are not grouped");
                    if (r < t->lo || r >= t->lo + t->n) return fail("broadcast issued outside its group: " + describe(*t, r));
                    if (r != t->lo) continue;             // resolved when visiting the group's first rank
                    bool ready = true;
                    for (int q = t->lo; q < t->lo + t->n && ready; ++q) {
                        const Op* uo = comm_op(q);
                        if (!uo) { ready = false; break; }
                        const CommTask* u = &R[q].H.comm_tasks[uo->begin];
                        if (u->kind != COMM_BCAST || u->lo != t->lo || u->n != t->n) { ready = false; break; }   // busy with another exchange
                        if (u->root != t->root || u->count != t->count)
                            return fail("broadcast mismatch: " + describe(*t, r) + " vs " + describe(*u, q));
                    }
                    if (!ready) continue;
                    const int lo = t->lo, cnt = t->n, root = t->root;
                    const int64_t count = t->count;
                    const double* src = buf(root, R[root].H.comm_tasks[comm_op(root)->begin]);
                    for (int q = lo; q < lo + cnt; ++q)
                        if (q != root) std::memcpy(buf(q, R[q].H.comm_tasks[comm_op(q)->begin]), src, static_cast<size_t>(count) * sizeof(double));
                    for (int q = lo; q < lo + cnt; ++q) { R[q].done[0] = 1; finish_if_done(q); }
                    progress = true;
                    continue;
                }
                // grouped point-to-point: the i-th pending send r -> q pairs with the i-th pending recv q <- r
                for (int i = 0; i < op->count; ++i) {
                    if (R[r].done[i] || tasks[i].kind != COMM_SEND) continue;
                    const int q = tasks[i].root;
                    bool earlier = false;                  // FIFO per pair
                    for (int i2 = 0; i2 < i; ++i2) if (!R[r].done[i2] && tasks[i2].kind == COMM_SEND && tasks[i2].root == q) earlier = true;
                    if (earlier) continue;
                    const Op* uo = comm_op(q);
                    if (!uo) continue;
                    const CommTask* ut = R[q].H.comm_tasks.data() + uo->begin;
                    if (ut[0].kind == COMM_BCAST) continue;
                    for (int i2 = 0; i2 < uo->count; ++i2) {
                        if (R[q].done[i2] || ut[i2].kind != COMM_RECV || ut[i2].root != r) continue;
                        if (ut[i2].count != tasks[i].count)
                            return fail("send/recv length mismatch: " + describe(tasks[i], r) + " vs " + describe(ut[i2], q));
                        std::memcpy(buf(q, ut[i2]), buf(r, tasks[i]), static_cast<size_t>(tasks[i].count) * sizeof(double));
                        R[q].done[i2] = 1; R[r].done[i] = 1;
                        progress = true;
                        break;
                    }
                }
            }
            for (int r = 0; r < world; ++r) if (comm_op(r)) { const size_t before = R[r].pc[0] + R[r].pc[1]; finish_if_done(r); if (R[r].pc[0] + R[r].pc[1] != before) progress = true; }
            if (!progress) {
                std::string msg = "deadlock:";
                for (int r = 0; r < world; ++r) msg += " " + describe_rank(r);
                return fail(msg);
            }
        }
        // read-back marks: what was announced as final must be what the factorization ends with, bit for bit, and a
        // marked front must be announced completely (its level no longer reads it back)
        for (int r = 0; r < world; ++r) {
            const HostPlan& H = R[r].H;
            std::vector<int64_t> covered(H.ns, 0);
            for (const ExportTask& t : H.export_tasks) {
                if (std::memcmp(R[r].early.data() + t.off, R[r].L.data() + t.off, static_cast<size_t>(t.count) * sizeof(double)) != 0)
                    return fail("a range announced by OP_EXPORT changed afterwards (rank " + std::to_string(r) + ", offset " + std::to_string(t.off) + ")");
                for (int s = 0; s < H.ns; ++s)
                    if (H.holds(s) && t.off >= H.panel_off[s] && t.off < H.panel_off[s] + static_cast<int64_t>(H.ld[s]) * H.w[s]) { covered[s] += t.count; break; }
            }
            for (int s = 0; s < H.ns; ++s) {
                const int64_t want = H.front_marked[s] ? static_cast<int64_t>(H.ld[s]) * H.w[s] : 0;
                if (covered[s] != want) return fail("front " + std::to_string(s) + ": OP_EXPORT covers " + std::to_string(covered[s]) + " of " + std::to_string(want) + " entries");
            }
            if (stats) stats[5] += static_cast<int64_t>(H.export_tasks.size());
        }
        int info = std::numeric_limits<int>::max();
        for (int r = 0; r < world; ++r) info = std::min(info, R[r].info);
        *info_out = info == std::numeric_limits<int>::max() ? 0 : info;
        if (Lx) {
            for (int64_t i = 0; i < S.nnzL; ++i) {
                int holders = 0; double v = 0;
                for (int r = 0; r < world; ++r) {
                    const int64_t q = R[r].H.l_src[i];
                    if (q >= 0) { ++holders; v += R[r].L[q]; }
                }
                if (holders != 1) return fail("L entry " + std::to_string(i) + " exported by " + std::to_string(holders) + " ranks");
                Lx[i] = v;
            }
        }
    } catch (const std::exception& e) { return fail(e.what()); }
    return 0;
}
