// Host-only part of the C-ABI (ingest + symbolic); see include/spchol.h for the contract and
// the reference interfaces each entry point replaces.
#include <charconv>
#include <climits>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <new>
#include <string>
#include <vector>

#include "../../include/spchol.h"
#include "capi_internal.h"
#include "ordering.h"
#include "plan.h"

namespace spchol {
static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
}  // namespace spchol
using namespace spchol;

extern "C" {

const char* spchol_last_error(void) { return g_error.c_str(); }
const char* spchol_version(void) { return "spchol-b200 0.1 (sm_100a)"; }

int spchol_triplet_to_csc(int64_t nz, const int* ti, const int* tj, const double* tx, int n,
                          int* Ap, int* Ai, double* Ax, int64_t* nnz_out) {
    if (n <= 0 || nz < 0 || !Ap || (nz > 0 && (!ti || !tj || !tx || !Ai || !Ax))) {
        set_error("spchol_triplet_to_csc: bad argument"); return SPCHOL_EINVAL;
    }
    const int64_t m = triplet_to_csc(nz, ti, tj, tx, n, Ap, Ai, Ax);
    if (m < 0) { set_error("spchol_triplet_to_csc: index out of range"); return SPCHOL_EINVAL; }
    if (m > INT_MAX) { set_error("spchol_triplet_to_csc: nnz exceeds int"); return SPCHOL_ERANGE; }
    if (nnz_out) *nnz_out = m;
    return SPCHOL_OK;
}

// Matrix Market coordinate reader with the reference's rules (include/mtx_reader.hpp:17-62):
// leading '%' lines skipped, then "rows cols nnz", then nnz 1-based (row col value) triples,
// each mirrored into the upper triangle.  The banner is not interpreted.
// The reference reads the file with `ifstream >>` (mtx_reader.hpp:17-62): leading '%' lines are skipped,
// then whitespace-separated tokens "rows cols nnz" and nnz triples "r c v" (1-based).  Same grammar here,
// parsed from one in-memory copy of the file (from_chars: ~15x the throughput of the stream extractors).
namespace {
struct Cursor {
    const char* p; const char* end;
    void skip_ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r' || *p == '\f' || *p == '\v')) ++p; }
    bool integer(long long& v) {
        skip_ws();
        if (p < end && *p == '+') ++p;
        const auto r = std::from_chars(p, end, v);
        if (r.ec != std::errc()) return false;
        p = r.ptr; return true;
    }
    bool real(double& v) {
        skip_ws();
        if (p < end && *p == '+') ++p;
        const auto r = std::from_chars(p, end, v);
        if (r.ec == std::errc::result_out_of_range) {            // strtod semantics for overflow / underflow
            std::string tok(p, r.ptr);
            v = std::strtod(tok.c_str(), nullptr);
        } else if (r.ec != std::errc()) return false;
        p = r.ptr; return true;
    }
};
}  // namespace

int spchol_mtx_read(const char* path, int* n_out, int64_t* nz_out, int* ti, int* tj, double* tx) {
    if (!path || !n_out || !nz_out) { set_error("spchol_mtx_read: null argument"); return SPCHOL_EINVAL; }
    std::FILE* f = std::fopen(path, "rb");
    if (!f) { set_error(std::string("Could not open file ") + path); return SPCHOL_EIO; }
    std::vector<char> buf;
    if (!ti) {
        // sizes only: the header sits in the first few KB unless the comment block is long
        buf.resize(1 << 16);
        size_t got = std::fread(buf.data(), 1, buf.size(), f);
        auto header_complete = [&]() {
            size_t q = 0;
            while (q < got && buf[q] == '%') { while (q < got && buf[q] != '\n') ++q; if (q < got) ++q; }
            int tokens = 0; bool in_tok = false;
            for (; q < got; ++q) {
                const bool ws = buf[q] == ' ' || buf[q] == '\n' || buf[q] == '\t' || buf[q] == '\r';
                if (!ws && !in_tok) { in_tok = true; }
                else if (ws && in_tok) { in_tok = false; if (++tokens == 3) return true; }
            }
            return false;
        };
        while (!header_complete() && got == buf.size()) {
            buf.resize(buf.size() * 2);
            got += std::fread(buf.data() + got, 1, buf.size() - got, f);
        }
        buf.resize(got);
    } else {
        std::fseek(f, 0, SEEK_END);
        const long sz = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        if (sz < 0) { std::fclose(f); set_error("spchol_mtx_read: cannot size the file"); return SPCHOL_EIO; }
        try { buf.resize(static_cast<size_t>(sz)); } catch (const std::bad_alloc&) { std::fclose(f); set_error("out of host memory"); return SPCHOL_ENOMEM; }
        const size_t got = std::fread(buf.data(), 1, buf.size(), f);
        buf.resize(got);
    }
    std::fclose(f);
    Cursor c{buf.data(), buf.data() + buf.size()};
    while (c.p < c.end && *c.p == '%') { while (c.p < c.end && *c.p != '\n') ++c.p; if (c.p < c.end) ++c.p; }
    long long nr = 0, nc = 0, nl = 0;
    if (!c.integer(nr) || !c.integer(nc) || !c.integer(nl) || nr <= 0 || nr != nc || nl < 0 || nr > INT_MAX) {
        set_error("spchol_mtx_read: bad size line (matrix must be square)"); return SPCHOL_EINVAL;
    }
    *n_out = static_cast<int>(nr);
    *nz_out = nl;
    if (!ti) return SPCHOL_OK;
    for (long long l = 0; l < nl; ++l) {
        long long r, col; double v;
        if (!c.integer(r) || !c.integer(col) || !c.real(v)) { set_error("spchol_mtx_read: truncated file"); return SPCHOL_EIO; }
        --r; --col;
        if (col < r) std::swap(r, col);
        ti[l] = static_cast<int>(r); tj[l] = static_cast<int>(col); tx[l] = v;
    }
    return SPCHOL_OK;
}

int spchol_etree(int n, const int* Ap, const int* Ai, int* parent) {
    if (n <= 0 || !Ap || !Ai || !parent) return SPCHOL_EINVAL;
    etree(n, Ap, Ai, parent); return 0;
}
int spchol_post_order(int n, const int* parent, int* post) {
    if (n <= 0 || !parent || !post) return SPCHOL_EINVAL;
    post_order(n, parent, post); return 0;
}
int spchol_col_count(int n, const int* Ap, const int* Ai, const int* parent, const int* post, int* cc) {
    if (n <= 0 || !Ap || !Ai || !parent || !post || !cc) return SPCHOL_EINVAL;
    col_count(n, Ap, Ai, parent, post, cc); return 0;
}
int spchol_ereach(int n, const int* Ap, const int* Ai, const double* Ax, int k, const int* parent,
                  int* s, int* w, double* x, int top) {
    return ereach(n, Ap, Ai, Ax, k, parent, s, w, x, top);
}
int spchol_compute_levels(int n, const int* parent, int* level_ptr, int* level_nodes) {
    return compute_levels(n, parent, level_ptr, level_nodes);
}
int spchol_compute_supernodes(int n, const int* parent, const int64_t* Lp, int* sn_id, int* sn_start) {
    return compute_supernodes(n, parent, Lp, sn_id, sn_start);
}
int spchol_atree(int n, const int* parent, const int* sn_id, int ns, const int* sn_start, int* sp) {
    atree_from_etree(n, parent, sn_id, ns, sn_start, sp); return 0;
}

int spchol_symbolic_create_ex(int n, const int* Ap, const int* Ai, int relax, spchol_symbolic** out) {
    if (!out || !Ap || !Ai) { set_error("spchol_symbolic_create: null argument"); return SPCHOL_EINVAL; }
    *out = nullptr;
    spchol_symbolic* s = new (std::nothrow) spchol_symbolic();
    if (!s) return SPCHOL_ENOMEM;
    try {
        RelaxParams rp; rp.enable = relax != 0;
        if (relax == 2) {      // batched use: latency bound, so trade explicit zeros for far fewer, wider fronts
            rp.nrelax[0] = 32; rp.nrelax[1] = 96; rp.nrelax[2] = 128;
            rp.zrelax[0] = 0.9; rp.zrelax[1] = 0.7; rp.zrelax[2] = 0.5;
            rp.max_m = 128;
        }
        if (relax == 3) rp.par_min_w = 1024;   // multi-GPU plans: keep wide sibling separators as fronts of their own
        analyze(n, Ap, Ai, rp, s->S);
    } catch (const std::bad_alloc&) { delete s; set_error("out of host memory"); return SPCHOL_ENOMEM;
    } catch (const std::exception& e) { delete s; set_error(e.what()); return SPCHOL_EINVAL; }
    *out = s;
    return SPCHOL_OK;
}
int spchol_symbolic_create(int n, const int* Ap, const int* Ai, spchol_symbolic** out) {
    return spchol_symbolic_create_ex(n, Ap, Ai, 1, out);
}
void spchol_symbolic_destroy(spchol_symbolic* s) { delete s; }
int spchol_symbolic_n(const spchol_symbolic* s) { return s ? s->S.n : 0; }
int64_t spchol_symbolic_nnzL(const spchol_symbolic* s) { return s ? s->S.nnzL : 0; }
double spchol_symbolic_flops(const spchol_symbolic* s) { return s ? s->S.flops : 0; }
int spchol_symbolic_parent(const spchol_symbolic* s, int* parent) {
    if (!s || !parent) return SPCHOL_EINVAL;
    std::memcpy(parent, s->S.parent.data(), sizeof(int) * s->S.n); return 0;
}
int spchol_symbolic_colcount(const spchol_symbolic* s, int* cc) {
    if (!s || !cc) return SPCHOL_EINVAL;
    std::memcpy(cc, s->S.colcount.data(), sizeof(int) * s->S.n); return 0;
}
int spchol_symbolic_Lp64(const spchol_symbolic* s, int64_t* Lp) {
    if (!s || !Lp) return SPCHOL_EINVAL;
    std::memcpy(Lp, s->S.Lp.data(), sizeof(int64_t) * (s->S.n + 1)); return 0;
}
int spchol_symbolic_Lp(const spchol_symbolic* s, int* Lp) {
    if (!s || !Lp) return SPCHOL_EINVAL;
    if (s->S.nnzL > INT_MAX) { set_error("nnz(L) exceeds the 32-bit CSC of the reference API"); return SPCHOL_ERANGE; }
    for (int j = 0; j <= s->S.n; ++j) Lp[j] = static_cast<int>(s->S.Lp[j]);
    return 0;
}
int spchol_symbolic_Li(const spchol_symbolic* s, int* Li) {
    if (!s || !Li) return SPCHOL_EINVAL;
    if (s->S.nnzL > INT_MAX) { set_error("nnz(L) exceeds the 32-bit CSC of the reference API"); return SPCHOL_ERANGE; }
    s->S.fill_Li(Li); return 0;
}
int spchol_symbolic_supernodes(const spchol_symbolic* s, int which, int* ns_out, int* sn_start, int* sn_parent) {
    if (!s || !ns_out) return SPCHOL_EINVAL;
    const Symbolic& S = s->S;
    const int ns = which ? S.ns : S.nsf;
    *ns_out = ns;
    const std::vector<int>& st = which ? S.sn_start : S.fsn_start;
    const std::vector<int>& pa = which ? S.sn_parent : S.fsn_parent;
    if (sn_start) std::memcpy(sn_start, st.data(), sizeof(int) * (ns + 1));
    if (sn_parent) std::memcpy(sn_parent, pa.data(), sizeof(int) * ns);
    return 0;
}
int spchol_symbolic_stats(const spchol_symbolic* s, int64_t* st) {
    if (!s || !st) return SPCHOL_EINVAL;
    const Symbolic& S = s->S;
    int maxw = 0, maxm = 0;
    std::vector<int> lvl(S.ns, 0);
    int nlev = 0;
    for (int k = 0; k < S.ns; ++k) {
        maxw = std::max(maxw, S.w(k)); maxm = std::max(maxm, S.m(k));
        if (S.sn_parent[k] >= 0) lvl[S.sn_parent[k]] = std::max(lvl[S.sn_parent[k]], lvl[k] + 1);
        nlev = std::max(nlev, lvl[k] + 1);
    }
    st[0] = S.nsf; st[1] = S.ns; st[2] = S.sum_m; st[3] = S.panel_entries; st[4] = S.padded_zeros;
    st[5] = maxw; st[6] = maxm; st[7] = nlev;
    return 0;
}

// Supernodal export format: where the panel of every front sits in the L buffer that
// spchol_factor_lbuf / spchol_lbuf_to_host return for a single-GPU plan, and which rows it has.
int spchol_symbolic_panel_layout(const spchol_symbolic* s, int* w, int* m, int* ld, int64_t* panel_off, int64_t* lbuf_entries) {
    if (!s) return SPCHOL_EINVAL;
    const Symbolic& S = s->S;
    int64_t off = 0;
    for (int k = 0; k < S.ns; ++k) {
        const int mk = S.m(k), wk = S.w(k), ldk = (mk + 1) & ~1;
        if (w) w[k] = wk;
        if (m) m[k] = mk;
        if (ld) ld[k] = ldk;
        if (panel_off) panel_off[k] = off;
        off += static_cast<int64_t>(ldk) * wk;
    }
    if (lbuf_entries) *lbuf_entries = off;
    return 0;
}
int spchol_symbolic_front_rows(const spchol_symbolic* s, int front, int* rows) {
    if (!s || !rows || front < 0 || front >= s->S.ns) { set_error("spchol_symbolic_front_rows: bad argument"); return SPCHOL_EINVAL; }
    const Symbolic& S = s->S;
    const int wk = S.w(front), nb = S.below_count(front);
    for (int c = 0; c < wk; ++c) rows[c] = S.sn_start[front] + c;
    std::memcpy(rows + wk, S.below(front), sizeof(int) * nb);
    return 0;
}

int64_t spchol_symbolic_nnzA(const spchol_symbolic* s) { return s ? s->S.nnzA : 0; }

// Supernodal panels (single-GPU layout, spchol_symbolic_panel_layout) -> CSC of L with 64-bit column pointers: the
// container the reference cannot offer (its p_/i_ are std::vector<int>, include/chol.hpp:52,55, so nnz(L) <= INT_MAX).
// Layout as the reference's L: diagonal first, rows ascending (:844-855); only the TRUE pattern is emitted, the
// explicit zeros of relaxed amalgamation are dropped.  Any of Lp / Li / Lx may be NULL.  Host side, one pass.
int spchol_symbolic_gather_csc64(const spchol_symbolic* sp, const double* lbuf, int64_t* Lp, int* Li, double* Lx) {
    if (!sp) { set_error("spchol_symbolic_gather_csc64: null symbolic"); return SPCHOL_EINVAL; }
    if (Lx && !lbuf) { set_error("spchol_symbolic_gather_csc64: values wanted but no panel buffer given"); return SPCHOL_EINVAL; }
    const Symbolic& S = sp->S;
    if (Lp) std::memcpy(Lp, S.Lp.data(), sizeof(int64_t) * (static_cast<size_t>(S.n) + 1));
    if (!Li && !Lx) return 0;
    try {
        const std::vector<int>& fbelow = S.fundamental_below();
        int64_t po = 0;
        for (int s = 0; s < S.ns; ++s) {
            const int w = S.w(s), m = S.m(s), ld = (m + 1) & ~1, first = S.sn_start[s];
            const int* b = S.below(s);
            const int nbr = m - w, send = first + w;
            for (int c = 0; c < w; ++c) {
                const int j = first + c, f = S.fsn_id[j];
                int64_t q = S.Lp[j];
                const double* col = lbuf ? lbuf + po + static_cast<int64_t>(c) * ld : nullptr;
                const int fend = S.fsn_start[f + 1];
                for (int r = j; r < fend; ++r, ++q) { if (Li) Li[q] = r; if (Lx) Lx[q] = col[r - first]; }
                const int* fb = fbelow.data() + S.fbelow_ptr[f];
                const int nfb = static_cast<int>(S.fbelow_ptr[f + 1] - S.fbelow_ptr[f]);
                int k = 0;
                for (int t = 0; t < nfb; ++t, ++q) {
                    const int r = fb[t];
                    if (Li) Li[q] = r;
                    if (!Lx) continue;
                    if (r < send) { Lx[q] = col[r - first]; continue; }
                    while (k < nbr && b[k] < r) ++k;
                    if (k >= nbr || b[k] != r) { set_error("spchol_symbolic_gather_csc64: pattern row missing in its front"); return SPCHOL_EINVAL; }
                    Lx[q] = col[w + k];
                }
            }
            po += static_cast<int64_t>(ld) * w;
        }
    } catch (const std::exception& e) { set_error(e.what()); return SPCHOL_EINVAL; }
    return 0;
}

static int dryrun_impl(const spchol_symbolic* s, int rank, int world, const int* dist_opts, int64_t* st,
                       int64_t* extra, int nextra, int64_t* comm_sig) {
    if (!s || !st) return SPCHOL_EINVAL;
    try {
        HostPlan H; PlanOptions opt;
        opt.want_csc = s->S.nnzL <= INT_MAX ? 1 : 0;
        opt.rank = rank; opt.world = world;
        if (dist_opts) {
            if (dist_opts[0] > 0) opt.dist_NB = dist_opts[0];
            if (dist_opts[1] > 0) opt.dist_min_w = dist_opts[1];
        }
        build_plan(s->S, opt, H);
        st[0] = H.ns; st[1] = H.nlevels; st[2] = H.n_small; st[3] = H.n_large; st[4] = H.lbuf_entries;
        st[5] = H.ws_entries; st[6] = static_cast<int64_t>(H.ops.size()); st[7] = H.max_m; st[8] = H.max_w;
        st[9] = 8 * (H.lbuf_entries + H.ws_entries) + 8 * static_cast<int64_t>(H.a_dest.size() + H.l_src.size()) +
                4 * static_cast<int64_t>(H.rel.size()) + 60 * static_cast<int64_t>(H.ns);
        if (extra) {
            for (int k = 0; k < nextra; ++k) extra[k] = 0;
            for (const Op& op : H.ops) {
                if (op.kind == OP_SMALL) ++extra[0];
                else if (op.kind == OP_EXTADD) ++extra[1];
                else if (op.kind == OP_POTRF) ++extra[2];
                else if (op.kind == OP_TRSM) ++extra[3];
                else if (op.kind == OP_GEMM) ++extra[4];
            }
            extra[5] = static_cast<int64_t>(H.gemm_flops);
            if (nextra > 7) { extra[6] = static_cast<int64_t>(H.comm_tasks.size()); extra[7] = H.comm_doubles; }
        }
        if (comm_sig) {
            comm_sig[0] = comm_sig[1] = comm_sig[2] = 0;
            uint64_t h = 1469598103934665603ull;
            auto mix = [&](uint64_t v) { h ^= v; h *= 1099511628211ull; };
            for (const CommTask& t : H.comm_tasks) {
                ++comm_sig[t.kind];
                // what every participant must agree on: kind of exchange, group/pair, root, length
                if (t.kind == COMM_BCAST) { mix(1); mix(t.lo); mix(t.n); mix(t.root); }
                else { mix(2); mix(std::min(t.root, rank)); mix(std::max(t.root, rank)); }
                mix(static_cast<uint64_t>(t.count));
            }
            comm_sig[3] = static_cast<int64_t>(h & 0x7fffffffffffffffull);
        }
    } catch (const std::exception& e) { set_error(e.what()); return SPCHOL_EINVAL; }
    return 0;
}

int spchol_plan_dryrun(const spchol_symbolic* s, int64_t* st, int64_t* extra) {
    return dryrun_impl(s, 0, 1, nullptr, st, extra, 6, nullptr);
}

int spchol_plan_dryrun_dist(const spchol_symbolic* s, int rank, int world, const int* dist_opts, int64_t* st,
                            int64_t* extra, int64_t* comm_sig) {
    if (world < 1 || rank < 0 || rank >= world) { set_error("spchol_plan_dryrun_dist: bad rank/world"); return SPCHOL_EINVAL; }
    return dryrun_impl(s, rank, world, dist_opts, st, extra, 8, comm_sig);
}

int spchol_map_fronts(const spchol_symbolic* s, int world, int dist_min_w, int* own_lo, int* own_n, double* rank_load) {
    if (!s || world < 1) return SPCHOL_EINVAL;
    const Symbolic& S = s->S;
    std::vector<int> w(S.ns), m(S.ns), lo, cnt;
    std::vector<double> load;
    for (int k = 0; k < S.ns; ++k) { w[k] = S.w(k); m[k] = S.m(k); }
    map_fronts(S.ns, S.sn_parent.data(), w.data(), m.data(), world, dist_min_w > 0 ? dist_min_w : PlanOptions().dist_min_w,
               lo, cnt, load);
    for (int k = 0; k < S.ns; ++k) { if (own_lo) own_lo[k] = lo[k]; if (own_n) own_n[k] = cnt[k]; }
    if (rank_load) for (int r = 0; r < world; ++r) rank_load[r] = load[r];
    return 0;
}

// ---- orderings (ordering.h) -------------------------------------------------------------------
int spchol_nd_perm_grid(int nx, int ny, int nz, int64_t* perm) {
    if (!perm || nx < 1 || ny < 1 || nz < 1) { set_error("spchol_nd_perm_grid: bad argument"); return SPCHOL_EINVAL; }
    try { nd_perm_grid(nx, ny, nz, perm); } catch (const std::exception& e) { set_error(e.what()); return SPCHOL_EINVAL; }
    return 0;
}

int spchol_order_amd(int n, const int* Ap, const int* Ai, int* perm) {
    if (n < 0 || !Ap || (!Ai && Ap[n] > 0) || (!perm && n > 0)) { set_error("spchol_order_amd: bad argument"); return SPCHOL_EINVAL; }
    try { order_amd(n, Ap, Ai, perm); }
    catch (const std::bad_alloc&) { set_error("out of host memory"); return SPCHOL_ENOMEM; }
    catch (const std::exception& e) { set_error(e.what()); return SPCHOL_EINVAL; }
    return 0;
}

int spchol_permute_upper(int n, const int* Ap, const int* Ai, const double* Ax, const int* perm,
                         int* Bp, int* Bi, double* Bx, int64_t* map) {
    if (n < 0 || !Ap || !perm || !Bp || (Ap[n] > 0 && (!Ai || !Bi))) { set_error("spchol_permute_upper: bad argument"); return SPCHOL_EINVAL; }
    try { permute_upper(n, Ap, Ai, Ax, perm, Bp, Bi, Bx, map); }
    catch (const std::bad_alloc&) { set_error("out of host memory"); return SPCHOL_ENOMEM; }
    catch (const std::exception& e) { set_error(e.what()); return SPCHOL_EINVAL; }
    return 0;
}

}  // extern "C"
