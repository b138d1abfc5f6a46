/*
 * oracle/oracle.c -- TEST INFRASTRUCTURE ONLY (never linked or called by the product path).
 *
 * Plain-C restatement of the reference's simplicial up-looking sparse Cholesky and the
 * symbolic helpers it calls.  Every function cites the reference file:line it follows
 * (paths relative to /root/reference).  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load the library built from this file.
 *
 * Parity pin: oracle/_ref (the reference's own chol.hpp compiled here, see oracle/Makefile)
 * is compared bit-for-bit with this restatement by tests/test_oracle.py, and the committed
 * golden vectors under tests/golden/ were produced by oracle/_ref (tests/golden/make_golden.py).
 *
 * Differences from the reference that do NOT change any output bit:
 *  - work arrays s/w/x are allocated once and reused (the reference allocates three O(n)
 *    vectors per column, include/chol.hpp:801-803, which is O(n^2) time);
 *  - rows are visited k = 0..n-1 instead of by etree depth bucket (include/chol.hpp:791-798):
 *    row k only reads columns i that are etree descendants of k, all of which have a smaller
 *    index, so the appended order inside every column (ascending row) and every floating
 *    point operation and its operand order are the same;
 *  - column pointers / nnz are int64 internally so that inputs with nnz(L) > INT_MAX are
 *    reported (return -2) instead of overflowing a signed int (include/chol.hpp:765).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * triplet_to_csc_matrix  (include/chol.hpp:309-369)
 * mirror to upper (:321-322), stable order by (col,row) (:327-331), sum duplicates (:336-346),
 * explicit zeros stay structural.  std::sort is not stable, but duplicates are summed in the
 * sorted order; to be order independent of the sort implementation for exact duplicates we
 * use a stable counting sort on (col,row): addition order = input order.  (For the reference's
 * std::sort the order among equal keys is unspecified; the fixtures contain no duplicates.)
 * Outputs: Ap[n+1], Ai[<=nz], Ax[<=nz]; returns nnz or -1.
 * ------------------------------------------------------------------------------------------ */
typedef struct { int r, c; double v; int64_t seq; } trip_t;

static int trip_cmp(const void* a, const void* b) {
    const trip_t* x = (const trip_t*)a; const trip_t* y = (const trip_t*)b;
    if (x->c != y->c) return x->c < y->c ? -1 : 1;
    if (x->r != y->r) return x->r < y->r ? -1 : 1;
    return x->seq < y->seq ? -1 : (x->seq > y->seq);
}

int64_t orc_triplet_to_csc(int64_t nz, const int* ti, const int* tj, const double* tx, int n,
                           int* Ap, int* Ai, double* Ax) {
    trip_t* e = (trip_t*)malloc((size_t)(nz > 0 ? nz : 1) * sizeof(trip_t));
    if (!e) return -1;
    for (int64_t k = 0; k < nz; ++k) {
        int i = ti[k], j = tj[k];
        if (i < 0 || i >= n || j < 0 || j >= n) { free(e); return -1; }
        if (j < i) { int t = i; i = j; j = t; }
        e[k].r = i; e[k].c = j; e[k].v = tx[k]; e[k].seq = k;
    }
    qsort(e, (size_t)nz, sizeof(trip_t), trip_cmp);
    for (int j = 0; j <= n; ++j) Ap[j] = 0;
    int64_t m = 0;
    for (int64_t k = 0; k < nz; ++k) {
        if (k > 0 && e[k].r == e[k - 1].r && e[k].c == e[k - 1].c) {
            Ax[m - 1] += e[k].v;
        } else {
            Ai[m] = e[k].r; Ax[m] = e[k].v; Ap[e[k].c + 1]++; ++m;
        }
    }
    for (int j = 0; j < n; ++j) Ap[j + 1] += Ap[j];
    free(e);
    return m;
}

/* ------------------------------------------------------------------------------------------
 * etree  (include/chol.hpp:378-410): Liu's algorithm with path compression on `ancestor`.
 * ------------------------------------------------------------------------------------------ */
int orc_etree(int n, const int* Ap, const int* Ai, int* parent) {
    int* anc = (int*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int));
    if (!anc) return -1;
    for (int k = 0; k < n; ++k) {
        parent[k] = -1; anc[k] = -1;
        for (int p = Ap[k]; p < Ap[k + 1]; ++p) {
            int i = Ai[p];
            if (i > k) continue;
            while (i != -1 && i < k) {
                int inext = anc[i];
                anc[i] = k;
                if (inext == -1) { parent[i] = k; break; }
                i = inext;
            }
        }
    }
    free(anc);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * post_order + tdfs  (include/chol.hpp:466-499, :445-460): children visited in increasing
 * index (lists built by pushing j = n-1..0 to the front).
 * ------------------------------------------------------------------------------------------ */
int orc_post_order(int n, const int* parent, int* post) {
    int* head = (int*)malloc((size_t)(3 * (size_t)(n > 0 ? n : 1)) * sizeof(int));
    if (!head) return -1;
    int* next = head + n; int* stack = next + n;
    for (int j = 0; j < n; ++j) { head[j] = -1; next[j] = -1; }
    for (int j = n - 1; j >= 0; --j) {
        int p = parent[j];
        if (p == -1) continue;
        next[j] = head[p]; head[p] = j;
    }
    int k = 0;
    for (int j = 0; j < n; ++j) {
        if (parent[j] != -1) continue;
        int top = 0; stack[0] = j;
        while (top >= 0) {
            int p = stack[top], child = head[p];
            if (child == -1) { --top; post[k++] = p; }
            else { head[p] = next[child]; stack[++top] = child; }
        }
    }
    free(head);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * col_count  (include/chol.hpp:567-622) with transpose_pattern (:507-534) and
 * process_edge (:537-559): Gilbert-Ng-Peyton skeleton counts (cs_counts, ata = 0).
 * ------------------------------------------------------------------------------------------ */
int orc_col_count(int n, const int* Ap, const int* Ai, const int* parent, const int* post,
                  int* colcount) {
    int nzA = Ap[n];
    int* ATp = (int*)calloc((size_t)n + 2, sizeof(int));
    int* ATi = (int*)malloc((size_t)(nzA > 0 ? nzA : 1) * sizeof(int));
    int* wk = (int*)malloc((size_t)(6 * (size_t)(n > 0 ? n : 1)) * sizeof(int));
    if (!ATp || !ATi || !wk) { free(ATp); free(ATi); free(wk); return -1; }
    int *first = wk, *maxfirst = wk + n, *prevleaf = wk + 2 * n, *anc = wk + 3 * n,
        *delta = wk + 4 * n, *nxt = wk + 5 * n;
    for (int j = 0; j < n; ++j)
        for (int p = Ap[j]; p < Ap[j + 1]; ++p) ATp[Ai[p] + 1]++;
    for (int j = 0; j < n; ++j) ATp[j + 1] += ATp[j];
    for (int j = 0; j < n; ++j) nxt[j] = ATp[j];
    for (int j = 0; j < n; ++j)
        for (int p = Ap[j]; p < Ap[j + 1]; ++p) ATi[nxt[Ai[p]]++] = j;
    for (int i = 0; i < n; ++i) { first[i] = -1; maxfirst[i] = -1; prevleaf[i] = -1; anc[i] = i; delta[i] = 0; }
    for (int k = 0; k < n; ++k) {                       /* :588-594 */
        int j = post[k];
        delta[j] = (first[j] == -1) ? 1 : 0;
        for (; j != -1 && first[j] == -1; j = parent[j]) first[j] = k;
    }
    for (int k = 0; k < n; ++k) {                       /* :597-611 */
        int j = post[k];
        if (parent[j] != -1) delta[parent[j]]--;
        for (int p = ATp[j]; p < ATp[j + 1]; ++p) {
            int i = ATi[p];
            if (i <= j || first[j] <= maxfirst[i]) continue;   /* process_edge :539 */
            maxfirst[i] = first[j];
            int jprev = prevleaf[i];
            delta[j]++;
            if (jprev != -1) {
                int q = jprev;
                while (q != anc[q]) q = anc[q];
                for (int s = jprev; s != q;) { int sp = anc[s]; anc[s] = q; s = sp; }
                delta[q]--;
            }
            prevleaf[i] = j;
        }
        if (parent[j] != -1) anc[j] = parent[j];
    }
    for (int j = 0; j < n; ++j) colcount[j] = delta[j];  /* :614-619 */
    for (int j = 0; j < n; ++j) { int pj = parent[j]; if (pj != -1) colcount[pj] += colcount[j]; }
    free(ATp); free(ATi); free(wk);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * ereach  (include/chol.hpp:689-715): pattern of row k of L in topological order in
 * s[top..n).  `w` marks (w[i] == k <=> visited for row k); caller presets w[k] = k when it
 * wants the walk to stop at k (chol :807, schol :926; tests/test_chol.cpp:27-57 does not).
 * If x != NULL the numeric overload (:725-728) also scatters A(i,k) into x[i].
 * `path` is a caller-provided stack of length n (the reference uses a std::vector per entry).
 * ------------------------------------------------------------------------------------------ */
int orc_ereach(int n, const int* Ap, const int* Ai, const double* Ax, int k, const int* parent,
               int* s, int* w, double* x, int top, int* path) {
    for (int p = Ap[k]; p < Ap[k + 1]; ++p) {
        int i = Ai[p];
        if (i > k) continue;
        if (x) x[i] = Ax[p];
        int len = 0;
        while (i != -1 && w[i] != k) { path[len++] = i; w[i] = k; i = parent[i]; }
        while (len > 0) s[--top] = path[--len];
    }
    (void)n;
    return top;
}

/* ------------------------------------------------------------------------------------------
 * schol  (include/chol.hpp:874-946): parent, column pointers from col_count, and the row
 * pattern of L built row by row (ereach, append j to every reached column, then the
 * diagonal).  Lp is int64 here; returns nnz(L), or -1 on allocation failure.
 * Call with Li == NULL to get only parent/Lp (sizes) first.
 * ------------------------------------------------------------------------------------------ */
int64_t orc_schol(int n, const int* Ap, const int* Ai, int* parent, int64_t* Lp, int* Li) {
    int* post = (int*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int));
    int* cc = (int*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int));
    if (!post || !cc) { free(post); free(cc); return -1; }
    orc_etree(n, Ap, Ai, parent);
    orc_post_order(n, parent, post);
    orc_col_count(n, Ap, Ai, parent, post, cc);
    int64_t nnz = 0;
    for (int j = 0; j < n; ++j) { Lp[j] = nnz; nnz += cc[j]; }
    Lp[n] = nnz;
    free(post); free(cc);
    if (!Li) return nnz;
    int64_t* c = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
    int* wk = (int*)malloc((size_t)(3 * (size_t)(n > 0 ? n : 1)) * sizeof(int));
    if (!c || !wk) { free(c); free(wk); return -1; }
    int *s = wk, *w = wk + n, *path = wk + 2 * n;
    for (int j = 0; j < n; ++j) { c[j] = Lp[j]; w[j] = -1; }
    for (int j = 0; j < n; ++j) {
        w[j] = j;
        int top = orc_ereach(n, Ap, Ai, NULL, j, parent, s, w, NULL, n, path);
        for (int t = top; t < n; ++t) Li[c[s[t]]++] = j;
        Li[c[j]++] = j;
    }
    free(c); free(wk);
    return nnz;
}

/* ------------------------------------------------------------------------------------------
 * chol  (include/chol.hpp:750-863): up-looking numeric factorization.
 * Returns 0 ok; k+1 (>0) if the pivot of column k is <= 0 (the reference returns
 * unexpected("A is not positive definite."), :849-850); -1 allocation failure.
 * Lp (int64, from orc_schol) is input; Li and Lx are filled.
 * ------------------------------------------------------------------------------------------ */
int orc_chol(int n, const int* Ap, const int* Ai, const double* Ax, const int* parent,
             const int64_t* Lp, int* Li, double* Lx) {
    int64_t* c = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
    int* wk = (int*)malloc((size_t)(3 * (size_t)(n > 0 ? n : 1)) * sizeof(int));
    double* x = (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double));
    if (!c || !wk || !x) { free(c); free(wk); free(x); return -1; }
    int *s = wk, *w = wk + n, *path = wk + 2 * n;
    int rc = 0;
    for (int j = 0; j < n; ++j) { c[j] = Lp[j]; w[j] = -1; }
    for (int k = 0; k < n; ++k) {
        x[k] = 0.0; w[k] = k;                                        /* :806-807 */
        int top = orc_ereach(n, Ap, Ai, Ax, k, parent, s, w, x, n, path);   /* :810 */
        double d = x[k]; x[k] = 0.0;                                 /* :816-817 */
        for (int t = top; t < n; ++t) {
            int i = s[t];
            double Lii = Lx[Lp[i]];                                  /* :824 */
            double lki = x[i] / Lii;                                 /* :827 */
            x[i] = 0.0;
            for (int64_t p = Lp[i] + 1; p < c[i]; ++p)               /* :835-839 */
                x[Li[p]] -= Lx[p] * lki;
            d -= lki * lki;                                          /* :841 */
            int64_t q = c[i]++;                                      /* :844-846 */
            Li[q] = k; Lx[q] = lki;
        }
        if (d <= 0.0) { rc = k + 1; break; }                         /* :849-850 */
        int64_t q = c[k]++;                                          /* :853-855 */
        Li[q] = k; Lx[q] = sqrt(d);
    }
    free(c); free(wk); free(x);
    return rc;
}

/* ------------------------------------------------------------------------------------------
 * compute_levels  (src/chol.cpp:7-40): depth from the root, buckets reversed so the deepest
 * level comes first, ascending node id inside a bucket.  Output: level_ptr[nlev+1] (caller
 * allocates n+1) and level_nodes[n]; returns nlev.
 * ------------------------------------------------------------------------------------------ */
int orc_compute_levels(int n, const int* parent, int* level_ptr, int* level_nodes) {
    if (n <= 0) return 0;
    int* depth = (int*)malloc((size_t)n * 2 * sizeof(int));
    if (!depth) return -1;
    int* path = depth + n;
    for (int j = 0; j < n; ++j) depth[j] = -1;
    for (int j = 0; j < n; ++j) {
        if (depth[j] != -1) continue;
        int v = j, len = 0;
        while (v != -1 && depth[v] == -1) { path[len++] = v; v = parent[v]; }
        int base = (v == -1) ? 0 : depth[v] + 1;
        for (int k = len - 1; k >= 0; --k) depth[path[k]] = base++;
    }
    int maxd = 0;
    for (int j = 0; j < n; ++j) if (depth[j] > maxd) maxd = depth[j];
    int nlev = maxd + 1;
    for (int l = 0; l <= nlev; ++l) level_ptr[l] = 0;
    for (int j = 0; j < n; ++j) level_ptr[(maxd - depth[j]) + 1]++;
    for (int l = 0; l < nlev; ++l) level_ptr[l + 1] += level_ptr[l];
    int* cur = path;
    for (int l = 0; l < nlev; ++l) cur[l] = level_ptr[l];
    for (int j = 0; j < n; ++j) level_nodes[cur[maxd - depth[j]]++] = j;
    free(depth);
    return nlev;
}

/* ------------------------------------------------------------------------------------------
 * compute_supernodes  (src/chol.cpp:42-100): fundamental supernodes; column j joins j-1's
 * supernode iff parent[j-1] == j and colcount[j] == colcount[j-1] - 1 (:75-85).
 * Outputs sn_id[n], sn_start[ns+1] (caller allocates n+1); returns ns.
 * ------------------------------------------------------------------------------------------ */
int orc_compute_supernodes(int n, const int* parent, const int64_t* Lp, int* sn_id, int* sn_start) {
    if (n <= 0) return 0;
    int ns = 0;
    sn_start[0] = 0; sn_id[0] = 0;
    for (int j = 1; j < n; ++j) {
        int same = 0;
        if (parent[j - 1] == j) {
            int64_t lenj = Lp[j + 1] - Lp[j], lenjm1 = Lp[j] - Lp[j - 1];
            if (lenj == lenjm1 - 1) same = 1;
        }
        if (!same) { ++ns; sn_start[ns] = j; }
        sn_id[j] = ns;
    }
    ++ns;
    sn_start[ns] = n;
    return ns;
}

/* ------------------------------------------------------------------------------------------
 * atree  (src/chol.cpp:102-135): assembly tree; parent(s) = smallest supernode id among the
 * rows >= end(s) in any column of s.
 * ------------------------------------------------------------------------------------------ */
int orc_atree(int n, const int64_t* Lp, const int* Li, const int* sn_id, int ns, const int* sn_start,
              int* super_parent) {
    (void)n;
    for (int s = 0; s < ns; ++s) {
        super_parent[s] = -1;
        int start = sn_start[s], end = sn_start[s + 1];
        for (int j = start; j < end; ++j)
            for (int64_t p = Lp[j]; p < Lp[j + 1]; ++p) {
                int row = Li[p];
                if (row >= end) {
                    int t = sn_id[row];
                    if (t != s && (super_parent[s] == -1 || t < super_parent[s])) super_parent[s] = t;
                }
            }
    }
    return 0;
}
