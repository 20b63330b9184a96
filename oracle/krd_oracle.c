/* TEST INFRASTRUCTURE ONLY.  CPU restatement ("port") of the reference's KR-distance path on
 * flat arrays.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm
 * may load this library; the product (gappa_b200/) never does.
 *
 * Parity status: PINNED — checked in tests/test_oracle.py against
 *   - genesis' own golden values (libs/genesis/test/src/placement/measures.cpp:82-109,
 *     libs/genesis/test/src/tree/mass_tree.cpp:53-85) and
 *   - outputs of the compiled reference (oracle/_ref/libkrd_ref.so) committed under tests/golden/.
 *
 * Flat data model (shared with include/krd_b200.h):
 *   tree   : n_edges edges; edge e joins parent_node[e] (root side, "primary") and
 *            child_node[e] (leaf side, "secondary"); branch_length[e]; nodes are 0..n_edges.
 *   masses : CSR over samples: off[n_samples+1]; per entry edge[m], pos[m] (distance from the
 *            primary node, "proximal_length"), mass[m].  A "canonical" sample is sorted by
 *            (edge asc, pos asc) with unique (edge,pos) — the iteration order of genesis'
 *            per-edge std::map<double,double> (tree/mass_tree/tree.hpp:199).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- tree helpers --------------------------------------------------------------------------- */

/* Edge postorder (children before parents), children visited in listing order.  Any valid
 * postorder gives the same distance up to rounding; genesis uses its link order
 * (tree/iterator/postorder.hpp:194-223).  Returns 0 or -1 if the arrays are not a rooted tree. */
int okrd_postorder(int n_edges, const int* parent_node, const int* child_node, int root, int* order)
{
    int n_nodes = n_edges + 1, e, n;
    int* head = (int*)malloc(sizeof(int) * (size_t)(n_nodes + 1));
    int* kid  = (int*)malloc(sizeof(int) * (size_t)(n_edges > 0 ? n_edges : 1));
    int* next = (int*)malloc(sizeof(int) * (size_t)n_nodes);
    int* stk  = (int*)malloc(sizeof(int) * (size_t)n_nodes);
    int sp = 0, out = 0, ok = 0;
    memset(head, 0, sizeof(int) * (size_t)(n_nodes + 1));
    for (e = 0; e < n_edges; ++e) {
        if (parent_node[e] < 0 || parent_node[e] >= n_nodes || child_node[e] < 0 || child_node[e] >= n_nodes) goto done;
        head[parent_node[e] + 1]++;
    }
    for (n = 0; n < n_nodes; ++n) head[n + 1] += head[n];
    memcpy(next, head, sizeof(int) * (size_t)n_nodes);
    for (e = 0; e < n_edges; ++e) kid[next[parent_node[e]]++] = e;
    memcpy(next, head, sizeof(int) * (size_t)n_nodes);
    /* stack holds edges; root handled via its children */
    {
        int node = root, up = -1;
        int* upedge = (int*)malloc(sizeof(int) * (size_t)n_nodes);
        stk[sp] = node; upedge[sp] = up; sp++;
        while (sp > 0) {
            node = stk[sp - 1];
            if (next[node] < head[node + 1]) {
                e = kid[next[node]++];
                if (sp >= n_nodes) { free(upedge); goto done; }
                stk[sp] = child_node[e]; upedge[sp] = e; sp++;
            } else {
                if (upedge[sp - 1] >= 0) order[out++] = upedge[sp - 1];
                sp--;
            }
        }
        free(upedge);
    }
    ok = (out == n_edges);
done:
    free(head); free(kid); free(next); free(stk);
    return ok ? 0 : -1;
}

/* ---- pair distance: genesis tree/mass_tree/emd.cpp:60-181 ----------------------------------- */

/* per-sample index: start[e]..start[e+1] into that sample's entries (canonical order) */
static void edge_starts(int n_edges, int64_t lo, int64_t hi, const int* edge, int64_t* start)
{
    int e = 0; int64_t m;
    for (m = lo; m < hi; ++m) { while (e <= edge[m]) start[e++] = m; }
    while (e <= n_edges) start[e++] = hi;
}

static double pair_emd(int n_edges, const int* parent_node, const int* child_node,
                       const double* bl, const int* order,
                       const int64_t* sa, const int64_t* sb, /* edge starts of both samples */
                       const double* pos, const double* mass, double p, double* node_mass)
{
    double work = 0.0; int k;
    memset(node_mass, 0, sizeof(double) * (size_t)(n_edges + 1));      /* emd.cpp:89 */
    for (k = 0; k < n_edges; ++k) {                                      /* emd.cpp:96-100 */
        int e = order[k];
        int64_t ia = sa[e + 1] - 1, ib = sb[e + 1] - 1;                  /* walk distal -> proximal */
        double cur_pos = bl[e];                                          /* emd.cpp:141-144 (equal bl) */
        double cur_mass = node_mass[child_node[e]];                      /* emd.cpp:145 */
        while (ia >= sa[e] || ib >= sb[e]) {                             /* emd.cpp:149-161 */
            double x, dm;
            if (ib < sb[e] || (ia >= sa[e] && pos[ia] > pos[ib])) { x = pos[ia]; dm = mass[ia]; ia--; }
            else if (ia < sa[e] || pos[ib] > pos[ia])             { x = pos[ib]; dm = 0.0 - mass[ib]; ib--; }
            else { x = pos[ia]; dm = mass[ia] - mass[ib]; ia--; ib--; }  /* emd.cpp:134-137: map merge */
            work += pow(fabs(cur_mass), p) * (cur_pos - x);
            cur_pos = x;
            cur_mass += dm;
        }
        work += pow(fabs(cur_mass), p) * cur_pos;                        /* emd.cpp:166 */
        node_mass[parent_node[e]] += cur_mass;                           /* emd.cpp:167 */
    }
    if (p > 1.0) work = pow(work, 1.0 / p);                              /* emd.cpp:176-178 */
    return work;
}

/* All pairs: emd.cpp:183-235.  out is n×n row-major, zero diagonal, symmetric.
 * n_pairs_limit > 0 computes only the first n_pairs_limit pairs of the linearised upper triangle
 * in (i asc, j asc) order (bounded CPU-baseline sample); others stay 0.
 * Returns 0, 1 (p <= 0: "Invalid exponent value p ... Has to be > 0.0."), 2 (bad tree). */
int okrd_matrix(int n_edges, const int* parent_node, const int* child_node, const double* bl, int root,
                int n_samples, const int64_t* off, const int* edge, const double* pos,
                const double* mass, double p, int threads, int64_t n_pairs_limit, double* out)
{
    int* order; int64_t* starts; int i; int64_t n_pairs, k;
    if (!(p > 0.0)) return 1;                                            /* emd.cpp:186-190 */
    order = (int*)malloc(sizeof(int) * (size_t)(n_edges > 0 ? n_edges : 1));
    if (okrd_postorder(n_edges, parent_node, child_node, root, order) != 0) { free(order); return 2; }
    starts = (int64_t*)malloc(sizeof(int64_t) * (size_t)n_samples * (size_t)(n_edges + 1));
    for (i = 0; i < n_samples; ++i)
        edge_starts(n_edges, off[i], off[i + 1], edge, starts + (size_t)i * (size_t)(n_edges + 1));
    memset(out, 0, sizeof(double) * (size_t)n_samples * (size_t)n_samples);   /* emd.cpp:193 */
    n_pairs = (int64_t)n_samples * (n_samples - 1) / 2;
    if (n_pairs_limit > 0 && n_pairs_limit < n_pairs) n_pairs = n_pairs_limit;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    #pragma omp parallel
    {
        double* nm = (double*)malloc(sizeof(double) * (size_t)(n_edges + 1));
        #pragma omp for schedule(dynamic)
        for (k = 0; k < n_pairs; ++k) {
            /* invert k -> (a,b), a<b, row-major upper triangle */
            int64_t a = 0, rem = k, row = n_samples - 1; int b;
            double d;
            while (rem >= row) { rem -= row; row--; a++; }
            b = (int)(a + 1 + rem);
            d = pair_emd(n_edges, parent_node, child_node, bl, order,
                         starts + (size_t)a * (size_t)(n_edges + 1), starts + (size_t)b * (size_t)(n_edges + 1),
                         pos, mass, p, nm);
            out[(size_t)a * n_samples + b] = d;
            out[(size_t)b * n_samples + a] = d;
        }
        free(nm);
    }
    free(order); free(starts);
    return 0;
}

/* ---- binify: tree/mass_tree/functions.cpp:222-265 ------------------------------------------- */

/* bin index q in [0,bins) of a position; functions.cpp:231-241 (the floor(...) part).
 * fmax(0.0, NaN) must give 0.0 like std::max(0.0, NaN) does in libstdc++ (comparison false). */
int okrd_bin_index(double pos, double bl, int bins)
{
    double nb = (double)bins;
    double v = pos / bl * nb;
    double pn = (0.0 < v) ? v : 0.0;             /* std::max(0.0, v): returns 0.0 unless 0.0 < v */
    double hi = nextafter(nb, 0.0);
    pn = (hi < pn) ? hi : pn;                    /* std::min(pn, hi): returns pn unless hi < pn */
    return (int)floor(pn);
}

double okrd_bin_center(int q, double bl, int bins)
{
    double nb = (double)bins;
    return ((double)q * bl / nb) + (bl / nb / 2.0);  /* functions.cpp:240 */
}

/* Binify canonical samples in place-compatible fashion: writes canonical output (edge asc, bin
 * centre asc).  Output arrays sized like the inputs.  Returns total entry count. */
int64_t okrd_binify(int n_samples, const int64_t* off, const int* edge, const double* pos,
                    const double* mass, const double* bl, int bins,
                    int64_t* out_off, int* out_edge, double* out_pos, double* out_mass)
{
    int64_t w = 0, m; int i;
    out_off[0] = 0;
    for (i = 0; i < n_samples; ++i) {
        int64_t first = w;
        for (m = off[i]; m < off[i + 1]; ++m) {
            int e = edge[m];
            int q = okrd_bin_index(pos[m], bl[e], bins);
            double c = okrd_bin_center(q, bl[e], bins);
            /* std::map keyed by the centre VALUE: equal centres merge (also across q if bl==0) */
            if (w > first && out_edge[w - 1] == e && out_pos[w - 1] == c) out_mass[w - 1] += mass[m];
            else { out_edge[w] = e; out_pos[w] = c; out_mass[w] = mass[m]; ++w; }
        }
        out_off[i + 1] = w;
    }
    return w;
}

/* ---- canonicalise: sort by (edge,pos), merge duplicates in input order (map[pos] += mass) ---- */

typedef struct { int edge; double pos; double mass; int64_t idx; } okrd_ent;
static int ent_cmp(const void* a, const void* b)
{
    const okrd_ent* x = (const okrd_ent*)a; const okrd_ent* y = (const okrd_ent*)b;
    if (x->edge != y->edge) return x->edge < y->edge ? -1 : 1;
    if (x->pos != y->pos) return x->pos < y->pos ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx ? 1 : 0);
}

int64_t okrd_canonicalize(int n_samples, const int64_t* off, const int* edge, const double* pos,
                          const double* mass, int64_t* out_off, int* out_edge, double* out_pos,
                          double* out_mass)
{
    int64_t w = 0, m; int i;
    out_off[0] = 0;
    for (i = 0; i < n_samples; ++i) {
        int64_t n = off[i + 1] - off[i], first = w;
        okrd_ent* t = (okrd_ent*)malloc(sizeof(okrd_ent) * (size_t)(n > 0 ? n : 1));
        for (m = 0; m < n; ++m) { t[m].edge = edge[off[i] + m]; t[m].pos = pos[off[i] + m]; t[m].mass = mass[off[i] + m]; t[m].idx = m; }
        qsort(t, (size_t)n, sizeof(okrd_ent), ent_cmp);
        for (m = 0; m < n; ++m) {
            if (w > first && out_edge[w - 1] == t[m].edge && out_pos[w - 1] == t[m].pos) out_mass[w - 1] += t[m].mass;
            else { out_edge[w] = t[m].edge; out_pos[w] = t[m].pos; out_mass[w] = t[m].mass; ++w; }
        }
        free(t);
        out_off[i + 1] = w;
    }
    return w;
}

int okrd_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
