/* TEST INFRASTRUCTURE ONLY.  Extended-precision ground truth of the pair distance, for the question the FP64 oracle
 * cannot answer: how far is genesis' own double-precision result (and ours) from the mathematically exact KR distance of
 * the given mass trees?  Same walk as pair_emd in krd_oracle.c (genesis tree/mass_tree/emd.cpp:60-181), with every
 * quantity carried in IEEE binary128 (__float128, libquadmath): the masses and positions are the doubles the other
 * implementations receive, their partial sums are exact (or correct to 2^-113) in this format, so ties of the cumulative
 * mass are exact zeros, and pow / the final sum are correct to ~1e-33.  Only tests/ load this library.
 * Parity status: pinned indirectly — it reproduces the port (which is pinned to genesis) to the port's own rounding on
 * every fixture (tests/test_oracle.py::test_truth_agrees_with_port). */
#include <quadmath.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

int okrd_postorder(int n_edges, const int* parent_node, const int* child_node, int root, int* order);   /* krd_oracle.c */

typedef __float128 q128;

static void edge_starts_q(int n_edges, int64_t lo, int64_t hi, const int* edge, int64_t* start)
{
    int e = 0; int64_t m;
    for (m = lo; m < hi; ++m) { while (e <= edge[m]) start[e++] = m; }
    while (e <= n_edges) start[e++] = hi;
}

static q128 pair_truth(int n_edges, const int* parent_node, const int* child_node, const double* bl, const int* order,
                       const int64_t* sa, const int64_t* sb, const double* pos, const double* mass, double p, q128* node_mass)
{
    q128 work = 0; int k;
    const q128 pq = (q128)p;
    memset(node_mass, 0, sizeof(q128) * (size_t)(n_edges + 1));
    for (k = 0; k < n_edges; ++k) {
        int e = order[k];
        int64_t ia = sa[e + 1] - 1, ib = sb[e + 1] - 1;
        q128 cur_pos = (q128)bl[e];
        q128 cur_mass = node_mass[child_node[e]];
        while (ia >= sa[e] || ib >= sb[e]) {
            q128 x, dm;
            if (ib < sb[e] || (ia >= sa[e] && pos[ia] > pos[ib])) { x = (q128)pos[ia]; dm = (q128)mass[ia]; ia--; }
            else if (ia < sa[e] || pos[ib] > pos[ia])             { x = (q128)pos[ib]; dm = -(q128)mass[ib]; ib--; }
            else { x = (q128)pos[ia]; dm = (q128)mass[ia] - (q128)mass[ib]; ia--; ib--; }
            work += powq(fabsq(cur_mass), pq) * (cur_pos - x);
            cur_pos = x;
            cur_mass += dm;
        }
        work += powq(fabsq(cur_mass), pq) * cur_pos;
        node_mass[parent_node[e]] += cur_mass;
    }
    if (p > 1.0) work = powq(work, 1.0Q / pq);
    return work;
}

/* Listed pairs (pair_a[k], pair_b[k]) of canonical samples -> out[k] (rounded to double at the very end) and, if
 * out_hi/out_lo are given, the value as an unevaluated double-double (hi + lo).  Returns 0, 1 (p <= 0), 2 (bad tree). */
int okrd_truth_pairs(int n_edges, const int* parent_node, const int* child_node, const double* bl, int root,
                     int n_samples, const int64_t* off, const int* edge, const double* pos, const double* mass, double p,
                     int64_t n_pairs, const int* pair_a, const int* pair_b, double* out, double* out_lo)
{
    int* order; int64_t k;
    if (!(p > 0.0)) return 1;
    order = (int*)malloc(sizeof(int) * (size_t)(n_edges > 0 ? n_edges : 1));
    if (okrd_postorder(n_edges, parent_node, child_node, root, order) != 0) { free(order); return 2; }
    #pragma omp parallel
    {
        q128* nm = (q128*)malloc(sizeof(q128) * (size_t)(n_edges + 1));
        int64_t* sa = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_edges + 1));
        int64_t* sb = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_edges + 1));
        #pragma omp for schedule(dynamic)
        for (k = 0; k < n_pairs; ++k) {
            const int a = pair_a[k], b = pair_b[k];
            q128 d;
            if (a < 0 || b < 0 || a >= n_samples || b >= n_samples) { out[k] = -1.0; continue; }
            edge_starts_q(n_edges, off[a], off[a + 1], edge, sa);
            edge_starts_q(n_edges, off[b], off[b + 1], edge, sb);
            d = pair_truth(n_edges, parent_node, child_node, bl, order, sa, sb, pos, mass, p, nm);
            out[k] = (double)d;
            if (out_lo) out_lo[k] = (double)(d - (q128)out[k]);
        }
        free(nm); free(sa); free(sb);
    }
    free(order);
    return 0;
}
