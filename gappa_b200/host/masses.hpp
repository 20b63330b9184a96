// Placements -> per-sample mass lists -> the CSR arrays that cross the C ABI.
//
// Host-side restatement, on flat arrays, of what JplaceInputOptions::mass_tree_set() does before the
// distance call (gappa src/options/jplace_input.cpp:129-171,295-332):
//   --point-mass / --ignore-multiplicities filters            jplace_input.cpp:138-151
//   placements -> normalised masses at proximal positions     genesis placement/function/operators.cpp:132-182
//   scaler = sum_pq (sum LWR)(sum multiplicity)               placement/function/masses.cpp:142-158
//   identical-topology check                                  tree/function/operators.cpp:170-215
//   averaged branch lengths + position rescale                tree/mass_tree/functions.cpp:271-335
//   diameter for --normalize                                  tree/common_tree/functions.cpp:178-205
// Summation orders are the reference's (pquery order for equal positions, sample order for the
// branch-length average) so that the arrays agree with its MassTrees bit for bit.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "jplace.hpp"

namespace krdhost {

struct MassEntry { int32_t edge; double pos; double mass; };

struct SampleMasses {
    std::vector<MassEntry> entries;   // sorted by (edge, pos), positions unique per edge
    double pendant_work = 0.0;        // sum LWR * mult * pendant / scaler (operators.cpp:167)
};

struct MassSet {            // what krd_set_tree / krd_set_masses take
    FlatTree tree;          // averaged branch lengths
    std::vector<int64_t> off;
    std::vector<int32_t> edge;
    std::vector<double> pos, mass;
    std::vector<double> pendant;
    int32_t n_samples() const { return (int32_t)off.size() - 1; }
};

// jplace_input.cpp:138-141 -> placement/function/functions.cpp:246-266 (keep the max-LWR placement:
// std::sort by LWR descending, the same library call so ties resolve identically) and :122-134
// (divide by the remaining sum; throws when it is zero).
inline void apply_point_mass(Sample& s)
{
    std::vector<Placement> kept;
    std::vector<int64_t> off{0};
    kept.reserve(s.n_pqueries());
    for (size_t q = 0; q < s.n_pqueries(); ++q) {
        auto b = s.placements.begin() + s.pq_off[q], e = s.placements.begin() + s.pq_off[q + 1];
        if (e - b > 1) std::sort(b, e, [](const Placement& l, const Placement& r) { return l.lwr > r.lwr; });
        if (e - b >= 1) {
            Placement p = *b;
            const double sum = p.lwr;
            if (sum == 0.0) throw std::overflow_error("Cannot normalize weight ratios if all of them are zero.");
            p.lwr /= sum;
            kept.push_back(p);
        }
        off.push_back((int64_t)kept.size());
    }
    s.placements.swap(kept);
    s.pq_off.swap(off);
}

// jplace_input.cpp:144-151: every name's multiplicity divided by the pquery's total.
inline void apply_ignore_multiplicities(Sample& s)
{
    for (size_t q = 0; q < s.n_pqueries(); ++q) {
        double tm = 0.0;
        for (int64_t k = s.mult_off[q]; k < s.mult_off[q + 1]; ++k) tm += s.mult[(size_t)k];
        for (int64_t k = s.mult_off[q]; k < s.mult_off[q + 1]; ++k) s.mult[(size_t)k] /= tm;
    }
}

inline SampleMasses sample_to_masses(const Sample& s, bool normalize = true)
{
    const size_t nq = s.n_pqueries();
    std::vector<double> mult(nq, 0.0);
    double scaler = 1.0;
    {
        double sum = 0.0;
        for (size_t q = 0; q < nq; ++q) {
            double m = 0.0;
            for (int64_t k = s.mult_off[q]; k < s.mult_off[q + 1]; ++k) m += s.mult[(size_t)k];
            mult[q] = m;
            double lw = 0.0;
            for (int64_t k = s.pq_off[q]; k < s.pq_off[q + 1]; ++k) lw += s.placements[(size_t)k].lwr;
            sum += lw * m;
        }
        if (normalize) scaler = sum;
    }
    struct Tmp { int32_t edge; double pos; double mass; int64_t seq; };
    std::vector<Tmp> tmp;
    tmp.reserve(s.placements.size());
    SampleMasses out;
    for (size_t q = 0; q < nq; ++q) {
        for (int64_t k = s.pq_off[q]; k < s.pq_off[q + 1]; ++k) {
            const Placement& p = s.placements[(size_t)k];
            const double bl = s.tree.bl[(size_t)p.edge];
            // same branch on both sides here, but the reference still evaluates prox / bl * bl
            double position = p.proximal / bl * bl;
            if (!std::isfinite(bl) || bl == 0.0) position = 0.0;
            tmp.push_back({p.edge, position, 1.0 * p.lwr * mult[q] / scaler, k});
            out.pendant_work += p.lwr * mult[q] * p.pendant / scaler;
        }
    }
    // std::map<double,double>::operator[] += in placement order  ==  stable sort + run sum
    std::sort(tmp.begin(), tmp.end(), [](const Tmp& a, const Tmp& b) {
        if (a.edge != b.edge) return a.edge < b.edge;
        if (a.pos != b.pos) return a.pos < b.pos;
        return a.seq < b.seq;
    });
    for (size_t i = 0; i < tmp.size();) {
        size_t j = i;
        double m = 0.0;
        while (j < tmp.size() && tmp[j].edge == tmp[i].edge && tmp[j].pos == tmp[i].pos) { m += tmp[j].mass; ++j; }
        out.entries.push_back({tmp[i].edge, tmp[i].pos, m});
        i = j;
    }
    return out;
}

inline bool same_topology(const FlatTree& a, const FlatTree& b)
{
    return a.n_edges == b.n_edges && a.parent == b.parent && a.child == b.child;
}

// tree/mass_tree/functions.cpp:271-335
inline void average_branch_lengths(std::vector<FlatTree>& trees, std::vector<SampleMasses>& masses)
{
    const size_t n = trees.size();
    if (n < 2) return;
    const size_t E = (size_t)trees[0].n_edges;
    std::vector<double> avg(E, 0.0);
    for (const auto& t : trees) {
        if ((size_t)t.n_edges != E) throw std::runtime_error("Cannot make average branch lengths, because trees have different size.");
        for (size_t e = 0; e < E; ++e) avg[e] += t.bl[e];
    }
    for (auto& a : avg) a /= (double)n;
    for (size_t i = 0; i < n; ++i) {
        auto& en = masses[i].entries;
        std::vector<MassEntry> merged;
        merged.reserve(en.size());
        for (size_t b = 0; b < en.size();) {
            const int32_t e = en[b].edge;
            size_t c = b;
            while (c < en.size() && en[c].edge == e) ++c;
            double scaler = avg[(size_t)e] / trees[i].bl[(size_t)e];
            if (!std::isfinite(scaler)) scaler = 1.0;
            // new_masses[pos * scaler] += mass, iterating the old map in ascending position
            const size_t first = merged.size();
            for (size_t k = b; k < c; ++k) {
                const double np = en[k].pos * scaler;
                size_t w = merged.size();
                while (w > first && merged[w - 1].pos > np) --w;           // monotone in practice: no shift
                if (w > first && merged[w - 1].pos == np) merged[w - 1].mass += en[k].mass;
                else merged.insert(merged.begin() + (long)w, MassEntry{e, np, en[k].mass});
            }
            b = c;
        }
        en.swap(merged);
        trees[i].bl = avg;
    }
}

// tree/common_tree/functions.cpp:178-205 with common_tree/distances.cpp:121-153: farthest node from
// the root (first maximum in node-index order), then the farthest distance from that node.
inline double diameter(const FlatTree& t)
{
    const int n_nodes = t.n_edges + 1;
    if (t.n_edges == 0) return 0.0;
    std::vector<std::vector<std::pair<int, int>>> adj((size_t)n_nodes);   // (neighbour, edge)
    for (int e = 0; e < t.n_edges; ++e) {
        adj[(size_t)t.parent[(size_t)e]].push_back({t.child[(size_t)e], e});
        adj[(size_t)t.child[(size_t)e]].push_back({t.parent[(size_t)e], e});
    }
    auto sweep = [&](int start) {
        std::vector<double> d((size_t)n_nodes, -1.0);
        std::vector<int> queue{start};
        d[(size_t)start] = 0.0;
        for (size_t q = 0; q < queue.size(); ++q) {
            const int u = queue[q];
            for (auto [v, e] : adj[(size_t)u])
                if (d[(size_t)v] < 0.0) { d[(size_t)v] = d[(size_t)u] + t.bl[(size_t)e]; queue.push_back(v); }
        }
        return d;
    };
    const auto d0 = sweep(t.root);
    const int far = (int)(std::max_element(d0.begin(), d0.end()) - d0.begin());
    const auto d1 = sweep(far);
    return *std::max_element(d1.begin(), d1.end());
}

// Everything after parsing: checks, averaging, CSR assembly.
inline MassSet assemble(std::vector<FlatTree>& trees, std::vector<SampleMasses>& masses)
{
    for (size_t i = 1; i < trees.size(); ++i)
        if (!same_topology(trees[i - 1], trees[i])) throw std::runtime_error("Sample reference trees do not have identical topology.");
    average_branch_lengths(trees, masses);
    MassSet ms;
    ms.tree = trees.empty() ? FlatTree{} : trees[0];
    ms.off.push_back(0);
    size_t total = 0;
    for (const auto& m : masses) total += m.entries.size();
    ms.edge.reserve(total); ms.pos.reserve(total); ms.mass.reserve(total);
    for (const auto& m : masses) {
        for (const auto& en : m.entries) { ms.edge.push_back(en.edge); ms.pos.push_back(en.pos); ms.mass.push_back(en.mass); }
        ms.off.push_back((int64_t)ms.edge.size());
        ms.pendant.push_back(m.pendant_work);
    }
    return ms;
}

}  // namespace krdhost
