// Newick reference tree of a jplace document -> flat arrays.
//
// Mirrors what the reference needs from its reader for the KR path, not its class design:
//   * the first node of the string is the root, trees are used as rooted
//     (genesis tree/formats/newick/reader.cpp:730-734);
//   * node/edge numbering equals the reference's (reader.cpp:642-770 walks its broker root-first and
//     hangs every new node on the most recently opened child slot, which yields a preorder that
//     visits children right-to-left): node index = preorder rank, edge index = child index - 1.
//     Keeping the same numbering makes mass arrays, averaged branch lengths and the diameter's
//     tie-break directly comparable with the reference;
//   * branch length = first ':' value through strtod (common_tree/newick_reader.hpp:248 uses
//     std::stod), default 1.0 when absent (:297);
//   * edge_num = first {tag} (jplace v2/v3) or first [comment] (jplace v1)
//     (placement/formats/newick_reader.hpp:81-130); edges without one get max+1, in edge order
//     (:183-205).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

namespace krdhost {

struct FlatTree {
    int32_t n_edges = 0;
    int32_t root = 0;
    std::vector<int32_t> parent;      // [E] node on the root side of edge e
    std::vector<int32_t> child;       // [E] node on the leaf side (= e + 1 in this numbering)
    std::vector<double>  bl;          // [E]
    std::vector<int64_t> edge_num;    // [E] jplace edge_num of edge e
    std::vector<std::string> node_name;  // [E + 1]
};

namespace detail {

struct NwNode {
    std::vector<int> kids;            // textual order
    std::string name;
    std::vector<std::string> values, tags, comments;
};

class NewickParser {
public:
    explicit NewickParser(const std::string& s) : s_(s) {}

    int parse(std::vector<NwNode>& nodes)
    {
        nodes_ = &nodes;
        skip_ws();
        int root = subtree();
        skip_ws();
        if (i_ >= s_.size() || s_[i_] != ';') fail("Tree does not finish with a semicolon");
        return root;
    }

private:
    const std::string& s_;
    size_t i_ = 0;
    std::vector<NwNode>* nodes_ = nullptr;

    [[noreturn]] void fail(const std::string& what) const
    {
        throw std::runtime_error("Invalid Newick tree: " + what + " at position " + std::to_string(i_) + ".");
    }
    void skip_ws() { while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r')) ++i_; }

    std::string until(char close)
    {
        size_t j = s_.find(close, i_);
        if (j == std::string::npos) fail(std::string("missing '") + close + "'");
        std::string r = s_.substr(i_, j - i_);
        i_ = j + 1;
        return r;
    }

    int subtree()
    {
        // Nodes are created iteratively to survive caterpillar trees (depth ~ number of edges).
        struct Frame { int node; };
        std::vector<Frame> st;
        int result = -1;
        for (;;) {
            skip_ws();
            int me = (int)nodes_->size();
            nodes_->emplace_back();
            if (!st.empty()) (*nodes_)[st.back().node].kids.push_back(me);
            if (i_ < s_.size() && s_[i_] == '(') { ++i_; st.push_back({me}); continue; }
            // leaf (or empty) element: read its decorations, then close parents as long as we see ')'
            decorations(me);
            for (;;) {
                skip_ws();
                if (i_ >= s_.size()) fail("unexpected end");
                if (s_[i_] == ',') { if (st.empty()) fail("comma outside parentheses"); ++i_; break; }
                if (s_[i_] == ')') {
                    if (st.empty()) fail("unbalanced ')'");
                    ++i_;
                    int inner = st.back().node;
                    st.pop_back();
                    decorations(inner);
                    if (st.empty()) { result = inner; }
                    continue;
                }
                if (st.empty()) return result < 0 ? me : result;
                fail("unexpected character");
            }
        }
    }

    void decorations(int node)
    {
        for (;;) {
            skip_ws();
            if (i_ >= s_.size()) return;
            char c = s_[i_];
            if (c == ':') {
                ++i_; skip_ws();
                size_t b = i_;
                while (i_ < s_.size() && std::string(":,;()[]{} \t\r\n").find(s_[i_]) == std::string::npos) ++i_;
                (*nodes_)[node].values.push_back(s_.substr(b, i_ - b));
            } else if (c == '[') { ++i_; (*nodes_)[node].comments.push_back(until(']')); }
            else if (c == '{') { ++i_; (*nodes_)[node].tags.push_back(until('}')); }
            else if (c == '\'' || c == '"') {
                ++i_;
                std::string nm;
                for (;;) {
                    if (i_ >= s_.size()) fail("unterminated quoted label");
                    if (s_[i_] == c) { if (i_ + 1 < s_.size() && s_[i_ + 1] == c) { nm += c; i_ += 2; continue; } ++i_; break; }
                    nm += s_[i_++];
                }
                (*nodes_)[node].name = nm;
            } else if (c == ',' || c == ')' || c == ';' || c == '(') return;
            else {
                size_t b = i_;
                while (i_ < s_.size() && std::string(":,;()[]{}").find(s_[i_]) == std::string::npos) ++i_;
                std::string nm = s_.substr(b, i_ - b);
                while (!nm.empty() && (nm.back() == ' ' || nm.back() == '\t' || nm.back() == '\n' || nm.back() == '\r')) nm.pop_back();
                (*nodes_)[node].name = nm;
            }
        }
    }
};

}  // namespace detail

// edge_num_from_comments: jplace version 1 wrote edge numbers as [n] instead of {n}
// (genesis placement/formats/jplace_reader.cpp:237-243).
inline FlatTree parse_newick(const std::string& text, bool edge_num_from_comments = false)
{
    std::vector<detail::NwNode> nodes;
    detail::NewickParser P(text);
    const int root = P.parse(nodes);

    FlatTree t;
    const int n_nodes = (int)nodes.size();
    t.n_edges = n_nodes - 1;
    t.root = 0;
    t.parent.assign(t.n_edges, 0); t.child.assign(t.n_edges, 0);
    t.bl.assign(t.n_edges, 1.0); t.edge_num.assign(t.n_edges, -1);
    t.node_name.assign(n_nodes, std::string());

    // preorder, children right-to-left
    std::vector<std::pair<int, int>> st;  // (parsed node, new parent index)
    st.push_back({root, -1});
    int next = 0;
    int64_t max_num = -1;
    int missing = 0;
    while (!st.empty()) {
        auto [pn, par] = st.back();
        st.pop_back();
        const int idx = next++;
        const detail::NwNode& nd = nodes[pn];
        t.node_name[idx] = nd.name;
        if (par >= 0) {
            const int e = idx - 1;
            t.parent[e] = par; t.child[e] = idx;
            if (!nd.values.empty()) {
                try { t.bl[e] = std::stod(nd.values[0]); }
                catch (...) { throw std::runtime_error("Invalid Newick tree: cannot read branch length '" + nd.values[0] + "'."); }
            }
            const auto& en = edge_num_from_comments ? nd.comments : nd.tags;
            if (en.size() > 1)
                throw std::invalid_argument(std::string("Edge at node '") + nd.name + "' contains more than one " +
                                            (edge_num_from_comments ? "comment value '[xyz]'" : "tag value '{xyz}'") +
                                            ". Expecting only one for the placement edge_num of this edge.");
            if (en.size() == 1) {
                try { t.edge_num[e] = (int64_t)std::stoull(en[0]); }
                catch (...) { throw std::invalid_argument("Edge at node '" + nd.name + "' has an edge_num that is not a number."); }
                if (t.edge_num[e] > max_num) max_num = t.edge_num[e];
            } else ++missing;
        }
        // push in textual order so that the LAST child is popped (visited) first
        for (int k : nd.kids) st.push_back({k, idx});
    }
    if (missing > 0) {
        if (t.n_edges >= 3 && (t.n_edges - missing) < t.n_edges - 3)
            throw std::runtime_error("Jplace document contains too many edges without an edge_num.");
        for (int e = 0; e < t.n_edges; ++e)
            if (t.edge_num[e] < 0) t.edge_num[e] = ++max_num;
    }
    return t;
}

}  // namespace krdhost
