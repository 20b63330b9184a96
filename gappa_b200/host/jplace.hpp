// jplace document -> placements in flat arrays (one pass, no DOM).
//
// Semantics follow the reference reader for everything the KR path consumes
// (genesis placement/formats/jplace_reader.cpp:74-115 file/gzip input, :159-203 version,
// :237-262 tree, :268-340 fields, :346-567 placements incl. the distal -> proximal conversion at
// :524-527, :573-700 names/multiplicities).  Keys may come in any order ("fields" usually FOLLOWS
// "placements"), so rows are stored raw and interpreted once the whole document was seen.
// Pquery names are not kept: the KR path only needs each name's multiplicity.
#pragma once
#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <unordered_map>

#include "json_scan.hpp"
#include "newick.hpp"

namespace krdhost {

struct Placement {
    int32_t edge = 0;            // edge index in FlatTree numbering
    double likelihood = 0.0;
    double lwr = 0.0;            // like_weight_ratio
    double proximal = 0.0;       // distance from the edge's parent node
    double pendant = 0.0;
};

struct Sample {
    FlatTree tree;
    std::vector<Placement> placements;     // all pqueries, concatenated
    std::vector<int64_t> pq_off;           // [n_pqueries + 1] into placements
    std::vector<double> mult;              // name multiplicities, concatenated
    std::vector<int64_t> mult_off;         // [n_pqueries + 1] into mult
    int version = -1;
    size_t n_pqueries() const { return pq_off.empty() ? 0 : pq_off.size() - 1; }
};

// Whole file into memory; gzread handles plain and gzip-compressed input alike
// (the reference sniffs the gzip magic, utils/io/input_source.hpp:67-78).
inline std::string read_maybe_gzip(const std::string& path)
{
    gzFile f = gzopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("Cannot open input file '" + path + "'.");
    gzbuffer(f, 1 << 20);
    std::string out;
    std::vector<char> buf(1 << 20);
    for (;;) {
        int n = gzread(f, buf.data(), (unsigned)buf.size());
        if (n < 0) { gzclose(f); throw std::runtime_error("Error reading input file '" + path + "'."); }
        if (n == 0) break;
        out.append(buf.data(), (size_t)n);
    }
    gzclose(f);
    return out;
}

namespace detail {

struct RawCell { JsonNumber num; bool is_number; JsonScanner::Type type; };

struct RawDoc {
    bool have_tree = false, have_fields = false, have_placements = false;
    std::string tree;
    std::vector<std::string> fields;
    int version = -1;
    std::vector<RawCell> cells;            // all placement rows, concatenated
    std::vector<int64_t> row_off{0};       // [n_rows + 1]
    std::vector<int64_t> pq_row_off{0};    // [n_pq + 1] into rows
    std::vector<double> mult;
    std::vector<int64_t> mult_off{0};
};

inline void parse_pquery(JsonScanner& js, RawDoc& d)
{
    if (js.type() != JsonScanner::Object)
        throw std::runtime_error(std::string("Jplace document contains a value of type '") + JsonScanner::type_name(js.type()) +
                                 "' instead of an object with a pquery at key 'placements'.");
    js.expect('{');
    bool have_p = false, have_n = false, have_nm = false, have_m = false, m_is_number = false;
    double m_val = 1.0;
    int64_t n_names = 0;        // for key 'n'
    bool n_valid_kind = true, n_all_strings = true;
    std::vector<double> nm_vals;
    if (!js.try_char('}')) {
        for (;;) {
            const std::string key = js.string();
            js.expect(':');
            if (key == "p" && !have_p) {
                have_p = true;
                if (js.type() != JsonScanner::Array)
                    throw std::runtime_error("Jplace document contains a pquery at key 'placements' that does not contain an array of placements at key 'p'.");
                js.expect('[');
                if (js.try_char(']'))
                    throw std::runtime_error("Jplace document contains a pquery at key 'placements' that does not contain any placements at key 'p'.");
                for (;;) {
                    if (js.type() != JsonScanner::Array)
                        throw std::runtime_error("Jplace document contains a pquery with invalid placement at key 'p'.");
                    js.expect('[');
                    if (!js.try_char(']')) {
                        for (;;) {
                            RawCell c{};
                            c.type = js.type();
                            c.is_number = (c.type == JsonScanner::Number);
                            if (c.is_number) c.num = js.number(); else js.skip_value();
                            d.cells.push_back(c);
                            if (js.try_char(',')) continue;
                            js.expect(']');
                            break;
                        }
                    }
                    d.row_off.push_back((int64_t)d.cells.size());
                    if (js.try_char(',')) continue;
                    js.expect(']');
                    break;
                }
            } else if (key == "n" && !have_n) {
                have_n = true;
                if (js.type() == JsonScanner::String) { js.skip_string(); n_names = 1; }
                else if (js.type() == JsonScanner::Array) {
                    js.expect('[');
                    if (!js.try_char(']')) {
                        for (;;) {
                            if (js.type() != JsonScanner::String) { n_all_strings = false; js.skip_value(); } else js.skip_string();
                            ++n_names;
                            if (js.try_char(',')) continue;
                            js.expect(']');
                            break;
                        }
                    }
                } else { n_valid_kind = false; js.skip_value(); }
            } else if (key == "m" && !have_m) {
                have_m = true;
                m_is_number = (js.type() == JsonScanner::Number);
                if (m_is_number) m_val = js.number().as_double(); else js.skip_value();
            } else if (key == "nm" && !have_nm) {
                have_nm = true;
                if (js.type() != JsonScanner::Array)
                    throw std::runtime_error("Jplace document contains a pquery with key 'nm' that is not array.");
                js.expect('[');
                if (js.try_char(']'))
                    throw std::runtime_error("Jplace document contains a pquery with key 'nm' that does not contain any values.");
                for (;;) {
                    if (js.type() != JsonScanner::Array)
                        throw std::runtime_error("Jplace document contains a pquery where key 'nm' has a non-array field.");
                    js.expect('[');
                    int cnt = 0;
                    if (!js.try_char(']')) {
                        for (;;) {
                            if (cnt == 0) {
                                if (js.type() != JsonScanner::String)
                                    throw std::runtime_error("Jplace document contains a pquery where key 'nm' has an array whose first value is not a string for the name.");
                                js.skip_string();
                            } else if (cnt == 1) {
                                if (js.type() != JsonScanner::Number)
                                    throw std::runtime_error("Jplace document contains a pquery where key 'nm' has an array whose second value is not a number for the multiplicity.");
                                nm_vals.push_back(js.number().as_double());
                            } else js.skip_value();
                            ++cnt;
                            if (js.try_char(',')) continue;
                            js.expect(']');
                            break;
                        }
                    }
                    if (cnt != 2)
                        throw std::runtime_error("Jplace document contains a pquery where key 'nm' has an array field with size != 2 (one for the name, one for the multiplicity).");
                    if (js.try_char(',')) continue;
                    js.expect(']');
                    break;
                }
            } else js.skip_value();
            if (js.try_char(',')) continue;
            js.expect('}');
            break;
        }
    }
    if (!have_p)
        throw std::runtime_error("Jplace document contains a pquery at key 'placements' that does not contain an array of placements at key 'p'.");
    if (have_n && have_nm) throw std::runtime_error("Jplace document contains a pquery with both an 'n' and an 'nm' key.");
    if (!have_n && !have_nm) throw std::runtime_error("Jplace document contains a pquery with neither an 'n' nor an 'nm' key.");
    if (have_m && !have_n) throw std::runtime_error("Jplace document contains a pquery with key 'm' but without 'n' key.");
    if (have_n) {
        if (have_m) {
            if (!m_is_number)
                throw std::runtime_error("Jplace document contains a pquery where key 'm' has a value is not a valid number for the multiplicity.");
            if (n_names != 1)
                throw std::runtime_error("Jplace document contains a pquery with key 'n' that is an array of size greater than one, while also having key 'm' for the multiplicity. This is not allowed.");
        }
        if (!n_valid_kind) throw std::runtime_error("Jplace document contains a pquery with key 'n' that is neither an array nor a string.");
        if (n_names == 0) throw std::runtime_error("Jplace document contains a pquery with key 'n' that does not contain any values.");
        if (!n_all_strings) throw std::runtime_error("Jplace document contains a pquery where key 'n' has a non-string field.");
        for (int64_t k = 0; k < n_names; ++k) d.mult.push_back(m_val);
    } else {
        d.mult.insert(d.mult.end(), nm_vals.begin(), nm_vals.end());
    }
    d.mult_off.push_back((int64_t)d.mult.size());
    d.pq_row_off.push_back((int64_t)d.row_off.size() - 1);
}

}  // namespace detail

inline Sample parse_jplace(const std::string& text, const std::string& source_name)
{
    using detail::RawDoc;
    RawDoc d;
    JsonScanner js(text.data(), text.data() + text.size(), source_name);
    if (js.type() != JsonScanner::Object) throw std::runtime_error("Json value is not a Json document in " + source_name + ".");
    js.expect('{');
    if (!js.try_char('}')) {
        for (;;) {
            const std::string key = js.string();
            js.expect(':');
            if (key == "tree" && !d.have_tree && js.type() == JsonScanner::String) { d.tree = js.string(); d.have_tree = true; }
            else if (key == "version" && d.version == -1) {
                if (js.type() == JsonScanner::String) { try { d.version = std::stoi(js.string()); } catch (...) { d.version = -1; } }
                else if (js.type() == JsonScanner::Number) { JsonNumber n = js.number(); d.version = (n.kind == JsonNumber::Unsigned) ? (int)n.u : -1; }
                else js.skip_value();
            } else if (key == "fields" && !d.have_fields && js.type() == JsonScanner::Array) {
                d.have_fields = true;
                js.expect('[');
                if (!js.try_char(']')) {
                    for (;;) {
                        if (js.type() != JsonScanner::String)
                            throw std::runtime_error(std::string("Jplace document contains a value of type '") + JsonScanner::type_name(js.type()) +
                                                     "' instead of a string with a field name at key 'fields'.");
                        d.fields.push_back(js.string());
                        if (js.try_char(',')) continue;
                        js.expect(']');
                        break;
                    }
                }
            } else if (key == "placements" && !d.have_placements && js.type() == JsonScanner::Array) {
                d.have_placements = true;
                js.expect('[');
                if (!js.try_char(']')) {
                    for (;;) {
                        detail::parse_pquery(js, d);
                        if (js.try_char(',')) continue;
                        js.expect(']');
                        break;
                    }
                }
            } else js.skip_value();
            if (js.try_char(',')) continue;
            js.expect('}');
            break;
        }
    }

    // ---- tree (jplace_reader.cpp:237-262) ----
    if (!d.have_tree) throw std::runtime_error("Jplace document does not contain a valid Newick tree at key 'tree'.");
    Sample s;
    s.version = d.version;
    s.tree = parse_newick(d.tree, d.version == 1);

    // ---- fields (jplace_reader.cpp:268-340) ----
    if (!d.have_fields) throw std::runtime_error("Jplace document does not contain field names at key 'fields'.");
    int f_edge = -1, f_lik = -1, f_lwr = -1, f_dist = -1, f_pend = -1, f_prox = -1;
    for (size_t i = 0; i < d.fields.size(); ++i) {
        const std::string& f = d.fields[i];
        int* slot = nullptr;
        if (f == "edge_num") slot = &f_edge; else if (f == "likelihood") slot = &f_lik;
        else if (f == "like_weight_ratio") slot = &f_lwr; else if (f == "distal_length") slot = &f_dist;
        else if (f == "pendant_length") slot = &f_pend; else if (f == "proximal_length") slot = &f_prox;
        if (slot) {
            if (*slot >= 0) throw std::runtime_error("Jplace document contains field name '" + f + "' more than once at key 'fields'.");
            *slot = (int)i;
        }
    }
    const std::pair<const char*, int> required[] = {{"edge_num", f_edge}, {"likelihood", f_lik}, {"like_weight_ratio", f_lwr}, {"pendant_length", f_pend}};
    for (const auto& r : required)
        if (r.second < 0) throw std::runtime_error(std::string("Jplace document does not contain necessary field '") + r.first + "' at key 'fields'.");
    if (f_dist < 0 && f_prox < 0)
        throw std::runtime_error("Jplace document does not contain one of the necessary fields 'distal_length' or 'proximal_length' at key 'fields'.");

    // ---- placements (jplace_reader.cpp:346-567) ----
    if (!d.have_placements) throw std::runtime_error("Jplace document does not contain pqueries at key 'placements'.");
    std::unordered_map<uint64_t, int32_t> by_num;
    by_num.reserve((size_t)s.tree.n_edges * 2);
    for (int32_t e = 0; e < s.tree.n_edges; ++e) {
        if (!by_num.emplace((uint64_t)s.tree.edge_num[e], e).second)
            throw std::runtime_error("Jplace document contains a tree where the edge_num tag '" + std::to_string(s.tree.edge_num[e]) +
                                     "' is used more than once, and hence cannot be used to uniquely identify edges of the placements. "
                                     "This indicates a severe issue with the program that created the jplace file.");
    }
    const size_t W = d.fields.size();
    const int64_t n_rows = (int64_t)d.row_off.size() - 1;
    s.placements.resize((size_t)n_rows);
    auto number_at = [&](int64_t row, int col, const char* name) -> double {
        const detail::RawCell& c = d.cells[(size_t)(d.row_off[(size_t)row] + col)];
        if (!c.is_number)
            throw std::runtime_error(std::string("Jplace document contains a pquery where the field ") + name + " is of type '" +
                                     JsonScanner::type_name(c.type) + "' instead of a number.");
        return c.num.as_double();
    };
    for (int64_t r = 0; r < n_rows; ++r) {
        if ((size_t)(d.row_off[(size_t)r + 1] - d.row_off[(size_t)r]) != W)
            throw std::runtime_error("Jplace document contains a placement fields array with different size than the fields name array.");
        Placement pl;
        const detail::RawCell& ce = d.cells[(size_t)(d.row_off[(size_t)r] + f_edge)];
        if (!ce.is_number) throw std::runtime_error("Jplace document contains a pquery where field 'edge_num' is not a number.");
        const uint64_t en = ce.num.as_size();
        auto it = by_num.find(en);
        if (it == by_num.end())
            throw std::runtime_error("Jplace document contains a pquery where field 'edge_num' has value '" + std::to_string(en) +
                                     "', which does not correspond to any edge_num in the given Newick tree of the document.");
        pl.edge = it->second;
        pl.likelihood = number_at(r, f_lik, "likelihood");
        pl.lwr = number_at(r, f_lwr, "like_weight_ratio");
        pl.pendant = number_at(r, f_pend, "pendant_length");
        double distal = -1.0;
        if (f_dist >= 0) distal = number_at(r, f_dist, "distal_length");
        if (f_prox >= 0) pl.proximal = number_at(r, f_prox, "proximal_length");
        if (distal >= 0.0 && pl.proximal == 0.0) pl.proximal = s.tree.bl[(size_t)pl.edge] - distal;
        // InvalidNumberBehaviour::kIgnore is the reference's default (jplace_reader.hpp:316): out-of-range
        // LWR / lengths pass through unchanged.
        s.placements[(size_t)r] = pl;
    }
    s.pq_off = std::move(d.pq_row_off);
    s.mult = std::move(d.mult);
    s.mult_off = std::move(d.mult_off);
    return s;
}

inline Sample read_jplace_file(const std::string& path) { return parse_jplace(read_maybe_gzip(path), path); }

}  // namespace krdhost
