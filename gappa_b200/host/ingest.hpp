// --jplace-path resolution and the parallel load of all samples into one MassSet.
//
// Path handling follows gappa src/options/file_input.cpp:83-156: files are taken as given,
// directories contribute their `*.jplace` / `*.jplace.gz` entries, the pooled list is sorted
// (reproducible sample order), and a sample's label is its base name without ".gz" and without the
// last extension.  Files are parsed in parallel (the reference: OpenMP dynamic over files,
// src/options/jplace_input.cpp:308-321).
#pragma once
#include <dirent.h>
#include <sys/stat.h>

#include <atomic>
#include <exception>
#include <mutex>
#include <thread>

#include "masses.hpp"

namespace krdhost {

inline bool is_dir_path(const std::string& p) { struct stat st; return ::stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode); }
inline bool is_file_path(const std::string& p) { struct stat st; return ::stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode); }

inline bool has_suffix(const std::string& s, const std::string& suf)
{
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}

inline std::vector<std::string> resolve_jplace_paths(const std::vector<std::string>& raw)
{
    std::vector<std::string> out;
    for (const auto& path : raw) {
        if (is_file_path(path)) out.push_back(path);
        else if (is_dir_path(path)) {
            std::string dir = path;
            while (!dir.empty() && dir.back() == '/') dir.pop_back();
            dir += '/';
            DIR* dp = opendir(path.c_str());
            if (!dp) throw std::runtime_error("Cannot open directory '" + path + "'.");
            while (dirent* de = readdir(dp)) {
                const std::string fn = de->d_name;
                if (fn == "." || fn == "..") continue;
                // regex ".*\.jplace(\.gz)?$"
                if (!(has_suffix(fn, ".jplace") || has_suffix(fn, ".jplace.gz"))) continue;
                if (is_file_path(dir + fn)) out.push_back(dir + fn);
            }
            closedir(dp);
        } else throw std::invalid_argument("--jplace-path: Not a valid file or directory: " + path);
    }
    if (out.empty()) throw std::invalid_argument("--jplace-path: No files found.");
    std::sort(out.begin(), out.end());
    return out;
}

inline std::string sample_label(const std::string& path)
{
    std::string fn = path;
    const size_t slash = fn.find_last_of("\\/");
    if (slash != std::string::npos) fn.erase(0, slash + 1);
    if (has_suffix(fn, ".gz")) fn.erase(fn.size() - 3);
    const size_t dot = fn.rfind('.');
    if (dot != 0 && dot != std::string::npos) fn.erase(dot);
    return fn;
}

struct IngestOptions {
    bool point_mass = false;
    bool ignore_multiplicities = false;
    int threads = 1;
};

inline MassSet load_mass_set(const std::vector<std::string>& files, const IngestOptions& opt)
{
    const size_t n = files.size();
    std::vector<FlatTree> trees(n);
    std::vector<SampleMasses> masses(n);
    std::atomic<size_t> next{0};
    std::exception_ptr err;
    std::mutex err_mu;
    auto worker = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= n) return;
            try {
                Sample s = read_jplace_file(files[i]);
                if (opt.point_mass) apply_point_mass(s);
                if (opt.ignore_multiplicities) apply_ignore_multiplicities(s);
                masses[i] = sample_to_masses(s, true);
                trees[i] = std::move(s.tree);
            } catch (...) {
                std::lock_guard<std::mutex> g(err_mu);
                if (!err) err = std::current_exception();
                next.store(n);
                return;
            }
        }
    };
    const int T = std::max(1, std::min<int>(opt.threads, (int)n));
    if (T == 1) worker();
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < T; ++t) pool.emplace_back(worker);
        for (auto& th : pool) th.join();
    }
    if (err) std::rethrow_exception(err);
    return assemble(trees, masses);
}

}  // namespace krdhost
