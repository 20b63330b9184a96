// C ABI of libkrd_b200.so — see include/krd_b200.h.  No exception leaves this file.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>

#include "krd_common.cuh"

namespace krd {

static thread_local std::string g_create_error;

krd_status fail(krd_handle* h, krd_status st, const std::string& msg)
{
    if (h) h->err = msg; else g_create_error = msg;
    return st;
}

krd_status cuda_fail(krd_handle* h, cudaError_t e, const char* what)
{
    std::string m = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
    cudaGetLastError();  // clear sticky-less errors
    return fail(h, KRD_ERR_CUDA, m);
}

// cudaMalloc of one of the large buffers with a message that says what did not fit (a bare "out of memory" is useless
// to someone who ran a cohort the CPU path can hold); the CUDA error state is cleared.
krd_status alloc_sized(krd_handle* h, void** p, size_t bytes, const char* what)
{
    const cudaError_t e = cudaMalloc(p, bytes ? bytes : 16);
    if (e == cudaSuccess) return KRD_OK;
    *p = nullptr;
    cudaGetLastError();
    size_t free_b = 0, total_b = 0;
    cudaMemGetInfo(&free_b, &total_b);
    cudaGetLastError();
    char msg[320];
    std::snprintf(msg, sizeof msg, "%s needs %.2f GB of device memory, %.2f GB of %.2f GB are free (%s)", what, bytes / 1e9,
                  free_b / 1e9, total_b / 1e9, cudaGetErrorString(e));
    return fail(h, KRD_ERR_CUDA, msg);
}

krd_status reserve(krd_handle* h, void** p, size_t bytes)
{
    if (bytes == 0) bytes = 16;
    auto it = h->caps.find(p);
    if (*p && it != h->caps.end() && it->second >= bytes) return KRD_OK;
    if (h->capturing) return fail(h, KRD_ERR_STATE, "buffer growth inside a captured phase");   // (the phase is re-issued launch by launch)
    if (*p) { cudaFree(*p); *p = nullptr; }
    h->caps[p] = 0;
    KRD_CUDA(h, cudaMalloc(p, bytes));
    h->caps[p] = bytes;
    return KRD_OK;
}

// A pipelined krd_set_masses may still be uploading blocks on the copy stream: before the state is dropped and the
// buffers are written again (or the caller's arrays look free), wait for it.
static void pipe_quiesce(krd_handle* h)
{
    if (h->piped && h->copy_stream) cudaStreamSynchronize(h->copy_stream);
}

static void free_masses(krd_handle* h)   // buffers are kept for reuse; only the state is dropped
{
    h->d_edge = nullptr; h->d_pos = nullptr; h->d_mass = nullptr; h->d_cell32 = nullptr; h->binned_input = false; h->cell_bytes = 4;
    h->have_masses = false; h->F_valid = false;
}

static krd_status set_masses_common(krd_handle* h, int32_t n_samples, const int64_t* sample_off)
{
    if (!h->have_tree) return fail(h, KRD_ERR_STATE, "krd_set_tree must be called before krd_set_masses");
    if (n_samples < 0 || !sample_off) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad sample arguments");
    if (sample_off[0] != 0) return fail(h, KRD_ERR_INVALID_ARGUMENT, "sample_off[0] must be 0");
    for (int i = 0; i < n_samples; ++i)
        if (sample_off[i + 1] < sample_off[i]) return fail(h, KRD_ERR_INVALID_ARGUMENT, "sample_off must be non-decreasing");
    return KRD_OK;
}

}  // namespace krd

using namespace krd;

extern "C" {

krd_status krd_create(const krd_config* cfg, krd_handle** out)
{
    if (!out) return KRD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        return fail(nullptr, KRD_ERR_NO_DEVICE, "no CUDA device available (this library has no CPU fallback)");
    }
    krd_config c{};
    if (cfg) c = *cfg;
    if (c.device < 0 || c.device >= n_dev) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "device ordinal out of range");
    if (c.bins < 0) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "bins must be >= 0");
    if (c.max_f_mb < 0) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "max_f_mb must be >= 0");
    cudaDeviceProp prop;
    if ((e = cudaSetDevice(c.device)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, c.device)) != cudaSuccess)
        return cuda_fail(nullptr, e, "cudaSetDevice");
    if (prop.major != 10) return fail(nullptr, KRD_ERR_NO_DEVICE, "device is not sm_100 (this library is built for B200 only)");
    krd_handle* h = new (std::nothrow) krd_handle();
    if (!h) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "out of host memory");
    h->cfg = c; h->device = c.device; h->num_sms = prop.multiProcessorCount;
    for (auto& ev : h->ev) cudaEventCreate(&ev);
    cudaEventCreateWithFlags(&h->ev_copy, cudaEventDisableTiming);
    *out = h;
    return KRD_OK;
}

void krd_destroy(krd_handle* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    pipe_destroy(h);
    free_masses(h);
    for (void* q : h->d_owned) cudaFree(q);
    cudaFree(h->d_off);
    cudaFree(h->d_out_tmp_sq); cudaFree(h->d_squash); cudaFree(h->d_kmeans); cudaFree(h->d_fix); cudaFree(h->d_Fg); cudaFree(h->d_norm); cudaFree(h->d_norm_part); cudaFree(h->d_top); cudaFree(h->d_lvl_edges); cudaFree(h->d_kid_off); cudaFree(h->d_kid_edge); cudaFree(h->d_bl);
    cudaFree(h->d_sid); cudaFree(h->d_slot); cudaFree(h->d_edge_dec); cudaFree(h->d_grow); cudaFree(h->d_cum); cudaFree(h->d_acc);
    for (cudaEvent_t ev : h->slab_ev) cudaEventDestroy(ev); cudaFree(h->d_seg_off); cudaFree(h->d_len); cudaFree(h->d_sqrt_len); cudaFree(h->d_F);
    cudaFree(h->d_cell); cudaFree(h->d_tmass); cudaFree(h->d_tile_ptr); cudaFree(h->d_M); cudaFree(h->d_D);
    for (void* q : h->d_scratch) cudaFree(q);
    cudaFree(h->d_flags);
    cudaFree(h->d_ws); for (auto& kv : h->scheds) cudaFree(kv.second.d); cudaFree(h->d_tiles); cudaFree(h->d_out_tmp); cudaFree(h->d_shard); cudaFree(h->d_gathered);
    for (auto& ev : h->ev) if (ev) cudaEventDestroy(ev);
    if (h->ev_copy) cudaEventDestroy(h->ev_copy);
    delete h;
}

const char* krd_last_error(const krd_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

krd_status krd_set_stream(krd_handle* h, void* cuda_stream)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    h->stream = static_cast<cudaStream_t>(cuda_stream);
    return KRD_OK;
}

krd_status krd_set_tree(krd_handle* h, int32_t n_edges, const int32_t* parent_node, const int32_t* child_node,
                        const double* branch_length, int32_t root_node)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (n_edges < 0 || (n_edges > 0 && (!parent_node || !child_node || !branch_length)))
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad tree arguments");
    KRD_CUDA(h, cudaSetDevice(h->device));
    h->E = n_edges; h->root = root_node;
    h->parent.assign(parent_node, parent_node + n_edges);
    h->child.assign(child_node, child_node + n_edges);
    h->bl.assign(branch_length, branch_length + n_edges);
    pipe_quiesce(h);
    h->have_tree = false; h->F_valid = false; h->have_masses = false; h->piped = false; ++h->tree_version;
    krd_status st = build_tree(h);
    if (st == KRD_OK) h->have_tree = true;
    return st;
}

// The compact form needs bins >= 1 and every cell id edge * bins + bin representable in the cell type.
static krd_status check_cell_range(krd_handle* h, int cell_bytes)
{
    if (cell_bytes == 0) return KRD_OK;
    const int64_t cells = (int64_t)h->E * h->cfg.bins;
    if (h->cfg.bins < 1 || cells > 0xFFFFFFF0ll)
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_set_masses_binned needs krd_config.bins >= 1 and n_edges * bins < 2^32");
    if (cell_bytes == 2 && cells > 65536)
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_set_masses_binned16 needs n_edges * bins <= 65536 (" + std::to_string(cells) +
                                                  " here): use krd_set_masses_binned");
    return KRD_OK;
}

// Host arrays -> HBM + K0.  cell_bytes 4 / 2: the compact binned form (cell = edge * bins + bin as uint32_t / uint16_t,
// 12 / 10 B per entry); 0: (edge, pos, mass), 20 B per entry.
static krd_status set_masses_host(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* edge,
                                  const double* pos, const double* mass, const void* cell32, int cell_bytes)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    const bool binned = cell_bytes != 0;
    krd_status st = set_masses_common(h, n_samples, sample_off);
    if (st != KRD_OK) return st;
    const int64_t M = sample_off[n_samples];
    if (M > 0 && (!mass || (binned ? !cell32 : (!edge || !pos)))) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null mass arrays");
    if ((st = check_cell_range(h, cell_bytes)) != KRD_OK) return st;
    KRD_CUDA(h, cudaSetDevice(h->device));
    pipe_quiesce(h);
    free_masses(h);
    if (pipe_wanted(h, n_samples)) return pipe_set_masses(h, n_samples, sample_off, edge, pos, mass, cell32, cell_bytes);
    if (h->piped || h->custom_tiles) { h->piped = false; h->custom_tiles = false; h->n_tiles_alloc = -1; }
    h->N = n_samples; h->M = M;
    const size_t Mx = (size_t)std::max<int64_t>(M, 1);
    if ((st = reserve(h, (void**)&h->d_off, (size_t)(n_samples + 1) * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[0], Mx * 4)) != KRD_OK) return st;
    if (!binned && (st = reserve(h, &h->d_owned[1], Mx * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[2], Mx * 8)) != KRD_OK) return st;
    h->d_mass = (double*)h->d_owned[2];
    KRD_CUDA(h, cudaMemcpyAsync(h->d_off, sample_off, (size_t)(n_samples + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    if (binned) {
        h->binned_input = true; h->d_cell32 = h->d_owned[0]; h->cell_bytes = cell_bytes;
        KRD_CUDA(h, cudaMemcpyAsync(h->d_owned[0], cell32, (size_t)M * cell_bytes, cudaMemcpyHostToDevice, h->stream));
    } else {
        h->d_edge = (int32_t*)h->d_owned[0]; h->d_pos = (double*)h->d_owned[1];
        KRD_CUDA(h, cudaMemcpyAsync(h->d_edge, edge, (size_t)M * 4, cudaMemcpyHostToDevice, h->stream));
        KRD_CUDA(h, cudaMemcpyAsync(h->d_pos, pos, (size_t)M * 8, cudaMemcpyHostToDevice, h->stream));
    }
    KRD_CUDA(h, cudaMemcpyAsync(h->d_mass, mass, (size_t)M * 8, cudaMemcpyHostToDevice, h->stream));
    st = prepare_masses(h);
    if (st != KRD_OK) return st;
    h->have_masses = true;
    h->stats.n_samples = n_samples; h->stats.n_edges = h->E; h->stats.n_masses = M;
    return KRD_OK;
}

krd_status krd_set_masses(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* edge,
                          const double* pos, const double* mass)
{
    return set_masses_host(h, n_samples, sample_off, edge, pos, mass, nullptr, 0);
}

krd_status krd_set_masses_binned(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint32_t* cell,
                                 const double* mass)
{
    return set_masses_host(h, n_samples, sample_off, nullptr, nullptr, mass, cell, 4);
}

krd_status krd_set_masses_binned16(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint16_t* cell,
                                   const double* mass)
{
    return set_masses_host(h, n_samples, sample_off, nullptr, nullptr, mass, cell, 2);
}

}  // extern "C"

namespace krd {
krd_status set_masses_device_impl(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* d_edge,
                                  const double* d_pos, const double* d_mass, const void* d_cell32, int cell_bytes)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    krd_status st = set_masses_common(h, n_samples, sample_off);
    if (st != KRD_OK) return st;
    const bool binned = cell_bytes != 0;
    if ((st = check_cell_range(h, cell_bytes)) != KRD_OK) return st;
    KRD_CUDA(h, cudaSetDevice(h->device));
    pipe_quiesce(h);
    free_masses(h);
    if (h->piped || h->custom_tiles) { h->piped = false; h->custom_tiles = false; h->n_tiles_alloc = -1; }
    h->N = n_samples; h->M = sample_off[n_samples];
    if ((st = reserve(h, (void**)&h->d_off, (size_t)(n_samples + 1) * 8)) != KRD_OK) return st;
    if (h->d_off_from) launch_rebase_off(h->d_off_from, n_samples + 1, h->d_off, h->stream);
    else KRD_CUDA(h, cudaMemcpyAsync(h->d_off, sample_off, (size_t)(n_samples + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    h->d_edge = const_cast<int32_t*>(d_edge); h->d_pos = const_cast<double*>(d_pos); h->d_mass = const_cast<double*>(d_mass);
    h->binned_input = binned; h->d_cell32 = d_cell32; h->cell_bytes = binned ? cell_bytes : 4;
    st = prepare_masses(h);
    if (st != KRD_OK) return st;
    h->have_masses = true;
    h->stats.n_samples = n_samples; h->stats.n_edges = h->E; h->stats.n_masses = h->M;
    return KRD_OK;
}
}  // namespace krd

extern "C" {

krd_status krd_set_masses_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* d_edge,
                                 const double* d_pos, const double* d_mass)
{
    return set_masses_device_impl(h, n_samples, sample_off, d_edge, d_pos, d_mass, nullptr, 0);
}

krd_status krd_set_masses_binned_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint32_t* d_cell,
                                        const double* d_mass)
{
    return set_masses_device_impl(h, n_samples, sample_off, nullptr, nullptr, d_mass, d_cell, 4);
}

krd_status krd_set_masses_binned16_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint16_t* d_cell,
                                          const double* d_mass)
{
    return set_masses_device_impl(h, n_samples, sample_off, nullptr, nullptr, d_mass, d_cell, 2);
}

static krd_status check_compute(krd_handle* h, double p, double divisor)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!(p > 0.0))  // also rejects NaN; genesis emd.cpp:186-190
        return fail(h, KRD_ERR_INVALID_EXPONENT, "Invalid exponent value p for earth mover's distance calculation. Has to be > 0.0.");
    if (!h->have_tree || !h->have_masses) return fail(h, KRD_ERR_STATE, "krd_set_tree and krd_set_masses must be called first");
    if (divisor == 0.0 || divisor != divisor) return fail(h, KRD_ERR_INVALID_ARGUMENT, "divisor must be a non-zero number");
    return KRD_OK;
}

static krd_status compute_impl(krd_handle* h, double p, double divisor, int rank, int world, double* d_out, double* d_packed)
{
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped) {
        if (world != 1 || d_packed) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }   // sharding works on the ordinary tile order
        else {
            krd_status ps = pipe_compute(h, p, divisor, d_out);
            if (ps != KRD_OK || h->piped) return ps;
            // (a late child was not in MassTree order: the handle has been flattened, continue the ordinary way)
        }
    }
    if (h->streamed) return compute_streamed(h, p, divisor, rank, world, d_out, d_packed);
    h->stats.launches = 0;
    h->slabs_total = 0;
    cudaEventRecord(h->ev[0], h->stream);
    if (!h->F_valid) {
        krd_status st = build_F(h);
        if (st != KRD_OK) return st;
        h->stats.launches += h->build_launches;
    }
    cudaEventRecord(h->ev[1], h->stream);
    krd_status st = run_pairs(h, p, divisor, rank, world, d_out, d_packed);
    if (st != KRD_OK) return st;
    cudaEventRecord(h->ev[3], h->stream);
    return KRD_OK;
}

krd_status krd_compute_device(krd_handle* h, double p, double divisor, double* d_out)
{
    krd_status st = check_compute(h, p, divisor);
    if (st != KRD_OK) return st;
    if (!d_out && h->N > 0) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    if (h->N == 0) return KRD_OK;
    return compute_impl(h, p, divisor, 0, 1, d_out, nullptr);
}

krd_status krd_compute(krd_handle* h, double p, double divisor, double* out)
{
    krd_status st = check_compute(h, p, divisor);
    if (st != KRD_OK) return st;
    if (h->N == 0) return KRD_OK;
    if (!out) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    KRD_CUDA(h, cudaSetDevice(h->device));    // (the scratch matrix below must live on THIS handle's device)
    const size_t bytes = (size_t)h->N * h->N * 8;
    if (bytes > h->out_tmp_capacity) {
        cudaFree(h->d_out_tmp); h->d_out_tmp = nullptr; h->out_tmp_capacity = 0;
        if ((st = alloc_sized(h, (void**)&h->d_out_tmp, bytes, "the n x n distance matrix")) != KRD_OK) return st;
        h->out_tmp_capacity = bytes;
    }
    st = compute_impl(h, p, divisor, 0, 1, h->d_out_tmp, nullptr);
    if (st != KRD_OK) return st;
    KRD_CUDA(h, cudaMemcpyAsync(out, h->d_out_tmp, bytes, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    return KRD_OK;
}

int64_t krd_shard_doubles(const krd_handle* h, int32_t world)
{
    if (!h || world < 1 || !h->have_masses) return -1;
    const int64_t T = (h->N + TILE - 1) / TILE;
    return tiles_per_rank(T * (T + 1) / 2, world) * TILE * TILE;
}

krd_status krd_compute_shard(krd_handle* h, double p, double divisor, int32_t rank, int32_t world, double* d_packed)
{
    krd_status st = check_compute(h, p, divisor);
    if (st != KRD_OK) return st;
    if (world < 1 || rank < 0 || rank >= world) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad rank/world");
    if (!d_packed) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    return compute_impl(h, p, divisor, rank, world, nullptr, d_packed);
}

krd_status krd_compute_tile_range(krd_handle* h, double p, double divisor, int64_t t0, int64_t t1, double* d_packed)
{
    krd_status st = check_compute(h, p, divisor);
    if (st != KRD_OK) return st;
    if (!d_packed) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    if (t0 < 0 || t1 < t0 || t1 > krd_tile_count(h->N)) return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_compute_tile_range: bad tile range");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }
    if (t0 == t1) return KRD_OK;
    h->force_t0 = t0; h->force_t1 = t1;
    st = compute_impl(h, p, divisor, 0, 1, nullptr, d_packed);
    h->force_t0 = h->force_t1 = -1;
    return st;
}

krd_status krd_compute_shard_host(krd_handle* h, double p, double divisor, int32_t rank, int32_t world, double* packed)
{
    krd_status st = check_compute(h, p, divisor);
    if (st != KRD_OK) return st;
    if (world < 1 || rank < 0 || rank >= world) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad rank/world");
    if (!packed) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    const size_t bytes = (size_t)krd_shard_doubles(h, world) * 8;
    if (bytes == 0) return KRD_OK;
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (bytes > h->out_tmp_capacity) {
        cudaFree(h->d_out_tmp); h->d_out_tmp = nullptr; h->out_tmp_capacity = 0;
        KRD_CUDA(h, cudaMalloc(&h->d_out_tmp, bytes));
        h->out_tmp_capacity = bytes;
    }
    st = compute_impl(h, p, divisor, rank, world, nullptr, h->d_out_tmp);
    if (st != KRD_OK) return st;
    KRD_CUDA(h, cudaMemcpyAsync(packed, h->d_out_tmp, bytes, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    return KRD_OK;
}

krd_status krd_unpack_shards_host(int32_t n_samples, int32_t world, const double* packed_all, double* out)
{
    if (n_samples < 0 || world < 1 || (n_samples > 0 && (!packed_all || !out))) return KRD_ERR_INVALID_ARGUMENT;
    const int64_t n_tiles = krd_tile_count(n_samples), per_rank = tiles_per_rank(n_tiles, world), N = n_samples;
    std::vector<int32_t> ti((size_t)n_tiles), tj((size_t)n_tiles);
    krd_tile_list(n_samples, ti.data(), tj.data());
    for (int64_t t = 0; t < n_tiles; ++t) {
        // rank r holds tiles [r*per_rank, (r+1)*per_rank) back to back, ranks concatenated in order: tile t at offset t
        (void)per_rank;
        const double* src = packed_all + (size_t)t * (TILE * TILE);
        const int64_t i0 = (int64_t)ti[(size_t)t] * TILE, j0 = (int64_t)tj[(size_t)t] * TILE;
        for (int64_t r = 0; r < TILE && i0 + r < N; ++r)
            for (int64_t c = 0; c < TILE && j0 + c < N; ++c) {
                const int64_t i = i0 + r, j = j0 + c;
                if (i < j) { const double v = src[r * TILE + c]; out[i * N + j] = v; out[j * N + i] = v; }
                else if (i == j) out[i * N + i] = 0.0;
            }
    }
    return KRD_OK;
}

krd_status krd_unpack_shards(krd_handle* h, int32_t world, const double* d_gathered, double* d_out)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (world < 1 || !d_gathered || !d_out) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad arguments");
    KRD_CUDA(h, cudaSetDevice(h->device));
    return unpack_shards(h, world, d_gathered, d_out);
}

krd_status krd_unpack_shards_rows(krd_handle* h, int32_t world, const double* d_gathered, int32_t row0, int32_t row1, double* d_out_rows)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (world < 1 || !d_gathered || !d_out_rows || row0 < 0 || row1 < row0 || row1 > h->N) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad arguments");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }
    return unpack_shard_rows(h, d_gathered, row0, row1, d_out_rows);
}

krd_status krd_get_stats(const krd_handle* hc, krd_stats* out)
{
    if (!hc || !out) return KRD_ERR_INVALID_ARGUMENT;
    krd_handle* h = const_cast<krd_handle*>(hc);
    cudaSetDevice(h->device);
    if (h->ev[3] && cudaEventQuery(h->ev[3]) == cudaSuccess && h->stats.launches > 0 && h->stats.pair_ctas > 0) {
        cudaEventElapsedTime(&h->stats.ms_build, h->ev[0], h->ev[1]);
        cudaEventElapsedTime(&h->stats.ms_pairs, h->ev[1], h->ev[2]);
        cudaEventElapsedTime(&h->stats.ms_epilogue, h->ev[2], h->ev[3]);
        h->stats.ms_slabs = 0.f;
        if (h->streamed && h->slabs_total > 0) {   // ev[1]..ev[2] spans all slabs: take the slab builds out of the pair time
            float sum = 0.f, t = 0.f;
            for (int b = 0; b < h->slabs_timed; ++b)
                if (cudaEventElapsedTime(&t, h->slab_ev[(size_t)2 * b], h->slab_ev[(size_t)2 * b + 1]) == cudaSuccess) sum += t;
            h->stats.ms_slabs = sum * (float)h->slabs_total / (float)std::max(h->slabs_timed, 1);
            h->stats.ms_pairs = std::max(0.f, h->stats.ms_pairs - h->stats.ms_slabs);
        }
    }
    cudaGetLastError();
    *out = h->stats;
    return KRD_OK;
}

krd_status krd_squash(krd_handle* h, double p, krd_merger* out)
{
    krd_status st = check_compute(h, p, 1.0);
    if (st != KRD_OK) return st;
    if (h->N < 2) return fail(h, KRD_ERR_TOO_FEW_SAMPLES, "Squash clustering needs at least 2 samples.");
    if (!out) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if ((st = pipe_flatten(h)) != KRD_OK) return st;
    return run_squash(h, p, out);
}

krd_status krd_kmeans_ex(krd_handle* h, int32_t k, int32_t max_iterations, const int32_t* init_centroids, const int32_t* init_assignments,
                         int32_t* assignments, double* distances, int32_t* iterations)
{
    krd_status st = check_compute(h, 1.0, 1.0);
    if (st != KRD_OK) return st;
    if (k > h->N) return fail(h, KRD_ERR_INVALID_ARGUMENT, "Cannot run Kmeans with more clusters (k) than data points.");   // kmeans.hpp:538-544
    if (k < 1) return fail(h, KRD_ERR_INVALID_ARGUMENT, "Cannot run Kmeans with k == 0.");                                 // kmeans.hpp:532-536
    if (max_iterations < 1) return fail(h, KRD_ERR_INVALID_ARGUMENT, "k-means: max_iterations must be >= 1");
    if (!init_centroids || !assignments || !distances) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null argument");
    for (int c = 0; c < k; ++c)
        if (init_centroids[c] < 0 || init_centroids[c] >= h->N) return fail(h, KRD_ERR_INVALID_ARGUMENT, "k-means: initial centroid is not a sample index");
    if (init_assignments)
        for (int d = 0; d < h->N; ++d)
            if (init_assignments[d] < 0 || init_assignments[d] >= k) return fail(h, KRD_ERR_INVALID_ARGUMENT, "k-means: initial assignment is not a cluster index");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if ((st = pipe_flatten(h)) != KRD_OK) return st;
    return run_kmeans(h, k, max_iterations, init_centroids, init_assignments, assignments, distances, iterations);
}

krd_status krd_kmeans(krd_handle* h, int32_t k, int32_t max_iterations, const int32_t* init_centroids, int32_t* assignments,
                      double* distances, int32_t* iterations)
{
    return krd_kmeans_ex(h, k, max_iterations, init_centroids, nullptr, assignments, distances, iterations);
}

krd_status krd_invalidate_F(krd_handle* h)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    h->F_valid = false; h->cum_valid = false;
    return KRD_OK;
}

krd_status krd_debug_copy_F(krd_handle* h, double* out_F, double* out_len)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped && !h->F_valid) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }
    if (!h->F_valid) { krd_status st = build_F(h); if (st != KRD_OK) return st; }
    if (out_F) KRD_CUDA(h, cudaMemcpyAsync(out_F, h->d_F, (size_t)h->S * h->ld * 8, cudaMemcpyDeviceToHost, h->stream));
    if (out_len) KRD_CUDA(h, cudaMemcpyAsync(out_len, h->d_len, (size_t)h->S * 8, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    return KRD_OK;
}

void* krd_host_alloc(int64_t bytes)
{
    void* p = nullptr;
    if (cudaMallocHost(&p, (size_t)std::max<int64_t>(bytes, 1)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}

void krd_host_free(void* p) { if (p) cudaFreeHost(p); }

}  // extern "C"
