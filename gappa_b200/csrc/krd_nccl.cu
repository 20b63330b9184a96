// The north-star's collective inside the product: one rank's share of the tiles, an NCCL all-gather of the packed tiles over
// NVLink / NVSwitch, and the unpack into the full (or a row block of the) matrix — behind the C ABI, for callers that own an
// NCCL communicator (one process per GPU) and for `gappa-b200 --gpus G --exchange nccl` (one process, one thread per GPU,
// communicators from krd_nccl_init_all).  NCCL is loaded at run time (dlopen): libkrd_b200.so itself has no link-time
// dependency on it, single-GPU users never touch it, and a box without NCCL gets a clear error instead of a load failure.
#include <dlfcn.h>

#include <mutex>

#include "krd_common.cuh"

namespace krd {

// the few NCCL entry points used, with the documented C signatures (nccl.h); ncclDataType_t ncclFloat64 = 8
typedef int (*fn_all_gather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_comm_init_all)(void**, int, const int*);
typedef int (*fn_comm_destroy)(void*);
typedef const char* (*fn_get_error_string)(int);
typedef int (*fn_group)(void);

struct NcclApi {
    void* lib = nullptr;
    fn_all_gather all_gather = nullptr;
    fn_comm_init_all comm_init_all = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_get_error_string error_string = nullptr;
    fn_group group_start = nullptr, group_end = nullptr;
    std::string err;
};

static NcclApi& nccl()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* override_path = std::getenv("KRD_NCCL_LIB");
        // a process that already carries an NCCL (torch bundles one) must use THAT copy: RTLD_NOLOAD finds it
        const char* names[] = {override_path, "libnccl.so.2", "libnccl.so"};
        for (const char* nme : names) {
            if (!nme) continue;
            if ((api.lib = dlopen(nme, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL))) break;
        }
        for (const char* nme : names) {
            if (api.lib || !nme) continue;
            api.lib = dlopen(nme, RTLD_NOW | RTLD_GLOBAL);
        }
        if (!api.lib) { api.err = std::string("NCCL is not available: ") + (dlerror() ? dlerror() : "libnccl.so.2 not found"); return; }
        api.all_gather = (fn_all_gather)dlsym(api.lib, "ncclAllGather");
        api.comm_init_all = (fn_comm_init_all)dlsym(api.lib, "ncclCommInitAll");
        api.comm_destroy = (fn_comm_destroy)dlsym(api.lib, "ncclCommDestroy");
        api.error_string = (fn_get_error_string)dlsym(api.lib, "ncclGetErrorString");
        api.group_start = (fn_group)dlsym(api.lib, "ncclGroupStart");
        api.group_end = (fn_group)dlsym(api.lib, "ncclGroupEnd");
        if (!api.all_gather || !api.comm_init_all || !api.comm_destroy) api.err = "NCCL library lacks ncclAllGather / ncclCommInitAll";
    });
    return api;
}

static krd_status nccl_fail(krd_handle* h, int rc, const char* what)
{
    NcclApi& a = nccl();
    return fail(h, KRD_ERR_CUDA, std::string("NCCL error in ") + what + ": " + (a.error_string ? a.error_string(rc) : "unknown"));
}

}  // namespace krd

using namespace krd;

extern "C" {

krd_status krd_compute_allgather(krd_handle* h, double p, double divisor, int32_t rank, int32_t world, void* nccl_comm,
                                 int32_t row0, int32_t row1, double* d_out)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!nccl_comm || !d_out) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null communicator / output");
    if (world < 1 || rank < 0 || rank >= world) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad rank/world");
    NcclApi& a = nccl();
    if (!a.err.empty()) return fail(h, KRD_ERR_STATE, a.err);
    const int64_t shard = krd_shard_doubles(h, world);
    if (shard < 0) return fail(h, KRD_ERR_STATE, "krd_set_tree and krd_set_masses must be called first");
    if (row0 < 0 || row1 < row0 || row1 > h->N) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad row range");
    KRD_CUDA(h, cudaSetDevice(h->device));
    krd_status st;
    if ((st = reserve(h, (void**)&h->d_shard, (size_t)std::max<int64_t>(shard, 1) * 8)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_gathered, (size_t)std::max<int64_t>(shard, 1) * 8 * (size_t)world)) != KRD_OK) return st;
    if ((st = krd_compute_shard(h, p, divisor, rank, world, h->d_shard)) != KRD_OK) return st;
    if (shard == 0) return KRD_OK;
    const int rc = a.all_gather(h->d_shard, h->d_gathered, (size_t)shard, 8 /* ncclFloat64 */, nccl_comm, h->stream);
    if (rc != 0) return nccl_fail(h, rc, "ncclAllGather");
    return (row0 == 0 && row1 == h->N) ? krd_unpack_shards(h, world, h->d_gathered, d_out)
                                       : krd_unpack_shards_rows(h, world, h->d_gathered, row0, row1, d_out);
}

krd_status krd_compute_allgather_host(krd_handle* h, double p, double divisor, int32_t rank, int32_t world, void* nccl_comm,
                                      int32_t row0, int32_t row1, double* out_rows)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!out_rows) return fail(h, KRD_ERR_INVALID_ARGUMENT, "null output");
    if (row0 < 0 || row1 < row0 || row1 > h->N) return fail(h, KRD_ERR_INVALID_ARGUMENT, "bad row range");
    KRD_CUDA(h, cudaSetDevice(h->device));
    const size_t bytes = (size_t)(row1 - row0) * (size_t)h->N * 8;
    if (bytes > h->out_tmp_capacity) {
        cudaFree(h->d_out_tmp); h->d_out_tmp = nullptr; h->out_tmp_capacity = 0;
        krd_status sa = alloc_sized(h, (void**)&h->d_out_tmp, bytes, "this rank's row block of the distance matrix");
        if (sa != KRD_OK) return sa;
        h->out_tmp_capacity = bytes;
    }
    krd_status st = krd_compute_allgather(h, p, divisor, rank, world, nccl_comm, row0, row1, h->d_out_tmp);
    if (st != KRD_OK) return st;
    if (bytes) KRD_CUDA(h, cudaMemcpyAsync(out_rows, h->d_out_tmp, bytes, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    return KRD_OK;
}

krd_status krd_nccl_init_all(int32_t n, const int32_t* devices, void** comms)
{
    if (n < 1 || !devices || !comms) return KRD_ERR_INVALID_ARGUMENT;
    NcclApi& a = nccl();
    if (!a.err.empty()) return fail(nullptr, KRD_ERR_STATE, a.err);
    const int rc = a.comm_init_all(comms, n, devices);
    if (rc != 0) return nccl_fail(nullptr, rc, "ncclCommInitAll");
    return KRD_OK;
}

void krd_nccl_destroy(int32_t n, void** comms)
{
    NcclApi& a = nccl();
    if (!a.comm_destroy || !comms) return;
    for (int i = 0; i < n; ++i) if (comms[i]) { a.comm_destroy(comms[i]); comms[i] = nullptr; }
}

}  // extern "C"
